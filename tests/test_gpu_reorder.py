"""The draw-list re-ordering (k_reorder.cu: Morton order of the triangles per draw, cache slots in first-use order, K1 walking
{slots, submission ID} records) only engages for lists of >= 262,144 triangles by default, i.e. in the full-size tests.  Here
the whole device-vs-oracle parity suite runs once more in a child process with the threshold at zero, so every scene —
instanced draws, clipped, binned, huge, textured, multisampled, translucent, split — goes through the re-ordered path too.
(The threshold is read once per process, hence the child.)"""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.gpu
def test_parity_suite_with_every_draw_list_reordered():
    env = dict(os.environ, REB_REORDER_MIN_TRIS="0")
    r = subprocess.run([sys.executable, "-m", "pytest", "tests/test_gpu_parity.py", "tests/test_gpu_round2.py", "-m", "gpu", "-q", "-x", "-p", "no:cacheprovider"],
                       cwd=ROOT, env=env, capture_output=True, text=True, timeout=1500)
    tail = "\n".join(r.stdout.splitlines()[-25:])
    assert r.returncode == 0, tail + r.stderr[-2000:]
    assert " passed" in tail and "failed" not in tail, tail
