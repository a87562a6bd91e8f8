"""bench.py's contract pieces that need no GPU: both arms describe the workload with the same string, the reference arm prints one
well-formed JSON line (rank 0) or nothing (other ranks), for the single-GPU frame and for the N > 1 view batch."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(extra, env_extra):
    env = dict(os.environ); env.update(env_extra)
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference"] + extra, capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0, r.stderr
    return r.stdout.strip()


def test_reference_arm_single_gpu_line():
    out = _run(["--config", "tiny", "--steps", "2", "--warmup", "1"], {"WORLD_SIZE": "1", "RANK": "0"})
    line = json.loads(out)
    assert line["impl"] == "reference" and line["metric"] == "Mtri/s" and line["unit"] == "Mtri/s" and line["higher_is_better"] is True
    assert line["value"] > 0 and line["cpu_baseline"]["kind"] == "port" and line["cpu_baseline"]["cores"] >= 1
    assert line["e2e"] == {"value": line["value"], "unit": "Mtri/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    sys.path.insert(0, ROOT)
    import argparse
    import bench
    from renderengine_b200 import scene as sc
    args = argparse.Namespace(shading="phong", samples=1, views_per_rank=64, config="tiny")
    assert line["config"]["workload"] == bench.workload_string(args, 1, sc.make_scene("tiny"), "frame")   # what the device arm prints too


def test_reference_arm_multi_gpu_ranks():
    assert _run(["--steps", "1", "--views-per-rank", "2"], {"WORLD_SIZE": "2", "RANK": "1"}) == ""       # only rank 0 works and prints
    line = json.loads(_run(["--steps", "1", "--views-per-rank", "2"], {"WORLD_SIZE": "2", "RANK": "0"}))
    assert line["impl"] == "reference" and line["n_gpus"] == 2 and "C5 view batch" in line["config"]["workload"] and line["value"] > 0
