"""Scene-ingest conventions (AssimpSceneLoader.cpp:93-361) restated in renderengine_b200/host/SceneIngest.h:
compiled with g++ and run on the CPU — it stages data into the facade's host-side systems and makes no device call."""
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_ingest_conventions(tmp_path):
    from renderengine_b200 import build
    lib = build.build()                                  # the facade links against libreb.so (never called here)
    exe = str(tmp_path / "test_ingest")
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-Wall", "-o", exe, os.path.join(ROOT, "tests", "cpp", "test_ingest.cpp"),
                           os.path.join(ROOT, "renderengine_b200", "host", "RenderEngine.cpp"), "-L" + os.path.dirname(lib), "-lreb",
                           "-Wl,-rpath," + os.path.dirname(lib)])
    r = subprocess.run([exe], capture_output=True, text=True, timeout=60)
    assert r.returncode == 0, r.stderr
    assert "ingest OK" in r.stdout
