"""Builds libreb.so (the C-ABI product library) in-tree with nvcc for sm_100a.

    python -m renderengine_b200.build
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "libreb.so")
SOURCES = ["k_vertex.cu", "k_setup.cu", "k_tile.cu", "k_translucent.cu", "reb_capi.cu", "reb_noalloc.cu", "k_reorder.cu"]
NVCC_FLAGS = ["-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "--fmad=false",
              "-Xcompiler", "-fPIC,-ffp-contract=off,-fno-fast-math", "-Xptxas", "-v", "--expt-relaxed-constexpr"]


def _stale():
    if not os.path.exists(OUT):
        return True
    t = os.path.getmtime(OUT)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [os.path.join(os.path.dirname(HERE), "include", "reb.h")]
    return any(os.path.getmtime(d) > t for d in deps)


def build_variant(name, defines):
    """Build renderengine_b200/variants/libreb_<name>.so with extra -D flags (kernel tuning experiments; load with REB_LIB=...)."""
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    vdir = os.path.join(HERE, "variants")
    os.makedirs(vdir, exist_ok=True)
    out = os.path.join(vdir, f"libreb_{name}.so")
    cmd = [nvcc] + [f for f in NVCC_FLAGS if f not in ("-Xptxas", "-v")] + ["-D" + d for d in defines] + ["-shared", "-o", out] + \
          [os.path.join(CSRC, s) for s in SOURCES]
    subprocess.check_call(cmd)
    return out


def build(force=False, verbose=False):
    if not (force or _stale()):
        return OUT
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    objs = []
    procs = []
    for src in SOURCES:
        obj = os.path.join(CSRC, src.replace(".cu", ".o"))
        objs.append(obj)
        cmd = [nvcc] + NVCC_FLAGS + ["-c", os.path.join(CSRC, src), "-o", obj]
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    log = []
    for src, p in procs:
        out, _ = p.communicate()
        log.append(out)
        if p.returncode != 0:
            sys.stderr.write(out)
            raise RuntimeError("nvcc failed on " + src)
    subprocess.check_call([nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", OUT] + objs)
    with open(os.path.join(CSRC, "ptxas.log"), "w") as f:
        f.write("\n".join(log))
    if verbose:
        print("\n".join(log))
    return OUT


def build_host(force=False):
    """Compile the C++ host facade + demo (renderengine_b200/host) with g++ against libreb.so -> host/reb_host_demo."""
    hdir = os.path.join(HERE, "host")
    out = os.path.join(hdir, "reb_host_demo")
    srcs = [os.path.join(hdir, f) for f in ("RenderEngine.cpp", "demo_main.cpp")]
    deps = srcs + [os.path.join(hdir, f) for f in ("RenderEngine.h", "reb_math.h", "demo_scene.h")] + [os.path.join(os.path.dirname(HERE), "include", "reb.h"), OUT]
    if not force and os.path.exists(out) and all(os.path.getmtime(d) <= os.path.getmtime(out) for d in deps):
        return out
    cmd = [os.environ.get("CXX", "g++"), "-O2", "-std=c++17", "-Wall", "-o", out] + srcs + ["-L" + HERE, "-lreb", "-Wl,-rpath," + HERE, "-Wl,-rpath,$ORIGIN/.."]
    subprocess.check_call(cmd)
    build_mgpu_demo(force)
    return out


def build_mgpu_demo(force=False):
    """host/reb_mgpu_demo: the C++ multi-GPU caller (one context per GPU + NCCL called directly), g++ against libreb.so, libcudart, libnccl."""
    hdir = os.path.join(HERE, "host")
    out = os.path.join(hdir, "reb_mgpu_demo")
    srcs = [os.path.join(hdir, f) for f in ("RenderEngine.cpp", "demo_mgpu.cpp")]
    deps = srcs + [os.path.join(hdir, f) for f in ("RenderEngine.h", "reb_math.h", "demo_scene.h")] + [os.path.join(os.path.dirname(HERE), "include", "reb.h"), OUT]
    if not force and os.path.exists(out) and all(os.path.getmtime(d) <= os.path.getmtime(out) for d in deps):
        return out
    cuda = os.environ.get("CUDA_HOME", "/usr/local/cuda")
    cmd = [os.environ.get("CXX", "g++"), "-O2", "-std=c++17", "-Wall", "-I" + os.path.join(cuda, "include"), "-o", out] + srcs + \
          ["-L" + HERE, "-lreb", "-L" + os.path.join(cuda, "lib64"), "-lcudart", "-lnccl", "-Wl,-rpath," + HERE, "-Wl,-rpath,$ORIGIN/..",
           "-Wl,-rpath," + os.path.join(cuda, "lib64")]
    subprocess.check_call(cmd)
    return out


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "checked":   # device-side asserts on computed indices: REB_LIB=<printed path> python -m pytest tests -m gpu
        print("built", build_variant("checked", ["REB_CHECKED"]))
        sys.exit(0)
    build(force=True, verbose=True)
    print("built", OUT)
    print("built", build_host(force=True))
