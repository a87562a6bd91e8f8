"""Per-kernel and whole-frame device times of a config (CUDA events, L2 flushed between frames).
    [REB_LIB=renderengine_b200/variants/libreb_x.so] python tools/kt.py C2 [frames]"""
import os, sys
sys.path.insert(0, '.')
import numpy as np, torch
from renderengine_b200 import capi, scene as sc
cfg = sys.argv[1] if len(sys.argv) > 1 else "C2"
N = int(sys.argv[2]) if len(sys.argv) > 2 else 30
s = sc.make_scene(cfg)
st = torch.cuda.Stream()
ctx = capi.Context(s["width"], s["height"], stream=st.cuda_stream, outputs=capi.REB_OUT_COLOR_SRGB8 | capi.REB_OUT_DEPTH_F32 | capi.REB_OUT_TRIID_U32)
ctx.upload_scene(s)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
S = s["shadow_size"]
with torch.cuda.stream(st):
    def frame():
        if S: ctx.render_shadow(s["draws"])
        ctx.render_forward(s["draws"])
    for _ in range(5): flush.add_(1); frame()
    ctx.sync()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(N)]
    for a, b in ev:
        flush.add_(1); a.record(st); frame(); b.record(st)
    ctx.sync()
    ts = sorted(a.elapsed_time(b) for a, b in ev)
    ctx.set_profiling(True)
    acc = {}
    for _ in range(10):
        flush.add_(1); frame(); ctx.sync()
        for k in ((0, 1) if S else (1,)):
            for n, v in ctx.kernel_times_ms(k).items():
                acc[("sh." if k == 0 else "fw.") + n] = acc.get(("sh." if k == 0 else "fw.") + n, 0) + v / 10
slow = [round(t, 2) for t in ts if t > 5 * ts[0]]
if slow:
    print("   SLOW FRAMES (ms):", len(slow), slow[:8])
print(os.environ.get("REB_LIB", "default"), cfg, f"frame median {1e3*ts[len(ts)//2]:.1f} us  min {1e3*ts[0]:.1f}  Mtri/s {s['n_triangles']/ts[len(ts)//2]/1e3:.0f}")
print("   " + "  ".join(f"{k}={1e3*v:.1f}" for k, v in acc.items()), " counters", ctx.counters())
