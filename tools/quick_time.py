import sys, time
sys.path.insert(0, '.')
import numpy as np, torch
from renderengine_b200 import capi, scene as sc
cfg = sys.argv[1] if len(sys.argv) > 1 else "C2"
s = sc.make_scene(cfg)
print(cfg, s["n_triangles"], s["n_vertices"], s["width"], s["height"])
st = torch.cuda.Stream()
ctx = capi.Context(s["width"], s["height"], stream=st.cuda_stream, outputs=capi.REB_OUT_COLOR_SRGB8|capi.REB_OUT_DEPTH_F32|capi.REB_OUT_TRIID_U32)
ctx.upload_scene(s)
with torch.cuda.stream(st):
    for mode in (("full", "forward") if s["shadow_size"] else ("forward",)):
        for _ in range(3):
            if mode == "full": ctx.render_shadow(s["draws"])
            ctx.render_forward(s["draws"])
        ctx.sync()
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        N = int(sys.argv[2]) if len(sys.argv) > 2 else 50
        e0.record(st)
        for _ in range(N):
            if mode == "full": ctx.render_shadow(s["draws"])
            ctx.render_forward(s["draws"])
        e1.record(st)
        ctx.sync()
        ms = e0.elapsed_time(e1) / N
        bs, bf = sc.algorithmic_bytes(s)
        B = bs + bf if mode == "full" else bf
        print(f"{mode}: {ms*1000:.1f} us/frame  {s['n_triangles']/ms/1e3:.1f} Mtri/s  {B/ms/1e6:.1f} GB/s  counters {ctx.counters()}")
