"""Renders N 4x multisampled frames (shadow + forward) of a config through the C ABI; the target of ncu launch lists.
    python tools/prof_msaa.py C2 3"""
import sys
sys.path.insert(0, '.')
import torch
from renderengine_b200 import capi, scene as sc
cfg = sys.argv[1] if len(sys.argv) > 1 else "C2"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 3
s = sc.make_scene(cfg)
ctx = capi.Context(s["width"], s["height"], outputs=capi.REB_OUT_COLOR_SRGB8 | capi.REB_OUT_DEPTH_F32 | capi.REB_OUT_TRIID_U32)
ctx.upload_scene(s)
ctx.set_sample_count(4)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for _ in range(n):
    flush.add_(1); torch.cuda.synchronize()
    ctx.render_shadow(s["draws"]); ctx.render_forward(s["draws"]); ctx.sync()
print("done", cfg, n, "launches", ctx.launch_count())
