"""Renders N frames of C2 with every second material translucent, optionally 4x multisampled; the target of ncu launch lists.
    python tools/prof_tlms.py 3 4"""
import sys
sys.path.insert(0, '.')
import numpy as np, torch
from renderengine_b200 import capi, scene as sc
n = int(sys.argv[1]) if len(sys.argv) > 1 else 3
samples = int(sys.argv[2]) if len(sys.argv) > 2 else 4
s = sc.to_translucent(sc.make_scene("C2"), every=2, alpha=0.45)
ctx = capi.Context(s["width"], s["height"], outputs=capi.REB_OUT_COLOR_SRGB8 | capi.REB_OUT_DEPTH_F32 | capi.REB_OUT_TRIID_U32)
ctx.upload_scene(s)
if samples != 1:
    ctx.set_sample_count(samples)
both = np.concatenate([s["draws"], s["translucent_draws"]])
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for _ in range(n):
    flush.add_(1); torch.cuda.synchronize()
    ctx.render_shadow(both); ctx.render_forward(s["draws"], s["translucent_draws"]); ctx.sync()
print("done", n, samples, "launches", ctx.launch_count())
