"""Renders frames of ONE rank of a sort-first split (row bands of the forward pass + row slab of the shadow pass) on one GPU:
the target of ncu launch lists for the split path.
    python tools/split_frame.py C4 8 6 [bands_per_rank] [frames]"""
import sys
sys.path.insert(0, '.')
import torch
from renderengine_b200 import capi, parallel, scene as sc
cfg, N, r = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
bpr = int(sys.argv[4]) if len(sys.argv) > 4 else 2
frames = int(sys.argv[5]) if len(sys.argv) > 5 else 3
s = sc.make_scene(cfg)
band = parallel.contiguous_band(s["height"], N * bpr)
ctx = capi.Context(s["width"], s["height"], outputs=capi.REB_OUT_COLOR_SRGB8 | capi.REB_OUT_DEPTH_F32 | capi.REB_OUT_TRIID_U32, rank=r, n_ranks=N, band_h=band)
ctx.upload_scene(s)
rows = parallel.slab_rows(s["shadow_size"], N)
if rows:
    ctx.set_shadow_partition(r, N, rows)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for _ in range(frames):
    flush.add_(1); torch.cuda.synchronize()
    ctx.render_shadow(s["draws"]); ctx.render_forward(s["draws"]); ctx.sync()
print("done", cfg, "rank", r, "of", N, "band_h", band, "launches", ctx.launch_count())
