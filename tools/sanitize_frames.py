"""Target of compute-sanitizer (memcheck / racecheck / initcheck): small frames through every kernel family of libreb.so —
plain frame, camera inside the scene (Appendix B.3 clipping + binned tile path), guard-band shadow pass, 4x MSAA, translucent
list, PBR + textures + skybox, sort-first split (rank 1 of 2), and the allocation-free shadow entry point.

    compute-sanitizer --tool memcheck python tools/sanitize_frames.py tiny
"""
import sys
sys.path.insert(0, '.')
import numpy as np
from renderengine_b200 import capi, scene as sc

cfg = sys.argv[1] if len(sys.argv) > 1 else "tiny"


def frame(s, translucent=None, samples=1, **kw):
    ctx = capi.Context(s["width"], s["height"], **kw)
    ctx.upload_scene(s)
    if samples != 1:
        ctx.set_sample_count(samples)
    for _ in range(2):
        if s["shadow_size"]:
            ctx.render_shadow(s["draws"])
        ctx.render_forward(s["draws"], translucent)
        ctx.sync()
    out = ctx.readback(capi.REB_COLOR_SRGB8)
    c = ctx.counters()
    ctx.close()
    return int(out.astype(np.uint64).sum()), c


s = sc.make_scene(cfg)
print("plain", frame(s))
v = sc.view_from(s, (0.3, 1.2, 0.4), (6.0, 0.0, 5.0))
si = dict(s); si["view"] = v
print("inside (clip + binned)", frame(si))
sz = dict(s)
p = np.array(s["shadow_proj"], np.float32).copy()
p[[0, 4, 8, 12]] *= 4000.0; p[[1, 5, 9, 13]] *= 4000.0
sz["shadow_proj"] = p
print("guard-band shadow", frame(sz))
print("msaa4", frame(s, samples=4))
print("msaa4 inside", frame(si, samples=4))
st = sc.to_translucent(s)
print("translucent", frame(st, translucent=st["translucent_draws"]))
sp = sc.to_pbr(s, n_texture_sets=2, tex_size=32)
sp["skybox"] = sc.make_skybox(16)
print("pbr + textures + sky", frame(sp))
print("split rank 1/2", frame(s, rank=1, n_ranks=2, band_h=16))

# allocation-free entry point against the context path
import torch
model = np.eye(4, dtype=np.float32).T.reshape(16)
vd = torch.from_numpy(np.ascontiguousarray(s["verts"], np.float32)).cuda()
idx = torch.from_numpy(np.ascontiguousarray(s["indices"], np.uint32).view(np.int32)).cuda()
S = s["shadow_size"]
out = torch.empty((S, S), dtype=torch.float32, device="cuda")
for proj in (s["shadow_proj"], p):
    capi.shadow_depth_static(vd.data_ptr(), 32, vd.shape[0], idx.data_ptr(), idx.numel(), model, s["shadow_view"], proj, S, out.data_ptr())
    torch.cuda.synchronize()
    print("static shadow", float(out.sum()))
print("done", cfg)
