"""How fast does the host issue frames?  Wall time per frame of the e2e call sequence (set_view, shadow transform, render_shadow,
render_forward, readback_async) WITHOUT waiting for the GPU, vs the same with the final sync.
    python tools/host_rate.py C2 200"""
import sys, time
sys.path.insert(0, '.')
import numpy as np, torch
from renderengine_b200 import capi, scene as sc
cfg = sys.argv[1] if len(sys.argv) > 1 else "C2"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 200
s = sc.make_scene(cfg)
st = torch.cuda.Stream()
ctx = capi.Context(s["width"], s["height"], stream=st.cuda_stream, outputs=capi.REB_OUT_COLOR_SRGB8 | capi.REB_OUT_DEPTH_F32 | capi.REB_OUT_TRIID_U32)
ctx.upload_scene(s)
pinned = [torch.empty((s["height"], s["width"], 4), dtype=torch.uint8).pin_memory() for _ in range(3)]
view = np.array(s["view"])
def frame(i):
    ctx.set_view(view)
    if s["shadow_size"]:
        ctx.shadow_set_transform(s["shadow_view"], s["shadow_proj"])
        ctx.render_shadow(s["draws"])
    ctx.render_forward(s["draws"])
    ctx.readback_async_into(capi.REB_COLOR_SRGB8, pinned[i % 3].data_ptr())
for i in range(5): frame(i)
ctx.sync(); torch.cuda.synchronize()
t0 = time.perf_counter()
for i in range(n): frame(i)
t_issue = time.perf_counter() - t0
ctx.sync(); torch.cuda.synchronize()
t_all = time.perf_counter() - t0
print(f"{cfg}: host issue {1e6 * t_issue / n:.1f} us / frame, with the final sync {1e6 * t_all / n:.1f} us / frame")
