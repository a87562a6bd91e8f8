import sys
sys.path.insert(0, '.')
import numpy as np
from renderengine_b200 import capi, scene as sc
from oracle import orc
cfg = sys.argv[1] if len(sys.argv) > 1 else "tiny"
s = sc.make_scene(cfg)
o = orc.Oracle(s); rs = o.render_shadow(); ref = o.render_forward(rs)
ctx = capi.Context(s["width"], s["height"]); ctx.upload_scene(s)
ctx.render_shadow(s["draws"]); ctx.render_forward(s["draws"]); ctx.sync()
gs = ctx.readback(capi.REB_SHADOW_F32)
print("counters", ctx.counters(), "oracle", o.counters)
d = gs.view(np.uint32) != rs.view(np.uint32)
print("shadow diff px", d.sum(), "of", d.size)
if d.sum():
    ys, xs = np.nonzero(d)
    print("bbox", ys.min(), ys.max(), xs.min(), xs.max())
    for y, x in list(zip(ys, xs))[:8]:
        print(y, x, gs[y, x], rs[y, x])
    print("gpu<1:", (gs < 1).sum(), "ref<1:", (rs < 1).sum())
t = ctx.readback(capi.REB_TRIID_U32)
print("tri diff", (t != ref["tri_id"]).sum())
