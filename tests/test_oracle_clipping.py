"""Appendix B.3 clipping in the oracle, checked against an independent construction: a ground plane that crosses the eye
plane and the guard band as ONE quad (two huge triangles, clipped) must cover the same pixels, at the same depth up to
interpolation rounding, as the same plane tessellated into a fine grid (where almost no triangle needs clipping)."""
import numpy as np

from oracle import orc
from renderengine_b200 import scene as sc

f32 = np.float32


def scene_with_ground(n_cells):
    s = sc.make_scene("tiny", ground=False)
    ex, ez = s["extent"]
    hx, hz, gy = ex + 1.5, ez + 1.5, -1.05
    xs = np.linspace(-hx, hx, n_cells + 1, dtype=f32); zs = np.linspace(-hz, hz, n_cells + 1, dtype=f32)
    gx, gz = np.meshgrid(xs, zs, indexing="ij")
    gv = np.zeros(((n_cells + 1) ** 2, 8), f32)
    gv[:, 0] = gx.reshape(-1); gv[:, 1] = gy; gv[:, 2] = gz.reshape(-1); gv[:, 4] = 1.0
    gv[:, 6] = (gx.reshape(-1) + hx) / (2 * hx); gv[:, 7] = (gz.reshape(-1) + hz) / (2 * hz)
    base = len(s["verts"])
    idx = []
    for i in range(n_cells):
        for j in range(n_cells):
            a = base + i * (n_cells + 1) + j; b = a + 1; c = a + (n_cells + 1) + 1; d = a + (n_cells + 1)
            idx += [a, b, c, a, c, d]
    n_nodes = s["transforms"].shape[0]
    s["first_ground_tri"] = s["indices"].size // 3
    s["draws"] = np.concatenate([s["draws"], np.array([[s["indices"].size, len(idx), sc.material_handle(0), n_nodes]], np.uint32)])
    s["verts"] = np.concatenate([s["verts"], gv]); s["indices"] = np.concatenate([s["indices"], np.array(idx, np.uint32)])
    s["transforms"] = np.concatenate([s["transforms"], np.eye(4, dtype=f32).reshape(1, 16)])
    s["n_triangles"] = s["indices"].size // 3; s["n_vertices"] = len(s["verts"])
    return s


def test_clipped_ground_quad_matches_tessellated_ground():
    views = [((0.3, 0.2, 0.4), (3.0, -0.2, 2.0)), ((0.0, -0.5, 0.0), (1.0, -0.6, 0.2)), ((0.4, -0.9, -0.3), (-2.0, -1.0, 1.5))]
    one, many = scene_with_ground(1), scene_with_ground(96)
    for eye, look in views:
        out = []
        for s in (one, many):
            o = orc.Oracle(s)
            o.set_view(sc.view_from(s, eye, look))
            sm = o.render_shadow()
            r = o.render_forward(sm)
            assert o.counters[0] == 0   # nothing dropped
            out.append((r["tri_id"] >= s["first_ground_tri"]) & (r["tri_id"] != 0xFFFFFFFF))
            out.append(r["depth"])
        g1, d1, g2, d2 = out
        assert g1.mean() > 0.1, "the view must show a good part of the ground"
        assert (g1 != g2).mean() < 2e-4, f"coverage differs at {(g1 != g2).sum()} pixels"
        both = g1 & g2
        assert np.abs(d1[both] - d2[both]).max() < 2e-5
