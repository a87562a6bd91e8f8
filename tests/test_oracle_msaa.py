"""CPU checks of the oracle's 4x multisampled base pass (Appendix B.15; Swapchain.cpp:110-123,212): sample placement by
exact vertex shifts, one shading per pixel and primitive, average resolve."""
import numpy as np

from oracle import orc
from renderengine_b200 import scene as sc


def _scene():
    s = sc.make_scene("small", width=200, height=120)
    return s


def test_interior_pixels_equal_the_one_sample_frame_and_edges_blend():
    s = _scene()
    o = orc.Oracle(s)
    sm = o.render_shadow()
    one = o.render_forward(sm)
    ms = o.render_forward_msaa4(sm)
    t = ms["tri_id"]
    same = (t[0] == t[1]) & (t[1] == t[2]) & (t[2] == t[3])
    interior = same & (t[0] == one["tri_id"])
    assert interior.mean() > 0.6 and (~same).mean() > 0.01
    # four agreeing samples resolve to exactly the colour of the single shader invocation at the pixel centre
    assert np.array_equal(ms["color_f32"][interior].view(np.uint32), one["color_f32"][interior].view(np.uint32))
    # silhouette pixels whose samples see only the centre's primitive or the background: the resolve is the pairwise
    # average of that primitive's colour and the clear colour (0, 0, 0, 1), reproduced here in float32
    T = one["tri_id"]
    only = (~same) & (T != 0xFFFFFFFF) & np.all((t == T) | (t == 0xFFFFFFFF), axis=0)
    assert only.sum() > 20
    c = one["color_f32"][only]
    clear = np.array([0, 0, 0, 1], np.float32)
    col = [np.where((t[k][only] == T[only])[:, None], c, clear[None, :]).astype(np.float32) for k in range(4)]
    want = ((col[0] + col[1]) + (col[2] + col[3])) * np.float32(0.25)
    assert np.array_equal(ms["color_f32"][only].view(np.uint32), want.view(np.uint32))
    assert np.all(ms["color_f32"][..., 3] == 1.0)


def test_sample_planes_are_shifted_renders():
    """Sample k of pixel (x, y) sits at (x + sx, y + sy): coverage of a plane differs from the centre-sampled frame only
    near primitive edges, and each plane's depth at agreeing interior pixels differs from the centre depth by the plane
    equation's slope times the sample offset (sign check on the ground plane, whose depth grows towards the horizon)."""
    s = _scene()
    o = orc.Oracle(s)
    one = o.render_forward(None)
    ms = o.render_forward_msaa4(None)
    ground = s["n_triangles"] - 2
    for k in range(4):
        assert (ms["tri_id"][k] != one["tri_id"]).mean() < 0.2       # small triangles at 200 x 120: only edge pixels change owner
    g = (one["tri_id"] >= ground) & np.all(ms["tri_id"] >= ground, axis=0) & np.all(ms["tri_id"] != 0xFFFFFFFF, axis=0)
    assert g.mean() > 0.05
    # sample 0 is at y + 0.125, sample 3 at y + 0.875: on the ground (camera above, looking down) depth decreases with y
    d0, d3 = ms["depth"][0][g], ms["depth"][3][g]
    assert (d0 > d3).mean() > 0.99


def test_skybox_resolves_with_geometry():
    s = _scene(); s["skybox"] = sc.make_skybox(16)
    o = orc.Oracle(s)
    ms = o.render_forward_msaa4(None)
    one = o.render_forward(None)
    t = ms["tri_id"]
    allsky = np.all(t == 0xFFFFFFFF, axis=0)
    assert np.array_equal(ms["color_f32"][allsky].view(np.uint32), one["color_f32"][allsky].view(np.uint32))
