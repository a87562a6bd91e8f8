"""CPU checks of the oracle's translucent pass (Appendix B.16; RenderScene.cpp:127-159,226-227,
GraphicsPipelineCache.cpp:305-315): ordered src-alpha blending with depth test + write, cull none."""
import numpy as np

from oracle import orc
from renderengine_b200 import scene as sc


def test_translucent_pixels_blend_over_the_opaque_frame():
    base = sc.make_scene("small")
    s = sc.to_translucent(base, every=2, alpha=0.5)
    assert len(s["translucent_draws"]) > 0 and len(s["draws"]) + len(s["translucent_draws"]) == len(base["draws"])
    o = orc.Oracle(s)
    sm = o.render_shadow()
    fr = o.render_forward(sm)
    # the opaque-only frame: same scene without the translucent list
    so = dict(s); so.pop("translucent_draws")
    op = orc.Oracle(so).render_forward(sm)
    n_opaque = int(sum(int(d[1]) // 3 for d in s["draws"]))
    t = fr["tri_id"]
    tl = (t != 0xFFFFFFFF) & (t >= n_opaque)
    assert tl.mean() > 0.03
    # where no translucent fragment passed, the frame is the opaque frame bit for bit
    assert np.array_equal(fr["color_f32"][~tl].view(np.uint32), op["color_f32"][~tl].view(np.uint32))
    assert np.array_equal(fr["depth"][~tl], op["depth"][~tl]) and np.array_equal(t[~tl], op["tri_id"][~tl])
    # where one did, depth moved nearer (LessOrEqual + depth write) and alpha is As * (1 - As) = 0.25
    assert (fr["depth"][tl] <= op["depth"][tl]).all()
    assert np.allclose(fr["color_f32"][tl][:, 3], 0.25)
    # a translucent sphere shows its back AND front faces: somewhere two fragments blended, i.e. the colour is not a single
    # 50/50 mix of one shaded colour with the opaque frame -> just check the blend darkened / changed the pixel
    assert (np.abs(fr["color_f32"][tl][:, :3] - op["color_f32"][tl][:, :3]).max(axis=1) > 1e-4).mean() > 0.9


def test_back_faces_are_rasterised_for_translucent_draws_only():
    """Cull none: put the camera INSIDE a translucent sphere's bounding region is not needed — a translucent sphere in front
    of the ground shows its far (back-facing) side first; with alpha = 1 the result must equal drawing it opaque (the nearest
    fragment wins either way), which pins the two-sided raster + ordering without any blending arithmetic."""
    base = sc.make_scene("small")
    s = sc.to_translucent(base, every=2, alpha=1.0)
    o = orc.Oracle(s)
    sm = o.render_shadow()
    fr = o.render_forward(sm)
    ref = orc.Oracle(base).render_forward(sm)          # everything opaque
    assert np.array_equal(fr["depth"].view(np.uint32), ref["depth"].view(np.uint32))
    same = np.abs(fr["color_f32"][..., :3] - ref["color_f32"][..., :3]).max(axis=2) < 1e-6
    assert same.mean() > 0.999


def test_translucent_draw_order_is_back_to_front():
    s = sc.to_translucent(sc.make_scene("small"), every=2)
    vw = np.frombuffer(np.asarray(s["view"], np.uint8).tobytes(), np.float32)
    Vi = np.linalg.inv(vw[:16].reshape(4, 4).T.astype(np.float64))
    T = np.asarray(s["transforms"], np.float32).reshape(-1, 16)
    d = [float(np.dot(Vi[:3, 2], Vi[:3, 3] - T[int(n), 12:15])) for n in s["translucent_draws"][:, 3]]
    assert all(a >= b - 1e-9 for a, b in zip(d, d[1:]))


def test_translucent_list_inside_the_multisampled_pass():
    """Appendix B.15 + B.16 together (the reference draws opaque + translucent + skybox in ONE multisampled render pass,
    RenderScene.cpp:217-228, Swapchain.cpp:110-123): every sample is its own blend target.  Properties: (1) with alpha = 1 the
    resolved frame equals the all-opaque multisampled frame (nearest fragment wins per sample either way); (2) with alpha = 0.5
    a pixel none of whose samples saw a translucent fragment equals the opaque-only multisampled frame bit for bit; (3) a pixel
    whose four samples agree (same translucent history) resolves to exactly the 1-sample translucent colour."""
    base = sc.make_scene("tiny")
    base["skybox"] = sc.make_skybox(8)
    sm = orc.Oracle(base).render_shadow()
    s1 = sc.to_translucent(base, every=2, alpha=1.0)
    fr = orc.Oracle(s1).render_forward_msaa4(sm)
    ref = orc.Oracle(base).render_forward_msaa4(sm)
    assert np.array_equal(fr["depth"].view(np.uint32), ref["depth"].view(np.uint32))
    assert (np.abs(fr["color_f32"][..., :3] - ref["color_f32"][..., :3]).max(axis=2) < 1e-6).mean() > 0.999

    s = sc.to_translucent(base, every=2, alpha=0.5)
    n_opaque = int(sum(int(d[1]) // 3 for d in s["draws"]))
    fr = orc.Oracle(s).render_forward_msaa4(sm)
    so = dict(s); so.pop("translucent_draws")
    op = orc.Oracle(so).render_forward_msaa4(sm)
    t = fr["tri_id"]                                             # (4, H, W)
    tl = ((t != 0xFFFFFFFF) & (t >= n_opaque)).any(axis=0)
    assert 0.03 < tl.mean() < 0.97
    assert np.array_equal(fr["color_f32"][~tl].view(np.uint32), op["color_f32"][~tl].view(np.uint32))
    assert np.array_equal(fr["depth"][:, ~tl], op["depth"][:, ~tl])
    assert (fr["depth"] <= op["depth"]).all()
    one = orc.Oracle(s).render_forward(sm)
    # interior pixels: all four samples carry the same last primitive as the 1-sample frame and saw the same fragments
    same = (t == one["tri_id"][None]).all(axis=0) & (op["tri_id"] == orc.Oracle(so).render_forward(sm)["tri_id"][None]).all(axis=0)
    eq = (fr["color_f32"].view(np.uint32) == one["color_f32"].view(np.uint32)).all(axis=2)
    assert same.mean() > 0.3 and eq[same].mean() > 0.97          # (a hidden intermediate fragment may differ per sample on a few pixels)
