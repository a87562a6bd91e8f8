"""Known-answer scenes for the fixed-function contract — the Vulkan rules the reference's pipeline
state asks for (GraphicsPipelineCache.cpp:280-297,330-340: viewport, cull back / front CCW, LessOrEqual; glm_includes.h:18-23 kClip).
Every check takes a `render(scene) -> dict(depth, tri_id, color_f32, color_srgb8[, shadow])` callable, so the SAME answers are
demanded from the CPU oracle (tests/test_raster_kats.py, not gpu) and from the device through the C ABI (-m gpu).
Vertices are given directly in clip space (view = proj = identity, w = 1): pixel centre (i, j) is ndc ((i + .5 - W/2) / (W/2), ...)."""
import numpy as np

from renderengine_b200 import scene as sc

f32 = np.float32
NONE = 0xFFFFFFFF


def ndc(px, extent):
    """NDC coordinate of sample position `px` (in pixels, pixel centres at k + 0.5) — exact for power-of-two extents."""
    return (f32(px) - f32(extent) / f32(2)) / (f32(extent) / f32(2))


def clip_scene(W, H, verts_xyz, tris, draws=None, n_materials=2, proj=None, shadow=None):
    """A scene whose vertices are already clip-space positions: identity node transform and view, `proj` (default identity)."""
    v = np.zeros((len(verts_xyz), 8), f32)
    v[:, :3] = np.asarray(verts_xyz, f32)
    v[:, 3:6] = [0, 0, -1]
    idx = np.asarray(tris, np.uint32).reshape(-1)
    if draws is None:
        draws = [(0, idx.size, sc.material_handle(0), 0)]
    mats = np.stack([sc.pack_material(np.array([0.5, 0.5, 0.5, 1], f32), np.array([0.25, 0.25, 0.25, 1], f32), 16.0) for _ in range(n_materials)])
    light = sc.pack_light(sc.LIGHT_DIRECTIONAL, np.array([0, 0, -50], f32), np.array([0, 0, 1], f32), [0.1] * 4, [1.0] * 4, [1.0] * 4)
    P = np.eye(4, dtype=f32) if proj is None else np.asarray(proj, f32)
    s = dict(config="kat", verts=v, indices=idx, transforms=np.eye(4, dtype=f32).reshape(1, 16), draws=np.asarray(draws, np.uint32).reshape(-1, 4),
             materials=mats, lights=np.stack([light]), view=sc.pack_view(np.eye(4, dtype=f32), P, np.array([0, 0, -5], f32)), width=W, height=H,
             shadow_size=0, max_descriptors=8192, n_triangles=idx.size // 3, n_vertices=len(v), textures=[])
    if shadow:
        s["shadow_size"] = shadow
        s["shadow_view"] = sc.to_flat(np.eye(4, dtype=f32)); s["shadow_proj"] = sc.to_flat(np.eye(4, dtype=f32))
    return s


def coverage(render, scene):
    return render(scene)["tri_id"] != NONE


# ---- 1. facing: frontFace = CCW, cull back (GraphicsPipelineCache.cpp:296-297; Appendix B.6) -------------------------------------------
def kat_tutorial_triangle_facing(render):
    """The Vulkan-tutorial triangle (0, -.5), (.5, .5), (-.5, .5) is clockwise on a y-down screen: with frontFace = CCW it is a BACK
    face and culled (the reference only shows geometry because kClip flips y, glm_includes.h:18-23); reversed, it is drawn."""
    W = H = 64
    v = [(0.0, -0.5, 0.5), (0.5, 0.5, 0.5), (-0.5, 0.5, 0.5)]
    assert not coverage(render, clip_scene(W, H, v, [(0, 1, 2)])).any()
    front = coverage(render, clip_scene(W, H, v, [(0, 2, 1)]))
    # pixel-centre count of that triangle: rows j = 16..47 (y_f in (16, 48]), half-width (j + .5 - 16) / 2 around x_f = 32
    want = np.zeros((H, W), bool)
    for j in range(H):
        for i in range(W):
            x, y = i + 0.5, j + 0.5
            hw = (y - 16.0) / 2.0
            want[j, i] = (16.0 <= y < 48.0) and (32.0 - hw <= x < 32.0 + hw) if y > 16.0 else False
    assert front.sum() > 400
    # interior agrees everywhere; only pixels whose centre lies exactly ON an edge depend on the fill rule (checked by the next KATs)
    on_edge = np.zeros((H, W), bool)
    for j in range(H):
        for i in range(W):
            x, y = i + 0.5, j + 0.5
            hw = (y - 16.0) / 2.0
            on_edge[j, i] = y in (16.0, 48.0) or abs(abs(x - 32.0) - hw) < 1e-9
    assert np.array_equal(front[~on_edge], want[~on_edge])


# ---- 2. top-left rule: a shared edge is covered exactly once (Appendix B.7) ---------------------------------------------------------------
def kat_shared_edges_exactly_once(render):
    """A rectangle whose corners AND interior fan vertex lie exactly on pixel centres, cut into a fan: every triangle rendered alone,
    the masks summed — 1 everywhere inside (no hole, no double hit), and the union is the top-left-rule rectangle
    [x0, x1) x [y0, y1) in pixel indices."""
    W, H = 32, 16
    x0, x1, y0, y1 = 3, 27, 2, 13                     # corners at pixel centres (k + .5)
    cx, cy = 11, 7                                    # fan centre on a pixel centre
    ring = [(x0, y0), (11, y0), (19, y0), (x1, y0), (x1, 6), (x1, y1), (20, y1), (9, y1), (x0, y1), (x0, 8)]   # CCW on screen = front after the winding flip below
    P = [(ndc(cx + 0.5, W), ndc(cy + 0.5, H), 0.5)] + [(ndc(a + 0.5, W), ndc(b + 0.5, H), 0.5) for a, b in ring]
    total = np.zeros((H, W), np.int32)
    n_front = 0
    for k in range(len(ring)):
        a, b = 1 + k, 1 + (k + 1) % len(ring)
        m = coverage(render, clip_scene(W, H, P, [(0, b, a)]))
        if not m.any():
            m = coverage(render, clip_scene(W, H, P, [(0, a, b)]))
        n_front += 1 if m.any() else 0
        total += m
    want = np.zeros((H, W), np.int32)
    want[y0:y1, x0:x1] = 1
    assert n_front == len(ring)
    assert np.array_equal(total, want), f"double-covered: {(total > 1).sum()}, holes: {((total == 0) & (want == 1)).sum()}, spill: {((total > 0) & (want == 0)).sum()}"


def kat_shared_edges_subpixel_mesh(render):
    """The same exactly-once property for a mesh with arbitrary sub-pixel vertices (random interior points, Delaunay-free grid
    triangulation): masks of the single triangles sum to the mask of the whole mesh drawn at once, which has no hole inside."""
    W, H = 48, 40
    rng = np.random.default_rng(11)
    n = 6
    gx, gy = np.meshgrid(np.linspace(-0.8, 0.8, n), np.linspace(-0.8, 0.8, n), indexing="xy")
    jit = rng.uniform(-0.1, 0.1, (n, n, 2)); jit[[0, -1], :, :] = 0; jit[:, [0, -1], :] = 0
    pts = np.stack([gx + jit[..., 0], gy + jit[..., 1], np.full_like(gx, 0.5)], -1).reshape(-1, 3)
    tris = []
    for r in range(n - 1):
        for c in range(n - 1):
            a, b, d, e = r * n + c, r * n + c + 1, (r + 1) * n + c, (r + 1) * n + c + 1
            tris += [(a, d, b), (b, d, e)]            # front-facing on the y-down screen
    whole = render(clip_scene(W, H, pts, tris))["tri_id"]
    total = np.zeros((H, W), np.int32)
    for t in tris:
        total += coverage(render, clip_scene(W, H, pts, [t]))
    assert total.max() == 1 and np.array_equal(total == 1, whole != NONE)
    inside = np.zeros((H, W), bool)
    lo_x, hi_x = 0.1 * W, 0.9 * W; lo_y, hi_y = 0.1 * H, 0.9 * H
    for j in range(H):
        for i in range(W):
            inside[j, i] = lo_x < i + 0.5 < hi_x and lo_y < j + 0.5 < hi_y
    assert (whole != NONE)[inside].all()             # no hole strictly inside the outline


# ---- 3. viewport (GraphicsPipelineCache.cpp:280-285; Appendix B.4): the NDC square is exactly the target --------------------------------
def kat_fullscreen_quad_covers_every_pixel_once(render):
    W, H = 40, 24
    v = [(-1, -1, 0.25), (-1, 1, 0.25), (1, 1, 0.25), (1, -1, 0.25)]
    a = coverage(render, clip_scene(W, H, v, [(0, 1, 2)])); b = coverage(render, clip_scene(W, H, v, [(0, 2, 3)]))
    both = render(clip_scene(W, H, v, [(0, 1, 2), (0, 2, 3)]))
    assert (both["tri_id"] != NONE).sum() == W * H
    assert (a.astype(int) + b.astype(int) == 1).all()           # the diagonal is nobody's twice
    assert (both["depth"] == f32(0.25)).all()                   # constant-depth plane interpolates exactly


# ---- 4. depth test LessOrEqual (GraphicsPipelineCache.cpp:330-340; Appendix B.9): at equal depth the LATER primitive wins -----------------
def kat_equal_depth_later_primitive_wins(render):
    W, H = 32, 32
    v = [(-1, -1, 0.5), (-1, 1, 0.5), (1, 1, 0.5), (1, -1, 0.5)] * 2
    tris = [(0, 1, 2), (0, 2, 3), (4, 5, 6), (4, 6, 7)]
    draws = [(0, 6, sc.material_handle(0), 0), (6, 6, sc.material_handle(1), 0)]
    out = render(clip_scene(W, H, v, tris, draws=draws))
    assert set(np.unique(out["tri_id"])) == {2, 3}
    nearer = list(v); nearer[0:4] = [(x, y, 0.49) for x, y, _ in v[0:4]]     # a strictly nearer earlier primitive still wins
    out = render(clip_scene(W, H, nearer, tris, draws=draws))
    assert set(np.unique(out["tri_id"])) == {0, 1}
    farther = list(v); farther[4:8] = [(x, y, 0.51) for x, y, _ in v[4:8]]   # a strictly farther later primitive loses
    out = render(clip_scene(W, H, farther, tris, draws=draws))
    assert set(np.unique(out["tri_id"])) == {0, 1}


# ---- 5. PCF known answer (shadow.glsl:24-56 + phong.glsl:39-81): fully shadowed = ambient * diffuse = 0.05 ----------------------------------
def kat_pcf_fully_shadowed_is_ambient_times_diffuse(render):
    """Light and shadow transform = identity (the light looks along +z); the camera's proj flips z so it sees the receiver (z = .8) in
    front of the occluder (z = .3).  Receiver pixels whose 3x3 taps all land under the occluder: shadow = 1, colour =
    light.ambient * material.diffuse = 0.1 * 0.5 (SURVEY Appendix D known answer); far outside: shadow = 0, lit."""
    W = H = 64
    S = 64
    rec = [(-1, -1, 0.8), (-1, 1, 0.8), (1, 1, 0.8), (1, -1, 0.8)]
    occ = [(-0.5, -0.5, 0.3), (-0.5, 0.5, 0.3), (0.5, 0.5, 0.3), (0.5, -0.5, 0.3)]
    flipz = np.eye(4, dtype=f32); flipz[2, 2] = -1; flipz[2, 3] = 1                 # clip.z = 1 - z
    s = clip_scene(W, H, rec + occ, [(0, 1, 2), (0, 2, 3), (4, 5, 6), (4, 6, 7)], proj=flipz, shadow=S)
    out = render(s)
    assert (out["shadow"][20:44, 20:44] == f32(0.3)).all() and (out["shadow"][:8, :8] == f32(0.8)).all()
    assert (out["tri_id"] <= 1).all()                                               # the camera sees only the receiver
    c = out["color_f32"]
    inner = c[20:44, 20:44, :3]
    assert np.allclose(inner, 0.05, rtol=2e-6, atol=0), (inner.min(), inner.max())
    assert (c[:8, :8, :3] > 0.2).all()                                              # lit: ambient + diffuse
    assert (out["color_srgb8"][20:44, 20:44, :3] == 63).all()                       # sRGB OETF(0.05) * 255 = 63.0 -> 63
    edge = c[16, 20:44, 0]                                                          # the row whose taps straddle the occluder's edge
    assert ((edge > 0.05 + 1e-4) & (edge < c[0, 0, 0] - 1e-4)).all()               # partial PCF coverage: strictly between


# ---- 6. kClip depth range (glm_includes.h:18-23): the reference's camera puts every fragment in [0.5, 1] ------------------------
def kat_default_camera_depth_range(render):
    s = sc.make_scene("tiny")
    s["shadow_size"] = 0
    out = render(s)
    drawn = out["tri_id"] != NONE
    assert drawn.mean() > 0.05
    assert (out["depth"][drawn] >= 0.5).all() and (out["depth"][drawn] < 1.0).all() and (out["depth"][~drawn] == 1.0).all()


ALL = [kat_tutorial_triangle_facing, kat_shared_edges_exactly_once, kat_shared_edges_subpixel_mesh, kat_fullscreen_quad_covers_every_pixel_once,
       kat_equal_depth_later_primitive_wins, kat_pcf_fully_shadowed_is_ambient_times_diffuse, kat_default_camera_depth_range]
