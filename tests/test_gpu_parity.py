"""Parity of the CUDA path (through the C ABI, libreb.so) against the CPU oracle on identical inputs.
Bit-exact: shadow map, depth, triangle ID.  Colour: <= 1 LSB per 8-bit channel, PSNR >= 50 dB on the
f32 HDR buffer (BASELINE.json north_star tolerances)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _psnr(a, b):
    a = a.astype(np.float64); b = b.astype(np.float64)
    mse = np.mean((a - b) ** 2)
    peak = max(1.0, float(np.max(np.abs(b))))
    return 1e9 if mse == 0 else 10 * np.log10(peak * peak / mse)


def render_gpu(scene, shadow=True, **kw):
    from renderengine_b200 import capi
    ctx = capi.Context(scene["width"], scene["height"], **kw)
    ctx.upload_scene(scene)
    out = {}
    if shadow and scene["shadow_size"]:
        ctx.render_shadow(scene["draws"])
    ctx.render_forward(scene["draws"])
    ctx.sync()
    if shadow and scene["shadow_size"]:
        out["shadow"] = ctx.readback(capi.REB_SHADOW_F32)
    out["depth"] = ctx.readback(capi.REB_DEPTH_F32)
    out["tri_id"] = ctx.readback(capi.REB_TRIID_U32)
    out["color_f32"] = ctx.readback(capi.REB_COLOR_F32)
    out["color_srgb8"] = ctx.readback(capi.REB_COLOR_SRGB8)
    out["counters"] = ctx.counters()
    out["ctx"] = ctx
    return out


def check_parity(scene, gpu, ref_shadow, ref):
    if ref_shadow is not None:
        assert np.array_equal(gpu["shadow"].view(np.uint32), ref_shadow.view(np.uint32)), "shadow map not bit-exact"
    assert np.array_equal(gpu["tri_id"], ref["tri_id"]), f"triangle IDs differ at {np.count_nonzero(gpu['tri_id'] != ref['tri_id'])} pixels"
    assert np.array_equal(gpu["depth"].view(np.uint32), ref["depth"].view(np.uint32)), "depth not bit-exact"
    d8 = np.abs(gpu["color_srgb8"].astype(np.int32) - ref["color_srgb8"].astype(np.int32))
    assert d8.max() <= 1, f"sRGB8 colour differs by {d8.max()} LSB at {np.count_nonzero(d8 > 1)} channel samples"
    assert _psnr(gpu["color_f32"], ref["color_f32"]) >= 50.0


@pytest.mark.parametrize("config", ["tiny", "small", "C1"])
def test_parity_configs(config):
    from oracle import orc
    from renderengine_b200 import scene as sc
    s = sc.make_scene(config)
    o = orc.Oracle(s)
    ref_shadow = o.render_shadow() if s["shadow_size"] else None
    ref = o.render_forward(ref_shadow)
    gpu = render_gpu(s)
    assert gpu["counters"][0] == 0 and o.counters[0] == 0   # nothing needed true clipping
    check_parity(s, gpu, ref_shadow, ref)
    assert (ref["tri_id"] != 0xFFFFFFFF).mean() > 0.05      # the frame is not empty


def test_parity_rotated_views_and_second_frame():
    """Several frames on one context (the visibility buffer is consumed and re-armed by the tile stage)."""
    from oracle import orc
    from renderengine_b200 import capi, scene as sc
    s = sc.make_scene("small")
    o = orc.Oracle(s)
    ref_shadow = o.render_shadow()
    ctx = capi.Context(s["width"], s["height"])
    ctx.upload_scene(s)
    ctx.render_shadow(s["draws"])
    for az in (0.0, 0.7, 2.1, 0.0):
        v = sc.view_for_azimuth(s, az)
        o.set_view(v); ctx.set_view(v)
        ref = o.render_forward(ref_shadow)
        ctx.render_forward(s["draws"])
        gpu = dict(tri_id=ctx.readback(capi.REB_TRIID_U32), depth=ctx.readback(capi.REB_DEPTH_F32),
                   color_f32=ctx.readback(capi.REB_COLOR_F32), color_srgb8=ctx.readback(capi.REB_COLOR_SRGB8))
        check_parity(s, gpu, None, ref)


def test_sort_first_partition_matches_oracle_bands():
    from oracle import orc
    from renderengine_b200 import scene as sc
    s = sc.make_scene("small")
    full = orc.Oracle(s)
    ref_shadow = full.render_shadow()
    whole = full.render_forward(ref_shadow)
    merged = np.full_like(whole["tri_id"], 0xFFFFFFFF)
    for rank in range(3):
        o = orc.Oracle(s, rank=rank, n_ranks=3, band_h=16)
        ref = o.render_forward(ref_shadow)
        gpu = render_gpu(s, rank=rank, n_ranks=3, band_h=16)
        check_parity(s, gpu, ref_shadow, ref)
        own = ((np.arange(s["height"]) // 16) % 3 == rank)
        merged[own] = gpu["tri_id"][own]
    assert np.array_equal(merged, whole["tri_id"])   # the bands tile the frame exactly


def _quads_scene(n_quads, width=320, height=200, seed=7):
    """n_quads large, overlapping, randomly tilted quads: every triangle takes the large / huge path."""
    from renderengine_b200 import scene as sc
    rng = np.random.default_rng(seed)
    s = sc.make_scene("tiny", width=width, height=height, shadow_size=256)
    verts, idx, draws, tr = [], [], [], []
    for q in range(n_quads):
        hx, hz = rng.uniform(1.0, 5.0, 2)
        y = rng.uniform(-1.0, 1.5)
        tilt = rng.uniform(-0.3, 0.3, 4)
        base = 4 * q
        quad = np.array([[-hx, y + tilt[0], -hz, 0, 1, 0, 0, 0], [-hx, y + tilt[1], hz, 0, 1, 0, 0, 1],
                         [hx, y + tilt[2], hz, 0, 1, 0, 1, 1], [hx, y + tilt[3], -hz, 0, 1, 0, 1, 0]], np.float32)
        verts.append(quad)
        idx.append(np.array([0, 1, 2, 0, 2, 3], np.uint32) + base)
        t = np.eye(4, dtype=np.float32); t[0, 3], t[2, 3] = rng.uniform(-2, 2, 2)
        tr.append(sc.to_flat(t))
        draws.append((6 * q, 6, sc.material_handle(q % 3), q))
    s["verts"] = np.concatenate(verts); s["indices"] = np.concatenate(idx)
    s["transforms"] = np.stack(tr); s["draws"] = np.array(sorted(draws, key=lambda d: (d[2], d[3])), np.uint32)
    s["n_triangles"] = 2 * n_quads; s["n_vertices"] = 4 * n_quads
    return s


@pytest.mark.parametrize("n_quads", [3, 40, 150])
def test_parity_large_and_huge_triangles(n_quads):
    """3 quads: broadcast ('huge') list only; 40: 80 huge triangles overflow the 64-entry list into the binned path;
    150: the same at scale, plus equal-depth ties between coplanar pieces."""
    from oracle import orc
    s = _quads_scene(n_quads)
    o = orc.Oracle(s)
    ref_shadow = o.render_shadow()
    ref = o.render_forward(ref_shadow)
    gpu = render_gpu(s)
    check_parity(s, gpu, ref_shadow, ref)
    if n_quads >= 40:
        assert gpu["counters"][1] > 0   # the binned path really ran


def test_parity_lowpoly_binned_path():
    """Level-1 icospheres at 1280x720: triangles of ~60 px go through the binned tile lists (K1 rasterises up to 30-px boxes itself)."""
    from oracle import orc
    from renderengine_b200 import scene as sc
    s = sc.make_scene("small", level=1, width=1280, height=720)
    o = orc.Oracle(s)
    ref_shadow = o.render_shadow()
    ref = o.render_forward(ref_shadow)
    gpu = render_gpu(s)
    check_parity(s, gpu, ref_shadow, ref)
    assert gpu["counters"][1] > 100


@pytest.mark.parametrize("level,width,height,shadow", [(2, 640, 360, 512), (2, 1280, 720, 1024), (1, 640, 360, 512), (3, 1920, 1080, 2048),
                                                        (1, 960, 540, 768), (1, 1280, 720, 1024), (0, 640, 360, 512)])
def test_parity_medium_triangles_half_rows(level, width, height, shadow):
    """Triangles of 10 - 120 px: K1's own raster takes bboxes up to 30 x 30 pixel centres as rows of one or two half rows
    (SMALL_LIMIT, reb_device.cuh), appended <= 15 entries per lane and round; larger ones stay binned.  Both sides of both limits."""
    from oracle import orc
    from renderengine_b200 import scene as sc
    s = sc.make_scene("small", level=level, width=width, height=height, shadow_size=shadow)
    o = orc.Oracle(s)
    ref_shadow = o.render_shadow()
    ref = o.render_forward(ref_shadow)
    gpu = render_gpu(s)
    check_parity(s, gpu, ref_shadow, ref)
    assert (ref["tri_id"] != 0xFFFFFFFF).mean() > 0.05


def test_parity_deindexed_cache(monkeypatch):
    """Forces the fallback layout of the post-transform vertex cache (one slot per triangle corner)."""
    from oracle import orc
    from renderengine_b200 import scene as sc
    monkeypatch.setenv("REB_FORCE_DEINDEXED", "1")
    s = sc.make_scene("small")
    o = orc.Oracle(s)
    ref_shadow = o.render_shadow()
    ref = o.render_forward(ref_shadow)
    gpu = render_gpu(s)
    check_parity(s, gpu, ref_shadow, ref)


def test_parity_instanced_draws_share_vertices():
    """Several scene nodes draw the SAME vertex/index range (the MeshAllocator API allows it even though the
    reference's loader never produces it): every draw gets its own slice of the post-transform cache."""
    from oracle import orc
    from renderengine_b200 import scene as sc
    s = sc.make_scene("small", ground=False)
    sv, st = sc.icosphere(3)
    n_s = len(s["draws"])
    s["verts"] = sv.copy(); s["indices"] = st.reshape(-1).copy()
    s["draws"] = np.array(sorted([(0, st.size, sc.material_handle(k % 5), k) for k in range(n_s)], key=lambda d: (d[2], d[3])), np.uint32)
    s["n_triangles"] = n_s * len(st); s["n_vertices"] = len(sv)
    o = orc.Oracle(s)
    ref_shadow = o.render_shadow()
    ref = o.render_forward(ref_shadow)
    gpu = render_gpu(s)
    check_parity(s, gpu, ref_shadow, ref)
    assert (ref["tri_id"] != 0xFFFFFFFF).mean() > 0.05


def test_parity_textured_materials():
    """Bindless diffuse (sRGB) + specular (linear) textures: RGBA16 UNORM, device-generated mip chain, isotropic trilinear,
    repeat addressing, implicit derivatives from the neighbouring pixel centres of the same primitive."""
    from oracle import orc
    from renderengine_b200 import scene as sc
    s = sc.make_scene("small_tex")
    o = orc.Oracle(s)
    ref_shadow = o.render_shadow()
    ref = o.render_forward(ref_shadow)
    gpu = render_gpu(s)
    check_parity(s, gpu, ref_shadow, ref)
    plain = orc.Oracle(sc.make_scene("small")).render_forward(ref_shadow)
    assert (ref["color_srgb8"] != plain["color_srgb8"]).any(axis=2).mean() > 0.2   # the textures really modulate the frame


def test_parity_textured_minification():
    """A larger texture on distant geometry exercises the upper mip levels and the trilinear blend."""
    from oracle import orc
    from renderengine_b200 import scene as sc
    s = sc.make_scene("small_tex", tex_size=256, width=320, height=180)
    o = orc.Oracle(s)
    ref_shadow = o.render_shadow()
    ref = o.render_forward(ref_shadow)
    gpu = render_gpu(s)
    check_parity(s, gpu, ref_shadow, ref)


def test_register_texture_rejects_bad_arguments():
    from renderengine_b200 import capi
    ctx = capi.Context(64, 64, max_descriptors=2)
    t = np.zeros((4, 4, 4), np.uint16)
    assert ctx.register_texture(t) == 0 and ctx.register_texture(t, srgb=False) == 1
    with pytest.raises(capi.RebError):
        ctx.register_texture(t)          # descriptor table full (max_descriptors = 2)


@pytest.mark.parametrize("config", ["C2", "C3"])
def test_parity_full_size_configs(config):
    """BASELINE.json's own sizes against the oracle: C2 (1M triangles, 1080p, 2048^2 PCF shadow) and C3 (10M triangles,
    1024 materials / 4096 textures, 4K).  The oracle renders these in seconds on the host cores."""
    from oracle import orc
    from renderengine_b200 import scene as sc
    s = sc.make_scene(config)
    o = orc.Oracle(s)
    ref_shadow = o.render_shadow()
    ref = o.render_forward(ref_shadow)
    gpu = render_gpu(s)
    assert gpu["counters"][0] == 0 and o.counters[0] == 0
    check_parity(s, gpu, ref_shadow, ref)
    # size-independent properties: every drawn pixel names a front-facing primitive of the submitted range and has depth < 1
    drawn = gpu["tri_id"] != 0xFFFFFFFF
    assert gpu["tri_id"][drawn].max() < s["n_triangles"] and (gpu["depth"][drawn] < 1.0).all() and (gpu["depth"][~drawn] == 1.0).all()


def test_fused_and_unfused_vertex_stage_agree(monkeypatch):
    """reb_render_shadow is recorded and submitted with the forward pass (fused two-pass vertex kernel); forcing the
    separate kernels, or flushing the recording early, must give the same frame."""
    from renderengine_b200 import capi, scene as sc
    s = sc.make_scene("small")
    a = render_gpu(s)
    monkeypatch.setenv("REB_NO_FUSED_VERTEX", "1")
    b = render_gpu(s)
    monkeypatch.delenv("REB_NO_FUSED_VERTEX")
    ctx = capi.Context(s["width"], s["height"])
    ctx.upload_scene(s)
    ctx.render_shadow(s["draws"]); ctx.flush()
    shadow_early = ctx.readback(capi.REB_SHADOW_F32)
    ctx.render_forward(s["draws"]); ctx.sync()
    for k in ("shadow", "depth", "tri_id", "color_srgb8", "color_f32"):
        assert np.array_equal(a[k], b[k]), k
    assert np.array_equal(shadow_early, a["shadow"]) and np.array_equal(ctx.readback(capi.REB_TRIID_U32), a["tri_id"])
    assert np.array_equal(ctx.readback(capi.REB_COLOR_SRGB8), a["color_srgb8"])


def test_async_readback_swapchain_matches_blocking_readback():
    """reb_readback_async: frame i's copy (copy stream, swapchain image i % 3) overlaps the frames after it; every host image must
    equal the blocking readback of the same view."""
    import torch
    from renderengine_b200 import capi, scene as sc
    s = sc.make_scene("small")
    ctx = capi.Context(s["width"], s["height"])
    ctx.upload_scene(s)
    views = [sc.view_for_azimuth(s, az) for az in (0.0, 0.5, 1.0, 1.5, 2.0)]
    want = []
    for v in views:
        ctx.set_view(v); ctx.render_shadow(s["draws"]); ctx.render_forward(s["draws"])
        want.append(ctx.readback(capi.REB_COLOR_SRGB8))
    host = [torch.empty((s["height"], s["width"], 4), dtype=torch.uint8).pin_memory() for _ in views]
    for v, h in zip(views, host):
        ctx.set_view(v); ctx.render_shadow(s["draws"]); ctx.render_forward(s["draws"])
        ctx.readback_async_into(capi.REB_COLOR_SRGB8, h.data_ptr())
    ctx.sync()
    for w, h in zip(want, host):
        assert np.array_equal(w.reshape(h.shape), h.numpy())
    assert not np.array_equal(want[0], want[1])


# ---- the reference's live PBR fragment path (surface_pbr.frag -> pbr.glsl BRDF_Lighting), SURVEY.md 8f row 1 -------------
@pytest.mark.parametrize("config", ["small", "C1"])
def test_parity_pbr_untextured(config):
    """80-B MaterialProperties / Light records, directional (shadowed) + point + spot lights, GGX / Smith / Schlick."""
    from oracle import orc
    from renderengine_b200 import scene as sc
    base = sc.make_scene(config) if config != "C1" else sc.make_scene(config, shadow_size=1024)
    s = sc.to_pbr(base)
    o = orc.Oracle(s)
    ref_shadow = o.render_shadow()
    ref = o.render_forward(ref_shadow)
    gpu = render_gpu(s)
    check_parity(s, gpu, ref_shadow, ref)
    assert ref["color_f32"][..., :3].max() > 1.5   # there are specular highlights above the sRGB range (HDR buffer exercised)


def test_parity_pbr_textured_normal_mapped():
    """Base-colour (sRGB), emissive, ORM and tangent-space normal maps through the bindless table; the perturbed normal
    feeds the PCF bias, so a tap can flip where `depth - bias` ties with the map to the last ulp: such pixels are bounded."""
    from oracle import orc
    from renderengine_b200 import scene as sc
    s = sc.to_pbr(sc.make_scene("small"), n_texture_sets=3)
    o = orc.Oracle(s)
    ref_shadow = o.render_shadow()
    ref = o.render_forward(ref_shadow)
    gpu = render_gpu(s)
    assert np.array_equal(gpu["shadow"].view(np.uint32), ref_shadow.view(np.uint32))
    assert np.array_equal(gpu["tri_id"], ref["tri_id"]) and np.array_equal(gpu["depth"].view(np.uint32), ref["depth"].view(np.uint32))
    d8 = np.abs(gpu["color_srgb8"].astype(np.int32) - ref["color_srgb8"].astype(np.int32)).max(axis=2)
    assert (d8 > 1).mean() <= 1e-4, f"{np.count_nonzero(d8 > 1)} pixels beyond 1 LSB (max {d8.max()})"
    assert _psnr(gpu["color_f32"], ref["color_f32"]) >= 50.0


@pytest.mark.parametrize("dbg", [(1, 0), (3, 0), (5, 0), (7, 0), (0, 1), (0, 2), (0, 3), (0, 4), (0, 5)])
def test_parity_pbr_debug_views(dbg):
    """view.debugInput / view.debugEquation (pbr.glsl:409-439; CameraViewSystem::SetViewDebug)."""
    from oracle import orc
    from renderengine_b200 import scene as sc
    s = sc.to_pbr(sc.make_scene("tiny"), debug_input=dbg[0], debug_equation=dbg[1])
    o = orc.Oracle(s)
    ref_shadow = o.render_shadow()
    ref = o.render_forward(ref_shadow)
    gpu = render_gpu(s)
    check_parity(s, gpu, ref_shadow, ref)


def test_pbr_rejects_phong_records():
    from renderengine_b200 import capi, scene as sc
    s = sc.make_scene("tiny")
    ctx = capi.Context(s["width"], s["height"])
    ctx.upload_scene(s)
    ctx.set_shading_model(1)   # 48-B Phong materials are too short for the PBR path
    with pytest.raises(RuntimeError):
        ctx.render_forward(s["draws"])


# ---- Appendix B.3: triangles crossing the eye plane / guard band are clipped (K1 rare path + homogeneous varyings) ------------
INSIDE_VIEWS = [((0.3, 0.2, 0.4), (3.0, -0.2, 2.0)), ((0.0, -0.5, 0.0), (1.0, -0.6, 0.2)), ((1.2, 0.1, 1.3), (-2.0, 0.0, -1.5)),
                ((0.4, -0.9, -0.3), (-2.0, -1.0, 1.5))]


@pytest.mark.parametrize("pbr", [False, True])
def test_parity_camera_inside_scene_clipping(pbr):
    from oracle import orc
    from renderengine_b200 import capi, scene as sc
    s = sc.make_scene("small")
    if pbr:
        s = sc.to_pbr(s, n_texture_sets=2)
    o = orc.Oracle(s)
    ref_shadow = o.render_shadow()
    ctx = capi.Context(s["width"], s["height"])
    ctx.upload_scene(s)
    ground = s["n_triangles"] - 2
    for eye, look in INSIDE_VIEWS:
        v = sc.view_from(s, eye, look)
        o.set_view(v); ctx.set_view(v)
        ref = o.render_forward(ref_shadow)
        ctx.render_shadow(s["draws"]); ctx.render_forward(s["draws"])
        gpu = dict(shadow=ctx.readback(capi.REB_SHADOW_F32), tri_id=ctx.readback(capi.REB_TRIID_U32), depth=ctx.readback(capi.REB_DEPTH_F32),
                   color_f32=ctx.readback(capi.REB_COLOR_F32), color_srgb8=ctx.readback(capi.REB_COLOR_SRGB8))
        assert ctx.counters()[0] == 0 and o.counters[0] == 0          # nothing dropped
        if pbr:   # normal-mapped materials: see test_parity_pbr_textured_normal_mapped
            assert np.array_equal(gpu["tri_id"], ref["tri_id"]) and np.array_equal(gpu["depth"].view(np.uint32), ref["depth"].view(np.uint32))
            d8 = np.abs(gpu["color_srgb8"].astype(np.int32) - ref["color_srgb8"].astype(np.int32)).max(axis=2)
            assert (d8 > 1).mean() <= 1e-4 and _psnr(gpu["color_f32"], ref["color_f32"]) >= 50.0
        else:
            check_parity(s, gpu, ref_shadow, ref)
    assert (ref["tri_id"] >= ground).mean() > 0.05                      # the clipped ground plane is on screen


def test_parity_shadow_pass_guard_band_clipping():
    """A light frustum zoomed 4000x into the scene pushes the ground plane and many sphere triangles far outside the
    2^20-px guard band of the shadow pass: they are clipped (K1<shadow> rare path), not dropped."""
    from oracle import orc
    from renderengine_b200 import scene as sc
    s = sc.make_scene("small")
    p = np.array(s["shadow_proj"], np.float32).copy()
    p[[0, 4, 8, 12]] *= 4000.0; p[[1, 5, 9, 13]] *= 4000.0   # rows x and y of the column-major matrix
    s["shadow_proj"] = p
    o = orc.Oracle(s)
    ref_shadow = o.render_shadow()
    ref = o.render_forward(ref_shadow)
    gpu = render_gpu(s)
    assert gpu["counters"][0] == 0 and o.counters[0] == 0
    check_parity(s, gpu, ref_shadow, ref)
    assert (ref_shadow < 1.0).mean() > 0.5


def test_parity_perspective_shadow_transform():
    """A shadow transform whose w is not 1 (perspective light frustum placed inside the scene): the forward pass takes the
    general divide-by-w path of shadow.glsl:31 (the orthographic fast path is a separate kernel variant), and the shadow
    pass clips geometry behind the light's eye plane."""
    from oracle import orc
    from renderengine_b200 import scene as sc
    s = sc.make_scene("small")
    s["shadow_view"] = sc.to_flat(sc.camera_view(np.array([0.5, 6.0, 0.8], np.float32), np.array([0.0, -1.0, 0.0], np.float32), np.array([0, 0, 1], np.float32)))
    s["shadow_proj"] = sc.to_flat(sc.perspective_clip(100.0, np.float32(1.0), 0.5, 40.0))
    o = orc.Oracle(s)
    ref_shadow = o.render_shadow()
    ref = o.render_forward(ref_shadow)
    gpu = render_gpu(s)
    assert gpu["counters"][0] == 0 and o.counters[0] == 0
    check_parity(s, gpu, ref_shadow, ref)
    assert 0.2 < (ref_shadow < 1.0).mean()


@pytest.mark.parametrize("pbr", [False, True])
def test_parity_skybox_background(pbr):
    """the cube-map background (skybox.vert / skybox.frag, Appendix B.14) behind Phong and PBR frames,
    for the framing camera, a camera inside the scene looking at the horizon, and straight up (face seams and corners)."""
    from oracle import orc
    from renderengine_b200 import capi, scene as sc
    s = sc.make_scene("small")
    if pbr:
        s = sc.to_pbr(s)
    s["skybox"] = sc.make_skybox(48)
    o = orc.Oracle(s)
    ref_shadow = o.render_shadow()
    ctx = capi.Context(s["width"], s["height"])
    ctx.upload_scene(s)
    views = [np.asarray(s["view"]), sc.view_from(s, (0.3, 1.2, 0.4), (6.0, 1.4, 5.0)), sc.view_from(s, (0.0, 1.5, 0.0), (0.05, 9.0, 0.02)),
             sc.view_from(s, (0.0, 1.0, 0.0), (5.0, 5.0, 5.0), fov=100.0)]
    for v in views:
        o.set_view(v); ctx.set_view(v)
        ref = o.render_forward(ref_shadow)
        ctx.render_shadow(s["draws"]); ctx.render_forward(s["draws"])
        gpu = dict(shadow=ctx.readback(capi.REB_SHADOW_F32), tri_id=ctx.readback(capi.REB_TRIID_U32), depth=ctx.readback(capi.REB_DEPTH_F32),
                   color_f32=ctx.readback(capi.REB_COLOR_F32), color_srgb8=ctx.readback(capi.REB_COLOR_SRGB8))
        check_parity(s, gpu, ref_shadow, ref)
        sky = ref["tri_id"] == 0xFFFFFFFF
        assert sky.mean() > 0.02
        # the background is the same individually rounded arithmetic on both sides: bit-exact
        assert np.array_equal(gpu["color_f32"][sky].view(np.uint32), ref["color_f32"][sky].view(np.uint32))
        assert (ref["color_f32"][sky][:, :3].max(axis=1) > 0.05).all()
    # removing the skybox restores the clear colour
    ctx.set_skybox(None)
    ctx.render_forward(s["draws"])
    c = ctx.readback(capi.REB_COLOR_SRGB8)
    assert (c[sky] == np.array([0, 0, 0, 255], np.uint8)).all()


def test_set_skybox_rejects_bad_arguments():
    from renderengine_b200 import capi
    ctx = capi.Context(64, 64)
    with pytest.raises(ValueError):
        ctx.set_skybox(np.zeros((5, 8, 8, 4), np.uint16))
    with pytest.raises(capi.RebError):
        ctx._ck(ctx.lib.reb_set_skybox(ctx.h, np.zeros(8, np.uint16).ctypes.data, 0, 4))


@pytest.mark.parametrize("variant", ["phong", "pbr_sky", "textured"])
def test_parity_msaa4(variant):
    """the base pass with 4 samples per pixel at the Vulkan standard locations, one shading
    per pixel and primitive, average resolve.  All four depth / triangle-ID planes bit-exact, resolved colour <= 1 LSB."""
    from oracle import orc
    from renderengine_b200 import capi, scene as sc
    s = sc.make_scene("small_tex" if variant == "textured" else "small")
    if variant == "pbr_sky":
        s = sc.to_pbr(s)
        s["skybox"] = sc.make_skybox(32)
    o = orc.Oracle(s)
    ref_shadow = o.render_shadow()
    ctx = capi.Context(s["width"], s["height"])
    ctx.upload_scene(s)
    ctx.set_sample_count(4)
    views = [np.asarray(s["view"]), sc.view_from(s, (0.3, 1.2, 0.4), (6.0, 0.0, 5.0))]   # framing camera; inside the scene (clipped ground)
    for v in views:
        o.set_view(v); ctx.set_view(v)
        ref = o.render_forward_msaa4(ref_shadow)
        ctx.render_shadow(s["draws"]); ctx.render_forward(s["draws"])
        assert np.array_equal(ctx.readback(capi.REB_SHADOW_F32).view(np.uint32), ref_shadow.view(np.uint32))
        tri = ctx.readback(capi.REB_MS_TRIID_U32); dep = ctx.readback(capi.REB_MS_DEPTH_F32)
        for k in range(4):
            assert np.array_equal(tri[k], ref["tri_id"][k]), f"sample {k}: triangle IDs differ at {np.count_nonzero(tri[k] != ref['tri_id'][k])} pixels"
            assert np.array_equal(dep[k].view(np.uint32), ref["depth"][k].view(np.uint32)), f"sample {k}: depth not bit-exact"
        assert np.array_equal(ctx.readback(capi.REB_TRIID_U32), ref["tri_id"][0])           # the 1-plane targets hold sample 0
        assert np.array_equal(ctx.readback(capi.REB_DEPTH_F32).view(np.uint32), ref["depth"][0].view(np.uint32))
        c8 = ctx.readback(capi.REB_COLOR_SRGB8); cf = ctx.readback(capi.REB_COLOR_F32)
        d8 = np.abs(c8.astype(np.int32) - ref["color_srgb8"].astype(np.int32))
        assert d8.max() <= 1, f"sRGB8 colour differs by {d8.max()} LSB at {np.count_nonzero(d8 > 1)} channel samples"
        assert _psnr(cf, ref["color_f32"]) >= 50.0
        t = ref["tri_id"]
        assert (~((t[0] == t[1]) & (t[1] == t[2]) & (t[2] == t[3]))).mean() > 0.01          # there are edge pixels to resolve
    # back to one sample: the ordinary contract again
    ctx.set_sample_count(1)
    ctx.render_shadow(s["draws"]); ctx.render_forward(s["draws"])
    one = o.render_forward(ref_shadow)
    assert np.array_equal(ctx.readback(capi.REB_TRIID_U32), one["tri_id"])
    assert np.abs(ctx.readback(capi.REB_COLOR_SRGB8).astype(np.int32) - one["color_srgb8"].astype(np.int32)).max() <= 1


def test_set_sample_count_rejects_unsupported_counts():
    from renderengine_b200 import capi
    ctx = capi.Context(64, 64)
    for n in (0, 2, 8):
        with pytest.raises(capi.RebError):
            ctx.set_sample_count(n)


def test_shadow_pass_row_slabs_compose_the_full_map():
    """reb_set_shadow_partition: every rank rasterises only its slab of shadow-map rows (draws outside it are culled per
    rank); the slabs put together are the oracle's map bit for bit, the rows a rank does not own keep the clear value."""
    from oracle import orc
    from renderengine_b200 import capi, parallel, scene as sc
    s = sc.make_scene("small")
    S, world = s["shadow_size"], 4
    rows = parallel.slab_rows(S, world)
    ref = orc.Oracle(s).render_shadow()
    merged = np.zeros_like(ref)
    for rank in range(world):
        ctx = capi.Context(s["width"], s["height"])
        ctx.upload_scene(s)
        ctx.set_shadow_partition(rank, world, rows)
        ctx.render_shadow(s["draws"]); ctx.sync()
        got = ctx.readback(capi.REB_SHADOW_F32)
        own = slice(rank * rows, (rank + 1) * rows)
        assert np.array_equal(got[own].view(np.uint32), ref[own].view(np.uint32))
        other = np.ones(S, bool); other[own] = False
        assert (got[other] == 1.0).all()
        merged[own] = got[own]
        assert ctx.counters()[0] == 0
    assert np.array_equal(merged.view(np.uint32), ref.view(np.uint32)) and (ref < 1.0).mean() > 0.05


@pytest.mark.parametrize("inside", [False, True])
def test_sort_first_contiguous_bands_with_draw_culling(inside):
    """One contiguous band per rank: draws whose projected bounds miss the band are skipped by K0 / K1 (k_cull_draws); the
    result is still the oracle's band bit for bit — also with the camera inside the scene, where bounding boxes cross the
    eye plane (kept, the clipper decides)."""
    from oracle import orc
    from renderengine_b200 import capi, parallel, scene as sc
    s = sc.make_scene("small")
    if inside:
        s = dict(s); s["view"] = sc.view_from(s, (0.3, 1.2, 0.4), (6.0, 0.0, 5.0))
    world = 4
    band = parallel.contiguous_band(s["height"], world)
    full = orc.Oracle(s)
    ref_shadow = full.render_shadow()
    whole = full.render_forward(ref_shadow)
    merged = np.full_like(whole["tri_id"], 0xFFFFFFFF)
    for rank in range(world):
        ref = orc.Oracle(s, rank=rank, n_ranks=world, band_h=band).render_forward(ref_shadow)
        gpu = render_gpu(s, rank=rank, n_ranks=world, band_h=band)
        check_parity(s, gpu, ref_shadow, ref)
        own = ((np.arange(s["height"]) // band) % world == rank)
        merged[own] = gpu["tri_id"][own]
    assert np.array_equal(merged, whole["tri_id"])


def test_msaa4_and_skybox_under_the_sort_first_split():
    """The widened paths compose with the split: a rank renders its bands of the 4x multisampled, skybox-backed frame."""
    from oracle import orc
    from renderengine_b200 import capi, parallel, scene as sc
    s = sc.make_scene("small")
    s["skybox"] = sc.make_skybox(32)
    world = 2
    band = parallel.contiguous_band(s["height"], world * 2)
    ref_shadow = orc.Oracle(s).render_shadow()
    whole = orc.Oracle(s).render_forward_msaa4(ref_shadow)
    merged = np.zeros_like(whole["color_srgb8"])
    for rank in range(world):
        ref = orc.Oracle(s, rank=rank, n_ranks=world, band_h=band).render_forward_msaa4(ref_shadow)
        ctx = capi.Context(s["width"], s["height"], rank=rank, n_ranks=world, band_h=band)
        ctx.upload_scene(s)
        ctx.set_sample_count(4)
        ctx.render_shadow(s["draws"]); ctx.render_forward(s["draws"])
        own = ((np.arange(s["height"]) // band) % world == rank)
        tri = ctx.readback(capi.REB_MS_TRIID_U32); dep = ctx.readback(capi.REB_MS_DEPTH_F32)
        for k in range(4):
            assert np.array_equal(tri[k][own], ref["tri_id"][k][own]) and np.array_equal(dep[k][own].view(np.uint32), ref["depth"][k][own].view(np.uint32))
        c8 = ctx.readback(capi.REB_COLOR_SRGB8)
        assert np.abs(c8[own].astype(np.int32) - whole["color_srgb8"][own].astype(np.int32)).max() <= 1
        merged[own] = c8[own]
    assert np.abs(merged.astype(np.int32) - whole["color_srgb8"].astype(np.int32)).max() <= 1


def test_set_shadow_partition_rejects_bad_arguments():
    from renderengine_b200 import capi
    ctx = capi.Context(64, 64)
    for args in ((2, 2, 32), (0, 0, 32), (0, 2, 0), (0, 2, 24)):      # rank >= n, n == 0, band 0, band not a multiple of 32
        with pytest.raises(capi.RebError):
            ctx.set_shadow_partition(*args)
    ctx.set_shadow_partition(1, 2, 64)


def test_parity_msaa4_binned_path_and_queue_regrow():
    """150 large overlapping quads under 4x multisampling: every sample pass goes through the huge list, the binned tile lists
    and (first frame) the tile-list regrow + frame replay; the sample planes must still be the oracle's bit for bit."""
    from oracle import orc
    from renderengine_b200 import capi
    s = _quads_scene(150)
    o = orc.Oracle(s)
    ref_shadow = o.render_shadow()
    ref = o.render_forward_msaa4(ref_shadow)
    ctx = capi.Context(s["width"], s["height"])
    ctx.upload_scene(s)
    ctx.set_sample_count(4)
    for frame in range(2):
        ctx.render_shadow(s["draws"]); ctx.render_forward(s["draws"])
        tri = ctx.readback(capi.REB_MS_TRIID_U32); dep = ctx.readback(capi.REB_MS_DEPTH_F32)
        for k in range(4):
            assert np.array_equal(tri[k], ref["tri_id"][k]) and np.array_equal(dep[k].view(np.uint32), ref["depth"][k].view(np.uint32))
        d8 = np.abs(ctx.readback(capi.REB_COLOR_SRGB8).astype(np.int32) - ref["color_srgb8"].astype(np.int32))
        assert d8.max() <= 1
    assert ctx.counters()[1] > 0


@pytest.mark.parametrize("variant", ["phong", "pbr_sky", "textured", "opaque_alpha"])
def test_parity_translucent_pass(variant):
    """the translucent list drawn after the opaque one — cull none, LessOrEqual + depth write,
    ordered src-alpha blending, skybox afterwards.  Depth and triangle IDs bit-exact, colour <= 1 LSB / PSNR >= 50 dB."""
    from oracle import orc
    from renderengine_b200 import capi, scene as sc
    s = sc.make_scene("small_tex" if variant == "textured" else "small")
    if variant == "pbr_sky":
        s = sc.to_pbr(s)
        s["skybox"] = sc.make_skybox(32)
    alpha = 1.0 if variant == "opaque_alpha" else 0.45
    first = sc.to_translucent(s, every=2, alpha=alpha)
    ref_shadow = orc.Oracle(dict(s)).render_shadow()      # the shadow list is the opaque list followed by the translucent one (RenderScene.cpp:191-208): all draws
    ctx = capi.Context(s["width"], s["height"])
    ctx.upload_scene(first)
    for az in (0.0, 2.1):
        v = sc.view_for_azimuth(s, az)
        st = sc.to_translucent(s, every=2, alpha=alpha, view=v)                                         # split + sorted back to front for this camera
        st["view"] = v
        assert len(st["translucent_draws"]) > 0
        ref = orc.Oracle(st).render_forward(ref_shadow)
        ctx.set_view(v)
        ctx.render_shadow(np.concatenate([st["draws"], st["translucent_draws"]]))
        ctx.render_forward(st["draws"], st["translucent_draws"])
        gpu = dict(tri_id=ctx.readback(capi.REB_TRIID_U32), depth=ctx.readback(capi.REB_DEPTH_F32),
                   color_f32=ctx.readback(capi.REB_COLOR_F32), color_srgb8=ctx.readback(capi.REB_COLOR_SRGB8))
        assert ctx.counters()[0] == 0                                                                   # nothing dropped
        assert np.array_equal(gpu["tri_id"], ref["tri_id"]), f"triangle IDs differ at {np.count_nonzero(gpu['tri_id'] != ref['tri_id'])} pixels"
        assert np.array_equal(gpu["depth"].view(np.uint32), ref["depth"].view(np.uint32))
        d8 = np.abs(gpu["color_srgb8"].astype(np.int32) - ref["color_srgb8"].astype(np.int32))
        assert d8.max() <= 1, f"sRGB8 colour differs by {d8.max()} LSB at {np.count_nonzero(d8 > 1)} channel samples"
        assert _psnr(gpu["color_f32"], ref["color_f32"]) >= 50.0
        n_opaque = int(sum(int(d[1]) // 3 for d in st["draws"]))
        assert ((ref["tri_id"] != 0xFFFFFFFF) & (ref["tri_id"] >= n_opaque)).mean() > 0.03              # translucent fragments are on screen
    # without the list the same context renders the opaque frame again
    ctx.render_forward(st["draws"])
    so = dict(st); so.pop("translucent_draws")
    ref = orc.Oracle(so).render_forward(ref_shadow)
    assert np.array_equal(ctx.readback(capi.REB_TRIID_U32), ref["tri_id"])
    assert np.abs(ctx.readback(capi.REB_COLOR_SRGB8).astype(np.int32) - ref["color_srgb8"].astype(np.int32)).max() <= 1


def test_parity_translucent_triangles_clipped_at_the_eye_plane():
    """Appendix B.3 on the translucent list: camera inside the scene, the translucent ground plane crosses the eye plane and
    the guard band — its clip children are rasterised (cull none applies per child) and blended with the parent's ID."""
    from oracle import orc
    from renderengine_b200 import capi, scene as sc
    s = sc.make_scene("small")
    n_mat = np.asarray(s["materials"]).shape[0]
    which = [1, 3, n_mat - 1]                                   # the last material is the ground plane's
    ref_shadow = orc.Oracle(dict(s)).render_shadow()
    ctx = capi.Context(s["width"], s["height"])
    ctx.upload_scene(sc.to_translucent(s, which=which))
    for eye, look in INSIDE_VIEWS:
        v = sc.view_from(s, eye, look)
        st = sc.to_translucent(s, which=which, view=v)
        st["view"] = v
        ground_id = int(sum(int(d[1]) // 3 for d in st["draws"])) + int(sum(int(d[1]) // 3 for d in st["translucent_draws"][:[int(d[1]) for d in st["translucent_draws"]].index(6)]))
        o = orc.Oracle(st)
        ref = o.render_forward(ref_shadow)
        ctx.set_view(v)
        ctx.render_shadow(np.concatenate([st["draws"], st["translucent_draws"]]))
        ctx.render_forward(st["draws"], st["translucent_draws"])
        gpu = dict(tri_id=ctx.readback(capi.REB_TRIID_U32), depth=ctx.readback(capi.REB_DEPTH_F32),
                   color_f32=ctx.readback(capi.REB_COLOR_F32), color_srgb8=ctx.readback(capi.REB_COLOR_SRGB8))
        assert ctx.counters()[0] == 0 and o.counters[0] == 0    # nothing dropped
        assert np.array_equal(gpu["tri_id"], ref["tri_id"]), f"triangle IDs differ at {np.count_nonzero(gpu['tri_id'] != ref['tri_id'])} pixels"
        assert np.array_equal(gpu["depth"].view(np.uint32), ref["depth"].view(np.uint32))
        d8 = np.abs(gpu["color_srgb8"].astype(np.int32) - ref["color_srgb8"].astype(np.int32))
        assert d8.max() <= 1, f"sRGB8 colour differs by {d8.max()} LSB at {np.count_nonzero(d8 > 1)} channel samples"
        assert _psnr(gpu["color_f32"], ref["color_f32"]) >= 50.0
    assert ((ref["tri_id"] == ground_id) | (ref["tri_id"] == ground_id + 1)).mean() > 0.05       # the clipped translucent ground is on screen


def test_translucent_pass_under_the_sort_first_split():
    """The translucent list composes with the row-band split: each rank builds fragment lists only for its own rows (draw /
    block culling applies to the opaque list; the translucent list is rasterised unculled) and the bands merge to the full frame."""
    from oracle import orc
    from renderengine_b200 import capi, parallel, scene as sc
    base = sc.make_scene("small")
    base["skybox"] = sc.make_skybox(32)
    s = sc.to_translucent(base, every=2)
    world = 2
    band = parallel.contiguous_band(s["height"], world * 2)
    ref_shadow = orc.Oracle(dict(base)).render_shadow()
    whole = orc.Oracle(s).render_forward(ref_shadow)
    merged8 = np.zeros_like(whole["color_srgb8"]); merged_id = np.zeros_like(whole["tri_id"]); merged_z = np.zeros_like(whole["depth"])
    for rank in range(world):
        ctx = capi.Context(s["width"], s["height"], rank=rank, n_ranks=world, band_h=band)
        ctx.upload_scene(s)
        ctx.render_shadow(np.concatenate([s["draws"], s["translucent_draws"]])); ctx.render_forward(s["draws"], s["translucent_draws"])
        own = ((np.arange(s["height"]) // band) % world == rank)
        merged8[own] = ctx.readback(capi.REB_COLOR_SRGB8)[own]
        merged_id[own] = ctx.readback(capi.REB_TRIID_U32)[own]; merged_z[own] = ctx.readback(capi.REB_DEPTH_F32)[own]
        assert ctx.counters()[0] == 0
    assert np.array_equal(merged_id, whole["tri_id"]) and np.array_equal(merged_z.view(np.uint32), whole["depth"].view(np.uint32))
    assert np.abs(merged8.astype(np.int32) - whole["color_srgb8"].astype(np.int32)).max() <= 1


def test_translucent_many_layers_have_no_per_pixel_limit():
    """The same translucent draws submitted 3 times: coincident fragments pass LessOrEqual every time, so sphere pixels blend
    6 layers (front + back face per repeat; 12 where two spheres overlap) in submission order — parity as usual.  Submitted 10 times a pixel holds 20-40
    fragments, more than the 16 a thread sorts in place: those lists are replayed by selection, nothing is dropped, parity again."""
    from oracle import orc
    from renderengine_b200 import capi, scene as sc
    s = sc.to_translucent(sc.make_scene("small"), every=2, alpha=0.2)
    ref_shadow = orc.Oracle(dict(sc.make_scene("small"))).render_shadow()
    ctx = capi.Context(s["width"], s["height"])
    ctx.upload_scene(s)
    ctx.render_shadow(np.concatenate([s["draws"], s["translucent_draws"]]))
    s5 = dict(s); s5["translucent_draws"] = np.concatenate([s["translucent_draws"]] * 3)
    ref = orc.Oracle(s5).render_forward(ref_shadow)
    ctx.render_forward(s5["draws"], s5["translucent_draws"])
    ctx.sync()
    assert ctx.counters()[0] == 0
    tid = ctx.readback(capi.REB_TRIID_U32)
    assert np.array_equal(tid, ref["tri_id"]), f"triangle IDs differ at {np.count_nonzero(tid != ref['tri_id'])} pixels"
    assert np.array_equal(ctx.readback(capi.REB_DEPTH_F32).view(np.uint32), ref["depth"].view(np.uint32))
    assert np.abs(ctx.readback(capi.REB_COLOR_SRGB8).astype(np.int32) - ref["color_srgb8"].astype(np.int32)).max() <= 1
    assert _psnr(ctx.readback(capi.REB_COLOR_F32), ref["color_f32"]) >= 50.0
    s10 = dict(s); s10["translucent_draws"] = np.concatenate([s["translucent_draws"]] * 10)
    ref10 = orc.Oracle(s10).render_forward(ref_shadow)
    ctx.render_forward(s10["draws"], s10["translucent_draws"])
    ctx.sync()
    assert ctx.counters()[0] == 0                             # nothing dropped
    assert np.array_equal(ctx.readback(capi.REB_TRIID_U32), ref10["tri_id"])
    assert np.array_equal(ctx.readback(capi.REB_DEPTH_F32).view(np.uint32), ref10["depth"].view(np.uint32))
    assert np.abs(ctx.readback(capi.REB_COLOR_SRGB8).astype(np.int32) - ref10["color_srgb8"].astype(np.int32)).max() <= 1
    assert _psnr(ctx.readback(capi.REB_COLOR_F32), ref10["color_f32"]) >= 50.0
    ctx.render_forward(s5["draws"], s5["translucent_draws"])  # and back to the short lists
    assert np.array_equal(ctx.readback(capi.REB_TRIID_U32), ref["tri_id"]) and ctx.counters()[0] == 0
