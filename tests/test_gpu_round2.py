"""Round-2 device tests through the C ABI: regressions for the advisor's findings, Phong with every light type and more lights than
ride in the kernel parameters, the full-size sort-first (C4) and view-batch (C5) configurations, run-twice determinism."""
import hashlib

import numpy as np
import pytest

from test_gpu_parity import _psnr, _quads_scene, check_parity, render_gpu

pytestmark = pytest.mark.gpu


class _DevArray:
    def __init__(self, ptr, shape, typestr):
        self.__cuda_array_interface__ = {"data": (ptr, False), "shape": shape, "typestr": typestr, "version": 2}


def _fill_shadow_map(ctx, host_map):
    """What an NCCL all-gather / broadcast does on a rank: the map's storage (reb_get_output) filled externally, then marked valid."""
    import torch
    from renderengine_b200 import capi
    S = ctx.shadow_size
    ptr, _ = ctx.output_ptr(capi.REB_SHADOW_F32)
    ctx.sync()
    t = torch.as_tensor(_DevArray(ptr, (S, S), "<f4"), device="cuda:0")
    t.copy_(torch.from_numpy(np.ascontiguousarray(host_map, np.float32)))
    torch.cuda.synchronize()
    ctx.shadow_mark_valid()


def _frame(ctx, f32=True):
    from renderengine_b200 import capi
    out = dict(tri_id=ctx.readback(capi.REB_TRIID_U32), depth=ctx.readback(capi.REB_DEPTH_F32), color_srgb8=ctx.readback(capi.REB_COLOR_SRGB8))
    if f32:
        out["color_f32"] = ctx.readback(capi.REB_COLOR_F32)
    return out


# ---- advisor finding 1: reb_set_skybox must not touch the multisample planes -----------------------------------------------------------
def test_skybox_installed_after_the_sample_count_keeps_the_sample_planes():
    """Renderer::SetSampleCount(4) before RenderScene::Init (whose Skybox::UploadToGPU calls reb_set_skybox): the sample planes the
    sample passes and the resolve use must survive the skybox upload (they were freed there and used afterwards)."""
    from oracle import orc
    from renderengine_b200 import capi, scene as sc
    s = sc.make_scene("small")
    sky = sc.make_skybox(32)
    ctx = capi.Context(s["width"], s["height"])
    ctx.set_sample_count(4)                 # first the sample count ...
    ctx.upload_scene(s)
    ctx.set_skybox(sky)                     # ... then the skybox, twice (replacing an installed one)
    ctx.set_skybox(sky)
    s2 = dict(s); s2["skybox"] = sky
    o = orc.Oracle(s2)
    ref_shadow = o.render_shadow()
    ref = o.render_forward_msaa4(ref_shadow)
    for _ in range(2):
        ctx.render_shadow(s["draws"]); ctx.render_forward(s["draws"])
        tri = ctx.readback(capi.REB_MS_TRIID_U32); dep = ctx.readback(capi.REB_MS_DEPTH_F32)
        for k in range(4):
            assert np.array_equal(tri[k], ref["tri_id"][k]) and np.array_equal(dep[k].view(np.uint32), ref["depth"][k].view(np.uint32))
        assert np.abs(ctx.readback(capi.REB_COLOR_SRGB8).astype(np.int32) - ref["color_srgb8"].astype(np.int32)).max() <= 1
    ctx.set_sample_count(1); ctx.set_sample_count(4)   # and the planes can still be re-sized / freed exactly once
    ctx.render_forward(s["draws"]); ctx.sync()
    ctx.close()


# ---- advisor finding 2: culled draws inside a mixed 256-triangle block ------------------------------------------------------------------
@pytest.mark.parametrize("world", [2, 4])
def test_split_context_renders_a_sequence_of_views(world):
    """One sort-first context, several camera views in a row (and a light that moves): a draw culled for this rank in view k shares
    256-triangle blocks with visible draws, and its cache slots must read as rejected — not as the previous view's vertices."""
    from oracle import orc
    from renderengine_b200 import capi, parallel, scene as sc
    s = sc.make_scene("small", level=2)    # 320 triangles per sphere: every 256-triangle block spans two draws
    band = parallel.contiguous_band(s["height"], world)
    full = orc.Oracle(s)
    views = [sc.view_for_azimuth(s, az) for az in (0.0, 1.3, 2.9, 4.4)] + [sc.view_from(s, (0.3, 1.2, 0.4), (6.0, 0.0, 5.0))]
    ctxs = [capi.Context(s["width"], s["height"], rank=r, n_ranks=world, band_h=band) for r in range(world)]
    for c in ctxs:
        c.upload_scene(s)
    rows = parallel.slab_rows(s["shadow_size"], world)
    for k, v in enumerate(views):
        full.set_view(v)
        ref_shadow = full.render_shadow()
        whole = full.render_forward(ref_shadow)
        merged_id = np.full_like(whole["tri_id"], 0xFFFFFFFF); merged_z = np.ones_like(whole["depth"]); merged_c = np.zeros_like(whole["color_srgb8"])
        for r, c in enumerate(ctxs):
            c.set_view(v)
            c.render_shadow(s["draws"]); c.render_forward(s["draws"])
            own = ((np.arange(s["height"]) // band) % world == r)
            g = _frame(c, f32=False)
            merged_id[own] = g["tri_id"][own]; merged_z[own] = g["depth"][own]; merged_c[own] = g["color_srgb8"][own]
            assert (g["tri_id"][~own] == 0xFFFFFFFF).all(), f"view {k}, rank {r}: pixels outside the rank's bands were written"
            assert c.counters()[0] == 0
        assert np.array_equal(merged_id, whole["tri_id"]), f"view {k}: {np.count_nonzero(merged_id != whole['tri_id'])} pixels differ"
        assert np.array_equal(merged_z.view(np.uint32), whole["depth"].view(np.uint32))
        assert np.abs(merged_c.astype(np.int32) - whole["color_srgb8"].astype(np.int32)).max() <= 1
    # the same for the shadow pass split into row slabs, light direction changing between frames
    for k, ld in enumerate([(-0.3, -1.0, -0.2), (0.6, -1.0, 0.1), (-0.1, -1.0, 0.7)]):
        ldn = np.asarray(ld, np.float32); ldn = ldn / np.linalg.norm(ldn)
        sv = sc.light_view_matrix(ldn)
        s2 = dict(s); s2["shadow_view"] = sc.to_flat(sv)
        ref = orc.Oracle(s2).render_shadow()
        got = np.zeros_like(ref)
        for r, c in enumerate(ctxs):
            c.set_shadow_partition(r, world, rows)
            c.shadow_set_transform(s2["shadow_view"], s2["shadow_proj"])
            c.render_shadow(s["draws"]); c.sync()
            got[r * rows:(r + 1) * rows] = c.readback(capi.REB_SHADOW_F32)[r * rows:(r + 1) * rows]
        assert np.array_equal(got.view(np.uint32), ref.view(np.uint32)), f"light {k}"


# ---- advisor finding 3: overflow replay of a frame with a translucent list, no f32 output enabled ---------------------------------------
def test_translucent_frame_replayed_after_queue_overflow_without_f32_output(monkeypatch):
    from oracle import orc
    from renderengine_b200 import capi, scene as sc
    monkeypatch.setenv("REB_TILE_LIST_CAP", "2048")   # the first frame overflows the tile-list queue: regrow + replay
    s = _quads_scene(150)
    s = sc.to_translucent(s, which=[1], alpha=0.4)
    ref_shadow = orc.Oracle(dict(_quads_scene(150))).render_shadow()
    ref = orc.Oracle(s).render_forward(ref_shadow)
    ctx = capi.Context(s["width"], s["height"], outputs=capi.REB_OUT_COLOR_SRGB8 | capi.REB_OUT_TRIID_U32)   # the bench's outputs: no f32 colour, no depth
    ctx.upload_scene(s)
    for frame in range(2):
        ctx.render_shadow(np.concatenate([s["draws"], s["translucent_draws"]]))
        ctx.render_forward(s["draws"], s["translucent_draws"])
        ctx.sync()
        if frame == 0:
            assert ctx.counters()[3] > 0            # the tile-list queue was regrown and the frame replayed
        assert np.array_equal(ctx.readback(capi.REB_TRIID_U32), ref["tri_id"]), frame
        d8 = np.abs(ctx.readback(capi.REB_COLOR_SRGB8).astype(np.int32) - ref["color_srgb8"].astype(np.int32))
        assert d8.max() <= 1, f"frame {frame}: {np.count_nonzero(d8 > 1)} channel samples beyond 1 LSB (max {d8.max()})"


# ---- a10: PhongLighting for every light type, and more lights than the kernel parameters carry (phong.glsl:39-81) ---------------------------
def _phong_lights(s, n_extra=0):
    from renderengine_b200 import scene as sc
    ex, ez = s["extent"]
    ldir = sc._norm([-0.3, -1.0, -0.2])
    f = np.float32
    lights = [sc.pack_light(sc.LIGHT_DIRECTIONAL, (-ldir * f(50)).astype(f), ldir, [0.05] * 4, [0.7] * 4, [0.8] * 4),
              sc.pack_light(sc.LIGHT_POINT, [f(-0.4 * ex), f(2.5), f(0.3 * ez)], [0, -1, 0], [0.02, 0.02, 0.05, 0], [0.6, 1.2, 2.0, 0], [0.5, 1.0, 2.0, 0]),
              sc.pack_light(sc.LIGHT_SPOT, [f(0.5 * ex), f(3.0), f(-0.2 * ez)], sc._norm([-0.2, -1.0, 0.1]), [0.0] * 4, [3.0, 1.2, 0.8, 0], [2.0, 1.0, 0.5, 0],
                            inner=0.92, outer=0.75)]
    rng = np.random.default_rng(5)
    for k in range(n_extra):
        kind = sc.LIGHT_POINT if k % 2 == 0 else sc.LIGHT_SPOT
        pos = np.array([rng.uniform(-ex, ex), rng.uniform(1.5, 4.0), rng.uniform(-ez, ez)], f)
        lights.append(sc.pack_light(kind, pos, sc._norm([rng.uniform(-0.4, 0.4), -1.0, rng.uniform(-0.4, 0.4)]), [0.01] * 4,
                                    np.r_[rng.uniform(0.3, 1.5, 3), 0].astype(f), np.r_[rng.uniform(0.3, 1.5, 3), 0].astype(f), inner=0.95, outer=0.6))
    return np.stack(lights)


@pytest.mark.parametrize("n_extra,config", [(0, "small"), (4, "small"), (9, "small_tex"), (3, "C1")])
def test_parity_phong_directional_point_spot_and_many_lights(n_extra, config):
    """3 lights = all of phong.glsl's branches (1 / distance attenuation, spot smooth-step, shadowed directional); 7 / 12 / 6 lights
    also take the global-memory light path past SHADE_CONST_LIGHTS = 4."""
    from oracle import orc
    from renderengine_b200 import scene as sc
    s = sc.make_scene(config) if config != "C1" else sc.make_scene(config, shadow_size=1024)
    s["lights"] = _phong_lights(s, n_extra)
    o = orc.Oracle(s)
    ref_shadow = o.render_shadow()
    ref = o.render_forward(ref_shadow)
    gpu = render_gpu(s)
    check_parity(s, gpu, ref_shadow, ref)
    one = dict(s); one["lights"] = s["lights"][:1]
    base = orc.Oracle(one).render_forward(ref_shadow)
    drawn = ref["tri_id"] != 0xFFFFFFFF
    assert (ref["color_f32"][drawn][:, :3] > base["color_f32"][drawn][:, :3] + 0.02).any(axis=1).mean() > 0.3   # the extra lights really light the frame


# ---- full-size C4 (sort-first, 7680 x 4320, 10.5 M triangles, 4096 textures) and C5 (the 2 M-triangle view-batch scene) -----------------------
def test_parity_full_size_c4_one_rank_of_eight():
    """BASELINE.json config 4 at full size: rank 3 of 8 (two interleaved bands per rank, per-draw + per-block culling, shadow row slab):
    its bands of the 8K frame and its slab of the shadow map against the oracle, bit for bit."""
    from oracle import orc
    from renderengine_b200 import capi, parallel, scene as sc
    s = sc.make_scene("C4")
    world, rank = 8, 3
    band = parallel.contiguous_band(s["height"], world * 2)
    full = orc.Oracle(s)
    ref_shadow = full.render_shadow()
    ref = orc.Oracle(s, rank=rank, n_ranks=world, band_h=band).render_forward(ref_shadow)
    outputs = capi.REB_OUT_COLOR_SRGB8 | capi.REB_OUT_DEPTH_F32 | capi.REB_OUT_TRIID_U32       # an 8K f32 colour target would be 531 MB more; sRGB8 carries the colour check
    ctx = capi.Context(s["width"], s["height"], rank=rank, n_ranks=world, band_h=band, outputs=outputs)
    ctx.upload_scene(s)
    rows = parallel.slab_rows(s["shadow_size"], world)
    ctx.set_shadow_partition(rank, world, rows)
    ctx.render_shadow(s["draws"]); ctx.sync()
    slab = ctx.readback(capi.REB_SHADOW_F32)[rank * rows:(rank + 1) * rows]
    assert np.array_equal(slab.view(np.uint32), ref_shadow[rank * rows:(rank + 1) * rows].view(np.uint32))
    # complete the map the way the all-gather would, then the rank's share of the forward pass
    _fill_shadow_map(ctx, ref_shadow)
    ctx.render_forward(s["draws"]); ctx.sync()
    assert ctx.counters()[0] == 0
    own = ((np.arange(s["height"]) // band) % world == rank)
    tri = ctx.readback(capi.REB_TRIID_U32)
    assert np.array_equal(tri[own], ref["tri_id"][own]), f"{np.count_nonzero(tri[own] != ref['tri_id'][own])} pixels differ"
    assert np.array_equal(ctx.readback(capi.REB_DEPTH_F32)[own].view(np.uint32), ref["depth"][own].view(np.uint32))
    d8 = np.abs(ctx.readback(capi.REB_COLOR_SRGB8)[own].astype(np.int32) - ref["color_srgb8"][own].astype(np.int32))
    assert d8.max() <= 1
    assert (tri[~own] == 0xFFFFFFFF).all() and (ref["tri_id"][own] != 0xFFFFFFFF).mean() > 0.1


def test_parity_full_size_c5_views_share_one_shadow_map():
    """BASELINE.json config 5's scene (2,007,042 triangles, 1080p): the shadow map rendered ONCE (RenderScene.cpp:171-175), then several
    of the batch's camera views against it — every view bit-exact against the oracle."""
    from oracle import orc
    from renderengine_b200 import capi, scene as sc
    s = sc.make_scene("C5")
    o = orc.Oracle(s)
    ref_shadow = o.render_shadow()
    ctx = capi.Context(s["width"], s["height"])
    ctx.upload_scene(s)
    ctx.render_shadow(s["draws"])
    for view_index in (0, 137, 401):
        v = sc.view_for_azimuth(s, 2 * np.pi * view_index / 512)
        o.set_view(v); ctx.set_view(v)
        ref = o.render_forward(ref_shadow)
        ctx.render_forward(s["draws"])
        gpu = _frame(ctx)
        if view_index == 0:
            gpu["shadow"] = ctx.readback(capi.REB_SHADOW_F32)
        check_parity(s, gpu, ref_shadow if view_index == 0 else None, ref)
        assert ctx.counters()[0] == 0


# ---- determinism — the same frame twice, every output hashed ----------------------------------------------------------------------
@pytest.mark.parametrize("variant", ["phong", "binned", "msaa4", "translucent", "clipped_pbr_sky"])
def test_render_twice_is_bit_identical(variant):
    """Atomics decide every pixel (64-bit min keys, list-head exchanges), queues fill in a different order every run: the OUTPUT must
    not depend on it.  Two contexts and two frames on one context, all outputs hashed."""
    from renderengine_b200 import capi, scene as sc
    s = sc.make_scene("small")
    tl = None
    if variant == "binned":
        s = _quads_scene(150)
    if variant == "translucent":
        s = sc.to_translucent(s, every=2)
        tl = s["translucent_draws"]
    if variant == "clipped_pbr_sky":
        s = sc.to_pbr(s, n_texture_sets=2)
        s["skybox"] = sc.make_skybox(32)
        s["view"] = sc.view_from(s, (0.3, 0.2, 0.4), (3.0, -0.2, 2.0))
    shadow_draws = s["draws"] if tl is None else np.concatenate([s["draws"], tl])
    digests = []
    for c in range(2):
        ctx = capi.Context(s["width"], s["height"])
        ctx.upload_scene(s)
        if variant == "msaa4":
            ctx.set_sample_count(4)
        for frame in range(2):
            ctx.render_shadow(shadow_draws); ctx.render_forward(s["draws"], tl)
            ctx.sync()
            h = hashlib.sha256()
            kinds = [capi.REB_SHADOW_F32, capi.REB_DEPTH_F32, capi.REB_TRIID_U32, capi.REB_COLOR_F32, capi.REB_COLOR_SRGB8]
            if variant == "msaa4":
                kinds += [capi.REB_MS_DEPTH_F32, capi.REB_MS_TRIID_U32]
            for kind in kinds:
                h.update(np.ascontiguousarray(ctx.readback(kind)).tobytes())
            digests.append(h.hexdigest())
    assert len(set(digests)) == 1, digests


# ---- VERDICT "missing" 6: the translucent list inside the multisampled base pass; no per-pixel layer limit ------------------------------
@pytest.mark.parametrize("variant", ["phong", "pbr_sky", "quads", "inside"])
def test_parity_translucent_list_inside_the_multisampled_pass(variant):
    """The reference draws opaque + translucent + skybox in ONE multisampled render pass (RenderScene.cpp:217-228,
    Swapchain.cpp:110-123,212).  Appendix B.15 + B.16 together: coverage / depth test / blend per sample, one shading per pixel and
    primitive at the pixel centre, skybox per sample, average resolve.  All four depth / ID planes bit-exact after the translucent
    list, resolved colour <= 1 LSB, HDR PSNR >= 50 dB.  `quads`: large overlapping translucent quads (the large-triangle work
    list of every sample pass); `inside`: camera inside the scene, the translucent ground plane is clipped (Appendix B.3)."""
    from oracle import orc
    from renderengine_b200 import capi, scene as sc
    if variant == "quads":
        base = _quads_scene(40)
        s = sc.to_translucent(base, which=[1], alpha=0.4)
    else:
        base = sc.make_scene("small")
        if variant == "pbr_sky":
            base = sc.to_pbr(base)
            base["skybox"] = sc.make_skybox(32)
        if variant == "inside":
            base["view"] = sc.view_from(base, (0.3, 1.2, 0.4), (6.0, 0.0, 5.0))
            ground = int(np.asarray(base["draws"], np.uint32)[:, 2].max() >> 4)   # the ground plane's material is the last one
            s = sc.to_translucent(base, which=[ground, 0], alpha=0.45)
        else:
            s = sc.to_translucent(base, every=2, alpha=0.45)
    assert len(s["translucent_draws"]) > 0
    all_draws = np.concatenate([s["draws"], s["translucent_draws"]])
    ref_shadow = orc.Oracle(dict(base)).render_shadow()
    o = orc.Oracle(s)
    ref = o.render_forward_msaa4(ref_shadow)
    n_opaque = int(sum(int(d[1]) // 3 for d in s["draws"]))
    ctx = capi.Context(s["width"], s["height"])
    ctx.upload_scene(s)
    ctx.set_sample_count(4)
    for frame in range(2):                               # the second frame runs over re-armed list heads
        ctx.render_shadow(all_draws)
        ctx.render_forward(s["draws"], s["translucent_draws"])
        ctx.sync()
        assert ctx.counters()[0] == 0
        tri = ctx.readback(capi.REB_MS_TRIID_U32); dep = ctx.readback(capi.REB_MS_DEPTH_F32)
        for k in range(4):
            assert np.array_equal(tri[k], ref["tri_id"][k]), f"frame {frame} sample {k}: {np.count_nonzero(tri[k] != ref['tri_id'][k])} triangle IDs differ"
            assert np.array_equal(dep[k].view(np.uint32), ref["depth"][k].view(np.uint32)), f"frame {frame} sample {k}: depth not bit-exact"
        assert np.array_equal(ctx.readback(capi.REB_TRIID_U32), ref["tri_id"][0])
        assert np.array_equal(ctx.readback(capi.REB_DEPTH_F32).view(np.uint32), ref["depth"][0].view(np.uint32))
        d8 = np.abs(ctx.readback(capi.REB_COLOR_SRGB8).astype(np.int32) - ref["color_srgb8"].astype(np.int32))
        assert d8.max() <= 1, f"frame {frame}: sRGB8 differs by {d8.max()} LSB at {np.count_nonzero(d8 > 1)} channel samples"
        assert _psnr(ctx.readback(capi.REB_COLOR_F32), ref["color_f32"]) >= 50.0
    t = ref["tri_id"]
    assert ((t != 0xFFFFFFFF) & (t >= n_opaque)).any(axis=0).mean() > 0.03      # translucent fragments are on screen
    # and back to the 1-sample contract with the same context
    ctx.set_sample_count(1)
    one = o.render_forward(ref_shadow)
    ctx.render_shadow(all_draws); ctx.render_forward(s["draws"], s["translucent_draws"]); ctx.sync()
    assert np.array_equal(ctx.readback(capi.REB_TRIID_U32), one["tri_id"])
    assert np.abs(ctx.readback(capi.REB_COLOR_SRGB8).astype(np.int32) - one["color_srgb8"].astype(np.int32)).max() <= 1


def test_parity_more_translucent_layers_than_the_sorted_fast_path_holds():
    """100 large translucent quads stacked over the frame: many pixels carry 20-40 fragments — beyond the TL_MAX_FRAGS = 16 a thread
    sorts in place, so the replay selects straight from the list.  Nothing is dropped (counters[0] == 0) and the frame equals the
    oracle's, which has no limit."""
    from oracle import orc
    from renderengine_b200 import capi, scene as sc
    base = _quads_scene(150)
    s = sc.to_translucent(base, which=[0, 1], alpha=0.3)      # 100 of the 150 quads
    ref_shadow = orc.Oracle(dict(base)).render_shadow()
    ref = orc.Oracle(s).render_forward(ref_shadow)
    ctx = capi.Context(s["width"], s["height"])
    ctx.upload_scene(s)
    ctx.render_shadow(np.concatenate([s["draws"], s["translucent_draws"]]))
    ctx.render_forward(s["draws"], s["translucent_draws"])
    ctx.sync()
    assert ctx.counters()[0] == 0
    assert np.array_equal(ctx.readback(capi.REB_TRIID_U32), ref["tri_id"])
    assert np.array_equal(ctx.readback(capi.REB_DEPTH_F32).view(np.uint32), ref["depth"].view(np.uint32))
    d8 = np.abs(ctx.readback(capi.REB_COLOR_SRGB8).astype(np.int32) - ref["color_srgb8"].astype(np.int32))
    assert d8.max() <= 1, f"{np.count_nonzero(d8 > 1)} channel samples beyond 1 LSB (max {d8.max()})"
    assert _psnr(ctx.readback(capi.REB_COLOR_F32), ref["color_f32"]) >= 50.0


def test_multisampled_pass_on_a_split_context():
    """Sort-first row bands + 4x multisampling: the one-pass multisampled K1 walks the compacted block list of the rank (CULL
    instantiation) and filters rows per sample; every rank's own rows of all four sample planes equal the unsplit oracle frame."""
    from oracle import orc
    from renderengine_b200 import capi, parallel, scene as sc
    s = sc.make_scene("small", level=2)
    world = 2
    band = parallel.contiguous_band(s["height"], 2 * world)      # two interleaved bands per rank
    o = orc.Oracle(s)
    ref_shadow = o.render_shadow()
    for v in (np.asarray(s["view"]), sc.view_from(s, (0.3, 1.2, 0.4), (6.0, 0.0, 5.0))):
        o.set_view(v)
        ref = o.render_forward_msaa4(ref_shadow)
        got_t = np.full_like(ref["tri_id"], 0xFFFFFFFF); got_z = np.ones_like(ref["depth"]); got_c = np.zeros_like(ref["color_srgb8"])
        for r in range(world):
            ctx = capi.Context(s["width"], s["height"], rank=r, n_ranks=world, band_h=band)
            ctx.upload_scene(s)
            ctx.set_sample_count(4)
            ctx.set_view(v)
            for _ in range(2):
                ctx.render_shadow(s["draws"]); ctx.render_forward(s["draws"]); ctx.sync()
            own = ((np.arange(s["height"]) // band) % world == r)
            got_t[:, own] = ctx.readback(capi.REB_MS_TRIID_U32)[:, own]
            got_z[:, own] = ctx.readback(capi.REB_MS_DEPTH_F32)[:, own]
            got_c[own] = ctx.readback(capi.REB_COLOR_SRGB8)[own]
            assert ctx.counters()[0] == 0
            ctx.close()
        assert np.array_equal(got_t, ref["tri_id"]), f"{np.count_nonzero(got_t != ref['tri_id'])} sample IDs differ"
        assert np.array_equal(got_z.view(np.uint32), ref["depth"].view(np.uint32))
        assert np.abs(got_c.astype(np.int32) - ref["color_srgb8"].astype(np.int32)).max() <= 1


def test_parity_full_size_c2_multisampled_with_translucent_list():
    """C2 at full size (1 M triangles, 1920x1080) with every second material translucent, 4x multisampled: both lists are long
    enough to be re-ordered (k_reorder.cu), the one-pass multisampled K1 and the per-sample translucent raster run at scale, the
    node pool holds four samples' fragments.  Sample planes bit-exact, colour <= 1 LSB, nothing dropped."""
    from oracle import orc
    from renderengine_b200 import capi, scene as sc
    base = sc.make_scene("C2")
    s = sc.to_translucent(base, every=2, alpha=0.45)
    all_draws = np.concatenate([s["draws"], s["translucent_draws"]])
    ref_shadow = orc.Oracle(dict(base)).render_shadow()
    ref = orc.Oracle(s).render_forward_msaa4(ref_shadow)
    ctx = capi.Context(s["width"], s["height"], outputs=capi.REB_OUT_COLOR_SRGB8 | capi.REB_OUT_DEPTH_F32 | capi.REB_OUT_TRIID_U32)
    ctx.upload_scene(s)
    ctx.set_sample_count(4)
    ctx.render_shadow(all_draws); ctx.render_forward(s["draws"], s["translucent_draws"]); ctx.sync()
    assert ctx.counters()[0] == 0
    assert np.array_equal(ctx.readback(capi.REB_SHADOW_F32).view(np.uint32), ref_shadow.view(np.uint32))
    tri = ctx.readback(capi.REB_MS_TRIID_U32); dep = ctx.readback(capi.REB_MS_DEPTH_F32)
    for k in range(4):
        assert np.array_equal(tri[k], ref["tri_id"][k]), f"sample {k}: {np.count_nonzero(tri[k] != ref['tri_id'][k])} triangle IDs differ"
        assert np.array_equal(dep[k].view(np.uint32), ref["depth"][k].view(np.uint32)), f"sample {k}: depth not bit-exact"
    d8 = np.abs(ctx.readback(capi.REB_COLOR_SRGB8).astype(np.int32) - ref["color_srgb8"].astype(np.int32))
    assert d8.max() <= 1, f"sRGB8 differs by {d8.max()} LSB at {np.count_nonzero(d8 > 1)} channel samples"
    n_opaque = int(sum(int(d[1]) // 3 for d in s["draws"]))
    assert ((ref["tri_id"] != 0xFFFFFFFF) & (ref["tri_id"] >= n_opaque)).any(axis=0).mean() > 0.05


@pytest.mark.parametrize("config,world", [("small", 3), ("quads", 2), ("C2", 8)])
def test_shadow_pass_split_by_triangles_min_of_the_ranks_maps(config, world):
    """reb_set_shadow_split: every rank rasterises its contiguous share of the triangle stream into a whole map; the element-wise
    minimum of the ranks' maps must be the oracle's map bit for bit.  `quads`: everything is large / huge (binned path + the row
    split of the broadcast list); `C2`: full size, re-ordered list, eight ranks; light moved between frames on the same contexts."""
    from oracle import orc
    from renderengine_b200 import capi, scene as sc
    s = _quads_scene(40) if config == "quads" else sc.make_scene(config)
    ctxs = []
    for r in range(world):
        c = capi.Context(64, 64)
        c.upload_geometry(s["verts"], s["indices"]); c.upload_transforms(s["transforms"])
        c.shadow_create(s["shadow_size"])
        c.set_shadow_split(r, world)
        ctxs.append(c)
    lights = [(-0.3, -1.0, -0.2), (0.6, -1.0, 0.1)] if config != "C2" else [None]
    for ld in lights:
        s2 = dict(s)
        if ld is not None:
            ldn = np.asarray(ld, np.float32); ldn = ldn / np.linalg.norm(ldn)
            s2["shadow_view"] = sc.to_flat(sc.light_view_matrix(ldn))
        ref = orc.Oracle(s2).render_shadow()
        got = np.ones_like(ref)
        share = []
        for c in ctxs:
            c.shadow_set_transform(s2["shadow_view"], s2["shadow_proj"])
            c.render_shadow(s["draws"]); c.flush(); c.sync()
            m = c.readback(capi.REB_SHADOW_F32)
            share.append(float((m < 1.0).mean()))
            got = np.minimum(got, m)
            assert c.counters()[0] == 0
        assert np.array_equal(got.view(np.uint32), ref.view(np.uint32)), f"{np.count_nonzero(got.view(np.uint32) != ref.view(np.uint32))} texels differ"
        assert min(share) < 0.9 * float((ref < 1.0).mean())                      # the work was split: some rank's map is visibly partial
    # back to an unsplit pass on the same context
    ctxs[0].set_shadow_split(0, 1)
    ctxs[0].render_shadow(s["draws"]); ctxs[0].flush(); ctxs[0].sync()
    assert np.array_equal(ctxs[0].readback(capi.REB_SHADOW_F32).view(np.uint32), ref.view(np.uint32))


def test_moving_nodes_keep_the_draw_list_preprocessing():
    """Re-uploading the same number of node transforms (objects moving between frames, SceneTree::UploadToGPU again) must not redo
    the draw-list preprocessing, and the frames must follow the new transforms: parity against the oracle for three placements."""
    from oracle import orc
    from renderengine_b200 import capi, scene as sc
    s = sc.make_scene("small")
    ctx = capi.Context(s["width"], s["height"])
    ctx.upload_scene(s)
    ctx.render_shadow(s["draws"]); ctx.render_forward(s["draws"]); ctx.sync()
    launches0 = None
    rng = np.random.default_rng(3)
    for k in range(3):
        t = np.array(s["transforms"], np.float32).reshape(-1, 16).copy()
        t[:-1, 12] += rng.uniform(-0.4, 0.4, t.shape[0] - 1).astype(np.float32)      # spheres slide in x and y; the ground stays
        t[:-1, 13] += rng.uniform(0.0, 0.5, t.shape[0] - 1).astype(np.float32)
        s2 = dict(s); s2["transforms"] = t
        o = orc.Oracle(s2)
        ref_shadow = o.render_shadow(); ref = o.render_forward(ref_shadow)
        before = ctx.launch_count()
        ctx.upload_transforms(t)
        ctx.render_shadow(s["draws"]); ctx.render_forward(s["draws"]); ctx.sync()
        n = ctx.launch_count() - before
        launches0 = n if launches0 is None else launches0
        assert n == launches0 and n <= 8, f"frame {k}: {n} launches (draw-list preprocessing re-run?)"
        assert np.array_equal(ctx.readback(capi.REB_SHADOW_F32).view(np.uint32), ref_shadow.view(np.uint32))
        assert np.array_equal(ctx.readback(capi.REB_TRIID_U32), ref["tri_id"])
        assert np.array_equal(ctx.readback(capi.REB_DEPTH_F32).view(np.uint32), ref["depth"].view(np.uint32))
        assert np.abs(ctx.readback(capi.REB_COLOR_SRGB8).astype(np.int32) - ref["color_srgb8"].astype(np.int32)).max() <= 1
