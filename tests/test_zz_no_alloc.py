"""reb_shadow_depth_static (include/reb.h): the allocation-free shadow depth pass of one draw, whose scratch is a static device
array.  Checked through the C ABI against the context path (reb_render_shadow) on the same input, bit-exact like every other
depth comparison of this suite, on calls in a row that share the scratch.  Collected last on purpose:
a failure here hides nothing else."""
import numpy as np
import pytest


def _model():
    from renderengine_b200 import scene as sc
    c, s = np.float32(np.cos(0.4)), np.float32(np.sin(0.4))
    m = np.eye(4, dtype=np.float32)
    m[0, 0], m[0, 2], m[2, 0], m[2, 2] = c * 1.3, s * 1.3, -s * 1.3, c * 1.3
    m[1, 1] = 1.1
    m[0, 3], m[1, 3], m[2, 3] = 0.7, -0.2, -0.4
    return sc.to_flat(m)


def _cases():
    """(name, verts [n, 8], indices, shadow view, shadow proj, map size): ortho light; 4000x zoom (guard-band clipping);
    perspective light inside the scene (eye-plane clipping)."""
    from renderengine_b200 import scene as sc
    s = sc.make_scene("small")
    t = sc.make_scene("tiny")
    zoom = np.array(s["shadow_proj"], np.float32).copy()
    zoom[[0, 4, 8, 12]] *= 4000.0; zoom[[1, 5, 9, 13]] *= 4000.0
    pview = sc.to_flat(sc.camera_view(np.array([0.5, 6.0, 0.8], np.float32), np.array([0.0, -1.0, 0.0], np.float32), np.array([0, 0, 1], np.float32)))
    pproj = sc.to_flat(sc.perspective_clip(100.0, np.float32(1.0), 0.5, 40.0))
    return [("ortho", s["verts"], s["indices"], s["shadow_view"], s["shadow_proj"], 512),
            ("tiny-after-small", t["verts"], t["indices"], t["shadow_view"], t["shadow_proj"], 128),   # fewer vertices than the call before: stale scratch beyond them
            ("guard-band", s["verts"], s["indices"], s["shadow_view"], zoom, 512),
            ("perspective", s["verts"], s["indices"], pview, pproj, 1024)]


def _context_path(verts, indices, model, view, proj, size):
    from renderengine_b200 import capi
    ctx = capi.Context(64, 64)
    ctx.upload_geometry(verts, indices)
    ctx.upload_transforms(np.asarray(model, np.float32).reshape(1, 16))
    ctx.shadow_create(size)
    ctx.shadow_set_transform(view, proj)
    ctx.render_shadow(np.array([[0, indices.size, 0, 0]], np.uint32))
    ctx.flush()
    ctx.sync()
    out = ctx.readback(capi.REB_SHADOW_F32)
    assert ctx.counters()[0] == 0
    ctx.close()
    return out


@pytest.mark.gpu
def test_static_scratch_shadow_pass_matches_the_context_path_on_calls_in_a_row():
    import torch
    from renderengine_b200 import capi
    model = _model()
    cases = _cases()
    # the caller owns inputs and outputs (allocated up front, as a no-allocation caller would); the entry point allocates nothing
    dev = torch.device("cuda:0")
    torch.cuda.set_device(0)
    bufs = []
    for name, verts, indices, view, proj, size in cases:
        v = torch.from_numpy(np.ascontiguousarray(verts, np.float32)).to(dev)
        i = torch.from_numpy(np.ascontiguousarray(indices, np.uint32).view(np.int32)).to(dev)
        out = torch.full((size, size), -7.0, dtype=torch.float32, device=dev)   # garbage: the call clears the map itself
        bufs.append((v, i, out))
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    stream = torch.cuda.Stream()
    for rep in range(2):                                   # the whole sequence twice: every call follows one that left its scratch behind
        free0 = torch.cuda.mem_get_info()[0]
        for (name, verts, indices, view, proj, size), (v, i, out) in zip(cases, bufs):
            capi.shadow_depth_static(v.data_ptr(), 32, v.shape[0], i.data_ptr(), i.numel(), model, view, proj, size, out.data_ptr(), stream=stream.cuda_stream)
        stream.synchronize()
        assert torch.cuda.memory_allocated() == before
        if rep == 1:                                       # (round 0 may load the module that holds the static array)
            assert torch.cuda.mem_get_info()[0] == free0, "device memory changed across allocation-free calls"
        for (name, verts, indices, view, proj, size), (v, i, out) in zip(cases, bufs):
            got = out.cpu().numpy()
            want = _context_path(verts, indices, model, view, proj, size)
            assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), \
                f"{name} (round {rep}): {np.count_nonzero(got.view(np.uint32) != want.view(np.uint32))} texels differ from reb_render_shadow"
            assert 0.02 < (want < 1.0).mean(), name       # the map is not empty
            out.fill_(-7.0)
        torch.cuda.synchronize()
    # an empty draw clears the map
    v, i, out = bufs[1]
    capi.shadow_depth_static(v.data_ptr(), 32, v.shape[0], 0, 0, model, cases[1][3], cases[1][4], 128, out.data_ptr())
    torch.cuda.synchronize()
    assert bool((out == 1.0).all())


def test_static_scratch_limits_are_rejected_with_a_clear_error():
    """Argument checks come before any CUDA call: no GPU needed.  Sizes beyond the static scratch are REB_E_OVERFLOW."""
    from renderengine_b200 import capi
    eye = np.eye(4, dtype=np.float32).reshape(16)
    with pytest.raises(capi.RebError, match="REB_STATIC_MAX_VERTICES") as e:
        capi.shadow_depth_static(0x1000, 32, capi.REB_STATIC_MAX_VERTICES + 1, 0x1000, 3, eye, eye, eye, 64, 0x1000)
    assert e.value.code == -4
    with pytest.raises(capi.RebError, match="REB_STATIC_MAX_TRIANGLES") as e:
        capi.shadow_depth_static(0x1000, 32, 3, 0x1000, 3 * (capi.REB_STATIC_MAX_TRIANGLES + 1), eye, eye, eye, 64, 0x1000)
    assert e.value.code == -4
    with pytest.raises(capi.RebError, match="size must be"):
        capi.shadow_depth_static(0x1000, 32, 3, 0x1000, 3, eye, eye, eye, 16385, 0x1000)
    with pytest.raises(capi.RebError, match="vertexStride"):
        capi.shadow_depth_static(0x1000, 10, 3, 0x1000, 3, eye, eye, eye, 64, 0x1000)
    with pytest.raises(capi.RebError, match="multiple of 3"):
        capi.shadow_depth_static(0x1000, 32, 3, 0x1000, 4, eye, eye, eye, 64, 0x1000)
