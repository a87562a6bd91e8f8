"""SURVEY.md §7 step-2 known-answer tests of the raster contract, on the CPU oracle here and on the device (-m gpu) — see tests/kats.py."""
import numpy as np
import pytest

import kats


def render_oracle(scene):
    from oracle import orc
    o = orc.Oracle(scene)
    shadow = o.render_shadow() if scene["shadow_size"] else None
    out = o.render_forward(shadow)
    out["shadow"] = shadow
    return out


def render_device(scene):
    from renderengine_b200 import capi
    ctx = capi.Context(scene["width"], scene["height"])
    ctx.upload_scene(scene)
    if scene["shadow_size"]:
        ctx.render_shadow(scene["draws"])
    ctx.render_forward(scene["draws"])
    ctx.sync()
    out = dict(depth=ctx.readback(capi.REB_DEPTH_F32), tri_id=ctx.readback(capi.REB_TRIID_U32), color_f32=ctx.readback(capi.REB_COLOR_F32),
               color_srgb8=ctx.readback(capi.REB_COLOR_SRGB8))
    out["shadow"] = ctx.readback(capi.REB_SHADOW_F32) if scene["shadow_size"] else None
    return out


@pytest.mark.parametrize("kat", kats.ALL, ids=lambda f: f.__name__)
def test_kat_oracle(kat):
    kat(render_oracle)


@pytest.mark.gpu
@pytest.mark.parametrize("kat", kats.ALL, ids=lambda f: f.__name__)
def test_kat_device(kat):
    kat(render_device)
