"""CPU-side checks of the drop-in boundary: libreb.so loads, exports every symbol include/reb.h declares,
the ctypes mirror agrees with the header, and the product fails loudly (no CPU fallback) without a GPU."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_functions():
    src = open(os.path.join(ROOT, "include", "reb.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(reb_[a-z_0-9]+)\s*\(", src)))


def test_header_declares_the_expected_entry_points():
    fns = _header_functions()
    for must in ("reb_create", "reb_destroy", "reb_upload_geometry", "reb_upload_transforms", "reb_upload_materials", "reb_upload_lights",
                 "reb_register_texture", "reb_shadow_create", "reb_shadow_set_transform", "reb_set_view", "reb_render_shadow",
                 "reb_render_forward", "reb_readback", "reb_sync"):
        assert must in fns


def test_library_exports_every_declared_symbol():
    from renderengine_b200 import build, capi
    build.build()
    lib = C.CDLL(capi.LIB_PATH)
    missing = [f for f in _header_functions() if not hasattr(lib, f)]
    assert not missing, f"declared in include/reb.h but not exported by libreb.so: {missing}"
    assert sorted(capi.SYMBOLS) == _header_functions(), "renderengine_b200/capi.py SYMBOLS is out of sync with include/reb.h"
    assert lib.reb_abi_version() == 1


def test_ctypes_struct_layout_matches_header():
    from renderengine_b200 import capi
    # reb_params: 10 x u32/i32 + pointer (8-byte aligned) = 48 bytes on LP64; reb_draw 16; reb_texture_desc 16
    assert C.sizeof(capi.RebParams) == 48
    assert C.sizeof(capi.RebTextureDesc) == 16


def test_no_cpu_fallback_without_a_device():
    """On a box without a GPU the product must refuse to create a context (never render on the CPU)."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from renderengine_b200 import capi
    with pytest.raises(capi.RebError) as e:
        capi.Context(64, 64)
    assert "no CUDA device" in str(e.value) or "failed" in str(e.value)


def test_create_rejects_bad_extents_before_touching_a_device():
    """Argument validation of reb_create needs no GPU: zero extents and extents above REB_MAX_EXTENT (K1 packs x | y << 16 and 16-bit
    ring fields) are REB_E_INVALID with a message."""
    from renderengine_b200 import capi
    for w, h in ((0, 64), (64, 0), (16385, 64), (64, 70000)):
        with pytest.raises(capi.RebError) as e:
            capi.Context(w, h)
        assert "(-1)" in str(e.value) and "extent" in str(e.value), str(e.value)


def test_product_does_not_import_the_oracle():
    """oracle/ is test infrastructure: nothing under renderengine_b200/ may import, link or dlopen it."""
    bad = []
    for dp, _, files in os.walk(os.path.join(ROOT, "renderengine_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp", ".inl")):
                txt = open(os.path.join(dp, f), errors="ignore").read()
                if re.search(r"from\s+oracle|import\s+oracle|liboracle|oracle\.h|orc_render", txt):
                    bad.append(os.path.join(dp, f))
    assert not bad, bad


def test_host_facade_compiles_and_links():
    from renderengine_b200 import build
    build.build()
    exe = build.build_host()
    assert os.path.exists(exe) and os.access(exe, os.X_OK)
