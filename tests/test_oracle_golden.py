"""Pins the CPU oracle's arithmetic against golden vectors produced by the REFERENCE's own code
(glm, Camera.cpp, BoundingBox.cpp, phong.glsl, shadow.glsl compiled as C++ — tests/golden/gen_golden.py).
Bit-exact unless stated."""
import ctypes as C

import numpy as np

from oracle import orc

fp, u8p = orc.fp, orc.u8p


def P(a):
    return a.ctypes.data_as(fp)


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def test_mat4_ops_bit_exact(golden):
    lib = orc.lib()
    A, B, V = golden["mat_a"], golden["mat_b"], golden["vec"]
    for i in range(len(A)):
        mm = np.zeros(16, np.float32); mv = np.zeros(4, np.float32); nm = np.zeros(9, np.float32)
        lib.orc_mat4_mul(P(A[i].copy()), P(B[i].copy()), P(mm))
        lib.orc_mat4_mul_vec4(P(A[i].copy()), P(V[i].copy()), P(mv))
        lib.orc_normal_matrix(P(A[i].copy()), P(nm))
        assert np.array_equal(bits(mm), bits(golden["mat_mul"][i]))
        assert np.array_equal(bits(mv), bits(golden["mat_vec"][i]))
        assert np.array_equal(bits(nm), bits(golden["normal_matrix"][i]))


def test_phong_lighting_bit_exact(golden):
    lib = orc.lib()
    g = golden
    for i in range(len(g["ph_out"])):
        out = np.zeros(4, np.float32)
        lib.orc_phong_lighting(g["ph_lights"][i].copy().ctypes.data_as(u8p), P(g["ph_props"][i].copy()), P(g["ph_normal"][i].copy()),
                               P(g["ph_frag"][i].copy()), P(g["ph_view"][i].copy()), C.c_float(float(g["ph_shadow"][i])), P(out))
        assert np.array_equal(bits(out), bits(g["ph_out"][i])), i


def test_compute_shadow_bit_exact(golden):
    lib = orc.lib()
    g = golden
    smap = np.ascontiguousarray(g["cs_map"])
    got = np.zeros(len(g["cs_out"]), np.float32)
    for i in range(len(got)):
        got[i] = lib.orc_compute_shadow(g["ph_lights"][i].copy().ctypes.data_as(u8p), P(g["ph_frag"][i].copy()), P(g["ph_normal"][i].copy()),
                                        P(g["cs_transform"][i].copy()), P(smap), smap.shape[0])
    assert np.array_equal(bits(got), bits(g["cs_out"]))
    assert len(np.unique(got)) > 3   # the vectors exercise partial PCF coverage, not just 0/1


def test_pbr_lighting_bit_exact():
    """BRDF_Lighting (pbr.glsl:368-451, the reference's live fragment path) for untextured materials: all three light
    types, 1-3 lights, PCF shadow, every debugInput / debugEquation view."""
    import os
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "ref_pbr_vectors.npz"))
    lib = orc.lib()
    lib.orc_pbr_lighting.argtypes = [u8p, u8p, C.c_uint32, fp, fp, fp, C.c_float, C.c_uint32, C.c_uint32, fp, fp, C.c_uint32, fp]
    S = g["smap"].shape[1]
    bad = 0
    for i in range(len(g["out"])):
        out = np.zeros(4, np.float32)
        m = np.ascontiguousarray(g["mats"][i]); l = np.ascontiguousarray(g["lights"][i]); sm = np.ascontiguousarray(g["smap"][i])
        lib.orc_pbr_lighting(m.ctypes.data_as(u8p), l.ctypes.data_as(u8p), int(g["n_lights"][i]), P(np.ascontiguousarray(g["frag"][i])),
                             P(np.ascontiguousarray(g["normal"][i])), P(np.ascontiguousarray(g["view"][i])), C.c_float(float(g["exposure"][i])),
                             int(g["dbg_in"][i]), int(g["dbg_eq"][i]), P(np.ascontiguousarray(g["transform"][i])), P(sm), S, P(out))
        if not np.array_equal(bits(out), bits(g["out"][i])):
            bad += 1
    assert bad == 0, f"{bad} of {len(g['out'])} PBR vectors differ from the reference shader"
