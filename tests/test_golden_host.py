"""The drop-in surface's HOST arithmetic against the reference's own code, bit for bit.

tests/golden/ref_vectors.npz (Camera.cpp / BoundingBox.cpp / glm) and ref_shadow_vectors.npz (ComputeShadowTransform replayed over
those pieces) were generated from oracle/_ref by the committed scripts.  Three consumers are pinned here:
  * the C++ facade (renderengine_b200/host: reb::Camera, reb::BoundingBox, lookAt / ortho / inverse / products, ComputeShadowTransform)
    through tests/cpp/test_golden_math.cpp;
  * the Python twins in renderengine_b200/scene.py that build every test / bench scene;
  * the oracle's composed vertex stages (primitive.vert:49-57, shadow_map.vert:48-52) against vs_primitive / vs_shadow.
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def _shadow_golden():
    return np.load(os.path.join(ROOT, "tests", "golden", "ref_shadow_vectors.npz"))


def test_cpp_facade_math_matches_reference_bit_exact(tmp_path, golden):
    g, s = golden, _shadow_golden()
    K, N, KS, NB = len(g["cam_eye"]), len(g["mat_a"]), len(s["eye"]), s["boxes"].shape[1]
    parts = [np.array([K, N, KS, NB], np.int32)]
    for k in ("cam_eye", "cam_look", "cam_up"):
        parts.append(g[k].astype(np.float32))
    parts.append(g["cam_dims"].astype(np.int32))
    for k in ("cam_view", "cam_proj", "cam_corners", "bb_min", "bb_max", "bb_out", "la_dir", "la_up", "la_out", "ortho_out", "inv_out",
              "mat_a", "mat_b", "vec", "mat_mul", "mat_vec"):
        parts.append(g[k].astype(np.float32))
    for k in ("light_dir", "eye", "look", "up"):
        parts.append(s[k].astype(np.float32))
    parts.append(s["dims"].astype(np.int32))
    for k in ("fov", "boxes", "transforms", "scene_min", "scene_max", "view", "proj", "transform"):
        parts.append(s[k].astype(np.float32))
    dump = str(tmp_path / "golden.bin")
    with open(dump, "wb") as f:
        for p in parts:
            f.write(np.ascontiguousarray(p).tobytes())
    from renderengine_b200 import build
    lib = build.build()                                  # the facade links against libreb.so (never called here)
    exe = str(tmp_path / "test_golden_math")
    subprocess.check_call(["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-Wall", "-o", exe, os.path.join(ROOT, "tests", "cpp", "test_golden_math.cpp"),
                           os.path.join(ROOT, "renderengine_b200", "host", "RenderEngine.cpp"), "-L" + os.path.dirname(lib), "-lreb",
                           "-Wl,-rpath," + os.path.dirname(lib)])
    r = subprocess.run([exe, dump], capture_output=True, text=True, timeout=60)
    assert r.returncode == 0, r.stderr
    assert "golden math OK" in r.stdout


def test_scene_py_camera_and_matrices_match_reference_bit_exact(golden):
    from renderengine_b200 import scene as sc
    g = golden
    for i in range(len(g["cam_eye"])):
        w, h = (int(v) for v in g["cam_dims"][i])
        view = sc.camera_view(g["cam_eye"][i], g["cam_look"][i], g["cam_up"][i])
        proj = sc.perspective_clip(45.0, np.float32(w) / np.float32(h), 0.01, 100.0)
        assert np.array_equal(bits(sc.to_flat(view)), bits(g["cam_view"][i])), i
        assert np.array_equal(bits(sc.to_flat(proj)), bits(g["cam_proj"][i])), i
        assert np.array_equal(bits(sc.frustum_corners(view, proj)), bits(g["cam_corners"][i].reshape(8, 3))), i
        mn, mx = sc.bbox_transform(g["bb_min"][i], g["bb_max"][i], sc.from_flat(g["cam_view"][i]))
        assert np.array_equal(bits(np.r_[mn, mx]), bits(g["bb_out"][i, 0])), i
        assert np.array_equal(bits(sc.to_flat(sc.look_at(np.zeros(3, np.float32), g["la_dir"][i], g["la_up"][i]))), bits(g["la_out"][i])), i
        b0, b1 = g["bb_min"][i], g["bb_max"][i]
        assert np.array_equal(bits(sc.to_flat(sc.ortho_clip(b0[0], b1[0], b0[1], b1[1], b0[2], b1[2]))), bits(g["ortho_out"][i])), i
        assert np.array_equal(bits(sc.to_flat(sc.mat4_inverse(sc.from_flat(g["cam_view"][i])))), bits(g["inv_out"][i])), i
    for i in range(0, len(g["mat_a"]), 4):
        assert np.array_equal(bits(sc.to_flat(sc.mat4_mul(sc.from_flat(g["mat_a"][i]), sc.from_flat(g["mat_b"][i])))), bits(g["mat_mul"][i])), i


def test_scene_py_shadow_transform_matches_reference_bit_exact():
    from renderengine_b200 import scene as sc
    s = _shadow_golden()
    for i in range(len(s["eye"])):
        w, h = (int(v) for v in s["dims"][i])
        view = sc.camera_view(s["eye"][i], s["look"][i], s["up"][i])
        proj = sc.perspective_clip(s["fov"][i], np.float32(w) / np.float32(h), 0.01, 100.0)
        boxes = [(b[:3], b[3:]) for b in s["boxes"][i]]
        sview, sproj = sc.compute_shadow_transform(s["light_dir"][i], view, proj, (s["scene_min"][i], s["scene_max"][i]), boxes, s["transforms"][i])
        assert np.array_equal(bits(sc.to_flat(sview)), bits(s["view"][i])), i
        assert np.array_equal(bits(sc.to_flat(sproj)), bits(s["proj"][i])), i
        assert np.array_equal(bits(sc.to_flat(sc.mat4_mul(sproj, sview))), bits(s["transform"][i])), i


def test_oracle_vertex_stages_match_reference_shaders_bit_exact(golden):
    """primitive.vert:49-57 and shadow_map.vert:48-52 as the oracle composes them (world position incl. the divide by w, clip
    position through proj * view formed once on the host, the normal through transpose(inverse(mat3(T))) + normalize)."""
    from oracle import orc
    lib = orc.lib()
    fp = orc.fp
    lib.orc_primitive_vert.argtypes = [fp, fp, fp, fp, fp, fp]
    lib.orc_shadow_vert.argtypes = [fp, fp, fp, fp, fp]
    g = golden

    def P(a):
        return np.ascontiguousarray(a, np.float32).ctypes.data_as(fp)
    for i in range(len(g["vs_pos"])):
        T, V, Pm, pos, nrm = (np.ascontiguousarray(g[k][i], np.float32) for k in ("vs_T", "vs_view", "vs_proj", "vs_pos", "vs_nrm"))
        out10 = np.zeros(10, np.float32); out4 = np.zeros(4, np.float32)
        lib.orc_primitive_vert(P(T), P(V), P(Pm), P(pos), P(nrm), out10.ctypes.data_as(fp))
        lib.orc_shadow_vert(P(T), P(V), P(Pm), P(pos), out4.ctypes.data_as(fp))
        assert np.array_equal(bits(out10), bits(g["vs_primitive"][i])), (i, out10, g["vs_primitive"][i])
        assert np.array_equal(bits(out4), bits(g["vs_shadow"][i])), i
