"""Golden vectors for the PBR fragment path, generated from the REFERENCE's own pbr.glsl + shadow.glsl compiled as C++
(oracle/_ref/libref.so: ref_pbr_lighting = BRDF_Lighting, pbr.glsl:368-451).  Run in the build container only; the output
tests/golden/ref_pbr_vectors.npz is committed.

    python tests/golden/gen_golden_pbr.py
"""
import ctypes as C
import os
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
lib = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libref.so"))
fp = C.POINTER(C.c_float)
u8p = C.POINTER(C.c_uint8)
INVALID = 0xFFFFFFFF


def make_pbr_light(rng, ltype):
    rec = np.zeros(80, np.uint8)
    f = rec.view(np.float32); u = rec.view(np.uint32)
    f[0:3] = rng.uniform(0.2, 1.0, 3)           # color @0
    f[4:7] = rng.uniform(-10, 10, 3)            # position @16
    d = rng.normal(size=3); d /= np.linalg.norm(d)
    f[8:11] = d                                 # direction @32
    f[11] = rng.uniform(0.5, 8.0)               # intensity @44
    f[12] = rng.uniform(0.05, 30.0)             # falloffRadius @48
    f[13] = 0.95; f[14] = rng.choice([0.80, 0.95])   # cosInner @52 / cosOuter @56 (equal: exercises the 1e-4 clamp)
    u[15] = 0                                   # shadowIndex @60
    u[16] = ltype                               # type @64
    return rec


def make_pbr_material(rng):
    rec = np.zeros(80, np.uint8)
    f = rec.view(np.float32); u = rec.view(np.uint32)
    f[0:4] = rng.uniform(0.05, 1.0, 4)          # baseColor
    f[4:8] = rng.uniform(0.0, 0.5, 4)           # emissive (+ exposure compensation in w)
    f[8] = rng.uniform(0.02, 0.08)              # f0
    f[9] = rng.choice([0.0, 1.0, rng.uniform()])   # metallic
    f[10] = rng.choice([0.0, 1.0, rng.uniform()])  # perceptualRoughness (0 exercises the 0.089 clamp)
    f[11] = rng.uniform(0.3, 1.0)               # ambientOcclusion
    u[12:17] = INVALID                          # untextured
    return rec


def main():
    rng = np.random.default_rng(4321)
    M, S = 768, 8
    mats = np.stack([make_pbr_material(rng) for _ in range(M)])
    nl = rng.integers(1, 4, M).astype(np.uint32)
    lights = np.stack([np.concatenate([make_pbr_light(rng, 1 + ((i + k) % 3)) for k in range(3)]) for i in range(M)])
    frag = rng.uniform(-4, 4, (M, 3)).astype(np.float32)
    normal = rng.normal(size=(M, 3)).astype(np.float32); normal /= np.linalg.norm(normal, axis=1, keepdims=True)
    vpos = rng.uniform(-8, 8, (M, 3)).astype(np.float32)
    exposure = rng.uniform(0.5, 2.0, M).astype(np.float32)
    dbg_in = np.where(np.arange(M) % 4 == 1, rng.integers(0, 8, M), 0).astype(np.uint32)
    dbg_eq = np.where(np.arange(M) % 4 == 2, rng.integers(0, 6, M), 0).astype(np.uint32)
    transform = np.zeros((M, 16), np.float32)
    transform[:, [0, 5, 10]] = rng.uniform(0.05, 0.2, (M, 3)); transform[:, 15] = 1.0
    transform[:, 12:15] = rng.uniform(-0.2, 0.2, (M, 3)); transform[:, 14] += 0.5
    smap = rng.uniform(0.3, 0.9, (M, S, S)).astype(np.float32)
    uv = np.zeros(2, np.float32)
    out = np.zeros((M, 4), np.float32)
    lib.ref_pbr_lighting.argtypes = [u8p, u8p, C.c_uint32, fp, fp, fp, fp, C.c_float, C.c_uint32, C.c_uint32, fp, fp, C.c_uint32, fp]
    P = lambda a: a.ctypes.data_as(fp)
    for i in range(M):
        lib.ref_pbr_lighting(mats[i].ctypes.data_as(u8p), lights[i].ctypes.data_as(u8p), int(nl[i]), P(frag[i]), P(uv), P(normal[i]), P(vpos[i]),
                             C.c_float(float(exposure[i])), int(dbg_in[i]), int(dbg_eq[i]), P(transform[i]), P(smap[i]), S, P(out[i]))
    assert np.isfinite(out).all()
    np.savez_compressed(os.path.join(HERE, "ref_pbr_vectors.npz"), mats=mats, n_lights=nl, lights=lights, frag=frag, normal=normal, view=vpos,
                        exposure=exposure, dbg_in=dbg_in, dbg_eq=dbg_eq, transform=transform, smap=smap, out=out)
    print("wrote ref_pbr_vectors.npz", out[:3])


if __name__ == "__main__":
    main()
