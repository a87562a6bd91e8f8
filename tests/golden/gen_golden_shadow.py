"""Golden vectors for the COMPOSED shadow transform (ShadowSystem.cpp:129-204), generated from the reference's own compiled
Camera.cpp / BoundingBox.cpp / glm through oracle/_ref/libref.so (ref_shadow_transform in oracle/ref_shim.cpp replays the
statements of ComputeShadowTransform over those pieces; ShadowSystem.cpp itself needs the Vulkan headers).  Run in the build
container only; output tests/golden/ref_shadow_vectors.npz is committed.

    make -C oracle ref && python tests/golden/gen_golden_shadow.py
"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
lib = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libref.so"))
fp = C.POINTER(C.c_float)


def P(a):
    return a.ctypes.data_as(fp)


def main():
    rng = np.random.default_rng(4321)
    K, NB = 24, 12
    lib.ref_shadow_transform.argtypes = [fp, fp, fp, fp, C.c_float, C.c_float, C.c_float, C.c_int, C.c_int, fp, fp, fp, fp, C.c_int, fp, fp, fp]
    ldir = rng.normal(size=(K, 3)); ldir /= np.linalg.norm(ldir, axis=1, keepdims=True)
    ldir[0] = [0, 0, -1]; ldir[1] = [0, -1, 0]; ldir[2] = [-0.3, -1.0, -0.2] / np.linalg.norm([-0.3, -1.0, -0.2])   # the |dot| > 0.9999 branch, straight down, the bench light
    ldir = ldir.astype(np.float32)
    eye = rng.uniform(-25, 25, (K, 3)).astype(np.float32); eye[:, 1] = np.abs(eye[:, 1]) + 2
    look = rng.uniform(-2, 2, (K, 3)).astype(np.float32)
    up = np.tile(np.array([0, 1, 0], np.float32), (K, 1))
    dims = np.array([(1280, 720), (1920, 1080), (3840, 2160), (800, 600)], np.int32)[np.arange(K) % 4]
    fov = rng.choice([45.0, 60.0, 30.0], K).astype(np.float32)
    boxes = np.zeros((K, NB, 6), np.float32); tr = np.zeros((K, NB, 16), np.float32)
    smin = np.zeros((K, 3), np.float32); smax = np.zeros((K, 3), np.float32)
    view = np.zeros((K, 16), np.float32); proj = np.zeros((K, 16), np.float32); xf = np.zeros((K, 16), np.float32)
    for i in range(K):
        lo = rng.uniform(-1.5, -0.5, (NB, 3)); hi = lo + rng.uniform(0.5, 3.0, (NB, 3))
        boxes[i, :, :3] = lo; boxes[i, :, 3:] = hi
        T = np.tile(np.eye(4, dtype=np.float32).reshape(1, 16), (NB, 1))
        T[:, 12:15] = rng.uniform(-8, 8, (NB, 3)); T[:, [0, 5, 10]] = rng.uniform(0.5, 2.0, (NB, 3))
        T[NB // 2:, 4] = rng.uniform(-0.3, 0.3, NB - NB // 2)     # some shear
        tr[i] = T
        w = np.stack([T[:, 12:15] + T[:, [0, 5, 10]] * boxes[i, :, :3] - 1.0, T[:, 12:15] + T[:, [0, 5, 10]] * boxes[i, :, 3:] + 1.0])   # a loose scene box (input only)
        smin[i] = w.min(axis=(0, 1)); smax[i] = w.max(axis=(0, 1))
        lib.ref_shadow_transform(P(ldir[i]), P(eye[i]), P(look[i]), P(up[i]), float(fov[i]), 0.01, 100.0, int(dims[i, 0]), int(dims[i, 1]),
                                 P(smin[i]), P(smax[i]), P(boxes[i]), P(tr[i]), NB, P(view[i]), P(proj[i]), P(xf[i]))
    np.savez_compressed(os.path.join(HERE, "ref_shadow_vectors.npz"), light_dir=ldir, eye=eye, look=look, up=up, dims=dims, fov=fov, boxes=boxes, transforms=tr,
                        scene_min=smin, scene_max=smax, view=view, proj=proj, transform=xf)
    print("wrote ref_shadow_vectors.npz:", K, "cases; finite:", bool(np.isfinite(proj).all()))


if __name__ == "__main__":
    main()
