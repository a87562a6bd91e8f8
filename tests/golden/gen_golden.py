"""Generate golden vectors from the REFERENCE's own code (oracle/_ref/libref.so, built by
oracle/Makefile from /root/reference sources: glm, Camera.cpp, BoundingBox.cpp, phong.glsl,
shadow.glsl).  Run in the build container only; output tests/golden/ref_vectors.npz is committed.

    python tests/golden/gen_golden.py
"""
import ctypes as C
import os
import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
lib = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libref.so"))
fp = C.POINTER(C.c_float)
u8p = C.POINTER(C.c_uint8)


def P(a):
    return a.ctypes.data_as(fp)


def make_light(rng, ltype):
    rec = np.zeros(112, np.uint8)
    f = rec.view(np.float32)
    rec.view(np.int32)[0] = ltype
    f[4:7] = rng.uniform(-10, 10, 3)            # position @16
    d = rng.normal(size=3); d /= np.linalg.norm(d)
    f[8:11] = d                                 # direction @32
    f[12:16] = rng.uniform(0, 0.3, 4)           # ambient @48
    f[16:20] = rng.uniform(0.2, 1, 4)           # diffuse @64
    f[20:24] = rng.uniform(0.2, 1, 4)           # specular @80
    f[24] = 0.95; f[25] = 0.80                  # inner/outer cutoff
    rec.view(np.uint32)[26] = 0                 # shadowIndex @104
    return rec


def main():
    rng = np.random.default_rng(1234)
    out = {}
    N = 256
    # --- glm matrix arithmetic -------------------------------------------------------------
    A = rng.normal(size=(N, 16)).astype(np.float32)
    B = rng.normal(size=(N, 16)).astype(np.float32)
    V = rng.normal(size=(N, 4)).astype(np.float32)
    mm = np.zeros((N, 16), np.float32); mv = np.zeros((N, 4), np.float32); nm = np.zeros((N, 9), np.float32)
    for i in range(N):
        lib.ref_mat4_mul(P(A[i]), P(B[i]), P(mm[i]))
        lib.ref_mat4_mul_vec4(P(A[i]), P(V[i]), P(mv[i]))
        lib.ref_normal_matrix(P(A[i]), P(nm[i]))
    out.update(mat_a=A, mat_b=B, vec=V, mat_mul=mm, mat_vec=mv, normal_matrix=nm)
    # --- vertex stages ---------------------------------------------------------------------
    T = np.tile(np.eye(4, dtype=np.float32).reshape(1, 16), (N, 1))
    T[:, 12:15] = rng.uniform(-5, 5, (N, 3))
    T[:, [0, 5, 10]] = rng.uniform(0.5, 2, (N, 3))
    T[: N // 2, 1] = rng.uniform(-0.3, 0.3, N // 2)   # some shear
    view = rng.normal(size=(N, 16)).astype(np.float32)
    proj = rng.normal(size=(N, 16)).astype(np.float32)
    pos = rng.uniform(-1, 1, (N, 3)).astype(np.float32)
    nrm = rng.normal(size=(N, 3)).astype(np.float32)
    pv = np.zeros((N, 10), np.float32); sv = np.zeros((N, 4), np.float32)
    for i in range(N):
        lib.ref_primitive_vert(P(T[i]), P(view[i]), P(proj[i]), P(pos[i]), P(nrm[i]), P(pv[i]))
        lib.ref_shadow_vert(P(T[i]), P(view[i]), P(proj[i]), P(pos[i]), P(sv[i]))
    out.update(vs_T=T.astype(np.float32), vs_view=view, vs_proj=proj, vs_pos=pos, vs_nrm=nrm, vs_primitive=pv, vs_shadow=sv)
    # --- PhongLighting (phong.glsl:39-81) ----------------------------------------------------
    M = 512
    lights = np.stack([make_light(rng, 1 + (i % 3)) for i in range(M)])
    props = rng.uniform(0, 1, (M, 9)).astype(np.float32)
    props[:, 8] = rng.choice([1, 8, 16, 32, 64, 17.5], M)
    normal = rng.normal(size=(M, 3)).astype(np.float32); normal /= np.linalg.norm(normal, axis=1, keepdims=True)
    frag = rng.uniform(-4, 4, (M, 3)).astype(np.float32)
    vpos = rng.uniform(-8, 8, (M, 3)).astype(np.float32)
    shadow = (rng.integers(0, 10, M) / 9.0).astype(np.float32)
    ph = np.zeros((M, 4), np.float32)
    lib.ref_phong_lighting.argtypes = [u8p, fp, fp, fp, fp, C.c_float, fp]
    for i in range(M):
        lib.ref_phong_lighting(lights[i].ctypes.data_as(u8p), P(props[i]), P(normal[i]), P(frag[i]), P(vpos[i]),
                               C.c_float(float(shadow[i])), P(ph[i]))
    out.update(ph_lights=lights, ph_props=props, ph_normal=normal, ph_frag=frag, ph_view=vpos, ph_shadow=shadow, ph_out=ph)
    # --- ComputeShadow (shadow.glsl:24-56) ----------------------------------------------------
    S = 64
    smap = rng.uniform(0.3, 0.9, (S, S)).astype(np.float32)
    tr = np.zeros((M, 16), np.float32)
    tr[:, 0] = 0.25; tr[:, 5] = 0.25; tr[:, 10] = 0.1; tr[:, 15] = 1.0
    tr[:, 12:14] = rng.uniform(-0.2, 0.2, (M, 2)); tr[:, 14] = rng.uniform(0.3, 0.9, M)
    tr[::7, 14] = 1.2   # some beyond the far plane
    cs = np.zeros(M, np.float32)
    lib.ref_compute_shadow.restype = C.c_float
    lib.ref_compute_shadow.argtypes = [u8p, fp, fp, fp, fp, C.c_uint32]
    for i in range(M):
        cs[i] = lib.ref_compute_shadow(lights[i].ctypes.data_as(u8p), P(frag[i]), P(normal[i]), P(tr[i]), P(smap), S)
    out.update(cs_transform=tr, cs_map=smap, cs_out=cs)
    # --- Camera (Camera.h / Camera.cpp) --------------------------------------------------------
    K = 16
    eye = rng.uniform(-20, 20, (K, 3)).astype(np.float32)
    look = rng.uniform(-2, 2, (K, 3)).astype(np.float32)
    up = np.tile(np.array([0, 1, 0], np.float32), (K, 1))
    cam_view = np.zeros((K, 16), np.float32); cam_proj = np.zeros((K, 16), np.float32); cam_corners = np.zeros((K, 24), np.float32)
    dims = [(1280, 720), (1920, 1080), (3840, 2160), (800, 600)]
    lib.ref_camera.argtypes = [fp, fp, fp, C.c_float, C.c_float, C.c_float, C.c_int, C.c_int, fp, fp, fp]
    cam_dims = np.zeros((K, 2), np.int32)
    for i in range(K):
        w, h = dims[i % 4]; cam_dims[i] = (w, h)
        lib.ref_camera(P(eye[i]), P(look[i]), P(up[i]), 45.0, 0.01, 100.0, w, h, P(cam_view[i]), P(cam_proj[i]), P(cam_corners[i]))
    out.update(cam_eye=eye, cam_look=look, cam_up=up, cam_dims=cam_dims, cam_view=cam_view, cam_proj=cam_proj, cam_corners=cam_corners)
    # --- BoundingBox / lookAt / ortho / inverse (pieces of ShadowSystem.cpp:129-204) -----------
    bmin = rng.uniform(-5, 0, (K, 3)).astype(np.float32); bmax = (bmin + rng.uniform(0.5, 6, (K, 3))).astype(np.float32)
    bt = np.zeros((K, 2, 6), np.float32)
    lib.ref_bbox_transform.argtypes = [fp, fp, fp, C.c_int, fp, fp]
    for i in range(K):
        for op in (0, 1):
            mn = np.zeros(3, np.float32); mx = np.zeros(3, np.float32)
            lib.ref_bbox_transform(P(bmin[i]), P(bmax[i]), P(cam_view[i]), op, P(mn), P(mx))
            bt[i, op, :3] = mn; bt[i, op, 3:] = mx
    ldir = rng.normal(size=(K, 3)).astype(np.float32); ldir /= np.linalg.norm(ldir, axis=1, keepdims=True); ldir = ldir.astype(np.float32)
    lup = rng.normal(size=(K, 3)).astype(np.float32)
    la = np.zeros((K, 16), np.float32); orth = np.zeros((K, 16), np.float32); inv = np.zeros((K, 16), np.float32)
    zero = np.zeros(3, np.float32)
    lib.ref_clip_ortho.argtypes = [C.c_float] * 6 + [fp]
    for i in range(K):
        lib.ref_look_at(P(zero), P(ldir[i]), P(lup[i]), P(la[i]))
        lib.ref_clip_ortho(float(bmin[i, 0]), float(bmax[i, 0]), float(bmin[i, 1]), float(bmax[i, 1]), float(bmin[i, 2]), float(bmax[i, 2]), P(orth[i]))
        lib.ref_mat4_inverse(P(cam_view[i]), P(inv[i]))
    out.update(bb_min=bmin, bb_max=bmax, bb_out=bt, la_dir=ldir, la_up=lup, la_out=la, ortho_out=orth, inv_out=inv)
    np.savez_compressed(os.path.join(HERE, "ref_vectors.npz"), **out)
    print("wrote ref_vectors.npz with", len(out), "arrays")


if __name__ == "__main__":
    main()
