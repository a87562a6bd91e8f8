"""Procedural scenes and host-side record packing for the raster hot path.

Everything here is host logic in numpy float32: icosphere grids, the reference's GPU-visible record
layouts and the camera / shadow-transform matrices that are INPUTS to the
hot path.  No device code and no oracle code is used here.

Reference conventions followed:
  Vertex 32-B AoS                      Source/Runtime/Renderer/MeshAllocator.h:14-24
  MeshDrawInfo / draw sort             MeshAllocator.h:26-39, RenderScene.cpp:105-125
  MaterialHandle bit packing           Source/Runtime/Renderer/MaterialDefines.h:88-96
  ViewProperties (160 B)               Source/Runtime/Renderer/ViewProperties.h:27-35
  Camera look-at / kClip*perspective   Source/Runtime/Renderer/Camera.h:93-117, glm_includes.h:18-23
  ComputeShadowTransform               Source/Runtime/Renderer/ShadowSystem.cpp:129-204 (F8: up = cross(dir, right))
  Phong Light (std430, 112 B)          Assets/Shaders/phong.glsl:20-37
"""
import os

import numpy as np

f32 = np.float32

LIGHT_DIRECTIONAL, LIGHT_POINT, LIGHT_SPOT = 1, 2, 3
INVALID_HANDLE = 0xFFFFFFFF
MATERIAL_STRIDE = 48
LIGHT_STRIDE = 112
VIEW_BYTES = 160

K_CLIP = np.array([[1, 0, 0, 0], [0, -1, 0, 0], [0, 0, 0.5, 0.5], [0, 0, 0, 1]], f32)   # row-major view of glm_vk::kClip


# ---------------------------------------------------------------------------------------------
# small float32 linear algebra, glm 1.0.2's operations in glm's order (matrices stored column-major as flat[16]; inside this
# module a matrix is a 4x4 numpy array indexed [row, col]).  tests/test_golden_host.py checks every function below bit for bit
# against vectors generated from the reference's own Camera.cpp / BoundingBox.cpp / glm (tests/golden/gen_golden*.py), so the
# scenes the tests and the bench render use exactly the matrices the reference would compute for the same camera and light.
# ---------------------------------------------------------------------------------------------
def _tanf(x):
    """glm::tan -> std::tan(float): the C library's tanf (numpy's float32 tan is a different implementation)."""
    global _LIBM
    try:
        if _LIBM is None:
            import ctypes
            _LIBM = ctypes.CDLL("libm.so.6")
            _LIBM.tanf.restype = ctypes.c_float; _LIBM.tanf.argtypes = [ctypes.c_float]
        return f32(_LIBM.tanf(float(x)))
    except OSError:
        return np.tan(f32(x), dtype=f32)


_LIBM = None


def _dot3(a, b):
    """glm compute_dot<vec3> (detail/func_geometric.inl:48-55): (x + y) + z."""
    a = np.asarray(a, f32); b = np.asarray(b, f32)
    return f32(f32(f32(a[0] * b[0]) + f32(a[1] * b[1])) + f32(a[2] * b[2]))


def _cross(x, y):
    """glm compute_cross (detail/func_geometric.inl:73-83)."""
    x = np.asarray(x, f32); y = np.asarray(y, f32)
    return np.array([x[1] * y[2] - y[1] * x[2], x[2] * y[0] - y[2] * x[0], x[0] * y[1] - y[0] * x[1]], f32)


def _norm(v):
    """glm normalize: v * inversesqrt(dot(v, v)), inversesqrt = 1 / sqrt (func_geometric.inl:98-105, func_exponential.inl:134-139)."""
    v = np.asarray(v, f32)
    return (v * f32(f32(1) / np.sqrt(_dot3(v, v), dtype=f32))).astype(f32)


def to_flat(m_rowmajor):
    """4x4 numpy (row, col) -> column-major flat[16] as glm stores it."""
    return np.ascontiguousarray(np.asarray(m_rowmajor, f32).T).reshape(16)


def from_flat(flat):
    return np.asarray(flat, f32).reshape(4, 4).T.copy()


def mat4_mul(a, b):
    """glm mat4 * mat4 (detail/type_mat4x4.inl:660-697): column c = ((A0 b.x + A1 b.y) + A2 b.z) + A3 b.w."""
    a = np.asarray(a, f32); b = np.asarray(b, f32)
    out = np.empty((4, 4), f32)
    for c in range(4):
        col = a[:, 0] * b[0, c]
        col = col + a[:, 1] * b[1, c]
        col = col + a[:, 2] * b[2, c]
        col = col + a[:, 3] * b[3, c]
        out[:, c] = col
    return out


def mat4_vec4(m, v):
    """glm mat4 * vec4 (detail/type_mat4x4.inl:537-583): (m0 x + m1 y) + (m2 z + m3 w)."""
    m = np.asarray(m, f32); v = np.asarray(v, f32)
    return ((m[:, 0] * v[0] + m[:, 1] * v[1]) + (m[:, 2] * v[2] + m[:, 3] * v[3])).astype(f32)


def mat4_inverse(a):
    """glm::inverse(mat4) (detail/func_matrix.inl:388-446), all in fp32."""
    a = np.asarray(a, f32)

    def M(c, r):
        return a[r, c]
    c00 = M(2, 2) * M(3, 3) - M(3, 2) * M(2, 3); c02 = M(1, 2) * M(3, 3) - M(3, 2) * M(1, 3); c03 = M(1, 2) * M(2, 3) - M(2, 2) * M(1, 3)
    c04 = M(2, 1) * M(3, 3) - M(3, 1) * M(2, 3); c06 = M(1, 1) * M(3, 3) - M(3, 1) * M(1, 3); c07 = M(1, 1) * M(2, 3) - M(2, 1) * M(1, 3)
    c08 = M(2, 1) * M(3, 2) - M(3, 1) * M(2, 2); c10 = M(1, 1) * M(3, 2) - M(3, 1) * M(1, 2); c11 = M(1, 1) * M(2, 2) - M(2, 1) * M(1, 2)
    c12 = M(2, 0) * M(3, 3) - M(3, 0) * M(2, 3); c14 = M(1, 0) * M(3, 3) - M(3, 0) * M(1, 3); c15 = M(1, 0) * M(2, 3) - M(2, 0) * M(1, 3)
    c16 = M(2, 0) * M(3, 2) - M(3, 0) * M(2, 2); c18 = M(1, 0) * M(3, 2) - M(3, 0) * M(1, 2); c19 = M(1, 0) * M(2, 2) - M(2, 0) * M(1, 2)
    c20 = M(2, 0) * M(3, 1) - M(3, 0) * M(2, 1); c22 = M(1, 0) * M(3, 1) - M(3, 0) * M(1, 1); c23 = M(1, 0) * M(2, 1) - M(2, 0) * M(1, 1)
    f0 = np.array([c00, c00, c02, c03], f32); f1 = np.array([c04, c04, c06, c07], f32); f2 = np.array([c08, c08, c10, c11], f32)
    f3 = np.array([c12, c12, c14, c15], f32); f4 = np.array([c16, c16, c18, c19], f32); f5 = np.array([c20, c20, c22, c23], f32)
    v0 = np.array([M(1, 0), M(0, 0), M(0, 0), M(0, 0)], f32); v1 = np.array([M(1, 1), M(0, 1), M(0, 1), M(0, 1)], f32)
    v2 = np.array([M(1, 2), M(0, 2), M(0, 2), M(0, 2)], f32); v3 = np.array([M(1, 3), M(0, 3), M(0, 3), M(0, 3)], f32)
    sa = np.array([1, -1, 1, -1], f32); sb = np.array([-1, 1, -1, 1], f32)
    inv = np.empty((4, 4), f32)                        # [row, col]
    inv[:, 0] = ((v1 * f0 - v2 * f1) + v3 * f2) * sa
    inv[:, 1] = ((v0 * f0 - v2 * f3) + v3 * f4) * sb
    inv[:, 2] = ((v0 * f1 - v1 * f3) + v3 * f5) * sa
    inv[:, 3] = ((v0 * f2 - v1 * f4) + v2 * f5) * sb
    d = a[:, 0] * inv[0, :]                              # m[0] * (first row of the adjugate)
    ood = f32(1) / f32(f32(d[0] + d[1]) + f32(d[2] + d[3]))
    return (inv * ood).astype(f32)


def look_at(eye, center, up):
    """glm::lookAtRH (ThirdParty/glm/glm/ext/matrix_transform.inl:153-174)."""
    eye, center, up = (np.asarray(a, f32) for a in (eye, center, up))
    f = _norm(center - eye)
    s = _norm(_cross(f, up))
    u = _cross(s, f)
    m = np.eye(4, dtype=f32)
    m[0, :3] = s; m[1, :3] = u; m[2, :3] = -f
    m[0, 3] = -_dot3(s, eye); m[1, 3] = -_dot3(u, eye); m[2, 3] = _dot3(f, eye)
    return m


def perspective_clip(fov_deg, aspect, near, far):
    """kClip * glm::perspectiveRH_ZO(radians(fov), aspect, near, far) (Camera.h:110-117; matrix_clip_space.inl:233-246)."""
    near, far, aspect = f32(near), f32(far), f32(aspect)
    t = _tanf(f32(f32(fov_deg) * f32(0.01745329251994329576923690768489)) / f32(2))
    m = np.zeros((4, 4), f32)
    m[0, 0] = f32(1) / f32(aspect * t)
    m[1, 1] = f32(1) / t
    m[2, 2] = far / f32(near - far)
    m[3, 2] = -1
    m[2, 3] = -f32(far * near) / f32(far - near)
    return mat4_mul(K_CLIP, m)


def ortho_clip(l, r, b, t, n, f):
    """kClip * glm::orthoRH_ZO (ShadowSystem.cpp:192-203; matrix_clip_space.inl:42-52)."""
    l, r, b, t, n, f = (f32(a) for a in (l, r, b, t, n, f))
    m = np.eye(4, dtype=f32)
    m[0, 0] = f32(2) / (r - l); m[1, 1] = f32(2) / (t - b); m[2, 2] = -f32(1) / (f - n)
    m[0, 3] = -(r + l) / (r - l); m[1, 3] = -(t + b) / (t - b); m[2, 3] = -n / (f - n)
    return mat4_mul(K_CLIP, m)


def camera_view(eye, look, up):
    """Camera::UpdateViewMatrix (Camera.h:93-108): re-orthogonalised up, then lookAt."""
    eye, look, up = (np.asarray(a, f32) for a in (eye, look, up))
    d = _norm(eye - look)
    r = _norm(_cross(up, d))
    u = _cross(d, r)
    return look_at(eye, look, u)


def bbox_corners(mn, mx):
    """BoundingBox::GetCorners (BoundingBox.h:33-45): x outermost, z innermost."""
    return np.array([[x, y, z] for x in (mn[0], mx[0]) for y in (mn[1], mx[1]) for z in (mn[2], mx[2])], f32)


def _vmin(a, b):
    return np.where(b < a, b, a).astype(f32)   # glm::min(x, y) = (y < x) ? y : x


def _vmax(a, b):
    return np.where(a < b, b, a).astype(f32)   # glm::max(x, y) = (x < y) ? y : x


_FLT_MAX = np.finfo(f32).max


def bbox_from_points(pts):
    """BoundingBox::FromPoints (BoundingBox.cpp:18-29)."""
    mn = np.full(3, _FLT_MAX, f32); mx = -mn
    for p in np.asarray(pts, f32):
        mn = _vmin(mn, p); mx = _vmax(mx, p)
    return mn, mx


def bbox_transform(mn, mx, m):
    """BoundingBox::Transform / operator* (BoundingBox.h:47-60, BoundingBox.cpp:3-16): 8 corners, divide by w, min / max."""
    omn = np.full(3, _FLT_MAX, f32); omx = -omn
    for c in bbox_corners(np.asarray(mn, f32), np.asarray(mx, f32)):
        p4 = mat4_vec4(m, np.array([c[0], c[1], c[2], 1], f32))
        p = (p4[:3] / p4[3]).astype(f32)
        omn = _vmin(omn, p); omx = _vmax(omx, p)
    return omn, omx


def frustum_corners(view, proj):
    """Camera::ComputeFrustrumCorners (Camera.cpp:3-25): inverse(view) * inverse(proj) applied to the NDC cube, divided by w."""
    t = mat4_mul(mat4_inverse(view), mat4_inverse(proj))
    out = []
    for x in (-1, 1):
        for y in (-1, 1):
            for z in (-1, 1):
                w = mat4_vec4(t, np.array([x, y, z, 1], f32))
                out.append((w[:3] / w[3]).astype(f32))
    return np.array(out, f32)


def light_view_matrix(direction):
    """ComputeDirectionalLightViewMatrix (ShadowSystem.cpp:129-147) with up = cross(dir, right) (F8)."""
    d = np.asarray(direction, f32)
    right = np.array([0, 0, 1], f32)
    if abs(_dot3(d, right)) > f32(0.9999):
        right = np.array([1, 0, 0], f32)
    up = _cross(d, right)
    return look_at(np.zeros(3, f32), d, up)


def compute_shadow_transform(light_dir, cam_view, cam_proj, scene_box, boxes, transforms):
    """ComputeShadowTransform (ShadowSystem.cpp:148-204). Returns (view, proj) as [row, col] numpy 4x4; proj = kClip * ortho."""
    cam_min, cam_max = bbox_from_points(frustum_corners(cam_view, cam_proj))
    sview = light_view_matrix(light_dir)
    cv_min, cv_max = bbox_transform(cam_min, cam_max, sview)
    sv_min, sv_max = bbox_transform(scene_box[0], scene_box[1], sview)
    cv_max[2] = sv_max[2]
    cw_min, cw_max = bbox_transform(cv_min, cv_max, mat4_inverse(sview))
    lb_min = np.full(3, _FLT_MAX, f32); lb_max = -lb_min
    for (bmn, bmx), t in zip(boxes, transforms):
        wmn, wmx = bbox_transform(bmn, bmx, from_flat(t))
        if np.all(wmn <= cw_max) and np.all(wmx >= cw_min):                 # BoundingBox::Intersects (BoundingBox.h:25-31)
            lb_min = _vmin(lb_min, wmn); lb_max = _vmax(lb_max, wmx)    # Union
    lv_min, lv_max = bbox_transform(lb_min, lb_max, sview)
    sproj = ortho_clip(lv_min[0], lv_max[0], lv_min[1], lv_max[1], lv_min[2], lv_max[2])
    return sview, sproj


# ---------------------------------------------------------------------------------------------
# record packing
# ---------------------------------------------------------------------------------------------
def material_handle(index, domain=0, model=1):
    """MaterialHandle raw bits: domain[0:2] model[2:4] index[4:32]; (Surface, Lit, 5) = 0x54."""
    return (domain & 3) | ((model & 3) << 2) | (index << 4)


def pack_view(view, proj, pos, exposure=1.0):
    rec = np.zeros(VIEW_BYTES, np.uint8)
    f = rec.view(f32)
    f[0:16] = to_flat(view); f[16:32] = to_flat(proj); f[32:35] = np.asarray(pos, f32); f[35] = exposure
    return rec


def pack_light(ltype, position, direction, ambient, diffuse, specular, inner=1.0, outer=0.0, shadow_index=0):
    rec = np.zeros(LIGHT_STRIDE, np.uint8)
    f = rec.view(f32)
    rec.view(np.int32)[0] = ltype
    f[4:7] = position; f[8:11] = direction
    f[12:16] = ambient; f[16:20] = diffuse; f[20:24] = specular
    f[24] = inner; f[25] = outer
    rec.view(np.uint32)[26] = shadow_index
    return rec


def pack_material(diffuse, specular, shininess, diffuse_tex=INVALID_HANDLE, specular_tex=INVALID_HANDLE):
    """48-B Phong record (builder-defined from phong.glsl:2-12)."""
    rec = np.zeros(MATERIAL_STRIDE, np.uint8)
    f = rec.view(f32); u = rec.view(np.uint32)
    f[0:4] = diffuse; f[4:8] = specular; f[8] = shininess
    u[9] = diffuse_tex; u[10] = specular_tex
    return rec


PBR_MATERIAL_STRIDE = 80
PBR_LIGHT_STRIDE = 80


def pack_pbr_light(ltype, color, position, direction, intensity, falloff_radius=10.0, cos_inner=1.0, cos_outer=0.0, shadow_index=0):
    """80-B std430 Light of the reference's live shader (LightSystem.h:22-33 / pbr.glsl:288-299)."""
    rec = np.zeros(PBR_LIGHT_STRIDE, np.uint8)
    f = rec.view(f32); u = rec.view(np.uint32)
    f[0:3] = color; f[4:7] = position; f[8:11] = direction
    f[11] = intensity; f[12] = falloff_radius; f[13] = cos_inner; f[14] = cos_outer
    u[15] = shadow_index; u[16] = ltype
    return rec


def pack_pbr_material(base_color, emissive, f0, metallic, roughness, occlusion, textures=(INVALID_HANDLE,) * 5):
    """80-B MaterialProperties (MaterialSystem.h:41-51 <-> Material, pbr.glsl:78-94); textures = baseColor, emissive, ORM, normals, AO."""
    rec = np.zeros(PBR_MATERIAL_STRIDE, np.uint8)
    f = rec.view(f32); u = rec.view(np.uint32)
    f[0:4] = base_color; f[4:8] = emissive; f[8] = f0; f[9] = metallic; f[10] = roughness; f[11] = occlusion
    u[12:17] = np.asarray(textures, np.uint32)
    return rec


def make_pbr_textures(n_sets, size, seed=1234):
    """n_sets x 4 procedural RGBA16 maps per set: [4k] base colour (sRGB), [4k+1] emissive (sRGB), [4k+2] occlusion /
    roughness / metallic (linear), [4k+3] tangent-space normals (linear, smooth bumps)."""
    rng = np.random.default_rng(seed + 991)
    yy, xx = np.mgrid[0:size, 0:size]
    out = []
    for k in range(n_sets):
        cell = 2 << (k % 3)
        checker = (((xx // cell) + (yy // cell)) & 1).astype(np.float64)[..., None]
        base = np.clip(checker * rng.uniform(0.3, 1.0, 3) + (1 - checker) * rng.uniform(0.05, 0.6, 3) + rng.uniform(-0.05, 0.05, (size, size, 1)), 0, 1)
        emis = np.clip((1 - checker) * rng.uniform(0.0, 0.4, 3), 0, 1)
        orm = np.clip(np.stack([0.7 + 0.3 * checker[..., 0], 0.25 + 0.6 * ((xx + yy) % 16) / 15.0, checker[..., 0] * (k % 2)], axis=2), 0, 1)
        fx, fy = 2 * np.pi * (1 + k % 3) / size, 2 * np.pi * (1 + (k + 1) % 3) / size
        nx, ny = 0.35 * np.cos(fx * xx), 0.35 * np.sin(fy * yy)
        nz = np.sqrt(np.maximum(1 - nx * nx - ny * ny, 0.1))
        nrm = np.stack([nx, ny, nz], axis=2) * 0.5 + 0.5
        one = np.ones((size, size, 1))
        for tex, srgb in ((base, 1), (emis, 1), (orm, 0), (nrm, 0)):
            out.append(dict(texels=np.round(np.concatenate([tex, one], axis=2) * 65535.0).astype(np.uint16), srgb=srgb))
    return out


def to_pbr(scene, seed=1234, n_texture_sets=0, tex_size=64, extra_lights=True, exposure=1.0, debug_input=0, debug_equation=0):
    """Re-dress a scene dict for the reference's live PBR fragment path: 80-B materials / lights, optional texture sets
    (every second material fully textured, incl. normal map), a point and a spot light next to the shadowed directional one."""
    rng = np.random.default_rng(seed + 5)
    s = dict(scene)
    n_mat = scene["materials"].shape[0]
    s["textures"] = make_pbr_textures(n_texture_sets, tex_size, seed) if n_texture_sets else []
    mats = []
    for k in range(n_mat):
        tex = [INVALID_HANDLE] * 5
        if n_texture_sets and k % 2 == 0:
            b = 4 * ((k // 2) % n_texture_sets)
            tex = [b, b + 1 if k % 4 == 0 else INVALID_HANDLE, b + 2, b + 3 if k % 3 != 1 else INVALID_HANDLE, INVALID_HANDLE]
        mats.append(pack_pbr_material(np.r_[rng.uniform(0.2, 1.0, 3), 1.0].astype(f32), np.r_[rng.uniform(0.0, 0.2, 3), f32(k % 3 == 0)].astype(f32),
                                      0.04, f32((k % 4 == 1) * 1.0 + (k % 4 == 2) * 0.5), f32(rng.uniform(0.15, 0.9)), f32(rng.uniform(0.6, 1.0)), tex))
    s["materials"] = np.stack(mats)
    ldir = _norm([-0.3, -1.0, -0.2])
    ex, ez = scene["extent"]
    lights = [pack_pbr_light(LIGHT_DIRECTIONAL, [1.0, 0.97, 0.9], (-ldir * f32(50)).astype(f32), ldir, 2.5)]
    if extra_lights:
        lights.append(pack_pbr_light(LIGHT_POINT, [0.3, 0.5, 1.0], [f32(-0.4 * ex), f32(2.5), f32(0.3 * ez)], [0, -1, 0], 40.0, falloff_radius=25.0))
        lights.append(pack_pbr_light(LIGHT_SPOT, [1.0, 0.4, 0.3], [f32(0.5 * ex), f32(3.0), f32(-0.2 * ez)], _norm([-0.2, -1.0, 0.1]), 60.0,
                                     falloff_radius=30.0, cos_inner=0.92, cos_outer=0.75))
    s["lights"] = np.stack(lights)
    v = np.array(scene["view"], copy=True)
    v.view(f32)[35] = exposure; v.view(np.uint32)[36] = debug_input; v.view(np.uint32)[37] = debug_equation
    s["view"] = v
    s["shading_model"] = 1
    return s


# ---------------------------------------------------------------------------------------------
# geometry
# ---------------------------------------------------------------------------------------------
_ICO_CACHE = {}


def icosphere(level, order="breadth"):
    """Unit icosphere: 10*4^L + 2 shared vertices, 20*4^L CCW-outward triangles.  order = "breadth" (default, the headline
    scenes): every subdivision emits all first-corner children, then all second-corner ones, ... — consecutive triangles
    are scattered over the sphere; "hierarchical": the four children of a triangle stay together, so any run of triangles
    is a compact patch (what exporters that write faces in construction order produce).  Same triangles, same vertices."""
    if (level, order) in _ICO_CACHE:
        return _ICO_CACHE[(level, order)]
    t = (1.0 + 5.0 ** 0.5) / 2.0
    v = np.array([[-1, t, 0], [1, t, 0], [-1, -t, 0], [1, -t, 0], [0, -1, t], [0, 1, t], [0, -1, -t], [0, 1, -t],
                  [t, 0, -1], [t, 0, 1], [-t, 0, -1], [-t, 0, 1]], np.float64)
    v /= np.linalg.norm(v, axis=1, keepdims=True)
    tri = np.array([[0, 11, 5], [0, 5, 1], [0, 1, 7], [0, 7, 10], [0, 10, 11], [1, 5, 9], [5, 11, 4], [11, 10, 2], [10, 7, 6],
                    [7, 1, 8], [3, 9, 4], [3, 4, 2], [3, 2, 6], [3, 6, 8], [3, 8, 9], [4, 9, 5], [2, 4, 11], [6, 2, 10],
                    [8, 6, 7], [9, 8, 1]], np.int64)
    for _ in range(level):
        e = np.concatenate([tri[:, [0, 1]], tri[:, [1, 2]], tri[:, [2, 0]]])
        es = np.sort(e, axis=1)
        key = es[:, 0] * (len(v) + 1) + es[:, 1]
        uniq, inv = np.unique(key, return_inverse=True)
        first = np.zeros(len(uniq), np.int64); first[inv[::-1]] = np.arange(len(key))[::-1]
        mid = v[es[first, 0]] + v[es[first, 1]]
        mid /= np.linalg.norm(mid, axis=1, keepdims=True)
        base = len(v)
        v = np.concatenate([v, mid])
        n = len(tri)
        m01, m12, m20 = base + inv[:n], base + inv[n:2 * n], base + inv[2 * n:]
        a, b, c = tri[:, 0], tri[:, 1], tri[:, 2]
        kids = [np.stack([a, m01, m20], 1), np.stack([b, m12, m01], 1), np.stack([c, m20, m12], 1), np.stack([m01, m12, m20], 1)]
        tri = np.stack(kids, 1).reshape(-1, 3) if order == "hierarchical" else np.concatenate(kids)
    pos = v.astype(f32)
    uv = np.stack([np.arctan2(v[:, 2], v[:, 0]) / (2 * np.pi) + 0.5, np.arcsin(np.clip(v[:, 1], -1, 1)) / np.pi + 0.5], 1).astype(f32)
    verts = np.concatenate([pos, pos, uv], axis=1).astype(f32)   # normal = position on the unit sphere
    _ICO_CACHE[(level, order)] = (verts, tri.astype(np.uint32))
    return _ICO_CACHE[(level, order)]


CONFIGS = {
    # name: grid (nx, nz), level, target, shadow size, materials   
    "C1": dict(grid=(10, 5), level=3, width=1280, height=720, shadow_size=0, n_materials=4),
    "C2": dict(grid=(7, 7), level=5, width=1920, height=1080, shadow_size=2048, n_materials=8),
    "C3": dict(grid=(16, 8), level=6, width=3840, height=2160, shadow_size=2048, n_materials=1024, n_textures=4096, tex_size=64),
    "C4": dict(grid=(16, 8), level=6, width=7680, height=4320, shadow_size=2048, n_materials=1024, n_textures=4096, tex_size=64),
    "C5": dict(grid=(14, 7), level=5, width=1920, height=1080, shadow_size=2048, n_materials=8),
    "tiny": dict(grid=(3, 2), level=2, width=160, height=96, shadow_size=128, n_materials=3),
    "small": dict(grid=(4, 3), level=3, width=640, height=360, shadow_size=512, n_materials=5),
    "small_tex": dict(grid=(4, 3), level=3, width=640, height=360, shadow_size=512, n_materials=5, n_textures=7, tex_size=32),
}


def make_textures(n, size, seed=1234):
    """n procedural R16G16B16A16_UNORM textures (checker / noise, 64x64 by default): even handles are
    sRGB-encoded colour maps, odd handles linear (specular) maps."""
    rng = np.random.default_rng(seed + 77)
    yy, xx = np.mgrid[0:size, 0:size]
    out = []
    for k in range(n):
        cell = 2 << (k % 3)
        checker = (((xx // cell) + (yy // cell)) & 1).astype(np.float64)
        base = rng.uniform(0.25, 1.0, 3); alt = rng.uniform(0.05, 0.6, 3)
        noise = rng.uniform(-0.08, 0.08, (size, size, 1))
        rgb = np.clip(checker[..., None] * base + (1 - checker[..., None]) * alt + noise, 0.0, 1.0)
        tex = np.concatenate([rgb, np.ones((size, size, 1))], axis=2)
        out.append(dict(texels=np.round(tex * 65535.0).astype(np.uint16), srgb=1 if k % 2 == 0 else 0))
    return out


def camera_for(extent_x, extent_z, width, height, azimuth=0.0, fov=45.0, near=0.01, far=100.0):
    """Eye on the (1,1,1) diagonal (rotated by `azimuth` rad about +Y) at a distance framing the grid."""
    radius = float(np.hypot(extent_x, extent_z))
    aspect = width / float(height)
    half_v = np.radians(fov) / 2
    half_h = np.arctan(np.tan(half_v) * aspect)
    dist = min(0.85 * radius / np.sin(min(half_v, half_h)), far - radius - 1.0)
    d = np.array([1.0, 1.0, 1.0]) / np.sqrt(3.0)
    ca, sa = np.cos(azimuth), np.sin(azimuth)
    d = np.array([ca * d[0] + sa * d[2], d[1], -sa * d[0] + ca * d[2]])
    eye = (d * dist).astype(f32)
    view = camera_view(eye, np.zeros(3, f32), np.array([0, 1, 0], f32))
    proj = perspective_clip(fov, f32(width) / f32(height), near, far)
    return eye, view, proj


def make_scene(config="C2", seed=1234, ground=True, azimuth=0.0, **overrides):
    """Build a scene dict for a named config. All arrays are host numpy."""
    cfg = dict(CONFIGS[config]) if isinstance(config, str) else dict(config)
    cfg.update(overrides)
    if os.environ.get("REB_SHADOW_SIZE") and cfg["shadow_size"]:   # e.g. 4096, the reference's default map (ShadowSystem.h)
        cfg["shadow_size"] = int(os.environ["REB_SHADOW_SIZE"])
    if os.environ.get("REB_SCENE_GRID"):    # "100x100": many small draws (one draw per sphere)
        cfg["grid"] = tuple(int(v) for v in os.environ["REB_SCENE_GRID"].split("x"))
    if os.environ.get("REB_SCENE_LEVEL"):   # triangle-size sweeps of a named config (tools/kt.py)
        cfg["level"] = int(os.environ["REB_SCENE_LEVEL"])
    nx, nz = cfg["grid"]; level = cfg["level"]
    W, H, S, M = cfg["width"], cfg["height"], cfg["shadow_size"], cfg["n_materials"]
    n_tex, tex_size = cfg.get("n_textures", 0), cfg.get("tex_size", 64)
    rng = np.random.default_rng(seed)
    sv, st = icosphere(level, cfg.get("tri_order", os.environ.get("REB_SCENE_TRI_ORDER", "breadth")))
    n_s = nx * nz
    nv, nt = len(sv), len(st)
    spacing = 2.5
    verts = np.tile(sv, (n_s, 1))
    indices = (st.reshape(1, -1).astype(np.uint64) + (np.arange(n_s, dtype=np.uint64) * nv).reshape(-1, 1)).astype(np.uint32)
    transforms = np.tile(np.eye(4, dtype=f32).reshape(1, 16), (n_s, 1))
    gx = (np.arange(nx) - (nx - 1) / 2.0) * spacing
    gz = (np.arange(nz) - (nz - 1) / 2.0) * spacing
    cx, cz = np.meshgrid(gx, gz, indexing="ij")
    transforms[:, 12] = cx.reshape(-1); transforms[:, 14] = cz.reshape(-1)
    boxes = [(np.full(3, -1, f32), np.full(3, 1, f32))] * n_s
    draws = [(k * nt * 3, nt * 3, material_handle(k % M), k) for k in range(n_s)]
    ex, ez = abs(gx[0]) + 1.0, abs(gz[0]) + 1.0
    n_mat = M
    if ground:
        gy = f32(-1.05)
        hx, hz = f32(ex + 1.5), f32(ez + 1.5)
        q = np.array([[-hx, gy, -hz, 0, 1, 0, 0, 0], [-hx, gy, hz, 0, 1, 0, 0, 1], [hx, gy, hz, 0, 1, 0, 1, 1], [hx, gy, -hz, 0, 1, 0, 1, 0]], f32)
        base = len(verts)
        verts = np.concatenate([verts, q])
        qi = np.array([0, 1, 2, 0, 2, 3], np.uint32) + np.uint32(base)
        draws.append((indices.size, 6, material_handle(M), n_s))
        indices = np.concatenate([indices.reshape(-1), qi])
        transforms = np.concatenate([transforms, np.eye(4, dtype=f32).reshape(1, 16)])
        boxes = boxes + [(np.array([-hx, gy, -hz], f32), np.array([hx, gy, hz], f32))]
        n_mat = M + 1
    indices = indices.reshape(-1)
    # RenderScene::SortOpaqueMeshes: by (materialHandle raw, sceneNodeID)
    draws = np.array(sorted(draws, key=lambda d: (d[2], d[3])), np.uint32)
    shin = np.array([8, 16, 32, 64], f32)
    def tex_of(k, which):   # material k samples texture 2k (sRGB colour) and, every other material, 2k+1 (linear specular)
        if not n_tex or (which == 1 and k % 2):
            return INVALID_HANDLE
        return (2 * k + which) % n_tex
    materials = np.stack([pack_material(np.r_[rng.uniform(0.2, 1.0, 3), 1.0].astype(f32), np.array([0.5, 0.5, 0.5, 1], f32),
                                        shin[k % 4], tex_of(k, 0), tex_of(k, 1)) for k in range(n_mat)])
    if ground:
        materials[M] = pack_material(np.array([0.8, 0.8, 0.8, 1], f32), np.array([0.2, 0.2, 0.2, 1], f32), 16.0, tex_of(M, 0))
    ldir = _norm([-0.3, -1.0, -0.2])
    lights = np.stack([pack_light(LIGHT_DIRECTIONAL, (-ldir * f32(50)).astype(f32), ldir, [0.1] * 4, [1.0] * 4, [1.0] * 4)])
    eye, view, proj = camera_for(ex, ez, W, H, azimuth)
    scene = dict(config=config if isinstance(config, str) else "custom", verts=verts, indices=indices, transforms=transforms,
                 draws=draws, materials=materials, lights=lights, view=pack_view(view, proj, eye), width=W, height=H,
                 shadow_size=S, max_descriptors=8192, n_triangles=int(indices.size // 3), n_vertices=int(len(verts)),
                 eye=eye, view_mat=view, proj_mat=proj, boxes=boxes, extent=(ex, ez), textures=make_textures(n_tex, tex_size, seed) if n_tex else [])
    if S:
        all_min = np.array([-ex - 1.5, -1.05, -ez - 1.5], f32) if ground else np.array([-ex, -1, -ez], f32)
        all_max = np.array([ex + 1.5, 1, ez + 1.5], f32) if ground else np.array([ex, 1, ez], f32)
        sview, sproj = compute_shadow_transform(ldir, view, proj, (all_min, all_max), boxes, transforms)
        scene["shadow_view"] = to_flat(sview); scene["shadow_proj"] = to_flat(sproj)
    return scene


def view_for_azimuth(scene, azimuth):
    """A new 160-B ViewProperties record orbiting the same scene (view-parallel batches, C5)."""
    ex, ez = scene["extent"]
    eye, view, proj = camera_for(ex, ez, scene["width"], scene["height"], azimuth)
    return pack_view(view, proj, eye)


def view_from(scene, eye, look, fov=45.0, near=0.01, far=100.0):
    """A 160-B ViewProperties record for an arbitrary camera (e.g. INSIDE the scene, so geometry crosses the eye plane
    and the guard band and takes the clipping path, Appendix B.3)."""
    eye = np.asarray(eye, f32)
    view = camera_view(eye, np.asarray(look, f32), np.array([0, 1, 0], f32))
    proj = perspective_clip(fov, f32(scene["width"]) / f32(scene["height"]), near, far)
    return pack_view(view, proj, eye)


def to_translucent(scene, every=3, alpha=0.45, view=None, which=None):
    """Make every `every`-th material (or the materials listed in `which`) translucent (AlphaMode::eBlend, alpha in the material's diffuse / base-colour alpha)
    and split the draw list the way RenderScene does: opaque draws stay in `draws` (sorted by material, node:
    RenderScene.cpp:105-125), translucent ones go to `translucent_draws` sorted back to front by
    dot(viewInverse[2], cameraPosition - nodeTranslation), then material, then node (RenderScene.cpp:127-159)."""
    s = dict(scene)
    mats = np.array(scene["materials"], np.uint8).copy()
    n_mat = mats.shape[0]
    tl = np.zeros(n_mat, bool)
    tl[np.arange(0, n_mat, every) if which is None else np.asarray(which, np.int64)] = True
    for m in np.nonzero(tl)[0]:
        mats[m, 12:16] = np.frombuffer(f32(alpha).tobytes(), np.uint8)
    s["materials"] = mats
    draws = np.asarray(scene["draws"], np.uint32)
    is_t = tl[(draws[:, 2] >> 4).astype(np.int64)]
    vw = np.frombuffer(np.asarray(view if view is not None else scene["view"], np.uint8).tobytes(), f32)
    V = vw[:16].reshape(4, 4).T.astype(np.float64)
    Vi = np.linalg.inv(V)
    cam, front = Vi[:3, 3], Vi[:3, 2]
    T = np.asarray(scene["transforms"], f32).reshape(-1, 16)
    def key(d):
        dist = float(np.dot(front, cam - T[int(d[3]), 12:15].astype(np.float64)))
        return (-dist, int(d[2]), int(d[3]))
    s["draws"] = draws[~is_t]
    s["translucent_draws"] = np.array(sorted((tuple(int(v) for v in d) for d in draws[is_t]), key=key), np.uint32).reshape(-1, 4)
    s["n_triangles"] = int(scene["n_triangles"])
    return s


def make_skybox(size=64, seed=1234):
    """Procedural cube map in the reference's layout (TextureCache.cpp:199-290: six RGBA16 UNORM faces, one mip, layers
    +X -X +Y -Y +Z -Z = right, left, top, bottom, front, back of Skybox.cpp:71-78): a vertical sky gradient that is
    continuous across the faces, a sun disc, and a faint per-face checker so filtering and face seams are exercised."""
    j, i = np.meshgrid(np.arange(size, dtype=np.float64), np.arange(size, dtype=np.float64), indexing="ij")
    sc = 2.0 * (i + 0.5) / size - 1.0
    tc = 2.0 * (j + 0.5) / size - 1.0
    one = np.ones_like(sc)
    dirs = [(one, -tc, -sc), (-one, -tc, sc), (sc, one, tc), (sc, -one, -tc), (sc, -tc, one), (-sc, -tc, -one)]
    sun = _norm(np.array([0.3, 1.0, 0.2]))
    rng = np.random.default_rng(seed)
    tint = rng.uniform(0.9, 1.0, size=(6, 3))
    faces = np.empty((6, size, size, 4), np.uint16)
    for f, (x, y, z) in enumerate(dirs):
        n = np.sqrt(x * x + y * y + z * z)
        x, y, z = x / n, y / n, z / n
        up = 0.5 * y + 0.5
        col = np.stack([0.25 + 0.35 * (1 - up), 0.35 + 0.35 * (1 - up), 0.55 + 0.4 * up], -1)
        s_dot = np.clip(x * sun[0] + y * sun[1] + z * sun[2], 0.0, 1.0)
        col += (s_dot[..., None] ** 64) * np.array([1.0, 0.9, 0.6])
        chk = (((i.astype(np.int64) >> 2) + (j.astype(np.int64) >> 2)) & 1).astype(np.float64)
        col *= (0.92 + 0.08 * chk)[..., None] * tint[f]
        faces[f, ..., :3] = np.round(np.clip(col, 0.0, 1.0) * 65535.0).astype(np.uint16)
        faces[f, ..., 3] = 65535
    return faces


def algorithmic_bytes(scene, hdr=False):
    """Compulsory HBM traffic per frame: returns (B_shadow, B_forward)."""
    V, T = scene["n_vertices"], scene["n_triangles"]
    W, H, S = scene["width"], scene["height"], scene["shadow_size"]
    b_shadow = 12 * V + 12 * T + 4 * S * S if S else 0
    b_forward = 32 * V + 12 * T + (min(36 * W * H, 4 * S * S) if S else 0) + 12 * W * H + (16 * W * H if hdr else 0)
    return b_shadow, b_forward
