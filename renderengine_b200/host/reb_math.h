// reb_math.h — the small fp32 linear algebra the host facade needs (column-major 4x4, right-handed,
// depth zero-to-one).  The reference gets these through glm (GLM_FORCE_DEPTH_ZERO_TO_ONE,
// Source/Runtime/glm_includes.h:4; kClip glm_includes.h:18-23); a caller that switches to this facade must get
// the SAME matrices bit for bit, so every function here performs glm 1.0.2's fp32 operations in glm's order
// (non-SIMD paths; cited per function, paths relative to ThirdParty/glm/glm/).  Compile without contraction
// (-ffp-contract=off; g++ on x86-64 does not fuse by default).  tests/cpp/test_golden_math.cpp checks Camera,
// BoundingBox, lookAt, kClip * ortho, inverse and mat4 products against vectors generated from the reference's
// own Camera.cpp / BoundingBox.cpp / glm (tests/golden/gen_golden.py) bit for bit.
#pragma once
#include <cmath>
#include <cstring>

namespace reb {

struct vec2 { float x = 0, y = 0; };
struct vec3 {
    float x = 0, y = 0, z = 0;
    vec3() = default;
    vec3(float a, float b, float c) : x(a), y(b), z(c) {}
    explicit vec3(float a) : x(a), y(a), z(a) {}
};
struct vec4 {
    float x = 0, y = 0, z = 0, w = 0;
    vec4() = default;
    vec4(float a, float b, float c, float d) : x(a), y(b), z(c), w(d) {}
    vec4(vec3 v, float d) : x(v.x), y(v.y), z(v.z), w(d) {}
};

inline vec3 operator+(vec3 a, vec3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline vec3 operator-(vec3 a, vec3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline vec3 operator-(vec3 a) { return {-a.x, -a.y, -a.z}; }
inline vec3 operator*(vec3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
inline vec3 operator*(float s, vec3 a) { return a * s; }
inline vec3 operator/(vec3 a, float s) { return {a.x / s, a.y / s, a.z / s}; }
// detail/func_geometric.inl:48-55 (dot: (x + y) + z), :73-83 (cross), :98-105 (normalize = v * inversesqrt(dot(v, v)),
// inversesqrt = 1 / sqrt, detail/func_exponential.inl:134-139)
inline float dot(vec3 a, vec3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
inline vec3 cross(vec3 a, vec3 b) { return {a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y}; }
inline vec3 normalize(vec3 a) { return a * (1.0f / std::sqrt(dot(a, a))); }
// detail/func_common.inl: min(x, y) = (y < x) ? y : x; max(x, y) = (x < y) ? y : x
inline float fmin_glm(float x, float y) { return (y < x) ? y : x; }
inline float fmax_glm(float x, float y) { return (x < y) ? y : x; }
inline vec3 vmin(vec3 a, vec3 b) { return {fmin_glm(a.x, b.x), fmin_glm(a.y, b.y), fmin_glm(a.z, b.z)}; }
inline vec3 vmax(vec3 a, vec3 b) { return {fmax_glm(a.x, b.x), fmax_glm(a.y, b.y), fmax_glm(a.z, b.z)}; }

// column-major: m[c][r], flat index c*4 + r — the layout of `mat4 transforms[]` (primitive.vert:29-31)
struct mat4 {
    float m[16];
    mat4() { std::memset(m, 0, sizeof m); m[0] = m[5] = m[10] = m[15] = 1.0f; }
    float& at(int c, int r) { return m[c * 4 + r]; }
    float at(int c, int r) const { return m[c * 4 + r]; }
    static mat4 identity() { return mat4(); }
    static mat4 translation(vec3 t) { mat4 r; r.m[12] = t.x; r.m[13] = t.y; r.m[14] = t.z; return r; }
    const float* data() const { return m; }
};

// detail/type_mat4x4.inl:660-697 (mul4x4, non-SIMD): column c = ((A0 * b.x + A1 * b.y) + A2 * b.z) + A3 * b.w
inline mat4 operator*(const mat4& a, const mat4& b) {
    mat4 r;
    for (int c = 0; c < 4; ++c)
        for (int row = 0; row < 4; ++row) {
            float s = a.at(0, row) * b.at(c, 0);
            s = s + a.at(1, row) * b.at(c, 1);
            s = s + a.at(2, row) * b.at(c, 2);
            s = s + a.at(3, row) * b.at(c, 3);
            r.at(c, row) = s;
        }
    return r;
}
// detail/type_mat4x4.inl:537-583: (m0 * v.x + m1 * v.y) + (m2 * v.z + m3 * v.w)
inline vec4 operator*(const mat4& a, vec4 v) {
    return {(a.m[0] * v.x + a.m[4] * v.y) + (a.m[8] * v.z + a.m[12] * v.w), (a.m[1] * v.x + a.m[5] * v.y) + (a.m[9] * v.z + a.m[13] * v.w),
            (a.m[2] * v.x + a.m[6] * v.y) + (a.m[10] * v.z + a.m[14] * v.w), (a.m[3] * v.x + a.m[7] * v.y) + (a.m[11] * v.z + a.m[15] * v.w)};
}

// right-handed look-at, camera looks down -z: ext/matrix_transform.inl:153-174 (lookAtRH)
inline mat4 lookAt(vec3 eye, vec3 center, vec3 up) {
    const vec3 f = normalize(center - eye), s = normalize(cross(f, up)), u = cross(s, f);
    mat4 r;
    r.at(0, 0) = s.x; r.at(1, 0) = s.y; r.at(2, 0) = s.z;
    r.at(0, 1) = u.x; r.at(1, 1) = u.y; r.at(2, 1) = u.z;
    r.at(0, 2) = -f.x; r.at(1, 2) = -f.y; r.at(2, 2) = -f.z;
    r.at(3, 0) = -dot(s, eye); r.at(3, 1) = -dot(u, eye); r.at(3, 2) = dot(f, eye);
    return r;
}
// right-handed perspective, depth in [0, 1]: ext/matrix_clip_space.inl:233-246 (perspectiveRH_ZO)
inline mat4 perspective(float fovyRadians, float aspect, float zNear, float zFar) {
    const float t = std::tan(fovyRadians / 2.0f);
    mat4 r;
    std::memset(r.m, 0, sizeof r.m);
    r.at(0, 0) = 1.0f / (aspect * t);
    r.at(1, 1) = 1.0f / t;
    r.at(2, 2) = zFar / (zNear - zFar);
    r.at(2, 3) = -1.0f;
    r.at(3, 2) = -(zFar * zNear) / (zFar - zNear);
    return r;
}
// right-handed orthographic, depth in [0, 1]: ext/matrix_clip_space.inl:42-52 (orthoRH_ZO)
inline mat4 ortho(float l, float r_, float b, float t, float n, float f) {
    mat4 r;
    r.at(0, 0) = 2.0f / (r_ - l); r.at(1, 1) = 2.0f / (t - b); r.at(2, 2) = -1.0f / (f - n);
    r.at(3, 0) = -(r_ + l) / (r_ - l); r.at(3, 1) = -(t + b) / (t - b); r.at(3, 2) = -n / (f - n);
    return r;
}
// glm_vk::kClip (glm_includes.h:18-23): y flip and z' = 0.5 z + 0.5 w (applied on top of ZO depth)
inline mat4 kClip() {
    mat4 r;
    r.at(1, 1) = -1.0f; r.at(2, 2) = 0.5f; r.at(3, 2) = 0.5f;
    return r;
}
inline float radians(float deg) { return deg * static_cast<float>(0.01745329251994329576923690768489); }   // detail/func_trigonometric.inl:9-14

// general 4x4 inverse: detail/func_matrix.inl:388-446 (compute_inverse<4, 4>) — 2x2 sub-determinants, cofactor columns with
// alternating signs, determinant = (d.x + d.y) + (d.z + d.w) over m[0] * first row of the adjugate, one reciprocal, all in fp32.
inline mat4 inverse(const mat4& a) {
    auto M = [&](int c, int r) { return a.m[c * 4 + r]; };
    const float c00 = M(2, 2) * M(3, 3) - M(3, 2) * M(2, 3), c02 = M(1, 2) * M(3, 3) - M(3, 2) * M(1, 3), c03 = M(1, 2) * M(2, 3) - M(2, 2) * M(1, 3);
    const float c04 = M(2, 1) * M(3, 3) - M(3, 1) * M(2, 3), c06 = M(1, 1) * M(3, 3) - M(3, 1) * M(1, 3), c07 = M(1, 1) * M(2, 3) - M(2, 1) * M(1, 3);
    const float c08 = M(2, 1) * M(3, 2) - M(3, 1) * M(2, 2), c10 = M(1, 1) * M(3, 2) - M(3, 1) * M(1, 2), c11 = M(1, 1) * M(2, 2) - M(2, 1) * M(1, 2);
    const float c12 = M(2, 0) * M(3, 3) - M(3, 0) * M(2, 3), c14 = M(1, 0) * M(3, 3) - M(3, 0) * M(1, 3), c15 = M(1, 0) * M(2, 3) - M(2, 0) * M(1, 3);
    const float c16 = M(2, 0) * M(3, 2) - M(3, 0) * M(2, 2), c18 = M(1, 0) * M(3, 2) - M(3, 0) * M(1, 2), c19 = M(1, 0) * M(2, 2) - M(2, 0) * M(1, 2);
    const float c20 = M(2, 0) * M(3, 1) - M(3, 0) * M(2, 1), c22 = M(1, 0) * M(3, 1) - M(3, 0) * M(1, 1), c23 = M(1, 0) * M(2, 1) - M(2, 0) * M(1, 1);
    const float f0[4] = {c00, c00, c02, c03}, f1[4] = {c04, c04, c06, c07}, f2[4] = {c08, c08, c10, c11};
    const float f3[4] = {c12, c12, c14, c15}, f4[4] = {c16, c16, c18, c19}, f5[4] = {c20, c20, c22, c23};
    const float v0[4] = {M(1, 0), M(0, 0), M(0, 0), M(0, 0)}, v1[4] = {M(1, 1), M(0, 1), M(0, 1), M(0, 1)};
    const float v2[4] = {M(1, 2), M(0, 2), M(0, 2), M(0, 2)}, v3[4] = {M(1, 3), M(0, 3), M(0, 3), M(0, 3)};
    const float sa[4] = {+1.0f, -1.0f, +1.0f, -1.0f}, sb[4] = {-1.0f, +1.0f, -1.0f, +1.0f};
    mat4 inv;
    for (int k = 0; k < 4; ++k) {
        inv.m[0 + k] = ((v1[k] * f0[k] - v2[k] * f1[k]) + v3[k] * f2[k]) * sa[k];
        inv.m[4 + k] = ((v0[k] * f0[k] - v2[k] * f3[k]) + v3[k] * f4[k]) * sb[k];
        inv.m[8 + k] = ((v0[k] * f1[k] - v1[k] * f3[k]) + v3[k] * f5[k]) * sa[k];
        inv.m[12 + k] = ((v0[k] * f2[k] - v1[k] * f4[k]) + v2[k] * f5[k]) * sb[k];
    }
    const float d0 = M(0, 0) * inv.m[0], d1 = M(0, 1) * inv.m[4], d2 = M(0, 2) * inv.m[8], d3 = M(0, 3) * inv.m[12];
    const float ood = 1.0f / ((d0 + d1) + (d2 + d3));
    for (int i = 0; i < 16; ++i) inv.m[i] = inv.m[i] * ood;
    return inv;
}

}  // namespace reb
