// reb_sky.cuh — skybox background (Appendix B.14 of DESIGN.md).
//
// The reference draws a unit cube last with gl_Position = (proj * mat4(mat3(view)) * (p, 1)).xyww, i.e. at depth
// exactly 1.0 under LessOrEqual (skybox.vert:32-37, RenderScene.cpp:226-229, Skybox.cpp:126-140), and colours it with
// texture(cube, p) (skybox.frag:28-31).  A cube lookup depends only on the direction of p, and the direction seen
// through a pixel centre is A^-1 (xn, yn, 1) (A = rows x, y, w of proj * mat4(mat3(view)), 3x3), so the twelve
// triangles are never rasterised: the tile stage evaluates the ray for every pixel whose depth is still 1.0.
// Cube map = six RGBA16 UNORM faces, one mip, linear filter (TextureCache.cpp:251-266), seamless across edges.
// Same individually rounded operations as oracle/oracle.cpp (sky_pixel / sky_lookup / cube_texel).
#pragma once
#include "reb_kernels.h"

namespace reb {

__device__ __forceinline__ int cube_face(float rx, float ry, float rz, float& sc, float& tc, float& ma) {
    const float ax = rx >= 0.0f ? rx : -rx, ay = ry >= 0.0f ? ry : -ry, az = rz >= 0.0f ? rz : -rz;
    if (az >= ax && az >= ay) { ma = az; if (rz >= 0.0f) { sc = rx; tc = -ry; return 4; } sc = -rx; tc = -ry; return 5; }
    if (ay >= ax) { ma = ay; if (ry >= 0.0f) { sc = rx; tc = rz; return 2; } sc = rx; tc = -rz; return 3; }
    ma = ax;
    if (rx >= 0.0f) { sc = -rz; tc = -ry; return 0; }
    sc = rz; tc = -ry; return 1;
}

__device__ __forceinline__ float4 cube_fetch(const uint16_t* __restrict__ faces, uint32_t w, uint32_t h, int f, int i, int j) {
    const uint2 t = __ldg(reinterpret_cast<const uint2*>(faces) + ((size_t)f * h + (size_t)j) * w + (size_t)i);
    const float k = 1.0f / 65535.0f;
    return make_float4((float)(t.x & 0xFFFFu) * k, (float)(t.x >> 16) * k, (float)(t.y & 0xFFFFu) * k, (float)(t.y >> 16) * k);
}

__device__ __forceinline__ float4 cube_texel(const uint16_t* __restrict__ faces, uint32_t w, uint32_t h, int f, int i, int j) {
    if (i >= 0 && j >= 0 && i < (int)w && j < (int)h) return cube_fetch(faces, w, h, f, i, j);
    // over the edge: lift the texel centre onto the extended face plane, re-select the face, take the nearest texel
    const float sc = __fdiv_rn(2.0f * ((float)i + 0.5f), (float)w) - 1.0f, tc = __fdiv_rn(2.0f * ((float)j + 0.5f), (float)h) - 1.0f;
    float rx, ry, rz;
    switch (f) {
        case 0: rx = 1.0f; ry = -tc; rz = -sc; break;
        case 1: rx = -1.0f; ry = -tc; rz = sc; break;
        case 2: rx = sc; ry = 1.0f; rz = tc; break;
        case 3: rx = sc; ry = -1.0f; rz = -tc; break;
        case 4: rx = sc; ry = -tc; rz = 1.0f; break;
        default: rx = -sc; ry = -tc; rz = -1.0f; break;
    }
    float s2, t2, ma;
    const int f2 = cube_face(rx, ry, rz, s2, t2, ma);
    const float s = 0.5f * __fdiv_rn(s2, ma) + 0.5f, t = 0.5f * __fdiv_rn(t2, ma) + 0.5f;
    int i2 = (int)floorf(s * (float)w), j2 = (int)floorf(t * (float)h);
    i2 = max(0, min((int)w - 1, i2));
    j2 = max(0, min((int)h - 1, j2));
    return cube_fetch(faces, w, h, f2, i2, j2);
}

// Colour of the skybox through the centre of pixel (x, y).  Kept out of line: it runs for background pixels only and
// must not weigh on the register budget of the shading path.
static __device__ __noinline__ float4 sky_pixel(const ShadeConst& sh, int x, int y) {
    const float xn = (((float)x + 0.5f) - sh.skyHw) * sh.skyRhw, yn = (((float)y + 0.5f) - sh.skyHh) * sh.skyRhh;
    const float* a = sh.skyAinv;
    const float rx = (a[0] * xn + a[3] * yn) + a[6], ry = (a[1] * xn + a[4] * yn) + a[7], rz = (a[2] * xn + a[5] * yn) + a[8];
    float sc, tc, ma;
    const int f = cube_face(rx, ry, rz, sc, tc, ma);
    const uint32_t w = sh.skyW, h = sh.skyH;
    const float s = 0.5f * __fdiv_rn(sc, ma) + 0.5f, t = 0.5f * __fdiv_rn(tc, ma) + 0.5f;
    const float u = s * (float)w - 0.5f, v = t * (float)h - 0.5f;
    const float fu0 = floorf(u), fv0 = floorf(v);
    const int i0 = (int)fu0, j0 = (int)fv0;
    const float fu = u - fu0, fv = v - fv0;
    const float4 t00 = cube_texel(sh.skyFaces, w, h, f, i0, j0), t10 = cube_texel(sh.skyFaces, w, h, f, i0 + 1, j0);
    const float4 t01 = cube_texel(sh.skyFaces, w, h, f, i0, j0 + 1), t11 = cube_texel(sh.skyFaces, w, h, f, i0 + 1, j0 + 1);
    const float gu = 1.0f - fu, gv = 1.0f - fv;
    float4 o;
    o.x = (t00.x * gu + t10.x * fu) * gv + (t01.x * gu + t11.x * fu) * fv;
    o.y = (t00.y * gu + t10.y * fu) * gv + (t01.y * gu + t11.y * fu) * fv;
    o.z = (t00.z * gu + t10.z * fu) * gv + (t01.z * gu + t11.z * fu) * fv;
    o.w = (t00.w * gu + t10.w * fu) * gv + (t01.w * gu + t11.w * fu) * fv;
    return o;
}

}  // namespace reb
