/*
 * oracle.cpp — CPU oracle for the RenderEngine raster hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this.  The product (renderengine_b200/) never links, imports or calls it.
 *
 * PARITY STATUS: "parity unpinned" at the rasteriser boundary — the reference has no tests,
 * goldens or CPU rasteriser and its Vulkan path cannot run here.  The
 * fixed-function rules below instantiate the Vulkan state the reference sets
 * (Source/VkRHI/RHI/GraphicsPipelineCache.cpp:257-365) per SURVEY.md Appendix B.  The shading
 * arithmetic restates Assets/Shaders/phong.glsl:39-81 and shadow.glsl:24-56 in glm's operation
 * order and IS pinned: tests/test_oracle_golden.py checks it bit-for-bit against vectors produced
 * by the reference's own GLSL compiled as C++ (oracle/ref_shim.cpp -> oracle/_ref/libref.so).
 *
 * Build: g++ -O2 -ffp-contract=off -fno-fast-math -shared -fPIC (oracle/Makefile).  Every fp32
 * operation on the depth / coverage / shadow-lookup path is written as an un-fused binary op in a
 * fixed order; the CUDA path must reproduce that order exactly.
 */
#include "oracle.h"

#include <algorithm>
#include <cmath>
#include <cstring>
#include <thread>
#include <vector>

namespace {

constexpr uint32_t LIGHT_TYPE_DIRECTIONAL = 1;   // phong.glsl:16
constexpr uint32_t LIGHT_TYPE_POINT = 2;         // pbr.glsl:283-285
constexpr uint32_t LIGHT_TYPE_SPOT = 3;          // phong.glsl:18
constexpr float GUARD_PX = 1048576.0f;           // |x_f|,|y_f| <= 2^20 px keeps 24.8 fixed point in int32
constexpr uint32_t TRI_NONE = 0xFFFFFFFFu;

struct V3 { float x, y, z; };
struct V4 { float x, y, z, w; };

// glm::mat4 * glm::vec4 — ThirdParty/glm/glm/detail/type_mat4x4.inl:562-573:
// (m0*v0 + m1*v1) + (m2*v2 + m3*v3), column-major m[c*4+r].
inline V4 mat4_mul(const float* m, V4 v) {
    float o[4];
    for (int r = 0; r < 4; ++r) {
        float mul0 = m[0 + r] * v.x;
        float mul1 = m[4 + r] * v.y;
        float add0 = mul0 + mul1;
        float mul2 = m[8 + r] * v.z;
        float mul3 = m[12 + r] * v.w;
        float add1 = mul2 + mul3;
        o[r] = add0 + add1;
    }
    return {o[0], o[1], o[2], o[3]};
}

// glm::mat4 * glm::mat4, non-aligned path — type_mat4x4.inl:679-694: ((A0*b.x + A1*b.y) + A2*b.z) + A3*b.w.
inline void mat4_mul_mat4(const float* a, const float* b, float* out) {
    float tmp[16];
    for (int c = 0; c < 4; ++c)
        for (int r = 0; r < 4; ++r) {
            float t = a[0 + r] * b[c * 4 + 0];
            t = t + a[4 + r] * b[c * 4 + 1];
            t = t + a[8 + r] * b[c * 4 + 2];
            t = t + a[12 + r] * b[c * 4 + 3];
            tmp[c * 4 + r] = t;
        }
    std::memcpy(out, tmp, sizeof tmp);
}

inline float dot3(V3 a, V3 b) {   // glm compute_dot<vec3>: func_geometric.inl:48-55
    float tx = a.x * b.x, ty = a.y * b.y, tz = a.z * b.z;
    return (tx + ty) + tz;
}
inline V3 sub3(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline V3 add3(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline V3 neg3(V3 a) { return {-a.x, -a.y, -a.z}; }
inline V3 normalize3(V3 v) {      // v * (1/sqrt(dot(v,v))) — func_geometric.inl:104, func_exponential.inl:138
    float inv = 1.0f / std::sqrt(dot3(v, v));
    return {v.x * inv, v.y * inv, v.z * inv};
}
inline float glm_max(float x, float y) { return (x < y) ? y : x; }   // func_common.inl
inline float glm_min(float x, float y) { return (y < x) ? y : x; }

// transpose(inverse(mat3(T))) — primitive.vert:55; glm inv3x3<false> func_matrix.inl:353-376.
// out9 is column-major N[c*3+r].
void normal_matrix(const float* T, float* N) {
    auto m = [&](int c, int r) { return T[c * 4 + r]; };
    float det = +m(0, 0) * (m(1, 1) * m(2, 2) - m(2, 1) * m(1, 2))
                - m(1, 0) * (m(0, 1) * m(2, 2) - m(2, 1) * m(0, 2))
                + m(2, 0) * (m(0, 1) * m(1, 2) - m(1, 1) * m(0, 2));
    float ood = 1.0f / det;
    float inv[3][3];   // inv[c][r]
    inv[0][0] = +(m(1, 1) * m(2, 2) - m(2, 1) * m(1, 2));
    inv[1][0] = -(m(1, 0) * m(2, 2) - m(2, 0) * m(1, 2));
    inv[2][0] = +(m(1, 0) * m(2, 1) - m(2, 0) * m(1, 1));
    inv[0][1] = -(m(0, 1) * m(2, 2) - m(2, 1) * m(0, 2));
    inv[1][1] = +(m(0, 0) * m(2, 2) - m(2, 0) * m(0, 2));
    inv[2][1] = -(m(0, 0) * m(2, 1) - m(2, 0) * m(0, 1));
    inv[0][2] = +(m(0, 1) * m(1, 2) - m(1, 1) * m(0, 2));
    inv[1][2] = -(m(0, 0) * m(1, 2) - m(1, 0) * m(0, 2));
    inv[2][2] = +(m(0, 0) * m(1, 1) - m(1, 0) * m(0, 1));
    for (int c = 0; c < 3; ++c)
        for (int r = 0; r < 3; ++r) inv[c][r] = inv[c][r] * ood;
    for (int c = 0; c < 3; ++c)
        for (int r = 0; r < 3; ++r) N[c * 3 + r] = inv[r][c];   // transpose
}
inline V3 mat3_mul(const float* N, V3 v) {   // type_mat3x3.inl:472-473: (m0*x + m1*y) + m2*z
    V3 o;
    o.x = (N[0] * v.x + N[3] * v.y) + N[6] * v.z;
    o.y = (N[1] * v.x + N[4] * v.y) + N[7] * v.z;
    o.z = (N[2] * v.x + N[5] * v.y) + N[8] * v.z;
    return o;
}

struct Light {          // std430 layout of phong.glsl:20-37
    int32_t type; V3 position; V3 direction; V4 ambient, diffuse, specular;
    float innerCutoff, outerCutoff; uint32_t shadowIndex;
};
Light load_light(const uint8_t* p) {
    Light l;
    std::memcpy(&l.type, p + 0, 4);
    std::memcpy(&l.position, p + 16, 12);
    std::memcpy(&l.direction, p + 32, 12);
    std::memcpy(&l.ambient, p + 48, 16);
    std::memcpy(&l.diffuse, p + 64, 16);
    std::memcpy(&l.specular, p + 80, 16);
    std::memcpy(&l.innerCutoff, p + 96, 4);
    std::memcpy(&l.outerCutoff, p + 100, 4);
    std::memcpy(&l.shadowIndex, p + 104, 4);
    return l;
}
struct PhongProps { V4 diffuse, specular; float shininess; };

// phong.glsl:39-81, statement for statement.
V4 phong_lighting(const Light& light, const PhongProps& props, V3 normal, V3 fragPos, V3 viewPos, float shadow) {
    V3 viewDir = normalize3(sub3(viewPos, fragPos));
    V3 lightDir = (uint32_t)light.type == LIGHT_TYPE_DIRECTIONAL ? neg3(light.direction) : sub3(light.position, fragPos);
    float lightDistance = std::sqrt(dot3(lightDir, lightDir));
    lightDir = {lightDir.x / lightDistance, lightDir.y / lightDistance, lightDir.z / lightDistance};

    V4 ambient = {light.ambient.x * props.diffuse.x, light.ambient.y * props.diffuse.y,
                  light.ambient.z * props.diffuse.z, light.ambient.w * props.diffuse.w};

    float k_d = glm_max(dot3(lightDir, normal), 0.0f);
    V4 diffuse = {(k_d * light.diffuse.x) * props.diffuse.x, (k_d * light.diffuse.y) * props.diffuse.y,
                  (k_d * light.diffuse.z) * props.diffuse.z, (k_d * light.diffuse.w) * props.diffuse.w};

    V3 h = normalize3(add3(lightDir, viewDir));
    float k_s = std::pow(glm_max(dot3(normal, h), 0.0f), props.shininess);
    V4 specular = {(k_s * light.specular.x) * props.specular.x, (k_s * light.specular.y) * props.specular.y,
                   (k_s * light.specular.z) * props.specular.z, (k_s * light.specular.w) * props.specular.w};

    float attenuation = 1.0f;
    if ((uint32_t)light.type != LIGHT_TYPE_DIRECTIONAL) attenuation = 1.0f / lightDistance;

    if ((uint32_t)light.type == LIGHT_TYPE_SPOT) {
        float cos_theta = dot3(neg3(lightDir), normalize3(light.direction));
        float epsilon = light.innerCutoff - light.outerCutoff;
        float intensity = glm_min(glm_max((cos_theta - light.outerCutoff) / epsilon, 0.0f), 1.0f);
        diffuse = {diffuse.x * intensity, diffuse.y * intensity, diffuse.z * intensity, diffuse.w * intensity};
        specular = {specular.x * intensity, specular.y * intensity, specular.z * intensity, specular.w * intensity};
    }
    float k = 1.0f - shadow;
    V4 r;
    r.x = (ambient.x + k * (diffuse.x + specular.x)) * attenuation;
    r.y = (ambient.y + k * (diffuse.y + specular.y)) * attenuation;
    r.z = (ambient.z + k * (diffuse.z + specular.z)) * attenuation;
    r.w = (ambient.w + k * (diffuse.w + specular.w)) * attenuation;
    return r;
}

// Nearest, clamp-to-edge fetch of the S x S f32 shadow map (ShadowSystem.cpp:18-35; Appendix B.13).
inline float shadow_fetch(const float* map, uint32_t S, float u, float v) {
    float fs = (float)S;
    float fu = std::floor(u * fs), fv = std::floor(v * fs);
    float hi = (float)(S - 1);
    // clamp in float first so NaN / huge values are well defined (NaN -> 0)
    int i = (fu >= 0.0f) ? ((fu <= hi) ? (int)fu : (int)(S - 1)) : 0;
    int j = (fv >= 0.0f) ? ((fv <= hi) ? (int)fv : (int)(S - 1)) : 0;
    return map[(size_t)j * S + i];
}

// shadow.glsl:24-56, statement for statement.
float compute_shadow(const Light& light, V3 fragPos, V3 normal, const float* transform, const float* map, uint32_t S) {
    if ((uint32_t)light.type != LIGHT_TYPE_DIRECTIONAL) return 0.0f;
    V4 flp = mat4_mul(transform, {fragPos.x, fragPos.y, fragPos.z, 1.0f});
    V3 lightDir = normalize3(sub3(light.position, fragPos));
    float currentDepth = flp.z / flp.w;
    if (currentDepth > 1.0f) return 0.0f;
    float mu = 0.5f * (flp.x / flp.w) + 0.5f;
    float mv = 0.5f * (flp.y / flp.w) + 0.5f;
    float bias = glm_max(0.05f * (1.0f - dot3(normal, lightDir)), 0.005f);
    float shadow = 0.0f;
    float texel = 1.0f / (float)S;
    for (int x = -1; x <= 1; ++x)
        for (int y = -1; y <= 1; ++y) {
            float depth = shadow_fetch(map, S, mu + (float)x * texel, mv + (float)y * texel);
            shadow += (currentDepth - bias > depth) ? 1.0f : 0.0f;
        }
    shadow /= 9.0f;
    return shadow;
}

// Exact sRGB OETF + 8-bit quantisation of the R8G8B8A8_SRGB target (Swapchain.cpp:10-21; Appendix B.12).
inline uint8_t quant8(float x) {
    if (!(x > 0.0f)) return 0;
    if (x >= 1.0f) return 255;
    return (uint8_t)(int)std::floor(x * 255.0f + 0.5f);
}
inline float srgb_oetf(float x) {
    if (!(x > 0.0f)) return 0.0f;
    if (x >= 1.0f) return 1.0f;
    return (x <= 0.0031308f) ? 12.92f * x : 1.055f * std::pow(x, 1.0f / 2.4f) - 0.055f;
}

// ---------------------------------------------------------------------------------------------
// Bindless 2-D textures.  Object semantics follow TextureCache.cpp:67-73,167-197 and
// Texture.cpp:50-136: R16G16B16A16_UNORM, floor(log2(max(w,h)))+1 mips made by successive 2x linear blits
// (= 2x2 box average for even extents, edge texels repeated for odd ones), sampler linear / linear /
// mip-linear with repeat addressing and LOD in [0, mips-1].  The 16x anisotropy the reference requests is
// implementation-defined in Vulkan; the oracle defines ISOTROPIC trilinear filtering with the LOD taken from
// the longer screen-space derivative (builder decision).  Colour textures are decoded with
// the exact sRGB EOTF after filtering, as pbr.glsl:20-26,100-108 does for the base colour; alpha stays linear.
// ---------------------------------------------------------------------------------------------
struct Texture {
    std::vector<std::vector<uint16_t>> mips;   // RGBA16 per level
    std::vector<uint32_t> w, h;
    bool srgb = false;
};

Texture build_texture(const orc_texture& t) {
    Texture x;
    x.srgb = t.srgb != 0;
    x.w.push_back(t.width); x.h.push_back(t.height);
    x.mips.emplace_back(t.texels, t.texels + (size_t)t.width * t.height * 4);
    while (x.w.back() > 1 || x.h.back() > 1) {
        const uint32_t sw = x.w.back(), sh = x.h.back(), dw = std::max(1u, sw / 2), dh = std::max(1u, sh / 2);
        const std::vector<uint16_t>& src = x.mips.back();
        std::vector<uint16_t> dst((size_t)dw * dh * 4);
        for (uint32_t j = 0; j < dh; ++j)
            for (uint32_t i = 0; i < dw; ++i) {
                const uint32_t i0 = std::min(2 * i, sw - 1), i1 = std::min(2 * i + 1, sw - 1), j0 = std::min(2 * j, sh - 1), j1 = std::min(2 * j + 1, sh - 1);
                for (int c = 0; c < 4; ++c) {
                    const uint32_t sum = (uint32_t)src[((size_t)j0 * sw + i0) * 4 + c] + src[((size_t)j0 * sw + i1) * 4 + c] +
                                         src[((size_t)j1 * sw + i0) * 4 + c] + src[((size_t)j1 * sw + i1) * 4 + c];
                    dst[((size_t)j * dw + i) * 4 + c] = (uint16_t)((sum + 2u) >> 2);
                }
            }
        x.w.push_back(dw); x.h.push_back(dh);
        x.mips.push_back(std::move(dst));
    }
    return x;
}

inline int wrap_repeat(int i, int n) { int m = i % n; return m < 0 ? m + n : m; }

V4 sample_bilinear(const Texture& t, int level, float u, float v) {
    const int w = (int)t.w[level], h = (int)t.h[level];
    const float x = u * (float)w - 0.5f, y = v * (float)h - 0.5f;
    const float fx0 = std::floor(x), fy0 = std::floor(y);
    const float fx = x - fx0, fy = y - fy0;
    // huge / non-finite coordinates are outside the bit-exact domain; keep the index arithmetic defined
    const int ix = (std::fabs(fx0) < 1.0e9f) ? (int)fx0 : 0, iy = (std::fabs(fy0) < 1.0e9f) ? (int)fy0 : 0;
    const int i0 = wrap_repeat(ix, w), i1 = wrap_repeat(ix + 1, w), j0 = wrap_repeat(iy, h), j1 = wrap_repeat(iy + 1, h);
    const uint16_t* p = t.mips[level].data();
    float o[4];
    for (int c = 0; c < 4; ++c) {
        const float t00 = (float)p[((size_t)j0 * w + i0) * 4 + c] * (1.0f / 65535.0f), t10 = (float)p[((size_t)j0 * w + i1) * 4 + c] * (1.0f / 65535.0f);
        const float t01 = (float)p[((size_t)j1 * w + i0) * 4 + c] * (1.0f / 65535.0f), t11 = (float)p[((size_t)j1 * w + i1) * 4 + c] * (1.0f / 65535.0f);
        const float a = t00 + fx * (t10 - t00), b = t01 + fx * (t11 - t01);
        o[c] = a + fy * (b - a);
    }
    return {o[0], o[1], o[2], o[3]};
}

inline float srgb_eotf(float c) { return (c <= 0.04045f) ? c / 12.92f : std::pow((c + 0.055f) / 1.055f, 2.4f); }

// texture(sampler2D, uv) with implicit derivatives (ddx, ddy = uv at the +x / +y neighbour pixel minus uv here)
V4 sample_texture(const Texture& t, float u, float v, float dudx, float dvdx, float dudy, float dvdy) {
    const float W = (float)t.w[0], H = (float)t.h[0];
    const float rx = std::sqrt((dudx * W) * (dudx * W) + (dvdx * H) * (dvdx * H));
    const float ry = std::sqrt((dudy * W) * (dudy * W) + (dvdy * H) * (dvdy * H));
    const float rho = glm_max(rx, ry);
    const int last = (int)t.mips.size() - 1;
    float lod = (rho > 0.0f) ? std::log2(rho) : 0.0f;
    if (!(lod > 0.0f)) lod = 0.0f;
    if (lod > (float)last) lod = (float)last;
    const int l0 = (int)std::floor(lod), l1 = std::min(l0 + 1, last);
    const float f = lod - (float)l0;
    V4 a = sample_bilinear(t, l0, u, v);
    if (f > 0.0f && l1 != l0) {
        const V4 b = sample_bilinear(t, l1, u, v);
        a = {a.x + f * (b.x - a.x), a.y + f * (b.y - a.y), a.z + f * (b.z - a.z), a.w + f * (b.w - a.w)};
    }
    if (t.srgb) a = {srgb_eotf(a.x), srgb_eotf(a.y), srgb_eotf(a.z), a.w};
    return a;
}

// ---------------------------------------------------------------------------------------------
// Geometry front end (Appendix B.1-6)
// ---------------------------------------------------------------------------------------------
struct Setup {
    int32_t X[3], Y[3];
    float z[3];
    float w[3];
    int64_t A2;      // < 0 for kept triangles
    uint32_t triId;
    int32_t x0, x1, y0, y1;   // inclusive pixel bbox (already clamped to the target)
};

enum SetupResult { KEEP = 0, REJECT = 1, NEEDS_CLIP = 2, CULLED = 3, NO_SAMPLE = 4 };

struct PassParams {
    const float* PV;     // proj*view (formed once with mat4_mul_mat4)
    bool divideWorldByW; // primitive.vert:51 (forward) vs shadow_map.vert:49 (shadow)
    uint32_t W, H;
    // Appendix B.15 (multisampling): sample k of a pixel sits at (x + sx, y + sy) instead of the centre.  Edge functions
    // and depth are translation-invariant, so sampling there equals sampling the centre of the primitive moved by
    // (128 - 256 sx, 128 - 256 sy) sub-pixel units — an exact integer shift of the snapped vertices.
    int32_t shiftX = 0, shiftY = 0;
    // Appendix B.16 (translucent draws, cull none): a triangle with A2 > 0 is rasterised with its vertices 1 and 2 swapped
    bool twoSided = false;
};
// Vulkan standard sample locations for 4 samples ((0.375, 0.125), (0.875, 0.375), (0.125, 0.625), (0.625, 0.875)) as shifts
static const int32_t MSAA4_SHIFT[4][2] = {{32, 96}, {-96, 32}, {96, -32}, {-32, -96}};

inline V4 vertex_stage(const PassParams& pp, const float* T, const float* pos, V3* worldOut) {
    V4 wH = mat4_mul(T, {pos[0], pos[1], pos[2], 1.0f});
    V3 world;
    if (pp.divideWorldByW) world = {wH.x / wH.w, wH.y / wH.w, wH.z / wH.w};
    else world = {wH.x, wH.y, wH.z};
    if (worldOut) *worldOut = world;
    return mat4_mul(pp.PV, {world.x, world.y, world.z, 1.0f});
}

inline SetupResult setup_triangle(const PassParams& pp, const V4 c[3], uint32_t triId, Setup& s, bool skipTrivialReject = false) {
    // B.2 trivial reject: all three outside one clip plane (the parent of clipped children already passed it)
    bool allL = !skipTrivialReject, allR = allL, allT = allL, allB = allL, allN = allL, allF = allL;
    for (int k = 0; k < 3; ++k) {
        allL = allL && (c[k].x < -c[k].w); allR = allR && (c[k].x > c[k].w);
        allT = allT && (c[k].y < -c[k].w); allB = allB && (c[k].y > c[k].w);
        allN = allN && (c[k].z < 0.0f);    allF = allF && (c[k].z > c[k].w);
    }
    if (allL || allR || allT || allB || allN || allF) return REJECT;
    // B.3: a vertex at or behind the eye plane needs true clipping (not part of the bit-exact domain)
    for (int k = 0; k < 3; ++k) if (!(c[k].w > 0.0f)) return NEEDS_CLIP;
    float hw = 0.5f * (float)pp.W, hh = 0.5f * (float)pp.H;
    for (int k = 0; k < 3; ++k) {
        float rw = 1.0f / c[k].w;                                            // B.4 (revision r1b): one IEEE reciprocal per vertex,
        float nx = c[k].x * rw, ny = c[k].y * rw, nz = c[k].z * rw;          // the same 1/w the varyings use (B.10)
        float xf = hw * nx + hw;
        float yf = hh * ny + hh;
        if (!(std::fabs(xf) <= GUARD_PX) || !(std::fabs(yf) <= GUARD_PX)) return NEEDS_CLIP;
        s.X[k] = (int32_t)std::lrintf(xf * 256.0f) + pp.shiftX;   // B.5 round-half-even, 8 sub-pixel bits (+ B.15 sample shift)
        s.Y[k] = (int32_t)std::lrintf(yf * 256.0f) + pp.shiftY;
        s.z[k] = nz;
        s.w[k] = c[k].w;
    }
    // B.6 facing: front = CCW = A2 < 0; cull back (A2 > 0) and degenerate (A2 == 0)
    s.A2 = (int64_t)(s.X[1] - s.X[0]) * (int64_t)(s.Y[2] - s.Y[0]) - (int64_t)(s.X[2] - s.X[0]) * (int64_t)(s.Y[1] - s.Y[0]);
    if (pp.twoSided && s.A2 > 0) {
        std::swap(s.X[1], s.X[2]); std::swap(s.Y[1], s.Y[2]); std::swap(s.z[1], s.z[2]); std::swap(s.w[1], s.w[2]);
        s.A2 = -s.A2;
    }
    if (s.A2 >= 0) return CULLED;
    int32_t minX = std::min(s.X[0], std::min(s.X[1], s.X[2])), maxX = std::max(s.X[0], std::max(s.X[1], s.X[2]));
    int32_t minY = std::min(s.Y[0], std::min(s.Y[1], s.Y[2])), maxY = std::max(s.Y[0], std::max(s.Y[1], s.Y[2]));
    // pixel centres 256*i+128 inside [min,max]
    s.x0 = std::max((minX + 127) >> 8, 0);
    s.x1 = std::min((maxX - 128) >> 8, (int32_t)pp.W - 1);
    s.y0 = std::max((minY + 127) >> 8, 0);
    s.y1 = std::min((maxY - 128) >> 8, (int32_t)pp.H - 1);
    if (s.x0 > s.x1 || s.y0 > s.y1) return NO_SAMPLE;
    s.triId = triId;
    return KEEP;
}

// ---------------------------------------------------------------------------------------------
// B.3 clipping (revision r1c): a triangle with a vertex at or behind the eye plane (w <= 0) or outside the guard
// band is clipped in clip space with Sutherland-Hodgman, fp32, fixed plane order (near z >= 0, then the four guard
// planes |x| <= Gx*w, |y| <= Gy*w with G = (GUARD_PX / 2) / (extent / 2), so a clipped vertex lands within 2^19 px
// (+ half the extent) of the origin), every intersection computed FROM THE INSIDE vertex towards the outside one:
// t = d_in / (d_in - d_out), v = v_in + t * (v_out - v_in).  The polygon (<= 8 vertices) is fanned from its first
// vertex; every child runs the ordinary B.4-B.8 setup and inherits the parent's triId.  Varyings of a clipped parent
// are interpolated in homogeneous form at shading time (homogeneous_weights below), so children carry positions only.
// ---------------------------------------------------------------------------------------------
inline float clip_dist(int plane, const V4& v, float Gx, float Gy) {
    switch (plane) {
        case 0: return v.z;
        case 1: return v.x + Gx * v.w;
        case 2: return Gx * v.w - v.x;
        case 3: return v.y + Gy * v.w;
        default: return Gy * v.w - v.y;
    }
}
inline int clip_polygon(const PassParams& pp, const V4 in[3], V4 out[9]) {
    const float Gx = (0.5f * GUARD_PX) / (0.5f * (float)pp.W), Gy = (0.5f * GUARD_PX) / (0.5f * (float)pp.H);
    V4 a[9], b[9];
    int n = 3;
    for (int k = 0; k < 3; ++k) a[k] = in[k];
    for (int plane = 0; plane < 5; ++plane) {
        int m = 0;
        for (int i = 0; i < n; ++i) {
            const V4& P = a[i]; const V4& Q = a[(i + 1) % n];
            const float dP = clip_dist(plane, P, Gx, Gy), dQ = clip_dist(plane, Q, Gx, Gy);
            const bool inP = dP >= 0.0f, inQ = dQ >= 0.0f;
            if (inP) b[m++] = P;
            if (inP != inQ) {
                const V4& I = inP ? P : Q; const V4& O = inP ? Q : P;
                const float dI = inP ? dP : dQ, dO = inP ? dQ : dP;
                const float t = dI / (dI - dO);
                b[m++] = {I.x + t * (O.x - I.x), I.y + t * (O.y - I.y), I.z + t * (O.z - I.z), I.w + t * (O.w - I.w)};
            }
        }
        n = m;
        if (n < 3) return 0;
        for (int k = 0; k < n; ++k) a[k] = b[k];
    }
    for (int k = 0; k < n; ++k) out[k] = a[k];
    return n;
}

// Perspective-correct weights of the PARENT triangle at pixel centre (x, y) from its clip-space vertices: lambda = adj([c0 c1 c2]) * (xn, yn, 1)
// over the (x, y, w) rows; attribute = ((l0 a0 + l1 a1) + l2 a2) * (1 / ((l0 + l1) + l2)).
inline void homogeneous_weights(const PassParams& pp, const V4 c[3], float px, float py, float l[3]) {
    const float hw = 0.5f * (float)pp.W, hh = 0.5f * (float)pp.H;
    const float xn = ((px + 0.5f) - hw) / hw, yn = ((py + 0.5f) - hh) / hh;
    for (int k = 0; k < 3; ++k) {
        const V4& u = c[(k + 1) % 3]; const V4& v = c[(k + 2) % 3];
        l[k] = ((u.y * v.w - u.w * v.y) * xn + (u.w * v.x - u.x * v.w) * yn) + (u.x * v.y - u.y * v.x);
    }
}

// B.7 edge functions with interior positive (for A2 < 0) and the top-left bias folded in.
struct Edges {
    int64_t A[3], B[3], C[3];   // F_k(P) = A*Px + B*Py + C  (unbiased)
    int64_t bias[3];            // 0 for top/left edges, 1 otherwise: covered iff F_k - bias_k >= 0
};
inline Edges make_edges(const Setup& s) {
    Edges e;
    for (int k = 0; k < 3; ++k) {
        int a = (k + 1) % 3, b = (k + 2) % 3;   // edge a->b opposes vertex k
        int64_t A = (int64_t)s.Y[b] - (int64_t)s.Y[a];
        int64_t B = -((int64_t)s.X[b] - (int64_t)s.X[a]);
        e.A[k] = A; e.B[k] = B;
        e.C[k] = -(A * (int64_t)s.X[a]) - B * (int64_t)s.Y[a];
        bool topLeft = (A > 0) || (A == 0 && B > 0);
        e.bias[k] = topLeft ? 0 : 1;
    }
    return e;
}

// B.8 depth at a covered sample; F1, F2 are the unbiased edge values of vertices 1 and 2.
// Revision r1b: the barycentrics are formed with ONE IEEE reciprocal per triangle, b = float(F) * (1 / |A2|),
// the way a rasteriser's plane-equation setup does it (the first revision divided per sample).
inline float depth_at(const Setup& s, int64_t F1, int64_t F2) {
    float inv = 1.0f / (float)(-s.A2);
    float b1 = (float)F1 * inv;
    float b2 = (float)F2 * inv;
    float z = (s.z[0] + b1 * (s.z[1] - s.z[0])) + b2 * (s.z[2] - s.z[0]);
    if (z == 0.0f) z = 0.0f;   // canonicalise -0
    return z;
}

// ---------------------------------------------------------------------------------------------
// PBR fragment path — the reference's LIVE shader: surface_pbr.frag:38-50 -> BRDF_Lighting, pbr.glsl:368-451,
// restated statement for statement with glm's operation order (normalize = v * (1 / sqrt(dot)), clamp = min(max()),
// vec3 products left to right).  Pinned bit-for-bit against oracle/_ref (tests/test_oracle_golden.py).
// ---------------------------------------------------------------------------------------------
struct PbrLight {        // std430 layout of pbr.glsl:288-299 (80 B)
    V3 color, position, direction; float intensity, falloffRadius, cosInnerAngle, cosOuterAngle; uint32_t shadowIndex, type;
};
PbrLight load_pbr_light(const uint8_t* p) {
    PbrLight l;
    std::memcpy(&l.color, p + 0, 12); std::memcpy(&l.position, p + 16, 12); std::memcpy(&l.direction, p + 32, 12);
    std::memcpy(&l.intensity, p + 44, 4); std::memcpy(&l.falloffRadius, p + 48, 4); std::memcpy(&l.cosInnerAngle, p + 52, 4);
    std::memcpy(&l.cosOuterAngle, p + 56, 4); std::memcpy(&l.shadowIndex, p + 60, 4); std::memcpy(&l.type, p + 64, 4);
    return l;
}
struct PbrMaterial {     // MaterialProperties, MaterialSystem.h:41-51 <-> Material, pbr.glsl:78-94 (80 B)
    V4 baseColor, emissive; float f0, metallic, perceptualRoughness, ambientOcclusion; uint32_t tex[5];
};
PbrMaterial load_pbr_material(const uint8_t* p) {
    PbrMaterial m;
    std::memcpy(&m.baseColor, p + 0, 16); std::memcpy(&m.emissive, p + 16, 16); std::memcpy(&m.f0, p + 32, 16); std::memcpy(m.tex, p + 48, 20);
    return m;
}
struct PbrRemapped { V3 normal; V4 baseColor; V3 diffuseColor; V4 emissive; V3 f0; float metallic, roughness, occlusion; };
inline float glm_abs(float x) { return x >= 0.0f ? x : -x; }   // glm/detail/func_common.inl compute_abs (keeps -0)
inline float glm_clamp(float x, float lo, float hi) { return glm_min(glm_max(x, lo), hi); }
inline V3 cross3(V3 x, V3 y) { return {x.y * y.z - y.y * x.z, x.z * y.x - y.z * x.x, x.x * y.y - y.x * x.y}; }   // glm/detail/func_geometric.inl:130-134
inline V3 scale3(V3 v, float s) { return {v.x * s, v.y * s, v.z * s}; }
constexpr float PBR_PI = 3.14159265359f;

// GetNormal, pbr.glsl:125-148; tn = texture(normals).xyz when hasMap; derivatives of fragPos / uv in screen x and y
V3 pbr_get_normal(bool hasMap, V3 tn, V3 q1, V3 q2, const float st1[2], const float st2[2], V3 fragNormal) {
    if (!hasMap) return fragNormal;
    V3 tangentNormal = {tn.x * 2.0f - 1.0f, tn.y * 2.0f - 1.0f, tn.z * 2.0f - 1.0f};
    V3 N = normalize3(fragNormal);
    V3 T = normalize3(sub3(scale3(q1, st2[1]), scale3(q2, st1[1])));
    V3 B = neg3(normalize3(cross3(N, T)));
    V3 r = {(T.x * tangentNormal.x + B.x * tangentNormal.y) + N.x * tangentNormal.z,     // mat3(T, B, N) * v
            (T.y * tangentNormal.x + B.y * tangentNormal.y) + N.y * tangentNormal.z,
            (T.z * tangentNormal.x + B.z * tangentNormal.y) + N.z * tangentNormal.z};
    return normalize3(r);
}

// BRDF_Lighting, pbr.glsl:368-451, after the texture fetches (which pass their results in `m`)
V4 pbr_lighting(const PbrRemapped& material, const std::vector<PbrLight>& lights, uint32_t lightCount, V3 fragPos, V3 viewPosition,
                float exposure, uint32_t debugInput, uint32_t debugEquation, bool haveShadow, const float* STR, const float* shadowMap, uint32_t S) {
    const V3 n = material.normal;
    const V3 v = normalize3(sub3(viewPosition, fragPos));
    V3 Fd = {0, 0, 0}, F = {0, 0, 0}, Fr = {0, 0, 0}; float D = 0.0f, G = 0.0f, NoV = 0.0f, NoL = 0.0f;   // brdf / pbr of the LAST light (debug views)
    V3 l0 = {0.0f, 0.0f, 0.0f};
    for (uint32_t li = 0; li < lightCount; ++li) {
        const PbrLight& light = lights[li];
        V3 l = normalize3(sub3(light.position, fragPos));                       // GetLightDirection, pbr.glsl:303-311
        if (light.type == LIGHT_TYPE_DIRECTIONAL) l = normalize3(neg3(light.direction));
        const V3 h = normalize3(add3(v, l));
        NoV = glm_abs(dot3(n, v)) + 1e-5f;
        const float NoH = glm_clamp(dot3(n, h), 0.0f, 1.0f);
        NoL = glm_clamp(dot3(n, l), 0.0f, 1.0f);
        const float HoL = glm_clamp(dot3(h, l), 0.0f, 1.0f);
        {   // BRDF, pbr.glsl:262-279
            const float a = material.roughness;
            { const float a2 = a * a; const float f = (NoH * a2 - NoH) * NoH + 1.0f; D = a2 / (PBR_PI * f * f); }   // D_GGX
            { const float p5 = std::pow((1.0f - HoL), 5.0f);                                                          // F_Schlick
              F = {material.f0.x + (1.0f - material.f0.x) * p5, material.f0.y + (1.0f - material.f0.y) * p5, material.f0.z + (1.0f - material.f0.z) * p5}; }
            { const float a2 = a * a;                                                                                 // G_SmithGGXCorrelated
              const float GGXL = NoV * std::sqrt((-NoL * a2 + NoL) * NoL + a2);
              const float GGXV = NoL * std::sqrt((-NoV * a2 + NoV) * NoV + a2);
              G = 0.5f / (GGXV + GGXL); }
            const float DG = D * G;
            Fr = {DG * F.x, DG * F.y, DG * F.z};
            const float lambert = 1.0f / PBR_PI;
            Fd = {(((1.0f - F.x) * material.occlusion) * material.diffuseColor.x) * lambert, (((1.0f - F.y) * material.occlusion) * material.diffuseColor.y) * lambert,
                  (((1.0f - F.z) * material.occlusion) * material.diffuseColor.z) * lambert};
        }
        const V3 FrFd = add3(Fr, Fd);
        float attenuation = 1.0f;                                               // EvaluatePunctualLight, pbr.glsl:335-352
        if (light.type == LIGHT_TYPE_POINT || light.type == LIGHT_TYPE_SPOT) {
            {   // ComputeSquareFalloff(pbr.l, falloffRadius): l is the NORMALISED direction here, as in the reference
                const float inverseFalloffRadius = 1.0f / glm_max(light.falloffRadius, 0.083f);
                const float distanceSquare = dot3(l, l);
                const float factor = distanceSquare * inverseFalloffRadius * inverseFalloffRadius;
                const float smoothFactor = glm_max(1.0f - factor * factor, 0.0f);
                attenuation = (smoothFactor * smoothFactor) / glm_max(distanceSquare, 1e-4f);
            }
            if (light.type == LIGHT_TYPE_SPOT) {
                const float spotScale = 1.0f / glm_max(light.cosInnerAngle - light.cosOuterAngle, 1e-4f);
                const float spotOffset = -light.cosOuterAngle * spotScale;
                const float cd = dot3(normalize3(neg3(light.direction)), l);
                const float att = glm_clamp(cd * spotScale + spotOffset, 0.0f, 1.0f);
                attenuation = attenuation * (att * att);
            }
        }
        const V3 lightResult = {(((FrFd.x * light.intensity) * attenuation) * NoL) * light.color.x, (((FrFd.y * light.intensity) * attenuation) * NoL) * light.color.y,
                                (((FrFd.z * light.intensity) * attenuation) * NoL) * light.color.z};
        float shadow = 0.0f;
        if (light.type == LIGHT_TYPE_DIRECTIONAL && haveShadow && light.shadowIndex == 0) {
            Light sl; sl.type = (int32_t)light.type; sl.position = light.position;   // ComputeShadow only reads type / position / shadowIndex
            shadow = compute_shadow(sl, fragPos, n, STR, shadowMap, S);
        }
        const float k = 1.0f - shadow;
        l0 = {l0.x + k * lightResult.x, l0.y + k * lightResult.y, l0.z + k * lightResult.z};
    }
    switch (debugInput) {                                                       // pbr.glsl:409-425
        case 1: return material.baseColor;
        case 2: return {material.diffuseColor.x, material.diffuseColor.y, material.diffuseColor.z, 1.0f};
        case 3: return {material.normal.x, material.normal.y, material.normal.z, 1.0f};
        case 4: return {material.occlusion, material.occlusion, material.occlusion, 1.0f};
        case 5: return {material.emissive.x * material.emissive.w, material.emissive.y * material.emissive.w, material.emissive.z * material.emissive.w, 1.0f};
        case 6: return {material.metallic, material.metallic, material.metallic, 1.0f};
        case 7: return {material.roughness, material.roughness, material.roughness, 1.0f};
        default: break;
    }
    switch (debugEquation) {                                                    // pbr.glsl:427-439
        case 1: return {Fd.x, Fd.y, Fd.z, 1.0f};
        case 2: return {F.x, F.y, F.z, 1.0f};
        case 3: { const float g = (G * 4.0f) * (glm_abs(NoL) * glm_abs(NoV)); return {g, g, g, 1.0f}; }
        case 4: return {D, D, D, 1.0f};
        case 5: return {Fr.x, Fr.y, Fr.z, 1.0f};
        default: break;
    }
    V3 color = l0;
    color = {color.x + material.emissive.x * material.emissive.w, color.y + material.emissive.y * material.emissive.w, color.z + material.emissive.z * material.emissive.w};
    color = scale3(color, exposure);
    const V3 ambient = {(0.03f * material.baseColor.x) * material.occlusion, (0.03f * material.baseColor.y) * material.occlusion, (0.03f * material.baseColor.z) * material.occlusion};
    color = add3(color, ambient);
    return {color.x, color.y, color.z, material.baseColor.w};
}

// RemapMaterial, pbr.glsl:186-206, from the already-fetched texture values (each defaults to "no texture")
PbrRemapped pbr_remap(const PbrMaterial& m, V3 normal, const V4* baseTex, const V4* emissiveTex, const V4* ormTex) {
    PbrRemapped r;
    r.normal = normal;
    r.baseColor = m.baseColor;
    if (baseTex) r.baseColor = {m.baseColor.x * baseTex->x, m.baseColor.y * baseTex->y, m.baseColor.z * baseTex->z, m.baseColor.w * baseTex->w};
    r.emissive = m.emissive;
    if (emissiveTex) r.emissive = {m.emissive.x * emissiveTex->x, m.emissive.y * emissiveTex->y, m.emissive.z * emissiveTex->z, m.emissive.w * emissiveTex->w};
    V3 orm = {m.ambientOcclusion, m.perceptualRoughness, m.metallic};
    if (ormTex) orm = {m.ambientOcclusion * ormTex->x, m.perceptualRoughness * ormTex->y, m.metallic * ormTex->z};
    r.occlusion = orm.x;
    { const float rr = glm_clamp(orm.y, 0.089f, 1.0f); r.roughness = rr * rr; }
    r.metallic = orm.z;
    r.diffuseColor = {(1.0f - r.metallic) * r.baseColor.x, (1.0f - r.metallic) * r.baseColor.y, (1.0f - r.metallic) * r.baseColor.z};
    const float f0s = m.f0 * (1.0f - r.metallic);
    r.f0 = {f0s + r.baseColor.x * r.metallic, f0s + r.baseColor.y * r.metallic, f0s + r.baseColor.z * r.metallic};
    return r;
}

struct DrawRec { uint32_t indexOffset, nbIndices, materialRaw, sceneNode; };

struct Frontend {
    std::vector<Setup> setups;       // submission order
    std::vector<uint64_t> triPrefix; // per draw: first triId
    uint64_t needsClip = 0;
};

unsigned thread_count(const orc_scene* s) {
    unsigned n = s->nThreads ? s->nThreads : std::thread::hardware_concurrency();
    return n ? n : 1;
}

template <class F> void parallel_for(unsigned nThreads, F f) {
    if (nThreads <= 1) { f(0u); return; }
    std::vector<std::thread> th;
    for (unsigned t = 0; t < nThreads; ++t) th.emplace_back(f, t);
    for (auto& x : th) x.join();
}

// drawList / nDrawsIn default to the scene's opaque list; idBase is added to the running primitive counter (the translucent
// list continues the numbering of the opaque one)
void run_frontend(const orc_scene* sc, const PassParams& pp, Frontend& fe, const uint32_t* drawList = nullptr, uint32_t nDrawsIn = 0, uint64_t idBase = 0) {
    const DrawRec* draws = reinterpret_cast<const DrawRec*>(drawList ? drawList : sc->draws);
    const uint32_t nDraws = drawList ? nDrawsIn : sc->nDraws;
    fe.triPrefix.resize(nDraws + 1);
    uint64_t total = 0;
    for (uint32_t d = 0; d < nDraws; ++d) { fe.triPrefix[d] = total; total += draws[d].nbIndices / 3; }
    fe.triPrefix[nDraws] = total;
    unsigned nT = thread_count(sc);
    std::vector<std::vector<Setup>> parts(nT);
    std::vector<uint64_t> clip(nT, 0);
    parallel_for(nT, [&](unsigned t) {
        uint64_t lo = total * t / nT, hi = total * (t + 1) / nT;
        uint32_t d = 0;
        for (uint64_t tri = lo; tri < hi; ++tri) {
            while (fe.triPrefix[d + 1] <= tri) ++d;
            const DrawRec& dr = draws[d];
            uint64_t local = tri - fe.triPrefix[d];
            const uint32_t* ix = sc->indices + dr.indexOffset + 3 * local;
            const float* T = sc->transforms + 16 * (size_t)dr.sceneNode;
            V4 c[3];
            for (int k = 0; k < 3; ++k) c[k] = vertex_stage(pp, T, sc->verts + 8 * (size_t)ix[k], nullptr);
            Setup s;
            SetupResult r = setup_triangle(pp, c, (uint32_t)(tri + idBase), s);
            if (r == KEEP) parts[t].push_back(s);
            else if (r == NEEDS_CLIP) {   // B.3: clip, fan, set the children up like any triangle (same triId, parent's place in the order)
                V4 poly[9];
                const int n = clip_polygon(pp, c, poly);
                for (int i = 1; i + 1 < n; ++i) {
                    const V4 cc[3] = {poly[0], poly[i], poly[i + 1]};
                    Setup cs;
                    SetupResult cr = setup_triangle(pp, cc, (uint32_t)(tri + idBase), cs, /*skipTrivialReject=*/true);
                    if (cr == KEEP) parts[t].push_back(cs);
                    else if (cr == NEEDS_CLIP) ++clip[t];   // cannot happen for finite inputs; counted, dropped
                }
            }
        }
    });
    size_t n = 0;
    for (auto& p : parts) n += p.size();
    fe.setups.reserve(n);
    for (auto& p : parts) fe.setups.insert(fe.setups.end(), p.begin(), p.end());
    for (auto c : clip) fe.needsClip += c;
}

inline bool row_owned(const orc_scene* sc, int y) {
    if (sc->nRanks <= 1) return true;
    return ((uint32_t)y / sc->bandH) % sc->nRanks == sc->rank;
}

// In-order depth test "z <= stored" (LessOrEqual, GraphicsPipelineCache.cpp:330-340): processing
// primitives in submission order makes the later primitive win ties, which is what the CUDA path's
// (depthBits<<32 | ~triId) minimum key encodes (Appendix B.9).
void rasterize(const orc_scene* sc, const PassParams& pp, const Frontend& fe, float* depth, uint32_t* triId,
               bool honourRanks, uint64_t* fragCount) {
    unsigned nT = thread_count(sc);
    std::vector<uint64_t> frags(nT, 0);
    const int H = (int)pp.H, W = (int)pp.W;
    parallel_for(nT, [&](unsigned t) {
        int r0 = (int)((int64_t)H * t / nT), r1 = (int)((int64_t)H * (t + 1) / nT);   // rows [r0,r1)
        for (int y = r0; y < r1; ++y)
            for (int x = 0; x < W; ++x) { depth[(size_t)y * W + x] = 1.0f; if (triId) triId[(size_t)y * W + x] = TRI_NONE; }
        for (const Setup& s : fe.setups) {
            int ya = std::max(s.y0, r0), yb = std::min(s.y1, r1 - 1);
            if (ya > yb) continue;
            Edges e = make_edges(s);
            for (int y = ya; y <= yb; ++y) {
                if (honourRanks && !row_owned(sc, y)) continue;
                int64_t Py = 256 * (int64_t)y + 128;
                for (int x = s.x0; x <= s.x1; ++x) {
                    int64_t Px = 256 * (int64_t)x + 128;
                    int64_t F0 = e.A[0] * Px + e.B[0] * Py + e.C[0];
                    int64_t F1 = e.A[1] * Px + e.B[1] * Py + e.C[1];
                    int64_t F2 = e.A[2] * Px + e.B[2] * Py + e.C[2];
                    if (F0 - e.bias[0] < 0 || F1 - e.bias[1] < 0 || F2 - e.bias[2] < 0) continue;
                    float z = depth_at(s, F1, F2);
                    if (!(z >= 0.0f && z <= 1.0f)) continue;   // B.8 depth clip
                    ++frags[t];
                    size_t p = (size_t)y * W + x;
                    if (z <= depth[p]) { depth[p] = z; if (triId) triId[p] = s.triId; }
                }
            }
        }
    });
    if (fragCount) for (auto f : frags) *fragCount += f;
}

}  // namespace

// ---------------------------------------------------------------------------------------------
// Skybox (Appendix B.14).  The reference draws a unit cube last with
// gl_Position = (proj * mat4(mat3(view)) * (p, 1)).xyww (skybox.vert:32-37): depth exactly 1.0, test LessOrEqual
// (RenderScene.cpp:226-229), so it shows wherever the depth buffer still holds 1.0; the fragment colour is
// texture(cube, p) (skybox.frag:28-31) and a cube lookup depends only on the direction of p.  The direction of the
// view ray through a pixel centre is p ~ A^-1 (xn, yn, 1) with A = rows x, y, w of proj * mat4(mat3(view))
// restricted to its 3x3 part (exact when the projection has no translation in x, y, w — every perspective matrix).
// A^-1 is formed in double (adjugate / determinant) and rounded to float once per frame.
// ---------------------------------------------------------------------------------------------
struct SkyParams { float Ainv[9]; float hw, hh, rhw, rhh; const uint16_t* faces; uint32_t w, h; };

inline void sky_setup(const float* view, const float* proj, uint32_t W, uint32_t H, SkyParams& sp) {
    // M = proj * mat4(mat3(view)), column-major; A = rows {0, 1, 3} x columns {0, 1, 2}
    double A[3][3];
    const int rows[3] = {0, 1, 3};
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) {
            double acc = 0.0;
            for (int k = 0; k < 3; ++k) acc += (double)proj[4 * k + rows[r]] * (double)view[4 * c + k];
            A[r][c] = acc;
        }
    const double c00 = A[1][1] * A[2][2] - A[1][2] * A[2][1], c01 = A[1][2] * A[2][0] - A[1][0] * A[2][2], c02 = A[1][0] * A[2][1] - A[1][1] * A[2][0];
    const double det = A[0][0] * c00 + A[0][1] * c01 + A[0][2] * c02;
    const double adj[3][3] = {
        {c00, A[0][2] * A[2][1] - A[0][1] * A[2][2], A[0][1] * A[1][2] - A[0][2] * A[1][1]},
        {c01, A[0][0] * A[2][2] - A[0][2] * A[2][0], A[0][2] * A[1][0] - A[0][0] * A[1][2]},
        {c02, A[0][1] * A[2][0] - A[0][0] * A[2][1], A[0][0] * A[1][1] - A[0][1] * A[1][0]}};
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) sp.Ainv[3 * c + r] = (float)(adj[r][c] / det);   // column-major 3x3
    sp.hw = 0.5f * (float)W; sp.hh = 0.5f * (float)H;
    sp.rhw = 1.0f / sp.hw; sp.rhh = 1.0f / sp.hh;
}

// Vulkan cube face selection (spec "Cube Map Face Selection"; ties: z over y over x): face, sc, tc, |ma|
inline int cube_face(float rx, float ry, float rz, float& sc, float& tc, float& ma) {
    const float ax = glm_abs(rx), ay = glm_abs(ry), az = glm_abs(rz);
    if (az >= ax && az >= ay) { ma = az; if (rz >= 0.0f) { sc = rx; tc = -ry; return 4; } sc = -rx; tc = -ry; return 5; }
    if (ay >= ax) { ma = ay; if (ry >= 0.0f) { sc = rx; tc = rz; return 2; } sc = rx; tc = -rz; return 3; }
    ma = ax;
    if (rx >= 0.0f) { sc = -rz; tc = -ry; return 0; }
    sc = rz; tc = -ry; return 1;
}
inline V4 cube_fetch(const uint16_t* faces, uint32_t w, uint32_t h, int f, int i, int j) {
    const uint16_t* t = faces + (((size_t)f * h + (size_t)j) * w + (size_t)i) * 4;
    const float k = 1.0f / 65535.0f;
    return {(float)t[0] * k, (float)t[1] * k, (float)t[2] * k, (float)t[3] * k};
}
// texel (i, j) of face f; coordinates outside the face are carried over the edge (seamless filtering): the texel
// centre is lifted to its 3-D position on the extended face plane, the face re-selected, the nearest texel taken
inline V4 cube_texel(const uint16_t* faces, uint32_t w, uint32_t h, int f, int i, int j) {
    if (i >= 0 && j >= 0 && i < (int)w && j < (int)h) return cube_fetch(faces, w, h, f, i, j);
    const float sc = (2.0f * ((float)i + 0.5f)) / (float)w - 1.0f, tc = (2.0f * ((float)j + 0.5f)) / (float)h - 1.0f;
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
    const float s = 0.5f * (s2 / ma) + 0.5f, t = 0.5f * (t2 / ma) + 0.5f;
    int i2 = (int)std::floor(s * (float)w), j2 = (int)std::floor(t * (float)h);
    i2 = i2 < 0 ? 0 : (i2 > (int)w - 1 ? (int)w - 1 : i2);
    j2 = j2 < 0 ? 0 : (j2 > (int)h - 1 ? (int)h - 1 : j2);
    return cube_fetch(faces, w, h, f2, i2, j2);
}
inline V4 sky_lookup(const uint16_t* faces, uint32_t w, uint32_t h, float rx, float ry, float rz) {
    float sc, tc, ma;
    const int f = cube_face(rx, ry, rz, sc, tc, ma);
    const float s = 0.5f * (sc / ma) + 0.5f, t = 0.5f * (tc / ma) + 0.5f;
    const float u = s * (float)w - 0.5f, v = t * (float)h - 0.5f;
    const float fu0 = std::floor(u), fv0 = std::floor(v);
    const int i0 = (int)fu0, j0 = (int)fv0;
    const float fu = u - fu0, fv = v - fv0;
    const V4 t00 = cube_texel(faces, w, h, f, i0, j0), t10 = cube_texel(faces, w, h, f, i0 + 1, j0);
    const V4 t01 = cube_texel(faces, w, h, f, i0, j0 + 1), t11 = cube_texel(faces, w, h, f, i0 + 1, j0 + 1);
    const float gu = 1.0f - fu, gv = 1.0f - fv;
    auto mix = [&](float a, float b, float c, float d) { return (a * gu + b * fu) * gv + (c * gu + d * fu) * fv; };
    return {mix(t00.x, t10.x, t01.x, t11.x), mix(t00.y, t10.y, t01.y, t11.y), mix(t00.z, t10.z, t01.z, t11.z), mix(t00.w, t10.w, t01.w, t11.w)};
}
inline V4 sky_pixel(const SkyParams& sp, int x, int y) {
    const float xn = (((float)x + 0.5f) - sp.hw) * sp.rhw, yn = (((float)y + 0.5f) - sp.hh) * sp.rhh;
    const float* a = sp.Ainv;
    const float rx = (a[0] * xn + a[3] * yn) + a[6], ry = (a[1] * xn + a[4] * yn) + a[7], rz = (a[2] * xn + a[5] * yn) + a[8];
    return sky_lookup(sp.faces, sp.w, sp.h, rx, ry, rz);
}

extern "C" {

void orc_sky_pixel(const uint8_t* view160, uint32_t W, uint32_t H, const uint16_t* faces, uint32_t w, uint32_t h, uint32_t x, uint32_t y, float* out4) {
    const float* view = reinterpret_cast<const float*>(view160);
    SkyParams sp; sky_setup(view, view + 16, W, H, sp); sp.faces = faces; sp.w = w; sp.h = h;
    const V4 c = sky_pixel(sp, (int)x, (int)y);
    out4[0] = c.x; out4[1] = c.y; out4[2] = c.z; out4[3] = c.w;
}
void orc_sky_lookup(const uint16_t* faces, uint32_t w, uint32_t h, const float* d, float* out4) {
    const V4 c = sky_lookup(faces, w, h, d[0], d[1], d[2]);
    out4[0] = c.x; out4[1] = c.y; out4[2] = c.z; out4[3] = c.w;
}
void orc_mat4_mul(const float* a, const float* b, float* out) { mat4_mul_mat4(a, b, out); }
void orc_mat4_mul_vec4(const float* m, const float* v, float* out) {
    V4 r = mat4_mul(m, {v[0], v[1], v[2], v[3]});
    out[0] = r.x; out[1] = r.y; out[2] = r.z; out[3] = r.w;
}
void orc_normal_matrix(const float* mat4, float* out9) { normal_matrix(mat4, out9); }

// The composed vertex stages exactly as the raster front end and the shading stage call them (vertex_stage + the per-node
// normal matrix), for pinning against the reference's shaders run through glm (tests/golden: vs_primitive / vs_shadow).
// primitive.vert:49-57: out10 = clip(4) world(3) normal(3); shadow_map.vert:48-52: out4 = clip.
void orc_primitive_vert(const float* transform, const float* view, const float* proj, const float* pos, const float* nrm, float* out10) {
    float PV[16];
    mat4_mul_mat4(proj, view, PV);
    PassParams pp{}; pp.PV = PV; pp.divideWorldByW = true; pp.W = pp.H = 1;
    V3 world;
    const V4 c = vertex_stage(pp, transform, pos, &world);
    float N[9];
    normal_matrix(transform, N);
    const V3 n = normalize3(mat3_mul(N, {nrm[0], nrm[1], nrm[2]}));
    out10[0] = c.x; out10[1] = c.y; out10[2] = c.z; out10[3] = c.w;
    out10[4] = world.x; out10[5] = world.y; out10[6] = world.z; out10[7] = n.x; out10[8] = n.y; out10[9] = n.z;
}
void orc_shadow_vert(const float* transform, const float* view, const float* proj, const float* pos, float* out4) {
    float PV[16];
    mat4_mul_mat4(proj, view, PV);
    PassParams pp{}; pp.PV = PV; pp.divideWorldByW = false; pp.W = pp.H = 1;
    const V4 c = vertex_stage(pp, transform, pos, nullptr);
    out4[0] = c.x; out4[1] = c.y; out4[2] = c.z; out4[3] = c.w;
}

void orc_phong_lighting(const uint8_t* light112, const float* p, const float* normal, const float* fragPos,
                        const float* viewPos, float shadow, float* out4) {
    Light l = load_light(light112);
    PhongProps props{{p[0], p[1], p[2], p[3]}, {p[4], p[5], p[6], p[7]}, p[8]};
    V4 r = phong_lighting(l, props, {normal[0], normal[1], normal[2]}, {fragPos[0], fragPos[1], fragPos[2]},
                          {viewPos[0], viewPos[1], viewPos[2]}, shadow);
    out4[0] = r.x; out4[1] = r.y; out4[2] = r.z; out4[3] = r.w;
}

float orc_compute_shadow(const uint8_t* light112, const float* fragPos, const float* normal, const float* transform16,
                         const float* shadowMap, uint32_t S) {
    Light l = load_light(light112);
    return compute_shadow(l, {fragPos[0], fragPos[1], fragPos[2]}, {normal[0], normal[1], normal[2]}, transform16, shadowMap, S);
}

void orc_srgb8(const float* rgba, uint8_t* out4) {
    out4[0] = quant8(srgb_oetf(rgba[0]));
    out4[1] = quant8(srgb_oetf(rgba[1]));
    out4[2] = quant8(srgb_oetf(rgba[2]));
    out4[3] = quant8(rgba[3]);
}

// Shadow depth pass: ShadowSystem::Render (ShadowSystem.cpp:319-376) + shadow_map.vert:48-52 with the
// pipeline state of ShadowSystem.cpp:37-53 (1 sample, cull back / CCW front per F5, LessOrEqual, clear 1.0).
int orc_render_shadow(const orc_scene* sc, float* shadowMap, uint64_t* counters) {
    if (!sc || !shadowMap || sc->shadowSize == 0) return -1;
    float SPV[16];
    mat4_mul_mat4(sc->shadowProj, sc->shadowView, SPV);   // shadow.proj * shadow.view (shadow_map.vert:51)
    PassParams pp{SPV, false, sc->shadowSize, sc->shadowSize};
    Frontend fe;
    run_frontend(sc, pp, fe);
    uint64_t frags = 0;
    rasterize(sc, pp, fe, shadowMap, nullptr, false, &frags);
    if (counters) { counters[0] = fe.needsClip; counters[1] = fe.setups.size(); counters[2] = frags; counters[3] = 0; }
    return 0;
}

// Forward pass: MaterialSystem::Draw (MaterialSystem.cpp:88-101) + primitive.vert:49-57 + the
// builder-defined Phong main() of Appendix B.11 over phong.glsl / shadow.glsl.
namespace {
// depth / triId hold nSamples planes of W x H (1 = the 1-sample contract, 4 = Appendix B.15)
int render_forward_impl(const orc_scene* sc, const float* shadowMap, float* depth, uint32_t* triId, float* colorF32,
                        uint8_t* colorSRGB8, uint64_t* counters, int nSamples) {
    if (!sc || !depth || !triId) return -1;
    const float* view = reinterpret_cast<const float*>(sc->view);
    const float* proj = view + 16;
    V3 viewPos = {view[32], view[33], view[34]};
    float PV[16], STR[16];
    mat4_mul_mat4(proj, view, PV);   // GetView().proj * GetView().view (primitive.vert:53)
    const bool haveShadow = sc->shadowSize > 0 && shadowMap != nullptr;
    if (haveShadow) mat4_mul_mat4(sc->shadowProj, sc->shadowView, STR);   // ShadowSystem.cpp:303
    PassParams pp{PV, true, sc->width, sc->height};   // unshifted: shading happens at pixel centres
    Frontend fe;
    uint64_t frags = 0;
    if (nSamples == 1) {
        run_frontend(sc, pp, fe);
        rasterize(sc, pp, fe, depth, triId, true, &frags);
    } else {
        for (int k = 0; k < nSamples; ++k) {
            PassParams ps = pp;
            ps.shiftX = MSAA4_SHIFT[k][0]; ps.shiftY = MSAA4_SHIFT[k][1];
            fe = Frontend();
            run_frontend(sc, ps, fe);
            rasterize(sc, ps, fe, depth + (size_t)k * sc->width * sc->height, triId + (size_t)k * sc->width * sc->height, true, &frags);
        }
    }
    if (counters) { counters[0] = fe.needsClip; counters[1] = fe.setups.size(); counters[2] = frags; counters[3] = 0; }
    if (!colorF32 && !colorSRGB8) return 0;

    const DrawRec* draws = reinterpret_cast<const DrawRec*>(sc->draws);
    std::vector<float> nmat((size_t)sc->nNodes * 9);
    for (uint32_t n = 0; n < sc->nNodes; ++n) normal_matrix(sc->transforms + 16 * (size_t)n, &nmat[9 * (size_t)n]);
    std::vector<Light> lights(sc->nLights);
    std::vector<PbrLight> pbrLights(sc->nLights);
    for (uint32_t l = 0; l < sc->nLights; ++l) {
        if (sc->shadingModel == 1) pbrLights[l] = load_pbr_light(sc->lights + (size_t)l * sc->lightStride);
        else lights[l] = load_light(sc->lights + (size_t)l * sc->lightStride);
    }
    std::vector<Texture> textures;
    for (uint32_t t = 0; t < sc->nTextures; ++t) textures.push_back(build_texture(sc->textures[t]));

    const bool haveSky = sc->skyFaces != nullptr && sc->skyW && sc->skyH;
    SkyParams sky{};
    if (haveSky) { sky_setup(view, proj, sc->width, sc->height, sky); sky.faces = sc->skyFaces; sky.w = sc->skyW; sky.h = sc->skyH; }

    // colour of triangle `tri` shaded at the centre of pixel (x, y) (the fragment shader runs once per pixel and primitive)
    // (tri is relative to the draw list being shaded: curDraws / curPrefix switch to the translucent list for Appendix B.16)
    const DrawRec* curDraws = draws;
    const std::vector<uint64_t>* curPrefix = &fe.triPrefix;
    auto shadeTri = [&](int x, int y, uint32_t tri) -> V4 {
        const std::vector<uint64_t>& prefix = *curPrefix;
        uint32_t d = (uint32_t)(std::upper_bound(prefix.begin(), prefix.end(), (uint64_t)tri) - prefix.begin()) - 1;
        const DrawRec& dr = curDraws[d];
        const uint32_t* ix = sc->indices + dr.indexOffset + 3 * ((uint64_t)tri - prefix[d]);
        const float* T = sc->transforms + 16 * (size_t)dr.sceneNode;
        const float* N = &nmat[9 * (size_t)dr.sceneNode];
        V4 c[3]; V3 world[3], nrm[3]; float uv[3][2];
        for (int k = 0; k < 3; ++k) {
            const float* v = sc->verts + 8 * (size_t)ix[k];
            c[k] = vertex_stage(pp, T, v, &world[k]);
            nrm[k] = normalize3(mat3_mul(N, {v[3], v[4], v[5]}));   // primitive.vert:55
            uv[k][0] = v[6]; uv[k][1] = v[7];
        }
        Setup s;
        const bool clippedParent = setup_triangle(pp, c, tri, s) == NEEDS_CLIP;   // B.3: the pixel was won by one of its clipped children
        Edges e{}; float inv = 0.0f;
        if (!clippedParent) { e = make_edges(s); inv = 1.0f / (float)(-s.A2); }
        const int64_t Px = 256 * (int64_t)x + 128, Py = 256 * (int64_t)y + 128;
        // B.10 perspective-correct weights at pixel (x + ox, y + oy): from the snapped vertices, or in homogeneous form for a clipped parent
        auto weights = [&](int ox, int oy, float qq[3]) {
            if (clippedParent) { homogeneous_weights(pp, c, (float)(x + ox), (float)(y + oy), qq); return; }
            const int64_t Qx = Px + 256 * ox, Qy = Py + 256 * oy;
            for (int k = 0; k < 3; ++k) {
                int64_t F = e.A[k] * Qx + e.B[k] * Qy + e.C[k];
                qq[k] = ((float)F * inv) * (1.0f / s.w[k]);
            }
        };
        float q[3];
        weights(0, 0, q);
        float den = (q[0] + q[1]) + q[2];
        float rden = 1.0f / den;             // r1b: one reciprocal per fragment, then multiplies
        auto interp = [&](float a0, float a1, float a2) { return ((q[0] * a0 + q[1] * a1) + q[2] * a2) * rden; };
        V3 fragPos = {interp(world[0].x, world[1].x, world[2].x), interp(world[0].y, world[1].y, world[2].y),
                      interp(world[0].z, world[1].z, world[2].z)};
        V3 normal = {interp(nrm[0].x, nrm[1].x, nrm[2].x), interp(nrm[0].y, nrm[1].y, nrm[2].y),
                     interp(nrm[0].z, nrm[1].z, nrm[2].z)};
        uint32_t matIndex = dr.materialRaw >> 4;   // MaterialHandle::GetIndex (MaterialDefines.h:88-96)
        const uint8_t* mr = sc->materials + (size_t)matIndex * sc->materialStride;
        if (sc->shadingModel == 1) {   // ---- reference's live PBR fragment path (surface_pbr.frag:38-50) ----
            const PbrMaterial pm = load_pbr_material(mr);
            auto live = [&](uint32_t h) { return h < sc->maxDescriptors && h < textures.size(); };
            const bool anyTex = live(pm.tex[0]) || live(pm.tex[1]) || live(pm.tex[2]) || live(pm.tex[3]);
            V3 nrmOut = normal;
            V4 tb, te, to; const V4 *pb = nullptr, *pe = nullptr, *po = nullptr;
            if (anyTex) {
                // attributes here and at the +x / +y neighbour pixel centres of the same primitive (implicit derivatives)
                float tuv[3][2]; V3 tpos[3];
                for (int n = 0; n < 3; ++n) {
                    float qq[3];
                    weights(n == 1 ? 1 : 0, n == 2 ? 1 : 0, qq);
                    float rdd = 1.0f / ((qq[0] + qq[1]) + qq[2]);
                    auto ip = [&](float a0, float a1, float a2) { return ((qq[0] * a0 + qq[1] * a1) + qq[2] * a2) * rdd; };
                    tuv[n][0] = ip(uv[0][0], uv[1][0], uv[2][0]); tuv[n][1] = ip(uv[0][1], uv[1][1], uv[2][1]);
                    tpos[n] = {ip(world[0].x, world[1].x, world[2].x), ip(world[0].y, world[1].y, world[2].y), ip(world[0].z, world[1].z, world[2].z)};
                }
                const float st1[2] = {tuv[1][0] - tuv[0][0], tuv[1][1] - tuv[0][1]}, st2[2] = {tuv[2][0] - tuv[0][0], tuv[2][1] - tuv[0][1]};
                auto samp = [&](uint32_t h) { return sample_texture(textures[h], tuv[0][0], tuv[0][1], st1[0], st1[1], st2[0], st2[1]); };
                if (live(pm.tex[3])) { V4 t = samp(pm.tex[3]); nrmOut = pbr_get_normal(true, {t.x, t.y, t.z}, sub3(tpos[1], tpos[0]), sub3(tpos[2], tpos[0]), st1, st2, normal); }
                if (live(pm.tex[0])) { tb = samp(pm.tex[0]); pb = &tb; }
                if (live(pm.tex[1])) { te = samp(pm.tex[1]); pe = &te; }
                if (live(pm.tex[2])) { to = samp(pm.tex[2]); po = &to; }
            }
            const PbrRemapped rm = pbr_remap(pm, nrmOut, pb, pe, po);
            float exposure; uint32_t dbgIn, dbgEq;
            std::memcpy(&exposure, sc->view + 140, 4); std::memcpy(&dbgIn, sc->view + 144, 4); std::memcpy(&dbgEq, sc->view + 148, 4);
            return pbr_lighting(rm, pbrLights, sc->nLights, fragPos, viewPos, exposure, dbgIn, dbgEq, haveShadow, STR, shadowMap, sc->shadowSize);
        }
        PhongProps props;
        uint32_t texHandle[2];
        std::memcpy(&props.diffuse, mr + 0, 16);
        std::memcpy(&props.specular, mr + 16, 16);
        std::memcpy(&props.shininess, mr + 32, 4);
        std::memcpy(texHandle, mr + 36, 8);          // diffuseTex, specularTex (phong.glsl:8-12)

        // bindless texture lookup (pattern of pbr.glsl:100-108): a handle is live iff < MAX_DESCRIPTOR_COUNT
        const bool useTex[2] = {texHandle[0] < sc->maxDescriptors && texHandle[0] < textures.size(),
                                texHandle[1] < sc->maxDescriptors && texHandle[1] < textures.size()};
        if (useTex[0] || useTex[1]) {
            // uv here and at the +x / +y neighbour pixel centres of the same primitive (implicit derivatives)
            float tuv[3][2];
            for (int n = 0; n < 3; ++n) {
                float qq[3];
                weights(n == 1 ? 1 : 0, n == 2 ? 1 : 0, qq);
                float rdd = 1.0f / ((qq[0] + qq[1]) + qq[2]);
                tuv[n][0] = ((qq[0] * uv[0][0] + qq[1] * uv[1][0]) + qq[2] * uv[2][0]) * rdd;
                tuv[n][1] = ((qq[0] * uv[0][1] + qq[1] * uv[1][1]) + qq[2] * uv[2][1]) * rdd;
            }
            const float dudx = tuv[1][0] - tuv[0][0], dvdx = tuv[1][1] - tuv[0][1];
            const float dudy = tuv[2][0] - tuv[0][0], dvdy = tuv[2][1] - tuv[0][1];
            if (useTex[0]) {
                V4 tcol = sample_texture(textures[texHandle[0]], tuv[0][0], tuv[0][1], dudx, dvdx, dudy, dvdy);
                props.diffuse = {props.diffuse.x * tcol.x, props.diffuse.y * tcol.y, props.diffuse.z * tcol.z, props.diffuse.w * tcol.w};
            }
            if (useTex[1]) {
                V4 tcol = sample_texture(textures[texHandle[1]], tuv[0][0], tuv[0][1], dudx, dvdx, dudy, dvdy);
                props.specular = {props.specular.x * tcol.x, props.specular.y * tcol.y, props.specular.z * tcol.z, props.specular.w * tcol.w};
            }
        }

        V4 acc = {0.0f, 0.0f, 0.0f, 0.0f};
        for (uint32_t l = 0; l < sc->nLights; ++l) {
            const Light& L = lights[l];
            float shadow = 0.0f;
            if (haveShadow && (uint32_t)L.type == LIGHT_TYPE_DIRECTIONAL && L.shadowIndex == 0)
                shadow = compute_shadow(L, fragPos, normal, STR, shadowMap, sc->shadowSize);
            V4 cl = phong_lighting(L, props, normal, fragPos, viewPos, shadow);
            acc = {acc.x + cl.x, acc.y + cl.y, acc.z + cl.z, acc.w + cl.w};
        }
        return {acc.x, acc.y, acc.z, props.diffuse.w};
    };
    const V4 clearColour = {0.0f, 0.0f, 0.0f, 1.0f};   // RenderScene.cpp:217-220

    unsigned nT = thread_count(sc);
    const int W = (int)sc->width, H = (int)sc->height;
    const size_t plane = (size_t)W * H;
    // translucent draws blend over the opaque frame in linear float and are drawn BEFORE the skybox (RenderScene.cpp:226-228)
    const bool haveTranslucent = sc->translucentDraws && sc->nTranslucentDraws;
    std::vector<float> hdrTmp;
    if (haveTranslucent && !colorF32) { hdrTmp.resize(plane * 4); colorF32 = hdrTmp.data(); }
    // Appendix B.15 + B.16 together: the multisampled target keeps one colour per SAMPLE until the resolve at the end of the pass
    std::vector<float> colS;
    if (haveTranslucent && nSamples == 4) colS.resize(4 * plane * 4);
    const bool skyNow = haveSky && !haveTranslucent;
    parallel_for(nT, [&](unsigned t) {
        int r0 = (int)((int64_t)H * t / nT), r1 = (int)((int64_t)H * (t + 1) / nT);
        for (int y = r0; y < r1; ++y)
            for (int x = 0; x < W; ++x) {
                size_t p = (size_t)y * W + x;
                V4 out;
                if (nSamples == 1) {
                    // the skybox is drawn last at depth 1.0 under LessOrEqual (Appendix B.14)
                    if (skyNow && depth[p] == 1.0f) out = sky_pixel(sky, x, y);
                    else if (triId[p] != TRI_NONE) out = shadeTri(x, y, triId[p]);
                    else out = clearColour;
                } else {
                    // Appendix B.15: every sample takes the colour its winning primitive was shaded with at the pixel centre
                    // (one shader invocation per pixel and primitive); the resolve is the average in linear float, (s0 + s1) + (s2 + s3)
                    V4 col[4];
                    for (int k = 0; k < 4; ++k) {
                        const uint32_t tri = triId[k * plane + p];
                        const bool skyHere = skyNow && depth[k * plane + p] == 1.0f;
                        int same = -1;
                        for (int j = 0; j < k && same < 0; ++j) {
                            const bool skyJ = skyNow && depth[j * plane + p] == 1.0f;
                            if (skyJ == skyHere && (skyHere || triId[j * plane + p] == tri)) same = j;
                        }
                        if (same >= 0) col[k] = col[same];
                        else if (skyHere) col[k] = sky_pixel(sky, x, y);
                        else if (tri != TRI_NONE) col[k] = shadeTri(x, y, tri);
                        else col[k] = clearColour;
                        if (!colS.empty()) { float* c = &colS[(k * plane + p) * 4]; c[0] = col[k].x; c[1] = col[k].y; c[2] = col[k].z; c[3] = col[k].w; }
                    }
                    // pairwise sum: a pixel whose four samples agree resolves to exactly the 1-sample colour
                    out = {((col[0].x + col[1].x) + (col[2].x + col[3].x)) * 0.25f, ((col[0].y + col[1].y) + (col[2].y + col[3].y)) * 0.25f,
                           ((col[0].z + col[1].z) + (col[2].z + col[3].z)) * 0.25f, ((col[0].w + col[1].w) + (col[2].w + col[3].w)) * 0.25f};
                }
                if (colorF32) { colorF32[4 * p + 0] = out.x; colorF32[4 * p + 1] = out.y; colorF32[4 * p + 2] = out.z; colorF32[4 * p + 3] = out.w; }
                if (colorSRGB8) orc_srgb8(&out.x, colorSRGB8 + 4 * p);
            }
    });
    if (!haveTranslucent) return 0;

    // ---- Appendix B.16: the translucent list, in order, over the opaque frame --------------------------------------------------
    // Same front end with cull none (a back-facing triangle is rasterised with vertices 1 and 2 swapped); every fragment that
    // passes LessOrEqual against the current depth is shaded (at its pixel centre, attributes in the triangle's own vertex
    // order), blended  C = Cs * As + Cd * (1 - As),  A = As * (1 - As)  (GraphicsPipelineCache.cpp:305-315) in linear float —
    // the target is encoded once at the end — and writes its depth and ID.  The skybox comes after (depth still 1.0).
    // With 4 samples (B.15) coverage and the depth test run per sample and a primitive's one colour per pixel (shaded at the
    // pixel centre) is blended into every sample it passes at: the samples are independent targets, so each one is replayed
    // over its own colour / depth / ID plane; sky per sample, then the average resolve.
    const uint64_t idBase = fe.triPrefix.back();
    curDraws = reinterpret_cast<const DrawRec*>(sc->translucentDraws);
    for (int k = 0; k < nSamples; ++k) {
        PassParams pt = pp; pt.twoSided = true;
        if (nSamples == 4) { pt.shiftX = MSAA4_SHIFT[k][0]; pt.shiftY = MSAA4_SHIFT[k][1]; }
        Frontend ft;
        run_frontend(sc, pt, ft, sc->translucentDraws, sc->nTranslucentDraws, idBase);
        curPrefix = &ft.triPrefix;
        float* colK = nSamples == 4 ? &colS[(size_t)k * plane * 4] : colorF32;
        float* depthK = depth + (size_t)k * plane;
        uint32_t* triK = triId + (size_t)k * plane;
        parallel_for(nT, [&](unsigned t) {
            int r0 = (int)((int64_t)H * t / nT), r1 = (int)((int64_t)H * (t + 1) / nT);
            for (const Setup& s : ft.setups) {
                int ya = std::max(s.y0, r0), yb = std::min(s.y1, r1 - 1);
                if (ya > yb) continue;
                Edges e = make_edges(s);
                for (int y = ya; y <= yb; ++y) {
                    if (!row_owned(sc, y)) continue;
                    int64_t Py = 256 * (int64_t)y + 128;
                    for (int x = s.x0; x <= s.x1; ++x) {
                        int64_t Px = 256 * (int64_t)x + 128;
                        int64_t F0 = e.A[0] * Px + e.B[0] * Py + e.C[0];
                        int64_t F1 = e.A[1] * Px + e.B[1] * Py + e.C[1];
                        int64_t F2 = e.A[2] * Px + e.B[2] * Py + e.C[2];
                        if (F0 - e.bias[0] < 0 || F1 - e.bias[1] < 0 || F2 - e.bias[2] < 0) continue;
                        float z = depth_at(s, F1, F2);
                        if (!(z >= 0.0f && z <= 1.0f)) continue;
                        size_t p = (size_t)y * W + x;
                        if (!(z <= depthK[p])) continue;
                        const V4 src = shadeTri(x, y, (uint32_t)(s.triId - idBase));
                        float* dst = colK + 4 * p;
                        const float a = src.w, ia = 1.0f - a;
                        dst[0] = src.x * a + dst[0] * ia; dst[1] = src.y * a + dst[1] * ia; dst[2] = src.z * a + dst[2] * ia;
                        dst[3] = a * ia;
                        depthK[p] = z; triK[p] = s.triId;
                    }
                }
            }
        });
    }
    parallel_for(nT, [&](unsigned t) {
        int r0 = (int)((int64_t)H * t / nT), r1 = (int)((int64_t)H * (t + 1) / nT);
        for (int y = r0; y < r1; ++y)
            for (int x = 0; x < W; ++x) {
                size_t p = (size_t)y * W + x;
                if (nSamples == 1) {
                    if (haveSky && depth[p] == 1.0f && row_owned(sc, y)) {
                        const V4 c = sky_pixel(sky, x, y);
                        colorF32[4 * p + 0] = c.x; colorF32[4 * p + 1] = c.y; colorF32[4 * p + 2] = c.z; colorF32[4 * p + 3] = c.w;
                    }
                } else {
                    float c4[4][4];
                    bool skyDone = false; V4 skyC{};
                    for (int k = 0; k < 4; ++k) {
                        const float* c = &colS[((size_t)k * plane + p) * 4];
                        if (haveSky && depth[(size_t)k * plane + p] == 1.0f) {
                            if (!skyDone) { skyC = sky_pixel(sky, x, y); skyDone = true; }
                            c4[k][0] = skyC.x; c4[k][1] = skyC.y; c4[k][2] = skyC.z; c4[k][3] = skyC.w;
                        } else { c4[k][0] = c[0]; c4[k][1] = c[1]; c4[k][2] = c[2]; c4[k][3] = c[3]; }
                    }
                    for (int ch = 0; ch < 4; ++ch) colorF32[4 * p + ch] = ((c4[0][ch] + c4[1][ch]) + (c4[2][ch] + c4[3][ch])) * 0.25f;
                }
                if (colorSRGB8) orc_srgb8(colorF32 + 4 * p, colorSRGB8 + 4 * p);
            }
    });
    return 0;
}
}  // namespace

int orc_render_forward(const orc_scene* sc, const float* shadowMap, float* depth, uint32_t* triId, float* colorF32,
                       uint8_t* colorSRGB8, uint64_t* counters) {
    return render_forward_impl(sc, shadowMap, depth, triId, colorF32, colorSRGB8, counters, 1);
}
int orc_render_forward_msaa4(const orc_scene* sc, const float* shadowMap, float* depth4, uint32_t* triId4, float* colorF32,
                             uint8_t* colorSRGB8, uint64_t* counters) {
    return render_forward_impl(sc, shadowMap, depth4, triId4, colorF32, colorSRGB8, counters, 4);
}

// BRDF_Lighting for one untextured 80-B material and nLights 80-B PBR lights (pins pbr_lighting / pbr_remap against oracle/_ref)
void orc_pbr_lighting(const uint8_t* material80, const uint8_t* lights80, uint32_t nLights, const float* fragPos, const float* normal, const float* viewPos,
                      float exposure, uint32_t debugInput, uint32_t debugEquation, const float* transform16, const float* shadowMap, uint32_t S, float* out4) {
    std::vector<PbrLight> L(nLights);
    for (uint32_t i = 0; i < nLights; ++i) L[i] = load_pbr_light(lights80 + 80 * (size_t)i);
    const PbrRemapped rm = pbr_remap(load_pbr_material(material80), {normal[0], normal[1], normal[2]}, nullptr, nullptr, nullptr);
    const float one = 1.0f;
    const V4 r = pbr_lighting(rm, L, nLights, {fragPos[0], fragPos[1], fragPos[2]}, {viewPos[0], viewPos[1], viewPos[2]}, exposure, debugInput, debugEquation,
                              true, transform16, (shadowMap && S) ? shadowMap : &one, (shadowMap && S) ? S : 1u);
    out4[0] = r.x; out4[1] = r.y; out4[2] = r.z; out4[3] = r.w;
}

}  // extern "C"
