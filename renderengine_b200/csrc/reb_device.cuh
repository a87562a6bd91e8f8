// reb_device.cuh — device-side types and exact-arithmetic helpers shared by the raster kernels.
//
// Bit-exactness contract: every fp32 operation on the coverage / depth /
// shadow-lookup path is an individually rounded IEEE op in a fixed order.  This translation unit is
// compiled with --fmad=false (no contraction), default -prec-div/-prec-sqrt (IEEE / and sqrtf).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace reb {

constexpr uint32_t TRI_NONE = 0xFFFFFFFFu;
constexpr unsigned long long VIS_CLEAR = 0xFFFFFFFFFFFFFFFFull;   // loses against every real fragment key
constexpr float GUARD_PX = 1048576.0f;                            // |x_f|,|y_f| <= 2^20 px
constexpr int TILE_W = 32, TILE_H = 8;                            // one CTA (256 threads) per screen tile
constexpr int HUGE_TILES = 256;                                   // bbox covering more tiles than this: broadcast list
constexpr int HUGE_CAP = 64;                                      // capacity of the broadcast list (overflow -> binned path)
constexpr int SMALL_MAX = 8;                                      // bbox <= 8x8 pixel centres: per-thread raster
constexpr uint32_t LIGHT_TYPE_DIRECTIONAL = 1, LIGHT_TYPE_SPOT = 3;   // phong.glsl:16-18

// One vkCmdDrawIndexed of the reference's draw loops, with the running primitive counter
// (Appendix B.9) precomputed on the host.
struct DrawDev { uint32_t indexOffset, firstTri, material, node; };

// A triangle that goes through the binned tile path: snapped vertices + depth + id (48 B, 3 x 16 B).
struct __align__(16) LargeRec {
    int32_t X0, Y0, X1, Y1;
    int32_t X2, Y2; float z0, z1;
    float z2; uint32_t triId; uint32_t pad0, pad1;
};

struct PassConst {
    float PV[16];                 // proj*view (forward) or shadowProj*shadowView (shadow), column-major
    uint32_t W, H;                // target extent of this pass
    uint32_t nTris, nDraws;
    const uint32_t* indices;
    const float4* pos4;           // (x, y, z, u)
    const float* transforms;      // 16 floats per node
    const DrawDev* draws;
    unsigned long long* vis;      // forward: W*H 64-bit keys
    uint32_t* depthBits;          // shadow: S*S depth bits (the shadow map itself)
    LargeRec* large; uint32_t largeCap;
    LargeRec* huge;               // HUGE_CAP records tested by every tile
    uint32_t tilesX, tilesY;
    uint32_t* status;             // [0] largeCount [1] needsClip [2] overflow flags [3] tile-list total [4] hugeCount
    uint32_t rank, nRanks, bandH; // sort-first row bands (forward only)
};

// status[2] bits
constexpr uint32_t OVF_LARGE = 1u, OVF_LIST = 2u;

struct V3 { float x, y, z; };

// glm::mat4 * glm::vec4: (m0*x + m1*y) + (m2*z + m3*w) — ThirdParty/glm/glm/detail/type_mat4x4.inl:562-573
__device__ __forceinline__ float4 mat4_mul_point(const float* __restrict__ m, float x, float y, float z) {
    float4 o;
    o.x = (m[0] * x + m[4] * y) + (m[8] * z + m[12] * 1.0f);
    o.y = (m[1] * x + m[5] * y) + (m[9] * z + m[13] * 1.0f);
    o.z = (m[2] * x + m[6] * y) + (m[10] * z + m[14] * 1.0f);
    o.w = (m[3] * x + m[7] * y) + (m[11] * z + m[15] * 1.0f);
    return o;
}

__device__ __forceinline__ float dot3(V3 a, V3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
__device__ __forceinline__ V3 sub3(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
__device__ __forceinline__ V3 add3(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
__device__ __forceinline__ V3 neg3(V3 a) { return {-a.x, -a.y, -a.z}; }
__device__ __forceinline__ V3 normalize3(V3 v) {
    float inv = 1.0f / sqrtf(dot3(v, v));
    return {v.x * inv, v.y * inv, v.z * inv};
}
__device__ __forceinline__ float glm_max(float x, float y) { return (x < y) ? y : x; }
__device__ __forceinline__ float glm_min(float x, float y) { return (y < x) ? y : x; }

// primitive.vert:50-53 (DIVIDE = true) / shadow_map.vert:49-51 (DIVIDE = false)
template <bool DIVIDE>
__device__ __forceinline__ float4 vertex_stage(const float* __restrict__ PV, const float* __restrict__ T, float px, float py, float pz, V3* world) {
    float4 wH = mat4_mul_point(T, px, py, pz);
    V3 w;
    if (DIVIDE) w = {wH.x / wH.w, wH.y / wH.w, wH.z / wH.w};
    else w = {wH.x, wH.y, wH.z};
    if (world) *world = w;
    return mat4_mul_point(PV, w.x, w.y, w.z);
}

enum SetupResult { SR_KEEP = 0, SR_REJECT = 1, SR_NEEDS_CLIP = 2, SR_CULLED = 3, SR_NO_SAMPLE = 4 };

struct Snapped {
    int32_t X[3], Y[3];
    float z[3], w[3];
    long long A2;
    int32_t x0, x1, y0, y1;
    int32_t spanX, spanY;   // unclamped sub-pixel extents
};

// Appendix B.2-B.6: trivial reject, viewport, snap (8 sub-pixel bits, round-half-even), facing, bbox.
__device__ __forceinline__ SetupResult setup_triangle(const float4 c[3], uint32_t W, uint32_t H, Snapped& s) {
    bool allL = true, allR = true, allT = true, allB = true, allN = true, allF = true;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        allL = allL && (c[k].x < -c[k].w); allR = allR && (c[k].x > c[k].w);
        allT = allT && (c[k].y < -c[k].w); allB = allB && (c[k].y > c[k].w);
        allN = allN && (c[k].z < 0.0f);    allF = allF && (c[k].z > c[k].w);
    }
    if (allL || allR || allT || allB || allN || allF) return SR_REJECT;
#pragma unroll
    for (int k = 0; k < 3; ++k) if (!(c[k].w > 0.0f)) return SR_NEEDS_CLIP;
    const float hw = 0.5f * (float)W, hh = 0.5f * (float)H;
    bool guard = false;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        float nx = c[k].x / c[k].w, ny = c[k].y / c[k].w, nz = c[k].z / c[k].w;
        float xf = hw * nx + hw;
        float yf = hh * ny + hh;
        guard = guard || !(fabsf(xf) <= GUARD_PX) || !(fabsf(yf) <= GUARD_PX);
        s.X[k] = __float2int_rn(xf * 256.0f);
        s.Y[k] = __float2int_rn(yf * 256.0f);
        s.z[k] = nz;
        s.w[k] = c[k].w;
    }
    if (guard) return SR_NEEDS_CLIP;
    s.A2 = (long long)(s.X[1] - s.X[0]) * (long long)(s.Y[2] - s.Y[0]) - (long long)(s.X[2] - s.X[0]) * (long long)(s.Y[1] - s.Y[0]);
    if (s.A2 >= 0) return SR_CULLED;
    int32_t minX = min(s.X[0], min(s.X[1], s.X[2])), maxX = max(s.X[0], max(s.X[1], s.X[2]));
    int32_t minY = min(s.Y[0], min(s.Y[1], s.Y[2])), maxY = max(s.Y[0], max(s.Y[1], s.Y[2]));
    s.spanX = maxX - minX; s.spanY = maxY - minY;
    s.x0 = max((minX + 127) >> 8, 0);
    s.x1 = min((maxX - 128) >> 8, (int32_t)W - 1);
    s.y0 = max((minY + 127) >> 8, 0);
    s.y1 = min((maxY - 128) >> 8, (int32_t)H - 1);
    if (s.x0 > s.x1 || s.y0 > s.y1) return SR_NO_SAMPLE;
    return SR_KEEP;
}

// Appendix B.8: depth from the unbiased edge values of vertices 1 and 2 (exact integers as floats).
__device__ __forceinline__ float depth_from(float fF1, float fF2, float area, float z0, float z1, float z2) {
    float b1 = fF1 / area;
    float b2 = fF2 / area;
    float z = (z0 + b1 * (z1 - z0)) + b2 * (z2 - z0);
    if (z == 0.0f) z = 0.0f;
    return z;
}

// binary search: last d with firstTri[d] <= tri
__device__ __forceinline__ uint32_t find_draw(const DrawDev* __restrict__ draws, uint32_t nDraws, uint32_t tri) {
    uint32_t lo = 0, hi = nDraws;
    while (hi - lo > 1) {
        uint32_t mid = (lo + hi) >> 1;
        if (__ldg(&draws[mid].firstTri) <= tri) lo = mid; else hi = mid;
    }
    return lo;
}

__device__ __forceinline__ bool row_owned(uint32_t y, uint32_t rank, uint32_t nRanks, uint32_t bandH) {
    return nRanks <= 1 || ((y / bandH) % nRanks) == rank;
}

}  // namespace reb
