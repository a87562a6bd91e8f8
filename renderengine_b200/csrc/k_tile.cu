// k_tile.cu — K3: the tile stage.  One CTA (256 threads) per TILE_W x TILE_H screen tile, one thread per
// sample.  The tile's visibility keys live in registers (one per thread), the tile's large-triangle
// list is staged through shared memory as precomputed edge equations, coverage uses the top-left
// rule with 64-bit edge functions and an early-Z key compare, then (forward pass) every sample is
// shaded in place: attribute reconstruction, bindless material fetch, Phong + 3x3 PCF.
//
// Reference: fixed-function raster/depth state GraphicsPipelineCache.cpp:280-340; fragment work
// phong.glsl:39-81 + shadow.glsl:24-56; clears RenderScene.cpp:217-220 / ShadowSystem.cpp:111-127.
#include "reb_shade.cuh"
#include <algorithm>

namespace reb {

constexpr int STAGE = 64;   // large-triangle records staged per round

struct EdgeRec {            // one staged large triangle: F_k(P) = A*Px + B*Py + C (interior positive)
    int32_t A[3], B[3];
    long long C[3];
    int32_t bias[3];
    float area, z0, dz1, dz2;
    uint32_t idKey;
};

__device__ __forceinline__ void stage_record(const LargeRec& r, EdgeRec& e) {
    const int32_t X[3] = {r.X0, r.X1, r.X2}, Y[3] = {r.Y0, r.Y1, r.Y2};
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const int a = (k + 1) % 3, b = (k + 2) % 3;
        const int32_t A = Y[b] - Y[a], B = -(X[b] - X[a]);
        e.A[k] = A; e.B[k] = B;
        e.C[k] = -((long long)A * X[a]) - (long long)B * Y[a];
        e.bias[k] = ((A > 0) || (A == 0 && B > 0)) ? 0 : 1;
    }
    const long long A2 = (long long)(X[1] - X[0]) * (long long)(Y[2] - Y[0]) - (long long)(X[2] - X[0]) * (long long)(Y[1] - Y[0]);
    e.area = (float)(-A2);
    e.z0 = r.z0; e.dz1 = r.z1 - r.z0; e.dz2 = r.z2 - r.z0;
    e.idKey = ~r.triId;
}

// Does the triangle touch any pixel centre of the tile whose first pixel is (tx0, ty0)?  (bbox + edge reject)
__device__ __forceinline__ bool tile_overlap(const LargeRec& r, const EdgeRec& e, int tx0, int ty0) {
    const int32_t minX = min(r.X0, min(r.X1, r.X2)), maxX = max(r.X0, max(r.X1, r.X2));
    const int32_t minY = min(r.Y0, min(r.Y1, r.Y2)), maxY = max(r.Y0, max(r.Y1, r.Y2));
    const long long cx0 = 256ll * tx0 + 128, cx1 = 256ll * (tx0 + TILE_W - 1) + 128;
    const long long cy0 = 256ll * ty0 + 128, cy1 = 256ll * (ty0 + TILE_H - 1) + 128;
    if (maxX < cx0 || minX > cx1 || maxY < cy0 || minY > cy1) return false;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const long long Fmax = (long long)e.A[k] * (e.A[k] > 0 ? cx1 : cx0) + (long long)e.B[k] * (e.B[k] > 0 ? cy1 : cy0) + e.C[k];
        if (Fmax < 0) return false;
    }
    return true;
}

// Merge a list of large triangles into this thread's key (depthBits<<32 | ~triId; minimum wins).
// idx == nullptr: records src[begin..end) are used directly and culled against the tile (broadcast list);
// otherwise src[idx[i]] for i in [begin, end) (binned list, already culled by the binner).
template <bool SHADOW>
__device__ __forceinline__ unsigned long long merge_large(const LargeRec* __restrict__ src, const uint32_t* __restrict__ idx, uint32_t begin,
                                                          uint32_t end, int tx0, int ty0, int x, int y, unsigned long long key,
                                                          EdgeRec* stage, uint32_t* live) {
    const long long Px = 256ll * x + 128, Py = 256ll * y + 128;
    for (uint32_t base = begin; base < end; base += STAGE) {
        const uint32_t n = min((uint32_t)STAGE, end - base);
        __syncthreads();
        if (threadIdx.x < n) {
            const LargeRec r = src[idx ? __ldg(idx + base + threadIdx.x) : base + threadIdx.x];
            stage_record(r, stage[threadIdx.x]);
            live[threadIdx.x] = idx ? 1u : (tile_overlap(r, stage[threadIdx.x], tx0, ty0) ? 1u : 0u);
        }
        __syncthreads();
        for (uint32_t i = 0; i < n; ++i) {
            if (!live[i]) continue;   // CTA-uniform
            const EdgeRec& e = stage[i];
            const long long F0 = (long long)e.A[0] * Px + (long long)e.B[0] * Py + e.C[0];
            const long long F1 = (long long)e.A[1] * Px + (long long)e.B[1] * Py + e.C[1];
            const long long F2 = (long long)e.A[2] * Px + (long long)e.B[2] * Py + e.C[2];
            if (((F0 - e.bias[0]) | (F1 - e.bias[1]) | (F2 - e.bias[2])) < 0) continue;
            const float b1 = (float)F1 / e.area, b2 = (float)F2 / e.area;
            float z = (e.z0 + b1 * e.dz1) + b2 * e.dz2;
            if (z == 0.0f) z = 0.0f;
            if (!(z >= 0.0f && z <= 1.0f)) continue;
            const unsigned long long k = SHADOW ? (unsigned long long)__float_as_uint(z)
                                                : (((unsigned long long)__float_as_uint(z) << 32) | e.idKey);
            if (k < key) key = k;   // early-Z + later-primitive-wins in one compare
        }
    }
    return key;
}

// Shadow pass tile stage (persistent CTAs striding over the tiles): only tiles touched by large triangles do work.
__global__ void __launch_bounds__(256) k_tile_shadow(const __grid_constant__ PassConst pc, uint32_t* __restrict__ tileCount,
                                                     const uint32_t* __restrict__ tileStart, const uint32_t* __restrict__ tileList, uint32_t listCap) {
    __shared__ EdgeRec stage[STAGE];
    __shared__ uint32_t live[STAGE];
    const uint32_t nLarge = pc.status[0], nHuge = min(pc.status[4], (uint32_t)HUGE_CAP);
    if (nLarge == 0 && nHuge == 0) return;
    const uint32_t nTiles = pc.tilesX * pc.tilesY;
    for (uint32_t tile = blockIdx.x; tile < nTiles; tile += gridDim.x) {
        uint32_t begin = 0, end = 0;
        if (nLarge) { begin = min(tileStart[tile], listCap); end = min(tileStart[tile + 1], listCap); }
        const int tx0 = (int)(tile % pc.tilesX) * TILE_W, ty0 = (int)(tile / pc.tilesX) * TILE_H;
        const int x = tx0 + (int)(threadIdx.x & 31u), y = ty0 + (int)(threadIdx.x >> 5);
        const bool inb = x < (int)pc.W && y < (int)pc.H;
        const size_t p = (size_t)y * pc.W + x;
        const unsigned long long cur = inb ? (unsigned long long)pc.depthBits[p] : 0ull;
        unsigned long long key = cur;
        if (nHuge) key = merge_large<true>(pc.huge, nullptr, 0, nHuge, tx0, ty0, x, y, key, stage, live);
        if (begin != end) key = merge_large<true>(pc.large, tileList, begin, end, tx0, ty0, x, y, key, stage, live);
        if (inb && key < cur) pc.depthBits[p] = (uint32_t)key;
        if (nLarge && threadIdx.x == 0) tileCount[tile] = 0;   // re-arm the bin counters for the next pass
    }
}

// Forward pass tile stage: consume (and reset) the visibility keys, merge large triangles, shade, write.
__global__ void __launch_bounds__(256) k_tile_forward(const __grid_constant__ PassConst pc, const __grid_constant__ ShadeConst sh,
                                                      uint32_t* __restrict__ tileCount, const uint32_t* __restrict__ tileStart,
                                                      const uint32_t* __restrict__ tileList, uint32_t listCap) {
    __shared__ EdgeRec stage[STAGE];
    __shared__ uint32_t live[STAGE];
    const uint32_t tile = blockIdx.x;
    const uint32_t nLarge = pc.status[0], nHuge = min(pc.status[4], (uint32_t)HUGE_CAP);
    if (nLarge && threadIdx.x == 0) tileCount[tile] = 0;
    const int tx0 = (int)(tile % pc.tilesX) * TILE_W, ty0 = (int)(tile / pc.tilesX) * TILE_H;
    if (!row_owned((uint32_t)ty0, pc.rank, pc.nRanks, pc.bandH)) return;
    const int x = tx0 + (int)(threadIdx.x & 31u), y = ty0 + (int)(threadIdx.x >> 5);
    const bool inb = x < (int)pc.W && y < (int)pc.H;
    const size_t p = (size_t)y * pc.W + x;
    unsigned long long key = VIS_CLEAR;
    if (inb) {
        key = pc.vis[p];
        if (key != VIS_CLEAR) pc.vis[p] = VIS_CLEAR;   // leave the buffer cleared for the next frame
    }
    if (nHuge) key = merge_large<false>(pc.huge, nullptr, 0, nHuge, tx0, ty0, x, y, key, stage, live);
    if (nLarge) {
        const uint32_t begin = min(tileStart[tile], listCap), end = min(tileStart[tile + 1], listCap);
        if (begin != end) key = merge_large<false>(pc.large, tileList, begin, end, tx0, ty0, x, y, key, stage, live);
    }
    if (!inb) return;

    float depth = 1.0f;                    // depth clear 1.0 (RenderScene.cpp:219)
    uint32_t tri = TRI_NONE;
    float4 col = make_float4(0.0f, 0.0f, 0.0f, 1.0f);   // colour clear (0,0,0,1) (RenderScene.cpp:218)
    if (key != VIS_CLEAR) {
        depth = __uint_as_float((uint32_t)(key >> 32));
        tri = ~(uint32_t)key;
        if (sh.outSRGB8 || sh.outF32) col = shade_pixel(pc, sh, x, y, tri);
    }
    if (sh.outDepth) sh.outDepth[p] = depth;
    if (sh.outTriId) sh.outTriId[p] = tri;
    if (sh.outSRGB8) reinterpret_cast<uint32_t*>(sh.outSRGB8)[p] = pack_srgb8(col);
    if (sh.outF32) reinterpret_cast<float4*>(sh.outF32)[p] = col;
}

void launch_tile_shadow(const PassConst& pc, uint32_t* tileCount, const uint32_t* tileStart, const uint32_t* tileList, uint32_t listCap,
                        int numSMs, cudaStream_t st) {
    const uint32_t nTiles = pc.tilesX * pc.tilesY;
    k_tile_shadow<<<std::min<uint32_t>(nTiles, (uint32_t)numSMs * 8u), 256, 0, st>>>(pc, tileCount, tileStart, tileList, listCap);
}

void launch_tile_forward(const PassConst& pc, const ShadeConst& sh, uint32_t* tileCount, const uint32_t* tileStart, const uint32_t* tileList,
                         uint32_t listCap, cudaStream_t st) {
    k_tile_forward<<<pc.tilesX * pc.tilesY, 256, 0, st>>>(pc, sh, tileCount, tileStart, tileList, listCap);
}

}  // namespace reb
