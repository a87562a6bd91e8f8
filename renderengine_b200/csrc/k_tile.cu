// k_tile.cu — K3: the tile stage.  One CTA (256 threads) per TILE_W x TILE_H screen tile on a 2-D grid, one
// thread per sample, each warp owning an 8 x 4 pixel block (compact footprint: neighbouring samples share
// triangles, so the post-transform-cache gathers of a warp fall into few sectors).  The tile's visibility
// keys live in registers (one per thread); the large triangles that touch the tile (broadcast list + binned
// list) are staged through shared memory as edge equations relative to the tile origin; coverage uses the
// top-left rule with 64-bit edge functions and an early-Z key compare; then (forward pass) every sample is
// shaded in place: attribute reconstruction, bindless material fetch, Phong + 3x3 PCF.
//
// Reference: fixed-function raster/depth state GraphicsPipelineCache.cpp:280-340; fragment work
// phong.glsl:39-81 + shadow.glsl:24-56; clears RenderScene.cpp:217-220 / ShadowSystem.cpp:111-127.
#include "reb_shade.cuh"
#include <algorithm>

namespace reb {

constexpr int STAGE = 64;   // large-triangle records staged per round (two warps' worth)

struct EdgeRec {            // one staged large triangle, relative to the tile's first pixel centre
    long long E[3];         // biased edge value at the tile origin (covered iff E + A*dx + B*dy >= 0 for all three)
    int32_t A[3], B[3];
    float area, z0, dz1, dz2;
    uint32_t idKey, bias12; // bias12: bit 0 = bias of edge 1, bit 1 = bias of edge 2 (to recover the unbiased values)
};

struct TileShared {
    EdgeRec stage[STAGE];
    uint32_t liveMask[2];
};

// Build the staged record of triangle r for the tile whose first pixel is (tx0, ty0); returns false when the
// triangle cannot touch any pixel centre of the tile (bbox + edge reject at the most-interior corner).
template <int ROWS>
__device__ __forceinline__ bool stage_record(const LargeRec& r, int tx0, int ty0, EdgeRec& e) {
    const int32_t X[3] = {r.X0, r.X1, r.X2}, Y[3] = {r.Y0, r.Y1, r.Y2};
    const int32_t minX = min(r.X0, min(r.X1, r.X2)), maxX = max(r.X0, max(r.X1, r.X2));
    const int32_t minY = min(r.Y0, min(r.Y1, r.Y2)), maxY = max(r.Y0, max(r.Y1, r.Y2));
    const long long cx0 = 256ll * tx0 + 128, cy0 = 256ll * ty0 + 128;
    const long long spanX = 256ll * (TILE_W - 1), spanY = 256ll * (ROWS - 1);
    bool live = !(maxX < cx0 || minX > cx0 + spanX || maxY < cy0 || minY > cy0 + spanY);
    uint32_t bias12 = 0;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const int a = (k + 1) % 3, b = (k + 2) % 3;
        const int32_t A = Y[b] - Y[a], B = -(X[b] - X[a]);
        const int32_t bias = ((A > 0) || (A == 0 && B > 0)) ? 0 : 1;   // top-left rule (Appendix B.7)
        const long long E = (long long)A * (cx0 - X[a]) + (long long)B * (cy0 - Y[a]) - bias;
        e.A[k] = A; e.B[k] = B; e.E[k] = E;
        const long long Emax = E + (A > 0 ? (long long)A * spanX : 0ll) + (B > 0 ? (long long)B * spanY : 0ll);
        live = live && (Emax >= 0);
        if (k > 0) bias12 |= (uint32_t)bias << (k - 1);
    }
    const long long A2 = (long long)(X[1] - X[0]) * (long long)(Y[2] - Y[0]) - (long long)(X[2] - X[0]) * (long long)(Y[1] - Y[0]);
    e.area = (float)(-A2);
    e.z0 = r.z0; e.dz1 = r.z1 - r.z0; e.dz2 = r.z2 - r.z0;
    e.idKey = ~r.triId; e.bias12 = bias12;
    return live;
}

// Merge a list of large triangles into this thread's key (depthBits<<32 | ~triId, or the bare depth bits for the
// shadow pass; minimum wins).  idx == nullptr: records src[begin..end) directly (broadcast list); otherwise
// src[idx[i]] for i in [begin, end) (binned list).  (lx, ly) = sample position inside the tile.
template <bool SHADOW>
__device__ __forceinline__ unsigned long long merge_large(const LargeRec* __restrict__ src, const uint32_t* __restrict__ idx, uint32_t begin,
                                                          uint32_t end, int tx0, int ty0, int lx, int ly, unsigned long long key, TileShared& sm) {
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const int32_t dx = 256 * lx, dy = 256 * ly;
    for (uint32_t base = begin; base < end; base += STAGE) {
        const uint32_t n = min((uint32_t)STAGE, end - base);
        __syncthreads();
        if (warp < 2) {
            bool live = false;
            if (threadIdx.x < n) {
                const LargeRec r = src[idx ? __ldcg(idx + base + threadIdx.x) : base + threadIdx.x];
                live = stage_record<TILE_H>(r, tx0, ty0, sm.stage[threadIdx.x]);
            }
            const unsigned m = __ballot_sync(0xFFFFFFFFu, live);
            if (lane == 0) sm.liveMask[warp] = m;
        }
        __syncthreads();
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            unsigned m = sm.liveMask[h];   // CTA-uniform
            while (m) {
                const EdgeRec& e = sm.stage[32 * h + __ffs(m) - 1];
                m &= m - 1;
                const long long F0 = e.E[0] + (long long)e.A[0] * dx + (long long)e.B[0] * dy;
                const long long F1 = e.E[1] + (long long)e.A[1] * dx + (long long)e.B[1] * dy;
                const long long F2 = e.E[2] + (long long)e.A[2] * dx + (long long)e.B[2] * dy;
                if ((F0 | F1 | F2) < 0) continue;
                const float z = depth_from((float)(F1 + (long long)(e.bias12 & 1u)), (float)(F2 + (long long)(e.bias12 >> 1)), e.area, e.z0, e.dz1, e.dz2);
                if (!(z >= 0.0f && z <= 1.0f)) continue;
                const unsigned long long k = SHADOW ? (unsigned long long)__float_as_uint(z)
                                                    : (((unsigned long long)__float_as_uint(z) << 32) | e.idKey);
                if (k < key) key = k;   // early-Z + later-primitive-wins in one compare
            }
        }
    }
    return key;
}

// sample position of this thread inside its tile: warp w owns the 8 x 4 block at (8 * (w & 3), 4 * (w >> 2))
__device__ __forceinline__ void tile_sample(int& lx, int& ly) {
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    lx = (int)(8u * (warp & 3u) + (lane & 7u));
    ly = (int)(4u * (warp >> 2) + (lane >> 3));
}

// Broadcast-list merge for a block of SUB vertically stacked tiles: every thread carries SUB samples (same
// column, rows ly + TILE_H * r), so the staging cost and the load latency are amortised over SUB samples.
template <int SUB>
__device__ __forceinline__ void merge_huge_shadow(const LargeRec* __restrict__ src, uint32_t count, int tx0, int ty0, int lx, int ly,
                                                  unsigned long long (&key)[SUB], TileShared& sm) {
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const int32_t dx = 256 * lx, dy = 256 * ly;
    for (uint32_t base = 0; base < count; base += STAGE) {
        const uint32_t n = min((uint32_t)STAGE, count - base);
        __syncthreads();
        if (warp < 2) {
            bool live = false;
            if (threadIdx.x < n) live = stage_record<SUB * TILE_H>(src[base + threadIdx.x], tx0, ty0, sm.stage[threadIdx.x]);
            const unsigned m = __ballot_sync(0xFFFFFFFFu, live);
            if (lane == 0) sm.liveMask[warp] = m;
        }
        __syncthreads();
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            unsigned m = sm.liveMask[h];   // CTA-uniform
            while (m) {
                const EdgeRec& e = sm.stage[32 * h + __ffs(m) - 1];
                m &= m - 1;
                long long F0 = e.E[0] + (long long)e.A[0] * dx + (long long)e.B[0] * dy;
                long long F1 = e.E[1] + (long long)e.A[1] * dx + (long long)e.B[1] * dy;
                long long F2 = e.E[2] + (long long)e.A[2] * dx + (long long)e.B[2] * dy;
                const long long s0 = (long long)e.B[0] * (256 * TILE_H), s1 = (long long)e.B[1] * (256 * TILE_H), s2 = (long long)e.B[2] * (256 * TILE_H);
#pragma unroll
                for (int r = 0; r < SUB; ++r, F0 += s0, F1 += s1, F2 += s2) {
                    if ((F0 | F1 | F2) < 0) continue;
                    const float z = depth_from((float)(F1 + (long long)(e.bias12 & 1u)), (float)(F2 + (long long)(e.bias12 >> 1)), e.area, e.z0, e.dz1, e.dz2);
                    if (!(z >= 0.0f && z <= 1.0f)) continue;
                    const unsigned long long k = (unsigned long long)__float_as_uint(z);
                    if (k < key[r]) key[r] = k;
                }
            }
        }
    }
}

// Shadow pass tile stage: only tiles touched by large triangles do work (the small ones were resolved by K1's
// atomics).  One CTA per column of SHADOW_SUB stacked tiles, SHADOW_SUB samples per thread, all loads up front.
constexpr int SHADOW_SUB = 4;
__global__ void __launch_bounds__(256) k_tile_shadow(const __grid_constant__ PassConst pc) {
    __shared__ TileShared sm;
    pdl_enter();
    const uint32_t nLarge = pc.status[0], nHuge = min(pc.status[4], (uint32_t)HUGE_CAP);
    uint32_t begin[SHADOW_SUB], end[SHADOW_SUB];
    bool anyList = false;
#pragma unroll
    for (int r = 0; r < SHADOW_SUB; ++r) {
        begin[r] = end[r] = 0;
        const uint32_t tyR = blockIdx.y * SHADOW_SUB + r;
        if (nLarge && tyR < pc.tilesY) {
            const uint32_t tile = tyR * pc.tilesX + blockIdx.x;
            const uint32_t cnt = __ldcg(pc.tileCount + tile);
            if (cnt) {
                const uint32_t start = __ldcg(pc.tileStart + tile);
                begin[r] = min(start, pc.listCap); end[r] = min(start + cnt, pc.listCap);
                if (threadIdx.x == 0) pc.tileCount[tile] = 0;   // re-arm the bin counter for the next pass (only this CTA touches it here)
                anyList = anyList || begin[r] != end[r];
            }
        }
    }
    if (nHuge == 0 && !anyList) return;   // CTA-uniform
    const int tx0 = (int)blockIdx.x * TILE_W, ty0 = (int)blockIdx.y * (TILE_H * SHADOW_SUB);
    int lx, ly;
    tile_sample(lx, ly);
    const int x = tx0 + lx;
    unsigned long long cur[SHADOW_SUB], key[SHADOW_SUB];
#pragma unroll
    for (int r = 0; r < SHADOW_SUB; ++r) {
        const int y = ty0 + TILE_H * r + ly;
        const bool inb = x < (int)pc.W && y < (int)pc.H;
        cur[r] = inb ? (unsigned long long)__ldcg(pc.depthBits + (size_t)y * pc.W + x) : 0ull;
        key[r] = cur[r];
    }
    if (nHuge) merge_huge_shadow<SHADOW_SUB>(pc.huge, nHuge, tx0, ty0, lx, ly, key, sm);
#pragma unroll
    for (int r = 0; r < SHADOW_SUB; ++r)
        if (begin[r] != end[r]) key[r] = merge_large<true>(pc.large, pc.tileList, begin[r], end[r], tx0, ty0 + TILE_H * r, lx, ly, key[r], sm);
#pragma unroll
    for (int r = 0; r < SHADOW_SUB; ++r) {
        const int y = ty0 + TILE_H * r + ly;
        if (key[r] < cur[r]) pc.depthBits[(size_t)y * pc.W + x] = (uint32_t)key[r];   // cur = 0 out of bounds, so never taken there
    }
}

// Forward pass tile stage: consume (and reset) the visibility keys, merge large triangles, shade, write.
#ifndef KT_MIN_BLOCKS
#define KT_MIN_BLOCKS 3
#endif
template <bool TEXTURED>
__global__ void __launch_bounds__(256, KT_MIN_BLOCKS) k_tile_forward(const __grid_constant__ PassConst pc, const __grid_constant__ ShadeConst sh) {
    __shared__ TileShared sm;
    pdl_enter();
    const uint32_t tile = blockIdx.y * pc.tilesX + blockIdx.x;
    const int tx0 = (int)blockIdx.x * TILE_W, ty0 = (int)blockIdx.y * TILE_H;
    int lx, ly;
    tile_sample(lx, ly);
    const int x = tx0 + lx, y = ty0 + ly;
    const bool inb = x < (int)pc.W && y < (int)pc.H;
    const size_t p = (size_t)y * pc.W + x;
    const bool owned = row_owned((uint32_t)ty0, pc.rank, pc.nRanks, pc.bandH);
    unsigned long long key = VIS_CLEAR;
    if (inb && owned) key = __ldcg(pc.vis + p);
    const uint32_t nLarge = pc.status[0], nHuge = min(pc.status[4], (uint32_t)HUGE_CAP);
    uint32_t begin = 0, end = 0;
    if (nLarge) {
        const uint32_t cnt = __ldcg(pc.tileCount + tile);
        if (cnt) {
            const uint32_t start = __ldcg(pc.tileStart + tile);
            begin = min(start, pc.listCap); end = min(start + cnt, pc.listCap);
            if (threadIdx.x == 0) pc.tileCount[tile] = 0;   // re-arm (only this CTA touches this counter here)
        }
    }
    if (!owned) return;
    if (inb && key != VIS_CLEAR) pc.vis[p] = VIS_CLEAR;   // leave the buffer cleared for the next frame
    if (nHuge) key = merge_large<false>(pc.huge, nullptr, 0, nHuge, tx0, ty0, lx, ly, key, sm);
    if (begin != end) key = merge_large<false>(pc.large, pc.tileList, begin, end, tx0, ty0, lx, ly, key, sm);
    if (!inb) return;

    float depth = 1.0f;                    // depth clear 1.0 (RenderScene.cpp:219)
    uint32_t tri = TRI_NONE;
    float4 col = make_float4(0.0f, 0.0f, 0.0f, 1.0f);   // colour clear (0,0,0,1) (RenderScene.cpp:218)
    if (key != VIS_CLEAR) {
        depth = __uint_as_float((uint32_t)(key >> 32));
        tri = ~(uint32_t)key;
        if (sh.outSRGB8 || sh.outF32) col = shade_pixel<TEXTURED>(pc, sh, x, y, tri);
    }
    if (sh.outDepth) sh.outDepth[p] = depth;
    if (sh.outTriId) sh.outTriId[p] = tri;
    if (sh.outSRGB8) reinterpret_cast<uint32_t*>(sh.outSRGB8)[p] = pack_srgb8(col);
    if (sh.outF32) reinterpret_cast<float4*>(sh.outF32)[p] = col;
}

void launch_tile_shadow(const PassConst& pc, cudaStream_t st) {
    launch_k(k_tile_shadow, dim3(pc.tilesX, (pc.tilesY + SHADOW_SUB - 1) / SHADOW_SUB), dim3(256), st, true, pc);
}

void launch_tile_forward(const PassConst& pc, const ShadeConst& sh, cudaStream_t st) {
    // the textured variant carries the uv derivatives + sampler state; untextured scenes keep the leaner kernel
    if (sh.nTextures && pc.vtxUV) launch_k(k_tile_forward<true>, dim3(pc.tilesX, pc.tilesY), dim3(256), st, true, pc, sh);
    else launch_k(k_tile_forward<false>, dim3(pc.tilesX, pc.tilesY), dim3(256), st, true, pc, sh);
}

}  // namespace reb
