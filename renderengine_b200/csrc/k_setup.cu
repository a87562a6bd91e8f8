// k_setup.cu — K1: vertex transform + triangle setup + small-triangle raster; compaction of the
// remaining ("large") triangles for the binned tile path; K2: tile binning (count / scan / scatter).
//
// Replaces, for the hot path, the vertex stage and fixed-function front end that the reference
// reaches through vkCmdDrawIndexed (MaterialSystem.cpp:99, ShadowSystem.cpp:369) with the state of
// GraphicsPipelineCache.cpp:257-365: shadow_map.vert:48-52 / primitive.vert:49-53, viewport, cull
// back + CCW front (F5), 1 sample per pixel, LessOrEqual depth.
#include "reb_device.cuh"
#include "reb_kernels.h"
#include <algorithm>

namespace reb {

// ---------------------------------------------------------------------------------------------
// K1: triangle setup + small-triangle raster.  256 triangles per CTA, two phases.
//
// Phase 1, one thread per triangle: 3 indices + 3 VtxA gathers (16 B each) from the post-transform
//   cache, clip-flag logic, facing (int64 shoelace), pixel-centre bbox, classification:
//     small  (spans <= SMALL_MAX px)  -> compacted into shared memory as 32-bit edge equations,
//     huge   (> HUGE_TILES tiles)     -> short broadcast list tested by every tile,
//     large  (everything else)        -> warp-aggregated append to the LargeRec queue (binned path).
//   Compaction = warp ballot + shuffle scan + one shared-memory pass over the 8 warp totals; the same
//   scan carries the number of pixel rows of every small triangle.
// Phase 2, one thread per (small triangle, pixel row) work item: back-face-culled and clipped lanes
//   no longer idle through the raster loop, and lanes of a warp walk rows of near-equal width.
//   Coverage = three incrementally updated 32-bit edge functions with the top-left bias folded in;
//   covered samples resolve with one 64-bit atomicMin (key = depthBits<<32 | ~triId; shadow pass:
//   32-bit atomicMin on the depth bits, which are the shadow map itself).
// ---------------------------------------------------------------------------------------------
constexpr int K1_THREADS = 256;
constexpr int K1_MAX_ROWS = K1_THREADS * (SMALL_MAX + 1);

struct K1Shared {
    int32_t Ax[3][K1_THREADS];      // 256 * A_k  (step of the biased edge function per pixel in x)
    int32_t By[3][K1_THREADS];      // 256 * B_k  (step per pixel in y)
    int32_t C[3][K1_THREADS];       // biased edge value at the bbox origin's pixel centre
    uint32_t xy[K1_THREADS];        // x0 | y0 << 16
    uint32_t whb[K1_THREADS];       // w | h << 8 | bias1 << 16 | bias2 << 17
    float z0[K1_THREADS], dz1[K1_THREADS], dz2[K1_THREADS], area[K1_THREADS];
    uint32_t idKey[K1_THREADS];
    uint32_t rowStart[K1_THREADS];
    uint32_t warpTot[K1_THREADS / 32];
    uint32_t drawBase;
    uint8_t itemTri[K1_MAX_ROWS];
};

template <bool SHADOW>
__global__ void __launch_bounds__(K1_THREADS) k_setup_raster(const __grid_constant__ PassConst pc) {
    __shared__ K1Shared sm;
    const uint32_t firstTri = blockIdx.x * K1_THREADS;
    const uint32_t tri = firstTri + threadIdx.x;
    const bool valid = tri < pc.nTris;
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;

    const uint32_t d = find_draw_cta<false>(pc.draws, pc.nDraws, firstTri, valid ? tri : firstTri, &sm.drawBase);
    bool keep = false, needsClip = false;
    VtxA a0, a1, a2;
    long long A2 = 0;
    int32_t x0 = 0, x1 = -1, y0 = 0, y1 = -1, spanX = 0, spanY = 0;
    if (valid) {
        const uint4 dr = __ldg(reinterpret_cast<const uint4*>(pc.draws + d));        // indexOffset, firstTri, material, node
        const int32_t cacheOffset = __ldg(&pc.draws[d].cacheOffset);
        const uint32_t* ix = pc.indices + dr.x + 3u * (tri - dr.y);
        const uint32_t i0 = __ldg(ix), i1 = __ldg(ix + 1), i2 = __ldg(ix + 2);
        const uint4* va = reinterpret_cast<const uint4*>(pc.vtxA);
        const uint4 q0 = __ldg(va + cache_slot(pc, cacheOffset, tri, 0, i0));
        const uint4 q1 = __ldg(va + cache_slot(pc, cacheOffset, tri, 1, i1));
        const uint4 q2 = __ldg(va + cache_slot(pc, cacheOffset, tri, 2, i2));
        a0 = {(int32_t)q0.x, (int32_t)q0.y, __uint_as_float(q0.z), q0.w};
        a1 = {(int32_t)q1.x, (int32_t)q1.y, __uint_as_float(q1.z), q1.w};
        a2 = {(int32_t)q2.x, (int32_t)q2.y, __uint_as_float(q2.z), q2.w};
        const bool reject = (a0.flags & a1.flags & a2.flags & VF_PLANES) != 0u;             // Appendix B.2
        needsClip = !reject && ((a0.flags | a1.flags | a2.flags) & VF_NEEDS_CLIP) != 0u;    // Appendix B.3
        if (!reject && !needsClip) {
            A2 = (long long)(a1.X - a0.X) * (long long)(a2.Y - a0.Y) - (long long)(a2.X - a0.X) * (long long)(a1.Y - a0.Y);
            if (A2 < 0) {                                                                   // Appendix B.6: front = CCW = A2 < 0
                const int32_t minX = min(a0.X, min(a1.X, a2.X)), maxX = max(a0.X, max(a1.X, a2.X));
                const int32_t minY = min(a0.Y, min(a1.Y, a2.Y)), maxY = max(a0.Y, max(a1.Y, a2.Y));
                spanX = maxX - minX; spanY = maxY - minY;
                x0 = max((minX + 127) >> 8, 0); x1 = min((maxX - 128) >> 8, (int32_t)pc.W - 1);
                y0 = max((minY + 127) >> 8, 0); y1 = min((maxY - 128) >> 8, (int32_t)pc.H - 1);
                keep = x0 <= x1 && y0 <= y1;
            }
        }
    }
    if (needsClip) atomicAdd(&pc.status[1], 1u);

    const bool isSmall = keep && spanX <= SMALL_MAX * 256 && spanY <= SMALL_MAX * 256;   // unclamped spans keep the 32-bit edge math exact
    bool isLarge = keep && !isSmall;
    if (isLarge) {
        // screen-filling triangles go to a short broadcast list that every tile tests (no per-tile atomics)
        const int ntiles = (x1 / TILE_W - x0 / TILE_W + 1) * (y1 / TILE_H - y0 / TILE_H + 1);
        if (ntiles > HUGE_TILES) {
            const uint32_t slot = atomicAdd(&pc.status[4], 1u);
            if (slot < (uint32_t)HUGE_CAP) {
                LargeRec r;
                r.X0 = a0.X; r.Y0 = a0.Y; r.X1 = a1.X; r.Y1 = a1.Y;
                r.X2 = a2.X; r.Y2 = a2.Y; r.z0 = a0.z; r.z1 = a1.z;
                r.z2 = a2.z; r.triId = tri; r.pad0 = 0; r.pad1 = 0;
                pc.huge[slot] = r;
                isLarge = false;
            }
        }
    }

    // warp-aggregated append of large triangles to the global queue
    const unsigned lmask = __ballot_sync(0xFFFFFFFFu, isLarge);
    if (lmask) {
        const int leader = __ffs(lmask) - 1;
        uint32_t base = 0;
        if ((int)lane == leader) base = atomicAdd(&pc.status[0], (uint32_t)__popc(lmask));
        base = __shfl_sync(0xFFFFFFFFu, base, leader);
        if (isLarge) {
            const uint32_t slot = base + __popc(lmask & ((1u << lane) - 1u));
            if (slot < pc.largeCap) {
                LargeRec r;
                r.X0 = a0.X; r.Y0 = a0.Y; r.X1 = a1.X; r.Y1 = a1.Y;
                r.X2 = a2.X; r.Y2 = a2.Y; r.z0 = a0.z; r.z1 = a1.z;
                r.z2 = a2.z; r.triId = tri; r.pad0 = 0; r.pad1 = 0;
                pc.large[slot] = r;
            } else {
                atomicOr(&pc.status[2], OVF_LARGE);
            }
        }
    }

    // CTA compaction of small triangles: packed scan of (count, rows) = 1 | h << 16
    const uint32_t h = isSmall ? (uint32_t)(y1 - y0 + 1) : 0u;
    const uint32_t mine = isSmall ? (1u | (h << 16)) : 0u;
    uint32_t incl = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const uint32_t n = __shfl_up_sync(0xFFFFFFFFu, incl, o); if (lane >= (unsigned)o) incl += n; }
    if (lane == 31) sm.warpTot[warp] = incl;
    __syncthreads();
    uint32_t base = 0, total = 0;
#pragma unroll
    for (int wgt = 0; wgt < K1_THREADS / 32; ++wgt) { const uint32_t t = sm.warpTot[wgt]; if (wgt < (int)warp) base += t; total += t; }
    const uint32_t nRows = total >> 16;
    if (nRows == 0) return;   // CTA-uniform
    if (isSmall) {
        const uint32_t excl = base + incl - mine;
        const uint32_t slot = excl & 0xFFFFu, row0 = excl >> 16;
        const int32_t X[3] = {a0.X, a1.X, a2.X}, Y[3] = {a0.Y, a1.Y, a2.Y};
        const int32_t Px = 256 * x0 + 128, Py = 256 * y0 + 128;
        uint32_t biasBits = 0;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const int a = (k + 1) % 3, b = (k + 2) % 3;     // edge a->b opposes vertex k
            const int32_t A = Y[b] - Y[a], B = -(X[b] - X[a]);
            const int32_t bias = ((A > 0) || (A == 0 && B > 0)) ? 0 : 1;   // top-left rule (Appendix B.7)
            sm.Ax[k][slot] = 256 * A;
            sm.By[k][slot] = 256 * B;
            sm.C[k][slot] = A * (Px - X[a]) + B * (Py - Y[a]) - bias;
            if (k > 0) biasBits |= (uint32_t)bias << (15 + k);
        }
        sm.xy[slot] = (uint32_t)x0 | ((uint32_t)y0 << 16);
        sm.whb[slot] = (uint32_t)(x1 - x0 + 1) | (h << 8) | biasBits;
        sm.z0[slot] = a0.z; sm.dz1[slot] = a1.z - a0.z; sm.dz2[slot] = a2.z - a0.z;
        sm.area[slot] = (float)(-A2);
        sm.idKey[slot] = ~tri;
        sm.rowStart[slot] = row0;
        for (uint32_t j = 0; j < h; ++j) sm.itemTri[row0 + j] = (uint8_t)slot;
    }
    __syncthreads();

    for (uint32_t item = threadIdx.x; item < nRows; item += K1_THREADS) {
        const uint32_t t = sm.itemTri[item];
        const uint32_t j = item - sm.rowStart[t];
        const uint32_t xy = sm.xy[t], whb = sm.whb[t];
        const uint32_t y = (xy >> 16) + j;
        if (!SHADOW && !row_owned(y, pc.rank, pc.nRanks, pc.bandH)) continue;
        const int32_t s0 = sm.Ax[0][t], s1 = sm.Ax[1][t], s2 = sm.Ax[2][t];
        int32_t F0 = sm.C[0][t] + (int32_t)j * sm.By[0][t];
        int32_t F1 = sm.C[1][t] + (int32_t)j * sm.By[1][t];
        int32_t F2 = sm.C[2][t] + (int32_t)j * sm.By[2][t];
        const uint32_t w = whb & 0xFFu;
        const size_t rowBase = (size_t)y * pc.W + (xy & 0xFFFFu);
        for (uint32_t i = 0; i < w; ++i, F0 += s0, F1 += s1, F2 += s2) {
            if ((F0 | F1 | F2) < 0) continue;
            const float z = depth_from((float)(F1 + (int32_t)((whb >> 16) & 1u)), (float)(F2 + (int32_t)((whb >> 17) & 1u)),
                                       sm.area[t], sm.z0[t], sm.dz1[t], sm.dz2[t]);
            if (!(z >= 0.0f && z <= 1.0f)) continue;
            if (SHADOW) atomicMin(pc.depthBits + rowBase + i, __float_as_uint(z));
            else atomicMin(pc.vis + rowBase + i, ((unsigned long long)__float_as_uint(z) << 32) | sm.idKey[t]);
        }
    }
}

// ---------------------------------------------------------------------------------------------
// K2: bin the large triangles into TILE_W x TILE_H screen tiles.  One warp per triangle, lanes
// stride over the tiles of its bbox; a tile is skipped when it lies entirely outside one edge
// (evaluated at the tile's most-interior corner for that edge).  SCATTER=false counts,
// SCATTER=true writes triangle slots into the per-tile lists laid out by the scan.
// ---------------------------------------------------------------------------------------------
template <bool SCATTER>
__global__ void __launch_bounds__(256) k_bin_large(const __grid_constant__ PassConst pc, uint32_t* __restrict__ tileCount,
                                                   const uint32_t* __restrict__ tileStart, uint32_t* __restrict__ tileList, uint32_t listCap) {
    const uint32_t nLarge = min(pc.status[0], pc.largeCap);
    if (nLarge == 0) return;
    const unsigned lane = threadIdx.x & 31u;
    const uint32_t warpsPerGrid = (gridDim.x * blockDim.x) >> 5;
    for (uint32_t t = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; t < nLarge; t += warpsPerGrid) {
        const LargeRec r = pc.large[t];
        const int32_t X[3] = {r.X0, r.X1, r.X2}, Y[3] = {r.Y0, r.Y1, r.Y2};
        const int32_t minX = min(X[0], min(X[1], X[2])), maxX = max(X[0], max(X[1], X[2]));
        const int32_t minY = min(Y[0], min(Y[1], Y[2])), maxY = max(Y[0], max(Y[1], Y[2]));
        const int x0 = max((minX + 127) >> 8, 0), x1 = min((maxX - 128) >> 8, (int)pc.W - 1);
        const int y0 = max((minY + 127) >> 8, 0), y1 = min((maxY - 128) >> 8, (int)pc.H - 1);
        const int tx0 = x0 / TILE_W, tx1 = x1 / TILE_W, ty0 = y0 / TILE_H, ty1 = y1 / TILE_H;
        const int tw = tx1 - tx0 + 1, n = tw * (ty1 - ty0 + 1);
        long long A[3], B[3], C[3];
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const int a = (k + 1) % 3, b = (k + 2) % 3;
            A[k] = (long long)Y[b] - Y[a];
            B[k] = -((long long)X[b] - X[a]);
            C[k] = -(A[k] * X[a]) - B[k] * Y[a];
        }
        for (int i = lane; i < n; i += 32) {
            const int tx = tx0 + i % tw, ty = ty0 + i / tw;
            if (pc.nRanks > 1 && !row_owned((uint32_t)(ty * TILE_H), pc.rank, pc.nRanks, pc.bandH)) continue;
            // pixel-centre extent of the tile in sub-pixel units
            const long long cx0 = 256ll * (tx * TILE_W) + 128, cx1 = 256ll * (tx * TILE_W + TILE_W - 1) + 128;
            const long long cy0 = 256ll * (ty * TILE_H) + 128, cy1 = 256ll * (ty * TILE_H + TILE_H - 1) + 128;
            bool outside = false;
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                const long long Fmax = A[k] * (A[k] > 0 ? cx1 : cx0) + B[k] * (B[k] > 0 ? cy1 : cy0) + C[k];
                outside = outside || (Fmax < 0);
            }
            if (outside) continue;
            const uint32_t tile = (uint32_t)ty * pc.tilesX + (uint32_t)tx;
            if (!SCATTER) {
                atomicAdd(&tileCount[tile], 1u);
            } else {
                const uint32_t slot = tileStart[tile] + atomicAdd(&tileCount[tile], 1u);
                if (slot < listCap) tileList[slot] = t;
            }
        }
    }
}

// Exclusive scan of the per-tile counts (single CTA): every thread owns a contiguous chunk, one block scan of
// the chunk sums.  Resets the counts so the scatter pass can reuse them as cursors; records the list total.
__global__ void __launch_bounds__(1024) k_scan_tiles(uint32_t* __restrict__ tileCount, uint32_t* __restrict__ tileStart, uint32_t nTiles,
                                                     uint32_t listCap, uint32_t* __restrict__ status) {
    if (status[0] == 0) return;   // nothing was binned: the tile stage ignores the lists
    __shared__ uint32_t warpSums[32];
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const uint32_t chunk = (nTiles + 1023u) / 1024u;
    const uint32_t lo = min(threadIdx.x * chunk, nTiles), hi = min(lo + chunk, nTiles);
    uint32_t sum = 0;
    for (uint32_t i = lo; i < hi; ++i) sum += tileCount[i];
    uint32_t incl = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { uint32_t n = __shfl_up_sync(0xFFFFFFFFu, incl, o); if (lane >= (unsigned)o) incl += n; }
    if (lane == 31) warpSums[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        uint32_t ws = warpSums[lane], wi = ws;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { uint32_t n = __shfl_up_sync(0xFFFFFFFFu, wi, o); if (lane >= (unsigned)o) wi += n; }
        warpSums[lane] = wi - ws;
    }
    __syncthreads();
    uint32_t run = warpSums[warp] + incl - sum;
    for (uint32_t i = lo; i < hi; ++i) { const uint32_t c = tileCount[i]; tileStart[i] = run; tileCount[i] = 0; run += c; }
    if (threadIdx.x == 1023) {
        tileStart[nTiles] = run;
        status[3] = run;
        if (run > listCap) atomicOr(&status[2], OVF_LIST);
    }
}

__global__ void k_fill_u32(uint32_t* __restrict__ p, uint32_t v, size_t n) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    uint4* p4 = reinterpret_cast<uint4*>(p);
    const size_t n4 = n / 4;
    const uint4 v4 = make_uint4(v, v, v, v);
    for (size_t k = i; k < n4; k += stride) p4[k] = v4;
    for (size_t k = n4 * 4 + i; k < n; k += stride) p[k] = v;
}

__global__ void k_fill_u32x4(uint4* __restrict__ p, uint4 v, size_t n4) {
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (size_t k = (size_t)blockIdx.x * blockDim.x + threadIdx.x; k < n4; k += stride) p[k] = v;
}

// AoS 32-B Vertex (MeshAllocator.h:14-24) -> two 16-B streams: pos4 = (x,y,z,u), nrm4 = (nx,ny,nz,v)
__global__ void k_split_vertices(const float4* __restrict__ aos, float4* __restrict__ pos4, float4* __restrict__ nrm4, size_t nVerts) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nVerts) return;
    const float4 a = aos[2 * i], b = aos[2 * i + 1];   // a = pos.xyz, n.x ; b = n.y, n.z, u, v
    pos4[i] = make_float4(a.x, a.y, a.z, b.z);
    nrm4[i] = make_float4(a.w, b.x, b.y, b.w);
}

// ---------------------------------------------------------------------------------------------
// launch wrappers
// ---------------------------------------------------------------------------------------------
void launch_setup_raster(const PassConst& pc, bool shadow, cudaStream_t st) {
    if (pc.nTris == 0) return;
    const uint32_t grid = (pc.nTris + 255u) / 256u;
    if (shadow) k_setup_raster<true><<<grid, 256, 0, st>>>(pc);
    else k_setup_raster<false><<<grid, 256, 0, st>>>(pc);
}

void launch_bin_count(const PassConst& pc, uint32_t* tileCount, uint32_t* tileStart, uint32_t* tileList, uint32_t listCap,
                      int numSMs, cudaStream_t st) {
    k_bin_large<false><<<numSMs * 4, 256, 0, st>>>(pc, tileCount, tileStart, tileList, listCap);
}
void launch_scan_tiles(const PassConst& pc, uint32_t* tileCount, uint32_t* tileStart, uint32_t listCap, cudaStream_t st) {
    k_scan_tiles<<<1, 1024, 0, st>>>(tileCount, tileStart, pc.tilesX * pc.tilesY, listCap, pc.status);
}
void launch_bin_scatter(const PassConst& pc, uint32_t* tileCount, uint32_t* tileStart, uint32_t* tileList, uint32_t listCap,
                        int numSMs, cudaStream_t st) {
    k_bin_large<true><<<numSMs * 4, 256, 0, st>>>(pc, tileCount, tileStart, tileList, listCap);
}

void launch_fill_u32(uint32_t* p, uint32_t v, size_t n, int numSMs, cudaStream_t st) {
    if (n == 0) return;
    size_t blocks = (n / 4 + 255) / 256;
    if (blocks < 1) blocks = 1;
    if (blocks > (size_t)numSMs * 16) blocks = (size_t)numSMs * 16;
    k_fill_u32<<<(unsigned)blocks, 256, 0, st>>>(p, v, n);
}

void launch_fill_u32x4(void* p, uint32_t a, uint32_t b, uint32_t c, uint32_t d, size_t n4, int numSMs, cudaStream_t st) {
    if (n4 == 0) return;
    size_t blocks = std::min<size_t>((n4 + 255) / 256, (size_t)numSMs * 16);
    k_fill_u32x4<<<(unsigned)blocks, 256, 0, st>>>(reinterpret_cast<uint4*>(p), make_uint4(a, b, c, d), n4);
}

void launch_split_vertices(const void* aos, float4* pos4, float4* nrm4, size_t nVerts, cudaStream_t st) {
    if (nVerts == 0) return;
    k_split_vertices<<<(unsigned)((nVerts + 255) / 256), 256, 0, st>>>(reinterpret_cast<const float4*>(aos), pos4, nrm4, nVerts);
}

}  // namespace reb
