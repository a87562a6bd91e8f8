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
// K1: one thread per triangle.
//   * indices / float4 positions fetched through the read-only path (LDG.128 gathers),
//   * both matrix stages, viewport, snap, cull (exact order of Appendix B),
//   * triangles whose bbox covers <= SMALL_MAX x SMALL_MAX pixel centres are rasterised here with
//     32-bit edge functions and resolved with one 64-bit atomicMin per covered sample
//     (key = depthBits<<32 | ~triId; shadow pass: 32-bit atomicMin on the depth bits),
//   * everything else is compacted (warp ballot + popc prefix, one atomicAdd per warp) into the
//     LargeRec queue consumed by the tile path.
// ---------------------------------------------------------------------------------------------
template <bool SHADOW>
__global__ void __launch_bounds__(256) k_setup_raster(const __grid_constant__ PassConst pc) {
    const uint32_t tri = blockIdx.x * blockDim.x + threadIdx.x;
    const bool valid = tri < pc.nTris;
    Snapped s;
    SetupResult res = SR_REJECT;
    if (valid) {
        const uint32_t d = find_draw(pc.draws, pc.nDraws, tri);
        const uint4 dr = __ldg(reinterpret_cast<const uint4*>(pc.draws) + d);   // indexOffset, firstTri, material, node
        const uint32_t* ix = pc.indices + dr.x + 3u * (tri - dr.y);
        const uint32_t i0 = __ldg(ix), i1 = __ldg(ix + 1), i2 = __ldg(ix + 2);
        const float4 p0 = __ldg(pc.pos4 + i0), p1 = __ldg(pc.pos4 + i1), p2 = __ldg(pc.pos4 + i2);
        const float4* Tq = reinterpret_cast<const float4*>(pc.transforms) + 4u * dr.w;
        float T[16];
#pragma unroll
        for (int k = 0; k < 4; ++k) { float4 q = __ldg(Tq + k); T[4 * k] = q.x; T[4 * k + 1] = q.y; T[4 * k + 2] = q.z; T[4 * k + 3] = q.w; }
        float4 c[3];
        c[0] = vertex_stage<!SHADOW>(pc.PV, T, p0.x, p0.y, p0.z, nullptr);
        c[1] = vertex_stage<!SHADOW>(pc.PV, T, p1.x, p1.y, p1.z, nullptr);
        c[2] = vertex_stage<!SHADOW>(pc.PV, T, p2.x, p2.y, p2.z, nullptr);
        res = setup_triangle(c, pc.W, pc.H, s);
    }
    if (res == SR_NEEDS_CLIP) atomicAdd(&pc.status[1], 1u);

    const bool keep = (res == SR_KEEP);
    const bool isSmall = keep && (s.spanX <= SMALL_MAX * 256) && (s.spanY <= SMALL_MAX * 256);   // unclamped: keeps 32-bit edge math exact
    bool isLarge = keep && !isSmall;
    if (isLarge) {
        // screen-filling triangles go to a short broadcast list that every tile tests (no per-tile atomics)
        const int ntiles = (s.x1 / TILE_W - s.x0 / TILE_W + 1) * (s.y1 / TILE_H - s.y0 / TILE_H + 1);
        if (ntiles > HUGE_TILES) {
            const uint32_t slot = atomicAdd(&pc.status[4], 1u);
            if (slot < (uint32_t)HUGE_CAP) {
                LargeRec r;
                r.X0 = s.X[0]; r.Y0 = s.Y[0]; r.X1 = s.X[1]; r.Y1 = s.Y[1];
                r.X2 = s.X[2]; r.Y2 = s.Y[2]; r.z0 = s.z[0]; r.z1 = s.z[1];
                r.z2 = s.z[2]; r.triId = tri; r.pad0 = 0; r.pad1 = 0;
                pc.huge[slot] = r;
                isLarge = false;
            }
        }
    }

    if (isSmall) {
        // 32-bit edge functions relative to the edge's first vertex; all magnitudes < 2^26 here.
        int32_t A[3], B[3], Xa[3], Ya[3], bias[3];
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const int a = (k + 1) % 3, b = (k + 2) % 3;
            A[k] = s.Y[b] - s.Y[a];
            B[k] = -(s.X[b] - s.X[a]);
            Xa[k] = s.X[a]; Ya[k] = s.Y[a];
            bias[k] = ((A[k] > 0) || (A[k] == 0 && B[k] > 0)) ? 0 : 1;   // top-left rule
        }
        const float area = (float)(-s.A2);
        const uint32_t idKey = ~tri;
        for (int y = s.y0; y <= s.y1; ++y) {
            if (!SHADOW && !row_owned((uint32_t)y, pc.rank, pc.nRanks, pc.bandH)) continue;
            const int32_t Py = 256 * y + 128;
            for (int x = s.x0; x <= s.x1; ++x) {
                const int32_t Px = 256 * x + 128;
                const int32_t F0 = A[0] * (Px - Xa[0]) + B[0] * (Py - Ya[0]);
                const int32_t F1 = A[1] * (Px - Xa[1]) + B[1] * (Py - Ya[1]);
                const int32_t F2 = A[2] * (Px - Xa[2]) + B[2] * (Py - Ya[2]);
                if (((F0 - bias[0]) | (F1 - bias[1]) | (F2 - bias[2])) < 0) continue;
                const float z = depth_from((float)F1, (float)F2, area, s.z[0], s.z[1], s.z[2]);
                if (!(z >= 0.0f && z <= 1.0f)) continue;
                const size_t p = (size_t)y * pc.W + x;
                if (SHADOW) atomicMin(pc.depthBits + p, __float_as_uint(z));
                else atomicMin(pc.vis + p, ((unsigned long long)__float_as_uint(z) << 32) | idKey);
            }
        }
    }

    // warp-aggregated compaction of large triangles
    const unsigned mask = __ballot_sync(0xFFFFFFFFu, isLarge);
    if (mask) {
        const unsigned lane = threadIdx.x & 31u;
        const int leader = __ffs(mask) - 1;
        uint32_t base = 0;
        if ((int)lane == leader) base = atomicAdd(&pc.status[0], (uint32_t)__popc(mask));
        base = __shfl_sync(0xFFFFFFFFu, base, leader);
        if (isLarge) {
            const uint32_t slot = base + __popc(mask & ((1u << lane) - 1u));
            if (slot < pc.largeCap) {
                LargeRec r;
                r.X0 = s.X[0]; r.Y0 = s.Y[0]; r.X1 = s.X[1]; r.Y1 = s.Y[1];
                r.X2 = s.X[2]; r.Y2 = s.Y[2]; r.z0 = s.z[0]; r.z1 = s.z[1];
                r.z2 = s.z[2]; r.triId = tri; r.pad0 = 0; r.pad1 = 0;
                pc.large[slot] = r;
            } else {
                atomicOr(&pc.status[2], OVF_LARGE);
            }
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
