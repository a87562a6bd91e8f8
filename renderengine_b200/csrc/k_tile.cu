// k_tile.cu — the tile stages.  Every WARP is an independent worker that owns an 8-px-wide block of samples (compact footprint:
// neighbouring samples share triangles, so the post-transform-cache gathers of a warp fall into few sectors).
//   K2 k_tile_binned : the binned path of triangles too large for K1's own raster — leaves at once for the usual frame.  Clips
//                      queued parents, lays out the per-tile lists, scatters the 48-B records (bin_phase, reb_bin.cuh), then
//                      rasterises every tile's list: the list is a contiguous run of records, streamed into the warp's landing
//                      buffer in shared memory by 1-D bulk asynchronous copies (cp.async.bulk + mbarrier; the next chunk lands
//                      while the current one is tested), converted — one record per lane — into edge equations relative to the
//                      block origin, tested by every lane with the top-left rule, 64-bit edge functions and a min-key compare
//                      against the key held in a register, and merged with one atomicMin per covered sample.
//   K3 shadow pass   : the broadcast list (triangles covering > 256 tiles, e.g. a ground plane) against every block; the surviving
//                      depth of every covered sample is merged into the map with one fire-and-forget atomicMin (the small
//                      triangles were resolved the same way by K1), so the map is never read here.
//   K3 forward pass  : persistent CTAs stride over the tiles; the visibility key of the NEXT tile is loaded before the current one is
//                      shaded (software prefetch); consumes + re-arms the visibility buffer, merges the broadcast list, then shades
//                      every sample in place: attribute reconstruction, bindless material fetch, Phong + 3x3 PCF.
//
// Reference: fixed-function raster/depth state GraphicsPipelineCache.cpp:280-340; fragment work
// phong.glsl:39-81 + shadow.glsl:24-56; clears RenderScene.cpp:217-220 / ShadowSystem.cpp:111-127.
#include "reb_bin.cuh"
#include "reb_shade.cuh"
#include "reb_shade_pbr.cuh"
#include "reb_sky.cuh"
#include <algorithm>
#include <string>
#include <cstdlib>

namespace reb {

constexpr int KT_THREADS = 256, KT_WARPS = KT_THREADS / 32;
constexpr int BLK_W = 8, BLK_ROWS = 4;   // a warp's lanes: 8 columns x 4 rows; SUB stacked copies 4 rows apart

// One warp's staging area: up to 32 large triangles (one per lane), SoA so the per-lane writes are conflict-free
// and the per-record reads are broadcasts.  Edge values are relative to the block's first pixel centre.
struct WarpStage {
    long long E[3][32];         // biased edge value at the block origin (covered iff E + A*dx + B*dy >= 0 for all three)
    int32_t A[3][32], B[3][32];
    float inv[32], z0[32], dz1[32], dz2[32];   // inv = 1 / |A2|
    uint32_t idKey[32], bias12[32];   // bias12: bit 0 = bias of edge 1, bit 1 = bias of edge 2 (to recover the unbiased values)
};

// Stage triangle r for the block whose first pixel is (bx0, by0) and that spans cols x rows pixels; returns false
// when the triangle cannot touch any pixel centre of the block (bbox + edge reject at the most-interior corner).
__device__ __forceinline__ bool stage_record(const LargeRec& r, int bx0, int by0, int cols, int rows, WarpStage& ws, unsigned lane) {
    const int32_t X[3] = {r.X0, r.X1, r.X2}, Y[3] = {r.Y0, r.Y1, r.Y2};
    const int32_t minX = min(r.X0, min(r.X1, r.X2)), maxX = max(r.X0, max(r.X1, r.X2));
    const int32_t minY = min(r.Y0, min(r.Y1, r.Y2)), maxY = max(r.Y0, max(r.Y1, r.Y2));
    const long long cx0 = 256ll * bx0 + 128, cy0 = 256ll * by0 + 128;
    const long long spanX = 256ll * (cols - 1), spanY = 256ll * (rows - 1);
    bool live = !(maxX < cx0 || minX > cx0 + spanX || maxY < cy0 || minY > cy0 + spanY);
    uint32_t bias12 = 0;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const int a = (k + 1) % 3, b = (k + 2) % 3;
        const int32_t A = Y[b] - Y[a], B = -(X[b] - X[a]);
        const int32_t bias = ((A > 0) || (A == 0 && B > 0)) ? 0 : 1;   // top-left rule (Appendix B.7)
        const long long E = (long long)A * (cx0 - X[a]) + (long long)B * (cy0 - Y[a]) - bias;
        ws.A[k][lane] = A; ws.B[k][lane] = B; ws.E[k][lane] = E;
        const long long Emax = E + (A > 0 ? (long long)A * spanX : 0ll) + (B > 0 ? (long long)B * spanY : 0ll);
        live = live && (Emax >= 0);
        if (k > 0) bias12 |= (uint32_t)bias << (k - 1);
    }
    const long long A2 = (long long)(X[1] - X[0]) * (long long)(Y[2] - Y[0]) - (long long)(X[2] - X[0]) * (long long)(Y[1] - Y[0]);
    ws.inv[lane] = 1.0f / (float)(-A2);
    ws.z0[lane] = r.z0; ws.dz1[lane] = r.z1 - r.z0; ws.dz2[lane] = r.z2 - r.z0;
    ws.idKey[lane] = ~r.triId; ws.bias12[lane] = bias12;
    return live;
}

// Inner test of one staged triangle against this lane's SUB samples (F* = biased edge values at sample 0, s* = their
// step to the next sample BLK_ROWS rows below).
template <bool SHADOW> struct KeyOf { typedef unsigned long long type; };
template <> struct KeyOf<true> { typedef uint32_t type; };   // shadow pass: the bare depth bits

template <bool SHADOW, int SUB>
__device__ __forceinline__ void test_samples(long long F0, long long F1, long long F2, long long s0, long long s1, long long s2, uint32_t bias12,
                                             float inv, float z0, float dz1, float dz2, uint32_t idKey, typename KeyOf<SHADOW>::type (&key)[SUB]) {
#pragma unroll
    for (int r = 0; r < SUB; ++r, F0 += s0, F1 += s1, F2 += s2) {
        if ((F0 | F1 | F2) < 0) continue;
        const float z = depth_from((float)(F1 + (long long)(bias12 & 1u)), (float)(F2 + (long long)(bias12 >> 1)), inv, z0, dz1, dz2);
        if (!(z >= 0.0f && z <= 1.0f)) continue;
        const typename KeyOf<SHADOW>::type k = SHADOW ? (typename KeyOf<SHADOW>::type)__float_as_uint(z)
                                                      : (typename KeyOf<SHADOW>::type)(((unsigned long long)__float_as_uint(z) << 32) | idKey);
        if (k < key[r]) key[r] = k;   // early-Z + later-primitive-wins in one compare
    }
}

// The same for a block that lies entirely inside the triangle (every edge function is >= 0 at the block's extreme centres, hence at
// all of them): no coverage test, only the two edge values the depth needs.
template <bool SHADOW, int SUB>
__device__ __forceinline__ void test_samples_inside(long long F1, long long F2, long long s1, long long s2, uint32_t bias12,
                                                    float inv, float z0, float dz1, float dz2, uint32_t idKey, typename KeyOf<SHADOW>::type (&key)[SUB]) {
    F1 += (long long)(bias12 & 1u); F2 += (long long)(bias12 >> 1);   // unbiased
#pragma unroll
    for (int r = 0; r < SUB; ++r, F1 += s1, F2 += s2) {
        const float z = depth_from((float)F1, (float)F2, inv, z0, dz1, dz2);
        if (!(z >= 0.0f && z <= 1.0f)) continue;
        const typename KeyOf<SHADOW>::type k = SHADOW ? (typename KeyOf<SHADOW>::type)__float_as_uint(z)
                                                      : (typename KeyOf<SHADOW>::type)(((unsigned long long)__float_as_uint(z) << 32) | idKey);
        if (k < key[r]) key[r] = k;
    }
}

// ---- 1-D bulk asynchronous copy (TMA engine, no tensor map) global -> shared, completion on an mbarrier --------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// A warp's landing buffers for the raw records of the tile lists it walks: two 32-record buffers (1536 B each), one mbarrier
// each.  The warp's tiles form one stream of <= 32-record chunks; chunk k + 1 (the next chunk of the same list, or the first
// chunk of the warp's next non-empty tile) is copied by the TMA engine while chunk k is converted and tested.
struct __align__(16) WarpRecs {
    LargeRec rec[2][32];
    uint64_t bar[2];
};
__device__ __forceinline__ void recs_init(WarpRecs& wr, unsigned lane) {
    if (lane == 0) { mbar_init(&wr.bar[0], 1u); mbar_init(&wr.bar[1], 1u); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
    __syncwarp();
}
// lane 0 arms barrier b and starts the copy of n records into buffer b
__device__ __forceinline__ void recs_issue(WarpRecs& wr, int b, const LargeRec* __restrict__ src, uint32_t n, unsigned lane) {
#ifndef REB_NO_TMA
    if (lane == 0) {
        fence_proxy_async_smem();   // the buffer's earlier generic-proxy reads are ordered before the async-proxy write
        const uint32_t bytes = n * (uint32_t)sizeof(LargeRec);
        mbar_expect_tx(&wr.bar[b], bytes);
        bulk_g2s(&wr.rec[b][0], src, bytes, &wr.bar[b]);
    }
#endif
}

constexpr int HUGE_DIRECT = 6;   // broadcast lists up to this long skip the per-block liveness pass

// The broadcast list (triangles covering > HUGE_TILES tiles) as screen-space edge equations, built once per CTA.
struct HugeStage {
    long long E[3][HUGE_CAP];       // biased edge value at the centre of pixel (0, 0)
    int32_t A[3][HUGE_CAP], B[3][HUGE_CAP];
    int32_t minX[HUGE_CAP], maxX[HUGE_CAP], minY[HUGE_CAP], maxY[HUGE_CAP];
    float inv[HUGE_CAP], z0[HUGE_CAP], dz1[HUGE_CAP], dz2[HUGE_CAP];
    uint32_t idKey[HUGE_CAP], bias12[HUGE_CAP];
};

__device__ __forceinline__ void stage_huge(const LargeRec* __restrict__ huge, uint32_t nHuge, HugeStage& hs) {
    const uint32_t i = threadIdx.x;
    if (i >= nHuge) return;
    const LargeRec r = huge[i];
    const int32_t X[3] = {r.X0, r.X1, r.X2}, Y[3] = {r.Y0, r.Y1, r.Y2};
    hs.minX[i] = min(r.X0, min(r.X1, r.X2)); hs.maxX[i] = max(r.X0, max(r.X1, r.X2));
    hs.minY[i] = min(r.Y0, min(r.Y1, r.Y2)); hs.maxY[i] = max(r.Y0, max(r.Y1, r.Y2));
    uint32_t bias12 = 0;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const int a = (k + 1) % 3, b = (k + 2) % 3;
        const int32_t A = Y[b] - Y[a], B = -(X[b] - X[a]);
        const int32_t bias = ((A > 0) || (A == 0 && B > 0)) ? 0 : 1;   // top-left rule (Appendix B.7)
        hs.A[k][i] = A; hs.B[k][i] = B;
        hs.E[k][i] = (long long)A * (128ll - X[a]) + (long long)B * (128ll - Y[a]) - bias;
        if (k > 0) bias12 |= (uint32_t)bias << (k - 1);
    }
    const long long A2 = (long long)(X[1] - X[0]) * (long long)(Y[2] - Y[0]) - (long long)(X[2] - X[0]) * (long long)(Y[1] - Y[0]);
    hs.inv[i] = 1.0f / (float)(-A2);
    hs.z0[i] = r.z0; hs.dz1[i] = r.z1 - r.z0; hs.dz2[i] = r.z2 - r.z0;
    hs.idKey[i] = ~r.triId; hs.bias12[i] = bias12;
}

// Merge the broadcast list into this lane's SUB keys; (x, y) = absolute pixel of sample 0, block = [bx0, bx0 + BLK_W) x
// [by0, by0 + BLK_ROWS * SUB).  One lane per record decides whether the record can touch the block at all.
template <bool SHADOW, int SUB>
__device__ __forceinline__ void merge_huge(const HugeStage& hs, uint32_t nHuge, int bx0, int by0, int x, int y, typename KeyOf<SHADOW>::type (&key)[SUB], unsigned lane) {
    const int32_t cx0 = 256 * bx0 + 128, cy0 = 256 * by0 + 128;
    const int32_t spanX = 256 * (BLK_W - 1), spanY = 256 * (BLK_ROWS * SUB - 1);
    if (nHuge <= (uint32_t)HUGE_DIRECT) {
        // short list (a ground plane is two triangles)
        if (!SHADOW) {
            // forward tile stage (one sample per lane): a warp-uniform bbox reject, then every lane tests its sample.  (The block
            // classification below costs this kernel registers it does not have: 80 + spills, tile stage 50 -> 60 us on C2.)
            for (uint32_t s = 0; s < nHuge; ++s) {
                if (hs.maxX[s] < cx0 || hs.minX[s] > cx0 + spanX || hs.maxY[s] < cy0 || hs.minY[s] > cy0 + spanY) continue;
                const long long F0 = hs.E[0][s] + mulw(hs.A[0][s], 256 * x) + mulw(hs.B[0][s], 256 * y);
                const long long F1 = hs.E[1][s] + mulw(hs.A[1][s], 256 * x) + mulw(hs.B[1][s], 256 * y);
                const long long F2 = hs.E[2][s] + mulw(hs.A[2][s], 256 * x) + mulw(hs.B[2][s], 256 * y);
                test_samples<SHADOW, SUB>(F0, F1, F2, mulw(hs.B[0][s], 256 * BLK_ROWS), mulw(hs.B[1][s], 256 * BLK_ROWS), mulw(hs.B[2][s], 256 * BLK_ROWS),
                                          hs.bias12[s], hs.inv[s], hs.z0[s], hs.dz1[s], hs.dz2[s], hs.idKey[s], key);
            }
            return;
        }
        // the block against every (triangle, edge) pair, one pair per lane (<= 3 * HUGE_DIRECT = 18 pairs): an edge function is
        // linear, so its extremes over the block's pixel centres are at the corners.  Outside one edge everywhere -> the triangle
        // is skipped (a screen-filling ground quad is two triangles: every block but the diagonal's is outside one of them);
        // inside all three everywhere -> no per-sample coverage test.  (Replaces a per-triangle bbox reject, which this subsumes.)
        bool edgeOut = false, edgeCut = false;
        if (lane < 3u * nHuge) {
            const uint32_t s = lane / 3u, k = lane - 3u * s;
            const int32_t A = hs.A[k][s], B = hs.B[k][s];
            const long long E0 = hs.E[k][s] + mulw(A, 256 * bx0) + mulw(B, 256 * by0);
            const long long dA = mulw(A, spanX), dB = mulw(B, spanY);
            edgeOut = E0 + (dA > 0 ? dA : 0ll) + (dB > 0 ? dB : 0ll) < 0;
            edgeCut = E0 + (dA < 0 ? dA : 0ll) + (dB < 0 ? dB : 0ll) < 0;
        }
        const unsigned mOut = __ballot_sync(0xFFFFFFFFu, edgeOut), mCut = __ballot_sync(0xFFFFFFFFu, edgeCut);
        for (uint32_t s = 0; s < nHuge; ++s) {
            const bool outside = ((mOut >> (3u * s)) & 7u) != 0u, inside = ((mCut >> (3u * s)) & 7u) == 0u;
            if (outside) continue;
            const long long F1 = hs.E[1][s] + mulw(hs.A[1][s], 256 * x) + mulw(hs.B[1][s], 256 * y);
            const long long F2 = hs.E[2][s] + mulw(hs.A[2][s], 256 * x) + mulw(hs.B[2][s], 256 * y);
            if (inside) {
                test_samples_inside<SHADOW, SUB>(F1, F2, mulw(hs.B[1][s], 256 * BLK_ROWS), mulw(hs.B[2][s], 256 * BLK_ROWS),
                                                 hs.bias12[s], hs.inv[s], hs.z0[s], hs.dz1[s], hs.dz2[s], hs.idKey[s], key);
                continue;
            }
            const long long F0 = hs.E[0][s] + mulw(hs.A[0][s], 256 * x) + mulw(hs.B[0][s], 256 * y);
            test_samples<SHADOW, SUB>(F0, F1, F2, mulw(hs.B[0][s], 256 * BLK_ROWS), mulw(hs.B[1][s], 256 * BLK_ROWS), mulw(hs.B[2][s], 256 * BLK_ROWS),
                                      hs.bias12[s], hs.inv[s], hs.z0[s], hs.dz1[s], hs.dz2[s], hs.idKey[s], key);
        }
        return;
    }
    for (uint32_t base = 0; base < nHuge; base += 32u) {
        const uint32_t i = base + lane;
        bool live = false;
        if (i < nHuge) {
            live = !(hs.maxX[i] < cx0 || hs.minX[i] > cx0 + spanX || hs.maxY[i] < cy0 || hs.minY[i] > cy0 + spanY);
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                const int32_t A = hs.A[k][i], B = hs.B[k][i];
                const long long Emax = hs.E[k][i] + mulw(A, 256 * bx0) + mulw(B, 256 * by0) + (A > 0 ? mulw(A, spanX) : 0ll) + (B > 0 ? mulw(B, spanY) : 0ll);
                live = live && (Emax >= 0);
            }
        }
        unsigned m = __ballot_sync(0xFFFFFFFFu, live);
        while (m) {
            const int s = (int)base + __ffs(m) - 1;
            m &= m - 1;
            const long long F0 = hs.E[0][s] + mulw(hs.A[0][s], 256 * x) + mulw(hs.B[0][s], 256 * y);
            const long long F1 = hs.E[1][s] + mulw(hs.A[1][s], 256 * x) + mulw(hs.B[1][s], 256 * y);
            const long long F2 = hs.E[2][s] + mulw(hs.A[2][s], 256 * x) + mulw(hs.B[2][s], 256 * y);
            test_samples<SHADOW, SUB>(F0, F1, F2, mulw(hs.B[0][s], 256 * BLK_ROWS), mulw(hs.B[1][s], 256 * BLK_ROWS), mulw(hs.B[2][s], 256 * BLK_ROWS), hs.bias12[s], hs.inv[s], hs.z0[s], hs.dz1[s], hs.dz2[s], hs.idKey[s], key);
        }
    }
}

// Shadow pass tile stage: only blocks touched by large triangles do work (the small ones were resolved by K1's
// atomics).  Persistent CTAs stride over 32 x 32 px regions (the status words and the broadcast list are fetched
// once per CTA, so a region costs no dependent memory round trip); warp = 8 x 16 px (SHADOW_SUB samples per lane).
// The map is only written.
#ifdef SHADOW_SUB_D
constexpr int SHADOW_SUB = SHADOW_SUB_D;
#else
constexpr int SHADOW_SUB = 4;
#endif
constexpr int SHADOW_REGION_H = BLK_ROWS * SHADOW_SUB * 2;

// Head of both tile-stage kernels: mirror the pass status words into mapped pinned memory for the host (overflow detection
// without a device->host copy node); the binned path of this pass has completed by now (k_tile_binned).
__device__ __forceinline__ void kt_head(const PassConst& pc, uint32_t& nHuge) {
    if (blockIdx.x == 0 && threadIdx.x < 8 && pc.hostStatus) pc.hostStatus[threadIdx.x] = __ldcg(pc.status + threadIdx.x);
    nHuge = min(__ldcg(pc.status + 4), (uint32_t)HUGE_CAP);
}

// ---------------------------------------------------------------------------------------------------------------------------------
// K2 `k_tile_binned`: the tile rasteriser of the binned path.  Leaves at once when K1 queued nothing (the usual frame of small
// triangles: one near-empty launch per pass).  Otherwise every warp of the persistent grid runs bin_phase (reb_bin.cuh: Appendix
// B.3 clipping of queued parents -> per-tile list layout -> scatter of the 48-B records; every phase's work is claimed from a
// global counter and nothing waits for a CTA that is not running), then every WARP strides over the 32 x 8 tiles with a non-empty list (lane = pixel column, 8 samples per lane): it streams
// the tile's records through its two landing buffers (cp.async.bulk + mbarrier; the next chunk — of this list or of the warp's next
// tile — lands while the current one is tested), tests them against its samples with the visibility keys in registers (top-left rule, early-Z and later-primitive-wins in one min-key compare) and merges the winner of every covered sample
// into the pass's target with ONE atomicMin — the 64-bit visibility key (forward) or the depth bits of the map (shadow), exactly
// what K1 does for the triangles it rasterises itself.  The shading / clear / broadcast-list stage (k_tile_forward, k_tile_shadow)
// therefore carries no list code, no staging memory and none of this kernel's registers.
// ---------------------------------------------------------------------------------------------------------------------------------
// the warp's next tile (stride nWarps) with a non-empty list, the list clamped to the capacity; t >= nTiles: none left
__device__ __forceinline__ uint32_t k2_next_tile(const PassConst& pc, uint32_t t, uint32_t stride, uint32_t nTiles, uint32_t& cnt, uint32_t& start) {
    cnt = 0; start = 0;
    for (; t < nTiles; t += stride) {
        const uint2 h = __ldcg(reinterpret_cast<const uint2*>(pc.tileHead + t));   // {count, start}
        if (!h.x || h.y >= pc.listCap) continue;            // empty, or an overflowed list (the host regrows and replays the frame)
        start = h.y; cnt = min(h.x, pc.listCap - h.y);
        REB_ASSERT((unsigned long long)start + cnt <= pc.listCap);
        break;
    }
    return t;
}

template <bool SHADOW>
__global__ void __launch_bounds__(KT_THREADS, 3) k_tile_binned(const __grid_constant__ PassConst pc) {
    __shared__ WarpStage stage[KT_WARPS];
    __shared__ WarpRecs recs[KT_WARPS];
    __shared__ uint32_t claim;
    pdl_enter();
    if ((__ldcg(pc.status + 0) | __ldcg(pc.status + 6)) == 0u) return;   // grid-uniform: nothing has touched these words since K1 ended
    bin_phase<SHADOW>(pc, &claim);
    if (min(ld_acquire_u32(pc.status + 0), pc.largeCap) == 0u) return;   // (final: this warp has seen every phase complete)
    typedef typename KeyOf<SHADOW>::type Key;
    constexpr Key KEY_NONE = (Key)~0ull;
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    WarpStage& ws = stage[warp];
    WarpRecs& wr = recs[warp];
    recs_init(wr, lane);
    // one WARP per 32 x 8 tile: lane = pixel column, TILE_H samples per lane (one per row); tiles strided over all warps of the grid
    const uint32_t nTiles = pc.tilesX * pc.tilesY, nWarps = (gridDim.x * KT_THREADS) >> 5;
    uint32_t cnt, start, ncnt, nstart;
    uint32_t tile = k2_next_tile(pc, blockIdx.x * KT_WARPS + warp, nWarps, nTiles, cnt, start);
    if (tile >= nTiles) return;
    uint32_t phase = 0u;
    int buf = 0;
    recs_issue(wr, 0, pc.tileRecs + start, min(32u, cnt), lane);
    while (tile < nTiles) {
        const uint32_t next = k2_next_tile(pc, tile + nWarps, nWarps, nTiles, ncnt, nstart);   // its loads overlap the copy in flight
        const uint32_t ty = tile / pc.tilesX, tx = tile - ty * pc.tilesX;
        const int bx0 = (int)tx * TILE_W, by0 = (int)ty * TILE_H;
        Key key[TILE_H];
#pragma unroll
        for (int r = 0; r < TILE_H; ++r) key[r] = KEY_NONE;
        for (uint32_t off = 0; off < cnt; off += 32u, buf ^= 1) {
            const uint32_t n = min(32u, cnt - off);
            // the chunk after this one: the rest of this list, else the head of the next tile's list
            if (off + 32u < cnt) recs_issue(wr, buf ^ 1, pc.tileRecs + start + off + 32u, min(32u, cnt - off - 32u), lane);
            else if (next < nTiles) recs_issue(wr, buf ^ 1, pc.tileRecs + nstart, min(32u, ncnt), lane);
            bool live = false;
#ifndef REB_NO_TMA
            mbar_wait(&wr.bar[buf], (phase >> buf) & 1u);
            phase ^= 1u << buf;
            if (lane < n) live = stage_record(wr.rec[buf][lane], bx0, by0, TILE_W, TILE_H, ws, lane);   // raw record -> edge equations (one per lane)
#else
            if (lane < n) {   // A/B build: one dependent 48-B load per lane instead of the bulk copy
                const uint4* q = reinterpret_cast<const uint4*>(pc.tileRecs + start + off + lane);
                union { LargeRec r; uint4 v[3]; } u;
                u.v[0] = __ldcg(q); u.v[1] = __ldcg(q + 1); u.v[2] = __ldcg(q + 2);
                live = stage_record(u.r, bx0, by0, TILE_W, TILE_H, ws, lane);
            }
#endif
            unsigned m = __ballot_sync(0xFFFFFFFFu, live);
            __syncwarp();
            const int32_t dx = 256 * (int)lane;
            while (m) {
                const int s = __ffs(m) - 1;
                m &= m - 1;
                const long long F0 = ws.E[0][s] + mulw(ws.A[0][s], dx);
                const long long F1 = ws.E[1][s] + mulw(ws.A[1][s], dx);
                const long long F2 = ws.E[2][s] + mulw(ws.A[2][s], dx);
                test_samples<SHADOW, TILE_H>(F0, F1, F2, mulw(ws.B[0][s], 256), mulw(ws.B[1][s], 256), mulw(ws.B[2][s], 256), ws.bias12[s], ws.inv[s], ws.z0[s], ws.dz1[s],
                                             ws.dz2[s], ws.idKey[s], key);
            }
            __syncwarp();
        }
        // merge the winners into the pass target: one atomicMin per covered sample, 32 consecutive samples per row
        const int x = bx0 + (int)lane;
        if (x < (int)pc.W) {
#pragma unroll
            for (int r = 0; r < TILE_H; ++r) {
                const int y = by0 + r;
                if (key[r] == KEY_NONE || y >= (int)pc.H) continue;
                if (SHADOW) atomicMin(pc.depthBits + (size_t)y * pc.W + x, (uint32_t)key[r]);
                else atomicMin(pc.vis + (size_t)y * pc.W + x, (unsigned long long)key[r]);
            }
        }
        tile = next; cnt = ncnt; start = nstart;
    }
}

__global__ void __launch_bounds__(KT_THREADS) k_tile_shadow(const __grid_constant__ PassConst pc) {
    __shared__ HugeStage hs;
    pdl_enter();
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    uint32_t nHuge;
    kt_head(pc, nHuge);
    if (nHuge == 0) return;
    stage_huge(pc.huge, nHuge, hs);
    __syncthreads();   // once per CTA, CTA-uniform
    const uint32_t regionsY = (pc.H + SHADOW_REGION_H - 1) / SHADOW_REGION_H, nRegions = pc.tilesX * regionsY;
    const uint32_t stepY = gridDim.x / pc.tilesX, stepX = gridDim.x - stepY * pc.tilesX;
    uint32_t ry = blockIdx.x / pc.tilesX, rx = blockIdx.x - ry * pc.tilesX;
    const int lx = (int)(lane & 7u), ly = (int)(lane >> 3);
    for (uint32_t region = blockIdx.x; region < nRegions; region += gridDim.x) {
        const int bx0 = (int)rx * TILE_W + BLK_W * (int)(warp & 3u);
        const int by0 = (int)ry * SHADOW_REGION_H + BLK_ROWS * SHADOW_SUB * (int)(warp >> 2);
        rx += stepX; ry += stepY;
        if (rx >= pc.tilesX) { rx -= pc.tilesX; ++ry; }
        if (bx0 >= (int)pc.W || by0 >= (int)pc.H) continue;   // warp-uniform
        if (pc.nRanks > 1 && !row_owned((uint32_t)by0, pc.rank, pc.nRanks, pc.bandH)) continue;   // shadow slabs (bandH % 32 == 0), warp-uniform
        const int x = bx0 + lx;
        uint32_t key[SHADOW_SUB];
#pragma unroll
        for (int r = 0; r < SHADOW_SUB; ++r) key[r] = 0xFFFFFFFFu;
        merge_huge<true, SHADOW_SUB>(hs, nHuge, bx0, by0, x, by0 + ly, key, lane);
#pragma unroll
        for (int r = 0; r < SHADOW_SUB; ++r) {
            const int y = by0 + BLK_ROWS * r + ly;
            if (key[r] != 0xFFFFFFFFu && x < (int)pc.W && y < (int)pc.H) atomicMin(pc.depthBits + (size_t)y * pc.W + x, key[r]);
        }
    }
}

// Forward pass tile stage: consume (and reset) the visibility keys, merge large triangles, shade, write.
// Persistent, warp-autonomous: a warp draws 32 x 8 tiles from a global counter (the draw for the tile after next is
// in flight while the current one is processed) and walks the tile's eight 8 x 4 blocks; the visibility keys of the
// next block are loaded before the current block is shaded.
#ifndef KT_MIN_BLOCKS
#define KT_MIN_BLOCKS 3
#endif
__device__ __forceinline__ unsigned long long kt_load_key(const PassConst& pc, uint32_t tx, uint32_t ty, uint32_t b, int lx, int ly) {
    const int x = (int)tx * TILE_W + BLK_W * (int)(b & 3u) + lx, y = (int)ty * TILE_H + BLK_ROWS * (int)(b >> 2) + ly;
    if (x >= (int)pc.W || y >= (int)pc.H) return VIS_CLEAR;   // callers visit owned tile rows only
    return __ldcg(pc.vis + (size_t)y * pc.W + x);
}

// One 8 x 4 block: re-arm vis, merge the broadcast list, shade, write.  All arguments but key are warp-uniform.
template <bool TEXTURED, bool PBR, bool AFFINE>
__device__ __forceinline__ void kt_block(const PassConst& pc, const ShadeConst& sh, unsigned long long key0, int bx0, int by0, int lx, int ly, uint32_t nHuge,
                                         const HugeStage& hs, unsigned lane) {
    if (bx0 >= (int)pc.W || by0 >= (int)pc.H) return;   // warp-uniform
    unsigned long long key[1] = {key0};
    const int x = bx0 + lx, y = by0 + ly;
    const bool inb = x < (int)pc.W && y < (int)pc.H;
    const size_t p = (size_t)y * pc.W + x;
    REB_ASSERT(!inb || p < (size_t)pc.W * pc.H);
    if (inb && key[0] != VIS_CLEAR) pc.vis[p] = VIS_CLEAR;   // leave the buffer cleared for the next frame
    if (nHuge) merge_huge<false, 1>(hs, nHuge, bx0, by0, x, y, key, lane);
    if (!inb) return;
    float depth = 1.0f;                    // depth clear 1.0 (RenderScene.cpp:219)
    uint32_t tri = TRI_NONE;
    float4 col = make_float4(0.0f, 0.0f, 0.0f, 1.0f);   // colour clear (0,0,0,1) (RenderScene.cpp:218)
    const bool hit = key[0] != VIS_CLEAR;
    if (hit) {
        depth = __uint_as_float((uint32_t)(key[0] >> 32));
        tri = ~(uint32_t)key[0];
    }
    if (sh.skyFaces && depth == 1.0f) col = sky_pixel(sh, x, y);   // the skybox is drawn last at depth 1.0 under LessOrEqual (Appendix B.14)
    else if (hit && (sh.outSRGB8 || sh.outF32)) col = PBR ? shade_pixel_pbr<TEXTURED, AFFINE>(pc, sh, x, y, tri) : shade_pixel<TEXTURED, AFFINE>(pc, sh, x, y, tri);
    if (sh.outDepth) sh.outDepth[p] = depth;
    if (sh.outTriId) sh.outTriId[p] = tri;
    if (sh.outSRGB8) reinterpret_cast<uint32_t*>(sh.outSRGB8)[p] = pack_srgb8(col);
    if (sh.outF32) reinterpret_cast<float4*>(sh.outF32)[p] = col;
}

#ifndef KT_PBR_MIN_BLOCKS
#define KT_PBR_MIN_BLOCKS 3
#endif
template <bool TEXTURED, bool PBR, bool AFFINE, bool PART>
__global__ void __launch_bounds__(KT_THREADS, PBR ? KT_PBR_MIN_BLOCKS : KT_MIN_BLOCKS) k_tile_forward(const __grid_constant__ PassConst pc, const __grid_constant__ ShadeConst sh) {
    __shared__ HugeStage hs;
    pdl_enter();
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const int lx = (int)(lane & 7u), ly = (int)(lane >> 3);
    uint32_t nHuge;
    kt_head(pc, nHuge);
    if (!PART) {
        // persistent CTAs, static striding over tiles, warp = block, next tile's key prefetched
        const uint32_t nTiles = pc.tilesX * pc.tilesY;
        uint32_t tile = blockIdx.x;
        const uint32_t stepY = gridDim.x / pc.tilesX, stepX = gridDim.x - stepY * pc.tilesX;   // the only divisions: once per CTA
        uint32_t tyN = tile / pc.tilesX, txN = tile - tyN * pc.tilesX;
        unsigned long long keyNext = tile < nTiles ? kt_load_key(pc, txN, tyN, warp, lx, ly) : VIS_CLEAR;
        if (nHuge) { stage_huge(pc.huge, nHuge, hs); __syncthreads(); }
        for (; tile < nTiles; tile += gridDim.x) {
            const unsigned long long key = keyNext;
            const uint32_t tx = txN, ty = tyN;
            txN += stepX; tyN += stepY;
            if (txN >= pc.tilesX) { txN -= pc.tilesX; ++tyN; }
            if (tile + gridDim.x < nTiles) keyNext = kt_load_key(pc, txN, tyN, warp, lx, ly);
            kt_block<TEXTURED, PBR, AFFINE>(pc, sh, key, (int)tx * TILE_W + BLK_W * (int)(warp & 3u), (int)ty * TILE_H + BLK_ROWS * (int)(warp >> 2), lx, ly, nHuge, hs, lane);
        }
    } else {
        // sort-first split: the same loop over the tile rows of this rank's bands only ("virtual" rows vy = the owned rows in
        // order; row vy is real row vy + (vy / bandRows) * (nRanks - 1) * bandRows + rank * bandRows), so a rank's tile stage
        // costs what its share of the frame costs — visiting every tile to skip most of them was a fixed ~0.2 ms at 8K
        uint32_t vtile = blockIdx.x;
        const uint32_t nVirtual = pc.tilesX * pc.vTilesY;
        const uint32_t stepY = gridDim.x / pc.tilesX, stepX = gridDim.x - stepY * pc.tilesX;
        const uint32_t bandSkip = (pc.nRanks - 1u) * pc.bandRows, bandBase = pc.rank * pc.bandRows;
        uint32_t vyN = vtile / pc.tilesX, txN = vtile - vyN * pc.tilesX;
        uint32_t tyN = vyN + (vyN / pc.bandRows) * bandSkip + bandBase;
        unsigned long long keyNext = vtile < nVirtual ? kt_load_key(pc, txN, tyN, warp, lx, ly) : VIS_CLEAR;
        if (nHuge) { stage_huge(pc.huge, nHuge, hs); __syncthreads(); }
        for (; vtile < nVirtual; vtile += gridDim.x) {
            const unsigned long long key = keyNext;
            const uint32_t tx = txN, ty = tyN;
            txN += stepX; vyN += stepY;
            if (txN >= pc.tilesX) { txN -= pc.tilesX; ++vyN; }
            tyN = vyN + (vyN / pc.bandRows) * bandSkip + bandBase;
            if (vtile + gridDim.x < nVirtual) keyNext = kt_load_key(pc, txN, tyN, warp, lx, ly);
            kt_block<TEXTURED, PBR, AFFINE>(pc, sh, key, (int)tx * TILE_W + BLK_W * (int)(warp & 3u), (int)ty * TILE_H + BLK_ROWS * (int)(warp >> 2), lx, ly, nHuge, hs, lane);
        }
    }
}

// Multisample resolve (Appendix B.15; the reference's base pass is multisampled with an average resolve,
// Swapchain.cpp:110-123,212).  The four sample planes were rasterised by four runs of K1..K3 with shifted vertices; here
// every pixel shades each DISTINCT primitive among its samples once at the pixel centre (one fragment-shader invocation per
// pixel and primitive), gives every sample the colour of its primitive (sky / clear colour for uncovered samples) and
// averages (s0 + s1) + (s2 + s3) in linear float.  Warp = 8 x 4 px block as in the tile stage (texture / shadow-map locality);
// the block's (pixel, primitive) pairs are compacted through shared memory so the shading rounds run with full lanes.
__device__ __forceinline__ uint32_t warp_incl_scan_u32(uint32_t v, unsigned lane) {
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const uint32_t n = __shfl_up_sync(0xFFFFFFFFu, v, o); if (lane >= (unsigned)o) v += n; }
    return v;
}
struct ResolveStage {                 // per warp: the block's shading tasks, compacted, and their colours
    uint32_t task[128];               // bits 0-4 owner lane, 5-6 owner slot, 7 sky; the primitive is triOf[]
    uint32_t triOf[128];
    float4 colour[32][4];             // [owner lane][owner slot]
};
#ifndef KR_MIN_BLOCKS
#define KR_MIN_BLOCKS 3
#endif
template <bool TEXTURED, bool PBR, bool AFFINE>
__global__ void __launch_bounds__(KT_THREADS, KR_MIN_BLOCKS) k_msaa_resolve(const __grid_constant__ PassConst pc, const __grid_constant__ ShadeConst sh,
                                                                const float* __restrict__ msDepth, const uint32_t* __restrict__ msTriId, float4* __restrict__ msColour) {
    __shared__ ResolveStage stage[KT_WARPS];
    pdl_enter();
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    ResolveStage& rs = stage[warp];
    const uint32_t ty = blockIdx.x / pc.tilesX, tx = blockIdx.x - ty * pc.tilesX;
    if (!row_owned(ty * TILE_H, pc.rank, pc.nRanks, pc.bandH)) return;
    const int bx0 = (int)tx * TILE_W + BLK_W * (int)(warp & 3u), by0 = (int)ty * TILE_H + BLK_ROWS * (int)(warp >> 2);
    if (bx0 >= (int)pc.W || by0 >= (int)pc.H) return;   // warp-uniform
    const int x = bx0 + (int)(lane & 7u), y = by0 + (int)(lane >> 3);
    const bool inb = x < (int)pc.W && y < (int)pc.H;
    const size_t p = (size_t)y * pc.W + x, plane = (size_t)pc.W * pc.H;
    // 1. this pixel's samples -> distinct (primitive | sky | clear) slots; src[k] = slot whose colour sample k takes
    uint32_t tri[4]; bool sky[4]; int src[4]; uint32_t nTasks = 0, slotTask[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        tri[k] = inb ? __ldg(msTriId + k * plane + p) : TRI_NONE;
        sky[k] = inb && sh.skyFaces && __ldg(msDepth + k * plane + p) == 1.0f;   // the skybox is drawn last at depth 1.0 under LessOrEqual
        if (sky[k]) tri[k] = TRI_NONE;
        src[k] = k;
#pragma unroll
        for (int j = k - 1; j >= 0; --j)
            if (sky[j] == sky[k] && tri[j] == tri[k]) src[k] = src[j];
        slotTask[k] = 0xFFFFFFFFu;
        if (src[k] == k && (sky[k] || tri[k] != TRI_NONE)) slotTask[k] = nTasks++;   // clear-colour slots need no shading
    }
    // 2. compact the warp's tasks so the shading rounds run with full lanes (small triangles: most pixels hold 2-3 primitives)
    const uint32_t incl = warp_incl_scan_u32(nTasks, lane);
    const uint32_t base = incl - nTasks, total = __shfl_sync(0xFFFFFFFFu, incl, 31);
#pragma unroll
    for (int k = 0; k < 4; ++k)
        if (slotTask[k] != 0xFFFFFFFFu) {
            rs.task[base + slotTask[k]] = lane | ((uint32_t)k << 5) | (sky[k] ? 128u : 0u);
            rs.triOf[base + slotTask[k]] = tri[k];
        }
    __syncwarp();
    // 3. shade: one fragment-shader invocation per pixel and primitive, at the pixel centre
    for (uint32_t t0 = 0; t0 < total; t0 += 32u) {
        const uint32_t t = t0 + lane;
        if (t < total) {
            const uint32_t task = rs.task[t], owner = task & 31u, slot = (task >> 5) & 3u;
            const int ox = bx0 + (int)(owner & 7u), oy = by0 + (int)(owner >> 3);
            float4 c;
            if (task & 128u) c = sky_pixel(sh, ox, oy);
            else c = PBR ? shade_pixel_pbr<TEXTURED, AFFINE>(pc, sh, ox, oy, rs.triOf[t]) : shade_pixel<TEXTURED, AFFINE>(pc, sh, ox, oy, rs.triOf[t]);
            rs.colour[owner][slot] = c;
        }
    }
    __syncwarp();
    if (!inb) return;
    // 4. resolve: every sample takes its slot's colour; (s0 + s1) + (s2 + s3), then 1/4
    float4 col[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const bool shaded = sky[src[k]] || tri[src[k]] != TRI_NONE;
        col[k] = shaded ? rs.colour[lane][src[k]] : make_float4(0.0f, 0.0f, 0.0f, 1.0f);   // colour clear (RenderScene.cpp:218)
    }
    if (msColour) {   // a translucent list follows inside this multisampled pass (Appendix B.16): every sample stays its own blend target
#pragma unroll
        for (int k = 0; k < 4; ++k) msColour[k * plane + p] = col[k];
        return;       // k_tl_resolve_ms resolves, draws the sky and encodes
    }
    float4 out;
    out.x = ((col[0].x + col[1].x) + (col[2].x + col[3].x)) * 0.25f;
    out.y = ((col[0].y + col[1].y) + (col[2].y + col[3].y)) * 0.25f;
    out.z = ((col[0].z + col[1].z) + (col[2].z + col[3].z)) * 0.25f;
    out.w = ((col[0].w + col[1].w) + (col[2].w + col[3].w)) * 0.25f;
    if (sh.outDepth) sh.outDepth[p] = __ldg(msDepth + p);      // depth / triangle ID targets: sample 0
    if (sh.outTriId) sh.outTriId[p] = __ldg(msTriId + p);
    if (sh.outSRGB8) reinterpret_cast<uint32_t*>(sh.outSRGB8)[p] = pack_srgb8(out);
    if (sh.outF32) reinterpret_cast<float4*>(sh.outF32)[p] = out;
}

void launch_msaa_resolve(const PassConst& pc, const ShadeConst& sh, const float* msDepth, const uint32_t* msTriId, float4* msColour, cudaStream_t st) {
    const bool tex = sh.nTextures && pc.vtxUV, pbr = sh.shadingModel == 1u;
    const bool affine = sh.STR[3] == 0.0f && sh.STR[7] == 0.0f && sh.STR[11] == 0.0f && sh.STR[15] == 1.0f;
    using Kern = void (*)(PassConst, ShadeConst, const float*, const uint32_t*, float4*);
    static const Kern kern[8] = {k_msaa_resolve<false, false, false>, k_msaa_resolve<true, false, false>, k_msaa_resolve<false, true, false>, k_msaa_resolve<true, true, false>,
                                 k_msaa_resolve<false, false, true>,  k_msaa_resolve<true, false, true>,  k_msaa_resolve<false, true, true>,  k_msaa_resolve<true, true, true>};
    launch_k(kern[(tex ? 1 : 0) | (pbr ? 2 : 0) | (affine ? 4 : 0)], dim3(pc.tilesX * pc.tilesY), dim3(KT_THREADS), st, true, pc, sh, msDepth, msTriId, msColour);
}

void launch_tile_binned(const PassConst& pc, bool shadow, int numSMs, cudaStream_t st) {
    static int perSM[2] = {0, 0};
    const int v = shadow ? 1 : 0;
    if (!perSM[v]) {
        if (shadow) cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSM[v], k_tile_binned<true>, KT_THREADS, 0);
        else cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSM[v], k_tile_binned<false>, KT_THREADS, 0);
        if (perSM[v] < 1) perSM[v] = 1;
    }
    // persistent; nothing in the kernel needs the CTAs to be resident at the same time (reb_bin.cuh)
    const uint32_t grid = std::max(1u, std::min<uint32_t>((pc.tilesX * pc.tilesY + KT_WARPS - 1) / KT_WARPS, (uint32_t)(numSMs * perSM[v])));
    if (shadow) launch_k(k_tile_binned<true>, dim3(grid), dim3(KT_THREADS), st, true, pc);
    else launch_k(k_tile_binned<false>, dim3(grid), dim3(KT_THREADS), st, true, pc);
}

void launch_tile_shadow(const PassConst& pc, int numSMs, cudaStream_t st) {
    static int perSM = 0;
    if (!perSM) { cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSM, k_tile_shadow, KT_THREADS, 0); if (perSM < 1) perSM = 1; }
    const uint32_t nRegions = pc.tilesX * ((pc.H + SHADOW_REGION_H - 1) / SHADOW_REGION_H);
    launch_k(k_tile_shadow, dim3(std::min<uint32_t>(nRegions, (uint32_t)(numSMs * perSM))), dim3(KT_THREADS), st, true, pc);
}

void launch_tile_forward(const PassConst& pc, const ShadeConst& sh, int numSMs, cudaStream_t st) {
    // the textured variants carry the uv derivatives + sampler state; untextured scenes keep the leaner kernels
    const bool tex = sh.nTextures && pc.vtxUV, pbr = sh.shadingModel == 1u;
    using Kern = void (*)(PassConst, ShadeConst);
    // shadow transform with last row exactly (0, 0, 0, 1) (every orthographic light): w == 1.0f exactly, shadow.glsl's three divisions are identities
    const bool affine = sh.STR[3] == 0.0f && sh.STR[7] == 0.0f && sh.STR[11] == 0.0f && sh.STR[15] == 1.0f;
#define KTF(T, P, A) k_tile_forward<T, P, A, false>, k_tile_forward<T, P, A, true>
    static const Kern kern[16] = {KTF(false, false, false), KTF(true, false, false), KTF(false, true, false), KTF(true, true, false),
                                  KTF(false, false, true),  KTF(true, false, true),  KTF(false, true, true),  KTF(true, true, true)};
#undef KTF
    const int v = 2 * ((tex ? 1 : 0) | (pbr ? 2 : 0) | (affine ? 4 : 0)) + (pc.nRanks > 1 ? 1 : 0);
    static int perSM[16] = {0};
    if (!perSM[v]) {
        int n = 0;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kern[v], KT_THREADS, 0);
        perSM[v] = n > 0 ? n : 1;
    }
    const uint32_t grid = std::max(1u, std::min<uint32_t>(pc.tilesX * pc.vTilesY, (uint32_t)(numSMs * perSM[v])));
    launch_k(kern[v], dim3(grid), dim3(KT_THREADS), st, true, pc, sh);
}

}  // namespace reb
