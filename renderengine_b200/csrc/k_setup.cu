// k_setup.cu — K1: triangle setup + small-triangle raster + large-triangle queue / per-tile counts; clears; AoS -> SoA split.
// (Clipping of queued parents, list layout and the record scatter run at the head of the tile stage: reb_bin.cuh.)
//
// Replaces, for the hot path, the fixed-function front end that the reference reaches through
// vkCmdDrawIndexed (MaterialSystem.cpp:99, ShadowSystem.cpp:369) with the state of
// GraphicsPipelineCache.cpp:257-365: viewport, cull back + CCW front (F5), 1 sample per pixel,
// LessOrEqual depth.  The vertex stage itself is k_vertex.cu.
#include "reb_bin.cuh"
#include "reb_kernels.h"
#include <algorithm>
#include <cstdlib>

namespace reb {

// ---------------------------------------------------------------------------------------------
// K1: triangle setup + small-triangle raster, WARP-AUTONOMOUS: every warp is an independent persistent
// worker with private shared-memory queues, so there is no CTA barrier anywhere in the steady state and
// the load latency of one warp is hidden by the raster arithmetic of the others.
//
// Per batch of 32 triangles (one per lane; batches are strided statically over all resident warps, the
// cache gathers of batch n+1 are issued before batch n is rasterised):
//   1. gather 3 VtxA (16 B each) through the flattened slot stream, clip-flag logic, facing (int64
//      shoelace), pixel-centre bbox, classification:
//        small (spans <= SMALL_LIMIT px) -> edge equations appended to the warp's 64-slot record ring, one
//                                         entry per bbox ROW (two half rows when the bbox is wider than
//                                         SMALL_MAX + 1) appended to the warp's row ring, <= SMALL_MAX + 1
//                                         entries per lane and round;
//        huge  (> HUGE_TILES tiles)    -> short broadcast list tested by every tile;
//        large (everything else)       -> warp-aggregated append to the LargeRec queue; the warp walks the
//                                         triangle's tiles and bumps the per-tile counts.
//   2. while >= 32 rows are pending: one lane per row computes the row's coverage mask (three 32-bit edge
//      functions stepped incrementally, top-left bias folded in, warp-uniform trip count) and appends one
//      entry per covered sample to the warp's pixel queue.  Rows of sparse batches accumulate across
//      batches, so lanes stay busy even when most triangles are culled.
//   3. one lane per covered sample: depth (two IEEE divides, Appendix B.8) and one atomicMin — 64-bit key
//      depthBits<<32 | ~triId into the visibility buffer (forward) or the 32-bit depth bits into the
//      shadow map itself (shadow).  Consecutive lanes hold consecutive samples of a row.
// ---------------------------------------------------------------------------------------------
constexpr int K1_THREADS = 128;                  // 4 autonomous warps per CTA (24.6 KB of queues)
constexpr int K1_WARPS = K1_THREADS / 32;
constexpr int K1_RING = 64;                      // triangle records per warp (32 being appended + <= 32 with pending rows)
constexpr int K1_ROWQ = 512;                     // row ring: < 32 pending + 32 * (SMALL_MAX + 1) new
constexpr int K1_PIXQ = 512;                     // pixel ring: < 32 pending + 32 * (SMALL_MAX + 1) new

struct K1Warp {
    int32_t Ax[3][K1_RING];         // 256 * A_k  (step of the edge function per pixel in x)
    int32_t By[3][K1_RING];         // 256 * B_k  (step per pixel in y)
    int32_t C[3][K1_RING];          // unbiased edge value at the bbox origin's pixel centre
    uint32_t xy[K1_RING];           // x0 | y0 << 16
    uint32_t whb[K1_RING];          // w | h << 8 | top-left bias of edge k << (16 + k)
    float z0[K1_RING], dz1[K1_RING], dz2[K1_RING], inv[K1_RING];   // inv = 1 / |A2|
    uint32_t idKey[K1_RING];
    uint16_t rowq[K1_ROWQ];         // slot | entry << 6   (entry = row, or 2 * row + half when the bbox is wider than K1_SPLIT)
    uint16_t pixq[K1_PIXQ];         // slot | i << 6 | j << 11
};
struct K1Shared {
    K1Warp w[K1_WARPS];
};

template <bool SHIFT>
__device__ __forceinline__ void k1_classify(const PassConst& pc, bool valid, const uint4 q[3], K1Tri& t) {
    t.isSmall = t.isLarge = t.needsClip = false;
    t.A2 = 0; t.x0 = t.y0 = 0; t.x1 = t.y1 = -1;
#pragma unroll
    for (int k = 0; k < 3; ++k)   // SHIFT: + B.15 sample shift (its own instantiation: the 1-sample kernels keep their register budget)
        t.a[k] = {(int32_t)q[k].x + (SHIFT ? pc.shiftX : 0), (int32_t)q[k].y + (SHIFT ? pc.shiftY : 0), __uint_as_float(q[k].z), q[k].w};
    if (!valid) return;
    const uint32_t f0 = t.a[0].flags, f1 = t.a[1].flags, f2 = t.a[2].flags;
    if ((f0 & f1 & f2 & VF_PLANES) != 0u) return;                                   // Appendix B.2 trivial reject
    if (((f0 | f1 | f2) & VF_NEEDS_CLIP) != 0u) { t.needsClip = true; return; }    // Appendix B.3: clipped by the whole warp below
    const VtxA &a0 = t.a[0], &a1 = t.a[1], &a2 = t.a[2];
    t.A2 = (long long)(a1.X - a0.X) * (long long)(a2.Y - a0.Y) - (long long)(a2.X - a0.X) * (long long)(a1.Y - a0.Y);
    if (t.A2 >= 0) return;                                                          // Appendix B.6: front = CCW = A2 < 0
    const int32_t minX = min(a0.X, min(a1.X, a2.X)), maxX = max(a0.X, max(a1.X, a2.X));
    const int32_t minY = min(a0.Y, min(a1.Y, a2.Y)), maxY = max(a0.Y, max(a1.Y, a2.Y));
    t.x0 = max((minX + 127) >> 8, 0); t.x1 = min((maxX - 128) >> 8, (int32_t)pc.W - 1);
    t.y0 = max((minY + 127) >> 8, 0); t.y1 = min((maxY - 128) >> 8, (int32_t)pc.H - 1);
    if (t.x0 > t.x1 || t.y0 > t.y1) return;                                         // no pixel centre inside the bbox
    // unclamped spans keep the 32-bit edge math of the row stage exact
    t.isSmall = (maxX - minX) <= SMALL_LIMIT * 256 && (maxY - minY) <= SMALL_LIMIT * 256;
    t.isLarge = !t.isSmall;
}

// huge -> broadcast list; large -> queue + per-tile counts.  Called by all 32 lanes of the warp.
__device__ __forceinline__ void k1_emit_large(const PassConst& pc, K1Tri& t, uint32_t tri, unsigned lane) {
    if (t.isLarge) {
        const int ntiles = (t.x1 / TILE_W - t.x0 / TILE_W + 1) * (t.y1 / TILE_H - t.y0 / TILE_H + 1);
        if (ntiles > HUGE_TILES) {
            const uint32_t slot = atomicAdd(&pc.status[4], 1u);
            if (slot < (uint32_t)HUGE_CAP) { pc.huge[slot] = k1_large_rec(t, tri); t.isLarge = false; }
        }
    }
    unsigned lmask = __ballot_sync(0xFFFFFFFFu, t.isLarge);
    if (!lmask) return;
    const int leader = __ffs(lmask) - 1;
    uint32_t base = 0;
    if ((int)lane == leader) base = atomicAdd(&pc.status[0], (uint32_t)__popc(lmask));
    base = __shfl_sync(0xFFFFFFFFu, base, leader);
    const uint32_t slot = base + __popc(lmask & ((1u << lane) - 1u));
    if (t.isLarge) {
        if (slot < pc.largeCap) pc.large[slot] = k1_large_rec(t, tri);
        else atomicOr(&pc.status[2], OVF_LARGE);
    }
    const LargeRec mine = k1_large_rec(t, tri);
    while (lmask) {   // warp-cooperative tile counting, one large triangle at a time
        const int src = __ffs(lmask) - 1;
        lmask &= lmask - 1;
        LargeRec r;
        r.X0 = __shfl_sync(0xFFFFFFFFu, mine.X0, src); r.Y0 = __shfl_sync(0xFFFFFFFFu, mine.Y0, src);
        r.X1 = __shfl_sync(0xFFFFFFFFu, mine.X1, src); r.Y1 = __shfl_sync(0xFFFFFFFFu, mine.Y1, src);
        r.X2 = __shfl_sync(0xFFFFFFFFu, mine.X2, src); r.Y2 = __shfl_sync(0xFFFFFFFFu, mine.Y2, src);
        r.z0 = r.z1 = r.z2 = 0.0f; r.triId = 0u; r.pad0 = r.pad1 = 0u;   // the count needs the footprint only
        const uint32_t s = __shfl_sync(0xFFFFFFFFu, slot, src);
        if (s < pc.largeCap) tile_walk<false>(pc, r, lane);
    }
}

constexpr uint32_t K1_SPLIT = SMALL_MAX + 1;   // widest (half) row of the row stage; row entries appended per lane and round
static_assert(31 + 32 * (SMALL_MAX + 1) <= 512, "row / pixel rings: < 32 pending + 32 * (SMALL_MAX + 1) new");

__device__ __forceinline__ void k1_store_small(K1Warp& ws, const K1Tri& t, uint32_t tri, uint32_t slot) {
    const int32_t X[3] = {t.a[0].X, t.a[1].X, t.a[2].X}, Y[3] = {t.a[0].Y, t.a[1].Y, t.a[2].Y};
    const int32_t Px = 256 * t.x0 + 128, Py = 256 * t.y0 + 128;
    const uint32_t h = (uint32_t)(t.y1 - t.y0 + 1);
    uint32_t biasBits = 0;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const int a = (k + 1) % 3, b = (k + 2) % 3;     // edge a->b opposes vertex k
        const int32_t A = Y[b] - Y[a], B = -(X[b] - X[a]);
        const uint32_t bias = ((A > 0) || (A == 0 && B > 0)) ? 0u : 1u;   // top-left rule (Appendix B.7)
        ws.Ax[k][slot] = 256 * A;
        ws.By[k][slot] = 256 * B;
        ws.C[k][slot] = A * (Px - X[a]) + B * (Py - Y[a]);
        biasBits |= bias << (16 + k);
    }
    ws.xy[slot] = (uint32_t)t.x0 | ((uint32_t)t.y0 << 16);
    ws.whb[slot] = (uint32_t)(t.x1 - t.x0 + 1) | (h << 8) | biasBits;
    ws.z0[slot] = t.a[0].z; ws.dz1[slot] = t.a[1].z - t.a[0].z; ws.dz2[slot] = t.a[2].z - t.a[0].z;
    ws.inv[slot] = 1.0f / (float)(-t.A2);
    ws.idKey[slot] = ~tri;
}

// Cache slots of the three corners of triangle `tri` (flatSlots: the draw list's index stream re-based into
// cache-slot space in submission order, built when the draw list changes; de-indexed mode needs no table).
// CULL (sort-first split): triangles of a 256-triangle block whose box touches no row of this rank come back as the sentinel
// and are gathered as an all-zero (degenerate, hence culled) triangle; K0 skipped the vertices of wholly invisible draws, so in a
// block that spans draws the triangle's own draw is looked up.
constexpr uint32_t SLOT_CULLED = 0xFFFFFFFFu;
// `tri` is the position in K1's stream; .w of the result is the triangle's SUBMISSION-order ID (Appendix B.9) — the same number
// unless the draw list was re-ordered (k_reorder.cu), in which case one 16-B record holds the slots and the ID.
template <bool CULL>
__device__ __forceinline__ uint4 k1_load_slots(const PassConst& pc, uint32_t tri) {
    if (tri >= pc.nTris) return make_uint4(0u, 0u, 0u, tri);
    if (CULL) {
        const uint32_t vis = __ldg(pc.blockVisible + (tri >> 8));
        if (vis == 0u) return make_uint4(SLOT_CULLED, 0u, 0u, tri);
        // a block that spans draws: K0 did not run the vertices of a culled draw, so its triangles must not be gathered
        // (a re-ordered list keeps every triangle inside its draw's range of positions, so the lookup by position holds)
        if (vis == BLOCK_MIXED && __ldg(pc.drawVisible + find_draw_block<false>(pc.draws, pc.nDraws, pc.triBlockDraw, tri)) == 0u) return make_uint4(SLOT_CULLED, 0u, 0u, tri);
    }
    if (pc.triStream) return __ldg(pc.triStream + tri);
    if (pc.deindexed) return make_uint4(3u * tri, 3u * tri + 1u, 3u * tri + 2u, tri);
    const uint32_t* f = pc.flatSlots + 3ull * tri;
    return make_uint4(__ldg(f), __ldg(f + 1), __ldg(f + 2), tri);
}
template <bool CULL>
__device__ __forceinline__ void k1_gather(const PassConst& pc, uint32_t tri, uint4 s, uint4 q[3]) {
    if (tri >= pc.nTris || (CULL && s.x == SLOT_CULLED)) { q[0] = q[1] = q[2] = make_uint4(0, 0, 0, 0); return; }
    const uint4* va = reinterpret_cast<const uint4*>(pc.vtxA);
    REB_ASSERT(s.x < pc.cacheSize && s.y < pc.cacheSize && s.z < pc.cacheSize && s.w < pc.nTris);
    q[0] = __ldg(va + s.x); q[1] = __ldg(va + s.y); q[2] = __ldg(va + s.z);
}

// A "tiny" triangle — pixel-centre bbox of at most 4 x 4 and at most 15 centres (so a lane never emits more than the 15 samples a
// row entry can, and the ring bound 31 + 32 * 15 holds) — is ONE row-queue entry covering its whole bbox: the coverage mask is
// 2-D, bit i + 4 * j.  A frame of 1-3 px triangles (C2, C3) then costs one entry, one slot of the row stage and no scan in the
// append per triangle instead of one per bbox row.
__device__ __forceinline__ bool k1_is_tiny(uint32_t w, uint32_t h) { return w <= 4u && h <= 4u && w * h <= 15u; }

// Row stage: `cnt` (<= 32) pending entries, one per lane -> coverage masks -> pixel queue entries.
template <bool SHADOW, bool CULL = false>
__device__ __forceinline__ void k1_rows(const PassConst& pc, K1Warp& ws, uint32_t rowHead, uint32_t cnt, uint32_t& pixTail, unsigned lane) {
    bool active = lane < cnt;
    uint32_t slot = 0, j = 0, whb = 0, xo = 0;
    bool tiny = false;
    if (active) {
        const uint32_t e = ws.rowq[(rowHead + lane) & (K1_ROWQ - 1)];
        slot = e & 0x3Fu;
        whb = ws.whb[slot];
        tiny = k1_is_tiny(whb & 0xFFu, (whb >> 8) & 0xFFu);
        const uint32_t wide = (whb & 0xFFu) > K1_SPLIT ? 1u : 0u;   // two half rows per row
        j = (e >> 6) >> wide; xo = ((e >> 6) & wide) * K1_SPLIT;      // tiny: entry 0 -> j = 0, xo = 0
        // (sort-first / shadow slabs: a tiny entry spans up to 4 rows, which may straddle a band border: its rows are filtered below)
        if (!tiny && (!SHADOW || CULL) && !row_owned((ws.xy[slot] >> 16) + j, pc.rank, pc.nRanks, pc.bandH)) active = false;
    }
    const uint32_t w = active ? min((whb & 0xFFu) - xo, K1_SPLIT) : 0u;
    const uint32_t h = active ? (tiny ? ((whb >> 8) & 0xFFu) : 1u) : 0u;
    const uint32_t wmax = __reduce_max_sync(0xFFFFFFFFu, w);
    const uint32_t hmax = __reduce_max_sync(0xFFFFFFFFu, h);
    uint32_t mask = 0;
    {
        const int32_t s0 = ws.Ax[0][slot], s1 = ws.Ax[1][slot], s2 = ws.Ax[2][slot];
        int32_t F0 = ws.C[0][slot] + (int32_t)j * ws.By[0][slot] + (int32_t)xo * s0 - (int32_t)((whb >> 16) & 1u);   // top-left bias folded in
        int32_t F1 = ws.C[1][slot] + (int32_t)j * ws.By[1][slot] + (int32_t)xo * s1 - (int32_t)((whb >> 17) & 1u);
        int32_t F2 = ws.C[2][slot] + (int32_t)j * ws.By[2][slot] + (int32_t)xo * s2 - (int32_t)((whb >> 18) & 1u);
        if (hmax <= 1u) {   // warp-uniform: row entries only (or 1-row tiny ones)
#pragma unroll 1
            for (uint32_t i = 0; i < wmax; ++i, F0 += s0, F1 += s1, F2 += s2) mask |= ((F0 | F1 | F2) >= 0 ? 1u : 0u) << i;
            mask &= (1u << w) - 1u;
        } else {
            // rows 1 .. hmax - 1 exist for tiny entries only (<= 4 columns); a row entry in the same round is finished by row 0
            const int32_t b0 = ws.By[0][slot], b1 = ws.By[1][slot], b2 = ws.By[2][slot];
            const uint32_t wmaxT = __reduce_max_sync(0xFFFFFFFFu, tiny ? w : 0u);
#pragma unroll 1
            for (uint32_t i = 0; i < wmax; ++i) mask |= (((F0 + (int32_t)i * s0) | (F1 + (int32_t)i * s1) | (F2 + (int32_t)i * s2)) >= 0 ? 1u : 0u) << i;
            mask &= (1u << w) - 1u;
            if (tiny) {
#pragma unroll 1
                for (uint32_t jj = 1; jj < hmax; ++jj) {
                    F0 += b0; F1 += b1; F2 += b2;
#pragma unroll 1
                    for (uint32_t i = 0; i < wmaxT; ++i)
                        mask |= (((F0 + (int32_t)i * s0) | (F1 + (int32_t)i * s1) | (F2 + (int32_t)i * s2)) >= 0 ? 1u : 0u) << (4u * jj + i);
                }
                mask &= ((1u << w) - 1u) * (0x1111u & ((1u << (4u * h)) - 1u));
                if ((!SHADOW || CULL) && pc.nRanks > 1) {   // rows of the bbox this rank does not own
                    const uint32_t y0 = ws.xy[slot] >> 16;
#pragma unroll
                    for (uint32_t jj = 0; jj < 4u; ++jj)
                        if (!row_owned(y0 + jj, pc.rank, pc.nRanks, pc.bandH)) mask &= ~(0xFu << (4u * jj));
                }
            }
        }
    }
    if (hmax <= 1u && tiny && (!SHADOW || CULL) && pc.nRanks > 1 && !row_owned(ws.xy[slot] >> 16, pc.rank, pc.nRanks, pc.bandH)) mask = 0u;   // 1-row tiny entry
    const uint32_t n = __popc(mask);
    const uint32_t incl = warp_incl_scan(n, lane);
    uint32_t pos = pixTail + incl - n;
    pixTail += __shfl_sync(0xFFFFFFFFu, incl, 31);
    const uint32_t tag = slot | (xo << 6) | (j << 11);   // i + xo < 32: the add cannot carry into j
    while (mask) {
        const uint32_t bpos = (uint32_t)__ffs(mask) - 1u;
        mask &= mask - 1u;
        const uint32_t i = tiny ? (bpos & 3u) : bpos, jj = tiny ? (bpos >> 2) : 0u;
        ws.pixq[pos++ & (K1_PIXQ - 1)] = (uint16_t)(tag + (i << 6) + (jj << 11));
    }
    __syncwarp();
}

// Sample stage: `cnt` (<= 32) queued covered samples, one per lane: depth + resolve.
template <bool SHADOW>
__device__ __forceinline__ void k1_pixels(const PassConst& pc, const K1Warp& ws, uint32_t pixHead, uint32_t cnt, unsigned lane) {
    if (lane >= cnt) return;
    const uint32_t e = ws.pixq[(pixHead + lane) & (K1_PIXQ - 1)];
    const uint32_t slot = e & 0x3Fu;
    const int32_t pi = (int32_t)((e >> 6) & 0x1Fu), pj = (int32_t)(e >> 11);
    const int32_t u1 = ws.C[1][slot] + pj * ws.By[1][slot] + pi * ws.Ax[1][slot];   // unbiased edge values (Appendix B.8)
    const int32_t u2 = ws.C[2][slot] + pj * ws.By[2][slot] + pi * ws.Ax[2][slot];
    const float z = depth_from((float)u1, (float)u2, ws.inv[slot], ws.z0[slot], ws.dz1[slot], ws.dz2[slot]);
    if (!(z >= 0.0f && z <= 1.0f)) return;
    const uint32_t xy = ws.xy[slot];
    const size_t px = (size_t)((xy >> 16) + (uint32_t)pj) * pc.W + (xy & 0xFFFFu) + (uint32_t)pi;
    REB_ASSERT((xy >> 16) + (uint32_t)pj < pc.H && (xy & 0xFFFFu) + (uint32_t)pi < pc.W && slot < (uint32_t)K1_RING);
    if (SHADOW) atomicMin(pc.depthBits + px, __float_as_uint(z));
    else atomicMin(pc.vis + px, ((unsigned long long)__float_as_uint(z) << 32) | ws.idKey[slot]);
}

#ifndef K1_MIN_BLOCKS
#define K1_MIN_BLOCKS 6
#endif
template <bool SHADOW, bool SHIFT, bool CULL>
__global__ void __launch_bounds__(K1_THREADS, K1_MIN_BLOCKS) k_setup_raster(const __grid_constant__ PassConst pc) {
    __shared__ K1Shared sm;
    pdl_enter();
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    K1Warp& ws = sm.w[warp];
    const uint32_t bl = pc.batchLog2, tpb = 1u << bl;
    // CULL (sort-first split): batches are numbered over the compacted list of the 256-triangle blocks this rank has to look at
    // (k_cull_blocks), 2^(8 - bl) batches per block, instead of over all triangles
    const uint32_t nBatches = CULL ? (__ldcg(pc.cullCounts) << (8u - bl)) : ((pc.nTris + tpb - 1u) >> bl);
    const uint32_t totalWarps = gridDim.x * K1_WARPS;
    const uint32_t laneTri = lane < tpb ? lane : 0xFFFFFFFFu;   // lanes past the batch see triangle index >= nTris (nothing to do)
    auto batch_tri = [&](uint32_t b) -> uint32_t {
        if (laneTri == 0xFFFFFFFFu) return 0xFFFFFFFFu;
        if (!CULL) return (b << bl) + laneTri;
        if (b >= nBatches) return 0xFFFFFFFFu;
        REB_ASSERT(__ldg(pc.triBlockList + (b >> (8u - bl))) < (pc.nTris + 255u) / 256u);
        return (__ldg(pc.triBlockList + (b >> (8u - bl))) << 8) + ((b & ((1u << (8u - bl)) - 1u)) << bl) + laneTri;
    };
#define K1_TRI(b) batch_tri(b)

    // static schedule: warp g takes batches g, g + totalWarps, ... (a shared work counter would serialise on one
    // L2 atomic unit: ~10 cycles per same-address atomic-with-return chip-wide, measured as the top stall)
    uint32_t batch = blockIdx.x * K1_WARPS + warp;
    uint4 q[3];
    bool curCulled = false;   // CULL: every triangle of the current batch belongs to a culled draw (warp-uniform)
    uint32_t curId;           // submission-order ID of this lane's triangle of the current batch
    {
        const uint32_t tri = K1_TRI(batch);
        const uint4 s0 = k1_load_slots<CULL>(pc, tri);
        if (CULL) curCulled = __all_sync(0xFFFFFFFFu, s0.x == SLOT_CULLED || tri >= pc.nTris);
        k1_gather<CULL>(pc, tri, s0, q);
        curId = s0.w;
    }
    uint4 ns = k1_load_slots<CULL>(pc, K1_TRI(batch + totalWarps));   // slot stream runs two batches ahead of the raster stages
    uint32_t ringPos = 0, rowHead = 0, rowTail = 0, pixHead = 0, pixTail = 0;   // warp-uniform
    while (batch < nBatches) {
        const uint32_t nbatch = batch + totalWarps;
        const uint32_t tri = K1_TRI(batch), ntri = K1_TRI(nbatch);
        const uint4 nns = k1_load_slots<CULL>(pc, K1_TRI(nbatch + totalWarps));   // in flight for a whole iteration
        if (CULL && curCulled) {   // nothing of this batch is ours: only advance the software pipeline
            curCulled = __all_sync(0xFFFFFFFFu, ns.x == SLOT_CULLED || ntri >= pc.nTris);
            k1_gather<CULL>(pc, ntri, ns, q);
            curId = ns.w; ns = nns; batch = nbatch;
            continue;
        }

        // ---- 1. classify, emit large, append small records + rows ----------------------------------------
        K1Tri t;
        k1_classify<SHIFT>(pc, tri < pc.nTris, q, t);
        k1_emit_large(pc, t, curId, lane);
        if (t.needsClip) {   // Appendix B.3, rare: queued for the tile stage's bin_phase (keeps the clipper out of this kernel's register budget)
            const uint32_t qi = atomicAdd(&pc.status[6], 1u);
            if (qi < pc.clipCap) pc.clipQueue[qi] = curId; else atomicOr(&pc.status[2], OVF_LARGE);
        }
        const unsigned smallMask = __ballot_sync(0xFFFFFFFFu, t.isSmall);
        const uint32_t mySlot = (ringPos + __popc(smallMask & ((1u << lane) - 1u))) & (K1_RING - 1);
        // row entries of this lane's triangle: one per bbox row, two (half rows) when the bbox is wider than K1_SPLIT
        const bool isTiny = t.isSmall && k1_is_tiny((uint32_t)(t.x1 - t.x0 + 1), (uint32_t)(t.y1 - t.y0 + 1));   // one entry for the whole bbox
        const uint32_t nEnt = !t.isSmall ? 0u : (isTiny ? 1u : (uint32_t)(t.y1 - t.y0 + 1) << ((t.x1 - t.x0 + 1) > (int32_t)K1_SPLIT ? 1 : 0));
        if (t.isSmall) k1_store_small(ws, t, curId, mySlot);
        ringPos += __popc(smallMask);
        if (CULL) curCulled = __all_sync(0xFFFFFFFFu, ns.x == SLOT_CULLED || ntri >= pc.nTris);
        k1_gather<CULL>(pc, ntri, ns, q);                            // in flight during the raster stages
        curId = ns.w; ns = nns;

        // ---- 2./3. rows -> samples -> resolve; <= K1_SPLIT row entries per lane and round keep the rings within capacity ----
        uint32_t done = 0;
        const bool oneEach = __all_sync(0xFFFFFFFFu, nEnt <= 1u);   // a batch of tiny (or culled) triangles: ballot instead of scan + loop
        do {
            const uint32_t c = min(nEnt - done, K1_SPLIT);
            if (oneEach) {
                const unsigned em = __ballot_sync(0xFFFFFFFFu, c != 0u);
                if (c) ws.rowq[(rowTail + __popc(em & ((1u << lane) - 1u))) & (K1_ROWQ - 1)] = (uint16_t)mySlot;
                rowTail += __popc(em);
            } else {
                const uint32_t cIncl = warp_incl_scan(c, lane);
                const uint32_t rowPos = rowTail + cIncl - c;
                for (uint32_t j = 0; j < c; ++j) ws.rowq[(rowPos + j) & (K1_ROWQ - 1)] = (uint16_t)(mySlot | ((done + j) << 6));
                rowTail += __shfl_sync(0xFFFFFFFFu, cIncl, 31);
            }
            done += c;
            REB_ASSERT(rowTail - rowHead <= (uint32_t)K1_ROWQ && pixTail - pixHead <= (uint32_t)K1_PIXQ);
            __syncwarp();
            while (rowTail - rowHead >= 32u) {
                k1_rows<SHADOW, CULL>(pc, ws, rowHead, 32u, pixTail, lane);
                rowHead += 32u;
                while (pixTail - pixHead >= 32u) { k1_pixels<SHADOW>(pc, ws, pixHead, 32u, lane); pixHead += 32u; }
            }
        } while (SMALL_LIMIT > SMALL_MAX && __any_sync(0xFFFFFFFFu, done < nEnt));
        // queued samples reference ring slots of this batch or the one before: drain them before the ring advances
        if (pixTail != pixHead) { k1_pixels<SHADOW>(pc, ws, pixHead, pixTail - pixHead, lane); pixHead = pixTail; }
        __syncwarp();
        batch = nbatch;
    }
    if (rowTail != rowHead) {   // drain the rows left over from sparse batches
        k1_rows<SHADOW, CULL>(pc, ws, rowHead, rowTail - rowHead, pixTail, lane);
        while (pixTail != pixHead) { const uint32_t c = min(32u, pixTail - pixHead); k1_pixels<SHADOW>(pc, ws, pixHead, c, lane); pixHead += c; }
    }

    // clipped parents and the per-tile lists are finished at the head of the tile stage (bin_phase, reb_bin.cuh)
#undef K1_TRI
}

// ---------------------------------------------------------------------------------------------------------------------------------
// K1 for the 4x multisampled base pass (Appendix B.15) in ONE pass over the triangles.  Sampling at (x + sx, y + sy) equals sampling
// the pixel centre of the triangle moved by s = (128 - 256 sx, 128 - 256 sy) sub-pixel units, and an edge function of the moved
// triangle is the unmoved one plus the constant -(A sx' + B sy'): so gather, clip flags, facing, edge equations, 1 / |A2| and the
// ring record are computed once per triangle and the row stage tests every pixel centre of the UNION bbox against four sets of
// constants; samples go to the visibility plane of their sample index.  Triangles beyond the small class are appended to the
// queue set of every sample with that sample's moved vertices, so K2 / the tile stage of each sample pass run unchanged.
// (The first version ran the whole K1 once per sample: 4 x 26 us on C2.)
// ---------------------------------------------------------------------------------------------------------------------------------
__constant__ int32_t c_msShift[4][2] = {{32, 96}, {-96, 32}, {96, -32}, {-32, -96}};   // Vulkan standard 4-sample locations as vertex shifts
constexpr int32_t MS_PAD = 96;          // largest |shift|: the union bbox is the bbox of the unmoved triangle grown by this much
constexpr int MS_PIXQ = 1024;           // sample ring (32-bit entries): < 32 pending + 32 * 16 new per sample
// the union bbox of a 1-3 px triangle is 3 x 3 or 4 x 4 centres: the one-entry class takes the full 4 x 4 here
__device__ __forceinline__ bool k1ms_is_tiny(uint32_t w, uint32_t h) { return w <= 4u && h <= 4u; }

__device__ __forceinline__ void k1ms_classify(const PassConst& pc, bool valid, const uint4 q[3], K1Tri& t) {
    t.isSmall = t.isLarge = t.needsClip = false;
    t.A2 = 0; t.x0 = t.y0 = 0; t.x1 = t.y1 = -1;
#pragma unroll
    for (int k = 0; k < 3; ++k) t.a[k] = {(int32_t)q[k].x, (int32_t)q[k].y, __uint_as_float(q[k].z), q[k].w};
    if (!valid) return;
    const uint32_t f0 = t.a[0].flags, f1 = t.a[1].flags, f2 = t.a[2].flags;
    if ((f0 & f1 & f2 & VF_PLANES) != 0u) return;
    if (((f0 | f1 | f2) & VF_NEEDS_CLIP) != 0u) { t.needsClip = true; return; }
    const VtxA &a0 = t.a[0], &a1 = t.a[1], &a2 = t.a[2];
    t.A2 = (long long)(a1.X - a0.X) * (long long)(a2.Y - a0.Y) - (long long)(a2.X - a0.X) * (long long)(a1.Y - a0.Y);   // translation-invariant
    if (t.A2 >= 0) return;
    const int32_t minX = min(a0.X, min(a1.X, a2.X)) - MS_PAD, maxX = max(a0.X, max(a1.X, a2.X)) + MS_PAD;
    const int32_t minY = min(a0.Y, min(a1.Y, a2.Y)) - MS_PAD, maxY = max(a0.Y, max(a1.Y, a2.Y)) + MS_PAD;
    t.x0 = max((minX + 127) >> 8, 0); t.x1 = min((maxX - 128) >> 8, (int32_t)pc.W - 1);
    t.y0 = max((minY + 127) >> 8, 0); t.y1 = min((maxY - 128) >> 8, (int32_t)pc.H - 1);
    if (t.x0 > t.x1 || t.y0 > t.y1) return;                     // no pixel centre under any sample's bbox
    t.isSmall = (maxX - minX) <= SMALL_LIMIT * 256 && (maxY - minY) <= SMALL_LIMIT * 256;   // the grown spans: ring and 32-bit bounds as in K1
    t.isLarge = !t.isSmall;
}

// the triangle as sample k's pass sees it: moved vertices, that pass's own bbox (k1_classify's), large by construction
__device__ __forceinline__ bool k1ms_moved(const PassConst& pc, const K1Tri& t, int k, K1Tri& m) {
    m = t;
#pragma unroll
    for (int v = 0; v < 3; ++v) { m.a[v].X += c_msShift[k][0]; m.a[v].Y += c_msShift[k][1]; }
    const int32_t minX = min(m.a[0].X, min(m.a[1].X, m.a[2].X)), maxX = max(m.a[0].X, max(m.a[1].X, m.a[2].X));
    const int32_t minY = min(m.a[0].Y, min(m.a[1].Y, m.a[2].Y)), maxY = max(m.a[0].Y, max(m.a[1].Y, m.a[2].Y));
    m.x0 = max((minX + 127) >> 8, 0); m.x1 = min((maxX - 128) >> 8, (int32_t)pc.W - 1);
    m.y0 = max((minY + 127) >> 8, 0); m.y1 = min((maxY - 128) >> 8, (int32_t)pc.H - 1);
    m.isLarge = t.isLarge && m.x0 <= m.x1 && m.y0 <= m.y1;
    m.isSmall = false;
    return m.isLarge;
}

// Sample stage: entry = slot | i << 6 | j << 11 | sample << 16
__device__ __forceinline__ void k1ms_pixels(const PassConst& pc, const MsTargets& ms, const K1Warp& ws, const uint32_t* pixq, uint32_t pixHead, uint32_t cnt, unsigned lane) {
    if (lane >= cnt) return;
    const uint32_t e = pixq[(pixHead + lane) & (MS_PIXQ - 1)];
    const uint32_t slot = e & 0x3Fu, k = e >> 16;
    const int32_t pi = (int32_t)((e >> 6) & 0x1Fu), pj = (int32_t)((e >> 11) & 0x1Fu);
    const int32_t sx = c_msShift[k][0], sy = c_msShift[k][1];
    const int32_t c1 = -((ws.Ax[1][slot] >> 8) * sx + (ws.By[1][slot] >> 8) * sy), c2 = -((ws.Ax[2][slot] >> 8) * sx + (ws.By[2][slot] >> 8) * sy);
    const int32_t u1 = ws.C[1][slot] + pj * ws.By[1][slot] + pi * ws.Ax[1][slot] + c1;   // unbiased edge values of the moved triangle (Appendix B.8)
    const int32_t u2 = ws.C[2][slot] + pj * ws.By[2][slot] + pi * ws.Ax[2][slot] + c2;
    const float z = depth_from((float)u1, (float)u2, ws.inv[slot], ws.z0[slot], ws.dz1[slot], ws.dz2[slot]);
    if (!(z >= 0.0f && z <= 1.0f)) return;
    const uint32_t xy = ws.xy[slot];
    const size_t px = (size_t)((xy >> 16) + (uint32_t)pj) * pc.W + (xy & 0xFFFFu) + (uint32_t)pi;
    REB_ASSERT((xy >> 16) + (uint32_t)pj < pc.H && (xy & 0xFFFFu) + (uint32_t)pi < pc.W && k < 4u);
    atomicMin(ms.vis[k] + px, ((unsigned long long)__float_as_uint(z) << 32) | ws.idKey[slot]);
}

// Row stage: as k1_rows, but every entry yields one coverage mask per sample; the sample queue is drained between samples, so it
// needs no more room than the 1-sample kernel's.
template <bool CULL>
__device__ __forceinline__ void k1ms_rows(const PassConst& pc, const MsTargets& ms, K1Warp& ws, uint32_t* pixq, uint32_t rowHead, uint32_t cnt, uint32_t& pixHead,
                                          uint32_t& pixTail, unsigned lane) {
    bool active = lane < cnt;
    uint32_t slot = 0, j = 0, whb = 0, xo = 0;
    bool tiny = false;
    if (active) {
        const uint32_t e = ws.rowq[(rowHead + lane) & (K1_ROWQ - 1)];
        slot = e & 0x3Fu;
        whb = ws.whb[slot];
        tiny = k1ms_is_tiny(whb & 0xFFu, (whb >> 8) & 0xFFu);
        const uint32_t wide = (whb & 0xFFu) > K1_SPLIT ? 1u : 0u;
        j = (e >> 6) >> wide; xo = ((e >> 6) & wide) * K1_SPLIT;
        if (!tiny && !row_owned((ws.xy[slot] >> 16) + j, pc.rank, pc.nRanks, pc.bandH)) active = false;
    }
    const uint32_t w = active ? min((whb & 0xFFu) - xo, K1_SPLIT) : 0u;
    const uint32_t h = active ? (tiny ? ((whb >> 8) & 0xFFu) : 1u) : 0u;
    const uint32_t wmax = __reduce_max_sync(0xFFFFFFFFu, w);
    const uint32_t hmax = __reduce_max_sync(0xFFFFFFFFu, h);
    const int32_t s0 = ws.Ax[0][slot], s1 = ws.Ax[1][slot], s2 = ws.Ax[2][slot];
    const int32_t b0 = ws.By[0][slot], b1 = ws.By[1][slot], b2 = ws.By[2][slot];
    const int32_t E0 = ws.C[0][slot] + (int32_t)j * b0 + (int32_t)xo * s0 - (int32_t)((whb >> 16) & 1u);   // top-left bias folded in
    const int32_t E1 = ws.C[1][slot] + (int32_t)j * b1 + (int32_t)xo * s1 - (int32_t)((whb >> 17) & 1u);
    const int32_t E2 = ws.C[2][slot] + (int32_t)j * b2 + (int32_t)xo * s2 - (int32_t)((whb >> 18) & 1u);
    uint32_t ownRows = 0xFu;   // tiny entries: rows of the bbox this rank owns
    if (tiny && pc.nRanks > 1) {
        const uint32_t y0 = ws.xy[slot] >> 16;
        ownRows = 0u;
#pragma unroll
        for (uint32_t jj = 0; jj < 4u; ++jj) if (row_owned(y0 + jj, pc.rank, pc.nRanks, pc.bandH)) ownRows |= 1u << jj;
    }
    const uint32_t tag = slot | (xo << 6) | (j << 11);
#pragma unroll 1
    for (uint32_t k = 0; k < 4u; ++k) {
        const int32_t sx = c_msShift[k][0], sy = c_msShift[k][1];
        const int32_t F0 = E0 - ((s0 >> 8) * sx + (b0 >> 8) * sy), F1 = E1 - ((s1 >> 8) * sx + (b1 >> 8) * sy), F2 = E2 - ((s2 >> 8) * sx + (b2 >> 8) * sy);
        uint32_t mask = 0;
#pragma unroll 1
        for (uint32_t jj = 0; jj < hmax; ++jj) {   // rows past the first exist for tiny entries only (<= 4 columns, bit i + 4 jj)
            if (jj > 0 && !tiny) break;
#pragma unroll 1
            for (uint32_t i = 0; i < wmax; ++i) {
                const int32_t v = (F0 + (int32_t)jj * b0 + (int32_t)i * s0) | (F1 + (int32_t)jj * b1 + (int32_t)i * s1) | (F2 + (int32_t)jj * b2 + (int32_t)i * s2);
                if (i < w && jj < h) mask |= (v >= 0 ? 1u : 0u) << (tiny ? 4u * jj + i : i);
            }
        }
        if (tiny) {
#pragma unroll
            for (uint32_t jj = 0; jj < 4u; ++jj) if (!((ownRows >> jj) & 1u)) mask &= ~(0xFu << (4u * jj));
        }
        const uint32_t n = __popc(mask);
        const uint32_t incl = warp_incl_scan(n, lane);
        uint32_t pos = pixTail + incl - n;
        pixTail += __shfl_sync(0xFFFFFFFFu, incl, 31);
        while (mask) {
            const uint32_t bpos = (uint32_t)__ffs(mask) - 1u;
            mask &= mask - 1u;
            const uint32_t i = tiny ? (bpos & 3u) : bpos, jj = tiny ? (bpos >> 2) : 0u;
            pixq[pos++ & (MS_PIXQ - 1)] = (tag + (i << 6) + (jj << 11)) | (k << 16);
        }
        __syncwarp();
        while (pixTail - pixHead >= 32u) { k1ms_pixels(pc, ms, ws, pixq, pixHead, 32u, lane); pixHead += 32u; }
        __syncwarp();
    }
}

template <bool CULL>
__global__ void __launch_bounds__(K1_THREADS, 5) k_setup_raster_ms(const __grid_constant__ PassConst pc, const __grid_constant__ MsTargets ms) {
    __shared__ K1Shared sm;
    __shared__ uint32_t pixq32[K1_WARPS][MS_PIXQ];
    pdl_enter();
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    K1Warp& ws = sm.w[warp];
    uint32_t* pixq = pixq32[warp];
    const uint32_t bl = pc.batchLog2, tpb = 1u << bl;
    const uint32_t nBatches = CULL ? (__ldcg(pc.cullCounts) << (8u - bl)) : ((pc.nTris + tpb - 1u) >> bl);
    const uint32_t totalWarps = gridDim.x * K1_WARPS;
    const uint32_t laneTri = lane < tpb ? lane : 0xFFFFFFFFu;
    auto batch_tri = [&](uint32_t b) -> uint32_t {
        if (laneTri == 0xFFFFFFFFu) return 0xFFFFFFFFu;
        if (!CULL) return (b << bl) + laneTri;
        if (b >= nBatches) return 0xFFFFFFFFu;
        return (__ldg(pc.triBlockList + (b >> (8u - bl))) << 8) + ((b & ((1u << (8u - bl)) - 1u)) << bl) + laneTri;
    };
    uint32_t ringPos = 0, rowHead = 0, rowTail = 0, pixHead = 0, pixTail = 0;
    // software pipeline as in K1: the slot stream runs two batches ahead, the cache gathers one batch ahead of the raster stages
    uint32_t batch = blockIdx.x * K1_WARPS + warp;
    uint4 q[3];
    uint32_t id;
    {
        const uint32_t tri0 = batch_tri(batch);
        const uint4 s0 = k1_load_slots<CULL>(pc, tri0);
        k1_gather<CULL>(pc, tri0, s0, q);
        id = s0.w;
    }
    uint4 ns = k1_load_slots<CULL>(pc, batch_tri(batch + totalWarps));
    for (; batch < nBatches; batch += totalWarps) {
        const uint32_t tri = batch_tri(batch), ntri = batch_tri(batch + totalWarps);
        const uint4 nns = k1_load_slots<CULL>(pc, batch_tri(batch + 2u * totalWarps));
        K1Tri t;
        k1ms_classify(pc, tri < pc.nTris, q, t);
        // ---- beyond the small class (rare in the dense meshes this kernel is for): every sample's queue set gets its own record
        if (__any_sync(0xFFFFFFFFu, t.isLarge || t.needsClip)) {
#pragma unroll 1
            for (int k = 0; k < 4; ++k) {
                PassConst pk = pc;
                pk.status = ms.status[k]; pk.large = ms.large[k]; pk.huge = ms.huge[k]; pk.tileHead = ms.tileHead[k]; pk.clipQueue = ms.clipQueue[k];
                K1Tri m;
                k1ms_moved(pc, t, k, m);
                k1_emit_large(pk, m, id, lane);
                if (t.needsClip) {
                    const uint32_t qi = atomicAdd(&pk.status[6], 1u);
                    if (qi < pc.clipCap) pk.clipQueue[qi] = id; else atomicOr(&pk.status[2], OVF_LARGE);
                }
            }
        }
        // ---- small: one ring record, row entries over the union bbox
        const unsigned smallMask = __ballot_sync(0xFFFFFFFFu, t.isSmall);
        const uint32_t mySlot = (ringPos + __popc(smallMask & ((1u << lane) - 1u))) & (K1_RING - 1);
        const bool isTiny = t.isSmall && k1ms_is_tiny((uint32_t)(t.x1 - t.x0 + 1), (uint32_t)(t.y1 - t.y0 + 1));
        const uint32_t nEnt = !t.isSmall ? 0u : (isTiny ? 1u : (uint32_t)(t.y1 - t.y0 + 1) << ((t.x1 - t.x0 + 1) > (int32_t)K1_SPLIT ? 1 : 0));
        if (t.isSmall) k1_store_small(ws, t, id, mySlot);
        ringPos += __popc(smallMask);
        k1_gather<CULL>(pc, ntri, ns, q);                            // in flight during the raster stages
        id = ns.w; ns = nns;
        uint32_t done = 0;
        do {
            const uint32_t c = min(nEnt - done, K1_SPLIT);
            const uint32_t cIncl = warp_incl_scan(c, lane);
            const uint32_t rowPos = rowTail + cIncl - c;
            for (uint32_t r = 0; r < c; ++r) ws.rowq[(rowPos + r) & (K1_ROWQ - 1)] = (uint16_t)(mySlot | ((done + r) << 6));
            rowTail += __shfl_sync(0xFFFFFFFFu, cIncl, 31);
            done += c;
            __syncwarp();
            while (rowTail - rowHead >= 32u) { k1ms_rows<CULL>(pc, ms, ws, pixq, rowHead, 32u, pixHead, pixTail, lane); rowHead += 32u; }
        } while (__any_sync(0xFFFFFFFFu, done < nEnt));
        // queued samples reference ring slots of this batch or the one before: drain them before the ring advances
        if (pixTail != pixHead) { k1ms_pixels(pc, ms, ws, pixq, pixHead, pixTail - pixHead, lane); pixHead = pixTail; }
        __syncwarp();
    }
    if (rowTail != rowHead) {
        k1ms_rows<CULL>(pc, ms, ws, pixq, rowHead, rowTail - rowHead, pixHead, pixTail, lane);
        while (pixTail != pixHead) { const uint32_t c = min(32u, pixTail - pixHead); k1ms_pixels(pc, ms, ws, pixq, pixHead, c, lane); pixHead += c; }
    }
}

// Re-bases a draw list's index stream into cache-slot space, in submission order (runs when the list changes).
__global__ void __launch_bounds__(256) k_flatten_slots(const __grid_constant__ PassConst pc, uint32_t* __restrict__ flat, uint32_t* __restrict__ triMaterial) {
    const uint32_t tri = blockIdx.x * 256u + threadIdx.x;
    if (tri >= pc.nTris) return;
    const uint32_t d = find_draw<false>(pc.draws, pc.nDraws, tri);
    const uint32_t* ix = pc.indices + pc.draws[d].indexOffset + 3u * (tri - pc.draws[d].firstTri);
    const int32_t off = pc.draws[d].cacheOffset;
#pragma unroll
    for (int k = 0; k < 3; ++k) flat[3ull * tri + k] = (uint32_t)((int32_t)ix[k] + off);
    triMaterial[tri] = pc.draws[d].material;
}

__global__ void k_fill_u32(uint32_t* __restrict__ p, uint32_t v, size_t n) {
    pdl_enter();
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
void launch_setup_raster(const PassConst& pc, bool shadow, int numSMs, cudaStream_t st) {
    if (pc.nTris == 0) return;
    using Kern = void (*)(PassConst);
    static const Kern kern[6] = {k_setup_raster<false, false, false>, k_setup_raster<true, false, false>, k_setup_raster<false, true, false>,
                                 k_setup_raster<false, false, true>,  k_setup_raster<true, false, true>,  k_setup_raster<false, true, true>};
    // +1: shadow pass; 2: a sample pass of a multisampled frame (Appendix B.15); +3: per-draw culling for the sort-first split
    const int v = (shadow ? 1 : ((pc.shiftX | pc.shiftY) ? 2 : 0)) + (pc.drawVisible ? 3 : 0);
    static int perSM[6] = {0, 0, 0, 0, 0, 0};
    if (!perSM[v]) {
        int n = 0;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kern[v], K1_THREADS, 0);
        perSM[v] = n > 0 ? n : 1;
    }
    // short draw lists: fewer triangles per batch until at least half the resident warps have one (16 k triangles of 25 px in
    // 512 batches kept 512 warps busy for 43 us while 3 000 idled; C1's 64 k triangles in 2 000 full batches are best left alone)
    const uint32_t maxWarps = (uint32_t)(numSMs * perSM[v]) * K1_WARPS;
    PassConst pk = pc;
    pk.batchLog2 = 5u;
    static const bool fullBatches = std::getenv("REB_K1_FULL_BATCHES") != nullptr;   // A/B switch
    while (!fullBatches && pk.batchLog2 > 0u && ((pc.nTris + (1u << pk.batchLog2) - 1u) >> pk.batchLog2) < maxWarps / 2u) --pk.batchLog2;
    const uint32_t nBatches = (pc.nTris + (1u << pk.batchLog2) - 1u) >> pk.batchLog2;
    const uint32_t grid = std::min<uint32_t>((nBatches + K1_WARPS - 1u) / K1_WARPS, (uint32_t)(numSMs * perSM[v]));
    launch_k(kern[v], dim3(grid), dim3(K1_THREADS), st, true, pk);
}

// Between the sample passes of a multisampled frame: re-arm the pass status words and the per-tile counters (K0's
// prologue does this once per frame).  The overflow word and the largest list length stay, so the host sees an overflow of
// any sample pass and sizes the regrow for the longest one.
__global__ void __launch_bounds__(256) k_pass_reset(const __grid_constant__ PassConst pc) {
    pdl_enter();
    if (blockIdx.x == 0 && threadIdx.x < STATUS_WORDS && threadIdx.x != 2 && threadIdx.x != 3) pc.status[threadIdx.x] = 0u;
    uint4* th = reinterpret_cast<uint4*>(pc.tileHead);
    for (uint32_t i = blockIdx.x * 256u + threadIdx.x; i < pc.tilesX * pc.tilesY; i += gridDim.x * 256u) th[i] = make_uint4(0u, 0u, 0u, 0u);
}
void launch_pass_reset(const PassConst& pc, int numSMs, cudaStream_t st) {
    launch_k(k_pass_reset, dim3((unsigned)numSMs), dim3(256), st, true, pc);
}

// Before the one-pass multisampled K1: arm the queue sets of samples 1-3 (K0's prologue armed sample 0's, the pass's own)
__global__ void __launch_bounds__(256) k_ms_reset(const __grid_constant__ MsTargets ms, uint32_t nTiles) {
    pdl_enter();
    for (int k = 1; k < 4; ++k) {
        if (blockIdx.x == 0 && threadIdx.x < STATUS_WORDS) ms.status[k][threadIdx.x] = 0u;
        uint4* th = reinterpret_cast<uint4*>(ms.tileHead[k]);
        for (uint32_t i = blockIdx.x * 256u + threadIdx.x; i < nTiles; i += gridDim.x * 256u) th[i] = make_uint4(0u, 0u, 0u, 0u);
    }
}
void launch_setup_raster_ms(const PassConst& pc, const MsTargets& ms, int numSMs, cudaStream_t st) {
    if (pc.nTris == 0) return;
    launch_k(k_ms_reset, dim3((unsigned)numSMs), dim3(256), st, true, ms, pc.tilesX * pc.tilesY);
    using Kern = void (*)(PassConst, MsTargets);
    static const Kern kern[2] = {k_setup_raster_ms<false>, k_setup_raster_ms<true>};
    const int v = pc.drawVisible ? 1 : 0;
    static int perSM[2] = {0, 0};
    if (!perSM[v]) {
        int n = 0;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kern[v], K1_THREADS, 0);
        perSM[v] = n > 0 ? n : 1;
    }
    const uint32_t maxWarps = (uint32_t)(numSMs * perSM[v]) * K1_WARPS;
    PassConst pk = pc;
    pk.batchLog2 = 5u;
    while (pk.batchLog2 > 0u && ((pc.nTris + (1u << pk.batchLog2) - 1u) >> pk.batchLog2) < maxWarps / 2u) --pk.batchLog2;
    const uint32_t nBatches = (pc.nTris + (1u << pk.batchLog2) - 1u) >> pk.batchLog2;
    const uint32_t grid = std::min<uint32_t>((nBatches + K1_WARPS - 1u) / K1_WARPS, (uint32_t)(numSMs * perSM[v]));
    launch_k(kern[v], dim3(grid), dim3(K1_THREADS), st, true, pk, ms);
}

void launch_flatten_slots(const PassConst& pc, uint32_t* flat, uint32_t* triMaterial, cudaStream_t st) {
    if (pc.nTris == 0) return;
    k_flatten_slots<<<(pc.nTris + 255u) / 256u, 256, 0, st>>>(pc, flat, triMaterial);
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
