// k_setup.cu — K1: triangle setup + small-triangle raster + large-triangle tile counting;
// K2: scatter of the large triangles into per-tile lists; clears; AoS -> SoA split.
//
// Replaces, for the hot path, the fixed-function front end that the reference reaches through
// vkCmdDrawIndexed (MaterialSystem.cpp:99, ShadowSystem.cpp:369) with the state of
// GraphicsPipelineCache.cpp:257-365: viewport, cull back + CCW front (F5), 1 sample per pixel,
// LessOrEqual depth.  The vertex stage itself is k_vertex.cu.
#include "reb_device.cuh"
#include "reb_kernels.h"
#include <algorithm>
#include <cstdlib>

namespace reb {

// ---------------------------------------------------------------------------------------------
// Tile walk of one large triangle by one warp: lanes stride over the TILE_W x TILE_H tiles of its
// bbox; a tile is skipped when it lies entirely outside one edge (edge function evaluated at the
// tile's most-interior pixel centre for that edge).  SCATTER = false counts (inside K1),
// SCATTER = true writes the triangle's queue slot into the per-tile lists (K2).
// All arguments are warp-uniform.
// ---------------------------------------------------------------------------------------------
template <bool SCATTER>
__device__ __noinline__ void tile_walk(const PassConst& pc, int32_t X0, int32_t Y0, int32_t X1, int32_t Y1, int32_t X2, int32_t Y2,
                                          uint32_t queueSlot, unsigned lane) {
    const int32_t X[3] = {X0, X1, X2}, Y[3] = {Y0, Y1, Y2};
    const int32_t minX = min(X0, min(X1, X2)), maxX = max(X0, max(X1, X2));
    const int32_t minY = min(Y0, min(Y1, Y2)), maxY = max(Y0, max(Y1, Y2));
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
    for (int i = (int)lane; i < n; i += 32) {
        const int tx = tx0 + i % tw, ty = ty0 + i / tw;
        if (pc.nRanks > 1 && !row_owned((uint32_t)(ty * TILE_H), pc.rank, pc.nRanks, pc.bandH)) continue;
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
            atomicAdd(&pc.tileCount[tile], 1u);
        } else {
            const uint32_t slot = pc.tileStart[tile] + atomicAdd(&pc.tileCount[tile], 1u);
            if (slot < pc.listCap) pc.tileList[slot] = queueSlot;
        }
    }
}

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
// Tail (once per CTA): the last CTA to finish lays out the per-tile lists (start = running cursor; no
// ordered scan is needed because list order inside a tile is irrelevant to the min-key resolve).
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

struct K1Tri {
    VtxA a[3];
    long long A2;
    int32_t x0, x1, y0, y1;
    bool isSmall, isLarge, needsClip;
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

__device__ __forceinline__ LargeRec k1_large_rec(const K1Tri& t, uint32_t tri) {
    LargeRec r;
    r.X0 = t.a[0].X; r.Y0 = t.a[0].Y; r.X1 = t.a[1].X; r.Y1 = t.a[1].Y;
    r.X2 = t.a[2].X; r.Y2 = t.a[2].Y; r.z0 = t.a[0].z; r.z1 = t.a[1].z;
    r.z2 = t.a[2].z; r.triId = tri; r.pad0 = 0; r.pad1 = 0;
    return r;
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
    while (lmask) {   // warp-cooperative tile counting, one large triangle at a time
        const int src = __ffs(lmask) - 1;
        lmask &= lmask - 1;
        const int32_t X0 = __shfl_sync(0xFFFFFFFFu, t.a[0].X, src), Y0 = __shfl_sync(0xFFFFFFFFu, t.a[0].Y, src);
        const int32_t X1 = __shfl_sync(0xFFFFFFFFu, t.a[1].X, src), Y1 = __shfl_sync(0xFFFFFFFFu, t.a[1].Y, src);
        const int32_t X2 = __shfl_sync(0xFFFFFFFFu, t.a[2].X, src), Y2 = __shfl_sync(0xFFFFFFFFu, t.a[2].Y, src);
        const uint32_t s = __shfl_sync(0xFFFFFFFFu, slot, src);
        if (s < pc.largeCap) tile_walk<false>(pc, X0, Y0, X1, Y1, X2, Y2, s, lane);
    }
}

// ---------------------------------------------------------------------------------------------
// Appendix B.3 (revision r1c), the rare path: a triangle with a vertex at / behind the eye plane or outside the guard band.
// All 32 lanes run this with the SAME triangle (warp-uniform, no divergence): re-run the vertex stage on its three source
// vertices (the cache holds no clip coordinates), clip in clip space with Sutherland-Hodgman (near z >= 0, then the four
// guard planes; intersections computed from the inside vertex: t = d_in / (d_in - d_out)), fan the polygon from its
// first vertex, and send every child through the ordinary viewport / snap / facing / bbox steps into the large-triangle
// path with the parent's triId.  Returns true when a child was queued into the binned path.
// ---------------------------------------------------------------------------------------------
template <bool SHADOW>
__device__ __noinline__ bool k1_clip_emit(const PassConst& pc, uint32_t tri, unsigned lane) {
    float4 a[9];
    const int n = clip_parent<SHADOW>(pc, tri, a);
    if (n < 3) return false;
    bool queued = false;
    VtxA v0 = viewport_snap(a[0], pc.W, pc.H);
    v0.X += pc.shiftX; v0.Y += pc.shiftY;   // B.15 sample shift
    for (int i = 1; i + 1 < n; ++i) {
        K1Tri t;
        t.a[0] = v0; t.a[1] = viewport_snap(a[i], pc.W, pc.H); t.a[2] = viewport_snap(a[i + 1], pc.W, pc.H);
        t.a[1].X += pc.shiftX; t.a[1].Y += pc.shiftY; t.a[2].X += pc.shiftX; t.a[2].Y += pc.shiftY;
        if (((t.a[0].flags | t.a[1].flags | t.a[2].flags) & VF_NEEDS_CLIP) != 0u) {   // cannot happen for finite inputs: counted, dropped
            if (lane == 0) atomicAdd(&pc.status[1], 1u);
            continue;
        }
        t.A2 = (long long)(t.a[1].X - t.a[0].X) * (long long)(t.a[2].Y - t.a[0].Y) - (long long)(t.a[2].X - t.a[0].X) * (long long)(t.a[1].Y - t.a[0].Y);
        if (t.A2 >= 0) continue;                                                        // Appendix B.6 applies to the child
        const int32_t minX = min(t.a[0].X, min(t.a[1].X, t.a[2].X)), maxX = max(t.a[0].X, max(t.a[1].X, t.a[2].X));
        const int32_t minY = min(t.a[0].Y, min(t.a[1].Y, t.a[2].Y)), maxY = max(t.a[0].Y, max(t.a[1].Y, t.a[2].Y));
        t.x0 = max((minX + 127) >> 8, 0); t.x1 = min((maxX - 128) >> 8, (int32_t)pc.W - 1);
        t.y0 = max((minY + 127) >> 8, 0); t.y1 = min((maxY - 128) >> 8, (int32_t)pc.H - 1);
        if (t.x0 > t.x1 || t.y0 > t.y1) continue;
        const LargeRec rec = k1_large_rec(t, tri);
        const int ntiles = (t.x1 / TILE_W - t.x0 / TILE_W + 1) * (t.y1 / TILE_H - t.y0 / TILE_H + 1);
        uint32_t slot = 0xFFFFFFFFu;   // 0xFFFFFFFF = went to the broadcast list
        if (lane == 0) {
            bool binned = true;
            if (ntiles > HUGE_TILES) {
                const uint32_t h = atomicAdd(&pc.status[4], 1u);
                if (h < (uint32_t)HUGE_CAP) { pc.huge[h] = rec; binned = false; }
            }
            if (binned) {
                slot = atomicAdd(&pc.status[0], 1u);
                if (slot < pc.largeCap) pc.large[slot] = rec;
                else { atomicOr(&pc.status[2], OVF_LARGE); slot = 0xFFFFFFFEu; }
            }
        }
        slot = __shfl_sync(0xFFFFFFFFu, slot, 0);
        if (slot < pc.largeCap) {
            queued = true;
            tile_walk<false>(pc, rec.X0, rec.Y0, rec.X1, rec.Y1, rec.X2, rec.Y2, slot, lane);
        }
    }
    return queued;
}

__device__ __forceinline__ uint32_t warp_incl_scan(uint32_t v, unsigned lane) {
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const uint32_t n = __shfl_up_sync(0xFFFFFFFFu, v, o); if (lane >= (unsigned)o) v += n; }
    return v;
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
// and are gathered as an all-zero (degenerate, hence culled) triangle; K0 skipped the vertices of wholly invisible draws.
constexpr uint32_t SLOT_CULLED = 0xFFFFFFFFu;
template <bool CULL>
__device__ __forceinline__ uint3 k1_load_slots(const PassConst& pc, uint32_t tri) {
    if (tri >= pc.nTris) return make_uint3(0u, 0u, 0u);
    if (CULL && __ldg(pc.blockVisible + (tri >> 8)) == 0u) return make_uint3(SLOT_CULLED, 0u, 0u);
    if (pc.deindexed) return make_uint3(3u * tri, 3u * tri + 1u, 3u * tri + 2u);
    const uint32_t* f = pc.flatSlots + 3ull * tri;
    return make_uint3(__ldg(f), __ldg(f + 1), __ldg(f + 2));
}
template <bool CULL>
__device__ __forceinline__ void k1_gather(const PassConst& pc, uint32_t tri, uint3 s, uint4 q[3]) {
    if (tri >= pc.nTris || (CULL && s.x == SLOT_CULLED)) { q[0] = q[1] = q[2] = make_uint4(0, 0, 0, 0); return; }
    const uint4* va = reinterpret_cast<const uint4*>(pc.vtxA);
    q[0] = __ldg(va + s.x); q[1] = __ldg(va + s.y); q[2] = __ldg(va + s.z);
}

// Row stage: `cnt` (<= 32) pending rows, one per lane -> coverage masks -> pixel queue entries.
template <bool SHADOW, bool CULL = false>
__device__ __forceinline__ void k1_rows(const PassConst& pc, K1Warp& ws, uint32_t rowHead, uint32_t cnt, uint32_t& pixTail, unsigned lane) {
    bool active = lane < cnt;
    uint32_t slot = 0, j = 0, whb = 0, xo = 0;
    if (active) {
        const uint32_t e = ws.rowq[(rowHead + lane) & (K1_ROWQ - 1)];
        slot = e & 0x3Fu;
        whb = ws.whb[slot];
        const uint32_t wide = (whb & 0xFFu) > K1_SPLIT ? 1u : 0u;   // two half rows per row
        j = (e >> 6) >> wide; xo = ((e >> 6) & wide) * K1_SPLIT;
        if ((!SHADOW || CULL) && !row_owned((ws.xy[slot] >> 16) + j, pc.rank, pc.nRanks, pc.bandH)) active = false;   // shadow slabs only in the split build
    }
    const uint32_t w = active ? min((whb & 0xFFu) - xo, K1_SPLIT) : 0u;
    const uint32_t wmax = __reduce_max_sync(0xFFFFFFFFu, w);
    uint32_t mask = 0;
    {
        const int32_t s0 = ws.Ax[0][slot], s1 = ws.Ax[1][slot], s2 = ws.Ax[2][slot];
        int32_t F0 = ws.C[0][slot] + (int32_t)j * ws.By[0][slot] + (int32_t)xo * s0 - (int32_t)((whb >> 16) & 1u);   // top-left bias folded in
        int32_t F1 = ws.C[1][slot] + (int32_t)j * ws.By[1][slot] + (int32_t)xo * s1 - (int32_t)((whb >> 17) & 1u);
        int32_t F2 = ws.C[2][slot] + (int32_t)j * ws.By[2][slot] + (int32_t)xo * s2 - (int32_t)((whb >> 18) & 1u);
#pragma unroll 1
        for (uint32_t i = 0; i < wmax; ++i, F0 += s0, F1 += s1, F2 += s2) mask |= ((F0 | F1 | F2) >= 0 ? 1u : 0u) << i;
    }
    mask &= (1u << w) - 1u;
    const uint32_t n = __popc(mask);
    const uint32_t incl = warp_incl_scan(n, lane);
    uint32_t pos = pixTail + incl - n;
    pixTail += __shfl_sync(0xFFFFFFFFu, incl, 31);
    const uint32_t tag = slot | (xo << 6) | (j << 11);   // i + xo < 32: the add cannot carry into j
    while (mask) {
        const uint32_t i = (uint32_t)__ffs(mask) - 1u;
        mask &= mask - 1u;
        ws.pixq[pos++ & (K1_PIXQ - 1)] = (uint16_t)(tag + (i << 6));
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
    const uint32_t nBatches = (pc.nTris + tpb - 1u) >> bl;
    const uint32_t totalWarps = gridDim.x * K1_WARPS;
    const uint32_t laneTri = lane < tpb ? lane : 0xFFFFFFFFu;   // lanes past the batch see triangle index >= nTris (nothing to do)
#define K1_TRI(b) (laneTri == 0xFFFFFFFFu ? 0xFFFFFFFFu : ((b) << bl) + laneTri)

    // static schedule: warp g takes batches g, g + totalWarps, ... (a shared work counter would serialise on one
    // L2 atomic unit: ~10 cycles per same-address atomic-with-return chip-wide, measured as the top stall)
    uint32_t batch = blockIdx.x * K1_WARPS + warp;
    uint4 q[3];
    bool curCulled = false;   // CULL: every triangle of the current batch belongs to a culled draw (warp-uniform)
    {
        const uint32_t tri = K1_TRI(batch);
        const uint3 s0 = k1_load_slots<CULL>(pc, tri);
        if (CULL) curCulled = __all_sync(0xFFFFFFFFu, s0.x == SLOT_CULLED || tri >= pc.nTris);
        k1_gather<CULL>(pc, tri, s0, q);
    }
    uint3 ns = k1_load_slots<CULL>(pc, K1_TRI(batch + totalWarps));   // slot stream runs two batches ahead of the raster stages
    uint32_t ringPos = 0, rowHead = 0, rowTail = 0, pixHead = 0, pixTail = 0;   // warp-uniform
    while (batch < nBatches) {
        const uint32_t nbatch = batch + totalWarps;
        const uint32_t tri = K1_TRI(batch), ntri = K1_TRI(nbatch);
        const uint3 nns = k1_load_slots<CULL>(pc, K1_TRI(nbatch + totalWarps));   // in flight for a whole iteration
        if (CULL && curCulled) {   // nothing of this batch is ours: only advance the software pipeline
            curCulled = __all_sync(0xFFFFFFFFu, ns.x == SLOT_CULLED || ntri >= pc.nTris);
            k1_gather<CULL>(pc, ntri, ns, q);
            ns = nns; batch = nbatch;
            continue;
        }

        // ---- 1. classify, emit large, append small records + rows ----------------------------------------
        K1Tri t;
        k1_classify<SHIFT>(pc, tri < pc.nTris, q, t);
        k1_emit_large(pc, t, tri, lane);
        if (t.needsClip) {   // Appendix B.3, rare: handed to k_clip_layout (keeps the clipper out of this kernel's register budget)
            const uint32_t qi = atomicAdd(&pc.status[6], 1u);
            if (qi < pc.clipCap) pc.clipQueue[qi] = tri; else atomicOr(&pc.status[2], OVF_LARGE);
        }
        const unsigned smallMask = __ballot_sync(0xFFFFFFFFu, t.isSmall);
        const uint32_t mySlot = (ringPos + __popc(smallMask & ((1u << lane) - 1u))) & (K1_RING - 1);
        // row entries of this lane's triangle: one per bbox row, two (half rows) when the bbox is wider than K1_SPLIT
        const uint32_t nEnt = t.isSmall ? (uint32_t)(t.y1 - t.y0 + 1) << ((t.x1 - t.x0 + 1) > (int32_t)K1_SPLIT ? 1 : 0) : 0u;
        if (t.isSmall) k1_store_small(ws, t, tri, mySlot);
        ringPos += __popc(smallMask);
        if (CULL) curCulled = __all_sync(0xFFFFFFFFu, ns.x == SLOT_CULLED || ntri >= pc.nTris);
        k1_gather<CULL>(pc, ntri, ns, q);                            // in flight during the raster stages
        ns = nns;

        // ---- 2./3. rows -> samples -> resolve; <= K1_SPLIT row entries per lane and round keep the rings within capacity ----
        uint32_t done = 0;
        do {
            const uint32_t c = min(nEnt - done, K1_SPLIT);
            const uint32_t cIncl = warp_incl_scan(c, lane);
            const uint32_t rowPos = rowTail + cIncl - c;
            for (uint32_t j = 0; j < c; ++j) ws.rowq[(rowPos + j) & (K1_ROWQ - 1)] = (uint16_t)(mySlot | ((done + j) << 6));
            rowTail += __shfl_sync(0xFFFFFFFFu, cIncl, 31);
            done += c;
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

    // the tile lists are laid out by k_clip_layout, after the clipped triangles (if any) have been counted too
#undef K1_TRI
}

// ---------------------------------------------------------------------------------------------
// K1c: clip + layout.  Warps stride over the queue of triangles K1 found to need true clipping (Appendix B.3; empty for
// ordinary frames) and send their children into the large-triangle path; then the per-tile lists are laid out: start =
// a running cursor (status[7]) bumped once per warp of 32 tiles — no ordered scan is needed because list order inside a
// tile is irrelevant to the min-key resolve — and the counts become K2's scatter cursors.  Without clipping no CTA
// depends on another and the whole grid shares the tiles (one CTA walking the 16 k tiles of a 2048^2 map took 64 us, the
// 65 k tiles of a 4096^2 map 250 us); with clipping the last CTA to finish its share of the queue does it alone.
// ---------------------------------------------------------------------------------------------
constexpr int KC_THREADS = 128;
template <bool SHADOW>
__global__ void __launch_bounds__(KC_THREADS) k_clip_layout(const __grid_constant__ PassConst pc) {
    __shared__ uint32_t isLast;
    pdl_enter();
    const unsigned lane = threadIdx.x & 31u;
    const uint32_t nClip = min(pc.status[6], pc.clipCap);
    uint32_t first = blockIdx.x * KC_THREADS, stride = gridDim.x * KC_THREADS;
    if (nClip == 0) {
        if (pc.status[0] == 0u) return;   // nothing binned (the usual frame of small triangles)
    } else {
        const uint32_t warpsPerGrid = (gridDim.x * KC_THREADS) >> 5;
        for (uint32_t i = (blockIdx.x * KC_THREADS + threadIdx.x) >> 5; i < nClip; i += warpsPerGrid) k1_clip_emit<SHADOW>(pc, __ldcg(pc.clipQueue + i), lane);
        __threadfence();
        __syncthreads();
        if (threadIdx.x == 0) isLast = (atomicAdd(&pc.status[5], 1u) == gridDim.x - 1u) ? 1u : 0u;
        __syncthreads();
        if (!isLast) return;
        __threadfence();
        first = 0u; stride = KC_THREADS;
    }
    const uint32_t nLarge = min(*reinterpret_cast<volatile uint32_t*>(&pc.status[0]), pc.largeCap);
    if (nLarge == 0) return;
    const uint32_t nTiles = pc.tilesX * pc.tilesY;
    for (uint32_t i0 = first; i0 < nTiles; i0 += stride) {
        const uint32_t i = i0 + threadIdx.x;
        const uint32_t c = i < nTiles ? __ldcg(pc.tileCount + i) : 0u;
        const uint32_t inc = warp_incl_scan(c, lane);
        uint32_t wbase = 0;
        if (lane == 31 && inc) {
            wbase = atomicAdd(&pc.status[7], inc);
            atomicMax(&pc.status[3], wbase + inc);   // = the final cursor once every warp is through; max over the sample passes of a multisampled frame (k_pass_reset keeps it)
            if (wbase + inc > pc.listCap) atomicOr(&pc.status[2], OVF_LIST);
        }
        wbase = __shfl_sync(0xFFFFFFFFu, wbase, 31);
        if (i < nTiles && c) { pc.tileStart[i] = wbase + inc - c; pc.tileCount[i] = 0u; }   // counts become the scatter cursors
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

// ---------------------------------------------------------------------------------------------
// K2: scatter the queued large triangles into the per-tile lists laid out by K1's tail.
// One warp per triangle (same tile walk as the count).  Does nothing when the queue is empty.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_bin_scatter(const __grid_constant__ PassConst pc) {
    pdl_enter();
    const uint32_t nLarge = min(pc.status[0], pc.largeCap);
    if (nLarge == 0) return;
    const unsigned lane = threadIdx.x & 31u;
    const uint32_t warpsPerGrid = (gridDim.x * blockDim.x) >> 5;
    for (uint32_t t = (blockIdx.x * blockDim.x + threadIdx.x) >> 5; t < nLarge; t += warpsPerGrid) {
        const LargeRec r = pc.large[t];
        tile_walk<true>(pc, r.X0, r.Y0, r.X1, r.Y1, r.X2, r.Y2, t, lane);
    }
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

void launch_clip_layout(const PassConst& pc, bool shadow, int numSMs, cudaStream_t st) {
    if (shadow) launch_k(k_clip_layout<true>, dim3(numSMs * 2), dim3(KC_THREADS), st, true, pc);
    else launch_k(k_clip_layout<false>, dim3(numSMs * 2), dim3(KC_THREADS), st, true, pc);
}

// Between the sample passes of a multisampled frame: re-arm the pass status words and the per-tile counters (K0's
// prologue does this once per frame).  The overflow word and the largest list length stay, so the host sees an overflow of
// any sample pass and sizes the regrow for the longest one.
__global__ void __launch_bounds__(256) k_pass_reset(const __grid_constant__ PassConst pc) {
    pdl_enter();
    if (blockIdx.x == 0 && threadIdx.x < 8 && threadIdx.x != 2 && threadIdx.x != 3) pc.status[threadIdx.x] = 0u;
    for (uint32_t i = blockIdx.x * 256u + threadIdx.x; i < pc.tilesX * pc.tilesY; i += gridDim.x * 256u) pc.tileCount[i] = 0u;
}
void launch_pass_reset(const PassConst& pc, int numSMs, cudaStream_t st) {
    launch_k(k_pass_reset, dim3((unsigned)numSMs), dim3(256), st, true, pc);
}

void launch_flatten_slots(const PassConst& pc, uint32_t* flat, uint32_t* triMaterial, cudaStream_t st) {
    if (pc.nTris == 0) return;
    k_flatten_slots<<<(pc.nTris + 255u) / 256u, 256, 0, st>>>(pc, flat, triMaterial);
}

void launch_bin_scatter(const PassConst& pc, int numSMs, cudaStream_t st) {
    launch_k(k_bin_scatter, dim3(numSMs * 4), dim3(256), st, true, pc);
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
