// reb_bin.cuh — the binned path of large triangles, shared by K1 (k_setup.cu: classification, queue append, per-tile counts) and
// K2 (k_tile.cu, k_tile_binned: Appendix B.3 clipping of the queued parents, scatter of the 48-B records into per-tile lists, tile
// raster).
//
// Per-tile lists are CONTIGUOUS arrays of LargeRec (not indices into the queue): the tile rasteriser streams a tile's list into
// shared memory with 1-D bulk asynchronous copies (cp.async.bulk + mbarrier) instead of one dependent gather per lane.
//
// Ordering without assuming that all CTAs of the grid are resident at once (a software grid barrier counting ARRIVALS over a grid
// sized by the occupancy API stalled for seconds on B200 in one frame out of three): every phase is a queue of work items CLAIMED
// from a global counter plus a counter of COMPLETED items; a warp leaves a phase when the queue is empty and then waits until
// completed == total — i.e. only for items that running warps have already claimed, never for a CTA that has not started:
//   clip    : queued parents (Appendix B.3); their children join the queue and the per-tile counts;
//   layout  : blocks of 32 tiles: list start = a running cursor bumped once per block (no ordered scan is needed: list order
//             inside a tile is irrelevant to the min-key resolve);
//   scatter : queued triangles: the 48-B record is copied into the list of every tile the triangle touches;
//   raster  : static striding over the tiles (k_tile.cu).
// Replaces the fixed-function clipper / binner behind vkCmdDrawIndexed (GraphicsPipelineCache.cpp:257-365; the reference
// delegates all of it to the Vulkan driver).
#pragma once
#include "reb_device.cuh"

namespace reb {

struct K1Tri {
    VtxA a[3];
    long long A2;
    int32_t x0, x1, y0, y1;
    bool isSmall, isLarge, needsClip;
};

__device__ __forceinline__ LargeRec k1_large_rec(const K1Tri& t, uint32_t tri) {
    LargeRec r;
    r.X0 = t.a[0].X; r.Y0 = t.a[0].Y; r.X1 = t.a[1].X; r.Y1 = t.a[1].Y;
    r.X2 = t.a[2].X; r.Y2 = t.a[2].Y; r.z0 = t.a[0].z; r.z1 = t.a[1].z;
    r.z2 = t.a[2].z; r.triId = tri; r.pad0 = 0; r.pad1 = 0;
    return r;
}

__device__ __forceinline__ uint32_t warp_incl_scan(uint32_t v, unsigned lane) {
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const uint32_t n = __shfl_up_sync(0xFFFFFFFFu, v, o); if (lane >= (unsigned)o) v += n; }
    return v;
}

__device__ __forceinline__ uint32_t ld_acquire_u32(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// ---------------------------------------------------------------------------------------------
// Tile walk of one large triangle by one warp: lanes stride over the TILE_W x TILE_H tiles of its
// bbox; a tile is skipped when it lies entirely outside one edge (edge function evaluated at the
// tile's most-interior pixel centre for that edge).  SCATTER = false counts (K1, the clipper),
// SCATTER = true copies the triangle's record into the per-tile lists.  All arguments are warp-uniform.
// ---------------------------------------------------------------------------------------------
template <bool SCATTER>
__device__ __noinline__ void tile_walk(const PassConst& pc, const LargeRec r, unsigned lane) {
    const int32_t X[3] = {r.X0, r.X1, r.X2}, Y[3] = {r.Y0, r.Y1, r.Y2};
    const int32_t minX = min(r.X0, min(r.X1, r.X2)), maxX = max(r.X0, max(r.X1, r.X2));
    const int32_t minY = min(r.Y0, min(r.Y1, r.Y2)), maxY = max(r.Y0, max(r.Y1, r.Y2));
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
        REB_ASSERT(tx >= 0 && ty >= 0 && (uint32_t)tx < pc.tilesX && (uint32_t)ty < pc.tilesY);
        TileHead* th = pc.tileHead + ((uint32_t)ty * pc.tilesX + (uint32_t)tx);
        if (!SCATTER) {
            atomicAdd(&th->count, 1u);
        } else {
            const uint32_t slot = __ldcg(&th->start) + atomicAdd(&th->pos, 1u);
            if (slot < pc.listCap) {   // the record itself, so a tile's list is one contiguous run of 48-B records
                uint4* d = reinterpret_cast<uint4*>(pc.tileRecs + slot);
                const uint4* s = reinterpret_cast<const uint4*>(&r);
                d[0] = s[0]; d[1] = s[1]; d[2] = s[2];
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------
// Appendix B.3 (revision r1c), the rare path: a triangle with a vertex at / behind the eye plane or outside the guard band.
// All 32 lanes run this with the SAME triangle (warp-uniform, no divergence): re-run the vertex stage on its three source
// vertices (the cache holds no clip coordinates), clip in clip space with Sutherland-Hodgman (near z >= 0, then the four
// guard planes; intersections computed from the inside vertex: t = d_in / (d_in - d_out)), fan the polygon from its
// first vertex, and send every child through the ordinary viewport / snap / facing / bbox steps into the large-triangle
// path with the parent's triId.
// ---------------------------------------------------------------------------------------------
template <bool SHADOW>
__device__ __noinline__ void clip_emit(const PassConst& pc, uint32_t tri, unsigned lane) {
    float4 a[9];
    const int n = clip_parent<SHADOW>(pc, tri, a);
    if (n < 3) return;
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
        if (slot < pc.largeCap) tile_walk<false>(pc, rec, lane);
    }
}

// Phase queues at CTA granularity (a same-address atomic costs ~5 ns chip-wide: one claim per warp and phase — 3,552 warps — was
// 50 us of a 60-us kernel): thread 0 claims `n` items for the CTA / publishes `n` completed items / waits for the phase to complete.
__device__ __forceinline__ uint32_t phase_claim(uint32_t* next, uint32_t n, uint32_t* sh) {
    __syncthreads();                       // the previous round's readers of *sh are done
    if (threadIdx.x == 0) *sh = atomicAdd(next, n);
    __syncthreads();
    return *sh;
}
__device__ __forceinline__ void phase_done(uint32_t* done, uint32_t n) {
    __syncthreads();                       // every warp's writes of this round precede thread 0's fence
    if (threadIdx.x == 0 && n) { __threadfence(); atomicAdd(done, n); }
}
__device__ __forceinline__ void phase_wait(const uint32_t* done, uint32_t total) {
    if (threadIdx.x == 0) { while (ld_acquire_u32(done) < total) __nanosleep(64); }
    __syncthreads();
}

// Clip + layout + scatter by every CTA of k_tile_binned, called (by all threads) when K1 queued something; see the ordering
// argument at the top.  status words: [8] / [9] clip items claimed / done, [11] / [12] tile blocks, [10] / [13] queued triangles.
template <bool SHADOW>
__device__ __noinline__ void bin_phase(const PassConst& pc, uint32_t* sh) {
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5, nw = blockDim.x >> 5;
    // 1. Appendix B.3: the parents K1 found to need true clipping, one per warp
    const uint32_t nClip = min(__ldcg(pc.status + 6), pc.clipCap);
    if (nClip) {
        for (;;) {
            const uint32_t base = phase_claim(pc.status + 8, nw, sh);
            if (base >= nClip) break;
            if (base + warp < nClip) clip_emit<SHADOW>(pc, __ldcg(pc.clipQueue + base + warp), lane);
            phase_done(pc.status + 9, min(nw, nClip - base));
        }
        phase_wait(pc.status + 9, nClip);
    }
    const uint32_t nLarge = min(ld_acquire_u32(pc.status + 0), pc.largeCap);   // final: queue and per-tile counts
    if (nLarge == 0) return;
    // 2. layout: 32 tiles per warp
    const uint32_t nTiles = pc.tilesX * pc.tilesY, nBlocks = (nTiles + 31u) >> 5;
    for (;;) {
        const uint32_t base = phase_claim(pc.status + 11, nw, sh);
        if (base >= nBlocks) break;
        const uint32_t i = (base + warp) * 32u + lane;
        const uint32_t c = i < nTiles ? __ldcg(&pc.tileHead[i].count) : 0u;
        const uint32_t inc = warp_incl_scan(c, lane);
        uint32_t wbase = 0;
        if (lane == 31 && inc) {
            wbase = atomicAdd(&pc.status[7], inc);
            atomicMax(&pc.status[3], wbase + inc);   // = the final cursor; max over the sample passes of a multisampled frame (k_pass_reset keeps it)
            if (wbase + inc > pc.listCap) atomicOr(&pc.status[2], OVF_LIST);
        }
        wbase = __shfl_sync(0xFFFFFFFFu, wbase, 31);
        if (c) pc.tileHead[i].start = wbase + inc - c;
        phase_done(pc.status + 12, min(nw, nBlocks - base));
    }
    phase_wait(pc.status + 12, nBlocks);
    // 3. scatter: a few queued triangles per warp and round (one when the queue is short, so that it spreads over as many warps as
    //    possible; up to 32 when it is long, to keep the claim counter cool)
    const uint32_t chunk = max(1u, min(32u, nLarge / (2u * nw * gridDim.x)));
    for (;;) {
        const uint32_t base = phase_claim(pc.status + 10, nw * chunk, sh);
        if (base >= nLarge) break;
        const uint32_t t0 = min(base + warp * chunk, nLarge), t1 = min(t0 + chunk, nLarge);
        for (uint32_t t = t0; t < t1; ++t) {
            const uint4* q = reinterpret_cast<const uint4*>(pc.large + t);
            union { LargeRec r; uint4 v[3]; } u;
            u.v[0] = __ldcg(q); u.v[1] = __ldcg(q + 1); u.v[2] = __ldcg(q + 2);
            tile_walk<true>(pc, u.r, lane);
        }
        phase_done(pc.status + 13, min(nw * chunk, nLarge - base));
    }
    phase_wait(pc.status + 13, nLarge);
}

}  // namespace reb
