// k_vertex.cu — K0: the vertex stage, run once per referenced vertex into a post-transform cache.
//
// Replaces the vertex shaders the reference runs through vkCmdDrawIndexed: shadow_map.vert:48-52
// (SHADOW) and primitive.vert:49-57 (forward), followed by the fixed-function viewport transform of
// GraphicsPipelineCache.cpp:280-285.  The hardware's post-transform vertex cache becomes an explicit
// SoA stream in HBM/L2: VtxA (snapped position + depth + clip flags, all that triangle setup needs)
// and, for the forward pass, VtxBC (world position, 1/w, normalised normal — what shading interpolates).
//
// One thread per cache slot.  Indexed mode: every draw owns the contiguous slot range that covers the
// vertex range its indices reference (host-side prefix over per-draw ranges), so slot -> (draw, vertex)
// is a CTA-level search plus an add, loads are 16-B coalesced (float4 SoA) and stores are 16/32-B
// coalesced.  De-indexed mode (draws referencing sparse / shared vertex ranges): slot = 3*tri + corner.
#include "reb_device.cuh"
#include "reb_kernels.h"
#include <algorithm>

namespace reb {

#ifndef K0_PER_THREAD_D
#define K0_PER_THREAD_D 1
#endif
// cache slots per thread.  Four (independent load chains in flight) was the round-1 optimum; since the source streams are read in slot
// order (k_reorder.cu) more, lighter threads win: C2 136.2 / 134.5 / 136.2 / 132.1 us per frame at 4 / 2 / 3 / 1, C1 48.2 -> 43.0 us
constexpr int K0_PER_THREAD = K0_PER_THREAD_D;
constexpr int K0_PER_THREAD_CULL = 4;            // sort-first split: most CTAs find nothing to do, fewer and fatter ones cost less (slowest rank of C4 at N = 8: 307 vs 316 us)

constexpr uint32_t K0_MIN_GRID = 296u;           // two CTAs per SM
struct K0In { float4 p, n; uint32_t node; bool valid; };

template <bool SHADOW, bool CULL = false>
__device__ __forceinline__ K0In k0_load(const PassConst& pc, uint32_t slot) {
    K0In in;
    in.valid = slot < pc.cacheSize;
    in.node = 0; in.p = in.n = make_float4(0.f, 0.f, 0.f, 0.f);
    if (!in.valid) return in;
    // sort-first split: no visible 256-triangle block of this rank references this 256-slot block
    if (CULL && __ldg(pc.slotBlockVisible + (slot >> 8)) != pc.cullStamp) { in.valid = false; return in; }
    // cacheBase is 3 * firstTri in de-indexed mode, so one table serves both layouts
    const uint32_t d = find_draw_block<true>(pc.draws, pc.nDraws, pc.slotBlockDraw, slot);
    uint32_t v = slot;
    if (pc.deindexed) {
        const uint32_t tri = slot / 3u, k = slot - 3u * tri;
        v = __ldg(pc.indices + __ldg(&pc.draws[d].indexOffset) + 3u * (tri - __ldg(&pc.draws[d].firstTri)) + k);
    } else if (!pc.posS) {
        v = (uint32_t)((int32_t)slot - __ldg(&pc.draws[d].cacheOffset));
    }
    // no row of this draw belongs to this rank: nothing is fetched, transformed or stored (K1 never gathers these slots: it skips
    // the draw's 256-triangle blocks, and resolves the draw per triangle in a block that spans two draws)
    if (CULL && __ldg(pc.drawVisible + d) == 0u) { in.valid = false; return in; }
    in.node = __ldg(&pc.draws[d].node);
    // re-ordered draw list: the source streams were copied into slot order (k_reorder.cu), the read is as linear as the write
    in.p = __ldg((pc.posS ? pc.posS : pc.pos4) + v);
    if (!SHADOW) in.n = __ldg((pc.posS ? pc.nrmS : pc.nrm4) + v);
    return in;
}

template <bool SHADOW>
__device__ __forceinline__ void k0_transform_store(const PassConst& pc, uint32_t slot, const K0In& in) {
    if (!in.valid) return;
    const float4* Tq = reinterpret_cast<const float4*>(pc.transforms) + 4u * in.node;
    float T[16];
#pragma unroll
    for (int k = 0; k < 4; ++k) { const float4 q = __ldg(Tq + k); T[4 * k] = q.x; T[4 * k + 1] = q.y; T[4 * k + 2] = q.z; T[4 * k + 3] = q.w; }
    V3 world;
    const float4 c = vertex_stage<!SHADOW>(pc.PV, T, in.p.x, in.p.y, in.p.z, &world);
    const VtxA a = viewport_snap(c, pc.W, pc.H);
    REB_ASSERT(slot < pc.cacheSize && in.node * 16u + 15u < 0xFFFFFFF0u);
    reinterpret_cast<uint4*>(pc.vtxA)[slot] = make_uint4((uint32_t)a.X, (uint32_t)a.Y, __float_as_uint(a.z), a.flags);
    if (!SHADOW) {
        const float4 n = in.n;
        const float4* Nq = reinterpret_cast<const float4*>(pc.nmats) + 3u * in.node;
        const float4 n0 = __ldg(Nq), n1 = __ldg(Nq + 1), n2 = __ldg(Nq + 2);
        V3 tn;   // glm mat3 * vec3: (m0*x + m1*y) + m2*z   (primitive.vert:55)
        tn.x = (n0.x * n.x + n0.w * n.y) + n1.z * n.z;
        tn.y = (n0.y * n.x + n1.x * n.y) + n1.w * n.z;
        tn.z = (n0.z * n.x + n1.y * n.y) + n2.x * n.z;
        tn = normalize3(tn);
        float4* o = reinterpret_cast<float4*>(pc.vtxBC + slot);
        o[0] = make_float4(world.x, world.y, world.z, 1.0f / c.w);   // Appendix B.10: q_k = b_k * (1 / w_k)
        o[1] = make_float4(tn.x, tn.y, tn.z, 0.0f);
        if (pc.vtxUV) pc.vtxUV[slot] = make_float2(in.p.w, in.n.w);   // (u, v) ride in the w lanes of the two SoA streams
    }
}

// Pass prologue (K0 is the first kernel of a pass and is stream-ordered after the previous use of these buffers): reset
// the pass status words and the per-tile list counters, so the pass needs no memset node and the tile stage no re-arm.
__device__ __forceinline__ void k0_prologue(const PassConst& pc) {
    if (blockIdx.x == 0 && threadIdx.x < STATUS_WORDS) pc.status[threadIdx.x] = 0u;
    uint4* th = reinterpret_cast<uint4*>(pc.tileHead);
    for (uint32_t i = blockIdx.x * 256u + threadIdx.x; i < pc.tilesX * pc.tilesY; i += gridDim.x * 256u) th[i] = make_uint4(0u, 0u, 0u, 0u);
}
// Fused clear of the shadow map to depth 1.0 (ShadowSystem.cpp:111-127): every CTA streams one contiguous slice.
__device__ __forceinline__ void k0_clear_shadow(const PassConst& pc) {
    if (pc.nRanks > 1 && (pc.W & 3u) == 0u) {
        // row slabs (reb_set_shadow_partition): only this rank's rows are rasterised, and the other rows are overwritten by the
        // all-gather that completes the map — clearing them on every rank was 7/8 of a 16.8-MB store stream at N = 8
        const uint32_t perBand = pc.bandH * (pc.W >> 2), period = pc.bandH * pc.nRanks;
        const uint32_t nBands = (pc.H + period - 1u) / period;                  // bands of this rank (the last may be cut by H)
        const size_t n4 = (size_t)nBands * perBand;
        const size_t per = (n4 + gridDim.x - 1) / gridDim.x;
        const size_t lo = per * blockIdx.x, hi = lo + per < n4 ? lo + per : n4;
        uint4* dst = reinterpret_cast<uint4*>(pc.depthBits);
        const uint4 one = make_uint4(0x3F800000u, 0x3F800000u, 0x3F800000u, 0x3F800000u);
        for (size_t i = lo + threadIdx.x; i < hi; i += 256u) {
            const uint32_t vb = (uint32_t)(i / perBand), r = (uint32_t)(i - (size_t)vb * perBand);
            const uint32_t y = (vb * pc.nRanks + pc.rank) * pc.bandH + r / (pc.W >> 2);
            if (y < pc.H) dst[(size_t)y * (pc.W >> 2) + r % (pc.W >> 2)] = one;
        }
        return;
    }
    const size_t nPix = (size_t)pc.W * pc.H, n4 = nPix / 4u;
    if (blockIdx.x == 0 && threadIdx.x < (nPix & 3u)) pc.depthBits[4u * n4 + threadIdx.x] = 0x3F800000u;
    const size_t per = (n4 + gridDim.x - 1) / gridDim.x;
    const size_t lo = per * blockIdx.x, hi = lo + per < n4 ? lo + per : n4;
    uint4* dst = reinterpret_cast<uint4*>(pc.depthBits);
    const uint4 one = make_uint4(0x3F800000u, 0x3F800000u, 0x3F800000u, 0x3F800000u);
    for (size_t i = lo + threadIdx.x; i < hi; i += 256u) dst[i] = one;
}

// Both vertex stages of a frame in one launch: used when the shadow and forward draw lists are identical (the
// reference's shadow list is the opaque list followed by the translucent one, RenderScene.cpp:191-208), so both
// post-transform caches have the same layout and every source vertex is fetched once.
__global__ void __launch_bounds__(256) k_vertex_both(const __grid_constant__ PassConst ps, const __grid_constant__ PassConst pf) {
    pdl_enter();
    const uint32_t base = blockIdx.x * (256u * K0_PER_THREAD) + threadIdx.x;
    k0_prologue(ps);
    k0_prologue(pf);
    k0_clear_shadow(ps);
    K0In in[K0_PER_THREAD];
#pragma unroll
    for (int k = 0; k < K0_PER_THREAD; ++k) in[k] = k0_load<false>(pf, base + 256u * k);
#pragma unroll
    for (int k = 0; k < K0_PER_THREAD; ++k) {
        k0_transform_store<true>(ps, base + 256u * k, in[k]);
        k0_transform_store<false>(pf, base + 256u * k, in[k]);
    }
}

template <bool SHADOW, bool CULL>
__global__ void __launch_bounds__(256) k_vertex(const __grid_constant__ PassConst pc) {
    constexpr int PER = CULL ? K0_PER_THREAD_CULL : K0_PER_THREAD;
    const uint32_t base = blockIdx.x * (256u * PER) + threadIdx.x;
    if (CULL) {
        // launched with PDL behind k_cull (an ordinary launch, so everything older has completed): the prologue and the map clear
        // touch nothing k_cull does and run while it does; only the vertex loop needs its marks
        asm volatile("griddepcontrol.launch_dependents;");
        k0_prologue(pc);
        if (SHADOW) k0_clear_shadow(pc);
        asm volatile("griddepcontrol.wait;" ::: "memory");
    } else {
        pdl_enter();
        k0_prologue(pc);
        if (SHADOW) k0_clear_shadow(pc);
    }
    K0In in[PER];
#pragma unroll
    for (int k = 0; k < PER; ++k) in[k] = k0_load<SHADOW, CULL>(pc, base + 256u * k);
#pragma unroll
    for (int k = 0; k < PER; ++k) k0_transform_store<SHADOW>(pc, base + 256u * k, in[k]);
}

// Per-draw referenced vertex range (min, max index), computed when a draw list changes (not per frame).
// ranges[2*d] = min, ranges[2*d+1] = max; initialised to (0xFFFFFFFF, 0) by the caller.
__global__ void __launch_bounds__(256) k_draw_ranges(const uint32_t* __restrict__ indices, const DrawDev* __restrict__ draws, uint32_t nDraws,
                                                     uint32_t* __restrict__ ranges) {
    const uint32_t d = blockIdx.y;
    if (d >= nDraws) return;
    const uint32_t off = draws[d].indexOffset, n = 3u * (draws[d + 1].firstTri - draws[d].firstTri);
    uint32_t lo = 0xFFFFFFFFu, hi = 0u;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint32_t v = __ldg(indices + off + i);
        lo = min(lo, v); hi = max(hi, v);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        lo = min(lo, __shfl_xor_sync(0xFFFFFFFFu, lo, o));
        hi = max(hi, __shfl_xor_sync(0xFFFFFFFFu, hi, o));
    }
    if ((threadIdx.x & 31u) == 0 && lo <= hi) { atomicMin(&ranges[2 * d], lo); atomicMax(&ranges[2 * d + 1], hi); }
}

// One mip level from the previous one: 2x2 box average in UNORM16 with round-to-nearest (a 2x linear blit, Texture.cpp:50-136)
__global__ void __launch_bounds__(256) k_mip_down(const ushort4* __restrict__ src, ushort4* __restrict__ dst, uint32_t sw, uint32_t sh, uint32_t dw, uint32_t dh) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x, j = blockIdx.y;
    if (i >= dw || j >= dh) return;
    const uint32_t i0 = min(2u * i, sw - 1u), i1 = min(2u * i + 1u, sw - 1u), j0 = min(2u * j, sh - 1u), j1 = min(2u * j + 1u, sh - 1u);
    const ushort4 a = src[(size_t)j0 * sw + i0], b = src[(size_t)j0 * sw + i1], c = src[(size_t)j1 * sw + i0], d = src[(size_t)j1 * sw + i1];
    ushort4 o;
    o.x = (unsigned short)(((uint32_t)a.x + b.x + c.x + d.x + 2u) >> 2);
    o.y = (unsigned short)(((uint32_t)a.y + b.y + c.y + d.y + 2u) >> 2);
    o.z = (unsigned short)(((uint32_t)a.z + b.z + c.z + d.z + 2u) >> 2);
    o.w = (unsigned short)(((uint32_t)a.w + b.w + c.w + d.w + 2u) >> 2);
    dst[(size_t)j * dw + i] = o;
}

void launch_mip_down(const void* src, void* dst, uint32_t sw, uint32_t sh, uint32_t dw, uint32_t dh, cudaStream_t st) {
    k_mip_down<<<dim3((dw + 255u) / 256u, dh), 256, 0, st>>>(reinterpret_cast<const ushort4*>(src), reinterpret_cast<ushort4*>(dst), sw, sh, dw, dh);
}

void launch_vertex(const PassConst& pc, bool shadow, bool pdl, cudaStream_t st) {
    if (pc.cacheSize == 0) return;
    // at least K0_MIN_GRID CTAs: the pass prologue and the shadow-map clear are sliced over the grid (a 2 k-vertex scene
    // otherwise streams the 16-MB clear through three CTAs: 107 us); CTAs past the cache only do their slice
    const uint32_t per = 256u * (uint32_t)(pc.drawVisible ? K0_PER_THREAD_CULL : K0_PER_THREAD);
    const uint32_t grid = std::max((pc.cacheSize + per - 1u) / per, K0_MIN_GRID);
    using Kern = void (*)(PassConst);
    static const Kern kern[4] = {k_vertex<false, false>, k_vertex<true, false>, k_vertex<false, true>, k_vertex<true, true>};
    launch_k(kern[(shadow ? 1 : 0) | (pc.drawVisible ? 2 : 0)], dim3(grid), dim3(256), st, pdl, pc);
}

// ---- per-draw culling for the sort-first split ("each GPU runs K1 on all triangles (or culls to its region)") ------
// Object-space bounds of every draw, reduced once when a draw list changes.  Floats are kept in an order-preserving
// unsigned encoding so the reduction is atomicMin / atomicMax: bounds[6 d + 0..2] = min xyz, [3..5] = max xyz.
__device__ __forceinline__ uint32_t enc_ordered(float f) { const uint32_t b = __float_as_uint(f); return (b & 0x80000000u) ? ~b : (b | 0x80000000u); }
__device__ __forceinline__ float dec_ordered(uint32_t e) { return __uint_as_float((e & 0x80000000u) ? (e & 0x7FFFFFFFu) : ~e); }
__global__ void __launch_bounds__(256) k_draw_bounds(const uint32_t* __restrict__ indices, const float4* __restrict__ pos4, const DrawDev* __restrict__ draws,
                                                     uint32_t nDraws, uint32_t* __restrict__ bounds) {
    const uint32_t d = blockIdx.y;
    if (d >= nDraws) return;
    const uint32_t off = draws[d].indexOffset, n = 3u * (draws[d + 1].firstTri - draws[d].firstTri);
    uint32_t lo[3] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu}, hi[3] = {0u, 0u, 0u};
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float4 p = __ldg(pos4 + __ldg(indices + off + i));
        const uint32_t e[3] = {enc_ordered(p.x), enc_ordered(p.y), enc_ordered(p.z)};
#pragma unroll
        for (int c = 0; c < 3; ++c) { lo[c] = min(lo[c], e[c]); hi[c] = max(hi[c], e[c]); }
    }
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        lo[c] = __reduce_min_sync(0xFFFFFFFFu, lo[c]); hi[c] = __reduce_max_sync(0xFFFFFFFFu, hi[c]);
        if ((threadIdx.x & 31u) == 0 && lo[c] <= hi[c]) { atomicMin(&bounds[6 * d + c], lo[c]); atomicMax(&bounds[6 * d + 3 + c], hi[c]); }
    }
}
void launch_draw_bounds(const uint32_t* indices, const float4* pos4, const DrawDev* draws, uint32_t nDraws, uint32_t* bounds, cudaStream_t st) {
    for (uint32_t d0 = 0; d0 < nDraws; d0 += 65535u) {
        const uint32_t nd = std::min(65535u, nDraws - d0);
        k_draw_bounds<<<dim3(32, nd), 256, 0, st>>>(indices, pos4, draws + d0, nd, bounds + 6 * (size_t)d0);
    }
}
// Can the box [mn, mx] (object space, node transform T) touch a row of this rank?  The eight corners go through the pass's
// own vertex stage; a corner at or behind the eye plane keeps it (the clipper decides); one pixel of slack.
template <bool SHADOW>
__device__ __forceinline__ uint32_t box_touches_rank(const PassConst& pc, const float* T, const float mn[3], const float mx[3]) {
    const float hw = 0.5f * (float)pc.W, hh = 0.5f * (float)pc.H;
    float x0 = 3.0e38f, x1 = -3.0e38f, y0 = 3.0e38f, y1 = -3.0e38f;
    for (int c = 0; c < 8; ++c) {
        V3 world;
        const float4 q = vertex_stage<!SHADOW>(pc.PV, T, (c & 1) ? mx[0] : mn[0], (c & 2) ? mx[1] : mn[1], (c & 4) ? mx[2] : mn[2], &world);
        if (!(q.w > 0.0f)) return 1u;       // also catches NaN
        const float xf = hw * (q.x / q.w) + hw, yf = hh * (q.y / q.w) + hh;
        if (!(fabsf(xf) < 1.0e30f) || !(fabsf(yf) < 1.0e30f)) return 1u;
        x0 = fminf(x0, xf); x1 = fmaxf(x1, xf); y0 = fminf(y0, yf); y1 = fmaxf(y1, yf);
    }
    if (x1 < -1.0f || x0 > (float)pc.W + 1.0f || y1 < -1.0f || y0 > (float)pc.H + 1.0f) return 0u;
    if (pc.nRanks <= 1) return 1u;
    const uint32_t r0 = (uint32_t)fmaxf(y0 - 1.0f, 0.0f), r1 = (uint32_t)fminf(fmaxf(y1 + 1.0f, 0.0f), (float)(pc.H - 1u));
    const uint32_t b0 = r0 / pc.bandH, b1 = r1 / pc.bandH;
    if (b1 - b0 + 1u >= pc.nRanks) return 1u;
    for (uint32_t q = b0; q <= b1; ++q) if (q % pc.nRanks == pc.rank) return 1u;
    return 0u;
}
__device__ __forceinline__ void load_node_transform(const PassConst& pc, uint32_t node, float T[16]) {
    const float4* Tq = reinterpret_cast<const float4*>(pc.transforms) + 4u * node;
#pragma unroll
    for (int k = 0; k < 4; ++k) { const float4 q = Tq[k]; T[4 * k] = q.x; T[4 * k + 1] = q.y; T[4 * k + 2] = q.z; T[4 * k + 3] = q.w; }
}
// Once per pass and frame, ahead of K0, ONE launch: per-draw visibility (K0 skips the vertices of the other draws; K1 asks it per
// triangle in blocks that span draws) and per 256-triangle block (meshes come out of the re-ordering / the importer in spatially
// coherent order, so a block is a small patch).  blockBox[2 tb]: min xyz, draw (0xFFFFFFFF: the block spans draws); [2 tb + 1]: max xyz.
// A block that stays is appended to the compacted list K1 walks (one atomicAdd per warp; list order varies from frame to frame,
// which no result depends on: every resolve is a minimum) and stamps the 256-slot blocks of the post-transform cache its triangles
// reference (table built once per draw list by k_block_slot_refs) for K0.  Nothing is cleared per frame: the marks are a frame
// stamp, and the list length alternates between two words (this launch zeroes the other one for the next frame).  A block thread
// derives its draw's visibility itself instead of waiting for the per-draw pass (two launches, each ~6 us of latency, became one).
template <bool SHADOW>
__global__ void __launch_bounds__(128) k_cull(const __grid_constant__ PassConst pc, const uint32_t* __restrict__ bounds, uint32_t* __restrict__ drawVisible,
                                              const float4* __restrict__ blockBox, uint32_t nBlocks, uint32_t* __restrict__ visible, const uint32_t* __restrict__ refs,
                                              uint32_t* __restrict__ slotBlockVisible, uint32_t* __restrict__ list, uint32_t* __restrict__ count, uint32_t* __restrict__ nextCount) {
    const uint32_t i = blockIdx.x * 128u + threadIdx.x;
    if (i == 0) *nextCount = 0u;
    auto draw_visible = [&](uint32_t d) -> uint32_t {
        const uint32_t* b = bounds + 6 * (size_t)d;
        if (b[0] > b[3]) return 0u;   // draw without primitives
        const float mn[3] = {dec_ordered(b[0]), dec_ordered(b[1]), dec_ordered(b[2])}, mx[3] = {dec_ordered(b[3]), dec_ordered(b[4]), dec_ordered(b[5])};
        float T[16];
        load_node_transform(pc, pc.draws[d].node, T);
        return box_touches_rank<SHADOW>(pc, T, mn, mx);
    };
    if (SHADOW && pc.splitN > 1u) {   // sort-last split: a contiguous range of the block stream, every draw "visible"
        if (i < pc.nDraws) drawVisible[i] = 1u;
    } else if (i < pc.nDraws) drawVisible[i] = draw_visible(i);
    uint32_t vis = 0u;
    if (SHADOW && pc.splitN > 1u) {
        const uint32_t lo = (uint32_t)((unsigned long long)nBlocks * pc.splitRank / pc.splitN), hi = (uint32_t)((unsigned long long)nBlocks * (pc.splitRank + 1u) / pc.splitN);
        vis = (i >= lo && i < hi) ? 1u : 0u;
        if (i < nBlocks) visible[i] = vis;
    } else if (i < nBlocks) {
        const float4 lo = blockBox[2 * (size_t)i], hi = blockBox[2 * (size_t)i + 1];
        const uint32_t d = __float_as_uint(lo.w);
        if (d == 0xFFFFFFFFu) vis = BLOCK_MIXED;   // the block spans draws: K1 asks per triangle whether its draw is visible
        else if (draw_visible(d) != 0u) {
            const float mn[3] = {lo.x, lo.y, lo.z}, mx[3] = {hi.x, hi.y, hi.z};
            float T[16];
            load_node_transform(pc, pc.draws[d].node, T);
            vis = box_touches_rank<SHADOW>(pc, T, mn, mx);
        }
        visible[i] = vis;
    }
    const unsigned m = __ballot_sync(0xFFFFFFFFu, vis != 0u);
    if (!m) return;
    const unsigned lane = threadIdx.x & 31u;
    uint32_t base = 0;
    if (lane == (unsigned)(__ffs(m) - 1)) base = atomicAdd(count, (uint32_t)__popc(m));
    base = __shfl_sync(0xFFFFFFFFu, base, __ffs(m) - 1);
    if (vis == 0u) return;
    list[base + __popc(m & ((1u << lane) - 1u))] = i;
    const uint32_t* r = refs + (size_t)i * BLOCK_REFS;
    const uint32_t n = r[0];
    REB_ASSERT(n == 0xFFFFFFFFu ? (r[1] <= r[2] && r[2] < (pc.cacheSize + 255u) / 256u) : n < (uint32_t)BLOCK_REFS);
    if (n == 0xFFFFFFFFu) { for (uint32_t b = r[1]; b <= r[2]; ++b) slotBlockVisible[b] = pc.cullStamp; }
    else for (uint32_t k = 0; k < n; ++k) slotBlockVisible[r[1 + k]] = pc.cullStamp;
}
// Once per draw list: the distinct 256-slot blocks the triangles of every 256-triangle block reference (with slots in first-use
// order of a Morton-ordered stream: its own run of new slots plus a few runs shared with neighbouring patches).  One CTA per
// block, one triangle per thread, a 64-entry shared hash set; more than BLOCK_REFS - 1 distinct blocks: the [min, max] range.
__global__ void __launch_bounds__(256) k_block_slot_refs(const __grid_constant__ PassConst pc, uint32_t* __restrict__ refs) {
    __shared__ uint32_t tab[64];
    __shared__ uint32_t cnt, lo, hi, full;
    const uint32_t tb = blockIdx.x, tri = tb * 256u + threadIdx.x;
    if (threadIdx.x < 64) tab[threadIdx.x] = 0xFFFFFFFFu;
    if (threadIdx.x == 0) { cnt = 0u; lo = 0xFFFFFFFFu; hi = 0u; full = 0u; }
    __syncthreads();
    if (tri < pc.nTris) {
        uint32_t sl[3];
        if (pc.triStream) { const uint4 r = __ldg(pc.triStream + tri); sl[0] = r.x; sl[1] = r.y; sl[2] = r.z; }
        else if (pc.deindexed) { sl[0] = 3u * tri; sl[1] = 3u * tri + 1u; sl[2] = 3u * tri + 2u; }
        else { const uint32_t* f = pc.flatSlots + 3ull * tri; sl[0] = f[0]; sl[1] = f[1]; sl[2] = f[2]; }
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const uint32_t v = sl[k] >> 8;
            if (k > 0 && v == (sl[k - 1] >> 8)) continue;
            atomicMin(&lo, v); atomicMax(&hi, v);
            uint32_t h = (v * 2654435761u) >> 26;
            for (int probe = 0; probe < 64; ++probe, h = (h + 1u) & 63u) {
                const uint32_t old = atomicCAS(&tab[h], 0xFFFFFFFFu, v);
                if (old == 0xFFFFFFFFu || old == v) break;
                if (probe == 63) full = 1u;
            }
        }
    }
    __syncthreads();
    uint32_t* r = refs + (size_t)tb * BLOCK_REFS;
    if (threadIdx.x < 64 && tab[threadIdx.x] != 0xFFFFFFFFu) {
        const uint32_t i = atomicAdd(&cnt, 1u);
        if (i < (uint32_t)BLOCK_REFS - 1u) r[1 + i] = tab[threadIdx.x];
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        if (full || cnt > (uint32_t)BLOCK_REFS - 1u) { r[0] = 0xFFFFFFFFu; r[1] = lo; r[2] = hi; }
        else r[0] = cnt;
    }
}
void launch_block_slot_refs(const PassConst& pc, uint32_t* refs, cudaStream_t st) {
    if (pc.nTris == 0) return;
    k_block_slot_refs<<<(pc.nTris + 255u) / 256u, 256, 0, st>>>(pc, refs);
}
// Object-space box of every 256-triangle block, once per draw list: one CTA per block, one triangle per thread.
__global__ void __launch_bounds__(256) k_block_bounds(const __grid_constant__ PassConst pc, float4* __restrict__ blockBox) {
    __shared__ float red[8][6];
    __shared__ uint32_t dFirst, dLast;
    const uint32_t tb = blockIdx.x, tri = tb * 256u + threadIdx.x;
    float lo[3] = {3.0e38f, 3.0e38f, 3.0e38f}, hi[3] = {-3.0e38f, -3.0e38f, -3.0e38f};
    if (tri < pc.nTris) {   // (with a re-ordered list `tri` is a position in K1's stream: blocks are blocks of that stream)
        const uint32_t d = find_draw_block<false>(pc.draws, pc.nDraws, pc.triBlockDraw, tri);
        const uint32_t* ix = pc.indices + pc.draws[d].indexOffset + 3u * (tri - pc.draws[d].firstTri);
        uint32_t sl[3] = {0u, 0u, 0u};
        if (pc.triStream) { const uint4 r = __ldg(pc.triStream + tri); sl[0] = r.x; sl[1] = r.y; sl[2] = r.z; }
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const float4 p = pc.triStream ? __ldg(pc.posS + sl[k]) : __ldg(pc.pos4 + __ldg(ix + k));
            lo[0] = fminf(lo[0], p.x); lo[1] = fminf(lo[1], p.y); lo[2] = fminf(lo[2], p.z);
            hi[0] = fmaxf(hi[0], p.x); hi[1] = fmaxf(hi[1], p.y); hi[2] = fmaxf(hi[2], p.z);
        }
        if (threadIdx.x == 0) dFirst = d;
        if (tri == min(pc.nTris, tb * 256u + 256u) - 1u) dLast = d;
    }
#pragma unroll
    for (int c = 0; c < 3; ++c)
        for (int o = 16; o > 0; o >>= 1) {
            lo[c] = fminf(lo[c], __shfl_xor_sync(0xFFFFFFFFu, lo[c], o));
            hi[c] = fmaxf(hi[c], __shfl_xor_sync(0xFFFFFFFFu, hi[c], o));
        }
    if ((threadIdx.x & 31u) == 0) for (int c = 0; c < 3; ++c) { red[threadIdx.x >> 5][c] = lo[c]; red[threadIdx.x >> 5][3 + c] = hi[c]; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < 8; ++w) for (int c = 0; c < 3; ++c) { lo[c] = fminf(lo[c], red[w][c]); hi[c] = fmaxf(hi[c], red[w][3 + c]); }
        blockBox[2 * (size_t)tb] = make_float4(lo[0], lo[1], lo[2], __uint_as_float(dFirst == dLast ? dFirst : 0xFFFFFFFFu));
        blockBox[2 * (size_t)tb + 1] = make_float4(hi[0], hi[1], hi[2], 0.0f);
    }
}
void launch_block_bounds(const PassConst& pc, float4* blockBox, cudaStream_t st) {
    if (pc.nTris == 0) return;
    k_block_bounds<<<(pc.nTris + 255u) / 256u, 256, 0, st>>>(pc, blockBox);
}
void launch_cull(const PassConst& pc, bool shadow, const uint32_t* bounds, uint32_t* drawVisible, const float4* blockBox, uint32_t nBlocks, uint32_t* visible,
                 const uint32_t* refs, uint32_t* slotBlockVisible, uint32_t* list, uint32_t* count, uint32_t* nextCount, cudaStream_t st) {
    const uint32_t n = std::max(nBlocks, pc.nDraws);
    if (n == 0) return;
    if (shadow) k_cull<true><<<(n + 127u) / 128u, 128, 0, st>>>(pc, bounds, drawVisible, blockBox, nBlocks, visible, refs, slotBlockVisible, list, count, nextCount);
    else k_cull<false><<<(n + 127u) / 128u, 128, 0, st>>>(pc, bounds, drawVisible, blockBox, nBlocks, visible, refs, slotBlockVisible, list, count, nextCount);
}

void launch_vertex_both(const PassConst& ps, const PassConst& pf, cudaStream_t st) {
    if (pf.cacheSize == 0) return;
    const uint32_t grid = std::max((pf.cacheSize + 256u * K0_PER_THREAD - 1u) / (256u * K0_PER_THREAD), K0_MIN_GRID);
    launch_k(k_vertex_both, dim3(grid), dim3(256), st, false, ps, pf);
}

void launch_draw_ranges(const uint32_t* indices, const DrawDev* draws, uint32_t nDraws, uint32_t* ranges, cudaStream_t st) {
    if (nDraws == 0) return;
    for (uint32_t d0 = 0; d0 < nDraws; d0 += 65535u) {
        const uint32_t n = nDraws - d0 < 65535u ? nDraws - d0 : 65535u;
        k_draw_ranges<<<dim3(32, n), 256, 0, st>>>(indices, draws + d0, n, ranges + 2 * (size_t)d0);
    }
}

}  // namespace reb
