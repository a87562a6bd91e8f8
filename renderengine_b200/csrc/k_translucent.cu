// k_translucent.cu — the translucent meshes of the base pass (Appendix B.16 of DESIGN.md).
//
// The reference draws its AlphaMode::eBlend meshes after the opaque ones in the same render pass, sorted back to front
// per mesh by the caller (RenderScene.cpp:127-159,226-227), with cull none, depth test LessOrEqual + depth write and
// src-alpha blending (GraphicsPipelineCache.cpp:305-315,330-340): per pixel the result depends on the ORDER of the
// fragments, which the visibility-buffer resolve of the opaque pass cannot express.  Two kernels:
//   k_tl_raster   the translucent triangles (cull none: vertices 1 and 2 swap when A2 > 0), small ones per lane, medium ones
//                 per warp, large / clipped ones per CTA and 256 x 8 tile of their box (k_tl_raster_large): every covered pixel whose
//                 depth passes against the OPAQUE depth appends {id, z} to that pixel's linked list (any order);
//   k_tl_resolve  per pixel: pull the list, sort by submission-order ID, replay it — LessOrEqual against the running
//                 depth, shade at the pixel centre with the ordinary shade_pixel* code, C = Cs*As + Cd*(1-As),
//                 A = As*(1-As) in linear float on the colour the opaque pass left, depth / ID write — then the skybox
//                 (drawn after the translucent list, RenderScene.cpp:228) where the depth is still 1.0, then the encode.
// Same individually rounded operations as oracle/oracle.cpp.  A pixel's list has no length limit (up to TL_MAX_FRAGS fragments are
// sorted in registers / local memory, longer lists are replayed by selection from the list itself); the frame holds at most `cap`
// list nodes (the excess is dropped and counted).  Inside a 4x multisampled pass (Appendix B.15) the raster kernels run once per
// sample with the sample's shift and list-head plane, and k_tl_resolve_ms merges the four lists of a pixel.
#include "reb_shade.cuh"
#include "reb_shade_pbr.cuh"
#include "reb_sky.cuh"

namespace reb {

constexpr uint32_t TL_NONE = 0xFFFFFFFFu;
constexpr int TL_MAX_FRAGS = 16;

constexpr int TL_SMALL = 8;      // bounding box <= 8 x 8 pixel centres: rasterised by the lane that set it up
constexpr int TL_MEDIUM = 64;    // <= 64 x 64: by its warp; larger (or needs clipping): listed for k_tl_raster_large
constexpr int TL_TILE_W = 256, TL_TILE_H = 8;   // work item of a large triangle: one row per warp, 8 column steps

struct TlSetup {                 // Appendix B.6-B.8 of one snapped triangle, front-facing after the B.16 swap
    int32_t A[3], B[3], bias[3]; long long E[3];   // E = unbiased edge value at the centre of pixel (0, 0)
    float inv, z0, dz1, dz2;
    int x0, x1, y0, y1;
};

__device__ __forceinline__ void tl_load(const PassConst& pc, uint32_t tri, VtxA a[3]) {
    uint32_t s3[3] = {3u * tri, 3u * tri + 1u, 3u * tri + 2u};
    if (!pc.deindexed) { const uint32_t* f = pc.flatSlots + 3ull * tri; s3[0] = __ldg(f); s3[1] = __ldg(f + 1); s3[2] = __ldg(f + 2); }
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const uint32_t slot = s3[k];
        REB_ASSERT(tri < pc.nTris && slot < pc.cacheSize);
        const uint4 q = __ldg(reinterpret_cast<const uint4*>(pc.vtxA) + slot);
        a[k] = {(int32_t)q.x + pc.shiftX, (int32_t)q.y + pc.shiftY, __uint_as_float(q.z), q.w};   // + the B.15 sample shift (0 in a 1-sample pass)
    }
}

__device__ __forceinline__ bool tl_setup(const PassConst& pc, VtxA a[3], TlSetup& s) {
    long long A2 = mulw(a[1].X - a[0].X, a[2].Y - a[0].Y) - mulw(a[2].X - a[0].X, a[1].Y - a[0].Y);
    if (A2 == 0) return false;
    if (A2 > 0) { const VtxA t = a[1]; a[1] = a[2]; a[2] = t; A2 = -A2; }                          // cull none (Appendix B.16)
    const int32_t minX = min(a[0].X, min(a[1].X, a[2].X)), maxX = max(a[0].X, max(a[1].X, a[2].X));
    const int32_t minY = min(a[0].Y, min(a[1].Y, a[2].Y)), maxY = max(a[0].Y, max(a[1].Y, a[2].Y));
    s.x0 = max((minX + 127) >> 8, 0); s.x1 = min((maxX - 128) >> 8, (int32_t)pc.W - 1);
    s.y0 = max((minY + 127) >> 8, 0); s.y1 = min((maxY - 128) >> 8, (int32_t)pc.H - 1);
    if (s.x0 > s.x1 || s.y0 > s.y1) return false;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const int p = (k + 1) % 3, q = (k + 2) % 3;
        s.A[k] = a[q].Y - a[p].Y; s.B[k] = -(a[q].X - a[p].X);
        s.bias[k] = ((s.A[k] > 0) || (s.A[k] == 0 && s.B[k] > 0)) ? 0 : 1;   // top-left rule (Appendix B.7)
        s.E[k] = mulw(s.A[k], 128 - a[p].X) + mulw(s.B[k], 128 - a[p].Y);
    }
    s.inv = 1.0f / (float)(-A2);
    s.z0 = a[0].z; s.dz1 = a[1].z - a[0].z; s.dz2 = a[2].z - a[0].z;
    return true;
}

// Coverage + depth of pixel (x, y); a fragment that can still pass (z <= opaque depth: the running depth only ever gets
// nearer than that) is appended to the pixel's list.  The node index comes from one atomicAdd per converged group of lanes.
__device__ __forceinline__ void tl_sample(const PassConst& pc, const TlSetup& s, int x, int y, uint32_t tri, const float* __restrict__ opaqueDepth,
                                          uint32_t* __restrict__ head, uint4* __restrict__ nodes, uint32_t cap, uint32_t* __restrict__ counters) {
    const long long F0 = s.E[0] + mulw(s.A[0], 256 * x) + mulw(s.B[0], 256 * y);
    const long long F1 = s.E[1] + mulw(s.A[1], 256 * x) + mulw(s.B[1], 256 * y);
    const long long F2 = s.E[2] + mulw(s.A[2], 256 * x) + mulw(s.B[2], 256 * y);
    if (((F0 - s.bias[0]) | (F1 - s.bias[1]) | (F2 - s.bias[2])) < 0) return;
    const float z = depth_from((float)F1, (float)F2, s.inv, s.z0, s.dz1, s.dz2);
    if (!(z >= 0.0f && z <= 1.0f)) return;                                                       // Appendix B.8 depth clip
    const size_t p = (size_t)y * pc.W + x;
    if (!(z <= __ldg(opaqueDepth + p))) return;
    const unsigned m = __activemask();
    const unsigned lane = threadIdx.x & 31u;
    const int leader = __ffs(m) - 1;
    uint32_t base = 0;
    if ((int)lane == leader) base = atomicAdd(&counters[0], (uint32_t)__popc(m));
    base = __shfl_sync(m, base, leader);
    const uint32_t idx = base + (uint32_t)__popc(m & ((1u << lane) - 1u));
    if (idx >= cap) { atomicAdd(&counters[1], 1u); return; }
    nodes[idx] = make_uint4(atomicExch(&head[p], idx), tri, __float_as_uint(z), 0u);
}

// Coverage / depth only: true when pixel (x, y) yields a fragment that can still pass; z returned.
__device__ __forceinline__ bool tl_covers(const PassConst& pc, const TlSetup& s, int x, int y, const float* __restrict__ opaqueDepth, float& z) {
    const long long F0 = s.E[0] + mulw(s.A[0], 256 * x) + mulw(s.B[0], 256 * y);
    const long long F1 = s.E[1] + mulw(s.A[1], 256 * x) + mulw(s.B[1], 256 * y);
    const long long F2 = s.E[2] + mulw(s.A[2], 256 * x) + mulw(s.B[2], 256 * y);
    if (((F0 - s.bias[0]) | (F1 - s.bias[1]) | (F2 - s.bias[2])) < 0) return false;
    z = depth_from((float)F1, (float)F2, s.inv, s.z0, s.dz1, s.dz2);
    if (!(z >= 0.0f && z <= 1.0f)) return false;                                                 // Appendix B.8 depth clip
    return z <= __ldg(opaqueDepth + (size_t)y * pc.W + x);
}

// Pass A: one lane per translucent triangle, 32 at a time per warp.  Small triangles (the common case for dense meshes) are
// rasterised by their lane — coverage of the <= 8 x 8 box into a 64-bit mask first, then ONE node-pool atomicAdd for the
// whole warp, then the nodes; medium ones by the whole warp, one after the other; large ones and the ones that need
// Appendix B.3 clipping are listed for pass B with the number of 256 x 8 target tiles they may touch.
__global__ void __launch_bounds__(256) k_tl_raster(const __grid_constant__ PassConst pc, const float* __restrict__ opaqueDepth, uint32_t* __restrict__ head,
                                                   uint4* __restrict__ nodes, uint32_t cap, uint32_t* __restrict__ counters, uint2* __restrict__ largeList) {
    pdl_enter();
    const unsigned lane = threadIdx.x & 31u;
    const uint32_t gwarp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, nWarps = (gridDim.x * blockDim.x) >> 5;
    for (uint32_t base = gwarp * 32u; base < pc.nTris; base += nWarps * 32u) {
        const uint32_t tri = base + lane;
        int cls = 0;                                   // 0 nothing, 1 small, 2 medium, 3 large / needs clipping
        uint32_t nTiles = 0;
        TlSetup s;
        if (tri < pc.nTris) {
            VtxA a[3];
            tl_load(pc, tri, a);
            if ((a[0].flags & a[1].flags & a[2].flags & VF_PLANES) != 0u) cls = 0;                // Appendix B.2
            else if (((a[0].flags | a[1].flags | a[2].flags) & VF_NEEDS_CLIP) != 0u) {
                cls = 3; nTiles = ((pc.W + TL_TILE_W - 1) / TL_TILE_W) * ((pc.H + TL_TILE_H - 1) / TL_TILE_H);   // children: anywhere on the target
            } else if (tl_setup(pc, a, s)) {
                const int w = s.x1 - s.x0 + 1, h = s.y1 - s.y0 + 1;
                cls = (w <= TL_SMALL && h <= TL_SMALL) ? 1 : (w <= TL_MEDIUM && h <= TL_MEDIUM) ? 2 : 3;
                nTiles = (uint32_t)((s.x1 / TL_TILE_W - s.x0 / TL_TILE_W + 1) * (s.y1 / TL_TILE_H - s.y0 / TL_TILE_H + 1));
            }
        }
        if (cls == 3) {   // one 64-bit atomic hands out the list slot (high word) and the first work item (low word) consistently
            const unsigned long long t = atomicAdd(reinterpret_cast<unsigned long long*>(counters + 4), (1ull << 32) | nTiles);
            REB_ASSERT((uint32_t)(t >> 32) < pc.nTris);
            largeList[(uint32_t)(t >> 32)] = make_uint2(tri, (uint32_t)t);
        }
        // ---- small: coverage mask, one allocation per warp, emit
        unsigned long long mask = 0ull;
        int w = 1;
        if (cls == 1) {
            w = s.x1 - s.x0 + 1;
            int bit = 0;
            for (int y = s.y0; y <= s.y1; ++y) {
                const bool mine = row_owned((uint32_t)y, pc.rank, pc.nRanks, pc.bandH);
                for (int x = s.x0; x <= s.x1; ++x, ++bit) {
                    float z;
                    if (mine && tl_covers(pc, s, x, y, opaqueDepth, z)) mask |= 1ull << bit;
                }
            }
        }
        const uint32_t cnt = (uint32_t)__popcll(mask);
        uint32_t incl = cnt;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) { const uint32_t v = __shfl_up_sync(0xFFFFFFFFu, incl, d); if ((int)lane >= d) incl += v; }
        const uint32_t total = __shfl_sync(0xFFFFFFFFu, incl, 31);
        if (total) {
            uint32_t wbase = 0;
            if (lane == 0) wbase = atomicAdd(&counters[0], total);
            uint32_t idx = __shfl_sync(0xFFFFFFFFu, wbase, 0) + incl - cnt;
            while (mask) {
                const int bit = __ffsll((long long)mask) - 1;
                mask &= mask - 1;
                const int y = s.y0 + bit / w, x = s.x0 + bit - (bit / w) * w;
                float z;
                tl_covers(pc, s, x, y, opaqueDepth, z);
                if (idx >= cap) atomicAdd(&counters[1], 1u);
                else nodes[idx] = make_uint4(atomicExch(&head[(size_t)y * pc.W + x], idx), tri, __float_as_uint(z), 0u);
                ++idx;
            }
        }
        // ---- medium: the whole warp, one triangle after the other
        unsigned mm = __ballot_sync(0xFFFFFFFFu, cls == 2);
        while (mm) {
            const int src = __ffs(mm) - 1;
            mm &= mm - 1;
            const uint32_t t2 = base + (uint32_t)src;
            VtxA a[3];
            tl_load(pc, t2, a);
            TlSetup m;
            tl_setup(pc, a, m);
            for (int y = m.y0; y <= m.y1; ++y) {
                if (!row_owned((uint32_t)y, pc.rank, pc.nRanks, pc.bandH)) continue;
                for (int x = m.x0 + (int)lane; x <= m.x1; x += 32) tl_sample(pc, m, x, y, t2, opaqueDepth, head, nodes, cap, counters);
            }
        }
    }
}

// The part of one snapped triangle (a parent or a clip child) inside target tile (tx, ty) by the whole CTA: one row per warp, lanes over columns.
__device__ __forceinline__ void tl_raster_tile(const PassConst& pc, VtxA a[3], uint32_t tri, int tx, int ty, const float* __restrict__ opaqueDepth, uint32_t* __restrict__ head,
                                               uint4* __restrict__ nodes, uint32_t cap, uint32_t* __restrict__ counters, unsigned lane, unsigned warp) {
    TlSetup s;
    if (!tl_setup(pc, a, s)) return;
    const int y = ty * TL_TILE_H + (int)warp;
    if (y < s.y0 || y > s.y1 || !row_owned((uint32_t)y, pc.rank, pc.nRanks, pc.bandH)) return;
    const int xa = max(s.x0, tx * TL_TILE_W), xb = min(s.x1, tx * TL_TILE_W + TL_TILE_W - 1);
    // all eight column steps in flight: coverage + depth first (independent loads), ONE node-pool atomicAdd for the row, then the nodes
    float z[TL_TILE_W / 32];
    unsigned covered = 0u;
#pragma unroll
    for (int k = 0; k < TL_TILE_W / 32; ++k) {
        const int x = xa + (int)lane + 32 * k;
        z[k] = 0.0f;
        if (x <= xb && tl_covers(pc, s, x, y, opaqueDepth, z[k])) covered |= 1u << k;
    }
    const uint32_t cnt = (uint32_t)__popc(covered);
    uint32_t incl = cnt;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const uint32_t v = __shfl_up_sync(0xFFFFFFFFu, incl, d); if ((int)lane >= d) incl += v; }
    const uint32_t total = __shfl_sync(0xFFFFFFFFu, incl, 31);
    if (!total) return;
    uint32_t wbase = 0;
    if (lane == 0) wbase = atomicAdd(&counters[0], total);
    uint32_t idx = __shfl_sync(0xFFFFFFFFu, wbase, 0) + incl - cnt;
#pragma unroll
    for (int k = 0; k < TL_TILE_W / 32; ++k) {
        if (!((covered >> k) & 1u)) continue;
        const int x = xa + (int)lane + 32 * k;
        if (idx >= cap) atomicAdd(&counters[1], 1u);
        else nodes[idx] = make_uint4(atomicExch(&head[(size_t)y * pc.W + x], idx), tri, __float_as_uint(z[k]), 0u);
        ++idx;
    }
}

// Pass B: the listed triangles, one CTA per (triangle, 256 x 8 tile of its bounding box).  Appendix B.3 for a translucent
// parent runs CTA-uniform (every thread clips the same triangle); the children of the fan go through the same snap / facing /
// raster steps and append the PARENT's ID (children of one fan share edges, which the top-left rule assigns to exactly one).
__global__ void __launch_bounds__(256) k_tl_raster_large(const __grid_constant__ PassConst pc, const float* __restrict__ opaqueDepth, uint32_t* __restrict__ head,
                                                         uint4* __restrict__ nodes, uint32_t cap, uint32_t* __restrict__ counters, const uint2* __restrict__ largeList) {
    pdl_enter();
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const uint32_t nItems = __ldcg(&counters[4]), nLarge = __ldcg(&counters[5]);
    for (uint32_t item = blockIdx.x; item < nItems; item += gridDim.x) {
        uint32_t lo = 0, hi = nLarge;                // last entry whose first work item is <= item
        while (hi - lo > 1) { const uint32_t mid = (lo + hi) >> 1; if (__ldcg(&largeList[mid].y) <= item) lo = mid; else hi = mid; }
        const uint2 e = __ldcg(largeList + lo);
        const uint32_t tri = e.x, local = item - e.y;
        VtxA a[3];
        tl_load(pc, tri, a);
        if (((a[0].flags | a[1].flags | a[2].flags) & VF_NEEDS_CLIP) == 0u) {
            TlSetup s;
            VtxA b[3] = {a[0], a[1], a[2]};
            if (!tl_setup(pc, b, s)) continue;
            const int tx0 = s.x0 / TL_TILE_W, tw = s.x1 / TL_TILE_W - tx0 + 1;
            tl_raster_tile(pc, a, tri, tx0 + (int)(local % (uint32_t)tw), s.y0 / TL_TILE_H + (int)(local / (uint32_t)tw), opaqueDepth, head, nodes, cap, counters, lane, warp);
            continue;
        }
        const uint32_t tw = (pc.W + TL_TILE_W - 1) / TL_TILE_W;
        float4 c[9];
        const int n = clip_parent<false>(pc, tri, c);
        for (int i = 1; i + 1 < n; ++i) {
            VtxA ch[3] = {viewport_snap(c[0], pc.W, pc.H), viewport_snap(c[i], pc.W, pc.H), viewport_snap(c[i + 1], pc.W, pc.H)};
#pragma unroll
            for (int k = 0; k < 3; ++k) { ch[k].X += pc.shiftX; ch[k].Y += pc.shiftY; }
            if (((ch[0].flags | ch[1].flags | ch[2].flags) & VF_NEEDS_CLIP) != 0u) {   // cannot happen for finite inputs: counted, dropped
                if (threadIdx.x == 0 && local == 0) atomicAdd(&counters[2], 1u);
                continue;
            }
            tl_raster_tile(pc, ch, tri, (int)(local % tw), (int)(local / tw), opaqueDepth, head, nodes, cap, counters, lane, warp);
        }
    }
}

// Ordered walk over a pixel's (unsorted) list without storing it: the fragment with the smallest submission-order ID above `prev`
// (one list traversal per step; only lists longer than TL_MAX_FRAGS and the per-sample lists of a multisampled pass use it).
struct TlCursor { uint32_t id; float z; };
__device__ __forceinline__ TlCursor tl_scan(const uint4* __restrict__ nodes, uint32_t cap, uint32_t h, bool any, uint32_t prev) {
    TlCursor c = {TL_NONE, 0.0f};
    for (uint32_t q = h; q != TL_NONE && q < cap;) {
        const uint4 nd = nodes[q];
        if ((any || nd.y > prev) && nd.y < c.id) { c.id = nd.y; c.z = __uint_as_float(nd.z); }
        q = nd.x;
    }
    return c;
}
__device__ __forceinline__ TlCursor tl_first(const uint4* __restrict__ nodes, uint32_t cap, uint32_t h) { return tl_scan(nodes, cap, h, true, 0u); }
__device__ __forceinline__ TlCursor tl_next(const uint4* __restrict__ nodes, uint32_t cap, uint32_t h, uint32_t prev) { return tl_scan(nodes, cap, h, false, prev); }

template <bool TEXTURED, bool PBR, bool AFFINE>
__global__ void __launch_bounds__(256, 2) k_tl_resolve(const __grid_constant__ PassConst pc, const __grid_constant__ ShadeConst sh, uint32_t* __restrict__ head,
                                                       const uint4* __restrict__ nodes, uint32_t cap, uint32_t idBase, float* __restrict__ hdr, float* __restrict__ depth,
                                                       uint32_t* __restrict__ counters) {
    pdl_enter();
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const uint32_t ty = blockIdx.x / pc.tilesX, tx = blockIdx.x - ty * pc.tilesX;
    if (!row_owned(ty * 8u, pc.rank, pc.nRanks, pc.bandH)) return;
    const int x = (int)tx * 32 + 8 * (int)(warp & 3u) + (int)(lane & 7u), y = (int)ty * 8 + 4 * (int)(warp >> 2) + (int)(lane >> 3);
    if (x >= (int)pc.W || y >= (int)pc.H) return;
    const size_t p = (size_t)y * pc.W + x;
    float4 col = reinterpret_cast<const float4*>(hdr)[p];
    float d = depth[p];
    uint32_t lastId = TL_NONE;
    uint32_t h = head[p];
    if (h != TL_NONE) {
        head[p] = TL_NONE;                       // re-armed for the next frame
        uint32_t id[TL_MAX_FRAGS]; float z[TL_MAX_FRAGS];
        int n = 0;
        bool longList = false;                   // more than TL_MAX_FRAGS fragments: replayed by selection straight from the list
        for (uint32_t q = h; q != TL_NONE && q < cap;) {
            const uint4 nd = nodes[q];
            if (n < TL_MAX_FRAGS) {              // insertion by submission-order ID
                int i = n++;
                while (i > 0 && id[i - 1] > nd.y) { id[i] = id[i - 1]; z[i] = z[i - 1]; --i; }
                id[i] = nd.y; z[i] = __uint_as_float(nd.z);
            } else {
                longList = true;
                break;
            }
            q = nd.x;
        }
        if (!longList) {
            for (int i = 0; i < n; ++i) {
                if (!(z[i] <= d)) continue;      // LessOrEqual against the running depth (depth write on)
                const float4 src = PBR ? shade_pixel_pbr<TEXTURED, AFFINE>(pc, sh, x, y, id[i]) : shade_pixel<TEXTURED, AFFINE>(pc, sh, x, y, id[i]);
                const float a = src.w, ia = 1.0f - a;
                col.x = src.x * a + col.x * ia; col.y = src.y * a + col.y * ia; col.z = src.z * a + col.z * ia;
                col.w = a * ia;
                d = z[i]; lastId = id[i];
            }
        } else {
            TlCursor c = tl_first(nodes, cap, h);
            while (c.id != TL_NONE) {
                if (c.z <= d) {
                    const float4 src = PBR ? shade_pixel_pbr<TEXTURED, AFFINE>(pc, sh, x, y, c.id) : shade_pixel<TEXTURED, AFFINE>(pc, sh, x, y, c.id);
                    const float a = src.w, ia = 1.0f - a;
                    col.x = src.x * a + col.x * ia; col.y = src.y * a + col.y * ia; col.z = src.z * a + col.z * ia;
                    col.w = a * ia;
                    d = c.z; lastId = c.id;
                }
                c = tl_next(nodes, cap, h, c.id);
            }
        }
    }
    if (sh.skyFaces && d == 1.0f) col = sky_pixel(sh, x, y);   // the skybox follows the translucent list (RenderScene.cpp:228)
    if (lastId != TL_NONE) {
        depth[p] = d;
        if (sh.outTriId) sh.outTriId[p] = lastId + idBase;
    }
    if (sh.outDepth && sh.outDepth != depth) sh.outDepth[p] = d;
    if (sh.outSRGB8) reinterpret_cast<uint32_t*>(sh.outSRGB8)[p] = pack_srgb8(col);
    if (sh.outF32) reinterpret_cast<float4*>(sh.outF32)[p] = col;
    else reinterpret_cast<float4*>(hdr)[p] = col;
}

// ---------------------------------------------------------------------------------------------------------------------------------
// The 1-sample resolve as three kernels (k_tl_resolve above does the same in one: 117-128 registers because the list replay and
// the whole fragment shader are live together, 20 % occupancy, and a warp shades as many rounds as its longest list):
//   k_tl_gather  per pixel: pull the list, order it by ID, replay ONLY the depth test (LessOrEqual + write) — which fragments pass is
//                known without shading — and emit the passing ones, in order, as {pixel, id} records into a compact array (one
//                atomicAdd per warp); leaves the final depth / ID and the pixel's span {first record, count}; light, full occupancy;
//   k_tl_shade   one thread per PASSING FRAGMENT of the whole frame: the ordinary shade_pixel* at the fragment's pixel centre, colour
//                stored; every lane busy, the shader's registers only;
//   k_tl_blend   per pixel: blend the span's colours in order over the opaque colour (same three individually rounded operations per
//                channel), then the skybox where the depth is still 1.0, then the encode.
// ---------------------------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_tl_gather(const __grid_constant__ PassConst pc, uint32_t* __restrict__ head, const uint4* __restrict__ nodes, uint32_t cap,
                                                   uint32_t idBase, float* __restrict__ depth, uint32_t* __restrict__ outTriId, float* __restrict__ outDepth,
                                                   uint2* __restrict__ span, uint2* __restrict__ frags, uint32_t* __restrict__ counters) {
    pdl_enter();
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const uint32_t ty = blockIdx.x / pc.tilesX, tx = blockIdx.x - ty * pc.tilesX;
    if (!row_owned(ty * 8u, pc.rank, pc.nRanks, pc.bandH)) return;
    const int x = (int)tx * 32 + 8 * (int)(warp & 3u) + (int)(lane & 7u), y = (int)ty * 8 + 4 * (int)(warp >> 2) + (int)(lane >> 3);
    const bool inb = x < (int)pc.W && y < (int)pc.H;
    const size_t p = inb ? (size_t)y * pc.W + x : 0;
    float d = inb ? depth[p] : 0.0f;
    uint32_t lastId = TL_NONE;
    uint32_t h = inb ? head[p] : TL_NONE;
    uint32_t pass[TL_MAX_FRAGS];                 // IDs of the passing fragments, in order (fast path)
    int nPass = 0;
    bool longList = false;
    if (h != TL_NONE) {
        head[p] = TL_NONE;                       // re-armed for the next frame
        uint32_t id[TL_MAX_FRAGS]; float z[TL_MAX_FRAGS];
        int n = 0;
        for (uint32_t q = h; q != TL_NONE && q < cap;) {
            const uint4 nd = nodes[q];
            if (n < TL_MAX_FRAGS) {
                int i = n++;
                while (i > 0 && id[i - 1] > nd.y) { id[i] = id[i - 1]; z[i] = z[i - 1]; --i; }
                id[i] = nd.y; z[i] = __uint_as_float(nd.z);
            } else { longList = true; break; }
            q = nd.x;
        }
        if (!longList) {
            for (int i = 0; i < n; ++i)
                if (z[i] <= d) { pass[nPass++] = id[i]; d = z[i]; lastId = id[i]; }
        } else {                                 // count first (selection walk), emit below with the same walk
            float dd = d;
            for (TlCursor c = tl_first(nodes, cap, h); c.id != TL_NONE; c = tl_next(nodes, cap, h, c.id))
                if (c.z <= dd) { ++nPass; dd = c.z; }
        }
    }
    // one allocation per warp
    uint32_t incl = (uint32_t)nPass;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const uint32_t v = __shfl_up_sync(0xFFFFFFFFu, incl, o); if ((int)lane >= o) incl += v; }
    const uint32_t total = __shfl_sync(0xFFFFFFFFu, incl, 31);
    uint32_t wbase = 0;
    if (total) {
        if (lane == 0) wbase = atomicAdd(&counters[6], total);
        wbase = __shfl_sync(0xFFFFFFFFu, wbase, 0);
    }
    if (!inb) return;
    const uint32_t base = wbase + incl - (uint32_t)nPass;
    if (nPass) {
        if (!longList) {
            for (int i = 0; i < nPass; ++i) if (base + i < cap) frags[base + i] = make_uint2((uint32_t)p, pass[i]);
        } else {
            uint32_t k = 0;
            for (TlCursor c = tl_first(nodes, cap, h); c.id != TL_NONE; c = tl_next(nodes, cap, h, c.id))
                if (c.z <= d) { if (base + k < cap) frags[base + k] = make_uint2((uint32_t)p, c.id); ++k; d = c.z; lastId = c.id; }
        }
        span[p] = make_uint2(base, (uint32_t)nPass);
    }
    if (lastId != TL_NONE) {
        depth[p] = d;
        if (outTriId) outTriId[p] = lastId + idBase;
    }
    if (outDepth && outDepth != depth) outDepth[p] = d;
}

template <bool TEXTURED, bool PBR, bool AFFINE>
__global__ void __launch_bounds__(256, 3) k_tl_shade(const __grid_constant__ PassConst pc, const __grid_constant__ ShadeConst sh, const uint2* __restrict__ frags,
                                                     float4* __restrict__ cols, uint32_t cap, const uint32_t* __restrict__ counters) {
    pdl_enter();
    const uint32_t n = min(__ldcg(&counters[6]), cap);
    for (uint32_t i = blockIdx.x * 256u + threadIdx.x; i < n; i += gridDim.x * 256u) {
        const uint2 f = __ldg(frags + i);
        const uint32_t pix = f.x & 0x0FFFFFFFu;   // (bits 28-31: the sample mask of a multisampled pass; extents are <= 16384, so a pixel index fits 28 bits)
        const int y = (int)(pix / pc.W), x = (int)(pix - (uint32_t)y * pc.W);
        cols[i] = PBR ? shade_pixel_pbr<TEXTURED, AFFINE>(pc, sh, x, y, f.y) : shade_pixel<TEXTURED, AFFINE>(pc, sh, x, y, f.y);
    }
}

__global__ void __launch_bounds__(256) k_tl_blend(const __grid_constant__ PassConst pc, const __grid_constant__ ShadeConst sh, uint2* __restrict__ span,
                                                  const float4* __restrict__ cols, uint32_t cap, float* __restrict__ hdr, const float* __restrict__ depth) {
    pdl_enter();
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const uint32_t ty = blockIdx.x / pc.tilesX, tx = blockIdx.x - ty * pc.tilesX;
    if (!row_owned(ty * 8u, pc.rank, pc.nRanks, pc.bandH)) return;
    const int x = (int)tx * 32 + 8 * (int)(warp & 3u) + (int)(lane & 7u), y = (int)ty * 8 + 4 * (int)(warp >> 2) + (int)(lane >> 3);
    if (x >= (int)pc.W || y >= (int)pc.H) return;
    const size_t p = (size_t)y * pc.W + x;
    float4 col = reinterpret_cast<const float4*>(hdr)[p];
    const uint2 sp = span[p];
    if (sp.y) {
        span[p] = make_uint2(0u, 0u);            // re-armed for the next frame
        for (uint32_t i = 0; i < sp.y && sp.x + i < cap; ++i) {
            const float4 src = __ldg(cols + sp.x + i);
            const float a = src.w, ia = 1.0f - a;
            col.x = src.x * a + col.x * ia; col.y = src.y * a + col.y * ia; col.z = src.z * a + col.z * ia;
            col.w = a * ia;
        }
    }
    if (sh.skyFaces && depth[p] == 1.0f) col = sky_pixel(sh, x, y);   // the skybox follows the translucent list (RenderScene.cpp:228)
    if (sh.outSRGB8) reinterpret_cast<uint32_t*>(sh.outSRGB8)[p] = pack_srgb8(col);
    if (sh.outF32) reinterpret_cast<float4*>(sh.outF32)[p] = col;
    else reinterpret_cast<float4*>(hdr)[p] = col;
}

// Appendix B.15 + B.16: the translucent list inside the 4x multisampled base pass.  Every sample is its own blend target: the
// opaque resolve left one linear colour per sample (msColour) and the four sample passes of the translucent raster one list per
// sample.  Per pixel the four lists are merged in ID order; a primitive is shaded ONCE at the pixel centre when any of its
// samples passes LessOrEqual against that sample's running depth, and blended into each passing sample; then the skybox per
// sample (depth still 1.0), the average resolve ((s0 + s1) + (s2 + s3)) / 4 and the encode.  The per-sample depth / ID planes
// are updated; the depth / ID targets take sample 0.
template <bool TEXTURED, bool PBR, bool AFFINE>
__global__ void __launch_bounds__(256, 2) k_tl_resolve_ms(const __grid_constant__ PassConst pc, const __grid_constant__ ShadeConst sh, uint32_t* __restrict__ head,
                                                          const uint4* __restrict__ nodes, uint32_t cap, uint32_t idBase, const float4* __restrict__ msColour,
                                                          float* __restrict__ msDepth, uint32_t* __restrict__ msTriId) {
    pdl_enter();
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const uint32_t ty = blockIdx.x / pc.tilesX, tx = blockIdx.x - ty * pc.tilesX;
    if (!row_owned(ty * 8u, pc.rank, pc.nRanks, pc.bandH)) return;
    const int x = (int)tx * 32 + 8 * (int)(warp & 3u) + (int)(lane & 7u), y = (int)ty * 8 + 4 * (int)(warp >> 2) + (int)(lane >> 3);
    if (x >= (int)pc.W || y >= (int)pc.H) return;
    const size_t p = (size_t)y * pc.W + x, plane = (size_t)pc.W * pc.H;
    float4 col[4]; float d[4]; uint32_t h[4]; TlCursor c[4]; uint32_t lastId[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        col[k] = msColour[k * plane + p];
        d[k] = msDepth[k * plane + p];
        h[k] = head[k * plane + p];
        lastId[k] = TL_NONE;
        if (h[k] != TL_NONE) head[k * plane + p] = TL_NONE;   // re-armed for the next frame
        c[k] = tl_first(nodes, cap, h[k]);
    }
    for (;;) {
        const uint32_t id = min(min(c[0].id, c[1].id), min(c[2].id, c[3].id));
        if (id == TL_NONE) break;
        bool pass[4]; bool any = false;
#pragma unroll
        for (int k = 0; k < 4; ++k) { pass[k] = c[k].id == id && c[k].z <= d[k]; any = any || pass[k]; }
        if (any) {
            const float4 src = PBR ? shade_pixel_pbr<TEXTURED, AFFINE>(pc, sh, x, y, id) : shade_pixel<TEXTURED, AFFINE>(pc, sh, x, y, id);
            const float a = src.w, ia = 1.0f - a;
#pragma unroll
            for (int k = 0; k < 4; ++k)
                if (pass[k]) {
                    col[k].x = src.x * a + col[k].x * ia; col[k].y = src.y * a + col[k].y * ia; col[k].z = src.z * a + col[k].z * ia;
                    col[k].w = a * ia;
                    d[k] = c[k].z; lastId[k] = id;
                }
        }
#pragma unroll
        for (int k = 0; k < 4; ++k)
            if (c[k].id == id) c[k] = tl_next(nodes, cap, h[k], id);
    }
    bool anySky = false;
#pragma unroll
    for (int k = 0; k < 4; ++k) anySky = anySky || (sh.skyFaces && d[k] == 1.0f);
    if (anySky) {
        const float4 s = sky_pixel(sh, x, y);
#pragma unroll
        for (int k = 0; k < 4; ++k) if (d[k] == 1.0f) col[k] = s;
    }
#pragma unroll
    for (int k = 0; k < 4; ++k)
        if (lastId[k] != TL_NONE) { msDepth[k * plane + p] = d[k]; msTriId[k * plane + p] = lastId[k] + idBase; }
    float4 out;
    out.x = ((col[0].x + col[1].x) + (col[2].x + col[3].x)) * 0.25f;
    out.y = ((col[0].y + col[1].y) + (col[2].y + col[3].y)) * 0.25f;
    out.z = ((col[0].z + col[1].z) + (col[2].z + col[3].z)) * 0.25f;
    out.w = ((col[0].w + col[1].w) + (col[2].w + col[3].w)) * 0.25f;
    if (sh.outDepth) sh.outDepth[p] = d[0];
    if (sh.outTriId) sh.outTriId[p] = lastId[0] != TL_NONE ? lastId[0] + idBase : msTriId[p];   // (the opaque resolve left the 1-plane targets to this kernel)
    if (sh.outSRGB8) reinterpret_cast<uint32_t*>(sh.outSRGB8)[p] = pack_srgb8(out);
    if (sh.outF32) reinterpret_cast<float4*>(sh.outF32)[p] = out;
}

// The multisampled resolve (k_tl_resolve_ms) as the same three kernels: the gather merges the pixel's four per-sample lists in ID order
// (four selection cursors, walked twice: count, then emit), replays the per-sample depth tests and emits one record per primitive
// that passes at any sample, {pixel | sample mask << 28, id}; k_tl_shade shades the records; the blend applies each colour to the
// samples of its mask, draws the sky per sample, resolves and encodes.
__global__ void __launch_bounds__(256) k_tl_gather_ms(const __grid_constant__ PassConst pc, uint32_t* __restrict__ head, const uint4* __restrict__ nodes, uint32_t cap,
                                                      uint32_t idBase, float* __restrict__ msDepth, uint32_t* __restrict__ msTriId, uint2* __restrict__ span,
                                                      uint2* __restrict__ frags, uint32_t* __restrict__ counters) {
    pdl_enter();
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const uint32_t ty = blockIdx.x / pc.tilesX, tx = blockIdx.x - ty * pc.tilesX;
    if (!row_owned(ty * 8u, pc.rank, pc.nRanks, pc.bandH)) return;
    const int x = (int)tx * 32 + 8 * (int)(warp & 3u) + (int)(lane & 7u), y = (int)ty * 8 + 4 * (int)(warp >> 2) + (int)(lane >> 3);
    const bool inb = x < (int)pc.W && y < (int)pc.H;
    const size_t p = inb ? (size_t)y * pc.W + x : 0, plane = (size_t)pc.W * pc.H;
    float d0[4]; uint32_t h[4];
    bool any = false;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        d0[k] = inb ? msDepth[k * plane + p] : 0.0f;
        h[k] = inb ? head[k * plane + p] : TL_NONE;
        if (h[k] != TL_NONE) { head[k * plane + p] = TL_NONE; any = true; }   // re-armed for the next frame
    }
    // the merge, run twice: emit == false counts the records, emit == true writes them (and the planes)
    auto merge = [&](bool emit, uint32_t base) -> uint32_t {
        float d[4] = {d0[0], d0[1], d0[2], d0[3]};
        uint32_t lastId[4] = {TL_NONE, TL_NONE, TL_NONE, TL_NONE};
        TlCursor c[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) c[k] = tl_first(nodes, cap, h[k]);
        uint32_t n = 0;
        for (;;) {
            const uint32_t id = min(min(c[0].id, c[1].id), min(c[2].id, c[3].id));
            if (id == TL_NONE) break;
            uint32_t mask = 0u;
#pragma unroll
            for (int k = 0; k < 4; ++k)
                if (c[k].id == id && c[k].z <= d[k]) { mask |= 1u << k; d[k] = c[k].z; lastId[k] = id; }
            if (mask) {
                if (emit && base + n < cap) frags[base + n] = make_uint2((uint32_t)p | (mask << 28), id);
                ++n;
            }
#pragma unroll
            for (int k = 0; k < 4; ++k)
                if (c[k].id == id) c[k] = tl_next(nodes, cap, h[k], id);
        }
        if (emit) {
#pragma unroll
            for (int k = 0; k < 4; ++k)
                if (lastId[k] != TL_NONE) { msDepth[k * plane + p] = d[k]; msTriId[k * plane + p] = lastId[k] + idBase; }
        }
        return n;
    };
    const uint32_t nRec = any ? merge(false, 0u) : 0u;
    uint32_t incl = nRec;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const uint32_t v = __shfl_up_sync(0xFFFFFFFFu, incl, o); if ((int)lane >= o) incl += v; }
    const uint32_t total = __shfl_sync(0xFFFFFFFFu, incl, 31);
    uint32_t wbase = 0;
    if (total) {
        if (lane == 0) wbase = atomicAdd(&counters[6], total);
        wbase = __shfl_sync(0xFFFFFFFFu, wbase, 0);
    }
    if (!inb || !nRec) return;
    const uint32_t base = wbase + incl - nRec;
    merge(true, base);
    span[p] = make_uint2(base, nRec);
}

__global__ void __launch_bounds__(256) k_tl_blend_ms(const __grid_constant__ PassConst pc, const __grid_constant__ ShadeConst sh, uint2* __restrict__ span,
                                                     const uint2* __restrict__ frags, const float4* __restrict__ cols, uint32_t cap, const float4* __restrict__ msColour,
                                                     const float* __restrict__ msDepth, const uint32_t* __restrict__ msTriId) {
    pdl_enter();
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const uint32_t ty = blockIdx.x / pc.tilesX, tx = blockIdx.x - ty * pc.tilesX;
    if (!row_owned(ty * 8u, pc.rank, pc.nRanks, pc.bandH)) return;
    const int x = (int)tx * 32 + 8 * (int)(warp & 3u) + (int)(lane & 7u), y = (int)ty * 8 + 4 * (int)(warp >> 2) + (int)(lane >> 3);
    if (x >= (int)pc.W || y >= (int)pc.H) return;
    const size_t p = (size_t)y * pc.W + x, plane = (size_t)pc.W * pc.H;
    float4 col[4]; float d[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) { col[k] = msColour[k * plane + p]; d[k] = msDepth[k * plane + p]; }
    const uint2 sp = span[p];
    if (sp.y) {
        span[p] = make_uint2(0u, 0u);
        for (uint32_t i = 0; i < sp.y && sp.x + i < cap; ++i) {
            const uint32_t mask = __ldg(&frags[sp.x + i].x) >> 28;
            const float4 src = __ldg(cols + sp.x + i);
            const float a = src.w, ia = 1.0f - a;
#pragma unroll
            for (int k = 0; k < 4; ++k)
                if ((mask >> k) & 1u) {
                    col[k].x = src.x * a + col[k].x * ia; col[k].y = src.y * a + col[k].y * ia; col[k].z = src.z * a + col[k].z * ia;
                    col[k].w = a * ia;
                }
        }
    }
    bool anySky = false;
#pragma unroll
    for (int k = 0; k < 4; ++k) anySky = anySky || (sh.skyFaces && d[k] == 1.0f);
    if (anySky) {
        const float4 s = sky_pixel(sh, x, y);
#pragma unroll
        for (int k = 0; k < 4; ++k) if (d[k] == 1.0f) col[k] = s;
    }
    float4 out;
    out.x = ((col[0].x + col[1].x) + (col[2].x + col[3].x)) * 0.25f;
    out.y = ((col[0].y + col[1].y) + (col[2].y + col[3].y)) * 0.25f;
    out.z = ((col[0].z + col[1].z) + (col[2].z + col[3].z)) * 0.25f;
    out.w = ((col[0].w + col[1].w) + (col[2].w + col[3].w)) * 0.25f;
    if (sh.outDepth) sh.outDepth[p] = d[0];
    if (sh.outTriId) sh.outTriId[p] = msTriId[p];
    if (sh.outSRGB8) reinterpret_cast<uint32_t*>(sh.outSRGB8)[p] = pack_srgb8(out);
    if (sh.outF32) reinterpret_cast<float4*>(sh.outF32)[p] = out;
}

void launch_tl_raster(const PassConst& pc, const float* opaqueDepth, uint32_t* head, uint4* nodes, uint32_t cap, uint32_t* counters, uint2* largeList, int numSMs,
                      cudaStream_t st) {
    if (pc.nTris == 0) return;
    const uint32_t grid = std::min<uint32_t>((pc.nTris + 255u) / 256u, (uint32_t)numSMs * 8u);
    launch_k(k_tl_raster, dim3(grid), dim3(256), st, true, pc, opaqueDepth, head, nodes, cap, counters, largeList);
    launch_k(k_tl_raster_large, dim3((uint32_t)numSMs * 4u), dim3(256), st, true, pc, opaqueDepth, head, nodes, cap, counters, (const uint2*)largeList);
}

void launch_tl_resolve(const PassConst& pc, const ShadeConst& sh, uint32_t* head, const uint4* nodes, uint32_t cap, uint32_t idBase, float* hdr, float* depth,
                       uint32_t* counters, cudaStream_t st) {
    const bool tex = sh.nTextures && pc.vtxUV, pbr = sh.shadingModel == 1u;
    const bool affine = sh.STR[3] == 0.0f && sh.STR[7] == 0.0f && sh.STR[11] == 0.0f && sh.STR[15] == 1.0f;
    using Kern = void (*)(PassConst, ShadeConst, uint32_t*, const uint4*, uint32_t, uint32_t, float*, float*, uint32_t*);
    static const Kern kern[8] = {k_tl_resolve<false, false, false>, k_tl_resolve<true, false, false>, k_tl_resolve<false, true, false>, k_tl_resolve<true, true, false>,
                                 k_tl_resolve<false, false, true>,  k_tl_resolve<true, false, true>,  k_tl_resolve<false, true, true>,  k_tl_resolve<true, true, true>};
    const uint32_t tilesX = (pc.W + 31u) / 32u, tilesY = (pc.H + 7u) / 8u;
    PassConst q = pc; q.tilesX = tilesX; q.tilesY = tilesY;
    launch_k(kern[(tex ? 1 : 0) | (pbr ? 2 : 0) | (affine ? 4 : 0)], dim3(tilesX * tilesY), dim3(256), st, true, q, sh, head, nodes, cap, idBase, hdr, depth, counters);
}

void launch_tl_resolve_ms(const PassConst& pc, const ShadeConst& sh, uint32_t* head, const uint4* nodes, uint32_t cap, uint32_t idBase, const float4* msColour,
                          float* msDepth, uint32_t* msTriId, cudaStream_t st) {
    const bool tex = sh.nTextures && pc.vtxUV, pbr = sh.shadingModel == 1u;
    const bool affine = sh.STR[3] == 0.0f && sh.STR[7] == 0.0f && sh.STR[11] == 0.0f && sh.STR[15] == 1.0f;
    using Kern = void (*)(PassConst, ShadeConst, uint32_t*, const uint4*, uint32_t, uint32_t, const float4*, float*, uint32_t*);
    static const Kern kern[8] = {k_tl_resolve_ms<false, false, false>, k_tl_resolve_ms<true, false, false>, k_tl_resolve_ms<false, true, false>, k_tl_resolve_ms<true, true, false>,
                                 k_tl_resolve_ms<false, false, true>,  k_tl_resolve_ms<true, false, true>,  k_tl_resolve_ms<false, true, true>,  k_tl_resolve_ms<true, true, true>};
    const uint32_t tilesX = (pc.W + 31u) / 32u, tilesY = (pc.H + 7u) / 8u;
    PassConst q = pc; q.tilesX = tilesX; q.tilesY = tilesY;
    launch_k(kern[(tex ? 1 : 0) | (pbr ? 2 : 0) | (affine ? 4 : 0)], dim3(tilesX * tilesY), dim3(256), st, true, q, sh, head, nodes, cap, idBase, msColour, msDepth, msTriId);
}

void launch_tl_resolve_split(const PassConst& pc, const ShadeConst& sh, uint32_t* head, const uint4* nodes, uint32_t cap, uint32_t idBase, float* hdr, float* depth,
                             uint32_t* counters, uint2* span, uint2* frags, float4* cols, int numSMs, cudaStream_t st) {
    const bool tex = sh.nTextures && pc.vtxUV, pbr = sh.shadingModel == 1u;
    const bool affine = sh.STR[3] == 0.0f && sh.STR[7] == 0.0f && sh.STR[11] == 0.0f && sh.STR[15] == 1.0f;
    const uint32_t tilesX = (pc.W + 31u) / 32u, tilesY = (pc.H + 7u) / 8u;
    PassConst q = pc; q.tilesX = tilesX; q.tilesY = tilesY;
    launch_k(k_tl_gather, dim3(tilesX * tilesY), dim3(256), st, true, q, head, nodes, cap, idBase, depth, sh.outTriId, sh.outDepth, span, frags, counters);
    using Kern = void (*)(PassConst, ShadeConst, const uint2*, float4*, uint32_t, const uint32_t*);
    static const Kern kern[8] = {k_tl_shade<false, false, false>, k_tl_shade<true, false, false>, k_tl_shade<false, true, false>, k_tl_shade<true, true, false>,
                                 k_tl_shade<false, false, true>,  k_tl_shade<true, false, true>,  k_tl_shade<false, true, true>,  k_tl_shade<true, true, true>};
    launch_k(kern[(tex ? 1 : 0) | (pbr ? 2 : 0) | (affine ? 4 : 0)], dim3((unsigned)numSMs * 6u), dim3(256), st, true, q, sh, (const uint2*)frags, cols, cap, (const uint32_t*)counters);
    launch_k(k_tl_blend, dim3(tilesX * tilesY), dim3(256), st, true, q, sh, span, (const float4*)cols, cap, hdr, (const float*)depth);
}

void launch_tl_resolve_ms_split(const PassConst& pc, const ShadeConst& sh, uint32_t* head, const uint4* nodes, uint32_t cap, uint32_t idBase, const float4* msColour,
                                float* msDepth, uint32_t* msTriId, uint32_t* counters, uint2* span, uint2* frags, float4* cols, int numSMs, cudaStream_t st) {
    const bool tex = sh.nTextures && pc.vtxUV, pbr = sh.shadingModel == 1u;
    const bool affine = sh.STR[3] == 0.0f && sh.STR[7] == 0.0f && sh.STR[11] == 0.0f && sh.STR[15] == 1.0f;
    const uint32_t tilesX = (pc.W + 31u) / 32u, tilesY = (pc.H + 7u) / 8u;
    PassConst q = pc; q.tilesX = tilesX; q.tilesY = tilesY;
    launch_k(k_tl_gather_ms, dim3(tilesX * tilesY), dim3(256), st, true, q, head, nodes, cap, idBase, msDepth, msTriId, span, frags, counters);
    using Kern = void (*)(PassConst, ShadeConst, const uint2*, float4*, uint32_t, const uint32_t*);
    static const Kern kern[8] = {k_tl_shade<false, false, false>, k_tl_shade<true, false, false>, k_tl_shade<false, true, false>, k_tl_shade<true, true, false>,
                                 k_tl_shade<false, false, true>,  k_tl_shade<true, false, true>,  k_tl_shade<false, true, true>,  k_tl_shade<true, true, true>};
    launch_k(kern[(tex ? 1 : 0) | (pbr ? 2 : 0) | (affine ? 4 : 0)], dim3((unsigned)numSMs * 6u), dim3(256), st, true, q, sh, (const uint2*)frags, cols, cap, (const uint32_t*)counters);
    launch_k(k_tl_blend_ms, dim3(tilesX * tilesY), dim3(256), st, true, q, sh, span, (const uint2*)frags, (const float4*)cols, cap, msColour, (const float*)msDepth, (const uint32_t*)msTriId);
}

}  // namespace reb
