// k_translucent.cu — the translucent meshes of the base pass (Appendix B.16 of DESIGN.md).
//
// The reference draws its AlphaMode::eBlend meshes after the opaque ones in the same render pass, sorted back to front
// per mesh by the caller (RenderScene.cpp:127-159,226-227), with cull none, depth test LessOrEqual + depth write and
// src-alpha blending (GraphicsPipelineCache.cpp:305-315,330-340): per pixel the result depends on the ORDER of the
// fragments, which the visibility-buffer resolve of the opaque pass cannot express.  Two kernels:
//   k_tl_raster   one CTA per translucent triangle (cull none: vertices 1 and 2 swap when A2 > 0): every covered pixel whose
//                 depth passes against the OPAQUE depth appends {id, z} to that pixel's linked list (any order);
//   k_tl_resolve  per pixel: pull the list, sort by submission-order ID, replay it — LessOrEqual against the running
//                 depth, shade at the pixel centre with the ordinary shade_pixel* code, C = Cs*As + Cd*(1-As),
//                 A = As*(1-As) in linear float on the colour the opaque pass left, depth / ID write — then the skybox
//                 (drawn after the translucent list, RenderScene.cpp:228) where the depth is still 1.0, then the encode.
// Same individually rounded operations as oracle/oracle.cpp.  Limits of this first version: a pixel keeps at most
// TL_MAX_FRAGS fragments and the frame at most `cap` list nodes (the excess is dropped and counted).
#include "reb_shade.cuh"
#include "reb_shade_pbr.cuh"
#include "reb_sky.cuh"

namespace reb {

constexpr uint32_t TL_NONE = 0xFFFFFFFFu;
constexpr int TL_MAX_FRAGS = 16;

// One snapped triangle (a parent or a clip child), rasterised by the whole CTA: warps over rows, lanes over columns.
__device__ __forceinline__ void tl_raster_tri(const PassConst& pc, VtxA a[3], uint32_t tri, const float* __restrict__ opaqueDepth, uint32_t* __restrict__ head,
                                              uint4* __restrict__ nodes, uint32_t cap, uint32_t* __restrict__ counters, unsigned lane, unsigned warp) {
    long long A2 = mulw(a[1].X - a[0].X, a[2].Y - a[0].Y) - mulw(a[2].X - a[0].X, a[1].Y - a[0].Y);
    if (A2 == 0) return;
    if (A2 > 0) { const VtxA t = a[1]; a[1] = a[2]; a[2] = t; A2 = -A2; }                          // cull none (Appendix B.16)
    const int32_t minX = min(a[0].X, min(a[1].X, a[2].X)), maxX = max(a[0].X, max(a[1].X, a[2].X));
    const int32_t minY = min(a[0].Y, min(a[1].Y, a[2].Y)), maxY = max(a[0].Y, max(a[1].Y, a[2].Y));
    const int x0 = max((minX + 127) >> 8, 0), x1 = min((maxX - 128) >> 8, (int32_t)pc.W - 1);
    const int y0 = max((minY + 127) >> 8, 0), y1 = min((maxY - 128) >> 8, (int32_t)pc.H - 1);
    if (x0 > x1 || y0 > y1) return;
    int32_t A[3], B[3], bias[3]; long long E[3];   // unbiased edge value at the centre of pixel (0, 0)
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const int p = (k + 1) % 3, q = (k + 2) % 3;
        A[k] = a[q].Y - a[p].Y; B[k] = -(a[q].X - a[p].X);
        bias[k] = ((A[k] > 0) || (A[k] == 0 && B[k] > 0)) ? 0 : 1;   // top-left rule (Appendix B.7)
        E[k] = mulw(A[k], 128 - a[p].X) + mulw(B[k], 128 - a[p].Y);
    }
    const float inv = 1.0f / (float)(-A2);
    const float z0 = a[0].z, dz1 = a[1].z - a[0].z, dz2 = a[2].z - a[0].z;
    for (int y = y0 + (int)warp; y <= y1; y += 8) {
        if (!row_owned((uint32_t)y, pc.rank, pc.nRanks, pc.bandH)) continue;
        for (int x = x0 + (int)lane; x <= x1; x += 32) {
            const long long F0 = E[0] + mulw(A[0], 256 * x) + mulw(B[0], 256 * y);
            const long long F1 = E[1] + mulw(A[1], 256 * x) + mulw(B[1], 256 * y);
            const long long F2 = E[2] + mulw(A[2], 256 * x) + mulw(B[2], 256 * y);
            if (((F0 - bias[0]) | (F1 - bias[1]) | (F2 - bias[2])) < 0) continue;
            const float z = depth_from((float)F1, (float)F2, inv, z0, dz1, dz2);
            if (!(z >= 0.0f && z <= 1.0f)) continue;                                              // Appendix B.8 depth clip
            const size_t p = (size_t)y * pc.W + x;
            if (!(z <= __ldg(opaqueDepth + p))) continue;     // the running depth only ever gets nearer than the opaque depth
            const uint32_t idx = atomicAdd(&counters[0], 1u);
            if (idx >= cap) { atomicAdd(&counters[1], 1u); continue; }
            nodes[idx] = make_uint4(atomicExch(&head[p], idx), tri, __float_as_uint(z), 0u);
        }
    }
}

// Appendix B.3 for a translucent parent (CTA-uniform: every thread clips the same triangle): the children of the fan go
// through the same snap / facing / raster steps and append the PARENT's ID (children of one fan share edges, which the
// top-left rule assigns to exactly one of them).
__device__ __noinline__ void tl_clip_raster(const PassConst& pc, uint32_t tri, const float* __restrict__ opaqueDepth, uint32_t* __restrict__ head,
                                            uint4* __restrict__ nodes, uint32_t cap, uint32_t* __restrict__ counters, unsigned lane, unsigned warp) {
    float4 c[9];
    const int n = clip_parent<false>(pc, tri, c);
    for (int i = 1; i + 1 < n; ++i) {
        VtxA a[3] = {viewport_snap(c[0], pc.W, pc.H), viewport_snap(c[i], pc.W, pc.H), viewport_snap(c[i + 1], pc.W, pc.H)};
        if (((a[0].flags | a[1].flags | a[2].flags) & VF_NEEDS_CLIP) != 0u) {   // cannot happen for finite inputs: counted, dropped
            if (threadIdx.x == 0) atomicAdd(&counters[2], 1u);
            continue;
        }
        tl_raster_tri(pc, a, tri, opaqueDepth, head, nodes, cap, counters, lane, warp);
    }
}

__global__ void __launch_bounds__(256) k_tl_raster(const __grid_constant__ PassConst pc, const float* __restrict__ opaqueDepth, uint32_t* __restrict__ head,
                                                   uint4* __restrict__ nodes, uint32_t cap, uint32_t* __restrict__ counters) {
    pdl_enter();
    const unsigned lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    for (uint32_t tri = blockIdx.x; tri < pc.nTris; tri += gridDim.x) {
        uint32_t slot[3];
#pragma unroll
        for (int k = 0; k < 3; ++k) slot[k] = pc.deindexed ? 3u * tri + (uint32_t)k : __ldg(pc.flatSlots + 3ull * tri + k);
        VtxA a[3];
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const uint4 q = __ldg(reinterpret_cast<const uint4*>(pc.vtxA) + slot[k]);
            a[k] = {(int32_t)q.x, (int32_t)q.y, __uint_as_float(q.z), q.w};
        }
        if ((a[0].flags & a[1].flags & a[2].flags & VF_PLANES) != 0u) continue;                       // Appendix B.2
        if (((a[0].flags | a[1].flags | a[2].flags) & VF_NEEDS_CLIP) != 0u) tl_clip_raster(pc, tri, opaqueDepth, head, nodes, cap, counters, lane, warp);
        else tl_raster_tri(pc, a, tri, opaqueDepth, head, nodes, cap, counters, lane, warp);
    }
}

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
        while (h != TL_NONE && h < cap) {
            const uint4 nd = nodes[h];
            if (n < TL_MAX_FRAGS) {              // insertion by submission-order ID
                int i = n++;
                while (i > 0 && id[i - 1] > nd.y) { id[i] = id[i - 1]; z[i] = z[i - 1]; --i; }
                id[i] = nd.y; z[i] = __uint_as_float(nd.z);
            } else {
                atomicAdd(&counters[1], 1u);
            }
            h = nd.x;
        }
        for (int i = 0; i < n; ++i) {
            if (!(z[i] <= d)) continue;          // LessOrEqual against the running depth (depth write on)
            const float4 src = PBR ? shade_pixel_pbr<TEXTURED, AFFINE>(pc, sh, x, y, id[i]) : shade_pixel<TEXTURED, AFFINE>(pc, sh, x, y, id[i]);
            const float a = src.w, ia = 1.0f - a;
            col.x = src.x * a + col.x * ia; col.y = src.y * a + col.y * ia; col.z = src.z * a + col.z * ia;
            col.w = a * ia;
            d = z[i]; lastId = id[i];
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

void launch_tl_raster(const PassConst& pc, const float* opaqueDepth, uint32_t* head, uint4* nodes, uint32_t cap, uint32_t* counters, int numSMs, cudaStream_t st) {
    if (pc.nTris == 0) return;
    const uint32_t grid = std::min<uint32_t>(pc.nTris, (uint32_t)numSMs * 8u);
    launch_k(k_tl_raster, dim3(grid), dim3(256), st, true, pc, opaqueDepth, head, nodes, cap, counters);
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

}  // namespace reb
