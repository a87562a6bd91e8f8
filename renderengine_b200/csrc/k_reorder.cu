// k_reorder.cu — spatially coherent order for the triangles and cache slots of a draw list, built ONCE when the list changes.
//
// The reference draws meshes in whatever index order the importer produced (AssimpSceneLoader.cpp:198-261); a GPU's
// post-transform cache and rasteriser only see locality if that order has some.  Here the raster front end walks triangles in
// its own order, so the order is a free choice as long as results do not depend on it: K1 carries the SUBMISSION-order ID of
// every triangle through the pipeline (Appendix B.9: later primitive wins at equal depth — the visibility key holds ~id), and
// all resolves are order-independent minima.  So, per draw (a triangle never leaves its draw: draw -> node / material lookups by
// position stay valid):
//   1. triangles are sorted along a 30-bit Morton curve of their object-space centroid inside the draw's bounds;
//   2. cache slots are renumbered in first-use order of that triangle stream (a slot no triangle uses goes last in its draw),
//      and the source vertex streams are copied into slot order: K0 reads and writes linearly, K1's three gathers per triangle
//      hit lines its neighbours in the batch just touched, and a 256-triangle block is a compact patch (what the per-block
//      culling of the sort-first split needs);
//   3. K1 walks a stream of 16-B records {slot0, slot1, slot2, submission ID}; flatSlots (indexed by submission ID, read by
//      shading and the translucent raster) is rewritten to the new slot numbers.
// Indexed cache mode only (the de-indexed fallback has no shared slots to make coherent).
#include "reb_device.cuh"
#include "reb_kernels.h"

#include <cub/device/device_radix_sort.cuh>
#include <algorithm>

namespace reb {
namespace {

__device__ __forceinline__ uint32_t part1by2(uint32_t v) {   // 10 bits -> every third bit
    v &= 0x3FFu;
    v = (v | (v << 16)) & 0x030000FFu;
    v = (v | (v << 8)) & 0x0300F00Fu;
    v = (v | (v << 4)) & 0x030C30C3u;
    v = (v | (v << 2)) & 0x09249249u;
    return v;
}
__device__ __forceinline__ float dec_ord(uint32_t e) { return __uint_as_float((e & 0x80000000u) ? (e & 0x7FFFFFFFu) : ~e); }

// key = draw << 32 | Morton(centroid within the draw's object-space bounds); value = submission-order triangle
__global__ void __launch_bounds__(256) k_tri_keys(const __grid_constant__ PassConst pc, const uint32_t* __restrict__ drawBounds, unsigned long long* __restrict__ keys,
                                                  uint32_t* __restrict__ vals) {
    const uint32_t tri = blockIdx.x * 256u + threadIdx.x;
    if (tri >= pc.nTris) return;
    const uint32_t d = find_draw_block<false>(pc.draws, pc.nDraws, pc.triBlockDraw, tri);
    const uint32_t* ix = pc.indices + pc.draws[d].indexOffset + 3u * (tri - pc.draws[d].firstTri);
    float c[3] = {0.0f, 0.0f, 0.0f};
#pragma unroll
    for (int k = 0; k < 3; ++k) { const float4 p = __ldg(pc.pos4 + __ldg(ix + k)); c[0] += p.x; c[1] += p.y; c[2] += p.z; }
    uint32_t q[3];
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const float lo = dec_ord(drawBounds[6 * (size_t)d + a]), hi = dec_ord(drawBounds[6 * (size_t)d + 3 + a]);
        const float ext = hi - lo;
        const float t = ext > 0.0f ? (c[a] * (1.0f / 3.0f) - lo) / ext : 0.0f;
        q[a] = (uint32_t)fminf(fmaxf(t * 1024.0f, 0.0f), 1023.0f);   // NaN -> 0
    }
    keys[tri] = ((unsigned long long)d << 32) | (part1by2(q[0]) | (part1by2(q[1]) << 1) | (part1by2(q[2]) << 2));
    vals[tri] = tri;
}

__global__ void __launch_bounds__(256) k_fill_ff(uint32_t* __restrict__ p, size_t n) {
    for (size_t i = (size_t)blockIdx.x * 256u + threadIdx.x; i < n; i += (size_t)gridDim.x * 256u) p[i] = 0xFFFFFFFFu;
}

// first position in the sorted triangle stream that references each slot
__global__ void __launch_bounds__(256) k_first_use(uint32_t nTris, const uint32_t* __restrict__ sortedTri, const uint32_t* __restrict__ flat, uint32_t* __restrict__ firstUse) {
    const uint32_t pos = blockIdx.x * 256u + threadIdx.x;
    if (pos >= nTris) return;
    REB_ASSERT(sortedTri[pos] < nTris);
    const uint32_t* f = flat + 3ull * sortedTri[pos];
    atomicMin(firstUse + f[0], pos); atomicMin(firstUse + f[1], pos); atomicMin(firstUse + f[2], pos);
}

// key = draw << 32 | first use; value = old slot
__global__ void __launch_bounds__(256) k_slot_keys(const __grid_constant__ PassConst pc, const uint32_t* __restrict__ firstUse, unsigned long long* __restrict__ keys,
                                                   uint32_t* __restrict__ vals) {
    const uint32_t slot = blockIdx.x * 256u + threadIdx.x;
    if (slot >= pc.cacheSize) return;
    const uint32_t d = find_draw_block<true>(pc.draws, pc.nDraws, pc.slotBlockDraw, slot);
    keys[slot] = ((unsigned long long)d << 32) | firstUse[slot];
    vals[slot] = slot;
}

// new slot -> source vertex streams in slot order, old slot -> new slot
__global__ void __launch_bounds__(256) k_slot_streams(const __grid_constant__ PassConst pc, const uint32_t* __restrict__ newToOld, uint32_t* __restrict__ remap,
                                                      float4* __restrict__ posS, float4* __restrict__ nrmS) {
    const uint32_t slot = blockIdx.x * 256u + threadIdx.x;
    if (slot >= pc.cacheSize) return;
    const uint32_t old = newToOld[slot];
    REB_ASSERT(old < pc.cacheSize);
    remap[old] = slot;
    const uint32_t d = find_draw_block<true>(pc.draws, pc.nDraws, pc.slotBlockDraw, old);
    const uint32_t v = (uint32_t)((int32_t)old - pc.draws[d].cacheOffset);
    posS[slot] = __ldg(pc.pos4 + v);
    nrmS[slot] = __ldg(pc.nrm4 + v);
}

__global__ void __launch_bounds__(256) k_tri_streams(uint32_t nTris, const uint32_t* __restrict__ sortedTri, const uint32_t* __restrict__ flat,
                                                     const uint32_t* __restrict__ remap, uint4* __restrict__ triStream) {
    const uint32_t pos = blockIdx.x * 256u + threadIdx.x;
    if (pos >= nTris) return;
    const uint32_t tri = sortedTri[pos];
    const uint32_t* f = flat + 3ull * tri;
    triStream[pos] = make_uint4(remap[f[0]], remap[f[1]], remap[f[2]], tri);
}

__global__ void __launch_bounds__(256) k_remap_flat(size_t n, uint32_t* __restrict__ flat, const uint32_t* __restrict__ remap) {
    const size_t i = (size_t)blockIdx.x * 256u + threadIdx.x;
    if (i < n) flat[i] = remap[flat[i]];
}

template <class T> cudaError_t tmp_alloc(T** p, size_t n) { return cudaMalloc(reinterpret_cast<void**>(p), n * sizeof(T)); }

}  // namespace

// Runs on `st` and synchronises it (draw-list preprocessing is not per-frame work).  pc: the pass with draws, lookup tables,
// flatSlots (= flat) filled; posS / nrmS: cacheSize records; triStream: nTris records.  Returns a CUDA error code.
cudaError_t reorder_draw_list(const PassConst& pc, const uint32_t* drawBounds, float4* posS, float4* nrmS, uint4* triStream, uint32_t* flat, cudaStream_t st) {
    const uint32_t nT = pc.nTris, nS = pc.cacheSize;
    if (nT == 0 || nS == 0) return cudaSuccess;
    const size_t nMax = std::max<size_t>(nT, nS);
    unsigned long long *keysA = nullptr, *keysB = nullptr;
    uint32_t *valsA = nullptr, *valsB = nullptr, *sortedTri = nullptr, *firstUse = nullptr, *remap = nullptr;
    void* tmp = nullptr;
    size_t tmpBytes = 0;
    cudaError_t e = cudaSuccess;
    auto done = [&](cudaError_t r) {
        cudaStreamSynchronize(st);
        cudaFree(keysA); cudaFree(keysB); cudaFree(valsA); cudaFree(valsB); cudaFree(sortedTri); cudaFree(firstUse); cudaFree(remap); cudaFree(tmp);
        return r;
    };
    if ((e = tmp_alloc(&keysA, nMax)) || (e = tmp_alloc(&keysB, nMax)) || (e = tmp_alloc(&valsA, nMax)) || (e = tmp_alloc(&valsB, nMax)) ||
        (e = tmp_alloc(&sortedTri, (size_t)nT)) || (e = tmp_alloc(&firstUse, (size_t)nS)) || (e = tmp_alloc(&remap, (size_t)nS))) return done(e);
    cub::DoubleBuffer<unsigned long long> dk(keysA, keysB);
    cub::DoubleBuffer<uint32_t> dv(valsA, valsB);
    if ((e = cub::DeviceRadixSort::SortPairs(nullptr, tmpBytes, dk, dv, (int)nMax, 0, 64, st))) return done(e);
    if ((e = cudaMalloc(&tmp, tmpBytes))) return done(e);
    // 1. triangles
    k_tri_keys<<<(nT + 255u) / 256u, 256, 0, st>>>(pc, drawBounds, dk.Current(), dv.Current());
    if ((e = cub::DeviceRadixSort::SortPairs(tmp, tmpBytes, dk, dv, (int)nT, 0, 64, st))) return done(e);
    if ((e = cudaMemcpyAsync(sortedTri, dv.Current(), (size_t)nT * 4, cudaMemcpyDeviceToDevice, st))) return done(e);
    // 2. slots in first-use order
    k_fill_ff<<<1024, 256, 0, st>>>(firstUse, nS);
    k_first_use<<<(nT + 255u) / 256u, 256, 0, st>>>(nT, sortedTri, flat, firstUse);
    k_slot_keys<<<(nS + 255u) / 256u, 256, 0, st>>>(pc, firstUse, dk.Current(), dv.Current());
    if ((e = cub::DeviceRadixSort::SortPairs(tmp, tmpBytes, dk, dv, (int)nS, 0, 64, st))) return done(e);
    k_slot_streams<<<(nS + 255u) / 256u, 256, 0, st>>>(pc, dv.Current(), remap, posS, nrmS);
    // 3. the K1 stream and the slot table shading reads
    k_tri_streams<<<(nT + 255u) / 256u, 256, 0, st>>>(nT, sortedTri, flat, remap, triStream);
    k_remap_flat<<<(unsigned)((3ull * nT + 255u) / 256u), 256, 0, st>>>(3ull * nT, flat, remap);
    return done(cudaGetLastError());
}

}  // namespace reb
