// reb_noalloc.cu — reb_shadow_depth_static: the directional shadow depth pass of ONE draw for callers that must not allocate
// (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  Same raster contract as the context
// path (shadow_map.vert:48-52 vertex stage, SURVEY.md Appendix B.2-B.8: clip flags, viewport snap to 24.8, CCW front / back
// culled, top-left rule, depth from the unbiased edge values, LessOrEqual resolve as a min over depth bits), bit-exact with
// reb_render_shadow on the same input.  No counterpart in the reference.
//
// The scratch the context path takes from cudaMalloc — the post-transform vertex cache (K0 -> K1) and the queue of triangles
// too large for one thread (K1 -> K2) — lives in ONE file-scope __device__ array of fixed size (NA_SCRATCH_BYTES, below): every
// GPU that loads libreb.so holds it until the process exits.  g_naScratch names a host-side stand-in in host code; the device
// address is taken once per device with cudaGetSymbolAddress (na_setup) and passed to the kernels.
//
// All calls on one device share the array: calls that overlap on the GPU (two streams, two threads) race.  It keeps the last
// call's contents; the only word that must start at zero — the queue length — is cleared by the first kernel of every call.
#include "../../include/reb.h"
#include "reb_device.cuh"

#include <mutex>

namespace reb {
namespace {

constexpr uint32_t NA_MAX_VERTS = REB_STATIC_MAX_VERTICES;    // 16 B of cache each: 4 MiB
constexpr uint32_t NA_MAX_TRIS = REB_STATIC_MAX_TRIANGLES;    // 4 B of queue each:  2 MiB
constexpr uint32_t NA_HEADER_U4 = 16;                         // 256 B: word 0 = queue length
constexpr size_t NA_SCRATCH_U4 = (size_t)NA_HEADER_U4 + NA_MAX_VERTS + NA_MAX_TRIS / 4;
constexpr size_t NA_SCRATCH_BYTES = NA_SCRATCH_U4 * 16;       // 6 MiB + 256 B
constexpr uint32_t NA_CLIP_BIT = 0x80000000u;                 // queue entry: the triangle needs Appendix B.3 clipping
constexpr int NA_INLINE_PX = 64;                              // pixel centres a single thread rasterises itself
constexpr int NA_MAX_DEVICES = 64;
static_assert(NA_SCRATCH_BYTES == 6u * 1024u * 1024u + 256u, "the size include/reb.h documents");
static_assert(NA_MAX_TRIS < NA_CLIP_BIT && NA_MAX_TRIS % 4 == 0, "queue entries carry a flag in bit 31");

static __device__ __align__(256) uint4 g_naScratch[NA_SCRATCH_U4];

struct NaConst {
    float PV[16];                 // shadowProj * shadowView, column-major
    float T[16];                  // model transform of the draw
    const uint8_t* verts; uint32_t stride, nVerts;   // position = the first three floats of every record
    const uint32_t* indices; uint32_t nTris;
    uint32_t S;                   // shadow map extent
    uint32_t* depthBits;          // S * S, the caller's output
    uint32_t* header; uint4* cache; uint32_t* queue;   // carved from the static scratch
};

__device__ __forceinline__ float4 na_clip_vertex(const NaConst& c, uint32_t v) {
    const float* p = reinterpret_cast<const float*>(c.verts + (size_t)v * c.stride);
    return vertex_stage<false>(c.PV, c.T, __ldg(p), __ldg(p + 1), __ldg(p + 2), nullptr);
}

// Rasterise one set-up triangle: samples first, first + step, ... of its pixel-centre bbox (row-major).  64-bit edge functions
// (the same integers K1's 32-bit row stage and the tile stage's 64-bit one produce), Appendix B.7 bias, B.8 depth.
__device__ __forceinline__ void na_raster(const NaConst& c, const VtxA a[3], uint32_t first, uint32_t step) {
    const long long A2 = mulw(a[1].X - a[0].X, a[2].Y - a[0].Y) - mulw(a[2].X - a[0].X, a[1].Y - a[0].Y);
    if (A2 >= 0) return;                                                                     // Appendix B.6
    const int32_t X[3] = {a[0].X, a[1].X, a[2].X}, Y[3] = {a[0].Y, a[1].Y, a[2].Y};
    const int32_t minX = min(X[0], min(X[1], X[2])), maxX = max(X[0], max(X[1], X[2]));
    const int32_t minY = min(Y[0], min(Y[1], Y[2])), maxY = max(Y[0], max(Y[1], Y[2]));
    const int32_t x0 = max((minX + 127) >> 8, 0), x1 = min((maxX - 128) >> 8, (int32_t)c.S - 1);
    const int32_t y0 = max((minY + 127) >> 8, 0), y1 = min((maxY - 128) >> 8, (int32_t)c.S - 1);
    if (x0 > x1 || y0 > y1) return;
    long long E[3]; int32_t A[3], B[3], bias[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const int p = (k + 1) % 3, q = (k + 2) % 3;                                          // edge p -> q opposes vertex k
        A[k] = Y[q] - Y[p]; B[k] = -(X[q] - X[p]);
        bias[k] = ((A[k] > 0) || (A[k] == 0 && B[k] > 0)) ? 0 : 1;
        E[k] = (long long)A[k] * (256ll * x0 + 128 - X[p]) + (long long)B[k] * (256ll * y0 + 128 - Y[p]);   // unbiased, at the bbox origin
    }
    const float inv = 1.0f / (float)(-A2);
    const float z0 = a[0].z, dz1 = a[1].z - a[0].z, dz2 = a[2].z - a[0].z;
    const uint32_t w = (uint32_t)(x1 - x0 + 1);
    const unsigned long long n = (unsigned long long)w * (uint32_t)(y1 - y0 + 1);
    for (unsigned long long i = first; i < n; i += step) {
        const long long dx = 256ll * (long long)(i % w), dy = 256ll * (long long)(i / w);
        const long long F0 = E[0] + A[0] * dx + B[0] * dy, F1 = E[1] + A[1] * dx + B[1] * dy, F2 = E[2] + A[2] * dx + B[2] * dy;
        if (((F0 - bias[0]) | (F1 - bias[1]) | (F2 - bias[2])) < 0) continue;
        const float z = depth_from((float)F1, (float)F2, inv, z0, dz1, dz2);
        if (!(z >= 0.0f && z <= 1.0f)) continue;
        atomicMin(c.depthBits + ((size_t)(y0 + (int32_t)(i / w)) * c.S + (size_t)(x0 + (int32_t)(i % w))), __float_as_uint(z));
    }
}

// Appendix B.3 for one parent (the steps of clip_parent / clip_emit, reb_device.cuh / reb_bin.cuh, on this entry point's inputs):
// vertex stage re-run, Sutherland-Hodgman against near z >= 0 and the four guard planes, fan from the first vertex, every child
// through viewport / snap / facing / bbox.
__device__ __noinline__ void na_clip_raster(const NaConst& c, uint32_t tri, uint32_t first, uint32_t step) {
    float4 a[9], b[9];
#pragma unroll
    for (int k = 0; k < 3; ++k) a[k] = na_clip_vertex(c, __ldg(c.indices + 3ull * tri + k));
    const float Gx = (0.5f * GUARD_PX) / (0.5f * (float)c.S), Gy = Gx;
    int n = 3;
    for (int plane = 0; plane < 5; ++plane) {
        int m = 0;
        for (int i = 0; i < n; ++i) {
            const float4 P = a[i], Q = a[i + 1 == n ? 0 : i + 1];
            const float dP = clip_dist(plane, P, Gx, Gy), dQ = clip_dist(plane, Q, Gx, Gy);
            const bool inP = dP >= 0.0f, inQ = dQ >= 0.0f;
            if (inP) b[m++] = P;
            if (inP != inQ) {
                const float4 I = inP ? P : Q, O = inP ? Q : P;
                const float dI = inP ? dP : dQ, dO = inP ? dQ : dP;
                const float t = dI / (dI - dO);
                b[m++] = make_float4(I.x + t * (O.x - I.x), I.y + t * (O.y - I.y), I.z + t * (O.z - I.z), I.w + t * (O.w - I.w));
            }
        }
        n = m;
        if (n < 3) return;
        for (int k = 0; k < n; ++k) a[k] = b[k];
    }
    VtxA v[3];
    v[0] = viewport_snap(a[0], c.S, c.S);
    for (int i = 1; i + 1 < n; ++i) {
        v[1] = viewport_snap(a[i], c.S, c.S); v[2] = viewport_snap(a[i + 1], c.S, c.S);
        if (((v[0].flags | v[1].flags | v[2].flags) & VF_NEEDS_CLIP) != 0u) continue;   // non-finite input: dropped, as the context path does
        na_raster(c, v, first, step);
    }
}

// 1. clear the caller's map to depth 1.0 (ShadowSystem.cpp:111-127), reset the queue length, run the vertex stage into the cache
__global__ void __launch_bounds__(256) k_na_vertex(const __grid_constant__ NaConst c) {
    const size_t tid = (size_t)blockIdx.x * 256u + threadIdx.x, nThreads = (size_t)gridDim.x * 256u;
    if (tid == 0) c.header[0] = 0u;
    for (size_t i = tid; i < (size_t)c.S * c.S; i += nThreads) c.depthBits[i] = 0x3F800000u;
    for (size_t v = tid; v < c.nVerts; v += nThreads) {
        const VtxA a = viewport_snap(na_clip_vertex(c, (uint32_t)v), c.S, c.S);
        c.cache[v] = make_uint4((uint32_t)a.X, (uint32_t)a.Y, __float_as_uint(a.z), a.flags);
    }
}

__device__ __forceinline__ bool na_gather(const NaConst& c, uint32_t tri, VtxA a[3]) {
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const uint32_t v = __ldg(c.indices + 3ull * tri + k);
        if (v >= c.nVerts) return false;                          // an index past the vertex buffer: the triangle is skipped
        const uint4 q = c.cache[v];
        a[k] = {(int32_t)q.x, (int32_t)q.y, __uint_as_float(q.z), q.w};
    }
    return true;
}

// 2. one thread per triangle: reject / cull, rasterise bboxes of <= NA_INLINE_PX centres in place, queue the rest
__global__ void __launch_bounds__(256) k_na_setup(const __grid_constant__ NaConst c) {
    const uint32_t tri = blockIdx.x * 256u + threadIdx.x;
    if (tri >= c.nTris) return;
    VtxA a[3];
    if (!na_gather(c, tri, a)) return;
    const uint32_t f0 = a[0].flags, f1 = a[1].flags, f2 = a[2].flags;
    if ((f0 & f1 & f2 & VF_PLANES) != 0u) return;                                                       // Appendix B.2
    if (((f0 | f1 | f2) & VF_NEEDS_CLIP) != 0u) { c.queue[atomicAdd(c.header, 1u)] = tri | NA_CLIP_BIT; return; }
    const int32_t minX = min(a[0].X, min(a[1].X, a[2].X)), maxX = max(a[0].X, max(a[1].X, a[2].X));
    const int32_t minY = min(a[0].Y, min(a[1].Y, a[2].Y)), maxY = max(a[0].Y, max(a[1].Y, a[2].Y));
    const long long w = (long long)min((maxX - 128) >> 8, (int32_t)c.S - 1) - max((minX + 127) >> 8, 0) + 1;
    const long long h = (long long)min((maxY - 128) >> 8, (int32_t)c.S - 1) - max((minY + 127) >> 8, 0) + 1;
    if (w <= 0 || h <= 0) return;
    if (w * h <= NA_INLINE_PX) na_raster(c, a, 0u, 1u);
    else c.queue[atomicAdd(c.header, 1u)] = tri;                                                        // <= nTris <= NA_MAX_TRIS entries
}

// 3. one CTA per queued triangle (persistent grid): 256 threads stride over its samples
__global__ void __launch_bounds__(256) k_na_queued(const __grid_constant__ NaConst c) {
    const uint32_t n = c.header[0];
    for (uint32_t q = blockIdx.x; q < n; q += gridDim.x) {
        const uint32_t e = c.queue[q], tri = e & ~NA_CLIP_BIT;
        if (e & NA_CLIP_BIT) { na_clip_raster(c, tri, threadIdx.x, 256u); continue; }
        VtxA a[3];
        if (na_gather(c, tri, a)) na_raster(c, a, threadIdx.x, 256u);
    }
}

struct NaDevice { uint4* scratch; int numSMs; };
NaDevice g_naDev[NA_MAX_DEVICES] = {};
std::mutex g_naMutex;

// once per device, with that device current: the device address behind the host-side stand-in g_naScratch
cudaError_t na_setup(int dev, NaDevice* out) {
    std::lock_guard<std::mutex> lock(g_naMutex);
    if (!g_naDev[dev].scratch) {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_naScratch);
        if (e != cudaSuccess) return e;
        int sms = 0;
        e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        if (e != cudaSuccess) return e;
        g_naDev[dev].numSMs = sms > 0 ? sms : 148;
        g_naDev[dev].scratch = static_cast<uint4*>(p);
    }
    *out = g_naDev[dev];
    return cudaSuccess;
}

// glm mat4 * mat4 (type_mat4x4.inl:679-694), the order reb_shadow_set_transform uses
void na_mat4_mul(const float* a, const float* b, float* out) {
    for (int col = 0; col < 4; ++col)
        for (int r = 0; r < 4; ++r) {
            float t = a[r] * b[col * 4];
            t = t + a[4 + r] * b[col * 4 + 1];
            t = t + a[8 + r] * b[col * 4 + 2];
            t = t + a[12 + r] * b[col * 4 + 3];
            out[col * 4 + r] = t;
        }
}

int na_fail(const char** errOut, int code, const char* msg) {
    if (errOut) *errOut = msg;
    return code;
}

}  // namespace
}  // namespace reb

using namespace reb;

extern "C" int reb_shadow_depth_static(const void* dVertices, uint32_t vertexStride, uint32_t nVertices, const uint32_t* dIndices, uint32_t nIndices,
                                       const float* model16, const float* shadowView16, const float* shadowProj16, uint32_t size, float* dDepthOut,
                                       void* cudaStream, const char** errOut) {
    if (errOut) *errOut = "";
    if (!model16 || !shadowView16 || !shadowProj16 || !dDepthOut) return na_fail(errOut, REB_E_INVALID, "reb_shadow_depth_static: null matrix or output pointer");
    if (size == 0 || size > REB_MAX_EXTENT) return na_fail(errOut, REB_E_INVALID, "reb_shadow_depth_static: size must be in 1..16384");
    if (vertexStride < 12 || (vertexStride & 3u)) return na_fail(errOut, REB_E_INVALID, "reb_shadow_depth_static: vertexStride must be a multiple of 4 and >= 12");
    if (nIndices % 3u) return na_fail(errOut, REB_E_INVALID, "reb_shadow_depth_static: nIndices must be a multiple of 3");
    if ((nVertices && !dVertices) || (nIndices && !dIndices)) return na_fail(errOut, REB_E_INVALID, "reb_shadow_depth_static: null vertex / index pointer");
    if (nVertices > NA_MAX_VERTS) return na_fail(errOut, REB_E_OVERFLOW, "reb_shadow_depth_static: more than REB_STATIC_MAX_VERTICES (262144) vertices: the static scratch cannot hold their post-transform cache");
    if (nIndices / 3u > NA_MAX_TRIS) return na_fail(errOut, REB_E_OVERFLOW, "reb_shadow_depth_static: more than REB_STATIC_MAX_TRIANGLES (524288) triangles: the static scratch cannot hold their queue");
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= NA_MAX_DEVICES) return na_fail(errOut, REB_E_NO_DEVICE, "reb_shadow_depth_static: no current CUDA device (ordinal 0..63)");
    NaDevice nd;
    if (na_setup(dev, &nd) != cudaSuccess) { cudaGetLastError(); return na_fail(errOut, REB_E_CUDA, "reb_shadow_depth_static: cudaGetSymbolAddress of the static scratch failed"); }
    NaConst c;
    na_mat4_mul(shadowProj16, shadowView16, c.PV);   // ShadowSystem.cpp:303
    for (int i = 0; i < 16; ++i) c.T[i] = model16[i];
    c.verts = static_cast<const uint8_t*>(dVertices); c.stride = vertexStride; c.nVerts = nVertices;
    c.indices = dIndices; c.nTris = nIndices / 3u;
    c.S = size; c.depthBits = reinterpret_cast<uint32_t*>(dDepthOut);
    c.header = reinterpret_cast<uint32_t*>(nd.scratch);
    c.cache = nd.scratch + NA_HEADER_U4;
    c.queue = reinterpret_cast<uint32_t*>(nd.scratch + NA_HEADER_U4 + NA_MAX_VERTS);
    cudaStream_t st = static_cast<cudaStream_t>(cudaStream);
    const unsigned persistent = (unsigned)nd.numSMs * 8u;
    k_na_vertex<<<persistent, 256, 0, st>>>(c);
    if (c.nTris) {
        k_na_setup<<<(c.nTris + 255u) / 256u, 256, 0, st>>>(c);
        k_na_queued<<<persistent, 256, 0, st>>>(c);
    }
    if (cudaGetLastError() != cudaSuccess) return na_fail(errOut, REB_E_CUDA, "reb_shadow_depth_static: kernel launch failed");
    return REB_OK;
}
