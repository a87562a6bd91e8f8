// reb_device.cuh — device-side types and exact-arithmetic helpers shared by the raster kernels.
//
// Bit-exactness contract: every fp32 operation on the coverage / depth /
// shadow-lookup path is an individually rounded IEEE op in a fixed order.  These translation units
// are compiled with --fmad=false (no contraction), default -prec-div/-prec-sqrt (IEEE / and sqrtf).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

// Checked build (-DREB_CHECKED, `python -m renderengine_b200.build checked` -> variants/libreb_checked.so, loaded with REB_LIB=...):
// device-side asserts on every index this code computes into a buffer it does not own the bounds of by construction — the GPU
// suite run against that library is this repo's stand-in for compute-sanitizer memcheck, which is closed on this pool.
#ifdef REB_CHECKED
#include <cassert>
#define REB_ASSERT(c) assert(c)
#else
#define REB_ASSERT(c) ((void)0)
#endif

namespace reb {

constexpr uint32_t TRI_NONE = 0xFFFFFFFFu;
constexpr unsigned long long VIS_CLEAR = 0xFFFFFFFFFFFFFFFFull;   // loses against every real fragment key
constexpr float GUARD_PX = 1048576.0f;                            // |x_f|,|y_f| <= 2^20 px
constexpr int TILE_W = 32, TILE_H = 8;                            // one CTA (256 threads) per screen tile
constexpr int HUGE_TILES = 256;                                   // bbox covering more tiles than this: broadcast list
constexpr int HUGE_CAP = 64;                                      // capacity of the broadcast list (overflow -> binned path)
#ifndef SMALL_MAX_D
#define SMALL_MAX_D 14
#endif
#ifndef MED_MAX_D
#define MED_MAX_D 29
#endif
// K1 rasterises a triangle itself when its bbox spans <= SMALL_LIMIT px (everything larger is binned for the tile stage).  The unit
// of work of its row stage is one row of <= SMALL_MAX + 1 pixel centres: a bbox wider than that contributes two half rows per
// row.  History: 8 was a cliff (C2 with the reference's 4096^2 map has 6-10 px triangles: 630 -> 226 us / frame at 14); 29
// keeps 15-30 px triangles (ordinary game meshes) out of the binned path.
constexpr int SMALL_MAX = SMALL_MAX_D;
constexpr int SMALL_LIMIT = MED_MAX_D > SMALL_MAX_D ? MED_MAX_D : SMALL_MAX_D;
static_assert(SMALL_LIMIT + 1 <= 2 * (SMALL_MAX + 1) && SMALL_LIMIT + 1 <= 32, "two half rows per row; 5-bit row / column numbers");
constexpr uint32_t LIGHT_TYPE_DIRECTIONAL = 1, LIGHT_TYPE_SPOT = 3;   // phong.glsl:16-18

// Post-transform vertex flags (VtxA.flags): bits 0-5 = outside x<-w, x>w, y<-w, y>w, z<0, z>w (Appendix B.2);
// bit 6 = !(w > 0); bit 7 = outside the guard band (Appendix B.3).
constexpr uint32_t VF_PLANES = 0x3Fu, VF_NEEDS_CLIP = 0xC0u, VF_W = 0x40u, VF_GUARD = 0x80u;

// One vkCmdDrawIndexed of the reference's draw loops (MeshDrawInfo, MeshAllocator.h:35-39) with the
// running primitive counter (Appendix B.9) and its slice of the post-transform vertex cache
// precomputed on the host.  The array holds nDraws + 1 records; the last one is a sentinel
// (firstTri = nTris, cacheBase = cacheSize).
struct __align__(16) DrawDev {
    uint32_t indexOffset, firstTri, material, node;
    int32_t cacheOffset;          // cache slot of vertex v of this draw = v + cacheOffset (indexed mode)
    uint32_t cacheBase, nCache, pad;
};

// Post-transform vertex cache.  VtxA is all the setup kernel needs (one 16-B gather per corner);
// VtxBC (one 32-B sector) carries what shading interpolates.
struct __align__(16) VtxA { int32_t X, Y; float z; uint32_t flags; };      // snapped 24.8 position, ndc depth
struct __align__(16) VtxBC { float wx, wy, wz, invw, nx, ny, nz, pad; };    // world position, 1/w_clip, normalised normal

// Bindless 2-D texture (TextureCache.cpp:67-73,167-197): RGBA16 UNORM, full mip chain in one allocation.
struct __align__(16) TexDev {
    const uint16_t* base;         // level 0 first; level l starts at texel off[l]
    uint32_t w, h, mips, srgb;
    uint32_t off[16];
};

// A triangle that goes through the binned tile path: snapped vertices + depth + id (48 B, 3 x 16 B).
struct __align__(16) LargeRec {
    int32_t X0, Y0, X1, Y1;
    int32_t X2, Y2; float z0, z1;
    float z2; uint32_t triId; uint32_t pad0, pad1;
};

// Per-tile list header of the binned path (one 16-B word group per 32 x 8 tile, zeroed by K0's prologue):
//   count = records the list will hold (K1 / the clipper count while classifying); start = first record inside tileRecs (layout
//   phase); pos = scatter cursor.
struct __align__(16) TileHead { uint32_t count, start, pos, pad; };
constexpr int STATUS_WORDS = 16;   // per pass

struct PassConst {
    float PV[16];                 // proj*view (forward) or shadowProj*shadowView (shadow), column-major
    uint32_t W, H;                // target extent of this pass
    uint32_t nTris, nDraws;
    uint32_t cacheSize;           // number of post-transform cache slots
    uint32_t deindexed;           // 1: cache slot = 3*tri + corner (fallback when draws share vertices sparsely)
    const uint32_t* indices;
    const float4* pos4;           // (x, y, z, u)
    const float4* nrm4;           // (nx, ny, nz, v)
    const float* transforms;      // 16 floats per node
    const float* nmats;           // 12 floats per node: transpose(inverse(mat3(T))), column-major 3x3 + pad
    const DrawDev* draws;         // nDraws + 1
    VtxA* vtxA;
    VtxBC* vtxBC;                 // forward only
    float2* vtxUV;                // forward only, and only when textures are registered
    unsigned long long* vis;      // forward: W*H 64-bit keys
    uint32_t* depthBits;          // shadow: S*S depth bits (the shadow map itself)
    LargeRec* large; uint32_t largeCap;
    uint32_t* clipQueue; uint32_t clipCap;   // triangles that need true clipping (Appendix B.3), queued by K1 for k_clip_layout
    LargeRec* huge;               // HUGE_CAP records tested by every tile
    uint32_t tilesX, tilesY;
    TileHead* tileHead;           // per tile: its large-triangle list (16 B, zeroed by the pass prologue)
    LargeRec* tileRecs; uint32_t listCap;   // per-tile lists: contiguous runs of 48-B records (listCap records), streamed by the tile stage with bulk async copies
    const uint32_t* flatSlots;    // 3 cache slots per triangle in submission order (indexed mode)
    const uint32_t* triMaterial;  // raw material handle of every triangle's draw (indexed mode; shading reads it instead of walking to the draw record)
    const uint32_t* triBlockDraw; // draw containing the first triangle of every 256-triangle block
    const uint32_t* slotBlockDraw;// draw containing the first cache slot of every 256-slot block
    uint32_t* status;             // [0] largeCount [1] needsClip [2] overflow flags [3] tile-list total [4] hugeCount [5] - [6] clip-queue length [7] list allocation cursor
                                  // k_tile_binned's phase queues: [8] / [9] clip items claimed / done, [11] / [12] tile blocks, [10] / [13] queued triangles
    uint32_t* hostStatus;         // mapped pinned mirror of status[0..7], written by the tile stage (no D2H copy node)
    uint32_t rank, nRanks, bandH; // sort-first row bands (forward only)
    // Appendix B.15 (multisampling): this raster pass samples at pixel centre - shift / 256, implemented as an exact integer
    // shift of the snapped vertices (edge functions and depth are translation-invariant).  0 for 1-sample passes.
    int32_t shiftX, shiftY;
    // K1: log2 of the triangles a warp takes per batch (5 = one per lane; launch_setup_raster lowers it for short draw lists
    // so that every resident warp gets a batch: the row and sample stages spread the work over the lanes anyway)
    uint32_t batchLog2;
    // per-draw visibility for this rank's rows (sort-first split): written by k_cull_draws before K0, NULL = no culling
    const uint32_t* drawVisible;
    const uint32_t* blockVisible;   // the same per 256-triangle block (K1), written by k_cull_blocks
    // tile rows of this rank's bands, in order ("virtual" rows the forward tile stage iterates): bandRows = bandH / TILE_H,
    // vTilesY = how many tile rows the rank owns; without a split bandRows = vTilesY = tilesY
    uint32_t bandRows, vTilesY;
    // draw list re-ordered along a Morton curve (k_reorder.cu; NULL = submission order): source vertex streams in cache-slot order
    // for K0, and K1's stream of {slot0, slot1, slot2, submission-order ID} records
    const float4* posS; const float4* nrmS; const uint4* triStream;
    // sort-first split, per pass and frame (k_cull_blocks): the 256-triangle blocks this rank has to look at, compacted (their
    // number is cullCounts[0]; K1 walks this list instead of all blocks), and which 256-slot blocks of the post-transform cache
    // those blocks reference (K0 skips the others).  NULL without culling.
    const uint32_t* triBlockList; const uint32_t* cullCounts; const uint32_t* slotBlockVisible;
    // sort-LAST split of the shadow pass (reb_set_shadow_split): this rank rasterises the splitRank-th of splitN contiguous ranges of
    // K1's block stream into the WHOLE map (no row filter: nRanks stays 1) and the ranks' maps are combined with an element-wise
    // minimum.  (A triangle of the broadcast list is seen by the one rank whose range holds it, which rasterises all of it.)  0 / 1 = off.
    uint32_t splitRank, splitN;
    uint32_t cullStamp;           // slotBlockVisible[b] == cullStamp: marked this frame (a stamp instead of a clear; cullCounts alternates between two words)
};
constexpr int BLOCK_REFS = 16;   // words per 256-triangle block in the slot-block reference table: [0] count (0xFFFFFFFF: a range [1]..[2]), then block ids

// status[2] bits
constexpr uint32_t OVF_LARGE = 1u, OVF_LIST = 2u;

// Programmatic dependent launch (PDL): every kernel of a pass lets its successor's launch overlap its own last
// wave (trigger at entry) and waits for its predecessor's completion + memory flush before touching global data.
// Both are no-ops when the kernel was launched without the programmatic-serialization attribute.
__device__ __forceinline__ void pdl_enter() {
    asm volatile("griddepcontrol.launch_dependents;");
    asm volatile("griddepcontrol.wait;" ::: "memory");
}

struct V3 { float x, y, z; };

__device__ __forceinline__ long long mulw(int32_t a, int32_t b) { return (long long)a * (long long)b; }   // one IMAD.WIDE

// glm::mat4 * glm::vec4: (m0*x + m1*y) + (m2*z + m3*w) — ThirdParty/glm/glm/detail/type_mat4x4.inl:562-573
__device__ __forceinline__ float4 mat4_mul_point(const float* __restrict__ m, float x, float y, float z) {
    float4 o;
    o.x = (m[0] * x + m[4] * y) + (m[8] * z + m[12] * 1.0f);
    o.y = (m[1] * x + m[5] * y) + (m[9] * z + m[13] * 1.0f);
    o.z = (m[2] * x + m[6] * y) + (m[10] * z + m[14] * 1.0f);
    o.w = (m[3] * x + m[7] * y) + (m[11] * z + m[15] * 1.0f);
    return o;
}

__device__ __forceinline__ float dot3(V3 a, V3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
__device__ __forceinline__ V3 sub3(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
__device__ __forceinline__ V3 add3(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
__device__ __forceinline__ V3 neg3(V3 a) { return {-a.x, -a.y, -a.z}; }
__device__ __forceinline__ V3 normalize3(V3 v) {
    float inv = 1.0f / sqrtf(dot3(v, v));
    return {v.x * inv, v.y * inv, v.z * inv};
}
__device__ __forceinline__ float glm_max(float x, float y) { return (x < y) ? y : x; }
__device__ __forceinline__ float glm_min(float x, float y) { return (y < x) ? y : x; }

// primitive.vert:50-53 (DIVIDE = true) / shadow_map.vert:49-51 (DIVIDE = false)
template <bool DIVIDE>
__device__ __forceinline__ float4 vertex_stage(const float* __restrict__ PV, const float* __restrict__ T, float px, float py, float pz, V3* world) {
    float4 wH = mat4_mul_point(T, px, py, pz);
    V3 w = {wH.x, wH.y, wH.z};
    if (DIVIDE && wH.w != 1.0f) w = {wH.x / wH.w, wH.y / wH.w, wH.z / wH.w};   // x / 1.0f == x exactly: affine nodes skip the divides
    if (world) *world = w;
    return mat4_mul_point(PV, w.x, w.y, w.z);
}

// Appendix B.2-B.5 for one vertex: clip-plane outcodes, viewport transform, snap to 24.8 (round-half-even).
__device__ __forceinline__ VtxA viewport_snap(float4 c, uint32_t W, uint32_t H) {
    VtxA a;
    uint32_t f = 0;
    f |= (c.x < -c.w) ? 1u : 0u;  f |= (c.x > c.w) ? 2u : 0u;
    f |= (c.y < -c.w) ? 4u : 0u;  f |= (c.y > c.w) ? 8u : 0u;
    f |= (c.z < 0.0f) ? 16u : 0u; f |= (c.z > c.w) ? 32u : 0u;
    if (!(c.w > 0.0f)) { a.X = 0; a.Y = 0; a.z = 0.0f; a.flags = f | VF_W; return a; }
    const float hw = 0.5f * (float)W, hh = 0.5f * (float)H;
    const float rw = 1.0f / c.w;                                   // Appendix B.4 (revision r1b): one IEEE reciprocal per vertex
    const float nx = c.x * rw, ny = c.y * rw, nz = c.z * rw;
    const float xf = hw * nx + hw;
    const float yf = hh * ny + hh;
    if (!(fabsf(xf) <= GUARD_PX) || !(fabsf(yf) <= GUARD_PX)) { a.X = 0; a.Y = 0; a.z = nz; a.flags = f | VF_GUARD; return a; }
    a.X = __float2int_rn(xf * 256.0f);
    a.Y = __float2int_rn(yf * 256.0f);
    a.z = nz;
    a.flags = f;
    return a;
}

// Appendix B.8 (revision r1b): depth from the unbiased edge values of vertices 1 and 2 (dz1 = z1 - z0, dz2 = z2 - z0);
// inv = 1.0f / float(|A2|), one IEEE reciprocal per triangle.
__device__ __forceinline__ float depth_from(float fF1, float fF2, float inv, float z0, float dz1, float dz2) {
    float b1 = fF1 * inv;
    float b2 = fF2 * inv;
    float z = (z0 + b1 * dz1) + b2 * dz2;
    if (z == 0.0f) z = 0.0f;
    return z;
}

// binary search: last d in [0, nDraws) with key(d) <= v, where key = firstTri (BY_CACHE = false) or cacheBase
template <bool BY_CACHE>
__device__ __forceinline__ uint32_t draw_key(const DrawDev* __restrict__ draws, uint32_t d) {
    return BY_CACHE ? __ldg(&draws[d].cacheBase) : __ldg(&draws[d].firstTri);
}
template <bool BY_CACHE>
__device__ __forceinline__ uint32_t find_draw(const DrawDev* __restrict__ draws, uint32_t nDraws, uint32_t v) {
    uint32_t lo = 0, hi = nDraws;
    while (hi - lo > 1) {
        uint32_t mid = (lo + hi) >> 1;
        if (draw_key<BY_CACHE>(draws, mid) <= v) lo = mid; else hi = mid;
    }
    return lo;
}
// Table-assisted variant: blockDraw[v >> 8] is the draw holding key (v & ~255) (host-built when the draw list
// changes), followed by a short forward walk.  Relies on the sentinel record draws[nDraws].
// Bit 31 of a table entry (BLOCK_ONE_DRAW, set by the host for the cache-slot table): the whole 256-element block lies inside that
// draw, so the walk — a dependent load of the next draw's key, 24 % of the vertex kernel's stall samples — is skipped.
constexpr uint32_t BLOCK_ONE_DRAW = 0x80000000u;
constexpr uint32_t BLOCK_MIXED = 2u;   // blockVisible value of a 256-triangle block that spans two or more draws (sort-first culling)
template <bool BY_CACHE>
__device__ __forceinline__ uint32_t find_draw_block(const DrawDev* __restrict__ draws, uint32_t nDraws, const uint32_t* __restrict__ blockDraw, uint32_t v) {
    const uint32_t e = __ldg(blockDraw + (v >> 8));
    uint32_t d = e & ~BLOCK_ONE_DRAW;
    if (e & BLOCK_ONE_DRAW) return d;
    while (d + 1 < nDraws && draw_key<BY_CACHE>(draws, d + 1) <= v) ++d;
    return d;
}

// What shading needs to know about triangle `tri`: the cache slots of its corners and its draw's material handle.  Indexed
// mode: four independent loads from the tables k_flatten_slots wrote (the material handle used to sit behind the blockDraw ->
// draw key -> draw record chain; packing all four words into one 16-byte record instead made the tile stage 2 us faster but
// K1, which streams the slots of every triangle, 3 us slower per pass on C2 and 22 us on C3); de-indexed mode derives the
// slots and walks to the draw.
struct TriRef { uint32_t s[3]; uint32_t materialRaw; };
__device__ __forceinline__ TriRef tri_ref(const PassConst& pc, uint32_t tri) {
    TriRef r;
    if (pc.deindexed) {
        r.s[0] = 3u * tri; r.s[1] = 3u * tri + 1u; r.s[2] = 3u * tri + 2u;
        const uint32_t d = find_draw_block<false>(pc.draws, pc.nDraws, pc.triBlockDraw, tri);
        r.materialRaw = __ldg(&pc.draws[d].material);
    } else {
        const uint32_t* f = pc.flatSlots + 3ull * tri;
        r.s[0] = __ldg(f); r.s[1] = __ldg(f + 1); r.s[2] = __ldg(f + 2); r.materialRaw = __ldg(pc.triMaterial + tri);
    }
    REB_ASSERT(tri < pc.nTris && r.s[0] < pc.cacheSize && r.s[1] < pc.cacheSize && r.s[2] < pc.cacheSize);
    return r;
}

// Appendix B.3 (revision r1c): the clip-space polygon of parent triangle `tri` — vertex stage re-run on its three source
// vertices (the cache holds no clip coordinates), Sutherland-Hodgman against near z >= 0 and then the four guard planes,
// intersections computed from the inside vertex: t = d_in / (d_in - d_out).  Returns the vertex count (0 = nothing left).
__device__ __forceinline__ float clip_dist(int plane, float4 v, float Gx, float Gy) {
    switch (plane) {
        case 0: return v.z;
        case 1: return v.x + Gx * v.w;
        case 2: return Gx * v.w - v.x;
        case 3: return v.y + Gy * v.w;
        default: return Gy * v.w - v.y;
    }
}
template <bool SHADOW>
__device__ __forceinline__ int clip_parent(const PassConst& pc, uint32_t tri, float4 a[9]) {
    const uint32_t d = find_draw<false>(pc.draws, pc.nDraws, tri);
    const uint32_t* ix = pc.indices + __ldg(&pc.draws[d].indexOffset) + 3u * (tri - __ldg(&pc.draws[d].firstTri));
    const float4* Tq = reinterpret_cast<const float4*>(pc.transforms) + 4u * __ldg(&pc.draws[d].node);
    float T[16];
#pragma unroll
    for (int k = 0; k < 4; ++k) { const float4 q = __ldg(Tq + k); T[4 * k] = q.x; T[4 * k + 1] = q.y; T[4 * k + 2] = q.z; T[4 * k + 3] = q.w; }
    float4 b[9];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const float4 p = __ldg(pc.pos4 + __ldg(ix + k));
        a[k] = vertex_stage<!SHADOW>(pc.PV, T, p.x, p.y, p.z, nullptr);
    }
    const float Gx = (0.5f * GUARD_PX) / (0.5f * (float)pc.W), Gy = (0.5f * GUARD_PX) / (0.5f * (float)pc.H);
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
        if (n < 3) return 0;
        for (int k = 0; k < n; ++k) a[k] = b[k];
    }
    return n;
}

__device__ __forceinline__ bool row_owned(uint32_t y, uint32_t rank, uint32_t nRanks, uint32_t bandH) {
    return nRanks <= 1 || ((y / bandH) % nRanks) == rank;
}

// cache slot of corner k of triangle `tri` whose vertex index is v
__device__ __forceinline__ uint32_t cache_slot(const PassConst& pc, int32_t cacheOffset, uint32_t tri, uint32_t k, uint32_t v) {
    return pc.deindexed ? 3u * tri + k : (uint32_t)((int32_t)v + cacheOffset);
}

}  // namespace reb
