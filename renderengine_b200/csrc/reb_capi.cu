// reb_capi.cu — the C ABI (include/reb.h) over the raster kernels.  Host code only: owns device
// memory, orders the kernels on one CUDA stream, converts the reference's host records into kernel
// parameters.  There is no CPU rendering path here: every render call launches the sm_100a kernels.
#include "../../include/reb.h"
#include "reb_kernels.h"

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

using namespace reb;

namespace {

std::string g_createError;
// Images of the sRGB8 "swapchain": the reference asks for minImageCount + 1 (Swapchain.cpp:55-59), three on common drivers, so a
// finished frame is always waiting when the copy engine becomes free.  (Measured against two images on C2: end to end 163.3 vs
// 164.9 us / frame, i.e. no difference — the copy itself runs at 0.90 of the copy-only ceiling while the kernels of the next frames
// compete for HBM; the third image is kept for the reference's frames-in-flight behaviour, not for speed.)
constexpr int SWAP_IMAGES = 3;

struct PassState {               // per pass kind: 0 = shadow, 1 = forward (opaque), 2 = the translucent list of the base pass
    std::vector<reb_draw> draws; // last submitted list (host copy, for change detection and replay)
    DrawDev* dDraws = nullptr; uint32_t drawCap = 0;   // nDraws + 1 records (sentinel last)
    uint32_t nDraws = 0, nTris = 0;
    // post-transform vertex cache of this pass (K0 output)
    VtxA* vtxA = nullptr; VtxBC* vtxBC = nullptr; size_t cacheCap = 0;
    float2* vtxUV = nullptr; size_t uvCap = 0;          // forward pass, only once textures are registered
    uint32_t* blockDraw = nullptr; size_t blockDrawCap = 0; uint32_t triBlocks = 0;   // [triBlocks] by triangle, then by cache slot
    uint32_t* flatSlots = nullptr; uint32_t* triMaterial = nullptr; size_t flatCap = 0;   // 3 cache slots + the material handle per triangle, submission order (indexed mode); flatCap in triangles
    uint32_t cacheSize = 0, deindexed = 0;
    uint32_t* drawBounds = nullptr; uint32_t* drawVisible = nullptr; uint32_t cullCap = 0;
    float4* blockBox = nullptr; uint32_t* blockVisible = nullptr; uint32_t blockCullCap = 0;
    uint32_t* blockRefs = nullptr; uint32_t* triBlockList = nullptr; uint32_t* cullCounts = nullptr;   // [triBlocks * BLOCK_REFS], [triBlocks], [4]
    uint32_t* slotBlockVisible = nullptr; uint32_t slotBlocks = 0, slotBlockCap = 0; uint32_t cullStamp = 0;   // marks of frame f carry stamp f; cullCounts[f & 1] is its list length   // the same per 256-triangle block   // sort-first split: per-draw object-space bounds (encoded), per-frame visibility
    // re-ordered list (k_reorder.cu): source streams in slot order, K1's {slots, submission ID} stream
    float4* posS = nullptr; float4* nrmS = nullptr; size_t streamCap = 0; uint4* triStream = nullptr; size_t triStreamCap = 0; bool reordered = false;
    bool valid = false;          // dDraws / cache layout match `draws` and the current geometry
    bool pending = false;        // enqueued since the last overflow check
};

}  // namespace

struct reb_ctx {
    int device = 0; int numSMs = 148;
    cudaStream_t stream = nullptr; bool ownStream = false;
    std::string err;
    uint32_t W = 0, H = 0, outputs = 0, maxDesc = 8192;
    uint32_t rank = 0, nRanks = 1, bandH = 64;
    uint32_t sRank = 0, sNRanks = 1, sBandH = 64;   // row split of the shadow map (reb_set_shadow_partition)
    uint32_t sSplitRank = 0, sSplitN = 1;           // sort-last split of the shadow pass (reb_set_shadow_split)
    // geometry (SoA streams)
    float4* pos4 = nullptr; float4* nrm4 = nullptr; uint32_t* indices = nullptr; uint64_t nVerts = 0, nIdx = 0;
    float* transforms = nullptr; float* nmats = nullptr; uint32_t nNodes = 0;
    uint8_t* materials = nullptr; uint32_t materialStride = 0, nMaterials = 0;
    uint8_t* lights = nullptr; uint32_t lightStride = 0, nLights = 0;
    float lightsHost[SHADE_CONST_LIGHTS][28] = {};   // host mirror of the first records (ShadeConst::lightsC)
    // bindless textures (TextureHandle = running index, Bindless.cpp:221-224)
    std::vector<TexDev> texHost; std::vector<void*> texAllocs; TexDev* dTextures = nullptr; size_t texTableCap = 0; bool texDirty = false;
    // view
    float view[16], proj[16], PV[16], viewPos[3]; bool haveView = false;
    float exposure = 1.0f; uint32_t debugInput = 0, debugEquation = 0;   // ViewProperties @140 / @144 / @148 (PBR path)
    uint32_t shadingModel = 0;                                           // REB_SHADING_PHONG / REB_SHADING_PBR
    uint16_t* skyFaces = nullptr; uint32_t skyW = 0, skyH = 0;          // skybox cube map (reb_set_skybox), six RGBA16 faces
    // shadow (one directional shadow map, F10)
    uint32_t S = 0; float* shadowMap = nullptr; float sview[16], sproj[16], SPV[16]; bool haveShadowXf = false, shadowValid = false;
    // outputs + visibility buffer
    uint8_t* outSRGB8 = nullptr; float* outF32 = nullptr; float* outDepth = nullptr; uint32_t* outTriId = nullptr;
    // SWAP_IMAGES-image "swapchain" for the sRGB8 target (Swapchain.cpp:55-59,200-237):
    // reb_readback_async copies image i on a copy stream while the next frames render into the other images
    uint8_t* swap[SWAP_IMAGES] = {}; int swapCur = 0; bool copyInFlight[SWAP_IMAGES] = {};
    cudaStream_t copyStream = nullptr; cudaEvent_t evRendered = nullptr; cudaEvent_t evCopied[SWAP_IMAGES] = {};
    void* asyncHost = nullptr; bool asyncOutstanding = false;   // last reb_readback_async target (re-copied if the frame is replayed)
    unsigned long long* vis = nullptr;
    // translucent list (Appendix B.16, k_translucent.cu): per-pixel list heads, node pool, device counters {nodes, dropped, needs-clip};
    // internal linear-colour / depth targets when the caller did not enable those outputs
    uint32_t* tlHead = nullptr; uint4* tlNodes = nullptr; uint32_t tlCap = 0; uint32_t* tlCounters = nullptr; uint2* tlLarge = nullptr; uint32_t tlLargeCap = 0;
    float* tlHdr = nullptr; float* tlDepth = nullptr; size_t tlPx = 0;   // tlPx = pixels x samples the lists are sized for
    uint2* tlSpan = nullptr; uint2* tlFrags = nullptr; float4* tlCols = nullptr;   // split resolve (1 sample): per-pixel span, passing fragments {pixel, id}, their colours
    float4* msColour = nullptr;   // 4 sample planes of linear colour: a translucent list inside a multisampled pass blends per sample
    bool tlFollows = false, tlPending = false; ShadeConst tlShade; uint32_t tlIdBase = 0; uint64_t tlDropped = 0;
    // multisampling (Appendix B.15): 1 or 4 samples; with 4, `vis` holds four planes and each sample pass leaves its depth / triangle ID plane here
    uint32_t samples = 1; float* msDepth = nullptr; uint32_t* msTriId = nullptr;
    // queue sets of samples 1-3 for the one-pass multisampled K1 (sample 0 uses the forward pass's own): large-triangle queue, clip queue,
    // broadcast list, per-tile heads; their status words are sets 3-5 of dStatus, their host mirrors sets 2-4 of hStatus
    LargeRec* msLarge[3] = {}; uint32_t* msClip[3] = {}; LargeRec* msHuge = nullptr; TileHead* msTileHead[3] = {}; uint32_t msLargeAlloc = 0, msClipCap = 0, msTileCap = 0;
    // queues
    LargeRec* large = nullptr; uint32_t largeCap = 0;
    uint32_t* clipQueue = nullptr;   // largeCap entries
    uint32_t largeAlloc = 0; bool largeWorstCase = false;
    LargeRec* huge = nullptr;        // 2 x HUGE_CAP
    TileHead* tileHead[2] = {nullptr, nullptr}; uint32_t tileCap[2] = {0, 0};   // per pass: both may be in flight behind one fused K0
    cudaEvent_t shadowReadyEvent = nullptr;   // reb_shadow_wait_event: waited for right before the next forward tile stage (one-shot)
    bool shadowDeferred = false; float recordedSPV[16];   // reb_render_shadow recorded, not yet submitted
    LargeRec* tileList = nullptr; uint32_t listCap = 0;   // per-tile lists of 48-B records (listCap records)
    uint32_t* dStatus = nullptr;     // 6 x STATUS_WORDS words: shadow, forward, translucent, samples 1-3 of a multisampled forward pass
    uint32_t* hStatus = nullptr;     // pinned + mapped mirror, written by the tile stage kernels
    uint32_t* hStatusDev = nullptr;  // device-side address of hStatus
    PassState pass[3];
    uint64_t launches = 0, regrows = 0;
    bool profiling = false; cudaEvent_t ev[2][7] = {}; float kernelMs[2][6] = {};   // per-kernel timing slots (diagnostics)
    uint64_t counters[4] = {0, 0, 0, 0};
};

namespace {

int fail(reb_ctx* c, int code, const std::string& msg) {
    if (c) c->err = msg; else g_createError = msg;
    return code;
}

#define CU(call)                                                                                         \
    do {                                                                                                 \
        cudaError_t e_ = (call);                                                                         \
        if (e_ != cudaSuccess)                                                                           \
            return fail(ctx, REB_E_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_));            \
    } while (0)

// glm mat4*mat4, non-aligned path (type_mat4x4.inl:679-694): ((A0*b.x + A1*b.y) + A2*b.z) + A3*b.w
void mat4_mul(const float* a, const float* b, float* out) {
    float tmp[16];
    for (int c = 0; c < 4; ++c)
        for (int r = 0; r < 4; ++r) {
            float t = a[r] * b[c * 4];
            t = t + a[4 + r] * b[c * 4 + 1];
            t = t + a[8 + r] * b[c * 4 + 2];
            t = t + a[12 + r] * b[c * 4 + 3];
            tmp[c * 4 + r] = t;
        }
    std::memcpy(out, tmp, sizeof tmp);
}

// transpose(inverse(mat3(T))) hoisted per node (primitive.vert:55; glm func_matrix.inl:353-376)
void normal_matrix12(const float* T, float* out12) {
    const float a00 = T[0], a01 = T[1], a02 = T[2];     // column 0
    const float a10 = T[4], a11 = T[5], a12 = T[6];     // column 1
    const float a20 = T[8], a21 = T[9], a22 = T[10];    // column 2
    const float det = +a00 * (a11 * a22 - a21 * a12) - a10 * (a01 * a22 - a21 * a02) + a20 * (a01 * a12 - a11 * a02);
    const float ood = 1.0f / det;
    float inv[3][3];
    inv[0][0] = +(a11 * a22 - a21 * a12); inv[1][0] = -(a10 * a22 - a20 * a12); inv[2][0] = +(a10 * a21 - a20 * a11);
    inv[0][1] = -(a01 * a22 - a21 * a02); inv[1][1] = +(a00 * a22 - a20 * a02); inv[2][1] = -(a00 * a21 - a20 * a01);
    inv[0][2] = +(a01 * a12 - a11 * a02); inv[1][2] = -(a00 * a12 - a10 * a02); inv[2][2] = +(a00 * a11 - a10 * a01);
    for (int c = 0; c < 3; ++c)
        for (int r = 0; r < 3; ++r) out12[c * 3 + r] = inv[r][c] * ood;   // scaled, then transposed
    out12[9] = out12[10] = out12[11] = 0.0f;
}

template <class T> int dev_alloc(reb_ctx* ctx, T** p, size_t count) {
    if (*p) { cudaFree(*p); *p = nullptr; }
    if (count == 0) return REB_OK;
    CU(cudaMalloc(reinterpret_cast<void**>(p), count * sizeof(T)));
    return REB_OK;
}

int alloc_targets(reb_ctx* ctx) {
    const size_t px = (size_t)ctx->W * ctx->H;
    int rc;
    if ((rc = dev_alloc(ctx, &ctx->vis, px * ctx->samples))) return rc;
    launch_fill_u32(reinterpret_cast<uint32_t*>(ctx->vis), 0xFFFFFFFFu, px * 2 * ctx->samples, ctx->numSMs, ctx->stream); ctx->launches++;
    if ((rc = dev_alloc(ctx, &ctx->msDepth, ctx->samples > 1 ? px * ctx->samples : 0))) return rc;
    if ((rc = dev_alloc(ctx, &ctx->msTriId, ctx->samples > 1 ? px * ctx->samples : 0))) return rc;
    if (ctx->outputs & REB_OUT_COLOR_SRGB8) {
        for (int i = 0; i < SWAP_IMAGES; ++i) {
            if ((rc = dev_alloc(ctx, &ctx->swap[i], px * 4))) return rc;
            launch_fill_u32(reinterpret_cast<uint32_t*>(ctx->swap[i]), 0xFF000000u, px, ctx->numSMs, ctx->stream); ctx->launches++;
        }
        ctx->swapCur = 0; ctx->outSRGB8 = ctx->swap[0];
    }
    if (ctx->outputs & REB_OUT_COLOR_F32) {
        if ((rc = dev_alloc(ctx, &ctx->outF32, px * 4))) return rc;
        launch_fill_u32x4(ctx->outF32, 0u, 0u, 0u, 0x3F800000u, px, ctx->numSMs, ctx->stream); ctx->launches++;   // (0,0,0,1)
    }
    if (ctx->outputs & REB_OUT_DEPTH_F32) {
        if ((rc = dev_alloc(ctx, &ctx->outDepth, px))) return rc;
        launch_fill_u32(reinterpret_cast<uint32_t*>(ctx->outDepth), 0x3F800000u, px, ctx->numSMs, ctx->stream); ctx->launches++;
    }
    if (ctx->outputs & REB_OUT_TRIID_U32) {
        if ((rc = dev_alloc(ctx, &ctx->outTriId, px))) return rc;
        launch_fill_u32(ctx->outTriId, 0xFFFFFFFFu, px, ctx->numSMs, ctx->stream); ctx->launches++;
    }
    return REB_OK;
}

int ensure_tiles(reb_ctx* ctx, int kind, uint32_t nTiles) {
    int rc;
    if (nTiles + 1 > ctx->tileCap[kind]) {
        CU(cudaStreamSynchronize(ctx->stream));
        if ((rc = dev_alloc(ctx, &ctx->tileHead[kind], (size_t)nTiles + 1))) return rc;
        CU(cudaMemsetAsync(ctx->tileHead[kind], 0, ((size_t)nTiles + 1) * sizeof(TileHead), ctx->stream));   // K0's prologue zeroes them every pass
        ctx->tileCap[kind] = nTiles + 1;
    }
    if (kind == 1 && ctx->samples == 4 && nTiles + 1 > ctx->msTileCap) {
        CU(cudaStreamSynchronize(ctx->stream));
        for (int k = 0; k < 3; ++k) if ((rc = dev_alloc(ctx, &ctx->msTileHead[k], (size_t)nTiles + 1))) return rc;
        if (!ctx->msHuge && (rc = dev_alloc(ctx, &ctx->msHuge, (size_t)3 * HUGE_CAP))) return rc;
        ctx->msTileCap = nTiles + 1;
    }
    return REB_OK;
}

// Large-triangle queue: one slot per triangle plus room for the children of clipped triangles (a clipped parent yields up
// to 6; the slack starts at 25 % and is raised to the worst case by overflow recovery).  The clip queue holds triangle ids.
int ensure_large(reb_ctx* ctx, uint32_t nTris) {
    const uint64_t need = (uint64_t)nTris + 64u + (ctx->largeWorstCase ? 6ull * nTris : nTris / 4u);
    if (need > 0xFFFFFFF0ull) return fail(ctx, REB_E_OVERFLOW, "large-triangle queue would exceed 2^32 records");
    int rc;
    if (need > ctx->largeAlloc) {
        CU(cudaStreamSynchronize(ctx->stream));
        if ((rc = dev_alloc(ctx, &ctx->large, (size_t)need))) return rc;
        ctx->largeAlloc = (uint32_t)need;
    }
    if (nTris > ctx->largeCap) {
        CU(cudaStreamSynchronize(ctx->stream));
        if ((rc = dev_alloc(ctx, &ctx->clipQueue, (size_t)nTris))) return rc;
        ctx->largeCap = nTris;
    }
    if (ctx->samples == 4 && (ctx->msLargeAlloc < ctx->largeAlloc || ctx->msClipCap < ctx->largeCap)) {   // the same for samples 1-3
        CU(cudaStreamSynchronize(ctx->stream));
        for (int k = 0; k < 3; ++k)
            if ((rc = dev_alloc(ctx, &ctx->msLarge[k], (size_t)ctx->largeAlloc)) || (rc = dev_alloc(ctx, &ctx->msClip[k], (size_t)ctx->largeCap))) return rc;
        ctx->msLargeAlloc = ctx->largeAlloc; ctx->msClipCap = ctx->largeCap;
    }
    return REB_OK;
}

void fill_pass(reb_ctx* ctx, int kind, PassConst& pc);

// Draw-list preprocessing (runs when a list changes, not per frame — the reference does the analogous work in
// RenderScene::Init: PopulateMeshDrawCalls / SortOpaqueMeshes, RenderScene.cpp:94-125): running primitive counters,
// the vertex range every draw references (device reduction) and from it the layout of the post-transform cache.
int set_draws(reb_ctx* ctx, int kind, const reb_draw* draws, uint32_t n) {
    PassState& ps = ctx->pass[kind];
    if (ps.valid && ps.draws.size() == n && (n == 0 || std::memcmp(ps.draws.data(), draws, n * sizeof(reb_draw)) == 0)) return REB_OK;
    ps.valid = false;
    std::vector<DrawDev> dd((size_t)n + 1);
    uint64_t tri = 0;
    for (uint32_t i = 0; i < n; ++i) {
        if ((uint64_t)draws[i].indexOffset + draws[i].nbIndices > ctx->nIdx) return fail(ctx, REB_E_INVALID, "draw index range exceeds the index buffer");
        if (draws[i].sceneNode >= ctx->nNodes) return fail(ctx, REB_E_INVALID, "draw references an unknown scene node");
        dd[i] = {draws[i].indexOffset, (uint32_t)tri, draws[i].materialRaw, draws[i].sceneNode, 0, 0u, 0u, 0u};
        tri += draws[i].nbIndices / 3;
    }
    if (tri >= 0x55555555ull) return fail(ctx, REB_E_INVALID, "more than 2^32/3 primitives in one pass");
    dd[n] = {0u, (uint32_t)tri, 0u, 0u, 0, 0u, 0u, 0u};
    CU(cudaStreamSynchronize(ctx->stream));
    if (n + 1 > ps.drawCap) {
        int rc = dev_alloc(ctx, &ps.dDraws, (size_t)std::max(n + 1, 64u));
        if (rc) return rc;
        ps.drawCap = std::max(n + 1, 64u);
    }
    CU(cudaMemcpyAsync(ps.dDraws, dd.data(), dd.size() * sizeof(DrawDev), cudaMemcpyHostToDevice, ctx->stream));
    // per-draw referenced vertex range
    std::vector<uint32_t> ranges((size_t)n * 2);
    if (n) {
        for (uint32_t i = 0; i < n; ++i) { ranges[2 * i] = 0xFFFFFFFFu; ranges[2 * i + 1] = 0u; }
        uint32_t* dRanges = nullptr;
        CU(cudaMalloc(reinterpret_cast<void**>(&dRanges), ranges.size() * 4));
        cudaError_t e = cudaMemcpyAsync(dRanges, ranges.data(), ranges.size() * 4, cudaMemcpyHostToDevice, ctx->stream);
        if (e == cudaSuccess) { launch_draw_ranges(ctx->indices, ps.dDraws, n, dRanges, ctx->stream); ctx->launches++; }
        if (e == cudaSuccess) e = cudaMemcpyAsync(ranges.data(), dRanges, ranges.size() * 4, cudaMemcpyDeviceToHost, ctx->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
        cudaFree(dRanges);
        if (e != cudaSuccess) return fail(ctx, REB_E_CUDA, std::string("draw range reduction: ") + cudaGetErrorString(e));
    }
    uint64_t cache = 0;
    for (uint32_t i = 0; i < n; ++i) {
        const uint32_t lo = ranges[2 * i], hi = ranges[2 * i + 1];
        if (lo > hi) { dd[i].cacheBase = (uint32_t)std::min<uint64_t>(cache, 0xFFFFFFFFull); continue; }   // draw without primitives
        if (hi >= ctx->nVerts) return fail(ctx, REB_E_INVALID, "an index of draw " + std::to_string(i) + " exceeds the vertex buffer");
        dd[i].cacheBase = (uint32_t)std::min<uint64_t>(cache, 0xFFFFFFFFull);
        dd[i].nCache = hi - lo + 1;
        dd[i].cacheOffset = (int32_t)((int64_t)cache - (int64_t)lo);
        cache += dd[i].nCache;
    }
    // Indexed cache when the draws' vertex ranges are (nearly) disjoint, as the reference's loader produces them
    // (AssimpSceneLoader.cpp:198-261 appends a fresh vertex block per mesh); otherwise one slot per triangle corner.
    const bool forceDeindexed = std::getenv("REB_FORCE_DEINDEXED") != nullptr;
    ps.deindexed = (forceDeindexed || cache > 2 * ctx->nVerts + 65536ull || cache > 3 * tri || cache >= 0x7FFFFFFFull) ? 1u : 0u;
    const uint64_t slots = ps.deindexed ? 3 * tri : cache;
    if (ps.deindexed) for (uint32_t i = 0; i <= n; ++i) { dd[i].cacheBase = 3u * dd[i].firstTri; dd[i].cacheOffset = 0; }
    else dd[n].cacheBase = (uint32_t)cache;
    CU(cudaMemcpyAsync(ps.dDraws, dd.data(), dd.size() * sizeof(DrawDev), cudaMemcpyHostToDevice, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));   // dd is a stack-lifetime staging buffer
    if (slots > ps.cacheCap) {
        int rc;
        if ((rc = dev_alloc(ctx, &ps.vtxA, (size_t)slots))) return rc;
        if (kind >= 1 && (rc = dev_alloc(ctx, &ps.vtxBC, (size_t)slots))) return rc;
        ps.cacheCap = (size_t)slots;
    }
    ps.cacheSize = (uint32_t)slots;
    {   // per-256-element block lookup tables: draw holding the first triangle / cache slot of the block
        const uint32_t tb = (uint32_t)((tri + 255) / 256), sb = (uint32_t)((slots + 255) / 256);
        std::vector<uint32_t> tbl((size_t)tb + sb);
        uint32_t d = 0;
        for (uint32_t b = 0; b < tb; ++b) { while (d + 1 < n && dd[d + 1].firstTri <= b * 256ull) ++d; tbl[b] = d; }
        d = 0;
        for (uint32_t b = 0; b < sb; ++b) {
            while (d + 1 < n && dd[d + 1].cacheBase <= b * 256ull) ++d;
            // dd[n] is the sentinel (cacheBase = cache size): a block that ends before the next draw begins lies inside draw d
            const bool oneDraw = n > 0 && (uint64_t)dd[d + 1].cacheBase >= (b + 1) * 256ull;
            tbl[(size_t)tb + b] = d | (oneDraw ? BLOCK_ONE_DRAW : 0u);
        }
        if (tbl.size() > ps.blockDrawCap) {
            int rc = dev_alloc(ctx, &ps.blockDraw, tbl.size());
            if (rc) return rc;
            ps.blockDrawCap = tbl.size();
        }
        if (!tbl.empty()) {
            CU(cudaMemcpyAsync(ps.blockDraw, tbl.data(), tbl.size() * 4, cudaMemcpyHostToDevice, ctx->stream));
            CU(cudaStreamSynchronize(ctx->stream));
        }
        ps.triBlocks = tb;
    }
    ps.nDraws = n; ps.nTris = (uint32_t)tri;
    if (n) {   // object-space bounds per draw, for the per-rank draw culling of the sort-first split
        if (n > ps.cullCap) {
            int rc;
            if ((rc = dev_alloc(ctx, &ps.drawBounds, (size_t)n * 6)) || (rc = dev_alloc(ctx, &ps.drawVisible, (size_t)n))) return rc;
            ps.cullCap = n;
        }
        std::vector<uint32_t> init((size_t)n * 6);
        for (uint32_t i = 0; i < n; ++i) for (int c = 0; c < 6; ++c) init[6 * (size_t)i + c] = c < 3 ? 0xFFFFFFFFu : 0u;
        CU(cudaMemcpyAsync(ps.drawBounds, init.data(), init.size() * 4, cudaMemcpyHostToDevice, ctx->stream));
        launch_draw_bounds(ctx->indices, ctx->pos4, ps.dDraws, n, ps.drawBounds, ctx->stream); ctx->launches++;
        CU(cudaStreamSynchronize(ctx->stream));   // init is a stack-lifetime staging buffer
        if (ps.triBlocks > ps.blockCullCap) {
            int rc;
            if ((rc = dev_alloc(ctx, &ps.blockBox, (size_t)ps.triBlocks * 2)) || (rc = dev_alloc(ctx, &ps.blockVisible, (size_t)ps.triBlocks)) ||
                (rc = dev_alloc(ctx, &ps.blockRefs, (size_t)ps.triBlocks * BLOCK_REFS)) || (rc = dev_alloc(ctx, &ps.triBlockList, (size_t)ps.triBlocks)) ||
                (rc = dev_alloc(ctx, &ps.cullCounts, (size_t)4))) return rc;
            CU(cudaMemsetAsync(ps.cullCounts, 0, 16, ctx->stream));
            ps.blockCullCap = ps.triBlocks;
        }
        ps.slotBlocks = (uint32_t)((slots + 255) / 256);
        if (ps.slotBlocks > ps.slotBlockCap) {
            int rc;
            if ((rc = dev_alloc(ctx, &ps.slotBlockVisible, (size_t)ps.slotBlocks))) return rc;
            CU(cudaMemsetAsync(ps.slotBlockVisible, 0, (size_t)ps.slotBlocks * 4, ctx->stream));   // no stamp yet
            ps.slotBlockCap = ps.slotBlocks;
        }
    }
    ps.reordered = false;
    if (!ps.deindexed && tri) {   // index stream re-based into cache-slot space
        if (tri > ps.flatCap) {
            int rc;
            if ((rc = dev_alloc(ctx, &ps.flatSlots, (size_t)(3 * tri))) || (rc = dev_alloc(ctx, &ps.triMaterial, (size_t)tri))) return rc;
            ps.flatCap = (size_t)tri;
        }
        PassConst pc;
        fill_pass(ctx, kind, pc);
        launch_flatten_slots(pc, ps.flatSlots, ps.triMaterial, ctx->stream); ctx->launches++;
        CU(cudaStreamSynchronize(ctx->stream));
        // long lists: triangles along a Morton curve per draw, slots in first-use order (k_reorder.cu).  Short lists stay in
        // submission order: a list that fits the caches has no locality to win and K1's 16-B stream records cost more than
        // 12 B of slots (C1, 64 k triangles: K1 12 -> 16 us).
        static const bool noReorder = std::getenv("REB_NO_REORDER") != nullptr;
        static const uint64_t minTris = std::getenv("REB_REORDER_MIN_TRIS") ? std::strtoull(std::getenv("REB_REORDER_MIN_TRIS"), nullptr, 10) : 262144ull;
        if (!noReorder && tri >= minTris) {
            int rc;
            if (slots > ps.streamCap) {
                if ((rc = dev_alloc(ctx, &ps.posS, (size_t)slots)) || (rc = dev_alloc(ctx, &ps.nrmS, (size_t)slots))) return rc;
                ps.streamCap = (size_t)slots;
            }
            if (tri > ps.triStreamCap) {
                if ((rc = dev_alloc(ctx, &ps.triStream, (size_t)tri))) return rc;
                ps.triStreamCap = (size_t)tri;
            }
            CU(reorder_draw_list(pc, ps.drawBounds, ps.posS, ps.nrmS, ps.triStream, ps.flatSlots, ctx->stream));
            ctx->launches += 9;
            ps.reordered = true;
        }
    }
    if (n) {   // per-block bounds for the sort-first culling: blocks of K1's stream, i.e. after a re-ordering
        PassConst pcb;
        fill_pass(ctx, kind, pcb);
        launch_block_bounds(pcb, ps.blockBox, ctx->stream);
        launch_block_slot_refs(pcb, ps.blockRefs, ctx->stream); ctx->launches += 2;
        CU(cudaStreamSynchronize(ctx->stream));
    }
    ps.draws.assign(draws, draws + n);
    ps.valid = true;
    return REB_OK;
}

void fill_pass(reb_ctx* ctx, int kind, PassConst& pc) {
    const PassState& ps = ctx->pass[kind];
    std::memset(&pc, 0, sizeof pc);
    std::memcpy(pc.PV, kind == 0 ? ctx->SPV : ctx->PV, 64);
    pc.W = kind == 0 ? ctx->S : ctx->W; pc.H = kind == 0 ? ctx->S : ctx->H;
    pc.nTris = ps.nTris; pc.nDraws = ps.nDraws;
    pc.cacheSize = ps.cacheSize; pc.deindexed = ps.deindexed;
    pc.indices = ctx->indices; pc.pos4 = ctx->pos4; pc.nrm4 = ctx->nrm4; pc.transforms = ctx->transforms; pc.nmats = ctx->nmats;
    pc.draws = ps.dDraws; pc.vtxA = ps.vtxA; pc.vtxBC = ps.vtxBC; pc.vtxUV = (kind >= 1 && !ctx->texHost.empty()) ? ps.vtxUV : nullptr;
    pc.vis = ctx->vis; pc.depthBits = reinterpret_cast<uint32_t*>(ctx->shadowMap);
    pc.large = ctx->large; pc.largeCap = ctx->largeAlloc; pc.clipQueue = ctx->clipQueue; pc.clipCap = ctx->largeCap; pc.huge = ctx->huge + (size_t)kind * HUGE_CAP;
    pc.tilesX = (pc.W + TILE_W - 1) / TILE_W; pc.tilesY = (pc.H + TILE_H - 1) / TILE_H;
    pc.tileHead = ctx->tileHead[kind == 2 ? 1 : kind]; pc.tileRecs = ctx->tileList; pc.listCap = ctx->listCap;
    pc.flatSlots = ps.flatSlots; pc.triMaterial = ps.triMaterial;
    if (ps.reordered) { pc.posS = ps.posS; pc.nrmS = ps.nrmS; pc.triStream = ps.triStream; }
    pc.triBlockDraw = ps.blockDraw; pc.slotBlockDraw = ps.blockDraw ? ps.blockDraw + ps.triBlocks : nullptr;
    pc.status = ctx->dStatus + STATUS_WORDS * kind;
    pc.hostStatus = ctx->hStatusDev ? ctx->hStatusDev + 8 * kind : nullptr;
    if (kind >= 1) { pc.rank = ctx->rank; pc.nRanks = ctx->nRanks; pc.bandH = ctx->bandH; }
    else { pc.rank = ctx->sRank; pc.nRanks = ctx->sNRanks; pc.bandH = ctx->sBandH; pc.splitRank = ctx->sSplitRank; pc.splitN = ctx->sSplitN; }
    pc.bandRows = pc.tilesY; pc.vTilesY = pc.tilesY;
    if (pc.nRanks > 1) {
        pc.bandRows = pc.bandH / TILE_H; pc.vTilesY = 0;
        for (uint32_t b = pc.rank; b * pc.bandRows < pc.tilesY; b += pc.nRanks) pc.vTilesY += std::min(pc.bandRows, pc.tilesY - b * pc.bandRows);
    }
    static const bool noCull = std::getenv("REB_NO_DRAW_CULL") != nullptr;
    pc.drawVisible = ((pc.nRanks > 1 || pc.splitN > 1) && !noCull && ps.nDraws) ? ps.drawVisible : nullptr;
    if (!pc.drawVisible) pc.splitN = 0;   // (REB_NO_DRAW_CULL: every rank rasterises everything; the minimum is the same)
    pc.blockVisible = pc.drawVisible ? ps.blockVisible : nullptr;
    if (pc.drawVisible) { pc.triBlockList = ps.triBlockList; pc.cullCounts = ps.cullCounts; pc.slotBlockVisible = ps.slotBlockVisible; }   // (run_pass picks the stamp and the length word)
}

// Everything a pass needs before its kernels can be enqueued: queue capacities, texture table, kernel parameters.
int prepare_pass(reb_ctx* ctx, int kind, PassConst& pc) {
    PassState& ps = ctx->pass[kind];
    int rc;
    if ((rc = ensure_large(ctx, ps.nTris))) return rc;
    if (kind >= 1 && !ctx->texHost.empty()) {
        if (ps.cacheSize > ps.uvCap) {
            CU(cudaStreamSynchronize(ctx->stream));
            if ((rc = dev_alloc(ctx, &ps.vtxUV, (size_t)ps.cacheSize))) return rc;
            ps.uvCap = ps.cacheSize;
        }
        if (ctx->texDirty) {   // descriptor table: handle -> (pointer, extent, mips)
            CU(cudaStreamSynchronize(ctx->stream));
            if (ctx->texHost.size() > ctx->texTableCap) {
                if ((rc = dev_alloc(ctx, &ctx->dTextures, ctx->texHost.size() * 2))) return rc;
                ctx->texTableCap = ctx->texHost.size() * 2;
            }
            CU(cudaMemcpyAsync(ctx->dTextures, ctx->texHost.data(), ctx->texHost.size() * sizeof(TexDev), cudaMemcpyHostToDevice, ctx->stream));
            CU(cudaStreamSynchronize(ctx->stream));
            ctx->texDirty = false;
        }
    }
    fill_pass(ctx, kind, pc);
    if (kind < 2 && (rc = ensure_tiles(ctx, kind, pc.tilesX * pc.tilesY))) return rc;
    fill_pass(ctx, kind, pc);
    return REB_OK;
}

// Enqueue the kernels of one pass: K0 vertex stage (unless the fused two-pass K0 already ran), K1 setup + small raster,
// K2 binned path (clip + list layout + scatter + TMA-staged tile raster; leaves at once when K1 queued nothing), K3 tile stage.
// Every kernel after the first is PDL-chained.
// Appendix B.14: inverse of A = rows x, y, w of proj * mat4(mat3(view)) restricted to 3x3 — maps (xn, yn, 1) to the
// direction of the view ray (skybox.vert:32-37).  Adjugate / determinant in double, rounded to float once per frame.
static void sky_ray_matrix(const float* view, const float* proj, float* out9) {
    double A[3][3];
    const int rows[3] = {0, 1, 3};
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) {
            double acc = 0.0;
            for (int k = 0; k < 3; ++k) acc += (double)proj[4 * k + rows[r]] * (double)view[4 * c + k];
            A[r][c] = acc;
        }
    const double c00 = A[1][1] * A[2][2] - A[1][2] * A[2][1], c01 = A[1][2] * A[2][0] - A[1][0] * A[2][2], c02 = A[1][0] * A[2][1] - A[1][1] * A[2][0];
    const double det = A[0][0] * c00 + A[0][1] * c01 + A[0][2] * c02;
    const double adj[3][3] = {
        {c00, A[0][2] * A[2][1] - A[0][1] * A[2][2], A[0][1] * A[1][2] - A[0][2] * A[1][1]},
        {c01, A[0][0] * A[2][2] - A[0][2] * A[2][0], A[0][2] * A[1][0] - A[0][0] * A[1][2]},
        {c02, A[0][1] * A[2][0] - A[0][0] * A[2][1], A[0][0] * A[1][1] - A[0][1] * A[1][0]}};
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 3; ++c) out9[3 * c + r] = (float)(adj[r][c] / det);
}

int run_pass(reb_ctx* ctx, int kind, const PassConst& pcIn, bool vertexDone) {
    PassState& ps = ctx->pass[kind];
    PassConst pcFrame = pcIn;
    if (pcFrame.drawVisible && ps.nTris) {   // exactly one k_cull launch per increment: its stamp, and which of the two list-length words it counts in
        ++ps.cullStamp;
        pcFrame.cullStamp = ps.cullStamp; pcFrame.cullCounts = ps.cullCounts + (ps.cullStamp & 1u);
    }
    const PassConst& pc = pcFrame;
    // slots: 0 clear, 1 vertex stage, 2 setup + small raster (+ tile counts/layout), 3 bin scatter, 4 tile stage
    const bool fusedShadow = vertexDone && kind == 0;   // slots 0-1 were stamped before the fused two-pass vertex kernel
    auto mark = [&](int i) { if (ctx->profiling && !(fusedShadow && i < 2)) cudaEventRecord(ctx->ev[kind][i], ctx->stream); };
    mark(0);
    if (!ps.nTris) {   // no K0 to run the pass prologue: reset status here; an empty shadow list still has to clear the map
        CU(cudaMemsetAsync(pc.status, 0, STATUS_WORDS * 4, ctx->stream));
        if (kind == 0) { launch_fill_u32(pc.depthBits, 0x3F800000u, (size_t)ctx->S * ctx->S, ctx->numSMs, ctx->stream); ctx->launches++; }
    }
    mark(1);
    if (ps.nTris && pc.drawVisible) {
        launch_cull(pc, kind == 0, ps.drawBounds, ps.drawVisible, ps.blockBox, ps.triBlocks, ps.blockVisible, ps.blockRefs, ps.slotBlockVisible, ps.triBlockList,
                    ps.cullCounts + (pc.cullStamp & 1u), ps.cullCounts + ((pc.cullStamp + 1u) & 1u), ctx->stream);
        ctx->launches++;
    }
    if (ps.nTris && !vertexDone) { launch_vertex(pc, kind == 0, /*pdl=*/pc.drawVisible != nullptr, ctx->stream); ctx->launches++; }   // (PDL behind k_cull: K0's prologue and clear overlap it)
    mark(2);
    ShadeConst sh;
    if (kind == 1) {
        if (ctx->swap[0]) {   // acquire the next swapchain image; wait for an async readback still draining it
            ctx->swapCur = (ctx->swapCur + 1) % SWAP_IMAGES; ctx->outSRGB8 = ctx->swap[ctx->swapCur];
            if (ctx->copyInFlight[ctx->swapCur]) { CU(cudaStreamWaitEvent(ctx->stream, ctx->evCopied[ctx->swapCur], 0)); ctx->copyInFlight[ctx->swapCur] = false; }
        }
        std::memset(&sh, 0, sizeof sh);
        std::memcpy(sh.STR, ctx->SPV, 64);
        std::memcpy(sh.viewPos, ctx->viewPos, 12);
        sh.nLights = ctx->nLights;
        sh.materials = ctx->materials; sh.materialStride = ctx->materialStride; sh.nMaterials = ctx->nMaterials;
        sh.lights = ctx->lights; sh.lightStride = ctx->lightStride;
        std::memcpy(sh.lightsC, ctx->lightsHost, sizeof sh.lightsC);
        const bool useShadow = ctx->shadowMap && ctx->shadowValid && ctx->haveShadowXf;
        sh.shadowMap = useShadow ? ctx->shadowMap : nullptr; sh.shadowSize = useShadow ? ctx->S : 0;
        sh.shadingModel = ctx->shadingModel; sh.exposure = ctx->exposure; sh.debugInput = ctx->debugInput; sh.debugEquation = ctx->debugEquation;
        sh.maxDescriptors = ctx->maxDesc; sh.nTextures = (uint32_t)ctx->texHost.size(); sh.textures = ctx->dTextures;
        sh.outSRGB8 = ctx->outSRGB8; sh.outF32 = ctx->outF32; sh.outDepth = ctx->outDepth; sh.outTriId = ctx->outTriId;
        ctx->tlShade = sh;   // (outputs as the caller enabled them; the translucent resolve writes these)
        if (ctx->tlFollows) {   // Appendix B.16: the opaque pass leaves linear colour + depth for the translucent list; sky and encode come after it
            sh.outSRGB8 = nullptr;
            sh.outF32 = ctx->outF32 ? ctx->outF32 : ctx->tlHdr;
            sh.outDepth = ctx->outDepth ? ctx->outDepth : ctx->tlDepth;
        }
        if (ctx->skyFaces) {
            sh.skyFaces = ctx->skyFaces; sh.skyW = ctx->skyW; sh.skyH = ctx->skyH;
            sky_ray_matrix(ctx->view, ctx->proj, sh.skyAinv);
            sh.skyHw = 0.5f * (float)ctx->W; sh.skyHh = 0.5f * (float)ctx->H;
            sh.skyRhw = 1.0f / sh.skyHw; sh.skyRhh = 1.0f / sh.skyHh;
            ShadeConst& t = ctx->tlShade;   // the same sky parameters for the translucent resolve, which draws it in that case
            t.skyFaces = sh.skyFaces; t.skyW = sh.skyW; t.skyH = sh.skyH; std::memcpy(t.skyAinv, sh.skyAinv, sizeof t.skyAinv);
            t.skyHw = sh.skyHw; t.skyHh = sh.skyHh; t.skyRhw = sh.skyRhw; t.skyRhh = sh.skyRhh;
            if (ctx->tlFollows) sh.skyFaces = nullptr;
        }
    }
    if (kind == 1 && ctx->samples == 4) {
        // Appendix B.15: four raster passes over the same post-transform cache, one per Vulkan standard sample location
        // ((0.375, 0.125), (0.875, 0.375), (0.125, 0.625), (0.625, 0.875)), each an exact shift of the snapped vertices by
        // (128 - 256 s) sub-pixel units; every pass leaves its depth / triangle-ID plane, the resolve kernel shades.
        static const int32_t shift[4][2] = {{32, 96}, {-96, 32}, {96, -32}, {-32, -96}};
        const size_t px = (size_t)ctx->W * ctx->H;
        // ONE K1 pass tests every triangle against the four sample positions and fills the four visibility planes and the four
        // queue sets; K2 and the (visibility-only) tile stage then run per sample on that sample's set
        MsTargets ms;
        for (int k = 0; k < 4; ++k) {
            ms.vis[k] = ctx->vis + (size_t)k * px;
            ms.status[k] = k ? ctx->dStatus + (size_t)(2 + k) * STATUS_WORDS : pc.status;
            ms.large[k] = k ? ctx->msLarge[k - 1] : pc.large;
            ms.clipQueue[k] = k ? ctx->msClip[k - 1] : pc.clipQueue;
            ms.huge[k] = k ? ctx->msHuge + (size_t)(k - 1) * HUGE_CAP : pc.huge;
            ms.tileHead[k] = k ? ctx->msTileHead[k - 1] : pc.tileHead;
        }
        if (ps.nTris) { launch_setup_raster_ms(pc, ms, ctx->numSMs, ctx->stream); ctx->launches += 2; }
        else CU(cudaMemsetAsync(ctx->dStatus + 3 * STATUS_WORDS, 0, 3 * STATUS_WORDS * 4, ctx->stream));   // no K1 to arm the sets of samples 1-3: their tile stages read the list lengths
        for (int k = 0; k < 4; ++k) {
            PassConst pk = pc;
            pk.vis = ms.vis[k]; pk.shiftX = shift[k][0]; pk.shiftY = shift[k][1];   // (the shift still applies to the children K2 clips)
            pk.status = ms.status[k]; pk.large = ms.large[k]; pk.clipQueue = ms.clipQueue[k]; pk.huge = ms.huge[k]; pk.tileHead = ms.tileHead[k];
            if (k && ctx->hStatusDev) pk.hostStatus = ctx->hStatusDev + 8 * (1 + k);
            ShadeConst sk = sh;
            sk.outSRGB8 = nullptr; sk.outF32 = nullptr; sk.skyFaces = nullptr;   // visibility only
            sk.outDepth = ctx->msDepth + (size_t)k * px; sk.outTriId = ctx->msTriId + (size_t)k * px;
            if (ps.nTris) { launch_tile_binned(pk, false, ctx->numSMs, ctx->stream); ctx->launches++; }
            launch_tile_forward(pk, sk, ctx->numSMs, ctx->stream); ctx->launches++;   // (folding these four passes into the resolve was measured: 4 x 21 + 118 us became 207 us)
        }
        mark(3); mark(4);
        if (ctx->shadowReadyEvent) { CU(cudaStreamWaitEvent(ctx->stream, ctx->shadowReadyEvent, 0)); ctx->shadowReadyEvent = nullptr; }
        launch_msaa_resolve(pc, sh, ctx->msDepth, ctx->msTriId, ctx->tlFollows ? ctx->msColour : nullptr, ctx->stream); ctx->launches++;
    } else {
        if (ps.nTris) { launch_setup_raster(pc, kind == 0, ctx->numSMs, ctx->stream); ctx->launches++; }
        mark(3);
        if (ps.nTris) { launch_tile_binned(pc, kind == 0, ctx->numSMs, ctx->stream); ctx->launches++; }   // K2: leaves at once when K1 queued nothing
        mark(4);
        if (kind == 0) { launch_tile_shadow(pc, ctx->numSMs, ctx->stream); ctx->launches++; }
        else {
            // the shadow map is being filled by an exchange on another stream (reb_shadow_wait_event): only this kernel reads it
            if (ctx->shadowReadyEvent) { CU(cudaStreamWaitEvent(ctx->stream, ctx->shadowReadyEvent, 0)); ctx->shadowReadyEvent = nullptr; }
            launch_tile_forward(pc, sh, ctx->numSMs, ctx->stream); ctx->launches++;
        }
    }
    mark(5);
    CU(cudaGetLastError());
    ps.pending = true;
    return REB_OK;
}

// The shadow pass is RECORDED by reb_render_shadow and submitted with the forward pass (the reference records both
// passes into one command buffer per frame, RenderScene.cpp:168-184, and submits once): when both draw lists are the
// same the two vertex stages run as one kernel.  Anything that needs the shadow map earlier flushes the recording.
int flush_deferred(reb_ctx* ctx) {
    if (!ctx->shadowDeferred) return REB_OK;
    ctx->shadowDeferred = false;
    PassConst pc;
    int rc = prepare_pass(ctx, 0, pc);
    if (rc) return rc;
    std::memcpy(pc.PV, ctx->recordedSPV, 64);
    return run_pass(ctx, 0, pc, false);
}

int submit_forward(reb_ctx* ctx) {
    PassConst pf, psh;
    int rc;
    if ((rc = prepare_pass(ctx, 1, pf))) return rc;
    const PassState &s0 = ctx->pass[0], &s1 = ctx->pass[1];
    const bool fuse = ctx->shadowDeferred && s0.nTris && s1.nTris && s0.deindexed == s1.deindexed && s0.cacheSize == s1.cacheSize &&
                      s0.draws.size() == s1.draws.size() && std::memcmp(s0.draws.data(), s1.draws.data(), s0.draws.size() * sizeof(reb_draw)) == 0 &&
                      ctx->nRanks <= 1 && ctx->sNRanks <= 1 && ctx->sSplitN <= 1 &&   // a split pass culls per rank: its K0 is not the other pass's
                      !std::getenv("REB_NO_FUSED_VERTEX");
    if (!fuse) {
        if ((rc = flush_deferred(ctx))) return rc;
        if ((rc = prepare_pass(ctx, 1, pf))) return rc;   // the flush may have grown shared queues
        return run_pass(ctx, 1, pf, false);
    }
    ctx->shadowDeferred = false;
    if ((rc = prepare_pass(ctx, 0, psh))) return rc;
    if ((rc = prepare_pass(ctx, 1, pf))) return rc;
    std::memcpy(psh.PV, ctx->recordedSPV, 64);
    if (ctx->profiling) { cudaEventRecord(ctx->ev[0][0], ctx->stream); cudaEventRecord(ctx->ev[0][1], ctx->stream); }
    launch_vertex_both(psh, pf, ctx->stream); ctx->launches++;
    if ((rc = run_pass(ctx, 0, psh, true))) return rc;
    return run_pass(ctx, 1, pf, true);
}

// After a stream sync: look at the status words of the passes enqueued since the last check; grow the
// tile-list queue and replay the frame if it overflowed (failure detection / recovery).
// Appendix B.16: the translucent list of the base pass, after the opaque kernels on the same stream (k_translucent.cu).
int run_translucent(reb_ctx* ctx) {
    PassConst pc;
    int rc;
    if ((rc = prepare_pass(ctx, 2, pc))) return rc;
    pc.tilesX = pc.tilesY = 0;                       // K0's prologue has no tile counters to re-arm for this list
    pc.drawVisible = nullptr; pc.blockVisible = nullptr; pc.triBlockList = nullptr; pc.cullCounts = nullptr; pc.slotBlockVisible = nullptr;
    pc.status = ctx->dStatus + 2 * STATUS_WORDS;
    pc.hostStatus = nullptr;
    ShadeConst sh = ctx->tlShade;
    float* hdr = ctx->outF32 ? ctx->outF32 : ctx->tlHdr;
    float* depth = ctx->outDepth ? ctx->outDepth : ctx->tlDepth;
    if (pc.nTris > ctx->tlLargeCap) {                 // list of the triangles pass B rasterises (worst case: all of them)
        CU(cudaStreamSynchronize(ctx->stream));
        if ((rc = dev_alloc(ctx, &ctx->tlLarge, (size_t)pc.nTris))) return rc;
        ctx->tlLargeCap = pc.nTris;
    }
    CU(cudaMemsetAsync(ctx->tlCounters, 0, 32, ctx->stream));
    if (ctx->samples == 4) {
        // Appendix B.15 + B.16: one raster pass of the translucent list per sample (the sample's exact vertex shift, its own opaque depth
        // plane and list-head plane; the node pool is shared), then the merge over the per-sample colours the opaque resolve left
        static const int32_t shift[4][2] = {{32, 96}, {-96, 32}, {96, -32}, {-32, -96}};
        const size_t px = (size_t)ctx->W * ctx->H;
        if (pc.nTris) { launch_vertex(pc, false, /*pdl=*/false, ctx->stream); ctx->launches++; }
        for (int k = 0; k < 4 && pc.nTris; ++k) {
            PassConst pk = pc;
            pk.shiftX = shift[k][0]; pk.shiftY = shift[k][1];
            if (k) CU(cudaMemsetAsync(ctx->tlCounters + 4, 0, 8, ctx->stream));   // the large-triangle work list restarts; the node counter runs on
            launch_tl_raster(pk, ctx->msDepth + (size_t)k * px, ctx->tlHead + (size_t)k * px, ctx->tlNodes, ctx->tlCap, ctx->tlCounters, ctx->tlLarge, ctx->numSMs, ctx->stream);
            ctx->launches += 2;
        }
        pc.tilesX = (pc.W + TILE_W - 1) / TILE_W; pc.tilesY = (pc.H + TILE_H - 1) / TILE_H;
        static const bool fusedMs = std::getenv("REB_TL_FUSED_RESOLVE") != nullptr;
        if (fusedMs || !ctx->tlFrags) { launch_tl_resolve_ms(pc, sh, ctx->tlHead, ctx->tlNodes, ctx->tlCap, ctx->tlIdBase, ctx->msColour, ctx->msDepth, ctx->msTriId, ctx->stream); ctx->launches++; }
        else {
            launch_tl_resolve_ms_split(pc, sh, ctx->tlHead, ctx->tlNodes, ctx->tlCap, ctx->tlIdBase, ctx->msColour, ctx->msDepth, ctx->msTriId, ctx->tlCounters, ctx->tlSpan, ctx->tlFrags,
                                       ctx->tlCols, ctx->numSMs, ctx->stream);
            ctx->launches += 3;
        }
        CU(cudaGetLastError());
        ctx->tlPending = true;
        return REB_OK;
    }
    if (pc.nTris) {
        launch_vertex(pc, false, /*pdl=*/false, ctx->stream);
        launch_tl_raster(pc, depth, ctx->tlHead, ctx->tlNodes, ctx->tlCap, ctx->tlCounters, ctx->tlLarge, ctx->numSMs, ctx->stream);
        ctx->launches += 3;
    }
    pc.tilesX = (pc.W + TILE_W - 1) / TILE_W; pc.tilesY = (pc.H + TILE_H - 1) / TILE_H;
    static const bool fusedResolve = std::getenv("REB_TL_FUSED_RESOLVE") != nullptr;   // A/B: the one-kernel resolve
    if (fusedResolve || !ctx->tlFrags) { launch_tl_resolve(pc, sh, ctx->tlHead, ctx->tlNodes, ctx->tlCap, ctx->tlIdBase, hdr, depth, ctx->tlCounters, ctx->stream); ctx->launches++; }
    else {
        launch_tl_resolve_split(pc, sh, ctx->tlHead, ctx->tlNodes, ctx->tlCap, ctx->tlIdBase, hdr, depth, ctx->tlCounters, ctx->tlSpan, ctx->tlFrags, ctx->tlCols, ctx->numSMs, ctx->stream);
        ctx->launches += 3;
    }
    CU(cudaGetLastError());
    ctx->tlPending = true;
    return REB_OK;
}

int check_and_recover(reb_ctx* ctx) {
    for (int iter = 0; iter < 8; ++iter) {
        bool overflow = false;
        uint32_t need = ctx->listCap;
        bool first = true;                           // counters[0] = drops of the passes checked now (not carried over from earlier frames)
        for (int k = 0; k < 2; ++k) {
            if (!ctx->pass[k].pending) continue;
            const int nSets = (k == 1 && ctx->samples == 4) ? 4 : 1;   // a multisampled forward pass has one status set per sample
            for (int q = 0; q < nSets; ++q) {
                const uint32_t* st = ctx->hStatus + 8 * (q ? 1 + q : k);
                ctx->counters[0] = (first ? 0 : ctx->counters[0]) + st[1];
                ctx->counters[1] = (first ? 0 : ctx->counters[1]) + st[0];    // binned triangles / tile-list entries: summed over the passes of the frame
                ctx->counters[2] = (first ? 0 : ctx->counters[2]) + st[3];
                first = false;
                if (st[2] & OVF_LARGE) { overflow = true; ctx->largeWorstCase = true; }   // clip children outgrew the slack: size for the worst case
                if (st[2] & OVF_LIST) { overflow = true; need = std::max(need, st[3] + st[3] / 4 + 1024u); }
            }
        }
        if (!overflow) {
            ctx->pass[0].pending = ctx->pass[1].pending = false;
            if (ctx->tlPending) {   // fragments / triangles the translucent pass had to drop (list capacity, Appendix B.3 clipping)
                uint32_t c[8] = {};   // nodes, dropped fragments, dropped clip children, -, work items of pass B, triangles of pass B
                CU(cudaMemcpy(c, ctx->tlCounters, 32, cudaMemcpyDeviceToHost));
                ctx->tlDropped = (uint64_t)c[1] + c[2];
                if (std::getenv("REB_TL_DEBUG")) std::fprintf(stderr, "[reb] translucent pass: %u list nodes, %u fragments dropped, %u clip children dropped, %u triangles / %u tiles in pass B\n", c[0], c[1], c[2], c[5], c[4]);
                ctx->counters[0] += ctx->tlDropped;
                ctx->tlPending = false;
            }
            if (iter > 0 && ctx->asyncOutstanding && ctx->asyncHost && ctx->outSRGB8) {   // the frame was replayed: refresh the async copy
                CU(cudaMemcpyAsync(ctx->asyncHost, ctx->outSRGB8, (size_t)ctx->W * ctx->H * 4, cudaMemcpyDeviceToHost, ctx->stream));
                CU(cudaStreamSynchronize(ctx->stream));
            }
            ctx->asyncOutstanding = false;
            return REB_OK;
        }
        int rc = dev_alloc(ctx, &ctx->tileList, (size_t)need);
        if (rc) return rc;
        ctx->listCap = need; ctx->regrows++; ctx->counters[3] = ctx->regrows;
        const bool p0 = ctx->pass[0].pending, p1 = ctx->pass[1].pending;
        for (int k = 0; k < 2; ++k) {
            if (!(k == 0 ? p0 : p1)) continue;
            PassConst pc;
            if ((rc = prepare_pass(ctx, k, pc))) return rc;
            if (k == 0) std::memcpy(pc.PV, ctx->recordedSPV, 64);
            // the replayed opaque pass must leave what the translucent list composites over (linear colour + depth, sky / encode deferred)
            if (k == 1) ctx->tlFollows = ctx->tlPending;
            rc = run_pass(ctx, k, pc, false);
            if (k == 1) ctx->tlFollows = false;
            if (rc) return rc;
        }
        if (p1 && ctx->tlPending && (rc = run_translucent(ctx))) return rc;   // the opaque frame was replayed under it
        CU(cudaStreamSynchronize(ctx->stream));
    }
    return fail(ctx, REB_E_OVERFLOW, "tile-list queue kept overflowing after 8 regrow attempts");
}

}  // namespace

#include "reb_capi_entry.inl"
