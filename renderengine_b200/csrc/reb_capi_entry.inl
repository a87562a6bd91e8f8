// reb_capi_entry.inl — extern "C" entry points (included by reb_capi.cu).

#define CHECK_CTX() do { if (!ctx) return REB_E_INVALID; } while (0)

extern "C" {

int reb_abi_version(void) { return REB_ABI_VERSION; }

const char* reb_last_error(reb_ctx* ctx) { return ctx ? ctx->err.c_str() : g_createError.c_str(); }

int reb_create(const reb_params* params, reb_ctx** out) {
    reb_ctx* ctx = nullptr;
    if (!params || !out || params->struct_size != sizeof(reb_params)) return fail(nullptr, REB_E_INVALID, "reb_create: bad params / struct_size");
    if (params->width == 0 || params->height == 0) return fail(nullptr, REB_E_INVALID, "reb_create: zero extent");
    if (params->width > REB_MAX_EXTENT || params->height > REB_MAX_EXTENT)   // K1 packs x0 | y0 << 16 and 16-bit ring fields; same limit as shadow maps / textures
        return fail(nullptr, REB_E_INVALID, "reb_create: extent must be in 1..16384");
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0) return fail(nullptr, REB_E_NO_DEVICE, "reb_create: no CUDA device (there is no CPU fallback)");
    if (params->device < 0 || params->device >= count) return fail(nullptr, REB_E_INVALID, "reb_create: device ordinal out of range");
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, params->device) != cudaSuccess) return fail(nullptr, REB_E_CUDA, "cudaGetDeviceProperties failed");
    if (prop.major != 10) return fail(nullptr, REB_E_NO_DEVICE, "reb_create: kernels are built for sm_100a only; device is not compute capability 10.x");
    ctx = new reb_ctx();
    *out = nullptr;
    auto bail = [&](int rc) { g_createError = ctx->err; reb_destroy(ctx); return rc; };
    ctx->device = params->device; ctx->numSMs = prop.multiProcessorCount;
    if (cudaSetDevice(ctx->device) != cudaSuccess) return bail(fail(ctx, REB_E_CUDA, "cudaSetDevice failed"));
    if (params->stream) ctx->stream = reinterpret_cast<cudaStream_t>(params->stream);
    else {
        if (cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess) return bail(fail(ctx, REB_E_CUDA, "cudaStreamCreate failed"));
        ctx->ownStream = true;
    }
    ctx->W = params->width; ctx->H = params->height;
    ctx->maxDesc = params->max_descriptors ? params->max_descriptors : 8192u;
    ctx->outputs = params->outputs ? params->outputs : (REB_OUT_COLOR_SRGB8 | REB_OUT_COLOR_F32 | REB_OUT_DEPTH_F32 | REB_OUT_TRIID_U32);
    ctx->nRanks = params->n_ranks ? params->n_ranks : 1u; ctx->rank = params->rank;
    ctx->bandH = params->band_h ? params->band_h : 64u;
    if (ctx->rank >= ctx->nRanks || ctx->bandH % TILE_H) return bail(fail(ctx, REB_E_INVALID, "reb_create: bad partition (rank < n_ranks, band_h multiple of 8)"));
    int rc;
    if ((rc = alloc_targets(ctx))) return bail(rc);
    if (cudaMalloc(reinterpret_cast<void**>(&ctx->dStatus), 6 * STATUS_WORDS * 4) != cudaSuccess || cudaMemset(ctx->dStatus, 0, 6 * STATUS_WORDS * 4) != cudaSuccess ||
        cudaMalloc(reinterpret_cast<void**>(&ctx->huge), 2 * HUGE_CAP * sizeof(LargeRec)) != cudaSuccess ||
        cudaHostAlloc(reinterpret_cast<void**>(&ctx->hStatus), 256, cudaHostAllocMapped) != cudaSuccess ||
        cudaHostGetDevicePointer(reinterpret_cast<void**>(&ctx->hStatusDev), ctx->hStatus, 0) != cudaSuccess)
        return bail(fail(ctx, REB_E_CUDA, "status buffers allocation failed"));
    std::memset(ctx->hStatus, 0, 256);
    ctx->listCap = 1u << 20;   // records of 48 B; grown on overflow (check_and_recover)
    if (const char* e = std::getenv("REB_TILE_LIST_CAP")) ctx->listCap = std::max(1024u, (uint32_t)std::strtoul(e, nullptr, 10));   // tests: force the regrow + replay path
    if (cudaMalloc(reinterpret_cast<void**>(&ctx->tileList), (size_t)ctx->listCap * sizeof(LargeRec)) != cudaSuccess) return bail(fail(ctx, REB_E_CUDA, "tile list allocation failed"));
    if (cudaStreamSynchronize(ctx->stream) != cudaSuccess) return bail(fail(ctx, REB_E_CUDA, "stream sync failed in reb_create"));
    *out = ctx;
    return REB_OK;
}

int reb_destroy(reb_ctx* ctx) {
    CHECK_CTX();
    cudaSetDevice(ctx->device);
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    void* ptrs[] = {ctx->pos4, ctx->nrm4, ctx->indices, ctx->transforms, ctx->nmats, ctx->materials, ctx->lights, ctx->shadowMap,
                    ctx->swap[0], ctx->swap[1], ctx->swap[2], ctx->outF32, ctx->outDepth, ctx->outTriId, ctx->vis, ctx->large, ctx->clipQueue, ctx->tileHead[0], ctx->tileHead[1],
                    ctx->tileList, ctx->dStatus, ctx->huge, ctx->pass[0].dDraws, ctx->pass[1].dDraws,
                    ctx->pass[0].vtxA, ctx->pass[1].vtxA, ctx->pass[0].vtxBC, ctx->pass[1].vtxBC,
                    ctx->pass[0].blockDraw, ctx->pass[1].blockDraw, ctx->pass[0].flatSlots, ctx->pass[1].flatSlots,
                    ctx->pass[0].drawBounds, ctx->pass[1].drawBounds, ctx->pass[0].drawVisible, ctx->pass[1].drawVisible,
                    ctx->pass[0].blockBox, ctx->pass[1].blockBox, ctx->pass[0].blockVisible, ctx->pass[1].blockVisible,
                    ctx->pass[2].dDraws, ctx->pass[2].vtxA, ctx->pass[2].vtxBC, ctx->pass[2].blockDraw, ctx->pass[2].flatSlots, ctx->pass[2].drawBounds,
                    ctx->pass[2].drawVisible, ctx->pass[2].blockBox, ctx->pass[2].blockVisible, ctx->pass[2].vtxUV,
                    ctx->pass[0].triMaterial, ctx->pass[1].triMaterial, ctx->pass[2].triMaterial, ctx->tlHead, ctx->tlNodes, ctx->tlCounters, ctx->tlHdr, ctx->tlDepth, ctx->tlLarge, ctx->msColour, ctx->tlSpan, ctx->tlFrags, ctx->tlCols,
                    ctx->pass[0].posS, ctx->pass[1].posS, ctx->pass[2].posS, ctx->pass[0].nrmS, ctx->pass[1].nrmS, ctx->pass[2].nrmS,
                    ctx->pass[0].triStream, ctx->pass[1].triStream, ctx->pass[2].triStream,
                    ctx->pass[0].blockRefs, ctx->pass[1].blockRefs, ctx->pass[2].blockRefs, ctx->pass[0].triBlockList, ctx->pass[1].triBlockList, ctx->pass[2].triBlockList,
                    ctx->msLarge[0], ctx->msLarge[1], ctx->msLarge[2], ctx->msClip[0], ctx->msClip[1], ctx->msClip[2], ctx->msHuge, ctx->msTileHead[0], ctx->msTileHead[1], ctx->msTileHead[2],
                    ctx->pass[0].cullCounts, ctx->pass[1].cullCounts, ctx->pass[2].cullCounts, ctx->pass[0].slotBlockVisible, ctx->pass[1].slotBlockVisible, ctx->pass[2].slotBlockVisible};
    for (void* p : ptrs) if (p) cudaFree(p);
    for (void* p : ctx->texAllocs) cudaFree(p);
    if (ctx->dTextures) cudaFree(ctx->dTextures);
    if (ctx->skyFaces) cudaFree(ctx->skyFaces);
    if (ctx->msDepth) cudaFree(ctx->msDepth);
    if (ctx->msTriId) cudaFree(ctx->msTriId);
    if (ctx->pass[1].vtxUV) cudaFree(ctx->pass[1].vtxUV);
    if (ctx->hStatus) cudaFreeHost(ctx->hStatus);
    if (ctx->copyStream) { cudaStreamSynchronize(ctx->copyStream); cudaStreamDestroy(ctx->copyStream); }
    if (ctx->evRendered) cudaEventDestroy(ctx->evRendered);
    for (int i = 0; i < SWAP_IMAGES; ++i) if (ctx->evCopied[i]) cudaEventDestroy(ctx->evCopied[i]);
    if (ctx->ownStream && ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
    return REB_OK;
}

int reb_upload_geometry(reb_ctx* ctx, const void* verts, uint64_t nVerts, const uint32_t* idx, uint64_t nIdx) {
    CHECK_CTX();
    if ((!verts && nVerts) || (!idx && nIdx)) return fail(ctx, REB_E_INVALID, "reb_upload_geometry: null buffer");
    if (nVerts >= 0xFFFFFFFFull || nIdx >= 0xFFFFFFFFull) return fail(ctx, REB_E_INVALID, "reb_upload_geometry: counts must fit u32");
    CU(cudaSetDevice(ctx->device));
    { int frc = flush_deferred(ctx); if (frc) return frc; }
    CU(cudaStreamSynchronize(ctx->stream));
    int rc;
    if ((rc = dev_alloc(ctx, &ctx->pos4, nVerts))) return rc;
    if ((rc = dev_alloc(ctx, &ctx->nrm4, nVerts))) return rc;
    if ((rc = dev_alloc(ctx, &ctx->indices, nIdx))) return rc;
    ctx->nVerts = nVerts; ctx->nIdx = nIdx;
    ctx->pass[0].valid = ctx->pass[1].valid = ctx->pass[2].valid = false;
    if (nVerts) {
        float4* aos = nullptr;
        CU(cudaMalloc(reinterpret_cast<void**>(&aos), nVerts * 32));
        CU(cudaMemcpyAsync(aos, verts, nVerts * 32, cudaMemcpyHostToDevice, ctx->stream));
        launch_split_vertices(aos, ctx->pos4, ctx->nrm4, nVerts, ctx->stream); ctx->launches++;
        CU(cudaStreamSynchronize(ctx->stream));
        CU(cudaFree(aos));
    }
    if (nIdx) {
        CU(cudaMemcpyAsync(ctx->indices, idx, nIdx * 4, cudaMemcpyHostToDevice, ctx->stream));
        CU(cudaStreamSynchronize(ctx->stream));
    }
    return REB_OK;
}

int reb_upload_transforms(reb_ctx* ctx, const float* m, uint32_t nNodes) {
    CHECK_CTX();
    if (!m && nNodes) return fail(ctx, REB_E_INVALID, "reb_upload_transforms: null buffer");
    CU(cudaSetDevice(ctx->device));
    { int frc = flush_deferred(ctx); if (frc) return frc; }
    CU(cudaStreamSynchronize(ctx->stream));
    int rc;
    if (nNodes != ctx->nNodes || !ctx->transforms) {
        if ((rc = dev_alloc(ctx, &ctx->transforms, (size_t)nNodes * 16))) return rc;
        if ((rc = dev_alloc(ctx, &ctx->nmats, (size_t)nNodes * 12))) return rc;
    }
    // draws are validated against the node count: a re-upload of the same number of nodes (moving objects) keeps the draw lists'
    // preprocessing — cache layout, tables, the re-ordering — which depends on geometry and draws only
    if (nNodes != ctx->nNodes) ctx->pass[0].valid = ctx->pass[1].valid = ctx->pass[2].valid = false;
    ctx->nNodes = nNodes;
    if (!nNodes) return REB_OK;
    std::vector<float> nm((size_t)nNodes * 12);
    for (uint32_t n = 0; n < nNodes; ++n) normal_matrix12(m + 16 * (size_t)n, &nm[12 * (size_t)n]);
    CU(cudaMemcpyAsync(ctx->transforms, m, (size_t)nNodes * 64, cudaMemcpyHostToDevice, ctx->stream));
    CU(cudaMemcpyAsync(ctx->nmats, nm.data(), nm.size() * 4, cudaMemcpyHostToDevice, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    return REB_OK;
}

static int upload_records(reb_ctx* ctx, uint8_t** dst, const void* records, uint32_t stride, uint32_t n, uint32_t minStride, const char* what) {
    if ((!records && n) || stride < minStride || stride % 16) return fail(ctx, REB_E_INVALID, std::string(what) + ": bad records / stride (16-B multiple required)");
    CU(cudaSetDevice(ctx->device));
    { int frc = flush_deferred(ctx); if (frc) return frc; }
    CU(cudaStreamSynchronize(ctx->stream));
    int rc;
    if ((rc = dev_alloc(ctx, dst, (size_t)stride * n))) return rc;
    if (n) {
        CU(cudaMemcpyAsync(*dst, records, (size_t)stride * n, cudaMemcpyHostToDevice, ctx->stream));
        CU(cudaStreamSynchronize(ctx->stream));
    }
    return REB_OK;
}

int reb_upload_materials(reb_ctx* ctx, const void* records, uint32_t stride, uint32_t n) {
    CHECK_CTX();
    int rc = upload_records(ctx, &ctx->materials, records, stride, n, 48, "reb_upload_materials");
    if (rc == REB_OK) { ctx->materialStride = stride; ctx->nMaterials = n; }
    return rc;
}

int reb_upload_lights(reb_ctx* ctx, const void* records, uint32_t stride, uint32_t n) {
    CHECK_CTX();
    int rc = upload_records(ctx, &ctx->lights, records, stride, n, 80, "reb_upload_lights");
    if (rc == REB_OK) {
        ctx->lightStride = stride; ctx->nLights = n;
        for (uint32_t i = 0; i < n && i < (uint32_t)SHADE_CONST_LIGHTS; ++i) std::memcpy(ctx->lightsHost[i], static_cast<const uint8_t*>(records) + (size_t)i * stride, std::min<uint32_t>(stride, 112u));
    }
    return rc;
}

int reb_register_texture(reb_ctx* ctx, const reb_texture_desc* desc, const void* texels, uint32_t* outHandle) {
    CHECK_CTX();
    if (!desc || !texels || !outHandle) return fail(ctx, REB_E_INVALID, "reb_register_texture: null argument");
    if (desc->format != 0) return fail(ctx, REB_E_UNSUPPORTED, "reb_register_texture: only format 0 (R16G16B16A16_UNORM, TextureCache.h:59) is supported");
    if (desc->width == 0 || desc->height == 0 || desc->width > 16384 || desc->height > 16384) return fail(ctx, REB_E_INVALID, "reb_register_texture: extent must be in 1..16384");
    if (ctx->texHost.size() >= ctx->maxDesc) return fail(ctx, REB_E_OVERFLOW, "reb_register_texture: descriptor table full (max_descriptors, Bindless.h:76)");
    CU(cudaSetDevice(ctx->device));
    TexDev t;
    std::memset(&t, 0, sizeof t);
    t.w = desc->width; t.h = desc->height; t.srgb = desc->srgb ? 1u : 0u;
    size_t total = 0;
    uint32_t w = t.w, h = t.h, lw[16], lh[16];
    for (;;) {   // floor(log2(max(w,h))) + 1 levels (TextureCache.cpp:67-73)
        lw[t.mips] = w; lh[t.mips] = h; t.off[t.mips] = (uint32_t)total; total += (size_t)w * h; ++t.mips;
        if ((w == 1 && h == 1) || t.mips == 16) break;
        w = std::max(1u, w / 2); h = std::max(1u, h / 2);
    }
    uint16_t* base = nullptr;
    CU(cudaMalloc(reinterpret_cast<void**>(&base), total * 8));
    cudaError_t e = cudaMemcpyAsync(base, texels, (size_t)t.w * t.h * 8, cudaMemcpyHostToDevice, ctx->stream);
    for (uint32_t l = 1; l < t.mips && e == cudaSuccess; ++l) {   // successive 2x linear blits (Texture.cpp:50-136)
        launch_mip_down(base + (size_t)t.off[l - 1] * 4, base + (size_t)t.off[l] * 4, lw[l - 1], lh[l - 1], lw[l], lh[l], ctx->stream); ctx->launches++;
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);   // the host buffer is consumed before returning
    if (e != cudaSuccess) { cudaFree(base); return fail(ctx, REB_E_CUDA, std::string("reb_register_texture: ") + cudaGetErrorString(e)); }
    t.base = base;
    ctx->texAllocs.push_back(base);
    ctx->texHost.push_back(t);
    ctx->texDirty = true;
    *outHandle = (uint32_t)ctx->texHost.size() - 1;
    return REB_OK;
}

int reb_shadow_create(reb_ctx* ctx, uint32_t size, uint32_t* outShadowId) {
    CHECK_CTX();
    if (size == 0 || size > 16384) return fail(ctx, REB_E_INVALID, "reb_shadow_create: size must be in 1..16384");
    if (ctx->shadowMap) return fail(ctx, REB_E_UNSUPPORTED, "reb_shadow_create: exactly one directional shadow map is supported (SURVEY F10)");
    CU(cudaSetDevice(ctx->device));
    int rc;
    if ((rc = dev_alloc(ctx, &ctx->shadowMap, (size_t)size * size))) return rc;
    ctx->S = size; ctx->shadowValid = false;
    launch_fill_u32(reinterpret_cast<uint32_t*>(ctx->shadowMap), 0x3F800000u, (size_t)size * size, ctx->numSMs, ctx->stream); ctx->launches++;
    if (outShadowId) *outShadowId = 0;
    return REB_OK;
}

int reb_shadow_set_transform(reb_ctx* ctx, uint32_t shadowId, const float* view16, const float* proj16) {
    CHECK_CTX();
    if (shadowId != 0 || !ctx->shadowMap || !view16 || !proj16) return fail(ctx, REB_E_INVALID, "reb_shadow_set_transform: unknown shadow id or null matrix");
    std::memcpy(ctx->sview, view16, 64); std::memcpy(ctx->sproj, proj16, 64);
    mat4_mul(ctx->sproj, ctx->sview, ctx->SPV);   // ShadowSystem.cpp:303 / shadow_map.vert:51
    ctx->haveShadowXf = true;
    return REB_OK;
}

int reb_set_view(reb_ctx* ctx, const void* vp) {
    CHECK_CTX();
    if (!vp) return fail(ctx, REB_E_INVALID, "reb_set_view: null");
    const float* f = static_cast<const float*>(vp);
    std::memcpy(ctx->view, f, 64); std::memcpy(ctx->proj, f + 16, 64); std::memcpy(ctx->viewPos, f + 32, 12);
    std::memcpy(&ctx->exposure, f + 35, 4); std::memcpy(&ctx->debugInput, f + 36, 4); std::memcpy(&ctx->debugEquation, f + 37, 4);   // view.glsl:1-9
    mat4_mul(ctx->proj, ctx->view, ctx->PV);      // primitive.vert:53
    ctx->haveView = true;
    return REB_OK;
}

static int ready_to_render(reb_ctx* ctx) {
    if (!ctx->pos4 || !ctx->indices) return fail(ctx, REB_E_INVALID, "render before reb_upload_geometry");
    if (!ctx->transforms) return fail(ctx, REB_E_INVALID, "render before reb_upload_transforms");
    return REB_OK;
}

int reb_render_shadow(reb_ctx* ctx, uint32_t shadowId, const reb_draw* draws, uint32_t n) {
    CHECK_CTX();
    if (shadowId != 0 || !ctx->shadowMap) return fail(ctx, REB_E_INVALID, "reb_render_shadow: unknown shadow id");
    if (!ctx->haveShadowXf) return fail(ctx, REB_E_INVALID, "reb_render_shadow before reb_shadow_set_transform");
    if (!draws && n) return fail(ctx, REB_E_INVALID, "reb_render_shadow: null draws");
    int rc;
    if ((rc = ready_to_render(ctx))) return rc;
    CU(cudaSetDevice(ctx->device));
    if ((rc = flush_deferred(ctx))) return rc;   // a previous recording that was never paired with a forward pass
    if ((rc = set_draws(ctx, 0, draws, n))) return rc;
    std::memcpy(ctx->recordedSPV, ctx->SPV, 64);
    ctx->shadowDeferred = true;                  // submitted by reb_render_forward / reb_flush / anything that needs the map
    ctx->shadowValid = true;
    return REB_OK;
}

static int render_base_pass(reb_ctx* ctx, const reb_draw* draws, uint32_t n, const reb_draw* translucent, uint32_t nTranslucent) {
    CHECK_CTX();
    if (!ctx->haveView) return fail(ctx, REB_E_INVALID, "reb_render_forward before reb_set_view");
    if (!draws && n) return fail(ctx, REB_E_INVALID, "reb_render_forward: null draws");
    int rc;
    if ((rc = ready_to_render(ctx))) return rc;
    if (!ctx->materials || !ctx->nMaterials) return fail(ctx, REB_E_INVALID, "reb_render_forward before reb_upload_materials");
    if (ctx->nLights && !ctx->lights) return fail(ctx, REB_E_INVALID, "reb_render_forward: lights missing");
    if (ctx->shadingModel == REB_SHADING_PBR && (ctx->materialStride < 80 || (ctx->nLights && ctx->lightStride < 80)))
        return fail(ctx, REB_E_INVALID, "reb_render_forward: the PBR model needs 80-B MaterialProperties and 80-B Light records");
    if (ctx->shadingModel == REB_SHADING_PHONG && ctx->nLights && ctx->lightStride < 112)
        return fail(ctx, REB_E_INVALID, "reb_render_forward: the Phong model needs 112-B Light records (phong.glsl:20-37)");
    CU(cudaSetDevice(ctx->device));
    if ((rc = set_draws(ctx, 1, draws, n))) return rc;
    ctx->tlFollows = nTranslucent > 0;
    if (!ctx->tlFollows) return submit_forward(ctx);
    // Appendix B.16: the translucent list of the same base pass (RenderScene.cpp:226-227)
    if (!translucent) return fail(ctx, REB_E_INVALID, "reb_render_forward2: null translucent draws");
    if ((rc = set_draws(ctx, 2, translucent, nTranslucent))) return rc;
    const size_t px = (size_t)ctx->W * ctx->H, spx = px * ctx->samples;   // one list head per sample
    if (ctx->tlPx != spx) {
        CU(cudaStreamSynchronize(ctx->stream));
        const uint32_t cap = (uint32_t)std::min<size_t>(std::max<size_t>(4 * spx, (size_t)1 << 20), (size_t)1 << 28);
        if ((rc = dev_alloc(ctx, &ctx->msColour, ctx->samples == 4 ? spx : 0))) return rc;
        if ((rc = dev_alloc(ctx, &ctx->tlHead, spx)) || (rc = dev_alloc(ctx, &ctx->tlNodes, (size_t)cap)) || (rc = dev_alloc(ctx, &ctx->tlCounters, (size_t)8)) ||
            (rc = dev_alloc(ctx, &ctx->tlHdr, ctx->outF32 ? 0 : px * 4)) || (rc = dev_alloc(ctx, &ctx->tlDepth, ctx->outDepth ? 0 : px))) return rc;
        launch_fill_u32(ctx->tlHead, 0xFFFFFFFFu, spx, ctx->numSMs, ctx->stream); ctx->launches++;
        // split resolve: span per pixel (zero = nothing), one record + one colour per passing fragment (per primitive and pixel with 4 samples)
        if ((rc = dev_alloc(ctx, &ctx->tlSpan, px)) || (rc = dev_alloc(ctx, &ctx->tlFrags, (size_t)cap)) || (rc = dev_alloc(ctx, &ctx->tlCols, (size_t)cap))) return rc;
        if (ctx->tlSpan) CU(cudaMemsetAsync(ctx->tlSpan, 0, px * sizeof(uint2), ctx->stream));
        ctx->tlCap = cap; ctx->tlPx = spx;
    }
    uint64_t nOpaque = 0;
    for (uint32_t i = 0; i < n; ++i) nOpaque += draws[i].nbIndices / 3;
    ctx->tlIdBase = (uint32_t)nOpaque;               // triangle IDs continue after the opaque list
    if ((rc = submit_forward(ctx))) { ctx->tlFollows = false; return rc; }
    rc = run_translucent(ctx);
    ctx->tlFollows = false;
    return rc;
}

int reb_render_forward(reb_ctx* ctx, const reb_draw* draws, uint32_t n) { return render_base_pass(ctx, draws, n, nullptr, 0); }

int reb_render_forward2(reb_ctx* ctx, const reb_draw* opaque, uint32_t nOpaque, const reb_draw* translucent, uint32_t nTranslucent) {
    return render_base_pass(ctx, opaque, nOpaque, translucent, nTranslucent);
}

int reb_set_shading_model(reb_ctx* ctx, uint32_t model) {
    CHECK_CTX();
    if (model != REB_SHADING_PHONG && model != REB_SHADING_PBR) return fail(ctx, REB_E_INVALID, "reb_set_shading_model: unknown model");
    ctx->shadingModel = model;
    return REB_OK;
}

int reb_set_skybox(reb_ctx* ctx, const void* facesRGBA16, uint32_t width, uint32_t height) {
    CHECK_CTX();
    CU(cudaSetDevice(ctx->device));
    { int rc = flush_deferred(ctx); if (rc) return rc; }
    if (!facesRGBA16) {                                   // remove the skybox: the background is the clear colour again
        if (ctx->skyFaces) { CU(cudaStreamSynchronize(ctx->stream)); cudaFree(ctx->skyFaces); }
        ctx->skyFaces = nullptr; ctx->skyW = ctx->skyH = 0;
        return REB_OK;
    }
    if (width == 0 || height == 0 || width > 16384 || height > 16384) return fail(ctx, REB_E_INVALID, "reb_set_skybox: face extent must be in 1..16384");
    const size_t bytes = (size_t)6 * width * height * 8;
    uint16_t* dev = nullptr;
    CU(cudaMalloc(reinterpret_cast<void**>(&dev), bytes));
    cudaError_t e = cudaMemcpyAsync(dev, facesRGBA16, bytes, cudaMemcpyHostToDevice, ctx->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);   // the host buffer is consumed before returning; earlier frames are done
    if (e != cudaSuccess) { cudaFree(dev); return fail(ctx, REB_E_CUDA, std::string("reb_set_skybox: ") + cudaGetErrorString(e)); }
    if (ctx->skyFaces) cudaFree(ctx->skyFaces);
    ctx->skyFaces = dev; ctx->skyW = width; ctx->skyH = height;
    return REB_OK;
}

int reb_set_sample_count(reb_ctx* ctx, uint32_t samples) {
    CHECK_CTX();
    if (samples != 1u && samples != 4u) return fail(ctx, REB_E_UNSUPPORTED, "reb_set_sample_count: 1 or 4 samples per pixel");
    if (samples == ctx->samples) return REB_OK;
    CU(cudaSetDevice(ctx->device));
    { int rc = flush_deferred(ctx); if (rc) return rc; }
    CU(cudaStreamSynchronize(ctx->stream));
    ctx->samples = samples;
    const size_t px = (size_t)ctx->W * ctx->H;
    int rc;
    if ((rc = dev_alloc(ctx, &ctx->vis, px * samples))) return rc;
    launch_fill_u32(reinterpret_cast<uint32_t*>(ctx->vis), 0xFFFFFFFFu, px * 2 * samples, ctx->numSMs, ctx->stream); ctx->launches++;
    if ((rc = dev_alloc(ctx, &ctx->msDepth, samples > 1 ? px * samples : 0))) return rc;
    if ((rc = dev_alloc(ctx, &ctx->msTriId, samples > 1 ? px * samples : 0))) return rc;
    return REB_OK;
}

int reb_flush(reb_ctx* ctx) {
    CHECK_CTX();
    CU(cudaSetDevice(ctx->device));
    return flush_deferred(ctx);
}

int reb_sync(reb_ctx* ctx) {
    CHECK_CTX();
    CU(cudaSetDevice(ctx->device));
    { int rc = flush_deferred(ctx); if (rc) return rc; }
    CU(cudaStreamSynchronize(ctx->stream));
    if (ctx->copyStream) CU(cudaStreamSynchronize(ctx->copyStream));
    for (bool& f : ctx->copyInFlight) f = false;
    return check_and_recover(ctx);
}

static int output_info(reb_ctx* ctx, reb_output_kind kind, void** p, uint64_t* bytes) {
    const uint64_t px = (uint64_t)ctx->W * ctx->H;
    switch (kind) {
        case REB_COLOR_SRGB8: *p = ctx->outSRGB8; *bytes = px * 4; break;
        case REB_COLOR_F32: *p = ctx->outF32; *bytes = px * 16; break;
        case REB_DEPTH_F32: *p = ctx->outDepth; *bytes = px * 4; break;
        case REB_TRIID_U32: *p = ctx->outTriId; *bytes = px * 4; break;
        case REB_SHADOW_F32: *p = ctx->shadowMap; *bytes = (uint64_t)ctx->S * ctx->S * 4; break;
        case REB_MS_DEPTH_F32: *p = ctx->msDepth; *bytes = px * 4 * ctx->samples; break;
        case REB_MS_TRIID_U32: *p = ctx->msTriId; *bytes = px * 4 * ctx->samples; break;
        default: return fail(ctx, REB_E_INVALID, "unknown output kind");
    }
    if (!*p) return fail(ctx, REB_E_INVALID, "requested output is not enabled on this context");
    return REB_OK;
}

int reb_get_output(reb_ctx* ctx, reb_output_kind kind, void** devPtr, uint64_t* bytes) {
    CHECK_CTX();
    if (ctx->shadowDeferred) { cudaSetDevice(ctx->device); int frc = flush_deferred(ctx); if (frc) return frc; }   // the caller may read the map on the stream
    void* p = nullptr; uint64_t b = 0;
    int rc = output_info(ctx, kind, &p, &b);
    if (rc) return rc;
    if (devPtr) *devPtr = p;
    if (bytes) *bytes = b;
    return REB_OK;
}

int reb_readback(reb_ctx* ctx, reb_output_kind kind, void* host) {
    CHECK_CTX();
    if (!host) return fail(ctx, REB_E_INVALID, "reb_readback: null host buffer");
    CU(cudaSetDevice(ctx->device));
    int rc;
    if (ctx->shadowDeferred || ctx->pass[0].pending || ctx->pass[1].pending) { if ((rc = reb_sync(ctx))) return rc; }
    void* p = nullptr; uint64_t b = 0;
    if ((rc = output_info(ctx, kind, &p, &b))) return rc;
    CU(cudaMemcpyAsync(host, p, b, cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    return REB_OK;
}

int reb_readback_async(reb_ctx* ctx, reb_output_kind kind, void* host) {
    CHECK_CTX();
    if (!host) return fail(ctx, REB_E_INVALID, "reb_readback_async: null host buffer");
    CU(cudaSetDevice(ctx->device));
    int rc;
    if ((rc = flush_deferred(ctx))) return rc;
    void* p = nullptr; uint64_t b = 0;
    if ((rc = output_info(ctx, kind, &p, &b))) return rc;
    if (kind != REB_COLOR_SRGB8) {   // single-buffered outputs: ordered copy on the render stream
        CU(cudaMemcpyAsync(host, p, b, cudaMemcpyDeviceToHost, ctx->stream));
        return REB_OK;
    }
    if (!ctx->copyStream) {
        CU(cudaStreamCreateWithFlags(&ctx->copyStream, cudaStreamNonBlocking));
        CU(cudaEventCreateWithFlags(&ctx->evRendered, cudaEventDisableTiming));
        for (int i = 0; i < SWAP_IMAGES; ++i) CU(cudaEventCreateWithFlags(&ctx->evCopied[i], cudaEventDisableTiming));
    }
    CU(cudaEventRecord(ctx->evRendered, ctx->stream));
    CU(cudaStreamWaitEvent(ctx->copyStream, ctx->evRendered, 0));
    CU(cudaMemcpyAsync(host, p, b, cudaMemcpyDeviceToHost, ctx->copyStream));
    CU(cudaEventRecord(ctx->evCopied[ctx->swapCur], ctx->copyStream));
    ctx->copyInFlight[ctx->swapCur] = true;
    ctx->asyncHost = host; ctx->asyncOutstanding = true;
    return REB_OK;
}

int reb_set_partition(reb_ctx* ctx, uint32_t rank, uint32_t nRanks, uint32_t bandH) {
    CHECK_CTX();
    if (nRanks == 0 || rank >= nRanks || bandH == 0 || bandH % TILE_H) return fail(ctx, REB_E_INVALID, "reb_set_partition: need rank < nRanks and bandH a multiple of 8");
    CU(cudaSetDevice(ctx->device));
    { int frc = flush_deferred(ctx); if (frc) return frc; }
    CU(cudaStreamSynchronize(ctx->stream));
    if (ctx->copyStream) CU(cudaStreamSynchronize(ctx->copyStream));   // an async readback may still be draining a swapchain image
    for (bool& f : ctx->copyInFlight) f = false; ctx->asyncOutstanding = false;
    ctx->rank = rank; ctx->nRanks = nRanks; ctx->bandH = bandH;
    return alloc_targets(ctx);   // re-clear: pixels outside this rank's bands keep the clear values
}

int reb_set_shadow_partition(reb_ctx* ctx, uint32_t rank, uint32_t nRanks, uint32_t bandH) {
    CHECK_CTX();
    if (nRanks == 0 || rank >= nRanks || bandH == 0 || bandH % 32u) return fail(ctx, REB_E_INVALID, "reb_set_shadow_partition: need rank < nRanks and bandH a multiple of 32");
    CU(cudaSetDevice(ctx->device));
    { int frc = flush_deferred(ctx); if (frc) return frc; }
    ctx->sRank = rank; ctx->sNRanks = nRanks; ctx->sBandH = bandH;
    ctx->sSplitRank = 0; ctx->sSplitN = 1;   // one split of the shadow pass at a time
    return REB_OK;
}

int reb_set_shadow_split(reb_ctx* ctx, uint32_t rank, uint32_t nRanks) {
    CHECK_CTX();
    if (nRanks == 0 || rank >= nRanks) return fail(ctx, REB_E_INVALID, "reb_set_shadow_split: need rank < nRanks");
    CU(cudaSetDevice(ctx->device));
    { int frc = flush_deferred(ctx); if (frc) return frc; }
    ctx->sSplitRank = rank; ctx->sSplitN = nRanks;
    ctx->sRank = 0; ctx->sNRanks = 1;
    return REB_OK;
}

int reb_shadow_mark_valid(reb_ctx* ctx, uint32_t shadowId) {
    CHECK_CTX();
    if (shadowId != 0 || !ctx->shadowMap) return fail(ctx, REB_E_INVALID, "reb_shadow_mark_valid: unknown shadow id");
    if (ctx->shadowDeferred) { cudaSetDevice(ctx->device); int frc = flush_deferred(ctx); if (frc) return frc; }
    ctx->shadowValid = true;
    return REB_OK;
}

int reb_shadow_wait_event(reb_ctx* ctx, uint32_t shadowId, void* cudaEvent) {
    CHECK_CTX();
    if (shadowId != 0 || !ctx->shadowMap || !cudaEvent) return fail(ctx, REB_E_INVALID, "reb_shadow_wait_event: unknown shadow id or null event");
    if (ctx->shadowDeferred) { cudaSetDevice(ctx->device); int frc = flush_deferred(ctx); if (frc) return frc; }
    ctx->shadowReadyEvent = reinterpret_cast<cudaEvent_t>(cudaEvent);
    ctx->shadowValid = true;
    return REB_OK;
}

int reb_get_counters(reb_ctx* ctx, uint64_t* c) {
    CHECK_CTX();
    if (!c) return REB_E_INVALID;
    for (int i = 0; i < 4; ++i) c[i] = ctx->counters[i];
    return REB_OK;
}

int reb_set_profiling(reb_ctx* ctx, int on) {
    CHECK_CTX();
    CU(cudaSetDevice(ctx->device));
    { int frc = flush_deferred(ctx); if (frc) return frc; }
    if (on && !ctx->ev[0][0])
        for (int k = 0; k < 2; ++k) for (int i = 0; i < 6; ++i) CU(cudaEventCreate(&ctx->ev[k][i]));
    ctx->profiling = on != 0;
    return REB_OK;
}

int reb_get_kernel_times(reb_ctx* ctx, int pass, float* ms6) {
    CHECK_CTX();
    if (pass < 0 || pass > 1 || !ms6 || !ctx->ev[0][0]) return fail(ctx, REB_E_INVALID, "reb_get_kernel_times: profiling not enabled / bad args");
    CU(cudaSetDevice(ctx->device));
    { int frc = flush_deferred(ctx); if (frc) return frc; }
    CU(cudaStreamSynchronize(ctx->stream));
    ms6[5] = 0.0f;
    for (int i = 0; i < 5; ++i) {
        float ms = 0.0f;
        if (cudaEventElapsedTime(&ms, ctx->ev[pass][i], ctx->ev[pass][i + 1]) != cudaSuccess) { ms = 0.0f; cudaGetLastError(); }
        ms6[i] = ms;
    }
    return REB_OK;
}

int reb_get_launch_count(reb_ctx* ctx, uint64_t* out) {
    CHECK_CTX();
    if (!out) return REB_E_INVALID;
    *out = ctx->launches;
    return REB_OK;
}

}  // extern "C"
