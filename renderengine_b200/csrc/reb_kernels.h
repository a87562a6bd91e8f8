// reb_kernels.h — host-visible launch wrappers of the raster kernels (internal to libreb.so).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <cstdlib>
#include "reb_device.cuh"

namespace reb {

constexpr int SHADE_CONST_LIGHTS = 4;   // Light records (112 B std430, phong.glsl:20-37) carried in the kernel parameters

struct ShadeConst {
    float STR[16];                 // shadowProj*shadowView (ShadowSystem.cpp:303)
    float viewPos[3]; uint32_t nLights;
    const uint8_t* materials; uint32_t materialStride; uint32_t nMaterials;
    const uint8_t* lights; uint32_t lightStride;
    float4 lightsC[SHADE_CONST_LIGHTS][7];   // copy of the first records of `lights`
    const float* shadowMap; uint32_t shadowSize;   // NULL / 0: shadow = 0
    uint32_t shadingModel;                         // 0 = Phong (phong.glsl), 1 = PBR (pbr.glsl, the reference's live fragment path)
    float exposure; uint32_t debugInput, debugEquation;   // View (view.glsl:1-9), read by the PBR path only
    uint32_t maxDescriptors; uint32_t nTextures;
    const TexDev* textures;
    uint8_t* outSRGB8; float* outF32; float* outDepth; uint32_t* outTriId;
    // skybox (reb_sky.cuh): six RGBA16 faces, A^-1 of the view-ray map (column-major 3x3), half extents and their reciprocals
    const uint16_t* skyFaces; uint32_t skyW, skyH;
    float skyAinv[9]; float skyHw, skyHh, skyRhw, skyRhh;
};

// <<<>>> with the programmatic-stream-serialization attribute (PDL) when `pdl` is set
template <class... KArgs, class... Args>
inline void launch_k(void (*kernel)(KArgs...), dim3 grid, dim3 block, cudaStream_t st, bool pdl, Args&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = 0; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    static const bool noPdl = std::getenv("REB_NO_PDL") != nullptr;   // A/B + debugging switch
    cfg.attrs = at; cfg.numAttrs = (pdl && !noPdl) ? 1u : 0u;
    cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

// one-pass K1 of the 4x multisampled base pass (k_setup.cu): per sample, where K1 puts what that sample's K2 / tile stage consume
struct MsTargets {
    unsigned long long* vis[4];
    uint32_t* status[4]; LargeRec* large[4]; LargeRec* huge[4]; TileHead* tileHead[4]; uint32_t* clipQueue[4];
};
void launch_setup_raster_ms(const PassConst& pc, const MsTargets& ms, int numSMs, cudaStream_t st);

void launch_vertex(const PassConst& pc, bool shadow, bool pdl, cudaStream_t st);
void launch_vertex_both(const PassConst& shadowPass, const PassConst& forwardPass, cudaStream_t st);   // identical cache layouts
void launch_mip_down(const void* src, void* dst, uint32_t sw, uint32_t sh, uint32_t dw, uint32_t dh, cudaStream_t st);
void launch_draw_ranges(const uint32_t* indices, const DrawDev* draws, uint32_t nDraws, uint32_t* ranges, cudaStream_t st);
void launch_draw_bounds(const uint32_t* indices, const float4* pos4, const DrawDev* draws, uint32_t nDraws, uint32_t* bounds, cudaStream_t st);
// sort-first split, one launch per pass and frame: per-draw + per-256-triangle-block row-band culling, the compacted list of visible
// blocks for K1 (its length in *count; *nextCount is zeroed for the next frame) and the frame-stamped slot-block marks for K0
void launch_cull(const PassConst& pc, bool shadow, const uint32_t* bounds, uint32_t* drawVisible, const float4* blockBox, uint32_t nBlocks, uint32_t* visible,
                 const uint32_t* refs, uint32_t* slotBlockVisible, uint32_t* list, uint32_t* count, uint32_t* nextCount, cudaStream_t st);
void launch_block_bounds(const PassConst& pc, float4* blockBox, cudaStream_t st);   // per 256-triangle block, once per draw list
void launch_block_slot_refs(const PassConst& pc, uint32_t* refs, cudaStream_t st);   // once per draw list: slot blocks referenced by every 256-triangle block
void launch_setup_raster(const PassConst& pc, bool shadow, int numSMs, cudaStream_t st);
void launch_flatten_slots(const PassConst& pc, uint32_t* flat, uint32_t* triMaterial, cudaStream_t st);
void launch_fill_u32(uint32_t* p, uint32_t v, size_t n, int numSMs, cudaStream_t st);
void launch_fill_u32x4(void* p, uint32_t a, uint32_t b, uint32_t c, uint32_t d, size_t n4, int numSMs, cudaStream_t st);
void launch_split_vertices(const void* aos, float4* pos4, float4* nrm4, size_t nVerts, cudaStream_t st);
void launch_tile_binned(const PassConst& pc, bool shadow, int numSMs, cudaStream_t st);   // K2: clip + list layout + record scatter + TMA-staged tile raster of the binned triangles
void launch_tile_shadow(const PassConst& pc, int numSMs, cudaStream_t st);
void launch_tile_forward(const PassConst& pc, const ShadeConst& sh, int numSMs, cudaStream_t st);
// Appendix B.16 (k_translucent.cu): per-pixel fragment lists of the translucent triangles, then the ordered blend
void launch_tl_raster(const PassConst& pc, const float* opaqueDepth, uint32_t* head, uint4* nodes, uint32_t cap, uint32_t* counters, uint2* largeList, int numSMs,
                      cudaStream_t st);
void launch_tl_resolve(const PassConst& pc, const ShadeConst& sh, uint32_t* head, const uint4* nodes, uint32_t cap, uint32_t idBase, float* hdr, float* depth,
                       uint32_t* counters, cudaStream_t st);
// the 1-sample resolve as three kernels: gather (order + depth replay -> passing fragments), shade (one thread per fragment), blend
void launch_tl_resolve_split(const PassConst& pc, const ShadeConst& sh, uint32_t* head, const uint4* nodes, uint32_t cap, uint32_t idBase, float* hdr, float* depth,
                             uint32_t* counters, uint2* span, uint2* frags, float4* cols, int numSMs, cudaStream_t st);
void launch_tl_resolve_ms_split(const PassConst& pc, const ShadeConst& sh, uint32_t* head, const uint4* nodes, uint32_t cap, uint32_t idBase, const float4* msColour,
                                float* msDepth, uint32_t* msTriId, uint32_t* counters, uint2* span, uint2* frags, float4* cols, int numSMs, cudaStream_t st);
// the same inside a 4x multisampled pass: merges the four per-sample lists of a pixel over the per-sample colours / depths
void launch_tl_resolve_ms(const PassConst& pc, const ShadeConst& sh, uint32_t* head, const uint4* nodes, uint32_t cap, uint32_t idBase, const float4* msColour,
                          float* msDepth, uint32_t* msTriId, cudaStream_t st);
// k_reorder.cu: Morton order of a draw list's triangles + first-use order of its cache slots, once per draw list (synchronises st)
cudaError_t reorder_draw_list(const PassConst& pc, const uint32_t* drawBounds, float4* posS, float4* nrmS, uint4* triStream, uint32_t* flat, cudaStream_t st);
void launch_pass_reset(const PassConst& pc, int numSMs, cudaStream_t st);   // between the sample passes of a multisampled frame
// Appendix B.15 resolve: per pixel, shade every distinct primitive of the 4 sample planes once at the pixel centre, average
void launch_msaa_resolve(const PassConst& pc, const ShadeConst& sh, const float* msDepth, const uint32_t* msTriId, float4* msColour, cudaStream_t st);   // msColour != NULL: leave per-sample colours for the translucent list

}  // namespace reb
