/*
 * reb.h — C ABI of the B200-native raster hot path (shadow depth pass + Phong/PCF forward pass).
 *
 * The reference (hugbed/RenderEngine) has no plugin / FFI layer: the boundary is a set of C++
 * member calls inside one process.  Each entry point below replaces the reference
 * call cited beside it; the C++ facade in renderengine_b200/host keeps the reference's class and
 * method names and forwards here.  All functions return 0 on success or a negative reb_status;
 * reb_last_error() gives the text (the reference throws std::runtime_error instead —
 * Source/VkRHI/RHI/RenderLoop.cpp:144, Source/Runtime/Renderer/TextureCache.cpp:68-70).
 *
 * Threading: a context is NOT thread-safe (reference: single-threaded frame loop,
 * RenderLoop.cpp:52-67; Device.h:12-13).  One context per GPU.  All work is asynchronous on the
 * context's CUDA stream; host input buffers are consumed (copied) before the call returns.
 * There is no CPU fallback: without a CUDA device reb_create fails.
 */
#ifndef REB_H
#define REB_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define REB_ABI_VERSION 1
#define REB_MAX_EXTENT 16384u   /* largest width / height of a render target, shadow map or texture */

typedef struct reb_ctx reb_ctx;

typedef enum reb_status {
    REB_OK = 0,
    REB_E_INVALID = -1,     /* bad argument / call order (std::invalid_argument in the reference) */
    REB_E_CUDA = -2,        /* CUDA runtime error */
    REB_E_NO_DEVICE = -3,   /* no usable sm_100 device */
    REB_E_OVERFLOW = -4,    /* an internal queue overflowed and could not be grown */
    REB_E_UNSUPPORTED = -5
} reb_status;

/* which buffers the forward pass writes (reb_params.outputs) */
#define REB_OUT_COLOR_SRGB8 1u   /* R8G8B8A8_SRGB target, Source/VkRHI/RHI/Swapchain.cpp:10-21 */
#define REB_OUT_COLOR_F32   2u   /* linear RGBA f32 before encode (HDR parity buffer) */
#define REB_OUT_DEPTH_F32   4u
#define REB_OUT_TRIID_U32   8u

typedef enum reb_output_kind {
    REB_COLOR_SRGB8 = 0, REB_COLOR_F32 = 1, REB_DEPTH_F32 = 2, REB_TRIID_U32 = 3, REB_SHADOW_F32 = 4,
    /* with reb_set_sample_count(ctx, 4): the four per-sample planes (sample-major, 4 x H x W) the resolve read */
    REB_MS_DEPTH_F32 = 5, REB_MS_TRIID_U32 = 6
} reb_output_kind;

/* Renderer ctor (Source/Runtime/Renderer/Renderer.cpp:38-51): runtime parameters that the
 * reference hard-codes (image extent Swapchain.cpp:200-237, Bindless.h:76 descriptor limit). */
typedef struct reb_params {
    uint32_t struct_size;      /* sizeof(reb_params) */
    int32_t  device;           /* CUDA ordinal */
    uint32_t width, height;    /* forward target extent */
    uint32_t max_descriptors;  /* 0 -> 8192; texture handle valid iff handle < max_descriptors (bindless.glsl:14) */
    uint32_t outputs;          /* REB_OUT_* mask; 0 -> all */
    uint32_t rank, n_ranks;    /* sort-first split: this context owns row bands b with b % n_ranks == rank */
    uint32_t band_h;           /* band height in pixels (multiple of 8); 0 -> 64 */
    uint32_t reserved;
    void*    stream;           /* cudaStream_t to run on, or NULL for a private stream */
} reb_params;

/* MeshDrawInfo (Source/Runtime/Renderer/MeshAllocator.h:26-39): one vkCmdDrawIndexed of the reference's
 * draw loops (MaterialSystem.cpp:94-100, ShadowSystem.cpp:358-370). materialRaw = MaterialHandle bits. */
typedef struct reb_draw { uint32_t indexOffset, nbIndices, materialRaw, sceneNode; } reb_draw;

/* TextureCache::LoadTexture (TextureCache.cpp:44-106): R16G16B16A16_UNORM, full mip chain, repeat, trilinear */
typedef struct reb_texture_desc { uint32_t width, height, format /* 0 = RGBA16_UNORM */, srgb; } reb_texture_desc;

int         reb_abi_version(void);
/* Renderer::Renderer / ~Renderer — Renderer.cpp:38-51 */
int         reb_create(const reb_params* params, reb_ctx** out);
int         reb_destroy(reb_ctx* ctx);
const char* reb_last_error(reb_ctx* ctx /* may be NULL: error of the last failed reb_create */);

/* MeshAllocator::AddVertex/AddIndex/UploadToGPU — MeshAllocator.h:45-61, MeshAllocator.cpp:5-35.
 * verts: 32-B AoS (pos@0 normal@12 uv@24); idx: absolute u32. Split into SoA streams on the device. */
int reb_upload_geometry(reb_ctx* ctx, const void* verts, uint64_t nVerts, const uint32_t* idx, uint64_t nIdx);
/* SceneTree::CreateNode/UploadToGPU — SceneTree.h:32, SceneTree.cpp:6-31 (column-major mat4 per node) */
int reb_upload_transforms(reb_ctx* ctx, const float* mat4ColMajor, uint32_t nNodes);
/* MaterialSystem::CreateMaterialInstance/UploadToGPU — MaterialSystem.cpp:49-59,68-86 (48-B Phong records) */
int reb_upload_materials(reb_ctx* ctx, const void* records, uint32_t stride, uint32_t n);
/* LightSystem::AddLight/UploadToGPU — LightSystem.h:47-55, LightSystem.cpp:17-34 (112-B std430 Phong Light) */
int reb_upload_lights(reb_ctx* ctx, const void* records, uint32_t stride, uint32_t n);
/* BindlessDescriptors::StoreTexture — Bindless.cpp:221-241: handle = running index */
int reb_register_texture(reb_ctx* ctx, const reb_texture_desc* desc, const void* texelsMip0, uint32_t* outHandle);
/* ShadowSystem::CreateShadowMap — ShadowSystem.cpp:229-238 */
int reb_shadow_create(reb_ctx* ctx, uint32_t size, uint32_t* outShadowId);
/* ShadowSystem::Update result (ShadowSystem.cpp:277-317): light view and proj (= kClip * ortho) */
int reb_shadow_set_transform(reb_ctx* ctx, uint32_t shadowId, const float* view16, const float* proj16);
/* CameraViewSystem::Update — CameraViewSystem.cpp:61-72: the 160-B ViewProperties block */
int reb_set_view(reb_ctx* ctx, const void* viewProperties160);

/* Which fragment path reb_render_forward runs.  REB_SHADING_PHONG (default): phong.glsl PhongLighting, 48-B Phong material
 * records, 112-B Phong Light records (the north star's path).  REB_SHADING_PBR: the reference's live fragment shader,
 * surface_pbr.frag:38-50 -> BRDF_Lighting pbr.glsl:368-451 — materials are the reference's 80-B MaterialProperties
 * (MaterialSystem.h:41-51: baseColor@0 emissive@16 f0@32 metallic@36 perceptualRoughness@40 ambientOcclusion@44
 * textures[5]@48 = baseColor, emissive, ORM, normals, AO), lights its 80-B Light (LightSystem.h:22-33 / pbr.glsl:288-299);
 * exposure / debugInput / debugEquation are read from the ViewProperties block (@140 / @144 / @148). */
#define REB_SHADING_PHONG 0u
#define REB_SHADING_PBR   1u
int reb_set_shading_model(reb_ctx* ctx, uint32_t model);

/* The whole base pass of RenderScene::RenderBasePass (RenderScene.cpp:210-242): the opaque list, then the translucent list
 * (AlphaMode::eBlend materials, MaterialSystem.cpp:42,133) in the order given — the caller sorts it back to front as
 * RenderScene::SortTranslucentMeshes does (RenderScene.cpp:127-159) — with cull none, depth test LessOrEqual + depth
 * write and src-alpha blending (GraphicsPipelineCache.cpp:305-315), then the skybox.  Triangle IDs of the translucent
 * list continue after the opaque list's.  reb_render_forward(ctx, d, n) == reb_render_forward2(ctx, d, n, NULL, 0).
 * Translucent triangles that cross the eye plane / guard band are clipped like opaque ones.
 * Limits of this version: a pixel keeps at most 16 translucent fragments and a frame max(4 * W * H, 2^20) of them;
 * fragments beyond either limit are dropped and counted in reb_get_counters()[0]. */
int reb_render_forward2(reb_ctx* ctx, const reb_draw* opaque, uint32_t nOpaque, const reb_draw* translucent, uint32_t nTranslucent);

/* Skybox::Skybox / Skybox::Render — Skybox.cpp:58-92,126-140; TextureCache::LoadCubeMapFaces — TextureCache.cpp:199-290;
 * skybox.vert:32-37, skybox.frag:28-31.  Installs the cube map the forward pass shows wherever the depth buffer still
 * holds 1.0 (the reference draws its cube last at depth 1.0 under LessOrEqual, RenderScene.cpp:226-229).
 * facesRGBA16: six width x height R16G16B16A16_UNORM images, contiguous, in Vulkan layer order +X -X +Y -Y +Z -Z (the
 * reference's right, left, top, bottom, front, back, Skybox.cpp:71-78); one mip, linear filter, seamless edges; the
 * texel values are written to the colour target as they are (no sRGB decode, skybox.frag:30).  The buffer is consumed
 * before the call returns.  NULL removes the skybox (background = clear colour (0,0,0,1)). */
int reb_set_skybox(reb_ctx* ctx, const void* facesRGBA16, uint32_t width, uint32_t height);

/* Multisampling of the base pass — Swapchain.cpp:110-123,212 (multisampled colour / depth attachments + resolve),
 * PhysicalDevice.cpp:172-186 (sample count), GraphicsPipelineCache.cpp:224,302.  samples = 1 (default; the 1-sample
 * parity contract) or 4: four samples per pixel at the Vulkan standard locations, depth-tested per sample, every
 * primitive shaded once per pixel at the pixel centre, average resolve in linear float before the sRGB encode.
 * The depth / triangle-ID outputs then hold sample 0; REB_MS_* expose all four planes.  The shadow pass stays 1 sample
 * (ShadowSystem.cpp:44). */
int reb_set_sample_count(reb_ctx* ctx, uint32_t samples);

/* ShadowSystem::Render(drawCommands) — ShadowSystem.cpp:319-376 */
int reb_render_shadow(reb_ctx* ctx, uint32_t shadowId, const reb_draw* draws, uint32_t n);
/* RenderScene::RenderBasePass -> MaterialSystem::Draw — RenderScene.cpp:210-242, MaterialSystem.cpp:88-101 */
int reb_render_forward(reb_ctx* ctx, const reb_draw* draws, uint32_t n);

/* Submit recorded work without waiting (the reference's CommandRingBuffer submit, RenderLoop.cpp:124-132).
 * reb_render_shadow only RECORDS the depth pass; it is submitted together with the next reb_render_forward (both
 * vertex stages then run as one kernel when the draw lists match), or by reb_flush / reb_sync / reb_readback /
 * reb_get_output.  Call reb_flush before consuming the shadow map on the stream yourself (e.g. an NCCL broadcast). */
int reb_flush(reb_ctx* ctx);

/* swapchain resolve / present (Swapchain.cpp:200-237, RenderLoop.cpp:124-132) become headless buffers */
int reb_get_output(reb_ctx* ctx, reb_output_kind kind, void** devPtr, uint64_t* bytes);
int reb_readback(reb_ctx* ctx, reb_output_kind kind, void* host);
int reb_sync(reb_ctx* ctx);
/* Present-style readback: enqueue the device->host copy of an output and return at once; the data is valid after the
 * next reb_sync.  REB_COLOR_SRGB8 is a three-image swapchain (Swapchain.cpp:55-59,200-237): its copy runs on a copy stream
 * while the next frames render into the other images.  host should be pinned memory for the copy to overlap. */
int reb_readback_async(reb_ctx* ctx, reb_output_kind kind, void* host);

/* multi-GPU helpers */
int reb_set_partition(reb_ctx* ctx, uint32_t rank, uint32_t nRanks, uint32_t bandH);
/* Row split of the shadow pass: this rank rasterises only the shadow-map rows ((y / bandH) % nRanks == rank) (bandH a
 * multiple of 32) and skips draws that cannot touch them; the other rows keep the clear value until the caller fills them
 * (e.g. an in-place NCCL all-gather of the slabs on the buffer reb_get_output(REB_SHADOW_F32) returns, after reb_flush)
 * and calls reb_shadow_mark_valid.  With reb_set_partition (or reb_params.rank / n_ranks) the forward pass likewise skips
 * draws whose projected bounds touch none of this rank's rows. */
int reb_set_shadow_partition(reb_ctx* ctx, uint32_t rank, uint32_t nRanks, uint32_t bandH);
/* Sort-LAST split of the shadow pass, the alternative to the row slabs above: this context rasterises the rank-th of nRanks
 * contiguous ranges of the draw list's triangle stream — equal shares of the work whatever the light sees where, nothing
 * culled and nothing rasterised twice at slab borders — into the WHOLE map (cleared to 1.0 first).  The complete map is the
 * element-wise minimum of the ranks' maps: after reb_flush, ncclAllReduce(ncclMin, ncclFloat) in place on the buffer
 * reb_get_output(REB_SHADOW_F32) returns (depths are non-negative floats), then reb_shadow_mark_valid / reb_shadow_wait_event.
 * 4 S^2 bytes are reduced per frame instead of 4 S^2 / n gathered.  Replaces a reb_set_shadow_partition; (0, 1) turns the split off.  No counterpart in the reference. */
int reb_set_shadow_split(reb_ctx* ctx, uint32_t rank, uint32_t nRanks);

/* marks the shadow map as valid after its storage (reb_get_output REB_SHADOW_F32) was filled externally,
 * e.g. by an NCCL broadcast from the rank that rendered it */
int reb_shadow_mark_valid(reb_ctx* ctx, uint32_t shadowId);
/* The same for an exchange that is still IN FLIGHT on another stream: cudaEvent (a cudaEvent_t the caller recorded behind the
 * NCCL broadcast / all-gather that fills the map) is waited for by the next forward pass right before its tile stage — the only
 * kernel that reads the map (shadow.glsl:24-56 runs in the fragment stage) — so that pass's vertex stage and triangle setup
 * overlap the transfer.  Marks the map valid; one-shot (consumed by the next reb_render_forward*).  The reference has no
 * counterpart (single GPU; it renders the map once, RenderScene.cpp:171-175, and samples it every frame). */
int reb_shadow_wait_event(reb_ctx* ctx, uint32_t shadowId, void* cudaEvent);

/* Shadow depth pass of ONE draw that ALLOCATES NOTHING (no cudaMalloc / cudaFree, no context): for callers under CUDA graph
 * capture, tight serving loops or harnesses that forbid allocation.  No counterpart in the reference.
 * Same raster contract as reb_render_shadow (bit-exact on the same input): the map is cleared to 1.0, then
 * every triangle (dIndices[3t..3t+2] into dVertices; position = the first three floats of each vertexStride-byte record, so the
 * reference's 32-B Vertex buffer fits as is) goes through model16 -> shadowProj16 * shadowView16 -> viewport -> raster.
 * dVertices / dIndices / dDepthOut (size * size floats) are DEVICE pointers on the current device; the three matrices are host
 * pointers (column-major); cudaStream is a cudaStream_t (NULL = default stream); the work is asynchronous on it.
 * Its scratch (post-transform vertex cache + large-triangle queue) is one static device array of 6 MiB + 256 B that every GPU
 * loading this library holds until exit; inputs beyond REB_STATIC_MAX_VERTICES / REB_STATIC_MAX_TRIANGLES are rejected with
 * REB_E_OVERFLOW.  All calls on a device share that array: calls that overlap on the GPU race — keep them on one stream.
 * errOut (may be NULL) receives a static string describing a failure. */
#define REB_STATIC_MAX_VERTICES  262144u
#define REB_STATIC_MAX_TRIANGLES 524288u
int reb_shadow_depth_static(const void* dVertices, uint32_t vertexStride, uint32_t nVertices, const uint32_t* dIndices, uint32_t nIndices,
                            const float* model16, const float* shadowView16, const float* shadowProj16, uint32_t size, float* dDepthOut,
                            void* cudaStream, const char** errOut);

/* diagnostics: counters[0] clip children that could not be set up (non-finite input) + translucent fragments dropped by a
 * capacity limit — 0 for every finite scene within the limits above (triangles crossing the eye plane are CLIPPED, not dropped),
 * [1] triangles sent to the binned tile path by the passes of the last frame, [2] their tile-list entries, [3] queue regrow count */
int reb_get_counters(reb_ctx* ctx, uint64_t* counters4);
/* number of kernels this library has launched on the context since creation */
int reb_get_launch_count(reb_ctx* ctx, uint64_t* out);

/* per-kernel device timing of the last enqueued pass (0 = shadow, 1 = forward), CUDA events on the
 * context stream: ms6 = { clear, vertex stage, setup + small raster, bin scatter, tile stage, 0 } */
int reb_set_profiling(reb_ctx* ctx, int on);
int reb_get_kernel_times(reb_ctx* ctx, int pass, float* ms6);

#ifdef __cplusplus
}
#endif
#endif /* REB_H */
