/*
 * oracle.h — C interface of the CPU oracle (TEST INFRASTRUCTURE, NOT PRODUCT).
 *
 * This is a plain C++ restatement of the reference's raster hot path (shadow depth pass +
 * Phong/PCF forward pass) used ONLY as the checker by tests/, __graft_entry__.smoke() and the
 * cpu_baseline / --impl reference legs of bench.py.  Nothing under renderengine_b200/ may call it.
 *
 * PARITY STATUS: "parity unpinned" at the rasteriser boundary.  The reference holds no tests,
 * golden vectors or CPU rasteriser for this path and a Vulkan
 * implementation cannot run on this box.  What IS pinned: the shading arithmetic
 * (phong.glsl / shadow.glsl) and the camera / bounding-box code are checked against the
 * reference's own sources compiled in oracle/_ref (see oracle/Makefile, tests/golden/).
 */
#ifndef ORACLE_H
#define ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct orc_texture {
    const uint16_t* texels;   /* mip 0, RGBA16 UNORM, row-major */
    uint32_t width, height;
    uint32_t srgb;            /* decode with the sRGB EOTF after filtering (pbr.glsl:20-26,104-105) */
    uint32_t pad;
} orc_texture;

typedef struct orc_scene {
    const float*    verts;        /* 32-B AoS: pos@0 normal@12 uv@24 (MeshAllocator.h:14-24) */
    uint64_t        nVerts;
    const uint32_t* indices;      /* absolute u32 (MeshAllocator.cpp:42) */
    uint64_t        nIndices;
    const float*    transforms;   /* column-major mat4 per node (SceneTree.h:69) */
    uint32_t        nNodes;
    uint32_t        nDraws;
    const uint32_t* draws;        /* 4 x u32: indexOffset, nbIndices, materialRaw, sceneNode (MeshAllocator.h:26-39) */
    const uint8_t*  materials;    /* 48-B Phong records: diffuse@0 specular@16 shininess@32 diffuseTex@36 specularTex@40 */
    uint32_t        materialStride;
    uint32_t        nMaterials;
    const uint8_t*  lights;       /* 112-B std430 Phong Light (phong.glsl:20-37) */
    uint32_t        lightStride;
    uint32_t        nLights;
    const uint8_t*  view;         /* 160-B ViewProperties (ViewProperties.h:27-35) */
    uint32_t        width, height;
    uint32_t        shadowSize;   /* S; 0 = no shadow map */
    uint32_t        maxDescriptors;
    const float*    shadowView;   /* mat4 */
    const float*    shadowProj;   /* mat4 (already kClip * ortho) */
    const orc_texture* textures;
    uint32_t        nTextures;
    uint32_t        nThreads;     /* 0 = hardware_concurrency */
    /* sort-first tile split: only pixels with ((y / bandH) % nRanks) == rank are owned.
       nRanks <= 1 disables. */
    uint32_t        rank, nRanks, bandH;
    uint32_t        shadingModel; /* 0 = Phong (48-B materials, 112-B lights), 1 = PBR (80-B MaterialProperties, 80-B Light: pbr.glsl) */
    /* skybox (Skybox.cpp:58-92, skybox.vert:32-37, skybox.frag:28-31): cube map of 6 RGBA16 UNORM faces, one mip
       (TextureCache.cpp:251-266), layers +X -X +Y -Y +Z -Z, each skyW x skyH row-major.  NULL = no skybox. */
    const uint16_t* skyFaces;
    uint32_t        skyW, skyH;
    /* translucent meshes (AlphaMode::eBlend, MaterialSystem.cpp:42,133): drawn after the opaque list in the same base pass
       (RenderScene.cpp:226-227), in the order given (the caller sorts them back to front, RenderScene.cpp:127-159), cull
       none, depth test LessOrEqual + depth write, src-alpha blending (GraphicsPipelineCache.cpp:305-315).  Same 4 x u32
       records as `draws`; triangle IDs continue after the opaque list's.  Appendix B.16. */
    const uint32_t* translucentDraws;
    uint32_t        nTranslucentDraws;
} orc_scene;

/* counters[0] = triangles dropped because they need true clipping (w<=0 or guard band),
   counters[1] = front-facing triangles rasterised, counters[2] = fragments passing coverage */
int orc_render_shadow(const orc_scene* s, float* shadowMap /* S*S */, uint64_t* counters /* 4 or NULL */);
int orc_render_forward(const orc_scene* s, const float* shadowMap /* S*S or NULL */,
                       float* depth, uint32_t* triId, float* colorF32 /* RGBA or NULL */,
                       uint8_t* colorSRGB8 /* RGBA or NULL */, uint64_t* counters);

/* Appendix B.15: the base pass with 4 samples per pixel at the Vulkan standard locations and an average resolve (the
   reference's actual base-pass state, Swapchain.cpp:110-123,212; PhysicalDevice.cpp:172-186).  depth4 / triId4 hold
   four W x H planes (sample-major); the colour outputs are the resolved image. */
int orc_render_forward_msaa4(const orc_scene* s, const float* shadowMap, float* depth4, uint32_t* triId4,
                             float* colorF32, uint8_t* colorSRGB8, uint64_t* counters);

/* unit-level entry points used to pin the restatement against oracle/_ref */
void orc_mat4_mul(const float* a, const float* b, float* out);
void orc_mat4_mul_vec4(const float* m, const float* v, float* out);
void orc_normal_matrix(const float* mat4, float* out9);
/* the composed vertex stages as the renderer calls them: primitive.vert:49-57 -> clip(4) world(3) normal(3); shadow_map.vert:48-52 -> clip */
void orc_primitive_vert(const float* transform16, const float* view16, const float* proj16, const float* pos3, const float* nrm3, float* out10);
void orc_shadow_vert(const float* transform16, const float* view16, const float* proj16, const float* pos3, float* out4);
void orc_phong_lighting(const uint8_t* light112, const float* props9 /* diffuse4 specular4 shininess */,
                        const float* normal, const float* fragPos, const float* viewPos, float shadow, float* out4);
float orc_compute_shadow(const uint8_t* light112, const float* fragPos, const float* normal,
                         const float* transform16, const float* shadowMap, uint32_t S);
void orc_srgb8(const float* rgba, uint8_t* out4);
/* Appendix B.14: skybox colour of pixel (x, y) for the 160-B view block; and the cube lookup on its own */
void orc_sky_pixel(const uint8_t* view160, uint32_t W, uint32_t H, const uint16_t* faces, uint32_t w, uint32_t h, uint32_t x, uint32_t y, float* out4);
void orc_sky_lookup(const uint16_t* faces, uint32_t w, uint32_t h, const float* dir3, float* out4);
void orc_pbr_lighting(const uint8_t* material80, const uint8_t* lights80, uint32_t nLights, const float* fragPos, const float* normal, const float* viewPos,
                      float exposure, uint32_t debugInput, uint32_t debugEquation, const float* transform16, const float* shadowMap, uint32_t S, float* out4);

#ifdef __cplusplus
}
#endif
#endif
