/*
 * ref_shim.cpp — C wrapper that exposes the REFERENCE's own code for pinning the oracle.
 * TEST INFRASTRUCTURE ONLY.  Compiled by oracle/Makefile against sources under /root/reference
 * (never copied): Renderer/Camera.{h,cpp}, BoundingBox.{h,cpp}, glm_includes.h, vendored glm, and
 * phong.glsl / shadow.glsl after the reversible literal/swizzle rewrite.
 * The ~20 lines of GLSL environment below (RegisterBuffer, GetResource, uGlobalTextures2D, texture,
 * textureSize) are ours; the nearest/clamp sampler follows ShadowSystem.cpp:18-35.
 */
#include <glm_includes.h>
#include <glm/gtc/matrix_inverse.hpp>
#include <Renderer/Camera.h>
#include <BoundingBox.h>
#include <cstring>
#include <cstdint>
#include <cmath>

namespace glsl {
using namespace glm;
typedef unsigned int uint;

struct Sampler2D { const float* data; uint32_t S; };
static Sampler2D uGlobalTextures2D[4];

inline vec4 texture(const Sampler2D& s, vec2 uv) {
    float fs = (float)s.S;
    float fu = std::floor(uv.x * fs), fv = std::floor(uv.y * fs);
    float hi = (float)(s.S - 1);
    int i = (fu >= 0.0f) ? ((fu <= hi) ? (int)fu : (int)(s.S - 1)) : 0;
    int j = (fv >= 0.0f) ? ((fv <= hi) ? (int)fv : (int)(s.S - 1)) : 0;
    return vec4(s.data[(size_t)j * s.S + i], 0.0f, 0.0f, 1.0f);
}
inline vec2 textureSize(const Sampler2D& s, int) { return vec2((float)s.S, (float)s.S); }

#define RegisterBuffer(layout_, access_, Name, body) struct Name body; static Name* g_##Name = nullptr
#define GetResource(Name, handle) (*g_##Name)

#include "phong.inc"
#include "shadow.inc"
}  // namespace glsl

// The reference's LIVE fragment path (surface_pbr.frag:38-50 -> pbr.glsl BRDF_Lighting) in its own namespace: pbr.glsl
// declares its own Light / LIGHT_TYPE_* and re-includes shadow.glsl for that Light.  Untextured materials only here
// (handles >= MAX_DESCRIPTOR_COUNT), so texture() of a material map and dFdx/dFdy are never reached.
namespace glsl_pbr {
using namespace glm;
typedef unsigned int uint;
#define MAX_DESCRIPTOR_COUNT 1024
struct View { mat4 view; mat4 proj; vec3 pos; float exposure; uint debugInput; uint debugEquation; };   // view.glsl:1-9
using glsl::Sampler2D;
using glsl::uGlobalTextures2D;
using glsl::texture;
using glsl::textureSize;
#define GetTexture2D(index) uGlobalTextures2D[(index) & 3u]
inline vec3 dFdx(vec3) { return vec3(0.0f); }
inline vec3 dFdy(vec3) { return vec3(0.0f); }
inline vec2 dFdx(vec2) { return vec2(0.0f); }
inline vec2 dFdy(vec2) { return vec2(0.0f); }
#include "pbr.inc"
}  // namespace glsl_pbr

extern "C" {

// BRDF_Lighting (pbr.glsl:368-451) for one untextured material record (80 B, MaterialSystem.h:41-51) and nLights
// 80-B std430 Light records (pbr.glsl:288-299); shadowMap may be NULL (then S = 0 and directional lights get a map of 1.0).
void ref_pbr_lighting(const uint8_t* material80, const uint8_t* lights80, uint32_t nLights, const float* fragPos, const float* uv,
                      const float* normal, const float* viewPos, float exposure, uint32_t debugInput, uint32_t debugEquation,
                      const float* transform16, const float* shadowMap, uint32_t S, float* out4) {
    static glsl_pbr::Material mats[1];
    static glsl_pbr::Light lts[16];
    static glsl_pbr::MaterialShadowData storage[1];
    static const float one = 1.0f;
    std::memcpy(&mats[0], material80, 80);
    for (uint32_t i = 0; i < nLights && i < 16; ++i) {
        const uint8_t* p = lights80 + 80 * (size_t)i;
        std::memcpy(&lts[i].color, p + 0, 12); std::memcpy(&lts[i].position, p + 16, 12); std::memcpy(&lts[i].direction, p + 32, 12);
        std::memcpy(&lts[i].intensity, p + 44, 4); std::memcpy(&lts[i].falloffRadius, p + 48, 4);
        std::memcpy(&lts[i].cosInnerAngle, p + 52, 4); std::memcpy(&lts[i].cosOuterAngle, p + 56, 4);
        std::memcpy(&lts[i].shadowIndex, p + 60, 4); std::memcpy(&lts[i].type, p + 64, 4);
        lts[i].shadowIndex = 0;
    }
    storage[0].transform = glm::make_mat4(transform16);
    storage[0].shadowMapTextureHandle = 0;
    glsl_pbr::g_MaterialBuffer = reinterpret_cast<glsl_pbr::MaterialBuffer*>(mats);
    glsl_pbr::g_LightBuffer = reinterpret_cast<glsl_pbr::LightBuffer*>(lts);
    glsl_pbr::g_MaterialShadowDataBuffer = reinterpret_cast<glsl_pbr::MaterialShadowDataBuffer*>(storage);
    if (shadowMap && S) glsl::uGlobalTextures2D[0] = {shadowMap, S}; else glsl::uGlobalTextures2D[0] = {&one, 1};
    glsl_pbr::View view;
    view.exposure = exposure; view.debugInput = debugInput; view.debugEquation = debugEquation;
    glm::vec4 r = glsl_pbr::BRDF_Lighting(glm::make_vec3(fragPos), glm::make_vec2(uv), glm::make_vec3(normal), glm::make_vec3(viewPos),
                                          0u, 0u, 0u, nLights, 0u, view);
    std::memcpy(out4, glm::value_ptr(r), 16);
}

void ref_mat4_mul(const float* a, const float* b, float* out) {
    glm::mat4 r = glm::make_mat4(a) * glm::make_mat4(b);
    std::memcpy(out, glm::value_ptr(r), 64);
}
void ref_mat4_mul_vec4(const float* m, const float* v, float* out) {
    glm::vec4 r = glm::make_mat4(m) * glm::make_vec4(v);
    std::memcpy(out, glm::value_ptr(r), 16);
}
// primitive.vert:55 — transpose(inverse(mat3(transform)))
void ref_normal_matrix(const float* m, float* out9) {
    glm::mat3 n = glm::transpose(glm::inverse(glm::mat3(glm::make_mat4(m))));
    std::memcpy(out9, glm::value_ptr(n), 36);
}
// primitive.vert:50-55 with glm; out = clip(4) world(3) normal(3)
void ref_primitive_vert(const float* transform, const float* view, const float* proj, const float* pos, const float* nrm, float* out10) {
    glm::mat4 T = glm::make_mat4(transform);
    glm::vec4 p = T * glm::vec4(pos[0], pos[1], pos[2], 1.0f);
    glm::vec3 fragPos = glm::vec3(p) / p.w;
    glm::vec4 clip = glm::make_mat4(proj) * glm::make_mat4(view) * glm::vec4(fragPos, 1.0f);
    glm::vec3 n = glm::normalize(glm::transpose(glm::inverse(glm::mat3(T))) * glm::vec3(nrm[0], nrm[1], nrm[2]));
    std::memcpy(out10, glm::value_ptr(clip), 16);
    std::memcpy(out10 + 4, glm::value_ptr(fragPos), 12);
    std::memcpy(out10 + 7, glm::value_ptr(n), 12);
}
// shadow_map.vert:49-51
void ref_shadow_vert(const float* transform, const float* view, const float* proj, const float* pos, float* out4) {
    glm::vec3 fragPos = glm::vec3(glm::make_mat4(transform) * glm::vec4(pos[0], pos[1], pos[2], 1.0f));
    glm::vec4 clip = glm::make_mat4(proj) * glm::make_mat4(view) * glm::vec4(fragPos, 1.0f);
    std::memcpy(out4, glm::value_ptr(clip), 16);
}

static glsl::Light make_light(const uint8_t* p) {
    glsl::Light l;
    std::memcpy(&l.type, p + 0, 4);
    std::memcpy(&l.position, p + 16, 12);
    std::memcpy(&l.direction, p + 32, 12);
    std::memcpy(&l.ambient, p + 48, 16);
    std::memcpy(&l.diffuse, p + 64, 16);
    std::memcpy(&l.specular, p + 80, 16);
    std::memcpy(&l.innerCutoff, p + 96, 4);
    std::memcpy(&l.outerCutoff, p + 100, 4);
    std::memcpy(&l.shadowIndex, p + 104, 4);
    return l;
}

void ref_phong_lighting(const uint8_t* light112, const float* p, const float* normal, const float* fragPos,
                        const float* viewPos, float shadow, float* out4) {
    glsl::PhongProperties props;
    props.diffuse = glm::vec4(p[0], p[1], p[2], p[3]);
    props.specular = glm::vec4(p[4], p[5], p[6], p[7]);
    props.shininess = p[8];
    glm::vec4 r = glsl::PhongLighting(make_light(light112), props, glm::make_vec3(normal), glm::make_vec3(fragPos),
                                      glm::make_vec3(viewPos), shadow);
    std::memcpy(out4, glm::value_ptr(r), 16);
}

float ref_compute_shadow(const uint8_t* light112, const float* fragPos, const float* normal, const float* transform16,
                         const float* shadowMap, uint32_t S) {
    static glsl::MaterialShadowData storage[1];
    storage[0].transform = glm::make_mat4(transform16);
    storage[0].shadowMapTextureHandle = 0;
    glsl::g_MaterialShadowDataBuffer = reinterpret_cast<glsl::MaterialShadowDataBuffer*>(storage);
    glsl::uGlobalTextures2D[0] = {shadowMap, S};
    glsl::Light l = make_light(light112);
    l.shadowIndex = 0;
    return glsl::ComputeShadow(l, glm::make_vec3(fragPos), glm::make_vec3(normal), 0u);
}

// Camera.h:8-130 / Camera.cpp:3-25: view, proj (= kClip * perspective), 8 frustum corners
void ref_camera(const float* eye, const float* lookat, const float* up, float fov, float nearP, float farP, int w, int h,
                float* view16, float* proj16, float* corners24) {
    Camera cam(glm::make_vec3(eye), glm::make_vec3(lookat), glm::make_vec3(up), fov, nearP, farP, w, h);
    cam.SetImageExtent(w, h);
    std::memcpy(view16, glm::value_ptr(cam.GetViewMatrix()), 64);
    std::memcpy(proj16, glm::value_ptr(cam.GetProjectionMatrix()), 64);
    auto c = cam.ComputeFrustrumCorners();
    for (int i = 0; i < 8; ++i) std::memcpy(corners24 + 3 * i, glm::value_ptr(c[i]), 12);
}
// BoundingBox.h:47-60 Transform and BoundingBox.cpp:3-16 operator*
void ref_bbox_transform(const float* mn, const float* mx, const float* m16, int useOperator, float* outMin, float* outMax) {
    BoundingBox b; b.min = glm::make_vec3(mn); b.max = glm::make_vec3(mx);
    BoundingBox r = useOperator ? (glm::make_mat4(m16) * b) : b.Transform(glm::make_mat4(m16));
    std::memcpy(outMin, glm::value_ptr(r.min), 12); std::memcpy(outMax, glm::value_ptr(r.max), 12);
}
// glm pieces used by ShadowSystem.cpp:129-204 (the file itself needs Vulkan headers and cannot compile)
void ref_look_at(const float* eye, const float* center, const float* up, float* out16) {
    glm::mat4 m = glm::lookAt(glm::make_vec3(eye), glm::make_vec3(center), glm::make_vec3(up));
    std::memcpy(out16, glm::value_ptr(m), 64);
}
void ref_clip_ortho(float l, float r, float b, float t, float n, float f, float* out16) {
    glm::mat4 m = glm_vk::kClip * glm::ortho(l, r, b, t, n, f);
    std::memcpy(out16, glm::value_ptr(m), 64);
}
void ref_mat4_inverse(const float* m16, float* out16) {
    glm::mat4 m = glm::inverse(glm::make_mat4(m16));
    std::memcpy(out16, glm::value_ptr(m), 64);
}

// ComputeDirectionalLightViewMatrix + ComputeShadowTransform (ShadowSystem.cpp:129-204).  That file needs the Vulkan headers and
// cannot be compiled here, so its statements are replayed over the reference's OWN compiled pieces — Camera::ComputeFrustrumCorners
// (Camera.cpp:3-25), BoundingBox::FromPoints / operator* / Transform / Intersects / Union (BoundingBox.{h,cpp}), glm::lookAt /
// inverse / ortho, glm_vk::kClip — in the reference's order.  One deviation, SURVEY.md F8: ShadowSystem.cpp:136 reads `up`
// uninitialised (`up = cross(lightDirection, up)`); the evident intent `cross(lightDirection, right)` is used.
void ref_shadow_transform(const float* lightDir, const float* eye, const float* lookat, const float* up, float fov, float nearP, float farP, int w, int h,
                          const float* sceneMin, const float* sceneMax, const float* boxes6, const float* transforms16, int nBoxes,
                          float* view16, float* proj16, float* transform16) {
    Camera camera(glm::make_vec3(eye), glm::make_vec3(lookat), glm::make_vec3(up), fov, nearP, farP, w, h);
    camera.SetImageExtent(w, h);
    const glm::vec3 lightDirection = glm::make_vec3(lightDir);
    glm::vec3 right(0.0f, 0.0f, 1.0f);
    if (std::abs(glm::dot(lightDirection, right)) > 0.9999f) right = glm::vec3(1.0f, 0.0f, 0.0f);
    const glm::vec3 lup = glm::cross(lightDirection, right);
    const glm::mat4 shadowView = glm::lookAt(glm::vec3(0.0f, 0.0f, 0.0f), lightDirection, lup);
    BoundingBox sceneBox; sceneBox.min = glm::make_vec3(sceneMin); sceneBox.max = glm::make_vec3(sceneMax);
    BoundingBox camBox_world = BoundingBox::FromPoints(camera.ComputeFrustrumCorners());
    BoundingBox camBox_view = shadowView * camBox_world;
    BoundingBox sceneBox_view = shadowView * sceneBox;
    camBox_view.max.z = sceneBox_view.max.z;
    camBox_world = camBox_view.Transform(glm::inverse(shadowView));
    BoundingBox lightBox_world;
    for (int i = 0; i < nBoxes; ++i) {
        BoundingBox box_local; box_local.min = glm::make_vec3(boxes6 + 6 * i); box_local.max = glm::make_vec3(boxes6 + 6 * i + 3);
        BoundingBox box_world = box_local.Transform(glm::make_mat4(transforms16 + 16 * i));
        if (box_world.Intersects(camBox_world)) lightBox_world = lightBox_world.Union(box_world);
    }
    BoundingBox lightBox_view = lightBox_world.Transform(shadowView);
    const glm::mat4 proj = glm_vk::kClip * glm::ortho(lightBox_view.min.x, lightBox_view.max.x, lightBox_view.min.y, lightBox_view.max.y, lightBox_view.min.z, lightBox_view.max.z);
    const glm::mat4 transform = proj * shadowView;   // ShadowSystem.cpp:303
    std::memcpy(view16, glm::value_ptr(shadowView), 64); std::memcpy(proj16, glm::value_ptr(proj), 64); std::memcpy(transform16, glm::value_ptr(transform), 64);
}

}  // extern "C"
