// RenderEngine.h — C++ host facade over the C ABI (include/reb.h).
//
// Keeps the call surface of the reference Runtime for the raster hot path — class and method names,
// argument meaning, ownership (systems own GPU resources, callers hold integer handles) and error
// behaviour (std::runtime_error / std::invalid_argument instead of status codes) — so that code written
// against hugbed/RenderEngine's RenderScene / SceneTree / MeshAllocator / LightSystem / ShadowSystem /
// MaterialSystem / CameraViewSystem / Camera keeps working with the draw path replaced by CUDA:
//
//   reference class (file under /root/reference/Source/Runtime)      facade
//   Renderer             Renderer/Renderer.h:23-66                   reb::Renderer      (headless: owns the reb_ctx)
//   RenderScene          Renderer/RenderScene.h:21-69                reb::RenderScene
//   SceneTree            Renderer/SceneTree.h:24-77                  reb::SceneTree
//   MeshAllocator        Renderer/MeshAllocator.h:41-84              reb::MeshAllocator
//   LightSystem          Renderer/LightSystem.h:37-66                reb::LightSystem   (Phong Light of phong.glsl:20-37)
//   ShadowSystem         Renderer/ShadowSystem.h:39-109              reb::ShadowSystem
//   MaterialSystem       Renderer/MaterialSystem.h:72-154            reb::MaterialSystem (PhongProperties of phong.glsl:2-12)
//   CameraViewSystem     Renderer/CameraViewSystem.h:17-40           reb::CameraViewSystem
//   Camera               Renderer/Camera.h:8-130                     reb::Camera
//   Skybox               Renderer/Skybox.h:17-55                     reb::Skybox        (cube-map background, reb_set_skybox)
//   BoundingBox          BoundingBox.h:7-65                          reb::BoundingBox
//
// Vulkan-only arguments (CommandRingBuffer&, RenderCommandEncoder&, vk::Extent2D) are dropped or replaced by
// plain integers; everything else is positional-compatible.  Not thread-safe (the reference frame loop is
// single-threaded, RenderLoop.cpp:52-67).
#pragma once
#include "../../include/reb.h"
#include "reb_math.h"

#include <array>
#include <cstdint>
#include <limits>
#include <memory>
#include <stdexcept>
#include <utility>
#include <vector>

namespace reb {

// ---- BoundingBox (BoundingBox.h:7-65) -----------------------------------------------------------------------------
struct BoundingBox {
    vec3 min = vec3(+(std::numeric_limits<float>::max)());
    vec3 max = vec3(-(std::numeric_limits<float>::max)());
    BoundingBox Union(const BoundingBox& o) const { BoundingBox b; b.min = vmin(min, o.min); b.max = vmax(max, o.max); return b; }
    BoundingBox Intersection(const BoundingBox& o) const { BoundingBox b; b.min = vmax(min, o.min); b.max = vmin(max, o.max); return b; }
    bool Intersects(const BoundingBox& b) const {
        return (min.x <= b.max.x && max.x >= b.min.x) && (min.y <= b.max.y && max.y >= b.min.y) && (min.z <= b.max.z && max.z >= b.min.z);
    }
    static BoundingBox FromPoints(const std::vector<vec3>& pts);
    std::array<vec3, 8> GetCorners() const;
    BoundingBox Transform(const mat4& transform) const;   // 8 corners, divide by w, min/max
};
BoundingBox operator*(const mat4& transform, const BoundingBox& box);

// ---- Camera (Camera.h:8-130) ----------------------------------------------------------------------------------------
class Camera {
public:
    Camera() = default;
    Camera(vec3 eye, vec3 lookat, vec3 upVector, float fieldOfView, float nearPlane, float farPlane, int imageWidth, int imageHeight);
    mat4 GetViewMatrix() const { return m_viewMatrix; }
    mat4 GetProjectionMatrix() const { return m_projMatrix; }
    vec3 GetEye() const { return m_eye; }
    vec3 GetUpVector() const { return m_upVector; }
    vec3 GetLookAt() const { return m_lookAt; }
    float GetFieldOfView() const { return m_fieldOfView; }
    float GetNearPlane() const { return m_nearPlane; }
    float GetFarPlane() const { return m_farPlane; }
    float GetExposure() const { return m_exposure; }
    std::vector<vec3> ComputeFrustrumCorners() const;   // Camera.cpp:3-25 (spelling as in the reference)
    void SetFieldOfView(float fov) { m_fieldOfView = fov; UpdateProjectionMatrix(); }
    void SetImageExtent(int width, int height) { m_imageWidth = width; m_imageHeight = height; UpdateProjectionMatrix(); }
    void SetCameraView(vec3 eye, vec3 lookat, vec3 up) { m_eye = eye; m_lookAt = lookat; m_upVector = up; UpdateViewMatrix(); }
    void SetExposure(float exposure) { m_exposure = exposure; }
    void Move(const vec3& direction, float speed) { m_eye = m_eye + speed * direction; m_lookAt = m_lookAt + speed * direction; UpdateViewMatrix(); }
    void LookAt(vec3 lookat, vec3 up) { m_lookAt = lookat; m_upVector = up; UpdateViewMatrix(); }

private:
    void UpdateViewMatrix();         // Camera.h:93-108: re-orthogonalised up, then lookAt
    void UpdateProjectionMatrix();   // Camera.h:110-117: kClip * perspective(radians(fov), w/h, near, far)
    mat4 m_viewMatrix, m_projMatrix;
    vec3 m_eye, m_lookAt, m_upVector{0.0f, 1.0f, 0.0f};
    float m_nearPlane = 0.01f, m_farPlane = 100.0f, m_fieldOfView = 45.0f, m_exposure = 1.0f;   // CameraViewSystem.cpp:11-15 defaults
    int m_imageWidth = 1, m_imageHeight = 1;
};

// ---- the host arithmetic of ShadowSystem::Update (ShadowSystem.cpp:129-204), free functions as in the reference's anonymous namespace ----
struct ShadowTransform { mat4 view, proj; };   // proj = kClip * ortho
mat4 ComputeDirectionalLightViewMatrix(vec3 lightDirection);
ShadowTransform ComputeShadowTransform(vec3 lightDirection, const Camera& camera, BoundingBox sceneBox, const std::vector<BoundingBox>& boxes,
                                       const std::vector<mat4>& transforms);

// ---- handles (BindlessDefines.h:6-8, SceneTree.h:15, LightSystem.h:35, ShadowSystem.h:37, MaterialDefines.h:26-97) ---------
enum class SceneNodeHandle : uint32_t { Invalid = (std::numeric_limits<uint32_t>::max)() };
enum class TextureHandle : uint32_t { Invalid = (std::numeric_limits<uint32_t>::max)() };
using LightID = uint32_t;
using ShadowID = uint32_t;

enum class MaterialShadingDomain : uint32_t { Surface = 0, Count };
enum class MaterialShadingModel : uint32_t { Unlit = 0, Lit = 1, Count };
class MaterialHandle {   // bits 0-1 domain, 2-3 model, 4-31 index; Invalid = 0xFFFFFFFF
public:
    constexpr MaterialHandle() : m_raw((std::numeric_limits<uint32_t>::max)()) {}
    MaterialHandle(MaterialShadingDomain domain, MaterialShadingModel model, uint32_t index)
        : m_raw((static_cast<uint32_t>(domain) & 3u) | ((static_cast<uint32_t>(model) & 3u) << 2) | (index << 4)) {
        if (index >= (1u << 28)) throw std::invalid_argument("MaterialHandle: index needs more than 28 bits");
    }
    static constexpr MaterialHandle Invalid() { return MaterialHandle(); }
    MaterialShadingDomain GetDomain() const { return static_cast<MaterialShadingDomain>(m_raw & 3u); }
    MaterialShadingModel GetModel() const { return static_cast<MaterialShadingModel>((m_raw >> 2) & 3u); }
    uint32_t GetIndex() const { return m_raw >> 4; }
    void SetIndex(uint32_t index) { m_raw = (m_raw & 0xFu) | (index << 4); }
    uint32_t GetRaw() const { return m_raw; }
    bool operator==(const MaterialHandle& o) const { return m_raw == o.m_raw; }
    bool operator!=(const MaterialHandle& o) const { return m_raw != o.m_raw; }
    bool operator<(const MaterialHandle& o) const { return m_raw < o.m_raw; }
private:
    uint32_t m_raw;
};

// ---- geometry records (MeshAllocator.h:14-39) ---------------------------------------------------------------------------
struct Vertex { vec3 pos; vec3 normal; vec2 uv; };
static_assert(sizeof(Vertex) == 32, "Vertex must keep the reference's 32-byte layout");
struct Mesh { uint64_t indexOffset = 0; uint64_t nbIndices = 0; MaterialHandle materialHandle = MaterialHandle::Invalid(); };
struct MeshDrawInfo { SceneNodeHandle sceneNodeID = SceneNodeHandle::Invalid; Mesh mesh; };

class Renderer;

// ---- SceneTree (SceneTree.h:24-77) ----------------------------------------------------------------------------------------
class SceneTree {
public:
    SceneNodeHandle CreateNode(mat4 transform, BoundingBox boundingBox, SceneNodeHandle parent = SceneNodeHandle::Invalid);
    size_t GetNodeCount() const { return m_transforms.size(); }
    void UploadToGPU(Renderer& renderer);
    BoundingBox ComputeWorldBoundingBox() const;
    mat4 GetTransform(SceneNodeHandle id) const { return m_transforms.at(static_cast<size_t>(id)); }
    const std::vector<mat4>& GetTransforms() const { return m_transforms; }
    const std::vector<BoundingBox>& GetBoundingBoxes() const { return m_boundingBoxes; }
    const BoundingBox& GetSceneBoundingBox() const { return m_sceneBoundingBox; }
private:
    std::vector<BoundingBox> m_boundingBoxes;
    std::vector<mat4> m_transforms;
    std::vector<SceneNodeHandle> m_parents;
    BoundingBox m_sceneBoundingBox;
};

// ---- MeshAllocator (MeshAllocator.h:41-84) ------------------------------------------------------------------------------
class MeshAllocator {
public:
    void GroupMeshes(SceneNodeHandle sceneNodeID, const std::vector<Mesh>& meshes);
    void UploadToGPU(Renderer& renderer);
    void ReserveVertices(size_t count) { m_vertices.reserve(m_vertices.size() + count); }
    void ReserveIndices(size_t count) { m_indices.reserve(m_indices.size() + count); }
    void AddVertex(Vertex vertex) { m_vertices.push_back(vertex); }
    void AddIndex(uint32_t index) { m_indices.push_back(index); }
    size_t GetVertexCount() const { return m_vertices.size(); }
    size_t GetIndexCount() const { return m_indices.size(); }
    const std::vector<Vertex>& GetVertices() const { return m_vertices; }   // (facade extension: host copies for tools/tests)
    const std::vector<uint32_t>& GetIndices() const { return m_indices; }
    template <class Func> void ForEachMesh(Func f) {
        for (const auto& entry : m_meshEntries)
            for (size_t i = entry.second.first; i < entry.second.first + entry.second.second; ++i) f(entry.first, m_meshes[i]);
    }
private:
    std::vector<std::pair<SceneNodeHandle, std::pair<size_t, size_t>>> m_meshEntries;   // node -> (offset, size) into m_meshes
    std::vector<Vertex> m_vertices;
    std::vector<uint32_t> m_indices;
    std::vector<Mesh> m_meshes;
};

// ---- LightSystem (LightSystem.h:12-66); record = the Phong Light of phong.glsl:20-37 in std430 (112 B) -----------------------
enum class LightType : uint32_t { Directional = 1, Point = 2, Spot = 3, Count };
struct alignas(16) Light {
    uint32_t type = static_cast<uint32_t>(LightType::Directional); uint32_t pad0[3] = {0, 0, 0};
    float position[3] = {0, 0, 0}; float pad1 = 0;
    float direction[3] = {0, -1, 0}; float pad2 = 0;
    float ambient[4] = {0, 0, 0, 0}, diffuse[4] = {1, 1, 1, 1}, specular[4] = {1, 1, 1, 1};
    float innerCutoff = 1.0f, outerCutoff = 0.0f; uint32_t shadowIndex = 0; uint32_t pad3 = 0;
};
static_assert(sizeof(Light) == 112, "Light must match the std430 layout of phong.glsl:20-37");

class LightSystem {
public:
    void ReserveLights(size_t count) { m_lights.reserve(m_lights.size() + count); }
    LightID AddLight(Light light) { m_lights.push_back(light); return static_cast<LightID>(m_lights.size() - 1); }
    void SetLightShadowID(LightID id, uint32_t shadowID) { m_lights.at(id).shadowIndex = shadowID; }
    const std::vector<Light>& GetLights() const { return m_lights; }
    const Light& GetLight(LightID id) const { return m_lights.at(id); }
    uint32_t GetLightCount() const { return static_cast<uint32_t>(m_lights.size()); }
    void UploadToGPU(Renderer& renderer);
private:
    std::vector<Light> m_lights;
};

// ---- MaterialSystem (MaterialSystem.h:41-154); record = PhongProperties + PhongTextures of phong.glsl:2-12 (48 B) --------------
struct MaterialProperties {
    float diffuse[4] = {1, 1, 1, 1}, specular[4] = {1, 1, 1, 1};
    float shininess = 32.0f;
    TextureHandle diffuseTexture = TextureHandle::Invalid, specularTexture = TextureHandle::Invalid;
    uint32_t padding = 0;
};
static_assert(sizeof(MaterialProperties) == 48, "Phong material record is 48 bytes (SURVEY.md a12)");
enum class AlphaMode { eOpaque, eMask, eBlend };
struct MaterialPipelineProperties { AlphaMode alphaMode = AlphaMode::eOpaque; };
struct MaterialInstanceInfo { MaterialProperties properties = {}; MaterialPipelineProperties pipelineProperties = {}; };

class MaterialSystem {
public:
    MaterialHandle CreateMaterialInstance(const MaterialInstanceInfo& materialInfo);   // MaterialSystem.cpp:49-59
    void UploadToGPU(Renderer& renderer);                                               // MaterialSystem.cpp:68-86
    void Draw(Renderer& renderer, const std::vector<MeshDrawInfo>& drawCalls) const;     // MaterialSystem.cpp:88-101 -> reb_render_forward
    // both Draw calls of RenderScene::RenderBasePass (RenderScene.cpp:226-227) as one base pass -> reb_render_forward2
    void Draw(Renderer& renderer, const std::vector<MeshDrawInfo>& opaque, const std::vector<MeshDrawInfo>& translucent) const;
    size_t GetMaterialInstanceCount() const { return m_properties.size(); }
    const std::vector<MaterialProperties>& GetProperties() const { return m_properties; }   // (facade extension)
    bool IsTranslucent(MaterialHandle h) const { return m_pipelineProperties.at(h.GetIndex()).alphaMode == AlphaMode::eBlend; }
private:
    std::vector<MaterialProperties> m_properties;
    std::vector<MaterialPipelineProperties> m_pipelineProperties;
};

// ---- PBR records: the reference's OWN GPU-visible structs for its live fragment path (surface_pbr.frag -> pbr.glsl) --------------
// Use with Renderer::SetShadingModel(REB_SHADING_PBR): pbr::LightSystem / pbr::MaterialSystem upload 80-B records.
namespace pbr {
struct alignas(16) Light {   // LightSystem.h:22-33 <-> pbr.glsl:288-299 (std430, 80 B)
    float color[3] = {1, 1, 1}; float pad0 = 0;
    float position[3] = {0, 0, 0}; float pad1 = 0;
    float direction[3] = {0, -1, 0}; float intensity = 1.0f;
    float falloffRadius = 1.0f, cosInnerAngle = 1.0f, cosOuterAngle = 0.0f; uint32_t shadowIndex = 0;
    uint32_t type = static_cast<uint32_t>(LightType::Directional); uint32_t pad2[3] = {0, 0, 0};
};
static_assert(sizeof(Light) == 80, "pbr::Light must match the std430 layout of pbr.glsl:288-299");
struct MaterialProperties {   // MaterialSystem.h:41-51 <-> Material, pbr.glsl:78-94 (80 B)
    float baseColor[4] = {1, 1, 1, 1}, emissive[4] = {0, 0, 0, 0};
    float f0 = 0.04f, metallic = 0.0f, perceptualRoughness = 0.5f, ambientOcclusion = 1.0f;
    TextureHandle baseColorTexture = TextureHandle::Invalid, emissiveTexture = TextureHandle::Invalid, occlusionMetallicRoughnessTexture = TextureHandle::Invalid,
                  normalsTexture = TextureHandle::Invalid, ambientOcclusionTexture = TextureHandle::Invalid;
    uint32_t padding[3] = {0, 0, 0};
};
static_assert(sizeof(MaterialProperties) == 80, "pbr::MaterialProperties must keep the reference's 80-byte layout");
class LightSystem {
public:
    LightID AddLight(Light light) { m_lights.push_back(light); return static_cast<LightID>(m_lights.size() - 1); }
    void SetLightShadowID(LightID id, uint32_t shadowID) { m_lights.at(id).shadowIndex = shadowID; }
    uint32_t GetLightCount() const { return static_cast<uint32_t>(m_lights.size()); }
    const std::vector<Light>& GetLights() const { return m_lights; }
    void UploadToGPU(Renderer& renderer);
private:
    std::vector<Light> m_lights;
};
class MaterialSystem {
public:
    MaterialHandle CreateMaterialInstance(const MaterialProperties& properties) {
        m_properties.push_back(properties);
        return MaterialHandle(MaterialShadingDomain::Surface, MaterialShadingModel::Lit, (uint32_t)m_properties.size() - 1);
    }
    const std::vector<MaterialProperties>& GetProperties() const { return m_properties; }   // (facade extension)
    void UploadToGPU(Renderer& renderer);
private:
    std::vector<MaterialProperties> m_properties;
};
}  // namespace pbr

// ---- ViewProperties (ViewProperties.h:27-35, 160 B) + CameraViewSystem (CameraViewSystem.h:17-40) ---------------------------
struct alignas(16) ViewProperties {
    float view[16]; float proj[16]; float position[3]; float exposure = 1.0f; uint32_t debugInput = 0, debugEquation = 0; uint32_t pad[2] = {0, 0};
};
static_assert(sizeof(ViewProperties) == 160, "ViewProperties must keep the reference's 160-byte layout");

class CameraViewSystem {
public:
    CameraViewSystem(uint32_t imageWidth, uint32_t imageHeight);
    void Update(Renderer& renderer);   // CameraViewSystem.cpp:61-72 -> reb_set_view
    const Camera& GetCamera() const { return m_camera; }
    Camera& GetCamera() { return m_camera; }
    const ViewProperties& GetViewProperties() const { return m_viewUniforms; }
private:
    ViewProperties m_viewUniforms;
    Camera m_camera;
};

class RenderScene;

// ---- ShadowSystem (ShadowSystem.h:39-109) -----------------------------------------------------------------------------------
class ShadowSystem {
public:
    ShadowSystem(uint32_t shadowMapExtent, Renderer& renderer, RenderScene& scene);
    ShadowID CreateShadowMap(LightID lightID);                                 // ShadowSystem.cpp:229-238
    void UploadToGPU();                                                        // allocates the depth map: reb_shadow_create
    void Update(const Camera& camera, BoundingBox sceneBoundingBox);           // ShadowSystem.cpp:277-317 -> reb_shadow_set_transform
    void Render(const std::vector<MeshDrawInfo>& drawCommands) const;          // ShadowSystem.cpp:319-376 -> reb_render_shadow
    size_t GetShadowCount() const { return m_lights.size(); }
    mat4 GetLightTransform(ShadowID id) const { return m_transforms.at(id); }  // proj * view (ShadowSystem.cpp:303)
    mat4 GetShadowView(ShadowID id) const { return m_views.at(id); }
    mat4 GetShadowProj(ShadowID id) const { return m_projs.at(id); }
    uint32_t GetExtent() const { return m_extent; }
private:
    uint32_t m_extent;
    Renderer* m_renderer;
    RenderScene* m_scene;
    std::vector<LightID> m_lights;
    std::vector<mat4> m_views, m_projs, m_transforms;
    std::vector<uint32_t> m_deviceIds;
};

// ---- Renderer: headless stand-in for Renderer + RenderLoop + Swapchain (Renderer.cpp:38-111) -------------------------------------
class Renderer {
public:
    Renderer(uint32_t width, uint32_t height, int device = 0, uint32_t outputs = 0, void* cudaStream = nullptr);
    ~Renderer();
    Renderer(const Renderer&) = delete;
    Renderer& operator=(const Renderer&) = delete;
    reb_ctx* GetContext() const { return m_ctx; }
    uint32_t GetWidth() const { return m_width; }
    uint32_t GetHeight() const { return m_height; }
    void Check(int status, const char* what) const;       // throws std::runtime_error with reb_last_error()
    void WaitIdle() const;                                 // reb_sync (vkDeviceWaitIdle analogue)
    void SetShadingModel(uint32_t model) const;            // REB_SHADING_PHONG (default) / REB_SHADING_PBR (surface_pbr.frag)
    void SetSampleCount(uint32_t samples) const;           // 1 (default) or 4: PhysicalDevice::GetMsaaSamples analogue (PhysicalDevice.cpp:172-186), Swapchain.cpp:110-123
    void Submit() const;                                   // reb_flush (CommandRingBuffer submit analogue, RenderLoop.cpp:124-132)
    void Readback(reb_output_kind kind, void* host) const; // swapchain image -> host (headless present)
    // multi-GPU (no counterpart in the reference: one vk::Device, Device.h:12-13); see host/demo_mgpu.cpp for a complete caller
    void SetPartition(uint32_t rank, uint32_t nRanks, uint32_t bandH) const;        // sort-first row bands of the frame
    void SetShadowPartition(uint32_t rank, uint32_t nRanks, uint32_t rows) const;   // shadow pass: row slabs, completed by an all-gather
    void SetShadowSplit(uint32_t rank, uint32_t nRanks) const;                      // shadow pass: shares of the triangle stream, completed by an all-reduce (min)
    float* GetShadowMapDevicePointer() const;                                       // the buffer those collectives run on
    void ShadowMapFilled() const;                                                   // reb_shadow_mark_valid
    void ShadowMapArrivesWith(void* cudaEvent) const;                               // reb_shadow_wait_event
    static std::vector<reb_draw> ToDraws(const std::vector<MeshDrawInfo>& drawCalls);
private:
    reb_ctx* m_ctx = nullptr;
    uint32_t m_width, m_height;
};

// ---- Skybox (Skybox.h:17-55, Skybox.cpp:58-140) --------------------------------------------------------------------------------------
// The reference loads six face images through TextureCache::LoadCubeMapFaces (TextureCache.cpp:199-290, stb_image -> RGBA16);
// image decoding is outside the hot path, so the faces arrive decoded: +X -X +Y -Y +Z -Z = right, left, top, bottom, front, back.
class Skybox {
public:
    explicit Skybox(Renderer& renderer) : m_renderer(&renderer) {}
    void SetCubeMapFaces(const uint16_t* facesRGBA16, uint32_t width, uint32_t height);   // staged until UploadToGPU
    void UploadToGPU();                                                                   // Skybox.cpp:107-124 -> reb_set_skybox
    void Remove();                                                                        // background = clear colour again
    bool IsLoaded() const { return m_uploaded; }
private:
    Renderer* m_renderer;
    std::vector<uint16_t> m_faces;
    uint32_t m_width = 0, m_height = 0;
    bool m_uploaded = false;
};

// ---- RenderScene (RenderScene.h:21-69, RenderScene.cpp:55-242) -------------------------------------------------------------------
class RenderScene {
public:
    RenderScene(Renderer& renderer, uint32_t shadowMapExtent = 4096);   // reference default 4096^2 (RenderScene.cpp:26)
    void Init();          // PopulateMeshDrawCalls, SortOpaqueMeshes, UploadToGPU (RenderScene.cpp:55-68)
    void UploadToGPU();   // RenderScene.cpp:80-92
    void Update();        // camera view block, shadow transform, translucent sort (RenderScene.cpp:161-166)
    void Render();        // shadow depth pass while dirty, then the base pass (RenderScene.cpp:168-184)
    void SetShadowsDirty() { m_areShadowsDirty = true; }   // north-star mode: re-render the shadow map every frame
    Renderer* GetRenderer() const { return m_renderer; }
    MeshAllocator* GetMeshAllocator() const { return m_meshAllocator.get(); }
    SceneTree* GetSceneTree() const { return m_sceneTree.get(); }
    LightSystem* GetLightSystem() const { return m_lightSystem.get(); }
    ShadowSystem* GetShadowSystem() const { return m_shadowSystem.get(); }
    CameraViewSystem* GetCameraViewSystem() const { return m_cameraViewSystem.get(); }
    MaterialSystem* GetMaterialSystem() const { return m_materialSystem.get(); }
    Skybox* GetSkybox() const { return m_skybox.get(); }                                  // RenderScene.h:41
    const std::vector<MeshDrawInfo>& GetOpaqueMeshes() const { return m_opaqueMeshes; }
    const std::vector<MeshDrawInfo>& GetTranslucentMeshes() const { return m_translucentMeshes; }
private:
    void PopulateMeshDrawCalls();
    void SortOpaqueMeshes();
    void SortTranslucentMeshes();
    void RenderShadowDepthPass() const;
    void RenderBasePass() const;
    Renderer* m_renderer;
    std::unique_ptr<MeshAllocator> m_meshAllocator;
    std::unique_ptr<SceneTree> m_sceneTree;
    std::unique_ptr<LightSystem> m_lightSystem;
    std::unique_ptr<ShadowSystem> m_shadowSystem;
    std::unique_ptr<CameraViewSystem> m_cameraViewSystem;
    std::unique_ptr<MaterialSystem> m_materialSystem;
    std::unique_ptr<Skybox> m_skybox;
    std::vector<MeshDrawInfo> m_opaqueMeshes, m_translucentMeshes;
    bool m_areShadowsDirty = true;
};

}  // namespace reb
