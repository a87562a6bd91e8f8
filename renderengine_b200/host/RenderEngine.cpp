// RenderEngine.cpp — host facade implementation: plain C++ over the C ABI of libreb.so (include/reb.h).
// No CUDA, no torch, no oracle code here; every render call ends in reb_render_shadow / reb_render_forward.
#include "RenderEngine.h"

#include <algorithm>
#include <cmath>
#include <string>

namespace reb {

// ---- BoundingBox ----------------------------------------------------------------------------------------------------
BoundingBox BoundingBox::FromPoints(const std::vector<vec3>& pts) {
    BoundingBox box;
    for (const vec3& p : pts) { box.min = vmin(box.min, p); box.max = vmax(box.max, p); }
    return box;
}
std::array<vec3, 8> BoundingBox::GetCorners() const {
    return {vec3(min.x, min.y, min.z), vec3(min.x, min.y, max.z), vec3(min.x, max.y, min.z), vec3(min.x, max.y, max.z),
            vec3(max.x, min.y, min.z), vec3(max.x, min.y, max.z), vec3(max.x, max.y, min.z), vec3(max.x, max.y, max.z)};
}
BoundingBox BoundingBox::Transform(const mat4& transform) const {
    BoundingBox box;
    for (const vec3& corner : GetCorners()) {
        const vec4 p4 = transform * vec4(corner, 1.0f);
        const vec3 p = vec3(p4.x, p4.y, p4.z) / p4.w;
        box.min = vmin(box.min, p); box.max = vmax(box.max, p);
    }
    return box;
}
BoundingBox operator*(const mat4& transform, const BoundingBox& box) { return box.Transform(transform); }

// ---- Camera ---------------------------------------------------------------------------------------------------------
Camera::Camera(vec3 eye, vec3 lookat, vec3 upVector, float fieldOfView, float nearPlane, float farPlane, int imageWidth, int imageHeight)
    : m_eye(eye), m_lookAt(lookat), m_upVector(upVector), m_nearPlane(nearPlane), m_farPlane(farPlane), m_fieldOfView(fieldOfView),
      m_imageWidth(imageWidth), m_imageHeight(imageHeight) {
    UpdateViewMatrix();
    UpdateProjectionMatrix();   // the reference leaves the projection unset until SetImageExtent; we build it here
}
void Camera::UpdateViewMatrix() {
    const vec3 directionVector = normalize(m_eye - m_lookAt);
    const vec3 rightVector = normalize(cross(m_upVector, directionVector));
    m_upVector = cross(directionVector, rightVector);
    m_viewMatrix = lookAt(m_eye, m_lookAt, m_upVector);
}
void Camera::UpdateProjectionMatrix() {
    m_projMatrix = kClip() * perspective(radians(GetFieldOfView()), m_imageWidth / (float)m_imageHeight, GetNearPlane(), GetFarPlane());
}
std::vector<vec3> Camera::ComputeFrustrumCorners() const {
    const mat4 transform = inverse(m_viewMatrix) * inverse(m_projMatrix);
    std::vector<vec3> points;
    for (float x : {-1.0f, 1.0f}) for (float y : {-1.0f, 1.0f}) for (float z : {-1.0f, 1.0f}) {
        const vec4 w = transform * vec4(x, y, z, 1.0f);
        points.push_back(vec3(w.x, w.y, w.z) / w.w);
    }
    return points;
}

// ---- Renderer ---------------------------------------------------------------------------------------------------------
Renderer::Renderer(uint32_t width, uint32_t height, int device, uint32_t outputs, void* cudaStream) : m_width(width), m_height(height) {
    reb_params p = {};
    p.struct_size = sizeof p; p.device = device; p.width = width; p.height = height; p.outputs = outputs; p.n_ranks = 1; p.stream = cudaStream;
    const int rc = reb_create(&p, &m_ctx);
    if (rc != REB_OK) throw std::runtime_error(std::string("reb_create failed: ") + reb_last_error(nullptr));
}
Renderer::~Renderer() { if (m_ctx) reb_destroy(m_ctx); }
void Renderer::Check(int status, const char* what) const {
    if (status == REB_OK) return;
    const std::string msg = std::string(what) + ": " + reb_last_error(m_ctx);
    if (status == REB_E_INVALID) throw std::invalid_argument(msg);
    throw std::runtime_error(msg);
}
void Renderer::WaitIdle() const { Check(reb_sync(m_ctx), "reb_sync"); }
void Renderer::SetSampleCount(uint32_t samples) const { Check(reb_set_sample_count(m_ctx, samples), "reb_set_sample_count"); }
void Renderer::SetShadingModel(uint32_t model) const { Check(reb_set_shading_model(m_ctx, model), "reb_set_shading_model"); }
void pbr::LightSystem::UploadToGPU(Renderer& renderer) {
    renderer.Check(reb_upload_lights(renderer.GetContext(), m_lights.data(), sizeof(Light), (uint32_t)m_lights.size()), "pbr::LightSystem::UploadToGPU");
}
void pbr::MaterialSystem::UploadToGPU(Renderer& renderer) {
    renderer.Check(reb_upload_materials(renderer.GetContext(), m_properties.data(), sizeof(MaterialProperties), (uint32_t)m_properties.size()),
                   "pbr::MaterialSystem::UploadToGPU");
}
void Skybox::SetCubeMapFaces(const uint16_t* facesRGBA16, uint32_t width, uint32_t height) {
    if (!facesRGBA16 || !width || !height) throw std::invalid_argument("Skybox::SetCubeMapFaces: six non-empty faces are required");   // TextureCache.cpp:201-202
    m_faces.assign(facesRGBA16, facesRGBA16 + (size_t)6 * width * height * 4);
    m_width = width; m_height = height;
}
void Skybox::UploadToGPU() {
    if (m_faces.empty()) return;   // no cube map staged: the base pass keeps the clear colour
    m_renderer->Check(reb_set_skybox(m_renderer->GetContext(), m_faces.data(), m_width, m_height), "Skybox::UploadToGPU");
    m_faces.clear(); m_faces.shrink_to_fit();
    m_uploaded = true;
}
void Skybox::Remove() {
    m_renderer->Check(reb_set_skybox(m_renderer->GetContext(), nullptr, 0, 0), "Skybox::Remove");
    m_uploaded = false;
}
void Renderer::Submit() const { Check(reb_flush(m_ctx), "reb_flush"); }
void Renderer::Readback(reb_output_kind kind, void* host) const { Check(reb_readback(m_ctx, kind, host), "reb_readback"); }
void Renderer::SetPartition(uint32_t rank, uint32_t nRanks, uint32_t bandH) const { Check(reb_set_partition(m_ctx, rank, nRanks, bandH), "reb_set_partition"); }
void Renderer::SetShadowPartition(uint32_t rank, uint32_t nRanks, uint32_t rows) const { Check(reb_set_shadow_partition(m_ctx, rank, nRanks, rows), "reb_set_shadow_partition"); }
void Renderer::SetShadowSplit(uint32_t rank, uint32_t nRanks) const { Check(reb_set_shadow_split(m_ctx, rank, nRanks), "reb_set_shadow_split"); }
float* Renderer::GetShadowMapDevicePointer() const {
    void* p = nullptr; uint64_t bytes = 0;
    Check(reb_get_output(m_ctx, REB_SHADOW_F32, &p, &bytes), "reb_get_output(REB_SHADOW_F32)");
    return static_cast<float*>(p);
}
void Renderer::ShadowMapFilled() const { Check(reb_shadow_mark_valid(m_ctx, 0), "reb_shadow_mark_valid"); }
void Renderer::ShadowMapArrivesWith(void* cudaEvent) const { Check(reb_shadow_wait_event(m_ctx, 0, cudaEvent), "reb_shadow_wait_event"); }
std::vector<reb_draw> Renderer::ToDraws(const std::vector<MeshDrawInfo>& drawCalls) {
    std::vector<reb_draw> d(drawCalls.size());
    for (size_t i = 0; i < drawCalls.size(); ++i)
        d[i] = {(uint32_t)drawCalls[i].mesh.indexOffset, (uint32_t)drawCalls[i].mesh.nbIndices, drawCalls[i].mesh.materialHandle.GetRaw(),
                static_cast<uint32_t>(drawCalls[i].sceneNodeID)};
    return d;
}

// ---- SceneTree --------------------------------------------------------------------------------------------------------
SceneNodeHandle SceneTree::CreateNode(mat4 transform, BoundingBox boundingBox, SceneNodeHandle parent) {
    const SceneNodeHandle id = static_cast<SceneNodeHandle>(m_transforms.size());
    m_transforms.push_back(transform);
    m_boundingBoxes.push_back(boundingBox);
    m_parents.push_back(parent);
    m_sceneBoundingBox = m_sceneBoundingBox.Union(boundingBox.Transform(transform));
    return id;
}
BoundingBox SceneTree::ComputeWorldBoundingBox() const {
    BoundingBox worldBox;
    for (size_t i = 0; i < m_transforms.size(); ++i) worldBox = worldBox.Union(m_boundingBoxes[i].Transform(m_transforms[i]));
    return worldBox;
}
void SceneTree::UploadToGPU(Renderer& renderer) {
    static_assert(sizeof(mat4) == 64, "mat4 is uploaded as 16 packed floats");
    renderer.Check(reb_upload_transforms(renderer.GetContext(), m_transforms.empty() ? nullptr : m_transforms[0].data(), (uint32_t)m_transforms.size()),
                   "SceneTree::UploadToGPU");
}

// ---- MeshAllocator ----------------------------------------------------------------------------------------------------
void MeshAllocator::GroupMeshes(SceneNodeHandle sceneNodeID, const std::vector<Mesh>& meshes) {
    m_meshEntries.push_back({sceneNodeID, {m_meshes.size(), meshes.size()}});
    m_meshes.insert(m_meshes.end(), meshes.begin(), meshes.end());
}
void MeshAllocator::UploadToGPU(Renderer& renderer) {
    renderer.Check(reb_upload_geometry(renderer.GetContext(), m_vertices.data(), m_vertices.size(), m_indices.data(), m_indices.size()),
                   "MeshAllocator::UploadToGPU");
}

// ---- LightSystem / MaterialSystem -----------------------------------------------------------------------------------------
void LightSystem::UploadToGPU(Renderer& renderer) {
    renderer.Check(reb_upload_lights(renderer.GetContext(), m_lights.data(), sizeof(Light), (uint32_t)m_lights.size()), "LightSystem::UploadToGPU");
}
MaterialHandle MaterialSystem::CreateMaterialInstance(const MaterialInstanceInfo& materialInfo) {
    const uint32_t index = (uint32_t)m_properties.size();
    m_properties.push_back(materialInfo.properties);
    m_pipelineProperties.push_back(materialInfo.pipelineProperties);
    return MaterialHandle(MaterialShadingDomain::Surface, MaterialShadingModel::Lit, index);
}
void MaterialSystem::UploadToGPU(Renderer& renderer) {
    renderer.Check(reb_upload_materials(renderer.GetContext(), m_properties.data(), sizeof(MaterialProperties), (uint32_t)m_properties.size()),
                   "MaterialSystem::UploadToGPU");
}
void MaterialSystem::Draw(Renderer& renderer, const std::vector<MeshDrawInfo>& drawCalls) const {
    const std::vector<reb_draw> d = Renderer::ToDraws(drawCalls);
    renderer.Check(reb_render_forward(renderer.GetContext(), d.data(), (uint32_t)d.size()), "MaterialSystem::Draw");
}
void MaterialSystem::Draw(Renderer& renderer, const std::vector<MeshDrawInfo>& opaque, const std::vector<MeshDrawInfo>& translucent) const {
    const std::vector<reb_draw> o = Renderer::ToDraws(opaque), t = Renderer::ToDraws(translucent);
    renderer.Check(reb_render_forward2(renderer.GetContext(), o.data(), (uint32_t)o.size(), t.data(), (uint32_t)t.size()), "MaterialSystem::Draw");
}

// ---- CameraViewSystem -------------------------------------------------------------------------------------------------
CameraViewSystem::CameraViewSystem(uint32_t imageWidth, uint32_t imageHeight)
    : m_camera(vec3(1.0f, 1.0f, 1.0f), vec3(0.0f, 0.0f, 0.0f), vec3(0.0f, 1.0f, 0.0f), 45.0f, 0.01f, 100.0f, (int)imageWidth, (int)imageHeight) {}
void CameraViewSystem::Update(Renderer& renderer) {
    const vec3 eye = m_camera.GetEye();
    std::memcpy(m_viewUniforms.view, m_camera.GetViewMatrix().data(), 64);
    std::memcpy(m_viewUniforms.proj, m_camera.GetProjectionMatrix().data(), 64);
    m_viewUniforms.position[0] = eye.x; m_viewUniforms.position[1] = eye.y; m_viewUniforms.position[2] = eye.z;
    m_viewUniforms.exposure = m_camera.GetExposure();
    renderer.Check(reb_set_view(renderer.GetContext(), &m_viewUniforms), "CameraViewSystem::Update");
}

// ---- ShadowSystem -----------------------------------------------------------------------------------------------------
// ComputeDirectionalLightViewMatrix (ShadowSystem.cpp:129-147).  The reference reads `up` uninitialised
// (`up = cross(lightDirection, up)`, `right` unused): the evident intent, cross(lightDirection, right), is used.
mat4 ComputeDirectionalLightViewMatrix(vec3 lightDirection) {
    vec3 right(0.0f, 0.0f, 1.0f);
    if (std::fabs(dot(lightDirection, right)) > 0.9999f) right = vec3(1.0f, 0.0f, 0.0f);
    const vec3 up = cross(lightDirection, right);
    return lookAt(vec3(0.0f, 0.0f, 0.0f), lightDirection, up);
}
// ComputeShadowTransform (ShadowSystem.cpp:148-204): light view, ortho fit of the scene boxes that intersect the camera box
// stretched towards the light, proj = kClip * ortho.  Pure host arithmetic; checked bit for bit against the reference's own
// Camera / BoundingBox / glm code by tests/cpp/test_golden_math.cpp.
ShadowTransform ComputeShadowTransform(vec3 lightDirection, const Camera& camera, BoundingBox sceneBox, const std::vector<BoundingBox>& boxes,
                                       const std::vector<mat4>& transforms) {
    BoundingBox camBox_world = BoundingBox::FromPoints(camera.ComputeFrustrumCorners());
    const mat4 shadowView = ComputeDirectionalLightViewMatrix(lightDirection);
    BoundingBox camBox_view = shadowView * camBox_world;
    const BoundingBox sceneBox_view = shadowView * sceneBox;
    camBox_view.max.z = sceneBox_view.max.z;   // keep every caster between the light and the camera frustum
    camBox_world = camBox_view.Transform(inverse(shadowView));
    BoundingBox lightBox_world;
    for (size_t i = 0; i < boxes.size(); ++i) {
        const BoundingBox box_world = boxes[i].Transform(transforms[i]);
        if (box_world.Intersects(camBox_world)) lightBox_world = lightBox_world.Union(box_world);
    }
    const BoundingBox lightBox_view = lightBox_world.Transform(shadowView);
    const mat4 proj = ortho(lightBox_view.min.x, lightBox_view.max.x, lightBox_view.min.y, lightBox_view.max.y, lightBox_view.min.z, lightBox_view.max.z);
    return {shadowView, kClip() * proj};
}

ShadowSystem::ShadowSystem(uint32_t shadowMapExtent, Renderer& renderer, RenderScene& scene) : m_extent(shadowMapExtent), m_renderer(&renderer), m_scene(&scene) {}

ShadowID ShadowSystem::CreateShadowMap(LightID lightID) {
    if (!m_lights.empty()) throw std::runtime_error("ShadowSystem::CreateShadowMap: the hot path supports one directional shadow map (SURVEY.md F10)");
    m_lights.push_back(lightID);
    m_views.emplace_back(); m_projs.emplace_back(); m_transforms.emplace_back();
    m_deviceIds.push_back(0xFFFFFFFFu);
    return (ShadowID)(m_lights.size() - 1);
}
void ShadowSystem::UploadToGPU() {
    for (size_t i = 0; i < m_lights.size(); ++i)
        if (m_deviceIds[i] == 0xFFFFFFFFu) m_renderer->Check(reb_shadow_create(m_renderer->GetContext(), m_extent, &m_deviceIds[i]), "ShadowSystem::UploadToGPU");
}
// Update (ShadowSystem.cpp:277-317): transform = proj * view (:303)
void ShadowSystem::Update(const Camera& camera, BoundingBox /*sceneBoundingBox: recomputed from the tree as in ShadowSystem.cpp:289*/) {
    if (GetShadowCount() == 0) return;
    const SceneTree* sceneTree = m_scene->GetSceneTree();
    const LightSystem* lightSystem = m_scene->GetLightSystem();
    const BoundingBox sceneBox = sceneTree->ComputeWorldBoundingBox();
    for (ShadowID id = 0; id < m_lights.size(); ++id) {
        const Light& light = lightSystem->GetLight(m_lights[id]);
        const ShadowTransform t = ComputeShadowTransform(vec3(light.direction[0], light.direction[1], light.direction[2]), camera, sceneBox,
                                                         sceneTree->GetBoundingBoxes(), sceneTree->GetTransforms());
        m_views[id] = t.view;
        m_projs[id] = t.proj;
        m_transforms[id] = m_projs[id] * m_views[id];
        if (m_deviceIds[id] != 0xFFFFFFFFu)
            m_renderer->Check(reb_shadow_set_transform(m_renderer->GetContext(), m_deviceIds[id], m_views[id].data(), m_projs[id].data()), "ShadowSystem::Update");
    }
}
void ShadowSystem::Render(const std::vector<MeshDrawInfo>& drawCommands) const {
    const std::vector<reb_draw> d = Renderer::ToDraws(drawCommands);
    for (uint32_t dev : m_deviceIds) m_renderer->Check(reb_render_shadow(m_renderer->GetContext(), dev, d.data(), (uint32_t)d.size()), "ShadowSystem::Render");
}

// ---- RenderScene ------------------------------------------------------------------------------------------------------
RenderScene::RenderScene(Renderer& renderer, uint32_t shadowMapExtent)
    : m_renderer(&renderer), m_meshAllocator(new MeshAllocator()), m_sceneTree(new SceneTree()), m_lightSystem(new LightSystem()),
      m_shadowSystem(new ShadowSystem(shadowMapExtent, renderer, *this)),
      m_cameraViewSystem(new CameraViewSystem(renderer.GetWidth(), renderer.GetHeight())), m_materialSystem(new MaterialSystem()), m_skybox(new Skybox(renderer)) {}

void RenderScene::Init() {
    PopulateMeshDrawCalls();
    SortOpaqueMeshes();
    UploadToGPU();
}
void RenderScene::UploadToGPU() {   // same order as RenderScene.cpp:80-92
    m_sceneTree->UploadToGPU(*m_renderer);
    m_meshAllocator->UploadToGPU(*m_renderer);
    m_lightSystem->UploadToGPU(*m_renderer);
    m_shadowSystem->UploadToGPU();
    m_materialSystem->UploadToGPU(*m_renderer);
    m_skybox->UploadToGPU();
}
void RenderScene::PopulateMeshDrawCalls() {
    m_opaqueMeshes.clear(); m_translucentMeshes.clear();
    m_meshAllocator->ForEachMesh([this](SceneNodeHandle sceneNodeID, Mesh mesh) {
        MeshDrawInfo info = {sceneNodeID, mesh};
        if (!m_materialSystem->IsTranslucent(mesh.materialHandle)) m_opaqueMeshes.push_back(info);
        else m_translucentMeshes.push_back(info);
    });
}
void RenderScene::SortOpaqueMeshes() {   // by material handle, then scene node (RenderScene.cpp:105-125)
    std::sort(m_opaqueMeshes.begin(), m_opaqueMeshes.end(), [](const MeshDrawInfo& a, const MeshDrawInfo& b) {
        if (a.mesh.materialHandle != b.mesh.materialHandle) return a.mesh.materialHandle < b.mesh.materialHandle;
        return a.sceneNodeID < b.sceneNodeID;
    });
}
void RenderScene::SortTranslucentMeshes() {   // back to front along the camera axis (RenderScene.cpp:127-159)
    if (m_translucentMeshes.empty()) return;
    const mat4 viewInverse = inverse(m_cameraViewSystem->GetCamera().GetViewMatrix());
    const vec3 cameraPosition(viewInverse.at(3, 0), viewInverse.at(3, 1), viewInverse.at(3, 2));
    const vec3 front(viewInverse.at(2, 0), viewInverse.at(2, 1), viewInverse.at(2, 2));
    std::sort(m_translucentMeshes.begin(), m_translucentMeshes.end(), [&](const MeshDrawInfo& a, const MeshDrawInfo& b) {
        const mat4 ta = m_sceneTree->GetTransform(a.sceneNodeID), tb = m_sceneTree->GetTransform(b.sceneNodeID);
        const float distA = dot(front, cameraPosition - vec3(ta.at(3, 0), ta.at(3, 1), ta.at(3, 2)));
        const float distB = dot(front, cameraPosition - vec3(tb.at(3, 0), tb.at(3, 1), tb.at(3, 2)));
        if (distA != distB) return distA > distB;
        if (a.mesh.materialHandle != b.mesh.materialHandle) return a.mesh.materialHandle < b.mesh.materialHandle;
        return a.sceneNodeID < b.sceneNodeID;
    });
}
void RenderScene::Update() {
    m_cameraViewSystem->Update(*m_renderer);
    m_shadowSystem->Update(m_cameraViewSystem->GetCamera(), m_sceneTree->GetSceneBoundingBox());
    SortTranslucentMeshes();
}
void RenderScene::Render() {
    if (m_areShadowsDirty) { RenderShadowDepthPass(); m_areShadowsDirty = false; }
    RenderBasePass();
}
void RenderScene::RenderShadowDepthPass() const {   // RenderScene.cpp:191-208: opaque ++ translucent
    if (m_shadowSystem->GetShadowCount() == 0 || (m_opaqueMeshes.empty() && m_translucentMeshes.empty())) return;
    std::vector<MeshDrawInfo> drawCalls(m_opaqueMeshes);
    drawCalls.insert(drawCalls.end(), m_translucentMeshes.begin(), m_translucentMeshes.end());
    m_shadowSystem->Render(drawCalls);
}
void RenderScene::RenderBasePass() const {
    // RenderScene.cpp:210-232 issues the opaque list, then the translucent list (sorted back to front by Update), then the
    // skybox inside one render pass: one reb_render_forward2 call.
    m_materialSystem->Draw(*m_renderer, m_opaqueMeshes, m_translucentMeshes);
}

}  // namespace reb
