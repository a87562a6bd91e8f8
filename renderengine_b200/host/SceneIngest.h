// SceneIngest.h — the data conventions of the reference's scene loader, applied to an already-parsed scene.
//
// The reference ingests scene files through Assimp (AssimpSceneLoader.cpp); the importer library itself is a file
// parser outside the raster hot path and is not part of this build.  What the hot path depends on is how the loader
// turns importer records into the GPU-visible streams, and that is restated here over plain structs shaped like
// Assimp's (aiNode / aiMesh / aiLight / aiMaterial / aiCamera), so any importer can feed it:
//
//   * node transforms arrive row-major and are transposed to column-major, children inherit parent * node
//     (AssimpSceneLoader.cpp:186-196);
//   * every mesh-bearing node becomes one SceneTree node; its meshes are appended to the global vertex / index arrays
//     with ABSOLUTE indices (index + running vertex count), uv.v negated "for vulkan", missing normals / uvs = 0,
//     object-space AABB from the positions (:198-268);
//   * lights: colour normalised by its largest channel when that exceeds 1, intensity = 10 * max / 683, position =
//     translation of the node's global transform, direction rotated by it, cones as cosines, point falloff
//     1 / attenuationConstant (or 1e6), directional lights take consecutive shadow indices (:93-151);
//   * materials: base colour, emissive, f0 = ((ior - 1) / (ior + 1))^2, metallic, roughness, alpha mode -> the 80-B
//     MaterialProperties record with all texture slots Invalid unless supplied (:270-361);
//   * camera: eye = node translation, look-at origin, up = node y axis, fov = horizontalFOV in degrees * 2; without a
//     camera an orbit view at 15 * the largest vertex distance on the (1,1,1) diagonal with up = +z (:153-178).
#pragma once
#include "RenderEngine.h"

#include <cmath>
#include <string>
#include <vector>

namespace reb {
namespace ingest {

struct MeshSource {                      // aiMesh
    std::vector<float> positions;        // 3 per vertex
    std::vector<float> normals;          // 3 per vertex, or empty (HasNormals() == false)
    std::vector<float> texCoords0;       // 3 per vertex as aiVector3D (u, v, w), or empty
    std::vector<uint32_t> faces;         // indicesPerFace per face
    uint32_t indicesPerFace = 3;         // mFaces->mNumIndices (the loader assumes it is the same for every face)
    uint32_t materialIndex = 0;
};
struct NodeSource {                      // aiNode
    std::string name;
    float transformRowMajor[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};   // aiMatrix4x4 a1..d4
    std::vector<uint32_t> meshes;
    std::vector<uint32_t> children;      // indices into Scene::nodes
    int parent = -1;
};
enum LightSourceType : uint32_t { Undefined = 0, Directional = 1, Point = 2, Spot = 3 };   // aiLightSourceType
struct LightSource {                     // aiLight
    std::string name;                    // name of the node that places it
    uint32_t type = Directional;
    float colorDiffuse[3] = {1, 1, 1};
    float direction[3] = {0, 0, -1};
    float angleInnerCone = 0.0f, angleOuterCone = 0.0f;   // radians
    float attenuationConstant = 0.0f;
};
struct MaterialSource {                  // aiMaterial properties the loader reads; has* = the key is present
    bool hasBaseColor = false; float baseColor[4] = {1, 1, 1, 1};
    bool hasEmissive = false; float emissive[4] = {0, 0, 0, 0};
    bool hasIor = false; float ior = 1.5f;
    bool hasMetallic = false; float metallic = 0.0f;
    bool hasRoughness = false; float roughness = 0.0f;
    std::string alphaMode;               // "", "OPAQUE", "MASK", "BLEND" (AI_MATKEY_GLTF_ALPHAMODE)
    TextureHandle baseColorTexture = TextureHandle::Invalid, emissiveTexture = TextureHandle::Invalid,
                  metallicRoughnessTexture = TextureHandle::Invalid, normalsTexture = TextureHandle::Invalid;   // already registered
};
struct CameraSource { std::string name; float horizontalFov = 0.7853982f; };   // aiCamera
struct Scene {
    std::vector<NodeSource> nodes;       // nodes[0] is the root
    std::vector<MeshSource> meshes;
    std::vector<LightSource> lights;
    std::vector<MaterialSource> materials;
    std::vector<CameraSource> cameras;
};

struct Result {
    std::vector<MaterialHandle> materials;       // per source material
    std::vector<AlphaMode> alphaModes;           // per source material
    std::vector<SceneNodeHandle> sceneNodes;     // one per mesh-bearing node, in traversal order
    BoundingBox boundingBox;                     // world box of everything loaded
    float maxVertexDist = 0.0f;
    uint32_t nbShadowCastingLights = 0;
};

inline mat4 ToColumnMajor(const float* rowMajor) {   // glm::transpose(glm::make_mat4(&a1)) (AssimpSceneLoader.cpp:190)
    mat4 m;
    for (int r = 0; r < 4; ++r)
        for (int c = 0; c < 4; ++c) m.at(c, r) = rowMajor[4 * r + c];
    return m;
}

// ComputeAiNodeGlobalTransform: product of the node's transform with every ancestor's, root first
inline mat4 GlobalTransform(const Scene& scene, int node) {
    mat4 t = mat4::identity();
    for (int n = node; n >= 0; n = scene.nodes[(size_t)n].parent) t = ToColumnMajor(scene.nodes[(size_t)n].transformRowMajor) * t;
    return t;
}
inline int FindNode(const Scene& scene, const std::string& name) {
    for (size_t i = 0; i < scene.nodes.size(); ++i)
        if (scene.nodes[i].name == name) return (int)i;
    return -1;
}

// AssimpSceneLoader::LoadMaterials (:270-361)
inline pbr::MaterialProperties ConvertMaterial(const MaterialSource& src, AlphaMode* alphaMode = nullptr) {
    pbr::MaterialProperties p;
    if (src.hasBaseColor) for (int c = 0; c < 4; ++c) p.baseColor[c] = src.baseColor[c];
    if (src.hasEmissive) for (int c = 0; c < 4; ++c) p.emissive[c] = src.emissive[c];
    if (src.hasIor) p.f0 = (float)std::pow((src.ior - 1.0f) / (src.ior + 1.0f), 2);
    if (src.hasMetallic) p.metallic = src.metallic;
    if (src.hasRoughness) p.perceptualRoughness = src.roughness;
    p.baseColorTexture = src.baseColorTexture; p.emissiveTexture = src.emissiveTexture;
    p.occlusionMetallicRoughnessTexture = src.metallicRoughnessTexture; p.normalsTexture = src.normalsTexture;
    p.ambientOcclusionTexture = TextureHandle::Invalid;
    if (alphaMode) *alphaMode = src.alphaMode == "MASK" ? AlphaMode::eMask : (src.alphaMode == "BLEND" ? AlphaMode::eBlend : AlphaMode::eOpaque);
    return p;
}

// AssimpSceneLoader::LoadLights (:93-151); *castsShadow = directional (the only kind that takes a shadow index)
inline pbr::Light ConvertLight(const LightSource& src, const mat4& nodeGlobalTransform, uint32_t* nbShadowCastingLights, bool* castsShadow = nullptr) {
    pbr::Light light;
    light.type = src.type;
    float intensity = std::fmax(src.colorDiffuse[0], std::fmax(src.colorDiffuse[1], src.colorDiffuse[2]));
    for (int c = 0; c < 3; ++c) light.color[c] = intensity > 1.0f ? src.colorDiffuse[c] / intensity : src.colorDiffuse[c];
    for (int c = 0; c < 3; ++c) light.position[c] = nodeGlobalTransform.m[12 + c];
    light.intensity = 10.0f * intensity / 683.0f;
    bool shadows = false;
    if (src.type == Directional || src.type == Spot) {
        const vec4 d = nodeGlobalTransform * vec4{src.direction[0], src.direction[1], src.direction[2], 0.0f};
        light.direction[0] = d.x; light.direction[1] = d.y; light.direction[2] = d.z;
        if (src.type == Directional) { light.shadowIndex = (*nbShadowCastingLights)++; shadows = true; }
        if (src.type == Spot) { light.cosInnerAngle = std::cos(src.angleInnerCone); light.cosOuterAngle = std::cos(src.angleOuterCone); }
    } else if (src.type == Point) {
        light.falloffRadius = src.attenuationConstant > 1.0e-6f ? 1.0f / src.attenuationConstant : 1.0e6f;
    }
    if (castsShadow) *castsShadow = shadows;
    return light;
}

// AssimpSceneLoader::LoadSceneNode (:198-268): one scene node, its meshes appended to the global arrays
inline SceneNodeHandle LoadSceneNode(const Scene& scene, const NodeSource& node, const mat4& transform, const std::vector<MaterialHandle>& materials,
                                     MeshAllocator& meshAllocator, SceneTree& sceneTree, Result& result) {
    BoundingBox box;
    std::vector<Mesh> meshes;
    for (uint32_t mi : node.meshes) {
        const MeshSource& src = scene.meshes.at(mi);
        Mesh mesh;
        mesh.indexOffset = meshAllocator.GetIndexCount();
        mesh.nbIndices = src.faces.size();
        mesh.materialHandle = materials.at(src.materialIndex);
        meshes.push_back(mesh);
        const size_t vertexIndexOffset = meshAllocator.GetVertexCount();
        const size_t nVertices = src.positions.size() / 3;
        const bool hasUV = !src.texCoords0.empty(), hasNormals = !src.normals.empty();
        meshAllocator.ReserveVertices(nVertices);
        for (size_t v = 0; v < nVertices; ++v) {
            Vertex vertex;
            vertex.pos = vec3(src.positions[3 * v], src.positions[3 * v + 1], src.positions[3 * v + 2]);
            vertex.uv.x = hasUV ? src.texCoords0[3 * v] : 0.0f;
            vertex.uv.y = -(hasUV ? src.texCoords0[3 * v + 1] : 0.0f);   // v axis flipped for vulkan (:234)
            vertex.normal = hasNormals ? vec3(src.normals[3 * v], src.normals[3 * v + 1], src.normals[3 * v + 2]) : vec3(0.0f);
            result.maxVertexDist = std::fmax(result.maxVertexDist, std::sqrt(dot(vertex.pos, vertex.pos)));
            box.min = vmin(box.min, vertex.pos); box.max = vmax(box.max, vertex.pos);
            meshAllocator.AddVertex(vertex);
        }
        meshAllocator.ReserveIndices(src.faces.size());
        for (uint32_t index : src.faces) meshAllocator.AddIndex(index + (uint32_t)vertexIndexOffset);   // absolute (:250)
    }
    const SceneNodeHandle id = sceneTree.CreateNode(transform, box);
    meshAllocator.GroupMeshes(id, meshes);
    result.boundingBox = result.boundingBox.Union(box.Transform(transform));
    result.sceneNodes.push_back(id);
    return id;
}

// AssimpSceneLoader::LoadNodeAndChildren (:186-196)
inline void LoadNodeAndChildren(const Scene& scene, uint32_t nodeIndex, const mat4& transform, const std::vector<MaterialHandle>& materials,
                                MeshAllocator& meshAllocator, SceneTree& sceneTree, Result& result) {
    const NodeSource& node = scene.nodes.at(nodeIndex);
    const mat4 newTransform = transform * ToColumnMajor(node.transformRowMajor);
    if (!node.meshes.empty()) LoadSceneNode(scene, node, newTransform, materials, meshAllocator, sceneTree, result);
    for (uint32_t child : node.children) LoadNodeAndChildren(scene, child, newTransform, materials, meshAllocator, sceneTree, result);
}

// AssimpSceneLoader::LoadCamera (:153-178)
inline void LoadCamera(const Scene& scene, float maxVertexDist, Camera& camera) {
    if (!scene.cameras.empty()) {
        const mat4 t = GlobalTransform(scene, FindNode(scene, scene.cameras[0].name));
        camera.SetCameraView(vec3(t.m[12], t.m[13], t.m[14]), vec3(0.0f), vec3(t.m[4], t.m[5], t.m[6]));
        camera.SetFieldOfView(scene.cameras[0].horizontalFov * 180.0f / 3.14159265358979323846f * 2.0f);
    } else {
        const float r = maxVertexDist * 15.0f;
        camera.SetCameraView(vec3(r, r, r), vec3(0.0f), vec3(0.0f, 0.0f, 1.0f));
    }
}

// AssimpSceneLoader::Load order: materials, scene nodes, lights, camera (:60-91).  Shadow maps are created by the
// caller for the lights reported in shadowCasters (ShadowSystem::CreateShadowMap + SetLightShadowID, :143-147).
inline Result Load(const Scene& scene, MeshAllocator& meshAllocator, SceneTree& sceneTree, pbr::MaterialSystem& materialSystem, pbr::LightSystem& lightSystem,
                   Camera& camera, std::vector<LightID>* shadowCasters = nullptr) {
    Result result;
    for (const MaterialSource& m : scene.materials) {
        AlphaMode mode;
        result.materials.push_back(materialSystem.CreateMaterialInstance(ConvertMaterial(m, &mode)));
        result.alphaModes.push_back(mode);
    }
    if (!scene.nodes.empty()) LoadNodeAndChildren(scene, 0, mat4::identity(), result.materials, meshAllocator, sceneTree, result);
    for (const LightSource& l : scene.lights) {
        bool shadows = false;
        const LightID id = lightSystem.AddLight(ConvertLight(l, GlobalTransform(scene, FindNode(scene, l.name)), &result.nbShadowCastingLights, &shadows));
        if (shadows && shadowCasters) shadowCasters->push_back(id);
    }
    LoadCamera(scene, result.maxVertexDist, camera);
    return result;
}

}  // namespace ingest
}  // namespace reb
