// CPU check of renderengine_b200/host/SceneIngest.h against the conventions of the reference's AssimpSceneLoader.cpp
// (cited per check).  Stages data into the facade's host-side systems only: no device call is made.
#include "../../renderengine_b200/host/SceneIngest.h"

#include <cstdio>
#include <cstdlib>

using namespace reb;

#define CHECK(c) do { if (!(c)) { std::fprintf(stderr, "FAILED %s:%d: %s\n", __FILE__, __LINE__, #c); std::exit(1); } } while (0)
static bool near(float a, float b, float eps = 1e-6f) { return std::fabs(a - b) <= eps * std::fmax(1.0f, std::fabs(b)); }

int main() {
    ingest::Scene src;
    // root (translate x+10) -> child "A" (translate y+2, one mesh), child "L" (rotation about z by 90 deg, places a light), "Cam"
    src.nodes.resize(4);
    src.nodes[0].name = "root"; src.nodes[0].children = {1, 2, 3};
    src.nodes[0].transformRowMajor[3] = 10.0f;                       // a4 = tx (row-major)
    src.nodes[1].name = "A"; src.nodes[1].parent = 0; src.nodes[1].meshes = {0, 1};
    src.nodes[1].transformRowMajor[7] = 2.0f;                        // b4 = ty
    src.nodes[2].name = "L"; src.nodes[2].parent = 0;
    { float* m = src.nodes[2].transformRowMajor; m[0] = 0; m[1] = -1; m[4] = 1; m[5] = 0; m[3] = 1.0f; m[7] = 5.0f; m[11] = -3.0f; }
    src.nodes[3].name = "Cam"; src.nodes[3].parent = 0;
    { float* m = src.nodes[3].transformRowMajor; m[3] = 4.0f; m[7] = 6.0f; m[11] = 8.0f; }

    ingest::MeshSource tri;
    tri.positions = {0, 0, 0, 1, 0, 0, 0, 2, 0};
    tri.normals = {0, 0, 1, 0, 0, 1, 0, 0, 1};
    tri.texCoords0 = {0.25f, 0.75f, 0, 1, 0, 0, 0, 1, 0};
    tri.faces = {0, 1, 2};
    tri.materialIndex = 1;
    ingest::MeshSource bare;                                          // no normals, no uvs
    bare.positions = {0, 0, 3, -1, 0, 3, 0, -1, 3, 5, 5, 5};
    bare.faces = {0, 1, 2, 1, 2, 3};
    bare.materialIndex = 0;
    src.meshes = {tri, bare};

    src.materials.resize(2);
    src.materials[1].hasBaseColor = true; src.materials[1].baseColor[0] = 0.5f; src.materials[1].baseColor[3] = 0.25f;
    src.materials[1].hasIor = true; src.materials[1].ior = 2.0f;
    src.materials[1].hasMetallic = true; src.materials[1].metallic = 0.75f;
    src.materials[1].hasRoughness = true; src.materials[1].roughness = 0.3f;
    src.materials[1].alphaMode = "BLEND";
    src.materials[1].normalsTexture = static_cast<TextureHandle>(7);

    ingest::LightSource sun; sun.name = "L"; sun.type = ingest::Directional; sun.colorDiffuse[0] = 1366.0f; sun.colorDiffuse[1] = 683.0f; sun.colorDiffuse[2] = 0.0f;
    sun.direction[0] = 1; sun.direction[1] = 0; sun.direction[2] = 0;
    ingest::LightSource bulb; bulb.name = "L"; bulb.type = ingest::Point; bulb.colorDiffuse[0] = 0.5f; bulb.colorDiffuse[1] = 0.25f; bulb.colorDiffuse[2] = 0.125f;
    bulb.attenuationConstant = 0.5f;
    ingest::LightSource spot; spot.name = "L"; spot.type = ingest::Spot; spot.angleInnerCone = 0.2f; spot.angleOuterCone = 0.6f; spot.direction[0] = 0; spot.direction[1] = 1; spot.direction[2] = 0;
    ingest::LightSource sun2 = sun; sun2.colorDiffuse[0] = 0.9f; sun2.colorDiffuse[1] = 0.1f;
    src.lights = {sun, bulb, spot, sun2};
    src.cameras.push_back({"Cam", 0.5f});

    MeshAllocator meshes; SceneTree tree; pbr::MaterialSystem materials; pbr::LightSystem lights; Camera camera;
    std::vector<LightID> casters;
    const ingest::Result r = ingest::Load(src, meshes, tree, materials, lights, camera, &casters);

    // --- materials (AssimpSceneLoader.cpp:270-361): defaults kept when a key is absent; f0 from the index of refraction
    CHECK(r.materials.size() == 2 && r.materials[1].GetIndex() == 1 && r.materials[1].GetRaw() == ((1u << 4) | (1u << 2)));
    const pbr::MaterialProperties& m0 = materials.GetProperties()[0];
    const pbr::MaterialProperties& m1 = materials.GetProperties()[1];
    CHECK(m0.baseColor[0] == 1.0f && m0.f0 == 0.04f && m0.metallic == 0.0f && m0.perceptualRoughness == 0.5f && m0.normalsTexture == TextureHandle::Invalid);
    CHECK(m1.baseColor[0] == 0.5f && m1.baseColor[3] == 0.25f && near(m1.f0, 1.0f / 9.0f) && m1.metallic == 0.75f && m1.perceptualRoughness == 0.3f);
    CHECK(m1.normalsTexture == static_cast<TextureHandle>(7) && m1.baseColorTexture == TextureHandle::Invalid && m1.ambientOcclusionTexture == TextureHandle::Invalid);
    CHECK(r.alphaModes[0] == AlphaMode::eOpaque && r.alphaModes[1] == AlphaMode::eBlend);

    // --- geometry (:198-268): one scene node for "A", transform = root * A (column-major), absolute indices, v flipped
    CHECK(tree.GetNodeCount() == 1 && r.sceneNodes.size() == 1);
    const mat4 T = tree.GetTransform(r.sceneNodes[0]);
    CHECK(T.m[12] == 10.0f && T.m[13] == 2.0f && T.m[14] == 0.0f && T.m[0] == 1.0f && T.m[15] == 1.0f);
    CHECK(meshes.GetVertexCount() == 7 && meshes.GetIndexCount() == 9);
    const Vertex& v0 = meshes.GetVertices()[0];
    CHECK(v0.uv.x == 0.25f && v0.uv.y == -0.75f && v0.normal.z == 1.0f);
    const Vertex& v3 = meshes.GetVertices()[3];
    CHECK(v3.pos.z == 3.0f && v3.uv.x == 0.0f && v3.uv.y == 0.0f && v3.normal.x == 0.0f && v3.normal.y == 0.0f && v3.normal.z == 0.0f);   // missing streams = 0
    const uint32_t want[9] = {0, 1, 2, 3, 4, 5, 4, 5, 6};
    for (int i = 0; i < 9; ++i) CHECK(meshes.GetIndices()[i] == want[i]);
    int nMesh = 0;
    meshes.ForEachMesh([&](SceneNodeHandle node, Mesh mesh) {
        CHECK(node == r.sceneNodes[0]);
        if (nMesh == 0) CHECK(mesh.indexOffset == 0 && mesh.nbIndices == 3 && mesh.materialHandle == r.materials[1]);
        if (nMesh == 1) CHECK(mesh.indexOffset == 3 && mesh.nbIndices == 6 && mesh.materialHandle == r.materials[0]);
        ++nMesh;
    });
    CHECK(nMesh == 2);
    CHECK(near(r.maxVertexDist, std::sqrt(75.0f)));
    CHECK(r.boundingBox.min.x == 9.0f && r.boundingBox.max.x == 15.0f && r.boundingBox.min.y == 1.0f && r.boundingBox.max.y == 7.0f && r.boundingBox.max.z == 5.0f);

    // --- lights (:93-151)
    CHECK(lights.GetLightCount() == 4 && r.nbShadowCastingLights == 2 && casters.size() == 2 && casters[0] == 0 && casters[1] == 3);
    const pbr::Light& l0 = lights.GetLights()[0];
    CHECK(l0.type == 1 && near(l0.color[0], 1.0f) && near(l0.color[1], 0.5f) && l0.color[2] == 0.0f && near(l0.intensity, 20.0f));   // 10 * 1366 / 683
    CHECK(l0.position[0] == 11.0f && l0.position[1] == 5.0f && l0.position[2] == -3.0f);        // root * L translation
    CHECK(near(l0.direction[0], 0.0f) && near(l0.direction[1], 1.0f) && l0.direction[2] == 0.0f && l0.shadowIndex == 0);   // +x rotated by 90 deg about z
    const pbr::Light& l1 = lights.GetLights()[1];
    CHECK(l1.type == 2 && l1.color[0] == 0.5f && near(l1.intensity, 10.0f * 0.5f / 683.0f) && near(l1.falloffRadius, 2.0f));     // colour kept when max <= 1
    const pbr::Light& l2 = lights.GetLights()[2];
    CHECK(l2.type == 3 && near(l2.cosInnerAngle, std::cos(0.2f)) && near(l2.cosOuterAngle, std::cos(0.6f)) && near(l2.direction[0], -1.0f) && near(l2.direction[1], 0.0f));
    CHECK(lights.GetLights()[3].shadowIndex == 1);

    // --- camera (:153-178)
    CHECK(camera.GetEye().x == 14.0f && camera.GetEye().y == 6.0f && camera.GetEye().z == 8.0f);
    // up = the node's y axis, re-orthogonalised against the view direction by Camera::UpdateViewMatrix (Camera.h:96-108)
    CHECK(std::fabs(dot(camera.GetUpVector(), normalize(camera.GetEye()))) < 1e-6f && camera.GetUpVector().y > 0.5f);
    CHECK(near(camera.GetFieldOfView(), 0.5f * 180.0f / 3.14159265f * 2.0f, 1e-5f));
    {   // no camera in the file: orbit view at 15 x the largest vertex distance, up = +z
        ingest::Scene s2 = src; s2.cameras.clear();
        MeshAllocator me; SceneTree tr; pbr::MaterialSystem ma; pbr::LightSystem li; Camera cam;
        const ingest::Result r2 = ingest::Load(s2, me, tr, ma, li, cam);
        CHECK(near(cam.GetEye().x, 15.0f * r2.maxVertexDist) && near(cam.GetEye().z, cam.GetEye().x) && cam.GetUpVector().z > 0.5f);
    }
    std::puts("ingest OK");
    return 0;
}
