// demo_scene.h — the procedural icosphere-grid scene of the host demos (reb_host_demo, reb_mgpu_demo), built through the facade
// the way AssimpSceneLoader.cpp:198-261 fills the reference's systems: one node + one mesh per sphere with its own vertex block,
// a ground quad, one directional light with a shadow map, a camera on the (1,1,1) diagonal framing the grid.
#pragma once
#include "RenderEngine.h"

#include <array>
#include <cmath>
#include <map>
#include <vector>

namespace reb {

struct Ico { std::vector<vec3> v; std::vector<std::array<uint32_t, 3>> t; };

inline Ico icosphere(int level) {
    const float g = (1.0f + std::sqrt(5.0f)) / 2.0f;
    Ico s;
    const float raw[12][3] = {{-1, g, 0}, {1, g, 0}, {-1, -g, 0}, {1, -g, 0}, {0, -1, g}, {0, 1, g}, {0, -1, -g}, {0, 1, -g}, {g, 0, -1}, {g, 0, 1}, {-g, 0, -1}, {-g, 0, 1}};
    for (auto& r : raw) s.v.push_back(normalize(vec3(r[0], r[1], r[2])));
    const uint32_t tri[20][3] = {{0, 11, 5}, {0, 5, 1}, {0, 1, 7}, {0, 7, 10}, {0, 10, 11}, {1, 5, 9}, {5, 11, 4}, {11, 10, 2}, {10, 7, 6}, {7, 1, 8},
                                 {3, 9, 4}, {3, 4, 2}, {3, 2, 6}, {3, 6, 8}, {3, 8, 9}, {4, 9, 5}, {2, 4, 11}, {6, 2, 10}, {8, 6, 7}, {9, 8, 1}};
    for (auto& t : tri) s.t.push_back({t[0], t[1], t[2]});
    for (int l = 0; l < level; ++l) {
        std::map<std::pair<uint32_t, uint32_t>, uint32_t> mid;
        auto midpoint = [&](uint32_t a, uint32_t b) {
            const auto key = std::make_pair(std::min(a, b), std::max(a, b));
            auto it = mid.find(key);
            if (it != mid.end()) return it->second;
            s.v.push_back(normalize(s.v[a] + s.v[b]));
            return mid[key] = (uint32_t)s.v.size() - 1;
        };
        std::vector<std::array<uint32_t, 3>> nt;
        for (auto& t : s.t) {
            const uint32_t a = midpoint(t[0], t[1]), b = midpoint(t[1], t[2]), c = midpoint(t[2], t[0]);
            nt.push_back({t[0], a, c}); nt.push_back({t[1], b, a}); nt.push_back({t[2], c, b}); nt.push_back({a, b, c});
        }
        s.t.swap(nt);
    }
    return s;
}

inline void BuildGridScene(RenderScene& scene, int nx, int nz, int level, uint32_t W, uint32_t H, uint32_t S, int translucentEvery = 0) {
    MeshAllocator& meshes = *scene.GetMeshAllocator();
    SceneTree& tree = *scene.GetSceneTree();
    MaterialSystem& materials = *scene.GetMaterialSystem();

    // materials: a few opaque Phong instances + the ground
    std::vector<MaterialHandle> mats;
    const float palette[5][3] = {{0.9f, 0.3f, 0.25f}, {0.25f, 0.8f, 0.35f}, {0.3f, 0.4f, 0.9f}, {0.85f, 0.8f, 0.3f}, {0.7f, 0.35f, 0.8f}};
    for (int m = 0; m < 5; ++m) {
        MaterialInstanceInfo info;
        for (int c = 0; c < 3; ++c) { info.properties.diffuse[c] = palette[m][c]; info.properties.specular[c] = 0.5f; }
        info.properties.shininess = (float)(8 << (m % 4));
        if (translucentEvery > 0 && m % translucentEvery == 0) { info.properties.diffuse[3] = 0.5f; info.pipelineProperties.alphaMode = AlphaMode::eBlend; }
        mats.push_back(materials.CreateMaterialInstance(info));
    }
    MaterialInstanceInfo groundInfo;
    for (int c = 0; c < 3; ++c) { groundInfo.properties.diffuse[c] = 0.8f; groundInfo.properties.specular[c] = 0.2f; }
    groundInfo.properties.shininess = 16.0f;
    const MaterialHandle groundMat = materials.CreateMaterialInstance(groundInfo);

    // geometry: one node + one mesh per sphere, each with its own vertex block (as AssimpSceneLoader appends them)
    const Ico ico = icosphere(level);
    const float spacing = 2.5f;
    BoundingBox unitBox; unitBox.min = vec3(-1.0f); unitBox.max = vec3(1.0f);
    for (int ix = 0; ix < nx; ++ix)
        for (int iz = 0; iz < nz; ++iz) {
            const vec3 centre((ix - (nx - 1) / 2.0f) * spacing, 0.0f, (iz - (nz - 1) / 2.0f) * spacing);
            const SceneNodeHandle node = tree.CreateNode(mat4::translation(centre), unitBox);
            const uint32_t base = (uint32_t)meshes.GetVertexCount();
            Mesh mesh; mesh.indexOffset = meshes.GetIndexCount(); mesh.nbIndices = ico.t.size() * 3; mesh.materialHandle = mats[(ix * nz + iz) % mats.size()];
            for (const vec3& p : ico.v) {
                Vertex v; v.pos = p; v.normal = p;
                v.uv.x = std::atan2(p.z, p.x) / 6.28318530718f + 0.5f; v.uv.y = std::asin(std::fmax(-1.0f, std::fmin(1.0f, p.y))) / 3.14159265359f + 0.5f;
                meshes.AddVertex(v);
            }
            for (auto& t : ico.t) for (int k = 0; k < 3; ++k) meshes.AddIndex(base + t[k]);
            meshes.GroupMeshes(node, {mesh});
        }
    const float ex = (nx - 1) / 2.0f * spacing + 1.0f, ez = (nz - 1) / 2.0f * spacing + 1.0f;
    {   // ground quad under the grid (shadow receiver)
        const float hx = ex + 1.5f, hz = ez + 1.5f, gy = -1.05f;
        BoundingBox gb; gb.min = vec3(-hx, gy, -hz); gb.max = vec3(hx, gy, hz);
        const SceneNodeHandle node = tree.CreateNode(mat4::identity(), gb);
        const uint32_t base = (uint32_t)meshes.GetVertexCount();
        const float q[4][2] = {{-hx, -hz}, {-hx, hz}, {hx, hz}, {hx, -hz}};
        for (auto& c : q) { Vertex v; v.pos = vec3(c[0], gy, c[1]); v.normal = vec3(0, 1, 0); v.uv.x = c[0] > 0; v.uv.y = c[1] > 0; meshes.AddVertex(v); }
        Mesh mesh; mesh.indexOffset = meshes.GetIndexCount(); mesh.nbIndices = 6; mesh.materialHandle = groundMat;
        for (uint32_t i : {0u, 1u, 2u, 0u, 2u, 3u}) meshes.AddIndex(base + i);
        meshes.GroupMeshes(node, {mesh});
    }

    // one directional light with a shadow map (AssimpSceneLoader.cpp:124-147 does this per directional light)
    Light light;
    const vec3 ldir = normalize(vec3(-0.3f, -1.0f, -0.2f));
    light.direction[0] = ldir.x; light.direction[1] = ldir.y; light.direction[2] = ldir.z;
    light.position[0] = -50.0f * ldir.x; light.position[1] = -50.0f * ldir.y; light.position[2] = -50.0f * ldir.z;
    for (int c = 0; c < 4; ++c) { light.ambient[c] = 0.1f; light.diffuse[c] = 1.0f; light.specular[c] = 1.0f; }
    const LightID lightId = scene.GetLightSystem()->AddLight(light);
    if (S) scene.GetLightSystem()->SetLightShadowID(lightId, scene.GetShadowSystem()->CreateShadowMap(lightId));

    // camera on the (1,1,1) diagonal framing the grid
    const float radius = std::sqrt(ex * ex + ez * ez), dist = std::fmin(0.85f * radius / std::sin(radians(22.5f)), 100.0f - radius - 1.0f);
    Camera& cam = scene.GetCameraViewSystem()->GetCamera();
    cam.SetImageExtent((int)W, (int)H);
    cam.SetCameraView(normalize(vec3(1, 1, 1)) * dist, vec3(0, 0, 0), vec3(0, 1, 0));
}

}  // namespace reb
