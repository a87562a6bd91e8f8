// demo_main.cpp — a host program written against the facade exactly as one would write it against the
// reference Runtime (compare Source/Samples/MainSample/main.cpp:52-83 + AssimpSceneLoader.cpp:198-261):
// build a procedural icosphere-grid scene through MeshAllocator / SceneTree / MaterialSystem /
// LightSystem / ShadowSystem, RenderScene::Init, then Update + Render per frame.
//
//   reb_host_demo <nx> <nz> <level> <width> <height> <shadowSize> <frames> [dump.bin]
//
// With a dump path it writes every hot-path input and the rendered buffers of the last frame, so the parity
// test can feed the same inputs to the test oracle (tests/test_host_facade.py).  Prints the wall time per frame.
#include "RenderEngine.h"

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <map>

using namespace reb;

namespace {

struct Ico { std::vector<vec3> v; std::vector<std::array<uint32_t, 3>> t; };

Ico icosphere(int level) {
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

template <class T> void put(FILE* f, const T* p, size_t n) { std::fwrite(p, sizeof(T), n, f); }

}  // namespace

int main(int argc, char** argv) {
    if (argc < 8) { std::fprintf(stderr, "usage: %s nx nz level width height shadowSize frames [dump.bin [skyFaceSize [translucentEvery]]]\n", argv[0]); return 2; }
    const int nx = std::atoi(argv[1]), nz = std::atoi(argv[2]), level = std::atoi(argv[3]);
    const uint32_t W = (uint32_t)std::atoi(argv[4]), H = (uint32_t)std::atoi(argv[5]), S = (uint32_t)std::atoi(argv[6]);
    const int frames = std::atoi(argv[7]);
    const char* dump = argc > 8 ? argv[8] : nullptr;
    const uint32_t skyN = argc > 9 ? (uint32_t)std::atoi(argv[9]) : 0u;
    const int translucentEvery = argc > 10 ? std::atoi(argv[10]) : 0;   // every k-th material is AlphaMode::eBlend with alpha 0.5
    try {
        Renderer renderer(W, H);
        RenderScene scene(renderer, S ? S : 4096);
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

        // optional cube-map background (Skybox.cpp:58-92): a procedural gradient, one tone per face
        std::vector<uint16_t> skyFaces((size_t)6 * skyN * skyN * 4);
        for (uint32_t face = 0; face < 6 && skyN; ++face)
            for (uint32_t j = 0; j < skyN; ++j)
                for (uint32_t i = 0; i < skyN; ++i) {
                    uint16_t* t = &skyFaces[(((size_t)face * skyN + j) * skyN + i) * 4];
                    t[0] = (uint16_t)(65535u * i / std::max(skyN - 1u, 1u)); t[1] = (uint16_t)(65535u * j / std::max(skyN - 1u, 1u));
                    t[2] = (uint16_t)(9000u * (face + 1u)); t[3] = 65535u;
                }
        if (skyN) scene.GetSkybox()->SetCubeMapFaces(skyFaces.data(), skyN, skyN);

        scene.Init();
        const auto t0 = std::chrono::steady_clock::now();
        for (int f = 0; f < frames; ++f) {
            scene.SetShadowsDirty();   // north-star mode: shadow map every frame
            scene.Update();
            scene.Render();
        }
        renderer.WaitIdle();
        const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count() / std::max(frames, 1);
        const size_t nTris = meshes.GetIndexCount() / 3;
        std::printf("reb_host_demo: %zu triangles, %ux%u, shadow %u^2: %.3f ms/frame (wall, %d frames)\n", nTris, W, H, S, ms, frames);

        if (dump) {
            FILE* f = std::fopen(dump, "wb");
            if (!f) { std::perror("dump"); return 1; }
            std::vector<MeshDrawInfo> all(scene.GetOpaqueMeshes());
            all.insert(all.end(), scene.GetTranslucentMeshes().begin(), scene.GetTranslucentMeshes().end());
            const std::vector<reb_draw> draws = Renderer::ToDraws(all);
            const uint32_t hdr[10] = {0x52454244u, W, H, S, (uint32_t)meshes.GetVertexCount(), (uint32_t)meshes.GetIndexCount(), (uint32_t)tree.GetNodeCount(),
                                      (uint32_t)draws.size(), (uint32_t)materials.GetMaterialInstanceCount(), scene.GetLightSystem()->GetLightCount()};
            put(f, hdr, 10);
            put(f, meshes.GetVertices().data(), meshes.GetVertices().size());
            put(f, meshes.GetIndices().data(), meshes.GetIndices().size());
            put(f, tree.GetTransforms().data(), tree.GetTransforms().size());
            put(f, draws.data(), draws.size());
            put(f, materials.GetProperties().data(), materials.GetProperties().size());
            put(f, scene.GetLightSystem()->GetLights().data(), scene.GetLightSystem()->GetLights().size());
            put(f, &scene.GetCameraViewSystem()->GetViewProperties(), 1);
            const mat4 sv = S ? scene.GetShadowSystem()->GetShadowView(0) : mat4(), sp = S ? scene.GetShadowSystem()->GetShadowProj(0) : mat4();
            put(f, sv.data(), 16); put(f, sp.data(), 16);
            std::vector<float> shadow((size_t)S * S), depth((size_t)W * H), colf((size_t)W * H * 4);
            std::vector<uint32_t> tri((size_t)W * H); std::vector<uint8_t> col8((size_t)W * H * 4);
            if (S) renderer.Readback(REB_SHADOW_F32, shadow.data());
            renderer.Readback(REB_DEPTH_F32, depth.data()); renderer.Readback(REB_TRIID_U32, tri.data());
            renderer.Readback(REB_COLOR_SRGB8, col8.data()); renderer.Readback(REB_COLOR_F32, colf.data());
            put(f, shadow.data(), shadow.size()); put(f, depth.data(), depth.size()); put(f, tri.data(), tri.size());
            put(f, col8.data(), col8.size()); put(f, colf.data(), colf.size());
            const uint32_t trailer[2] = {skyN, (uint32_t)scene.GetTranslucentMeshes().size()};   // the last trailer[1] draws are the translucent list
            put(f, trailer, 2);
            if (skyN) put(f, skyFaces.data(), skyFaces.size());
            std::fclose(f);
        }
    } catch (const std::exception& e) {
        std::fprintf(stderr, "reb_host_demo: %s\n", e.what());
        return 1;
    }
    return 0;
}
