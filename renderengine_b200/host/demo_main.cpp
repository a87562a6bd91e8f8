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
#include "demo_scene.h"

#include <chrono>
#include <cstdio>
#include <cstdlib>

using namespace reb;

namespace {

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
        BuildGridScene(scene, nx, nz, level, W, H, S, translucentEvery);
        MeshAllocator& meshes = *scene.GetMeshAllocator();
        SceneTree& tree = *scene.GetSceneTree();
        MaterialSystem& materials = *scene.GetMaterialSystem();

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
