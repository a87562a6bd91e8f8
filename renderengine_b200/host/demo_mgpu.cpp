// demo_mgpu.cpp — a C++ multi-GPU caller of libreb.so: one process, one context (Renderer + RenderScene of the facade) per GPU,
// NCCL called directly on the device pointers the C ABI hands out.  The reference is single-GPU (one vk::Device, Device.h:12-13);
// this is the host code a maintainer would write to spread its frame loop (RenderScene.cpp:168-184) over the GPUs of one node.
//
//   reb_mgpu_demo <nGpus> <nx> <nz> <level> <width> <height> <shadowSize> <frames>
//
// Two schemes, each checked byte for byte against GPU 0 rendering the same frame(s) alone:
//   sort-first   every GPU owns one contiguous band of pixel rows (reb_set_partition) and one slab of shadow-map rows
//                (reb_set_shadow_partition); per frame: shadow slab -> reb_flush -> in-place ncclAllGather on the pointer
//                reb_get_output(REB_SHADOW_F32) returns -> reb_shadow_mark_valid -> forward band -> rows gathered on the host.
//                Run a second time with the shadow pass split by triangles instead (reb_set_shadow_split + ncclAllReduce min).
//   view batch   GPU g renders camera g of a ring of cameras; the shadow map is rendered once on GPU 0, ncclBroadcast on a
//                communication stream of its own, and handed to the forward pass with reb_shadow_wait_event so that the vertex
//                stage and triangle setup of every receiver overlap the transfer.
// Exit code 0 = every scheme reproduced the single-GPU frames exactly.
#include "RenderEngine.h"
#include "demo_scene.h"

#include <cuda_runtime.h>
#include <nccl.h>

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>

using namespace reb;

#define CK_CUDA(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { std::fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e_)); std::exit(1); } } while (0)
#define CK_NCCL(x) do { ncclResult_t r_ = (x); if (r_ != ncclSuccess) { std::fprintf(stderr, "%s: %s\n", #x, ncclGetErrorString(r_)); std::exit(1); } } while (0)

namespace {

struct Gpu {
    int device = 0;
    cudaStream_t stream = nullptr, commStream = nullptr;
    cudaEvent_t shadowRendered = nullptr, shadowArrived = nullptr;
    std::unique_ptr<Renderer> renderer;
    std::unique_ptr<RenderScene> scene;
    float* shadow = nullptr;   // device pointer of the context's shadow map
    vec3 eye0;                 // the framing camera the scene was built with
    std::vector<uint8_t> frame;
};

void make_gpu(Gpu& g, int device, int nx, int nz, int level, uint32_t W, uint32_t H, uint32_t S) {
    g.device = device;
    CK_CUDA(cudaSetDevice(device));
    CK_CUDA(cudaStreamCreateWithFlags(&g.stream, cudaStreamNonBlocking));
    CK_CUDA(cudaStreamCreateWithFlags(&g.commStream, cudaStreamNonBlocking));
    CK_CUDA(cudaEventCreateWithFlags(&g.shadowRendered, cudaEventDisableTiming));
    CK_CUDA(cudaEventCreateWithFlags(&g.shadowArrived, cudaEventDisableTiming));
    g.renderer.reset(new Renderer(W, H, device, REB_OUT_COLOR_SRGB8, g.stream));
    g.scene.reset(new RenderScene(*g.renderer, S));
    BuildGridScene(*g.scene, nx, nz, level, W, H, S);
    g.scene->Init();
    g.scene->Update();
    g.shadow = g.renderer->GetShadowMapDevicePointer();
    g.frame.resize((size_t)W * H * 4);
    g.eye0 = g.scene->GetCameraViewSystem()->GetCamera().GetEye();
}

void orbit(Gpu& g, int view, int nViews) {   // camera `view` of a ring around the grid (same height and distance as the framing camera)
    Camera& cam = g.scene->GetCameraViewSystem()->GetCamera();
    const vec3 p0 = g.eye0;
    const float r = std::sqrt(p0.x * p0.x + p0.z * p0.z), a0 = std::atan2(p0.z, p0.x), a = a0 + 6.28318530718f * (float)view / (float)nViews;
    cam.SetCameraView(vec3(r * std::cos(a), p0.y, r * std::sin(a)), vec3(0, 0, 0), vec3(0, 1, 0));
}

std::vector<MeshDrawInfo> shadow_list(const RenderScene& s) {   // RenderScene.cpp:191-208: opaque ++ translucent
    std::vector<MeshDrawInfo> d(s.GetOpaqueMeshes());
    d.insert(d.end(), s.GetTranslucentMeshes().begin(), s.GetTranslucentMeshes().end());
    return d;
}

size_t diff_bytes(const std::vector<uint8_t>& a, const std::vector<uint8_t>& b) {
    size_t n = 0;
    for (size_t i = 0; i < a.size(); ++i) n += a[i] != b[i];
    return n;
}

double now_ms() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

}  // namespace

int main(int argc, char** argv) {
    if (argc < 9) { std::fprintf(stderr, "usage: %s nGpus nx nz level width height shadowSize frames\n", argv[0]); return 2; }
    int N = std::atoi(argv[1]);
    const int nx = std::atoi(argv[2]), nz = std::atoi(argv[3]), level = std::atoi(argv[4]);
    const uint32_t W = (uint32_t)std::atoi(argv[5]), H = (uint32_t)std::atoi(argv[6]), S = (uint32_t)std::atoi(argv[7]);
    const int frames = std::max(1, std::atoi(argv[8]));
    int have = 0;
    CK_CUDA(cudaGetDeviceCount(&have));
    if (N <= 0 || N > have) { std::fprintf(stderr, "reb_mgpu_demo: %d GPUs asked, %d visible: using %d\n", N, have, have); N = have; }
    if (N < 1) { std::fprintf(stderr, "reb_mgpu_demo: no CUDA device\n"); return 1; }
    try {
        std::vector<Gpu> gpus((size_t)N);
        std::vector<int> devs((size_t)N);
        for (int g = 0; g < N; ++g) { devs[g] = g; make_gpu(gpus[g], g, nx, nz, level, W, H, S); }
        std::vector<ncclComm_t> comms((size_t)N);
        CK_NCCL(ncclCommInitAll(comms.data(), N, devs.data()));
        Gpu solo;   // the single-GPU frame everything is compared with: an unpartitioned context on GPU 0
        make_gpu(solo, 0, nx, nz, level, W, H, S);
        const size_t nTris = solo.scene->GetMeshAllocator()->GetIndexCount() / 3;
        std::vector<uint8_t> want((size_t)W * H * 4), got((size_t)W * H * 4);
        size_t bad = 0;

        // ---- single GPU --------------------------------------------------------------------------------------------------------
        CK_CUDA(cudaSetDevice(0));
        solo.scene->SetShadowsDirty(); solo.scene->Update(); solo.scene->Render();   // first frame: draw-list preprocessing (once per list), module loading
        solo.renderer->WaitIdle();
        double t0 = now_ms();
        for (int f = 0; f < frames; ++f) { solo.scene->SetShadowsDirty(); solo.scene->Update(); solo.scene->Render(); }
        solo.renderer->WaitIdle();
        const double soloMs = (now_ms() - t0) / frames;
        solo.renderer->Readback(REB_COLOR_SRGB8, want.data());

        // ---- sort-first: row bands + shadow slabs + all-gather ---------------------------------------------------------------------
        const uint32_t bandH = ((H + (uint32_t)N - 1) / (uint32_t)N + 7u) / 8u * 8u;          // one contiguous band per GPU
        const bool slabs = N > 1 && S % (32u * (uint32_t)N) == 0;                             // reb_set_shadow_partition's granularity
        const uint32_t slabRows = slabs ? S / (uint32_t)N : S;
        std::printf("reb_mgpu_demo: %zu triangles %ux%u shadow %u^2, %d GPU(s)\n", nTris, W, H, S, N);
        for (int mode = 0; mode < 2; ++mode) {   // 0: shadow pass in row slabs + all-gather; 1: shares of the triangle stream + all-reduce (min)
            const bool byTriangles = mode == 1;
            for (int g = 0; g < N; ++g) {
                CK_CUDA(cudaSetDevice(g));
                gpus[g].renderer->SetPartition((uint32_t)g, (uint32_t)N, bandH);
                if (byTriangles) gpus[g].renderer->SetShadowSplit((uint32_t)g, (uint32_t)N);
                else if (slabs) gpus[g].renderer->SetShadowPartition((uint32_t)g, (uint32_t)N, slabRows);
            }
            const bool exchange = byTriangles ? N > 1 : slabs;
            double splitMs = 0.0;
            for (int f = 0; f < frames + 1; ++f) {   // frame 0 warms NCCL up
                if (f == 1) { for (int g = 0; g < N; ++g) gpus[g].renderer->WaitIdle(); t0 = now_ms(); }
                for (int g = 0; g < N; ++g) {
                    CK_CUDA(cudaSetDevice(g));
                    gpus[g].scene->Update();
                    gpus[g].scene->GetShadowSystem()->Render(shadow_list(*gpus[g].scene));
                    gpus[g].renderer->Submit();                                                    // reb_flush: this GPU's part of the map is on the stream
                }
                if (exchange) {
                    CK_NCCL(ncclGroupStart());
                    for (int g = 0; g < N; ++g) {
                        if (byTriangles) CK_NCCL(ncclAllReduce(gpus[g].shadow, gpus[g].shadow, (size_t)S * S, ncclFloat, ncclMin, comms[g], gpus[g].stream));
                        else CK_NCCL(ncclAllGather(gpus[g].shadow + (size_t)g * slabRows * S, gpus[g].shadow, (size_t)slabRows * S, ncclFloat, comms[g], gpus[g].stream));
                    }
                    CK_NCCL(ncclGroupEnd());
                }
                for (int g = 0; g < N; ++g) {
                    CK_CUDA(cudaSetDevice(g));
                    if (exchange) gpus[g].renderer->ShadowMapFilled();
                    gpus[g].scene->GetMaterialSystem()->Draw(*gpus[g].renderer, gpus[g].scene->GetOpaqueMeshes(), gpus[g].scene->GetTranslucentMeshes());
                }
            }
            for (int g = 0; g < N; ++g) { CK_CUDA(cudaSetDevice(g)); gpus[g].renderer->WaitIdle(); }
            splitMs = (now_ms() - t0) / frames;
            for (int g = 0; g < N; ++g) {            // compose: rows of band g come from GPU g
                CK_CUDA(cudaSetDevice(g));
                gpus[g].renderer->Readback(REB_COLOR_SRGB8, gpus[g].frame.data());
                for (uint32_t y = 0; y < H; ++y)
                    if ((y / bandH) % (uint32_t)N == (uint32_t)g) std::memcpy(&got[(size_t)y * W * 4], &gpus[g].frame[(size_t)y * W * 4], (size_t)W * 4);
            }
            const size_t d = diff_bytes(got, want);
            bad += d;
            std::printf("  sort-first  : %zu differing bytes vs the single-GPU frame; %.3f ms/frame (wall) vs %.3f ms single GPU; shadow %s\n", d, splitMs, soloMs,
                        byTriangles ? (exchange ? "split by triangles + ncclAllReduce(min)" : "split off (one GPU)") : slabs ? "slabs + ncclAllGather" : "rendered whole on every GPU");
            for (int g = 0; g < N; ++g) { CK_CUDA(cudaSetDevice(g)); gpus[g].renderer->SetShadowSplit(0u, 1u); }
        }

        // ---- view batch: shadow once on GPU 0 + broadcast overlapped with the receivers' geometry stages -----------------------------
        for (int g = 0; g < N; ++g) {
            CK_CUDA(cudaSetDevice(g));
            gpus[g].renderer->SetPartition(0u, 1u, 64u);
            gpus[g].renderer->SetShadowPartition(0u, 1u, 64u);
        }
        const mat4 sv = gpus[0].scene->GetShadowSystem()->GetShadowView(0), sp = gpus[0].scene->GetShadowSystem()->GetShadowProj(0);   // the batch shares GPU 0's light frustum
        double batchMs = 0.0;
        for (int f = 0; f < frames + 1; ++f) {
            if (f == 1) { for (int g = 0; g < N; ++g) gpus[g].renderer->WaitIdle(); t0 = now_ms(); }
            CK_CUDA(cudaSetDevice(0));
            gpus[0].renderer->Check(reb_shadow_set_transform(gpus[0].renderer->GetContext(), 0, sv.data(), sp.data()), "reb_shadow_set_transform");
            gpus[0].scene->GetShadowSystem()->Render(shadow_list(*gpus[0].scene));
            gpus[0].renderer->Submit();
            CK_CUDA(cudaEventRecord(gpus[0].shadowRendered, gpus[0].stream));
            CK_CUDA(cudaStreamWaitEvent(gpus[0].commStream, gpus[0].shadowRendered, 0));
            CK_NCCL(ncclGroupStart());
            for (int g = 0; g < N; ++g) CK_NCCL(ncclBroadcast(gpus[g].shadow, gpus[g].shadow, (size_t)S * S, ncclFloat, 0, comms[g], gpus[g].commStream));
            CK_NCCL(ncclGroupEnd());
            for (int g = 0; g < N; ++g) {
                CK_CUDA(cudaSetDevice(g));
                reb_ctx* c = gpus[g].renderer->GetContext();
                CK_CUDA(cudaEventRecord(gpus[g].shadowArrived, gpus[g].commStream));
                orbit(gpus[g], g, N);
                gpus[g].scene->GetCameraViewSystem()->Update(*gpus[g].renderer);
                gpus[g].renderer->Check(reb_shadow_set_transform(c, 0, sv.data(), sp.data()), "reb_shadow_set_transform");
                gpus[g].renderer->ShadowMapArrivesWith(gpus[g].shadowArrived);
                gpus[g].scene->GetMaterialSystem()->Draw(*gpus[g].renderer, gpus[g].scene->GetOpaqueMeshes(), gpus[g].scene->GetTranslucentMeshes());
            }
            for (int g = 0; g < N; ++g) { CK_CUDA(cudaSetDevice(g)); gpus[g].renderer->WaitIdle(); }   // the next frame's shadow pass overwrites the map
        }
        batchMs = (now_ms() - t0) / frames;
        size_t dBatch = 0;
        for (int g = 0; g < N; ++g) {
            CK_CUDA(cudaSetDevice(g));
            gpus[g].renderer->Readback(REB_COLOR_SRGB8, gpus[g].frame.data());
            CK_CUDA(cudaSetDevice(0));           // the same view alone on GPU 0, its own shadow pass with the shared light frustum
            orbit(solo, g, N);
            solo.scene->GetCameraViewSystem()->Update(*solo.renderer);
            solo.renderer->Check(reb_shadow_set_transform(solo.renderer->GetContext(), 0, sv.data(), sp.data()), "reb_shadow_set_transform");
            solo.scene->GetShadowSystem()->Render(shadow_list(*solo.scene));
            solo.scene->GetMaterialSystem()->Draw(*solo.renderer, solo.scene->GetOpaqueMeshes(), solo.scene->GetTranslucentMeshes());
            solo.renderer->Readback(REB_COLOR_SRGB8, want.data());
            dBatch += diff_bytes(gpus[g].frame, want);
        }
        bad += dBatch;
        std::printf("  view batch  : %zu differing bytes over %d views vs GPU 0 alone; %.3f ms per %d-view frame (wall); shadow once + ncclBroadcast behind reb_shadow_wait_event\n",
                    dBatch, N, batchMs, N);
        for (int g = 0; g < N; ++g) ncclCommDestroy(comms[g]);
        std::printf("  (wall-clock of ONE host thread driving all GPUs in turn through the facade - per-frame host work included; device-timed\n"
                    "   scaling is bench.py's job: one process per GPU)\n");
        std::printf("reb_mgpu_demo: %s\n", bad == 0 ? "PASS" : "FAIL");
        return bad == 0 ? 0 : 1;
    } catch (const std::exception& e) {
        std::fprintf(stderr, "reb_mgpu_demo: %s\n", e.what());
        return 1;
    }
}
