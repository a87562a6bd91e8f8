// test_golden_math.cpp — the host facade's arithmetic (reb_math.h, reb::Camera, reb::BoundingBox, ComputeShadowTransform) against
// vectors produced by the REFERENCE's own code (Camera.cpp, BoundingBox.cpp, glm — tests/golden/gen_golden.py and
// gen_golden_shadow.py).  Bit for bit: a caller that switches to the facade must get the matrices the reference computes.
// Input: the flat float32 / int32 dump tests/test_golden_host.py writes from the committed .npz files.  CPU only, no device call.
#include "../../renderengine_b200/host/RenderEngine.h"

#include <cstdio>
#include <cstring>
#include <vector>

using namespace reb;

static std::vector<unsigned char> g_buf;
static size_t g_off = 0;
static const float* takef(size_t n) { const float* p = reinterpret_cast<const float*>(g_buf.data() + g_off); g_off += 4 * n; return p; }
static const int* takei(size_t n) { const int* p = reinterpret_cast<const int*>(g_buf.data() + g_off); g_off += 4 * n; return p; }
static int g_bad = 0;
static void same(const char* what, int i, const float* got, const float* want, int n) {
    if (std::memcmp(got, want, 4 * (size_t)n) != 0) {
        if (g_bad < 12) {
            std::fprintf(stderr, "%s[%d] differs:", what, i);
            for (int k = 0; k < n; ++k) if (std::memcmp(got + k, want + k, 4)) std::fprintf(stderr, " [%d] %.9g != %.9g", k, got[k], want[k]);
            std::fprintf(stderr, "\n");
        }
        ++g_bad;
    }
}
static vec3 v3(const float* p) { return vec3(p[0], p[1], p[2]); }
static mat4 m4(const float* p) { mat4 m; std::memcpy(m.m, p, 64); return m; }

int main(int argc, char** argv) {
    if (argc < 2) return 2;
    FILE* f = std::fopen(argv[1], "rb");
    if (!f) return 2;
    std::fseek(f, 0, SEEK_END); g_buf.resize((size_t)std::ftell(f)); std::fseek(f, 0, SEEK_SET);
    if (std::fread(g_buf.data(), 1, g_buf.size(), f) != g_buf.size()) return 2;
    std::fclose(f);
    const int K = takei(1)[0], N = takei(1)[0], KS = takei(1)[0], NB = takei(1)[0];
    // ---- Camera (Camera.h:93-117, Camera.cpp:3-25) ---------------------------------------------------------------------
    const float *eye = takef(3 * K), *look = takef(3 * K), *up = takef(3 * K);
    const int* dims = takei(2 * K);
    const float *camView = takef(16 * K), *camProj = takef(16 * K), *camCorners = takef(24 * K);
    for (int i = 0; i < K; ++i) {
        Camera cam(v3(eye + 3 * i), v3(look + 3 * i), v3(up + 3 * i), 45.0f, 0.01f, 100.0f, dims[2 * i], dims[2 * i + 1]);
        same("Camera::GetViewMatrix", i, cam.GetViewMatrix().data(), camView + 16 * i, 16);
        same("Camera::GetProjectionMatrix", i, cam.GetProjectionMatrix().data(), camProj + 16 * i, 16);
        const std::vector<vec3> c = cam.ComputeFrustrumCorners();
        for (int k = 0; k < 8; ++k) { const float g[3] = {c[k].x, c[k].y, c[k].z}; same("Camera::ComputeFrustrumCorners", 8 * i + k, g, camCorners + 24 * i + 3 * k, 3); }
        Camera moved;   // the setters reach the same state
        moved.SetCameraView(v3(eye + 3 * i), v3(look + 3 * i), v3(up + 3 * i)); moved.SetImageExtent(dims[2 * i], dims[2 * i + 1]);
        same("Camera::SetCameraView", i, moved.GetViewMatrix().data(), camView + 16 * i, 16);
        same("Camera::SetImageExtent", i, moved.GetProjectionMatrix().data(), camProj + 16 * i, 16);
    }
    // ---- BoundingBox (BoundingBox.h:47-60, BoundingBox.cpp:3-16), lookAt, kClip * ortho, inverse ---------------------------------
    const float *bbMin = takef(3 * K), *bbMax = takef(3 * K), *bbOut = takef(12 * K);
    const float *laDir = takef(3 * K), *laUp = takef(3 * K), *laOut = takef(16 * K), *orthoOut = takef(16 * K), *invOut = takef(16 * K);
    for (int i = 0; i < K; ++i) {
        BoundingBox b; b.min = v3(bbMin + 3 * i); b.max = v3(bbMax + 3 * i);
        const mat4 V = m4(camView + 16 * i);
        const BoundingBox t0 = b.Transform(V), t1 = V * b;
        const float g0[6] = {t0.min.x, t0.min.y, t0.min.z, t0.max.x, t0.max.y, t0.max.z}, g1[6] = {t1.min.x, t1.min.y, t1.min.z, t1.max.x, t1.max.y, t1.max.z};
        same("BoundingBox::Transform", i, g0, bbOut + 12 * i, 6);
        same("mat4 * BoundingBox", i, g1, bbOut + 12 * i + 6, 6);
        same("lookAt", i, lookAt(vec3(0.0f, 0.0f, 0.0f), v3(laDir + 3 * i), v3(laUp + 3 * i)).data(), laOut + 16 * i, 16);
        const float* mn = bbMin + 3 * i; const float* mx = bbMax + 3 * i;
        same("kClip * ortho", i, (kClip() * ortho(mn[0], mx[0], mn[1], mx[1], mn[2], mx[2])).data(), orthoOut + 16 * i, 16);
        same("inverse", i, inverse(V).data(), invOut + 16 * i, 16);
    }
    // ---- glm products (type_mat4x4.inl) -------------------------------------------------------------------------------------
    const float *A = takef(16 * N), *B = takef(16 * N), *Vv = takef(4 * N), *MM = takef(16 * N), *MV = takef(4 * N);
    for (int i = 0; i < N; ++i) {
        same("mat4 * mat4", i, (m4(A + 16 * i) * m4(B + 16 * i)).data(), MM + 16 * i, 16);
        const vec4 r = m4(A + 16 * i) * vec4(Vv[4 * i], Vv[4 * i + 1], Vv[4 * i + 2], Vv[4 * i + 3]);
        const float g[4] = {r.x, r.y, r.z, r.w};
        same("mat4 * vec4", i, g, MV + 4 * i, 4);
    }
    // ---- ComputeShadowTransform (ShadowSystem.cpp:129-204) --------------------------------------------------------------------
    const float *ld = takef(3 * KS), *se = takef(3 * KS), *sl = takef(3 * KS), *su = takef(3 * KS);
    const int* sd = takei(2 * KS);
    const float *sf = takef(KS), *boxes = takef((size_t)6 * NB * KS), *tr = takef((size_t)16 * NB * KS), *smin = takef(3 * KS), *smax = takef(3 * KS);
    const float *sView = takef(16 * KS), *sProj = takef(16 * KS), *sXf = takef(16 * KS);
    for (int i = 0; i < KS; ++i) {
        Camera cam(v3(se + 3 * i), v3(sl + 3 * i), v3(su + 3 * i), sf[i], 0.01f, 100.0f, sd[2 * i], sd[2 * i + 1]);
        std::vector<BoundingBox> bx(NB); std::vector<mat4> tm(NB);
        for (int k = 0; k < NB; ++k) {
            bx[k].min = v3(boxes + 6 * ((size_t)i * NB + k)); bx[k].max = v3(boxes + 6 * ((size_t)i * NB + k) + 3);
            tm[k] = m4(tr + 16 * ((size_t)i * NB + k));
        }
        BoundingBox sb; sb.min = v3(smin + 3 * i); sb.max = v3(smax + 3 * i);
        const ShadowTransform t = ComputeShadowTransform(v3(ld + 3 * i), cam, sb, bx, tm);
        same("ComputeShadowTransform.view", i, t.view.data(), sView + 16 * i, 16);
        same("ComputeShadowTransform.proj", i, t.proj.data(), sProj + 16 * i, 16);
        same("proj * view (ShadowSystem.cpp:303)", i, (t.proj * t.view).data(), sXf + 16 * i, 16);
    }
    if (g_off != g_buf.size()) { std::fprintf(stderr, "dump size mismatch: read %zu of %zu\n", g_off, g_buf.size()); return 3; }
    if (g_bad) { std::fprintf(stderr, "%d vectors differ from the reference\n", g_bad); return 1; }
    std::printf("golden math OK: %d cameras, %d products, %d shadow transforms bit-exact\n", K, N, KS);
    return 0;
}
