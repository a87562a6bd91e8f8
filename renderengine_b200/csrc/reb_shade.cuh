// reb_shade.cuh — per-pixel attribute reconstruction + Phong + PCF, device side.
//
// PhongLighting  : Assets/Shaders/phong.glsl:39-81 (statement order preserved)
// ComputeShadow  : Assets/Shaders/shadow.glsl:24-56 (9 nearest/clamp taps, ShadowSystem.cpp:18-35)
// main()         : builder-defined, modelled on pbr.glsl:388-408
// varyings       : primitive.vert:49-57, perspective-correct (Appendix B.10)
#pragma once
#include "reb_device.cuh"
#include "reb_kernels.h"

namespace reb {

struct LightDev {
    uint32_t type; V3 position; V3 direction; float4 ambient, diffuse, specular;
    float innerCutoff, outerCutoff; uint32_t shadowIndex;
};

__device__ __forceinline__ LightDev load_light(const uint8_t* __restrict__ base) {
    const float4* q = reinterpret_cast<const float4*>(base);   // std430: 7 x 16 B
    const float4 a = __ldg(q), b = __ldg(q + 1), c = __ldg(q + 2), g = __ldg(q + 6);
    LightDev l;
    l.type = __float_as_uint(a.x);
    l.position = {b.x, b.y, b.z};
    l.direction = {c.x, c.y, c.z};
    l.ambient = __ldg(q + 3); l.diffuse = __ldg(q + 4); l.specular = __ldg(q + 5);
    l.innerCutoff = g.x; l.outerCutoff = g.y; l.shadowIndex = __float_as_uint(g.z);
    return l;
}

__device__ __forceinline__ float4 phong_lighting(const LightDev& light, float4 pDiffuse, float4 pSpecular, float shininess,
                                                 V3 normal, V3 fragPos, V3 viewPos, float shadow) {
    const V3 viewDir = normalize3(sub3(viewPos, fragPos));
    V3 lightDir = light.type == LIGHT_TYPE_DIRECTIONAL ? neg3(light.direction) : sub3(light.position, fragPos);
    const float lightDistance = sqrtf(dot3(lightDir, lightDir));
    lightDir = {lightDir.x / lightDistance, lightDir.y / lightDistance, lightDir.z / lightDistance};

    const float4 ambient = make_float4(light.ambient.x * pDiffuse.x, light.ambient.y * pDiffuse.y, light.ambient.z * pDiffuse.z, light.ambient.w * pDiffuse.w);

    const float k_d = glm_max(dot3(lightDir, normal), 0.0f);
    float4 diffuse = make_float4((k_d * light.diffuse.x) * pDiffuse.x, (k_d * light.diffuse.y) * pDiffuse.y,
                                 (k_d * light.diffuse.z) * pDiffuse.z, (k_d * light.diffuse.w) * pDiffuse.w);

    const V3 h = normalize3(add3(lightDir, viewDir));
    const float k_s = powf(glm_max(dot3(normal, h), 0.0f), shininess);
    float4 specular = make_float4((k_s * light.specular.x) * pSpecular.x, (k_s * light.specular.y) * pSpecular.y,
                                  (k_s * light.specular.z) * pSpecular.z, (k_s * light.specular.w) * pSpecular.w);

    float attenuation = 1.0f;
    if (light.type != LIGHT_TYPE_DIRECTIONAL) attenuation = 1.0f / lightDistance;

    if (light.type == LIGHT_TYPE_SPOT) {
        const float cos_theta = dot3(neg3(lightDir), normalize3(light.direction));
        const float epsilon = light.innerCutoff - light.outerCutoff;
        const float intensity = glm_min(glm_max((cos_theta - light.outerCutoff) / epsilon, 0.0f), 1.0f);
        diffuse = make_float4(diffuse.x * intensity, diffuse.y * intensity, diffuse.z * intensity, diffuse.w * intensity);
        specular = make_float4(specular.x * intensity, specular.y * intensity, specular.z * intensity, specular.w * intensity);
    }
    const float k = 1.0f - shadow;
    float4 r;
    r.x = (ambient.x + k * (diffuse.x + specular.x)) * attenuation;
    r.y = (ambient.y + k * (diffuse.y + specular.y)) * attenuation;
    r.z = (ambient.z + k * (diffuse.z + specular.z)) * attenuation;
    r.w = (ambient.w + k * (diffuse.w + specular.w)) * attenuation;
    return r;
}

__device__ __forceinline__ float shadow_fetch(const float* __restrict__ map, uint32_t S, float u, float v) {
    const float fs = (float)S;
    const float fu = floorf(u * fs), fv = floorf(v * fs);
    const float hi = (float)(S - 1);
    const int i = (fu >= 0.0f) ? ((fu <= hi) ? (int)fu : (int)(S - 1)) : 0;
    const int j = (fv >= 0.0f) ? ((fv <= hi) ? (int)fv : (int)(S - 1)) : 0;
    return __ldg(map + (size_t)j * S + i);
}

__device__ __forceinline__ float compute_shadow(const LightDev& light, V3 fragPos, V3 normal, const float* __restrict__ STR,
                                                const float* __restrict__ map, uint32_t S) {
    if (light.type != LIGHT_TYPE_DIRECTIONAL) return 0.0f;
    const float4 flp = mat4_mul_point(STR, fragPos.x, fragPos.y, fragPos.z);
    const V3 lightDir = normalize3(sub3(light.position, fragPos));
    const float currentDepth = flp.z / flp.w;
    if (currentDepth > 1.0f) return 0.0f;
    const float mu = 0.5f * (flp.x / flp.w) + 0.5f;
    const float mv = 0.5f * (flp.y / flp.w) + 0.5f;
    const float bias = glm_max(0.05f * (1.0f - dot3(normal, lightDir)), 0.005f);
    const float texel = 1.0f / (float)S;
    const float ref = currentDepth - bias;
    float shadow = 0.0f;
#pragma unroll
    for (int x = -1; x <= 1; ++x)
#pragma unroll
        for (int y = -1; y <= 1; ++y) {
            const float depth = shadow_fetch(map, S, mu + (float)x * texel, mv + (float)y * texel);
            shadow += (ref > depth) ? 1.0f : 0.0f;
        }
    shadow /= 9.0f;
    return shadow;
}

__device__ __forceinline__ uint32_t quant8(float x) {
    if (!(x > 0.0f)) return 0u;
    if (x >= 1.0f) return 255u;
    return (uint32_t)(int)floorf(x * 255.0f + 0.5f);
}
__device__ __forceinline__ float srgb_oetf(float x) {
    if (!(x > 0.0f)) return 0.0f;
    if (x >= 1.0f) return 1.0f;
    return (x <= 0.0031308f) ? 12.92f * x : 1.055f * powf(x, 1.0f / 2.4f) - 0.055f;
}
__device__ __forceinline__ uint32_t pack_srgb8(float4 c) {
    return quant8(srgb_oetf(c.x)) | (quant8(srgb_oetf(c.y)) << 8) | (quant8(srgb_oetf(c.z)) << 16) | (quant8(c.w) << 24);
}

// Shade the sample at pixel (x, y) won by primitive `tri`.
__device__ __forceinline__ float4 shade_pixel(const PassConst& pc, const ShadeConst& sh, int x, int y, uint32_t tri) {
    const uint32_t d = find_draw(pc.draws, pc.nDraws, tri);
    const uint4 dr = __ldg(reinterpret_cast<const uint4*>(pc.draws) + d);
    const uint32_t* ix = pc.indices + dr.x + 3u * (tri - dr.y);
    uint32_t vi[3] = {__ldg(ix), __ldg(ix + 1), __ldg(ix + 2)};
    const float4* Tq = reinterpret_cast<const float4*>(pc.transforms) + 4u * dr.w;
    float T[16];
#pragma unroll
    for (int k = 0; k < 4; ++k) { float4 q = __ldg(Tq + k); T[4 * k] = q.x; T[4 * k + 1] = q.y; T[4 * k + 2] = q.z; T[4 * k + 3] = q.w; }
    const float4* Nq = reinterpret_cast<const float4*>(sh.nmats) + 3u * dr.w;
    const float4 n0 = __ldg(Nq), n1 = __ldg(Nq + 1), n2 = __ldg(Nq + 2);
    const float NM[9] = {n0.x, n0.y, n0.z, n0.w, n1.x, n1.y, n1.z, n1.w, n2.x};

    float4 c[3]; V3 world[3], nrm[3]; float tu[3], tv[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const float4 p = __ldg(pc.pos4 + vi[k]);
        const float4 n = __ldg(sh.nrm4 + vi[k]);
        c[k] = vertex_stage<true>(pc.PV, T, p.x, p.y, p.z, &world[k]);
        V3 tn;   // glm mat3 * vec3: (m0*x + m1*y) + m2*z
        tn.x = (NM[0] * n.x + NM[3] * n.y) + NM[6] * n.z;
        tn.y = (NM[1] * n.x + NM[4] * n.y) + NM[7] * n.z;
        tn.z = (NM[2] * n.x + NM[5] * n.y) + NM[8] * n.z;
        nrm[k] = normalize3(tn);
        tu[k] = p.w; tv[k] = n.w;
    }
    Snapped s;
    setup_triangle(c, pc.W, pc.H, s);
    const long long Px = 256ll * x + 128, Py = 256ll * y + 128;
    const float area = (float)(-s.A2);
    float q[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const int a = (k + 1) % 3, b = (k + 2) % 3;
        const long long A = (long long)s.Y[b] - s.Y[a];
        const long long B = -((long long)s.X[b] - s.X[a]);
        const long long C = -(A * s.X[a]) - B * s.Y[a];
        const long long F = A * Px + B * Py + C;
        const float bk = (float)F / area;
        q[k] = bk * (1.0f / s.w[k]);
    }
    const float den = (q[0] + q[1]) + q[2];
#define REB_INTERP(a0, a1, a2) (((q[0] * (a0) + q[1] * (a1)) + q[2] * (a2)) / den)
    const V3 fragPos = {REB_INTERP(world[0].x, world[1].x, world[2].x), REB_INTERP(world[0].y, world[1].y, world[2].y),
                        REB_INTERP(world[0].z, world[1].z, world[2].z)};
    const V3 normal = {REB_INTERP(nrm[0].x, nrm[1].x, nrm[2].x), REB_INTERP(nrm[0].y, nrm[1].y, nrm[2].y),
                       REB_INTERP(nrm[0].z, nrm[1].z, nrm[2].z)};
    const float u = REB_INTERP(tu[0], tu[1], tu[2]), v = REB_INTERP(tv[0], tv[1], tv[2]);
#undef REB_INTERP
    (void)u; (void)v;

    // bindless material fetch: materials[materialIndex] (pbr.glsl:379 pattern; MaterialDefines.h:88-96)
    uint32_t mi = dr.z >> 4;
    if (mi >= sh.nMaterials) mi = sh.nMaterials - 1u;
    const float4* mq = reinterpret_cast<const float4*>(sh.materials + (size_t)mi * sh.materialStride);
    const float4 pDiffuse = __ldg(mq), pSpecular = __ldg(mq + 1);
    const float4 mExtra = __ldg(mq + 2);   // shininess, diffuseTex, specularTex, pad
    const float shininess = mExtra.x;

    const V3 viewPos = {sh.viewPos[0], sh.viewPos[1], sh.viewPos[2]};
    float4 acc = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    for (uint32_t l = 0; l < sh.nLights; ++l) {
        const LightDev L = load_light(sh.lights + (size_t)l * sh.lightStride);
        float shadow = 0.0f;
        if (sh.shadowMap != nullptr && L.type == LIGHT_TYPE_DIRECTIONAL && L.shadowIndex == 0u)
            shadow = compute_shadow(L, fragPos, normal, sh.STR, sh.shadowMap, sh.shadowSize);
        const float4 cl = phong_lighting(L, pDiffuse, pSpecular, shininess, normal, fragPos, viewPos, shadow);
        acc = make_float4(acc.x + cl.x, acc.y + cl.y, acc.z + cl.z, acc.w + cl.w);
    }
    return make_float4(acc.x, acc.y, acc.z, pDiffuse.w);
}

}  // namespace reb
