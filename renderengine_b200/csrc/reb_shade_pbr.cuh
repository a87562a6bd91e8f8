// reb_shade_pbr.cuh — the reference's LIVE fragment path on the device: surface_pbr.frag:38-50 -> BRDF_Lighting
// (pbr.glsl:368-451): GGX / Smith-correlated / Schlick BRDF (pbr.glsl:237-279), punctual lights with square falloff and
// spot cone (pbr.glsl:303-352), bindless material record (MaterialSystem.h:41-51 <-> pbr.glsl:78-94) with base-colour,
// emissive, occlusion-roughness-metallic and tangent-space normal textures (pbr.glsl:100-157), 3x3 PCF shadow
// (shadow.glsl:24-56), debug views (pbr.glsl:409-439), exposure + ambient (pbr.glsl:441-450).
//
// Arithmetic: the same individually rounded IEEE operations, in the same order, as the oracle's restatement
// (oracle/oracle.cpp pbr_lighting / pbr_remap, pinned bit-for-bit against the reference shader) — these translation
// units are compiled with --fmad=false — except pow(1 - HoL, 5) (a multiply chain) and the texture filter, which are
// continuous and held to the colour tolerance.  Everything that feeds the nine PCF comparisons is exact.
#pragma once
#include "reb_shade.cuh"

namespace reb {

struct PbrLightDev { V3 color, position, direction; float intensity, falloffRadius, cosInner, cosOuter; uint32_t shadowIndex, type; };

__device__ __forceinline__ PbrLightDev pbr_light_from(const float4 a, const float4 b, const float4 c, const float4 d, const float4 e) {
    PbrLightDev l;   // std430, 80 B: color@0 position@16 direction@32 intensity@44 falloffRadius@48 cosInner@52 cosOuter@56 shadowIndex@60 type@64
    l.color = {a.x, a.y, a.z}; l.position = {b.x, b.y, b.z}; l.direction = {c.x, c.y, c.z}; l.intensity = c.w;
    l.falloffRadius = d.x; l.cosInner = d.y; l.cosOuter = d.z; l.shadowIndex = __float_as_uint(d.w); l.type = __float_as_uint(e.x);
    return l;
}
__device__ __forceinline__ PbrLightDev pbr_load_light(const ShadeConst& sh, uint32_t i) {
    if (i < (uint32_t)SHADE_CONST_LIGHTS) { const float4* q = sh.lightsC[i]; return pbr_light_from(q[0], q[1], q[2], q[3], q[4]); }
    const float4* q = reinterpret_cast<const float4*>(sh.lights + (size_t)i * sh.lightStride);
    return pbr_light_from(__ldg(q), __ldg(q + 1), __ldg(q + 2), __ldg(q + 3), __ldg(q + 4));
}

__device__ __forceinline__ float glm_clampf(float x, float lo, float hi) { return glm_min(glm_max(x, lo), hi); }
__device__ __forceinline__ float glm_absf(float x) { return x >= 0.0f ? x : -x; }
__device__ __forceinline__ V3 cross3(V3 x, V3 y) { return {x.y * y.z - y.y * x.z, x.z * y.x - y.z * x.x, x.x * y.y - y.x * x.y}; }
__device__ __forceinline__ V3 scale3(V3 v, float s) { return {v.x * s, v.y * s, v.z * s}; }

struct PbrRemapped { V3 normal; float4 baseColor; V3 diffuseColor; float4 emissive; V3 f0; float metallic, roughness, occlusion; };

// BRDF_Lighting after the texture fetches (pbr.glsl:368-451)
template <bool AFFINE>
__device__ __forceinline__ float4 pbr_lighting(const ShadeConst& sh, const PbrRemapped& m, V3 fragPos, V3 viewPos) {
    const float PI = 3.14159265359f;
    const V3 n = m.normal;
    const V3 v = normalize3(sub3(viewPos, fragPos));
    V3 Fd = {0.f, 0.f, 0.f}, F = {0.f, 0.f, 0.f}, Fr = {0.f, 0.f, 0.f};
    float D = 0.0f, G = 0.0f, NoV = 0.0f, NoL = 0.0f;
    V3 l0 = {0.0f, 0.0f, 0.0f};
    for (uint32_t li = 0; li < sh.nLights; ++li) {
        const PbrLightDev light = pbr_load_light(sh, li);
        V3 l = normalize3(sub3(light.position, fragPos));
        if (light.type == LIGHT_TYPE_DIRECTIONAL) l = normalize3(neg3(light.direction));
        const V3 h = normalize3(add3(v, l));
        NoV = glm_absf(dot3(n, v)) + 1e-5f;
        const float NoH = glm_clampf(dot3(n, h), 0.0f, 1.0f);
        NoL = glm_clampf(dot3(n, l), 0.0f, 1.0f);
        const float HoL = glm_clampf(dot3(h, l), 0.0f, 1.0f);
        const float a = m.roughness, a2 = a * a;
        { const float f = (NoH * a2 - NoH) * NoH + 1.0f; D = a2 / (PI * f * f); }                       // D_GGX
        { const float x = 1.0f - HoL, x2 = x * x, p5 = (x2 * x2) * x;                                    // F_Schlick
          F = {m.f0.x + (1.0f - m.f0.x) * p5, m.f0.y + (1.0f - m.f0.y) * p5, m.f0.z + (1.0f - m.f0.z) * p5}; }
        { const float GGXL = NoV * sqrtf((-NoL * a2 + NoL) * NoL + a2);                                  // G_SmithGGXCorrelated
          const float GGXV = NoL * sqrtf((-NoV * a2 + NoV) * NoV + a2);
          G = 0.5f / (GGXV + GGXL); }
        const float DG = D * G;
        Fr = {DG * F.x, DG * F.y, DG * F.z};
        const float lambert = 1.0f / PI;
        Fd = {(((1.0f - F.x) * m.occlusion) * m.diffuseColor.x) * lambert, (((1.0f - F.y) * m.occlusion) * m.diffuseColor.y) * lambert,
              (((1.0f - F.z) * m.occlusion) * m.diffuseColor.z) * lambert};
        const V3 FrFd = add3(Fr, Fd);
        float attenuation = 1.0f;                                                                        // EvaluatePunctualLight
        if (light.type == 2u || light.type == LIGHT_TYPE_SPOT) {
            const float inverseFalloffRadius = 1.0f / glm_max(light.falloffRadius, 0.083f);
            const float distanceSquare = dot3(l, l);   // of the NORMALISED direction, as in the reference (pbr.glsl:340)
            const float factor = distanceSquare * inverseFalloffRadius * inverseFalloffRadius;
            const float smoothFactor = glm_max(1.0f - factor * factor, 0.0f);
            attenuation = (smoothFactor * smoothFactor) / glm_max(distanceSquare, 1e-4f);
            if (light.type == LIGHT_TYPE_SPOT) {
                const float spotScale = 1.0f / glm_max(light.cosInner - light.cosOuter, 1e-4f);
                const float spotOffset = -light.cosOuter * spotScale;
                const float cd = dot3(normalize3(neg3(light.direction)), l);
                const float att = glm_clampf(cd * spotScale + spotOffset, 0.0f, 1.0f);
                attenuation = attenuation * (att * att);
            }
        }
        const V3 lightResult = {(((FrFd.x * light.intensity) * attenuation) * NoL) * light.color.x, (((FrFd.y * light.intensity) * attenuation) * NoL) * light.color.y,
                                (((FrFd.z * light.intensity) * attenuation) * NoL) * light.color.z};
        float shadow = 0.0f;
        if (light.type == LIGHT_TYPE_DIRECTIONAL && sh.shadowMap != nullptr && light.shadowIndex == 0u) {
            LightDev sl; sl.type = light.type; sl.position = light.position;   // ComputeShadow reads only these (shadow.glsl:24-56)
            shadow = compute_shadow<AFFINE>(sl, fragPos, n, sh.STR, sh.shadowMap, sh.shadowSize);
        }
        const float k = 1.0f - shadow;
        l0 = {l0.x + k * lightResult.x, l0.y + k * lightResult.y, l0.z + k * lightResult.z};
    }
    switch (sh.debugInput) {
        case 1: return m.baseColor;
        case 2: return make_float4(m.diffuseColor.x, m.diffuseColor.y, m.diffuseColor.z, 1.0f);
        case 3: return make_float4(m.normal.x, m.normal.y, m.normal.z, 1.0f);
        case 4: return make_float4(m.occlusion, m.occlusion, m.occlusion, 1.0f);
        case 5: return make_float4(m.emissive.x * m.emissive.w, m.emissive.y * m.emissive.w, m.emissive.z * m.emissive.w, 1.0f);
        case 6: return make_float4(m.metallic, m.metallic, m.metallic, 1.0f);
        case 7: return make_float4(m.roughness, m.roughness, m.roughness, 1.0f);
        default: break;
    }
    switch (sh.debugEquation) {
        case 1: return make_float4(Fd.x, Fd.y, Fd.z, 1.0f);
        case 2: return make_float4(F.x, F.y, F.z, 1.0f);
        case 3: { const float g = (G * 4.0f) * (glm_absf(NoL) * glm_absf(NoV)); return make_float4(g, g, g, 1.0f); }
        case 4: return make_float4(D, D, D, 1.0f);
        case 5: return make_float4(Fr.x, Fr.y, Fr.z, 1.0f);
        default: break;
    }
    V3 color = {l0.x + m.emissive.x * m.emissive.w, l0.y + m.emissive.y * m.emissive.w, l0.z + m.emissive.z * m.emissive.w};
    color = scale3(color, sh.exposure);
    const V3 ambient = {(0.03f * m.baseColor.x) * m.occlusion, (0.03f * m.baseColor.y) * m.occlusion, (0.03f * m.baseColor.z) * m.occlusion};
    return make_float4(color.x + ambient.x, color.y + ambient.y, color.z + ambient.z, m.baseColor.w);
}

// Shade the sample at pixel (x, y) won by primitive `tri` with the PBR model (same attribute reconstruction as
// shade_pixel: Appendix B.10 from the cached snapped vertices).
template <bool TEXTURED, bool AFFINE>
__device__ __forceinline__ float4 shade_pixel_pbr(const PassConst& pc, const ShadeConst& sh, int x, int y, uint32_t tri) {
    const TriRef tr = tri_ref(pc, tri);
    const uint32_t materialRaw = tr.materialRaw;
    int32_t X[3], Y[3]; float invw[3]; V3 world[3], nrm[3]; uint32_t slots[3]; uint32_t vflags = 0u;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const uint32_t slot = tr.s[k];
        slots[k] = slot;
        const uint4 a = __ldg(reinterpret_cast<const uint4*>(pc.vtxA) + slot);
        const float4* bc = reinterpret_cast<const float4*>(pc.vtxBC + slot);
        const float4 b = __ldg(bc), c = __ldg(bc + 1);
        X[k] = (int32_t)a.x; Y[k] = (int32_t)a.y; vflags |= a.w;
        world[k] = {b.x, b.y, b.z}; invw[k] = b.w;
        nrm[k] = {c.x, c.y, c.z};
    }
    const bool clippedParent = (vflags & VF_NEEDS_CLIP) != 0u;   // Appendix B.3: won by a clipped child, weights in homogeneous form
    const long long A2 = mulw(X[1] - X[0], Y[2] - Y[0]) - mulw(X[2] - X[0], Y[1] - Y[0]);
    const int32_t Px = 256 * x + 128, Py = 256 * y + 128;
    const float inv = 1.0f / (float)(-A2);
    // perspective-correct weights at (Qx, Qy): q_k / sum q
    auto weights = [&](int32_t Qx, int32_t Qy, float (&q)[3], float& rden) {
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const int a = (k + 1) % 3, b = (k + 2) % 3;
            const long long F = mulw(Y[b] - Y[a], Qx - X[a]) + mulw(-(X[b] - X[a]), Qy - Y[a]);
            q[k] = ((float)F * inv) * invw[k];
        }
        if (clippedParent) {
            const float4 l = homogeneous_weights(pc, world[0], world[1], world[2], (float)((Qx - 128) >> 8), (float)((Qy - 128) >> 8));
            q[0] = l.x; q[1] = l.y; q[2] = l.z;
        }
        rden = 1.0f / ((q[0] + q[1]) + q[2]);
    };
    float q[3], rden;
    weights(Px, Py, q, rden);
#define REB_INTERP(a0, a1, a2) (((q[0] * (a0) + q[1] * (a1)) + q[2] * (a2)) * rden)
    const V3 fragPos = {REB_INTERP(world[0].x, world[1].x, world[2].x), REB_INTERP(world[0].y, world[1].y, world[2].y),
                        REB_INTERP(world[0].z, world[1].z, world[2].z)};
    const V3 normal = {REB_INTERP(nrm[0].x, nrm[1].x, nrm[2].x), REB_INTERP(nrm[0].y, nrm[1].y, nrm[2].y),
                       REB_INTERP(nrm[0].z, nrm[1].z, nrm[2].z)};

    // bindless material fetch: materials[materialIndex] (pbr.glsl:379), 80-B MaterialProperties
    uint32_t mi = materialRaw >> 4;
    if (mi >= sh.nMaterials) mi = sh.nMaterials - 1u;
    const float4* mq = reinterpret_cast<const float4*>(sh.materials + (size_t)mi * sh.materialStride);
    const float4 mBase = __ldg(mq), mEmis = __ldg(mq + 1), mScal = __ldg(mq + 2) /* f0, metallic, perceptualRoughness, ambientOcclusion */;
    const uint4 mTex = __ldg(reinterpret_cast<const uint4*>(mq + 3));   // baseColor, emissive, ORM, normals texture handles (AO handle unused by the shader)

    PbrRemapped m;
    m.normal = normal;
    m.baseColor = mBase; m.emissive = mEmis;
    V3 orm = {mScal.w, mScal.z, mScal.y};   // occlusion, perceptual roughness, metallic
    const uint32_t texLimit = min(sh.maxDescriptors, sh.nTextures);   // live iff handle < MAX_DESCRIPTOR_COUNT (bindless.glsl:14) and registered
    if (TEXTURED && pc.vtxUV != nullptr && (mTex.x < texLimit || mTex.y < texLimit || mTex.z < texLimit || mTex.w < texLimit)) {
        // uv and world position here and at the +x / +y neighbour pixel centres of the same primitive (implicit derivatives, as the oracle)
        float2 uvv[3];
#pragma unroll
        for (int k = 0; k < 3; ++k) uvv[k] = __ldg(pc.vtxUV + slots[k]);
        float tu[3], tv[3]; V3 tp[3];
        tu[0] = REB_INTERP(uvv[0].x, uvv[1].x, uvv[2].x); tv[0] = REB_INTERP(uvv[0].y, uvv[1].y, uvv[2].y); tp[0] = fragPos;
#pragma unroll
        for (int nb = 1; nb < 3; ++nb) {
            float qq[3], rdd;
            weights(Px + (nb == 1 ? 256 : 0), Py + (nb == 2 ? 256 : 0), qq, rdd);
#define REB_INTERP2(a0, a1, a2) (((qq[0] * (a0) + qq[1] * (a1)) + qq[2] * (a2)) * rdd)
            tu[nb] = REB_INTERP2(uvv[0].x, uvv[1].x, uvv[2].x); tv[nb] = REB_INTERP2(uvv[0].y, uvv[1].y, uvv[2].y);
            tp[nb] = {REB_INTERP2(world[0].x, world[1].x, world[2].x), REB_INTERP2(world[0].y, world[1].y, world[2].y), REB_INTERP2(world[0].z, world[1].z, world[2].z)};
#undef REB_INTERP2
        }
        const float dudx = tu[1] - tu[0], dvdx = tv[1] - tv[0], dudy = tu[2] - tu[0], dvdy = tv[2] - tv[0];
        if (mTex.w < texLimit) {   // GetNormal, pbr.glsl:125-148
            const float4 t = sample_texture(sh.textures, mTex.w, tu[0], tv[0], dudx, dvdx, dudy, dvdy);
            const V3 tn = {t.x * 2.0f - 1.0f, t.y * 2.0f - 1.0f, t.z * 2.0f - 1.0f};
            const V3 q1 = sub3(tp[1], tp[0]), q2 = sub3(tp[2], tp[0]);
            const V3 N = normalize3(normal);
            const V3 T = normalize3(sub3(scale3(q1, dvdy), scale3(q2, dvdx)));
            const V3 B = neg3(normalize3(cross3(N, T)));
            m.normal = normalize3({(T.x * tn.x + B.x * tn.y) + N.x * tn.z, (T.y * tn.x + B.y * tn.y) + N.y * tn.z, (T.z * tn.x + B.z * tn.y) + N.z * tn.z});
        }
        if (mTex.x < texLimit) {
            const float4 c = sample_texture(sh.textures, mTex.x, tu[0], tv[0], dudx, dvdx, dudy, dvdy);
            m.baseColor = make_float4(mBase.x * c.x, mBase.y * c.y, mBase.z * c.z, mBase.w * c.w);
        }
        if (mTex.y < texLimit) {
            const float4 c = sample_texture(sh.textures, mTex.y, tu[0], tv[0], dudx, dvdx, dudy, dvdy);
            m.emissive = make_float4(mEmis.x * c.x, mEmis.y * c.y, mEmis.z * c.z, mEmis.w * c.w);
        }
        if (mTex.z < texLimit) {
            const float4 c = sample_texture(sh.textures, mTex.z, tu[0], tv[0], dudx, dvdx, dudy, dvdy);
            orm = {orm.x * c.x, orm.y * c.y, orm.z * c.z};
        }
    }
#undef REB_INTERP
    // RemapMaterial, pbr.glsl:186-206
    m.occlusion = orm.x;
    { const float rr = glm_clampf(orm.y, 0.089f, 1.0f); m.roughness = rr * rr; }
    m.metallic = orm.z;
    m.diffuseColor = {(1.0f - m.metallic) * m.baseColor.x, (1.0f - m.metallic) * m.baseColor.y, (1.0f - m.metallic) * m.baseColor.z};
    const float f0s = mScal.x * (1.0f - m.metallic);
    m.f0 = {f0s + m.baseColor.x * m.metallic, f0s + m.baseColor.y * m.metallic, f0s + m.baseColor.z * m.metallic};
    const V3 viewPos = {sh.viewPos[0], sh.viewPos[1], sh.viewPos[2]};
    return pbr_lighting<AFFINE>(sh, m, fragPos, viewPos);
}

}  // namespace reb
