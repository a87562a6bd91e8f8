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

__device__ __forceinline__ LightDev const_light(const ShadeConst& sh, uint32_t i) {
    const float4* q = sh.lightsC[i];
    LightDev l;
    l.type = __float_as_uint(q[0].x);
    l.position = {q[1].x, q[1].y, q[1].z};
    l.direction = {q[2].x, q[2].y, q[2].z};
    l.ambient = q[3]; l.diffuse = q[4]; l.specular = q[5];
    l.innerCutoff = q[6].x; l.outerCutoff = q[6].y; l.shadowIndex = __float_as_uint(q[6].z);
    return l;
}

// PhongLighting (phong.glsl:39-81).  Colour is held to <= 1 LSB of the sRGB8 target / PSNR >= 50 dB on the f32
// buffer (BASELINE.json), not to bit-exactness, so this block — and only this block plus the sRGB encode — uses the
// SFU approximations (MUFU rsqrt / lg2 / ex2, relative error ~1e-6); everything that feeds a discontinuous
// decision (coverage, depth, the PCF comparisons) stays IEEE-exact.
__device__ __forceinline__ V3 fast_normalize3(V3 v) {
    const float inv = rsqrtf(dot3(v, v));
    return {v.x * inv, v.y * inv, v.z * inv};
}
__device__ __forceinline__ float fast_pow(float x, float y) {     // x >= 0
    if (y == 0.0f) return 1.0f;
    return exp2f(y * __log2f(x));                                 // x == 0 -> exp2(-inf) = 0
}
__device__ __forceinline__ float4 phong_lighting(const LightDev& light, float4 pDiffuse, float4 pSpecular, float shininess,
                                                 V3 normal, V3 fragPos, V3 viewPos, float shadow) {
    const V3 viewDir = fast_normalize3(sub3(viewPos, fragPos));
    V3 lightDir = light.type == LIGHT_TYPE_DIRECTIONAL ? neg3(light.direction) : sub3(light.position, fragPos);
    const float invDist = rsqrtf(dot3(lightDir, lightDir));
    lightDir = {lightDir.x * invDist, lightDir.y * invDist, lightDir.z * invDist};

    const float4 ambient = make_float4(light.ambient.x * pDiffuse.x, light.ambient.y * pDiffuse.y, light.ambient.z * pDiffuse.z, light.ambient.w * pDiffuse.w);

    const float k_d = fmaxf(dot3(lightDir, normal), 0.0f);
    float4 diffuse = make_float4((k_d * light.diffuse.x) * pDiffuse.x, (k_d * light.diffuse.y) * pDiffuse.y,
                                 (k_d * light.diffuse.z) * pDiffuse.z, (k_d * light.diffuse.w) * pDiffuse.w);

    const V3 h = fast_normalize3(add3(lightDir, viewDir));
    const float k_s = fast_pow(fmaxf(dot3(normal, h), 0.0f), shininess);
    float4 specular = make_float4((k_s * light.specular.x) * pSpecular.x, (k_s * light.specular.y) * pSpecular.y,
                                  (k_s * light.specular.z) * pSpecular.z, (k_s * light.specular.w) * pSpecular.w);

    float attenuation = 1.0f;
    if (light.type != LIGHT_TYPE_DIRECTIONAL) attenuation = invDist;

    if (light.type == LIGHT_TYPE_SPOT) {
        const float cos_theta = dot3(neg3(lightDir), fast_normalize3(light.direction));
        const float epsilon = light.innerCutoff - light.outerCutoff;
        const float intensity = fminf(fmaxf(__fdividef(cos_theta - light.outerCutoff, epsilon), 0.0f), 1.0f);
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

// AFFINE: the host found STR's last row to be exactly (0, 0, 0, 1) — every directional light's orthographic transform — so
// fragPosLightSpace.w is exactly 1.0f for any finite point and the three divisions of shadow.glsl:31 are exact identities.
template <bool AFFINE>
__device__ __forceinline__ float compute_shadow(const LightDev& light, V3 fragPos, V3 normal, const float* __restrict__ STR,
                                                const float* __restrict__ map, uint32_t S) {
    if (light.type != LIGHT_TYPE_DIRECTIONAL) return 0.0f;
    const float4 flp = mat4_mul_point(STR, fragPos.x, fragPos.y, fragPos.z);
    const V3 lightDir = normalize3(sub3(light.position, fragPos));
    const float currentDepth = AFFINE ? flp.z : flp.z / flp.w;
    if (currentDepth > 1.0f) return 0.0f;
    const float mu = 0.5f * (AFFINE ? flp.x : flp.x / flp.w) + 0.5f;
    const float mv = 0.5f * (AFFINE ? flp.y : flp.y / flp.w) + 0.5f;
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
    return shadow * (1.0f / 9.0f);   // shadow /= 9.0 (shadow.glsl:54); feeds colour only (tolerance, not bit-exactness)
}

// ---- bindless textures: isotropic trilinear, repeat, RGBA16 UNORM, optional sRGB decode after filtering.
// Filtering is continuous in every input, so (like PhongLighting) it is held to the colour tolerance, not to bit-exactness.
__device__ __forceinline__ int wrap_repeat(int i, int n) { int m = i % n; return m < 0 ? m + n : m; }
__device__ __forceinline__ float4 unorm16x4(uint2 t) {
    const float k = 1.0f / 65535.0f;
    return make_float4((float)(t.x & 0xFFFFu) * k, (float)(t.x >> 16) * k, (float)(t.y & 0xFFFFu) * k, (float)(t.y >> 16) * k);
}
__device__ __forceinline__ float4 lerp4(float4 a, float4 b, float f) {
    return make_float4(a.x + f * (b.x - a.x), a.y + f * (b.y - a.y), a.z + f * (b.z - a.z), a.w + f * (b.w - a.w));
}
__device__ __forceinline__ float4 sample_bilinear(const uint2* __restrict__ p, int w, int h, float u, float v) {
    const float x = u * (float)w - 0.5f, y = v * (float)h - 0.5f;
    const float fx0 = floorf(x), fy0 = floorf(y);
    const float fx = x - fx0, fy = y - fy0;
    const int ix = (fabsf(fx0) < 1.0e9f) ? (int)fx0 : 0, iy = (fabsf(fy0) < 1.0e9f) ? (int)fy0 : 0;
    const int i0 = wrap_repeat(ix, w), i1 = wrap_repeat(ix + 1, w), j0 = wrap_repeat(iy, h), j1 = wrap_repeat(iy + 1, h);
    const float4 t00 = unorm16x4(__ldg(p + (size_t)j0 * w + i0)), t10 = unorm16x4(__ldg(p + (size_t)j0 * w + i1));
    const float4 t01 = unorm16x4(__ldg(p + (size_t)j1 * w + i0)), t11 = unorm16x4(__ldg(p + (size_t)j1 * w + i1));
    return lerp4(lerp4(t00, t10, fx), lerp4(t01, t11, fx), fy);
}
__device__ __forceinline__ float srgb_eotf(float c) { return (c <= 0.04045f) ? c * (1.0f / 12.92f) : exp2f(2.4f * __log2f((c + 0.055f) * (1.0f / 1.055f))); }
__device__ __forceinline__ float4 sample_texture(const TexDev* __restrict__ table, uint32_t handle, float u, float v, float dudx, float dvdx, float dudy, float dvdy) {
    const uint4* q = reinterpret_cast<const uint4*>(table + handle);   // 96-B descriptor: {base, w, h} {mips, srgb, off[0..1]} ...
    const uint4 d0 = __ldg(q), d1 = __ldg(q + 1);
    const uint2* base = reinterpret_cast<const uint2*>(((unsigned long long)d0.y << 32) | d0.x);
    const uint32_t tw = d0.z, th = d0.w, mips = d1.x, srgb = d1.y;
    const uint32_t* offs = reinterpret_cast<const uint32_t*>(table + handle) + 6;   // off[] starts at byte 24
    const float W = (float)tw, H = (float)th;
    const float rx = sqrtf((dudx * W) * (dudx * W) + (dvdx * H) * (dvdx * H));
    const float ry = sqrtf((dudy * W) * (dudy * W) + (dvdy * H) * (dvdy * H));
    const float rho = fmaxf(rx, ry);
    const int last = (int)mips - 1;
    float lod = (rho > 0.0f) ? __log2f(rho) : 0.0f;
    if (!(lod > 0.0f)) lod = 0.0f;
    if (lod > (float)last) lod = (float)last;
    const int l0 = (int)floorf(lod), l1 = min(l0 + 1, last);
    const float f = lod - (float)l0;
    float4 a = sample_bilinear(base + __ldg(offs + l0), max((int)(tw >> l0), 1), max((int)(th >> l0), 1), u, v);
    if (f > 0.0f && l1 != l0)
        a = lerp4(a, sample_bilinear(base + __ldg(offs + l1), max((int)(tw >> l1), 1), max((int)(th >> l1), 1), u, v), f);
    if (srgb) a = make_float4(srgb_eotf(a.x), srgb_eotf(a.y), srgb_eotf(a.z), a.w);
    return a;
}

// sRGB8 encode of the R8G8B8A8_SRGB target (Swapchain.cpp:14): clamp, OETF, round(x * 255).  Colour is held to <= 1 LSB,
// so this uses the SFU lg2 / ex2 and explicit FMAs (see the tolerance note above phong_lighting).
__device__ __forceinline__ float ex2_approx(float x) { float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ uint32_t quant8(float x01) { return (uint32_t)__float2int_rd(__fmaf_rn(x01, 255.0f, 0.5f)); }   // x01 in [0, 1]
__device__ __forceinline__ float srgb_oetf(float x) {
    x = __saturatef(x);                                            // NaN -> 0
    const float hi = __fmaf_rn(1.055f, ex2_approx((1.0f / 2.4f) * __log2f(x)), -0.055f);   // x = 0: ex2(-inf) = 0, selected away below
    return (x <= 0.0031308f) ? 12.92f * x : hi;
}
__device__ __forceinline__ uint32_t pack_srgb8(float4 c) {
    return quant8(srgb_oetf(c.x)) | (quant8(srgb_oetf(c.y)) << 8) | (quant8(srgb_oetf(c.z)) << 16) | (quant8(__saturatef(c.w)) << 24);
}

// Appendix B.3 / B.10 (revision r1c): perspective-correct weights of a CLIPPED parent triangle at pixel centre (px, py), in
// homogeneous form from its clip-space vertices c_k = PV * (world_k, 1) (exactly what K0 computed):
// lambda = adj([c0 c1 c2]) * (xn, yn, 1) over the (x, y, w) rows.  Rare path, kept out of line.
static __device__ __noinline__ float4 homogeneous_weights(const PassConst& pc, V3 w0, V3 w1, V3 w2, float px, float py) {
    const float hw = 0.5f * (float)pc.W, hh = 0.5f * (float)pc.H;
    const float xn = ((px + 0.5f) - hw) / hw, yn = ((py + 0.5f) - hh) / hh;
    const float4 c0 = mat4_mul_point(pc.PV, w0.x, w0.y, w0.z), c1 = mat4_mul_point(pc.PV, w1.x, w1.y, w1.z), c2 = mat4_mul_point(pc.PV, w2.x, w2.y, w2.z);
    float4 l;
    l.x = ((c1.y * c2.w - c1.w * c2.y) * xn + (c1.w * c2.x - c1.x * c2.w) * yn) + (c1.x * c2.y - c1.y * c2.x);
    l.y = ((c2.y * c0.w - c2.w * c0.y) * xn + (c2.w * c0.x - c2.x * c0.w) * yn) + (c2.x * c0.y - c2.y * c0.x);
    l.z = ((c0.y * c1.w - c0.w * c1.y) * xn + (c0.w * c1.x - c0.x * c1.w) * yn) + (c0.x * c1.y - c0.y * c1.x);
    l.w = 0.0f;
    return l;
}

// Shade the sample at pixel (x, y) won by primitive `tri`: re-derive the barycentrics from the cached snapped
// vertices (exactly the integers triangle setup used), interpolate perspective-correct (Appendix B.10), then the
// builder-defined Phong main() (Appendix B.11) over phong.glsl + shadow.glsl.
template <bool TEXTURED, bool AFFINE>
__device__ __forceinline__ float4 shade_pixel(const PassConst& pc, const ShadeConst& sh, int x, int y, uint32_t tri) {
    const TriRef tr = tri_ref(pc, tri);
    const uint32_t materialRaw = tr.materialRaw;
    int32_t X[3], Y[3]; float invw[3]; V3 world[3], nrm[3]; uint32_t vflags = 0u;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const uint32_t slot = tr.s[k];
        const uint4 a = __ldg(reinterpret_cast<const uint4*>(pc.vtxA) + slot);
        const float4* bc = reinterpret_cast<const float4*>(pc.vtxBC + slot);
        const float4 b = __ldg(bc), c = __ldg(bc + 1);
        X[k] = (int32_t)a.x; Y[k] = (int32_t)a.y; vflags |= a.w;
        world[k] = {b.x, b.y, b.z}; invw[k] = b.w;
        nrm[k] = {c.x, c.y, c.z};
    }
    const bool clippedParent = (vflags & VF_NEEDS_CLIP) != 0u;   // the sample was won by a clipped child: no snapped parent vertices
    const long long A2 = mulw(X[1] - X[0], Y[2] - Y[0]) - mulw(X[2] - X[0], Y[1] - Y[0]);
    const int32_t Px = 256 * x + 128, Py = 256 * y + 128;   // |X|, |Y| <= 2^28 (guard band), so every difference below fits int32
    const float inv = 1.0f / (float)(-A2);      // Appendix B.8/B.10 revision r1b: one reciprocal per triangle ...
    float q[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const int a = (k + 1) % 3, b = (k + 2) % 3;
        const long long F = mulw(Y[b] - Y[a], Px - X[a]) + mulw(-(X[b] - X[a]), Py - Y[a]);
        const float bk = (float)F * inv;
        q[k] = bk * invw[k];
    }
    if (clippedParent) { const float4 l = homogeneous_weights(pc, world[0], world[1], world[2], (float)x, (float)y); q[0] = l.x; q[1] = l.y; q[2] = l.z; }
    const float rden = 1.0f / ((q[0] + q[1]) + q[2]);   // ... and one per fragment
#define REB_INTERP(a0, a1, a2) (((q[0] * (a0) + q[1] * (a1)) + q[2] * (a2)) * rden)
    const V3 fragPos = {REB_INTERP(world[0].x, world[1].x, world[2].x), REB_INTERP(world[0].y, world[1].y, world[2].y),
                        REB_INTERP(world[0].z, world[1].z, world[2].z)};
    const V3 normal = {REB_INTERP(nrm[0].x, nrm[1].x, nrm[2].x), REB_INTERP(nrm[0].y, nrm[1].y, nrm[2].y),
                       REB_INTERP(nrm[0].z, nrm[1].z, nrm[2].z)};
#undef REB_INTERP

    // bindless material fetch: materials[materialIndex] (pbr.glsl:379 pattern; MaterialDefines.h:88-96)
    uint32_t mi = materialRaw >> 4;
    if (mi >= sh.nMaterials) mi = sh.nMaterials - 1u;
    const float4* mq = reinterpret_cast<const float4*>(sh.materials + (size_t)mi * sh.materialStride);
    float4 pDiffuse = __ldg(mq), pSpecular = __ldg(mq + 1);
    const float4 mExtra = __ldg(mq + 2);   // shininess, diffuseTex, specularTex, pad
    const float shininess = mExtra.x;
    const uint32_t texD = __float_as_uint(mExtra.y), texS = __float_as_uint(mExtra.z);
    const uint32_t texLimit = min(sh.maxDescriptors, sh.nTextures);   // live iff handle < MAX_DESCRIPTOR_COUNT (bindless.glsl:14) and registered
    if (TEXTURED && (texD < texLimit || texS < texLimit) && pc.vtxUV != nullptr) {
        // uv here and at the +x / +y neighbour pixel centres of the same primitive (implicit derivatives, as the oracle)
        float2 uvv[3];
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const uint32_t slot = tr.s[k];
            uvv[k] = __ldg(pc.vtxUV + slot);
        }
        float tu[3], tv[3];
#pragma unroll
        for (int n = 0; n < 3; ++n) {
            const int32_t Qx = Px + (n == 1 ? 256 : 0), Qy = Py + (n == 2 ? 256 : 0);
            float qq[3];
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                const int a = (k + 1) % 3, b = (k + 2) % 3;
                const long long F = mulw(Y[b] - Y[a], Qx - X[a]) + mulw(-(X[b] - X[a]), Qy - Y[a]);
                qq[k] = ((float)F * inv) * invw[k];
            }
            if (clippedParent) { const float4 l = homogeneous_weights(pc, world[0], world[1], world[2], (float)(x + (n == 1)), (float)(y + (n == 2))); qq[0] = l.x; qq[1] = l.y; qq[2] = l.z; }
            const float rdd = 1.0f / ((qq[0] + qq[1]) + qq[2]);
            tu[n] = ((qq[0] * uvv[0].x + qq[1] * uvv[1].x) + qq[2] * uvv[2].x) * rdd;
            tv[n] = ((qq[0] * uvv[0].y + qq[1] * uvv[1].y) + qq[2] * uvv[2].y) * rdd;
        }
        const float dudx = tu[1] - tu[0], dvdx = tv[1] - tv[0], dudy = tu[2] - tu[0], dvdy = tv[2] - tv[0];
        if (texD < texLimit) {
            const float4 c = sample_texture(sh.textures, texD, tu[0], tv[0], dudx, dvdx, dudy, dvdy);
            pDiffuse = make_float4(pDiffuse.x * c.x, pDiffuse.y * c.y, pDiffuse.z * c.z, pDiffuse.w * c.w);
        }
        if (texS < texLimit) {
            const float4 c = sample_texture(sh.textures, texS, tu[0], tv[0], dudx, dvdx, dudy, dvdy);
            pSpecular = make_float4(pSpecular.x * c.x, pSpecular.y * c.y, pSpecular.z * c.z, pSpecular.w * c.w);
        }
    }

    const V3 viewPos = {sh.viewPos[0], sh.viewPos[1], sh.viewPos[2]};
    float4 acc = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    for (uint32_t l = 0; l < sh.nLights; ++l) {
        // the first SHADE_CONST_LIGHTS records ride in the kernel parameters (constant bank: no loads, no registers held)
        const LightDev L = l < (uint32_t)SHADE_CONST_LIGHTS ? const_light(sh, l) : load_light(sh.lights + (size_t)l * sh.lightStride);
        float shadow = 0.0f;
        if (sh.shadowMap != nullptr && L.type == LIGHT_TYPE_DIRECTIONAL && L.shadowIndex == 0u)
            shadow = compute_shadow<AFFINE>(L, fragPos, normal, sh.STR, sh.shadowMap, sh.shadowSize);
        const float4 cl = phong_lighting(L, pDiffuse, pSpecular, shininess, normal, fragPos, viewPos, shadow);
        acc = make_float4(acc.x + cl.x, acc.y + cl.y, acc.z + cl.z, acc.w + cl.w);
    }
    return make_float4(acc.x, acc.y, acc.z, pDiffuse.w);
}

}  // namespace reb
