// ORACLE — TEST INFRASTRUCTURE, NOT PRODUCT CODE.
// CPU restatement of the reference's OptiX device programs and launch schedule, literal on purpose:
//   __raygen__renderFrame        sv4/deviceProgram.cu:389-588   -> oracleRaygen
//   __closesthit__radiance       sv4/deviceProgram.cu:590-703   -> oracleClosestHitProgram
//   __miss__radiance             sv4/deviceProgram.cu:251-280   -> oracleMissProgram
//   SampleLights / SampleShadow  sv4/deviceProgram.cu:301-384
//   SampleRenderer::render()     sv4/SimplePathtracer.cpp:77-217 -> oracleRenderFrame
//   orig variant display/AOVs    HelloPathtracing_original/deviceProgram.cu:452-474
// Shading math comes from the selected backend (reference headers or shading_port.h).
#pragma once
#include "scene_core.h"
#include "../include/b200pt.h"

static const int RAY_STATE_FLAGS_DONE = 1 << 0;
static const int RAY_STATE_FLAGS_SECONDARY_RAY = 1 << 1;
static const int RAY_STATE_FLAGS_RR_KILLED = 1 << 8;     // opt-in Russian roulette (not in the reference; TODO at sv4/deviceProgram.cu:498-500)

struct RadiancePRD {                    // sv4/deviceProgram.cu:59-87
    float3 radiance, alpha, origin, direction, normal, albedo;
    float bsdfPdf;
    float3 pathThroughput;
    float rayEta;
    float3 rayAbsorption;
    BSDFType rayType;
    int depth, stateFlags;
    unsigned int seed;
    Random rand;
};

struct OCounters { uint64_t samples = 0, radianceRays = 0, shadowRays = 0; };

struct OFrame {                          // host images, W*H each
    float4* accum;  uint32_t* frame;  float4* normal;  float4* albedo;  float4* color;
};

static inline float3 reinhardToneMap(const float3& color, const float white) {   // sv4/deviceProgram.cu:124-129
    const float luminance = 0.2126f * color.x + 0.7152f * color.y + 0.0722f * color.z;
    return (color * 1.0f) / (1.0f + luminance / white);
}

// tex2D<float4> on an RGBA8 array: wrap addressing, normalized coordinates, linear filter, normalized-float
// read (descriptor at sv4/SimplePathtracer.cpp:782-793).  Stated here in exact fp32 (CUDA's filter formula
// from the programming guide, without the TEX unit's 8-bit weight quantisation); the CUDA path evaluates the
// same expression on a linear RGBA8 buffer.
static inline float4 oracleTex2D(const OTexture& tex, float s, float t) {
    const float x = (s - floorf(s)) * (float)tex.w - 0.5f;
    const float y = (t - floorf(t)) * (float)tex.h - 0.5f;
    const float fx = floorf(x), fy = floorf(y);
    const float a = x - fx, b = y - fy;
    int i0 = (int)fx, j0 = (int)fy;
    int i1 = i0 + 1, j1 = j0 + 1;
    i0 = ((i0 % tex.w) + tex.w) % tex.w;  i1 = ((i1 % tex.w) + tex.w) % tex.w;
    j0 = ((j0 % tex.h) + tex.h) % tex.h;  j1 = ((j1 % tex.h) + tex.h) % tex.h;
    const uint32_t p00 = tex.px[(size_t)j0 * tex.w + i0], p10 = tex.px[(size_t)j0 * tex.w + i1];
    const uint32_t p01 = tex.px[(size_t)j1 * tex.w + i0], p11 = tex.px[(size_t)j1 * tex.w + i1];
    float out[4];
    for (int c = 0; c < 4; ++c) {
        const float c00 = (float)((p00 >> (8 * c)) & 255u) * (1.0f / 255.0f), c10 = (float)((p10 >> (8 * c)) & 255u) * (1.0f / 255.0f);
        const float c01 = (float)((p01 >> (8 * c)) & 255u) * (1.0f / 255.0f), c11 = (float)((p11 >> (8 * c)) & 255u) * (1.0f / 255.0f);
        out[c] = ((1.0f - a) * (1.0f - b)) * c00 + (a * (1.0f - b)) * c10 + ((1.0f - a) * b) * c01 + (a * b) * c11;
    }
    return make_float4(out[0], out[1], out[2], out[3]);
}

struct OTraceCtx {
    const OScene* scene;
    const b200pt_config* cfg;
    OCounters* counters;
};

static inline bool oracleTraceOcclusion(const OTraceCtx& C, const float3& o, const float3& d) {
    C.counters->shadowRays++;
    const float tminShadow = C.cfg->shadow_tmin > 0.f ? C.cfg->shadow_tmin : C.cfg->tmin;      // 0.01 in every variant (orig/deviceProgram.cu:268,310)
    return oracleAnyHit(*C.scene, o, d, tminShadow, C.cfg->tmax, C.cfg->shadow_mode == B200PT_SHADOW_CULL_BACK);
}

// sv4/deviceProgram.cu:301-341 (occludedWanted=false) and :344-384 (occludedWanted=true)
static inline float3 oracleSampleLights(const OTraceCtx& C, bool occludedWanted, const Material& material, float3 albedo,
                                        const float etaI, const float etaO, const float3& surfacePos,
                                        const float3& surfaceNormal, const float3& wo, Random& rand) {
    const float kProbeSamples = 1.f, kBsdfSamples = 1.f;
    float3 sum = make_float3(0.0f);
    for (int i = 0; i < kProbeSamples; ++i) {
        float3 skyColor;  float skyPdf;  float3 wi;
        ProbeSample(C.scene->probe, wi, skyColor, skyPdf, rand);
        const bool occluded = oracleTraceOcclusion(C, surfacePos, wi);
        if (occluded == occludedWanted) {
            float bsdfPdf = BSDFPdf(material, etaI, etaO, surfacePos, surfaceNormal, wo, wi);
            float3 f = BSDFEval(material, albedo, etaI, etaO, surfacePos, surfaceNormal, wo, wi);
            if (bsdfPdf > 0.0f) {
                int N = kProbeSamples + kBsdfSamples;
                float cbsdf = kBsdfSamples / N;
                float csky = float(kProbeSamples) / N;
                float weight = csky * skyPdf / (cbsdf * bsdfPdf + csky * skyPdf);
                if (weight > 0.0f) {
                    float3 val = weight * skyColor * f * abs(dot(wi, surfaceNormal)) / skyPdf * (1.0f / kProbeSamples);
                    sum += val;
                }
            }
        }
    }
    return sum;
}

// sv4/deviceProgram.cu:251-280
static inline void oracleMissProgram(const OTraceCtx& C, const float3& ray_dir, RadiancePRD* prd) {
    // Opt-in extension (cfg.mis_on_miss): the reference's commented-out block :257-267, with cbsdf = csky = 0.5 as in SampleLights
    // (its own `kBsdfSamples / N` is an integer division).  Only secondary rays: the primary ray's miss is the backplate.
    if (C.cfg->mis_on_miss && (prd->stateFlags & RAY_STATE_FLAGS_SECONDARY_RAY) != 0) {
        const float skyPdf = ProbePdf(C.scene->probe, ray_dir);
        const float cbsdf = 0.5f, csky = 0.5f;
        const float weight = cbsdf * prd->bsdfPdf / (cbsdf * prd->bsdfPdf + csky * skyPdf);
        prd->radiance += weight * make_float3(ProbeEval(C.scene->probe, ProbeDirToUV(ray_dir))) * prd->pathThroughput;
    }
    prd->albedo = make_float3(0.f);
    prd->normal = make_float3(0.f);
    prd->stateFlags |= RAY_STATE_FLAGS_DONE;
}

// sv4/deviceProgram.cu:590-703
static inline void oracleClosestHitProgram(const OTraceCtx& C, const OHit& hit, const float3& ray_org,
                                           const float3& ray_dir, RadiancePRD* prd) {
    const OScene& S = *C.scene;
    const int meshID = S.meshOf[hit.gid];
    const OMesh& mesh = S.meshes[meshID];
    const int prim_idx = hit.gid - (int)mesh.firstGid;
    const uint3 index = mesh.index[prim_idx];
    const float3 v0 = mesh.vertex[index.x], v1 = mesh.vertex[index.y], v2 = mesh.vertex[index.z];
    const float3 N_0 = normalize(cross(v1 - v0, v2 - v0));
    float3 N = faceforward(N_0, -ray_dir, N_0);
    const float t = hit.t;
    const float3 P = ray_org + t * ray_dir;
    const Material& material = mesh.material;
    float outEta;  float3 outAbsorption;

    if ((material.flags & MATERIAL_FLAG_SHADOW_CATCHER) != 0 && (prd->stateFlags & RAY_STATE_FLAGS_SECONDARY_RAY) != 0) {
        prd->origin = P;  prd->direction = ray_dir;  --prd->depth;
        return;
    }
    prd->normal = N;
    prd->albedo = material.color;
    if (mesh.textureID >= 0 && !mesh.texcoord.empty()) {
        const float u = hit.u, v = hit.v;
        const float2 tc = (1.f - u - v) * mesh.texcoord[index.x] + u * mesh.texcoord[index.y] + v * mesh.texcoord[index.z];
        float4 fromTexture = oracleTex2D(S.textures[mesh.textureID], tc.x, tc.y);
        prd->albedo = make_float3(fromTexture);
    }
    float& bsdfPdf = prd->bsdfPdf;
    if (prd->rayEta == 1.0f) { outEta = material.GetIndexOfRefraction();  outAbsorption = material.absorption; }
    else { outEta = 1.0f;  outAbsorption = make_float3(0.0f); }

    if ((material.flags & MATERIAL_FLAG_SHADOW_CATCHER) == 0) {
        float3 lightSample = oracleSampleLights(C, false, material, prd->albedo, prd->rayEta, outEta, P, N, -ray_dir, prd->rand);
        prd->radiance += prd->pathThroughput * lightSample;
        prd->alpha = make_float3(1.0f);
    } else {
        float3 shadowSample = oracleSampleLights(C, true, material, prd->albedo, prd->rayEta, outEta, P, N, -ray_dir, prd->rand);
        prd->alpha += prd->pathThroughput * shadowSample;
    }
    if ((prd->stateFlags & RAY_STATE_FLAGS_SECONDARY_RAY) == 0) prd->radiance += material.emission;

    float3 u, v;
    BasisFromVector(N, &u, &v);
    float3 bsdfDir;  BSDFType bsdfType;
    BSDFSample(material, prd->rayEta, outEta, P, u, v, N, -ray_dir, bsdfDir, bsdfPdf, bsdfType, prd->rand);
    if (bsdfPdf <= 0.0f) { prd->stateFlags |= RAY_STATE_FLAGS_DONE;  return; }
    float3 f = BSDFEval(material, prd->albedo, prd->rayEta, outEta, P, N, -ray_dir, bsdfDir);
    if (dot(bsdfDir, N) <= 0.0f) { prd->rayEta = outEta;  prd->rayAbsorption = outAbsorption; }
    prd->pathThroughput *= f * abs(dot(N, bsdfDir)) / bsdfPdf;
    prd->rayType = bsdfType;
    prd->direction = bsdfDir;
    prd->origin = P;
    prd->stateFlags |= RAY_STATE_FLAGS_SECONDARY_RAY;
    // Opt-in extension (cfg.russian_roulette): from the third segment on, continue with probability p = min(1, max(T));
    // one extra Randf per decision, so the RNG stream differs from the reference's from here on.  This segment's radiance is kept.
    if (C.cfg->russian_roulette && prd->depth + 1 >= 2) {
        const float3 T = prd->pathThroughput;
        const float p = fminf(1.0f, fmaxf(T.x, fmaxf(T.y, T.z)));
        const float r = prd->rand.Randf();
        if (!(r < p)) prd->stateFlags |= RAY_STATE_FLAGS_RR_KILLED;
        else prd->pathThroughput = prd->pathThroughput / p;
    }
}
