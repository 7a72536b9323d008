// b200pt — device shading library: the reference's Disney BSDF, probe sampling, RNG and display transform
// re-stated for the wavefront kernels.  Compiled with -fmad=false and IEEE div/sqrt so that every basic
// fp32 operation rounds exactly like the host oracle; only libm-level functions (sinf, cosf, acosf,
// atan2f, logf, powf) can differ, by <= 2 ulp.  Where the reference mixes double literals into float
// expressions (e.g. `1.0 / sqrt(..)`, `.3 * c.x`, `0.5 + ...`) the double arithmetic is kept: it defines the bits.
// reference files: cuda/random.h, sv4/maths.h, sv4/sample.h, sv4/Disney.cuh, sv4/Probe.cuh, cuda/helpers.h.
#pragma once
#include "common.cuh"

#define DEV __device__ __forceinline__

// ---- float3 algebra with sutil/vec_math.h rounding ---------------------------------------------------
DEV float3 f3(float x, float y, float z) { return make_float3(x, y, z); }
DEV float3 f3(float s) { return make_float3(s, s, s); }
DEV float3 operator+(float3 a, float3 b) { return f3(a.x + b.x, a.y + b.y, a.z + b.z); }
DEV float3 operator-(float3 a, float3 b) { return f3(a.x - b.x, a.y - b.y, a.z - b.z); }
DEV float3 operator-(float3 a) { return f3(-a.x, -a.y, -a.z); }
DEV float3 operator*(float3 a, float3 b) { return f3(a.x * b.x, a.y * b.y, a.z * b.z); }
DEV float3 operator*(float3 a, float s) { return f3(a.x * s, a.y * s, a.z * s); }
DEV float3 operator*(float s, float3 a) { return f3(a.x * s, a.y * s, a.z * s); }
DEV float3 operator/(float3 a, float s) { const float inv = 1.0f / s; return a * inv; }       // vec_math.h:483
DEV float dot3(float3 a, float3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
DEV float3 cross3(float3 a, float3 b) { return f3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
DEV float3 normalize3(float3 v) { const float inv = 1.0f / sqrtf(dot3(v, v)); return v * inv; }   // vec_math.h:545
DEV float clampf(float f, float a, float b) { return fmaxf(a, fminf(f, b)); }                   // vec_math.h:119
DEV float3 clamp3(float3 v, float a, float b) { return f3(clampf(v.x, a, b), clampf(v.y, a, b), clampf(v.z, a, b)); }
DEV float lerpf(float a, float b, float t) { return a + t * (b - a); }
DEV float3 lerp3(float3 a, float3 b, float t) { return a + t * (b - a); }
DEV float sqrf(float a) { return a * a; }

#define kPi (3.141592653589793f)
#define k2Pi (3.141592653589793f * 2.0f)
#define kInvPi (1.0f / kPi)
#define kInv2Pi (1.0f / k2Pi)

// ---- seeds: cuda/random.h:34-59,96-99 ---------------------------------------------------------------
DEV unsigned tea4(unsigned v0, unsigned v1) {
    unsigned s = 0;
#pragma unroll
    for (int n = 0; n < 4; ++n) {
        s += 0x9e3779b9u;
        v0 += ((v1 << 4) + 0xa341316cu) ^ (v1 + s) ^ ((v1 >> 5) + 0xc8013ea4u);
        v1 += ((v0 << 4) + 0xad90777du) ^ (v0 + s) ^ ((v0 >> 5) + 0x7e95761eu);
    }
    return v0;
}
DEV float lcgRnd(unsigned& prev) {
    prev = 1664525u * prev + 1013904223u;
    return (float)(prev & 0x00FFFFFFu) / (float)0x01000000;
}

// ---- class Random: sv4/maths.h:170-226 ---------------------------------------------------------------
struct Rng {
    unsigned s1, s2;
    DEV void seed(unsigned seed) { s1 = 315645664u + seed;  s2 = s1 ^ 0x13ab45feu; }
    DEV unsigned next() {
        s1 = (s2 ^ ((s1 << 5) | (s1 >> 27))) ^ (s1 * s2);
        s2 = s1 ^ ((s2 << 12) | (s2 >> 20));
        return s1;
    }
    DEV float randf() { return clampf((float)next() * (1.0f / 4294967296.0f), 0.f, 0.999999f); }
    DEV float randf01() { const float t = randf(); return (1.0f - t) * 0.0f + t * 1.0f; }        // Randf(0,1), maths.h:212-216
};
DEV void sample2D(Rng& r, float& u1, float& u2) { u1 = r.randf01();  u2 = r.randf01(); }   // sample.h:254-260

// ---- frames / hemisphere sampling: sv4/maths.h:94-108,144-156,242-275 --------------------------------
DEV void basisFromVector(float3 w, float3& u, float3& v) {
    if (fabsf(w.x) > fabsf(w.y)) {
        const float invLen = (float)(1.0 / (double)sqrtf(w.x * w.x + w.z * w.z));
        u = f3(-w.z * invLen, 0.0f, w.x * invLen);
    } else {
        const float invLen = (float)(1.0 / (double)sqrtf(w.y * w.y + w.z * w.z));
        u = f3(0.0f, w.z * invLen, -w.y * invLen);
    }
    v = cross3(w, u);
}
DEV float3 safeNormalize(float3 a) {
    const float m = dot3(a, a);
    if (m > 0.0f) return a * (float)(1.0 / (double)sqrtf(m));
    return f3(0.0f);
}
DEV float3 uniformSampleHemisphere(Rng& r) {
    const float z = r.randf01();
    const float w = sqrtf(1.0f - z * z);
    const float phi = k2Pi * r.randf01();
    return f3(cosf(phi) * w, sinf(phi) * w, z);
}
DEV float3 cosineSampleHemisphere(float u1, float u2) {
    const float r = sqrtf(u1);
    const float theta = k2Pi * u2;
    const float sx = r * cosf(theta), sy = r * sinf(theta);
    const float z = sqrtf(fmaxf(0.0f, 1.0f - sx * sx - sy * sy));
    return f3(sx, sy, z);
}

// code-size knobs (tuning builds): out-of-line copies of the small helpers with IEEE divisions / square roots
#ifdef SHADE_NI_GGX
#define SHADE_NI_GGX_ATTR __device__ __noinline__
#else
#define SHADE_NI_GGX_ATTR DEV
#endif
#ifdef SHADE_NI_FRESNEL
#define SHADE_NI_FRESNEL_ATTR __device__ __noinline__
#else
#define SHADE_NI_FRESNEL_ATTR DEV
#endif
// ---- Disney BSDF: sv4/Disney.cuh ---------------------------------------------------------------------
enum { eReflected = 0, eTransmitted = 1, eSpecular = 2 };

DEV bool refractDir(float3 wi, float3 n, float eta, float3& wt) {                  // :36-49
    const float cosI = dot3(n, wi);
    const float sin2I = fmaxf(0.0f, 1.0f - cosI * cosI);
    const float sin2T = eta * eta * sin2I;
    if (sin2T >= 1.0f) return false;
    const float cosT = sqrtf(1.0f - sin2T);
    wt = eta * (-wi) + (eta * cosI - cosT) * n;
    return true;
}
DEV float schlickFresnel(float u) { const float m = clampf(1.0f - u, 0.0f, 1.0f); const float m2 = m * m; return m2 * m2 * m; }
DEV float GTR1(float NDotH, float a) {
    if (a >= 1.0f) return kInvPi;
    const float a2 = a * a;
    const float t = 1.0f + (a2 - 1.0f) * NDotH * NDotH;
    return (a2 - 1.0f) / (kPi * logf(a2) * t);
}
DEV float GTR2(float NDotH, float a) {
    const float a2 = a * a;
    const float t = 1.0f + (a2 - 1.0f) * NDotH * NDotH;
    return a2 / (kPi * t * t);
}
SHADE_NI_GGX_ATTR float smithGGX(float NDotv, float alphaG) {
    const float a = alphaG * alphaG, b = NDotv * NDotv;
    return 1.0f / (NDotv + sqrtf(a + b - a * b));
}
// Fr (:81-98).  Both media of a vertex are fixed while it is shaded, so the two divisions that involve only the indices are done ONCE per vertex
// (EtaPair) instead of in every one of the up to five Fr evaluations: the same operands, so the same bits.
struct EtaPair { float sq, inv; };                               // sqr(etaI / etaT), etaT / etaI
DEV EtaPair makeEtaPair(float etaI, float etaT) { EtaPair e;  e.sq = sqrf(etaI / etaT);  e.inv = etaT / etaI;  return e; }
SHADE_NI_FRESNEL_ATTR float fresnelDielectric(float VDotN, const EtaPair e) {
    const float sinT2 = e.sq * (1.0f - VDotN * VDotN);
    if (sinT2 > 1.0f) return 1.0f;
    const float LDotN = sqrtf(1.0f - sinT2);
    const float eta = e.inv;
    const float r1 = (VDotN - eta * LDotN) / (VDotN + eta * LDotN);
    const float r2 = (LDotN - eta * VDotN) / (LDotN + eta * VDotN);
    return 0.5f * (sqrf(r1) + sqrf(r2));
}
DEV float fresnelDielectric(float VDotN, float etaI, float etaT) { return fresnelDielectric(VDotN, makeEtaPair(etaI, etaT)); }

// FNV = Fr(n.V, etaI, etaO): needed by BSDFSample's transmission branch and by BSDFPdf of every direction above the surface - the same value each time,
// so shadeKernel computes it once per vertex and passes it in
DEV float bsdfPdf(const Material& mat, float FNV, float3 n, float3 V, float3 L) {   // :152-193
    if (dot3(L, n) <= 0.0f) {
        const float brdfPdf = kInv2Pi * mat.subsurface * 0.5f;
        return lerpf(brdfPdf, 0.0f, mat.transmission);
    }
    const float F = FNV;
    const float a = fmaxf(0.001f, mat.roughness);
    const float3 half = safeNormalize(L + V);
    const float cosThetaHalf = fabsf(dot3(half, n));
    const float pdfHalf = GTR2(cosThetaHalf, a) * cosThetaHalf;
    const float pdfSpec = 0.25f * pdfHalf / fmaxf(1.e-6f, dot3(L, half));
    const float pdfDiff = fabsf(dot3(L, n)) * kInvPi * (1.0f - mat.subsurface);
    const float bsdf = pdfSpec * F;
    const float brdf = lerpf(pdfDiff, pdfSpec, 0.5f);
    return lerpf(brdf, bsdf, mat.transmission);
}

DEV float3 ggxReflect(const Material& mat, float3 U, float3 V, float3 N, float3 view, float r1, float r2) {   // :209-228, :281-305
    const float a = fmaxf(0.001f, mat.roughness);
    const float phiHalf = r1 * k2Pi;
    const float cosThetaHalf = sqrtf((1.0f - r2) / (1.0f + (sqrf(a) - 1.0f) * r2));
    const float sinThetaHalf = sqrtf(fmaxf(0.0f, 1.0f - sqrf(cosThetaHalf)));
    const float sinPhiHalf = sinf(phiHalf);
    const float cosPhiHalf = cosf(phiHalf);
    float3 half = U * (sinThetaHalf * cosPhiHalf) + V * (sinThetaHalf * sinPhiHalf) + N * cosThetaHalf;
    if (dot3(half, view) <= 0.0f) half = half * -1.0f;
    return 2.0f * dot3(view, half) * half - view;
}

// The direction-drawing half of DisneySample (:197-315).  needPdf = the reference goes on to `pdf = DisneyPdf(...)` for the drawn
// direction (every branch but the two refraction returns, which set pdf themselves); shadeKernel evaluates that pdf in the same
// rolled loop as the NEE direction's, so the kernel holds ONE copy of bsdfPdf/bsdfEval instead of two (instruction-cache footprint).
DEV void bsdfSampleDir(const Material& mat, float etaI, float etaO, float FNV, float3 U, float3 V, float3 N, float3 view,
                       float3& light, float& pdf, int& type, bool& needPdf, Rng& rand) {
    needPdf = false;
    // The GGX half-vector reflection is reached from two branches (transmission lobe after the Fresnel test, :209-228; specular lobe, :281-305).  Both only
    // set `ggx` and share ONE copy of ggxReflect below: lanes of a warp that took different branches then run it together instead of one copy after the
    // other.  The random numbers are drawn inside the branches, in the reference's order.
    bool ggx = false;
    float r1 = 0.f, r2 = 0.f;
    if (rand.randf() < mat.transmission) {
        const float F = FNV;
        if (rand.randf() < F) {
            sample2D(rand, r1, r2);
            ggx = true;
        } else {
            const float eta = etaI / etaO;
            if (refractDir(view, N, eta, light)) { type = eSpecular;  pdf = (1.0f - F) * mat.transmission;  return; }
            pdf = 0.0f;
            return;
        }
    } else {
        sample2D(rand, r1, r2);
        if (rand.randf() < 0.5f) {
            if (rand.randf() < mat.subsurface) {
                const float3 d = uniformSampleHemisphere(rand);
                light = U * d.x + V * d.y - N * d.z;
                type = eTransmitted;
            } else {
                const float3 d = cosineSampleHemisphere(r1, r2);
                light = U * d.x + V * d.y + N * d.z;
                type = eReflected;
            }
        } else ggx = true;
    }
    if (ggx) { light = ggxReflect(mat, U, V, N, view, r1, r2);  type = eReflected; }
    needPdf = true;
}
DEV void bsdfSample(const Material& mat, float etaI, float etaO, float3 U, float3 V, float3 N, float3 view,
                    float3& light, float& pdf, int& type, Rng& rand) {            // :197-315
    bool needPdf;
    const float FNV = fresnelDielectric(dot3(N, view), etaI, etaO);
    bsdfSampleDir(mat, etaI, etaO, FNV, U, V, N, view, light, pdf, type, needPdf, rand);
    if (needPdf) pdf = bsdfPdf(mat, FNV, N, view, light);
}

DEV float3 bsdfEval(const Material& mat, float3 albedo, const EtaPair eta, float FNV, float3 N, float3 V, float3 L) {   // :318-427
    const float NDotL = dot3(N, L), NDotV = dot3(N, V);
    const float3 H = normalize3(L + V);
    const float NDotH = dot3(N, H), LDotH = dot3(L, H);
    const float3 Cdlin = albedo;
    const float Cdlum = (float)(.3 * (double)Cdlin.x + .6 * (double)Cdlin.y + .1 * (double)Cdlin.z);
    const float3 Ctint = Cdlum > 0.0f ? Cdlin / Cdlum : f3(1.0f);
    const float3 Cspec0 = lerp3((float)((double)mat.specular * .08) * lerp3(f3(1.0f), Ctint, mat.specularTint), Cdlin, mat.metallic);
    float3 bsdf = f3(0.0f), brdf = f3(0.0f);
    // Above the surface both lobes need the same GTR2 and Smith terms (same arguments), and most materials run both lobes (the reference's default has
    // transmission 0.4): computed once here.  `above` is the reference's else-branch of `NDotL <= 0` (so a NaN goes where it goes there).
    const bool above = !(NDotL <= 0.0f);
    float a = 0.0f, Ds = 0.0f, Gs = 0.0f;
    if (above) {
        a = fmaxf(0.001f, mat.roughness);
        Ds = GTR2(NDotH, a);
        Gs = smithGGX(NDotV, a) * smithGGX(NDotL, a);
    }
    if (mat.transmission > 0.0f) {
        // one copy of Fr for the two sides: below the surface it is evaluated at N.V (:332-336), above at L.H (:338-345)
        const float F = above ? fresnelDielectric(LDotH, eta) : FNV;               // below: Fr(N.V), the per-vertex value
        if (!above) {
            bsdf = f3(mat.transmission * (1.0f - F) / fabsf(NDotL) * (1.0f - mat.metallic));
        } else {
            const float3 Fs = lerp3(Cspec0, f3(1.0f), F);
            bsdf = Gs * Fs * Ds;
        }
    }
    if (mat.transmission < 1.0f) {
        if (!above) {
            if (mat.subsurface > 0.0f) {
                const float3 s = f3(sqrtf(mat.color.x), sqrtf(mat.color.y), sqrtf(mat.color.z));
                const float FL = schlickFresnel(fabsf(NDotL)), FV = schlickFresnel(NDotV);
                const float Fd = (1.0f - 0.5f * FL) * (1.0f - 0.5f * FV);
                brdf = kInvPi * s * mat.subsurface * Fd * (1.0f - mat.metallic);
            }
        } else {
            const float FH = schlickFresnel(LDotH);
            const float3 Fs = lerp3(Cspec0, f3(1.f), FH);
            const float FL = schlickFresnel(NDotL), FV = schlickFresnel(NDotV);
            const float Fd90 = (float)(0.5 + (double)(2.0f * LDotH * LDotH * mat.roughness));
            const float Fd = lerpf(1.0f, Fd90, FL) * lerpf(1.0f, Fd90, FV);
            // Clearcoat lobe (:411-417).  With clearcoat == 0 - every material the reference constructs (Material.h:13-38) - the term is exactly +0: Gr, Fc
            // and Dr are finite and positive for N.L > 0 (each Smith denominator is >= 0.25; GTR1's t lies in [a^2, 1], a <= 0.1), and x + (+0) == x, so the
            // logf, the three divisions and the two square roots are skipped without changing a bit of the result.
            float3 coat = f3(0.0f);
            if (mat.clearcoat != 0.0f) {
                const float Dr = GTR1(NDotH, lerpf(.1f, .001f, mat.clearcoatGloss));
                const float Fc = lerpf(.04f, 1.0f, FH);
                const float Gr = smithGGX(NDotL, .25f) * smithGGX(NDotV, .25f);
                coat = f3(mat.clearcoat * Gr * Fc * Dr);
            }
            brdf = kInvPi * Fd * Cdlin * (1.0f - mat.metallic) * (1.0f - mat.subsurface) + Gs * Fs * Ds + coat;
        }
    }
    return lerp3(brdf, bsdf, mat.transmission);
}

DEV float bsdfPdf(const Material& mat, float etaI, float etaO, float3 n, float3 V, float3 L) { return bsdfPdf(mat, fresnelDielectric(dot3(n, V), etaI, etaO), n, V, L); }
DEV float3 bsdfEval(const Material& mat, float3 albedo, float etaI, float etaO, float3 N, float3 V, float3 L) {
    return bsdfEval(mat, albedo, makeEtaPair(etaI, etaO), fresnelDielectric(dot3(N, V), etaI, etaO), N, V, L);
}

// ---- environment probe: sv4/Probe.cuh ----------------------------------------------------------------
DEV float2 probeDirToUV(float3 dir) {                                               // :38-46
    const float theta = acosf(clampf(dir.y, -1.0f, 1.0f));
    const float phi = (dir.x == 0.0f && dir.z == 0.0f) ? 0.0f : atan2f(dir.z, dir.x);
    return make_float2((kPi + phi) * kInvPi * 0.5f, theta * kInvPi);
}
DEV float3 probeUVToDir(float u, float v) {                                         // :48-58
    const float theta = v * kPi, phi = u * 2.0f * kPi;
    return f3(-sinf(theta) * cosf(phi), cosf(theta), -sinf(theta) * sinf(phi));
}
DEV int clampi(int f, int a, int b) { return max(a, min(f, b)); }
DEV float4 probeEval(const Probe& P, float2 uv) {                                    // :61-67
    const int px = clampi((int)(uv.x * P.width), 0, P.width - 1);
    const int py = clampi((int)(uv.y * P.height), 0, P.height - 1);
    return __ldg(reinterpret_cast<const float4*>(P.data) + (py * P.width + px));
}
DEV float probePdf(const Probe& P, float3 d) {                                       // :69-93
    const float2 uv = probeDirToUV(d);
    const int col = clampi((int)(uv.x * P.width), 0, P.width - 1);
    const int row = clampi((int)(uv.y * P.height), 0, P.height - 1);
    float pdf = __ldg(P.pdfValuesX + (row * P.width + col)) * __ldg(P.pdfValuesY + row);
    const float sinTheta = sinf(uv.y * kPi);
    if (fabsf(sinTheta) < 0.0001f) pdf = 0.0f;
    else pdf *= (float)P.width * (float)P.height / (2.0f * kPi * kPi * sinTheta);
    return pdf;
}
DEV int lowerBound(const float* __restrict__ a, int lower, int upper, float value) {   // :119-136
    while (lower < upper) {
        const int mid = lower + (upper - lower) / 2;
        if (__ldg(a + mid) < value) lower = mid + 1;
        else upper = mid;
    }
    return lower;
}
// Guide tables (built by b200pt_set_probe*): guide[k] = LowerBound(cdf, k / G) for k = 0..G.  LowerBound is monotone in `value`, so
// for value in [k/G, (k+1)/G) the answer lies in [guide[k], guide[k+1]] and a search of that short range with the SAME comparison
// returns the SAME index as the reference's search of the whole array (each CDF is non-decreasing) — in ~3 dependent loads
// instead of log2(n) ~ 10-11.  G is a power of two, so k/G and value*G are exact.
#ifndef B200PT_GUIDE_Y
#define B200PT_GUIDE_Y 1024
#endif
#ifndef B200PT_GUIDE_X
#define B200PT_GUIDE_X 256
#endif
DEV int guidedLowerBound(const float* __restrict__ a, int base, const int* __restrict__ guide, int G, float value) {
    const int k = min(G - 1, max(0, (int)(value * (float)G)));
    const int lo = __ldg(guide + k), hi = __ldg(guide + k + 1);
    return lowerBound(a, base + lo, base + hi, value) - base;
}
// the same search on copies of cdfValuesY and its guide table in shared memory (shadeKernel stages them per CTA when they fit): same data, same
// comparisons, same index - the row search then costs shared-memory latency instead of ~4 dependent L1/L2 round trips
DEV int lowerBoundShared(const float* a, int lower, int upper, float value) {
    while (lower < upper) {
        const int mid = lower + (upper - lower) / 2;
        if (a[mid] < value) lower = mid + 1;
        else upper = mid;
    }
    return lower;
}
DEV void probeSample(const Probe& P, const int* __restrict__ guide, float3& dir, float3& color, float& pdf, Rng& rand,
                     const float* sCdfY = nullptr, const int* sGuideY = nullptr) {   // :138-169
    float r1, r2;
    sample2D(rand, r1, r2);
    int row, col;
    if (guide && sCdfY) {
        const int k = min(B200PT_GUIDE_Y - 1, max(0, (int)(r1 * (float)B200PT_GUIDE_Y)));
        row = lowerBoundShared(sCdfY, sGuideY[k], sGuideY[k + 1], r1);
        col = guidedLowerBound(P.cdfValuesX, row * P.width, guide + (B200PT_GUIDE_Y + 1) + (size_t)row * (B200PT_GUIDE_X + 1), B200PT_GUIDE_X, r2);
    } else if (guide) {
        row = guidedLowerBound(P.cdfValuesY, 0, guide, B200PT_GUIDE_Y, r1);
        col = guidedLowerBound(P.cdfValuesX, row * P.width, guide + (B200PT_GUIDE_Y + 1) + (size_t)row * (B200PT_GUIDE_X + 1), B200PT_GUIDE_X, r2);
    } else {
        row = lowerBound(P.cdfValuesY, 0, P.height, r1);
        col = lowerBound(P.cdfValuesX, row * P.width, (row + 1) * P.width, r2) - row * P.width;
    }
    const float4 c = __ldg(reinterpret_cast<const float4*>(P.data) + (row * P.width + col));
    color = f3(c.x, c.y, c.z);
    pdf = __ldg(P.pdfValuesX + row * P.width + col) * __ldg(P.pdfValuesY + row);
    const float u = col / (float)P.width, v = row / (float)P.height;
    const float theta = v * kPi, phi = u * 2.0f * kPi;                      // ProbeUVToDir's angles (:48-58)
    const float sinTheta = sinf(theta);                                     // the pdf's sin(theta) (:160) and ProbeUVToDir's are the same value: once
    if (sinTheta == 0.0f) pdf = 0.0f;
    else pdf *= (float)(P.width * P.height) / (2.0f * kPi * kPi * sinTheta);
    dir = f3(-sinTheta * cosf(phi), cosf(theta), -sinTheta * sinf(phi));
}

// ---- display transform: cuda/helpers.h:35-62, sv4/deviceProgram.cu:124-129 ---------------------------
DEV float toSRGB1(float c) {
    const float powed = powf(c, 1.0f / 2.4f);
    return c < 0.0031308f ? 12.92f * c : 1.055f * powed - 0.055f;
}
DEV unsigned quantize8(float x) {
    x = clampf(x, 0.0f, 1.0f);
    return min((unsigned)(x * 256.0f), 255u);
}
DEV uint32_t makeColor(float3 c) {
    const float3 cc = clamp3(c, 0.0f, 1.0f);
    return quantize8(toSRGB1(cc.x)) | (quantize8(toSRGB1(cc.y)) << 8) | (quantize8(toSRGB1(cc.z)) << 16) | (255u << 24);
}
DEV float3 reinhard(float3 color, float white) {
    const float lum = 0.2126f * color.x + 0.7152f * color.y + 0.0722f * color.z;
    return (color * 1.0f) / (1.0f + lum / white);
}
