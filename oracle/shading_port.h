// ORACLE — TEST INFRASTRUCTURE, NOT PRODUCT CODE.
// CPU restatement ("port") of the reference's shading library, in plain C++ with IEEE fp32 math.
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may use it.
//
// Pinning: every function here is checked bit-for-bit (on the same host libm) against the reference's
// own headers host-compiled into oracle/_ref/liboracle_ref.so (see oracle/Makefile, oracle/ref_backend.h)
// and against the golden vectors that build wrote to tests/golden/ (tests/golden/make_golden.py).
// The OptiX boundary itself (closed libnvoptix arithmetic) has no reference test: "parity unpinned" there.
//
// Each function cites the reference lines it follows, relative to /root/reference/
// (sv4/ = HelloPathtracing_sv4_vmv23/).  Mixed float/double arithmetic of the reference (unsuffixed
// literals such as `1.0 / sqrt(..)` or `.3 * c.x`) is reproduced deliberately: it defines the bits.
#pragma once
#include <stdlib.h>
#include <math.h>
#include <cmath>
#include <cstdint>
#include <vector_types.h>
#include <vector_functions.h>
#include "../include/ref_api/Material.h"
#include "../include/ref_api/Probe.h"

// ---------------------------------------------------------------------------------------------------
// float3 algebra with the rounding behaviour of sutil/vec_math.h
// ---------------------------------------------------------------------------------------------------
static inline float3 make_float3(float s) { return make_float3(s, s, s); }
static inline float3 make_float3(const float4& a) { return make_float3(a.x, a.y, a.z); }
static inline float3 operator+(const float3& a, const float3& b) { return make_float3(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline float3 operator-(const float3& a, const float3& b) { return make_float3(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline float3 operator-(const float3& a) { return make_float3(-a.x, -a.y, -a.z); }
static inline float3 operator*(const float3& a, const float3& b) { return make_float3(a.x * b.x, a.y * b.y, a.z * b.z); }
static inline float3 operator*(const float3& a, float s) { return make_float3(a.x * s, a.y * s, a.z * s); }
static inline float3 operator*(float s, const float3& a) { return make_float3(a.x * s, a.y * s, a.z * s); }
static inline float3 operator/(const float3& a, const float3& b) { return make_float3(a.x / b.x, a.y / b.y, a.z / b.z); }
// vec_math.h:483-487: division by a scalar multiplies by the reciprocal
static inline float3 operator/(const float3& a, float s) { float inv = 1.0f / s; return a * inv; }
static inline void operator+=(float3& a, const float3& b) { a.x += b.x; a.y += b.y; a.z += b.z; }
static inline void operator*=(float3& a, const float3& b) { a.x *= b.x; a.y *= b.y; a.z *= b.z; }
static inline void operator*=(float3& a, float s) { a.x *= s; a.y *= s; a.z *= s; }
static inline void operator/=(float3& a, float s) { float inv = 1.0f / s; a *= inv; }
static inline float2 operator*(float s, const float2& a) { return make_float2(a.x * s, a.y * s); }
static inline float2 operator+(const float2& a, const float2& b) { return make_float2(a.x + b.x, a.y + b.y); }
static inline float2 operator-(const float2& a, float b) { return make_float2(a.x - b, a.y - b); }
static inline float dot(const float3& a, const float3& b) { return a.x * b.x + a.y * b.y + a.z * b.z; }   // :526
static inline float3 cross(const float3& a, const float3& b) {                                           // :533
    return make_float3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
static inline float length(const float3& v) { return sqrtf(dot(v, v)); }                                  // :539
static inline float3 normalize(const float3& v) { float invLen = 1.0f / sqrtf(dot(v, v)); return v * invLen; }   // :545
static inline float clamp(float f, float a, float b) { return fmaxf(a, fminf(f, b)); }                    // :119
static inline int clamp(int f, int a, int b) { return f < a ? a : (f > b ? b : f); }                      // :828 (max(a,min(f,b)))
static inline float3 clamp(const float3& v, float a, float b) { return make_float3(clamp(v.x, a, b), clamp(v.y, a, b), clamp(v.z, a, b)); }
static inline float3 clamp(const float3& v, const float3& a, const float3& b) {
    return make_float3(clamp(v.x, a.x, b.x), clamp(v.y, a.y, b.y), clamp(v.z, a.z, b.z));
}
static inline float lerp(float a, float b, float t) { return a + t * (b - a); }                           // :98
static inline float3 lerp(const float3& a, const float3& b, float t) { return a + t * (b - a); }          // :500
static inline float3 faceforward(const float3& n, const float3& i, const float3& nref) {                 // :567
    return n * copysignf(1.0f, dot(i, nref));
}

// ---------------------------------------------------------------------------------------------------
// constants — sv4/maths.h:30-33
// ---------------------------------------------------------------------------------------------------
#define kPi (3.141592653589793f)
#define k2Pi (3.141592653589793f * 2.0f)
#define kInvPi (1.0f / kPi)
#define kInv2Pi (1.0f / k2Pi)

// ---------------------------------------------------------------------------------------------------
// seed hash + LCG — cuda/random.h:34-49, 53-59, 96-99
// ---------------------------------------------------------------------------------------------------
template <unsigned int N>
static inline unsigned int tea(unsigned int val0, unsigned int val1) {
    unsigned int a = val0, b = val1, sum = 0;
    for (unsigned int round = 0; round < N; ++round) {
        sum += 0x9e3779b9u;
        a += ((b << 4) + 0xa341316cu) ^ (b + sum) ^ ((b >> 5) + 0xc8013ea4u);
        b += ((a << 4) + 0xad90777du) ^ (a + sum) ^ ((a >> 5) + 0x7e95761eu);
    }
    return a;
}
static inline unsigned int lcg(unsigned int& prev) {
    prev = 1664525u * prev + 1013904223u;
    return prev & 0x00FFFFFFu;
}
static inline float rnd(unsigned int& prev) { return (float)lcg(prev) / (float)0x01000000; }

// ---------------------------------------------------------------------------------------------------
// class Random — sv4/maths.h:170-226
// ---------------------------------------------------------------------------------------------------
class Random {
public:
    inline Random(int seed = 0) {
        seed1 = 315645664 + seed;
        seed2 = seed1 ^ 0x13ab45fe;
    }
    inline unsigned int Rand() {
        seed1 = (seed2 ^ ((seed1 << 5) | (seed1 >> 27))) ^ (seed1 * seed2);
        seed2 = seed1 ^ ((seed2 << 12) | (seed2 >> 20));
        return seed1;
    }
    inline float Randf() {
        unsigned int value = Rand();
        unsigned int limit = 0xffffffffu;
        return clamp((float)value * (1.0f / (float)limit), 0.f, 0.999999f);
    }
    inline float Randf(float lo, float hi) {
        float t = Randf();
        return (1.0f - t) * lo + t * hi;
    }
    unsigned int seed1, seed2;
};

// sv4/sample.h:254-260 (USE_RANDOM branch — the only live one)
static inline void Sample2D(Random& rand, float& u1, float& u2) {
    u1 = rand.Randf(0.0f, 1.0f);
    u2 = rand.Randf(0.0f, 1.0f);
}

// ---------------------------------------------------------------------------------------------------
// frames and hemisphere sampling — sv4/maths.h:73-108, 144-168, 242-275
// ---------------------------------------------------------------------------------------------------
static inline float maxf(float a, float b) { return a > b ? a : b; }
static inline float sqr(float a) { return a * a; }

static inline void BasisFromVector(const float3& w, float3* u, float3* v) {      // :94-108 (double reciprocal)
    if (fabs(w.x) > fabs(w.y)) {
        float invLen = 1.0 / sqrt(w.x * w.x + w.z * w.z);
        *u = make_float3(-w.z * invLen, 0.0f, w.x * invLen);
    } else {
        float invLen = 1.0 / sqrt(w.y * w.y + w.z * w.z);
        *u = make_float3(0.0f, w.z * invLen, -w.y * invLen);
    }
    *v = cross(w, *u);
}
static inline float3 SafeNormalize(const float3& a, const float3& fallback = make_float3(0.0f)) {   // :144-156
    float m = dot(a, a);
    if (m > 0.0) return a * (float)(1.0 / sqrt(m));
    return fallback;
}
static inline float Luminance(const Color& c) { return c.x * 0.3f + c.y * 0.6f + c.z * 0.1f; }      // :165-168

static inline float3 UniformSampleHemisphere(Random& rand) {                     // :242-254
    float z = rand.Randf(0.0f, 1.0f);
    float w = sqrt(1.0f - z * z);
    float phi = k2Pi * rand.Randf(0.0f, 1.0f);
    float x = cosf(phi) * w;
    float y = sinf(phi) * w;
    return make_float3(x, y, z);
}
static inline float2 UniformSampleDisc(float u1, float u2) {                     // :256-262
    float r = sqrt(u1);
    float theta = k2Pi * u2;
    return make_float2(r * cos(theta), r * sin(theta));
}
static inline float3 CosineSampleHemisphere(float u1, float u2) {                // :271-277
    float2 s = UniformSampleDisc(u1, u2);
    float z = sqrt(maxf(0.0f, 1.0f - s.x * s.x - s.y * s.y));
    return make_float3(s.x, s.y, z);
}

// ---------------------------------------------------------------------------------------------------
// Disney BSDF — sv4/Disney.cuh
// ---------------------------------------------------------------------------------------------------
enum BSDFType { eReflected, eTransmitted, eSpecular };                           // :29-34

static inline bool Refract(const float3& wi, const float3& n, float eta, float3& wt) {              // :36-49
    float cosI = dot(n, wi);
    float sin2I = fmax(0.0f, float(1.0f - cosI * cosI));
    float sin2T = eta * eta * sin2I;
    if (sin2T >= 1) return false;                       // total internal reflection
    float cosT = sqrtf(1.0f - sin2T);
    wt = eta * -wi + (eta * cosI - cosT) * float3(n);
    return true;
}
static inline float SchlickFresnel(float u) {                                    // :51-56
    float m = clamp(1 - u, 0.0f, 1.0f);
    float m2 = m * m;
    return m2 * m2 * m;
}
static inline float GTR1(float NDotH, float a) {                                 // :58-64
    if (a >= 1) return kInvPi;
    float a2 = a * a;
    float t = 1 + (a2 - 1) * NDotH * NDotH;
    return (a2 - 1) / (kPi * logf(a2) * t);
}
static inline float GTR2(float NDotH, float a) {                                 // :66-71
    float a2 = a * a;
    float t = 1.0f + (a2 - 1.0f) * NDotH * NDotH;
    return a2 / (kPi * t * t);
}
static inline float SmithGGX(float NDotv, float alphaG) {                        // :73-78
    float a = alphaG * alphaG;
    float b = NDotv * NDotv;
    return 1 / (NDotv + sqrtf(a + b - a * b));
}
static inline float Fr(float VDotN, float etaI, float etaT) {                    // :81-98
    float sinT2 = sqr(etaI / etaT) * (1.0f - VDotN * VDotN);
    if (sinT2 > 1.0f) return 1.0f;                      // total internal reflection
    float LDotN = sqrt(1.0f - sinT2);
    float eta = etaT / etaI;
    float r1 = (VDotN - eta * LDotN) / (VDotN + eta * LDotN);
    float r2 = (LDotN - eta * VDotN) / (LDotN + eta * VDotN);
    return 0.5f * (sqr(r1) + sqr(r2));
}

static inline float BSDFPdf(const Material& mat, float etaI, float etaO, const float3& P, const float3& n,
                            const float3& V, const float3& L) {                  // :152-193
    (void)P;
    if (dot(L, n) <= 0.0f) {
        float bsdfPdf = 0.0f;
        float brdfPdf = kInv2Pi * mat.subsurface * 0.5f;
        return lerp(brdfPdf, bsdfPdf, mat.transmission);
    }
    float F = Fr(dot(n, V), etaI, etaO);
    const float a = fmax(0.001f, mat.roughness);
    const float3 half = SafeNormalize(L + V);
    const float cosThetaHalf = abs(dot(half, n));
    const float pdfHalf = GTR2(cosThetaHalf, a) * cosThetaHalf;
    float pdfSpec = 0.25f * pdfHalf / fmax(1.e-6f, dot(L, half));
    float pdfDiff = abs(dot(L, n)) * kInvPi * (1.0f - mat.subsurface);
    float bsdfPdf = pdfSpec * F;
    float brdfPdf = lerp(pdfDiff, pdfSpec, 0.5f);
    return lerp(brdfPdf, bsdfPdf, mat.transmission);
}

// GGX half-vector construction shared by the two specular branches (:209-228 and :281-305)
static inline float3 ggxReflect_(const Material& mat, const float3& U, const float3& V, const float3& N,
                                 const float3& view, float r1, float r2) {
    const float a = fmax(0.001f, mat.roughness);
    const float phiHalf = r1 * k2Pi;
    const float cosThetaHalf = sqrtf((1.0f - r2) / (1.0f + (sqr(a) - 1.0f) * r2));
    const float sinThetaHalf = sqrtf(fmax(0.0f, 1.0f - sqr(cosThetaHalf)));
    const float sinPhiHalf = sinf(phiHalf);
    const float cosPhiHalf = cosf(phiHalf);
    float3 half = U * (sinThetaHalf * cosPhiHalf) + V * (sinThetaHalf * sinPhiHalf) + N * cosThetaHalf;
    if (dot(half, view) <= 0.0f) half *= -1.0f;
    return 2.0f * dot(view, half) * half - view;
}

static inline void BSDFSample(const Material& mat, float etaI, float etaO, const float3& P, const float3& U,
                              const float3& V, const float3& N, const float3& view, float3& light, float& pdf,
                              BSDFType& type, Random& rand) {                    // :197-315
    if (rand.Randf() < mat.transmission) {
        float F = Fr(dot(N, view), etaI, etaO);
        if (rand.Randf() < F) {
            float r1, r2;
            Sample2D(rand, r1, r2);
            type = eReflected;
            light = ggxReflect_(mat, U, V, N, view, r1, r2);
        } else {
            float eta = etaI / etaO;
            if (Refract(view, N, eta, light)) {
                type = eSpecular;
                pdf = (1.0f - F) * mat.transmission;
                return;
            }
            pdf = 0.0f;
            return;
        }
    } else {
        float r1, r2;
        Sample2D(rand, r1, r2);
        if (rand.Randf() < 0.5f) {
            if (rand.Randf() < mat.subsurface) {
                const float3 d = UniformSampleHemisphere(rand);
                light = U * d.x + V * d.y - N * d.z;      // below the surface
                type = eTransmitted;
            } else {
                const float3 d = CosineSampleHemisphere(r1, r2);
                light = U * d.x + V * d.y + N * d.z;
                type = eReflected;
            }
        } else {
            light = ggxReflect_(mat, U, V, N, view, r1, r2);
            type = eReflected;
        }
    }
    pdf = BSDFPdf(mat, etaI, etaO, P, N, view, light);
}

static inline float3 BSDFEval(const Material& mat, float3 albedo, float etaI, float etaO, const float3& P,
                              const float3& N, const float3& V, const float3& L) {   // :318-427
    (void)P;
    float NDotL = dot(N, L);
    float NDotV = dot(N, V);
    float3 H = normalize(L + V);
    float NDotH = dot(N, H);
    float LDotH = dot(L, H);

    float3 Cdlin = albedo;
    float Cdlum = .3 * Cdlin.x + .6 * Cdlin.y + .1 * Cdlin.z;                     // double arithmetic
    float3 Ctint = Cdlum > 0.0f ? Cdlin / Cdlum : make_float3(1.0f);
    float3 Cspec0 = lerp((float)(mat.specular * .08) * lerp(make_float3(1.0f), Ctint, mat.specularTint), Cdlin, mat.metallic);

    float3 bsdf = make_float3(0.0f);
    float3 brdf = make_float3(0.0f);

    if (mat.transmission > 0.0f) {
        if (NDotL <= 0) {
            float F = Fr(NDotV, etaI, etaO);
            bsdf = make_float3(mat.transmission * (1.0f - F) / abs(NDotL) * (1.0f - mat.metallic));
        } else {
            float a = fmax(0.001f, mat.roughness);
            float Ds = GTR2(NDotH, a);
            float FH = Fr(LDotH, etaI, etaO);
            float3 Fs = lerp(Cspec0, make_float3(1.0f), FH);
            float Gs = SmithGGX(NDotV, a) * SmithGGX(NDotL, a);
            bsdf = Gs * Fs * Ds;
        }
    }
    if (mat.transmission < 1.0f) {
        if (NDotL <= 0) {
            if (mat.subsurface > 0.0f) {
                float3 s = make_float3(sqrt(mat.color.x), sqrt(mat.color.y), sqrt(mat.color.z));
                float FL = SchlickFresnel(abs(NDotL)), FV = SchlickFresnel(NDotV);
                float Fd = (1.0f - 0.5f * FL) * (1.0f - 0.5f * FV);
                brdf = kInvPi * s * mat.subsurface * Fd * (1.0f - mat.metallic);
            }
        } else {
            float a = fmax(0.001f, mat.roughness);
            float Ds = GTR2(NDotH, a);
            float FH = SchlickFresnel(LDotH);
            float3 Fs = lerp(Cspec0, make_float3(1.f), FH);
            float Gs = SmithGGX(NDotV, a) * SmithGGX(NDotL, a);
            float FL = SchlickFresnel(NDotL), FV = SchlickFresnel(NDotV);
            float Fd90 = 0.5 + 2.0f * LDotH * LDotH * mat.roughness;             // double add
            float Fd = lerp(1.0f, Fd90, FL) * lerp(1.0f, Fd90, FV);
            float Dr = GTR1(NDotH, lerp(.1f, .001f, mat.clearcoatGloss));
            float Fc = lerp(.04f, 1.0f, FH);
            float Gr = SmithGGX(NDotL, .25f) * SmithGGX(NDotV, .25f);
            brdf = kInvPi * Fd * Cdlin * (1.0f - mat.metallic) * (1.0f - mat.subsurface) + Gs * Fs * Ds +
                   make_float3(mat.clearcoat * Gr * Fc * Dr);
        }
    }
    return lerp(brdf, bsdf, mat.transmission);
}

// ---------------------------------------------------------------------------------------------------
// Environment probe — sv4/Probe.cuh
// ---------------------------------------------------------------------------------------------------
static inline float2 ProbeDirToUV(const float3& dir) {                           // :38-46
    float theta = acosf(clamp(dir.y, -1.0f, 1.0f));
    float phi = (dir.x == 0.0f && dir.z == 0.0f) ? 0.0f : atan2(dir.z, dir.x);
    float u = (kPi + phi) * kInvPi * 0.5f;
    float v = theta * kInvPi;
    return make_float2(u, v);
}
static inline float3 ProbeUVToDir(const float2& uv) {                            // :48-58
    float theta = uv.y * kPi;
    float phi = uv.x * 2.0f * kPi;
    float x = -sinf(theta) * cosf(phi);
    float y = cosf(theta);
    float z = -sinf(theta) * sinf(phi);
    return make_float3(x, y, z);
}
static inline float4 ProbeEval(const Probe& image, const float2& uv) {           // :61-67
    int px = clamp(int(uv.x * image.width), 0, image.width - 1);
    int py = clamp(int(uv.y * image.height), 0, image.height - 1);
    return image.data[py * image.width + px];
}
static inline float ProbePdf(const Probe& image, const float3& d) {              // :69-93
    float2 uv = ProbeDirToUV(d);
    int col = clamp(int(uv.x * image.width), 0, image.width - 1);
    int row = clamp(int(uv.y * image.height), 0, image.height - 1);
    float pdf = image.pdfValuesX[row * image.width + col] * image.pdfValuesY[row];
    float sinTheta = sinf(uv.y * kPi);
    if (fabsf(sinTheta) < 0.0001f) pdf = 0.0f;
    else pdf *= float(image.width) * float(image.height) / (2.0f * kPi * kPi * sinTheta);
    return pdf;
}
static inline int LowerBound(const float* array, int lower, int upper, const float value) {   // :119-136
    while (lower < upper) {
        int mid = lower + (upper - lower) / 2;
        if (array[mid] < value) lower = mid + 1;
        else upper = mid;
    }
    return lower;
}
static inline void ProbeSample(const Probe& image, float3& dir, float3& color, float& pdf, Random& rand) {   // :138-169
    float r1, r2;
    Sample2D(rand, r1, r2);
    int row = LowerBound(image.cdfValuesY, 0, image.height, r1);
    int col = LowerBound(image.cdfValuesX, row * image.width, (row + 1) * image.width, r2) - row * image.width;
    color = make_float3(image.data[row * image.width + col]);
    pdf = image.pdfValuesX[row * image.width + col] * image.pdfValuesY[row];
    float u = col / float(image.width);
    float v = row / float(image.height);
    float sinTheta = sinf(v * kPi);
    if (sinTheta == 0.0f) pdf = 0.0f;
    else pdf *= image.width * image.height / (2.0f * kPi * kPi * sinTheta);
    dir = ProbeUVToDir(make_float2(u, v));
}
