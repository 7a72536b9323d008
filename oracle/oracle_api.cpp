// ORACLE — TEST INFRASTRUCTURE, NOT PRODUCT CODE.
// C entry points (ctypes) over the CPU oracle.  Compiled twice by oracle/Makefile:
//   oracle/liboracle_port.so        shading = oracle/shading_port.h (our CPU restatement)           kind "port"
//   oracle/_ref/liboracle_ref.so    shading = the reference's own headers (-DORACLE_REF)            kind "reference"
// Only tests/, __graft_entry__.smoke() and bench.py (cpu_baseline leg, --impl reference) may load these.
#ifdef ORACLE_REF
#include "ref_backend.h"
#else
#include "display_port.h"
#include "../include/ref_api/LaunchParams.h"
#endif
#include "raygen_core.h"
#include <omp.h>
#include <chrono>

extern "C" {

const char* oracle_backend() { return ORACLE_BACKEND_NAME; }
int oracle_max_threads() { return omp_get_max_threads(); }
void oracle_set_threads(int n) { if (n > 0) omp_set_num_threads(n); }
int oracle_sizeof(int which) {
    switch (which) { case 0: return (int)sizeof(LaunchParams); case 1: return (int)sizeof(Material);
                     case 2: return (int)sizeof(Probe); case 3: return (int)sizeof(TriangleMeshSBTData);
                     case 4: return (int)sizeof(Random); case 5: return (int)sizeof(b200pt_config); }
    return -1;
}
// default-constructed Material of the sv4 variant (sv4/Material.h:13-38), as raw bytes
void oracle_default_material(void* out104) { Material m; memcpy(out104, &m, sizeof(Material)); }

// ---------------------------------------------------------------- integer / RNG known-answer hooks
void oracle_tea4(int n, const uint32_t* v0, const uint32_t* v1, uint32_t* out) {
    for (int i = 0; i < n; ++i) out[i] = tea<4>(v0[i], v1[i]);
}
void oracle_lcg_stream(uint32_t seed, int n, uint32_t* state_out, float* rnd_out) {
    unsigned int s = seed;
    for (int i = 0; i < n; ++i) { rnd_out[i] = rnd(s);  state_out[i] = s; }
}
void oracle_random_stream(uint32_t seed, int n, int as_float, uint32_t* u_out, float* f_out, uint32_t* state_out2) {
    Random r(seed);
    for (int i = 0; i < n; ++i) {
        if (as_float) f_out[i] = r.Randf(); else u_out[i] = r.Rand();
    }
    state_out2[0] = r.seed1;  state_out2[1] = r.seed2;
}

// ---------------------------------------------------------------- shading known-answer hooks
void oracle_basis(int n, const float* N, float* U, float* V) {
    for (int i = 0; i < n; ++i) {
        float3 u, v;
        BasisFromVector(make_float3(N[3 * i], N[3 * i + 1], N[3 * i + 2]), &u, &v);
        U[3 * i] = u.x; U[3 * i + 1] = u.y; U[3 * i + 2] = u.z;  V[3 * i] = v.x; V[3 * i + 1] = v.y; V[3 * i + 2] = v.z;
    }
}
static inline float3 ld3(const float* p, int i) { return make_float3(p[3 * i], p[3 * i + 1], p[3 * i + 2]); }
static inline void st3(float* p, int i, const float3& v) { p[3 * i] = v.x; p[3 * i + 1] = v.y; p[3 * i + 2] = v.z; }

void oracle_bsdf_sample(int n, const void* materials104, const float* etaI, const float* etaO, const float* N,
                        const float* view, const uint32_t* seeds, float* L, float* pdf, int32_t* type, uint32_t* rng_after) {
    const Material* mats = (const Material*)materials104;
    for (int i = 0; i < n; ++i) {
        Random r(seeds[i]);
        float3 n3 = ld3(N, i), u, v, l = make_float3(0.f);
        BasisFromVector(n3, &u, &v);
        float p = 0.f;  BSDFType ty = eReflected;
        BSDFSample(mats[i], etaI[i], etaO[i], make_float3(0.f), u, v, n3, ld3(view, i), l, p, ty, r);
        st3(L, i, l);  pdf[i] = p;  type[i] = (int)ty;  rng_after[2 * i] = r.seed1;  rng_after[2 * i + 1] = r.seed2;
    }
}
void oracle_bsdf_eval(int n, const void* materials104, const float* albedo, const float* etaI, const float* etaO,
                      const float* N, const float* V, const float* L, float* f_out, float* pdf_out) {
    const Material* mats = (const Material*)materials104;
    for (int i = 0; i < n; ++i) {
        float3 f = BSDFEval(mats[i], ld3(albedo, i), etaI[i], etaO[i], make_float3(0.f), ld3(N, i), ld3(V, i), ld3(L, i));
        st3(f_out, i, f);
        pdf_out[i] = BSDFPdf(mats[i], etaI[i], etaO[i], make_float3(0.f), ld3(N, i), ld3(V, i), ld3(L, i));
    }
}

static Probe mkProbe(int w, int h, const float* data, const float* pdfX, const float* cdfX, const float* pdfY, const float* cdfY) {
    Probe p;
    p.width = w;  p.height = h;  p.data = (Color*)data;  p.offset = make_float3(0.f, 0.f, 0.f);
    p.pdfValuesX = (float*)pdfX;  p.cdfValuesX = (float*)cdfX;  p.pdfValuesY = (float*)pdfY;  p.cdfValuesY = (float*)cdfY;
    return p;
}
void oracle_build_cdf(int w, int h, const float* data_rgba, float* pdfX, float* cdfX, float* pdfY, float* cdfY) {
    ProbeData pd;                                      // sv4/Probe.h:29-77
    pd.width = w;  pd.height = h;  pd.data = (Color*)data_rgba;
    pd.BuildCDF();
    memcpy(pdfX, pd.pdfValuesX, sizeof(float) * (size_t)w * h);  memcpy(cdfX, pd.cdfValuesX, sizeof(float) * (size_t)w * h);
    memcpy(pdfY, pd.pdfValuesY, sizeof(float) * h);  memcpy(cdfY, pd.cdfValuesY, sizeof(float) * h);
    delete[] pd.pdfValuesX;  delete[] pd.cdfValuesX;  delete[] pd.pdfValuesY;  delete[] pd.cdfValuesY;
}
void oracle_probe_sample(int w, int h, const float* data, const float* pdfX, const float* cdfX, const float* pdfY,
                         const float* cdfY, int n, const uint32_t* seeds, float* dir, float* color, float* pdf) {
    Probe p = mkProbe(w, h, data, pdfX, cdfX, pdfY, cdfY);
    for (int i = 0; i < n; ++i) {
        Random r(seeds[i]);
        float3 d, c;  float q;
        ProbeSample(p, d, c, q, r);
        st3(dir, i, d);  st3(color, i, c);  pdf[i] = q;
    }
}
void oracle_probe_eval_dir(int w, int h, const float* data, const float* pdfX, const float* cdfX, const float* pdfY,
                           const float* cdfY, int n, const float* dir, float* uv, float* color, float* pdf) {
    Probe p = mkProbe(w, h, data, pdfX, cdfX, pdfY, cdfY);
    for (int i = 0; i < n; ++i) {
        float2 q = ProbeDirToUV(ld3(dir, i));
        uv[2 * i] = q.x;  uv[2 * i + 1] = q.y;
        float4 c = ProbeEval(p, q);
        st3(color, i, make_float3(c.x, c.y, c.z));
        pdf[i] = ProbePdf(p, ld3(dir, i));
    }
}

// display transforms: mode 0 = make_color(c) (orig), 1 = make_color(reinhard(c*exposure, 1)) (sv3/sv4)
void oracle_display(int n, const float* rgb, int mode, float exposure, uint32_t* out) {
    for (int i = 0; i < n; ++i) {
        float3 c = ld3(rgb, i);
        uchar4 px = mode == B200PT_DISPLAY_REINHARD ? make_color(reinhardToneMap(c * exposure, 1.0f)) : make_color(c);
        out[i] = (uint32_t)px.x | ((uint32_t)px.y << 8) | ((uint32_t)px.z << 16) | ((uint32_t)px.w << 24);
    }
}
// computeFinalPixelColorsKernel, sv4/toneMap.cu:41-58
void oracle_tonemap_sqrt(int n, const float* rgba, uint32_t* out) {
    for (int i = 0; i < n; ++i) {
        float f[4];
        for (int c = 0; c < 4; ++c) f[c] = fminf(1.f, fmaxf(0.f, sqrtf(rgba[4 * i + c])));
        uint32_t px = 0;
        px |= (uint32_t)(f[0] * 255.9f) << 0;
        px |= (uint32_t)(f[1] * 255.9f) << 8;
        px |= (uint32_t)(f[2] * 255.9f) << 16;
        px |= (uint32_t)255 << 24;
        out[i] = px;
    }
}
// sutil::Camera::UVWFrame, sutil/Camera.cpp:34-45
void oracle_camera_uvw(const float* eye, const float* lookat, const float* up, float fovY, float aspect,
                       float* U, float* V, float* W) {
    float3 w = ld3(lookat, 0) - ld3(eye, 0);
    float wlen = length(w);
    float3 u = normalize(cross(w, ld3(up, 0)));
    float3 v = normalize(cross(u, w));
    float vlen = wlen * tanf(0.5f * fovY * 3.14159265358979323846f / 180.0f);
    v *= vlen;
    float ulen = vlen * aspect;
    u *= ulen;
    st3(U, 0, u);  st3(V, 0, v);  st3(W, 0, w);
}

}  // extern "C"

#ifdef ORACLE_REF
// The reference's HDR reader: loadProbe calls stbi_loadf(file, &w, &h, &channels, 4) (sv4/main.cpp:154-165) from the stb_image.h
// vendored in the reference tree (support/stb/), compiled here in place.  Used only by tests/golden/make_golden.py to mint the
// fixture that pins the product's own Radiance-HDR readers (host/Probe.cpp, optixpathtracer_b200/probe.py).
#define STB_IMAGE_IMPLEMENTATION
#define STBI_ONLY_HDR
#include <support/stb/stb_image.h>
extern "C" int oracle_ref_load_hdr(const char* path, int* w, int* h, float* out_rgba, size_t capacity_floats) {
    int ch = 0;
    float* px = stbi_loadf(path, w, h, &ch, 4);
    if (!px) return -1;
    const size_t n = (size_t)*w * *h * 4;
    if (out_rgba && capacity_floats >= n) memcpy(out_rgba, px, n * sizeof(float));
    stbi_image_free(px);
    return 0;
}
#endif

#include "oracle_scene_api.inl"
