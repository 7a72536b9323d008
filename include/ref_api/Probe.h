// Environment probe types, layout-compatible with the reference.
//   struct Probe       — device-side view, 64 bytes (reference: sv4/Probe.cuh:6-21)
//   struct ProbeData   — host-side image + importance tables (reference: sv4/Probe.h:7-89)
// BuildCDF keeps the reference's *sequential fp32 accumulation order*, which defines the bits of
// the tables; it therefore stays a plain host loop.
#pragma once
#include <vector_types.h>
#include <cstddef>
#include <string>

typedef float4 Color;

struct Probe {
    int    width;        // @0
    int    height;       // @4
    Color* data;         // @8    width*height float4 texels (row-major, row = latitude)
    float3 offset;       // @16   world-space skybox offset (carried, never read by the path)
    float* pdfValuesX;   // @32   per-texel pdf inside its row
    float* cdfValuesX;   // @40   per-row inclusive cdf
    float* pdfValuesY;   // @48   per-row pdf
    float* cdfValuesY;   // @56   inclusive cdf over rows
};
static_assert(sizeof(Probe) == 64, "Probe must stay 64 bytes (reference ABI)");
static_assert(offsetof(Probe, data) == 8 && offsetof(Probe, offset) == 16 &&
              offsetof(Probe, pdfValuesX) == 32 && offsetof(Probe, cdfValuesY) == 56, "Probe offsets");

struct ProbeData {
    int    width  = 0;
    int    height = 0;
    Color* data   = nullptr;     // caller-owned
    float3 offset = {0.f, 0.f, 0.f};
    bool   valid  = false;
    float* pdfValuesX = nullptr;
    float* cdfValuesX = nullptr;
    float* pdfValuesY = nullptr;
    float* cdfValuesY = nullptr;

    // reference: sv4/Probe.h:29-77.  Luminance weights 0.3/0.6/0.1 (sv4/maths.h:165-168).
    inline void BuildCDF() {
        const size_t n = (size_t)width * (size_t)height;
        pdfValuesX = new float[n];  cdfValuesX = new float[n];
        pdfValuesY = new float[height];  cdfValuesY = new float[height];
        float runY = 0.0f;
        for (int row = 0; row < height; ++row) {
            float runX = 0.0f;
            float* px = pdfValuesX + (size_t)row * width;
            float* cx = cdfValuesX + (size_t)row * width;
            const Color* texel = data + (size_t)row * width;
            for (int col = 0; col < width; ++col) {
                const float lum = texel[col].x * 0.3f + texel[col].y * 0.6f + texel[col].z * 0.1f;
                runX += lum;
                px[col] = lum;
                cx[col] = runX;
            }
            const float inv = 1.0f / runX;
            for (int col = 0; col < width; ++col) { px[col] *= inv;  cx[col] *= inv; }
            runY += runX;
            pdfValuesY[row] = runX;
            cdfValuesY[row] = runY;
        }
        for (int row = 0; row < height; ++row) {
            cdfValuesY[row] /= float(runY);
            pdfValuesY[row] /= float(runY);
        }
        valid = true;
    }
};

// host/Probe.cpp — loadProbe / loadColor (sv4/main.cpp:154-180).  readHDR is our own Radiance .hdr reader with stbi_loadf(..., 4)
// conventions; the returned array is new[]-allocated and owned by the caller (as probe.data is in the reference).
Color* readHDR(const std::string& hdrFile, int& width, int& height);
void loadProbe(ProbeData& probe, const std::string& hdrFile);
void loadColor(ProbeData& probe, const float3& color, int width = 3840, int height = 2160);
