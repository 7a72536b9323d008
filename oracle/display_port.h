// ORACLE — TEST INFRASTRUCTURE, NOT PRODUCT CODE.
// Port backend only: CPU restatement of the display transform in cuda/helpers.h:35-62
// (toSRGB, quantizeUnsigned8Bits, make_color).  In the reference backend the reference's own
// cuda/helpers.h is host-compiled instead (oracle/ref_backend.h).
#pragma once
#include "shading_port.h"

static inline float3 toSRGB(const float3& c) {                                   // helpers.h:35-44
    float invGamma = 1.0f / 2.4f;
    float3 powed = make_float3(powf(c.x, invGamma), powf(c.y, invGamma), powf(c.z, invGamma));
    return make_float3(c.x < 0.0031308f ? 12.92f * c.x : 1.055f * powed.x - 0.055f,
                       c.y < 0.0031308f ? 12.92f * c.y : 1.055f * powed.y - 0.055f,
                       c.z < 0.0031308f ? 12.92f * c.z : 1.055f * powed.z - 0.055f);
}
static inline unsigned char quantizeUnsigned8Bits(float x) {                     // helpers.h:51-56
    x = clamp(x, 0.0f, 1.0f);
    unsigned int q = (unsigned int)(x * 256.0f);
    return (unsigned char)(q < 255u ? q : 255u);
}
static inline uchar4 make_color(const float3& c) {                               // helpers.h:58-63
    float3 srgb = toSRGB(clamp(c, 0.0f, 1.0f));
    return make_uchar4(quantizeUnsigned8Bits(srgb.x), quantizeUnsigned8Bits(srgb.y),
                       quantizeUnsigned8Bits(srgb.z), 255u);
}
#define ORACLE_BACKEND_NAME "port"
