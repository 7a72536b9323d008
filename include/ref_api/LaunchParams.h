// Launch ABI, bit-compatible with the reference's sv4 `LaunchParams.h`
// (reference: HelloPathtracing_sv4_vmv23/LaunchParams.h:24-106; sizes/offsets in SURVEY.md §8b).
// `traversable` stays an opaque 64-bit handle (here: the b200pt context's BVH handle) so this header
// does not need optix.h.
#pragma once
#include <vector_types.h>
#include <cstddef>
#include <cstdint>
#include "Probe.h"
#include "Material.h"

enum { RAY_TYPE_RADIANCE = 0, RAY_TYPE_OCCLUSION = 1, RAY_TYPE_COUNT = 2 };

typedef unsigned long long OptixTraversableHandle;   // opaque in the reference too

struct ParallelogramLight {          // carried for layout; the path never reads it (LaunchParams.h:30-36)
    float3 corner, v1, v2, normal, emission;
};

struct TriangleMeshSBTData {         // LaunchParams.h:38-47 — one per mesh
    float3* vertex;
    float3* normal;
    float2* texcoord;
    uint3*  index;
    bool    hasTexture;
    unsigned long long texture;      // cudaTextureObject_t
    Material material;
};

struct LaunchParams {
    struct {
        float4* accum_buffer;        // @0
        uchar4* frame_buffer;        // @8
        float4* color_buffer;        // @16
        float4* normal_buffer;       // @24
        float4* albedo_buffer;       // @32
        int2    size;                // @40
        unsigned int subframe_index; // @48
        // eccentricity-ring sample budget (sv4 foveation)
        uint3   factor;              // @52
        int     fillSize;            // @64
        uint2   c;                   // @72  gaze point in pixels
        float   r_inner, r_outer;    // @80, @84
        uint2   offset;              // @88
        unsigned int redraw;         // @96
    } frame;
    struct { float3 eye, U, V, W; } camera;   // @104
    unsigned int samples_per_launch;          // @152
    OptixTraversableHandle traversable;       // @160
    ParallelogramLight light;                 // @168
    Probe probe;                              // @232
    int2  viewportSize;                       // @296
    float white;                              // @304
};

static_assert(sizeof(ParallelogramLight) == 60, "ParallelogramLight");
static_assert(sizeof(TriangleMeshSBTData) == 152, "TriangleMeshSBTData must stay 152 bytes");
static_assert(offsetof(TriangleMeshSBTData, hasTexture) == 32 && offsetof(TriangleMeshSBTData, texture) == 40 &&
              offsetof(TriangleMeshSBTData, material) == 48, "TriangleMeshSBTData offsets");
static_assert(sizeof(LaunchParams) == 312, "LaunchParams must stay 312 bytes");
static_assert(offsetof(LaunchParams, frame.size) == 40 && offsetof(LaunchParams, frame.subframe_index) == 48 &&
              offsetof(LaunchParams, frame.factor) == 52 && offsetof(LaunchParams, frame.fillSize) == 64 &&
              offsetof(LaunchParams, frame.c) == 72 && offsetof(LaunchParams, frame.r_inner) == 80 &&
              offsetof(LaunchParams, frame.offset) == 88 && offsetof(LaunchParams, frame.redraw) == 96,
              "LaunchParams.frame offsets");
static_assert(offsetof(LaunchParams, camera) == 104 && offsetof(LaunchParams, samples_per_launch) == 152 &&
              offsetof(LaunchParams, traversable) == 160 && offsetof(LaunchParams, light) == 168 &&
              offsetof(LaunchParams, probe) == 232 && offsetof(LaunchParams, viewportSize) == 296 &&
              offsetof(LaunchParams, white) == 304, "LaunchParams offsets");
