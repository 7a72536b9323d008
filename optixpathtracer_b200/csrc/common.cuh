// b200pt — shared declarations of the CUDA implementation (sm_100a).
// Data layout in HBM:
//   BvhNode   32 B  binary record: two quantised child boxes + two links (build, refit, parity hooks)
//   BvhNode4  64 B  the 4-wide node traceKernel walks (two 256-bit loads per step)
//   TriRecord 48 B  {v0.xyz,e1.x | e1.yz,e2.xy | e2.z, gid, mesh, -}  in BVH (Morton) order — the "Woop-sized" record;
//                   untransformed edges so the device test is bit-identical to the oracle's and so the shading
//                   normal cross(e1,e2) (sv4/deviceProgram.cu:598-605) needs no second gather
//   queues    32 B ray {o.xyz, slot | d.xyz, -}, 48 B hit {o, slot | d, t | u, v, triSlot}, 48 B shadow ray {P, slot | wi, target | contribution}
//   path state: one 32-byte PathState per sample slot + three float4 accumulators.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <string>
#include <vector>
#include <stdexcept>
#include "../../include/b200pt.h"
#include "../../include/ref_api/LaunchParams.h"

// SM count of the current device (cudaDevAttrMultiProcessorCount, cached per device; 148 on B200): sizes every persistent grid
#define B200PT_SMEM_CARVEOUT 28       // percent of the SM's shared memory asked for by traceKernel / shadeKernel (-> the 64 KB configuration; trace.cu launchTrace)
int numSMs();

// One 32-byte record per INNER node, holding both children: their boxes quantised (conservatively, 2 cells of padding) to a
// 16-bit grid over the scene AABB, and their links.  link >= 0: record index of the child's own node; < 0: leaf code.
//   q[0] = c0.lo.x | c0.hi.x << 16,  q[1] = c0.lo.y | c0.hi.y << 16,  q[2] = c0.lo.z | c0.hi.z << 16,  q[3..5]: child 1
struct BvhNode { unsigned q[6]; int link[2]; };
static_assert(sizeof(BvhNode) == 32, "BvhNode must be 32 bytes");
// The tree the traversal kernels read: the binary tree above collapsed to 4 children per node (64 bytes = two 256-bit loads).
// Child c: q[3c..3c+2] in the BvhNode word layout, link[c].  Unused slots: inverted box (lo 65535, hi 0: never hit), link EMPTY.
// B200PT_WIDE = children per node of that tree: 4 (default, 64 B = two 256-bit loads) or 8 (tuning builds, 128 B = four: measured slower, profiles/README.md r02f).
#ifndef B200PT_WIDE
#define B200PT_WIDE 4
#endif
struct BvhNode4 { unsigned q[3 * B200PT_WIDE]; int link[B200PT_WIDE]; };
static_assert(sizeof(BvhNode4) == 16 * B200PT_WIDE, "wide node = 16 bytes per child");
#define B200PT_LINK_EMPTY 0x7fffffff
#define B200PT_GRID_ORIGIN 2.0      // grid coordinate of the scene AABB's low corner ...
#define B200PT_GRID_SPAN 65531.0    // ... and the span of the scene AABB in cells: padded boxes stay inside [0, 65535]
#define B200PT_GRID_PAD 2
// leaf code: -(1 + first*8 + (count-1)), count in 1..8
__host__ __device__ inline int leafEncode(unsigned first, unsigned count) { return -(int)(1u + first * 8u + (count - 1u)); }
__host__ __device__ inline void leafDecode(int link, unsigned& first, unsigned& count) {
    unsigned c = (unsigned)(-(link + 1));
    first = c >> 3;  count = (c & 7u) + 1u;
}
#ifndef B200PT_MAX_LEAF
#define B200PT_MAX_LEAF 8
#endif
#define B200PT_STACK 96

struct MeshInfo {                  // device-side stand-in for the per-mesh SBT record (sv4/LaunchParams.h:38-47)
    Material material;             // 104 B, reference layout
    unsigned firstTri;             // global id of the mesh's triangle 0
    unsigned firstVertex;          // offset into the global vertex / texcoord arrays
    int      textureID;            // -1 = none
    int      hasTexcoord;
    unsigned texOffset;            // first texel of the texture in the RGBA8 pool
    int      texW, texH;
    int      hasNormal;            // per-vertex normals uploaded (read by the Whitted integrator only)
};
static_assert(sizeof(MeshInfo) == 136, "MeshInfo");

struct SceneDev {                  // everything the kernels need about the scene (passed by value)
    const uint4* nodes;            // BvhNode records viewed as uint4 x2 (binary tree: parity hooks, refit)
    const uint4* nodes4;           // BvhNode4 records viewed as uint4 x4 (what traceKernel walks)
    int rootLink4;
    int rootLink;                  // link of the root (EMPTY sentinel 0x7fffffff for an empty scene)
    double gridLo[3], gridScale[3];   // world -> grid: g = GRID_ORIGIN + (x - gridLo) * gridScale
    const float4* tris;            // TriRecord viewed as float4 x3, BVH order
    const MeshInfo* meshes;
    const uint3* indices;          // global index buffer, LOCAL vertex ids (per mesh)
    const float2* texcoords;       // global, per vertex (zeros where a mesh has none)
    const float3* vertices;        // global, per vertex (Whitted integrator: surfPos, :121-124)
    const float3* normals;         // global, per vertex (zeros where a mesh has none)
    const uint32_t* texels;        // RGBA8 pool
    unsigned nTris, nNodes;
    unsigned nNodes4, nMeshes, nTexels, nVertices;   // array sizes (B200PT_CHECKED builds bounds-check against them)
};

struct PassParams {                // one optixLaunch of the reference = one pass of the wavefront
    int launchW, launchH;
    unsigned factorX, factorY;
    int fillSize;
    unsigned cX, cY;
    float rInner, rOuter;
    unsigned offX, offY;
    unsigned redraw;
    unsigned subframe;
    int spl;                       // samples_per_launch of this pass
    unsigned slotBase;             // first sample slot of this pass in the pool
    unsigned launchBase;           // first launch-index record of this pass
};

struct FrameParams {               // shared by every pass of a frame
    int W, H;
    float3 eye, U, V, Wv;
    Probe probe;                   // device pointers
    const int* probeGuide;         // guide tables of the probe's CDFs (shade_math.cuh; nullptr: plain binary search)
    float tmin, tmax;
    int maxDepth, shadowMode, display, accumMode, writeAovs, russianRoulette, misOnMiss;
    float exposure;
    int tileRank, tileCount;
    float4* accum;  uint32_t* frame;  float4* normal;  float4* albedo;  float4* color;
};

// Per-path state, one 32-byte record (= one DRAM sector) per sample slot: shadeKernel reads and rewrites it as a whole.
struct PathState {
    unsigned rng1, rng2;           // Random {seed1, seed2}
    float    thrX, thrY, thrZ;     // pathThroughput
    float    eta;                  // rayEta
    int      state;                // depth | flags << 8
    float    bsdfPdf;              // only with mis_on_miss: pdf of the BSDF sample that produced the current segment
};
static_assert(sizeof(PathState) == 32, "PathState must be one 32-byte sector");

struct PathPool {                  // path state indexed by sample slot
    unsigned cap;                  // slots allocated (= capacity of every ray queue)
    PathState* st;
    float4* direct;                // radiance added at depth 0
    float4* indirect;              // radiance added at depth >= 1, in bounce order
    float4* alpha;                 // prd.alpha
    float4* aovNormal;             // only with write_aovs
    float4* aovAlbedo;
};

// Wavefront queues of one frame in flight.
//   extension rays  rayO {o.xyz, slot} rayD {d.xyz, -}      double-buffered, written by raygen / shade, read by trace and shade
//   hits            hit {t, u, v, triSlot}                  written by traceKernel at the ray's own index (triSlot < 0: the ray left the scene)
//   shadow rays     shO {P.xyz, slot} shD {wi.xyz, target} shP {T * lightSample, -}: the NEE contribution travels with its occlusion ray
// The order of the extension-ray queue is inherited from bounce to bounce (a ray's successor is appended in the order of its parent), so the
// 8x4-pixel tiles of raygen keep neighbouring queue entries close in space at every depth; shadeKernel compacts the hits of a queue segment
// into dense warps WITHOUT reordering them (a ray that leaves the scene needs no shading: __miss__radiance only sets DONE, sv4/deviceProgram.cu:251-280).
struct Queues {
    float4* rayO[2];  float4* rayD[2];
    float4* hit;
    float4* shO;  float4* shD;  float4* shP;
    unsigned* counters;                      // per trace launch d: counters[QC_STRIDE*d + QC_*]
    unsigned cap;                            // entries allocated per queue
};
// trace launch d of a frame handles the extension rays of bounce d and the shadow rays emitted by shade pass d-1; QC_HEAD / QC_SHEAD are the
// work tickets of trace launch d / shade pass d
enum { QC_RAYS = 0, QC_SHADOWS = 1, QC_HEAD = 2, QC_SHEAD = 3, QC_STRIDE = 4, QC_MAX_LAUNCHES = 208 };

// ---- B200PT_CHECKED builds (tools/build_variant.sh checked -DB200PT_CHECKED): compute-sanitizer is closed on the GPU pool, so the kernels carry
// their own bounds checks: every index into a queue, the path pool, the traversal stack, the node / triangle / mesh / texel arrays goes through
// CHK(condition, code); a violation is counted (per .cu file, first code kept), the access is skipped or redirected to element 0, and
// b200pt_destroy prints the totals.  In normal builds CHK(c, k) is the constant `true` and costs nothing.
#ifdef B200PT_CHECKED
#define CHK_DECL static __device__ unsigned g_chkCount, g_chkFirst; \
    static __device__ __noinline__ void chkFail(unsigned code) { if (atomicAdd(&g_chkCount, 1u) == 0u) g_chkFirst = code; }
#define CHK(cond, code) ((cond) ? true : (chkFail(code), false))
#else
#define CHK_DECL
#define CHK(cond, code) true
#endif
void checkedReportTrace(unsigned* count, unsigned* first);   // trace.cu / shade.cu: this file's violation count (0, 0 in normal builds)
void checkedReportShade(unsigned* count, unsigned* first);

struct TraceStats { unsigned long long nodes, tris, rays, hits; };      // hits: closest-hit rays that found a surface ([0]) / occluded shadow rays ([1])

#define CUDA_OK(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) \
    throw std::runtime_error(std::string(#call) + ": " + cudaGetErrorString(e_)); } while (0)

// path flags (sv4/deviceProgram.cu:54-56 + internal)
#define PF_DONE       1
#define PF_SECONDARY  2
#define PF_ALPHA_ONE  4

// ---- launchers implemented in the .cu files ------------------------------------------------------------
// bvh_build.cu
struct BvhBuildResult {
    BvhNode* nodes = nullptr;  int rootLink = 0x7fffffff;  BvhNode4* nodes4 = nullptr;  int rootLink4 = 0x7fffffff;  unsigned nNodes4 = 0, maxDepth4 = 0;  double gridLo[3] = {0, 0, 0}, gridScale[3] = {1, 1, 1};
    float4* tris = nullptr;  unsigned nNodes = 0, nTris = 0, nLeaves = 0, maxLeaf = 0, maxDepth = 0;
    float sahCost = 0.f, sahCostLbvh = 0.f, buildMs = 0.f;
    unsigned long long* mortonByGid = nullptr;  uint32_t* sortedGid = nullptr;  uint32_t* finalGid = nullptr;  unsigned refineLevels = 0;  float sceneBounds[6] = {0, 0, 0, 0, 0, 0};
};
void bvhBuild(const float3* d_vertices, const uint3* d_indices, const unsigned* d_triMesh, const MeshInfo* d_meshes,
              unsigned nTris, cudaStream_t stream, BvhBuildResult& out);
void bvhFree(BvhBuildResult& r);
// new vertex positions, same topology: triangle records + bottom-up bounds + re-quantisation (bvh_build.cu)
void bvhRefit(const float3* d_vertices, const uint3* d_indices, const unsigned* d_triMesh, const MeshInfo* d_meshes, cudaStream_t stream,
              BvhBuildResult& io, float* refitMs);

// trace.cu  (compiled with FMA contraction; the triangle test itself uses explicit _rn intrinsics)
// ctr = the QC_* counter block of this launch; cur = which extension-ray buffer it reads; closest = 0: shadow rays only (the frame's last launch)
void launchTrace(const SceneDev& scene, const Queues& Q, int cur, int closest, unsigned* ctr, float tmin, float tminShadow, float tmax, int cullBack,
                 const PathPool& pool, TraceStats* stats /* [0] closest, [1] any; nullptr = no counting */, cudaStream_t stream);
void launchTraceRaysDebug(const SceneDev& scene, const float4* rayO, const float4* rayD, unsigned n, float tmin, float tmax,
                          int anyHit, int cullBack, float4* hitOut, cudaStream_t stream);

// shade.cu  (compiled with -fmad=false: basic fp32 arithmetic bit-identical to the IEEE host oracle)
void launchRaygen(const FrameParams& F, const PassParams& P, int k0, int k1, const PathPool& pool, const Queues& Q,
                  cudaStream_t stream);
void launchShade(const FrameParams& F, const SceneDev& scene, const PathPool& pool, const Queues& Q, int cur, int launch,
                 cudaStream_t stream);
void launchGather(const FrameParams& F, const PassParams& P, int k0, int k1, const PathPool& pool, float4* sumRad,
                  float4* sumAlpha, float4* sumNormal, float4* sumAlbedo, cudaStream_t stream);
// the pools a fused finalize pass sums the samples from: one per band set of the frame (n = 1: the whole frame; n = 0: not fused, read gatherKernel's records).
// Band b of a pass belongs to set ((b % (tileCount * n)) - tileRank) / tileCount - the partition enqueuePart gives the sets (api_render.inl).
struct FusedPools { PathPool p[4];  int n; };
void launchFinalize(const FrameParams& F, const PassParams& P, const float4* sumRad, const float4* sumAlpha,
                    const float4* sumNormal, const float4* sumAlbedo, const FusedPools& fused, cudaStream_t stream);
void launchResolveSum(const FrameParams& F, int nPasses, cudaStream_t stream);
void launchResolveTotal(const FrameParams& F, const float4* total, int nPasses, cudaStream_t stream);
struct WhittedParams { int W, H;  float3 position, direction, horizontal, vertical, lightPos; };
void launchWhittedRaygen(const WhittedParams& P, const Queues& Q, const PathPool& pool, unsigned* counters, cudaStream_t stream);
void launchWhittedShade(const WhittedParams& P, const SceneDev& scene, const PathPool& pool, const Queues& Q, unsigned* shadowCounter, cudaStream_t stream);
void launchWhittedFinish(const WhittedParams& P, const PathPool& pool, uint32_t* colorBuffer, cudaStream_t stream);
void launchDebugBsdf(unsigned n, const Material* mats, const float* etaI, const float* etaO, const float3* N, const float3* view, const float3* L,
                     const float3* albedo, const unsigned* seeds, float3* basisU, float3* basisV, float3* sampleL, float* samplePdf, int* sampleType,
                     uint2* sampleRng, float3* evalF, float* evalPdf, cudaStream_t stream);
void launchDebugProbe(const Probe& P, const int* guide, unsigned n, const unsigned* seeds, float3* dir, float3* color, float* pdf, unsigned m,
                      const float3* dirs, float2* uv, float3* evalColor, float* evalPdf, cudaStream_t stream);
size_t probeGuideSize(int H);
void launchProbeGuide(const float* cdfX, const float* cdfY, int W, int H, int* guide, cudaStream_t stream);
void launchProbeCdf(const float4* data, int W, int H, float* pdfX, float* cdfX, float* pdfY, float* cdfY, float* rowTotal, cudaStream_t stream);
void launchProbeCdfNoAlloc(const float4* data, int W, int H, float* pdfX, float* cdfX, float* pdfY, float* cdfY, cudaStream_t stream);   // scratch: static __device__ array in shade.cu
void launchTonemapSqrt(int w, int h, const float4* in, uint32_t* out, cudaStream_t stream);
void launchPrimaryRays(const FrameParams& F, const PassParams& P, float4* rayO, float4* rayD, cudaStream_t stream);
void launchResolvePrimary(const SceneDev& scene, const float4* hit, unsigned n, const float4* rayO, int32_t* prim, int32_t* mesh,
                          float* t, float* u, float* v, cudaStream_t stream);
