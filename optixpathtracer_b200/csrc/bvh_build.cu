// b200pt — LBVH builder (replaces optixAccelBuild + optixAccelCompact, sv4/SimplePathtracer.cpp:670-737).
// Stages, all on the GPU:
//   1. triBoundsKernel   per-triangle AABB + scene AABB                (HBM stream: 12 B index + 36 B verts in, 32 B out)
//   2. mortonKernel      63-bit Morton code of the AABB centre          (bit-exact vs oracle_morton)
//   3. radix sort        LSD, 8-bit digits, 64-bit key + 32-bit id, stable (== std::stable_sort order)
//   4. karrasKernel      Karras 2012 binary radix tree over the sorted codes (index tie-break for duplicates)
//   5. refitKernel       bottom-up AABBs + SAH cost of the plain LBVH (atomic arrival flags); reported, not traversed
//   6. bvh_sah.cuh       binned-SAH treelet refinement, level by level, over the LBVH's ranges -> one 32-byte record per
//                        inner node holding both child boxes, quantised to a 16-bit grid (the tree the traversal kernels read)
//   7. emitTrisKernel    48-byte triangle records in final BVH order
// Integer stages (2,3) are the ones north_star asks to be bit-exact against a host reference.
#include "common.cuh"
#include <cfloat>
#include <chrono>

#define SORT_THREADS 256
#define SORT_ITEMS 16
#define SORT_TILE (SORT_THREADS * SORT_ITEMS)
#define SAH_CI 1.2f
#define SAH_CT 1.0f
#include "bvh_sah.cuh"

// ---------------------------------------------------------------- float <-> ordered uint for atomic min/max
__device__ __forceinline__ unsigned fOrder(float f) { unsigned u = __float_as_uint(f); return (u & 0x80000000u) ? ~u : (u | 0x80000000u); }
__host__ __device__ inline float fUnorder(unsigned u) {
    u = (u & 0x80000000u) ? (u & 0x7fffffffu) : ~u;
#ifdef __CUDA_ARCH__
    return __uint_as_float(u);
#else
    float f;  memcpy(&f, &u, 4);  return f;
#endif
}

__global__ void triBoundsKernel(const float3* __restrict__ vtx, const uint3* __restrict__ idx, const unsigned* __restrict__ triMesh,
                                const MeshInfo* __restrict__ meshes, unsigned n, float4* __restrict__ triLo, float4* __restrict__ triHi,
                                unsigned* sceneBox /* 6 ordered uints */) {
    const unsigned g = blockIdx.x * blockDim.x + threadIdx.x;
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    if (g < n) {
        const unsigned fv = meshes[triMesh[g]].firstVertex;
        const uint3 id = idx[g];
        const float3 a = vtx[fv + id.x], b = vtx[fv + id.y], c = vtx[fv + id.z];
        lo[0] = fminf(fminf(a.x, b.x), c.x);  hi[0] = fmaxf(fmaxf(a.x, b.x), c.x);
        lo[1] = fminf(fminf(a.y, b.y), c.y);  hi[1] = fmaxf(fmaxf(a.y, b.y), c.y);
        lo[2] = fminf(fminf(a.z, b.z), c.z);  hi[2] = fmaxf(fmaxf(a.z, b.z), c.z);
        triLo[g] = make_float4(lo[0], lo[1], lo[2], 0.f);
        triHi[g] = make_float4(hi[0], hi[1], hi[2], 0.f);
    }
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        float l = lo[a], h = hi[a];
        for (int off = 16; off; off >>= 1) { l = fminf(l, __shfl_xor_sync(0xffffffffu, l, off));  h = fmaxf(h, __shfl_xor_sync(0xffffffffu, h, off)); }
        if ((threadIdx.x & 31) == 0) { atomicMin(sceneBox + a, fOrder(l));  atomicMax(sceneBox + 3 + a, fOrder(h)); }
    }
}

__device__ __forceinline__ unsigned long long expand21(unsigned v) {
    unsigned long long x = v & 0x1fffffull;
    x = (x | x << 32) & 0x1f00000000ffffull;
    x = (x | x << 16) & 0x1f0000ff0000ffull;
    x = (x | x << 8) & 0x100f00f00f00f00full;
    x = (x | x << 4) & 0x10c30c30c30c30c3ull;
    x = (x | x << 2) & 0x1249249249249249ull;
    return x;
}

__global__ void mortonKernel(const float4* __restrict__ triLo, const float4* __restrict__ triHi, unsigned n,
                             const unsigned* __restrict__ sceneBox, unsigned long long* __restrict__ codes, unsigned* __restrict__ ids) {
    const unsigned g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= n) return;
    const float4 l = triLo[g], h = triHi[g];
    const float tl[3] = {l.x, l.y, l.z}, th[3] = {h.x, h.y, h.z};
    unsigned cell[3];
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const float slo = fUnorder(sceneBox[a]), shi = fUnorder(sceneBox[3 + a]);
        const float e = __fsub_rn(shi, slo);
        const float scale = e > 0.f ? __fdiv_rn(1.0f, e) : 0.f;
        const float c = __fmul_rn(0.5f, __fadd_rn(tl[a], th[a]));
        const float q = __fmul_rn(__fmul_rn(__fsub_rn(c, slo), scale), 2097152.0f);
        const int qi = (int)q;
        cell[a] = (unsigned)min(max(qi, 0), 2097151);
    }
    codes[g] = (expand21(cell[0]) << 2) | (expand21(cell[1]) << 1) | expand21(cell[2]);
    ids[g] = g;
}

// ---------------------------------------------------------------- radix sort (one 8-bit digit per pass)
__global__ void __launch_bounds__(SORT_THREADS)
sortHistKernel(const unsigned long long* __restrict__ keys, unsigned n, int shift, unsigned* __restrict__ blockHist, unsigned numBlocks) {
    __shared__ unsigned hist[256];
    hist[threadIdx.x] = 0;
    __syncthreads();
    const unsigned base = blockIdx.x * SORT_TILE;
    for (int r = 0; r < SORT_ITEMS; ++r) {
        const unsigned i = base + r * SORT_THREADS + threadIdx.x;
        if (i < n) atomicAdd(&hist[(unsigned)(keys[i] >> shift) & 255u], 1u);
    }
    __syncthreads();
    blockHist[threadIdx.x * numBlocks + blockIdx.x] = hist[threadIdx.x];      // digit-major
}

// exclusive scan of `n` uints by a single CTA (small arrays: histograms, survivor flags)
__global__ void __launch_bounds__(1024) scanExclusiveKernel(const unsigned* __restrict__ in, unsigned* __restrict__ out, unsigned n, unsigned* total) {
    __shared__ unsigned part[1024];
    const unsigned chunk = (n + 1023) / 1024;
    const unsigned b = threadIdx.x * chunk, e = min(n, b + chunk);
    unsigned s = 0;
    for (unsigned i = b; i < e; ++i) s += in[i];
    part[threadIdx.x] = s;
    __syncthreads();
    for (unsigned off = 1; off < 1024; off <<= 1) {
        unsigned v = threadIdx.x >= off ? part[threadIdx.x - off] : 0;
        __syncthreads();
        part[threadIdx.x] += v;
        __syncthreads();
    }
    unsigned run = threadIdx.x ? part[threadIdx.x - 1] : 0;
    for (unsigned i = b; i < e; ++i) { const unsigned v = in[i];  out[i] = run;  run += v; }
    if (threadIdx.x == 1023 && total) *total = part[1023];
}

__global__ void __launch_bounds__(SORT_THREADS)
sortScatterKernel(const unsigned long long* __restrict__ keysIn, const unsigned* __restrict__ valsIn,
                  unsigned long long* __restrict__ keysOut, unsigned* __restrict__ valsOut, unsigned n, int shift,
                  const unsigned* __restrict__ blockOffset, unsigned numBlocks) {
    __shared__ unsigned digitBase[256];
    __shared__ unsigned roundBase[256];
    __shared__ unsigned warpCount[SORT_THREADS / 32][256];
    const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    digitBase[threadIdx.x] = blockOffset[threadIdx.x * numBlocks + blockIdx.x];
    const unsigned base = blockIdx.x * SORT_TILE;
    for (int r = 0; r < SORT_ITEMS; ++r) {
        for (int w = 0; w < SORT_THREADS / 32; ++w) warpCount[w][threadIdx.x] = 0;
        __syncthreads();
        const unsigned i = base + r * SORT_THREADS + threadIdx.x;
        const bool valid = i < n;
        unsigned long long key = 0;  unsigned val = 0, d = 0xffffffffu;
        if (valid) { key = keysIn[i];  val = valsIn[i];  d = (unsigned)(key >> shift) & 255u; }
        const unsigned peers = __match_any_sync(0xffffffffu, d);
        const unsigned rankInWarp = __popc(peers & ((1u << lane) - 1u));
        if (valid && rankInWarp == 0) warpCount[warp][d] = __popc(peers);
        __syncthreads();
        {   // thread `threadIdx.x` owns digit threadIdx.x: exclusive prefix over the warps (stable order)
            unsigned run = 0;
            for (int w = 0; w < SORT_THREADS / 32; ++w) { const unsigned c = warpCount[w][threadIdx.x];  warpCount[w][threadIdx.x] = run;  run += c; }
            roundBase[threadIdx.x] = digitBase[threadIdx.x];
            digitBase[threadIdx.x] += run;
        }
        __syncthreads();
        if (valid) {
            const unsigned pos = roundBase[d] + warpCount[warp][d] + rankInWarp;
            keysOut[pos] = key;  valsOut[pos] = val;
        }
        __syncthreads();
    }
}

static void radixSort64(unsigned long long*& keys, unsigned*& vals, unsigned long long*& keysTmp, unsigned*& valsTmp, unsigned n,
                        unsigned* blockHist, unsigned* blockOffset, cudaStream_t s) {
    const unsigned numBlocks = (n + SORT_TILE - 1) / SORT_TILE;
    for (int pass = 0; pass < 8; ++pass) {
        const int shift = pass * 8;
        sortHistKernel<<<numBlocks, SORT_THREADS, 0, s>>>(keys, n, shift, blockHist, numBlocks);
        scanExclusiveKernel<<<1, 1024, 0, s>>>(blockHist, blockOffset, 256u * numBlocks, nullptr);
        sortScatterKernel<<<numBlocks, SORT_THREADS, 0, s>>>(keys, vals, keysTmp, valsTmp, n, shift, blockOffset, numBlocks);
        std::swap(keys, keysTmp);  std::swap(vals, valsTmp);
    }
}

// ---------------------------------------------------------------- Karras 2012 hierarchy
// node ids: internal i in [0, n-1), leaf j -> (n-1)+j.  delta = common-prefix length with index tie-break.
__device__ __forceinline__ int kDelta(const unsigned long long* __restrict__ codes, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    const unsigned long long a = codes[i], b = codes[j];
    if (a != b) return __clzll((long long)(a ^ b));
    return 64 + __clz(i ^ j);
}

__global__ void karrasKernel(const unsigned long long* __restrict__ codes, int n, int* __restrict__ left, int* __restrict__ right,
                             int* __restrict__ parent, int2* __restrict__ range) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    const int d = (kDelta(codes, n, i, i + 1) - kDelta(codes, n, i, i - 1)) >= 0 ? 1 : -1;
    const int dmin = kDelta(codes, n, i, i - d);
    int lmax = 2;
    while (kDelta(codes, n, i, i + lmax * d) > dmin) lmax *= 2;
    int l = 0;
    for (int t = lmax / 2; t >= 1; t /= 2)
        if (kDelta(codes, n, i, i + (l + t) * d) > dmin) l += t;
    const int j = i + l * d;
    const int dnode = kDelta(codes, n, i, j);
    int s = 0, t = l;
    do {
        t = (t + 1) >> 1;
        if (kDelta(codes, n, i, i + (s + t) * d) > dnode) s += t;
    } while (t > 1);
    const int gamma = i + s * d + min(d, 0);
    const int lo = min(i, j), hi = max(i, j);
    const int lc = (lo == gamma) ? (n - 1) + gamma : gamma;
    const int rc = (hi == gamma + 1) ? (n - 1) + gamma + 1 : gamma + 1;
    left[i] = lc;  right[i] = rc;
    parent[lc] = i;  parent[rc] = i;
    range[i] = make_int2(lo, hi);
    if (i == 0) parent[0] = -1;
}

__device__ __forceinline__ float boxArea(float4 lo, float4 hi) {
    const float dx = hi.x - lo.x, dy = hi.y - lo.y, dz = hi.z - lo.z;
    return 2.f * (dx * dy + dy * dz + dz * dx);
}

// bottom-up: boxes, SAH cost with and without leaf collapse.  boxLo.w carries the collapsed-tree cost,
// boxHi.w the plain LBVH cost (one triangle per leaf).
__global__ void refitKernel(const float4* __restrict__ triLo, const float4* __restrict__ triHi, const unsigned* __restrict__ sortedGid,
                            int n, const int* __restrict__ left, const int* __restrict__ right, const int* __restrict__ parent,
                            const int2* __restrict__ range, float4* boxLo, float4* boxHi, unsigned* flags, unsigned char* collapse) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    const unsigned g = sortedGid[j];
    float4 lo = triLo[g], hi = triHi[g];
    const float a = boxArea(lo, hi);
    lo.w = SAH_CT * a;  hi.w = SAH_CT * a;
    int node = (n - 1) + j;
    boxLo[node] = lo;  boxHi[node] = hi;
    __threadfence();
    int p = parent[node];
    while (p >= 0) {
        if (atomicAdd(flags + p, 1u) == 0u) return;           // first child to arrive stops; the second one continues
        __threadfence();
        const int lc = left[p], rc = right[p];
        const float4 llo = __ldcg(boxLo + lc), lhi = __ldcg(boxHi + lc), rlo = __ldcg(boxLo + rc), rhi = __ldcg(boxHi + rc);
        float4 plo = make_float4(fminf(llo.x, rlo.x), fminf(llo.y, rlo.y), fminf(llo.z, rlo.z), 0.f);
        float4 phi = make_float4(fmaxf(lhi.x, rhi.x), fmaxf(lhi.y, rhi.y), fmaxf(lhi.z, rhi.z), 0.f);
        const float area = boxArea(plo, phi);
        const int2 rg = range[p];
        const int count = rg.y - rg.x + 1;
        const float cInt = SAH_CI * area + llo.w + rlo.w;
        const float cLeaf = SAH_CT * area * (float)count;
        const bool col = count <= B200PT_MAX_LEAF && cLeaf <= cInt;
        collapse[p] = col ? 1 : 0;
        plo.w = col ? cLeaf : cInt;
        phi.w = SAH_CI * area + lhi.w + rhi.w;
        boxLo[p] = plo;  boxHi[p] = phi;
        __threadfence();
        p = parent[p];
    }
}

// ---------------------------------------------------------------- multi-block exclusive scan (flags -> ranks)
#define SCAN_TILE 4096
__global__ void __launch_bounds__(1024) scanTileSumKernel(const unsigned* __restrict__ in, unsigned n, unsigned* __restrict__ tileSum) {
    __shared__ unsigned part[32];
    const unsigned base = blockIdx.x * SCAN_TILE + threadIdx.x * 4;
    unsigned s = 0;
    for (int k = 0; k < 4; ++k) if (base + k < n) s += in[base + k];
    for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x < 32) {
        unsigned v = part[threadIdx.x];
        for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (threadIdx.x == 0) tileSum[blockIdx.x] = v;
    }
}
__global__ void __launch_bounds__(1024) scanTileApplyKernel(const unsigned* __restrict__ in, unsigned n, const unsigned* __restrict__ tileOffset,
                                                            unsigned* __restrict__ out) {
    __shared__ unsigned part[32];
    const unsigned base = blockIdx.x * SCAN_TILE + threadIdx.x * 4;
    unsigned v[4], s = 0;
    for (int k = 0; k < 4; ++k) { v[k] = base + k < n ? in[base + k] : 0;  s += v[k]; }
    unsigned inc = s;
    for (int o = 1; o < 32; o <<= 1) { const unsigned t = __shfl_up_sync(0xffffffffu, inc, o);  if ((threadIdx.x & 31) >= o) inc += t; }
    if ((threadIdx.x & 31) == 31) part[threadIdx.x >> 5] = inc;
    __syncthreads();
    if (threadIdx.x < 32) {
        unsigned w = part[threadIdx.x], winc = w;
        for (int o = 1; o < 32; o <<= 1) { const unsigned t = __shfl_up_sync(0xffffffffu, winc, o);  if (threadIdx.x >= o) winc += t; }
        part[threadIdx.x] = winc - w;
    }
    __syncthreads();
    unsigned run = tileOffset[blockIdx.x] + part[threadIdx.x >> 5] + inc - s;
    for (int k = 0; k < 4; ++k) { if (base + k < n) out[base + k] = run;  run += v[k]; }
}

// ---------------------------------------------------------------- treelet job selection (see bvh_sah.cuh)
// rank = exclusive prefix of the current level's item-start flags over sorted triangle positions (nullptr: level 1,
// every triangle is an item).  A node roots a job when it holds <= JOB_T items and its parent holds more.
__device__ __forceinline__ unsigned itemsIn(const unsigned* __restrict__ rank, int lo, int hi) { return rank ? rank[hi + 1] - rank[lo] : (unsigned)(hi - lo + 1); }

__global__ void selectJobsKernel(int n, const int* __restrict__ parent, const int2* __restrict__ range, const unsigned* __restrict__ rank,
                                 unsigned* __restrict__ flagNext, const unsigned* __restrict__ rankNext, SahJob* __restrict__ jobs) {
    const int v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= 2 * n - 1) return;
    int2 rg;
    if (v < n - 1) rg = range[v];  else rg = make_int2(v - (n - 1), v - (n - 1));
    const unsigned cnt = itemsIn(rank, rg.x, rg.y);
    if (cnt == 0 || cnt > JOB_T) return;
    const int p = parent[v];
    if (p >= 0) { const int2 pr = range[p];  if (itemsIn(rank, pr.x, pr.y) <= JOB_T) return; }
    if (!rankNext) flagNext[rg.x] = 1u;
    else jobs[rankNext[rg.x]] = SahJob{rank ? rank[rg.x] : (unsigned)rg.x, cnt, (unsigned)rg.x};
}

// SAH cost of the final tree over its (dequantised, i.e. padded) child boxes: sum of Ci*A per inner child and
// Ct*count*A per leaf child; the host adds the root's own term and divides by the root area.
__global__ void sahCostKernel(const BvhNode* __restrict__ nodes, unsigned nRecords, double invScaleX, double invScaleY, double invScaleZ,
                              double* cost) {
    const unsigned r = blockIdx.x * blockDim.x + threadIdx.x;
    float c = 0.f;
    if (r < nRecords) {
        const BvhNode n = nodes[r];
#pragma unroll
        for (int k = 0; k < 2; ++k) {
            const unsigned q0 = n.q[3 * k], q1 = n.q[3 * k + 1], q2 = n.q[3 * k + 2];
            const float dx = (float)((double)((q0 >> 16) - (q0 & 0xffffu)) * invScaleX);
            const float dy = (float)((double)((q1 >> 16) - (q1 & 0xffffu)) * invScaleY);
            const float dz = (float)((double)((q2 >> 16) - (q2 & 0xffffu)) * invScaleZ);
            const float a = 2.f * (dx * dy + dy * dz + dz * dx);
            if (n.link[k] >= 0) c += SAH_CI * a;
            else { unsigned first, count;  leafDecode(n.link[k], first, count);  c += SAH_CT * a * (float)count; }
        }
    }
    for (int o = 16; o; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    if ((threadIdx.x & 31) == 0 && c != 0.f) atomicAdd(cost, (double)c);
}

// 48-byte triangle records in BVH order: {v0, e1 = v1-v0, e2 = v2-v0, gid, mesh}
__global__ void emitTrisKernel(const float3* __restrict__ vtx, const uint3* __restrict__ idx, const unsigned* __restrict__ triMesh,
                               const MeshInfo* __restrict__ meshes, const unsigned* __restrict__ sortedGid, unsigned n, float4* __restrict__ tris) {
    const unsigned s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    const unsigned g = sortedGid[s];
    const unsigned m = triMesh[g];
    const unsigned fv = meshes[m].firstVertex;
    const uint3 id = idx[g];
    const float3 a = vtx[fv + id.x], b = vtx[fv + id.y], c = vtx[fv + id.z];
    const float e1x = __fsub_rn(b.x, a.x), e1y = __fsub_rn(b.y, a.y), e1z = __fsub_rn(b.z, a.z);
    const float e2x = __fsub_rn(c.x, a.x), e2y = __fsub_rn(c.y, a.y), e2z = __fsub_rn(c.z, a.z);
    tris[3 * (size_t)s] = make_float4(a.x, a.y, a.z, e1x);
    tris[3 * (size_t)s + 1] = make_float4(e1y, e1z, e2x, e2y);
    tris[3 * (size_t)s + 2] = make_float4(e2z, __int_as_float((int)g), __int_as_float((int)m), 0.f);
}

__global__ void scatterCodesKernel(const unsigned long long* __restrict__ sortedCodes, const unsigned* __restrict__ sortedGid, unsigned n,
                                   unsigned long long* __restrict__ byGid) {
    const unsigned s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s < n) byGid[sortedGid[s]] = sortedCodes[s];
}

// ---------------------------------------------------------------- host driver
template <typename T> static T* dalloc(size_t n) { T* p = nullptr;  CUDA_OK(cudaMalloc((void**)&p, std::max<size_t>(n, 1) * sizeof(T)));  return p; }

float bvhCollapse4(cudaStream_t s, BvhBuildResult& io);      // returns its own wall time in ms

void bvhFree(BvhBuildResult& r) {
    cudaFree(r.nodes);  cudaFree(r.nodes4);  cudaFree(r.tris);  cudaFree(r.mortonByGid);  cudaFree(r.sortedGid);  cudaFree(r.finalGid);
    r = BvhBuildResult();
}

void bvhBuild(const float3* d_vertices, const uint3* d_indices, const unsigned* d_triMesh, const MeshInfo* d_meshes,
              unsigned nTris, cudaStream_t s, BvhBuildResult& out) {
    out = BvhBuildResult();
    out.nTris = nTris;
    if (nTris == 0) {                                        // empty scene: the root link is the EMPTY sentinel
        out.nodes = dalloc<BvhNode>(1);
        CUDA_OK(cudaMemset(out.nodes, 0, sizeof(BvhNode)));
        out.tris = dalloc<float4>(3);  out.nNodes = 0;  out.rootLink = 0x7fffffff;
        return;
    }
    const int n = (int)nTris;
    cudaEvent_t e0, e1;
    CUDA_OK(cudaEventCreate(&e0));  CUDA_OK(cudaEventCreate(&e1));
    CUDA_OK(cudaEventRecord(e0, s));
    const unsigned B = 256, G = (nTris + B - 1) / B;
    float4* triLo = dalloc<float4>(nTris);  float4* triHi = dalloc<float4>(nTris);
    unsigned* sceneBox = dalloc<unsigned>(8);
    const unsigned boxInit[8] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0u, 0u, 0u, 0u, 0u};
    CUDA_OK(cudaMemcpyAsync(sceneBox, boxInit, sizeof(boxInit), cudaMemcpyHostToDevice, s));
    triBoundsKernel<<<G, B, 0, s>>>(d_vertices, d_indices, d_triMesh, d_meshes, nTris, triLo, triHi, sceneBox);

    unsigned long long* keys = dalloc<unsigned long long>(nTris);  unsigned long long* keysTmp = dalloc<unsigned long long>(nTris);
    unsigned* vals = dalloc<unsigned>(nTris);  unsigned* valsTmp = dalloc<unsigned>(nTris);
    mortonKernel<<<G, B, 0, s>>>(triLo, triHi, nTris, sceneBox, keys, vals);
    const unsigned numBlocks = (nTris + SORT_TILE - 1) / SORT_TILE;
    unsigned* blockHist = dalloc<unsigned>(256u * numBlocks);  unsigned* blockOffset = dalloc<unsigned>(256u * numBlocks);
    radixSort64(keys, vals, keysTmp, valsTmp, nTris, blockHist, blockOffset, s);
    out.sortedGid = vals;                                    // kept: sorted order for the parity hook
    out.mortonByGid = keysTmp;                               // reuse the spare key buffer for codes in input order
    scatterCodesKernel<<<G, B, 0, s>>>(keys, vals, nTris, out.mortonByGid);

    const size_t nAll = 2 * (size_t)n;                       // internal [0,n-1) + leaves [n-1, 2n-1)
    int* left = dalloc<int>(n);  int* right = dalloc<int>(n);  int* parent = dalloc<int>(nAll);  int2* range = dalloc<int2>(n);
    float4* boxLo = dalloc<float4>(nAll);  float4* boxHi = dalloc<float4>(nAll);
    unsigned* flags = dalloc<unsigned>(n);  unsigned char* collapse = dalloc<unsigned char>(n);
    unsigned* scalars = dalloc<unsigned>(4);                 // [0] pairs allocated, [1] -, [2] leaves, [3] max leaf size
    CUDA_OK(cudaMemsetAsync(flags, 0, sizeof(unsigned) * n, s));
    CUDA_OK(cudaMemsetAsync(scalars, 0, 4 * sizeof(unsigned), s));
    CUDA_OK(cudaMemsetAsync(parent, 0xff, sizeof(int) * nAll, s));
    if (n > 1) karrasKernel<<<(n - 1 + B - 1) / B, B, 0, s>>>(keys, n, left, right, parent, range);
    refitKernel<<<G, B, 0, s>>>(triLo, triHi, vals, n, left, right, parent, range, boxLo, boxHi, flags, collapse);

    // ---- binned-SAH treelet refinement, level by level (bvh_sah.cuh)
    out.nodes = dalloc<BvhNode>((size_t)n);                  // at most n-1 inner nodes
    int* rootLinkDev = dalloc<int>(1);
    {   // quantisation grid over the scene AABB (needs the bounds on the host: one small read-back)
        unsigned hb[8];
        CUDA_OK(cudaMemcpyAsync(hb, sceneBox, sizeof(hb), cudaMemcpyDeviceToHost, s));
        CUDA_OK(cudaStreamSynchronize(s));
        for (int a = 0; a < 3; ++a) {
            const double lo = (double)fUnorder(hb[a]), hi = (double)fUnorder(hb[3 + a]);
            out.gridLo[a] = lo;
            out.gridScale[a] = hi > lo ? B200PT_GRID_SPAN / (hi - lo) : 1.0;
        }
    }
    out.tris = dalloc<float4>(3 * (size_t)nTris);
    out.finalGid = dalloc<unsigned>(nTris);
    // per device (a host with one thread per GPU builds on several devices in one process): cheap, so set on every build
    CUDA_OK(cudaFuncSetAttribute(sahJobKernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(SahShared)));
    const unsigned nFlag = nTris + 1, nTiles = (nFlag + SCAN_TILE - 1) / SCAN_TILE;
    unsigned* flagNext = dalloc<unsigned>(nFlag);  unsigned* rankA = dalloc<unsigned>(nFlag);  unsigned* rankB = dalloc<unsigned>(nFlag);
    unsigned* tileSum = dalloc<unsigned>(nTiles);  unsigned* tileOff = dalloc<unsigned>(nTiles);
    static const float ciLeaf = getenv("B200PT_SAH_CI") ? (float)atof(getenv("B200PT_SAH_CI")) : SAH_CI;   // tuning knob
    const unsigned* rank = nullptr;                          // level 1: every triangle is an item
    float4* itemLo = nullptr;  float4* itemHi = nullptr;  int* itemLink = nullptr;  int* itemDepth = nullptr;
    unsigned levels = 0, depthHost = 0;
    const unsigned GV = (unsigned)((nAll + B - 1) / B);
    while (true) {
        unsigned* rankNext = (rank == rankA) ? rankB : rankA;
        CUDA_OK(cudaMemsetAsync(flagNext, 0, sizeof(unsigned) * nFlag, s));
        selectJobsKernel<<<GV, B, 0, s>>>(n, parent, range, rank, flagNext, nullptr, nullptr);
        scanTileSumKernel<<<nTiles, 1024, 0, s>>>(flagNext, nFlag, tileSum);
        scanExclusiveKernel<<<1, 1024, 0, s>>>(tileSum, tileOff, nTiles, nullptr);
        scanTileApplyKernel<<<nTiles, 1024, 0, s>>>(flagNext, nFlag, tileOff, rankNext);
        unsigned nJobs = 0;
        CUDA_OK(cudaMemcpyAsync(&nJobs, rankNext + nTris, sizeof(unsigned), cudaMemcpyDeviceToHost, s));
        CUDA_OK(cudaStreamSynchronize(s));
        if (nJobs == 0) throw std::runtime_error("bvhBuild: treelet selection found no job");
        SahJob* jobs = dalloc<SahJob>(nJobs);
        selectJobsKernel<<<GV, B, 0, s>>>(n, parent, range, rank, flagNext, rankNext, jobs);
        float4* nextLo = dalloc<float4>(nJobs);  float4* nextHi = dalloc<float4>(nJobs);  int* nextLink = dalloc<int>(nJobs);  int* nextDepth = dalloc<int>(nJobs);
        SahLevelIn in{itemLo, itemHi, itemLink, itemDepth, triLo, triHi, vals, levels == 0 ? 1 : 0, ciLeaf};
        SahLevelOut lo{nextLo, nextHi, nextLink, nextDepth, out.finalGid, out.nodes, rootLinkDev,
                       {out.gridLo[0], out.gridLo[1], out.gridLo[2]}, {out.gridScale[0], out.gridScale[1], out.gridScale[2]}, scalars + 0, scalars + 2};
        const bool last = nJobs == 1;
        sahJobKernel<<<nJobs, JOB_THREADS, sizeof(SahShared), s>>>(jobs, in, lo, last ? 1 : 0);
        ++levels;
        if (last) CUDA_OK(cudaMemcpyAsync(&depthHost, nextDepth, sizeof(unsigned), cudaMemcpyDeviceToHost, s));
        CUDA_OK(cudaStreamSynchronize(s));
        CUDA_OK(cudaGetLastError());
        cudaFree(jobs);  cudaFree(itemLo);  cudaFree(itemHi);  cudaFree(itemLink);  cudaFree(itemDepth);
        itemLo = nextLo;  itemHi = nextHi;  itemLink = nextLink;  itemDepth = nextDepth;  rank = rankNext;
        if (last) break;
        if (levels > 80) throw std::runtime_error("bvhBuild: treelet refinement did not converge");
    }
    cudaFree(itemLo);  cudaFree(itemHi);  cudaFree(itemLink);  cudaFree(itemDepth);
    emitTrisKernel<<<G, B, 0, s>>>(d_vertices, d_indices, d_triMesh, d_meshes, out.finalGid, nTris, out.tris);
    unsigned hs[4];
    CUDA_OK(cudaMemcpyAsync(hs, scalars, sizeof(hs), cudaMemcpyDeviceToHost, s));
    CUDA_OK(cudaStreamSynchronize(s));
    out.nNodes = hs[0];
    CUDA_OK(cudaMemcpy(&out.rootLink, rootLinkDev, sizeof(int), cudaMemcpyDeviceToHost));
    double* costDev = dalloc<double>(1);
    CUDA_OK(cudaMemsetAsync(costDev, 0, sizeof(double), s));
    if (out.nNodes) sahCostKernel<<<(out.nNodes + B - 1) / B, B, 0, s>>>(out.nodes, out.nNodes, 1.0 / out.gridScale[0], 1.0 / out.gridScale[1], 1.0 / out.gridScale[2], costDev);
    CUDA_OK(cudaEventRecord(e1, s));
    float4 rootLo, rootHi;  unsigned hostBox[8];  double costHost = 0.0;
    CUDA_OK(cudaMemcpyAsync(&rootLo, boxLo, sizeof(float4), cudaMemcpyDeviceToHost, s));
    CUDA_OK(cudaMemcpyAsync(&rootHi, boxHi, sizeof(float4), cudaMemcpyDeviceToHost, s));
    CUDA_OK(cudaMemcpyAsync(hostBox, sceneBox, sizeof(hostBox), cudaMemcpyDeviceToHost, s));
    CUDA_OK(cudaMemcpyAsync(&costHost, costDev, sizeof(double), cudaMemcpyDeviceToHost, s));
    CUDA_OK(cudaStreamSynchronize(s));
    CUDA_OK(cudaEventElapsedTime(&out.buildMs, e0, e1));
    cudaEventDestroy(e0);  cudaEventDestroy(e1);
    out.maxDepth = depthHost + 1;  out.nLeaves = hs[2];  out.maxLeaf = hs[3];  out.refineLevels = levels;
    const float dx = rootHi.x - rootLo.x, dy = rootHi.y - rootLo.y, dz = rootHi.z - rootLo.z;
    const float rootArea = 2.f * (dx * dy + dy * dz + dz * dx);
    if (out.rootLink >= 0) costHost += SAH_CI * rootArea;  else costHost += SAH_CT * rootArea * (double)nTris;   // the root's own term
    out.sahCost = rootArea > 0.f ? (float)(costHost / rootArea) : 0.f;
    out.sahCostLbvh = rootArea > 0.f ? rootHi.w / rootArea : 0.f;
    for (int a = 0; a < 6; ++a) out.sceneBounds[a] = fUnorder(hostBox[a]);
    cudaFree(triLo);  cudaFree(triHi);  cudaFree(sceneBox);  cudaFree(keys);  cudaFree(valsTmp);  cudaFree(blockHist);  cudaFree(blockOffset);
    cudaFree(left);  cudaFree(right);  cudaFree(parent);  cudaFree(range);  cudaFree(boxLo);  cudaFree(boxHi);  cudaFree(flags);
    cudaFree(collapse);  cudaFree(scalars);  cudaFree(flagNext);  cudaFree(rankA);  cudaFree(rankB);  cudaFree(tileSum);  cudaFree(tileOff);
    cudaFree(costDev);  cudaFree(rootLinkDev);
    if (out.maxDepth > B200PT_STACK - 2) throw std::runtime_error("BVH deeper than the traversal stack");
    out.buildMs += bvhCollapse4(s, out);
}

// ================================================================ collapse: binary records -> 4-wide nodes
// Top-down, one launch per level of the 4-wide tree: a frontier item (binary record r, output node i) takes r's two children and
// twice replaces the inner child with the largest surface area (in grid cells) by that child's own two children.
__device__ __forceinline__ float cellArea(const unsigned* q) {
    const float dx = (float)((q[0] >> 16) - (q[0] & 0xffffu)), dy = (float)((q[1] >> 16) - (q[1] & 0xffffu)), dz = (float)((q[2] >> 16) - (q[2] & 0xffffu));
    return dx * dy + dy * dz + dz * dx;
}
__global__ void collapse4Kernel(const BvhNode* __restrict__ n2, BvhNode4* __restrict__ n4, const int2* __restrict__ frontIn,
                                const unsigned* __restrict__ countIn, int2* __restrict__ frontOut, unsigned* countOut, unsigned* nodeCounter) {
    const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= *countIn) return;
    const int2 it = frontIn[i];
    unsigned q[B200PT_WIDE][3];
    int link[B200PT_WIDE];
    {
        const BvhNode a = n2[it.x];
        for (int k = 0; k < 3; ++k) { q[0][k] = a.q[k];  q[1][k] = a.q[3 + k]; }
        link[0] = a.link[0];  link[1] = a.link[1];
    }
    int n = 2;
    for (int iter = 0; iter < B200PT_WIDE - 2; ++iter) {
        int best = -1;
        float bestA = -1.f;
        for (int c = 0; c < n; ++c)
            if (link[c] >= 0) { const float A = cellArea(q[c]);  if (A > bestA) { bestA = A;  best = c; } }
        if (best < 0) break;
        const BvhNode b = n2[link[best]];
        for (int k = 0; k < 3; ++k) { q[best][k] = b.q[k];  q[n][k] = b.q[3 + k]; }
        link[best] = b.link[0];  link[n] = b.link[1];
        ++n;
    }
    for (int c = 0; c < n; ++c)
        if (link[c] >= 0) {
            const unsigned idx = atomicAdd(nodeCounter, 1u), pos = atomicAdd(countOut, 1u);
            frontOut[pos] = make_int2(link[c], (int)idx);
            link[c] = (int)idx;
        }
    for (int c = n; c < B200PT_WIDE; ++c) { q[c][0] = q[c][1] = q[c][2] = 0x0000ffffu;  link[c] = B200PT_LINK_EMPTY; }
    unsigned w[4 * B200PT_WIDE];                                  // record: 3 words per child, then the links
#pragma unroll
    for (int c = 0; c < B200PT_WIDE; ++c) { w[3 * c] = q[c][0];  w[3 * c + 1] = q[c][1];  w[3 * c + 2] = q[c][2];  w[3 * B200PT_WIDE + c] = (unsigned)link[c]; }
    uint4* dst = reinterpret_cast<uint4*>(n4) + B200PT_WIDE * (size_t)it.y;
#pragma unroll
    for (int k = 0; k < B200PT_WIDE; ++k) dst[k] = make_uint4(w[4 * k], w[4 * k + 1], w[4 * k + 2], w[4 * k + 3]);
}
// (re)derives io.nodes4 from io.nodes; io.nodes4 is allocated on first use (as many nodes as binary records at most)
float bvhCollapse4(cudaStream_t s, BvhBuildResult& io) {
    io.rootLink4 = io.rootLink;  io.nNodes4 = 0;  io.maxDepth4 = 0;
    if (io.rootLink < 0 || io.rootLink == B200PT_LINK_EMPTY || !io.nNodes) return 0.f;
    CUDA_OK(cudaStreamSynchronize(s));
    const auto t0 = std::chrono::steady_clock::now();      // leaf root / empty scene: no inner nodes
    const unsigned nr = io.nNodes;
    if (!io.nodes4) io.nodes4 = dalloc<BvhNode4>(nr);
    int2* front[2] = {dalloc<int2>(nr), dalloc<int2>(nr)};
    unsigned* ctr = dalloc<unsigned>(4);                     // [0],[1] frontier sizes, [2] node counter
    const unsigned init[4] = {1u, 0u, 1u, 0u};
    const int2 root = make_int2(io.rootLink, 0);
    CUDA_OK(cudaMemcpyAsync(ctr, init, sizeof(init), cudaMemcpyHostToDevice, s));
    CUDA_OK(cudaMemcpyAsync(front[0], &root, sizeof(root), cudaMemcpyHostToDevice, s));
    unsigned width = 1, depth = 0;
    int cur = 0;
    while (width) {                                           // one small read-back per level (the 4-wide tree is ~log4 deep)
        CUDA_OK(cudaMemsetAsync(ctr + (cur ^ 1), 0, sizeof(unsigned), s));
        collapse4Kernel<<<(width + 127) / 128, 128, 0, s>>>(io.nodes, io.nodes4, front[cur], ctr + cur, front[cur ^ 1], ctr + (cur ^ 1), ctr + 2);
        CUDA_OK(cudaMemcpyAsync(&width, ctr + (cur ^ 1), sizeof(unsigned), cudaMemcpyDeviceToHost, s));
        CUDA_OK(cudaStreamSynchronize(s));
        cur ^= 1;  ++depth;
    }
    CUDA_OK(cudaMemcpy(&io.nNodes4, ctr + 2, sizeof(unsigned), cudaMemcpyDeviceToHost));
    io.rootLink4 = 0;  io.maxDepth4 = depth;
    cudaFree(front[0]);  cudaFree(front[1]);  cudaFree(ctr);
    if ((B200PT_WIDE - 1) * io.maxDepth4 + B200PT_WIDE > B200PT_STACK) throw std::runtime_error("wide BVH deeper than the traversal stack");
    return std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();
}

// ================================================================ refit: new vertex positions, same topology
// (what OptiX calls OPTIX_BUILD_OPERATION_UPDATE; the reference always rebuilds, sv4/SimplePathtracer.cpp:670-737)
//   1. refitTrisKernel    new 48-byte triangle records in the existing leaf order + per-slot AABB + scene AABB
//   2. refitParentsKernel parent of every record (records carry child links only)
//   3. refitBoundsKernel  bottom-up: a thread per record folds its leaf children, the second child to arrive at a record
//                         unions the pair and carries it to the parent (Karras-style atomic visit counter)
//   4. refitQuantKernel   exact child boxes -> the records' 16-bit grid cells over the NEW scene box
__global__ void refitTrisKernel(const float3* __restrict__ vtx, const uint3* __restrict__ idx, const unsigned* __restrict__ triMesh,
                                const MeshInfo* __restrict__ meshes, const unsigned* __restrict__ finalGid, unsigned n,
                                float4* __restrict__ tris, float4* __restrict__ slotLo, float4* __restrict__ slotHi, unsigned* sceneBox) {
    const unsigned s = blockIdx.x * blockDim.x + threadIdx.x;
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    if (s < n) {
        const unsigned g = finalGid[s];
        const unsigned m = triMesh[g];
        const unsigned fv = meshes[m].firstVertex;
        const uint3 id = idx[g];
        const float3 a = vtx[fv + id.x], b = vtx[fv + id.y], c = vtx[fv + id.z];
        const float e1x = __fsub_rn(b.x, a.x), e1y = __fsub_rn(b.y, a.y), e1z = __fsub_rn(b.z, a.z);
        const float e2x = __fsub_rn(c.x, a.x), e2y = __fsub_rn(c.y, a.y), e2z = __fsub_rn(c.z, a.z);
        tris[3 * (size_t)s] = make_float4(a.x, a.y, a.z, e1x);
        tris[3 * (size_t)s + 1] = make_float4(e1y, e1z, e2x, e2y);
        tris[3 * (size_t)s + 2] = make_float4(e2z, __int_as_float((int)g), __int_as_float((int)m), 0.f);
        lo[0] = fminf(fminf(a.x, b.x), c.x);  hi[0] = fmaxf(fmaxf(a.x, b.x), c.x);
        lo[1] = fminf(fminf(a.y, b.y), c.y);  hi[1] = fmaxf(fmaxf(a.y, b.y), c.y);
        lo[2] = fminf(fminf(a.z, b.z), c.z);  hi[2] = fmaxf(fmaxf(a.z, b.z), c.z);
        slotLo[s] = make_float4(lo[0], lo[1], lo[2], 0.f);
        slotHi[s] = make_float4(hi[0], hi[1], hi[2], 0.f);
    }
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        for (int o = 16; o; o >>= 1) {
            lo[a] = fminf(lo[a], __shfl_xor_sync(0xffffffffu, lo[a], o));
            hi[a] = fmaxf(hi[a], __shfl_xor_sync(0xffffffffu, hi[a], o));
        }
        if ((threadIdx.x & 31) == 0 && lo[a] <= hi[a]) { atomicMin(sceneBox + a, fOrder(lo[a]));  atomicMax(sceneBox + 3 + a, fOrder(hi[a])); }
    }
}
__global__ void refitParentsKernel(const BvhNode* __restrict__ nodes, unsigned nRecords, int* __restrict__ parent) {
    const unsigned r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= nRecords) return;
#pragma unroll
    for (int k = 0; k < 2; ++k) { const int l = nodes[r].link[k];  if (l >= 0) parent[l] = (int)(2 * r + k); }
}
// childBox: 12 floats per record {c0.lo, c0.hi, c1.lo, c1.hi}; visit: one counter per record, zeroed
__global__ void refitBoundsKernel(const BvhNode* __restrict__ nodes, unsigned nRecords, const int* __restrict__ parent,
                                  const float4* __restrict__ slotLo, const float4* __restrict__ slotHi, float* childBox, unsigned* visit) {
    const unsigned r0 = blockIdx.x * blockDim.x + threadIdx.x;
    if (r0 >= nRecords) return;
    for (int k = 0; k < 2; ++k) {
        const int l = nodes[r0].link[k];
        if (l >= 0) continue;
        unsigned first, count;
        leafDecode(l, first, count);
        float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
        for (unsigned t = first; t < first + count; ++t) {
            const float4 a = slotLo[t], b = slotHi[t];
            lo[0] = fminf(lo[0], a.x);  lo[1] = fminf(lo[1], a.y);  lo[2] = fminf(lo[2], a.z);
            hi[0] = fmaxf(hi[0], b.x);  hi[1] = fmaxf(hi[1], b.y);  hi[2] = fmaxf(hi[2], b.z);
        }
        unsigned r = r0;
        int which = k;
        for (;;) {
            volatile float* cb = childBox + 12 * (size_t)r + 6 * which;
            cb[0] = lo[0];  cb[1] = lo[1];  cb[2] = lo[2];  cb[3] = hi[0];  cb[4] = hi[1];  cb[5] = hi[2];
            __threadfence();
            if (atomicAdd(visit + r, 1u) == 0u) break;                  // the sibling has not arrived yet: it will carry on
            __threadfence();
            const volatile float* ob = childBox + 12 * (size_t)r + 6 * (which ^ 1);
            lo[0] = fminf(lo[0], ob[0]);  lo[1] = fminf(lo[1], ob[1]);  lo[2] = fminf(lo[2], ob[2]);
            hi[0] = fmaxf(hi[0], ob[3]);  hi[1] = fmaxf(hi[1], ob[4]);  hi[2] = fmaxf(hi[2], ob[5]);
            const int p = parent[r];
            if (p < 0) break;                                           // r is the root record
            r = (unsigned)p >> 1;  which = p & 1;
        }
    }
}
__global__ void refitQuantKernel(BvhNode* __restrict__ nodes, unsigned nRecords, const float* __restrict__ childBox,
                                 double gx, double gy, double gz, double sx, double sy, double sz) {
    const unsigned r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= nRecords) return;
    const double gl[3] = {gx, gy, gz}, gs[3] = {sx, sy, sz};
    const float* cb = childBox + 12 * (size_t)r;
    unsigned q[6];
#pragma unroll
    for (int c = 0; c < 2; ++c)
#pragma unroll
        for (int a = 0; a < 3; ++a) {                              // same outward rounding as the SAH write-out (bvh_sah.cuh)
            const double l = B200PT_GRID_ORIGIN + ((double)cb[6 * c + a] - gl[a]) * gs[a];
            const double h = B200PT_GRID_ORIGIN + ((double)cb[6 * c + 3 + a] - gl[a]) * gs[a];
            const unsigned ql = (unsigned)min(65535, max(0, (int)floor(l) - B200PT_GRID_PAD));
            const unsigned qh = (unsigned)min(65535, max(0, (int)ceil(h) + B200PT_GRID_PAD));
            q[3 * c + a] = ql | (qh << 16);
        }
    uint4* dst = reinterpret_cast<uint4*>(nodes) + 2 * (size_t)r;
    const uint4 old1 = dst[1];
    dst[0] = make_uint4(q[0], q[1], q[2], q[3]);
    dst[1] = make_uint4(q[4], q[5], old1.z, old1.w);
}

void bvhRefit(const float3* d_vertices, const uint3* d_indices, const unsigned* d_triMesh, const MeshInfo* d_meshes, cudaStream_t s,
              BvhBuildResult& io, float* refitMs) {
    const unsigned n = io.nTris, nr = io.nNodes;
    if (!n) { if (refitMs) *refitMs = 0.f;  return; }
    const unsigned B = 256, G = (n + B - 1) / B, GR = (std::max(nr, 1u) + B - 1) / B;
    cudaEvent_t e0, e1;
    CUDA_OK(cudaEventCreate(&e0));  CUDA_OK(cudaEventCreate(&e1));
    CUDA_OK(cudaEventRecord(e0, s));
    float4* slotLo = dalloc<float4>(n);
    float4* slotHi = dalloc<float4>(n);
    unsigned* sceneBox = dalloc<unsigned>(8);
    int* parent = dalloc<int>(nr);
    float* childBox = dalloc<float>(12 * (size_t)nr);
    unsigned* visit = dalloc<unsigned>(nr);
    const unsigned initBox[8] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0u, 0u, 0u, 0u, 0u};
    CUDA_OK(cudaMemcpyAsync(sceneBox, initBox, sizeof(initBox), cudaMemcpyHostToDevice, s));
    refitTrisKernel<<<G, B, 0, s>>>(d_vertices, d_indices, d_triMesh, d_meshes, io.finalGid, n, io.tris, slotLo, slotHi, sceneBox);
    unsigned hb[8];
    CUDA_OK(cudaMemcpyAsync(hb, sceneBox, sizeof(hb), cudaMemcpyDeviceToHost, s));
    if (nr) {
        CUDA_OK(cudaMemsetAsync(parent, 0xff, sizeof(int) * (size_t)nr, s));
        CUDA_OK(cudaMemsetAsync(visit, 0, sizeof(unsigned) * (size_t)nr, s));
        refitParentsKernel<<<GR, B, 0, s>>>(io.nodes, nr, parent);
        refitBoundsKernel<<<GR, B, 0, s>>>(io.nodes, nr, parent, slotLo, slotHi, childBox, visit);
    }
    CUDA_OK(cudaStreamSynchronize(s));
    for (int a = 0; a < 3; ++a) {
        const double lo = (double)fUnorder(hb[a]), hi = (double)fUnorder(hb[3 + a]);
        io.gridLo[a] = lo;
        io.gridScale[a] = hi > lo ? B200PT_GRID_SPAN / (hi - lo) : 1.0;
        io.sceneBounds[a] = (float)lo;  io.sceneBounds[3 + a] = (float)hi;
    }
    if (nr) refitQuantKernel<<<GR, B, 0, s>>>(io.nodes, nr, childBox, io.gridLo[0], io.gridLo[1], io.gridLo[2], io.gridScale[0], io.gridScale[1], io.gridScale[2]);
    double* costDev = dalloc<double>(1);
    CUDA_OK(cudaMemsetAsync(costDev, 0, sizeof(double), s));
    if (nr) sahCostKernel<<<GR, B, 0, s>>>(io.nodes, nr, 1.0 / io.gridScale[0], 1.0 / io.gridScale[1], 1.0 / io.gridScale[2], costDev);
    CUDA_OK(cudaEventRecord(e1, s));
    double costHost = 0.0;
    CUDA_OK(cudaMemcpyAsync(&costHost, costDev, sizeof(double), cudaMemcpyDeviceToHost, s));
    CUDA_OK(cudaStreamSynchronize(s));
    CUDA_OK(cudaGetLastError());
    float ms = 0.f;
    CUDA_OK(cudaEventElapsedTime(&ms, e0, e1));
    if (refitMs) *refitMs = ms;
    cudaEventDestroy(e0);  cudaEventDestroy(e1);
    const double dx = (double)io.sceneBounds[3] - io.sceneBounds[0], dy = (double)io.sceneBounds[4] - io.sceneBounds[1], dz = (double)io.sceneBounds[5] - io.sceneBounds[2];
    const double rootArea = 2.0 * (dx * dy + dy * dz + dz * dx);
    if (io.rootLink >= 0) costHost += SAH_CI * rootArea;  else costHost += SAH_CT * rootArea * (double)n;
    io.sahCost = rootArea > 0.0 ? (float)(costHost / rootArea) : 0.f;
    cudaFree(slotLo);  cudaFree(slotHi);  cudaFree(sceneBox);  cudaFree(parent);  cudaFree(childBox);  cudaFree(visit);  cudaFree(costDev);
    const float cms = bvhCollapse4(s, io);                   // the 4-wide nodes are copies of the binary records' boxes
    if (refitMs) *refitMs += cms;
}
