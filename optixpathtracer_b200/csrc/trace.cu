// b200pt — BVH traversal kernels (replace the closed optixTrace of sv4/deviceProgram.cu:207,232).
//   traceClosestKernel : RADIANCE rays  — closest hit in (tmin,tmax), ties -> lowest global triangle id
//   traceAnyKernel     : OCCLUSION rays — any hit (optionally only triangles front-facing the ray), early out;
//                        visible rays add their pending NEE contribution to the path's radiance accumulator
// Persistent threads: a fixed grid of B200PT_NUM_SMS x k CTAs; each warp pulls batches of 32 rays from a
// device-side queue head, so queue lengths never return to the host.  Node pairs are fetched as 4x128-bit
// loads (two 32-byte siblings, 64-byte aligned), triangles as 3x128-bit loads through the read-only path.
// Roofline: HBM/L2 gather bound; algorithmic bytes per ray = 32*nodes + 48*tris + 32 (ray) + 16 (hit).
// The ray/triangle test is the exact fp32 operation sequence of oracle/scene_core.h (no FMA contraction,
// enforced with __f*_rn intrinsics), which makes the chosen triangle and (t,u,v) BVH-independent.
#include "common.cuh"

#define TRACE_BLOCK 128
#define TRACE_BLOCKS_PER_SM 8

__device__ __forceinline__ float4 ldg4(const float4* p) { return __ldg(p); }

struct TriHit { float t, u, v; };

// Moeller-Trumbore, unfused.  det > 0 <=> triangle front-faces the ray.
__device__ __forceinline__ bool triTest(const float4 a, const float4 b, const float4 c, const float3 o, const float3 d,
                                        float tmin, float tmax, bool cullBack, TriHit& h) {
    const float v0x = a.x, v0y = a.y, v0z = a.z, e1x = a.w, e1y = b.x, e1z = b.y, e2x = b.z, e2y = b.w, e2z = c.x;
    const float px = __fsub_rn(__fmul_rn(d.y, e2z), __fmul_rn(d.z, e2y));
    const float py = __fsub_rn(__fmul_rn(d.z, e2x), __fmul_rn(d.x, e2z));
    const float pz = __fsub_rn(__fmul_rn(d.x, e2y), __fmul_rn(d.y, e2x));
    const float det = __fadd_rn(__fadd_rn(__fmul_rn(e1x, px), __fmul_rn(e1y, py)), __fmul_rn(e1z, pz));
    if (det == 0.0f) return false;
    if (cullBack && det < 0.0f) return false;
    const float sx = __fsub_rn(o.x, v0x), sy = __fsub_rn(o.y, v0y), sz = __fsub_rn(o.z, v0z);
    const float U = __fadd_rn(__fadd_rn(__fmul_rn(sx, px), __fmul_rn(sy, py)), __fmul_rn(sz, pz));
    const float qx = __fsub_rn(__fmul_rn(sy, e1z), __fmul_rn(sz, e1y));
    const float qy = __fsub_rn(__fmul_rn(sz, e1x), __fmul_rn(sx, e1z));
    const float qz = __fsub_rn(__fmul_rn(sx, e1y), __fmul_rn(sy, e1x));
    const float V = __fadd_rn(__fadd_rn(__fmul_rn(d.x, qx), __fmul_rn(d.y, qy)), __fmul_rn(d.z, qz));
    if (det > 0.0f) { if (U < 0.0f || V < 0.0f || __fadd_rn(U, V) > det) return false; }
    else            { if (U > 0.0f || V > 0.0f || __fadd_rn(U, V) < det) return false; }
    const float Tn = __fadd_rn(__fadd_rn(__fmul_rn(e2x, qx), __fmul_rn(e2y, qy)), __fmul_rn(e2z, qz));
    const float t = __fdiv_rn(Tn, det);
    if (!(t > tmin && t < tmax)) return false;
    h.t = t;  h.u = __fdiv_rn(U, det);  h.v = __fdiv_rn(V, det);
    return true;
}

// slab test against one 32-byte node (boxes are padded at build time; far side widened by 2 ulp)
__device__ __forceinline__ bool boxTest(const float4 n0, const float4 n1, const float3 o, const float3 inv,
                                        float tmin, float tmax, float& tnear) {
    const float ax = (n0.x - o.x) * inv.x, bx = (n1.x - o.x) * inv.x;
    const float ay = (n0.y - o.y) * inv.y, by = (n1.y - o.y) * inv.y;
    const float az = (n0.z - o.z) * inv.z, bz = (n1.z - o.z) * inv.z;
    const float t0 = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fmaxf(fminf(az, bz), tmin));
    const float t1 = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fminf(fmaxf(az, bz), tmax));
    tnear = t0;
    return t0 <= t1 * 1.0000004f;
}

struct Best { float t, u, v; int slot; unsigned gid; };

template <bool ANY, bool COUNT>
__device__ __forceinline__ bool traverse(const SceneDev& S, const float3 o, const float3 d, float tmin, float tmax,
                                         bool cullBack, Best& best, unsigned& nNodes, unsigned& nTris) {
    const float3 inv = make_float3(1.0f / d.x, 1.0f / d.y, 1.0f / d.z);
    int stack[B200PT_STACK];
    int sp = 0;
    best.t = tmax;  best.slot = -1;  best.gid = 0xffffffffu;  best.u = best.v = 0.f;
    int cur = __float_as_int(ldg4(S.nodes).w);            // root record's link
    const int EMPTY = 0x7fffffff;
    while (cur != EMPTY) {
        if (cur >= 0) {
            const float4* p = S.nodes + 2 * (size_t)cur;
            const float4 l0 = ldg4(p), l1 = ldg4(p + 1), r0 = ldg4(p + 2), r1 = ldg4(p + 3);
            if (COUNT) nNodes += 2;
            float tl, tr;
            const bool hl = boxTest(l0, l1, o, inv, tmin, best.t, tl);
            const bool hr = boxTest(r0, r1, o, inv, tmin, best.t, tr);
            const int cl = __float_as_int(l0.w), cr = __float_as_int(r0.w);
            if (hl && hr) {
                const bool swap = tr < tl;
                stack[sp++] = swap ? cl : cr;
                cur = swap ? cr : cl;
            } else if (hl) cur = cl;
            else if (hr) cur = cr;
            else cur = sp ? stack[--sp] : EMPTY;
        } else {
            unsigned first, count;
            leafDecode(cur, first, count);
            for (unsigned k = 0; k < count; ++k) {
                const float4* tp = S.tris + 3 * (size_t)(first + k);
                const float4 a = ldg4(tp), b = ldg4(tp + 1), c = ldg4(tp + 2);
                if (COUNT) nTris += 1;
                TriHit h;
                if (triTest(a, b, c, o, d, tmin, tmax, cullBack, h)) {
                    if (ANY) return true;
                    const unsigned gid = (unsigned)__float_as_int(c.y);
                    if (h.t < best.t || (h.t == best.t && gid < best.gid)) {
                        best.t = h.t;  best.u = h.u;  best.v = h.v;  best.slot = (int)(first + k);  best.gid = gid;
                    }
                }
            }
            cur = sp ? stack[--sp] : EMPTY;
        }
    }
    return best.slot >= 0;
}

__device__ __forceinline__ void flushStats(TraceStats* stats, unsigned nNodes, unsigned nTris, unsigned nRays) {
    for (int off = 16; off; off >>= 1) {
        nNodes += __shfl_xor_sync(0xffffffffu, nNodes, off);
        nTris += __shfl_xor_sync(0xffffffffu, nTris, off);
        nRays += __shfl_xor_sync(0xffffffffu, nRays, off);
    }
    if ((threadIdx.x & 31) == 0) {
        atomicAdd(&stats->nodes, (unsigned long long)nNodes);
        atomicAdd(&stats->tris, (unsigned long long)nTris);
        atomicAdd(&stats->rays, (unsigned long long)nRays);
    }
}

template <bool COUNT>
__global__ void __launch_bounds__(TRACE_BLOCK, TRACE_BLOCKS_PER_SM)
traceClosestKernel(SceneDev S, const float4* __restrict__ rayO, const float4* __restrict__ rayD, float4* __restrict__ hit,
                   const unsigned* __restrict__ countPtr, unsigned* head, float tmin, float tmax, TraceStats* stats) {
    const unsigned count = *countPtr;
    const unsigned lane = threadIdx.x & 31;
    unsigned nNodes = 0, nTris = 0, nRays = 0;
    while (true) {
        unsigned base = 0;
        if (lane == 0) base = atomicAdd(head, 32u);
        base = __shfl_sync(0xffffffffu, base, 0);
        if (base >= count) break;
        const unsigned i = base + lane;
        if (i < count) {
            const float4 o4 = rayO[i], d4 = rayD[i];
            Best b;
            traverse<false, COUNT>(S, make_float3(o4.x, o4.y, o4.z), make_float3(d4.x, d4.y, d4.z), tmin, tmax, false, b, nNodes, nTris);
            hit[i] = make_float4(b.t, b.u, b.v, __int_as_float(b.slot));
            if (COUNT) nRays++;
        }
    }
    if (COUNT) flushStats(stats, nNodes, nTris, nRays);
}

template <bool COUNT>
__global__ void __launch_bounds__(TRACE_BLOCK, TRACE_BLOCKS_PER_SM)
traceAnyKernel(SceneDev S, const float4* __restrict__ shO, const float4* __restrict__ shD,
               const unsigned* __restrict__ countPtr, unsigned* head, float tmin, float tmax, int cullBack,
               PathPool pool, TraceStats* stats) {
    const unsigned count = *countPtr;
    const unsigned lane = threadIdx.x & 31;
    unsigned nNodes = 0, nTris = 0, nRays = 0;
    while (true) {
        unsigned base = 0;
        if (lane == 0) base = atomicAdd(head, 32u);
        base = __shfl_sync(0xffffffffu, base, 0);
        if (base >= count) break;
        const unsigned i = base + lane;
        if (i < count) {
            const float4 o4 = shO[i], d4 = shD[i];
            Best b;
            const bool occluded = traverse<true, COUNT>(S, make_float3(o4.x, o4.y, o4.z), make_float3(d4.x, d4.y, d4.z), tmin, tmax,
                                                        cullBack != 0, b, nNodes, nTris);
            if (COUNT) nRays++;
            // prd.radiance += pathThroughput * lightSample (sv4/deviceProgram.cu:659) resolved here.  d4.w carries the
            // target: 0 = direct (depth 0), 1 = indirect, 2 = alpha when OCCLUDED (shadow catcher, :663-664).
            // Single writer per slot per launch => deterministic.
            const int target = __float_as_int(d4.w);
            if (occluded == (target == 2)) {
                const unsigned slot = (unsigned)__float_as_int(o4.w);
                const float4 pend = pool.pending[slot];
                float4* acc = (target == 0 ? pool.direct : (target == 1 ? pool.indirect : pool.alpha)) + slot;
                float4 r = *acc;
                r.x = __fadd_rn(r.x, pend.x);  r.y = __fadd_rn(r.y, pend.y);  r.z = __fadd_rn(r.z, pend.z);
                *acc = r;
            }
        }
    }
    if (COUNT) flushStats(stats, nNodes, nTris, nRays);
}

__global__ void traceRaysDebugKernel(SceneDev S, const float4* rayO, const float4* rayD, unsigned n, float tmin, float tmax,
                                     int anyHit, int cullBack, float4* hitOut) {
    const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 o4 = rayO[i], d4 = rayD[i];
    Best b;
    unsigned a = 0, c = 0;
    if (anyHit) {
        const bool occ = traverse<true, false>(S, make_float3(o4.x, o4.y, o4.z), make_float3(d4.x, d4.y, d4.z), tmin, tmax, cullBack != 0, b, a, c);
        hitOut[i] = make_float4(0.f, 0.f, 0.f, __int_as_float(occ ? 1 : 0));
    } else {
        traverse<false, false>(S, make_float3(o4.x, o4.y, o4.z), make_float3(d4.x, d4.y, d4.z), tmin, tmax, false, b, a, c);
        hitOut[i] = make_float4(b.t, b.u, b.v, __int_as_float(b.slot));
    }
}

void launchTraceClosest(const SceneDev& scene, const float4* rayO, const float4* rayD, float4* hit, const unsigned* count,
                        unsigned* head, float tmin, float tmax, TraceStats* stats, cudaStream_t stream) {
    const dim3 grid(B200PT_NUM_SMS * TRACE_BLOCKS_PER_SM), block(TRACE_BLOCK);
    if (stats) traceClosestKernel<true><<<grid, block, 0, stream>>>(scene, rayO, rayD, hit, count, head, tmin, tmax, stats);
    else traceClosestKernel<false><<<grid, block, 0, stream>>>(scene, rayO, rayD, hit, count, head, tmin, tmax, nullptr);
}

void launchTraceAny(const SceneDev& scene, const float4* shO, const float4* shD, const unsigned* count, unsigned* head,
                    float tmin, float tmax, int cullBack, const PathPool& pool, TraceStats* stats, cudaStream_t stream) {
    const dim3 grid(B200PT_NUM_SMS * TRACE_BLOCKS_PER_SM), block(TRACE_BLOCK);
    if (stats) traceAnyKernel<true><<<grid, block, 0, stream>>>(scene, shO, shD, count, head, tmin, tmax, cullBack, pool, stats);
    else traceAnyKernel<false><<<grid, block, 0, stream>>>(scene, shO, shD, count, head, tmin, tmax, cullBack, pool, nullptr);
}

void launchTraceRaysDebug(const SceneDev& scene, const float4* rayO, const float4* rayD, unsigned n, float tmin, float tmax,
                          int anyHit, int cullBack, float4* hitOut, cudaStream_t stream) {
    if (!n) return;
    traceRaysDebugKernel<<<(n + 127) / 128, 128, 0, stream>>>(scene, rayO, rayD, n, tmin, tmax, anyHit, cullBack, hitOut);
}
