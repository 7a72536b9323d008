// b200pt — BVH traversal kernels (replace the closed optixTrace of sv4/deviceProgram.cu:207,232).
//   RADIANCE rays  — closest hit in (tmin,tmax), ties -> lowest global triangle id; the result {t,u,v,triSlot} is written at the ray's own queue index
//   OCCLUSION rays — any hit (optionally only triangles front-facing the ray), early out;
//                    visible rays add the NEE contribution that travels with them to the path's radiance accumulator
// Both kinds are traced by the same launch of traceKernel (one launch per bounce).
// Persistent warps (a fixed grid of numSMs() x k CTAs).  Each warp stages chunks of 32 rays from the
// device-side queues into shared memory (one ticket atomic per chunk, cp.async, double-buffered and prefetched);
// a lane whose ray terminates takes the next staged ray as soon as enough lanes are idle (per-lane replacement
// decided by a warp vote), so a warp never idles behind its slowest ray.  Every loop iteration each lane takes up to
// NODE_UNROLL inner-node steps if it can; a lane that reaches a leaf postpones it and keeps walking; once enough lanes
// are blocked on a second leaf (warp vote) the warp intersects all pending leaves together.
// A traversal step is TWO 256-bit loads (LDG.E.256, new on sm_100): the 64-byte record of a 4-wide node holds four child
// boxes, quantised to a 16-bit grid over the scene AABB, and four links; a 48-byte triangle is 3x128-bit.  All fetches go
// through the read-only path.  The kernel sits between the issue rate and the L1 wavefront rate (one wavefront per lane per
// divergent load).  Algorithmic bytes per ray = 64*nodes + 48*tris + 32 (ray) + 16 (hit).
// The ray/triangle test is the exact fp32 operation sequence of oracle/scene_core.h (no FMA contraction,
// enforced with __f*_rn intrinsics), which makes the chosen triangle and (t,u,v) BVH-independent.
#include "common.cuh"
#include <atomic>
CHK_DECL          // B200PT_CHECKED builds: violation counter of this file (codes 10-19)

#ifndef TRACE_BLOCK
#define TRACE_BLOCK 128
#endif
#define TRACE_WARPS (TRACE_BLOCK / 32)
#ifndef TRACE_BLOCKS_PER_SM
#define TRACE_BLOCKS_PER_SM 7      // 72 registers per thread, 28 warps/SM: measured best for the 4-wide step (5: 3.25 ms, 6: 3.11, 7: 2.94, 8: 3.02, 9: 3.11 per frame)
#endif
#define FULL 0xffffffffu
#define EMPTY 0x7fffffff          // "nothing left to visit"; also the root link of an empty scene
#define NO_LEAF 0                 // leaf codes are < 0
#ifndef TRACE_LEAF_PASSES
#define TRACE_LEAF_PASSES 0     // 0: the leaves a lane holds (postponed + blocking [+ next on the stack]) go through ONE flattened triangle loop;
#endif                          // n > 0: n separate passes of one leaf each (1 = the original scheme).  Per frame: 1: 2.75 ms, 2: 2.71, flattened: 2.62
#ifndef TRACE_LEAF_MERGE
#define TRACE_LEAF_MERGE 2      // leaves merged into the flattened loop (2 or 3)
#endif
#ifndef NODE_ROUNDS
#define NODE_ROUNDS 8           // vote rounds before the staging logic runs again
#endif
#ifndef NODE_UNROLL
#define NODE_UNROLL 3           // node steps per lane between two warp votes
#endif

static int g_refillThreshold = -1, g_leafThreshold = -1;   // lanes that must be idle before a warp refills; tuned on B200

__device__ __forceinline__ float4 ldg4(const float4* p) { return __ldg(p); }

struct Node8 { unsigned w[8]; };   // BvhNode: w[0..5] = 12 quantised bounds (two children), w[6], w[7] = links
__device__ __forceinline__ Node8 ldgNode(const uint4* p) {        // ONE 256-bit load per traversal step
    Node8 n;
    asm("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
        : "=r"(n.w[0]), "=r"(n.w[1]), "=r"(n.w[2]), "=r"(n.w[3]), "=r"(n.w[4]), "=r"(n.w[5]), "=r"(n.w[6]), "=r"(n.w[7]) : "l"(p));
    return n;
}

// Ray in grid coordinates.  A quantised plane q is decoded with one PRMT into the float 2^23 + q (exact), and the slab
// distance is ONE FMA: t = (2^23 + q) * inv + c with c = -(2^23 + og) * inv, og = ray origin in cells, inv = 1 / (d * cells per
// unit).  c is formed with a single rounding (fma of og*inv onto the exact 2^23*inv), so a plane is off by at most ~0.51 cell
// (0.5 from c, 0.01 from og); the same `inv` multiplies both terms, so its own rounding only rescales t by 1 +- 2^-23.  Every
// stored box carries GRID_PAD = 2 cells of padding, which covers this.  (fp32 only: plain FP64 is slow on this part.)
#ifndef TRACE_FAST_RCP
#define TRACE_FAST_RCP 1         // rcp.approx for the slab reciprocals: measured -1.1 % trace time, hits unchanged (boxes only cull)
#endif
#ifndef TRACE_BOX_SLACK
#define TRACE_BOX_SLACK 1
#endif
#ifndef TRACE_SORT_CEX
#define TRACE_SORT_CEX 3          // compare-exchanges of the child sort: 5 = full order, 4 = nearest + farthest exact, 3 = nearest only.  Measured on B200
                                  // (trace ms per frame, node visits per ray): 5: 2.945, 14.6; 4: 2.900, 14.7; 3: 2.866, 14.8 - the instructions saved outweigh the order lost
#endif
__device__ __forceinline__ float rcpApprox(float x) { float r;  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));  return r; }
struct GridRay { float ix, iy, iz, cx, cy, cz;  unsigned nx, ny, nz; };    // n*: PRMT selector of the NEAR plane's half-word on that axis
__device__ __forceinline__ GridRay makeGridRay(const SceneDev& S, const float3 o, const float3 d) {
    GridRay g;
    const float sx = (float)S.gridScale[0], sy = (float)S.gridScale[1], sz = (float)S.gridScale[2];
    const float ox = fmaf(o.x - (float)S.gridLo[0], sx, (float)B200PT_GRID_ORIGIN);
    const float oy = fmaf(o.y - (float)S.gridLo[1], sy, (float)B200PT_GRID_ORIGIN);
    const float oz = fmaf(o.z - (float)S.gridLo[2], sz, (float)B200PT_GRID_ORIGIN);
    // a zero direction component must not turn the slab into NaNs (the axis would be ignored and the ray would walk the whole
    // tree along it): clamp |d * s| to 2^-100, so inv stays finite and the 0.51-cell bound above still holds
    const float tiny = 7.8886090522101181e-31f;
    float dx = d.x * sx, dy = d.y * sy, dz = d.z * sz;
    if (!(fabsf(dx) >= tiny)) dx = copysignf(tiny, dx);
    if (!(fabsf(dy) >= tiny)) dy = copysignf(tiny, dy);
    if (!(fabsf(dz) >= tiny)) dz = copysignf(tiny, dz);
#if TRACE_FAST_RCP
    g.ix = rcpApprox(dx);  g.iy = rcpApprox(dy);  g.iz = rcpApprox(dz);      // any error of inv only rescales t (see above)
#else
    g.ix = 1.0f / dx;  g.iy = 1.0f / dy;  g.iz = 1.0f / dz;
#endif
    g.nx = dx < 0.f ? 0x7632u : 0x7610u;  g.ny = dy < 0.f ? 0x7632u : 0x7610u;  g.nz = dz < 0.f ? 0x7632u : 0x7610u;
    g.cx = -fmaf(ox, g.ix, 8388608.0f * g.ix);  g.cy = -fmaf(oy, g.iy, 8388608.0f * g.iy);  g.cz = -fmaf(oz, g.iz, 8388608.0f * g.iz);
    return g;
}
// 2^23 + one half-word of w: selector 0x7610 takes the low half (lo plane), 0x7632 the high half (hi plane)
__device__ __forceinline__ float qSel(unsigned w, unsigned sel) { return __uint_as_float(__byte_perm(w, 0x4B000000u, sel)); }
// slab test of one child: words {lo.x|hi.x<<16, lo.y|hi.y<<16, lo.z|hi.z<<16}.  The ray's direction signs pick the near and
// the far plane of each axis directly (far selector = near ^ 0x22), so no per-axis min/max is needed.
__device__ __forceinline__ bool boxTest(unsigned w0, unsigned w1, unsigned w2, const GridRay& g, float tmin, float tmax, float& tnear) {
    const float ax = fmaf(qSel(w0, g.nx), g.ix, g.cx), bx = fmaf(qSel(w0, g.nx ^ 0x22u), g.ix, g.cx);
    const float ay = fmaf(qSel(w1, g.ny), g.iy, g.cy), by = fmaf(qSel(w1, g.ny ^ 0x22u), g.iy, g.cy);
    const float az = fmaf(qSel(w2, g.nz), g.iz, g.cz), bz = fmaf(qSel(w2, g.nz ^ 0x22u), g.iz, g.cz);
    const float t0 = fmaxf(fmaxf(ax, ay), fmaxf(az, tmin));
    const float t1 = fminf(fminf(bx, by), fminf(bz, tmax));
    tnear = t0;
#if TRACE_BOX_SLACK
    return t0 <= t1 * 1.0000004f;
#else
    return t0 <= t1;
#endif
}

__device__ __forceinline__ void prefetchL1(const void* p) { asm volatile("prefetch.global.L1 [%0];" :: "l"(p)); }
#ifndef TRACE_PREFETCH
#define TRACE_PREFETCH 0          // bit 0: child pair of a pushed node, bit 1: triangles of a postponed leaf.  Measured on B200: no gain (bit 0: -3 %), so off
#endif
// warms L1 for what a link refers to: the child pair of an inner node, or the first triangle lines of a leaf
__device__ __forceinline__ void prefetchLink(const SceneDev& S, int link) {
    if (link >= 0) prefetchL1(S.nodes + 2 * (size_t)link);
    else { const unsigned c = (unsigned)(-(link + 1));  prefetchL1(S.tris + 3 * (size_t)(c >> 3)); }
}

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

// The same operation sequence as triTest up to the divisions, without early outs: the traversal kernel runs it for every
// lane of a leaf phase in lock step and divides only for the lanes that pass.  `ok` is triTest's accept condition before the
// t-range test (comparisons written in the same negated form, so NaNs fall through identically).
__device__ __forceinline__ bool triEval(const float4 a, const float4 b, const float4 c, const float3 o, const float3 d, bool cullBack,
                                        float& det, float& U, float& V, float& Tn) {
    const float v0x = a.x, v0y = a.y, v0z = a.z, e1x = a.w, e1y = b.x, e1z = b.y, e2x = b.z, e2y = b.w, e2z = c.x;
    const float px = __fsub_rn(__fmul_rn(d.y, e2z), __fmul_rn(d.z, e2y));
    const float py = __fsub_rn(__fmul_rn(d.z, e2x), __fmul_rn(d.x, e2z));
    const float pz = __fsub_rn(__fmul_rn(d.x, e2y), __fmul_rn(d.y, e2x));
    det = __fadd_rn(__fadd_rn(__fmul_rn(e1x, px), __fmul_rn(e1y, py)), __fmul_rn(e1z, pz));
    const float sx = __fsub_rn(o.x, v0x), sy = __fsub_rn(o.y, v0y), sz = __fsub_rn(o.z, v0z);
    U = __fadd_rn(__fadd_rn(__fmul_rn(sx, px), __fmul_rn(sy, py)), __fmul_rn(sz, pz));
    const float qx = __fsub_rn(__fmul_rn(sy, e1z), __fmul_rn(sz, e1y));
    const float qy = __fsub_rn(__fmul_rn(sz, e1x), __fmul_rn(sx, e1z));
    const float qz = __fsub_rn(__fmul_rn(sx, e1y), __fmul_rn(sy, e1x));
    V = __fadd_rn(__fadd_rn(__fmul_rn(d.x, qx), __fmul_rn(d.y, qy)), __fmul_rn(d.z, qz));
    Tn = __fadd_rn(__fadd_rn(__fmul_rn(e2x, qx), __fmul_rn(e2y, qy)), __fmul_rn(e2z, qz));
    const float UV = __fadd_rn(U, V);
    const bool inPos = !(U < 0.0f || V < 0.0f || UV > det), inNeg = !(U > 0.0f || V > 0.0f || UV < det);
    return det != 0.0f && !(cullBack && det < 0.0f) && (det > 0.0f ? inPos : inNeg);
}

struct Best { float t, u, v; int slot; unsigned gid; };

// Traversal stack: a per-thread array of links in local memory (the hardware interleaves it by lane, so lanes at the same depth share a line).
// Measured on B200 and rejected (profiles/README.md, r01h): the top entry in a register with predicated PTX ld.local/st.local instead of nvcc's
// branch around the LDL (2.85 against 2.76 ms per frame); entry distances stored beside the links and stale entries culled on pop (node visits
// per ray 14.8 -> 14.0, but 3.28 ms: the second array doubles the stack's L1 traffic).
#define STACK_DECL_ int stack[B200PT_STACK];  int sp = 0;
#define PUSH_IF_(x, c) { if ((c) && CHK(sp < B200PT_STACK, 11)) stack[sp++] = (x); }
#define POP_IF_(dst, c) { if (c) dst = sp ? stack[--sp] : EMPTY; }
#define POP_(dst) POP_IF_(dst, true)

// One step through a 4-wide node (two 256-bit loads): test the four child boxes, order the hit ones by entry distance with a
// 5-exchange sorting network, continue with the nearest and push the others, farthest first.
#define CEX_(ta, la, tb, lb) { const bool s_ = tb < ta;  const float tt_ = s_ ? tb : ta;  const int ll_ = s_ ? lb : la;  tb = s_ ? ta : tb;  lb = s_ ? la : lb;  ta = tt_;  la = ll_; }
#if B200PT_WIDE == 4
#define NODE_STEP4()                                                                             \
    {                                                                                            \
        const uint4* np_ = S.nodes4 + 4 * (size_t)(CHK((unsigned)cur < S.nNodes4, 12) ? cur : 0);   \
        const Node8 A_ = ldgNode(np_), B_ = ldgNode(np_ + 2);                                    \
        float t0_, t1_, t2_, t3_;                                                                \
        const bool h0_ = boxTest(A_.w[0], A_.w[1], A_.w[2], g, tmin, best.t, t0_);               \
        const bool h1_ = boxTest(A_.w[3], A_.w[4], A_.w[5], g, tmin, best.t, t1_);               \
        const bool h2_ = boxTest(A_.w[6], A_.w[7], B_.w[0], g, tmin, best.t, t2_);               \
        const bool h3_ = boxTest(B_.w[1], B_.w[2], B_.w[3], g, tmin, best.t, t3_);               \
        const float inf_ = __int_as_float(0x7f800000);                                           \
        t0_ = h0_ ? t0_ : inf_;  t1_ = h1_ ? t1_ : inf_;  t2_ = h2_ ? t2_ : inf_;  t3_ = h3_ ? t3_ : inf_;   \
        int l0_ = (int)B_.w[4], l1_ = (int)B_.w[5], l2_ = (int)B_.w[6], l3_ = (int)B_.w[7];      \
        CEX_(t0_, l0_, t1_, l1_)  CEX_(t2_, l2_, t3_, l3_)  CEX_(t0_, l0_, t2_, l2_)                    \
        if (TRACE_SORT_CEX >= 4) CEX_(t1_, l1_, t3_, l3_)                                        \
        if (TRACE_SORT_CEX >= 5) CEX_(t1_, l1_, t2_, l2_)                                        \
        PUSH_IF_(l3_, t3_ < inf_)  PUSH_IF_(l2_, t2_ < inf_)  PUSH_IF_(l1_, t1_ < inf_)             \
        { const bool any_ = t0_ < inf_;  cur = any_ ? l0_ : cur;  POP_IF_(cur, !any_) }                 \
    }
#else
// The 8-wide variant (tuning builds, -DB200PT_WIDE=8): four 256-bit loads, eight box tests; the nearest hit child is visited next (a chain of seven
// compare-exchanges leaves it in slot 0), the other hit children are pushed in slot order.
// Measured on B200 (profiles/README.md, r02f): node visits per ray 14.6 -> 10.2, identical hits, trace 2.60 -> 3.07 ms per frame: the kernel is issue-bound and
// 10.2 x 8 box tests cost more instructions than 14.6 x 4.  Kept as a variant, not the default.
#define NODE_STEP4()                                                                             \
    {                                                                                            \
        const uint4* np_ = S.nodes4 + 8 * (size_t)(CHK((unsigned)cur < S.nNodes4, 12) ? cur : 0);   \
        const Node8 A_ = ldgNode(np_), B_ = ldgNode(np_ + 2), C_ = ldgNode(np_ + 4), D_ = ldgNode(np_ + 6);   \
        const float inf_ = __int_as_float(0x7f800000);                                           \
        float t0_, t1_, t2_, t3_, t4_, t5_, t6_, t7_;                                            \
        t0_ = boxTest(A_.w[0], A_.w[1], A_.w[2], g, tmin, best.t, t0_) ? t0_ : inf_;             \
        t1_ = boxTest(A_.w[3], A_.w[4], A_.w[5], g, tmin, best.t, t1_) ? t1_ : inf_;             \
        t2_ = boxTest(A_.w[6], A_.w[7], B_.w[0], g, tmin, best.t, t2_) ? t2_ : inf_;             \
        t3_ = boxTest(B_.w[1], B_.w[2], B_.w[3], g, tmin, best.t, t3_) ? t3_ : inf_;             \
        t4_ = boxTest(B_.w[4], B_.w[5], B_.w[6], g, tmin, best.t, t4_) ? t4_ : inf_;             \
        t5_ = boxTest(B_.w[7], C_.w[0], C_.w[1], g, tmin, best.t, t5_) ? t5_ : inf_;             \
        t6_ = boxTest(C_.w[2], C_.w[3], C_.w[4], g, tmin, best.t, t6_) ? t6_ : inf_;             \
        t7_ = boxTest(C_.w[5], C_.w[6], C_.w[7], g, tmin, best.t, t7_) ? t7_ : inf_;             \
        int l0_ = (int)D_.w[0], l1_ = (int)D_.w[1], l2_ = (int)D_.w[2], l3_ = (int)D_.w[3];      \
        int l4_ = (int)D_.w[4], l5_ = (int)D_.w[5], l6_ = (int)D_.w[6], l7_ = (int)D_.w[7];      \
        CEX_(t0_, l0_, t1_, l1_)  CEX_(t0_, l0_, t2_, l2_)  CEX_(t0_, l0_, t3_, l3_)  CEX_(t0_, l0_, t4_, l4_)   \
        CEX_(t0_, l0_, t5_, l5_)  CEX_(t0_, l0_, t6_, l6_)  CEX_(t0_, l0_, t7_, l7_)             \
        PUSH_IF_(l7_, t7_ < inf_)  PUSH_IF_(l6_, t6_ < inf_)  PUSH_IF_(l5_, t5_ < inf_)  PUSH_IF_(l4_, t4_ < inf_)   \
        PUSH_IF_(l3_, t3_ < inf_)  PUSH_IF_(l2_, t2_ < inf_)  PUSH_IF_(l1_, t1_ < inf_)          \
        { const bool any_ = t0_ < inf_;  cur = any_ ? l0_ : cur;  POP_IF_(cur, !any_) }                 \
    }
#endif

// Reference single-ray traversal (parity hooks only; the production kernels below inline the same steps).
template <bool ANY, bool COUNT>
__device__ __forceinline__ bool traverse(const SceneDev& S, const float3 o, const float3 d, float tmin, float tmax,
                                         bool cullBack, Best& best, unsigned& nNodes, unsigned& nTris) {
    const GridRay g = makeGridRay(S, o, d);
    STACK_DECL_
    best.t = tmax;  best.slot = -1;  best.gid = 0xffffffffu;  best.u = best.v = 0.f;
    int cur = S.rootLink4;
    while (cur != EMPTY) {
        if (cur >= 0) {
            if (COUNT) nNodes += 1;
            NODE_STEP4()
        } else {
            unsigned first, count;
            leafDecode(cur, first, count);
            for (unsigned k = 0; k < count; ++k) {
                const float4* tp = S.tris + 3 * (size_t)(CHK(first + k < S.nTris, 13) ? first + k : 0u);
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
            POP_(cur)
        }
    }
    return best.slot >= 0;
}

__device__ __forceinline__ void flushStats(TraceStats* stats, unsigned nNodes, unsigned nTris, unsigned nRays, unsigned nHits) {
    for (int off = 16; off; off >>= 1) {
        nNodes += __shfl_xor_sync(FULL, nNodes, off);
        nTris += __shfl_xor_sync(FULL, nTris, off);
        nRays += __shfl_xor_sync(FULL, nRays, off);
        nHits += __shfl_xor_sync(FULL, nHits, off);
    }
    if ((threadIdx.x & 31) == 0) {
        atomicAdd(&stats->nodes, (unsigned long long)nNodes);
        atomicAdd(&stats->tris, (unsigned long long)nTris);
        atomicAdd(&stats->rays, (unsigned long long)nRays);
        atomicAdd(&stats->hits, (unsigned long long)nHits);
    }
}

#ifndef TRACE_STREAM_BYPASS_L1
#define TRACE_STREAM_BYPASS_L1 1
#endif
// cp.async (LDGSTS): 16 bytes global -> shared without a register round trip; completion via wait_all.
__device__ __forceinline__ void cpAsync16(void* smemDst, const void* gmemSrc) {
    const unsigned d = (unsigned)__cvta_generic_to_shared(smemDst);
#if TRACE_STREAM_BYPASS_L1
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(d), "l"(gmemSrc) : "memory");   // rays are read once: keep them out of L1 (L2 only)
#else
    asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" :: "r"(d), "l"(gmemSrc) : "memory");
#endif
}
__device__ __forceinline__ void cpAsyncCommit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cpAsyncWaitAll() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// One launch traces BOTH queues of a bounce: the extension rays (closest hit -> hit[i] = {t,u,v,triSlot}) and the
// shadow rays the previous shade pass emitted (any hit; rayO.w = path slot, rayD.w = target: 0 direct, 1 indirect,
// 2 alpha-when-occluded, see shade.cu).  Work is handed out in chunks of 32 rays by one ticket counter; chunk ids
// below ceil(countC/32) are extension rays, the rest shadow rays, so the short any-hit rays fill the tail of the
// long closest-hit ones.  The next chunk is prefetched (ticket atomic, then cp.async into the spare stage buffer)
// while the warp is still traversing the current one.
template <bool COUNT>
__global__ void __launch_bounds__(TRACE_BLOCK, TRACE_BLOCKS_PER_SM)
traceKernel(SceneDev S, const float4* __restrict__ rayO, const float4* __restrict__ rayD, float4* __restrict__ hit,
            const float4* __restrict__ shO, const float4* __restrict__ shD, const float4* __restrict__ shP,
            unsigned* ctr, int closest, float tminC, float tminS, float tmax, int cullBack, PathPool pool,
            TraceStats* stats, int refillThreshold, int leafThreshold, int nodeMin) {
    __shared__ float4 stageO[TRACE_WARPS][2][32], stageD[TRACE_WARPS][2][32];
    const unsigned countC = closest ? ctr[QC_RAYS] : 0u, countS = ctr[QC_SHADOWS];
    unsigned* const head = ctr + QC_HEAD;
    const unsigned chunksC = (countC + 31u) >> 5, chunksTotal = chunksC + ((countS + 31u) >> 5);
    const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const unsigned ltMask = (1u << lane) - 1u;
    const int rootLink = S.rootLink4;
    unsigned nNodes[2] = {0, 0}, nTris[2] = {0, 0}, nRays[2] = {0, 0}, nHits[2] = {0, 0}, myNodes = 0, myTris = 0;

    // warp-uniform staging state: rays [stageNext, stageEnd) of the current chunk are still unassigned
    unsigned stageNext = 0, stageEnd = 0, stageFirst = 0;  int stageAny = 0, buf = 0;
    // prefetch state: 0 idle, 1 ticket in flight, 2 chunk landing in buffer buf^1, 3 queue exhausted
    int pf = 0;  unsigned ticket = 0, pfFirst = 0, pfEnd = 0;  int pfAny = 0, pfAge = 0;

    // per-lane ray state
    bool active = false, isAny = false;
    float tmin = tminC;                    // per lane: occlusion rays have their own tmin (0.01 in every variant, also where radiance rays use 0.001)
    unsigned rayIdx = 0;
    float3 o = make_float3(0, 0, 0), d = o;
    GridRay g = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    float wO = 0.f, wD = 0.f;
    Best best;  best.t = tmax;  best.slot = -1;  best.gid = 0xffffffffu;  best.u = best.v = 0.f;
    float bestDet = 1.f;
    int cur = EMPTY, leaf = NO_LEAF;
    STACK_DECL_

    while (true) {
        // ---------------------------------------------------------------- prefetch the next chunk (warp-uniform)
        if (pf == 0 && stageEnd - stageNext <= 8u) {
            if (lane == 0) ticket = atomicAdd(head, 1u);          // not read until a later iteration: no stall here
            pf = 1;  pfAge = 0;
        } else if (pf == 1 && ++pfAge >= 3) {
            const unsigned c = __shfl_sync(FULL, ticket, 0);
            if (c >= chunksTotal) pf = 3;
            else {
                pfAny = c >= chunksC;
                const unsigned first = (pfAny ? c - chunksC : c) << 5, cnt = pfAny ? countS : countC;
                pfFirst = first;  pfEnd = min(first + 32u, cnt);
                if (first + lane < pfEnd) {
                    cpAsync16(&stageO[warp][buf ^ 1][lane], (pfAny ? shO : rayO) + first + lane);
                    cpAsync16(&stageD[warp][buf ^ 1][lane], (pfAny ? shD : rayD) + first + lane);
                }
                cpAsyncCommit();
                pf = 2;
            }
        }
        // ---------------------------------------------------------------- refill idle lanes (warp-synchronous)
        unsigned idle = __ballot_sync(FULL, !active);
        if (idle == FULL || ((int)__popc(idle) >= refillThreshold && !(pf == 3 && stageNext >= stageEnd))) {
#pragma unroll 1
            for (int round = 0; round < 2 && idle; ++round) {
                if (stageNext >= stageEnd) {                      // current chunk used up: switch to the prefetched one
                    if (pf == 0) { if (lane == 0) ticket = atomicAdd(head, 1u);  pf = 1; }
                    if (pf == 1) {
                        const unsigned c = __shfl_sync(FULL, ticket, 0);
                        if (c >= chunksTotal) pf = 3;
                        else {
                            pfAny = c >= chunksC;
                            const unsigned first = (pfAny ? c - chunksC : c) << 5, cnt = pfAny ? countS : countC;
                            pfFirst = first;  pfEnd = min(first + 32u, cnt);
                            if (first + lane < pfEnd) {
                                cpAsync16(&stageO[warp][buf ^ 1][lane], (pfAny ? shO : rayO) + first + lane);
                                cpAsync16(&stageD[warp][buf ^ 1][lane], (pfAny ? shD : rayD) + first + lane);
                            }
                            cpAsyncCommit();
                            pf = 2;
                        }
                    }
                    if (pf == 3) break;
                    cpAsyncWaitAll();
                    __syncwarp();
                    buf ^= 1;  stageFirst = pfFirst;  stageNext = pfFirst;  stageEnd = pfEnd;  stageAny = pfAny;  pf = 0;
                }
                const unsigned avail = stageEnd - stageNext;
                const unsigned r = __popc(idle & ltMask);
                if (!active && r < avail) {
                    const unsigned j = stageNext + r;
                    const unsigned sj_ = CHK(j - stageFirst < 32u, 15) ? j - stageFirst : 0u;
                    const float4 o4 = stageO[warp][buf][sj_], d4 = stageD[warp][buf][sj_];
                    o = make_float3(o4.x, o4.y, o4.z);  d = make_float3(d4.x, d4.y, d4.z);  wO = o4.w;  wD = d4.w;
                    g = makeGridRay(S, o, d);
                    rayIdx = j;  active = true;  isAny = stageAny != 0;  tmin = isAny ? tminS : tminC;
                    best.t = tmax;  best.slot = -1;  best.gid = 0xffffffffu;  best.u = best.v = 0.f;
                    cur = rootLink;  leaf = NO_LEAF;  sp = 0;
                    if (COUNT) { nRays[isAny]++;  myNodes = 0;  myTris = 0; }
                }
                stageNext += min(avail, (unsigned)__popc(idle));
                __syncwarp();
                idle = __ballot_sync(FULL, !active);
            }
            if (idle == FULL) break;                               // queues drained and every lane finished
        }

        // ---------------------------------------------------------------- inner-node steps: one per lane per round, up to
        // NODE_ROUNDS rounds before the staging logic above runs again; a round ends early once enough lanes are blocked on a leaf
        bool occluded = false;
        unsigned nodeMask, blockedMask;
#pragma unroll 1
        for (int round = 0; ; ++round) {
#pragma unroll
            for (int sub_ = 0; sub_ < NODE_UNROLL; ++sub_)
            if (active && (unsigned)cur < (unsigned)EMPTY) {
                if (COUNT) myNodes += 1;
                NODE_STEP4()
                {                                                   // first leaf: postpone it and keep walking
                    const bool first_ = cur < 0 && leaf == NO_LEAF;
                    leaf = first_ ? cur : leaf;
                    POP_IF_(cur, first_)
                }
            }
            const bool canNode = active && (unsigned)cur < (unsigned)EMPTY;
            nodeMask = __ballot_sync(FULL, canNode);
            blockedMask = __ballot_sync(FULL, active && !canNode);
            if (nodeMask == 0u || (int)__popc(blockedMask) >= leafThreshold || round + 1 >= NODE_ROUNDS) break;
            if ((int)__popc(nodeMask) <= nodeMin && blockedMask != 0u) { nodeMask = 0u;  break; }      // too few walkers left: intersect the pending leaves now
        }
        // ---------------------------------------------------------------- leaves, once enough lanes are blocked on one
        if (nodeMask == 0u || (int)__popc(blockedMask) >= leafThreshold) {
            if (active && leaf == NO_LEAF && cur < 0) { leaf = cur;  POP_(cur) }   // root leaf
#if TRACE_LEAF_PASSES == 0
            // both leaves of a blocked lane in ONE flattened triangle loop: the warp iterates max(cntA + cntB) times, not max(cntA) + max(cntB)
            if (active && leaf != NO_LEAF) {
                unsigned firstA, cntA, firstB = 0, cntB = 0;
                leafDecode(leaf, firstA, cntA);
                if (cur < 0) { leafDecode(cur, firstB, cntB);  POP_(cur) }
#if TRACE_LEAF_MERGE >= 3
                unsigned firstC = 0, cntC = 0;
                if (cur < 0) { leafDecode(cur, firstC, cntC);  POP_(cur) }
                const unsigned cntAB = cntA + cntB, total = cntAB + cntC;
#else
                const unsigned total = cntA + cntB;
#endif
                const bool cull = isAny && cullBack != 0;
                for (unsigned k = 0; k < total; ++k) {
#if TRACE_LEAF_MERGE >= 3
                    const unsigned slot_ = k < cntA ? firstA + k : (k < cntAB ? firstB + (k - cntA) : firstC + (k - cntAB));
#else
                    const unsigned slot_ = k < cntA ? firstA + k : firstB + (k - cntA);
#endif
                    const float4* tp = S.tris + 3 * (size_t)(CHK(slot_ < S.nTris, 13) ? slot_ : 0u);
                    const float4 a = ldg4(tp), b = ldg4(tp + 1), c = ldg4(tp + 2);
                    if (COUNT) myTris += 1;
                    float det, U, V, Tn;
                    if (triEval(a, b, c, o, d, cull, det, U, V, Tn) && (Tn != 0.0f || tmin < 0.0f)) {
                        const float t = __fdiv_rn(Tn, det);
                        if (t > tmin && t < tmax) {
                            if (isAny) { occluded = true;  break; }
                            const unsigned gid = (unsigned)__float_as_int(c.y);
                            if (t < best.t || (t == best.t && gid < best.gid)) {
                                best.t = t;  best.u = U;  best.v = V;  bestDet = det;  best.slot = (int)slot_;  best.gid = gid;
                            }
                        }
                    }
                }
                leaf = NO_LEAF;
                if (occluded) cur = EMPTY;
                else if (cur < 0) { leaf = cur;  POP_(cur) }
            }
#else
#pragma unroll 1
            for (int pass_ = 0; pass_ < TRACE_LEAF_PASSES; ++pass_)
            if (active && leaf != NO_LEAF) {
                unsigned first, cnt;
                leafDecode(leaf, first, cnt);
                const bool cull = isAny && cullBack != 0;
                for (unsigned k = 0; k < cnt; ++k) {
                    const float4* tp = S.tris + 3 * (size_t)(CHK(first + k < S.nTris, 13) ? first + k : 0u);
                    const float4 a = ldg4(tp), b = ldg4(tp + 1), c = ldg4(tp + 2);
                    if (COUNT) myTris += 1;
                    float det, U, V, Tn;
                    // Tn == 0 (ray origin exactly in the triangle's plane: common for bounce rays leaving axis-aligned geometry) gives t = +-0,
                    // which t > tmin rejects for every tmin >= 0; skipping it here keeps ~0.4 M zero-numerator divisions per launch out of the
                    // IEEE division's slow path (3 % of the kernel's warp instructions in the r01h source-level profile)
                    if (triEval(a, b, c, o, d, cull, det, U, V, Tn) && (Tn != 0.0f || tmin < 0.0f)) {
                        const float t = __fdiv_rn(Tn, det);       // the only division in the loop; u, v are divided at retirement
                        if (t > tmin && t < tmax) {
                            if (isAny) { occluded = true;  break; }
                            const unsigned gid = (unsigned)__float_as_int(c.y);
                            if (t < best.t || (t == best.t && gid < best.gid)) {
                                best.t = t;  best.u = U;  best.v = V;  bestDet = det;  best.slot = (int)(first + k);  best.gid = gid;
                            }
                        }
                    }
                }
                leaf = NO_LEAF;
                if (occluded) cur = EMPTY;
                else if (cur < 0) { leaf = cur;  POP_(cur) }   // (EMPTY is positive: never taken for it)
            }
#endif
        }
        // ---------------------------------------------------------------- retire finished rays
        if (active && cur == EMPTY && leaf == NO_LEAF) {
            if (!isAny) {
                // best.u/best.v hold the undivided U, V of the winning triangle: same quotients as dividing at hit time
                const bool got = best.slot >= 0;
                if (CHK(rayIdx < countC && rayIdx < pool.cap, 10)) hit[rayIdx] = make_float4(best.t, got ? __fdiv_rn(best.u, bestDet) : 0.f, got ? __fdiv_rn(best.v, bestDet) : 0.f, __int_as_float(best.slot));
            } else {
                // prd.radiance += pathThroughput * lightSample (sv4/deviceProgram.cu:659) resolved here.  Target 2 = alpha
                // when OCCLUDED (shadow catcher, :663-664).  Single writer per slot per launch => deterministic.
                const int target = __float_as_int(wD);
                if (occluded == (target == 2) && CHK(rayIdx < countS && (unsigned)__float_as_int(wO) < pool.cap && (unsigned)target <= 2u, 14)) {
                    const unsigned slot = (unsigned)__float_as_int(wO);
                    float4* acc = (target == 0 ? pool.direct : (target == 1 ? pool.indirect : pool.alpha)) + slot;
#if TRACE_STREAM_BYPASS_L1
                    const float4 pend = __ldcg(shP + rayIdx);                 // touched once per launch: L2 only, L1 stays with the tree
                    float4 r = __ldcg(acc);
#else
                    const float4 pend = shP[rayIdx];
                    float4 r = *acc;
#endif
                    r.x = __fadd_rn(r.x, pend.x);  r.y = __fadd_rn(r.y, pend.y);  r.z = __fadd_rn(r.z, pend.z);
                    *acc = r;
                }
            }
            if (COUNT) { nNodes[isAny] += myNodes;  nTris[isAny] += myTris;  nHits[isAny] += isAny ? (occluded ? 1u : 0u) : (best.slot >= 0 ? 1u : 0u); }
            active = false;
        }
    }
    if (COUNT) { flushStats(stats, nNodes[0], nTris[0], nRays[0], nHits[0]);  flushStats(stats + 1, nNodes[1], nTris[1], nRays[1], nHits[1]); }
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

void checkedReportTrace(unsigned* count, unsigned* first) {
    *count = 0;  *first = 0;
#ifdef B200PT_CHECKED
    cudaMemcpyFromSymbol(count, g_chkCount, sizeof(unsigned));  cudaMemcpyFromSymbol(first, g_chkFirst, sizeof(unsigned));
#endif
}

static int refillThreshold() {
    if (g_refillThreshold < 0) {
        const char* e = getenv("B200PT_REFILL");                  // tuning knob; the default is the measured optimum
        g_refillThreshold = e ? atoi(e) : 6;
        if (g_refillThreshold < 1) g_refillThreshold = 1;
        if (g_refillThreshold > 32) g_refillThreshold = 32;
    }
    return g_refillThreshold;
}
static int nodeMinThreshold() {
    static int v = -1;
    if (v < 0) { const char* e = getenv("B200PT_NODEMIN");  v = e ? atoi(e) : 8;  if (v < 0) v = 0;  if (v > 31) v = 31; }   // tuning knob; measured per frame: 0: 2.63 ms, 4-16: 2.59-2.60
    return v;
}
static int leafThreshold() {
    if (g_leafThreshold < 0) {
        const char* e = getenv("B200PT_LEAF");                    // lanes blocked on a leaf before the warp intersects triangles
        g_leafThreshold = e ? atoi(e) : 14;          // measured with the flattened two-leaf loop: 10: 2.61 ms, 12: 2.60, 14: 2.58, 16: 2.59
        if (g_leafThreshold < 1) g_leafThreshold = 1;
        if (g_leafThreshold > 32) g_leafThreshold = 32;
    }
    return g_leafThreshold;
}

void launchTrace(const SceneDev& scene, const Queues& Q, int cur, int closest, unsigned* ctr, float tmin, float tminShadow, float tmax, int cullBack,
                 const PathPool& pool, TraceStats* stats, cudaStream_t stream) {
    static const int ctasPerSM = [] { const char* e = getenv("B200PT_TRACE_CTAS");  const int v = e ? atoi(e) : TRACE_BLOCKS_PER_SM;  return v < 1 ? 1 : (v > TRACE_BLOCKS_PER_SM ? TRACE_BLOCKS_PER_SM : v); }();   // tuning knob: fewer persistent CTAs per SM leave room for a concurrent shadeKernel of another frame
    const dim3 grid((unsigned)numSMs() * ctasPerSM), block(TRACE_BLOCK);
    // Shared-memory carve-out: 7 CTAs x (8 KB of ray staging + 1 KB reserved) fit the 64 KB configuration, which leaves the most L1 for the tree and the
    // stacks.  Left to itself the driver picks a larger carve-out here; measured on B200 (trace ms per frame): default 2.601, 25-30 % (= 64 KB) 2.558,
    // 50 % 2.611, 100 % 2.738.  shadeKernel asks for the same configuration (shade.cu): kernels with different carve-outs do not share an SM, and with
    // frames in flight they should (shade at 0 % or 10 %: 596 Msamples/s in flight instead of 603).  Set once per device (B200PT_TRACE_CARVEOUT overrides).
    static const int carveout = [] { const char* e = getenv("B200PT_TRACE_CARVEOUT");  return e ? atoi(e) : B200PT_SMEM_CARVEOUT; }();
    static std::atomic<unsigned long long> carved{0ull};
    int dev = 0;
    if (cudaGetDevice(&dev) == cudaSuccess && dev >= 0 && dev < 64 && !(carved.load() >> dev & 1ull)) {
        cudaFuncSetAttribute(traceKernel<false>, cudaFuncAttributePreferredSharedMemoryCarveout, carveout);
        cudaFuncSetAttribute(traceKernel<true>, cudaFuncAttributePreferredSharedMemoryCarveout, carveout);
        carved.fetch_or(1ull << dev);
    }
    if (stats) traceKernel<true><<<grid, block, 0, stream>>>(scene, Q.rayO[cur], Q.rayD[cur], Q.hit, Q.shO, Q.shD, Q.shP, ctr, closest, tmin, tminShadow, tmax, cullBack, pool, stats, refillThreshold(), leafThreshold(), nodeMinThreshold());
    else traceKernel<false><<<grid, block, 0, stream>>>(scene, Q.rayO[cur], Q.rayD[cur], Q.hit, Q.shO, Q.shD, Q.shP, ctr, closest, tmin, tminShadow, tmax, cullBack, pool, nullptr, refillThreshold(), leafThreshold(), nodeMinThreshold());
}

void launchTraceRaysDebug(const SceneDev& scene, const float4* rayO, const float4* rayD, unsigned n, float tmin, float tmax,
                          int anyHit, int cullBack, float4* hitOut, cudaStream_t stream) {
    if (!n) return;
    traceRaysDebugKernel<<<(n + 127) / 128, 128, 0, stream>>>(scene, rayO, rayD, n, tmin, tmax, anyHit, cullBack, hitOut);
}
