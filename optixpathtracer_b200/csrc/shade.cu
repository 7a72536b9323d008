// b200pt — wavefront stages around the traversal kernels (replace the OptiX megakernel programs):
//   raygenKernel    __raygen__renderFrame prologue, sv4/deviceProgram.cu:389-475 (seeding, ring cull, primary rays)
//   shadeKernel     __closesthit__radiance + __miss__radiance + the raygen loop body, :251-280, :477-514, :590-703
//   gatherKernel    per-launch-index sums over samples in sample order (:516-517), chunk by chunk
//   finalizeKernel  accumulation + display transform + block replication, :521-587 (orig: :452-474)
//   tonemapSqrtKernel  computeFinalPixelColorsKernel, sv4/toneMap.cu:41-58
// All HBM-streaming / gather bound; compiled with -fmad=false (see shade_math.cuh).
#include "shade_math.cuh"
#include <atomic>
CHK_DECL          // B200PT_CHECKED builds: violation counter of this file (codes 20-39)

#ifndef SHADE_HOIST_THR
#define SHADE_HOIST_THR 1
#endif
#ifndef SHADE_BLOCK
#define SHADE_BLOCK 128
#endif
#ifndef SHADE_MIN_BLOCKS
#define SHADE_MIN_BLOCKS 7          // CTAs per SM for shadeKernel (72 registers); measured per frame: 6: 1.55 ms, 7: 1.44, 8: 1.47, 9: 1.52, 10: 1.56, 12: 1.69
#endif

// warp-aggregated append: one atomicAdd per warp, entries of a warp contiguous and in lane order
DEV unsigned enqueueSlot(bool want, unsigned* counter) {
    const unsigned mask = __ballot_sync(0xffffffffu, want);
    if (!mask) return 0u;
    const unsigned lane = threadIdx.x & 31;
    unsigned base = 0;
    if (lane == (unsigned)(__ffs(mask) - 1)) base = atomicAdd(counter, (unsigned)__popc(mask));
    base = __shfl_sync(0xffffffffu, base, __ffs(mask) - 1);
    return base + __popc(mask & ((1u << lane) - 1u));
}
DEV void enqueueRay(bool want, float4 o, float4 d, float4* qo, float4* qd, unsigned* counter, unsigned cap) {
    const unsigned q = enqueueSlot(want, counter);
    if (want && CHK(q < cap, 20)) { qo[q] = o;  qd[q] = d; }
}
DEV void enqueueShadow(bool want, float4 o, float4 d, float4 p, const Queues& Q, unsigned* counter) {
    const unsigned q = enqueueSlot(want, counter);
    if (want && CHK(q < Q.cap, 21)) { Q.shO[q] = o;  Q.shD[q] = d;  Q.shP[q] = p; }
}
// Both appends of a shade step with ONE 64-bit atomicAdd: the continuation-ray counter and the shadow-ray counter of a launch are neighbours
// (QC_RAYS = 0, QC_SHADOWS = 1 of a 16-byte aligned group), so {rays, shadows << 32} added to the pair returns both bases in one round trip
// (the ray count cannot carry into the high word: it is bounded by the queue capacity).
#ifndef SHADE_PAIR_ATOMIC
#define SHADE_PAIR_ATOMIC 1          // measured on B200: shade 1.222 -> 1.165 ms per frame (one atomic round trip per step instead of two)
#endif
DEV void enqueuePair(bool wantRay, bool wantShadow, float4 o, float4 d, float4 sd, float4 sp, float4* qo, float4* qd, const Queues& Q, unsigned* pairCounter) {
    static_assert(QC_RAYS == 0 && QC_SHADOWS == 1 && QC_STRIDE % 2 == 0, "the two counters must form an aligned 64-bit pair");
    const unsigned mR = __ballot_sync(0xffffffffu, wantRay), mS = __ballot_sync(0xffffffffu, wantShadow);
    if (!(mR | mS)) return;
    const unsigned lane = threadIdx.x & 31, lt = (1u << lane) - 1u;
    unsigned long long old = 0ull;
    if (lane == 0) old = atomicAdd(reinterpret_cast<unsigned long long*>(pairCounter), (unsigned long long)__popc(mR) | ((unsigned long long)__popc(mS) << 32));
    old = __shfl_sync(0xffffffffu, old, 0);
    const unsigned qr = (unsigned)old + __popc(mR & lt), qs = (unsigned)(old >> 32) + __popc(mS & lt);
    if (wantRay && CHK(qr < Q.cap, 20)) { qo[qr] = o;  qd[qr] = d; }
    if (wantShadow && CHK(qs < Q.cap, 21)) { Q.shO[qs] = o;  Q.shD[qs] = sd;  Q.shP[qs] = sp; }
}
// Block-level append with the block's rays grouped by direction octant (SHADE_SORT_OCTANT): the up-to-SHADE_BLOCK continuation rays of one
// grid-stride step leave neighbouring surface points; grouped by the sign octant of their direction, the 32 consecutive queue entries a trace
// warp takes as one chunk also agree on the order in which they visit a node's children.  Deterministic (warp-major order inside an octant);
// a path's result does not depend on its position in the queue, so images are unchanged bit for bit.  Must be reached by every thread of the block.
#ifndef SHADE_SORT_OCTANT
#define SHADE_SORT_OCTANT 0
#endif
DEV void enqueueRayByOctant(bool want, float4 o, float4 d, float4* qo, float4* qd, unsigned* counter, unsigned cap) {
    __shared__ unsigned warpCount[SHADE_BLOCK / 32][8];
    __shared__ unsigned binBase[8];
    __shared__ unsigned blockBase;
    const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const unsigned key = want ? ((d.x < 0.f ? 1u : 0u) | (d.y < 0.f ? 2u : 0u) | (d.z < 0.f ? 4u : 0u)) : 8u;
    if (lane < 8) warpCount[warp][lane] = 0;
    __syncwarp();
    const unsigned peers = __match_any_sync(0xffffffffu, key);
    const unsigned rank = __popc(peers & ((1u << lane) - 1u));
    if (want && rank == 0) warpCount[warp][key] = __popc(peers);
    __syncthreads();
    if (threadIdx.x < 8) {                                      // thread b: bin b's offset inside the block's batch needs every bin's total
        unsigned tot[8];
#pragma unroll
        for (int b = 0; b < 8; ++b) { tot[b] = 0;  for (int w = 0; w < SHADE_BLOCK / 32; ++w) tot[b] += warpCount[w][b]; }
        unsigned run = 0, mine = 0;
#pragma unroll
        for (int b = 0; b < 8; ++b) { if (b == (int)threadIdx.x) mine = run;  run += tot[b]; }
        binBase[threadIdx.x] = mine;
        if (threadIdx.x == 0) blockBase = run ? atomicAdd(counter, run) : 0u;
    }
    __syncthreads();
    if (want) {
        unsigned before = 0;
        for (unsigned w = 0; w < warp; ++w) before += warpCount[w][key];
        const unsigned q = blockBase + binBase[key] + before + rank;
        if (CHK(q < cap, 20)) { qo[q] = o;  qd[q] = d; }
    }
    __syncthreads();                                            // the tables are rewritten by the next step
}

// the 32-byte path record as one 256-bit access (sm_100: LDG.E.256 / STG.E.256)
#ifndef SHADE_STATE_PLAIN
#define SHADE_STATE_PLAIN 0
#endif
#ifndef SHADE_SKIP_ZERO_EMISSION
#define SHADE_SKIP_ZERO_EMISSION 1   // measured on B200: shade 1.161 -> 1.145 ms per frame (64 MB less DRAM traffic at bounce 0)
#endif
DEV PathState ldState(const PathState* p) {
    PathState s;
#if SHADE_STATE_PLAIN
    const uint4 a = *reinterpret_cast<const uint4*>(p), b = *(reinterpret_cast<const uint4*>(p) + 1);
    s.rng1 = a.x;  s.rng2 = a.y;  s.thrX = __uint_as_float(a.z);  s.thrY = __uint_as_float(a.w);
    s.thrZ = __uint_as_float(b.x);  s.eta = __uint_as_float(b.y);  s.state = (int)b.z;  s.bsdfPdf = __uint_as_float(b.w);
#else
    asm volatile("ld.global.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(s.rng1), "=r"(s.rng2), "=f"(s.thrX), "=f"(s.thrY), "=f"(s.thrZ), "=f"(s.eta), "=r"(s.state), "=f"(s.bsdfPdf) : "l"(p) : "memory");
#endif
    return s;
}
DEV void stState(PathState* p, const PathState& s) {
#if SHADE_STATE_PLAIN
    uint4* q = reinterpret_cast<uint4*>(p);
    q[0] = make_uint4(s.rng1, s.rng2, __float_as_uint(s.thrX), __float_as_uint(s.thrY));
    q[1] = make_uint4(__float_as_uint(s.thrZ), __float_as_uint(s.eta), (unsigned)s.state, __float_as_uint(s.bsdfPdf));
#else
    asm volatile("st.global.v8.u32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
                 :: "l"(p), "r"(s.rng1), "r"(s.rng2), "f"(s.thrX), "f"(s.thrY), "f"(s.thrZ), "f"(s.eta), "r"(s.state), "f"(s.bsdfPdf) : "memory");
#endif
}

// launch-index bookkeeping shared by raygen / gather / finalize ------------------------------------------------
struct LaunchIdx { unsigned lx, ly, ix, iy; bool live; unsigned seed0; };

DEV LaunchIdx decodeLaunch(const FrameParams& F, const PassParams& P, unsigned li) {
    LaunchIdx L;
    L.lx = li % (unsigned)P.launchW;  L.ly = li / (unsigned)P.launchW;
    L.seed0 = tea4(L.ly * (unsigned)F.W + L.lx, P.subframe);                     // :407, unscaled index, full-frame w
    L.ix = L.lx * P.factorX + P.offX;  L.iy = L.ly * P.factorY + P.offY;          // :419 (uint wrap-around intended)
    const float dx = (float)L.ix - (float)P.cX, dy = (float)L.iy - (float)P.cY, dz = 0.0f - 0.0f;
    const float range = sqrtf(dx * dx + dy * dy + dz * dz);                       // :421
    L.live = !(range < P.rInner || range > P.rOuter);                             // :423
    if (F.tileCount > 1 && (int)((L.ly / 32u) % (unsigned)F.tileCount) != F.tileRank) L.live = false;
    return L;
}

// primary ray of sample k of a launch index; on return `seed` has been advanced past this sample's jitter
DEV float3 primaryDir(const FrameParams& F, const LaunchIdx& L, unsigned& seed) {
    const float jx = lcgRnd(seed);                                                // :459
    const float jy = lcgRnd(seed);
    const float dx = 2.0f * (((float)L.ix + jx) / (float)F.W) - 1.0f;             // :463-466
    const float dy = 2.0f * (((float)L.iy + jy) / (float)F.H) - 1.0f;
    return normalize3(dx * F.U + dy * F.V + F.Wv);                                // :471
}

DEV unsigned lcgSkip(unsigned seed, int steps) {
    for (int s = 0; s < steps; ++s) seed = 1664525u * seed + 1013904223u;
    return seed;
}

__global__ void __launch_bounds__(SHADE_BLOCK)
raygenKernel(FrameParams F, PassParams P, int k0, int k1, PathPool pool, Queues Q) {
    // Threads walk the launch grid in 8x4-pixel tiles (one warp = one tile when kc = 1), not in scanlines: the rays a warp
    // appends to the queue together — and the bounce rays that inherit their order — then start close to each other.
    const unsigned kc = (unsigned)(k1 - k0);
    const unsigned tilesX = ((unsigned)P.launchW + 7u) >> 3, tilesY = ((unsigned)P.launchH + 3u) >> 2;
    const unsigned total = tilesX * tilesY * 32u * kc;
    const unsigned stride = gridDim.x * blockDim.x;
    for (unsigned base = blockIdx.x * blockDim.x; base < total; base += stride) {     // warp-uniform trip count
        const unsigned t = base + threadIdx.x;
        bool want = false;
        float4 o4 = make_float4(0, 0, 0, 0), d4 = o4;
        const unsigned cell = t / kc, tile = cell >> 5, within = cell & 31u;
        const unsigned lx = (tile % tilesX) * 8u + (within & 7u), ly = (tile / tilesX) * 4u + (within >> 3);
        if (t < total && lx < (unsigned)P.launchW && ly < (unsigned)P.launchH) {
            const unsigned li = ly * (unsigned)P.launchW + lx, k = (unsigned)k0 + t % kc;
            const LaunchIdx L = decodeLaunch(F, P, li);
            if (L.live && CHK(P.slotBase + li * kc + (k - (unsigned)k0) < pool.cap, 22)) {
                const unsigned slot = P.slotBase + li * kc + (k - (unsigned)k0);       // chunk-local slot
                unsigned seed = lcgSkip(L.seed0, 2 * (int)k);
                Rng r;  r.seed(seed);                                             // prd.rand = Random(seed), :445
                const float3 dir = primaryDir(F, L, seed);
                PathState st;                                                     // throughput 1, rayEta 1 (:447-448), depth 0, no flags
                st.rng1 = r.s1;  st.rng2 = r.s2;  st.thrX = st.thrY = st.thrZ = 1.f;  st.eta = 1.f;  st.state = 0;  st.bsdfPdf = 0.f;
                stState(pool.st + slot, st);
                pool.direct[slot] = make_float4(0, 0, 0, 0);
                pool.indirect[slot] = make_float4(0, 0, 0, 0);
                // prd.alpha and the AOV records are NOT zeroed here: every primary ray passes through shadeKernel's first pass exactly once, which writes them
                // whatever it finds (a surface: alpha 1 or, on a shadow catcher, 0, and the hit's normal / albedo; a miss: zeros) - 48 B per sample less to write
                o4 = make_float4(F.eye.x, F.eye.y, F.eye.z, __int_as_float((int)slot));
                d4 = make_float4(dir.x, dir.y, dir.z, 0.f);
                want = true;
            }
        }
        enqueueRay(want, o4, d4, Q.rayO[0], Q.rayD[0], Q.counters + QC_RAYS, Q.cap);
    }
}

// exact-fp32 bilinear fetch, wrap addressing, normalized coordinates (descriptor: sv4/SimplePathtracer.cpp:782-793)
DEV float3 sampleTexture(const SceneDev& S, const MeshInfo* M, float s, float t) {
    const int W = M->texW, H = M->texH;
    const float x = (s - floorf(s)) * (float)W - 0.5f, y = (t - floorf(t)) * (float)H - 0.5f;
    const float fx = floorf(x), fy = floorf(y);
    const float a = x - fx, b = y - fy;
    int i0 = (int)fx, j0 = (int)fy, i1 = i0 + 1, j1 = j0 + 1;
    // wrap: x is in [-0.5, W - 0.5], so i0 is in [-1, W - 1] and i1 in [0, W] (same for j): one conditional add replaces the modulo
    i0 += i0 < 0 ? W : 0;  i1 -= i1 >= W ? W : 0;  j0 += j0 < 0 ? H : 0;  j1 -= j1 >= H ? H : 0;
    if (!CHK((size_t)M->texOffset + (size_t)max(j0, j1) * W + max(i0, i1) < S.nTexels && min(i0, j0) >= 0, 26)) return f3(0.f);
    const uint32_t* px = S.texels + M->texOffset;
    const uint32_t p00 = __ldg(px + (size_t)j0 * W + i0), p10 = __ldg(px + (size_t)j0 * W + i1);
    const uint32_t p01 = __ldg(px + (size_t)j1 * W + i0), p11 = __ldg(px + (size_t)j1 * W + i1);
    float out[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        const float c00 = (float)((p00 >> (8 * c)) & 255u) * (1.0f / 255.0f), c10 = (float)((p10 >> (8 * c)) & 255u) * (1.0f / 255.0f);
        const float c01 = (float)((p01 >> (8 * c)) & 255u) * (1.0f / 255.0f), c11 = (float)((p11 >> (8 * c)) & 255u) * (1.0f / 255.0f);
        out[c] = ((1.0f - a) * (1.0f - b)) * c00 + (a * (1.0f - b)) * c10 + ((1.0f - a) * b) * c01 + (a * b) * c11;
    }
    return f3(out[0], out[1], out[2]);
}

// Shades the hits of trace launch `launch`.
// SHADE_COMPACT 0 (default): one lane per queue entry in a static grid-stride loop; the lanes of rays that left the scene idle
// (__miss__radiance only sets DONE, sv4/deviceProgram.cu:251-280; with mis_on_miss they add the environment term).
// SHADE_COMPACT 1 (measured, not used): warps take tickets of 32 queue entries and squeeze the rays that hit something into dense groups of
// 32 with ballots, keeping their queue order; leftovers carry over to the next ticket.  27 % fewer warp-level passes through the Disney code,
// and yet 1.61 ms per frame against 1.41: the kernel is bound by the latency of its dependent loads (ray -> path record -> probe tables ->
// triangle -> material), not by issue slots, and the ticket + flag + ballot stage in front of every group lengthens exactly that chain.
// (A dense hit queue written by traceKernel itself in retirement order was measured too: shade 1.35 ms, but the queue order - and with it the
// spatial coherence the next bounce's rays inherit - is lost, and traceKernel goes from 2.59 to 2.84 ms.  profiles/README.md, r02a-r02c.)
#ifndef SHADE_COMPACT
#define SHADE_COMPACT 0
#endif
#ifndef SHADE_SMEM_CDFY
#define SHADE_SMEM_CDFY 0         // rows of cdfValuesY staged in shared memory per CTA (0 = off)
#endif
__global__ void __launch_bounds__(SHADE_BLOCK, SHADE_MIN_BLOCKS)
shadeKernel(FrameParams F, SceneDev S, PathPool pool, Queues Q, int cur, int launch) {
    const unsigned count = Q.counters[QC_STRIDE * launch + QC_RAYS];
    const int nxt = cur ^ 1;
    unsigned* nextCounter = Q.counters + QC_STRIDE * (launch + 1) + QC_RAYS;
#if !SHADE_PAIR_ATOMIC || (SHADE_SORT_OCTANT && !SHADE_COMPACT)
    unsigned* shadowCounter = Q.counters + QC_STRIDE * (launch + 1) + QC_SHADOWS;    // (enqueuePair reaches it through nextCounter's 64-bit pair)
#endif
    const float4* __restrict__ rayOq = Q.rayO[cur];
    const float4* __restrict__ rayDq = Q.rayD[cur];
#if SHADE_SMEM_CDFY
    // ProbeSample's row search from shared memory: cdfValuesY (H floats) and its guide table, staged once per CTA when the probe has <= SHADE_SMEM_CDFY rows
    __shared__ float sCdfYBuf[SHADE_SMEM_CDFY];
    __shared__ int sGuideYBuf[B200PT_GUIDE_Y + 1];
    const bool stageY = F.probeGuide != nullptr && F.probe.height <= SHADE_SMEM_CDFY;
    if (stageY) {
        for (int k = threadIdx.x; k < F.probe.height; k += SHADE_BLOCK) sCdfYBuf[k] = __ldg(F.probe.cdfValuesY + k);
        for (int k = threadIdx.x; k <= B200PT_GUIDE_Y; k += SHADE_BLOCK) sGuideYBuf[k] = __ldg(F.probeGuide + k);
    }
    __syncthreads();
    const float* const sCdfY = stageY ? sCdfYBuf : nullptr;
    const int* const sGuideY = stageY ? sGuideYBuf : nullptr;
#else
    const float* const sCdfY = nullptr;
    const int* const sGuideY = nullptr;
#endif
#if SHADE_COMPACT
    __shared__ unsigned list[SHADE_BLOCK / 32][64];
    unsigned* const head = Q.counters + QC_STRIDE * launch + QC_SHEAD;
    const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5, ltMask = (1u << lane) - 1u;
    unsigned fill = 0, ticket = 0;
    bool more = true;
    if (lane == 0) ticket = atomicAdd(head, 32u);
    for (;;) {
        while (fill < 32u && more) {                                          // append the hits of the next ticket's 32 queue entries, in order
            const unsigned base = __shfl_sync(0xffffffffu, ticket, 0);
            if (base >= count) { more = false;  break; }
            const unsigned j = base + lane;
            const bool take = j < count && (F.misOnMiss || __float_as_int(Q.hit[j].w) >= 0);
            if (lane == 0) ticket = atomicAdd(head, 32u);                      // the next ticket is in flight while this group is shaded
            const unsigned m = __ballot_sync(0xffffffffu, take);
            if (take) list[warp][fill + __popc(m & ltMask)] = j;
            fill += __popc(m);
        }
        if (fill == 0u) break;
        __syncwarp();
        const unsigned n = min(fill, 32u);
        const bool mine = lane < n;
        const unsigned i = mine ? list[warp][lane] : 0u;
        const unsigned carry = (fill > 32u && lane < fill - 32u) ? list[warp][32u + lane] : 0u;
        __syncwarp();
        if (fill > 32u && lane < fill - 32u) list[warp][lane] = carry;
        fill -= n;
        __syncwarp();
        {
#else
    const unsigned stride = gridDim.x * blockDim.x;
    for (unsigned base = blockIdx.x * blockDim.x; base < count; base += stride) {
        {
        const unsigned i = base + threadIdx.x;
        const bool mine = i < count;
#endif
        bool wantRay = false, wantShadow = false;
        float4 po = make_float4(0, 0, 0, 0), rd = po, sd = po, sp = po;     // po = {P, slot}: origin record of both the continuation and the shadow ray
        if (mine && CHK(i < Q.cap && (unsigned)__float_as_int(rayOq[i].w) < pool.cap && __float_as_int(Q.hit[i].w) < (int)S.nTris, 23)) {
            const float4 o4 = rayOq[i], d4 = rayDq[i], h = Q.hit[i];         // {o, slot} {d, -} {t, u, v, triSlot}
            const unsigned slot = (unsigned)__float_as_int(o4.w);
            const int triSlot = __float_as_int(h.w);
            PathState st = ldState(pool.st + slot);
            int depth = st.state & 0xff, flags = st.state >> 8;
            const int segDepth = depth;                                           // prd.depth while this segment is shaded
            const float3 org = f3(o4.x, o4.y, o4.z), dir = f3(d4.x, d4.y, d4.z);
            if (triSlot < 0) {
                flags |= PF_DONE;                                                 // __miss__radiance, :251-280 (shaded only with mis_on_miss)
                if (depth == 0) {                                                 // a primary miss: the records raygen leaves to this pass (alpha 0, no AOV features)
                    pool.alpha[slot] = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (F.writeAovs) { pool.aovNormal[slot] = make_float4(0.f, 0.f, 0.f, 0.f);  pool.aovAlbedo[slot] = make_float4(0.f, 0.f, 0.f, 0.f); }
                }
                if (F.misOnMiss && (flags & PF_SECONDARY)) {                      // opt-in extension (:257-267), same rule as oracle/render_core.h
                    const float skyPdf = probePdf(F.probe, dir);
                    const float bp = st.bsdfPdf;
                    const float weight = 0.5f * bp / (0.5f * bp + 0.5f * skyPdf);
                    const float4 sky = probeEval(F.probe, probeDirToUV(dir));
                    const float3 add = weight * f3(sky.x, sky.y, sky.z) * f3(st.thrX, st.thrY, st.thrZ);
                    float4 acc = pool.indirect[slot];
                    acc.x += add.x;  acc.y += add.y;  acc.z += add.z;
                    pool.indirect[slot] = acc;
                }
            } else {
                const float4 ta = __ldg(S.tris + 3 * (size_t)triSlot), tb = __ldg(S.tris + 3 * (size_t)triSlot + 1),
                             tc = __ldg(S.tris + 3 * (size_t)triSlot + 2);
                // The probe sample of this vertex (SampleLights, :301-384) depends only on the path's RNG state: start its chain of
                // dependent table lookups now, while the triangle / mesh / material loads above are still in flight.  The RNG state
                // is committed further down, only where the reference really draws the sample.
                Rng rng;  rng.s1 = st.rng1;  rng.s2 = st.rng2;
                float3 wi, skyColor;  float skyPdf;
                probeSample(F.probe, F.probeGuide, wi, skyColor, skyPdf, rng, sCdfY, sGuideY);
                const float3 e1 = f3(ta.w, tb.x, tb.y), e2 = f3(tb.z, tb.w, tc.x);
                const unsigned gid = (unsigned)__float_as_int(tc.y);
                const MeshInfo* M = S.meshes + (CHK((unsigned)__float_as_int(tc.z) < S.nMeshes && gid < S.nTris, 24) ? __float_as_int(tc.z) : 0);
                const Material& mat = M->material;
                const float3 N0 = normalize3(cross3(e1, e2));                     // :603
                const float3 wo = -dir;
                const float3 N = N0 * copysignf(1.0f, dot3(wo, N0));              // faceforward, :605
                const float3 Pp = org + h.x * dir;                                // :609
                const bool catcher = (mat.flags & MATERIAL_FLAG_SHADOW_CATCHER) != 0;
                if (catcher && (flags & PF_SECONDARY)) {                          // pass-through, :617-622 (--depth there, ++depth in the raygen loop: depth stays)
                    wantRay = true;
                    po = make_float4(Pp.x, Pp.y, Pp.z, o4.w);  rd = make_float4(d4.x, d4.y, d4.z, 0.f);
                } else {
                    po = make_float4(Pp.x, Pp.y, Pp.z, o4.w);
                    float3 albedo = mat.color;                                    // :625-641
                    if (M->textureID >= 0 && M->hasTexcoord) {
                        const unsigned* ip = reinterpret_cast<const unsigned*>(S.indices + gid);
                        const uint3 id = make_uint3(__ldg(ip), __ldg(ip + 1), __ldg(ip + 2));
                        const bool tcOk = CHK(M->firstVertex + max(id.x, max(id.y, id.z)) < S.nVertices, 25);      // (a violation reads vertex 0 instead)
                        const unsigned fv = tcOk ? M->firstVertex : 0u;
                        const float2 t0 = __ldg(S.texcoords + fv + (tcOk ? id.x : 0u)), t1 = __ldg(S.texcoords + fv + (tcOk ? id.y : 0u)),
                                     t2 = __ldg(S.texcoords + fv + (tcOk ? id.z : 0u));
                        const float w0 = 1.f - h.y - h.z;
                        albedo = sampleTexture(S, M, w0 * t0.x + h.y * t1.x + h.z * t2.x, w0 * t0.y + h.y * t1.y + h.z * t2.y);
                    }
                    if (depth == 0 && F.writeAovs) {
                        pool.aovNormal[slot] = make_float4(N.x, N.y, N.z, 0.f);
                        pool.aovAlbedo[slot] = make_float4(albedo.x, albedo.y, albedo.z, 0.f);
                    }
                    float3 T = f3(st.thrX, st.thrY, st.thrZ);
                    float rayEta = st.eta;
                    const float outEta = rayEta == 1.0f ? (mat.eta == 0.0f ? 2.0f / (1.0f - sqrtf(0.08f * mat.specular)) - 1.0f : mat.eta) : 1.0f;   // :644-654
                    const bool last = depth >= F.maxDepth;                        // its radiance is dropped at :495
                    // prd.alpha = 1 at every non-catcher hit (:660).  Once it is 1 nothing can change it again (only a PRIMARY catcher hit emits an
                    // alpha-targeted shadow ray, and that one resolves before this pass), so the 16-byte scattered store happens once per path.
                    if (!catcher && !(flags & PF_ALPHA_ONE)) { pool.alpha[slot] = make_float4(1.f, 1.f, 1.f, 0.f);  flags |= PF_ALPHA_ONE; }
                    else if (catcher && depth == 0) pool.alpha[slot] = make_float4(0.f, 0.f, 0.f, 0.f);   // a primary catcher hit starts from prd.alpha = 0 (:441; raygen does not write it)
                    if (last && !catcher) {
                        flags |= PF_DONE;
                    } else {
                        // The continuation direction is drawn first (the RNG order is the reference's: the probe sample was drawn above,
                        // DisneySample comes next, and nothing in between consumes random numbers) ...
                        // Fr(N.wo) and the two index ratios are the same in DisneySample, DisneyPdf and DisneyEval of this vertex: once
                        const EtaPair etaPair = makeEtaPair(rayEta, outEta);
                        const float FNV = fresnelDielectric(dot3(N, wo), etaPair);
                        float3 bsdfDir = f3(0.f);  float pdf = 0.0f;  int type = 0;  bool dirNeedsPdf = false;
                        if (!last) {
                            float3 u, v;
                            basisFromVector(N, u, v);                             // :671-677
                            bsdfSampleDir(mat, rayEta, outEta, FNV, u, v, N, wo, bsdfDir, pdf, type, dirNeedsPdf, rng);
                        }
                        // ... so that DisneyPdf / DisneyEval of the two directions of this vertex — k = 0: the NEE direction wi (SampleLights,
                        // :301-384), k = 1: the drawn direction (:683-695) — run through ONE rolled copy of the code.
                        float bp = 0.0f;  float3 fNee = f3(0.f), fDir = f3(0.f);  bool evalNee = false;
#pragma unroll 1
                        for (int k = 0; k < 2; ++k) {
                            const float3 L = k ? bsdfDir : wi;
                            float p = pdf;
                            if (k == 0 || dirNeedsPdf) p = bsdfPdf(mat, FNV, N, wo, L);
                            bool ev;
                            if (k == 0) ev = p > 0.0f && 0.5f * skyPdf / (0.5f * p + 0.5f * skyPdf) > 0.0f;
                            else ev = !last && p > 0.0f;
                            float3 f = f3(0.f);
                            if (ev) f = bsdfEval(mat, albedo, etaPair, FNV, N, wo, L);
                            if (k == 0) { bp = p;  fNee = f;  evalNee = ev; }
                            else { pdf = p;  fDir = f; }
                        }
                        float3 pend = f3(0.f);
                        bool valid = false;
                        if (evalNee) {
                            const float weight = 0.5f * skyPdf / (0.5f * bp + 0.5f * skyPdf);
                            const float3 val = weight * skyColor * fNee * fabsf(dot3(wi, N)) / skyPdf * 1.0f;
                            pend = T * val;
                            valid = !(pend.x == 0.f && pend.y == 0.f && pend.z == 0.f);
                        }
                        if (last) {                                               // catcher hit by the primary ray at max depth
                            flags |= PF_DONE;
                        } else {
                            if (pdf <= 0.0f) {
                                // :679-682: the raygen loop breaks before it adds this segment's prd.radiance (:495-507), so the NEE term is dropped;
                                // prd.alpha of a catcher hit was already updated by SampleShadow (:663-664) and survives
                                flags |= PF_DONE;  valid = valid && catcher;
                            } else {
                                // primary-only emission, :667-669.  x + (+-0) == x bit for bit for every x the sum can hold (it never holds -0: it starts
                                // at +0, and a round-to-nearest sum is -0 only if both terms are), so an all-zero emission needs no read-modify-write of the record
                                if (!(flags & PF_SECONDARY) && (!SHADE_SKIP_ZERO_EMISSION || mat.emission.x != 0.0f || mat.emission.y != 0.0f || mat.emission.z != 0.0f)) {
                                    float4 dacc = pool.direct[slot];
                                    dacc.x += mat.emission.x;  dacc.y += mat.emission.y;  dacc.z += mat.emission.z;
                                    pool.direct[slot] = dacc;
                                }
                                const float3 f = fDir;
                                if (dot3(bsdfDir, N) <= 0.0f) rayEta = outEta;    // :688-692
                                T = T * (f * fabsf(dot3(N, bsdfDir)) / pdf);      // :695
                                flags |= PF_SECONDARY;
                                bool rrKill = false;
                                if (F.russianRoulette && depth + 1 >= 2) {        // opt-in extension, same rule as oracle/render_core.h
                                    const float p = fminf(1.0f, fmaxf(T.x, fmaxf(T.y, T.z)));
                                    const float r = rng.randf();
                                    if (!(r < p)) rrKill = true;
                                    else T = T / p;
                                }
                                st.thrX = T.x;  st.thrY = T.y;  st.thrZ = T.z;  st.eta = rayEta;
                                st.rng1 = rng.s1;  st.rng2 = rng.s2;
                                st.bsdfPdf = pdf;
                                const int nd = depth + 1;
                                if (rrKill || (nd >= F.maxDepth && (flags & PF_ALPHA_ONE) && !F.misOnMiss)) flags |= PF_DONE;   // roulette, or the next segment could only be dropped
                                else { wantRay = true;  rd = make_float4(bsdfDir.x, bsdfDir.y, bsdfDir.z, 0.f); }
                                depth = nd;
                            }
                        }
                        if (valid) {
                            // target: 0 = directLight (:502-504), 1 = indirectLight, 2 = alpha if OCCLUDED (SampleShadow, :663-664)
                            wantShadow = true;
                            const int target = catcher ? 2 : (segDepth == 0 ? 0 : 1);
                            sd = make_float4(wi.x, wi.y, wi.z, __int_as_float(target));
                            sp = make_float4(pend.x, pend.y, pend.z, 0.f);
                        }
                    }
                }
            }
            // the record is read again only by the pass that shades this path's next ray: a path that ends here (miss, depth cap, pdf <= 0, roulette) leaves none
            if (wantRay) { st.state = (depth & 0xff) | (flags << 8);  stState(pool.st + slot, st); }
        }
#if SHADE_SORT_OCTANT && !SHADE_COMPACT
        enqueueRayByOctant(wantRay, po, rd, Q.rayO[nxt], Q.rayD[nxt], nextCounter, Q.cap);
        enqueueShadow(wantShadow, po, sd, sp, Q, shadowCounter);
#elif SHADE_PAIR_ATOMIC
        enqueuePair(wantRay, wantShadow, po, rd, sd, sp, Q.rayO[nxt], Q.rayD[nxt], Q, nextCounter);
#else
        enqueueRay(wantRay, po, rd, Q.rayO[nxt], Q.rayD[nxt], nextCounter, Q.cap);
        enqueueShadow(wantShadow, po, sd, sp, Q, shadowCounter);
#endif
        }
    }
}

// samples k0..k1-1 of launch index li added to the running sums, in sample order (result += direct + indirect, :516-517)
DEV void gatherSamples(const FrameParams& F, const PassParams& P, int k0, int k1, const PathPool& pool, unsigned li,
                       float3& result, float3& alpha, float3& normal, float3& albedo) {
    for (int k = k0; k < k1; ++k) {
        const unsigned slot = P.slotBase + li * (unsigned)(k1 - k0) + (unsigned)(k - k0);
        if (!CHK(slot < pool.cap, 27)) break;
        const float4 d = pool.direct[slot], i = pool.indirect[slot], a = pool.alpha[slot];
        result = result + (f3(d.x, d.y, d.z) + f3(i.x, i.y, i.z));
        alpha = alpha + f3(a.x, a.y, a.z);
        if (F.writeAovs) {
            const float4 nn = pool.aovNormal[slot], aa = pool.aovAlbedo[slot];
            normal = normal + f3(nn.x, nn.y, nn.z);  albedo = albedo + f3(aa.x, aa.y, aa.z);
        }
    }
}

// Per-launch-index sums over the samples of one chunk, in sample order (result += direct+indirect, :516-517).
__global__ void __launch_bounds__(SHADE_BLOCK)
gatherKernel(FrameParams F, PassParams P, int k0, int k1, PathPool pool, float4* sumRad, float4* sumAlpha,
             float4* sumNormal, float4* sumAlbedo) {
    const unsigned n = (unsigned)P.launchW * (unsigned)P.launchH;
    const unsigned li = blockIdx.x * blockDim.x + threadIdx.x;
    if (li >= n) return;
    const LaunchIdx L = decodeLaunch(F, P, li);
    if (!L.live) return;
    const unsigned rec = P.launchBase + li;
    float3 result = f3(0.f), alpha = f3(0.f), normal = f3(0.f), albedo = f3(0.f);
    if (k0 > 0) {
        const float4 r = sumRad[rec], a = sumAlpha[rec];
        result = f3(r.x, r.y, r.z);  alpha = f3(a.x, a.y, a.z);
        if (F.writeAovs) { const float4 nn = sumNormal[rec], aa = sumAlbedo[rec];  normal = f3(nn.x, nn.y, nn.z);  albedo = f3(aa.x, aa.y, aa.z); }
    }
    gatherSamples(F, P, k0, k1, pool, li, result, alpha, normal, albedo);
    sumRad[rec] = make_float4(result.x, result.y, result.z, 0.f);
    sumAlpha[rec] = make_float4(alpha.x, alpha.y, alpha.z, 0.f);
    if (F.writeAovs) { sumNormal[rec] = make_float4(normal.x, normal.y, normal.z, 0.f);  sumAlbedo[rec] = make_float4(albedo.x, albedo.y, albedo.z, 0.f); }
}

DEV uint32_t displayColor(const FrameParams& F, float3 accum) {
    if (F.display == B200PT_DISPLAY_REINHARD) return makeColor(reinhard(accum * F.exposure, 1.0f));   // :557,568
    return makeColor(accum);                                                                            // orig :467
}

// Accumulation + display transform + fillSize^2 block replication (:521-587).  36 B/pixel streaming.
__global__ void __launch_bounds__(SHADE_BLOCK)
finalizeKernel(FrameParams F, PassParams P, const float4* sumRad, const float4* sumAlpha, const float4* sumNormal,
               const float4* sumAlbedo, FusedPools fused) {
    const unsigned n = (unsigned)P.launchW * (unsigned)P.launchH;
    const unsigned li = blockIdx.x * blockDim.x + threadIdx.x;
    if (li >= n) return;
    const LaunchIdx L = decodeLaunch(F, P, li);
    if (!L.live) return;
    const unsigned rec = P.launchBase + li;
    const float spl = (float)P.spl;
    // backplate of the LAST sample survives the loop (:475, :538)
    unsigned seed = lcgSkip(L.seed0, 2 * (P.spl - 1));
    const float3 lastDir = primaryDir(F, L, seed);
    const float4 bp4 = probeEval(F.probe, probeDirToUV(lastDir));
    const float3 backplate = f3(bp4.x, bp4.y, bp4.z);
    // The per-launch-index sums: read back from gatherKernel's records, or - a frame whose samples sit in ONE chunk of the path pool (fused) - summed here
    // straight from the pool in the same order, which saves the records' round trip through HBM (64 B written + 64 B read per launch index, 128 with AOVs)
    float3 result = f3(0.f), alphaSum = f3(0.f), normal = f3(0.f), albedo = f3(0.f);
    if (fused.n > 0) {
        int set = 0;
        if (fused.n > 1) {                                                            // the band set that rendered this launch row (enqueuePart's partition)
            const unsigned tc = (unsigned)max(F.tileCount, 1), tr = F.tileCount > 1 ? (unsigned)F.tileRank : 0u;
            set = (int)((((L.ly / 32u) % (tc * (unsigned)fused.n)) - tr) / tc);
        }
        gatherSamples(F, P, 0, P.spl, fused.p[set], li, result, alphaSum, normal, albedo);
    }
    else {
        const float4 r4 = sumRad[rec], a4 = sumAlpha[rec];
        result = f3(r4.x, r4.y, r4.z);  alphaSum = f3(a4.x, a4.y, a4.z);
        if (F.writeAovs) { const float4 nn = sumNormal[rec], aa = sumAlbedo[rec];  normal = f3(nn.x, nn.y, nn.z);  albedo = f3(aa.x, aa.y, aa.z); }
    }
    const float3 alpha = alphaSum / spl;                                              // :523
    if (F.writeAovs) { normal = normal / spl;  albedo = albedo / spl; }               // :521-522
    const float3 color = (backplate * spl) * (f3(1.0f) - alpha) + result;             // :538
    const float3 fresh = color / spl;                                                 // :540
    for (int i = 0; i < P.fillSize; ++i) {
        for (int j = 0; j < P.fillSize; ++j) {
            unsigned ix = L.lx * P.factorX + (unsigned)i + P.offX;
            unsigned iy = L.ly * P.factorY + (unsigned)j + P.offY;
            ix = min(ix, (unsigned)(F.W - 1));  iy = min(iy, (unsigned)(F.H - 1));      // :534
            const unsigned image_index = iy * (unsigned)F.W + ix;
            if (!CHK(image_index < (unsigned)F.W * (unsigned)F.H, 28)) continue;
            float3 acc = fresh;
            if (F.accumMode == B200PT_ACCUM_SUM) {
                const float3 c = P.subframe > 0 ? clamp3(acc, 0.0f, 10.0f) : acc;
                const float4 prev = F.accum[image_index];
                F.accum[image_index] = make_float4(prev.x + c.x, prev.y + c.y, prev.z + c.z, prev.w + 1.0f);
                continue;
            }
            if (P.subframe > 0 && !P.redraw) {                                          // :545-552
                acc = clamp3(acc, 0.0f, 10.0f);
                const float a = 1.0f / (float)(P.subframe + 1);
                const float4 prev = F.accum[image_index];
                acc = lerp3(f3(prev.x, prev.y, prev.z), acc, a);
            }
            F.accum[image_index] = make_float4(acc.x, acc.y, acc.z, 1.0f);              // :553
            F.frame[image_index] = displayColor(F, acc);
            if (F.writeAovs) {                                                          // orig :472-474
                F.normal[image_index] = make_float4(normal.x, normal.y, normal.z, 1.0f);
                F.color[image_index] = make_float4(acc.x, acc.y, acc.z, 1.0f);
                F.albedo[image_index] = make_float4(albedo.x, albedo.y, albedo.z, 1.0f);
            }
        }
    }
}

// ACCUM_SUM epilogue (multi-GPU sample split, after the NCCL reduce): mean + display transform.
__global__ void resolveSumKernel(FrameParams F, int nPasses) {
    const unsigned n = (unsigned)F.W * (unsigned)F.H;
    const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 s = F.accum[i];
    const float3 mean = f3(s.x, s.y, s.z) / (float)nPasses;
    F.accum[i] = make_float4(mean.x, mean.y, mean.z, 1.0f);
    F.frame[i] = displayColor(F, mean);
}

// the same from a separate buffer of totals, refreshing only the frame buffer (progressive display of a sample split: the accumulation
// buffer keeps this rank's partial sum)
__global__ void resolveTotalKernel(FrameParams F, const float4* __restrict__ total, int nPasses) {
    const unsigned n = (unsigned)F.W * (unsigned)F.H;
    const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 s = total[i];
    F.frame[i] = displayColor(F, f3(s.x, s.y, s.z) / (float)nPasses);
}

// computeFinalPixelColorsKernel, sv4/toneMap.cu:41-58.  20 B/pixel streaming; one thread per pixel, 1-D
// (the reference's 32x32 tiling buys nothing for a pure stream).
__global__ void tonemapSqrtKernel(uint32_t* __restrict__ out, const float4* __restrict__ in, unsigned n) {
    const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 f = in[i];
    const float x = fminf(1.f, fmaxf(0.f, sqrtf(f.x))), y = fminf(1.f, fmaxf(0.f, sqrtf(f.y))), z = fminf(1.f, fmaxf(0.f, sqrtf(f.z)));
    out[i] = ((uint32_t)(x * 255.9f)) | ((uint32_t)(y * 255.9f) << 8) | ((uint32_t)(z * 255.9f) << 16) | (255u << 24);
}

// ---- parity hook: the first radiance ray of sample 0 for every launch index ---------------------------------
__global__ void primaryRaysKernel(FrameParams F, PassParams P, float4* rayO, float4* rayD) {
    const unsigned n = (unsigned)P.launchW * (unsigned)P.launchH;
    const unsigned li = blockIdx.x * blockDim.x + threadIdx.x;
    if (li >= n) return;
    const LaunchIdx L = decodeLaunch(F, P, li);
    unsigned seed = L.seed0;
    const float3 dir = primaryDir(F, L, seed);
    rayO[li] = make_float4(F.eye.x, F.eye.y, F.eye.z, L.live ? 1.f : 0.f);
    rayD[li] = make_float4(dir.x, dir.y, dir.z, 0.f);
}

__global__ void resolvePrimaryKernel(SceneDev S, const float4* hit, unsigned n, const float4* rayO, int32_t* prim, int32_t* mesh,
                                     float* t, float* u, float* v) {
    const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 h = hit[i];
    const int slot = __float_as_int(h.w);
    const bool live = rayO == nullptr || rayO[i].w != 0.f;
    if (!live || slot < 0) { prim[i] = -1;  mesh[i] = live ? -1 : -2;  t[i] = 0.f;  u[i] = 0.f;  v[i] = 0.f;  return; }
    const float4 c = S.tris[3 * (size_t)slot + 2];
    const int m = __float_as_int(c.z);
    mesh[i] = m;  prim[i] = (int)((unsigned)__float_as_int(c.y) - S.meshes[m].firstTri);
    t[i] = h.x;  u[i] = h.y;  v[i] = h.z;
}

// guide tables for probeSample (shade_math.cuh): [0, GY] for cdfValuesY, then GX + 1 entries per row of cdfValuesX
__global__ void probeGuideKernel(const float* __restrict__ cdfX, const float* __restrict__ cdfY, int W, int H, int* __restrict__ guide) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t nY = B200PT_GUIDE_Y + 1, nX = (size_t)H * (B200PT_GUIDE_X + 1);
    if (i < nY) guide[i] = lowerBound(cdfY, 0, H, (float)i / (float)B200PT_GUIDE_Y);
    else if (i < nY + nX) {
        const int row = (int)((i - nY) / (B200PT_GUIDE_X + 1)), k = (int)((i - nY) % (B200PT_GUIDE_X + 1));
        guide[i] = lowerBound(cdfX, row * W, (row + 1) * W, (float)k / (float)B200PT_GUIDE_X) - row * W;
    }
}
size_t probeGuideSize(int H) { return (size_t)(B200PT_GUIDE_Y + 1) + (size_t)H * (B200PT_GUIDE_X + 1); }
void launchProbeGuide(const float* cdfX, const float* cdfY, int W, int H, int* guide, cudaStream_t s) {
    const size_t n = probeGuideSize(H);
    probeGuideKernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(cdfX, cdfY, W, H, guide);
}

// ---- ProbeData::BuildCDF on the device (sv4/Probe.h:29-77) ---------------------------------------------------
// The reference accumulates each row's luminance SEQUENTIALLY in fp32, and that order defines the bits of the tables.  One warp
// owns one row and keeps that order: 32 texels are loaded coalesced, then the running total walks over the 32 lanes' weights in
// lane order (every lane carries the same total; lane k keeps the value after its own weight).  Rows are independent, so the
// image is processed H rows at a time; the column pass is the same walk over the H row totals by a single warp.
__global__ void probeRowCdfKernel(const float4* __restrict__ data, int W, int H, float* __restrict__ pdfX, float* __restrict__ cdfX,
                                  float* __restrict__ rowTotal) {
    const int row = (int)((blockIdx.x * blockDim.x + threadIdx.x) >> 5), lane = threadIdx.x & 31;
    if (row >= H) return;
    const float4* src = data + (size_t)row * W;
    float* p = pdfX + (size_t)row * W;
    float* c = cdfX + (size_t)row * W;
    float total = 0.0f;                                                    // totalWeightX, Probe.h:41
    for (int base = 0; base < W; base += 32) {
        const int i = base + lane, n = min(32, W - base);
        float wgt = 0.0f;
        if (i < W) { const float4 t = src[i];  wgt = t.x * 0.3f + t.y * 0.6f + t.z * 0.1f; }           // Luminance, maths.h:165-168 (no FMA: -fmad=false)
        float mine = 0.0f;
#pragma unroll
        for (int k = 0; k < 32; ++k) {
            const float v = __shfl_sync(0xffffffffu, wgt, k);
            if (k < n) total = total + v;                                  // :47
            if (k == lane) mine = total;
        }
        if (i < W) { p[i] = wgt;  c[i] = mine; }                           // :49-50
    }
    const float inv = 1.0f / total;                                        // :53
    for (int i = lane; i < W; i += 32) { p[i] = p[i] * inv;  c[i] = c[i] * inv; }     // :56-60 (each lane re-reads what it wrote)
    if (lane == 0) rowTotal[row] = total;
}
__global__ void probeColCdfKernel(const float* __restrict__ rowTotal, int H, float* __restrict__ pdfY, float* __restrict__ cdfY) {
    const int lane = threadIdx.x & 31;
    float total = 0.0f;                                                    // totalWeightY, Probe.h:36
    for (int base = 0; base < H; base += 32) {
        const int j = base + lane, n = min(32, H - base);
        const float wgt = j < H ? rowTotal[j] : 0.0f;
        float mine = 0.0f;
#pragma unroll
        for (int k = 0; k < 32; ++k) {
            const float v = __shfl_sync(0xffffffffu, wgt, k);
            if (k < n) total = total + v;                                  // :63
            if (k == lane) mine = total;
        }
        if (j < H) { pdfY[j] = wgt;  cdfY[j] = mine; }                     // :65-66
    }
    for (int j = lane; j < H; j += 32) { cdfY[j] = cdfY[j] / total;  pdfY[j] = pdfY[j] / total; }   // :70-74
}
void launchProbeCdf(const float4* data, int W, int H, float* pdfX, float* cdfX, float* pdfY, float* cdfY, float* rowTotal, cudaStream_t s) {
    if (W <= 0 || H <= 0) return;
    probeRowCdfKernel<<<(H + 3) / 4, 128, 0, s>>>(data, W, H, pdfX, cdfX, rowTotal);
    probeColCdfKernel<<<1, 32, 0, s>>>(rowTotal, H, pdfY, cdfY);
}

// ---- the same BuildCDF without any allocation (b200pt_build_cdf_noalloc, no reference counterpart) ----
// The only scratch of the two kernels above is the per-row total (H floats).  Here it lives in a file-scope device array of a fixed maximum
// size, so callers that must not allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc) can build the tables:
//   * every GPU that loads the library holds these B200PT_NOALLOC_MAX_ROWS * 4 bytes = 64 KiB until the process exits;
//   * the array is shared by every call on that GPU and keeps the last call's contents.  Nothing has to start at zero: probeRowCdfKernel writes
//     rowTotal[0, H) in full before probeColCdfKernel (next on the same stream) reads exactly that range, stale entries beyond H are never read;
//   * two calls that overlap on the GPU (different streams, or different contexts on one device) race on the array: serialise them.
// In host code `g_noallocRowTotal` is only a host-side stand-in for the symbol, NOT a device address; the device address is resolved with
// cudaGetSymbolAddress once per device, with that device current, and that pointer is what the kernels get.
static __device__ __align__(256) float g_noallocRowTotal[B200PT_NOALLOC_MAX_ROWS];

static float* noallocRowTotalAddress() {                     // the current device's copy of the array
    static float* cache[64];                                 // one slot per device ordinal; 0 = not resolved yet
    int dev = 0;
    CUDA_OK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) throw std::runtime_error("b200pt_build_cdf_noalloc: device ordinal out of range");
    if (!cache[dev]) {
        float* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress((void**)&p, g_noallocRowTotal);
        if (e != cudaSuccess || !p) throw std::runtime_error(std::string("b200pt_build_cdf_noalloc: cudaGetSymbolAddress failed: ") + cudaGetErrorString(e));
        cache[dev] = p;
    }
    return cache[dev];
}
void launchProbeCdfNoAlloc(const float4* data, int W, int H, float* pdfX, float* cdfX, float* pdfY, float* cdfY, cudaStream_t s) {
    if (H > B200PT_NOALLOC_MAX_ROWS) throw std::runtime_error("b200pt_build_cdf_noalloc: height exceeds B200PT_NOALLOC_MAX_ROWS");
    launchProbeCdf(data, W, H, pdfX, cdfX, pdfY, cdfY, noallocRowTotalAddress(), s);
}

void checkedReportShade(unsigned* count, unsigned* first) {
    *count = 0;  *first = 0;
#ifdef B200PT_CHECKED
    cudaMemcpyFromSymbol(count, g_chkCount, sizeof(unsigned));  cudaMemcpyFromSymbol(first, g_chkFirst, sizeof(unsigned));
#endif
}

// ---- launchers ---------------------------------------------------------------------------------------------
static inline unsigned gridFor(unsigned n) { return (n + SHADE_BLOCK - 1) / SHADE_BLOCK; }
static inline unsigned persistentGrid() {
    static const int ctasPerSM = [] { const char* e = getenv("B200PT_SHADE_CTAS");  const int v = e ? atoi(e) : SHADE_MIN_BLOCKS;  return v < 1 ? 1 : (v > 16 ? 16 : v); }();   // tuning knob (see launchTrace)
    return (unsigned)numSMs() * ctasPerSM;
}

void launchRaygen(const FrameParams& F, const PassParams& P, int k0, int k1, const PathPool& pool, const Queues& Q, cudaStream_t s) {
    raygenKernel<<<persistentGrid(), SHADE_BLOCK, 0, s>>>(F, P, k0, k1, pool, Q);
}
void launchShade(const FrameParams& F, const SceneDev& scene, const PathPool& pool, const Queues& Q, int cur, int depth, cudaStream_t s) {
    // the same shared-memory carve-out as traceKernel (see launchTrace: kernels of frames in flight share an SM only with equal configurations); the
    // kernel's own time does not depend on it (1.143-1.151 ms per frame from 0 to 28 %)
    static const int carveout = [] { const char* e = getenv("B200PT_SHADE_CARVEOUT");  return e ? atoi(e) : B200PT_SMEM_CARVEOUT; }();
    static std::atomic<unsigned long long> carved{0ull};
    int dev = 0;
    if (cudaGetDevice(&dev) == cudaSuccess && dev >= 0 && dev < 64 && !(carved.load() >> dev & 1ull)) {
        cudaFuncSetAttribute(shadeKernel, cudaFuncAttributePreferredSharedMemoryCarveout, carveout);
        carved.fetch_or(1ull << dev);
    }
    shadeKernel<<<persistentGrid(), SHADE_BLOCK, 0, s>>>(F, scene, pool, Q, cur, depth);
}
void launchGather(const FrameParams& F, const PassParams& P, int k0, int k1, const PathPool& pool, float4* sumRad, float4* sumAlpha,
                  float4* sumNormal, float4* sumAlbedo, cudaStream_t s) {
    gatherKernel<<<gridFor((unsigned)P.launchW * P.launchH), SHADE_BLOCK, 0, s>>>(F, P, k0, k1, pool, sumRad, sumAlpha, sumNormal, sumAlbedo);
}
void launchFinalize(const FrameParams& F, const PassParams& P, const float4* sumRad, const float4* sumAlpha, const float4* sumNormal,
                    const float4* sumAlbedo, const FusedPools& fused, cudaStream_t s) {
    finalizeKernel<<<gridFor((unsigned)P.launchW * P.launchH), SHADE_BLOCK, 0, s>>>(F, P, sumRad, sumAlpha, sumNormal, sumAlbedo, fused);
}
void launchResolveSum(const FrameParams& F, int nPasses, cudaStream_t s) {
    resolveSumKernel<<<gridFor((unsigned)F.W * F.H), SHADE_BLOCK, 0, s>>>(F, nPasses);
}
void launchResolveTotal(const FrameParams& F, const float4* total, int nPasses, cudaStream_t s) {
    resolveTotalKernel<<<gridFor((unsigned)F.W * F.H), SHADE_BLOCK, 0, s>>>(F, total, nPasses);
}
void launchTonemapSqrt(int w, int h, const float4* in, uint32_t* out, cudaStream_t s) {
    const unsigned n = (unsigned)w * (unsigned)h;
    if (n) tonemapSqrtKernel<<<(n + 255) / 256, 256, 0, s>>>(out, in, n);
}
void launchPrimaryRays(const FrameParams& F, const PassParams& P, float4* rayO, float4* rayD, cudaStream_t s) {
    primaryRaysKernel<<<gridFor((unsigned)P.launchW * P.launchH), SHADE_BLOCK, 0, s>>>(F, P, rayO, rayD);
}
void launchResolvePrimary(const SceneDev& scene, const float4* hit, unsigned n, const float4* rayO, int32_t* prim, int32_t* mesh,
                          float* t, float* u, float* v, cudaStream_t s) {
    if (n) resolvePrimaryKernel<<<gridFor(n), SHADE_BLOCK, 0, s>>>(scene, hit, n, rayO, prim, mesh, t, u, v);
}

// ---- parity hooks: the device shading library on caller-supplied inputs (tests/golden/bsdf.npz, probe.npz) -----------------------
// One thread per item; the same inlined functions shadeKernel calls, in the reference's call order:
//   BasisFromVector (maths.h:94-108), BSDFSample with a fresh Random(seed) (Disney.cuh:197-315), BSDFEval / BSDFPdf for a given L (:152-193, :318-427).
__global__ void debugBsdfKernel(unsigned n, const Material* __restrict__ mats, const float* __restrict__ etaI, const float* __restrict__ etaO,
                                const float3* __restrict__ N, const float3* __restrict__ view, const float3* __restrict__ L, const float3* __restrict__ albedo,
                                const unsigned* __restrict__ seeds, float3* basisU, float3* basisV, float3* sampleL, float* samplePdf, int* sampleType,
                                uint2* sampleRng, float3* evalF, float* evalPdf) {
    const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const Material mat = mats[i];
    float3 u, v;
    basisFromVector(N[i], u, v);
    basisU[i] = u;  basisV[i] = v;
    Rng r;  r.seed(seeds[i]);
    float3 l = f3(0.f);  float pdf = 0.f;  int type = eReflected;
    bsdfSample(mat, etaI[i], etaO[i], u, v, N[i], view[i], l, pdf, type, r);
    sampleL[i] = l;  samplePdf[i] = pdf;  sampleType[i] = type;  sampleRng[i] = make_uint2(r.s1, r.s2);
    evalF[i] = bsdfEval(mat, albedo[i], etaI[i], etaO[i], N[i], view[i], L[i]);
    evalPdf[i] = bsdfPdf(mat, etaI[i], etaO[i], N[i], view[i], L[i]);
}
// ProbeSample with a fresh Random(seed) (Probe.cuh:138-169; guide != nullptr: through the guide tables, as shadeKernel does), and
// ProbeDirToUV / ProbeEval / ProbePdf of given directions (:38-46, :61-67, :69-93)
__global__ void debugProbeKernel(Probe P, const int* __restrict__ guide, unsigned n, const unsigned* __restrict__ seeds, float3* dir, float3* color,
                                 float* pdf, unsigned m, const float3* __restrict__ dirs, float2* uv, float3* evalColor, float* evalPdf) {
    const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {
        Rng r;  r.seed(seeds[i]);
        float3 d, c;  float q;
        probeSample(P, guide, d, c, q, r);
        dir[i] = d;  color[i] = c;  pdf[i] = q;
    }
    if (i < m) {
        const float2 q = probeDirToUV(dirs[i]);
        uv[i] = q;
        const float4 c = probeEval(P, q);
        evalColor[i] = f3(c.x, c.y, c.z);
        evalPdf[i] = probePdf(P, dirs[i]);
    }
}
void launchDebugBsdf(unsigned n, const Material* mats, const float* etaI, const float* etaO, const float3* N, const float3* view, const float3* L,
                     const float3* albedo, const unsigned* seeds, float3* basisU, float3* basisV, float3* sampleL, float* samplePdf, int* sampleType,
                     uint2* sampleRng, float3* evalF, float* evalPdf, cudaStream_t s) {
    if (n) debugBsdfKernel<<<(n + 127) / 128, 128, 0, s>>>(n, mats, etaI, etaO, N, view, L, albedo, seeds, basisU, basisV, sampleL, samplePdf, sampleType, sampleRng, evalF, evalPdf);
}
void launchDebugProbe(const Probe& P, const int* guide, unsigned n, const unsigned* seeds, float3* dir, float3* color, float* pdf, unsigned m,
                      const float3* dirs, float2* uv, float3* evalColor, float* evalPdf, cudaStream_t s) {
    const unsigned k = n > m ? n : m;
    if (k) debugProbeKernel<<<(k + 127) / 128, 128, 0, s>>>(P, guide, n, seeds, dir, color, pdf, m, dirs, uv, evalColor, evalPdf);
}

// ---- Whitted-style integrator (04HelloRaytracing/deviceProgram.cu:66-243) on the same trace kernels ---------------------------
// raygen (:196-243): one ray through each pixel centre.  Pixel i is ray i and pool slot i.
__global__ void whittedRaygenKernel(WhittedParams P, float4* __restrict__ rayO, float4* __restrict__ rayD, PathPool pool, unsigned* counters) {
    const unsigned n = (unsigned)P.W * (unsigned)P.H;
    const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0) counters[QC_RAYS] = n;
    if (i >= n) return;
    const int ix = (int)(i % (unsigned)P.W), iy = (int)(i / (unsigned)P.W);
    const float sx = ((float)ix + .5f) / (float)P.W, sy = ((float)iy + .5f) / (float)P.H;                 // :213-214
    const float3 dir = normalize3(P.direction + (sx - 0.5f) * P.horizontal + (sy - 0.5f) * P.vertical);    // :217-219
    rayO[i] = make_float4(P.position.x, P.position.y, P.position.z, __int_as_float((int)i));
    rayD[i] = make_float4(dir.x, dir.y, dir.z, 0.f);
    pool.direct[i] = make_float4(0.f, 0.f, 0.f, 0.f);               // lightVisibility (:145): the shadow trace adds the ray's contribution (1) when unoccluded
    PathState st;                                                   // __miss__radiance (:172-175) until a hit says otherwise: eta < 0 marks "no surface"
    st.rng1 = st.rng2 = 0u;  st.thrX = st.thrY = st.thrZ = 0.f;  st.eta = -1.0f;  st.state = 0;  st.bsdfPdf = 0.f;
    stState(pool.st + i, st);
}
// __closesthit__radiance up to the shadow trace (:66-165); parks diffuseColor and cosDN in the pixel's path record for the finish kernel
__global__ void __launch_bounds__(SHADE_BLOCK)
whittedShadeKernel(WhittedParams P, SceneDev S, PathPool pool, Queues Q, unsigned* shadowCounter) {
    const unsigned n = (unsigned)P.W * (unsigned)P.H;
    const unsigned stride = gridDim.x * blockDim.x;
    for (unsigned base = blockIdx.x * blockDim.x; base < n; base += stride) {
        const unsigned i = base + threadIdx.x;                                                     // ray i = pixel i = path slot i
        bool wantShadow = false;
        float4 so = make_float4(0, 0, 0, 0), sd = so;
        if (i < n) {
            const float4 d4 = Q.rayD[0][i], h = Q.hit[i];
            const int triSlot = __float_as_int(h.w);
            if (triSlot >= 0) {                                                                     // a miss keeps the record whittedRaygen wrote
                const float4 tc = __ldg(S.tris + 3 * (size_t)triSlot + 2);
                const unsigned gid = (unsigned)__float_as_int(tc.y);
                const MeshInfo* M = S.meshes + __float_as_int(tc.z);
                const unsigned* ip = reinterpret_cast<const unsigned*>(S.indices + gid);
                const uint3 id = make_uint3(__ldg(ip), __ldg(ip + 1), __ldg(ip + 2));
                const float u = h.y, v = h.z, w0 = 1.f - u - v;
                const float3 A = S.vertices[M->firstVertex + id.x], B = S.vertices[M->firstVertex + id.y], C = S.vertices[M->firstVertex + id.z];
                float3 Ng = cross3(B - A, C - A);                                                       // :85
                float3 Ns = M->hasNormal ? (w0 * S.normals[M->firstVertex + id.x] + u * S.normals[M->firstVertex + id.y]
                                            + v * S.normals[M->firstVertex + id.z]) : Ng;               // :86-90
                const float3 rayDir = f3(d4.x, d4.y, d4.z);
                if (dot3(rayDir, Ng) > 0.f) Ng = -Ng;                                                   // :97-102
                Ng = normalize3(Ng);
                if (dot3(Ng, Ns) < 0.f) Ns = Ns - 2.f * dot3(Ng, Ns) * Ng;
                Ns = normalize3(Ns);
                float3 diffuse = M->material.color;                                                     // sbtData.color (:108)
                if (M->textureID >= 0 && M->hasTexcoord) {                                              // :109-116
                    const float2 t0 = __ldg(S.texcoords + M->firstVertex + id.x), t1 = __ldg(S.texcoords + M->firstVertex + id.y),
                                 t2 = __ldg(S.texcoords + M->firstVertex + id.z);
                    diffuse = diffuse * sampleTexture(S, M, w0 * t0.x + u * t1.x + v * t2.x, w0 * t0.y + u * t1.y + v * t2.y);
                }
                const float3 surfPos = w0 * A + u * B + v * C;                                          // :121-124
                const float3 lightDir = P.lightPos - surfPos;                                           // :125-126
                const float cosDN = 0.1f + .8f * fabsf(dot3(rayDir, Ns));                               // :158-160
                PathState st;
                st.rng1 = st.rng2 = 0u;  st.thrX = diffuse.x;  st.thrY = diffuse.y;  st.thrZ = diffuse.z;  st.eta = cosDN;  st.state = 0;  st.bsdfPdf = 0.f;
                stState(pool.st + i, st);
                const float3 o = surfPos + 1e-3f * Ng;                                                  // :134
                wantShadow = true;
                so = make_float4(o.x, o.y, o.z, __int_as_float((int)i));
                sd = make_float4(lightDir.x, lightDir.y, lightDir.z, __int_as_float(0));                // target 0: pool.direct
            }
        }
        enqueueShadow(wantShadow, so, sd, make_float4(1.f, 1.f, 1.f, 0.f), Q, shadowCounter);
    }
}
// the rest of __closesthit__radiance (:162-163) and the raygen epilogue (:232-242)
__global__ void whittedFinishKernel(WhittedParams P, PathPool pool, uint32_t* __restrict__ colorBuffer) {
    const unsigned n = (unsigned)P.W * (unsigned)P.H;
    const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const PathState t = ldState(pool.st + i);
    float3 prd = f3(0.f);
    if (t.eta >= 0.f) {
        const float4 vis4 = pool.direct[i];
        const float3 vis = f3(vis4.x, vis4.y, vis4.z);
        prd = (f3(.1f) + (f3(.2f) + .8f * vis) * t.eta) * f3(t.thrX, t.thrY, t.thrZ);
    }
    const int r = (int)(255.99f * prd.x), g = (int)(255.99f * prd.y), b = (int)(255.99f * prd.z);
    colorBuffer[i] = 0xff000000u | ((unsigned)r << 0) | ((unsigned)g << 8) | ((unsigned)b << 16);
}
void launchWhittedRaygen(const WhittedParams& P, const Queues& Q, const PathPool& pool, unsigned* counters, cudaStream_t s) {
    const unsigned n = (unsigned)P.W * (unsigned)P.H;
    whittedRaygenKernel<<<gridFor(n), SHADE_BLOCK, 0, s>>>(P, Q.rayO[0], Q.rayD[0], pool, counters);
}
void launchWhittedShade(const WhittedParams& P, const SceneDev& scene, const PathPool& pool, const Queues& Q, unsigned* shadowCounter, cudaStream_t s) {
    whittedShadeKernel<<<persistentGrid(), SHADE_BLOCK, 0, s>>>(P, scene, pool, Q, shadowCounter);
}
void launchWhittedFinish(const WhittedParams& P, const PathPool& pool, uint32_t* colorBuffer, cudaStream_t s) {
    const unsigned n = (unsigned)P.W * (unsigned)P.H;
    whittedFinishKernel<<<gridFor(n), SHADE_BLOCK, 0, s>>>(P, pool, colorBuffer);
}

