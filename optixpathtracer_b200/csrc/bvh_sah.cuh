// b200pt — binned-SAH treelet refinement of the LBVH (included by bvh_build.cu).
//
// The Karras tree orders triangles along the Morton curve and gives a hierarchy of *ranges*; its splits are
// spatial medians.  Refinement keeps the ranges and replaces the splits:
//   level 1: every maximal LBVH subtree with <= JOB_T triangles is a treelet; one CTA rebuilds it top-down with
//            a 16-bin SAH sweep entirely in shared memory (leaves of 1..8 triangles chosen by SAH cost) and
//            appends its nodes to the node pool as one contiguous block;
//   level k: the treelet roots of level k-1 are the items; every maximal LBVH subtree holding <= JOB_T items is
//            rebuilt the same way (leaf = one item); repeated until a single job produces the root.
// For 262 k triangles that is 3 levels (~1.5 k + 4 + 1 jobs); for 20 M, 3-4.  The result is a full SAH-quality
// BVH2 whose inner nodes are written as 32-byte records holding both (quantised) child boxes.
#pragma once

#define JOB_T 512                       // max items per treelet job
#define JOB_THREADS 256
#define JOB_WARPS (JOB_THREADS / 32)
#ifndef SAH_BINS
#define SAH_BINS 16
#endif
#define LINK_LOCAL 0x40000000           // job-local pair index (relocated on write-out)
#define REC_ROOT 0xffffu

struct SahJob { unsigned itemLo, count, triLo; };

struct SahTask { unsigned short b, e, rec, pad; };

struct SahLevelIn {                     // items of this level
    const float4* itemLo;               // level >= 2: unpadded boxes / links of the previous level's roots
    const float4* itemHi;
    const int*    itemLink;
    const int*    itemDepth;            // depth of the subtree under each item
    const float4* triLo;                // level 1: per-gid triangle boxes + Morton order
    const float4* triHi;
    const unsigned* sortedGid;
    int isTri;
    float ciLeaf;                       // inner-node cost used in the leaf-vs-split decision (tuned to the traversal kernel)
};
struct SahLevelOut {
    float4* nextLo;  float4* nextHi;  int* nextLink;  int* nextDepth;   // indexed by job
    unsigned* finalGid;                 // level 1: BVH-order triangle ids
    BvhNode* nodes;                     // node pool: one 32-byte record per inner node
    int* rootLink;                      // written by the last level's job
    double gridLo[3], gridScale[3];     // quantisation grid (see BvhNode)
    unsigned* pairCounter;              // pairs allocated so far
    unsigned* stats;                    // [0] leaves, [1] max leaf size
};

__device__ __forceinline__ unsigned fOrd(float f) { unsigned u = __float_as_uint(f); return (u & 0x80000000u) ? ~u : (u | 0x80000000u); }
__device__ __forceinline__ float fUnord(unsigned u) { return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u); }
__device__ __forceinline__ float halfArea(float lx, float ly, float lz, float hx, float hy, float hz) {
    const float dx = hx - lx, dy = hy - ly, dz = hz - lz;
    return dx * dy + dy * dz + dz * dx;
}
__device__ __forceinline__ float warpMin(float v) { for (int o = 16; o; o >>= 1) v = fminf(v, __shfl_xor_sync(0xffffffffu, v, o)); return v; }
__device__ __forceinline__ float warpMax(float v) { for (int o = 16; o; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o)); return v; }

struct SahBins { unsigned lo[3][SAH_BINS][3], hi[3][SAH_BINS][3], cnt[3][SAH_BINS]; };   // ordered-uint boxes

struct SahShared {
    float4 lo[JOB_T], hi[JOB_T];
    float4 rec[4 * JOB_T];              // output records: pair p -> rec[4p..4p+3] = {L.lo|link, L.hi, R.lo|link, R.hi}
    float4 rootRec[2];
    SahTask tasks[2][JOB_T];
    unsigned short idx[JOB_T], tmp[JOB_T];
    unsigned short mid[JOB_T];          // per task: split position (0 = leaf)
    unsigned short scan[JOB_T];
    SahBins bins[JOB_WARPS];
    unsigned nTasks[2], pairCount, base, leafCount, maxLeaf, depth, iter;
};

__device__ __forceinline__ float4* recPtr(SahShared& S, unsigned rec) { return rec == REC_ROOT ? S.rootRec : &S.rec[2 * rec]; }

// One warp decides and applies the split of one task.
__device__ void sahProcessTask(SahShared& S, const SahTask t, unsigned taskIndex, const SahLevelIn& in, const SahJob& job,
                               unsigned warp, unsigned lane) {
    const unsigned b = t.b, e = t.e, c = e - b;
    // ---- bounds of the node and of its centroids
    float bl[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, bh[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    float cl[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, ch[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (unsigned k = b + lane; k < e; k += 32) {
        const float4 l = S.lo[S.idx[k]], h = S.hi[S.idx[k]];
        const float lv[3] = {l.x, l.y, l.z}, hv[3] = {h.x, h.y, h.z};
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            bl[a] = fminf(bl[a], lv[a]);  bh[a] = fmaxf(bh[a], hv[a]);
            const float cen = 0.5f * (lv[a] + hv[a]);
            cl[a] = fminf(cl[a], cen);  ch[a] = fmaxf(ch[a], cen);
        }
    }
#pragma unroll
    for (int a = 0; a < 3; ++a) { bl[a] = warpMin(bl[a]);  bh[a] = warpMax(bh[a]);  cl[a] = warpMin(cl[a]);  ch[a] = warpMax(ch[a]); }
    float4* R = recPtr(S, t.rec);
    const float area = halfArea(bl[0], bl[1], bl[2], bh[0], bh[1], bh[2]);

    auto makeLeaf = [&]() {
        int link;
        if (in.isTri) {
            link = leafEncode(job.triLo + b, c);
            if (lane == 0) { atomicAdd(&S.leafCount, 1u);  atomicMax(&S.maxLeaf, c); }
        } else link = in.itemLink[job.itemLo + S.idx[b]];
        if (lane == 0) {
            atomicMax(&S.depth, S.iter + (in.isTri ? 0u : (unsigned)in.itemDepth[job.itemLo + S.idx[b]]));
            R[0] = make_float4(bl[0], bl[1], bl[2], __int_as_float(link));
            R[1] = make_float4(bh[0], bh[1], bh[2], 0.f);
            S.mid[taskIndex] = 0;
        }
    };
    if (c == 1 || (!in.isTri && c == 0)) { makeLeaf();  return; }

    // ---- bin the items on every non-degenerate axis
    float scale[3];
#pragma unroll
    for (int a = 0; a < 3; ++a) scale[a] = ch[a] > cl[a] ? (float)SAH_BINS * (1.0f - 1e-6f) / (ch[a] - cl[a]) : 0.f;
    SahBins& B = S.bins[warp];
    for (unsigned k = lane; k < 3 * SAH_BINS; k += 32) {
        const unsigned a = k / SAH_BINS, j = k % SAH_BINS;
        B.cnt[a][j] = 0;
#pragma unroll
        for (int q = 0; q < 3; ++q) { B.lo[a][j][q] = 0xffffffffu;  B.hi[a][j][q] = 0u; }
    }
    __syncwarp();
    for (unsigned k = b + lane; k < e; k += 32) {
        const float4 l = S.lo[S.idx[k]], h = S.hi[S.idx[k]];
        const float lv[3] = {l.x, l.y, l.z}, hv[3] = {h.x, h.y, h.z};
        const unsigned ol[3] = {fOrd(l.x), fOrd(l.y), fOrd(l.z)}, oh[3] = {fOrd(h.x), fOrd(h.y), fOrd(h.z)};
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            if (scale[a] == 0.f) continue;
            const int j = min(SAH_BINS - 1, (int)((0.5f * (lv[a] + hv[a]) - cl[a]) * scale[a]));
            atomicAdd(&B.cnt[a][j], 1u);
#pragma unroll
            for (int q = 0; q < 3; ++q) { atomicMin(&B.lo[a][j][q], ol[q]);  atomicMax(&B.hi[a][j][q], oh[q]); }
        }
    }
    __syncwarp();
    // ---- sweep: candidate = (axis, split) with split in 1..15; lanes take candidates lane and lane+32
    float bestCost = FLT_MAX;  int bestCand = -1;
    for (int cand = lane; cand < 3 * (SAH_BINS - 1); cand += 32) {
        const int a = cand / (SAH_BINS - 1), sp = cand % (SAH_BINS - 1) + 1;
        if (scale[a] == 0.f) continue;
        unsigned nl = 0, nr = 0;
        float ll[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, lh[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
        float rl[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, rh[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
        for (int j = 0; j < SAH_BINS; ++j) {
            const unsigned n = B.cnt[a][j];
            if (!n) continue;
            if (j < sp) { nl += n;  for (int q = 0; q < 3; ++q) { ll[q] = fminf(ll[q], fUnord(B.lo[a][j][q]));  lh[q] = fmaxf(lh[q], fUnord(B.hi[a][j][q])); } }
            else        { nr += n;  for (int q = 0; q < 3; ++q) { rl[q] = fminf(rl[q], fUnord(B.lo[a][j][q]));  rh[q] = fmaxf(rh[q], fUnord(B.hi[a][j][q])); } }
        }
        if (!nl || !nr) continue;
        const float cost = halfArea(ll[0], ll[1], ll[2], lh[0], lh[1], lh[2]) * (float)nl + halfArea(rl[0], rl[1], rl[2], rh[0], rh[1], rh[2]) * (float)nr;
        if (cost < bestCost) { bestCost = cost;  bestCand = cand; }
    }
    for (int o = 16; o; o >>= 1) {          // argmin, ties -> lowest candidate (deterministic)
        const float oc = __shfl_xor_sync(0xffffffffu, bestCost, o);
        const int od = __shfl_xor_sync(0xffffffffu, bestCand, o);
        if (od >= 0 && (bestCand < 0 || oc < bestCost || (oc == bestCost && od < bestCand))) { bestCost = oc;  bestCand = od; }
    }
    // ---- leaf or split
    if (in.isTri && c <= B200PT_MAX_LEAF) {
        const float leafCost = SAH_CT * area * (float)c;
        const float splitCost = bestCand >= 0 ? in.ciLeaf * area + SAH_CT * bestCost : FLT_MAX;
        if (leafCost <= splitCost) { makeLeaf();  return; }
    }
    unsigned nl;
    if (bestCand >= 0) {
        const int a = bestCand / (SAH_BINS - 1), sp = bestCand % (SAH_BINS - 1) + 1;
        nl = 0;
        for (int j = 0; j < sp; ++j) nl += B.cnt[a][j];
        unsigned doneL = 0, doneR = 0;
        for (unsigned k0 = b; k0 < e; k0 += 32) {
            const unsigned k = k0 + lane;
            bool valid = k < e, left = false;
            unsigned short id = 0;
            if (valid) {
                id = S.idx[k];
                const float4 l = S.lo[id], h = S.hi[id];
                const float cen = 0.5f * ((a == 0 ? l.x : a == 1 ? l.y : l.z) + (a == 0 ? h.x : a == 1 ? h.y : h.z));
                left = min(SAH_BINS - 1, (int)((cen - cl[a]) * scale[a])) < sp;
            }
            const unsigned lm = __ballot_sync(0xffffffffu, valid && left), rm = __ballot_sync(0xffffffffu, valid && !left);
            const unsigned lt = (1u << lane) - 1u;
            if (valid) S.tmp[left ? b + doneL + __popc(lm & lt) : b + nl + doneR + __popc(rm & lt)] = id;
            doneL += __popc(lm);  doneR += __popc(rm);
        }
        __syncwarp();
        for (unsigned k = b + lane; k < e; k += 32) S.idx[k] = S.tmp[k];
    } else nl = c / 2;                      // all centroids coincide: split the (Morton-ordered) range in half
    if (lane == 0) {
        R[0] = make_float4(bl[0], bl[1], bl[2], 0.f);                  // link filled in phase B
        R[1] = make_float4(bh[0], bh[1], bh[2], 0.f);
        S.mid[taskIndex] = (unsigned short)(b + nl);
    }
}

__global__ void __launch_bounds__(JOB_THREADS)
sahJobKernel(const SahJob* __restrict__ jobs, SahLevelIn in, SahLevelOut out, int lastLevel) {
    extern __shared__ __align__(16) unsigned char smemRaw[];
    SahShared& S = *reinterpret_cast<SahShared*>(smemRaw);
    const SahJob job = jobs[blockIdx.x];
    const unsigned tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (unsigned k = tid; k < job.count; k += JOB_THREADS) {
        if (in.isTri) { const unsigned g = in.sortedGid[job.itemLo + k];  S.lo[k] = in.triLo[g];  S.hi[k] = in.triHi[g]; }
        else { S.lo[k] = in.itemLo[job.itemLo + k];  S.hi[k] = in.itemHi[job.itemLo + k]; }
        S.idx[k] = (unsigned short)k;
    }
    if (tid == 0) {
        S.tasks[0][0] = SahTask{0, (unsigned short)job.count, (unsigned short)REC_ROOT, 0};
        S.nTasks[0] = 1;  S.nTasks[1] = 0;  S.pairCount = 0;  S.leafCount = 0;  S.maxLeaf = 0;  S.depth = 0;  S.iter = 0;
    }
    __syncthreads();
    int cur = 0;
    while (true) {
        const unsigned n = S.nTasks[cur];
        if (n == 0) break;
        for (unsigned t = warp; t < n; t += JOB_WARPS) sahProcessTask(S, S.tasks[cur][t], t, in, job, warp, lane);
        __syncthreads();
        // phase B: number the splitting tasks, allocate their child pairs, emit the next level's tasks
        for (unsigned t = tid; t < n; t += JOB_THREADS) S.scan[t] = S.mid[t] ? 1 : 0;
        __syncthreads();
        if (warp == 0) {                    // exclusive scan of <= JOB_T flags by one warp
            unsigned run = 0;
            for (unsigned k0 = 0; k0 < n; k0 += 32) {
                const unsigned k = k0 + lane;
                const unsigned f = k < n ? S.scan[k] : 0;
                const unsigned m = __ballot_sync(0xffffffffu, f != 0);
                if (k < n) S.scan[k] = (unsigned short)(run + __popc(m & ((1u << lane) - 1u)));
                run += __popc(m);
            }
            if (lane == 0) S.nTasks[cur ^ 1] = 2 * run;
        }
        __syncthreads();
        for (unsigned t = tid; t < n; t += JOB_THREADS) {
            if (!S.mid[t]) continue;
            const SahTask tk = S.tasks[cur][t];
            const unsigned p = S.pairCount + S.scan[t];
            float4* R = recPtr(S, tk.rec);
            R[0].w = __int_as_float((int)(p | LINK_LOCAL));
            S.tasks[cur ^ 1][2 * S.scan[t]] = SahTask{tk.b, S.mid[t], (unsigned short)(2 * p), 0};
            S.tasks[cur ^ 1][2 * S.scan[t] + 1] = SahTask{S.mid[t], tk.e, (unsigned short)(2 * p + 1), 0};
        }
        __syncthreads();
        if (tid == 0) { S.pairCount += S.nTasks[cur ^ 1] / 2;  S.nTasks[cur] = 0;  S.iter++; }
        __syncthreads();
        cur ^= 1;
    }
    // ---- write-out: reserve a contiguous block of the node pool, relocate local links, quantise both child boxes of every
    // inner node into ONE 32-byte record (BvhNode, common.cuh)
    if (tid == 0) {
        S.base = S.pairCount ? atomicAdd(out.pairCounter, S.pairCount) : 0u;
        if (in.isTri) { atomicAdd(out.stats + 0, S.leafCount);  atomicMax(out.stats + 1, S.maxLeaf); }
    }
    __syncthreads();
    const unsigned base = S.base;
    auto relocate = [&](int link) { return (link >= 0 && (link & LINK_LOCAL)) ? (int)(base + (unsigned)(link & ~LINK_LOCAL)) : link; };
    for (unsigned p = tid; p < S.pairCount; p += JOB_THREADS) {
        unsigned q[6];
        int links[2];
#pragma unroll
        for (int c = 0; c < 2; ++c) {
            const float4 lo = S.rec[4 * p + 2 * c], hi = S.rec[4 * p + 2 * c + 1];
            links[c] = relocate(__float_as_int(lo.w));
            const float l[3] = {lo.x, lo.y, lo.z}, h[3] = {hi.x, hi.y, hi.z};
            unsigned ql[3], qh[3];
#pragma unroll
            for (int a = 0; a < 3; ++a) {                      // conservative: floor/ceil in double, then GRID_PAD cells outward
                const double gl = B200PT_GRID_ORIGIN + ((double)l[a] - out.gridLo[a]) * out.gridScale[a];
                const double gh = B200PT_GRID_ORIGIN + ((double)h[a] - out.gridLo[a]) * out.gridScale[a];
                ql[a] = (unsigned)min(65535, max(0, (int)floor(gl) - B200PT_GRID_PAD));
                qh[a] = (unsigned)min(65535, max(0, (int)ceil(gh) + B200PT_GRID_PAD));
            }
            q[3 * c + 0] = ql[0] | (qh[0] << 16);  q[3 * c + 1] = ql[1] | (qh[1] << 16);  q[3 * c + 2] = ql[2] | (qh[2] << 16);
        }
        uint4* dst = reinterpret_cast<uint4*>(out.nodes) + 2 * (size_t)(base + p);
        dst[0] = make_uint4(q[0], q[1], q[2], q[3]);
        dst[1] = make_uint4(q[4], q[5], (unsigned)links[0], (unsigned)links[1]);
    }
    if (tid == 0) {
        const float4 lo = S.rootRec[0], hi = S.rootRec[1];
        const int link = relocate(__float_as_int(lo.w));
        if (lastLevel) *out.rootLink = link;
        out.nextLo[blockIdx.x] = make_float4(lo.x, lo.y, lo.z, 0.f);
        out.nextHi[blockIdx.x] = make_float4(hi.x, hi.y, hi.z, 0.f);
        out.nextLink[blockIdx.x] = link;
        out.nextDepth[blockIdx.x] = (int)S.depth;
    }
    if (in.isTri)
        for (unsigned k = tid; k < job.count; k += JOB_THREADS) out.finalGid[job.triLo + k] = in.sortedGid[job.itemLo + S.idx[k]];
}
