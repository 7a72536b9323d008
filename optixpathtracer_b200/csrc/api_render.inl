// b200pt — frame state + SampleRenderer::render() orchestration (included by api.cu).


// Folds the per-launch queue counters of one group into the frame totals {radiance rays, shadow rays, samples}
// and clears them for the next group (one tiny launch per group instead of two per bounce).
__global__ void tallyCountersKernel(unsigned* counters, int nLaunches, unsigned* totals) {
    unsigned rays = 0, shadows = 0;
    for (int d = threadIdx.x; d < nLaunches; d += 32) {
        if (d < nLaunches - 1) rays += counters[QC_STRIDE * d + QC_RAYS];      // the last launch traces shadow rays only
        shadows += counters[QC_STRIDE * d + QC_SHADOWS];
    }
    for (int o = 16; o; o >>= 1) { rays += __shfl_xor_sync(0xffffffffu, rays, o);  shadows += __shfl_xor_sync(0xffffffffu, shadows, o); }
    __syncwarp();
    const unsigned samples = counters[QC_RAYS];
    __syncwarp();
    if (threadIdx.x == 0) { totals[0] += rays;  totals[1] += shadows;  totals[2] += samples; }
    for (int i = threadIdx.x; i < QC_STRIDE * nLaunches; i += 32) counters[i] = 0;
}

static size_t poolLimit() {
    const char* e = getenv("B200PT_POOL_SLOTS");
    size_t v = e ? (size_t)atoll(e) : (size_t)8 << 20;
    return std::max<size_t>(v, 1024);
}

static void ensurePool(Lane& c, size_t slots, bool aovs) {
    if (slots > c.poolCap) {
        c.st.ensure(slots);  c.direct.ensure(slots);  c.indirect.ensure(slots);  c.alpha.ensure(slots);
        for (int i = 0; i < 2; ++i) { c.rayO[i].ensure(slots);  c.rayD[i].ensure(slots); }
        c.hit.ensure(slots);  c.shO.ensure(slots);  c.shD.ensure(slots);  c.shP.ensure(slots);
        c.poolCap = slots;
    }
    if (aovs) { c.aovN.ensure(c.poolCap);  c.aovA.ensure(c.poolCap); }
    c.counters.ensure(QC_STRIDE * QC_MAX_LAUNCHES + 8);                      // per-launch queue counters, then 8 frame totals
    c.stats.ensure(2);
}
static PathPool poolOf(Lane& L) { return PathPool{(unsigned)L.poolCap, L.st.p, L.direct.p, L.indirect.p, L.alpha.p, L.aovN.p, L.aovA.p}; }
static Queues queuesOf(Lane& L) {
    Queues Q;
    for (int i = 0; i < 2; ++i) { Q.rayO[i] = L.rayO[i].p;  Q.rayD[i] = L.rayD[i].p; }
    Q.hit = L.hit.p;  Q.shO = L.shO.p;  Q.shD = L.shD.p;  Q.shP = L.shP.p;  Q.counters = L.counters.p;  Q.cap = (unsigned)L.poolCap;
    return Q;
}
// lane k > 0 gets its stream and `done` event on first use
static Lane& laneOf(b200pt_ctx* c, int k) {
    Lane& L = c->lanes[k];
    if (!L.stream) CUDA_OK(cudaStreamCreateWithFlags(&L.stream, cudaStreamNonBlocking));
    if (!L.done) CUDA_OK(cudaEventCreateWithFlags(&L.done, cudaEventDisableTiming));
    return L;
}

// NVTX ranges (SURVEY §5 tracing row): one per render() call and one per stage of a wavefront, around the host-side submission of that stage's
// launches - what a timeline tool (Nsight Systems) needs to attribute the kernels of overlapping frames.  Header-only nvtx3: a no-op costing a few
// nanoseconds per call when no tool is attached.
static const char* const kStageName[] = {"b200pt raygen", "b200pt trace", "b200pt trace (shadow tail)", "b200pt shade", "b200pt accumulate", "b200pt frame"};
struct NvtxRange { explicit NvtxRange(const char* name) { nvtxRangePushA(name); }  ~NvtxRange() { nvtxRangePop(); } };

struct StageTimer {          // NVTX range per stage; CUDA-event pairs per stage, resolved after the frame (profile=1 only)
    b200pt_ctx* c;  bool on;
    void begin(int stage, cudaEvent_t& a, cudaEvent_t& b) {
        if (stage != 5 /* ST_TOTAL: spans two functions, see render() */) nvtxRangePushA(kStageName[stage]);
        if (!on) return;
        while (c->evPool.size() < c->evNext + 2) { cudaEvent_t e;  CUDA_OK(cudaEventCreate(&e));  c->evPool.push_back(e); }
        a = c->evPool[c->evNext++];  b = c->evPool[c->evNext++];
        c->evUsed.push_back({stage, {a, b}});
        CUDA_OK(cudaEventRecord(a, c->stream));
    }
    void end(cudaEvent_t b, bool stageRange = true) { if (on) CUDA_OK(cudaEventRecord(b, c->stream));  if (stageRange) nvtxRangePop(); }
};
enum { ST_RAYGEN = 0, ST_TRACE, ST_TRACE_TAIL, ST_SHADE, ST_ACCUM, ST_TOTAL };

static void collectCounters(b200pt_ctx* c) {
    if (!c->countersPending) return;
    c->last.radiance_rays = 0;  c->last.shadow_rays = 0;  c->last.samples = 0;
    std::vector<int> lanes(1, c->countersLane);
    lanes.insert(lanes.end(), c->countersExtra.begin(), c->countersExtra.end());
    for (int k : lanes) {                                                    // a split frame: one set of totals per part
        unsigned h[8];
        CUDA_OK(cudaMemcpy(h, c->lanes[k].counters.p + QC_STRIDE * QC_MAX_LAUNCHES, sizeof(h), cudaMemcpyDeviceToHost));
        c->last.radiance_rays += h[0];  c->last.shadow_rays += h[1];  c->last.samples += h[2];
    }
    Lane& L = c->lanes[c->countersLane];
    if (c->profiled & 2) {
        TraceStats st[2];
        CUDA_OK(cudaMemcpy(st, L.stats.p, sizeof(st), cudaMemcpyDeviceToHost));
        c->last.nodes_fetched = st[0].nodes;  c->last.tris_fetched = st[0].tris;
        c->last.nodes_fetched_shadow = st[1].nodes;  c->last.tris_fetched_shadow = st[1].tris;
        c->last.closest_hits = st[0].hits;  c->last.occluded_shadow_rays = st[1].hits;
    }
    if (c->profiled & 1) {
        float acc[6] = {0, 0, 0, 0, 0, 0};
        for (auto& u : c->evUsed) { float ms = 0.f;  cudaEventElapsedTime(&ms, u.second.first, u.second.second);  acc[u.first] += ms; }
        c->last.ms_raygen = acc[ST_RAYGEN];  c->last.ms_trace_closest = acc[ST_TRACE];  c->last.ms_trace_any = acc[ST_TRACE_TAIL];
        c->last.ms_shade = acc[ST_SHADE];  c->last.ms_accumulate = acc[ST_ACCUM];  c->last.ms_total = acc[ST_TOTAL];
    }
    c->countersPending = false;
}

static PassDesc makePass(const LaunchParams& P, int launchW, int launchH) {
    PassDesc d;
    PassParams& q = d.P;
    q.launchW = launchW;  q.launchH = launchH;  q.factorX = P.frame.factor.x;  q.factorY = P.frame.factor.y;
    q.fillSize = P.frame.fillSize;  q.cX = P.frame.c.x;  q.cY = P.frame.c.y;  q.rInner = P.frame.r_inner;  q.rOuter = P.frame.r_outer;
    q.offX = P.frame.offset.x;  q.offY = P.frame.offset.y;  q.redraw = P.frame.redraw;  q.subframe = P.frame.subframe_index;
    q.spl = (int)P.samples_per_launch;  q.slotBase = 0;  q.launchBase = 0;
    d.nLaunch = (size_t)std::max(launchW, 0) * (size_t)std::max(launchH, 0);
    return d;
}

// The launch schedule of SampleRenderer::render(): mutates P exactly as the reference leaves its public
// launchParams (sv4/SimplePathtracer.cpp:85-131 uniform, :136-214 foveated; orig/SimplePathtracer.cpp:73-97).
// Deviation: fields are uploaded AFTER they are set, not one frame late.
static void buildSchedule(LaunchParams& P, const b200pt_config& cfg, std::vector<PassDesc>& passes) {
    const int W = P.frame.size.x, H = P.frame.size.y;
    if (cfg.schedule == B200PT_SCHEDULE_AS_GIVEN) {
        passes.push_back(makePass(P, cfg.launch_w ? cfg.launch_w : W, cfg.launch_h ? cfg.launch_h : H));
        if (cfg.increment_subframe) P.frame.subframe_index++;
    } else if (cfg.schedule == B200PT_SCHEDULE_UNIFORM) {
        if (!cfg.accumulate) P.frame.subframe_index = 0;
        P.frame.factor = make_uint3(1, 1, 1);  P.frame.fillSize = 1;  P.frame.r_outer = 1000000000;  P.frame.r_inner = 0;
        if (cfg.uniform_spl > 0) P.samples_per_launch = cfg.uniform_spl;
        P.frame.offset = make_uint2(0, 0);  P.frame.redraw = 0;
        P.viewportSize.x = W;  P.viewportSize.y = H;
        const unsigned temp = P.frame.subframe_index;
        passes.push_back(makePass(P, W, H));
        P.frame.subframe_index = temp;
        if (cfg.increment_subframe) P.frame.subframe_index++;
    } else {
        const int inner = cfg.inner_radius, outer = cfg.outer_radius;
        P.frame.factor = make_uint3(4, 4, 1);  P.frame.fillSize = 4;  P.frame.r_outer = 1000000000;  P.frame.r_inner = (float)outer;
        P.samples_per_launch = 1;  P.frame.offset = make_uint2(0, 0);  P.frame.redraw = 0;
        passes.push_back(makePass(P, W / 4, H / 4));
        const unsigned temp = P.frame.subframe_index;
        P.frame.subframe_index = 0;
        P.frame.factor = make_uint3(2, 2, 1);  P.frame.fillSize = 2;  P.frame.r_outer = (float)(outer + 2);  P.frame.r_inner = (float)inner;
        P.samples_per_launch = 2;
        P.frame.offset = make_uint2(P.frame.c.x - (unsigned)(outer + 2), P.frame.c.y - (unsigned)(outer + 2));
        P.frame.redraw = 1;
        passes.push_back(makePass(P, (int)P.frame.r_outer, (int)P.frame.r_outer));
        P.frame.factor = make_uint3(1, 1, 1);  P.frame.fillSize = 1;  P.frame.r_outer = (float)(inner + 1);  P.frame.r_inner = 0;
        P.samples_per_launch = 8;
        P.frame.offset = make_uint2(P.frame.c.x - (unsigned)(inner + 1), P.frame.c.y - (unsigned)(inner + 1));
        P.frame.redraw = 1;
        passes.push_back(makePass(P, (int)(P.frame.r_outer * 2), (int)(P.frame.r_outer * 2)));
        P.frame.subframe_index = temp;
        if (cfg.increment_subframe) P.frame.subframe_index++;
    }
}

static FrameParams frameOf(const LaunchParams& P, const b200pt_config& cfg, const b200pt_ctx* c = nullptr) {
    FrameParams F;
    F.W = P.frame.size.x;  F.H = P.frame.size.y;
    F.eye = P.camera.eye;  F.U = P.camera.U;  F.V = P.camera.V;  F.Wv = P.camera.W;
    F.probe = P.probe;
    // the guide tables belong to the probe uploaded by b200pt_set_probe*; a caller-substituted probe falls back to the plain search
    F.probeGuide = (c && c->probeGuide.p && P.probe.cdfValuesX == c->probeCdfX.p && P.probe.cdfValuesY == c->probeCdfY.p && P.probe.width == c->probeW && P.probe.height == c->probeH) ? c->probeGuide.p : nullptr;
    F.tmin = cfg.tmin;  F.tmax = cfg.tmax;  F.maxDepth = cfg.max_depth;  F.shadowMode = cfg.shadow_mode;  F.display = cfg.display;
    F.accumMode = cfg.accum_mode;  F.writeAovs = cfg.write_aovs;  F.russianRoulette = cfg.russian_roulette;  F.misOnMiss = cfg.mis_on_miss;  F.exposure = cfg.exposure;
    F.tileRank = cfg.tile_rank;  F.tileCount = cfg.tile_count;
    F.accum = P.frame.accum_buffer;  F.frame = (uint32_t*)P.frame.frame_buffer;  F.normal = P.frame.normal_buffer;
    F.albedo = P.frame.albedo_buffer;  F.color = P.frame.color_buffer;
    return F;
}

struct Chunk { int pass, k0, k1;  unsigned slotBase; };

// A shadow-catcher pass-through (sv4/deviceProgram.cu:617-622) re-traces a segment without consuming depth, so a path that crosses catcher
// geometry k times needs max_depth + 1 + k trace launches; the wavefront allows max(8, max_depth + 1) crossings per path - one per segment plus the
// primary ray's, the case of a single catcher plane - (launches on empty queues cost ~2 us each); a path that crosses more often is cut off there.
#define B200PT_CATCHER_EXTRA_LAUNCHES 8

// Enqueues one wavefront (raygen, the bounce loop, counters, per-launch-index sums) of `job` on lane `laneIdx`: the whole frame (nParts = 1)
// or the launch rows of band set `part` of `nParts` (a frame split into interleaved 32-row band sets that run concurrently on their own lanes;
// the same partition as the multi-GPU tile split, so the pixels are those of the unsplit frame bit for bit).  The sums land in sumLane's buffers.
static void enqueuePart(b200pt_ctx* c, const b200pt_config& cfg, FrameJob& job, int laneIdx, int part, int nParts, Lane& sumLane) {
    std::vector<PassDesc>& passes = job.passes;
    FrameParams F = job.F;
    if (nParts > 1) {                                                            // band b belongs to the part iff b % (tc * nParts) == tr + tc * part
        const int tc = std::max(F.tileCount, 1), tr = F.tileCount > 1 ? F.tileRank : 0;
        F.tileRank = tr + tc * part;  F.tileCount = tc * nParts;
    }
    const SceneDev S = sceneOf(c);
    Lane& L = laneOf(c, laneIdx);
    cudaStream_t s = L.stream;
    if (laneIdx > 0 && !L.forked) {                                              // first ordered after whatever the primary stream held when the batch began
        CUDA_OK(cudaEventRecord(c->forkEv, c->stream));
        CUDA_OK(cudaStreamWaitEvent(s, c->forkEv, 0));
        L.forked = true;
    }
    // a part writes its sums into ANOTHER lane's buffers: wait until the previous frame's finalize passes have read them
    if (nParts > 1 && c->lastLane >= 0 && c->lastLane != laneIdx && c->lanes[c->lastLane].done) CUDA_OK(cudaStreamWaitEvent(s, c->lanes[c->lastLane].done, 0));

    const size_t limit = poolLimit();
    std::vector<std::vector<Chunk>> groups(1);
    size_t used = 0, maxUsed = 0;
    for (int pi = 0; pi < (int)passes.size(); ++pi) {
        const size_t nL = passes[pi].nLaunch;
        if (!nL) continue;
        const int kmax = (int)std::max<size_t>(1, limit / nL);
        for (int k0 = 0; k0 < passes[pi].P.spl; k0 += kmax) {
            const int k1 = std::min(passes[pi].P.spl, k0 + kmax);
            const size_t need = nL * (size_t)(k1 - k0);
            if (used && used + need > limit) { groups.emplace_back();  used = 0; }
            groups.back().push_back({pi, k0, k1, (unsigned)used});
            used += need;  maxUsed = std::max(maxUsed, used);
        }
    }
    if (maxUsed >= ((size_t)1 << 31)) throw std::runtime_error("b200pt_render: pool too large");
    ensurePool(L, std::max<size_t>(maxUsed, 1), cfg.write_aovs != 0);

    PathPool pool = poolOf(L);
    Queues Q = queuesOf(L);
    // One group whose chunks are whole passes: the per-launch-index records are not needed - the finalize passes read the pool(s) (which nothing
    // reuses before them: a lane's next wavefront is enqueued behind its finalize passes).  B200PT_FUSE_GATHER=0 keeps the separate gather pass (A/B).
    static const bool fuseAllowed = [] { const char* e = getenv("B200PT_FUSE_GATHER");  return !e || atoi(e) != 0; }();
    // (every band set of a split frame sees the same passes and the same pool limit, so they all decide alike and share the passes' slot bases)
    bool fuse = fuseAllowed && nParts <= 4 && groups.size() == 1;
    for (auto& ch : groups[0]) fuse = fuse && ch.k0 == 0 && ch.k1 == passes[ch.pass].P.spl;
    if (fuse) {
        job.fused.n = nParts;  job.fused.p[part] = pool;
        for (auto& ch : groups[0]) passes[ch.pass].P.slotBase = ch.slotBase;
    }

    const bool prof = (cfg.profile & 1) != 0, count = (cfg.profile & 2) != 0;   // 1: stage events, 2: traversal counters
    if (prof) { c->evUsed.clear();  c->evNext = 0; }
    StageTimer T{c, prof};
    cudaEvent_t a = nullptr, b = nullptr, ta = nullptr;
    unsigned& launches = job.launches;
    unsigned* totals = L.counters.p + QC_STRIDE * QC_MAX_LAUNCHES;
    CUDA_OK(cudaMemsetAsync(L.counters.p, 0, sizeof(unsigned) * (QC_STRIDE * QC_MAX_LAUNCHES + 8), s));
    if (count) CUDA_OK(cudaMemsetAsync(L.stats.p, 0, 2 * sizeof(TraceStats), s));
    T.begin(ST_TOTAL, ta, job.totalEnd);
    // the dropped last segment is traced only if alpha can still change (a catcher in the scene) or the miss carries radiance (mis_on_miss)
    const int bounces = cfg.max_depth + ((c->hasCatcher || cfg.max_depth == 0) ? 1 : 0) + (c->hasCatcher ? std::max(B200PT_CATCHER_EXTRA_LAUNCHES, cfg.max_depth + 1) : 0);
    const int cullBack = cfg.shadow_mode == B200PT_SHADOW_CULL_BACK;
    const float shadowTmin = cfg.shadow_tmin > 0.f ? cfg.shadow_tmin : cfg.tmin;
    // Frames in flight are kept staggered: this wavefront starts only when the wavefront enqueued just before it (on another lane) has finished its pass
    // `stagger`.  Left alone, the kernels of concurrent frames are dispatched in launch order, the younger frames catch up with the oldest and all of them finish
    // together: a blocking caller then waits ~3 frame times for one call in three (9.3 / 0.3 / 0.4 ms at 1080p) and a run that stops pays for the frames it
    // never asks for.  With the head start the oldest frame stays ahead and the calls return at an even pace (3.2 / 3.9 / 2.8 ms) at 99 % of the throughput
    // (measured: 2 passes; 3 evens the calls out completely but costs 3 %, holding EVERY pass back costs 1.5-3 %).  Only the blocking render() asks for it
    // (c->staggerWaves): a batch submitted with b200pt_render_async has nobody waiting for single frames.  B200PT_STAGGER=0: off.
    static const int stagger = [] { const char* e = getenv("B200PT_STAGGER");  return e ? atoi(e) : 2; }();
    const bool wholeFrame = nParts == 1 && groups.size() == 1 && laneIdx > 0;
    if (stagger > 0 && c->staggerWaves && wholeFrame && c->wavePrevLane > 0 && c->wavePrevLane != laneIdx && c->lanes[c->wavePrevLane].headStart)
        CUDA_OK(cudaStreamWaitEvent(s, c->lanes[c->wavePrevLane].headStart, 0));
    if (wholeFrame && !L.headStart) CUDA_OK(cudaEventCreateWithFlags(&L.headStart, cudaEventDisableTiming));
    for (auto& group : groups) {
        if (group.empty()) continue;
        T.begin(ST_RAYGEN, a, b);
        for (auto& ch : group) {
            PassParams P = passes[ch.pass].P;  P.slotBase = ch.slotBase;
            launchRaygen(F, P, ch.k0, ch.k1, pool, Q, s);  ++launches;
        }
        T.end(b);
        int cur = 0;
        // trace launch d: extension rays of bounce d + the shadow rays shade pass d-1 emitted; launch `bounces` only shadows
        for (int d = 0; d <= bounces; ++d) {
            T.begin(d < bounces ? ST_TRACE : ST_TRACE_TAIL, a, b);
            launchTrace(S, Q, cur, d < bounces ? 1 : 0, Q.counters + QC_STRIDE * d, cfg.tmin, shadowTmin, cfg.tmax, cullBack, pool, count ? L.stats.p : nullptr, s);
            T.end(b);  ++launches;  ++job.nClosest;
            if (d == bounces) { if (wholeFrame && d <= stagger) CUDA_OK(cudaEventRecord(L.headStart, s));  break; }
            T.begin(ST_SHADE, a, b);
            launchShade(F, S, pool, Q, cur, d, s);
            T.end(b);  ++launches;
            if (wholeFrame && d == stagger) CUDA_OK(cudaEventRecord(L.headStart, s));
            cur ^= 1;
        }
        tallyCountersKernel<<<1, 32, 0, s>>>(Q.counters, bounces + 1, totals);  ++launches;
        if (fuse) continue;
        T.begin(ST_ACCUM, a, b);
        for (auto& ch : group) {
            PassParams P = passes[ch.pass].P;  P.slotBase = ch.slotBase;
            launchGather(F, P, ch.k0, ch.k1, pool, sumLane.sumRad.p, sumLane.sumAlpha.p, sumLane.sumNormal.p, sumLane.sumAlbedo.p, s);  ++launches;
        }
        T.end(b);
    }
    CUDA_OK(cudaGetLastError());
    if (part == 0) c->wavePrevLane = wholeFrame ? laneIdx : -1;
}

// Enqueues one frame's wavefront on `lanes` (one lane: the whole frame; several: the frame split into that many band sets).  `params` is mutated
// exactly as the reference's render() leaves its public launchParams; nothing outside the lanes is written.
static void enqueueWavefront(b200pt_ctx* c, LaunchParams& params, const b200pt_config& cfg, const std::vector<int>& lanes, FrameJob& job) {
    job = FrameJob();
    job.lane = lanes[0];
    job.extraLanes.assign(lanes.begin() + 1, lanes.end());
    buildSchedule(params, cfg, job.passes);
    job.F = frameOf(params, cfg, c);
    size_t nRec = 0;                                                             // launch-index records
    for (auto& p : job.passes) { if (p.P.spl <= 0) throw std::runtime_error("b200pt_render: samples_per_launch must be > 0");  p.P.launchBase = (unsigned)nRec;  nRec += p.nLaunch; }
    Lane& L0 = laneOf(c, job.lane);
    L0.sumRad.ensure(std::max<size_t>(nRec, 1));  L0.sumAlpha.ensure(std::max<size_t>(nRec, 1));
    if (cfg.write_aovs) { L0.sumNormal.ensure(std::max<size_t>(nRec, 1));  L0.sumAlbedo.ensure(std::max<size_t>(nRec, 1)); }
    const int nParts = (int)lanes.size();
    for (int part = 0; part < nParts; ++part) {
        enqueuePart(c, cfg, job, lanes[part], part, nParts, L0);
        if (part > 0) CUDA_OK(cudaEventRecord(c->lanes[lanes[part]].done, c->lanes[lanes[part]].stream));   // the finalize passes wait for it
    }
}
static void enqueueWavefront(b200pt_ctx* c, LaunchParams& params, const b200pt_config& cfg, int laneIdx, FrameJob& job) {
    enqueueWavefront(c, params, cfg, std::vector<int>(1, laneIdx), job);
}

// The finalize passes of an enqueued wavefront: frames fold into the accumulation / frame / AOV buffers in submission order, whatever lane they ran on.
static void enqueueFinalize(b200pt_ctx* c, const b200pt_config& cfg, FrameJob& job) {
    Lane& L = c->lanes[job.lane];
    cudaStream_t s = L.stream;
    if (c->lastLane >= 0 && c->lastLane != job.lane) CUDA_OK(cudaStreamWaitEvent(s, c->lanes[c->lastLane].done, 0));
    for (int k : job.extraLanes) CUDA_OK(cudaStreamWaitEvent(s, c->lanes[k].done, 0));      // the other parts of a split frame
    StageTimer T{c, (cfg.profile & 1) != 0};
    cudaEvent_t a = nullptr, b = nullptr;
    T.begin(ST_ACCUM, a, b);
    // one-sided tile split (api_peer.inl): this rank's bands of the RGBA8 frame are written by finalizeKernel straight into root's frame buffer over
    // NVLink, once root has consumed the previous frame; a one-thread kernel behind the passes tells root that they have landed
    FrameParams F = job.F;
    const int role = peerPutBegin(c, cfg, F, s);
    int nLive = 0, done = 0;
    for (auto& p : job.passes) nLive += p.nLaunch ? 1 : 0;
    for (auto& p : job.passes) {                                                 // pass order = the reference's launch order (overwrite semantics)
        if (!p.nLaunch) continue;
        launchFinalize(F, p.P, L.sumRad.p, L.sumAlpha.p, L.sumNormal.p, L.sumAlbedo.p, job.fused, s);  ++job.launches;
        // several passes (the foveated schedule) written by several ranks into ONE frame: pass k + 1 of any rank may overwrite pass k of another, so every
        // rank waits here until all of them have finished pass k (epoch flags in root's memory; nothing synchronises with a host)
        if (role && ++done < nLive) peerPassBarrier(c, role, done, s);
    }
    if (role) peerPutEnd(c, role, s);
    T.end(b);
    T.end(job.totalEnd, false);
    CUDA_OK(cudaGetLastError());
    CUDA_OK(cudaEventRecord(L.done, s));
    c->lastLane = job.lane;  c->countersLane = job.lane;  c->countersExtra = job.extraLanes;  L.busy = job.lane > 0;
    c->profiled = cfg.profile & 3;
    c->last = b200pt_counters{};
    c->last.kernel_launches = job.launches;  c->last.n_trace_closest_launches = job.nClosest;  c->last.n_trace_any_launches = 0;  c->last.fused_finalize = job.fused.n > 0 ? 1u : 0u;
    c->countersPending = true;
}

static void renderImpl(b200pt_ctx* c, LaunchParams* params, const b200pt_config* cfgp, bool sync) {
    if (!c || !params || !cfgp) throw std::runtime_error("b200pt_render: null argument");
    if (params->frame.size.x == 0) return;                                       // SimplePathtracer.cpp:81
    const b200pt_config cfg = *cfgp;
    NvtxRange frameRange(sync ? "b200pt_render" : "b200pt_render_async");
    CUDA_OK(cudaSetDevice(c->device));
    if (!c->built) throw std::runtime_error("b200pt_render: b200pt_build_accel has not been called");
    if (!params->probe.data || !params->probe.cdfValuesX) throw std::runtime_error("b200pt_render: no probe set (call b200pt_set_probe)");
    if (!params->frame.accum_buffer || !params->frame.frame_buffer) throw std::runtime_error("b200pt_render: frame buffers not allocated (call b200pt_resize)");
    if (cfg.max_depth < 0 || cfg.max_depth > (c->hasCatcher ? (QC_MAX_LAUNCHES - 12) / 2 : QC_MAX_LAUNCHES - 8)) throw std::runtime_error("b200pt_render: max_depth out of range");
    c->countersPending = false;                                                  // totals of an un-synchronised previous frame are dropped

    // Lanes: instrumented frames run on lane 0 (the primary stream); everything else rotates over cfg.frames_in_flight lanes.  With one lane
    // there is nothing to overlap and nothing to speculate on.
    Speculation& sp = c->spec;
    const int nLanes = (cfg.profile & 3) ? 1 : std::min(std::max(cfg.frames_in_flight, 1), B200PT_MAX_FRAMES_IN_FLIGHT);
    auto nextLane = [&]() { return nLanes > 1 ? 1 + (int)(c->seq++ % (unsigned)nLanes) : 0; };
    if (!sync || nLanes == 1) {                                                  // asynchronous submission, or strictly one frame at a time
        sp.drop();  sp.havePrev = false;
        FrameJob job;
        // one frame at a time, split into interleaved band sets on lanes 1..nSplit: the drain phase of every persistent launch and the thin late bounces
        // of one part are filled by the other parts' kernels, which is what frames in flight buy for a caller that can run frames ahead
        const int nSplit = (nLanes == 1 && !(cfg.profile & 3)) ? std::min(std::max(cfg.intra_frame_split, 1), B200PT_MAX_FRAMES_IN_FLIGHT) : 1;
        if (nSplit > 1) {
            std::vector<int> lanes;
            for (int k = 1; k <= nSplit; ++k) lanes.push_back(k);
            enqueueWavefront(c, *params, cfg, lanes, job);
        } else enqueueWavefront(c, *params, cfg, nextLane(), job);
        enqueueFinalize(c, cfg, job);
        if (sync) { CUDA_OK(cudaStreamSynchronize(c->lanes[job.lane].stream));  collectCounters(c); }   // CUDA_SYNC_CHECK, SimplePathtracer.cpp:214
        return;
    }

    // ---- blocking call with lanes to spare: consume the oldest speculative wavefront if this call is the one that was predicted
    // (not in ACCUM_SUM mode: a frame of a sample split is a partial sum nobody looks at on its own, and the collective that follows every call competes with
    // the wavefronts for the GPU - measured on 2 x B200: e2e 1163 without the head start, 1122 with it)
    struct StaggerScope { b200pt_ctx* c;  StaggerScope(b200pt_ctx* c_, bool on) : c(c_) { c->staggerWaves = on; }  ~StaggerScope() { c->staggerWaves = false; } }
        staggerScope(c, cfg.accum_mode != B200PT_ACCUM_SUM);
    const unsigned thisIndex = params->frame.subframe_index;
    // "the previous call again, apart from subframe_index" (progressive accumulation) / "... and with exactly the predicted index"
    bool sameButIndex = false;
    if (sp.havePrediction && memcmp(&cfg, &sp.predictedCfg, sizeof(cfg)) == 0) {
        LaunchParams a, b;
        memcpy(&a, params, sizeof(LaunchParams));  memcpy(&b, &sp.predicted, sizeof(LaunchParams));
        a.frame.subframe_index = 0;  b.frame.subframe_index = 0;
        sameButIndex = memcmp(&a, &b, sizeof(LaunchParams)) == 0;
    }
    const bool predictedRight = sameButIndex && thisIndex == sp.predicted.frame.subframe_index;
    sp.hits = predictedRight ? sp.hits + 1 : 0;
    FrameJob job;
    if (predictedRight && !sp.ahead.empty()) {
        job = std::move(sp.ahead.front().job);
        memcpy(params, &sp.ahead.front().output, sizeof(LaunchParams));
        sp.ahead.pop_front();
    } else {
        sp.ahead.clear();                                                        // wrong guesses are simply left behind on their lanes
        enqueueWavefront(c, *params, cfg, nextLane(), job);
    }
    enqueueFinalize(c, cfg, job);
    const b200pt_counters mine = c->last;
    const int myLane = job.lane;

    // ---- predict the next calls: the launchParams as a call leaves them, subframe_index advanced by the stride seen between the last two calls
    // stride of the index: kept while the guesses are right; re-learnt from the last two calls when only the index was off (a caller striding over
    // sub-frames, e.g. one rank of a sample split; 0 for a caller that keeps re-rendering one sub-frame); 1 after any other change (the reference's
    // loop restarts at index 0 and counts up, sv4/main.cpp:319-320,392)
    if (!predictedRight) {
        sp.stride = (sameButIndex && sp.havePrev) ? thisIndex - sp.prevIndex : 1u;
        if ((int)sp.stride < 0) sp.stride = 1u;                                  // the index went backwards: an accumulation restarted, it counts up from here
    }
    const unsigned stride = sp.stride;
    sp.prevIndex = thisIndex;  sp.havePrev = true;
    memcpy(&sp.predicted, params, sizeof(LaunchParams));
    sp.predicted.frame.subframe_index = thisIndex + stride;
    sp.predictedCfg = cfg;  sp.havePrediction = true;
    // a right guess: keep nLanes - 1 predicted frames' wavefronts in flight - one only when a frame is heavy (>= 16 M samples: 32 spp at 1080p, 8 spp at 4K): such a
    // frame fills the GPU almost by itself (two in flight already give the overlap: +6 %), and every frame run ahead for a call that never comes costs as much as
    // a real one (Lost-Empire-class config, 8 calls of 40 ms: e2e 1335 Msamples/s with two frames ahead against 1560 without speculation)
    size_t frameSamples = 0;
    for (auto& p : job.passes) frameSamples += p.nLaunch * (size_t)std::max(p.P.spl, 1);
    const int depth = frameSamples >= ((size_t)16 << 20) ? 1 : nLanes - 1;
    if (sp.hits >= 1) {
        while ((int)sp.ahead.size() < depth) {
            SpecFrame f;
            if (sp.ahead.empty()) memcpy(&f.input, &sp.predicted, sizeof(LaunchParams));
            else {
                memcpy(&f.input, &sp.ahead.back().output, sizeof(LaunchParams));
                f.input.frame.subframe_index = sp.ahead.back().input.frame.subframe_index + stride;
            }
            memcpy(&f.output, &f.input, sizeof(LaunchParams));
            sp.ahead.push_back(std::move(f));
            enqueueWavefront(c, sp.ahead.back().output, cfg, nextLane(), sp.ahead.back().job);
        }
    } else sp.ahead.clear();
    CUDA_OK(cudaStreamSynchronize(c->lanes[myLane].stream));                      // CUDA_SYNC_CHECK, SimplePathtracer.cpp:214 - this frame only
    c->last = mine;  c->countersLane = myLane;  c->countersExtra.clear();  c->countersPending = true;
    collectCounters(c);
}

extern "C" {

int b200pt_resize(b200pt_ctx* c, int w, int h, LaunchParams* params) {
    if (!c || !params) return fail("b200pt_resize: null argument");
    if (w == 0 || h == 0) return 0;                                              // minimised window, SimplePathtracer.cpp:233
    if (w < 0 || h < 0) return fail("b200pt_resize: negative size");
    API_BEGIN
    enterCtxMutating(c);
    CUDA_OK(cudaStreamSynchronize(c->stream));
    const size_t n = (size_t)w * h;
    destroyPeer(c);                                                              // an exported / mapped frame does not survive a resize: export and attach again
    freeAccum(c);
    c->frameBuf.release();  c->normalBuf.release();  c->albedoBuf.release();  c->colorBuf.release();
    c->frameBuf.ensure(n);  c->normalBuf.ensure(n);  c->albedoBuf.ensure(n);  c->colorBuf.ensure(n);
    allocAccum(c, n);                                                            // ncclMemAlloc + registration once a communicator exists (api_multi.inl)
    // The reference's CUDABuffer::resize leaves the new buffers uninitialised; a render() that keeps accumulating across a resize (subframe_index > 0)
    // would then average against whatever the allocator handed out.  Zeroed here, so that sequence is at least deterministic.
    CUDA_OK(cudaMemsetAsync(c->frameBuf.p, 0, sizeof(uint32_t) * n, c->stream));  CUDA_OK(cudaMemsetAsync(c->accumBuf.p, 0, sizeof(float4) * n, c->stream));
    CUDA_OK(cudaMemsetAsync(c->normalBuf.p, 0, sizeof(float4) * n, c->stream));  CUDA_OK(cudaMemsetAsync(c->albedoBuf.p, 0, sizeof(float4) * n, c->stream));
    CUDA_OK(cudaMemsetAsync(c->colorBuf.p, 0, sizeof(float4) * n, c->stream));
    CUDA_OK(cudaStreamSynchronize(c->stream));
    c->W = w;  c->H = h;
    params->frame.size = make_int2(w, h);
    params->frame.frame_buffer = (uchar4*)c->frameBuf.p;  params->frame.accum_buffer = c->accumBuf.p;
    params->frame.normal_buffer = c->normalBuf.p;  params->frame.color_buffer = c->colorBuf.p;  params->frame.albedo_buffer = c->albedoBuf.p;
    API_END
}

int b200pt_set_probe(b200pt_ctx* c, int w, int h, const float* data, const float* pdfX, const float* cdfX, const float* pdfY,
                     const float* cdfY, const float* offset, LaunchParams* params) {
    if (!c || !params) return fail("b200pt_set_probe: null argument");
    if (w <= 0 || h <= 0 || !data || !pdfX || !cdfX || !pdfY || !cdfY) return fail("Probe Data is not valid");   // Probe.h:104-105
    API_BEGIN
    enterCtxMutating(c);
    const size_t n = (size_t)w * h;
    c->probeData.ensure(4 * n);  c->probePdfX.ensure(n);  c->probeCdfX.ensure(n);  c->probePdfY.ensure(h);  c->probeCdfY.ensure(h);
    CUDA_OK(cudaMemcpyAsync(c->probeData.p, data, 16 * n, cudaMemcpyHostToDevice, c->stream));
    CUDA_OK(cudaMemcpyAsync(c->probePdfX.p, pdfX, 4 * n, cudaMemcpyHostToDevice, c->stream));
    CUDA_OK(cudaMemcpyAsync(c->probeCdfX.p, cdfX, 4 * n, cudaMemcpyHostToDevice, c->stream));
    CUDA_OK(cudaMemcpyAsync(c->probePdfY.p, pdfY, 4 * (size_t)h, cudaMemcpyHostToDevice, c->stream));
    CUDA_OK(cudaMemcpyAsync(c->probeCdfY.p, cdfY, 4 * (size_t)h, cudaMemcpyHostToDevice, c->stream));
    c->probeGuide.ensure(probeGuideSize(h));
    launchProbeGuide(c->probeCdfX.p, c->probeCdfY.p, w, h, c->probeGuide.p, c->stream);
    c->probeW = w;  c->probeH = h;
    CUDA_OK(cudaStreamSynchronize(c->stream));
    params->probe.width = w;  params->probe.height = h;  params->probe.data = (Color*)c->probeData.p;
    params->probe.pdfValuesX = c->probePdfX.p;  params->probe.cdfValuesX = c->probeCdfX.p;
    params->probe.pdfValuesY = c->probePdfY.p;  params->probe.cdfValuesY = c->probeCdfY.p;
    params->probe.offset = offset ? make_float3(offset[0], offset[1], offset[2]) : make_float3(0.f, 0.f, 0.f);
    API_END
}

// Second integrator on the same traversal kernels: the Whitted-style frame of 04HelloRaytracing (deviceProgram.cu:66-243):
// raygen -> traceKernel (closest, tmin 0, tmax 1e20) -> whittedShade -> traceKernel (any hit, tmin 1e-3, tmax 1 - 1e-3, no culling)
// -> finish.  Output: W*H RGBA8 into `deviceColorBuffer`, or the context's frame buffer when NULL.
int b200pt_render_whitted(b200pt_ctx* c, const b200pt_whitted_params* wp, uint32_t* deviceColorBuffer) {
    if (!c || !wp) return fail("b200pt_render_whitted: null argument");
    API_BEGIN
    enterCtxMutating(c);
    if (!c->built) throw std::runtime_error("b200pt_render_whitted: b200pt_build_accel has not been called");
    if (c->W <= 0 || c->H <= 0) throw std::runtime_error("b200pt_render_whitted: frame buffers not allocated (call b200pt_resize)");
    const size_t n = (size_t)c->W * c->H;
    if (n >= ((size_t)1 << 31)) throw std::runtime_error("b200pt_render_whitted: frame too large");
    Lane& L = c->lanes[0];
    ensurePool(L, n, false);
    WhittedParams P;
    P.W = c->W;  P.H = c->H;
    P.position = make_float3(wp->position[0], wp->position[1], wp->position[2]);
    P.direction = make_float3(wp->direction[0], wp->direction[1], wp->direction[2]);
    P.horizontal = make_float3(wp->horizontal[0], wp->horizontal[1], wp->horizontal[2]);
    P.vertical = make_float3(wp->vertical[0], wp->vertical[1], wp->vertical[2]);
    P.lightPos = make_float3(wp->light_pos[0], wp->light_pos[1], wp->light_pos[2]);
    const SceneDev S = sceneOf(c);
    cudaStream_t s = c->stream;
    PathPool pool = poolOf(L);
    Queues Q = queuesOf(L);
    c->countersPending = false;
    CUDA_OK(cudaMemsetAsync(L.counters.p, 0, sizeof(unsigned) * (QC_STRIDE * QC_MAX_LAUNCHES + 8), s));
    unsigned* c0 = Q.counters;
    unsigned* c1 = Q.counters + QC_STRIDE;
    launchWhittedRaygen(P, Q, pool, c0, s);
    launchTrace(S, Q, 0, 1, c0, 0.f, 0.f, 1e20f, 0, pool, nullptr, s);                          // closest hits; no shadow rays yet
    launchWhittedShade(P, S, pool, Q, c1 + QC_SHADOWS, s);
    launchTrace(S, Q, 0, 0, c1, 1e-3f, 1e-3f, 1.f - 1e-3f, 0, pool, nullptr, s);                // shadow rays only
    launchWhittedFinish(P, pool, deviceColorBuffer ? deviceColorBuffer : (uint32_t*)c->frameBuf.p, s);
    CUDA_OK(cudaGetLastError());
    CUDA_OK(cudaStreamSynchronize(s));
    std::vector<unsigned> hc(2 * QC_STRIDE);
    CUDA_OK(cudaMemcpy(hc.data(), L.counters.p, sizeof(unsigned) * hc.size(), cudaMemcpyDeviceToHost));
    c->last = b200pt_counters{};
    c->last.samples = n;  c->last.radiance_rays = hc[QC_RAYS];  c->last.shadow_rays = hc[QC_STRIDE + QC_SHADOWS];
    c->last.kernel_launches = 5;  c->last.n_trace_closest_launches = 1;  c->last.n_trace_any_launches = 1;
    API_END
}

// loadProbe/loadColor + BuildCDF + setProbe in one call (sv4/main.cpp:154-180, Probe.h:29-77, SimplePathtracer.cpp:295-311): only the
// texels cross PCIe; the four tables are built on the device, bit-identical to the host's sequential BuildCDF.
int b200pt_set_probe_image(b200pt_ctx* c, int w, int h, const float* data, const float* offset, LaunchParams* params) {
    if (!c || !params) return fail("b200pt_set_probe_image: null argument");
    if (w <= 0 || h <= 0 || !data) return fail("Probe Data is not valid");
    API_BEGIN
    enterCtxMutating(c);
    const size_t n = (size_t)w * h;
    c->probeData.ensure(4 * n);  c->probePdfX.ensure(n);  c->probeCdfX.ensure(n);  c->probePdfY.ensure(h);  c->probeCdfY.ensure(h);
    c->probeGuide.ensure(probeGuideSize(h));
    c->probeRowTotal.ensure(h);
    CUDA_OK(cudaMemcpyAsync(c->probeData.p, data, 16 * n, cudaMemcpyHostToDevice, c->stream));
    launchProbeCdf((const float4*)c->probeData.p, w, h, c->probePdfX.p, c->probeCdfX.p, c->probePdfY.p, c->probeCdfY.p, c->probeRowTotal.p, c->stream);
    launchProbeGuide(c->probeCdfX.p, c->probeCdfY.p, w, h, c->probeGuide.p, c->stream);
    CUDA_OK(cudaGetLastError());
    c->probeW = w;  c->probeH = h;
    CUDA_OK(cudaStreamSynchronize(c->stream));
    params->probe.width = w;  params->probe.height = h;  params->probe.data = (Color*)c->probeData.p;
    params->probe.pdfValuesX = c->probePdfX.p;  params->probe.cdfValuesX = c->probeCdfX.p;
    params->probe.pdfValuesY = c->probePdfY.p;  params->probe.cdfValuesY = c->probeCdfY.p;
    params->probe.offset = offset ? make_float3(offset[0], offset[1], offset[2]) : make_float3(0.f, 0.f, 0.f);
    API_END
}
// the device-resident tables of the current probe, for hosts that want ProbeData's arrays back (and for the parity tests)
int b200pt_download_probe(b200pt_ctx* c, float* pdfX, float* cdfX, float* pdfY, float* cdfY) {
    if (!c) return fail("b200pt_download_probe: null context");
    if (!c->probeW || !c->probeH) return fail("b200pt_download_probe: no probe set");
    API_BEGIN
    enterCtx(c);
    const size_t n = (size_t)c->probeW * c->probeH;
    if (pdfX) CUDA_OK(cudaMemcpyAsync(pdfX, c->probePdfX.p, 4 * n, cudaMemcpyDeviceToHost, c->stream));
    if (cdfX) CUDA_OK(cudaMemcpyAsync(cdfX, c->probeCdfX.p, 4 * n, cudaMemcpyDeviceToHost, c->stream));
    if (pdfY) CUDA_OK(cudaMemcpyAsync(pdfY, c->probePdfY.p, 4 * (size_t)c->probeH, cudaMemcpyDeviceToHost, c->stream));
    if (cdfY) CUDA_OK(cudaMemcpyAsync(cdfY, c->probeCdfY.p, 4 * (size_t)c->probeH, cudaMemcpyDeviceToHost, c->stream));
    CUDA_OK(cudaStreamSynchronize(c->stream));
    API_END
}

int b200pt_render(b200pt_ctx* c, LaunchParams* params, const b200pt_config* cfg) {
    API_BEGIN
    renderImpl(c, params, cfg, true);
    API_END
}
int b200pt_render_async(b200pt_ctx* c, LaunchParams* params, const b200pt_config* cfg) {
    API_BEGIN
    renderImpl(c, params, cfg, false);
    API_END
}
int b200pt_synchronize(b200pt_ctx* c) {
    if (!c) return fail("b200pt_synchronize: null context");
    API_BEGIN
    enterCtx(c);                                                                 // orders the primary stream after every lane in flight
    CUDA_OK(cudaStreamSynchronize(c->stream));
    collectCounters(c);
    API_END
}
int b200pt_join(b200pt_ctx* c) {
    if (!c) return fail("b200pt_join: null context");
    API_BEGIN
    enterCtx(c);
    API_END
}

}  // extern "C"
