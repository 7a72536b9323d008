// b200pt — downloads, epilogues and parity hooks (included by api.cu).
extern "C" {

int b200pt_download_pixels(b200pt_ctx* c, uint32_t* host) {
    if (!c || !host) return fail("b200pt_download_pixels: null argument");
    API_BEGIN
    enterCtx(c);
    CUDA_OK(cudaMemcpyAsync(host, c->frameBuf.p, sizeof(uint32_t) * (size_t)c->W * c->H, cudaMemcpyDeviceToHost, c->stream));
    CUDA_OK(cudaStreamSynchronize(c->stream));
    API_END
}
// downloadPixels for frames in flight: the copy is enqueued right behind the finalize pass of the most recently submitted frame, on that
// frame's own stream, and the frame's `done` event moves behind it - so the next frame (whose finalize overwrites the frame buffer) waits for
// the copy while every other lane keeps tracing.  `host` should be pinned; it is valid after b200pt_synchronize.
int b200pt_download_pixels_async(b200pt_ctx* c, uint32_t* host) {
    if (!c || !host) return fail("b200pt_download_pixels_async: null argument");
    API_BEGIN
    CUDA_OK(cudaSetDevice(c->device));
    const int k = c->lastLane < 0 ? 0 : c->lastLane;
    Lane& L = c->lanes[k];
    if (!L.stream) throw std::runtime_error("b200pt_download_pixels_async: no frame has been rendered");
    CUDA_OK(cudaMemcpyAsync(host, c->frameBuf.p, sizeof(uint32_t) * (size_t)c->W * c->H, cudaMemcpyDeviceToHost, L.stream));
    if (L.done) CUDA_OK(cudaEventRecord(L.done, L.stream));
    L.busy = k > 0;                                  // the copy is work the primary stream has not been ordered after: b200pt_synchronize / b200pt_join must wait for it
    API_END
}
int b200pt_download_accum(b200pt_ctx* c, float* host) {
    if (!c || !host) return fail("b200pt_download_accum: null argument");
    API_BEGIN
    enterCtx(c);
    CUDA_OK(cudaMemcpyAsync(host, c->accumBuf.p, sizeof(float4) * (size_t)c->W * c->H, cudaMemcpyDeviceToHost, c->stream));
    CUDA_OK(cudaStreamSynchronize(c->stream));
    API_END
}
int b200pt_upload_accum(b200pt_ctx* c, const float* host) {
    if (!c || !host) return fail("b200pt_upload_accum: null argument");
    API_BEGIN
    enterCtxMutating(c);
    CUDA_OK(cudaMemcpyAsync(c->accumBuf.p, host, sizeof(float4) * (size_t)c->W * c->H, cudaMemcpyHostToDevice, c->stream));
    CUDA_OK(cudaStreamSynchronize(c->stream));
    API_END
}
int b200pt_download_aov(b200pt_ctx* c, int which, float* host) {
    if (!c || !host || which < 0 || which > 2) return fail("b200pt_download_aov: bad argument");
    API_BEGIN
    enterCtx(c);
    const float4* src = which == 0 ? c->normalBuf.p : (which == 1 ? c->albedoBuf.p : c->colorBuf.p);
    CUDA_OK(cudaMemcpyAsync(host, src, sizeof(float4) * (size_t)c->W * c->H, cudaMemcpyDeviceToHost, c->stream));
    CUDA_OK(cudaStreamSynchronize(c->stream));
    API_END
}
int b200pt_device_buffers(b200pt_ctx* c, void** accum, void** frame, void** stream) {
    if (!c) return fail("b200pt_device_buffers: null context");
    if (accum) *accum = c->accumBuf.p;
    if (frame) *frame = c->frameBuf.p;
    if (stream) *stream = (void*)c->stream;
    return 0;
}
int b200pt_clear_accum(b200pt_ctx* c) {
    if (!c) return fail("b200pt_clear_accum: null context");
    API_BEGIN
    enterCtxMutating(c);
    if (c->accumBuf.p) CUDA_OK(cudaMemsetAsync(c->accumBuf.p, 0, sizeof(float4) * (size_t)c->W * c->H, c->stream));
    if (c->frameBuf.p) CUDA_OK(cudaMemsetAsync(c->frameBuf.p, 0, sizeof(uint32_t) * (size_t)c->W * c->H, c->stream));
    API_END
}
int b200pt_resolve_sum(b200pt_ctx* c, const LaunchParams* params, const b200pt_config* cfg, int nPasses) {
    if (!c || !params || !cfg || nPasses <= 0) return fail("b200pt_resolve_sum: bad argument");
    API_BEGIN
    enterCtxMutating(c);
    launchResolveSum(frameOf(*params, *cfg), nPasses, c->stream);
    CUDA_OK(cudaStreamSynchronize(c->stream));
    API_END
}

// BuildCDF on caller-owned device buffers with the kernels' scratch in a static __device__ array (shade.cu): no allocation anywhere on this path.
int b200pt_build_cdf_noalloc(b200pt_ctx* c, int w, int h, const void* data, void* pdfX, void* cdfX, void* pdfY, void* cdfY, void* stream) {
    if (!c) return fail("b200pt_build_cdf_noalloc: null context");
    if (w <= 0 || h <= 0 || !data || !pdfX || !cdfX || !pdfY || !cdfY) return fail("Probe Data is not valid");
    if (h > B200PT_NOALLOC_MAX_ROWS)
        return fail("b200pt_build_cdf_noalloc: height " + std::to_string(h) + " exceeds the static scratch (B200PT_NOALLOC_MAX_ROWS = " +
                    std::to_string(B200PT_NOALLOC_MAX_ROWS) + " rows); use b200pt_set_probe_image");
    if ((long long)w * h >= (1ll << 31)) return fail("b200pt_build_cdf_noalloc: probe too large");
    API_BEGIN
    CUDA_OK(cudaSetDevice(c->device));
    launchProbeCdfNoAlloc((const float4*)data, w, h, (float*)pdfX, (float*)cdfX, (float*)pdfY, (float*)cdfY, stream ? (cudaStream_t)stream : c->stream);
    CUDA_OK(cudaPeekAtLastError());
    API_END
}

int b200pt_tonemap_sqrt(int w, int h, const void* in, void* out, void* stream) {
    if (w < 0 || h < 0 || ((size_t)w * h && (!in || !out))) return fail("b200pt_tonemap_sqrt: bad argument");
    API_BEGIN
    launchTonemapSqrt(w, h, (const float4*)in, (uint32_t*)out, (cudaStream_t)stream);
    CUDA_OK(cudaGetLastError());
    API_END
}
// the reference's one C-linkage export (sv4/toneMap.h:4): async launch on the default stream, no error return
struct int2;
void computeFinalPixelColors(const int2& fbSize, float4* denoisedBuffer, uint32_t* finalColorBuffer) {
    launchTonemapSqrt(fbSize.x, fbSize.y, denoisedBuffer, finalColorBuffer, 0);
}
int b200pt_tonemap_sqrt_host(b200pt_ctx* c, int w, int h, const float* in, uint32_t* out) {
    if (!c || w < 0 || h < 0) return fail("b200pt_tonemap_sqrt_host: bad argument");
    API_BEGIN
    const size_t n = (size_t)w * h;
    if (!n) return 0;
    enterCtx(c);
    DevBuf<float4> din;  DevBuf<uint32_t> dout;
    din.ensure(n);  dout.ensure(n);
    CUDA_OK(cudaMemcpyAsync(din.p, in, 16 * n, cudaMemcpyHostToDevice, c->stream));
    launchTonemapSqrt(w, h, din.p, dout.p, c->stream);
    CUDA_OK(cudaMemcpyAsync(out, dout.p, 4 * n, cudaMemcpyDeviceToHost, c->stream));
    CUDA_OK(cudaStreamSynchronize(c->stream));
    din.release();  dout.release();
    API_END
}

int b200pt_trace_primary(b200pt_ctx* c, const LaunchParams* params, const b200pt_config* cfg, int32_t* prim, int32_t* mesh,
                         float* t, float* u, float* v) {
    if (!c || !params || !cfg || !prim || !mesh || !t || !u || !v) return fail("b200pt_trace_primary: null argument");
    API_BEGIN
    if (!c->built) throw std::runtime_error("b200pt_trace_primary: b200pt_build_accel has not been called");
    enterCtx(c);
    LaunchParams P = *params;
    std::vector<PassDesc> passes;
    b200pt_config one = *cfg;
    if (one.schedule == B200PT_SCHEDULE_FOVEATED) one.schedule = B200PT_SCHEDULE_AS_GIVEN;
    buildSchedule(P, one, passes);
    const PassDesc& pd = passes[0];
    const size_t n = pd.nLaunch;
    if (!n) return 0;
    DevBuf<float4> ro, rd, hit;  DevBuf<int32_t> dp, dm;  DevBuf<float> dt, du, dv;
    ro.ensure(n);  rd.ensure(n);  hit.ensure(n);  dp.ensure(n);  dm.ensure(n);  dt.ensure(n);  du.ensure(n);  dv.ensure(n);
    const FrameParams F = frameOf(P, *cfg);
    const SceneDev S = sceneOf(c);
    launchPrimaryRays(F, pd.P, ro.p, rd.p, c->stream);
    launchTraceRaysDebug(S, ro.p, rd.p, (unsigned)n, cfg->tmin, cfg->tmax, 0, 0, hit.p, c->stream);
    launchResolvePrimary(S, hit.p, (unsigned)n, ro.p, dp.p, dm.p, dt.p, du.p, dv.p, c->stream);
    CUDA_OK(cudaMemcpyAsync(prim, dp.p, 4 * n, cudaMemcpyDeviceToHost, c->stream));
    CUDA_OK(cudaMemcpyAsync(mesh, dm.p, 4 * n, cudaMemcpyDeviceToHost, c->stream));
    CUDA_OK(cudaMemcpyAsync(t, dt.p, 4 * n, cudaMemcpyDeviceToHost, c->stream));
    CUDA_OK(cudaMemcpyAsync(u, du.p, 4 * n, cudaMemcpyDeviceToHost, c->stream));
    CUDA_OK(cudaMemcpyAsync(v, dv.p, 4 * n, cudaMemcpyDeviceToHost, c->stream));
    CUDA_OK(cudaStreamSynchronize(c->stream));
    CUDA_OK(cudaGetLastError());
    ro.release();  rd.release();  hit.release();  dp.release();  dm.release();  dt.release();  du.release();  dv.release();
    API_END
}

int b200pt_trace_rays(b200pt_ctx* c, const float* org, const float* dir, size_t n, float tmin, float tmax, int anyHit, int shadowMode,
                      int32_t* prim, int32_t* mesh, float* t, float* u, float* v) {
    if (!c || (n && (!org || !dir || !prim))) return fail("b200pt_trace_rays: null argument");
    API_BEGIN
    if (!c->built) throw std::runtime_error("b200pt_trace_rays: b200pt_build_accel has not been called");
    if (!n) return 0;
    enterCtx(c);
    std::vector<float4> ho(n), hd(n);
    for (size_t i = 0; i < n; ++i) { ho[i] = make_float4(org[3 * i], org[3 * i + 1], org[3 * i + 2], 1.f);  hd[i] = make_float4(dir[3 * i], dir[3 * i + 1], dir[3 * i + 2], 0.f); }
    DevBuf<float4> ro, rd, hit;  DevBuf<int32_t> dp, dm;  DevBuf<float> dt, du, dv;
    ro.ensure(n);  rd.ensure(n);  hit.ensure(n);  dp.ensure(n);  dm.ensure(n);  dt.ensure(n);  du.ensure(n);  dv.ensure(n);
    CUDA_OK(cudaMemcpyAsync(ro.p, ho.data(), 16 * n, cudaMemcpyHostToDevice, c->stream));
    CUDA_OK(cudaMemcpyAsync(rd.p, hd.data(), 16 * n, cudaMemcpyHostToDevice, c->stream));
    const SceneDev S = sceneOf(c);
    launchTraceRaysDebug(S, ro.p, rd.p, (unsigned)n, tmin, tmax, anyHit, shadowMode == B200PT_SHADOW_CULL_BACK, hit.p, c->stream);
    if (anyHit) {
        std::vector<float4> hh(n);
        CUDA_OK(cudaMemcpyAsync(hh.data(), hit.p, 16 * n, cudaMemcpyDeviceToHost, c->stream));
        CUDA_OK(cudaStreamSynchronize(c->stream));
        for (size_t i = 0; i < n; ++i) { int f;  memcpy(&f, &hh[i].w, 4);  prim[i] = f; }
    } else {
        if (!mesh || !t || !u || !v) throw std::runtime_error("b200pt_trace_rays: null output");
        launchResolvePrimary(S, hit.p, (unsigned)n, nullptr, dp.p, dm.p, dt.p, du.p, dv.p, c->stream);
        CUDA_OK(cudaMemcpyAsync(prim, dp.p, 4 * n, cudaMemcpyDeviceToHost, c->stream));
        CUDA_OK(cudaMemcpyAsync(mesh, dm.p, 4 * n, cudaMemcpyDeviceToHost, c->stream));
        CUDA_OK(cudaMemcpyAsync(t, dt.p, 4 * n, cudaMemcpyDeviceToHost, c->stream));
        CUDA_OK(cudaMemcpyAsync(u, du.p, 4 * n, cudaMemcpyDeviceToHost, c->stream));
        CUDA_OK(cudaMemcpyAsync(v, dv.p, 4 * n, cudaMemcpyDeviceToHost, c->stream));
        CUDA_OK(cudaStreamSynchronize(c->stream));
    }
    CUDA_OK(cudaGetLastError());
    ro.release();  rd.release();  hit.release();  dp.release();  dm.release();  dt.release();  du.release();  dv.release();
    API_END
}

// Device-level known-answer hooks for the shading library (tests/golden/bsdf.npz, probe.npz): host arrays in, host arrays out.
int b200pt_debug_bsdf(b200pt_ctx* c, size_t n, const Material* materials, const float* etaI, const float* etaO, const float* N, const float* view,
                      const float* L, const float* albedo, const uint32_t* seeds, float* basisU, float* basisV, float* sampleL, float* samplePdf,
                      int32_t* sampleType, uint32_t* sampleRng, float* evalF, float* evalPdf) {
    if (!c || (n && (!materials || !etaI || !etaO || !N || !view || !L || !albedo || !seeds || !basisU || !basisV || !sampleL || !samplePdf || !sampleType || !sampleRng || !evalF || !evalPdf)))
        return fail("b200pt_debug_bsdf: null argument");
    API_BEGIN
    if (!n) return 0;
    enterCtx(c);
    cudaStream_t s = c->stream;
    DevBuf<Material> dm;  DevBuf<float> dEtaI, dEtaO, dPdf, dEvalPdf;  DevBuf<float3> dN, dView, dL, dAlb, dU, dV, dSL, dF;  DevBuf<unsigned> dSeeds;  DevBuf<int> dType;  DevBuf<uint2> dRng;
    dm.ensure(n);  dEtaI.ensure(n);  dEtaO.ensure(n);  dPdf.ensure(n);  dEvalPdf.ensure(n);  dN.ensure(n);  dView.ensure(n);  dL.ensure(n);  dAlb.ensure(n);
    dU.ensure(n);  dV.ensure(n);  dSL.ensure(n);  dF.ensure(n);  dSeeds.ensure(n);  dType.ensure(n);  dRng.ensure(n);
    auto up = [&](void* d, const void* h, size_t bytes) { CUDA_OK(cudaMemcpyAsync(d, h, bytes, cudaMemcpyHostToDevice, s)); };
    auto down = [&](void* h, const void* d, size_t bytes) { CUDA_OK(cudaMemcpyAsync(h, d, bytes, cudaMemcpyDeviceToHost, s)); };
    up(dm.p, materials, n * sizeof(Material));  up(dEtaI.p, etaI, 4 * n);  up(dEtaO.p, etaO, 4 * n);  up(dN.p, N, 12 * n);  up(dView.p, view, 12 * n);
    up(dL.p, L, 12 * n);  up(dAlb.p, albedo, 12 * n);  up(dSeeds.p, seeds, 4 * n);
    launchDebugBsdf((unsigned)n, dm.p, dEtaI.p, dEtaO.p, dN.p, dView.p, dL.p, dAlb.p, dSeeds.p, dU.p, dV.p, dSL.p, dPdf.p, dType.p, dRng.p, dF.p, dEvalPdf.p, s);
    CUDA_OK(cudaGetLastError());
    down(basisU, dU.p, 12 * n);  down(basisV, dV.p, 12 * n);  down(sampleL, dSL.p, 12 * n);  down(samplePdf, dPdf.p, 4 * n);  down(sampleType, dType.p, 4 * n);
    down(sampleRng, dRng.p, 8 * n);  down(evalF, dF.p, 12 * n);  down(evalPdf, dEvalPdf.p, 4 * n);
    CUDA_OK(cudaStreamSynchronize(s));
    dm.release();  dEtaI.release();  dEtaO.release();  dPdf.release();  dEvalPdf.release();  dN.release();  dView.release();  dL.release();  dAlb.release();
    dU.release();  dV.release();  dSL.release();  dF.release();  dSeeds.release();  dType.release();  dRng.release();
    API_END
}
int b200pt_debug_probe(b200pt_ctx* c, const LaunchParams* params, int useGuide, size_t n, const uint32_t* seeds, float* sampleDir, float* sampleColor,
                       float* samplePdf, size_t m, const float* dirs, float* uv, float* evalColor, float* evalPdf) {
    if (!c || !params) return fail("b200pt_debug_probe: null argument");
    if ((n && (!seeds || !sampleDir || !sampleColor || !samplePdf)) || (m && (!dirs || !uv || !evalColor || !evalPdf))) return fail("b200pt_debug_probe: null array");
    if (!params->probe.data || !params->probe.cdfValuesX) return fail("b200pt_debug_probe: no probe set (call b200pt_set_probe)");
    API_BEGIN
    if (!n && !m) return 0;
    enterCtx(c);
    cudaStream_t s = c->stream;
    b200pt_config cfg;
    memset(&cfg, 0, sizeof(cfg));
    const FrameParams F = frameOf(*params, cfg, c);                 // resolves the guide tables exactly as render() does
    if (useGuide && !F.probeGuide) throw std::runtime_error("b200pt_debug_probe: the probe in *params is not the one uploaded by b200pt_set_probe*, no guide tables");
    DevBuf<unsigned> dSeeds;  DevBuf<float3> dDir, dCol, dDirs, dECol;  DevBuf<float> dPdf, dEPdf;  DevBuf<float2> dUv;
    dSeeds.ensure(std::max<size_t>(n, 1));  dDir.ensure(std::max<size_t>(n, 1));  dCol.ensure(std::max<size_t>(n, 1));  dPdf.ensure(std::max<size_t>(n, 1));
    dDirs.ensure(std::max<size_t>(m, 1));  dECol.ensure(std::max<size_t>(m, 1));  dEPdf.ensure(std::max<size_t>(m, 1));  dUv.ensure(std::max<size_t>(m, 1));
    if (n) CUDA_OK(cudaMemcpyAsync(dSeeds.p, seeds, 4 * n, cudaMemcpyHostToDevice, s));
    if (m) CUDA_OK(cudaMemcpyAsync(dDirs.p, dirs, 12 * m, cudaMemcpyHostToDevice, s));
    launchDebugProbe(F.probe, useGuide ? F.probeGuide : nullptr, (unsigned)n, dSeeds.p, dDir.p, dCol.p, dPdf.p, (unsigned)m, dDirs.p, dUv.p, dECol.p, dEPdf.p, s);
    CUDA_OK(cudaGetLastError());
    if (n) { CUDA_OK(cudaMemcpyAsync(sampleDir, dDir.p, 12 * n, cudaMemcpyDeviceToHost, s));  CUDA_OK(cudaMemcpyAsync(sampleColor, dCol.p, 12 * n, cudaMemcpyDeviceToHost, s));
             CUDA_OK(cudaMemcpyAsync(samplePdf, dPdf.p, 4 * n, cudaMemcpyDeviceToHost, s)); }
    if (m) { CUDA_OK(cudaMemcpyAsync(uv, dUv.p, 8 * m, cudaMemcpyDeviceToHost, s));  CUDA_OK(cudaMemcpyAsync(evalColor, dECol.p, 12 * m, cudaMemcpyDeviceToHost, s));
             CUDA_OK(cudaMemcpyAsync(evalPdf, dEPdf.p, 4 * m, cudaMemcpyDeviceToHost, s)); }
    CUDA_OK(cudaStreamSynchronize(s));
    dSeeds.release();  dDir.release();  dCol.release();  dPdf.release();  dDirs.release();  dECol.release();  dEPdf.release();  dUv.release();
    API_END
}

int b200pt_debug_morton(b200pt_ctx* c, uint64_t* codes, uint32_t* order, float* bounds) {
    if (!c || !codes || !order || !bounds) return fail("b200pt_debug_morton: null argument");
    API_BEGIN
    if (!c->built) throw std::runtime_error("b200pt_debug_morton: b200pt_build_accel has not been called");
    enterCtx(c);
    const size_t n = c->bvh.nTris;
    if (n) {
        CUDA_OK(cudaMemcpy(codes, c->bvh.mortonByGid, 8 * n, cudaMemcpyDeviceToHost));
        CUDA_OK(cudaMemcpy(order, c->bvh.sortedGid, 4 * n, cudaMemcpyDeviceToHost));
    }
    for (int a = 0; a < 6; ++a) bounds[a] = c->bvh.sceneBounds[a];
    API_END
}

int b200pt_bvh_info_get(b200pt_ctx* c, b200pt_bvh_info* o) {
    if (!c || !o) return fail("b200pt_bvh_info_get: null argument");
    if (!c->built) return fail("b200pt_bvh_info_get: b200pt_build_accel has not been called");
    memset(o, 0, sizeof(*o));
    o->n_triangles = c->bvh.nTris;  o->n_nodes = c->bvh.nNodes;  o->n_leaves = c->bvh.nLeaves;  o->max_leaf_size = c->bvh.maxLeaf;
    o->max_depth = c->bvh.maxDepth;  o->sah_cost = c->bvh.sahCost;  o->sah_cost_lbvh = c->bvh.sahCostLbvh;  o->build_ms = c->bvh.buildMs;
    o->node_bytes = (uint64_t)c->bvh.nNodes4 * sizeof(BvhNode4);  o->n_wide_nodes = c->bvh.nNodes4;  o->wide_depth = c->bvh.maxDepth4;  o->tri_bytes = (uint64_t)c->bvh.nTris * 48;
    o->refit_ms = c->refitMs;
    o->root_link = c->bvh.rootLink;  o->refine_levels = c->bvh.refineLevels;
    for (int a = 0; a < 3; ++a) { o->grid_lo[a] = c->bvh.gridLo[a];  o->grid_scale[a] = c->bvh.gridScale[a]; }
    return 0;
}

int b200pt_download_bvh(b200pt_ctx* c, void* nodes32, size_t maxNodes, uint32_t* triGid, size_t maxTris) {
    if (!c || !nodes32 || !triGid) return fail("b200pt_download_bvh: null argument");
    API_BEGIN
    if (!c->built) throw std::runtime_error("b200pt_download_bvh: b200pt_build_accel has not been called");
    if (maxNodes < c->bvh.nNodes || maxTris < c->bvh.nTris) throw std::runtime_error("b200pt_download_bvh: buffers too small");
    enterCtx(c);
    if (c->bvh.nNodes) CUDA_OK(cudaMemcpy(nodes32, c->bvh.nodes, sizeof(BvhNode) * (size_t)c->bvh.nNodes, cudaMemcpyDeviceToHost));
    if (c->bvh.nTris) CUDA_OK(cudaMemcpy(triGid, c->bvh.finalGid, 4 * (size_t)c->bvh.nTris, cudaMemcpyDeviceToHost));
    API_END
}

int b200pt_download_bvh4(b200pt_ctx* c, void* nodes64, size_t maxNodes) {
    if (!c || !nodes64) return fail("b200pt_download_bvh4: null argument");
    API_BEGIN
    if (!c->built) throw std::runtime_error("b200pt_download_bvh4: b200pt_build_accel has not been called");
    if (maxNodes < c->bvh.nNodes4) throw std::runtime_error("b200pt_download_bvh4: buffer too small");
    enterCtx(c);
    if (c->bvh.nNodes4) CUDA_OK(cudaMemcpy(nodes64, c->bvh.nodes4, sizeof(BvhNode4) * (size_t)c->bvh.nNodes4, cudaMemcpyDeviceToHost));
    API_END
}

int b200pt_get_counters(b200pt_ctx* c, b200pt_counters* o) {
    if (!c || !o) return fail("b200pt_get_counters: null argument");
    API_BEGIN
    enterCtx(c);
    if (c->countersPending) { CUDA_OK(cudaStreamSynchronize(c->stream));  collectCounters(c); }
    *o = c->last;
    API_END
}

}  // extern "C"
