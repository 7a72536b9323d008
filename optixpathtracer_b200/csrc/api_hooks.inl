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
