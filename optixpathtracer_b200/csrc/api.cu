// b200pt — the C-ABI (include/b200pt.h) over the CUDA kernels.  Host orchestration only: scene staging,
// the LBVH build, the wavefront loop of one render() and the parity hooks.  No CPU fallback anywhere.
#include "common.cuh"
#include <algorithm>
#include <deque>
#include <cstring>
#include <cstdlib>
#include <nvtx3/nvToolsExt.h>

static thread_local std::string g_lastError;
static int fail(const std::string& msg) { g_lastError = msg;  return -1; }
#define API_BEGIN try {
#define API_END } catch (const std::exception& e) { return fail(e.what()); } return 0;

template <typename T> struct DevBuf {
    T* p = nullptr;  size_t n = 0;
    void ensure(size_t count) { if (count > n) { release();  CUDA_OK(cudaMalloc((void**)&p, count * sizeof(T)));  n = count; } }
    void release() { if (p) cudaFree(p);  p = nullptr;  n = 0; }
};

struct HostMesh { size_t firstVertex, nVertices, firstTri, nTris;  bool hasTexcoord, hasNormal;  int textureID;  Material material{Material::NoInit{}}; };
struct HostTexture { size_t offset;  int w, h; };

// Everything one frame in flight owns: its stream, the path pool, the ray queues, the per-launch counters and the per-launch-index sums.
// With cfg.frames_in_flight <= 1 every frame runs on lane 0, the context's primary stream; otherwise frames rotate over lanes 1..frames_in_flight,
// so that the drain phases and the thin late bounces of one frame overlap the wide early bounces of the next - and so that the primary stream, where
// downloads, uploads and the caller's own work are ordered, never queues behind a wavefront that is still speculative.  Frames still fold into the accumulation buffer in submission order:
// each frame's finalize pass waits for the `done` event of the frame before it.
#define B200PT_MAX_FRAMES_IN_FLIGHT 4
#define B200PT_MAX_LANES (B200PT_MAX_FRAMES_IN_FLIGHT + 1)      // lane 0 = the primary stream: frames that run one at a time; lanes 1.. = frames in flight
struct Lane {
    cudaStream_t stream = nullptr;
    size_t poolCap = 0;
    DevBuf<PathState> st;  DevBuf<float4> direct, indirect, alpha, aovN, aovA;
    DevBuf<float4> rayO[2], rayD[2], hit, shO, shD, shP;
    DevBuf<unsigned> counters;          // QC_COUNT queue counters + 8 running totals
    DevBuf<float4> sumRad, sumAlpha, sumNormal, sumAlbedo;
    DevBuf<TraceStats> stats;           // [0] closest, [1] any
    cudaEvent_t done = nullptr;         // recorded after the last kernel of the lane's latest frame
    cudaEvent_t headStart = nullptr;    // recorded behind pass B200PT_STAGGER of the lane's latest wavefront: the next wavefront (on another lane) starts behind it
    bool busy = false;                  // has work the primary stream has not been ordered after yet
    bool forked = false;                // already ordered after the primary stream's earlier work (since the last join)
    void release() {
        st.release();  direct.release();  indirect.release();  alpha.release();  aovN.release();  aovA.release();
        for (int i = 0; i < 2; ++i) { rayO[i].release();  rayD[i].release(); }
        hit.release();  shO.release();  shD.release();  shP.release();  counters.release();
        sumRad.release();  sumAlpha.release();  sumNormal.release();  sumAlbedo.release();  stats.release();
        poolCap = 0;
    }
};

struct PassDesc { PassParams P;  size_t nLaunch; };
// One frame's wavefront as enqueued on a lane: everything up to, but not including, its finalize passes (the only stage that touches the
// shared accumulation / frame / AOV buffers).
// `lane` runs the wavefront - or, for a frame split into interleaved band sets (cfg.intra_frame_split), part 0 of it, with the other parts on `extraLanes`;
// the finalize passes always run on `lane`, whose sumRad/sumAlpha buffers collect the per-launch-index sums of every part.
// `fused.n > 0`: every sample of the frame (of each band set) sits in one chunk of its lane's pool, so no gather pass ran: the finalize passes sum the samples from
// those pools themselves.
struct FrameJob { int lane = -1;  std::vector<int> extraLanes;  std::vector<PassDesc> passes;  FrameParams F;  unsigned launches = 0, nClosest = 0;  cudaEvent_t totalEnd = nullptr;
                  FusedPools fused{}; };
// Speculative next sub-frame (blocking b200pt_render with cfg.frames_in_flight > 1).  Progressive accumulation calls render() again and again with
// the same launchParams and the next subframe_index; once a call has been predicted right, render() enqueues the wavefront of the
// PREDICTED next call(s) on the other lanes before it blocks on the current frame.  The next call either matches the prediction byte for byte
// (LaunchParams and config) - then only the finalize passes remain to be enqueued - or the speculative wavefront is dropped: it has touched
// nothing but its own lane.  Pixels are those of plain blocking calls, bit for bit; a moving camera just never gets a prediction right.
struct SpecFrame { LaunchParams input, output;  FrameJob job; };   // a predicted call: its launchParams on entry / as render() leaves them, and its wavefront
struct Speculation {
    std::deque<SpecFrame> ahead;        // wavefronts already enqueued for the next calls, oldest first (at most frames_in_flight - 1)
    bool havePrev = false;  unsigned prevIndex = 0, stride = 1;   // subframe_index of the previous blocking call (prediction = last index + stride)
    bool havePrediction = false;  LaunchParams predicted;  b200pt_config predictedCfg;  int hits = 0;   // consecutive calls that matched their prediction
    void drop() { ahead.clear();  havePrediction = false;  hits = 0; }
};

struct MultiState;                    // api_multi.inl: NCCL communicator + registrations of one rank
// One-sided tile split (api_peer.inl): what a context knows about an exported frame (root) or about its mapping of root's frame (the other ranks)
struct PeerState {
    unsigned* flags = nullptr;            // exporter: [r] = last frame epoch rank r has put, [PEER_ACK] = last epoch root has consumed (its own device memory)
    bool exported = false;  unsigned acquired = 0, released = 0;
    uint32_t* remoteFrame = nullptr;  unsigned* remoteFlags = nullptr;    // attached side: root's frame buffer and flag block, mapped over NVLink
    bool viaIpc = false;  int rank = -1, nranks = 0, W = 0, H = 0;  unsigned put = 0;
    unsigned* err = nullptr;              // device word: 0 = fine, else the code of the wait that timed out
    unsigned long long timeoutNs = 5000000000ull;
};

struct b200pt_ctx {
    int device = 0;
    MultiState* multi = nullptr;
    PeerState* peer = nullptr;
    bool accumFromNccl = false;       // accumBuf.p came from ncclMemAlloc (freed with ncclMemFree)
    cudaStream_t stream = nullptr;
    bool stale = false;           // vertices updated since the last build/refit
    float refitMs = 0.f;
    // --- scene staging (host) and device copies
    std::vector<float3> hVertices, hNormals;  std::vector<float2> hTexcoords;  std::vector<uint3> hIndices;  std::vector<unsigned> hTriMesh;
    std::vector<HostMesh> hMeshes;  std::vector<HostTexture> hTextures;  std::vector<uint32_t> hTexels;
    DevBuf<float3> dVertices, dNormals;  DevBuf<float2> dTexcoords;  DevBuf<uint3> dIndices;  DevBuf<unsigned> dTriMesh;
    DevBuf<MeshInfo> dMeshes;  DevBuf<uint32_t> dTexels;
    BvhBuildResult bvh;  bool built = false;  bool hasCatcher = false;
    // --- frame
    int W = 0, H = 0;
    DevBuf<uint32_t> frameBuf;  DevBuf<float4> accumBuf, normalBuf, albedoBuf, colorBuf;
    DevBuf<float> probeData, probePdfX, probeCdfX, probePdfY, probeCdfY, probeRowTotal;  DevBuf<int> probeGuide;  int probeW = 0, probeH = 0;
    // --- wavefront state: one Lane per frame in flight (lane 0 runs on `stream`, the reference's single stream)
    Lane lanes[B200PT_MAX_LANES];
    int wavePrevLane = -1;              // lane of the most recently enqueued WAVEFRONT (speculative ones included): the next one is staggered behind it
    bool staggerWaves = false;          // set by the blocking render() around its enqueues: batches submitted with b200pt_render_async are not held back
    int lastLane = -1;                  // lane of the most recently enqueued frame (its `done` event orders the next frame's accumulation)
    int countersLane = 0;               // lane whose device counters `last` will be read from ...
    std::vector<int> countersExtra;     // ... plus the lanes of the other parts of a split frame
    unsigned seq = 0;                   // round-robin lane pick of b200pt_render_async
    cudaEvent_t forkEv = nullptr;
    Speculation spec;
    b200pt_counters last{};
    bool countersPending = false;  int profiled = 0;
    std::vector<cudaEvent_t> evPool;  std::vector<std::pair<int, std::pair<cudaEvent_t, cudaEvent_t>>> evUsed;  size_t evNext = 0;
};

static void freeAccum(b200pt_ctx* c);                 // api_multi.inl
static void allocAccum(b200pt_ctx* c, size_t n);
static void destroyMulti(b200pt_ctx* c);
static void destroyPeer(b200pt_ctx* c);               // api_peer.inl
static int peerPutBegin(b200pt_ctx* c, const b200pt_config& cfg, FrameParams& F, cudaStream_t s);
static void peerPassBarrier(b200pt_ctx* c, int role, int k, cudaStream_t s);
static void peerPutEnd(b200pt_ctx* c, int role, cudaStream_t s);

// Orders the primary stream after every frame in flight on the other lanes (no host synchronisation): what the entry points that touch
// shared state (frame buffers, scene, probe, counters) do first, and what b200pt_join exposes to callers that time or copy on that stream.
static void joinLanes(b200pt_ctx* c) {
    for (int k = 1; k < B200PT_MAX_LANES; ++k) {
        Lane& L = c->lanes[k];
        if (L.busy) CUDA_OK(cudaStreamWaitEvent(c->stream, L.done, 0));
        L.busy = false;  L.forked = false;
    }
}
static void enterCtx(b200pt_ctx* c) { CUDA_OK(cudaSetDevice(c->device));  joinLanes(c); }
// entry points that change what a frame reads or writes (scene, probe, frame buffers, lane 0's pool) also drop the speculative next sub-frame
static void enterCtxMutating(b200pt_ctx* c) {
    enterCtx(c);
    // Every lane, not only those still listed in spec.ahead: a mispredicted or dropped speculative wavefront keeps running on its lane after it
    // has been forgotten, and it reads the scene / probe / pool that the caller is about to change.
    for (int k = 1; k < B200PT_MAX_LANES; ++k) if (c->lanes[k].stream) CUDA_OK(cudaStreamSynchronize(c->lanes[k].stream));
    c->spec.drop();
}

int numSMs() {
    static int cache[64];                                     // 0 = not queried yet
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (!cache[dev]) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
        cache[dev] = n;
    }
    return cache[dev];
}

extern "C" {

const char* b200pt_last_error(void) { return g_lastError.c_str(); }

int b200pt_config_preset(b200pt_config* c, const char* name) {
    if (!c || !name) return fail("b200pt_config_preset: null argument");
    memset(c, 0, sizeof(*c));
    c->tmax = 1e16f;  c->exposure = 1.0f;
    c->shadow_tmin = 0.01f;       // every variant traces occlusion rays from 0.01 (orig/deviceProgram.cu:268,310; sv3 :40; sv4), also orig whose radiance rays start at 0.001 (:420)
    const std::string n(name);
    if (n == "orig") {            // orig/deviceProgram.cu:198,429,467; orig/SimplePathtracer.cpp:73-97
        c->schedule = B200PT_SCHEDULE_UNIFORM;  c->max_depth = 8;  c->tmin = 0.001f;  c->shadow_mode = B200PT_SHADOW_ANY_HIT;
        c->display = B200PT_DISPLAY_SRGB;  c->uniform_spl = 0;  c->accumulate = 1;  c->write_aovs = 1;  c->increment_subframe = 0;
    } else if (n == "sv3_uniform") {   // sv3/deviceProgram.cu:243,511,582; sv3/SimplePathtracer.cpp:81-128
        c->schedule = B200PT_SCHEDULE_UNIFORM;  c->max_depth = 4;  c->tmin = 0.01f;  c->shadow_mode = B200PT_SHADOW_CULL_BACK;
        c->display = B200PT_DISPLAY_REINHARD;  c->exposure = 8.0f;  c->uniform_spl = 32;  c->accumulate = 1;  c->increment_subframe = 1;
    } else if (n == "sv4_uniform") {   // sv4/SimplePathtracer.cpp:85-131
        c->schedule = B200PT_SCHEDULE_UNIFORM;  c->max_depth = 4;  c->tmin = 0.01f;  c->shadow_mode = B200PT_SHADOW_CULL_BACK;
        c->display = B200PT_DISPLAY_REINHARD;  c->exposure = 4.0f;  c->uniform_spl = 8;  c->accumulate = 0;  c->increment_subframe = 1;
    } else if (n == "sv4_fov") {       // sv4/SimplePathtracer.cpp:20-21,132-216
        c->schedule = B200PT_SCHEDULE_FOVEATED;  c->max_depth = 4;  c->tmin = 0.01f;  c->shadow_mode = B200PT_SHADOW_CULL_BACK;
        c->display = B200PT_DISPLAY_REINHARD;  c->exposure = 4.0f;  c->inner_radius = 157;  c->outer_radius = 515;  c->increment_subframe = 1;
    } else return fail("b200pt_config_preset: unknown preset '" + n + "'");
    return 0;
}

b200pt_ctx* b200pt_create(int device) {
    try {
        int count = 0;
        cudaError_t e = cudaGetDeviceCount(&count);
        if (e != cudaSuccess || count == 0) { fail(std::string("b200pt_create: no CUDA device (") + cudaGetErrorString(e) + "); there is no CPU fallback");  return nullptr; }
        if (device < 0 || device >= count) { fail("b200pt_create: device index out of range");  return nullptr; }
        CUDA_OK(cudaSetDevice(device));
        CUDA_OK(cudaFree(0));
        b200pt_ctx* c = new b200pt_ctx();
        c->device = device;
        CUDA_OK(cudaStreamCreate(&c->stream));                 // the reference's single stream, SimplePathtracer.cpp:342
        c->lanes[0].stream = c->stream;
        CUDA_OK(cudaEventCreateWithFlags(&c->forkEv, cudaEventDisableTiming));
        return c;
    } catch (const std::exception& ex) { fail(ex.what());  return nullptr; }
}

void b200pt_destroy(b200pt_ctx* c) {
    if (!c) return;
    cudaSetDevice(c->device);
    for (int k = 0; k < B200PT_MAX_LANES; ++k) if (c->lanes[k].stream) cudaStreamSynchronize(c->lanes[k].stream);
#ifdef B200PT_CHECKED
    {   // the checked build's verdict for everything this process ran on the device so far
        unsigned nt = 0, ft = 0, ns = 0, fs = 0;
        checkedReportTrace(&nt, &ft);  checkedReportShade(&ns, &fs);
        fprintf(stderr, "B200PT_CHECKED: trace.cu %u violations (first code %u), shade.cu %u violations (first code %u), last CUDA error: %s\n", nt, ft, ns, fs,
                cudaGetErrorString(cudaGetLastError()));
    }
#endif
    bvhFree(c->bvh);
    c->dVertices.release();  c->dNormals.release();  c->dTexcoords.release();  c->dIndices.release();  c->dTriMesh.release();  c->dMeshes.release();  c->dTexels.release();
    destroyPeer(c);
    freeAccum(c);  destroyMulti(c);
    c->frameBuf.release();  c->normalBuf.release();  c->albedoBuf.release();  c->colorBuf.release();
    c->probeData.release();  c->probePdfX.release();  c->probeCdfX.release();  c->probePdfY.release();  c->probeCdfY.release();  c->probeGuide.release();  c->probeRowTotal.release();
    for (int k = 0; k < B200PT_MAX_LANES; ++k) {
        Lane& L = c->lanes[k];
        if (L.stream) cudaStreamSynchronize(L.stream);
        L.release();
        if (L.done) cudaEventDestroy(L.done);
        if (L.headStart) cudaEventDestroy(L.headStart);
        if (k > 0 && L.stream) cudaStreamDestroy(L.stream);
    }
    if (c->forkEv) cudaEventDestroy(c->forkEv);
    for (auto ev : c->evPool) cudaEventDestroy(ev);
    cudaStreamDestroy(c->stream);
    delete c;
}

int b200pt_add_texture(b200pt_ctx* c, const uint32_t* rgba8, int w, int h) {
    if (!c || !rgba8 || w <= 0 || h <= 0) return fail("b200pt_add_texture: bad argument");
    HostTexture t{c->hTexels.size(), w, h};
    c->hTexels.insert(c->hTexels.end(), rgba8, rgba8 + (size_t)w * h);
    c->hTextures.push_back(t);
    return (int)c->hTextures.size() - 1;
}

int b200pt_add_mesh(b200pt_ctx* c, const float* vtx, size_t nv, const float* normal, const float* texcoord, const uint32_t* idx,
                    size_t ntri, const Material* material, int textureID) {
    if (!c || !material || (nv && !vtx) || (ntri && !idx)) return fail("b200pt_add_mesh: bad argument");
    if (textureID >= (int)c->hTextures.size()) return fail("b200pt_add_mesh: texture id out of range");
    for (size_t i = 0; i < 3 * ntri; ++i) if (idx[i] >= nv) return fail("b200pt_add_mesh: vertex index out of range");
    HostMesh m;
    m.firstVertex = c->hVertices.size();  m.nVertices = nv;  m.firstTri = c->hIndices.size();  m.nTris = ntri;
    m.hasTexcoord = texcoord != nullptr;  m.hasNormal = normal != nullptr;  m.textureID = textureID;  m.material = *material;
    if (normal) {                                             // never read by the path tracer's shaders; the Whitted integrator interpolates them
        const float3* n3 = reinterpret_cast<const float3*>(normal);
        c->hNormals.resize(m.firstVertex, make_float3(0.f, 0.f, 0.f));
        c->hNormals.insert(c->hNormals.end(), n3, n3 + nv);
    }
    const float3* v3 = reinterpret_cast<const float3*>(vtx);
    c->hVertices.insert(c->hVertices.end(), v3, v3 + nv);
    if (texcoord) { const float2* t2 = reinterpret_cast<const float2*>(texcoord);  c->hTexcoords.insert(c->hTexcoords.end(), t2, t2 + nv); }
    else c->hTexcoords.insert(c->hTexcoords.end(), nv, make_float2(0.f, 0.f));
    const uint3* i3 = reinterpret_cast<const uint3*>(idx);
    c->hIndices.insert(c->hIndices.end(), i3, i3 + ntri);
    c->hTriMesh.insert(c->hTriMesh.end(), ntri, (unsigned)c->hMeshes.size());
    if (material->flags & MATERIAL_FLAG_SHADOW_CATCHER) c->hasCatcher = true;
    c->hMeshes.push_back(m);
    c->built = false;
    return (int)c->hMeshes.size() - 1;
}

// Animated geometry: replace one mesh's vertex positions (same count, same indices).  Follow with b200pt_refit_accel (cheap,
// topology kept) or b200pt_build_accel (full rebuild).
int b200pt_update_vertices(b200pt_ctx* c, int meshID, const float* vtx, size_t nv) {
    if (!c || !vtx) return fail("b200pt_update_vertices: null argument");
    if (meshID < 0 || meshID >= (int)c->hMeshes.size()) return fail("b200pt_update_vertices: mesh id out of range");
    const HostMesh& m = c->hMeshes[meshID];
    if (nv != m.nVertices) return fail("b200pt_update_vertices: vertex count differs from b200pt_add_mesh");
    API_BEGIN
    memcpy(c->hVertices.data() + m.firstVertex, vtx, nv * sizeof(float3));
    if (c->built) {                                           // device copy is live: patch it in place
        enterCtxMutating(c);
        CUDA_OK(cudaMemcpyAsync(c->dVertices.p + m.firstVertex, vtx, nv * sizeof(float3), cudaMemcpyHostToDevice, c->stream));
        CUDA_OK(cudaStreamSynchronize(c->stream));
        c->stale = true;
    }
    API_END
}
int b200pt_refit_accel(b200pt_ctx* c) {
    if (!c) return fail("b200pt_refit_accel: null context");
    if (!c->built) return fail("b200pt_refit_accel: b200pt_build_accel has not been called");
    API_BEGIN
    enterCtxMutating(c);
    bvhRefit(c->dVertices.p, c->dIndices.p, c->dTriMesh.p, c->dMeshes.p, c->stream, c->bvh, &c->refitMs);
    c->stale = false;
    API_END
}

int b200pt_build_accel(b200pt_ctx* c, uint64_t* handle_out) {
    if (!c) return fail("b200pt_build_accel: null context");
    API_BEGIN
    enterCtxMutating(c);
    if (c->hIndices.size() >= (1u << 28)) throw std::runtime_error("b200pt_build_accel: more than 2^28 triangles");
    std::vector<MeshInfo> infos(c->hMeshes.size());
    for (size_t i = 0; i < infos.size(); ++i) {
        const HostMesh& m = c->hMeshes[i];
        MeshInfo& d = infos[i];
        memset(&d, 0, sizeof(d));
        d.material = m.material;  d.firstTri = (unsigned)m.firstTri;  d.firstVertex = (unsigned)m.firstVertex;
        d.textureID = m.textureID;  d.hasTexcoord = m.hasTexcoord ? 1 : 0;  d.hasNormal = m.hasNormal ? 1 : 0;
        if (m.textureID >= 0) { const HostTexture& t = c->hTextures[m.textureID];  d.texOffset = (unsigned)t.offset;  d.texW = t.w;  d.texH = t.h; }
    }
    auto up = [&](auto& buf, const auto& vec) {
        buf.ensure(std::max<size_t>(vec.size(), 1));
        if (!vec.empty()) CUDA_OK(cudaMemcpyAsync(buf.p, vec.data(), vec.size() * sizeof(vec[0]), cudaMemcpyHostToDevice, c->stream));
    };
    if (!c->hNormals.empty()) c->hNormals.resize(c->hVertices.size(), make_float3(0.f, 0.f, 0.f));
    up(c->dNormals, c->hNormals);
    up(c->dVertices, c->hVertices);  up(c->dTexcoords, c->hTexcoords);  up(c->dIndices, c->hIndices);  up(c->dTriMesh, c->hTriMesh);
    up(c->dMeshes, infos);  up(c->dTexels, c->hTexels);
    bvhFree(c->bvh);
    bvhBuild(c->dVertices.p, c->dIndices.p, c->dTriMesh.p, c->dMeshes.p, (unsigned)c->hIndices.size(), c->stream, c->bvh);
    CUDA_OK(cudaStreamSynchronize(c->stream));
    CUDA_OK(cudaGetLastError());
    c->built = true;  c->stale = false;  c->refitMs = 0.f;
    if (handle_out) *handle_out = (uint64_t)(uintptr_t)c->bvh.nodes;
    API_END
}

}  // extern "C"

static SceneDev sceneOf(b200pt_ctx* c) {
    SceneDev S;
    S.nodes = (const uint4*)c->bvh.nodes;  S.rootLink = c->bvh.rootLink;
    S.nodes4 = (const uint4*)c->bvh.nodes4;  S.rootLink4 = c->bvh.rootLink4;
    for (int a = 0; a < 3; ++a) { S.gridLo[a] = c->bvh.gridLo[a];  S.gridScale[a] = c->bvh.gridScale[a]; }
    S.tris = c->bvh.tris;  S.meshes = c->dMeshes.p;  S.indices = c->dIndices.p;
    S.vertices = c->dVertices.p;  S.normals = c->dNormals.p;
    S.texcoords = c->dTexcoords.p;  S.texels = c->dTexels.p;  S.nTris = c->bvh.nTris;  S.nNodes = c->bvh.nNodes;
    S.nNodes4 = c->bvh.nNodes4;  S.nMeshes = (unsigned)c->hMeshes.size();  S.nTexels = (unsigned)c->hTexels.size();  S.nVertices = (unsigned)c->hVertices.size();
    return S;
}

#include "api_render.inl"
#include "api_hooks.inl"
#include "api_multi.inl"
#include "api_peer.inl"
