// b200pt — multi-GPU below Python (included by api.cu): one context per rank, the scene replicated, ONE collective per frame partition
// (SURVEY §8e): a fp32 SUM reduce of the accumulation buffer for the sample split, a gather of row bands for the tile split.
// The collectives are NCCL's, issued by this library on the context's primary stream, on buffers allocated with ncclMemAlloc and registered
// with the communicator (zero-copy for NVLS / user-buffer registration).  libnccl.so.2 is opened at run time: libb200pt.so itself does not
// link it, so single-GPU hosts need no NCCL, and inside a torch process the loader hands back the NCCL torch already mapped (same SONAME).
#include <dlfcn.h>
#include <nccl.h>

struct NcclApi {
    void* lib = nullptr;
    ncclResult_t (*GetVersion)(int*) = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    ncclResult_t (*MemAlloc)(void**, size_t) = nullptr;
    ncclResult_t (*MemFree)(void*) = nullptr;
    ncclResult_t (*CommRegister)(const ncclComm_t, void*, size_t, void**) = nullptr;
    ncclResult_t (*CommDeregister)(const ncclComm_t, void*) = nullptr;
    ncclResult_t (*Reduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
};
static NcclApi loadNcclApi() {
    NcclApi a;
    const char* names[] = {"libnccl.so.2", "libnccl.so"};
    for (const char* n : names) { a.lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL);  if (a.lib) break; }
    if (a.lib) {
        bool ok = true;
        auto sym = [&](const char* name, bool required) -> void* { void* p = dlsym(a.lib, name);  if (!p && required) ok = false;  return p; };
#define B200PT_NCCL_SYM(field, name, req) a.field = reinterpret_cast<decltype(a.field)>(sym(name, req))
        B200PT_NCCL_SYM(GetVersion, "ncclGetVersion", true);  B200PT_NCCL_SYM(GetUniqueId, "ncclGetUniqueId", true);
        B200PT_NCCL_SYM(CommInitRank, "ncclCommInitRank", true);  B200PT_NCCL_SYM(CommDestroy, "ncclCommDestroy", true);
        B200PT_NCCL_SYM(GetErrorString, "ncclGetErrorString", true);  B200PT_NCCL_SYM(Reduce, "ncclReduce", true);
        B200PT_NCCL_SYM(Send, "ncclSend", true);  B200PT_NCCL_SYM(Recv, "ncclRecv", true);
        B200PT_NCCL_SYM(GroupStart, "ncclGroupStart", true);  B200PT_NCCL_SYM(GroupEnd, "ncclGroupEnd", true);
        B200PT_NCCL_SYM(MemAlloc, "ncclMemAlloc", false);  B200PT_NCCL_SYM(MemFree, "ncclMemFree", false);           // NCCL >= 2.19
        B200PT_NCCL_SYM(CommRegister, "ncclCommRegister", false);  B200PT_NCCL_SYM(CommDeregister, "ncclCommDeregister", false);
#undef B200PT_NCCL_SYM
        if (!ok) { dlclose(a.lib);  a.lib = nullptr; }
    }
    return a;
}
static NcclApi& ncclApi() {
    static NcclApi a = loadNcclApi();                          // initialised once, thread-safe (one host thread per GPU may call in concurrently)
    if (!a.lib) throw std::runtime_error("libnccl.so.2 could not be loaded (multi-GPU entry points need NCCL on the library path)");
    return a;
}
#define NCCL_OK(call) do { ncclResult_t r_ = (call); if (r_ != ncclSuccess) \
    throw std::runtime_error(std::string(#call) + ": " + ncclApi().GetErrorString(r_)); } while (0)

struct MultiState {
    ncclComm_t comm = nullptr;
    int nranks = 1, rank = 0;
    void* regAccum = nullptr;  void* regFrame = nullptr;      // ncclCommRegister handles
    float4* total = nullptr;  size_t totalN = 0;  bool totalFromNccl = false;   // root's receive buffer of b200pt_reduce_sum(keep_local = 1)
};

// The accumulation buffer: cudaMalloc without a communicator, ncclMemAlloc + registration with one (the buffer the last
// accumulation kernel writes IS the collective's send buffer).
static void freeAccum(b200pt_ctx* c) {
    MultiState* m = c->multi;
    if (m && m->comm && m->regAccum && ncclApi().CommDeregister) { ncclApi().CommDeregister(m->comm, m->regAccum);  m->regAccum = nullptr; }
    if (m && m->comm && m->regFrame && ncclApi().CommDeregister) { ncclApi().CommDeregister(m->comm, m->regFrame);  m->regFrame = nullptr; }
    if (c->accumFromNccl && c->accumBuf.p) { ncclApi().MemFree(c->accumBuf.p);  c->accumBuf.p = nullptr;  c->accumBuf.n = 0; }
    else c->accumBuf.release();
    c->accumFromNccl = false;
}
static void allocAccum(b200pt_ctx* c, size_t n) {
    freeAccum(c);
    MultiState* m = c->multi;
    if (m && m->comm && ncclApi().MemAlloc && ncclApi().MemFree) {
        void* p = nullptr;
        if (ncclApi().MemAlloc(&p, n * sizeof(float4)) == ncclSuccess && p) { c->accumBuf.p = (float4*)p;  c->accumBuf.n = n;  c->accumFromNccl = true; }
    }
    if (!c->accumBuf.p) c->accumBuf.ensure(n);
    if (m && m->comm && ncclApi().CommRegister) {                              // best effort: registration only speeds the collective up
        if (ncclApi().CommRegister(m->comm, c->accumBuf.p, n * sizeof(float4), &m->regAccum) != ncclSuccess) m->regAccum = nullptr;
        if (c->frameBuf.p && ncclApi().CommRegister(m->comm, c->frameBuf.p, n * sizeof(uint32_t), &m->regFrame) != ncclSuccess) m->regFrame = nullptr;
    }
}
static void destroyMulti(b200pt_ctx* c) {
    MultiState* m = c->multi;
    if (!m) return;
    if (m->comm) {
        if (m->regAccum && ncclApi().CommDeregister) ncclApi().CommDeregister(m->comm, m->regAccum);
        if (m->regFrame && ncclApi().CommDeregister) ncclApi().CommDeregister(m->comm, m->regFrame);
        if (m->total) { if (m->totalFromNccl) ncclApi().MemFree(m->total);  else cudaFree(m->total); }
        ncclApi().CommDestroy(m->comm);
    }
    delete m;
    c->multi = nullptr;
}

// drops the communicator; an accumulation buffer that lives in NCCL-allocated memory goes back to a plain allocation, contents kept
static void leaveComm(b200pt_ctx* c) {
    if (!c->multi) return;
    if (c->accumFromNccl && c->accumBuf.p) {
        const size_t n = c->accumBuf.n;
        std::vector<float4> keep(n);
        CUDA_OK(cudaMemcpy(keep.data(), c->accumBuf.p, n * sizeof(float4), cudaMemcpyDeviceToHost));
        freeAccum(c);                                                          // deregisters + ncclMemFree while the communicator still exists
        destroyMulti(c);
        c->accumBuf.ensure(n);
        CUDA_OK(cudaMemcpy(c->accumBuf.p, keep.data(), n * sizeof(float4), cudaMemcpyHostToDevice));
    } else destroyMulti(c);                                                    // (deregisters what it registered)
}

extern "C" {

int b200pt_comm_unique_id(void* id128) {
    if (!id128) return fail("b200pt_comm_unique_id: null argument");
    API_BEGIN
    static_assert(sizeof(ncclUniqueId) == B200PT_UNIQUE_ID_BYTES, "ncclUniqueId size");
    ncclUniqueId id;
    NCCL_OK(ncclApi().GetUniqueId(&id));
    memcpy(id128, &id, sizeof(id));
    API_END
}

int b200pt_comm_init(b200pt_ctx* c, int nranks, int rank, const void* id128, LaunchParams* params) {
    if (!c || !id128) return fail("b200pt_comm_init: null argument");
    if (nranks < 1 || rank < 0 || rank >= nranks) return fail("b200pt_comm_init: rank out of range");
    API_BEGIN
    enterCtxMutating(c);
    CUDA_OK(cudaStreamSynchronize(c->stream));
    const bool callerPointsAtOurs = params && c->accumBuf.p && params->frame.accum_buffer == c->accumBuf.p;
    leaveComm(c);                                                              // re-initialisation: drop the old communicator first (the buffer's contents survive)
    ncclUniqueId id;
    memcpy(&id, id128, sizeof(id));
    MultiState* m = new MultiState();
    m->nranks = nranks;  m->rank = rank;
    c->multi = m;
    try { NCCL_OK(ncclApi().CommInitRank(&m->comm, nranks, id, rank)); }
    catch (...) { delete m;  c->multi = nullptr;  throw; }
    if (c->W > 0 && c->H > 0) {                                                // frame buffers exist already: move the accumulation buffer into NCCL-allocated memory
        const size_t n = (size_t)c->W * c->H;
        std::vector<float4> keep(n);
        CUDA_OK(cudaMemcpy(keep.data(), c->accumBuf.p, n * sizeof(float4), cudaMemcpyDeviceToHost));
        allocAccum(c, n);
        CUDA_OK(cudaMemcpy(c->accumBuf.p, keep.data(), n * sizeof(float4), cudaMemcpyHostToDevice));
        if (callerPointsAtOurs) params->frame.accum_buffer = c->accumBuf.p;
    }
    API_END
}

int b200pt_comm_destroy(b200pt_ctx* c, LaunchParams* params) {
    if (!c) return fail("b200pt_comm_destroy: null context");
    API_BEGIN
    enterCtxMutating(c);
    CUDA_OK(cudaStreamSynchronize(c->stream));
    const bool callerPointsAtOurs = params && c->accumBuf.p && params->frame.accum_buffer == c->accumBuf.p;
    leaveComm(c);
    if (callerPointsAtOurs) params->frame.accum_buffer = c->accumBuf.p;       // the buffer moved back to a plain allocation
    API_END
}

// Sample split epilogue: SUM-reduce the fp32 accumulation buffer onto `root` and, there, divide by n_passes and run the display transform -
// all enqueued on the primary stream behind every frame in flight; no host synchronisation.
//   keep_local = 0: in place; on root the accumulation buffer ends up holding the MEAN (alpha 1), like b200pt_resolve_sum (end of a job);
//   keep_local = 1: progressive display - every rank's buffer keeps its own partial SUM, root receives the total into a buffer of its own and
//                   only the frame buffer is refreshed, so the ranks can go on accumulating and reduce again after the next frames.
int b200pt_reduce_sum(b200pt_ctx* c, const LaunchParams* params, const b200pt_config* cfg, int root, int nPasses, int keepLocal) {
    if (!c || !params || !cfg || nPasses <= 0) return fail("b200pt_reduce_sum: bad argument");
    if (!c->multi || !c->multi->comm) return fail("b200pt_reduce_sum: no communicator (call b200pt_comm_init)");
    if (root < 0 || root >= c->multi->nranks) return fail("b200pt_reduce_sum: root out of range");
    API_BEGIN
    enterCtx(c);                                                               // stream-level join only: the primary stream waits for every frame in flight, the host does not
    float4* buf = params->frame.accum_buffer;                                  // (speculative wavefronts never touch the accumulation buffer: only finalize passes do)
    const size_t n = (size_t)params->frame.size.x * params->frame.size.y;
    if (!buf || !n) throw std::runtime_error("b200pt_reduce_sum: frame buffers not allocated (call b200pt_resize)");
    float4* dst = buf;
    if (keepLocal && c->multi->rank == root) {
        if (c->multi->totalN < n) {
            if (c->multi->total) { if (c->multi->totalFromNccl) ncclApi().MemFree(c->multi->total);  else cudaFree(c->multi->total); }
            c->multi->total = nullptr;  c->multi->totalFromNccl = false;
            void* p = nullptr;
            if (ncclApi().MemAlloc && ncclApi().MemFree && ncclApi().MemAlloc(&p, n * sizeof(float4)) == ncclSuccess && p) c->multi->totalFromNccl = true;
            else CUDA_OK(cudaMalloc(&p, n * sizeof(float4)));
            c->multi->total = (float4*)p;  c->multi->totalN = n;
        }
        dst = c->multi->total;
    }
    NCCL_OK(ncclApi().Reduce(buf, dst, 4 * n, ncclFloat, ncclSum, root, c->multi->comm, c->stream));
    if (c->multi->rank == root) {
        FrameParams F = frameOf(*params, *cfg);
        if (keepLocal) launchResolveTotal(F, dst, nPasses, c->stream);
        else launchResolveSum(F, nPasses, c->stream);
    }
    CUDA_OK(cudaGetLastError());
    API_END
}

// Tile split epilogue (uniform schedules: launch rows are pixel rows): every rank sends the 32-row bands it rendered
// (band % nranks == rank) to `root`, which receives them straight into its own frame / accumulation buffer.  what: bit 0 = RGBA8 frame
// buffer, bit 1 = float4 accumulation buffer.  One grouped ncclSend/ncclRecv batch on the primary stream; no host synchronisation.
int b200pt_gather_bands(b200pt_ctx* c, const LaunchParams* params, int root, int what) {
    if (!c || !params || !(what & 3)) return fail("b200pt_gather_bands: bad argument");
    if (!c->multi || !c->multi->comm) return fail("b200pt_gather_bands: no communicator (call b200pt_comm_init)");
    if (root < 0 || root >= c->multi->nranks) return fail("b200pt_gather_bands: root out of range");
    API_BEGIN
    enterCtx(c);                                                               // stream-level join only (see b200pt_reduce_sum)
    const int W = params->frame.size.x, H = params->frame.size.y, n = c->multi->nranks, me = c->multi->rank;
    uint32_t* frame = (uint32_t*)params->frame.frame_buffer;
    float4* accum = params->frame.accum_buffer;
    if (!frame || !accum || W <= 0 || H <= 0) throw std::runtime_error("b200pt_gather_bands: frame buffers not allocated (call b200pt_resize)");
    NCCL_OK(ncclApi().GroupStart());
    for (int b = 0; b * 32 < H; ++b) {
        const int owner = b % n;
        if (owner == root || (me != owner && me != root)) continue;
        const size_t first = (size_t)b * 32 * W, count = (size_t)(std::min(H, b * 32 + 32) - b * 32) * W;
        if (what & 1) {
            if (me == owner) NCCL_OK(ncclApi().Send(frame + first, 4 * count, ncclUint8, root, c->multi->comm, c->stream));
            else NCCL_OK(ncclApi().Recv(frame + first, 4 * count, ncclUint8, owner, c->multi->comm, c->stream));
        }
        if (what & 2) {
            if (me == owner) NCCL_OK(ncclApi().Send(accum + first, 4 * count, ncclFloat, root, c->multi->comm, c->stream));
            else NCCL_OK(ncclApi().Recv(accum + first, 4 * count, ncclFloat, owner, c->multi->comm, c->stream));
        }
    }
    NCCL_OK(ncclApi().GroupEnd());
    API_END
}

int b200pt_comm_info(b200pt_ctx* c, int* nranks, int* rank, int* ncclVersion, int* accumRegistered) {
    if (!c) return fail("b200pt_comm_info: null context");
    API_BEGIN
    if (nranks) *nranks = c->multi ? c->multi->nranks : 1;
    if (rank) *rank = c->multi ? c->multi->rank : 0;
    if (ncclVersion) { *ncclVersion = 0;  if (c->multi) ncclApi().GetVersion(ncclVersion); }
    if (accumRegistered) *accumRegistered = (c->accumFromNccl ? 1 : 0) | ((c->multi && c->multi->regAccum) ? 2 : 0);
    API_END
}

}  // extern "C"
