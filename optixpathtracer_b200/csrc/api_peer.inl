// One-sided tile split: the frame partition's exchange step done by the producing kernel instead of a collective.
//   root         b200pt_peer_export  -> a 192-byte handle (CUDA IPC handles of its RGBA8 frame buffer and of a small flag block, plus the raw pointers
//                                       for ranks that live in the same process), handed to the other ranks by the caller (any byte transport);
//   other ranks  b200pt_peer_attach  -> map both over NVLink (cudaIpcOpenMemHandle, or peer access + the raw pointer inside one process).
// From then on a frame rendered with cfg.tile_count > 1 writes this rank's 32-row bands from finalizeKernel STRAIGHT into root's frame buffer - the
// transfer is the kernel's own stores, band by band as they are computed - and a one-thread kernel behind it publishes the frame's epoch in root's flag
// block (st.release.sys).  Root orders its primary stream behind all ranks' epochs with b200pt_peer_acquire (a one-CTA kernel spinning on
// ld.acquire.sys), consumes the frame (download, display), and hands the buffer back with b200pt_peer_release; a rank's next finalize waits for
// that acknowledgement, so a fast rank never overwrites a frame root is still reading.  No NCCL, no staging copy, no host synchronisation; every spin
// gives up after a time-out (B200PT_PEER_TIMEOUT_MS, default 5000) and records which wait failed (b200pt_peer_status).
// Frames with several finalize passes (the foveated schedule: later passes overwrite earlier ones) add a barrier between the passes through the same flags:
// every rank, root included, publishes "pass k done" and waits for all the others before its pass k + 1 may write.
// The reference has no counterpart (sv4/main.cpp:289-290); pixels are the same bytes b200pt_gather_bands delivers.
#include <unistd.h>

#define PEER_MAX_RANKS 64
#define PEER_ACK PEER_MAX_RANKS
#define PEER_FLAG_WORDS (PEER_MAX_RANKS + 1)
#define PEER_MAGIC 0x62327074u

struct PeerHandle {                       // the 192 bytes b200pt_peer_export writes
    unsigned magic;  int device;  long long pid;  int W, H, ipcOk, reserved;
    unsigned long long framePtr, flagsPtr;
    cudaIpcMemHandle_t frameIpc, flagsIpc;
};
static_assert(sizeof(PeerHandle) <= B200PT_PEER_HANDLE_BYTES, "PeerHandle must fit the public handle size");

__device__ __forceinline__ unsigned ldAcquireSys(const unsigned* p) { unsigned v;  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");  return v; }
__device__ __forceinline__ void stReleaseSys(unsigned* p, unsigned v) { asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ unsigned long long globalTimerNs() { unsigned long long t;  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));  return t; }

// everything enqueued before this kernel on its stream (the finalize passes' remote stores) is complete when it runs; the fence + release store
// publish the epoch after them at system scope
__global__ void peerSignalKernel(unsigned* flag, unsigned epoch) { __threadfence_system();  stReleaseSys(flag, epoch); }
// thread t waits until flags[first + t] has reached `epoch` (wrap-safe compare); `skip` = a slot that is not waited for (root's own)
__global__ void peerWaitKernel(const unsigned* flags, int first, int count, int skip, unsigned epoch, unsigned long long timeoutNs, unsigned* err, unsigned code) {
    const int r = first + (int)threadIdx.x;
    if ((int)threadIdx.x >= count || r == skip) return;
    const unsigned long long t0 = globalTimerNs();
    while ((int)(ldAcquireSys(flags + r) - epoch) < 0) {
        if (globalTimerNs() - t0 > timeoutNs) { atomicCAS(err, 0u, code + (unsigned)r);  return; }
        __nanosleep(256);
    }
}

static PeerState* peerState(b200pt_ctx* c) {
    if (!c->peer) {
        c->peer = new PeerState();
        if (const char* e = getenv("B200PT_PEER_TIMEOUT_MS")) { const long ms = atol(e);  if (ms > 0) c->peer->timeoutNs = (unsigned long long)ms * 1000000ull; }
        CUDA_OK(cudaMalloc((void**)&c->peer->err, sizeof(unsigned)));
        CUDA_OK(cudaMemset(c->peer->err, 0, sizeof(unsigned)));
        // load the two kernels now: CUDA loads kernels lazily and a load may synchronise the context, which must not happen later, when a stream of
        // this context (or of another rank sharing the GPU) is already spinning in peerWaitKernel
        cudaFuncAttributes fa;
        CUDA_OK(cudaFuncGetAttributes(&fa, peerSignalKernel));
        CUDA_OK(cudaFuncGetAttributes(&fa, peerWaitKernel));
    }
    return c->peer;
}
static void peerUnmap(b200pt_ctx* c) {
    PeerState* p = c->peer;
    if (!p || !p->remoteFrame) return;
    if (p->viaIpc) { cudaIpcCloseMemHandle(p->remoteFrame);  cudaIpcCloseMemHandle(p->remoteFlags); }
    p->remoteFrame = nullptr;  p->remoteFlags = nullptr;  p->viaIpc = false;  p->rank = -1;  p->nranks = 0;  p->put = 0;
}
static void destroyPeer(b200pt_ctx* c) {
    PeerState* p = c->peer;
    if (!p) return;
    for (int k = 0; k < B200PT_MAX_LANES; ++k) if (c->lanes[k].stream) cudaStreamSynchronize(c->lanes[k].stream);   // no put or wait may still be running
    peerUnmap(c);
    if (p->flags) cudaFree(p->flags);
    if (p->err) cudaFree(p->err);
    delete p;
    c->peer = nullptr;
}

// enqueueFinalize's hooks (api_render.inl) ------------------------------------------------------------------------------------------------
// Epochs: a rank that has put f frames publishes 8 * f; inside frame f + 1 it publishes 8 * f + k after its k-th finalize pass when the frame has several
// (the foveated schedule: a later pass overwrites pixels of an earlier one, so no rank may start pass k + 1 before every rank has finished pass k).
// role: 0 = not part of a one-sided partition for this frame, 1 = attached rank (writes root's frame), 2 = the exporting root (writes its own frame, takes
// part in the pass barriers).
static int peerRole(b200pt_ctx* c, const b200pt_config& cfg, const FrameParams& F) {
    PeerState* p = c->peer;
    if (!p || cfg.tile_count <= 1 || cfg.accum_mode == B200PT_ACCUM_SUM || F.W != p->W || F.H != p->H) return 0;
    if (p->remoteFrame) return 1;
    return (p->exported && cfg.tile_count == p->nranks && cfg.tile_rank == p->rank) ? 2 : 0;
}
static int peerPutBegin(b200pt_ctx* c, const b200pt_config& cfg, FrameParams& F, cudaStream_t s) {
    const int role = peerRole(c, cfg, F);
    PeerState* p = c->peer;
    if (role == 1) {
        // root must have consumed the frame before this one (epoch 8 * put) - the first frame passes at once
        peerWaitKernel<<<1, 32, 0, s>>>(p->remoteFlags, PEER_ACK, 1, -1, 8u * p->put, p->timeoutNs, p->err, 1000u);
        F.frame = p->remoteFrame;
    }
    return role;
}
// between two finalize passes of a frame: publish "pass k done", wait until every rank has
static void peerPassBarrier(b200pt_ctx* c, int role, int k, cudaStream_t s) {
    PeerState* p = c->peer;
    unsigned* flags = role == 1 ? p->remoteFlags : p->flags;
    const unsigned epoch = 8u * p->put + (unsigned)k;
    peerSignalKernel<<<1, 1, 0, s>>>(flags + p->rank, epoch);
    peerWaitKernel<<<1, PEER_MAX_RANKS, 0, s>>>(flags, 0, p->nranks, p->rank, epoch, p->timeoutNs, p->err, 2000u);
}
static void peerPutEnd(b200pt_ctx* c, int role, cudaStream_t s) {
    PeerState* p = c->peer;
    ++p->put;
    peerSignalKernel<<<1, 1, 0, s>>>((role == 1 ? p->remoteFlags : p->flags) + p->rank, 8u * p->put);
}

extern "C" {

int b200pt_peer_export(b200pt_ctx* c, const LaunchParams* params, int rank, int nranks, void* handle) {
    if (!c || !params || !handle) return fail("b200pt_peer_export: null argument");
    if (nranks < 2 || nranks > PEER_MAX_RANKS || rank < 0 || rank >= nranks) return fail("b200pt_peer_export: rank out of range (2 <= n_ranks <= 64)");
    API_BEGIN
    enterCtxMutating(c);
    CUDA_OK(cudaStreamSynchronize(c->stream));
    if (!c->frameBuf.p || c->W <= 0 || c->H <= 0) throw std::runtime_error("b200pt_peer_export: frame buffers not allocated (call b200pt_resize)");
    if ((void*)params->frame.frame_buffer != (void*)c->frameBuf.p) throw std::runtime_error("b200pt_peer_export: params->frame.frame_buffer is not this context's frame buffer");
    PeerState* p = peerState(c);
    if (!p->flags) CUDA_OK(cudaMalloc((void**)&p->flags, PEER_FLAG_WORDS * sizeof(unsigned)));
    CUDA_OK(cudaMemset(p->flags, 0, PEER_FLAG_WORDS * sizeof(unsigned)));      // a new export starts a new epoch count: every rank attaches again
    CUDA_OK(cudaMemset(p->err, 0, sizeof(unsigned)));
    peerUnmap(c);                                                              // (a context is either the root of a partition or attached to one)
    p->exported = true;  p->acquired = 0;  p->released = 0;  p->put = 0;  p->W = c->W;  p->H = c->H;  p->rank = rank;  p->nranks = nranks;
    PeerHandle h;
    memset(&h, 0, sizeof(h));
    h.magic = PEER_MAGIC;  h.device = c->device;  h.pid = (long long)getpid();  h.W = c->W;  h.H = c->H;
    h.framePtr = (unsigned long long)(uintptr_t)c->frameBuf.p;  h.flagsPtr = (unsigned long long)(uintptr_t)p->flags;
    h.ipcOk = (cudaIpcGetMemHandle(&h.frameIpc, c->frameBuf.p) == cudaSuccess && cudaIpcGetMemHandle(&h.flagsIpc, p->flags) == cudaSuccess) ? 1 : 0;
    if (!h.ipcOk) cudaGetLastError();                                          // (ranks inside this process can still attach through the raw pointers)
    memset(handle, 0, B200PT_PEER_HANDLE_BYTES);
    memcpy(handle, &h, sizeof(h));
    API_END
}

int b200pt_peer_attach(b200pt_ctx* c, const void* handle, int rank, int nranks) {
    if (!c || !handle) return fail("b200pt_peer_attach: null argument");
    if (nranks < 2 || nranks > PEER_MAX_RANKS || rank < 0 || rank >= nranks) return fail("b200pt_peer_attach: rank out of range (2 <= n_ranks <= 64)");
    API_BEGIN
    enterCtxMutating(c);
    CUDA_OK(cudaStreamSynchronize(c->stream));
    PeerHandle h;
    memcpy(&h, handle, sizeof(h));
    if (h.magic != PEER_MAGIC) throw std::runtime_error("b200pt_peer_attach: not a handle written by b200pt_peer_export");
    if (h.W != c->W || h.H != c->H) throw std::runtime_error("b200pt_peer_attach: the exported frame has a different size (resize both contexts first)");
    PeerState* p = peerState(c);
    if (p->exported && h.pid == (long long)getpid() && h.flagsPtr == (unsigned long long)(uintptr_t)p->flags)
        throw std::runtime_error("b200pt_peer_attach: this context exported that frame (root does not attach)");
    peerUnmap(c);
    p->exported = false;
    if (h.pid == (long long)getpid()) {                                        // one thread per GPU inside one process: peer access + the raw pointers
        if (h.device != c->device) {
            int can = 0;
            CUDA_OK(cudaDeviceCanAccessPeer(&can, c->device, h.device));
            if (!can) throw std::runtime_error("b200pt_peer_attach: no peer access between the two devices");
            const cudaError_t e = cudaDeviceEnablePeerAccess(h.device, 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) CUDA_OK(e);
            cudaGetLastError();
        }
        p->remoteFrame = (uint32_t*)(uintptr_t)h.framePtr;  p->remoteFlags = (unsigned*)(uintptr_t)h.flagsPtr;  p->viaIpc = false;
    } else {
        if (!h.ipcOk) throw std::runtime_error("b200pt_peer_attach: the exporting process could not create CUDA IPC handles");
        void* f = nullptr;  void* g = nullptr;
        CUDA_OK(cudaIpcOpenMemHandle(&f, h.frameIpc, cudaIpcMemLazyEnablePeerAccess));
        const cudaError_t e = cudaIpcOpenMemHandle(&g, h.flagsIpc, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) { cudaIpcCloseMemHandle(f);  CUDA_OK(e); }
        p->remoteFrame = (uint32_t*)f;  p->remoteFlags = (unsigned*)g;  p->viaIpc = true;
    }
    p->rank = rank;  p->nranks = nranks;  p->W = h.W;  p->H = h.H;  p->put = 0;
    CUDA_OK(cudaMemset(p->err, 0, sizeof(unsigned)));
    API_END
}

int b200pt_peer_detach(b200pt_ctx* c) {
    if (!c) return fail("b200pt_peer_detach: null context");
    API_BEGIN
    enterCtxMutating(c);
    CUDA_OK(cudaStreamSynchronize(c->stream));
    destroyPeer(c);
    API_END
}

int b200pt_peer_acquire(b200pt_ctx* c) {
    if (!c) return fail("b200pt_peer_acquire: null context");
    if (!c->peer || !c->peer->exported) return fail("b200pt_peer_acquire: nothing exported (call b200pt_peer_export)");
    API_BEGIN
    enterCtx(c);                                                               // stream-level join: root's own bands of this frame are behind the primary stream too
    PeerState* p = c->peer;
    ++p->acquired;
    peerWaitKernel<<<1, PEER_MAX_RANKS, 0, c->stream>>>(p->flags, 0, p->nranks, p->rank, 8u * p->acquired, p->timeoutNs, p->err, 1u);
    CUDA_OK(cudaGetLastError());
    API_END
}

int b200pt_peer_release(b200pt_ctx* c) {
    if (!c) return fail("b200pt_peer_release: null context");
    if (!c->peer || !c->peer->exported) return fail("b200pt_peer_release: nothing exported (call b200pt_peer_export)");
    API_BEGIN
    CUDA_OK(cudaSetDevice(c->device));
    PeerState* p = c->peer;
    if (p->released == p->acquired) throw std::runtime_error("b200pt_peer_release: no acquired frame to release");
    ++p->released;
    peerSignalKernel<<<1, 1, 0, c->stream>>>(p->flags + PEER_ACK, 8u * p->released);   // behind whatever the caller enqueued on the primary stream since the acquire
    CUDA_OK(cudaGetLastError());
    API_END
}

int b200pt_peer_status(b200pt_ctx* c, int* state, int* attachedRank, unsigned* epoch) {
    if (!c) return fail("b200pt_peer_status: null context");
    API_BEGIN
    CUDA_OK(cudaSetDevice(c->device));
    unsigned e = 0;
    if (c->peer) {
        for (int k = 0; k < B200PT_MAX_LANES; ++k) if (c->lanes[k].stream) CUDA_OK(cudaStreamSynchronize(c->lanes[k].stream));
        CUDA_OK(cudaMemcpy(&e, c->peer->err, sizeof(e), cudaMemcpyDeviceToHost));
    }
    if (state) *state = (int)e;
    if (attachedRank) *attachedRank = (c->peer && c->peer->remoteFrame) ? c->peer->rank : -1;
    if (epoch) *epoch = c->peer ? (c->peer->exported ? c->peer->acquired : c->peer->put) : 0u;
    API_END
}

}  // extern "C"
