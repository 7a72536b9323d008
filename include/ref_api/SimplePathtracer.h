// class SampleRenderer — the reference's host renderer class (HelloPathtracing_sv4_vmv23/SimplePathtracer.h:45-188,
// SimplePathtracer.cpp:42-311) re-implemented as a header-only forwarder over the C-ABI of include/b200pt.h.
// Same member names, argument meaning and error behaviour: every failure throws std::runtime_error (the reference throws
// sutil::Exception, which derives from it); `launchParams` and `stream` are public and live, and render() re-reads
// launchParams (camera, frame.c, subframe_index, ...) on every call, exactly as main.cpp expects (main.cpp:310-320,392).
// No OptiX, no GL: render(CUDAOutputBuffer&) becomes render(uint32_t* deviceFrameBuffer).
#pragma once
#include <cstdint>
#include <cmath>
#include <stdexcept>
#include <string>
#include "../b200pt.h"
#include "Camera.h"
#include "LaunchParams.h"
#include "Model.h"
#include "Probe.h"

class SampleRenderer {
public:
    // ctor: uploads every mesh/texture and builds the acceleration structure (SimplePathtracer.cpp:42-75).
    // `model` is borrowed and must outlive the renderer (SimplePathtracer.h:173).  `preset` selects the variant's literals:
    // "sv4_fov" (the sv4 default, FOV_ON), "sv4_uniform", "sv3_uniform", "orig".
    explicit SampleRenderer(const Model* model, int device = 0, const char* preset = "sv4_fov") : model(model) {
        launchParams = LaunchParams{};
        if (b200pt_config_preset(&config, preset) != 0) fail();
        ctx = b200pt_create(device);
        if (!ctx) fail();
        try {
            for (const Texture* t : model->textures)
                if (b200pt_add_texture(ctx, t->pixel, t->resolution.x, t->resolution.y) < 0) fail();
            for (const TriangleMesh* m : model->meshes) {
                const bool hasN = m->normal.size() == m->vertex.size() && !m->normal.empty();
                const bool hasT = m->texcoord.size() == m->vertex.size() && !m->texcoord.empty();
                if (b200pt_add_mesh(ctx, reinterpret_cast<const float*>(m->vertex.data()), m->vertex.size(),
                                    hasN ? reinterpret_cast<const float*>(m->normal.data()) : nullptr,
                                    hasT ? reinterpret_cast<const float*>(m->texcoord.data()) : nullptr,
                                    reinterpret_cast<const uint32_t*>(m->index.data()), m->index.size(), &m->material,
                                    m->diffuseTextureID) < 0) fail();
            }
            uint64_t handle = 0;
            if (b200pt_build_accel(ctx, &handle) != 0) fail();
            launchParams.traversable = handle;
            void* s = nullptr;
            b200pt_device_buffers(ctx, nullptr, nullptr, &s);
            stream = s;
        } catch (...) { b200pt_destroy(ctx);  ctx = nullptr;  throw; }
    }
    ~SampleRenderer() { if (ctx) b200pt_destroy(ctx); }
    SampleRenderer(const SampleRenderer&) = delete;
    SampleRenderer& operator=(const SampleRenderer&) = delete;

    // render one frame (SimplePathtracer.cpp:77-217): no-op before the first resize; uniform or 3-pass foveated schedule per
    // `config`; increments launchParams.frame.subframe_index as the variant does; blocks until the GPU is idle.
    void render() {
        if (launchParams.frame.size.x == 0) return;
        if (b200pt_render(ctx, &launchParams, &config) != 0) fail();
    }
    // headless stand-in for render(sutil::CUDAOutputBuffer<uint32_t>&) (:219-229): render into a caller-owned DEVICE buffer
    // of W*H RGBA8.  As in the reference, frame_buffer keeps pointing at it afterwards.
    void render(uint32_t* deviceFrameBuffer) {
        launchParams.frame.frame_buffer = reinterpret_cast<uchar4*>(deviceFrameBuffer);
        render();
    }
    // Extensions with no reference counterpart (its render() always ends in CUDA_SYNC_CHECK): submit a frame without waiting, with
    // config.frames_in_flight (1..4) frames overlapping on the device - they still fold into the accumulation buffer in submission order,
    // so the pixels are those of consecutive render() calls; downloadPixelsAsync queues the copy behind the frame just submitted (pinned
    // h_pixels, valid after synchronize()).
    void renderAsync() {
        if (launchParams.frame.size.x == 0) return;
        if (b200pt_render_async(ctx, &launchParams, &config) != 0) fail();
    }
    void downloadPixelsAsync(uint32_t h_pixels[]) { if (b200pt_download_pixels_async(ctx, h_pixels) != 0) fail(); }
    void synchronize() { if (b200pt_synchronize(ctx) != 0) fail(); }
    // resize frame buffer to given resolution (:231-277); zero sizes are ignored; does not reset subframe_index.
    void resize(const int2& newSize) {
        if (newSize.x == 0 || newSize.y == 0) return;
        if (b200pt_resize(ctx, newSize.x, newSize.y, &launchParams) != 0) fail();
    }
    // download the rendered color buffer (:279-283): W*H RGBA8, R in the low byte, from the internal frame buffer.
    void downloadPixels(uint32_t h_pixels[]) { if (b200pt_download_pixels(ctx, h_pixels) != 0) fail(); }
    // set camera to render with (:285-292): aspect is overridden by W/H.
    void setCamera(const sutil::Camera& camera) {
        lastSetCamera = camera;
        lastSetCamera.setAspectRatio(float(launchParams.frame.size.x) / float(launchParams.frame.size.y));
        lastSetCamera.UVWFrame(launchParams.camera.U, launchParams.camera.V, launchParams.camera.W);
        launchParams.camera.eye = lastSetCamera.eye();
    }
    // (:295-311) requires probe.valid, else throws; the host arrays stay caller-owned.
    void setProbe(const ProbeData& probe) {
        if (!probe.valid) throw std::runtime_error("Probe Data is not valid");
        const float off[3] = {probe.offset.x, probe.offset.y, probe.offset.z};
        if (b200pt_set_probe(ctx, probe.width, probe.height, reinterpret_cast<const float*>(probe.data), probe.pdfValuesX,
                             probe.cdfValuesX, probe.pdfValuesY, probe.cdfValuesY, off, &launchParams) != 0) fail();
    }
    // loadProbe/loadColor without the host BuildCDF: uploads probe.data only and builds the four tables on the GPU (same bits as
    // ProbeData::BuildCDF, Probe.h:29-77); probe.valid is not required.
    void setProbeImage(const ProbeData& probe) {
        if (!probe.data || probe.width <= 0 || probe.height <= 0) throw std::runtime_error("Probe Data is not valid");
        const float off[3] = {probe.offset.x, probe.offset.y, probe.offset.z};
        if (b200pt_set_probe_image(ctx, probe.width, probe.height, reinterpret_cast<const float*>(probe.data), off, &launchParams) != 0) fail();
    }
    // Second integrator on the same traversal kernels: the 04HelloRaytracing frame (04HelloRaytracing/deviceProgram.cu:66-243) into the
    // frame buffer; camera fields as 04HelloRaytracing/SampleRenderer.cpp:125-138 fills them.
    void renderWhitted(const sutil::Camera& camera, const float3& lightPos = float3{-907.108f, 2205.875f, -400.0267f}) {
        auto sub = [](const float3& a, const float3& b) { return float3{a.x - b.x, a.y - b.y, a.z - b.z}; };
        auto cross = [](const float3& a, const float3& b) { return float3{a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; };
        auto scaled = [](const float3& a, float s) { return float3{a.x * s, a.y * s, a.z * s}; };
        auto normalized = [&](const float3& a) { return scaled(a, 1.0f / std::sqrt(a.x * a.x + a.y * a.y + a.z * a.z)); };
        const float3 position = camera.eye(), direction = normalized(sub(camera.lookat(), camera.eye()));
        const float cosFovy = 0.66f, aspect = launchParams.frame.size.x / float(launchParams.frame.size.y);
        const float3 horizontal = scaled(normalized(cross(direction, camera.up())), cosFovy * aspect);
        const float3 vertical = scaled(normalized(cross(horizontal, direction)), cosFovy);
        b200pt_whitted_params wp = {{position.x, position.y, position.z}, {direction.x, direction.y, direction.z},
                                    {horizontal.x, horizontal.y, horizontal.z}, {vertical.x, vertical.y, vertical.z},
                                    {lightPos.x, lightPos.y, lightPos.z}};
        if (b200pt_render_whitted(ctx, &wp, nullptr) != 0) fail();
    }
    // Multi-GPU (no reference counterpart: sv4/main.cpp:289-290 "for nvlink configuration ... currently no support"): one SampleRenderer per GPU
    // (one host thread or process each), the scene replicated, ONE collective per frame partition, issued by the library (NCCL) on `stream`.
    //   sample split: config.accum_mode = B200PT_ACCUM_SUM, rank r renders sub-frames r, r+N, ..., then reduceSum(passes) -> image on `root`
    //   tile split  : config.tile_rank = r, config.tile_count = N (uniform schedules), render the SAME sub-frame, then gatherBands() -> image on `root`
    static void commUniqueId(unsigned char id[B200PT_UNIQUE_ID_BYTES]) { if (b200pt_comm_unique_id(id) != 0) fail(); }
    void commInit(int nRanks, int rank, const unsigned char id[B200PT_UNIQUE_ID_BYTES]) { if (b200pt_comm_init(ctx, nRanks, rank, id, &launchParams) != 0) fail(); }
    void commDestroy() { if (b200pt_comm_destroy(ctx, &launchParams) != 0) fail(); }
    void reduceSum(int nPasses, int root = 0, bool keepLocal = false) { if (b200pt_reduce_sum(ctx, &launchParams, &config, root, nPasses, keepLocal ? 1 : 0) != 0) fail(); }
    void gatherBands(int root = 0, bool frame = true, bool accum = false) { if (b200pt_gather_bands(ctx, &launchParams, root, (frame ? 1 : 0) | (accum ? 2 : 0)) != 0) fail(); }
    //   one-sided tile split: no communicator - root peerExport()s its frame buffer once and hands the 192 bytes to the other ranks, which peerAttach();
    //   from then on their finalize pass stores their bands straight into root's frame over NVLink.  Per frame on root: render, peerAcquire(),
    //   consume the frame (downloadPixels), peerRelease()
    void peerExport(int rank, int nRanks, unsigned char handle[B200PT_PEER_HANDLE_BYTES]) { if (b200pt_peer_export(ctx, &launchParams, rank, nRanks, handle) != 0) fail(); }
    void peerAttach(const unsigned char handle[B200PT_PEER_HANDLE_BYTES], int rank, int nRanks) { if (b200pt_peer_attach(ctx, handle, rank, nRanks) != 0) fail(); }
    void peerDetach() { if (b200pt_peer_detach(ctx) != 0) fail(); }
    void peerAcquire() { if (b200pt_peer_acquire(ctx) != 0) fail(); }
    void peerRelease() { if (b200pt_peer_release(ctx) != 0) fail(); }
    int peerState() { int st = 0;  if (b200pt_peer_status(ctx, &st, nullptr, nullptr) != 0) fail();  return st; }
    void clearAccum() { if (b200pt_clear_accum(ctx) != 0) fail(); }
    // animated geometry: new positions for mesh `meshID` (same count), then refit (topology kept) or a full rebuild
    void updateVertices(int meshID, const std::vector<float3>& vertex) {
        if (b200pt_update_vertices(ctx, meshID, reinterpret_cast<const float*>(vertex.data()), vertex.size()) != 0) fail();
    }
    void refitAccel() { if (b200pt_refit_accel(ctx) != 0) fail(); }
    void rebuildAccel() { uint64_t h = 0;  if (b200pt_build_accel(ctx, &h) != 0) fail();  launchParams.traversable = h; }
    // extras behind the same C-ABI
    void downloadAccum(float* h_rgba) { if (b200pt_download_accum(ctx, h_rgba) != 0) fail(); }
    b200pt_counters counters() { b200pt_counters c{};  if (b200pt_get_counters(ctx, &c) != 0) fail();  return c; }
    b200pt_bvh_info bvhInfo() { b200pt_bvh_info b{};  if (b200pt_bvh_info_get(ctx, &b) != 0) fail();  return b; }

    bool denoiserOn = true;              // carried for source compatibility; the reference's denoiser is a dead stub
    LaunchParams launchParams;           // public and live, like the reference (SimplePathtracer.h:111-187)
    void* stream = nullptr;              // CUstream of the context (public in the reference; main.cpp:284,300)
    b200pt_config config;                // the variant's compile-time switches as run-time fields
    b200pt_ctx* ctx = nullptr;
    sutil::Camera lastSetCamera;
    const Model* model;

private:
    [[noreturn]] static void fail() { throw std::runtime_error(b200pt_last_error()); }
};
