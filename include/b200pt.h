/* b200pt — C-ABI of the B200-native path tracer (libb200pt.so).
 *
 * This is the drop-in boundary for the reference's render hot path.  The reference has no FFI layer;
 * its boundary is the C++ class `SampleRenderer` (HelloPathtracing_sv4_vmv23/SimplePathtracer.h:45-188).
 * Every entry point below names the reference member / call site it replaces.  The C++ class in
 * include/ref_api/SimplePathtracer.h forwards 1:1 to these functions; Python binds them with ctypes.
 *
 * Conventions: plain pointers and sizes only (no torch / STL types); every function returns 0 on
 * success and -1 on failure, with the message retrievable through b200pt_last_error() (thread-local).
 * One context per GPU, externally synchronised (the reference is single-threaded).
 * There is NO CPU fallback: creating a context without a usable CUDA device fails.
 */
#ifndef B200PT_H
#define B200PT_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

struct LaunchParams;               /* include/ref_api/LaunchParams.h — 312-byte reference layout */
struct Material;                   /* include/ref_api/Material.h     — 104-byte reference layout */
typedef struct b200pt_ctx b200pt_ctx;

/* ---- render configuration: the reference's compile-time switches as run-time fields --------------
 * (SURVEY.md §5 "config / flag system").  Presets reproduce each variant's literals. */
enum { B200PT_SCHEDULE_AS_GIVEN = 0,   /* one launch, every LaunchParams field as the caller set it      */
       B200PT_SCHEDULE_UNIFORM  = 1,   /* FOV_OFF branch, sv4/SimplePathtracer.cpp:85-131                */
       B200PT_SCHEDULE_FOVEATED = 2 }; /* 3-pass sv4 schedule, sv4/SimplePathtracer.cpp:132-216          */
enum { B200PT_SHADOW_ANY_HIT = 0,      /* orig: TERMINATE_ON_FIRST_HIT, orig/deviceProgram.cu:198        */
       B200PT_SHADOW_CULL_BACK = 1 };  /* sv3/sv4: CULL_BACK_FACING_TRIANGLES, sv4/deviceProgram.cu:240  */
enum { B200PT_DISPLAY_SRGB = 0,        /* make_color(accum), orig/deviceProgram.cu:467                   */
       B200PT_DISPLAY_REINHARD = 1 };  /* make_color(reinhard(accum*exposure, 1)), sv4/deviceProgram.cu:557-568 */
enum { B200PT_ACCUM_LERP = 0,          /* running mean lerp(prev, clamp(cur,0,10), 1/(n+1)), sv4/deviceProgram.cu:545-553 */
       B200PT_ACCUM_SUM  = 1 };        /* multi-GPU sample split: accum_buffer holds the SUM of clamped passes
                                          (associative across ranks); finish with b200pt_resolve_sum()   */

typedef struct b200pt_config {
    int   schedule;           /* B200PT_SCHEDULE_*                                                       */
    int   max_depth;          /* loop breaks when depth >= max_depth: 4 (sv3/sv4), 8 (orig), 3 (sv/sv2)  */
    float tmin, tmax;         /* 0.01 / 1e16 (sv3/sv4), 0.001 / 1e16 (orig)                              */
    int   shadow_mode;        /* B200PT_SHADOW_*                                                         */
    int   display;            /* B200PT_DISPLAY_*                                                        */
    float exposure;           /* 4 = 2^2 (sv4), 8 = 2^3 (sv3); ignored for DISPLAY_SRGB                  */
    int   uniform_spl;        /* SCHEDULE_UNIFORM: samples_per_launch written by render(); 0 = keep caller's */
    int   accumulate;         /* SCHEDULE_UNIFORM: 0 = force subframe_index 0 (sv4 FOV_OFF), 1 = progressive (sv3/orig) */
    int   inner_radius, outer_radius;   /* SCHEDULE_FOVEATED ring radii, 157 / 515                       */
    int   launch_w, launch_h; /* SCHEDULE_AS_GIVEN launch dimensions; 0 = frame.size                     */
    int   write_aovs;         /* 1 = also write normal/color/albedo buffers (orig/deviceProgram.cu:472-474) */
    int   increment_subframe; /* 1 = render() does subframe_index++ (sv3/sv4); 0 = caller does (orig)    */
    int   accum_mode;         /* B200PT_ACCUM_*                                                          */
    int   russian_roulette;   /* 0 = off (reference behaviour).  1 = opt-in extension: from the third segment on a path survives
                                 with probability min(1, max(throughput)); one extra Randf per decision (RNG stream differs) */
    int   tile_rank, tile_count; /* multi-GPU tile split: this context renders launch rows whose 32-row band
                                    index % tile_count == tile_rank; tile_count <= 1 = whole frame        */
    int   profile;            /* bit 0: per-stage CUDA-event timings; bit 1: traversal counters (slower kernels) */
    int   mis_on_miss;        /* 0 = off (reference behaviour: a BSDF-sampled ray that leaves the scene adds nothing, sv4/deviceProgram.cu:251-280).
                                 1 = opt-in extension, the reference's commented-out MIS (:257-267): a secondary ray that misses adds
                                 w * ProbeEval(dir) * throughput with w = p_bsdf / (p_bsdf + ProbePdf(dir)), the partner of the NEE weight */
    float shadow_tmin;        /* tmin of the occlusion (NEE) rays: 0.01 in every variant — `orig` traces radiance rays from 0.001 (orig/deviceProgram.cu:420)
                                 but occlusion rays from 0.01 (:268, :310); <= 0 means "same as tmin" */
    int   frames_in_flight;   /* how many frames may be in flight at once (1..4; 0 = 1).  Each frame in flight owns a stream, a path pool and a queue set;
                                 its wavefront runs concurrently with its neighbours' (one frame's drain phases and thin late bounces overlap the next
                                 frame's wide early ones) and frames still fold into the accumulation buffer in submission order, so the pixels are
                                 those of the one-at-a-time sequence, bit for bit.
                                 b200pt_render_async: consecutive calls overlap.
                                 b200pt_render (blocking): once a call has looked like progressive accumulation (the previous call's launchParams,
                                 subframe_index advanced by a constant stride), the wavefronts of the predicted next frames_in_flight - 1 calls are
                                 enqueued before the call blocks on its own frame; a call that matches its prediction byte for byte (launchParams and
                                 config) only adds its finalize passes, any other call - a camera move, a new gaze point - drops the speculative
                                 wavefronts, which have written nothing outside their own lanes.  Entry points that change the scene, the probe or the
                                 frame buffers drop them too.  Instrumented frames (profile != 0) always run one at a time on the primary stream. */
    int   intra_frame_split;  /* 0/1 = off.  2..4: with frames_in_flight <= 1 (strictly one frame at a time, e.g. a camera or gaze point that moves every frame, so
                                 that nothing can be rendered ahead) the frame is split into that many interleaved sets of 32-row bands, each a wavefront on
                                 its own stream; the sets' kernels fill each other's drain phases and thin late bounces.  Same partition as tile_rank/tile_count,
                                 same seeds per launch index: pixels are bit-identical to the unsplit frame. */
} b200pt_config;

/* Fill *cfg with a named preset: "orig", "sv3_uniform", "sv4_uniform", "sv4_fov". */
int b200pt_config_preset(b200pt_config* cfg, const char* name);

/* ---- lifetime: replaces SampleRenderer ctor plumbing (sv4/SimplePathtracer.cpp:42-75, 313-343) ---- */
b200pt_ctx* b200pt_create(int device);            /* NULL on failure (see b200pt_last_error)            */
void        b200pt_destroy(b200pt_ctx* ctx);
const char* b200pt_last_error(void);

/* ---- scene upload: replaces the per-mesh uploads + build inputs of buildAccel()
 *      (sv4/SimplePathtracer.cpp:603-664), createTextures() (:749-800) and the SBT records (:574-595). */
int b200pt_add_texture(b200pt_ctx* ctx, const uint32_t* rgba8, int width, int height);   /* returns textureID >= 0, or -1 */
int b200pt_add_mesh(b200pt_ctx* ctx,
                    const float* vertex_xyz, size_t n_vertices,
                    const float* normal_xyz,            /* may be NULL (never read by the path)          */
                    const float* texcoord_uv,           /* may be NULL                                   */
                    const uint32_t* index_xyz, size_t n_triangles,
                    const struct Material* material,
                    int diffuse_texture_id);            /* -1 = none; returns meshID >= 0, or -1         */
/* LBVH build: Morton codes -> radix sort -> Karras hierarchy -> refit -> SAH leaf collapse + treelet
 * refinement -> 32-byte two-child records (16-bit quantised boxes), collapsed to 64-byte 4-wide nodes, + 48-byte triangle records.  Replaces optixAccelBuild/optixAccelCompact
 * (sv4/SimplePathtracer.cpp:670-737).  *handle_out receives the value for LaunchParams.traversable. */
int b200pt_build_accel(b200pt_ctx* ctx, uint64_t* handle_out);

/* Animated geometry (rebuild-per-frame / refit; the reference builds once, sv4/SimplePathtracer.cpp:670-737).
 * b200pt_update_vertices replaces one mesh's positions (same vertex count as b200pt_add_mesh; indices, normals, texcoords kept).
 * Then either b200pt_refit_accel — topology kept: new triangle records, bottom-up bounds, re-quantisation over the new scene box,
 * O(N) with no sort (OptiX's OPTIX_BUILD_OPERATION_UPDATE) — or b200pt_build_accel again for a full rebuild.  Hits stay exact
 * either way (the triangle test does not depend on the tree); only traversal cost degrades with large deformations under refit. */
int b200pt_update_vertices(b200pt_ctx* ctx, int mesh_id, const float* vertex_xyz, size_t n_vertices);
int b200pt_refit_accel(b200pt_ctx* ctx);

/* ---- frame state: SampleRenderer::resize / setProbe (sv4/SimplePathtracer.cpp:231-277, 295-311) ---- */
int b200pt_resize(b200pt_ctx* ctx, int width, int height, struct LaunchParams* params /* buffers + size written */);
int b200pt_set_probe(b200pt_ctx* ctx, int width, int height, const float* data_rgba,
                     const float* pdfX, const float* cdfX, const float* pdfY, const float* cdfY,
                     const float* offset_xyz, struct LaunchParams* params /* probe pointers written */);

/* ProbeData::BuildCDF on the device WITHOUT any allocation.  No counterpart in the reference: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same kernels and the same
 * bits as b200pt_set_probe_image; the only scratch they need (one float per probe row) lives in a file-scope `static __device__` array of
 * B200PT_NOALLOC_MAX_ROWS floats (64 KiB, held by every GPU that loads the library until the process exits).  The caller passes DEVICE pointers:
 * width*height float4 texels in; pdfX, cdfX (width*height floats each) and pdfY, cdfY (height floats each) out.  The launches go to `stream`
 * (NULL = the context's primary stream) and are not waited for.  Calls no cudaMalloc / cudaFree / allocating library routine; the first call on a
 * device resolves the array's address (cudaGetSymbolAddress) - make one call before starting a graph capture.  height > B200PT_NOALLOC_MAX_ROWS
 * is rejected with an error.  Every call on a GPU shares the array: calls that overlap on the GPU race on it, serialise them (one stream). */
#define B200PT_NOALLOC_MAX_ROWS 16384
int b200pt_build_cdf_noalloc(b200pt_ctx* ctx, int width, int height, const void* d_data_rgba_f32,
                             void* d_pdfX, void* d_cdfX, void* d_pdfY, void* d_cdfY, void* stream);

/* ---- second integrator on the same traversal kernels: 04HelloRaytracing (deviceProgram.cu:66-243) -----
 * One primary ray per pixel centre (tmin 0, tmax 1e20), interpolated shading normal (geometric normal when the mesh has none),
 * diffuse colour x texture, one shadow ray to a point light (origin + 1e-3 Ng, unnormalised direction, t in (1e-3, 1 - 1e-3),
 * first hit terminates, no culling), colour = (0.1 + (0.2 + 0.8 vis) (0.1 + 0.8 |d.Ns|)) * diffuse, RGBA8 = int(255.99 c).
 * Camera fields as SampleRenderer::setCamera fills them (04HelloRaytracing/SampleRenderer.cpp:125-138). */
typedef struct b200pt_whitted_params {
    float position[3], direction[3], horizontal[3], vertical[3];
    float light_pos[3];                 /* the reference hard-codes (-907.108, 2205.875, -400.0267), deviceProgram.cu:125 */
} b200pt_whitted_params;
int b200pt_render_whitted(b200pt_ctx* ctx, const b200pt_whitted_params* params,
                          uint32_t* device_color_buffer /* NULL = the context's frame buffer (b200pt_download_pixels) */);

/* loadProbe/loadColor + ProbeData::BuildCDF + setProbe in one call (sv4/main.cpp:154-180, sv4/Probe.h:29-77,
 * sv4/SimplePathtracer.cpp:295-311): uploads width*height float4 texels and builds pdfValuesX/cdfValuesX/pdfValuesY/cdfValuesY ON THE
 * DEVICE, bit-identical to the host's sequential fp32 BuildCDF (one warp per row keeps the accumulation order). */
int b200pt_set_probe_image(b200pt_ctx* ctx, int width, int height, const float* data_rgba, const float* offset_xyz,
                           struct LaunchParams* params /* probe pointers written */);
/* Copies the current probe's device tables to the host (any pointer may be NULL): width*height, width*height, height, height floats. */
int b200pt_download_probe(b200pt_ctx* ctx, float* pdfX, float* cdfX, float* pdfY, float* cdfY);

/* ---- the hot path: SampleRenderer::render() (sv4/SimplePathtracer.cpp:77-217) ---------------------
 * Reads the caller's live LaunchParams (camera, frame.c, subframe_index, ...), runs the wavefront
 * integrator for the configured schedule, writes accum/frame buffers, updates *params exactly as the
 * reference's render() leaves its public `launchParams`, and blocks until the GPU is idle. */
int b200pt_render(b200pt_ctx* ctx, struct LaunchParams* params, const b200pt_config* cfg);
/* Same, without the trailing device synchronisation; with cfg->frames_in_flight > 1 consecutive calls overlap on the device (see above). */
int b200pt_render_async(b200pt_ctx* ctx, struct LaunchParams* params, const b200pt_config* cfg);
/* Blocks until every frame submitted so far has finished. */
int b200pt_synchronize(b200pt_ctx* ctx);
/* Orders the context's primary stream (b200pt_device_buffers) after every frame in flight, without blocking the host: work the caller then
 * enqueues on that stream - an event, a copy, an NCCL reduce of the accumulation buffer - sees all submitted frames.  Every other entry
 * point that touches the frame buffers, the scene or the probe does this itself. */
int b200pt_join(b200pt_ctx* ctx);

/* SampleRenderer::downloadPixels (sv4/SimplePathtracer.cpp:279-283): W*H RGBA8, R in the low byte. */
int b200pt_download_pixels(b200pt_ctx* ctx, uint32_t* host_rgba8);
/* The same copy for frames in flight (b200pt_render_async): enqueued behind the most recently submitted frame on that frame's stream, and the
 * next frame's write of the frame buffer waits for it; `host_rgba8` (pinned memory) is valid after b200pt_synchronize. */
int b200pt_download_pixels_async(b200pt_ctx* ctx, uint32_t* host_rgba8);
int b200pt_download_accum(b200pt_ctx* ctx, float* host_rgba_f32);                /* W*H float4             */
/* Restore a progressive state saved with b200pt_download_accum (checkpoint/resume of accum_buffer). */
int b200pt_upload_accum(b200pt_ctx* ctx, const float* host_rgba_f32);
int b200pt_download_aov(b200pt_ctx* ctx, int which /*0 normal,1 albedo,2 color*/, float* host_rgba_f32);
/* Device pointers of the internal buffers (for NCCL reduce/gather by the multi-GPU driver). */
int b200pt_device_buffers(b200pt_ctx* ctx, void** accum_f4, void** frame_rgba8, void** stream);
/* ACCUM_SUM epilogue on rank 0: accum = sum / n_passes, then the display transform into frame_buffer. */
int b200pt_resolve_sum(b200pt_ctx* ctx, const struct LaunchParams* params, const b200pt_config* cfg, int n_passes);
int b200pt_clear_accum(b200pt_ctx* ctx);

/* ---- multi-GPU: one context per rank (one process or one thread per GPU), the scene replicated, ONE collective per frame partition --------
 * The reference has no multi-GPU path (sv4/main.cpp:289-290: "for nvlink configuration ... currently no support"; the SDK's buffer modes are
 * sutil/CUDAOutputBuffer.h:45-51).  These entry points are what a maintainer's multi-GPU driver binds: the collectives are NCCL's, issued by
 * the library on the context's primary stream behind every frame in flight; libnccl.so.2 is opened at run time (no link-time dependency).
 *   sample split  rank r renders sub-frames r, r+N, ... with cfg.accum_mode = B200PT_ACCUM_SUM, then b200pt_reduce_sum
 *   tile split    rank r renders with cfg.tile_rank = r, cfg.tile_count = N (uniform schedules), then b200pt_gather_bands                      */
#define B200PT_UNIQUE_ID_BYTES 128
/* ncclGetUniqueId on the calling rank; the host distributes the 128 bytes to the other ranks with its own plumbing (MPI, torch.distributed, a file). */
int b200pt_comm_unique_id(void* id_out_128);
/* ncclCommInitRank (collective over all ranks).  From here on the accumulation buffer lives in ncclMemAlloc'd memory registered with the
 * communicator - an existing one is moved (contents kept) and params->frame.accum_buffer, if it pointed at it, is rewritten (params may be NULL). */
int b200pt_comm_init(b200pt_ctx* ctx, int n_ranks, int rank, const void* id_128, struct LaunchParams* params);
/* Drops the communicator; the accumulation buffer moves back to a plain allocation (contents kept) and params->frame.accum_buffer follows it. */
int b200pt_comm_destroy(b200pt_ctx* ctx, struct LaunchParams* params);
/* n_ranks, rank, NCCL version code, accum_registered: bit 0 = ncclMemAlloc'd, bit 1 = ncclCommRegister'ed.  Any pointer may be NULL. */
int b200pt_comm_info(b200pt_ctx* ctx, int* n_ranks, int* rank, int* nccl_version, int* accum_registered);
/* Sample split epilogue: fp32 SUM ncclReduce of params->frame.accum_buffer onto `root`, followed there by sum / n_passes and the display
 * transform into frame_buffer.  keep_local = 0: in place, root's accumulation buffer ends up holding the mean (b200pt_resolve_sum; end of a
 * job).  keep_local = 1: progressive display - every rank keeps its partial sum, root receives the total into a buffer of its own and only
 * refreshes frame_buffer, so the ranks go on accumulating and reduce again later (n_passes = passes summed so far over all ranks).
 * Enqueued on the primary stream; b200pt_synchronize or a download waits for it. */
int b200pt_reduce_sum(b200pt_ctx* ctx, const struct LaunchParams* params, const b200pt_config* cfg, int root, int n_passes, int keep_local);
/* Tile split epilogue: the 32-row bands each rank rendered (band % n_ranks == rank) are sent to `root`, which receives them in place.
 * what: bit 0 = RGBA8 frame buffer, bit 1 = float4 accumulation buffer.  One grouped ncclSend / ncclRecv batch on the primary stream. */
int b200pt_gather_bands(b200pt_ctx* ctx, const struct LaunchParams* params, int root, int what);

/* ---- one-sided tile split over NVLink peer memory: the exchange done by the producing kernel, no collective ---------------------------------
 * No reference counterpart (same note as above).  Root exports its RGBA8 frame buffer once; the other ranks map it (CUDA IPC between processes,
 * peer access inside one process) and from then on every frame they render with cfg.tile_count > 1 (and not ACCUM_SUM) writes its 32-row bands
 * from the finalize pass straight into root's frame buffer, followed by a release-store of the frame's epoch into root's flag block.
 *   root, once:         b200pt_peer_export(ctx, params, root_rank, n_ranks, handle)  -> hand the 192 bytes to every other rank
 *   other ranks, once:  b200pt_peer_attach(ctx, handle, rank, n_ranks)
 *   every frame:        all ranks b200pt_render[_async] with cfg.tile_rank / tile_count;
 *                       root: b200pt_peer_acquire (primary stream waits until every rank's bands of this frame have landed), consume the frame
 *                       (b200pt_download_pixels...), b200pt_peer_release (the ranks' next finalize pass waits for it: no frame is overwritten early)
 * Epochs count frames: after the export / attach every rank must render the same sequence of tile-split frames (as any frame partition requires).
 * Nothing synchronises with the host.  A wait that sees no progress for B200PT_PEER_TIMEOUT_MS (environment, default 5000) gives up and records
 * its code (b200pt_peer_status: 0 = fine, 1 + r = root timed out waiting for rank r, 1000 + 64 = a rank timed out waiting for root's release, 2000 + r = a pass barrier timed out waiting for rank r).
 * Schedules with several passes (the foveated presets: a later pass overwrites pixels of an earlier one) work too: the partition is by 32-row bands of
 * every pass's launch rows (each rank gets an equal share of fovea, mid ring and periphery), and between two passes every rank publishes "pass k done"
 * and waits for all the others (the same flag block), so the overwrite order of a single GPU is kept; such a frame is byte-identical to one GPU when it
 * does not read the accumulation buffer (subframe_index 0 or redraw: the interactive mode), and root (cfg.tile_rank / tile_count = the exported rank /
 * n_ranks) takes part in the barriers.
 * The accumulation buffer of a band stays on the rank that renders it (progressive accumulation needs no exchange); a resize drops the export /
 * the mapping (export and attach again).  Pixels are byte-identical to b200pt_gather_bands' and to a single-GPU frame.
 * Ranks that share ONE GPU inside ONE process (tests): CUDA loads kernels lazily and a load synchronises the context, so every such rank must have
 * rendered a frame and attached before any of them starts waiting; ranks on different GPUs or in different processes need no such care. */
#define B200PT_PEER_HANDLE_BYTES 192
int b200pt_peer_export(b200pt_ctx* ctx, const struct LaunchParams* params, int rank, int n_ranks, void* handle_192);
int b200pt_peer_attach(b200pt_ctx* ctx, const void* handle_192, int rank, int n_ranks);
int b200pt_peer_detach(b200pt_ctx* ctx);
int b200pt_peer_acquire(b200pt_ctx* ctx);
int b200pt_peer_release(b200pt_ctx* ctx);
/* state: see above; attached_rank: this context's rank in a partition it attached to (-1: none); epoch: frames acquired (root) / put (others).
 * Waits for the context's streams.  Any pointer may be NULL. */
int b200pt_peer_status(b200pt_ctx* ctx, int* state, int* attached_rank, unsigned* epoch);

/* computeFinalPixelColors (sv4/toneMap.cu:41-70, toneMap.h:4): sqrt-gamma float4 -> RGBA8, device pointers. */
int b200pt_tonemap_sqrt(int width, int height, const void* d_float4_in, void* d_rgba8_out, void* stream);
/* Host-buffer convenience around the same kernel (H2D + kernel + D2H). */
int b200pt_tonemap_sqrt_host(b200pt_ctx* ctx, int width, int height, const float* h_float4_in, uint32_t* h_rgba8_out);

/* ---- parity hooks (tests only; none of these is on the timed path) ---------------------------------*/
/* Primary-ray closest hit for every launch index (jitter of sample 0, i.e. exactly the first radiance
 * optixTrace of sv4/deviceProgram.cu:481): mesh = build-input index, prim = index within that mesh. */
int b200pt_trace_primary(b200pt_ctx* ctx, const struct LaunchParams* params, const b200pt_config* cfg,
                         int32_t* prim, int32_t* mesh, float* t, float* u, float* v);
/* Arbitrary rays: o/d are n*3 floats.  any_hit != 0 -> occlusion query with cfg->shadow_mode (result in prim: 1/0). */
int b200pt_trace_rays(b200pt_ctx* ctx, const float* origin_xyz, const float* dir_xyz, size_t n,
                      float tmin, float tmax, int any_hit, int shadow_mode,
                      int32_t* prim, int32_t* mesh, float* t, float* u, float* v);
/* Device-level known-answer hooks for the shading library, on caller-supplied inputs (n items each; float3 arrays are n*3 floats):
 * BasisFromVector(N) (sv4/maths.h:94-108); BSDFSample with a fresh Random(seeds[i]) (sv4/Disney.cuh:197-315: direction, pdf, lobe type and the
 * generator's state afterwards - the number of draws is data dependent); BSDFEval / BSDFPdf for the given L (:318-427, :152-193).
 * These run the very functions shadeKernel inlines, so a lobe-selection or pdf bug shows up here item by item instead of hiding in an image. */
int b200pt_debug_bsdf(b200pt_ctx* ctx, size_t n, const struct Material* materials, const float* etaI, const float* etaO,
                      const float* N_xyz, const float* view_xyz, const float* L_xyz, const float* albedo_rgb, const uint32_t* seeds,
                      float* basisU_xyz, float* basisV_xyz, float* sampleL_xyz, float* samplePdf, int32_t* sampleType, uint32_t* sampleRng_2,
                      float* evalF_rgb, float* evalPdf);
/* ProbeSample with a fresh Random(seeds[i]) (sv4/Probe.cuh:138-169) on the probe in *params (b200pt_set_probe*): use_guide != 0 searches the
 * CDFs through the guide tables (what render() does), 0 with the reference's plain LowerBound (:119-136) - both must return the same texel;
 * and ProbeDirToUV / ProbeEval / ProbePdf (:38-46, :61-67, :69-93) of m given directions. */
int b200pt_debug_probe(b200pt_ctx* ctx, const struct LaunchParams* params, int use_guide, size_t n, const uint32_t* seeds,
                       float* sampleDir_xyz, float* sampleColor_rgb, float* samplePdf,
                       size_t m, const float* dirs_xyz, float* uv, float* evalColor_rgb, float* evalPdf);
/* Integer BVH-build stages, bit-exact against a host reference: 63-bit Morton code (21 bits/axis) per
 * global triangle (input order) and the stable radix-sort permutation.  scene_bounds = 6 floats (lo xyz, hi xyz). */
int b200pt_debug_morton(b200pt_ctx* ctx, uint64_t* codes, uint32_t* sorted_order, float* scene_bounds);
typedef struct b200pt_bvh_info {
    uint64_t n_triangles, n_nodes, n_leaves;
    uint32_t max_leaf_size, max_depth;
    float    sah_cost;           /* SAH cost of the final tree (Ct = 1, Ci = 1.2) */
    float    sah_cost_lbvh;      /* before refinement */
    float    build_ms;
    uint64_t node_bytes, tri_bytes;      /* node_bytes = the 4-wide nodes */
    int32_t  root_link;          /* >= 0: record index of the root's node; < 0: leaf code; 0x7fffffff: empty scene */
    uint32_t refine_levels;      /* treelet-refinement levels run by the build */
    double   grid_lo[3], grid_scale[3];   /* node boxes are quantised: g = 2 + (x - grid_lo) * grid_scale, 16 bits per bound */
    float    refit_ms;           /* device time of the last b200pt_refit_accel (0 if none since the build) */
    uint32_t n_wide_nodes;       /* 64-byte 4-wide nodes the traversal kernels read (the binary records collapsed) */
    uint32_t wide_depth, reserved;
} b200pt_bvh_info;
int b200pt_bvh_info_get(b200pt_ctx* ctx, b200pt_bvh_info* out);
/* Copies of the flattened BVH for structural validation in tests: n_nodes 32-byte records, one per inner node, each
 * {6 x uint32: 12 quantised 16-bit bounds of child 0 then child 1 (lo.x|hi.x<<16, lo.y|hi.y<<16, lo.z|hi.z<<16), int32 link0, link1};
 * tri_global_id = triangle ids in leaf order. */
int b200pt_download_bvh(b200pt_ctx* ctx, void* nodes32, size_t max_nodes, uint32_t* tri_global_id, size_t max_tris);
/* The 4-wide tree the traversal kernels read: n_wide_nodes 64-byte records {12 x uint32: child c's box in words 3c..3c+2 (same
 * packing as above), int32 link[4]}; unused slots: box lo 65535 / hi 0, link 0x7fffffff; the root is record 0. */
int b200pt_download_bvh4(b200pt_ctx* ctx, void* nodes64, size_t max_nodes);

typedef struct b200pt_counters {      /* last render()                                                  */
    uint64_t samples;                 /* camera paths started                                           */
    uint64_t radiance_rays, shadow_rays;
    uint64_t nodes_fetched, tris_fetched;      /* 64-byte 4-wide nodes / 48-byte triangle records (profile bit 1) */
    uint64_t nodes_fetched_shadow, tris_fetched_shadow;
    float    ms_total, ms_raygen, ms_trace_closest, ms_trace_any, ms_shade, ms_accumulate;   /* profile=1 */
    uint32_t kernel_launches;         /* kernels launched by the last render()                          */
    uint32_t n_trace_closest_launches, n_trace_any_launches;
    uint32_t fused_finalize;          /* 1: the frame's finalize passes summed the samples straight from the path pool (no gather pass) */
    uint64_t closest_hits;            /* radiance rays that hit a surface = entries shadeKernel shades (profile bit 1)  */
    uint64_t occluded_shadow_rays;    /* (profile bit 1)                                                                */
} b200pt_counters;
int b200pt_get_counters(b200pt_ctx* ctx, b200pt_counters* out);

#ifdef __cplusplus
}
#endif
#endif /* B200PT_H */
