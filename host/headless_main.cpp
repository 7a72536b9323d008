// Headless driver: the reference application's call sequence (HelloPathtracing_sv4_vmv23/main.cpp:182-427) without the
// GLFW/OpenGL window (no graphics stack on this box): loadOBJ -> probe -> camera -> SampleRenderer -> setCamera -> resize ->
// setProbe -> loop { gaze -> render -> fps log } -> downloadPixels -> PPM.  The gaze point, which the reference takes from the
// mouse cursor (main.cpp:310-314), follows a scripted Lissajous track.  fps is appended every 0.5 s in the reference's
// log format (sutil/sutil.cpp:760-772: one value per line).
//
//   b200pt_headless [--obj file.obj | --boxes N] [--preset sv4_fov|sv4_uniform|sv3_uniform|orig] [--size WxH] [--frames N] [--spl N] [--in-flight N [--async]]
//                   [--camera sponza|lost_empire|san_miguel|cornell] [--sky r,g,b] [--out image.ppm|image.png] [--log fps.tsv]
//                   [--hdr probe.hdr] [--gpu-cdf] [--whitted] [--dump-model file.bin] [--dump-probe file.bin] [--no-render]
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <future>
#include <memory>
#include <random>
#include <string>
#include <thread>
#include <vector>

#include "../include/ref_api/SimplePathtracer.h"

// Minimal PNG writer (8-bit RGB, zlib "stored" blocks): the reference saves screenshots through stb_image_write (sutil.cpp:723-739),
// which is third-party code of the reference and not copied.  rows: top row first.
static void writePNG(const std::string& path, int W, int H, const std::vector<unsigned char>& rgbTopDown) {
    auto be32 = [](std::vector<unsigned char>& v, uint32_t x) { for (int s = 24; s >= 0; s -= 8) v.push_back((unsigned char)(x >> s)); };
    static uint32_t table[256];
    if (!table[1]) for (uint32_t i = 0; i < 256; ++i) { uint32_t c = i;  for (int k = 0; k < 8; ++k) c = (c & 1) ? 0xEDB88320u ^ (c >> 1) : c >> 1;  table[i] = c; }
    auto crc = [&](const unsigned char* p, size_t n) { uint32_t c = 0xffffffffu;  for (size_t i = 0; i < n; ++i) c = table[(c ^ p[i]) & 0xff] ^ (c >> 8);  return c ^ 0xffffffffu; };
    std::vector<unsigned char> raw;                                          // filter byte 0 + RGB per row
    raw.reserve((size_t)H * (3 * (size_t)W + 1));
    for (int y = 0; y < H; ++y) { raw.push_back(0);  raw.insert(raw.end(), rgbTopDown.begin() + (size_t)y * 3 * W, rgbTopDown.begin() + (size_t)(y + 1) * 3 * W); }
    std::vector<unsigned char> z = {0x78, 0x01};
    uint32_t a = 1, b = 0;
    for (size_t off = 0; off < raw.size() || off == 0; off += 65535) {
        const size_t len = std::min<size_t>(65535, raw.size() - off);
        z.push_back(off + len >= raw.size() ? 1 : 0);
        z.push_back((unsigned char)(len & 255));  z.push_back((unsigned char)(len >> 8));
        z.push_back((unsigned char)(~len & 255));  z.push_back((unsigned char)((~len >> 8) & 255));
        z.insert(z.end(), raw.begin() + off, raw.begin() + off + len);
        for (size_t i = 0; i < len; ++i) { a = (a + raw[off + i]) % 65521u;  b = (b + a) % 65521u; }
        if (raw.empty()) break;
    }
    be32(z, (b << 16) | a);
    FILE* f = fopen(path.c_str(), "wb");
    if (!f) throw std::runtime_error("cannot write " + path);
    const unsigned char sig[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
    fwrite(sig, 1, 8, f);
    auto chunk = [&](const char* type, const std::vector<unsigned char>& data) {
        std::vector<unsigned char> c(type, type + 4);
        c.insert(c.end(), data.begin(), data.end());
        std::vector<unsigned char> hdr;  be32(hdr, (uint32_t)data.size());
        std::vector<unsigned char> tail;  be32(tail, crc(c.data(), c.size()));
        fwrite(hdr.data(), 1, 4, f);  fwrite(c.data(), 1, c.size(), f);  fwrite(tail.data(), 1, 4, f);
    };
    std::vector<unsigned char> ihdr;
    be32(ihdr, (uint32_t)W);  be32(ihdr, (uint32_t)H);
    const unsigned char rest[5] = {8, 2, 0, 0, 0};
    ihdr.insert(ihdr.end(), rest, rest + 5);
    chunk("IHDR", ihdr);  chunk("IDAT", z);  chunk("IEND", {});
    fclose(f);
}

static sutil::Camera cameraPreset(const std::string& name) {                                 // main.cpp:211-240
    if (name == "sponza") return sutil::Camera({-1293.07f, 154.681f, 1.0f}, {0.f, 200.f, 0.f}, {0.f, 1.f, 0.f}, 35.f, 1.f);
    if (name == "san_miguel") return sutil::Camera({26.f, 8.f, -2.f}, {0.f, 0.f, 0.f}, {0.f, 1.f, 0.f}, 35.f, 1.f);
    if (name == "lost_empire") return sutil::Camera({-70.f, 40.f, 100.f}, {0.f, 0.f, 0.f}, {0.f, 1.f, 0.f}, 45.f, 1.f);
    return sutil::Camera({0.f, 1.f, 3.2f}, {0.f, 1.f, 0.f}, {0.f, 1.f, 0.f}, 40.f, 1.f);
}

static Model* boxScene(int n, unsigned seed) {          // addBox-only scene (Model.cpp:218-290): a floor and n random boxes
    Model* m = new Model;
    std::mt19937 rng(seed);
    std::uniform_real_distribution<float> u(0.f, 1.f);
    Material floor = Material::origDefaults();
    addBox(m, floor, {0.f, -0.05f, 0.f}, {3.f, 0.05f, 3.f});
    for (int i = 0; i < n; ++i) {
        Material mat = (i % 3 == 0) ? Material() : Material::origDefaults();
        mat.color = {0.2f + 0.7f * u(rng), 0.2f + 0.7f * u(rng), 0.2f + 0.7f * u(rng)};
        mat.emission = {0.f, 0.f, 0.f};
        const float s = 0.08f + 0.25f * u(rng);
        addBox(m, mat, {-1.5f + 3.f * u(rng), s, -1.5f + 3.f * u(rng)}, {s, s, s});
    }
    return m;
}

static void dumpModel(const Model* m, const char* path) {    // flat binary for the loadOBJ parity test (tests/test_cpp_host.py)
    FILE* f = fopen(path, "wb");
    if (!f) throw std::runtime_error(std::string("cannot write ") + path);
    const int nm = (int)m->meshes.size(), nt = (int)m->textures.size();
    fwrite(&nm, 4, 1, f);  fwrite(&nt, 4, 1, f);
    for (auto mesh : m->meshes) {
        const int c[5] = {(int)mesh->vertex.size(), (int)mesh->normal.size(), (int)mesh->texcoord.size(), (int)mesh->index.size(), mesh->diffuseTextureID};
        fwrite(c, 4, 5, f);
        fwrite(&mesh->material, sizeof(Material), 1, f);
        fwrite(mesh->vertex.data(), 12, mesh->vertex.size(), f);
        fwrite(mesh->normal.data(), 12, mesh->normal.size(), f);
        fwrite(mesh->texcoord.data(), 8, mesh->texcoord.size(), f);
        fwrite(mesh->index.data(), 12, mesh->index.size(), f);
    }
    for (auto t : m->textures) {
        fwrite(&t->resolution, 8, 1, f);
        fwrite(t->pixel, 4, (size_t)t->resolution.x * t->resolution.y, f);
    }
    fclose(f);
}

// --ranks N: the multi-GPU driver below Python.  One host thread and one SampleRenderer per GPU in this process, the scene replicated, the frame
// partitioned by samples (--split sample: rank r renders sub-frames r, r+N, ..., ONE fp32 SUM reduce + resolve on rank 0) or by 32-row bands
// (--split tile: every frame gathered to rank 0; --split peer: the same partition, the other ranks' finalize passes write into rank 0's frame buffer).  The collectives are the library's (b200pt_reduce_sum / b200pt_gather_bands over NCCL); the
// 128-byte NCCL id is shared through this process' memory.  Prints the checksum of rank 0's final image, comparable with a single-GPU run.
static int multiMain(const Model* model, const ProbeData& probe, const sutil::Camera& camera, const std::string& preset, int W, int H, int frames, int spl,
                     int ranks, int mode /* 0 sample, 1 tile (gather), 2 tile, one-sided over peer memory */, int devices, const std::string& out) {
    const bool tileSplit = mode != 0, oneSided = mode == 2;
    unsigned char uid[B200PT_UNIQUE_ID_BYTES];
    if (!oneSided) SampleRenderer::commUniqueId(uid);
    unsigned char peerHandle[B200PT_PEER_HANDLE_BYTES];
    std::promise<void> exported;                                   // rank 0 fulfils it once peerHandle is valid
    std::shared_future<void> exportedF = exported.get_future().share();
    std::vector<std::string> errors(ranks);
    std::vector<uint32_t> px((size_t)W * H);
    std::vector<double> seconds(ranks, 0.0);
    unsigned long long samples0 = 0;
    std::atomic<int> ready{0};
    std::atomic<bool> failed{false};
    auto worker = [&](int rank) {
        try {
            SampleRenderer sample(model, devices > 0 ? rank % devices : rank, preset.c_str());   // --devices 1: every rank on GPU 0 (one-sided split only: NCCL wants one GPU per rank)
            sample.setCamera(camera);
            sample.resize(int2{W, H});
            sample.setProbe(probe);
            sample.launchParams.samples_per_launch = spl;
            if (!oneSided) sample.commInit(ranks, rank, uid);
            else if (ranks > 1) {
                // one throw-away frame first: every kernel of a frame is loaded (CUDA loads kernels lazily, and a load synchronises the context - with
                // ranks sharing one GPU that must not happen while another rank's stream is already waiting for this rank's bands)
                sample.config.tile_rank = rank;  sample.config.tile_count = ranks;
                sample.render();
                if (rank == 0) {
                    try { sample.peerExport(0, ranks, peerHandle);  exported.set_value(); }
                    catch (...) { exported.set_exception(std::current_exception());  throw; }
                } else { exportedF.get();  sample.peerAttach(peerHandle, rank, ranks); }
            }
            sample.config.frames_in_flight = tileSplit ? 1 : 3;
            if (tileSplit) { sample.config.tile_rank = rank;  sample.config.tile_count = ranks; }
            else sample.config.accum_mode = B200PT_ACCUM_SUM;
            sample.clearAccum();
            sample.synchronize();
            // every rank has finished its set-up before any rank starts a frame: set-up calls synchronise the whole device (cudaFree, copies on the
            // legacy stream), which ranks sharing one GPU (--devices) must not do while another rank's stream already waits for their bands
            ready.fetch_add(1);
            while (ready.load() < ranks && !failed.load()) std::this_thread::yield();
            const auto t0 = std::chrono::steady_clock::now();
            if (tileSplit) {
                for (int f = 0; f < frames; ++f) {                                  // every rank renders its bands of the SAME sub-frame
                    sample.launchParams.frame.subframe_index = f;
                    sample.setCamera(camera);
                    sample.renderAsync();
                    if (!oneSided) { sample.gatherBands(0, true, true);  sample.synchronize(); }
                    else if (rank == 0 && ranks > 1) { sample.peerAcquire();  sample.downloadPixels(px.data());  sample.peerRelease(); }
                }
                sample.synchronize();
                if (oneSided && ranks > 1 && sample.peerState() != 0) throw std::runtime_error("one-sided tile split: a wait timed out (b200pt_peer_status)");
            } else {
                const int total = (frames / ranks) * ranks;
                for (int f = rank; f < total; f += ranks) {                         // rank r renders sub-frames r, r+N, ...
                    sample.launchParams.frame.subframe_index = f;
                    sample.setCamera(camera);
                    sample.renderAsync();
                }
                sample.reduceSum(total, 0, false);
                sample.synchronize();
            }
            seconds[rank] = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
            if (rank == 0) { sample.downloadPixels(px.data());  samples0 = sample.counters().samples; }
            if (!oneSided) sample.commDestroy();
        } catch (const std::exception& e) { errors[rank] = e.what();  failed.store(true); }
    };
    std::vector<std::thread> threads;
    for (int r = 0; r < ranks; ++r) threads.emplace_back(worker, r);
    for (auto& t : threads) t.join();
    for (int r = 0; r < ranks; ++r) if (!errors[r].empty()) { fprintf(stderr, "rank %d: %s\n", r, errors[r].c_str());  return 1; }
    double slowest = 0.0;
    for (double s : seconds) slowest = std::max(slowest, s);
    const int done = tileSplit ? frames : (frames / ranks) * ranks;
    printf("%d ranks, %s split: %d frames in %.3f s (slowest rank, first-use set-up included): %.2f frames/s; rank 0's last frame: %llu samples\n", ranks,
           oneSided ? "tile (one-sided)" : tileSplit ? "tile" : "sample", done, slowest, done / slowest, samples0);
    unsigned long long sum = 0;
    for (uint32_t p : px) sum += p & 0xffffffu;
    printf("pixel checksum %llu\n", sum);
    if (!out.empty()) {
        std::vector<unsigned char> rgb((size_t)W * H * 3);
        for (int y = 0; y < H; ++y)
            for (int x = 0; x < W; ++x) {
                const uint32_t p = px[(size_t)(H - 1 - y) * W + x];
                unsigned char* q = &rgb[((size_t)y * W + x) * 3];
                q[0] = (unsigned char)p;  q[1] = (unsigned char)(p >> 8);  q[2] = (unsigned char)(p >> 16);
            }
        writePNG(out, W, H, rgb);
    }
    return 0;
}

int main(int argc, char** argv) {
    std::string obj, preset = "sv4_fov", cam = "cornell", out, log, dump, hdr, dumpProbe;
    int W = 1280, H = 720, frames = 16, boxes = 64, spl = 1, inFlight = 1, ranks = 1;
    std::string split = "sample";
    bool splitGiven = false;
    int devices = 0;
    float3 sky{1.f, 1.f, 1.f};
    bool noRender = false, gpuCdf = false, whitted = false, async = false;
    for (int i = 1; i < argc; ++i) {
        const std::string a = argv[i];
        auto next = [&]() -> const char* { if (i + 1 >= argc) { fprintf(stderr, "missing value for %s\n", a.c_str());  exit(2); } return argv[++i]; };
        if (a == "--obj") obj = next();
        else if (a == "--boxes") boxes = atoi(next());
        else if (a == "--preset") preset = next();
        else if (a == "--camera") cam = next();
        else if (a == "--size") { if (sscanf(next(), "%dx%d", &W, &H) != 2) { fprintf(stderr, "--size WxH\n");  return 2; } }
        else if (a == "--frames") frames = atoi(next());
        else if (a == "--spl") spl = atoi(next());
        else if (a == "--ranks") ranks = atoi(next());             // one thread + one renderer per GPU, collectives in the library (multiMain)
        else if (a == "--devices") devices = atoi(next());        // ranks share `devices` GPUs (rank % devices); default: one GPU per rank
        else if (a == "--split") { split = next();  splitGiven = true; }   // sample | tile | peer (given explicitly, also runs with --ranks 1: a one-rank communicator)
        else if (a == "--in-flight") inFlight = atoi(next());      // b200pt_config.frames_in_flight: render() starts the predicted next sub-frames before it blocks
        else if (a == "--async") async = true;                     // with --in-flight N: submit frames through renderAsync() and synchronize once at the end
        else if (a == "--sky") { if (sscanf(next(), "%f,%f,%f", &sky.x, &sky.y, &sky.z) != 3) { fprintf(stderr, "--sky r,g,b\n");  return 2; } }
        else if (a == "--out") out = next();
        else if (a == "--log") log = next();
        else if (a == "--dump-model") dump = next();
        else if (a == "--hdr") hdr = next();
        else if (a == "--gpu-cdf") gpuCdf = true;
        else if (a == "--whitted") whitted = true;
        else if (a == "--dump-probe") dumpProbe = next();
        else if (a == "--no-render") noRender = true;
        else { fprintf(stderr, "unknown argument %s\n", a.c_str());  return 2; }
    }
    try {
        std::unique_ptr<Model> model(obj.empty() ? boxScene(boxes, 0x5702A) : loadOBJ(obj));             // main.cpp:188
        printf("model: %zu meshes, %zu textures, %zu triangles\n", model->meshes.size(), model->textures.size(), model->numTriangles());
        if (!dump.empty()) dumpModel(model.get(), dump.c_str());
        ProbeData probe;
        if (!hdr.empty()) {                                                                                 // main.cpp:201-202
            if (gpuCdf) probe.data = readHDR(hdr, probe.width, probe.height);                               // tables built by setProbeImage
            else loadProbe(probe, hdr);
        } else if (!noRender || !dumpProbe.empty()) loadColor(probe, sky, 512, 256);                        // main.cpp:205-206
        if (!dumpProbe.empty()) {                                  // flat binary for the reader parity test (tests/test_cpp_host.py)
            FILE* fp = fopen(dumpProbe.c_str(), "wb");
            if (!fp) throw std::runtime_error("cannot write " + dumpProbe);
            const int wh[2] = {probe.width, probe.height};
            fwrite(wh, 4, 2, fp);
            fwrite(probe.data, 16, (size_t)probe.width * probe.height, fp);
            fclose(fp);
        }
        if (noRender) return 0;
        sutil::Camera camera = cameraPreset(cam);                                                           // :213-217
        if (ranks > 1 || splitGiven) {
            if (split != "sample" && split != "tile" && split != "peer") throw std::runtime_error("--split sample|tile|peer");
            if (gpuCdf || !probe.valid) throw std::runtime_error("--ranks needs a host-built probe (no --gpu-cdf)");
            const int rc = multiMain(model.get(), probe, camera, preset, W, H, frames, spl, ranks, split == "tile" ? 1 : split == "peer" ? 2 : 0, devices, out);
            delete[] probe.data;  delete[] probe.pdfValuesX;  delete[] probe.cdfValuesX;  delete[] probe.pdfValuesY;  delete[] probe.cdfValuesY;
            return rc;
        }
        SampleRenderer sample(model.get(), 0, preset.c_str());                                              // :252
        sample.setCamera(camera);                                                                           // :253 (before resize, like the reference)
        sample.resize(int2{W, H});                                                                          // :256-262
        if (gpuCdf) sample.setProbeImage(probe);  else sample.setProbe(probe);                              // :269
        sample.launchParams.samples_per_launch = spl;       // read by the `orig` preset (orig/main.cpp sets it); the sv presets write their own
        const b200pt_bvh_info bi = sample.bvhInfo();
        printf("bvh: %llu nodes, %llu leaves, SAH %.2f (LBVH %.2f), build %.2f ms\n", (unsigned long long)bi.n_nodes,
               (unsigned long long)bi.n_leaves, bi.sah_cost, bi.sah_cost_lbvh, bi.build_ms);
        FILE* flog = log.empty() ? nullptr : fopen(log.c_str(), "a");
        auto t0 = std::chrono::steady_clock::now(), tWindow = t0;
        int windowFrames = 0;
        unsigned long long samples = 0, rays = 0;
        for (int f = 0; f < frames; ++f) {                                                                  // :308-394
            const float ph = 0.37f * f;
            sample.launchParams.frame.c = uint2{(unsigned)(W * (0.5f + 0.35f * std::sin(2.f * ph))),  // gaze = cursor, :310-314
                                                     (unsigned)(H * (0.5f + 0.35f * std::sin(3.f * ph + 0.5f)))};
            if (f == 0) sample.launchParams.frame.subframe_index = 0;                                       // camera_changed starts true, :319-320
            sample.setCamera(camera);                                                                       // handleCameraUpdate, :328
            if (whitted) sample.renderWhitted(camera);
            else if (async) { sample.config.frames_in_flight = inFlight;  sample.renderAsync();  if (f + 1 == frames) sample.synchronize();  else continue; }
            else { sample.config.frames_in_flight = inFlight;  sample.render(); }                           // :335
            const b200pt_counters c = sample.counters();
            const unsigned long long mult = (!whitted && async) ? (unsigned long long)frames : 1ull;            // asynchronous: counters of the last frame stand for all
            samples += mult * c.samples;  rays += mult * (c.radiance_rays + c.shadow_rays);
            ++windowFrames;
            const auto now = std::chrono::steady_clock::now();
            const double dt = std::chrono::duration<double>(now - tWindow).count();
            if (dt > 0.5) {                                                                                 // displayStats, sutil.cpp:741-772
                if (flog) fprintf(flog, "%f\n", windowFrames / dt);
                tWindow = now;  windowFrames = 0;
            }
        }
        const double total = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        printf("%d frames in %.3f s: %.2f fps, %.1f Msamples/s, %.1f Mrays/s\n", frames, total, frames / total, samples / total / 1e6, rays / total / 1e6);
        if (flog) { fprintf(flog, "%f\n", frames / total);  fclose(flog); }
        std::vector<uint32_t> px((size_t)W * H);
        sample.downloadPixels(px.data());
        unsigned long long sum = 0;
        for (uint32_t p : px) sum += p & 0xffffffu;
        printf("pixel checksum %llu\n", sum);
        if (!out.empty()) {
            std::vector<unsigned char> rgb((size_t)W * H * 3);                                              // image row 0 is the bottom row
            for (int y = 0; y < H; ++y)
                for (int x = 0; x < W; ++x) {
                    const uint32_t p = px[(size_t)(H - 1 - y) * W + x];
                    unsigned char* q = &rgb[((size_t)y * W + x) * 3];
                    q[0] = (unsigned char)p;  q[1] = (unsigned char)(p >> 8);  q[2] = (unsigned char)(p >> 16);
                }
            if (out.size() > 4 && out.compare(out.size() - 4, 4, ".png") == 0) writePNG(out, W, H, rgb);   // sutil::saveImage, sutil.cpp:723-739
            else {
                FILE* fo = fopen(out.c_str(), "wb");
                if (!fo) throw std::runtime_error("cannot write " + out);
                fprintf(fo, "P6\n%d %d\n255\n", W, H);
                fwrite(rgb.data(), 1, rgb.size(), fo);
                fclose(fo);
            }
        }
        delete[] probe.data;  delete[] probe.pdfValuesX;  delete[] probe.cdfValuesX;  delete[] probe.pdfValuesY;  delete[] probe.cdfValuesY;
    } catch (const std::runtime_error& e) {                                                                // main.cpp:422-425
        fprintf(stderr, "Caught exception: %s\n", e.what());
        return 1;
    }
    return 0;
}
