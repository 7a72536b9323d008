#!/usr/bin/env python
"""bench.py — the reference's headline workload on B200 (BASELINE.json configs[1]):
Sponza-class scene (~262 k triangles, procedural), 1920x1080, 1 spp per frame accumulating, 8 bounces
(`orig` preset: depth 8, tmin 0.001), Disney BSDF + environment NEE.  One step = one frame = one pass of the hot
path (SampleRenderer::render()) over W*H camera samples.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

value   = whole-job Msamples/s with everything resident in HBM: K frames submitted back to back through
          b200pt_render_async with --in-flight frames in flight, one CUDA-event pair around the batch on the renderer's
          primary stream, max over ranks (value_one_at_a_time: one frame at a time, L2 flushed between frames).
e2e     = the same metric through the reference-facing call sequence with HOST buffers: poke launchParams,
          render() (blocking), downloadPixels() into host memory — every step.
roofline= the dominant kernel (traceKernel, one launch per bounce): algorithmic bytes (64 B/4-wide node + 48 B/triangle + 32 B ray +
          16 B hit, counted by the kernel itself in a counters replay of the same frames) / its CUDA-event time.
N > 1   = sample split: rank r renders sub-frames r, r+N, ... into an fp32 SUM buffer, one NCCL
          reduce to rank 0 at the end + divide/tone-map there (b200pt_reduce_sum: the library's own collective on its
          ncclMemAlloc'd accumulation buffer; torch.distributed only carries the 128-byte NCCL id, barriers and the
          max-over-ranks of the timings).  Weak scaling: one frame per rank per step.
tile_split = ONE frame split by 32-row bands over the N ranks and gathered to rank 0 (b200pt_gather_bands): frame
          latency, strong scaling, checked bit-identical against the single-GPU frame.
c5      = BASELINE configs[4]: 20 M triangles, 3840x2160, 256 spp = spl 8 x 32 passes split over the ranks, one reduce.
--impl reference = the CPU oracle (reference shading headers host-compiled when oracle/_ref exists, else the
          port) on all host cores, on a bounded sample of the same frame; rank 0 only.
"""
import argparse
import copy
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

W, H = 1920, 1080
SCENE_TRIS = 262144
PRESET = "orig"
SPL = 1


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=64)          # configs[1]: 1 spp/frame x 64 frames
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--width", type=int, default=W)
    ap.add_argument("--height", type=int, default=H)
    ap.add_argument("--tris", type=int, default=SCENE_TRIS)
    ap.add_argument("--spl", type=int, default=SPL)
    ap.add_argument("--preset", default=PRESET)
    ap.add_argument("--scene", default="sponza", choices=["sponza", "lost_empire", "city", "cornell"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--in-flight", type=int, default=3, help="frames in flight in the timed region (b200pt_config.frames_in_flight, 1..4)")
    ap.add_argument("--no-c5", action="store_true", help="skip the 20 M-triangle 4K 256-spp object (BASELINE configs[4])")
    ap.add_argument("--c5-tris", type=int, default=20000000)
    ap.add_argument("--c5-passes", type=int, default=32, help="passes of 8 spp in total (32 = 256 spp), split over the ranks")
    ap.add_argument("--no-tile-split", action="store_true")
    return ap.parse_args()


def build_workload(a):
    from optixpathtracer_b200 import analytic_sky, preset, preset_camera, scenes
    if a.scene == "sponza":
        model, cam = scenes.sponza_class(a.tris), preset_camera("sponza")
    elif a.scene == "lost_empire":
        model, cam = scenes.lost_empire_class(a.tris), preset_camera("lost_empire")
    elif a.scene == "city":
        model, cam = scenes.city_class(a.tris), preset_camera("sponza")
    else:
        model, cam = scenes.cornell(), preset_camera("cornell")
    probe = analytic_sky(2048, 1024)
    cfg = preset(a.preset, uniform_spl=a.spl)
    cfg.accumulate = 1
    return model, cam, probe, cfg


def reference_optix_status():
    """north_star asks for the reference's OptiX-on-B200 time beside ours.  The reference cannot be built without the NVIDIA OptiX SDK 7.5
    headers (un-vendored: CMake/FindOptiX.cmake:32-76 expects them next to the source tree), so one look for optix.h decides."""
    for root in ("/usr/include", "/usr/local/include", "/usr/local/cuda/include", "/opt"):
        for sub in ("", "optix", "optix/include", "OptiX/include", "nvidia/optix/include"):
            if os.path.isfile(os.path.join(root, sub, "optix.h")):
                return {"available": False, "why": f"optix.h found under {os.path.join(root, sub)} but the reference's OptiX build is not wired into this bench"}
    return {"available": False, "why": "unavailable in this environment: no optix.h on the box (the OptiX SDK 7.5 headers are an un-vendored dependency of the reference)"}


def config_json(a, model, extra=None):
    c = {"workload": f"{a.scene}-class procedural scene, {model.num_triangles()} triangles, {a.width}x{a.height}, "
                     f"{a.spl} spp/frame progressive, preset {a.preset} (max depth {8 if a.preset == 'orig' else 4}), Disney BSDF + probe NEE",
         "triangles": model.num_triangles(), "width": a.width, "height": a.height, "spp_per_step": a.spl, "preset": a.preset,
         "probe": "analytic sun+sky 2048x1024",
         "l2": "inputs larger than L2: every frame streams its own path pool + ray queues (92 B + 112 B per sample slot, "
               f"~{a.width * a.height * a.spl * 204 / 1e6:.0f} MB per frame against the 126 MB L2) and consecutive frames run on different pools; the "
               "one-at-a-time figure (value_one_at_a_time) and the stage timings are taken with a 256 MiB L2 flush before every frame"}
    if extra:
        c.update(extra)
    return c


class ClockSampler:
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.path = tempfile.mktemp(suffix=".csv")
        self.proc = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "20", "-i", str(index)],
                                         stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if not self.proc:
            return out
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for line in open(self.path):
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if sm:
            out = {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons), "samples": len(sm)}
        try:
            os.unlink(self.path)
        except OSError:
            pass
        return out


def cpu_baseline(a, model, cam, probe, cfg, seconds, want_ref=True):
    """The oracle on the host cores, bounded sample: a band of launch rows of the same frame (sub-frame 0)."""
    from optixpathtracer_b200 import host_params
    from oracle import binding
    kind = "reference" if (want_ref and os.path.exists(binding.REF_SO)) else "port"
    orc = binding.Oracle(kind)
    try:                                               # all host threads this process may use; torchrun exports OMP_NUM_THREADS=1, which
        avail = len(os.sched_getaffinity(0))           # would silently turn the CPU arm into a single-thread run
    except AttributeError:
        avail = os.cpu_count() or 1
    orc.set_threads(avail)
    cores = orc.max_threads()
    sc = orc.scene(model, probe)
    p = host_params(a.width, a.height, cam, samples_per_launch=a.spl)
    mid = a.height // 2
    rows = max(cores, 16)
    _, cnt, sec = sc.render_rows(p, cfg, mid - rows // 2, mid - rows // 2 + rows)          # calibration
    rate = cnt["samples"] / max(sec, 1e-6)
    rows2 = int(min(a.height, max(rows, seconds * rate / (a.width * a.spl))))
    r0 = max(0, mid - rows2 // 2)
    _, cnt, sec = sc.render_rows(p, cfg, r0, r0 + rows2)
    return {"value": cnt["samples"] / sec / 1e6, "unit": "Msamples/s", "cores": cores, "kind": kind,
            "sample": f"launch rows [{r0},{r0 + rows2}) of the {a.width}x{a.height} frame at {a.spl} spp, sub-frame 0 ({cnt['samples']} samples, "
                      f"{cnt['radiance_rays'] + cnt['shadow_rays']} rays, {sec:.1f} s)",
            "mrays_per_s": (cnt["radiance_rays"] + cnt["shadow_rays"]) / sec / 1e6, "seconds": sec, "samples": cnt["samples"]}


def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    model, cam, probe, cfg = build_workload(a)
    per_step = max(2.0, min(20.0, 150.0 / max(1, a.steps + a.warmup)))
    vals, last = [], None
    for i in range(a.warmup + a.steps):
        last = cpu_baseline(a, model, cam, probe, cfg, per_step)
        if i >= a.warmup:
            vals.append((last["samples"], last["seconds"]))
    tot_s, tot_t = sum(v[0] for v in vals), sum(v[1] for v in vals)
    value = tot_s / tot_t / 1e6
    line = {"impl": "reference", "metric": "Msamples/s", "value": value, "unit": "Msamples/s", "n_gpus": a.gpus, "steps": a.steps,
            "warmup": a.warmup, "ms_per_step": 1e3 * tot_t / max(1, a.steps), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config_json(a, model),
            "cpu_baseline": {"value": value, "unit": "Msamples/s", "cores": last["cores"], "kind": last["kind"], "sample": last["sample"]},
            "e2e": {"value": value, "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    emit(line)


# ------------------------------------------------------------------------------------------ the B200 arm
def _peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        p = json.load(open(path))
        return float(p["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (burst copy bandwidth)"
    except Exception:
        return 7700.0, "B200_PROFILING.md fallback (nominal HBM3e ~7.7 TB/s); MEASURED_PEAKS.json absent"


MULTI_E2E_SPECULATE = os.environ.get("B200PT_MULTI_E2E", "spec") == "spec"


def _share_unique_id(torch, dist, rank):
    """The 128-byte NCCL id of the library's own communicator travels over torch.distributed (plumbing only)."""
    from optixpathtracer_b200 import SampleRenderer
    uid = torch.from_numpy(SampleRenderer.commUniqueId().copy()) if rank == 0 else torch.zeros(128, dtype=torch.uint8)
    uid = uid.cuda()
    dist.broadcast(uid, src=0)
    return uid.cpu().numpy()


def _max_over_ranks(torch, dist, world, x):
    if world == 1:
        return float(x)
    t = torch.tensor([x], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def _measure_l2_peak(torch):
    """Copy bandwidth (read + write bytes) of a working set that stays in the 126 MB L2: 2 x 24 MiB, best of 5 batches of 40 copies.  The
    denominator of roofline.l2_frac; the HBM figure comes from MEASURED_PEAKS.json."""
    n = 24 << 20
    a_, b_ = torch.empty(n, dtype=torch.uint8, device="cuda"), torch.empty(n, dtype=torch.uint8, device="cuda")
    for _ in range(10):
        b_.copy_(a_)
    best = 0.0
    for _ in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(40):
            b_.copy_(a_)
        e1.record()
        torch.cuda.synchronize()
        best = max(best, 40 * 2 * n / (e0.elapsed_time(e1) * 1e-3) / 1e9)
    return best


def _ncu_summary():
    """Per-kernel figures of the committed `ncu --set full` capture (profiles/traffic.json, written by tools/summarise_profile.py)."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            return json.load(f)
    except (OSError, ValueError):
        return {}


def _tile_split_section(a, torch, dist, world, rank, local, r, cam, stream):
    """ONE frame of the bench workload split by 32-row bands over the ranks (cfg.tile_rank / tile_count: same seeds per launch index as the
    unsplit frame), gathered to rank 0 with the library's grouped ncclSend / ncclRecv: device time of render + gather per frame (max over ranks),
    the same with downloadPixels() on rank 0 inside a wall clock, and a bit-identity check against the unsplit frame rendered on rank 0."""
    cfg = r.config
    keep = (cfg.accum_mode, cfg.tile_rank, cfg.tile_count, cfg.frames_in_flight, cfg.intra_frame_split, cfg.profile)
    cfg.accum_mode, cfg.tile_rank, cfg.tile_count, cfg.frames_in_flight, cfg.intra_frame_split, cfg.profile = 0, rank, world, 1, 0, 0
    host_px = torch.empty((a.height, a.width), dtype=torch.int32, pin_memory=True).numpy().view(np.uint32)
    frames = 16

    def one(i, download):
        r.launchParams.frame.subframe_index = i
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        e0.record(stream)
        r.render(sync=False)
        if world > 1:
            r.gatherBands(root=0, frame=True, accum=False)
        else:
            r.join()
        e1.record(stream)
        if download and rank == 0:
            r.downloadPixels(host_px)
        else:
            r.synchronize()
        wall = time.perf_counter() - t0
        torch.cuda.synchronize()
        return e0.elapsed_time(e1), wall * 1e3

    r.clearAccum()
    for i in range(3):
        one(i, True)
    dev_ms = sum(one(i, False)[0] for i in range(frames)) / frames
    wall_ms = sum(one(i, True)[1] for i in range(frames)) / frames
    dev_ms = _max_over_ranks(torch, dist, world, dev_ms)
    wall_ms = _max_over_ranks(torch, dist, world, wall_ms)
    # bit identity: sub-frame 0 split + gathered (frame and accumulation buffer) against the unsplit frame on rank 0
    r.clearAccum()
    r.launchParams.frame.subframe_index = 0
    r.render(sync=False)
    if world > 1:
        r.gatherBands(root=0, frame=True, accum=True)
    r.synchronize()
    identical = None
    if rank == 0:
        got_px, got_acc = r.downloadPixels().copy(), r.downloadAccum().copy()
        cfg.tile_rank, cfg.tile_count = 0, 1
        r.clearAccum()
        r.launchParams.frame.subframe_index = 0
        r.render()
        identical = bool(np.array_equal(got_px, r.downloadPixels()) and np.array_equal(got_acc.view(np.uint32), r.downloadAccum().view(np.uint32)))
    if world > 1:
        dist.barrier()
    one_sided = None
    if world > 1:
        # The same frames with the one-sided exchange (b200pt_peer_*): rank 0 exports its frame buffer, the other ranks map it (CUDA IPC; the 192-byte
        # handle travels over torch.distributed) and their finalize passes write their bands straight into it over NVLink; rank 0 acquires /
        # downloads / releases.  No collective and no staging copy in the frame.
        cfg.tile_rank, cfg.tile_count = rank, world
        why = ""
        h = torch.zeros(192, dtype=torch.uint8, device="cuda")
        try:
            if rank == 0:
                h = torch.from_numpy(r.peerExport(0, world).copy()).cuda()
        except Exception as e:                                                   # (every rank still takes part in the broadcast and the vote below)
            why = str(e)
        dist.broadcast(h, src=0)
        try:
            if rank != 0:
                r.peerAttach(h.cpu().numpy(), rank, world)
        except Exception as e:
            why = str(e)
        mapped = torch.tensor([0 if why else 1], device="cuda")
        dist.all_reduce(mapped, op=dist.ReduceOp.MIN)
    if world > 1 and int(mapped.item()) == 0:
        # e.g. CUDA IPC closed between the ranks' containers: the gather above is the tile split's figure then
        r.peerDetach()
        one_sided = {"unavailable": why or "another rank could not map rank 0's frame buffer (b200pt_peer_attach)"}
    elif world > 1:

        def one_p2p(i, download):
            r.launchParams.frame.subframe_index = i
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            dist.barrier()
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            e0.record(stream)
            r.render(sync=False)
            if rank == 0:
                r.peerAcquire()
            else:
                r.join()
            e1.record(stream)
            if download and rank == 0:
                r.downloadPixels(host_px)
            else:
                r.synchronize()
            wall = time.perf_counter() - t0
            if rank == 0:
                r.peerRelease()
            torch.cuda.synchronize()
            return e0.elapsed_time(e1), wall * 1e3

        r.clearAccum()
        for i in range(3):
            one_p2p(i, True)
        p_dev = _max_over_ranks(torch, dist, world, sum(one_p2p(i, False)[0] for i in range(frames)) / frames)
        p_wall = _max_over_ranks(torch, dist, world, sum(one_p2p(i, True)[1] for i in range(frames)) / frames)
        r.clearAccum()
        one_p2p(0, True)
        same = None
        if rank == 0:
            got_px = r.downloadPixels().copy()
            cfg.tile_rank, cfg.tile_count = 0, 1
            r.clearAccum()
            r.launchParams.frame.subframe_index = 0
            r.render()
            same = bool(np.array_equal(got_px, r.downloadPixels()))
        st = r.peerStatus()
        bad = _max_over_ranks(torch, dist, world, float(st["state"]))
        dist.barrier()
        r.peerDetach()
        one_sided = {"frame_ms": p_dev, "frames_per_s": 1e3 / p_dev, "with_download_ms": p_wall, "bit_identical_to_single_gpu": same, "timeouts": int(bad),
                     "note": "the same frames with b200pt_peer_export / attach / acquire / release instead of the gather: every other rank's finalizeKernel "
                             "stores its bands straight into rank 0's frame buffer over NVLink (CUDA IPC mapping), a release-store publishes the frame's "
                             "epoch, rank 0's stream waits on the epochs with a one-CTA kernel; no NCCL call and no staging copy inside the frame"}
    cfg.accum_mode, cfg.tile_rank, cfg.tile_count, cfg.frames_in_flight, cfg.intra_frame_split, cfg.profile = keep
    return {"frame_ms": dev_ms, "frames_per_s": 1e3 / dev_ms, "value": a.width * a.height * a.spl / dev_ms / 1e3, "unit": "Msamples/s",
            "with_download_ms": wall_ms, "scaling": "strong", "n_gpus": world, "bit_identical_to_single_gpu": identical, "one_sided": one_sided,
            "gather_bytes_to_root": int(4 * a.width * a.height * (world - 1) / world) if world > 1 else 0,
            "note": f"one {a.width}x{a.height} frame per step, 32-row bands interleaved over {world} rank(s) (band b on rank b % N), "
                    "b200pt_render_async + b200pt_gather_bands (RGBA8 bands to rank 0, grouped ncclSend/ncclRecv on the primary stream), CUDA events "
                    f"around both, max over ranks, mean of {frames} frames one at a time; with_download_ms adds downloadPixels() on rank 0 (wall clock)"}


def _c5_section(a, torch, dist, world, rank, local, peak):
    """BASELINE configs[4]: procedural 20 M-triangle scene, 3840x2160, 256 spp = 8 spp per launch x 32 passes, depth 8, the passes split over the
    ranks (rank r renders sub-frames r, r+N, ...), ONE fp32 reduce of the 132.7 MB SUM buffer + resolve on rank 0.  Everything between the two
    CUDA events is on the renderer's primary stream; max over ranks.  Also this scene's traversal roofline: its BVH (~2 GB) does not fit the L2."""
    from optixpathtracer_b200 import SampleRenderer, analytic_sky, preset, preset_camera, scenes
    from optixpathtracer_b200.structs import ACCUM_SUM
    W4, H4, SPL = 3840, 2160, 8
    passes = max(world, (a.c5_passes // world) * world)
    mine = passes // world
    t0 = time.time()
    model = scenes.city_class(a.c5_tris)
    cfg = preset("orig", uniform_spl=SPL)
    cfg.accumulate, cfg.accum_mode, cfg.frames_in_flight, cfg.profile = 1, ACCUM_SUM, a.in_flight, 0
    r = SampleRenderer(model, device=local, config=cfg)
    r.resize((W4, H4))
    r.setCamera(preset_camera("sponza"))
    r.setProbe(analytic_sky(2048, 1024))
    if world > 1:
        r.commInit(world, rank, _share_unique_id(torch, dist, rank))
    setup_s = time.time() - t0
    info = r.bvhInfo()
    _, _, sp = r.deviceBuffers()
    stream = torch.cuda.ExternalStream(sp, device=local)

    def finish():
        if world > 1:
            r.reduceSum(passes, root=0)
        else:
            r.resolveSum(passes)

    r.launchParams.frame.subframe_index = rank                              # warm-up: one pass + the collective at this size
    r.render(sync=False)
    finish()
    r.synchronize()
    r.clearAccum()
    r.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
    e0.record(stream)
    for i in range(mine):
        r.launchParams.frame.subframe_index = i * world + rank
        r.render(sync=False)
    r.join()
    e1.record(stream)
    finish()
    e2.record(stream)
    r.synchronize()
    torch.cuda.synchronize()
    ms = _max_over_ranks(torch, dist, world, e0.elapsed_time(e2))
    reduce_ms = _max_over_ranks(torch, dist, world, e1.elapsed_time(e2))
    finite = bool(np.isfinite(r.downloadAccum()).all()) if rank == 0 else None
    # traversal roofline of this scene: stage timings (profile 1) and a counters replay (profile 2) of one pass
    cfg.accum_mode, cfg.frames_in_flight = 0, 1
    out = {}
    if rank == 0:
        cfg.profile = 1
        r.launchParams.frame.subframe_index = 0
        r.render()
        c1 = r.counters()
        cfg.profile = 2
        r.launchParams.frame.subframe_index = 0
        r.render()
        c2 = r.counters()
        cfg.profile = 0
        trace_ms = c1.ms_trace_closest + c1.ms_trace_any
        alg = 64 * (c2.nodes_fetched + c2.nodes_fetched_shadow) + 48 * (c2.tris_fetched + c2.tris_fetched_shadow) + 48 * c2.radiance_rays + 36 * c2.shadow_rays
        ncu = _ncu_summary().get("c5_traceKernel", {})
        gbs = alg / (trace_ms * 1e-3) / 1e9
        out["roofline"] = {"kernel": "traceKernel", "bound": "l1/l2 gather + issue (the rays of a frame touch a cache-sized part of the tree: see dram_frac)",
                           "achieved": gbs, "peak": peak, "unit": "GB/s", "frac": gbs / peak,
                           "traffic": ncu.get("dram_bytes_per_launch"), "dram_frac": ncu.get("dram_frac_of_peak"), "l2_hit_pct": ncu.get("l2_hit_pct"),
                           "traffic_source": ncu.get("source"), "nodes_per_ray": c2.nodes_fetched / max(1, c2.radiance_rays),
                           "tris_per_ray": c2.tris_fetched / max(1, c2.radiance_rays), "trace_ms_per_pass": trace_ms, "shade_ms_per_pass": c1.ms_shade,
                           "bvh_bytes": int(info.node_bytes + info.tri_bytes)}
    r.close()
    del model
    if world > 1:
        dist.barrier()
    samples = float(W4) * H4 * SPL * passes
    out.update({"workload": f"city-class procedural scene, {int(info.n_triangles)} triangles, {W4}x{H4}, {SPL * passes} spp = {SPL} spp/launch x {passes} passes, "
                            "preset orig (max depth 8), Disney BSDF + probe NEE",
                "value": samples / (ms * 1e-3) / 1e6, "unit": "Msamples/s", "n_gpus": world, "seconds": ms * 1e-3, "passes_per_rank": mine,
                "reduce_ms": reduce_ms, "reduce_bytes": 16 * W4 * H4, "scaling": "strong", "finite": finite, "setup_seconds": round(setup_s, 2),
                "bvh_build_ms": round(float(info.build_ms), 1),
                "note": "strong scaling: the 256-spp job is fixed and its passes are split over the ranks; the clock covers every pass of the slowest rank, "
                        "the fp32 SUM reduce (b200pt_reduce_sum on the ncclMemAlloc'd accumulation buffer) and the resolve on rank 0; scene generation, "
                        "upload and the LBVH build (setup_seconds, once per rank, in parallel) are outside it"})
    return out


def run_b200(a):
    import torch
    import torch.distributed as dist
    from optixpathtracer_b200 import SampleRenderer
    from optixpathtracer_b200.structs import ACCUM_SUM

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    if world > 1:                                             # (NCCL's own log is left alone: fd 1 already points at stderr, see main())
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    model, cam, probe, cfg = build_workload(a)
    t0 = time.time()
    r = SampleRenderer(model, device=local, config=cfg)
    ctor_s = time.time() - t0
    r.resize((a.width, a.height))
    r.setCamera(cam)
    r.setProbe(probe)
    info = r.bvhInfo()
    n_px = a.width * a.height
    _, _, stream_ptr = r.deviceBuffers()
    stream = torch.cuda.ExternalStream(stream_ptr, device=local)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")           # > 126 MB L2

    # N > 1: sample split.  The library opens its own NCCL communicator (b200pt_comm_init) and moves the accumulation buffer into
    # ncclMemAlloc'd, registered memory: the buffer the finalize kernels add into IS the send buffer of the reduce.
    comm_info = None
    if world > 1:
        cfg.accum_mode = ACCUM_SUM
        r.commInit(world, rank, _share_unique_id(torch, dist, rank))
        comm_info = r.commInfo()
    cfg.profile = 0                                                              # timed region: no instrumentation (stage events cost ~3 us each between kernels)

    def frame_index(step):                                                      # rank r renders sub-frames r, r+N, ...
        return step * world + rank

    stage = dict(raygen=0.0, trace_closest=0.0, trace_any=0.0, shade=0.0, accumulate=0.0, total=0.0)
    n_closest = n_any = launches = 0
    rays = dict(radiance=0, shadow=0, samples=0)

    def tally(c):
        nonlocal n_closest, n_any, launches
        n_closest += c.n_trace_closest_launches
        n_any += c.n_trace_any_launches
        launches += c.kernel_launches
        rays["radiance"] += c.radiance_rays
        rays["shadow"] += c.shadow_rays
        rays["samples"] += c.samples

    def one_step(step, timed):
        """One frame at a time: L2 flushed, then a CUDA-event pair around the frame (stage timings, the one-at-a-time figure)."""
        with torch.cuda.stream(stream):
            flush.fill_(step & 0xff)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        r.launchParams.frame.subframe_index = frame_index(step)
        e0.record(stream)
        r.render(sync=False)
        r.join()
        e1.record(stream)
        r.synchronize()
        if timed:
            c = r.counters()
            if cfg.profile & 1:
                for k in stage:
                    stage[k] += getattr(c, "ms_" + k)
            tally(c)
        return e0.elapsed_time(e1)

    # ---- the timed region: K frames submitted back to back through b200pt_render_async with `in_flight` frames in flight (each on its own
    # stream / path pool / queue set; they fold into the accumulation buffer in submission order), one CUDA-event pair around the whole
    # batch on the renderer's primary stream.  No L2 flush inside it: every frame streams its own pool + queue state (see config.l2).
    cfg.frames_in_flight = a.in_flight
    for i in range(a.warmup):
        r.launchParams.frame.subframe_index = frame_index(i)
        r.render(sync=False)
    r.synchronize()
    if world > 1:
        for _ in range(2):                                                      # warm-up of the collective itself: NCCL sets up its channels for this
            r.reduceSum(max(1, a.warmup) * world, root=0, keep_local=True)              # message size on first use, which is not a cost of the render path
        r.synchronize()
        r.clearAccum()
        r.synchronize()
        dist.barrier()
    torch.cuda.synchronize()
    sampler = ClockSampler(local) if rank == 0 else None
    wall0 = time.time()
    e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
    e0.record(stream)
    for i in range(a.steps):
        r.launchParams.frame.subframe_index = frame_index(a.warmup + i)
        r.render(sync=False)
    r.join()                                                                    # primary stream ordered after every frame in flight
    e1.record(stream)
    if world > 1:                                                                # one fp32 reduce of the SUM buffer + resolve on rank 0, on the primary stream
        r.reduceSum(a.steps * world, root=0)
    e2.record(stream)
    r.synchronize()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e2)
    reduce_ms = e1.elapsed_time(e2) if world > 1 else 0.0
    if world > 1:
        dist.barrier()
    wall = time.time() - wall0
    clocks = sampler.stop() if sampler else None
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())

    # ---- e2e: the reference-facing call sequence with HOST buffers, every step, one frame at a time.  The same a.steps frames as the timed
    # region, so their ray / sample / launch counters (read after each step, outside the clock) are the timed region's.
    #   N = 1: setCamera + render() (blocking) + downloadPixels() - the reference's sequence (sv4/main.cpp:328-335 + SimplePathtracer.cpp:279-283).
    #   N > 1: the sample split END TO END: every rank setCamera + render (blocking, predicted next sub-frames in flight as at N = 1), then the fp32 reduce of the SUM buffers onto rank 0 with the resolve there
    #          (b200pt_reduce_sum, keep_local: the ranks keep their partial sums and go on accumulating - progressive display) and downloadPixels()
    #          of the combined image on rank 0, all inside the clock; the collective keeps the ranks in lockstep.
    cfg.profile = 0
    host_px = torch.empty((a.height, a.width), dtype=torch.int32, pin_memory=True).numpy().view(np.uint32)   # pinned host frame buffer

    def e2e_step(i, index, camera):
        r.setCamera(camera)                                                     # host -> LaunchParams (312 B, re-read by every render)
        r.launchParams.frame.subframe_index = index
        if world == 1:
            r.render()                                                          # blocking, as the reference's render()
            r.downloadPixels(host_px)                                           # D2H W*H RGBA8
        else:
            if MULTI_E2E_SPECULATE:
                r.render()                                                      # blocking, with the predicted next sub-frames in flight on the other lanes
            else:
                r.render(sync=False)
            r.reduceSum((i + 1) * world, root=0, keep_local=True)
            if rank == 0:
                r.downloadPixels(host_px)
            else:
                r.synchronize()

    def e2e_loop(steps, count, moving=False):
        if world > 1:
            r.clearAccum()
        for i in range(3):                                                      # warm-up on sub-frame indices that do NOT lead into the timed ones: the first
            e2e_step(i, frame_index(100000 + i), cam)                           # timed call is then a wrong guess, the speculative wavefronts of the warm-up are
        r.synchronize()                                                         # dropped, and the timed region starts without work done before its clock
        if world > 1:
            r.clearAccum()
            r.synchronize()
            dist.barrier()
        torch.cuda.synchronize()                                                # every lane drained, dropped wavefronts included
        sec = 0.0
        for i in range(steps):
            camera = cam
            if moving:                                                          # a caller whose launchParams change every frame (gaze / camera follows the cursor,
                camera = copy.copy(cam)                                         # sv4/main.cpp:310-314): nothing can be rendered ahead
                camera.eye = np.asarray((cam.eye[0] + 0.25 * (i + 1), cam.eye[1], cam.eye[2]), np.float32)
            t0 = time.perf_counter()
            e2e_step(i, frame_index(a.warmup + i), camera)
            dt = time.perf_counter() - t0
            sec += dt
            if os.environ.get("B200PT_BENCH_TRACE"):
                print(f"e2e step {i}: {dt * 1e3:.3f} ms", file=sys.stderr)
            if count:
                tally(r.counters())                                             # outside the clock
        return _max_over_ranks(torch, dist, world, sec)

    e2e_serial_steps = min(a.steps, 16)
    e2e_serial_value = e2e_async_value = e2e_moving_value = None
    if world == 1:
        # strictly one frame at a time first (frames_in_flight = 1: no speculative next sub-frame), on a bounded number of steps ...
        cfg.frames_in_flight = 1
        e2e_serial_s = e2e_loop(e2e_serial_steps, False)
        # ... a camera that moves every frame: speculation never hits; the frame is split into two concurrent band sets instead (intra_frame_split)
        cfg.intra_frame_split = 2
        e2e_moving_s = e2e_loop(e2e_serial_steps, False, moving=True)
        cfg.intra_frame_split = 0
        # ... then the same blocking calls with the library allowed to start the predicted next sub-frame while it blocks (frames_in_flight > 1)
        cfg.frames_in_flight = a.in_flight
        e2e_s = e2e_loop(a.steps, True)
    elif MULTI_E2E_SPECULATE:
        cfg.frames_in_flight = a.in_flight
        e2e_s = e2e_loop(a.steps, True)
    else:
        cfg.frames_in_flight, cfg.intra_frame_split = 1, 2                      # every step completes before the next starts: split the frame instead
        e2e_s = e2e_loop(a.steps, True)
        cfg.intra_frame_split = 0
    if world > 1:
        tot = torch.tensor([rays["samples"], rays["radiance"], rays["shadow"], launches], dtype=torch.float64, device="cuda")
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
        all_samples, all_rad, all_sh, all_launches = (float(x) for x in tot.tolist())
    else:
        all_samples, all_rad, all_sh, all_launches = rays["samples"], rays["radiance"], rays["shadow"], launches
    value = all_samples / (ms * 1e-3) / 1e6
    mrays = (all_rad + all_sh) / (ms * 1e-3) / 1e6
    e2e_value = all_samples / e2e_s / 1e6
    if world == 1:
        e2e_serial_value = (all_samples / a.steps) * e2e_serial_steps / e2e_serial_s / 1e6
        e2e_moving_value = (all_samples / a.steps) * e2e_serial_steps / e2e_moving_s / 1e6

        # ---- the same end-to-end step through the asynchronous entry points: setCamera + b200pt_render_async + b200pt_download_pixels_async
        # into a ring of pinned host frames every step, `in_flight` frames in flight, one b200pt_synchronize at the end (inside the clock)
        cfg.frames_in_flight = a.in_flight
        ring = [torch.empty((a.height, a.width), dtype=torch.int32, pin_memory=True).numpy().view(np.uint32) for _ in range(a.in_flight + 1)]
        for i in range(a.in_flight + 1):
            r.launchParams.frame.subframe_index = frame_index(i)
            r.render(sync=False)
            r.downloadPixelsAsync(ring[i])
        r.synchronize()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for i in range(a.steps):
            r.setCamera(cam)
            r.launchParams.frame.subframe_index = frame_index(a.warmup + i)
            r.render(sync=False)
            r.downloadPixelsAsync(ring[i % len(ring)])
        r.synchronize()
        e2e_async_value = all_samples / (time.perf_counter() - t0) / 1e6
    cfg.frames_in_flight = 1

    # ---- one frame at a time, L2 flushed before each (what `value` was before frames could overlap; explains the in-flight gain)
    serial_steps = min(a.steps, 16)
    serial_ms = sum(one_step(a.warmup + i, False) for i in range(serial_steps))
    serial_value = (all_samples / a.steps) * serial_steps / (serial_ms * 1e-3) / 1e6      # whole job: every rank renders its own frame per step

    # ---- per-stage CUDA-event timings: the same frames again with a CUDA-event pair around every kernel of the frame
    cfg.profile = 1
    timed_launches, timed_closest, timed_rays = launches, n_closest, dict(rays)
    n_closest = 0
    stage_steps = min(a.steps, 8)
    stage_ms = 0.0
    for i in range(stage_steps):
        stage_ms += one_step(a.warmup + i, True)
    stage_closest = n_closest
    launches, n_closest, rays = timed_launches, timed_closest, timed_rays

    # ---- rooflines: algorithmic bytes from a counters replay of the same frames (the kernels count what they fetch)
    cfg.profile = 2
    nodes = tris = crays = hits = 0
    nodes_s = tris_s = srays = 0
    replay = min(a.steps, 4)
    fused_finalize = False
    for i in range(replay):
        r.launchParams.frame.subframe_index = frame_index(a.warmup + i)
        r.render()
        c = r.counters()
        nodes += c.nodes_fetched
        tris += c.tris_fetched
        crays += c.radiance_rays
        hits += c.closest_hits
        nodes_s += c.nodes_fetched_shadow
        tris_s += c.tris_fetched_shadow
        srays += c.shadow_rays
        fused_finalize = bool(c.fused_finalize)
    cfg.profile = 0

    peak, peak_src = _peaks()
    l2_peak = _measure_l2_peak(torch) if rank == 0 else 0.0

    def side_object(fn, *args):
        """The additional objects must not cost the contract line: on one GPU a failure inside one of them is reported in its place.  (On several
        ranks an exception is left to torchrun - a rank that skipped a collective would leave the others waiting.)"""
        if world > 1:
            return fn(*args)
        try:
            return fn(*args)
        except Exception as e:                                                   # noqa: BLE001
            return {"error": f"{type(e).__name__}: {e}"}

    tile = None if a.no_tile_split else side_object(_tile_split_section, a, torch, dist, world, rank, local, r, cam, stream)
    r.close()
    c5 = None if a.no_c5 else side_object(_c5_section, a, torch, dist, world, rank, local, peak)

    if rank == 0:
        tj = _ncu_summary()
        kern = tj.get("kernels", {})

        def ncu_of(name):
            k = dict(kern.get(name, {}))
            if not k and name == "traceKernel" and "dram_bytes_per_launch" in tj:      # round-1 layout of traffic.json
                k = {x: tj[x] for x in tj if x not in ("kernels",)}
            out = {x: k[x] for x in ("ipc_per_sm", "active_threads_per_warp_inst", "l1_wavefront_pct", "l1_hit_pct", "l2_hit_pct", "dram_pct_of_peak",
                                     "warps_active_pct", "long_scoreboard_per_issue", "duration_us", "launches_captured") if x in k}
            dram, lts, dur = k.get("dram_bytes_per_launch"), k.get("lts_bytes_per_launch"), k.get("duration_us")
            fr = {"traffic": dram, "dram_frac": None, "l2_bytes_per_launch": lts, "l2_gbs": None, "l2_frac": None, "l2_peak_gbs": l2_peak}
            if dram is not None and dur:
                fr["dram_frac"] = dram / (dur * 1e-6) / 1e9 / peak
            if lts is not None and dur:
                fr["l2_gbs"] = lts / (dur * 1e-6) / 1e9
                fr["l2_frac"] = fr["l2_gbs"] / l2_peak if l2_peak else None
            return out, fr

        src = tj.get("source")
        per_frame = 1.0 / replay
        # (1) traceKernel: one launch per bounce traces the extension rays AND the previous pass's shadow rays
        bytes_closest_per_frame = (64 * nodes + 48 * tris + (32 + 16) * crays) * per_frame
        bytes_any_per_frame = (64 * nodes_s + 48 * tris_s + (32 + 4) * srays) * per_frame
        launches_per_frame = stage_closest / stage_steps
        st = {k: v / stage_steps for k, v in stage.items()}
        trace_ms = st["trace_closest"] + st["trace_any"]
        ms_per_launch = trace_ms / max(1e-9, launches_per_frame)
        bytes_per_launch = (bytes_closest_per_frame + bytes_any_per_frame) / launches_per_frame
        achieved = bytes_per_launch / (ms_per_launch * 1e-3) / 1e9
        t_ncu, t_fr = ncu_of("traceKernel")
        roof_trace = {"bound": "issue rate + L1 wavefronts (BVH resident in L1/L2; not HBM: see dram_frac)", "kernel": "traceKernel",
                      "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, **t_fr, "traffic_source": src,
                      "ncu": t_ncu, "peak_source": peak_src, "algorithmic_gbs": achieved,
                      "algorithmic_bytes_per_launch": bytes_per_launch, "launches_per_step": launches_per_frame, "avg_launch_ms": ms_per_launch,
                      "nodes_per_ray": nodes / max(1, crays), "tris_per_ray": tris / max(1, crays),
                      "nodes_per_shadow_ray": nodes_s / max(1, srays), "tris_per_shadow_ray": tris_s / max(1, srays),
                      "bytes_per_ray": (64 * nodes + 48 * tris) / max(1, crays) + 48,
                      "note": "achieved = ALGORITHMIC bytes / launch time: 64 B x 4-wide node records + 48 B x triangle records fetched (counted by the kernel in a counters "
                              "replay of the same frames) + 32 B ray in + 16 B hit (4 B flag for shadow rays) out, against the measured HBM copy peak (SURVEY 8d's formula). "
                              f"The BVH is {(info.node_bytes + info.tri_bytes) / 1e6:.0f} MB (L2: 126 MB), so these bytes are served by L1/L2: dram_frac = measured DRAM bytes per launch / "
                              "its duration / HBM peak, l2_frac = measured L2 bytes (32 B x lts__t_sectors) / duration / l2_peak_gbs = the read + write bandwidth of an L2-resident 2 x 24 MiB "
                              "copy measured in this run (a conservative denominator: the microarchitecture guide puts the L2 slice cap at ~6300 B/cycle chip-wide); the limiter is the "
                              "issue rate at about half of the lanes active (ncu.active_threads_per_warp_inst, ncu.ipc_per_sm), not bandwidth"}
        # (2) shadeKernel: per queue entry 48 B in (ray o, d, hit); per surface hit the 32-byte path record read + written, the 48-byte triangle, the 136-byte
        # mesh record (material), ~60 B of probe tables (guide entries, ~3+3 CDF probes, texel, two pdfs); out: 32 B per continuation ray, 48 B per shadow ray
        cont = max(0.0, crays - replay * (rays["samples"] / max(1, a.steps)))
        shade_bytes = (48 * crays + (64 + 48 + 136 + 60) * hits + 32 * cont + 48 * srays) * per_frame
        s_ncu, s_fr = ncu_of("shadeKernel")
        shade_gbs = shade_bytes / (st["shade"] * 1e-3) / 1e9
        roof_shade = {"bound": "latency of dependent gathers (ray -> path record -> probe CDF search -> triangle -> material), hbm second", "kernel": "shadeKernel",
                      "achieved": shade_gbs, "peak": peak, "unit": "GB/s", "frac": shade_gbs / peak, **s_fr, "traffic_source": src, "ncu": s_ncu,
                      "algorithmic_bytes_per_step": shade_bytes, "launches_per_step": launches_per_frame - 1, "ms_per_step": st["shade"],
                      "hits_per_sample": hits / max(1.0, replay * rays["samples"] / max(1, a.steps)),
                      "bytes_per_hit": shade_bytes / max(1.0, hits * per_frame),
                      "note": "algorithmic bytes per step = 48 B x queue entries + 308 B x surface hits (64 path record r/w + 48 triangle + 136 mesh/material + ~60 probe) "
                              "+ 32 B x continuation rays + 48 B x shadow rays, / the CUDA-event time of the frame's shade launches; the material records and probe guide "
                              "tables are cache-resident, so DRAM traffic (traffic, dram_frac) is below the algorithmic figure"}
        # (3) gather + finalize: 48 B per sample (direct, indirect, alpha) in, 2 x 32 B per launch index through the per-launch-index sums, 36 B per pixel (accum r/w + RGBA8)
        aov = 1 if cfg.write_aovs else 0                                            # the `orig` preset also writes the normal / albedo / colour AOVs (orig/deviceProgram.cu:472-474)
        rec_bytes = 0 if fused_finalize else 64 + aov * 64                         # per-launch-index records of a separate gather pass (written + read back)
        acc_bytes = (48.0 + 32.0 * aov) * rays["samples"] / max(1, a.steps) + (rec_bytes + 36 + aov * 48) * n_px
        g_ncu, g_fr = ncu_of("gatherKernel")
        f_ncu, f_fr = ncu_of("finalizeKernel")
        acc_gbs = acc_bytes / (st["accumulate"] * 1e-3) / 1e9
        roof_acc = {"bound": "hbm", "kernel": "finalizeKernel (gather fused)" if fused_finalize else "gatherKernel + finalizeKernel", "fused": fused_finalize, "achieved": acc_gbs, "peak": peak, "unit": "GB/s", "frac": acc_gbs / peak,
                    "traffic": (g_fr["traffic"] or 0) + (f_fr["traffic"] or 0) if (g_fr["traffic"] is not None or f_fr["traffic"] is not None) else None,
                    "dram_frac_gather": g_fr["dram_frac"], "dram_frac_finalize": f_fr["dram_frac"], "ncu_gather": g_ncu, "ncu_finalize": f_ncu, "traffic_source": src,
                    "algorithmic_bytes_per_step": acc_bytes, "ms_per_step": st["accumulate"],
                    "note": "48 B per sample (direct, indirect, alpha) + 36 B per pixel (16 B accumulation read, 16 B written, 4 B RGBA8); with AOVs (preset orig) + 32 B per "
                            "sample and + 48 B per pixel; a frame whose samples sit in one chunk of the path pool (fused: true) sums the samples inside finalizeKernel, "
                            "otherwise gatherKernel writes per-launch-index records that finalizeKernel reads back (+ 64 B per launch index, + 64 with AOVs); / the "
                            "CUDA-event time of the launch(es) (tens of microseconds: launch ramp and tail are a visible share of a kernel this short); "
                            "dram_frac_* = measured DRAM bytes / duration / HBM peak of each kernel under ncu"}
        parallelism = "single GPU" if world == 1 else (f"sample split over {world} GPUs (rank r renders sub-frames r, r+N, ...), one NCCL fp32 reduce of the SUM buffer "
                                                        "+ resolve on rank 0 inside the timed region (b200pt_reduce_sum)")
        e2e_note = (f"setCamera + render() (blocking) + downloadPixels() into pinned host memory every step, the reference's call sequence; cfg.frames_in_flight = {a.in_flight}: while "
                    "render() blocks on frame i the library has already started the wavefront of the predicted next call (same launchParams, next subframe_index - the progressive "
                    "accumulation this config is; a wrong guess is dropped, pixels are bit-identical either way); one_at_a_time_value = the same calls with frames_in_flight = 1; "
                    "moving_camera_value = the camera moves every step, so nothing can be rendered ahead (frames_in_flight = 1, intra_frame_split = 2): the figure for interactive "
                    f"use; scene/BVH upload is one-off in the ctor ({ctor_s:.2f} s incl. LBVH build {float(info.build_ms):.2f} ms)") if world == 1 else (
                    f"the sample split end to end, every step, on {world} ranks: setCamera + render() on every rank, b200pt_reduce_sum (fp32 SUM of the {16 * n_px / 1e6:.1f} MB "
                    "accumulation buffers onto rank 0 + resolve there, keep_local: the ranks keep accumulating) and downloadPixels() of the combined image on rank 0, all inside the "
                    "wall clock (max over ranks); " + (f"render() is the blocking call with cfg.frames_in_flight = {a.in_flight}, as at N = 1: while a rank reduces and rank 0 downloads, "
                    "the library already runs the predicted next sub-frames of that rank (index stride = the rank count); B200PT_MULTI_E2E=serial gives the figure with strictly "
                    "one frame at a time per rank" if MULTI_E2E_SPECULATE else "one frame at a time per rank (b200pt_render_async, frames_in_flight = 1, intra_frame_split = 2)"))
        line = {
            "metric": "Msamples/s", "value": value, "unit": "Msamples/s", "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": ms / a.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": config_json(a, model),
            "parallelism": parallelism,
            "bvh": {"nodes": int(info.n_nodes), "leaves": int(info.n_leaves), "sah": round(float(info.sah_cost), 2),
                    "sah_lbvh": round(float(info.sah_cost_lbvh), 2), "build_ms": round(float(info.build_ms), 3),
                    "bytes": int(info.node_bytes + info.tri_bytes)},
            "ctor_seconds": round(ctor_s, 3),
            "comm": comm_info,
            "frames_in_flight": a.in_flight,
            "value_one_at_a_time": serial_value,
            "value_note": f"value = {a.steps} frames submitted back to back through b200pt_render_async with {a.in_flight} frames in flight (b200pt_config.frames_in_flight: "
                          "each frame in flight has its own stream, path pool and queues; one frame's drain phases and thin late bounces overlap the next frame's wide early "
                          "bounces; frames fold into the accumulation buffer in submission order, bit-identical to one at a time), one CUDA-event pair around the batch - "
                          f"the fill and the drain of the {a.in_flight}-lane pipeline are inside the clock, so a short run (e.g. 20 steps) reads a few percent lower than 64; "
                          f"value_one_at_a_time = {serial_steps} of the same frames with a blocking wait and an L2 flush between them",
            "mrays_per_s": mrays, "radiance_rays_per_sample": all_rad / max(1.0, all_samples), "shadow_rays_per_sample": all_sh / max(1.0, all_samples),
            "e2e": {"value": e2e_value, "unit": "Msamples/s", "h2d_bytes_per_step": 312 * world, "d2h_bytes_per_step": 4 * n_px,
                    "one_at_a_time_value": e2e_serial_value, "moving_camera_value": e2e_moving_value,
                    "async_value": e2e_async_value,
                    "async_note": f"the same step through b200pt_render_async + b200pt_download_pixels_async ({a.in_flight} frames in flight, ring of pinned host frames, "
                                  "one synchronize at the end, wall clock): what a caller that adopts the asynchronous entry points gets (N = 1 only)",
                    "note": e2e_note},
            "gpu_launches": int(all_launches),
            "reference_optix": reference_optix_status(),
            "roofline": roof_trace, "roofline_shade": roof_shade, "roofline_accumulate": roof_acc,
            "stage_ms_per_step": st,
            "stage_note": f"per-stage CUDA-event times of {stage_steps} of the timed frames rendered again with an event pair around every kernel; the ~40 extra "
                          "event records per frame cost ~0.1 ms, so the stage total sits slightly above the uninstrumented ms_per_step",
            "reduce_ms": reduce_ms, "wall_seconds_timed_region": wall, "clocks": clocks,
            "tile_split": tile, "c5": c5,
        }
        if not a.no_cpu_baseline and world == 1:
            cb = cpu_baseline(a, model, cam, probe, cfg, a.cpu_seconds)
            line["cpu_baseline"] = {k: cb[k] for k in ("value", "unit", "cores", "kind", "sample", "mrays_per_s")}
        emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


_JSON_OUT = None


def emit(line):
    """The one JSON line goes to the REAL stdout; everything else any library prints to fd 1 (NCCL's version banner, torchrun
    notices) was pointed at stderr by main()."""
    out = _JSON_OUT or sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def main():
    global _JSON_OUT
    a = parse()
    if not (a.impl != "reference" and a.gpus > 1 and "WORLD_SIZE" not in os.environ):
        sys.stdout.flush()
        _JSON_OUT = os.fdopen(os.dup(1), "w")
        os.dup2(2, 1)                                     # fd 1 -> stderr for the rest of the process (C libraries included)
    if a.impl == "reference":
        run_reference(a)
        return
    if a.gpus > 1 and "WORLD_SIZE" not in os.environ:
        # convenience: re-launch under torchrun exactly as the driver would
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={a.gpus}", "--master-addr", "127.0.0.1",
               "--master-port", str(29500 + os.getpid() % 2000), os.path.abspath(__file__)] + sys.argv[1:]
        raise SystemExit(subprocess.call(cmd))
    run_b200(a)


if __name__ == "__main__":
    main()
