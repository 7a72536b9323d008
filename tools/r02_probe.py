#!/usr/bin/env python
"""Round-2 tuning probe (not a bench line): on the bench workload prints
  * frames in flight 1..3 back to back through b200pt_render_async (no host sync),
  * strictly one frame at a time (blocking render) with cfg.intra_frame_split = 1..4: device time per frame (CUDA events, L2 flushed) and the
    wall time of render() + downloadPixels(),
  * the per-stage CUDA-event split of one instrumented frame and the traversal counters,
and checks that every variant leaves the same accumulation buffer (bit for bit).
Run with B200PT_LIB=build/variants/libb200pt_<name>.so to probe a tuning build."""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from optixpathtracer_b200 import SampleRenderer  # noqa: E402


def main():
    a = bench.parse()
    model, cam, probe, cfg = bench.build_workload(a)
    cfg.profile = 0
    r = SampleRenderer(model, device=0, config=cfg)
    r.resize((a.width, a.height))
    r.setCamera(cam)
    r.setProbe(probe)
    _, _, sp = r.deviceBuffers()
    stream = torch.cuda.ExternalStream(sp, device=0)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    host_px = torch.empty((a.height, a.width), dtype=torch.int32, pin_memory=True).numpy().view(np.uint32)
    tag = os.path.basename(os.environ.get("B200PT_LIB", "default"))
    n_samples = a.width * a.height * a.spl

    def reset():
        r.clearAccum()
        r.synchronize()

    def run_async(frames):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for f in range(frames):
            r.launchParams.frame.subframe_index = f
            r.render(sync=False)
        r.join()
        e1.record(stream)
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / frames

    ref = None
    for n in (1, 2, 3):
        cfg.frames_in_flight, cfg.intra_frame_split = n, 0
        reset()
        run_async(4)
        reset()
        ms = run_async(a.steps)
        acc = r.downloadAccum()
        if ref is None:
            ref = acc
            import hashlib
            print(f"[{tag}] accumulation buffer sha1 {hashlib.sha1(acc.tobytes()).hexdigest()[:16]} (hits are tree-independent: equal across builds)", flush=True)
        same = np.array_equal(ref.view(np.uint32), acc.view(np.uint32))
        print(f"[{tag}] in flight {n}: {n_samples / ms / 1e3:.1f} Msamples/s, {ms:.3f} ms/frame, accum identical: {same}", flush=True)

    cfg.frames_in_flight = 1
    for split in (1, 2, 3, 4):
        cfg.intra_frame_split = split
        reset()
        dev_ms, wall = 0.0, 0.0
        for f in range(-3, a.steps):
            with torch.cuda.stream(stream):
                flush.fill_(f & 0xff)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            if f == 0:
                reset()
            r.launchParams.frame.subframe_index = max(f, 0)
            t0 = time.perf_counter()
            e0.record(stream)
            r.render(sync=False)
            r.join()
            e1.record(stream)
            r.synchronize()
            r.downloadPixels(host_px)
            dt = time.perf_counter() - t0
            if f >= 0:
                dev_ms += e0.elapsed_time(e1)
                wall += dt
        acc = r.downloadAccum()
        same = np.array_equal(ref.view(np.uint32), acc.view(np.uint32))
        print(f"[{tag}] one at a time, split {split}: device {dev_ms / a.steps:.3f} ms/frame ({n_samples * a.steps / dev_ms / 1e3:.1f} Msamples/s), "
              f"render+download wall {1e3 * wall / a.steps:.3f} ms ({n_samples * a.steps / wall / 1e6:.1f} Msamples/s), accum identical: {same}", flush=True)

    cfg.intra_frame_split = 0
    cfg.profile = 3
    stage = {}
    nodes = tris = rays = 0
    for f in range(4):
        with torch.cuda.stream(stream):
            flush.fill_(f)
        r.launchParams.frame.subframe_index = f
        r.render()
        c = r.counters()
        for k in ("raygen", "trace_closest", "trace_any", "shade", "accumulate", "total"):
            stage[k] = stage.get(k, 0.0) + getattr(c, "ms_" + k) / 4
        nodes += c.nodes_fetched
        tris += c.tris_fetched
        rays += c.radiance_rays
    print(f"[{tag}] stages (ms/frame, counters build): " + ", ".join(f"{k} {v:.3f}" for k, v in stage.items()) +
          f"; nodes/ray {nodes / max(rays, 1):.2f}, tris/ray {tris / max(rays, 1):.2f}", flush=True)
    cfg.profile = 1
    stage = {}
    for f in range(4):
        with torch.cuda.stream(stream):
            flush.fill_(f)
        r.launchParams.frame.subframe_index = f
        r.render()
        c = r.counters()
        for k in ("raygen", "trace_closest", "trace_any", "shade", "accumulate", "total"):
            stage[k] = stage.get(k, 0.0) + getattr(c, "ms_" + k) / 4
    print(f"[{tag}] stages (ms/frame): " + ", ".join(f"{k} {v:.3f}" for k, v in stage.items()), flush=True)
    r.close()


if __name__ == "__main__":
    main()
