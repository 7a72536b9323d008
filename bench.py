#!/usr/bin/env python
"""bench.py — the reference's headline workload on B200 (BASELINE.json configs[1]):
Sponza-class scene (~262 k triangles, procedural), 1920x1080, 1 spp per frame accumulating, 8 bounces
(`orig` preset: depth 8, tmin 0.001), Disney BSDF + environment NEE.  One step = one frame = one pass of the hot
path (SampleRenderer::render()) over W*H camera samples.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

value   = whole-job Msamples/s with everything resident in HBM (CUDA events on the renderer's stream, L2 flushed
          between steps, max over ranks).
e2e     = the same metric through the reference-facing call sequence with HOST buffers: poke launchParams,
          render() (blocking), downloadPixels() into host memory — every step.
roofline= the dominant kernel (traceClosestKernel): algorithmic bytes (32 B/node + 48 B/triangle + 32 B ray +
          16 B hit, counted by the kernel itself in a counters replay of the same frames) / its CUDA-event time.
N > 1   = sample split: rank r renders sub-frames r, r+N, ... into an fp32 SUM buffer, one NCCL
          reduce to rank 0 at the end + divide/tone-map there.  Weak scaling: one frame per rank per step.
--impl reference = the CPU oracle (reference shading headers host-compiled when oracle/_ref exists, else the
          port) on all host cores, on a bounded sample of the same frame; rank 0 only.
"""
import argparse
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
    ap.add_argument("--steps", type=int, default=16)
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


def config_json(a, model, extra=None):
    c = {"workload": f"{a.scene}-class procedural scene, {model.num_triangles()} triangles, {a.width}x{a.height}, "
                     f"{a.spl} spp/frame progressive, preset {a.preset} (max depth {8 if a.preset == 'orig' else 4}), Disney BSDF + probe NEE",
         "triangles": model.num_triangles(), "width": a.width, "height": a.height, "spp_per_step": a.spl, "preset": a.preset,
         "probe": "analytic sun+sky 2048x1024", "l2": "flushed between timed steps (256 MiB write); working set also exceeds the 126 MB L2"}
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
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100", "-i", str(index)],
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
    print(json.dumps(line))
