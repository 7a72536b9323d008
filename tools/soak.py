#!/usr/bin/env python
"""Soak of the blocking render() with speculative sub-frames: N frames at 1080p with random camera moves, index jumps, downloads and an occasional
probe change, compared at the end (accumulation buffer of a final static run of 8 frames) with a renderer that never speculated.  Tuning / robustness tool."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from optixpathtracer_b200 import SampleRenderer, preset, preset_camera, scenes  # noqa: E402
from optixpathtracer_b200.probe import analytic_sky  # noqa: E402


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 1500
    model, probe = scenes.sponza_class(60000), analytic_sky(256, 128)
    out = []
    for in_flight in (1, 3):
        rng = np.random.default_rng(7)
        cfg = preset("orig", uniform_spl=1)
        cfg.frames_in_flight = in_flight
        r = SampleRenderer(model, device=0, config=cfg)
        r.resize((960, 540))
        cam = preset_camera("sponza")
        r.setCamera(cam)
        r.setProbe(probe)
        idx, t0 = 0, time.time()
        for k in range(n):
            u = rng.random()
            if u < 0.05:
                cam.eye = (cam.eye[0] + float(rng.normal()), cam.eye[1], cam.eye[2])
                r.setCamera(cam)
                idx = 0
            elif u < 0.07:
                idx += int(rng.integers(2, 9))
            elif u < 0.08:
                r.setProbe(probe)
            r.launchParams.frame.subframe_index = idx
            idx += 1
            r.render()
            if u > 0.9:
                r.downloadPixels()
        for k in range(8):
            r.launchParams.frame.subframe_index = idx + k
            r.render()
        out.append(r.downloadAccum().copy())
        print(f"frames_in_flight {in_flight}: {n + 8} frames in {time.time() - t0:.2f} s")
        r.close()
    same = np.array_equal(out[0], out[1])
    print("accumulation buffers equal:", same)
    sys.exit(0 if same else 1)


if __name__ == "__main__":
    main()
