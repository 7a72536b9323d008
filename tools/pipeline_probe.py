#!/usr/bin/env python
"""How much do frames in flight buy?  Renders the bench workload K frames back to back through b200pt_render_async with
cfg.frames_in_flight = 1..4 (no host synchronisation, no L2 flush between frames) and prints Msamples/s for each.  Tuning tool, not a bench line."""
import os
import sys

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

    def run(frames):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for f in range(frames):
            r.launchParams.frame.subframe_index = f
            r.render(sync=False)
        r.join()
        e1.record(stream)
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        return a.width * a.height * a.spl * frames / (ms * 1e-3) / 1e6, ms / frames

    for n in (1, 2, 3, 4):
        cfg.frames_in_flight = n
        run(4)
        v, ms = run(a.steps)
        print(f"frames in flight {n}: {v:.1f} Msamples/s, {ms:.3f} ms per frame ({a.steps} frames back to back)")
    r.close()


if __name__ == "__main__":
    main()
