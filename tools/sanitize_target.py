#!/usr/bin/env python
"""Target of the compute-sanitizer passes (tools/sanitize.sh): small scenes, every scheduling mode of the frame loop.
  1. smoke(): Cornell 64x64, 4 spp, checked against the oracle;
  2. 20 blocking render() calls with cfg.frames_in_flight = 3 on a 12 k-triangle Sponza-class scene at 160x90 - progressive accumulation (speculative next
     sub-frames in flight), a camera move in the middle (the speculative wavefronts are abandoned while still running), a downloadPixels every few frames;
  3. the same frames one at a time with the frame split into two concurrent band sets (intra_frame_split = 2), a foveated three-pass frame,
     b200pt_render_async + b200pt_download_pixels_async, a refit, the Whitted integrator, the device-side BuildCDF (both entry points)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry  # noqa: E402
from optixpathtracer_b200 import SampleRenderer, preset, preset_camera, scenes  # noqa: E402
from optixpathtracer_b200.probe import analytic_sky  # noqa: E402


def main():
    entry.smoke()
    model, probe = scenes.sponza_class(12000), analytic_sky(128, 64)
    cam = preset_camera("sponza")
    cfg = preset("orig", uniform_spl=1)
    cfg.frames_in_flight = 3
    r = SampleRenderer(model, device=0, config=cfg)
    r.resize((160, 90))
    r.setCamera(cam)
    r.setProbe(probe)
    idx = 0
    for k in range(20):
        if k == 10:
            cam.eye = np.asarray((cam.eye[0] + 2.0, cam.eye[1], cam.eye[2]), np.float32)
            r.setCamera(cam)
            idx = 0
        r.launchParams.frame.subframe_index = idx
        idx += 1
        r.render()
        if k % 4 == 3:
            r.downloadPixels()
    ref = r.downloadAccum().copy()
    # one at a time, split into band sets; must reproduce the last 10 frames bit for bit
    cfg.frames_in_flight, cfg.intra_frame_split = 1, 2
    r.clearAccum()
    for k in range(10):
        r.launchParams.frame.subframe_index = k
        r.render()
    assert np.array_equal(ref.view(np.uint32), r.downloadAccum().view(np.uint32)), "split frames differ from frames in flight"
    # asynchronous entry points
    cfg.frames_in_flight, cfg.intra_frame_split = 3, 0
    host = np.zeros((90, 160), np.uint32)
    for k in range(6):
        r.launchParams.frame.subframe_index = 10 + k
        r.render(sync=False)
        r.downloadPixelsAsync(host)
    r.synchronize()
    # refit + render, Whitted frame
    v = np.asarray(model.meshes[0].vertex, np.float32) + np.float32(3.0)
    r.updateVertices(0, v)
    r.refitAccel()
    r.render()
    r.renderWhitted(cam)
    r.downloadPixels()
    # foveated three-pass frame
    fov = preset("sv4_fov", inner_radius=12, outer_radius=40)
    r.launchParams.frame.c.x, r.launchParams.frame.c.y = 80, 45
    r.render(fov)
    # device-side BuildCDF through both entry points
    r.setProbeImage(probe.data)
    tables = r.downloadProbe()
    try:
        import torch
        d = torch.from_numpy(probe.data).cuda()
        outs = [torch.empty((64, 128), device="cuda"), torch.empty((64, 128), device="cuda"), torch.empty(64, device="cuda"), torch.empty(64, device="cuda")]
        torch.cuda.synchronize()
        r.buildCdfNoAlloc(128, 64, d.data_ptr(), *[o.data_ptr() for o in outs])
        r.synchronize()
        assert all(np.array_equal(o.cpu().numpy().view(np.uint32), t.view(np.uint32)) for o, t in zip(outs, tables))
    except ImportError:
        pass
    r.close()
    print("sanitize target: OK")


if __name__ == "__main__":
    main()
