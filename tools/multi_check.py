#!/usr/bin/env python
"""Multi-GPU correctness check of the C-ABI collectives (run under torchrun on an N-GPU box; N >= 2):
     python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/multi_check.py
  * sample split: rank r renders sub-frames r, r+N, ... in ACCUM_SUM mode, b200pt_reduce_sum onto rank 0 -> equals the single-GPU running mean of
    the same sub-frames up to fp32 summation order, and the tone-mapped frame to 1/255;
  * tile split: rank r renders bands b % N == r of ONE sub-frame, b200pt_gather_bands -> frame and accumulation buffer on rank 0 bit-identical to
    the single-GPU frame;
  * one-sided tile split (b200pt_peer_*): the same partition with the other ranks' finalize passes writing into rank 0's frame over NVLink;
  * the same for the three-pass foveated schedule at 3840x2160 (pass barriers between the ranks), with its frame time against one GPU.
torch.distributed only carries the 128-byte NCCL unique id (and the final verdict); the collectives are the library's own."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from optixpathtracer_b200 import SampleRenderer, preset, preset_camera, scenes  # noqa: E402
from optixpathtracer_b200.probe import analytic_sky  # noqa: E402
from optixpathtracer_b200.structs import ACCUM_SUM  # noqa: E402


def share_unique_id(rank):
    uid = torch.from_numpy(SampleRenderer.commUniqueId().copy()) if rank == 0 else torch.zeros(128, dtype=torch.uint8)
    uid = uid.cuda()
    dist.broadcast(uid, src=0)
    return uid.cpu().numpy()


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    W, H, FR = 640, 360, 2 * world
    model = scenes.sponza_class(30000)
    probe = analytic_sky(256, 128)
    cam = preset_camera("sponza")
    ok = True

    # ---- sample split
    cfg = preset("orig", uniform_spl=1)
    cfg.accumulate = 1
    cfg.accum_mode = ACCUM_SUM
    cfg.frames_in_flight = 3
    r = SampleRenderer(model, device=local, config=cfg)
    r.resize((W, H))
    r.setCamera(cam)
    r.setProbe(probe)
    r.commInit(world, rank, share_unique_id(rank))
    info = r.commInfo()
    r.clearAccum()
    for f in range(rank, FR, world):
        r.launchParams.frame.subframe_index = f
        r.render(sync=False)
    r.reduceSum(FR, root=0)
    r.synchronize()
    if rank == 0:
        got_acc, got_px = r.downloadAccum().copy(), r.downloadPixels().copy()
        cfg1 = preset("orig", uniform_spl=1)
        cfg1.accumulate = 1
        one = SampleRenderer(model, device=local, config=cfg1)
        one.resize((W, H))
        one.setCamera(cam)
        one.setProbe(probe)
        for f in range(FR):
            one.launchParams.frame.subframe_index = f
            one.render()
        want_acc, want_px = one.downloadAccum(), one.downloadPixels()
        err = float(np.abs(got_acc[..., :3] - want_acc[..., :3]).max() / max(1e-6, float(np.abs(want_acc[..., :3]).max())))
        d8 = np.abs(got_px.view(np.uint8).astype(int) - want_px.view(np.uint8).astype(int))
        s_ok = err < 2e-5 and (d8 > 1).mean() < 1e-3
        print(f"sample split over {world} ranks: max rel err of the mean {err:.2e}, pixels off by > 1/255: {(d8 > 1).mean():.2e}, comm {info} -> {'OK' if s_ok else 'FAIL'}", flush=True)
        ok &= s_ok
        one.close()
    # re-initialise the communicator on a live context (the accumulation buffer moves, contents kept), then drop it: the context must
    # keep rendering into a valid buffer that launchParams follows
    before = r.downloadAccum().copy()
    r.commInit(world, rank, share_unique_id(rank))
    same_after_reinit = np.array_equal(before.view(np.uint32), r.downloadAccum().view(np.uint32))
    r.commDestroy()
    same_after_destroy = np.array_equal(before.view(np.uint32), r.downloadAccum().view(np.uint32))
    r.config.accum_mode = 0
    r.launchParams.frame.subframe_index = 0
    r.render()
    alive = bool(np.isfinite(r.downloadAccum()).all()) and r.commInfo()["n_ranks"] == 1
    c_ok = same_after_reinit and same_after_destroy and alive
    print(f"rank {rank}: comm re-init keeps the buffer: {same_after_reinit}, comm destroy keeps it: {same_after_destroy}, renders afterwards: {alive}", flush=True)
    flag = torch.tensor([1 if c_ok else 0], device="cuda")
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    ok &= bool(int(flag.item()))
    r.close()

    # ---- tile split
    cfg = preset("sv3_uniform", uniform_spl=2)
    cfg.tile_rank, cfg.tile_count = rank, world
    r = SampleRenderer(model, device=local, config=cfg)
    r.resize((W, H))
    r.setCamera(cam)
    r.setProbe(probe)
    r.commInit(world, rank, share_unique_id(rank))
    r.clearAccum()
    r.launchParams.frame.subframe_index = 0
    r.render(sync=False)
    r.gatherBands(root=0, frame=True, accum=True)
    r.synchronize()
    if rank == 0:
        got_acc, got_px = r.downloadAccum().copy(), r.downloadPixels().copy()
        one = SampleRenderer(model, device=local, config=preset("sv3_uniform", uniform_spl=2))
        one.resize((W, H))
        one.setCamera(cam)
        one.setProbe(probe)
        one.launchParams.frame.subframe_index = 0
        one.render()
        t_ok = np.array_equal(got_px, one.downloadPixels()) and np.array_equal(got_acc.view(np.uint32), one.downloadAccum().view(np.uint32))
        print(f"tile split over {world} ranks: gathered frame + accumulation buffer bit-identical to one GPU: {t_ok}", flush=True)
        ok &= bool(t_ok)
        one.close()
    dist.barrier()

    # ---- tile split, one-sided: no communicator at all - rank 0 exports its frame, the others map it (CUDA IPC) and their finalize passes write
    # their bands into it; four progressive frames with the ranks running ahead of rank 0's acquire / download / release
    r.commDestroy()
    r.config.frames_in_flight = 3
    r.clearAccum()
    h = torch.from_numpy(r.peerExport(0, world).copy()).cuda() if rank == 0 else torch.zeros(192, dtype=torch.uint8, device="cuda")
    dist.broadcast(h, src=0)
    if rank != 0:
        r.peerAttach(h.cpu().numpy(), rank, world)
    frames = []
    for f in range(4):
        r.launchParams.frame.subframe_index = f
        r.render(sync=False)
        if rank == 0:
            r.peerAcquire()
            frames.append(r.downloadPixels().copy())
            r.peerRelease()
    st = r.peerStatus()
    if rank == 0:
        one = SampleRenderer(model, device=local, config=preset("sv3_uniform", uniform_spl=2))
        one.resize((W, H))
        one.setCamera(cam)
        one.setProbe(probe)
        p_ok = True
        for f in range(4):
            one.launchParams.frame.subframe_index = f
            one.render()
            p_ok &= bool(np.array_equal(frames[f], one.downloadPixels()))
        print(f"one-sided tile split over {world} ranks: 4 progressive frames byte-identical to one GPU: {p_ok}, status {st}", flush=True)
        ok &= p_ok
        one.close()
    flag = torch.tensor([1 if st["state"] == 0 else 0], device="cuda")
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    ok &= bool(int(flag.item()))
    dist.barrier()
    r.close()

    # ---- the foveated schedule (three passes, later passes overwrite earlier ones) split one-sided: pass barriers between the ranks.  3840x2160, the
    # gaze moves every frame, subframe_index 0 (the reference's interactive mode); byte-identical to one GPU, and timed (device, max over ranks)
    W4, H4 = 3840, 2160
    fcfg = preset("sv4_fov")
    fcfg.tile_rank, fcfg.tile_count = rank, world
    r = SampleRenderer(model, device=local, config=fcfg)
    r.resize((W4, H4))
    r.setCamera(cam)
    r.setProbe(probe)
    r.clearAccum()
    h = torch.from_numpy(r.peerExport(0, world).copy()).cuda() if rank == 0 else torch.zeros(192, dtype=torch.uint8, device="cuda")
    dist.broadcast(h, src=0)
    if rank != 0:
        r.peerAttach(h.cpu().numpy(), rank, world)
    _, _, sp = r.deviceBuffers()
    stream = torch.cuda.ExternalStream(sp, device=local)
    gazes = [(1920 + 37 * k, 1080 + 23 * k) for k in range(20)]
    got, ms = [], 0.0
    for k, g in enumerate(gazes):
        r.launchParams.frame.c.x, r.launchParams.frame.c.y = g
        r.launchParams.frame.subframe_index = 0
        dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        r.render(sync=False)
        if rank == 0:
            r.peerAcquire()
        else:
            r.join()
        e1.record(stream)
        r.synchronize()
        torch.cuda.synchronize()
        if k >= 4:
            ms += e0.elapsed_time(e1)
        if rank == 0:
            if k < 3:
                got.append(r.downloadPixels().copy())
            r.peerRelease()
    t = torch.tensor([ms / (len(gazes) - 4)], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    st = r.peerStatus()
    if rank == 0:
        one = SampleRenderer(model, device=local, config=preset("sv4_fov"))
        one.resize((W4, H4))
        one.setCamera(cam)
        one.setProbe(probe)
        f_ok, one_ms = True, 0.0
        _, _, sp1 = one.deviceBuffers()
        s1 = torch.cuda.ExternalStream(sp1, device=local)
        for k, g in enumerate(gazes):
            one.launchParams.frame.c.x, one.launchParams.frame.c.y = g
            one.launchParams.frame.subframe_index = 0
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(s1)
            one.render(sync=False)
            one.join()
            e1.record(s1)
            one.synchronize()
            torch.cuda.synchronize()
            if k >= 4:
                one_ms += e0.elapsed_time(e1)
            if k < 3:
                f_ok &= bool(np.array_equal(got[k], one.downloadPixels()))
        one_ms /= len(gazes) - 4
        print(f"one-sided FOVEATED tile split over {world} ranks (3840x2160, sv4 preset, gaze moves every frame): 3 frames byte-identical to one GPU: {f_ok}; "
              f"{float(t.item()):.3f} ms per frame against {one_ms:.3f} ms on one GPU ({one_ms / float(t.item()):.2f}x), status {st}", flush=True)
        ok &= f_ok
        one.close()
    flag = torch.tensor([1 if st["state"] == 0 else 0], device="cuda")
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    ok &= bool(int(flag.item()))
    dist.barrier()
    r.close()
    v = torch.tensor([1 if ok else 0], device="cuda")
    dist.broadcast(v, src=0)
    dist.barrier()
    dist.destroy_process_group()
    if rank == 0:
        print("MULTI_CHECK " + ("PASSED" if ok else "FAILED"), flush=True)
    sys.exit(0 if int(v.item()) else 1)


if __name__ == "__main__":
    main()
