"""N > 1 host logic at world_size 2 over gloo on the CPU.  The CPU oracle stands in for the device renderer —
it implements the same b200pt_config fields (accum_mode = ACCUM_SUM, tile_rank/tile_count) — so what is tested here is
the partitioning, the one collective of each split, and the resolve; the CUDA renderer's own SUM/tile modes are covered
by tests/test_gpu_parity.py."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from optixpathtracer_b200 import host_params, loadColor, preset, preset_camera, scenes
from optixpathtracer_b200 import multi
from optixpathtracer_b200.structs import ACCUM_SUM

W, H, FRAMES = 48, 80, 4


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _scene():
    from oracle import binding
    model = scenes.cornell("diffuse")
    probe = loadColor((1.0, 1.0, 1.0), 16, 8)
    return binding.Oracle("port").scene(model, probe, brute_force=True)


def _worker(rank, world, port, mode, out_path):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sc = _scene()
    cam = preset_camera("cornell")
    if mode == "sample":
        cfg = preset("sv3_uniform", uniform_spl=2, accum_mode=ACCUM_SUM)
        accum = np.zeros((H, W, 4), np.float32)
        for f in multi.subframes_of_rank(rank, world, FRAMES):
            p = host_params(W, H, cam, subframe_index=f)
            sc.render(p, cfg, accum=accum)
        t = torch.from_numpy(accum)
        multi.reduce_sum_to_root(t)
        if rank == 0:
            np.save(out_path, t.numpy())
    else:
        cfg = preset("sv3_uniform", uniform_spl=2, tile_rank=rank, tile_count=world)
        p = host_params(W, H, cam, subframe_index=0)
        accum, frame, _, _ = sc.render(p, cfg)
        full = multi.gather_bands_to_root(torch.from_numpy(frame.astype(np.int64)))
        if rank == 0:
            np.save(out_path, full.numpy())
    dist.barrier()
    dist.destroy_process_group()


def test_partition_helpers():
    assert multi.subframes_of_rank(1, 4, 10) == [1, 5, 9]
    rows = [multi.band_rows(r, 3, 200) for r in range(3)]
    assert sorted(np.concatenate(rows).tolist()) == list(range(200))
    assert rows[1][0] == 32 and rows[2][0] == 64 and rows[0][32] == 96


def test_sample_split_reduce_matches_single_rank(tmp_path):
    out = str(tmp_path / "sum.npy")
    mp.spawn(_worker, args=(2, _free_port(), "sample", out), nprocs=2, join=True)
    got = np.load(out)
    sc = _scene()
    cam = preset_camera("cornell")
    cfg = preset("sv3_uniform", uniform_spl=2, accum_mode=ACCUM_SUM)
    want = np.zeros((H, W, 4), np.float32)
    for f in range(FRAMES):
        sc.render(host_params(W, H, cam, subframe_index=f), cfg, accum=want)
    assert np.array_equal(got[..., 3], np.full((H, W), FRAMES, np.float32))          # every pixel got all passes
    assert np.allclose(got[..., :3], want[..., :3], rtol=1e-6, atol=1e-7)             # fp32 summation order only
    # and the resolved mean equals the reference's running-mean accumulation within fp32 rounding
    cfg2 = preset("sv3_uniform", uniform_spl=2)
    mean = np.zeros((H, W, 4), np.float32)
    for f in range(FRAMES):
        sc.render(host_params(W, H, cam, subframe_index=f), cfg2, accum=mean)
    assert np.allclose(got[..., :3] / FRAMES, mean[..., :3], rtol=2e-5, atol=1e-6)


def test_tile_split_gather_is_bit_identical(tmp_path):
    out = str(tmp_path / "tiles.npy")
    mp.spawn(_worker, args=(2, _free_port(), "tile", out), nprocs=2, join=True)
    got = np.load(out).astype(np.uint32)
    sc = _scene()
    _, want, _, _ = sc.render(host_params(W, H, preset_camera("cornell"), subframe_index=0), preset("sv3_uniform", uniform_spl=2))
    assert np.array_equal(got, want)
