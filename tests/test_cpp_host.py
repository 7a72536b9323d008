"""The C++ host side above the C-ABI (include/ref_api/*.h, host/): same surface as the reference's Model.h / SimplePathtracer.h.

CPU: the C++ loadOBJ/addBox agree with the Python mirror field by field; the driver fails loudly without a GPU.
GPU: the C++ SampleRenderer (header-only forwarder) renders the Cornell OBJ bit-identically to the Python mirror.
"""
import os
import struct
import subprocess

import numpy as np
import pytest

from optixpathtracer_b200 import SampleRenderer, loadColor, loadOBJ, preset, preset_camera, scenes
from optixpathtracer_b200.structs import Material

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST = os.path.join(ROOT, "host")
EXE = os.path.join(HOST, "b200pt_headless")


@pytest.fixture(scope="module")
def exe():
    lib = os.path.join(ROOT, "optixpathtracer_b200", "csrc", "libb200pt.so")
    if not os.path.exists(lib):
        subprocess.check_call(["make", "-s", "-C", os.path.dirname(lib), "-j4"])
    subprocess.check_call(["make", "-s", "-C", HOST])
    return EXE


def read_dump(path):
    raw = open(path, "rb").read()
    nm, nt = struct.unpack_from("ii", raw, 0)
    off, meshes = 8, []
    for _ in range(nm):
        nv, nn, ntc, ni, tex = struct.unpack_from("5i", raw, off)
        off += 20
        mat = raw[off:off + 104]
        off += 104
        v = np.frombuffer(raw, np.float32, nv * 3, off).reshape(-1, 3); off += nv * 12
        n = np.frombuffer(raw, np.float32, nn * 3, off).reshape(-1, 3); off += nn * 12
        tc = np.frombuffer(raw, np.float32, ntc * 2, off).reshape(-1, 2); off += ntc * 8
        idx = np.frombuffer(raw, np.uint32, ni * 3, off).reshape(-1, 3); off += ni * 12
        meshes.append(dict(vertex=v, normal=n, texcoord=tc, index=idx, material=mat, tex=tex))
    return meshes, nt


def test_cpp_loadobj_matches_python_mirror(exe, tmp_path):
    scenes.cornell("disney", directory=str(tmp_path))
    obj = os.path.join(str(tmp_path), "cornell.obj")
    assert os.path.exists(obj)
    dump = os.path.join(str(tmp_path), "model.bin")
    subprocess.check_call([exe, "--obj", obj, "--dump-model", dump, "--no-render"])
    meshes, nt = read_dump(dump)
    ref = loadOBJ(obj)
    assert len(meshes) == len(ref.meshes) and nt == len(ref.textures)
    import ctypes
    for got, want in zip(meshes, ref.meshes):
        assert np.array_equal(got["vertex"], want.vertex) and np.array_equal(got["index"], want.index)
        assert got["material"] == ctypes.string_at(ctypes.byref(want.material), 104)
        assert got["tex"] == want.diffuseTextureID


def test_cpp_loadobj_missing_file_throws(exe):
    p = subprocess.run([exe, "--obj", "/nonexistent/scene.obj", "--no-render"], capture_output=True, text=True)
    assert p.returncode == 1 and "Could not read OBJ model" in p.stderr


def test_cpp_driver_fails_loudly_without_gpu(exe):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    p = subprocess.run([exe, "--boxes", "2", "--frames", "1"], capture_output=True, text=True)
    assert p.returncode == 1 and "no CPU fallback" in p.stderr


@pytest.mark.gpu
def test_cpp_renderer_equals_python_mirror(exe, tmp_path):
    scenes.cornell("disney", directory=str(tmp_path))
    obj = os.path.join(str(tmp_path), "cornell.obj")
    out = os.path.join(str(tmp_path), "out.ppm")
    subprocess.check_call([exe, "--obj", obj, "--preset", "sv4_uniform", "--size", "96x64", "--frames", "2", "--camera", "cornell",
                           "--sky", "1,1,1", "--out", out])
    raw = open(out, "rb").read()
    assert raw.startswith(b"P6\n96 64\n255\n")
    got = np.frombuffer(raw, np.uint8, 96 * 64 * 3, len(b"P6\n96 64\n255\n")).reshape(64, 96, 3)[::-1]
    r = SampleRenderer(loadOBJ(obj), device=0, config=preset("sv4_uniform"))
    r.setCamera(preset_camera("cornell"))
    r.resize((96, 64))
    r.setProbe(loadColor((1.0, 1.0, 1.0), 512, 256))
    for _ in range(2):
        r.setCamera(preset_camera("cornell"))
        r.render()
    want = r.downloadPixels().view(np.uint8).reshape(64, 96, 4)[..., :3]
    r.close()
    assert np.array_equal(got, want)
    # the same frames as a PNG (the reference's screenshot format, sutil.cpp:723-739): decode with zlib and compare
    import zlib
    png = os.path.join(str(tmp_path), "out.png")
    subprocess.check_call([exe, "--obj", obj, "--preset", "sv4_uniform", "--size", "96x64", "--frames", "2", "--camera", "cornell",
                           "--sky", "1,1,1", "--out", png])
    raw = open(png, "rb").read()
    assert raw[:8] == b"\x89PNG\r\n\x1a\n"
    pos, idat, ihdr = 8, b"", None
    while pos < len(raw):
        n, typ = struct.unpack(">I4s", raw[pos:pos + 8])
        body = raw[pos + 8:pos + 8 + n]
        assert zlib.crc32(typ + body) == struct.unpack(">I", raw[pos + 8 + n:pos + 12 + n])[0]
        if typ == b"IHDR":
            ihdr = struct.unpack(">IIBBBBB", body)
        if typ == b"IDAT":
            idat += body
        pos += 12 + n
    assert ihdr == (96, 64, 8, 2, 0, 0, 0)
    rows = np.frombuffer(zlib.decompress(idat), np.uint8).reshape(64, 1 + 96 * 3)
    assert np.all(rows[:, 0] == 0) and np.array_equal(rows[:, 1:].reshape(64, 96, 3)[::-1], want)


@pytest.mark.gpu
def test_cpp_frames_in_flight_give_the_same_image(exe, tmp_path):
    """The C++ host with 3 frames in flight - blocking render() with speculative next sub-frames (--in-flight 3) and renderAsync()
    (--in-flight 3 --async) - writes the same pixels as plain blocking render() calls."""
    scenes.cornell("disney", directory=str(tmp_path))
    obj = os.path.join(str(tmp_path), "cornell.obj")
    imgs = []
    for k, extra in ((1, []), (3, []), (3, ["--async"])):
        out = os.path.join(str(tmp_path), f"out{len(imgs)}.ppm")
        subprocess.check_call([exe, "--obj", obj, "--preset", "sv3_uniform", "--spl", "2", "--size", "96x64", "--frames", "7", "--camera", "cornell",
                               "--sky", "1,1,1", "--in-flight", str(k), "--out", out] + extra)
        imgs.append(open(out, "rb").read())
    assert imgs[0] == imgs[1] == imgs[2] and len(imgs[0]) > 96 * 64 * 3


@pytest.mark.gpu
def test_cpp_multi_gpu_driver_one_rank(exe, tmp_path):
    """The multi-GPU driver below Python (host/headless_main.cpp multiMain: one thread + one SampleRenderer per GPU, b200pt_comm_init, the library's
    NCCL collectives) on the ranks this box has: one.  A tile split over one rank + gather is the plain sequence bit for bit; a sample split's
    resolved SUM equals the running mean to the last RGBA8 step (fp32 summation order).  tools/multi_check.py and bench.py --gpus N cover N > 1."""
    import re
    scenes.cornell("disney", directory=str(tmp_path))
    obj = os.path.join(str(tmp_path), "cornell.obj")
    base = [exe, "--obj", obj, "--preset", "sv3_uniform", "--size", "96x64", "--frames", "6", "--camera", "cornell", "--sky", "1,1,1"]

    def checksum(extra):
        p = subprocess.run(base + extra, capture_output=True, text=True)
        assert p.returncode == 0, p.stderr
        return int(re.search(r"pixel checksum (\d+)", p.stdout).group(1)), p.stdout

    plain, _ = checksum([])
    tile, out = checksum(["--ranks", "1", "--split", "tile"])
    assert "1 ranks, tile split" in out and tile == plain
    samp, out = checksum(["--ranks", "1", "--split", "sample"])
    # the checksum sums the packed 24-bit pixels: one RGBA8 step of a blue channel moves it by 65536, so a few pixels off by one step stay below 1e-3
    assert "1 ranks, sample split" in out and abs(samp - plain) <= 1e-3 * plain
    # one-sided tile split: two rank threads sharing this box's one GPU (no NCCL in this mode), rank 1's finalize pass writes into rank 0's frame
    peer, out = checksum(["--ranks", "2", "--split", "peer", "--devices", "1"])
    assert "2 ranks, tile (one-sided) split" in out and peer == plain


# ---- Radiance .hdr readers (loadProbe, sv4/main.cpp:154-165) against the reference's own stb_image (tests/golden/hdr.npz) ----
HDR_CASES = ["rle", "flat", "narrow", "wide_rle"]


@pytest.mark.parametrize("case", HDR_CASES)
def test_hdr_readers_match_reference_stb(exe, tmp_path, case):
    from conftest import golden
    from optixpathtracer_b200.probe import read_hdr
    g = golden("hdr")
    blob, want = g[case + "_file"].tobytes(), g[case + "_pixels"]
    got = read_hdr(blob)                                                              # Python mirror
    assert got.shape == want.shape and np.array_equal(got.view(np.uint32), want.view(np.uint32))
    path = os.path.join(str(tmp_path), case + ".hdr")
    open(path, "wb").write(blob)
    dump = os.path.join(str(tmp_path), "probe.bin")
    subprocess.check_call([exe, "--boxes", "1", "--hdr", path, "--dump-probe", dump, "--no-render"])   # C++ host (host/Probe.cpp)
    raw = open(dump, "rb").read()
    w, h = struct.unpack_from("ii", raw, 0)
    px = np.frombuffer(raw, np.float32, w * h * 4, 8).reshape(h, w, 4)
    assert (h, w, 4) == want.shape and np.array_equal(px.view(np.uint32), want.view(np.uint32))


def test_hdr_reader_rejects_bad_files(exe, tmp_path):
    from optixpathtracer_b200.probe import read_hdr, write_hdr
    good = write_hdr(None, np.full((4, 16, 3), 0.5, np.float32))
    for blob in (b"P6\n1 1\n255\n...", good.replace(b"FORMAT=32-bit_rle_rgbe", b"FORMAT=32-bit_rle_xyze"),
                 good.replace(b"-Y 4 +X 16", b"+Y 4 +X 16"), good[:-5]):
        with pytest.raises(RuntimeError):
            read_hdr(blob)
        path = os.path.join(str(tmp_path), "bad.hdr")
        open(path, "wb").write(blob)
        p = subprocess.run([exe, "--boxes", "1", "--hdr", path, "--no-render"], capture_output=True, text=True)
        assert p.returncode == 1 and "Could not load HDR probe" in p.stderr
    p = subprocess.run([exe, "--boxes", "1", "--hdr", "/nonexistent/sky.hdr", "--no-render"], capture_output=True, text=True)
    assert p.returncode == 1 and "Could not load HDR probe" in p.stderr


@pytest.mark.gpu
def test_cpp_hdr_probe_with_device_cdf_equals_host_cdf(exe, tmp_path):
    """--gpu-cdf (SampleRenderer::setProbeImage, tables built on the GPU) renders the same pixels as loadProbe + setProbe."""
    from optixpathtracer_b200.probe import analytic_sky, write_hdr
    path = os.path.join(str(tmp_path), "sky.hdr")
    write_hdr(path, analytic_sky(256, 128).data)
    outs = []
    for extra in ([], ["--gpu-cdf"]):
        out = os.path.join(str(tmp_path), "o%d.ppm" % len(outs))
        subprocess.check_call([exe, "--boxes", "12", "--preset", "orig", "--size", "96x64", "--frames", "2", "--hdr", path, "--out", out] + extra)
        outs.append(open(out, "rb").read())
    assert outs[0] == outs[1] and len(outs[0]) > 96 * 64 * 3


@pytest.mark.gpu
def test_cpp_whitted_frame_equals_python_mirror(exe, tmp_path):
    """SampleRenderer::renderWhitted (C++ forwarder, camera set up as 04HelloRaytracing/SampleRenderer.cpp:125-138) gives the same
    RGBA8 frame as the Python mirror."""
    scenes.cornell("disney", directory=str(tmp_path))
    obj = os.path.join(str(tmp_path), "cornell.obj")
    out = os.path.join(str(tmp_path), "w.ppm")
    subprocess.check_call([exe, "--obj", obj, "--size", "96x64", "--frames", "1", "--camera", "cornell", "--whitted", "--out", out])
    raw = open(out, "rb").read()
    got = np.frombuffer(raw, np.uint8, 96 * 64 * 3, len(b"P6\n96 64\n255\n")).reshape(64, 96, 3)[::-1]
    r = SampleRenderer(loadOBJ(obj), device=0, config=preset("sv4_fov"))
    r.resize((96, 64))
    r.renderWhitted(preset_camera("cornell"))
    want = r.downloadPixels().view(np.uint8).reshape(64, 96, 4)[..., :3]
    r.close()
    same = np.all(got == want, axis=-1)
    assert same.mean() > 0.995 and np.abs(got.astype(int) - want.astype(int)).max() <= 1 + 254 * (same.mean() < 1.0)
    assert got.max() > 50
