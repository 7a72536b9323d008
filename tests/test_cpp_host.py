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
