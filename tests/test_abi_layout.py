"""Launch-ABI layouts against SURVEY.md §8b (measured from the reference headers)."""
import ctypes as C
import os
import subprocess

from optixpathtracer_b200 import structs as S
from conftest import ROOT


def test_sizes():
    assert C.sizeof(S.LaunchParams) == 312
    assert C.sizeof(S.Material) == 104
    assert C.sizeof(S.Probe) == 64
    assert C.sizeof(S.TriangleMeshSBTData) == 152
    assert C.sizeof(S.ParallelogramLight) == 60


def test_launchparams_offsets():
    f = S._Frame
    want = dict(accum_buffer=0, frame_buffer=8, color_buffer=16, normal_buffer=24, albedo_buffer=32, size=40, subframe_index=48,
                factor=52, fillSize=64, c=72, r_inner=80, r_outer=84, offset=88, redraw=96)
    for k, v in want.items():
        assert getattr(f, k).offset == v, k
    lp = S.LaunchParams
    want = dict(frame=0, camera=104, samples_per_launch=152, traversable=160, light=168, probe=232, viewportSize=296, white=304)
    for k, v in want.items():
        assert getattr(lp, k).offset == v, k
    assert S._Camera.U.offset == 12 and S._Camera.V.offset == 24 and S._Camera.W.offset == 36


def test_material_and_probe_offsets():
    m = S.Material
    want = dict(emission=0, color=12, absorption=24, eta=36, metallic=40, subsurface=44, specular=48, roughness=52, specularTint=56,
                anisotropic=60, sheen=64, sheenTint=68, clearcoat=72, clearcoatGloss=76, transmission=80, bump=84, bumpTile=88, flags=100)
    for k, v in want.items():
        assert getattr(m, k).offset == v, k
    p = S.Probe
    for k, v in dict(width=0, height=4, data=8, offset=16, pdfValuesX=32, cdfValuesX=40, pdfValuesY=48, cdfValuesY=56).items():
        assert getattr(p, k).offset == v, k
    t = S.TriangleMeshSBTData
    for k, v in dict(vertex=0, normal=8, texcoord=16, index=24, hasTexture=32, texture=40, material=48).items():
        assert getattr(t, k).offset == v, k


def test_c_headers_static_asserts_compile():
    """include/ref_api/*.h carry the same numbers as static_asserts; a host compile checks them."""
    src = '#include "ref_api/LaunchParams.h"\n#include "b200pt.h"\nint main(){return 0;}\n'
    r = subprocess.run(["g++", "-std=c++17", "-fsyntax-only", "-I", os.path.join(ROOT, "include"), "-I", "/usr/local/cuda/include",
                        "-x", "c++", "-"], input=src.encode(), capture_output=True)
    assert r.returncode == 0, r.stderr.decode()


def test_oracle_sees_same_layouts(port):
    assert [port.sizeof(i) for i in range(6)] == [312, 104, 64, 152, 8, C.sizeof(S.Config)]
    assert bytes(port.default_material()) == bytes(S.Material())


def test_reference_headers_agree(ref):
    """The reference's own LaunchParams.h / Material.h, host-compiled: same sizes, same default Material."""
    assert [ref.sizeof(i) for i in range(5)] == [312, 104, 64, 152, 8]
    assert bytes(ref.default_material()) == bytes(S.Material())
