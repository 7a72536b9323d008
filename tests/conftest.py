import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def golden(name):
    return np.load(os.path.join(GOLDEN, name + ".npz"))


@pytest.fixture(scope="session")
def port():
    """CPU oracle, our restatement (oracle/shading_port.h)."""
    from oracle import binding
    binding.build(ref=False)
    return binding.Oracle("port")


@pytest.fixture(scope="session")
def ref():
    """CPU oracle over the reference's own headers; only where oracle/_ref was built (needs /root/reference)."""
    from oracle import binding
    if os.path.isdir("/root/reference"):
        binding.build(ref=True)
    if not os.path.exists(binding.REF_SO):
        pytest.skip("oracle/_ref/liboracle_ref.so not built (no /root/reference here)")
    return binding.Oracle("reference")


def materials_from_bytes(raw):
    import ctypes
    from optixpathtracer_b200.structs import Material
    out = []
    for row in raw:
        m = Material()
        ctypes.memmove(ctypes.byref(m), row.tobytes(), 104)
        out.append(m)
    return out


def probe_from_golden(g):
    from optixpathtracer_b200.probe import ProbeData
    p = ProbeData(g["data"])
    p.pdfX, p.cdfX, p.pdfY, p.cdfY, p.valid = g["pdfX"], g["cdfX"], g["pdfY"], g["cdfY"], True
    return p
