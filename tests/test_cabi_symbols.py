"""The C-ABI library loads and exports every symbol include/b200pt.h declares (no compute calls, no GPU)."""
import ctypes as C
import os
import re

import pytest

from conftest import ROOT
from optixpathtracer_b200 import _lib, structs


@pytest.fixture(scope="module")
def lib():
    import __graft_entry__ as g
    g.build()
    return _lib.lib()


def header_symbols():
    text = open(os.path.join(ROOT, "include", "b200pt.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b200pt_[a-z_0-9]+)\s*\(", text)))


def test_every_declared_symbol_is_exported(lib):
    names = header_symbols()
    assert len(names) >= 25
    for n in names:
        assert hasattr(lib, n), n
    assert set(names) <= set(_lib.SYMBOLS)
    assert hasattr(lib, "computeFinalPixelColors")       # the reference's own extern "C" export, sv4/toneMap.h:4


def test_presets_match_python_twin(lib):
    for name in ("orig", "sv3_uniform", "sv4_uniform", "sv4_fov"):
        c = structs.Config()
        assert lib.b200pt_config_preset(C.byref(c), name.encode()) == 0
        assert bytes(c) == bytes(structs.preset(name)), name
    c = structs.Config()
    assert lib.b200pt_config_preset(C.byref(c), b"nope") == -1
    assert b"unknown preset" in lib.b200pt_last_error()


def test_no_cpu_fallback(lib):
    """Without a CUDA device, context creation must fail loudly instead of falling back."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    assert not lib.b200pt_create(0)
    assert b"no CPU fallback" in lib.b200pt_last_error()
    from optixpathtracer_b200 import SampleRenderer, scenes
    with pytest.raises(_lib.B200ptError):
        SampleRenderer(scenes.cornell())


def test_product_does_not_import_oracle():
    """Only tests/, smoke() and bench.py may touch oracle/."""
    pkg = os.path.join(ROOT, "optixpathtracer_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".inl", ".cpp", ".h")):
                src = open(os.path.join(dirpath, f), errors="replace").read()
                assert not re.search(r"^\s*(from|import)\s+oracle|#include\s+[\"<].*oracle/", src, flags=re.M), f


def test_header_constants_match_python_twin():
    """Sizes the Python mirror hard-codes must be the ones include/b200pt.h declares (handle of the one-sided tile split, NCCL id, no-alloc scratch)."""
    import re
    from optixpathtracer_b200.renderer import SampleRenderer
    from optixpathtracer_b200.structs import NOALLOC_MAX_ROWS
    text = open(os.path.join(ROOT, "include", "b200pt.h")).read()

    def const(name):
        return int(re.search(rf"#define\s+{name}\s+(\d+)", text).group(1))

    assert const("B200PT_PEER_HANDLE_BYTES") == SampleRenderer.PEER_HANDLE_BYTES == 192
    assert const("B200PT_UNIQUE_ID_BYTES") == 128
    assert const("B200PT_NOALLOC_MAX_ROWS") == NOALLOC_MAX_ROWS
