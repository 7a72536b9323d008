"""Mint tests/golden/hdr.npz: Radiance .hdr files decoded by the REFERENCE's own reader.

Runs only in the build container (needs /root/reference): oracle/_ref/liboracle_ref.so compiles the stb_image.h vendored in the
reference tree (support/stb/) in place, and `oracle_ref_load_hdr` calls stbi_loadf(file, &w, &h, &ch, 4) exactly as loadProbe does
(sv4/main.cpp:154-165).  The fixture holds the file bytes (written by optixpathtracer_b200.probe.write_hdr — the reference ships no
.hdr asset) and the reference's decoded floats; it pins the product's readers (host/Probe.cpp, probe.read_hdr) bit for bit.
Usage:  python tests/golden/make_golden_hdr.py
"""
import ctypes as C
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from optixpathtracer_b200.probe import analytic_sky, write_hdr      # noqa: E402
from oracle import binding                                          # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def main():
    binding.build(ref=True)
    L = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "liboracle_ref.so"))
    L.oracle_ref_load_hdr.argtypes = [C.c_char_p, C.POINTER(C.c_int), C.POINTER(C.c_int), C.c_void_p, C.c_size_t]
    rng = np.random.default_rng(20231105)
    out = {}
    for name, w, h, rle in (("rle", 96, 24, True), ("flat", 40, 9, False), ("narrow", 5, 3, True), ("wide_rle", 300, 7, True)):
        img = analytic_sky(w, h).data[..., :3].copy()
        img[h // 2:, : w // 2] = 0.25                                   # constant block -> long runs
        img[0, 0] = 0.0                                                 # exponent 0
        img *= rng.choice([1e-3, 1.0, 1e3], size=(h, w, 1)).astype(np.float32)
        with tempfile.TemporaryDirectory() as d:
            path = os.path.join(d, name + ".hdr")
            blob = write_hdr(path, img, rle)
            ww, hh = C.c_int(), C.c_int()
            assert L.oracle_ref_load_hdr(path.encode(), C.byref(ww), C.byref(hh), None, 0) == 0
            px = np.empty((hh.value, ww.value, 4), np.float32)
            assert L.oracle_ref_load_hdr(path.encode(), C.byref(ww), C.byref(hh), px.ctypes.data, px.size) == 0
        out[name + "_file"] = np.frombuffer(blob, np.uint8)
        out[name + "_pixels"] = px
    np.savez_compressed(os.path.join(OUT, "hdr.npz"), **out)
    print("hdr.npz written:", {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
