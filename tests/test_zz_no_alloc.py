"""b200pt_build_cdf_noalloc (include/b200pt.h): the device-side BuildCDF with its scratch in a static __device__ array, for callers that cannot
allocate.  No counterpart in the reference.  Collected last on purpose (
a failure here hides nothing after it).

Checked through the C-ABI against the EXISTING entry point for the same operation (b200pt_set_probe_image + b200pt_download_probe, whose scratch
comes from cudaMalloc), with the tolerance the other BuildCDF tests use: every table bit for bit (tests/test_gpu_parity.py,
test_device_build_cdf_is_bit_identical_to_sequential_host).  Two calls in a row share the static array, which keeps the last call's contents: the
second call uses a probe with FEWER rows than the first (stale entries beyond its height must not leak in), then the first probe again."""
import numpy as np
import pytest

from optixpathtracer_b200 import SampleRenderer, scenes
from optixpathtracer_b200._lib import B200ptError
from optixpathtracer_b200.probe import analytic_sky
from optixpathtracer_b200.structs import NOALLOC_MAX_ROWS

pytestmark = pytest.mark.gpu


def _tables_existing(r, data):
    """The existing entry point: uploads the texels, builds the four tables on the device (scratch from cudaMalloc), reads them back."""
    r.setProbeImage(data)
    return [t.copy() for t in r.downloadProbe()]


def _tables_noalloc(r, torch, data):
    """The new entry point on caller-owned device buffers; torch is only this TEST's way of owning device memory."""
    h, w = data.shape[:2]
    d_data = torch.from_numpy(np.ascontiguousarray(data, np.float32)).cuda()
    outs = [torch.full((h, w), -7.0, dtype=torch.float32, device="cuda"), torch.full((h, w), -7.0, dtype=torch.float32, device="cuda"),
            torch.full((h,), -7.0, dtype=torch.float32, device="cuda"), torch.full((h,), -7.0, dtype=torch.float32, device="cuda")]
    torch.cuda.synchronize()                                   # inputs resident before the library's stream reads them
    r.buildCdfNoAlloc(w, h, d_data.data_ptr(), *[o.data_ptr() for o in outs])
    r.synchronize()                                            # the launches went to the renderer's primary stream
    return [o.cpu().numpy() for o in outs]


def test_build_cdf_noalloc_matches_existing_entry_point_twice_in_a_row():
    import torch
    rng = np.random.default_rng(23)
    big = analytic_sky(512, 256).data                          # 256 rows
    small = np.concatenate([rng.uniform(0.0, 4.0, (37, 100, 3)), np.ones((37, 100, 1))], axis=2).astype(np.float32)   # 37 ragged rows, W % 32 != 0
    r = SampleRenderer(scenes.cornell(), device=0)
    want = {"big": _tables_existing(r, big), "small": _tables_existing(r, small)}
    for name, data in (("big", big), ("small", small), ("big", big)):        # consecutive calls on the shared scratch
        got = _tables_noalloc(r, torch, data)
        for g, w_, what in zip(got, want[name], ("pdfX", "cdfX", "pdfY", "cdfY")):
            assert g.shape == w_.shape and np.array_equal(g.view(np.uint32), w_.view(np.uint32)), (name, what)
    # limits and error behaviour: beyond the static scratch -> a clear error, nothing launched
    d = torch.zeros(16, device="cuda")
    with pytest.raises(B200ptError, match="B200PT_NOALLOC_MAX_ROWS"):
        r.buildCdfNoAlloc(1, NOALLOC_MAX_ROWS + 1, d.data_ptr(), d.data_ptr(), d.data_ptr(), d.data_ptr(), d.data_ptr())
    with pytest.raises(B200ptError, match="not valid"):
        r.buildCdfNoAlloc(4, 4, 0, d.data_ptr(), d.data_ptr(), d.data_ptr(), d.data_ptr())
    r.close()


def test_build_cdf_noalloc_at_the_maximum_height():
    """height == B200PT_NOALLOC_MAX_ROWS uses the whole array; compared with the existing entry point bit for bit."""
    import torch
    rng = np.random.default_rng(29)
    data = np.concatenate([rng.uniform(0.0, 2.0, (NOALLOC_MAX_ROWS, 8, 3)), np.ones((NOALLOC_MAX_ROWS, 8, 1))], axis=2).astype(np.float32)
    r = SampleRenderer(scenes.cornell(), device=0)
    want = _tables_existing(r, data)
    got = _tables_noalloc(r, torch, data)
    for g, w_ in zip(got, want):
        assert np.array_equal(g.view(np.uint32), w_.view(np.uint32))
    r.close()
