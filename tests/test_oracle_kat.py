"""The CPU oracle port (oracle/shading_port.h) against golden vectors recorded from the reference's own
host-compiled headers (tests/golden/make_golden.py).  Everything here is BIT-exact: same libm, IEEE fp32."""
import numpy as np

from conftest import golden, materials_from_bytes, probe_from_golden
from optixpathtracer_b200 import host_params, preset, preset_camera, scenes
from optixpathtracer_b200.probe import ProbeData


def bits(a):
    a = np.ascontiguousarray(a)
    return a.view(np.uint32) if a.dtype == np.float32 else a


def assert_bits(a, b, what=""):
    assert np.array_equal(bits(a), bits(b)), f"{what}: {int((bits(a) != bits(b)).sum())} of {bits(a).size} words differ"


def test_tea_lcg_random(port):
    g = golden("kat_rng")
    assert np.array_equal(port.tea4(g["v0"], g["v1"]), g["tea4"])
    assert port.tea4([5], [0])[0] == 1535087949          # SURVEY §8c pin: then rnd -> 0.918797
    st, f = port.lcg_stream(int(g["lcg_seed"]), 64)
    assert np.array_equal(st, g["lcg_state"]) and np.array_equal(bits(f), bits(g["lcg_f"]))
    assert abs(float(f[0]) - 0.918797) < 1e-6
    u, s = port.random_stream(int(g["rand_seed_u"]), 64, False)
    assert np.array_equal(u, g["rand_u"]) and np.array_equal(s, g["rand_state_u"])
    fl, s = port.random_stream(int(g["rand_seed_f"]), 64, True)
    assert_bits(fl, g["rand_f"], "Randf")
    assert np.array_equal(s, g["rand_state_f"])
    assert fl.min() >= 0.0 and fl.max() <= 0.999999


def test_disney_bsdf(port):
    g = golden("bsdf")
    mats = materials_from_bytes(g["materials"])
    U, V = port.basis(g["N"])
    assert_bits(U, g["basisU"], "BasisFromVector u")
    assert_bits(V, g["basisV"], "BasisFromVector v")
    L, pdf, ty, rng = port.bsdf_sample(mats, g["etaI"], g["etaO"], g["N"], g["view"], g["seeds"])
    assert np.array_equal(ty, g["sampleType"])
    assert np.array_equal(rng, g["sampleRng"])            # RNG consumption order is data dependent: 2..7 draws
    assert_bits(pdf, g["samplePdf"], "BSDFSample pdf")
    ok = g["samplePdf"] > 0                               # light is unspecified when pdf == 0 (TIR)
    assert_bits(L[ok], g["sampleL"][ok], "BSDFSample dir")
    f, p = port.bsdf_eval(mats, g["albedo"], g["etaI"], g["etaO"], g["N"], g["view"], g["L"])
    assert_bits(f, g["evalF"], "BSDFEval")
    assert_bits(p, g["evalPdf"], "BSDFPdf")
    assert len({int(t) for t in ty}) == 3                 # all three lobes exercised


def test_probe(port):
    g = golden("probe")
    pdfX, cdfX, pdfY, cdfY = port.build_cdf(g["data"])
    for a, b, n in ((pdfX, g["pdfX"], "pdfX"), (cdfX, g["cdfX"], "cdfX"), (pdfY, g["pdfY"], "pdfY"), (cdfY, g["cdfY"], "cdfY")):
        assert_bits(a, b, n)
    host = ProbeData(g["data"]).BuildCDF()                # the numpy BuildCDF the product's host side uses
    for a, b, n in ((host.pdfX, g["pdfX"], "pdfX"), (host.cdfX, g["cdfX"], "cdfX"), (host.pdfY, g["pdfY"], "pdfY"), (host.cdfY, g["cdfY"], "cdfY")):
        assert_bits(a, b, "numpy " + n)
    probe = probe_from_golden(g)
    d, c, p = port.probe_sample(probe, g["seeds"])
    assert_bits(d, g["sampleDir"], "ProbeSample dir")
    assert_bits(c, g["sampleColor"], "ProbeSample color")
    assert_bits(p, g["samplePdf"], "ProbeSample pdf")
    uv, ec, ep = port.probe_eval_dir(probe, g["dirs"])
    assert_bits(uv, g["uv"], "ProbeDirToUV")
    assert_bits(ec, g["evalColor"], "ProbeEval")
    assert_bits(ep, g["evalPdf"], "ProbePdf")


def test_display_transforms(port):
    g = golden("display")
    assert np.array_equal(port.display(g["rgb"], 0, 1.0), g["srgb"])
    assert np.array_equal(port.display(g["rgb"], 1, 4.0), g["reinhard4"])
    assert np.array_equal(port.display(g["rgb"], 1, 8.0), g["reinhard8"])
    rgba = np.concatenate([g["rgb"], np.ones((len(g["rgb"]), 1), np.float32)], 1)
    assert np.array_equal(port.tonemap_sqrt(rgba), g["sqrt"])


def test_camera(port):
    g = golden("camera")
    for name in ("cornell", "sponza", "lost_empire"):
        cam = preset_camera(name)
        got = np.stack(port.camera_uvw(cam.eye, cam.lookat, cam.up, float(cam.fovY), 16.0 / 9.0))
        assert_bits(got, g[name], name)
        cam.setAspectRatio(16.0 / 9.0)                    # the host-side numpy camera agrees to fp32 rounding
        np.testing.assert_allclose(np.stack(cam.UVWFrame()), g[name], rtol=2e-6, atol=1e-5)


def _render_cornell(oracle, variant, pname, probe_data):
    model = scenes.cornell(variant)
    probe = ProbeData(probe_data).BuildCDF()
    sc = oracle.scene(model, probe, brute_force=True)
    cfg = preset(pname, uniform_spl=4)
    cfg.accumulate = 1
    p = host_params(48, 48, preset_camera("cornell"), samples_per_launch=4)
    accum = frame = None
    for sub in range(2):
        p.frame.subframe_index = sub
        accum, frame, cnt, _ = sc.render(p, cfg, accum, frame)
    p0 = host_params(48, 48, preset_camera("cornell"), samples_per_launch=1)
    return accum, frame, sc.trace_primary(p0, cfg), cnt


def test_cornell_render_matches_reference_backend(port):
    g = golden("cornell_render")
    for variant, pname in (("disney", "sv4_uniform"), ("diffuse", "orig")):
        accum, frame, (prim, mesh, t, u, v), cnt = _render_cornell(port, variant, pname, g["probe"])
        assert_bits(accum, g[f"{variant}_accum"], variant + " accum")
        assert np.array_equal(frame, g[f"{variant}_frame"])
        assert np.array_equal(prim, g[f"{variant}_prim"]) and np.array_equal(mesh, g[f"{variant}_mesh"])
        assert_bits(t, g[f"{variant}_t"], variant + " t")
        assert cnt["samples"] == 48 * 48 * 4 and cnt["radiance_rays"] > cnt["samples"]


def test_port_equals_reference_live(port, ref):
    """Where the reference tree is present: the two backends agree bit-for-bit on a fresh workload too."""
    rng = np.random.default_rng(99)
    v0 = rng.integers(0, 2**32, 1000, dtype=np.uint64).astype(np.uint32)
    assert np.array_equal(port.tea4(v0, v0[::-1].copy()), ref.tea4(v0, v0[::-1].copy()))
    data = rng.uniform(0.01, 3, (8, 16, 4)).astype(np.float32)
    for a, b in zip(port.build_cdf(data), ref.build_cdf(data)):
        assert_bits(a, b)
    g = golden("cornell_render")
    a1 = _render_cornell(port, "disney", "sv3_uniform", g["probe"])[0]
    a2 = _render_cornell(ref, "disney", "sv3_uniform", g["probe"])[0]
    assert_bits(a1, a2, "sv3-shading render")


def test_bvh_oracle_equals_brute_force(port):
    """The oracle's BVH is only an accelerator for the brute-force contract."""
    model = scenes.sponza_class(4000)
    a, b = port.scene(model), port.scene(model, brute_force=True)
    rng = np.random.default_rng(3)
    n = 3000
    org = rng.uniform(-1500, 1500, (n, 3)).astype(np.float32) * np.array([1, 0.3, 0.4], np.float32) + np.array([0, 400, 0], np.float32)
    d = rng.normal(size=(n, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    ra, rb = a.trace_rays(org, d, 0.01, 1e16), b.trace_rays(org, d, 0.01, 1e16)
    assert (ra[0] >= 0).sum() > n // 3
    for x, y in zip(ra, rb):
        assert np.array_equal(bits(x), bits(y))
    oa, ob = a.trace_rays(org, d, 0.01, 1e16, True, 1)[0], b.trace_rays(org, d, 0.01, 1e16, True, 1)[0]
    assert np.array_equal(oa, ob)
