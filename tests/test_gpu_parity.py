"""Parity of the CUDA path (through the C-ABI) against the CPU oracle and the golden vectors.
Bars: Morton codes + sort order bit-exact; primary prim IDs >= 99.99 % with t to 1e-5 relative;
accumulated radiance within fp32 libm noise of the oracle (tolerances stated per test)."""
import numpy as np
import pytest

from conftest import golden, materials_from_bytes, probe_from_golden
from optixpathtracer_b200 import SampleRenderer, host_params, loadColor, preset, preset_camera, scenes
from optixpathtracer_b200.probe import ProbeData, analytic_sky
from optixpathtracer_b200.renderer import computeFinalPixelColors

pytestmark = pytest.mark.gpu


def make_renderer(model, probe, size, cam, cfg):
    r = SampleRenderer(model, device=0, config=cfg)
    r.resize(size)
    r.setCamera(cam)
    r.setProbe(probe)
    return r


def rel_mismatch(got, want, rtol):
    err = np.abs(got - want)
    return float((err > rtol * (1.0 + np.abs(want))).mean())


@pytest.fixture(scope="module")
def sponza_small():
    return scenes.sponza_class(30000)


# ------------------------------------------------------------------------------------------------ BVH build
@pytest.mark.parametrize("scene", ["cornell", "sponza", "lost_empire", "duplicates"])
def test_morton_and_sort_bit_exact(port, scene, sponza_small):
    if scene == "duplicates":                     # many identical centroids: exercises the stable tie order
        model = scenes.cornell()
        m = model.meshes[-1]
        m.index = np.concatenate([m.index] * 40)
    else:
        model = {"cornell": scenes.cornell, "sponza": lambda: sponza_small, "lost_empire": lambda: scenes.lost_empire_class(20000)}[scene]()
    r = SampleRenderer(model)
    codes, order, bounds = r.debugMorton()
    ocodes, oorder, obounds = port.scene(model).morton()
    assert np.array_equal(bounds, obounds)                 # values (the sign of a zero bound is not significant)
    assert np.array_equal(codes, ocodes)
    assert np.array_equal(order, oorder)
    assert np.all(np.diff(codes[order].astype(np.uint64)) >= 0)
    r.close()


def test_bvh_structure(sponza_small):
    """Every triangle in exactly one leaf; every child box (dequantised) contains its triangles / grandchildren and lies inside its
    parent's; records are 32 bytes with both children."""
    r = SampleRenderer(sponza_small)
    info = r.bvhInfo()
    nodes, tri_ids = r.downloadBvh()
    n = sponza_small.num_triangles()
    assert info.n_triangles == n and np.array_equal(np.sort(tri_ids), np.arange(n)) and info.n_nodes == len(nodes)
    glo, gs = np.array(info.grid_lo[:]), np.array(info.grid_scale[:])

    def boxes(rec, k):                      # child k of record rec -> world-space (lo, hi), conservative
        q0, q1, q2 = (int(x) for x in nodes[rec, 3 * k:3 * k + 3])
        qlo = np.array([q0 & 0xffff, q1 & 0xffff, q2 & 0xffff], np.float64)
        qhi = np.array([q0 >> 16, q1 >> 16, q2 >> 16], np.float64)
        return glo + (qlo - 2.0) / gs, glo + (qhi - 2.0) / gs

    # world-space bounds of every triangle, in leaf order
    V = np.concatenate([m.vertex[m.index].reshape(-1, 3, 3) for m in sponza_small.meshes])[tri_ids]
    tlo, thi = V.min(1), V.max(1)
    link = nodes[:, 6:8].view(np.int32)
    seen = np.zeros(n, np.int32)
    leaves = 0
    assert info.root_link >= 0
    stack = [(int(info.root_link), np.full(3, -np.inf), np.full(3, np.inf))]
    while stack:
        rec, plo, phi = stack.pop()
        for k in range(2):
            lo, hi = boxes(rec, k)
            # parents and children are quantised independently (floor/ceil + 2 cells of padding each), so a child box may stick out of
            # its parent's by a few cells at most; exact containment is checked via the triangles below
            assert np.all(lo >= plo - 6.0 / gs) and np.all(hi <= phi + 6.0 / gs)
            l = int(link[rec, k])
            if l < 0:
                c = -(l + 1)
                first, count = c >> 3, (c & 7) + 1
                seen[first:first + count] += 1
                leaves += 1
                assert np.all(tlo[first:first + count] >= lo) and np.all(thi[first:first + count] <= hi)
            else:
                assert 0 <= l < len(nodes)
                stack.append((l, lo, hi))
    assert np.all(seen == 1) and leaves == info.n_leaves
    assert info.max_leaf_size <= 8 and info.sah_cost > 0 and info.sah_cost_lbvh > 0 and info.refine_levels >= 1

    # the 4-wide tree the traversal kernels walk: the same leaves, each exactly once; every child box contains its triangles;
    # unused slots are inverted boxes with the EMPTY link; at most as many nodes as binary records
    wide = r.downloadBvh4()
    assert info.n_wide_nodes == len(wide) and 0 < len(wide) <= len(nodes) and info.node_bytes == 64 * len(wide)
    assert 3 * info.wide_depth + 4 <= 96
    wl = wide[:, 12:16].view(np.int32)
    q = wide[:, :12].astype(np.int64).reshape(-1, 4, 3)
    wlo = glo + ((q & 0xffff) - 2.0) / gs
    whi = glo + ((q >> 16) - 2.0) / gs
    empty = wl == 0x7fffffff
    assert np.all((q[empty] & 0xffff) == 0xffff) and np.all((q[empty] >> 16) == 0)
    assert np.all((~empty).sum(1) >= 2)
    seen4 = np.zeros(n, np.int32)
    visited = np.zeros(len(wide), np.int32)
    todo = [0]
    while todo:
        rec = todo.pop()
        visited[rec] += 1
        for k in range(4):
            l = int(wl[rec, k])
            if l == 0x7fffffff:
                continue
            if l < 0:
                c = -(l + 1)
                first, count = c >> 3, (c & 7) + 1
                seen4[first:first + count] += 1
                assert np.all(tlo[first:first + count] >= wlo[rec, k]) and np.all(thi[first:first + count] <= whi[rec, k])
            else:
                assert 0 < l < len(wide)
                todo.append(l)
    assert np.all(seen4 == 1) and np.all(visited == 1)
    r.close()


# ------------------------------------------------------------------------------------------------ traversal
def test_primary_hits_cornell_golden():
    g = golden("cornell_render")
    for variant, pname in (("disney", "sv4_uniform"), ("diffuse", "orig")):
        cfg = preset(pname)
        r = make_renderer(scenes.cornell(variant), ProbeData(g["probe"]).BuildCDF(), (48, 48), preset_camera("cornell"), cfg)
        r.launchParams.samples_per_launch = 1
        prim, mesh, t, u, v = r.tracePrimary()
        assert np.array_equal(prim, g[f"{variant}_prim"]) and np.array_equal(mesh, g[f"{variant}_mesh"])
        assert np.array_equal(t.view(np.uint32), g[f"{variant}_t"].view(np.uint32))       # same fp32 op sequence => same bits
        r.close()


def test_primary_hits_sponza_class(port, sponza_small):
    """north_star geometry bar: prim IDs agree on >= 99.99 % of pixels, t to 1e-5 relative."""
    probe = loadColor((1, 1, 1), 16, 8)
    cfg = preset("orig")
    r = make_renderer(sponza_small, probe, (480, 270), preset_camera("sponza"), cfg)
    r.launchParams.samples_per_launch = 1
    prim, mesh, t, u, v = r.tracePrimary()
    sc = port.scene(sponza_small, probe)
    p = host_params(480, 270, preset_camera("sponza"))
    oprim, omesh, ot, ou, ov = sc.trace_primary(p, cfg)
    same = (prim == oprim) & (mesh == omesh)
    assert same.mean() >= 0.9999, same.mean()
    hit = same & (prim >= 0)
    assert hit.mean() > 0.5
    assert np.all(np.abs(t[hit] - ot[hit]) <= 1e-5 * np.abs(ot[hit]))
    assert np.allclose(u[hit], ou[hit], atol=1e-4) and np.allclose(v[hit], ov[hit], atol=1e-4)
    r.close()


def test_random_rays_closest_and_any(port, sponza_small):
    r = SampleRenderer(sponza_small)
    sc = port.scene(sponza_small)
    rng = np.random.default_rng(11)
    n = 20000
    org = (rng.uniform(-1, 1, (n, 3)) * np.array([1800, 600, 700]) + np.array([0, 650, 0])).astype(np.float32)
    d = rng.normal(size=(n, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    d[:50, 0] = 0.0                                   # axis-parallel components: inf / NaN slabs
    d[:50] /= np.linalg.norm(d[:50], axis=1, keepdims=True)
    prim, mesh, t, u, v = r.traceRays(org, d, 0.01, 1e16)
    oprim, omesh, ot, ou, ov = sc.trace_rays(org, d, 0.01, 1e16)
    same = (prim == oprim) & (mesh == omesh)
    assert same.mean() >= 0.9999
    assert np.array_equal(t[same].view(np.uint32), ot[same].view(np.uint32))
    for mode in (0, 1):
        occ = r.traceRays(org, d, 0.01, 1e16, True, mode)[0]
        oocc = sc.trace_rays(org, d, 0.01, 1e16, True, mode)[0]
        assert (occ == oocc).mean() >= 0.9999
    assert r.traceRays(org[:0], d[:0], 0.01, 1e16)[0].shape == (0,)      # empty input
    r.close()


def test_empty_and_single_triangle_scene():
    from optixpathtracer_b200 import Model
    from optixpathtracer_b200.scenes import _mesh
    from optixpathtracer_b200 import Material
    empty = Model()
    r = SampleRenderer(empty)
    o = np.zeros((4, 3), np.float32)
    d = np.tile(np.array([[0, 0, -1]], np.float32), (4, 1))
    assert np.all(r.traceRays(o, d, 0.01, 1e16)[0] == -1)
    r.close()
    one = Model()
    one.meshes.append(_mesh([[-1, -1, -2], [1, -1, -2], [0, 1, -2]], [[0, 1, 2]], Material()))
    r = SampleRenderer(one)
    prim, mesh, t, u, v = r.traceRays(o, d, 0.01, 1e16)
    assert np.all(prim == 0) and np.allclose(t, 2.0)
    r.close()


# ------------------------------------------------------------------------------------------------ shading library, item by item
def ulp_diff(a, b):
    """Distance in units in the last place between two float32 arrays (same sign assumed where it matters; +0 == -0)."""
    a, b = np.ascontiguousarray(a, np.float32), np.ascontiguousarray(b, np.float32)
    ia, ib = a.view(np.int32).astype(np.int64), b.view(np.int32).astype(np.int64)
    ia = np.where(ia < 0, -(ia & 0x7fffffff), ia)
    ib = np.where(ib < 0, -(ib & 0x7fffffff), ib)
    return np.abs(ia - ib)


def test_device_bsdf_kat():
    """The CUDA shading library on the reference-minted fixture (tests/golden/bsdf.npz, from the reference's own Disney.cuh / maths.h host-compiled):
    BasisFromVector bit-equal (double 1.0/sqrt kept); BSDFSample lobe type and the generator's state afterwards EQUAL (data-dependent draw count);
    direction and pdf bit-equal on the refraction branch (no sinf/cosf); elsewhere the direction is a sum of three basis vectors weighted by
    products of sinf/cosf (CUDA's differ from glibc's by <= 2 ulp each): <= 8 ulp of its largest component (measured: 90 % of the directions
    bit-equal, worst 4.7 ulp) and the pdf of that direction <= 32 ulp (measured 12); BSDFEval / BSDFPdf of a GIVEN direction: bit-equal
    (logf in GTR1 happens to agree on the fixture).  sv4/Disney.cuh:152-427."""
    g = golden("bsdf")
    mats = materials_from_bytes(g["materials"])
    r = SampleRenderer(scenes.cornell())
    out = r.debugBsdf(mats, g["etaI"], g["etaO"], g["N"], g["view"], g["L"], g["albedo"], g["seeds"])
    r.close()
    assert np.array_equal(out["basisU"].view(np.uint32), g["basisU"].view(np.uint32))
    assert np.array_equal(out["basisV"].view(np.uint32), g["basisV"].view(np.uint32))
    assert np.array_equal(out["sampleType"], g["sampleType"])
    assert np.array_equal(out["sampleRng"], g["sampleRng"])
    ok = g["samplePdf"] > 0                                             # the direction is unspecified when pdf == 0 (total internal reflection)
    assert np.array_equal(out["samplePdf"] > 0, ok)
    spec = ok & (g["sampleType"] == 2)                                  # eSpecular: Refract(), no trigonometry
    assert spec.sum() > 5
    assert np.array_equal(out["sampleL"][spec].view(np.uint32), g["sampleL"][spec].view(np.uint32))
    assert np.array_equal(out["samplePdf"][spec].view(np.uint32), g["samplePdf"][spec].view(np.uint32))
    scale = np.abs(g["sampleL"][ok]).max(axis=1, keepdims=True)         # ulp of the direction's largest component
    derr = np.abs(out["sampleL"][ok] - g["sampleL"][ok]) / (scale * 2.0 ** -23)
    assert derr.max() <= 8.0, derr.max()
    assert (derr.max(axis=1) == 0).mean() > 0.8
    assert ulp_diff(out["samplePdf"][ok], g["samplePdf"][ok]).max() <= 32
    assert ulp_diff(out["evalPdf"], g["evalPdf"]).max() <= 1
    assert np.array_equal(np.isfinite(out["evalF"]), np.isfinite(g["evalF"]))
    fin = np.isfinite(g["evalF"]).all(axis=1)
    fscale = np.maximum(np.abs(g["evalF"][fin]).max(axis=1, keepdims=True), 1e-30)
    ferr = np.abs(out["evalF"][fin] - g["evalF"][fin]) / (fscale * 2.0 ** -23)
    assert ferr.max() <= 1.0, ferr.max()


@pytest.mark.parametrize("use_guide", [True, False])
def test_device_probe_kat(use_guide):
    """ProbeSample / ProbeDirToUV / ProbeEval / ProbePdf on the device against tests/golden/probe.npz (sv4/Probe.cuh:38-169): the CDF search must
    land on the SAME texel with and without the guide tables (colour bit-equal: it is a table entry), direction / pdf / uv within 4 ulp
    (sinf, cosf, acosf, atan2f)."""
    g = golden("probe")
    probe = probe_from_golden(g)
    r = make_renderer(scenes.cornell(), probe, (16, 16), preset_camera("cornell"), preset("sv4_uniform"))
    out = r.debugProbe(g["seeds"], g["dirs"], use_guide=use_guide)
    r.close()
    assert np.array_equal(out["sampleColor"].view(np.uint32), g["sampleColor"].view(np.uint32))
    assert np.array_equal(out["evalColor"].view(np.uint32), g["evalColor"].view(np.uint32))
    derr = np.abs(out["sampleDir"] - g["sampleDir"]) / 2.0 ** -23       # unit vectors (measured: 1 ulp)
    assert derr.max() <= 4.0, derr.max()
    assert ulp_diff(out["samplePdf"], g["samplePdf"]).max() <= 8        # measured 2
    assert np.abs(out["uv"] - g["uv"]).max() <= 4 * 2.0 ** -24          # measured 2 * 2^-24
    assert ulp_diff(out["evalPdf"], g["evalPdf"]).max() <= 16           # measured 5 (sinf of acosf)


def test_device_probe_kat_large_probe_guided_equals_plain(port):
    """2048 x 1024 analytic sky (the bench probe): 100 k ProbeSample draws through the guide tables and through the reference's plain LowerBound
    return the same texel, direction and pdf bit for bit, and both equal the oracle's choice of texel."""
    probe = analytic_sky(2048, 1024)
    r = make_renderer(scenes.cornell(), probe, (16, 16), preset_camera("cornell"), preset("sv4_uniform"))
    seeds = np.random.default_rng(17).integers(0, 2 ** 32, 100000, dtype=np.uint64).astype(np.uint32)
    dirs = np.zeros((0, 3), np.float32)
    a, b = r.debugProbe(seeds, dirs, use_guide=True), r.debugProbe(seeds, dirs, use_guide=False)
    r.close()
    for k in ("sampleDir", "sampleColor", "samplePdf"):
        assert np.array_equal(a[k].view(np.uint32), b[k].view(np.uint32)), k
    d, c, p = port.probe_sample(probe, seeds)
    assert np.array_equal(a["sampleColor"].view(np.uint32), c.view(np.uint32))


def _catcher_scene():
    """A shadow-catcher ground plane (MATERIAL_FLAG_SHADOW_CATCHER) at y = 0 under two boxes, with an ordinary floor at y = -1 BELOW it: secondary
    rays cross the catcher on the way down and again on the way up (pass-through, sv4/deviceProgram.cu:617-622), primary rays that see the
    catcher collect the shadow term in alpha (:663-664)."""
    from optixpathtracer_b200 import Material, Model, addBox
    from optixpathtracer_b200.scenes import _mesh
    from optixpathtracer_b200.structs import MATERIAL_FLAG_SHADOW_CATCHER
    m = Model()
    addBox(m, Material.orig_defaults(color=(0.8, 0.3, 0.2)), (-0.6, 0.5, 0.0), (0.4, 0.5, 0.4))
    addBox(m, Material(color=(0.2, 0.5, 0.9), emission=(0, 0, 0)), (0.7, 0.35, 0.3), (0.35, 0.35, 0.35))
    catcher = Material.orig_defaults(color=(0.7, 0.7, 0.7))
    catcher.flags = MATERIAL_FLAG_SHADOW_CATCHER
    m.meshes.append(_mesh([[-4, 0, -4], [4, 0, -4], [4, 0, 4], [-4, 0, 4]], [[0, 2, 1], [0, 3, 2]], catcher))
    m.meshes.append(_mesh([[-6, -1, -6], [6, -1, -6], [6, -1, 6], [-6, -1, 6]], [[0, 2, 1], [0, 3, 2]], Material.orig_defaults(color=(0.5, 0.6, 0.4))))
    return m


@pytest.mark.parametrize("pname,spl", [("sv4_uniform", 4), ("orig", 2)])
def test_shadow_catcher_against_oracle(port, pname, spl):
    """Shadow-catcher materials: pass-throughs do not consume depth, so paths that cross the catcher need more trace launches than
    max_depth + 1; the alpha shadow term survives a BSDFSample with pdf <= 0.  Accumulation buffer and pixel alpha against the oracle."""
    from optixpathtracer_b200 import Camera
    model = _catcher_scene()
    probe = analytic_sky(128, 64)
    cfg = preset(pname, uniform_spl=spl)
    cam = Camera(eye=(0.0, 2.2, 4.5), lookat=(0.0, 0.3, 0.0), up=(0, 1, 0), fovY=40.0)
    W, H = 128, 96
    r = make_renderer(model, probe, (W, H), cam, cfg)
    r.launchParams.samples_per_launch = spl
    r.render()
    got = r.downloadAccum()
    cnt = r.counters()
    sc = port.scene(model, probe, brute_force=True)
    p = host_params(W, H, cam, samples_per_launch=spl)
    want, wframe, ocnt, _ = sc.render(p, cfg, np.zeros((H, W, 4), np.float32))
    assert cnt.samples == ocnt["samples"]
    assert rel_mismatch(got[..., :3], want[..., :3], 2e-4) < 1e-2
    assert float(np.sqrt(np.mean((np.clip(got[..., :3], 0, 10) - np.clip(want[..., :3], 0, 10)) ** 2))) < 2e-2
    # every radiance segment the reference traces and keeps is traced here too: pass-throughs included
    assert cnt.radiance_rays > 0.6 * ocnt["radiance_rays"]
    r.close()


# ------------------------------------------------------------------------------------------------ integrator
def test_cornell_render_against_golden():
    """Two accumulated sub-frames of 4 spp, reference-shading golden image.  Identical RNG streams and fp32
    op order; only libm-level functions differ (<= 2 ulp), so all but a handful of pixels agree to 1e-4."""
    g = golden("cornell_render")
    for variant, pname in (("disney", "sv4_uniform"), ("diffuse", "orig")):
        cfg = preset(pname, uniform_spl=4)
        cfg.accumulate = 1
        r = make_renderer(scenes.cornell(variant), ProbeData(g["probe"]).BuildCDF(), (48, 48), preset_camera("cornell"), cfg)
        r.launchParams.samples_per_launch = 4
        for sub in range(2):
            r.launchParams.frame.subframe_index = sub
            r.render()
        accum, frame = r.downloadAccum(), r.downloadPixels()
        want = g[f"{variant}_accum"]
        assert rel_mismatch(accum[..., :3], want[..., :3], 1e-4) < 5e-3, variant
        rmse = float(np.sqrt(np.mean((accum[..., :3] - want[..., :3]) ** 2)))
        assert rmse < 1e-3, (variant, rmse)               # north_star image bar
        assert np.all(accum[..., 3] == 1.0)
        d8 = np.abs(frame.view(np.uint8).astype(int) - g[f"{variant}_frame"].view(np.uint8).astype(int))
        assert (d8 > 1).mean() < 5e-3
        r.close()


@pytest.mark.parametrize("pname,spl", [("sv4_uniform", 8), ("sv3_uniform", 3), ("orig", 2)])
def test_sponza_class_render_against_oracle(port, sponza_small, pname, spl):
    probe = analytic_sky(256, 128)
    cfg = preset(pname, uniform_spl=spl)
    cam = preset_camera("sponza")
    r = make_renderer(sponza_small, probe, (160, 90), cam, cfg)
    r.launchParams.samples_per_launch = spl
    r.launchParams.frame.subframe_index = 3
    r.render()
    got = r.downloadAccum()
    cnt = r.counters()
    sc = port.scene(sponza_small, probe)
    p = host_params(160, 90, cam, samples_per_launch=spl, subframe_index=3)
    p.frame.accum_buffer = 0
    want, wframe, ocnt, _ = sc.render(p, cfg, np.zeros((90, 160, 4), np.float32))
    assert cnt.samples == ocnt["samples"] == 160 * 90 * spl
    # the reference traces one more (dropped) segment + shadow ray per surviving path; we skip exactly those
    assert cnt.radiance_rays <= ocnt["radiance_rays"] and cnt.radiance_rays > 0.6 * ocnt["radiance_rays"]
    assert p.frame.subframe_index == r.launchParams.frame.subframe_index
    if pname != "sv4_uniform":      # accumulate=1: lerp against the (zero) previous buffer happens on both sides
        pass
    assert rel_mismatch(got[..., :3], want[..., :3], 2e-4) < 1e-2, pname
    assert float(np.sqrt(np.mean((np.clip(got[..., :3], 0, 10) - np.clip(want[..., :3], 0, 10)) ** 2))) < 2e-2
    r.close()


def test_aovs_orig(port):
    model = scenes.cornell("diffuse")
    probe = loadColor((1, 1, 1), 16, 8)
    cfg = preset("orig")
    r = make_renderer(model, probe, (64, 64), preset_camera("cornell"), cfg)
    r.launchParams.samples_per_launch = 2
    r.render()
    sc = port.scene(model, probe, brute_force=True)
    p = host_params(64, 64, preset_camera("cornell"), samples_per_launch=2)
    accum, frame, cnt, sec, aov = sc.render(p, cfg, aovs=True)
    assert np.allclose(r.downloadAOV("normal"), aov["normal"], atol=1e-5)
    assert np.allclose(r.downloadAOV("albedo"), aov["albedo"], atol=1e-5)
    assert rel_mismatch(r.downloadAOV("color")[..., :3], aov["color"][..., :3], 1e-4) < 5e-3
    assert r.launchParams.frame.subframe_index == 0      # orig: the caller increments (orig/main.cpp)
    r.close()


def test_foveated_three_pass_schedule(port, sponza_small):
    """sv4 foveated frame: one fused wavefront over the three rings, finalised in launch order."""
    probe = analytic_sky(128, 64)
    cam = preset_camera("sponza")
    W, H = 256, 144
    cfg = preset("sv4_fov", inner_radius=20, outer_radius=60)
    r = make_renderer(sponza_small, probe, (W, H), cam, cfg)
    sc = port.scene(sponza_small, probe)
    p = host_params(W, H, cam)
    accum = np.zeros((H, W, 4), np.float32)
    frame = np.zeros((H, W), np.uint32)
    r.clearAccum()
    for gaze in ((128, 72), (100, 70)):                  # keep every ring inside the frame: no border-clamp write races
        r.launchParams.frame.c.x, r.launchParams.frame.c.y = gaze
        p.frame.c.x, p.frame.c.y = gaze
        r.render()
        accum, frame, ocnt, _ = sc.render(p, cfg, accum, frame)
        cnt = r.counters()
        assert cnt.samples == ocnt["samples"]
        for f in ("fillSize", "redraw", "r_inner", "r_outer", "subframe_index"):
            assert getattr(r.launchParams.frame, f) == getattr(p.frame, f), f
        assert (r.launchParams.frame.offset.x, r.launchParams.samples_per_launch) == (p.frame.offset.x, p.samples_per_launch)
    got = r.downloadAccum()
    assert rel_mismatch(got[..., :3], accum[..., :3], 2e-4) < 1e-2
    d8 = np.abs(r.downloadPixels().view(np.uint8).astype(int) - frame.view(np.uint8).astype(int))
    assert (d8 > 1).mean() < 1e-2
    r.close()


def test_pool_chunking_is_invisible(sponza_small, monkeypatch):
    """Splitting the samples of a launch over several wavefront chunks must not change a single bit."""
    probe = analytic_sky(64, 32)
    cfg = preset("sv3_uniform", uniform_spl=6)
    outs = []
    for slots in ("0", "40000"):
        if slots != "0":
            monkeypatch.setenv("B200PT_POOL_SLOTS", slots)
        r = make_renderer(sponza_small, probe, (96, 54), preset_camera("sponza"), cfg)
        r.render()
        outs.append(r.downloadAccum())
        r.close()
    assert np.array_equal(outs[0].view(np.uint32), outs[1].view(np.uint32))


def test_render_is_deterministic_and_progressive(sponza_small):
    probe = analytic_sky(64, 32)
    cfg = preset("sv3_uniform", uniform_spl=2)
    r = make_renderer(sponza_small, probe, (96, 54), preset_camera("sponza"), cfg)
    r.render()
    a = r.downloadAccum()
    r.launchParams.frame.subframe_index = 0
    r.render()
    assert np.array_equal(a.view(np.uint32), r.downloadAccum().view(np.uint32))
    assert r.launchParams.frame.subframe_index == 1
    for _ in range(3):
        r.render()
    assert r.launchParams.frame.subframe_index == 4
    b = r.downloadAccum()
    # sub-frame 0 is stored unclamped, every later one is clamped to 10 before the lerp (sv4/deviceProgram.cu:545-552)
    assert np.isfinite(b).all() and b[..., :3].max() <= max(10.0, float(a[..., :3].max())) + 1e-3
    r.close()


@pytest.mark.parametrize("pname", ["orig", "sv4_fov"])
def test_frames_in_flight_fold_in_submission_order(sponza_small, pname):
    """b200pt_render_async with cfg.frames_in_flight = 3: consecutive frames overlap on the device (three lanes: stream + pool + queues each)
    but fold into the accumulation buffer in submission order, so accumulation, pixels and AOVs equal the one-at-a-time sequence bit for bit;
    a blocking render() in the middle of the batch and the copies that follow it see every frame submitted before them."""
    probe = analytic_sky(128, 64)
    cam = preset_camera("sponza")

    def run(in_flight):
        cfg = preset(pname)
        if pname == "orig":
            cfg.uniform_spl = 1
        cfg.frames_in_flight = in_flight
        r = make_renderer(sponza_small, probe, (200, 120), cam, cfg)
        if pname == "sv4_fov":
            r.launchParams.frame.c.x, r.launchParams.frame.c.y = 100, 60
            r.config.inner_radius, r.config.outer_radius = 20, 50
        for k in range(7):
            if pname == "orig":
                r.launchParams.frame.subframe_index = k
            r.render(sync=(in_flight == 1 or k == 3))
        acc, px = r.downloadAccum(), r.downloadPixels()             # no explicit synchronize: the copies order themselves after the lanes
        cnt = r.counters()
        r.close()
        return acc, px, cnt

    acc1, px1, cnt1 = run(1)
    acc3, px3, cnt3 = run(3)
    assert np.array_equal(acc1, acc3) and np.array_equal(px1, px3)
    assert (cnt1.samples, cnt1.radiance_rays, cnt1.shadow_rays) == (cnt3.samples, cnt3.radiance_rays, cnt3.shadow_rays)   # of the last frame


@pytest.mark.parametrize("pname", ["orig", "orig_stride2", "sv3_uniform", "sv4_fov"])
def test_blocking_render_with_speculative_next_subframe(sponza_small, pname):
    """Blocking render() with cfg.frames_in_flight = 3: after two calls that looked like progressive accumulation (same launchParams, next
    subframe_index) the library starts the predicted next sub-frame's wavefront before it blocks.  Right or wrong guesses, the accumulation buffer
    and the pixels after every call equal those of plain one-at-a-time calls: static camera, a camera move, a jump of the sub-frame index, a probe
    change and a resize in the middle of the sequence."""
    probe, probe2 = analytic_sky(128, 64), loadColor((0.7, 0.8, 1.0), 64, 32)
    cam, cam2 = preset_camera("sponza"), preset_camera("sponza")
    cam2.eye = (cam2.eye[0] + 0.5, cam2.eye[1], cam2.eye[2])

    step = 2 if pname == "orig_stride2" else 1                     # one rank of a 2-GPU sample split strides over the sub-frames
    pname = "orig" if pname == "orig_stride2" else pname

    def run(in_flight):
        cfg = preset(pname)
        if pname != "sv4_fov":
            cfg.uniform_spl = 1
        cfg.frames_in_flight = in_flight
        r = make_renderer(sponza_small, probe, (160, 96), cam, cfg)
        if pname == "sv4_fov":
            r.launchParams.frame.c.x, r.launchParams.frame.c.y = 80, 48
            r.config.inner_radius, r.config.outer_radius = 16, 40
        out, idx = [], 0
        for k in range(22):
            if k == 8:
                r.setCamera(cam2)                                   # a wrong guess: the speculative wavefront is dropped
            if k == 12:
                idx += 5                                            # the caller jumps ahead
            if k == 15:
                r.setProbe(probe2)
            if k == 18:
                r.resize((144, 92))                                 # (the foveal rings must stay inside the frame: where the reference clamps block
                r.setCamera(cam2)                                   # replication to the border, :534, several launch indices write one pixel - a race in any build)
            if pname == "orig":
                r.launchParams.frame.subframe_index = idx           # orig: the caller owns the index (increment_subframe = 0)
            idx += step
            r.render()
            if k % 3 == 0 or k >= 18:                                # (AOVs / pixels / counters of the frame just rendered)
                out.append((r.downloadAccum().copy(), r.downloadPixels().copy(), r.launchParams.frame.subframe_index, r.counters().radiance_rays))
        r.close()
        return out

    a, b = run(1), run(3)
    assert len(a) == len(b)
    for (acc1, px1, i1, n1), (acc3, px3, i3, n3) in zip(a, b):
        assert i1 == i3 and n1 == n3
        assert np.array_equal(acc1, acc3) and np.array_equal(px1, px3)


def test_tonemap_sqrt_exact():
    g = golden("display")
    rgba = np.concatenate([g["rgb"], np.ones((len(g["rgb"]), 1), np.float32)], 1).reshape(1, -1, 4)
    r = SampleRenderer(scenes.cornell())
    out = computeFinalPixelColors(r, rgba)
    assert np.array_equal(out.ravel(), g["sqrt"])
    assert computeFinalPixelColors(r, np.zeros((0, 0, 4), np.float32)).size == 0
    r.close()


def test_display_transform_against_golden():
    """make_color / Reinhard of the finalize kernel: powf differs by <= 2 ulp from glibc, so allow 1 LSB."""
    g = golden("display")
    # feed colours through the ACCUM_SUM resolve path: accum <- rgb, n_passes = 1
    n = len(g["rgb"])
    model = scenes.cornell()
    rgba = np.concatenate([g["rgb"], np.ones((n, 1), np.float32)], 1)
    for mode, key, expo in ((0, "srgb", 1.0), (1, "reinhard4", 4.0), (1, "reinhard8", 8.0)):
        cfg = preset("sv4_uniform")
        cfg.display, cfg.exposure = mode, expo
        r = SampleRenderer(model, config=cfg)
        r.resize((n, 1))
        r.uploadAccum(rgba)
        r.resolveSum(1, cfg)
        px = r.downloadPixels().ravel()
        d8 = np.abs(px.view(np.uint8).astype(int) - g[key].view(np.uint8).astype(int))
        assert d8.max() <= 1 and (d8 > 0).mean() < 0.02, key
        r.close()


def test_error_behaviour():
    from optixpathtracer_b200 import _lib
    r = SampleRenderer(scenes.cornell())
    r.render()                                            # size 0: silent no-op like the reference (:81)
    r.resize((0, 10))                                     # ignored (:233)
    assert r.launchParams.frame.size.x == 0
    r.resize((32, 32))
    r.setCamera(preset_camera("cornell"))
    with pytest.raises(_lib.B200ptError):
        r.render()                                        # no probe yet
    bad = ProbeData(np.ones((4, 8, 4), np.float32))
    with pytest.raises(RuntimeError):
        r.setProbe(bad)                                   # Probe.h:104-105 "Probe Data is not valid"
    r.setProbe(bad.BuildCDF())
    r.render()
    assert r.downloadPixels().shape == (32, 32)
    r.close()


# ------------------------------------------------------------------------------------------------ multi-GPU modes (§8e)
def test_tile_split_is_bit_identical_to_full_frame(sponza_small):
    """Two tile-split contexts (as two ranks would run them) assemble to exactly the single-GPU frame."""
    from optixpathtracer_b200 import multi
    probe = analytic_sky(64, 32)
    size, cam = (96, 100), preset_camera("sponza")
    full = make_renderer(sponza_small, probe, size, cam, preset("sv3_uniform", uniform_spl=2))
    full.render()
    want_px, want_acc = full.downloadPixels().copy(), full.downloadAccum().copy()
    full.close()
    got_px, got_acc = np.zeros_like(want_px), np.zeros_like(want_acc)
    for rank in range(2):
        r = make_renderer(sponza_small, probe, size, cam, preset("sv3_uniform", uniform_spl=2, tile_rank=rank, tile_count=2))
        r.clearAccum()
        r.render()
        rows = multi.band_rows(rank, 2, size[1])
        got_px[rows], got_acc[rows] = r.downloadPixels()[rows], r.downloadAccum()[rows]
        r.close()
    assert np.array_equal(got_px, want_px) and np.array_equal(got_acc.view(np.uint32), want_acc.view(np.uint32))


def test_sample_split_sum_mode_matches_running_mean(sponza_small):
    """ACCUM_SUM over sub-frames {0,2} + {1,3} on two contexts, summed and resolved == the 4-frame running mean (fp32 order only)."""
    from optixpathtracer_b200.structs import ACCUM_SUM
    probe = analytic_sky(64, 32)
    size, cam = (96, 54), preset_camera("sponza")
    ref = make_renderer(sponza_small, probe, size, cam, preset("sv3_uniform", uniform_spl=2))
    for _ in range(4):
        ref.render()
    want = ref.downloadAccum()
    ref.close()
    total = np.zeros_like(want)
    for rank in range(2):
        r = make_renderer(sponza_small, probe, size, cam, preset("sv3_uniform", uniform_spl=2, accum_mode=ACCUM_SUM))
        r.clearAccum()
        for f in (rank, rank + 2):
            r.launchParams.frame.subframe_index = f
            r.render()
        total += r.downloadAccum()
        r.close()
    assert np.array_equal(total[..., 3], np.full(total.shape[:2], 4.0, np.float32))
    res = make_renderer(sponza_small, probe, size, cam, preset("sv3_uniform", uniform_spl=2, accum_mode=ACCUM_SUM))
    res.uploadAccum(total)
    res.resolveSum(4)
    got = res.downloadAccum()
    res.close()
    assert rel_mismatch(got[..., :3], want[..., :3], 2e-5) == 0.0


def test_russian_roulette_opt_in_matches_oracle_rule(port, sponza_small):
    """cfg.russian_roulette (north_star extension, off by default): the CUDA path applies the same rule with the same
    extra Randf draw as the oracle, so the two still agree path by path; and it really terminates paths early."""
    probe = analytic_sky(256, 128)
    cam = preset_camera("sponza")
    out = {}
    for rr in (0, 1):
        cfg = preset("orig", uniform_spl=2, russian_roulette=rr)
        r = make_renderer(sponza_small, probe, (160, 90), cam, cfg)
        r.launchParams.samples_per_launch = 2
        r.render()
        got, cnt = r.downloadAccum(), r.counters()
        r.close()
        p = host_params(160, 90, cam, samples_per_launch=2)
        want, _, ocnt, _ = port.scene(sponza_small, probe).render(p, cfg, np.zeros((90, 160, 4), np.float32))
        assert rel_mismatch(got[..., :3], want[..., :3], 2e-4) < 1e-2
        out[rr] = (cnt.radiance_rays, ocnt["radiance_rays"])
    assert out[1][0] < 0.9 * out[0][0] and out[1][1] < 0.9 * out[0][1]


# ------------------------------------------------------------------------------------------------ probe tables on the device
def test_device_build_cdf_is_bit_identical_to_sequential_host(port):
    """b200pt_set_probe_image builds ProbeData::BuildCDF's four tables on the GPU (sv4/Probe.h:29-77).  Bar: bit-exact against
    (a) the golden tables minted from the reference's own Probe.h, (b) the oracle's sequential loop on ragged and large sizes."""
    r = SampleRenderer(scenes.cornell())
    g = golden("probe")
    r.setProbeImage(g["data"])
    for got, name in zip(r.downloadProbe(), ("pdfX", "cdfX", "pdfY", "cdfY")):
        assert np.array_equal(got.view(np.uint32).ravel(), g[name].view(np.uint32).ravel()), name
    rng = np.random.default_rng(77)
    cases = [rng.uniform(0.0, 50.0, (h, w, 4)).astype(np.float32) for (w, h) in ((1, 1), (33, 5), (257, 63), (1000, 37))]
    cases.append(analytic_sky(2048, 1024).data)
    cases.append(loadColor((1.0, 1.0, 1.0), 3840, 2160).data)              # the reference's loadColor size (main.cpp:167-180)
    for data in cases:
        r.setProbeImage(data)
        want = port.build_cdf(data)
        for got, w_, name in zip(r.downloadProbe(), want, ("pdfX", "cdfX", "pdfY", "cdfY")):
            assert np.array_equal(got.view(np.uint32).ravel(), np.asarray(w_).view(np.uint32).ravel()), (data.shape, name)
    # a frame rendered with the device-built probe equals one rendered with the host-built probe
    sky = analytic_sky(256, 128)
    cfg = preset("orig")
    cam = preset_camera("cornell")
    a = make_renderer(scenes.cornell(), sky, (64, 64), cam, cfg)
    a.launchParams.samples_per_launch = 2
    a.render()
    b = SampleRenderer(scenes.cornell(), device=0, config=cfg)
    b.resize((64, 64));  b.setCamera(cam);  b.setProbeImage(sky.data)
    b.launchParams.samples_per_launch = 2
    b.render()
    assert np.array_equal(a.downloadAccum(), b.downloadAccum())
    with pytest.raises(RuntimeError):
        r.setProbeImage(np.zeros((4, 4, 3), np.float32))
    for x in (r, a, b):
        x.close()


def test_mis_on_miss_opt_in_matches_oracle_rule(port, sponza_small):
    """cfg.mis_on_miss (opt-in, off by default; sv4/deviceProgram.cu:257-267): same ProbePdf / weight / ProbeEval arithmetic as the
    oracle, path by path, and it adds the sky energy that BSDF-sampled rays carry."""
    probe = analytic_sky(256, 128)
    cam = preset_camera("sponza")
    means = {}
    for mis in (0, 1):
        cfg = preset("orig", uniform_spl=2, mis_on_miss=mis)
        r = make_renderer(sponza_small, probe, (160, 90), cam, cfg)
        r.launchParams.samples_per_launch = 2
        r.render()
        got = r.downloadAccum()
        r.close()
        p = host_params(160, 90, cam, samples_per_launch=2)
        want, _, _, _ = port.scene(sponza_small, probe).render(p, cfg, np.zeros((90, 160, 4), np.float32))
        assert rel_mismatch(got[..., :3], want[..., :3], 2e-4) < 1e-2
        means[mis] = float(got[..., :3].mean())
    assert means[1] > 1.02 * means[0]


# ------------------------------------------------------------------------------------------------ second integrator
@pytest.mark.parametrize("smooth", [False, True])
def test_whitted_frame_against_oracle(port, smooth):
    """b200pt_render_whitted = the 04HelloRaytracing frame (deviceProgram.cu:66-243) on the production trace kernels: primary ray,
    interpolated normal, textured diffuse colour, one any-hit shadow ray to the point light.  Bar: RGBA8 identical to the CPU
    restatement except where a ray/triangle tie or a grazing shadow ray decides differently (< 0.1 % of pixels)."""
    model = scenes.sponza_class(30000)
    if smooth:
        scenes.add_smooth_normals(model)
    W, H = 320, 180
    cam = preset_camera("sponza")
    r = SampleRenderer(model, device=0, config=preset("orig"))
    r.resize((W, H))
    r.renderWhitted(cam)
    got = r.downloadPixels()
    cnt = r.counters()
    from optixpathtracer_b200.renderer import WHITTED_LIGHT_POS, whitted_camera
    want = port.scene(model, None).render_whitted(W, H, whitted_camera(cam, W, H), WHITTED_LIGHT_POS)
    same = got == want
    assert same.mean() > 0.999, same.mean()
    diff = np.abs(got.view(np.uint8).astype(int) - want.view(np.uint8).astype(int)).reshape(H, W, 4).max(-1)
    assert (diff > 1).mean() < 1e-3
    hit = (want & 0xffffff) != 0
    assert 0.5 < hit.mean() and cnt.radiance_rays == W * H and 0 < cnt.shadow_rays <= W * H
    lit = ((want >> 8) & 0xff)[hit]
    assert lit.max() > 100 and np.unique(want).size > 500            # shaded, shadowed and textured values all occur
    # a frame into a caller-owned device buffer is not covered here (needs a device allocator); the frame buffer path is
    r.close()


def test_whitted_requires_resize_and_accel():
    r = SampleRenderer(scenes.cornell())
    with pytest.raises(RuntimeError):
        r.renderWhitted(preset_camera("cornell"))                       # no resize yet
    r.close()


# ------------------------------------------------------------------------------------------------ animated geometry
def test_refit_and_rebuild_after_vertex_update(port):
    """b200pt_update_vertices + b200pt_refit_accel / b200pt_build_accel.  Bar: after deforming every mesh, rays traced through the
    refitted tree and through a rebuilt tree both give the hits of an oracle built on the deformed scene (prim ids >= 99.99 %, t
    bit-equal on agreeing rays — hits are tree-independent), every refitted box still contains its triangles, and a rendered
    frame is bit-identical between refit and rebuild."""
    import copy
    base = scenes.sponza_class(30000)
    r = SampleRenderer(base, device=0, config=preset("orig"))
    moved = copy.deepcopy(base)
    for k, mesh in enumerate(moved.meshes):
        v = np.asarray(mesh.vertex, np.float32).copy()
        v[:, 1] += (40.0 * np.sin(v[:, 0] * 0.01 + 0.3 * k)).astype(np.float32)      # a travelling wave
        v[:, 0] += (25.0 * np.cos(v[:, 2] * 0.02)).astype(np.float32)
        mesh.vertex = v
        r.updateVertices(k, v)
    rng = np.random.default_rng(5)
    n = 20000
    org = (rng.uniform(-1, 1, (n, 3)) * np.array([1800, 600, 700]) + np.array([0, 650, 0])).astype(np.float32)
    d = rng.normal(size=(n, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    oprim, omesh, ot, _, _ = port.scene(moved).trace_rays(org, d, 0.01, 1e16)
    probe = analytic_sky(128, 64)
    cam = preset_camera("sponza")
    frames = {}
    for how in ("refit", "rebuild"):
        r.refitAccel() if how == "refit" else r.rebuildAccel()
        info = r.bvhInfo()
        assert (info.refit_ms > 0) == (how == "refit")
        prim, mesh, t, _, _ = r.traceRays(org, d, 0.01, 1e16)
        same = (prim == oprim) & (mesh == omesh)
        assert same.mean() >= 0.9999, (how, same.mean())
        assert np.array_equal(t[same].view(np.uint32), ot[same].view(np.uint32))
        check_bvh_boxes_contain_triangles(r, moved)
        r.resize((96, 54));  r.setCamera(cam);  r.setProbe(probe)
        r.launchParams.samples_per_launch = 2
        r.launchParams.frame.subframe_index = 0
        r.render()
        frames[how] = r.downloadAccum()
    assert np.array_equal(frames["refit"], frames["rebuild"])
    with pytest.raises(RuntimeError):
        r.updateVertices(0, np.zeros((3, 3), np.float32))               # wrong vertex count
    with pytest.raises(RuntimeError):
        r.updateVertices(10 ** 6, np.zeros((3, 3), np.float32))
    r.close()


def test_async_download_after_join_is_ordered(sponza_small):
    """render_async, then an entry point that joins the lanes (b200pt_join / counters), THEN b200pt_download_pixels_async: the copy
    is new work on the frame's lane, and b200pt_synchronize must wait for it ('valid after b200pt_synchronize')."""
    import torch
    probe = analytic_sky(128, 64)
    cfg = preset("orig", uniform_spl=1)
    cfg.frames_in_flight = 3
    r = make_renderer(sponza_small, probe, (640, 360), preset_camera("sponza"), cfg)
    host = torch.zeros((360, 640), dtype=torch.int32).pin_memory().numpy().view(np.uint32)
    for k in range(5):
        r.launchParams.frame.subframe_index = k
        r.render(sync=False)
        r.join()                                         # clears the lanes' busy flags
        host[:] = 0
        r.downloadPixelsAsync(host)
        r.synchronize()
        assert np.array_equal(host, r.downloadPixels()), k
        assert (host >> 24 == 255).all()
    r.close()


def test_speculative_wavefront_dropped_before_refit(port):
    """a blocking render() that has speculative next sub-frames in flight, then a camera move (mis-prediction: the wavefronts are
    abandoned but keep running), then IMMEDIATELY update_vertices + refit + render: the mutating entry points must drain every lane first.
    The frame after the refit equals the frame of a renderer that never speculated."""
    import copy
    base = scenes.sponza_class(30000)
    probe = analytic_sky(128, 64)
    cam, cam2 = preset_camera("sponza"), preset_camera("sponza")
    cam2.eye = (cam2.eye[0] + 3.0, cam2.eye[1], cam2.eye[2])
    moved = [np.asarray(m.vertex, np.float32) + np.float32(7.5) for m in base.meshes]

    def run(in_flight):
        cfg = preset("orig", uniform_spl=1)
        cfg.frames_in_flight = in_flight
        r = make_renderer(copy.deepcopy(base), probe, (320, 180), cam, cfg)
        for k in range(6):                               # progressive accumulation: from the third call on the next sub-frames are speculated
            r.launchParams.frame.subframe_index = k
            r.render()
        r.setCamera(cam2)                                # wrong guess for the speculative wavefronts: abandoned, still running
        r.launchParams.frame.subframe_index = 0
        r.render()
        for k, v in enumerate(moved):                    # scene changes right behind it
            r.updateVertices(k, v)
        r.refitAccel()
        r.launchParams.frame.subframe_index = 1
        r.render()
        out = r.downloadAccum().copy()
        r.close()
        return out

    assert np.array_equal(run(1), run(3))


@pytest.mark.parametrize("pname", ["orig", "sv4_fov"])
def test_intra_frame_split_is_bit_identical(sponza_small, pname):
    """cfg.intra_frame_split = 2..4 (one frame at a time, run as concurrent sets of 32-row bands on separate lanes): accumulation buffer, pixels, AOVs and the
    frame's counters equal the unsplit frame's, also across accumulating sub-frames and for the foveated three-pass schedule."""
    probe = analytic_sky(128, 64)
    cam = preset_camera("sponza")

    def run(split):
        cfg = preset(pname)
        if pname == "orig":
            cfg.uniform_spl = 2
        cfg.frames_in_flight, cfg.intra_frame_split = 1, split
        r = make_renderer(sponza_small, probe, (200, 150), cam, cfg)
        if pname == "sv4_fov":
            r.launchParams.frame.c.x, r.launchParams.frame.c.y = 100, 75
            r.config.inner_radius, r.config.outer_radius = 20, 50
        cnts = []
        for k in range(3):
            if pname == "orig":
                r.launchParams.frame.subframe_index = k
            r.render()
            c = r.counters()
            cnts.append((c.samples, c.radiance_rays, c.shadow_rays))
        out = (r.downloadAccum().copy(), r.downloadPixels().copy(), r.downloadAOV("normal").copy() if pname == "orig" else None, cnts)
        r.close()
        return out

    want = run(0)
    for split in (2, 3, 4):
        got = run(split)
        assert np.array_equal(got[0].view(np.uint32), want[0].view(np.uint32)) and np.array_equal(got[1], want[1]), split
        assert got[3] == want[3], split
        if pname == "orig":
            assert np.array_equal(got[2], want[2])


def test_collectives_on_a_one_rank_communicator(sponza_small):
    """The library-side NCCL path on the ranks the test box has (one): b200pt_comm_init moves the accumulation buffer into ncclMemAlloc'd memory with its
    contents and launchParams follows it; b200pt_reduce_sum (in place) equals b200pt_resolve_sum; keep_local leaves the partial sum untouched and only
    refreshes the frame buffer; b200pt_gather_bands is a no-op; b200pt_comm_destroy keeps the buffer.  N > 1: tools/multi_check.py, bench.py --gpus N."""
    from optixpathtracer_b200.structs import ACCUM_SUM
    probe = analytic_sky(128, 64)
    cam = preset_camera("sponza")

    def fresh():
        cfg = preset("orig", uniform_spl=1)
        cfg.accumulate, cfg.accum_mode, cfg.frames_in_flight = 1, ACCUM_SUM, 3
        r = make_renderer(sponza_small, probe, (160, 96), cam, cfg)
        r.clearAccum()
        for k in range(4):
            r.launchParams.frame.subframe_index = k
            r.render(sync=False)
        r.synchronize()
        return r

    a = fresh()
    sums = a.downloadAccum().copy()
    a.resolveSum(4)
    want_acc, want_px = a.downloadAccum().copy(), a.downloadPixels().copy()
    a.close()

    b = fresh()
    before_ptr = b.launchParams.frame.accum_buffer
    b.commInit(1, 0, SampleRenderer.commUniqueId())
    info = b.commInfo()
    assert info["n_ranks"] == 1 and info["rank"] == 0 and info["nccl_version"] >= 21900
    if info["accum_nccl_alloc"]:
        assert b.launchParams.frame.accum_buffer != before_ptr                      # the buffer moved and launchParams followed it
    assert np.array_equal(b.downloadAccum().view(np.uint32), sums.view(np.uint32))  # ... with its contents
    b.reduceSum(4, root=0, keep_local=True)                                         # progressive display: sums stay, frame refreshed
    b.synchronize()
    assert np.array_equal(b.downloadAccum().view(np.uint32), sums.view(np.uint32))
    assert np.array_equal(b.downloadPixels(), want_px)
    b.gatherBands(root=0, frame=True, accum=True)                                   # one rank owns every band
    b.reduceSum(4, root=0)                                                          # end of job: in place, accum = mean
    b.synchronize()
    assert np.array_equal(b.downloadAccum().view(np.uint32), want_acc.view(np.uint32)) and np.array_equal(b.downloadPixels(), want_px)
    b.commDestroy()
    assert b.commInfo()["n_ranks"] == 1 and not b.commInfo()["accum_nccl_alloc"]
    assert np.array_equal(b.downloadAccum().view(np.uint32), want_acc.view(np.uint32))
    with pytest.raises(RuntimeError):
        b.reduceSum(4)                                                              # no communicator any more
    b.launchParams.frame.subframe_index = 0
    b.render()                                                                      # still renders into a valid buffer
    assert np.isfinite(b.downloadAccum()).all()
    b.close()


def check_bvh_boxes_contain_triangles(r, model):
    """Every leaf's triangles lie inside the (dequantised) child box that refers to them, and child boxes nest in their parents'."""
    info = r.bvhInfo()
    nodes, gid = r.downloadBvh()
    lo0, sc = np.array(info.grid_lo), np.array(info.grid_scale)
    tri_lo, tri_hi = [], []
    for mesh in model.meshes:
        v = np.asarray(mesh.vertex, np.float64)[np.asarray(mesh.index, np.int64)]
        tri_lo.append(v.min(1));  tri_hi.append(v.max(1))
    tri_lo, tri_hi = np.concatenate(tri_lo), np.concatenate(tri_hi)
    q = nodes[:, :6].astype(np.int64).reshape(-1, 2, 3)
    blo = (q & 0xffff).astype(np.float64)
    bhi = (q >> 16).astype(np.float64)
    wlo = (blo - 2.0) / sc + lo0                                          # world-space child boxes, (records, child, axis)
    whi = (bhi - 2.0) / sc + lo0
    links = nodes[:, 6:8].view(np.int32)
    for k in range(2):
        leaf = links[:, k] < 0
        code = -(links[leaf, k].astype(np.int64) + 1)
        first, count = code >> 3, (code & 7) + 1
        for j in range(8):
            sel = count > j
            g = gid[first[sel] + j]
            assert np.all(tri_lo[g] >= wlo[leaf][sel][:, k] - 1e-9) and np.all(tri_hi[g] <= whi[leaf][sel][:, k] + 1e-9)
        inner = ~leaf
        ch = links[inner, k]
        assert np.all(np.minimum(blo[ch, 0], blo[ch, 1]) >= blo[inner][:, k] - 4) and np.all(np.maximum(bhi[ch, 0], bhi[ch, 1]) <= bhi[inner][:, k] + 4)


def test_one_sided_tile_split_two_contexts_one_device(sponza_small):
    """b200pt_peer_*: context B (tile_rank 1 of 2) attaches to the frame context A (tile_rank 0) exported and its finalize pass writes its bands
    straight into A's frame buffer; A acquires, downloads, releases.  Three progressive frames with frames in flight: every acquired frame equals
    the unsplit single-context frame byte for byte, and nothing times out.  (Same-process path: raw pointers; two processes / two GPUs go through
    CUDA IPC - tools/multi_check.py, bench.py --gpus N.)"""
    probe = analytic_sky(64, 32)
    size, cam = (96, 100), preset_camera("sponza")
    full = make_renderer(sponza_small, probe, size, cam, preset("sv3_uniform", uniform_spl=2))
    want = []
    for i in range(3):
        full.launchParams.frame.subframe_index = i
        full.render()
        want.append(full.downloadPixels().copy())
    full.close()
    ctx = []
    for rank in range(2):
        cfg = preset("sv3_uniform", uniform_spl=2, tile_rank=rank, tile_count=2)
        cfg.frames_in_flight = 3
        r = make_renderer(sponza_small, probe, size, cam, cfg)
        r.clearAccum()
        ctx.append(r)
    a, b = ctx
    with pytest.raises(Exception, match="nothing exported"):
        a.peerAcquire()
    handle = a.peerExport(0, 2)
    with pytest.raises(Exception, match="root does not attach"):
        a.peerAttach(handle, 1, 2)
    with pytest.raises(Exception, match="not a handle"):
        b.peerAttach(np.zeros(192, np.uint8), 1, 2)
    b.peerAttach(handle, 1, 2)
    with pytest.raises(Exception, match="no acquired frame"):
        a.peerRelease()
    for i in range(3):
        for r in ctx:
            r.launchParams.frame.subframe_index = i
        b.render(sync=False)                                   # B runs ahead; its frame i + 1 must not land before A released frame i
        a.render(sync=False)
        a.peerAcquire()
        got = a.downloadPixels().copy()
        a.peerRelease()
        assert np.array_equal(got, want[i]), f"frame {i}"
    assert a.peerStatus() == dict(state=0, attached_rank=-1, epoch=3)
    assert b.peerStatus() == dict(state=0, attached_rank=1, epoch=3)
    # B's own frame buffer was not written (its bands went to A), and after a detach it is again
    assert not b.downloadPixels().any()
    b.peerDetach()
    b.launchParams.frame.subframe_index = 3
    b.render()
    assert b.downloadPixels().any()
    for r in ctx:
        r.close()


def test_one_sided_tile_split_times_out_instead_of_hanging(sponza_small, monkeypatch):
    """A root that acquires a frame nobody puts gives up after B200PT_PEER_TIMEOUT_MS and reports which rank it waited for."""
    monkeypatch.setenv("B200PT_PEER_TIMEOUT_MS", "100")
    probe = analytic_sky(64, 32)
    cfg = preset("sv3_uniform", uniform_spl=1, tile_rank=0, tile_count=3)
    a = make_renderer(sponza_small, probe, (64, 64), preset_camera("sponza"), cfg)
    a.peerExport(0, 3)
    a.render(sync=False)
    a.peerAcquire()
    st = a.peerStatus()
    assert st["state"] in (2, 3) and st["epoch"] == 1                  # 1 + r for r = 1 or 2, whichever gave up first
    a.close()


def test_sum_mode_blocking_render_with_speculation_is_bit_identical(sponza_small):
    """One rank of a sample split as bench.py's N > 1 end-to-end loop drives it: ACCUM_SUM, sub-frames striding by the rank count, BLOCKING render() with
    frames in flight (the predicted next sub-frames run while the caller reduces / downloads).  The SUM buffer after every call and the resolved image equal
    those of strictly one-at-a-time calls bit for bit."""
    from optixpathtracer_b200.structs import ACCUM_SUM
    probe, cam = analytic_sky(128, 64), preset_camera("sponza")

    def run(in_flight):
        cfg = preset("orig", uniform_spl=1)
        cfg.accumulate, cfg.accum_mode, cfg.frames_in_flight = 1, ACCUM_SUM, in_flight
        r = make_renderer(sponza_small, probe, (160, 96), cam, cfg)
        r.clearAccum()
        out = []
        for k in range(8):
            r.setCamera(cam)
            r.launchParams.frame.subframe_index = 1 + 2 * k               # rank 1 of 2
            r.render()
            out.append(r.downloadAccum().copy())
        r.resolveSum(8)
        out.append(r.downloadAccum().copy())
        px = r.downloadPixels().copy()
        r.close()
        return out, px

    want, want_px = run(1)
    got, got_px = run(3)
    for k, (a, b) in enumerate(zip(got, want)):
        assert np.array_equal(a.view(np.uint32), b.view(np.uint32)), f"call {k}"
    assert np.array_equal(got_px, want_px)


def test_one_sided_tile_split_foveated_pass_barriers(sponza_small):
    """The sv4 foveated schedule (three passes, a later pass overwrites pixels of an earlier one) split over two contexts that write ONE frame: every rank
    waits after pass k until all ranks have finished pass k (epoch flags in the exporting context's memory), so the frame root acquires equals the unsplit
    frame byte for byte.  Gaze moves every frame and subframe_index stays 0 (the interactive mode: the frame does not read the accumulation buffer)."""
    probe = analytic_sky(128, 64)
    cam = preset_camera("sponza")
    W, H = 256, 160
    gazes = ((128, 80), (100, 70), (150, 90))                 # every ring inside the frame: no border-clamp write races

    def fov(rank, count):
        cfg = preset("sv4_fov", inner_radius=20, outer_radius=60, tile_rank=rank, tile_count=count)
        r = make_renderer(sponza_small, probe, (W, H), cam, cfg)
        r.clearAccum()
        return r

    def shoot(r, gaze):
        r.launchParams.frame.c.x, r.launchParams.frame.c.y = gaze
        r.launchParams.frame.subframe_index = 0
        r.render(sync=False)

    full = fov(0, 1)
    want = []
    for g in gazes:
        shoot(full, g)
        want.append(full.downloadPixels().copy())
    full.close()
    a, b = fov(0, 2), fov(1, 2)
    b.peerAttach(a.peerExport(0, 2), 1, 2)
    for k, g in enumerate(gazes):
        shoot(b, g)
        shoot(a, g)
        a.peerAcquire()
        got = a.downloadPixels().copy()
        a.peerRelease()
        assert np.array_equal(got, want[k]), f"frame {k}: {(got != want[k]).sum()} pixels differ"
    assert a.peerStatus()["state"] == 0 and b.peerStatus()["state"] == 0
    a.close()
    b.close()
