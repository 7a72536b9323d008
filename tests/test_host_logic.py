"""Host-side mirror of the reference interface: loadOBJ / addBox / scenes / probe / schedule bookkeeping."""
import numpy as np
import pytest

from optixpathtracer_b200 import Material, Model, addBox, host_params, loadColor, loadOBJ, preset, preset_camera, scenes
from optixpathtracer_b200.probe import analytic_sky


def test_cornell_through_loadobj(tmp_path):
    model = scenes.cornell(directory=str(tmp_path))
    assert model.num_triangles() == 32
    assert [len(m.index) for m in model.meshes] == [6, 2, 2, 2, 20]          # one mesh per (shape, material id)
    light = model.meshes[3]
    assert light.material.emission.tolist() == [12.0, 12.0, 12.0]
    # loadOBJ overwrites only color + emission: everything else is the sv4 constructor default (Model.cpp:189-190)
    assert abs(light.material.transmission - 0.4) < 1e-7 and abs(light.material.metallic - 0.5) < 1e-7
    for m in model.meshes:
        assert m.index.max() < len(m.vertex)                                 # vertices de-duplicated per mesh
    assert len(model.meshes[0].vertex) == 12                                 # 3 quads, 4 shared corners each
    with pytest.raises(RuntimeError):
        loadOBJ(str(tmp_path / "missing.obj"))


def test_addbox_matches_reference_layout():
    model = Model()
    mat = Material(color=(0.2, 0.3, 0.4))
    mesh = addBox(model, mat, (1.0, 2.0, 3.0), (0.5, 1.0, 1.5))
    assert mesh.vertex.shape == (36, 3) and mesh.index.shape == (12, 3) and mesh.normal.shape == (36, 3)
    assert np.array_equal(mesh.index.ravel(), np.arange(36))
    assert np.allclose(mesh.vertex.min(0), [0.5, 1.0, 1.5]) and np.allclose(mesh.vertex.max(0), [1.5, 3.0, 4.5])
    assert np.allclose(mesh.vertex[0], [0.5, 1.0, 4.5])                      # A = (-ex+px, -ey+py, ez+pz), Model.cpp:220
    assert bytes(mesh.material) == bytes(mat)


def test_procedural_scene_sizes():
    s = scenes.sponza_class()
    assert abs(s.num_triangles() - 262144) <= 0.011 * 262144
    ext = np.concatenate([m.vertex for m in s.meshes])
    assert ext.min() >= -2000 and ext.max() <= 2000
    assert any(m.diffuseTextureID >= 0 for m in s.meshes) and len(s.textures) >= 1
    le = scenes.lost_empire_class()
    assert abs(le.num_triangles() - 225000) <= 0.011 * 225000
    a, b = scenes.sponza_class(20000), scenes.sponza_class(20000)
    assert all(np.array_equal(x.vertex, y.vertex) for x, y in zip(a.meshes, b.meshes))   # deterministic
    c = scenes.city_class(300000)
    assert abs(c.num_triangles() - 300000) <= 0.05 * 300000


def test_probe_tables():
    p = loadColor((1, 1, 1), 64, 32)
    assert p.valid and p.cdfX.shape == (32, 64)
    assert np.allclose(p.cdfX[:, -1], 1.0, atol=1e-5) and abs(p.cdfY[-1] - 1.0) < 1e-6
    assert np.all(np.diff(p.cdfY) >= 0) and np.all(np.diff(p.cdfX, axis=1) >= 0)
    sky = analytic_sky(128, 64)
    assert sky.data.shape == (64, 128, 4) and np.isfinite(sky.data).all() and sky.data.max() > 10


def test_oracle_schedule_mutates_params_like_the_reference(port):
    """SampleRenderer::render() leaves pass-3 values in the public launchParams (sv4/SimplePathtracer.cpp:189-212)."""
    model = scenes.cornell()
    probe = loadColor((1, 1, 1), 16, 8)
    sc = port.scene(model, probe, brute_force=True)
    p = host_params(64, 64, preset_camera("cornell"))
    p.frame.c.x, p.frame.c.y = 32, 32
    cfg = preset("sv4_fov", inner_radius=8, outer_radius=20)
    p.frame.subframe_index = 5
    accum, frame, cnt, _ = sc.render(p, cfg)
    assert p.frame.subframe_index == 6
    assert (p.frame.fillSize, p.samples_per_launch, p.frame.redraw) == (1, 8, 1)
    assert (p.frame.offset.x, p.frame.offset.y) == (32 - 9, 32 - 9) and p.frame.r_outer == 9.0
    # budget: periphery 1 spp per 4x4 block, mid ring 2 spp per 2x2, fovea 8 spp per pixel
    assert 0 < cnt["samples"] < 64 * 64 * 8
    cfg_u = preset("sv4_uniform")
    p.frame.subframe_index = 7
    sc.render(p, cfg_u)
    assert p.frame.subframe_index == 1 and p.samples_per_launch == 8          # FOV_OFF forces subframe 0, then ++


def test_russian_roulette_is_unbiased_in_the_oracle(port):
    """The opt-in roulette (oracle/render_core.h, mirrored in csrc/shade.cu) must not change the expected image."""
    from optixpathtracer_b200 import host_params, loadColor, preset, preset_camera, scenes
    import numpy as np
    model = scenes.cornell("diffuse")
    sc = port.scene(model, loadColor((1.0, 1.0, 1.0), 16, 8), brute_force=True)
    means = []
    for rr in (0, 1):
        cfg = preset("orig", uniform_spl=256, russian_roulette=rr)
        acc, _, cnt, _ = sc.render(host_params(24, 24, preset_camera("cornell"), samples_per_launch=256), cfg)
        means.append((acc[..., :3].mean(), cnt["radiance_rays"]))
    assert means[1][1] < means[0][1]
    assert abs(means[1][0] - means[0][0]) < 0.02 * means[0][0]


def test_mis_on_miss_completes_the_estimator(port):
    """cfg.mis_on_miss (opt-in; the reference's commented-out block, sv4/deviceProgram.cu:257-267).  An open floor under a constant
    sky has a closed form: L = integral of f * cos over the upper hemisphere.  NEE alone carries only its MIS weight (the reference's
    live path, biased dark); with the miss term the two weights sum to one and the estimate matches a quadrature of BSDFEval."""
    import numpy as np
    from optixpathtracer_b200 import Material, Model, TriangleMesh, Camera, host_params, loadColor, preset
    mat = Material(color=(0.7, 0.6, 0.5), emission=(0, 0, 0), metallic=0.0, transmission=0.0, subsurface=0.0, specular=0.5,
                   roughness=0.1, clearcoat=0.0, specularTint=0.0, eta=1.5)
    mesh = TriangleMesh()
    mesh.vertex = np.array([[-500, 0, -500], [500, 0, -500], [500, 0, 500], [-500, 0, 500]], np.float32)
    mesh.index = np.array([[0, 2, 1], [0, 3, 2]], np.uint32)                 # normal +y
    mesh.material = mat
    model = Model()
    model.meshes.append(mesh)
    # view along the normal and a narrow lobe: BSDF samples below the horizon (which end the path with pdf 0 and, as in the
    # reference, take that vertex's NEE term with them, :679-682) stay under 1 %
    eye = np.array([0.0, 5.0, 0.0])
    cam = Camera(tuple(eye), (0.0, 0.0, 0.0), (0.0, 0.0, 1.0), 0.5, 1.0)      # half a degree: one view direction for all pixels
    sc = port.scene(model, loadColor((1.0, 1.0, 1.0), 512, 256), brute_force=True)
    means = {}
    for mis in (0, 1):
        cfg = preset("orig", uniform_spl=512, mis_on_miss=mis)
        acc, _, _, _ = sc.render(host_params(16, 16, cam, samples_per_launch=512), cfg)
        means[mis] = acc[..., :3].reshape(-1, 3).mean(0)
    # quadrature on a theta-phi grid (fine in theta: the specular lobe is a few degrees wide)
    n_th, n_phi = 4000, 64
    th = (np.arange(n_th) + 0.5) / n_th * (np.pi / 2)
    phi = (np.arange(n_phi) + 0.5) / n_phi * 2 * np.pi
    th, phi = np.meshgrid(th, phi, indexing="ij")
    mu, s = np.cos(th), np.sin(th)
    L = np.stack([s * np.cos(phi), mu, s * np.sin(phi)], -1).reshape(-1, 3).astype(np.float32)
    n = len(L)
    V = np.tile((eye / np.linalg.norm(eye)).astype(np.float32), (n, 1))
    N = np.tile(np.array([0, 1, 0], np.float32), (n, 1))
    f, _ = port.bsdf_eval([mat] * n, np.tile(np.array(mat_color(mat), np.float32), (n, 1)), np.ones(n, np.float32),
                          np.full(n, 1.5, np.float32), N, V, L)
    want = (f * (mu * s).reshape(-1, 1)).mean(0) * (np.pi / 2) * (2 * np.pi)
    assert np.all(np.abs(means[1] - want) < 0.02 * want), (means, want)
    assert np.all(means[0] < 0.9 * want)                                      # NEE alone misses a visible share of the energy


def mat_color(m):
    return (m.color.x, m.color.y, m.color.z)


def test_bench_reference_arm_prints_one_contract_line():
    """`bench.py --impl reference` (the CPU arm: oracle on the host cores) prints exactly one JSON line with the contract's keys; tiny frame, no GPU."""
    import json
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0", "--width", "96", "--height", "54",
                          "--tris", "3000"], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "Msamples/s" and d["unit"] == "Msamples/s" and d["higher_is_better"] is True
    assert d["value"] > 0 and d["cpu_baseline"]["kind"] in ("reference", "port") and d["cpu_baseline"]["cores"] >= 1
    assert d["e2e"] == {"value": d["value"], "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["gpu_launches"] == 0 and "workload" in d["config"]
