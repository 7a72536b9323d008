"""Generate the committed golden vectors from the REFERENCE's own code.

Runs only in the build container (needs /root/reference): loads oracle/_ref/liboracle_ref.so, i.e. the
reference's cuda/random.h, sv4/maths.h, sv4/sample.h, sv4/Disney.cuh, sv4/Probe.cuh, sv4/Probe.h and
cuda/helpers.h host-compiled in place, and records their outputs on seeded inputs.  The fixtures pin
(a) the CPU restatement in oracle/shading_port.h (`-m "not gpu"` tests, bit-exact) and (b) the CUDA kernels
(`-m gpu` tests).  Usage:  python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from optixpathtracer_b200 import Material, host_params, preset, preset_camera, scenes      # noqa: E402
from optixpathtracer_b200.probe import ProbeData                                            # noqa: E402
from oracle import binding                                                                  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def random_materials(rng, n):
    mats = []
    for i in range(n):
        kind = i % 4
        if kind == 0:
            m = Material(color=tuple(rng.uniform(0.05, 1, 3)), emission=(0, 0, 0))
        elif kind == 1:
            m = Material.orig_defaults(color=tuple(rng.uniform(0.05, 1, 3)))
        elif kind == 2:
            m = Material(color=tuple(rng.uniform(0.05, 1, 3)), transmission=1.0, metallic=0.0, roughness=float(rng.uniform(0.02, 0.5)), eta=1.5)
        else:
            m = Material(color=tuple(rng.uniform(0, 1, 3)), metallic=float(rng.uniform()), subsurface=float(rng.uniform()),
                         specular=float(rng.uniform()), roughness=float(rng.uniform(0.0, 1.0)), specularTint=float(rng.uniform()),
                         clearcoat=float(rng.uniform()), clearcoatGloss=float(rng.uniform()), transmission=float(rng.uniform()),
                         eta=float(rng.choice([0.0, 1.33, 1.5])))
        mats.append(m)
    return mats


def unit(v):
    return (v / np.linalg.norm(v, axis=-1, keepdims=True)).astype(np.float32)


def small_probe(rng, w=16, h=8):
    data = rng.uniform(0.05, 4.0, (h, w, 4)).astype(np.float32)
    data[..., 3] = 1.0
    data[2, 5, :3] = 60.0                       # a bright texel ("sun")
    return data


def main():
    binding.build(ref=True)
    ref = binding.Oracle("reference")
    rng = np.random.default_rng(20231104)

    # ---- integer / RNG known answers
    v0 = rng.integers(0, 2**32, 256, dtype=np.uint64).astype(np.uint32)
    v1 = rng.integers(0, 2**32, 256, dtype=np.uint64).astype(np.uint32)
    v0[:4] = [0, 5, 1, 0xffffffff]
    v1[:4] = [0, 0, 1, 0xffffffff]
    lcg_state, lcg_f = ref.lcg_stream(1535087949, 64)
    ru, rstate_u = ref.random_stream(12345, 64, False)
    rf, rstate_f = ref.random_stream(0xdeadbeef, 64, True)
    np.savez_compressed(os.path.join(OUT, "kat_rng.npz"), v0=v0, v1=v1, tea4=ref.tea4(v0, v1), lcg_seed=np.uint32(1535087949),
                        lcg_state=lcg_state, lcg_f=lcg_f, rand_seed_u=np.uint32(12345), rand_u=ru, rand_state_u=rstate_u,
                        rand_seed_f=np.uint32(0xdeadbeef), rand_f=rf, rand_state_f=rstate_f)

    # ---- Disney BSDF
    n = 512
    mats = random_materials(rng, n)
    N = unit(rng.normal(size=(n, 3)))
    view = unit(N + 0.8 * unit(rng.normal(size=(n, 3))))          # mostly above the surface, some below
    L = unit(rng.normal(size=(n, 3)))
    etaI = np.where(rng.uniform(size=n) < 0.7, 1.0, 1.5).astype(np.float32)
    etaO = np.where(etaI == 1.0, 1.5, 1.0).astype(np.float32)
    albedo = rng.uniform(0, 1, (n, 3)).astype(np.float32)
    albedo[:3] = 0.0
    seeds = rng.integers(0, 2**32, n, dtype=np.uint64).astype(np.uint32)
    U, V = ref.basis(N)
    sL, spdf, stype, srng = ref.bsdf_sample(mats, etaI, etaO, N, view, seeds)
    f, pdf = ref.bsdf_eval(mats, albedo, etaI, etaO, N, view, L)
    mat_bytes = np.frombuffer(b"".join(bytes(m) for m in mats), np.uint8).reshape(n, 104)
    np.savez_compressed(os.path.join(OUT, "bsdf.npz"), materials=mat_bytes, N=N, view=view, L=L, etaI=etaI, etaO=etaO, albedo=albedo,
                        seeds=seeds, basisU=U, basisV=V, sampleL=sL, samplePdf=spdf, sampleType=stype, sampleRng=srng, evalF=f, evalPdf=pdf)

    # ---- probe
    data = small_probe(rng)
    pdfX, cdfX, pdfY, cdfY = ref.build_cdf(data)
    probe = ProbeData(data)
    probe.pdfX, probe.cdfX, probe.pdfY, probe.cdfY, probe.valid = pdfX, cdfX, pdfY, cdfY, True
    pseeds = rng.integers(0, 2**32, 512, dtype=np.uint64).astype(np.uint32)
    pdir, pcol, ppdf = ref.probe_sample(probe, pseeds)
    dirs = unit(rng.normal(size=(256, 3)))
    dirs[0] = [0, 1, 0]
    dirs[1] = [0, -1, 0]
    uv, ecol, epdf = ref.probe_eval_dir(probe, dirs)
    np.savez_compressed(os.path.join(OUT, "probe.npz"), data=data, pdfX=pdfX, cdfX=cdfX, pdfY=pdfY, cdfY=cdfY, seeds=pseeds,
                        sampleDir=pdir, sampleColor=pcol, samplePdf=ppdf, dirs=dirs, uv=uv, evalColor=ecol, evalPdf=epdf)

    # ---- display transforms
    rgb = np.concatenate([rng.uniform(0, 2.5, (500, 3)), rng.uniform(0, 0.01, (100, 3)),
                          [[0, 0, 0], [1, 1, 1], [10, 10, 10], [-1, 0.5, 2], [0.0031308, 0.0031307, 0.5]]]).astype(np.float32)
    rgba = np.concatenate([rgb, np.ones((len(rgb), 1), np.float32)], 1)
    np.savez_compressed(os.path.join(OUT, "display.npz"), rgb=rgb, srgb=ref.display(rgb, 0, 1.0), reinhard4=ref.display(rgb, 1, 4.0),
                        reinhard8=ref.display(rgb, 1, 8.0), sqrt=ref.tonemap_sqrt(rgba))

    # ---- camera
    cams = {}
    for name in ("cornell", "sponza", "lost_empire"):
        cam = preset_camera(name)
        cams[name] = np.stack(ref.camera_uvw(cam.eye, cam.lookat, cam.up, float(cam.fovY), 16.0 / 9.0))
    np.savez_compressed(os.path.join(OUT, "camera.npz"), **cams)

    # ---- small end-to-end renders (reference shading headers + oracle traversal contract, brute force)
    probe_c = ProbeData(small_probe(np.random.default_rng(7), 32, 16)).BuildCDF()
    renders = {}
    for variant, pname in (("disney", "sv4_uniform"), ("diffuse", "orig")):
        model = scenes.cornell(variant)
        sc = ref.scene(model, probe_c, brute_force=True)
        cfg = preset(pname, uniform_spl=4)
        p = host_params(48, 48, preset_camera("cornell"), samples_per_launch=4)
        accum = frame = None
        for sub in range(2):
            p.frame.subframe_index = sub
            cfg.accumulate = 1
            accum, frame, cnt, _ = sc.render(p, cfg, accum, frame)
        renders[f"{variant}_accum"], renders[f"{variant}_frame"] = accum, frame
        p0 = host_params(48, 48, preset_camera("cornell"), samples_per_launch=1)
        prim, mesh, t, u, v = sc.trace_primary(p0, cfg)
        renders[f"{variant}_prim"], renders[f"{variant}_mesh"], renders[f"{variant}_t"] = prim, mesh, t
    np.savez_compressed(os.path.join(OUT, "cornell_render.npz"), probe=probe_c.data, **renders)
    print("golden vectors written to", OUT)


if __name__ == "__main__":
    main()
