"""ORACLE — TEST INFRASTRUCTURE, NOT PRODUCT CODE.

ctypes loader for the CPU oracle (oracle/liboracle_port.so, oracle/_ref/liboracle_ref.so).
Only tests/, __graft_entry__.smoke() and bench.py (cpu_baseline leg, --impl reference) may import this.
"""
import ctypes as C
import os
import subprocess

import numpy as np

from optixpathtracer_b200.structs import Config, LaunchParams, Material

_HERE = os.path.dirname(os.path.abspath(__file__))
PORT_SO = os.path.join(_HERE, "liboracle_port.so")
REF_SO = os.path.join(_HERE, "_ref", "liboracle_ref.so")


def build(ref=None):
    """Compile the port (always) and the reference-header backend (when /root/reference exists)."""
    subprocess.check_call(["make", "-s", "-C", _HERE, "port"])
    if ref is None:
        ref = os.path.isdir("/root/reference")
    if ref:
        subprocess.check_call(["make", "-s", "-C", _HERE, "ref"])


def _p(a, t=None):
    if a is None:
        return None
    return a.ctypes.data_as(C.c_void_p)


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _u32(a):
    return np.ascontiguousarray(a, dtype=np.uint32)


class Oracle:
    """One loaded oracle backend ('port' or 'reference')."""

    def __init__(self, kind="port"):
        path = PORT_SO if kind == "port" else REF_SO
        if not os.path.exists(path):
            if kind == "port":
                build(ref=False)
            else:
                raise FileNotFoundError(path)
        self.lib = L = C.CDLL(path)
        L.oracle_backend.restype = C.c_char_p
        L.oracle_scene_create.restype = C.c_void_p
        L.oracle_render.restype = C.c_double
        L.oracle_render_rows.restype = C.c_double
        L.oracle_scene_num_triangles.restype = C.c_uint64
        self.kind = L.oracle_backend().decode()
        assert self.kind == kind, (self.kind, kind)

    # ---- info
    def max_threads(self):
        return self.lib.oracle_max_threads()

    def set_threads(self, n):
        self.lib.oracle_set_threads(int(n))

    def sizeof(self, which):
        return self.lib.oracle_sizeof(which)

    def default_material(self):
        m = Material()
        self.lib.oracle_default_material(C.byref(m))
        return m

    # ---- integer / rng
    def tea4(self, v0, v1):
        v0, v1 = _u32(v0), _u32(v1)
        out = np.empty_like(v0)
        self.lib.oracle_tea4(len(v0), _p(v0), _p(v1), _p(out))
        return out

    def lcg_stream(self, seed, n):
        st, f = np.empty(n, np.uint32), np.empty(n, np.float32)
        self.lib.oracle_lcg_stream(C.c_uint32(seed), n, _p(st), _p(f))
        return st, f

    def random_stream(self, seed, n, as_float):
        u, f, st = np.zeros(n, np.uint32), np.zeros(n, np.float32), np.zeros(2, np.uint32)
        self.lib.oracle_random_stream(C.c_uint32(seed), n, int(as_float), _p(u), _p(f), _p(st))
        return (f if as_float else u), st

    # ---- shading
    def basis(self, N):
        N = _f32(N)
        U, V = np.empty_like(N), np.empty_like(N)
        self.lib.oracle_basis(len(N), _p(N), _p(U), _p(V))
        return U, V

    @staticmethod
    def _mats(materials):
        arr = (Material * len(materials))(*materials)
        return arr

    def bsdf_sample(self, materials, etaI, etaO, N, view, seeds):
        n = len(materials)
        mats = self._mats(materials)
        etaI, etaO, N, view, seeds = _f32(etaI), _f32(etaO), _f32(N), _f32(view), _u32(seeds)
        L, pdf = np.empty((n, 3), np.float32), np.empty(n, np.float32)
        ty, rng = np.empty(n, np.int32), np.empty((n, 2), np.uint32)
        self.lib.oracle_bsdf_sample(n, mats, _p(etaI), _p(etaO), _p(N), _p(view), _p(seeds), _p(L), _p(pdf), _p(ty), _p(rng))
        return L, pdf, ty, rng

    def bsdf_eval(self, materials, albedo, etaI, etaO, N, V, L):
        n = len(materials)
        mats = self._mats(materials)
        albedo, etaI, etaO, N, V, L = map(_f32, (albedo, etaI, etaO, N, V, L))
        f, pdf = np.empty((n, 3), np.float32), np.empty(n, np.float32)
        self.lib.oracle_bsdf_eval(n, mats, _p(albedo), _p(etaI), _p(etaO), _p(N), _p(V), _p(L), _p(f), _p(pdf))
        return f, pdf

    def build_cdf(self, data):
        data = _f32(data)
        h, w = data.shape[:2]
        pdfX, cdfX = np.empty((h, w), np.float32), np.empty((h, w), np.float32)
        pdfY, cdfY = np.empty(h, np.float32), np.empty(h, np.float32)
        self.lib.oracle_build_cdf(w, h, _p(data), _p(pdfX), _p(cdfX), _p(pdfY), _p(cdfY))
        return pdfX, cdfX, pdfY, cdfY

    def probe_sample(self, probe, seeds):
        seeds = _u32(seeds)
        n = len(seeds)
        d, c, p = np.empty((n, 3), np.float32), np.empty((n, 3), np.float32), np.empty(n, np.float32)
        h, w = probe.data.shape[:2]
        self.lib.oracle_probe_sample(w, h, _p(probe.data), _p(probe.pdfX), _p(probe.cdfX), _p(probe.pdfY), _p(probe.cdfY),
                                     n, _p(seeds), _p(d), _p(c), _p(p))
        return d, c, p

    def probe_eval_dir(self, probe, dirs):
        dirs = _f32(dirs)
        n = len(dirs)
        uv, c, p = np.empty((n, 2), np.float32), np.empty((n, 3), np.float32), np.empty(n, np.float32)
        h, w = probe.data.shape[:2]
        self.lib.oracle_probe_eval_dir(w, h, _p(probe.data), _p(probe.pdfX), _p(probe.cdfX), _p(probe.pdfY), _p(probe.cdfY),
                                       n, _p(dirs), _p(uv), _p(c), _p(p))
        return uv, c, p

    def display(self, rgb, mode, exposure):
        rgb = _f32(rgb).reshape(-1, 3)
        out = np.empty(len(rgb), np.uint32)
        self.lib.oracle_display(len(rgb), _p(rgb), int(mode), C.c_float(exposure), _p(out))
        return out

    def tonemap_sqrt(self, rgba):
        rgba = _f32(rgba).reshape(-1, 4)
        out = np.empty(len(rgba), np.uint32)
        self.lib.oracle_tonemap_sqrt(len(rgba), _p(rgba), _p(out))
        return out

    def camera_uvw(self, eye, lookat, up, fovY, aspect):
        U, V, W = (np.empty(3, np.float32) for _ in range(3))
        self.lib.oracle_camera_uvw(_p(_f32(eye)), _p(_f32(lookat)), _p(_f32(up)), C.c_float(fovY), C.c_float(aspect),
                                   _p(U), _p(V), _p(W))
        return U, V, W

    # ---- scenes
    def scene(self, model, probe=None, brute_force=False):
        return OracleScene(self, model, probe, brute_force)


class OracleScene:
    """CPU twin of a renderer context: same Model / ProbeData inputs as optixpathtracer_b200.SampleRenderer."""

    def __init__(self, oracle, model, probe=None, brute_force=False):
        self.o, self.lib = oracle, oracle.lib
        self.h = C.c_void_p(self.lib.oracle_scene_create())
        for tex in model.textures:
            px = _u32(tex.pixel)
            self.lib.oracle_scene_add_texture(self.h, _p(px), int(tex.resolution[0]), int(tex.resolution[1]))
        for mesh in model.meshes:
            v, idx = _f32(mesh.vertex), _u32(mesh.index)
            tc = _f32(mesh.texcoord) if mesh.texcoord is not None and len(mesh.texcoord) else None
            mid = self.lib.oracle_scene_add_mesh(self.h, _p(v), C.c_size_t(len(v)), _p(tc), _p(idx), C.c_size_t(len(idx)),
                                                 C.byref(mesh.material), int(mesh.diffuseTextureID))
            if getattr(mesh, "normal", None) is not None and len(mesh.normal) == len(v) and len(v):
                self.lib.oracle_scene_set_mesh_normals(self.h, int(mid), _p(_f32(mesh.normal)))
        self.lib.oracle_scene_build(self.h, int(brute_force))
        if probe is not None:
            self.set_probe(probe)

    def __del__(self):
        try:
            self.lib.oracle_scene_destroy(self.h)
        except Exception:
            pass

    def set_probe(self, probe):
        h, w = probe.data.shape[:2]
        self.lib.oracle_scene_set_probe(self.h, w, h, _p(probe.data), _p(probe.pdfX), _p(probe.cdfX), _p(probe.pdfY), _p(probe.cdfY))

    def num_triangles(self):
        return int(self.lib.oracle_scene_num_triangles(self.h))

    def morton(self):
        n = self.num_triangles()
        codes, order, bounds = np.empty(n, np.uint64), np.empty(n, np.uint32), np.empty(6, np.float32)
        self.lib.oracle_morton(self.h, _p(codes), _p(order), _p(bounds))
        return codes, order, bounds

    def render(self, params: LaunchParams, cfg: Config, accum=None, frame=None, aovs=False):
        """SampleRenderer::render() on host images.  Returns (accum, frame, counters, seconds[, aovs])."""
        W, H = params.frame.size.x, params.frame.size.y
        if accum is None:
            accum = np.zeros((H, W, 4), np.float32)
        if frame is None:
            frame = np.zeros((H, W), np.uint32)
        extra = [np.zeros((H, W, 4), np.float32) for _ in range(3)] if aovs else [None] * 3
        cnt = np.zeros(3, np.uint64)
        sec = self.lib.oracle_render(self.h, C.byref(params), C.byref(cfg), _p(accum), _p(frame), _p(extra[0]), _p(extra[1]),
                                     _p(extra[2]), _p(cnt))
        counters = dict(samples=int(cnt[0]), radiance_rays=int(cnt[1]), shadow_rays=int(cnt[2]))
        if aovs:
            return accum, frame, counters, sec, dict(normal=extra[0], albedo=extra[1], color=extra[2])
        return accum, frame, counters, sec

    def render_rows(self, params, cfg, row0, row1):
        W, H = params.frame.size.x, params.frame.size.y
        accum, frame, cnt = np.zeros((H, W, 4), np.float32), np.zeros((H, W), np.uint32), np.zeros(3, np.uint64)
        sec = self.lib.oracle_render_rows(self.h, C.byref(params), C.byref(cfg), _p(accum), _p(frame), int(row0), int(row1), _p(cnt))
        return accum, dict(samples=int(cnt[0]), radiance_rays=int(cnt[1]), shadow_rays=int(cnt[2])), sec

    def trace_primary(self, params, cfg, launch_w=None, launch_h=None):
        lw = launch_w or params.frame.size.x
        lh = launch_h or params.frame.size.y
        n = lw * lh
        prim, mesh = np.empty(n, np.int32), np.empty(n, np.int32)
        t, u, v = (np.empty(n, np.float32) for _ in range(3))
        self.lib.oracle_trace_primary(self.h, C.byref(params), C.byref(cfg), lw, lh, _p(prim), _p(mesh), _p(t), _p(u), _p(v))
        return prim, mesh, t, u, v

    def trace_rays(self, org, dirs, tmin, tmax, any_hit=False, shadow_mode=0):
        org, dirs = _f32(org), _f32(dirs)
        n = len(org)
        prim, mesh = np.zeros(n, np.int32), np.zeros(n, np.int32)
        t, u, v = (np.zeros(n, np.float32) for _ in range(3))
        self.lib.oracle_trace_rays(self.h, _p(org), _p(dirs), C.c_size_t(n), C.c_float(tmin), C.c_float(tmax), int(any_hit),
                                   int(shadow_mode), _p(prim), _p(mesh), _p(t), _p(u), _p(v))
        return prim, mesh, t, u, v

    def render_whitted(self, width, height, cam12, light_pos):
        """04HelloRaytracing frame: cam12 = position, direction, horizontal, vertical (4 x 3 floats)."""
        cam = _f32(np.asarray(cam12, np.float32).reshape(12))
        light = _f32(np.asarray(light_pos, np.float32).reshape(3))
        out = np.zeros((height, width), np.uint32)
        self.lib.oracle_render_whitted(self.h, int(width), int(height), _p(cam), _p(light), _p(out))
        return out
