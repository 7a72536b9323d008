"""Python mirror of the reference's host renderer class, forwarding 1:1 to the C-ABI (include/b200pt.h).

reference: class SampleRenderer, HelloPathtracing_sv4_vmv23/SimplePathtracer.h:45-188 and
SimplePathtracer.cpp:42-75 (ctor), :77-217 (render), :231-277 (resize), :279-283 (downloadPixels),
:285-292 (setCamera), :295-311 (setProbe).  `launchParams` is a public, live LaunchParams that callers poke
directly (main.cpp:310-320,392) and that every render() re-reads.
"""
import ctypes as C

import numpy as np

from . import _lib
from .structs import BvhInfo, Config, Counters, LaunchParams, float3, preset


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def host_params(width, height, camera=None, samples_per_launch=1, subframe_index=0):
    """A LaunchParams filled the way resize()+setCamera() would, without device buffers (CPU-side tools)."""
    p = LaunchParams()
    p.frame.size.x, p.frame.size.y = int(width), int(height)
    p.frame.subframe_index = subframe_index
    p.frame.factor.x = p.frame.factor.y = p.frame.factor.z = 1
    p.frame.fillSize = 1
    p.frame.r_inner, p.frame.r_outer = 0.0, 1e9
    p.samples_per_launch = samples_per_launch
    if camera is not None:
        set_camera(p, camera)
    return p


def set_camera(p, camera):
    """SimplePathtracer.cpp:285-292 — aspect is overridden by the frame size."""
    with np.errstate(invalid="ignore", divide="ignore"):      # before resize() the reference divides 0/0 here too (main.cpp:253 precedes :256)
        camera.setAspectRatio(np.float32(p.frame.size.x) / np.float32(p.frame.size.y))
    U, V, W = camera.UVWFrame()
    p.camera.U, p.camera.V, p.camera.W = float3(*U), float3(*V), float3(*W)
    p.camera.eye = float3(*camera.eye)


WHITTED_LIGHT_POS = (-907.108, 2205.875, -400.0267)          # 04HelloRaytracing/deviceProgram.cu:125


def whitted_camera(camera, width, height):
    """04HelloRaytracing/SampleRenderer.cpp:125-138 -> (4, 3) float32: position, direction, horizontal, vertical."""
    f = np.float32

    def norm(v):
        v = v.astype(np.float32)
        d = f(v[0] * v[0] + v[1] * v[1]) + f(v[2] * v[2])
        return (v * (f(1.0) / np.sqrt(d, dtype=np.float32))).astype(np.float32)

    def cross(a, b):
        return np.array([a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0]], np.float32)

    eye, lookat, up = (np.asarray(v, np.float32) for v in (camera.eye, camera.lookat, camera.up))
    direction = norm(lookat - eye)
    cos_fovy = f(0.66)
    with np.errstate(invalid="ignore", divide="ignore"):
        aspect = f(width) / f(height)
    horizontal = ((cos_fovy * aspect) * norm(cross(direction, up))).astype(np.float32)
    vertical = (cos_fovy * norm(cross(horizontal, direction))).astype(np.float32)
    return np.stack([eye, direction, horizontal, vertical]).astype(np.float32)


class SampleRenderer:
    def __init__(self, model, device=0, config="sv4_fov"):
        """ctor: uploads every mesh/texture and builds the acceleration structure (LBVH) on the GPU."""
        self._L = _lib.lib()
        self.model = model                     # borrowed, like the reference (SimplePathtracer.h:173)
        self.config = preset(config) if isinstance(config, str) else config
        self.launchParams = LaunchParams()
        self._h = self._L.b200pt_create(int(device))
        if not self._h:
            raise _lib.B200ptError(self._L.b200pt_last_error().decode())
        self._h = C.c_void_p(self._h)
        for tex in model.textures:
            px = np.ascontiguousarray(tex.pixel, np.uint32)
            if self._L.b200pt_add_texture(self._h, _ptr(px), int(tex.resolution[0]), int(tex.resolution[1])) < 0:
                _lib.check(-1)
        for mesh in model.meshes:
            v = np.ascontiguousarray(mesh.vertex, np.float32)
            idx = np.ascontiguousarray(mesh.index, np.uint32)
            n = np.ascontiguousarray(mesh.normal, np.float32) if len(mesh.normal) else None
            tc = np.ascontiguousarray(mesh.texcoord, np.float32) if len(mesh.texcoord) else None
            if self._L.b200pt_add_mesh(self._h, _ptr(v), len(v), _ptr(n), _ptr(tc), _ptr(idx), len(idx),
                                       C.byref(mesh.material), int(mesh.diffuseTextureID)) < 0:
                _lib.check(-1)
        handle = C.c_uint64(0)
        _lib.check(self._L.b200pt_build_accel(self._h, C.byref(handle)))
        self.launchParams.traversable = handle.value

    def close(self):
        if getattr(self, "_h", None):
            self._L.b200pt_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- reference interface ---------------------------------------------------------------------
    def render(self, config=None, sync=True):
        if self.launchParams.frame.size.x == 0:
            return
        cfg = config or self.config
        fn = self._L.b200pt_render if sync else self._L.b200pt_render_async
        _lib.check(fn(self._h, C.byref(self.launchParams), C.byref(cfg)))

    def synchronize(self):
        _lib.check(self._L.b200pt_synchronize(self._h))

    def join(self):
        """Order the renderer's primary stream after every frame in flight (config.frames_in_flight > 1), without blocking the host."""
        _lib.check(self._L.b200pt_join(self._h))

    def resize(self, newSize):
        w, h = int(newSize[0]), int(newSize[1])
        if w == 0 or h == 0:
            return
        _lib.check(self._L.b200pt_resize(self._h, w, h, C.byref(self.launchParams)))

    def downloadPixels(self, h_pixels=None):
        w, h = self.launchParams.frame.size.x, self.launchParams.frame.size.y
        if h_pixels is None:
            h_pixels = np.empty((h, w), np.uint32)
        _lib.check(self._L.b200pt_download_pixels(self._h, _ptr(h_pixels)))
        return h_pixels

    def downloadPixelsAsync(self, h_pixels):
        """downloadPixels behind the most recently submitted frame (render(sync=False)); h_pixels: pinned (h, w) uint32, valid after synchronize()."""
        _lib.check(self._L.b200pt_download_pixels_async(self._h, _ptr(h_pixels)))

    def setCamera(self, camera):
        self.lastSetCamera = camera
        set_camera(self.launchParams, camera)

    def setProbe(self, probe):
        if not probe.valid:
            raise RuntimeError("Probe Data is not valid")          # Probe.h:104-105
        off = np.asarray(probe.offset, np.float32)
        _lib.check(self._L.b200pt_set_probe(self._h, probe.width, probe.height, _ptr(probe.data), _ptr(probe.pdfX),
                                            _ptr(probe.cdfX), _ptr(probe.pdfY), _ptr(probe.cdfY), _ptr(off),
                                            C.byref(self.launchParams)))

    def setProbeImage(self, data, offset=(0.0, 0.0, 0.0)):
        """loadProbe/loadColor + BuildCDF + setProbe with the CDF tables built on the GPU (bit-identical to
        ProbeData.BuildCDF, sv4/Probe.h:29-77); `data` is (H, W, 4) float32."""
        data = np.ascontiguousarray(data, np.float32)
        if data.ndim != 3 or data.shape[2] != 4:
            raise RuntimeError("Probe Data is not valid")
        off = np.asarray(offset, np.float32)
        _lib.check(self._L.b200pt_set_probe_image(self._h, data.shape[1], data.shape[0], _ptr(data), _ptr(off),
                                                  C.byref(self.launchParams)))

    def buildCdfNoAlloc(self, width, height, d_data, d_pdfX, d_cdfX, d_pdfY, d_cdfY, stream=None):
        """BuildCDF on caller-owned DEVICE buffers (integer addresses) without any allocation (b200pt_build_cdf_noalloc): height <=
        structs.NOALLOC_MAX_ROWS; asynchronous on `stream` (None = the renderer's primary stream)."""
        _lib.check(self._L.b200pt_build_cdf_noalloc(self._h, int(width), int(height), C.c_void_p(d_data), C.c_void_p(d_pdfX), C.c_void_p(d_cdfX),
                                                    C.c_void_p(d_pdfY), C.c_void_p(d_cdfY), C.c_void_p(stream) if stream else None))

    def downloadProbe(self):
        """(pdfX, cdfX, pdfY, cdfY) of the current probe, read back from the device."""
        w, h = self.launchParams.probe.width, self.launchParams.probe.height
        pdfX, cdfX = np.empty((h, w), np.float32), np.empty((h, w), np.float32)
        pdfY, cdfY = np.empty(h, np.float32), np.empty(h, np.float32)
        _lib.check(self._L.b200pt_download_probe(self._h, _ptr(pdfX), _ptr(cdfX), _ptr(pdfY), _ptr(cdfY)))
        return pdfX, cdfX, pdfY, cdfY

    def renderWhitted(self, camera, light_pos=WHITTED_LIGHT_POS):
        """Second integrator on the same traversal kernels: the 04HelloRaytracing frame (deviceProgram.cu:66-243) into the
        frame buffer (downloadPixels).  Camera set up as 04HelloRaytracing/SampleRenderer.cpp:125-138."""
        cam = whitted_camera(camera, self.launchParams.frame.size.x, self.launchParams.frame.size.y)
        buf = np.concatenate([cam.reshape(12), np.asarray(light_pos, np.float32)]).astype(np.float32)
        _lib.check(self._L.b200pt_render_whitted(self._h, _ptr(buf), None))

    # ---- animated geometry: refit or rebuild per frame ----------------------------------------------
    def updateVertices(self, meshID, vertex):
        """New positions for one mesh (same count); follow with refitAccel() or rebuildAccel()."""
        v = np.ascontiguousarray(vertex, np.float32)
        _lib.check(self._L.b200pt_update_vertices(self._h, int(meshID), _ptr(v), len(v)))

    def refitAccel(self):
        _lib.check(self._L.b200pt_refit_accel(self._h))

    def rebuildAccel(self):
        handle = C.c_uint64(0)
        _lib.check(self._L.b200pt_build_accel(self._h, C.byref(handle)))
        self.launchParams.traversable = handle.value

    # ---- extras behind the same C-ABI -----------------------------------------------------------
    def downloadAccum(self):
        w, h = self.launchParams.frame.size.x, self.launchParams.frame.size.y
        out = np.empty((h, w, 4), np.float32)
        _lib.check(self._L.b200pt_download_accum(self._h, _ptr(out)))
        return out

    def uploadAccum(self, rgba):
        rgba = np.ascontiguousarray(rgba, np.float32)
        _lib.check(self._L.b200pt_upload_accum(self._h, _ptr(rgba)))

    def downloadAOV(self, which):
        w, h = self.launchParams.frame.size.x, self.launchParams.frame.size.y
        out = np.empty((h, w, 4), np.float32)
        _lib.check(self._L.b200pt_download_aov(self._h, {"normal": 0, "albedo": 1, "color": 2}[which], _ptr(out)))
        return out

    def clearAccum(self):
        _lib.check(self._L.b200pt_clear_accum(self._h))

    def resolveSum(self, n_passes, config=None):
        cfg = config or self.config
        _lib.check(self._L.b200pt_resolve_sum(self._h, C.byref(self.launchParams), C.byref(cfg), int(n_passes)))

    # ---- multi-GPU below Python: NCCL collectives issued by the library (include/b200pt.h) ----------
    @staticmethod
    def commUniqueId():
        """ncclGetUniqueId -> 128 bytes (numpy uint8) for rank 0 to hand to the other ranks."""
        L = _lib.lib()
        buf = np.zeros(128, np.uint8)
        _lib.check(L.b200pt_comm_unique_id(_ptr(buf)))
        return buf

    def commInit(self, n_ranks, rank, unique_id):
        uid = np.ascontiguousarray(unique_id, np.uint8)
        assert uid.size == 128
        _lib.check(self._L.b200pt_comm_init(self._h, int(n_ranks), int(rank), _ptr(uid), C.byref(self.launchParams)))

    def commDestroy(self):
        _lib.check(self._L.b200pt_comm_destroy(self._h, C.byref(self.launchParams)))

    def commInfo(self):
        n, r, v, reg = C.c_int(), C.c_int(), C.c_int(), C.c_int()
        _lib.check(self._L.b200pt_comm_info(self._h, C.byref(n), C.byref(r), C.byref(v), C.byref(reg)))
        return dict(n_ranks=n.value, rank=r.value, nccl_version=v.value, accum_nccl_alloc=bool(reg.value & 1), accum_registered=bool(reg.value & 2))

    def reduceSum(self, n_passes, root=0, config=None, keep_local=False):
        """Sample split: fp32 SUM reduce of the accumulation buffer onto `root` + resolve there (asynchronous on the primary stream).
        keep_local: progressive display - the ranks keep their partial sums, root only refreshes its frame buffer."""
        cfg = config or self.config
        _lib.check(self._L.b200pt_reduce_sum(self._h, C.byref(self.launchParams), C.byref(cfg), int(root), int(n_passes), int(bool(keep_local))))

    def gatherBands(self, root=0, frame=True, accum=False):
        """Tile split: this rank's 32-row bands go to `root` (asynchronous on the primary stream)."""
        _lib.check(self._L.b200pt_gather_bands(self._h, C.byref(self.launchParams), int(root), (1 if frame else 0) | (2 if accum else 0)))

    # ---- one-sided tile split over NVLink peer memory (b200pt_peer_*): the finalize pass of every other rank writes its bands into root's frame
    PEER_HANDLE_BYTES = 192

    def peerExport(self, rank, n_ranks):
        """Root, once: exports the RGBA8 frame buffer; returns the 192-byte handle (numpy uint8) every other rank passes to peerAttach."""
        h = np.zeros(self.PEER_HANDLE_BYTES, np.uint8)
        _lib.check(self._L.b200pt_peer_export(self._h, C.byref(self.launchParams), int(rank), int(n_ranks), h.ctypes.data))
        return h

    def peerAttach(self, handle, rank, n_ranks):
        """Other ranks, once: maps root's frame; from now on tile-split frames (cfg.tile_count > 1) are written there by finalizeKernel."""
        h = np.ascontiguousarray(handle, np.uint8)
        if h.size != self.PEER_HANDLE_BYTES:
            raise ValueError("peerAttach: the handle must be the 192 bytes peerExport returned")
        _lib.check(self._L.b200pt_peer_attach(self._h, h.ctypes.data, int(rank), int(n_ranks)))

    def peerDetach(self):
        _lib.check(self._L.b200pt_peer_detach(self._h))

    def peerAcquire(self):
        """Root, every frame: the primary stream waits until every rank's bands of the next frame have landed (no host synchronisation)."""
        _lib.check(self._L.b200pt_peer_acquire(self._h))

    def peerRelease(self):
        """Root, every frame, after the frame has been consumed on the primary stream: the ranks may write the next frame."""
        _lib.check(self._L.b200pt_peer_release(self._h))

    def peerStatus(self):
        st, r, e = C.c_int(), C.c_int(), C.c_uint()
        _lib.check(self._L.b200pt_peer_status(self._h, C.byref(st), C.byref(r), C.byref(e)))
        return dict(state=st.value, attached_rank=r.value, epoch=e.value)

    def deviceBuffers(self):
        a, f, s = C.c_void_p(), C.c_void_p(), C.c_void_p()
        _lib.check(self._L.b200pt_device_buffers(self._h, C.byref(a), C.byref(f), C.byref(s)))
        return a.value, f.value, s.value

    def counters(self):
        c = Counters()
        _lib.check(self._L.b200pt_get_counters(self._h, C.byref(c)))
        return c

    def bvhInfo(self):
        b = BvhInfo()
        _lib.check(self._L.b200pt_bvh_info_get(self._h, C.byref(b)))
        return b

    def tracePrimary(self, config=None):
        cfg = config or self.config
        lw = cfg.launch_w or self.launchParams.frame.size.x
        lh = cfg.launch_h or self.launchParams.frame.size.y
        n = lw * lh
        prim, mesh = np.empty(n, np.int32), np.empty(n, np.int32)
        t, u, v = (np.empty(n, np.float32) for _ in range(3))
        _lib.check(self._L.b200pt_trace_primary(self._h, C.byref(self.launchParams), C.byref(cfg), _ptr(prim), _ptr(mesh),
                                                _ptr(t), _ptr(u), _ptr(v)))
        return prim, mesh, t, u, v

    def traceRays(self, org, dirs, tmin, tmax, any_hit=False, shadow_mode=0):
        org, dirs = np.ascontiguousarray(org, np.float32), np.ascontiguousarray(dirs, np.float32)
        n = len(org)
        prim, mesh = np.zeros(n, np.int32), np.zeros(n, np.int32)
        t, u, v = (np.zeros(n, np.float32) for _ in range(3))
        _lib.check(self._L.b200pt_trace_rays(self._h, _ptr(org), _ptr(dirs), n, tmin, tmax, int(any_hit), int(shadow_mode),
                                             _ptr(prim), _ptr(mesh), _ptr(t), _ptr(u), _ptr(v)))
        return prim, mesh, t, u, v

    def debugBsdf(self, materials, etaI, etaO, N, view, L, albedo, seeds):
        """Device-level KAT hook: BasisFromVector, BSDFSample (fresh Random(seed)), BSDFEval, BSDFPdf on the given inputs."""
        from .structs import Material
        n = len(materials)
        mats = (Material * n)(*materials)
        f32 = lambda a: np.ascontiguousarray(a, np.float32)
        etaI, etaO, N, view, L, albedo = map(f32, (etaI, etaO, N, view, L, albedo))
        seeds = np.ascontiguousarray(seeds, np.uint32)
        out = dict(basisU=np.empty((n, 3), np.float32), basisV=np.empty((n, 3), np.float32), sampleL=np.empty((n, 3), np.float32),
                   samplePdf=np.empty(n, np.float32), sampleType=np.empty(n, np.int32), sampleRng=np.empty((n, 2), np.uint32),
                   evalF=np.empty((n, 3), np.float32), evalPdf=np.empty(n, np.float32))
        _lib.check(self._L.b200pt_debug_bsdf(self._h, n, C.cast(mats, C.c_void_p), _ptr(etaI), _ptr(etaO), _ptr(N), _ptr(view), _ptr(L), _ptr(albedo),
                                             _ptr(seeds), *[_ptr(out[k]) for k in ("basisU", "basisV", "sampleL", "samplePdf", "sampleType", "sampleRng", "evalF", "evalPdf")]))
        return out

    def debugProbe(self, seeds, dirs, use_guide=True):
        """Device-level KAT hook on the current probe: ProbeSample (fresh Random(seed)) and ProbeDirToUV / ProbeEval / ProbePdf of `dirs`."""
        seeds = np.ascontiguousarray(seeds, np.uint32)
        dirs = np.ascontiguousarray(dirs, np.float32)
        n, m = len(seeds), len(dirs)
        out = dict(sampleDir=np.empty((n, 3), np.float32), sampleColor=np.empty((n, 3), np.float32), samplePdf=np.empty(n, np.float32),
                   uv=np.empty((m, 2), np.float32), evalColor=np.empty((m, 3), np.float32), evalPdf=np.empty(m, np.float32))
        _lib.check(self._L.b200pt_debug_probe(self._h, C.byref(self.launchParams), int(bool(use_guide)), n, _ptr(seeds), _ptr(out["sampleDir"]),
                                              _ptr(out["sampleColor"]), _ptr(out["samplePdf"]), m, _ptr(dirs), _ptr(out["uv"]), _ptr(out["evalColor"]),
                                              _ptr(out["evalPdf"])))
        return out

    def debugMorton(self):
        n = self.model.num_triangles()
        codes, order, bounds = np.empty(n, np.uint64), np.empty(n, np.uint32), np.empty(6, np.float32)
        _lib.check(self._L.b200pt_debug_morton(self._h, _ptr(codes), _ptr(order), _ptr(bounds)))
        return codes, order, bounds

    def downloadBvh(self):
        info = self.bvhInfo()
        nodes = np.empty((int(info.n_nodes), 8), np.uint32)
        tri_ids = np.empty(int(info.n_triangles), np.uint32)
        _lib.check(self._L.b200pt_download_bvh(self._h, _ptr(nodes), len(nodes), _ptr(tri_ids), len(tri_ids)))
        return nodes, tri_ids

    def downloadBvh4(self):
        """(n_wide_nodes, 16) uint32: the 4-wide nodes the traversal kernels read (include/b200pt.h)."""
        info = self.bvhInfo()
        nodes = np.empty((int(info.n_wide_nodes), 16), np.uint32)
        _lib.check(self._L.b200pt_download_bvh4(self._h, _ptr(nodes), len(nodes)))
        return nodes


def computeFinalPixelColors(renderer, rgba_f32):
    """toneMap.cu:41-70 through the host-buffer convenience entry point."""
    rgba = np.ascontiguousarray(rgba_f32, np.float32)
    h, w = rgba.shape[:2]
    out = np.empty((h, w), np.uint32)
    _lib.check(renderer._L.b200pt_tonemap_sqrt_host(renderer._h, w, h, _ptr(rgba), _ptr(out)))
    return out
