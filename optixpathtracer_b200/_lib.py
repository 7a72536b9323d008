"""Loader for the C-ABI shared library (csrc/libb200pt.so).  There is no CPU fallback: if the library is
missing or cannot be loaded, every product entry point raises."""
import ctypes as C
import os

from .structs import BvhInfo, Config, Counters, LaunchParams, Material

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("B200PT_LIB") or os.path.join(_HERE, "csrc", "libb200pt.so")   # B200PT_LIB: kernel-tuning builds only

# every symbol include/b200pt.h declares (tests/test_cabi_symbols.py checks the header against this list)
SYMBOLS = [
    "b200pt_config_preset", "b200pt_create", "b200pt_destroy", "b200pt_last_error", "b200pt_add_texture",
    "b200pt_add_mesh", "b200pt_update_vertices", "b200pt_refit_accel", "b200pt_build_accel", "b200pt_resize", "b200pt_set_probe", "b200pt_set_probe_image", "b200pt_download_probe", "b200pt_render",
    "b200pt_render_whitted", "b200pt_render_async", "b200pt_synchronize", "b200pt_join", "b200pt_download_pixels", "b200pt_download_pixels_async", "b200pt_download_accum",
    "b200pt_upload_accum", "b200pt_download_aov", "b200pt_device_buffers", "b200pt_resolve_sum", "b200pt_clear_accum",
    "b200pt_comm_unique_id", "b200pt_comm_init", "b200pt_comm_destroy", "b200pt_comm_info", "b200pt_reduce_sum", "b200pt_gather_bands",
    "b200pt_peer_export", "b200pt_peer_attach", "b200pt_peer_detach", "b200pt_peer_acquire", "b200pt_peer_release", "b200pt_peer_status",
    "b200pt_build_cdf_noalloc", "b200pt_tonemap_sqrt", "b200pt_tonemap_sqrt_host", "b200pt_trace_primary", "b200pt_trace_rays",
    "b200pt_debug_bsdf", "b200pt_debug_probe", "b200pt_debug_morton", "b200pt_bvh_info_get", "b200pt_download_bvh", "b200pt_download_bvh4", "b200pt_get_counters",
    "computeFinalPixelColors",
]

_lib = None


class B200ptError(RuntimeError):
    pass


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise B200ptError(f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                          "(nvcc, sm_100a). There is no CPU fallback.")
    L = C.CDLL(LIB_PATH)
    vp, i32, f32 = C.c_void_p, C.c_int, C.c_float
    L.b200pt_create.restype = vp
    L.b200pt_create.argtypes = [i32]
    L.b200pt_destroy.argtypes = [vp]
    L.b200pt_destroy.restype = None
    L.b200pt_last_error.restype = C.c_char_p
    L.b200pt_config_preset.argtypes = [C.POINTER(Config), C.c_char_p]
    L.b200pt_add_texture.argtypes = [vp, vp, i32, i32]
    L.b200pt_add_mesh.argtypes = [vp, vp, C.c_size_t, vp, vp, vp, C.c_size_t, C.POINTER(Material), i32]
    L.b200pt_build_accel.argtypes = [vp, C.POINTER(C.c_uint64)]
    L.b200pt_resize.argtypes = [vp, i32, i32, C.POINTER(LaunchParams)]
    L.b200pt_set_probe.argtypes = [vp, i32, i32, vp, vp, vp, vp, vp, vp, C.POINTER(LaunchParams)]
    L.b200pt_set_probe_image.argtypes = [vp, i32, i32, vp, vp, C.POINTER(LaunchParams)]
    L.b200pt_download_probe.argtypes = [vp, vp, vp, vp, vp]
    L.b200pt_render_whitted.argtypes = [vp, vp, vp]
    L.b200pt_download_bvh4.argtypes = [vp, vp, C.c_size_t]
    L.b200pt_update_vertices.argtypes = [vp, i32, vp, C.c_size_t]
    L.b200pt_refit_accel.argtypes = [vp]
    for name in ("b200pt_render", "b200pt_render_async"):
        getattr(L, name).argtypes = [vp, C.POINTER(LaunchParams), C.POINTER(Config)]
    L.b200pt_synchronize.argtypes = [vp]
    L.b200pt_join.argtypes = [vp]
    L.b200pt_download_pixels.argtypes = [vp, vp]
    L.b200pt_download_pixels_async.argtypes = [vp, vp]
    L.b200pt_download_accum.argtypes = [vp, vp]
    L.b200pt_upload_accum.argtypes = [vp, vp]
    L.b200pt_download_aov.argtypes = [vp, i32, vp]
    L.b200pt_device_buffers.argtypes = [vp, C.POINTER(vp), C.POINTER(vp), C.POINTER(vp)]
    L.b200pt_resolve_sum.argtypes = [vp, C.POINTER(LaunchParams), C.POINTER(Config), i32]
    L.b200pt_clear_accum.argtypes = [vp]
    L.b200pt_build_cdf_noalloc.argtypes = [vp, i32, i32, vp, vp, vp, vp, vp, vp]
    L.b200pt_comm_unique_id.argtypes = [vp]
    L.b200pt_comm_init.argtypes = [vp, i32, i32, vp, C.POINTER(LaunchParams)]
    L.b200pt_comm_destroy.argtypes = [vp, C.POINTER(LaunchParams)]
    L.b200pt_comm_info.argtypes = [vp, C.POINTER(i32), C.POINTER(i32), C.POINTER(i32), C.POINTER(i32)]
    L.b200pt_reduce_sum.argtypes = [vp, C.POINTER(LaunchParams), C.POINTER(Config), i32, i32, i32]
    L.b200pt_gather_bands.argtypes = [vp, C.POINTER(LaunchParams), i32, i32]
    L.b200pt_peer_export.argtypes = [vp, C.POINTER(LaunchParams), i32, i32, vp]
    L.b200pt_peer_attach.argtypes = [vp, vp, i32, i32]
    L.b200pt_peer_detach.argtypes = [vp]
    L.b200pt_peer_acquire.argtypes = [vp]
    L.b200pt_peer_release.argtypes = [vp]
    L.b200pt_peer_status.argtypes = [vp, C.POINTER(i32), C.POINTER(i32), C.POINTER(C.c_uint)]
    L.b200pt_tonemap_sqrt.argtypes = [i32, i32, vp, vp, vp]
    L.b200pt_tonemap_sqrt_host.argtypes = [vp, i32, i32, vp, vp]
    L.b200pt_trace_primary.argtypes = [vp, C.POINTER(LaunchParams), C.POINTER(Config), vp, vp, vp, vp, vp]
    L.b200pt_trace_rays.argtypes = [vp, vp, vp, C.c_size_t, f32, f32, i32, i32, vp, vp, vp, vp, vp]
    L.b200pt_debug_morton.argtypes = [vp, vp, vp, vp]
    L.b200pt_debug_bsdf.argtypes = [vp, C.c_size_t] + [vp] * 16
    L.b200pt_debug_probe.argtypes = [vp, C.POINTER(LaunchParams), i32, C.c_size_t, vp, vp, vp, vp, C.c_size_t, vp, vp, vp, vp]
    L.b200pt_bvh_info_get.argtypes = [vp, C.POINTER(BvhInfo)]
    L.b200pt_download_bvh.argtypes = [vp, vp, C.c_size_t, vp, C.c_size_t]
    L.b200pt_get_counters.argtypes = [vp, C.POINTER(Counters)]
    _lib = L
    return L


def check(rc):
    if rc != 0:
        raise B200ptError(lib().b200pt_last_error().decode(errors="replace"))
