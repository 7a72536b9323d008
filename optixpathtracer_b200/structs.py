"""ctypes mirrors of the reference's launch ABI (bit-for-bit layouts).

reference: HelloPathtracing_sv4_vmv23/LaunchParams.h:24-106 (LaunchParams 312 B, TriangleMeshSBTData 152 B),
Material.h:11-70 (Material 104 B), Probe.cuh:6-21 (Probe 64 B).  The C side re-states the same layouts in
include/ref_api/*.h with static_asserts; tests/test_abi_layout.py checks both against SURVEY.md §8b.
"""
import ctypes as C


class float2(C.Structure):
    _fields_ = [("x", C.c_float), ("y", C.c_float)]


class float3(C.Structure):
    _fields_ = [("x", C.c_float), ("y", C.c_float), ("z", C.c_float)]

    def __init__(self, x=0.0, y=0.0, z=0.0):
        super().__init__(float(x), float(y), float(z))

    def tolist(self):
        return [self.x, self.y, self.z]


class int2(C.Structure):
    _fields_ = [("x", C.c_int), ("y", C.c_int)]


class uint2(C.Structure):          # CUDA's uint2 is 8-byte aligned; containers below pad explicitly
    _fields_ = [("x", C.c_uint), ("y", C.c_uint)]


class uint3(C.Structure):
    _fields_ = [("x", C.c_uint), ("y", C.c_uint), ("z", C.c_uint)]


class Material(C.Structure):
    """sv4/Material.h:11-70.  Default = the sv4 constructor values (:13-38)."""
    _fields_ = [
        ("emission", float3), ("color", float3), ("absorption", float3),
        ("eta", C.c_float), ("metallic", C.c_float), ("subsurface", C.c_float), ("specular", C.c_float),
        ("roughness", C.c_float), ("specularTint", C.c_float), ("anisotropic", C.c_float), ("sheen", C.c_float),
        ("sheenTint", C.c_float), ("clearcoat", C.c_float), ("clearcoatGloss", C.c_float),
        ("transmission", C.c_float), ("bump", C.c_float), ("bumpTile", float3), ("flags", C.c_int),
    ]

    def __init__(self, **kw):
        super().__init__()
        d = dict(emission=(1, 1, 1), color=(1, 0, 0), absorption=(1, 1, 1), eta=1.4, metallic=0.5, subsurface=0.0,
                 specular=1.0, roughness=1.0, specularTint=1.0, anisotropic=0.0, sheen=0.0, sheenTint=0.0,
                 clearcoat=0.0, clearcoatGloss=1.0, transmission=0.4, bump=0.0, bumpTile=(1, 1, 1), flags=0)
        d.update(kw)
        for k, v in d.items():
            setattr(self, k, float3(*v) if isinstance(v, (tuple, list)) else v)

    @classmethod
    def orig_defaults(cls, **kw):
        """HelloPathtracing_original/Material.h:13-36."""
        d = dict(emission=(0, 0, 0), color=(0.6, 0.6, 0.6), absorption=(0, 0, 0), eta=0.0, metallic=0.0,
                 subsurface=0.0, specular=0.5, roughness=1.0, specularTint=0.0, clearcoat=0.0, clearcoatGloss=1.0,
                 transmission=0.0, bump=0.0, bumpTile=(10, 10, 10), flags=0)
        d.update(kw)
        return cls(**d)

    @classmethod
    def sv3_defaults(cls, **kw):
        """HelloPathtracing_sv3/Material.h:13-37 (the green tint 1.6 is the reference's literal)."""
        d = dict(emission=(0, 0, 0), color=(0.6, 1.6, 0.6), absorption=(0, 0, 0), eta=1.4, metallic=0.0,
                 subsurface=0.0, specular=0.5, roughness=1.0, specularTint=0.0, clearcoat=0.0, clearcoatGloss=1.0,
                 transmission=0.0, bump=0.0, bumpTile=(10, 10, 10), flags=0)
        d.update(kw)
        return cls(**d)


MATERIAL_FLAG_SHADOW_CATCHER = 1
NOALLOC_MAX_ROWS = 16384          # B200PT_NOALLOC_MAX_ROWS, include/b200pt.h


class Probe(C.Structure):
    _fields_ = [
        ("width", C.c_int), ("height", C.c_int), ("data", C.c_void_p), ("offset", float3),
        ("pdfValuesX", C.c_void_p), ("cdfValuesX", C.c_void_p), ("pdfValuesY", C.c_void_p), ("cdfValuesY", C.c_void_p),
    ]


class ParallelogramLight(C.Structure):
    _fields_ = [("corner", float3), ("v1", float3), ("v2", float3), ("normal", float3), ("emission", float3)]


class _Frame(C.Structure):
    _fields_ = [
        ("accum_buffer", C.c_void_p), ("frame_buffer", C.c_void_p), ("color_buffer", C.c_void_p),
        ("normal_buffer", C.c_void_p), ("albedo_buffer", C.c_void_p),
        ("size", int2), ("subframe_index", C.c_uint), ("factor", uint3), ("fillSize", C.c_int),
        ("_pad_c", C.c_int),           # uint2 c is 8-byte aligned -> @72
        ("c", uint2), ("r_inner", C.c_float), ("r_outer", C.c_float), ("offset", uint2), ("redraw", C.c_uint),
    ]


class _Camera(C.Structure):
    _fields_ = [("eye", float3), ("U", float3), ("V", float3), ("W", float3)]


class LaunchParams(C.Structure):
    _fields_ = [
        ("frame", _Frame), ("camera", _Camera), ("samples_per_launch", C.c_uint),
        ("traversable", C.c_uint64), ("light", ParallelogramLight), ("probe", Probe),
        ("viewportSize", int2), ("white", C.c_float),
    ]


class TriangleMeshSBTData(C.Structure):
    _fields_ = [("vertex", C.c_void_p), ("normal", C.c_void_p), ("texcoord", C.c_void_p), ("index", C.c_void_p),
                ("hasTexture", C.c_bool), ("texture", C.c_uint64), ("material", Material)]


# --- include/b200pt.h -----------------------------------------------------------------------------------
SCHEDULE_AS_GIVEN, SCHEDULE_UNIFORM, SCHEDULE_FOVEATED = 0, 1, 2
SHADOW_ANY_HIT, SHADOW_CULL_BACK = 0, 1
DISPLAY_SRGB, DISPLAY_REINHARD = 0, 1
ACCUM_LERP, ACCUM_SUM = 0, 1


class Config(C.Structure):
    """b200pt_config (include/b200pt.h)."""
    _fields_ = [
        ("schedule", C.c_int), ("max_depth", C.c_int), ("tmin", C.c_float), ("tmax", C.c_float),
        ("shadow_mode", C.c_int), ("display", C.c_int), ("exposure", C.c_float), ("uniform_spl", C.c_int),
        ("accumulate", C.c_int), ("inner_radius", C.c_int), ("outer_radius", C.c_int),
        ("launch_w", C.c_int), ("launch_h", C.c_int), ("write_aovs", C.c_int), ("increment_subframe", C.c_int),
        ("accum_mode", C.c_int), ("russian_roulette", C.c_int), ("tile_rank", C.c_int), ("tile_count", C.c_int),
        ("profile", C.c_int), ("mis_on_miss", C.c_int), ("shadow_tmin", C.c_float), ("frames_in_flight", C.c_int), ("intra_frame_split", C.c_int),
    ]


_PRESETS = {
    # occlusion rays start at 0.01 in every variant, radiance rays at 0.001 in orig only (orig/deviceProgram.cu:268,310 vs :420)
    # orig/deviceProgram.cu:198,429,467 ; orig/SimplePathtracer.cpp:73-97 (render() leaves subframe_index to main)
    "orig": dict(schedule=SCHEDULE_UNIFORM, max_depth=8, tmin=0.001, shadow_tmin=0.01, tmax=1e16, shadow_mode=SHADOW_ANY_HIT,
                 display=DISPLAY_SRGB, exposure=1.0, uniform_spl=0, accumulate=1, write_aovs=1, increment_subframe=0),
    # sv3/deviceProgram.cu:243,511,582 ; sv3/SimplePathtracer.cpp:81-128
    "sv3_uniform": dict(schedule=SCHEDULE_UNIFORM, max_depth=4, tmin=0.01, shadow_tmin=0.01, tmax=1e16, shadow_mode=SHADOW_CULL_BACK,
                        display=DISPLAY_REINHARD, exposure=8.0, uniform_spl=32, accumulate=1, increment_subframe=1),
    # sv4/SimplePathtracer.cpp:85-131
    "sv4_uniform": dict(schedule=SCHEDULE_UNIFORM, max_depth=4, tmin=0.01, shadow_tmin=0.01, tmax=1e16, shadow_mode=SHADOW_CULL_BACK,
                        display=DISPLAY_REINHARD, exposure=4.0, uniform_spl=8, accumulate=0, increment_subframe=1),
    # sv4/SimplePathtracer.cpp:20-21,132-216
    "sv4_fov": dict(schedule=SCHEDULE_FOVEATED, max_depth=4, tmin=0.01, shadow_tmin=0.01, tmax=1e16, shadow_mode=SHADOW_CULL_BACK,
                    display=DISPLAY_REINHARD, exposure=4.0, inner_radius=157, outer_radius=515, increment_subframe=1),
}


def preset(name, **overrides):
    """Python twin of b200pt_config_preset(); tests check the two agree."""
    cfg = Config()
    for k, v in {**_PRESETS[name], **overrides}.items():
        setattr(cfg, k, v)
    return cfg


class BvhInfo(C.Structure):
    _fields_ = [("n_triangles", C.c_uint64), ("n_nodes", C.c_uint64), ("n_leaves", C.c_uint64),
                ("max_leaf_size", C.c_uint32), ("max_depth", C.c_uint32), ("sah_cost", C.c_float),
                ("sah_cost_lbvh", C.c_float), ("build_ms", C.c_float), ("node_bytes", C.c_uint64),
                ("tri_bytes", C.c_uint64), ("root_link", C.c_int32), ("refine_levels", C.c_uint32),
                ("grid_lo", C.c_double * 3), ("grid_scale", C.c_double * 3), ("refit_ms", C.c_float), ("n_wide_nodes", C.c_uint32), ("wide_depth", C.c_uint32), ("reserved", C.c_uint32)]


class Counters(C.Structure):
    _fields_ = [("samples", C.c_uint64), ("radiance_rays", C.c_uint64), ("shadow_rays", C.c_uint64),
                ("nodes_fetched", C.c_uint64), ("tris_fetched", C.c_uint64),
                ("nodes_fetched_shadow", C.c_uint64), ("tris_fetched_shadow", C.c_uint64),
                ("ms_total", C.c_float), ("ms_raygen", C.c_float), ("ms_trace_closest", C.c_float),
                ("ms_trace_any", C.c_float), ("ms_shade", C.c_float), ("ms_accumulate", C.c_float),
                ("kernel_launches", C.c_uint32), ("n_trace_closest_launches", C.c_uint32),
                ("n_trace_any_launches", C.c_uint32), ("fused_finalize", C.c_uint32),
                ("closest_hits", C.c_uint64), ("occluded_shadow_rays", C.c_uint64)]


assert C.sizeof(LaunchParams) == 312 and C.sizeof(Material) == 104 and C.sizeof(Probe) == 64
assert C.sizeof(TriangleMeshSBTData) == 152 and C.sizeof(ParallelogramLight) == 60
