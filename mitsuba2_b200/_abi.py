"""ctypes mirror of include/mtsb200.h and the loader of the CUDA library.

The product path has no CPU fallback: if ``libmtsb200.so`` is missing or fails to load,
:func:`lib` raises.  (The oracle under ``oracle/`` is test infrastructure and is never
loaded from here.)
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libmtsb200.so")

MEM_HOST, MEM_DEVICE = 0, 1
SHAPE_SPHERE, SHAPE_DISK, SHAPE_CYLINDER, SHAPE_RECTANGLE = 1, 2, 3, 4
TEX_CONSTANT, TEX_BITMAP, TEX_REGULAR, TEX_IRREGULAR, TEX_UNIFORM, TEX_MESH_ATTRIBUTE = 0, 1, 2, 3, 4, 5
MODE_RGB, MODE_SPECTRAL = 0, 1
BSDF_DIFFUSE, BSDF_ROUGHCONDUCTOR, BSDF_DIELECTRIC, BSDF_CONDUCTOR, BSDF_PLASTIC, BSDF_ROUGHDIELECTRIC, BSDF_ROUGHPLASTIC = 0, 1, 2, 3, 4, 5, 6
MF_BECKMANN, MF_GGX = 0, 1
EMITTER_AREA, EMITTER_ENVMAP, EMITTER_CONSTANT, EMITTER_POINT = 0, 1, 2, 3
RFILTER_BOX, RFILTER_TENT, RFILTER_GAUSSIAN, RFILTER_MITCHELL, RFILTER_CATMULLROM, RFILTER_LANCZOS = 0, 1, 2, 3, 4, 5

c_float_p = C.POINTER(C.c_float)
c_u32_p = C.POINTER(C.c_uint32)
c_u64_p = C.POINTER(C.c_uint64)


class TextureDesc(C.Structure):
    _fields_ = [("type", C.c_int32), ("value", C.c_float * 3), ("width", C.c_uint32), ("height", C.c_uint32),
                ("channels", C.c_uint32), ("data", c_float_p), ("to_uv", C.c_float * 9),
                ("lambda_min", C.c_float), ("lambda_max", C.c_float), ("d65_scale", C.c_float)]


class BsdfDesc(C.Structure):
    _fields_ = [("type", C.c_int32), ("reflectance_tex", C.c_int32), ("eta_tex", C.c_int32), ("k_tex", C.c_int32),
                ("specular_reflectance_tex", C.c_int32), ("specular_transmittance_tex", C.c_int32),
                ("alpha_u", C.c_float), ("alpha_v", C.c_float), ("distribution", C.c_int32),
                ("sample_visible", C.c_int32), ("eta", C.c_float), ("two_sided", C.c_int32), ("nonlinear", C.c_int32),
                ("specular_sampling_weight", C.c_float), ("opacity_tex_plus1", C.c_int32)]


class EmitterDesc(C.Structure):
    _fields_ = [("type", C.c_int32), ("radiance", C.c_float * 3), ("mesh_id", C.c_int32), ("width", C.c_uint32),
                ("height", C.c_uint32), ("rgba", c_float_p), ("to_world", C.c_float * 16), ("scale", C.c_float),
                ("position", C.c_float * 3), ("coeff_scale", c_float_p), ("radiance_tex", C.c_int32)]


class MeshDesc(C.Structure):
    _fields_ = [("faces", c_u32_p), ("positions", c_float_p), ("normals", c_float_p), ("texcoords", c_float_p),
                ("face_count", C.c_uint32), ("vertex_count", C.c_uint32), ("bsdf_id", C.c_int32),
                ("emitter_id", C.c_int32), ("mem", C.c_int32)]


class ShapeDesc(C.Structure):
    _fields_ = [("type", C.c_int32), ("flip_normals", C.c_int32), ("to_world", C.c_float * 16),
                ("bsdf_id", C.c_int32), ("emitter_id", C.c_int32)]


class SensorDesc(C.Structure):
    _fields_ = [("to_world", C.c_float * 16), ("sample_to_camera", C.c_float * 16), ("near_clip", C.c_float),
                ("far_clip", C.c_float), ("film_width", C.c_uint32), ("film_height", C.c_uint32),
                ("rfilter", C.c_int32), ("rfilter_radius", C.c_float), ("rfilter_stddev", C.c_float), ("rfilter_c", C.c_float)]


class IntegratorDesc(C.Structure):
    _fields_ = [("max_depth", C.c_int32), ("rr_depth", C.c_int32)]


class SceneDesc(C.Structure):
    _fields_ = [("n_meshes", C.c_uint32), ("n_bsdfs", C.c_uint32), ("n_textures", C.c_uint32),
                ("n_emitters", C.c_uint32), ("n_sensors", C.c_uint32),
                ("meshes", C.POINTER(MeshDesc)), ("bsdfs", C.POINTER(BsdfDesc)), ("textures", C.POINTER(TextureDesc)),
                ("emitters", C.POINTER(EmitterDesc)), ("sensors", C.POINTER(SensorDesc)),
                ("integrator", IntegratorDesc), ("mode", C.c_int32),
                ("n_shapes", C.c_uint32), ("shapes", C.POINTER(ShapeDesc))]


class Stats(C.Structure):
    _fields_ = [("samples", C.c_uint64), ("closest_rays", C.c_uint64), ("shadow_rays", C.c_uint64),
                ("segments", C.c_uint64), ("kernel_launches", C.c_uint64), ("ms_total", C.c_float),
                ("ms_closest", C.c_float), ("ms_any", C.c_float), ("ms_shade", C.c_float), ("ms_raygen", C.c_float),
                ("ms_film", C.c_float), ("ms_build", C.c_float), ("bvh_nodes", C.c_uint32), ("bvh_depth", C.c_uint32)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


# every symbol include/mtsb200.h declares: (name, restype, argtypes)
vp = C.c_void_p
SYMBOLS = [
    ("mtsb_last_error", C.c_char_p, []),
    ("mtsb_abi_version", C.c_int, []),
    ("mtsb_ctx_create", C.c_int, [C.c_int, C.POINTER(vp)]),
    ("mtsb_ctx_destroy", C.c_int, [vp]),
    ("mtsb_cancel", C.c_int, [vp]),
    ("mtsb_set_profiling", C.c_int, [vp, C.c_int]),
    ("mtsb_set_timeout", C.c_int, [vp, C.c_float]),
    ("mtsb_scene_create", C.c_int, [vp, C.POINTER(SceneDesc), C.POINTER(vp)]),
    ("mtsb_scene_destroy", C.c_int, [vp]),
    ("mtsb_scene_stats", C.c_int, [vp, C.POINTER(Stats)]),
    ("mtsb_scene_update_vertices", C.c_int, [vp, C.c_uint32, vp, C.c_int]),
    ("mtsb_scene_update_texture", C.c_int, [vp, C.c_uint32, vp, C.c_int]),
    ("mtsb_intersect_closest", C.c_int, [vp, C.c_uint64, vp, vp, vp, C.c_int, C.POINTER(Stats)]),
    ("mtsb_intersect_any", C.c_int, [vp, C.c_uint64, vp, vp, vp, C.c_int, C.POINTER(Stats)]),
    ("mtsb_intersect_naive", C.c_int, [vp, C.c_uint64, vp, vp, vp, C.c_int]),
    ("mtsb_intersect_closest_noalloc", C.c_int, [vp, C.c_uint64, vp, vp, vp, C.c_int, C.POINTER(Stats)]),
    ("mtsb_noalloc_max_rays", C.c_uint64, []),
    ("mtsb_render", C.c_int, [vp, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint64, C.POINTER(Stats)]),
    ("mtsb_film_clear", C.c_int, [vp, C.c_uint32]),
    ("mtsb_film_read", C.c_int, [vp, C.c_uint32, vp, C.c_int]),
    ("mtsb_film_enable_moments", C.c_int, [vp, C.c_uint32, C.c_int]),
    ("mtsb_film_read_moments", C.c_int, [vp, C.c_uint32, vp, C.c_int]),
    ("mtsb_film_accumulate", C.c_int, [vp, C.c_uint32, vp, C.c_int]),
    ("mtsb_film_develop", C.c_int, [vp, C.c_uint32, vp, C.c_int]),
    ("mtsb_film_device_ptr", C.c_int, [vp, C.c_uint32, C.POINTER(vp), C.POINTER(C.c_uint64)]),
    ("mtsb_render_grad", C.c_int, [vp, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint64, vp,
                                   C.c_int32, vp, C.c_int32, vp, C.c_int, C.POINTER(Stats)]),
    ("mtsb_grad_config", C.c_int, [vp, C.c_int, C.c_float]),
    ("mtsb_scene_update_config", C.c_int, [vp, C.c_int, C.c_uint32]),
    ("mtsb_pcg32_floats", C.c_int, [vp, C.c_uint64, vp, vp, C.c_uint32, vp]),
    ("mtsb_sampler_seed", C.c_int, [vp, C.c_uint64, C.c_uint64, vp, vp]),
    ("mtsb_sensor_sample_ray", C.c_int, [vp, C.c_uint32, C.c_uint64, vp, vp, vp]),
    ("mtsb_bsdf_sample", C.c_int, [vp, C.c_uint32, C.c_uint64, vp, vp, vp, vp, vp]),
    ("mtsb_bsdf_eval_pdf", C.c_int, [vp, C.c_uint32, C.c_uint64, vp, vp, vp, vp, vp]),
    ("mtsb_sample_emitter_direction", C.c_int, [vp, C.c_uint64, vp, vp, vp]),
    ("mtsb_film_put", C.c_int, [vp, C.c_uint32, C.c_uint64, vp, vp]),
]

_lib = None


def lib():
    """Load libmtsb200.so (built by ``__graft_entry__.build()``); raises if it is missing."""
    global _lib
    if _lib is None:
        path = os.environ.get("MTSB_LIB", LIB_PATH)       # MTSB_LIB: alternative build of the same library (kernel experiments)
        if not os.path.exists(path):
            raise RuntimeError(
                f"mitsuba2_b200: CUDA library {path} not found -- run `python -c 'import __graft_entry__ as g; "
                "g.build()'` (nvcc, sm_100a). There is no CPU fallback.")
        l = C.CDLL(path)
        for name, res, args in SYMBOLS:
            fn = getattr(l, name)
            fn.restype = res
            fn.argtypes = args
        _lib = l
    return _lib


class MtsbError(RuntimeError):
    pass


def check(rc):
    if rc != 0:
        msg = lib().mtsb_last_error()
        raise MtsbError(f"mtsb200 error {rc}: {msg.decode() if msg else '?'}")
