"""ctypes binding of libb200pt.so (include/b200pt.h).  Thin: every function here is one C-ABI call.

The library is built in-tree by ``__graft_entry__.build()`` / ``make -C pathtrace_b200/csrc``; importing this
module fails loudly when it is missing -- there is no Python or CPU fallback for the render path.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("B200PT_LIB") or os.path.join(_HERE, "libb200pt.so")  # B200PT_LIB: A/B builds

OK, ERR_INVALID, ERR_CUDA, ERR_NO_DEVICE, ERR_STATE = 0, 1, 2, 3, 4
PRIM_SPHERE, PRIM_RECT, PRIM_BOX = 0, 1, 2
PLANE_XY, PLANE_XZ, PLANE_YZ = 0, 1, 2
INST_MEDIUM, INST_BVH_DUP, INST_RECT_FLIP = 1, 2, 4
MAT_LAMBERTIAN, MAT_METAL, MAT_DIELECTRIC, MAT_DIFFUSE_LIGHT, MAT_ISOTROPIC = 0, 1, 2, 3, 4
TEX_CONSTANT, TEX_CHECKER, TEX_PERLIN = 0, 1, 2
INTEGRATOR_RPT, INTEGRATOR_INEEPT = 0, 4


class Instance(C.Structure):
    _fields_ = [("fwd", C.c_float * 12), ("prim_kind", C.c_int32), ("plane", C.c_int32), ("flags", C.c_uint32),
                ("material", C.c_int32), ("params", C.c_float * 6), ("density", C.c_float), ("light_index", C.c_int32)]


class Material(C.Structure):
    _fields_ = [("kind", C.c_int32), ("texture", C.c_int32), ("color", C.c_float * 3), ("ior", C.c_float),
                ("power", C.c_float), ("two_sided", C.c_int32)]


class Texture(C.Structure):
    _fields_ = [("kind", C.c_int32), ("color", C.c_float * 3), ("alpha", C.c_float), ("scale", C.c_float),
                ("odd", C.c_int32), ("even", C.c_int32)]


class Camera(C.Structure):
    _fields_ = [("origin", C.c_float * 3), ("lower_left", C.c_float * 3), ("horizontal", C.c_float * 3),
                ("vertical", C.c_float * 3), ("u", C.c_float * 3), ("v", C.c_float * 3), ("w", C.c_float * 3),
                ("lens_radius", C.c_float)]


class Image(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("rgba", C.POINTER(C.c_float))]


class SceneDesc(C.Structure):
    _fields_ = [("instances", C.POINTER(Instance)), ("n_instances", C.c_int32),
                ("materials", C.POINTER(Material)), ("n_materials", C.c_int32),
                ("textures", C.POINTER(Texture)), ("n_textures", C.c_int32),
                ("lights", C.POINTER(C.c_int32)), ("n_lights", C.c_int32),
                ("background", C.c_int32), ("camera", Camera),
                ("perlin_ranvec", C.POINTER(C.c_float)), ("perlin_perm", C.POINTER(C.c_int32)),
                ("images", C.POINTER(Image)), ("n_images", C.c_int32)]


class Config(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("exposure", C.c_float), ("gamma", C.c_float),
                ("integrator", C.c_int32), ("max_bounces", C.c_int32), ("samples", C.c_int32),
                ("light_samples", C.c_int32), ("russian_roulette", C.c_int32),
                ("only_direct_illumination", C.c_int32), ("normal_offset", C.c_float),
                ("block_width", C.c_int32), ("block_height", C.c_int32), ("threads", C.c_int32),
                ("seed", C.c_uint64), ("sample_offset", C.c_int64), ("device", C.c_int32),
                ("batch_samples", C.c_int32), ("stats_clamp", C.c_float), ("profile", C.c_int32),
                ("trace_probability", C.c_float), ("trace_max_paths", C.c_int32),
                ("physically_corrected", C.c_int32), ("device_count", C.c_int32), ("device_ids", C.c_int32 * 8),
                ("async_render", C.c_int32)]


class Progress(C.Structure):
    _fields_ = [("samples_done", C.c_int64), ("samples_total", C.c_int64), ("rays_extension", C.c_int64),
                ("rays_shadow", C.c_int64), ("rays_reference", C.c_int64), ("seconds", C.c_double),
                ("done", C.c_int32)]


class Profile(C.Structure):
    _fields_ = [("ms", C.c_double * 6), ("launches", C.c_int64 * 6),
                ("ext_rays", C.c_uint64), ("ext_boxes", C.c_uint64), ("ext_prims", C.c_uint64),
                ("nee_rays", C.c_uint64), ("nee_boxes", C.c_uint64), ("nee_prims", C.c_uint64),
                ("nee_traced", C.c_uint64), ("nee_queued", C.c_uint64)]


KERNEL_NAMES = ("raygen", "extend", "shade", "nee", "accumulate", "sort")

HIT_DTYPE = np.dtype([("instance", "<i4"), ("t", "<f4"), ("p", "<f4", 3), ("n", "<f4", 3), ("u", "<f4"), ("v", "<f4")])
assert HIT_DTYPE.itemsize == 40

# name -> (restype, argtypes): every symbol include/b200pt.h declares
SIGNATURES = {
    "b200pt_last_error": (C.c_char_p, []),
    "b200pt_version": (C.c_char_p, []),
    "b200pt_config_from_json": (C.c_int, [C.c_char_p, C.POINTER(Config), C.c_char_p, C.c_size_t, C.c_char_p, C.c_size_t]),
    "b200pt_scene_create_from_json": (C.c_int, [C.c_char_p, C.c_int32, C.c_int32, C.c_int32, C.POINTER(C.c_void_p)]),
    "b200pt_scene_create_from_file": (C.c_int, [C.c_char_p, C.c_int32, C.c_int32, C.c_int32, C.POINTER(C.c_void_p)]),
    "b200pt_scene_create": (C.c_int, [C.POINTER(SceneDesc), C.c_int32, C.POINTER(C.c_void_p)]),
    "b200pt_scene_get_desc": (C.c_int, [C.c_void_p, C.POINTER(SceneDesc)]),
    "b200pt_scene_bvh_info": (C.c_int, [C.c_void_p, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_float)]),
    "b200pt_scene_destroy": (None, [C.c_void_p]),
    "b200pt_trace_batch": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p]),
    "b200pt_trace_batch_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_int32,
                                            C.POINTER(C.c_float)]),
    "b200pt_trace_count": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.POINTER(C.c_uint64)]),
    "b200pt_philox4x32_10": (C.c_int, [C.c_int32, C.c_void_p, C.c_size_t, C.c_void_p]),
    "b200pt_render_start": (C.c_int, [C.c_void_p, C.POINTER(Config), C.POINTER(C.c_void_p)]),
    "b200pt_render_poll": (C.c_int, [C.c_void_p, C.c_int32, C.POINTER(Progress)]),
    "b200pt_render_wait": (C.c_int, [C.c_void_p, C.POINTER(Progress)]),
    "b200pt_render_read": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(C.c_int64)]),
    "b200pt_render_read_stats": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p]),
    "b200pt_render_device_buffer": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_size_t)]),
    "b200pt_render_write": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64]),
    "b200pt_render_tonemap": (C.c_int, [C.c_void_p, C.c_int64, C.c_void_p, C.POINTER(C.c_float), C.POINTER(C.c_float)]),
    "b200pt_tonemap_static": (C.c_int, [C.c_int32, C.c_void_p, C.c_int32, C.c_int32, C.c_int64, C.c_float, C.c_void_p,
                                        C.c_void_p, C.POINTER(C.c_float), C.POINTER(C.c_float)]),
    "b200pt_render_write_ppm": (C.c_int, [C.c_void_p, C.c_int64, C.c_char_p]),
    "b200pt_render_paths": (C.c_int, [C.c_void_p, C.c_int32, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.c_void_p,
                                      C.c_void_p, C.c_void_p, C.c_void_p]),
    "b200pt_render_write_paths": (C.c_int, [C.c_void_p, C.c_char_p, C.c_char_p, C.POINTER(C.c_int32)]),
    "b200pt_camera_project": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p]),
    "b200pt_config_string": (C.c_int, [C.c_char_p, C.c_char_p, C.c_char_p, C.c_size_t]),
    "b200pt_render_profile": (C.c_int, [C.c_void_p, C.POINTER(Profile)]),
    "b200pt_render_destroy": (None, [C.c_void_p]),
    "b200pt_trim_cache": (None, []),
}

_lib = None


def lib():
    """Load libb200pt.so (once).  Raises when the CUDA extension has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(or `make -C pathtrace_b200/csrc`).  There is no CPU fallback.")
        handle = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)
            fn.restype = res
            fn.argtypes = args
        _lib = handle
    return _lib


class B200ptError(RuntimeError):
    def __init__(self, code, message):
        super().__init__(f"b200pt error {code}: {message}")
        self.code = code


def check(code):
    if code != OK:
        raise B200ptError(code, lib().b200pt_last_error().decode("utf-8", "replace"))
