"""ctypes mirror of include/prt_b200.h (struct layouts and prototypes).

The shared library `libprt_b200.so` is built in-tree by `__graft_entry__.build()` /
`pooraytracer_b200/csrc/Makefile`. Loading fails loudly when it is missing: there is no
Python or CPU fallback for any compute entry point.
"""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("PRT_B200_LIB") or os.path.join(HERE, "libprt_b200.so")  # override: A/B builds of the same ABI

PRT_OK = 0
PRT_ERR_INVALID_ARGUMENT = -1
PRT_ERR_NO_DEVICE = -2
PRT_ERR_CUDA = -3
PRT_ERR_OUT_OF_MEMORY = -4
PRT_ERR_NOT_COMMITTED = -5
PRT_ERR_UNSUPPORTED = -6

# Source/Material.h:48-51
MAT_LAMBERTIAN, MAT_PHONG, MAT_MIRROR, MAT_COOK_TORRANCE, MAT_DIFFUSE_LIGHT, MAT_DEBUG, MAT_EMPTY = range(7)

ACCEL_REFERENCE_BVH2 = 0
ACCEL_SAH_BVH2 = 1
ACCEL_SAH_BVH4 = 2
ACCEL_LBVH4_DEVICE = 3
ACCEL_SAH_BVH4Q = 4
ACCEL_AUTO = 5
NOALLOC_MAX_RAYS = 1048576  # PRT_NOALLOC_MAX_RAYS
REDUCE_NONE, REDUCE_NCCL, REDUCE_PEER = 0, 1, 2  # prt_reduce_kind

RENDER_EXACT_VISIBILITY = 1
RENDER_COUNTERS = 2
RENDER_STAGE_TIMING = 4
RENDER_PIXEL_JITTER = 8
RENDER_UNBIASED_LIGHT_PICK = 16
RENDER_LOBE_PROBABILITY = 32
RENDER_MIS = 64

LOAD_STRTOD = 1

c_double3 = C.c_double * 3


class PrtMaterial(C.Structure):
    _fields_ = [
        ("type", C.c_int32),
        ("kd_tex", C.c_int32),
        ("kd", c_double3),
        ("ks", c_double3),
        ("ns", C.c_double),
        ("emission", c_double3),
        ("alpha_x", C.c_double),
        ("alpha_y", C.c_double),
        ("eta", c_double3),
        ("k", c_double3),
    ]


class PrtTexture(C.Structure):
    _fields_ = [
        ("width", C.c_int32),
        ("height", C.c_int32),
        ("channels", C.c_int32),
        ("reserved", C.c_int32),
        ("texels", C.c_void_p),
    ]


class PrtSceneDesc(C.Structure):
    _fields_ = [
        ("n_triangles", C.c_int64),
        ("vertices", C.c_void_p),
        ("normals", C.c_void_p),
        ("texcoords", C.c_void_p),
        ("n_meshes", C.c_int32),
        ("mesh_first", C.c_void_p),
        ("mesh_material", C.c_void_p),
        ("n_materials", C.c_int32),
        ("materials", C.POINTER(PrtMaterial)),
        ("n_textures", C.c_int32),
        ("textures", C.POINTER(PrtTexture)),
    ]


class PrtBuildOpts(C.Structure):
    _fields_ = [("accel", C.c_int32), ("reserved", C.c_int32 * 7)]


class PrtCamera(C.Structure):
    _fields_ = [
        ("image_width", C.c_int32),
        ("image_height", C.c_int32),
        ("fovy", C.c_double),
        ("eye", c_double3),
        ("look_at", c_double3),
        ("up", c_double3),
    ]


class PrtRenderParams(C.Structure):
    _fields_ = [
        ("samples_per_pixel", C.c_int32),
        ("max_depth", C.c_int32),
        ("sample_lights", C.c_int32),
        ("spp_begin", C.c_int32),
        ("spp_end", C.c_int32),
        ("flags", C.c_uint32),
        ("russian_roulette", C.c_double),
        ("background", c_double3),
        ("rng_key", C.c_uint64),
    ]


class PrtStats(C.Structure):
    _fields_ = [
        ("extend_rays", C.c_uint64),
        ("shadow_rays", C.c_uint64),
        ("paths", C.c_uint64),
        ("node_visits", C.c_uint64),
        ("tri_tests", C.c_uint64),
        ("kernel_launches", C.c_uint64),
        ("gpu_ms", C.c_double),
        ("extend_ms", C.c_double),
        ("shade_ms", C.c_double),
        ("shadow_ms", C.c_double),
        ("node_bytes", C.c_int32),
        ("tri_bytes", C.c_int32),
        ("n_nodes", C.c_int64),
        ("n_triangles", C.c_int64),
        ("build_ms", C.c_double),
        ("ref_stack_overflows", C.c_uint64),
    ]


# name -> (restype, argtypes); every symbol include/prt_b200.h declares
PROTOTYPES = {
    "prt_abi_version": (C.c_int, []),
    "prt_context_create": (C.c_int, [C.c_int, C.POINTER(C.c_void_p)]),
    "prt_context_destroy": (None, [C.c_void_p]),
    "prt_last_error": (C.c_char_p, []),
    "prt_scene_create": (C.c_int, [C.c_void_p, C.POINTER(PrtSceneDesc), C.POINTER(C.c_void_p)]),
    "prt_scene_commit": (C.c_int, [C.c_void_p, C.POINTER(PrtBuildOpts)]),
    "prt_scene_destroy": (None, [C.c_void_p]),
    "prt_scene_clone": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(C.c_void_p)]),
    "prt_scene_reference_order": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(C.c_int64)]),
    "prt_trace_closest": (C.c_int, [C.c_void_p, C.c_int64] + [C.c_void_p] * 8),
    "prt_trace_any": (C.c_int, [C.c_void_p, C.c_int64] + [C.c_void_p] * 5),
    "prt_trace_closest_device": (C.c_int, [C.c_void_p, C.c_int64] + [C.c_void_p] * 9),
    "prt_trace_any_device": (C.c_int, [C.c_void_p, C.c_int64] + [C.c_void_p] * 6),
    "prt_trace_closest_device_noalloc": (C.c_int, [C.c_void_p, C.c_int64] + [C.c_void_p] * 9),
    "prt_sample_lights": (C.c_int, [C.c_void_p, C.c_int64] + [C.c_void_p] * 7),
    "prt_render": (C.c_int, [C.c_void_p, C.POINTER(PrtCamera), C.POINTER(PrtRenderParams), C.c_void_p]),
    "prt_render_device": (C.c_int, [C.c_void_p, C.POINTER(PrtCamera), C.POINTER(PrtRenderParams), C.c_void_p, C.c_void_p]),
    "prt_resolve_device": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p]),
    "prt_resolve": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p]),
    "prt_resolve_thresholds": (C.c_int, [C.c_void_p]),
    "prt_set_counters": (C.c_int, [C.c_void_p, C.c_int]),
    "prt_get_stats": (C.c_int, [C.c_void_p, C.POINTER(PrtStats)]),
    "prt_synchronize": (C.c_int, [C.c_void_p]),
    "prt_multi_create": (C.c_int, [C.c_void_p, C.c_int32, C.POINTER(PrtSceneDesc), C.POINTER(PrtBuildOpts), C.POINTER(C.c_void_p)]),
    "prt_multi_destroy": (None, [C.c_void_p]),
    "prt_multi_device_count": (C.c_int, [C.c_void_p]),
    "prt_multi_reduce_kind": (C.c_int, [C.c_void_p]),
    "prt_multi_render": (C.c_int, [C.c_void_p, C.POINTER(PrtCamera), C.POINTER(PrtRenderParams), C.c_void_p]),
    "prt_multi_resolve": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p]),
    "prt_multi_get_stats": (C.c_int, [C.c_void_p, C.POINTER(PrtStats), C.POINTER(C.c_double)]),
    "prt_multi_trace_closest": (C.c_int, [C.c_void_p, C.c_int64] + [C.c_void_p] * 8),
    "prt_multi_trace_any": (C.c_int, [C.c_void_p, C.c_int64] + [C.c_void_p] * 5),
    "prt_model_load": (C.c_int, [C.c_char_p, C.c_char_p, C.c_uint32, C.POINTER(C.c_void_p)]),
    "prt_model_destroy": (None, [C.c_void_p]),
    "prt_model_desc": (C.POINTER(PrtSceneDesc), [C.c_void_p]),
    "prt_model_camera": (C.c_int, [C.c_void_p, C.POINTER(PrtCamera)]),
    "prt_model_mesh_name": (C.c_char_p, [C.c_void_p, C.c_int32]),
    "prt_model_material_name": (C.c_char_p, [C.c_void_p, C.c_int32]),
    "prt_model_warnings": (C.c_char_p, [C.c_void_p]),
    "prt_write_png": (C.c_int, [C.c_char_p, C.c_int32, C.c_int32, C.c_void_p]),
    "prt_write_hdr": (C.c_int, [C.c_char_p, C.c_int32, C.c_int32, C.c_void_p]),
    "prt_image_load": (C.c_int, [C.c_char_p, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_void_p)]),
    "prt_image_free": (None, [C.c_void_p]),
}

_lib = None


def load_library(path: str | None = None) -> C.CDLL:
    """Load libprt_b200.so and attach prototypes. Raises if the extension was not built."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    p = path or LIB_PATH
    if not os.path.exists(p):
        raise RuntimeError(
            f"pooraytracer_b200: native library {p} is missing - run `python -c 'import __graft_entry__ as g; g.build()'` "
            "(there is no CPU fallback)")
    lib = C.CDLL(p)
    for name, (res, args) in PROTOTYPES.items():
        fn = getattr(lib, name)  # AttributeError = symbol missing = broken build
        fn.restype = res
        fn.argtypes = args
    if lib.prt_abi_version() != 2:
        raise RuntimeError("pooraytracer_b200: ABI version mismatch")
    if path is None:
        _lib = lib
    return lib
