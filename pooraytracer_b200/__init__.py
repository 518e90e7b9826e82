"""pooraytracer_b200 — B200-native (sm_100a) path-tracing backend for Pooraytracer.

Only what the render hot path needs: `csrc/` (CUDA kernels + the C-ABI library
`libprt_b200.so`, declared in include/prt_b200.h), a thin ctypes binding (`api`), and the
procedural scene / ray generators (`scenes`) that stand in for the reference's absent scene files.
"""
from . import _capi, scenes  # noqa: F401
from .api import Context, Model, MultiScene, Scene, PrtError, load_image, make_camera, make_params, make_scene_desc, write_hdr, write_png  # noqa: F401

__all__ = ["Context", "Model", "MultiScene", "Scene", "load_image", "write_png", "write_hdr", "PrtError", "make_camera", "make_params", "make_scene_desc", "scenes"]
