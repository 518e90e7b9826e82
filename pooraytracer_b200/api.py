"""Host-side Python binding of the C ABI (include/prt_b200.h).

This is plumbing for tests and bench.py: numpy arrays in, numpy arrays out, every call going
through `libprt_b200.so`. The product's host logic (BVH build, flattening, wavefront scheduling)
lives in the C++/CUDA library, not here.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _capi as capi
from .scenes import SceneData


class PrtError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"prt error {code}: {msg}")
        self.code = code


def _check(lib, rc):
    if rc != capi.PRT_OK:
        raise PrtError(rc, lib.prt_last_error().decode("utf-8", "replace"))


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def make_scene_desc(scene: SceneData):
    """Returns (PrtSceneDesc, keepalive list). Shared by the product binding and the oracle harness."""
    keep = []
    v = np.ascontiguousarray(scene.vertices, dtype=np.float64); keep.append(v)
    uv = None if scene.texcoords is None else np.ascontiguousarray(scene.texcoords, dtype=np.float64); keep.append(uv)
    nrm = None if scene.normals is None else np.ascontiguousarray(scene.normals, dtype=np.float64); keep.append(nrm)
    first = np.ascontiguousarray(scene.mesh_first, dtype=np.int64); keep.append(first)
    mm = np.ascontiguousarray(scene.mesh_material, dtype=np.int32); keep.append(mm)
    mats = (capi.PrtMaterial * max(1, len(scene.materials)))()
    for i, m in enumerate(scene.materials):
        mats[i].type = m.type
        mats[i].kd_tex = m.kd_tex
        mats[i].kd[:] = m.kd
        mats[i].ks[:] = m.ks
        mats[i].ns = m.ns
        mats[i].emission[:] = m.emission
        mats[i].alpha_x = m.alpha_x
        mats[i].alpha_y = m.alpha_y
        mats[i].eta[:] = m.eta
        mats[i].k[:] = m.k
    keep.append(mats)
    texs = (capi.PrtTexture * max(1, len(scene.textures)))()
    for i, t in enumerate(scene.textures):
        if t is None:
            texs[i].width = texs[i].height = texs[i].channels = 0
            texs[i].texels = None
            continue
        t = np.ascontiguousarray(t, dtype=np.uint8); keep.append(t)
        texs[i].height, texs[i].width = t.shape[0], t.shape[1]
        texs[i].channels = 1 if t.ndim == 2 else t.shape[2]
        texs[i].texels = t.ctypes.data
    keep.append(texs)
    d = capi.PrtSceneDesc()
    d.n_triangles = v.shape[0]
    d.vertices = v.ctypes.data
    d.normals = None if nrm is None else nrm.ctypes.data
    d.texcoords = None if uv is None else uv.ctypes.data
    d.n_meshes = len(mm)
    d.mesh_first = first.ctypes.data
    d.mesh_material = mm.ctypes.data
    d.n_materials = len(scene.materials)
    d.materials = mats
    d.n_textures = len(scene.textures)
    d.textures = texs
    return d, keep


def make_camera(cam: dict, width=None, height=None) -> capi.PrtCamera:
    c = capi.PrtCamera()
    c.image_width = int(width or cam["width"])
    c.image_height = int(height or cam["height"])
    c.fovy = float(cam["fovy"])
    c.eye[:] = cam["eye"]
    c.look_at[:] = cam["look_at"]
    c.up[:] = cam["up"]
    return c


def make_params(spp, max_depth, rr=0.8, sample_lights=True, background=(0.0, 0.0, 0.0), rng_key=0,
                spp_begin=0, spp_end=0, flags=0) -> capi.PrtRenderParams:
    p = capi.PrtRenderParams()
    p.samples_per_pixel = int(spp)
    p.max_depth = int(max_depth)
    p.sample_lights = 1 if sample_lights else 0
    p.spp_begin, p.spp_end = int(spp_begin), int(spp_end)
    p.flags = int(flags)
    p.russian_roulette = float(rr)
    p.background[:] = background
    p.rng_key = int(rng_key)
    return p


class Context:
    """One context = one GPU (device < 0: host-only, for scene building / introspection)."""

    def __init__(self, device: int = 0):
        self.lib = capi.load_library()
        h = C.c_void_p()
        _check(self.lib, self.lib.prt_context_create(int(device), C.byref(h)))
        self.handle = h
        self.device = device

    def synchronize(self):
        _check(self.lib, self.lib.prt_synchronize(self.handle))

    def close(self):
        if getattr(self, "handle", None):
            self.lib.prt_context_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Model:
    """Scene files -> flat arrays: Model::Model + Camera::SetViewParametersByXmlFile
    (Source/Model.cpp:53-394, Source/Camera.cpp:339-389) through prt_model_load."""

    def __init__(self, directory: str, name: str, flags: int = 0):
        self.lib = capi.load_library()
        h = C.c_void_p()
        _check(self.lib, self.lib.prt_model_load(directory.encode(), name.encode(), int(flags), C.byref(h)))
        self.handle = h
        self.desc = self.lib.prt_model_desc(h).contents
        self.n_triangles = int(self.desc.n_triangles)

    def _array(self, ptr, shape, dtype):
        n = int(np.prod(shape))
        if not ptr or n == 0:
            return np.zeros(shape, dtype=dtype)
        buf = (C.c_char * (n * np.dtype(dtype).itemsize)).from_address(ptr)
        return np.frombuffer(buf, dtype=dtype).reshape(shape).copy()

    @property
    def vertices(self):
        return self._array(self.desc.vertices, (self.n_triangles, 3, 3), np.float64)

    @property
    def normals(self):
        return self._array(self.desc.normals, (self.n_triangles, 3, 3), np.float64)

    @property
    def texcoords(self):
        return self._array(self.desc.texcoords, (self.n_triangles, 3, 2), np.float64)

    @property
    def mesh_first(self):
        return self._array(self.desc.mesh_first, (self.desc.n_meshes + 1,), np.int64)

    @property
    def mesh_material(self):
        return self._array(self.desc.mesh_material, (self.desc.n_meshes,), np.int32)

    @property
    def mesh_names(self):
        return [self.lib.prt_model_mesh_name(self.handle, i).decode() for i in range(self.desc.n_meshes)]

    @property
    def material_names(self):
        return [self.lib.prt_model_material_name(self.handle, i).decode() for i in range(self.desc.n_materials)]

    @property
    def materials(self):
        return [capi.PrtMaterial.from_buffer_copy(self.desc.materials[i]) for i in range(self.desc.n_materials)]  # copies: outlive the model

    def texture(self, i):
        t = self.desc.textures[i]
        if not t.texels:
            return None
        return self._array(t.texels, (t.height, t.width, t.channels), np.uint8)

    @property
    def warnings(self):
        return [w for w in self.lib.prt_model_warnings(self.handle).decode("utf-8", "replace").split("\n") if w]

    def camera(self) -> capi.PrtCamera:
        c = capi.PrtCamera()
        _check(self.lib, self.lib.prt_model_camera(self.handle, C.byref(c)))
        return c

    def close(self):
        if getattr(self, "handle", None):
            self.lib.prt_model_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def write_png(path: str, rgb8: np.ndarray):
    lib = capi.load_library()
    a = np.ascontiguousarray(rgb8, dtype=np.uint8)
    _check(lib, lib.prt_write_png(path.encode(), a.shape[1], a.shape[0], _ptr(a)))


def write_hdr(path: str, rgb32f: np.ndarray):
    lib = capi.load_library()
    a = np.ascontiguousarray(rgb32f, dtype=np.float32)
    _check(lib, lib.prt_write_hdr(path.encode(), a.shape[1], a.shape[0], _ptr(a)))


def load_image(path: str) -> np.ndarray:
    """stbi_load(path, &w, &h, &c, 0) -> uint8 [h, w, c]"""
    lib = capi.load_library()
    w, h, c, px = C.c_int32(), C.c_int32(), C.c_int32(), C.c_void_p()
    _check(lib, lib.prt_image_load(path.encode(), C.byref(w), C.byref(h), C.byref(c), C.byref(px)))
    try:
        n = w.value * h.value * c.value
        out = np.frombuffer((C.c_char * n).from_address(px.value), dtype=np.uint8).reshape(h.value, w.value, c.value).copy()
    finally:
        lib.prt_image_free(px)
    return out


class Scene:
    def __init__(self, ctx: Context, data, accel: int | None = capi.ACCEL_SAH_BVH2):
        """data: a SceneData (procedural arrays) or a Model (loaded scene files)"""
        self.ctx, self.lib, self.data = ctx, ctx.lib, data
        h = C.c_void_p()
        if isinstance(data, Model):
            _check(self.lib, self.lib.prt_scene_create(ctx.handle, C.byref(data.desc), C.byref(h)))
        else:
            desc, keep = make_scene_desc(data)
            _check(self.lib, self.lib.prt_scene_create(ctx.handle, C.byref(desc), C.byref(h)))
            del keep
        self.handle = h
        if accel is not None:
            self.commit(accel)

    def clone(self, ctx: "Context") -> "Scene":
        """prt_scene_clone: a replica of this committed scene on ctx's GPU (device-to-device copies, no host rebuild)"""
        h = C.c_void_p()
        _check(self.lib, self.lib.prt_scene_clone(self.handle, ctx.handle, C.byref(h)))
        other = Scene.__new__(Scene)
        other.ctx, other.lib, other.data, other.handle = ctx, self.lib, self.data, h
        return other

    def commit(self, accel: int, max_leaf: int = 0):
        """prt_scene_commit; max_leaf = triangles per leaf of the SAH / LBVH trees (0: the default, 4)"""
        o = capi.PrtBuildOpts()
        o.accel = int(accel)
        o.reserved[0] = int(max_leaf)
        _check(self.lib, self.lib.prt_scene_commit(self.handle, C.byref(o)))

    def close(self):
        if getattr(self, "handle", None):
            self.lib.prt_scene_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- host introspection ----------------------------------------------------------------------
    def reference_order(self):
        n = self.data.n_triangles
        rank = np.empty(n, dtype=np.int64)
        nl = C.c_int64(0)
        _check(self.lib, self.lib.prt_scene_reference_order(self.handle, _ptr(rank), None, C.byref(nl)))
        lo = np.empty(nl.value, dtype=np.int64)
        _check(self.lib, self.lib.prt_scene_reference_order(self.handle, None, _ptr(lo), C.byref(nl)))
        return rank, lo

    # -- ray queries -----------------------------------------------------------------------------
    @staticmethod
    def _rays(o, d, tmin, tmax):
        o = np.ascontiguousarray(o, dtype=np.float64).reshape(-1, 3)
        d = np.ascontiguousarray(d, dtype=np.float64).reshape(-1, 3)
        n = o.shape[0]
        tmin = np.ascontiguousarray(np.broadcast_to(np.asarray(tmin, dtype=np.float64), (n,)))
        tmax = np.ascontiguousarray(np.broadcast_to(np.asarray(tmax, dtype=np.float64), (n,)))
        return n, o, d, tmin, tmax

    def trace_closest(self, o, d, tmin=1e-4, tmax=np.inf):
        n, o, d, tmin, tmax = self._rays(o, d, tmin, tmax)
        prim = np.empty(n, dtype=np.int32)
        t = np.zeros(n); a = np.zeros(n); b = np.zeros(n)
        _check(self.lib, self.lib.prt_trace_closest(self.handle, n, _ptr(o), _ptr(d), _ptr(tmin), _ptr(tmax),
                                                    _ptr(prim), _ptr(t), _ptr(a), _ptr(b)))
        return prim, t, a, b

    def trace_any(self, o, d, tmin=1e-3, tmax=np.inf):
        n, o, d, tmin, tmax = self._rays(o, d, tmin, tmax)
        hit = np.empty(n, dtype=np.uint8)
        _check(self.lib, self.lib.prt_trace_any(self.handle, n, _ptr(o), _ptr(d), _ptr(tmin), _ptr(tmax), _ptr(hit)))
        return hit

    def sample_lights(self, origins, u):
        origins = np.ascontiguousarray(origins, dtype=np.float64).reshape(-1, 3)
        u = np.ascontiguousarray(u, dtype=np.float64).reshape(-1, 4)
        n = origins.shape[0]
        pos = np.empty((n, 3)); nrm = np.empty((n, 3)); front = np.empty(n, np.uint8)
        pdf = np.empty(n); prim = np.empty(n, np.int32)
        _check(self.lib, self.lib.prt_sample_lights(self.handle, n, _ptr(origins), _ptr(u), _ptr(pos), _ptr(nrm),
                                                    _ptr(front), _ptr(pdf), _ptr(prim)))
        return pos, nrm, front, pdf, prim

    # -- rendering -------------------------------------------------------------------------------
    def render(self, camera: capi.PrtCamera, params: capi.PrtRenderParams) -> np.ndarray:
        img = np.empty((camera.image_height, camera.image_width, 3), dtype=np.float64)
        _check(self.lib, self.lib.prt_render(self.handle, C.byref(camera), C.byref(params), _ptr(img)))
        return img

    def render_device(self, camera, params, device_ptr: int, stream: int | None = None):
        _check(self.lib, self.lib.prt_render_device(self.handle, C.byref(camera), C.byref(params),
                                                    C.c_void_p(device_ptr), C.c_void_p(stream or 0)))

    def resolve_device(self, device_ptr: int, width: int, height: int, want_float=True):
        rgb8 = np.empty((height, width, 3), dtype=np.uint8)
        f32 = np.empty((height, width, 3), dtype=np.float32) if want_float else None
        _check(self.lib, self.lib.prt_resolve_device(self.handle, C.c_void_p(device_ptr), width, height, _ptr(rgb8), _ptr(f32)))
        return rgb8, f32

    def resolve(self, rgb: np.ndarray, want_float=True):
        """prt_resolve: host float64 [H, W, 3] -> (uint8 [H, W, 3], float32 [H, W, 3])"""
        rgb = np.ascontiguousarray(rgb, dtype=np.float64)
        h, w, _ = rgb.shape
        rgb8 = np.empty((h, w, 3), dtype=np.uint8)
        f32 = np.empty((h, w, 3), dtype=np.float32) if want_float else None
        _check(self.lib, self.lib.prt_resolve(self.handle, _ptr(rgb), w, h, _ptr(rgb8), _ptr(f32)))
        return rgb8, f32

    def set_counters(self, enable: bool):
        _check(self.lib, self.lib.prt_set_counters(self.handle, 1 if enable else 0))

    def stats(self) -> capi.PrtStats:
        s = capi.PrtStats()
        _check(self.lib, self.lib.prt_get_stats(self.handle, C.byref(s)))
        return s


class MultiScene:
    """prt_multi: one scene replicated on several GPUs of the box, samples sharded, one reduction of the frames to the
    first GPU (peer loads over NVLink, or NCCL: `reduce_kind`). Harness mirror of include/prt_b200.h's prt_multi_* calls."""

    def __init__(self, devices, data, accel: int = capi.ACCEL_SAH_BVH4, max_leaf: int = 0):
        self.lib = capi.load_library()
        self.data = data
        devs = np.ascontiguousarray(list(devices), dtype=np.int32)
        o = capi.PrtBuildOpts()
        o.accel = int(accel)
        o.reserved[0] = int(max_leaf)
        h = C.c_void_p()
        if isinstance(data, Model):
            _check(self.lib, self.lib.prt_multi_create(_ptr(devs), len(devs), C.byref(data.desc), C.byref(o), C.byref(h)))
        else:
            desc, keep = make_scene_desc(data)
            _check(self.lib, self.lib.prt_multi_create(_ptr(devs), len(devs), C.byref(desc), C.byref(o), C.byref(h)))
            del keep
        self.handle = h
        self.n = len(devs)

    @property
    def reduce_kind(self) -> str:
        return {capi.REDUCE_NONE: "none", capi.REDUCE_NCCL: "nccl", capi.REDUCE_PEER: "peer"}[self.lib.prt_multi_reduce_kind(self.handle)]

    def close(self):
        if getattr(self, "handle", None):
            self.lib.prt_multi_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def render(self, camera: capi.PrtCamera, params: capi.PrtRenderParams, out: np.ndarray | None = None) -> np.ndarray:
        img = out if out is not None else np.empty((camera.image_height, camera.image_width, 3), dtype=np.float64)
        _check(self.lib, self.lib.prt_multi_render(self.handle, C.byref(camera), C.byref(params), _ptr(img)))
        return img

    def resolve(self, width: int, height: int, want_float=True):
        rgb8 = np.empty((height, width, 3), dtype=np.uint8)
        f32 = np.empty((height, width, 3), dtype=np.float32) if want_float else None
        _check(self.lib, self.lib.prt_multi_resolve(self.handle, _ptr(rgb8), _ptr(f32)))
        return rgb8, f32

    def stats(self):
        s = capi.PrtStats()
        ms = C.c_double(0.0)
        _check(self.lib, self.lib.prt_multi_get_stats(self.handle, C.byref(s), C.byref(ms)))
        return s, ms.value

    def trace_closest(self, o, d, tmin=1e-4, tmax=np.inf):
        n, o, d, tmin, tmax = Scene._rays(o, d, tmin, tmax)
        prim = np.empty(n, dtype=np.int32)
        t = np.zeros(n); a = np.zeros(n); b = np.zeros(n)
        _check(self.lib, self.lib.prt_multi_trace_closest(self.handle, n, _ptr(o), _ptr(d), _ptr(tmin), _ptr(tmax),
                                                          _ptr(prim), _ptr(t), _ptr(a), _ptr(b)))
        return prim, t, a, b

    def trace_any(self, o, d, tmin=1e-3, tmax=np.inf):
        n, o, d, tmin, tmax = Scene._rays(o, d, tmin, tmax)
        hit = np.empty(n, dtype=np.uint8)
        _check(self.lib, self.lib.prt_multi_trace_any(self.handle, n, _ptr(o), _ptr(d), _ptr(tmin), _ptr(tmax), _ptr(hit)))
        return hit
