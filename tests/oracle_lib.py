"""TEST INFRASTRUCTURE: Python harness for the CPU checkers under oracle/.

  kind="ref"  oracle/_ref/libprt_ref.so      — unmodified reference sources (oracle/ref_driver.cpp)
  kind="orc"  oracle/libprt_oracle.so        — our CPU restatement (oracle/prt_oracle.cpp)

Both export the interface of oracle/oracle_api.h under their prefix. Nothing in
pooraytracer_b200/ imports this module.
"""
from __future__ import annotations

import ctypes as C
import os
import tempfile

import numpy as np

from pooraytracer_b200 import _capi as capi
from pooraytracer_b200.api import make_scene_desc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PATHS = {
    "ref": os.path.join(ROOT, "oracle", "_ref", "libprt_ref.so"),
    "ref_shipped": os.path.join(ROOT, "oracle", "_ref", "libprt_ref_shipped.so"),
    "orc": os.path.join(ROOT, "oracle", "libprt_oracle.so"),
}

PER_TRIANGLE_MATERIAL, UV_TAG, DEFINE_UB = 1, 2, 4
RNG_COUNTER, RNG_THREAD_LOCAL, RNG_LIBC = 0, 1, 2


class RenderInfo(C.Structure):
    _fields_ = [("world_hit_calls", C.c_uint64), ("shadow_calls", C.c_uint64), ("paths", C.c_uint64),
                ("seconds", C.c_double), ("threads", C.c_int32), ("reserved", C.c_int32)]


def available(kind: str) -> bool:
    return os.path.exists(PATHS[kind])


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class OracleLib:
    _cache = {}

    def __init__(self, kind: str):
        self.kind = kind
        self.prefix = "orc" if kind == "orc" else "ref"
        if kind not in OracleLib._cache:
            OracleLib._cache[kind] = C.CDLL(PATHS[kind])
        self.lib = OracleLib._cache[kind]

    def fn(self, name, restype=C.c_int):
        f = getattr(self.lib, f"{self.prefix}_{name}")
        f.restype = restype
        return f


class OracleScene:
    """A scene inside one of the CPU checkers."""

    def __init__(self, kind: str, data, flags: int = 0):
        self.o = OracleLib(kind)
        self.data = data
        desc, keep = make_scene_desc(data)
        self._tmp = None
        paths = (C.c_char_p * max(1, len(data.textures)))()
        if data.textures:
            # the reference only loads textures from files (Source/Texture.cpp:10-21): write PNGs
            from PIL import Image
            self._tmp = tempfile.TemporaryDirectory(prefix="prt_tex_")
            for i, t in enumerate(data.textures):
                if t is None:
                    paths[i] = None
                    continue
                p = os.path.join(self._tmp.name, f"tex{i}.png")
                Image.fromarray(t).save(p)
                paths[i] = p.encode()
        h = C.c_void_p()
        rc = self.o.fn("scene_create")(C.byref(desc), paths, C.c_int(flags), C.byref(h))
        if rc != 0:
            raise RuntimeError(f"{kind}_scene_create failed: {rc}")
        self.handle = h

    def close(self):
        if getattr(self, "handle", None):
            self.o.fn("scene_destroy", None)(self.handle)
            self.handle = None
        if self._tmp is not None:
            self._tmp.cleanup()
            self._tmp = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def trace_closest(self, o, d, tmin=1e-4, tmax=np.inf, nthreads=8):
        o = np.ascontiguousarray(o, dtype=np.float64).reshape(-1, 3)
        d = np.ascontiguousarray(d, dtype=np.float64).reshape(-1, 3)
        n = o.shape[0]
        tmin = np.ascontiguousarray(np.broadcast_to(np.asarray(tmin, dtype=np.float64), (n,)))
        tmax = np.ascontiguousarray(np.broadcast_to(np.asarray(tmax, dtype=np.float64), (n,)))
        prim = np.empty(n, np.int32); t = np.zeros(n); a = np.zeros(n); b = np.zeros(n)
        rc = self.o.fn("trace_closest")(self.handle, C.c_int64(n), _ptr(o), _ptr(d), _ptr(tmin), _ptr(tmax),
                                        _ptr(prim), _ptr(t), _ptr(a), _ptr(b), C.c_int(nthreads))
        assert rc == 0, rc
        return prim, t, a, b

    def reference_order(self):
        n = self.data.n_triangles
        rank = np.empty(n, np.int64)
        nl = C.c_int64(0)
        rc = self.o.fn("reference_order")(self.handle, _ptr(rank), None, C.byref(nl))
        assert rc == 0, rc
        lo = np.empty(nl.value, np.int64)
        rc = self.o.fn("reference_order")(self.handle, None, _ptr(lo), C.byref(nl))
        assert rc == 0, rc
        return rank, lo

    def sample_lights(self, origins, u):
        origins = np.ascontiguousarray(origins, dtype=np.float64).reshape(-1, 3)
        u = np.ascontiguousarray(u, dtype=np.float64).reshape(-1, 4)
        n = origins.shape[0]
        pos = np.empty((n, 3)); nrm = np.empty((n, 3)); front = np.empty(n, np.uint8)
        pdf = np.empty(n); prim = np.empty(n, np.int32)
        rc = self.o.fn("sample_lights")(self.handle, C.c_int64(n), _ptr(origins), _ptr(u), _ptr(pos), _ptr(nrm),
                                        _ptr(front), _ptr(pdf), _ptr(prim))
        assert rc == 0, rc
        return pos, nrm, front, pdf, prim

    def render(self, camera, params, rng_mode=RNG_COUNTER, nthreads=8):
        img = np.zeros((camera.image_height, camera.image_width, 3))
        info = RenderInfo()
        rc = self.o.fn("render")(self.handle, C.byref(camera), C.byref(params), C.c_int(rng_mode), C.c_int(nthreads),
                                 _ptr(img), C.byref(info))
        assert rc == 0, rc
        return img, info

    def pixel_samples(self, camera, params, px, py):
        sb, se = params.spp_begin, params.spp_end
        if sb == 0 and se == 0:
            se = params.samples_per_pixel
        out = np.zeros((se - sb, 3))
        rc = self.o.fn("pixel_samples")(self.handle, C.byref(camera), C.byref(params), C.c_int32(px), C.c_int32(py), _ptr(out))
        assert rc == 0, rc
        return out

    def triangle_record(self, prim):
        out = np.zeros(20)
        rc = self.o.fn("triangle_record")(self.handle, C.c_int64(prim), _ptr(out))
        assert rc == 0, rc
        return out

    def material_scatter(self, material, ray_dir, normal, tangent, uv, u):
        u = np.ascontiguousarray(u, dtype=np.float64)
        att = np.zeros(3); od = np.zeros(3)
        nd = C.c_int32(0); sc = C.c_int32(0)
        rc = self.o.fn("material_scatter")(
            self.handle, C.c_int32(material), _ptr(np.ascontiguousarray(ray_dir, dtype=np.float64)),
            _ptr(np.ascontiguousarray(normal, dtype=np.float64)), _ptr(np.ascontiguousarray(tangent, dtype=np.float64)),
            _ptr(np.ascontiguousarray(uv, dtype=np.float64)), _ptr(u), C.c_int32(len(u)), _ptr(att), _ptr(od),
            C.byref(nd), C.byref(sc))
        assert rc == 0, rc
        return att, od, nd.value, bool(sc.value)

    def material_eval(self, material, ray_dir, normal, tangent, uv, wi_world, u):
        u = np.ascontiguousarray(u, dtype=np.float64)
        f = np.zeros(3)
        c = lambda a: _ptr(np.ascontiguousarray(a, dtype=np.float64))  # noqa: E731
        rc = self.o.fn("material_eval")(self.handle, C.c_int32(material), c(ray_dir), c(normal), c(tangent), c(uv), c(wi_world),
                                        _ptr(u), C.c_int32(len(u)), _ptr(f))
        assert rc == 0, rc
        return f

    def texture_value(self, texture, uv):
        uv = np.ascontiguousarray(uv, dtype=np.float64).reshape(-1, 2)
        out = np.zeros((uv.shape[0], 3))
        rc = self.o.fn("texture_value")(self.handle, C.c_int32(texture), C.c_int64(uv.shape[0]), _ptr(uv), _ptr(out))
        assert rc == 0, rc
        return out


def camera_ray(kind, camera, i, j):
    o = OracleLib(kind)
    org = np.zeros(3); d = np.zeros(3)
    rc = o.fn("camera_ray")(C.byref(camera), C.c_int32(i), C.c_int32(j), _ptr(org), _ptr(d))
    assert rc == 0, rc
    return org, d
