"""Procedural stand-ins for the reference's scenes and the ray-query microbench inputs.

The reference renders `example-scenes-cg24/{cornell-box,veach-mis,bathroom2}` (main.cpp:13-17), but
the scene files are not in the repository (`*.obj` is git-ignored there), so every configuration of
BASELINE.json is generated here — deterministically, from an integer hash (splitmix64), never from
library RNG state — as the flat triangle/mesh/material arrays `prt_scene_create` consumes. Material
NAMES follow the reference's name→type table (Source/Model.cpp:16-51) so that the same scenes
written as OBJ/MTL/XML go through the loader path unchanged.

Conventions: right-handed, +y up; a light's geometric normal must face the receiver (NEE is
one-sided: Source/Camera.cpp:154, Source/Triangle.cpp:89-90).
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field

import numpy as np

from . import _capi as capi

_GOLDEN = np.uint64(0x9E3779B97F4A7C15)


def _mix64(z: np.ndarray) -> np.ndarray:
    z = z.astype(np.uint64, copy=True)
    with np.errstate(over="ignore"):
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
    return z ^ (z >> np.uint64(31))


def hash_uniform(n: int, seed: int, stream: int = 0, offset: int = 0) -> np.ndarray:
    """n doubles in [0,1): u[i] = mix64(mix64(seed, stream) + (offset+i+1)*GOLDEN) >> 11 * 2^-53."""
    with np.errstate(over="ignore"):
        base = _mix64(np.array([np.uint64(seed) * np.uint64(0x100000001B3) + np.uint64(stream)], dtype=np.uint64))[0]
        idx = (np.arange(offset + 1, offset + n + 1, dtype=np.uint64)) * _GOLDEN + base
    return (_mix64(idx) >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)


@dataclass
class Material:
    name: str
    type: int
    kd: tuple = (0.5, 0.5, 0.5)
    ks: tuple = (0.0, 0.0, 0.0)
    ns: float = 1.0
    emission: tuple = (0.0, 0.0, 0.0)
    kd_tex: int = -1
    alpha_x: float = 0.3
    alpha_y: float = 0.3
    eta: tuple = (0.1, 0.5, 1.5)   # Source/Model.cpp:309 (Au)
    k: tuple = (4.0, 0.02, 0.3)


@dataclass
class SceneData:
    name: str
    vertices: np.ndarray                 # [n,3,3] float64
    texcoords: np.ndarray | None         # [n,3,2] float64
    mesh_first: np.ndarray               # [m+1] int64
    mesh_material: np.ndarray            # [m] int32
    mesh_names: list
    materials: list
    textures: list = field(default_factory=list)   # uint8 [H,W,C] (or [H,W])
    camera: dict = field(default_factory=dict)
    normals: np.ndarray | None = None

    @property
    def n_triangles(self) -> int:
        return int(self.vertices.shape[0])


class _Builder:
    def __init__(self, name):
        self.name = name
        self.v, self.uv = [], []
        self.first = [0]
        self.mesh_mat, self.mesh_names = [], []
        self.materials, self.textures = [], []
        self._count = 0

    def material(self, m: Material) -> int:
        self.materials.append(m)
        return len(self.materials) - 1

    def texture(self, img: np.ndarray) -> int:
        self.textures.append(np.ascontiguousarray(img, dtype=np.uint8))
        return len(self.textures) - 1

    def mesh(self, name: str, mat: int, v: np.ndarray, uv: np.ndarray | None = None):
        v = np.asarray(v, dtype=np.float64).reshape(-1, 3, 3)
        if uv is None:
            uv = np.tile(np.array([[0.0, 0.0], [1.0, 0.0], [1.0, 1.0]]), (v.shape[0], 1, 1))
        uv = np.asarray(uv, dtype=np.float64).reshape(-1, 3, 2)
        assert uv.shape[0] == v.shape[0]
        self.v.append(v)
        self.uv.append(apply_degenerate_uv_rule(uv))
        self._count += v.shape[0]
        self.first.append(self._count)
        self.mesh_mat.append(mat)
        self.mesh_names.append(name)

    def build(self, camera: dict) -> SceneData:
        return SceneData(
            name=self.name,
            vertices=np.ascontiguousarray(np.concatenate(self.v, axis=0)),
            texcoords=np.ascontiguousarray(np.concatenate(self.uv, axis=0)),
            mesh_first=np.array(self.first, dtype=np.int64),
            mesh_material=np.array(self.mesh_mat, dtype=np.int32),
            mesh_names=self.mesh_names,
            materials=self.materials,
            textures=self.textures,
            camera=camera,
        )


def apply_degenerate_uv_rule(uv: np.ndarray) -> np.ndarray:
    """Source/Model.cpp:170-175: a triangle with any two equal texcoords gets (0,0),(1,0),(1,1)."""
    uv = uv.copy()
    e01 = np.all(uv[:, 0] == uv[:, 1], axis=1)
    e12 = np.all(uv[:, 1] == uv[:, 2], axis=1)
    e02 = np.all(uv[:, 0] == uv[:, 2], axis=1)
    bad = e01 | e12 | e02
    uv[bad] = np.array([[0.0, 0.0], [1.0, 0.0], [1.0, 1.0]])
    return uv


# ------------------------------------------------------------------------------------------------
# geometry helpers (all return (vertices [n,3,3], texcoords [n,3,2]))
# ------------------------------------------------------------------------------------------------
def quad(p0, p1, p2, p3):
    """Two triangles (p0,p1,p2), (p0,p2,p3); geometric normal = (p1-p0) x (p2-p0)."""
    p0, p1, p2, p3 = (np.asarray(p, dtype=np.float64) for p in (p0, p1, p2, p3))
    v = np.stack([np.stack([p0, p1, p2]), np.stack([p0, p2, p3])])
    uv = np.array([[[0, 0], [1, 0], [1, 1]], [[0, 0], [1, 1], [0, 1]]], dtype=np.float64)
    return v, uv


def grid(origin, du, dv, nu, nv, height=None, uv_scale=(1.0, 1.0)):
    """nu x nv cells over origin + s*du + t*dv, optional displacement along normalize(du x dv)."""
    origin, du, dv = (np.asarray(p, dtype=np.float64) for p in (origin, du, dv))
    s = np.arange(nu + 1, dtype=np.float64) / nu
    t = np.arange(nv + 1, dtype=np.float64) / nv
    S, T = np.meshgrid(s, t, indexing="ij")
    P = origin + S[..., None] * du + T[..., None] * dv
    if height is not None:
        n = np.cross(du, dv)
        n = n / np.linalg.norm(n)
        P = P + np.asarray(height, dtype=np.float64)[..., None] * n
    UV = np.stack([S * uv_scale[0], T * uv_scale[1]], axis=-1)
    a, b, c, d = P[:-1, :-1], P[1:, :-1], P[1:, 1:], P[:-1, 1:]
    ua, ub, uc, ud = UV[:-1, :-1], UV[1:, :-1], UV[1:, 1:], UV[:-1, 1:]
    v = np.stack([np.stack([a, b, c], axis=2), np.stack([a, c, d], axis=2)], axis=2).reshape(-1, 3, 3)
    uv = np.stack([np.stack([ua, ub, uc], axis=2), np.stack([ua, uc, ud], axis=2)], axis=2).reshape(-1, 3, 2)
    return v, uv


def box(center, half, rot_y_deg=0.0):
    """Axis-aligned box rotated about +y; outward normals."""
    cx, cy, cz = center
    hx, hy, hz = half
    c, s = math.cos(math.radians(rot_y_deg)), math.sin(math.radians(rot_y_deg))

    def P(x, y, z):
        return (cx + c * x + s * z, cy + y, cz - s * x + c * z)

    corners = {(i, j, k): P(i * hx, j * hy, k * hz) for i in (-1, 1) for j in (-1, 1) for k in (-1, 1)}
    faces = [
        [(1, -1, -1), (1, 1, -1), (1, 1, 1), (1, -1, 1)],      # +x
        [(-1, -1, 1), (-1, 1, 1), (-1, 1, -1), (-1, -1, -1)],  # -x
        [(-1, 1, -1), (-1, 1, 1), (1, 1, 1), (1, 1, -1)],      # +y
        [(-1, -1, 1), (-1, -1, -1), (1, -1, -1), (1, -1, 1)],  # -y
        [(-1, -1, 1), (1, -1, 1), (1, 1, 1), (-1, 1, 1)],      # +z
        [(1, -1, -1), (-1, -1, -1), (-1, 1, -1), (1, 1, -1)],  # -z
    ]
    vs, uvs = [], []
    for f in faces:
        v, uv = quad(*(corners[k] for k in f))
        vs.append(v)
        uvs.append(uv)
    return np.concatenate(vs), np.concatenate(uvs)


def uv_sphere(center, radius, n_lon, n_lat):
    """Outward-facing UV sphere: 2*n_lon*(n_lat-1) triangles."""
    center = np.asarray(center, dtype=np.float64)
    th = np.pi * np.arange(n_lat + 1, dtype=np.float64) / n_lat          # 0 = +y pole
    ph = 2.0 * np.pi * np.arange(n_lon + 1, dtype=np.float64) / n_lon
    TH, PH = np.meshgrid(th, ph, indexing="ij")
    P = center + radius * np.stack([np.sin(TH) * np.cos(PH), np.cos(TH), np.sin(TH) * np.sin(PH)], axis=-1)
    UV = np.stack([PH / (2.0 * np.pi), 1.0 - TH / np.pi], axis=-1)
    vs, uvs = [], []
    for i in range(n_lat):
        a, b = P[i, :-1], P[i, 1:]
        c, d = P[i + 1, 1:], P[i + 1, :-1]
        ua, ub, uc, ud = UV[i, :-1], UV[i, 1:], UV[i + 1, 1:], UV[i + 1, :-1]
        if i != 0:            # (a, b, d) degenerates at the top pole (a == b)
            vs.append(np.stack([a, b, d], axis=1)); uvs.append(np.stack([ua, ub, ud], axis=1))
        if i != n_lat - 1:    # (b, c, d) degenerates at the bottom pole (c == d)
            vs.append(np.stack([b, c, d], axis=1)); uvs.append(np.stack([ub, uc, ud], axis=1))
    v = np.concatenate(vs)
    uv = np.concatenate(uvs)
    # orientation: make every geometric normal point away from the centre
    n = np.cross(v[:, 1] - v[:, 0], v[:, 2] - v[:, 0])
    flip = np.einsum("ij,ij->i", n, v.mean(axis=1) - center) < 0
    v[flip] = v[flip][:, [0, 2, 1]]
    uv[flip] = uv[flip][:, [0, 2, 1]]
    return v, uv


def cylinder(center, radius, half_height, n_seg):
    cx, cy, cz = center
    ang = 2.0 * np.pi * np.arange(n_seg + 1) / n_seg
    vs, uvs = [], []
    for i in range(n_seg):
        a0, a1 = ang[i], ang[i + 1]
        p = lambda a, y: (cx + radius * math.cos(a), cy + y, cz + radius * math.sin(a))
        v, uv = quad(p(a1, -half_height), p(a0, -half_height), p(a0, half_height), p(a1, half_height))
        vs.append(v); uvs.append(uv)
        top = np.array([[(cx, cy + half_height, cz), p(a1, half_height), p(a0, half_height)]])
        vs.append(top); uvs.append(np.array([[[0.5, 0.5], [1, 0], [0, 0]]], dtype=np.float64))
    return np.concatenate(vs), np.concatenate(uvs)


# ------------------------------------------------------------------------------------------------
# textures
# ------------------------------------------------------------------------------------------------
def checker_texture(w, h, cells, c0, c1, seed=7, channels=3):
    y, x = np.mgrid[0:h, 0:w]
    m = ((x * cells // w) + (y * cells // h)) % 2
    noise = (hash_uniform(w * h, seed, 11).reshape(h, w) * 24.0).astype(np.int32)
    img = np.where(m[..., None] == 0, np.array(c0, dtype=np.int32), np.array(c1, dtype=np.int32)) + noise[..., None]
    img = np.clip(img, 0, 255).astype(np.uint8)
    if channels == 1:
        return np.ascontiguousarray(img[..., 0])
    if channels == 4:
        return np.ascontiguousarray(np.concatenate([img, np.full((h, w, 1), 255, np.uint8)], axis=-1))
    return np.ascontiguousarray(img)


def wood_texture(w, h, seed=3):
    y, x = np.mgrid[0:h, 0:w].astype(np.float64)
    rings = 0.5 + 0.5 * np.sin(x * 0.11 + 6.0 * np.sin(y * 0.013) + hash_uniform(1, seed, 2)[0] * 6.28)
    base = np.stack([110 + 90 * rings, 60 + 60 * rings, 25 + 30 * rings], axis=-1)
    noise = hash_uniform(w * h, seed, 5).reshape(h, w, 1) * 20.0
    return np.clip(base + noise, 0, 255).astype(np.uint8)


# ------------------------------------------------------------------------------------------------
# render scenes
# ------------------------------------------------------------------------------------------------
def cornell_box(width=1024, height=1024, ball_lon=128, ball_lat=64) -> SceneData:
    """Cornell box in the reference's material vocabulary (Model.cpp:27-32): white floor/ceiling/back
    wall + tall box, red left wall, green right wall, a tessellated diffuse ball, a ceiling light.
    Room = [-1,1] x [0,2] x [-1,1], open towards +z."""
    b = _Builder("cornell-box")
    white = b.material(Material("DiffuseWhite", capi.MAT_LAMBERTIAN, kd=(0.725, 0.71, 0.68)))
    red = b.material(Material("LeftWall", capi.MAT_LAMBERTIAN, kd=(0.63, 0.065, 0.05)))
    green = b.material(Material("RightWall", capi.MAT_LAMBERTIAN, kd=(0.14, 0.45, 0.091)))
    ball = b.material(Material("DiffuseBall", capi.MAT_LAMBERTIAN, kd=(0.55, 0.55, 0.75)))
    light = b.material(Material("Light", capi.MAT_DIFFUSE_LIGHT, emission=(17.0, 12.0, 4.0)))
    b.mesh("floor", white, *quad((-1, 0, 1), (1, 0, 1), (1, 0, -1), (-1, 0, -1)))
    b.mesh("ceiling", white, *quad((-1, 2, -1), (1, 2, -1), (1, 2, 1), (-1, 2, 1)))
    b.mesh("backWall", white, *quad((-1, 0, -1), (1, 0, -1), (1, 2, -1), (-1, 2, -1)))
    b.mesh("leftWall", red, *quad((-1, 0, 1), (-1, 0, -1), (-1, 2, -1), (-1, 2, 1)))
    b.mesh("rightWall", green, *quad((1, 0, -1), (1, 0, 1), (1, 2, 1), (1, 2, -1)))
    b.mesh("tallBox", white, *box((-0.35, 0.6, -0.3), (0.3, 0.6, 0.3), rot_y_deg=18.0))
    b.mesh("ball", ball, *uv_sphere((0.42, 0.35, 0.3), 0.35, ball_lon, ball_lat))
    # light just under the ceiling, normal pointing down (-y)
    b.mesh("light", light, *quad((-0.25, 1.98, -0.25), (0.25, 1.98, -0.25), (0.25, 1.98, 0.25), (-0.25, 1.98, 0.25)))
    cam = dict(width=width, height=height, fovy=30.0, eye=(0.0, 1.0, 4.8), look_at=(0.0, 1.0, 0.0), up=(0.0, 1.0, 0.0))
    return b.build(cam)


def veach_mis(width=1280, height=720, light_lon=32, light_lat=16) -> SceneData:
    """veach-mis stand-in (Model.cpp:17-25): four spherical lights of growing radius over four tilted
    Phong plates of growing shininess, Phong floor/back wall with Ns = 1 (the NEE-able Phong)."""
    b = _Builder("veach-mis")
    m0 = b.material(Material("material0", capi.MAT_PHONG, kd=(0.4, 0.4, 0.4), ks=(0.0, 0.0, 0.0), ns=1.0))
    plates = [b.material(Material(f"material{i+1}", capi.MAT_PHONG, kd=(0.07, 0.09, 0.13), ks=(0.9, 0.9, 0.9), ns=ns))
              for i, ns in enumerate((12.0, 50.0, 200.0, 1000.0))]
    radii = (0.05, 0.12, 0.3, 0.7)
    lights = [b.material(Material(f"light{i+1}", capi.MAT_DIFFUSE_LIGHT,
                                  emission=tuple(c * (0.7 / r) ** 2 for c in col)))
              for i, (r, col) in enumerate(zip(radii, ((1.2, 1.2, 1.2), (1.2, 1.0, 0.6), (0.6, 1.2, 0.7), (0.7, 0.8, 1.3))))]
    b.mesh("floor", m0, *quad((-8, 0, 6), (8, 0, 6), (8, 0, -6), (-8, 0, -6)))
    b.mesh("backWall", m0, *quad((-8, 0, -6), (8, 0, -6), (8, 10, -6), (-8, 10, -6)))
    # plates: long in x, tilted about x so that they mirror the lights towards the camera
    for i, tilt in enumerate((38.0, 28.0, 18.0, 9.0)):
        zc, yc = 1.6 - 1.25 * i, 0.35 + 0.32 * i
        c, s = math.cos(math.radians(tilt)), math.sin(math.radians(tilt))
        hw, hd = 4.0, 0.5
        p = lambda x, d: (x, yc + d * s, zc - d * c)
        b.mesh(f"plate{i+1}", plates[i], *quad(p(-hw, -hd), p(hw, -hd), p(hw, hd), p(-hw, hd)))
    for i, r in enumerate(radii):
        b.mesh(f"light{i+1}", lights[i], *uv_sphere((-3.6 + 2.4 * i, 3.6, -1.5), r, light_lon, light_lat))
    cam = dict(width=width, height=height, fovy=28.0, eye=(0.0, 4.5, 13.0), look_at=(0.0, 1.6, -1.0), up=(0.0, 1.0, 0.0))
    return b.build(cam)


def bathroom_like(width=1280, height=720, detail=320) -> SceneData:
    """bathroom2 stand-in (Model.cpp:34-50): a closed room of displaced, image-textured Lambertian
    walls, mirrors (`Mirror`, `Bin`), an `Empty` pane (`quad1`) and a window light.
    `detail` = grid resolution of the big surfaces; the default 320 gives 1 109 234 triangles (SURVEY.md C4 asks for a
    room of at least 1 M; round 1 measured with detail=256 = 709 730 triangles while its documents said 1.07 M)."""
    b = _Builder("bathroom2")
    t_wall = b.texture(checker_texture(256, 256, 16, (200, 200, 190), (150, 165, 175), seed=5))
    t_floor = b.texture(wood_texture(512, 256, seed=3))
    t_label = b.texture(checker_texture(64, 64, 8, (230, 230, 230), (40, 40, 40), seed=9, channels=1))
    t_towel = b.texture(checker_texture(128, 128, 32, (180, 60, 60), (210, 190, 170), seed=13, channels=4))
    wall = b.material(Material("Wall", capi.MAT_LAMBERTIAN, kd_tex=t_wall))
    wall_r = b.material(Material("WallRight", capi.MAT_LAMBERTIAN, kd=(0.75, 0.72, 0.68)))
    floor = b.material(Material("WoodFloor", capi.MAT_LAMBERTIAN, kd_tex=t_floor))
    ceramic = b.material(Material("Ceramic", capi.MAT_LAMBERTIAN, kd=(0.8, 0.8, 0.82)))
    towel = b.material(Material("Towel", capi.MAT_LAMBERTIAN, kd_tex=t_towel))
    label = b.material(Material("Label", capi.MAT_LAMBERTIAN, kd_tex=t_label))
    mirror = b.material(Material("Mirror", capi.MAT_MIRROR))
    bin_m = b.material(Material("Bin", capi.MAT_MIRROR))
    empty = b.material(Material("quad1", capi.MAT_EMPTY))
    plastic = b.material(Material("DarkPlastic", capi.MAT_LAMBERTIAN, kd=(0.08, 0.08, 0.09)))
    trims = b.material(Material("Trims", capi.MAT_LAMBERTIAN, kd=(0.6, 0.55, 0.5)))
    light = b.material(Material("Light", capi.MAT_DIFFUSE_LIGHT, emission=(24.0, 22.0, 19.0)))

    n = int(detail)

    def bumps(nu, nv, seed, amp, freq):
        s = np.arange(nu + 1)[:, None] / nu
        t = np.arange(nv + 1)[None, :] / nv
        ph = hash_uniform(4, seed, 1) * 6.28
        h = amp * (np.sin(freq * 6.28 * s + ph[0]) * np.sin(freq * 6.28 * t + ph[1])) \
            + 0.25 * amp * np.sin(3.1 * freq * 6.28 * (s + t) + ph[2])
        return h + (hash_uniform((nu + 1) * (nv + 1), seed, 2).reshape(nu + 1, nv + 1) - 0.5) * amp * 0.2

    # room [-2,2] x [0,2.6] x [-3,3]; normals of the displaced grids face inward
    b.mesh("floor", floor, *grid((-2, 0, 3), (4, 0, 0), (0, 0, -6), n, n, bumps(n, n, 21, 0.004, 12), uv_scale=(4, 6)))
    b.mesh("ceiling", wall_r, *grid((-2, 2.6, -3), (4, 0, 0), (0, 0, 6), n // 2, n // 2, bumps(n // 2, n // 2, 22, 0.003, 5)))
    b.mesh("backWall", wall, *grid((-2, 0, -3), (4, 0, 0), (0, 2.6, 0), n, n, bumps(n, n, 23, 0.006, 16), uv_scale=(4, 3)))
    b.mesh("leftWall", wall, *grid((-2, 0, 3), (0, 0, -6), (0, 2.6, 0), n, n, bumps(n, n, 24, 0.006, 24), uv_scale=(6, 3)))
    b.mesh("rightWall", wall_r, *grid((2, 0, -3), (0, 0, 6), (0, 2.6, 0), n, n, bumps(n, n, 25, 0.006, 24), uv_scale=(6, 3)))
    b.mesh("frontWall", wall, *grid((2, 0, 3), (-4, 0, 0), (0, 2.6, 0), n // 2, n // 2, bumps(n // 2, n // 2, 26, 0.006, 8), uv_scale=(4, 3)))
    # furniture
    b.mesh("bathtub", ceramic, *uv_sphere((-1.0, 0.2, -1.8), 0.85, n, n // 2))
    b.mesh("sink", ceramic, *uv_sphere((1.35, 0.95, -2.2), 0.35, n // 2, n // 4))
    b.mesh("cabinet", trims, *box((1.35, 0.4, -2.3), (0.5, 0.4, 0.45)))
    tn = max(8, n // 2)
    s = np.arange(tn + 1)[:, None] / tn
    t = np.arange(tn + 1)[None, :] / tn
    folds = 0.03 * np.sin(18.0 * s * 3.14159) * (0.3 + t) + 0.01 * np.sin(40.0 * t)
    b.mesh("towel", towel, *grid((1.97, 0.9, 0.2), (0, 0, 0.9), (0, 0.9, 0), tn, tn, folds.astype(np.float64), uv_scale=(2, 2)))
    b.mesh("mirror", mirror, *quad((0.7, 1.2, -2.97), (1.9, 1.2, -2.97), (1.9, 2.2, -2.97), (0.7, 2.2, -2.97)))
    b.mesh("bin", bin_m, *cylinder((-1.5, 0.3, 1.6), 0.22, 0.3, max(16, n // 4)))
    b.mesh("label", label, *quad((-1.6, 0.25, 1.83), (-1.4, 0.25, 1.83), (-1.4, 0.4, 1.83), (-1.6, 0.4, 1.83)))
    b.mesh("shelf", plastic, *box((-1.95, 1.5, 0.5), (0.05, 0.02, 0.6)))
    for i in range(6):
        b.mesh(f"bottle{i}", plastic if i % 2 else ceramic, *uv_sphere((-1.9, 1.6, 0.05 + 0.18 * i), 0.06, max(12, n // 8), max(6, n // 16)))
    # an absorbing `Empty` mat on the floor; window: emissive quad on the left wall (normal +x)
    b.mesh("quad1", empty, *quad((0.2, 0.012, 1.5), (1.2, 0.012, 1.5), (1.2, 0.012, 0.5), (0.2, 0.012, 0.5)))
    b.mesh("window", light, *quad((-1.97, 1.2, 0.6), (-1.97, 1.2, -0.6), (-1.97, 2.2, -0.6), (-1.97, 2.2, 0.6)))
    b.mesh("lamp", light, *quad((-0.3, 2.58, -0.3), (0.3, 2.58, -0.3), (0.3, 2.58, 0.3), (-0.3, 2.58, 0.3)))
    cam = dict(width=width, height=height, fovy=55.0, eye=(0.6, 1.45, 2.7), look_at=(-0.2, 1.0, -2.0), up=(0.0, 1.0, 0.0))
    return b.build(cam)


def tiny_scene(kind="mixed") -> SceneData:
    """A few dozen triangles touching every material type: the unit-test scene."""
    b = _Builder("tiny-" + kind)
    tex = b.texture(checker_texture(16, 8, 4, (220, 40, 40), (40, 60, 220), seed=2))
    white = b.material(Material("DiffuseWhite", capi.MAT_LAMBERTIAN, kd=(0.7, 0.7, 0.7)))
    texm = b.material(Material("Wood", capi.MAT_LAMBERTIAN, kd_tex=tex))
    ph_d = b.material(Material("material0", capi.MAT_PHONG, kd=(0.5, 0.4, 0.3), ks=(0.0, 0.0, 0.0), ns=1.0))
    ph_s = b.material(Material("material1", capi.MAT_PHONG, kd=(0.2, 0.2, 0.3), ks=(0.8, 0.8, 0.8), ns=40.0))
    mir = b.material(Material("Mirror", capi.MAT_MIRROR))
    emp = b.material(Material("quad1", capi.MAT_EMPTY))
    light = b.material(Material("Light", capi.MAT_DIFFUSE_LIGHT, emission=(12.0, 11.0, 9.0)))
    light2 = b.material(Material("light1", capi.MAT_DIFFUSE_LIGHT, emission=(2.0, 3.0, 6.0)))
    box_m, wall_m, ball_m = white, ph_s, ph_s
    if kind == "cook":
        # CookTorrance (Source/Material.h:368-521): the Au constants of Source/Model.cpp:309, an
        # anisotropic Cu and a k = 0 dielectric-like Al2O3 (the commented-out alternatives there)
        ball_m = b.material(Material("CookGold", capi.MAT_COOK_TORRANCE, kd=(0.8, 0.6, 0.2)))
        wall_m = b.material(Material("CookCopper", capi.MAT_COOK_TORRANCE, kd=(0.7, 0.4, 0.3), alpha_x=0.1, alpha_y=0.5,
                                     eta=(0.3, 0.7, 1.1), k=(4.4, 3.1, 2.5)))
        box_m = b.material(Material("CookAlumina", capi.MAT_COOK_TORRANCE, kd=(0.9, 0.9, 0.9), alpha_x=0.6, alpha_y=0.05,
                                    eta=(1.75, 1.76, 1.78), k=(0.0, 0.0, 0.0)))
    if kind == "debug":
        # DebugMaterial (Source/Material.h:523-535, Model.cpp:317-319): HasEmission() is true and it
        # emits its albedo, so main.cpp:40 also lists its meshes among the lights
        ball_m = b.material(Material("DebugBall", capi.MAT_DEBUG, kd=(0.9, 0.3, 0.1)))
        box_m = b.material(Material("DebugBox", capi.MAT_DEBUG, kd=(0.1, 0.5, 1.5)))
    b.mesh("floor", texm, *grid((-1, 0, 1), (2, 0, 0), (0, 0, -2), 3, 3))
    b.mesh("ceiling", white, *quad((-1, 2, -1), (1, 2, -1), (1, 2, 1), (-1, 2, 1)))
    b.mesh("backWall", ph_d, *quad((-1, 0, -1), (1, 0, -1), (1, 2, -1), (-1, 2, -1)))
    b.mesh("leftWall", wall_m, *quad((-1, 0, 1), (-1, 0, -1), (-1, 2, -1), (-1, 2, 1)))
    b.mesh("rightWall", mir, *quad((1, 0, -1), (1, 0, 1), (1, 2, 1), (1, 2, -1)))
    b.mesh("box", box_m, *box((-0.3, 0.35, -0.2), (0.3, 0.35, 0.3), rot_y_deg=25.0))
    b.mesh("ball", ball_m, *uv_sphere((0.45, 0.3, 0.35), 0.3, 12, 6))
    b.mesh("pane", emp, *quad((-0.9, 0.2, 0.6), (-0.5, 0.2, 0.6), (-0.5, 0.6, 0.6), (-0.9, 0.6, 0.6)))
    b.mesh("light", light, *quad((-0.3, 1.98, -0.3), (0.3, 1.98, -0.3), (0.3, 1.98, 0.3), (-0.3, 1.98, 0.3)))
    b.mesh("light2", light2, *uv_sphere((-0.6, 1.4, -0.6), 0.12, 8, 4))
    cam = dict(width=48, height=32, fovy=32.0, eye=(0.0, 1.0, 4.6), look_at=(0.0, 1.0, 0.0), up=(0.0, 1.0, 0.0))
    return b.build(cam)


# ------------------------------------------------------------------------------------------------
# ray-query microbench inputs (BASELINE.json configs[4])
# ------------------------------------------------------------------------------------------------
def terrain(nx=1000, ny=500, seed=1) -> SceneData:
    """Noisy height field over [0,1]^2: 2*nx*ny triangles (1000x500 -> 1 M, 2500x2000 -> 10 M)."""
    b = _Builder(f"terrain-{nx}x{ny}")
    m = b.material(Material("DiffuseWhite", capi.MAT_LAMBERTIAN, kd=(0.6, 0.6, 0.6)))
    s = np.arange(nx + 1)[:, None] / nx
    t = np.arange(ny + 1)[None, :] / ny
    h = 0.08 * np.sin(9.0 * s) * np.cos(7.0 * t) + 0.03 * np.sin(31.0 * s + 17.0 * t)
    h = h + (hash_uniform((nx + 1) * (ny + 1), seed, 3).reshape(nx + 1, ny + 1) - 0.5) * (0.6 / max(nx, ny))
    b.mesh("terrain", m, *grid((0, 0, 0), (1, 0, 0), (0, 0, 1), nx, ny, -h))
    cam = dict(width=4096, height=4096, fovy=40.0, eye=(0.5, 0.9, 1.9), look_at=(0.5, 0.0, 0.5), up=(0.0, 1.0, 0.0))
    return b.build(cam)


def soup(n=1_000_000, seed=1, size=None) -> SceneData:
    """n random small triangles in the unit cube."""
    b = _Builder(f"soup-{n}")
    m = b.material(Material("DiffuseWhite", capi.MAT_LAMBERTIAN, kd=(0.6, 0.6, 0.6)))
    if size is None:
        size = 1.5 / max(1.0, n ** (1.0 / 3.0))
    c = hash_uniform(3 * n, seed, 1).reshape(n, 1, 3)
    e = (hash_uniform(9 * n, seed, 2).reshape(n, 3, 3) - 0.5) * size
    b.mesh("soup", m, c + e)
    cam = dict(width=4096, height=4096, fovy=40.0, eye=(0.5, 0.5, 2.6), look_at=(0.5, 0.5, 0.5), up=(0.0, 1.0, 0.0))
    return b.build(cam)


def few_triangles(n: int, seed=5) -> SceneData:
    """n random triangles in the unit cube, one Lambertian mesh (degenerate sizes for the BVH builders)"""
    b = _Builder(f"few-{n}")
    mat = b.material(Material("DiffuseWhite", capi.MAT_LAMBERTIAN, kd=(0.7, 0.7, 0.7)))
    c = hash_uniform(3 * n, seed, 0).reshape(n, 1, 3)
    e = (hash_uniform(9 * n, seed, 1).reshape(n, 3, 3) - 0.5) * 0.8
    b.mesh("few", mat, c + e)
    return b.build(dict(width=64, height=64, fovy=40.0, eye=(0.5, 0.5, 3.0), look_at=(0.5, 0.5, 0.5), up=(0.0, 1.0, 0.0)))


def repeated_triangle(n: int) -> SceneData:
    """the same triangle n times: every box centre (hence every Morton code) equal, every hit an equal-t tie"""
    b = _Builder(f"repeat-{n}")
    mat = b.material(Material("DiffuseWhite", capi.MAT_LAMBERTIAN, kd=(0.7, 0.7, 0.7)))
    tri = np.array([[[0.1, 0.1, 0.5], [0.9, 0.2, 0.4], [0.4, 0.9, 0.6]]])
    b.mesh("repeat", mat, np.repeat(tri, n, axis=0))
    return b.build(dict(width=64, height=64, fovy=40.0, eye=(0.5, 0.5, 3.0), look_at=(0.5, 0.5, 0.5), up=(0.0, 1.0, 0.0)))


def random_rays(scene: SceneData, n: int, seed=2):
    """Origins uniform in the 1.5x-inflated bounding box, directions uniform on the sphere."""
    lo = scene.vertices.reshape(-1, 3).min(axis=0)
    hi = scene.vertices.reshape(-1, 3).max(axis=0)
    c, h = 0.5 * (lo + hi), 0.75 * (hi - lo) + 1e-3
    o = c + (hash_uniform(3 * n, seed, 1).reshape(n, 3) * 2.0 - 1.0) * h
    z = hash_uniform(n, seed, 2) * 2.0 - 1.0
    ph = hash_uniform(n, seed, 3) * (2.0 * np.pi)
    r = np.sqrt(np.maximum(0.0, 1.0 - z * z))
    d = np.stack([r * np.cos(ph), r * np.sin(ph), z], axis=1)
    return np.ascontiguousarray(o), np.ascontiguousarray(d)


def camera_rays(camera: dict, width=None, height=None):
    """Pixel-centre rays of a pinhole camera, unnormalised directions (Source/Camera.cpp:75-117)."""
    W = int(width or camera["width"]); H = int(height or camera["height"])
    eye = np.array(camera["eye"], dtype=np.float64); look = np.array(camera["look_at"], dtype=np.float64)
    up = np.array(camera["up"], dtype=np.float64)
    focal = np.linalg.norm(eye - look)
    hh = math.tan(math.radians(camera["fovy"]) / 2.0)
    vh = 2.0 * hh * focal
    vw = vh * (W / H)
    w = (eye - look) / focal
    u = np.cross(up, w); u /= np.linalg.norm(u)
    v = np.cross(w, u)
    du, dv = vw * u / W, -vh * v / H
    p00 = eye - focal * w - vw * u / 2.0 + vh * v / 2.0 + 0.5 * (du + dv)
    i, j = np.meshgrid(np.arange(W, dtype=np.float64), np.arange(H, dtype=np.float64), indexing="xy")
    d = (p00 + i[..., None] * du + j[..., None] * dv - eye).reshape(-1, 3)
    o = np.broadcast_to(eye, d.shape).copy()
    return o, np.ascontiguousarray(d)


# ------------------------------------------------------------------------------------------------
# scene files (example-scenes-cg24 layout: <dir>/<name>.obj + .mtl + .xml + textures)
# ------------------------------------------------------------------------------------------------
def png_bytes(img: np.ndarray, filters=(0, 1, 2, 3, 4), level=6) -> bytes:
    """PNG encoder for uint8 [H,W] / [H,W,C] (C = 1..4): real deflate (dynamic Huffman blocks) and a
    per-row rotation through the five filter types, so that reading it back exercises all of a
    decoder."""
    import struct
    import zlib
    a = np.ascontiguousarray(img, dtype=np.uint8)
    if a.ndim == 2:
        a = a[:, :, None]
    h, w, c = a.shape
    ctype = {1: 0, 2: 4, 3: 2, 4: 6}[c]
    rows = a.reshape(h, w * c).astype(np.int32)
    raw = bytearray()
    prev = np.zeros(w * c, dtype=np.int32)
    for y in range(h):
        cur = rows[y]
        ft = filters[y % len(filters)]
        left = np.concatenate([np.zeros(c, np.int32), cur[:-c]])
        upleft = np.concatenate([np.zeros(c, np.int32), prev[:-c]])
        if ft == 0:
            out = cur
        elif ft == 1:
            out = cur - left
        elif ft == 2:
            out = cur - prev
        elif ft == 3:
            out = cur - ((left + prev) >> 1)
        else:
            p = left + prev - upleft
            pa, pb, pc = np.abs(p - left), np.abs(p - prev), np.abs(p - upleft)
            pred = np.where((pa <= pb) & (pa <= pc), left, np.where(pb <= pc, prev, upleft))
            out = cur - pred
        raw.append(ft)
        raw += (out & 255).astype(np.uint8).tobytes()
        prev = cur

    def chunk(t, d):
        return struct.pack(">I", len(d)) + t + d + struct.pack(">I", zlib.crc32(t + d) & 0xFFFFFFFF)

    return (b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, ctype, 0, 0, 0))
            + chunk(b"IDAT", zlib.compress(bytes(raw), level)) + chunk(b"IEND", b""))


def write_scene_files(sc: SceneData, directory: str, name: str, anonymous_groups=False) -> None:
    """Writes `sc` the way the cg24 example scenes are laid out, numbers with repr() (17 significant
    digits: strtod reads back the very same doubles)."""
    import os
    os.makedirs(directory, exist_ok=True)
    tex_names = []
    for i, t in enumerate(sc.textures):
        fn = f"{name}_tex{i}.png"
        tex_names.append(fn)
        if t is not None:
            with open(os.path.join(directory, fn), "wb") as f:
                f.write(png_bytes(t))
    with open(os.path.join(directory, name + ".mtl"), "w") as f:
        for m in sc.materials:
            f.write(f"newmtl {m.name}\n")
            f.write("Kd " + " ".join(repr(float(x)) for x in m.kd) + "\n")
            f.write("Ks " + " ".join(repr(float(x)) for x in m.ks) + "\n")
            f.write(f"Ns {float(m.ns)!r}\n")
            if m.kd_tex >= 0:
                f.write(f"map_Kd {tex_names[m.kd_tex]}\n")
            f.write("\n")
    v = np.asarray(sc.vertices, dtype=np.float64).reshape(-1, 3)
    uv = None if sc.texcoords is None else np.asarray(sc.texcoords, dtype=np.float64).reshape(-1, 2)
    with open(os.path.join(directory, name + ".obj"), "w") as f:
        f.write(f"# {sc.name}\nmtllib {name}.mtl\n")
        for m in range(len(sc.mesh_material)):
            b, e = int(sc.mesh_first[m]), int(sc.mesh_first[m + 1])
            f.write("g\n" if anonymous_groups else f"g {sc.mesh_names[m]}\n")
            f.write(f"usemtl {sc.materials[int(sc.mesh_material[m])].name}\n")
            for k in range(3 * b, 3 * e):
                f.write(f"v {float(v[k, 0])!r} {float(v[k, 1])!r} {float(v[k, 2])!r}\n")
            if uv is not None:
                for k in range(3 * b, 3 * e):
                    f.write(f"vt {float(uv[k, 0])!r} {float(uv[k, 1])!r}\n")
            for t in range(b, e):
                i = 3 * t + 1
                # alternate absolute and relative (negative) indices
                if uv is not None:
                    if t % 2 == 0:
                        f.write(f"f {i}/{i} {i + 1}/{i + 1} {i + 2}/{i + 2}\n")
                    else:
                        n = 3 * e
                        f.write(f"f {i - 1 - n}/{i - 1 - n} {i - n}/{i - n} {i + 1 - n}/{i + 1 - n}\n")
                else:
                    f.write(f"f {i} {i + 1} {i + 2}\n")
    cam = sc.camera
    with open(os.path.join(directory, name + ".xml"), "w") as f:
        f.write('<?xml version="1.0" encoding="utf-8"?>\n')
        f.write(f'<camera type="perspective" width="{int(cam["width"])}" height="{int(cam["height"])}" fovy="{float(cam["fovy"])!r}">\n')
        for tag, key in (("eye", "eye"), ("lookat", "look_at"), ("up", "up")):
            x, y, z = (float(c) for c in cam[key])
            f.write(f'    <{tag} x="{x!r}" y="{y!r}" z="{z!r}"/>\n')
        f.write("</camera>\n<!-- light radiance by material name -->\n")
        for m in sc.materials:
            if m.type == capi.MAT_DIFFUSE_LIGHT:
                f.write(f'<light mtlname="{m.name}" radiance="{float(m.emission[0])!r},{float(m.emission[1])!r},{float(m.emission[2])!r}"/>\n')
