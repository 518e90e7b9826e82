"""TEST INFRASTRUCTURE: the deterministic inputs shared by tests/golden/make_golden.py (which ran
them through the compiled reference) and the tests (which run them through the oracle port and the
CUDA path)."""
from __future__ import annotations

import numpy as np

from pooraytracer_b200 import scenes
from pooraytracer_b200.scenes import hash_uniform


def tiny():
    return scenes.tiny_scene()


def cornell_small():
    return scenes.cornell_box(64, 64, ball_lon=24, ball_lat=12)


def veach_small():
    return scenes.veach_mis(64, 36, light_lon=8, light_lat=4)


def bathroom_small():
    return scenes.bathroom_like(64, 36, detail=16)


def cook():
    return scenes.tiny_scene("cook")


def debug():
    return scenes.tiny_scene("debug")


SCENES = {"tiny": tiny, "cornell": cornell_small, "veach": veach_small, "bathroom": bathroom_small}


FAR_SCALE, FAR_OFFSET = 1.0e4, 1.0e6


def far():
    """cornell_small scaled by 1e4 and moved to (1e6, 1e6, 1e6): |o| / extent ~ 50, coordinates carry ~2e-10 of
    absolute rounding. The adversarial rays of ray_batch (vertices, edge midpoints, in-plane, axis-parallel) are
    rebuilt from the TRANSFORMED vertices, so they still aim at the stored vertex bits. Pins on_boundary
    (pooraytracer_b200/csrc/traverse.cuh) away from unit scale."""
    sc = cornell_small()
    sc.name = "far"
    sc.vertices = sc.vertices * FAR_SCALE + FAR_OFFSET
    cam = dict(sc.camera)
    cam["eye"] = tuple(np.array(cam["eye"]) * FAR_SCALE + FAR_OFFSET)
    cam["look_at"] = tuple(np.array(cam["look_at"]) * FAR_SCALE + FAR_OFFSET)
    sc.camera = cam
    return sc


def ray_batch(sc, n_random=3000, cam_w=40, cam_h=30, seed=2, scale=1.0, offset=0.0):
    """random rays + pixel-centre camera rays + adversarial rays (axis-parallel directions with exact
    zeros, rays aimed at vertices and at edge midpoints, rays in triangle planes). scale / offset move the
    fixed-coordinate axis-parallel rays with a transformed scene (far())."""
    o1, d1 = scenes.random_rays(sc, n_random, seed=seed)
    o2, d2 = scenes.camera_rays(sc.camera, cam_w, cam_h)
    v = sc.vertices.reshape(-1, 3, 3)
    k = min(200, v.shape[0])
    pick = (hash_uniform(k, seed, 9) * v.shape[0]).astype(np.int64)
    eye = np.array(sc.camera["eye"], dtype=np.float64)
    tgt = np.concatenate([v[pick, 0], 0.5 * (v[pick, 0] + v[pick, 1]), (v[pick, 0] + v[pick, 1] + v[pick, 2]) / 3.0])
    o3 = np.broadcast_to(eye, tgt.shape).copy()
    d3 = tgt - eye
    # axis-parallel rays from points on a grid (exact zeros in two direction components)
    g = np.linspace(-0.9, 0.9, 7)
    gx, gy = np.meshgrid(g, g + 1.0)
    o4 = np.stack([gx.ravel(), gy.ravel(), np.full(gx.size, 5.0)], axis=1)
    d4 = np.tile(np.array([[0.0, 0.0, -1.0]]), (gx.size, 1))
    o5 = np.stack([np.full(gx.size, -5.0), gy.ravel(), gx.ravel()], axis=1)
    d5 = np.tile(np.array([[2.0, 0.0, 0.0]]), (gx.size, 1))
    o6 = np.stack([gx.ravel(), np.full(gx.size, 7.0), gy.ravel() - 1.0], axis=1)
    d6 = np.tile(np.array([[0.0, -0.5, 0.0]]), (gx.size, 1))
    # rays lying in the plane of the first triangles (denominator ~ 0)
    e = v[pick[:50], 1] - v[pick[:50], 0]
    o7 = v[pick[:50], 0] - 2.0 * e
    d7 = e
    o4, o5, o6 = (x * scale + offset for x in (o4, o5, o6))
    d4, d5, d6 = (x * scale for x in (d4, d5, d6))
    o = np.concatenate([o1, o2, o3, o4, o5, o6, o7])
    d = np.concatenate([d1, d2, d3, d4, d5, d6, d7])
    return np.ascontiguousarray(o), np.ascontiguousarray(d)


def rays_for(key):
    """(scene, origins, directions) of the ray fixture tests/golden/rays_<key>.npz"""
    if key == "ties":
        return tie_scene(), *tie_rays()
    if key == "far":
        sc = far()
        return sc, *ray_batch(sc, scale=FAR_SCALE, offset=FAR_OFFSET)
    sc = SCENES[key]()
    return sc, *ray_batch(sc)


def tie_scene():
    """Two coplanar quads built from triangles sharing edges, dyadic coordinates: rays through shared
    edges / vertices meet two or more triangles at exactly the same t."""
    b = scenes._Builder("ties")
    from pooraytracer_b200 import _capi as capi
    m0 = b.material(scenes.Material("DiffuseWhite", capi.MAT_LAMBERTIAN, kd=(0.5, 0.5, 0.5)))
    m1 = b.material(scenes.Material("LeftWall", capi.MAT_LAMBERTIAN, kd=(0.7, 0.2, 0.2)))
    v, uv = scenes.grid((-1, -1, 0), (2, 0, 0), (0, 2, 0), 4, 4)
    b.mesh("gridA", m0, v, uv)
    v2, uv2 = scenes.grid((-1, -1, -1), (2, 0, 0), (0, 2, 0), 2, 2)
    b.mesh("gridB", m1, v2, uv2)
    b.mesh("fan", m0, np.array([[[0, 0, 0.5], [0.5, 0, 0.5], [0, 0.5, 0.5]], [[0, 0, 0.5], [0, 0.5, 0.5], [-0.5, 0, 0.5]],
                                [[0, 0, 0.5], [-0.5, 0, 0.5], [0, -0.5, 0.5]], [[0, 0, 0.5], [0, -0.5, 0.5], [0.5, 0, 0.5]]], dtype=np.float64))
    cam = dict(width=33, height=33, fovy=45.0, eye=(0.0, 0.0, 4.0), look_at=(0.0, 0.0, 0.0), up=(0.0, 1.0, 0.0))
    return b.build(cam)


def tie_rays():
    g = np.arange(-8, 9) / 8.0  # dyadic grid hitting grid lines, diagonals and vertices exactly
    gx, gy = np.meshgrid(g, g)
    o = np.stack([gx.ravel(), gy.ravel(), np.full(gx.size, 4.0)], axis=1)
    d = np.tile(np.array([[0.0, 0.0, -1.0]]), (gx.size, 1))
    o2 = np.tile(np.array([[0.0, 0.0, 4.0]]), (gx.size, 1))
    d2 = np.stack([gx.ravel(), gy.ravel(), np.full(gx.size, -4.0)], axis=1)
    return np.ascontiguousarray(np.concatenate([o, o2])), np.ascontiguousarray(np.concatenate([d, d2]))


def light_inputs(sc, n=2000, seed=4):
    o, _ = scenes.random_rays(sc, n, seed=seed)
    u = hash_uniform(4 * n, seed, 7).reshape(n, 4)
    u = np.floor(u * 2147483648.0) / 2147483648.0  # the generator's 31-bit grid
    u[:8, 1] = [0.0, 1.0 - 2.0 ** -31, 0.25, 0.5, 0.75, 2.0 ** -31, 0.999, 0.001]
    u[:4, 2] = [0.0, 1.0 - 2.0 ** -31, 0.0, 0.5]
    u[:4, 3] = [0.0, 0.0, 1.0 - 2.0 ** -31, 0.5]
    return o, u


def scatter_inputs(n=64, seed=6):
    """(ray_dir, normal, tangent, uv, u[12]) tuples for Material::Scatter with scripted uniforms."""
    r = hash_uniform(n * 24, seed, 3).reshape(n, 24)
    out = []
    for i in range(n):
        nrm = r[i, 0:3] * 2 - 1
        nrm /= np.linalg.norm(nrm)
        t = np.cross(nrm, [0.3, 0.5, 0.8]); t /= np.linalg.norm(t)
        rd = r[i, 3:6] * 2 - 1
        u = np.floor(r[i, 8:20] * 2147483648.0) / 2147483648.0
        out.append((rd * (0.5 + 2 * r[i, 6]), nrm, t, r[i, 20:22] * 1.4 - 0.2, u))
    return out


def eval_inputs(n=64, seed=8):
    """(ray_dir, normal, tangent, uv, wi_world, u[2]) tuples for Material::Eval as Camera::RayColor calls it
    (unnormalised ray directions, unit light directions on either side of the surface)."""
    r = hash_uniform(n * 8, seed, 5).reshape(n, 8)
    out = []
    for i, (rd, nrm, tng, uv, u) in enumerate(scatter_inputs(n, seed)):
        wi = r[i, 0:3] * 2 - 1
        wi /= np.linalg.norm(wi)
        if i % 9 == 0:
            wi = np.cross(nrm, tng)  # grazing: exactly in the tangent plane
        out.append((rd, nrm, tng, uv, wi, u[:2]))
    return out


# scenes that only the render / material fixtures use (no ray fixtures)
EXTRA_SCENES = {"cook": cook, "debug": debug}

RENDER_CASES = {
    # name: (scene key, width, height, spp, depth, sample_lights, rng_key, background)
    "tiny_nee": ("tiny", 48, 32, 4, 8, True, 5, (0.0, 0.0, 0.0)),
    "tiny_plain": ("tiny", 24, 16, 3, 5, False, 9, (0.1, 0.2, 0.3)),
    "cornell_nee": ("cornell", 40, 40, 2, 10, True, 1, (0.0, 0.0, 0.0)),
    "veach_nee": ("veach", 48, 27, 3, 12, True, 3, (0.0, 0.0, 0.0)),
    "bathroom_nee": ("bathroom", 48, 27, 2, 12, True, 7, (0.0, 0.0, 0.0)),
    "cook_nee": ("cook", 48, 32, 4, 8, True, 13, (0.0, 0.0, 0.0)),
    "cook_plain": ("cook", 24, 16, 3, 6, False, 17, (0.2, 0.3, 0.1)),
    "debug_nee": ("debug", 48, 32, 4, 8, True, 19, (0.0, 0.0, 0.0)),
    "debug_plain": ("debug", 24, 16, 3, 6, False, 23, (0.05, 0.05, 0.1)),
}


def scene_by_key(key):
    return (SCENES.get(key) or EXTRA_SCENES[key])()
