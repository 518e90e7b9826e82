"""Generates tests/golden/*.npz by running the inputs of tests/cases.py through the UNMODIFIED
reference sources (oracle/_ref/libprt_ref.so, built from /root/reference by oracle/Makefile).

Run in the build container only (needs /root/reference):   python tests/golden/make_golden.py
The fixtures pin the oracle port (tests/test_oracle_pinning.py) and, on the GPU box where the
reference cannot be built, stand in for it.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import cases  # noqa: E402
import oracle_lib as ol  # noqa: E402
from pooraytracer_b200.api import make_camera, make_params  # noqa: E402


def material_rows(R, sc):
    """Scatter and Eval of every material with scripted uniforms -> (scatter rows, eval rows)"""
    rows, erows = [], []
    for mi in range(len(sc.materials)):
        for (rd, nrm, tng, uv, u) in cases.scatter_inputs():
            att, od, nd, ok = R.material_scatter(mi, rd, nrm, tng, uv, u)
            rows.append(np.concatenate([[mi], att, od, [nd, float(ok)]]))
        for (rd, nrm, tng, uv, wi, u) in cases.eval_inputs():
            erows.append(np.concatenate([[mi], R.material_eval(mi, rd, nrm, tng, uv, wi, u)]))
    return np.array(rows), np.array(erows)


def render_fixture(name):
    key, w, h, spp, depth, nee, rkey, bg = cases.RENDER_CASES[name]
    sc = cases.scene_by_key(key)
    R = ol.OracleScene("ref", sc, ol.DEFINE_UB)
    cam = make_camera(sc.camera, w, h)
    prm = make_params(spp, depth, sample_lights=nee, rng_key=rkey, background=bg)
    img, info = R.render(cam, prm, ol.RNG_COUNTER, nthreads=8)
    px = [(w // 2, h // 2), (3, 5), (w - 2, h - 3), (w // 3, 2 * h // 3)]
    samples = np.stack([R.pixel_samples(cam, prm, x, y) for x, y in px])
    np.savez_compressed(os.path.join(HERE, f"render_{name}.npz"), image=img, pixels=np.array(px), samples=samples,
                        world_hit_calls=info.world_hit_calls, shadow_calls=info.shadow_calls)
    print("render", name, img.shape, float(img.mean()), info.world_hit_calls)
    R.close()


def cook_fixtures():
    """CookTorrance: Scatter / Eval with scripted uniforms + two renders"""
    sc = cases.cook()
    R = ol.OracleScene("ref", sc, ol.DEFINE_UB)
    rows, erows = material_rows(R, sc)
    np.savez_compressed(os.path.join(HERE, "material_cook.npz"), scatter=rows, eval=erows)
    print("material_cook", rows.shape, erows.shape, "scattered", int(rows[:, -1].sum()), "eval nonzero", int((erows[:, 1:] != 0).any(1).sum()))
    R.close()
    for name in ("cook_nee", "cook_plain"):
        render_fixture(name)


def ref_output_stage(img):
    """(rgb8 [H,W,3], raw .hdr file bytes) that the reference's Camera::WriteColorAttachment writes for img"""
    import ctypes as C
    import tempfile
    import imgutil
    lib = ol.OracleLib("ref").lib
    img = np.ascontiguousarray(img, dtype=np.float64)
    h, w, _ = img.shape
    with tempfile.TemporaryDirectory(prefix="prt_out_") as d:
        path = os.path.join(d, "frame.png")
        rc = lib.ref_write_color_attachment(img.ctypes.data_as(C.c_void_p), C.c_int32(w), C.c_int32(h), path.encode(), C.c_int(1))
        assert rc == 0
        return imgutil.png_rgb8(open(path, "rb").read()), open(os.path.join(d, "frame.hdr"), "rb").read()


def output_stage_fixture():
    """f2: the bytes Camera::WriteColorAttachment (Camera.cpp:279-331) produces. The 254 places where
    the 8-bit quantiser steps are located by bisection THROUGH the reference (each probe is a PNG it wrote), then
    the fixture holds every double within 6 ulps of each step, the special values, and a noisy HDR-range frame."""
    k = np.arange(1, 255)
    lo = np.zeros(254, np.uint64)
    hi = np.full(254, np.float64(1.0).view(np.uint64))
    while (hi - lo > 1).any():
        mid = lo + (hi - lo) // 2
        img = np.zeros((1, 256, 3)); img[0, :254, 0] = mid.view(np.float64)
        q = ref_output_stage(img)[0][0, :254, 0].astype(int)
        up = q >= k
        hi = np.where(up & (hi - lo > 1), mid, hi)
        lo = np.where(~up & (hi - lo > 1), mid, lo)
    steps = hi
    near = (steps[:, None].astype(np.int64) + np.arange(-6, 7)[None, :]).astype(np.uint64).view(np.float64).ravel()
    special = np.array([np.nan, -np.nan, np.inf, -np.inf, -1.0, -0.0, 0.0, 5e-324, 1e-300, 0.0031308, np.nextafter(0.0031308, 1), np.nextafter(0.0031308, 0),
                        0.5, 0.9999, 1.0, np.nextafter(1.0, 2), 5.0, 17.0, 1e300, -1e-300, 2.0 ** -130, 3.5e38, 1e39])
    noise = np.exp(cases.hash_uniform(3 * 64 * 48, 21, 1) * 14.0 - 10.0)  # 4.5e-5 .. 55
    vals = np.concatenate([near, special, noise])
    w = 96
    h = -(-len(vals) // (3 * w))
    img = np.zeros(h * w * 3)
    img[:len(vals)] = vals
    img = img.reshape(h, w, 3)
    with np.errstate(all="ignore"):
        rgb8, hdr = ref_output_stage(img)
    np.savez_compressed(os.path.join(HERE, "output_stage.npz"), image=img, rgb8=rgb8, hdr_file=np.frombuffer(hdr, np.uint8), steps=steps.view(np.float64))
    print("output_stage", img.shape, "steps", steps.view(np.float64)[[0, 1, 127, 253]], "bytes used", len(np.unique(rgb8)))


def jpeg_fixture():
    """f1: 60 of the JPEG files of tests/test_scene_io.py:jpeg_cases() (written by PIL here) with the texels the stb_image.h
    vendored by the reference decodes them to (stbi_load inside oracle/_ref/libprt_ref.so, compiled from the reference's
    own Texture.cpp). The fixture pins pooraytracer_b200/csrc/jpeg_decode.cpp on boxes without oracle/_ref."""
    import io
    import tempfile
    from PIL import Image
    import test_scene_io as tsi
    out = {}
    all_cases = tsi.jpeg_cases()
    pick = all_cases[::max(1, len(all_cases) // 60)][:60]
    with tempfile.TemporaryDirectory(prefix="prt_jpg_") as d:
        for name, kw, img in pick:
            p = os.path.join(d, "t.jpg")
            try:
                Image.fromarray(img).save(p, "JPEG", **kw)
            except Exception:
                continue
            out["file_" + name] = np.frombuffer(open(p, "rb").read(), np.uint8)
            out["stb_" + name] = tsi._stb_load(p)
    np.savez_compressed(os.path.join(HERE, "jpeg_fixture.npz"), **out)
    print("jpeg_fixture", len(out) // 2, "files", sum(v.nbytes for k, v in out.items() if k.startswith("file_")), "bytes of JPEG")


def main():
    assert ol.available("ref"), "build oracle/_ref first (make -C oracle ref)"
    if len(sys.argv) > 1 and sys.argv[1] == "jpeg":
        return jpeg_fixture()
    if len(sys.argv) > 1 and sys.argv[1] == "output":
        return output_stage_fixture()
    if len(sys.argv) > 1 and sys.argv[1] == "cook":
        return cook_fixtures()
    if len(sys.argv) > 1 and sys.argv[1] == "debug":  # DebugMaterial renders only
        for name in ("debug_nee", "debug_plain"):
            render_fixture(name)
        return
    tag = ol.PER_TRIANGLE_MATERIAL | ol.UV_TAG
    only = sys.argv[2:] if len(sys.argv) > 2 and sys.argv[1] == "rays" else None  # e.g. `make_golden.py rays far`
    for key in list(cases.SCENES) + ["ties", "far"]:
        if only is not None and key not in only:
            continue
        sc, o, d = cases.rays_for(key)
        R = ol.OracleScene("ref", sc, tag)
        prim, t, a, b = R.trace_closest(o, d, 1e-4, np.inf)
        rank, lo = R.reference_order()
        out = dict(prim=prim, t=t, alpha=a, beta=b, world_rank=rank, light_order=lo)
        # bounded interval (the shadow-ray style query)
        tmax = np.where(prim >= 0, t, 1.0) * np.where(np.arange(len(t)) % 3 == 0, 1.0, 0.5 + (np.arange(len(t)) % 7) * 0.25)
        p2, t2, _, _ = R.trace_closest(o, d, 1e-3, tmax)
        out.update(tmax=tmax, prim_bounded=p2, t_bounded=t2)
        recs = np.stack([R.triangle_record(p) for p in range(0, sc.n_triangles, max(1, sc.n_triangles // 64))])
        out["tri_records"] = recs
        if len(lo):
            lo_o, lo_u = cases.light_inputs(sc)
            pos, nrm, front, pdf, lprim = R.sample_lights(lo_o, lo_u)
            out.update(l_pos=pos, l_nrm=nrm, l_front=front, l_pdf=pdf, l_prim=lprim)
        np.savez_compressed(os.path.join(HERE, f"rays_{key}.npz"), **out)
        print("rays", key, len(prim), int((prim >= 0).sum()), "hits")
        R.close()
    if only is not None:
        return

    # materials + textures on the tiny scene (every material type, one texture of each channel count
    # lives in the bathroom scene)
    sc = cases.tiny()
    for flags, name in ((0, "raw"), (ol.DEFINE_UB, "defined")):
        R = ol.OracleScene("ref", sc, flags)
        rows = []
        for mi in range(len(sc.materials)):
            for (rd, nrm, tng, uv, u) in cases.scatter_inputs():
                att, od, nd, ok = R.material_scatter(mi, rd, nrm, tng, uv, u)
                rows.append(np.concatenate([[mi], att, od, [nd, float(ok)]]))
        np.savez_compressed(os.path.join(HERE, f"scatter_tiny_{name}.npz"), rows=np.array(rows))
        R.close()
    scb = cases.bathroom_small()
    R = ol.OracleScene("ref", scb, 0)
    uv = np.concatenate([cases.hash_uniform(800, 12, 1).reshape(-1, 2) * 1.4 - 0.2,
                         np.array([[0, 0], [1, 1], [0, 1], [1, 0], [0.5, 0.5], [1.0, 0.999999], [-3, 7]], dtype=np.float64)])
    tex = {f"tex{i}": R.texture_value(i, uv) for i in range(len(scb.textures))}
    np.savez_compressed(os.path.join(HERE, "texture_bathroom.npz"), uv=uv, **tex)
    R.close()

    for name in cases.RENDER_CASES:
        if not name.startswith("cook"):
            render_fixture(name)
    cook_fixtures()
    output_stage_fixture()
    jpeg_fixture()


if __name__ == "__main__":
    main()
