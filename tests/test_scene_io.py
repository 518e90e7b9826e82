"""Scene files in, image files out: prt_model_load / prt_image_load / prt_write_*.

CPU tests: the loader returns the arrays the procedural generator wrote (strtod mode: bit for bit;
tinyobjloader's own number parser: within a few ulp and exactly where its arithmetic is exact), the
reference's Model.cpp rules (name table, face-0 material, degenerate-UV fix, anonymous groups,
<light radiance>), the image decoders and writers. GPU tests: a scene loaded from files renders the
very image of the same scene handed over as arrays, and the prt_main drop-in writes it.
"""
import os
import struct
import subprocess
import zlib

import numpy as np
import pytest

import cases
import oracle_lib as ol
from pooraytracer_b200 import api
from pooraytracer_b200 import Context, Model, PrtError, Scene, _capi as capi, load_image, make_camera, make_params, scenes, write_hdr, write_png

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def built():
    import __graft_entry__ as g
    if not os.path.exists(capi.LIB_PATH):
        g.build()
    return True


def _write(tmp_path, sc, **kw):
    d = str(tmp_path / "scene")
    scenes.write_scene_files(sc, d, "s", **kw)
    return d


@pytest.mark.parametrize("key", ["tiny", "cornell", "veach", "bathroom"])
def test_loader_round_trip_is_bit_exact_with_strtod(built, tmp_path, key):
    sc = cases.SCENES[key]()
    m = Model(_write(tmp_path, sc), "s", capi.LOAD_STRTOD)
    assert m.n_triangles == sc.n_triangles
    assert np.array_equal(m.vertices, sc.vertices)
    assert np.array_equal(m.texcoords, sc.texcoords)
    assert np.array_equal(m.mesh_first, sc.mesh_first) and np.array_equal(m.mesh_material, sc.mesh_material)
    assert m.mesh_names == list(sc.mesh_names) and m.material_names == [x.name for x in sc.materials]
    assert not np.any(m.normals)  # no vn in the file: zeros, the Triangle.cpp:21-29 fallback input
    for got, want in zip(m.materials, sc.materials):
        assert got.type == want.type  # Model.cpp:16-51 name table
        if want.type == capi.MAT_DIFFUSE_LIGHT:
            assert tuple(got.emission) == tuple(float(x) for x in want.emission)  # <light radiance="..">
        if want.type in (capi.MAT_LAMBERTIAN, capi.MAT_PHONG):
            assert (got.kd_tex >= 0) == (want.kd_tex >= 0)
            if want.kd_tex >= 0:
                t = sc.textures[want.kd_tex]
                assert np.array_equal(m.texture(got.kd_tex).reshape(t.shape), t)
            else:
                assert tuple(got.kd) == tuple(float(x) for x in want.kd)
        if want.type == capi.MAT_PHONG:
            assert tuple(got.ks) == tuple(float(x) for x in want.ks) and got.ns == float(want.ns)
    c, w = m.camera(), sc.camera
    assert (c.image_width, c.image_height, c.fovy) == (w["width"], w["height"], float(w["fovy"]))
    assert tuple(c.eye) == tuple(map(float, w["eye"])) and tuple(c.look_at) == tuple(map(float, w["look_at"])) and tuple(c.up) == tuple(map(float, w["up"]))
    assert m.warnings == []


def test_tinyobj_number_parser(built, tmp_path):
    """tinyobjloader's tryParseDouble accumulates decimal digits in floating point: exact for short
    dyadic decimals, a few ulp off elsewhere (0.15 -> 0.1 + 5 * 0.01)."""
    d = tmp_path / "n"
    d.mkdir()
    vals = ["0.5", "-2.25", "1e2", "3", ".75", "-.125", "0.15", "12.000001", "1.5e-3", "+7.0E+1", "0.1", "123456789.125"]
    while len(vals) % 3:
        vals.append("0")
    (d / "s.mtl").write_text("newmtl DiffuseWhite\nKd 0.15 0.5 1\n")
    lines = ["mtllib s.mtl", "g a", "usemtl DiffuseWhite"]
    for i in range(0, len(vals), 3):
        lines.append("v " + " ".join(vals[i:i + 3]))
    n = len(vals) // 3
    lines += [f"f {1 + (i % n)} {1 + ((i + 1) % n)} {1 + ((i + 2) % n)}" for i in range(n)]
    (d / "s.obj").write_text("\n".join(lines) + "\n")
    got = Model(str(d), "s", 0).vertices.reshape(-1, 3)[::3].ravel()[:len(vals)]  # first vertex of face i = vertex i
    exact = np.array([float(v) for v in vals])

    def tinyobj(s):  # the algorithm, restated once more in Python
        sign = -1.0 if s[0] == "-" else 1.0
        s = s.lstrip("+-")
        mant, _, ex = s.lower().partition("e")
        ip, _, fp = mant.partition(".")
        m = 0.0
        for ch in ip:
            m = m * 10 + int(ch)
        lut = [1.0, 0.1, 0.01, 0.001, 0.0001, 0.00001, 0.000001, 0.0000001]
        for k, ch in enumerate(fp, start=1):
            m += int(ch) * (lut[k] if k < len(lut) else 10.0 ** -k)
        e = int(ex) if ex else 0
        return sign * (np.ldexp(m * 5.0 ** e, e) if e else m)

    assert np.array_equal(got, np.array([tinyobj(v) for v in vals]))
    assert np.allclose(got, exact, rtol=1e-15, atol=0)
    assert got[6] == 0.1 + 5 * 0.01 and got[6] != 0.15  # not correctly rounded, like the reference's loader
    assert np.array_equal(Model(str(d), "s", capi.LOAD_STRTOD).vertices.reshape(-1, 3)[::3].ravel()[:len(vals)], exact)
    assert Model(str(d), "s", 0).materials[0].kd[0] == 0.1 + 5 * 0.01  # the MTL goes through the same parser


def test_model_rules(built, tmp_path):
    d = tmp_path / "r"
    d.mkdir()
    (d / "s.mtl").write_text(
        "newmtl Light\nKd 1 1 1\n\nnewmtl Mystery\nKd 0.25 0.5 0.75\nKs 1 1 1\nNs 50\n\nnewmtl material3\nKd 0.5 0.5 0.5\nKs 0.25 0.25 0.25\nNs 32\n"
        "newmtl Mirror\n\nnewmtl quad1\n\nnewmtl Wood\nmap_Kd -s 1 1 1 missing.png\n\nnewmtl Mystery\nKd 0 0 0\n")
    (d / "s.xml").write_text('<?xml version="1.0"?>\n<!-- c -->\n<camera type="perspective" width="320" height="200" fovy="39.5">\n'
                             '  <eye x="1" y="2" z="3"/> <lookat x="0" y="0.5" z="-1"/>\n</camera>\n<light mtlname="Light" radiance="17.5, 12 ,4e0"/>\n<light mtlname="x"/>\n')
    (d / "s.obj").write_text("\n".join([
        "mtllib s.mtl", "v 0 0 0", "v 1 0 0", "v 1 1 0", "v 0 1 0", "v 0 0 1", "v 2 0 1", "v 2 1 1", "v 0 1 1", "vt 0 0", "vt 1 0", "vt 1 1", "vt 0.5 0.5",
        "vn 0 0 1",
        "g", "usemtl Mystery", "f 1/1/1 2/2/1 3/3/1", "usemtl Mirror", "f 1 3 4",     # face-0 material rules the shape
        "g   ", "usemtl Light", "f 1/1 2/1 3/2",                                      # coinciding uv -> (0,0),(1,0),(1,1)
        "g Group_1 extra", "usemtl material3", "f 5 6 7 8",                            # quad: shorter diagonal 5-7 vs 6-8
        "o named object", "usemtl quad1", "f -4 -3 -2",
        "g", "usemtl Wood", "f 1//1 2//1 3//1", "g", "g emptygroup", ""]))
    m = Model(str(d), "s", 0)
    assert m.mesh_names == ["Group_0", "Group_1", "Group_1 extra", "named object", "Group_2"]
    assert list(m.mesh_first) == [0, 2, 3, 5, 6, 7]
    names = m.material_names
    assert [names[i] for i in m.mesh_material] == ["Mystery", "Light", "material3", "quad1", "Wood"]
    assert m.mesh_material[0] == names.index("Mystery")  # the FIRST of two materials of one name
    types = {n: t.type for n, t in zip(names, m.materials)}
    assert types["Light"] == capi.MAT_DIFFUSE_LIGHT and types["Mirror"] == capi.MAT_MIRROR and types["quad1"] == capi.MAT_EMPTY
    assert types["material3"] == capi.MAT_PHONG and types["Wood"] == capi.MAT_LAMBERTIAN and types["Mystery"] == capi.MAT_LAMBERTIAN
    assert tuple(m.materials[names.index("Light")].emission) == (17.5, 12.0, 4.0)
    assert m.materials[names.index("Mirror")].ns == 1.0 and tuple(m.materials[names.index("Mirror")].kd) == (0.0, 0.0, 0.0)  # tinyobj defaults
    uv = m.texcoords
    assert np.array_equal(uv[0], [[0, 0], [1, 0], [1, 1]]) and np.array_equal(uv[2], [[0, 0], [1, 0], [1, 1]])
    assert np.array_equal(m.normals[0], [[0, 0, 1]] * 3) and not np.any(m.normals[1])
    v = m.vertices
    # |5-7|^2 = 5 < |6-8|^2 = 5 is false -> [0,1,3],[1,2,3]
    assert np.array_equal(v[3], [[0, 0, 1], [2, 0, 1], [0, 1, 1]]) and np.array_equal(v[4], [[2, 0, 1], [2, 1, 1], [0, 1, 1]])
    assert np.array_equal(v[5], [[0, 0, 1], [2, 0, 1], [2, 1, 1]])  # relative indices
    wood = m.materials[names.index("Wood")]
    assert wood.kd_tex == 0 and m.texture(0) is None  # failed load: cyan at render time (Texture.cpp:24)
    w = "\n".join(m.warnings)
    assert "missing.png" in w and "Same Name" in w and "<light>" in w
    c = m.camera()
    assert (c.image_width, c.image_height, c.fovy, tuple(c.eye), tuple(c.look_at), tuple(c.up)) == (320, 200, 39.5, (1, 2, 3), (0, 0.5, -1), (0, 1, 0))
    # the input file was not rewritten (the reference renames anonymous groups in place)
    assert (d / "s.obj").read_text().count("\ng\n") == 3


def test_loader_errors(built, tmp_path):
    with pytest.raises(PrtError):
        Model(str(tmp_path), "nothing")
    d = tmp_path / "e"
    d.mkdir()
    (d / "s.mtl").write_text("newmtl Light\n")
    (d / "s.obj").write_text("mtllib s.mtl\nv 0 0 0\nv 1 0 0\nv 0 1 0\nusemtl Light\nf 1 2 3\n")
    with pytest.raises(PrtError, match="radiance"):  # lightRadianceMap.at() throws in the reference
        Model(str(d), "s")
    (d / "s.obj").write_text("v 0 0 0\nv 1 0 0\nv 0 1 0\nf 1 2 3\n")
    with pytest.raises(PrtError, match="no material"):
        Model(str(d), "s")


def _png(path, img, **kw):
    with open(path, "wb") as f:
        f.write(scenes.png_bytes(img, **kw))


def test_png_decoder(built, tmp_path):
    rng = np.random.default_rng(5)
    for c in (1, 2, 3, 4):
        img = rng.integers(0, 256, size=(37, 53, c), dtype=np.uint8)
        img[5:20, 3:40] = img[5, 3]  # long runs: back-references with large lengths
        for level in (0, 1, 9):      # stored, fixed / dynamic Huffman
            p = str(tmp_path / f"t{c}_{level}.png")
            _png(p, img, level=level)
            assert np.array_equal(load_image(p), img)
    # palette + tRNS, 4-bit grey, 16-bit RGB: written by hand
    def chunk(t, d):
        return struct.pack(">I", len(d)) + t + d + struct.pack(">I", zlib.crc32(t + d) & 0xFFFFFFFF)

    def png(w, h, depth, ctype, rows, extra=b""):
        raw = b"".join(b"\x00" + r for r in rows)
        return b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, depth, ctype, 0, 0, 0)) + extra + chunk(b"IDAT", zlib.compress(raw)) + chunk(b"IEND", b"")

    pal = bytes([255, 0, 0, 0, 255, 0, 0, 0, 255, 9, 8, 7])
    (tmp_path / "pal.png").write_bytes(png(4, 2, 2, 3, [bytes([0b00011011]), bytes([0b11100100])], chunk(b"PLTE", pal) + chunk(b"tRNS", bytes([10, 20]))))
    got = load_image(str(tmp_path / "pal.png"))
    want = np.array([[[255, 0, 0, 10], [0, 255, 0, 20], [0, 0, 255, 255], [9, 8, 7, 255]], [[9, 8, 7, 255], [0, 0, 255, 255], [0, 255, 0, 20], [255, 0, 0, 10]]], dtype=np.uint8)
    assert np.array_equal(got, want)
    (tmp_path / "g4.png").write_bytes(png(3, 1, 4, 0, [bytes([0x0f, 0x80])]))
    assert np.array_equal(load_image(str(tmp_path / "g4.png")).ravel(), [0, 255, 0x88])
    (tmp_path / "rgb16.png").write_bytes(png(1, 1, 16, 2, [bytes([0x12, 0x34, 0xab, 0xcd, 0xff, 0x00])]))
    assert np.array_equal(load_image(str(tmp_path / "rgb16.png")).ravel(), [0x12, 0xab, 0xff])
    (tmp_path / "bad.png").write_bytes(b"\x89PNG\r\n\x1a\nxxxx")
    with pytest.raises(PrtError):
        load_image(str(tmp_path / "bad.png"))
    (tmp_path / "x.jpg").write_bytes(b"\xff\xd8\xff\xe0abcd")
    with pytest.raises(PrtError, match="JPEG"):
        load_image(str(tmp_path / "x.jpg"))


def jpeg_cases():
    """(name, PIL save options, picture) of the JPEG fixture set: baseline / progressive, every chroma layout PIL writes,
    grey, restart intervals, sizes that are not multiples of the MCU, smooth / noisy / hard-edged content"""
    import itertools
    rng = np.random.default_rng(5)
    out = []
    for (w, h), kind, sub, prog, q, grey, rst in itertools.product([(1, 1), (7, 5), (17, 33), (64, 48), (129, 64)], ["smooth", "noise", "checker"],
                                                                   [0, 1, 2, "4:4:0", "4:1:1"], [False, True], [35, 100], [False, True], [0, 3]):
        if grey and sub != 0:
            continue
        y, x = np.mgrid[0:h, 0:w]
        if kind == "smooth":
            img = np.stack([128 + 100 * np.sin(x / 7.0) * np.cos(y / 11.0), 128 + 90 * np.cos(x / 5.0 + y / 9.0), 30 + (x * 3 + y * 5) % 200], -1)
        elif kind == "noise":
            img = rng.integers(0, 256, (h, w, 3))
        else:
            img = ((x // 4 + y // 4) % 2 * 255)[..., None].repeat(3, -1) * np.array([1, 0.5, 0.2])
        img = np.clip(img, 0, 255).astype(np.uint8)
        kw = dict(quality=q, progressive=prog)
        if not grey:
            kw["subsampling"] = sub
        if rst:
            kw["restart_marker_blocks"] = rst
        out.append((f"{w}x{h}_{kind}_{str(sub).replace(':', '')}_{int(prog)}_{q}_{int(grey)}_{rst}", kw, img[..., 0] if grey else img))
    return out


def _stb_load(path):
    """stbi_load(path, &w, &h, &c, 0) of the stb_image.h the reference vendors, as compiled into oracle/_ref by the
    reference's own Texture.cpp (test infrastructure)"""
    import ctypes as C
    lib = ol.OracleLib("ref").lib
    lib.stbi_load.restype = C.POINTER(C.c_ubyte)
    lib.stbi_load.argtypes = [C.c_char_p, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int), C.c_int]
    w, h, c = C.c_int(), C.c_int(), C.c_int()
    ptr = lib.stbi_load(path.encode(), C.byref(w), C.byref(h), C.byref(c), 0)
    assert ptr, path
    a = np.ctypeslib.as_array(ptr, shape=(h.value, w.value, c.value)).copy()
    lib.stbi_image_free(ptr)
    return a


def test_jpeg_decoder_matches_stb_image(built, tmp_path):
    """f1: ImageTexture loads whatever stbi_load reads (Source/Texture.cpp:10-21), and a JPEG decoder's
    output is an implementation choice (IDCT, chroma upsampling, colour conversion): jpeg_decode.cpp must return stb's
    bytes. Live against the vendored stb_image where oracle/_ref is present (every file of jpeg_cases(), ~500), and
    always against the committed fixture (tests/golden/jpeg_fixture.npz: 60 files with the bytes stb decoded)."""
    from PIL import Image
    g = np.load(os.path.join(ROOT, "tests", "golden", "jpeg_fixture.npz"))
    names = [n[5:] for n in g.files if n.startswith("file_")]
    assert len(names) >= 40
    for n in names:
        p = str(tmp_path / "f.jpg")
        open(p, "wb").write(g["file_" + n].tobytes())
        got = load_image(p)
        want = g["stb_" + n]
        assert np.array_equal(got.reshape(want.shape), want), n
    if not ol.available("ref"):
        return
    checked = 0
    for name, kw, img in jpeg_cases():
        p = str(tmp_path / "t.jpg")
        try:
            Image.fromarray(img).save(p, "JPEG", **kw)
        except Exception:
            continue  # an option this PIL build does not know
        want = _stb_load(p)
        got = load_image(p)
        assert got.size == want.size and np.array_equal(got.reshape(want.shape), want), name
        checked += 1
    assert checked > 300


def test_jpeg_texture_in_a_scene(built, tmp_path):
    """a map_Kd that names a .jpg is loaded (round 1 reported it and rendered cyan): the model's texture holds stb's texels"""
    from PIL import Image
    sc = cases.bathroom_small()
    d = _write(tmp_path, sc)
    mtl = [f for f in os.listdir(d) if f.endswith(".mtl")][0]
    text = open(os.path.join(d, mtl)).read()
    pngs = [f for f in os.listdir(d) if f.endswith(".png")]
    first = sorted(pngs)[0]
    rgb = Image.open(os.path.join(d, first)).convert("RGB")
    rgb.save(os.path.join(d, "wall.jpg"), "JPEG", quality=92, subsampling=2)
    assert first in text
    open(os.path.join(d, mtl), "w").write(text.replace(first, "wall.jpg"))
    m = Model(d, "s")
    assert not any("wall.jpg" in w for w in m.warnings)
    tex = [t for t in (m.texture(i) for i in range(m.desc.n_textures)) if t is not None and t.shape[2] == 3]
    want = load_image(os.path.join(d, "wall.jpg"))
    assert any(t.shape == want.shape and np.array_equal(t, want) for t in tex)
    if ol.available("ref"):
        assert np.array_equal(want, _stb_load(os.path.join(d, "wall.jpg")))


def test_image_writers(built, tmp_path):
    rng = np.random.default_rng(7)
    img = rng.integers(0, 256, size=(70, 333, 3), dtype=np.uint8)  # > 65535 bytes: several stored blocks
    p = str(tmp_path / "o.png")
    write_png(p, img)
    assert np.array_equal(load_image(p), img)
    # an independent reader: zlib + filter-0 rows
    b = open(p, "rb").read()
    assert b[:8] == b"\x89PNG\r\n\x1a\n"
    pos, idat = 8, b""
    while pos < len(b):
        n, t = struct.unpack(">I", b[pos:pos + 4])[0], b[pos + 4:pos + 8]
        assert zlib.crc32(b[pos + 4:pos + 8 + n]) & 0xFFFFFFFF == struct.unpack(">I", b[pos + 8 + n:pos + 12 + n])[0]
        if t == b"IDAT":
            idat += b[pos + 8:pos + 8 + n]
        pos += 12 + n
    raw = np.frombuffer(zlib.decompress(idat), dtype=np.uint8).reshape(70, 1 + 333 * 3)
    assert not raw[:, 0].any() and np.array_equal(raw[:, 1:].reshape(70, 333, 3), img)
    # HDR: stb_image_write's float -> RGBE
    f = (rng.random((5, 9, 3)) * np.array([1e-3, 1.0, 300.0])).astype(np.float32)
    f[0, 0] = 0.0
    f[0, 1] = [1.0, 0.5, 0.25]
    h = str(tmp_path / "o.hdr")
    write_hdr(h, f)
    b = open(h, "rb").read()
    assert b.startswith(b"#?RADIANCE\n") and b"FORMAT=32-bit_rle_rgbe\n" in b and b"\n\n-Y 5 +X 9\n" in b
    px = np.frombuffer(b[b.index(b"+X 9\n") + 5:], dtype=np.uint8).reshape(5, 9, 4)
    mx = f.max(axis=2)
    m, e = np.frexp(mx)
    with np.errstate(divide="ignore", invalid="ignore"):
        norm = (m.astype(np.float32) * np.float32(256.0) / mx).astype(np.float32)
    want = np.where(mx[..., None] < 1e-32, 0, np.concatenate([(f * norm[..., None]).astype(np.uint8), (e + 128).astype(np.uint8)[..., None]], axis=2))
    assert np.array_equal(px, want.astype(np.uint8))
    assert tuple(px[0, 1]) == (128, 64, 32, 129)


# ---- GPU ------------------------------------------------------------------------------------------
@pytest.fixture(scope="module")
def gpu_ctx():
    import __graft_entry__ as g
    if not os.path.exists(capi.LIB_PATH):
        g.build()
    return Context(0)


@pytest.mark.gpu
@pytest.mark.parametrize("key", ["tiny", "bathroom"])
def test_loaded_scene_renders_like_the_array_scene(gpu_ctx, tmp_path, key):
    sc = cases.SCENES[key]()
    m = Model(_write(tmp_path, sc, anonymous_groups=(key == "tiny")), "s", capi.LOAD_STRTOD)
    cam, prm = m.camera(), make_params(3, 8, rng_key=4)
    a = Scene(gpu_ctx, m).render(cam, prm)
    b = Scene(gpu_ctx, sc).render(make_camera(sc.camera), prm)
    assert np.array_equal(a, b)
    # the reference's own golden for this scene, through the file loader
    name = {"tiny": "tiny_nee", "bathroom": "bathroom_nee"}[key]
    _, w, h, spp, depth, nee, rkey, bg = cases.RENDER_CASES[name]
    g = np.load(os.path.join(ROOT, "tests", "golden", f"render_{name}.npz"))
    cam.image_width, cam.image_height = w, h
    img = Scene(gpu_ctx, m).render(cam, make_params(spp, depth, sample_lights=nee, rng_key=rkey, background=bg))
    rel = np.abs(img - g["image"]) / np.maximum(np.abs(g["image"]), 1e-3)
    assert float((rel.max(axis=-1) < 1e-9).mean()) >= 0.99  # TOLERANCE: as tests/test_gpu_parity.py


@pytest.mark.gpu
def test_prt_main_drop_in(gpu_ctx, tmp_path):
    """the C++ replacement of the reference's main.cpp: files in, PNG + HDR out"""
    sc = cases.tiny()
    d = _write(tmp_path, sc)
    exe = os.path.join(ROOT, "pooraytracer_b200", "prt_main")
    out = str(tmp_path / "img")
    r = subprocess.run([exe, d, "s", "--spp", "6", "--depth", "8", "--chunk", "4", "--seed", "4", "--strtod", "--out", out],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "Shapes/Meshes Nums: 10" in r.stdout
    png = load_image(out + ".png")
    s = Scene(gpu_ctx, sc)
    cam = make_camera(sc.camera)
    img = s.render(cam, make_params(6, 8, rng_key=4))  # prt_main renders through prt_multi_render: one call, passes cut by the library
    rgb8, f32 = s.resolve(img)
    assert np.array_equal(png, rgb8)
    hdr = open(out + ".hdr", "rb").read()
    assert hdr.startswith(b"#?RADIANCE") and len(hdr) > cam.image_width * cam.image_height * 4
    import imgutil
    api.write_hdr(out + "_check.hdr", f32)
    assert np.array_equal(imgutil.hdr_rgbe(hdr), imgutil.hdr_rgbe(open(out + "_check.hdr", "rb").read()))
    assert "job: load" in r.stdout and "1 GPU" in r.stdout
