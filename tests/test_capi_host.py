"""CPU tests of the C-ABI library: it loads, exports every symbol include/prt_b200.h declares, does
the host-side work (reference BVH order, flattening) without a GPU, and REFUSES compute calls
without one (no CPU fallback)."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import cases
import oracle_lib as ol
from pooraytracer_b200 import Context, PrtError, Scene, _capi as capi, make_camera, make_params, scenes

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_every_declared_symbol_is_exported(built):
    hdr = open(os.path.join(ROOT, "include", "prt_b200.h")).read()
    declared = set(re.findall(r"^(?:int|void|const char\*|const prt_scene_desc\*)\s+(prt_[a-z_0-9]+)\(", hdr, flags=re.M))
    assert len(declared) >= 18
    lib = C.CDLL(capi.LIB_PATH)
    for name in sorted(declared):
        assert hasattr(lib, name), f"{name} declared in include/prt_b200.h but not exported"
    assert declared == set(capi.PROTOTYPES), "ctypes mirror out of sync with the header"
    assert lib.prt_abi_version() == 2


def test_struct_sizes_match_header(built):
    # sizes a C compiler gives the header's structs (computed by hand from the field lists)
    assert C.sizeof(capi.PrtMaterial) == 8 + 8 * (3 + 3 + 1 + 3 + 2 + 3 + 3)
    assert C.sizeof(capi.PrtCamera) == 8 + 8 * 10
    assert C.sizeof(capi.PrtRenderParams) == 24 + 8 * 5
    assert C.sizeof(capi.PrtStats) == 8 * 6 + 8 * 4 + 8 + 16 + 8 + 8


@pytest.mark.parametrize("key", ["tiny", "cornell", "veach", "bathroom", "ties"])
def test_reference_order_matches_oracle(built, key):
    sc = cases.tie_scene() if key == "ties" else cases.SCENES[key]()
    ctx = Context(-1)
    s = Scene(ctx, sc, accel=capi.ACCEL_SAH_BVH2)
    rank, lo = s.reference_order()
    g = np.load(os.path.join(ROOT, "tests", "golden", f"rays_{key}.npz"))
    assert np.array_equal(rank, g["world_rank"]) and np.array_equal(lo, g["light_order"])
    st = s.stats()
    assert st.n_triangles == sc.n_triangles and st.tri_bytes == 128 and st.node_bytes == 64 and st.n_nodes >= 1
    s.commit(capi.ACCEL_REFERENCE_BVH2)
    assert s.stats().n_nodes >= sc.n_triangles // 2


def test_reference_order_large_mesh(built):
    sc = scenes.terrain(120, 90)
    s = Scene(Context(-1), sc)
    O = ol.OracleScene("orc", sc, 0)
    assert np.array_equal(s.reference_order()[0], O.reference_order()[0])


def test_compute_calls_fail_loudly_without_gpu(built):
    sc = cases.tiny()
    s = Scene(Context(-1), sc)
    o, d = scenes.random_rays(sc, 8)
    with pytest.raises(PrtError) as e:
        s.trace_closest(o, d)
    assert e.value.code == capi.PRT_ERR_NO_DEVICE
    with pytest.raises(PrtError) as e:
        s.trace_any(o, d)
    assert e.value.code == capi.PRT_ERR_NO_DEVICE
    with pytest.raises(PrtError) as e:
        s.render(make_camera(sc.camera), make_params(1, 1))
    assert e.value.code == capi.PRT_ERR_NO_DEVICE
    with pytest.raises(PrtError) as e:
        s.sample_lights(o, np.zeros((8, 4)))
    assert e.value.code == capi.PRT_ERR_NO_DEVICE
    # the GPU-built BVH has no host version either
    with pytest.raises(PrtError) as e:
        Scene(Context(-1), sc, accel=capi.ACCEL_LBVH4_DEVICE)
    assert e.value.code == capi.PRT_ERR_NO_DEVICE


def test_whole_box_handle_and_noalloc_entry_fail_loudly_without_gpu(built):
    """prt_multi_create brings the other GPUs' contexts up on threads of its own: without a GPU every one of them must
    fail into ONE clean error (no hang, no crash on the way out); the entry point without allocations refuses a host-only
    scene like every other compute call."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is visible")
    from pooraytracer_b200 import MultiScene
    for devices in ([0], [0, 1], [0, 1, 2, 3, 4, 5, 6, 7]):
        with pytest.raises(PrtError) as e:
            MultiScene(devices, cases.tiny())
        assert e.value.code == capi.PRT_ERR_NO_DEVICE
    s = Scene(Context(-1), cases.tiny())
    rc = s.lib.prt_trace_closest_device_noalloc(s.handle, 8, 1, 1, 1, 1, 1, None, None, None, None)
    assert rc == capi.PRT_ERR_NO_DEVICE
    assert s.lib.prt_multi_reduce_kind(None) == capi.REDUCE_NONE


def test_bad_descriptions_are_rejected(built):
    sc = cases.tiny()
    bad = scenes.SceneData(name="bad", vertices=sc.vertices, texcoords=sc.texcoords, mesh_first=sc.mesh_first.copy(),
                           mesh_material=sc.mesh_material.copy(), mesh_names=sc.mesh_names, materials=sc.materials,
                           textures=sc.textures, camera=sc.camera)
    bad.mesh_material[0] = 99
    with pytest.raises(PrtError) as e:
        Scene(Context(-1), bad)
    assert e.value.code == capi.PRT_ERR_INVALID_ARGUMENT
    bad.mesh_material[0] = 0
    bad.mesh_first[-1] -= 1
    with pytest.raises(PrtError):
        Scene(Context(-1), bad)


def test_accel_auto_picks_by_mesh_size(built):
    """PRT_ACCEL_AUTO: 128-byte FP32 nodes for small meshes, 64-byte quantised ones from 262144 triangles up."""
    small = Scene(Context(-1), cases.tiny(), accel=capi.ACCEL_AUTO)
    assert small.stats().node_bytes == 128
    big = scenes.terrain(400, 330)
    assert big.n_triangles >= 262144
    assert Scene(Context(-1), big, accel=capi.ACCEL_AUTO).stats().node_bytes == 64
    assert Scene(Context(-1), big, accel=capi.ACCEL_SAH_BVH4).stats().node_bytes == 128


def test_empty_scene_host(built):
    empty = scenes.SceneData(name="empty", vertices=np.zeros((0, 3, 3)), texcoords=None, mesh_first=np.zeros(1, np.int64),
                             mesh_material=np.zeros(0, np.int32), mesh_names=[], materials=[], textures=[], camera=cases.tiny().camera)
    s = Scene(Context(-1), empty)
    rank, lo = s.reference_order()
    assert len(rank) == 0 and len(lo) == 0


def test_product_never_touches_the_oracle():
    """the shipped package must not import, link or load anything under oracle/"""
    pkg = os.path.join(ROOT, "pooraytracer_b200")
    for dirpath, _, files in os.walk(pkg):
        if "build" in dirpath:
            continue
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")) or f == "Makefile":
                txt = open(os.path.join(dirpath, f), errors="replace").read()
                assert "oracle_lib" not in txt and "libprt_oracle" not in txt and "libprt_ref" not in txt, f
                assert "/root/reference" not in txt or f.endswith((".cpp", ".cu", ".cuh", ".h")), f


def test_resolve_thresholds_are_the_reference_quantiser_steps(built):
    """f2: the 254 steps of trunc(clamp(LinearToSRGB(c), 0, 0.9999) * 255) (Camera.cpp:216-221,297-300). The library
    finds them on the host with glibc's pow; tests/golden/output_stage.npz holds the steps located by bisection
    through PNGs the compiled reference wrote. Bit-equal, all 254."""
    from pooraytracer_b200 import _capi
    lib = _capi.load_library()
    thr = np.zeros(254)
    assert lib.prt_resolve_thresholds(thr.ctypes.data_as(C.c_void_p)) == 0
    g = np.load(os.path.join(ROOT, "tests", "golden", "output_stage.npz"))
    assert np.array_equal(thr, g["steps"])
    assert (np.diff(thr) > 0).all() and thr[0] > 0 and thr[-1] < 1


def test_hdr_decoder_reads_the_reference_file():
    """tests/imgutil.hdr_rgbe on the file stbi_write_hdr produced (run-length scanlines): mantissas / exponents
    reproduce float32(image) to RGBE precision"""
    import imgutil
    g = np.load(os.path.join(ROOT, "tests", "golden", "output_stage.npz"))
    rgbe = imgutil.hdr_rgbe(g["hdr_file"].tobytes())
    img = np.where(np.isnan(g["image"]), 0.0, g["image"])
    assert rgbe.shape[:2] == img.shape[:2]
    with np.errstate(all="ignore"):
        val = rgbe[..., :3].astype(np.float64) * np.exp2(rgbe[..., 3:4].astype(np.float64) - 136.0)
        mx = np.max(img.astype(np.float32), axis=-1, keepdims=True).astype(np.float64)
    ok = np.isfinite(mx[..., 0]) & (mx[..., 0] >= 1e-32)
    assert ok.sum() > 1000
    assert (np.abs(val - np.clip(img, 0, None))[ok] <= (mx / 128.0 + 1e-30)[ok[..., None].repeat(1, -1)[..., 0]]).all()


def test_hdr_writer_payload_equals_reference(tmp_path):
    """prt_write_hdr (host code) on float32(golden image): the RGBE quadruples stbi_write_hdr wrote for the reference
    (Camera.cpp:308-327), including the inf / negative / denormal pixels"""
    import imgutil
    from pooraytracer_b200 import api
    g = np.load(os.path.join(ROOT, "tests", "golden", "output_stage.npz"))
    with np.errstate(all="ignore"):
        f32 = np.where(np.isnan(g["image"]), 0.0, g["image"]).astype(np.float32)
    path = str(tmp_path / "x.hdr")
    api.write_hdr(path, f32)
    assert np.array_equal(imgutil.hdr_rgbe(open(path, "rb").read()), imgutil.hdr_rgbe(g["hdr_file"].tobytes()))
