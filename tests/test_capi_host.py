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
