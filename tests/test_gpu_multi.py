"""The whole-box layer behind the C ABI (prt_multi_*, pooraytracer_b200/csrc/multi.cu): scene replicated,
samples sharded in C++, ONE reduction of the FP64 frames to the first GPU (peer loads over NVLink by default, ncclReduce
under PRT_MULTI_REDUCE=nccl: both are run), output stage after the reduce. Runs on however many GPUs
the box shows: with one, the handle must be the single-GPU path bit for bit; with more, the reduced image equals the one-GPU
image up to the order of the final FP64 additions (TOLERANCE: 1e-13 relative), and ray-query batches split into contiguous
ranges return the one-GPU answers exactly. Replaces /root/reference/main.cpp:36-52 for a C / C++ caller."""
import os
import subprocess

import numpy as np
import pytest

import cases
from pooraytracer_b200 import MultiScene, Scene, _capi as capi, make_camera, make_params, scenes

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def n_gpus():
    import torch
    return torch.cuda.device_count()


def test_one_gpu_handle_is_the_single_gpu_path(gpu_ctx):
    sc = cases.tiny()
    cam = make_camera(sc.camera)
    prm = make_params(6, 8, rng_key=4)
    one = Scene(gpu_ctx, sc, accel=capi.ACCEL_SAH_BVH4)
    box = MultiScene([0], sc)
    assert box.reduce_kind == "none"
    img = box.render(cam, prm)
    assert np.array_equal(img, one.render(cam, prm))
    st, reduce_ms = box.stats()
    assert st.paths == cam.image_width * cam.image_height * 6 and st.extend_rays > 0 and st.ref_stack_overflows == 0
    rgb8, f32 = box.resolve(cam.image_width, cam.image_height)
    want8, want32 = one.resolve(img)
    assert np.array_equal(rgb8, want8) and np.array_equal(f32, want32)
    o, d = cases.ray_batch(sc)
    for a, b in zip(box.trace_closest(o, d), one.trace_closest(o, d)):
        assert np.array_equal(a, b)
    assert np.array_equal(box.trace_any(o, d, 1e-3, 2.0), one.trace_any(o, d, 1e-3, 2.0))
    # a sample sub-range, as a caller that splits the job itself would ask for
    part = box.render(cam, make_params(6, 8, rng_key=4, spp_begin=2, spp_end=5))
    assert np.array_equal(part, one.render(cam, make_params(6, 8, rng_key=4, spp_begin=2, spp_end=5)))


def test_scene_clone_is_a_faithful_replica(gpu_ctx):
    """prt_scene_clone (how prt_multi_create replicates the scene): device-to-device copies of every buffer, pointer table
    rebound. On one GPU the replica lives in a second context of the same device; with more, on the last device. Same
    image, same hits, bit for bit; the original may be destroyed first."""
    from pooraytracer_b200 import Context
    sc = cases.bathroom_small()  # textures, mirrors, several material classes
    cam = make_camera(sc.camera, 48, 27)
    prm = make_params(3, 8, rng_key=5)
    for accel in (capi.ACCEL_SAH_BVH4, capi.ACCEL_LBVH4_DEVICE, capi.ACCEL_SAH_BVH4Q, capi.ACCEL_REFERENCE_BVH2):
        one = Scene(gpu_ctx, sc, accel=accel)
        want = one.render(cam, prm)
        o, d = cases.ray_batch(sc)
        want_hits = one.trace_closest(o, d)
        ctx2 = Context(n_gpus() - 1)
        rep = one.clone(ctx2)
        one.close()
        assert np.array_equal(rep.render(cam, prm), want), accel
        for a, b in zip(rep.trace_closest(o, d), want_hits):
            assert np.array_equal(a, b)
        st = rep.stats()
        assert st.n_triangles == sc.n_triangles and st.ref_stack_overflows == 0
        rep.close()


def test_bad_arguments_are_refused():
    sc = cases.tiny()
    with pytest.raises(Exception):
        MultiScene([0, 0], sc)  # a device listed twice
    with pytest.raises(Exception):
        MultiScene([], sc)
    with pytest.raises(Exception):
        MultiScene([99], sc)


@pytest.mark.parametrize("kind", ["peer", "nccl"])
@pytest.mark.parametrize("spp", [1, 5, 16])
def test_all_visible_gpus_reduce_to_the_one_gpu_image(gpu_ctx, spp, kind, monkeypatch):
    """spp 1: more GPUs than samples (the idle ones only join the reduce); 5: uneven slices; 16: even.
    kind: k_sum_peers on the first GPU (the default on an NVLink box) or ncclReduce."""
    n = n_gpus()
    if n < 2:
        pytest.skip("one GPU visible: a reduce needs two")
    monkeypatch.setenv("PRT_MULTI_REDUCE", kind)
    sc = cases.cornell_small()
    cam = make_camera(sc.camera, 96, 96)
    prm = make_params(spp, 10, rng_key=11)
    one = Scene(gpu_ctx, sc, accel=capi.ACCEL_SAH_BVH4).render(cam, prm)
    box = MultiScene(list(range(n)), sc)
    assert box.reduce_kind == kind
    img = box.render(cam, prm)
    assert np.allclose(img, one, rtol=1e-13, atol=1e-15)
    st, reduce_ms = box.stats()
    assert st.paths == 96 * 96 * spp and reduce_ms > 0
    rgb8, _ = box.resolve(96, 96)
    assert (rgb8 == Scene(gpu_ctx, sc).resolve(one)[0]).mean() > 0.9999
    # determinism of the reduce
    assert np.array_equal(box.render(cam, prm), img)


def test_default_reduce_is_peer_loads_and_odd_frames_work(gpu_ctx, monkeypatch):
    """without PRT_MULTI_REDUCE an NVLink box gets the peer-load kernel (no communicator to initialise); a frame with an
    odd number of doubles (5 x 3 x 3) exercises the kernel's scalar tail"""
    n = n_gpus()
    if n < 2:
        pytest.skip("one GPU visible")
    monkeypatch.delenv("PRT_MULTI_REDUCE", raising=False)
    sc = cases.tiny()
    box = MultiScene(list(range(n)), sc)
    assert box.reduce_kind == "peer"
    cam = make_camera(sc.camera, 5, 3)
    prm = make_params(2 * n + 1, 6, rng_key=3)
    one = Scene(gpu_ctx, sc, accel=capi.ACCEL_SAH_BVH4).render(cam, prm)
    img = box.render(cam, prm)
    assert np.allclose(img, one, rtol=1e-13, atol=1e-15)
    assert np.array_equal(box.render(cam, prm), img)


def test_ray_batches_split_across_gpus(gpu_ctx):
    n = n_gpus()
    if n < 2:
        pytest.skip("one GPU visible")
    sc = scenes.terrain(256, 128)
    o, d = scenes.random_rays(sc, 100_003)
    one = Scene(gpu_ctx, sc, accel=capi.ACCEL_SAH_BVH4)
    box = MultiScene(list(range(n)), sc)
    for a, b in zip(box.trace_closest(o, d), one.trace_closest(o, d)):
        assert np.array_equal(a, b)
    assert np.array_equal(box.trace_any(o, d, 1e-3, 0.4), one.trace_any(o, d, 1e-3, 0.4))


def test_prt_main_on_all_gpus(gpu_ctx, tmp_path):
    """the drop-in for main.cpp with --gpus N: same PNG as one GPU (bytes may differ only where an ulp of the reduce
    order crosses a quantiser step: none on this frame)"""
    n = n_gpus()
    if n < 2:
        pytest.skip("one GPU visible")
    sc = cases.tiny()
    d = str(tmp_path / "scene")
    os.makedirs(d)
    scenes.write_scene_files(sc, d, "s")
    exe = os.path.join(ROOT, "pooraytracer_b200", "prt_main")
    outs = []
    for g in (1, n):
        out = str(tmp_path / f"img{g}")
        r = subprocess.run([exe, d, "s", "--spp", "8", "--depth", "8", "--seed", "4", "--gpus", str(g), "--out", out], capture_output=True, text=True, timeout=600)
        assert r.returncode == 0, r.stdout + r.stderr
        outs.append(open(out + ".png", "rb").read())
    import imgutil
    a, b = imgutil.png_rgb8(outs[0]), imgutil.png_rgb8(outs[1])
    assert (a == b).mean() > 0.9999
