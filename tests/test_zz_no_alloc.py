"""prt_trace_closest_device_noalloc — the one entry point that allocates nothing (scratch in a static device array).

Collected last on purpose. Checked against the existing entry point
prt_trace_closest_device on the same device arrays, bit for bit (the tolerance of every ray-query test here is
exact equality), on two calls in a row whose second call is SHORTER than the first and uses other rays, so that
anything the shared array kept from the first call would show."""
import numpy as np
import pytest

import cases
from pooraytracer_b200 import Scene, _capi as capi, scenes

pytestmark = pytest.mark.gpu


def _device_rays(torch, o, d, tmin=1e-4, tmax=float("inf")):
    n = len(o)
    return dict(n=n, o=torch.from_numpy(np.ascontiguousarray(o)).cuda(), d=torch.from_numpy(np.ascontiguousarray(d)).cuda(),
                tmin=torch.full((n,), tmin, dtype=torch.float64, device="cuda"),
                tmax=torch.full((n,), tmax, dtype=torch.float64, device="cuda"))


def _outputs(torch, n):
    # poisoned: a ray the kernel skipped would keep the poison
    return dict(prim=torch.full((n,), -7, dtype=torch.int32, device="cuda"), t=torch.full((n,), -1.0, dtype=torch.float64, device="cuda"),
                a=torch.full((n,), -1.0, dtype=torch.float64, device="cuda"), b=torch.full((n,), -1.0, dtype=torch.float64, device="cuda"))


def _call(fn, scene, r, out):
    rc = fn(scene.handle, r["n"], r["o"].data_ptr(), r["d"].data_ptr(), r["tmin"].data_ptr(), r["tmax"].data_ptr(),
            out["prim"].data_ptr(), out["t"].data_ptr(), out["a"].data_ptr(), out["b"].data_ptr(), None)
    assert rc == 0, scene.lib.prt_last_error().decode()


def _same(torch, got, want):
    for k in ("prim", "t", "a", "b"):
        assert torch.equal(got[k], want[k]), k


@pytest.mark.parametrize("sort_min", [None, "1"])
def test_noalloc_equals_the_allocating_entry_point_twice_in_a_row(gpu_ctx, monkeypatch, sort_min):
    import torch
    if sort_min is None:
        monkeypatch.delenv("PRT_SORT_MIN", raising=False)
    else:
        monkeypatch.setenv("PRT_SORT_MIN", sort_min)  # ray binning: keys, order, histogram and cursors in the static array too
    sc = scenes.terrain(96, 64)
    scene = Scene(gpu_ctx, sc)
    lib = scene.lib
    o1, d1 = scenes.random_rays(sc, 60_000)
    o2, d2 = scenes.camera_rays(sc.camera, 160, 100)
    batches = [_device_rays(torch, o1, d1), _device_rays(torch, o2, d2)]
    assert batches[1]["n"] < batches[0]["n"]
    want = []
    for r in batches:  # the existing entry point first: its result is the yardstick (and the kernels are loaded)
        out = _outputs(torch, r["n"])
        torch.cuda.synchronize()
        _call(lib.prt_trace_closest_device, scene, r, out)
        gpu_ctx.synchronize()
        want.append(out)
        assert int((out["prim"] >= 0).sum()) > r["n"] // 10 and int((out["prim"] == -7).sum()) == 0
    # set-up alone (n = 0), then nothing may be taken from the device allocator
    assert lib.prt_trace_closest_device_noalloc(scene.handle, 0, None, None, None, None, None, None, None, None, None) == 0
    outs = [_outputs(torch, r["n"]) for r in batches]
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    for r, out in zip(batches, outs):  # two calls in a row on the shared array, the second shorter
        _call(lib.prt_trace_closest_device_noalloc, scene, r, out)
    gpu_ctx.synchronize()
    assert torch.cuda.mem_get_info()[0] >= free_before, "device memory was allocated"
    for out, w in zip(outs, want):
        _same(torch, out, w)
    # and once more in the other order, into fresh outputs
    for r, w in list(zip(batches, want))[::-1]:
        out = _outputs(torch, r["n"])
        torch.cuda.synchronize()
        _call(lib.prt_trace_closest_device_noalloc, scene, r, out)
        gpu_ctx.synchronize()
        _same(torch, out, w)


def test_noalloc_refuses_what_the_static_scratch_cannot_hold(gpu_ctx):
    scene = Scene(gpu_ctx, cases.tiny())
    lib = scene.lib
    rc = lib.prt_trace_closest_device_noalloc(scene.handle, capi.NOALLOC_MAX_RAYS + 1, 1, 1, 1, 1, 1, None, None, None, None)
    assert rc == capi.PRT_ERR_INVALID_ARGUMENT
    msg = lib.prt_last_error().decode()
    assert "1048577" in msg and "1048576" in msg
    assert lib.prt_trace_closest_device_noalloc(scene.handle, -1, None, None, None, None, None, None, None, None, None) == capi.PRT_ERR_INVALID_ARGUMENT
    assert lib.prt_trace_closest_device_noalloc(scene.handle, 8, None, None, None, None, None, None, None, None, None) == capi.PRT_ERR_INVALID_ARGUMENT


def test_noalloc_at_the_limit(gpu_ctx, monkeypatch):
    """PRT_NOALLOC_MAX_RAYS rays exactly, binned: the last byte of the keys plane is inside the array."""
    import torch
    monkeypatch.setenv("PRT_SORT_MIN", "1")
    sc = scenes.terrain(64, 48)
    scene = Scene(gpu_ctx, sc)
    o, d = scenes.random_rays(sc, capi.NOALLOC_MAX_RAYS)
    r = _device_rays(torch, o, d)
    want, got = _outputs(torch, r["n"]), _outputs(torch, r["n"])
    torch.cuda.synchronize()
    _call(scene.lib.prt_trace_closest_device, scene, r, want)
    _call(scene.lib.prt_trace_closest_device_noalloc, scene, r, got)
    gpu_ctx.synchronize()
    _same(torch, got, want)
