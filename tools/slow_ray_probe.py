"""Runs one workload on a -DPRT_DEBUG_SLOW_RAY=<fetches> build (tools/build_variants.sh "slowray:-DPRT_DEBUG_SLOW_RAY=1500ull",
PRT_B200_LIB=.../lib_slowray.so): the library prints every ray whose walk fetched more records than that.
usage: python tools/slow_ray_probe.py bath|soup|terrain"""
import sys, os
sys.path.insert(0, '.')
import torch, numpy as np
from pooraytracer_b200 import Context, Scene, _capi as capi, scenes, make_camera, make_params
ctx = Context(0)
which = sys.argv[1]
if which == "bath":
    sc = scenes.bathroom_like()
    s = Scene(ctx, sc, accel=capi.ACCEL_SAH_BVH4)
    s.set_counters(True)
    cam = make_camera(sc.camera, 1280, 720)
    acc = torch.zeros(1280 * 720 * 3, dtype=torch.float64, device="cuda")
    st = torch.cuda.Stream()
    prm = make_params(3000, 50, rr=0.8, rng_key=20251104, spp_begin=9 * 16, spp_end=10 * 16)
    s.render_device(cam, prm, acc.data_ptr(), st.cuda_stream)
    print("gpu_ms", s.stats().gpu_ms)
else:
    sc = scenes.soup(1_000_000) if which == "soup" else scenes.terrain(2500, 2000)
    s = Scene(ctx, sc, accel=capi.ACCEL_SAH_BVH4)
    s.set_counters(True)
    o, d = scenes.random_rays(sc, 1 << 22)
    s.trace_closest(o, d, 1e-4, np.inf)
    print("gpu_ms", s.stats().gpu_ms)
