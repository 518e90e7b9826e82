"""Per-step device time of bathroom2 (how the 101 .. 204 ms spread of profiles/r1_s13_grazing_rays.txt was seen)."""
import sys, time, os
sys.path.insert(0, '.')
import torch
from pooraytracer_b200 import Context, Scene, _capi as capi, scenes, make_camera, make_params
ctx = Context(0)
sc = scenes.bathroom_like()
s = Scene(ctx, sc, accel=capi.ACCEL_SAH_BVH4)
cam = make_camera(sc.camera, 1280, 720)
acc = torch.zeros(1280 * 720 * 3, dtype=torch.float64, device="cuda")
st = torch.cuda.Stream()
for flags in (0, capi.RENDER_STAGE_TIMING):
    for step in (0, 2, 3, 9, 3, 9):
        prm = make_params(3000, 50, rr=0.8, rng_key=20251104, spp_begin=step * 16, spp_end=(step + 1) * 16, flags=flags)
        s.render_device(cam, prm, acc.data_ptr(), st.cuda_stream)
        x = s.stats()
        print(step, "flags", flags, "gpu_ms", round(x.gpu_ms, 2), "ext/shade/shadow", round(x.extend_ms, 2), round(x.shade_ms, 2), round(x.shadow_ms, 2), "rays", x.extend_rays, x.shadow_rays, flush=True)
os.environ["PRT_TAIL_PATHS"] = "0"
for step in (3, 9):
    prm = make_params(3000, 50, rr=0.8, rng_key=20251104, spp_begin=step * 16, spp_end=(step + 1) * 16)
    s.render_device(cam, prm, acc.data_ptr(), st.cuda_stream)
    x = s.stats()
    print(step, "no tail: gpu_ms", round(x.gpu_ms, 2), flush=True)
