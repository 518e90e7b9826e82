import sys, time, os
sys.path.insert(0, '.')
from pooraytracer_b200 import Context, Scene, _capi as capi, scenes
ctx = Context(0)
for name, sc in (("bathroom", scenes.bathroom_like()), ("terrain10M", scenes.terrain(2500, 2000))):
    s = Scene(ctx, sc, accel=None)
    for rep in range(2):
        t0 = time.time(); s.commit(capi.ACCEL_LBVH4_DEVICE); t = time.time() - t0
        print(name, "rep", rep, "commit_s", round(t, 3), "build_ms", round(s.stats().build_ms, 2), "n_nodes", s.stats().n_nodes, flush=True)
