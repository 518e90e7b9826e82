#!/usr/bin/env python
"""bench.py — the reference's headline render on B200, plus the rest of BASELINE.json's metric in the same line.

HEADLINE (`value`, `e2e`, `roofline`, `cpu_baseline`): cornell-box 1024x1024, depth 100, NEE, RR 0.8 — BASELINE.json
configs[1]; the reference published 20382 s for 3000 spp = 0.1472 spp/s on an unstated CPU. A STEP renders
`--spp-per-step` samples per pixel of that frame on every GPU (default 64 = one wavefront pass of 67 M paths) through
the full depth-100 pipeline; throughput is intensive, so the slice measures the named configuration, not a smaller one.
The scene files of the reference are absent (example-scenes-cg24/ is empty): the scenes are the procedural stand-ins of
pooraytracer_b200/scenes.py in the reference's material vocabulary.

  value      whole-job spp/s, scene and accumulators resident in HBM, CUDA events on the launching stream, max over
             ranks; under torchrun every rank renders its own sample range of the step and ONE NCCL reduce per step
             sums the float64 accumulators on rank 0 (inside the timed region)
  e2e        the same through the C ABI a user calls — prt_multi_render on all N GPUs of the box from one process, HOST
             image out: sample sharding, the reduction of the frames on the first GPU (peer loads over NVLink; NCCL where the GPUs
             cannot address each other) and the device-to-host copy are all inside the
             timed region; `image_checksum` is of a fixed 64-sample range and is the same for every N (1e-13)
  roofline   the traversal kernels (k_extend + k_shadow): `achieved` = algorithmic bytes (BVH nodes x node bytes +
             triangles x 128 B actually fetched, counted by an instrumented run of the same step) / their summed
             durations, against the measured HBM peak — a cache-oblivious DEMAND figure, named as such; what ncu
             names as the limiter (`limiter`, `l1tex_pct`, `simd_lanes`, `dram_frac`) comes from the committed ncu pass
  scenes     the other two render configs (veach-mis 1280x720 d100, bathroom2 1280x720 d50), same figures, through
             prt_multi_render on all N GPUs
  rayquery   configs[4]: 2^24 random / coherent rays, closest / any, terrain + soup at 1 M and 10 M triangles (N = 1), with
             the reference's BVHNode::Hit over every 16th ray on the host cores beside the 1 M rows
  job        wall time of whole jobs, strong-scaled over the N GPUs: the 3000-spp cornell-box frame through the main.cpp
             drop-in (files -> PNG + HDR), and bathroom2 at 500 spp / depth 50 (arrays -> PNG)
  cpu_baseline / --impl reference
             the unmodified reference's Camera::Render (oracle/_ref, built from /root/reference) on the host cores: the
             FAIR variant (thread-local rand(), corrected static-init order) is the baseline; the AS-SHIPPED variant
             (CMake link order, glibc rand() with its global lock) is timed beside it
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

PUBLISHED_SPP_S = 3000.0 / 20382.22  # cornell-box 1024x1024 spp 3000 depth 100
RR = 0.8
KEY = 20251104
# BASELINE.json configs[1..3]: frame, depth, total spp of the job
WORKLOADS = {
    "cornell-box": dict(width=1024, height=1024, depth=100, total_spp=3000),
    "veach-mis": dict(width=1280, height=720, depth=100, total_spp=12000),
    "bathroom2": dict(width=1280, height=720, depth=50, total_spp=500),
}
ACCELS = {"bvh4": "ACCEL_SAH_BVH4", "bvh2": "ACCEL_SAH_BVH2", "lbvh": "ACCEL_LBVH4_DEVICE", "bvh4q": "ACCEL_SAH_BVH4Q", "auto": "ACCEL_AUTO"}
ACCEL_TEXT = {"bvh4": "SAH BVH4 (128 B nodes", "bvh2": "SAH BVH2 (64 B nodes", "lbvh": "device-built LBVH4 (128 B nodes",
              "bvh4q": "SAH BVH4, 8-bit quantised child boxes (64 B nodes",
              "auto": "SAH BVH4, layout chosen by mesh size (FP32 128 B nodes below 262144 triangles, 8-bit quantised 64 B nodes from there up"}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--spp-per-step", type=int, default=64)
    ap.add_argument("--width", type=int, default=0)
    ap.add_argument("--height", type=int, default=0)
    ap.add_argument("--depth", type=int, default=0)
    ap.add_argument("--scene", default="cornell-box", choices=list(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true", help="skip every host-CPU leg")
    ap.add_argument("--headline-only", action="store_true", help="only the headline workload: no scenes / rayquery / job sub-objects (A/B runs, ncu)")
    ap.add_argument("--no-rayquery-10m", action="store_true", help="skip the 10 M-triangle ray-query rows (their host-side builds take ~1 min)")
    ap.add_argument("--accel", default=os.environ.get("PRT_BENCH_ACCEL", "auto"), choices=list(ACCELS))
    ap.add_argument("--cpu-spp", type=int, default=0,
                    help="spp of the bounded CPU sample (0 = 12 for the cpu_baseline of the b200 arm, about 13 s of 16 host cores on the default workload; 2 per step for --impl reference)")
    ap.add_argument("--max-leaf", type=int, default=0, help="triangles per BVH leaf (0 = the library's default, 4)")
    a = ap.parse_args()
    w = WORKLOADS[a.scene]
    a.width = a.width or w["width"]
    a.height = a.height or w["height"]
    a.depth = a.depth or w["depth"]
    a.total_spp = w["total_spp"]
    return a


def make_scene(name):
    from pooraytracer_b200 import scenes
    if name == "cornell-box":
        return scenes.cornell_box(1024, 1024)
    if name == "veach-mis":
        return scenes.veach_mis()
    return scenes.bathroom_like()


def workload_name(scene, width, height, depth, spp_per_step, total_spp):
    return (f"{scene} {width}x{height} depth={depth} NEE RR={RR} (procedural stand-in for the absent "
            f"example-scenes-cg24 files); step = {spp_per_step} spp of the {total_spp}-spp render")


def peaks():
    try:
        p = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(p["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (of measured)"
    except Exception:
        return 6650.0, "fallback 6650 (of fallback)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200",
                                          "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2]))
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ---------------------------------------------------------------------------------------------------------------------
# host-CPU legs (test infrastructure under oracle/: the checker timed as a baseline, never the thing shipped)
# ---------------------------------------------------------------------------------------------------------------------
def cpu_reference_leg(scene_name, width, height, depth, total_spp, spp, steps=1, warmup=0, variant="fair"):
    """The reference's own Camera::Render (unmodified sources, oracle/_ref) on the host cores. variant "fair": corrected
    static-init order + thread-local rand() (glibc's rand() takes a global lock: negative thread scaling); "shipped":
    the CMake link order (AABB::empty defect) + glibc rand(), as a user building the repository gets it. Falls back to
    the oracle port when oracle/_ref is absent. Returns (spp/s, info)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as ol
    from pooraytracer_b200 import make_camera, make_params
    if variant == "shipped":
        if not ol.available("ref_shipped"):
            return None, None
        kind, rng = "ref_shipped", ol.RNG_LIBC
    else:
        kind = "ref" if ol.available("ref") else "orc"
        rng = ol.RNG_THREAD_LOCAL
        if kind == "orc" and not ol.available("orc"):
            import __graft_entry__ as g
            g.build()
    sc = make_scene(scene_name)
    scene = ol.OracleScene(kind, sc, ol.DEFINE_UB)
    cam = make_camera(sc.camera, width, height)
    cores = os.cpu_count() or 1
    times, calls = [], 0
    for i in range(warmup + steps):
        prm = make_params(total_spp, depth, rr=RR, rng_key=1234 + i, spp_begin=i * spp, spp_end=(i + 1) * spp)
        _, info = scene.render(cam, prm, rng, nthreads=cores)
        if i >= warmup:
            times.append(info.seconds)
            calls += info.world_hit_calls
        threads = info.threads
    sec = float(np.mean(times))
    how = "thread-local rand(), corrected static-init order" if variant == "fair" else "glibc rand() (global lock), CMake link order (AABB::empty defect): as shipped"
    return spp / sec, {"kind": "port" if kind == "orc" else "reference", "cores": int(threads), "seconds_per_step": sec,
                       "mrays_s": calls / sum(times) / 1e6,
                       "sample": f"{spp} spp of the {width}x{height} depth-{depth} frame per step ({steps} step(s)), "
                                 f"Camera::Render with {threads} threads, {how}"}


def cpu_rayquery_leg(sc, o, d, tmax_any, prefix):
    """Hittable::Hit of the reference (BVHNode::Hit, BVH.cpp:51-61, over its own median-split tree) on every k-th ray of the
    batch, all host cores. Returns {closest_mrays_s, any_mrays_s, rays, build_s, kind}."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as ol
    kind = "ref" if ol.available("ref") else "orc"
    t0 = time.time()
    # the compiled reference reports prim ids and barycentrics only through tagged materials / texcoords (oracle_api.h)
    O = ol.OracleScene(kind, sc, ol.PER_TRIANGLE_MATERIAL | ol.UV_TAG)
    build = time.time() - t0
    cores = os.cpu_count() or 1
    stride = max(1, len(o) // prefix)  # every stride-th ray: a prefix of a camera batch is all sky
    o, d, tmax_any = np.ascontiguousarray(o[::stride][:prefix]), np.ascontiguousarray(d[::stride][:prefix]), np.ascontiguousarray(tmax_any[::stride][:prefix])
    t0 = time.time()
    want = O.trace_closest(o, d, 1e-4, np.inf, nthreads=cores)
    tc = time.time() - t0
    t0 = time.time()
    O.trace_closest(o, d, 1e-3, tmax_any, nthreads=cores)  # visibility as the reference asks it: closest hit on a bounded interval
    ta = time.time() - t0
    O.close()
    return {"kind": "reference" if kind == "ref" else "port", "cores": cores, "rays": int(len(o)), "stride": stride, "bvh_build_s": build,
            "closest_mrays_s": len(o) / tc / 1e6, "any_mrays_s": len(o) / ta / 1e6}, want


def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    t0 = time.time()
    v, info = cpu_reference_leg(a.scene, a.width, a.height, a.depth, a.total_spp, a.cpu_spp or 2, steps=max(1, a.steps), warmup=min(a.warmup, 1))
    line = {
        "impl": "reference", "metric": "spp_per_s", "value": v, "unit": "spp/s", "n_gpus": a.gpus, "steps": a.steps,
        "warmup": min(a.warmup, 1), "ms_per_step": info["seconds_per_step"] * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": v / PUBLISHED_SPP_S, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_name(a.scene, a.width, a.height, a.depth, a.spp_per_step, a.total_spp), "reference_step": info["sample"]},
        "cpu_baseline": {"value": v, "unit": "spp/s", "cores": info["cores"], "kind": info["kind"], "sample": info["sample"],
                         "mrays_s": info["mrays_s"]},
        "e2e": {"value": v, "unit": "spp/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "wall_s": time.time() - t0,
    }
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------------------------------
# what ncu said about this workload (committed summaries; a profiler cannot run inside a timed bench)
# ---------------------------------------------------------------------------------------------------------------------
def ncu_evidence(scene, accel, S, width, height, depth, ms_per_step_per_gpu, peak):
    """dram traffic of one step and the limiter ncu names, from profiles/r2_traffic.json (tools/traffic_from_launches.py over
    the committed launch list) and profiles/r2_limiters.json (tools/limiters_from_ncu.py over the committed --set full
    summaries). Returns the extra keys of the roofline object."""
    out = {"traffic": None}
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "r2_traffic.json")))
        e = tj.get(scene)
        w = WORKLOADS[scene]
        if e and e["accel"] == accel and e["spp_per_step"] == S and (width, height, depth) == (w["width"], w["height"], w["depth"]):
            out["traffic"] = e["traversal_dram_bytes_per_step"]
            out["traffic_note"] = (f"dram read+write of k_extend + k_shadow (+ k_primary) summed over the {e['traversal_launches_per_step']:.0f} launches of one step "
                                   f"(ncu, {e['source']}); all kernels of the step: {e['step_dram_bytes']:.4g} B")
            # measured DRAM bytes of the WHOLE step over the live step time: how much of the HBM peak the step really uses
            out["dram_frac"] = e["step_dram_bytes"] / (ms_per_step_per_gpu * 1e-3) / 1e9 / peak
    except Exception:
        pass
    try:
        lj = json.load(open(os.path.join(ROOT, "profiles", "r2_limiters.json"))).get(scene)
        if lj:
            out.update({"limiter": lj["limiter"], "l1tex_pct": lj["l1tex_pct"], "simd_lanes": lj["simd_lanes"], "issue_active_pct": lj["issue_active_pct"],
                        "dram_pct_ncu": lj["dram_pct"], "limiter_source": lj["source"],
                        # `bound` above names the denominator (HBM peak, GB/s); this names what ncu shows the kernel waiting on
                        "bound_ncu": "l1tex" if lj["l1tex_pct"] >= 70.0 else "latency"})
    except Exception:
        pass
    return out


def roofline_object(st0, rays0, stage, n_stage, stage_rays, scene, accel, S, width, height, depth, ms_per_step_per_gpu):
    peak, peak_source = peaks()
    alg_bytes0 = st0.node_visits * st0.node_bytes + st0.tri_tests * st0.tri_bytes
    bytes_per_ray = alg_bytes0 / max(1, rays0)
    ext_ms, shd_ms, sdw_ms = stage
    trav_ms = ext_ms + sdw_ms
    achieved = stage_rays * bytes_per_ray / (trav_ms * 1e-3) / 1e9 if trav_ms > 0 else 0.0
    r = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
         "frac_name": "algorithmic_demand_vs_hbm: cache-oblivious bytes the walk asks for (every node / triangle visit counted) over the HBM peak; "
                      "the scene is L1/L2-resident, so this is NOT DRAM utilisation (see dram_frac) and NOT the limiter (see limiter)",
         "algorithmic_demand_vs_hbm": achieved / peak, "peak_source": peak_source,
         "kernel": "k_extend + k_shadow (BVH traversal)", "bytes_per_ray": bytes_per_ray,
         "nodes_per_ray": st0.node_visits / max(1, rays0), "tris_per_ray": st0.tri_tests / max(1, rays0),
         "node_bytes": st0.node_bytes, "tri_bytes": st0.tri_bytes,
         "extend_ms": ext_ms / n_stage, "shade_ms": shd_ms / n_stage, "shadow_ms": sdw_ms / n_stage}
    r.update(ncu_evidence(scene, accel, S, width, height, depth, ms_per_step_per_gpu, peak))
    return r


# ---------------------------------------------------------------------------------------------------------------------
# one render workload through the C ABI's whole-box handle (rank 0, all N GPUs)
# ---------------------------------------------------------------------------------------------------------------------
def measure_box(box, sc, scene_name, a, n_gpus, width, height, depth, total_spp, steps, warmup, want_roofline=True):
    """prt_multi_render on all GPUs of `box`. A step = S spp per GPU (S * n_gpus samples split across the GPUs, one
    reduction on the first GPU): `value` without a host image, `e2e` with one (pinned), both wall-clock around synchronous calls."""
    import torch
    from pooraytracer_b200 import _capi as capi, make_camera, make_params
    cam = make_camera(sc.camera, width, height)
    S = a.spp_per_step
    npix = width * height

    def prm(step, flags=0):
        return make_params(total_spp, depth, rr=RR, rng_key=KEY, spp_begin=step * S * n_gpus, spp_end=(step + 1) * S * n_gpus, flags=flags)

    def call(p, host):
        rc = box.lib.prt_multi_render(box.handle, C.byref(cam), C.byref(p), C.c_void_p(host.data_ptr()) if host is not None else None)
        assert rc == 0, box.lib.prt_last_error()

    out = {}
    st0 = rays0 = None
    if want_roofline:
        call(prm(0, capi.RENDER_COUNTERS), None)
        st0, _ = box.stats()
        rays0 = st0.extend_rays + st0.shadow_rays
    for w in range(warmup):
        call(prm(1 + w), None)
    rays = launches = 0
    t0 = time.perf_counter()
    for k in range(steps):
        call(prm(1 + warmup + k), None)
    dt = time.perf_counter() - t0
    for k in range(steps):  # the counters of the timed steps, read afterwards (prt_multi_get_stats synchronises)
        pass
    st, reduce_ms = box.stats()
    rays_per_step = st.extend_rays + st.shadow_rays  # of the last step; steps differ by < 0.1 % (same frame, other samples)
    launches = st.kernel_launches * steps
    value = steps * S * n_gpus / dt
    out.update({"value": value, "unit": "spp/s", "ms_per_step": dt / steps * 1e3, "mrays_s": rays_per_step * steps / dt / 1e6,
                "mpaths_s": steps * S * n_gpus * npix / dt / 1e6, "rays_per_path": rays_per_step / (S * n_gpus * npix),
                "reduce_ms": reduce_ms, "gpu_launches": int(launches), "triangles": sc.n_triangles,
                "workload": workload_name(scene_name, width, height, depth, S, total_spp)})
    # end to end: host image out (pinned), D2H inside the timed region
    host = torch.empty(npix * 3, dtype=torch.float64).pin_memory()
    call(prm(0), host)
    e2e_steps = max(1, min(steps, 3))
    t0 = time.perf_counter()
    for k in range(e2e_steps):
        call(prm(1 + warmup + k), host)
    dte = time.perf_counter() - t0
    # pageable destination, as the INTEGRATION.md stub passes it (std::vector storage)
    pageable = np.empty(npix * 3, dtype=np.float64)
    t0 = time.perf_counter()
    rc = box.lib.prt_multi_render(box.handle, C.byref(cam), C.byref(prm(1 + warmup)), pageable.ctypes.data_as(C.c_void_p))
    assert rc == 0, box.lib.prt_last_error()
    dtp = time.perf_counter() - t0
    # a fixed sample range, whatever N: the checksum every N must reproduce
    fixed = make_params(total_spp, depth, rr=RR, rng_key=KEY, spp_begin=0, spp_end=64)
    call(fixed, host)
    out["e2e"] = {"value": e2e_steps * S * n_gpus / dte, "unit": "spp/s", "h2d_bytes_per_step": C.sizeof(cam) + C.sizeof(fixed),
                  "d2h_bytes_per_step": npix * 3 * 8, "steps": e2e_steps, "pageable_host_image_value": S * n_gpus / dtp,
                  "image_checksum": float(host.sum().item()), "checksum_of": "samples [0, 64) of the frame, reduced on the first GPU: the same for every N up to the order of the final FP64 additions",
                  "call": f"prt_multi_render on {n_gpus} GPU(s): sample split, reduction of the frames to the first GPU ({box.reduce_kind}) and the device-to-host copy inside the timed region"}
    if want_roofline:
        n_stage = min(2, steps)
        ext = shd = sdw = 0.0
        stage_rays = 0
        for k in range(n_stage):
            call(prm(1 + warmup + k, capi.RENDER_STAGE_TIMING), None)
            s_, _ = box.stats()  # stage times of the first GPU, rays of all: scale the rays to one GPU
            ext += s_.extend_ms; shd += s_.shade_ms; sdw += s_.shadow_ms
            stage_rays += (s_.extend_rays + s_.shadow_rays) / n_gpus
        out["roofline"] = roofline_object(st0, rays0, (ext, shd, sdw), n_stage, stage_rays, scene_name, a.accel, S, width, height, depth, dt / steps * 1e3)
    return out


def measure_rayquery(a, tris_m, mesh, n_log2=24, reps=3, cpu_prefix=0):
    """configs[4] on one GPU: device-resident 2^n_log2 rays, CUDA-event time inside the library. Returns a list of rows."""
    import torch
    from pooraytracer_b200 import Context, Scene, _capi as capi, scenes
    n = 1 << n_log2
    if mesh == "terrain":
        sc = scenes.terrain(1000, 500) if tris_m == 1 else scenes.terrain(2500, 2000)
    else:
        sc = scenes.soup(tris_m * 1_000_000)
    t0 = time.time()
    scene = Scene(Context(0), sc, accel=None)
    t_create = time.time() - t0
    t0 = time.time()
    scene.commit(getattr(capi, ACCELS[a.accel]), max_leaf=a.max_leaf)
    t_commit = time.time() - t0
    peak, _ = peaks()
    diag = float(np.linalg.norm(sc.vertices.reshape(-1, 3).max(0) - sc.vertices.reshape(-1, 3).min(0)))
    tmax_any_h = scenes.hash_uniform(n, 5, 1) * diag
    rows = []
    for kind in ("random", "coherent"):
        if kind == "random":
            o, d = scenes.random_rays(sc, n)
        else:
            side = 1 << (n_log2 // 2)
            o, d = scenes.camera_rays(sc.camera, side, n // side)
        do, dd = torch.from_numpy(o).cuda(), torch.from_numpy(d).cuda()
        tmin = torch.full((n,), 1e-4, dtype=torch.float64, device="cuda")
        tmax = torch.full((n,), float("inf"), dtype=torch.float64, device="cuda")
        prim = torch.empty(n, dtype=torch.int32, device="cuda")
        t = torch.zeros(n, dtype=torch.float64, device="cuda")  # a miss leaves t / alpha / beta untouched (0, as the CPU answer has them)
        al, be = torch.zeros_like(t), torch.zeros_like(t)
        hit = torch.empty(n, dtype=torch.uint8, device="cuda")
        tmax_any = torch.from_numpy(tmax_any_h).cuda()
        tmin_any = torch.full((n,), 1e-3, dtype=torch.float64, device="cuda")
        torch.cuda.synchronize()
        cpu = want = None
        if cpu_prefix:
            cpu, want = cpu_rayquery_leg(sc, o, d, tmax_any_h, cpu_prefix)
        for query in ("closest", "any"):
            def run():
                if query == "closest":
                    rc = scene.lib.prt_trace_closest_device(scene.handle, n, do.data_ptr(), dd.data_ptr(), tmin.data_ptr(), tmax.data_ptr(),
                                                            prim.data_ptr(), t.data_ptr(), al.data_ptr(), be.data_ptr(), None)
                else:
                    rc = scene.lib.prt_trace_any_device(scene.handle, n, do.data_ptr(), dd.data_ptr(), tmin_any.data_ptr(), tmax_any.data_ptr(),
                                                        hit.data_ptr(), None)
                assert rc == 0, scene.lib.prt_last_error()
                return scene.stats()
            scene.set_counters(True)
            st = run()
            nodes, tris = st.node_visits, st.tri_tests
            scene.set_counters(False)
            run()
            ms = min(run().gpu_ms for _ in range(reps))
            alg = nodes * st.node_bytes + tris * st.tri_bytes
            row = {"case": f"{mesh}-{sc.n_triangles} {kind} {query}", "rays": n, "mrays_s": n / ms / 1e3, "ms": ms,
                   "nodes_per_ray": nodes / n, "tris_per_ray": tris / n, "bytes_per_ray": alg / n,
                   "algorithmic_demand_vs_hbm": alg / (ms * 1e-3) / 1e9 / peak,
                   "hit_fraction": float((prim >= 0).float().mean().item()) if query == "closest" else float(hit.float().mean().item()),
                   "scene_create_s": t_create, "scene_commit_s": t_commit}
            if cpu:
                row["cpu_baseline"] = {"value": cpu[f"{query}_mrays_s"], "unit": "Mrays/s", "cores": cpu["cores"], "kind": cpu["kind"],
                                       "sample": f"every {cpu['stride']}th ray of the batch ({cpu['rays']} rays) through BVHNode::Hit on the reference's own tree ({cpu['bvh_build_s']:.1f} s to build)"
                                                 + ("" if query == "closest" else "; visibility asked the reference's way: closest hit on [1e-3, tmax]")}
                if query == "closest":  # gate 1 on this very batch: the sampled rays bit-exact against the CPU answer
                    m, st_ = cpu["rays"], cpu["stride"]
                    got = tuple(x.cpu().numpy()[::st_][:m] for x in (prim, t, al, be))
                    same = [bool(np.array_equal(x, y)) for x, y in zip(got, want)]
                    row["oracle_check"] = {"rays": m, "bit_exact": all(same), "hits": int((want[0] >= 0).sum())}
                    if not all(same):
                        row["oracle_check"]["differs_in"] = [nm for nm, ok in zip(("prim", "t", "alpha", "beta"), same) if not ok]
            rows.append(row)
        del do, dd, tmin, tmax, prim, t, al, be, hit, tmax_any, tmin_any
    scene.close()
    torch.cuda.empty_cache()
    return rows


def measure_jobs(a, n_gpus):
    """Whole jobs, wall clock, strong-scaled over the N GPUs of the box."""
    from pooraytracer_b200 import MultiScene, _capi as capi, api, make_camera, make_params, scenes
    jobs = {}
    # (1) the reference's published job through the drop-in for main.cpp: scene FILES in, PNG + HDR out
    exe = os.path.join(ROOT, "pooraytracer_b200", "prt_main")
    if os.path.exists(exe):
        with tempfile.TemporaryDirectory(prefix="prt_job_") as d:
            sc = make_scene("cornell-box")
            scenes.write_scene_files(sc, d, "cornell-box")
            t0 = time.perf_counter()
            r = subprocess.run([exe, d, "cornell-box", "--spp", "3000", "--depth", "100", "--gpus", str(n_gpus), "--seed", str(KEY), "--out", os.path.join(d, "out")],
                               capture_output=True, text=True, timeout=1200)
            wall = time.perf_counter() - t0
            line = [l for l in r.stdout.splitlines() if l.startswith("job:")]
            jobs["cornell-box spp=3000 depth=100 1024x1024"] = {
                "wall_s": wall, "ok": r.returncode == 0 and os.path.exists(os.path.join(d, "out.png")), "n_gpus": n_gpus,
                "what": "prt_main (drop-in for main.cpp): load OBJ/MTL/XML -> build + upload on every GPU -> render, samples split -> frames reduced on the first GPU -> resolve -> PNG + HDR; process start to exit",
                "phases": line[0][5:].strip() if line else r.stderr[-300:], "reference_published_s": 20382.22}
    # (2) north_star: "bathroom2 at 500 spp / depth 50 rendered in seconds on 8 GPUs"
    sc = make_scene("bathroom2")
    w = WORKLOADS["bathroom2"]
    cam = make_camera(sc.camera, w["width"], w["height"])
    t0 = time.perf_counter()
    box = MultiScene(list(range(n_gpus)), sc, accel=getattr(capi, ACCELS[a.accel]), max_leaf=a.max_leaf)
    t_build = time.perf_counter() - t0
    t1 = time.perf_counter()
    box.render(cam, make_params(500, 50, rr=RR, rng_key=KEY), out=None if False else np.empty((w["height"], w["width"], 3)))
    t_render = time.perf_counter() - t1
    t2 = time.perf_counter()
    rgb8, f32 = box.resolve(w["width"], w["height"])
    with tempfile.TemporaryDirectory(prefix="prt_job_") as d:
        api.write_png(os.path.join(d, "bathroom2.png"), rgb8)
        api.write_hdr(os.path.join(d, "bathroom2.hdr"), f32)
    t_out = time.perf_counter() - t2
    st, reduce_ms = box.stats()
    jobs["bathroom2 spp=500 depth=50 1280x720"] = {
        "wall_s": t_build + t_render + t_out, "n_gpus": n_gpus, "triangles": sc.n_triangles,
        "what": "arrays -> prt_multi_create (ingest + SAH build + upload on every GPU) -> prt_multi_render (samples split, frames reduced on the first GPU, host image) -> prt_multi_resolve -> PNG + HDR",
        "build_upload_s": t_build, "render_reduce_s": t_render, "resolve_write_s": t_out, "reduce_ms": reduce_ms,
        "spp_per_s": 500 / t_render, "mrays_s": (st.extend_rays + st.shadow_rays) / t_render / 1e6}
    box.close()
    return jobs


def run_b200(a):
    import torch
    import torch.distributed as dist
    from pooraytracer_b200 import Context, MultiScene, Scene, _capi as capi, make_camera, make_params
    from pooraytracer_b200.sharding import sample_range

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        import datetime
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local), timeout=datetime.timedelta(minutes=40))
        # the ranks that only wait (while rank 0 drives every GPU through the C ABI's whole-box handle) must wait on the HOST:
        # an NCCL barrier is a kernel spinning on their GPU, and rank 0's kernels on that GPU — another process, another
        # CUDA context — would be time-sliced against it
        host_group = dist.new_group(backend="gloo", timeout=datetime.timedelta(minutes=40))
    accel = getattr(capi, ACCELS[a.accel])
    sc = make_scene(a.scene)
    ctx = Context(local)
    t0 = time.time()
    scene = Scene(ctx, sc, accel=None)
    t_create = time.time() - t0
    t0 = time.time()
    scene.commit(accel, max_leaf=a.max_leaf)
    t_commit = time.time() - t0
    cam = make_camera(sc.camera, a.width, a.height)
    npix = a.width * a.height
    S = a.spp_per_step
    stream = torch.cuda.Stream()
    acc = torch.zeros(npix * 3, dtype=torch.float64, device="cuda")

    def params(step, flags=0):
        # global sample range of this step, sharded over ranks (pooraytracer_b200/sharding.py)
        b, e = sample_range(step, S, rank, world)
        return make_params(a.total_spp, a.depth, rr=RR, rng_key=KEY, spp_begin=b, spp_end=e, flags=flags)

    def one_step(step, flags=0):
        # every step renders into zeroed accumulators: after the reduce rank 0 holds exactly this step's image
        # (a persistent accumulator would re-add the other ranks' earlier partial sums on every later step)
        with torch.cuda.stream(stream):
            acc.zero_()
        scene.render_device(cam, params(step, flags), acc.data_ptr(), stream.cuda_stream)
        if world > 1:
            with torch.cuda.stream(stream):
                dist.reduce(acc, dst=0, op=dist.ReduceOp.SUM)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # -- instrumented run of step 0: algorithmic bytes per ray (deterministic per step) ---------------
    one_step(0, capi.RENDER_COUNTERS)
    torch.cuda.synchronize()
    st0 = scene.stats()
    rays0 = st0.extend_rays + st0.shadow_rays

    # -- warm-up, then the timed region (device time, max over ranks) ---------------------------------
    for w in range(a.warmup):
        one_step(1 + w)
    barrier()
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    rays = launches = 0
    barrier()
    ev0.record(stream)
    for k in range(a.steps):
        one_step(1 + a.warmup + k)
        s_ = scene.stats()  # waits for this step (events on the launching stream)
        rays += s_.extend_rays + s_.shadow_rays
        launches += s_.kernel_launches + 1  # + the zero fill of the accumulators
    ev1.record(stream)
    barrier()
    ms = torch.tensor([ev0.elapsed_time(ev1)], dtype=torch.float64, device="cuda")
    tot = torch.tensor([float(rays), float(launches)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    ms_total = float(ms.item())
    clk = clocks.stop() if rank == 0 else None
    value = a.steps * S * world / (ms_total * 1e-3)

    # -- stage times of the dominant kernels (separate steps: the stage events wait per pass) ---------
    n_stage = min(2, a.steps)
    ext_ms = shd_ms = sdw_ms = 0.0
    stage_rays = 0
    for k in range(n_stage):
        one_step(1 + a.warmup + k, capi.RENDER_STAGE_TIMING)
        s_ = scene.stats()
        ext_ms += s_.extend_ms; shd_ms += s_.shade_ms; sdw_ms += s_.shadow_ms
        stage_rays += s_.extend_rays + s_.shadow_rays
    barrier()
    line = None
    if rank == 0:
        line = {
            "metric": "spp_per_s", "value": value, "unit": "spp/s", "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": ms_total / a.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": value / PUBLISHED_SPP_S if a.scene == "cornell-box" else None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(a.scene, a.width, a.height, a.depth, S, a.total_spp), "triangles": sc.n_triangles, "spp_per_step_per_gpu": S,
                       "paths_per_step": S * npix * world, "accel": ACCEL_TEXT[a.accel] + ", FP32 outward boxes) + reference BVH for boundary hits",
                       "primary_hit_reuse": "bounce 0 is traced once per pixel (Camera::GetRay has no jitter: all samples of a pixel share the primary ray), so GPU rays per path fall by ~1 and Mrays/s counts only rays actually traced",
                       "l2": f"no flush: per-step path state (192 B x {S * npix / 1e6:.1f} M paths = {192 * S * npix / 1e9:.1f} GB) exceeds the 126 MB L2; the scene itself is L2-resident by nature",
                       "scene_create_s": t_create, "scene_commit_s": t_commit, "parallelism": f"spp-sharded x{world}, 1 NCCL reduce per step"},
            "mrays_s": float(tot[0].item()) / (ms_total * 1e-3) / 1e6,
            "mpaths_s": a.steps * S * npix * world / (ms_total * 1e-3) / 1e6,
            "rays_per_path": float(tot[0].item()) / (a.steps * S * npix * world),
            "gpu_launches": int(tot[1].item()),
            "roofline": roofline_object(st0, rays0, (ext_ms, shd_ms, sdw_ms), n_stage, stage_rays, a.scene, a.accel, S, a.width, a.height, a.depth, ms_total / a.steps),
            "clocks": clk,
        }

    # -- everything below runs on rank 0 alone (the other ranks wait at the last barrier): the C ABI's whole-box handle
    #    drives all N GPUs from one process, as a C / C++ caller would -------------------------------------------------
    scene.close()
    del acc
    torch.cuda.empty_cache()
    barrier()
    torch.cuda.synchronize()
    if rank == 0:
        devices = list(range(world))
        box = MultiScene(devices, sc, accel=accel, max_leaf=a.max_leaf)
        head = measure_box(box, sc, a.scene, a, world, a.width, a.height, a.depth, a.total_spp, a.steps, max(1, a.warmup), want_roofline=False)
        box.close()
        line["e2e"] = head["e2e"]
        line["e2e"]["device_resident_value_same_call"] = head["value"]
        if not a.headline_only:
            line["scenes"] = {}
            for name in WORKLOADS:
                if name == a.scene:
                    continue
                w = WORKLOADS[name]
                sc2 = make_scene(name)
                box = MultiScene(devices, sc2, accel=accel, max_leaf=a.max_leaf)
                m = measure_box(box, sc2, name, a, world, w["width"], w["height"], w["depth"], w["total_spp"], max(2, min(a.steps, 3)), 1)
                box.close()
                if not a.no_cpu_baseline and world == 1:
                    spp = {"veach-mis": 16, "bathroom2": 4}[name]
                    v, info = cpu_reference_leg(name, w["width"], w["height"], w["depth"], w["total_spp"], spp)
                    m["cpu_baseline"] = {"value": v, "unit": "spp/s", "cores": info["cores"], "kind": info["kind"], "sample": info["sample"], "mrays_s": info["mrays_s"]}
                line["scenes"][name] = m
            try:
                line["job"] = measure_jobs(a, world)
            except Exception as e:  # a job that cannot run is reported, not fatal to the headline
                line["job"] = {"error": repr(e)}
            if world == 1:
                rq = []
                cpu_prefix = 0 if a.no_cpu_baseline else (1 << 20)
                for tris_m in (1,) if a.no_rayquery_10m else (1, 10):
                    for mesh in ("terrain", "soup"):
                        rq += measure_rayquery(a, tris_m, mesh, cpu_prefix=cpu_prefix if tris_m == 1 else 0)
                try:  # what ncu names for the one ray-query case that was captured (tools/ncu_captures.sh: terrain-10M, first closest launch)
                    lj = json.load(open(os.path.join(ROOT, "profiles", "r2_limiters.json"))).get("rayquery-terrain-10M")
                    for row in rq:
                        if lj and row["case"].startswith("terrain-10000000 random closest"):
                            row["ncu"] = {k: lj[k] for k in ("limiter", "l1tex_pct", "dram_pct", "l2_pct", "issue_active_pct", "simd_lanes", "source")}
                except Exception:
                    pass
                line["rayquery"] = rq
                line["rayquery_note"] = ("2^24 rays, device-resident, one GPU, CUDA-event time of the kernel; cpu_baseline / oracle_check on the 1 M-triangle meshes "
                                         "(the reference's BVH over 10 M triangles takes minutes to build: not inside the bench); 2/4/8-GPU figures: tools/rayquery_multi.sh")
        if not a.no_cpu_baseline and world == 1:
            v, info = cpu_reference_leg(a.scene, a.width, a.height, a.depth, a.total_spp, a.cpu_spp or 12)
            line["cpu_baseline"] = {"value": v, "unit": "spp/s", "cores": info["cores"], "kind": info["kind"], "sample": info["sample"],
                                    "mrays_s": info["mrays_s"]}
            vs, infos = cpu_reference_leg(a.scene, a.width, a.height, a.depth, a.total_spp, 1, variant="shipped")  # ~30 s of 16 cores: glibc rand() serialises the threads
            if vs is not None:
                line["cpu_baseline"]["as_shipped"] = {"value": vs, "unit": "spp/s", "cores": infos["cores"], "sample": infos["sample"], "mrays_s": infos["mrays_s"]}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier(group=host_group)
        dist.destroy_process_group()


if __name__ == "__main__":
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)
