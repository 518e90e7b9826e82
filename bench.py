#!/usr/bin/env python
"""bench.py — the reference's headline render on B200: cornell-box 1024x1024, depth 100, NEE, RR 0.8
(BASELINE.json configs[1]; the reference published 20382 s for 3000 spp = 0.1472 spp/s on an
unstated CPU).

A STEP renders a slice of `--spp-per-step` samples per pixel of that frame (default 64 = one
wavefront pass of 67 M paths, 12.9 GB of path records) through the full depth-100 pipeline; 3000 / 64 steps make the whole
image, and throughput is intensive, so the slice measures the named configuration, not a smaller
one. The scene files of the reference are absent (example-scenes-cg24/ is empty), so the scene is the
procedural cornell-box of pooraytracer_b200/scenes.py in the reference's material vocabulary.

  value      whole-job spp/s (samples per pixel of the 1024x1024 frame per second), scene and
             accumulators resident in HBM, CUDA events on the launching stream, max over ranks
  e2e        the same through prt_render() with a HOST image buffer (the call a user of the C ABI
             makes): camera + parameters in, W*H*3 float64 out, D2H inside the timed region
  roofline   extend + shadow kernels: algorithmic bytes (nodes * 64 B + triangles * 128 B actually
             fetched, counted by an instrumented run of the same step) / their summed durations
  cpu_baseline / --impl reference
             the unmodified reference's Camera::Render (oracle/_ref, built from /root/reference)
             on the host cores, thread-local rand(), on a bounded sample of the same frame

Multi-GPU (torchrun, one rank per GPU): the scene is replicated, every rank renders its own
sample range of every step, and one NCCL reduce per step sums the float64 accumulators on rank 0.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

PUBLISHED_SPP_S = 3000.0 / 20382.22  # cornell-box 1024x1024 spp 3000 depth 100
WIDTH = HEIGHT = 1024
DEPTH = 100
RR = 0.8
TOTAL_SPP = 3000


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--spp-per-step", type=int, default=64)
    ap.add_argument("--width", type=int, default=WIDTH)
    ap.add_argument("--height", type=int, default=HEIGHT)
    ap.add_argument("--depth", type=int, default=DEPTH)
    ap.add_argument("--scene", default="cornell-box", choices=["cornell-box", "veach-mis", "bathroom2"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--accel", default=os.environ.get("PRT_BENCH_ACCEL", "bvh4"), choices=["bvh2", "bvh4", "lbvh", "bvh4q"])
    ap.add_argument("--cpu-spp", type=int, default=0,
                    help="spp of the bounded CPU sample (0 = 12 for the cpu_baseline of the b200 arm, about 13 s of 16 host cores on the default workload; 2 per step for --impl reference)")
    ap.add_argument("--max-leaf", type=int, default=0, help="triangles per BVH leaf (0 = the library's default, 4)")
    return ap.parse_args()


def make_scene(name):
    from pooraytracer_b200 import scenes
    if name == "cornell-box":
        return scenes.cornell_box(WIDTH, HEIGHT)
    if name == "veach-mis":
        return scenes.veach_mis()
    return scenes.bathroom_like()


def workload_name(a):
    return (f"{a.scene} {a.width}x{a.height} depth={a.depth} NEE RR={RR} (procedural stand-in for the absent "
            f"example-scenes-cg24 files); step = {a.spp_per_step} spp of the {TOTAL_SPP}-spp render")


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


def cpu_reference_leg(a, spp, steps=1, warmup=0):
    """The reference's own Camera::Render (unmodified sources, oracle/_ref) on the host cores, with
    the thread-local rand() interposition (glibc's rand() takes a global lock: negative thread
    scaling). Falls back to the oracle port when oracle/_ref is absent. Returns (spp/s, info)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as ol
    from pooraytracer_b200 import make_camera, make_params
    kind = "ref" if ol.available("ref") else "orc"
    if kind == "orc" and not ol.available("orc"):
        import __graft_entry__ as g
        g.build()
    sc = make_scene(a.scene)
    scene = ol.OracleScene(kind, sc, ol.DEFINE_UB)
    cam = make_camera(sc.camera, a.width, a.height)
    cores = os.cpu_count() or 1
    times, calls = [], 0
    for i in range(warmup + steps):
        prm = make_params(TOTAL_SPP, a.depth, rr=RR, rng_key=1234 + i, spp_begin=i * spp, spp_end=(i + 1) * spp)
        _, info = scene.render(cam, prm, ol.RNG_THREAD_LOCAL, nthreads=cores)
        if i >= warmup:
            times.append(info.seconds)
            calls += info.world_hit_calls
        threads = info.threads
    sec = float(np.mean(times))
    return spp / sec, {"kind": "reference" if kind == "ref" else "port", "cores": int(threads), "seconds_per_step": sec,
                       "mrays_s": calls / sum(times) / 1e6,
                       "sample": f"{spp} spp of the {a.width}x{a.height} depth-{a.depth} frame per step ({steps} step(s)), "
                                 f"Camera::Render with {threads} threads, thread-local rand(), corrected static-init order"}


def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    t0 = time.time()
    v, info = cpu_reference_leg(a, a.cpu_spp or 2, steps=max(1, a.steps), warmup=min(a.warmup, 1))
    line = {
        "impl": "reference", "metric": "spp_per_s", "value": v, "unit": "spp/s", "n_gpus": a.gpus, "steps": a.steps,
        "warmup": min(a.warmup, 1), "ms_per_step": info["seconds_per_step"] * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": v / PUBLISHED_SPP_S, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_name(a), "reference_step": info["sample"]},
        "cpu_baseline": {"value": v, "unit": "spp/s", "cores": info["cores"], "kind": info["kind"], "sample": info["sample"],
                         "mrays_s": info["mrays_s"]},
        "e2e": {"value": v, "unit": "spp/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "wall_s": time.time() - t0,
    }
    print(json.dumps(line), flush=True)


def run_b200(a):
    import torch
    import torch.distributed as dist
    from pooraytracer_b200 import Context, Scene, _capi as capi, make_camera, make_params
    from pooraytracer_b200.sharding import sample_range

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    sc = make_scene(a.scene)
    ctx = Context(local)
    t0 = time.time()
    scene = Scene(ctx, sc, accel=None)
    t_create = time.time() - t0
    t0 = time.time()
    scene.commit({"bvh4": capi.ACCEL_SAH_BVH4, "bvh2": capi.ACCEL_SAH_BVH2, "lbvh": capi.ACCEL_LBVH4_DEVICE, "bvh4q": capi.ACCEL_SAH_BVH4Q}[a.accel], max_leaf=a.max_leaf)
    t_commit = time.time() - t0
    cam = make_camera(sc.camera, a.width, a.height)
    npix = a.width * a.height
    S = a.spp_per_step
    stream = torch.cuda.Stream()
    acc = torch.zeros(npix * 3, dtype=torch.float64, device="cuda")

    def params(step, flags=0):
        # global sample range of this step, sharded over ranks (pooraytracer_b200/sharding.py)
        b, e = sample_range(step, S, rank, world)
        return make_params(TOTAL_SPP, a.depth, rr=RR, rng_key=20251104, spp_begin=b, spp_end=e, flags=flags)

    def one_step(step, flags=0):
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
    st = scene.stats()
    rays0 = st.extend_rays + st.shadow_rays
    alg_bytes0 = st.node_visits * st.node_bytes + st.tri_tests * st.tri_bytes
    bytes_per_ray = alg_bytes0 / max(1, rays0)
    nodes_per_ray, tris_per_ray = st.node_visits / max(1, rays0), st.tri_tests / max(1, rays0)

    # -- warm-up, then the timed region (device time, max over ranks) ---------------------------------
    for w in range(a.warmup):
        one_step(1 + w)
    barrier()
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    rays = launches = 0
    ext_ms = shd_ms = sdw_ms = 0.0
    barrier()
    ev0.record(stream)
    for k in range(a.steps):
        one_step(1 + a.warmup + k)
        s_ = scene.stats()  # waits for this step (events on the launching stream)
        rays += s_.extend_rays + s_.shadow_rays
        launches += s_.kernel_launches
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
    stage_rays = 0
    for k in range(n_stage):
        one_step(1 + a.warmup + k, capi.RENDER_STAGE_TIMING)
        s_ = scene.stats()
        ext_ms += s_.extend_ms; shd_ms += s_.shade_ms; sdw_ms += s_.shadow_ms
        stage_rays += s_.extend_rays + s_.shadow_rays
    trav_ms = ext_ms + sdw_ms
    achieved = stage_rays * bytes_per_ray / (trav_ms * 1e-3) / 1e9 if trav_ms > 0 else 0.0
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))

    # -- end to end through prt_render with a pinned HOST image ---------------------------------------
    host_img = torch.empty(npix * 3, dtype=torch.float64).pin_memory()
    e2e_steps = max(1, min(a.steps, 3))
    p = params(0)  # one untimed call: the library's host-image staging is allocated on first use
    assert scene.lib.prt_render(scene.handle, C.byref(cam), C.byref(p), C.c_void_p(host_img.data_ptr())) == 0, scene.lib.prt_last_error()
    barrier()
    t0 = time.perf_counter()
    for k in range(e2e_steps):
        p = params(1 + a.warmup + k)
        rc = scene.lib.prt_render(scene.handle, C.byref(cam), C.byref(p), C.c_void_p(host_img.data_ptr()))
        assert rc == 0, scene.lib.prt_last_error()
    barrier()
    e2e_t = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(e2e_t, op=dist.ReduceOp.MAX)
    e2e_value = e2e_steps * S * world / float(e2e_t.item())
    checksum = float(host_img.sum().item())

    # DRAM traffic of the traversal kernels over one step, from the committed ncu pass of this very
    # workload (profiles/r1_traffic.json; a profiler cannot run inside a timed bench): same basis as
    # `achieved`, which is the step's algorithmic bytes over the step's traversal time
    traffic = traffic_note = None
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "r1_traffic.json")))
        if (tj["workload"] == a.scene and tj["accel"] == a.accel and tj["spp_per_step"] == S and a.width == WIDTH and a.height == HEIGHT
                and a.depth == DEPTH):
            traffic = tj["traversal_dram_bytes_per_step"]
            traffic_note = (f"dram read+write of k_extend + k_shadow summed over the {tj['traversal_launches_per_step']:.0f} launches of one step "
                            f"(ncu, profiles/r1_traffic.json); algorithmic bytes of the same step: {bytes_per_ray * stage_rays / max(1, n_stage):.4g} "
                            "(cache-oblivious: the 2.5 MB scene is L1/L2-resident, DRAM carries the path records)")
    except Exception:
        pass

    if rank == 0:
        line = {
            "metric": "spp_per_s", "value": value, "unit": "spp/s", "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": ms_total / a.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": value / PUBLISHED_SPP_S if a.scene == "cornell-box" else None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(a), "triangles": sc.n_triangles, "spp_per_step_per_gpu": S,
                       "paths_per_step": S * npix * world, "accel": {"bvh4": "SAH BVH4 (128 B nodes", "bvh2": "SAH BVH2 (64 B nodes", "lbvh": "device-built LBVH4 (128 B nodes", "bvh4q": "SAH BVH4, 8-bit quantised child boxes (64 B nodes"}[a.accel] + ", FP32 outward boxes) + reference BVH for boundary hits",
                       "l2": f"no flush: per-step path state (192 B x {S * npix / 1e6:.1f} M paths = {192 * S * npix / 1e9:.1f} GB) exceeds the 126 MB L2; the scene itself is L2-resident by nature",
                       "scene_create_s": t_create, "scene_commit_s": t_commit, "parallelism": f"spp-sharded x{world}, 1 NCCL reduce per step"},
            "mrays_s": float(tot[0].item()) / (ms_total * 1e-3) / 1e6,
            "mpaths_s": a.steps * S * npix * world / (ms_total * 1e-3) / 1e6,
            "rays_per_path": float(tot[0].item()) / (a.steps * S * npix * world),
            "e2e": {"value": e2e_value, "unit": "spp/s", "h2d_bytes_per_step": C.sizeof(cam) + C.sizeof(params(0)),
                    "d2h_bytes_per_step": npix * 3 * 8, "steps": e2e_steps, "image_checksum": checksum},
            "gpu_launches": int(tot[1].item()),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "traffic_note": traffic_note,
                         "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "fallback 6650 (of fallback)",
                         "kernel": "k_extend + k_shadow (BVH traversal)", "bytes_per_ray": bytes_per_ray,
                         "nodes_per_ray": nodes_per_ray, "tris_per_ray": tris_per_ray, "node_bytes": st.node_bytes, "tri_bytes": st.tri_bytes,
                         "extend_ms": ext_ms / n_stage, "shade_ms": shd_ms / n_stage, "shadow_ms": sdw_ms / n_stage},
            "clocks": clk,
        }
        if not a.no_cpu_baseline and world == 1:
            v, info = cpu_reference_leg(a, a.cpu_spp or 12)
            line["cpu_baseline"] = {"value": v, "unit": "spp/s", "cores": info["cores"], "kind": info["kind"], "sample": info["sample"],
                                    "mrays_s": info["mrays_s"]}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)
