#!/usr/bin/env python
"""Ray-query microbench (BASELINE.json configs[4]): 2^24 random / coherent rays against the procedural
`terrain` and `soup` meshes (1 M and 10 M triangles), closest-hit and any-hit, device-resident inputs,
CUDA-event timing inside the library (prt_stats.gpu_ms). One JSON line per case.

  python tools/rayquery_bench.py --tris 1 --rays 24 [--mesh terrain|soup] [--check N]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/rayquery_bench.py ...

Under torchrun the scene is replicated, rank r traces the r-th contiguous range of the batch
(pooraytracer_b200/sharding.py:ray_range), there is no collective on the data path, and rank 0 prints the
aggregate rate = all rays / the slowest rank's device time, with the per-GPU rates beside it.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tris", type=int, default=1, help="millions of triangles (1 or 10)")
    ap.add_argument("--rays", type=int, default=24, help="log2 of the ray count")
    ap.add_argument("--mesh", default="terrain")
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--check", type=int, default=0, help="compare the first N rays with the CPU oracle")
    ap.add_argument("--max-leaf", type=int, default=0)
    ap.add_argument("--split", default="contiguous", choices=["contiguous", "blocks"],
                    help="under torchrun: rank r traces the r-th contiguous range of the batch (SURVEY 8e) or every N-th block of 65536 rays (balances a camera batch whose top rows are sky)")
    ap.add_argument("--accel", default=os.environ.get("PRT_BENCH_ACCEL", "auto"), choices=["auto", "bvh2", "bvh4", "lbvh", "bvh4q"])
    a = ap.parse_args()
    import torch
    from pooraytracer_b200 import Context, Scene, _capi as capi, scenes

    rank, local, world = int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    if world > 1:
        import torch.distributed as dist
        from pooraytracer_b200.sharding import ray_range
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    n_all = 1 << a.rays
    lo_, hi_ = ray_range(n_all, rank, world) if world > 1 else (0, n_all)
    if world > 1 and a.split == "blocks":
        B = 65536
        mine = np.concatenate([np.arange(b, min(b + B, n_all)) for b in range(rank * B, n_all, world * B)])
    else:
        mine = np.arange(lo_, hi_)
    n = len(mine)
    if a.mesh == "terrain":
        sc = scenes.terrain(1000, 500) if a.tris == 1 else scenes.terrain(2500, 2000)
    else:
        sc = scenes.soup(a.tris * 1_000_000)
    t0 = time.time()
    ctx = Context(local)
    scene = Scene(ctx, sc, accel=None)
    t_create = time.time() - t0
    t0 = time.time()
    import ctypes as C
    o_ = capi.PrtBuildOpts()
    o_.accel = {"bvh4": capi.ACCEL_SAH_BVH4, "bvh2": capi.ACCEL_SAH_BVH2, "lbvh": capi.ACCEL_LBVH4_DEVICE, "bvh4q": capi.ACCEL_SAH_BVH4Q, "auto": capi.ACCEL_AUTO}[a.accel]
    o_.reserved[0] = a.max_leaf
    rc = scene.lib.prt_scene_commit(scene.handle, C.byref(o_))
    assert rc == 0, scene.lib.prt_last_error()
    t_commit = time.time() - t0
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
    peak = float(peaks.get("hbm_gbs", 6650.0))
    diag = float(np.linalg.norm(sc.vertices.reshape(-1, 3).max(0) - sc.vertices.reshape(-1, 3).min(0)))
    for kind in ("random", "coherent"):
        if kind == "random":
            o, d = scenes.random_rays(sc, n_all)
        else:
            side = 1 << (a.rays // 2)
            o, d = scenes.camera_rays(sc.camera, side, n_all // side)
        o, d = np.ascontiguousarray(o[mine]), np.ascontiguousarray(d[mine])
        do, dd = torch.from_numpy(o).cuda(), torch.from_numpy(d).cuda()
        tmin = torch.full((n,), 1e-4, dtype=torch.float64, device="cuda")
        tmax = torch.full((n,), float("inf"), dtype=torch.float64, device="cuda")
        prim = torch.empty(n, dtype=torch.int32, device="cuda")
        t = torch.zeros(n, dtype=torch.float64, device="cuda")  # a miss leaves t / alpha / beta untouched
        al, be = torch.zeros_like(t), torch.zeros_like(t)
        hit = torch.empty(n, dtype=torch.uint8, device="cuda")
        tmax_any = torch.from_numpy(np.ascontiguousarray((scenes.hash_uniform(n_all, 5, 1) * diag)[mine])).cuda()
        tmin_any = torch.full((n,), 1e-3, dtype=torch.float64, device="cuda")
        torch.cuda.synchronize()
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
            ms = min(run().gpu_ms for _ in range(a.reps))
            hits = float((prim >= 0).float().sum().item()) if query == "closest" else float(hit.float().sum().item())
            per_gpu = None
            if world > 1:
                # aggregate = all rays / the slowest rank; counters and hits summed over the ranks
                v = torch.tensor([ms, float(nodes), float(tris), hits, float(n)], dtype=torch.float64, device="cuda")
                g = [torch.empty_like(v) for _ in range(world)]
                dist.all_gather(g, v)
                g = torch.stack(g).cpu().numpy()
                per_gpu = [float(r[4] / r[0] / 1e3) for r in g]
                ms, nodes, tris, hits, n_tot = float(g[:, 0].max()), int(g[:, 1].sum()), int(g[:, 2].sum()), float(g[:, 3].sum()), int(g[:, 4].sum())
                if rank != 0:
                    continue
            else:
                n_tot = n
            alg = nodes * st.node_bytes + tris * st.tri_bytes
            line = {"case": f"{a.mesh}-{sc.n_triangles} {kind} {query}", "n_gpus": world, "rays": n_tot, "mrays_s": n_tot / ms / 1e3, "ms": ms,
                    "nodes_per_ray": nodes / n_tot, "tris_per_ray": tris / n_tot, "bytes_per_ray": alg / n_tot,
                    # cache-oblivious demand (every node / triangle fetch counted as if it came from HBM) against the HBM peak:
                    # above 1 on coherent batches, where L1 / L2 serve the fetches; the limiter ncu names is in profiles/r2_limiters.json
                    "algorithmic_gb_s": alg / (ms * 1e-3) / 1e9, "hbm_peak_gb_s": peak * world,
                    "algorithmic_demand_vs_hbm": alg / (ms * 1e-3) / 1e9 / (peak * world), "accel": a.accel, "node_bytes": st.node_bytes, "split": a.split if world > 1 else None,
                    "hit_fraction": hits / n_tot,
                    "scene_create_s": t_create, "scene_commit_s": t_commit, "n_nodes": st.n_nodes, "build_ms": st.build_ms}
            if a.check and query == "closest" and kind == "random":
                sys.path.insert(0, os.path.join(ROOT, "tests"))
                import oracle_lib as ol
                O = ol.OracleScene("orc", sc, 0)
                m = a.check
                want = O.trace_closest(o[:m], d[:m], 1e-4, np.inf, nthreads=os.cpu_count() or 4)
                got = (prim[:m].cpu().numpy(), t[:m].cpu().numpy(), al[:m].cpu().numpy(), be[:m].cpu().numpy())
                line["oracle_check"] = {"rays": m, "bit_exact": bool(all(np.array_equal(x, y) for x, y in zip(got, want)))}
            if per_gpu:
                line["mrays_s_per_gpu"] = per_gpu
            print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
