"""CPU test of the N>1 path: 2 ranks over gloo, sample-sharded rendering + one reduce.
The renderer here is the CPU oracle (a checker standing in for the GPU, which this box lacks); what
is under test is the host-side sharding logic of pooraytracer_b200/sharding.py and the property it
relies on: shards of the counter-based sample space sum to the single-rank image."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, out_path):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import cases
    import oracle_lib as ol
    from pooraytracer_b200 import make_camera, make_params
    from pooraytracer_b200.sharding import reduce_image, sample_range, split_samples
    sc = cases.tiny()
    O = ol.OracleScene("orc", sc, ol.DEFINE_UB)
    cam = make_camera(sc.camera, 24, 16)
    total = 5
    # strong scaling: one render split over the ranks
    b, e = split_samples(total, rank, world)
    img, _ = O.render(cam, make_params(total, 6, rng_key=3, spp_begin=b, spp_end=e), ol.RNG_COUNTER, nthreads=1)
    acc = torch.from_numpy(img.copy())
    reduce_image(acc, dst=0)
    # weak scaling: two steps, every rank renders 2 samples per step into the same accumulator
    acc2 = torch.zeros_like(acc)
    for step in range(2):
        b, e = sample_range(step, 2, rank, world)
        part, _ = O.render(cam, make_params(8, 6, rng_key=3, spp_begin=b, spp_end=e), ol.RNG_COUNTER, nthreads=1)
        acc2 += torch.from_numpy(part)
    reduce_image(acc2, dst=0)
    if rank == 0:
        full, _ = O.render(cam, make_params(total, 6, rng_key=3), ol.RNG_COUNTER, nthreads=1)
        full2, _ = O.render(cam, make_params(8, 6, rng_key=3), ol.RNG_COUNTER, nthreads=1)
        np.savez(out_path, a=acc.numpy(), full=full, a2=acc2.numpy(), full2=full2)
    dist.barrier()
    dist.destroy_process_group()


def test_two_ranks_shard_and_reduce(built, tmp_path):
    out = str(tmp_path / "out.npz")
    port = 29500 + (os.getpid() % 2000)
    mp.spawn(_worker, args=(2, port, out), nprocs=2, join=True)
    r = np.load(out)
    assert np.allclose(r["a"], r["full"], rtol=1e-13, atol=1e-15)
    assert np.allclose(r["a2"], r["full2"], rtol=1e-13, atol=1e-15)
    assert r["full"].mean() > 0.01


def test_sample_ranges_partition():
    from pooraytracer_b200.sharding import sample_range, split_samples
    for world in (1, 2, 4, 8):
        for total in (16, 500, 3000, 12000, 7):
            got = [split_samples(total, r, world) for r in range(world)]
            assert got[0][0] == 0 and got[-1][1] == total
            assert all(got[i][1] == got[i + 1][0] for i in range(world - 1))
            sizes = [e - b for b, e in got]
            assert max(sizes) - min(sizes) <= 1
        seen = []
        for step in range(3):
            for r in range(world):
                seen.append(sample_range(step, 16, r, world))
        flat = sorted(seen)
        assert flat[0][0] == 0 and all(flat[i][1] == flat[i + 1][0] for i in range(len(flat) - 1))


def _ray_worker(rank, world, port, out_path):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import cases
    import oracle_lib as ol
    from pooraytracer_b200.sharding import ray_range
    sc = cases.tiny()
    o, d = cases.ray_batch(sc)
    n = (o.shape[0] // 2) * 2 - 1  # odd: the two ranges differ in size
    o, d = o[:n], d[:n]
    O = ol.OracleScene("orc", sc, 0)
    b, e = ray_range(n, rank, world)
    prim, t, al, be = O.trace_closest(o[b:e], d[b:e], 1e-4, np.inf, nthreads=1)
    # no collective on the data path: every rank hands back its own range (here: gathered for the check)
    mine = torch.full((n, 2), -7.0, dtype=torch.float64)
    mine[b:e, 0] = torch.from_numpy(prim.astype(np.float64))
    mine[b:e, 1] = torch.from_numpy(np.where(np.isfinite(t), t, -1.0))
    parts = [torch.empty_like(mine) for _ in range(world)] if rank == 0 else None
    dist.gather(mine, parts, dst=0)
    if rank == 0:
        full = O.trace_closest(o, d, 1e-4, np.inf, nthreads=1)
        got = torch.full((n, 2), -7.0, dtype=torch.float64)
        for r, p in enumerate(parts):
            rb, re_ = ray_range(n, r, world)
            got[rb:re_] = p[rb:re_]
        np.savez(out_path, got=got.numpy(), prim=full[0], t=np.where(np.isfinite(full[1]), full[1], -1.0))
    dist.barrier()
    dist.destroy_process_group()


def test_two_ranks_shard_a_ray_batch(built, tmp_path):
    """Ray queries over N ranks: contiguous, balanced ranges whose union is the batch, each ray's answer
    independent of the range it was traced in (bit-exact against the single-rank trace)."""
    from pooraytracer_b200.sharding import ray_range
    for world in (1, 2, 3, 8):
        got = [ray_range(1 << 24, r, world) for r in range(world)]
        assert got[0][0] == 0 and got[-1][1] == 1 << 24 and all(got[i][1] == got[i + 1][0] for i in range(world - 1))
    out = str(tmp_path / "rays.npz")
    port = 31500 + (os.getpid() % 2000)
    mp.spawn(_ray_worker, args=(2, port, out), nprocs=2, join=True)
    r = np.load(out)
    assert np.array_equal(r["got"][:, 0], r["prim"].astype(np.float64))
    assert np.array_equal(r["got"][:, 1], r["t"])
    assert (r["prim"] >= 0).any()
