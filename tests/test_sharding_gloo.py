"""N>1 host logic on CPU: gloo ranks.  (1) bench.py's counter-based generator gives every sharding the same
global matrix; (2) per-shard histograms all-reduced with SUM equal the global histogram exactly in the integer
representation the data-parallel learner exchanges (fixed-point int64), which is why N-GPU trees == 1-GPU trees."""
import os
import socket
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import bench
    cfg = dict(bench.HIGGS)
    rows = 300_000
    local = rows // world
    X, y = bench.gen_block(cfg, rank * local, local)
    # (1) shards of the global matrix
    gathered = [torch.zeros(local, cfg["features"]) for _ in range(world)] if rank == 0 else None
    dist.gather(torch.from_numpy(X), gathered, dst=0)
    # (2) exact integer histograms: bins by a shared rule, gradients in 2^k fixed point
    bins = np.clip(((X[:, 0] + 4.0) * 32).astype(np.int64), 0, 255)
    g = (y - 0.5).astype(np.float32) * np.float32(0.731)
    gfix = np.rint(g.astype(np.float64) * 2.0 ** 30).astype(np.int64)
    hist = torch.from_numpy(np.bincount(bins, weights=None, minlength=256).astype(np.int64) * 0)
    np.add.at(hist.numpy(), bins, gfix)
    cnt = torch.tensor([local], dtype=torch.int64)
    dist.all_reduce(hist, op=dist.ReduceOp.SUM)
    dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
    if rank == 0:
        Xg, yg = bench.gen_block(cfg, 0, rows)
        ok_shards = bool(np.array_equal(torch.cat(gathered).numpy(), Xg))
        bg = np.clip(((Xg[:, 0] + 4.0) * 32).astype(np.int64), 0, 255)
        gg = np.rint(((yg - 0.5).astype(np.float32) * np.float32(0.731)).astype(np.float64) * 2.0 ** 30).astype(np.int64)
        ref = np.zeros(256, dtype=np.int64)
        np.add.at(ref, bg, gg)
        ok_hist = bool(np.array_equal(hist.numpy(), ref)) and int(cnt.item()) == rows
        with open(os.path.join(out_dir, "result.txt"), "w") as fh:
            fh.write("%d %d" % (ok_shards, ok_hist))
    dist.barrier()
    dist.destroy_process_group()


def _exchange_worker(rank, world, port, out_dir):
    """The peer-memory exchange of exchange.cu (xchg_reduce_kernel) with gloo standing in for NVLink: the slot range
    is cut into 16-byte units, rank k owns the k-th chunk of ceil(units / R) units, reads that chunk from every rank's
    pool, adds, and writes the sum into every rank's pool -- in place, shares disjoint."""
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rng = np.random.default_rng(100 + rank)
    slot_stride, slots = 2 * 8192 + 8, 5  # one tile of float-mode words + tail, five fresh slots of a wave
    words = slot_stride * slots
    pool = torch.from_numpy(rng.integers(-2 ** 40, 2 ** 40, size=words, dtype=np.int64))
    expect = pool.clone()
    dist.all_reduce(expect, op=dist.ReduceOp.SUM)
    units = words // 2
    chunk = (units + world - 1) // world
    u0, u1 = rank * chunk, min(units, (rank + 1) * chunk)
    # "load my share from all pools": every rank publishes its pool, I read only [u0, u1)
    pools = [torch.zeros(words, dtype=torch.int64) for _ in range(world)]
    dist.all_gather(pools, pool)
    mine = sum(p[2 * u0:2 * u1] for p in pools)
    # "store the sum into all pools": every rank receives every owner's share and places it at the owner's range
    shares = [torch.zeros(2 * chunk, dtype=torch.int64) for _ in range(world)]
    padded = torch.zeros(2 * chunk, dtype=torch.int64)
    padded[:2 * (u1 - u0)] = mine
    dist.all_gather(shares, padded)
    for k in range(world):
        a, b = k * chunk, min(units, (k + 1) * chunk)
        pool[2 * a:2 * b] = shares[k][:2 * (b - a)]
    ok = bool(torch.equal(pool, expect))
    covered = sum(min(units, (k + 1) * chunk) - min(units, k * chunk) for k in range(world)) == units
    with open(os.path.join(out_dir, "xchg_%d.txt" % rank), "w") as fh:
        fh.write("%d %d" % (ok, covered))
    dist.barrier()
    dist.destroy_process_group()


def test_two_shot_exchange_ownership_covers_and_sums(tmp_path):
    for world in (2, 3):  # 3: the chunks do not divide the range evenly
        port = _free_port()
        mp.spawn(_exchange_worker, args=(world, port, str(tmp_path)), nprocs=world, join=True)
        for r in range(world):
            ok, covered = open(os.path.join(str(tmp_path), "xchg_%d.txt" % r)).read().split()
            assert ok == "1" and covered == "1", "world %d rank %d" % (world, r)


def test_two_rank_sharding_and_exact_histogram_reduce(tmp_path):
    port = _free_port()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    ok_shards, ok_hist = open(os.path.join(str(tmp_path), "result.txt")).read().split()
    assert ok_shards == "1", "shards of gen_block do not tile the global matrix"
    assert ok_hist == "1", "all-reduced shard histograms differ from the global histogram"
