"""N>1 host logic on CPU: two gloo ranks.  (1) bench.py's counter-based generator gives every sharding the same
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


def test_two_rank_sharding_and_exact_histogram_reduce(tmp_path):
    port = _free_port()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    ok_shards, ok_hist = open(os.path.join(str(tmp_path), "result.txt")).read().split()
    assert ok_shards == "1", "shards of gen_block do not tile the global matrix"
    assert ok_hist == "1", "all-reduced shard histograms differ from the global histogram"
