"""Run under torchrun (N ranks, one GPU each): the data-parallel learner over row shards must grow exactly the
trees the single-GPU learner grows on the full data (exact integer histograms => order / shard independent).
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 scripts/check_multi_gpu.py
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from lambdagap_b200 import engine
import bench

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
cfg = dict(bench.HIGGS)
rows = 400_000
Xg, yg = bench.gen_block(cfg, 0, rows)
sample = Xg[:100_000]
ok = True
for quant in (1, 0):
    params = "objective=binary num_leaves=63 learning_rate=0.1 min_data_in_leaf=20 verbose=-1"
    if quant:
        params += " use_quantized_grad=true stochastic_rounding=false"
    single = None
    if rank == 0:  # single-GPU run on the full data, before the network exists
        ds = engine.Dataset(Xg, yg, params="max_bin=255 verbose=-1 gpu_device_id=%d" % local, bin_sample=sample, num_total_rows=rows)
        b = engine.Booster(ds, params + " device_type=cuda gpu_device_id=%d" % local)
        for _ in range(4):
            b.update()
        single = b.model_to_string().split("feature_importances")[0]
        b.free()
        ds.free()
    uid = [engine.network_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(uid, src=0)
    engine.network_init_nccl(uid[0], world, rank, local)
    per = rows // world
    Xs, ys = Xg[rank * per:(rank + 1) * per], yg[rank * per:(rank + 1) * per]
    ds = engine.Dataset(Xs, ys, params="max_bin=255 verbose=-1 gpu_device_id=%d" % local, bin_sample=sample, num_total_rows=rows)
    b = engine.Booster(ds, params + " device_type=cuda tree_learner=data gpu_device_id=%d" % local)
    for _ in range(4):
        b.update()
    multi = b.model_to_string().split("feature_importances")[0]
    b.free()
    ds.free()
    engine.network_free()
    if rank == 0:
        same = single.split("parameters:")[0] == multi.split("parameters:")[0]
        print("quant=%d  %d-GPU model == 1-GPU model: %s" % (quant, world, same), flush=True)
        if not same:
            a, c = single.split("\n"), multi.split("\n")
            for x, y_ in zip(a, c):
                if x != y_ and not x.startswith('tree_sizes'):
                    print("  1gpu:", x[:160])
                    print("  Ngpu:", y_[:160])
                    break
        ok = ok and same
    dist.barrier()
dist.destroy_process_group()
if rank == 0:
    print("MULTI_GPU_CHECK", "PASS" if ok else "FAIL", flush=True)
    sys.exit(0 if ok else 1)
