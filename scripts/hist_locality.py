"""What would physically regrouped row records buy the histogram kernel?  Measured, not estimated (asked for in round 1's
review): the kernel over the SAME number of rows, once as one contiguous block of the record array (what a node's rows
would be right after a regroup), once as an ascending random subset of density 1/2, 1/4, ... (what a node's rows are
`depth` levels below the last regroup, or below the root without one).  Beside it: the cost of moving the records of all
rows once (a device gather of 32-byte records = the regroup pass itself).  Higgs shape, quantized and float.

  python scripts/hist_locality.py [--rows N]
Prints one JSON object; not a bench.
"""
import argparse
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from lambdagap_b200 import capi, engine  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--rows", type=int, default=10_500_000)
args = ap.parse_args()

cfg = dict(bench.HIGGS)
X, y = bench.gen_block(cfg, 0, args.rows)
n = len(y)
dev = torch.device("cuda", 0)
out = {"rows": n}
for quant in (1, 0):
    ds = engine.Dataset(X, y, params="max_bin=255 verbose=-1")
    params = cfg["params"] + " device_type=cuda verbose=-1"
    if quant:
        params += " use_quantized_grad=true stochastic_rounding=false"
    bst = engine.Booster(ds, params)
    bst.update()  # one tree: the learner holds gradients and records
    lrn = bst.learner()
    words = lrn.histogram_words()
    hist = torch.zeros(words, dtype=torch.int64, device=dev)
    res = {}
    g = torch.Generator(device=dev)
    g.manual_seed(1)
    for level in range(0, 7):
        m = n >> level  # rows of a node `level` levels below the root of the regroup
        if m < 100_000:
            break
        if level == 0:
            sets = {"contiguous": torch.arange(m, dtype=torch.int32, device=dev)}
        else:
            perm = torch.randperm(n, generator=g, device=dev)[:m]
            sets = {"contiguous": torch.arange(m, dtype=torch.int32, device=dev),
                    "scattered": torch.sort(perm).values.to(torch.int32)}
        row = {"rows": m}
        for name, idx in sets.items():
            for _ in range(3):
                hist.zero_()
                lrn.construct_histogram(idx.data_ptr(), m, hist.data_ptr())
            torch.cuda.synchronize()
            ts = []
            for _ in range(5):
                hist.zero_()
                # flush L2 between repetitions (the 126 MB L2 would keep a small node's records)
                flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev).fill_(1)
                del flush
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                capi.synchronize(0)
                torch.cuda.synchronize()
                a.record()
                lrn.construct_histogram(idx.data_ptr(), m, hist.data_ptr())
                capi.synchronize(0)
                b.record()
                torch.cuda.synchronize()
                ts.append(a.elapsed_time(b) * 1e3)
            row[name + "_us"] = round(float(np.median(ts)), 1)
            row[name + "_ps_per_row"] = round(float(np.median(ts)) * 1e6 / m, 1)
        res["density_1/%d" % (1 << level)] = row
    # the regroup pass itself: gather + write of one 32-byte (quantized) or 64-byte (float) record per row
    rec_bytes = 32 if quant else 64
    rec = torch.zeros((n, rec_bytes // 4), dtype=torch.int32, device=dev)
    perm = torch.randperm(n, generator=g, device=dev)
    dst = torch.empty_like(rec)
    for _ in range(2):
        torch.index_select(rec, 0, perm, out=dst)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(5):
        torch.index_select(rec, 0, perm, out=dst)
    b.record()
    torch.cuda.synchronize()
    res["regroup_all_rows_us"] = round(a.elapsed_time(b) * 1e3 / 5, 1)
    res["record_bytes"] = rec_bytes
    out["quantized" if quant else "float"] = res
    del rec, dst, hist
    bst.free()
print(json.dumps(out))
