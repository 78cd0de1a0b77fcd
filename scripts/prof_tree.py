"""Profiling driver (not a bench): the bench's Higgs-shape workload, a few boosting iterations, per-iteration wall time
and the event-timed stage split.  Run with GBL_USE_GRAPH=0 under ncu so that every kernel of a tree is a launch of its own.

  python scripts/prof_tree.py [--rows N] [--iters K] [--quantized 0|1] [--workload higgs|reg] [--prof 0|1]
"""
import argparse
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from lambdagap_b200 import capi, engine  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--rows", type=int, default=10_500_000)
ap.add_argument("--iters", type=int, default=4)
ap.add_argument("--quantized", type=int, default=0)
ap.add_argument("--workload", default="higgs")
ap.add_argument("--prof", type=int, default=0)
ap.add_argument("--extra", default="")
args = ap.parse_args()

cfg = dict(bench.HIGGS if args.workload == "higgs" else bench.REG100M)
X, y = bench.gen_block(cfg, 0, args.rows)
ds = engine.Dataset(X, y, params="max_bin=255 verbose=-1")
params = cfg["params"] + " device_type=cuda verbose=-1 " + args.extra
if args.quantized:
    params += " use_quantized_grad=true stochastic_rounding=false"
bst = engine.Booster(ds, params)
lrn = bst.learner()
if args.prof:
    lrn.set_profiling(True)
    lrn.profile(reset=True)
for it in range(args.iters):
    capi.synchronize(0)
    t0 = time.perf_counter()
    bst.update()
    capi.synchronize(0)
    dt = time.perf_counter() - t0
    line = "iter %d: %.3f ms" % (it, 1e3 * dt)
    if args.prof:
        p = lrn.profile(reset=True)
        line += "  hist %.3f find %.3f part %.3f other %.3f | waves %d expansions %d hist_rows %.2fM part_rows %.2fM hist_launches %d" % (
            p["hist_ms"], p["find_ms"], p["partition_ms"], p["other_ms"], p["waves"], p["expansions"], p["hist_rows"] / 1e6,
            p["partition_rows"] / 1e6, p["hist_launches"])
    print(line, flush=True)
