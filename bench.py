#!/usr/bin/env python
"""bench.py -- boosting iterations/sec of the histogram-GBDT hot path on B200 (BASELINE.json metric).

  python bench.py --gpus N --steps K --warmup W [--impl reference] [--workload higgs|reg100m|epsilon|mslr]

A "step" is one boosting iteration (gradients -> one leaf-wise tree of num_leaves=255 -> score update).
  N=1  headline workload = BASELINE.json configs[1]: Higgs-shape synthetic binary classification, 10.5M rows x 28 dense
       features, max_bin=255, num_leaves=255 (float gradients).  Inputs are larger than L2 (294 MB of bins + 672 MB of
       row records + 84 MB gradients per step), so no L2 flush is needed between steps.
  N>1  the same per-GPU shape sharded by rows over N ranks (data-parallel learner): weak scaling, `value` =
       iterations/sec of the whole N x 10.5M-row job.  The line also carries `parity` (a 400K-row side run: the N-GPU
       trees equal the 1-GPU trees and every rank holds the same model).
  every N: `reg100m` = BASELINE.json configs[4], 100M x 100 regression with the data-parallel learner at N GPUs
       (strong scaling: the 100M rows are divided over the ranks), generated on the device.
  N=1 also: `quantized_mode` (the bit-exact mode on the headline shape), `other_shapes` (configs[3] Epsilon shape,
       configs[2] MSLR shape).
Prints ONE JSON line on rank 0.  `value` = device-resident loop timed with CUDA events on the learner's stream;
`e2e` = the same iteration driven through the C API with HOST gradient / hessian buffers in and the score vector
out (LGBM_BoosterUpdateOneIterCustom + LGBM_BoosterGetTrainScoreRaw), copies inside the timed region.
`--impl reference` times the UNMODIFIED reference CPU learner (oracle/_ref/lib_lightgbm.so, built from
/root/reference by oracle/Makefile.ref) on the host cores on the same workload.
"""
import argparse
import hashlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "boosting_iters_per_sec"
UNIT = "iters/s"

HIGGS = dict(name="higgs_shape_binary_10.5Mx28", rows=10_500_000, features=28, objective="binary", seed=42,
             params="objective=binary num_leaves=255 learning_rate=0.1 min_data_in_leaf=20 max_bin=255")
REG100M = dict(name="reg100m_shape_regression_100Mx100", rows=100_000_000, features=100, objective="regression",
               seed=45, params="objective=regression num_leaves=255 learning_rate=0.1 min_data_in_leaf=20 max_bin=255")
EPSILON = dict(name="epsilon_shape_binary_400Kx2000_quantized", rows=400_000, features=2000, objective="binary", seed=44,
               params="objective=binary num_leaves=255 learning_rate=0.1 min_data_in_leaf=20 max_bin=255 "
                      "use_quantized_grad=true num_grad_quant_bins=4 stochastic_rounding=false")
MSLR = dict(name="mslr_shape_lambdarank_3.77Mx136", queries=31_531, features=136, objective="lambdarank", seed=43,
            params="objective=lambdarank lambdarank_target=ndcg lambdarank_truncation_level=30 num_leaves=255 "
                   "learning_rate=0.1 min_data_in_leaf=20 max_bin=255")
SM_COUNT = 148


def gen_block(cfg, start, rows):
    """Rows [start, start+rows) of the synthetic matrix; counter-based per 1M-row block so that any sharding of
    the same global matrix yields the same rows."""
    f = cfg["features"]
    X = np.empty((rows, f), dtype=np.float32)
    y = np.empty(rows, dtype=np.float32)
    blk = 1_000_000
    wrng = np.random.default_rng(cfg["seed"])
    w = wrng.normal(size=f)
    pos = 0
    b = start // blk
    while pos < rows:
        b0 = b * blk
        rng = np.random.default_rng([cfg["seed"], b])
        xb = rng.normal(size=(blk, f)).astype(np.float32)
        if f >= 4:  # a few heavy-tailed columns -> unequal bins
            xb[:, 1] = np.exp(xb[:, 1])
            xb[:, 3] = np.exp(0.5 * xb[:, 3])
        noise = rng.normal(size=blk)
        lo = max(start, b0) - b0
        hi = min(start + rows, b0 + blk) - b0
        take = hi - lo
        X[pos:pos + take] = xb[lo:hi]
        s = xb[lo:hi].astype(np.float64) @ w / np.sqrt(f) + noise[lo:hi]
        if cfg["objective"] == "binary":
            y[pos:pos + take] = (s > 0.3).astype(np.float32)
        else:
            y[pos:pos + take] = (s + 0.1 * np.sin(3 * xb[lo:hi, 0])).astype(np.float32)
        pos += take
        b += 1
    return X, y


def gen_device(cfg, start, rows, device):
    """The same idea on the GPU (torch's counter-based generator, one seed per 1M-row block, so any sharding of the
    global matrix yields the same rows): returns (X float32 [rows, f] on `device`, y float32 on the host).  Used for
    the shapes whose host generation would take minutes (100M x 100)."""
    import torch
    f = cfg["features"]
    dev = torch.device("cuda", device)
    X = torch.empty((rows, f), dtype=torch.float32, device=dev)
    y = torch.empty(rows, dtype=torch.float32, device=dev)
    gw = torch.Generator(device=dev)
    gw.manual_seed(cfg["seed"])
    w = torch.randn(f, generator=gw, device=dev, dtype=torch.float32)
    blk = 1_000_000
    pos, b = 0, start // blk
    while pos < rows:
        b0 = b * blk
        g = torch.Generator(device=dev)
        g.manual_seed(cfg["seed"] * 1_000_003 + b)
        xb = torch.randn((blk, f), generator=g, device=dev, dtype=torch.float32)
        xb[:, 1::2] = torch.special.ndtr(xb[:, 1::2])  # odd columns ~ U(0,1): a normal / uniform mix
        noise = torch.randn(blk, generator=g, device=dev, dtype=torch.float32)
        lo = max(start, b0) - b0
        hi = min(start + rows, b0 + blk) - b0
        take = hi - lo
        X[pos:pos + take] = xb[lo:hi]
        s = xb[lo:hi] @ w / float(np.sqrt(f)) + noise[lo:hi]
        if cfg["objective"] == "binary":
            y[pos:pos + take] = (s > 0.3).to(torch.float32)
        else:
            y[pos:pos + take] = s + 0.1 * torch.sin(3 * xb[lo:hi, 0])
        pos += take
        b += 1
        del xb, noise
    return X, y.cpu().numpy()


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""

    def __init__(self, device):
        self.device = device
        self.proc = None
        self.lines = []   # (arrival time, csv line)
        self.t0 = None    # timed region (time.time()), set by mark()
        self.t1 = None
        self.load_end = None  # end of the same load continued after the timed region

    def mark(self, begin):
        if begin:
            self.t0 = time.time()
        else:
            self.t1 = time.time()

    def start(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.device), "--query-gpu=" + q,
                                          "--format=csv,noheader,nounits", "-lms", "20"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append((time.time(), line.strip()))

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        # samples that arrived inside the timed region; a region shorter than one nvidia-smi query (tens of ms)
        # can miss all of them, then the warm-up steps right before it (same load, same sampler) stand in
        inside = [ln for (t, ln) in self.lines if self.t0 is not None and self.t1 is not None and self.t0 <= t <= self.t1 + 0.05]
        window = "timed region"
        if not inside:
            end = self.load_end if self.load_end is not None else float("inf")
            inside = [ln for (t, ln) in self.lines if self.t0 is not None and self.t0 <= t <= end + 0.05]
            window = "timed region + the same load continued (timed region shorter than one nvidia-smi sample)"
        if not inside:
            inside = [ln for (_, ln) in self.lines]
            window = "warm-up + timed region + the same load continued"
        for ln in inside:
            parts = [p.strip() for p in ln.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0]))
                mx.append(float(parts[1]))
            except ValueError:
                continue
            for name, val in zip(names, parts[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "window": window, "reasons": sorted(reasons)}


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as fh:
            return json.load(fh).get("hbm_gbs", 6650.0), "MEASURED_PEAKS.json hbm_gbs (of measured)"
    return 6650.0, "B200_PROFILING.md fallback 6.65 TB/s (of fallback)"


def ncu_traffic(workload, quantized, world):
    """DRAM bytes per launch of the histogram kernel (tree-averaged) from the committed ncu launch list of THIS
    workload and mode (profiles/hist_traffic.json, keyed "<workload>/<float|quant>"); anything else -> None."""
    if world != 1:
        return None, "no ncu capture of multi-rank runs (ncu is never wrapped around a multi-rank command)"
    p = os.path.join(ROOT, "profiles", "hist_traffic.json")
    if not os.path.exists(p):
        return None, "no capture committed"
    with open(p) as fh:
        table = json.load(fh)
    key = "%s/%s" % (workload, "quant" if quantized else "float")
    ent = table.get(key)
    if not isinstance(ent, dict):
        return None, "no capture of %s committed" % key
    return ent.get("dram_bytes_per_launch"), ent.get("source", "profiles/hist_traffic.json")


# ------------------------------------------------------------------------------------------------ reference arm
def run_reference(args, cfg, rows, rank, world):
    if rank != 0:
        return None
    from oracle import ref_lgbm as R
    if not R.available():
        return {"impl": "reference", "unavailable": "oracle/_ref/lib_lightgbm.so was not built (no /root/reference at build time)"}
    cores = os.cpu_count() or 1
    X, y = gen_block(cfg, 0, rows)
    t0 = time.time()
    ds = R.RefDataset(X, y, params="max_bin=255 verbose=-1 num_threads=%d" % cores)
    t_ds = time.time() - t0
    bst = R.RefBooster(ds, cfg["params"] + " device_type=cpu verbose=-1 num_threads=%d" % cores)
    for _ in range(args.warmup):
        bst.update()
    t0 = time.time()
    for _ in range(args.steps):
        bst.update()
    dt = time.time() - t0
    val = args.steps / dt
    sample = "full %d x %d matrix, reference CPU learner (device_type=cpu, %d threads), %d warm-up + %d timed iterations; dataset build %.1f s excluded" % (
        rows, cfg["features"], cores, args.warmup, args.steps, t_ds)
    return {"metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic", "impl": "reference",
            "config": {"workload": cfg["name"], "rows": rows, "features": cfg["features"], "params": cfg["params"]},
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "reference", "sample": sample},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}


def cpu_baseline_leg(cfg, rows, X, y, budget_s=25.0):
    """Reference CPU learner on the box's host cores, bounded to ~budget_s of iterations (rank 0, N=1 only)."""
    from oracle import ref_lgbm as R
    if not R.available():
        return None
    cores = os.cpu_count() or 1
    t0 = time.time()
    ds = R.RefDataset(X, y, params="max_bin=255 verbose=-1 num_threads=%d" % cores)
    t_ds = time.time() - t0
    bst = R.RefBooster(ds, cfg["params"] + " device_type=cpu verbose=-1 num_threads=%d" % cores)
    bst.update()  # warm-up (first iteration is excluded)
    iters, t0 = 0, time.time()
    while iters < 3 or (time.time() - t0 < budget_s and iters < 40):
        bst.update()
        iters += 1
    dt = time.time() - t0
    bst.free()
    ds.free()
    return {"value": iters / dt, "unit": UNIT, "cores": cores, "kind": "reference",
            "sample": "full %d x %d matrix, unmodified reference CPU learner (oracle/_ref), %d threads, 1 warm-up + %d "
                      "timed iterations in %.1f s; dataset build %.1f s excluded" % (rows, cfg["features"], cores, iters, dt, t_ds)}


# ------------------------------------------------------------------------------------------------ timing helpers
class Ctx:
    """Process-wide state of our arm: rank / world, torch.distributed handles, device."""

    def __init__(self, rank, world, device):
        self.rank, self.world, self.device = rank, world, device
        self.dist = None
        self.torch = None

    def barrier(self):
        from lambdagap_b200 import capi
        capi.synchronize(self.device)
        if self.world > 1:
            self.dist.barrier()
        capi.synchronize(self.device)

    def max_over_ranks(self, x):
        if self.world == 1:
            return x
        t = self.torch.tensor([x], dtype=self.torch.float64, device="cuda")
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())


def time_steps(ctx, bst, steps, warmup):
    """`steps` boosting iterations timed with CUDA events on the learner's stream after `warmup` untimed ones,
    bracketed by a barrier + synchronize on both sides; max over ranks.  Returns (ms total, launches)."""
    from lambdagap_b200 import capi
    lrn = bst.learner()
    for _ in range(warmup):
        bst.update()
    ctx.barrier()
    launches0 = capi.kernel_launch_count()
    lrn.timer_record(0)
    for _ in range(steps):
        bst.update()
    lrn.timer_record(1)
    ms = lrn.timer_elapsed_ms(0, 1)
    ctx.barrier()
    return ctx.max_over_ranks(ms), capi.kernel_launch_count() - launches0


def profile_steps(bst, steps):
    """The same steps with CUDA-event stage timing (host-driven waves instead of the CUDA graph)."""
    lrn = bst.learner()
    lrn.set_profiling(True)
    lrn.profile(reset=True)
    for _ in range(steps):
        bst.update()
    prof = lrn.profile(reset=True)
    lrn.set_profiling(False)
    return prof


def hist_roofline(prof, features, quantized, const_hessian, ms_per_step, sm_mhz, workload, world):
    """Roofline of the histogram kernel over the launches of the profiled trees.  Algorithmic bytes of a call on n rows:
    n * (F * 1 + gh + 4 [index list]) + histogram write-out; the shared-memory atomic bound is reported
    beside it: the kernel issues `atoms` warp-wide ATOMS per (row, 32-feature tile) at 2 cycles each per SM."""
    peak, peak_src = measured_peaks()
    gh_bytes = 2 if quantized else 8
    hist_entry = 8 if quantized else 16
    tiles = (features + 31) // 32
    launches = max(1, prof["hist_launches"])
    alg_bytes = prof["hist_rows"] * (features + gh_bytes + 4) + prof["hist_launches"] * (features * 256 * hist_entry)
    hist_s = prof["hist_ms"] / 1e3
    achieved = alg_bytes / hist_s / 1e9 if hist_s > 0 else 0.0
    atoms = 1 if quantized else (3 if const_hessian else 4)
    clk = (sm_mhz or 1965.0) * 1e6
    atomic_floor_s = prof["hist_rows"] * tiles * atoms * 2.0 / (SM_COUNT * clk)
    stage_total = prof["hist_ms"] + prof["find_ms"] + prof["partition_ms"] + prof["other_ms"] + prof["collective_ms"]
    trees = max(1, prof["trees"])
    traffic, traffic_src = ncu_traffic(workload, quantized, world)
    stage_ms_per_tree = stage_total / trees
    host_gap = ms_per_step - stage_ms_per_tree
    return {"bound": "hbm", "kernel": "hist_quant_kernel" if quantized else "hist_float_kernel",
            "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "peak_source": peak_src,
            "traffic": traffic, "traffic_source": traffic_src,
            "algorithmic_bytes_per_launch": alg_bytes / launches,
            "avg_launch_us": 1e3 * prof["hist_ms"] / launches,
            "launches_per_step": prof["hist_launches"] / trees,
            "shared_atomic_bound": {"atoms_per_row_tile": atoms, "cycles_per_atoms": 2,
                                    "floor_us_per_launch": 1e6 * atomic_floor_s / launches,
                                    "frac_of_floor": atomic_floor_s / hist_s if hist_s > 0 else None,
                                    "note": "ATOMS.ADD is the only native shared-memory atomic on sm_100a; "
                                            "frac_of_floor = atomic floor time / measured kernel time"},
            "stage_share": {k: (prof[k + "_ms"] / stage_total if stage_total > 0 else None)
                            for k in ("hist", "find", "partition", "other", "collective")},
            "stage_ms_per_tree": stage_ms_per_tree,
            # step time (graph replay, device events) minus the event-timed stage sum of the profiled pass: a host-side
            # stall between launches shows up here (round 1 shipped a 6 ms per-iteration host scan unnoticed)
            "host_gap_ms": host_gap, "host_bound": bool(host_gap > 0.15 * ms_per_step)}


# ------------------------------------------------------------------------------------------------ side legs
def parity_single_gpu(device, rows=400_000, iters=4):
    """Rank 0, BEFORE the network exists: 1-GPU models (float, quantized) of the parity side run."""
    from lambdagap_b200 import engine
    cfg = dict(HIGGS)
    Xg, yg = gen_block(cfg, 0, rows)
    sample = Xg[:100_000]
    out = {}
    for quant in (0, 1):
        params = "objective=binary num_leaves=63 learning_rate=0.1 min_data_in_leaf=20 verbose=-1"
        if quant:
            params += " use_quantized_grad=true stochastic_rounding=false"
        ds = engine.Dataset(Xg, yg, params="max_bin=255 verbose=-1 gpu_device_id=%d" % device, bin_sample=sample, num_total_rows=rows)
        b = engine.Booster(ds, params + " device_type=cuda gpu_device_id=%d" % device)
        for _ in range(iters):
            b.update()
        out[quant] = b.model_to_string().split("feature_importances")[0].split("parameters:")[0]
        b.free()
        ds.free()
    return out


def parity_multi_gpu(ctx, single, rows=400_000, iters=4):
    """All ranks, network up: the same side run on row shards.  equals_1gpu: rank 0's N-GPU model text equals its
    1-GPU model; ranks_agree: every rank holds the same model (hash all-gathered)."""
    from lambdagap_b200 import engine
    cfg = dict(HIGGS)
    Xg, yg = gen_block(cfg, 0, rows)
    sample = Xg[:100_000]
    per = rows // ctx.world
    lo = ctx.rank * per
    hi = rows if ctx.rank == ctx.world - 1 else lo + per
    res = {"rows": rows, "iterations": iters, "num_leaves": 63, "modes": {}}
    for quant in (0, 1):
        params = "objective=binary num_leaves=63 learning_rate=0.1 min_data_in_leaf=20 verbose=-1"
        if quant:
            params += " use_quantized_grad=true stochastic_rounding=false"
        ds = engine.Dataset(Xg[lo:hi], yg[lo:hi], params="max_bin=255 verbose=-1 gpu_device_id=%d" % ctx.device,
                            bin_sample=sample, num_total_rows=rows)
        b = engine.Booster(ds, params + " device_type=cuda tree_learner=data gpu_device_id=%d" % ctx.device)
        for _ in range(iters):
            b.update()
        multi = b.model_to_string().split("feature_importances")[0].split("parameters:")[0]
        b.free()
        ds.free()
        digest = hashlib.sha256(multi.encode()).hexdigest()
        digests = [None] * ctx.world
        ctx.dist.all_gather_object(digests, digest)
        mode = {"ranks_agree": len(set(digests)) == 1}
        if ctx.rank == 0:
            mode["equals_1gpu"] = single is not None and single[quant] == multi
        res["modes"]["quantized" if quant else "float"] = mode
    if ctx.rank == 0:
        res["ranks_agree"] = all(m["ranks_agree"] for m in res["modes"].values())
        res["equals_1gpu"] = all(m["equals_1gpu"] for m in res["modes"].values())
    return res


def reg100m_leg(ctx, steps, warmup, rows_total):
    """BASELINE.json configs[4]: 100M x 100 regression, rows divided over the ranks (strong scaling), matrix generated
    on the device block by block (the host would need minutes for 10^10 normals)."""
    import torch
    from lambdagap_b200 import capi, engine
    cfg = dict(REG100M)
    cfg["rows"] = rows_total
    per = rows_total // ctx.world
    start = ctx.rank * per
    local = rows_total - start if ctx.rank == ctx.world - 1 else per
    t0 = time.time()
    Xs, _ = gen_device(cfg, 0, 200_000, ctx.device)
    sample = Xs.cpu().numpy()
    del Xs
    X, y = gen_device(cfg, start, local, ctx.device)
    torch.cuda.synchronize()
    t_gen = time.time() - t0
    t0 = time.time()
    ds = engine.Dataset.from_pointer(X.data_ptr(), local, cfg["features"], label=y,
                                     params="max_bin=255 verbose=-1 gpu_device_id=%d" % ctx.device,
                                     bin_sample=sample, num_total_rows=rows_total)
    del X
    torch.cuda.empty_cache()  # the learner allocates with cudaMalloc: hand torch's cached blocks back first
    t_ds = time.time() - t0
    params = cfg["params"] + " device_type=cuda verbose=-1 gpu_device_id=%d" % ctx.device
    if ctx.world > 1:
        params += " tree_learner=data"
    bst = engine.Booster(ds, params)
    ms, _ = time_steps(ctx, bst, steps, warmup)
    prof = profile_steps(bst, steps)
    capi.synchronize(ctx.device)
    roof = hist_roofline(prof, cfg["features"], False, True, ms / steps, None, "reg100m", ctx.world)
    out = {"workload": cfg["name"], "scaling": "strong", "rows_total": rows_total, "rows_per_gpu": local,
           "features": cfg["features"], "params": cfg["params"] + (" tree_learner=data" if ctx.world > 1 else ""),
           "value": steps / (ms / 1e3), "unit": UNIT, "ms_per_step": ms / steps, "steps": steps, "warmup": warmup,
           "data": "synthetic, generated on the device (torch Philox, one seed per 1M-row block)",
           "datagen_s": round(t_gen, 2), "dataset_build_s": round(t_ds, 2),
           "roofline": {k: roof[k] for k in ("kernel", "achieved", "peak", "unit", "frac", "avg_launch_us", "launches_per_step",
                                             "algorithmic_bytes_per_launch", "shared_atomic_bound", "stage_share",
                                             "stage_ms_per_tree")},
           "schedule": {"waves_per_tree": prof["waves"] / max(1, prof["trees"]),
                        "expansions_per_tree": prof["expansions"] / max(1, prof["trees"])}}
    bst.free()
    ds.free()
    return out


def gen_epsilon(device):
    import torch
    cfg = EPSILON
    dev = torch.device("cuda", device)
    g = torch.Generator(device=dev)
    g.manual_seed(cfg["seed"])
    n, f = cfg["rows"], cfg["features"]
    X = torch.randn((n, f), generator=g, device=dev, dtype=torch.float32)
    X /= X.norm(dim=1, keepdim=True)
    w = torch.zeros(f, device=dev)
    sel = torch.randperm(f, generator=g, device=dev)[:50]
    w[sel] = torch.randn(50, generator=g, device=dev)
    y = ((X @ w + 0.01 * torch.randn(n, generator=g, device=dev)) > 0).to(torch.float32)
    return X, y.cpu().numpy(), None


def gen_mslr(device):
    """MSLR-WEB30K shape: 31 531 queries, clipped lognormal sizes (~120 documents on average, max 1 251), 136 features of
    which ~30 % are zero-inflated, labels 0..4 correlated with a linear score."""
    import torch
    cfg = MSLR
    rng = np.random.default_rng(cfg["seed"])
    sizes = np.clip(np.round(rng.lognormal(mean=np.log(95.0), sigma=0.7, size=cfg["queries"])), 1, 1251).astype(np.int32)
    n, f = int(sizes.sum()), cfg["features"]
    dev = torch.device("cuda", device)
    g = torch.Generator(device=dev)
    g.manual_seed(cfg["seed"])
    X = torch.randn((n, f), generator=g, device=dev, dtype=torch.float32)
    zero_cols = torch.arange(0, f, 3, device=dev)[:int(0.3 * f)]
    mask = torch.rand((n, zero_cols.numel()), generator=g, device=dev) < 0.6
    X[:, zero_cols] = torch.where(mask, torch.zeros((), device=dev), X[:, zero_cols])
    w = torch.randn(f, generator=g, device=dev) / float(np.sqrt(f))
    s = X @ w + 0.8 * torch.randn(n, generator=g, device=dev)
    # P(label) ~ (.52, .32, .13, .02, .01)
    q = torch.quantile(s[:2_000_000], torch.tensor([0.52, 0.84, 0.97, 0.99], device=dev))
    y = torch.bucketize(s, q).to(torch.float32)
    return X, y.cpu().numpy(), sizes


def shape_leg(ctx, which, steps, warmup):
    """One of the other BASELINE shapes at full size on one GPU: step time, stage split, and the roofline of the kernel
    that shape stresses (Epsilon: the split scan over 2 x B x 8 bytes of histogram per node; MSLR: the histogram pass)."""
    import torch
    from lambdagap_b200 import capi, engine
    cfg = EPSILON if which == "epsilon" else MSLR
    t0 = time.time()
    X, y, group = gen_epsilon(ctx.device) if which == "epsilon" else gen_mslr(ctx.device)
    n, f = X.shape
    sample = X[:min(n, 200_000)].cpu().numpy()
    torch.cuda.synchronize()
    t_gen = time.time() - t0
    t0 = time.time()
    ds = engine.Dataset.from_pointer(X.data_ptr(), n, f, label=y, group=group,
                                     params="max_bin=255 verbose=-1 gpu_device_id=%d" % ctx.device, bin_sample=sample,
                                     num_total_rows=n)
    del X
    torch.cuda.empty_cache()
    t_ds = time.time() - t0
    bst = engine.Booster(ds, cfg["params"] + " device_type=cuda verbose=-1 gpu_device_id=%d" % ctx.device)
    ms, _ = time_steps(ctx, bst, steps, warmup)
    prof = profile_steps(bst, steps)
    capi.synchronize(ctx.device)
    quant = "use_quantized_grad=true" in cfg["params"]
    roof = hist_roofline(prof, f, quant, False, ms / steps, None, which, 1)
    out = {"workload": cfg["name"], "rows": n, "features": f, "params": cfg["params"], "value": steps / (ms / 1e3),
           "unit": UNIT, "ms_per_step": ms / steps, "steps": steps, "warmup": warmup, "datagen_s": round(t_gen, 2),
           "dataset_build_s": round(t_ds, 2),
           "hist_roofline": {k: roof[k] for k in ("kernel", "achieved", "peak", "unit", "frac", "avg_launch_us",
                                                  "stage_share", "stage_ms_per_tree")}}
    if which == "epsilon":
        # split scan: every expanded node scans both children, each over F x 256 entries of 8 bytes (packed int
        # histogram); the larger child also reads the smaller one's entries (parent - smaller in place)
        peak, _ = measured_peaks()
        scan_bytes = prof["expansions"] * 3.0 * f * 256 * 8
        ach = scan_bytes / (prof["find_ms"] / 1e3) / 1e9 if prof["find_ms"] > 0 else 0.0
        out["scan_roofline"] = {"kernel": "find_best_kernel", "bound": "hbm", "achieved": ach, "peak": peak, "unit": "GB/s",
                                "frac": ach / peak, "bytes_per_node": 3.0 * f * 256 * 8,
                                "find_ms_per_tree": prof["find_ms"] / max(1, prof["trees"])}
    bst.free()
    ds.free()
    return out


def _json_leg(cmd, timeout_s, env=None):
    try:
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout_s, env=env)
        for line in reversed(r.stdout.strip().split("\n")):
            if line.startswith("{"):
                return json.loads(line)
        return {"error": ("rc=%d " % r.returncode) + (r.stderr or r.stdout)[-300:]}
    except subprocess.TimeoutExpired:
        return {"error": "timed out after %d s" % timeout_s}
    except Exception as ex:
        return {"error": str(ex)[:300]}


def dropin_leg(rows, timeout_s=240, workload="higgs"):
    """The reference library with the B200 objective and learner compiled in behind its own interfaces (integration/_ref):
    the reference's GBDT, score array and model on the host, device_type=cuda -> B200Objective + B200TreeLearner.  Own
    process.  workload "mslr": the MSLR shape with lambdarank (beside it the same library with the reference's host
    objective, whose pairwise gradients are what the device objective replaces)."""
    if not os.path.exists(os.path.join(ROOT, "integration", "_ref", "lib_lightgbm.so")):
        return {"unavailable": "integration/_ref/lib_lightgbm.so not built (integration/Makefile needs /root/reference)"}
    script = os.path.join(ROOT, "scripts", "dropin_speed.py")
    if workload == "mslr":
        out = _json_leg([sys.executable, script, "0", "", "mslr"], timeout_s)
        if isinstance(out, dict) and "value" in out:
            env = dict(os.environ, GBL_DROPIN_DEVICE_OBJECTIVE="0")
            host = _json_leg([sys.executable, script, "0", "", "mslr"], timeout_s, env=env)
            out["with_host_objective"] = host.get("value") if isinstance(host, dict) else None
        return out
    return _json_leg([sys.executable, script, str(rows)], timeout_s)


def reference_cuda_leg(rows, timeout_s=240):
    script = os.path.join(ROOT, "scripts", "ref_cuda_leg.py")
    if not os.path.exists(os.path.join(ROOT, "oracle", "_ref_cuda", "lib_lightgbm.so")):
        return {"unavailable": "oracle/_ref_cuda/lib_lightgbm.so not built (oracle/Makefile.ref_cuda needs /root/reference)"}
    try:
        r = subprocess.run([sys.executable, script, "--steps", "10", "--warmup", "3", "--rows", str(rows)], capture_output=True,
                           text=True, timeout=timeout_s)
        for line in reversed(r.stdout.strip().split("\n")):
            if line.startswith("{"):
                return json.loads(line)
        return {"error": ("rc=%d " % r.returncode) + (r.stderr or r.stdout)[-300:]}
    except subprocess.TimeoutExpired:
        return {"error": "timed out after %d s" % timeout_s}
    except Exception as ex:
        return {"error": str(ex)[:300]}


# ------------------------------------------------------------------------------------------------ our arm
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="auto", choices=["auto", "higgs", "reg100m", "epsilon", "mslr"])
    ap.add_argument("--rows", type=int, default=0, help="override the row count (debugging only; noted in config)")
    ap.add_argument("--quantized", type=int, default=0, help="1: use_quantized_grad=true (bit-exact mode)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-secondary", action="store_true", help="headline only: skip the reg100m / other-shape / parity legs")
    ap.add_argument("--reg-rows", type=int, default=100_000_000, help="rows of the reg100m leg (debugging only)")
    args = ap.parse_args()
    warmup_requested = args.warmup
    if args.impl == "ours" and args.warmup < 3:
        # the timing rules ask for at least 3 warm-up steps; the line reports both numbers
        args.warmup = 3

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        args.gpus = world
    workload = args.workload
    if workload == "auto":
        workload = "higgs"

    if args.impl == "reference":
        cfg = dict(HIGGS if workload != "reg100m" else REG100M)
        if args.rows:
            cfg["rows"] = args.rows
        out = run_reference(args, cfg, cfg["rows"], rank, world)
        if out is not None:
            print(json.dumps(out), flush=True)
        return

    import torch
    import torch.distributed as dist
    from lambdagap_b200 import capi, engine
    if capi.device_count() < 1:
        raise RuntimeError("bench.py: no sm_100 device visible; this library has no CPU fallback")
    device = local_rank
    torch.cuda.set_device(device)
    ctx = Ctx(rank, world, device)
    ctx.torch, ctx.dist = torch, dist
    secondary = not args.no_secondary
    parity_single = None
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", device))
        if secondary and rank == 0:
            parity_single = parity_single_gpu(device)  # before the network exists: a plain single-GPU learner
        dist.barrier()
        uid = [engine.network_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        engine.network_init_nccl(uid[0], world, rank, device)

    # ---- stand-alone workloads other than the headline
    if workload in ("epsilon", "mslr"):
        if world > 1:
            raise SystemExit("--workload %s is a single-GPU leg" % workload)
        leg = shape_leg(ctx, workload, args.steps, args.warmup)
        leg.update({"metric": METRIC, "n_gpus": 1, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                    "data": "synthetic", "config": {"workload": leg["workload"]}})
        print(json.dumps(leg), flush=True)
        return
    if workload == "reg100m":
        leg = reg100m_leg(ctx, args.steps, args.warmup, args.rows or args.reg_rows)
        if rank == 0:
            leg.update({"metric": METRIC, "n_gpus": world, "higher_is_better": True, "vs_baseline": None,
                        "config": {"workload": leg["workload"], "rows_total": leg["rows_total"], "parallelism": "dp%d" % world}})
            print(json.dumps(leg), flush=True)
        if world > 1:
            dist.barrier()
            engine.network_free()
            dist.destroy_process_group()
        return

    cfg = dict(HIGGS)
    if args.rows:
        cfg["rows"] = args.rows
    if args.quantized:
        cfg["params"] += " use_quantized_grad=true stochastic_rounding=false"
    # weak scaling: every rank holds cfg["rows"] rows of the global N x rows matrix
    local_rows = cfg["rows"]
    global_rows = cfg["rows"] * world
    row_start = rank * local_rows

    # ---- data (host) -> device-resident binned dataset (dataset build is outside the timed region)
    t0 = time.time()
    X, y = gen_block(cfg, row_start, local_rows)
    sample, _ = gen_block(cfg, 0, 200_000)  # identical on every rank -> identical bin mappers
    t_gen = time.time() - t0
    t0 = time.time()
    ds = engine.Dataset(X, y, params="max_bin=255 verbose=-1 gpu_device_id=%d" % device, bin_sample=sample,
                        num_total_rows=global_rows)
    t_ds = time.time() - t0
    params = cfg["params"] + " device_type=cuda verbose=-1 gpu_device_id=%d" % device
    if world > 1:
        params += " tree_learner=data"
    bst = engine.Booster(ds, params)

    # ---- value: device-resident iterations, CUDA events on the learner's stream, max over ranks
    clocks = ClockSampler(device)
    clocks.start()
    for _ in range(args.warmup):
        bst.update()
    ctx.barrier()
    clocks.mark(True)
    ms, launches = time_steps(ctx, bst, args.steps, 0)
    clocks.mark(False)
    value = args.steps / (ms / 1e3)

    # ---- roofline of the dominant kernel (histogram construction): a second pass over the same K steps with
    # CUDA-event stage timing switched on (the learner then drives its waves from the host instead of replaying
    # its CUDA graph, which perturbs the step time -- hence `value` above is timed without it)
    prof = profile_steps(bst, args.steps)
    # The timed region lasts tens of milliseconds, one nvidia-smi sample takes longer: the SAME load keeps running
    # (the profiling pass above, then untimed extra iterations, the same number on every rank) for ~0.6 s in total so
    # that the sampler sees it
    extra = int(min(400, max(8, 600.0 / max(1e-3, ms / args.steps) - 2 * args.steps)))
    for _ in range(extra):
        bst.update()
    capi.synchronize(device)
    clocks.load_end = time.time()
    clk = clocks.stop()
    clk["extra_untimed_steps_for_sampling"] = extra
    F = cfg["features"]
    roofline = hist_roofline(prof, F, bool(args.quantized), False, ms / args.steps, clk.get("sm_mhz"), "higgs", world)

    # ---- e2e: the same iteration through the C API with HOST buffers (custom-objective entry point): pinned host
    # gradients / hessians in (H2D), one tree, score update, full score vector out (D2H) -- all inside the timed
    # region; the host-side gradient computation between steps (the caller's objective) is not timed.
    e2e = None
    try:
        cparams = cfg["params"].replace("objective=%s" % cfg["objective"], "objective=custom")
        if world > 1:
            cparams += " tree_learner=data"  # every rank feeds the gradients of its own row shard
        cb = engine.Booster(ds, cparams + " device_type=cuda verbose=-1 gpu_device_id=%d" % device)
        n = local_rows
        g_pin = torch.empty(n, dtype=torch.float32).pin_memory()
        h_pin = torch.empty(n, dtype=torch.float32).pin_memory()
        s_pin = torch.empty(n, dtype=torch.float64).pin_memory()
        g_np, h_np, s_np = g_pin.numpy(), h_pin.numpy(), s_pin.numpy()
        s_np[:] = 0.0
        import ctypes
        nlen = ctypes.c_int64(0)

        def host_objective():
            if cfg["objective"] == "binary":
                p = 1.0 / (1.0 + np.exp(-s_np))
                g_np[:] = (p - y)
                h_np[:] = p * (1.0 - p)
            else:
                g_np[:] = s_np - y
                h_np[:] = 1.0

        total = 0.0
        e2e_steps = min(args.steps, 10)
        for it in range(2 + e2e_steps):
            host_objective()
            ctx.barrier()
            t0 = time.perf_counter()
            cb.update_custom(g_np, h_np)
            engine._check(engine.lib().LGBM_BoosterGetTrainScoreRaw(cb.handle, ctypes.byref(nlen), s_np.ctypes.data_as(ctypes.c_void_p)))
            dt = time.perf_counter() - t0
            if it >= 2:
                total += dt
        total = ctx.max_over_ranks(total)  # slowest rank
        e2e = {"value": e2e_steps / total, "unit": UNIT, "h2d_bytes_per_step": 8 * n * world, "d2h_bytes_per_step": 8 * n * world,
               "steps": e2e_steps, "api": "LGBM_BoosterUpdateOneIterCustom(host grad, host hess) + LGBM_BoosterGetTrainScoreRaw(host)",
               "ms_per_step": 1e3 * total / e2e_steps}
        cb.free()
        del g_pin, h_pin, s_pin
    except Exception as ex:  # a secondary leg must not take the headline with it
        if world > 1:
            raise  # ranks must stay in step: no partial legs in the data-parallel run
        e2e = {"error": str(ex)[:300]}

    # ---- secondary: the same workload with quantized gradients (the bit-exact mode: int8 gradients, integer
    # histograms), same timing method; reported beside the headline, not instead of it
    quant_leg = None
    if world == 1 and not args.quantized:
        try:
            qb = engine.Booster(ds, params + " use_quantized_grad=true stochastic_rounding=false")
            qms, _ = time_steps(ctx, qb, args.steps, args.warmup)
            qp = profile_steps(qb, args.steps)
            qr = hist_roofline(qp, F, True, False, qms / args.steps, clk.get("sm_mhz"), "higgs", world)
            quant_leg = {"params": "use_quantized_grad=true stochastic_rounding=false num_grad_quant_bins=4",
                         "value": args.steps / (qms / 1e3), "unit": UNIT, "ms_per_step": qms / args.steps, "roofline": qr}
            qb.free()
        except Exception as ex:
            quant_leg = {"error": str(ex)[:300]}

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            cpu = cpu_baseline_leg(cfg, local_rows, X, y)
        except Exception as ex:
            cpu = {"error": str(ex)[:300]}

    # ---- everything below needs the headline's memory back
    bst.free()
    ds.free()
    del X, y

    # ---- context, never instead of --impl reference: the reference's own CUDA tree learner (the kernel to beat) on the
    # same box and workload, in a process of its own (oracle/_ref_cuda, scripts/ref_cuda_leg.py)
    ref_cuda = None
    if rank == 0 and world == 1 and secondary and not args.quantized:
        ref_cuda = reference_cuda_leg(cfg["rows"])
    dropin = None
    if rank == 0 and world == 1 and secondary and not args.quantized:
        dropin = dropin_leg(cfg["rows"])

    parity = None
    reg = None
    shapes = None
    if secondary:
        if world > 1:
            parity = parity_multi_gpu(ctx, parity_single)
        try:
            reg = reg100m_leg(ctx, min(args.steps, 5), 3, args.reg_rows)
        except Exception as ex:
            if world > 1:
                raise
            reg = {"error": str(ex)[:300]}
        if world == 1:
            shapes = {}
            for which in ("epsilon", "mslr"):
                try:
                    shapes[which] = shape_leg(ctx, which, min(args.steps, 8), 3)
                except Exception as ex:
                    shapes[which] = {"error": str(ex)[:300]}
            if isinstance(shapes.get("mslr"), dict) and "value" in shapes["mslr"] and not args.quantized:
                shapes["mslr"]["dropin"] = dropin_leg(0, workload="mslr")

    if rank == 0:
        out = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
               "warmup_requested": warmup_requested,
               "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
               "vs_baseline": None, "dtype": "int64 fixed-point (f32 gradients)" if not args.quantized else "int8 gradients / int32 histograms",
               "data": "synthetic",
               "config": {"workload": cfg["name"], "rows_per_gpu": local_rows, "rows_total": global_rows,
                          "features": F, "params": cfg["params"], "parallelism": "dp%d" % world,
                          "l2_policy": "inputs larger than L2 (no flush)", "dataset_build_s": round(t_ds, 2),
                          "datagen_s": round(t_gen, 2)},
               "clocks": clk, "gpu_launches": int(launches), "roofline": roofline, "e2e": e2e, "cpu_baseline": cpu,
               "schedule": {"waves_per_tree": prof["waves"] / max(1, prof["trees"]),
                            "expansions_per_tree": prof["expansions"] / max(1, prof["trees"]),
                            "splits_per_tree": prof["splits"] / max(1, prof["trees"])},
               "quantized_mode": quant_leg, "parity": parity, "reg100m": reg, "other_shapes": shapes,
               "reference_cuda": ref_cuda, "dropin": dropin}
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.barrier()
        engine.network_free()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
