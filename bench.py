#!/usr/bin/env python
"""bench.py -- boosting iterations/sec of the histogram-GBDT hot path on B200 (BASELINE.json metric).

  python bench.py --gpus N --steps K --warmup W [--impl reference] [--workload higgs|reg100m]

A "step" is one boosting iteration (gradients -> one leaf-wise tree of num_leaves=255 -> score update).
  N=1  workload = BASELINE.json configs[1]: Higgs-shape synthetic binary classification, 10.5M rows x 28 dense
       features, max_bin=255, num_leaves=255 (float gradients).  Inputs are larger than L2 (294 MB of bins + 336 MB
       gather mirror + 84 MB gradients per step), so no L2 flush is needed between steps.
  N>1  the same per-GPU shape sharded by rows over N ranks (data-parallel learner, one NCCL all-reduce of the
       smaller-leaf histogram per split): weak scaling, `value` = iterations/sec of the whole N x 10.5M-row job.
       --workload reg100m selects BASELINE.json configs[4] instead (100M x 100 regression, strong scaling).
Prints ONE JSON line on rank 0.  `value` = device-resident loop timed with CUDA events on the learner's stream;
`e2e` = the same iteration driven through the C API with HOST gradient / hessian buffers in and the score vector
out (LGBM_BoosterUpdateOneIterCustom + LGBM_BoosterGetTrainScoreRaw), copies inside the timed region.
`--impl reference` times the UNMODIFIED reference CPU learner (oracle/_ref/lib_lightgbm.so, built from
/root/reference by oracle/Makefile.ref) on the host cores on the same workload.
"""
import argparse
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


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""

    def __init__(self, device):
        self.device = device
        self.proc = None
        self.lines = []   # (arrival time, csv line)
        self.t0 = None    # timed region (time.time()), set by mark()
        self.t1 = None
        self.load_end = None  # end of the same load continued after the timed region

    def samples_since_start(self):
        return len([1 for (t, _) in self.lines if self.t0 is not None and t >= self.t0])

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


def ncu_traffic():
    """Per-launch DRAM bytes of the histogram kernel from the committed ncu --set full capture, if any."""
    p = os.path.join(ROOT, "profiles", "hist_traffic.json")
    if os.path.exists(p):
        with open(p) as fh:
            return json.load(fh)
    return None


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


# ------------------------------------------------------------------------------------------------ our arm
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="auto", choices=["auto", "higgs", "reg100m"])
    ap.add_argument("--rows", type=int, default=0, help="override the row count (debugging only; noted in config)")
    ap.add_argument("--quantized", type=int, default=0, help="1: use_quantized_grad=true (bit-exact mode)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else max(args.warmup, 1)

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        args.gpus = world
    workload = args.workload
    if workload == "auto":
        workload = "higgs"
    cfg = dict(HIGGS if workload == "higgs" else REG100M)
    strong = workload == "reg100m"
    if args.rows:
        cfg["rows"] = args.rows
    if args.quantized:
        cfg["params"] += " use_quantized_grad=true stochastic_rounding=false"
    # weak scaling (higgs): every rank holds cfg["rows"] rows; strong scaling (reg100m): rows are divided
    local_rows = cfg["rows"] // world if strong else cfg["rows"]
    global_rows = cfg["rows"] if strong else cfg["rows"] * world
    row_start = rank * local_rows

    if args.impl == "reference":
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
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", device))
        uid = [engine.network_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        engine.network_init_nccl(uid[0], world, rank, device)

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
    lrn = bst.learner()

    def barrier():
        capi.synchronize(device)
        if world > 1:
            dist.barrier()
        capi.synchronize(device)

    # ---- value: device-resident iterations, CUDA events on the learner's stream, max over ranks
    clocks = ClockSampler(device)
    clocks.start()
    for _ in range(args.warmup):
        bst.update()
    barrier()
    clocks.mark(True)
    launches0 = capi.kernel_launch_count()
    lrn.timer_record(0)
    for _ in range(args.steps):
        bst.update()
    lrn.timer_record(1)
    ms = lrn.timer_elapsed_ms(0, 1)
    barrier()
    clocks.mark(False)
    launches = capi.kernel_launch_count() - launches0
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    value = args.steps / (ms / 1e3)

    # ---- roofline of the dominant kernel (histogram construction): a second pass over the same K steps with
    # CUDA-event stage timing switched on (the learner then drives its waves from the host instead of replaying
    # its CUDA graph, which perturbs the step time -- hence `value` above is timed without it)
    lrn.set_profiling(True)
    lrn.profile(reset=True)
    for _ in range(args.steps):
        bst.update()
    prof = lrn.profile(reset=True)
    lrn.set_profiling(False)
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
    gh_bytes = 2 if args.quantized else 8
    hist_entry = 8 if args.quantized else 16
    # algorithmic bytes of one histogram call on n rows: n * (F*1 + gh + 4 [index list]) + histogram write-out
    alg_bytes = prof["hist_rows"] * (F * 1 + gh_bytes + 4) + prof["hist_launches"] * (F * 256 * hist_entry)
    hist_s = prof["hist_ms"] / 1e3
    achieved = alg_bytes / hist_s / 1e9 if hist_s > 0 else 0.0
    peak, peak_src = measured_peaks()
    traffic = ncu_traffic()
    stage_total = prof["hist_ms"] + prof["find_ms"] + prof["partition_ms"] + prof["other_ms"] + prof["collective_ms"]
    roofline = {"bound": "hbm", "kernel": "hist_quant_kernel" if args.quantized else "hist_float_kernel",
                "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "peak_source": peak_src,
                "traffic": traffic["dram_bytes_per_launch"] if traffic else None,
                "algorithmic_bytes_per_launch": alg_bytes / max(1, prof["hist_launches"]),
                "avg_launch_us": 1e3 * prof["hist_ms"] / max(1, prof["hist_launches"]),
                "launches_per_step": prof["hist_launches"] / max(1, prof["trees"]),
                "stage_share": {k: (prof[k + "_ms"] / stage_total if stage_total > 0 else None)
                                for k in ("hist", "find", "partition", "other", "collective")}}

    # ---- e2e: the same iteration through the C API with HOST buffers (custom-objective entry point): pinned host
    # gradients / hessians in (H2D), one tree, score update, full score vector out (D2H) -- all inside the timed
    # region; the host-side gradient computation between steps (the caller's objective) is not timed.
    e2e = None
    try:
        if True:
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
                barrier()
                t0 = time.perf_counter()
                cb.update_custom(g_np, h_np)
                engine._check(engine.lib().LGBM_BoosterGetTrainScoreRaw(cb.handle, ctypes.byref(nlen), s_np.ctypes.data_as(ctypes.c_void_p)))
                dt = time.perf_counter() - t0
                if it >= 2:
                    total += dt
            if world > 1:  # slowest rank
                tt = torch.tensor([total], dtype=torch.float64, device="cuda")
                dist.all_reduce(tt, op=dist.ReduceOp.MAX)
                total = float(tt.item())
            e2e = {"value": e2e_steps / total, "unit": UNIT, "h2d_bytes_per_step": 8 * n * world, "d2h_bytes_per_step": 8 * n * world,
                   "steps": e2e_steps, "api": "LGBM_BoosterUpdateOneIterCustom(host grad, host hess) + LGBM_BoosterGetTrainScoreRaw(host)",
                   "ms_per_step": 1e3 * total / e2e_steps}
            cb.free()
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
        ql = qb.learner()
        for _ in range(args.warmup):
            qb.update()
        capi.synchronize(device)
        ql.timer_record(0)
        for _ in range(args.steps):
            qb.update()
        ql.timer_record(1)
        qms = ql.timer_elapsed_ms(0, 1)
        ql.set_profiling(True)
        ql.profile(reset=True)
        for _ in range(args.steps):
            qb.update()
        qp = ql.profile(reset=True)
        ql.set_profiling(False)
        q_alg = qp["hist_rows"] * (F * 1 + 2 + 4) + qp["hist_launches"] * (F * 256 * 8)
        q_ach = q_alg / (qp["hist_ms"] / 1e3) / 1e9 if qp["hist_ms"] > 0 else 0.0
        quant_leg = {"params": "use_quantized_grad=true stochastic_rounding=false num_grad_quant_bins=4",
                     "value": args.steps / (qms / 1e3), "unit": UNIT, "ms_per_step": qms / args.steps,
                     "roofline": {"kernel": "hist_quant_kernel", "achieved": q_ach, "peak": peak, "unit": "GB/s",
                                  "frac": q_ach / peak, "avg_launch_us": 1e3 * qp["hist_ms"] / max(1, qp["hist_launches"])}}
        qb.free()
      except Exception as ex:
        quant_leg = {"error": str(ex)[:300]}

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            cpu = cpu_baseline_leg(cfg, local_rows, X, y)
        except Exception as ex:
            cpu = {"error": str(ex)[:300]}

    if rank == 0:
        out = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
               "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong" if strong else "weak",
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
               "quantized_mode": quant_leg}
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.barrier()
        engine.network_free()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
