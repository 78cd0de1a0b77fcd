"""The kernel to beat on the same box: the reference's OWN CUDA tree learner (device_type=cuda of the unmodified
reference built with -DUSE_CUDA for sm_100 by oracle/Makefile.ref_cuda) on bench.py's headline workload.  Runs in a
process of its own (bench.py calls it with a timeout) and prints one JSON line.  Measurement infrastructure only.
    python scripts/ref_cuda_leg.py [--steps K] [--warmup W] [--rows N]"""
import argparse
import ctypes
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

LIB = os.path.join(ROOT, "oracle", "_ref_cuda", "lib_lightgbm.so")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--rows", type=int, default=0)
    ap.add_argument("--extra", default="")
    args = ap.parse_args()
    if not os.path.exists(LIB):
        print(json.dumps({"unavailable": "oracle/_ref_cuda/lib_lightgbm.so was not built (oracle/Makefile.ref_cuda)"}))
        return
    cfg = dict(bench.HIGGS)
    if args.rows:
        cfg["rows"] = args.rows
    X, y = bench.gen_block(cfg, 0, cfg["rows"])
    lib = ctypes.CDLL(LIB)
    lib.LGBM_GetLastError.restype = ctypes.c_char_p

    def check(ret):
        if ret != 0:
            raise RuntimeError(lib.LGBM_GetLastError().decode())

    cores = os.cpu_count() or 1
    n, f = X.shape
    ds = ctypes.c_void_p()
    t0 = time.time()
    check(lib.LGBM_DatasetCreateFromMat(X.ctypes.data_as(ctypes.c_void_p), 0, ctypes.c_int32(n), ctypes.c_int32(f), 1,
                                        ("max_bin=255 verbose=-1 device_type=cuda num_threads=%d" % cores).encode(), None,
                                        ctypes.byref(ds)))
    check(lib.LGBM_DatasetSetField(ds, b"label", y.ctypes.data_as(ctypes.c_void_p), ctypes.c_int(n), 0))
    t_ds = time.time() - t0
    params = cfg["params"] + " device_type=cuda verbose=-1 num_threads=%d %s" % (cores, args.extra)
    bst = ctypes.c_void_p()
    t0 = time.time()
    check(lib.LGBM_BoosterCreate(ds, params.encode(), ctypes.byref(bst)))
    t_init = time.time() - t0
    fin = ctypes.c_int(0)
    for _ in range(args.warmup):
        check(lib.LGBM_BoosterUpdateOneIter(bst, ctypes.byref(fin)))
    t0 = time.time()
    for _ in range(args.steps):
        check(lib.LGBM_BoosterUpdateOneIter(bst, ctypes.byref(fin)))
    dt = time.time() - t0
    print(json.dumps({"value": args.steps / dt, "unit": "iters/s", "ms_per_step": 1e3 * dt / args.steps, "steps": args.steps,
                      "warmup": args.warmup, "rows": n, "features": f, "params": params,
                      "what": "unmodified reference, its own CUDA tree learner (src/treelearner/cuda/*, -DUSE_CUDA, sm_100), "
                              "LGBM_BoosterUpdateOneIter wall clock",
                      "dataset_build_s": round(t_ds, 2), "booster_init_s": round(t_init, 2)}), flush=True)


if __name__ == "__main__":
    main()
