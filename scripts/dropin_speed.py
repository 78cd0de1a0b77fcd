"""Speed of the drop-in library (integration/_ref/lib_lightgbm.so: the reference's GBDT on the host + the B200 objective and tree learner)
on bench.py's headline workload, LGBM_BoosterUpdateOneIter wall clock.  Prints one JSON line.  Own process (see
tests/dropin_runner.py for why)."""
import ctypes
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

lib = ctypes.CDLL(os.path.join(ROOT, "integration", "_ref", "lib_lightgbm.so"))
lib.LGBM_GetLastError.restype = ctypes.c_char_p


def check(r):
    if r != 0:
        raise RuntimeError(lib.LGBM_GetLastError().decode())


# usage: dropin_speed.py [rows [extra params [higgs|mslr]]]
workload = sys.argv[3] if len(sys.argv) > 3 else "higgs"
extra = sys.argv[2] if len(sys.argv) > 2 else ""
group = None
if workload == "mslr":
    cfg = dict(bench.MSLR)
    Xd, y, group = bench.gen_mslr(0)
    X = Xd.cpu().numpy()
    del Xd
else:
    cfg = dict(bench.HIGGS)
    rows = int(sys.argv[1]) if len(sys.argv) > 1 and int(sys.argv[1]) > 0 else cfg["rows"]
    X, y = bench.gen_block(cfg, 0, rows)
n, f = X.shape
cores = os.cpu_count() or 1
ds = ctypes.c_void_p()
check(lib.LGBM_DatasetCreateFromMat(X.ctypes.data_as(ctypes.c_void_p), 0, ctypes.c_int32(n), ctypes.c_int32(f), 1,
                                    ("max_bin=255 verbose=-1 num_threads=%d" % cores).encode(), None, ctypes.byref(ds)))
check(lib.LGBM_DatasetSetField(ds, b"label", y.ctypes.data_as(ctypes.c_void_p), ctypes.c_int(n), 0))
if group is not None:
    import numpy as np
    group = np.ascontiguousarray(group, dtype=np.int32)
    check(lib.LGBM_DatasetSetField(ds, b"group", group.ctypes.data_as(ctypes.c_void_p), ctypes.c_int(len(group)), 2))
bst = ctypes.c_void_p()
params = cfg["params"] + " device_type=cuda verbose=-1 num_threads=%d %s" % (cores, extra)
check(lib.LGBM_BoosterCreate(ds, params.encode(), ctypes.byref(bst)))
fin = ctypes.c_int(0)
for _ in range(3):
    check(lib.LGBM_BoosterUpdateOneIter(bst, ctypes.byref(fin)))
steps = 10
t0 = time.time()
for _ in range(steps):
    check(lib.LGBM_BoosterUpdateOneIter(bst, ctypes.byref(fin)))
dt = time.time() - t0
print(json.dumps({"value": steps / dt, "unit": "iters/s", "ms_per_step": 1e3 * dt / steps, "rows": n, "params": params,
                  "what": "the reference library with B200Objective + B200TreeLearner compiled in (integration/_ref): its own GBDT, score array and model on the host, gradients and tree on the device; LGBM_BoosterUpdateOneIter wall clock"}))
check(lib.LGBM_BoosterFree(bst))  # releases the learner (prints the GBL_DROPIN_TIMING split)
