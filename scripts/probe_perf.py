"""Perf probe (not a bench): random bins of a given shape, random gradients, N trees; prints ms/tree."""
import sys
import time

import numpy as np

sys.path.insert(0, ".")
from lambdagap_b200 import capi

n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 10_500_000
f = int(sys.argv[2]) if len(sys.argv) > 2 else 28
leaves = int(sys.argv[3]) if len(sys.argv) > 3 else 255
quant = int(sys.argv[4]) if len(sys.argv) > 4 else 0
trees = int(sys.argv[5]) if len(sys.argv) > 5 else 5
rng = np.random.default_rng(0)
infos = [dict(num_bin=255, missing_type=0, default_bin=0, most_freq_bin=0, real_index=i) for i in range(f)]
t0 = time.time()
X = rng.normal(size=(n, f)).astype(np.float32)
w = rng.normal(size=f)
y = (X @ w + rng.normal(size=n) * 2 > 0).astype(np.float32)
cols = []
for i in range(f):
    qs = np.quantile(X[::50, i], np.linspace(0, 1, 256)[1:-1])
    cols.append(np.searchsorted(qs, X[:, i]).astype(np.uint8))
print("data+bins %.1fs" % (time.time() - t0), flush=True)
ds = capi.Dataset.from_bins(infos, cols)
cfg = capi.learner_config(num_leaves=leaves, use_quantized_grad=quant, stochastic_rounding=0)
lrn = capi.Learner(ds, cfg)
score = np.zeros(n)
gbuf = capi.DeviceBuffer(4 * n)
hbuf = capi.DeviceBuffer(4 * n)
import os
if os.environ.get('GBL_PROF'):
    lrn.set_profiling(True)
for t in range(trees):
    p = 1 / (1 + np.exp(-score))
    g = (p - y).astype(np.float32)
    h = (p * (1 - p)).astype(np.float32)
    gbuf.upload(g)
    hbuf.upload(h)
    capi.synchronize()
    l0 = capi.kernel_launch_count()
    t0 = time.time()
    res, recs = lrn.train(gbuf, hbuf)
    dt = time.time() - t0
    print("tree %d: %.2f ms, splits=%d launches=%d" % (t, dt * 1e3, res["num_splits"], capi.kernel_launch_count() - l0), flush=True)
    if os.environ.get('GBL_PROF'):
        print("profile", lrn.profile(reset=True), flush=True)
    sizes = [n] + [min(r["left_count"], r["right_count"]) for r in recs]
    print("hist_rows", " ".join(str(x) for x in sizes[:len(recs)]), flush=True)
    assign = lrn.leaf_assignment()
    vals = np.zeros(leaves)
    vals[0] = res["root_output"]
    for k, r in enumerate(recs):
        vals[r["leaf"]] = r["left_output"]
        vals[k + 1] = r["right_output"]
    score += 0.1 * vals[assign]
