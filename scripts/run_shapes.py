"""Runs the other BASELINE.json shapes end to end on one GPU at full size (synthetic), a few iterations each, with
size-independent sanity properties: leaf counts sum to N, scores finite, training loss decreases, stage profile."""
import sys
import time

import numpy as np

sys.path.insert(0, ".")
from lambdagap_b200 import capi, engine
from tests import helpers as H

which = sys.argv[1] if len(sys.argv) > 1 else "epsilon"
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 5
rng = np.random.default_rng(44)
t0 = time.time()
if which == "epsilon":
    n, f = 400_000, 2000
    X = rng.normal(size=(n, f)).astype(np.float32)
    X /= np.linalg.norm(X, axis=1, keepdims=True)
    w = np.zeros(f)
    w[rng.choice(f, 50, replace=False)] = rng.normal(size=50)
    y = (X @ w + 0.01 * rng.normal(size=n) > 0).astype(np.float32)
    group = None
    params = "objective=binary num_leaves=255 learning_rate=0.1 use_quantized_grad=true stochastic_rounding=false"
elif which == "mslr":
    # 31 531 queries, clipped lognormal sizes with mean ~120 (max 1 251) -> ~3.77 M documents
    X, y, group = H.make_ranking(31531, 136, seed=43, min_len=1, max_len=1251, zero_frac=0.3, median_len=95)
    n, f = X.shape
    params = "objective=lambdarank lambdarank_target=%s lambdarank_truncation_level=30 num_leaves=255 learning_rate=0.1" % (
        sys.argv[3] if len(sys.argv) > 3 else "ndcg")
elif which == "higgs":
    # the bench shape with extra parameters from the command line, e.g. "bagging_fraction=0.9 bagging_freq=1" or
    # "metric=auc,binary_logloss" (then the training metrics are evaluated after every iteration and timed)
    n, f = 10_500_000, 28
    X = rng.normal(size=(n, f)).astype(np.float32)
    X[:, 1] = np.exp(X[:, 1])
    w = rng.normal(size=f)
    y = (X @ w + rng.normal(size=n) > 0).astype(np.float32)
    group = None
    params = "objective=binary num_leaves=255 learning_rate=0.1 " + (sys.argv[3] if len(sys.argv) > 3 else "")
else:
    raise SystemExit("epsilon | mslr | higgs")
print(which, X.shape, "gen %.1fs" % (time.time() - t0), flush=True)
t0 = time.time()
ds = engine.Dataset(X, y, group=group, params="max_bin=255 verbose=-1")
print("dataset %.1fs" % (time.time() - t0), flush=True)
bst = engine.Booster(ds, params + " device_type=cuda verbose=-1")
lrn = bst.learner()
for it in range(iters):
    capi.synchronize()
    t0 = time.time()
    bst.update()
    dt = time.time() - t0
    ev = ""
    if "metric=" in params:
        t1 = time.time()
        vals = bst.eval(0)
        ev = "  eval %s = %s in %.2f ms" % (bst.eval_names(), np.round(vals, 6), (time.time() - t1) * 1e3)
    print("iter %d: %.2f ms%s" % (it, dt * 1e3, ev), flush=True)
lrn.set_profiling(True)
bst.update()
print("profile", lrn.profile(reset=True))
s = bst.train_scores(raw=True)
assert np.all(np.isfinite(s))
from oracle import ref_lgbm as R
trees = R.parse_model_trees(bst.model_to_string())
for t in trees:
    assert int(t["leaf_count"].sum()) == n, (t["leaf_count"].sum(), n)
print("leaves per tree", [t["num_leaves"] for t in trees][:8], "score range", s.min(), s.max())
if which == "epsilon":
    p = 1 / (1 + np.exp(-s))
    print("train logloss", float(-(y * np.log(p) + (1 - y) * np.log(1 - p)).mean()), "acc", float(((p > 0.5) == (y > 0.5)).mean()))
print("OK", which)
