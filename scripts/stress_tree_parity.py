"""Repeats one tree many times (fresh learner every few trees, quantized and float) against the oracle's C
restatement: catches schedule-dependent results (races) that a single parity run can miss.
Usage: python scripts/stress_tree_parity.py [reps] [rows]"""
import sys
import numpy as np
sys.path.insert(0, ".")
from lambdagap_b200 import capi
from oracle import oracle as O

def build(seed, n, f):
    rng = np.random.default_rng(seed)
    infos, cols = [], []
    for j in range(f):
        nb = [255, 64, 12, 2][j % 4]
        mt = j % 3
        col = rng.integers(0, nb, size=n).astype(np.uint8)
        if j % 5 == 0:
            col[rng.random(n) < 0.6] = nb // 2
        cnt = np.bincount(col, minlength=nb)
        default_bin = int(rng.integers(0, nb)) if mt == 1 else 0
        mfb = int(np.argmax(cnt))
        if mfb != default_bin and cnt[mfb] / n < 0.7:
            mfb = default_bin
        infos.append(dict(num_bin=nb, missing_type=mt if nb > 2 else [0, 0, 2][mt], default_bin=default_bin,
                          most_freq_bin=mfb, real_index=f - 1 - j, upper_bounds=np.arange(nb, dtype=np.float64)))
        cols.append(col)
    g = rng.normal(size=n).astype(np.float32)
    h = rng.uniform(0.05, 1.0, size=n).astype(np.float32)
    return infos, cols, g, h

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 10
n = int(sys.argv[2]) if len(sys.argv) > 2 else 60000
bad = 0
infos, cols, g, h = build(17, n, 33)
kw = dict(num_leaves=63, min_data_in_leaf=25, lambda_l2=0.5)
ref = {}
for quant in (1, 0):
    ref[quant] = O.train_tree(cols, infos, g, h, O.make_config(use_quantized_grad=quant, **kw))
for rep in range(reps):
    for quant in (1, 0):
        root, recs, leaf_of_row = ref[quant]
        dds = capi.Dataset.from_bins(infos, cols)
        lrn = capi.Learner(dds, capi.learner_config(use_quantized_grad=quant, stochastic_rounding=0, **kw))
        for rep2 in range(3):
            res, ours = lrn.train_host(g, h)
            ok = res["num_splits"] == len(recs)
            for k, (a, b) in enumerate(zip(ours, recs)):
                badk = [key for key in ("leaf", "feature", "threshold_bin", "default_left", "left_count", "right_count") if a[key] != b[key]]
                if badk:
                    ok = False
                    print("rep", rep, rep2, "quant", quant, "split", k, badk, {x: a[x] for x in ("leaf", "feature", "threshold_bin", "gain")},
                          {x: b[x] for x in ("leaf", "feature", "threshold_bin", "gain")}, flush=True)
                    break
            bad += (not ok)
        lrn.free()
        dds.free()
print("mismatching trees:", bad, "of", reps * 6)
