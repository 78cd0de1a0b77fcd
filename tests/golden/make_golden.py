"""Generates the golden fixtures of tests/golden/*.npz from the UNMODIFIED reference CPU library
(oracle/_ref/lib_lightgbm.so, built from /root/reference by oracle/Makefile.ref).  Run in the build container:

    python tests/golden/make_golden.py

Each fixture holds a small seeded input (raw matrix, labels, query sizes), what the reference derived from it
(per-feature BinMapper: bin upper bounds, most-frequent / default bin, missing type; the mapper bin of every cell)
and what the reference trained on it (every tree of the model string, raw training scores).  The fixtures pin
(a) the oracle's C restatement and the host BinMapper on CPU (tests/test_oracle_cpu.py) and (b) the CUDA path on the
GPU box even where oracle/_ref is absent (tests/test_golden_gpu.py).
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import ref_lgbm as R  # noqa: E402
from tests import helpers as H  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
ITERS = 3

CASES = {
    "binary_float": dict(kind="binary", params="objective=binary num_leaves=15 learning_rate=0.1 min_data_in_leaf=20"),
    "binary_quant": dict(kind="binary", params="objective=binary num_leaves=15 learning_rate=0.1 min_data_in_leaf=20 "
                                               "use_quantized_grad=true stochastic_rounding=false num_grad_quant_bins=4"),
    "regression_l2": dict(kind="regression", params="objective=regression num_leaves=31 learning_rate=0.05 "
                                                    "min_data_in_leaf=10 lambda_l2=1.0 lambda_l1=0.2 max_depth=6"),
    "rank_ndcg": dict(kind="rank", params="objective=lambdarank lambdarank_target=ndcg lambdarank_truncation_level=5 "
                                          "num_leaves=15 min_data_in_leaf=10"),
    "rank_lambdagap_x": dict(kind="rank", params="objective=lambdarank lambdarank_target=lambdagap-x "
                                                 "lambdarank_truncation_level=5 num_leaves=15 min_data_in_leaf=10"),
    "rank_arpk": dict(kind="rank", params="objective=lambdarank lambdarank_target=arpk lambdarank_truncation_level=5 "
                                          "num_leaves=15 min_data_in_leaf=10"),
    "rank_ll_ndcg_pp": dict(kind="rank", params="objective=lambdarank lambdarank_target=lambdaloss-ndcg-plus-plus "
                                                "lambdagap_weight=1.5 lambdarank_truncation_level=5 num_leaves=15 "
                                                "min_data_in_leaf=10"),
    "rank_precision": dict(kind="rank", params="objective=lambdarank lambdarank_target=precision "
                                               "lambdarank_truncation_level=5 num_leaves=15 min_data_in_leaf=10"),
}

TREE_INT = ("split_feature", "decision_type", "left_child", "right_child", "leaf_count", "internal_count")
TREE_FLT = ("split_gain", "threshold", "leaf_value", "internal_value")


def make_inputs(kind, seed):
    if kind == "rank":
        X, y, group = H.make_ranking(70, 8, seed=seed, min_len=6, max_len=60)
        return X, y, group
    if kind == "binary":
        X, y = H.make_binary(3000, 8, seed=seed, nan_frac=0.05, zero_frac=0.3, discrete_cols=1)
        return X, y, None
    rng = np.random.default_rng(seed)
    X = rng.normal(size=(3000, 8)).astype(np.float32)
    X[rng.random(3000) < 0.4, 2] = 0.0
    y = (2 * X[:, 0] + np.sin(3 * X[:, 1]) + X[:, 2] * X[:, 3] + 0.1 * rng.normal(size=3000)).astype(np.float32)
    return X, y, None


def pack_trees(trees):
    out = {"num_leaves": np.array([t["num_leaves"] for t in trees], dtype=np.int64)}
    m = max(2, int(out["num_leaves"].max()))
    for key in TREE_INT:
        a = np.full((len(trees), m), -999, dtype=np.int64)
        for i, t in enumerate(trees):
            v = t.get(key, np.zeros(0, dtype=np.int64))
            a[i, :len(v)] = v
        out["tree_" + key] = a
    for key in TREE_FLT:
        a = np.full((len(trees), m), np.nan, dtype=np.float64)
        for i, t in enumerate(trees):
            v = t.get(key, np.zeros(0))
            a[i, :len(v)] = v
        out["tree_" + key] = a
    return out


def main():
    for idx, (name, case) in enumerate(sorted(CASES.items())):
        X, y, group = make_inputs(case["kind"], 100 + idx)
        ds = R.RefDataset(X, y, group=group, params="max_bin=255 verbose=-1 num_threads=4 force_col_wise=true")
        bst = R.RefBooster(ds, case["params"] + " device_type=cpu force_col_wise=true deterministic=true num_threads=4 verbose=-1")
        for _ in range(ITERS):
            bst.update()
        trees = R.parse_model_trees(bst.model_string())
        scores = bst.train_scores_raw() if case["kind"] == "binary" else bst.train_scores()
        nf = ds.num_used_features()
        infos = [ds.feature_info(f) for f in range(nf)]
        meta = np.array([[i["num_bin"], i["missing_type"], i["default_bin"], i["most_freq_bin"], i["real_index"]]
                         for i in infos], dtype=np.int32)
        ub = np.concatenate([i["upper_bounds"] for i in infos])
        bins = np.stack([ds.feature_bins(f) for f in range(nf)])
        out = dict(X=X, y=y, group=group if group is not None else np.zeros(0, dtype=np.int32), meta=meta,
                   upper_bounds=ub, bins=bins, scores=scores, params=np.array(case["params"]), kind=np.array(case["kind"]),
                   iters=np.array(ITERS))
        out.update(pack_trees(trees))
        np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
        print(name, X.shape, "trees", [t["num_leaves"] for t in trees])
        bst.free()
        ds.free()


if __name__ == "__main__":
    main()
