"""Shared test helpers: synthetic data, host Tree bookkeeping from split records, reference harness glue."""
import numpy as np

from oracle import ref_lgbm as R


def make_binary(n, f, seed=0, heavy_tail=True, nan_frac=0.0, zero_frac=0.0, discrete_cols=0):
    rng = np.random.default_rng(seed)
    X = rng.normal(size=(n, f)).astype(np.float32)
    if heavy_tail and f >= 4:
        X[:, 1] = np.exp(X[:, 1])
        X[:, 3] = np.exp(2 * X[:, 3])
    w = rng.normal(size=f)
    logits = X @ w / np.sqrt(f) + 0.5 * rng.normal(size=n)
    y = (logits > np.median(logits)).astype(np.float32)
    if zero_frac > 0:
        for c in range(0, f, 3):
            X[rng.random(n) < zero_frac, c] = 0.0
    for c in range(discrete_cols):
        X[:, f - 1 - c] = rng.integers(0, 6 + 3 * c, size=n).astype(np.float32)
    if nan_frac > 0:
        for c in range(2, f, 4):
            X[rng.random(n) < nan_frac, c] = np.nan
    return X, y


def ref_infos_and_bins(ref_ds):
    infos, cols = [], []
    for i in range(ref_ds.num_used_features()):
        info = ref_ds.feature_info(i)
        infos.append(info)
        cols.append(ref_ds.feature_bins(i))
    return infos, cols


class HostTree:
    """Tree bookkeeping of the reference (include/LightGBM/tree.h:543-583, src/io/tree.cpp:60-74) from split records."""

    def __init__(self, max_leaves):
        self.num_leaves = 1
        n = max_leaves
        self.left_child = np.zeros(n - 1, dtype=np.int64)
        self.right_child = np.zeros(n - 1, dtype=np.int64)
        self.split_feature = np.zeros(n - 1, dtype=np.int64)
        self.split_feature_inner = np.zeros(n - 1, dtype=np.int64)
        self.threshold_bin = np.zeros(n - 1, dtype=np.int64)
        self.threshold = np.zeros(n - 1, dtype=np.float64)
        self.decision_type = np.zeros(n - 1, dtype=np.int64)
        self.split_gain = np.zeros(n - 1, dtype=np.float64)
        self.internal_value = np.zeros(n - 1, dtype=np.float64)
        self.internal_weight = np.zeros(n - 1, dtype=np.float64)
        self.internal_count = np.zeros(n - 1, dtype=np.int64)
        self.leaf_parent = np.full(n, -1, dtype=np.int64)
        self.leaf_value = np.zeros(n, dtype=np.float64)
        self.leaf_weight = np.zeros(n, dtype=np.float64)
        self.leaf_count = np.zeros(n, dtype=np.int64)

    def split(self, rec, infos, min_gain_to_split=0.0):
        leaf = rec["leaf"]
        new = self.num_leaves - 1
        parent = self.leaf_parent[leaf]
        if parent >= 0:
            if self.left_child[parent] == ~leaf:
                self.left_child[parent] = new
            else:
                self.right_child[parent] = new
        info = infos[rec["feature"]]
        self.split_feature_inner[new] = rec["feature"]
        self.split_feature[new] = rec["real_feature"]
        self.split_gain[new] = np.float32(rec["gain"] + min_gain_to_split)
        self.left_child[new] = ~leaf
        self.right_child[new] = ~self.num_leaves
        self.leaf_parent[leaf] = new
        self.leaf_parent[self.num_leaves] = new
        self.internal_weight[new] = rec["left_sum_hessian"] + rec["right_sum_hessian"]
        self.internal_value[new] = self.leaf_value[leaf]
        self.internal_count[new] = rec["left_count"] + rec["right_count"]
        self.leaf_value[leaf] = 0.0 if np.isnan(rec["left_output"]) else rec["left_output"]
        self.leaf_weight[leaf] = rec["left_sum_hessian"]
        self.leaf_count[leaf] = rec["left_count"]
        r = self.num_leaves
        self.leaf_value[r] = 0.0 if np.isnan(rec["right_output"]) else rec["right_output"]
        self.leaf_weight[r] = rec["right_sum_hessian"]
        self.leaf_count[r] = rec["right_count"]
        self.threshold_bin[new] = rec["threshold_bin"]
        self.threshold[new] = info["upper_bounds"][rec["threshold_bin"]]
        self.decision_type[new] = (2 if rec["default_left"] else 0) | (info["missing_type"] << 2)
        self.num_leaves += 1


def build_tree(result, records, infos, max_leaves, min_gain_to_split=0.0):
    t = HostTree(max_leaves)
    t.leaf_value[0] = result["root_output"]
    for rec in records:
        t.split(rec, infos, min_gain_to_split)
    return t


def assert_tree_matches(ours, ref, shrinkage=1.0, bias=0.0, rtol=1e-5, exact_counts=True, what=""):
    """ours: HostTree (unshrunk); ref: dict from ref_lgbm.parse_model_trees."""
    nl = ref["num_leaves"]
    assert ours.num_leaves == nl, "%s num_leaves %d vs ref %d" % (what, ours.num_leaves, nl)
    if nl == 1:
        return
    ni = nl - 1
    np.testing.assert_array_equal(ours.split_feature[:ni], ref["split_feature"], err_msg=what + " split_feature")
    np.testing.assert_array_equal(ours.left_child[:ni], ref["left_child"], err_msg=what + " left_child")
    np.testing.assert_array_equal(ours.right_child[:ni], ref["right_child"], err_msg=what + " right_child")
    np.testing.assert_array_equal(ours.decision_type[:ni], ref["decision_type"], err_msg=what + " decision_type")
    np.testing.assert_allclose(ours.threshold[:ni], ref["threshold"], rtol=1e-15, atol=0, err_msg=what + " threshold")
    if exact_counts:
        np.testing.assert_array_equal(ours.leaf_count[:nl], ref["leaf_count"], err_msg=what + " leaf_count")
        np.testing.assert_array_equal(ours.internal_count[:ni], ref["internal_count"], err_msg=what + " internal_count")
    np.testing.assert_allclose(ours.leaf_value[:nl] * shrinkage + bias, ref["leaf_value"], rtol=rtol, atol=1e-12,
                               err_msg=what + " leaf_value")
    np.testing.assert_allclose(ours.split_gain[:ni], ref["split_gain"], rtol=max(rtol, 1e-5), err_msg=what + " split_gain (model text prints 6 digits)")


def make_ranking(num_queries, f, seed=0, min_len=2, max_len=60, zero_frac=0.3, max_label=4, median_len=None):
    """MSLR-shaped synthetic ranking data: query sizes ~ clipped lognormal, labels 0..max_label correlated with a
    linear score, some zero-inflated columns."""
    rng = np.random.default_rng(seed)
    med = max(3, max_len / 3) if median_len is None else median_len
    sizes = np.clip(rng.lognormal(mean=np.log(med), sigma=0.7, size=num_queries).astype(np.int64), min_len, max_len)
    n = int(sizes.sum())
    X = rng.normal(size=(n, f)).astype(np.float32)
    for c in range(0, f, 3):
        X[rng.random(n) < zero_frac, c] = 0.0
    w = rng.normal(size=f)
    s = X @ w / np.sqrt(f) + rng.normal(size=n)
    qs = np.quantile(s, [0.52, 0.84, 0.97, 0.99][:max_label])
    y = np.searchsorted(qs, s).astype(np.float32)
    return X, y, sizes.astype(np.int32)


def load_golden(name):
    import os
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", name + ".npz")
    z = np.load(path)
    meta = z["meta"]
    ub = z["upper_bounds"]
    infos, off = [], 0
    for row in meta:
        nb = int(row[0])
        infos.append(dict(num_bin=nb, missing_type=int(row[1]), default_bin=int(row[2]), most_freq_bin=int(row[3]),
                          real_index=int(row[4]), upper_bounds=ub[off:off + nb].copy()))
        off += nb
    trees = []
    for i, nl in enumerate(z["num_leaves"]):
        nl = int(nl)
        t = {"num_leaves": nl}
        for key in ("split_feature", "decision_type", "left_child", "right_child", "internal_count", "split_gain",
                    "threshold", "internal_value"):
            t[key] = z["tree_" + key][i, :nl - 1]
        for key in ("leaf_count", "leaf_value"):
            t[key] = z["tree_" + key][i, :nl]
        trees.append(t)
    return dict(X=z["X"], y=z["y"], group=z["group"] if len(z["group"]) else None, infos=infos, bins=z["bins"],
                scores=z["scores"], params=str(z["params"]), kind=str(z["kind"]), iters=int(z["iters"]), trees=trees)


def params_to_dict(params):
    return dict(tok.split("=", 1) for tok in params.split())
