"""The BASELINE.json configurations at their FULL sizes (synthetic data of each named shape), checked through the
unmodified reference where it finishes in seconds on the box's host cores (quantized mode: bit-exact trees) and through
size-independent properties elsewhere: leaf counts partition the rows, scores equal the model's own prediction,
pairwise lambdas cancel within every query, metrics agree with the restatement."""
import numpy as np
import pytest

from lambdagap_b200 import capi, engine
from oracle import oracle_metrics as M
from oracle import ref_lgbm as R
from tests import helpers as H

pytestmark = pytest.mark.gpu


def _assert_same_trees(ours, ref, leaf_rtol, empty_bin_ties=0.0):
    """empty_bin_ties: fraction of a tree's thresholds that may differ from the reference's although the split is the
    same one (float mode only).  When a node holds no row in a bin, the two thresholds on either side of that bin
    separate its rows identically: same children, same counts, same sums.  With exact integer sums that is an exact
    tie here (resolved like the reference's sequential scan resolves ties); in the reference the empty most-frequent
    bin carries the rounding residue of `leaf total - sum of the other bins` (Dataset::FixHistogram), so its two gains
    differ in the last bits and either may win.  Everything that depends on the rows -- feature, default direction,
    children, every leaf / internal count, leaf values -- is still asserted equal."""
    ot, rt = R.parse_model_trees(ours), R.parse_model_trees(ref)
    assert len(ot) == len(rt) > 0
    for k, (a, b) in enumerate(zip(ot, rt)):
        assert a["num_leaves"] == b["num_leaves"], "tree %d" % k
        for key in ("split_feature", "left_child", "right_child", "decision_type", "leaf_count", "internal_count"):
            np.testing.assert_array_equal(a[key], b[key], err_msg="tree %d %s" % (k, key))
        if empty_bin_ties > 0.0:
            differ = np.mean(a["threshold"] != b["threshold"])
            assert differ <= empty_bin_ties, "tree %d: %.1f %% of the thresholds differ" % (k, 100 * differ)
        else:
            np.testing.assert_array_equal(a["threshold"], b["threshold"], err_msg="tree %d threshold" % k)
        np.testing.assert_allclose(a["leaf_value"], b["leaf_value"], rtol=leaf_rtol, atol=1e-12, err_msg="tree %d" % k)
    return ot


def test_higgs_shape_full_size_quantized_is_bit_exact(gpu):
    """Config 2: 10.5 M rows x 28 dense features, max_bin=255, num_leaves=255.  Quantized gradients make every
    histogram an integer sum, so the trees must equal the reference CPU learner's exactly at full size too."""
    n, f, iters = 10_500_000, 28, 3
    rng = np.random.default_rng(42)
    X = rng.standard_normal(size=(n, f), dtype=np.float32)
    X[:, 1] = np.exp(X[:, 1])
    X[:, 5] = np.exp(2 * X[:, 5])
    w = rng.normal(size=f)
    y = (X @ w.astype(np.float32) + rng.standard_normal(size=n, dtype=np.float32) > 0).astype(np.float32)
    params = ("objective=binary num_leaves=255 learning_rate=0.1 min_data_in_leaf=20 use_quantized_grad=true "
              "stochastic_rounding=false metric=auc,binary_logloss verbose=-1")
    ods = engine.Dataset(X, y, params="max_bin=255 verbose=-1")
    ob = engine.Booster(ods, params + " device_type=cuda")
    for _ in range(iters):
        ob.update()
    ours = ob.model_to_string()
    rds = R.RefDataset(X, y, params="max_bin=255 verbose=-1 force_col_wise=true")
    rb = R.RefBooster(rds, params + " device_type=cpu force_col_wise=true deterministic=true")
    for _ in range(iters):
        rb.update()
    trees = _assert_same_trees(ours, rb.model_string(), leaf_rtol=1e-9)
    for t in trees:
        assert t["num_leaves"] == 255 and int(t["leaf_count"].sum()) == n
    # the device-resident scores are the model's own predictions, and the device metrics agree with the reference's
    s = ob.train_scores(raw=True)
    sample = rng.choice(n, size=200_000, replace=False)
    np.testing.assert_allclose(s[sample], ob.predict(X[sample], raw_score=True), rtol=1e-9, atol=1e-12)
    ours_eval, ref_eval = ob.eval(0), rb.eval(0)
    assert ob.eval_names() == rb.eval_names() == ["auc", "binary_logloss"]
    np.testing.assert_allclose(ours_eval, ref_eval, rtol=1e-7)
    assert ours_eval[0] == M.auc(s, y)  # AUC of 10.5 M scores: exact integer arithmetic


def test_epsilon_shape_full_size_quantized_is_bit_exact(gpu):
    """Config 4: 400 K rows x 2 000 dense features (63 feature tiles, the split-scan stress), quantized gradients."""
    n, f, iters = 400_000, 2000, 2
    rng = np.random.default_rng(44)
    X = rng.standard_normal(size=(n, f), dtype=np.float32)
    X /= np.linalg.norm(X, axis=1, keepdims=True)
    w = np.zeros(f, dtype=np.float32)
    w[rng.choice(f, 50, replace=False)] = rng.normal(size=50)
    y = (X @ w + 0.01 * rng.standard_normal(size=n, dtype=np.float32) > 0).astype(np.float32)
    params = ("objective=binary num_leaves=255 learning_rate=0.1 use_quantized_grad=true num_grad_quant_bins=4 "
              "stochastic_rounding=false verbose=-1")
    ods = engine.Dataset(X, y, params="max_bin=255 verbose=-1")
    ob = engine.Booster(ods, params + " device_type=cuda")
    for _ in range(iters):
        ob.update()
    rds = R.RefDataset(X, y, params="max_bin=255 verbose=-1 force_col_wise=true")
    rb = R.RefBooster(rds, params + " device_type=cpu force_col_wise=true deterministic=true")
    for _ in range(iters):
        rb.update()
    trees = _assert_same_trees(ob.model_to_string(), rb.model_string(), leaf_rtol=1e-9)
    for t in trees:
        assert int(t["leaf_count"].sum()) == n


def test_mslr_shape_full_size_properties(gpu):
    """Config 3: ~3.77 M documents x 136 features in 31 531 queries, lambdarank, num_leaves=255.  Float mode at this size
    is a tolerance case, so the checks are size-independent properties: every tree's leaves partition the rows, the
    scores are the model's predictions, the lambdas of every query cancel (each pair adds +l to one document and -l to
    the other; the per-query rescaling keeps that), and NDCG@10 on the device equals the restatement."""
    X, y, group = H.make_ranking(31531, 136, seed=43, min_len=1, max_len=1251, zero_frac=0.3, median_len=95)
    n = X.shape[0]
    assert 3_000_000 < n < 4_500_000
    qb = np.concatenate(([0], np.cumsum(group))).astype(np.int64)
    params = ("objective=lambdarank lambdarank_target=lambdagap-x lambdarank_truncation_level=30 num_leaves=255 "
              "learning_rate=0.1 metric=ndcg eval_at=10 verbose=-1")
    ods = engine.Dataset(X, y, group=group, params="max_bin=255 verbose=-1")
    ob = engine.Booster(ods, params + " device_type=cuda")
    ndcg = []
    for _ in range(3):
        ob.update()
        ndcg.append(ob.eval(0)[0])
    assert ndcg[-1] > ndcg[0]  # training NDCG@10 improves
    trees = R.parse_model_trees(ob.model_to_string())
    for t in trees:
        assert t["num_leaves"] == 255 and int(t["leaf_count"].sum()) == n
    s = ob.train_scores(raw=True)
    rng = np.random.default_rng(1)
    sample = rng.choice(n, size=100_000, replace=False)
    np.testing.assert_allclose(s[sample], ob.predict(X[sample], raw_score=True), rtol=1e-9, atol=1e-12)
    np.testing.assert_allclose(ndcg[-1], M.ndcg(s, y, qb, [10])[0], rtol=1e-12)
    # gradients of the current scores through the C ABI objective: per-query sums vanish
    obj_cfg = capi.ObjectiveConfig()
    obj_cfg.kind = 2
    obj_cfg.lambdarank_target = 13  # lambdagap-x
    obj_cfg.lambdarank_truncation_level = 30
    obj_cfg.lambdarank_norm = 1
    obj_cfg.sigmoid = 1.0
    obj_cfg.lambdagap_weight = 1.0
    gains = np.array([float((1 << i) - 1) for i in range(31)])
    import ctypes
    obj_cfg.num_label_gain = len(gains)
    obj_cfg.label_gain = gains.ctypes.data_as(ctypes.POINTER(ctypes.c_double))
    handle = ctypes.c_void_p()
    qb32 = qb.astype(np.int32)
    capi.check(capi.lib().gbl_objective_create(ctypes.byref(obj_cfg), ctypes.c_int64(n), y.ctypes.data_as(ctypes.c_void_p), None,
                                               qb32.ctypes.data_as(ctypes.c_void_p), ctypes.c_int32(len(group)),
                                               ctypes.c_int32(0), ctypes.byref(handle)))
    sd = capi.DeviceBuffer.from_numpy(s)
    gd, hd = capi.DeviceBuffer(4 * n), capi.DeviceBuffer(4 * n)
    capi.check(capi.lib().gbl_objective_get_gradients(handle, ctypes.c_void_p(sd.data_ptr()), ctypes.c_void_p(gd.data_ptr()),
                                                      ctypes.c_void_p(hd.data_ptr()), None))
    capi.synchronize()
    g = gd.download(np.float32, n).astype(np.float64)
    h = hd.download(np.float32, n)
    capi.check(capi.lib().gbl_objective_free(handle))
    assert np.all(np.isfinite(g)) and np.all(h >= 0)
    per_query = np.add.reduceat(g, qb[:-1])
    scale = np.add.reduceat(np.abs(g), qb[:-1]) + 1e-30
    assert np.max(np.abs(per_query) / scale) < 1e-4  # fp32 accumulation of +l / -l in different orders


def test_higgs_shape_full_size_float_matches_reference(gpu):
    """The configuration bench.py's headline is quoted on, at its own size and in its own mode: bench.py's Higgs-shape
    matrix (10.5 M x 28, seed 42), float gradients, num_leaves=255, 3 iterations against the unmodified reference CPU
    learner.  SURVEY.md 8(c): identical chosen splits (feature, threshold, default direction, children, leaf counts),
    leaf values within 1e-5."""
    import bench
    cfg = dict(bench.HIGGS)
    X, y = bench.gen_block(cfg, 0, cfg["rows"])
    n, iters = X.shape[0], 3
    params = cfg["params"] + " verbose=-1"
    ods = engine.Dataset(X, y, params="max_bin=255 verbose=-1")
    ob = engine.Booster(ods, params + " device_type=cuda")
    for _ in range(iters):
        ob.update()
    ours = ob.model_to_string()
    s = ob.train_scores(raw=True)
    ob.free()
    ods.free()
    rds = R.RefDataset(X, y, params="max_bin=255 verbose=-1 force_col_wise=true")
    rb = R.RefBooster(rds, params + " device_type=cpu force_col_wise=true deterministic=true")
    for _ in range(iters):
        rb.update()
    trees = _assert_same_trees(ours, rb.model_string(), leaf_rtol=1e-5)
    for t in trees:
        assert t["num_leaves"] == 255 and int(t["leaf_count"].sum()) == n
    np.testing.assert_allclose(s, rb.train_scores_raw(), rtol=1e-5, atol=1e-9)


@pytest.mark.parametrize("target", ["ndcg", "lambdagap-x"])
def test_mslr_shape_full_size_matches_reference(gpu, target):
    """Config 3 at full size against the reference itself: ~3.77 M documents x 136 features in 31 531 queries,
    lambdarank with num_leaves=255, 2 iterations.  The lambda gradients replay the CPU's fp32 pair order, so the chosen
    splits are expected to be identical; leaf values and scores within 1e-5."""
    X, y, group = H.make_ranking(31531, 136, seed=43, min_len=1, max_len=1251, zero_frac=0.3, median_len=95)
    n, iters = X.shape[0], 2
    params = ("objective=lambdarank lambdarank_target=%s lambdarank_truncation_level=30 num_leaves=255 learning_rate=0.1 "
              "min_data_in_leaf=20 verbose=-1" % target)
    ods = engine.Dataset(X, y, group=group, params="max_bin=255 verbose=-1")
    ob = engine.Booster(ods, params + " device_type=cuda")
    for _ in range(iters):
        ob.update()
    ours = ob.model_to_string()
    s = ob.train_scores(raw=True)
    ob.free()
    ods.free()
    rds = R.RefDataset(X, y, group=group, params="max_bin=255 verbose=-1 force_col_wise=true")
    rb = R.RefBooster(rds, params + " device_type=cpu force_col_wise=true deterministic=true")
    for _ in range(iters):
        rb.update()
    # ~30 % of the columns are zero-inflated: deep nodes without a single zero row produce the empty-bin tie described
    # in _assert_same_trees (seen: 2 of 254 thresholds of one tree, +1e-35 here against -1e-35 in the reference)
    trees = _assert_same_trees(ours, rb.model_string(), leaf_rtol=1e-5, empty_bin_ties=0.02)
    for t in trees:
        assert int(t["leaf_count"].sum()) == n
    np.testing.assert_allclose(s, rb.train_scores_raw(), rtol=1e-5, atol=1e-9)
