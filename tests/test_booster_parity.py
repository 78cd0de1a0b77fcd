"""End-to-end GPU parity through the reference's own C API names: LGBM_DatasetCreateFromMat -> LGBM_BoosterCreate
-> LGBM_BoosterUpdateOneIter on liblgbm_b200.so (device_type=cuda) versus the unmodified reference
lib_lightgbm.so (device_type=cpu) on identical inputs and parameters (SURVEY.md section 8c protocol)."""
import numpy as np
import pytest

from lambdagap_b200 import engine
from oracle import ref_lgbm as R
from tests import helpers as H

pytestmark = pytest.mark.gpu


def _train_both(X, y, params, iters, group=None, weight=None, ds_params="max_bin=255"):
    common = "%s verbose=-1 num_threads=4" % ds_params
    rds = R.RefDataset(X, y, group=group, weight=weight, params=common + " force_col_wise=true")
    rb = R.RefBooster(rds, params + " device_type=cpu force_col_wise=true deterministic=true num_threads=4 verbose=-1")
    ods = engine.Dataset(X, y, group=group, weight=weight, params=common)
    ob = engine.Booster(ods, params + " device_type=cuda num_threads=4 verbose=-1")
    for _ in range(iters):
        rb.update()
        ob.update()
    return rb, ob


def _compare_models(rb, ob, rtol, exact_counts=True):
    rt = R.parse_model_trees(rb.model_string())
    ot = R.parse_model_trees(ob.model_to_string())
    assert len(rt) == len(ot)
    for k, (a, b) in enumerate(zip(ot, rt)):
        assert a["num_leaves"] == b["num_leaves"], "tree %d" % k
        for key in ("split_feature", "left_child", "right_child", "decision_type"):
            np.testing.assert_array_equal(a[key], b[key], err_msg="tree %d %s" % (k, key))
        np.testing.assert_array_equal(a["threshold"], b["threshold"], err_msg="tree %d threshold" % k)
        if exact_counts:
            np.testing.assert_array_equal(a["leaf_count"], b["leaf_count"], err_msg="tree %d leaf_count" % k)
            np.testing.assert_array_equal(a["internal_count"], b["internal_count"])
        np.testing.assert_allclose(a["leaf_value"], b["leaf_value"], rtol=rtol, atol=1e-10, err_msg="tree %d leaf_value" % k)
        np.testing.assert_allclose(a["split_gain"], b["split_gain"], rtol=max(rtol, 1e-5), err_msg="tree %d gain" % k)


@pytest.mark.parametrize("quant", [False, True])
def test_binary_boosting(gpu, quant):
    X, y = H.make_binary(30000, 20, seed=21, nan_frac=0.05, zero_frac=0.2)
    params = "objective=binary num_leaves=31 learning_rate=0.1 min_data_in_leaf=20"
    if quant:
        params += " use_quantized_grad=true stochastic_rounding=false"
    rb, ob = _train_both(X, y, params, 10)
    # binary logloss gradients go through exp(): device and host libm differ in the last ulp, so float mode is the
    # tolerance case (1e-5 as stated by the north star); quantized mode re-discretises to identical integers
    _compare_models(rb, ob, rtol=1e-5)
    np.testing.assert_allclose(ob.train_scores(raw=True), rb.train_scores_raw(), rtol=1e-5, atol=1e-8)


def test_regression_l2(gpu):
    rng = np.random.default_rng(5)
    X = rng.normal(size=(20000, 15)).astype(np.float32)
    y = (X[:, 0] * 2 + np.sin(X[:, 1] * 3) + X[:, 2] * X[:, 3] + 0.1 * rng.normal(size=20000)).astype(np.float32)
    rb, ob = _train_both(X, y, "objective=regression num_leaves=63 learning_rate=0.05 min_data_in_leaf=10 lambda_l2=1.0", 12)
    _compare_models(rb, ob, rtol=1e-6)
    np.testing.assert_allclose(ob.train_scores(raw=True), rb.train_scores_raw(), rtol=1e-6, atol=1e-8)


def test_weighted_and_regularised(gpu):
    X, y = H.make_binary(15000, 10, seed=9)
    w = np.random.default_rng(1).uniform(0.5, 2.0, size=len(y)).astype(np.float32)
    params = ("objective=binary num_leaves=15 lambda_l1=0.5 lambda_l2=2.0 min_gain_to_split=0.1 max_depth=5 "
              "max_delta_step=0.7 path_smooth=2.0 min_sum_hessian_in_leaf=5.0")
    rb, ob = _train_both(X, y, params, 6, weight=w)
    _compare_models(rb, ob, rtol=1e-5)


def test_predict_matches_reference_on_our_model(gpu):
    """Drop-in at the model-file level: the reference loads OUR model string and predicts the same values."""
    X, y = H.make_binary(8000, 8, seed=2, nan_frac=0.1)
    rb, ob = _train_both(X, y, "objective=binary num_leaves=15", 5)
    ours = ob.predict(X[:500].astype(np.float64), raw_score=True)
    ref_loaded = R.RefBooster.from_model_string(ob.model_to_string())
    np.testing.assert_allclose(ours, ref_loaded.predict(X[:500].astype(np.float64), raw_score=True), rtol=1e-12, atol=1e-12)


def test_quant_train_renew_leaf(gpu):
    """GradientDiscretizer::RenewIntGradTreeOutput (gradient_discretizer.cpp:209-260): same integer tree, leaf outputs
    recomputed from the true gradient sums (fp64 sums in a different order -> tolerance)."""
    X, y = H.make_binary(30000, 20, seed=23, nan_frac=0.05, zero_frac=0.2)
    params = ("objective=binary num_leaves=31 learning_rate=0.1 min_data_in_leaf=20 use_quantized_grad=true "
              "stochastic_rounding=false quant_train_renew_leaf=true lambda_l1=0.1 lambda_l2=1.0")
    rb, ob = _train_both(X, y, params, 8)
    _compare_models(rb, ob, rtol=1e-5)
    np.testing.assert_allclose(ob.train_scores(raw=True), rb.train_scores_raw(), rtol=1e-5, atol=1e-8)
    # and it is not a no-op: without renewal the leaf values of the second tree differ visibly (the first tree's
    # gradients take two values that quantize exactly for balanced labels)
    rb2, ob2 = _train_both(X, y, params.replace("quant_train_renew_leaf=true", "quant_train_renew_leaf=false"), 2)
    a = R.parse_model_trees(ob.model_to_string())[1]["leaf_value"]
    b = R.parse_model_trees(ob2.model_to_string())[1]["leaf_value"]
    assert a.shape != b.shape or np.max(np.abs(a - b)) > 1e-4


def test_long_quantized_boosting(gpu):
    """Many iterations: late trees have flat gains and long chains (the shape that stresses the wave schedule);
    quantized mode must stay bit-identical in structure to the reference all the way."""
    X, y = H.make_binary(40000, 12, seed=31, nan_frac=0.03, zero_frac=0.2)
    params = ("objective=binary num_leaves=63 learning_rate=0.2 min_data_in_leaf=20 use_quantized_grad=true "
              "stochastic_rounding=false")
    rb, ob = _train_both(X, y, params, 120)
    _compare_models(rb, ob, rtol=1e-5)


def test_bagging_binary(gpu):
    """Row bagging in the boosting loop (BaggingSampleStrategy, src/boosting/bagging.hpp): same bag every iteration
    as the reference (block-wise generators), tree on the bag, out-of-bag rows scored by walking the tree."""
    X, y = H.make_binary(30000, 15, seed=41, nan_frac=0.05, zero_frac=0.2)
    params = "objective=binary num_leaves=31 learning_rate=0.1 min_data_in_leaf=20 bagging_fraction=0.7 bagging_freq=2 bagging_seed=11"
    rb, ob = _train_both(X, y, params, 9)
    _compare_models(rb, ob, rtol=1e-5)
    np.testing.assert_allclose(ob.train_scores(raw=True), rb.train_scores_raw(), rtol=1e-5, atol=1e-8)


def test_example_lambdarank_conf(gpu):
    """The parameters of the reference's examples/lambdarank/train.conf (config #1 of BASELINE.json) on a synthetic
    ranking set: lambdarank + bagging_fraction=0.9 / bagging_freq=1, min_sum_hessian_in_leaf=5."""
    X, y, group = H.make_ranking(300, 40, seed=8, min_len=5, max_len=80)
    params = ("objective=lambdarank num_leaves=31 learning_rate=0.1 feature_fraction=1.0 bagging_freq=1 bagging_fraction=0.9 "
              "min_data_in_leaf=50 min_sum_hessian_in_leaf=5.0")
    rb, ob = _train_both(X, y, params, 12, group=group)
    _compare_models(rb, ob, rtol=1e-5)


@pytest.mark.parametrize("quant", [False, True])
def test_feature_fraction_by_tree(gpu, quant):
    """ColSampler::ResetByTree (src/treelearner/col_sampler.hpp:77-90): every tree sees the feature subset the reference
    draws (Random::Sample with feature_fraction_seed, one draw at Init and one before every tree); the mask changes
    from tree to tree without a new graph."""
    X, y = H.make_binary(30000, 24, seed=23, nan_frac=0.03, zero_frac=0.2)
    params = "objective=binary num_leaves=31 learning_rate=0.1 min_data_in_leaf=20 feature_fraction=0.6 feature_fraction_seed=11"
    if quant:
        params += " use_quantized_grad=true stochastic_rounding=false"
    rb, ob = _train_both(X, y, params, 8)
    _compare_models(rb, ob, rtol=1e-5)
    used = {int(f) for t in R.parse_model_trees(ob.model_to_string()) for f in t["split_feature"]}
    assert 8 <= len(used) <= 24
    # the derived seed (seed -> feature_fraction_seed, src/io/config.cpp:257-268) and a small fraction
    rb, ob = _train_both(X, y, "objective=binary num_leaves=15 seed=7 feature_fraction=0.2 min_data_in_leaf=20", 5)
    _compare_models(rb, ob, rtol=1e-5)


def test_reset_parameter_and_model_queries(gpu, tmp_path):
    """LGBM_BoosterResetParameter between iterations (a learning-rate schedule, then a different tree size and
    regularisation), feature importances, leaf values and the saved model file, against the reference."""
    X, y = H.make_binary(20000, 12, seed=29)
    params = "objective=binary num_leaves=15 learning_rate=0.2 min_data_in_leaf=20 bagging_fraction=0.8 bagging_freq=1 bagging_seed=5"
    common = "max_bin=255 verbose=-1 num_threads=4"
    rds = R.RefDataset(X, y, params=common + " force_col_wise=true")
    rb = R.RefBooster(rds, params + " device_type=cpu force_col_wise=true deterministic=true num_threads=4 verbose=-1")
    ods = engine.Dataset(X, y, params=common)
    ob = engine.Booster(ods, params + " device_type=cuda num_threads=4 verbose=-1")
    schedule = ["learning_rate=0.2", "learning_rate=0.1", "learning_rate=0.05", "learning_rate=0.05 num_leaves=31 lambda_l2=2.0",
                "learning_rate=0.05 num_leaves=31 lambda_l2=2.0 bagging_fraction=0.7", "min_data_in_leaf=100"]  # (0.5 would switch the reference CPU learner to its subset mode, which aborts on this schedule)
    for step in schedule:
        rb.reset_parameter(step)
        ob.reset_parameter(step)
        rb.update()
        ob.update()
    _compare_models(rb, ob, rtol=1e-5)
    np.testing.assert_array_equal(ob.feature_importance(0), rb.feature_importance(0))
    np.testing.assert_allclose(ob.feature_importance(1), rb.feature_importance(1), rtol=1e-5)
    trees = R.parse_model_trees(ob.model_to_string())
    assert ob.leaf_value(3, 2) == trees[3]["leaf_value"][2]
    path = str(tmp_path / "model.txt")
    ob.save_model(path)
    assert open(path).read() == ob.model_to_string()
    with pytest.raises(engine.LightGBMError, match="Cannot change"):
        ob.reset_parameter("metric=auc")


def test_balanced_bagging(gpu):
    """pos_bagging_fraction / neg_bagging_fraction (BalancedBaggingHelper, src/boosting/bagging.hpp:245-270): the label
    picks the fraction a row's draw is compared with; drawn on the device, trees equal the reference's."""
    X, y = H.make_binary(40000, 14, seed=31, zero_frac=0.1)
    params = ("objective=binary num_leaves=31 learning_rate=0.1 min_data_in_leaf=20 pos_bagging_fraction=0.6 "
              "neg_bagging_fraction=0.9 bagging_freq=1 bagging_seed=9")
    rb, ob = _train_both(X, y, params, 8)
    _compare_models(rb, ob, rtol=1e-5, exact_counts=True)
    np.testing.assert_allclose(ob.train_scores(raw=True), rb.train_scores_raw(), rtol=1e-5, atol=1e-8)


def test_bagging_by_query(gpu):
    """bagging_by_query (src/boosting/bagging.hpp:54-99): whole queries are drawn (generators per block of 1024 queries),
    their documents form the bag.  The first tree equals the reference's.  Later trees cannot: the reference fills only
    the in-bag part of bag_data_indices_ in this mode (:88-98), so its out-of-bag score update (gbdt.cpp:505-513) walks
    stale indices and the training scores of unsampled queries go wrong from the second iteration on.  That is not
    reproduced; instead every iteration's bag is checked against the restated draw (it depends on the generators only)
    and the scores against the model's own predictions for ALL rows."""
    from tests.test_wave_schedule import _reference_bags
    X, y, q = H.make_ranking(3000, 12, seed=33, min_len=2, max_len=60)  # 3000 queries: three generator blocks
    params = ("objective=lambdarank num_leaves=15 learning_rate=0.1 min_data_in_leaf=20 bagging_by_query=true "
              "bagging_fraction=0.7 bagging_freq=1 bagging_seed=4")
    rb, ob = _train_both(X, y, params, 1, group=q)
    _compare_models(rb, ob, rtol=1e-5)
    for _ in range(5):
        ob.update()
    counts = [int(t["leaf_count"].sum()) for t in R.parse_model_trees(ob.model_to_string())]
    want = [int(q[b].sum()) for b in _reference_bags(len(q), 4, 0.7, 6)]
    assert counts == want and len(set(counts)) > 1
    np.testing.assert_allclose(ob.train_scores(raw=True), ob.predict(X, raw_score=True), rtol=1e-9, atol=1e-12)


def test_quantized_stochastic_rounding(gpu):
    """use_quantized_grad with its default stochastic rounding: the per-row random offsets are the reference's host
    mt19937 arrays (GradientDiscretizer::Init, gradient_discretizer.cpp:16-41; one generator pair per thread block, so
    both sides run with the same num_threads) rotated by the per-iteration start (:116); trees stay bit-identical."""
    X, y = H.make_binary(30000, 16, seed=37, nan_frac=0.03, zero_frac=0.2)
    params = "objective=binary num_leaves=31 learning_rate=0.1 min_data_in_leaf=20 use_quantized_grad=true seed=3"
    rb, ob = _train_both(X, y, params, 10)
    _compare_models(rb, ob, rtol=1e-5)
