"""The drop-in the north star asks for, compiled and run: integration/_ref/lib_lightgbm.so is the REFERENCE library
(its GBDT, Config, Dataset, objectives, sampling strategies, Tree, model I/O, C API -- built from /root/reference by
integration/Makefile) with its TreeLearner factory re-pointed so that `device_type=cuda` creates the B200 learner
(integration/b200_tree_learner.cpp, a subclass of the reference's own TreeLearner, include/LightGBM/tree_learner.h:27-115).
Here the reference's own LGBM_BoosterCreate(... device_type=cuda) in that library must grow the trees that the
unmodified reference grows with device_type=cpu (oracle/_ref) on the same inputs: bit-exact with quantized gradients,
identical splits and 1e-5 on values in float mode.  The drop-in runs in a process of its own (tests/dropin_runner.py)."""
import os
import subprocess
import sys

import numpy as np
import pytest

from lambdagap_b200 import capi
from oracle import ref_lgbm as R
from tests import helpers as H

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DROPIN = os.path.join(ROOT, "integration", "_ref", "lib_lightgbm.so")


def _run_dropin(tmp_path, X, y, params, iters, group=None, dataset_params="max_bin=255 verbose=-1", env=None, weight=None,
                reset=None, rollback_after=None):
    if not os.path.exists(DROPIN):
        pytest.fail("integration/_ref/lib_lightgbm.so missing: run __graft_entry__.build() where /root/reference exists")
    src, dst = str(tmp_path / "in.npz"), str(tmp_path / "out.npz")
    arrays = dict(X=X, y=y, params=np.array(params), dataset_params=np.array(dataset_params), iters=np.array(iters))
    if group is not None:
        arrays["group"] = group
    if weight is not None:
        arrays["weight"] = weight
    if reset is not None:
        arrays["reset_after"] = np.array(reset[0])
        arrays["reset_params"] = np.array(reset[1])
    if rollback_after is not None:
        arrays["rollback_after"] = np.array(rollback_after)
    np.savez(src, **arrays)
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "dropin_runner.py"), src, dst], capture_output=True, text=True,
                       timeout=600, env=dict(os.environ, **(env or {})))
    assert r.returncode == 0, "drop-in run failed:\n" + r.stdout[-2000:] + r.stderr[-4000:]
    z = np.load(dst)
    return str(z["model"]), z["scores"]


def _run_reference(X, y, params, iters, group=None, dataset_params="max_bin=255 verbose=-1", weight=None):
    ds = R.RefDataset(X, y, group=group, weight=weight, params=dataset_params + " force_col_wise=true")
    b = R.RefBooster(ds, params + " device_type=cpu force_col_wise=true deterministic=true")
    for _ in range(iters):
        b.update()
    return b.model_string(), b.train_scores()


def _assert_models(ours, ref, leaf_rtol, count_exact=True):
    ot, rt = R.parse_model_trees(ours), R.parse_model_trees(ref)
    assert len(ot) == len(rt) > 0
    for k, (a, b) in enumerate(zip(ot, rt)):
        assert a["num_leaves"] == b["num_leaves"], "tree %d" % k
        keys = ["split_feature", "left_child", "right_child", "decision_type"]
        if count_exact:
            keys += ["leaf_count", "internal_count"]
        for key in keys:
            np.testing.assert_array_equal(a[key], b[key], err_msg="tree %d %s" % (k, key))
        np.testing.assert_array_equal(a["threshold"], b["threshold"], err_msg="tree %d threshold" % k)
        np.testing.assert_allclose(a["leaf_value"], b["leaf_value"], rtol=leaf_rtol, atol=1e-12, err_msg="tree %d" % k)


CASES = {
    "binary_quantized": ("objective=binary num_leaves=63 learning_rate=0.1 min_data_in_leaf=20 use_quantized_grad=true "
                         "stochastic_rounding=false verbose=-1 num_threads=4", 1e-9),
    "binary_quantized_stochastic": ("objective=binary num_leaves=31 learning_rate=0.1 use_quantized_grad=true "
                                    "stochastic_rounding=true seed=7 verbose=-1 num_threads=4", 1e-9),
    "binary_float": ("objective=binary num_leaves=63 learning_rate=0.1 min_data_in_leaf=20 verbose=-1 num_threads=4", 1e-5),
    "regression_l2": ("objective=regression num_leaves=63 learning_rate=0.1 lambda_l2=0.5 verbose=-1 num_threads=4", 1e-5),
    "regression_l1_renew": ("objective=regression_l1 num_leaves=31 learning_rate=0.1 verbose=-1 num_threads=4", 1e-5),
    "bagging": ("objective=binary num_leaves=31 learning_rate=0.1 bagging_fraction=0.7 bagging_freq=1 bagging_seed=5 "
                "verbose=-1 num_threads=4", 1e-5),
    # GOSS starts sampling at iteration 1 / learning_rate (goss.hpp:32-34): learning_rate=0.5 and six iterations, so that
    # four of them are sampled.  top_rate + other_rate <= 0.5 makes the reference train on a row-subset copy and hand
    # Train the bag's gradients packed in bag order (goss.hpp:101-108, gbdt.cpp:392-400); above 0.5 it passes the index list
    "goss": ("objective=binary num_leaves=31 learning_rate=0.5 data_sample_strategy=goss top_rate=0.2 other_rate=0.1 "
             "verbose=-1 num_threads=4", 1e-5, 6),
    "goss_index_list": ("objective=binary num_leaves=31 learning_rate=0.5 data_sample_strategy=goss top_rate=0.4 "
                        "other_rate=0.3 verbose=-1 num_threads=4", 1e-5, 6),
    # monotone constraints, method basic: +1 on column 0, -1 on column 1, +1 on column 4 (20 columns)
    "monotone_float": ("objective=binary num_leaves=63 learning_rate=0.1 verbose=-1 num_threads=4 "
                       "monotone_constraints=1,-1,0,0,1,0,0,0,0,0,0,0,0,0,0,0,0,0,0,0", 1e-5),
    "monotone_quantized_penalty": ("objective=binary num_leaves=63 learning_rate=0.1 verbose=-1 num_threads=4 "
                                   "use_quantized_grad=true stochastic_rounding=false monotone_penalty=1.5 lambda_l2=1.0 "
                                   "monotone_constraints=1,-1,0,0,1,0,0,0,0,0,0,0,0,0,0,0,0,0,0,0", 1e-9),
    # L1 / max_delta_step clip many candidate outputs to the same value, which multiplies exact gain ties between the two
    # scan directions; quantized gradients make both sides resolve them on identical integer sums
    "monotone_regression_l1l2": ("objective=regression num_leaves=31 learning_rate=0.1 verbose=-1 num_threads=4 lambda_l1=0.1 "
                                 "lambda_l2=0.5 max_delta_step=0.8 use_quantized_grad=true stochastic_rounding=false "
                                 "monotone_constraints=-1,0,1,0,0,0,0,0,0,0,0,0,0,0,0,0,0,0,0,0", 1e-9),
    "feature_fraction": ("objective=binary num_leaves=31 learning_rate=0.1 feature_fraction=0.6 feature_fraction_seed=9 "
                         "verbose=-1 num_threads=4", 1e-5),
    # the reference's other boosting modes drive the same TreeLearner: DART drops and re-weights trees on the host score
    # array between iterations (the device objective re-reads it every iteration), random forest trains every tree on
    # the initial gradients with a fresh bag (host objective, as in the reference's own CUDA build)
    "dart": ("objective=binary boosting=dart drop_rate=0.5 skip_drop=0.2 drop_seed=4 num_leaves=31 learning_rate=0.1 "
             "verbose=-1 num_threads=4", 1e-5, 6),
    "rf": ("objective=binary boosting=rf bagging_fraction=0.7 bagging_freq=1 bagging_seed=3 feature_fraction=0.8 "
           "num_leaves=31 verbose=-1 num_threads=4", 1e-5, 4),
}


@pytest.mark.parametrize("name", sorted(CASES))
def test_dropin_equals_reference_cpu(gpu, tmp_path, name):
    params, rtol = CASES[name][:2]
    X, y = H.make_binary(120_000, 20, seed=3, nan_frac=0.02, zero_frac=0.2)
    if "regression" in name:
        y = (X[:, 0] * 0.5 + np.nan_to_num(X[:, 2]) + 0.1 * np.random.default_rng(1).normal(size=len(y))).astype(np.float32)
    iters = CASES[name][2] if len(CASES[name]) > 2 else 4
    ours, s_ours = _run_dropin(tmp_path, X, y, params + " device_type=cuda", iters)
    ref, s_ref = _run_reference(X, y, params, iters)
    _assert_models(ours, ref, rtol)
    np.testing.assert_allclose(s_ours, s_ref, rtol=max(rtol, 1e-7), atol=1e-9)


@pytest.mark.parametrize("route", ["host_objective", "device_objective_fills_host"])
@pytest.mark.parametrize("name", ["binary_quantized", "regression_l2"])
def test_dropin_gradient_routes(gpu, tmp_path, name, route):
    """By default the drop-in computes regression / binary / lambdarank gradients on the device and trains from the device
    copies (integration/b200_objective.cpp) -- that is what every other test in this file runs.  The two other routes must
    give the same trees: the reference's host objective + upload, and the device objective that also fills GBDT's vectors."""
    params, rtol = CASES[name][:2]
    X, y = H.make_binary(120_000, 20, seed=3, nan_frac=0.02, zero_frac=0.2)
    if "regression" in name:
        y = (X[:, 0] * 0.5 + np.nan_to_num(X[:, 2]) + 0.1 * np.random.default_rng(1).normal(size=len(y))).astype(np.float32)
    env = {"GBL_DROPIN_DEVICE_OBJECTIVE": "0"} if route == "host_objective" else {"GBL_DROPIN_FILL_HOST_GRADIENTS": "1"}
    ours, s_ours = _run_dropin(tmp_path, X, y, params + " device_type=cuda", 4, env=env)
    ref, s_ref = _run_reference(X, y, params, 4)
    _assert_models(ours, ref, rtol)
    np.testing.assert_allclose(s_ours, s_ref, rtol=max(rtol, 1e-7), atol=1e-9)


@pytest.mark.parametrize("case", ["binary_unbalance", "binary_scale_pos_weight", "regression", "lambdarank"])
def test_dropin_device_objective_with_weights(gpu, tmp_path, case):
    """Row weights, is_unbalance / scale_pos_weight label weights and a non-default sigmoid through the device objectives
    of the drop-in (B200Objective: binary_objective.hpp:73-136, regression_objective.hpp:112-127, rank_objective.hpp:98-105)."""
    rng = np.random.default_rng(17)
    group = None
    if case == "lambdarank":
        X, y, group = H.make_ranking(1200, 16, seed=5)
        params = ("objective=lambdarank lambdarank_truncation_level=8 sigmoid=1.5 num_leaves=31 learning_rate=0.1 "
                  "min_data_in_leaf=20 verbose=-1 num_threads=4")
    else:
        X, y = H.make_binary(80_000, 12, seed=8, nan_frac=0.02)
        if case == "regression":
            y = (X[:, 0] - 0.5 * np.nan_to_num(X[:, 3]) + 0.1 * rng.normal(size=len(y))).astype(np.float32)
            params = "objective=regression num_leaves=31 learning_rate=0.1 verbose=-1 num_threads=4"
        elif case == "binary_unbalance":
            y = (y * (rng.random(len(y)) < 0.3)).astype(np.float32)  # about 15 % positives
            params = "objective=binary is_unbalance=true sigmoid=0.8 num_leaves=31 learning_rate=0.1 verbose=-1 num_threads=4"
        else:
            params = "objective=binary scale_pos_weight=2.5 num_leaves=31 learning_rate=0.1 verbose=-1 num_threads=4"
    w = rng.uniform(0.5, 2.0, size=len(y)).astype(np.float32)
    ours, s_ours = _run_dropin(tmp_path, X, y, params + " device_type=cuda", 4, group=group, weight=w)
    ref, s_ref = _run_reference(X, y, params, 4, group=group, weight=w)
    _assert_models(ours, ref, 1e-5)
    np.testing.assert_allclose(s_ours, s_ref, rtol=1e-5, atol=1e-9)


def test_dropin_reset_parameter_mid_training(gpu, tmp_path):
    """LGBM_BoosterResetParameter between iterations (a learning-rate schedule, then a change the device learner reads:
    min_data_in_leaf and lambda_l2): TreeLearner::ResetConfig on the B200 learner, the device objective kept."""
    X, y = H.make_binary(100_000, 16, seed=12, nan_frac=0.02)
    params = "objective=binary num_leaves=31 learning_rate=0.2 verbose=-1 num_threads=4"
    new = "learning_rate=0.05 min_data_in_leaf=200 lambda_l2=3.0"
    ours, s_ours = _run_dropin(tmp_path, X, y, params + " device_type=cuda", 5, reset=(2, new))
    ds = R.RefDataset(X, y, params="max_bin=255 verbose=-1 force_col_wise=true")
    b = R.RefBooster(ds, params + " device_type=cpu force_col_wise=true deterministic=true")
    for it in range(5):
        if it == 2:
            b.reset_parameter(new)
        b.update()
    _assert_models(ours, b.model_string(), 1e-5)
    np.testing.assert_allclose(s_ours, b.train_scores(), rtol=1e-5, atol=1e-9)


@pytest.mark.parametrize("kind", ["binary", "lambdarank"])
def test_dropin_validation_metrics(gpu, tmp_path, kind):
    """A validation set and metrics through the reference's own LGBM_BoosterAddValidData / LGBM_BoosterGetEval in the
    drop-in library: training and validation metric values after 5 iterations equal the reference's."""
    if kind == "binary":
        X, y = H.make_binary(60_000, 12, seed=31, nan_frac=0.02)
        VX, vy = H.make_binary(20_000, 12, seed=32, nan_frac=0.02)
        group = vgroup = None
        params = "objective=binary metric=auc,binary_logloss num_leaves=31 learning_rate=0.1 verbose=-1 num_threads=4"
    else:
        X, y, group = H.make_ranking(900, 14, seed=33)
        VX, vy, vgroup = H.make_ranking(300, 14, seed=34)
        params = ("objective=lambdarank lambdarank_target=lambdagap-s-plus metric=ndcg eval_at=1,5 num_leaves=15 "
                  "learning_rate=0.1 min_data_in_leaf=20 verbose=-1 num_threads=4")
    src, dst = str(tmp_path / "in.npz"), str(tmp_path / "out.npz")
    arrays = dict(X=X, y=y, params=np.array(params + " device_type=cuda"), dataset_params=np.array("max_bin=255 verbose=-1"),
                  iters=np.array(5), valid_X=VX, valid_y=vy)
    if group is not None:
        arrays["group"] = group
        arrays["valid_group"] = vgroup
    np.savez(src, **arrays)
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "dropin_runner.py"), src, dst], capture_output=True, text=True,
                       timeout=600)
    assert r.returncode == 0, "drop-in run failed:\n" + r.stdout[-2000:] + r.stderr[-4000:]
    z = np.load(dst)
    ds = R.RefDataset(X, y, group=group, params="max_bin=255 verbose=-1 force_col_wise=true")
    vds = R.RefDataset(VX, vy, group=vgroup, params="max_bin=255 verbose=-1", reference=ds)
    b = R.RefBooster(ds, params + " device_type=cpu force_col_wise=true deterministic=true")
    b.add_valid(vds)
    for _ in range(5):
        b.update()
    _assert_models(str(z["model"]), b.model_string(), 1e-5)
    np.testing.assert_allclose(z["train_eval"], b.eval(0), rtol=1e-6, atol=1e-9)
    np.testing.assert_allclose(z["valid_eval"], b.eval(1), rtol=1e-6, atol=1e-9)
    assert len(z["valid_eval"]) == 2


def test_dropin_rollback_one_iter(gpu, tmp_path):
    """LGBM_BoosterRollbackOneIter after the third iteration: the host score array loses that tree (GBDT::RollbackOneIter,
    gbdt.cpp:447-466) and the next gradients -- computed on the device from the uploaded scores -- must see it."""
    X, y = H.make_binary(80_000, 14, seed=41, nan_frac=0.02)
    params = "objective=binary num_leaves=31 learning_rate=0.2 verbose=-1 num_threads=4"
    ours, s_ours = _run_dropin(tmp_path, X, y, params + " device_type=cuda", 5, rollback_after=2)
    ds = R.RefDataset(X, y, params="max_bin=255 verbose=-1 force_col_wise=true")
    b = R.RefBooster(ds, params + " device_type=cpu force_col_wise=true deterministic=true")
    for it in range(5):
        b.update()
        if it == 2:
            R._check(R.lib().LGBM_BoosterRollbackOneIter(b.handle))
    _assert_models(ours, b.model_string(), 1e-5)
    np.testing.assert_allclose(s_ours, b.train_scores(), rtol=1e-5, atol=1e-9)


@pytest.mark.parametrize("route", ["device_objective", "device_objective_weighted", "host_objective"])
def test_dropin_multiclass(gpu, tmp_path, route):
    """num_class trees per iteration: GBDT calls TreeLearner::Train once per class with offset gradient pointers.  The
    softmax gradients come from the device objective (class-major blocks, one Train per block from the device copy) or,
    with GBL_DROPIN_DEVICE_OBJECTIVE=0, from the reference's host objective."""
    rng = np.random.default_rng(5)
    X = rng.normal(size=(60_000, 12)).astype(np.float32)
    y = (np.argmax(X[:, :4] + 0.3 * rng.normal(size=(len(X), 4)), axis=1)).astype(np.float32)
    w = rng.uniform(0.5, 2.0, size=len(y)).astype(np.float32) if route.endswith("weighted") else None
    params = "objective=multiclass num_class=4 num_leaves=31 learning_rate=0.1 verbose=-1 num_threads=4"
    env = {"GBL_DROPIN_DEVICE_OBJECTIVE": "0"} if route == "host_objective" else None
    ours, s_ours = _run_dropin(tmp_path, X, y, params + " device_type=cuda", 3, weight=w, env=env)
    ref, s_ref = _run_reference(X, y, params, 3, weight=w)
    _assert_models(ours, ref, 1e-5)
    np.testing.assert_allclose(s_ours, s_ref, rtol=1e-5, atol=1e-9)


@pytest.mark.parametrize("target", ["ndcg", "lambdagap-x"])
def test_dropin_lambdarank(gpu, tmp_path, target):
    """The fork's lambdarank targets: device gradients (integration/b200_objective.cpp) and the B200 learner behind the
    reference's GBDT, against the reference's own objective and CPU learner."""
    X, y, group = H.make_ranking(1500, 24, seed=11)
    params = ("objective=lambdarank lambdarank_target=%s lambdarank_truncation_level=10 num_leaves=31 learning_rate=0.1 "
              "min_data_in_leaf=20 verbose=-1 num_threads=4" % target)
    ours, s_ours = _run_dropin(tmp_path, X, y, params + " device_type=cuda", 4, group=group)
    ref, s_ref = _run_reference(X, y, params, 4, group=group)
    _assert_models(ours, ref, 1e-5)
    np.testing.assert_allclose(s_ours, s_ref, rtol=1e-5, atol=1e-9)


def test_dropin_num_gpu_shards_rows_inside_one_booster(gpu, tmp_path):
    """SURVEY 8(e) route B: device_type=cuda num_gpu=2 in ONE Booster (the reference refuses num_gpu != 1,
    src/treelearner/tree_learner.cpp:46-50).  Integer histogram sums make the 2-GPU trees equal the CPU's."""
    if capi.device_count() < 2:
        pytest.skip("needs two GPUs (run with gpurun --gpus 2)")
    X, y = H.make_binary(200_000, 20, seed=4)
    for extra, rtol in ((" use_quantized_grad=true stochastic_rounding=false", 1e-9), ("", 1e-5)):
        params = "objective=binary num_leaves=63 learning_rate=0.1 verbose=-1 num_threads=4" + extra
        ours, s_ours = _run_dropin(tmp_path, X, y, params + " device_type=cuda num_gpu=2", 4)
        ref, s_ref = _run_reference(X, y, params, 4)
        _assert_models(ours, ref, rtol)
        np.testing.assert_allclose(s_ours, s_ref, rtol=max(rtol, 1e-7), atol=1e-9)


def test_dropin_refuses_what_the_learner_does_not_implement(gpu, tmp_path):
    X, y = H.make_binary(5_000, 6, seed=1)
    src, dst = str(tmp_path / "in.npz"), str(tmp_path / "out.npz")
    np.savez(src, X=X, y=y, params=np.array("objective=binary device_type=cuda extra_trees=true verbose=-1"),
             dataset_params=np.array("max_bin=255 verbose=-1"), iters=np.array(1))
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tests", "dropin_runner.py"), src, dst], capture_output=True, text=True,
                       timeout=300)
    assert r.returncode != 0 and "extra_trees is not supported" in (r.stdout + r.stderr)


@pytest.mark.parametrize("mode", ["float", "quantized", "small_groups"])
def test_dropin_categorical(gpu, tmp_path, mode):
    """Categorical splits (FindBestThresholdCategoricalInner / ...IntInner, feature_histogram.cpp:142-700): one-vs-rest for a
    3-category column (num_bin <= max_cat_to_onehot), the sorted many-vs-many search for 40- and 200-category columns, next
    to numerical columns.  The split's bin set travels as a bitset; Tree::SplitCategorical writes it into the model."""
    rng = np.random.default_rng(21)
    n = 150_000
    X = rng.normal(size=(n, 10)).astype(np.float32)
    c0 = rng.integers(0, 3, size=n)
    c1 = rng.integers(0, 40, size=n)
    c2 = (rng.zipf(1.3, size=n) % 200)
    e1 = rng.normal(size=40)
    e2 = rng.normal(size=200)
    X[:, 0], X[:, 1], X[:, 2] = c0, c1, c2
    logit = 0.8 * (c0 == 1) + e1[c1] + 0.7 * e2[c2] + X[:, 3] - 0.5 * X[:, 4] + 0.5 * rng.normal(size=n)
    y = (logit > np.median(logit)).astype(np.float32)
    dsp = "max_bin=255 verbose=-1 categorical_feature=0,1,2"
    params = "objective=binary num_leaves=63 learning_rate=0.1 verbose=-1 num_threads=4 " + dsp
    rtol = 1e-5
    if mode == "quantized":
        params += " use_quantized_grad=true stochastic_rounding=false"
        rtol = 1e-9
    if mode == "small_groups":
        params += " min_data_per_group=20 cat_smooth=2 cat_l2=1 max_cat_threshold=16 min_data_in_leaf=10"
    ours, s_ours = _run_dropin(tmp_path, X, y, params + " device_type=cuda", 4, dataset_params=dsp)
    ref, s_ref = _run_reference(X, y, params, 4, dataset_params=dsp)
    assert "cat_threshold" in ref and "num_cat=0" not in ref.split("Tree=0")[1].split("Tree=1")[0], "the case must contain categorical splits"
    ot, rt = R.parse_model_trees(ours), R.parse_model_trees(ref)
    _assert_models(ours, ref, rtol)
    # the categorical thresholds themselves: the cat_boundaries / cat_threshold lines of every tree
    def cat_lines(text):
        return [ln for ln in text.split("\n") if ln.startswith(("cat_boundaries=", "cat_threshold=", "num_cat="))]
    assert cat_lines(ours) == cat_lines(ref)
    np.testing.assert_allclose(s_ours, s_ref, rtol=max(rtol, 1e-7), atol=1e-9)


def test_dropin_num_gpu_with_categorical_and_monotone(gpu, tmp_path):
    """Route B with the round-2 split types: two GPUs inside one Booster, a categorical column and monotone constraints,
    quantized gradients (integer histogram sums make the sharded result equal the CPU's bit for bit)."""
    if capi.device_count() < 2:
        pytest.skip("needs two GPUs (run with gpurun --gpus 2)")
    rng = np.random.default_rng(8)
    n = 160_000
    X = rng.normal(size=(n, 8)).astype(np.float32)
    c = rng.integers(0, 30, size=n)
    X[:, 0] = c
    e = rng.normal(size=30)
    logit = e[c] + X[:, 1] - 0.7 * X[:, 2] + 0.4 * rng.normal(size=n)
    y = (logit > np.median(logit)).astype(np.float32)
    dsp = "max_bin=255 verbose=-1 categorical_feature=0"
    params = ("objective=binary num_leaves=63 learning_rate=0.1 verbose=-1 num_threads=4 use_quantized_grad=true "
              "stochastic_rounding=false monotone_constraints=0,1,-1,0,0,0,0,0 " + dsp)
    ours, s_ours = _run_dropin(tmp_path, X, y, params + " device_type=cuda num_gpu=2", 4, dataset_params=dsp)
    ref, s_ref = _run_reference(X, y, params, 4, dataset_params=dsp)
    _assert_models(ours, ref, 1e-9)
    np.testing.assert_allclose(s_ours, s_ref, rtol=1e-7, atol=1e-9)
