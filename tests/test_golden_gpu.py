"""GPU: the CUDA path through the C API against the committed golden fixtures (generated from the unmodified
reference by tests/golden/make_golden.py) -- works on the GPU box even without oracle/_ref -- and against the
oracle's C restatement on seeded inputs of a size it finishes in seconds."""
import numpy as np
import pytest

from lambdagap_b200 import capi, engine
from oracle import oracle as O
from tests import helpers as H
from tests.test_oracle_cpu import GOLDEN, _oracle_config

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", GOLDEN)
def test_cuda_path_reproduces_golden(gpu, name):
    fx = H.load_golden(name)
    ds = engine.Dataset(fx["X"], fx["y"], group=fx["group"], params="max_bin=255 verbose=-1")
    bst = engine.Booster(ds, fx["params"] + " device_type=cuda verbose=-1 num_threads=4")
    for _ in range(fx["iters"]):
        bst.update()
    from oracle import ref_lgbm as R  # parser only (pure Python)
    ours = R.parse_model_trees(bst.model_to_string())
    assert len(ours) == len(fx["trees"])
    quant = "use_quantized_grad=true" in fx["params"]
    for k, (a, b) in enumerate(zip(ours, fx["trees"])):
        assert a["num_leaves"] == b["num_leaves"]
        for key in ("split_feature", "left_child", "right_child", "decision_type", "leaf_count", "internal_count"):
            np.testing.assert_array_equal(a[key], b[key], err_msg="%s tree %d %s" % (name, k, key))
        np.testing.assert_array_equal(a["threshold"], b["threshold"])
        np.testing.assert_allclose(a["leaf_value"], b["leaf_value"], rtol=1e-12 if quant and k == 0 else 1e-5, atol=1e-10)
    np.testing.assert_allclose(bst.train_scores(raw=True), fx["scores"], rtol=1e-5, atol=1e-8)
    bst.free()
    ds.free()


@pytest.mark.parametrize("quant", [True, False])
def test_cuda_tree_equals_oracle_tree(gpu, quant):
    """One tree, identical gradients: CUDA learner through the C ABI vs the sequential C restatement."""
    rng = np.random.default_rng(17)
    n, f = 60000, 33
    infos, cols = [], []
    for j in range(f):
        nb = [255, 64, 12, 2][j % 4]
        mt = j % 3
        col = rng.integers(0, nb, size=n).astype(np.uint8)
        if j % 5 == 0:
            col[rng.random(n) < 0.6] = nb // 2  # a dominant bin
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
    kw = dict(num_leaves=63, min_data_in_leaf=25, lambda_l2=0.5)
    root, recs, leaf_of_row = O.train_tree(cols, infos, g, h, O.make_config(use_quantized_grad=int(quant), **kw))
    dds = capi.Dataset.from_bins(infos, cols)
    lrn = capi.Learner(dds, capi.learner_config(use_quantized_grad=int(quant), stochastic_rounding=0, **kw))
    res, ours = lrn.train_host(g, h)
    assert res["num_splits"] == len(recs) > 30
    for k, (a, b) in enumerate(zip(ours, recs)):
        for key in ("leaf", "feature", "threshold_bin", "default_left", "left_count", "right_count"):
            assert a[key] == b[key], "split %d %s: %s vs %s" % (k, key, a[key], b[key])
        np.testing.assert_allclose([a["left_output"], a["right_output"], a["gain"]],
                                   [b["left_output"], b["right_output"], b["gain"]], rtol=1e-12 if quant else 1e-6)
    np.testing.assert_array_equal(lrn.leaf_assignment(), leaf_of_row)
    lrn.free()
    dds.free()


def test_cuda_lambdarank_gradients_equal_oracle(gpu):
    """Raw lambda / hessian arrays (not just trees) for every target at non-trivial scores."""
    X, y, group = H.make_ranking(300, 4, seed=5, min_len=6, max_len=200)
    n = len(y)
    qb = np.concatenate([[0], np.cumsum(group)]).astype(np.int32)
    score = np.random.default_rng(2).normal(size=n)
    sdev = capi.DeviceBuffer.from_numpy(score)
    gdev, hdev = capi.DeviceBuffer(4 * n), capi.DeviceBuffer(4 * n)
    import ctypes
    for t, target in enumerate(O.RANK_TARGETS):
        cfg = capi.ObjectiveConfig(kind=capi.OBJ_LAMBDARANK, lambdarank_target=t, lambdarank_truncation_level=5,
                                   lambdarank_norm=1, sigmoid=1.0, lambdagap_weight=1.5, is_unbalance=0,
                                   scale_pos_weight=1.0, num_label_gain=0)
        h = ctypes.c_void_p()
        capi.check(capi.lib().gbl_objective_create(ctypes.byref(cfg), ctypes.c_int64(n), y.ctypes.data_as(ctypes.c_void_p),
                                                   None, qb.ctypes.data_as(ctypes.c_void_p), ctypes.c_int32(len(group)),
                                                   ctypes.c_int32(0), ctypes.byref(h)))
        capi.check(capi.lib().gbl_objective_get_gradients(h, ctypes.c_void_p(sdev.ptr), ctypes.c_void_p(gdev.ptr),
                                                          ctypes.c_void_p(hdev.ptr), None))
        capi.synchronize()
        g_gpu, h_gpu = gdev.download(np.float32, n), hdev.download(np.float32, n)
        capi.lib().gbl_objective_free(h)
        ro = O.RankObjective(target, 5, True, 1.0, 1.5)
        g_cpu, h_cpu = ro.gradients(score, y, qb)
        ro.free()
        scale = max(1e-12, float(np.abs(g_cpu).max()))
        np.testing.assert_allclose(g_gpu, g_cpu, rtol=1e-5, atol=1e-6 * scale, err_msg=target)
        np.testing.assert_allclose(h_gpu, h_cpu, rtol=1e-5, atol=1e-6 * scale, err_msg=target)
