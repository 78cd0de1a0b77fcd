"""The wave schedule (lambdagap_b200/csrc/learner.cu) must be invisible in the result: whatever is expanded
speculatively, the tree equals the one the reference's strictly sequential best-first loop grows
(serial_tree_learner.cpp:210-245).  Checked against the oracle's C restatement (sequential by construction) for
every speculation setting, for shapes that stress the schedule (chains, depth limits, tiny leaves, more leaves than
the data supports), and repeatedly (a schedule-dependent result is a race)."""
import os

import numpy as np
import pytest

from lambdagap_b200 import capi
from oracle import oracle as O

pytestmark = pytest.mark.gpu

KEYS = ("leaf", "feature", "threshold_bin", "default_left", "left_count", "right_count")


def _data(seed, n, f, signal=0.0):
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
    g = rng.normal(size=n)
    if signal:
        # a chain-shaped optimum: one feature carries a strong monotone signal, so the best leaf is always the newest
        g += signal * (cols[0].astype(np.float64) / 255.0) ** 3
    h = rng.uniform(0.05, 1.0, size=n)
    return infos, cols, g.astype(np.float32), h.astype(np.float32)


def _compare(infos, cols, g, h, quant, kw, reps=1):
    root, recs, leaf_of_row = O.train_tree(cols, infos, g, h, O.make_config(use_quantized_grad=int(quant), **kw))
    dds = capi.Dataset.from_bins(infos, cols)
    lrn = capi.Learner(dds, capi.learner_config(use_quantized_grad=int(quant), stochastic_rounding=0, **kw))
    for rep in range(reps):
        res, ours = lrn.train_host(g, h)
        assert res["num_splits"] == len(recs), "rep %d" % rep
        for k, (a, b) in enumerate(zip(ours, recs)):
            for key in KEYS:
                assert a[key] == b[key], "rep %d split %d %s: %s vs %s" % (rep, k, key, a[key], b[key])
            np.testing.assert_allclose([a["left_output"], a["right_output"], a["gain"]],
                                       [b["left_output"], b["right_output"], b["gain"]], rtol=1e-12 if quant else 1e-6)
    np.testing.assert_array_equal(lrn.leaf_assignment(), leaf_of_row)
    prof = lrn.profile()
    lrn.free()
    dds.free()
    return prof, len(recs)


@pytest.fixture
def spec_env():
    old = os.environ.get("GBL_SPEC_LIMIT")
    yield
    if old is None:
        os.environ.pop("GBL_SPEC_LIMIT", None)
    else:
        os.environ["GBL_SPEC_LIMIT"] = old


@pytest.mark.parametrize("quant", [True, False])
@pytest.mark.parametrize("spec", ["default", "0", "7"])
def test_result_does_not_depend_on_speculation(gpu, spec_env, quant, spec):
    """GBL_SPEC_LIMIT=0 is the reference's own order (one certain node per wave); 7 switches from speculative to
    certain expansion in the middle of the tree; default speculates all the way."""
    if spec == "default":
        os.environ.pop("GBL_SPEC_LIMIT", None)
    else:
        os.environ["GBL_SPEC_LIMIT"] = spec
    infos, cols, g, h = _data(17, 60000, 33)
    prof, ns = _compare(infos, cols, g, h, quant, dict(num_leaves=63, min_data_in_leaf=25, lambda_l2=0.5))
    assert prof["splits"] == ns
    if spec == "0":
        assert prof["expansions"] == ns and prof["waves"] == ns  # nothing speculative, one node per wave
    else:
        assert prof["waves"] < ns  # several nodes per wave


@pytest.mark.parametrize("quant", [True, False])
def test_repeated_trees_are_identical(gpu, quant):
    """Same inputs, same learner, many trees: any difference would be a race between CTAs of a wave."""
    infos, cols, g, h = _data(17, 60000, 33)
    _compare(infos, cols, g, h, quant, dict(num_leaves=63, min_data_in_leaf=25, lambda_l2=0.5), reps=25)


@pytest.mark.parametrize("case", ["chain", "flat_gains", "max_depth", "tiny_leaves", "too_many_leaves", "two_leaves", "no_split"])
def test_schedule_edge_shapes(gpu, case):
    if case == "chain":
        infos, cols, g, h = _data(5, 40000, 9, signal=6.0)
        kw = dict(num_leaves=40, min_data_in_leaf=50)
    elif case == "flat_gains":
        # pure-noise gradients: every gain is small and about the same (late boosting iterations look like this), so
        # children routinely beat their parents' gains and whole subtrees tie in priority; the tree grows long chains
        infos, cols, g, h = _data(12, 120000, 10)
        kw = dict(num_leaves=255, min_data_in_leaf=20)
    elif case == "max_depth":
        infos, cols, g, h = _data(6, 50000, 12)
        kw = dict(num_leaves=255, min_data_in_leaf=20, max_depth=5)
    elif case == "tiny_leaves":
        infos, cols, g, h = _data(7, 3000, 10)
        kw = dict(num_leaves=127, min_data_in_leaf=1, min_sum_hessian_in_leaf=1e-3)
    elif case == "too_many_leaves":
        infos, cols, g, h = _data(8, 400, 6)
        kw = dict(num_leaves=255, min_data_in_leaf=20)
    elif case == "two_leaves":
        infos, cols, g, h = _data(9, 20000, 8)
        kw = dict(num_leaves=2, min_data_in_leaf=20)
    else:
        infos, cols, g, h = _data(10, 30, 5)
        kw = dict(num_leaves=31, min_data_in_leaf=20)
    # float mode is exercised where gains are well separated.  In leaves without missing rows the two scan directions
    # of a feature describe the same partition; their gains then differ only by rounding (~1e-14), differently for
    # fixed-point and for the reference's sequential double sums, so which of the two equivalent splits is reported
    # (threshold / default_left) is not pinned for very small leaves or noise-only gradients.  Quantized mode is
    # bit-exact everywhere.
    for quant in ((True,) if case in ("tiny_leaves", "too_many_leaves", "flat_gains") else (True, False)):
        prof, ns = _compare(infos, cols, g, h, quant, kw)
        if case == "no_split":
            assert ns == 0
        if case == "chain":
            assert prof["waves"] >= 8
        if case == "flat_gains":
            assert ns == 254 and prof["waves"] < 120 and prof["expansions"] < 4 * ns  # bounded speculation


def test_many_leaves_replay_in_global_scratch(gpu):
    """num_leaves large enough that the replay's compact node arrays no longer fit shared memory
    (finalize_kernel falls back to a global scratch) and the speculation budget is in the thousands."""
    infos, cols, g, h = _data(21, 60000, 12, signal=2.0)
    prof, ns = _compare(infos, cols, g, h, True, dict(num_leaves=1200, min_data_in_leaf=3, min_sum_hessian_in_leaf=1e-3))
    assert ns > 600


def test_bagging_subset(gpu):
    """SetBaggingData (tree_learner.h:79): the tree is grown on the bagged rows only -- the same tree the sequential
    restatement grows on the extracted subset -- and the score update touches only those rows."""
    infos, cols, g, h = _data(31, 50000, 10, signal=1.5)
    rng = np.random.default_rng(3)
    bag = np.sort(rng.choice(50000, size=31000, replace=False)).astype(np.int32)
    kw = dict(num_leaves=31, min_data_in_leaf=40)
    sub_cols = [c[bag] for c in cols]
    root, recs, leaf_of_row = O.train_tree(sub_cols, infos, g[bag], h[bag], O.make_config(use_quantized_grad=0, **kw))
    dds = capi.Dataset.from_bins(infos, cols)
    lrn = capi.Learner(dds, capi.learner_config(use_quantized_grad=0, stochastic_rounding=0, **kw))
    lrn.set_bagging_data(bag)
    for rep in range(2):
        res, ours = lrn.train_host(g, h)
        assert res["num_splits"] == len(recs)
        for k, (a, b) in enumerate(zip(ours, recs)):
            for key in KEYS:
                assert a[key] == b[key], "split %d %s: %s vs %s" % (k, key, a[key], b[key])
        assign = lrn.leaf_assignment()
        np.testing.assert_array_equal(assign[bag], leaf_of_row)
        out = np.ones(50000, dtype=bool)
        out[bag] = False
        assert np.all(assign[out] == -1)  # rows outside the bag belong to no leaf
    lrn.set_bagging_data(None)  # back to all rows
    res, ours = lrn.train_host(g, h)
    assert np.all(lrn.leaf_assignment() >= 0)
    lrn.free()
    dds.free()


def test_packed_columns_round_trip(gpu):
    """Column store: features with at most 16 bins are kept 4-bit packed (two rows per byte, even row in the low
    nibble, src/io/dense_bin.hpp:69-82); both views of the dataset must return the bins that went in, for an odd row
    count as well, and splits on packed features must partition like the sequential restatement."""
    infos, cols, g, h = _data(41, 30001, 9)
    dds = capi.Dataset.from_bins(infos, cols)
    for f in range(len(cols)):
        np.testing.assert_array_equal(dds.feature_bins(f), cols[f], err_msg="column store, feature %d" % f)
        np.testing.assert_array_equal(dds.feature_bins(f, rowmajor=True), cols[f], err_msg="row mirror, feature %d" % f)
    dds.free()
    # only low-cardinality features: every split reads a packed column
    keep = [f for f in range(len(cols)) if infos[f]["num_bin"] <= 16]
    sub_infos = [dict(infos[f], real_index=i) for i, f in enumerate(keep)]
    sub_cols = [cols[f] for f in keep]
    prof, ns = _compare(sub_infos, sub_cols, g, h, True, dict(num_leaves=31, min_data_in_leaf=20))
    assert ns >= 5


def _reference_bags(n, seed, fraction, rounds):
    """BaggingSampleStrategy::BaggingHelper (src/boosting/bagging.hpp:224-243) restated with integers: one LCG per block
    of 1024 rows (include/LightGBM/utils/random.h), state carried from round to round."""
    nblk = (n + 1023) // 1024
    x = (seed + np.arange(nblk, dtype=np.int64)).astype(np.uint32)
    bags = []
    for _ in range(rounds):
        draws = np.zeros((nblk, 1024), dtype=np.float32)
        for k in range(1024):
            live = (np.arange(nblk) * 1024 + k) < n  # the last block draws only for the rows it has
            nx = (x.astype(np.uint64) * np.uint64(214013) + np.uint64(2531011)).astype(np.uint32)
            x = np.where(live, nx, x)
            draws[:, k] = ((x >> np.uint32(16)) & np.uint32(0x7FFF)).astype(np.float32) / np.float32(32768.0)
        flat = draws.reshape(-1)[:n]
        bags.append(np.flatnonzero(flat.astype(np.float64) < fraction).astype(np.int32))
    return bags


def test_bagging_drawn_on_device(gpu):
    """gbl_learner_bagging_dev draws the reference's bag (block-wise LCGs, state carried between calls) on the device:
    the tree equals the one grown on the same bag handed over from the host, rows outside the bag belong to no leaf,
    and the bag size changes from call to call without invalidating anything."""
    n = 50000 + 333  # the last block of 1024 rows is partial
    infos, cols, g, h = _data(32, n, 10, signal=1.5)
    kw = dict(num_leaves=31, min_data_in_leaf=40)
    seed, fraction = 7, 0.62
    bags = _reference_bags(n, seed, fraction, 3)
    assert len({len(b) for b in bags}) > 1
    dds = capi.Dataset.from_bins(infos, cols)
    dev = capi.Learner(dds, capi.learner_config(use_quantized_grad=0, stochastic_rounding=0, **kw))
    host = capi.Learner(dds, capi.learner_config(use_quantized_grad=0, stochastic_rounding=0, **kw))
    for rnd, bag in enumerate(bags):
        assert dev.bagging_dev(seed, fraction) == len(bag)
        host.set_bagging_data(bag)
        res_d, recs_d = dev.train_host(g, h)
        res_h, recs_h = host.train_host(g, h)
        assert res_d["num_splits"] == res_h["num_splits"] > 5
        for a, b in zip(recs_d, recs_h):
            assert a == b, "round %d" % rnd
        assign = dev.leaf_assignment()
        np.testing.assert_array_equal(assign, host.leaf_assignment())
        inside = np.zeros(n, dtype=bool)
        inside[bag] = True
        assert np.all(assign[inside] >= 0) and np.all(assign[~inside] == -1)
    # re-seeding restarts the sequence
    assert dev.bagging_dev(seed, fraction, reset=True) == len(bags[0])
    # a fraction of 1 puts every row in the bag
    assert dev.bagging_dev(seed, 1.0) == n
    res, _ = dev.train_host(g, h)
    assert np.all(dev.leaf_assignment() >= 0)
    dev.free()
    host.free()
    dds.free()
