"""GPU parity of the device tree learner against the UNMODIFIED reference CPU learner (oracle/_ref), one tree at a
time with identical gradients fed to both (LGBM_BoosterUpdateOneIterCustom, src/c_api.cpp:2087-2100)."""
import numpy as np
import pytest

from lambdagap_b200 import capi
from oracle import ref_lgbm as R
from tests import helpers as H

pytestmark = pytest.mark.gpu


def _grads(y, seed):
    rng = np.random.default_rng(seed)
    s = rng.normal(scale=0.7, size=len(y))
    p = 1.0 / (1.0 + np.exp(-s))
    g = (p - y).astype(np.float32)
    h = (p * (1 - p)).astype(np.float32)
    return g, h


def _run_pair(X, y, g, h, num_leaves, extra_ref="", quant=False, cfg_kw=None, lr=0.1):
    params_ds = "max_bin=255 verbose=-1 num_threads=4 force_col_wise=true"
    ds = R.RefDataset(X, y, params=params_ds)
    params = ("objective=custom num_leaves=%d learning_rate=%g min_data_in_leaf=20 num_threads=4 force_col_wise=true "
              "deterministic=true verbose=-1 %s" % (num_leaves, lr, extra_ref))
    if quant:
        params += " use_quantized_grad=true stochastic_rounding=false num_grad_quant_bins=4"
    bst = R.RefBooster(ds, params)
    bst.update_custom(g, h)
    ref_tree = R.parse_model_trees(bst.model_string())[0]

    infos, cols = H.ref_infos_and_bins(ds)
    dds = capi.Dataset.from_bins(infos, cols)
    kw = dict(num_leaves=num_leaves, min_data_in_leaf=20, use_quantized_grad=int(quant), stochastic_rounding=0)
    kw.update(cfg_kw or {})
    lrn = capi.Learner(dds, capi.learner_config(**kw))
    res, recs = lrn.train_host(g, h)
    ours = H.build_tree(res, recs, infos, num_leaves)
    out = dict(ref=ref_tree, ours=ours, res=res, recs=recs, lrn=lrn, dds=dds, infos=infos, cols=cols, ref_ds=ds, bst=bst)
    return out


@pytest.mark.parametrize("const_hess", [False, True])
@pytest.mark.parametrize("features", [33, 36, 45, 48, 70])
def test_ragged_last_tile_histograms(gpu, features, const_hess):
    """Feature counts that leave a ragged last 32-feature tile (1, 4, 13, 16 features in it; 70 = two full tiles + 6), with
    and without the constant-hessian count plane: root histogram and parent-minus-sibling, every feature of the last tile
    and one of a full tile, against numpy; and the tree against the reference."""
    X, y = H.make_binary(40000, features, seed=features, zero_frac=0.2, nan_frac=0.03)
    g, h = _grads(y, 5)
    if const_hess:
        h = np.ones_like(h)
    kw = dict(is_constant_hessian=1) if const_hess else {}
    for num_leaves in (2, 3, 31):
        o = _run_pair(X, y, g, h, num_leaves, cfg_kw=kw)
        H.assert_tree_matches(o["ours"], o["ref"], shrinkage=0.1, rtol=1e-5)
        if num_leaves == 31:
            continue
        lrn, recs = o["lrn"], o["recs"]
        assign = lrn.leaf_assignment()
        if num_leaves == 2:
            leaf = 0 if recs[0]["left_count"] >= recs[0]["right_count"] else 1
            m = np.ones(len(y), dtype=bool)
        else:
            leaf = 1 if recs[1]["leaf"] == 0 else 0
            m = assign == leaf
        first_of_last = (features - 1) // 32 * 32
        for f in [3] + list(range(first_of_last, features)):
            bins = o["cols"][f]
            nb = o["infos"][f]["num_bin"]
            hist = lrn.histogram(leaf, f)
            eg = np.bincount(bins[m], weights=g[m].astype(np.float64), minlength=nb)[:nb]
            eh = np.bincount(bins[m], weights=h[m].astype(np.float64), minlength=nb)[:nb]
            np.testing.assert_allclose(hist[:, 0], eg, rtol=1e-6, atol=1e-5, err_msg="feature %d" % f)
            np.testing.assert_allclose(hist[:, 1], eh, rtol=1e-6, atol=1e-5, err_msg="feature %d" % f)


@pytest.mark.parametrize("quant", [True, False])
@pytest.mark.parametrize("case", ["plain", "nan", "zero", "discrete"])
def test_single_tree(gpu, quant, case):
    kw = dict(plain={}, nan=dict(nan_frac=0.1), zero=dict(zero_frac=0.4), discrete=dict(discrete_cols=3))[case]
    X, y = H.make_binary(20000, 12, seed=3, **kw)
    g, h = _grads(y, 5)
    o = _run_pair(X, y, g, h, 31, quant=quant)
    H.assert_tree_matches(o["ours"], o["ref"], shrinkage=0.1, rtol=1e-12 if quant else 1e-6, what="%s quant=%s" % (case, quant))
    # partition: leaf assignment must reproduce the leaf counts and be consistent with the index segments
    assign = o["lrn"].leaf_assignment()
    nl = o["ref"]["num_leaves"]
    np.testing.assert_array_equal(np.bincount(assign, minlength=nl)[:nl], o["ref"]["leaf_count"])
    for leaf in (0, nl - 1):
        idx = o["lrn"].leaf_indices(leaf)
        assert np.all(np.diff(idx) > 0), "stable partition keeps ascending order"
        assert np.all(assign[idx] == leaf)


@pytest.mark.parametrize("quant", [True, False])
def test_histograms_exact(gpu, quant):
    """Root histogram (2 leaves: the parent's slot is never subtracted) and parent-minus-sibling subtraction
    (3 leaves) against a direct numpy accumulation of the same integer / float gradients."""
    X, y = H.make_binary(30000, 40, seed=11, zero_frac=0.3)
    g, h = _grads(y, 7)
    for num_leaves in (2, 3):
        o = _run_pair(X, y, g, h, num_leaves, quant=quant)
        lrn, recs = o["lrn"], o["recs"]
        assert len(recs) == num_leaves - 1
        assign = lrn.leaf_assignment()
        if quant:
            q = lrn.discretized().reshape(-1, 2).astype(np.float64)  # [hess, grad]
            wg, wh = q[:, 1], q[:, 0]
        else:
            wg, wh = g.astype(np.float64), h.astype(np.float64)
        if num_leaves == 2:
            leaf = 0 if recs[0]["left_count"] >= recs[0]["right_count"] else 1  # keeps the parent's slot
            m = np.ones(len(y), dtype=bool)
        else:
            leaf = 1 if recs[1]["leaf"] == 0 else 0  # the child of split 0 that was not split again
            m = assign == leaf
        for f in (0, 5, 31, 32, 39):
            bins = o["cols"][f]
            nb = o["infos"][f]["num_bin"]
            hist = lrn.histogram(leaf, f)
            eg = np.bincount(bins[m], weights=wg[m], minlength=nb)[:nb]
            eh = np.bincount(bins[m], weights=wh[m], minlength=nb)[:nb]
            if quant:
                np.testing.assert_array_equal(hist[:, 0], eg)
                np.testing.assert_array_equal(hist[:, 1], eh)
            else:
                # float mode: exact 2^-k fixed point of each addend -> |err| <= n_bin * 2^-31 * max|g|
                np.testing.assert_allclose(hist[:, 0], eg, rtol=1e-6, atol=1e-5)
                np.testing.assert_allclose(hist[:, 1], eh, rtol=1e-6, atol=1e-5)
