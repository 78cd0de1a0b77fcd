"""CPU-only tests (`-m "not gpu"`): the oracle's C restatement and the host-side logic against the golden
fixtures generated from the unmodified reference (tests/golden/make_golden.py), plus the C-ABI export check.
No GPU compute is called here."""
import os
import re

import numpy as np
import pytest

from lambdagap_b200 import capi, engine
from oracle import oracle as O
from tests import helpers as H

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = ["binary_float", "binary_quant", "regression_l2", "rank_ndcg", "rank_lambdagap_x", "rank_arpk",
          "rank_ll_ndcg_pp", "rank_precision"]


def _oracle_config(p, kind):
    return O.make_config(num_leaves=int(p.get("num_leaves", 31)), max_depth=int(p.get("max_depth", -1)),
                         min_data_in_leaf=int(p.get("min_data_in_leaf", 20)),
                         use_quantized_grad=int(p.get("use_quantized_grad", "false") == "true"),
                         num_grad_quant_bins=int(p.get("num_grad_quant_bins", 4)),
                         is_constant_hessian=int(kind == "regression"),
                         lambda_l1=float(p.get("lambda_l1", 0.0)), lambda_l2=float(p.get("lambda_l2", 0.0)))


@pytest.mark.parametrize("name", GOLDEN)
def test_oracle_reproduces_reference_trees(name):
    """The C restatement, fed the reference's own bins, grows the reference's trees."""
    fx = H.load_golden(name)
    p = H.params_to_dict(fx["params"])
    cfg = _oracle_config(p, fx["kind"])
    lr = float(p.get("learning_rate", 0.1))
    rank = None
    if fx["kind"] == "rank":
        rank = O.RankObjective(p["lambdarank_target"], int(p["lambdarank_truncation_level"]), True, 1.0,
                               float(p.get("lambdagap_weight", 1.0)))
    objective = {"binary": "binary", "regression": "regression", "rank": "lambdarank"}[fx["kind"]]
    trees, score = O.boost(list(fx["bins"]), fx["infos"], fx["y"], objective, cfg, fx["iters"], lr, group=fx["group"], rank=rank)
    assert len(trees) == len(fx["trees"])
    quant = bool(cfg.use_quantized_grad)
    for k, ((root, recs, bias), ref) in enumerate(zip(trees, fx["trees"])):
        ours = H.build_tree({"root_output": root}, _as_records(recs), fx["infos"], cfg.num_leaves)
        H.assert_tree_matches(ours, ref, shrinkage=lr, bias=bias, rtol=1e-9 if not quant else 1e-12, what="%s tree %d" % (name, k))
    np.testing.assert_allclose(score, fx["scores"], rtol=1e-9, atol=1e-12)


def _as_records(recs):
    out = []
    for r in recs:
        d = dict(r)
        out.append(d)
    return out


@pytest.mark.parametrize("name", ["binary_float", "regression_l2", "rank_ndcg"])
def test_host_bin_mapper_matches_reference(name):
    """BinMapper::FindBin restated in lambdagap_b200/host/bin_mapper.cpp: boundaries bit-exact vs the reference's."""
    fx = H.load_golden(name)
    X = fx["X"].astype(np.float64)
    n = X.shape[0]
    for info in fx["infos"]:
        col = X[:, info["real_index"]]
        vals = col[(np.abs(col) > 1e-35) | np.isnan(col)]
        got = engine.find_bin(vals, n, max_bin=255, min_data_in_bin=3, min_split_data=20)
        assert not got["is_trivial"]
        for key in ("num_bin", "missing_type", "default_bin", "most_freq_bin"):
            assert got[key] == info[key], key
        # the bound stored for the NaN bin itself is a placeholder that no code path reads (ValueToBin stops one
        # bin earlier, a threshold is never the last bin); the reference happens to store 2.0 there, we store NaN
        real = info["num_bin"] - (1 if info["missing_type"] == 2 else 0)
        np.testing.assert_array_equal(got["upper_bounds"][:real], info["upper_bounds"][:real])
        # and ValueToBin through numpy's searchsorted on those bounds reproduces the reference's bins
        ub = info["upper_bounds"]
        nb = info["num_bin"] - (1 if info["missing_type"] == 2 else 0)
        v = np.where(np.isnan(col), 0.0, col)
        bins = np.minimum(np.searchsorted(ub[:nb - 1], v, side="left"), nb - 1)
        if info["missing_type"] == 2:
            bins = np.where(np.isnan(col), info["num_bin"] - 1, bins)
        np.testing.assert_array_equal(bins.astype(np.uint8), fx["bins"][fx["infos"].index(info)])


def test_oracle_histogram_is_plain_accumulation():
    rng = np.random.default_rng(0)
    n = 5000
    col = rng.integers(0, 37, size=n).astype(np.uint8)
    g = rng.normal(size=n).astype(np.float32)
    h = rng.uniform(0.1, 1, size=n).astype(np.float32)
    info = dict(num_bin=37, missing_type=0, default_bin=0, most_freq_bin=5)
    idx = np.sort(rng.choice(n, size=1700, replace=False)).astype(np.int32)
    hist = O.histogram_float(col, info, idx, g, h)
    m = col[idx] != 5
    eg = np.bincount(col[idx][m], weights=g[idx][m].astype(np.float64), minlength=37)
    np.testing.assert_allclose(hist[:, 0], eg, rtol=1e-12, atol=1e-12)
    assert hist[5, 0] == 0 and hist[5, 1] == 0  # the most-frequent bin is reconstructed by FixHistogram, not summed


def test_oracle_discretize_matches_formula():
    rng = np.random.default_rng(1)
    g = rng.normal(size=1000).astype(np.float32)
    h = rng.uniform(0, 0.25, size=1000).astype(np.float32)
    q, gs, hs = O.discretize(g, h, 4, False)
    q = q.reshape(-1, 2)
    assert gs == np.abs(g).max().astype(np.float64) / 2 and hs == np.abs(h).max().astype(np.float64) / 4
    eg = np.trunc(g.astype(np.float64) * (1.0 / gs) + np.where(g >= 0, 0.5, -0.5)).astype(np.int8)
    np.testing.assert_array_equal(q[:, 1], eg)
    np.testing.assert_array_equal(q[:, 0], np.trunc(h.astype(np.float64) * (1.0 / hs) + 0.5).astype(np.int8))


def _declared(header, macro):
    text = open(os.path.join(ROOT, "include", header)).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    text = "\n".join(ln for ln in text.split("\n") if not ln.lstrip().startswith("#"))
    return re.findall(macro + r"[^;(]*?\b(\w+)\s*\(", text)


def test_c_abi_library_exports_every_declared_symbol():
    """Loads liblambdagap_b200.so / liblgbm_b200.so on this GPU-less machine and checks the headers' symbols."""
    names = _declared("lambdagap_b200.h", "GBL_API")
    assert len(names) > 40
    lib = capi.lib()
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing
    assert sorted(set(names)) == sorted(set(capi.EXPORTED_SYMBOLS)), set(names) ^ set(capi.EXPORTED_SYMBOLS)
    hnames = _declared("lgbm_b200_c_api.h", "LGBM_B200_EXPORT")
    hl = engine.lib()
    missing = [n for n in hnames if not hasattr(hl, n)]
    assert not missing, missing
    assert sorted(set(hnames)) == sorted(set(engine.LGBM_EXPORTED_SYMBOLS))


def test_no_cpu_fallback_fails_loudly():
    """Without an sm_100 device every compute entry point must fail with a clear error, never fall back."""
    if capi.device_count() > 0:
        pytest.skip("a GPU is present")
    infos = [dict(num_bin=4, missing_type=0, default_bin=0, most_freq_bin=0, real_index=0)]
    with pytest.raises(capi.GblError):
        capi.Dataset.from_bins(infos, [np.zeros(16, dtype=np.uint8)])
    X = np.random.default_rng(0).normal(size=(100, 3)).astype(np.float32)
    with pytest.raises(engine.LightGBMError):
        engine.Dataset(X, (X[:, 0] > 0).astype(np.float32), params="verbose=-1")


def test_config_rejects_what_the_learner_does_not_implement():
    """Parameters that would change what is trained are rejected by the parser itself (LGBM_B200_DumpConfig needs no
    device, so the error cannot be the missing GPU), under their aliases too."""
    for bad in ("boosting=dart", "linear_tree=true", "feature_fraction_bynode=0.5", "max_bin=1023", "device_type=cpu",
                "monotone_constraints=1,0", "mc=1,0", "tree_learner=voting", "data_sample_strategy=goss", "boosting=goss",
                "num_class=3", "fc=0.5,1", "cat_feature=0", "objective=huber", "extra_trees=true"):
        with pytest.raises(engine.LightGBMError, match="not supported|no CPU path|one byte per cell"):
            engine.dump_config(bad + " objective=binary device_type=cuda verbose=-1")  # the first occurrence of a key wins
    engine.dump_config("objective=binary device_type=cuda verbose=-1 feature_fraction=0.5 bagging_fraction=0.5 bagging_freq=1")


def test_metric_restatement_matches_reference():
    """oracle/oracle_metrics.py against the unmodified reference: the reference evaluates its metrics on its own scores
    (LGBM_BoosterGetEval), the restatement on the same scores (LGBM_BoosterGetPredict).  AUC is exact (integer
    arithmetic in doubles); the means differ only in summation order."""
    from oracle import oracle_metrics as M
    from oracle import ref_lgbm as R
    if not R.available():
        pytest.skip("oracle/_ref not built")
    rng = np.random.default_rng(0)
    n, f, nv = 6000, 8, 2500
    X = rng.normal(size=(n, f)).astype(np.float32)
    y = (X[:, 0] + 0.5 * X[:, 1] * X[:, 2] + 0.5 * rng.normal(size=n) > 0).astype(np.float32)
    Xv = rng.normal(size=(nv, f)).astype(np.float32)
    yv = (Xv[:, 0] + 0.5 * Xv[:, 1] * Xv[:, 2] + 0.5 * rng.normal(size=nv) > 0).astype(np.float32)
    ds = R.RefDataset(X, y, params="max_bin=63 verbose=-1")
    dv = R.RefDataset(Xv, yv, params="max_bin=63 verbose=-1", reference=ds)
    common = " num_leaves=7 verbose=-1 num_threads=2 min_data_in_leaf=200"
    b = R.RefBooster(ds, "objective=binary metric=binary_logloss,auc,binary_error" + common)
    b.add_valid(dv)
    for _ in range(4):
        b.update()
    assert b.eval_names() == ["binary_logloss", "auc", "binary_error"]
    for idx, lab in ((0, y), (1, yv)):
        e = b.eval(idx)
        p = b.train_scores() if idx == 0 else b.valid_scores(1)  # few distinct values: large tie groups
        np.testing.assert_allclose(e[0], M.binary_logloss(p, lab), rtol=1e-13)
        assert e[1] == M.auc(p, lab)
        assert e[2] == M.binary_error(p, lab)
    b = R.RefBooster(ds, "objective=regression metric=l2,rmse,l1,auc" + common)
    b.add_valid(dv)
    for _ in range(4):
        b.update()
    e, p = b.eval(1), b.valid_scores(1)
    np.testing.assert_allclose(e[:3], [M.l2(p, yv), M.rmse(p, yv), M.l1(p, yv)], rtol=1e-13)
    assert e[3] == M.auc(p, yv)
    q = rng.integers(1, 60, size=150)
    q[3] = 400
    qb = np.concatenate(([0], np.cumsum(q)))
    Xr = rng.normal(size=(qb[-1], f)).astype(np.float32)
    yr = np.clip((Xr[:, 0] * 1.2 + rng.normal(size=qb[-1])).round(), 0, 4).astype(np.float32)
    yr[qb[5]:qb[6]] = 0  # a query without relevant documents counts NDCG = 1
    dr = R.RefDataset(Xr, yr, group=q.astype(np.int32), params="max_bin=63 verbose=-1")
    b = R.RefBooster(dr, "objective=lambdarank metric=ndcg eval_at=1,3,5,10,500 num_leaves=7 verbose=-1 num_threads=2")
    for _ in range(3):
        b.update()
    assert b.eval_names() == ["ndcg@1", "ndcg@3", "ndcg@5", "ndcg@10", "ndcg@500"]
    np.testing.assert_allclose(b.eval(0), M.ndcg(b.train_scores(), yr, qb, [1, 3, 5, 10, 500]), rtol=1e-14)


def _ref_param_block(params, group=None):
    from oracle import ref_lgbm as R
    rng = np.random.default_rng(0)
    X = rng.normal(size=(400, 4)).astype(np.float32)
    y = rng.integers(0, 2, size=400).astype(np.float32)
    ds = R.RefDataset(X, y, group=group, params="verbose=-1")
    text = R.RefBooster(ds, params + " verbose=-1").model_string()
    block = text[text.index("parameters:"):]
    out = {}
    for line in block.split("\n"):
        if line.startswith("[") and ": " in line:
            k, v = line[1:-1].split(": ", 1)
            out[k] = v
        elif line.startswith("["):
            out[line[1:-2].rstrip(":")] = ""
    return out


@pytest.mark.parametrize("params,group", [
    ("objective=binary seed=7 metric=auc,binary,auc", None),
    ("objective=regression_l2 metric=mse,l2_root,mae,l2", None),
    ("objective=mse", None),
    ("objective=lambdarank ndcg_eval_at=5,1,3 seed=123", [100] * 4),
    ("objective=lambdarank metric=lambdarank", [100] * 4),
    ("application=binary num_leaf=7 min_data=9 subsample=0.8 subsample_freq=2 bagging_fraction_seed=77", None),
    ("objective=binary max_depth=3", None),                       # a depth limit without num_leaves caps num_leaves at 2^depth
    ("objective=binary max_depth=3 num_leaves=20", None),
    ("objective=binary path_smooth=1.5 min_data_in_leaf=1", None),  # path smoothing needs min_data_in_leaf >= 2
    ("objective=binary min_data_in_leaf=0 min_sum_hessian_in_leaf=0", None),
])
def test_host_config_matches_reference(params, group):
    """Config::Set of the host layer (aliases, metric list with aliases resolved and duplicates dropped, the objective's
    default metric, sorted eval_at, the seeds derived from `seed`) against the parameter block the reference writes into
    its model file (Config::SaveMembersToString).  Needs no GPU: LGBM_B200_DumpConfig is a tap on the parser."""
    from oracle import ref_lgbm as R
    if not R.available():
        pytest.skip("oracle/_ref not built")
    ref = _ref_param_block(params, None if group is None else np.array(group, dtype=np.int32))
    ours = engine.dump_config(params + " device_type=cuda")
    for key in ("objective", "metric", "eval_at", "data_random_seed", "bagging_seed", "feature_fraction_seed", "num_leaves",
                "bagging_freq", "min_data_in_leaf"):
        assert ours[key] == ref[key], "%s: ours %r, reference %r" % (key, ours[key], ref[key])


@pytest.mark.parametrize("ds_params", [
    "max_bin=255", "max_bin=15", "max_bin=63 min_data_in_bin=10", "max_bin=255 zero_as_missing=true",
    "max_bin=255 use_missing=false", "max_bin=7 min_data_in_bin=1", "max_bin=255 feature_pre_filter=false",
])
def test_host_bin_mapper_edge_columns_against_reference(ds_params):
    """BinMapper::FindBin (host restatement, no GPU) against the unmodified reference on columns chosen to hit its
    branches: few distinct values, heavy zeros, NaN-rich, negative-only, constants, huge ranges, values straddling zero,
    under several dataset parameters (max_bin, min_data_in_bin, zero_as_missing, use_missing, feature_pre_filter)."""
    from oracle import ref_lgbm as R
    if not R.available():
        pytest.skip("oracle/_ref not built")
    rng = np.random.default_rng(5)
    n = 4000
    cols = [
        rng.normal(size=n),                                              # continuous
        rng.integers(0, 5, size=n).astype(np.float64),                   # few distinct values incl. zero
        np.where(rng.random(n) < 0.9, 0.0, rng.normal(size=n)),          # 90 % zeros
        np.where(rng.random(n) < 0.3, np.nan, rng.normal(size=n)),       # NaN-rich
        -np.abs(rng.normal(size=n)) - 0.5,                               # negative only
        np.exp(rng.normal(size=n) * 4),                                  # huge positive range
        np.round(rng.normal(size=n), 1),                                 # a few hundred distinct values around zero
        np.where(rng.random(n) < 0.5, np.nan, 0.0),                      # NaN or zero only
        np.full(n, 3.25),                                                # constant
        rng.integers(-3, 4, size=n).astype(np.float64) * 1e-3,           # tiny magnitudes around zero
        np.where(rng.random(n) < 0.02, rng.normal(size=n), 0.0),         # almost all zeros (pre-filter candidate)
    ]
    X = np.stack(cols, axis=1)
    y = rng.integers(0, 2, size=n).astype(np.float32)
    ds = R.RefDataset(X, y, params=ds_params + " verbose=-1 min_data_in_leaf=20")
    ref_by_col = {}
    for i in range(ds.num_used_features()):
        info = ds.feature_info(i)
        ref_by_col[info["real_index"]] = info
    kw = dict(max_bin=255, min_data_in_bin=3, pre_filter=True, use_missing=True, zero_as_missing=False)
    for tok in ds_params.split():
        k, v = tok.split("=")
        if k == "max_bin":
            kw["max_bin"] = int(v)
        elif k == "min_data_in_bin":
            kw["min_data_in_bin"] = int(v)
        elif k == "zero_as_missing":
            kw["zero_as_missing"] = v == "true"
        elif k == "use_missing":
            kw["use_missing"] = v == "true"
        elif k == "feature_pre_filter":
            kw["pre_filter"] = v == "true"
    for c in range(X.shape[1]):
        col = X[:, c]
        vals = col[(np.abs(col) > 1e-35) | np.isnan(col)]
        got = engine.find_bin(vals, n, min_split_data=20, **kw)
        if c not in ref_by_col:
            assert got["is_trivial"], "column %d: the reference dropped it as trivial" % c
            continue
        info = ref_by_col[c]
        assert not got["is_trivial"], "column %d" % c
        for key in ("num_bin", "missing_type", "default_bin", "most_freq_bin"):
            assert got[key] == info[key], "column %d %s: %s vs %s" % (c, key, got[key], info[key])
        real = info["num_bin"] - (1 if info["missing_type"] == 2 else 0)
        np.testing.assert_array_equal(got["upper_bounds"][:real], info["upper_bounds"][:real], err_msg="column %d" % c)


def test_metric_restatement_reproduces_golden():
    """oracle/oracle_metrics.py on the committed fixtures (tests/golden/metrics_*.npz, written by
    tests/golden/make_golden_metrics.py from the reference's own scores and metric values): no reference library needed."""
    from oracle import oracle_metrics as M
    z = np.load(os.path.join(ROOT, "tests", "golden", "metrics_binary.npz"))
    assert list(z["names"]) == ["binary_logloss", "auc", "binary_error"]
    np.testing.assert_allclose(M.binary_logloss(z["prob"], z["label"]), z["values"][0], rtol=1e-13)
    assert M.auc(z["prob"], z["label"]) == z["values"][1]
    assert M.binary_error(z["prob"], z["label"]) == z["values"][2]
    z = np.load(os.path.join(ROOT, "tests", "golden", "metrics_regression.npz"))
    assert list(z["names"]) == ["l2", "rmse", "l1", "auc"]
    np.testing.assert_allclose([M.l2(z["score"], z["label"]), M.rmse(z["score"], z["label"]), M.l1(z["score"], z["label"])],
                               z["values"][:3], rtol=1e-13)
    assert M.auc(z["score"], z["label"]) == z["values"][3]
    z = np.load(os.path.join(ROOT, "tests", "golden", "metrics_rank.npz"))
    qb = np.concatenate(([0], np.cumsum(z["group"])))
    np.testing.assert_allclose(M.ndcg(z["score"], z["label"], qb, list(z["eval_at"])), z["values"], rtol=1e-14)
