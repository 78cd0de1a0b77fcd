"""Evaluation metrics on device-resident scores: the gbl_metric_* kernels against the
oracle's restatement on identical scores, and LGBM_BoosterAddValidData / LGBM_BoosterGetEval end to end against the
unmodified reference on identical data and parameters."""
import numpy as np
import pytest

from lambdagap_b200 import capi, engine
from oracle import oracle_metrics as M
from oracle import ref_lgbm as R

pytestmark = pytest.mark.gpu


def _dev(a):
    return capi.DeviceBuffer.from_numpy(np.ascontiguousarray(a, dtype=np.float64))


@pytest.mark.parametrize("case", ["continuous", "ties", "constant", "one_class", "signed_zero"])
def test_auc_is_exact(gpu, case):
    """AUC is integer arithmetic (tie-group counts): the device value equals the restatement bit for bit."""
    rng = np.random.default_rng(11)
    n = 200003
    y = (rng.random(n) < 0.3).astype(np.float32)
    s = rng.normal(size=n) + 0.8 * y
    if case == "ties":
        s = np.round(s * 8) / 8  # ~60 distinct values: tie groups of thousands of rows
    elif case == "constant":
        s = np.full(n, 0.25)
    elif case == "one_class":
        y[:] = 1
    elif case == "signed_zero":
        s = np.round(s)  # -0.0 and +0.0 both occur and compare equal in the reference
        assert np.any(np.signbit(s) & (s == 0)) and np.any(~np.signbit(s) & (s == 0))
    m = capi.Metric("auc", y)
    sd = _dev(s)
    got = m.eval(sd.data_ptr())[0]
    assert got == M.auc(s, y)
    assert m.eval(sd.data_ptr())[0] == got  # scratch is reused: a second evaluation gives the same value
    m.free()


@pytest.mark.parametrize("kind", ["l2", "rmse", "l1", "binary_logloss", "binary_error"])
@pytest.mark.parametrize("weighted", [False, True])
def test_pointwise_metrics(gpu, kind, weighted):
    rng = np.random.default_rng(12)
    n = 150001
    w = rng.uniform(0.5, 2.0, size=n).astype(np.float32) if weighted else None
    if kind.startswith("binary"):
        y = (rng.random(n) < 0.4).astype(np.float32)
        raw = rng.normal(size=n) * 3
        raw[:5] = [800.0, -800.0, 40.0, -40.0, 0.0]  # saturated probabilities hit the kEpsilon branches
        sigmoid = 1.3
        with np.errstate(over="ignore"):
            prob = 1.0 / (1.0 + np.exp(-sigmoid * raw))
        want = getattr(M, kind)(prob, y, w)
        m = capi.Metric(kind, y, weight=w, sigmoid=sigmoid)
        buf = _dev(raw)  # keep the buffer alive across the call
        got = m.eval(buf.data_ptr())[0]
        m.free()
        m = capi.Metric(kind, y, weight=w, sigmoid=0.0)  # no objective attached: the scores are probabilities
        buf = _dev(prob)
        got_p = m.eval(buf.data_ptr())[0]
        np.testing.assert_allclose(got_p, want, rtol=1e-12)
    else:
        y = rng.normal(size=n).astype(np.float32)
        s = y + rng.normal(size=n) * 0.3
        want = getattr(M, kind)(s, y, w)
        m = capi.Metric(kind, y, weight=w)
        buf = _dev(s)
        got = m.eval(buf.data_ptr())[0]
    m.free()
    np.testing.assert_allclose(got, want, rtol=1e-12)


def test_ndcg_matches_restatement(gpu):
    """Ragged queries (1 document .. 3000), a query without relevant documents, eval_at beyond the query length, tied
    scores (stable order decides): per-query DCG is added in rank order like DCGCalculator::CalDCG."""
    rng = np.random.default_rng(13)
    q = rng.integers(1, 200, size=700)
    q[7] = 3000
    q[8] = 1
    qb = np.concatenate(([0], np.cumsum(q))).astype(np.int32)
    n = int(qb[-1])
    y = np.clip(rng.normal(size=n).round() + 1, 0, 4).astype(np.float32)
    y[qb[20]:qb[21]] = 0
    s = rng.normal(size=n) + 0.5 * y
    s[qb[30]:qb[31]] = np.round(s[qb[30]:qb[31]])  # ties inside a query
    s[qb[31]:qb[32]] = 0.0
    for eval_at in ([1, 2, 3, 4, 5], [10], [3, 50, 4000]):
        m = capi.Metric("ndcg", y, query_boundaries=qb, eval_at=eval_at)
        assert m.num_outputs == len(eval_at)
        buf = _dev(s)
        got = m.eval(buf.data_ptr())
        m.free()
        np.testing.assert_allclose(got, M.ndcg(s, y, qb, eval_at), rtol=1e-14)


def test_device_metrics_reproduce_golden(gpu):
    """The committed fixtures hold the reference's own scores and the metric values it computed on them
    (tests/golden/make_golden_metrics.py): the device metrics must give the same numbers -- AUC exactly."""
    import os
    golden = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
    z = np.load(os.path.join(golden, "metrics_binary.npz"))
    buf = _dev(z["prob"])
    for k, name in enumerate(z["names"]):
        m = capi.Metric(str(name), z["label"], sigmoid=0.0)  # the fixture's scores are probabilities already
        got = m.eval(buf.data_ptr())[0]
        m.free()
        if name == "binary_logloss":
            np.testing.assert_allclose(got, z["values"][k], rtol=1e-11)
        else:
            assert got == z["values"][k], name
    z = np.load(os.path.join(golden, "metrics_regression.npz"))
    buf = _dev(z["score"])
    for k, name in enumerate(z["names"]):
        m = capi.Metric(str(name), z["label"])
        got = m.eval(buf.data_ptr())[0]
        m.free()
        if name == "auc":
            assert got == z["values"][k]
        else:
            np.testing.assert_allclose(got, z["values"][k], rtol=1e-11)
    z = np.load(os.path.join(golden, "metrics_rank.npz"))
    qb = np.concatenate(([0], np.cumsum(z["group"]))).astype(np.int32)
    buf = _dev(z["score"])
    m = capi.Metric("ndcg", z["label"], query_boundaries=qb, eval_at=[int(k) for k in z["eval_at"]])
    np.testing.assert_allclose(m.eval(buf.data_ptr()), z["values"], rtol=1e-13)
    m.free()


def test_metric_argument_errors(gpu):
    y = np.zeros(10, dtype=np.float32)
    with pytest.raises(capi.GblError, match="query information"):
        capi.Metric("ndcg", y)
    with pytest.raises(capi.GblError, match="4096"):
        capi.Metric("ndcg", y, query_boundaries=[0, 10], eval_at=[5000])
    with pytest.raises(capi.GblError, match="label"):
        capi.Metric("ndcg", np.full(10, 40, dtype=np.float32), query_boundaries=[0, 10])
    with pytest.raises(capi.GblError, match="weighted AUC"):
        capi.Metric("auc", y, weight=np.ones(10, dtype=np.float32))


def _both(X, y, Xv, yv, params, iters, group=None, vgroup=None):
    ds_params = "max_bin=255 verbose=-1 num_threads=4"
    rds = R.RefDataset(X, y, group=group, params=ds_params + " force_col_wise=true")
    rdv = R.RefDataset(Xv, yv, group=vgroup, params=ds_params, reference=rds)
    rb = R.RefBooster(rds, params + " device_type=cpu force_col_wise=true deterministic=true num_threads=4 verbose=-1")
    rb.add_valid(rdv)
    ods = engine.Dataset(X, y, group=group, params=ds_params)
    odv = engine.Dataset(Xv, yv, group=vgroup, params=ds_params, reference=ods)
    ob = engine.Booster(ods, params + " device_type=cuda num_threads=4 verbose=-1")
    ob.add_valid(odv)
    evals = []
    for _ in range(iters):
        rb.update()
        ob.update()
        evals.append((rb.eval(0), ob.eval(0), rb.eval(1), ob.eval(1)))
    return rb, ob, evals


def test_valid_data_and_eval_binary(gpu):
    """Per-iteration eval of the training and a validation set: same names, values within the float-mode tolerance
    (identical trees, leaf values within 1e-5); validation scores are built by walking each new tree on the device."""
    rng = np.random.default_rng(14)
    n, nv, f = 40000, 15000, 12
    w = rng.normal(size=f)
    X = rng.normal(size=(n, f)).astype(np.float32)
    Xv = rng.normal(size=(nv, f)).astype(np.float32)
    X[rng.random(n) < 0.05, 3] = np.nan
    Xv[rng.random(nv) < 0.05, 3] = np.nan
    Xv[:50, 0] = 1e6  # values beyond the training range land in the last bin
    y = (np.nan_to_num(X) @ w + rng.normal(size=n) > 0).astype(np.float32)
    yv = (np.nan_to_num(Xv) @ w + rng.normal(size=nv) > 0).astype(np.float32)
    params = "objective=binary metric=binary_logloss,auc,binary_error num_leaves=31 learning_rate=0.1 min_data_in_leaf=20"
    rb, ob, evals = _both(X, y, Xv, yv, params, 8)
    assert ob.eval_names() == rb.eval_names() == ["binary_logloss", "auc", "binary_error"]
    for r0, o0, r1, o1 in evals:
        np.testing.assert_allclose(o0, r0, rtol=2e-5)
        np.testing.assert_allclose(o1, r1, rtol=2e-5)
    np.testing.assert_allclose(ob.valid_scores(1), rb.valid_scores(1), rtol=1e-5, atol=1e-8)
    with pytest.raises(engine.LightGBMError, match="before the first"):
        ob.add_valid(ob._valid[0])


def test_valid_data_and_eval_lambdarank(gpu):
    rng = np.random.default_rng(15)
    f = 10
    w = rng.normal(size=f)

    def make(nq):
        q = rng.integers(5, 80, size=nq).astype(np.int32)
        n = int(q.sum())
        X = rng.normal(size=(n, f)).astype(np.float32)
        y = np.clip((X @ w * 0.6 + rng.normal(size=n)).round() + 1, 0, 4).astype(np.float32)
        return X, y, q

    X, y, q = make(600)
    Xv, yv, qv = make(250)
    params = "objective=lambdarank metric=ndcg eval_at=1,5,10 num_leaves=15 learning_rate=0.1 min_data_in_leaf=20"
    rb, ob, evals = _both(X, y, Xv, yv, params, 6, group=q, vgroup=qv)
    assert ob.eval_names() == rb.eval_names() == ["ndcg@1", "ndcg@5", "ndcg@10"]
    for r0, o0, r1, o1 in evals:
        np.testing.assert_allclose(o0, r0, rtol=1e-4)
        np.testing.assert_allclose(o1, r1, rtol=1e-4)


def test_default_metric_follows_objective(gpu):
    rng = np.random.default_rng(16)
    X = rng.normal(size=(5000, 5)).astype(np.float32)
    y = (X[:, 0] + 0.1 * rng.normal(size=5000)).astype(np.float32)
    ds = engine.Dataset(X, y, params="max_bin=63 verbose=-1")
    b = engine.Booster(ds, "objective=regression num_leaves=7 verbose=-1 device_type=cuda")
    assert b.eval_names() == ["l2"]
    b.update()
    s = b.train_scores(raw=True)
    np.testing.assert_allclose(b.eval(0), [M.l2(s, y)], rtol=1e-12)
    b2 = engine.Booster(ds, "objective=regression metric=rmse,l1,l2 num_leaves=7 verbose=-1 device_type=cuda")
    assert b2.eval_names() == ["rmse", "l1", "l2"]
    with pytest.raises(engine.LightGBMError, match="not evaluated"):
        engine.Booster(ds, "objective=regression metric=mape num_leaves=7 verbose=-1 device_type=cuda")
