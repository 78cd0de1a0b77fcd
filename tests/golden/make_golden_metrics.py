"""Generates tests/golden/metrics_*.npz from the UNMODIFIED reference CPU library (oracle/_ref/lib_lightgbm.so):

    python tests/golden/make_golden_metrics.py

Each fixture holds labels (and query sizes), the scores the reference itself produced after a few boosting iterations
(LGBM_BoosterGetPredict) and the metric values the reference computed on exactly those scores (LGBM_BoosterGetEval).
They pin oracle/oracle_metrics.py on CPU (tests/test_oracle_cpu.py) and the device metrics on the GPU box even where
oracle/_ref is absent (tests/test_metrics.py::test_device_metrics_reproduce_golden).
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import ref_lgbm as R  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))


def main():
    rng = np.random.default_rng(2024)
    n, f = 5000, 8
    X = rng.normal(size=(n, f)).astype(np.float32)
    y = (X[:, 0] + 0.5 * X[:, 1] * X[:, 2] + 0.5 * rng.normal(size=n) > 0).astype(np.float32)
    common = " num_leaves=7 verbose=-1 num_threads=2 min_data_in_leaf=150"
    ds = R.RefDataset(X, y, params="max_bin=63 verbose=-1")
    # binary objective: GetPredict returns probabilities (few distinct values -> large tie groups for the AUC)
    b = R.RefBooster(ds, "objective=binary sigmoid=1.0 metric=binary_logloss,auc,binary_error" + common)
    for _ in range(4):
        b.update()
    np.savez_compressed(os.path.join(HERE, "metrics_binary.npz"), label=y, prob=b.train_scores(), names=np.array(b.eval_names()),
                        values=b.eval(0))
    # regression objective on the same labels: raw scores, point-wise regression metrics + AUC
    b = R.RefBooster(ds, "objective=regression metric=l2,rmse,l1,auc" + common)
    for _ in range(4):
        b.update()
    np.savez_compressed(os.path.join(HERE, "metrics_regression.npz"), label=y, score=b.train_scores(),
                        names=np.array(b.eval_names()), values=b.eval(0))
    # ranking: ragged queries, one without relevant documents, eval_at beyond most query lengths
    q = rng.integers(1, 50, size=160)
    q[3] = 300
    qb = np.concatenate(([0], np.cumsum(q)))
    Xr = rng.normal(size=(qb[-1], f)).astype(np.float32)
    yr = np.clip((Xr[:, 0] * 1.2 + rng.normal(size=qb[-1])).round(), 0, 4).astype(np.float32)
    yr[qb[5]:qb[6]] = 0
    dr = R.RefDataset(Xr, yr, group=q.astype(np.int32), params="max_bin=63 verbose=-1")
    b = R.RefBooster(dr, "objective=lambdarank metric=ndcg eval_at=1,3,5,10,100 num_leaves=7 verbose=-1 num_threads=2")
    for _ in range(3):
        b.update()
    np.savez_compressed(os.path.join(HERE, "metrics_rank.npz"), label=yr, group=q.astype(np.int32), score=b.train_scores(),
                        names=np.array(b.eval_names()), values=b.eval(0), eval_at=np.array([1, 3, 5, 10, 100]))
    print("wrote metrics_binary.npz, metrics_regression.npz, metrics_rank.npz")


if __name__ == "__main__":
    main()
