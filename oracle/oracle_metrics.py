"""oracle/oracle_metrics.py -- TEST INFRASTRUCTURE ONLY (imported by tests/ alone; the product never does).

CPU restatement of the reference's evaluation metrics, sequential like the reference's own loops so that the sums
are added in the reference's order.  Pinned against the unmodified reference (oracle/_ref) by
tests/test_oracle_cpu.py::test_metric_restatement_matches_reference: the reference evaluates a metric on its own
scores (LGBM_BoosterGetEval), this file evaluates it on the same scores (LGBM_BoosterGetPredict).

  auc             AUCMetric::Eval              /root/reference/src/metric/binary_metric.hpp:192-247
  binary_logloss  BinaryLoglossMetric          /root/reference/src/metric/binary_metric.hpp:104-125
  binary_error    BinaryErrorMetric            /root/reference/src/metric/binary_metric.hpp:131-150
  l2 / rmse / l1  RegressionMetric             /root/reference/src/metric/regression_metric.hpp:58-95,119-183
  ndcg@k          NDCGMetric::Eval             /root/reference/src/metric/rank_metric.hpp:86-163
                  DCGCalculator::CalMaxDCG/CalDCG  /root/reference/src/metric/dcg_calculator.cpp:96-156
"""
import math

import numpy as np

K_EPSILON = float(np.float32(1e-15))


def default_label_gain():
    return [0.0] + [float((1 << i) - 1) for i in range(1, 31)]  # dcg_calculator.cpp:33-43


def auc(score, label):
    """Unweighted AUC over tie groups, in descending score order (binary_metric.hpp:192-247)."""
    score = np.asarray(score, dtype=np.float64)
    pos = np.asarray(label) > 0
    order = np.argsort(-score, kind="stable")
    s = score[order]
    p = pos[order]
    n = len(s)
    starts = np.flatnonzero(np.concatenate(([True], s[1:] != s[:-1])))
    ends = np.concatenate((starts[1:], [n]))
    cpos = np.concatenate(([0], np.cumsum(p)))
    accum = 0.0
    sum_pos = 0.0
    for a, b in zip(starts, ends):  # every addend is a half-integer: the double sums are exact
        cur_pos = float(cpos[b] - cpos[a])
        cur_neg = float((b - a)) - cur_pos
        accum += cur_neg * (cur_pos * 0.5 + sum_pos)
        sum_pos += cur_pos
    if sum_pos > 0.0 and sum_pos != float(n):
        return accum / (sum_pos * (float(n) - sum_pos))
    return 1.0


def binary_logloss(prob, label, weight=None):
    prob = np.asarray(prob, dtype=np.float64)
    label = np.asarray(label)
    loss = np.full(len(prob), -math.log(K_EPSILON))
    neg = label <= 0
    ok = neg & (1.0 - prob > K_EPSILON)
    loss[ok] = -np.log(1.0 - prob[ok])
    ok = (~neg) & (prob > K_EPSILON)
    loss[ok] = -np.log(prob[ok])
    return _mean(loss, weight)


def binary_error(prob, label, weight=None):
    prob = np.asarray(prob, dtype=np.float64)
    label = np.asarray(label)
    loss = np.where(prob <= 0.5, label > 0, label <= 0).astype(np.float64)
    return _mean(loss, weight)


def l2(score, label, weight=None):
    d = np.asarray(score, dtype=np.float64) - np.asarray(label, dtype=np.float32).astype(np.float64)
    return _mean(d * d, weight)


def rmse(score, label, weight=None):
    return math.sqrt(l2(score, label, weight))


def l1(score, label, weight=None):
    d = np.asarray(score, dtype=np.float64) - np.asarray(label, dtype=np.float32).astype(np.float64)
    return _mean(np.abs(d), weight)


def _mean(loss, weight):
    if weight is None:
        return math.fsum(loss) / float(len(loss))
    w = np.asarray(weight, dtype=np.float32).astype(np.float64)
    sw = 0.0
    for x in np.asarray(weight, dtype=np.float32):  # `sum_weights_ += weights_[i]`, sequential
        sw += float(x)
    return math.fsum(loss * w) / sw


def ndcg(score, label, query_boundaries, eval_at, label_gain=None):
    """Mean NDCG@k over queries; a query without relevant documents counts 1 (rank_metric.hpp:104-120)."""
    score = np.asarray(score, dtype=np.float64)
    label = np.asarray(label).astype(np.int64)
    gain = np.asarray(default_label_gain() if label_gain is None else label_gain, dtype=np.float64)
    discount = 1.0 / np.log2(2.0 + np.arange(10000, dtype=np.float64))
    eval_at = list(eval_at)
    total = [[] for _ in eval_at]
    for q in range(len(query_boundaries) - 1):
        a, b = int(query_boundaries[q]), int(query_boundaries[q + 1])
        cnt = b - a
        lab = label[a:b]
        # CalMaxDCG: labels from the top down
        ideal = np.sort(lab)[::-1]
        max_dcg = []
        cur, left = 0.0, 0
        for k in eval_at:
            ck = min(k, cnt)
            for j in range(left, ck):
                cur += discount[j] * gain[ideal[j]]
            max_dcg.append(cur)
            left = ck
        if not max_dcg[0] > 0.0:
            for t in total:
                t.append(1.0)
            continue
        order = np.argsort(-score[a:b], kind="stable")  # std::stable_sort by score, descending
        cur, left = 0.0, 0
        for i, k in enumerate(eval_at):
            ck = min(k, cnt)
            for j in range(left, ck):
                cur += gain[lab[order[j]]] * discount[j]
            inv = 1.0 / max_dcg[i] if max_dcg[i] > 0.0 else -1.0
            total[i].append(cur * inv)
            left = ck
    nq = float(len(query_boundaries) - 1)
    return np.array([math.fsum(t) / nq for t in total])
