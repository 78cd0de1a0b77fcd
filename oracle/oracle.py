"""oracle/oracle.py -- TEST INFRASTRUCTURE ONLY.

ctypes front of the C restatement (oracle/liboracle.so, built from oracle/oracle_gbdt.c + oracle_rank.c by
`make -C oracle`) plus a small boosting loop assembled from its pieces the way GBDT::TrainOneIter
(/root/reference/src/boosting/gbdt.cpp:346-454) assembles the reference's.  Only tests/, __graft_entry__.smoke()
and bench.py's cpu_baseline leg may import this; the product never does.
"""
import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "liboracle.so")

RANK_TARGETS = ["ndcg", "lambdaloss-ndcg", "lambdaloss-ndcg-plus-plus", "bndcg", "lambdaloss-bndcg",
                "lambdaloss-bndcg-plus-plus", "precision", "arpk", "lambdaloss-arp1", "lambdaloss-arp2", "ranknet",
                "bin-ranknet", "lambdagap-s", "lambdagap-x", "lambdagap-s-plus", "lambdagap-x-plus",
                "lambdagap-s-plus-plus", "lambdagap-x-plus-plus"]


class Meta(ctypes.Structure):
    _fields_ = [("num_bin", ctypes.c_int32), ("missing_type", ctypes.c_int32), ("default_bin", ctypes.c_int32),
                ("most_freq_bin", ctypes.c_int32), ("real_index", ctypes.c_int32)]


class Config(ctypes.Structure):
    _fields_ = [("num_leaves", ctypes.c_int32), ("max_depth", ctypes.c_int32), ("min_data_in_leaf", ctypes.c_int32),
                ("use_quantized_grad", ctypes.c_int32), ("num_grad_quant_bins", ctypes.c_int32),
                ("is_constant_hessian", ctypes.c_int32), ("min_sum_hessian_in_leaf", ctypes.c_double),
                ("lambda_l1", ctypes.c_double), ("lambda_l2", ctypes.c_double), ("min_gain_to_split", ctypes.c_double),
                ("max_delta_step", ctypes.c_double), ("path_smooth", ctypes.c_double)]


class Split(ctypes.Structure):
    _fields_ = [("leaf", ctypes.c_int32), ("feature", ctypes.c_int32), ("real_feature", ctypes.c_int32),
                ("threshold_bin", ctypes.c_int32), ("default_left", ctypes.c_int32), ("left_count", ctypes.c_int32),
                ("right_count", ctypes.c_int32), ("pad", ctypes.c_int32), ("left_output", ctypes.c_double),
                ("right_output", ctypes.c_double), ("left_sum_gradient", ctypes.c_double),
                ("left_sum_hessian", ctypes.c_double), ("right_sum_gradient", ctypes.c_double),
                ("right_sum_hessian", ctypes.c_double), ("gain", ctypes.c_double), ("left_ig", ctypes.c_int64),
                ("left_ih", ctypes.c_int64), ("right_ig", ctypes.c_int64), ("right_ih", ctypes.c_int64)]


_lib = None


def available():
    return os.path.exists(LIB_PATH)


def lib():
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(LIB_PATH)
        _lib.orc_rank_create.restype = ctypes.c_void_p
        _lib.orc_max_dcg_at_k.restype = ctypes.c_double
        _lib.orc_max_bdcg_at_k.restype = ctypes.c_double
    return _lib


def _p(a):
    return a.ctypes.data_as(ctypes.c_void_p) if a is not None else None


def make_config(**kw):
    c = Config(num_leaves=31, max_depth=-1, min_data_in_leaf=20, use_quantized_grad=0, num_grad_quant_bins=4,
               is_constant_hessian=0, min_sum_hessian_in_leaf=1e-3, lambda_l1=0.0, lambda_l2=0.0,
               min_gain_to_split=0.0, max_delta_step=0.0, path_smooth=0.0)
    for k, v in kw.items():
        if not hasattr(c, k):
            raise KeyError(k)
        setattr(c, k, v)
    return c


def make_metas(infos):
    arr = (Meta * len(infos))()
    for i, d in enumerate(infos):
        arr[i] = Meta(int(d["num_bin"]), int(d["missing_type"]), int(d["default_bin"]), int(d["most_freq_bin"]),
                      int(d.get("real_index", i)))
    return arr


def histogram_float(col, info, idx, g, h):
    m = make_metas([info])
    out = np.zeros((info["num_bin"], 2), dtype=np.float64)
    idx = None if idx is None else np.ascontiguousarray(idx, dtype=np.int32)
    n = len(col) if idx is None else len(idx)
    lib().orc_histogram_float(_p(np.ascontiguousarray(col, dtype=np.uint8)), m, _p(idx), ctypes.c_int(n),
                              _p(np.ascontiguousarray(g, dtype=np.float32)), _p(np.ascontiguousarray(h, dtype=np.float32)),
                              _p(out))
    return out


def discretize(g, h, num_bins=4, is_constant_hessian=False, rand_g=None, rand_h=None, rand_start=0):
    g = np.ascontiguousarray(g, dtype=np.float32)
    h = np.ascontiguousarray(h, dtype=np.float32)
    out = np.zeros(2 * len(g), dtype=np.int8)
    gs, hs = ctypes.c_double(0), ctypes.c_double(0)
    rg = None if rand_g is None else np.ascontiguousarray(rand_g, dtype=np.float64)
    rh = None if rand_h is None else np.ascontiguousarray(rand_h, dtype=np.float64)
    lib().orc_discretize(_p(g), _p(h), ctypes.c_int(len(g)), ctypes.c_int(num_bins), ctypes.c_int(int(is_constant_hessian)),
                         _p(rg), _p(rh), ctypes.c_int(rand_start), _p(out), ctypes.byref(gs), ctypes.byref(hs))
    return out, gs.value, hs.value


def train_tree(cols, infos, g, h, config):
    """SerialTreeLearner::Train restated.  Returns (root_output, [split dicts], leaf_of_row)."""
    cols = [np.ascontiguousarray(c, dtype=np.uint8) for c in cols]
    n, f = len(cols[0]), len(cols)
    g = np.ascontiguousarray(g, dtype=np.float32)
    h = np.ascontiguousarray(h, dtype=np.float32)
    ptrs = (ctypes.c_void_p * f)(*[c.ctypes.data for c in cols])
    metas = make_metas(infos)
    gh8, gs, hs = None, 0.0, 0.0
    if config.use_quantized_grad:
        gh8, gs, hs = discretize(g, h, config.num_grad_quant_bins, bool(config.is_constant_hessian))
    splits = (Split * max(1, config.num_leaves))()
    ns = ctypes.c_int32(0)
    root = ctypes.c_double(0)
    leaf_of_row = np.zeros(n, dtype=np.int32)
    r = lib().orc_train_tree(ptrs, metas, ctypes.c_int(f), ctypes.c_int(n), _p(g), _p(h), _p(gh8), ctypes.c_double(gs),
                             ctypes.c_double(hs), ctypes.byref(config), splits, ctypes.byref(ns), ctypes.byref(root),
                             _p(leaf_of_row))
    assert r == 0
    recs = [{name: getattr(splits[k], name) for name, _ in Split._fields_} for k in range(ns.value)]
    return root.value, recs, leaf_of_row


class RankObjective:
    def __init__(self, target="ndcg", truncation_level=30, norm=True, sigmoid=1.0, lambdagap_weight=1.0):
        self.h = ctypes.c_void_p(lib().orc_rank_create(ctypes.c_int(RANK_TARGETS.index(target)), ctypes.c_int(truncation_level),
                                                       ctypes.c_int(int(norm)), ctypes.c_double(sigmoid),
                                                       ctypes.c_double(lambdagap_weight), None, ctypes.c_int(0)))

    def gradients(self, score, label, query_boundaries, weight=None):
        score = np.ascontiguousarray(score, dtype=np.float64)
        label = np.ascontiguousarray(label, dtype=np.float32)
        qb = np.ascontiguousarray(query_boundaries, dtype=np.int32)
        w = None if weight is None else np.ascontiguousarray(weight, dtype=np.float32)
        g = np.zeros(len(score), dtype=np.float32)
        h = np.zeros(len(score), dtype=np.float32)
        lib().orc_lambdarank_gradients(self.h, _p(score), _p(label), _p(w), _p(qb), ctypes.c_int(len(qb) - 1), _p(g), _p(h))
        return g, h

    def free(self):
        if self.h:
            lib().orc_rank_free(self.h)
            self.h = None


def binary_gradients(score, label, weight=None, sigmoid=1.0):
    score = np.ascontiguousarray(score, dtype=np.float64)
    label = np.ascontiguousarray(label, dtype=np.float32)
    w = None if weight is None else np.ascontiguousarray(weight, dtype=np.float32)
    g = np.zeros(len(score), dtype=np.float32)
    h = np.zeros(len(score), dtype=np.float32)
    lib().orc_binary_gradients(_p(score), _p(label), _p(w), ctypes.c_int(len(score)), ctypes.c_double(sigmoid),
                               ctypes.c_double(1.0), ctypes.c_double(1.0), _p(g), _p(h))
    return g, h


def l2_gradients(score, label, weight=None):
    score = np.ascontiguousarray(score, dtype=np.float64)
    label = np.ascontiguousarray(label, dtype=np.float32)
    w = None if weight is None else np.ascontiguousarray(weight, dtype=np.float32)
    g = np.zeros(len(score), dtype=np.float32)
    h = np.zeros(len(score), dtype=np.float32)
    lib().orc_l2_gradients(_p(score), _p(label), _p(w), ctypes.c_int(len(score)), _p(g), _p(h))
    return g, h


def boost(cols, infos, label, objective, config, iterations, learning_rate=0.1, group=None, weight=None, rank=None):
    """GBDT::TrainOneIter loop restated: BoostFromAverage, gradients, tree, shrinkage, score update.
    Returns (list of (root_output, splits, init_score_added), final raw scores)."""
    n = len(label)
    label = np.asarray(label, dtype=np.float32)
    score = np.zeros(n, dtype=np.float64)
    init = 0.0
    if objective == "binary":
        pavg = float((label > 0).sum()) / n if weight is None else float(((label > 0) * weight).sum() / weight.sum())
        pavg = min(pavg, 1.0 - 1e-15)
        pavg = max(pavg, 1e-15)
        init = float(np.log(pavg / (1.0 - pavg)))
    elif objective == "regression":
        init = float(label.astype(np.float64).sum() / n) if weight is None else float(
            (label.astype(np.float64) * weight).sum() / weight.astype(np.float64).sum())
    if abs(init) > 1e-15:
        score += init
    qb = None if group is None else np.concatenate([[0], np.cumsum(group)]).astype(np.int32)
    trees = []
    for it in range(iterations):
        if objective == "binary":
            g, h = binary_gradients(score, label, weight)
        elif objective == "regression":
            g, h = l2_gradients(score, label, weight)
        else:
            g, h = rank.gradients(score, label, qb, weight)
        root, recs, leaf_of_row = train_tree(cols, infos, g, h, config)
        if not recs:
            break
        vals = np.zeros(len(recs) + 1, dtype=np.float64)
        vals[0] = root
        for k, r in enumerate(recs):
            vals[r["leaf"]] = r["left_output"]
            vals[k + 1] = r["right_output"]
        score += (vals * learning_rate)[leaf_of_row]
        trees.append((root, recs, init if it == 0 else 0.0))
    return trees, score
