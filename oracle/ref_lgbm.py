"""oracle/ref_lgbm.py -- TEST INFRASTRUCTURE ONLY.

ctypes harness around the UNMODIFIED reference CPU library built by oracle/Makefile.ref
(oracle/_ref/lib_lightgbm.so; the reference's own C API, include/LightGBM/c_api.h) plus the read-only taps
of oracle/ref_tap.cpp.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this module; the product (lambdagap_b200/) never does.
"""
import ctypes
import os
import re

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_REF_DIR = os.path.join(_HERE, "_ref")
C_API_DTYPE_FLOAT32 = 0  # include/LightGBM/c_api.h:30-33
C_API_DTYPE_FLOAT64 = 1
C_API_DTYPE_INT32 = 2

_lib = None
_tap = None


def available():
    return os.path.exists(os.path.join(_REF_DIR, "lib_lightgbm.so"))


def lib():
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(os.path.join(_REF_DIR, "lib_lightgbm.so"), mode=ctypes.RTLD_GLOBAL)
        _lib.LGBM_GetLastError.restype = ctypes.c_char_p
    return _lib


def tap():
    global _tap
    if _tap is None:
        lib()
        _tap = ctypes.CDLL(os.path.join(_REF_DIR, "libref_tap.so"))
    return _tap


def _check(ret):
    if ret != 0:
        raise RuntimeError("reference error: %s" % lib().LGBM_GetLastError().decode())


def _params(d):
    if isinstance(d, (bytes, str)):
        return d if isinstance(d, bytes) else d.encode()
    return " ".join("%s=%s" % (k, v) for k, v in d.items()).encode()


class RefDataset:
    """LGBM_DatasetCreateFromMat (src/c_api.cpp:1303) + LGBM_DatasetSetField (:1760)."""

    def __init__(self, X, label=None, group=None, weight=None, params=None, reference=None):
        X = np.ascontiguousarray(X)
        assert X.dtype in (np.float32, np.float64)
        self.handle = ctypes.c_void_p()
        self.n, self.f = X.shape
        dt = C_API_DTYPE_FLOAT32 if X.dtype == np.float32 else C_API_DTYPE_FLOAT64
        _check(lib().LGBM_DatasetCreateFromMat(
            X.ctypes.data_as(ctypes.c_void_p), ctypes.c_int(dt), ctypes.c_int32(self.n), ctypes.c_int32(self.f),
            ctypes.c_int(1), ctypes.c_char_p(_params(params or {})), reference.handle if reference is not None else None,
            ctypes.byref(self.handle)))
        if label is not None:
            self.set_field("label", np.asarray(label, dtype=np.float32))
        if weight is not None:
            self.set_field("weight", np.asarray(weight, dtype=np.float32))
        if group is not None:
            self.set_field("group", np.asarray(group, dtype=np.int32))

    def set_field(self, name, arr):
        arr = np.ascontiguousarray(arr)
        dt = {np.dtype(np.float32): C_API_DTYPE_FLOAT32, np.dtype(np.float64): C_API_DTYPE_FLOAT64,
              np.dtype(np.int32): C_API_DTYPE_INT32}[arr.dtype]
        _check(lib().LGBM_DatasetSetField(self.handle, ctypes.c_char_p(name.encode()),
                                          arr.ctypes.data_as(ctypes.c_void_p), ctypes.c_int(len(arr)), ctypes.c_int(dt)))

    # ---- taps (oracle/ref_tap.cpp) ----
    def num_used_features(self):
        return tap().RefTap_NumFeatures(self.handle)

    def feature_info(self, inner):
        real, nb, mt, db, mfb = (ctypes.c_int() for _ in range(5))
        ub = np.zeros(4096, dtype=np.float64)
        r = tap().RefTap_FeatureInfo(self.handle, ctypes.c_int(inner), ctypes.byref(real), ctypes.byref(nb),
                                     ctypes.byref(mt), ctypes.byref(db), ctypes.byref(mfb),
                                     ub.ctypes.data_as(ctypes.c_void_p), ctypes.c_int(len(ub)))
        assert r == 0, r
        return dict(real_index=real.value, num_bin=nb.value, missing_type=mt.value, default_bin=db.value,
                    most_freq_bin=mfb.value, upper_bounds=ub[:nb.value].copy())

    def feature_bins(self, inner):
        out = np.zeros(self.n, dtype=np.uint8)
        r = tap().RefTap_FeatureBins(self.handle, ctypes.c_int(inner), out.ctypes.data_as(ctypes.c_void_p))
        assert r == 0, r
        return out

    def free(self):
        if self.handle:
            lib().LGBM_DatasetFree(self.handle)
            self.handle = None


class RefBooster:
    """LGBM_BoosterCreate (src/c_api.cpp:1953) / LGBM_BoosterUpdateOneIter (:2076) / GetPredict / SaveModelToString."""

    def __init__(self, dataset, params):
        self.ds = dataset
        self.handle = ctypes.c_void_p()
        _check(lib().LGBM_BoosterCreate(dataset.handle, ctypes.c_char_p(_params(params)), ctypes.byref(self.handle)))

    def update(self):
        fin = ctypes.c_int(0)
        _check(lib().LGBM_BoosterUpdateOneIter(self.handle, ctypes.byref(fin)))
        return fin.value

    def update_custom(self, grad, hess):
        fin = ctypes.c_int(0)
        grad = np.ascontiguousarray(grad, dtype=np.float32)
        hess = np.ascontiguousarray(hess, dtype=np.float32)
        _check(lib().LGBM_BoosterUpdateOneIterCustom(self.handle, grad.ctypes.data_as(ctypes.c_void_p),
                                                     hess.ctypes.data_as(ctypes.c_void_p), ctypes.byref(fin)))
        return fin.value

    def reset_parameter(self, params):
        """LGBM_BoosterResetParameter (src/c_api.cpp:2053)."""
        _check(lib().LGBM_BoosterResetParameter(self.handle, ctypes.c_char_p(_params(params))))

    def feature_importance(self, importance_type=0, num_iteration=-1):
        n = ctypes.c_int(0)
        _check(lib().LGBM_BoosterGetNumFeature(self.handle, ctypes.byref(n)))
        out = np.zeros(n.value, dtype=np.float64)
        _check(lib().LGBM_BoosterFeatureImportance(self.handle, ctypes.c_int(num_iteration), ctypes.c_int(importance_type),
                                                   out.ctypes.data_as(ctypes.c_void_p)))
        return out

    def add_valid(self, valid_set):
        """LGBM_BoosterAddValidData (src/c_api.cpp:2040)."""
        _check(lib().LGBM_BoosterAddValidData(self.handle, valid_set.handle))
        self._valid = getattr(self, "_valid", []) + [valid_set]

    def eval_names(self):
        n = ctypes.c_int(0)
        _check(lib().LGBM_BoosterGetEvalCounts(self.handle, ctypes.byref(n)))
        bufs = [ctypes.create_string_buffer(255) for _ in range(n.value)]
        ptrs = (ctypes.c_char_p * n.value)(*map(ctypes.addressof, bufs))
        out_n = ctypes.c_int(0)
        need = ctypes.c_size_t(0)
        _check(lib().LGBM_BoosterGetEvalNames(self.handle, ctypes.c_int(n.value), ctypes.byref(out_n), ctypes.c_size_t(255),
                                              ctypes.byref(need), ptrs))
        return [b.value.decode() for b in bufs]

    def eval(self, data_idx):
        """LGBM_BoosterGetEval (src/c_api.cpp:2200): metric values on the training data (0) or a validation set."""
        n = ctypes.c_int(0)
        _check(lib().LGBM_BoosterGetEvalCounts(self.handle, ctypes.byref(n)))
        out = np.zeros(max(n.value, 1), dtype=np.float64)
        got = ctypes.c_int(0)
        _check(lib().LGBM_BoosterGetEval(self.handle, ctypes.c_int(data_idx), ctypes.byref(got),
                                         out.ctypes.data_as(ctypes.c_void_p)))
        return out[:got.value]

    def valid_scores(self, data_idx):
        n = ctypes.c_int64(0)
        _check(lib().LGBM_BoosterGetNumPredict(self.handle, ctypes.c_int(data_idx), ctypes.byref(n)))
        out = np.zeros(n.value, dtype=np.float64)
        _check(lib().LGBM_BoosterGetPredict(self.handle, ctypes.c_int(data_idx), ctypes.byref(n),
                                            out.ctypes.data_as(ctypes.c_void_p)))
        return out

    def train_scores(self):
        n = ctypes.c_int64(0)
        _check(lib().LGBM_BoosterGetNumPredict(self.handle, ctypes.c_int(0), ctypes.byref(n)))
        out = np.zeros(n.value, dtype=np.float64)
        _check(lib().LGBM_BoosterGetPredict(self.handle, ctypes.c_int(0), ctypes.byref(n),
                                            out.ctypes.data_as(ctypes.c_void_p)))
        return out

    def train_scores_raw(self):
        """Raw (unconverted) training scores: GetPredict applies ConvertOutput, so invert the sigmoid for binary."""
        s = self.train_scores()
        if self._objective() == "binary":
            s = np.log(s / (1.0 - s))
        return s

    def _objective(self):
        for line in self.model_string().split("\n"):
            if line.startswith("objective="):
                return line.split("=", 1)[1].split()[0]
        return ""

    @classmethod
    def from_model_string(cls, text):
        self = cls.__new__(cls)
        self.ds = None
        self.handle = ctypes.c_void_p()
        it = ctypes.c_int(0)
        _check(lib().LGBM_BoosterLoadModelFromString(ctypes.c_char_p(text.encode()), ctypes.byref(it), ctypes.byref(self.handle)))
        return self

    def predict(self, X, raw_score=False):
        X = np.ascontiguousarray(X, dtype=np.float64)
        out = np.zeros(X.shape[0], dtype=np.float64)
        n = ctypes.c_int64(0)
        _check(lib().LGBM_BoosterPredictForMat(self.handle, X.ctypes.data_as(ctypes.c_void_p), ctypes.c_int(C_API_DTYPE_FLOAT64),
                                               ctypes.c_int32(X.shape[0]), ctypes.c_int32(X.shape[1]), ctypes.c_int(1),
                                               ctypes.c_int(1 if raw_score else 0), ctypes.c_int(0), ctypes.c_int(-1),
                                               ctypes.c_char_p(b"num_threads=4"), ctypes.byref(n), out.ctypes.data_as(ctypes.c_void_p)))
        return out

    def model_string(self):
        need = ctypes.c_int64(0)
        buf = ctypes.create_string_buffer(1 << 20)
        _check(lib().LGBM_BoosterSaveModelToString(self.handle, ctypes.c_int(0), ctypes.c_int(-1), ctypes.c_int(0),
                                                   ctypes.c_int64(len(buf)), ctypes.byref(need), buf))
        if need.value > len(buf):
            buf = ctypes.create_string_buffer(need.value)
            _check(lib().LGBM_BoosterSaveModelToString(self.handle, ctypes.c_int(0), ctypes.c_int(-1), ctypes.c_int(0),
                                                       ctypes.c_int64(len(buf)), ctypes.byref(need), buf))
        return buf.value.decode()

    def free(self):
        if self.handle:
            lib().LGBM_BoosterFree(self.handle)
            self.handle = None


_TREE_INT = ("split_feature", "decision_type", "left_child", "right_child", "leaf_count", "internal_count")
_TREE_FLT = ("split_gain", "threshold", "leaf_value", "leaf_weight", "internal_value", "internal_weight")


def parse_model_trees(text):
    """Parses the `Tree=k` blocks of a LightGBM model string (src/io/tree.cpp:339-395 Tree::ToString)."""
    trees = []
    for block in re.split(r"\nTree=\d+\n", text)[1:]:
        block = block.split("\n\n")[0]
        t = {}
        for line in block.strip().split("\n"):
            if "=" not in line:
                continue
            k, v = line.split("=", 1)
            if k in ("num_leaves", "num_cat"):
                t[k] = int(v)
            elif k == "shrinkage":
                t[k] = float(v)
            elif k in _TREE_INT:
                t[k] = np.array([int(x) for x in v.split()], dtype=np.int64)
            elif k in _TREE_FLT:
                t[k] = np.array([float(x) for x in v.split()], dtype=np.float64)
        trees.append(t)
    return trees
