"""Python front of the host library (lambdagap_b200/lib/liblgbm_b200.so): the reference's C API names
(include/lgbm_b200_c_api.h <-> include/LightGBM/c_api.h) through ctypes, the way the reference's own
python-package/lightgbm/basic.py binds lib_lightgbm.so.  No compute here; no CPU fallback anywhere below.
"""
import ctypes
import os

import numpy as np

from . import capi

_HERE = os.path.dirname(os.path.abspath(__file__))
HOST_LIB_PATH = os.path.join(capi.LIB_DIR, "liblgbm_b200.so")
C_API_DTYPE_FLOAT32, C_API_DTYPE_FLOAT64, C_API_DTYPE_INT32 = 0, 1, 2

LGBM_EXPORTED_SYMBOLS = [
    "LGBM_GetLastError", "LGBM_DatasetCreateFromMat", "LGBM_DatasetSetField", "LGBM_DatasetGetNumData",
    "LGBM_DatasetGetNumFeature", "LGBM_DatasetFree", "LGBM_BoosterCreate", "LGBM_BoosterFree",
    "LGBM_BoosterUpdateOneIter", "LGBM_BoosterUpdateOneIterCustom", "LGBM_BoosterGetCurrentIteration",
    "LGBM_BoosterNumberOfTotalModel", "LGBM_BoosterGetNumPredict", "LGBM_BoosterGetPredict",
    "LGBM_BoosterPredictForMat", "LGBM_BoosterSaveModelToString", "LGBM_NetworkGetNcclUniqueId",
    "LGBM_NetworkInitWithNccl", "LGBM_NetworkFree", "LGBM_BoosterGetTrainScoreRaw",
    "LGBM_DatasetCreateFromMatWithSample", "LGBM_BoosterGetLearnerHandle", "LGBM_B200_FindBin",
    "LGBM_B200_DumpConfig", "LGBM_BoosterResetParameter", "LGBM_BoosterGetNumClasses", "LGBM_BoosterNumModelPerIteration",
    "LGBM_BoosterGetNumFeature", "LGBM_BoosterGetLeafValue", "LGBM_BoosterFeatureImportance", "LGBM_BoosterSaveModel",
    "LGBM_BoosterAddValidData", "LGBM_BoosterGetEvalCounts", "LGBM_BoosterGetEvalNames", "LGBM_BoosterGetEval",
]

_lib = None


def lib():
    global _lib
    if _lib is None:
        capi.lib()  # the CUDA library first: fails loudly when it is missing
        if not os.path.exists(HOST_LIB_PATH):
            raise capi.GblError("%s not found: run __graft_entry__.build() (make -C lambdagap_b200/host)" % HOST_LIB_PATH)
        _lib = ctypes.CDLL(HOST_LIB_PATH)
        _lib.LGBM_GetLastError.restype = ctypes.c_char_p
    return _lib


class LightGBMError(RuntimeError):
    pass


def _check(ret):
    if ret != 0:
        raise LightGBMError(lib().LGBM_GetLastError().decode())


def _params(p):
    if p is None:
        return b""
    if isinstance(p, bytes):
        return p
    if isinstance(p, str):
        return p.encode()
    return " ".join("%s=%s" % (k, v) for k, v in p.items()).encode()


class Dataset:
    def __init__(self, X, label=None, group=None, weight=None, params=None, reference=None, bin_sample=None,
                 num_total_rows=None):
        X = np.ascontiguousarray(X)
        if X.dtype not in (np.float32, np.float64):
            X = X.astype(np.float32)
        self.n, self.f = X.shape
        self.handle = ctypes.c_void_p()
        dt = C_API_DTYPE_FLOAT32 if X.dtype == np.float32 else C_API_DTYPE_FLOAT64
        if bin_sample is not None:
            S = np.ascontiguousarray(bin_sample, dtype=X.dtype)
            _check(lib().LGBM_DatasetCreateFromMatWithSample(
                X.ctypes.data_as(ctypes.c_void_p), ctypes.c_int(dt), ctypes.c_int32(self.n), ctypes.c_int32(self.f),
                ctypes.c_int(1), S.ctypes.data_as(ctypes.c_void_p), ctypes.c_int32(S.shape[0]),
                ctypes.c_int64(num_total_rows if num_total_rows is not None else self.n),
                ctypes.c_char_p(_params(params)), ctypes.byref(self.handle)))
        else:
            _check(lib().LGBM_DatasetCreateFromMat(
                X.ctypes.data_as(ctypes.c_void_p), ctypes.c_int(dt), ctypes.c_int32(self.n), ctypes.c_int32(self.f),
                ctypes.c_int(1), ctypes.c_char_p(_params(params)), reference.handle if reference is not None else None,
                ctypes.byref(self.handle)))
        if label is not None:
            self.set_field("label", np.asarray(label, dtype=np.float32))
        if weight is not None:
            self.set_field("weight", np.asarray(weight, dtype=np.float32))
        if group is not None:
            self.set_field("group", np.asarray(group, dtype=np.int32))

    @classmethod
    def from_pointer(cls, ptr, nrow, ncol, label=None, group=None, params=None, bin_sample=None, num_total_rows=None):
        """Row-major float32 matrix given by address (host or DEVICE memory: the C-ABI layer copies with
        cudaMemcpyDefault); the bin mappers come from `bin_sample` (a host matrix), which is required here."""
        self = cls.__new__(cls)
        self.n, self.f = int(nrow), int(ncol)
        self.handle = ctypes.c_void_p()
        S = np.ascontiguousarray(bin_sample, dtype=np.float32)
        _check(lib().LGBM_DatasetCreateFromMatWithSample(
            ctypes.c_void_p(int(ptr)), ctypes.c_int(C_API_DTYPE_FLOAT32), ctypes.c_int32(self.n), ctypes.c_int32(self.f),
            ctypes.c_int(1), S.ctypes.data_as(ctypes.c_void_p), ctypes.c_int32(S.shape[0]),
            ctypes.c_int64(num_total_rows if num_total_rows is not None else self.n),
            ctypes.c_char_p(_params(params)), ctypes.byref(self.handle)))
        if label is not None:
            self.set_field("label", np.asarray(label, dtype=np.float32))
        if group is not None:
            self.set_field("group", np.asarray(group, dtype=np.int32))
        return self

    def set_field(self, name, arr):
        arr = np.ascontiguousarray(arr)
        dt = {np.dtype(np.float32): C_API_DTYPE_FLOAT32, np.dtype(np.float64): C_API_DTYPE_FLOAT64,
              np.dtype(np.int32): C_API_DTYPE_INT32}[arr.dtype]
        _check(lib().LGBM_DatasetSetField(self.handle, ctypes.c_char_p(name.encode()),
                                          arr.ctypes.data_as(ctypes.c_void_p), ctypes.c_int(len(arr)), ctypes.c_int(dt)))

    def free(self):
        if self.handle:
            lib().LGBM_DatasetFree(self.handle)
            self.handle = None


def dump_config(params):
    """Config::Set of the host layer applied to a parameter string -> dict of the `[key: value]` lines (no GPU needed)."""
    buf = ctypes.create_string_buffer(4096)
    n = ctypes.c_int64(0)
    _check(lib().LGBM_B200_DumpConfig(ctypes.c_char_p(_params(params)), ctypes.c_int64(len(buf)), ctypes.byref(n), buf))
    out = {}
    for line in buf.value.decode().split("\n"):
        if line.startswith("["):
            k, v = line[1:-1].split(": ", 1) if ": " in line else (line[1:-2], "")
            out[k] = v
    return out


class Booster:
    def __init__(self, train_set, params):
        self.train_set = train_set
        self.handle = ctypes.c_void_p()
        _check(lib().LGBM_BoosterCreate(train_set.handle, ctypes.c_char_p(_params(params)), ctypes.byref(self.handle)))

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
        _check(lib().LGBM_BoosterResetParameter(self.handle, ctypes.c_char_p(_params(params))))

    def feature_importance(self, importance_type=0, num_iteration=-1):
        n = ctypes.c_int(0)
        _check(lib().LGBM_BoosterGetNumFeature(self.handle, ctypes.byref(n)))
        out = np.zeros(n.value, dtype=np.float64)
        _check(lib().LGBM_BoosterFeatureImportance(self.handle, ctypes.c_int(num_iteration), ctypes.c_int(importance_type),
                                                   out.ctypes.data_as(ctypes.c_void_p)))
        return out

    def save_model(self, filename):
        _check(lib().LGBM_BoosterSaveModel(self.handle, ctypes.c_int(0), ctypes.c_int(-1), ctypes.c_int(0),
                                           ctypes.c_char_p(filename.encode())))

    def leaf_value(self, tree_idx, leaf_idx):
        out = ctypes.c_double(0)
        _check(lib().LGBM_BoosterGetLeafValue(self.handle, ctypes.c_int(tree_idx), ctypes.c_int(leaf_idx), ctypes.byref(out)))
        return out.value

    def add_valid(self, valid_set):
        """LGBM_BoosterAddValidData; `valid_set` must have been created with reference=<the training set>."""
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
        """LGBM_BoosterGetEval: metric values on the training data (0) or the data_idx-th validation set."""
        n = ctypes.c_int(0)
        _check(lib().LGBM_BoosterGetEvalCounts(self.handle, ctypes.byref(n)))
        out = np.zeros(max(n.value, 1), dtype=np.float64)
        got = ctypes.c_int(0)
        _check(lib().LGBM_BoosterGetEval(self.handle, ctypes.c_int(data_idx), ctypes.byref(got),
                                         out.ctypes.data_as(ctypes.c_void_p)))
        return out[:got.value]

    def valid_scores(self, data_idx, raw=False):
        n = ctypes.c_int64(0)
        _check(lib().LGBM_BoosterGetNumPredict(self.handle, ctypes.c_int(data_idx), ctypes.byref(n)))
        out = np.zeros(n.value, dtype=np.float64)
        _check(lib().LGBM_BoosterGetPredict(self.handle, ctypes.c_int(data_idx), ctypes.byref(n),
                                            out.ctypes.data_as(ctypes.c_void_p)))
        return out

    def learner(self):
        """The C-ABI learner behind this booster, wrapped for timers / profiling (capi.Learner methods)."""
        h = ctypes.c_void_p()
        _check(lib().LGBM_BoosterGetLearnerHandle(self.handle, ctypes.byref(h)))
        lrn = capi.Learner.__new__(capi.Learner)
        lrn.handle = h
        lrn.ds = None
        lrn.cfg = None
        return lrn

    def current_iteration(self):
        out = ctypes.c_int(0)
        _check(lib().LGBM_BoosterGetCurrentIteration(self.handle, ctypes.byref(out)))
        return out.value

    def train_scores(self, raw=True):
        n = ctypes.c_int64(0)
        _check(lib().LGBM_BoosterGetNumPredict(self.handle, ctypes.c_int(0), ctypes.byref(n)))
        out = np.zeros(n.value, dtype=np.float64)
        if raw:
            _check(lib().LGBM_BoosterGetTrainScoreRaw(self.handle, ctypes.byref(n), out.ctypes.data_as(ctypes.c_void_p)))
        else:
            _check(lib().LGBM_BoosterGetPredict(self.handle, ctypes.c_int(0), ctypes.byref(n),
                                                out.ctypes.data_as(ctypes.c_void_p)))
        return out

    def predict(self, X, raw_score=False):
        X = np.ascontiguousarray(X)
        if X.dtype not in (np.float32, np.float64):
            X = X.astype(np.float64)
        dt = C_API_DTYPE_FLOAT32 if X.dtype == np.float32 else C_API_DTYPE_FLOAT64
        out = np.zeros(X.shape[0], dtype=np.float64)
        n = ctypes.c_int64(0)
        _check(lib().LGBM_BoosterPredictForMat(self.handle, X.ctypes.data_as(ctypes.c_void_p), ctypes.c_int(dt),
                                               ctypes.c_int32(X.shape[0]), ctypes.c_int32(X.shape[1]), ctypes.c_int(1),
                                               ctypes.c_int(1 if raw_score else 0), ctypes.c_int(0), ctypes.c_int(-1),
                                               ctypes.c_char_p(b""), ctypes.byref(n), out.ctypes.data_as(ctypes.c_void_p)))
        return out

    def model_to_string(self):
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


def find_bin(values, total_sample_cnt, max_bin=255, min_data_in_bin=3, min_split_data=20, pre_filter=True,
             use_missing=True, zero_as_missing=False):
    """Host BinMapper::FindBin on one column's sampled non-zero values (no GPU involved)."""
    v = np.ascontiguousarray(values, dtype=np.float64).copy()
    sc = (ctypes.c_int32 * 5)()
    ub = np.zeros(max_bin + 2, dtype=np.float64)
    _check(lib().LGBM_B200_FindBin(v.ctypes.data_as(ctypes.c_void_p), ctypes.c_int(len(v)), ctypes.c_int64(total_sample_cnt),
                                   ctypes.c_int(max_bin), ctypes.c_int(min_data_in_bin), ctypes.c_int(min_split_data),
                                   ctypes.c_int(int(pre_filter)), ctypes.c_int(int(use_missing)),
                                   ctypes.c_int(int(zero_as_missing)), sc, ub.ctypes.data_as(ctypes.c_void_p),
                                   ctypes.c_int(len(ub))))
    return dict(num_bin=sc[0], missing_type=sc[1], default_bin=sc[2], most_freq_bin=sc[3], is_trivial=bool(sc[4]),
                upper_bounds=ub[:sc[0]].copy())


def network_init_nccl(unique_id, num_machines, rank, device_id):
    buf = (ctypes.c_uint8 * 128).from_buffer_copy(bytes(unique_id))
    _check(lib().LGBM_NetworkInitWithNccl(buf, ctypes.c_int(num_machines), ctypes.c_int(rank), ctypes.c_int(device_id)))


def network_unique_id():
    buf = (ctypes.c_uint8 * 128)()
    _check(lib().LGBM_NetworkGetNcclUniqueId(buf))
    return bytes(buf)


def network_free():
    _check(lib().LGBM_NetworkFree())
