"""ctypes binding of the C ABI declared in include/lambdagap_b200.h (lambdagap_b200/lib/liblambdagap_b200.so).

Thin by design: one Python method per C entry point, no compute on this side.  The library has no CPU
fallback; loading it without the built .so, or calling it without an sm_100 device, raises.
"""
import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# GBL_LIB_DIR: an alternative directory holding both libraries (A/B timing of two builds in one process launch each)
LIB_DIR = os.environ.get("GBL_LIB_DIR") or os.path.join(_HERE, "lib")
LIB_PATH = os.path.join(LIB_DIR, "liblambdagap_b200.so")

MISSING_NONE, MISSING_ZERO, MISSING_NAN = 0, 1, 2
OBJ_REGRESSION_L2, OBJ_BINARY, OBJ_LAMBDARANK = 0, 1, 2


class GblError(RuntimeError):
    pass


class FeatureMeta(ctypes.Structure):
    _fields_ = [("num_bin", ctypes.c_int32), ("missing_type", ctypes.c_int32), ("default_bin", ctypes.c_int32),
                ("most_freq_bin", ctypes.c_int32), ("real_index", ctypes.c_int32), ("monotone_type", ctypes.c_int32),
                ("penalty", ctypes.c_double), ("bin_type", ctypes.c_int32), ("reserved", ctypes.c_int32)]


class LearnerConfig(ctypes.Structure):
    _fields_ = [("num_leaves", ctypes.c_int32), ("max_depth", ctypes.c_int32), ("min_data_in_leaf", ctypes.c_int32),
                ("use_quantized_grad", ctypes.c_int32), ("num_grad_quant_bins", ctypes.c_int32),
                ("stochastic_rounding", ctypes.c_int32), ("quant_train_renew_leaf", ctypes.c_int32),
                ("is_constant_hessian", ctypes.c_int32), ("min_sum_hessian_in_leaf", ctypes.c_double),
                ("lambda_l1", ctypes.c_double), ("lambda_l2", ctypes.c_double), ("min_gain_to_split", ctypes.c_double),
                ("max_delta_step", ctypes.c_double), ("path_smooth", ctypes.c_double)]


class SplitRecord(ctypes.Structure):
    _fields_ = [("leaf", ctypes.c_int32), ("feature", ctypes.c_int32), ("real_feature", ctypes.c_int32),
                ("threshold_bin", ctypes.c_uint32), ("default_left", ctypes.c_int32), ("left_count", ctypes.c_int32),
                ("right_count", ctypes.c_int32), ("reserved", ctypes.c_int32), ("left_output", ctypes.c_double),
                ("right_output", ctypes.c_double), ("left_sum_gradient", ctypes.c_double),
                ("left_sum_hessian", ctypes.c_double), ("right_sum_gradient", ctypes.c_double),
                ("right_sum_hessian", ctypes.c_double), ("gain", ctypes.c_double),
                ("left_sum_gradient_and_hessian", ctypes.c_int64), ("right_sum_gradient_and_hessian", ctypes.c_int64),
                ("cat_bitset", ctypes.c_uint32 * 8), ("num_cat_bins", ctypes.c_int32), ("reserved2", ctypes.c_int32)]


class TreeResult(ctypes.Structure):
    _fields_ = [("num_splits", ctypes.c_int32), ("reserved", ctypes.c_int32), ("root_output", ctypes.c_double),
                ("root_sum_gradient", ctypes.c_double), ("root_sum_hessian", ctypes.c_double),
                ("grad_scale", ctypes.c_double), ("hess_scale", ctypes.c_double)]


class Profile(ctypes.Structure):
    _fields_ = [("hist_ms", ctypes.c_double), ("find_ms", ctypes.c_double), ("partition_ms", ctypes.c_double),
                ("other_ms", ctypes.c_double), ("collective_ms", ctypes.c_double), ("hist_launches", ctypes.c_int64),
                ("hist_rows", ctypes.c_int64), ("trees", ctypes.c_int64),
                ("partition_rows", ctypes.c_int64), ("waves", ctypes.c_int64), ("expansions", ctypes.c_int64),
                ("splits", ctypes.c_int64)]


class ObjectiveConfig(ctypes.Structure):
    _fields_ = [("kind", ctypes.c_int32), ("lambdarank_target", ctypes.c_int32),
                ("lambdarank_truncation_level", ctypes.c_int32), ("lambdarank_norm", ctypes.c_int32),
                ("sigmoid", ctypes.c_double), ("lambdagap_weight", ctypes.c_double), ("is_unbalance", ctypes.c_int32),
                ("num_class", ctypes.c_int32), ("scale_pos_weight", ctypes.c_double), ("num_label_gain", ctypes.c_int32),
                ("reserved2", ctypes.c_int32), ("label_gain", ctypes.POINTER(ctypes.c_double))]


class MetricConfig(ctypes.Structure):
    _fields_ = [("kind", ctypes.c_int32), ("num_eval_at", ctypes.c_int32), ("eval_at", ctypes.POINTER(ctypes.c_int32)),
                ("num_label_gain", ctypes.c_int32), ("reserved", ctypes.c_int32),
                ("label_gain", ctypes.POINTER(ctypes.c_double)), ("sigmoid", ctypes.c_double)]


METRIC_KINDS = {"l2": 0, "rmse": 1, "l1": 2, "binary_logloss": 3, "binary_error": 4, "auc": 5, "ndcg": 6}

# every symbol include/lambdagap_b200.h declares (tests check that the .so exports all of them)
EXPORTED_SYMBOLS = [
    "gbl_last_error", "gbl_device_count", "gbl_kernel_launch_count",
    "gbl_dataset_create_from_bins", "gbl_dataset_create_from_mat_f32", "gbl_dataset_free", "gbl_dataset_num_data",
    "gbl_dataset_get_feature_bins", "gbl_dataset_get_feature_bins_rowmajor",
    "gbl_comm_get_unique_id", "gbl_comm_create", "gbl_comm_free", "gbl_comm_allreduce_host_f64", "gbl_comm_info",
    "gbl_learner_create", "gbl_learner_free", "gbl_learner_reset_config", "gbl_learner_set_feature_mask", "gbl_learner_set_monotone", "gbl_learner_set_categorical",
    "gbl_learner_set_bagging_data", "gbl_learner_bagging_dev", "gbl_learner_bagging_balanced_dev", "gbl_learner_bagging_by_query_dev", "gbl_learner_set_quant_random", "gbl_learner_set_quant_random_start",
    "gbl_learner_train", "gbl_learner_train_host", "gbl_learner_train_async", "gbl_learner_train_finish",
    "gbl_learner_add_prediction_to_score", "gbl_learner_add_prediction_to_score_dev",
    "gbl_learner_add_prediction_out_of_bag_dev", "gbl_learner_add_prediction_to_dataset_dev", "gbl_learner_get_leaf_values", "gbl_learner_get_leaf_assignment", "gbl_learner_get_leaf_assignment_u16", "gbl_learner_get_leaf_indices", "gbl_learner_get_histogram",
    "gbl_learner_get_discretized", "gbl_learner_histogram_words", "gbl_learner_construct_histogram",
    "gbl_learner_stream", "gbl_learner_set_profiling", "gbl_learner_get_profile", "gbl_learner_timer_record",
    "gbl_learner_timer_elapsed_ms",
    "gbl_objective_create", "gbl_objective_free", "gbl_objective_get_gradients", "gbl_objective_boost_from_score",
    "gbl_objective_is_constant_hessian", "gbl_objective_set_label_counts",
    "gbl_metric_create", "gbl_metric_free", "gbl_metric_num_outputs", "gbl_metric_eval",
    "gbl_device_malloc", "gbl_device_free", "gbl_host_register", "gbl_host_unregister", "gbl_memcpy_h2d", "gbl_memcpy_d2h", "gbl_device_fill_f64",
    "gbl_device_synchronize", "gbl_partition_rows_noalloc",
]

_lib = None


def lib():
    """Loads the CUDA library; fails loudly when it was not built (no fallback exists)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise GblError("%s not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(make -C lambdagap_b200/csrc); there is no CPU fallback" % LIB_PATH)
        _lib = ctypes.CDLL(LIB_PATH)
        _lib.gbl_last_error.restype = ctypes.c_char_p
        _lib.gbl_kernel_launch_count.restype = ctypes.c_int64
        _lib.gbl_learner_stream.restype = ctypes.c_void_p
        _lib.gbl_learner_stream.argtypes = [ctypes.c_void_p]
    return _lib


def check(ret):
    if ret != 0:
        raise GblError(lib().gbl_last_error().decode())


def device_count():
    n = ctypes.c_int(0)
    check(lib().gbl_device_count(ctypes.byref(n)))
    return n.value


def kernel_launch_count():
    return int(lib().gbl_kernel_launch_count())


def _ptr(a):
    return a.ctypes.data_as(ctypes.c_void_p) if a is not None else None


def _dev(p):
    """Accepts a raw int device address or anything with .data_ptr() (a torch tensor)."""
    if p is None:
        return None
    if hasattr(p, "data_ptr"):
        return ctypes.c_void_p(p.data_ptr())
    return ctypes.c_void_p(int(p))


def make_metas(infos):
    """infos: iterable of dicts with num_bin, missing_type, default_bin, most_freq_bin, real_index[, penalty]."""
    arr = (FeatureMeta * len(infos))()
    for i, d in enumerate(infos):
        arr[i].num_bin = int(d["num_bin"])
        arr[i].missing_type = int(d["missing_type"])
        arr[i].default_bin = int(d["default_bin"])
        arr[i].most_freq_bin = int(d["most_freq_bin"])
        arr[i].real_index = int(d.get("real_index", i))
        arr[i].penalty = float(d.get("penalty", 1.0))
    return arr


class DeviceBuffer:
    """Raw device allocation through the C ABI (no torch needed)."""

    def __init__(self, nbytes, device=0):
        self.device = device
        self.nbytes = int(nbytes)
        p = ctypes.c_void_p()
        check(lib().gbl_device_malloc(ctypes.c_int32(device), ctypes.c_size_t(max(1, self.nbytes)), ctypes.byref(p)))
        self.ptr = p.value

    @classmethod
    def from_numpy(cls, a, device=0):
        a = np.ascontiguousarray(a)
        b = cls(a.nbytes, device)
        b.upload(a)
        return b

    def upload(self, a):
        a = np.ascontiguousarray(a)
        assert a.nbytes <= self.nbytes
        check(lib().gbl_memcpy_h2d(ctypes.c_int32(self.device), ctypes.c_void_p(self.ptr), _ptr(a),
                                   ctypes.c_size_t(a.nbytes)))

    def download(self, dtype, count):
        out = np.empty(count, dtype=dtype)
        check(lib().gbl_memcpy_d2h(ctypes.c_int32(self.device), _ptr(out), ctypes.c_void_p(self.ptr),
                                   ctypes.c_size_t(out.nbytes)))
        return out

    def data_ptr(self):
        return self.ptr

    def free(self):
        if self.ptr:
            lib().gbl_device_free(ctypes.c_int32(self.device), ctypes.c_void_p(self.ptr))
            self.ptr = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class Dataset:
    def __init__(self, handle, num_data, num_features, infos, device):
        self.handle = handle
        self.num_data = num_data
        self.num_features = num_features
        self.infos = infos
        self.device = device

    @classmethod
    def from_bins(cls, infos, columns, device=0):
        """columns: list of uint8 arrays (mapper bins), one per used feature."""
        cols = [np.ascontiguousarray(c, dtype=np.uint8) for c in columns]
        n = len(cols[0])
        metas = make_metas(infos)
        ptrs = (ctypes.c_void_p * len(cols))(*[c.ctypes.data for c in cols])
        h = ctypes.c_void_p()
        check(lib().gbl_dataset_create_from_bins(metas, ctypes.c_int32(len(cols)), ctypes.c_int64(n), ptrs,
                                                 ctypes.c_int32(device), ctypes.byref(h)))
        return cls(h, n, len(cols), list(infos), device)

    @classmethod
    def from_mat(cls, infos, X, upper_bounds, device=0):
        """X: row-major float32 [n, total_columns]; upper_bounds: concatenated float64 bin upper bounds."""
        X = np.ascontiguousarray(X, dtype=np.float32)
        ub = np.ascontiguousarray(upper_bounds, dtype=np.float64)
        metas = make_metas(infos)
        h = ctypes.c_void_p()
        check(lib().gbl_dataset_create_from_mat_f32(metas, ctypes.c_int32(len(infos)), ctypes.c_int64(X.shape[0]),
                                                    _ptr(X), ctypes.c_int32(X.shape[1]), _ptr(ub),
                                                    ctypes.c_int32(device), ctypes.byref(h)))
        return cls(h, X.shape[0], len(infos), list(infos), device)

    def feature_bins(self, f, rowmajor=False):
        out = np.empty(self.num_data, dtype=np.uint8)
        fn = lib().gbl_dataset_get_feature_bins_rowmajor if rowmajor else lib().gbl_dataset_get_feature_bins
        check(fn(self.handle, ctypes.c_int32(f), _ptr(out)))
        return out

    def free(self):
        if self.handle:
            lib().gbl_dataset_free(self.handle)
            self.handle = None


class Comm:
    def __init__(self, unique_id, num_ranks, rank, device):
        self.handle = ctypes.c_void_p()
        self.num_ranks, self.rank = num_ranks, rank
        buf = (ctypes.c_uint8 * 128).from_buffer_copy(bytes(unique_id))
        check(lib().gbl_comm_create(buf, ctypes.c_int32(num_ranks), ctypes.c_int32(rank), ctypes.c_int32(device),
                                    ctypes.byref(self.handle)))

    @staticmethod
    def unique_id():
        buf = (ctypes.c_uint8 * 128)()
        check(lib().gbl_comm_get_unique_id(buf))
        return bytes(buf)

    def free(self):
        if self.handle:
            lib().gbl_comm_free(self.handle)
            self.handle = None


def learner_config(**kw):
    c = LearnerConfig()
    c.num_leaves = 31
    c.max_depth = -1
    c.min_data_in_leaf = 20
    c.use_quantized_grad = 0
    c.num_grad_quant_bins = 4
    c.stochastic_rounding = 1
    c.quant_train_renew_leaf = 0
    c.is_constant_hessian = 0
    c.min_sum_hessian_in_leaf = 1e-3
    c.lambda_l1 = 0.0
    c.lambda_l2 = 0.0
    c.min_gain_to_split = 0.0
    c.max_delta_step = 0.0
    c.path_smooth = 0.0
    for k, v in kw.items():
        if not hasattr(c, k):
            raise KeyError(k)
        setattr(c, k, v)
    return c


class Learner:
    def __init__(self, dataset, config, comm=None):
        self.ds = dataset
        self.cfg = config
        self.handle = ctypes.c_void_p()
        check(lib().gbl_learner_create(dataset.handle, ctypes.byref(config), comm.handle if comm else None,
                                       ctypes.byref(self.handle)))
        self._records = (SplitRecord * max(1, config.num_leaves))()
        self._result = TreeResult()

    def reset_config(self, config):
        check(lib().gbl_learner_reset_config(self.handle, ctypes.byref(config)))
        self.cfg = config
        self._records = (SplitRecord * max(1, config.num_leaves))()

    def set_feature_mask(self, mask):
        m = None if mask is None else np.ascontiguousarray(mask, dtype=np.int8)
        check(lib().gbl_learner_set_feature_mask(self.handle, _ptr(m)))

    def set_bagging_data(self, indices):
        if indices is None:
            check(lib().gbl_learner_set_bagging_data(self.handle, None, ctypes.c_int32(0)))
        else:
            a = np.ascontiguousarray(indices, dtype=np.int32)
            check(lib().gbl_learner_set_bagging_data(self.handle, _ptr(a), ctypes.c_int32(len(a))))

    def bagging_dev(self, seed, fraction, reset=False):
        """Draws the reference's bag on the device (gbl_learner_bagging_dev); returns the bag size."""
        out = ctypes.c_int32(0)
        check(lib().gbl_learner_bagging_dev(self.handle, ctypes.c_int32(seed), ctypes.c_double(fraction),
                                            ctypes.c_int32(1 if reset else 0), ctypes.byref(out)))
        return out.value

    def set_quant_random(self, rg, rh):
        rg = np.ascontiguousarray(rg, dtype=np.float64)
        rh = np.ascontiguousarray(rh, dtype=np.float64)
        check(lib().gbl_learner_set_quant_random(self.handle, _ptr(rg), _ptr(rh), ctypes.c_int64(len(rg))))

    def set_quant_random_start(self, start):
        check(lib().gbl_learner_set_quant_random_start(self.handle, ctypes.c_int32(start)))

    def _collect(self):
        n = self._result.num_splits
        recs = []
        for k in range(n):
            r = self._records[k]
            rec = {name: getattr(r, name) for name, _ in SplitRecord._fields_}
            rec["cat_bitset"] = tuple(int(w) for w in r.cat_bitset)  # a plain value: records compare with ==
            recs.append(rec)
        res = {name: getattr(self._result, name) for name, _ in TreeResult._fields_}
        return res, recs

    def train(self, grad_dev, hess_dev):
        check(lib().gbl_learner_train(self.handle, _dev(grad_dev), _dev(hess_dev), ctypes.byref(self._result),
                                      self._records))
        return self._collect()

    def train_host(self, grad, hess):
        g = np.ascontiguousarray(grad, dtype=np.float32)
        h = np.ascontiguousarray(hess, dtype=np.float32)
        check(lib().gbl_learner_train_host(self.handle, _ptr(g), _ptr(h), ctypes.byref(self._result), self._records))
        return self._collect()

    def train_async(self, grad_dev, hess_dev):
        check(lib().gbl_learner_train_async(self.handle, _dev(grad_dev), _dev(hess_dev)))

    def train_finish(self):
        check(lib().gbl_learner_train_finish(self.handle, ctypes.byref(self._result), self._records))
        return self._collect()

    def add_prediction_to_score(self, leaf_values, score_dev):
        lv = np.ascontiguousarray(leaf_values, dtype=np.float64)
        check(lib().gbl_learner_add_prediction_to_score(self.handle, _ptr(lv), ctypes.c_int32(len(lv)), _dev(score_dev)))

    def add_prediction_to_score_dev(self, shrinkage, bias, score_dev):
        check(lib().gbl_learner_add_prediction_to_score_dev(self.handle, ctypes.c_double(shrinkage),
                                                            ctypes.c_double(bias), _dev(score_dev)))

    def leaf_assignment(self):
        out = np.empty(self.ds.num_data, dtype=np.int32)
        check(lib().gbl_learner_get_leaf_assignment(self.handle, _ptr(out)))
        return out

    def leaf_values(self, num_leaves):
        out = np.zeros(num_leaves, dtype=np.float64)
        check(lib().gbl_learner_get_leaf_values(self.handle, _ptr(out), ctypes.c_int32(num_leaves)))
        return out

    def leaf_indices(self, leaf):
        cnt = ctypes.c_int32(0)
        out = np.empty(self.ds.num_data, dtype=np.int32)
        check(lib().gbl_learner_get_leaf_indices(self.handle, ctypes.c_int32(leaf), _ptr(out), ctypes.byref(cnt)))
        return out[:cnt.value].copy()

    def histogram(self, leaf, feature):
        nb = self.ds.infos[feature]["num_bin"]
        out = np.zeros((nb, 2), dtype=np.float64)
        check(lib().gbl_learner_get_histogram(self.handle, ctypes.c_int32(leaf), ctypes.c_int32(feature), _ptr(out)))
        return out

    def discretized(self):
        out = np.empty(2 * self.ds.num_data, dtype=np.int8)
        check(lib().gbl_learner_get_discretized(self.handle, _ptr(out)))
        return out

    def histogram_words(self):
        w = ctypes.c_int64(0)
        check(lib().gbl_learner_histogram_words(self.handle, ctypes.byref(w)))
        return w.value

    def construct_histogram(self, indices_dev, num_indices, out_dev):
        check(lib().gbl_learner_construct_histogram(self.handle, _dev(indices_dev), ctypes.c_int64(num_indices),
                                                    _dev(out_dev)))

    def decode_histogram(self, raw_words):
        """Internal [tile][256][32] layout -> (num_features, 256, 2) int64 (grad, hess) integer sums."""
        F = self.ds.num_features
        tiles = (F + 31) // 32
        quant = bool(self.cfg.use_quantized_grad)
        raw = np.asarray(raw_words, dtype=np.int64)
        out = np.zeros((F, 256, 2), dtype=np.int64)
        if quant:
            a = raw[:tiles * 256 * 32].reshape(tiles, 256, 32)
            g = a >> 32
            h = a & 0xffffffff
        else:
            a = raw[:tiles * 256 * 32 * 2].reshape(tiles, 256, 32, 2)
            g = a[..., 0]
            h = a[..., 1]
        for f in range(F):
            out[f, :, 0] = g[f // 32, :, f % 32]
            out[f, :, 1] = h[f // 32, :, f % 32]
        return out

    def stream(self):
        return lib().gbl_learner_stream(self.handle)

    def timer_record(self, slot):
        check(lib().gbl_learner_timer_record(self.handle, ctypes.c_int32(slot)))

    def timer_elapsed_ms(self, a, b):
        out = ctypes.c_double(0)
        check(lib().gbl_learner_timer_elapsed_ms(self.handle, ctypes.c_int32(a), ctypes.c_int32(b), ctypes.byref(out)))
        return out.value

    def set_profiling(self, on):
        check(lib().gbl_learner_set_profiling(self.handle, ctypes.c_int32(1 if on else 0)))

    def profile(self, reset=False):
        p = Profile()
        check(lib().gbl_learner_get_profile(self.handle, ctypes.byref(p), ctypes.c_int32(1 if reset else 0)))
        return {name: getattr(p, name) for name, _ in Profile._fields_}

    def free(self):
        if self.handle:
            lib().gbl_learner_free(self.handle)
            self.handle = None


class Metric:
    """gbl_metric_*: one evaluation metric over a device-resident score vector (src/metric/*.hpp)."""

    def __init__(self, kind, label, weight=None, query_boundaries=None, eval_at=(1, 2, 3, 4, 5), label_gain=None,
                 sigmoid=0.0, device=0):
        label = np.ascontiguousarray(label, dtype=np.float32)
        cfg = MetricConfig()
        cfg.kind = METRIC_KINDS[kind]
        self._eval_at = np.ascontiguousarray(eval_at, dtype=np.int32)
        if label_gain is None:
            label_gain = [float((1 << i) - 1) for i in range(31)]
        self._gain = np.ascontiguousarray(label_gain, dtype=np.float64)
        cfg.num_eval_at = len(self._eval_at)
        cfg.eval_at = self._eval_at.ctypes.data_as(ctypes.POINTER(ctypes.c_int32))
        cfg.num_label_gain = len(self._gain)
        cfg.label_gain = self._gain.ctypes.data_as(ctypes.POINTER(ctypes.c_double))
        cfg.sigmoid = sigmoid
        w = None if weight is None else np.ascontiguousarray(weight, dtype=np.float32)
        qb = None if query_boundaries is None else np.ascontiguousarray(query_boundaries, dtype=np.int32)
        self.handle = ctypes.c_void_p()
        check(lib().gbl_metric_create(ctypes.byref(cfg), ctypes.c_int64(len(label)), _ptr(label),
                                      _ptr(w) if w is not None else None, _ptr(qb) if qb is not None else None,
                                      ctypes.c_int32(0 if qb is None else len(qb) - 1), ctypes.c_int32(device),
                                      ctypes.byref(self.handle)))
        n = ctypes.c_int32(0)
        check(lib().gbl_metric_num_outputs(self.handle, ctypes.byref(n)))
        self.num_outputs = n.value

    def eval(self, score_dev, stream=None):
        out = np.zeros(self.num_outputs, dtype=np.float64)
        check(lib().gbl_metric_eval(self.handle, _dev(score_dev), ctypes.c_void_p(stream), _ptr(out)))
        return out

    def free(self):
        if self.handle:
            check(lib().gbl_metric_free(self.handle))
            self.handle = None


def partition_rows_noalloc(dataset, feature, threshold_bin, default_left, indices_dev, num_indices, out_indices_dev,
                           out_left_count_dev, stream=None):
    """gbl_partition_rows_noalloc: stable split of a row list with library-resident scratch (no allocation)."""
    check(lib().gbl_partition_rows_noalloc(dataset.handle, ctypes.c_int32(feature), ctypes.c_uint32(threshold_bin),
                                           ctypes.c_int32(1 if default_left else 0), _dev(indices_dev),
                                           ctypes.c_int32(num_indices), _dev(out_indices_dev), _dev(out_left_count_dev),
                                           ctypes.c_void_p(stream) if stream else None))


def synchronize(device=0):
    check(lib().gbl_device_synchronize(ctypes.c_int32(device)))
