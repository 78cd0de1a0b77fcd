"""MulticlassSoftmax::GetGradients (src/objective/multiclass_objective.hpp:84-133, Common::Softmax
include/LightGBM/utils/common.h:580-594) on the device (GBL_OBJ_MULTICLASS), called through the C ABI, against a numpy
restatement of the reference's loop: class-major score / gradient / hessian blocks, fp64 softmax, results cast to float."""
import ctypes

import numpy as np
import pytest

from lambdagap_b200 import capi

pytestmark = pytest.mark.gpu


def _reference(score, label, weight, K):
    n = len(label)
    s = score.reshape(K, n)
    e = np.exp(s - s.max(axis=0, keepdims=True))
    p = e / e.sum(axis=0, keepdims=True)
    onehot = (np.arange(K)[:, None] == label.astype(np.int64)[None, :])
    factor = np.float64(K) / np.float64(np.float32(K) - np.float32(1.0))  # static_cast<double>(K) / (K - 1.0f), :31
    g = np.where(onehot, p - np.float32(1.0), p)
    h = factor * p * (np.float32(1.0) - p)
    if weight is not None:
        g = g * weight[None, :]
        h = h * weight[None, :]
    return g.astype(np.float32).ravel(), h.astype(np.float32).ravel()


@pytest.mark.parametrize("weighted", [False, True])
@pytest.mark.parametrize("K", [2, 3, 7])
def test_multiclass_gradients_equal_reference_loop(gpu, K, weighted):
    L = capi.lib()
    rng = np.random.default_rng(K)
    n = 50_001
    label = rng.integers(0, K, size=n).astype(np.float32)
    weight = rng.uniform(0.25, 3.0, size=n).astype(np.float32) if weighted else None
    score = (rng.normal(scale=3.0, size=K * n)).astype(np.float64)
    score[:K] = 700.0 * np.arange(K)  # far apart: exp() of large negative differences
    cfg = capi.ObjectiveConfig()
    cfg.kind = 3
    cfg.num_class = K
    cfg.sigmoid = 1.0
    cfg.scale_pos_weight = 1.0
    obj = ctypes.c_void_p()
    capi.check(L.gbl_objective_create(ctypes.byref(cfg), ctypes.c_int64(n), label.ctypes.data_as(ctypes.c_void_p),
                                      weight.ctypes.data_as(ctypes.c_void_p) if weighted else None, None, ctypes.c_int32(0),
                                      ctypes.c_int32(0), ctypes.byref(obj)))
    bufs = []
    for nbytes in (8 * K * n, 4 * K * n, 4 * K * n):
        p = ctypes.c_void_p()
        capi.check(L.gbl_device_malloc(ctypes.c_int32(0), ctypes.c_size_t(nbytes), ctypes.byref(p)))
        bufs.append(p)
    s_dev, g_dev, h_dev = bufs
    capi.check(L.gbl_memcpy_h2d(ctypes.c_int32(0), s_dev, score.ctypes.data_as(ctypes.c_void_p), ctypes.c_size_t(8 * K * n)))
    capi.check(L.gbl_objective_get_gradients(obj, s_dev, g_dev, h_dev, None))
    g = np.zeros(K * n, dtype=np.float32)
    h = np.zeros(K * n, dtype=np.float32)
    capi.check(L.gbl_memcpy_d2h(ctypes.c_int32(0), g.ctypes.data_as(ctypes.c_void_p), g_dev, ctypes.c_size_t(4 * K * n)))
    capi.check(L.gbl_memcpy_d2h(ctypes.c_int32(0), h.ctypes.data_as(ctypes.c_void_p), h_dev, ctypes.c_size_t(4 * K * n)))
    for p in bufs:
        capi.check(L.gbl_device_free(ctypes.c_int32(0), p))
    capi.check(L.gbl_objective_free(obj))
    eg, eh = _reference(score, label, weight, K)
    # exp() of the device's libm and the host's differ in the last ulp of a double; after the cast to float almost every
    # entry is identical, the rest one float ulp apart
    np.testing.assert_allclose(g, eg, rtol=3e-7, atol=1e-12)
    np.testing.assert_allclose(h, eh, rtol=3e-7, atol=1e-12)
    assert np.mean(g == eg) > 0.999 and np.mean(h == eh) > 0.999


def test_multiclass_rejects_labels_outside_the_classes(gpu):
    L = capi.lib()
    label = np.array([0, 1, 3], dtype=np.float32)
    cfg = capi.ObjectiveConfig()
    cfg.kind = 3
    cfg.num_class = 3
    cfg.sigmoid = 1.0
    obj = ctypes.c_void_p()
    ret = L.gbl_objective_create(ctypes.byref(cfg), ctypes.c_int64(3), label.ctypes.data_as(ctypes.c_void_p), None, None,
                                 ctypes.c_int32(0), ctypes.c_int32(0), ctypes.byref(obj))
    assert ret != 0
    L.gbl_last_error.restype = ctypes.c_char_p
    assert b"Label must be in [0, 3)" in L.gbl_last_error()
