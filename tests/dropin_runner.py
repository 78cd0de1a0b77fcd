"""Runs the drop-in library (integration/_ref/lib_lightgbm.so: the reference with the B200 tree learner compiled in
behind its own TreeLearner interface) in a process of its own -- the oracle's lib_lightgbm.so exports the same C++
symbols and must not share a process with it.  Usage (from tests/test_dropin_integration.py):
    python tests/dropin_runner.py IN.npz OUT.npz
IN.npz: X, y, [group], [weight], params (str), dataset_params (str), iters (int), [reset_after (int), reset_params (str)], [rollback_after (int)].  OUT.npz: model (str), scores (raw train scores)."""
import ctypes
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "integration", "_ref", "lib_lightgbm.so")


def main():
    src, dst = sys.argv[1], sys.argv[2]
    z = np.load(src, allow_pickle=False)
    X = np.ascontiguousarray(z["X"], dtype=np.float32)
    y = np.ascontiguousarray(z["y"], dtype=np.float32)
    lib = ctypes.CDLL(LIB)
    lib.LGBM_GetLastError.restype = ctypes.c_char_p

    def check(ret):
        if ret != 0:
            raise RuntimeError("drop-in library error: %s" % lib.LGBM_GetLastError().decode())

    ds = ctypes.c_void_p()
    n, f = X.shape
    check(lib.LGBM_DatasetCreateFromMat(X.ctypes.data_as(ctypes.c_void_p), ctypes.c_int(0), ctypes.c_int32(n), ctypes.c_int32(f),
                                        ctypes.c_int(1), ctypes.c_char_p(str(z["dataset_params"]).encode()), None, ctypes.byref(ds)))
    check(lib.LGBM_DatasetSetField(ds, b"label", y.ctypes.data_as(ctypes.c_void_p), ctypes.c_int(n), ctypes.c_int(0)))
    if "group" in z.files:
        g = np.ascontiguousarray(z["group"], dtype=np.int32)
        check(lib.LGBM_DatasetSetField(ds, b"group", g.ctypes.data_as(ctypes.c_void_p), ctypes.c_int(len(g)), ctypes.c_int(2)))
    if "weight" in z.files:
        w = np.ascontiguousarray(z["weight"], dtype=np.float32)
        check(lib.LGBM_DatasetSetField(ds, b"weight", w.ctypes.data_as(ctypes.c_void_p), ctypes.c_int(n), ctypes.c_int(0)))
    bst = ctypes.c_void_p()
    check(lib.LGBM_BoosterCreate(ds, ctypes.c_char_p(str(z["params"]).encode()), ctypes.byref(bst)))
    vds = None
    if "valid_X" in z.files:  # a validation set binned with the training set's mappers, LGBM_BoosterAddValidData
        VX = np.ascontiguousarray(z["valid_X"], dtype=np.float32)
        vy = np.ascontiguousarray(z["valid_y"], dtype=np.float32)
        vds = ctypes.c_void_p()
        check(lib.LGBM_DatasetCreateFromMat(VX.ctypes.data_as(ctypes.c_void_p), ctypes.c_int(0), ctypes.c_int32(VX.shape[0]),
                                            ctypes.c_int32(VX.shape[1]), ctypes.c_int(1),
                                            ctypes.c_char_p(str(z["dataset_params"]).encode()), ds, ctypes.byref(vds)))
        check(lib.LGBM_DatasetSetField(vds, b"label", vy.ctypes.data_as(ctypes.c_void_p), ctypes.c_int(len(vy)), ctypes.c_int(0)))
        if "valid_group" in z.files:
            vg = np.ascontiguousarray(z["valid_group"], dtype=np.int32)
            check(lib.LGBM_DatasetSetField(vds, b"group", vg.ctypes.data_as(ctypes.c_void_p), ctypes.c_int(len(vg)), ctypes.c_int(2)))
        check(lib.LGBM_BoosterAddValidData(bst, vds))
    fin = ctypes.c_int(0)
    reset_after = int(z["reset_after"]) if "reset_after" in z.files else -1
    for it in range(int(z["iters"])):
        if it == reset_after:  # LGBM_BoosterResetParameter (src/c_api.cpp:2053), e.g. a learning-rate schedule
            check(lib.LGBM_BoosterResetParameter(bst, ctypes.c_char_p(str(z["reset_params"]).encode())))
        check(lib.LGBM_BoosterUpdateOneIter(bst, ctypes.byref(fin)))
        if "rollback_after" in z.files and it == int(z["rollback_after"]):  # LGBM_BoosterRollbackOneIter (src/c_api.cpp:2113)
            check(lib.LGBM_BoosterRollbackOneIter(bst))
    need = ctypes.c_int64(0)
    buf = ctypes.create_string_buffer(1 << 24)
    check(lib.LGBM_BoosterSaveModelToString(bst, ctypes.c_int(0), ctypes.c_int(-1), ctypes.c_int(0), ctypes.c_int64(len(buf)),
                                            ctypes.byref(need), buf))
    if need.value > len(buf):
        buf = ctypes.create_string_buffer(need.value)
        check(lib.LGBM_BoosterSaveModelToString(bst, ctypes.c_int(0), ctypes.c_int(-1), ctypes.c_int(0), ctypes.c_int64(len(buf)),
                                                ctypes.byref(need), buf))
    cnt = ctypes.c_int64(0)
    check(lib.LGBM_BoosterGetNumPredict(bst, ctypes.c_int(0), ctypes.byref(cnt)))
    scores = np.zeros(cnt.value, dtype=np.float64)
    check(lib.LGBM_BoosterGetPredict(bst, ctypes.c_int(0), ctypes.byref(cnt), scores.ctypes.data_as(ctypes.c_void_p)))
    extra = {}
    if vds is not None:
        nm = ctypes.c_int(0)
        check(lib.LGBM_BoosterGetEvalCounts(bst, ctypes.byref(nm)))
        for idx, key in ((0, "train_eval"), (1, "valid_eval")):
            vals = np.zeros(max(nm.value, 1), dtype=np.float64)
            got = ctypes.c_int(0)
            check(lib.LGBM_BoosterGetEval(bst, ctypes.c_int(idx), ctypes.byref(got), vals.ctypes.data_as(ctypes.c_void_p)))
            extra[key] = vals[:got.value]
    np.savez(dst, model=np.array(buf.value.decode()), scores=scores, **extra)
    check(lib.LGBM_BoosterFree(bst))
    check(lib.LGBM_DatasetFree(ds))


if __name__ == "__main__":
    main()
