"""gbl_partition_rows_noalloc: the split partition for callers that are not allowed to allocate device memory
(a host application whose own pool owns all of HBM, or a latency-critical loop that cannot take cudaMalloc's
device-wide synchronisation).  Its scratch is a static device array inside the library, shared by every call and
still holding the previous call's numbers -- so this test makes two calls in a row, the second on a different and
smaller problem, and checks each against what the regular entry points produce for the same split
(gbl_learner_train_host + gbl_learner_get_leaf_indices: the learner's own partition with cudaMalloc'ed scratch).
Row indices are integers: the comparison is exact.  Collected last on purpose (file name) so that a run that
stops at the first failure has already run everything else."""
import numpy as np
import pytest

from lambdagap_b200 import capi

pytestmark = pytest.mark.gpu


def _dataset(n, f, seed):
    rng = np.random.default_rng(seed)
    infos, cols = [], []
    for j in range(f):
        nb = 255 if j % 3 else 40
        mt = [capi.MISSING_NONE, capi.MISSING_ZERO, capi.MISSING_NAN][j % 3]
        infos.append(dict(num_bin=nb, missing_type=mt, default_bin=5 if mt == capi.MISSING_ZERO else 0,
                          most_freq_bin=5 if mt == capi.MISSING_ZERO else 0, real_index=j))
        cols.append(rng.integers(0, nb, size=n).astype(np.uint8))
    return infos, cols


def _reference_split(ds, n, seed):
    """One split through the existing entry points; returns (record, left rows, right rows)."""
    rng = np.random.default_rng(seed)
    lrn = capi.Learner(ds, capi.learner_config(num_leaves=2, min_data_in_leaf=1))
    g = rng.normal(size=n).astype(np.float32)
    h = rng.uniform(0.5, 1.5, size=n).astype(np.float32)
    res, recs = lrn.train_host(g, h)
    assert res["num_splits"] == 1
    left, right = lrn.leaf_indices(0), lrn.leaf_indices(1)
    lrn.free()
    return recs[0], left, right


def test_noalloc_partition_matches_learner_twice(gpu):
    n, f = 300_000, 9
    infos, cols = _dataset(n, f, 1)
    ds = capi.Dataset.from_bins(infos, cols)
    out = capi.DeviceBuffer(4 * n)
    cnt = capi.DeviceBuffer(4)

    # call 1: all rows (indices = NULL means 0..n-1), the split the learner chose for the root
    rec, left, right = _reference_split(ds, n, 11)
    capi.partition_rows_noalloc(ds, rec["feature"], rec["threshold_bin"], rec["default_left"], None, n, out, cnt)
    capi.synchronize()
    got = out.download(np.int32, n)
    assert int(cnt.download(np.int32, 1)[0]) == len(left) == rec["left_count"]
    np.testing.assert_array_equal(got[:len(left)], left)
    np.testing.assert_array_equal(got[len(left):], right)

    # call 2, right after: a different split applied to an explicit, smaller row list (the left child of call 1);
    # the scratch still holds call 1's per-CTA counts and total, which must not leak into this result
    rec2, _, _ = _reference_split(ds, n, 12)
    sub = np.ascontiguousarray(left, dtype=np.int32)
    idx = capi.DeviceBuffer.from_numpy(sub)
    capi.partition_rows_noalloc(ds, rec2["feature"], rec2["threshold_bin"], rec2["default_left"], idx, len(sub), out, cnt)
    capi.synchronize()
    got2 = out.download(np.int32, len(sub))
    bins = cols[rec2["feature"]][sub].astype(np.int64)
    info = infos[rec2["feature"]]
    missing = {capi.MISSING_NONE: -1, capi.MISSING_ZERO: info["default_bin"], capi.MISSING_NAN: info["num_bin"] - 1}[info["missing_type"]]
    go_left = np.where(bins == missing, bool(rec2["default_left"]), bins <= rec2["threshold_bin"])
    assert int(cnt.download(np.int32, 1)[0]) == int(go_left.sum())
    np.testing.assert_array_equal(got2, np.concatenate([sub[go_left], sub[~go_left]]))
    # and the same split over all rows equals the learner's own partition for it
    capi.partition_rows_noalloc(ds, rec2["feature"], rec2["threshold_bin"], rec2["default_left"], None, n, out, cnt)
    capi.synchronize()
    _, left2, right2 = _reference_split(ds, n, 12)
    got3 = out.download(np.int32, n)
    np.testing.assert_array_equal(got3, np.concatenate([left2, right2]))

    # sizes beyond the fixed scratch are refused with a clear error, nothing is launched
    with pytest.raises(capi.GblError, match="at most"):
        capi.partition_rows_noalloc(ds, 0, 3, True, None, (1 << 24) + 1, out, cnt)
    # an in-place call (input list == output buffer) is refused too: the scatter pass reads the input while writing the output
    with pytest.raises(capi.GblError, match="overlap"):
        capi.partition_rows_noalloc(ds, rec2["feature"], rec2["threshold_bin"], rec2["default_left"], idx, len(sub), idx, cnt)
    ds.free()
