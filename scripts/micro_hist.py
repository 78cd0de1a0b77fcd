"""Micro-benchmark of the histogram kernel alone through gbl_learner_construct_histogram: back-to-back launches on
row subsets of different sizes (sorted random subsets = sparse leaves), CUDA-event timed on the learner stream."""
import os
import sys
import numpy as np
sys.path.insert(0, ".")
from lambdagap_b200 import capi

n, f = 10_500_000, 28
quant = int(sys.argv[1]) if len(sys.argv) > 1 else 0
rng = np.random.default_rng(0)
infos = [dict(num_bin=255, missing_type=0, default_bin=0, most_freq_bin=0, real_index=i) for i in range(f)]
cols = [rng.integers(0, 255, size=n).astype(np.uint8) for _ in range(f)]
ds = capi.Dataset.from_bins(infos, cols)
lrn = capi.Learner(ds, capi.learner_config(num_leaves=2, use_quantized_grad=quant, stochastic_rounding=0))
g = rng.normal(size=n).astype(np.float32)
h = rng.uniform(0.1, 1, size=n).astype(np.float32)
lrn.train_host(g, h)
words = lrn.histogram_words()
out = capi.DeviceBuffer(8 * words)
out.upload(np.zeros(words, dtype=np.int64))
gh = 2 if quant else 8
for m in (n, 5_000_000, 2_000_000, 500_000, 100_000, 20_000, 4_000, 512):
    if m == n:
        idx = None
    else:
        idx = capi.DeviceBuffer.from_numpy(np.sort(rng.choice(n, size=m, replace=False)).astype(np.int32))
    reps = 20 if m > 1_000_000 else 200
    reps = int(os.environ.get("MICRO_REPS", reps))  # 1 under ncu
    for _ in range(3):
        lrn.construct_histogram(idx, m, out)
    lrn.timer_record(0)
    for _ in range(reps):
        lrn.construct_histogram(idx, m, out)
    lrn.timer_record(1)
    us = 1e3 * lrn.timer_elapsed_ms(0, 1) / reps
    alg = m * (f + gh + (4 if idx is not None else 0))
    print("rows %9d: %8.2f us/launch  algorithmic %7.1f MB -> %7.1f GB/s (%.1f%% of 6548)" % (m, us, alg / 1e6, alg / us / 1e3, 100 * alg / us / 1e3 / 6548.2), flush=True)
