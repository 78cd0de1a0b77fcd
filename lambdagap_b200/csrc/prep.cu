// prep.cu -- per-tree preparation kernels: DataPartition::Init, gradient range scan, fixed-point conversion
// (float mode) and GradientDiscretizer::DiscretizeGradients (quantized mode).
//
// Reference: src/treelearner/data_partition.hpp:49-64 (iota / bagged indices),
//            src/treelearner/gradient_discretizer.cpp:68-161 (max |g|, max |h|, scales, int8 rounding).
#include "kernels.cuh"

namespace gbl {

__global__ void __launch_bounds__(256) iota_kernel(int* __restrict__ out, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = i;
}

__global__ void __launch_bounds__(256) copy_int_kernel(int* __restrict__ out, const int* __restrict__ in, int n,
                                                       const int* __restrict__ n_dev) {
  if (n_dev != nullptr) n = *n_dev;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = in[i];
}

// One CTA: resets the tree state for a new tree (SerialTreeLearner::BeforeTrain,
// src/treelearner/serial_tree_learner.cpp:288-338) and sets the root up as the first wave's histogram job.
__global__ void __launch_bounds__(256) tree_begin_kernel(LearnerDev d, int root_count, int root_buf,
                                                         const int* __restrict__ root_count_dev) {
  const int tid = threadIdx.x;
  if (root_count_dev != nullptr) root_count = *root_count_dev;
  for (int l = tid; l < d.p.num_leaves; l += blockDim.x) {
    d.leaf_values[l] = 0.0;
    d.leaf_node[l] = l == 0 ? 0 : -1;
  }
  for (int n = tid; n < d.max_nodes; n += blockDim.x) {
    d.bkey[n] = 0ull;
    d.expanded[n] = 0;
  }
  for (int j = tid; j < 2 * d.max_cand; j += blockDim.x) d.job_done[j] = 0u;
  // FeatureHistogram::is_splittable_ of the root slot starts true (feature_histogram.hpp:1356)
  for (int f = tid; f < d.p.num_features; f += blockDim.x) d.splittable[f] = 1;
  if (tid == 0) {
    TreeState& t = *d.ts;
    t.num_nodes = 1;
    t.num_expanded = 0;
    t.num_slots = 1;
    t.num_cand = 1;  // the root "wave": one histogram job, one scan job
    t.part_units = 0;
    t.hist_units = (root_count + kHistUnit - 1) / kHistUnit;
    t.scan_jobs = 1;
    t.done = 0;
    t.last_wave = 0;
    t.waves = 0;
    t.num_splits = 0;
    t.num_segments = 1;
    t.seg_units = 0;
    t.error = 0;
    t.rp_nl = 1;
    t.rp_ns = 0;
    t.scan_done_counter = 0;
    t.max_abs_g = 0.f;
    t.max_abs_h = 0.f;
    t.root_output = 0.0;
    t.hist_rows_total = root_count;
    t.hist_launches = 1;
    t.part_rows_total = 0;
    Node& R = d.nodes[0];
    R.begin = 0;
    R.count = root_count;
    R.global_count = root_count;
    R.depth = 0;
    R.buf = root_buf;
    R.slot = 0;
    R.parent = -1;
    R.left = -1;
    R.ref_leaf = 0;
    R.in_tree = 0;
    R.sum_g = 0.0;
    R.sum_h = 0.0;
    R.ig = 0;
    R.ih = 0;
    R.output = 0.0;
    R.cmin = -1.7976931348623157e308;  // BasicConstraint(): [-max double, +max double] (monotone_constraints.hpp:52)
    R.cmax = 1.7976931348623157e308;
    R.bval = -INFINITY;
    R.best.gain = -INFINITY;  // SplitInfo::Reset (split_info.hpp:132-136)
    R.best.feature = -1;
    R.best.real_feature = -1;
    Expand& w = d.wave[0];
    w.node = 0;
    w.hist_node = 0;
    w.hist_slot = 0;
    w.skip = 0;
    w.part_unit_begin = 0;
    w.hist_unit_begin = 0;
    w.left_total = 0;
    d.wave[1].hist_unit_begin = t.hist_units;
    d.wave[1].part_unit_begin = 0;
    // data-parallel: the local root count rides in the tail of slot 0 through the histogram all-reduce
    d.hist_pool[d.slot_stride - kTailWords + 0] = root_count;
    d.hist_pool[d.slot_stride - kTailWords + 1] = 0;
  }
}

// max |g|, max |h| over all rows (gradient_discretizer.cpp:72-100).  Non-negative floats order like their bit
// patterns, so atomicMax on the int view is exact.
__global__ void __launch_bounds__(256) maxabs_kernel(const float* __restrict__ g, const float* __restrict__ h, int n,
                                                     TreeState* ts) {
  float mg = 0.f, mh = 0.f;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    mg = fmaxf(mg, fabsf(g[i]));
    mh = fmaxf(mh, fabsf(h[i]));
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    mg = fmaxf(mg, __shfl_xor_sync(0xffffffffu, mg, o));
    mh = fmaxf(mh, __shfl_xor_sync(0xffffffffu, mh, o));
  }
  __shared__ float sg[8], sh[8];
  if ((threadIdx.x & 31) == 0) {
    sg[threadIdx.x >> 5] = mg;
    sh[threadIdx.x >> 5] = mh;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < 8; ++w) {
      mg = fmaxf(mg, sg[w]);
      mh = fmaxf(mh, sh[w]);
    }
    atomicMax(reinterpret_cast<int*>(&ts->max_abs_g), __float_as_int(mg));
    atomicMax(reinterpret_cast<int*>(&ts->max_abs_h), __float_as_int(mh));
  }
}

__device__ inline double pow2_scale_for(float max_abs) {
  if (!(max_abs > 0.f) || isinf(max_abs)) return 1.0;
  const int e = ilogbf(max_abs) + 1;  // max_abs = m * 2^e, m in [0.5, 1)
  return ldexp(1.0, 30 - e);          // |v| * scale <= 2^30
}

__global__ void scales_kernel(LearnerDev d) {
  TreeState& t = *d.ts;
  if (d.p.quantized) {
    // gradient_discretizer.cpp:105-114
    const double max_g = static_cast<double>(t.max_abs_g);
    const double max_h = static_cast<double>(t.max_abs_h);
    t.q_grad_scale = max_g / static_cast<double>(d.p.num_grad_quant_bins / 2);
    t.q_hess_scale = d.p.is_constant_hessian ? max_h : max_h / static_cast<double>(d.p.num_grad_quant_bins);
    t.q_inv_grad_scale = 1.0 / t.q_grad_scale;
    t.q_inv_hess_scale = 1.0 / t.q_hess_scale;
  } else {
    t.g_scale = pow2_scale_for(t.max_abs_g);
    t.h_scale = pow2_scale_for(t.max_abs_h);
    t.g_inv = 1.0 / t.g_scale;
    t.h_inv = 1.0 / t.h_scale;
  }
}

// copies the dataset's row-major bin mirror into the learner's row records: tile t of row r (32 bytes) goes to
// rec + r * rec_stride + t * tile_stride (16 bytes per thread)
__global__ void __launch_bounds__(256) build_records_kernel(uint8_t* __restrict__ rec, int rec_stride, int tile_stride,
                                                            const uint8_t* __restrict__ row_bins, int row_stride, int n) {
  const int per_row = row_stride / 16;
  const long long total = static_cast<long long>(n) * per_row;
  for (long long e = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; e < total;
       e += static_cast<long long>(gridDim.x) * blockDim.x) {
    const long long r = e / per_row;
    const int c = static_cast<int>(e - r * per_row);  // 16-byte chunk of the row: tile c / 2, half c & 1
    const uint4 v = *reinterpret_cast<const uint4*>(row_bins + r * row_stride + c * 16);
    *reinterpret_cast<uint4*>(rec + r * rec_stride + (c >> 1) * tile_stride + (c & 1) * 16) = v;
  }
}

void launch_build_records(uint8_t* rec, int rec_stride, int tile_stride, const uint8_t* row_bins, int row_stride, int n,
                          cudaStream_t s) {
  GBL_LAUNCH(build_records_kernel, kNumSM * 16, 256, 0, s, rec, rec_stride, tile_stride, row_bins, row_stride, n);
}

void launch_iota(int* out, int n, cudaStream_t s) {
  GBL_LAUNCH(iota_kernel, (n + 255) / 256, 256, 0, s, out, n);
}

void launch_tree_begin(const LearnerDev& d, const int* bag_indices, int bag_count, const int* bag_count_dev,
                       cudaStream_t s) {
  const int root_count = bag_indices != nullptr ? bag_count : d.num_data;
  if (bag_indices == nullptr) bag_count_dev = nullptr;
  // without bagging the root's segment is the learner's constant 0..N-1 buffer: nothing to initialise per tree
  GBL_LAUNCH(tree_begin_kernel, 1, 256, 0, s, d, root_count, bag_indices != nullptr ? 0 : 2, bag_count_dev);
  if (bag_indices != nullptr) {
    const int cap = bag_count_dev != nullptr ? d.num_data : bag_count;
    GBL_LAUNCH(copy_int_kernel, (cap + 255) / 256, 256, 0, s, d.idx[0], bag_indices, bag_count, bag_count_dev);
  }
}

void launch_maxabs(const LearnerDev& d, cudaStream_t s) {
  const int grid = std::min((d.num_data + 255) / 256, kNumSM * 8);
  GBL_LAUNCH(maxabs_kernel, grid, 256, 0, s, d.grad, d.hess, d.num_data, d.ts);
}

// scales only: the conversion itself (float mode: to fixed point; quantized: DiscretizeGradients) happens inside the
// root histogram pass, which reads every row anyway (histogram.cu)
void launch_scales_and_convert(const LearnerDev& d, cudaStream_t s) {
  GBL_LAUNCH(scales_kernel, 1, 1, 0, s, d);
}

}  // namespace gbl
