// partition.cu -- DataPartition::Split as a stable segmented two-way partition of row indices, the leaf
// bookkeeping that follows a split, and the per-leaf score update.
//
// Reference: DataPartition::Split                 src/treelearner/data_partition.hpp:101-121
//            ParallelPartitionRunner::Run<false>   include/LightGBM/utils/threading.h:118-185 (stable: left block
//                                                  keeps order, right block keeps order, left first)
//            DenseBin::SplitInner                  src/io/dense_bin.hpp:316-394 (which rows go left)
//            SerialTreeLearner::SplitInner         src/treelearner/serial_tree_learner.cpp:766-922
//            SerialTreeLearner::AddPredictionToScore  src/treelearner/serial_tree_learner.h:100-115
//
// In mapper-bin terms DenseBin::SplitInner reduces to:  missing rows (zero bin for MissingType::Zero, last bin
// for MissingType::NaN) follow default_left; every other row goes left iff bin <= threshold (the most-frequent
// bin is an ordinary bin here).  Index segments ping-pong between two buffers, so no copy-back is needed.
#include <cstdlib>

#include "kernels.cuh"

namespace gbl {

constexpr int kPartThreads = 256;
constexpr int kPartGrid = kNumSM * 4;
constexpr int kPartMinChunk = 2048;

int partition_grid() { return kPartGrid; }

struct SplitRule {
  const uint8_t* col;
  int threshold;
  int missing_bin;  // -1 = none
  bool default_left;
};

__device__ inline SplitRule load_rule(const LearnerDev& d) {
  const TreeState& t = *d.ts;
  const FeatMeta m = d.meta[t.split_feature];
  SplitRule r;
  r.col = d.col_bins + static_cast<int64_t>(t.split_feature) * d.col_stride;
  r.threshold = static_cast<int>(t.split_threshold);
  r.missing_bin = -1;
  if (m.missing_type == GBL_MISSING_ZERO) r.missing_bin = m.default_bin;
  if (m.missing_type == GBL_MISSING_NAN) r.missing_bin = m.num_bin - 1;
  r.default_left = t.split_default_left != 0;
  return r;
}

__device__ inline bool goes_left(const SplitRule& r, int row) {
  const int bin = r.col[row];
  return bin == r.missing_bin ? r.default_left : bin <= r.threshold;
}

__device__ inline int part_chunk(int n) {
  int c = (n + kPartGrid - 1) / kPartGrid;
  c = max(c, kPartMinChunk);
  return (c + kPartThreads - 1) / kPartThreads * kPartThreads;
}

// pass 1: left count of every CTA's chunk
__global__ void __launch_bounds__(kPartThreads) part_count_kernel(LearnerDev d) {
  const TreeState& t = *d.ts;
  if (t.done || !t.do_split) return;
  const LeafState& L = d.leaves[t.split_leaf];
  const int n = L.count;
  const int chunk = part_chunk(n);
  const int c_begin = blockIdx.x * chunk;
  if (c_begin >= n) {
    if (threadIdx.x == 0) d.block_counts[blockIdx.x] = 0;
    return;
  }
  const int c_end = min(n, c_begin + chunk);
  const int* __restrict__ src = d.idx[L.buf] + L.begin;
  const SplitRule rule = load_rule(d);
  int cnt = 0;
  for (int i = c_begin + threadIdx.x; i < c_end; i += kPartThreads) cnt += goes_left(rule, src[i]) ? 1 : 0;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
  __shared__ int s[kPartThreads / 32];
  if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = cnt;
  __syncthreads();
  if (threadIdx.x == 0) {
    int tot = 0;
    for (int w = 0; w < kPartThreads / 32; ++w) tot += s[w];
    d.block_counts[blockIdx.x] = tot;
  }
}

// pass 2: stable scatter + (last CTA) leaf bookkeeping for the next round
__global__ void __launch_bounds__(kPartThreads) part_scatter_kernel(LearnerDev d, int round) {
  TreeState& t = *d.ts;
  if (t.done || !t.do_split) return;
  __shared__ int s_red[kPartThreads / 32];
  __shared__ int s_prefix, s_total;
  __shared__ int s_wl[kPartThreads / 32];
  __shared__ int s_is_last;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const LeafState L = d.leaves[t.split_leaf];
  const int n = L.count;
  const int chunk = part_chunk(n);
  // prefix over CTA counts (kPartGrid <= 1024 values)
  int before = 0, total = 0;
  for (int b = tid; b < kPartGrid; b += kPartThreads) {
    const int cval = d.block_counts[b];
    total += cval;
    if (b < static_cast<int>(blockIdx.x)) before += cval;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    before += __shfl_xor_sync(0xffffffffu, before, o);
    total += __shfl_xor_sync(0xffffffffu, total, o);
  }
  if (lane == 0) s_red[warp] = before;
  __syncthreads();
  if (tid == 0) {
    int v = 0;
    for (int w = 0; w < kPartThreads / 32; ++w) v += s_red[w];
    s_prefix = v;
  }
  __syncthreads();
  if (lane == 0) s_red[warp] = total;
  __syncthreads();
  if (tid == 0) {
    int v = 0;
    for (int w = 0; w < kPartThreads / 32; ++w) v += s_red[w];
    s_total = v;
  }
  __syncthreads();
  const int left_total = s_total;
  const int c_begin = blockIdx.x * chunk;
  if (c_begin < n) {
    const int c_end = min(n, c_begin + chunk);
    const int* __restrict__ src = d.idx[L.buf] + L.begin;
    int* __restrict__ dst = d.idx[L.buf ^ 1] + L.begin;
    const SplitRule rule = load_rule(d);
    int left_run = s_prefix;                  // lefts written before this tile
    int right_run = c_begin - s_prefix;       // rights written before this tile
    for (int base = c_begin; base < c_end; base += kPartThreads) {
      const int i = base + tid;
      int row = 0;
      bool valid = i < c_end, left = false;
      if (valid) {
        row = src[i];
        left = goes_left(rule, row);
      }
      const unsigned bl = __ballot_sync(0xffffffffu, valid && left);
      const unsigned bv = __ballot_sync(0xffffffffu, valid);
      if (lane == 0) s_wl[warp] = __popc(bl);
      __syncthreads();
      int wl_before = 0, tile_left = 0;
#pragma unroll
      for (int w = 0; w < kPartThreads / 32; ++w) {
        const int cw = s_wl[w];
        if (w < warp) wl_before += cw;
        tile_left += cw;
      }
      const unsigned lt = (1u << lane) - 1u;
      if (valid) {
        if (left) {
          dst[left_run + wl_before + __popc(bl & lt)] = row;
        } else {
          const int rights_before = warp * 32 - wl_before + __popc((bv & ~bl) & lt);
          dst[left_total + right_run + rights_before] = row;
        }
      }
      const int tile_rows = min(kPartThreads, c_end - base);
      left_run += tile_left;
      right_run += tile_rows - tile_left;
      __syncthreads();
    }
  }
  // ---- last CTA: SerialTreeLearner::SplitInner bookkeeping (serial_tree_learner.cpp:776-907)
  __threadfence();
  __syncthreads();
  if (tid == 0) {
    const unsigned prev = atomicAdd(&t.part_done_counter, 1u);
    s_is_last = (prev == gridDim.x - 1);
  }
  __syncthreads();
  if (!s_is_last) return;
  __threadfence();
  const int p = t.split_leaf;
  const int r = t.num_leaves_cur;
  const int F = d.p.num_features;
  const BestSplit bs = L.best;
  const int nl = left_total, nr = n - left_total;
  const bool multi = d.p.num_ranks > 1;
  // smaller / larger: true counts on one GPU (serial_tree_learner.cpp:855); the split's own (rank-invariant)
  // counts when rows are sharded -- with exact integer histograms the choice does not change any result.
  const bool left_smaller = multi ? (bs.left_count < bs.right_count) : (nl < nr);
  const int parent_slot = L.slot;
  const int new_slot = r;  // == round + 1, fresh
  // children inherit the parent's is_splittable flags (serial_tree_learner.cpp:394-399)
  for (int f = tid; f < F; f += kPartThreads) {
    d.splittable[static_cast<long long>(new_slot) * F + f] = d.splittable[static_cast<long long>(parent_slot) * F + f];
  }
  if (tid == 0) {
    LeafState& A = d.leaves[p];
    LeafState& B = d.leaves[r];
    B.begin = L.begin + nl;
    B.count = nr;
    B.global_count = nr;
    B.depth = L.depth + 1;
    B.buf = L.buf ^ 1;
    B.sum_g = bs.right_sum_g;
    B.sum_h = bs.right_sum_h;
    B.ig = bs.right_ig;
    B.ih = bs.right_ih;
    B.output = bs.right_output;
    B.best.gain = -INFINITY;
    B.best.feature = -1;
    B.best.real_feature = -1;
    A.count = nl;
    A.global_count = nl;
    A.depth = L.depth + 1;
    A.buf = L.buf ^ 1;
    A.sum_g = bs.left_sum_g;
    A.sum_h = bs.left_sum_h;
    A.ig = bs.left_ig;
    A.ih = bs.left_ih;
    A.output = bs.left_output;
    A.slot = left_smaller ? new_slot : parent_slot;
    B.slot = left_smaller ? parent_slot : new_slot;
    d.leaf_values[p] = bs.left_output;
    d.leaf_values[r] = bs.right_output;
    gbl_split_record& rec = d.records[t.num_splits];
    rec.left_count = nl;
    rec.right_count = nr;
    t.smaller = left_smaller ? p : r;
    t.larger = left_smaller ? r : p;
    const int ns = left_smaller ? nl : nr, ng = left_smaller ? nr : nl;
    long long* tail = d.hist_pool + static_cast<long long>(new_slot) * d.slot_stride + d.slot_stride - kTailWords;
    tail[0] = ns;
    tail[1] = ng;
    // BeforeFindBestSplit shortcut when the global counts are already known (single GPU)
    int skip = 0;
    if (!multi) {
      if (d.p.max_depth > 0 && A.depth >= d.p.max_depth) skip = 1;
      if (nl < d.p.min_data_in_leaf * 2 && nr < d.p.min_data_in_leaf * 2) skip = 1;
    }
    t.skip_hist = skip;
    if (!skip && r + 1 < d.p.num_leaves) {
      t.hist_rows_total += ns;
      t.hist_launches += 1;
    }
    t.split_left_total = nl;
    t.num_leaves_cur = r + 1;
    t.num_splits += 1;
    t.do_split = 0;
    t.part_done_counter = 0;
  }
  (void)round;
}

// score[row] += scale * leaf_value[leaf] + bias for every leaf segment; grid.y = leaf
__global__ void __launch_bounds__(256) add_score_kernel(LearnerDev d, const double* __restrict__ leaf_values,
                                                        double scale, double bias, double* __restrict__ score) {
  const int leaf = blockIdx.y;
  if (leaf >= d.ts->num_leaves_cur) return;
  const LeafState& L = d.leaves[leaf];
  const double v = scale * leaf_values[leaf] + bias;
  const int* __restrict__ src = d.idx[L.buf] + L.begin;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < L.count; i += gridDim.x * blockDim.x) score[src[i]] += v;
}

__global__ void __launch_bounds__(256) leaf_assign_kernel(LearnerDev d, int* __restrict__ out) {
  const int leaf = blockIdx.y;
  if (leaf >= d.ts->num_leaves_cur) return;
  const LeafState& L = d.leaves[leaf];
  const int* __restrict__ src = d.idx[L.buf] + L.begin;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < L.count; i += gridDim.x * blockDim.x) out[src[i]] = leaf;
}

void launch_partition(const LearnerDev& d, int round, cudaStream_t s) {
  GBL_LAUNCH(part_count_kernel, kPartGrid, kPartThreads, 0, s, d);
  GBL_LAUNCH(part_scatter_kernel, kPartGrid, kPartThreads, 0, s, d, round);
}

void launch_add_score(const LearnerDev& d, const double* leaf_values_dev, double scale, double bias, double* score,
                      cudaStream_t s) {
  dim3 grid(64, d.p.num_leaves);
  GBL_LAUNCH(add_score_kernel, grid, 256, 0, s, d, leaf_values_dev, scale, bias, score);
}

void launch_leaf_assignment(const LearnerDev& d, int* out, cudaStream_t s) {
  dim3 grid(64, d.p.num_leaves);
  GBL_LAUNCH(leaf_assign_kernel, grid, 256, 0, s, d, out);
}

}  // namespace gbl
