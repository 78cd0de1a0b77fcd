// partition.cu -- DataPartition::Split as a stable segmented two-way partition of row indices for every node of the
// current wave at once, the node bookkeeping that follows a split, and the per-leaf score update.
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
//
// A wave splits many nodes with three launches.  Work is cut into units of kPartUnit rows of one node:
//   part_count   : one gather of the split feature's bin per row -> go-left bit per row (kept in HBM, 1 bit/row)
//                  and the number of left rows per unit;
//   part_plan    : one CTA: prefix over the units, then per node: child nodes, smaller/larger child, histogram slot,
//                  BeforeFindBestSplit gates and the histogram work list of the wave;
//   part_scatter : reads the bits back (no second gather) and writes both children in stable order; also clears the
//                  wave's fresh histogram slots and hands the parent's is_splittable flags to them.
#include "kernels.cuh"

namespace gbl {

constexpr int kPartThreads = 256;
constexpr int kPartWarps = kPartThreads / 32;
constexpr int kPartGrid = kNumSM * 8;
constexpr int kUnitWords = kPartUnit / 32;
constexpr int kPlanThreads = 1024;

struct SplitRule {
  const uint8_t* col;
  int bits;  // 8 or 4 (packed column)
  int threshold;
  int missing_bin;  // -1 = none
  bool default_left;
  const unsigned* cat;  // categorical split: bitset of the bins that go left (DenseBin::SplitCategoricalInner,
                        // dense_bin.hpp:450-483: in mapper-bin terms "left iff the bin is in the set"), else nullptr
};

__device__ inline SplitRule load_rule(const LearnerDev& d, const BestSplit& bs) {
  const FeatMeta m = d.meta[bs.feature];
  SplitRule r;
  r.col = d.col_bins + m.col_off;
  r.bits = m.bits;
  r.threshold = static_cast<int>(bs.threshold);
  r.missing_bin = -1;
  if (m.missing_type == GBL_MISSING_ZERO) r.missing_bin = m.default_bin;
  if (m.missing_type == GBL_MISSING_NAN) r.missing_bin = m.num_bin - 1;
  r.default_left = bs.default_left != 0;
  r.cat = bs.num_cat > 0 ? bs.cat_bits : nullptr;
  return r;
}

__device__ inline bool goes_left(const SplitRule& r, int row) {
  const int bin = r.bits == 4 ? ((__ldg(r.col + (row >> 1)) >> ((row & 1) << 2)) & 0xf) : __ldg(r.col + row);
  if (r.cat != nullptr) return (__ldg(r.cat + (bin >> 5)) >> (bin & 31)) & 1u;
  return bin == r.missing_bin ? r.default_left : bin <= r.threshold;
}

// wave entry owning partition unit u: the last entry whose first unit is <= u (entries without rows own nothing)
__device__ inline int part_owner(const Expand* __restrict__ w, int n, int u) {
  int lo = 0, hi = n;  // first entry with part_unit_begin > u lies in (lo, hi]
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (w[mid].part_unit_begin > u) {
      hi = mid;
    } else {
      lo = mid + 1;
    }
  }
  return lo - 1;
}

// ------------------------------------------------------------------------------------------------ pass 1
__global__ void __launch_bounds__(kPartThreads) part_count_kernel(LearnerDev d) {
  const TreeState& t = *d.ts;
  if (t.done) return;
  const int W = t.part_units, C = t.num_cand;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __shared__ int s_cnt[kPartWarps];
  for (int u = blockIdx.x; u < W; u += gridDim.x) {
    const int ci = part_owner(d.wave, C, u);
    const Expand w = d.wave[ci];
    const Node& nd = d.nodes[w.node];
    const int c_begin = (u - w.part_unit_begin) * kPartUnit;
    const int c_end = min(nd.count, c_begin + kPartUnit);
    const int* __restrict__ src = d.idx[nd.buf] + nd.begin;
    const SplitRule rule = load_rule(d, nd.best);
    unsigned* __restrict__ bits = d.unit_bits + static_cast<size_t>(u) * kUnitWords;
    // all row loads first, then all bin gathers, then the ballots: 8 independent loads in flight per thread
    constexpr int kIters = kPartUnit / kPartThreads;
    int rows[kIters];
    int bins[kIters];
#pragma unroll
    for (int k = 0; k < kIters; ++k) {
      const int i = c_begin + k * kPartThreads + warp * 32 + lane;
      rows[k] = i < c_end ? src[i] : -1;
    }
    if (rule.bits == 4) {  // CTA-uniform
#pragma unroll
      for (int k = 0; k < kIters; ++k)
        bins[k] = rows[k] >= 0 ? ((__ldg(rule.col + (rows[k] >> 1)) >> ((rows[k] & 1) << 2)) & 0xf) : -1;
    } else {
#pragma unroll
      for (int k = 0; k < kIters; ++k) bins[k] = rows[k] >= 0 ? static_cast<int>(__ldg(rule.col + rows[k])) : -1;
    }
    int cnt = 0;
#pragma unroll
    for (int k = 0; k < kIters; ++k) {
      bool left;
      if (rule.cat != nullptr) {  // CTA-uniform
        left = bins[k] >= 0 && ((__ldg(rule.cat + (bins[k] >> 5)) >> (bins[k] & 31)) & 1u) != 0u;
      } else {
        left = bins[k] >= 0 && (bins[k] == rule.missing_bin ? rule.default_left : bins[k] <= rule.threshold);
      }
      const unsigned bl = __ballot_sync(0xffffffffu, left);
      if (lane == 0 && c_begin + k * kPartThreads + warp * 32 < c_end) {
        bits[k * kPartWarps + warp] = bl;
        cnt += __popc(bl);
      }
    }
    if (lane == 0) s_cnt[warp] = cnt;
    __syncthreads();
    if (threadIdx.x == 0) {
      int tot = 0;
#pragma unroll
      for (int x = 0; x < kPartWarps; ++x) tot += s_cnt[x];
      d.unit_left[u] = tot;
    }
    __syncthreads();
  }
}

// ------------------------------------------------------------------------------------------------ plan
// exclusive prefix over the CTA's threads; `carry` accumulates the running total across calls
__device__ inline int plan_excl_scan(int v, int* s_warp, int& total) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int x = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int n = __shfl_up_sync(0xffffffffu, x, o);
    if (lane >= o) x += n;
  }
  __syncthreads();
  if (lane == 31) s_warp[warp] = x;
  __syncthreads();
  // the 32 warp totals: every warp scans them once with shuffles (lane k holds warp k's total)
  static_assert(kPlanThreads / 32 == 32, "one lane per warp total");
  int wt = s_warp[lane];
  int ws = wt;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int n = __shfl_up_sync(0xffffffffu, ws, o);
    if (lane >= o) ws += n;
  }
  total = __shfl_sync(0xffffffffu, ws, 31);
  const int before = __shfl_sync(0xffffffffu, ws - wt, warp);
  return before + x - v;
}

__global__ void __launch_bounds__(kPlanThreads) part_plan_kernel(LearnerDev d) {
  TreeState& t = *d.ts;
  if (t.done) return;
  __shared__ int s_warp[kPlanThreads / 32];
  __shared__ long long s_hist_rows, s_part_rows;
  const int tid = threadIdx.x;
  const int W = t.part_units, C = t.num_cand;
  const int node_base = t.num_nodes, slot_base = t.num_slots;
  const bool multi = d.p.num_ranks > 1;
  const bool last_wave = t.last_wave != 0;
  if (tid == 0) {
    s_hist_rows = 0;
    s_part_rows = 0;
  }
  // ---- exclusive prefix of the per-unit left counts (in place); unit_left[W] = total.  Eight consecutive units per
  // thread and one block scan per 8 192 units: a level of the Higgs shape (5 127 units) is one pass instead of six.
  int carry = 0;
  constexpr int kPer = 8;
  for (int base = 0; base < W; base += kPlanThreads * kPer) {
    const int u0 = base + tid * kPer;
    int v[kPer];
#pragma unroll
    for (int k = 0; k < kPer; ++k) v[k] = u0 + k < W ? d.unit_left[u0 + k] : 0;
    int sum = 0;
#pragma unroll
    for (int k = 0; k < kPer; ++k) sum += v[k];
    int tot;
    int run = carry + plan_excl_scan(sum, s_warp, tot);
#pragma unroll
    for (int k = 0; k < kPer; ++k) {
      if (u0 + k < W) d.unit_left[u0 + k] = run;
      run += v[k];
    }
    carry += tot;
  }
  if (tid == 0) d.unit_left[W] = carry;
  __syncthreads();
  // ---- children of every expanded node (SerialTreeLearner::SplitInner, serial_tree_learner.cpp:776-907)
  for (int ci = tid; ci < C; ci += kPlanThreads) {
    Expand& w = d.wave[ci];
    Node& P = d.nodes[w.node];
    const int ub = w.part_unit_begin, ue = d.wave[ci + 1].part_unit_begin;
    const int nl = d.unit_left[ue] - d.unit_left[ub];
    const int nr = P.count - nl;
    const BestSplit bs = P.best;
    // smaller / larger: true counts on one GPU (serial_tree_learner.cpp:855); the split's own (rank-invariant)
    // counts when rows are sharded -- with exact integer histograms the choice does not change any result.
    const bool left_smaller = multi ? (bs.left_count < bs.right_count) : (nl < nr);
    const int a = node_base + 2 * ci, b = a + 1;
    const int fresh = slot_base + ci;
    Node& A = d.nodes[a];
    Node& B = d.nodes[b];
    A.begin = P.begin;
    A.count = nl;
    A.global_count = nl;
    A.sum_g = bs.left_sum_g;
    A.sum_h = bs.left_sum_h;
    A.ig = bs.left_ig;
    A.ih = bs.left_ih;
    A.output = bs.left_output;
    A.slot = left_smaller ? fresh : P.slot;
    B.begin = P.begin + nl;
    B.count = nr;
    B.global_count = nr;
    B.sum_g = bs.right_sum_g;
    B.sum_h = bs.right_sum_h;
    B.ig = bs.right_ig;
    B.ih = bs.right_ih;
    B.output = bs.right_output;
    B.slot = left_smaller ? P.slot : fresh;
    // BasicLeafConstraints::Update (monotone_constraints.hpp:487-503): both children inherit the leaf's interval; a
    // split on a constrained feature tightens it at the mid-point of the two outputs (left child = `leaf`, right = new)
    A.cmin = B.cmin = P.cmin;
    A.cmax = B.cmax = P.cmax;
    if (d.p.use_mc) {
      const int mt = d.meta[bs.feature].monotone_type;
      const double mid = (bs.left_output + bs.right_output) / 2.0f;
      if (mt < 0) {
        A.cmin = fmax(mid, A.cmin);
        B.cmax = fmin(mid, B.cmax);
      } else if (mt > 0) {
        A.cmax = fmin(mid, A.cmax);
        B.cmin = fmax(mid, B.cmin);
      }
    }
    Node* ch[2] = {&A, &B};
    for (int k = 0; k < 2; ++k) {
      Node& X = *ch[k];
      X.depth = P.depth + 1;
      X.buf = child_buf(P.buf);
      X.parent = w.node;
      X.left = -1;
      X.ref_leaf = -1;
      X.in_tree = 0;
      X.bval = -INFINITY;
      X.best.gain = -INFINITY;
      X.best.feature = -1;
      X.best.real_feature = -1;
    }
    P.left = a;
    d.expanded[w.node] = 1;
    w.left_total = nl;
    w.hist_node = left_smaller ? a : b;
    w.hist_slot = fresh;
    const int ns = left_smaller ? nl : nr, ng = left_smaller ? nr : nl;
    long long* tail = d.hist_pool + static_cast<long long>(fresh) * d.slot_stride + d.slot_stride - kTailWords;
    tail[0] = ns;
    tail[1] = ng;
    // BeforeFindBestSplit shortcut when the global counts are already known (single GPU)
    int skip = last_wave ? 1 : 0;
    if (!multi) {
      if (d.p.max_depth > 0 && P.depth + 1 >= d.p.max_depth) skip = 1;
      if (nl < d.p.min_data_in_leaf * 2 && nr < d.p.min_data_in_leaf * 2) skip = 1;
    }
    w.skip = skip;
    w.hist_unit_begin = skip ? 0 : (ns + kHistUnit - 1) / kHistUnit;  // unit count, prefixed below
    if (!skip) atomicAdd(reinterpret_cast<unsigned long long*>(&s_hist_rows), static_cast<unsigned long long>(ns));
    atomicAdd(reinterpret_cast<unsigned long long*>(&s_part_rows), static_cast<unsigned long long>(P.count));
  }
  __syncthreads();
  // ---- histogram work list: prefix of the unit counts
  carry = 0;
  for (int base = 0; base < C; base += kPlanThreads) {
    const int ci = base + tid;
    const int v = ci < C ? d.wave[ci].hist_unit_begin : 0;
    int tot;
    const int ex = plan_excl_scan(v, s_warp, tot);
    if (ci < C) d.wave[ci].hist_unit_begin = carry + ex;
    carry += tot;
  }
  if (tid == 0) {
    d.wave[C].hist_unit_begin = carry;
    t.hist_units = carry;
    t.num_nodes = node_base + 2 * C;
    t.num_slots = slot_base + C;
    t.num_expanded += C;
    t.scan_jobs = last_wave ? 0 : 2 * C;
    t.waves += 1;
    t.part_rows_total += s_part_rows;
    if (carry > 0) {
      t.hist_rows_total += s_hist_rows;
      t.hist_launches += 1;
    }
  }
}

// ------------------------------------------------------------------------------------------------ pass 2
__global__ void __launch_bounds__(kPartThreads) part_scatter_kernel(LearnerDev d) {
  const TreeState& t = *d.ts;
  if (t.done) return;
  const int W = t.part_units, C = t.num_cand;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  __shared__ unsigned s_word[kUnitWords];
  __shared__ int s_pref[kUnitWords];
  const unsigned lt = (1u << lane) - 1u;
  for (int u = blockIdx.x; u < W; u += gridDim.x) {
    const int ci = part_owner(d.wave, C, u);
    const Expand w = d.wave[ci];
    const Node& nd = d.nodes[w.node];
    const int c_begin = (u - w.part_unit_begin) * kPartUnit;
    const int c_end = min(nd.count, c_begin + kPartUnit);
    const int nwords = (c_end - c_begin + 31) >> 5;
    const int lefts_before = d.unit_left[u] - d.unit_left[w.part_unit_begin];
    const int* __restrict__ src = d.idx[nd.buf] + nd.begin;
    int* __restrict__ dst = d.idx[child_buf(nd.buf)] + nd.begin;
    const unsigned* __restrict__ bits = d.unit_bits + static_cast<size_t>(u) * kUnitWords;
    if (tid < kUnitWords) {  // two warps: exclusive prefix of the words' popcounts
      const unsigned word = tid < nwords ? bits[tid] : 0u;
      s_word[tid] = word;
      const int pc = __popc(word);
      int x = pc;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int n = __shfl_up_sync(0xffffffffu, x, o);
        if (lane >= o) x += n;
      }
      s_pref[tid] = x - pc;
    }
    __syncthreads();
    // total of the first warp's words, added to the second warp's prefixes on the fly
    const int first_total = s_pref[31] + __popc(s_word[31]);
    const int right_base = w.left_total + (c_begin - lefts_before);
#pragma unroll
    for (int k = 0; k < kUnitWords / kPartWarps; ++k) {
      const int wi = k * kPartWarps + warp;
      const int i = c_begin + wi * 32 + lane;
      if (wi < nwords && i < c_end) {
        const int row = src[i];
        const unsigned word = s_word[wi];
        const int lb = s_pref[wi] + (wi >= 32 ? first_total : 0) + __popc(word & lt);
        if ((word >> lane) & 1u) {
          dst[lefts_before + lb] = row;
        } else {
          dst[right_base + (wi * 32 + lane - lb)] = row;
        }
      }
    }
    __syncthreads();
  }
  // ---- fresh histogram slots of the wave start at zero (their tail words were set by the plan kernel), and
  // children inherit the parent's is_splittable flags (serial_tree_learner.cpp:394-399)
  if (t.last_wave) return;
  const long long words_per = d.slot_stride - kTailWords;  // multiple of 8192
  const long long pairs_per = words_per >> 1;
  const long long gtid = static_cast<long long>(blockIdx.x) * kPartThreads + tid;
  const long long gsize = static_cast<long long>(gridDim.x) * kPartThreads;
  const int slot0 = d.wave[0].hist_slot;
  for (long long e = gtid; e < pairs_per * C; e += gsize) {
    const long long s = e / pairs_per, o = e - s * pairs_per;
    if (d.wave[s].skip) continue;
    reinterpret_cast<longlong2*>(d.hist_pool + (slot0 + s) * d.slot_stride)[o] = make_longlong2(0, 0);
  }
  const int F = d.p.num_features;
  for (long long e = gtid; e < static_cast<long long>(C) * F; e += gsize) {
    const int ci = static_cast<int>(e / F), f = static_cast<int>(e - static_cast<long long>(ci) * F);
    const Expand w = d.wave[ci];
    const Node& nd = d.nodes[w.node];
    // the parent's slot (and its flags) stays with the larger child; the fresh slot gets a copy
    d.splittable[static_cast<long long>(w.hist_slot) * F + f] = d.splittable[static_cast<long long>(nd.slot) * F + f];
  }
}

// ------------------------------------------------------------------------------------------------ score update
// Row segments = leaves of the grown tree (a final leaf that was expanded speculatively is the union of several);
// work units of kSegUnit rows, prefix in seg_unit_begin (built by the replay).
__device__ inline int seg_owner(const int* __restrict__ begin, int n, int u) {
  int lo = 0, hi = n;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (begin[mid] > u) {
      hi = mid;
    } else {
      lo = mid + 1;
    }
  }
  return lo - 1;
}

// score[row] += scale * leaf_value[leaf] + bias
__global__ void __launch_bounds__(256) add_score_kernel(LearnerDev d, const double* __restrict__ leaf_values,
                                                        double scale, double bias, double* __restrict__ score) {
  const TreeState& t = *d.ts;
  const int U = t.seg_units, S = t.num_segments;
  for (int u = blockIdx.x; u < U; u += gridDim.x) {
    const int si = seg_owner(d.seg_unit_begin, S, u);
    const Node& nd = d.nodes[d.seg_node[si]];
    const double v = scale * leaf_values[nd.ref_leaf] + bias;
    const int c_begin = (u - d.seg_unit_begin[si]) * kSegUnit;
    const int c_end = min(nd.count, c_begin + kSegUnit);
    const int* __restrict__ src = d.idx[nd.buf] + nd.begin;
    for (int i = c_begin + threadIdx.x; i < c_end; i += 256) score[src[i]] += v;
  }
}

// Two-step score update for trees of at most 65535 leaves.  The scattered read-modify-write of 8-byte scores above
// moves a 32-byte sector per row in both directions; writing a 2-byte leaf id per row instead is a plain store into an
// array that stays in L2 (2 bytes x num_data), and the scores are then updated by one streaming pass.
__global__ void __launch_bounds__(256) leaf_mark_kernel(LearnerDev d, unsigned short* __restrict__ out) {
  const TreeState& t = *d.ts;
  const int U = t.seg_units, S = t.num_segments;
  for (int u = blockIdx.x; u < U; u += gridDim.x) {
    const int si = seg_owner(d.seg_unit_begin, S, u);
    const Node& nd = d.nodes[d.seg_node[si]];
    const unsigned short leaf = static_cast<unsigned short>(nd.ref_leaf);
    const int c_begin = (u - d.seg_unit_begin[si]) * kSegUnit;
    const int c_end = min(nd.count, c_begin + kSegUnit);
    const int* __restrict__ src = d.idx[nd.buf] + nd.begin;
    for (int i = c_begin + threadIdx.x; i < c_end; i += 256) out[src[i]] = leaf;
  }
}

// score[row] += scale * leaf_value[leaf(row)] + bias, four rows per thread; 0xFFFF = row outside the tree's rows (bagging)
__global__ void __launch_bounds__(256) add_score_dense_kernel(const unsigned short* __restrict__ leaf,
                                                              const double* __restrict__ leaf_values, int num_leaves,
                                                              double scale, double bias, double* __restrict__ score,
                                                              int n) {
  extern __shared__ double s_val[];
  for (int l = threadIdx.x; l < num_leaves; l += 256) s_val[l] = scale * leaf_values[l] + bias;
  __syncthreads();
  const int n4 = n >> 2;
  for (int q = blockIdx.x * 256 + threadIdx.x; q < n4; q += gridDim.x * 256) {
    const ushort4 l = reinterpret_cast<const ushort4*>(leaf)[q];
    double2 a = reinterpret_cast<double2*>(score)[2 * q];
    double2 b = reinterpret_cast<double2*>(score)[2 * q + 1];
    if (l.x != 0xFFFF) a.x += s_val[l.x];
    if (l.y != 0xFFFF) a.y += s_val[l.y];
    if (l.z != 0xFFFF) b.x += s_val[l.z];
    if (l.w != 0xFFFF) b.y += s_val[l.w];
    reinterpret_cast<double2*>(score)[2 * q] = a;
    reinterpret_cast<double2*>(score)[2 * q + 1] = b;
  }
  if (blockIdx.x == 0) {
    for (int i = (n4 << 2) + threadIdx.x; i < n; i += 256) {
      const unsigned short l = leaf[i];
      if (l != 0xFFFF) score[i] += s_val[l];
    }
  }
}

__global__ void __launch_bounds__(256) leaf_assign_kernel(LearnerDev d, int* __restrict__ out) {
  const TreeState& t = *d.ts;
  const int U = t.seg_units, S = t.num_segments;
  for (int u = blockIdx.x; u < U; u += gridDim.x) {
    const int si = seg_owner(d.seg_unit_begin, S, u);
    const Node& nd = d.nodes[d.seg_node[si]];
    const int c_begin = (u - d.seg_unit_begin[si]) * kSegUnit;
    const int c_end = min(nd.count, c_begin + kSegUnit);
    const int* __restrict__ src = d.idx[nd.buf] + nd.begin;
    for (int i = c_begin + threadIdx.x; i < c_end; i += 256) out[src[i]] = nd.ref_leaf;
  }
}

// score[row] += scale * value(leaf(row)) + bias for rows that were NOT part of the tree's training subset (bagging, or
// the rows of a validation dataset binned with the same mappers; rows == nullptr: every row):
// the row walks the final tree on its bins (Tree::AddPredictionToScore with used_data_indices, src/io/tree.cpp:193-230;
// the decision of Tree::NumericalDecisionInner, include/LightGBM/tree.h:329-350, in mapper-bin terms).  The tree is
// the device-resident one: nodes the replay marked in_tree are inner nodes, the first node that is not is the leaf.
__global__ void __launch_bounds__(256) add_score_rows_kernel(LearnerDev d, const uint8_t* __restrict__ row_bins,
                                                             int row_stride, const int* __restrict__ rows, int num_rows,
                                                             const double* __restrict__ leaf_values, double scale,
                                                             double bias, double* __restrict__ score) {
  if (d.ts->num_splits == 0) return;  // a tree without splits changes no score
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < num_rows; i += gridDim.x * blockDim.x) {
    const int row = rows != nullptr ? rows[i] : i;
    const uint8_t* bins = row_bins + static_cast<size_t>(row) * row_stride;
    int n = 0;
    while (d.nodes[n].in_tree) {
      const Node& nd = d.nodes[n];
      const FeatMeta& m = d.meta[nd.best.feature];
      const int bin = bins[nd.best.feature];
      int missing_bin = -1;
      if (m.missing_type == GBL_MISSING_ZERO) missing_bin = m.default_bin;
      if (m.missing_type == GBL_MISSING_NAN) missing_bin = m.num_bin - 1;
      bool left = bin == missing_bin ? nd.best.default_left != 0 : bin <= static_cast<int>(nd.best.threshold);
      if (nd.best.num_cat > 0) left = ((nd.best.cat_bits[bin >> 5] >> (bin & 31)) & 1u) != 0u;  // Tree::CategoricalDecisionInner
      n = nd.left + (left ? 0 : 1);
    }
    score[row] += scale * leaf_values[d.nodes[n].ref_leaf] + bias;
  }
}

void launch_add_score_rows(const LearnerDev& d, const uint8_t* row_bins, int row_stride, const int* rows, int num_rows,
                           const double* leaf_values_dev, double scale, double bias, double* score, cudaStream_t s) {
  if (num_rows <= 0) return;
  GBL_LAUNCH(add_score_rows_kernel, kPartGrid, 256, 0, s, d, row_bins, row_stride, rows, num_rows, leaf_values_dev, scale,
             bias, score);
}

void launch_partition(const LearnerDev& d, cudaStream_t s) {
  GBL_LAUNCH(part_count_kernel, kPartGrid, kPartThreads, 0, s, d);
  GBL_LAUNCH(part_plan_kernel, 1, kPlanThreads, 0, s, d);
  GBL_LAUNCH(part_scatter_kernel, kPartGrid, kPartThreads, 0, s, d);
}

void launch_add_score(const LearnerDev& d, const double* leaf_values_dev, double scale, double bias, double* score,
                      unsigned short* leaf_scratch, bool partial_rows, cudaStream_t s) {
  const int nl = d.p.num_leaves;
  // the streaming pass wants a 16-byte aligned score vector; anything else takes the scattered update
  if (leaf_scratch == nullptr || nl > 4096 || (reinterpret_cast<uintptr_t>(score) & 15) != 0) {
    GBL_LAUNCH(add_score_kernel, kPartGrid, 256, 0, s, d, leaf_values_dev, scale, bias, score);
    return;
  }
  if (partial_rows) GBL_CUDA(cudaMemsetAsync(leaf_scratch, 0xFF, sizeof(unsigned short) * d.num_data, s));
  GBL_LAUNCH(leaf_mark_kernel, kPartGrid, 256, 0, s, d, leaf_scratch);
  GBL_LAUNCH(add_score_dense_kernel, kNumSM * 8, 256, sizeof(double) * nl, s, leaf_scratch, leaf_values_dev, nl, scale, bias,
             score, d.num_data);
}

void launch_leaf_mark(const LearnerDev& d, unsigned short* out, bool partial_rows, cudaStream_t s) {
  if (partial_rows) GBL_CUDA(cudaMemsetAsync(out, 0xFF, sizeof(unsigned short) * d.num_data, s));
  GBL_LAUNCH(leaf_mark_kernel, kPartGrid, 256, 0, s, d, out);
}

void launch_leaf_assignment(const LearnerDev& d, int* out, cudaStream_t s) {
  GBL_LAUNCH(leaf_assign_kernel, kPartGrid, 256, 0, s, d, out);
}

}  // namespace gbl
