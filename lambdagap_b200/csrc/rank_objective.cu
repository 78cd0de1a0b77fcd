// rank_objective.cu -- lambdarank gradients for all 18 `lambdarank_target` values of the LambdaGap fork, one CTA
// per query.
//
// Reference: LambdarankNDCG (src/objective/rank_objective.hpp:174-640): GetGradientsForOneQuery :253-524,
// GetSigmoid / ConstructSigmoidTable :526-552, Init :225-251; RankingObjective::GetGradients :82-121 (per-row
// weights :98-105); DCGCalculator (src/metric/dcg_calculator.cpp:14-96).  The reference's own CUDA kernel
// (src/objective/cuda/cuda_rank_objective.cu:19-188) knows only the plain NDCG target and sorts fp32 copies of the
// scores; this one sorts the fp64 scores themselves.
//
// Design.  (1) The CTA sorts (score desc, original index asc) with a shared-memory bitonic network -- the index
// tie-break makes it equal to std::stable_sort.  (2) *Gather* formulation: one thread per document walks that
// document's pairs in exactly the order the CPU's double loop touches it -- first as the `j` member for ascending
// `i`, then as the `i` member for ascending `j` -- so the fp32 `+=` sequence of every lambda / hessian is the
// CPU's, with no atomics.  Every pair is therefore evaluated twice (once per member); pair math is fp64 as in the
// reference.  Only the fp64 `sum_lambdas` reduction order differs (block reduction instead of pair order).
// (3) The 2^20-entry sigmoid table, the discount table and the per-query inverse max (B)DCG are built on the host
// with the reference's formulas (libm exp / log2) and uploaded once, so table look-ups are bit-identical.
#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <vector>

#include "objective.cuh"

namespace gbl {

enum RankTarget {
  kNDCG = 0, kLambdaLossNDCG, kLambdaLossNDCGpp, kBNDCG, kLambdaLossBNDCG, kLambdaLossBNDCGpp, kPrecision, kARPK,
  kLambdaLossARP1, kLambdaLossARP2, kRankNet, kBinRankNet, kLambdaGapS, kLambdaGapX, kLambdaGapSp, kLambdaGapXp,
  kLambdaGapSpp, kLambdaGapXpp, kNumRankTargets
};

constexpr int kSigmoidBins = 1024 * 1024;  // rank_objective.hpp:631
constexpr int kMaxPosition = 10000;        // dcg_calculator.cpp:17

struct RankParams {
  int target;
  int truncation_level;
  int norm;
  double sigmoid;
  double lambdagap_weight;
  double min_sigmoid_input, max_sigmoid_input, sigmoid_table_idx_factor;
  const double* sigmoid_table;
  const double* discount;
  const double* label_gain;
  int num_label_gain;
  const double* inverse_max_dcgs;
  const double* inverse_max_bdcgs;
  const int* query_boundaries;
  const float* label;
  const float* weight;
};

// Query size classes: one launch per class, shared memory laid out for the class's longest query, so that the many short
// queries keep many CTAs per SM and only the few very long ones pay for the big arrays.
constexpr int kNumClasses = 7;
static const int kClassLimit[kNumClasses] = {32, 64, 128, 256, 512, 1024, 16384};
static const int kClassThreads[kNumClasses] = {32, 64, 128, 128, 256, 512, 1024};

struct RankState {
  RankParams p{};
  double* sigmoid_table = nullptr;
  double* discount = nullptr;
  double* label_gain = nullptr;
  double* inv_dcg = nullptr;
  double* inv_bdcg = nullptr;
  int* query_boundaries = nullptr;
  int* class_queries[kNumClasses] = {};  // query ids per size class
  int class_count[kNumClasses] = {};
  int class_pad[kNumClasses] = {};  // array length (power of two) per class
  // the three size classes run side by side (the few very long queries would otherwise hold the GPU alone)
  cudaStream_t aux[2] = {nullptr, nullptr};
  cudaEvent_t fork = nullptr;
  cudaEvent_t join[2] = {nullptr, nullptr};
};

__device__ inline bool is_binary_target(int t) {
  return t == kPrecision || t == kBNDCG || t == kLambdaLossBNDCG || t == kLambdaLossBNDCGpp || t == kARPK ||
         t == kBinRankNet || t == kLambdaGapS || t == kLambdaGapX || t == kLambdaGapSp || t == kLambdaGapXp ||
         t == kLambdaGapSpp || t == kLambdaGapXpp;
}
__device__ inline bool is_unsorted_target(int t) {
  return t == kBinRankNet || t == kRankNet || t == kLambdaLossARP1 || t == kLambdaLossARP2;
}
__device__ inline bool is_truncated_i(int t) { return t <= kPrecision; }  // rank_objective.hpp:304-317

__device__ inline double get_discount(const RankParams& p, int k) { return p.discount[min(k, kMaxPosition - 1)]; }

// rank_objective.hpp:526-537
__device__ inline double get_sigmoid(const RankParams& p, double score) {
  if (score <= p.min_sigmoid_input) return p.sigmoid_table[0];
  if (score >= p.max_sigmoid_input) return p.sigmoid_table[kSigmoidBins - 1];
  return p.sigmoid_table[static_cast<size_t>((score - p.min_sigmoid_input) * p.sigmoid_table_idx_factor)];
}

// [start, end) of the inner loop for outer rank i (rank_objective.hpp:325-351)
__device__ inline void j_range(int t, int i, int k, int cnt, int* start, int* end) {
  switch (t) {
    case kPrecision:
      *start = k;
      *end = cnt;
      break;
    case kARPK:
    case kLambdaGapSp:
    case kLambdaGapXp:
    case kLambdaGapSpp:
    case kLambdaGapXpp:
      *start = max(i + 1, k);
      *end = cnt;
      break;
    case kLambdaGapS:
      *start = i + k;
      *end = min(*start + 1, cnt);
      break;
    case kLambdaGapX:
      *start = i + k;
      *end = cnt;
      break;
    default:
      *start = i + 1;
      *end = cnt;
  }
}

// p_lambda / p_hessian of the pair (ranks i < j) and which member has the lower label; false = the reference skips it
template <bool SMEM_LABEL>
__device__ inline bool pair_terms(const RankParams& p, int qid, int i, int j, const double* s_score, const float* s_label,
                                  const int* s_idx, const float* g_label, bool do_norm, int k, double* out_lambda,
                                  double* out_hessian, int* out_low_rank) {
  const double score_i = s_score[i], score_j = s_score[j];
  if (score_i == -INFINITY || score_j == -INFINITY) return false;  // kMinScore, :320,354
  const float label_i = SMEM_LABEL ? s_label[i] : g_label[s_idx[i]];
  const float label_j = SMEM_LABEL ? s_label[j] : g_label[s_idx[j]];
  if (label_i == label_j) return false;
  const int t = p.target;
  if (label_i > 0 && label_j > 0 && is_binary_target(t)) return false;
  int high_rank, low_rank;
  float high_label_f, low_label_f;
  if (label_i > label_j) {
    high_rank = i;
    low_rank = j;
    high_label_f = label_i;
    low_label_f = label_j;
  } else {
    high_rank = j;
    low_rank = i;
    high_label_f = label_j;
    low_label_f = label_i;
  }
  const double high_score = s_score[high_rank], low_score = s_score[low_rank];
  const double delta_score = high_score - low_score;
  double delta_pair;
  const int rank_diff = j - i;
  switch (t) {
    case kNDCG: {
      const double dcg_gap = p.label_gain[static_cast<int>(high_label_f)] - p.label_gain[static_cast<int>(low_label_f)];
      const double paired_discount = fabs(get_discount(p, high_rank) - get_discount(p, low_rank));
      delta_pair = dcg_gap * paired_discount * p.inverse_max_dcgs[qid];
      break;
    }
    case kLambdaLossNDCG: {
      const double dcg_gap = p.label_gain[static_cast<int>(high_label_f)] - p.label_gain[static_cast<int>(low_label_f)];
      const double paired_discount = get_discount(p, rank_diff) - get_discount(p, rank_diff + 1);
      delta_pair = dcg_gap * paired_discount * p.inverse_max_dcgs[qid];
      break;
    }
    case kLambdaLossNDCGpp: {
      const double dcg_gap = p.label_gain[static_cast<int>(high_label_f)] - p.label_gain[static_cast<int>(low_label_f)];
      const double pd_rank = fabs(get_discount(p, high_rank) - get_discount(p, low_rank));
      const double pd_loss = get_discount(p, rank_diff) - get_discount(p, rank_diff + 1);
      delta_pair = dcg_gap * (pd_rank + p.lambdagap_weight * pd_loss) * p.inverse_max_dcgs[qid];
      break;
    }
    case kBNDCG:
      delta_pair = fabs(get_discount(p, high_rank) - get_discount(p, low_rank)) * p.inverse_max_bdcgs[qid];
      break;
    case kLambdaLossBNDCG:
      delta_pair = (get_discount(p, rank_diff) - get_discount(p, rank_diff + 1)) * p.inverse_max_bdcgs[qid];
      break;
    case kLambdaLossBNDCGpp: {
      const double pd_rank = fabs(get_discount(p, high_rank) - get_discount(p, low_rank));
      const double pd_loss = get_discount(p, rank_diff) - get_discount(p, rank_diff + 1);
      delta_pair = (pd_rank + p.lambdagap_weight * pd_loss) * p.inverse_max_bdcgs[qid];
      break;
    }
    case kPrecision:
    case kLambdaGapS:
    case kLambdaGapX:
    case kBinRankNet:
    case kRankNet:
      delta_pair = 1;
      break;
    case kLambdaGapSp:
      delta_pair = (rank_diff == k) * p.lambdagap_weight + (i < k);
      break;
    case kLambdaGapXp:
      delta_pair = (rank_diff >= k) * p.lambdagap_weight + (i < k);
      break;
    case kLambdaGapSpp:
      delta_pair = (rank_diff == k) * p.lambdagap_weight + (j + 1 - k) - (i >= k) * (i + 1 - k);
      break;
    case kLambdaGapXpp:
      delta_pair = (rank_diff >= k) * p.lambdagap_weight + (j + 1 - k) - (i >= k) * (i + 1 - k);
      break;
    case kARPK:
      delta_pair = (j + 1 - k) - (i >= k) * (i + 1 - k);
      break;
    case kLambdaLossARP1:
      delta_pair = static_cast<double>(high_label_f);
      break;
    default:  // kLambdaLossARP2
      delta_pair = static_cast<double>(high_label_f) - static_cast<double>(low_label_f);
      break;
  }
  if (delta_pair == 0) return false;
  if (do_norm) delta_pair /= (0.01f + fabs(delta_score));
  double p_lambda = get_sigmoid(p, delta_score);
  double p_hessian = p_lambda * (1.0f - p_lambda);
  p_lambda *= -p.sigmoid * delta_pair;
  p_hessian *= p.sigmoid * p.sigmoid * delta_pair;
  *out_lambda = p_lambda;
  *out_hessian = p_hessian;
  *out_low_rank = low_rank;
  return true;
}

// One pair (ranks i < j) seen from the document at rank `me` (== i or j).  Accumulates into (lam, hes) exactly
// what the reference adds to that document (rank_objective.hpp:353-513); returns -2 * p_lambda (0 if skipped).
template <bool SMEM_LABEL>
__device__ inline double eval_pair(const RankParams& p, int qid, int i, int j, int me, const double* s_score,
                                   const float* s_label, const int* s_idx, const float* g_label, bool do_norm, int k,
                                   float* lam, float* hes) {
  double p_lambda, p_hessian;
  int low_rank;
  if (!pair_terms<SMEM_LABEL>(p, qid, i, j, s_score, s_label, s_idx, g_label, do_norm, k, &p_lambda, &p_hessian, &low_rank)) return 0.0;
  if (me == low_rank) {
    *lam -= static_cast<float>(p_lambda);
  } else {
    *lam += static_cast<float>(p_lambda);
  }
  *hes += static_cast<float>(p_hessian);
  return -2 * p_lambda;
}


// std::nth_element of libstdc++ (introselect: median-of-3 Hoare partition, insertion sort below 4 elements) on the
// index array with comparator score[a] > score[b], run by one thread.  The `precision` target orders documents
// with nth_element only (rank_objective.hpp:268-271), whose arrangement is implementation-defined; replaying the
// library's algorithm reproduces the reference binary's arrangement, ties included.  (If the 2*lg(n) depth limit
// is ever exhausted the library switches to heap-select; that path is replaced by an insertion sort of the range.)
__device__ inline bool idx_before(const double* sc, int a, int b) { return sc[a] > sc[b]; }
__device__ void device_nth_element(int* v, int n, int nth, const double* sc) {
  if (n <= 0 || nth >= n) return;
  int first = 0, last = n;
  int depth = 0;
  for (int m = n; m > 1; m >>= 1) ++depth;
  depth *= 2;
  while (last - first > 3) {
    if (depth == 0) break;
    --depth;
    // __unguarded_partition_pivot
    const int mid = first + (last - first) / 2;
    {
      const int a = first + 1, b = mid, c = last - 1;
      int pick;
      if (idx_before(sc, v[a], v[b])) {
        pick = idx_before(sc, v[b], v[c]) ? b : (idx_before(sc, v[a], v[c]) ? c : a);
      } else {
        pick = idx_before(sc, v[a], v[c]) ? a : (idx_before(sc, v[b], v[c]) ? c : b);
      }
      const int tmp = v[first];
      v[first] = v[pick];
      v[pick] = tmp;
    }
    int lo = first + 1, hi = last;
    const int pivot = v[first];
    while (true) {
      while (idx_before(sc, v[lo], pivot)) ++lo;
      --hi;
      while (idx_before(sc, pivot, v[hi])) --hi;
      if (!(lo < hi)) break;
      const int tmp = v[lo];
      v[lo] = v[hi];
      v[hi] = tmp;
      ++lo;
    }
    const int cut = lo;
    if (cut <= nth) {
      first = cut;
    } else {
      last = cut;
    }
  }
  // __insertion_sort(first, last)
  for (int i = first + 1; i < last; ++i) {
    const int val = v[i];
    if (idx_before(sc, val, v[first])) {
      for (int m = i; m > first; --m) v[m] = v[m - 1];
      v[first] = val;
    } else {
      int m = i;
      while (idx_before(sc, val, v[m - 1])) {
        v[m] = v[m - 1];
        --m;
      }
      v[m] = val;
    }
  }
}

// smem layout: double score[n_pad]; int idx[n_pad]; float label[n_pad] (if SMEM_LABEL); double red[32];
// double raw_score[n_pad] (precision target only)
// pair_rows > 0: pair-parallel form (see below), with three more shared arrays behind the ones above:
// float lam[n_pad]; float hes[n_pad]; float2 pair[pair_rows][n_pad]
template <bool SMEM_LABEL>
__global__ void lambdarank_kernel(RankParams p, const int* __restrict__ queries, int n_pad,
                                  const double* __restrict__ score, float* __restrict__ grad,
                                  float* __restrict__ hess, int pair_rows) {
  extern __shared__ unsigned char smem_raw[];
  double* s_score = reinterpret_cast<double*>(smem_raw);
  int* s_idx = reinterpret_cast<int*>(s_score + n_pad);
  float* s_label = reinterpret_cast<float*>(s_idx + n_pad);
  // n_pad is a power of two >= 2, so every sub-array below stays 8-byte aligned
  double* s_red = reinterpret_cast<double*>(smem_raw + static_cast<size_t>(n_pad) * (8 + 4 + (SMEM_LABEL ? 4 : 0)));
  __shared__ double s_bw[2];

  const int qid = queries[blockIdx.x];
  const int start = p.query_boundaries[qid];
  const int cnt = p.query_boundaries[qid + 1] - start;
  const int tid = threadIdx.x, nt = blockDim.x;
  const double* q_score = score + start;
  const float* q_label = p.label + start;
  const int t = p.target;
  const int k = p.truncation_level;

  // ---- load + sort.  The arrays are laid out for the class-wide n_pad; the sort network only spans this query's own
  // power of two (a 130-document query of the 129..1024 class sorts 256 elements, not 1024)
  int np = 2;
  while (np < cnt) np <<= 1;
  // pair form: the labels are read here, beside the scores, into the (still unused) lam array and permuted from shared
  // memory after the sort, instead of a dependent global gather behind it
  const bool stage_label = SMEM_LABEL && pair_rows > 0;
  float* s_stage = reinterpret_cast<float*>(smem_raw + static_cast<size_t>(n_pad) * (8 + 4 + (SMEM_LABEL ? 4 : 0)) + 32 * 8 +
                                            (t == kPrecision ? static_cast<size_t>(n_pad) * 8 : 0));
  for (int i = tid; i < np; i += nt) {
    s_score[i] = i < cnt ? q_score[i] : -INFINITY;
    s_idx[i] = i;
    if (stage_label && i < cnt) s_stage[i] = q_label[i];
  }
  __syncthreads();
  const bool sorted = !is_unsorted_target(t);
  if (t == kPrecision) {
    double* s_raw = s_red + 32;
    for (int i = tid; i < cnt; i += nt) s_raw[i] = s_score[i];
    __syncthreads();
    if (tid == 0) device_nth_element(s_idx, cnt, min(k - 1, cnt - 1), s_raw);
    __syncthreads();
    for (int i = tid; i < cnt; i += nt) s_score[i] = s_raw[s_idx[i]];
    __syncthreads();
  } else if (sorted) {
    // bitonic network on (score desc, index asc): a total order, hence == std::stable_sort by score desc (:289-291)
    for (int size = 2; size <= np; size <<= 1) {
      for (int stride = size >> 1; stride > 0; stride >>= 1) {
        for (int x = tid; x < (np >> 1); x += nt) {
          const int lo = (x / stride) * (stride << 1) + (x % stride);
          const int hi = lo + stride;
          const bool up = ((lo & size) == 0);  // ascending position = "comes first" order
          const double sa = s_score[lo], sb = s_score[hi];
          const int ia = s_idx[lo], ib = s_idx[hi];
          const bool a_first = sa > sb || (sa == sb && ia < ib);
          if (a_first != up) {
            s_score[lo] = sb;
            s_score[hi] = sa;
            s_idx[lo] = ib;
            s_idx[hi] = ia;
          }
        }
        __syncthreads();
      }
    }
  }
  if (SMEM_LABEL) {
    for (int i = tid; i < cnt; i += nt) s_label[i] = stage_label ? s_stage[s_idx[i]] : q_label[s_idx[i]];
  }
  // best / worst score (:268-298)
  if (tid == 0) {
    double best, worst;
    if (sorted && t != kPrecision) {  // kPrecision: minmax_element, like the unsorted targets
      best = s_score[0];
      int worst_idx = cnt - 1;
      if (worst_idx > 0 && s_score[worst_idx] == -INFINITY) worst_idx -= 1;
      worst = s_score[worst_idx];
    } else {
      best = -INFINITY;
      worst = INFINITY;
      for (int i = 0; i < cnt; ++i) {  // std::minmax_element over the raw scores
        best = fmax(best, s_score[i]);
        worst = fmin(worst, s_score[i]);
      }
    }
    s_bw[0] = best;
    s_bw[1] = worst;
  }
  __syncthreads();
  const bool do_norm = p.norm && s_bw[0] != s_bw[1];
  const int i_end = is_truncated_i(t) ? min(cnt - 1, k) : cnt - 1;

  double my_sum = 0.0;
  if (pair_rows > 0) {
    // ---- pair-parallel form.  The gather below lets the thread of a top-ranked document walk all of its pairs alone (a
    // truncated target pairs the first k ranks with everybody: thread 0 of a 120-document query evaluates 119 pairs while
    // most threads evaluate 30) and evaluates every pair twice, once per member.  Here the outer ranks are taken in chunks
    // of pair_rows: (1) all threads evaluate the chunk's pairs once each, in parallel, into shared memory -- the signed
    // float lambda as seen from member i, and the float hessian; (2) one thread per document adds the chunk's
    // contributions to its running fp32 sums in the CPU's order: as the j member for ascending i, then (if it is in the
    // chunk) as the i member for ascending j.  Chunks ascend, so the += sequence of every document is still the CPU's.
    const size_t base = static_cast<size_t>(n_pad) * (8 + 4 + (SMEM_LABEL ? 4 : 0)) + 32 * 8 +
                        (t == kPrecision ? static_cast<size_t>(n_pad) * 8 : 0);
    float* s_lam = reinterpret_cast<float*>(smem_raw + base);
    float* s_hes = s_lam + n_pad;
    float2* s_pair = reinterpret_cast<float2*>(s_hes + n_pad);
    for (int r = tid; r < cnt; r += nt) {
      s_lam[r] = 0.f;
      s_hes[r] = 0.f;
    }
    const bool a_from_k = t == kPrecision || t == kARPK || t == kLambdaGapSp || t == kLambdaGapXp || t == kLambdaGapSpp || t == kLambdaGapXpp;
    for (int i0 = 0; i0 < i_end; i0 += pair_rows) {
      const int i1 = min(i_end, i0 + pair_rows);
      __syncthreads();  // the previous chunk's pairs have been consumed (and lam / hes are initialised)
      for (int i = i0; i < i1; ++i) {
        int js, je;
        j_range(t, i, k, cnt, &js, &je);
        float2* row = s_pair + static_cast<size_t>(i - i0) * cnt;
        for (int j = js + tid; j < je; j += nt) {
          double pl, ph;
          int low;
          float2 v = make_float2(0.f, 0.f);
          if (pair_terms<SMEM_LABEL>(p, qid, i, j, s_score, s_label, s_idx, q_label, do_norm, k, &pl, &ph, &low)) {
            const float fl = static_cast<float>(pl);
            v.x = low == i ? -fl : fl;  // what member i adds; member j adds the opposite
            v.y = static_cast<float>(ph);
            my_sum += -2 * pl;
          }
          row[j] = v;
        }
      }
      __syncthreads();
      for (int r = tid; r < cnt; r += nt) {
        float lam = s_lam[r], hes = s_hes[r];
        // (a) as the j member: outer ranks i < r of this chunk whose inner range contains r, ascending i
        int i_lo = 0, i_hi = min(r, i_end);
        if (a_from_k) {
          if (r < k) i_hi = 0;
        } else if (t == kLambdaGapS) {
          i_lo = max(0, r - k);
          i_hi = r >= k ? min(i_hi, r - k + 1) : 0;
        } else if (t == kLambdaGapX) {
          i_hi = min(i_hi, r - k + 1);
        }
        i_lo = max(i_lo, i0);
        i_hi = min(i_hi, i1);
#pragma unroll 4
        for (int i = i_lo; i < i_hi; ++i) {
          const float2 v = s_pair[static_cast<size_t>(i - i0) * cnt + r];
          lam -= v.x;
          hes += v.y;
        }
        // (b) as the i member: inner ranks j of outer rank r, ascending j
        if (r >= i0 && r < i1) {
          int js, je;
          j_range(t, r, k, cnt, &js, &je);
          const float2* row = s_pair + static_cast<size_t>(r - i0) * cnt;
#pragma unroll 8
          for (int j = js; j < je; ++j) {
            const float2 v = row[j];
            lam += v.x;
            hes += v.y;
          }
        }
        s_lam[r] = lam;
        s_hes[r] = hes;
      }
    }
    __syncthreads();
    for (int r = tid; r < cnt; r += nt) {
      const int d = s_idx[r];
      grad[start + d] = s_lam[r];
      hess[start + d] = s_hes[r];
    }
  } else
  // ---- gather: thread per document (rank r), pairs in the CPU's order
  for (int r = tid; r < cnt; r += nt) {
    float lam = 0.f, hes = 0.f;
    // (a) as the j member: outer ranks i < r whose inner range contains r, ascending i
    {
      int i_lo = 0, i_hi = min(r, i_end);
      if (t == kPrecision || t == kARPK || t == kLambdaGapSp || t == kLambdaGapXp || t == kLambdaGapSpp || t == kLambdaGapXpp) {
        if (r < k) i_hi = 0;
      } else if (t == kLambdaGapS) {
        i_lo = max(0, r - k);
        i_hi = r >= k ? min(i_hi, r - k + 1) : 0;
      } else if (t == kLambdaGapX) {
        i_hi = min(i_hi, r - k + 1);
      }
      for (int i = i_lo; i < i_hi; ++i) {
        eval_pair<SMEM_LABEL>(p, qid, i, r, r, s_score, s_label, s_idx, q_label, do_norm, k, &lam, &hes);
      }
    }
    // (b) as the i member: inner ranks j of outer rank r, ascending j
    if (r < i_end) {
      int js, je;
      j_range(t, r, k, cnt, &js, &je);
      for (int j = js; j < je; ++j) {
        my_sum += eval_pair<SMEM_LABEL>(p, qid, r, j, r, s_score, s_label, s_idx, q_label, do_norm, k, &lam, &hes);
      }
    }
    const int d = s_idx[r];
    grad[start + d] = lam;
    hess[start + d] = hes;
  }
  // ---- sum_lambdas -> norm factor (:516-522), then per-row weights (:98-105)
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) my_sum += __shfl_xor_sync(0xffffffffu, my_sum, o);
  if ((tid & 31) == 0) s_red[tid >> 5] = my_sum;
  __syncthreads();
  double sum_lambdas = 0.0;
  for (int w = 0; w < (nt + 31) / 32; ++w) sum_lambdas += s_red[w];
  const bool scale = p.norm && sum_lambdas > 0;
  const double norm_factor = scale ? log2(1 + sum_lambdas) / sum_lambdas : 1.0;
  if (scale || p.weight != nullptr) {
    for (int r = tid; r < cnt; r += nt) {
      const int d = start + s_idx[r];
      float lam = grad[d], hes = hess[d];
      if (scale) {
        lam = static_cast<float>(lam * norm_factor);
        hes = static_cast<float>(hes * norm_factor);
      }
      if (p.weight != nullptr) {
        lam = static_cast<float>(lam * p.weight[d]);
        hes = static_cast<float>(hes * p.weight[d]);
      }
      grad[d] = lam;
      hess[d] = hes;
    }
  }
}

// ---------------------------------------------------------------------------------------------- host
static double cal_max_dcg_at_k(int k, const float* label, int num_data, const std::vector<double>& gain,
                               const std::vector<double>& discount) {
  // DCGCalculator::CalMaxDCGAtK (dcg_calculator.cpp:56-82)
  double ret = 0.0f;
  std::vector<int> label_cnt(gain.size(), 0);
  for (int i = 0; i < num_data; ++i) ++label_cnt[static_cast<int>(label[i])];
  int top_label = static_cast<int>(gain.size()) - 1;
  if (k > num_data) k = num_data;
  for (int j = 0; j < k; ++j) {
    while (top_label > 0 && label_cnt[top_label] <= 0) top_label -= 1;
    if (top_label < 0) break;
    ret += discount[j] * gain[top_label];
    label_cnt[top_label] -= 1;
  }
  return ret;
}

static double cal_max_bdcg_at_k(int k, const float* label, int num_data, const std::vector<double>& discount) {
  // DCGCalculator::CalMaxBDCGAtK (dcg_calculator.cpp:84-96)
  double ret = 0.0f;
  int relevant = 0;
  for (int i = 0; i < num_data; ++i) {
    if (label[i] > 0) ++relevant;
  }
  if (k > num_data) k = num_data;
  if (k > relevant) k = relevant;
  for (int j = 0; j < k; ++j) ret += discount[j];
  return ret;
}

template <typename T>
static T* upload(const std::vector<T>& v) {
  T* d = nullptr;
  GBL_CUDA(cudaMalloc(&d, sizeof(T) * std::max<size_t>(1, v.size())));
  if (!v.empty()) GBL_CUDA(cudaMemcpy(d, v.data(), sizeof(T) * v.size(), cudaMemcpyHostToDevice));
  return d;
}


void rank_objective_init(Objective* o, const gbl_objective_config* cfg, const float* label, const float* /*weight*/,
                         const int32_t* qb, int32_t num_queries) {
  if (qb == nullptr || num_queries <= 0) fatal("Ranking tasks require query information");
  if (cfg->lambdarank_target < 0 || cfg->lambdarank_target >= kNumRankTargets) fatal("Unknown lambdarank target %d", cfg->lambdarank_target);
  if (!(cfg->sigmoid > 0.0)) fatal("Sigmoid param %f should be greater than zero", cfg->sigmoid);
  if (cfg->lambdarank_truncation_level <= 0) fatal("lambdarank_truncation_level must be positive");
  if (qb[0] != 0 || qb[num_queries] != o->num_data) fatal("query boundaries do not cover the data");
  RankState* rs = new RankState();
  o->rank = rs;
  // label gains: DCGCalculator::DefaultLabelGain (dcg_calculator.cpp:33-43)
  std::vector<double> gain;
  if (cfg->label_gain != nullptr && cfg->num_label_gain > 0) {
    gain.assign(cfg->label_gain, cfg->label_gain + cfg->num_label_gain);
  } else {
    gain.push_back(0.0f);
    for (int i = 1; i < 31; ++i) gain.push_back(static_cast<double>((1 << i) - 1));
  }
  std::vector<double> discount(kMaxPosition);
  for (int i = 0; i < kMaxPosition; ++i) discount[i] = 1.0 / std::log2(2.0 + i);  // dcg_calculator.cpp:50-53
  // DCGCalculator::CheckLabel / CheckMetadata (dcg_calculator.cpp:158-189)
  for (int64_t i = 0; i < o->num_data; ++i) {
    const float l = label[i];
    if (std::fabs(l - static_cast<int>(l)) > 1e-15f) fatal("label should be int type (met %f) for ranking task", l);
    if (l < 0) fatal("Label should be non-negative (met %f) for ranking task", l);
    if (static_cast<size_t>(l) >= gain.size()) fatal("Label %zu is not less than the number of label mappings (%zu)", static_cast<size_t>(l), gain.size());
  }
  std::vector<double> inv_dcg(num_queries), inv_bdcg(num_queries);
  std::vector<int> cls[kNumClasses];
  int max_len[kNumClasses] = {};
  const int k = cfg->lambdarank_truncation_level;
  for (int q = 0; q < num_queries; ++q) {
    const int n = qb[q + 1] - qb[q];
    if (n > kMaxPosition) fatal("Number of rows %d exceeds upper limit of %d for a query", n, kMaxPosition);
    if (n <= 0) fatal("empty query %d", q);
    double v = cal_max_dcg_at_k(k, label + qb[q], n, gain, discount);
    if (v > 0.0) v = 1.0f / v;
    inv_dcg[q] = v;
    double b = cal_max_bdcg_at_k(k, label + qb[q], n, discount);
    if (b > 0.0) b = 1.0f / b;
    inv_bdcg[q] = b;
    int c = 0;
    while (n > kClassLimit[c]) ++c;
    cls[c].push_back(q);
    max_len[c] = std::max(max_len[c], n);
  }
  // ConstructSigmoidTable (:539-552)
  double min_in = -50.0 / cfg->sigmoid / 2;
  double max_in = -min_in;
  const double factor = kSigmoidBins / (max_in - min_in);
  std::vector<double> table(kSigmoidBins);
  for (size_t i = 0; i < static_cast<size_t>(kSigmoidBins); ++i) {
    const double s = i / factor + min_in;
    table[i] = 1.0f / (1.0f + std::exp(s * cfg->sigmoid));
  }
  rs->sigmoid_table = upload(table);
  rs->discount = upload(discount);
  rs->label_gain = upload(gain);
  rs->inv_dcg = upload(inv_dcg);
  rs->inv_bdcg = upload(inv_bdcg);
  std::vector<int> qbv(qb, qb + num_queries + 1);
  rs->query_boundaries = upload(qbv);
  for (int c = 0; c < kNumClasses; ++c) {
    rs->class_count[c] = static_cast<int>(cls[c].size());
    rs->class_queries[c] = upload(cls[c]);
    int pad = 2;
    while (pad < max_len[c]) pad <<= 1;
    rs->class_pad[c] = pad;
  }
  RankParams& p = rs->p;
  p.target = cfg->lambdarank_target;
  p.truncation_level = k;
  p.norm = cfg->lambdarank_norm ? 1 : 0;
  p.sigmoid = cfg->sigmoid;
  p.lambdagap_weight = cfg->lambdagap_weight;
  p.min_sigmoid_input = min_in;
  p.max_sigmoid_input = max_in;
  p.sigmoid_table_idx_factor = factor;
  p.sigmoid_table = rs->sigmoid_table;
  p.discount = rs->discount;
  p.label_gain = rs->label_gain;
  p.num_label_gain = static_cast<int>(gain.size());
  p.inverse_max_dcgs = rs->inv_dcg;
  p.inverse_max_bdcgs = rs->inv_bdcg;
  p.query_boundaries = rs->query_boundaries;
  p.label = o->label;
  p.weight = o->weight;
  o->init_score = 0.0;  // ranking objectives do not boost from average (ObjectiveFunction::BoostFromScore default)
  GBL_CUDA(cudaFuncSetAttribute(lambdarank_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  GBL_CUDA(cudaFuncSetAttribute(lambdarank_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
}

void rank_objective_free(Objective* o) {
  RankState* rs = o->rank;
  if (rs == nullptr) return;
  cudaFree(rs->sigmoid_table);
  cudaFree(rs->discount);
  cudaFree(rs->label_gain);
  cudaFree(rs->inv_dcg);
  cudaFree(rs->inv_bdcg);
  cudaFree(rs->query_boundaries);
  for (int c = 0; c < kNumClasses; ++c) cudaFree(rs->class_queries[c]);
  for (int a = 0; a < 2; ++a) {
    if (rs->aux[a]) cudaStreamDestroy(rs->aux[a]);
    if (rs->join[a]) cudaEventDestroy(rs->join[a]);
  }
  if (rs->fork) cudaEventDestroy(rs->fork);
  delete rs;
  o->rank = nullptr;
}

void rank_objective_get_gradients(Objective* o, const double* score, float* grad, float* hess, cudaStream_t s) {
  RankState* rs = o->rank;
  if (rs->fork == nullptr) {
    GBL_CUDA(cudaEventCreateWithFlags(&rs->fork, cudaEventDisableTiming));
    for (int a = 0; a < 2; ++a) {
      GBL_CUDA(cudaStreamCreateWithFlags(&rs->aux[a], cudaStreamNonBlocking));
      GBL_CUDA(cudaEventCreateWithFlags(&rs->join[a], cudaEventDisableTiming));
    }
  }
  GBL_CUDA(cudaEventRecord(rs->fork, s));
  // longest queries first; the classes go round the three streams so that one class's tail overlaps the next class
  cudaStream_t where[3] = {rs->aux[0], rs->aux[1], s};
  bool used[3] = {false, false, false};
  int turn = 0;
  for (int c = kNumClasses - 1; c >= 0; --c) {
    if (rs->class_count[c] == 0) continue;
    const int w = turn++ % 3;
    cudaStream_t cs = where[w];
    if (cs != s && !used[w]) GBL_CUDA(cudaStreamWaitEvent(cs, rs->fork, 0));
    used[w] = true;
    const int pad = rs->class_pad[c];
    const bool smem_label = pad <= 8192;
    size_t smem = static_cast<size_t>(pad) * (8 + 4 + (smem_label ? 4 : 0)) + 32 * 8;
    if (rs->p.target == kPrecision) {
      if (!smem_label) fatal("lambdarank_target=precision supports queries of up to 8192 documents on this device");
      smem += static_cast<size_t>(pad) * 8;
    }
    // pair-parallel form when a chunk of outer ranks fits shared memory beside the arrays: lam / hes + rows x pad float2
    static const int pair_mode = [] {
      const char* e = std::getenv("GBL_RANK_PAIRS");
      return e == nullptr ? 1 : std::atoi(e);
    }();
    int pair_rows = 0;
    if (pair_mode != 0) {
      const size_t limit = (smem_label ? 200 : 220) * 1024;
      const size_t extra = static_cast<size_t>(pad) * 8;  // lam + hes
      const size_t per_row = static_cast<size_t>(pad) * 8;
      // a CTA's share of the SM's shared memory when about 1200 threads are resident; the longest class takes what fits
      const size_t share = pad <= 1024 ? static_cast<size_t>(kClassThreads[c]) * 192 : limit;
      const size_t cap = rs->p.target <= kPrecision ? std::min(pad, rs->p.truncation_level) : pad;  // is_truncated_i
      if (smem + extra + per_row <= limit) {
        const size_t room = std::min(limit, std::max(share, smem + extra + per_row)) - smem - extra;
        pair_rows = static_cast<int>(std::max<size_t>(1, std::min<size_t>(cap, room / per_row)));
        smem += extra + static_cast<size_t>(pair_rows) * per_row;
      }
    }
    if (smem_label) {
      GBL_LAUNCH(lambdarank_kernel<true>, rs->class_count[c], kClassThreads[c], smem, cs, rs->p, rs->class_queries[c], pad,
                 score, grad, hess, pair_rows);
    } else {
      GBL_LAUNCH(lambdarank_kernel<false>, rs->class_count[c], kClassThreads[c], smem, cs, rs->p, rs->class_queries[c], pad,
                 score, grad, hess, pair_rows);
    }
  }
  for (int w = 0; w < 2; ++w) {
    if (!used[w]) continue;
    GBL_CUDA(cudaEventRecord(rs->join[w], where[w]));
    GBL_CUDA(cudaStreamWaitEvent(s, rs->join[w], 0));
  }
}

}  // namespace gbl
