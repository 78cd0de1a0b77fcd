// metric.cu -- evaluation metrics on device-resident scores: with the learner and the
// score vectors in HBM, a per-iteration eval would otherwise copy every score to the host each iteration.
//
// Reference: RegressionMetric<L2/RMSE/L1>::Eval   src/metric/regression_metric.hpp:58-95,119-183
//            BinaryMetric<logloss/error>::Eval    src/metric/binary_metric.hpp:60-97,104-150
//            AUCMetric::Eval                      src/metric/binary_metric.hpp:192-247
//            NDCGMetric::Init/Eval                src/metric/rank_metric.hpp:33-75,86-150
//            DCGCalculator::CalMaxDCG / CalDCG    src/metric/dcg_calculator.cpp:96-156
//
// What is reproduced exactly and what within rounding:
//   * AUC (unweighted) is integer arithmetic in disguise: every addend cur_neg * (cur_pos * 0.5 + sum_pos) is a
//     half-integer below 2^53, so the reference's double sums are exact and this file computes the same number
//     from exact 64-bit integer sums: 2 * accum = sum over negative rows of (#positives with a larger score +
//     #positives with a score >= the row's score).  Rows are sorted by score (cub::DeviceRadixSort, the one library
//     routine here -- a key sort is plumbing, not the path under test), tie groups found by binary search.
//   * NDCG: the per-query DCG adds label_gain * discount in rank order like CalDCG does (bit-identical per query; the
//     rank of a document is counted, not sorted: #docs with a larger score, or an equal score and a smaller index
//     = std::stable_sort's order); the sum over queries is a fixed-order tree, the reference's is per-thread chunks,
//     so the mean agrees to ~1e-15 relative.
//   * point-wise losses: per-row value as the reference computes it (fp64), fixed-order tree sum.
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

#include <cmath>
#include <vector>

#include "common.cuh"

namespace gbl {

constexpr int kMaxPosition = 10000;   // DCGCalculator::kMaxPosition (include/LightGBM/metric.h)
constexpr int kMaxEvalAt = 4096;      // positions kept per query in shared memory
constexpr int kRedThreads = 256;
constexpr int kRedBlocks = kNumSM * 4;

struct Metric {
  gbl_metric_config cfg{};
  int device = 0;
  int64_t n = 0;
  float* label = nullptr;
  float* weight = nullptr;
  double sum_weights = 0.0;
  // reductions
  double* partial = nullptr;  // [kRedBlocks] or [num_queries * nk]
  double* result = nullptr;   // device, [max(1, nk)]
  double* result_pinned = nullptr;
  // ndcg
  int nq = 0, nk = 0, max_k = 0;
  int* qb = nullptr;
  int* eval_at = nullptr;
  double* inv_max_dcg = nullptr;  // [nq][nk]
  double* label_gain = nullptr;
  double* discount = nullptr;
  std::vector<int> eval_at_host;
  // auc
  unsigned long long* keys_in = nullptr;
  unsigned long long* keys_out = nullptr;
  unsigned* pos_in = nullptr;
  unsigned* pos_out = nullptr;
  unsigned* pos_prefix = nullptr;  // [n + 1]
  unsigned long long* accum2 = nullptr;
  void* cub_temp = nullptr;
  size_t cub_temp_bytes = 0;
};

// ------------------------------------------------------------------------------------------------ reductions
// sum of v over the CTA in a fixed order (warp shuffles, then the warp totals); valid in thread 0
__device__ inline double block_sum(double v, double* s_warp) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) s_warp[warp] = v;
  __syncthreads();
  double t = 0.0;
  if (threadIdx.x == 0) {
    for (int w = 0; w < (blockDim.x + 31) / 32; ++w) t += s_warp[w];
  }
  __syncthreads();
  return t;
}

// out[c] = sum_r in[r * ncol + c], one CTA per column, fixed order
__global__ void __launch_bounds__(kRedThreads) column_sum_kernel(const double* __restrict__ in, int nrow, int ncol,
                                                                 double* __restrict__ out) {
  __shared__ double s_warp[kRedThreads / 32];
  const int c = blockIdx.x;
  double v = 0.0;
  for (int r = threadIdx.x; r < nrow; r += kRedThreads) v += in[static_cast<size_t>(r) * ncol + c];
  const double t = block_sum(v, s_warp);
  if (threadIdx.x == 0) out[c] = t;
}

// ------------------------------------------------------------------------------------------------ point-wise
// BinaryLoglossMetric::LossOnPoint (binary_metric.hpp:108-119); kEpsilon is the float 1e-15f
__device__ inline double logloss_on_point(float label, double prob) {
  const double eps = static_cast<double>(1e-15f);
  if (label <= 0) {
    if (1.0 - prob > eps) return -log(1.0 - prob);
  } else {
    if (prob > eps) return -log(prob);
  }
  return -log(eps);
}

__global__ void __launch_bounds__(kRedThreads) pointwise_kernel(int kind, const double* __restrict__ score,
                                                                const float* __restrict__ label,
                                                                const float* __restrict__ weight, int64_t n,
                                                                double sigmoid, double* __restrict__ partial) {
  __shared__ double s_warp[kRedThreads / 32];
  double v = 0.0;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * kRedThreads + threadIdx.x; i < n;
       i += static_cast<int64_t>(gridDim.x) * kRedThreads) {
    const double s = score[i];
    const float y = label[i];
    double loss;
    switch (kind) {
      case GBL_METRIC_L2:
      case GBL_METRIC_RMSE:
        loss = (s - y) * (s - y);
        break;
      case GBL_METRIC_L1:
        loss = fabs(s - y);
        break;
      case GBL_METRIC_BINARY_LOGLOSS: {
        // BinaryLogloss::ConvertOutput (binary_objective.hpp:167-169) when an objective is attached
        const double p = sigmoid > 0.0 ? 1.0 / (1.0 + exp(-sigmoid * s)) : s;
        loss = logloss_on_point(y, p);
        break;
      }
      default: {  // GBL_METRIC_BINARY_ERROR (binary_metric.hpp:138-144)
        const double p = sigmoid > 0.0 ? 1.0 / (1.0 + exp(-sigmoid * s)) : s;
        loss = p <= 0.5 ? (y > 0 ? 1.0 : 0.0) : (y <= 0 ? 1.0 : 0.0);
        break;
      }
    }
    v += weight != nullptr ? loss * weight[i] : loss;
  }
  const double t = block_sum(v, s_warp);
  if (threadIdx.x == 0) partial[blockIdx.x] = t;
}

// ------------------------------------------------------------------------------------------------ AUC
// order-preserving key of a double; -0.0 joins +0.0 because the reference groups ties with operator!=
__device__ inline unsigned long long score_key(double s) {
  if (s == 0.0) s = 0.0;
  const unsigned long long b = static_cast<unsigned long long>(__double_as_longlong(s));
  return (b >> 63) ? ~b : (b | 0x8000000000000000ull);
}

__global__ void __launch_bounds__(256) auc_keys_kernel(const double* __restrict__ score, const float* __restrict__ label,
                                                       int64_t n, unsigned long long* __restrict__ keys,
                                                       unsigned* __restrict__ pos) {
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * 256 + threadIdx.x; i <= n; i += static_cast<int64_t>(gridDim.x) * 256) {
    if (i < n) {
      keys[i] = score_key(score[i]);
      pos[i] = label[i] > 0 ? 1u : 0u;
    } else {
      pos[i] = 0u;  // so that the exclusive scan over n + 1 entries ends with the total
    }
  }
}

// keys sorted descending, pos = label > 0 in that order, prefix[i] = positives in [0, i), prefix[n] = all positives.
// For the negative row at i with tie group [a, b):  AUCMetric adds cur_neg * (cur_pos * 0.5 + sum_pos) per group,
// i.e. (prefix[b] - prefix[a]) * 0.5 + prefix[a] per negative row of the group = (prefix[a] + prefix[b]) / 2.
__global__ void __launch_bounds__(256) auc_accum_kernel(const unsigned long long* __restrict__ keys,
                                                        const unsigned* __restrict__ pos,
                                                        const unsigned* __restrict__ prefix, int64_t n,
                                                        unsigned long long* __restrict__ accum2) {
  unsigned long long v = 0;
  for (int64_t i = static_cast<int64_t>(blockIdx.x) * 256 + threadIdx.x; i < n; i += static_cast<int64_t>(gridDim.x) * 256) {
    if (pos[i] != 0u) continue;
    const unsigned long long k = keys[i];
    int64_t a = i, b = i + 1;
    if (i > 0 && keys[i - 1] == k) {  // first index with key == k (keys descend: everything before it is larger)
      int64_t lo = 0, hi = i - 1;     // answer in [lo, hi]
      while (lo < hi) {
        const int64_t mid = (lo + hi) >> 1;
        if (keys[mid] > k) {
          lo = mid + 1;
        } else {
          hi = mid;
        }
      }
      a = lo;
    }
    if (i + 1 < n && keys[i + 1] == k) {  // one past the last index with key == k
      int64_t lo = i + 2, hi = n;         // answer in [lo, hi]
      while (lo < hi) {
        const int64_t mid = (lo + hi) >> 1;
        if (keys[mid] == k) {
          lo = mid + 1;
        } else {
          hi = mid;
        }
      }
      b = lo;
    }
    v += static_cast<unsigned long long>(prefix[a]) + static_cast<unsigned long long>(prefix[b]);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  if ((threadIdx.x & 31) == 0 && v != 0) atomicAdd(accum2, v);  // exact integers: any order
}

// ------------------------------------------------------------------------------------------------ NDCG
// One CTA per query.  out[q][j] = DCG@eval_at[j] * inverse_max_dcg[q][j], or 1 for a query without relevant
// documents (rank_metric.hpp:104-120).
__global__ void __launch_bounds__(256) ndcg_kernel(const double* __restrict__ score, const float* __restrict__ label,
                                                   const int* __restrict__ qb, int nq, const int* __restrict__ eval_at,
                                                   int nk, int max_k, const double* __restrict__ inv_max_dcg,
                                                   const double* __restrict__ label_gain,
                                                   const double* __restrict__ discount, double* __restrict__ out) {
  extern __shared__ int s_label[];  // label of the document at each of the first max_k ranks
  for (int q = blockIdx.x; q < nq; q += gridDim.x) {
    const int begin = qb[q], cnt = qb[q + 1] - qb[q];
    double* o = out + static_cast<size_t>(q) * nk;
    const double* inv = inv_max_dcg + static_cast<size_t>(q) * nk;
    if (inv[0] <= 0.0) {
      for (int j = threadIdx.x; j < nk; j += blockDim.x) o[j] = 1.0;
      continue;
    }
    const double* s = score + begin;
    const int top = min(max_k, cnt);
    for (int i = threadIdx.x; i < cnt; i += blockDim.x) {
      const double si = s[i];
      int rank = 0;
      for (int j = 0; j < cnt; ++j) {
        const double sj = s[j];
        rank += (sj > si || (sj == si && j < i)) ? 1 : 0;
      }
      if (rank < top) s_label[rank] = static_cast<int>(label[begin + i]);
    }
    __syncthreads();
    if (threadIdx.x == 0) {  // DCGCalculator::CalDCG's accumulation, in rank order
      double cur = 0.0;
      int cur_left = 0;
      for (int j = 0; j < nk; ++j) {
        const int cur_k = min(eval_at[j], cnt);
        for (int r = cur_left; r < cur_k; ++r) cur += label_gain[s_label[r]] * discount[r];
        o[j] = cur * inv[j];
        cur_left = cur_k;
      }
    }
    __syncthreads();
  }
}

// ------------------------------------------------------------------------------------------------ host side
static void metric_free(Metric* m) {
  if (m == nullptr) return;
  DeviceGuard guard(m->device);
  cudaFree(m->label);
  cudaFree(m->weight);
  cudaFree(m->partial);
  cudaFree(m->result);
  if (m->result_pinned) cudaFreeHost(m->result_pinned);
  cudaFree(m->qb);
  cudaFree(m->eval_at);
  cudaFree(m->inv_max_dcg);
  cudaFree(m->label_gain);
  cudaFree(m->discount);
  cudaFree(m->keys_in);
  cudaFree(m->keys_out);
  cudaFree(m->pos_in);
  cudaFree(m->pos_out);
  cudaFree(m->pos_prefix);
  cudaFree(m->accum2);
  cudaFree(m->cub_temp);
  delete m;
}

template <class T>
static T* to_device(const T* host, size_t count) {
  T* p = nullptr;
  GBL_CUDA(cudaMalloc(&p, sizeof(T) * std::max<size_t>(count, 1)));
  if (count > 0) GBL_CUDA(cudaMemcpy(p, host, sizeof(T) * count, cudaMemcpyHostToDevice));
  return p;
}

static void init_ndcg(Metric* m, const gbl_metric_config* cfg, const float* label, const int32_t* qb, int32_t nq) {
  if (qb == nullptr || nq <= 0) fatal("The NDCG metric requires query information");  // rank_metric.hpp:47-49
  if (cfg->num_eval_at <= 0 || cfg->eval_at == nullptr) fatal("metric ndcg: eval_at is empty");
  if (cfg->num_label_gain <= 0 || cfg->label_gain == nullptr) fatal("metric ndcg: label_gain is empty");
  m->nq = nq;
  m->nk = cfg->num_eval_at;
  m->eval_at_host.assign(cfg->eval_at, cfg->eval_at + cfg->num_eval_at);
  for (int k : m->eval_at_host) {
    if (k <= 0) fatal("metric ndcg: eval_at positions must be positive");
    if (k > kMaxEvalAt) fatal("metric ndcg: eval_at=%d exceeds the %d positions this library keeps per query", k, kMaxEvalAt);
    m->max_k = std::max(m->max_k, k);
  }
  if (qb[nq] != m->n) fatal("metric ndcg: query boundaries end at %d, num_data is %lld", qb[nq], (long long)m->n);
  std::vector<double> gain(cfg->label_gain, cfg->label_gain + cfg->num_label_gain);
  std::vector<double> discount(kMaxPosition);
  for (int i = 0; i < kMaxPosition; ++i) discount[i] = 1.0 / std::log2(2.0 + i);  // dcg_calculator.cpp:50-53
  // DCGCalculator::CheckLabel (dcg_calculator.cpp:172-186) + CalMaxDCG (:96-128) per query + the inversion of
  // NDCGMetric::Init (rank_metric.hpp:62-73)
  std::vector<double> inv(static_cast<size_t>(nq) * m->nk);
  std::vector<int> label_cnt(gain.size());
  for (int q = 0; q < nq; ++q) {
    const int begin = qb[q], cnt = qb[q + 1] - qb[q];
    if (cnt < 0) fatal("metric ndcg: query boundaries must not decrease");
    if (cnt > kMaxPosition) fatal("Number of rows %d exceeds upper limit of %d for a query", cnt, kMaxPosition);
    std::fill(label_cnt.begin(), label_cnt.end(), 0);
    for (int i = 0; i < cnt; ++i) {
      const float l = label[begin + i];
      if (std::fabs(l - static_cast<int>(l)) > 1e-15f) fatal("label should be int type (met %f) for ranking task", l);
      if (l < 0 || static_cast<size_t>(l) >= gain.size()) fatal("Label %d is not less than the number of label mappings (%zu)", static_cast<int>(l), gain.size());
      ++label_cnt[static_cast<int>(l)];
    }
    double cur = 0.0;
    int cur_left = 0;
    int top_label = static_cast<int>(gain.size()) - 1;
    for (int j = 0; j < m->nk; ++j) {
      const int cur_k = std::min(m->eval_at_host[j], cnt);
      for (int r = cur_left; r < cur_k; ++r) {
        while (top_label > 0 && label_cnt[top_label] <= 0) top_label -= 1;
        if (top_label < 0) break;
        cur += discount[r] * gain[top_label];
        label_cnt[top_label] -= 1;
      }
      inv[static_cast<size_t>(q) * m->nk + j] = cur > 0.0 ? 1.0 / cur : -1.0;
      cur_left = cur_k;
    }
  }
  m->qb = to_device(qb, static_cast<size_t>(nq) + 1);
  m->eval_at = to_device(m->eval_at_host.data(), m->eval_at_host.size());
  m->inv_max_dcg = to_device(inv.data(), inv.size());
  m->label_gain = to_device(gain.data(), gain.size());
  m->discount = to_device(discount.data(), discount.size());
  GBL_CUDA(cudaMalloc(&m->partial, sizeof(double) * inv.size()));
}

static void init_auc(Metric* m) {
  if (m->weight != nullptr) fatal("metric auc: weighted AUC is not supported by this library");
  const size_t n = static_cast<size_t>(m->n);
  GBL_CUDA(cudaMalloc(&m->keys_in, sizeof(unsigned long long) * n));
  GBL_CUDA(cudaMalloc(&m->keys_out, sizeof(unsigned long long) * n));
  GBL_CUDA(cudaMalloc(&m->pos_in, sizeof(unsigned) * (n + 1)));
  GBL_CUDA(cudaMalloc(&m->pos_out, sizeof(unsigned) * (n + 1)));
  GBL_CUDA(cudaMalloc(&m->pos_prefix, sizeof(unsigned) * (n + 1)));
  GBL_CUDA(cudaMalloc(&m->accum2, sizeof(unsigned long long)));
  size_t sort_bytes = 0, scan_bytes = 0;
  GBL_CUDA(cub::DeviceRadixSort::SortPairsDescending(nullptr, sort_bytes, m->keys_in, m->keys_out, m->pos_in, m->pos_out,
                                                     static_cast<int>(m->n)));
  GBL_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, m->pos_out, m->pos_prefix, static_cast<int>(m->n + 1)));
  m->cub_temp_bytes = std::max(sort_bytes, scan_bytes);
  GBL_CUDA(cudaMalloc(&m->cub_temp, m->cub_temp_bytes));
}

static Metric* metric_create(const gbl_metric_config* cfg, int64_t num_data, const float* label, const float* weight,
                             const int32_t* qb, int32_t nq, int device) {
  if (cfg == nullptr || label == nullptr) fatal("gbl_metric_create: null argument");
  if (num_data <= 0 || num_data >= (1LL << 31) - 1) fatal("gbl_metric_create: num_data=%lld out of range", (long long)num_data);
  if (cfg->kind < GBL_METRIC_L2 || cfg->kind > GBL_METRIC_NDCG) fatal("gbl_metric_create: unknown metric kind %d", cfg->kind);
  int count = 0;
  GBL_CUDA(cudaGetDeviceCount(&count));
  if (device < 0 || device >= count) fatal("gbl_metric_create: device %d not available (%d devices)", device, count);
  DeviceGuard guard(device);
  Metric* m = new Metric();
  try {
    m->cfg = *cfg;
    m->cfg.eval_at = nullptr;
    m->cfg.label_gain = nullptr;
    m->device = device;
    m->n = num_data;
    m->label = to_device(label, static_cast<size_t>(num_data));
    if (weight != nullptr) {
      m->weight = to_device(weight, static_cast<size_t>(num_data));
      double sw = 0.0;  // `sum_weights_ += weights_[i]`, sequential (binary_metric.hpp:44-49)
      for (int64_t i = 0; i < num_data; ++i) sw += weight[i];
      m->sum_weights = sw;
    } else {
      m->sum_weights = static_cast<double>(num_data);
    }
    const int nres = cfg->kind == GBL_METRIC_NDCG ? cfg->num_eval_at : 1;
    GBL_CUDA(cudaMalloc(&m->result, sizeof(double) * std::max(nres, 1)));
    GBL_CUDA(cudaMallocHost(&m->result_pinned, sizeof(double) * std::max(nres, 2)));
    if (cfg->kind == GBL_METRIC_NDCG) {
      init_ndcg(m, cfg, label, qb, nq);
    } else if (cfg->kind == GBL_METRIC_AUC) {
      init_auc(m);
    } else {
      GBL_CUDA(cudaMalloc(&m->partial, sizeof(double) * kRedBlocks));
    }
  } catch (...) {
    metric_free(m);
    throw;
  }
  return m;
}

static int metric_num_outputs(const Metric* m) { return m->cfg.kind == GBL_METRIC_NDCG ? m->nk : 1; }

static void metric_eval(Metric* m, const double* score, cudaStream_t s, double* out) {
  if (score == nullptr || out == nullptr) fatal("gbl_metric_eval: null argument");
  DeviceGuard guard(m->device);
  const int kind = m->cfg.kind;
  if (kind == GBL_METRIC_NDCG) {
    const int grid = std::min(m->nq, kNumSM * 8);
    GBL_LAUNCH(ndcg_kernel, grid, 256, sizeof(int) * m->max_k, s, score, m->label, m->qb, m->nq, m->eval_at, m->nk, m->max_k,
               m->inv_max_dcg, m->label_gain, m->discount, m->partial);
    GBL_LAUNCH(column_sum_kernel, m->nk, kRedThreads, 0, s, m->partial, m->nq, m->nk, m->result);
    GBL_CUDA(cudaMemcpyAsync(m->result_pinned, m->result, sizeof(double) * m->nk, cudaMemcpyDeviceToHost, s));
    GBL_CUDA(cudaStreamSynchronize(s));
    // rank_metric.hpp:152-160: divided by the sum of query weights (= the number of queries without weights)
    for (int j = 0; j < m->nk; ++j) out[j] = m->result_pinned[j] / static_cast<double>(m->nq);
    return;
  }
  if (kind == GBL_METRIC_AUC) {
    const int n = static_cast<int>(m->n);
    const int grid = std::min((n + 256) / 256, kNumSM * 8);
    GBL_LAUNCH(auc_keys_kernel, grid, 256, 0, s, score, m->label, m->n, m->keys_in, m->pos_in);
    GBL_CUDA(cub::DeviceRadixSort::SortPairsDescending(m->cub_temp, m->cub_temp_bytes, m->keys_in, m->keys_out, m->pos_in,
                                                       m->pos_out, n, 0, 64, s));
    GBL_CUDA(cudaMemsetAsync(m->pos_out + n, 0, sizeof(unsigned), s));
    GBL_CUDA(cub::DeviceScan::ExclusiveSum(m->cub_temp, m->cub_temp_bytes, m->pos_out, m->pos_prefix, n + 1, s));
    GBL_CUDA(cudaMemsetAsync(m->accum2, 0, sizeof(unsigned long long), s));
    GBL_LAUNCH(auc_accum_kernel, grid, 256, 0, s, m->keys_out, m->pos_out, m->pos_prefix, m->n, m->accum2);
    unsigned long long* hp = reinterpret_cast<unsigned long long*>(m->result_pinned);
    GBL_CUDA(cudaMemcpyAsync(hp, m->accum2, sizeof(unsigned long long), cudaMemcpyDeviceToHost, s));
    unsigned* total = reinterpret_cast<unsigned*>(hp + 1);
    GBL_CUDA(cudaMemcpyAsync(total, m->pos_prefix + n, sizeof(unsigned), cudaMemcpyDeviceToHost, s));
    GBL_CUDA(cudaStreamSynchronize(s));
    const double accum = static_cast<double>(hp[0]) * 0.5;  // exact: accum2 < 2^53
    const double sum_pos = static_cast<double>(*total);
    double auc = 1.0;  // binary_metric.hpp:241-245
    if (sum_pos > 0.0 && sum_pos != m->sum_weights) auc = accum / (sum_pos * (m->sum_weights - sum_pos));
    out[0] = auc;
    return;
  }
  GBL_LAUNCH(pointwise_kernel, kRedBlocks, kRedThreads, 0, s, kind, score, m->label, m->weight, m->n, m->cfg.sigmoid,
             m->partial);
  GBL_LAUNCH(column_sum_kernel, 1, kRedThreads, 0, s, m->partial, kRedBlocks, 1, m->result);
  GBL_CUDA(cudaMemcpyAsync(m->result_pinned, m->result, sizeof(double), cudaMemcpyDeviceToHost, s));
  GBL_CUDA(cudaStreamSynchronize(s));
  const double mean = m->result_pinned[0] / m->sum_weights;
  out[0] = kind == GBL_METRIC_RMSE ? std::sqrt(mean) : mean;  // RMSEMetric::AverageLoss (regression_metric.hpp:127-130)
}

}  // namespace gbl

using namespace gbl;

extern "C" {

int gbl_metric_create(const gbl_metric_config* config, int64_t num_data, const float* label_host, const float* weight_host,
                      const int32_t* query_boundaries_host, int32_t num_queries, int32_t device_id, gbl_metric** out) {
  GBL_API_BEGIN();
  if (out == nullptr) fatal("gbl_metric_create: null argument");
  *out = reinterpret_cast<gbl_metric*>(
      metric_create(config, num_data, label_host, weight_host, query_boundaries_host, num_queries, device_id));
  GBL_API_END();
}

int gbl_metric_free(gbl_metric* metric) {
  GBL_API_BEGIN();
  metric_free(reinterpret_cast<Metric*>(metric));
  GBL_API_END();
}

int gbl_metric_num_outputs(const gbl_metric* metric, int32_t* out) {
  GBL_API_BEGIN();
  if (metric == nullptr || out == nullptr) fatal("gbl_metric_num_outputs: null argument");
  *out = metric_num_outputs(reinterpret_cast<const Metric*>(metric));
  GBL_API_END();
}

int gbl_metric_eval(gbl_metric* metric, const double* score_dev, void* stream, double* out_host) {
  GBL_API_BEGIN();
  if (metric == nullptr) fatal("gbl_metric_eval: null metric");
  metric_eval(reinterpret_cast<Metric*>(metric), score_dev, reinterpret_cast<cudaStream_t>(stream), out_host);
  GBL_API_END();
}

}  // extern "C"
