// objective.cu -- per-row gradient / hessian kernels and the objective C ABI.
//
// Reference: RegressionL2loss::GetGradients   src/objective/regression_objective.hpp:112-127
//            BinaryLogloss::Init/GetGradients src/objective/binary_objective.hpp:58-136, BoostFromScore :139-166
//            MulticlassSoftmax::GetGradients  src/objective/multiclass_objective.hpp:84-133, Common::Softmax
//                                             include/LightGBM/utils/common.h:580-594
//            (lambdarank lives in rank_objective.cu)
// Arithmetic follows the reference: fp64 per row, results cast to score_t = float.
#include <cmath>
#include <vector>

#include "objective.cuh"

namespace gbl {

__global__ void __launch_bounds__(256) l2_gradients_kernel(const double* __restrict__ score,
                                                           const float* __restrict__ label,
                                                           const float* __restrict__ weight, int n,
                                                           float* __restrict__ grad, float* __restrict__ hess) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    if (weight == nullptr) {
      grad[i] = static_cast<float>(score[i] - label[i]);
      hess[i] = 1.0f;
    } else {
      grad[i] = static_cast<float>(static_cast<float>(score[i] - label[i]) * weight[i]);
      hess[i] = weight[i];
    }
  }
}

__global__ void __launch_bounds__(256) binary_gradients_kernel(const double* __restrict__ score,
                                                               const float* __restrict__ label,
                                                               const float* __restrict__ weight, int n, double sigmoid,
                                                               double w_neg, double w_pos, float* __restrict__ grad,
                                                               float* __restrict__ hess) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const int is_pos = label[i] > 0 ? 1 : 0;
    const int lab = is_pos ? 1 : -1;
    const double label_weight = is_pos ? w_pos : w_neg;
    const double response = -lab * sigmoid / (1.0 + exp(lab * sigmoid * score[i]));
    const double abs_response = fabs(response);
    if (weight == nullptr) {
      grad[i] = static_cast<float>(response * label_weight);
      hess[i] = static_cast<float>(abs_response * (sigmoid - abs_response) * label_weight);
    } else {
      grad[i] = static_cast<float>(response * label_weight * weight[i]);
      hess[i] = static_cast<float>(abs_response * (sigmoid - abs_response) * label_weight * weight[i]);
    }
  }
}

// Softmax over the row's K class scores (class-major arrays, entry k * n + i), then p - [label == k] and
// factor * p * (1 - p) with factor = K / (K - 1) (multiclass_objective.hpp:31).  One thread per row; the K scores are
// read three times (maximum, sum of exponentials, results) instead of being held in a per-thread array of unknown size.
__global__ void __launch_bounds__(256) multiclass_gradients_kernel(const double* __restrict__ score,
                                                                   const float* __restrict__ label,
                                                                   const float* __restrict__ weight, int n, int num_class,
                                                                   double factor, float* __restrict__ grad,
                                                                   float* __restrict__ hess) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    double wmax = score[i];
    for (int k = 1; k < num_class; ++k) wmax = fmax(score[static_cast<size_t>(n) * k + i], wmax);
    double wsum = 0.0f;
    for (int k = 0; k < num_class; ++k) wsum += exp(score[static_cast<size_t>(n) * k + i] - wmax);
    const int lab = static_cast<int>(label[i]);
    for (int k = 0; k < num_class; ++k) {
      const size_t idx = static_cast<size_t>(n) * k + i;
      const double p = exp(score[idx] - wmax) / wsum;
      if (weight == nullptr) {
        grad[idx] = lab == k ? static_cast<float>(p - 1.0f) : static_cast<float>(p);
        hess[idx] = static_cast<float>(factor * p * (1.0f - p));
      } else {
        grad[idx] = lab == k ? static_cast<float>((p - 1.0f) * weight[i]) : static_cast<float>(p * weight[i]);
        hess[idx] = static_cast<float>((factor * p * (1.0f - p)) * weight[i]);
      }
    }
  }
}

// BinaryLogloss::Init, binary_objective.hpp:73-101: need_train and the is_unbalance label weights from the numbers of
// positive / negative rows.  With sharded rows the reference sums the two counts over machines first (:75-77); the
// data-parallel host layer does the same and hands the global counts in through gbl_objective_set_label_counts.
static void set_binary_label_weights(Objective* o, long long pos, long long neg) {
  o->need_train = !(pos == 0 || neg == 0);
  o->label_weight[0] = 1.0;
  o->label_weight[1] = 1.0;
  if (o->cfg.is_unbalance && pos > 0 && neg > 0) {
    if (pos > neg) {
      o->label_weight[1] = 1.0;
      o->label_weight[0] = static_cast<double>(pos) / neg;
    } else {
      o->label_weight[1] = static_cast<double>(neg) / pos;
      o->label_weight[0] = 1.0;
    }
  }
  o->label_weight[1] *= o->cfg.scale_pos_weight;
}

static void init_binary(Objective* o, const float* label, const float* weight) {
  // BinaryLogloss::Init, binary_objective.hpp:58-101
  long long pos = 0, neg = 0;
  for (int64_t i = 0; i < o->num_data; ++i) (label[i] > 0 ? pos : neg)++;
  set_binary_label_weights(o, pos, neg);
  // BoostFromScore, :139-166
  double suml = 0.0, sumw = 0.0;
  if (weight != nullptr) {
    for (int64_t i = 0; i < o->num_data; ++i) {
      suml += (label[i] > 0) * weight[i];
      sumw += weight[i];
    }
  } else {
    sumw = static_cast<double>(o->num_data);
    suml = static_cast<double>(pos);
  }
  double pavg = suml / sumw;
  pavg = std::min(pavg, 1.0 - GBL_KEPS);
  pavg = std::max<double>(pavg, GBL_KEPS);
  o->init_score = std::log(pavg / (1.0f - pavg)) / o->cfg.sigmoid;
}

static void init_l2(Objective* o, const float* label, const float* weight) {
  // RegressionL2loss::BoostFromScore, regression_objective.hpp:160-177
  double suml = 0.0, sumw = 0.0;
  if (weight != nullptr) {
    for (int64_t i = 0; i < o->num_data; ++i) {
      suml += static_cast<double>(label[i]) * weight[i];
      sumw += weight[i];
    }
  } else {
    sumw = static_cast<double>(o->num_data);
    for (int64_t i = 0; i < o->num_data; ++i) suml += label[i];
  }
  o->init_score = suml / sumw;
}

static Objective* objective_create(const gbl_objective_config* cfg, int64_t num_data, const float* label,
                                   const float* weight, const int32_t* qb, int32_t num_queries, int device) {
  if (cfg == nullptr || label == nullptr) fatal("gbl_objective_create: null argument");
  if (num_data <= 0 || num_data >= (1LL << 31)) fatal("objective: num_data out of range");
  if (cfg->kind == GBL_OBJ_BINARY && !(cfg->sigmoid > 0.0)) fatal("Sigmoid parameter %f should be greater than zero", cfg->sigmoid);
  if (cfg->kind == GBL_OBJ_BINARY && cfg->is_unbalance && std::fabs(cfg->scale_pos_weight - 1.0f) > 1e-6)
    fatal("Cannot set is_unbalance and scale_pos_weight at the same time");
  DeviceGuard guard(device);
  Objective* o = new Objective();
  o->cfg = *cfg;
  o->cfg.label_gain = nullptr;
  o->device = device;
  o->num_data = num_data;
  try {
    GBL_CUDA(cudaMalloc(&o->label, sizeof(float) * num_data));
    GBL_CUDA(cudaMemcpy(o->label, label, sizeof(float) * num_data, cudaMemcpyHostToDevice));
    if (weight != nullptr) {
      GBL_CUDA(cudaMalloc(&o->weight, sizeof(float) * num_data));
      GBL_CUDA(cudaMemcpy(o->weight, weight, sizeof(float) * num_data, cudaMemcpyHostToDevice));
    }
    switch (cfg->kind) {
      case GBL_OBJ_REGRESSION_L2:
        init_l2(o, label, weight);
        break;
      case GBL_OBJ_BINARY:
        init_binary(o, label, weight);
        break;
      case GBL_OBJ_LAMBDARANK:
        rank_objective_init(o, cfg, label, weight, qb, num_queries);
        break;
      case GBL_OBJ_MULTICLASS:
        if (cfg->num_class < 2) fatal("objective: num_class=%d (multiclass needs at least two classes)", cfg->num_class);
        for (int64_t i = 0; i < num_data; ++i) {  // MulticlassSoftmax::Init, multiclass_objective.hpp:62-66
          const int l = static_cast<int>(label[i]);
          if (l < 0 || l >= cfg->num_class) fatal("Label must be in [0, %d), but found %d in label", cfg->num_class, l);
        }
        break;
      default:
        fatal("objective: unknown kind %d", cfg->kind);
    }
  } catch (...) {
    objective_free(o);
    throw;
  }
  return o;
}

void objective_free(Objective* o) {
  if (o == nullptr) return;
  DeviceGuard guard(o->device);
  cudaFree(o->label);
  cudaFree(o->weight);
  rank_objective_free(o);
  delete o;
}

static void objective_get_gradients(Objective* o, const double* score, float* grad, float* hess, cudaStream_t s) {
  if (score == nullptr || grad == nullptr || hess == nullptr) fatal("gbl_objective_get_gradients: null argument");
  DeviceGuard guard(o->device);
  const int n = static_cast<int>(o->num_data);
  const int grid = std::min((n + 255) / 256, kNumSM * 8);
  switch (o->cfg.kind) {
    case GBL_OBJ_REGRESSION_L2:
      GBL_LAUNCH(l2_gradients_kernel, grid, 256, 0, s, score, o->label, o->weight, n, grad, hess);
      break;
    case GBL_OBJ_BINARY:
      if (!o->need_train) return;  // binary_objective.hpp:106-108
      GBL_LAUNCH(binary_gradients_kernel, grid, 256, 0, s, score, o->label, o->weight, n, o->cfg.sigmoid,
                 o->label_weight[0], o->label_weight[1], grad, hess);
      break;
    case GBL_OBJ_LAMBDARANK:
      rank_objective_get_gradients(o, score, grad, hess, s);
      break;
    case GBL_OBJ_MULTICLASS:
      GBL_LAUNCH(multiclass_gradients_kernel, grid, 256, 0, s, score, o->label, o->weight, n, o->cfg.num_class,
                 static_cast<double>(o->cfg.num_class) / (o->cfg.num_class - 1.0f), grad, hess);
      break;
    default:
      fatal("objective: unknown kind");
  }
}

}  // namespace gbl

using namespace gbl;

extern "C" {

int gbl_objective_create(const gbl_objective_config* config, int64_t num_data, const float* label_host,
                         const float* weight_host, const int32_t* query_boundaries_host, int32_t num_queries,
                         int32_t device_id, gbl_objective** out) {
  GBL_API_BEGIN();
  if (out == nullptr) fatal("gbl_objective_create: null argument");
  *out = reinterpret_cast<gbl_objective*>(
      objective_create(config, num_data, label_host, weight_host, query_boundaries_host, num_queries, device_id));
  GBL_API_END();
}

int gbl_objective_set_label_counts(gbl_objective* obj, int64_t num_positive, int64_t num_negative) {
  GBL_API_BEGIN();
  Objective* o = reinterpret_cast<Objective*>(obj);
  if (o == nullptr) fatal("gbl_objective_set_label_counts: null objective");
  if (num_positive < 0 || num_negative < 0) fatal("gbl_objective_set_label_counts: negative count");
  if (o->cfg.kind == GBL_OBJ_BINARY) set_binary_label_weights(o, num_positive, num_negative);
  GBL_API_END();
}

int gbl_objective_free(gbl_objective* obj) {
  GBL_API_BEGIN();
  objective_free(reinterpret_cast<Objective*>(obj));
  GBL_API_END();
}

int gbl_objective_get_gradients(gbl_objective* obj, const double* score_dev, float* gradients_dev, float* hessians_dev,
                                void* stream) {
  GBL_API_BEGIN();
  if (obj == nullptr) fatal("gbl_objective_get_gradients: null objective");
  objective_get_gradients(reinterpret_cast<Objective*>(obj), score_dev, gradients_dev, hessians_dev,
                          reinterpret_cast<cudaStream_t>(stream));
  GBL_API_END();
}

int gbl_objective_boost_from_score(gbl_objective* obj, double* out_init_score) {
  GBL_API_BEGIN();
  if (obj == nullptr || out_init_score == nullptr) fatal("gbl_objective_boost_from_score: null argument");
  *out_init_score = reinterpret_cast<Objective*>(obj)->init_score;
  GBL_API_END();
}

int gbl_objective_is_constant_hessian(gbl_objective* obj, int32_t* out) {
  GBL_API_BEGIN();
  if (obj == nullptr || out == nullptr) fatal("gbl_objective_is_constant_hessian: null argument");
  const Objective* o = reinterpret_cast<Objective*>(obj);
  *out = (o->cfg.kind == GBL_OBJ_REGRESSION_L2 && o->weight == nullptr) ? 1 : 0;  // regression_objective.hpp:152-158
  GBL_API_END();
}

}  // extern "C"
