// objective.cuh -- objective state shared by objective.cu and rank_objective.cu
#pragma once
#include "common.cuh"

namespace gbl {

struct RankState;

struct Objective {
  gbl_objective_config cfg{};
  int device = 0;
  int64_t num_data = 0;
  float* label = nullptr;   // device
  float* weight = nullptr;  // device or nullptr
  double init_score = 0.0;
  bool need_train = true;
  double label_weight[2] = {1.0, 1.0};
  RankState* rank = nullptr;
};

void objective_free(Objective* o);
// rank_objective.cu
void rank_objective_init(Objective* o, const gbl_objective_config* cfg, const float* label_host,
                         const float* weight_host, const int32_t* query_boundaries_host, int32_t num_queries);
void rank_objective_free(Objective* o);
void rank_objective_get_gradients(Objective* o, const double* score, float* grad, float* hess, cudaStream_t s);

}  // namespace gbl
