// rank_objective.cu -- placeholder until the lambdarank kernels land (fails loudly, never falls back).
#include "objective.cuh"

namespace gbl {
void rank_objective_init(Objective*, const gbl_objective_config*, const float*, const float*, const int32_t*, int32_t) {
  fatal("lambdarank objective: not built yet");
}
void rank_objective_free(Objective*) {}
void rank_objective_get_gradients(Objective*, const double*, float*, float*, cudaStream_t) {
  fatal("lambdarank objective: not built yet");
}
}  // namespace gbl
