// integration/b200_factory.cpp -- TreeLearner::CreateTreeLearner for the library built by integration/Makefile.
//
// Takes the place of the reference's factory translation unit (src/treelearner/tree_learner.cpp:15-57), which is left
// out of that build.  The CPU and OpenCL ("gpu") learners are created exactly as there; the difference is the
// `device_type == "cuda"` case, which the reference serves with CUDASingleGPUTreeLearner (and refuses for num_gpu != 1 or
// any parallel learner type, :44-55).  Here it returns the B200 learner:
//   tree_learner=serial, num_gpu=1   one GPU (config->gpu_device_id)
//   num_gpu=N > 1                    rows sharded over GPUs 0..N-1 inside this one learner (histogram sums over NVLink)
//   tree_learner=data, num_gpu=1     kept for the reference's multi-machine set-up: refused, one Booster drives all GPUs here
#include <LightGBM/tree_learner.h>
#include <LightGBM/utils/log.h>

#include <string>

#include "b200_tree_learner.h"
#include "gpu_tree_learner.h"
#include "linear_tree_learner.h"
#include "parallel_tree_learner.h"
#include "serial_tree_learner.h"

namespace LightGBM {

namespace {
// one learner family (SerialTreeLearner on the CPU, GPUTreeLearner for OpenCL) behind the four learner types
template <typename BASE>
TreeLearner* MakeHostLearner(const std::string& type, const Config* config) {
  if (type == "serial") {
    if (config->linear_tree) return new LinearTreeLearner<BASE>(config);
    return new BASE(config);
  }
  if (type == "feature") return new FeatureParallelTreeLearner<BASE>(config);
  if (type == "data") return new DataParallelTreeLearner<BASE>(config);
  if (type == "voting") return new VotingParallelTreeLearner<BASE>(config);
  return nullptr;
}
}  // namespace

TreeLearner* TreeLearner::CreateTreeLearner(const std::string& learner_type, const std::string& device_type,
                                            const Config* config, const bool /*boosting_on_cuda*/) {
  if (device_type == "cpu") return MakeHostLearner<SerialTreeLearner>(learner_type, config);
  if (device_type == "gpu") return MakeHostLearner<GPUTreeLearner>(learner_type, config);
  if (device_type == "cuda") {
    if (learner_type == "serial" || learner_type == "data") {
      if (learner_type == "data" && config->num_gpu <= 1)
        Log::Fatal("device_type=cuda tree_learner=data: set num_gpu=N; one Booster shards the rows over N GPUs itself");
      return new B200TreeLearner(config, config->num_gpu);
    }
    Log::Fatal("device_type=cuda supports tree_learner=serial (and num_gpu=N for data-parallel training on N GPUs)");
  }
  return nullptr;
}

}  // namespace LightGBM
