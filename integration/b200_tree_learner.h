// integration/b200_tree_learner.h -- the B200 tree learner as a subclass of the reference's own TreeLearner.
//
// This file and b200_tree_learner.cpp / b200_factory.cpp are compiled AGAINST THE REFERENCE'S HEADERS (nothing of the
// reference is copied): integration/Makefile builds integration/_ref/lib_lightgbm.so from the reference's sources
// where they lie, with src/treelearner/tree_learner.cpp (the factory) replaced by b200_factory.cpp.  In that library
//   device_type=cuda                        -> B200TreeLearner on one GPU            (reference: CUDASingleGPUTreeLearner)
//   device_type=cuda num_gpu=N (N > 1)      -> the same class over N row shards, one GPU each, inside one Booster
//                                              (reference: Log::Fatal, src/treelearner/tree_learner.cpp:46-50)
// while GBDT, Config, Dataset construction, objectives, metrics, sampling strategies (bagging, GOSS), Tree, model I/O
// and c_api.cpp are the reference's own, unchanged.  Everything that touches a GPU goes through the C ABI of
// include/lambdagap_b200.h (liblambdagap_b200.so).
//
// Interface implemented: include/LightGBM/tree_learner.h:27-115.  Boosting stays on the host in this build (the
// reference compiles its device-resident GBDT paths only under USE_CUDA): Train receives host gradients, copies them
// in, and AddPredictionToScore updates the host score array -- what CUDASingleGPUTreeLearner does when
// boosting_on_cuda is false (src/treelearner/cuda/cuda_single_gpu_tree_learner.cpp:99-105).
#ifndef LAMBDAGAP_B200_INTEGRATION_B200_TREE_LEARNER_H_
#define LAMBDAGAP_B200_INTEGRATION_B200_TREE_LEARNER_H_

#include <LightGBM/config.h>
#include <LightGBM/dataset.h>
#include <LightGBM/tree.h>
#include <LightGBM/tree_learner.h>

#include <memory>
#include <random>
#include <set>
#include <vector>

#include "col_sampler.hpp"  // the reference's own ColSampler (src/treelearner/col_sampler.hpp)
#include "b200_objective.h"
#include "lambdagap_b200.h"

namespace LightGBM {

class B200TreeLearner : public TreeLearner {
 public:
  // num_gpu == 1: one learner on config->gpu_device_id; num_gpu > 1: rows sharded over GPUs 0..num_gpu-1
  B200TreeLearner(const Config* config, int num_gpu);
  ~B200TreeLearner() override;

  void Init(const Dataset* train_data, bool is_constant_hessian) override;
  void ResetIsConstantHessian(bool is_constant_hessian) override;
  void ResetTrainingData(const Dataset* train_data, bool is_constant_hessian) override;
  void ResetConfig(const Config* config) override;
  void SetForcedSplit(const Json* forced_split_json) override;
  Tree* Train(const score_t* gradients, const score_t* hessians, bool is_first_tree) override;
  Tree* FitByExistingTree(const Tree* old_tree, const score_t* gradients, const score_t* hessians) const override;
  Tree* FitByExistingTree(const Tree* old_tree, const std::vector<int>& leaf_pred, const score_t* gradients,
                          const score_t* hessians) const override;
  void SetBaggingData(const Dataset* subset, const data_size_t* used_indices, data_size_t num_data) override;
  void AddPredictionToScore(const Tree* tree, double* out_score) const override;
  void RenewTreeOutput(Tree* tree, const ObjectiveFunction* obj, std::function<double(const label_t*, int)> residual_getter,
                       data_size_t total_num_data, const data_size_t* bag_indices, data_size_t bag_cnt,
                       const double* train_score) const override;

 private:
  struct Shard {  // one GPU: a contiguous slice of the rows
    int device = 0;
    data_size_t begin = 0, count = 0;
    gbl_dataset* dataset = nullptr;
    gbl_comm* comm = nullptr;
    gbl_learner* learner = nullptr;
    std::vector<gbl_split_record> records;
    gbl_tree_result result{};
    std::string error;  // message of a failed call on this shard's thread
  };
  void Release();
  void CheckConfig() const;
  gbl_learner_config MakeConfig() const;
  void BuildDeviceData();
  void CreateLearners();
  void InitQuantRandom();
  void PushFeatureMask();
  void PushMonotone(bool force);
  // runs fn(shard) on one host thread per shard (collective C-ABI calls must be issued concurrently)
  template <typename F>
  void ForEachShard(F fn) const;
  void FetchLeafAssignment() const;

  const Config* config_;
  int num_gpu_;
  const Dataset* train_data_ = nullptr;
  data_size_t num_data_ = 0;
  bool is_constant_hessian_ = false;
  mutable std::vector<Shard> shards_;
  std::unique_ptr<ColSampler> col_sampler_;
  // GradientDiscretizer random state (src/treelearner/gradient_discretizer.cpp:16-41,116)
  bool quant_random_ready_ = false;
  std::mt19937 quant_start_engine_;
  std::uniform_int_distribution<data_size_t> quant_start_dist_;
  // rows the last tree was trained on (SetBaggingData), and the leaf of every row after the last Train
  bool monotone_on_ = false;
  double monotone_penalty_ = 0.0;
  int cat_params_[3] = {0, 0, 0};
  double cat_l2_ = -1.0, cat_smooth_ = -1.0;
  bool bagging_ = false;
  // announced to DeviceGradients (b200_objective.h): Train takes the gradients the B200 objective left on these GPUs
  DeviceLayout layout_;
  bool announced_ = false;
  // GOSS with top_rate + other_rate <= 0.5 trains on a row-subset copy (goss.hpp:101-108) and GBDT then hands Train the
  // gradients of the bag's rows only, packed in bag order (gbdt.cpp:392-400): they are spread back to row positions here
  bool packed_gradients_ = false;
  std::vector<data_size_t> bag_rows_;
  std::vector<score_t> spread_gradients_, spread_hessians_;
  mutable bool leaf_assignment_valid_ = false;
  mutable std::vector<int32_t> leaf_of_row_;
  // Page-locked 2-byte leaf ids (trees of at most 65535 leaves), two buffers used in turn: the device-to-host copy of a
  // tree lands in the buffer the host read two trees ago.  A DMA write into lines the CPU caches still hold runs at a
  // third of the PCIe rate (measured: 1.25 ms instead of 0.49 ms for 21 MB), and with the gradients computed on the
  // device nothing else sweeps the caches between two trees; evicting the lines by hand (clflush) cost 1 ms per tree.
  mutable std::vector<uint16_t> leaf16_buf_[2];
  mutable int leaf16_turn_ = 0;
  const std::vector<uint16_t>& leaf16() const { return leaf16_buf_[leaf16_turn_]; }
  std::set<const score_t*> registered_;    // gradient / hessian buffers of GBDT that have been page-locked
};

}  // namespace LightGBM

#endif  // LAMBDAGAP_B200_INTEGRATION_B200_TREE_LEARNER_H_
