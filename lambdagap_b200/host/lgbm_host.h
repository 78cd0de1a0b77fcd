// lgbm_host.h -- host side above the C ABI (include/lambdagap_b200.h), in the reference's own language (C++).
//
// It mirrors the reference's operator interfaces for the hot path -- same class names, method names, argument
// meaning and error behaviour -- so that a user of the reference finds what they expect:
//   Config            include/LightGBM/config.h (the fields the path reads; key=value strings, aliases)
//   BinMapper         include/LightGBM/bin.h:86-260, src/io/bin.cpp:78-508            (numerical features)
//   Dataset           include/LightGBM/dataset.h, src/io/dataset_loader.cpp:593-760    (dense matrices)
//   Tree              include/LightGBM/tree.h, src/io/tree.cpp                          (numerical splits)
//   ObjectiveFunction include/LightGBM/objective_function.h:19-122                      (regression, binary, lambdarank)
//   TreeLearner       include/LightGBM/tree_learner.h:27-115                            (B200TreeLearner)
//   GBDT              src/boosting/gbdt.{h,cpp}                                         (TrainOneIter loop)
// and exports the reference's C API names for that subset (host/c_api.cpp <-> src/c_api.cpp).
// Everything that touches the GPU goes through the C ABI only; this library links no CUDA symbol itself.
#ifndef LAMBDAGAP_B200_HOST_LGBM_HOST_H_
#define LAMBDAGAP_B200_HOST_LGBM_HOST_H_

#include <cmath>
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <limits>
#include <memory>
#include <stdexcept>
#include <string>
#include <unordered_map>
#include <vector>

#include "lambdagap_b200.h"

namespace lgbm_b200 {

typedef int32_t data_size_t;  // include/LightGBM/meta.h:26-52
typedef float score_t;
typedef float label_t;
const double kEpsilon = 1e-15f;
const double kZeroThreshold = 1e-35f;
const double kSparseThreshold = 0.7;

// ---- Log (include/LightGBM/utils/log.h:78-185): Fatal throws std::runtime_error, caught at the C API boundary
struct Log {
  static int level;  // <0 fatal only, 0 warning, 1 info, 2 debug  (verbosity)
  [[noreturn]] static void Fatal(const char* fmt, ...);
  static void Warning(const char* fmt, ...);
  static void Info(const char* fmt, ...);
  static void Debug(const char* fmt, ...);
};
void GblCheck(int ret);  // non-zero C-ABI status -> Log::Fatal(gbl_last_error())

// ---- Config
struct Config {
  // core
  std::string objective = "regression";
  std::string boosting = "gbdt";
  std::string tree_learner = "serial";
  std::string device_type = "cuda";
  int num_iterations = 100;
  double learning_rate = 0.1;
  int num_leaves = 31;
  int num_threads = 0;
  int seed = 0;
  bool deterministic = false;
  int num_gpu = 1;
  int gpu_device_id = 0;
  // learning control
  int max_depth = -1;
  int min_data_in_leaf = 20;
  double min_sum_hessian_in_leaf = 1e-3;
  double bagging_fraction = 1.0;
  int bagging_freq = 0;
  int bagging_seed = 3;
  double pos_bagging_fraction = 1.0;
  double neg_bagging_fraction = 1.0;
  bool bagging_by_query = false;
  double feature_fraction = 1.0;
  int feature_fraction_seed = 2;
  double feature_fraction_bynode = 1.0;
  bool extra_trees = false;
  double max_delta_step = 0.0;
  double lambda_l1 = 0.0;
  double lambda_l2 = 0.0;
  double linear_lambda = 0.0;
  double min_gain_to_split = 0.0;
  double path_smooth = 0.0;
  bool linear_tree = false;
  int verbosity = 1;
  bool use_quantized_grad = false;
  int num_grad_quant_bins = 4;
  bool quant_train_renew_leaf = false;
  bool stochastic_rounding = true;
  // dataset
  int max_bin = 255;
  int min_data_in_bin = 3;
  int bin_construct_sample_cnt = 200000;
  int data_random_seed = 1;
  bool use_missing = true;
  bool zero_as_missing = false;
  bool feature_pre_filter = true;
  bool enable_bundle = true;
  // objective
  bool boost_from_average = true;
  bool is_unbalance = false;
  double scale_pos_weight = 1.0;
  double sigmoid = 1.0;
  bool reg_sqrt = false;
  std::string lambdarank_target = "ndcg";
  int lambdarank_truncation_level = 30;
  bool lambdarank_norm = true;
  double lambdagap_weight = 1.0;
  std::vector<double> label_gain;
  // metric
  std::string metric = "";
  std::vector<std::string> metric_types;  // Config::metric after GetMetricType (src/io/config.cpp:152-163)
  std::vector<int> eval_at;
  bool is_provide_training_metric = false;

  // Config::Str2Map + Config::Set (src/io/config.cpp, config_auto.cpp alias table)
  void Set(const std::string& parameters);
  std::string ToString() const;  // "[key: value]" lines of the model file's parameters section (subset)
  static std::unordered_map<std::string, std::string> Str2Map(const std::string& parameters);
};

// ---- BinMapper
enum MissingType { kMissingNone = 0, kMissingZero = 1, kMissingNaN = 2 };

class BinMapper {
 public:
  // BinMapper::FindBin (src/io/bin.cpp:313-508), numerical features only.  `values` holds the sampled non-zero
  // values (NaN allowed) and is reordered in place; total_sample_cnt counts the sampled rows (zeros implied).
  void FindBin(double* values, int num_sample_values, size_t total_sample_cnt, int max_bin, int min_data_in_bin,
               int min_split_data, bool pre_filter, bool use_missing, bool zero_as_missing);
  // BinMapper::ValueToBin (include/LightGBM/bin.h:612-649)
  uint32_t ValueToBin(double value) const;
  double BinToValue(uint32_t bin) const { return bin_upper_bound_[bin]; }
  int num_bin() const { return num_bin_; }
  MissingType missing_type() const { return missing_type_; }
  bool is_trivial() const { return is_trivial_; }
  double sparse_rate() const { return sparse_rate_; }
  uint32_t GetDefaultBin() const { return default_bin_; }
  uint32_t GetMostFreqBin() const { return most_freq_bin_; }
  const std::vector<double>& bin_upper_bound() const { return bin_upper_bound_; }
  std::string bin_info_string() const;  // "[min:max]" of the model file's feature_infos

 private:
  int num_bin_ = 1;
  MissingType missing_type_ = kMissingNone;
  std::vector<double> bin_upper_bound_;
  bool is_trivial_ = true;
  double sparse_rate_ = 1.0;
  double min_val_ = 0.0, max_val_ = 0.0;
  uint32_t default_bin_ = 0, most_freq_bin_ = 0;
};

// LightGBM's LCG (include/LightGBM/utils/random.h:20-120)
class Random {
 public:
  explicit Random(int seed) : x(static_cast<unsigned int>(seed)) {}
  int NextInt(int lower_bound, int upper_bound) { return (RandInt32()) % (upper_bound - lower_bound) + lower_bound; }
  float NextFloat() { return static_cast<float>(RandInt16()) / (32768.0f); }
  std::vector<int> Sample(int N, int K);

 private:
  int RandInt16() {
    x = (214013 * x + 2531011);
    return static_cast<int>((x >> 16) & 0x7FFF);
  }
  int RandInt32() {
    x = (214013 * x + 2531011);
    return static_cast<int>(x & 0x7FFFFFFF);
  }
  unsigned int x = 123456789;
};

// ---- Metadata + Dataset (dense numerical matrices)
struct Metadata {
  std::vector<label_t> label;
  std::vector<label_t> weights;
  std::vector<data_size_t> query_boundaries;  // num_queries + 1 entries when present
  int num_queries() const { return query_boundaries.empty() ? 0 : static_cast<int>(query_boundaries.size()) - 1; }
};

class Dataset {
 public:
  ~Dataset();
  // LGBM_DatasetCreateFromMat (src/c_api.cpp:1303-1405): sample rows, find bins on the host, bin every cell on
  // the device.  `reference` != nullptr aligns the bin mappers with another dataset (validation data).
  static Dataset* CreateFromMat(const void* data, int data_type, int32_t nrow, int32_t ncol, int is_row_major,
                                const Config& config, const Dataset* reference);
  // Same, but the bin mappers come from an explicit sample matrix (every row of it) instead of rows sampled from
  // `data`: the role LGBM_DatasetCreateFromSampledColumn (src/c_api.cpp:1063-1100) plays for distributed loading,
  // where every rank must derive identical mappers before binning its own row shard.
  static Dataset* CreateFromMatWithSample(const void* data, int data_type, int32_t nrow, int32_t ncol, int is_row_major,
                                          const void* sample, int32_t sample_nrow, int64_t num_total_rows,
                                          const Config& config);
  static Dataset* CreateFromMatWithSample(const void* data, int data_type, int32_t nrow, int32_t ncol, int is_row_major,
                                          const void* sample, int32_t sample_nrow, int64_t num_total_rows,
                                          const Config& config, const Dataset* reference);
  void SetField(const char* name, const void* data, int num_element, int type);
  data_size_t num_data() const { return num_data_; }
  int num_features() const { return static_cast<int>(used_feature_map_inv_.size()); }
  int num_total_features() const { return num_total_features_; }
  int RealFeatureIndex(int inner) const { return used_feature_map_inv_[inner]; }
  int InnerFeatureIndex(int col) const { return used_feature_map_[col]; }
  const BinMapper* FeatureBinMapper(int inner) const { return bin_mappers_[inner].get(); }
  double RealThreshold(int inner, uint32_t threshold) const { return bin_mappers_[inner]->BinToValue(threshold); }
  const Metadata& metadata() const { return metadata_; }
  std::vector<std::string> feature_names() const;
  std::vector<std::string> feature_infos() const;
  gbl_dataset* device_handle() const { return device_; }
  int device_id() const { return device_id_; }

 private:
  data_size_t num_data_ = 0;
  int num_total_features_ = 0;
  std::vector<int> used_feature_map_;      // column -> inner or -1
  std::vector<int> used_feature_map_inv_;  // inner -> column
  std::vector<std::unique_ptr<BinMapper>> bin_mappers_;        // per inner feature
  std::vector<std::shared_ptr<BinMapper>> all_bin_mappers_;    // per column (trivial ones kept for feature_infos)
  Metadata metadata_;
  gbl_dataset* device_ = nullptr;
  int device_id_ = 0;
  friend class GBDT;
};

// ---- Tree
class Tree {
 public:
  explicit Tree(int max_leaves);
  // Tree::Split (src/io/tree.cpp:60-74, include/LightGBM/tree.h:543-583)
  int Split(int leaf, int feature, int real_feature, uint32_t threshold_bin, double threshold_double,
            double left_value, double right_value, int left_cnt, int right_cnt, double left_weight,
            double right_weight, float gain, MissingType missing_type, bool default_left);
  void Shrinkage(double rate);
  void AddBias(double val);
  void AsConstantTree(double val, int count = 0);
  void SetLeafOutput(int leaf, double output) { leaf_value_[leaf] = MaybeRoundToZero(output); }
  double LeafOutput(int leaf) const { return leaf_value_[leaf]; }
  int num_leaves() const { return num_leaves_; }
  double Predict(const double* feature_values) const;  // raw feature values, Tree::Predict
  std::string ToString() const;                          // Tree::ToString (src/io/tree.cpp:339-395)
  const std::vector<double>& leaf_values() const { return leaf_value_; }
  int split_feature(int node) const { return split_feature_[node]; }
  float split_gain(int node) const { return split_gain_[node]; }

 private:
  static double MaybeRoundToZero(double v) { return std::fabs(v) > kZeroThreshold ? v : 0.0; }
  int max_leaves_;
  int num_leaves_ = 1;
  std::vector<int> left_child_, right_child_, split_feature_inner_, split_feature_;
  std::vector<uint32_t> threshold_in_bin_;
  std::vector<double> threshold_;
  std::vector<int8_t> decision_type_;
  std::vector<float> split_gain_;
  std::vector<int> leaf_parent_;
  std::vector<double> leaf_value_, leaf_weight_;
  std::vector<int> leaf_count_;
  std::vector<double> internal_value_, internal_weight_;
  std::vector<int> internal_count_;
  std::vector<int> leaf_depth_;
  double shrinkage_ = 1.0;
};

// ---- ObjectiveFunction: device-resident objectives behind the reference's interface
class ObjectiveFunction {
 public:
  // ObjectiveFunction::CreateObjectiveFunction (src/objective/objective_function.cpp:19-103)
  static ObjectiveFunction* CreateObjectiveFunction(const std::string& type, const Config& config);
  ~ObjectiveFunction();
  void Init(const Metadata& metadata, data_size_t num_data, int device_id);
  // score / gradients / hessians are DEVICE pointers (boosting on the device, cf. src/boosting/gbdt.cpp:849-858)
  void GetGradients(const double* score, score_t* gradients, score_t* hessians, void* stream) const;
  double BoostFromScore(int class_id) const;
  // with a communicator: label / weight sums are reduced over ranks first (BinaryLogloss::BoostFromScore does
  // Network::GlobalSyncUpBySum, binary_objective.hpp:157-160), so every rank starts from the global init score
  double BoostFromScoreGlobal(const Metadata& metadata, gbl_comm* comm) const;
  // with a communicator: positive / negative row counts summed over ranks (binary_objective.hpp:75-77)
  void SyncLabelCounts(const Metadata& metadata, gbl_comm* comm) const;
  bool IsConstantHessian() const;
  const char* GetName() const { return name_.c_str(); }
  std::string ToString() const;
  void ConvertOutput(const double* input, double* output) const;

 private:
  ObjectiveFunction() {}
  std::string name_;
  Config config_;
  gbl_objective* handle_ = nullptr;
};

// ---- Metric: evaluation on device-resident scores behind the reference's interface (include/LightGBM/metric.h)
class Metric {
 public:
  // Metric::CreateMetric (src/metric/metric.cpp:20-133): l2, rmse, l1, binary_logloss, binary_error, auc, ndcg;
  // nullptr for "custom" / unknown names like the reference
  static Metric* CreateMetric(const std::string& type, const Config& config);
  ~Metric();
  // label / weights / query boundaries of `metadata` are copied to the device at the first Eval
  void Init(const Metadata& metadata, data_size_t num_data, int device_id);
  const std::vector<std::string>& GetName() const { return name_; }
  double factor_to_bigger_better() const { return bigger_better_ ? 1.0 : -1.0; }
  // score: DEVICE pointer; kernels run on `stream` (the learner's), then the few result bytes are awaited
  std::vector<double> Eval(const double* score, const ObjectiveFunction* objective, void* stream) const;

 private:
  Metric() {}
  std::string type_;
  Config config_;
  std::vector<std::string> name_;
  bool bigger_better_ = false;
  const Metadata* metadata_ = nullptr;
  data_size_t num_data_ = 0;
  int device_id_ = 0;
  mutable gbl_metric* handle_ = nullptr;
  mutable bool handle_with_objective_ = false;
};

// ---- TreeLearner
class TreeLearner {
 public:
  virtual ~TreeLearner() {}
  virtual void Init(const Dataset* train_data, bool is_constant_hessian) = 0;
  virtual void ResetIsConstantHessian(bool is_constant_hessian) = 0;
  virtual void ResetConfig(const Config* config) = 0;
  // gradients / hessians: device pointers
  virtual Tree* Train(const score_t* gradients, const score_t* hessians, bool is_first_tree) = 0;
  virtual void SetBaggingData(const Dataset* subset, const data_size_t* used_indices, data_size_t num_data) = 0;
  // out_score: device pointer
  virtual void AddPredictionToScore(const Tree* tree, double* out_score) const = 0;
  // TreeLearner::CreateTreeLearner (src/treelearner/tree_learner.cpp:15-57)
  static TreeLearner* CreateTreeLearner(const std::string& learner_type, const std::string& device_type,
                                        const Config* config, gbl_comm* comm);
};

class B200TreeLearner : public TreeLearner {
 public:
  B200TreeLearner(const Config* config, gbl_comm* comm) : config_(config), comm_(comm) {}
  ~B200TreeLearner() override;
  void Init(const Dataset* train_data, bool is_constant_hessian) override;
  void ResetIsConstantHessian(bool is_constant_hessian) override;
  void ResetConfig(const Config* config) override;
  Tree* Train(const score_t* gradients, const score_t* hessians, bool is_first_tree) override;
  void SetBaggingData(const Dataset* subset, const data_size_t* used_indices, data_size_t num_data) override;
  void AddPredictionToScore(const Tree* tree, double* out_score) const override;
  gbl_learner* handle() const { return handle_; }
  // BaggingSampleStrategy::Bagging drawn on the device (the reference's bag, bit for bit); returns the bag size
  data_size_t BaggingOnDevice(int bagging_seed, double bagging_fraction, bool reset);
  // bagging_by_query on the device; the boundaries are uploaded when not null
  data_size_t BaggingByQueryOnDevice(int bagging_seed, double bagging_fraction, const data_size_t* query_boundaries,
                                     int num_queries, bool reset);
  // BalancedBaggingHelper on the device; `label` is uploaded when not null
  data_size_t BalancedBaggingOnDevice(int bagging_seed, double pos_fraction, double neg_fraction, const label_t* label, bool reset);
  // device-resident pipeline used by GBDT: enqueue the tree, update the scores from the leaf outputs that are
  // still on the device, and only then wait for the split records
  void TrainAsync(const score_t* gradients, const score_t* hessians);
  void AddPredictionToScoreAsync(double shrinkage, double* out_score) const;
  // rows outside the bag (GBDT::UpdateScore, src/boosting/gbdt.cpp:505-513); no-op without bagging
  void AddPredictionOutOfBagAsync(double shrinkage, double* out_score) const;
  Tree* FinishTrain();

 private:
  gbl_learner_config MakeConfig() const;
  void InitQuantRandom();
  // ColSampler (src/treelearner/col_sampler.hpp:20-90), by-tree sampling: SetTrainingData / SetConfig / ResetByTree
  void ColSamplerSetup(bool reseed);
  void ColSamplerResetByTree();
  Random col_random_{2};
  int col_seed_ = 2;
  std::vector<int> valid_feature_indices_;   // Dataset::ValidFeatureIndices(): real indices of the used features
  std::vector<int8_t> is_feature_used_;      // per inner feature
  int used_cnt_bytree_ = 0;
  bool need_reset_bytree_ = false;
  gbl_learner_config last_config_{};
  bool have_last_config_ = false;
  double last_feature_fraction_ = 1.0;
  const Config* config_;
  gbl_comm* comm_;
  const Dataset* train_data_ = nullptr;
  bool is_constant_hessian_ = false;
  gbl_learner* handle_ = nullptr;
  std::vector<gbl_split_record> records_;
  // GradientDiscretizer random state (src/treelearner/gradient_discretizer.cpp:16-41,116)
  bool quant_random_ready_ = false;
  std::shared_ptr<struct QuantRandomStart> quant_start_;  // shared_ptr: the type is private to tree_learner.cpp
};

// ---- GBDT
class GBDT {
 public:
  GBDT() {}
  ~GBDT();
  // GBDT::Init (src/boosting/gbdt.cpp:53-162)
  void Init(const Config* config, const Dataset* train_data, ObjectiveFunction* objective, gbl_comm* comm);
  // GBDT::TrainOneIter (src/boosting/gbdt.cpp:346-454); host gradients for objective=custom, nullptr otherwise.
  bool TrainOneIter(const score_t* gradients, const score_t* hessians);
  // GBDT::AddValidDataset (src/boosting/gbdt.cpp:184-226); before the first iteration only (the device keeps the
  // last tree, not the whole model).  The metrics stay owned by the caller (the Booster), as in the reference.
  void AddValidDataset(const Dataset* valid_data, const std::vector<const Metric*>& valid_metrics);
  void AddTrainingMetrics(const std::vector<const Metric*>& training_metrics) { training_metrics_ = training_metrics; }
  // GBDT::GetEvalAt (src/boosting/gbdt.cpp:620-640): 0 = training data, i > 0 = the i-th validation set
  std::vector<double> GetEvalAt(int data_idx) const;
  int64_t GetNumPredictAt(int data_idx) const;
  void GetPredictAt(int data_idx, double* out_result, int64_t* out_len);  // GBDT::GetPredictAt: converted outputs
  void GetTrainScore(double* out) const;                                  // raw scores (device -> host)
  std::string SaveModelToString(int start_iteration, int num_iteration) const;
  void PredictForMat(const void* data, int data_type, int32_t nrow, int32_t ncol, int is_row_major, bool raw_score,
                     double* out) const;
  int NumberOfTotalModel() const { return static_cast<int>(models_.size()); }
  int MaxFeatureIdx() const { return max_feature_idx_; }
  // GBDT::FeatureImportance (src/boosting/gbdt_model_text.cpp:627-661): 0 = split counts, 1 = total gains
  std::vector<double> FeatureImportance(int num_iteration, int importance_type) const;
  // GBDT::ResetConfig (src/boosting/gbdt.cpp:776-830) for the parameters this layer implements
  void ResetConfig(const Config* config);
  int GetCurrentIteration() const { return iter_; }
  const std::vector<std::unique_ptr<Tree>>& models() const { return models_; }
  TreeLearner* tree_learner() { return tree_learner_.get(); }
  double* device_score() const { return score_dev_; }

 private:
  double BoostFromAverage();
  // BaggingSampleStrategy::Bagging (src/boosting/bagging.hpp:31-139), row-level bagging
  void Bagging(int iter);
  std::vector<Random> bagging_rands_;     // one generator per block of 1024 rows (bagging.hpp:170-174); host path only
  std::vector<data_size_t> bag_data_indices_;
  bool bagging_ready_ = false;
  double bag_fraction_used_ = 1.0, bag_pos_used_ = 1.0, bag_neg_used_ = 1.0;
  bool balanced_bagging_ = false;
  bool BaggingEnabled(const Config* config, bool* balanced) const;
  data_size_t NumPositiveData() const;  // counted once, on first need
  mutable data_size_t num_pos_ = -1;
  bool bagging_off_ = false;            // "no bagging" decided once, so GBDT::Bagging is O(1) per iteration
  int bag_freq_used_ = 0;
  data_size_t bag_data_cnt_ = 0;
  bool need_re_bagging_ = false;
  gbl_comm* comm_ = nullptr;
  const Config* config_ = nullptr;
  const Dataset* train_data_ = nullptr;
  ObjectiveFunction* objective_function_ = nullptr;
  std::unique_ptr<TreeLearner> tree_learner_;
  std::vector<std::unique_ptr<Tree>> models_;
  int iter_ = 0;
  int num_data_ = 0;
  double shrinkage_rate_ = 0.1;
  int max_feature_idx_ = 0;
  int device_id_ = 0;
  bool has_init_score_ = false;
  double* score_dev_ = nullptr;
  std::vector<const Dataset*> valid_data_;
  std::vector<double*> valid_score_dev_;  // ScoreUpdater of each validation set (score_updater.hpp:22-125), in HBM
  std::vector<std::vector<const Metric*>> valid_metrics_;
  std::vector<const Metric*> training_metrics_;
  void* Stream() const;
  float* grad_dev_ = nullptr;
  float* hess_dev_ = nullptr;
};

}  // namespace lgbm_b200

#endif  // LAMBDAGAP_B200_HOST_LGBM_HOST_H_
