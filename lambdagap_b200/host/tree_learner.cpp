// tree_learner.cpp -- B200TreeLearner: the reference's TreeLearner interface over the C ABI, plus the
// ObjectiveFunction wrapper.
//
// Reference: TreeLearner (include/LightGBM/tree_learner.h:27-115), SerialTreeLearner::Train / SplitInner
// (src/treelearner/serial_tree_learner.cpp:179-245,766-922), GradientDiscretizer::Init random arrays
// (src/treelearner/gradient_discretizer.cpp:16-41,116), ObjectiveFunction factory
// (src/objective/objective_function.cpp:19-103).
#include <algorithm>
#include <cstring>
#include <random>

#include "lgbm_host.h"

namespace lgbm_b200 {

// per-iteration start offset of the stochastic-rounding arrays (gradient_discretizer.cpp:23-24,116)
struct QuantRandomStart {
  std::mt19937 eng;
  std::uniform_int_distribution<data_size_t> dist;
};

TreeLearner* TreeLearner::CreateTreeLearner(const std::string& learner_type, const std::string& device_type,
                                            const Config* config, gbl_comm* comm) {
  if (device_type != "cuda" && device_type != "gpu") {
    Log::Fatal("device_type=%s: only the B200 (cuda) tree learner exists in this library", device_type.c_str());
  }
  if (learner_type == "serial") {
    if (comm != nullptr) Log::Fatal("tree_learner=serial cannot be combined with a communicator; use tree_learner=data");
    return new B200TreeLearner(config, nullptr);
  }
  if (learner_type == "data" || learner_type == "data_parallel") {
    if (comm == nullptr) Log::Fatal("tree_learner=data needs an initialised communicator (LGBM_NetworkInitWithNccl)");
    return new B200TreeLearner(config, comm);
  }
  Log::Fatal("Unknown tree learner type %s", learner_type.c_str());
}

B200TreeLearner::~B200TreeLearner() {
  if (handle_ != nullptr) gbl_learner_free(handle_);
}

gbl_learner_config B200TreeLearner::MakeConfig() const {
  gbl_learner_config c;
  c.num_leaves = config_->num_leaves;
  c.max_depth = config_->max_depth;
  c.min_data_in_leaf = config_->min_data_in_leaf;
  c.use_quantized_grad = config_->use_quantized_grad ? 1 : 0;
  c.num_grad_quant_bins = config_->num_grad_quant_bins;
  c.stochastic_rounding = config_->stochastic_rounding ? 1 : 0;
  c.quant_train_renew_leaf = config_->quant_train_renew_leaf ? 1 : 0;
  c.is_constant_hessian = is_constant_hessian_ ? 1 : 0;
  c.min_sum_hessian_in_leaf = config_->min_sum_hessian_in_leaf;
  c.lambda_l1 = config_->lambda_l1;
  c.lambda_l2 = config_->lambda_l2;
  c.min_gain_to_split = config_->min_gain_to_split;
  c.max_delta_step = config_->max_delta_step;
  c.path_smooth = config_->path_smooth;
  return c;
}

void B200TreeLearner::InitQuantRandom() {
  // GradientDiscretizer::Init: one mt19937 pair per block of rows, seeded seed + block and seed + block +
  // num_threads; the values therefore depend on num_threads exactly like the reference's.
  const data_size_t n = train_data_->num_data();
  const int num_threads = std::max(1, config_->num_threads);
  int num_blocks = std::min<int>(num_threads, (n + 512 - 1) / 512);
  data_size_t block_size = n;
  if (num_blocks > 1) {
    block_size = (n + num_blocks - 1) / num_blocks;
    block_size = (block_size + 31) / 32 * 32;  // SIZE_ALIGNED
  }
  std::vector<double> rg(n), rh(n);
  for (int b = 0; b < num_blocks; ++b) {
    const data_size_t start = b * block_size;
    const data_size_t end = std::min(start + block_size, n);
    std::mt19937 ge(config_->seed + b);
    std::uniform_real_distribution<double> gd(0.0f, 1.0f);
    std::mt19937 he(config_->seed + b + num_threads);
    std::uniform_real_distribution<double> hd(0.0f, 1.0f);
    for (data_size_t i = start; i < end; ++i) {
      rg[i] = gd(ge);
      rh[i] = hd(he);
    }
  }
  GblCheck(gbl_learner_set_quant_random(handle_, rg.data(), rh.data(), n));
  quant_start_ = std::make_shared<QuantRandomStart>();
  quant_start_->eng = std::mt19937(config_->seed);
  quant_start_->dist = std::uniform_int_distribution<data_size_t>(0, n);
  quant_random_ready_ = true;
}

void B200TreeLearner::Init(const Dataset* train_data, bool is_constant_hessian) {
  train_data_ = train_data;
  is_constant_hessian_ = is_constant_hessian;
  if (handle_ != nullptr) {
    gbl_learner_free(handle_);
    handle_ = nullptr;
  }
  gbl_learner_config c = MakeConfig();
  GblCheck(gbl_learner_create(train_data->device_handle(), &c, comm_, &handle_));
  std::memset(&last_config_, 0, sizeof(last_config_));
  last_config_ = c;
  have_last_config_ = true;
  last_feature_fraction_ = config_->feature_fraction;
  records_.resize(config_->num_leaves);
  quant_random_ready_ = false;
  if (config_->use_quantized_grad && config_->stochastic_rounding) InitQuantRandom();
  ColSamplerSetup(true);  // SerialTreeLearner::Init -> col_sampler_.SetTrainingData (serial_tree_learner.cpp:59)
}

// ColSampler::SetTrainingData / SetConfig (col_sampler.hpp:39-75): the generator is created from
// feature_fraction_seed once and re-created only when the seed changes; both end with one ResetByTree.
void B200TreeLearner::ColSamplerSetup(bool reseed) {
  if (reseed || col_seed_ != config_->feature_fraction_seed) {
    col_seed_ = config_->feature_fraction_seed;
    col_random_ = Random(col_seed_);
  }
  valid_feature_indices_.clear();
  for (int i = 0; i < train_data_->num_features(); ++i) valid_feature_indices_.push_back(train_data_->RealFeatureIndex(i));
  std::sort(valid_feature_indices_.begin(), valid_feature_indices_.end());
  is_feature_used_.assign(train_data_->num_features(), 1);
  const int total = static_cast<int>(valid_feature_indices_.size());
  if (config_->feature_fraction >= 1.0f) {
    need_reset_bytree_ = false;
    used_cnt_bytree_ = total;
    GblCheck(gbl_learner_set_feature_mask(handle_, nullptr));
  } else {
    need_reset_bytree_ = true;
    // ColSampler::GetCnt: max(RoundInt(total * fraction), min(1, total))
    used_cnt_bytree_ = std::max(static_cast<int>(total * config_->feature_fraction + 0.5f), std::min(1, total));
  }
  ColSamplerResetByTree();
}

// ColSampler::ResetByTree (col_sampler.hpp:77-90): Random::Sample picks used_cnt_bytree_ of the valid features
void B200TreeLearner::ColSamplerResetByTree() {
  if (!need_reset_bytree_) return;
  std::fill(is_feature_used_.begin(), is_feature_used_.end(), 0);
  const std::vector<int> used = col_random_.Sample(static_cast<int>(valid_feature_indices_.size()), used_cnt_bytree_);
  for (int i : used) is_feature_used_[train_data_->InnerFeatureIndex(valid_feature_indices_[i])] = 1;
  GblCheck(gbl_learner_set_feature_mask(handle_, is_feature_used_.data()));
}

void B200TreeLearner::ResetIsConstantHessian(bool is_constant_hessian) {
  is_constant_hessian_ = is_constant_hessian;
  gbl_learner_config c = MakeConfig();
  GblCheck(gbl_learner_reset_config(handle_, &c));
}

void B200TreeLearner::ResetConfig(const Config* config) {
  const double old_fraction = config_ == config ? last_feature_fraction_ : config_->feature_fraction;
  config_ = config;
  gbl_learner_config c = MakeConfig();
  // a learning-rate schedule resets the parameters before every iteration: only touch the device learner (which drops
  // its captured tree) when a field it reads has changed
  if (!have_last_config_ || std::memcmp(&c, &last_config_, sizeof(c)) != 0) {
    GblCheck(gbl_learner_reset_config(handle_, &c));
    last_config_ = c;
    have_last_config_ = true;
  }
  records_.resize(config_->num_leaves);
  if (col_seed_ != config_->feature_fraction_seed || old_fraction != config_->feature_fraction)
    ColSamplerSetup(false);  // col_sampler_.SetConfig (serial_tree_learner.cpp:168)
  last_feature_fraction_ = config_->feature_fraction;
}

void B200TreeLearner::SetBaggingData(const Dataset* /*subset*/, const data_size_t* used_indices, data_size_t num_data) {
  GblCheck(gbl_learner_set_bagging_data(handle_, used_indices, num_data));
}

data_size_t B200TreeLearner::BaggingOnDevice(int bagging_seed, double bagging_fraction, bool reset) {
  int32_t cnt = 0;
  GblCheck(gbl_learner_bagging_dev(handle_, bagging_seed, bagging_fraction, reset ? 1 : 0, &cnt));
  return cnt;
}

data_size_t B200TreeLearner::BaggingByQueryOnDevice(int bagging_seed, double bagging_fraction, const data_size_t* query_boundaries,
                                                    int num_queries, bool reset) {
  int32_t cnt = 0;
  GblCheck(gbl_learner_bagging_by_query_dev(handle_, bagging_seed, bagging_fraction, query_boundaries, num_queries, reset ? 1 : 0, &cnt));
  return cnt;
}

data_size_t B200TreeLearner::BalancedBaggingOnDevice(int bagging_seed, double pos_fraction, double neg_fraction,
                                                     const label_t* label, bool reset) {
  int32_t cnt = 0;
  GblCheck(gbl_learner_bagging_balanced_dev(handle_, bagging_seed, pos_fraction, neg_fraction, label, reset ? 1 : 0, &cnt));
  return cnt;
}

Tree* B200TreeLearner::Train(const score_t* gradients, const score_t* hessians, bool /*is_first_tree*/) {
  TrainAsync(gradients, hessians);
  return FinishTrain();
}

void B200TreeLearner::TrainAsync(const score_t* gradients, const score_t* hessians) {
  ColSamplerResetByTree();  // SerialTreeLearner::BeforeTrain (serial_tree_learner.cpp:293)
  if (quant_random_ready_) {
    GblCheck(gbl_learner_set_quant_random_start(handle_, quant_start_->dist(quant_start_->eng)));
  }
  GblCheck(gbl_learner_train_async(handle_, gradients, hessians));
}

void B200TreeLearner::AddPredictionToScoreAsync(double shrinkage, double* out_score) const {
  GblCheck(gbl_learner_add_prediction_to_score_dev(handle_, shrinkage, 0.0, out_score));
}

void B200TreeLearner::AddPredictionOutOfBagAsync(double shrinkage, double* out_score) const {
  GblCheck(gbl_learner_add_prediction_out_of_bag_dev(handle_, shrinkage, 0.0, out_score));
}

Tree* B200TreeLearner::FinishTrain() {
  gbl_tree_result res;
  GblCheck(gbl_learner_train_finish(handle_, &res, records_.data()));
  std::unique_ptr<Tree> tree(new Tree(config_->num_leaves));
  tree->SetLeafOutput(0, res.root_output);  // serial_tree_learner.cpp:205-208
  for (int k = 0; k < res.num_splits; ++k) {
    const gbl_split_record& r = records_[k];
    const BinMapper* m = train_data_->FeatureBinMapper(r.feature);
    // SerialTreeLearner::SplitInner -> Tree::Split, serial_tree_learner.cpp:798-810
    tree->Split(r.leaf, r.feature, r.real_feature, r.threshold_bin, train_data_->RealThreshold(r.feature, r.threshold_bin),
                r.left_output, r.right_output, r.left_count, r.right_count, r.left_sum_hessian, r.right_sum_hessian,
                static_cast<float>(r.gain + config_->min_gain_to_split), m->missing_type(), r.default_left != 0);
  }
  if (config_->use_quantized_grad && config_->quant_train_renew_leaf && res.num_splits > 0) {
    // serial_tree_learner.cpp:238-241: the device recomputed every leaf output from the true gradient sums
    std::vector<double> values(res.num_splits + 1);
    GblCheck(gbl_learner_get_leaf_values(handle_, values.data(), res.num_splits + 1));
    for (int leaf = 0; leaf <= res.num_splits; ++leaf) tree->SetLeafOutput(leaf, values[leaf]);
  }
  if (res.num_splits == 0) Log::Warning("No further splits with positive gain");
  return tree.release();
}

void B200TreeLearner::AddPredictionToScore(const Tree* tree, double* out_score) const {
  if (tree->num_leaves() <= 1) return;  // serial_tree_learner.h:101-103
  GblCheck(gbl_learner_add_prediction_to_score(handle_, tree->leaf_values().data(), tree->num_leaves(), out_score));
}

// ------------------------------------------------------------------------------------------- objectives
static const char* kRankTargets[] = {
    // rank_objective.hpp:22-41, same order as the reference's target list
    "ndcg", "lambdaloss-ndcg", "lambdaloss-ndcg-plus-plus", "bndcg", "lambdaloss-bndcg", "lambdaloss-bndcg-plus-plus",
    "precision", "arpk", "lambdaloss-arp1", "lambdaloss-arp2", "ranknet", "bin-ranknet", "lambdagap-s", "lambdagap-x",
    "lambdagap-s-plus", "lambdagap-x-plus", "lambdagap-s-plus-plus", "lambdagap-x-plus-plus"};

int RankTargetIndex(const std::string& name) {
  for (int i = 0; i < 18; ++i) {
    if (name == kRankTargets[i]) return i;
  }
  return -1;
}

ObjectiveFunction* ObjectiveFunction::CreateObjectiveFunction(const std::string& type, const Config& config) {
  if (type == "custom") return nullptr;
  if (type != "regression" && type != "binary" && type != "lambdarank") {
    Log::Fatal("Unknown objective type name: %s (the B200 library provides regression, binary, lambdarank)", type.c_str());
  }
  if (type == "lambdarank" && RankTargetIndex(config.lambdarank_target) < 0) {
    Log::Fatal("Unknown lambdarank_target: %s", config.lambdarank_target.c_str());
  }
  if (type == "regression" && config.reg_sqrt) Log::Fatal("reg_sqrt is not supported by the B200 library");
  ObjectiveFunction* o = new ObjectiveFunction();
  o->name_ = type;
  o->config_ = config;
  return o;
}

ObjectiveFunction::~ObjectiveFunction() {
  if (handle_ != nullptr) gbl_objective_free(handle_);
}

void ObjectiveFunction::Init(const Metadata& metadata, data_size_t num_data, int device_id) {
  if (static_cast<data_size_t>(metadata.label.size()) != num_data) Log::Fatal("Objective: label not set (use LGBM_DatasetSetField)");
  gbl_objective_config c;
  c.kind = name_ == "regression" ? GBL_OBJ_REGRESSION_L2 : (name_ == "binary" ? GBL_OBJ_BINARY : GBL_OBJ_LAMBDARANK);
  c.lambdarank_target = std::max(0, RankTargetIndex(config_.lambdarank_target));
  c.lambdarank_truncation_level = config_.lambdarank_truncation_level;
  c.lambdarank_norm = config_.lambdarank_norm ? 1 : 0;
  c.sigmoid = config_.sigmoid;
  c.lambdagap_weight = config_.lambdagap_weight;
  c.is_unbalance = config_.is_unbalance ? 1 : 0;
  c.num_class = 0;
  c.scale_pos_weight = config_.scale_pos_weight;
  c.num_label_gain = static_cast<int>(config_.label_gain.size());
  c.reserved2 = 0;
  c.label_gain = config_.label_gain.empty() ? nullptr : config_.label_gain.data();
  if (name_ == "lambdarank" && metadata.query_boundaries.empty()) Log::Fatal("Ranking tasks require query information");
  GblCheck(gbl_objective_create(&c, num_data, metadata.label.data(),
                                metadata.weights.empty() ? nullptr : metadata.weights.data(),
                                metadata.query_boundaries.empty() ? nullptr : metadata.query_boundaries.data(),
                                metadata.num_queries(), device_id, &handle_));
}

void ObjectiveFunction::GetGradients(const double* score, score_t* gradients, score_t* hessians, void* stream) const {
  GblCheck(gbl_objective_get_gradients(handle_, score, gradients, hessians, stream));
}

double ObjectiveFunction::BoostFromScore(int /*class_id*/) const {
  double v = 0.0;
  GblCheck(gbl_objective_boost_from_score(handle_, &v));
  return v;
}

void ObjectiveFunction::SyncLabelCounts(const Metadata& md, gbl_comm* comm) const {
  // BinaryLogloss::Init, binary_objective.hpp:75-77: cnt_positive / cnt_negative are summed over machines before the
  // is_unbalance label weights are derived -- otherwise every rank would weight its rows by its own shard's ratio
  if (name_ != "binary" || comm == nullptr) return;
  double cnt[2] = {0.0, 0.0};
  for (label_t l : md.label) cnt[l > 0 ? 0 : 1] += 1.0;
  GblCheck(gbl_comm_allreduce_host_f64(comm, cnt, 2, 0));
  GblCheck(gbl_objective_set_label_counts(handle_, static_cast<int64_t>(cnt[0]), static_cast<int64_t>(cnt[1])));
}

double ObjectiveFunction::BoostFromScoreGlobal(const Metadata& md, gbl_comm* comm) const {
  if (name_ == "lambdarank") return 0.0;
  double sums[2] = {0.0, 0.0};  // suml, sumw
  const size_t n = md.label.size();
  const bool has_w = !md.weights.empty();
  for (size_t i = 0; i < n; ++i) {
    const double w = has_w ? md.weights[i] : 1.0;
    const double l = name_ == "binary" ? (md.label[i] > 0 ? 1.0 : 0.0) : static_cast<double>(md.label[i]);
    sums[0] += l * w;
    sums[1] += w;
  }
  GblCheck(gbl_comm_allreduce_host_f64(comm, sums, 2, 0));
  if (name_ == "binary") {
    double pavg = sums[0] / sums[1];
    pavg = std::min(pavg, 1.0 - kEpsilon);
    pavg = std::max<double>(pavg, kEpsilon);
    return std::log(pavg / (1.0f - pavg)) / config_.sigmoid;
  }
  return sums[0] / sums[1];
}

bool ObjectiveFunction::IsConstantHessian() const {
  int32_t v = 0;
  GblCheck(gbl_objective_is_constant_hessian(handle_, &v));
  return v != 0;
}

std::string ObjectiveFunction::ToString() const {
  // BinaryLogloss::ToString (binary_objective.hpp:176-181), RegressionL2loss (:139-146), LambdarankNDCG
  if (name_ == "binary") {
    char buf[64];
    snprintf(buf, sizeof(buf), "binary sigmoid:%g", config_.sigmoid);
    return buf;
  }
  return name_;
}

void ObjectiveFunction::ConvertOutput(const double* input, double* output) const {
  if (name_ == "binary") {
    output[0] = 1.0f / (1.0f + std::exp(-config_.sigmoid * input[0]));
  } else {
    output[0] = input[0];
  }
}

}  // namespace lgbm_b200
