// gbdt.cpp -- the boosting loop above the tree learner, device-resident (scores, gradients and hessians never
// leave HBM; per iteration only the split records of the new tree come back).
//
// Reference: GBDT::Init (src/boosting/gbdt.cpp:53-162), BoostFromAverage (:319-343), Boosting (:222-237),
// TrainOneIter (:346-454), UpdateScore (:495-524), GetPredictAt (:686-720), SaveModelToString
// (src/boosting/gbdt_model_text.cpp:311-400), ScoreUpdater (src/boosting/score_updater.hpp:22-125; the device
// variant src/boosting/cuda/cuda_score_updater.cpp:12-80).
#include <algorithm>
#include <cstring>
#include <sstream>

#include "lgbm_host.h"

namespace lgbm_b200 {

GBDT::~GBDT() {
  if (score_dev_) gbl_device_free(device_id_, score_dev_);
  if (grad_dev_) gbl_device_free(device_id_, grad_dev_);
  if (hess_dev_) gbl_device_free(device_id_, hess_dev_);
  for (double* p : valid_score_dev_) gbl_device_free(device_id_, p);
}

void* GBDT::Stream() const {
  B200TreeLearner* bl = dynamic_cast<B200TreeLearner*>(tree_learner_.get());
  return bl != nullptr ? gbl_learner_stream(bl->handle()) : nullptr;
}

void GBDT::AddValidDataset(const Dataset* valid_data, const std::vector<const Metric*>& valid_metrics) {
  if (valid_data == nullptr) Log::Fatal("GBDT::AddValidDataset: no data");
  // Dataset::CheckAlign (gbdt.cpp:186-188): same used features and bin mappers as the training data
  if (valid_data->num_features() != train_data_->num_features() ||
      valid_data->num_total_features() != train_data_->num_total_features()) {
    Log::Fatal("Cannot add validation data, since it has different bin mappers with training data");
  }
  for (int f = 0; f < train_data_->num_features(); ++f) {
    const BinMapper* a = train_data_->FeatureBinMapper(f);
    const BinMapper* b = valid_data->FeatureBinMapper(f);
    if (train_data_->RealFeatureIndex(f) != valid_data->RealFeatureIndex(f) || a->num_bin() != b->num_bin() ||
        a->missing_type() != b->missing_type() || a->bin_upper_bound().size() != b->bin_upper_bound().size() ||
        std::memcmp(a->bin_upper_bound().data(), b->bin_upper_bound().data(),  // bitwise: the NaN bin's bound is NaN
                    sizeof(double) * a->bin_upper_bound().size()) != 0) {
      Log::Fatal("Cannot add validation data, since it has different bin mappers with training data");
    }
  }
  if (valid_data->device_id() != device_id_) Log::Fatal("Cannot add validation data that lives on another device");
  if (iter_ > 0 || !models_.empty()) {
    // gbdt.cpp:198-203 replays the model on the new data; only the last tree is device-resident here
    Log::Fatal("Validation data must be added before the first boosting iteration");
  }
  void* p = nullptr;
  GblCheck(gbl_device_malloc(device_id_, sizeof(double) * valid_data->num_data(), &p));
  GblCheck(gbl_device_fill_f64(device_id_, static_cast<double*>(p), 0.0, valid_data->num_data()));
  valid_data_.push_back(valid_data);
  valid_score_dev_.push_back(static_cast<double*>(p));
  valid_metrics_.push_back(valid_metrics);
}

std::vector<double> GBDT::GetEvalAt(int data_idx) const {
  if (data_idx < 0 || data_idx > static_cast<int>(valid_score_dev_.size())) Log::Fatal("Check failed: data_idx >= 0 && data_idx <= valid sets");
  std::vector<double> ret;
  const std::vector<const Metric*>& metrics = data_idx == 0 ? training_metrics_ : valid_metrics_[data_idx - 1];
  const double* score = data_idx == 0 ? score_dev_ : valid_score_dev_[data_idx - 1];
  for (const Metric* m : metrics) {
    for (double v : m->Eval(score, objective_function_, Stream())) ret.push_back(v);  // GBDT::EvalOneMetric
  }
  return ret;
}

void GBDT::Init(const Config* config, const Dataset* train_data, ObjectiveFunction* objective, gbl_comm* comm) {
  if (train_data == nullptr) Log::Fatal("GBDT::Init: no training data");
  config_ = config;
  train_data_ = train_data;
  objective_function_ = objective;
  comm_ = comm;
  num_data_ = train_data->num_data();
  shrinkage_rate_ = config->learning_rate;
  device_id_ = train_data->device_id();
  max_feature_idx_ = train_data->num_total_features() - 1;
  iter_ = 0;
  bool is_constant_hessian = false;
  if (objective_function_ != nullptr) {
    objective_function_->Init(train_data->metadata(), num_data_, device_id_);
    if (comm != nullptr) objective_function_->SyncLabelCounts(train_data->metadata(), comm);
    is_constant_hessian = objective_function_->IsConstantHessian();
  }
  tree_learner_.reset(TreeLearner::CreateTreeLearner(comm != nullptr ? "data" : config->tree_learner == "serial" ? "serial" : config->tree_learner,
                                                     config->device_type, config, comm));
  tree_learner_->Init(train_data, is_constant_hessian);
  void* p = nullptr;
  GblCheck(gbl_device_malloc(device_id_, sizeof(double) * num_data_, &p));
  score_dev_ = static_cast<double*>(p);
  GblCheck(gbl_device_fill_f64(device_id_, score_dev_, 0.0, num_data_));
  GblCheck(gbl_device_malloc(device_id_, sizeof(float) * num_data_, &p));
  grad_dev_ = static_cast<float*>(p);
  GblCheck(gbl_device_malloc(device_id_, sizeof(float) * num_data_, &p));
  hess_dev_ = static_cast<float*>(p);
}

// BaggingSampleStrategy (src/boosting/bagging.hpp): every row draws NextFloat() from the generator of its block of
// 1024 rows (seeded bagging_seed + block, state carried from iteration to iteration) and is in the bag when the draw
// is below bagging_fraction; the in-bag rows are handed to the learner in ascending order.  Block-wise generators
// make the result independent of the number of threads, so this sequential version draws the reference's bag.
// bagging.hpp:146-147: plain bagging, or balanced bagging when a positive / negative fraction is set and the objective
// counts positive rows (binary only: ObjectiveFunction::NumPositiveData is 0 elsewhere)
// The labels are walked only when a positive / negative fraction asks for balanced bagging, and then once
// (BinaryLogloss::Init counts them once too, binary_objective.hpp:58-72); the per-iteration guard below is O(1) like
// the reference's (bagging.hpp:29-33 tests counters and flags only).
data_size_t GBDT::NumPositiveData() const {
  if (num_pos_ < 0) {
    data_size_t num_pos = 0;
    if (objective_function_ != nullptr && std::string(objective_function_->GetName()) == "binary") {
      for (label_t l : train_data_->metadata().label) num_pos += l > 0 ? 1 : 0;
    }
    num_pos_ = num_pos;
  }
  return num_pos_;
}

bool GBDT::BaggingEnabled(const Config* config, bool* balanced) const {
  *balanced = (config->pos_bagging_fraction < 1.0 || config->neg_bagging_fraction < 1.0) && NumPositiveData() > 0;
  return (config->bagging_fraction < 1.0 || *balanced) && config->bagging_freq > 0;
}

void GBDT::Bagging(int iter) {
  if (!bagging_ready_) {
    if (bagging_off_) return;  // decided at the first iteration / in ResetConfig, not per iteration
    bool balanced = false;
    if (!BaggingEnabled(config_, &balanced)) {
      bagging_off_ = true;
      return;
    }
    balanced_bagging_ = balanced;
  }
  B200TreeLearner* bl = dynamic_cast<B200TreeLearner*>(tree_learner_.get());
  constexpr int kBlock = 1024;  // bagging_rand_block_
  if (!bagging_ready_) {
    // ResetSampleConfig (bagging.hpp:139-201)
    bagging_ready_ = true;
    bag_fraction_used_ = config_->bagging_fraction;
    bag_pos_used_ = config_->pos_bagging_fraction;
    bag_neg_used_ = config_->neg_bagging_fraction;
    bag_freq_used_ = config_->bagging_freq;
    if (balanced_bagging_) {
      const data_size_t num_pos = NumPositiveData();
      bag_data_cnt_ = static_cast<data_size_t>(num_pos * config_->pos_bagging_fraction) +
                      static_cast<data_size_t>((num_data_ - num_pos) * config_->neg_bagging_fraction);
    } else {
      bag_data_cnt_ = static_cast<data_size_t>(config_->bagging_fraction * num_data_);
    }
    if (bl == nullptr) {
      bag_data_indices_.resize(num_data_);
      for (int i = 0; i < (num_data_ + kBlock - 1) / kBlock; ++i) bagging_rands_.emplace_back(config_->bagging_seed + i);
    }
    need_re_bagging_ = true;
  }
  if (!((bag_data_cnt_ < num_data_ && iter % config_->bagging_freq == 0) || need_re_bagging_)) return;
  const bool first = need_re_bagging_;
  need_re_bagging_ = false;
  if (bl != nullptr) {
    // the same draw on the device: one thread per block generator, compaction by a prefix scan; nothing but the
    // bag size crosses PCIe
    if (config_->bagging_by_query) {
      // bagging.hpp:54-99: queries are drawn, their documents form the bag.  The reference then computes gradients for
      // the sampled queries only (gbdt.cpp:229-232); computing them for every query gives the bagged rows the same values
      const Metadata& md = train_data_->metadata();
      if (md.query_boundaries.empty()) Log::Fatal("bagging_by_query needs query information");
      bag_data_cnt_ = bl->BaggingByQueryOnDevice(config_->bagging_seed, config_->bagging_fraction,
                                                 first ? md.query_boundaries.data() : nullptr, md.num_queries(), first);
    } else if (balanced_bagging_) {
      bag_data_cnt_ = bl->BalancedBaggingOnDevice(config_->bagging_seed, config_->pos_bagging_fraction, config_->neg_bagging_fraction,
                                                  first ? train_data_->metadata().label.data() : nullptr, first);
    } else {
      bag_data_cnt_ = bl->BaggingOnDevice(config_->bagging_seed, config_->bagging_fraction, first);
    }
    Log::Debug("Re-bagging, using %d data to train", bag_data_cnt_);
    return;
  }
  if (config_->bagging_by_query) Log::Fatal("bagging_by_query needs the B200 tree learner");
  data_size_t cnt = 0;
  const std::vector<label_t>& label = train_data_->metadata().label;
  for (data_size_t i = 0; i < num_data_; ++i) {
    // bagging.hpp:233 / :255-262: `NextFloat() < fraction` (float promoted to double)
    const double limit = !balanced_bagging_ ? config_->bagging_fraction
                                            : (label[i] > 0 ? config_->pos_bagging_fraction : config_->neg_bagging_fraction);
    if (static_cast<double>(bagging_rands_[i / kBlock].NextFloat()) < limit) bag_data_indices_[cnt++] = i;
  }
  bag_data_cnt_ = cnt;
  Log::Debug("Re-bagging, using %d data to train", bag_data_cnt_);
  tree_learner_->SetBaggingData(nullptr, bag_data_indices_.data(), bag_data_cnt_);
}

double GBDT::BoostFromAverage() {
  // gbdt.cpp:319-343 (single GPU: no Network::GlobalSyncUpByMean)
  if (models_.empty() && !has_init_score_ && objective_function_ != nullptr && config_->boost_from_average) {
    // single GPU: the objective's own value; sharded rows: label sums reduced over ranks first, so that the start
    // score (and with it every tree) is the one a single GPU would compute on the full data
    const double init_score = comm_ != nullptr ? objective_function_->BoostFromScoreGlobal(train_data_->metadata(), comm_)
                                               : objective_function_->BoostFromScore(0);
    if (std::fabs(init_score) > kEpsilon) {
      // ScoreUpdater::AddScore(val): one fill since all scores are still zero
      GblCheck(gbl_device_fill_f64(device_id_, score_dev_, init_score, num_data_));
      // gbdt.cpp:331-334: the validation scores start from the same value
      for (size_t i = 0; i < valid_score_dev_.size(); ++i) {
        GblCheck(gbl_device_fill_f64(device_id_, valid_score_dev_[i], init_score, valid_data_[i]->num_data()));
      }
      Log::Info("Start training from score %lf", init_score);
      return init_score;
    }
  }
  return 0.0;
}

bool GBDT::TrainOneIter(const score_t* gradients, const score_t* hessians) {
  double init_score = 0.0;
  if (gradients == nullptr || hessians == nullptr) {
    if (objective_function_ == nullptr) Log::Fatal("No objective function provided");
    init_score = BoostFromAverage();
    // GBDT::Boosting, enqueued on the learner's stream so that no host synchronisation is needed before Train
    B200TreeLearner* bl = dynamic_cast<B200TreeLearner*>(tree_learner_.get());
    objective_function_->GetGradients(score_dev_, grad_dev_, hess_dev_, bl ? gbl_learner_stream(bl->handle()) : nullptr);
    if (bl == nullptr) GblCheck(gbl_device_synchronize(device_id_));
  } else {
    if (objective_function_ != nullptr) Log::Fatal("Check failed: (objective_function_) == (nullptr): pass objective=custom to use custom gradients");
    GblCheck(gbl_memcpy_h2d(device_id_, grad_dev_, gradients, sizeof(float) * num_data_));
    GblCheck(gbl_memcpy_h2d(device_id_, hess_dev_, hessians, sizeof(float) * num_data_));
  }
  Bagging(iter_);  // GBDT::TrainOneIter -> data_sample_strategy_->Bagging(iter_, ...), gbdt.cpp:383
  const bool is_first_tree = models_.empty();
  std::unique_ptr<Tree> new_tree;
  bool score_updated = false;
  if (B200TreeLearner* bl = dynamic_cast<B200TreeLearner*>(tree_learner_.get())) {
    // tree + score update enqueued back to back; the host waits once, for the split records.  A tree without
    // splits has all-zero device leaf outputs, so the update is a no-op in that case (UpdateScore is skipped for
    // such trees in the reference, gbdt.cpp:408-419).
    bl->TrainAsync(grad_dev_, hess_dev_);
    bl->AddPredictionToScoreAsync(shrinkage_rate_, score_dev_);
    bl->AddPredictionOutOfBagAsync(shrinkage_rate_, score_dev_);
    // GBDT::UpdateScore (gbdt.cpp:519-522): every validation row walks the new tree, still on the device
    for (size_t i = 0; i < valid_score_dev_.size(); ++i) {
      GblCheck(gbl_learner_add_prediction_to_dataset_dev(bl->handle(), valid_data_[i]->device_handle(), shrinkage_rate_, 0.0,
                                                         valid_score_dev_[i]));
    }
    new_tree.reset(bl->FinishTrain());
    score_updated = true;
  } else {
    new_tree.reset(tree_learner_->Train(grad_dev_, hess_dev_, is_first_tree));
  }
  bool should_continue = false;
  if (new_tree->num_leaves() > 1) {
    should_continue = true;
    new_tree->Shrinkage(shrinkage_rate_);
    if (!score_updated) tree_learner_->AddPredictionToScore(new_tree.get(), score_dev_);  // UpdateScore
    if (std::fabs(init_score) > kEpsilon) new_tree->AddBias(init_score);
  } else {
    if (models_.empty()) {
      if (objective_function_ != nullptr && !config_->boost_from_average && !has_init_score_) {
        init_score = objective_function_->BoostFromScore(0);
        GblCheck(gbl_device_fill_f64(device_id_, score_dev_, init_score, num_data_));
        for (size_t i = 0; i < valid_score_dev_.size(); ++i) {
          GblCheck(gbl_device_fill_f64(device_id_, valid_score_dev_[i], init_score, valid_data_[i]->num_data()));
        }
      }
      new_tree->AsConstantTree(init_score, num_data_);
    } else {
      new_tree->AsConstantTree(0, num_data_);
    }
  }
  models_.push_back(std::move(new_tree));
  if (!should_continue) {
    Log::Warning("Stopped training because there are no more leaves that meet the split requirements");
    if (models_.size() > 1) models_.pop_back();
    return true;
  }
  ++iter_;
  return false;
}

std::vector<double> GBDT::FeatureImportance(int num_iteration, int importance_type) const {
  int num_used_model = static_cast<int>(models_.size());
  if (num_iteration > 0) num_used_model = std::min(num_iteration, num_used_model);
  std::vector<double> imp(max_feature_idx_ + 1, 0.0);
  if (importance_type != 0 && importance_type != 1) Log::Fatal("Unknown importance type: only support split=0 and gain=1");
  for (int i = 0; i < num_used_model; ++i) {
    for (int n = 0; n < models_[i]->num_leaves() - 1; ++n) {
      if (models_[i]->split_gain(n) > 0) imp[models_[i]->split_feature(n)] += importance_type == 0 ? 1.0 : models_[i]->split_gain(n);
    }
  }
  return imp;
}

void GBDT::ResetConfig(const Config* config) {
  config_ = config;
  shrinkage_rate_ = config->learning_rate;
  if (tree_learner_ != nullptr) tree_learner_->ResetConfig(config);
  // BaggingSampleStrategy::ResetSampleConfig (bagging.hpp:139-215): unchanged fraction / frequency keep the generators
  // and the current bag; anything else starts over (generators re-seeded, next iteration re-bags)
  bool balanced = false;
  const bool enabled = BaggingEnabled(config, &balanced);
  bagging_off_ = false;  // re-decided at the next iteration
  if (enabled && bagging_ready_ && bag_fraction_used_ == config->bagging_fraction && bag_freq_used_ == config->bagging_freq &&
      bag_pos_used_ == config->pos_bagging_fraction && bag_neg_used_ == config->neg_bagging_fraction)
    return;
  const bool was_bagging = bagging_ready_;
  bagging_ready_ = false;
  bagging_rands_.clear();
  if (was_bagging && !enabled) tree_learner_->SetBaggingData(nullptr, nullptr, 0);  // back to all rows
}

int64_t GBDT::GetNumPredictAt(int data_idx) const {
  if (data_idx < 0 || data_idx > static_cast<int>(valid_data_.size())) Log::Fatal("Check failed: data_idx >= 0 && data_idx <= valid sets");
  return data_idx == 0 ? num_data_ : valid_data_[data_idx - 1]->num_data();
}

void GBDT::GetTrainScore(double* out) const {
  GblCheck(gbl_memcpy_d2h(device_id_, out, score_dev_, sizeof(double) * num_data_));
}

void GBDT::GetPredictAt(int data_idx, double* out_result, int64_t* out_len) {
  const int64_t n = GetNumPredictAt(data_idx);
  if (data_idx == 0) {
    GetTrainScore(out_result);
  } else {
    GblCheck(gbl_device_synchronize(device_id_));
    GblCheck(gbl_memcpy_d2h(device_id_, out_result, valid_score_dev_[data_idx - 1], sizeof(double) * n));
  }
  if (objective_function_ != nullptr) {
    for (int64_t i = 0; i < n; ++i) {
      double v;
      objective_function_->ConvertOutput(out_result + i, &v);
      out_result[i] = v;
    }
  }
  *out_len = n;
}

std::string GBDT::SaveModelToString(int start_iteration, int num_iteration) const {
  std::stringstream ss;
  ss << "tree" << '\n';
  ss << "version=v4" << '\n';
  ss << "num_class=1" << '\n';
  ss << "num_tree_per_iteration=1" << '\n';
  ss << "label_index=0" << '\n';
  ss << "max_feature_idx=" << max_feature_idx_ << '\n';
  if (objective_function_ != nullptr) ss << "objective=" << objective_function_->ToString() << '\n';
  auto join = [](const std::vector<std::string>& v) {
    std::string s;
    for (size_t i = 0; i < v.size(); ++i) {
      if (i) s.push_back(' ');
      s += v[i];
    }
    return s;
  };
  ss << "feature_names=" << join(train_data_->feature_names()) << '\n';
  ss << "feature_infos=" << join(train_data_->feature_infos()) << '\n';
  int num_used = static_cast<int>(models_.size());
  start_iteration = std::max(start_iteration, 0);
  start_iteration = std::min(start_iteration, num_used);
  if (num_iteration > 0) num_used = std::min(start_iteration + num_iteration, num_used);
  std::vector<std::string> tree_strs;
  std::string sizes;
  for (int i = start_iteration; i < num_used; ++i) {
    std::string t = "Tree=" + std::to_string(i - start_iteration) + '\n' + models_[i]->ToString() + '\n';
    if (!sizes.empty()) sizes.push_back(' ');
    sizes += std::to_string(t.size());
    tree_strs.push_back(std::move(t));
  }
  ss << "tree_sizes=" << sizes << '\n' << '\n';
  for (const auto& t : tree_strs) ss << t;
  ss << "end of trees" << "\n";
  // feature importances (split counts), gbdt_model_text.cpp:368-389
  std::vector<size_t> imp(max_feature_idx_ + 1, 0);
  for (int i = start_iteration; i < num_used; ++i) {
    for (int n = 0; n < models_[i]->num_leaves() - 1; ++n) imp[models_[i]->split_feature(n)] += 1;
  }
  std::vector<std::pair<size_t, std::string>> pairs;
  const auto names = train_data_->feature_names();
  for (size_t i = 0; i < imp.size(); ++i) {
    if (imp[i] > 0) pairs.emplace_back(imp[i], names[i]);
  }
  std::stable_sort(pairs.begin(), pairs.end(),
                   [](const std::pair<size_t, std::string>& a, const std::pair<size_t, std::string>& b) { return a.first > b.first; });
  ss << '\n' << "feature_importances:" << '\n';
  for (const auto& p : pairs) ss << p.second << "=" << p.first << '\n';
  ss << "\nparameters:" << '\n' << config_->ToString() << "\n" << "end of parameters" << '\n';
  return ss.str();
}

void GBDT::PredictForMat(const void* data, int data_type, int32_t nrow, int32_t ncol, int is_row_major, bool raw_score,
                         double* out) const {
  // Predictor over raw feature values (src/application/predictor.hpp), host side: prediction is outside the hot path
  if (ncol <= max_feature_idx_) Log::Fatal("The number of features in data (%d) is not the same as it was in training data (%d)", ncol, max_feature_idx_ + 1);
#pragma omp parallel for schedule(static)
  for (int32_t r = 0; r < nrow; ++r) {
    std::vector<double> row(ncol);
    for (int c = 0; c < ncol; ++c) {
      const int64_t off = is_row_major ? static_cast<int64_t>(r) * ncol + c : static_cast<int64_t>(c) * nrow + r;
      row[c] = data_type == 0 ? static_cast<double>(static_cast<const float*>(data)[off]) : static_cast<const double*>(data)[off];
    }
    double s = 0.0;
    for (const auto& t : models_) s += t->Predict(row.data());
    if (!raw_score && objective_function_ != nullptr) {
      double v;
      objective_function_->ConvertOutput(&s, &v);
      s = v;
    }
    out[r] = s;
  }
}

}  // namespace lgbm_b200
