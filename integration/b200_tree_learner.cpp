// integration/b200_tree_learner.cpp -- see b200_tree_learner.h.  Compiled against the reference's headers.
//
// What each method replaces in the reference:
//   Init / ResetTrainingData   SerialTreeLearner::Init (src/treelearner/serial_tree_learner.cpp:29-85) +
//                              Dataset::CreateCUDAColumnData (src/io/dataset.cpp:1767-): the binned columns go to HBM once
//   Train                      SerialTreeLearner::Train (:179-245) / CUDASingleGPUTreeLearner::Train
//                              (src/treelearner/cuda/cuda_single_gpu_tree_learner.cpp:158-344); the split records that come
//                              back are applied with the reference's own Tree::Split, as SplitInner does (:798-810)
//   SetBaggingData             SerialTreeLearner::SetBaggingData (:247-285)
//   AddPredictionToScore       SerialTreeLearner::AddPredictionToScore (serial_tree_learner.h:100-115)
//   RenewTreeOutput            SerialTreeLearner::RenewTreeOutput (:924-966)
#include "b200_tree_learner.h"

#include "b200_objective.h"

#include <LightGBM/bin.h>
#include <LightGBM/objective_function.h>
#include <LightGBM/utils/common.h>
#include <LightGBM/utils/log.h>
#include <LightGBM/utils/openmp_wrapper.h>

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>

namespace LightGBM {

namespace {
// GBL_DROPIN_TIMING=1: wall-clock split of the learner's share of an iteration, printed when the learner is released
struct StageClock {
  bool on = std::getenv("GBL_DROPIN_TIMING") != nullptr && std::atoi(std::getenv("GBL_DROPIN_TIMING")) != 0;
  double train_device = 0, train_host = 0, fetch = 0, add_score = 0, first_call = 0, last_return = 0;
  int trees = 0;
  bool warm = false;
  static double Now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
  void Report() const {
    if (!on || trees == 0) return;
    const double span = last_return - first_call;
    // stderr, not Log: the split is wanted at verbosity -1 too
    std::fprintf(stderr, "[B200 drop-in timing] %d trees after 3 warm-up trees, %.3f ms per iteration: Train (gradient upload unless "
                 "they are in HBM already, device tree) %.3f, Tree::Split replay %.3f, leaf fetch %.3f, host score update %.3f, "
                 "everything outside the learner (GBDT, objective: score upload + device gradients or host gradients) %.3f\n",
                 trees, 1e3 * span / trees, 1e3 * train_device / trees, 1e3 * train_host / trees, 1e3 * fetch / trees,
                 1e3 * add_score / trees, 1e3 * (span - train_device - train_host - fetch - add_score) / trees);
  }
};
StageClock g_clock;

void Check(int ret) {  // C-ABI status -> Log::Fatal, the reference's error convention (log.h:117-140)
  if (ret != 0) Log::Fatal("%s", gbl_last_error());
}
}  // namespace

B200TreeLearner::B200TreeLearner(const Config* config, int num_gpu) : config_(config), num_gpu_(std::max(1, num_gpu)) {
  CheckConfig();
}

B200TreeLearner::~B200TreeLearner() { Release(); }

void B200TreeLearner::Release() {
  g_clock.Report();
  g_clock.trees = 0;
  if (announced_) DeviceGradients::RemoveConsumer(layout_);
  announced_ = false;
  for (const score_t* p : registered_) gbl_host_unregister(const_cast<score_t*>(p));
  registered_.clear();
  for (std::vector<uint16_t>& b : leaf16_buf_) {
    if (!b.empty()) gbl_host_unregister(b.data());
    b.clear();
  }
  for (Shard& s : shards_) {
    if (s.learner != nullptr) gbl_learner_free(s.learner);
    s.learner = nullptr;
  }
  for (Shard& s : shards_) {
    if (s.comm != nullptr) gbl_comm_free(s.comm);
    if (s.dataset != nullptr) gbl_dataset_free(s.dataset);
    s.comm = nullptr;
    s.dataset = nullptr;
  }
  shards_.clear();
}

// Everything the device learner does not implement is refused here, loudly, instead of being trained differently.
void B200TreeLearner::CheckConfig() const {
  if (config_->linear_tree) Log::Fatal("B200 tree learner: linear_tree is not supported");
  // monotone constraints: the reference's default method ("basic": one output interval per leaf) is implemented on the
  // device; "intermediate" / "advanced" re-tighten OTHER leaves after a split and are not
  if (!config_->monotone_constraints.empty() && config_->monotone_constraints_method != std::string("basic"))
    Log::Fatal("B200 tree learner: monotone_constraints_method=%s is not supported (basic is)",
               config_->monotone_constraints_method.c_str());
  if (config_->extra_trees) Log::Fatal("B200 tree learner: extra_trees is not supported");
  if (config_->feature_fraction_bynode < 1.0) Log::Fatal("B200 tree learner: feature_fraction_bynode < 1 is not supported");
  if (!config_->interaction_constraints.empty()) Log::Fatal("B200 tree learner: interaction_constraints are not supported");
  if (config_->cegb_tradeoff < 1.0 || config_->cegb_penalty_split > 0.0 || !config_->cegb_penalty_feature_lazy.empty() ||
      !config_->cegb_penalty_feature_coupled.empty())
    Log::Fatal("B200 tree learner: cost-effective gradient boosting (cegb_*) is not supported");
  if (!config_->forcedsplits_filename.empty()) Log::Fatal("B200 tree learner: forced splits are not supported");
}

gbl_learner_config B200TreeLearner::MakeConfig() const {
  gbl_learner_config c;
  std::memset(&c, 0, sizeof(c));
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

template <typename F>
void B200TreeLearner::ForEachShard(F fn) const {
  if (shards_.size() == 1) {
    fn(shards_[0]);
    return;
  }
  // one host thread per GPU: the data-parallel learner's calls are collective (NCCL set-up, peer-memory barriers)
  std::vector<std::thread> threads;
  for (Shard& s : shards_) {
    s.error.clear();
    threads.emplace_back([&s, &fn]() {
      try {
        fn(s);
      } catch (const std::exception& ex) {
        s.error = ex.what();
      }
    });
  }
  for (std::thread& t : threads) t.join();
  for (const Shard& s : shards_) {
    if (!s.error.empty()) Log::Fatal("%s", s.error.c_str());
  }
}

// The binned training data: one byte per cell (BinIterator::Get returns the mapper bin, what the device layout stores),
// one column per used feature, cut into one contiguous row slice per GPU.
void B200TreeLearner::BuildDeviceData() {
  const int nf = train_data_->num_features();
  if (nf <= 0) Log::Fatal("B200 tree learner: the dataset has no usable feature");
  std::vector<gbl_feature_meta> metas(nf);
  for (int f = 0; f < nf; ++f) {
    const BinMapper* m = train_data_->FeatureBinMapper(f);
    const bool categorical = m->bin_type() == BinType::CategoricalBin;
    if (m->num_bin() > 256) Log::Fatal("B200 tree learner: feature %d has %d bins; at most 256 (max_bin <= 255) are supported",
                                       train_data_->RealFeatureIndex(f), m->num_bin());
    metas[f].num_bin = m->num_bin();
    metas[f].missing_type = static_cast<int>(m->missing_type());
    metas[f].default_bin = static_cast<int>(m->GetDefaultBin());
    metas[f].most_freq_bin = static_cast<int>(m->GetMostFreqBin());
    metas[f].real_index = train_data_->RealFeatureIndex(f);
    // FeatureMetainfo::monotone_type (feature_histogram.hpp:1440-1444): indexed by the ORIGINAL column
    metas[f].monotone_type = (config_->monotone_constraints.empty() || categorical)
                                 ? 0
                                 : static_cast<int>(config_->monotone_constraints[train_data_->RealFeatureIndex(f)]);
    metas[f].bin_type = categorical ? 1 : 0;
    // feature_contri (FeatureMetainfo::penalty, feature_histogram.hpp:1435-1439)
    const int real = train_data_->RealFeatureIndex(f);
    metas[f].penalty = config_->feature_contri.empty() ? 1.0 : config_->feature_contri[real];
  }
  std::vector<std::vector<uint8_t>> cols(nf, std::vector<uint8_t>(num_data_));
#pragma omp parallel for num_threads(OMP_NUM_THREADS()) schedule(dynamic)
  for (int f = 0; f < nf; ++f) {
    std::unique_ptr<BinIterator> it(train_data_->FeatureIterator(f));
    it->Reset(0);
    uint8_t* out = cols[f].data();
    for (data_size_t i = 0; i < num_data_; ++i) out[i] = static_cast<uint8_t>(it->Get(i));
  }
  const int R = num_gpu_;
  shards_.resize(R);
  const data_size_t per = num_data_ / R;
  for (int r = 0; r < R; ++r) {
    Shard& s = shards_[r];
    s.device = R == 1 ? (config_->gpu_device_id < 0 ? 0 : config_->gpu_device_id) : r;
    s.begin = r * per;
    s.count = r == R - 1 ? num_data_ - s.begin : per;
    if (s.count <= 0) Log::Fatal("B200 tree learner: %d rows cannot be sharded over %d GPUs", num_data_, R);
  }
  ForEachShard([&](Shard& s) {
    std::vector<const uint8_t*> ptrs(nf);
    for (int f = 0; f < nf; ++f) ptrs[f] = cols[f].data() + s.begin;
    if (gbl_dataset_create_from_bins(metas.data(), nf, s.count, ptrs.data(), s.device, &s.dataset) != 0)
      throw std::runtime_error(gbl_last_error());
  });
}

void B200TreeLearner::CreateLearners() {
  const gbl_learner_config c = MakeConfig();
  const int R = static_cast<int>(shards_.size());
  uint8_t id[128];
  if (R > 1) Check(gbl_comm_get_unique_id(id));
  ForEachShard([&](Shard& s) {
    if (R > 1 && s.comm == nullptr) {
      const int rank = static_cast<int>(&s - shards_.data());
      if (gbl_comm_create(id, R, rank, s.device, &s.comm) != 0) throw std::runtime_error(gbl_last_error());
    }
    if (gbl_learner_create(s.dataset, &c, s.comm, &s.learner) != 0) throw std::runtime_error(gbl_last_error());
    s.records.resize(config_->num_leaves);
  });
  PushMonotone(true);
}

void B200TreeLearner::InitQuantRandom() {
  // GradientDiscretizer::Init (gradient_discretizer.cpp:16-41): one mt19937 pair per block of rows, seeded seed + block
  // and seed + block + num_threads, so the values depend on the thread count exactly like the reference's
  quant_random_ready_ = false;
  if (!(config_->use_quantized_grad && config_->stochastic_rounding)) return;
  if (shards_.size() > 1)
    Log::Warning("stochastic_rounding with num_gpu > 1: every GPU rotates its own slice of the random arrays, so the "
                 "discretised gradients differ from a single-GPU run (set stochastic_rounding=false for identical trees)");
  const data_size_t n = num_data_;
  const int num_threads = OMP_NUM_THREADS();
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
  for (Shard& s : shards_) Check(gbl_learner_set_quant_random(s.learner, rg.data() + s.begin, rh.data() + s.begin, s.count));
  quant_start_engine_ = std::mt19937(config_->seed);
  quant_start_dist_ = std::uniform_int_distribution<data_size_t>(0, n);
  quant_random_ready_ = true;
}

// `!monotone_constraints.empty()` selects the reference's USE_MC code path for every feature (feature_histogram.hpp:246-260)
void B200TreeLearner::PushMonotone(bool force) {
  const bool on = !config_->monotone_constraints.empty();
  const bool same_cat = config_->max_cat_threshold == cat_params_[0] && config_->max_cat_to_onehot == cat_params_[1] &&
                        config_->min_data_per_group == cat_params_[2] && config_->cat_l2 == cat_l2_ && config_->cat_smooth == cat_smooth_;
  if (!force && on == monotone_on_ && config_->monotone_penalty == monotone_penalty_ && same_cat) return;
  cat_params_[0] = config_->max_cat_threshold;
  cat_params_[1] = config_->max_cat_to_onehot;
  cat_params_[2] = config_->min_data_per_group;
  cat_l2_ = config_->cat_l2;
  cat_smooth_ = config_->cat_smooth;
  monotone_on_ = on;
  monotone_penalty_ = config_->monotone_penalty;
  for (Shard& s : shards_) {
    Check(gbl_learner_set_monotone(s.learner, on ? 1 : 0, monotone_penalty_));
    // categorical split search parameters (config.h:480-501)
    Check(gbl_learner_set_categorical(s.learner, config_->max_cat_threshold, config_->cat_l2, config_->cat_smooth,
                                      config_->max_cat_to_onehot, config_->min_data_per_group));
  }
}

void B200TreeLearner::PushFeatureMask() {
  const bool sampled = config_->feature_fraction < 1.0;
  for (Shard& s : shards_)
    Check(gbl_learner_set_feature_mask(s.learner, sampled ? col_sampler_->is_feature_used_bytree().data() : nullptr));
}

void B200TreeLearner::Init(const Dataset* train_data, bool is_constant_hessian) {
  Release();
  train_data_ = train_data;
  num_data_ = train_data->num_data();
  is_constant_hessian_ = is_constant_hessian;
  bagging_ = false;
  leaf_assignment_valid_ = false;
  BuildDeviceData();
  CreateLearners();
  InitQuantRandom();
  col_sampler_.reset(new ColSampler(config_));
  col_sampler_->SetTrainingData(train_data_);  // serial_tree_learner.cpp:59
  PushFeatureMask();
  layout_ = DeviceLayout();
  layout_.num_data = num_data_;
  if (shards_.size() <= static_cast<size_t>(DeviceLayout::kMaxParts)) {
    layout_.num_parts = static_cast<int>(shards_.size());
    for (size_t r = 0; r < shards_.size(); ++r) {
      layout_.begin[r] = shards_[r].begin;
      layout_.count[r] = shards_[r].count;
      layout_.device[r] = shards_[r].device;
    }
    DeviceGradients::AddConsumer(layout_);
    announced_ = true;
  }
  Log::Info("B200 tree learner: %d rows x %d features on %d GPU%s", num_data_, train_data_->num_features(),
            static_cast<int>(shards_.size()), shards_.size() > 1 ? "s (data-parallel, rows sharded)" : "");
}

void B200TreeLearner::ResetTrainingData(const Dataset* train_data, bool is_constant_hessian) {
  Init(train_data, is_constant_hessian);
}

void B200TreeLearner::ResetIsConstantHessian(bool is_constant_hessian) {
  if (is_constant_hessian == is_constant_hessian_) return;
  is_constant_hessian_ = is_constant_hessian;
  const gbl_learner_config c = MakeConfig();
  ForEachShard([&](Shard& s) {
    if (gbl_learner_reset_config(s.learner, &c) != 0) throw std::runtime_error(gbl_last_error());
  });
}

void B200TreeLearner::ResetConfig(const Config* config) {
  const gbl_learner_config before = MakeConfig();
  config_ = config;
  CheckConfig();
  const gbl_learner_config c = MakeConfig();
  // a learning-rate schedule resets the parameters before every iteration: the device learner (which drops its
  // captured tree on a reset) is only touched when a field it reads has changed
  if (std::memcmp(&before, &c, sizeof(c)) != 0) {
    ForEachShard([&](Shard& s) {
      if (gbl_learner_reset_config(s.learner, &c) != 0) throw std::runtime_error(gbl_last_error());
      s.records.resize(config_->num_leaves);
    });
    InitQuantRandom();
  }
  PushMonotone(false);
  if (col_sampler_ != nullptr) {
    col_sampler_->SetConfig(config_);  // serial_tree_learner.cpp:168
    PushFeatureMask();
  }
}

void B200TreeLearner::SetForcedSplit(const Json* forced_split_json) {
  if (forced_split_json != nullptr && !forced_split_json->is_null())
    Log::Fatal("B200 tree learner: forced splits are not supported");
}

Tree* B200TreeLearner::Train(const score_t* gradients, const score_t* hessians, bool /*is_first_tree*/) {
  // SerialTreeLearner::BeforeTrain: column sampling by tree (serial_tree_learner.cpp:293)
  const double t_begin = StageClock::Now();
  if (g_clock.trees == 0) g_clock.first_call = t_begin;
  col_sampler_->ResetByTree();
  if (config_->feature_fraction < 1.0) PushFeatureMask();
  int32_t quant_start = 0;
  if (quant_random_ready_) quant_start = quant_start_dist_(quant_start_engine_);  // gradient_discretizer.cpp:116
  // GBDT keeps its gradient / hessian vectors for the whole training (gbdt.cpp:849-865): page-lock them the first time
  // they are seen so that the per-tree upload runs at the PCIe rate instead of through a staging buffer
  if (packed_gradients_ && bagging_) {
    // position i of the packed arrays belongs to row bag_rows_[i]; rows outside the bag are never read by the learner
    if (spread_gradients_.size() != static_cast<size_t>(num_data_)) {
      spread_gradients_.assign(num_data_, 0.0f);
      spread_hessians_.assign(num_data_, 0.0f);
    }
    const data_size_t cnt = static_cast<data_size_t>(bag_rows_.size());
#pragma omp parallel for num_threads(OMP_NUM_THREADS()) schedule(static)
    for (data_size_t i = 0; i < cnt; ++i) {
      spread_gradients_[bag_rows_[i]] = gradients[i];
      spread_hessians_[bag_rows_[i]] = hessians[i];
    }
    gradients = spread_gradients_.data();
    hessians = spread_hessians_.data();
  }
  // gradients the B200 objective of this library left in HBM (b200_objective.h): train from them, no upload
  DeviceGradients on_device;
  const bool from_device =
      announced_ && DeviceGradients::Take(gradients, hessians, layout_, !(packed_gradients_ && bagging_), &on_device);
  if (!from_device) {
    for (const score_t* p : {gradients, hessians}) {
      if (registered_.insert(p).second && registered_.size() <= 256)
        gbl_host_register(const_cast<score_t*>(p), sizeof(score_t) * static_cast<size_t>(num_data_));
    }
  }
  ForEachShard([&](Shard& s) {
    // one GPU: the reference's own offset; shards rotate inside their slice (see the warning in InitQuantRandom)
    const int32_t start = shards_.size() == 1 ? quant_start : quant_start % std::max<data_size_t>(1, s.count);
    if (quant_random_ready_ && gbl_learner_set_quant_random_start(s.learner, start) != 0) throw std::runtime_error(gbl_last_error());
    if (from_device) {
      const int part = static_cast<int>(&s - shards_.data());
      if (gbl_learner_train(s.learner, on_device.dev_gradients[part], on_device.dev_hessians[part], &s.result,
                            s.records.data()) != 0)
        throw std::runtime_error(gbl_last_error());
      return;
    }
    // host gradients of this shard's rows -> HBM, one tree, split records back
    if (gbl_learner_train_host(s.learner, gradients + s.begin, hessians + s.begin, &s.result, s.records.data()) != 0)
      throw std::runtime_error(gbl_last_error());
  });
  leaf_assignment_valid_ = false;
  const double t_device = StageClock::Now();
  g_clock.train_device += t_device - t_begin;
  const Shard& s0 = shards_[0];  // every rank of the data-parallel learner returns the same records (global counts)
  std::unique_ptr<Tree> tree(new Tree(config_->num_leaves, false, false));
  tree->SetLeafOutput(0, s0.result.root_output);  // serial_tree_learner.cpp:205-208
  for (int k = 0; k < s0.result.num_splits; ++k) {
    const gbl_split_record& r = s0.records[k];
    if (r.num_cat_bins > 0) {
      // SerialTreeLearner::SplitInner, categorical branch (serial_tree_learner.cpp:811-848): the bins that go left as a
      // bitset over inner bins, and the same set translated to category values (Dataset::RealThreshold)
      std::vector<uint32_t> bins;
      for (uint32_t b = 0; b < 256; ++b) {
        if ((r.cat_bitset[b >> 5] >> (b & 31)) & 1u) bins.push_back(b);
      }
      std::vector<uint32_t> cat_bitset_inner = Common::ConstructBitset(bins.data(), static_cast<int>(bins.size()));
      std::vector<int> threshold_int(bins.size());
      for (size_t i = 0; i < bins.size(); ++i) threshold_int[i] = static_cast<int>(train_data_->RealThreshold(r.feature, bins[i]));
      std::vector<uint32_t> cat_bitset = Common::ConstructBitset(threshold_int.data(), static_cast<int>(threshold_int.size()));
      tree->SplitCategorical(r.leaf, r.feature, r.real_feature, cat_bitset_inner.data(), static_cast<int>(cat_bitset_inner.size()),
                             cat_bitset.data(), static_cast<int>(cat_bitset.size()), r.left_output, r.right_output, r.left_count,
                             r.right_count, r.left_sum_hessian, r.right_sum_hessian,
                             static_cast<float>(r.gain + config_->min_gain_to_split),
                             train_data_->FeatureBinMapper(r.feature)->missing_type());
      continue;
    }
    // SerialTreeLearner::SplitInner -> Tree::Split (serial_tree_learner.cpp:798-810)
    tree->Split(r.leaf, r.feature, r.real_feature, r.threshold_bin, train_data_->RealThreshold(r.feature, r.threshold_bin),
                r.left_output, r.right_output, r.left_count, r.right_count, r.left_sum_hessian, r.right_sum_hessian,
                static_cast<float>(r.gain + config_->min_gain_to_split),
                train_data_->FeatureBinMapper(r.feature)->missing_type(), r.default_left != 0);
  }
  if (config_->use_quantized_grad && config_->quant_train_renew_leaf && s0.result.num_splits > 0) {
    // serial_tree_learner.cpp:238-241: GradientDiscretizer::RenewIntGradTreeOutput ran on the device
    std::vector<double> values(s0.result.num_splits + 1);
    Check(gbl_learner_get_leaf_values(s0.learner, values.data(), s0.result.num_splits + 1));
    for (int leaf = 0; leaf <= s0.result.num_splits; ++leaf) tree->SetLeafOutput(leaf, values[leaf]);
  }
  if (s0.result.num_splits == 0) Log::Warning("No further splits with positive gain, best gain: -inf");
  g_clock.train_host += StageClock::Now() - t_device;
  return tree.release();
}

Tree* B200TreeLearner::FitByExistingTree(const Tree*, const score_t*, const score_t*) const {
  Log::Fatal("B200 tree learner: FitByExistingTree (refit) is not supported");
  return nullptr;
}

Tree* B200TreeLearner::FitByExistingTree(const Tree*, const std::vector<int>&, const score_t*, const score_t*) const {
  Log::Fatal("B200 tree learner: FitByExistingTree (refit) is not supported");
  return nullptr;
}

void B200TreeLearner::SetBaggingData(const Dataset* subset, const data_size_t* used_indices, data_size_t num_data) {
  // The reference may hand over a row-subset copy of the dataset (bagging_fraction <= 0.5, bagging.hpp:118-133); the
  // device keeps the full binned data, so the index list alone defines the bag -- same rows, same trees.
  leaf_assignment_valid_ = false;
  bagging_ = used_indices != nullptr && num_data > 0 && num_data < num_data_;
  packed_gradients_ = bagging_ && subset != nullptr;
  if (packed_gradients_) bag_rows_.assign(used_indices, used_indices + num_data);
  if (!bagging_) {
    for (Shard& s : shards_) Check(gbl_learner_set_bagging_data(s.learner, nullptr, 0));
    return;
  }
  // ascending global row ids (bagging.hpp / goss.hpp fill them in row order) -> per-shard local ids
  const data_size_t* p = used_indices;
  const data_size_t* end = used_indices + num_data;
  for (Shard& s : shards_) {
    const data_size_t* q = std::lower_bound(p, end, s.begin + s.count);
    std::vector<data_size_t> local(p, q);
    for (data_size_t& v : local) v -= s.begin;
    Check(gbl_learner_set_bagging_data(s.learner, local.data(), static_cast<data_size_t>(local.size())));
    p = q;
  }
}

void B200TreeLearner::FetchLeafAssignment() const {
  if (leaf_assignment_valid_) return;
  if (config_->num_leaves <= 65535) {
    // 2-byte leaf ids into a page-locked buffer: a quarter of the bytes of the int32 form at the full PCIe rate
    leaf16_turn_ ^= 1;
    std::vector<uint16_t>& buf = leaf16_buf_[leaf16_turn_];
    if (buf.size() != static_cast<size_t>(num_data_)) {
      if (!buf.empty()) gbl_host_unregister(buf.data());
      buf.assign(num_data_, 0xFFFF);
      gbl_host_register(buf.data(), sizeof(uint16_t) * buf.size());  // failure = slower copies, nothing else
    }
    ForEachShard([&](Shard& s) {
      if (gbl_learner_get_leaf_assignment_u16(s.learner, buf.data() + s.begin) != 0) throw std::runtime_error(gbl_last_error());
    });
  } else {
    leaf_of_row_.resize(num_data_);
    ForEachShard([&](Shard& s) {
      if (gbl_learner_get_leaf_assignment(s.learner, leaf_of_row_.data() + s.begin) != 0) throw std::runtime_error(gbl_last_error());
    });
  }
  leaf_assignment_valid_ = true;
}

void B200TreeLearner::AddPredictionToScore(const Tree* tree, double* out_score) const {
  if (tree->num_leaves() <= 1) return;  // serial_tree_learner.h:101-103
  // out_score is the host ScoreUpdater's array in this build: leaf of every trained row from the device partition,
  // then score[row] += leaf_output (rows outside the bag carry -1; GBDT::UpdateScore walks the tree for them itself)
  const double t_begin = StageClock::Now();
  FetchLeafAssignment();
  const double t_fetch = StageClock::Now();
  g_clock.fetch += t_fetch - t_begin;
  if (config_->num_leaves <= 65535) {
    const uint16_t* leaf = leaf16().data();
#pragma omp parallel for num_threads(OMP_NUM_THREADS()) schedule(static)
    for (data_size_t i = 0; i < num_data_; ++i) {
      if (leaf[i] != 0xFFFF) out_score[i] += tree->LeafOutput(leaf[i]);
    }
  } else {
    const int32_t* leaf = leaf_of_row_.data();
#pragma omp parallel for num_threads(OMP_NUM_THREADS()) schedule(static)
    for (data_size_t i = 0; i < num_data_; ++i) {
      if (leaf[i] >= 0) out_score[i] += tree->LeafOutput(leaf[i]);
    }
  }
  g_clock.last_return = StageClock::Now();
  g_clock.add_score += g_clock.last_return - t_fetch;
  g_clock.trees += 1;
  if (!g_clock.warm && g_clock.trees == 3) {  // graph capture, page-locking and first-touch costs stay out of the split
    g_clock.warm = true;
    g_clock.trees = 0;
    g_clock.train_device = g_clock.train_host = g_clock.fetch = g_clock.add_score = 0;
  }
}

void B200TreeLearner::RenewTreeOutput(Tree* tree, const ObjectiveFunction* obj,
                                      std::function<double(const label_t*, int)> residual_getter, data_size_t total_num_data,
                                      const data_size_t* bag_indices, data_size_t bag_cnt, const double* /*train_score*/) const {
  if (obj == nullptr || !obj->IsRenewTreeOutput()) return;
  // SerialTreeLearner::RenewTreeOutput (serial_tree_learner.cpp:924-966): per leaf, the rows of the leaf
  FetchLeafAssignment();
  const int nl = tree->num_leaves();
  if (config_->num_leaves <= 65535) {  // one representation below: widen the 2-byte ids
    leaf_of_row_.resize(num_data_);
    for (data_size_t i = 0; i < num_data_; ++i) leaf_of_row_[i] = leaf16()[i] == 0xFFFF ? -1 : static_cast<int32_t>(leaf16()[i]);
  }
  std::vector<data_size_t> begin(nl + 1, 0);
  for (data_size_t i = 0; i < num_data_; ++i) {
    if (leaf_of_row_[i] >= 0 && leaf_of_row_[i] < nl) ++begin[leaf_of_row_[i] + 1];
  }
  for (int l = 0; l < nl; ++l) begin[l + 1] += begin[l];
  std::vector<data_size_t> rows(begin[nl]);
  std::vector<data_size_t> fill(begin.begin(), begin.end() - 1);
  for (data_size_t i = 0; i < num_data_; ++i) {
    const int l = leaf_of_row_[i];
    if (l >= 0 && l < nl) rows[fill[l]++] = i;  // ascending row order inside every leaf, like DataPartition
  }
  const data_size_t* bag_mapper = nullptr;
  if (total_num_data != num_data_) {
    CHECK_EQ(bag_cnt, num_data_);
    bag_mapper = bag_indices;
  }
#pragma omp parallel for num_threads(OMP_NUM_THREADS()) schedule(static)
  for (int l = 0; l < nl; ++l) {
    const data_size_t cnt = begin[l + 1] - begin[l];
    if (cnt > 0) {
      const double out = obj->RenewTreeOutput(static_cast<double>(tree->LeafOutput(l)), residual_getter, rows.data() + begin[l],
                                              bag_mapper, cnt);
      tree->SetLeafOutput(l, out);
    }
  }
}

}  // namespace LightGBM
