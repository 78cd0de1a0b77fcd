// integration/b200_objective.cpp -- see b200_objective.h.  Compiled against the reference's headers.
#include "b200_objective.h"

#include <LightGBM/dataset.h>
#include <LightGBM/utils/log.h>

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <vector>

namespace LightGBM {

// The reference's own factory (src/objective/objective_function.cpp:19-113 and :115-158), compiled unchanged by
// integration/Makefile with -DCreateObjectiveFunction=CreateHostObjectiveFunction: the two static member functions keep
// their code and get these symbol names.  They are static members, i.e. plain functions, declared here by symbol.
ObjectiveFunction* CreateHostObjective(const std::string& type, const Config& config) __asm__(
    "_ZN8LightGBM17ObjectiveFunction27CreateHostObjectiveFunctionERKNSt7__cxx1112basic_stringIcSt11char_traitsIcESaIcEEERKNS_6ConfigE");
ObjectiveFunction* CreateHostObjectiveFromString(const std::string& str) __asm__(
    "_ZN8LightGBM17ObjectiveFunction27CreateHostObjectiveFunctionERKNSt7__cxx1112basic_stringIcSt11char_traitsIcESaIcEEE");

namespace {

void Check(int ret) {
  if (ret != 0) Log::Fatal("%s", gbl_last_error());
}

bool EnvFlag(const char* name, bool fallback) {
  const char* e = std::getenv(name);
  return e == nullptr ? fallback : std::atoi(e) != 0;
}

// rank_objective.hpp:22-41 / :182-201, the order of the reference's target list (= gbl_objective_config.lambdarank_target)
const char* const kRankTargets[] = {
    "ndcg", "lambdaloss-ndcg", "lambdaloss-ndcg-plus-plus", "bndcg", "lambdaloss-bndcg", "lambdaloss-bndcg-plus-plus",
    "precision", "arpk", "lambdaloss-arp1", "lambdaloss-arp2", "ranknet", "bin-ranknet", "lambdagap-s", "lambdagap-x",
    "lambdagap-s-plus", "lambdagap-x-plus", "lambdagap-s-plus-plus", "lambdagap-x-plus-plus"};

int RankTargetIndex(const std::string& name) {
  for (int i = 0; i < 18; ++i) {
    if (name == kRankTargets[i]) return i;
  }
  return -1;
}

std::mutex g_registry_mutex;
std::vector<DeviceGradients> g_registry;

}  // namespace

// ------------------------------------------------------------------------------------------- DeviceGradients
void DeviceGradients::Publish(const DeviceGradients& entry) {
  std::lock_guard<std::mutex> lock(g_registry_mutex);
  for (DeviceGradients& e : g_registry) {
    if (e.host_gradients == entry.host_gradients) {
      e = entry;
      return;
    }
  }
  g_registry.push_back(entry);
}

void DeviceGradients::Withdraw(const score_t* host_gradients) {
  std::lock_guard<std::mutex> lock(g_registry_mutex);
  g_registry.erase(std::remove_if(g_registry.begin(), g_registry.end(),
                                  [&](const DeviceGradients& e) { return e.host_gradients == host_gradients; }),
                   g_registry.end());
}

bool DeviceGradients::Take(const score_t* host_gradients, const score_t* host_hessians, data_size_t num_data, int device_id,
                           DeviceGradients* out) {
  std::lock_guard<std::mutex> lock(g_registry_mutex);
  for (DeviceGradients& e : g_registry) {
    if (e.host_gradients != host_gradients) continue;
    if (!e.fresh || e.host_hessians != host_hessians || e.num_data != num_data || e.device_id != device_id) return false;
    e.fresh = false;
    *out = e;
    return true;
  }
  return false;
}

// ------------------------------------------------------------------------------------------- factory
bool B200Objective::Supported(const std::string& type, const Config& config) {
  if (config.device_type != "cuda") return false;
  if (!EnvFlag("GBL_DROPIN_DEVICE_OBJECTIVE", true)) return false;
  // the reference keeps these on the host as well (objective_function.cpp:22-24): GOSS reads and rescales the gradient
  // vectors between GetGradients and Train; random forest computes them once
  if (config.data_sample_strategy == "goss" || config.boosting == "rf") return false;
  if (type == "regression") return !config.reg_sqrt;
  if (type == "binary") return true;
  if (type == "lambdarank") return RankTargetIndex(config.lambdarank_target) >= 0;
  return false;
}

ObjectiveFunction* ObjectiveFunction::CreateObjectiveFunction(const std::string& type, const Config& config) {
  ObjectiveFunction* host = CreateHostObjective(type, config);
  if (host == nullptr || !B200Objective::Supported(type, config)) return host;
  return new B200Objective(host, type, config);
}

// model files: prediction only, nothing to compute on the device
ObjectiveFunction* ObjectiveFunction::CreateObjectiveFunction(const std::string& str) {
  return CreateHostObjectiveFromString(str);
}

// ------------------------------------------------------------------------------------------- B200Objective
B200Objective::B200Objective(ObjectiveFunction* host, const std::string& type, const Config& config)
    : host_(host), type_(type), config_(config) {
  device_id_ = std::max(0, config.gpu_device_id);
  // with num_gpu > 1 the learner's shards on the other GPUs read the host vectors
  fill_host_ = config.num_gpu > 1 || EnvFlag("GBL_DROPIN_FILL_HOST_GRADIENTS", false);
}

B200Objective::~B200Objective() { ReleaseDevice(); }

void B200Objective::ReleaseDevice() {
  if (published_ != nullptr) DeviceGradients::Withdraw(published_);
  published_ = nullptr;
  if (registered_score_ != nullptr) gbl_host_unregister(const_cast<double*>(registered_score_));
  registered_score_ = nullptr;
  if (dev_ != nullptr) gbl_objective_free(dev_);
  dev_ = nullptr;
  for (void* p : {static_cast<void*>(score_dev_), static_cast<void*>(gradients_dev_), static_cast<void*>(hessians_dev_)}) {
    if (p != nullptr) gbl_device_free(device_id_, p);
  }
  score_dev_ = nullptr;
  gradients_dev_ = nullptr;
  hessians_dev_ = nullptr;
  on_device_ = false;
}

void B200Objective::Init(const Metadata& metadata, data_size_t num_data) {
  host_->Init(metadata, num_data);  // the reference's checks, label statistics, BoostFromScore inputs
  ReleaseDevice();
  num_data_ = num_data;
  // position-debiased lambdarank keeps per-position bias factors that GetGradients updates (rank_objective.hpp:91-120,
  // 123-172): host only
  if (type_ == "lambdarank" && (metadata.positions() != nullptr || metadata.query_boundaries() == nullptr)) return;
  if (num_data <= 0 || metadata.label() == nullptr) return;
  gbl_objective_config c;
  std::memset(&c, 0, sizeof(c));
  c.kind = type_ == "regression" ? GBL_OBJ_REGRESSION_L2 : (type_ == "binary" ? GBL_OBJ_BINARY : GBL_OBJ_LAMBDARANK);
  c.lambdarank_target = std::max(0, RankTargetIndex(config_.lambdarank_target));
  c.lambdarank_truncation_level = config_.lambdarank_truncation_level;
  c.lambdarank_norm = config_.lambdarank_norm ? 1 : 0;
  c.sigmoid = config_.sigmoid;
  c.lambdagap_weight = config_.lambdagap_weight;
  c.is_unbalance = config_.is_unbalance ? 1 : 0;
  c.scale_pos_weight = config_.scale_pos_weight;
  c.num_label_gain = static_cast<int>(config_.label_gain.size());
  c.label_gain = config_.label_gain.empty() ? nullptr : config_.label_gain.data();
  Check(gbl_objective_create(&c, num_data, metadata.label(), metadata.weights(),
                             type_ == "lambdarank" ? metadata.query_boundaries() : nullptr,
                             type_ == "lambdarank" ? metadata.num_queries() : 0, device_id_, &dev_));
  const size_t n = static_cast<size_t>(num_data);
  Check(gbl_device_malloc(device_id_, n * sizeof(double), reinterpret_cast<void**>(&score_dev_)));
  Check(gbl_device_malloc(device_id_, n * sizeof(float), reinterpret_cast<void**>(&gradients_dev_)));
  Check(gbl_device_malloc(device_id_, n * sizeof(float), reinterpret_cast<void**>(&hessians_dev_)));
  on_device_ = true;
}

void B200Objective::GetGradients(const double* score, score_t* gradients, score_t* hessians) const {
  if (!on_device_) {
    host_->GetGradients(score, gradients, hessians);
    return;
  }
  const size_t n = static_cast<size_t>(num_data_);
  // GBDT's ScoreUpdater keeps one score array for the whole training (score_updater.hpp:24-48): page-lock it once
  if (registered_score_ != score) {
    if (registered_score_ != nullptr) gbl_host_unregister(const_cast<double*>(registered_score_));
    gbl_host_register(const_cast<double*>(score), n * sizeof(double));  // failure = slower copies, nothing else
    registered_score_ = score;
  }
  Check(gbl_memcpy_h2d(device_id_, score_dev_, score, n * sizeof(double)));
  Check(gbl_objective_get_gradients(dev_, score_dev_, gradients_dev_, hessians_dev_, nullptr));
  if (fill_host_) {
    Check(gbl_memcpy_d2h(device_id_, gradients, gradients_dev_, n * sizeof(float)));
    Check(gbl_memcpy_d2h(device_id_, hessians, hessians_dev_, n * sizeof(float)));
  } else {
    Check(gbl_device_synchronize(device_id_));  // the learner trains on its own stream
  }
  if (published_ != nullptr && published_ != gradients) DeviceGradients::Withdraw(published_);
  DeviceGradients entry;
  entry.host_gradients = gradients;
  entry.host_hessians = hessians;
  entry.dev_gradients = gradients_dev_;
  entry.dev_hessians = hessians_dev_;
  entry.num_data = num_data_;
  entry.device_id = device_id_;
  entry.fresh = true;
  DeviceGradients::Publish(entry);
  published_ = gradients;
}

// bagging by query computes the gradients of the sampled queries only (gbdt.cpp:236-243): the reference's code, on the host
void B200Objective::GetGradients(const double* score, const data_size_t num_sampled_queries,
                                 const data_size_t* sampled_query_indices, score_t* gradients, score_t* hessians) const {
  if (published_ != nullptr) DeviceGradients::Withdraw(published_);
  published_ = nullptr;
  host_->GetGradients(score, num_sampled_queries, sampled_query_indices, gradients, hessians);
}

}  // namespace LightGBM
