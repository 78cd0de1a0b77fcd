// integration/b200_objective.cpp -- see b200_objective.h.  Compiled against the reference's headers.
#include "b200_objective.h"

#include <LightGBM/dataset.h>
#include <LightGBM/utils/log.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <stdexcept>
#include <string>
#include <thread>
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
std::vector<DeviceLayout> g_consumers;

}  // namespace

// ------------------------------------------------------------------------------------------- DeviceGradients
bool DeviceLayout::operator==(const DeviceLayout& o) const {
  if (num_data != o.num_data || num_parts != o.num_parts) return false;
  for (int p = 0; p < num_parts; ++p) {
    if (begin[p] != o.begin[p] || count[p] != o.count[p] || device[p] != o.device[p]) return false;
  }
  return true;
}

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

bool DeviceGradients::Take(const score_t* host_gradients, const score_t* host_hessians, const DeviceLayout& layout,
                           bool usable, DeviceGradients* out) {
  std::lock_guard<std::mutex> lock(g_registry_mutex);
  for (DeviceGradients& e : g_registry) {
    // which class block of the entry's vectors do the pointers address?
    const data_size_t n = e.layout.num_data;
    if (host_gradients < e.host_gradients || n <= 0) continue;
    const size_t off = static_cast<size_t>(host_gradients - e.host_gradients);
    if (off % static_cast<size_t>(n) != 0 || off / static_cast<size_t>(n) >= static_cast<size_t>(e.num_class)) continue;
    const int k = static_cast<int>(off / static_cast<size_t>(n));
    if (((e.fresh >> k) & 1ull) == 0) return false;
    e.fresh &= ~(1ull << k);
    if (usable && host_hessians == e.host_hessians + off && e.layout == layout) {
      *out = e;
      for (int p = 0; p < e.layout.num_parts; ++p) {
        out->dev_gradients[p] = e.dev_gradients[p] + static_cast<size_t>(k) * e.layout.count[p];
        out->dev_hessians[p] = e.dev_hessians[p] + static_cast<size_t>(k) * e.layout.count[p];
      }
      return true;
    }
    if (!e.host_filled) {
      Log::Fatal("B200 drop-in: the gradients of this iteration exist on the device only and the learner cannot train from "
                 "that copy (row layout or sampling mode changed); set GBL_DROPIN_FILL_HOST_GRADIENTS=1");
    }
    return false;
  }
  return false;
}

void DeviceGradients::AddConsumer(const DeviceLayout& layout) {
  std::lock_guard<std::mutex> lock(g_registry_mutex);
  g_consumers.push_back(layout);
}

void DeviceGradients::RemoveConsumer(const DeviceLayout& layout) {
  std::lock_guard<std::mutex> lock(g_registry_mutex);
  for (size_t i = 0; i < g_consumers.size(); ++i) {
    if (g_consumers[i] == layout) {
      g_consumers.erase(g_consumers.begin() + i);
      return;
    }
  }
}

bool DeviceGradients::FindConsumer(data_size_t num_data, int num_parts, int device_id, DeviceLayout* out) {
  std::lock_guard<std::mutex> lock(g_registry_mutex);
  for (auto it = g_consumers.rbegin(); it != g_consumers.rend(); ++it) {
    if (it->num_data == num_data && it->num_parts == num_parts && (num_parts > 1 || it->device[0] == device_id)) {
      *out = *it;
      return true;
    }
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
  if (type == "multiclass") return config.num_class >= 2 && config.num_class <= 64;  // one `fresh` bit per class
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
  force_fill_host_ = EnvFlag("GBL_DROPIN_FILL_HOST_GRADIENTS", false);
  num_class_ = type == "multiclass" ? config.num_class : 1;
}

B200Objective::~B200Objective() { ReleaseDevice(); }

template <typename F>
void B200Objective::ForEachPart(F fn) const {
  const int R = layout_.num_parts;
  if (R == 1) {
    fn(0);
    return;
  }
  // one host thread per GPU: the uploads run on all PCIe links at once
  std::vector<std::string> errors(R);
  std::vector<std::thread> threads;
  for (int p = 0; p < R; ++p) {
    threads.emplace_back([&, p] {
      try {
        fn(p);
      } catch (const std::exception& ex) {
        errors[p] = ex.what();
        if (errors[p].empty()) errors[p] = "error";
      }
    });
  }
  for (std::thread& t : threads) t.join();
  for (const std::string& e : errors) {
    if (!e.empty()) Log::Fatal("%s", e.c_str());
  }
}

void B200Objective::ReleaseDevice() const {
  if (published_ != nullptr) DeviceGradients::Withdraw(published_);
  published_ = nullptr;
  if (registered_score_ != nullptr) gbl_host_unregister(const_cast<double*>(registered_score_));
  registered_score_ = nullptr;
  for (int p = 0; p < layout_.num_parts; ++p) {
    Part& part = parts_[p];
    if (part.objective != nullptr) gbl_objective_free(part.objective);
    for (void* ptr : {static_cast<void*>(part.score), static_cast<void*>(part.gradients), static_cast<void*>(part.hessians)}) {
      if (ptr != nullptr) gbl_device_free(layout_.device[p], ptr);
    }
    part = Part();
  }
  layout_ = DeviceLayout();
  on_device_ = false;
  decided_ = false;
}

void B200Objective::Init(const Metadata& metadata, data_size_t num_data) {
  host_->Init(metadata, num_data);  // the reference's checks, label statistics, BoostFromScore inputs
  ReleaseDevice();
  num_data_ = num_data;
  metadata_ = &metadata;  // owned by the Dataset, which outlives the objective's use of it (objective_function.h:24-29)
}

// The objective is initialised before GBDT creates its tree learner (c_api.cpp:174-186, gbdt.cpp:100-112), so the device
// side is set up at the first GetGradients, when the learner has announced on which GPUs the rows live.
void B200Objective::EnsureDevice() const {
  if (decided_) return;
  decided_ = true;
  on_device_ = false;
  const Metadata& md = *metadata_;
  if (num_data_ <= 0 || md.label() == nullptr) return;
  // position-debiased lambdarank keeps per-position bias factors that GetGradients updates (rank_objective.hpp:91-120,
  // 123-172): host only
  if (type_ == "lambdarank" && (md.positions() != nullptr || md.query_boundaries() == nullptr)) return;
  DeviceLayout lay;
  fill_host_ = force_fill_host_;
  const int device_id = std::max(0, config_.gpu_device_id);
  if (!DeviceGradients::FindConsumer(num_data_, std::max(1, config_.num_gpu), device_id, &lay)) {
    // no learner takes device copies: one GPU computes, the host vectors are filled
    lay = DeviceLayout();
    lay.num_data = num_data_;
    lay.num_parts = 1;
    lay.begin[0] = 0;
    lay.count[0] = num_data_;
    lay.device[0] = device_id;
    fill_host_ = true;
  }
  // a query's documents may straddle two GPUs' row ranges: lambdarank stays on the host when the rows are sharded
  if (lay.num_parts > 1 && type_ == "lambdarank") return;
  layout_ = lay;
  gbl_objective_config c;
  std::memset(&c, 0, sizeof(c));
  c.kind = type_ == "regression" ? GBL_OBJ_REGRESSION_L2
                                 : (type_ == "binary" ? GBL_OBJ_BINARY : (type_ == "multiclass" ? GBL_OBJ_MULTICLASS : GBL_OBJ_LAMBDARANK));
  c.num_class = type_ == "multiclass" ? num_class_ : 0;
  c.lambdarank_target = std::max(0, RankTargetIndex(config_.lambdarank_target));
  c.lambdarank_truncation_level = config_.lambdarank_truncation_level;
  c.lambdarank_norm = config_.lambdarank_norm ? 1 : 0;
  c.sigmoid = config_.sigmoid;
  c.lambdagap_weight = config_.lambdagap_weight;
  c.is_unbalance = config_.is_unbalance ? 1 : 0;
  c.scale_pos_weight = config_.scale_pos_weight;
  c.num_label_gain = static_cast<int>(config_.label_gain.size());
  c.label_gain = config_.label_gain.empty() ? nullptr : config_.label_gain.data();
  // BinaryLogloss::Init derives need_train and the is_unbalance weights from the label counts of ALL rows
  // (binary_objective.hpp:73-101); a part sees only its own
  int64_t positive = 0, negative = 0;
  if (type_ == "binary" && lay.num_parts > 1) {
    const label_t* label = md.label();
#pragma omp parallel for schedule(static) reduction(+ : positive, negative)
    for (data_size_t i = 0; i < num_data_; ++i) {
      if (label[i] > 0) {
        ++positive;
      } else {
        ++negative;
      }
    }
  }
  const bool rank = type_ == "lambdarank";
  ForEachPart([&](int p) {
    Part& part = parts_[p];
    const data_size_t b = lay.begin[p];
    const size_t n = static_cast<size_t>(lay.count[p]);
    const int dev = lay.device[p];
    auto check = [](int ret) {
      if (ret != 0) throw std::runtime_error(gbl_last_error());
    };
    check(gbl_objective_create(&c, lay.count[p], md.label() + b, md.weights() != nullptr ? md.weights() + b : nullptr,
                               rank ? md.query_boundaries() : nullptr, rank ? md.num_queries() : 0, dev, &part.objective));
    if (type_ == "binary" && lay.num_parts > 1) check(gbl_objective_set_label_counts(part.objective, positive, negative));
    const size_t nk = n * static_cast<size_t>(num_class_);
    check(gbl_device_malloc(dev, nk * sizeof(double), reinterpret_cast<void**>(&part.score)));
    check(gbl_device_malloc(dev, nk * sizeof(float), reinterpret_cast<void**>(&part.gradients)));
    check(gbl_device_malloc(dev, nk * sizeof(float), reinterpret_cast<void**>(&part.hessians)));
  });
  on_device_ = true;
  Log::Info("B200 objective: %s gradients on %d GPU%s%s", type_.c_str(), lay.num_parts, lay.num_parts > 1 ? "s" : "",
            fill_host_ ? ", host vectors filled as well" : ", kept in HBM for the learner");
  if (EnvFlag("GBL_DROPIN_TIMING", false))
    std::fprintf(stderr, "[B200 drop-in] objective %s: device, %d part(s), %d class block(s), host vectors filled: %s\n",
                 type_.c_str(), lay.num_parts, num_class_, fill_host_ ? "yes" : "no");
}

void B200Objective::GetGradients(const double* score, score_t* gradients, score_t* hessians) const {
  EnsureDevice();
  if (!on_device_) {
    host_->GetGradients(score, gradients, hessians);
    return;
  }
  // GBDT's ScoreUpdater keeps one score array for the whole training (score_updater.hpp:24-48): page-lock it once
  if (registered_score_ != score) {
    if (registered_score_ != nullptr) gbl_host_unregister(const_cast<double*>(registered_score_));
    gbl_host_register(const_cast<double*>(score), static_cast<size_t>(num_data_) * num_class_ * sizeof(double));  // failure = slower copies
    registered_score_ = score;
  }
  const bool fill = fill_host_;
  ForEachPart([&](int p) {
    const Part& part = parts_[p];
    const data_size_t b = layout_.begin[p];
    const size_t n = static_cast<size_t>(layout_.count[p]);
    const int dev = layout_.device[p];
    auto check = [](int ret) {
      if (ret != 0) throw std::runtime_error(gbl_last_error());
    };
    const size_t total = static_cast<size_t>(num_data_);
    for (int k = 0; k < num_class_; ++k)  // class block k of the host array -> class block k of the part
      check(gbl_memcpy_h2d(dev, part.score + k * n, score + k * total + b, n * sizeof(double)));
    check(gbl_objective_get_gradients(part.objective, part.score, part.gradients, part.hessians, nullptr));
    if (fill) {
      for (int k = 0; k < num_class_; ++k) {
        check(gbl_memcpy_d2h(dev, gradients + k * total + b, part.gradients + k * n, n * sizeof(float)));
        check(gbl_memcpy_d2h(dev, hessians + k * total + b, part.hessians + k * n, n * sizeof(float)));
      }
    }
    check(gbl_device_synchronize(dev));  // the learner trains on its own stream
  });
  if (published_ != nullptr && published_ != gradients) DeviceGradients::Withdraw(published_);
  DeviceGradients entry;
  entry.host_gradients = gradients;
  entry.host_hessians = hessians;
  entry.layout = layout_;
  for (int p = 0; p < layout_.num_parts; ++p) {
    entry.dev_gradients[p] = parts_[p].gradients;
    entry.dev_hessians[p] = parts_[p].hessians;
  }
  entry.num_class = num_class_;
  entry.fresh = num_class_ >= 64 ? ~0ull : ((1ull << num_class_) - 1ull);
  entry.host_filled = fill;
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
