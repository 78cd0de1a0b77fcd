// c_api.cpp -- the reference's C API names for the hot-path subset (see include/lgbm_b200_c_api.h).
// Reference: src/c_api.cpp (Booster class :169-195, :410-423; API_BEGIN/API_END :38-52).
#include <algorithm>
#include <cstdio>
#include <cstring>
#include <mutex>

#include "lgbm_b200_c_api.h"
#include "lgbm_host.h"

using namespace lgbm_b200;

namespace {

thread_local char g_last_error[1024] = "Everything is fine";
thread_local gbl_comm* g_comm = nullptr;  // Network state is thread_local in the reference (src/network/network.cpp:17-27)
thread_local int g_num_machines = 1;

void SetLastError(const char* msg) {
  std::strncpy(g_last_error, msg, sizeof(g_last_error) - 1);
  g_last_error[sizeof(g_last_error) - 1] = '\0';
}

#define API_BEGIN() try {
#define API_END()                                                  \
  }                                                                \
  catch (std::exception & ex) { SetLastError(ex.what()); return -1; } \
  catch (...) { SetLastError("unknown exception"); return -1; }    \
  return 0;

// Booster (src/c_api.cpp:143-900): owns config, objective and boosting; one writer at a time
struct Booster {
  Config config;
  std::unique_ptr<ObjectiveFunction> objective;
  std::unique_ptr<GBDT> boosting;
  std::vector<std::unique_ptr<Metric>> train_metric;
  std::vector<std::vector<std::unique_ptr<Metric>>> valid_metrics;
  std::mutex mutex;
  Booster(const Dataset* train_data, const char* parameters) {
    config.Set(parameters ? parameters : "");
    // Booster ctor :185-191: a parallel learner without machines falls back to serial
    if (g_num_machines <= 1 && config.tree_learner != "serial") {
      Log::Warning("Only find one worker, will switch to serial tree learner");
      config.tree_learner = "serial";
    }
    if (g_num_machines > 1 && config.tree_learner == "serial") config.tree_learner = "data";
    // num_gpu > 1 in ONE Booster is implemented by the drop-in library (integration/, B200TreeLearner over row shards);
    // this stand-alone layer drives one GPU per Booster and scales with one rank per GPU (LGBM_NetworkInitWithNccl)
    if (config.num_gpu > 1 && g_num_machines <= 1)
      Log::Warning("num_gpu=%d is ignored by this library: this Booster trains on one GPU.  Use one rank per GPU "
                   "(LGBM_NetworkInitWithNccl + tree_learner=data) or integration/_ref/lib_lightgbm.so, which shards the "
                   "rows over num_gpu GPUs inside one Booster", config.num_gpu);
    objective.reset(ObjectiveFunction::CreateObjectiveFunction(config.objective, config));
    boosting.reset(new GBDT());
    boosting->Init(&config, train_data, objective.get(), g_num_machines > 1 ? g_comm : nullptr);
    // Booster::CreateObjectiveAndMetrics (src/c_api.cpp:216-227): one metric object per configured metric
    for (const std::string& type : config.metric_types) {
      std::unique_ptr<Metric> m(Metric::CreateMetric(type, config));
      if (m == nullptr) continue;
      m->Init(train_data->metadata(), train_data->num_data(), train_data->device_id());
      train_metric.push_back(std::move(m));
    }
    boosting->AddTrainingMetrics(Wrap(train_metric));
  }
  static std::vector<const Metric*> Wrap(const std::vector<std::unique_ptr<Metric>>& v) {  // Common::ConstPtrInVectorWrapper
    std::vector<const Metric*> out;
    for (const auto& m : v) out.push_back(m.get());
    return out;
  }
  // Booster::AddValidData (src/c_api.cpp:394-408)
  void AddValidData(const Dataset* valid_data) {
    std::lock_guard<std::mutex> lock(mutex);
    std::vector<std::unique_ptr<Metric>> metrics;
    for (const std::string& type : config.metric_types) {
      std::unique_ptr<Metric> m(Metric::CreateMetric(type, config));
      if (m == nullptr) continue;
      m->Init(valid_data->metadata(), valid_data->num_data(), valid_data->device_id());
      metrics.push_back(std::move(m));
    }
    boosting->AddValidDataset(valid_data, Wrap(metrics));
    valid_metrics.push_back(std::move(metrics));
  }
};

}  // namespace

extern "C" {

const char* LGBM_GetLastError(void) { return g_last_error; }

int LGBM_DatasetCreateFromMat(const void* data, int data_type, int32_t nrow, int32_t ncol, int is_row_major,
                              const char* parameters, const DatasetHandle reference, DatasetHandle* out) {
  API_BEGIN();
  Config config;
  config.Set(parameters ? parameters : "");
  *out = Dataset::CreateFromMat(data, data_type, nrow, ncol, is_row_major, config,
                                reinterpret_cast<const Dataset*>(reference));
  API_END();
}

int LGBM_DatasetCreateFromMatWithSample(const void* data, int data_type, int32_t nrow, int32_t ncol, int is_row_major,
                                        const void* sample, int32_t sample_nrow, int64_t num_total_row,
                                        const char* parameters, DatasetHandle* out) {
  API_BEGIN();
  Config config;
  config.Set(parameters ? parameters : "");
  if (sample == nullptr || sample_nrow <= 0) Log::Fatal("LGBM_DatasetCreateFromMatWithSample: empty sample");
  *out = Dataset::CreateFromMatWithSample(data, data_type, nrow, ncol, is_row_major, sample, sample_nrow, num_total_row,
                                          config);
  API_END();
}

int LGBM_DatasetSetField(DatasetHandle handle, const char* field_name, const void* field_data, int num_element,
                         int type) {
  API_BEGIN();
  reinterpret_cast<Dataset*>(handle)->SetField(field_name, field_data, num_element, type);
  API_END();
}

int LGBM_DatasetGetNumData(DatasetHandle handle, int32_t* out) {
  API_BEGIN();
  *out = reinterpret_cast<Dataset*>(handle)->num_data();
  API_END();
}

int LGBM_DatasetGetNumFeature(DatasetHandle handle, int32_t* out) {
  API_BEGIN();
  *out = reinterpret_cast<Dataset*>(handle)->num_total_features();
  API_END();
}

int LGBM_DatasetFree(DatasetHandle handle) {
  API_BEGIN();
  delete reinterpret_cast<Dataset*>(handle);
  API_END();
}

int LGBM_BoosterCreate(const DatasetHandle train_data, const char* parameters, BoosterHandle* out) {
  API_BEGIN();
  *out = new Booster(reinterpret_cast<const Dataset*>(train_data), parameters);
  API_END();
}

int LGBM_BoosterFree(BoosterHandle handle) {
  API_BEGIN();
  delete reinterpret_cast<Booster*>(handle);
  API_END();
}

int LGBM_BoosterUpdateOneIter(BoosterHandle handle, int* is_finished) {
  API_BEGIN();
  Booster* b = reinterpret_cast<Booster*>(handle);
  std::lock_guard<std::mutex> lock(b->mutex);
  *is_finished = b->boosting->TrainOneIter(nullptr, nullptr) ? 1 : 0;
  API_END();
}

int LGBM_BoosterUpdateOneIterCustom(BoosterHandle handle, const float* grad, const float* hess, int* is_finished) {
  API_BEGIN();
  Booster* b = reinterpret_cast<Booster*>(handle);
  std::lock_guard<std::mutex> lock(b->mutex);
  *is_finished = b->boosting->TrainOneIter(grad, hess) ? 1 : 0;
  API_END();
}

int LGBM_BoosterAddValidData(BoosterHandle handle, const DatasetHandle valid_data) {
  API_BEGIN();
  reinterpret_cast<Booster*>(handle)->AddValidData(reinterpret_cast<const Dataset*>(valid_data));
  API_END();
}

int LGBM_BoosterGetEvalCounts(BoosterHandle handle, int* out_len) {
  API_BEGIN();
  Booster* b = reinterpret_cast<Booster*>(handle);
  int ret = 0;  // Booster::GetEvalCounts (src/c_api.cpp:836-843)
  for (const auto& m : b->train_metric) ret += static_cast<int>(m->GetName().size());
  *out_len = ret;
  API_END();
}

int LGBM_BoosterGetEvalNames(BoosterHandle handle, const int len, int* out_len, const size_t buffer_len,
                             size_t* out_buffer_len, char** out_strs) {
  API_BEGIN();
  Booster* b = reinterpret_cast<Booster*>(handle);
  *out_buffer_len = 0;  // Booster::GetEvalNames (src/c_api.cpp:845-860)
  int idx = 0;
  for (const auto& m : b->train_metric) {
    for (const auto& name : m->GetName()) {
      if (idx < len && buffer_len > 0) {
        std::memcpy(out_strs[idx], name.c_str(), std::min(name.size() + 1, buffer_len));
        out_strs[idx][buffer_len - 1] = '\0';
      }
      *out_buffer_len = std::max(name.size() + 1, *out_buffer_len);
      ++idx;
    }
  }
  *out_len = idx;
  API_END();
}

int LGBM_BoosterGetEval(BoosterHandle handle, int data_idx, int* out_len, double* out_results) {
  API_BEGIN();
  Booster* b = reinterpret_cast<Booster*>(handle);
  std::lock_guard<std::mutex> lock(b->mutex);
  const std::vector<double> res = b->boosting->GetEvalAt(data_idx);  // LGBM_BoosterGetEval (src/c_api.cpp:2200-2212)
  *out_len = static_cast<int>(res.size());
  for (size_t i = 0; i < res.size(); ++i) out_results[i] = res[i];
  API_END();
}

int LGBM_BoosterResetParameter(BoosterHandle handle, const char* parameters) {
  API_BEGIN();
  Booster* b = reinterpret_cast<Booster*>(handle);
  std::lock_guard<std::mutex> lock(b->mutex);
  // Booster::ResetConfig (src/c_api.cpp:317-392): what may not change after the booster exists
  const auto params = Config::Str2Map(parameters ? parameters : "");
  if (params.count("num_class")) Log::Fatal("Cannot change num_class during training");
  if (params.count("boosting")) Log::Fatal("Cannot change boosting during training");
  if (params.count("metric")) Log::Fatal("Cannot change metric during training");
  if (params.count("objective")) Log::Fatal("Cannot change objective during training (not supported by the B200 host layer)");
  if (params.count("use_quantized_grad") || params.count("device_type") || params.count("tree_learner"))
    Log::Fatal("Cannot change use_quantized_grad / device_type / tree_learner during training");
  b->config.Set(parameters ? parameters : "");
  b->boosting->ResetConfig(&b->config);
  API_END();
}

int LGBM_BoosterGetNumClasses(BoosterHandle /*handle*/, int* out_len) {
  API_BEGIN();
  *out_len = 1;
  API_END();
}

int LGBM_BoosterNumModelPerIteration(BoosterHandle /*handle*/, int* out_tree_per_iteration) {
  API_BEGIN();
  *out_tree_per_iteration = 1;
  API_END();
}

int LGBM_BoosterGetNumFeature(BoosterHandle handle, int* out_len) {
  API_BEGIN();
  *out_len = reinterpret_cast<Booster*>(handle)->boosting->MaxFeatureIdx() + 1;
  API_END();
}

int LGBM_BoosterGetLeafValue(BoosterHandle handle, int tree_idx, int leaf_idx, double* out_val) {
  API_BEGIN();
  Booster* b = reinterpret_cast<Booster*>(handle);
  const auto& models = b->boosting->models();
  if (tree_idx < 0 || tree_idx >= static_cast<int>(models.size())) Log::Fatal("tree_idx %d out of range", tree_idx);
  if (leaf_idx < 0 || leaf_idx >= models[tree_idx]->num_leaves()) Log::Fatal("leaf_idx %d out of range", leaf_idx);
  *out_val = models[tree_idx]->LeafOutput(leaf_idx);
  API_END();
}

int LGBM_BoosterFeatureImportance(BoosterHandle handle, int num_iteration, int importance_type, double* out_results) {
  API_BEGIN();
  Booster* b = reinterpret_cast<Booster*>(handle);
  const std::vector<double> imp = b->boosting->FeatureImportance(num_iteration, importance_type);
  for (size_t i = 0; i < imp.size(); ++i) out_results[i] = imp[i];
  API_END();
}

int LGBM_BoosterSaveModel(BoosterHandle handle, int start_iteration, int num_iteration, int /*feature_importance_type*/,
                          const char* filename) {
  API_BEGIN();
  Booster* b = reinterpret_cast<Booster*>(handle);
  const std::string model = b->boosting->SaveModelToString(start_iteration, num_iteration);
  FILE* fh = std::fopen(filename, "wb");
  if (fh == nullptr) Log::Fatal("Model file %s is not available for writes", filename);
  const size_t n = std::fwrite(model.data(), 1, model.size(), fh);
  std::fclose(fh);
  if (n != model.size()) Log::Fatal("Could not write the model to %s", filename);
  API_END();
}

int LGBM_BoosterGetCurrentIteration(BoosterHandle handle, int* out_iteration) {
  API_BEGIN();
  *out_iteration = reinterpret_cast<Booster*>(handle)->boosting->GetCurrentIteration();
  API_END();
}

int LGBM_BoosterNumberOfTotalModel(BoosterHandle handle, int* out_models) {
  API_BEGIN();
  *out_models = reinterpret_cast<Booster*>(handle)->boosting->NumberOfTotalModel();
  API_END();
}

int LGBM_BoosterGetNumPredict(BoosterHandle handle, int data_idx, int64_t* out_len) {
  API_BEGIN();
  *out_len = reinterpret_cast<Booster*>(handle)->boosting->GetNumPredictAt(data_idx);
  API_END();
}

int LGBM_BoosterGetPredict(BoosterHandle handle, int data_idx, int64_t* out_len, double* out_result) {
  API_BEGIN();
  Booster* b = reinterpret_cast<Booster*>(handle);
  std::lock_guard<std::mutex> lock(b->mutex);
  b->boosting->GetPredictAt(data_idx, out_result, out_len);
  API_END();
}

int LGBM_B200_FindBin(double* values, int num_values, int64_t total_sample_cnt, int max_bin, int min_data_in_bin,
                      int min_split_data, int pre_filter, int use_missing, int zero_as_missing, int32_t out_scalars[5],
                      double* out_upper_bounds, int cap) {
  API_BEGIN();
  BinMapper m;
  m.FindBin(values, num_values, static_cast<size_t>(total_sample_cnt), max_bin, min_data_in_bin, min_split_data,
            pre_filter != 0, use_missing != 0, zero_as_missing != 0);
  out_scalars[0] = m.num_bin();
  out_scalars[1] = static_cast<int>(m.missing_type());
  out_scalars[2] = static_cast<int>(m.GetDefaultBin());
  out_scalars[3] = static_cast<int>(m.GetMostFreqBin());
  out_scalars[4] = m.is_trivial() ? 1 : 0;
  if (m.num_bin() > cap) Log::Fatal("LGBM_B200_FindBin: %d bins exceed the output capacity %d", m.num_bin(), cap);
  for (int b = 0; b < m.num_bin(); ++b) out_upper_bounds[b] = m.BinToValue(b);
  API_END();
}

int LGBM_B200_DumpConfig(const char* parameters, int64_t buffer_len, int64_t* out_len, char* out_str) {
  API_BEGIN();
  // test tap: what Config::Set made of a parameter string, in the `[key: value]` form of the model file's parameter
  // block (src/io/config_auto.cpp Config::SaveMembersToString), so that it can be compared with the reference's
  Config c;
  c.Set(parameters ? parameters : "");
  std::string metrics;
  for (size_t i = 0; i < c.metric_types.size(); ++i) metrics += (i ? "," : "") + c.metric_types[i];
  std::string eval_at;
  for (size_t i = 0; i < c.eval_at.size(); ++i) eval_at += (i ? "," : "") + std::to_string(c.eval_at[i]);
  const std::string s = "[objective: " + c.objective + "]\n[metric: " + metrics + "]\n[eval_at: " + eval_at + "]\n[seed: " +
                        std::to_string(c.seed) + "]\n[data_random_seed: " + std::to_string(c.data_random_seed) +
                        "]\n[bagging_seed: " + std::to_string(c.bagging_seed) + "]\n[feature_fraction_seed: " +
                        std::to_string(c.feature_fraction_seed) + "]\n[num_leaves: " + std::to_string(c.num_leaves) +
                        "]\n[min_data_in_leaf: " + std::to_string(c.min_data_in_leaf) + "]\n[bagging_freq: " +
                        std::to_string(c.bagging_freq) + "]\n";
  *out_len = static_cast<int64_t>(s.size()) + 1;
  if (*out_len <= buffer_len) std::memcpy(out_str, s.c_str(), *out_len);
  API_END();
}

int LGBM_BoosterGetLearnerHandle(BoosterHandle handle, void** out_learner) {
  API_BEGIN();
  Booster* b = reinterpret_cast<Booster*>(handle);
  B200TreeLearner* bl = dynamic_cast<B200TreeLearner*>(b->boosting->tree_learner());
  if (bl == nullptr) Log::Fatal("booster has no B200 tree learner");
  *out_learner = bl->handle();
  API_END();
}

int LGBM_BoosterGetTrainScoreRaw(BoosterHandle handle, int64_t* out_len, double* out_result) {
  API_BEGIN();
  Booster* b = reinterpret_cast<Booster*>(handle);
  std::lock_guard<std::mutex> lock(b->mutex);
  b->boosting->GetTrainScore(out_result);
  *out_len = b->boosting->GetNumPredictAt(0);
  API_END();
}

int LGBM_BoosterPredictForMat(BoosterHandle handle, const void* data, int data_type, int32_t nrow, int32_t ncol,
                              int is_row_major, int predict_type, int start_iteration, int num_iteration,
                              const char* /*parameter*/, int64_t* out_len, double* out_result) {
  API_BEGIN();
  Booster* b = reinterpret_cast<Booster*>(handle);
  if (predict_type != C_API_PREDICT_NORMAL && predict_type != C_API_PREDICT_RAW_SCORE)
    Log::Fatal("predict_type %d is not supported by the B200 host layer (normal, raw score)", predict_type);
  if (start_iteration > 0 || (num_iteration > 0 && num_iteration < b->boosting->NumberOfTotalModel()))
    Log::Fatal("partial-model prediction is not supported by the B200 host layer");
  b->boosting->PredictForMat(data, data_type, nrow, ncol, is_row_major, predict_type == C_API_PREDICT_RAW_SCORE, out_result);
  *out_len = nrow;
  API_END();
}

int LGBM_BoosterSaveModelToString(BoosterHandle handle, int start_iteration, int num_iteration,
                                  int /*feature_importance_type*/, int64_t buffer_len, int64_t* out_len, char* out_str) {
  API_BEGIN();
  Booster* b = reinterpret_cast<Booster*>(handle);
  const std::string model = b->boosting->SaveModelToString(start_iteration, num_iteration);
  *out_len = static_cast<int64_t>(model.size()) + 1;
  if (*out_len <= buffer_len) std::memcpy(out_str, model.c_str(), *out_len);
  API_END();
}

int LGBM_NetworkGetNcclUniqueId(uint8_t out_id[128]) {
  API_BEGIN();
  GblCheck(gbl_comm_get_unique_id(out_id));
  API_END();
}

int LGBM_NetworkInitWithNccl(const uint8_t id[128], int num_machines, int rank, int device_id) {
  API_BEGIN();
  if (g_comm != nullptr) Log::Fatal("Network already initialised on this thread");
  if (num_machines > 1) {
    GblCheck(gbl_comm_create(id, num_machines, rank, device_id, &g_comm));
    g_num_machines = num_machines;
  }
  API_END();
}

int LGBM_NetworkFree(void) {
  API_BEGIN();
  if (g_comm != nullptr) gbl_comm_free(g_comm);
  g_comm = nullptr;
  g_num_machines = 1;
  API_END();
}

}  // extern "C"
