// metric.cpp -- the reference's Metric interface over the device metrics of the C ABI (gbl_metric_*).
//
// Reference: Metric::CreateMetric (src/metric/metric.cpp:20-133), Metric (include/LightGBM/metric.h:25-70),
// NDCGMetric (src/metric/rank_metric.hpp:20-166), BinaryMetric / AUCMetric (src/metric/binary_metric.hpp),
// RegressionMetric (src/metric/regression_metric.hpp); DCGCalculator::DefaultEvalAt / DefaultLabelGain
// (src/metric/dcg_calculator.cpp:19-43).
#include "lgbm_host.h"

namespace lgbm_b200 {

namespace {
int MetricKind(const std::string& type) {
  if (type == "l2") return GBL_METRIC_L2;
  if (type == "rmse") return GBL_METRIC_RMSE;
  if (type == "l1") return GBL_METRIC_L1;
  if (type == "binary_logloss") return GBL_METRIC_BINARY_LOGLOSS;
  if (type == "binary_error") return GBL_METRIC_BINARY_ERROR;
  if (type == "auc") return GBL_METRIC_AUC;
  if (type == "ndcg") return GBL_METRIC_NDCG;
  return -1;
}
}  // namespace

Metric* Metric::CreateMetric(const std::string& type, const Config& config) {
  if (type == "custom") return nullptr;
  const int kind = MetricKind(type);
  if (kind < 0) {
    // the reference returns nullptr for names it does not know; a metric it knows but this library does not
    // evaluate must not silently disappear from the eval list
    Log::Fatal("metric=%s is not evaluated by the B200 host layer (l2, rmse, l1, binary_logloss, binary_error, auc, ndcg)",
               type.c_str());
  }
  Metric* m = new Metric();
  m->type_ = type;
  m->config_ = config;
  m->bigger_better_ = (kind == GBL_METRIC_AUC || kind == GBL_METRIC_NDCG);
  if (kind == GBL_METRIC_NDCG) {
    if (m->config_.eval_at.empty()) m->config_.eval_at = {1, 2, 3, 4, 5};  // DCGCalculator::DefaultEvalAt
    if (m->config_.label_gain.empty()) {                                    // DCGCalculator::DefaultLabelGain
      m->config_.label_gain.push_back(0.0);
      for (int i = 1; i < 31; ++i) m->config_.label_gain.push_back(static_cast<double>((1 << i) - 1));
    }
  }
  return m;
}

Metric::~Metric() {
  if (handle_ != nullptr) gbl_metric_free(handle_);
}

void Metric::Init(const Metadata& metadata, data_size_t num_data, int device_id) {
  metadata_ = &metadata;
  num_data_ = num_data;
  device_id_ = device_id;
  name_.clear();
  if (type_ == "ndcg") {
    for (int k : config_.eval_at) name_.emplace_back(std::string("ndcg@") + std::to_string(k));
    if (metadata.query_boundaries.empty()) Log::Fatal("The NDCG metric requires query information");
  } else {
    name_.emplace_back(type_);
  }
  if (static_cast<data_size_t>(metadata.label.size()) != num_data) Log::Fatal("Metric::Init: the dataset has no label");
}

std::vector<double> Metric::Eval(const double* score, const ObjectiveFunction* objective, void* stream) const {
  if (metadata_ == nullptr) Log::Fatal("Metric::Eval before Metric::Init");
  const bool with_objective = objective != nullptr;
  if (handle_ != nullptr && handle_with_objective_ != with_objective) {
    gbl_metric_free(handle_);
    handle_ = nullptr;
  }
  if (handle_ == nullptr) {
    gbl_metric_config c{};
    c.kind = MetricKind(type_);
    std::vector<int32_t> eval_at(config_.eval_at.begin(), config_.eval_at.end());
    c.num_eval_at = static_cast<int32_t>(eval_at.size());
    c.eval_at = eval_at.data();
    c.num_label_gain = static_cast<int32_t>(config_.label_gain.size());
    c.label_gain = config_.label_gain.data();
    // BinaryMetric::Eval converts raw scores with the objective's ConvertOutput when there is one
    // (binary_metric.hpp:76-95); the only objective these metrics are paired with here is `binary`
    c.sigmoid = 0.0;
    if (with_objective && (c.kind == GBL_METRIC_BINARY_LOGLOSS || c.kind == GBL_METRIC_BINARY_ERROR)) {
      if (std::string(objective->GetName()) == "binary") c.sigmoid = config_.sigmoid;
    }
    GblCheck(gbl_metric_create(&c, num_data_, metadata_->label.data(),
                               metadata_->weights.empty() ? nullptr : metadata_->weights.data(),
                               metadata_->query_boundaries.empty() ? nullptr : metadata_->query_boundaries.data(),
                               metadata_->num_queries(), device_id_, &handle_));
    handle_with_objective_ = with_objective;
  }
  int32_t n = 0;
  GblCheck(gbl_metric_num_outputs(handle_, &n));
  std::vector<double> out(n);
  GblCheck(gbl_metric_eval(handle_, score, stream, out.data()));
  return out;
}

}  // namespace lgbm_b200
