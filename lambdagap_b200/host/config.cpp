// config.cpp -- Log + Config for the host layer.
// Reference: Config::Str2Map / KV2Map / Set (src/io/config.cpp:20-120,255-312), alias table
// (src/io/config_auto.cpp:10-180), objective aliases (src/io/config.cpp:70-110).
#include <algorithm>
#include <cmath>
#include <cstring>
#include <sstream>

#include "lgbm_host.h"

namespace lgbm_b200 {

int Log::level = 1;

static void vlog(const char* tag, const char* fmt, va_list ap) {
  char buf[2048];
  vsnprintf(buf, sizeof(buf), fmt, ap);
  fprintf(stderr, "[LightGBM-B200] [%s] %s\n", tag, buf);
  fflush(stderr);
}

void Log::Fatal(const char* fmt, ...) {
  char buf[2048];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  if (level >= 0) fprintf(stderr, "[LightGBM-B200] [Fatal] %s\n", buf);
  throw std::runtime_error(buf);  // include/LightGBM/utils/log.h:139
}
void Log::Warning(const char* fmt, ...) {
  if (level < 0) return;
  va_list ap;
  va_start(ap, fmt);
  vlog("Warning", fmt, ap);
  va_end(ap);
}
void Log::Info(const char* fmt, ...) {
  if (level < 1) return;
  va_list ap;
  va_start(ap, fmt);
  vlog("Info", fmt, ap);
  va_end(ap);
}
void Log::Debug(const char* fmt, ...) {
  if (level < 2) return;
  va_list ap;
  va_start(ap, fmt);
  vlog("Debug", fmt, ap);
  va_end(ap);
}

void GblCheck(int ret) {
  if (ret != 0) Log::Fatal("%s", gbl_last_error());
}

static const std::unordered_map<std::string, std::string>& AliasTable() {
  static const std::unordered_map<std::string, std::string> t = {
      {"objective_type", "objective"}, {"app", "objective"}, {"application", "objective"}, {"loss", "objective"},
      {"boosting_type", "boosting"}, {"boost", "boosting"},
      {"num_iteration", "num_iterations"}, {"n_iter", "num_iterations"}, {"num_tree", "num_iterations"},
      {"num_trees", "num_iterations"}, {"num_round", "num_iterations"}, {"num_rounds", "num_iterations"},
      {"nrounds", "num_iterations"}, {"num_boost_round", "num_iterations"}, {"n_estimators", "num_iterations"},
      {"max_iter", "num_iterations"},
      {"shrinkage_rate", "learning_rate"}, {"eta", "learning_rate"},
      {"num_leaf", "num_leaves"}, {"max_leaves", "num_leaves"}, {"max_leaf", "num_leaves"},
      {"max_leaf_nodes", "num_leaves"},
      {"tree", "tree_learner"}, {"tree_type", "tree_learner"}, {"tree_learner_type", "tree_learner"},
      {"num_thread", "num_threads"}, {"nthread", "num_threads"}, {"nthreads", "num_threads"}, {"n_jobs", "num_threads"},
      {"device", "device_type"},
      {"random_seed", "seed"}, {"random_state", "seed"},
      {"min_data_per_leaf", "min_data_in_leaf"}, {"min_data", "min_data_in_leaf"},
      {"min_child_samples", "min_data_in_leaf"}, {"min_samples_leaf", "min_data_in_leaf"},
      {"min_sum_hessian_per_leaf", "min_sum_hessian_in_leaf"}, {"min_sum_hessian", "min_sum_hessian_in_leaf"},
      {"min_hessian", "min_sum_hessian_in_leaf"}, {"min_child_weight", "min_sum_hessian_in_leaf"},
      {"sub_row", "bagging_fraction"}, {"subsample", "bagging_fraction"}, {"bagging", "bagging_fraction"},
      {"subsample_freq", "bagging_freq"},
      {"sub_feature", "feature_fraction"}, {"colsample_bytree", "feature_fraction"},
      {"sub_feature_bynode", "feature_fraction_bynode"}, {"colsample_bynode", "feature_fraction_bynode"},
      {"max_tree_output", "max_delta_step"}, {"max_leaf_output", "max_delta_step"},
      {"reg_alpha", "lambda_l1"}, {"l1_regularization", "lambda_l1"},
      {"reg_lambda", "lambda_l2"}, {"lambda", "lambda_l2"}, {"l2_regularization", "lambda_l2"},
      {"min_split_gain", "min_gain_to_split"},
      {"verbose", "verbosity"},
      {"max_bins", "max_bin"},
      {"subsample_for_bin", "bin_construct_sample_cnt"},
      {"data_seed", "data_random_seed"},
      {"unbalance", "is_unbalance"}, {"unbalanced_sets", "is_unbalance"},
      {"metrics", "metric"}, {"metric_types", "metric"},
      {"ndcg_eval_at", "eval_at"}, {"ndcg_at", "eval_at"}, {"map_eval_at", "eval_at"}, {"map_at", "eval_at"},
      {"linear_trees", "linear_tree"},
      {"pos_sub_row", "pos_bagging_fraction"}, {"pos_subsample", "pos_bagging_fraction"},
      {"pos_bagging", "pos_bagging_fraction"}, {"neg_sub_row", "neg_bagging_fraction"},
      {"neg_subsample", "neg_bagging_fraction"}, {"neg_bagging", "neg_bagging_fraction"},
      {"bagging_fraction_seed", "bagging_seed"}, {"extra_tree", "extra_trees"},
      {"is_enable_bundle", "enable_bundle"}, {"bundle", "enable_bundle"},
      {"training_metric", "is_provide_training_metric"}, {"is_training_metric", "is_provide_training_metric"},
      {"train_metric", "is_provide_training_metric"},
      {"mc", "monotone_constraints"},
      {"monotone_constraint", "monotone_constraints"},
      {"monotonic_cst", "monotone_constraints"},
      {"feature_contrib", "feature_contri"},
      {"fc", "feature_contri"},
      {"fp", "feature_contri"},
      {"feature_penalty", "feature_contri"},
      {"fs", "forcedsplits_filename"},
      {"forced_splits_filename", "forcedsplits_filename"},
      {"forced_splits_file", "forcedsplits_filename"},
      {"forced_splits", "forcedsplits_filename"},
      {"cat_feature", "categorical_feature"},
      {"categorical_column", "categorical_feature"},
      {"cat_column", "categorical_feature"},
      {"categorical_features", "categorical_feature"},
      {"num_classes", "num_class"},
  };
  return t;
}

std::unordered_map<std::string, std::string> Config::Str2Map(const std::string& parameters) {
  std::unordered_map<std::string, std::string> params;
  std::string tok;
  std::stringstream ss(parameters);
  while (ss >> tok) {
    const size_t eq = tok.find('=');
    if (eq == std::string::npos || eq == 0) {
      if (!tok.empty()) Log::Warning("Unknown parameter %s", tok.c_str());
      continue;
    }
    std::string key = tok.substr(0, eq);
    const std::string value = tok.substr(eq + 1);
    auto it = AliasTable().find(key);
    if (it != AliasTable().end()) key = it->second;
    if (params.count(key) == 0) params[key] = value;  // first occurrence wins (Config::KV2Map)
  }
  return params;
}

static bool ParseBool(const std::string& key, std::string v) {
  std::transform(v.begin(), v.end(), v.begin(), ::tolower);
  if (v == "false" || v == "-" || v == "0") return false;
  if (v == "true" || v == "+" || v == "1") return true;
  Log::Fatal("Parameter %s should be \"true\"/\"+\" or \"false\"/\"-\", got \"%s\"", key.c_str(), v.c_str());
}

static std::string ParseObjectiveAlias(const std::string& type) {
  // src/io/config.cpp:70-110 (ParseObjectiveAlias)
  if (type == "regression" || type == "regression_l2" || type == "mean_squared_error" || type == "mse" ||
      type == "l2" || type == "l2_root" || type == "root_mean_squared_error" || type == "rmse") {
    return "regression";
  }
  if (type == "none" || type == "null" || type == "custom" || type == "na") return "custom";
  return type;
}

static std::string ParseMetricAlias(const std::string& type) {
  // include/LightGBM/config.h:1316-1344 (the metrics this library evaluates, plus "custom")
  if (type == "regression" || type == "regression_l2" || type == "l2" || type == "mean_squared_error" || type == "mse") return "l2";
  if (type == "l2_root" || type == "root_mean_squared_error" || type == "rmse") return "rmse";
  if (type == "regression_l1" || type == "l1" || type == "mean_absolute_error" || type == "mae") return "l1";
  if (type == "binary_logloss" || type == "binary") return "binary_logloss";
  if (type == "ndcg" || type == "lambdarank" || type == "rank_xendcg" || type == "xendcg" || type == "xe_ndcg" ||
      type == "xe_ndcg_mart" || type == "xendcg_mart")
    return "ndcg";
  if (type == "none" || type == "null" || type == "custom" || type == "na") return "custom";
  return type;
}

void Config::Set(const std::string& parameters) {
  auto params = Str2Map(parameters);
  auto get = [&params](const char* key) -> const std::string* {
    auto it = params.find(key);
    return it == params.end() ? nullptr : &it->second;
  };
  // generate seeds by seed (src/io/config.cpp:257-268)
  if (const std::string* v = get("seed")) {
    seed = std::stoi(*v);
    unsigned int x = static_cast<unsigned int>(seed);
    auto next_short = [&x](int lo, int hi) {
      x = (214013 * x + 2531011);
      return static_cast<int>((x >> 16) & 0x7FFF) % (hi - lo) + lo;
    };
    const int int_max = std::numeric_limits<int16_t>::max();
    data_random_seed = next_short(0, int_max);
    bagging_seed = next_short(0, int_max);
    next_short(0, int_max);  // drop_seed
    feature_fraction_seed = next_short(0, int_max);
  }
#define GET_INT(name) if (const std::string* v = get(#name)) name = std::stoi(*v)
#define GET_DBL(name) if (const std::string* v = get(#name)) name = std::stod(*v)
#define GET_BOOL(name) if (const std::string* v = get(#name)) name = ParseBool(#name, *v)
#define GET_STR(name) if (const std::string* v = get(#name)) name = *v
  GET_STR(objective);
  objective = ParseObjectiveAlias(objective);
  GET_STR(boosting);
  GET_STR(tree_learner);
  GET_STR(device_type);
  GET_INT(num_iterations);
  GET_DBL(learning_rate);
  GET_INT(num_leaves);
  GET_INT(num_threads);
  GET_BOOL(deterministic);
  GET_INT(num_gpu);
  GET_INT(gpu_device_id);
  GET_INT(max_depth);
  GET_INT(min_data_in_leaf);
  GET_DBL(min_sum_hessian_in_leaf);
  GET_DBL(bagging_fraction);
  GET_INT(bagging_freq);
  GET_INT(bagging_seed);
  GET_DBL(pos_bagging_fraction);
  GET_DBL(neg_bagging_fraction);
  GET_BOOL(bagging_by_query);
  GET_DBL(feature_fraction);
  GET_INT(feature_fraction_seed);
  GET_DBL(feature_fraction_bynode);
  GET_BOOL(extra_trees);
  GET_DBL(max_delta_step);
  GET_DBL(lambda_l1);
  GET_DBL(lambda_l2);
  GET_DBL(min_gain_to_split);
  GET_DBL(path_smooth);
  GET_BOOL(linear_tree);
  GET_INT(verbosity);
  GET_BOOL(use_quantized_grad);
  GET_INT(num_grad_quant_bins);
  GET_BOOL(quant_train_renew_leaf);
  GET_BOOL(stochastic_rounding);
  GET_INT(max_bin);
  GET_INT(min_data_in_bin);
  GET_INT(bin_construct_sample_cnt);
  GET_INT(data_random_seed);
  GET_BOOL(use_missing);
  GET_BOOL(zero_as_missing);
  GET_BOOL(feature_pre_filter);
  GET_BOOL(enable_bundle);
  GET_BOOL(boost_from_average);
  GET_BOOL(is_unbalance);
  GET_DBL(scale_pos_weight);
  GET_DBL(sigmoid);
  GET_BOOL(reg_sqrt);
  GET_STR(lambdarank_target);
  GET_INT(lambdarank_truncation_level);
  GET_BOOL(lambdarank_norm);
  GET_DBL(lambdagap_weight);
  GET_STR(metric);
#undef GET_INT
#undef GET_DBL
#undef GET_BOOL
#undef GET_STR
  auto parse_list = [](const std::string& s, std::vector<double>* out) {
    out->clear();
    std::string cur;
    for (char ch : s + ",") {
      if (ch == ',') {
        if (!cur.empty()) out->push_back(std::stod(cur));
        cur.clear();
      } else {
        cur.push_back(ch);
      }
    }
  };
  if (const std::string* v = get("label_gain")) parse_list(*v, &label_gain);
  if (const std::string* v = get("eval_at")) {
    std::vector<double> tmp;
    parse_list(*v, &tmp);
    eval_at.clear();
    for (double d : tmp) eval_at.push_back(static_cast<int>(d));
  }
  std::sort(eval_at.begin(), eval_at.end());  // src/io/config.cpp:287-288
  if (const std::string* v = get("is_provide_training_metric")) is_provide_training_metric = ParseBool("is_provide_training_metric", *v);
  // GetMetricType (src/io/config.cpp:152-163): the listed metrics (aliases resolved, duplicates dropped), or the
  // objective's own metric when none is given
  {
    auto parse_metrics = [this](const std::string& value) {
      metric_types.clear();
      std::string cur;
      for (char ch : value + ",") {
        if (ch == ',') {
          if (!cur.empty()) {
            const std::string type = ParseMetricAlias(cur);
            if (std::find(metric_types.begin(), metric_types.end(), type) == metric_types.end()) metric_types.push_back(type);
          }
          cur.clear();
        } else if (ch != ' ') {
          cur.push_back(ch);
        }
      }
    };
    const std::string* v = get("metric");
    if (v != nullptr) parse_metrics(*v);
    if (metric_types.empty() && (v == nullptr || v->empty())) parse_metrics(objective);
  }
  Log::level = verbosity;

  // ---- what this learner does not implement fails loudly instead of silently training something else
  if (objective != "regression" && objective != "binary" && objective != "lambdarank" && objective != "custom")
    Log::Fatal("objective=%s is not supported by the B200 learner (regression, binary, lambdarank, custom)", objective.c_str());
  if (boosting != "gbdt" && boosting != "gbrt") Log::Fatal("boosting=%s is not supported by the B200 learner (gbdt only)", boosting.c_str());
  if (tree_learner != "serial" && tree_learner != "data" && tree_learner != "data_parallel")
    Log::Fatal("tree_learner=%s is not supported by the B200 learner (serial, data)", tree_learner.c_str());
  if (device_type != "cuda" && device_type != "gpu")
    Log::Fatal("device_type=%s: this library has no CPU path; use device_type=cuda", device_type.c_str());
  if (linear_tree) Log::Fatal("linear_tree is not supported by the B200 learner");
  if (extra_trees) Log::Fatal("extra_trees is not supported by the B200 learner");
  if (feature_fraction_bynode < 1.0) Log::Fatal("feature_fraction_bynode < 1 is not supported by the B200 learner");
  if (feature_fraction <= 0.0 || feature_fraction > 1.0) Log::Fatal("feature_fraction should be in (0.0, 1.0]");
  if (bagging_freq > 0 && (bagging_fraction < 1.0 || pos_bagging_fraction < 1.0 || neg_bagging_fraction < 1.0)) {
    if (use_quantized_grad)
      Log::Fatal("bagging with use_quantized_grad is not supported by the B200 learner (the reference's bagged quantized "
                 "LeafSplits::Init indexes the gradients by position, leaf_splits.hpp:185)");
    if (tree_learner != "serial") Log::Fatal("bagging with tree_learner=%s is not supported by the B200 learner", tree_learner.c_str());
  }
  if (max_bin > 255) Log::Fatal("max_bin=%d: the B200 learner stores one byte per cell (max_bin <= 255)", max_bin);
  if (num_leaves < 2) Log::Fatal("num_leaves must be at least 2");
  for (const char* k : {"monotone_constraints", "interaction_constraints", "forcedsplits_filename", "categorical_feature",
                        "cegb_tradeoff", "cegb_penalty_split", "cegb_penalty_feature_lazy", "cegb_penalty_feature_coupled",
                        "forcedbins_filename", "max_bin_by_feature", "feature_contri", "init_score_filename"}) {
    if (get(k) != nullptr && !get(k)->empty()) Log::Fatal("parameter %s is not supported by the B200 learner", k);
  }
  // parameters that would change what is trained must not be ignored silently
  if (const std::string* v = get("data_sample_strategy")) {
    if (*v != "bagging") Log::Fatal("data_sample_strategy=%s is not supported by the B200 learner (bagging only)", v->c_str());
  }
  if (const std::string* v = get("num_class")) {
    if (std::stoi(*v) != 1) Log::Fatal("num_class=%s: multiclass training is not supported by the B200 learner", v->c_str());
  }
  // ---- Config::CheckParamConflict (src/io/config.cpp:375-460), the parts that touch this learner
  if (max_depth > 0 && (get("num_leaves") == nullptr || get("num_leaves")->empty())) {
    const double full_num_leaves = std::pow(2, max_depth);  // a depth limit without num_leaves caps num_leaves
    if (full_num_leaves < num_leaves) num_leaves = static_cast<int>(full_num_leaves);
  }
  if (path_smooth > 1e-15f && min_data_in_leaf < 2) {
    min_data_in_leaf = 2;
    Log::Warning("min_data_in_leaf has been increased to 2 because this is required when path smoothing is active.");
  }
  if (min_data_in_leaf <= 0 && min_sum_hessian_in_leaf <= 1e-15f) {
    Log::Warning("Cannot set both min_data_in_leaf and min_sum_hessian_in_leaf to 0. Will set min_data_in_leaf to 1.");
    min_data_in_leaf = 1;
  }
}

std::string Config::ToString() const {
  std::stringstream ss;
  ss << "[boosting: " << boosting << "]\n";
  ss << "[objective: " << objective << "]\n";
  ss << "[metric: " << metric << "]\n";
  ss << "[tree_learner: " << tree_learner << "]\n";
  ss << "[device_type: " << device_type << "]\n";
  ss << "[num_iterations: " << num_iterations << "]\n";
  ss << "[learning_rate: " << learning_rate << "]\n";
  ss << "[num_leaves: " << num_leaves << "]\n";
  ss << "[max_depth: " << max_depth << "]\n";
  ss << "[min_data_in_leaf: " << min_data_in_leaf << "]\n";
  ss << "[min_sum_hessian_in_leaf: " << min_sum_hessian_in_leaf << "]\n";
  ss << "[lambda_l1: " << lambda_l1 << "]\n";
  ss << "[lambda_l2: " << lambda_l2 << "]\n";
  ss << "[min_gain_to_split: " << min_gain_to_split << "]\n";
  ss << "[max_delta_step: " << max_delta_step << "]\n";
  ss << "[path_smooth: " << path_smooth << "]\n";
  ss << "[use_quantized_grad: " << (use_quantized_grad ? 1 : 0) << "]\n";
  ss << "[num_grad_quant_bins: " << num_grad_quant_bins << "]\n";
  ss << "[max_bin: " << max_bin << "]\n";
  ss << "[lambdarank_target: " << lambdarank_target << "]\n";
  ss << "[lambdarank_truncation_level: " << lambdarank_truncation_level << "]\n";
  ss << "[lambdagap_weight: " << lambdagap_weight << "]";
  return ss.str();
}

}  // namespace lgbm_b200
