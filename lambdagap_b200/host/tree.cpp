// tree.cpp -- host Tree model: the output contract of TreeLearner::Train.
// Reference: Tree::Split (include/LightGBM/tree.h:543-583, src/io/tree.cpp:60-74), Shrinkage / AddBias
// (tree.h:186-231), NumericalDecision (tree.h:340-358), Tree::ToString (src/io/tree.cpp:339-395; number
// formatting of CommonC::ArrayToString: "{:g}" for floats, "{:.17g}" for high-precision doubles).
#include <cstring>
#include <sstream>

#include "lgbm_host.h"

namespace lgbm_b200 {

namespace {
const int8_t kCategoricalMask = 1;
const int8_t kDefaultLeftMask = 2;

template <typename T>
std::string JoinInts(const std::vector<T>& v, size_t n) {
  std::string s;
  for (size_t i = 0; i < n && i < v.size(); ++i) {
    if (i) s.push_back(' ');
    s += std::to_string(static_cast<long long>(v[i]));
  }
  return s;
}
template <typename T>
std::string JoinG(const std::vector<T>& v, size_t n, bool high_precision) {
  std::string s;
  char buf[64];
  for (size_t i = 0; i < n && i < v.size(); ++i) {
    if (i) s.push_back(' ');
    snprintf(buf, sizeof(buf), high_precision ? "%.17g" : "%g", static_cast<double>(v[i]));
    s += buf;
  }
  return s;
}
}  // namespace

Tree::Tree(int max_leaves) : max_leaves_(max_leaves) {
  const int n = max_leaves < 2 ? 2 : max_leaves;
  left_child_.resize(n - 1);
  right_child_.resize(n - 1);
  split_feature_inner_.resize(n - 1);
  split_feature_.resize(n - 1);
  threshold_in_bin_.resize(n - 1);
  threshold_.resize(n - 1);
  decision_type_.resize(n - 1, 0);
  split_gain_.resize(n - 1);
  leaf_parent_.resize(n);
  leaf_value_.resize(n);
  leaf_weight_.resize(n);
  leaf_count_.resize(n);
  internal_value_.resize(n - 1);
  internal_weight_.resize(n - 1);
  internal_count_.resize(n - 1);
  leaf_depth_.resize(n);
  leaf_depth_[0] = 0;
  leaf_value_[0] = 0.0;
  leaf_weight_[0] = 0.0;
  leaf_parent_[0] = -1;
}

int Tree::Split(int leaf, int feature, int real_feature, uint32_t threshold_bin, double threshold_double,
                double left_value, double right_value, int left_cnt, int right_cnt, double left_weight,
                double right_weight, float gain, MissingType missing_type, bool default_left) {
  const int node = num_leaves_ - 1;
  const int parent = leaf_parent_[leaf];
  if (parent >= 0) {
    if (left_child_[parent] == ~leaf) {
      left_child_[parent] = node;
    } else {
      right_child_[parent] = node;
    }
  }
  split_feature_inner_[node] = feature;
  split_feature_[node] = real_feature;
  split_gain_[node] = gain;
  left_child_[node] = ~leaf;
  right_child_[node] = ~num_leaves_;
  leaf_parent_[leaf] = node;
  leaf_parent_[num_leaves_] = node;
  internal_weight_[node] = left_weight + right_weight;
  internal_value_[node] = leaf_value_[leaf];
  internal_count_[node] = left_cnt + right_cnt;
  leaf_value_[leaf] = std::isnan(left_value) ? 0.0f : left_value;
  leaf_weight_[leaf] = left_weight;
  leaf_count_[leaf] = left_cnt;
  leaf_value_[num_leaves_] = std::isnan(right_value) ? 0.0f : right_value;
  leaf_weight_[num_leaves_] = right_weight;
  leaf_count_[num_leaves_] = right_cnt;
  leaf_depth_[num_leaves_] = leaf_depth_[leaf] + 1;
  leaf_depth_[leaf]++;
  int8_t dt = 0;
  if (default_left) dt |= kDefaultLeftMask;
  dt |= static_cast<int8_t>(static_cast<int>(missing_type) << 2);
  decision_type_[node] = dt;
  threshold_in_bin_[node] = threshold_bin;
  threshold_[node] = threshold_double;
  ++num_leaves_;
  return num_leaves_ - 1;
}

void Tree::Shrinkage(double rate) {
  for (int i = 0; i < num_leaves_ - 1; ++i) {
    leaf_value_[i] = MaybeRoundToZero(leaf_value_[i] * rate);
    internal_value_[i] = MaybeRoundToZero(internal_value_[i] * rate);
  }
  leaf_value_[num_leaves_ - 1] = MaybeRoundToZero(leaf_value_[num_leaves_ - 1] * rate);
  shrinkage_ *= rate;
}

void Tree::AddBias(double val) {
  for (int i = 0; i < num_leaves_ - 1; ++i) {
    leaf_value_[i] = MaybeRoundToZero(leaf_value_[i] + val);
    internal_value_[i] = MaybeRoundToZero(internal_value_[i] + val);
  }
  leaf_value_[num_leaves_ - 1] = MaybeRoundToZero(leaf_value_[num_leaves_ - 1] + val);
  shrinkage_ = 1.0f;
}

void Tree::AsConstantTree(double val, int count) {
  num_leaves_ = 1;
  shrinkage_ = 1.0f;
  leaf_value_[0] = val;
  leaf_count_[0] = count;
}

double Tree::Predict(const double* feature_values) const {
  if (num_leaves_ <= 1) return leaf_value_[0];
  int node = 0;
  while (node >= 0) {
    double fval = feature_values[split_feature_[node]];
    const int missing_type = (decision_type_[node] >> 2) & 3;
    if (std::isnan(fval) && missing_type != kMissingNaN) fval = 0.0f;
    const bool is_zero = fval >= -kZeroThreshold && fval <= kZeroThreshold;
    if ((missing_type == kMissingZero && is_zero) || (missing_type == kMissingNaN && std::isnan(fval))) {
      node = (decision_type_[node] & kDefaultLeftMask) ? left_child_[node] : right_child_[node];
    } else if (fval <= threshold_[node]) {
      node = left_child_[node];
    } else {
      node = right_child_[node];
    }
  }
  return leaf_value_[~node];
}

std::string Tree::ToString() const {
  std::stringstream ss;
  const size_t ni = static_cast<size_t>(num_leaves_ - 1), nl = static_cast<size_t>(num_leaves_);
  std::vector<int> dt(decision_type_.begin(), decision_type_.end());
  ss << "num_leaves=" << num_leaves_ << '\n';
  ss << "num_cat=0" << '\n';
  ss << "split_feature=" << JoinInts(split_feature_, ni) << '\n';
  ss << "split_gain=" << JoinG(split_gain_, ni, false) << '\n';
  ss << "threshold=" << JoinG(threshold_, ni, true) << '\n';
  ss << "decision_type=" << JoinInts(dt, ni) << '\n';
  ss << "left_child=" << JoinInts(left_child_, ni) << '\n';
  ss << "right_child=" << JoinInts(right_child_, ni) << '\n';
  ss << "leaf_value=" << JoinG(leaf_value_, nl, true) << '\n';
  ss << "leaf_weight=" << JoinG(leaf_weight_, nl, true) << '\n';
  ss << "leaf_count=" << JoinInts(leaf_count_, nl) << '\n';
  ss << "internal_value=" << JoinG(internal_value_, ni, false) << '\n';
  ss << "internal_weight=" << JoinG(internal_weight_, ni, false) << '\n';
  ss << "internal_count=" << JoinInts(internal_count_, ni) << '\n';
  ss << "is_linear=0" << '\n';
  ss << "shrinkage=" << JoinG(std::vector<double>{shrinkage_}, 1, false) << '\n';
  ss << '\n';
  return ss.str();
}

}  // namespace lgbm_b200
