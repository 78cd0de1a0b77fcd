// bin_mapper.cpp -- value -> bin mapping for numerical features, bit-exact with the reference.
//
// Restates (does not copy) BinMapper::FindBin (src/io/bin.cpp:313-508) with its helpers GreedyFindBin (:78-157),
// FindBinWithZeroAsOneBin (:244-300) and NeedFilter (:54-76), BinMapper::ValueToBin
// (include/LightGBM/bin.h:612-649) and Random::Sample (include/LightGBM/utils/random.h:68-98).
// Bin boundaries are midpoints of neighbouring distinct sample values pushed one ulp up (nextafter), zero gets a
// bin of its own [-1e-35, 1e-35], NaN (when present and use_missing) the last bin.
#include <algorithm>
#include <cmath>
#include <set>
#include <sstream>
#include <iomanip>

#include "lgbm_host.h"

namespace lgbm_b200 {

namespace {

inline double NextUp(double a) { return std::nextafter(a, std::numeric_limits<double>::infinity()); }
// Common::CheckDoubleEqualOrdered (include/LightGBM/utils/common.h:856-859): b <= a + 1ulp
inline bool SameOrdered(double a, double b) { return b <= NextUp(a); }

struct Distinct {
  std::vector<double> value;
  std::vector<int> count;
  int size() const { return static_cast<int>(value.size()); }
};

// Equal-frequency style boundaries over `n` distinct values (GreedyFindBin).
std::vector<double> GreedyBounds(const double* value, const int* count, int n, int max_bin, size_t total_cnt,
                                 int min_data_in_bin) {
  std::vector<double> bounds;
  const double inf = std::numeric_limits<double>::infinity();
  if (n <= max_bin) {
    // few distinct values: one boundary between neighbours once a bin holds min_data_in_bin samples
    int in_bin = 0;
    for (int i = 0; i + 1 < n; ++i) {
      in_bin += count[i];
      if (in_bin >= min_data_in_bin) {
        const double mid = NextUp((value[i] + value[i + 1]) / 2.0);
        if (bounds.empty() || !SameOrdered(bounds.back(), mid)) {
          bounds.push_back(mid);
          in_bin = 0;
        }
      }
    }
    bounds.push_back(inf);
    return bounds;
  }
  if (min_data_in_bin > 0) {
    max_bin = std::min(max_bin, static_cast<int>(total_cnt / min_data_in_bin));
    max_bin = std::max(max_bin, 1);
  }
  double mean_size = static_cast<double>(total_cnt) / max_bin;
  // values that alone fill a bin get one; the rest share what is left
  int bins_left = max_bin;
  int samples_left = static_cast<int>(total_cnt);
  std::vector<bool> heavy(n, false);
  for (int i = 0; i < n; ++i) {
    if (count[i] >= mean_size) {
      heavy[i] = true;
      --bins_left;
      samples_left -= count[i];
    }
  }
  mean_size = static_cast<double>(samples_left) / bins_left;
  std::vector<double> upper(max_bin, inf), lower(max_bin, inf);
  int nb = 0;
  lower[0] = value[0];
  int in_bin = 0;
  for (int i = 0; i + 1 < n; ++i) {
    if (!heavy[i]) samples_left -= count[i];
    in_bin += count[i];
    const bool close_bin = heavy[i] || in_bin >= mean_size ||
                           (heavy[i + 1] && in_bin >= std::max(1.0, mean_size * 0.5f));
    if (close_bin) {
      upper[nb] = value[i];
      ++nb;
      lower[nb] = value[i + 1];
      if (nb >= max_bin - 1) break;
      in_bin = 0;
      if (!heavy[i]) {
        --bins_left;
        mean_size = samples_left / static_cast<double>(bins_left);
      }
    }
  }
  ++nb;
  for (int i = 0; i + 1 < nb; ++i) {
    const double mid = NextUp((upper[i] + lower[i + 1]) / 2.0);
    if (bounds.empty() || !SameOrdered(bounds.back(), mid)) bounds.push_back(mid);
  }
  bounds.push_back(inf);
  return bounds;
}

// Negative side, a dedicated zero bin, positive side (FindBinWithZeroAsOneBin).
std::vector<double> BoundsWithZeroBin(const Distinct& d, int max_bin, size_t total_sample_cnt, int min_data_in_bin) {
  const int n = d.size();
  int neg_cnt = 0, zero_cnt = 0, pos_cnt = 0;
  for (int i = 0; i < n; ++i) {
    if (d.value[i] <= -kZeroThreshold) {
      neg_cnt += d.count[i];
    } else if (d.value[i] > kZeroThreshold) {
      pos_cnt += d.count[i];
    } else {
      zero_cnt += d.count[i];
    }
  }
  int num_neg = n;  // distinct values strictly below the zero band
  for (int i = 0; i < n; ++i) {
    if (d.value[i] > -kZeroThreshold) {
      num_neg = i;
      break;
    }
  }
  std::vector<double> bounds;
  if (num_neg > 0 && max_bin > 1) {
    int neg_bins = static_cast<int>(static_cast<double>(neg_cnt) / (total_sample_cnt - zero_cnt) * (max_bin - 1));
    neg_bins = std::max(1, neg_bins);
    bounds = GreedyBounds(d.value.data(), d.count.data(), num_neg, neg_bins, neg_cnt, min_data_in_bin);
    if (!bounds.empty()) bounds.back() = -kZeroThreshold;
  }
  int pos_start = -1;
  for (int i = num_neg; i < n; ++i) {
    if (d.value[i] > kZeroThreshold) {
      pos_start = i;
      break;
    }
  }
  const int pos_bins = max_bin - 1 - static_cast<int>(bounds.size());
  if (pos_start >= 0 && pos_bins > 0) {
    std::vector<double> right = GreedyBounds(d.value.data() + pos_start, d.count.data() + pos_start, n - pos_start,
                                             pos_bins, pos_cnt, min_data_in_bin);
    bounds.push_back(kZeroThreshold);
    bounds.insert(bounds.end(), right.begin(), right.end());
  } else {
    bounds.push_back(std::numeric_limits<double>::infinity());
  }
  if (bounds.size() > static_cast<size_t>(max_bin)) Log::Fatal("BinMapper: more bins than max_bin");
  return bounds;
}

// No threshold leaves at least filter_cnt samples on both sides -> the feature cannot be split (NeedFilter).
bool CannotSplit(const std::vector<int>& cnt_in_bin, int total_cnt, int filter_cnt) {
  int left = 0;
  for (size_t i = 0; i + 1 < cnt_in_bin.size(); ++i) {
    left += cnt_in_bin[i];
    if (left >= filter_cnt && total_cnt - left >= filter_cnt) return false;
  }
  return true;
}

}  // namespace

void BinMapper::FindBin(double* values, int num_sample_values, size_t total_sample_cnt, int max_bin,
                        int min_data_in_bin, int min_split_data, bool pre_filter, bool use_missing,
                        bool zero_as_missing) {
  // drop NaNs to the front-compacted prefix
  int kept = 0;
  for (int i = 0; i < num_sample_values; ++i) {
    if (!std::isnan(values[i])) values[kept++] = values[i];
  }
  int na_cnt = 0;
  if (!use_missing) {
    missing_type_ = kMissingNone;
  } else if (zero_as_missing) {
    missing_type_ = kMissingZero;
  } else if (kept == num_sample_values) {
    missing_type_ = kMissingNone;
  } else {
    missing_type_ = kMissingNaN;
    na_cnt = num_sample_values - kept;
  }
  num_sample_values = kept;
  default_bin_ = 0;
  const int zero_cnt = static_cast<int>(total_sample_cnt - num_sample_values - na_cnt);

  std::stable_sort(values, values + num_sample_values);
  // distinct values (neighbours within one ulp merge, keeping the larger), with the implied zeros spliced in
  Distinct d;
  auto push = [&d](double v, int c) {
    d.value.push_back(v);
    d.count.push_back(c);
  };
  if (num_sample_values == 0 || (values[0] > 0.0f && zero_cnt > 0)) push(0.0f, zero_cnt);
  if (num_sample_values > 0) push(values[0], 1);
  for (int i = 1; i < num_sample_values; ++i) {
    if (!SameOrdered(values[i - 1], values[i])) {
      if (values[i - 1] < 0.0f && values[i] > 0.0f) push(0.0f, zero_cnt);
      push(values[i], 1);
    } else {
      d.value.back() = values[i];
      ++d.count.back();
    }
  }
  if (num_sample_values > 0 && values[num_sample_values - 1] < 0.0f && zero_cnt > 0) push(0.0f, zero_cnt);
  min_val_ = d.value.front();
  max_val_ = d.value.back();

  if (missing_type_ == kMissingNaN) {
    bin_upper_bound_ = BoundsWithZeroBin(d, max_bin - 1, total_sample_cnt - na_cnt, min_data_in_bin);
    bin_upper_bound_.push_back(std::numeric_limits<double>::quiet_NaN());
  } else {
    bin_upper_bound_ = BoundsWithZeroBin(d, max_bin, total_sample_cnt, min_data_in_bin);
    if (missing_type_ == kMissingZero && bin_upper_bound_.size() == 2) missing_type_ = kMissingNone;
  }
  num_bin_ = static_cast<int>(bin_upper_bound_.size());
  std::vector<int> cnt_in_bin(num_bin_, 0);
  {
    int b = 0;
    for (int i = 0; i < d.size(); ++i) {
      while (d.value[i] > bin_upper_bound_[b] && b < num_bin_ - 1) ++b;
      cnt_in_bin[b] += d.count[i];
    }
    if (missing_type_ == kMissingNaN) cnt_in_bin[num_bin_ - 1] = na_cnt;
  }
  if (num_bin_ > max_bin) Log::Fatal("BinMapper: num_bin %d exceeds max_bin %d", num_bin_, max_bin);

  is_trivial_ = num_bin_ <= 1;
  if (!is_trivial_ && pre_filter && CannotSplit(cnt_in_bin, static_cast<int>(total_sample_cnt), min_split_data)) {
    is_trivial_ = true;
  }
  if (!is_trivial_) {
    default_bin_ = ValueToBin(0);
    // ArrayArgs<int>::ArgMax: first maximum
    int arg = 0;
    for (int i = 1; i < num_bin_; ++i) {
      if (cnt_in_bin[i] > cnt_in_bin[arg]) arg = i;
    }
    most_freq_bin_ = static_cast<uint32_t>(arg);
    const double max_sparse_rate = static_cast<double>(cnt_in_bin[most_freq_bin_]) / total_sample_cnt;
    // a most-frequent bin other than the zero bin only pays off when the feature is sparse enough
    if (most_freq_bin_ != default_bin_ && max_sparse_rate < kSparseThreshold) most_freq_bin_ = default_bin_;
    sparse_rate_ = static_cast<double>(cnt_in_bin[most_freq_bin_]) / total_sample_cnt;
  } else {
    sparse_rate_ = 1.0f;
  }
}

uint32_t BinMapper::ValueToBin(double value) const {
  if (std::isnan(value)) {
    if (missing_type_ == kMissingNaN) return num_bin_ - 1;
    value = 0.0f;
  }
  int l = 0;
  int r = num_bin_ - 1;
  if (missing_type_ == kMissingNaN) r -= 1;
  while (l < r) {
    const int m = (r + l - 1) / 2;
    if (value <= bin_upper_bound_[m]) {
      r = m;
    } else {
      l = m + 1;
    }
  }
  return l;
}

std::string BinMapper::bin_info_string() const {
  std::stringstream ss;
  ss << std::setprecision(std::numeric_limits<double>::digits10 + 2);
  ss << '[' << min_val_ << ':' << max_val_ << ']';
  return ss.str();
}

std::vector<int> Random::Sample(int N, int K) {
  std::vector<int> ret;
  ret.reserve(K);
  if (K > N || K <= 0) return ret;
  if (K == N) {
    for (int i = 0; i < N; ++i) ret.push_back(i);
  } else if (K > 1 && K > (N / std::log2(K))) {
    for (int i = 0; i < N; ++i) {
      const double prob = (K - ret.size()) / static_cast<double>(N - i);
      if (NextFloat() < prob) ret.push_back(i);
    }
  } else {
    std::set<int> chosen;
    for (int r = N - K; r < N; ++r) {
      const int v = NextInt(0, r + 1);
      if (!chosen.insert(v).second) chosen.insert(r);
    }
    ret.assign(chosen.begin(), chosen.end());
  }
  return ret;
}

}  // namespace lgbm_b200
