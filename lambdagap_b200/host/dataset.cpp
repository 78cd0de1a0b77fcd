// dataset.cpp -- dense-matrix Dataset construction.
//
// Reference flow restated: LGBM_DatasetCreateFromMats (src/c_api.cpp:1322-1405) samples
// bin_construct_sample_cnt rows (CreateSampleIndices, :982-986), keeps the non-zero sample values of every
// column, DatasetLoader::ConstructFromSampleData (src/io/dataset_loader.cpp:593-760) finds one BinMapper per
// column and drops trivial columns, then every row is pushed through ValueToBin.  Here the last step runs on
// the GPU (gbl_dataset_create_from_mat_f32) for float32 row-major input.
#include <cstring>

#include "lgbm_host.h"

namespace lgbm_b200 {

namespace {
const int kDtypeFloat32 = 0, kDtypeFloat64 = 1, kDtypeInt32 = 2;  // include/LightGBM/c_api.h:30-33

inline double Cell(const void* data, int data_type, int64_t nrow, int64_t ncol, int is_row_major, int64_t r, int64_t c) {
  const int64_t off = is_row_major ? r * ncol + c : c * nrow + r;
  return data_type == kDtypeFloat32 ? static_cast<double>(static_cast<const float*>(data)[off])
                                    : static_cast<const double*>(data)[off];
}
}  // namespace

Dataset::~Dataset() {
  if (device_ != nullptr) gbl_dataset_free(device_);
}

Dataset* Dataset::CreateFromMat(const void* data, int data_type, int32_t nrow, int32_t ncol, int is_row_major,
                                const Config& config, const Dataset* reference) {
  return CreateFromMatWithSample(data, data_type, nrow, ncol, is_row_major, nullptr, 0, nrow, config,
                                 reference);
}

Dataset* Dataset::CreateFromMatWithSample(const void* data, int data_type, int32_t nrow, int32_t ncol, int is_row_major,
                                          const void* sample, int32_t sample_nrow, int64_t num_total_rows,
                                          const Config& config) {
  return CreateFromMatWithSample(data, data_type, nrow, ncol, is_row_major, sample, sample_nrow, num_total_rows, config,
                                 nullptr);
}

Dataset* Dataset::CreateFromMatWithSample(const void* data, int data_type, int32_t nrow, int32_t ncol, int is_row_major,
                                          const void* sample, int32_t sample_nrow, int64_t num_total_rows,
                                          const Config& config, const Dataset* reference) {
  if (data == nullptr) Log::Fatal("LGBM_DatasetCreateFromMat: data is null");
  if (data_type != kDtypeFloat32 && data_type != kDtypeFloat64) Log::Fatal("Unknown data type in RowFunctionFromDenseMatric");
  if (nrow <= 0 || ncol <= 0) Log::Fatal("LGBM_DatasetCreateFromMat: empty matrix");
  std::unique_ptr<Dataset> ds(new Dataset());
  ds->num_data_ = nrow;
  ds->num_total_features_ = ncol;
  ds->device_id_ = config.gpu_device_id < 0 ? 0 : config.gpu_device_id;

  if (reference == nullptr) {
    std::vector<int> sample_indices;
    const void* src = data;
    int32_t src_rows = nrow;
    if (sample != nullptr) {  // explicit sample: every row of it
      src = sample;
      src_rows = sample_nrow;
      for (int i = 0; i < sample_nrow; ++i) sample_indices.push_back(i);
    } else {
      Random rand(config.data_random_seed);
      const int sample_cnt = nrow < config.bin_construct_sample_cnt ? nrow : config.bin_construct_sample_cnt;
      sample_indices = rand.Sample(nrow, sample_cnt);
    }
    const size_t total_sample_size = sample_indices.size();
    // dataset_loader.cpp:623-624
    const data_size_t filter_cnt = static_cast<data_size_t>(
        static_cast<double>(config.min_data_in_leaf * total_sample_size) / num_total_rows);
    ds->all_bin_mappers_.resize(ncol);
#pragma omp parallel for schedule(guided)
    for (int c = 0; c < ncol; ++c) {
      std::vector<double> vals;
      vals.reserve(total_sample_size);
      for (size_t i = 0; i < total_sample_size; ++i) {
        const double v = Cell(src, data_type, src_rows, ncol, is_row_major, sample_indices[i], c);
        if (std::fabs(v) > kZeroThreshold || std::isnan(v)) vals.push_back(v);
      }
      std::shared_ptr<BinMapper> m(new BinMapper());
      m->FindBin(vals.data(), static_cast<int>(vals.size()), total_sample_size, config.max_bin, config.min_data_in_bin,
                 filter_cnt, config.feature_pre_filter, config.use_missing, config.zero_as_missing);
      ds->all_bin_mappers_[c] = m;
    }
  } else {
    if (reference->num_total_features_ != ncol) Log::Fatal("Validation data has %d columns, training data %d", ncol, reference->num_total_features_);
    ds->all_bin_mappers_ = reference->all_bin_mappers_;
  }
  ds->used_feature_map_.assign(ncol, -1);
  for (int c = 0; c < ncol; ++c) {
    if (!ds->all_bin_mappers_[c]->is_trivial()) {
      ds->used_feature_map_[c] = static_cast<int>(ds->used_feature_map_inv_.size());
      ds->used_feature_map_inv_.push_back(c);
      ds->bin_mappers_.emplace_back(new BinMapper(*ds->all_bin_mappers_[c]));
    }
  }
  const int nf = ds->num_features();
  if (nf == 0) Log::Fatal("There are no meaningful features which satisfy the provided configuration");

  std::vector<gbl_feature_meta> metas(nf);
  std::vector<double> bounds;
  for (int f = 0; f < nf; ++f) {
    const BinMapper* m = ds->bin_mappers_[f].get();
    metas[f].num_bin = m->num_bin();
    metas[f].missing_type = static_cast<int>(m->missing_type());
    metas[f].default_bin = static_cast<int>(m->GetDefaultBin());
    metas[f].most_freq_bin = static_cast<int>(m->GetMostFreqBin());
    metas[f].real_index = ds->used_feature_map_inv_[f];
    metas[f].monotone_type = 0;  // monotone constraints are refused by this host layer's Config
    metas[f].penalty = 1.0;
    bounds.insert(bounds.end(), m->bin_upper_bound().begin(), m->bin_upper_bound().end());
  }
  if (data_type == kDtypeFloat32 && is_row_major) {
    GblCheck(gbl_dataset_create_from_mat_f32(metas.data(), nf, nrow, static_cast<const float*>(data), ncol,
                                             bounds.data(), ds->device_id_, &ds->device_));
  } else {
    // float64 / column-major input: bin on the host with the same mappers, then upload the byte columns
    std::vector<std::vector<uint8_t>> cols(nf, std::vector<uint8_t>(nrow));
#pragma omp parallel for schedule(static)
    for (int f = 0; f < nf; ++f) {
      const BinMapper* m = ds->bin_mappers_[f].get();
      const int c = ds->used_feature_map_inv_[f];
      for (int32_t r = 0; r < nrow; ++r) {
        cols[f][r] = static_cast<uint8_t>(m->ValueToBin(Cell(data, data_type, nrow, ncol, is_row_major, r, c)));
      }
    }
    std::vector<const uint8_t*> ptrs(nf);
    for (int f = 0; f < nf; ++f) ptrs[f] = cols[f].data();
    GblCheck(gbl_dataset_create_from_bins(metas.data(), nf, nrow, ptrs.data(), ds->device_id_, &ds->device_));
  }
  return ds.release();
}

void Dataset::SetField(const char* name, const void* data, int num_element, int type) {
  // Dataset::SetFloatField / SetIntField (src/io/dataset.cpp:977-1040), Metadata::SetLabel/SetWeights/SetQuery
  const std::string n(name);
  if (n == "label" || n == "target") {
    if (type != kDtypeFloat32) Log::Fatal("Input data type error or field %s should be float32", name);
    if (num_element != num_data_) Log::Fatal("Length of label is not same with #data");
    const float* p = static_cast<const float*>(data);
    metadata_.label.assign(p, p + num_element);
  } else if (n == "weight" || n == "weights") {
    if (type != kDtypeFloat32) Log::Fatal("Input data type error or field %s should be float32", name);
    if (num_element == 0) {
      metadata_.weights.clear();
      return;
    }
    if (num_element != num_data_) Log::Fatal("Length of weights is not same with #data");
    const float* p = static_cast<const float*>(data);
    metadata_.weights.assign(p, p + num_element);
  } else if (n == "group" || n == "query") {
    if (type != kDtypeInt32) Log::Fatal("Input data type error or field %s should be int32", name);
    const int32_t* p = static_cast<const int32_t*>(data);
    int64_t sum = 0;
    for (int i = 0; i < num_element; ++i) sum += p[i];
    if (sum != num_data_) Log::Fatal("Sum of query counts (%lld) differs from the length of #data (%d)", (long long)sum, num_data_);
    metadata_.query_boundaries.resize(num_element + 1);
    metadata_.query_boundaries[0] = 0;
    for (int i = 0; i < num_element; ++i) metadata_.query_boundaries[i + 1] = metadata_.query_boundaries[i] + p[i];
  } else {
    Log::Fatal("Field %s is not supported by the B200 host layer (label, weight, group)", name);
  }
}

std::vector<std::string> Dataset::feature_names() const {
  std::vector<std::string> out;
  for (int i = 0; i < num_total_features_; ++i) out.push_back("Column_" + std::to_string(i));
  return out;
}

std::vector<std::string> Dataset::feature_infos() const {
  std::vector<std::string> out;
  for (int i = 0; i < num_total_features_; ++i) {
    out.push_back(used_feature_map_[i] < 0 ? std::string("none") : all_bin_mappers_[i]->bin_info_string());
  }
  return out;
}

}  // namespace lgbm_b200
