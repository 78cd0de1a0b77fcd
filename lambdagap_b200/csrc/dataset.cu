// dataset.cu -- GPU-resident binned feature storage (src/io scope of the north star).
//
// Reference behaviour restated: BinMapper::ValueToBin (include/LightGBM/bin.h:612-649) decides the bin of a
// cell; FeatureGroup::PushData / DenseBin (include/LightGBM/feature_group.h:253-267, src/io/dense_bin.hpp)
// store it.  Here the *mapper* bin itself is stored (no most-frequent-bin elision), one byte per cell,
// twice: column-major (canonical) and as a row-major mirror tiled by 32 features for index gathers.
#include <cmath>
#include <cstring>
#include <vector>

#include "common.cuh"

namespace gbl {

// ---------------------------------------------------------------------------------------------- kernels
// 32x32 byte tile transpose: col_bins[f][r] -> row_bins[r][f].  One CTA = 32 features x 128 rows.
__global__ void __launch_bounds__(256) transpose_to_rows_kernel(const uint8_t* __restrict__ col_bins, int64_t col_stride,
                                                                uint8_t* __restrict__ row_bins, int row_stride,
                                                                int64_t num_data, int num_features) {
  __shared__ uint32_t tile[32][33];  // [feature][row word]
  const int tile_f = blockIdx.y;
  const int64_t row0 = static_cast<int64_t>(blockIdx.x) * 128;
  const int lane = threadIdx.x & 31;
  const int warp = threadIdx.x >> 5;
  // load: each warp reads 4 features x 128 rows as 32-bit words (lane = word within the 128-byte row span)
  for (int ff = warp; ff < 32; ff += 8) {
    const int f = tile_f * 32 + ff;
    uint32_t w = 0;
    const int64_t r = row0 + lane * 4;
    if (f < num_features && r < num_data) {
      // col_stride is a multiple of 128 and the allocation is padded, so the 4-byte read is in bounds
      w = *reinterpret_cast<const uint32_t*>(col_bins + f * col_stride + r);
    }
    tile[ff][lane] = w;
  }
  __syncthreads();
  // store: thread -> (row, 4-feature group); 128 rows x 8 groups = 1024 items
  for (int item = threadIdx.x; item < 128 * 8; item += 256) {
    const int r = item >> 3;
    const int g = item & 7;
    const int64_t row = row0 + r;
    if (row >= num_data) continue;
    uint32_t out = 0;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const uint32_t w = tile[g * 4 + k][r >> 2];
      out |= ((w >> ((r & 3) * 8)) & 0xffu) << (8 * k);
    }
    *reinterpret_cast<uint32_t*>(row_bins + row * row_stride + tile_f * 32 + g * 4) = out;
  }
}

// BinMapper::ValueToBin for numerical features, one thread per cell of a row-major float matrix chunk.
// values: [rows x num_total_columns] float32; bounds: concatenated upper bounds; out: col_bins.
__global__ void __launch_bounds__(256) value_to_bin_kernel(const float* __restrict__ values, int64_t row_begin,
                                                           int64_t rows, int num_total_columns,
                                                           const FeatMeta* __restrict__ meta,
                                                           const int* __restrict__ bound_offset,
                                                           const double* __restrict__ bounds, int num_features,
                                                           uint8_t* __restrict__ col_bins, int64_t col_stride) {
  // thread -> (feature, row) with rows fastest so that column writes coalesce
  const int f = blockIdx.y;
  const int64_t r = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (r >= rows) return;
  const FeatMeta m = meta[f];
  double value = static_cast<double>(values[r * num_total_columns + m.real_index]);
  int bin;
  bool done = false;
  if (isnan(value)) {
    if (m.missing_type == GBL_MISSING_NAN) {
      bin = m.num_bin - 1;
      done = true;
    } else {
      value = 0.0;
    }
  }
  if (!done) {
    const double* ub = bounds + bound_offset[f];
    int l = 0;
    int rr = m.num_bin - 1;
    if (m.missing_type == GBL_MISSING_NAN) rr -= 1;
    while (l < rr) {
      const int mid = (rr + l - 1) / 2;
      if (value <= ub[mid]) {
        rr = mid;
      } else {
        l = mid + 1;
      }
    }
    bin = l;
  }
  col_bins[f * col_stride + row_begin + r] = static_cast<uint8_t>(bin);
}

// ---------------------------------------------------------------------------------------------- host side
static void check_metas(const gbl_feature_meta* metas, int num_features) {
  if (metas == nullptr || num_features <= 0) fatal("dataset: no features");
  for (int f = 0; f < num_features; ++f) {
    if (metas[f].num_bin < 1 || metas[f].num_bin > kMaxBin) {
      fatal("dataset: feature %d has num_bin=%d; this learner supports 1..%d bins per feature (max_bin<=255)", f,
            metas[f].num_bin, kMaxBin);
    }
    if (metas[f].missing_type < 0 || metas[f].missing_type > 2) fatal("dataset: bad missing_type for feature %d", f);
  }
}

static Dataset* alloc_dataset(const gbl_feature_meta* metas, int num_features, int64_t num_data, int device) {
  check_metas(metas, num_features);
  if (num_data <= 0 || num_data >= (1LL << 31)) fatal("dataset: num_data=%lld out of range", (long long)num_data);
  int count = 0;
  GBL_CUDA(cudaGetDeviceCount(&count));
  if (device < 0 || device >= count) fatal("dataset: device %d not available (%d devices)", device, count);
  DeviceGuard guard(device);
  Dataset* ds = new Dataset();
  ds->device = device;
  ds->num_data = num_data;
  ds->num_features = num_features;
  ds->num_tiles = (num_features + kTile - 1) / kTile;
  ds->row_stride = ds->num_tiles * kTile;
  // build-time staging: one byte per cell, uniform column stride (freed once the packed store exists)
  ds->stage_stride = round_up(num_data, 128) + 128;
  GBL_CUDA(cudaMalloc(&ds->stage, static_cast<size_t>(ds->stage_stride) * num_features));
  GBL_CUDA(cudaMemset(ds->stage, 0, static_cast<size_t>(ds->stage_stride) * num_features));
  GBL_CUDA(cudaMalloc(&ds->row_bins, static_cast<size_t>(ds->row_stride) * (num_data + 8)));
  ds->meta_host = new FeatMeta[num_features];
  for (int f = 0; f < num_features; ++f) {
    FeatMeta& m = ds->meta_host[f];
    m.num_bin = metas[f].num_bin;
    m.missing_type = metas[f].missing_type;
    m.default_bin = metas[f].default_bin;
    m.most_freq_bin = metas[f].most_freq_bin;
    m.real_index = metas[f].real_index;
    m.offset = metas[f].most_freq_bin == 0 ? 1 : 0;
    m.penalty = metas[f].penalty;
    m.monotone_type = metas[f].monotone_type;
    m.is_categorical = metas[f].bin_type == 1 ? 1 : 0;
    // the reference packs two rows per byte when a feature has at most 16 bins (bin.cpp:615-625 -> DenseBin<.., true>)
    m.bits = metas[f].num_bin <= 16 ? 4 : 8;
    m.col_off = ds->col_bytes;
    ds->col_bytes += round_up(m.bits == 4 ? (num_data + 1) / 2 : num_data, 128) + 128;
  }
  GBL_CUDA(cudaMalloc(&ds->meta_dev, sizeof(FeatMeta) * num_features));
  GBL_CUDA(cudaMemcpy(ds->meta_dev, ds->meta_host, sizeof(FeatMeta) * num_features, cudaMemcpyHostToDevice));
  return ds;
}

static void build_row_mirror(Dataset* ds) {
  dim3 grid(static_cast<unsigned>((ds->num_data + 127) / 128), ds->num_tiles);
  GBL_LAUNCH(transpose_to_rows_kernel, grid, 256, 0, 0, ds->stage, ds->stage_stride, ds->row_bins, ds->row_stride,
             ds->num_data, ds->num_features);
  GBL_CUDA(cudaDeviceSynchronize());
}

// staging (one byte per cell) -> the column store: 8-bit columns copied, 4-bit columns two rows per byte with the
// even row in the low nibble (dense_bin.hpp:69-82)
__global__ void __launch_bounds__(256) pack_columns_kernel(const uint8_t* __restrict__ stage, int64_t stage_stride,
                                                           const FeatMeta* __restrict__ meta, int64_t n,
                                                           uint8_t* __restrict__ store) {
  const int f = blockIdx.y;
  const FeatMeta m = meta[f];
  const uint8_t* src = stage + f * stage_stride;
  uint8_t* dst = store + m.col_off;
  const int64_t stride = static_cast<int64_t>(gridDim.x) * blockDim.x;
  if (m.bits == 4) {
    const int64_t bytes = (n + 1) / 2;
    for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < bytes; i += stride) {
      const unsigned lo = src[2 * i] & 0xf;
      const unsigned hi = 2 * i + 1 < n ? (src[2 * i + 1] & 0xf) : 0u;
      dst[i] = static_cast<uint8_t>(lo | (hi << 4));
    }
  } else {
    for (int64_t i = static_cast<int64_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) dst[i] = src[i];
  }
}

static void build_column_store(Dataset* ds) {
  GBL_CUDA(cudaMalloc(&ds->col_bins, static_cast<size_t>(ds->col_bytes)));
  GBL_CUDA(cudaMemset(ds->col_bins, 0, static_cast<size_t>(ds->col_bytes)));
  dim3 grid(kNumSM * 2, ds->num_features);
  GBL_LAUNCH(pack_columns_kernel, grid, 256, 0, 0, ds->stage, ds->stage_stride, ds->meta_dev, ds->num_data, ds->col_bins);
  GBL_CUDA(cudaDeviceSynchronize());
  cudaFree(ds->stage);
  ds->stage = nullptr;
}

static void free_dataset(Dataset* ds) {
  if (ds == nullptr) return;
  DeviceGuard guard(ds->device);
  cudaFree(ds->col_bins);
  cudaFree(ds->stage);
  cudaFree(ds->row_bins);
  cudaFree(ds->meta_dev);
  delete[] ds->meta_host;
  delete ds;
}

}  // namespace gbl

using namespace gbl;

extern "C" {

int gbl_dataset_create_from_bins(const gbl_feature_meta* metas, int32_t num_features, int64_t num_data,
                                 const uint8_t* const* columns_host, int32_t device_id, gbl_dataset** out) {
  GBL_API_BEGIN();
  if (out == nullptr || columns_host == nullptr) fatal("gbl_dataset_create_from_bins: null argument");
  Dataset* ds = alloc_dataset(metas, num_features, num_data, device_id);
  try {
    DeviceGuard guard(device_id);
    for (int f = 0; f < num_features; ++f) {
      GBL_CUDA(cudaMemcpy(ds->stage + f * ds->stage_stride, columns_host[f], static_cast<size_t>(num_data),
                          cudaMemcpyHostToDevice));
    }
    build_row_mirror(ds);
    build_column_store(ds);
  } catch (...) {
    free_dataset(ds);
    throw;
  }
  *out = reinterpret_cast<gbl_dataset*>(ds);
  GBL_API_END();
}

int gbl_dataset_create_from_mat_f32(const gbl_feature_meta* metas, int32_t num_features, int64_t num_data,
                                    const float* values_host, int32_t num_total_columns,
                                    const double* upper_bounds_host, int32_t device_id, gbl_dataset** out) {
  GBL_API_BEGIN();
  if (out == nullptr || values_host == nullptr || upper_bounds_host == nullptr)
    fatal("gbl_dataset_create_from_mat_f32: null argument");
  Dataset* ds = alloc_dataset(metas, num_features, num_data, device_id);
  int* off_dev = nullptr;
  double* bounds_dev = nullptr;
  float* chunk_dev = nullptr;
  try {
    DeviceGuard guard(device_id);
    std::vector<int> offs(num_features);
    int total = 0;
    for (int f = 0; f < num_features; ++f) {
      if (metas[f].real_index < 0 || metas[f].real_index >= num_total_columns)
        fatal("dataset: real_index of feature %d out of range", f);
      offs[f] = total;
      total += metas[f].num_bin;
    }
    GBL_CUDA(cudaMalloc(&off_dev, sizeof(int) * num_features));
    GBL_CUDA(cudaMemcpy(off_dev, offs.data(), sizeof(int) * num_features, cudaMemcpyHostToDevice));
    GBL_CUDA(cudaMalloc(&bounds_dev, sizeof(double) * total));
    GBL_CUDA(cudaMemcpy(bounds_dev, upper_bounds_host, sizeof(double) * total, cudaMemcpyHostToDevice));
    // stream the matrix through a bounded staging buffer (<= 1 GiB)
    const int64_t rows_per_chunk =
        std::max<int64_t>(1, std::min<int64_t>(num_data, (1LL << 30) / (sizeof(float) * num_total_columns)));
    GBL_CUDA(cudaMalloc(&chunk_dev, sizeof(float) * rows_per_chunk * num_total_columns));
    for (int64_t r0 = 0; r0 < num_data; r0 += rows_per_chunk) {
      const int64_t rows = std::min(rows_per_chunk, num_data - r0);
      // cudaMemcpyDefault: the matrix may also live in device (or managed) memory -- a caller that produced it on the
      // GPU hands it over without a round trip through the host
      GBL_CUDA(cudaMemcpy(chunk_dev, values_host + r0 * num_total_columns, sizeof(float) * rows * num_total_columns,
                          cudaMemcpyDefault));
      dim3 grid(static_cast<unsigned>((rows + 255) / 256), num_features);
      GBL_LAUNCH(value_to_bin_kernel, grid, 256, 0, 0, chunk_dev, r0, rows, num_total_columns, ds->meta_dev, off_dev,
                 bounds_dev, num_features, ds->stage, ds->stage_stride);
    }
    GBL_CUDA(cudaDeviceSynchronize());
    build_row_mirror(ds);
    build_column_store(ds);
  } catch (...) {
    cudaFree(off_dev);
    cudaFree(bounds_dev);
    cudaFree(chunk_dev);
    free_dataset(ds);
    throw;
  }
  cudaFree(off_dev);
  cudaFree(bounds_dev);
  cudaFree(chunk_dev);
  *out = reinterpret_cast<gbl_dataset*>(ds);
  GBL_API_END();
}

int gbl_dataset_free(gbl_dataset* ds) {
  GBL_API_BEGIN();
  free_dataset(reinterpret_cast<Dataset*>(ds));
  GBL_API_END();
}

int gbl_dataset_num_data(const gbl_dataset* ds, int64_t* out) {
  GBL_API_BEGIN();
  if (ds == nullptr || out == nullptr) fatal("gbl_dataset_num_data: null argument");
  *out = reinterpret_cast<const Dataset*>(ds)->num_data;
  GBL_API_END();
}

int gbl_dataset_get_feature_bins(const gbl_dataset* ds_, int32_t feature, uint8_t* out_host) {
  GBL_API_BEGIN();
  const Dataset* ds = reinterpret_cast<const Dataset*>(ds_);
  if (ds == nullptr || out_host == nullptr) fatal("gbl_dataset_get_feature_bins: null argument");
  if (feature < 0 || feature >= ds->num_features) fatal("gbl_dataset_get_feature_bins: feature out of range");
  DeviceGuard guard(ds->device);
  const FeatMeta& m = ds->meta_host[feature];
  if (m.bits == 4) {  // unpack on the host
    std::vector<uint8_t> packed(static_cast<size_t>((ds->num_data + 1) / 2));
    GBL_CUDA(cudaMemcpy(packed.data(), ds->col_bins + m.col_off, packed.size(), cudaMemcpyDeviceToHost));
    for (int64_t i = 0; i < ds->num_data; ++i) out_host[i] = static_cast<uint8_t>(read_col_bin(packed.data(), 4, static_cast<int>(i)));
  } else {
    GBL_CUDA(cudaMemcpy(out_host, ds->col_bins + m.col_off, static_cast<size_t>(ds->num_data), cudaMemcpyDeviceToHost));
  }
  GBL_API_END();
}

int gbl_dataset_get_feature_bins_rowmajor(const gbl_dataset* ds_, int32_t feature, uint8_t* out_host) {
  GBL_API_BEGIN();
  const Dataset* ds = reinterpret_cast<const Dataset*>(ds_);
  if (ds == nullptr || out_host == nullptr) fatal("gbl_dataset_get_feature_bins_rowmajor: null argument");
  if (feature < 0 || feature >= ds->num_features) fatal("gbl_dataset_get_feature_bins_rowmajor: feature out of range");
  DeviceGuard guard(ds->device);
  GBL_CUDA(cudaMemcpy2D(out_host, 1, ds->row_bins + feature, ds->row_stride, 1, static_cast<size_t>(ds->num_data),
                        cudaMemcpyDeviceToHost));
  GBL_API_END();
}

}  // extern "C"
