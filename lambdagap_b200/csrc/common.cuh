// common.cuh -- shared device/host declarations of the B200-native tree learner (sm_100a only).
//
// Layout conventions:
//   * bins are *mapper* bins (what BinIterator::Get returns in the reference), one byte per cell;
//   * canonical storage is column-major, one column per feature at FeatMeta::col_off, 4-bit packed when the feature
//     has at most 16 bins (the reference's DenseBin<uint8_t, true>, src/io/dense_bin.hpp:69-82) -- read by the partition;
//   * a row-major mirror  row_bins[N][row_stride]  (row_stride = 32 * num_tiles): one 32-byte sector per
//     (row, 32-feature tile); every learner copies it into its own row records, which also carry the row's
//     gradient pair (learner.cu: record_layout) -- that is what the histogram kernel gathers;
//   * histograms live in HBM as  hist[slot][tile][256 bins][32 lanes]  of exact integers:
//       float mode     : {int64 grad_fix, int64 hess_fix} fixed-point (power-of-two scale per tree)
//       quantized mode : int64 = (int32 grad_sum << 32) | uint32 hess_sum   (the reference's 32-bit packing,
//                        src/treelearner/feature_histogram.hpp:215-216)
#pragma once
#include <cuda_runtime.h>

#include <atomic>
#include <cstdarg>
#include <cstdint>
#include <cstdio>
#include <stdexcept>
#include <string>

#include "lambdagap_b200.h"

namespace gbl {

constexpr int kMaxBin = 256;       // bins per feature held by one histogram row block
constexpr int kTile = 32;          // features per tile = lanes per warp
constexpr int kTailWords = 8;      // int64 words appended to every histogram slot (data-parallel counts)
constexpr int kNumSM = 148;        // B200
constexpr int kPartUnit = 2048;    // rows per partition work unit
constexpr int kHistUnit = 1024;    // rows per histogram work unit
constexpr int kSegUnit = 4096;     // rows per score-update work unit

// include/LightGBM/meta.h:54-56: const double kEpsilon = 1e-15f; kMinScore = -inf
#define GBL_KEPS (static_cast<double>(1e-15f))

struct GblError : public std::runtime_error {
  explicit GblError(const std::string& m) : std::runtime_error(m) {}
};

void set_last_error(const char* msg);
[[noreturn]] void fatal(const char* fmt, ...);
extern std::atomic<long long> g_kernel_launches;
bool debug_sync_enabled();  // GBL_DEBUG_SYNC=1: synchronise after every launch (never inside a graph capture)

#define GBL_CUDA(call)                                                                              \
  do {                                                                                              \
    cudaError_t e__ = (call);                                                                       \
    if (e__ != cudaSuccess) ::gbl::fatal("CUDA error %s at %s:%d: %s", cudaGetErrorName(e__), __FILE__, \
                                          __LINE__, cudaGetErrorString(e__));                      \
  } while (0)

#define GBL_LAUNCH(kernel, grid, block, smem, stream, ...)                   \
  do {                                                                       \
    kernel<<<(grid), (block), (smem), (stream)>>>(__VA_ARGS__);              \
    ::gbl::g_kernel_launches.fetch_add(1, std::memory_order_relaxed);        \
    GBL_CUDA(cudaGetLastError());                                            \
    if (::gbl::debug_sync_enabled()) GBL_CUDA(cudaStreamSynchronize(stream)); \
  } while (0)

#define GBL_API_BEGIN() try {
#define GBL_API_END()                                   \
  }                                                     \
  catch (const std::exception& ex) {                    \
    ::gbl::set_last_error(ex.what());                   \
    return -1;                                          \
  }                                                     \
  catch (...) {                                         \
    ::gbl::set_last_error("unknown exception");        \
    return -1;                                          \
  }                                                     \
  return 0;

inline int64_t round_up(int64_t x, int64_t m) { return (x + m - 1) / m * m; }

// ------------------------------------------------------------------------------------------------ device structs
struct FeatMeta {
  int num_bin;
  int missing_type;
  int default_bin;
  int most_freq_bin;
  int real_index;
  int offset;  // most_freq_bin == 0 (FeatureMetainfo::offset, feature_histogram.hpp:1429-1433)
  int bits;    // 8, or 4 when num_bin <= 16: two rows per byte, even row in the low nibble (dense_bin.hpp:69-82)
  int monotone_type;  // FeatureMetainfo::monotone_type (-1 / 0 / +1)
  int is_categorical; // BinMapper::bin_type() == CategoricalBin
  double penalty;
  long long col_off;  // byte offset of this feature's column in Dataset::col_bins
};

// SplitInfo (src/treelearner/split_info.hpp:22-54) restricted to numerical splits
struct BestSplit {
  double gain;
  int feature;       // inner index, -1 = none
  int real_feature;  // original column, -1 = none
  unsigned threshold;
  int default_left;
  int left_count;
  int right_count;
  double left_output;
  double right_output;
  double left_sum_g;
  double left_sum_h;
  double right_sum_g;
  double right_sum_h;
  long long left_ig;  // exact integer sums (float mode: fixed point; quantized: int grad / hess sums)
  long long left_ih;
  long long right_ig;
  long long right_ih;
  // categorical split: bins that go left (SplitInfo::cat_threshold as a bitset over mapper bins); num_cat == 0: numerical
  unsigned cat_bits[8];
  int num_cat;
  int pad;
};

// One node of the tree under construction.  LeafSplits (src/treelearner/leaf_splits.hpp:22-275) + the
// DataPartition segment (data_partition.hpp:24-47) + best_split_per_leaf_ of the node while it is a leaf.
// Nodes are created in pairs when their parent is expanded; ids grow with creation time (parent id < child id).
struct Node {
  int begin;
  int count;         // local rows
  int global_count;  // rows over all ranks
  int depth;
  int buf;           // which index buffer holds the segment
  int slot;          // histogram slot
  int parent;        // -1 for the root
  int left;          // left child id (right child = left + 1); -1 while the node is a leaf of the grown tree
  int ref_leaf;      // leaf index in the reference's numbering (set by the replay at the end of the tree)
  int in_tree;       // 1 = this node's split belongs to the final tree (set by the replay)
  int pad0;
  int pad1;
  double sum_g;
  double sum_h;
  long long ig;
  long long ih;
  double output;  // LeafSplits::weight(), parent_output of the children
  double cmin, cmax;  // monotone constraints (method basic): the interval this leaf's output must stay in
  double bval;    // bottleneck priority: min of best.gain over the path root..node (<= 0 -> never split)
  BestSplit best;
};

// One node being expanded in the current wave (see learner.cu for the schedule).
struct Expand {
  int node;             // node being split
  int hist_node;        // child whose histogram is built (the smaller one); root wave: the root itself
  int hist_slot;        // slot that histogram goes to
  int skip;             // BeforeFindBestSplit said no: children get no histogram / no scan
  int part_unit_begin;  // first partition work unit of this node (prefix over the wave)
  int hist_unit_begin;  // first histogram work unit
  int left_total;       // rows going left (filled by the plan kernel)
  int pad;
};

struct TreeState {
  int num_nodes;
  int num_expanded;
  int num_slots;
  int num_cand;       // nodes expanded by the current wave
  int part_units;     // partition work units of the current wave
  int hist_units;     // histogram work units of the current wave
  int scan_jobs;      // (node, feature-set) scan jobs of the current wave: 2 per expanded node, 1 for the root
  int done;           // no further wave
  int last_wave;      // the current wave only applies its split (no histogram / scan needed afterwards)
  int waves;
  int num_splits;     // splits of the final tree (set by the replay)
  int num_segments;   // row segments = leaves of the grown tree (>= leaves of the final tree)
  int seg_units;
  int error;          // internal consistency failure (reported by train_finish)
  int rp_nl;          // replay of the reference's loop: current number of leaves ...
  int rp_ns;          // ... and of splits replayed so far (state kept across waves, see split.cu)
  unsigned scan_done_counter;
  unsigned pad;
  int pad2[2];
  // float mode fixed-point scales (powers of two): value_fix = value * g_scale ; value = fix * g_inv
  double g_scale, g_inv, h_scale, h_inv;
  // quantized mode (GradientDiscretizer): grad_scale/hess_scale and inverses
  double q_grad_scale, q_hess_scale, q_inv_grad_scale, q_inv_hess_scale;
  float max_abs_g, max_abs_h;
  double root_output;
  long long hist_rows_total;  // sum of rows over this tree's histogram launches (algorithmic-bytes accounting)
  long long hist_launches;
  long long part_rows_total;  // rows moved by this tree's partitions
};

struct LearnerParams {  // by-value kernel argument
  int num_leaves;
  int max_depth;
  int min_data_in_leaf;
  int quantized;
  int num_grad_quant_bins;
  int is_constant_hessian;
  int num_features;
  int num_tiles;
  int num_ranks;
  int use_mc;  // config.monotone_constraints is not empty: the reference's USE_MC path for every feature
  double monotone_penalty;
  // categorical split search (config.h:480-501)
  int max_cat_threshold, max_cat_to_onehot, min_data_per_group, num_cat_features;
  double cat_l2, cat_smooth;
  double min_sum_hessian_in_leaf;
  double lambda_l1, lambda_l2, min_gain_to_split, max_delta_step, path_smooth;
};

// ------------------------------------------------------------------------------------------------ host objects
struct Dataset {
  int device = 0;
  int64_t num_data = 0;
  int num_features = 0;
  int num_tiles = 0;
  int row_stride = 0;      // bytes between rows of the mirror
  uint8_t* col_bins = nullptr;  // packed column store (FeatMeta::col_off / bits)
  int64_t col_bytes = 0;
  uint8_t* stage = nullptr;     // build-time only: one byte per cell, column-major
  int64_t stage_stride = 0;
  uint8_t* row_bins = nullptr;
  FeatMeta* meta_dev = nullptr;
  FeatMeta* meta_host = nullptr;
};

// bin of `row` in a column of the store
__host__ __device__ inline int read_col_bin(const uint8_t* col, int bits, int row) {
  if (bits == 4) return (col[row >> 1] >> ((row & 1) << 2)) & 0xf;
  return col[row];
}

struct DeviceGuard {
  int prev = -1;
  explicit DeviceGuard(int dev) {
    GBL_CUDA(cudaGetDevice(&prev));
    if (prev != dev) GBL_CUDA(cudaSetDevice(dev));
  }
  ~DeviceGuard() {
    if (prev >= 0) cudaSetDevice(prev);
  }
};

}  // namespace gbl
