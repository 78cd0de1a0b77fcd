// kernels.cuh -- launcher declarations shared by learner.cu and the kernel translation units.
#pragma once
#include "common.cuh"

namespace gbl {

// Everything a per-tree kernel needs to find its inputs; passed by value (fits the 4 KB parameter space).
struct LearnerDev {
  // dataset
  const uint8_t* col_bins;
  const uint8_t* row_bins;
  int64_t col_stride;
  int row_stride;
  int num_data;  // local rows
  const FeatMeta* meta;
  // per-tree state
  TreeState* ts;
  LeafState* leaves;
  int* idx[2];             // ping-pong index buffers (DataPartition::indices_)
  // gradients
  const float* grad;       // inputs of this tree
  const float* hess;
  int2* gh_fix;            // float mode: fixed-point pair per row
  short* gh_q;             // quantized mode: int8 [hess, grad] pair per row (gradient_discretizer.cpp:124-127)
  // histograms
  long long* hist_pool;    // [num_leaves slots][slot_stride int64]
  long long slot_stride;   // int64 words per slot (including kTailWords)
  unsigned char* splittable;  // [num_leaves slots][num_features]  FeatureHistogram::is_splittable_
  const signed char* feature_mask;  // col sampler by-tree mask or nullptr
  BestSplit* cand;         // [2][num_features] per-feature candidates of the current round
  gbl_split_record* records;  // [num_leaves - 1]
  int* block_counts;       // partition scratch
  double* leaf_values;     // [num_leaves] outputs of the last tree (device-resident score update)
  // stochastic rounding
  const double* rand_g;
  const double* rand_h;
  int rand_start;
  LearnerParams p;
};

// prep.cu
void launch_iota(int* out, int n, cudaStream_t s);
void launch_tree_begin(const LearnerDev& d, const int* bag_indices, int bag_count, cudaStream_t s);
void launch_maxabs(const LearnerDev& d, cudaStream_t s);
void launch_scales_and_convert(const LearnerDev& d, cudaStream_t s);

// histogram.cu
void histogram_setup();
void launch_histogram(const LearnerDev& d, int slot, const int* explicit_indices, int explicit_count,
                      long long* explicit_out, cudaStream_t s);

// split.cu
void launch_root_init(const LearnerDev& d, cudaStream_t s);
void launch_find_best(const LearnerDev& d, int round, cudaStream_t s);

// partition.cu
int partition_grid();
void launch_partition(const LearnerDev& d, int round, cudaStream_t s);
void launch_add_score(const LearnerDev& d, const double* leaf_values_dev, double scale, double bias, double* score,
                      cudaStream_t s);
void launch_leaf_assignment(const LearnerDev& d, int* out, cudaStream_t s);

}  // namespace gbl
