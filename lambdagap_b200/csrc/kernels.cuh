// kernels.cuh -- launcher declarations shared by learner.cu and the kernel translation units.
#pragma once
#include "common.cuh"

namespace gbl {

// Peer-memory exchange of the data-parallel learner (exchange.cu); all null on one GPU.
struct PeerX {
  long long* const* pool;   // [num_ranks] every rank's histogram pool (own pool at [rank]), mapped through CUDA IPC
  unsigned* const* flags;   // [num_ranks] every rank's barrier flag array ([num_ranks] words each)
  float* const* maxbuf;     // [num_ranks] every rank's (max |g|, max |h|) deposit array ([num_ranks][2] floats each)
  unsigned* epoch;          // own barrier sequence number
  unsigned* dead;           // own sticky "a barrier timed out" flag
  unsigned* done;           // own count of exchange-kernel CTAs that have finished their share
  unsigned* go;             // own "every rank has arrived" word: CTA 0 of the exchange kernel releases the others with it
  int rank;
  int pad;
  long long timeout_cycles;  // bound of every barrier wait (SM clocks)
};

// Everything a per-tree kernel needs to find its inputs; passed by value (fits the 4 KB parameter space).
struct LearnerDev {
  PeerX px;
  // dataset
  const uint8_t* col_bins;
  uint8_t* rec;            // learner-owned row records: [row][32 bins per tile ...][gradient pair], see learner.cu
  int rec_stride;          // bytes between row records
  int tile_stride;         // bytes between the 32-feature tiles of a record (32, or 64 when each tile carries a pair)
  int gh_off;              // byte offset of the gradient pair inside a tile block
  int num_data;  // local rows
  const FeatMeta* meta;
  const int* cat_features;  // [p.num_cat_features] inner indices of the categorical features (nullptr when there are none)
  // per-tree state
  TreeState* ts;
  Node* nodes;             // [max_nodes]
  unsigned long long* bkey;   // [max_nodes] bit pattern of Node::bval where positive, else 0 (compact copy for the selection)
  unsigned char* expanded;    // [max_nodes] 1 once the node has been split
  Expand* wave;            // [max_cand + 1] nodes expanded by the current wave
  int* idx[3];             // [0], [1]: ping-pong index buffers (DataPartition::indices_); [2]: 0..N-1, never written
                           // after the learner is created -- the root's segment when there is no bagging
  // gradients
  const float* grad;       // inputs of this tree
  const float* hess;
  short* gh_q;             // quantized mode: compact copy of the int8 [hess, grad] pairs (gradient_discretizer.cpp:124-127)
  // histograms
  long long* hist_pool;    // [max_slots][slot_stride int64]
  long long slot_stride;   // int64 words per slot (including kTailWords)
  unsigned char* splittable;  // [max_slots][num_features]  FeatureHistogram::is_splittable_
  const signed char* feature_mask;  // col sampler by-tree mask or nullptr
  BestSplit* cand;         // [2 * max_cand][num_features] per-feature candidates of the current wave
  unsigned* job_done;      // [2 * max_cand] features finished per scan job
  gbl_split_record* records;  // [num_leaves - 1]
  // partition scratch
  int* unit_left;          // [max_part_units + 1] rows going left per work unit, then their exclusive prefix
  unsigned* unit_bits;     // [max_part_units][kPartUnit / 32] go-left flags of every row of the unit
  // result of the replay
  double* leaf_values;     // [num_leaves] outputs of the last tree in the reference's leaf numbering
  int* leaf_node;          // [num_leaves] node of every final leaf
  int* split_node;         // [num_leaves] node split by the k-th split of the final tree
  unsigned char* replay_scratch;  // compact node arrays of the replay when they do not fit shared memory, else null
  int* seg_node;           // [max_slots] nodes that hold rows (leaves of the grown tree)
  int* seg_unit_begin;     // [max_slots + 1] prefix of score-update work units
  // stochastic rounding
  const double* rand_g;
  const double* rand_h;
  int rand_start;
  // capacities
  int max_nodes;
  int max_slots;
  int max_cand;            // nodes per wave
  int max_expand;          // expansions per tree
  int spec_limit;          // expansions after which only certain nodes are expanded (one per wave)
  int max_part_units;
  int spec_min;            // damped waves still take at least this many ranks beyond the expanded nodes
  float spec_rho;          // fraction of the remaining split budget whose ranks a damped wave considers (1 = exact rule)
  LearnerParams p;
};

// GradientDiscretizer::DiscretizeGradients for one row (gradient_discretizer.cpp:116-160).  Casts truncate toward
// zero.  Low byte = hess, high byte = grad (int8 pair read as one int16, dense_bin.hpp:181,196).
__device__ __forceinline__ short discretize_row(const LearnerDev& d, int i, double inv_g, double inv_h) {
  const double gradient = static_cast<double>(d.grad[i]);
  double rg = 0.5, rh = 0.5;
  if (d.rand_g != nullptr) {
    const int pos = static_cast<int>((static_cast<long long>(i) + d.rand_start) % d.num_data);
    rg = d.rand_g[pos];
    rh = d.rand_h[pos];
  }
  const signed char qg = gradient >= 0.0 ? static_cast<signed char>(static_cast<int>(gradient * inv_g + rg))
                                          : static_cast<signed char>(static_cast<int>(gradient * inv_g - rg));
  signed char qh;
  if (d.p.is_constant_hessian) {
    qh = 1;
  } else {
    qh = static_cast<signed char>(static_cast<int>(static_cast<double>(d.hess[i]) * inv_h + rh));
  }
  return static_cast<short>((static_cast<int>(qg) << 8) | (static_cast<int>(qh) & 0xff));
}

// index buffer that receives the children of a node whose segment lives in buffer `buf`
__host__ __device__ inline int child_buf(int buf) { return buf == 2 ? 0 : (buf ^ 1); }

// prep.cu
void launch_iota(int* out, int n, cudaStream_t s);
void launch_build_records(uint8_t* rec, int rec_stride, int tile_stride, const uint8_t* row_bins, int row_stride, int n,
                          cudaStream_t s);
// bag_count_dev != nullptr: the bag size is read on the device (device-drawn bags change size without a new graph)
void launch_tree_begin(const LearnerDev& d, const int* bag_indices, int bag_count, const int* bag_count_dev, cudaStream_t s);
// exchange.cu
void launch_xchg_wave(const LearnerDev& d, int is_root, cudaStream_t s);  // barrier + in-place sum over ranks + barrier
bool xchg_fused_enabled();
int xchg_launches_per_wave();  // 3 (default), 2 or 1: GBL_XCHG_FUSED  // GBL_XCHG_FUSED=1: the exchange as one kernel instead of three launches
void launch_xchg_max(const LearnerDev& d, cudaStream_t s);
// bagging.cu
void launch_bag_seed(unsigned* state, int nblk, int seed, cudaStream_t s);
void launch_bag_draw_by_query(unsigned* state, int nblk_q, int nq, double fraction, unsigned* qbits, int* qblk_cnt,
                              const int* row_query, int nblk, int n, unsigned* bits, int* blk_cnt, int* blk_off, int* counts,
                              int* bag, int* oob, cudaStream_t s);
void launch_bag_draw(unsigned* state, int nblk, int n, double fraction, const float* label, double pos_fraction,
                     double neg_fraction, unsigned* bits, int* blk_cnt, int* blk_off, int* counts, int* bag, int* oob,
                     cudaStream_t s);
void launch_maxabs(const LearnerDev& d, cudaStream_t s);
void launch_scales_and_convert(const LearnerDev& d, cudaStream_t s);

// histogram.cu
void histogram_setup();
// wave mode (explicit_count < 0): every node of the current wave; tap mode: the given row list into explicit_out
void launch_histogram(const LearnerDev& d, const int* explicit_indices, int explicit_count, long long* explicit_out,
                      cudaStream_t s);

// split.cu
void launch_root_init(const LearnerDev& d, cudaStream_t s);
// scans the children of the current wave (or the root), reduces, and selects the next wave; `cond` is the
// conditional handle of the while-node when running inside the CUDA graph (use_cond != 0)
void launch_find_best(const LearnerDev& d, int is_root, cudaGraphConditionalHandle cond, int use_cond, cudaStream_t s);
constexpr size_t kFinalizeSmemMax = 200 * 1024;
size_t finalize_scratch_bytes(const LearnerDev& d);
void launch_finalize(const LearnerDev& d, unsigned char* global_scratch, cudaStream_t s);

void launch_renew_sums(const LearnerDev& d, double* unit_sums, double* leaf_sums, cudaStream_t s);
void launch_renew_outputs(const LearnerDev& d, const double* leaf_sums, cudaStream_t s);

// partition.cu
void launch_partition(const LearnerDev& d, cudaStream_t s);
void launch_add_score(const LearnerDev& d, const double* leaf_values_dev, double scale, double bias, double* score,
                      unsigned short* leaf_scratch, bool partial_rows, cudaStream_t s);
void launch_leaf_assignment(const LearnerDev& d, int* out, cudaStream_t s);
void launch_leaf_mark(const LearnerDev& d, unsigned short* out, bool partial_rows, cudaStream_t s);  // 0xFFFF = not in the tree
void launch_add_score_rows(const LearnerDev& d, const uint8_t* row_bins, int row_stride, const int* rows, int num_rows,
                           const double* leaf_values_dev, double scale, double bias, double* score, cudaStream_t s);

}  // namespace gbl
