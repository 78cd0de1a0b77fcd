/*
 * lambdagap_b200.h -- C ABI of the B200-native (sm_100a) tree-learner hot path for the
 * adaliaramon/LambdaGap LightGBM fork.
 *
 * This is the drop-in boundary: plain C, opaque handles, plain pointers and sizes, `int` status
 * (0 = ok, -1 = error, message via gbl_last_error(), mirroring API_BEGIN/API_END + LGBM_GetLastError of the
 * reference, src/c_api.cpp:38-52).  Nothing here allocates or frees memory the caller owns.  All citations are
 * relative to the reference tree (/root/reference).  A maintainer binds these entry points from a `TreeLearner`
 * / `ObjectiveFunction` subclass registered under device_type=cuda.
 *
 * Pointer conventions: `*_dev` = CUDA device pointer on the handle's device, `*_host` = host pointer.
 * Threading: one handle is used by one host thread at a time (the reference serialises Booster calls,
 * src/c_api.cpp:410-423); different handles may be used concurrently from different threads.
 * There is no CPU fallback: every entry point fails with -1 when no sm_100 device is usable.
 */
#ifndef LAMBDAGAP_B200_H_
#define LAMBDAGAP_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GBL_API __attribute__((visibility("default")))

/* MissingType, include/LightGBM/bin.h:28-32 */
#define GBL_MISSING_NONE 0
#define GBL_MISSING_ZERO 1
#define GBL_MISSING_NAN 2

/* objective kinds, src/objective/objective_function.cpp:19-103 */
#define GBL_OBJ_REGRESSION_L2 0
#define GBL_OBJ_BINARY 1
#define GBL_OBJ_LAMBDARANK 2
#define GBL_OBJ_MULTICLASS 3 /* MulticlassSoftmax, src/objective/multiclass_objective.hpp:25-170 */

/* metric kinds, src/metric/metric.cpp:20-133 (Metric::CreateMetric) */
#define GBL_METRIC_L2 0
#define GBL_METRIC_RMSE 1
#define GBL_METRIC_L1 2
#define GBL_METRIC_BINARY_LOGLOSS 3
#define GBL_METRIC_BINARY_ERROR 4
#define GBL_METRIC_AUC 5
#define GBL_METRIC_NDCG 6

typedef struct gbl_dataset gbl_dataset;
typedef struct gbl_learner gbl_learner;
typedef struct gbl_objective gbl_objective;
typedef struct gbl_metric gbl_metric;
typedef struct gbl_comm gbl_comm;

/* Per used ("inner") feature: what FeatureMetainfo (src/treelearner/feature_histogram.hpp:25-38) and
 * BinMapper (include/LightGBM/bin.h:86-260) carry for a numerical feature. */
typedef struct gbl_feature_meta {
  int32_t num_bin;       /* BinMapper::num_bin(), <= 256 */
  int32_t missing_type;  /* GBL_MISSING_* */
  int32_t default_bin;   /* BinMapper::GetDefaultBin(): bin of value 0.0 */
  int32_t most_freq_bin; /* BinMapper::GetMostFreqBin() */
  int32_t real_index;    /* Dataset::RealFeatureIndex(inner): original column, used for split tie-breaks */
  int32_t monotone_type; /* FeatureMetainfo::monotone_type: -1 / 0 / +1 (config.monotone_constraints), 0 by default */
  double penalty;        /* FeatureMetainfo::penalty (feature_contri), 1.0 by default */
  int32_t bin_type;      /* BinMapper::bin_type(): 0 = numerical, 1 = categorical (include/LightGBM/bin.h:22-25) */
  int32_t reserved;
} gbl_feature_meta;

/* The Config fields the hot path reads (include/LightGBM/config.h). */
typedef struct gbl_learner_config {
  int32_t num_leaves;              /* config.h:285 */
  int32_t max_depth;               /* <=0: no limit */
  int32_t min_data_in_leaf;        /* 20 */
  int32_t use_quantized_grad;      /* config.h:625 */
  int32_t num_grad_quant_bins;     /* 4 */
  int32_t stochastic_rounding;     /* 1; random arrays are supplied by gbl_learner_set_quant_random */
  int32_t quant_train_renew_leaf;  /* 0 */
  int32_t is_constant_hessian;     /* ObjectiveFunction::IsConstantHessian() */
  double min_sum_hessian_in_leaf;  /* 1e-3 */
  double lambda_l1;
  double lambda_l2;
  double min_gain_to_split;
  double max_delta_step;
  double path_smooth;
} gbl_learner_config;

/* One performed split = the arguments of Tree::Split (src/io/tree.cpp:60-74) as SerialTreeLearner::SplitInner
 * passes them (src/treelearner/serial_tree_learner.cpp:798-810).  Record k splits `leaf`; the left child keeps
 * id `leaf`, the right child gets id k+1 (include/LightGBM/tree.h:543-583).  The records come in the order of the
 * reference's leaf-wise loop (serial_tree_learner.cpp:210-245: always the leaf with the largest gain next), whatever
 * order the device expanded the nodes in. */
typedef struct gbl_split_record {
  int32_t leaf;
  int32_t feature;       /* inner feature index */
  int32_t real_feature;  /* original column */
  uint32_t threshold_bin;
  int32_t default_left;
  int32_t left_count;    /* true row counts after the partition (global over ranks) */
  int32_t right_count;
  int32_t reserved;
  double left_output;
  double right_output;
  double left_sum_gradient;
  double left_sum_hessian;
  double right_sum_gradient;
  double right_sum_hessian;
  double gain;           /* SplitInfo::gain (min_gain_to_split not yet re-added) */
  int64_t left_sum_gradient_and_hessian;  /* quantized mode: packed int32 grad << 32 | uint32 hess */
  int64_t right_sum_gradient_and_hessian;
  /* categorical split (SplitInfo::cat_threshold, split_info.hpp:45-47) as a bitset over the feature's bins: bin b goes
   * left iff bit b is set; num_cat_bins = number of set bits, 0 for a numerical split */
  uint32_t cat_bitset[8];
  int32_t num_cat_bins;
  int32_t reserved2;
} gbl_split_record;

/* Result header of one trained tree. */
typedef struct gbl_tree_result {
  int32_t num_splits;     /* number of valid records (num_leaves of the tree = num_splits + 1) */
  int32_t reserved;
  double root_output;     /* Tree::SetLeafOutput(0, ...) value, serial_tree_learner.cpp:205-208 */
  double root_sum_gradient;
  double root_sum_hessian;
  double grad_scale;      /* quantized mode: GradientDiscretizer::grad_scale()/hess_scale() */
  double hess_scale;
} gbl_tree_result;

/* ---------------------------------------------------------------- errors / device */
/* replaces LGBM_GetLastError (src/c_api.cpp:54) for this library; thread-local string */
GBL_API const char* gbl_last_error(void);
/* number of usable sm_100 devices (0 => every other call fails loudly) */
GBL_API int gbl_device_count(int* out_count);
/* Total kernels launched by this library in this process (bench.py's gpu_launches). */
GBL_API int64_t gbl_kernel_launch_count(void);

/* ---------------------------------------------------------------- src/io: binned storage */
/* Device-resident binned dataset from host mapper bins (one uint8 column of `num_data` rows per used
 * feature, the value BinIterator::Get returns, src/io/dense_bin.hpp:24-50).  Replaces
 * Dataset::CreateCUDAColumnData (src/io/dataset.cpp:1767-1899) hooked from Dataset::FinishLoad (:443-463). */
GBL_API int gbl_dataset_create_from_bins(const gbl_feature_meta* metas, int32_t num_features, int64_t num_data,
                                         const uint8_t* const* columns_host, int32_t device_id, gbl_dataset** out);
/* Same, but bins are computed on the device from raw values: BinMapper::ValueToBin (include/LightGBM/bin.h:612-649)
 * for every cell of a row-major float32 host matrix [num_data x num_total_columns]; `upper_bounds_host` holds
 * the concatenated bin_upper_bound_ arrays (num_bin doubles per used feature), found on the host by
 * BinMapper::FindBin (src/io/bin.cpp:313-508).  Replaces the PushOneRow loop of
 * LGBM_DatasetCreateFromMats (src/c_api.cpp:1322-1405). */
GBL_API int gbl_dataset_create_from_mat_f32(const gbl_feature_meta* metas, int32_t num_features, int64_t num_data,
                                            const float* values_host, int32_t num_total_columns,
                                            const double* upper_bounds_host, int32_t device_id, gbl_dataset** out);
GBL_API int gbl_dataset_free(gbl_dataset* ds);
GBL_API int gbl_dataset_num_data(const gbl_dataset* ds, int64_t* out);
/* test tap: copies the mapper bins of one feature back to the host (column-major store) */
GBL_API int gbl_dataset_get_feature_bins(const gbl_dataset* ds, int32_t feature, uint8_t* out_host);
/* test tap: same cell values read from the row-major gather mirror */
GBL_API int gbl_dataset_get_feature_bins_rowmajor(const gbl_dataset* ds, int32_t feature, uint8_t* out_host);

/* ---------------------------------------------------------------- src/network: NCCL communicator */
/* 128-byte NCCL unique id created on rank 0 and sent to the other ranks by the host (torch.distributed / MPI /
 * sockets).  Replaces Network::Init + Linkers (src/network/network.cpp:30-58, linkers_socket.cpp). */
GBL_API int gbl_comm_get_unique_id(uint8_t out_id[128]);
GBL_API int gbl_comm_create(const uint8_t id[128], int32_t num_ranks, int32_t rank, int32_t device_id, gbl_comm** out);
GBL_API int gbl_comm_free(gbl_comm* comm);
/* Network::GlobalSyncUpBySum / ByMax / ByMean on small host vectors (include/LightGBM/network.h:180-260), used by
 * the host layer for init scores and label statistics.  op: 0 = sum, 1 = max. */
GBL_API int gbl_comm_allreduce_host_f64(gbl_comm* comm, double* inout_host, int32_t n, int32_t op);
GBL_API int gbl_comm_info(const gbl_comm* comm, int32_t* out_num_ranks, int32_t* out_rank);

/* ---------------------------------------------------------------- src/treelearner: the tree learner */
/* TreeLearner::Init (include/LightGBM/tree_learner.h:37).  `comm` NULL = single GPU (SerialTreeLearner
 * semantics); non-NULL = data-parallel learner over row shards (DataParallelTreeLearner,
 * src/treelearner/data_parallel_tree_learner.cpp:123-463) with `ds` holding this rank's rows. */
GBL_API int gbl_learner_create(const gbl_dataset* ds, const gbl_learner_config* config, gbl_comm* comm,
                               gbl_learner** out);
GBL_API int gbl_learner_free(gbl_learner* learner);
/* TreeLearner::ResetConfig (tree_learner.h:49) */
GBL_API int gbl_learner_reset_config(gbl_learner* learner, const gbl_learner_config* config);
/* ColSampler::is_feature_used_bytree (src/treelearner/col_sampler.hpp); NULL = all features */
/* Monotone constraints, method "basic" (src/treelearner/monotone_constraints.hpp:465-514: every leaf carries one
 * [min, max] output interval, inherited from its parent and tightened at a split on a constrained feature to the
 * mid-point of the two child outputs).  `enabled` mirrors `!config.monotone_constraints.empty()`: the reference then
 * takes its USE_MC code path for EVERY feature (outputs clamped to the leaf's interval, gain from the clamped outputs,
 * feature_histogram.hpp:740-797), whatever the feature's own type; the per-feature types come with the dataset
 * (gbl_feature_meta.monotone_type).  `monotone_penalty`: config.h:532, applied to the gain of splits on constrained
 * features by depth (monotone_constraints.hpp:357-366).  Call before training; drops the captured tree. */
GBL_API int gbl_learner_set_monotone(gbl_learner* learner, int32_t enabled, double monotone_penalty);
/* Categorical split search parameters (config.h:480-501; FindBestThresholdCategoricalInner,
 * src/treelearner/feature_histogram.cpp:142-380).  Defaults of the reference: 32, 10.0, 10.0, 4, 100. */
GBL_API int gbl_learner_set_categorical(gbl_learner* learner, int32_t max_cat_threshold, double cat_l2, double cat_smooth,
                                        int32_t max_cat_to_onehot, int32_t min_data_per_group);
GBL_API int gbl_learner_set_feature_mask(gbl_learner* learner, const int8_t* is_feature_used_host);
/* TreeLearner::SetBaggingData (tree_learner.h:75): sorted row subset used for the next trees; NULL/0 = all rows */
GBL_API int gbl_learner_set_bagging_data(gbl_learner* learner, const int32_t* used_indices_host, int32_t num_used);
/* BaggingSampleStrategy::Bagging (src/boosting/bagging.hpp:31-139) drawn ON THE DEVICE: every block of 1024 rows owns
 * a generator seeded bagging_seed + block whose state persists between calls (reset != 0 or a new seed re-seeds),
 * each row of the block draws NextFloat() and joins the bag below bagging_fraction -- the reference's integers, so the
 * bag equals the reference's whatever its thread count.  Replaces the per-iteration host draw + index upload of
 * gbl_learner_set_bagging_data; the bag size is returned (it is read back, 8 bytes).  A bag that happens to hold every
 * row trains on every row.  Same restrictions as gbl_learner_set_bagging_data. */
GBL_API int gbl_learner_bagging_dev(gbl_learner* learner, int32_t bagging_seed, double bagging_fraction, int32_t reset,
                                    int32_t* out_bag_count);
/* BalancedBaggingHelper (src/boosting/bagging.hpp:245-270): the same draw, with pos_bagging_fraction for rows whose label
 * is positive and neg_bagging_fraction for the others.  label_host[num_data] is copied to the device when given (first
 * call, or when the labels changed); NULL reuses the previous copy. */
/* bagging_by_query (src/boosting/bagging.hpp:54-99): the draw runs over queries (generators per block of 1024 QUERIES) and
 * a row is in the bag when its query is.  query_boundaries_host[num_queries + 1] is copied when given; NULL reuses it. */
GBL_API int gbl_learner_bagging_by_query_dev(gbl_learner* learner, int32_t bagging_seed, double bagging_fraction,
                                             const int32_t* query_boundaries_host, int32_t num_queries, int32_t reset,
                                             int32_t* out_bag_count);
GBL_API int gbl_learner_bagging_balanced_dev(gbl_learner* learner, int32_t bagging_seed, double pos_bagging_fraction,
                                             double neg_bagging_fraction, const float* label_host, int32_t reset,
                                             int32_t* out_bag_count);
/* GradientDiscretizer's host-generated uniform [0,1) arrays for stochastic rounding
 * (src/treelearner/gradient_discretizer.cpp:16-41) and the per-iteration start offset (:116).  */
GBL_API int gbl_learner_set_quant_random(gbl_learner* learner, const double* grad_random_host,
                                         const double* hess_random_host, int64_t n);
GBL_API int gbl_learner_set_quant_random_start(gbl_learner* learner, int32_t random_values_use_start);
/* TreeLearner::Train (tree_learner.h:61): grows one tree from device gradients/hessians (score_t = float).
 * `records_host` must hold num_leaves-1 records.  Synchronises the learner's stream before returning.
 * Environment (read when the learner is created): GBL_USE_GRAPH=0 drives the tree from the host instead of replaying
 * one CUDA graph; GBL_SPEC_LIMIT=n caps speculative expansions (0 = the reference's own order, one split per round);
 * GBL_HIST_POOL_MB bounds the histogram pool.  None of them changes a result. */
GBL_API int gbl_learner_train(gbl_learner* learner, const float* gradients_dev, const float* hessians_dev,
                              gbl_tree_result* result_host, gbl_split_record* records_host);
/* Same with host gradients (the path taken when boosting is not on the device, cf.
 * src/treelearner/cuda/cuda_single_gpu_tree_learner.cpp:99-105). */
GBL_API int gbl_learner_train_host(gbl_learner* learner, const float* gradients_host, const float* hessians_host,
                                   gbl_tree_result* result_host, gbl_split_record* records_host);
/* Asynchronous form for a device-resident boosting loop: enqueues the tree on the learner's stream and
 * returns; gbl_learner_train_finish waits and copies the records.  Leaf outputs of the enqueued tree stay on
 * the device so that gbl_learner_add_prediction_to_score_dev can follow without a host round trip. */
GBL_API int gbl_learner_train_async(gbl_learner* learner, const float* gradients_dev, const float* hessians_dev);
GBL_API int gbl_learner_train_finish(gbl_learner* learner, gbl_tree_result* result_host,
                                     gbl_split_record* records_host);
/* TreeLearner::AddPredictionToScore (src/treelearner/serial_tree_learner.h:100-115): score[row] += leaf value
 * for every row of the last trained tree's partition; `leaf_values_host[num_leaves]` are the (shrunk) outputs. */
GBL_API int gbl_learner_add_prediction_to_score(gbl_learner* learner, const double* leaf_values_host,
                                                int32_t num_leaves, double* score_dev);
/* Device-resident variant: score[row] += shrinkage * leaf_output(last tree) + bias, no host data needed. */
GBL_API int gbl_learner_add_prediction_to_score_dev(gbl_learner* learner, double shrinkage, double bias,
                                                    double* score_dev);
/* test taps */
/* leaf id of every (local) row after the last tree: DataPartition (src/treelearner/data_partition.hpp) */
/* With bagging data set: the score update of the rows OUTSIDE the bag -- each walks the last tree on its bins
 * (GBDT::UpdateScore -> ScoreUpdater::AddScore(tree, out-of-bag indices), src/boosting/gbdt.cpp:505-513,
 * Tree::AddPredictionToScore src/io/tree.cpp:193-230).  Asynchronous on the learner's stream; no-op without bagging. */
GBL_API int gbl_learner_add_prediction_out_of_bag_dev(gbl_learner* learner, double shrinkage, double bias,
                                                      double* score_dev);
/* Scores of ANOTHER dataset binned with the training data's mappers (validation data): every row walks the last tree
 * (ScoreUpdater::AddScore(tree) on a validation set, src/boosting/score_updater.hpp:79-86 -> Tree::AddPredictionToScore,
 * src/io/tree.cpp:153-191).  Asynchronous on the learner's stream. */
GBL_API int gbl_learner_add_prediction_to_dataset_dev(gbl_learner* learner, const gbl_dataset* other, double shrinkage,
                                                      double bias, double* score_dev);
/* leaf outputs of the last tree in the reference's leaf numbering, as the device holds them (before shrinkage); with
 * quant_train_renew_leaf these are the renewed outputs of GradientDiscretizer::RenewIntGradTreeOutput
 * (src/treelearner/gradient_discretizer.cpp:209-260), which the split records do not carry */
GBL_API int gbl_learner_get_leaf_values(gbl_learner* learner, double* out_host, int32_t num_leaves);
GBL_API int gbl_learner_get_leaf_assignment(gbl_learner* learner, int32_t* out_host);
/* The same as 2-byte leaf ids (trees of at most 65535 leaves; 0xFFFF = row outside the tree's rows), written by the
 * kernel the device-resident score update uses -- a quarter of the bytes for a host that keeps the scores itself
 * (SerialTreeLearner::AddPredictionToScore on a host ScoreUpdater, serial_tree_learner.h:100-115). */
GBL_API int gbl_learner_get_leaf_assignment_u16(gbl_learner* learner, uint16_t* out_host);
/* index segment of one leaf, in partition order (DataPartition::GetIndexOnLeaf) */
GBL_API int gbl_learner_get_leaf_indices(gbl_learner* learner, int32_t leaf, int32_t* out_host, int32_t* out_count);
/* histogram of `leaf`'s slot for `feature` as (gradient, hessian) pairs per mapper bin [num_bin][2]:
 * float mode -> doubles; quantized mode -> integer sums stored as doubles (exact) */
GBL_API int gbl_learner_get_histogram(gbl_learner* learner, int32_t leaf, int32_t feature, double* out_host);
/* quantized mode: the int8 [hess, grad] pairs of the last tree (gradient_discretizer.cpp:124-127) */
GBL_API int gbl_learner_get_discretized(gbl_learner* learner, int8_t* out_host);
/* Standalone histogram construction of an arbitrary sorted row subset (test + bench tap of the dominant
 * kernel): Dataset::ConstructHistograms (include/LightGBM/dataset.h:735-766).  indices_dev NULL = all rows.
 * Uses the gradients of the last gbl_learner_train* call.  `out_hist_dev` is a zeroed device buffer of
 * gbl_learner_histogram_words() int64 words in the library's HBM layout [tile][256 bins][32 features] of
 * exact integer sums: float mode {int64 grad_fix, int64 hess_fix} pairs, quantized mode the reference's packed
 * (int32 grad << 32 | uint32 hess).  Asynchronous on the learner's stream. */
GBL_API int gbl_learner_histogram_words(gbl_learner* learner, int64_t* out_words);
GBL_API int gbl_learner_construct_histogram(gbl_learner* learner, const int32_t* indices_dev, int64_t num_indices,
                                            int64_t* out_hist_dev);
GBL_API void* gbl_learner_stream(gbl_learner* learner);

/* Per-stage device time of the trees trained since profiling was switched on, measured with CUDA events on the
 * learner's stream around every kernel group (profiling disables CUDA-graph replay and adds two event records
 * per stage and round).  Feeds bench.py's `roofline` object; the reference's analogue is the USE_TIMETAG
 * stage table (include/LightGBM/utils/common.h:984-1068). */
typedef struct gbl_profile {
  double hist_ms;        /* ConstructHistograms */
  double find_ms;        /* FindBestSplitsFromHistograms (+ arg-max) */
  double partition_ms;   /* DataPartition::Split + bookkeeping */
  double other_ms;       /* BeforeTrain, gradient preparation, root sums */
  double collective_ms;  /* NCCL all-reduces (data-parallel learner) */
  int64_t hist_launches; /* histogram kernel launches that had rows to process */
  int64_t hist_rows;     /* sum of leaf rows over those launches */
  int64_t trees;
  int64_t partition_rows; /* rows moved by DataPartition::Split */
  int64_t waves;          /* batched rounds (all leaves that may still be needed are split together) */
  int64_t expansions;     /* nodes split on the device, including speculative ones the final trees do not use */
  int64_t splits;         /* splits of the final trees */
} gbl_profile;
GBL_API int gbl_learner_set_profiling(gbl_learner* learner, int32_t enabled);
/* CUDA-event stopwatch on the learner's stream (the stream every kernel of this library is launched on):
 * record slot 0..7, then elapsed milliseconds between two recorded slots (synchronises on the later one). */
GBL_API int gbl_learner_timer_record(gbl_learner* learner, int32_t slot);
GBL_API int gbl_learner_timer_elapsed_ms(gbl_learner* learner, int32_t slot_begin, int32_t slot_end, double* out_ms);
GBL_API int gbl_learner_get_profile(gbl_learner* learner, gbl_profile* out, int32_t reset);

/* No counterpart in the reference.  The same stable row partition the learner performs for a split
 * (left block first, both blocks in their original order; missing rows follow default_left), for callers that may
 * not allocate device memory: hosts whose own pool allocator owns all of HBM, and latency-critical callers that
 * cannot take cudaMalloc's device-wide synchronisation.  The per-CTA scratch lives in a static device array inside
 * the library (16.25 KB per GPU, held until process exit), so the call performs no allocation of any kind; the
 * caller supplies every buffer: `indices_dev` (NULL = rows 0..num_indices-1), `out_indices_dev[num_indices]`,
 * `out_left_count_dev[1]`, all on the dataset's device.  `indices_dev` and `out_indices_dev` must not overlap (the
 * partition is not in place; an overlapping call is refused with -1) and every index must be a row of the dataset --
 * the caller guarantees that, the indices are not range-checked.  Limits and rules: at most 16,777,216 rows per call (more is
 * refused with -1); all calls on one GPU share the scratch, so calls that overlap on the GPU race -- issue them
 * one after another; the first call made on a device also resolves the scratch address and should come before any
 * stream capture.  Asynchronous on `stream` (NULL = default stream). */
GBL_API int gbl_partition_rows_noalloc(const gbl_dataset* ds, int32_t feature, uint32_t threshold_bin,
                                       int32_t default_left, const int32_t* indices_dev, int32_t num_indices,
                                       int32_t* out_indices_dev, int32_t* out_left_count_dev, void* stream);

/* ---------------------------------------------------------------- src/objective */
typedef struct gbl_objective_config {
  int32_t kind;                        /* GBL_OBJ_* */
  int32_t lambdarank_target;           /* index into the reference's target list, rank_objective.hpp:22-41 */
  int32_t lambdarank_truncation_level; /* 30 */
  int32_t lambdarank_norm;             /* 1 */
  double sigmoid;                      /* 1.0 */
  double lambdagap_weight;             /* 1.0 */
  int32_t is_unbalance;                /* binary */
  int32_t num_class;                   /* multiclass (GBL_OBJ_MULTICLASS): >= 2; scores / gradients / hessians are
                                          class-major arrays of num_class * num_data entries like the reference's
                                          (multiclass_objective.hpp:84-86); 0 for every other kind */
  double scale_pos_weight;             /* binary, 1.0 */
  int32_t num_label_gain;
  int32_t reserved2;
  const double* label_gain;            /* [num_label_gain], default 2^i - 1 (dcg_calculator.cpp:33-43) */
} gbl_objective_config;

/* ObjectiveFunction::Init(metadata, num_data) (include/LightGBM/objective_function.h:29).  label/weight/
 * query_boundaries are host arrays copied to the device; weight and query_boundaries may be NULL. */
GBL_API int gbl_objective_create(const gbl_objective_config* config, int64_t num_data, const float* label_host,
                                 const float* weight_host, const int32_t* query_boundaries_host,
                                 int32_t num_queries, int32_t device_id, gbl_objective** out);
/* BinaryLogloss::Init sums the numbers of positive / negative rows over machines before it derives need_train and the
 * is_unbalance label weights (src/objective/binary_objective.hpp:73-101, Network::GlobalSyncUpBySum at :75-77).  The
 * objective is created from one rank's rows; the data-parallel host layer reduces the two counts over ranks and hands the
 * global numbers in here.  No effect on other objectives. */
GBL_API int gbl_objective_set_label_counts(gbl_objective* obj, int64_t num_positive, int64_t num_negative);
GBL_API int gbl_objective_free(gbl_objective* obj);
/* ObjectiveFunction::GetGradients(score, gradients, hessians) (objective_function.h:37), device pointers */
GBL_API int gbl_objective_get_gradients(gbl_objective* obj, const double* score_dev, float* gradients_dev,
                                        float* hessians_dev, void* stream);
/* ObjectiveFunction::BoostFromScore(class_id) (objective_function.h:66) */
GBL_API int gbl_objective_boost_from_score(gbl_objective* obj, double* out_init_score);
/* ObjectiveFunction::IsConstantHessian() */
GBL_API int gbl_objective_is_constant_hessian(gbl_objective* obj, int32_t* out);

/* ---------------------------------------------------------------- src/metric (evaluation on device-resident scores) */
typedef struct gbl_metric_config {
  int32_t kind;            /* GBL_METRIC_* */
  int32_t num_eval_at;     /* ndcg */
  const int32_t* eval_at;  /* ndcg: positions, ascending as Config sorts them (src/io/config.cpp:287-288); <= 4096 */
  int32_t num_label_gain;  /* ndcg */
  int32_t reserved;
  const double* label_gain; /* ndcg: [num_label_gain], default 2^i - 1 (dcg_calculator.cpp:33-43) */
  double sigmoid;          /* binary_logloss / binary_error: > 0 = scores are raw and converted with
                              1 / (1 + exp(-sigmoid * score)) (BinaryLogloss::ConvertOutput, binary_objective.hpp:167-169),
                              the case `objective != nullptr` of BinaryMetric::Eval; <= 0 = scores are probabilities */
} gbl_metric_config;

/* Metric::Init(metadata, num_data) (include/LightGBM/metric.h:32): label / weight / query_boundaries are host arrays
 * copied to the device; weight and query_boundaries may be NULL (ndcg needs the boundaries, auc takes no weights). */
GBL_API int gbl_metric_create(const gbl_metric_config* config, int64_t num_data, const float* label_host,
                              const float* weight_host, const int32_t* query_boundaries_host, int32_t num_queries,
                              int32_t device_id, gbl_metric** out);
GBL_API int gbl_metric_free(gbl_metric* metric);
/* number of values Eval returns: one per eval_at position for ndcg (NDCGMetric::GetName, rank_metric.hpp:34-36), else 1 */
GBL_API int gbl_metric_num_outputs(const gbl_metric* metric, int32_t* out);
/* Metric::Eval(score, objective) (include/LightGBM/metric.h:47) on a DEVICE score vector: kernels on `stream`
 * (NULL = the default stream), then a wait for the few result bytes.  out_host[gbl_metric_num_outputs()]. */
GBL_API int gbl_metric_eval(gbl_metric* metric, const double* score_dev, void* stream, double* out_host);

/* ---------------------------------------------------------------- device memory helpers for host languages
 * that have no CUDA runtime binding (the adapter keeps scores/gradients device-resident like
 * src/boosting/cuda/cuda_score_updater.cpp:12-80) */
GBL_API int gbl_device_malloc(int32_t device_id, size_t bytes, void** out_dev);
GBL_API int gbl_device_free(int32_t device_id, void* ptr_dev);
/* Page-lock a host buffer the caller keeps passing to this library (the reference's gradient / hessian vectors, which
 * GBDT allocates once, src/boosting/gbdt.cpp:849-865): copies from pageable memory are staged and run at a fifth of the
 * PCIe rate.  Failure is reported but harmless (the copies then take the slow path). */
GBL_API int gbl_host_register(void* ptr_host, size_t bytes);
GBL_API int gbl_host_unregister(void* ptr_host);
GBL_API int gbl_memcpy_h2d(int32_t device_id, void* dst_dev, const void* src_host, size_t bytes);
GBL_API int gbl_memcpy_d2h(int32_t device_id, void* dst_host, const void* src_dev, size_t bytes);
GBL_API int gbl_device_fill_f64(int32_t device_id, double* dst_dev, double value, int64_t n);
GBL_API int gbl_device_synchronize(int32_t device_id);

#ifdef __cplusplus
}
#endif
#endif /* LAMBDAGAP_B200_H_ */
