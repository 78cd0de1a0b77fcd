/*
 * lgbm_b200_c_api.h -- the subset of the reference's C API (include/LightGBM/c_api.h) that the B200 host
 * library (lambdagap_b200/lib/liblgbm_b200.so) exports under the SAME names and signatures, so that a program
 * written against lib_lightgbm.so for the histogram-GBDT hot path (dense matrices; objectives regression /
 * binary / lambdarank incl. the fork's 18 lambdarank_target values; device_type=cuda) can switch libraries.
 * Each entry cites the reference definition it replaces.  Return 0 on success, -1 on error
 * (LGBM_GetLastError), exactly like API_BEGIN/API_END (src/c_api.cpp:38-52).
 */
#ifndef LGBM_B200_C_API_H_
#define LGBM_B200_C_API_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* DatasetHandle; /* include/LightGBM/c_api.h:27 */
typedef void* BoosterHandle; /* :28 */

#define C_API_DTYPE_FLOAT32 (0) /* :30 */
#define C_API_DTYPE_FLOAT64 (1)
#define C_API_DTYPE_INT32 (2)
#define C_API_PREDICT_NORMAL (0)    /* :35 */
#define C_API_PREDICT_RAW_SCORE (1) /* :36 */

#define LGBM_B200_EXPORT __attribute__((visibility("default")))

/* src/c_api.cpp:54 */
LGBM_B200_EXPORT const char* LGBM_GetLastError(void);
/* src/c_api.cpp:1303 (c_api.h:409) */
LGBM_B200_EXPORT int LGBM_DatasetCreateFromMat(const void* data, int data_type, int32_t nrow, int32_t ncol,
                                               int is_row_major, const char* parameters,
                                               const DatasetHandle reference, DatasetHandle* out);
/* Analogue of LGBM_DatasetCreateFromSampledColumn (src/c_api.cpp:1063-1100) for dense matrices: bin mappers are
 * found from `sample` (every row, same dtype / layout as data), then `data` (this rank's shard of `num_total_row`
 * rows overall) is binned.  Every rank that passes the same sample derives identical mappers. */
LGBM_B200_EXPORT int LGBM_DatasetCreateFromMatWithSample(const void* data, int data_type, int32_t nrow, int32_t ncol,
                                                         int is_row_major, const void* sample, int32_t sample_nrow,
                                                         int64_t num_total_row, const char* parameters,
                                                         DatasetHandle* out);
/* src/c_api.cpp:1760 (c_api.h:552) */
LGBM_B200_EXPORT int LGBM_DatasetSetField(DatasetHandle handle, const char* field_name, const void* field_data,
                                          int num_element, int type);
/* src/c_api.cpp:1850,1858 */
LGBM_B200_EXPORT int LGBM_DatasetGetNumData(DatasetHandle handle, int32_t* out);
LGBM_B200_EXPORT int LGBM_DatasetGetNumFeature(DatasetHandle handle, int32_t* out);
/* src/c_api.cpp:1711 */
LGBM_B200_EXPORT int LGBM_DatasetFree(DatasetHandle handle);
/* src/c_api.cpp:1953 (c_api.h:656) */
LGBM_B200_EXPORT int LGBM_BoosterCreate(const DatasetHandle train_data, const char* parameters, BoosterHandle* out);
/* src/c_api.cpp:1994 */
LGBM_B200_EXPORT int LGBM_BoosterFree(BoosterHandle handle);
/* src/c_api.cpp:2076 (c_api.h:769) */
LGBM_B200_EXPORT int LGBM_BoosterUpdateOneIter(BoosterHandle handle, int* is_finished);
/* src/c_api.cpp:2087 (c_api.h:801) */
LGBM_B200_EXPORT int LGBM_BoosterUpdateOneIterCustom(BoosterHandle handle, const float* grad, const float* hess,
                                                     int* is_finished);
/* src/c_api.cpp:2112,2119 */
/* include/LightGBM/c_api.h:729 -- the data must be binned with the training data's mappers (create it with
 * `reference` = the training dataset) and must be added before the first iteration */
LGBM_B200_EXPORT int LGBM_BoosterAddValidData(BoosterHandle handle, const DatasetHandle valid_data);
/* include/LightGBM/c_api.h:846,861,918 -- metrics are evaluated on the device-resident scores; data_idx 0 = training
 * data, i = the i-th validation set */
LGBM_B200_EXPORT int LGBM_BoosterGetEvalCounts(BoosterHandle handle, int* out_len);
LGBM_B200_EXPORT int LGBM_BoosterGetEvalNames(BoosterHandle handle, const int len, int* out_len, const size_t buffer_len,
                                              size_t* out_buffer_len, char** out_strs);
LGBM_B200_EXPORT int LGBM_BoosterGetEval(BoosterHandle handle, int data_idx, int* out_len, double* out_results);

/* include/LightGBM/c_api.h:747 -- learning_rate, num_leaves, min_data_in_leaf, regularisation, bagging / feature
 * sampling parameters may change between iterations; objective, metric, boosting, device may not */
LGBM_B200_EXPORT int LGBM_BoosterResetParameter(BoosterHandle handle, const char* parameters);
/* include/LightGBM/c_api.h:756,828,904,1523,1551,1471 */
LGBM_B200_EXPORT int LGBM_BoosterGetNumClasses(BoosterHandle handle, int* out_len);
LGBM_B200_EXPORT int LGBM_BoosterNumModelPerIteration(BoosterHandle handle, int* out_tree_per_iteration);
LGBM_B200_EXPORT int LGBM_BoosterGetNumFeature(BoosterHandle handle, int* out_len);
LGBM_B200_EXPORT int LGBM_BoosterGetLeafValue(BoosterHandle handle, int tree_idx, int leaf_idx, double* out_val);
LGBM_B200_EXPORT int LGBM_BoosterFeatureImportance(BoosterHandle handle, int num_iteration, int importance_type,
                                                   double* out_results);
LGBM_B200_EXPORT int LGBM_BoosterSaveModel(BoosterHandle handle, int start_iteration, int num_iteration,
                                           int feature_importance_type, const char* filename);

LGBM_B200_EXPORT int LGBM_BoosterGetCurrentIteration(BoosterHandle handle, int* out_iteration);
LGBM_B200_EXPORT int LGBM_BoosterNumberOfTotalModel(BoosterHandle handle, int* out_models);
/* src/c_api.cpp:2200,2209 */
LGBM_B200_EXPORT int LGBM_BoosterGetNumPredict(BoosterHandle handle, int data_idx, int64_t* out_len);
LGBM_B200_EXPORT int LGBM_BoosterGetPredict(BoosterHandle handle, int data_idx, int64_t* out_len, double* out_result);
/* src/c_api.cpp:2483 */
LGBM_B200_EXPORT int LGBM_BoosterPredictForMat(BoosterHandle handle, const void* data, int data_type, int32_t nrow,
                                               int32_t ncol, int is_row_major, int predict_type, int start_iteration,
                                               int num_iteration, const char* parameter, int64_t* out_len,
                                               double* out_result);
/* src/c_api.cpp:2640 */
LGBM_B200_EXPORT int LGBM_BoosterSaveModelToString(BoosterHandle handle, int start_iteration, int num_iteration,
                                                   int feature_importance_type, int64_t buffer_len, int64_t* out_len,
                                                   char* out_str);
/* NCCL counterpart of LGBM_NetworkInitWithFunctions (src/c_api.cpp:2787-2795): binds this host thread / process
 * as rank `rank` of `num_machines` using a 128-byte NCCL unique id distributed by the caller
 * (LGBM_NetworkGetNcclUniqueId on rank 0).  Boosters created afterwards with tree_learner=data shard rows. */
LGBM_B200_EXPORT int LGBM_NetworkGetNcclUniqueId(uint8_t out_id[128]);
LGBM_B200_EXPORT int LGBM_NetworkInitWithNccl(const uint8_t id[128], int num_machines, int rank, int device_id);
/* src/c_api.cpp:2781 */
LGBM_B200_EXPORT int LGBM_NetworkFree(void);
/* not in the reference (test tap, runs on the host without a GPU): BinMapper::FindBin (src/io/bin.cpp:313-508) on one
 * column's sampled non-zero values.  out_scalars = {num_bin, missing_type, default_bin, most_freq_bin, is_trivial};
 * out_upper_bounds receives num_bin doubles (capacity `cap`). */
LGBM_B200_EXPORT int LGBM_B200_FindBin(double* values, int num_values, int64_t total_sample_cnt, int max_bin,
                                       int min_data_in_bin, int min_split_data, int pre_filter, int use_missing,
                                       int zero_as_missing, int32_t out_scalars[5], double* out_upper_bounds, int cap);
/* not in the reference: the C-ABI learner handle behind a booster (gbl_learner*), for timers and profiling */
/* test tap (no counterpart in the reference's C API): the parameter string as Config::Set understood it, as `[key: value]`
 * lines like the parameter block of a model file; needs no GPU */
LGBM_B200_EXPORT int LGBM_B200_DumpConfig(const char* parameters, int64_t buffer_len, int64_t* out_len, char* out_str);
LGBM_B200_EXPORT int LGBM_BoosterGetLearnerHandle(BoosterHandle handle, void** out_learner);
/* not in the reference: raw (unconverted) training scores, device -> host */
LGBM_B200_EXPORT int LGBM_BoosterGetTrainScoreRaw(BoosterHandle handle, int64_t* out_len, double* out_result);

#ifdef __cplusplus
}
#endif
#endif /* LGBM_B200_C_API_H_ */
