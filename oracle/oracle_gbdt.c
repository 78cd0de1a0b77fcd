/*
 * oracle_gbdt.c -- TEST INFRASTRUCTURE ONLY.  CPU restatement, in plain C, of the reference's algorithm for the
 * histogram-GBDT hot path.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may use it; the
 * product (lambdagap_b200/) never links or calls it.
 *
 * Parity status: PINNED.  tests/test_oracle_cpu.py checks this file against golden fixtures generated from the
 * UNMODIFIED reference CPU library (oracle/_ref/lib_lightgbm.so, built from /root/reference by oracle/Makefile.ref)
 * with tests/golden/make_golden.py; on machines where oracle/_ref exists the GPU tests also compare against the
 * reference directly.
 *
 * It follows the reference's *sequential* formulation (row-order fp64 sums, sequential threshold loops with their
 * continue / break gates, FixHistogram, parent-minus-sibling subtraction), deliberately different from the
 * block-parallel CUDA formulation, so that agreement between the two is evidence and not a tautology.
 * Bins are mapper bins (BinIterator::Get); the most-frequent bin is skipped while accumulating and reconstructed
 * by FixHistogram exactly like the reference's stored-bin-0 elision.
 *
 * Reference (all under /root/reference):
 *   histogram        src/io/dense_bin.hpp:99-222, src/io/dataset.cpp:1262-1454
 *   FixHistogram     src/io/dataset.cpp:1488-1545
 *   Subtract         src/treelearner/feature_histogram.hpp:96-145
 *   threshold scan   src/treelearner/feature_histogram.hpp:830-1057 (float), :1059-1350 (quantized)
 *   gain / output    src/treelearner/feature_histogram.hpp:711-828
 *   tree loop        src/treelearner/serial_tree_learner.cpp:179-245, 340-384, 477-622, 766-922
 *   partition        src/io/dense_bin.hpp:316-394, src/treelearner/data_partition.hpp:101-121
 *   discretizer      src/treelearner/gradient_discretizer.cpp:68-199
 *   objectives       src/objective/regression_objective.hpp:112-127, binary_objective.hpp:58-166,
 *                    rank_objective.hpp:225-552, src/metric/dcg_calculator.cpp:33-96
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define ORC_API __attribute__((visibility("default")))
#define K_EPSILON ((double)1e-15f)
#define K_MIN_SCORE (-INFINITY)

typedef struct {
  int32_t num_bin, missing_type, default_bin, most_freq_bin, real_index;
} orc_meta;

typedef struct {
  int32_t num_leaves, max_depth, min_data_in_leaf, use_quantized_grad, num_grad_quant_bins, is_constant_hessian;
  double min_sum_hessian_in_leaf, lambda_l1, lambda_l2, min_gain_to_split, max_delta_step, path_smooth;
} orc_config;

typedef struct {
  int32_t leaf, feature, real_feature, threshold_bin, default_left, left_count, right_count, pad;
  double left_output, right_output, left_sum_g, left_sum_h, right_sum_g, right_sum_h, gain;
  int64_t left_ig, left_ih, right_ig, right_ih;
} orc_split;

/* ------------------------------------------------------------------------------------------- gain arithmetic */
static int sign_of(double x) { return (x > 0.0) - (x < 0.0); }
static double threshold_l1(double s, double l1) {
  const double reg_s = fmax(0.0, fabs(s) - l1);
  return sign_of(s) * reg_s;
}
typedef struct {
  double l1, l2, max_delta_step, path_smooth;
  int use_l1, use_max_output, use_smoothing;
} gain_cfg;
static gain_cfg make_gain_cfg(const orc_config* c) {
  gain_cfg g;
  g.l1 = c->lambda_l1;
  g.l2 = c->lambda_l2;
  g.max_delta_step = c->max_delta_step;
  g.path_smooth = c->path_smooth;
  g.use_l1 = c->lambda_l1 > 0.0;
  g.use_max_output = c->max_delta_step > 0.0;
  g.use_smoothing = c->path_smooth > K_EPSILON;
  return g;
}
static double leaf_output(double sg, double sh, const gain_cfg* c, int use_l1, int use_max_output, int use_smoothing,
                          int num_data, double parent_output) {
  double ret = use_l1 ? -threshold_l1(sg, c->l1) / (sh + c->l2) : -sg / (sh + c->l2);
  if (use_max_output && c->max_delta_step > 0 && fabs(ret) > c->max_delta_step) ret = sign_of(ret) * c->max_delta_step;
  if (use_smoothing) {
    ret = ret * (num_data / c->path_smooth) / (num_data / c->path_smooth + 1) + parent_output / (num_data / c->path_smooth + 1);
  }
  return ret;
}
static double leaf_gain_given_output(double sg, double sh, const gain_cfg* c, double output) {
  if (c->use_l1) {
    const double s = threshold_l1(sg, c->l1);
    return -(2.0 * s * output + (sh + c->l2) * output * output);
  }
  return -(2.0 * sg * output + (sh + c->l2) * output * output);
}
static double leaf_gain(double sg, double sh, const gain_cfg* c, int num_data, double parent_output) {
  if (!c->use_max_output && !c->use_smoothing) {
    if (c->use_l1) {
      const double s = threshold_l1(sg, c->l1);
      return (s * s) / (sh + c->l2);
    }
    return (sg * sg) / (sh + c->l2);
  }
  return leaf_gain_given_output(sg, sh, c, leaf_output(sg, sh, c, c->use_l1, c->use_max_output, c->use_smoothing, num_data, parent_output));
}
static double split_gains(double lg, double lh, double rg, double rh, const gain_cfg* c, int lc, int rc, double po) {
  return leaf_gain(lg, lh, c, lc, po) + leaf_gain(rg, rh, c, rc, po);
}
static int round_int(double x) { return (int)(x + 0.5f); }

/* ------------------------------------------------------------------------------------------- histograms */
/* Histogram of one feature over the rows idx[0..n) (idx NULL = all rows).  The most-frequent bin is not
 * accumulated (the reference stores it as "0" and skips it); fp64 sums in row order. */
ORC_API void orc_histogram_float(const uint8_t* col, const orc_meta* m, const int32_t* idx, int n, const float* g,
                                 const float* h, double* hist /* [num_bin][2] */) {
  memset(hist, 0, sizeof(double) * 2 * m->num_bin);
  for (int i = 0; i < n; ++i) {
    const int r = idx ? idx[i] : i;
    const int b = col[r];
    if (b == m->most_freq_bin) continue;
    hist[2 * b] += g[r];
    hist[2 * b + 1] += h[r];
  }
}
ORC_API void orc_histogram_int(const uint8_t* col, const orc_meta* m, const int32_t* idx, int n, const int8_t* gh /* [hess, grad] */,
                               int64_t* hist /* [num_bin][2] = grad, hess */) {
  memset(hist, 0, sizeof(int64_t) * 2 * m->num_bin);
  for (int i = 0; i < n; ++i) {
    const int r = idx ? idx[i] : i;
    const int b = col[r];
    if (b == m->most_freq_bin) continue;
    hist[2 * b] += gh[2 * r + 1];
    hist[2 * b + 1] += (uint8_t)gh[2 * r];
  }
}

/* GradientDiscretizer::DiscretizeGradients; rand_g / rand_h NULL = stochastic_rounding=false */
ORC_API void orc_discretize(const float* g, const float* h, int n, int num_bins, int is_constant_hessian,
                            const double* rand_g, const double* rand_h, int rand_start, int8_t* out,
                            double* grad_scale, double* hess_scale) {
  double max_g = fabs(g[0]), max_h = fabs(h[0]);
  for (int i = 0; i < n; ++i) {
    if (fabs(g[i]) > max_g) max_g = fabs(g[i]);
    if (fabs(h[i]) > max_h) max_h = fabs(h[i]);
  }
  const double gs = max_g / (double)(num_bins / 2);
  const double hs = is_constant_hessian ? max_h : max_h / (double)num_bins;
  const double ig = 1.0f / gs, ih = 1.0f / hs;
  for (int i = 0; i < n; ++i) {
    const double gradient = g[i];
    double rg = 0.5, rh = 0.5;
    if (rand_g) {
      const int pos = (int)(((int64_t)i + rand_start) % n);
      rg = rand_g[pos];
      rh = rand_h[pos];
    }
    out[2 * i + 1] = gradient >= 0.0f ? (int8_t)(gradient * ig + rg) : (int8_t)(gradient * ig - rg);
    out[2 * i] = is_constant_hessian ? (int8_t)1 : (int8_t)(h[i] * ih + rh);
  }
  *grad_scale = gs;
  *hess_scale = hs;
}

/* ------------------------------------------------------------------------------------------- threshold scan */
typedef struct {
  double gain;
  int feature, real_feature, threshold, default_left, left_count, right_count;
  double left_output, right_output, left_sum_g, left_sum_h, right_sum_g, right_sum_h;
  int64_t left_ig, left_ih, right_ig, right_ih;
} split_info;

static void split_reset(split_info* s) {
  memset(s, 0, sizeof(*s));
  s->gain = K_MIN_SCORE;
  s->feature = -1;
  s->real_feature = -1;
  s->default_left = 1;
}
/* SplitInfo::operator> */
static int split_better(const split_info* a, const split_info* b) {
  if (a->gain != b->gain) return a->gain > b->gain;
  const int fa = a->real_feature == -1 ? INT32_MAX : a->real_feature;
  const int fb = b->real_feature == -1 ? INT32_MAX : b->real_feature;
  return fa < fb;
}

/* One pass of FindBestThresholdSequentially (float).  hist is [num_bin][2] indexed by mapper bin. */
static void scan_float(const double* hist, const orc_meta* m, const orc_config* cfg, const gain_cfg* gc, double sum_gradient,
                       double sum_hessian, int num_data, double min_gain_shift, double parent_output, int reverse,
                       int skip_default, int na_missing, int* is_splittable, split_info* out) {
  const int offset = m->most_freq_bin == 0;
  double best_lg = NAN, best_lh = NAN, best_gain = K_MIN_SCORE;
  int best_left_count = 0, best_threshold = m->num_bin;
  const double cnt_factor = num_data / sum_hessian;
  if (reverse) {
    double rg = 0.0f, rh = K_EPSILON;
    int rc = 0;
    int t = m->num_bin - 1 - offset - na_missing;
    const int t_end = 1 - offset;
    for (; t >= t_end; --t) {
      const int b = t + offset;
      if (skip_default && b == m->default_bin) continue;
      const double grad = hist[2 * b], hess = hist[2 * b + 1];
      const int cnt = round_int(hess * cnt_factor);
      rg += grad;
      rh += hess;
      rc += cnt;
      if (rc < cfg->min_data_in_leaf || rh < cfg->min_sum_hessian_in_leaf) continue;
      const int lc = num_data - rc;
      if (lc < cfg->min_data_in_leaf) break;
      const double lh = sum_hessian - rh;
      if (lh < cfg->min_sum_hessian_in_leaf) break;
      const double lg = sum_gradient - rg;
      const double gain = split_gains(lg, lh, rg, rh, gc, lc, rc, parent_output);
      if (gain <= min_gain_shift) continue;
      *is_splittable = 1;
      if (gain > best_gain) {
        best_left_count = lc;
        best_lg = lg;
        best_lh = lh;
        best_threshold = b - 1;
        best_gain = gain;
      }
    }
  } else {
    double lg = 0.0f, lh = K_EPSILON;
    int lc = 0;
    int t = 0;
    const int t_end = m->num_bin - 2 - offset;
    if (na_missing && offset == 1) {
      lg = sum_gradient;
      lh = sum_hessian - K_EPSILON;
      lc = num_data;
      for (int i = 0; i < m->num_bin - offset; ++i) {
        const int b = i + offset;
        lg -= hist[2 * b];
        lh -= hist[2 * b + 1];
        lc -= round_int(hist[2 * b + 1] * cnt_factor);
      }
      t = -1;
    }
    for (; t <= t_end; ++t) {
      const int b = t + offset;
      if (skip_default && b == m->default_bin) continue;
      if (t >= 0) {
        lg += hist[2 * b];
        lh += hist[2 * b + 1];
        lc += round_int(hist[2 * b + 1] * cnt_factor);
      }
      if (lc < cfg->min_data_in_leaf || lh < cfg->min_sum_hessian_in_leaf) continue;
      const int rc = num_data - lc;
      if (rc < cfg->min_data_in_leaf) break;
      const double rh = sum_hessian - lh;
      if (rh < cfg->min_sum_hessian_in_leaf) break;
      const double rg = sum_gradient - lg;
      const double gain = split_gains(lg, lh, rg, rh, gc, lc, rc, parent_output);
      if (gain <= min_gain_shift) continue;
      *is_splittable = 1;
      if (gain > best_gain) {
        best_left_count = lc;
        best_lg = lg;
        best_lh = lh;
        best_threshold = b;
        best_gain = gain;
      }
    }
  }
  if (*is_splittable && best_gain > out->gain + min_gain_shift) {
    out->threshold = best_threshold;
    out->left_output = leaf_output(best_lg, best_lh, gc, gc->use_l1, gc->use_max_output, gc->use_smoothing, best_left_count, parent_output);
    out->left_count = best_left_count;
    out->left_sum_g = best_lg;
    out->left_sum_h = best_lh - K_EPSILON;
    out->right_output = leaf_output(sum_gradient - best_lg, sum_hessian - best_lh, gc, gc->use_l1, gc->use_max_output,
                                    gc->use_smoothing, num_data - best_left_count, parent_output);
    out->right_count = num_data - best_left_count;
    out->right_sum_g = sum_gradient - best_lg;
    out->right_sum_h = sum_hessian - best_lh - K_EPSILON;
    out->gain = best_gain - min_gain_shift;
    out->default_left = reverse;
  }
}

/* One pass of FindBestThresholdSequentiallyInt.  hist is [num_bin][2] = (grad sum, hess sum) integers. */
static void scan_int(const int64_t* hist, const orc_meta* m, const orc_config* cfg, const gain_cfg* gc, int64_t tot_g, int64_t tot_h,
                     double grad_scale, double hess_scale, int num_data, double min_gain_shift, double parent_output, int reverse,
                     int skip_default, int na_missing, int* is_splittable, split_info* out) {
  const int offset = m->most_freq_bin == 0;
  int64_t best_lg = 0, best_lh = 0;
  double best_gain = K_MIN_SCORE;
  int best_threshold = m->num_bin;
  const double cnt_factor = (double)num_data / (double)(uint32_t)tot_h;
  if (reverse) {
    int64_t rg = 0, rh = 0;
    int t = m->num_bin - 1 - offset - na_missing;
    const int t_end = 1 - offset;
    for (; t >= t_end; --t) {
      const int b = t + offset;
      if (skip_default && b == m->default_bin) continue;
      rg += hist[2 * b];
      rh += hist[2 * b + 1];
      const uint32_t irh = (uint32_t)rh;
      const int rc = round_int(irh * cnt_factor);
      const double srh = irh * hess_scale;
      if (rc < cfg->min_data_in_leaf || srh < cfg->min_sum_hessian_in_leaf) continue;
      const int lc = num_data - rc;
      if (lc < cfg->min_data_in_leaf) break;
      const int64_t lgi = tot_g - rg, lhi = tot_h - rh;
      const double slh = (uint32_t)lhi * hess_scale;
      if (slh < cfg->min_sum_hessian_in_leaf) break;
      const double srg = (double)(int32_t)rg * grad_scale, slg = (double)(int32_t)lgi * grad_scale;
      const double gain = split_gains(slg, slh + K_EPSILON, srg, srh + K_EPSILON, gc, lc, rc, parent_output);
      if (gain <= min_gain_shift) continue;
      *is_splittable = 1;
      if (gain > best_gain) {
        best_lg = lgi;
        best_lh = lhi;
        best_threshold = b - 1;
        best_gain = gain;
      }
    }
  } else {
    int64_t lg = 0, lh = 0;
    int t = 0;
    const int t_end = m->num_bin - 2 - offset;
    if (na_missing && offset == 1) {
      lg = tot_g;
      lh = tot_h;
      for (int i = 0; i < m->num_bin - offset; ++i) {
        lg -= hist[2 * (i + offset)];
        lh -= hist[2 * (i + offset) + 1];
      }
      t = -1;
    }
    for (; t <= t_end; ++t) {
      const int b = t + offset;
      if (skip_default && b == m->default_bin) continue;
      if (t >= 0) {
        lg += hist[2 * b];
        lh += hist[2 * b + 1];
      }
      const uint32_t ilh = (uint32_t)lh;
      const int lc = round_int((double)ilh * cnt_factor);
      const double slh = (double)ilh * hess_scale;
      if (lc < cfg->min_data_in_leaf || slh < cfg->min_sum_hessian_in_leaf) continue;
      const int rc = num_data - lc;
      if (rc < cfg->min_data_in_leaf) break;
      const int64_t rgi = tot_g - lg, rhi = tot_h - lh;
      const double srh = (double)(uint32_t)rhi * hess_scale;
      if (srh < cfg->min_sum_hessian_in_leaf) break;
      const double srg = (double)(int32_t)rgi * grad_scale, slg = (double)(int32_t)lg * grad_scale;
      const double gain = split_gains(slg, slh + K_EPSILON, srg, srh + K_EPSILON, gc, lc, rc, parent_output);
      if (gain <= min_gain_shift) continue;
      *is_splittable = 1;
      if (gain > best_gain) {
        best_lg = lg;
        best_lh = lh;
        best_threshold = b;
        best_gain = gain;
      }
    }
  }
  if (*is_splittable && best_gain > out->gain + min_gain_shift) {
    const int64_t rgi = tot_g - best_lg, rhi = tot_h - best_lh;
    const double blg = (double)(int32_t)best_lg * grad_scale, blh = (double)(uint32_t)best_lh * hess_scale;
    const double brg = (double)(int32_t)rgi * grad_scale, brh = (double)(uint32_t)rhi * hess_scale;
    const int lc = round_int((double)(uint32_t)best_lh * cnt_factor), rc = round_int((double)(uint32_t)rhi * cnt_factor);
    out->threshold = best_threshold;
    out->left_output = leaf_output(blg, blh, gc, gc->use_l1, gc->use_max_output, gc->use_smoothing, lc, parent_output);
    out->left_count = lc;
    out->left_sum_g = blg;
    out->left_sum_h = blh;
    out->left_ig = best_lg;
    out->left_ih = best_lh;
    out->right_output = leaf_output(brg, brh, gc, gc->use_l1, gc->use_max_output, gc->use_smoothing, rc, parent_output);
    out->right_count = rc;
    out->right_sum_g = brg;
    out->right_sum_h = brh;
    out->right_ig = rgi;
    out->right_ih = rhi;
    out->gain = best_gain - min_gain_shift;
    out->default_left = reverse;
  }
}

/* FeatureHistogram::FindBestThreshold[Int] for one feature: dispatch of FuncForNumricalL3 (:272-447) */
static void find_best_threshold(const void* hist, int quant, const orc_meta* m, const orc_config* cfg, double sum_g, double sum_h,
                                int64_t ig, int64_t ih, double grad_scale, double hess_scale, int num_data, double parent_output,
                                int* is_splittable, split_info* out) {
  const gain_cfg gc = make_gain_cfg(cfg);
  split_reset(out);
  *is_splittable = 0;
  const int two = m->num_bin > 2 && m->missing_type != 0;
  const int skip_default = two && m->missing_type == 1;
  const int na_missing = two && m->missing_type == 2;
  if (quant) {
    const double sg = (double)(int32_t)ig * grad_scale, sh = (double)(uint32_t)ih * hess_scale;
    const double mgs = leaf_gain(sg, sh, &gc, num_data, parent_output) + cfg->min_gain_to_split;
    scan_int((const int64_t*)hist, m, cfg, &gc, ig, ih, grad_scale, hess_scale, num_data, mgs, parent_output, 1, skip_default, na_missing, is_splittable, out);
    if (two) scan_int((const int64_t*)hist, m, cfg, &gc, ig, ih, grad_scale, hess_scale, num_data, mgs, parent_output, 0, skip_default, na_missing, is_splittable, out);
  } else {
    const double sh = sum_h + 2 * K_EPSILON;
    const double mgs = leaf_gain(sum_g, sh, &gc, num_data, parent_output) + cfg->min_gain_to_split;
    scan_float((const double*)hist, m, cfg, &gc, sum_g, sh, num_data, mgs, parent_output, 1, skip_default, na_missing, is_splittable, out);
    if (two) scan_float((const double*)hist, m, cfg, &gc, sum_g, sh, num_data, mgs, parent_output, 0, skip_default, na_missing, is_splittable, out);
  }
  if (!two && m->missing_type == 2) out->default_left = 0;
  /* gain *= penalty (1.0) */
}

/* ------------------------------------------------------------------------------------------- tree learner */
typedef struct {
  int begin, count, depth;
  double sum_g, sum_h, output;
  int64_t ig, ih;
  split_info best;
} leaf_t;

/* SerialTreeLearner::Train.  cols[f] = mapper bins of feature f for all N rows.  gh_int8 is only read in quantized
 * mode (int8 [hess, grad] pairs + scales).  leaf_of_row (may be NULL) receives the final leaf of every row. */
ORC_API int orc_train_tree(const uint8_t* const* cols, const orc_meta* metas, int F, int N, const float* g, const float* h,
                           const int8_t* gh_int8, double grad_scale, double hess_scale, const orc_config* cfg,
                           orc_split* out_splits, int32_t* out_num_splits, double* out_root_output, int32_t* leaf_of_row) {
  const int quant = cfg->use_quantized_grad;
  const int NL = cfg->num_leaves;
  const gain_cfg gc = make_gain_cfg(cfg);
  const size_t hist_stride = (size_t)F * 256 * 2;
  double* pool_f = quant ? NULL : (double*)calloc((size_t)NL * hist_stride, sizeof(double));
  int64_t* pool_i = quant ? (int64_t*)calloc((size_t)NL * hist_stride, sizeof(int64_t)) : NULL;
  int* slot_of_leaf = (int*)malloc(sizeof(int) * NL);
  unsigned char* splittable = (unsigned char*)malloc((size_t)NL * F);
  int32_t* idx = (int32_t*)malloc(sizeof(int32_t) * N);
  int32_t* tmp = (int32_t*)malloc(sizeof(int32_t) * N);
  leaf_t* leaves = (leaf_t*)calloc(NL, sizeof(leaf_t));
  if ((!pool_f && !pool_i) || !slot_of_leaf || !splittable || !idx || !tmp || !leaves) return -1;
  for (int i = 0; i < N; ++i) idx[i] = i;
  for (int l = 0; l < NL; ++l) {
    split_reset(&leaves[l].best);
    slot_of_leaf[l] = l;
  }
  memset(splittable, 1, (size_t)NL * F);
  /* root sums: LeafSplits::Init */
  leaf_t* root = &leaves[0];
  root->begin = 0;
  root->count = N;
  root->depth = 0;
  if (quant) {
    double sg = 0.0f, sh = 0.0f;
    int64_t ig = 0, ih = 0;
    for (int i = 0; i < N; ++i) {
      sg += gh_int8[2 * i + 1] * grad_scale;
      sh += gh_int8[2 * i] * hess_scale;
      ig += gh_int8[2 * i + 1];
      ih += (uint8_t)gh_int8[2 * i];
    }
    root->sum_g = sg;
    root->sum_h = sh;
    root->ig = ig;
    root->ih = ih;
  } else {
    double sg = 0.0f, sh = 0.0f;
    for (int i = 0; i < N; ++i) {
      sg += g[i];
      sh += h[i];
    }
    root->sum_g = sg;
    root->sum_h = sh;
  }
  root->output = leaf_output(root->sum_g, root->sum_h, &gc, 1, 1, 0, N, 0.0);
  *out_root_output = root->output;
  int num_leaves = 1, num_splits = 0;
  int left_leaf = 0, right_leaf = -1;
  for (int split = 0; split < NL - 1; ++split) {
    /* BeforeFindBestSplit */
    int do_find = 1;
    if (cfg->max_depth > 0 && leaves[left_leaf].depth >= cfg->max_depth) do_find = 0;
    const int n_left = leaves[left_leaf].count, n_right = right_leaf >= 0 ? leaves[right_leaf].count : 0;
    if (n_right < cfg->min_data_in_leaf * 2 && n_left < cfg->min_data_in_leaf * 2) do_find = 0;
    if (!do_find) {
      leaves[left_leaf].best.gain = K_MIN_SCORE;
      if (right_leaf >= 0) leaves[right_leaf].best.gain = K_MIN_SCORE;
    } else {
      int smaller, larger;
      if (right_leaf < 0) {
        smaller = left_leaf;
        larger = -1;
      } else if (n_left < n_right) {
        smaller = left_leaf;
        larger = right_leaf;
        /* parent (left id) histogram moves to the larger (right) leaf; smaller gets a fresh slot */
        const int parent_slot = slot_of_leaf[left_leaf];
        slot_of_leaf[right_leaf] = parent_slot;
        slot_of_leaf[left_leaf] = right_leaf; /* fresh: slot index == new leaf id */
        memcpy(splittable + (size_t)slot_of_leaf[left_leaf] * F, splittable + (size_t)parent_slot * F, F);
      } else {
        smaller = right_leaf;
        larger = left_leaf;
        slot_of_leaf[right_leaf] = right_leaf;
        memcpy(splittable + (size_t)right_leaf * F, splittable + (size_t)slot_of_leaf[left_leaf] * F, F);
      }
      split_info best_s, best_l;
      split_reset(&best_s);
      split_reset(&best_l);
      for (int f = 0; f < F; ++f) {
        const orc_meta* m = &metas[f];
        const size_t so = (size_t)slot_of_leaf[smaller] * hist_stride + (size_t)f * 512;
        unsigned char* flag_s = splittable + (size_t)slot_of_leaf[smaller] * F + f;
        if (larger >= 0 && !*flag_s) { /* parent unsplittable: serial_tree_learner.cpp:394-399 (flags were inherited) */
          continue;
        }
        /* smaller leaf: construct + FixHistogram */
        const leaf_t* S = &leaves[smaller];
        int is_spl = 0;
        split_info cand;
        if (quant) {
          int64_t* hs = pool_i + so;
          orc_histogram_int(cols[f], m, idx + S->begin, S->count, gh_int8, hs);
          if (m->most_freq_bin > 0) {
            int64_t fg = S->ig, fh = S->ih;
            for (int b = 0; b < m->num_bin; ++b) {
              if (b != m->most_freq_bin) {
                fg -= hs[2 * b];
                fh -= hs[2 * b + 1];
              }
            }
            hs[2 * m->most_freq_bin] = fg;
            hs[2 * m->most_freq_bin + 1] = fh;
          }
          find_best_threshold(hs, 1, m, cfg, 0, 0, S->ig, S->ih, grad_scale, hess_scale, S->count, S->output, &is_spl, &cand);
        } else {
          double* hs = pool_f + so;
          orc_histogram_float(cols[f], m, idx + S->begin, S->count, g, h, hs);
          if (m->most_freq_bin > 0) {
            hs[2 * m->most_freq_bin] = S->sum_g;
            hs[2 * m->most_freq_bin + 1] = S->sum_h;
            for (int b = 0; b < m->num_bin; ++b) {
              if (b != m->most_freq_bin) {
                hs[2 * m->most_freq_bin] -= hs[2 * b];
                hs[2 * m->most_freq_bin + 1] -= hs[2 * b + 1];
              }
            }
          }
          find_best_threshold(hs, 0, m, cfg, S->sum_g, S->sum_h, 0, 0, 0, 0, S->count, S->output, &is_spl, &cand);
        }
        *flag_s = (unsigned char)is_spl;
        cand.feature = f;
        cand.real_feature = m->real_index;
        if (split_better(&cand, &best_s)) best_s = cand;
        if (larger < 0) continue;
        /* larger leaf: parent - smaller */
        const leaf_t* Lg = &leaves[larger];
        const size_t lo = (size_t)slot_of_leaf[larger] * hist_stride + (size_t)f * 512;
        unsigned char* flag_l = splittable + (size_t)slot_of_leaf[larger] * F + f;
        if (quant) {
          int64_t* hl = pool_i + lo;
          const int64_t* hs = pool_i + so;
          for (int b = 0; b < 2 * m->num_bin; ++b) hl[b] -= hs[b];
          find_best_threshold(hl, 1, m, cfg, 0, 0, Lg->ig, Lg->ih, grad_scale, hess_scale, Lg->count, Lg->output, &is_spl, &cand);
        } else {
          double* hl = pool_f + lo;
          const double* hs = pool_f + so;
          for (int b = 0; b < 2 * m->num_bin; ++b) hl[b] -= hs[b];
          find_best_threshold(hl, 0, m, cfg, Lg->sum_g, Lg->sum_h, 0, 0, 0, 0, Lg->count, Lg->output, &is_spl, &cand);
        }
        *flag_l = (unsigned char)is_spl;
        cand.feature = f;
        cand.real_feature = m->real_index;
        if (split_better(&cand, &best_l)) best_l = cand;
      }
      leaves[smaller].best = best_s;
      if (larger >= 0) leaves[larger].best = best_l;
    }
    /* best leaf: ArrayArgs::ArgMax over all num_leaves slots (unsplit ones are reset) */
    int best_leaf = 0;
    for (int l = 1; l < num_leaves; ++l) {
      if (split_better(&leaves[l].best, &leaves[best_leaf].best)) best_leaf = l;
    }
    const split_info bs = leaves[best_leaf].best;
    if (bs.gain <= 0.0) break;
    /* DataPartition::Split: stable, left first */
    const orc_meta* m = &metas[bs.feature];
    leaf_t P = leaves[best_leaf];
    int missing_bin = -1;
    if (m->missing_type == 1) missing_bin = m->default_bin;
    if (m->missing_type == 2) missing_bin = m->num_bin - 1;
    int nl = 0, nr = 0;
    for (int i = 0; i < P.count; ++i) {
      const int r = idx[P.begin + i];
      const int b = cols[bs.feature][r];
      const int go_left = b == missing_bin ? bs.default_left : b <= bs.threshold;
      if (go_left) {
        idx[P.begin + nl++] = r;
      } else {
        tmp[nr++] = r;
      }
    }
    memcpy(idx + P.begin + nl, tmp, sizeof(int32_t) * nr);
    const int right_id = num_leaves;
    orc_split* rec = &out_splits[num_splits];
    memset(rec, 0, sizeof(*rec));
    rec->leaf = best_leaf;
    rec->feature = bs.feature;
    rec->real_feature = bs.real_feature;
    rec->threshold_bin = bs.threshold;
    rec->default_left = bs.default_left;
    rec->left_count = nl;
    rec->right_count = nr;
    rec->left_output = bs.left_output;
    rec->right_output = bs.right_output;
    rec->left_sum_g = bs.left_sum_g;
    rec->left_sum_h = bs.left_sum_h;
    rec->right_sum_g = bs.right_sum_g;
    rec->right_sum_h = bs.right_sum_h;
    rec->gain = bs.gain;
    rec->left_ig = bs.left_ig;
    rec->left_ih = bs.left_ih;
    rec->right_ig = bs.right_ig;
    rec->right_ih = bs.right_ih;
    leaf_t* A = &leaves[best_leaf];
    leaf_t* B = &leaves[right_id];
    A->count = nl;
    A->depth = P.depth + 1;
    A->sum_g = bs.left_sum_g;
    A->sum_h = bs.left_sum_h;
    A->ig = bs.left_ig;
    A->ih = bs.left_ih;
    A->output = bs.left_output;
    B->begin = P.begin + nl;
    B->count = nr;
    B->depth = P.depth + 1;
    B->sum_g = bs.right_sum_g;
    B->sum_h = bs.right_sum_h;
    B->ig = bs.right_ig;
    B->ih = bs.right_ih;
    B->output = bs.right_output;
    split_reset(&B->best);
    ++num_leaves;
    ++num_splits;
    left_leaf = best_leaf;
    right_leaf = right_id;
  }
  *out_num_splits = num_splits;
  if (leaf_of_row) {
    for (int l = 0; l < num_leaves; ++l) {
      for (int i = 0; i < leaves[l].count; ++i) leaf_of_row[idx[leaves[l].begin + i]] = l;
    }
  }
  free(pool_f);
  free(pool_i);
  free(slot_of_leaf);
  free(splittable);
  free(idx);
  free(tmp);
  free(leaves);
  return 0;
}

/* ------------------------------------------------------------------------------------------- objectives */
ORC_API void orc_l2_gradients(const double* score, const float* label, const float* weight, int n, float* g, float* h) {
  for (int i = 0; i < n; ++i) {
    if (!weight) {
      g[i] = (float)(score[i] - label[i]);
      h[i] = 1.0f;
    } else {
      g[i] = (float)((float)(score[i] - label[i]) * weight[i]);
      h[i] = weight[i];
    }
  }
}
ORC_API void orc_binary_gradients(const double* score, const float* label, const float* weight, int n, double sigmoid,
                                  double w_neg, double w_pos, float* g, float* h) {
  for (int i = 0; i < n; ++i) {
    const int is_pos = label[i] > 0;
    const int lab = is_pos ? 1 : -1;
    const double lw = is_pos ? w_pos : w_neg;
    const double response = -lab * sigmoid / (1.0f + exp(lab * sigmoid * score[i]));
    const double abs_response = fabs(response);
    if (!weight) {
      g[i] = (float)(response * lw);
      h[i] = (float)(abs_response * (sigmoid - abs_response) * lw);
    } else {
      g[i] = (float)(response * lw * weight[i]);
      h[i] = (float)(abs_response * (sigmoid - abs_response) * lw * weight[i]);
    }
  }
}
