/*
 * oracle_rank.c -- TEST INFRASTRUCTURE ONLY.  CPU restatement in plain C of the fork's lambdarank objective with
 * its 18 `lambdarank_target` values: LambdarankNDCG::GetGradientsForOneQuery
 * (/root/reference/src/objective/rank_objective.hpp:253-524), GetSigmoid / ConstructSigmoidTable (:526-552),
 * Init (:225-251), RankingObjective::GetGradients weights (:98-105), DCGCalculator
 * (/root/reference/src/metric/dcg_calculator.cpp:33-96).  Sequential double loop over pairs, fp32 `+=` into the
 * lambdas, fp64 sum_lambdas -- the reference's order, unlike the CUDA kernel's per-document gather.
 * Parity status: PINNED by tests/test_oracle_cpu.py against fixtures generated from the unmodified reference.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define ORC_API __attribute__((visibility("default")))
#define SIGMOID_BINS (1024 * 1024)
#define MAX_POSITION 10000

enum {
  T_NDCG = 0, T_LL_NDCG, T_LL_NDCG_PP, T_BNDCG, T_LL_BNDCG, T_LL_BNDCG_PP, T_PRECISION, T_ARPK, T_LL_ARP1, T_LL_ARP2,
  T_RANKNET, T_BIN_RANKNET, T_GAP_S, T_GAP_X, T_GAP_S_P, T_GAP_X_P, T_GAP_S_PP, T_GAP_X_PP
};

typedef struct {
  int target, truncation_level, norm, num_label_gain;
  double sigmoid, lambdagap_weight;
  double min_in, max_in, idx_factor;
  double* table;
  double discount[MAX_POSITION];
  double label_gain[64];
} orc_rank;

ORC_API orc_rank* orc_rank_create(int target, int truncation_level, int norm, double sigmoid, double lambdagap_weight,
                                  const double* label_gain, int num_label_gain) {
  orc_rank* r = (orc_rank*)calloc(1, sizeof(orc_rank));
  r->target = target;
  r->truncation_level = truncation_level;
  r->norm = norm;
  r->sigmoid = sigmoid;
  r->lambdagap_weight = lambdagap_weight;
  if (label_gain && num_label_gain > 0) {
    r->num_label_gain = num_label_gain;
    memcpy(r->label_gain, label_gain, sizeof(double) * num_label_gain);
  } else {
    r->num_label_gain = 31;
    r->label_gain[0] = 0.0f;
    for (int i = 1; i < 31; ++i) r->label_gain[i] = (double)((1 << i) - 1);
  }
  for (int i = 0; i < MAX_POSITION; ++i) r->discount[i] = 1.0 / log2(2.0 + i);
  r->min_in = -50.0 / sigmoid / 2;
  r->max_in = -r->min_in;
  r->idx_factor = SIGMOID_BINS / (r->max_in - r->min_in);
  r->table = (double*)malloc(sizeof(double) * SIGMOID_BINS);
  for (size_t i = 0; i < SIGMOID_BINS; ++i) {
    const double s = i / r->idx_factor + r->min_in;
    r->table[i] = 1.0f / (1.0f + exp(s * sigmoid));
  }
  return r;
}
ORC_API void orc_rank_free(orc_rank* r) {
  if (r) {
    free(r->table);
    free(r);
  }
}

static double get_sigmoid(const orc_rank* r, double score) {
  if (score <= r->min_in) return r->table[0];
  if (score >= r->max_in) return r->table[SIGMOID_BINS - 1];
  return r->table[(size_t)((score - r->min_in) * r->idx_factor)];
}
static double disc(const orc_rank* r, int k) { return r->discount[k < MAX_POSITION ? k : MAX_POSITION - 1]; }

ORC_API double orc_max_dcg_at_k(const orc_rank* r, int k, const float* label, int n) {
  double ret = 0.0f;
  int cnt[64];
  memset(cnt, 0, sizeof(cnt));
  for (int i = 0; i < n; ++i) ++cnt[(int)label[i]];
  int top = r->num_label_gain - 1;
  if (k > n) k = n;
  for (int j = 0; j < k; ++j) {
    while (top > 0 && cnt[top] <= 0) top -= 1;
    if (top < 0) break;
    ret += r->discount[j] * r->label_gain[top];
    cnt[top] -= 1;
  }
  return ret;
}
ORC_API double orc_max_bdcg_at_k(const orc_rank* r, int k, const float* label, int n) {
  double ret = 0.0f;
  int rel = 0;
  for (int i = 0; i < n; ++i) rel += label[i] > 0;
  if (k > n) k = n;
  if (k > rel) k = rel;
  for (int j = 0; j < k; ++j) ret += r->discount[j];
  return ret;
}

/* stable merge sort of indices by score descending == std::stable_sort with (score[a] > score[b]) */
static void stable_sort_desc(int* v, int n, const double* score, int* tmp) {
  for (int w = 1; w < n; w <<= 1) {
    for (int lo = 0; lo < n; lo += 2 * w) {
      int mid = lo + w < n ? lo + w : n, hi = lo + 2 * w < n ? lo + 2 * w : n;
      int a = lo, b = mid, o = lo;
      while (a < mid && b < hi) tmp[o++] = (score[v[b]] > score[v[a]]) ? v[b++] : v[a++];
      while (a < mid) tmp[o++] = v[a++];
      while (b < hi) tmp[o++] = v[b++];
    }
    memcpy(v, tmp, sizeof(int) * n);
  }
}

/* libstdc++ std::nth_element (introselect) with comparator score[a] > score[b]; the heap-select fallback of an
 * exhausted depth limit is replaced by an insertion sort of the remaining range (never reached in the tests). */
static int before(const double* sc, int a, int b) { return sc[a] > sc[b]; }
static void nth_element_desc(int* v, int n, int nth, const double* sc) {
  if (n <= 0 || nth >= n) return;
  int first = 0, last = n, depth = 0;
  for (int m = n; m > 1; m >>= 1) ++depth;
  depth *= 2;
  while (last - first > 3) {
    if (depth == 0) break;
    --depth;
    const int mid = first + (last - first) / 2;
    const int a = first + 1, b = mid, c = last - 1;
    int pick;
    if (before(sc, v[a], v[b])) {
      pick = before(sc, v[b], v[c]) ? b : (before(sc, v[a], v[c]) ? c : a);
    } else {
      pick = before(sc, v[a], v[c]) ? a : (before(sc, v[b], v[c]) ? c : b);
    }
    int t = v[first];
    v[first] = v[pick];
    v[pick] = t;
    int lo = first + 1, hi = last;
    const int pivot = v[first];
    for (;;) {
      while (before(sc, v[lo], pivot)) ++lo;
      --hi;
      while (before(sc, pivot, v[hi])) --hi;
      if (!(lo < hi)) break;
      t = v[lo];
      v[lo] = v[hi];
      v[hi] = t;
      ++lo;
    }
    if (lo <= nth) {
      first = lo;
    } else {
      last = lo;
    }
  }
  for (int i = first + 1; i < last; ++i) {
    const int val = v[i];
    if (before(sc, val, v[first])) {
      for (int m = i; m > first; --m) v[m] = v[m - 1];
      v[first] = val;
    } else {
      int m = i;
      while (before(sc, val, v[m - 1])) {
        v[m] = v[m - 1];
        --m;
      }
      v[m] = val;
    }
  }
}

static int is_binary_target(int t) {
  return t == T_PRECISION || t == T_BNDCG || t == T_LL_BNDCG || t == T_LL_BNDCG_PP || t == T_ARPK || t == T_BIN_RANKNET ||
         t == T_GAP_S || t == T_GAP_X || t == T_GAP_S_P || t == T_GAP_X_P || t == T_GAP_S_PP || t == T_GAP_X_PP;
}

static void one_query(const orc_rank* r, int cnt, const float* label, const double* score, double inv_dcg, double inv_bdcg,
                      float* lambdas, float* hessians) {
  const int t = r->target, k = r->truncation_level;
  for (int i = 0; i < cnt; ++i) lambdas[i] = hessians[i] = 0.0f;
  int* sorted = (int*)malloc(sizeof(int) * cnt * 2);
  for (int i = 0; i < cnt; ++i) sorted[i] = i;
  double best_score, worst_score;
  if (t == T_PRECISION || t == T_BIN_RANKNET || t == T_RANKNET || t == T_LL_ARP1 || t == T_LL_ARP2) {
    if (t == T_PRECISION) nth_element_desc(sorted, cnt, k - 1 < cnt ? k - 1 : cnt - 1, score);
    best_score = worst_score = score[0];
    for (int i = 1; i < cnt; ++i) {
      if (score[i] > best_score) best_score = score[i];
      if (score[i] < worst_score) worst_score = score[i];
    }
  } else {
    stable_sort_desc(sorted, cnt, score, sorted + cnt);
    best_score = score[sorted[0]];
    int wi = cnt - 1;
    if (wi > 0 && score[sorted[wi]] == -INFINITY) wi -= 1;
    worst_score = score[sorted[wi]];
  }
  double sum_lambdas = 0.0;
  const int i_end = t <= T_PRECISION ? (cnt - 1 < k ? cnt - 1 : k) : cnt - 1;
  for (int i = 0; i < i_end; ++i) {
    if (score[sorted[i]] == -INFINITY) continue;
    int start, end;
    switch (t) {
      case T_PRECISION: start = k; end = cnt; break;
      case T_ARPK: case T_GAP_S_P: case T_GAP_X_P: case T_GAP_S_PP: case T_GAP_X_PP:
        start = i + 1 > k ? i + 1 : k; end = cnt; break;
      case T_GAP_S: start = i + k; end = start + 1 < cnt ? start + 1 : cnt; break;
      case T_GAP_X: start = i + k; end = cnt; break;
      default: start = i + 1; end = cnt;
    }
    for (int j = start; j < end; ++j) {
      if (score[sorted[j]] == -INFINITY) continue;
      const float li = label[sorted[i]], lj = label[sorted[j]];
      if (li == lj) continue;
      if (li > 0 && lj > 0 && is_binary_target(t)) continue;
      int high_rank, low_rank;
      if (li > lj) { high_rank = i; low_rank = j; } else { high_rank = j; low_rank = i; }
      const int high = sorted[high_rank], low = sorted[low_rank];
      const double delta_score = score[high] - score[low];
      double delta_pair;
      const int d = j - i;
      const double hg = r->label_gain[(int)label[high]], lg = r->label_gain[(int)label[low]];
      switch (t) {
        case T_NDCG: delta_pair = (hg - lg) * fabs(disc(r, high_rank) - disc(r, low_rank)) * inv_dcg; break;
        case T_LL_NDCG: delta_pair = (hg - lg) * (disc(r, d) - disc(r, d + 1)) * inv_dcg; break;
        case T_LL_NDCG_PP:
          delta_pair = (hg - lg) * (fabs(disc(r, high_rank) - disc(r, low_rank)) + r->lambdagap_weight * (disc(r, d) - disc(r, d + 1))) * inv_dcg;
          break;
        case T_BNDCG: delta_pair = fabs(disc(r, high_rank) - disc(r, low_rank)) * inv_bdcg; break;
        case T_LL_BNDCG: delta_pair = (disc(r, d) - disc(r, d + 1)) * inv_bdcg; break;
        case T_LL_BNDCG_PP:
          delta_pair = (fabs(disc(r, high_rank) - disc(r, low_rank)) + r->lambdagap_weight * (disc(r, d) - disc(r, d + 1))) * inv_bdcg;
          break;
        case T_PRECISION: case T_GAP_S: case T_GAP_X: case T_BIN_RANKNET: case T_RANKNET: delta_pair = 1; break;
        case T_GAP_S_P: delta_pair = (d == k) * r->lambdagap_weight + (i < k); break;
        case T_GAP_X_P: delta_pair = (d >= k) * r->lambdagap_weight + (i < k); break;
        case T_GAP_S_PP: delta_pair = (d == k) * r->lambdagap_weight + (j + 1 - k) - (i >= k) * (i + 1 - k); break;
        case T_GAP_X_PP: delta_pair = (d >= k) * r->lambdagap_weight + (j + 1 - k) - (i >= k) * (i + 1 - k); break;
        case T_ARPK: delta_pair = (j + 1 - k) - (i >= k) * (i + 1 - k); break;
        case T_LL_ARP1: delta_pair = (double)label[high]; break;
        default: delta_pair = (double)label[high] - (double)label[low]; break;
      }
      if (delta_pair == 0) continue;
      if (r->norm && best_score != worst_score) delta_pair /= (0.01f + fabs(delta_score));
      double p_lambda = get_sigmoid(r, delta_score);
      double p_hessian = p_lambda * (1.0f - p_lambda);
      p_lambda *= -r->sigmoid * delta_pair;
      p_hessian *= r->sigmoid * r->sigmoid * delta_pair;
      lambdas[low] -= (float)p_lambda;
      hessians[low] += (float)p_hessian;
      lambdas[high] += (float)p_lambda;
      hessians[high] += (float)p_hessian;
      sum_lambdas -= 2 * p_lambda;
    }
  }
  if (r->norm && sum_lambdas > 0) {
    const double nf = log2(1 + sum_lambdas) / sum_lambdas;
    for (int i = 0; i < cnt; ++i) {
      lambdas[i] = (float)(lambdas[i] * nf);
      hessians[i] = (float)(hessians[i] * nf);
    }
  }
  free(sorted);
}

ORC_API void orc_lambdarank_gradients(const orc_rank* r, const double* score, const float* label, const float* weight,
                                      const int32_t* query_boundaries, int num_queries, float* g, float* h) {
  for (int q = 0; q < num_queries; ++q) {
    const int start = query_boundaries[q], cnt = query_boundaries[q + 1] - start;
    double inv_dcg = orc_max_dcg_at_k(r, r->truncation_level, label + start, cnt);
    if (inv_dcg > 0.0) inv_dcg = 1.0f / inv_dcg;
    double inv_bdcg = orc_max_bdcg_at_k(r, r->truncation_level, label + start, cnt);
    if (inv_bdcg > 0.0) inv_bdcg = 1.0f / inv_bdcg;
    one_query(r, cnt, label + start, score + start, inv_dcg, inv_bdcg, g + start, h + start);
    if (weight) {
      for (int j = 0; j < cnt; ++j) {
        g[start + j] = (float)(g[start + j] * weight[start + j]);
        h[start + j] = (float)(h[start + j] * weight[start + j]);
      }
    }
  }
}
