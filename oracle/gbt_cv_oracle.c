/*
 * oracle/gbt_cv_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU restatement of the arithmetic behind gentun's fitness evaluation:
 *   XgboostModel.cross_validate  (/root/reference/gentun/models/xgboost_models.py:28-37)
 * which calls xgb.DMatrix(...) (:32) and xgb.cv(params, d_train, num_boost_round,
 * early_stopping_rounds, nfold) (:33-36) and returns the last row of 'test-<metric>-mean' (:37).
 *
 * The arithmetic itself lives in the third-party dependency `xgboost` (pinned xgboost==0.72.1 in
 * /root/reference/requirements.txt:26), which is NOT vendored under /root/reference and is not
 * installed on this box.  This file restates the published algorithm of xgboost's modern
 * (>= 2.0) `tree_method=hist` CPU path, the path gentun reaches with its parameter dict
 * (xgboost_models.py:16-22 sets no tree_method/seed/max_bin/base_score) on any xgboost that runs
 * under Python 3.12.  Upstream file names are quoted per function below.
 *
 * PARITY UNPINNED: the reference's own tests hold no golden vector for this path
 * (tests/test_xgboost_model.py:23-24 prints one number, asserts nothing) and xgboost cannot be run
 * here, so this oracle is anchored on the reference call sites only.  The fold construction is the
 * one piece pinned bit-exactly (against numpy's frozen legacy MT19937 stream).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this file.  The product (gentun_b200/) never does.
 *
 * Sampling genes (subsample, colsample_bytree, colsample_bylevel): upstream draws from a
 * process-global std::mt19937 whose consumption order depends on thread count; that is not
 * restatable.  This oracle (and the engine) use the counter-based generator defined in
 * orc_mix64()/orc_u01() below; parity against real xgboost for sampled individuals is
 * statistical only.
 */
#include <float.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define ORC_MAX_BIN 256
#define ORC_KFACTOR 8
#define ORC_RT_EPS 1e-6f

/* gene order follows the genome dict of XgboostIndividual
 * (/root/reference/gentun/individuals.py:162-175) */
enum { G_ETA = 0, G_MCW, G_DEPTH, G_GAMMA, G_MDS, G_SUBSAMPLE, G_COLTREE, G_COLLEVEL, G_LAMBDA, G_ALPHA, G_SPW, G_COUNT };

enum { ORC_OBJ_SQUARED = 0, ORC_OBJ_LOGISTIC = 1 };
enum { ORC_METRIC_RMSE = 0, ORC_METRIC_LOGLOSS = 1 };

typedef struct {
  int32_t objective;             /* reg:linear / reg:squarederror = 0, binary:logistic = 1 */
  int32_t metric;                /* rmse = 0, logloss = 1 */
  int32_t nfold;                 /* individuals.py:159 default 5 */
  int32_t num_boost_round;       /* individuals.py:160 default 5000 */
  int32_t early_stopping_rounds; /* individuals.py:160 default 100 */
  int32_t nthread;               /* 0 = all */
} orc_opts;

/* ------------------------------------------------------------------------------------------ */
/* X1: folds.  xgboost python-package/xgboost/training.py: mknfold():
 *   np.random.seed(seed=0); idx = np.random.permutation(N); out_idset = np.array_split(idx, nfold)
 * numpy's legacy RandomState: mt19937_seed + Fisher-Yates with random_interval (masked rejection). */
typedef struct { uint32_t mt[624]; int pos; } orc_mt;

static void orc_mt_seed(orc_mt* s, uint32_t seed) {
  for (int i = 0; i < 624; ++i) {
    s->mt[i] = seed;
    seed = 1812433253u * (seed ^ (seed >> 30)) + (uint32_t)i + 1u;
  }
  s->pos = 624;
}
static uint32_t orc_mt_next(orc_mt* s) {
  if (s->pos == 624) {
    uint32_t* m = s->mt;
    int k;
    for (k = 0; k < 624 - 397; ++k) {
      uint32_t y = (m[k] & 0x80000000u) | (m[k + 1] & 0x7fffffffu);
      m[k] = m[k + 397] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
    }
    for (; k < 623; ++k) {
      uint32_t y = (m[k] & 0x80000000u) | (m[k + 1] & 0x7fffffffu);
      m[k] = m[k + (397 - 624)] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
    }
    uint32_t y = (m[623] & 0x80000000u) | (m[0] & 0x7fffffffu);
    m[623] = m[396] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
    s->pos = 0;
  }
  uint32_t y = s->mt[s->pos++];
  y ^= y >> 11;
  y ^= (y << 7) & 0x9d2c5680u;
  y ^= (y << 15) & 0xefc60000u;
  y ^= y >> 18;
  return y;
}
static uint64_t orc_interval(orc_mt* s, uint64_t max) {
  if (max == 0) return 0;
  uint64_t mask = max, v;
  mask |= mask >> 1; mask |= mask >> 2; mask |= mask >> 4;
  mask |= mask >> 8; mask |= mask >> 16; mask |= mask >> 32;
  if (max <= 0xffffffffull) {
    while ((v = (orc_mt_next(s) & mask)) > max) {}
  } else {
    for (;;) {
      uint64_t hi = orc_mt_next(s), lo = orc_mt_next(s);
      v = ((hi << 32) | lo) & mask;
      if (v <= max) break;
    }
  }
  return v;
}

/* perm: n int64 (the permutation); fold_begin: nfold+1 offsets into perm (np.array_split sizes) */
void orc_make_folds(int64_t n, int32_t nfold, uint32_t seed, int64_t* perm, int64_t* fold_begin) {
  orc_mt s;
  orc_mt_seed(&s, seed);
  for (int64_t i = 0; i < n; ++i) perm[i] = i;
  for (int64_t i = n - 1; i >= 1; --i) {
    int64_t j = (int64_t)orc_interval(&s, (uint64_t)i);
    int64_t t = perm[j]; perm[j] = perm[i]; perm[i] = t;
  }
  int64_t base = n / nfold, rem = n % nfold, off = 0;
  for (int k = 0; k < nfold; ++k) { fold_begin[k] = off; off += base + (k < rem ? 1 : 0); }
  fold_begin[nfold] = off;
}

/* ------------------------------------------------------------------------------------------ */
/* X2: weighted quantile sketch.  xgboost src/common/quantile.h (WQSummary, QuantileSketchTemplate)
 * and src/common/quantile.cc (HostSketchContainer ctor, AllReduce, MakeCuts), hist_util.h AddCutPoint.
 * DType = RType = float, as upstream (WQuantileSketch<bst_float, bst_float>). */
typedef struct { float rmin, rmax, wmin, value; } orc_entry;
typedef struct { orc_entry* data; size_t size; } orc_summary;
typedef struct { float value, weight; } orc_qentry;

typedef struct {
  orc_qentry* queue; size_t qcap, qtail;
  size_t nlevel, limit_size;
  orc_summary* level; size_t nlevel_alloc; orc_entry* level_space;
  orc_summary temp;
} orc_sketch;

static float ent_rmin_next(const orc_entry* e) { return e->rmin + e->wmin; }
static float ent_rmax_prev(const orc_entry* e) { return e->rmax - e->wmin; }

static void sum_copy(orc_summary* dst, const orc_summary* src) {
  if (src->size) memcpy(dst->data, src->data, src->size * sizeof(orc_entry));
  dst->size = src->size;
}

/* WQSummary::SetPrune */
static void sum_set_prune(orc_summary* dst, const orc_summary* src, size_t maxsize) {
  if (src->size <= maxsize) { sum_copy(dst, src); return; }
  const float begin = src->data[0].rmax;
  const float range = src->data[src->size - 1].rmin - src->data[0].rmax;
  const size_t n = maxsize - 1;
  dst->data[0] = src->data[0];
  dst->size = 1;
  size_t i = 1, lastidx = 0;
  for (size_t k = 1; k < n; ++k) {
    float dx2 = 2 * (((float)k * range) / (float)n + begin);
    while (i < src->size - 1 && dx2 >= src->data[i + 1].rmax + src->data[i + 1].rmin) ++i;
    if (i == src->size - 1) break;
    if (dx2 < ent_rmin_next(&src->data[i]) + ent_rmax_prev(&src->data[i + 1])) {
      if (i != lastidx) { dst->data[dst->size++] = src->data[i]; lastidx = i; }
    } else {
      if (i + 1 != lastidx) { dst->data[dst->size++] = src->data[i + 1]; lastidx = i + 1; }
    }
  }
  if (lastidx != src->size - 1) dst->data[dst->size++] = src->data[src->size - 1];
}

/* WQSummary::FixError */
static void sum_fix_error(orc_summary* s) {
  float prev_rmin = 0, prev_rmax = 0;
  for (size_t i = 0; i < s->size; ++i) {
    orc_entry* e = &s->data[i];
    if (e->rmin < prev_rmin) e->rmin = prev_rmin; else prev_rmin = e->rmin;
    if (e->rmax < prev_rmax) e->rmax = prev_rmax;
    float rn = ent_rmin_next(e);
    if (e->rmax < rn) e->rmax = rn;
    prev_rmax = e->rmax;
  }
}

/* WQSummary::SetCombine */
static void sum_set_combine(orc_summary* dst, const orc_summary* sa, const orc_summary* sb) {
  if (sa->size == 0) { sum_copy(dst, sb); return; }
  if (sb->size == 0) { sum_copy(dst, sa); return; }
  const orc_entry *a = sa->data, *a_end = sa->data + sa->size;
  const orc_entry *b = sb->data, *b_end = sb->data + sb->size;
  float aprev_rmin = 0, bprev_rmin = 0;
  orc_entry* d = dst->data;
  while (a != a_end && b != b_end) {
    if (a->value == b->value) {
      d->rmin = a->rmin + b->rmin; d->rmax = a->rmax + b->rmax; d->wmin = a->wmin + b->wmin; d->value = a->value;
      aprev_rmin = ent_rmin_next(a); bprev_rmin = ent_rmin_next(b);
      ++d; ++a; ++b;
    } else if (a->value < b->value) {
      d->rmin = a->rmin + bprev_rmin; d->rmax = a->rmax + ent_rmax_prev(b); d->wmin = a->wmin; d->value = a->value;
      aprev_rmin = ent_rmin_next(a);
      ++d; ++a;
    } else {
      d->rmin = b->rmin + aprev_rmin; d->rmax = b->rmax + ent_rmax_prev(a); d->wmin = b->wmin; d->value = b->value;
      bprev_rmin = ent_rmin_next(b);
      ++d; ++b;
    }
  }
  if (a != a_end) {
    float brmax = (b_end - 1)->rmax;
    do { d->rmin = a->rmin + bprev_rmin; d->rmax = a->rmax + brmax; d->wmin = a->wmin; d->value = a->value; ++d; ++a; } while (a != a_end);
  }
  if (b != b_end) {
    float armax = (a_end - 1)->rmax;
    do { d->rmin = b->rmin + aprev_rmin; d->rmax = b->rmax + armax; d->wmin = b->wmin; d->value = b->value; ++d; ++b; } while (b != b_end);
  }
  dst->size = (size_t)(d - dst->data);
  sum_fix_error(dst);
}

static int qentry_cmp(const void* pa, const void* pb) {
  float a = ((const orc_qentry*)pa)->value, b = ((const orc_qentry*)pb)->value;
  return (a > b) - (a < b);
}

/* QuantileSketchTemplate::Queue::MakeSummary */
static void sk_make_summary(orc_sketch* s, orc_summary* out) {
  qsort(s->queue, s->qtail, sizeof(orc_qentry), qentry_cmp);
  out->size = 0;
  float wsum = 0;
  for (size_t i = 0; i < s->qtail;) {
    size_t j = i + 1;
    float w = s->queue[i].weight;
    while (j < s->qtail && s->queue[j].value == s->queue[i].value) { w += s->queue[j].weight; ++j; }
    orc_entry e = { wsum, wsum + w, w, s->queue[i].value };
    out->data[out->size++] = e;
    wsum += w;
    i = j;
  }
}

/* QuantileSketchTemplate::LimitSizeLevel */
static void sk_limit_size_level(size_t maxn, double eps, size_t* nlevel, size_t* limit_size) {
  *nlevel = 1;
  for (;;) {
    *limit_size = (size_t)ceil((double)*nlevel / eps) + 1;
    if (maxn < *limit_size) *limit_size = maxn;
    size_t n = (size_t)1 << *nlevel;
    if (n * *limit_size >= maxn) break;
    ++*nlevel;
  }
}

static void sk_init(orc_sketch* s, size_t maxn, double eps) {
  memset(s, 0, sizeof(*s));
  sk_limit_size_level(maxn, eps, &s->nlevel, &s->limit_size);
  s->qcap = s->limit_size * 2;           /* HostSketchContainer ctor: inqueue.queue.resize(limit_size * 2) */
  s->queue = (orc_qentry*)malloc(sizeof(orc_qentry) * (s->qcap ? s->qcap : 1));
  s->nlevel_alloc = s->nlevel + 3;
  s->level_space = (orc_entry*)malloc(sizeof(orc_entry) * s->limit_size * s->nlevel_alloc + 16);
  s->level = (orc_summary*)calloc(s->nlevel_alloc, sizeof(orc_summary));
  for (size_t l = 0; l < s->nlevel_alloc; ++l) { s->level[l].data = s->level_space + l * s->limit_size; s->level[l].size = 0; }
  s->temp.data = (orc_entry*)malloc(sizeof(orc_entry) * (s->limit_size * 2 + 1));
  s->temp.size = 0;
}
static void sk_free(orc_sketch* s) { free(s->queue); free(s->level_space); free(s->level); free(s->temp.data); }

/* QuantileSketchTemplate::PushTemp */
static int sk_used_levels = 0; /* unused, kept for debugging */
static void sk_push_temp(orc_sketch* s, size_t* nlevels_in_use) {
  for (size_t l = 1;; ++l) {
    if (l + 1 > *nlevels_in_use) *nlevels_in_use = l + 1;  /* InitLevel(l + 1) */
    if (l >= s->nlevel_alloc) { fprintf(stderr, "oracle sketch: level overflow\n"); abort(); }
    if (s->level[l].size == 0) {
      sum_set_prune(&s->level[l], &s->temp, s->limit_size);
      break;
    } else {
      sum_set_prune(&s->level[0], &s->temp, s->limit_size);
      sum_set_combine(&s->temp, &s->level[0], &s->level[l]);
      if (s->temp.size > s->limit_size) {
        s->level[l].size = 0;
      } else {
        sum_copy(&s->level[l], &s->temp);
        break;
      }
    }
  }
}

/* Sketch one column stream (values in DMatrix-slice row order, unit weights) and emit cut points.
 * Returns number of cuts (<= max_bin) written to cuts[]; *min_val as in MakeCuts. */
static int sk_column_cuts(const float* vals, size_t n, int max_bin, float* cuts, float* min_val) {
  (void)sk_used_levels;
  if (n == 0) { *min_val = 1e-5f; cuts[0] = 1e-5f + (fabsf(1e-5f) + 1e-5f); return 1; }
  size_t n_bins = n < (size_t)max_bin ? n : (size_t)max_bin;
  if (n_bins < 1) n_bins = 1;
  double eps = 1.0 / (double)((float)n_bins * (float)ORC_KFACTOR);
  orc_sketch s;
  sk_init(&s, n, eps);
  size_t levels_in_use = 0;  /* level.size() upstream */
  /* QuantileSketchTemplate::Push */
  for (size_t r = 0; r < n; ++r) {
    float x = vals[r];
    if (s.qtail == s.qcap && s.queue[s.qtail - 1].value != x) {
      sk_make_summary(&s, &s.temp);
      s.qtail = 0;
      sk_push_temp(&s, &levels_in_use);
    }
    if (s.qtail == 0 || s.queue[s.qtail - 1].value != x) {
      s.queue[s.qtail].value = x; s.queue[s.qtail].weight = 1.0f; ++s.qtail;
    } else {
      s.queue[s.qtail - 1].weight += 1.0f;
    }
  }
  /* QuantileSketchTemplate::GetSummary */
  orc_summary out;
  out.data = (orc_entry*)malloc(sizeof(orc_entry) * (s.limit_size * 2 + 1));
  out.size = 0;
  sk_make_summary(&s, &out);
  if (levels_in_use != 0) {
    sum_set_prune(&s.level[0], &out, s.limit_size);
    for (size_t l = 1; l < levels_in_use; ++l) {
      if (s.level[l].size == 0) continue;
      if (s.level[0].size == 0) {
        sum_copy(&s.level[0], &s.level[l]);
      } else {
        sum_set_combine(&out, &s.level[0], &s.level[l]);
        sum_set_prune(&s.level[0], &out, s.limit_size);
      }
    }
    sum_copy(&out, &s.level[0]);
  } else if (out.size > s.limit_size) {
    sum_set_prune(&s.temp, &out, s.limit_size);
    sum_copy(&out, &s.temp);
  }
  /* SketchContainerImpl::AllReduce (single worker): reduced.SetPrune(out, min(n, max_bin*kFactor)) */
  size_t inter = n < (size_t)max_bin * ORC_KFACTOR ? n : (size_t)max_bin * ORC_KFACTOR;
  orc_summary reduced;
  reduced.data = (orc_entry*)malloc(sizeof(orc_entry) * (out.size + 1));
  reduced.size = 0;
  sum_set_prune(&reduced, &out, inter);
  /* MakeCuts: a.SetPrune(reduced, max_num_bins + 1) */
  size_t max_num_bins = inter < (size_t)max_bin ? inter : (size_t)max_bin;
  orc_summary a;
  a.data = (orc_entry*)malloc(sizeof(orc_entry) * (reduced.size + 1));
  a.size = 0;
  sum_set_prune(&a, &reduced, max_num_bins + 1);
  float mval = a.data[0].value;
  *min_val = mval - fabsf(mval) - 1e-5f;
  /* AddCutPoint */
  size_t required = a.size < max_num_bins ? a.size : max_num_bins;
  int nc = 0;
  for (size_t i = 1; i < required; ++i) {
    float cpt = a.data[i].value;
    if (i == 1 || cpt > cuts[nc - 1]) cuts[nc++] = cpt;
  }
  float cpt = a.size > 0 ? a.data[a.size - 1].value : *min_val;
  float last = cpt + (fabsf(cpt) + 1e-5f);
  cuts[nc++] = last;
  free(a.data); free(reduced.data); free(out.data);
  sk_free(&s);
  return nc;
}

/* Public: cuts for the rows `rows[0..nrows)` (in that order) of X (row-major, ldx floats per row).
 * cuts is f*ORC_MAX_BIN floats, ncuts is f ints, min_vals f floats. */
void orc_sketch_cuts(const float* X, int64_t ldx, int32_t f, const int64_t* rows, int64_t nrows,
                     float* cuts, int32_t* ncuts, float* min_vals) {
#pragma omp parallel for schedule(dynamic, 1)
  for (int j = 0; j < f; ++j) {
    float* col = (float*)malloc(sizeof(float) * (size_t)(nrows > 0 ? nrows : 1));
    for (int64_t r = 0; r < nrows; ++r) col[r] = X[rows[r] * ldx + j];
    ncuts[j] = sk_column_cuts(col, (size_t)nrows, ORC_MAX_BIN, cuts + (size_t)j * ORC_MAX_BIN, &min_vals[j]);
    free(col);
  }
}

/* X3: xgboost src/common/hist_util.h HistogramCuts::SearchBin: std::upper_bound, clamp to last. */
static inline uint8_t orc_search_bin(const float* cuts, int nc, float x) {
  int lo = 0, hi = nc;
  while (lo < hi) { int mid = (lo + hi) >> 1; if (cuts[mid] > x) hi = mid; else lo = mid + 1; }
  if (lo == nc) lo = nc - 1;
  return (uint8_t)lo;
}

void orc_bin_rows(const float* X, int64_t ldx, int32_t f, const int64_t* rows, int64_t nrows,
                  const float* cuts, const int32_t* ncuts, uint8_t* bins_rowmajor /* nrows*f */) {
#pragma omp parallel for schedule(static)
  for (int64_t r = 0; r < nrows; ++r)
    for (int j = 0; j < f; ++j)
      bins_rowmajor[r * f + j] = orc_search_bin(cuts + (size_t)j * ORC_MAX_BIN, ncuts[j], X[rows[r] * ldx + j]);
}

/* ------------------------------------------------------------------------------------------ */
/* Counter-based generator shared (by definition, not by code) with the engine. */
static inline uint64_t orc_mix64(uint64_t z) {
  z += 0x9e3779b97f4a7c15ull;
  z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
  z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
  return z ^ (z >> 31);
}
/* tag: 1 = row subsample, 2 = colsample_bytree, 3 = colsample_bylevel */
static inline uint64_t orc_stream_key(uint32_t tag, uint32_t fold, uint32_t round, uint32_t level) {
  return orc_mix64(((uint64_t)tag << 60) ^ ((uint64_t)fold << 48) ^ ((uint64_t)level << 40) ^ (uint64_t)round);
}
static inline uint64_t orc_draw(uint64_t key, uint64_t index) { return orc_mix64(key ^ (index * 0xd1342543de82ef95ull)); }
static inline float orc_u01(uint64_t h) { return (float)(h >> 40) * (1.0f / 16777216.0f); }

/* choose the n candidates with the smallest (draw, index); result ascending by feature index */
static void orc_sample_features(uint64_t key, const uint8_t* in_mask, int f, int n_keep, uint8_t* out_mask) {
  uint64_t* k = (uint64_t*)malloc(sizeof(uint64_t) * (size_t)f);
  for (int j = 0; j < f; ++j) k[j] = orc_draw(key, (uint64_t)j);
  for (int j = 0; j < f; ++j) {
    out_mask[j] = 0;
    if (!in_mask[j]) continue;
    int rank = 0;
    for (int i = 0; i < f; ++i)
      if (in_mask[i] && (k[i] < k[j] || (k[i] == k[j] && i < j))) ++rank;
    out_mask[j] = rank < n_keep;
  }
  free(k);
}

/* ------------------------------------------------------------------------------------------ */
/* split arithmetic: xgboost src/tree/param.h (ThresholdL1, CalcWeight, CalcGainGivenWeight) and
 * src/tree/split_evaluator.h TreeEvaluator::SplitEvaluator (no constraints). */
typedef struct {
  float eta, min_child_weight, gamma, max_delta_step, subsample, colsample_bytree, colsample_bylevel, lambda, alpha, spw;
  int max_depth;
} orc_params;

static inline double thr_l1(double w, float alpha) {
  if (w > +alpha) return w - alpha;
  if (w < -alpha) return w + alpha;
  return 0.0;
}
static inline float calc_weight(const orc_params* p, double G, double H) {
  if (H < p->min_child_weight || H <= 0.0) return 0.0f;
  double dw = -thr_l1(G, p->alpha) / (H + p->lambda);
  if (p->max_delta_step != 0.0f && fabs(dw) > p->max_delta_step) dw = copysign((double)p->max_delta_step, dw);
  return (float)dw;
}
static inline float calc_gain_given_weight(const orc_params* p, double G, double H, float w) {
  if (H <= 0) return 0.0f;
  if (p->max_delta_step == 0.0f) {
    double t = thr_l1(G, p->alpha);
    return (float)((t * t) / (H + p->lambda));
  }
  float g = (float)G, h = (float)H;
  return -(2.0f * g * w + (h + p->lambda) * (w * w));
}
static inline float calc_gain(const orc_params* p, double G, double H) { return calc_gain_given_weight(p, G, H, calc_weight(p, G, H)); }
static inline float calc_split_gain(const orc_params* p, double GL, double HL, double GR, double HR) {
  float wl = calc_weight(p, GL, HL), wr = calc_weight(p, GR, HR);
  return calc_gain_given_weight(p, GL, HL, wl) + calc_gain_given_weight(p, GR, HR, wr);
}

/* xgboost src/common/math.h Sigmoid */
static inline float orc_sigmoid(float x) {
  x = fminf(-x, 88.7f);
  float denom = expf(x) + 1.0f + 1e-16f;
  return 1.0f / denom;
}

/* ------------------------------------------------------------------------------------------ */
typedef struct {
  int64_t n; int32_t f; int64_t ldx; const float* X; const float* y;
  int32_t nfold;
  int64_t* perm; int64_t* fold_begin;
  /* per fold */
  int64_t* ntrain; int64_t* ntest;
  int64_t** train_rows; int64_t** test_rows;   /* original row ids, DMatrix.slice order */
  float** cuts; int32_t** ncuts; float** min_vals;
  uint8_t** bins_col;                          /* [f][ntrain] column-major train bins (partition) */
  uint8_t** bins_row;                          /* [ntrain][f] row-major train bins (histograms: GHistIndexMatrix is row-major) */
} orc_ctx;

void orc_ctx_destroy(orc_ctx* c) {
  if (!c) return;
  for (int k = 0; k < c->nfold; ++k) {
    if (c->train_rows) free(c->train_rows[k]);
    if (c->test_rows) free(c->test_rows[k]);
    if (c->cuts) free(c->cuts[k]);
    if (c->ncuts) free(c->ncuts[k]);
    if (c->min_vals) free(c->min_vals[k]);
    if (c->bins_col) free(c->bins_col[k]);
    if (c->bins_row) free(c->bins_row[k]);
  }
  free(c->perm); free(c->fold_begin); free(c->ntrain); free(c->ntest);
  free(c->train_rows); free(c->test_rows); free(c->cuts); free(c->ncuts); free(c->min_vals); free(c->bins_col); free(c->bins_row);
  free(c);
}

/* Everything xgb.cv derives from (data, nfold) alone: mknfold slices, per-slice cuts, GHistIndexMatrix. */
orc_ctx* orc_ctx_create(const float* X, int64_t n, int32_t f, int64_t ldx, const float* y, int32_t nfold, int32_t nthread) {
#ifdef _OPENMP
  if (nthread > 0) omp_set_num_threads(nthread);
#else
  (void)nthread;
#endif
  orc_ctx* c = (orc_ctx*)calloc(1, sizeof(orc_ctx));
  c->n = n; c->f = f; c->ldx = ldx; c->X = X; c->y = y; c->nfold = nfold;
  c->perm = (int64_t*)malloc(sizeof(int64_t) * (size_t)n);
  c->fold_begin = (int64_t*)malloc(sizeof(int64_t) * (size_t)(nfold + 1));
  orc_make_folds(n, nfold, 0u, c->perm, c->fold_begin);
  c->ntrain = (int64_t*)calloc((size_t)nfold, sizeof(int64_t));
  c->ntest = (int64_t*)calloc((size_t)nfold, sizeof(int64_t));
  c->train_rows = (int64_t**)calloc((size_t)nfold, sizeof(void*));
  c->test_rows = (int64_t**)calloc((size_t)nfold, sizeof(void*));
  c->cuts = (float**)calloc((size_t)nfold, sizeof(void*));
  c->ncuts = (int32_t**)calloc((size_t)nfold, sizeof(void*));
  c->min_vals = (float**)calloc((size_t)nfold, sizeof(void*));
  c->bins_col = (uint8_t**)calloc((size_t)nfold, sizeof(void*));
  c->bins_row = (uint8_t**)calloc((size_t)nfold, sizeof(void*));
  for (int k = 0; k < nfold; ++k) {
    int64_t nt = c->fold_begin[k + 1] - c->fold_begin[k];
    int64_t ntr = n - nt;
    c->ntest[k] = nt; c->ntrain[k] = ntr;
    c->test_rows[k] = (int64_t*)malloc(sizeof(int64_t) * (size_t)(nt > 0 ? nt : 1));
    c->train_rows[k] = (int64_t*)malloc(sizeof(int64_t) * (size_t)(ntr > 0 ? ntr : 1));
    memcpy(c->test_rows[k], c->perm + c->fold_begin[k], sizeof(int64_t) * (size_t)nt);
    int64_t o = 0;  /* np.concatenate of the other chunks, in chunk order */
    for (int i = 0; i < nfold; ++i) {
      if (i == k) continue;
      int64_t m = c->fold_begin[i + 1] - c->fold_begin[i];
      memcpy(c->train_rows[k] + o, c->perm + c->fold_begin[i], sizeof(int64_t) * (size_t)m);
      o += m;
    }
    c->cuts[k] = (float*)calloc((size_t)f * ORC_MAX_BIN, sizeof(float));
    c->ncuts[k] = (int32_t*)calloc((size_t)f, sizeof(int32_t));
    c->min_vals[k] = (float*)calloc((size_t)f, sizeof(float));
    orc_sketch_cuts(X, ldx, f, c->train_rows[k], ntr, c->cuts[k], c->ncuts[k], c->min_vals[k]);
    c->bins_col[k] = (uint8_t*)malloc((size_t)f * (size_t)(ntr > 0 ? ntr : 1));
    const int64_t* tr = c->train_rows[k];
    uint8_t* bc = c->bins_col[k];
#pragma omp parallel for schedule(static)
    for (int64_t r0 = 0; r0 < ntr; r0 += 4096) {   /* row blocks: every row of X is touched once */
      const int64_t r1 = r0 + 4096 < ntr ? r0 + 4096 : ntr;
      for (int j = 0; j < f; ++j) {
        const float* cj = c->cuts[k] + (size_t)j * ORC_MAX_BIN;
        int nc = c->ncuts[k][j];
        for (int64_t r = r0; r < r1; ++r) bc[(size_t)j * ntr + r] = orc_search_bin(cj, nc, X[tr[r] * ldx + j]);
      }
    }
    c->bins_row[k] = (uint8_t*)malloc((size_t)f * (size_t)(ntr > 0 ? ntr : 1));
    uint8_t* br = c->bins_row[k];
#pragma omp parallel for schedule(static)
    for (int64_t r = 0; r < ntr; ++r)
      for (int j = 0; j < f; ++j) br[(size_t)r * f + j] = bc[(size_t)j * ntr + r];
  }
  return c;
}

/* accessors used by the parity tests */
int64_t orc_ctx_ntrain(const orc_ctx* c, int k) { return c->ntrain[k]; }
int64_t orc_ctx_ntest(const orc_ctx* c, int k) { return c->ntest[k]; }
const int64_t* orc_ctx_train_rows(const orc_ctx* c, int k) { return c->train_rows[k]; }
const int64_t* orc_ctx_test_rows(const orc_ctx* c, int k) { return c->test_rows[k]; }
const float* orc_ctx_cuts(const orc_ctx* c, int k) { return c->cuts[k]; }
const int32_t* orc_ctx_ncuts(const orc_ctx* c, int k) { return c->ncuts[k]; }
const uint8_t* orc_ctx_bins_col(const orc_ctx* c, int k) { return c->bins_col[k]; }
const int64_t* orc_ctx_perm(const orc_ctx* c) { return c->perm; }

/* ------------------------------------------------------------------------------------------ */
/* Tree dump record, heap indexed (node i -> children 2i+1, 2i+2). feat: -2 absent, -1 leaf, >=0 split.
 * value: leaf value of a leaf; for a split node the loss_chg of the best OTHER candidate (0 if none).
 * gain: loss_chg of the best candidate -- also for a node that was evaluated and stayed a leaf (0 if never evaluated). */
typedef struct { int32_t feat; int32_t bin; float value; float gain; } orc_node_dump;
#define ORC_MAX_NODES 2047

typedef struct {
  double G, H; float root_gain, weight;
  int64_t begin, count;    /* segment in the row-set array */
  int32_t state;           /* 0 absent, 1 open, 2 split, 3 leaf */
  int32_t feat, bin; float loss_chg, second_chg; double GL, HL;
} orc_node;

typedef struct { double g, h; } orc_gh;

/* optional phase timers (ORC_PROFILE=1 prints them at the end of orc_cv) */
static double orc_t_hist, orc_t_eval, orc_t_part, orc_t_sub, orc_t_grad, orc_t_pred;
static double orc_now(void) {
#ifdef _OPENMP
  return omp_get_wtime();
#else
  return 0.0;
#endif
}

/* Histogram accumulation is row-parallel in the way upstream's BuildHist is (common/hist_util.cc: blocks of rows,
 * partial histograms, ordered reduction), but with a FIXED block size so that the result does not depend on the thread
 * count: a node's rows are cut into blocks of ORC_HIST_BLOCK rows, every block is one task that adds its rows in
 * row-set order into a partial histogram, and the partials of a node are summed in block order.  A node
 * of at most ORC_HIST_BLOCK rows is therefore accumulated exactly as a sequential loop would. */
#define ORC_HIST_BLOCK 16384
/* histogram buffers are recycled (a 200 KB calloc per node is an mmap + page faults + munmap every time); only the
 * serial parts of the tree loop take from / return to the pool */
static orc_gh** orc_pool; static size_t orc_pool_n, orc_pool_cap, orc_pool_hsz;
static orc_gh* orc_pool_get(size_t hsz) {
  if (hsz != orc_pool_hsz) { while (orc_pool_n) free(orc_pool[--orc_pool_n]); orc_pool_hsz = hsz; }
  if (orc_pool_n) return orc_pool[--orc_pool_n];
  return (orc_gh*)malloc(hsz * sizeof(orc_gh));
}
static void orc_pool_put(orc_gh* p) {
  if (!p) return;
  if (orc_pool_n == orc_pool_cap) { orc_pool_cap = orc_pool_cap ? 2 * orc_pool_cap : 64; orc_pool = (orc_gh**)realloc(orc_pool, orc_pool_cap * sizeof(void*)); }
  orc_pool[orc_pool_n++] = p;
}
static void orc_pool_release(void) { while (orc_pool_n) free(orc_pool[--orc_pool_n]); free(orc_pool); orc_pool = NULL; orc_pool_cap = 0; }
typedef struct { int32_t node; int64_t begin, count; orc_gh* dst; } orc_hist_task;

static void orc_build_hists(const orc_ctx* c, int k, const float* g, const float* h, const int64_t* cur,
                            const orc_node* nodes, const int* build_ids, int n_build, orc_gh** hist_of_node) {
  const int f = c->f;
  const uint8_t* br = c->bins_row[k];
  const size_t hsz = (size_t)f * ORC_MAX_BIN;
  int64_t n_tasks = 0, n_extra = 0;
  for (int q = 0; q < n_build; ++q) {
    int64_t nb = (nodes[build_ids[q]].count + ORC_HIST_BLOCK - 1) / ORC_HIST_BLOCK; if (nb < 1) nb = 1;
    n_tasks += nb; n_extra += nb - 1;
  }
  orc_hist_task* tasks = (orc_hist_task*)malloc(sizeof(orc_hist_task) * (size_t)(n_tasks > 0 ? n_tasks : 1));
  orc_gh* extra = n_extra ? (orc_gh*)malloc((size_t)n_extra * hsz * sizeof(orc_gh)) : NULL;
  int64_t t = 0, x = 0;
  for (int q = 0; q < n_build; ++q) {
    const orc_node* B = &nodes[build_ids[q]];
    hist_of_node[build_ids[q]] = orc_pool_get(hsz);
    int64_t nb = (B->count + ORC_HIST_BLOCK - 1) / ORC_HIST_BLOCK; if (nb < 1) nb = 1;
    for (int64_t b = 0; b < nb; ++b) {
      int64_t o = b * ORC_HIST_BLOCK, cnt = B->count - o < ORC_HIST_BLOCK ? B->count - o : ORC_HIST_BLOCK;
      tasks[t].node = build_ids[q]; tasks[t].begin = B->begin + o; tasks[t].count = cnt < 0 ? 0 : cnt;
      tasks[t].dst = b == 0 ? hist_of_node[build_ids[q]] : extra + (size_t)(x++) * hsz;
      ++t;
    }
  }
  /* row-wise accumulation (hist_util.cc RowsWiseBuildHistKernel): one pass over a block's rows, all features of a
   * row together; each cell still receives its rows in row-set order */
#pragma omp parallel for schedule(dynamic, 1)
  for (int64_t ti = 0; ti < n_tasks; ++ti) {
    orc_gh* H = tasks[ti].dst;
    memset(H, 0, hsz * sizeof(orc_gh));
    const int64_t* rs = cur + tasks[ti].begin; const int64_t cnt = tasks[ti].count;
    for (int64_t i = 0; i < cnt; ++i) {
      if (i + 8 < cnt) __builtin_prefetch(br + (size_t)rs[i + 8] * f);
      const int64_t r = rs[i];
      const double gi = g[r], hi = h[r];
      const uint8_t* row = br + (size_t)r * f;
      for (int j = 0; j < f; ++j) { orc_gh* cell = H + (size_t)j * ORC_MAX_BIN + row[j]; cell->g += gi; cell->h += hi; }
    }
  }
  if (n_extra) {   /* partials of a node, in block order */
#pragma omp parallel for schedule(static)
    for (int j = 0; j < f; ++j)
      for (int64_t ti = 0; ti < n_tasks; ++ti) {
        orc_gh* H = hist_of_node[tasks[ti].node];
        if (tasks[ti].dst == H) continue;
        const orc_gh* src = tasks[ti].dst + (size_t)j * ORC_MAX_BIN; orc_gh* dj = H + (size_t)j * ORC_MAX_BIN;
        for (int b = 0; b < ORC_MAX_BIN; ++b) { dj[b].g += src[b].g; dj[b].h += src[b].h; }
      }
    free(extra);
  }
  free(tasks);
}

/* one boosting round for one fold: grow a depthwise hist tree (updater_quantile_hist.cc UpdateTree,
 * hist/evaluate_splits.h HistEvaluator, hist/histogram.h subtraction trick, common/partition_builder.h). */
static void orc_grow_tree(const orc_ctx* c, int k, const orc_params* p, int round,
                          const float* g, const float* h, orc_node* nodes,
                          int64_t* rows_a, int64_t* rows_b, orc_gh* hist_pool /* ORC_MAX_NODES? no: per level */,
                          orc_gh** hist_of_node) {
  (void)hist_pool;
  const int f = c->f; const int64_t ntr = c->ntrain[k];
  const uint8_t* bc = c->bins_col[k];
  const int32_t* ncuts = c->ncuts[k];
  const size_t hsz = (size_t)f * ORC_MAX_BIN;
  for (int i = 0; i < ORC_MAX_NODES; ++i) { nodes[i].state = 0; hist_of_node[i] = NULL; }
  /* column sampling for this tree (ColumnSampler::Init) */
  uint8_t* all_mask = (uint8_t*)malloc((size_t)f); memset(all_mask, 1, (size_t)f);
  uint8_t* tree_mask = (uint8_t*)malloc((size_t)f);
  uint8_t* level_mask = (uint8_t*)malloc((size_t)f);
  if (p->colsample_bytree == 1.0f) memcpy(tree_mask, all_mask, (size_t)f);
  else {
    int nk = (int)(p->colsample_bytree * (float)f); if (nk < 1) nk = 1;
    orc_sample_features(orc_stream_key(2, (uint32_t)k, (uint32_t)round, 0), all_mask, f, nk, tree_mask);
  }
  int n_tree = 0; for (int j = 0; j < f; ++j) n_tree += tree_mask[j];

#pragma omp parallel for schedule(static)
  for (int64_t r = 0; r < ntr; ++r) rows_a[r] = r;
  int64_t* cur = rows_a; int64_t* nxt = rows_b;
  int* build_ids = (int*)malloc(sizeof(int) * 1024);
  /* root */
  orc_node* root = &nodes[0];
  root->state = 1; root->begin = 0; root->count = ntr; root->loss_chg = 0.0f; root->second_chg = 0.0f;
  {
    build_ids[0] = 0;
    double t0 = orc_now();
    orc_build_hists(c, k, g, h, cur, nodes, build_ids, 1, hist_of_node);
    orc_t_hist += orc_now() - t0;
    const orc_gh* H0 = hist_of_node[0];
    double G = 0, Hs = 0;   /* InitRoot: dense data -> sum of the first feature's histogram */
    for (int b = 0; b < ncuts[0]; ++b) { G += H0[b].g; Hs += H0[b].h; }
    root->G = G; root->H = Hs;
    root->weight = calc_weight(p, G, Hs);
    root->root_gain = calc_gain(p, G, Hs);
  }
  const int D = p->max_depth;
  for (int depth = 0; depth < D; ++depth) {
    int first = (1 << depth) - 1, last = (1 << (depth + 1)) - 1;
    /* bylevel feature set (ColumnSampler::GetFeatureSet(depth)) */
    if (p->colsample_bylevel == 1.0f) memcpy(level_mask, tree_mask, (size_t)f);
    else {
      int nk = (int)(p->colsample_bylevel * (float)n_tree); if (nk < 1) nk = 1;
      orc_sample_features(orc_stream_key(3, (uint32_t)k, (uint32_t)round, (uint32_t)depth), tree_mask, f, nk, level_mask);
    }
    int any_open = 0;
    double t_ph = orc_now();
    /* EvaluateSplits for all open nodes on this level (nodes are independent of each other) */
#pragma omp parallel for schedule(dynamic, 1) reduction(| : any_open)
    for (int nid = first; nid < last; ++nid) {
      orc_node* nd = &nodes[nid];
      if (nd->state != 1) continue;
      any_open |= 1;
      const orc_gh* Hn = hist_of_node[nid];
      float best_chg = 0.0f; int best_f = -1, best_b = -1; double best_GL = 0, best_HL = 0;
      float second_chg = 0.0f;   /* best loss_chg among the OTHER candidates (test diagnostics only: how close the decision was) */
      for (int j = 0; j < f; ++j) {
        if (!level_mask[j]) continue;
        const orc_gh* hj = Hn + (size_t)j * ORC_MAX_BIN;
        double GL = 0, HL = 0;
        for (int b = 0; b < ncuts[j]; ++b) {   /* EnumerateSplit<+1> */
          GL += hj[b].g; HL += hj[b].h;
          double GR = nd->G - GL, HR = nd->H - HL;
          if (HL >= p->min_child_weight && HR >= p->min_child_weight) {
            float chg = calc_split_gain(p, GL, HL, GR, HR) - nd->root_gain;
            /* SplitEntry::NeedReplace: features ascend, so strictly greater wins; inf is rejected */
            if (!isinf(chg) && chg > best_chg) { second_chg = best_chg; best_chg = chg; best_f = j; best_b = b; best_GL = GL; best_HL = HL; }
            else if (!isinf(chg) && chg > second_chg) second_chg = chg;
          }
        }
      }
      /* CPUExpandEntry::IsValid */
      int valid = best_f >= 0 && best_chg > ORC_RT_EPS;
      if (valid) {
        double HR = nd->H - best_HL;
        if (best_HL == 0 || HR == 0) valid = 0;
        if (best_chg < p->gamma) valid = 0;
      }
      nd->loss_chg = best_chg; nd->second_chg = second_chg;
      if (!valid) { nd->state = 3; continue; }
      nd->state = 2; nd->feat = best_f; nd->bin = best_b; nd->GL = best_GL; nd->HL = best_HL;
    }
    orc_t_eval += orc_now() - t_ph; t_ph = orc_now();
    if (!any_open) break;
    /* ApplyTreeSplit + UpdatePosition (stable partition); rows of finished nodes stay where they are */
    int children_evaluated = (depth + 1 < D);  /* Driver::IsChildValid */
#pragma omp parallel for schedule(static)
    for (int64_t q = 0; q < (ntr + 65535) / 65536; ++q) {
      int64_t o = q * 65536, m = ntr - o < 65536 ? ntr - o : 65536;
      memcpy(nxt + o, cur + o, sizeof(int64_t) * (size_t)m);
    }
    /* partition_builder.h: blocks of rows are counted in parallel, then scattered to their stable positions */
    int64_t n_pt = 0;
    for (int nid = first; nid < last; ++nid)
      if (nodes[nid].state == 2) n_pt += (nodes[nid].count + ORC_HIST_BLOCK - 1) / ORC_HIST_BLOCK;
    orc_hist_task* pt = (orc_hist_task*)malloc(sizeof(orc_hist_task) * (size_t)(n_pt > 0 ? n_pt : 1));
    int64_t* pt_left = (int64_t*)malloc(sizeof(int64_t) * (size_t)(n_pt > 0 ? n_pt : 1));
    n_pt = 0;
    for (int nid = first; nid < last; ++nid) {
      if (nodes[nid].state != 2) continue;
      for (int64_t o = 0; o < nodes[nid].count; o += ORC_HIST_BLOCK) {
        pt[n_pt].node = nid; pt[n_pt].begin = nodes[nid].begin + o;
        pt[n_pt].count = nodes[nid].count - o < ORC_HIST_BLOCK ? nodes[nid].count - o : ORC_HIST_BLOCK; pt[n_pt].dst = NULL;
        ++n_pt;
      }
    }
#pragma omp parallel for schedule(dynamic, 1)
    for (int64_t ti = 0; ti < n_pt; ++ti) {
      const orc_node* nd = &nodes[pt[ti].node];
      const uint8_t* col = bc + (size_t)nd->feat * ntr;
      int64_t nl = 0;
      for (int64_t i = 0; i < pt[ti].count; ++i) nl += (col[cur[pt[ti].begin + i]] <= nd->bin);
      pt_left[ti] = nl;
    }
    for (int64_t ti = 0; ti < n_pt;) {   /* per node: children, and each block's offset among the node's lefts */
      const int nid = pt[ti].node;
      orc_node* nd = &nodes[nid];
      int64_t tj = ti, nl = 0;
      while (tj < n_pt && pt[tj].node == nid) { int64_t v = pt_left[tj]; pt_left[tj] = nl; nl += v; ++tj; }
      orc_node* L = &nodes[2 * nid + 1]; orc_node* R = &nodes[2 * nid + 2];
      L->G = nd->GL; L->H = nd->HL; R->G = nd->G - nd->GL; R->H = nd->H - nd->HL;
      L->weight = calc_weight(p, L->G, L->H); R->weight = calc_weight(p, R->G, R->H);
      L->root_gain = calc_gain(p, L->G, L->H); R->root_gain = calc_gain(p, R->G, R->H);
      L->state = R->state = children_evaluated ? 1 : 3;
      L->loss_chg = R->loss_chg = 0.0f; L->second_chg = R->second_chg = 0.0f;
      L->begin = nd->begin; L->count = nl; R->begin = nd->begin + nl; R->count = nd->count - nl;
      ti = tj;
    }
#pragma omp parallel for schedule(dynamic, 1)
    for (int64_t ti = 0; ti < n_pt; ++ti) {
      const orc_node* nd = &nodes[pt[ti].node];
      const uint8_t* col = bc + (size_t)nd->feat * ntr;
      const int64_t nl_node = nodes[2 * pt[ti].node + 1].count;
      int64_t il = nd->begin + pt_left[ti];
      int64_t ir = nd->begin + nl_node + ((pt[ti].begin - nd->begin) - pt_left[ti]);
      for (int64_t i = 0; i < pt[ti].count; ++i) {
        int64_t r = cur[pt[ti].begin + i];
        if (col[r] <= nd->bin) nxt[il++] = r; else nxt[ir++] = r;
      }
    }
    free(pt); free(pt_left);
    { int64_t* t = cur; cur = nxt; nxt = t; }
    orc_t_part += orc_now() - t_ph; t_ph = orc_now();
    if (!children_evaluated) break;
    /* BuildHistogram: build the child with the smaller hessian sum, the sibling by subtraction */
    int n_build = 0;
    for (int nid = first; nid < last; ++nid) {
      if (nodes[nid].state != 2) continue;
      int li = 2 * nid + 1, ri = 2 * nid + 2;
      build_ids[n_build++] = nodes[ri].H < nodes[li].H ? ri : li;
    }
    orc_build_hists(c, k, g, h, cur, nodes, build_ids, n_build, hist_of_node);
    orc_t_hist += orc_now() - t_ph; t_ph = orc_now();
    for (int q = 0; q < n_build; ++q) {
      int bi = build_ids[q], nid = (bi - 1) >> 1, si = bi == 2 * nid + 1 ? 2 * nid + 2 : 2 * nid + 1;
      hist_of_node[si] = orc_pool_get(hsz);
    }
#pragma omp parallel for schedule(dynamic, 1)
    for (int q = 0; q < n_build; ++q) {
      int bi = build_ids[q], nid = (bi - 1) >> 1, si = bi == 2 * nid + 1 ? 2 * nid + 2 : 2 * nid + 1;
      orc_gh* Hs = hist_of_node[si];
      const orc_gh* Hb = hist_of_node[bi];
      const orc_gh* Hp = hist_of_node[nid];
      for (size_t q2 = 0; q2 < hsz; ++q2) { Hs[q2].g = Hp[q2].g - Hb[q2].g; Hs[q2].h = Hp[q2].h - Hb[q2].h; }
    }
    for (int q = 0; q < n_build; ++q) {
      int nid = (build_ids[q] - 1) >> 1;
      orc_pool_put(hist_of_node[nid]); hist_of_node[nid] = NULL;
    }
    orc_t_sub += orc_now() - t_ph;
  }
  for (int i = 0; i < ORC_MAX_NODES; ++i) { if (hist_of_node[i]) { orc_pool_put(hist_of_node[i]); hist_of_node[i] = NULL; } }
  /* the final row sets are in `cur`; make sure rows_a holds them for the caller */
  if (cur != rows_a) memcpy(rows_a, cur, sizeof(int64_t) * (size_t)ntr);
  free(all_mask); free(tree_mask); free(level_mask); free(build_ids);
}

typedef struct {
  double fitness;          /* what XgboostModel.cross_validate returns */
  int32_t rounds_run;      /* boosting rounds actually trained per fold */
  int32_t best_iteration;
  int32_t stopped_early;
} orc_result;

/* xgb.cv for one gene set.  trace_means: optional [num_boost_round] test-metric mean per round.
 * trees: optional [nfold][trees_rounds][ORC_MAX_NODES] dumps of the first `trees_rounds` rounds.
 * final_rows: optional [nfold][ntrain_max] final row sets of the LAST tree dumped (round trees_rounds-1). */
int orc_cv(const orc_ctx* c, const double* genes, const orc_opts* o, orc_result* res,
           double* trace_means, orc_node_dump* trees, int32_t trees_rounds,
           int64_t* final_rows, int64_t final_rows_stride) {
#ifdef _OPENMP
  if (o->nthread > 0) omp_set_num_threads(o->nthread);
#endif
  orc_params p;
  p.eta = (float)genes[G_ETA]; p.min_child_weight = (float)genes[G_MCW]; p.max_depth = (int)genes[G_DEPTH];
  p.gamma = (float)genes[G_GAMMA]; p.max_delta_step = (float)genes[G_MDS]; p.subsample = (float)genes[G_SUBSAMPLE];
  p.colsample_bytree = (float)genes[G_COLTREE]; p.colsample_bylevel = (float)genes[G_COLLEVEL];
  p.lambda = (float)genes[G_LAMBDA]; p.alpha = (float)genes[G_ALPHA]; p.spw = (float)genes[G_SPW];
  if (p.max_depth < 1 || p.max_depth > 10) return -1;
  const int nfold = c->nfold; const int f = c->f;
  /* per-fold state */
  float** ptr = (float**)calloc((size_t)nfold, sizeof(void*));   /* train margins */
  float** pte = (float**)calloc((size_t)nfold, sizeof(void*));   /* test margins */
  float** ytr = (float**)calloc((size_t)nfold, sizeof(void*));
  float** yte = (float**)calloc((size_t)nfold, sizeof(void*));
  int64_t ntr_max = 0;
  for (int k = 0; k < nfold; ++k) if (c->ntrain[k] > ntr_max) ntr_max = c->ntrain[k];
  float* g = (float*)malloc(sizeof(float) * (size_t)(ntr_max + 1));
  float* h = (float*)malloc(sizeof(float) * (size_t)(ntr_max + 1));
  int64_t* rows_a = (int64_t*)malloc(sizeof(int64_t) * (size_t)(ntr_max + 1));
  int64_t* rows_b = (int64_t*)malloc(sizeof(int64_t) * (size_t)(ntr_max + 1));
  orc_node* nodes = (orc_node*)calloc(ORC_MAX_NODES, sizeof(orc_node));
  orc_gh** hist_of_node = (orc_gh**)calloc(ORC_MAX_NODES, sizeof(void*));
  float* leaf_of_node = (float*)malloc(sizeof(float) * ORC_MAX_NODES);

  for (int k = 0; k < nfold; ++k) {
    int64_t ntr = c->ntrain[k], nte = c->ntest[k];
    ptr[k] = (float*)malloc(sizeof(float) * (size_t)(ntr + 1)); pte[k] = (float*)malloc(sizeof(float) * (size_t)(nte + 1));
    ytr[k] = (float*)malloc(sizeof(float) * (size_t)(ntr + 1)); yte[k] = (float*)malloc(sizeof(float) * (size_t)(nte + 1));
    for (int64_t r = 0; r < ntr; ++r) ytr[k][r] = c->y[c->train_rows[k][r]];
    for (int64_t r = 0; r < nte; ++r) yte[k][r] = c->y[c->test_rows[k][r]];
    /* base_score: xgboost src/objective/init_estimation.cc FitIntercept::InitEstimation + tree/fit_stump.cc:
     * gradient at margin 0, one Newton step -sum_g / max(sum_h, kRtEps), then PredTransform. */
    double sg = 0, sh = 0;
    for (int64_t r = 0; r < ntr; ++r) {
      float w = (ytr[k][r] == 1.0f) ? p.spw : 1.0f;
      float gi, hi;
      if (o->objective == ORC_OBJ_LOGISTIC) { float pr = orc_sigmoid(0.0f); gi = (pr - ytr[k][r]) * w; hi = fmaxf(pr * (1.0f - pr), 1e-16f) * w; }
      else { gi = (0.0f - ytr[k][r]) * w; hi = 1.0f * w; }
      sg += gi; sh += hi;
    }
    float stump = (float)(-sg / fmax(sh, (double)ORC_RT_EPS));
    float base_margin;
    if (o->objective == ORC_OBJ_LOGISTIC) {
      float bs = orc_sigmoid(stump);                 /* PredTransform */
      base_margin = -logf(1.0f / bs - 1.0f);         /* LogisticRegression::ProbToMargin */
    } else base_margin = stump;
    for (int64_t r = 0; r < ntr; ++r) ptr[k][r] = base_margin;
    for (int64_t r = 0; r < nte; ++r) pte[k][r] = base_margin;
  }

  double best_score = 0; int best_iter = 0; int stopped = 0; int rounds_run = 0; double last_mean = 0;
  double* means = (double*)malloc(sizeof(double) * (size_t)(o->num_boost_round + 1));
  for (int it = 0; it < o->num_boost_round; ++it) {
    double fold_metric_sum = 0;
    for (int k = 0; k < nfold; ++k) {
      int64_t ntr = c->ntrain[k], nte = c->ntest[k];
      /* X4: RegLossObj::GetGradient (src/objective/regression_obj.cu) */
      uint64_t skey = orc_stream_key(1, (uint32_t)k, (uint32_t)it, 0);
#pragma omp parallel for schedule(static)
      for (int64_t r = 0; r < ntr; ++r) {
        float yy = ytr[k][r];
        float w = (yy == 1.0f) ? p.spw : 1.0f;
        float gi, hi;
        if (o->objective == ORC_OBJ_LOGISTIC) { float pr = orc_sigmoid(ptr[k][r]); gi = (pr - yy) * w; hi = fmaxf(pr * (1.0f - pr), 1e-16f) * w; }
        else { gi = (ptr[k][r] - yy) * w; hi = 1.0f * w; }
        if (p.subsample < 1.0f && !(orc_u01(orc_draw(skey, (uint64_t)r)) < p.subsample)) { gi = 0.0f; hi = 0.0f; }
        g[r] = gi; h[r] = hi;
      }
      orc_grow_tree(c, k, &p, it, g, h, nodes, rows_a, rows_b, NULL, hist_of_node);
      /* leaf values: ApplyTreeSplit passes weight * learning_rate; root-only tree: eta * weight */
      for (int i = 0; i < ORC_MAX_NODES; ++i) leaf_of_node[i] = (nodes[i].state == 3) ? p.eta * nodes[i].weight : 0.0f;
      if (trees && it < trees_rounds) {
        orc_node_dump* td = trees + ((size_t)k * (size_t)trees_rounds + (size_t)it) * ORC_MAX_NODES;
        for (int i = 0; i < ORC_MAX_NODES; ++i) {
          /* split nodes carry the runner-up candidate's loss_chg in `value` (the engine's dump has 0 there) */
          if (nodes[i].state == 2) { td[i].feat = nodes[i].feat; td[i].bin = nodes[i].bin; td[i].value = nodes[i].second_chg; td[i].gain = nodes[i].loss_chg; }
          else if (nodes[i].state == 3) { td[i].feat = -1; td[i].bin = 0; td[i].value = leaf_of_node[i]; td[i].gain = nodes[i].loss_chg; }
          else { td[i].feat = -2; td[i].bin = 0; td[i].value = 0; td[i].gain = 0; }
        }
        if (final_rows && it == trees_rounds - 1) memcpy(final_rows + (size_t)k * (size_t)final_rows_stride, rows_a, sizeof(int64_t) * (size_t)ntr);
      }
      /* X8: UpdatePredictionCache -- every train row gets the leaf it ended in */
#pragma omp parallel for schedule(dynamic, 1)
      for (int i = 0; i < ORC_MAX_NODES; ++i) {
        if (nodes[i].state != 3) continue;
        /* every leaf owns the segment [begin, begin+count) of the final row set */
        float lv = leaf_of_node[i];
        for (int64_t q = 0; q < nodes[i].count; ++q) ptr[k][rows_a[nodes[i].begin + q]] += lv;
      }
      /* X8: test rows traverse the new tree on raw values: fvalue < split_cond goes left
       * (src/predictor/cpu_predictor.cc, RegTree::GetNext without missing values) */
      const float* cutsk = c->cuts[k];
      double msum = 0;
#pragma omp parallel for schedule(static) reduction(+ : msum)
      for (int64_t r = 0; r < nte; ++r) {
        const float* xr = c->X + c->test_rows[k][r] * c->ldx;
        int nid = 0;
        while (nodes[nid].state == 2) {
          float thr = cutsk[(size_t)nodes[nid].feat * ORC_MAX_BIN + nodes[nid].bin];
          nid = (xr[nodes[nid].feat] < thr) ? 2 * nid + 1 : 2 * nid + 2;
        }
        pte[k][r] += leaf_of_node[nid];
        /* X9: src/metric/elementwise_metric.cu EvalRowRMSE / EvalRowLogLoss, float per row, double sum */
        float yy = yte[k][r], e;
        if (o->metric == ORC_METRIC_LOGLOSS) {
          float py = (o->objective == ORC_OBJ_LOGISTIC) ? orc_sigmoid(pte[k][r]) : pte[k][r];
          const float eps = 1e-16f; float pneg = 1.0f - py;
          if (py < eps) e = -yy * logf(eps) - (1.0f - yy) * logf(1.0f - eps);
          else if (pneg < eps) e = -yy * logf(1.0f - eps) - (1.0f - yy) * logf(eps);
          else e = -yy * logf(py) - (1.0f - yy) * logf(pneg);
        } else {
          float pv = (o->objective == ORC_OBJ_LOGISTIC) ? orc_sigmoid(pte[k][r]) : pte[k][r];
          float d = yy - pv; e = d * d;
        }
        msum += e;
      }
      double fm = (o->metric == ORC_METRIC_RMSE) ? sqrt(msum / (double)nte) : msum / (double)nte;
      fold_metric_sum += fm;
    }
    /* aggcv: np.mean over folds; callback.EarlyStopping on the test mean (minimise, strict) */
    double mean = fold_metric_sum / (double)nfold;
    means[it] = mean; last_mean = mean; rounds_run = it + 1;
    if (trace_means) trace_means[it] = mean;
    if (o->early_stopping_rounds > 0) {
      if (it == 0) { best_score = mean; best_iter = 0; }
      else if (mean < best_score) { best_score = mean; best_iter = it; }
      else if (it - best_iter >= o->early_stopping_rounds) { stopped = 1; break; }
    }
  }
  if (o->early_stopping_rounds <= 0) { best_iter = rounds_run - 1; }
  if (getenv("ORC_PROFILE"))
    fprintf(stderr, "orc_cv phases (s): hist %.3f eval %.3f partition %.3f subtract %.3f\n", orc_t_hist, orc_t_eval, orc_t_part, orc_t_sub);
  orc_t_hist = orc_t_eval = orc_t_part = orc_t_sub = orc_t_grad = orc_t_pred = 0;
  res->rounds_run = rounds_run; res->best_iteration = best_iter; res->stopped_early = stopped;
  /* cv() truncates the table to best_iteration+1 rows when stopping fired; gentun takes the last row */
  res->fitness = stopped ? means[best_iter] : last_mean;
  if (f <= 0) res->fitness = NAN;
  for (int k = 0; k < nfold; ++k) { free(ptr[k]); free(pte[k]); free(ytr[k]); free(yte[k]); }
  free(ptr); free(pte); free(ytr); free(yte); free(g); free(h); free(rows_a); free(rows_b);
  free(nodes); free(hist_of_node); free(leaf_of_node); free(means);
  orc_pool_release();
  return 0;
}

/* The reference's per-fit path: xgb.DMatrix + xgb.cv rebuild slices, cuts and the bin matrix for
 * every individual (xgboost_models.py:32-36).  Used as the timed CPU baseline. */
int orc_cross_validate(const float* X, int64_t n, int32_t f, int64_t ldx, const float* y,
                       const double* genes, const orc_opts* o, orc_result* res) {
  orc_ctx* c = orc_ctx_create(X, n, f, ldx, y, o->nfold, o->nthread);
  int rc = orc_cv(c, genes, o, res, NULL, NULL, 0, NULL, 0);
  orc_ctx_destroy(c);
  return rc;
}

/* threads a parallel region really gets when `nthread` is asked for (0 = the OpenMP default): launchers such as
 * torch.distributed.run export OMP_NUM_THREADS=1, so the CPU baseline passes its core count explicitly and reports this */
int orc_threads_in_use(int nthread) {
#ifdef _OPENMP
  if (nthread > 0) omp_set_num_threads(nthread);
  int n = 1;
#pragma omp parallel
  {
#pragma omp single
    n = omp_get_num_threads();
  }
  return n;
#else
  (void)nthread;
  return 1;
#endif
}

int orc_max_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}
