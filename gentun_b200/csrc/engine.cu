// Host side of the engine and its C ABI (include/gentun_b200.h).
//
// Stands where XgboostModel.cross_validate (/root/reference/gentun/models/xgboost_models.py:28-37)
// calls into xgboost: gt_engine_create ~ xgb.DMatrix (:32), gt_engine_eval ~ xgb.cv (:33-36) + the
// 'test-<metric>-mean' last-row read (:37), for a whole batch of gene sets
// (Population.get_fittest, gentun/populations.py:55-58, evaluates them one by one).
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <exception>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <utility>
#include <vector>

#include "../../include/gentun_b200.h"
#include "common.cuh"
#include "kernels.h"

namespace {

thread_local std::string g_err = "";

int fail(int code, const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  g_err = buf;
  return code;
}

#define GT_CUDA(expr)                                                                          \
  do {                                                                                         \
    cudaError_t e__ = (expr);                                                                  \
    if (e__ != cudaSuccess) return fail(GT_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e__), __FILE__, __LINE__); \
  } while (0)

// ---- numpy legacy RandomState stream (np.random.seed(s); np.random.permutation(n)), as used by
// xgboost training.py mknfold.  MT19937 + masked-rejection interval + Fisher-Yates from the top.
// GENTUN_B200_TRACE=1: wall-clock milliseconds of the set-up phases on stderr (each mark drains the stream first)
struct PhaseTrace {
  bool on; cudaStream_t s; const char* what; std::chrono::steady_clock::time_point t;
  PhaseTrace(const char* what_, cudaStream_t s_) : on(getenv("GENTUN_B200_TRACE") != nullptr), s(s_), what(what_), t(std::chrono::steady_clock::now()) {}
  void mark(const char* phase) {
    if (!on) return;
    cudaStreamSynchronize(s);
    const auto now = std::chrono::steady_clock::now();
    fprintf(stderr, "[gentun_b200] %s: %-28s %9.2f ms\n", what, phase, std::chrono::duration<double, std::milli>(now - t).count());
    t = now;
  }
};

#ifdef GT_SKETCH_PROF
extern "C" void dump_sketch_prof();
#endif

struct Mt19937 {
  uint32_t s[624];
  int at;
  explicit Mt19937(uint32_t seed) {
    for (int i = 0; i < 624; ++i) { s[i] = seed; seed = 1812433253u * (seed ^ (seed >> 30)) + (uint32_t)(i + 1); }
    at = 624;
  }
  void refill() {
    for (int k = 0; k < 624; ++k) {
      uint32_t y = (s[k] & 0x80000000u) | (s[(k + 1) % 624] & 0x7fffffffu);
      s[k] = s[(k + 397) % 624] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
    }
    at = 0;
  }
  uint32_t next() {
    if (at == 624) refill();
    uint32_t y = s[at++];
    y ^= y >> 11; y ^= (y << 7) & 0x9d2c5680u; y ^= (y << 15) & 0xefc60000u; y ^= y >> 18;
    return y;
  }
  uint64_t below_or_equal(uint64_t top) {
    if (top == 0) return 0;
    uint64_t mask = top;
    for (int sh = 1; sh <= 32; sh <<= 1) mask |= mask >> sh;
    uint64_t v;
    if (top <= 0xffffffffull) {
      do { v = next() & mask; } while (v > top);
    } else {
      do { uint64_t hi = next(); uint64_t lo = next(); v = ((hi << 32) | lo) & mask; } while (v > top);
    }
    return v;
  }
};

void make_perm_host(int64_t n, uint32_t seed, int64_t* perm) {   // np.random.seed(seed); np.random.permutation(n)
  Mt19937 rng(seed);
  for (int64_t i = 0; i < n; ++i) perm[i] = i;
  for (int64_t i = n - 1; i > 0; --i) std::swap(perm[i], perm[(int64_t)rng.below_or_equal((uint64_t)i)]);
}

void make_folds_host(int64_t n, int32_t nfold, uint32_t seed, int64_t* perm, int64_t* fold_begin) {
  if (perm) make_perm_host(n, seed, perm);
  int64_t q = n / nfold, r = n % nfold, at = 0;   // np.array_split
  for (int k = 0; k < nfold; ++k) { fold_begin[k] = at; at += q + (k < r ? 1 : 0); }
  fold_begin[nfold] = at;
}

template <typename T>
struct DevBuf {
  T* p = nullptr;
  size_t count = 0;
  ~DevBuf() { release(); }
  void release() { if (p) cudaFree(p); p = nullptr; count = 0; }
  cudaError_t ensure(size_t c) {
    if (c <= count && p) return cudaSuccess;
    release();
    cudaError_t e = cudaMalloc(&p, std::max<size_t>(c, 1) * sizeof(T));
    if (e == cudaSuccess) count = c;
    return e;
  }
};

struct FoldSet {
  int nfold = 0;
  std::vector<int64_t> fold_begin;
  std::vector<int64_t> ntrain, ntest;
  std::vector<std::vector<uint32_t>> rows;   // fold order: train slice then test slice
  std::vector<DevBuf<uint32_t>> d_rows;
  std::vector<DevBuf<uint8_t>> d_bins;
  std::vector<DevBuf<uint8_t>> d_bins_col;
  std::vector<DevBuf<float>> d_y;
  std::vector<DevBuf<int>> d_count_hist;      // per fold: train rows per (feature, bin)
  DevBuf<float> d_cuts;                       // [nfold][f][256]
  DevBuf<int> d_ncuts;                        // [nfold][f]
  DevBuf<gt::FoldDev> d_folds;
  DevBuf<int> d_grp_maxnc;
  std::vector<int> h_ncuts;
  // label statistics of each train slice for the base-score stump (FitIntercept::InitEstimation)
  std::vector<double> sum_not1_sq, sum_not1_lg;   // sum over labels != 1 of (0 - y) / (0.5 - y)
  std::vector<int64_t> n1, nnot1;
  double prepare_ms = 0;
  // the shared-memory-resident regime (small.cu): per fold and feature, the train slice's rows sorted by bin
  bool small_ok = false;
  DevBuf<uint32_t> d_small_lists;             // [nfold][f][nt_stride] (bin << 16 | row)
};

// device buffers of gt_engine_eval, kept between calls (cudaMalloc / cudaFree of several GB per call
// cost more host time than a whole wave of kernels)
struct EvalBuffers {
  DevBuf<gt::JobParams> d_jp; DevBuf<gt::JobState> d_js; DevBuf<gt::IndState> d_is;
  DevBuf<gt::Node> d_nodes; DevBuf<gt::TreeNode> d_tree; DevBuf<float> d_pred; DevBuf<int2> d_gq;
  DevBuf<uint32_t> d_ridx; DevBuf<gt::HistEntry> d_hist; DevBuf<gt::HistItem> d_hitems; DevBuf<int> d_zlist;
  DevBuf<gt::PairRec> d_pairs; DevBuf<gt::PartItem> d_pitems; DevBuf<int> d_pleft; DevBuf<uint32_t> d_pflags;
  DevBuf<double> d_mpart; DevBuf<double> d_trace; DevBuf<unsigned long long> d_counters;
  DevBuf<double> d_fit; DevBuf<int> d_rounds; DevBuf<gt_tree_node> d_trees; DevBuf<gt_node_rows> d_nrows;
};

struct SmallBuffers {
  DevBuf<gt::JobParams> d_jp; DevBuf<gt::IndState> d_is; DevBuf<gt::Node> d_nodes; DevBuf<gt::TreeNode> d_tree;
  DevBuf<uint32_t> d_lists; DevBuf<double> d_fold_s; DevBuf<unsigned> d_bar; DevBuf<int> d_team_ind; DevBuf<int> d_queue;
  DevBuf<int> d_order; DevBuf<double> d_trace; DevBuf<double> d_fit; DevBuf<int> d_rounds;
};

}  // namespace

struct gt_engine {
  int device = 0;
  cudaStream_t stream = nullptr;
  int64_t n = 0, ldx = 0;
  int f = 0, fpad = 0, groups = 0;
  const float* d_X = nullptr;
  const float* d_y = nullptr;
  DevBuf<float> own_X, own_y;
  std::vector<float> h_y;
  // xgb.cv's row permutation depends on n alone (seed 0), not on nfold: a helper thread draws it (a serial Mersenne
  // Twister shuffle, 0.2 s at 10 M rows) while gt_engine_create copies the matrix to the device
  std::vector<int64_t> perm;
  std::thread perm_thread;
  int64_t mem_budget = 0;
  int collect_timing = 0;
  std::map<int, std::unique_ptr<FoldSet>> foldsets;
  EvalBuffers bufs;
  SmallBuffers small;
  int force_streaming = 0;   // never take the shared-memory-resident path (tests compare the two regimes)
  gt_stats stats{};
};

namespace {

void limit_size_level(size_t maxn, double eps, size_t* nlevel, size_t* limit) {   // quantile.h LimitSizeLevel
  *nlevel = 1;
  for (;;) {
    *limit = (size_t)std::ceil((double)*nlevel / eps) + 1;
    *limit = std::min(maxn, *limit);
    if (((size_t)1 << *nlevel) * *limit >= maxn) break;
    ++*nlevel;
  }
}

int prepare_impl(gt_engine* e, int nfold, FoldSet** out) {
  auto it = e->foldsets.find(nfold);
  if (it != e->foldsets.end()) { *out = it->second.get(); return GT_OK; }
  if (nfold < 2 || nfold > 32 || nfold > e->n) return fail(GT_ERR_ARG, "nfold=%d out of range for n=%lld", nfold, (long long)e->n);
  GT_CUDA(cudaSetDevice(e->device));
  std::unique_ptr<FoldSet> fs(new FoldSet());
  fs->nfold = nfold;
  const int64_t n = e->n;
  const int f = e->f;
  fs->fold_begin.resize(nfold + 1);
  PhaseTrace trace("prepare", e->stream);
  if (e->perm_thread.joinable()) e->perm_thread.join();
  if ((int64_t)e->perm.size() != n) { e->perm.resize(n); make_perm_host(n, 0u, e->perm.data()); }
  make_folds_host(n, nfold, 0u, nullptr, fs->fold_begin.data());
  const std::vector<int64_t>& perm = e->perm;
  trace.mark("fold permutation (host)");
  fs->ntrain.resize(nfold); fs->ntest.resize(nfold); fs->rows.resize(nfold);
  fs->d_rows.resize(nfold); fs->d_bins.resize(nfold); fs->d_bins_col.resize(nfold); fs->d_y.resize(nfold); fs->d_count_hist.resize(nfold);
  fs->sum_not1_sq.assign(nfold, 0); fs->sum_not1_lg.assign(nfold, 0); fs->n1.assign(nfold, 0); fs->nnot1.assign(nfold, 0);
  cudaEvent_t ev0, ev1;
  GT_CUDA(cudaEventCreate(&ev0)); GT_CUDA(cudaEventCreate(&ev1));
  struct EventGuard { cudaEvent_t a, b; ~EventGuard() { cudaEventDestroy(a); cudaEventDestroy(b); } } ev_guard{ev0, ev1};
  GT_CUDA(cudaEventRecord(ev0, e->stream));
  std::vector<const uint32_t*> h_row_ptrs(nfold);
  std::vector<long long> h_ntrain(nfold);
  std::vector<int> h_limit(nfold);
  size_t lmax = 1, nlevel_max = 1;
  // every fold's row order (DMatrix.slice(np.concatenate(other chunks)), then the test chunk) and label sums: independent
  // per fold and bound by the host's memory latency, so one thread each
  auto fold_rows_and_sums = [&](int k) {
    auto& rows = fs->rows[k];
    rows.resize(n);
    size_t o = 0;
    for (int c = 0; c < nfold; ++c) {
      if (c == k) continue;
      for (int64_t q = fs->fold_begin[c]; q < fs->fold_begin[c + 1]; ++q) rows[o++] = (uint32_t)perm[q];
    }
    for (int64_t q = fs->fold_begin[k]; q < fs->fold_begin[k + 1]; ++q) rows[o++] = (uint32_t)perm[q];
    int64_t n1 = 0, nnot1 = 0;
    double ssq = 0, slg = 0;
    for (int64_t q = 0; q < fs->ntrain[k]; ++q) {
      float y = e->h_y[rows[q]];
      if (y == 1.0f) n1++;
      else { nnot1++; ssq += (double)(0.0f - y); slg += (double)(0.5f - y); }
    }
    fs->n1[k] = n1; fs->nnot1[k] = nnot1; fs->sum_not1_sq[k] = ssq; fs->sum_not1_lg[k] = slg;
  };
  for (int k = 0; k < nfold; ++k) {
    int64_t nt = fs->fold_begin[k + 1] - fs->fold_begin[k];
    fs->ntest[k] = nt; fs->ntrain[k] = n - nt;
  }
  if (n >= (1 << 16)) {
    std::vector<std::thread> pool;
    int started = 1;                                     // folds [1, started) have a thread; the rest run here
    try {
      for (int k = 1; k < nfold; ++k) { pool.emplace_back(fold_rows_and_sums, k); started = k + 1; }
    } catch (const std::exception&) {}
    fold_rows_and_sums(0);
    for (int k = started; k < nfold; ++k) fold_rows_and_sums(k);
    for (auto& th : pool) th.join();
  } else {
    for (int k = 0; k < nfold; ++k) fold_rows_and_sums(k);
  }
  for (int k = 0; k < nfold; ++k) {
    GT_CUDA(fs->d_rows[k].ensure(n));
    GT_CUDA(cudaMemcpyAsync(fs->d_rows[k].p, fs->rows[k].data(), sizeof(uint32_t) * n, cudaMemcpyHostToDevice, e->stream));
    h_row_ptrs[k] = fs->d_rows[k].p;
    h_ntrain[k] = fs->ntrain[k];
    size_t ntr = (size_t)fs->ntrain[k];
    size_t n_bins = std::max<size_t>(1, std::min<size_t>(ntr, GT_MAX_BIN));
    double eps = 1.0 / (double)((float)n_bins * 8.0f);
    size_t nl, lim;
    limit_size_level(std::max<size_t>(ntr, 1), eps, &nl, &lim);
    h_limit[k] = (int)lim;
    lmax = std::max(lmax, lim); nlevel_max = std::max(nlevel_max, nl);
  }
  trace.mark("row orders + label sums");
  // ---- sketch -> cuts
  GT_CUDA(fs->d_cuts.ensure((size_t)nfold * f * GT_MAX_BIN));
  GT_CUDA(fs->d_ncuts.ensure((size_t)nfold * f));
  GT_CUDA(cudaMemsetAsync(fs->d_cuts.p, 0, sizeof(float) * (size_t)nfold * f * GT_MAX_BIN, e->stream));
  {
    gt::SketchPlan plan;
    plan.buf_cap = (int)(2 * lmax + 2);
    plan.nlevel_alloc = (int)nlevel_max + 3;
    int kp = 2; while (kp < (int)(2 * lmax)) kp <<= 1;
    plan.keys_per_stream = kp;
    plan.entries_per_stream = 3ll * plan.buf_cap + (long long)plan.nlevel_alloc * (long long)lmax;
    const int streams = nfold * f;
    size_t per_stream = (size_t)plan.keys_per_stream * 8 + (size_t)plan.entries_per_stream * 16;
    size_t free_b = 0, total_b = 0;
    GT_CUDA(cudaMemGetInfo(&free_b, &total_b));
    size_t room = (size_t)(0.6 * (double)free_b);
    int per_launch = (int)std::max<size_t>(1, std::min<size_t>((size_t)streams, room / std::max<size_t>(per_stream, 1)));
    DevBuf<unsigned long long> keys;
    DevBuf<float4> ents;
    DevBuf<const uint32_t*> d_row_ptrs;
    DevBuf<long long> d_ntrain;
    DevBuf<int> d_limit;
    GT_CUDA(keys.ensure((size_t)per_launch * plan.keys_per_stream));
    GT_CUDA(ents.ensure((size_t)per_launch * (size_t)plan.entries_per_stream));
    GT_CUDA(d_row_ptrs.ensure(nfold)); GT_CUDA(d_ntrain.ensure(nfold)); GT_CUDA(d_limit.ensure(nfold));
    GT_CUDA(cudaMemcpyAsync(d_row_ptrs.p, h_row_ptrs.data(), sizeof(void*) * nfold, cudaMemcpyHostToDevice, e->stream));
    GT_CUDA(cudaMemcpyAsync(d_ntrain.p, h_ntrain.data(), sizeof(long long) * nfold, cudaMemcpyHostToDevice, e->stream));
    GT_CUDA(cudaMemcpyAsync(d_limit.p, h_limit.data(), sizeof(int) * nfold, cudaMemcpyHostToDevice, e->stream));
    if (int crc = gt::configure_sketch()) return fail(GT_ERR_CUDA, "cudaFuncSetAttribute failed: %s", cudaGetErrorString((cudaError_t)crc));
    for (int s0 = 0; s0 < streams; s0 += per_launch) {
      int cnt = std::min(per_launch, streams - s0);
      gt::launch_sketch(e->d_X, e->ldx, f, nfold, d_row_ptrs.p, d_ntrain.p, plan, d_limit.p, keys.p, ents.p,
                        fs->d_cuts.p, fs->d_ncuts.p, s0, cnt, e->stream);
      GT_CUDA(cudaGetLastError());
    }
    GT_CUDA(cudaStreamSynchronize(e->stream));   // arenas are released on scope exit
  }
  trace.mark("sketch -> cuts");
#ifdef GT_SKETCH_PROF
  dump_sketch_prof();
#endif
  // ---- bins + labels in fold order
  std::vector<gt::FoldDev> h_folds(nfold);
  for (int k = 0; k < nfold; ++k) {
    GT_CUDA(fs->d_bins[k].ensure((size_t)n * e->fpad));
    GT_CUDA(fs->d_y[k].ensure(n));
    gt::launch_bin(e->d_X, e->ldx, f, e->fpad, fs->d_rows[k].p, n, fs->d_cuts.p + (size_t)k * f * GT_MAX_BIN,
                   fs->d_ncuts.p + (size_t)k * f, fs->d_bins[k].p, e->stream);
    GT_CUDA(cudaGetLastError());
    GT_CUDA(fs->d_bins_col[k].ensure((size_t)n * f));
    gt::launch_bins_to_columns(fs->d_bins[k].p, f, e->fpad, n, fs->d_bins_col[k].p, e->stream);
    GT_CUDA(cudaGetLastError());
    gt::launch_gather_y(e->d_y, fs->d_rows[k].p, n, fs->d_y[k].p, e->stream);
    GT_CUDA(cudaGetLastError());
    const size_t hist_cells = (size_t)e->groups * gt::kBins * gt::kLanes;
    GT_CUDA(fs->d_count_hist[k].ensure(hist_cells));
    GT_CUDA(cudaMemsetAsync(fs->d_count_hist[k].p, 0, sizeof(int) * hist_cells, e->stream));
    gt::launch_count_hist(fs->d_bins[k].p, f, e->fpad, fs->ntrain[k], fs->d_count_hist[k].p, e->stream);
    GT_CUDA(cudaGetLastError());
    h_folds[k].count_hist = fs->d_count_hist[k].p;
    h_folds[k].bins = fs->d_bins[k].p; h_folds[k].bins_col = fs->d_bins_col[k].p; h_folds[k].y = fs->d_y[k].p;
    h_folds[k].cuts = fs->d_cuts.p + (size_t)k * f * GT_MAX_BIN; h_folds[k].ncuts = fs->d_ncuts.p + (size_t)k * f;
    h_folds[k].ntrain = fs->ntrain[k]; h_folds[k].ntest = fs->ntest[k];
  }
  fs->h_ncuts.resize((size_t)nfold * f);
  GT_CUDA(cudaMemcpyAsync(fs->h_ncuts.data(), fs->d_ncuts.p, sizeof(int) * (size_t)nfold * f, cudaMemcpyDeviceToHost, e->stream));
  GT_CUDA(cudaStreamSynchronize(e->stream));
  trace.mark("binning + count histograms");
  std::vector<int> h_maxnc((size_t)nfold * e->groups, 0);
  for (int k = 0; k < nfold; ++k)
    for (int j = 0; j < f; ++j) {
      int& m = h_maxnc[(size_t)k * e->groups + j / gt::kLanes];
      m = std::max(m, fs->h_ncuts[(size_t)k * f + j]);
    }
  GT_CUDA(fs->d_grp_maxnc.ensure(h_maxnc.size()));
  GT_CUDA(cudaMemcpyAsync(fs->d_grp_maxnc.p, h_maxnc.data(), sizeof(int) * h_maxnc.size(), cudaMemcpyHostToDevice, e->stream));
  GT_CUDA(fs->d_folds.ensure(nfold));
  GT_CUDA(cudaMemcpyAsync(fs->d_folds.p, h_folds.data(), sizeof(gt::FoldDev) * nfold, cudaMemcpyHostToDevice, e->stream));
  {
    int64_t ntmax = 0;
    for (int k = 0; k < nfold; ++k) ntmax = std::max(ntmax, fs->ntrain[k]);
    if (ntmax <= gt::kSmallMaxTrain && n <= gt::kSmallMaxRows) {
      // small data: every feature's train rows sorted by bin (ties by row), the order small.cu keeps per node
      const int64_t nt_stride = (ntmax + 31) / 32 * 32;
      std::vector<uint8_t> h_bins((size_t)n * e->fpad);
      std::vector<uint32_t> h_lists((size_t)nfold * f * nt_stride, 0u);
      for (int k = 0; k < nfold; ++k) {
        GT_CUDA(cudaMemcpyAsync(h_bins.data(), fs->d_bins[k].p, h_bins.size(), cudaMemcpyDeviceToHost, e->stream));
        GT_CUDA(cudaStreamSynchronize(e->stream));
        const int64_t ntr = fs->ntrain[k];
        for (int j = 0; j < f; ++j) {
          uint32_t* out = h_lists.data() + ((size_t)k * f + j) * nt_stride;
          int64_t count[257] = {0};
          for (int64_t r = 0; r < ntr; ++r) ++count[h_bins[(size_t)r * e->fpad + j] + 1];
          for (int b = 0; b < 256; ++b) count[b + 1] += count[b];
          for (int64_t r = 0; r < ntr; ++r) {
            const uint32_t b = h_bins[(size_t)r * e->fpad + j];
            out[count[b]++] = (b << 16) | (uint32_t)r;
          }
        }
      }
      GT_CUDA(fs->d_small_lists.ensure(h_lists.size()));
      GT_CUDA(cudaMemcpyAsync(fs->d_small_lists.p, h_lists.data(), sizeof(uint32_t) * h_lists.size(), cudaMemcpyHostToDevice, e->stream));
      GT_CUDA(cudaStreamSynchronize(e->stream));
      fs->small_ok = true;
    }
  }
  GT_CUDA(cudaEventRecord(ev1, e->stream));
  GT_CUDA(cudaStreamSynchronize(e->stream));
  trace.mark("tables (+ small-data lists)");
  float ms = 0;
  GT_CUDA(cudaEventElapsedTime(&ms, ev0, ev1));
  fs->prepare_ms = ms;
  e->stats.prepare_ms = ms;
  *out = fs.get();
  e->foldsets[nfold] = std::move(fs);
  return GT_OK;
}

struct Genes { double v[GT_NUM_GENES]; };

float base_margin_for(const FoldSet& fs, int k, int objective, float spw) {
  // FitIntercept::InitEstimation: gradient at margin 0, one Newton step, PredTransform, ProbToMargin
  double sg, sh;
  if (objective == GT_OBJ_LOGISTIC) {
    float g1 = (0.5f - 1.0f) * spw, h1 = 0.25f * spw;
    sg = fs.sum_not1_lg[k] + (double)fs.n1[k] * (double)g1;
    sh = (double)fs.nnot1[k] * 0.25 + (double)fs.n1[k] * (double)h1;
  } else {
    float g1 = (0.0f - 1.0f) * spw, h1 = 1.0f * spw;
    sg = fs.sum_not1_sq[k] + (double)fs.n1[k] * (double)g1;
    sh = (double)fs.nnot1[k] + (double)fs.n1[k] * (double)h1;
  }
  float stump = (float)(-sg / std::max(sh, 1e-6));
  if (objective == GT_OBJ_LOGISTIC) {
    float x = std::fmin(-stump, 88.7f);
    float bs = 1.0f / (std::exp(x) + 1.0f + 1e-16f);
    return -std::log(1.0f / bs - 1.0f);
  }
  return stump;
}

size_t hist_cells_for_depth(int D, int groups) {
  // ping-pong level buffers: the largest even level and the largest odd level that get histograms
  int last = D - 1;                      // deepest level with histograms
  int even = last - (last & 1), odd = last - 1 + (last & 1);
  size_t cells = (size_t)1 << even;
  if (odd >= 1) cells += (size_t)1 << odd;
  return cells * (size_t)groups * gt::kBins * gt::kLanes;
}

// The shared-memory-resident regime: ONE cooperative launch evaluates every individual (small.cu).
int eval_small(gt_engine* e, FoldSet* fs, int n_ind, const double* genes, const gt_fit_opts* o, double* fitness_out,
               int32_t* rounds_out, const gt_dump* dump) {
  const int nfold = o->nfold;
  const int64_t n = e->n;
  int64_t ntmax = 0;
  for (int k = 0; k < nfold; ++k) ntmax = std::max(ntmax, fs->ntrain[k]);
  const int nt_stride = (int)((ntmax + 31) / 32 * 32);
  const int smem = gt::small_smem_bytes((int)n, nt_stride);
  const int max_teams = gt::small_max_teams(nfold, smem);
  if (max_teams < 1) return fail(GT_ERR_CUDA, "the shared-memory-resident kernel cannot hold a team of %d folds on this device", nfold);
  const int n_teams = std::min(n_ind, max_teams), n_ctas = n_teams * nfold;
  // most expensive first: long fits (small eta, deep trees) must not be the last ones started
  std::vector<int> order(n_ind);
  for (int i = 0; i < n_ind; ++i) order[i] = i;
  std::stable_sort(order.begin(), order.end(), [&](int a, int b) {
    const double* ga = genes + (size_t)a * GT_NUM_GENES; const double* gb = genes + (size_t)b * GT_NUM_GENES;
    const double ca = ga[2] / std::max(ga[0], 1e-3), cb = gb[2] / std::max(gb[0], 1e-3);
    return ca > cb;
  });
  std::vector<gt::JobParams> h_jp((size_t)n_ind * nfold);
  std::vector<gt::IndState> h_is(n_ind);
  for (int q = 0; q < n_ind; ++q) {
    const double* g = genes + (size_t)q * GT_NUM_GENES;
    memset(&h_is[q], 0, sizeof(gt::IndState));
    h_is[q].active = 1; h_is[q].out_index = q;
    for (int k = 0; k < nfold; ++k) {
      gt::JobParams& p = h_jp[(size_t)q * nfold + k];
      memset(&p, 0, sizeof(p));
      p.eta = (float)g[0]; p.mcw = (float)g[1]; p.max_depth = (int)g[2]; p.gamma = (float)g[3]; p.mds = (float)g[4];
      p.subsample = (float)g[5]; p.coltree = (float)g[6]; p.collevel = (float)g[7]; p.lambda = (float)g[8];
      p.alpha = (float)g[9]; p.spw = (float)g[10];
      p.fold = k; p.ind = q;
      p.base_margin = base_margin_for(*fs, k, o->objective, p.spw);
    }
  }
  SmallBuffers& B = e->small;
  const bool want_trace = dump && dump->trace_means;
  GT_CUDA(B.d_jp.ensure(h_jp.size())); GT_CUDA(B.d_is.ensure(n_ind));
  GT_CUDA(B.d_nodes.ensure((size_t)n_ctas * gt::kNodeCap)); GT_CUDA(B.d_tree.ensure((size_t)n_ctas * gt::kNodeCap));
  GT_CUDA(B.d_lists.ensure((size_t)n_ctas * 2 * e->f * nt_stride));
  GT_CUDA(B.d_fold_s.ensure((size_t)n_teams * 2 * nfold)); GT_CUDA(B.d_bar.ensure(n_teams)); GT_CUDA(B.d_team_ind.ensure((size_t)n_teams * 2));
  GT_CUDA(B.d_queue.ensure(1)); GT_CUDA(B.d_order.ensure(n_ind)); GT_CUDA(B.d_fit.ensure(n_ind)); GT_CUDA(B.d_rounds.ensure(n_ind));
  if (want_trace) GT_CUDA(B.d_trace.ensure((size_t)n_ind * o->num_boost_round));
  GT_CUDA(cudaMemcpyAsync(B.d_jp.p, h_jp.data(), sizeof(gt::JobParams) * h_jp.size(), cudaMemcpyHostToDevice, e->stream));
  GT_CUDA(cudaMemcpyAsync(B.d_is.p, h_is.data(), sizeof(gt::IndState) * n_ind, cudaMemcpyHostToDevice, e->stream));
  GT_CUDA(cudaMemcpyAsync(B.d_order.p, order.data(), sizeof(int) * n_ind, cudaMemcpyHostToDevice, e->stream));
  GT_CUDA(cudaMemsetAsync(B.d_bar.p, 0, sizeof(unsigned) * n_teams, e->stream));
  GT_CUDA(cudaMemsetAsync(B.d_queue.p, 0, sizeof(int), e->stream));
  GT_CUDA(cudaMemsetAsync(B.d_nodes.p, 0, sizeof(gt::Node) * (size_t)n_ctas * gt::kNodeCap, e->stream));
  GT_CUDA(cudaMemsetAsync(B.d_fit.p, 0xff, sizeof(double) * n_ind, e->stream));   // NaN until written
  GT_CUDA(cudaMemsetAsync(B.d_rounds.p, 0, sizeof(int) * n_ind, e->stream));
  if (want_trace) GT_CUDA(cudaMemsetAsync(B.d_trace.p, 0xff, sizeof(double) * (size_t)n_ind * o->num_boost_round, e->stream));
  gt::SmallWave w;
  memset(&w, 0, sizeof(w));
  w.n_inds = n_ind; w.nfold = nfold; w.n = (int)n; w.f = e->f; w.fpad = e->fpad; w.nt_stride = nt_stride;
  w.objective = o->objective; w.metric = o->metric; w.num_boost_round = o->num_boost_round; w.early_stopping_rounds = o->early_stopping_rounds;
  w.folds = fs->d_folds.p; w.base_lists = fs->d_small_lists.p; w.jp = B.d_jp.p; w.is = B.d_is.p; w.nodes = B.d_nodes.p; w.tree = B.d_tree.p;
  w.lists = B.d_lists.p; w.fold_s = B.d_fold_s.p; w.team_bar = B.d_bar.p; w.team_ind = B.d_team_ind.p; w.queue = B.d_queue.p;
  w.order = B.d_order.p; w.trace = want_trace ? B.d_trace.p : nullptr; w.fitness_out = B.d_fit.p; w.rounds_out = B.d_rounds.p;
  cudaEvent_t e0, e1;
  GT_CUDA(cudaEventCreate(&e0)); GT_CUDA(cudaEventCreate(&e1));
  struct EventGuard { cudaEvent_t a, b; ~EventGuard() { cudaEventDestroy(a); cudaEventDestroy(b); } } ev_guard{e0, e1};
  GT_CUDA(cudaEventRecord(e0, e->stream));
  const int lrc = gt::launch_small_cv(w, n_teams, smem, e->stream);
  if (lrc != 0) return fail(GT_ERR_CUDA, "cooperative launch of the shared-memory-resident kernel failed: %s", cudaGetErrorString((cudaError_t)lrc));
  GT_CUDA(cudaEventRecord(e1, e->stream));
  GT_CUDA(cudaMemcpyAsync(fitness_out, B.d_fit.p, sizeof(double) * n_ind, cudaMemcpyDeviceToHost, e->stream));
  std::vector<int> h_rounds(n_ind);
  GT_CUDA(cudaMemcpyAsync(h_rounds.data(), B.d_rounds.p, sizeof(int) * n_ind, cudaMemcpyDeviceToHost, e->stream));
  GT_CUDA(cudaStreamSynchronize(e->stream));
  if (rounds_out) for (int i = 0; i < n_ind; ++i) rounds_out[i] = h_rounds[i];
  if (want_trace) GT_CUDA(cudaMemcpy(dump->trace_means, B.d_trace.p, sizeof(double) * (size_t)n_ind * o->num_boost_round, cudaMemcpyDeviceToHost));
  float ms = 0;
  GT_CUDA(cudaEventElapsedTime(&ms, e0, e1));
  e->stats.eval_ms = ms; e->stats.hist_ms = 0; e->stats.hist_launches = 0; e->stats.kernel_launches = 1;
  e->stats.waves = (n_ind + n_teams - 1) / n_teams;   // how many individuals the busiest team takes from the queue
  e->stats.jobs = (int64_t)n_ind * nfold; e->stats.hist_row_features = 0; e->stats.hist_alg_bytes = 0;
  return GT_OK;
}

int eval_impl(gt_engine* e, int n_ind, const double* genes, const gt_fit_opts* o, double* fitness_out,
              int32_t* rounds_out, const gt_dump* dump) {
  if (!e || !genes || !o || !fitness_out || n_ind < 0) return fail(GT_ERR_ARG, "null argument");
  if (o->objective != GT_OBJ_SQUAREDERROR && o->objective != GT_OBJ_LOGISTIC) return fail(GT_ERR_ARG, "unknown objective %d", o->objective);
  if (o->metric != GT_METRIC_RMSE && o->metric != GT_METRIC_LOGLOSS) return fail(GT_ERR_ARG, "unknown metric %d", o->metric);
  if (o->num_boost_round < 1) return fail(GT_ERR_ARG, "num_boost_round must be >= 1");
  for (int i = 0; i < n_ind; ++i) {
    const double* g = genes + (size_t)i * GT_NUM_GENES;
    for (int q = 0; q < GT_NUM_GENES; ++q) if (!std::isfinite(g[q])) return fail(GT_ERR_ARG, "individual %d: gene %d is not finite", i, q);
    if (g[2] != std::floor(g[2]) || g[2] < 1 || g[2] > gt::kMaxDepth) return fail(GT_ERR_ARG, "individual %d: max_depth=%g outside 1..%d", i, g[2], gt::kMaxDepth);
    if (g[0] < 0) return fail(GT_ERR_ARG, "individual %d: eta < 0", i);
    if (g[1] < 0 || g[3] < 0 || g[4] < 0 || g[8] < 0 || g[9] < 0) return fail(GT_ERR_ARG, "individual %d: negative regulariser", i);
    for (int q = 5; q <= 7; ++q) if (g[q] < 0 || g[q] > 1) return fail(GT_ERR_ARG, "individual %d: sampling gene %d=%g outside [0,1]", i, q, g[q]);
    if (g[10] < 0) return fail(GT_ERR_ARG, "individual %d: scale_pos_weight < 0", i);
  }
  if (n_ind == 0) return GT_OK;
  GT_CUDA(cudaSetDevice(e->device));
  FoldSet* fs = nullptr;
  int rc = prepare_impl(e, o->nfold, &fs);
  if (rc != GT_OK) return rc;
  // small data: the whole fit of a job runs inside one persistent CTA (tree / row-set dumps exist only in the
  // streaming kernels, so a call that asks for them takes those)
  if (fs->small_ok && !e->force_streaming && !(dump && (dump->trees || dump->row_sets)))
    return eval_small(e, fs, n_ind, genes, o, fitness_out, rounds_out, dump);
  const int nfold = o->nfold;
  const int64_t n = e->n;
  int64_t ntmax = 0;
  for (int k = 0; k < nfold; ++k) ntmax = std::max(ntmax, fs->ntrain[k]);
  const int64_t nt_stride = (ntmax + 31) / 32 * 32;
  const int hist_cap = (int)(nt_stride / gt::kHistRows + 514);
  const int part_cap = (int)(nt_stride / gt::kPartRows + 514);
  const int pair_cap = 512;
  const int pred_chunks = (int)((n + gt::kPredRows - 1) / gt::kPredRows);
  if (pred_chunks > 65535) return fail(GT_ERR_ARG, "n=%lld rows exceeds this build's limit of %lld", (long long)n, 65535ll * gt::kPredRows);

  // order by depth so that a wave's level loop is not held open by one deep individual
  std::vector<int> order(n_ind);
  for (int i = 0; i < n_ind; ++i) order[i] = i;
  std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return genes[(size_t)a * GT_NUM_GENES + 2] > genes[(size_t)b * GT_NUM_GENES + 2]; });

  auto job_bytes = [&](int D) -> size_t {
    size_t b = (size_t)n * 4 + (size_t)nt_stride * 8 + (size_t)nt_stride * 8;
    b += hist_cells_for_depth(D, e->groups) * sizeof(gt::HistEntry);
    b += (size_t)gt::kNodeCap * (sizeof(gt::Node) + sizeof(gt::TreeNode));
    b += (size_t)hist_cap * (sizeof(gt::HistItem) + 4) + (size_t)pair_cap * sizeof(gt::PairRec);
    b += (size_t)part_cap * (sizeof(gt::PartItem) + 4 + gt::kPartRows / 8) + (size_t)pred_chunks * 8;
    b += sizeof(gt::JobParams) + sizeof(gt::JobState);
    return b;
  };
  size_t free_b = 0, total_b = 0;
  GT_CUDA(cudaMemGetInfo(&free_b, &total_b));
  size_t held = 0;   // buffers kept from the previous call are reusable
  held += e->bufs.d_pred.count * 4 + e->bufs.d_gq.count * 8 + e->bufs.d_ridx.count * 4 + e->bufs.d_hist.count * sizeof(gt::HistEntry);
  size_t budget = e->mem_budget > 0 ? (size_t)e->mem_budget : (size_t)(0.85 * (double)(free_b + held));
  std::vector<std::pair<int, int>> waves;   // [begin, end) in `order`
  {
    int b = 0;
    while (b < n_ind) {
      size_t used = 0; int q = b;
      while (q < n_ind && (q - b + 1) * nfold <= 16384 && (q - b + 1) <= gt::kMaxWaveInds) {
        size_t need = job_bytes((int)genes[(size_t)order[q] * GT_NUM_GENES + 2]) * (size_t)nfold;
        if (used + need > budget) break;
        used += need; ++q;
      }
      if (q == b) return fail(GT_ERR_NOMEM, "device memory budget %zu B cannot hold one individual (%zu B)", budget,
                              job_bytes((int)genes[(size_t)order[b] * GT_NUM_GENES + 2]) * (size_t)nfold);
      waves.emplace_back(b, q);
      b = q;
    }
  }
  size_t max_jobs = 0, max_inds = 0, max_hist = 0;
  for (auto& wv : waves) {
    size_t inds = wv.second - wv.first, hc = 0;
    for (int q = wv.first; q < wv.second; ++q) hc += hist_cells_for_depth((int)genes[(size_t)order[q] * GT_NUM_GENES + 2], e->groups) * nfold;
    max_inds = std::max(max_inds, inds); max_jobs = std::max(max_jobs, inds * nfold); max_hist = std::max(max_hist, hc);
  }
  EvalBuffers& B = e->bufs;
  auto &d_jp = B.d_jp; auto &d_js = B.d_js; auto &d_is = B.d_is; auto &d_nodes = B.d_nodes; auto &d_tree = B.d_tree;
  auto &d_pred = B.d_pred; auto &d_gq = B.d_gq; auto &d_ridx = B.d_ridx; auto &d_hist = B.d_hist; auto &d_hitems = B.d_hitems;
  auto &d_zlist = B.d_zlist; auto &d_pairs = B.d_pairs; auto &d_pitems = B.d_pitems; auto &d_pleft = B.d_pleft; auto &d_pflags = B.d_pflags;
  auto &d_mpart = B.d_mpart; auto &d_trace = B.d_trace; auto &d_counters = B.d_counters; auto &d_fit = B.d_fit;
  auto &d_rounds = B.d_rounds; auto &d_trees = B.d_trees; auto &d_nrows = B.d_nrows;
  GT_CUDA(d_jp.ensure(max_jobs)); GT_CUDA(d_js.ensure(max_jobs)); GT_CUDA(d_is.ensure(max_inds));
  GT_CUDA(d_nodes.ensure(max_jobs * gt::kNodeCap)); GT_CUDA(d_tree.ensure(max_jobs * gt::kNodeCap));
  GT_CUDA(d_pred.ensure(max_jobs * (size_t)n)); GT_CUDA(d_gq.ensure(max_jobs * (size_t)nt_stride));
  GT_CUDA(d_ridx.ensure(max_jobs * 2 * (size_t)nt_stride)); GT_CUDA(d_hist.ensure(max_hist));
  GT_CUDA(d_hitems.ensure(max_jobs * hist_cap)); GT_CUDA(d_zlist.ensure(max_jobs * hist_cap));
  GT_CUDA(d_pairs.ensure(max_jobs * pair_cap)); GT_CUDA(d_pitems.ensure(max_jobs * part_cap));
  GT_CUDA(d_pleft.ensure(max_jobs * part_cap)); GT_CUDA(d_pflags.ensure(max_jobs * part_cap * (gt::kPartRows / 32)));
  GT_CUDA(d_mpart.ensure(max_jobs * pred_chunks)); GT_CUDA(d_counters.ensure(8));
  GT_CUDA(d_fit.ensure(n_ind)); GT_CUDA(d_rounds.ensure(n_ind));
  const bool want_trace = dump && dump->trace_means;
  const int tree_rounds = (dump && dump->trees) ? std::min(dump->tree_rounds, o->num_boost_round) : 0;
  const bool want_rows = dump && dump->row_sets && dump->node_rows && dump->rows_round >= 0;
  if (want_trace) GT_CUDA(d_trace.ensure(max_inds * (size_t)o->num_boost_round));
  if (tree_rounds) GT_CUDA(d_trees.ensure(max_jobs * (size_t)tree_rounds * GT_MAX_NODES));
  if (want_rows) GT_CUDA(d_nrows.ensure(max_jobs * (size_t)GT_MAX_NODES));
  if (want_rows && dump->row_stride < nt_stride) return fail(GT_ERR_ARG, "row_stride %lld < %lld", (long long)dump->row_stride, (long long)nt_stride);
  GT_CUDA(cudaMemsetAsync(d_counters.p, 0, 8 * sizeof(unsigned long long), e->stream));
  GT_CUDA(cudaMemsetAsync(d_fit.p, 0xff, sizeof(double) * n_ind, e->stream));   // NaN until written
  GT_CUDA(cudaMemsetAsync(d_rounds.p, 0, sizeof(int) * n_ind, e->stream));

  rc = gt::configure_kernels();
  if (rc != 0) return fail(GT_ERR_CUDA, "cudaFuncSetAttribute failed: %s", cudaGetErrorString((cudaError_t)rc));

  std::vector<cudaEvent_t> ev;   // hist timing pairs
  auto get_event = [&](size_t idx) -> cudaEvent_t {
    while (ev.size() <= idx) { cudaEvent_t x; cudaEventCreate(&x); ev.push_back(x); }
    return ev[idx];
  };
  size_t ev_used = 0;
  double hist_ms = 0;
  long long launches = 0, hist_launches = 0;
  cudaEvent_t e0, e1;
  GT_CUDA(cudaEventCreate(&e0)); GT_CUDA(cudaEventCreate(&e1));
  struct EventGuard {   // every early return below releases the events
    cudaEvent_t a, b; std::vector<cudaEvent_t>* more;
    ~EventGuard() { cudaEventDestroy(a); cudaEventDestroy(b); for (auto x : *more) cudaEventDestroy(x); }
  } ev_guard{e0, e1, &ev};
  GT_CUDA(cudaEventRecord(e0, e->stream));
  auto drain_events = [&]() {
    for (size_t q = 0; q + 1 < ev_used; q += 2) { float ms = 0; cudaEventElapsedTime(&ms, ev[q], ev[q + 1]); hist_ms += ms; }
    ev_used = 0;
  };

  for (auto& wv : waves) {
    const int inds = wv.second - wv.first, jobs = inds * nfold;
    std::vector<gt::JobParams> h_jp(jobs);
    std::vector<gt::JobState> h_js(jobs);
    std::vector<gt::IndState> h_is(inds);
    size_t hoff = 0;
    int Dmax = 1;
    int lvl_jobs[gt::kMaxDepth + 2] = {0};   // jobs that still grow their tree at each level (jobs are ordered deepest first)
    for (int q = 0; q < inds; ++q) {
      const int src = order[wv.first + q];
      const double* g = genes + (size_t)src * GT_NUM_GENES;
      gt::IndState& is = h_is[q];
      memset(&is, 0, sizeof(is));
      is.active = 1; is.out_index = src;
      const int D = (int)g[2];
      Dmax = std::max(Dmax, D);
      for (int d = 0; d < D; ++d) lvl_jobs[d] += nfold;
      for (int k = 0; k < nfold; ++k) {
        gt::JobParams& p = h_jp[q * nfold + k];
        memset(&p, 0, sizeof(p));
        p.eta = (float)g[0]; p.mcw = (float)g[1]; p.max_depth = D; p.gamma = (float)g[3]; p.mds = (float)g[4];
        p.subsample = (float)g[5]; p.coltree = (float)g[6]; p.collevel = (float)g[7]; p.lambda = (float)g[8];
        p.alpha = (float)g[9]; p.spw = (float)g[10];
        p.fold = k; p.ind = q;
        p.base_margin = base_margin_for(*fs, k, o->objective, p.spw);
        p.uniform_h = o->objective == GT_OBJ_SQUAREDERROR && p.subsample >= 1.0f && (fs->n1[k] == 0 || p.spw == 1.0f);
        const int last = D - 1;
        const int even = last - (last & 1), odd = last - 1 + (last & 1);
        p.hist_off[0] = hoff; hoff += ((size_t)1 << even) * e->groups * gt::kBins * gt::kLanes;
        p.hist_off[1] = hoff; if (odd >= 1) hoff += ((size_t)1 << odd) * e->groups * gt::kBins * gt::kLanes;
        gt::JobState& s = h_js[q * nfold + k];
        memset(&s, 0, sizeof(s));
        s.active = 1;
      }
    }
    GT_CUDA(cudaMemcpyAsync(d_jp.p, h_jp.data(), sizeof(gt::JobParams) * jobs, cudaMemcpyHostToDevice, e->stream));
    GT_CUDA(cudaMemcpyAsync(d_js.p, h_js.data(), sizeof(gt::JobState) * jobs, cudaMemcpyHostToDevice, e->stream));
    GT_CUDA(cudaMemcpyAsync(d_is.p, h_is.data(), sizeof(gt::IndState) * inds, cudaMemcpyHostToDevice, e->stream));
    GT_CUDA(cudaMemsetAsync(d_nodes.p, 0, sizeof(gt::Node) * (size_t)jobs * gt::kNodeCap, e->stream));
    GT_CUDA(cudaMemsetAsync(d_tree.p, 0, sizeof(gt::TreeNode) * (size_t)jobs * gt::kNodeCap, e->stream));
    unsigned long long act = (unsigned long long)inds;
    GT_CUDA(cudaMemcpyAsync(d_counters.p + 2, &act, sizeof(act), cudaMemcpyHostToDevice, e->stream));
    if (want_trace) GT_CUDA(cudaMemsetAsync(d_trace.p, 0xff, sizeof(double) * (size_t)inds * o->num_boost_round, e->stream));

    gt::Wave w;
    memset(&w, 0, sizeof(w));
    w.n_jobs = jobs; w.n_inds = inds; w.nfold = nfold; w.n = n; w.nt_stride = nt_stride;
    w.f = e->f; w.fpad = e->fpad; w.groups = e->groups;
    w.objective = o->objective; w.metric = o->metric; w.num_boost_round = o->num_boost_round;
    w.early_stopping_rounds = o->early_stopping_rounds;
    w.hist_cap = hist_cap; w.pair_cap = pair_cap; w.part_cap = part_cap; w.pred_chunks = pred_chunks;
    for (int d = 0; d < gt::kMaxDepth + 2; ++d) w.lvl_jobs[d] = lvl_jobs[d];
    w.folds = fs->d_folds.p; w.grp_maxnc = fs->d_grp_maxnc.p;
    w.jp = d_jp.p; w.js = d_js.p; w.is = d_is.p; w.nodes = d_nodes.p; w.tree = d_tree.p; w.pred = d_pred.p;
    w.gq = d_gq.p; w.ridx = d_ridx.p; w.hist = d_hist.p; w.hitems = d_hitems.p; w.zlist = d_zlist.p;
    w.pairs = d_pairs.p; w.pitems = d_pitems.p; w.pleft = d_pleft.p; w.pflags = d_pflags.p; w.mpart = d_mpart.p;
    w.trace = want_trace ? d_trace.p : nullptr; w.counters = d_counters.p;

    gt::launch_init_wave(w, e->stream); ++launches;
    gt::launch_round_end(w, 1, d_fit.p, d_rounds.p, e->stream); ++launches;
    GT_CUDA(cudaGetLastError());
    const bool can_stop_early = o->early_stopping_rounds > 0 && o->early_stopping_rounds < o->num_boost_round;
    for (int round = 0; round < o->num_boost_round; ++round) {
      gt::launch_grad(w, e->stream); ++launches;
      for (int d = 0; d < Dmax; ++d) {
        gt::launch_plan_hist(w, d, e->stream);
        gt::launch_hist_zero(w, d, e->stream);
        if (e->collect_timing) cudaEventRecord(get_event(ev_used++), e->stream);
        launches += gt::launch_hist(w, d, e->stream);
        if (e->collect_timing) cudaEventRecord(get_event(ev_used++), e->stream);
        launches += gt::launch_scan(w, d, e->stream);
        launches += 2; ++hist_launches;
        if (d + 1 < Dmax) {
          gt::launch_plan_part(w, d, e->stream);
          gt::launch_part_count(w, d, e->stream);
          gt::launch_part_scatter(w, d, e->stream);
          launches += 3;
        }
      }
      if (round < tree_rounds) { gt::launch_dump_trees(w, round, tree_rounds, d_trees.p, e->stream); ++launches; }
      if (want_rows && round == dump->rows_round) {
        gt::launch_dump_node_rows(w, d_nrows.p, e->stream); ++launches;
        GT_CUDA(cudaStreamSynchronize(e->stream));
        for (int q = 0; q < inds; ++q)
          for (int k = 0; k < nfold; ++k) {
            const size_t dst = ((size_t)h_is[q].out_index * nfold + k);
            const int j = q * nfold + k;
            for (int b = 0; b < 2; ++b)
              GT_CUDA(cudaMemcpy(dump->row_sets + (dst * 2 + b) * (size_t)dump->row_stride, d_ridx.p + ((size_t)j * 2 + b) * nt_stride,
                                 sizeof(uint32_t) * (size_t)nt_stride, cudaMemcpyDeviceToHost));
            GT_CUDA(cudaMemcpy(dump->node_rows + dst * GT_MAX_NODES, d_nrows.p + (size_t)j * GT_MAX_NODES,
                               sizeof(gt_node_rows) * GT_MAX_NODES, cudaMemcpyDeviceToHost));
          }
      }
      gt::launch_predict(w, e->stream);
      gt::launch_round_end(w, 0, d_fit.p, d_rounds.p, e->stream);
      launches += 2;
      if (e->collect_timing && ev_used >= 4096) { GT_CUDA(cudaStreamSynchronize(e->stream)); drain_events(); }
      if (can_stop_early && (round % 8 == 7)) {
        unsigned long long left = 0;
        GT_CUDA(cudaMemcpyAsync(&left, d_counters.p + 2, sizeof(left), cudaMemcpyDeviceToHost, e->stream));
        GT_CUDA(cudaStreamSynchronize(e->stream));
        if (left == 0) break;
      }
    }
    GT_CUDA(cudaStreamSynchronize(e->stream));
    GT_CUDA(cudaGetLastError());
    if (e->collect_timing) drain_events();
    if (want_trace) {
      std::vector<double> h((size_t)inds * o->num_boost_round);
      GT_CUDA(cudaMemcpy(h.data(), d_trace.p, sizeof(double) * h.size(), cudaMemcpyDeviceToHost));
      for (int q = 0; q < inds; ++q)
        memcpy(dump->trace_means + (size_t)h_is[q].out_index * o->num_boost_round, h.data() + (size_t)q * o->num_boost_round,
               sizeof(double) * o->num_boost_round);
    }
    if (tree_rounds) {
      for (int q = 0; q < inds; ++q)
        for (int k = 0; k < nfold; ++k)
          GT_CUDA(cudaMemcpy(dump->trees + (((size_t)h_is[q].out_index * nfold + k) * dump->tree_rounds) * GT_MAX_NODES,
                             d_trees.p + ((size_t)(q * nfold + k) * tree_rounds) * GT_MAX_NODES,
                             sizeof(gt_tree_node) * (size_t)tree_rounds * GT_MAX_NODES, cudaMemcpyDeviceToHost));
    }
  }
  GT_CUDA(cudaEventRecord(e1, e->stream));
  GT_CUDA(cudaMemcpyAsync(fitness_out, d_fit.p, sizeof(double) * n_ind, cudaMemcpyDeviceToHost, e->stream));
  std::vector<int> h_rounds(n_ind);
  GT_CUDA(cudaMemcpyAsync(h_rounds.data(), d_rounds.p, sizeof(int) * n_ind, cudaMemcpyDeviceToHost, e->stream));
  unsigned long long h_counters[8];
  GT_CUDA(cudaMemcpyAsync(h_counters, d_counters.p, sizeof(h_counters), cudaMemcpyDeviceToHost, e->stream));
  GT_CUDA(cudaStreamSynchronize(e->stream));
  if (rounds_out) for (int i = 0; i < n_ind; ++i) rounds_out[i] = h_rounds[i];
  float ms = 0;
  GT_CUDA(cudaEventElapsedTime(&ms, e0, e1));
  e->stats.eval_ms = ms;
  e->stats.hist_ms = hist_ms;
  e->stats.hist_launches = hist_launches;
  e->stats.kernel_launches = launches;
  e->stats.waves = (int64_t)waves.size();
  e->stats.jobs = (int64_t)n_ind * nfold;
  // algorithmic bytes of the histogram launches: rows x (F bin bytes + 8 B (g,h) + 4 B row id)
  // + one F x 256 x 16 B histogram written per built node
  e->stats.hist_row_features = (double)h_counters[0] * (double)e->f;
  e->stats.hist_alg_bytes = (double)h_counters[0] * ((double)e->f + 12.0) + (double)h_counters[1] * (double)e->f * 256.0 * 16.0;
  if (h_counters[7])
    return fail(GT_ERR_STATE, "%s work-item list overflowed its capacity: results of this call are void", h_counters[7] == 1 ? "histogram" : "partition");
  // A non-finite fitness (xgboost would hand back nan / inf for that individual) is left in fitness_out for the caller
  // to see; the other individuals' values stand.
  return GT_OK;
}

}  // namespace

namespace gt {
void set_error_message(const char* msg) { g_err = msg ? msg : ""; }
}  // namespace gt

namespace {

// Host matrix as the caller holds it: element type, strides in elements (one of them 1).
struct HostMatrix { const void* p; int dtype; int64_t row_stride, col_stride; };

// Pinned staging buffers, kept for the life of the process (allocating pinned memory costs milliseconds per call and
// the GA makes an engine per dataset, not per individual).  One ingestion at a time uses them.
struct PinnedStage {
  std::mutex mu;
  void* buf[2] = {nullptr, nullptr};
  static constexpr size_t kBytes = 16u << 20;
  cudaError_t ensure() {
    for (int b = 0; b < 2; ++b)
      if (!buf[b]) { cudaError_t e = cudaHostAlloc(&buf[b], kBytes, cudaHostAllocDefault); if (e != cudaSuccess) return e; }
    return cudaSuccess;
  }
};
PinnedStage g_stage;

// X (host, any supported type / layout) -> e->own_X (device, float, row-major [n][f]): tiles of rows go through the two
// pinned buffers, so the host-side gather of tile t+1 overlaps the DMA and the cast kernel of tile t.
int ingest_matrix(gt_engine* e, const HostMatrix& m) {
  const int64_t n = e->n; const int f = e->f;
  const size_t esz = m.dtype == GT_DTYPE_F64 ? 8 : 4;
  const bool colmajor = m.row_stride == 1 && m.col_stride != 1;
  std::lock_guard<std::mutex> lock(g_stage.mu);
  GT_CUDA(g_stage.ensure());
  const int64_t row_elems = colmajor ? f : m.row_stride;   // elements a tile holds per row
  int64_t R = (int64_t)(PinnedStage::kBytes / ((size_t)row_elems * esz)) / 32 * 32;
  if (R < 32) return fail(GT_ERR_ARG, "row stride %lld too large to stage", (long long)m.row_stride);
  if (R > n) R = n;
  DevBuf<unsigned char> d_tile[2];
  cudaEvent_t done[2];
  for (int b = 0; b < 2; ++b) { GT_CUDA(d_tile[b].ensure(PinnedStage::kBytes)); GT_CUDA(cudaEventCreateWithFlags(&done[b], cudaEventDisableTiming)); }
  int rc = GT_OK;
  int64_t t = 0;
  for (int64_t r0 = 0; r0 < n && rc == GT_OK; r0 += R, ++t) {
    const int b = (int)(t & 1);
    const int64_t rows = std::min<int64_t>(R, n - r0);
    if (t >= 2 && cudaEventSynchronize(done[b]) != cudaSuccess) { rc = fail(GT_ERR_CUDA, "staging event failed"); break; }
    unsigned char* pin = (unsigned char*)g_stage.buf[b];
    size_t bytes;
    // the host-side gather into the pinned tile is the slow leg for large matrices (one core copies ~6 GB/s): a few
    // threads share it (pieces of 64 KB at least)
    std::vector<std::pair<const unsigned char*, unsigned char*>> piece_ptr;
    std::vector<size_t> piece_len;
    auto add_piece = [&](const unsigned char* src, unsigned char* dst, size_t len) {
      const size_t step = std::max<size_t>(64u << 10, (len + 7) / 8);
      for (size_t o = 0; o < len; o += step) { piece_ptr.emplace_back(src + o, dst + o); piece_len.push_back(std::min(step, len - o)); }
    };
    if (colmajor) {   // tile = [f][rows]
      for (int c = 0; c < f; ++c)
        add_piece((const unsigned char*)m.p + ((size_t)c * m.col_stride + (size_t)r0) * esz, pin + (size_t)c * rows * esz, (size_t)rows * esz);
      bytes = (size_t)f * rows * esz;
    } else {          // tile = [rows][row_stride]
      bytes = ((size_t)(rows - 1) * m.row_stride + f) * esz;
      add_piece((const unsigned char*)m.p + (size_t)r0 * m.row_stride * esz, pin, bytes);
    }
    {
      const unsigned hw = std::thread::hardware_concurrency();
      const int n_thr = (int)std::max(1u, std::min(std::min(8u, hw ? hw / 2 : 1u), (unsigned)(bytes >> 20)));   // tiles below a few MB: one thread
      auto work = [&](int t) { for (size_t q = (size_t)t; q < piece_ptr.size(); q += (size_t)n_thr) memcpy(piece_ptr[q].second, piece_ptr[q].first, piece_len[q]); };
      std::vector<std::thread> pool;
      int started = 1;                                   // shares [1, started) have a thread; the rest are copied here
      try {
        for (int t = 1; t < n_thr; ++t) { pool.emplace_back(work, t); started = t + 1; }
      } catch (const std::exception&) {}
      work(0);
      for (int t = started; t < n_thr; ++t) work(t);
      for (auto& th : pool) th.join();
    }
    cudaError_t ce = cudaMemcpyAsync(d_tile[b].p, pin, bytes, cudaMemcpyHostToDevice, e->stream);
    if (ce != cudaSuccess) { rc = fail(GT_ERR_CUDA, "H2D of a staged tile failed: %s", cudaGetErrorString(ce)); break; }
    gt::launch_ingest_tile(d_tile[b].p, m.dtype, colmajor ? 1 : 0, colmajor ? rows : m.row_stride, (int)rows, f, e->own_X.p, e->ldx, r0, e->stream);
    cudaEventRecord(done[b], e->stream);
  }
  cudaError_t ce = cudaStreamSynchronize(e->stream);
  for (int b = 0; b < 2; ++b) cudaEventDestroy(done[b]);
  if (rc == GT_OK && ce != cudaSuccess) rc = fail(GT_ERR_CUDA, "ingestion failed: %s", cudaGetErrorString(ce));
  return rc;
}

int create_impl(const HostMatrix* hm, const float* X_dev_or_null, int64_t n, int32_t f, int64_t ldx, const void* y, int y_dtype,
                const gt_data_opts* opts, gt_engine** out) {
  if (!y || !out || n < 2 || f < 1) return fail(GT_ERR_ARG, "bad data arguments");
  if (f > 256) return fail(GT_ERR_ARG, "f=%d exceeds 256 features", f);
  if (n >= (1ll << 31)) return fail(GT_ERR_ARG, "n=%lld exceeds 2^31 rows", (long long)n);
  int count = 0;
  cudaError_t ce = cudaGetDeviceCount(&count);
  if (ce != cudaSuccess || count == 0)
    return fail(GT_ERR_CUDA, "no CUDA device available (%s); this engine has no CPU path", cudaGetErrorString(ce));
  std::unique_ptr<gt_engine> e(new gt_engine());
  e->device = opts ? opts->device : 0;
  if (e->device < 0 || e->device >= count) return fail(GT_ERR_ARG, "device %d out of range (%d devices)", e->device, count);
  e->mem_budget = opts ? opts->mem_budget_bytes : 0;
  e->collect_timing = opts ? opts->reserved[0] : 0;
  e->force_streaming = (opts ? opts->reserved[1] : 0) || getenv("GENTUN_B200_NO_SMALL") != nullptr;
  e->n = n; e->f = f; e->ldx = ldx;
  e->fpad = (f + gt::kLanes - 1) / gt::kLanes * gt::kLanes;
  e->groups = e->fpad / gt::kLanes;
  GT_CUDA(cudaSetDevice(e->device));
  GT_CUDA(cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking));
  struct StreamGuard {   // the stream dies with a failed create
    gt_engine* e; bool armed = true;
    ~StreamGuard() {
      if (!armed) return;
      if (e->perm_thread.joinable()) e->perm_thread.join();
      if (e->stream) { cudaStreamDestroy(e->stream); e->stream = nullptr; }
    }
  } guard{e.get()};
  PhaseTrace trace("create", e->stream);
  e->h_y.resize(n);
  if (!hm) {
    e->d_X = X_dev_or_null; e->d_y = (const float*)y;
    GT_CUDA(cudaMemcpy(e->h_y.data(), y, sizeof(float) * n, cudaMemcpyDeviceToHost));
  } else {
    if (y_dtype == GT_DTYPE_F32) memcpy(e->h_y.data(), y, sizeof(float) * n);
    else if (y_dtype == GT_DTYPE_F64) for (int64_t i = 0; i < n; ++i) e->h_y[i] = (float)((const double*)y)[i];
    else if (y_dtype == GT_DTYPE_I64) for (int64_t i = 0; i < n; ++i) e->h_y[i] = (float)((const int64_t*)y)[i];
    else return fail(GT_ERR_ARG, "unknown label type %d", y_dtype);
    GT_CUDA(e->own_X.ensure((size_t)n * ldx));
    GT_CUDA(e->own_y.ensure(n));
    GT_CUDA(cudaMemcpyAsync(e->own_y.p, e->h_y.data(), sizeof(float) * n, cudaMemcpyHostToDevice, e->stream));
    trace.mark("labels + allocation");
    if (n >= (1 << 16)) {
      gt_engine* ep = e.get();
      ep->perm.resize(n);
      try {
        ep->perm_thread = std::thread([ep, n]() { make_perm_host(n, 0u, ep->perm.data()); });
      } catch (const std::exception&) {   // no thread to be had: gt_engine_prepare draws the permutation itself
        ep->perm.clear();
      }
    }
    int rc = ingest_matrix(e.get(), *hm);
    if (rc != GT_OK) { if (e->perm_thread.joinable()) e->perm_thread.join(); return rc; }
    e->d_X = e->own_X.p; e->d_y = e->own_y.p;
    trace.mark("matrix ingestion");
  }
  {
    DevBuf<unsigned long long> d_bad;
    unsigned long long h_bad[2] = {0, 0};
    GT_CUDA(d_bad.ensure(2));
    GT_CUDA(cudaMemsetAsync(d_bad.p, 0, sizeof(h_bad), e->stream));
    gt::launch_count_nonfinite(e->d_X, n, f, ldx, e->d_y, d_bad.p, e->stream);
    GT_CUDA(cudaMemcpyAsync(h_bad, d_bad.p, sizeof(h_bad), cudaMemcpyDeviceToHost, e->stream));
    GT_CUDA(cudaStreamSynchronize(e->stream));
    if (h_bad[0] || h_bad[1])
      return fail(GT_ERR_ARG, "%llu values of X and %llu labels are NaN or inf: xgb.DMatrix would treat NaN as missing and refuse inf; "
                  "this engine has no missing-value path", h_bad[0], h_bad[1]);
  }
  trace.mark("non-finite check");
  guard.armed = false;
  *out = e.release();
  return GT_OK;
}

}  // namespace

extern "C" {

int gt_engine_create(const float* X, int64_t n, int32_t f, int64_t ldx, const float* y, const gt_data_opts* opts,
                     gt_engine** out) {
  if (!X || ldx < f) return fail(GT_ERR_ARG, "bad data arguments");
  if (opts && opts->x_on_device) return create_impl(nullptr, X, n, f, ldx, y, GT_DTYPE_F32, opts, out);
  const HostMatrix hm{X, GT_DTYPE_F32, ldx, 1};
  return create_impl(&hm, nullptr, n, f, f, y, GT_DTYPE_F32, opts, out);
}

int gt_engine_create_from(const void* X, int32_t x_dtype, int64_t n, int32_t f, int64_t x_row_stride, int64_t x_col_stride,
                          const void* y, int32_t y_dtype, const gt_data_opts* opts, gt_engine** out) {
  if (!X) return fail(GT_ERR_ARG, "bad data arguments");
  if (opts && opts->x_on_device) return fail(GT_ERR_ARG, "gt_engine_create_from takes host arrays");
  if (x_dtype != GT_DTYPE_F32 && x_dtype != GT_DTYPE_F64) return fail(GT_ERR_ARG, "unknown matrix element type %d", x_dtype);
  const bool rowmajor = x_col_stride == 1 && x_row_stride >= f, colmajor = x_row_stride == 1 && x_col_stride >= n;
  if (!rowmajor && !colmajor) return fail(GT_ERR_ARG, "matrix must be row-major or column-major (strides %lld, %lld)", (long long)x_row_stride, (long long)x_col_stride);
  const HostMatrix hm{X, x_dtype, rowmajor ? x_row_stride : 1, rowmajor ? 1 : x_col_stride};
  return create_impl(&hm, nullptr, n, f, f, y, y_dtype, opts, out);
}

int gt_engine_prepare(gt_engine* e, int32_t nfold) {
  if (!e) return fail(GT_ERR_ARG, "null engine");
  FoldSet* fs = nullptr;
  return prepare_impl(e, nfold, &fs);
}

int gt_engine_eval(gt_engine* e, int32_t n_ind, const double* genes, const gt_fit_opts* opts, double* fitness_out,
                   int32_t* rounds_out) {
  return eval_impl(e, n_ind, genes, opts, fitness_out, rounds_out, nullptr);
}

int gt_engine_eval_ex(gt_engine* e, int32_t n_ind, const double* genes, const gt_fit_opts* opts, double* fitness_out,
                      int32_t* rounds_out, const gt_dump* dump) {
  return eval_impl(e, n_ind, genes, opts, fitness_out, rounds_out, dump);
}

#ifdef GT_SCAN_STATS
extern "C" void dump_scan_stats();
#endif
#ifdef GT_SMALL_PROF
extern "C" void dump_small_prof();
#endif
void gt_engine_destroy(gt_engine* e) {
#ifdef GT_SCAN_STATS
  dump_scan_stats();
#endif
  if (!e) return;
#ifdef GT_SMALL_PROF
  dump_small_prof();
#endif
  if (e->perm_thread.joinable()) e->perm_thread.join();
  cudaSetDevice(e->device);
  if (e->stream) { cudaStreamSynchronize(e->stream); }
  e->foldsets.clear();
  if (e->stream) cudaStreamDestroy(e->stream);
  delete e;
}

int gt_engine_fold_sizes(gt_engine* e, int32_t nfold, int32_t k, int64_t* ntrain, int64_t* ntest) {
  if (!e) return fail(GT_ERR_ARG, "null engine");
  FoldSet* fs = nullptr;
  int rc = prepare_impl(e, nfold, &fs);
  if (rc != GT_OK) return rc;
  if (k < 0 || k >= nfold) return fail(GT_ERR_ARG, "fold %d out of range", k);
  if (ntrain) *ntrain = fs->ntrain[k];
  if (ntest) *ntest = fs->ntest[k];
  return GT_OK;
}

int gt_engine_get_fold_rows(gt_engine* e, int32_t nfold, int32_t k, int64_t* train_rows, int64_t* test_rows) {
  if (!e) return fail(GT_ERR_ARG, "null engine");
  FoldSet* fs = nullptr;
  int rc = prepare_impl(e, nfold, &fs);
  if (rc != GT_OK) return rc;
  if (k < 0 || k >= nfold) return fail(GT_ERR_ARG, "fold %d out of range", k);
  // read the device copy back so the test sees what the kernels see
  std::vector<uint32_t> h(e->n);
  GT_CUDA(cudaMemcpy(h.data(), fs->d_rows[k].p, sizeof(uint32_t) * e->n, cudaMemcpyDeviceToHost));
  if (train_rows) for (int64_t i = 0; i < fs->ntrain[k]; ++i) train_rows[i] = h[i];
  if (test_rows) for (int64_t i = 0; i < fs->ntest[k]; ++i) test_rows[i] = h[fs->ntrain[k] + i];
  return GT_OK;
}

int gt_engine_get_cuts(gt_engine* e, int32_t nfold, int32_t k, float* cuts, int32_t* ncuts) {
  if (!e) return fail(GT_ERR_ARG, "null engine");
  FoldSet* fs = nullptr;
  int rc = prepare_impl(e, nfold, &fs);
  if (rc != GT_OK) return rc;
  if (k < 0 || k >= nfold) return fail(GT_ERR_ARG, "fold %d out of range", k);
  if (cuts) GT_CUDA(cudaMemcpy(cuts, fs->d_cuts.p + (size_t)k * e->f * GT_MAX_BIN, sizeof(float) * (size_t)e->f * GT_MAX_BIN, cudaMemcpyDeviceToHost));
  if (ncuts) GT_CUDA(cudaMemcpy(ncuts, fs->d_ncuts.p + (size_t)k * e->f, sizeof(int) * e->f, cudaMemcpyDeviceToHost));
  return GT_OK;
}

int gt_engine_get_bins(gt_engine* e, int32_t nfold, int32_t k, uint8_t* bins) {
  if (!e || !bins) return fail(GT_ERR_ARG, "null argument");
  FoldSet* fs = nullptr;
  int rc = prepare_impl(e, nfold, &fs);
  if (rc != GT_OK) return rc;
  if (k < 0 || k >= nfold) return fail(GT_ERR_ARG, "fold %d out of range", k);
  GT_CUDA(cudaMemcpy2D(bins, e->f, fs->d_bins[k].p, e->fpad, e->f, e->n, cudaMemcpyDeviceToHost));
  return GT_OK;
}

int gt_engine_get_stats(gt_engine* e, gt_stats* out) {
  if (!e || !out) return fail(GT_ERR_ARG, "null argument");
  *out = e->stats;
  return GT_OK;
}

int gt_make_folds(int64_t n, int32_t nfold, uint32_t seed, int64_t* perm, int64_t* fold_begin) {
  if (n < 1 || nfold < 1 || !perm || !fold_begin) return fail(GT_ERR_ARG, "bad arguments");
  make_folds_host(n, nfold, seed, perm, fold_begin);
  return GT_OK;
}

const char* gt_last_error(void) { return g_err.c_str(); }
int gt_version(void) { return 100; }

}  // extern "C"
