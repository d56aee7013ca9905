// X2/X3 of the hot path: per-fold weighted quantile sketch -> cut points, and value -> uint8 bin.
//
// Replaces what every xgboost Booster does lazily inside xgb.cv (call site
// /root/reference/gentun/models/xgboost_models.py:33-36): HostSketchContainer / WQuantileSketch
// (upstream src/common/quantile.{h,cc}) over the rows of each fold's train slice in slice order,
// then GHistIndexMatrix binning (upstream src/common/hist_util.h SearchBin).  The sketch is an
// order-dependent streaming algorithm; it is reproduced stream by stream (one CTA per
// fold x feature), with the queue fill expressed as a run-length scan, the queue sort as a
// CTA-wide merge sort in shared memory, and the summary merge/prune hierarchy as CTA-parallel
// passes (one thread walks it for streams of 2^24 rows or more, where fp32 ranks stop being exact).
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>

#include "common.cuh"
#include "kernels.h"

namespace gt {

namespace {

constexpr int kSketchThreads = 1024;

#ifdef GT_SKETCH_PROF   // experiment builds: cycles of every CTA's thread 0 by phase, summed
__device__ unsigned long long g_sketch_prof[8];
#define SK_PROF_DECL long long prof_t = clock64()
#define SK_PROF(i) do { if (threadIdx.x == 0) { const long long now__ = clock64(); atomicAdd(&g_sketch_prof[i], (unsigned long long)(now__ - prof_t)); prof_t = now__; } } while (0)
#else
#define SK_PROF_DECL
#define SK_PROF(i)
#endif

struct Summary { float4* d; int size; };   // float4 = (rmin, rmax, wmin, value)

__device__ __forceinline__ float rmin_next(const float4& e) { return __fadd_rn(e.x, e.z); }
__device__ __forceinline__ float rmax_prev(const float4& e) { return __fsub_rn(e.y, e.z); }

__device__ void s_copy(Summary& dst, const Summary& src) {
  for (int i = 0; i < src.size; ++i) dst.d[i] = src.d[i];
  dst.size = src.size;
}

// WQSummary::SetPrune
__device__ void s_prune(Summary& dst, const Summary& src, int maxsize) {
  if (src.size <= maxsize) { s_copy(dst, src); return; }
  const float begin = src.d[0].y;
  const float range = __fsub_rn(src.d[src.size - 1].x, src.d[0].y);
  const int n = maxsize - 1;
  const float fn = (float)n;
  dst.d[0] = src.d[0];
  int out = 1;
  int i = 1, lastidx = 0;
  float4 cur = src.d[1];       // src[i]
  float4 nxt = src.d[i + 1 < src.size ? i + 1 : i];   // src[i+1]
  for (int k = 1; k < n; ++k) {
    float dx2 = __fmul_rn(2.0f, __fadd_rn(__fdiv_rn(__fmul_rn((float)k, range), fn), begin));
    while (i < src.size - 1 && dx2 >= __fadd_rn(nxt.y, nxt.x)) {
      ++i; cur = nxt;
      if (i + 1 < src.size) nxt = src.d[i + 1];
    }
    if (i == src.size - 1) break;
    if (dx2 < __fadd_rn(rmin_next(cur), rmax_prev(nxt))) {
      if (i != lastidx) { dst.d[out++] = cur; lastidx = i; }
    } else {
      if (i + 1 != lastidx) { dst.d[out++] = nxt; lastidx = i + 1; }
    }
  }
  if (lastidx != src.size - 1) dst.d[out++] = src.d[src.size - 1];
  dst.size = out;
}

// WQSummary::SetCombine followed by FixError
__device__ void s_combine(Summary& dst, const Summary& sa, const Summary& sb) {
  if (sa.size == 0) { s_copy(dst, sb); return; }
  if (sb.size == 0) { s_copy(dst, sa); return; }
  int ia = 0, ib = 0, o = 0;
  float aprev = 0.f, bprev = 0.f;
  float4 a = sa.d[0], b = sb.d[0];
  while (ia < sa.size && ib < sb.size) {
    float4 r;
    if (a.w == b.w) {
      r = make_float4(__fadd_rn(a.x, b.x), __fadd_rn(a.y, b.y), __fadd_rn(a.z, b.z), a.w);
      aprev = rmin_next(a); bprev = rmin_next(b);
      ++ia; ++ib;
      if (ia < sa.size) a = sa.d[ia];
      if (ib < sb.size) b = sb.d[ib];
    } else if (a.w < b.w) {
      r = make_float4(__fadd_rn(a.x, bprev), __fadd_rn(a.y, rmax_prev(b)), a.z, a.w);
      aprev = rmin_next(a);
      ++ia;
      if (ia < sa.size) a = sa.d[ia];
    } else {
      r = make_float4(__fadd_rn(b.x, aprev), __fadd_rn(b.y, rmax_prev(a)), b.z, b.w);
      bprev = rmin_next(b);
      ++ib;
      if (ib < sb.size) b = sb.d[ib];
    }
    dst.d[o++] = r;
  }
  if (ia < sa.size) {
    float brmax = sb.d[sb.size - 1].y;
    for (; ia < sa.size; ++ia) { float4 e = sa.d[ia]; dst.d[o++] = make_float4(__fadd_rn(e.x, bprev), __fadd_rn(e.y, brmax), e.z, e.w); }
  }
  if (ib < sb.size) {
    float armax = sa.d[sa.size - 1].y;
    for (; ib < sb.size; ++ib) { float4 e = sb.d[ib]; dst.d[o++] = make_float4(__fadd_rn(e.x, aprev), __fadd_rn(e.y, armax), e.z, e.w); }
  }
  dst.size = o;
  float prev_rmin = 0.f, prev_rmax = 0.f;   // FixError
  for (int i = 0; i < o; ++i) {
    float4 e = dst.d[i];
    bool dirty = false;
    if (e.x < prev_rmin) { e.x = prev_rmin; dirty = true; } else prev_rmin = e.x;
    if (e.y < prev_rmax) { e.y = prev_rmax; dirty = true; }
    float rn = __fadd_rn(e.x, e.z);
    if (e.y < rn) { e.y = rn; dirty = true; }
    prev_rmax = e.y;
    if (dirty) dst.d[i] = e;
  }
}

__device__ __forceinline__ uint32_t ordered_bits(float v) {
  uint32_t b = __float_as_uint(v);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float unordered_bits(uint32_t o) {
  uint32_t b = (o & 0x80000000u) ? (o & 0x7fffffffu) : ~o;
  return __uint_as_float(b);
}

// Dynamic shared memory of k_sketch_cuts: one region, used by the sort (two padded blocks of keys) or as a table of
// search keys.
constexpr int kSortBlock = 8192;                        // keys sorted in shared memory at a time: 8 per thread
constexpr int kSortPer = kSortBlock / kSketchThreads;
constexpr int kSortPadded = kSortBlock + kSortBlock / 8;   // one idle slot per 8 keys: neighbouring threads' windows fall on different banks
constexpr int kTabCap = 43008;                          // floats of a search table (168 KiB): 2 x limit_size for every n < 2^24
constexpr int kSketchSmem = kTabCap * 4;
static_assert(2 * kSortPadded * 8 <= kSketchSmem, "the sort's two blocks must fit the dynamic shared memory");
static_assert(kSortPer == 8, "the register stage sorts 8 keys per thread");

__device__ __forceinline__ int sort_slot(int i) { return i + (i >> 3); }

// Merge path: how many of the first `diag` outputs of merge(A, B) come from A (equal keys are identical, so the
// tie rule is immaterial).  KEY(array, index) reads one key.
#define GT_MERGE_PATH(A_AT, B_AT, lenA, lenB, diag, result)                              \
  do {                                                                                    \
    int lo__ = max(0, (diag) - (lenB)), hi__ = min((diag), (lenA));                       \
    while (lo__ < hi__) {                                                                 \
      const int mid__ = (lo__ + hi__) >> 1;                                               \
      if (A_AT(mid__) <= B_AT((diag) - 1 - mid__)) lo__ = mid__ + 1; else hi__ = mid__;   \
    }                                                                                     \
    (result) = lo__;                                                                      \
  } while (0)

// CTA-wide sort of `count` 64-bit keys in global memory (every key below ~0).  Merge sort: blocks of 8192 keys are
// sorted in shared memory (8 keys per thread in registers, then ten rounds of pairwise run merges, each thread
// producing 8 consecutive outputs found by merge path), then the sorted blocks are merged pairwise in global memory,
// ping-ponging between keys and scratch (count keys of room).  The result ends in keys.
__device__ void cta_sort(unsigned long long* keys, int count, unsigned long long* scratch, unsigned long long* sm) {
  const int tid = threadIdx.x;
  unsigned long long* buf0 = sm;
  unsigned long long* buf1 = sm + kSortPadded;
  for (int base = 0; base < count; base += kSortBlock) {
    const int real = min(kSortBlock, count - base);
    unsigned long long r[kSortPer];
#pragma unroll
    for (int q = 0; q < kSortPer; ++q) {                 // thread t: keys t, t + 1024, ... (coalesced), any order will do
      const int i = q * kSketchThreads + tid;
      r[q] = i < real ? keys[base + i] : ~0ull;
    }
#pragma unroll
    for (int round = 0; round < kSortPer; ++round) {     // odd-even transposition sort of the thread's 8 keys
#pragma unroll
      for (int q = round & 1; q + 1 < kSortPer; q += 2) {
        const unsigned long long lo = r[q] < r[q + 1] ? r[q] : r[q + 1], hi = r[q] < r[q + 1] ? r[q + 1] : r[q];
        r[q] = lo; r[q + 1] = hi;
      }
    }
#pragma unroll
    for (int q = 0; q < kSortPer; ++q) buf0[sort_slot(tid * kSortPer + q)] = r[q];
    __syncthreads();
    unsigned long long* src = buf0;
    unsigned long long* dst = buf1;
    for (int W = kSortPer; W < kSortBlock; W <<= 1) {
      const int o0 = tid * kSortPer;
      const int pair0 = o0 & ~(2 * W - 1), diag = o0 - pair0;
      const unsigned long long* A = src;
#define A_AT(i) A[sort_slot(pair0 + (i))]
#define B_AT(i) A[sort_slot(pair0 + W + (i))]
      int ai;
      GT_MERGE_PATH(A_AT, B_AT, W, W, diag, ai);
      int bi = diag - ai;
      unsigned long long a = ai < W ? A_AT(ai) : ~0ull, b = bi < W ? B_AT(bi) : ~0ull;
#pragma unroll
      for (int q = 0; q < kSortPer; ++q) {
        const bool take_a = bi >= W || (ai < W && a <= b);
        r[q] = take_a ? a : b;
        if (take_a) { ++ai; a = ai < W ? A_AT(ai) : ~0ull; } else { ++bi; b = bi < W ? B_AT(bi) : ~0ull; }
      }
#undef A_AT
#undef B_AT
#pragma unroll
      for (int q = 0; q < kSortPer; ++q) dst[sort_slot(o0 + q)] = r[q];
      __syncthreads();
      unsigned long long* t = src; src = dst; dst = t;
    }
    for (int i = tid; i < real; i += kSketchThreads) keys[base + i] = src[sort_slot(i)];
    __syncthreads();
  }
  // sorted blocks -> one sorted array: pairwise merges of runs of W keys (the last run may be shorter, or alone)
  unsigned long long* from = keys;
  unsigned long long* to = scratch;
  constexpr int S = 32;                                  // outputs per thread and step; divides every run length
  for (int W = kSortBlock; W < count; W <<= 1) {
    const int nseg = (count + S - 1) / S;
    for (int sg = tid; sg < nseg; sg += kSketchThreads) {
      const int o0 = sg * S;
      const int pair0 = (o0 / (2 * W)) * (2 * W), diag = o0 - pair0;
      const int lenA = min(W, count - pair0), lenB = max(0, min(W, count - pair0 - W));
      const unsigned long long* A = from + pair0;
      const unsigned long long* B = A + W;
#define A_AT(i) A[i]
#define B_AT(i) B[i]
      int ai;
      GT_MERGE_PATH(A_AT, B_AT, lenA, lenB, diag, ai);
#undef A_AT
#undef B_AT
      int bi = diag - ai;
      const int nout = min(S, count - o0);
      unsigned long long a = ai < lenA ? A[ai] : ~0ull, b = bi < lenB ? B[bi] : ~0ull;
      for (int q = 0; q < nout; ++q) {
        const bool take_a = bi >= lenB || (ai < lenA && a <= b);
        to[o0 + q] = take_a ? a : b;
        if (take_a) { ++ai; a = ai < lenA ? A[ai] : ~0ull; } else { ++bi; b = bi < lenB ? B[bi] : ~0ull; }
      }
    }
    __syncthreads();
    unsigned long long* t = from; from = to; to = t;
  }
  if (from != keys) {
    for (int i = tid; i < count; i += kSketchThreads) keys[i] = from[i];
    __syncthreads();
  }
}

// Queue::MakeSummary on sorted keys (value in the high word, weight in the low word)
__device__ void make_summary(const unsigned long long* keys, int qtail, Summary& out) {
  int o = 0;
  float wsum = 0.f;
  int i = 0;
  while (i < qtail) {
    unsigned long long kk = keys[i];
    float v = unordered_bits((uint32_t)(kk >> 32));
    float w = (float)(uint32_t)kk;
    int j = i + 1;
    while (j < qtail) {
      unsigned long long k2 = keys[j];
      float v2 = unordered_bits((uint32_t)(k2 >> 32));
      if (!(v2 == v)) break;
      w = __fadd_rn(w, (float)(uint32_t)k2);
      ++j;
    }
    out.d[o++] = make_float4(wsum, __fadd_rn(wsum, w), w, v);
    wsum = __fadd_rn(wsum, w);
    i = j;
  }
  out.size = o;
}

// ---------------------------------------------------------------------------------------------------
// CTA-cooperative versions of the summary operations.  Every thread of the CTA calls them with the same
// arguments; they end with a barrier.  Ranks are sums of unit weights, i.e. integers below 2^24 (the
// caller falls back to the one-thread versions beyond that), so float rank arithmetic is exact and the
// sequential pointer walks of SetPrune / SetCombine can be replaced by binary searches.
struct CtaScratch {
  int warp_tot[32];
  int warp_tot2[32];
  int total, total2;
  int flag;
  float* tab;        // kTabCap floats of dynamic shared memory: the search keys of the summary being searched
  int cnt[2][24 * 32];   // per (row of 1024 elements, warp) counts of the interleaved passes, then their exclusive prefixes
};
constexpr int kRowsMax = 24;   // rows of 1024 elements the interleaved passes handle (20 for every n < 2^24); longer lists take the blocked versions

// exclusive scan of one int per thread over the CTA (blockDim.x == 1024); *total gets the CTA sum
__device__ __forceinline__ int cta_excl_scan(int v, int* warp_tot, int* total) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  int inc = v;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) { int t = __shfl_up_sync(0xffffffffu, inc, d); if (lane >= d) inc += t; }
  __syncthreads();   // protects warp_tot from the previous use
  if (lane == 31) warp_tot[wid] = inc;
  __syncthreads();
  if (wid == 0) {
    int t = warp_tot[lane];
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { int u = __shfl_up_sync(0xffffffffu, t, d); if (lane >= d) t += u; }
    warp_tot[lane] = t;
  }
  __syncthreads();
  const int base = wid ? warp_tot[wid - 1] : 0;
  *total = warp_tot[31];
  return base + inc - v;
}

__device__ void p_copy(float4* dst, const float4* src, int n) {
  for (int i = threadIdx.x; i < n; i += blockDim.x) dst[i] = src[i];
  __syncthreads();
}

// Queue::MakeSummary on sorted keys: one entry per distinct value, ranks from the weight prefix.  Every warp takes a
// contiguous run of the keys 32 at a time (coalesced); starts of distinct values and weights are prefix-summed inside
// the warp, the warps' totals across the CTA.
__device__ int p_make_summary(const unsigned long long* keys, int qtail, float4* out, CtaScratch& sc) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = (int)blockDim.x >> 5;
  const int chunk = ((qtail + nw - 1) / nw + 31) / 32 * 32;
  const int lo = min(qtail, wid * chunk), hi = min(qtail, lo + chunk);
  int starts = 0, wsum = 0;                              // warp-uniform
  for (int i0 = lo; i0 < hi; i0 += 32) {
    const int i = i0 + lane;
    const bool in = i < hi;
    const unsigned long long k = in ? keys[i] : 0ull;
    const unsigned long long kp = in && i > 0 ? keys[i - 1] : 0ull;
    const bool st = in && (i == 0 || !(unordered_bits((uint32_t)(kp >> 32)) == unordered_bits((uint32_t)(k >> 32))));
    starts += __popc(__ballot_sync(0xffffffffu, st));
    wsum += __reduce_add_sync(0xffffffffu, in ? (int)(uint32_t)k : 0);
  }
  int tot_g, tot_w;
  int g = cta_excl_scan(lane == 0 ? starts : 0, sc.warp_tot, &tot_g);
  int w = cta_excl_scan(lane == 0 ? wsum : 0, sc.warp_tot2, &tot_w);
  g = __shfl_sync(0xffffffffu, g, 0); w = __shfl_sync(0xffffffffu, w, 0);
  for (int i0 = lo; i0 < hi; i0 += 32) {
    const int i = i0 + lane;
    const bool in = i < hi;
    const unsigned long long k = in ? keys[i] : 0ull;
    const unsigned long long kp = in && i > 0 ? keys[i - 1] : 0ull;
    const unsigned long long kn = in && i + 1 < qtail ? keys[i + 1] : 0ull;
    const float v = unordered_bits((uint32_t)(k >> 32));
    const bool st = in && (i == 0 || !(unordered_bits((uint32_t)(kp >> 32)) == v));
    const bool en = in && (i == qtail - 1 || !(unordered_bits((uint32_t)(kn >> 32)) == v));
    const int wt = in ? (int)(uint32_t)k : 0;
    int incw = wt;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { const int t = __shfl_up_sync(0xffffffffu, incw, d); if (lane >= d) incw += t; }
    const unsigned bal = __ballot_sync(0xffffffffu, st);
    const int idx = g + __popc(bal & (0xffffffffu >> (31 - lane))) - 1;    // the distinct value this key belongs to
    if (st) { out[idx].x = (float)(w + incw - wt); out[idx].w = v; }
    if (en) out[idx].y = (float)(w + incw);
    g += __popc(bal);
    w += __shfl_sync(0xffffffffu, incw, 31);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < tot_g; i += blockDim.x) { float4 e = out[i]; e.z = __fsub_rn(e.y, e.x); out[i] = e; }
  __syncthreads();
  return tot_g;
}

// WQSummary::SetPrune.  For query k the sequential walk stops at the first i with dx2 < rmin[i+1] + rmax[i+1];
// both dx2 (in k) and that sum (in i) are non-decreasing, so i(k) is an upper-bound search.
__device__ int p_prune_blocked(float4* dst, const float4* src, int m, int maxsize, CtaScratch& sc) {
  if (m <= maxsize) { p_copy(dst, src, m); return m; }
  const float begin = src[0].y;
  const float range = __fsub_rn(src[m - 1].x, src[0].y);
  const int n = maxsize - 1;
  const float fn = (float)n;
  const int nq = n - 1;                                  // queries k = 1 .. n-1
  const int per = (nq + (int)blockDim.x - 1) / (int)blockDim.x;
  const int klo = 1 + min(nq, (int)threadIdx.x * per), khi = min(n, klo + per);
  // the keys the walk compares with, rmin + rmax of every entry, once into shared memory: a query is then a binary
  // search there instead of ~15 dependent loads from global memory
  const bool tabbed = m <= kTabCap;
  float* tab = sc.tab;
  if (tabbed) {
    __syncthreads();                                     // the table's previous user is done
    for (int i = threadIdx.x; i < m; i += blockDim.x) { const float4 e = src[i]; tab[i] = __fadd_rn(e.y, e.x); }
    __syncthreads();
  }
  auto choose = [&](int k) -> int {                      // chosen source index, or -1 once the walk has hit the end
    const float dx2 = __fmul_rn(2.0f, __fadd_rn(__fdiv_rn(__fmul_rn((float)k, range), fn), begin));
    int lo = 2, hi = m;                                  // count j in [2, m-1] with rmin[j] + rmax[j] <= dx2
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      float key;
      if (tabbed) key = tab[mid]; else { const float4 e = src[mid]; key = __fadd_rn(e.y, e.x); }
      if (dx2 >= key) lo = mid + 1; else hi = mid;
    }
    const int i = lo - 1;
    if (i >= m - 1) return -1;
    const float4 a = src[i], b = src[i + 1];
    return dx2 < __fadd_rn(__fadd_rn(a.x, a.z), __fsub_rn(b.y, b.z)) ? i : i + 1;
  };
  const int prev0 = klo > 1 && klo < n ? choose(klo - 1) : 0;   // lastidx seen by this thread's first query
  constexpr int kKeep = 24;                              // chosen indices kept from the counting pass (20 queries per thread at most for n < 2^24)
  int kept[kKeep];
  int prev = prev0, emits = 0;
  for (int k = klo; k < khi; ++k) {
    const int c = choose(k);
    if (k - klo < kKeep) kept[k - klo] = c;
    if (c < 0) break;
    emits += (c != prev);
    prev = c;
  }
  int total;
  int pos = 1 + cta_excl_scan(emits, sc.warp_tot, &total);
  prev = prev0;
  int last_c = -2;                                       // chosen index of the last valid query in this range
  for (int k = klo; k < khi; ++k) {
    const int c = k - klo < kKeep ? kept[k - klo] : choose(k);
    if (c < 0) break;
    if (c != prev) dst[pos++] = src[c];
    prev = c;
    last_c = c;
  }
  if (threadIdx.x == 0) { dst[0] = src[0]; sc.flag = 0; }
  __syncthreads();
  if (last_c >= 0) atomicMax(&sc.flag, last_c);          // chosen indices never decrease: the max is the final lastidx
  __syncthreads();
  int size = 1 + total;
  if (sc.flag != m - 1) { if (threadIdx.x == 0) dst[size] = src[m - 1]; ++size; }
  __syncthreads();
  return size;
}

// first index whose value is not below v; the values come from the shared-memory table when there is one
__device__ __forceinline__ int lower_bound_value(const float4* s, const float* tab, int n, float v) {
  int lo = 0, hi = n;
  while (lo < hi) { const int mid = (lo + hi) >> 1; if ((tab ? tab[mid] : s[mid].w) < v) lo = mid + 1; else hi = mid; }
  return lo;
}

// WQSummary::SetCombine: a merge by value in which equal values fuse; every output entry depends only on its
// own element and on its neighbours in the other list.  Returns the size, or -1 if FixError would have had to
// change something (impossible with integer ranks; the caller then redoes the combine with one thread).
__device__ int p_combine_blocked(float4* dst, const float4* A, int na, const float4* B, int nb, CtaScratch& sc) {
  if (na == 0) { p_copy(dst, B, nb); return nb; }
  if (nb == 0) { p_copy(dst, A, na); return na; }
  const int T = (int)blockDim.x;
  const int pa = (na + T - 1) / T, alo = min(na, (int)threadIdx.x * pa), ahi = min(na, alo + pa);
  const int pb = (nb + T - 1) / T, blo = min(nb, (int)threadIdx.x * pb), bhi = min(nb, blo + pb);
  // both lists' values once into shared memory: the searches below run there
  const float* tA = nullptr;
  const float* tB = nullptr;
  if (na + nb <= kTabCap) {
    __syncthreads();                                     // the table's previous user is done
    for (int i = threadIdx.x; i < na; i += T) sc.tab[i] = A[i].w;
    for (int i = threadIdx.x; i < nb; i += T) sc.tab[na + i] = B[i].w;
    __syncthreads();
    tA = sc.tab; tB = sc.tab + na;
  }
  constexpr int kKeep = 24;                              // search results kept from the counting pass (20 elements per thread and list at most for n < 2^24)
  int ja[kKeep], ib[kKeep];
  int ea = 0, eb = 0;
  for (int i = alo; i < ahi; ++i) {
    const float v = tA ? tA[i] : A[i].w;
    const int j = lower_bound_value(B, tB, nb, v);
    if (i - alo < kKeep) ja[i - alo] = j;
    ea += (j < nb && (tB ? tB[j] : B[j].w) == v);
  }
  for (int j = blo; j < bhi; ++j) {
    const float v = tB ? tB[j] : B[j].w;
    const int i = lower_bound_value(A, tA, na, v);
    if (j - blo < kKeep) ib[j - blo] = i;
    eb += (i < na && (tA ? tA[i] : A[i].w) == v);
  }
  int tot_a, tot_b;
  int Ea = cta_excl_scan(ea, sc.warp_tot, &tot_a);
  int Eb = cta_excl_scan(eb, sc.warp_tot2, &tot_b);
  const float armax = A[na - 1].y, brmax = B[nb - 1].y;
  for (int i = alo; i < ahi; ++i) {
    const float4 a = A[i];
    const int j = i - alo < kKeep ? ja[i - alo] : lower_bound_value(B, tB, nb, a.w);
    const bool eq = j < nb && (tB ? tB[j] : B[j].w) == a.w;
    float4 r;
    if (eq) {
      const float4 b = B[j];
      r = make_float4(__fadd_rn(a.x, b.x), __fadd_rn(a.y, b.y), __fadd_rn(a.z, b.z), a.w);
    } else {
      const float bprev = j > 0 ? __fadd_rn(B[j - 1].x, B[j - 1].z) : 0.f;               // RMinNext of the last smaller b
      const float bnext = j < nb ? __fsub_rn(B[j].y, B[j].z) : brmax;                     // RMaxPrev of the next b, or b's last rmax
      r = make_float4(__fadd_rn(a.x, bprev), __fadd_rn(a.y, bnext), a.z, a.w);
    }
    dst[i + j - Ea] = r;
    Ea += eq;
  }
  for (int j = blo; j < bhi; ++j) {
    const float4 b = B[j];
    const int i = j - blo < kKeep ? ib[j - blo] : lower_bound_value(A, tA, na, b.w);
    const bool eq = i < na && (tA ? tA[i] : A[i].w) == b.w;
    if (!eq) {
      const float aprev = i > 0 ? __fadd_rn(A[i - 1].x, A[i - 1].z) : 0.f;
      const float anext = i < na ? __fsub_rn(A[i].y, A[i].z) : armax;
      dst[j + i - Eb] = make_float4(__fadd_rn(b.x, aprev), __fadd_rn(b.y, anext), b.z, b.w);
    }
    Eb += eq;
  }
  if (threadIdx.x == 0) sc.flag = 0;
  __syncthreads();
  const int size = na + nb - tot_a;
  int bad = 0;                                                                            // FixError would act iff one of these holds
  for (int i = threadIdx.x; i < size; i += T) {
    const float4 e = dst[i];
    if (e.y < __fadd_rn(e.x, e.z)) bad = 1;
    if (i > 0) { const float4 p = dst[i - 1]; if (e.x < p.x || e.y < p.y) bad = 1; }
  }
  if (bad) sc.flag = 1;
  __syncthreads();
  const int res = sc.flag ? -1 : size;
  __syncthreads();
  return res;
}

// exclusive prefix over the per-(row, warp) counts in sc.cnt[which][0 .. rows * 32), in place; returns the total
__device__ __forceinline__ int scan_row_counts(CtaScratch& sc, int which, int rows, int* warp_tot) {
  __syncthreads();
  const int t = threadIdx.x;
  const int v = t < rows * 32 ? sc.cnt[which][t] : 0;
  int total;
  const int ex = cta_excl_scan(v, warp_tot, &total);
  if (t < rows * 32) sc.cnt[which][t] = ex;
  __syncthreads();
  return total;
}

// SetPrune with the queries dealt to the threads round-robin (query k = 1 + row * 1024 + thread): the entries that
// neighbouring lanes read and the outputs they write are neighbours in memory.  Same result as p_prune_blocked.
__device__ int p_prune(float4* dst, const float4* src, int m, int maxsize, CtaScratch& sc) {
  if (m <= maxsize) { p_copy(dst, src, m); return m; }
  const int T = (int)blockDim.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int n = maxsize - 1, nq = n - 1;
  const int rows = (nq + T - 1) / T;
  if (rows > kRowsMax || m > kTabCap || T != 1024) return p_prune_blocked(dst, src, m, maxsize, sc);
  const float begin = src[0].y;
  const float range = __fsub_rn(src[m - 1].x, src[0].y);
  const float fn = (float)n;
  float* tab = sc.tab;
  __syncthreads();                                       // the table's previous user is done
  for (int i = tid; i < m; i += T) { const float4 e = src[i]; tab[i] = __fadd_rn(e.y, e.x); }
  __syncthreads();
  auto choose = [&](int k) -> int {                      // chosen source index, or -1 once the walk has hit the end
    const float dx2 = __fmul_rn(2.0f, __fadd_rn(__fdiv_rn(__fmul_rn((float)k, range), fn), begin));
    int lo = 2, hi = m;                                  // count j in [2, m-1] with rmin[j] + rmax[j] <= dx2
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if (dx2 >= tab[mid]) lo = mid + 1; else hi = mid;
    }
    const int i = lo - 1;
    if (i >= m - 1) return -1;
    const float4 a = src[i], b = src[i + 1];
    return dx2 < __fadd_rn(__fadd_rn(a.x, a.z), __fsub_rn(b.y, b.z)) ? i : i + 1;
  };
  // query k is emitted iff its choice differs from query k-1's (lastidx starts at 0); choices never decrease and
  // stay -1 once the walk has ended
  int kept[kRowsMax];
  unsigned emask = 0u;
  int last_c = -2;
#pragma unroll 1
  for (int q = 0; q < rows; ++q) {
    const int k = 1 + q * T + tid;
    const int c = k < n ? choose(k) : -1;
    kept[q] = c;
    int pc = __shfl_up_sync(0xffffffffu, c, 1);
    if (lane == 0) pc = k == 1 ? 0 : (k - 1 < n ? choose(k - 1) : -1);
    const bool emit = c >= 0 && c != pc;
    const unsigned bal = __ballot_sync(0xffffffffu, emit);
    if (lane == 0) sc.cnt[0][q * 32 + wid] = __popc(bal);
    emask |= (emit ? 1u : 0u) << q;
    if (c >= 0) last_c = c;
  }
  const int total = scan_row_counts(sc, 0, rows, sc.warp_tot);
#pragma unroll 1
  for (int q = 0; q < rows; ++q) {
    const bool emit = (emask >> q) & 1u;
    const unsigned bal = __ballot_sync(0xffffffffu, emit);
    if (emit) dst[1 + sc.cnt[0][q * 32 + wid] + __popc(bal & ((1u << lane) - 1u))] = src[kept[q]];
  }
  if (tid == 0) { dst[0] = src[0]; sc.flag = 0; }
  __syncthreads();
  if (last_c >= 0) atomicMax(&sc.flag, last_c);          // chosen indices never decrease: the max is the final lastidx
  __syncthreads();
  int size = 1 + total;
  if (sc.flag != m - 1) { if (tid == 0) dst[size] = src[m - 1]; ++size; }
  __syncthreads();
  return size;
}

// SetCombine with the elements dealt to the threads round-robin (element = row * 1024 + thread), see p_prune.
__device__ int p_combine(float4* dst, const float4* A, int na, const float4* B, int nb, CtaScratch& sc) {
  if (na == 0) { p_copy(dst, B, nb); return nb; }
  if (nb == 0) { p_copy(dst, A, na); return na; }
  const int T = (int)blockDim.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int rows_a = (na + T - 1) / T, rows_b = (nb + T - 1) / T;
  if (rows_a > kRowsMax || rows_b > kRowsMax || na + nb > kTabCap || T != 1024) return p_combine_blocked(dst, A, na, B, nb, sc);
  __syncthreads();                                       // the table's previous user is done
  for (int i = tid; i < na; i += T) sc.tab[i] = A[i].w;
  for (int i = tid; i < nb; i += T) sc.tab[na + i] = B[i].w;
  __syncthreads();
  const float* tA = sc.tab;
  const float* tB = sc.tab + na;
  int ja[kRowsMax], ib[kRowsMax];
  unsigned eqa = 0u, eqb = 0u;
#pragma unroll 1
  for (int q = 0; q < rows_a; ++q) {
    const int i = q * T + tid;
    int j = 0; bool eq = false;
    if (i < na) { const float v = tA[i]; j = lower_bound_value(B, tB, nb, v); eq = j < nb && tB[j] == v; }
    ja[q] = j;
    const unsigned bal = __ballot_sync(0xffffffffu, eq);
    if (lane == 0) sc.cnt[0][q * 32 + wid] = __popc(bal);
    eqa |= (eq ? 1u : 0u) << q;
  }
#pragma unroll 1
  for (int q = 0; q < rows_b; ++q) {
    const int j = q * T + tid;
    int i = 0; bool eq = false;
    if (j < nb) { const float v = tB[j]; i = lower_bound_value(A, tA, na, v); eq = i < na && tA[i] == v; }
    ib[q] = i;
    const unsigned bal = __ballot_sync(0xffffffffu, eq);
    if (lane == 0) sc.cnt[1][q * 32 + wid] = __popc(bal);
    eqb |= (eq ? 1u : 0u) << q;
  }
  const int tot_a = scan_row_counts(sc, 0, rows_a, sc.warp_tot);
  scan_row_counts(sc, 1, rows_b, sc.warp_tot2);
  const float armax = A[na - 1].y, brmax = B[nb - 1].y;
#pragma unroll 1
  for (int q = 0; q < rows_a; ++q) {
    const int i = q * T + tid;
    const bool eq = (eqa >> q) & 1u;
    const unsigned bal = __ballot_sync(0xffffffffu, eq);
    if (i < na) {
      const int Ea = sc.cnt[0][q * 32 + wid] + __popc(bal & ((1u << lane) - 1u));   // equal pairs among A's earlier elements
      const float4 a = A[i];
      const int j = ja[q];
      float4 r;
      if (eq) {
        const float4 b = B[j];
        r = make_float4(__fadd_rn(a.x, b.x), __fadd_rn(a.y, b.y), __fadd_rn(a.z, b.z), a.w);
      } else {
        const float bprev = j > 0 ? __fadd_rn(B[j - 1].x, B[j - 1].z) : 0.f;               // RMinNext of the last smaller b
        const float bnext = j < nb ? __fsub_rn(B[j].y, B[j].z) : brmax;                     // RMaxPrev of the next b, or b's last rmax
        r = make_float4(__fadd_rn(a.x, bprev), __fadd_rn(a.y, bnext), a.z, a.w);
      }
      dst[i + j - Ea] = r;
    }
  }
#pragma unroll 1
  for (int q = 0; q < rows_b; ++q) {
    const int j = q * T + tid;
    const bool eq = (eqb >> q) & 1u;
    const unsigned bal = __ballot_sync(0xffffffffu, eq);
    if (j < nb && !eq) {
      const int Eb = sc.cnt[1][q * 32 + wid] + __popc(bal & ((1u << lane) - 1u));
      const float4 b = B[j];
      const int i = ib[q];
      const float aprev = i > 0 ? __fadd_rn(A[i - 1].x, A[i - 1].z) : 0.f;
      const float anext = i < na ? __fsub_rn(A[i].y, A[i].z) : armax;
      dst[j + i - Eb] = make_float4(__fadd_rn(b.x, aprev), __fadd_rn(b.y, anext), b.z, b.w);
    }
  }
  if (tid == 0) sc.flag = 0;
  __syncthreads();
  const int size = na + nb - tot_a;
  int bad = 0;                                                                            // FixError would act iff one of these holds
  for (int i = tid; i < size; i += T) {
    const float4 e = dst[i];
    if (e.y < __fadd_rn(e.x, e.z)) bad = 1;
    if (i > 0) { const float4 p = dst[i - 1]; if (e.x < p.x || e.y < p.y) bad = 1; }
  }
  if (bad) sc.flag = 1;
  __syncthreads();
  const int res = sc.flag ? -1 : size;
  __syncthreads();
  return res;
}

__global__ void __launch_bounds__(kSketchThreads)
k_sketch_cuts(const float* __restrict__ X, long long ldx, int f, int nfold,
              const uint32_t* const* __restrict__ fold_rows, const long long* __restrict__ fold_ntrain,
              SketchPlan plan, const int* __restrict__ fold_limit, unsigned long long* __restrict__ key_arena,
              float4* __restrict__ entry_arena,
              float* __restrict__ cuts_out, int* __restrict__ ncuts_out, int stream0) {
  const int stream = blockIdx.x + stream0;
  const int k = stream / f, j = stream % f;
  const long long n = fold_ntrain[k];
  const uint32_t* rows = fold_rows[k];
  const int L = fold_limit[k], qcap = 2 * L;
  unsigned long long* keys = key_arena + (size_t)blockIdx.x * plan.keys_per_stream;
  float4* ent = entry_arena + (size_t)blockIdx.x * plan.entries_per_stream;
  const int buf_cap = plan.buf_cap;   // >= 2 * L + 2 for every fold
  Summary temp = { ent, 0 }, out = { ent + buf_cap, 0 }, red = { ent + 2 * buf_cap, 0 };
  float4* level_space = ent + 3 * buf_cap;
  float* cuts = cuts_out + (size_t)stream * kBins;

  extern __shared__ __align__(16) unsigned char s_dyn[];   // kSketchSmem bytes: sort block / search table
  unsigned long long* sk = reinterpret_cast<unsigned long long*>(s_dyn);
  __shared__ float sv[kSketchThreads];
  __shared__ int warp_tot[kSketchThreads / 32];
  __shared__ int level_size[40];
  __shared__ int levels_in_use;
  __shared__ CtaScratch sc;
  const bool parallel_ok = n < (1ll << 24);   // ranks stay exact integers in fp32
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;

  if (tid == 0) { levels_in_use = 0; for (int l = 0; l < 40; ++l) level_size[l] = 0; sc.tab = reinterpret_cast<float*>(s_dyn); }
  for (int i = tid; i < plan.keys_per_stream; i += blockDim.x) keys[i] = 0ull;
  __syncthreads();

  if (n == 0) {   // empty column: MakeCuts' "Empty column" branch
    if (tid == 0) { cuts[0] = __fadd_rn(1e-5f, __fadd_rn(fabsf(1e-5f), 1e-5f)); ncuts_out[stream] = 1; }
    return;
  }

  long long pos = 0;
  long long runs_done = 0, chunk_base = 0;
  float lastv = 0.f;
  SK_PROF_DECL;
  while (pos < n) {
    // ---- queue fill: QuantileSketchTemplate::Push as a run-length scan over one tile of rows
    long long i = pos + tid;
    bool have = i < n;
    float v = have ? X[(size_t)rows[i] * (size_t)ldx + j] : 0.f;
    sv[tid] = v;
    __syncthreads();
    float prev = tid ? sv[tid - 1] : lastv;
    int flag = have && (i == 0 || v != prev);
    int inc = flag;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) { int t = __shfl_up_sync(0xffffffffu, inc, d); if (lane >= d) inc += t; }
    if (lane == 31) warp_tot[wid] = inc;
    __syncthreads();
    if (wid == 0) {
      int t = warp_tot[lane];
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) { int u = __shfl_up_sync(0xffffffffu, t, d); if (lane >= d) t += u; }
      warp_tot[lane] = t;
    }
    __syncthreads();
    if (wid) inc += warp_tot[wid - 1];
    long long slot = runs_done + inc - 1 - chunk_base;
    bool in_chunk = have && slot < qcap;
    if (in_chunk) atomicAdd(&keys[slot], (flag ? ((unsigned long long)ordered_bits(v) << 32) : 0ull) + 1ull);
    int consumed = __syncthreads_count(in_chunk);
    int flags_consumed = __syncthreads_count(in_chunk && flag);
    int overflow = __syncthreads_or(have && !in_chunk);
    if (consumed > 0) lastv = sv[consumed - 1];
    pos += consumed;
    runs_done += flags_consumed;
    __syncthreads();   // sv is rewritten by the next tile
    if (overflow) {
      SK_PROF(0);
      // ---- queue full and the next value differs: MakeSummary + PushTemp
      cta_sort(keys, qcap, reinterpret_cast<unsigned long long*>(red.d), sk);
      SK_PROF(1);
      if (parallel_ok) {
        temp.size = p_make_summary(keys, qcap, temp.d, sc);
        SK_PROF(2);
        for (int l = 1;; ++l) {   // QuantileSketchTemplate::PushTemp, all threads in step
          if (tid == 0 && l + 1 > levels_in_use) levels_in_use = l + 1;
          float4* lvd = level_space + (size_t)l * L;
          const int lsz = level_size[l];
          __syncthreads();
          if (lsz == 0) {
            const int sz = p_prune(lvd, temp.d, temp.size, L, sc);
            if (tid == 0) level_size[l] = sz;
            __syncthreads();
            SK_PROF(3);
            break;
          }
          const int l0sz = p_prune(level_space, temp.d, temp.size, L, sc);
          SK_PROF(3);
          int csz = p_combine(temp.d, level_space, l0sz, lvd, lsz, sc);
          SK_PROF(4);
          if (csz < 0) {   // never with integer ranks; keep the reference behaviour anyway
            if (tid == 0) { Summary a0 = { level_space, l0sz }, b0 = { lvd, lsz }; s_combine(temp, a0, b0); sc.total = temp.size; }
            __syncthreads();
            csz = sc.total;
            __syncthreads();
          }
          temp.size = csz;
          if (csz > L) {
            if (tid == 0) level_size[l] = 0;
            __syncthreads();
          } else {
            p_copy(lvd, temp.d, csz);
            if (tid == 0) level_size[l] = csz;
            __syncthreads();
            break;
          }
        }
      } else if (tid == 0) {
        make_summary(keys, qcap, temp);
        for (int l = 1;; ++l) {   // QuantileSketchTemplate::PushTemp
          if (l + 1 > levels_in_use) levels_in_use = l + 1;
          Summary lv = { level_space + (size_t)l * L, level_size[l] };
          if (lv.size == 0) {
            s_prune(lv, temp, L);
            level_size[l] = lv.size;
            break;
          }
          Summary l0 = { level_space, 0 };
          s_prune(l0, temp, L);
          s_combine(temp, l0, lv);
          if (temp.size > L) {
            level_size[l] = 0;
          } else {
            s_copy(lv, temp);
            level_size[l] = lv.size;
            break;
          }
        }
      }
      __syncthreads();
      for (int t = tid; t < qcap; t += blockDim.x) keys[t] = 0ull;     // the queue's slots, ready for the next fill
      chunk_base += qcap;
      __syncthreads();
      SK_PROF(5);
    }
  }
  SK_PROF(0);
  // ---- QuantileSketchTemplate::GetSummary, SketchContainerImpl::AllReduce (one worker), MakeCuts
  const int qtail = (int)(runs_done - chunk_base);
  cta_sort(keys, qtail, reinterpret_cast<unsigned long long*>(red.d), sk);
  if (parallel_ok) {
    out.size = p_make_summary(keys, qtail, out.d, sc);
    const int used = levels_in_use;
    __syncthreads();
    if (used != 0) {
      int l0sz = p_prune(level_space, out.d, out.size, L, sc);
      for (int l = 1; l < used; ++l) {
        const int lsz = level_size[l];
        if (lsz == 0) continue;
        float4* lvd = level_space + (size_t)l * L;
        if (l0sz == 0) {
          p_copy(level_space, lvd, lsz);
          l0sz = lsz;
        } else {
          int csz = p_combine(out.d, level_space, l0sz, lvd, lsz, sc);
          if (csz < 0) {
            if (tid == 0) { Summary a0 = { level_space, l0sz }, b0 = { lvd, lsz }; s_combine(out, a0, b0); sc.total = out.size; }
            __syncthreads();
            csz = sc.total;
            __syncthreads();
          }
          l0sz = p_prune(level_space, out.d, csz, L, sc);
        }
      }
      p_copy(out.d, level_space, l0sz);
      out.size = l0sz;
    } else if (out.size > L) {
      const int sz = p_prune(temp.d, out.d, out.size, L, sc);
      p_copy(out.d, temp.d, sz);
      out.size = sz;
    }
    const long long inter_p = n < (long long)kBins * 8 ? n : (long long)kBins * 8;
    red.size = p_prune(red.d, out.d, out.size, (int)inter_p, sc);
    const int mnb = inter_p < kBins ? (int)inter_p : kBins;
    temp.size = p_prune(temp.d, red.d, red.size, mnb + 1, sc);
  }
  if (tid == 0) {
    if (!parallel_ok) {
    make_summary(keys, qtail, out);
    Summary l0 = { level_space, 0 };
    if (levels_in_use != 0) {
      s_prune(l0, out, L);
      for (int l = 1; l < levels_in_use; ++l) {
        if (level_size[l] == 0) continue;
        Summary lv = { level_space + (size_t)l * L, level_size[l] };
        if (l0.size == 0) {
          s_copy(l0, lv);
        } else {
          s_combine(out, l0, lv);
          s_prune(l0, out, L);
        }
      }
      s_copy(out, l0);
    } else if (out.size > L) {
      s_prune(temp, out, L);
      s_copy(out, temp);
    }
    }
    const long long inter = n < (long long)kBins * 8 ? n : (long long)kBins * 8;
    const int max_num_bins = inter < kBins ? (int)inter : kBins;
    if (!parallel_ok) {
      s_prune(red, out, (int)inter);
      s_prune(temp, red, max_num_bins + 1);
    }
    // hist_util.h AddCutPoint + the trailing sentinel of MakeCuts
    const int required = temp.size < max_num_bins ? temp.size : max_num_bins;
    int nc = 0;
    for (int i2 = 1; i2 < required; ++i2) {
      float cpt = temp.d[i2].w;
      if (i2 == 1 || cpt > cuts[nc - 1]) cuts[nc++] = cpt;
    }
    float cpt = temp.d[temp.size - 1].w;
    cuts[nc++] = __fadd_rn(cpt, __fadd_rn(fabsf(cpt), 1e-5f));
    ncuts_out[stream] = nc;
  }
  SK_PROF(6);
}

// HistogramCuts::SearchBin: index of the first cut strictly greater than x, clamped to the last cut.
__global__ void k_bin_u8(const float* __restrict__ X, long long ldx, int f, int fpad,
                         const uint32_t* __restrict__ rows, long long n,
                         const float* __restrict__ cuts, const int* __restrict__ ncuts, uint8_t* __restrict__ bins) {
  const int words = fpad >> 2;
  const long long total = n * words;
  for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
    long long r = t / words;
    int w = (int)(t % words);
    const float* xr = X + (size_t)rows[r] * (size_t)ldx;
    uint32_t packed = 0;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      int j = w * 4 + q;
      uint32_t b = 0;
      if (j < f) {
        const float* c = cuts + (size_t)j * kBins;
        int nc = ncuts[j];
        float x = xr[j];
        int lo = 0, hi = nc;
        while (lo < hi) { int mid = (lo + hi) >> 1; if (c[mid] > x) hi = mid; else lo = mid + 1; }
        if (lo == nc) lo = nc - 1;
        b = (uint32_t)lo;
      }
      packed |= b << (8 * q);
    }
    reinterpret_cast<uint32_t*>(bins)[t] = packed;
  }
}

// feature-major copy of the bin matrix: out[j][r] = bins[r][j].  A CTA moves a tile of 128 rows through shared
// memory: row-major words in (coalesced), 128 consecutive rows of one feature out per warp store.
constexpr int kColTileRows = 128;
__global__ void __launch_bounds__(256) k_bins_to_columns(const uint8_t* __restrict__ bins, int f, int fpad, long long n, uint8_t* __restrict__ out) {
  extern __shared__ __align__(16) uint8_t s_tile[];      // [128][fpad + 4]: the 4-byte skew spreads a column over the banks
  const int pitch = fpad + 4, words = fpad >> 2;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const long long tiles = (n + kColTileRows - 1) / kColTileRows;
  for (long long tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
    const long long r0 = tile * kColTileRows;
    const int rows = (int)min((long long)kColTileRows, n - r0);
    const uint32_t* src = reinterpret_cast<const uint32_t*>(bins + (size_t)r0 * fpad);
    for (int t = threadIdx.x; t < rows * words; t += 256) {
      const int r = t / words, w = t - r * words;
      *reinterpret_cast<uint32_t*>(s_tile + r * pitch + 4 * w) = src[t];
    }
    __syncthreads();
    for (int j = wid; j < f; j += 8) {                  // lane l: rows 4l .. 4l + 3 of feature j
      const int r = 4 * lane;
      uint32_t v = 0;
#pragma unroll
      for (int q = 0; q < 4; ++q) if (r + q < rows) v |= (uint32_t)s_tile[(r + q) * pitch + j] << (8 * q);
      uint8_t* dst = out + (size_t)j * (size_t)n + (size_t)r0 + r;
      if (r + 3 < rows && ((n & 3) == 0)) *reinterpret_cast<uint32_t*>(dst) = v;     // r0 is a multiple of 128: aligned when n is a multiple of 4
      else for (int q = 0; q < 4; ++q) if (r + q < rows) dst[q] = (uint8_t)(v >> (8 * q));
    }
    __syncthreads();
  }
}

__global__ void k_gather_y(const float* __restrict__ y, const uint32_t* __restrict__ rows, long long n, float* __restrict__ out) {
  for (long long t = blockIdx.x * (long long)blockDim.x + threadIdx.x; t < n; t += (long long)gridDim.x * blockDim.x) out[t] = y[rows[t]];
}

}  // namespace

void launch_sketch(const float* X, long long ldx, int f, int nfold, const uint32_t* const* fold_rows,
                   const long long* fold_ntrain, const SketchPlan& plan, const int* fold_limit,
                   unsigned long long* key_arena, float4* entry_arena, float* cuts, int* ncuts, int stream0,
                   int count, cudaStream_t s) {
  k_sketch_cuts<<<count, kSketchThreads, kSketchSmem, s>>>(X, ldx, f, nfold, fold_rows, fold_ntrain, plan, fold_limit,
                                                           key_arena, entry_arena, cuts, ncuts, stream0);
}

#ifdef GT_SKETCH_PROF
extern "C" void dump_sketch_prof() {
  unsigned long long h[8];
  cudaMemcpyFromSymbol(h, g_sketch_prof, sizeof(h));
  const char* names[] = {"queue fill", "sort", "make summary", "prune", "combine", "copy + clear", "final merge + cuts"};
  unsigned long long tot = 0; for (int i = 0; i < 7; ++i) tot += h[i];
  for (int i = 0; i < 7; ++i) fprintf(stderr, "sketch prof %-18s %6.2f%%  %llu cycles\n", names[i], 100.0 * h[i] / (double)tot, h[i]);
  memset(h, 0, sizeof(h));
  cudaMemcpyToSymbol(g_sketch_prof, h, sizeof(h));
}
#endif

int configure_sketch() {   // once per device before the first launch_sketch: the kernel's dynamic shared memory is above 48 KB
  return (int)cudaFuncSetAttribute(k_sketch_cuts, cudaFuncAttributeMaxDynamicSharedMemorySize, kSketchSmem);
}

void launch_bin(const float* X, long long ldx, int f, int fpad, const uint32_t* rows, long long n,
                const float* cuts, const int* ncuts, uint8_t* bins, cudaStream_t s) {
  long long total = n * (fpad / 4);
  int blocks = (int)((total + 255) / 256 < 148 * 16 ? (total + 255) / 256 : 148 * 16);
  if (blocks < 1) blocks = 1;
  k_bin_u8<<<blocks, 256, 0, s>>>(X, ldx, f, fpad, rows, n, cuts, ncuts, bins);
}

void launch_bins_to_columns(const uint8_t* bins, int f, int fpad, long long n, uint8_t* out, cudaStream_t s) {
  const long long tiles = (n + kColTileRows - 1) / kColTileRows;
  const int smem = kColTileRows * (fpad + 4);            // 33 KB at 256 features: below the 48 KB that need no opt-in
  k_bins_to_columns<<<(int)std::min<long long>(tiles, 148 * 8), 256, smem, s>>>(bins, f, fpad, n, out);
}

// xgb.DMatrix treats NaN as "missing" and refuses inf; this engine has no missing-value path, so both are counted
// here and refused at gt_engine_create instead of flowing into the sketch as ordinary numbers.
static __global__ void k_count_nonfinite(const float* X, long long n, int f, long long ldx, const float* y, unsigned long long* out) {
  unsigned long long bad_x = 0, bad_y = 0;
  const long long total = n * f;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / f;
    const int c = (int)(i - r * f);
    if (!isfinite(X[r * ldx + c])) ++bad_x;
  }
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    if (!isfinite(y[i])) ++bad_y;
  bad_x = __reduce_add_sync(0xffffffffu, (unsigned)bad_x);
  bad_y = __reduce_add_sync(0xffffffffu, (unsigned)bad_y);
  if ((threadIdx.x & 31) == 0) {
    if (bad_x) atomicAdd(&out[0], bad_x);
    if (bad_y) atomicAdd(&out[1], bad_y);
  }
}
void launch_count_nonfinite(const float* X, long long n, int f, long long ldx, const float* y, unsigned long long* out, cudaStream_t s) {
  k_count_nonfinite<<<148 * 8, 256, 0, s>>>(X, n, f, ldx, y, out);
}
// Ingestion of the caller's matrix as it lies in host memory (gt_engine_create_from): a tile of `rows` rows arrives
// on the device in the caller's element type, either row-major (tile[r * ld + c]) or column-major (tile[c * ld + r],
// how pandas keeps a float64 DataFrame); it is converted to float (the cast xgb.DMatrix applies) and written row-major.
template <typename T, bool COLMAJOR>
static __global__ void __launch_bounds__(256) k_ingest_tile(const T* __restrict__ tile, long long ld, int rows, int f,
                                                            float* __restrict__ X, long long ldx, long long r0) {
  __shared__ float s_t[32][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;   // 32 x 8
  const int rb = blockIdx.x * 32, cb = blockIdx.y * 32;
  if (COLMAJOR) {
    for (int q = ty; q < 32; q += 8) {   // read along rows (contiguous in a column-major tile)
      const int c = cb + q, r = rb + tx;
      s_t[q][tx] = (c < f && r < rows) ? (float)tile[(long long)c * ld + r] : 0.f;
    }
    __syncthreads();
    for (int q = ty; q < 32; q += 8) {   // write along columns
      const int r = rb + q, c = cb + tx;
      if (r < rows && c < f) X[(r0 + r) * ldx + c] = s_t[tx][q];
    }
  } else {
    for (int q = ty; q < 32; q += 8) {
      const int r = rb + q, c = cb + tx;
      if (r < rows && c < f) X[(r0 + r) * ldx + c] = (float)tile[(long long)r * ld + c];
    }
  }
}
void launch_ingest_tile(const void* tile, int dtype, int colmajor, long long ld, int rows, int f, float* X, long long ldx,
                        long long r0, cudaStream_t s) {
  const dim3 grid((rows + 31) / 32, (f + 31) / 32);
  if (dtype == GT_DTYPE_F64) {
    if (colmajor) k_ingest_tile<double, true><<<grid, 256, 0, s>>>((const double*)tile, ld, rows, f, X, ldx, r0);
    else k_ingest_tile<double, false><<<grid, 256, 0, s>>>((const double*)tile, ld, rows, f, X, ldx, r0);
  } else {
    if (colmajor) k_ingest_tile<float, true><<<grid, 256, 0, s>>>((const float*)tile, ld, rows, f, X, ldx, r0);
    else k_ingest_tile<float, false><<<grid, 256, 0, s>>>((const float*)tile, ld, rows, f, X, ldx, r0);
  }
}
// train rows per (feature, bin) of one fold, in the cell order of a node histogram ([group][bin][feature lane])
// rows per (feature, bin) of a fold's train slice, in the histogram layout [group][bin][lane].  A CTA counts one
// feature group of a run of rows in shared memory (lane = feature: conflict-free banks), then adds its table to the
// global one.
constexpr int kCountRowsPerCta = 16384;
static __global__ void __launch_bounds__(256) k_count_hist(const uint8_t* __restrict__ bins, int f, int fpad, long long ntrain, int* __restrict__ out) {
  __shared__ int s_cnt[kBins * kLanes];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int g = blockIdx.y;
  const bool real = g * kLanes + lane < f;               // the padding lanes of the last group stay zero
  for (int i = threadIdx.x; i < kBins * kLanes; i += 256) s_cnt[i] = 0;
  __syncthreads();
  const long long r0 = (long long)blockIdx.x * kCountRowsPerCta, r1 = min(ntrain, r0 + kCountRowsPerCta);
  // a warp takes 4 rows per step: lane l loads word (l & 7) of row (l >> 3), then every lane collects the byte of its
  // own feature from each of the four rows
  for (long long r = r0 + 4 * wid; r < r1; r += 32) {
    const long long rr = r + (lane >> 3);
    const uint32_t word = rr < r1 ? *reinterpret_cast<const uint32_t*>(bins + (size_t)rr * fpad + (size_t)g * kLanes + 4 * (lane & 7)) : 0u;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const uint32_t wq = __shfl_sync(0xffffffffu, word, q * 8 + (lane >> 2));
      if (real && r + q < r1) atomicAdd(&s_cnt[((wq >> (8 * (lane & 3))) & 255u) * kLanes + lane], 1);
    }
  }
  __syncthreads();
  int* dst = out + (size_t)g * kBins * kLanes;
  for (int i = threadIdx.x; i < kBins * kLanes; i += 256) { const int c = s_cnt[i]; if (c) atomicAdd(&dst[i], c); }
}
void launch_count_hist(const uint8_t* bins, int f, int fpad, long long ntrain, int* out /* zeroed */, cudaStream_t s) {
  if (ntrain < 1) return;
  dim3 grid((unsigned)((ntrain + kCountRowsPerCta - 1) / kCountRowsPerCta), (unsigned)(fpad / kLanes));
  k_count_hist<<<grid, 256, 0, s>>>(bins, f, fpad, ntrain, out);
}
void launch_gather_y(const float* y, const uint32_t* rows, long long n, float* out, cudaStream_t s) {
  int blocks = (int)((n + 255) / 256 < 148 * 8 ? (n + 255) / 256 : 148 * 8);
  if (blocks < 1) blocks = 1;
  k_gather_y<<<blocks, 256, 0, s>>>(y, rows, n, out);
}

}  // namespace gt

// =====================================================================================================
// Allocation-free cut points for one column (gt_column_cuts_noalloc, include/gentun_b200.h).
//
// Who this is for: code that must not touch the device allocator when it asks for cut points -- a
// serving loop that re-bins a drifting feature while other streams are busy (cudaMalloc / cudaFree
// synchronise the whole device), a process running under a fixed memory reservation where a late
// allocation can fail, or a library hook that is only allowed to enqueue work.  gt_engine_prepare
// allocates its sketch arenas with cudaMalloc; this entry point runs the same k_sketch_cuts kernel
// for a single stream with every byte of scratch (row ids, sort keys, summary levels) taken from the
// fixed array below, which each GPU that loads the library keeps resident until the process exits.
// The price is a hard size limit (kNoAllocMaxRows values per call) and no concurrency: all calls share
// the one array, so two calls that overlap on the GPU (different streams, or different host threads
// on different streams) corrupt each other.  Calls on one stream are safe back to back: the kernel
// clears its sort keys and level sizes on entry and the set-up kernel rewrites the header and row ids.
namespace gt {
namespace {

constexpr long long kNoAllocMaxRows = 65536;
// LimitSizeLevel at 65536 rows and eps = 1/2048 gives limit_size 8193 over 4 levels; smaller inputs need less
constexpr int kNaLimitMax = 8193;
constexpr int kNaLevelAlloc = 4 + 3;
constexpr int kNaBufCap = 2 * kNaLimitMax + 2;
constexpr int kNaKeys = 32768;                                                       // power of two >= 2 * 8193
constexpr long long kNaEntries = 3ll * kNaBufCap + (long long)kNaLevelAlloc * kNaLimitMax;
constexpr size_t kNaRowsOff = 256;                                                   // header: rows pointer, count, limit
constexpr size_t kNaKeysOff = kNaRowsOff + (size_t)kNoAllocMaxRows * sizeof(uint32_t);
constexpr size_t kNaEntsOff = kNaKeysOff + (size_t)kNaKeys * sizeof(unsigned long long);
constexpr size_t kNaBytes = kNaEntsOff + (size_t)kNaEntries * sizeof(float4);       // about 2.2 MB
static_assert(kNaKeysOff % 256 == 0 && kNaEntsOff % 256 == 0, "scratch sections must stay 256-byte aligned");

static __device__ __align__(256) unsigned char g_cut_scratch[kNaBytes];

// device code may name the array directly; host code needs the address from cudaGetSymbolAddress
__global__ void k_noalloc_setup(long long n, int limit) {
  uint32_t* rows = reinterpret_cast<uint32_t*>(g_cut_scratch + kNaRowsOff);
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) rows[i] = (uint32_t)i;
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    *reinterpret_cast<const uint32_t**>(g_cut_scratch) = rows;     // the kernel's fold_rows[0]
    *reinterpret_cast<long long*>(g_cut_scratch + 8) = n;          // fold_ntrain[0]
    *reinterpret_cast<int*>(g_cut_scratch + 16) = limit;           // fold_limit[0]
  }
}

unsigned char* g_scratch_on_device[64] = {nullptr};   // per device, filled at first use

void na_limit_size_level(size_t maxn, double eps, size_t* nlevel, size_t* limit) {
  *nlevel = 1;
  for (;;) {
    *limit = (size_t)ceil((double)*nlevel / eps) + 1;
    if (maxn < *limit) *limit = maxn;
    if (((size_t)1 << *nlevel) * *limit >= maxn) break;
    ++*nlevel;
  }
}

int na_fail(int code, const char* what, const char* detail = "") {
  char buf[512];
  snprintf(buf, sizeof(buf), "gt_column_cuts_noalloc: %s%s", what, detail);
  set_error_message(buf);
  return code;
}

}  // namespace
}  // namespace gt

extern "C" int gt_column_cuts_noalloc(const float* d_values, int64_t n, float* d_cuts, int32_t* d_ncuts, void* stream) {
  using namespace gt;
  char msg[256];
  if (!d_values || !d_cuts || !d_ncuts) return na_fail(GT_ERR_ARG, "null pointer");
  if (n < 1 || n > kNoAllocMaxRows) {
    snprintf(msg, sizeof(msg), "n=%lld is outside 1..%lld, the capacity of the fixed scratch array", (long long)n, kNoAllocMaxRows);
    return na_fail(GT_ERR_ARG, msg);
  }
  int dev = 0;
  cudaError_t ce = cudaGetDevice(&dev);
  if (ce != cudaSuccess) return na_fail(GT_ERR_CUDA, "cudaGetDevice failed: ", cudaGetErrorString(ce));
  if (dev < 0 || dev >= 64) return na_fail(GT_ERR_ARG, "device ordinal above 63 is not supported");
  if (!g_scratch_on_device[dev]) {
    // once per device, with that device current: the host-side name of a __device__ array is not a device address
    void* p = nullptr;
    ce = cudaGetSymbolAddress(&p, g_cut_scratch);
    if (ce != cudaSuccess || !p) return na_fail(GT_ERR_CUDA, "cudaGetSymbolAddress failed: ", cudaGetErrorString(ce));
    ce = cudaFuncSetAttribute(k_sketch_cuts, cudaFuncAttributeMaxDynamicSharedMemorySize, kSketchSmem);   // no allocation: a launch attribute
    if (ce != cudaSuccess) return na_fail(GT_ERR_CUDA, "cudaFuncSetAttribute failed: ", cudaGetErrorString(ce));
    g_scratch_on_device[dev] = static_cast<unsigned char*>(p);
  }
  unsigned char* base = g_scratch_on_device[dev];
  const size_t n_bins = (size_t)(n < GT_MAX_BIN ? n : GT_MAX_BIN);
  const double eps = 1.0 / (double)((float)n_bins * 8.0f);
  size_t nlevel, limit;
  na_limit_size_level((size_t)n, eps, &nlevel, &limit);
  SketchPlan plan;
  plan.buf_cap = (int)(2 * limit + 2);
  plan.nlevel_alloc = (int)nlevel + 3;
  int kp = 2; while (kp < (int)(2 * limit)) kp <<= 1;
  plan.keys_per_stream = kp;
  plan.entries_per_stream = 3ll * plan.buf_cap + (long long)plan.nlevel_alloc * (long long)limit;
  if ((long long)limit > kNaLimitMax || plan.nlevel_alloc > kNaLevelAlloc || kp > kNaKeys || plan.entries_per_stream > kNaEntries)
    return na_fail(GT_ERR_ARG, "this n needs more scratch than the fixed array holds");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int blocks = (int)((n + 255) / 256 < 64 ? (n + 255) / 256 : 64);
  k_noalloc_setup<<<blocks, 256, 0, s>>>((long long)n, (int)limit);
  k_sketch_cuts<<<1, kSketchThreads, kSketchSmem, s>>>(d_values, 1, 1, 1, reinterpret_cast<const uint32_t* const*>(base),
                                             reinterpret_cast<const long long*>(base + 8), plan, reinterpret_cast<const int*>(base + 16),
                                             reinterpret_cast<unsigned long long*>(base + kNaKeysOff),
                                             reinterpret_cast<float4*>(base + kNaEntsOff), d_cuts, d_ncuts, 0);
  ce = cudaGetLastError();
  if (ce != cudaSuccess) return na_fail(GT_ERR_CUDA, "kernel launch failed: ", cudaGetErrorString(ce));
  return GT_OK;
}

extern "C" int64_t gt_column_cuts_noalloc_max_rows(void) { return gt::kNoAllocMaxRows; }
