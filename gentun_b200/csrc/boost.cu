// X4-X9 of the hot path: one boosting round of depth-wise histogram tree growth for a whole wave of
// (individual x fold) jobs at once, plus the per-round CV metric and early-stopping bookkeeping.
//
// Replaces the inner loop of xgb.cv as gentun calls it
// (/root/reference/gentun/models/xgboost_models.py:33-36): for every round, for every fold,
// Booster.update (objective gradient -> per-level histogram -> split enumeration -> row partition ->
// leaf values -> prediction cache) and Booster.eval_set + aggcv + EarlyStopping.
// Arithmetic follows upstream xgboost >= 2.0 tree_method=hist: src/tree/param.h (CalcWeight,
// CalcGainGivenWeight, ThresholdL1), src/tree/split_evaluator.h, src/tree/hist/evaluate_splits.h
// (EnumerateSplit<+1>, CPUExpandEntry::IsValid), src/objective/regression_obj.cu,
// src/metric/elementwise_metric.cu, python-package/xgboost/callback.py (EarlyStopping).
//
// Numeric design: gradients are quantised once per round to fixed point (power-of-two scale,
// |q| <= 2^19) so that all histogram arithmetic is integer: shared-memory partials are int32
// (native ATOMS.ADD), global cells are int64, parent - child subtraction is exact and the result
// does not depend on summation order.
#include "common.cuh"
#include "kernels.h"
#include "split.cuh"
#include <cstdio>

namespace gt {
namespace {

constexpr int kHistSmem = 2 * kBins * kLanes * (int)sizeof(int);   // g plane + h plane = 64 KiB
#ifndef GT_JOB_FASTEST_LEVELS
#define GT_JOB_FASTEST_LEVELS 2
#endif
constexpr int kJobFastestLevels = GT_JOB_FASTEST_LEVELS;   // k_hist: levels up to this one run the same item of many jobs together


// ------------------------------------------------------------------ round set-up / tear-down
__device__ void prepare_next_round(const Wave& w, int j) {
  JobState& js = w.js[j];
  const JobParams& jp = w.jp[j];
  float gmax = __uint_as_float(js.gmax_bits), hmax = __uint_as_float(js.hmax_bits);
  int e;
  float sg = 1.0f, sh = 1.0f;
  if (gmax > 0.f && isfinite(gmax)) { frexpf(gmax, &e); sg = ldexpf(1.0f, kQuantBits - e); }
  if (hmax > 0.f && isfinite(hmax)) { frexpf(hmax, &e); sh = ldexpf(1.0f, kQuantBitsH - e); }
  js.sg = sg; js.sh = sh;
  js.inv_sg = 1.0 / (double)sg; js.inv_sh = 1.0 / (double)sh;
  js.gmax_bits = 0u; js.hmax_bits = 0u;
  Node& r = w.nodes[(size_t)j * kNodeCap];
  r.G = 0; r.H = 0; r.begin = 0; r.cnt = (uint32_t)w.folds[jp.fold].ntrain;
  r.leaf = 0.f; r.root_gain = 0.f; r.loss_chg = 0.f; r.feat = -1; r.bin = 0; r.state = kOpen; r.buf = 0;
  r.narrow = r.cnt <= (uint32_t)kNarrowRows; r.absG = 0ull;
#ifdef GT_DBG_NO_NARROW
  r.narrow = 0;
#endif
  TreeNode& t = w.tree[(size_t)j * kNodeCap];
  t.leaf = 0.f; t.feat = -1; t.bin = 0; t.state = kOpen;
}

// pred = base margin; record max |g|, max h of round 0
__global__ void __launch_bounds__(256) k_init_wave(Wave w) {
  const int j = blockIdx.y;
  const JobParams jp = w.jp[j];
  const FoldDev fd = w.folds[jp.fold];
  float* pred = w.pred + (size_t)j * w.n;
  const long long r0 = (long long)blockIdx.x * kPredRows;
  float gm = 0.f, hm = 0.f;
  for (int i = threadIdx.x; i < kPredRows; i += blockDim.x) {
    long long r = r0 + i;
    if (r >= w.n) break;
    pred[r] = jp.base_margin;
    if (r < fd.ntrain) {
      float g, h;
      grad_pair(w.objective, jp.base_margin, fd.y[r], jp.spw, g, h);
      gm = fmaxf(gm, fabsf(g)); hm = fmaxf(hm, fabsf(h));
    }
  }
  for (int d = 16; d; d >>= 1) { gm = fmaxf(gm, __shfl_xor_sync(0xffffffffu, gm, d)); hm = fmaxf(hm, __shfl_xor_sync(0xffffffffu, hm, d)); }
  if ((threadIdx.x & 31) == 0) {
    if (gm > 0.f) atomicMax(&w.js[j].gmax_bits, __float_as_uint(gm));
    if (hm > 0.f) atomicMax(&w.js[j].hmax_bits, __float_as_uint(hm));
  }
}

// X4: objective gradient, row subsampling, fixed-point quantisation, root sums, identity row set
__global__ void __launch_bounds__(256) k_grad(Wave w) {
  const int j = blockIdx.y;
  const JobState& js = w.js[j];
  // a new round: the per-level pair lists start empty (done by one thread whether or not job 0 is still training)
  if (blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0) { w.counters[5] = 0ull; w.counters[6] = 0ull; }
  if (!js.active) return;
  const JobParams jp = w.jp[j];
  const FoldDev fd = w.folds[jp.fold];
  const int round = w.is[jp.ind].round;
  const float* pred = w.pred + (size_t)j * w.n;
  int2* gq = w.gq + (size_t)j * w.nt_stride;
  uint32_t* ridx = w.ridx + (size_t)j * 2 * w.nt_stride;
  Node* nodes = w.nodes + (size_t)j * kNodeCap;
  if (blockIdx.x == 0)
    for (int i = 1 + threadIdx.x; i < kNodeCap; i += blockDim.x) nodes[i].state = kAbsent;

  const uint64_t skey = stream_key(1u, (uint32_t)jp.fold, (uint32_t)round, 0u);
  const bool sub = jp.subsample < 1.0f;
  long long sG = 0, sH = 0;
  const long long r0 = (long long)blockIdx.x * kPredRows;
  for (int i = threadIdx.x; i < kPredRows; i += blockDim.x) {
    long long r = r0 + i;
    if (r >= fd.ntrain) break;
    float g, h;
    grad_pair(w.objective, pred[r], fd.y[r], jp.spw, g, h);
    if (sub && !(u01(draw64(skey, (uint64_t)r)) < jp.subsample)) { g = 0.f; h = 0.f; }
    int qg = __float2int_rn(g * js.sg), qh = __float2int_rn(h * js.sh);
    gq[r] = make_int2(qg, qh);
    ridx[r] = (uint32_t)r;
    sG += qg; sH += qh;
  }
  for (int d = 16; d; d >>= 1) { sG += __shfl_xor_sync(0xffffffffu, sG, d); sH += __shfl_xor_sync(0xffffffffu, sH, d); }
  __shared__ long long s_g[8], s_h[8];
  const int wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) { s_g[wid] = sG; s_h[wid] = sH; }
  __syncthreads();
  if (threadIdx.x == 0) {
    long long a = 0, b = 0;
    for (int q = 0; q < 8; ++q) { a += s_g[q]; b += s_h[q]; }
    if (a) atomicAdd(reinterpret_cast<unsigned long long*>(&nodes[0].G), (unsigned long long)a);
    if (b) atomicAdd(reinterpret_cast<unsigned long long*>(&nodes[0].H), (unsigned long long)b);
  }
}


__global__ void __launch_bounds__(256) k_plan_hist(Wave w, int level) {
  const int j = blockIdx.x;
  JobState& js = w.js[j];
  const JobParams jp = w.jp[j];
  __shared__ uint32_t s_mask[kMaskWords];
  __shared__ uint32_t s_in[kMaskWords];
  if (threadIdx.x == 0) { js.n_hist_items = 0; js.n_pairs = 0; js.n_zero = 0; js.n_part_items = 0; }
  if (j == 0 && threadIdx.x == 0) w.counters[5 + ((level + 1) & 1)] = 0ull;   // the next level's pair list starts empty
  if (!js.active || level >= jp.max_depth) return;
  const int round = w.is[jp.ind].round;
  // feature sets (colsample_bytree at level 0, colsample_bylevel at every level)
  if (level == 0) {
    for (int q = threadIdx.x; q < kMaskWords; q += blockDim.x) {
      uint32_t m = 0u;
      for (int b = 0; b < 32; ++b) if (q * 32 + b < w.f) m |= 1u << b;
      s_in[q] = m;
    }
    __syncthreads();
    int n_tree = w.f;
    if (jp.coltree != 1.0f) {
      n_tree = (int)(jp.coltree * (float)w.f); if (n_tree < 1) n_tree = 1;
      sample_mask(stream_key(2u, (uint32_t)jp.fold, (uint32_t)round, 0u), s_in, w.f, n_tree, s_mask);
      for (int q = threadIdx.x; q < kMaskWords; q += blockDim.x) js.tmask[q] = s_mask[q];
    } else {
      for (int q = threadIdx.x; q < kMaskWords; q += blockDim.x) js.tmask[q] = s_in[q];
    }
    if (threadIdx.x == 0) js.n_tree_feats = n_tree;
    __syncthreads();
  }
  for (int q = threadIdx.x; q < kMaskWords; q += blockDim.x) s_in[q] = js.tmask[q];
  __syncthreads();
  if (jp.collevel != 1.0f) {
    int n_lvl = (int)(jp.collevel * (float)js.n_tree_feats); if (n_lvl < 1) n_lvl = 1;
    sample_mask(stream_key(3u, (uint32_t)jp.fold, (uint32_t)round, (uint32_t)level), s_in, w.f, n_lvl, s_mask);
    for (int q = threadIdx.x; q < kMaskWords; q += blockDim.x) js.lmask[q] = s_mask[q];
  } else {
    for (int q = threadIdx.x; q < kMaskWords; q += blockDim.x) js.lmask[q] = s_in[q];
  }
  // work items: histogram chunks of the nodes to build, (built, sibling) pairs to evaluate
  Node* nodes = w.nodes + (size_t)j * kNodeCap;
  HistItem* items = w.hitems + (size_t)j * w.hist_cap;
  int* zl = w.zlist + (size_t)j * w.hist_cap;
  __shared__ int s_scan[3][256];
  __shared__ unsigned long long s_rows;
  __shared__ int s_pair_base;
  if (threadIdx.x == 0) s_rows = 0ull;

  // each thread owns up to two parents (2 * 256 >= 2^(kMaxDepth - 1))
  const int n_par = level == 0 ? 1 : (1 << (level - 1));
  const int first = level == 0 ? 0 : level_first(level - 1);
  int bnode[2], chunks[2], shr[2], single[2];
  for (int s2 = 0; s2 < 2; ++s2) {
    const int pi = 2 * threadIdx.x + s2;
    bnode[s2] = -1; chunks[s2] = 0; shr[s2] = 0; single[s2] = 0;
    if (pi >= n_par) continue;
    if (level == 0) {
      bnode[s2] = 0;
    } else {
      const int p = first + pi;
      if (nodes[p].state != kSplit) continue;
      const int l = 2 * p + 1, r = 2 * p + 2;
      const bool lo = nodes[l].state == kOpen, ro = nodes[r].state == kOpen;
      if (!lo && !ro) continue;                       // both children were settled as leaves when the parent split
      single[s2] = !(lo && ro);
      bnode[s2] = single[s2] ? (lo ? l : r) : (nodes[l].cnt <= nodes[r].cnt ? l : r);
    }
    const uint32_t cnt = nodes[bnode[s2]].cnt;
    chunks[s2] = (int)((cnt + kHistRows - 1) / kHistRows);
    if (chunks[s2] < 1) chunks[s2] = 1;
    // several CTAs build the node: they accumulate with atomics into a cleared histogram.  A node built by one
    // CTA is owned by it: plain stores, nothing to clear.
    shr[s2] = chunks[s2] > 1;
  }
  __syncthreads();   // s_rows initialised
  int v[3] = {chunks[0] + chunks[1], (bnode[0] >= 0) + (bnode[1] >= 0), shr[0] + shr[1]};
  for (int q = 0; q < 3; ++q) s_scan[q][threadIdx.x] = v[q];
  __syncthreads();
  for (int d = 1; d < 256; d <<= 1) {   // Hillis-Steele inclusive scan of the three counters
    int t[3];
    for (int q = 0; q < 3; ++q) t[q] = threadIdx.x >= d ? s_scan[q][threadIdx.x - d] : 0;
    __syncthreads();
    for (int q = 0; q < 3; ++q) s_scan[q][threadIdx.x] += t[q];
    __syncthreads();
  }
  if (threadIdx.x == 255) s_pair_base = (int)atomicAdd(&w.counters[5 + (level & 1)], (unsigned long long)s_scan[1][255]);
  __syncthreads();
  PairRec* pairs = w.pairs + s_pair_base;
  int ni = s_scan[0][threadIdx.x] - v[0], np = s_scan[1][threadIdx.x] - v[1], nz = s_scan[2][threadIdx.x] - v[2];
  unsigned long long rows_mine = 0;
  for (int s2 = 0; s2 < 2; ++s2) {
    if (bnode[s2] < 0) continue;
    const Node& bn = nodes[bnode[s2]];
    for (int c = 0; c < chunks[s2]; ++c, ++ni) {
      if (ni >= w.hist_cap) { w.counters[7] = 1ull; break; }   // never expected (the cap is an upper bound); reported, not ignored
      const uint32_t b = (uint32_t)c * kHistRows;
      const uint32_t cnt = bn.cnt > b ? (bn.cnt - b < (uint32_t)kHistRows ? bn.cnt - b : (uint32_t)kHistRows) : 0u;
      items[ni] = HistItem{bnode[s2], bn.begin + b, cnt, shr[s2] | ((int)bn.buf << 1)};
    }
    if (shr[s2]) { if (nz < w.hist_cap) zl[nz] = bnode[s2]; else w.counters[7] = 1ull; ++nz; }
    pairs[np] = level == 0 ? PairRec{-1, 0, 1, j} : PairRec{(bnode[s2] - 1) >> 1, (bnode[s2] & 1) ? 0 : 1, single[s2], j};
    ++np;
    rows_mine += bn.cnt;
  }
  if (rows_mine) atomicAdd(&s_rows, rows_mine);
  __syncthreads();
  if (threadIdx.x == 255) {
    js.n_hist_items = min(s_scan[0][255], w.hist_cap);
    if (s_scan[0][255] > hist_items_bound(w, level)) w.counters[7] = 1ull;   // items the grid would not reach
    js.n_pairs = s_scan[1][255];
    js.n_zero = min(s_scan[2][255], w.hist_cap);
    if (s_rows) atomicAdd(&w.counters[0], s_rows);
    if (s_scan[1][255]) atomicAdd(&w.counters[1], (unsigned long long)s_scan[1][255]);
  }
}

__device__ __forceinline__ HistEntry* node_hist(const Wave& w, const JobParams& jp, int level, int node) {
  return w.hist + ((level & 1) ? jp.hist_off[1] : jp.hist_off[0]) + (size_t)(node - level_first(level)) * w.groups * (kBins * kLanes);
}

// A node keeps its histogram as int32 pairs ("narrow": half the bytes, and a 32-bit split scan) when no cell can
// exceed 31 bits.  |cell| <= sum of |g_q| over the node's rows <= the same sum over its parent's rows, which is known
// exactly when the node gets its rows (Node::absG: k_hist for the root and every built child, parent - built for a
// sibling); without that, rows * 2^19 bounds it.  Only nodes built by one CTA qualify (plain stores).  Else int64 pairs.
__device__ __forceinline__ bool child_narrow(const Node& parent, uint32_t cnt, long long H) {
#ifdef GT_DBG_NO_NARROW
  return false;
#endif
  if (cnt <= (uint32_t)kNarrowRows) return true;
  return cnt <= (uint32_t)kHistRows && parent.absG <= 0x7fffffffull && H <= 0x7fffffffll;
}
__device__ __forceinline__ longlong2 load_cell(const HistEntry* slot, bool narrow, size_t idx) {
  if (narrow) { const int2 v = reinterpret_cast<const int2*>(slot)[idx]; return make_longlong2((long long)v.x, (long long)v.y); }
  return reinterpret_cast<const longlong2*>(slot)[idx];
}
__device__ __forceinline__ void store_cell(HistEntry* slot, bool narrow, size_t idx, const longlong2 e) {
  if (narrow) reinterpret_cast<int2*>(slot)[idx] = make_int2((int)e.x, (int)e.y);
  else reinterpret_cast<longlong2*>(slot)[idx] = e;
}

__global__ void __launch_bounds__(256) k_hist_zero(Wave w, int level) {
  const int j = blockIdx.y;
  const JobState& js = w.js[j];
  const JobParams jp = w.jp[j];
  const int* zl = w.zlist + (size_t)j * w.hist_cap;
  const int cells = w.groups * kBins * kLanes;
  for (int z = blockIdx.x; z < js.n_zero; z += gridDim.x) {
    longlong2* h = reinterpret_cast<longlong2*>(node_hist(w, jp, level, zl[z]));
    if (level == 0 && jp.uniform_h) {       // see k_hist: hessian plane = row count x the one quantised hessian
      const int* cnt = w.folds[jp.fold].count_hist;
      const long long hq = (long long)__float2int_rn(1.0f * js.sh);
      for (int i = threadIdx.x; i < cells; i += blockDim.x) h[i] = make_longlong2(0, (long long)cnt[i] * hq);
    } else {
      for (int i = threadIdx.x; i < cells; i += blockDim.x) h[i] = make_longlong2(0, 0);
    }
  }
}

#ifdef GT_DBG_CLEAR_HIST   // debugging aid: every histogram slot of the level starts as zeros, so a read of a slot that
                           // nothing wrote this level shows up as a changed result instead of depending on stale bytes
__global__ void __launch_bounds__(256) k_dbg_clear_level(Wave w, int level) {
  const int j = blockIdx.y;
  const JobParams jp = w.jp[j];
  if (level >= jp.max_depth) return;
  longlong2* h = reinterpret_cast<longlong2*>(node_hist(w, jp, level, level_first(level) + blockIdx.x));
  for (int i = threadIdx.x; i < w.groups * kBins * kLanes; i += blockDim.x) h[i] = make_longlong2(0x0123456789abcdefll, 0x0fedcba987654321ll);
}
#endif

// ------------------------------------------------------------------ X5: histogram
// One CTA = (job, work item, feature group of 32).  A warp takes 16 rows per step: lane = (row, 16-byte
// half of the group's 32 bin bytes), loaded with one 128-bit read.  The 16 bytes are pre-rotated by the
// row slot so that at unrolled step s the 32 lanes of a warp address 32 different features; with the
// shared histogram laid out [bin][feature] the bank is the feature, so every ATOMS is conflict-free.
__device__ __forceinline__ uint4 ld_stream(const uint4* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}

// UNITS = 16-byte units of a row handled by one CTA: 2 = one feature group (32 B of the row, 64 KiB of
// shared histogram, 3 CTAs/SM), 4 = a pair of groups (the whole 64-byte line of the row, 128 KiB, 1 CTA/SM).
template <int UNITS>
struct HistStage { uint4 b[UNITS]; int2 gv; };   // gv: gradient pair of THIS lane's row; sub-batches take theirs by shuffle

__device__ __forceinline__ void red_shared(uint32_t addr, int v) {
  asm volatile("red.shared.add.s32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}
__device__ __forceinline__ void red_shared_hi(uint32_t addr, int v) {   // hessian plane sits 32 KiB above
  asm volatile("red.shared.add.s32 [%0+32768], %1;" ::"r"(addr), "r"(v) : "memory");
}

// 32 rows of a warp's stripe = UNITS sub-batches of 32/UNITS rows; lane = (row slot, 16-byte unit)
template <int UNITS>
__device__ __forceinline__ void hist_load(HistStage<UNITS>& st, uint32_t myrow, int rslot, const uint8_t* __restrict__ bins_unit,
                                          const int2* __restrict__ gq, int fpad) {
#pragma unroll
  for (int sb = 0; sb < UNITS; ++sb) {
    const uint32_t r = __shfl_sync(0xffffffffu, myrow, sb * (32 / UNITS) + rslot);
    // rows past the end of the item read row 0 and add a zero gradient: no predicate in the hot loop
    const bool ok = r != 0xffffffffu;
    const uint32_t a = ok ? r : 0u;
    st.b[sb] = ld_stream(reinterpret_cast<const uint4*>(bins_unit + (unsigned long long)a * (unsigned)fpad));   // u32 x u32 + u64: one IMAD.WIDE
  }
  // one coalesced gather for the warp's 32 rows (each lane its own row) instead of one per sub-batch
  st.gv = myrow != 0xffffffffu ? gq[myrow] : make_int2(0, 0);
}

template <bool WITH_H>
__device__ __forceinline__ void hist_accumulate(const uint4 v, const int2 g2, bool q1, bool q2, int rb, const uint32_t (&cs)[16]) {
  // rotate the 16 bytes right by `rot` bytes: two conditional word rotations, then a funnel shift
  const uint32_t a0 = q1 ? v.y : v.x, a1 = q1 ? v.z : v.y, a2 = q1 ? v.w : v.z, a3 = q1 ? v.x : v.w;
  const uint32_t u0 = q2 ? a2 : a0, u1 = q2 ? a3 : a1, u2 = q2 ? a0 : a2, u3 = q2 ? a1 : a3;
  uint32_t r[4];
  r[0] = __funnelshift_r(u0, u1, rb);
  r[1] = __funnelshift_r(u1, u2, rb);
  r[2] = __funnelshift_r(u2, u3, rb);
  r[3] = __funnelshift_r(u3, u0, rb);
  // addresses of eight steps first, then their sixteen reductions: an in-flight RED still holds its address
  // register, so reusing one or two registers for every step stalls the next address computation
#pragma unroll
  for (int half = 0; half < 2; ++half) {
    uint32_t cell[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const int s = half * 8 + k;
      const uint32_t bin = __byte_perm(r[s >> 2], 0u, 0x4440u + (uint32_t)(s & 3));
      cell[k] = bin * (kLanes * 4u) + cs[s];
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      red_shared(cell[k], g2.x);
      if (WITH_H) red_shared_hi(cell[k], g2.y);
    }
  }
}

// Shared partial sums are int32.  A cell can never exceed the sum of |q| over the rows the CTA has
// accumulated since its last flush, so the CTA tracks that sum (and the hessian sum) and flushes to
// the global histogram before another epoch of kEpochRows rows could overflow.
// How a CTA's partial sums reach the node's global histogram:
enum FlushMode {
  kFlushAtomic,       // node shared by several CTAs: 64-bit atomic adds into cells cleared by k_hist_zero
  kFlushStoreNarrow,  // node owned by this CTA, <= kNarrowRows rows: int32 pairs, one plain store (never flushes early)
  kFlushStoreWide,    // node owned by this CTA, first flush: int64 pairs, plain store
  kFlushAddWide       // node owned by this CTA, later flushes: add to what it stored before
};
template <int UNITS>
__device__ __forceinline__ void hist_flush(const Wave& w, const JobParams& jp, int level, const HistItem& it, int g_first, int* sh,
                                           FlushMode mode, bool rezero) {
  constexpr int kThreads = UNITS == 2 ? 256 : 512;
  constexpr int kGroupsPerCta = UNITS / 2;
  HistEntry* slot = node_hist(w, jp, level, it.node);
  const size_t c0 = (size_t)g_first * (kBins * kLanes);   // first cell of this CTA's groups inside the node
  for (int i = threadIdx.x; i < kGroupsPerCta * kBins * kLanes; i += kThreads) {
    const int gsel = i / (kBins * kLanes), cell = i % (kBins * kLanes);
    int* pg = sh + gsel * (2 * kBins * kLanes) + cell;
    const int g = pg[0], h = pg[kBins * kLanes];
    if (mode == kFlushAtomic) {
      if (g) atomicAdd(reinterpret_cast<unsigned long long*>(&slot[c0 + i].g), (unsigned long long)(long long)g);
      if (h) atomicAdd(reinterpret_cast<unsigned long long*>(&slot[c0 + i].h), (unsigned long long)(long long)h);
    } else if (mode == kFlushStoreNarrow) {
      reinterpret_cast<int2*>(slot)[c0 + i] = make_int2(g, h);   // cells are int32 pairs packed from the slot base
    } else {
      longlong2* cellp = reinterpret_cast<longlong2*>(slot) + c0 + i;
      longlong2 v = make_longlong2((long long)g, (long long)h);
      if (mode == kFlushAddWide) { const longlong2 o = *cellp; v.x += o.x; v.y += o.y; }
      *cellp = v;
    }
    if (rezero) { pg[0] = 0; pg[kBins * kLanes] = 0; }
  }
}

template <int UNITS>
__global__ void __launch_bounds__(UNITS == 2 ? 256 : 512, UNITS == 2 ? 2 : 1) k_hist(Wave w, int level, int g0, int ncg) {
  constexpr int kThreads = UNITS == 2 ? 256 : 512;
  constexpr int kGroupsPerCta = UNITS / 2;
  constexpr uint32_t kStride = (kThreads / 32) * 32u;          // rows per CTA iteration
  constexpr int kEpochIters = kEpochRows / (int)kStride;       // CTA iterations between overflow checks
  // Grid at the top levels: x = (job, feature-group slice), y = work item, and x enumerates the level's jobs fold-major
  // (all individuals of fold 0, then fold 1, ...).  CTAs are dispatched x-fastest, so the CTAs in flight at one time
  // are the SAME item index of many jobs of ONE fold: with row sets still near the identity they read the same bin
  // rows, which then come from L2 for all but the first of them (level-0 DRAM reads 9.0 -> 2.3 GB on the C3 shape).
  // this launch covers the feature groups [g0, g0 + ncg * kGroupsPerCta) in ncg slices of kGroupsPerCta groups
  int j, item_idx, cg;
  if (level <= kJobFastestLevels) {
    const int lvl_jobs = (int)gridDim.x / ncg, lvl_inds = lvl_jobs / w.nfold;
    const int jx = (int)blockIdx.x / ncg;
    cg = (int)blockIdx.x % ncg;
    j = (jx % lvl_inds) * w.nfold + jx / lvl_inds;
    item_idx = blockIdx.y;
  } else {     // deeper down the jobs' row sets have nothing in common any more: one job's items side by side (measured 2 % faster there)
    j = blockIdx.y;
    item_idx = (int)blockIdx.x / ncg; cg = (int)blockIdx.x % ncg;
  }
  const JobState& js = w.js[j];
  if (item_idx >= js.n_hist_items) return;
  const int g_first = g0 + cg * kGroupsPerCta;
  const JobParams& jp = w.jp[j];
  const HistItem it = w.hitems[(size_t)j * w.hist_cap + item_idx];
  extern __shared__ int sh[];   // [group in CTA][g plane | h plane][bin][feature lane]
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  // row ids of the first two iterations are requested before the shared histogram is cleared, so that the
  // clear overlaps the first of the three dependent loads (item -> row id -> bin line)
  const uint32_t* ridx = w.ridx + ((size_t)j * 2 + (size_t)((it.flags >> 1) & 1)) * w.nt_stride + it.begin;
  const uint32_t cnt = it.cnt;
  const uint32_t my0 = (uint32_t)wid * 32u + (uint32_t)lane;
  uint32_t row_b = my0 < cnt ? ridx[my0] : 0xffffffffu;                              // iteration 0
  uint32_t row_c = my0 + kStride < cnt ? ridx[my0 + kStride] : 0xffffffffu;          // iteration 1
  for (int i = threadIdx.x; i < kGroupsPerCta * 2 * kBins * kLanes / 4; i += kThreads) reinterpret_cast<int4*>(sh)[i] = make_int4(0, 0, 0, 0);
  __syncthreads();
  const int rslot = lane / UNITS, unit = lane % UNITS;
  const int rot = UNITS == 2 ? rslot : rslot * 2 + (unit >> 1);   // 16 distinct rotations per bank half
  const bool q1 = (rot >> 2) & 1, q2 = (rot >> 3) & 1;
  const int rb = (rot & 3) * 8;
  const uint8_t* bins_unit = w.folds[jp.fold].bins + (size_t)g_first * kLanes + unit * 16;
  const int2* gq = w.gq + (size_t)j * w.nt_stride;
  const int fpad = w.fpad;
  const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(sh);
  uint32_t cs[16];   // shared address of this lane's feature column (bin 0, g plane) at unrolled step s
#pragma unroll
  for (int s = 0; s < 16; ++s) {
    const uint32_t fl = (uint32_t)((unit & 1) * 16 + ((s + rot) & 15));
    const uint32_t v = sbase + (uint32_t)(unit >> 1) * (2u * kBins * kLanes * 4u) + fl * 4u;
    asm volatile("mov.u32 %0, %1;" : "=r"(cs[s]) : "r"(v));   // opaque: keep it in a register, never rematerialise
  }
  const bool shared_node = it.flags & 1;
  // Root of a job whose rows all carry the same hessian (squared error, no row subsampling, no weighted label): the
  // hessian plane of its histogram is the fold's row COUNT per (feature, bin) times that one value -- the same for
  // every such job of the fold, counted once when the folds were binned.  k_hist_zero writes it; only the gradient
  // plane is accumulated here, i.e. half the shared-memory reductions of the level that has the most rows.
  const bool with_h = !(level == 0 && shared_node && jp.uniform_h);
  Node& node = w.nodes[(size_t)j * kNodeCap + it.node];
  const bool never_early = !shared_node && node.narrow;   // owned narrow node: no cell of it can leave 31 bits
  bool stored = false;                                                     // owned wide node: has an early flush happened
  // three-deep software pipeline: row ids two iterations ahead, bin bytes + gradients one ahead
  HistStage<UNITS> cur, nxt;
  hist_load<UNITS>(cur, row_b, rslot, bins_unit, gq, fpad);
  row_b = row_c;
  const uint32_t n_iter = (cnt + kStride - 1) / kStride;   // CTA-uniform trip count
  unsigned acc_g = 0u, acc_h = 0u, abs_rows = 0u;   // abs_rows: this lane's share of the node's sum of |g_q| (< 2^27 per item)
  unsigned run_g = 0u, run_h = 0u;                  // this warp's sums since the CTA last looked at its cells
  constexpr unsigned kWarps = kThreads / 32;
  // what a warp may add before the CTA has to look: an equal share of what the int32 range leaves above the largest
  // cell seen at the last look, less a worst-case epoch
  unsigned lim_g = (0x7fffffffu - ((unsigned)kEpochRows << kQuantBits)) / kWarps;
  unsigned lim_h = (0x7fffffffu - ((unsigned)kEpochRows << kQuantBitsH)) / kWarps;
  __shared__ unsigned s_cellmax[kWarps];
  for (uint32_t itn = 0; itn < n_iter; ++itn) {
    const uint32_t p2 = my0 + (itn + 2) * kStride;
    row_c = p2 < cnt ? ridx[p2] : 0xffffffffu;
    if (itn + 1 < n_iter) hist_load<UNITS>(nxt, row_b, rslot, bins_unit, gq, fpad);
    if (with_h) {          // CTA-uniform
#pragma unroll
      for (int sb = 0; sb < UNITS; ++sb) {
        const int src = sb * (32 / UNITS) + rslot;
        const int2 g2 = make_int2(__shfl_sync(0xffffffffu, cur.gv.x, src), __shfl_sync(0xffffffffu, cur.gv.y, src));
        hist_accumulate<true>(cur.b[sb], g2, q1, q2, rb, cs);
      }
    } else {
#pragma unroll
      for (int sb = 0; sb < UNITS; ++sb) {
        const int src = sb * (32 / UNITS) + rslot;
        const int2 g2 = make_int2(__shfl_sync(0xffffffffu, cur.gv.x, src), 0);
        hist_accumulate<false>(cur.b[sb], g2, q1, q2, rb, cs);
      }
    }
    acc_g += (unsigned)abs(cur.gv.x); acc_h += (unsigned)cur.gv.y;   // every lane holds one row of the iteration
    cur = nxt;
    row_b = row_c;
    if ((itn + 1) % kEpochIters == 0 && itn + 1 < n_iter) {
      // epoch boundary: will another epoch of worst-case rows still fit in int32?  Every cell is bounded by the sum,
      // over the CTA's warps, of the |g_q| (h_q) each has accumulated since the last flush; rows are dealt to the
      // warps round-robin, so each warp holds its own sum against an equal share of what the int32 range leaves
      // after a worst-case epoch, and one barrier collects the verdict.
      run_g += __reduce_add_sync(0xffffffffu, acc_g);
      run_h += __reduce_add_sync(0xffffffffu, acc_h);
      abs_rows += acc_g;
      acc_g = 0u; acc_h = 0u;
      const bool must = __syncthreads_or((run_g >= lim_g || run_h >= lim_h) && !never_early);
#ifdef GT_HIST_NO_LOOK
      if (must) {
        hist_flush<UNITS>(w, jp, level, it, g_first, sh, shared_node ? kFlushAtomic : stored ? kFlushAddWide : kFlushStoreWide, true);
        stored = true;
        run_g = 0u; run_h = 0u;
        __syncthreads();
      }
#else
      if (must) {
        // The running sums bound every cell as if all rows had fallen into one bin.  Look at the cells themselves:
        // with rows spread over many bins the largest one is far below the bound, and the CTA carries on from there
        // instead of flushing (a flush is ~65 000 global atomics, a look is one pass over shared memory).
        unsigned m = 0u;
        for (int i = threadIdx.x; i < kGroupsPerCta * 2 * kBins * kLanes / 4; i += kThreads) {
          const int4 v = reinterpret_cast<const int4*>(sh)[i];
          m = max(max(m, (unsigned)abs(v.x)), max((unsigned)abs(v.y), max((unsigned)abs(v.z), (unsigned)abs(v.w))));
        }
        m = __reduce_max_sync(0xffffffffu, m);
        if (lane == 0) s_cellmax[wid] = m;
        __syncthreads();
        m = 0u;
#pragma unroll
        for (unsigned q = 0; q < kWarps; ++q) m = max(m, s_cellmax[q]);
        if (m >= 0x40000000u) {
          hist_flush<UNITS>(w, jp, level, it, g_first, sh, shared_node ? kFlushAtomic : stored ? kFlushAddWide : kFlushStoreWide, true);
          stored = true;
          m = 0u;
        }
        lim_g = (0x7fffffffu - m - ((unsigned)kEpochRows << kQuantBits)) / kWarps;
        lim_h = (0x7fffffffu - m - ((unsigned)kEpochRows << kQuantBitsH)) / kWarps;
        run_g = 0u; run_h = 0u;
        __syncthreads();
      }
#endif
    }
  }
  // the node's exact sum of |g_q| decides the cell format of its children (child_narrow)
  if (g_first == 0) {
    const unsigned warp_abs = __reduce_add_sync(0xffffffffu, abs_rows + acc_g);
    if (lane == 0 && warp_abs) atomicAdd(&node.absG, (unsigned long long)warp_abs);
  }
  __syncthreads();
  hist_flush<UNITS>(w, jp, level, it, g_first, sh,
                    shared_node ? kFlushAtomic : never_early ? kFlushStoreNarrow : stored ? kFlushAddWide : kFlushStoreWide, false);
}

// ------------------------------------------------------------------ X6: split enumeration
#ifdef GT_SCAN_STATS   // experiment builds only: how often stage 2 takes which path
__device__ unsigned long long g_scan_stats[4];
#define SCAN_STAT(i, v) atomicAdd(&g_scan_stats[i], (unsigned long long)(v))
#else
#define SCAN_STAT(i, v)
#endif


__host__ __device__ inline int scan_chunks(int groups) { return groups <= 2 ? 8 : groups <= 4 ? 4 : 2; }

struct LaneRes { float e1, e2, Gt_f, Ht_f; int b1; long long GL1, HL1; };


template <typename T>
struct ChildAcc {
  T GL, HL;        // running prefix
  T hmin, hmax;    // a candidate keeps min_child_weight on both sides  <=>  hmin <= HL <= hmax
  T Ht;            // node total: the right side is empty  <=>  HL == Ht
  T GL1, HL1;      // prefix at the best estimate
  float Gt_f, Ht_f, Htl_f, e1, e2;   // Htl_f = Ht_f + lambda
  int b1;
};

template <typename T>
__device__ __forceinline__ void acc_init(ChildAcc<T>& c, const ChildTot& t, const GainK& k, long long hmin_q, float isg, float ish,
                                         long long GL0, long long HL0) {
  long long lo = hmin_q, hi = t.Ht - hmin_q;
  if (sizeof(T) == 4) {   // 32-bit lanes: prefixes stay below 2^31, so an out-of-range bound just means "never a candidate"
    lo = lo > 0x7fffffffll ? 0x7fffffffll : lo;
    hi = hi < -1 ? -1 : hi > 0x7fffffffll ? 0x7fffffffll : hi;
  }
  c.GL = (T)GL0; c.HL = (T)HL0; c.hmin = (T)lo; c.hmax = (T)hi; c.Ht = (T)t.Ht;
  c.GL1 = 0; c.HL1 = 0; c.b1 = -1;
  c.Gt_f = (float)t.Gt * isg; c.Ht_f = (float)t.Ht * ish; c.Htl_f = c.Ht_f + k.lambda;
  c.e1 = c.e2 = -3.0e38f;
}

template <typename T>
__device__ __forceinline__ void acc_result(const ChildAcc<T>& c, LaneRes& r) {
  r.e1 = c.e1; r.e2 = c.e2; r.Gt_f = c.Gt_f; r.Ht_f = c.Ht_f; r.b1 = c.b1;
  r.GL1 = (long long)c.GL1; r.HL1 = (long long)c.HL1;
}

// one bin of one child; every bin is estimated and then masked: no divergent control flow in the hot loop.
// (The estimate may contract multiplies and adds -- only the exact stage mirrors the reference's expression order.)
template <typename T, bool MDS>
__device__ __forceinline__ void child_step(ChildAcc<T>& c, const GainK& k, T ex, T ey, int b, bool in_range, float isg, float ish) {
  c.GL += ex; c.HL += ey;
  const bool cand = in_range & ((ex | ey) != 0) & (c.HL >= c.hmin) & (c.HL <= c.hmax);
  const float glq = (float)c.GL, hlq = (float)c.HL;
  const float gl = glq * isg, gr = __fmaf_rn(-glq, isg, c.Gt_f);
  const float dl = __fmaf_rn(hlq, ish, k.lambda), dr = __fmaf_rn(-hlq, ish, c.Htl_f);
  const float a = c.HL > 0 ? gain_term_k<MDS>(k, gl, dl) : 0.f;
  const float r = c.HL < c.Ht ? gain_term_k<MDS>(k, gr, dr) : 0.f;
  const float est = cand ? a + r : -3.0e38f;
  const bool top = est > c.e1;
  c.e2 = fmaxf(c.e2, fminf(c.e1, est));
  c.e1 = fmaxf(c.e1, est);
  c.b1 = top ? b : c.b1;
  c.GL1 = top ? c.GL : c.GL1;
  c.HL1 = top ? c.HL : c.HL1;
}


template <int WIDTH>
__device__ __forceinline__ void warp_merge(Cand& best) {
  for (int d = WIDTH; d; d >>= 1) {
    const float ochg = __shfl_xor_sync(0xffffffffu, best.chg, d);
    const int of = __shfl_xor_sync(0xffffffffu, best.f, d);
    const int ob = __shfl_xor_sync(0xffffffffu, best.b, d);
    const long long oGL = __shfl_xor_sync(0xffffffffu, best.GL, d);
    const long long oHL = __shfl_xor_sync(0xffffffffu, best.HL, d);
    if (better(ochg, of, ob, best.chg, best.f, best.b)) { best.chg = ochg; best.f = of; best.b = ob; best.GL = oGL; best.HL = oHL; }
  }
}

// Stage 2, slow path: several candidates of ONE feature column (feature f_col, cells at goff_col) have estimates within
// the margin of the best.  The whole warp takes the column: lane i re-reads bins b_lo + i, b_lo + 32 + i, ..., a warp
// scan rebuilds the exact prefixes, and every candidate whose estimate reaches `cut` is evaluated exactly.  All
// arguments are warp-uniform.  Equal gains keep the earlier bin, as the sequential scan would.
__device__ __noinline__ Cand column_exact(const JobParams& jp, const ChildTot& t, float Gt_f, float Ht_f, int side, float cut, int f_col,
                                             int b_lo, int hi, long long GL0, long long HL0, const HistEntry* hbuilt, bool nar_b,
                                             const HistEntry* hpar, bool nar_p, size_t goff_col, long long hmin_q, float isg, float ish,
                                             double inv_sg, double inv_sh) {
  const int lane = threadIdx.x & 31;
  const GainK k{jp.alpha, jp.lambda, jp.mds};
  constexpr int kBlocks = kBins / 32;
  longlong2 e[kBlocks];
#pragma unroll
  for (int q = 0; q < kBlocks; ++q) {
    const int b = b_lo + q * 32 + lane;
    e[q] = make_longlong2(0, 0);
    if (b < hi) {
      e[q] = load_cell(hbuilt, nar_b, goff_col + (size_t)b * kLanes);
      if (side == 1) { const longlong2 pe = load_cell(hpar, nar_p, goff_col + (size_t)b * kLanes); e[q].x = pe.x - e[q].x; e[q].y = pe.y - e[q].y; }
    }
  }
  Cand best{0.0f, -1, -1, 0, 0};
  long long GLc = GL0, HLc = HL0;
#pragma unroll
  for (int q = 0; q < kBlocks; ++q) {
    if (b_lo + q * 32 < hi) {   // warp-uniform
      long long pg = e[q].x, ph = e[q].y;
      for (int d = 1; d < 32; d <<= 1) {
        const long long og = __shfl_up_sync(0xffffffffu, pg, d), oh = __shfl_up_sync(0xffffffffu, ph, d);
        if (lane >= d) { pg += og; ph += oh; }
      }
      const long long GL = GLc + pg, HL = HLc + ph;
      const int b = b_lo + q * 32 + lane;
      if (b < hi && (e[q].x != 0 || e[q].y != 0) && HL >= hmin_q && (t.Ht - HL) >= hmin_q) {
        const float glf = (float)GL * isg, hlf = (float)HL * ish;
        const float est = k.mds == 0.0f ? split_est<false>(k, glf, hlf, Gt_f - glf, Ht_f - hlf) : split_est<true>(k, glf, hlf, Gt_f - glf, Ht_f - hlf);
        if (est >= cut) exact_candidate(best, jp, t, GL, HL, f_col, b, inv_sg, inv_sh);
      }
      GLc += __shfl_sync(0xffffffffu, pg, 31);
      HLc += __shfl_sync(0xffffffffu, ph, 31);
    }
  }
  warp_merge<16>(best);
  return best;
}

// Stage 2 for one child, called by all 32 lanes of a warp.  amax is the child's best estimate over (at least) this warp's
// lanes; [b_lo, hi) is the lane's bin range and (GL0, HL0) the prefix in front
// of it.  Returns the lane's exact best candidate.
template <bool COOP>
__device__ __forceinline__ Cand lane_exact(const JobParams& jp, const ChildTot& t, const LaneRes& c, int side, float amax, int f,
                                           int b_lo, int hi, long long GL0, long long HL0, const HistEntry* hbuilt, bool nar_b,
                                           const HistEntry* hpar, bool nar_p, size_t goff, long long hmin_q, float isg, float ish,
                                           double inv_sg, double inv_sh) {
  const int lane = threadIdx.x & 31;
  Cand best{0.0f, -1, -1, 0, 0};
  // estimate error: a few fp32 roundings on its two terms (the right-hand sums come from an fp32 subtraction from
  // the node totals), plus the root gain.  Both terms are >= 0, so no candidate's terms exceed the best estimate.
  const float margin = 3.2e-5f * (fmaxf(amax, 0.f) + fabsf(t.root_gain)) + 1e-12f;
  // a candidate can win only with est close to the best estimate and with a positive gain (est > root gain)
  const float cut = fmaxf(amax, t.root_gain) - 2.0f * margin;
  const bool alive = c.e1 >= cut;
  if (lane == 0) SCAN_STAT(0, 1);
  SCAN_STAT(1, alive ? 1 : 0);
  // the usual case: one candidate of this lane survives; evaluate it exactly from the saved prefix
  if (alive && c.e2 < cut) exact_candidate(best, jp, t, c.GL1, c.HL1, f, c.b1, inv_sg, inv_sh);
  if (COOP) {
    // several near-ties in a lane: the warp re-walks that lane's column together (256 bins per lane in the many-pairs
    // kernel: a per-lane walk would leave the pair's other warp waiting for tens of microseconds)
    unsigned walk = __ballot_sync(0xffffffffu, alive && c.e2 >= cut);
    while (walk) {
      const int src = __ffs(walk) - 1;
      if (lane == 0) SCAN_STAT(2, 1);
      walk &= walk - 1;
      const int hi_s = __shfl_sync(0xffffffffu, hi, src);
      const long long GL0_s = __shfl_sync(0xffffffffu, GL0, src), HL0_s = __shfl_sync(0xffffffffu, HL0, src);
      const Cand col = column_exact(jp, t, c.Gt_f, c.Ht_f, side, cut, f - lane + src, b_lo, hi_s, GL0_s, HL0_s, hbuilt, nar_b, hpar, nar_p,
                                    goff - (size_t)lane + (size_t)src, hmin_q, isg, ish, inv_sg, inv_sh);
      if (lane == src) best = col;
    }
  } else if (alive && c.e2 >= cut) {
    // several near-ties in this lane: walk its (at most 128) bins again in scan order (an empty bin repeats the
    // previous candidate, which never wins under strict >)
    const GainK k{jp.alpha, jp.lambda, jp.mds};
    long long GL = GL0, HL = HL0;
    for (int b = b_lo; b < hi; ++b) {
      longlong2 e = load_cell(hbuilt, nar_b, goff + (size_t)b * kLanes);
      if (side == 1) { const longlong2 pe = load_cell(hpar, nar_p, goff + (size_t)b * kLanes); e.x = pe.x - e.x; e.y = pe.y - e.y; }
      GL += e.x; HL += e.y;
      if ((e.x != 0 || e.y != 0) && HL >= hmin_q && (t.Ht - HL) >= hmin_q) {
        const float glf = (float)GL * isg, hlf = (float)HL * ish;
        const float est = k.mds == 0.0f ? split_est<false>(k, glf, hlf, c.Gt_f - glf, c.Ht_f - hlf)
                                        : split_est<true>(k, glf, hlf, c.Gt_f - glf, c.Ht_f - hlf);
        if (est >= cut) exact_candidate(best, jp, t, GL, HL, f, b, inv_sg, inv_sh);
      }
    }
  }
  return best;
}


// ---- stage 1 proper.  A lane walks the bins [b0, b1) of its feature for both children of a pair.  The latency of
// the histogram reads is covered by a per-lane ring in shared memory filled with cp.async (kScanAhead bins in
// flight per lane, no registers held); sibling = parent - built is formed, stored and scanned in the same step.
#ifndef GT_SCAN_AHEAD
#define GT_SCAN_AHEAD 3
#endif
#ifndef GT_DEEP_CTAS
#define GT_DEEP_CTAS 2
#endif
#ifndef GT_SCAN_CTAS
#define GT_SCAN_CTAS 2
#endif
constexpr int kScanAhead = GT_SCAN_AHEAD;
constexpr int kScanRing = kScanAhead + 1;   // the slot refilled in an iteration is the one consumed in the previous iteration
static_assert((kScanRing & (kScanRing - 1)) == 0, "ring slots are addressed with a mask");
constexpr int kRingWarpBytes = kScanRing * 2 * kLanes * 16;   // [slot][built | parent][lane] 16-byte cells

__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async8(uint32_t dst, const void* src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// A cell out of the ring.  32-bit walks (T = int) take the low words of an int64 cell: they are used only when both
// children's sums fit 32 bits, and then parent - built is right modulo 2^32 whatever the parent's own magnitude.
template <typename T>
__device__ __forceinline__ void ring_cell(const uint4 v, bool narrow, T& x, T& y) {
  if (narrow) { x = (T)(int)v.x; y = (T)(int)v.y; return; }
  if (sizeof(T) == 4) { x = (T)(int)v.x; y = (T)(int)v.z; return; }
  x = (T)(long long)(((unsigned long long)v.y << 32) | v.x);
  y = (T)(long long)(((unsigned long long)v.w << 32) | v.z);
}

struct WalkArgs {
  const char* src_b; const char* src_p;   // this lane's bin-0 cell of the built child / the parent
  HistEntry* hsib;
  size_t goff;
  int b0, b1, hi;                         // bins to walk (warp-uniform); this lane's feature is a candidate in bins < hi
  long long GLb0, HLb0, GLs0, HLs0;       // prefix in front of b0: built child, sibling
  bool nar_b, nar_p, nar_s, has_sib, write_sib;
};

template <typename T, bool MDS>
__device__ __forceinline__ void walk_stage1(const GainK k, const ChildTot& tb, const ChildTot& ts, long long hmin_q, float isg, float ish,
                                            const WalkArgs& a, uint4* ring, uint32_t ring_s, LaneRes& rb, LaneRes& rs) {
  ChildAcc<T> cb, cs;
  acc_init(cb, tb, k, hmin_q, isg, ish, a.GLb0, a.HLb0);
  acc_init(cs, ts, k, hmin_q, isg, ish, a.GLs0, a.HLs0);
  const int b0 = a.b0, b1 = a.b1, hi = a.hi;
  const bool nar_b = sizeof(T) == 4 ? true : a.nar_b, nar_p = a.nar_p, has_sib = a.has_sib;
  const unsigned step_b = kLanes * (nar_b ? 8 : 16), step_p = kLanes * (nar_p ? 8 : 16);
  const char* src_b = a.src_b;
  const char* src_p = a.src_p;
  auto request = [&](int b, int slot) {
    if (b < b1) {
      const uint32_t dst = ring_s + (uint32_t)slot * (2 * kLanes * 16);
      if (nar_b) cp_async8(dst, src_b + (size_t)b * step_b); else cp_async16(dst, src_b + (size_t)b * step_b);
      if (has_sib) {
        if (nar_p) cp_async8(dst + kLanes * 16, src_p + (size_t)b * step_p); else cp_async16(dst + kLanes * 16, src_p + (size_t)b * step_p);
      }
    }
    cp_async_commit();   // a group per bin, empty past the end, so the wait below always counts the same
  };
#pragma unroll
  for (int d = 0; d < kScanAhead; ++d) request(b0 + d, d);
  for (int b = b0, n = 0; b < b1; ++b, ++n) {
    const int slot = n & (kScanRing - 1), slot_in = (n + kScanAhead) & (kScanRing - 1);
    cp_async_wait<kScanAhead - 1>();
    T bx, by, px = 0, py = 0;
    ring_cell(ring[slot * (2 * kLanes)], nar_b, bx, by);
    if (has_sib) ring_cell(ring[slot * (2 * kLanes) + kLanes], nar_p, px, py);
    request(b + kScanAhead, slot_in);
    const T sx = px - bx, sy = py - by;
    if (a.write_sib) store_cell(a.hsib, a.nar_s, a.goff + (size_t)b * kLanes, make_longlong2((long long)sx, (long long)sy));
    // a bin that is empty for every feature of the warp (common in small nodes) changes no prefix and offers no
    // candidate: skip both evaluations.  (If the parent's cell is zero, so are both children's.)
    const bool occupied = has_sib ? ((px | py | bx | by) != 0) : ((bx | by) != 0);
    if (__any_sync(0xffffffffu, occupied)) {
      const bool in_range = b < hi;
      child_step<T, MDS>(cb, k, bx, by, b, in_range, isg, ish);
      if (has_sib) child_step<T, MDS>(cs, k, sx, sy, b, in_range, isg, ish);
    }
  }
  cp_async_wait<0>();
  acc_result(cb, rb);
  acc_result(cs, rs);
}

__device__ __forceinline__ void walk_dispatch(const GainK k, const ChildTot& tb, const ChildTot& ts, long long hmin_q, float isg, float ish,
                                              const WalkArgs& a, uint4* ring, uint32_t ring_s, LaneRes& rb, LaneRes& rs) {
  const bool all32 = a.nar_b && (!a.has_sib || a.nar_s);   // both children's sums fit 32 bits (the parent's need not)
  if (k.mds == 0.0f) {
    if (all32) walk_stage1<int, false>(k, tb, ts, hmin_q, isg, ish, a, ring, ring_s, rb, rs);
    else walk_stage1<long long, false>(k, tb, ts, hmin_q, isg, ish, a, ring, ring_s, rb, rs);
  } else {
    if (all32) walk_stage1<int, true>(k, tb, ts, hmin_q, isg, ish, a, ring, ring_s, rb, rs);
    else walk_stage1<long long, true>(k, tb, ts, hmin_q, isg, ish, a, ring, ring_s, rb, rs);
  }
}

template <typename T>
__device__ __forceinline__ void cell_as(const longlong2 e, T& x, T& y) { x = (T)e.x; y = (T)e.y; }

// Stage 1 of the chunked kernel for one lane: bins [b_lo, hi) of its feature, both children, prefix (GLx0, HLx0) in
// front.  A chunk is at most 128 bins and the CTA runs at full occupancy, so a one-bin register prefetch is enough here
// (the shared-memory ring of walk_stage1 was measured slower in this kernel: it costs the second resident CTA).
template <typename T, bool MDS>
__device__ __forceinline__ void chunk_stage1(const GainK k, const ChildTot& tb, const ChildTot& ts, long long hmin_q, float isg, float ish,
                                             long long GLb0, long long HLb0, long long GLs0, long long HLs0, const HistEntry* hbuilt,
                                             bool nar_b, const HistEntry* hpar, bool nar_p, bool has_sib, size_t goff, int b_lo, int hi,
                                             LaneRes& rb, LaneRes& rs) {
  ChildAcc<T> cb, cs;
  acc_init(cb, tb, k, hmin_q, isg, ish, GLb0, HLb0);
  acc_init(cs, ts, k, hmin_q, isg, ish, GLs0, HLs0);
  // the loop body is long and in-order, so the next bin's cells are requested before this bin is evaluated
  longlong2 eb = make_longlong2(0, 0), pe = make_longlong2(0, 0);
  if (b_lo < hi) {
    eb = load_cell(hbuilt, nar_b, goff + (size_t)b_lo * kLanes);
    if (has_sib) pe = load_cell(hpar, nar_p, goff + (size_t)b_lo * kLanes);
  }
  for (int b = b_lo; b < hi; ++b) {
    longlong2 eb_n = make_longlong2(0, 0), pe_n = make_longlong2(0, 0);
    if (b + 1 < hi) {
      eb_n = load_cell(hbuilt, nar_b, goff + (size_t)(b + 1) * kLanes);
      if (has_sib) pe_n = load_cell(hpar, nar_p, goff + (size_t)(b + 1) * kLanes);
    }
    // a bin that is empty for every feature of the warp (common in small nodes) changes no prefix and offers no
    // candidate: skip both evaluations.  (If the parent's cell is zero, so are both children's.)
    const bool occupied = has_sib ? ((pe.x | pe.y | eb.x | eb.y) != 0) : ((eb.x | eb.y) != 0);
    if (__any_sync(__activemask(), occupied)) {
      T bx, by, sx, sy;
      cell_as(eb, bx, by);
      cell_as(make_longlong2(pe.x - eb.x, pe.y - eb.y), sx, sy);
      child_step<T, MDS>(cb, k, bx, by, b, true, isg, ish);
      if (has_sib) child_step<T, MDS>(cs, k, sx, sy, b, true, isg, ish);
    }
    eb = eb_n; pe = pe_n;
  }
  acc_result(cb, rb);
  acc_result(cs, rs);
}

// One CTA per (job, parent).  Warps = feature groups x bin chunks, lane = feature; every warp scans BOTH
// children of the parent: the built child's cells are read once, the sibling's are formed on the fly as
// parent - built (and stored when a later level needs them as a parent).
// Pass 1: chunk sums.  Pass 2: with the chunk prefix from shared memory each lane runs EnumerateSplit<+1>
// over its bins (stage 1 above), then stage 2.
__device__ __forceinline__ void scan_pair(const Wave& w, int level, const PairRec pr) {
  const int j = pr.job;
  const JobState& js = w.js[j];
  const JobParams& jp = w.jp[j];
  const int G = w.groups, NCH = scan_chunks(G), CB = kBins / NCH;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int grp = warp / NCH, ch = warp % NCH;
  Node* nodes = w.nodes + (size_t)j * kNodeCap;
  TreeNode* tree = w.tree + (size_t)j * kNodeCap;
  const bool has_sib = level > 0 && !pr.single;
  const int bnode = level == 0 ? 0 : 2 * pr.parent + 1 + pr.build_right;       // histogram built by k_hist
  const int snode = level == 0 ? 0 : 2 * pr.parent + 2 - pr.build_right;       // its sibling
  const FoldDev& fd = w.folds[jp.fold];
  const int f = grp * kLanes + lane;
  const int nc = f < w.f ? fd.ncuts[f] : 0;
  const bool factive = f < w.f && ((js.lmask[f >> 5] >> (f & 31)) & 1u);
  const int maxnc = w.grp_maxnc[jp.fold * G + grp];
  const bool write_sib = has_sib && (level + 1 < jp.max_depth);
  const size_t goff = (size_t)grp * (kBins * kLanes) + lane;
  const HistEntry* hbuilt = node_hist(w, jp, level, bnode);
  const HistEntry* hpar = has_sib ? node_hist(w, jp, level - 1, pr.parent) : nullptr;
  HistEntry* hsib = node_hist(w, jp, level, snode);
  const bool nar_b = nodes[bnode].narrow, nar_s = nodes[snode].narrow;
  const bool nar_p = has_sib ? (bool)nodes[pr.parent].narrow : false;
  const double inv_sg = js.inv_sg, inv_sh = js.inv_sh;
  const float isg = (float)inv_sg, ish = (float)inv_sh;
  const int b_lo = ch * CB, b_hi = min(b_lo + CB, maxnc);

  ChildTot tb{nodes[bnode].G, nodes[bnode].H, nodes[bnode].root_gain};
  const ChildTot ts{nodes[snode].G, nodes[snode].H, nodes[snode].root_gain};
  if (level == 0) tb.root_gain = calc_gain(jp, (double)tb.Gt * inv_sg, (double)tb.Ht * inv_sh);

  __shared__ longlong2 s_sum[16][2][kLanes];   // chunk sums per warp: built, sibling
  __shared__ float s_est[16][2];
  __shared__ Cand s_c[16][2];
  {
    longlong2 ab = make_longlong2(0, 0), as = make_longlong2(0, 0);
#pragma unroll 8
    for (int b = b_lo; b < b_hi; ++b) {
      const longlong2 eb = load_cell(hbuilt, nar_b, goff + (size_t)b * kLanes);
      ab.x += eb.x; ab.y += eb.y;
      if (has_sib) {
        const longlong2 pe = load_cell(hpar, nar_p, goff + (size_t)b * kLanes);
        const longlong2 es = make_longlong2(pe.x - eb.x, pe.y - eb.y);
        if (write_sib) store_cell(hsib, nar_s, goff + (size_t)b * kLanes, es);
        as.x += es.x; as.y += es.y;
      }
    }
    s_sum[warp][0][lane] = ab;
    s_sum[warp][1][lane] = as;
  }
  __syncthreads();
  long long GLb0 = 0, HLb0 = 0, GLs0 = 0, HLs0 = 0;
  for (int k = 0; k < ch; ++k) {
    const longlong2 tb0 = s_sum[warp - ch + k][0][lane], ts0 = s_sum[warp - ch + k][1][lane];
    GLb0 += tb0.x; HLb0 += tb0.y; GLs0 += ts0.x; HLs0 += ts0.y;
  }
  // hl >= mcw  <=>  HL_q >= mcw * sh (exact: sh is a power of two)
  const long long hmin_q = (long long)ceil((double)jp.mcw * (double)js.sh);
  const int hi = factive ? min(b_hi, nc) : b_lo;
  LaneRes rb, rs;
  {
    const GainK k{jp.alpha, jp.lambda, jp.mds};
    const bool all32 = nar_b && (!has_sib || nar_s);   // both children's sums fit 32 bits (the parent's need not)
    if (k.mds == 0.0f) {
      if (all32) chunk_stage1<int, false>(k, tb, ts, hmin_q, isg, ish, GLb0, HLb0, GLs0, HLs0, hbuilt, nar_b, hpar, nar_p, has_sib, goff, b_lo, hi, rb, rs);
      else chunk_stage1<long long, false>(k, tb, ts, hmin_q, isg, ish, GLb0, HLb0, GLs0, HLs0, hbuilt, nar_b, hpar, nar_p, has_sib, goff, b_lo, hi, rb, rs);
    } else {
      if (all32) chunk_stage1<int, true>(k, tb, ts, hmin_q, isg, ish, GLb0, HLb0, GLs0, HLs0, hbuilt, nar_b, hpar, nar_p, has_sib, goff, b_lo, hi, rb, rs);
      else chunk_stage1<long long, true>(k, tb, ts, hmin_q, isg, ish, GLb0, HLb0, GLs0, HLs0, hbuilt, nar_b, hpar, nar_p, has_sib, goff, b_lo, hi, rb, rs);
    }
  }
  float v0 = rb.e1, v2 = rs.e1;
  for (int d = 16; d; d >>= 1) {
    v0 = fmaxf(v0, __shfl_xor_sync(0xffffffffu, v0, d)); v2 = fmaxf(v2, __shfl_xor_sync(0xffffffffu, v2, d));
  }
  if (lane == 0) { s_est[warp][0] = v0; s_est[warp][1] = v2; }
  __syncthreads();
  v0 = v2 = -3.0e38f;
  for (int k = 0; k < G * NCH; ++k) { v0 = fmaxf(v0, s_est[k][0]); v2 = fmaxf(v2, s_est[k][1]); }
  Cand bestc[2];
#pragma unroll
  for (int side = 0; side < 2; ++side) {
    Cand best{0.0f, -1, -1, 0, 0};
    if (side == 0 || has_sib)
      best = lane_exact<false>(jp, side == 0 ? tb : ts, side == 0 ? rb : rs, side, side == 0 ? v0 : v2, f, b_lo, hi,
                        side == 0 ? GLb0 : GLs0, side == 0 ? HLb0 : HLs0, hbuilt, nar_b, hpar, nar_p, goff, hmin_q, isg, ish, inv_sg, inv_sh);
    warp_merge<16>(best);
    bestc[side] = best;
  }
  if (lane == 0) { s_c[warp][0] = bestc[0]; s_c[warp][1] = bestc[1]; }
  __syncthreads();
  // warp 0 finishes the built child, warp 1 the sibling: lanes merge the warps' candidates with shuffles,
  // then lanes 0 and 1 each set up one grandchild
  const int side = warp;
  if (side > 1 || (side == 1 && !has_sib)) return;
  Cand best = lane < G * NCH ? s_c[lane][side] : Cand{0.0f, -1, -1, 0, 0};
  warp_merge<8>(best);
  best.chg = __shfl_sync(0xffffffffu, best.chg, 0); best.f = __shfl_sync(0xffffffffu, best.f, 0);
  best.b = __shfl_sync(0xffffffffu, best.b, 0);
  best.GL = __shfl_sync(0xffffffffu, best.GL, 0); best.HL = __shfl_sync(0xffffffffu, best.HL, 0);
  if (lane > 1) return;
  if (side == 1 && lane == 0) nodes[snode].absG = nodes[pr.parent].absG - nodes[bnode].absG;   // exact: both are sums over rows
  settle_node(jp, level, nodes, tree, side == 0 ? bnode : snode, side == 0 ? tb : ts, best, inv_sg, inv_sh, hmin_q, lane);
}

// Levels with few pairs use k_scan (a whole CTA per pair, bins split over warps); levels with at least
// kDeepPairs pairs use k_scan_deep (one warp per pair and feature group).  Both are launched every level
// and decide from the device-side pair count.
#ifndef GT_DEEP_PAIRS
#define GT_DEEP_PAIRS 2048
#endif
constexpr int kDeepPairs = GT_DEEP_PAIRS;

__global__ void __launch_bounds__(512, GT_SCAN_CTAS) k_scan(Wave w, int level) {
  const int n_pairs = (int)w.counters[5 + (level & 1)];
  if (n_pairs >= kDeepPairs) return;
  for (int pair_idx = blockIdx.x; pair_idx < n_pairs; pair_idx += gridDim.x) {
    scan_pair(w, level, w.pairs[pair_idx]);
    __syncthreads();   // shared scratch is reused by the next pair
  }
}

// ---- the many-pairs variant.  With thousands of (job, parent) pairs in a level there is no need to split one pair's
// bins over warps: one warp owns (pair, feature group) and walks all bins of its 32 features ONCE, so every
// histogram cell is read exactly once and no chunk sums are needed.
constexpr int kDeepWarps = 8;
// One CTA = one pair, one warp per feature group: the only barriers are between the warps of that pair, and the
// hardware CTA scheduler balances pairs of different cost (narrow / wide cells, single children, empty bins).
__global__ void __launch_bounds__(kDeepWarps * 32, GT_DEEP_CTAS) k_scan_deep(Wave w, int level) {
  extern __shared__ uint4 s_ring[];                       // [warp][slot][built | parent][lane]
  __shared__ Cand s_c[kDeepWarps][2];
  const int n_pairs = (int)w.counters[5 + (level & 1)];
  if (n_pairs < kDeepPairs || (int)blockIdx.x >= n_pairs) return;
  const int G = w.groups;                                  // == warps of this CTA (G <= 8 by kMaskWords)
  const int grp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint4* ring = s_ring + (size_t)grp * (kRingWarpBytes / 16) + lane;
  const uint32_t ring_s = (uint32_t)__cvta_generic_to_shared(ring);
  const PairRec pr = w.pairs[blockIdx.x];
  const int j = pr.job;
  const JobState& js = w.js[j];
  const JobParams& jp = w.jp[j];
  Node* nodes = w.nodes + (size_t)j * kNodeCap;
  const bool has_sib = level > 0 && !pr.single;
  const int bnode = level == 0 ? 0 : 2 * pr.parent + 1 + pr.build_right;
  const int snode = level == 0 ? 0 : 2 * pr.parent + 2 - pr.build_right;
  const int f = grp * kLanes + lane;
  const int nc = f < w.f ? w.folds[jp.fold].ncuts[f] : 0;
  const bool factive = f < w.f && ((js.lmask[f >> 5] >> (f & 31)) & 1u);
  const int hi = factive ? nc : 0;
  const size_t goff = (size_t)grp * (kBins * kLanes) + lane;
  const bool nar_b = nodes[bnode].narrow, nar_p = has_sib ? (bool)nodes[pr.parent].narrow : false;
  const HistEntry* hbuilt = node_hist(w, jp, level, bnode);
  const HistEntry* hpar = has_sib ? node_hist(w, jp, level - 1, pr.parent) : nullptr;
  const long long hmin_q = (long long)ceil((double)jp.mcw * (double)js.sh);
  LaneRes rb, rs;
  {
    WalkArgs a;
    a.has_sib = has_sib;
    a.b0 = 0; a.b1 = w.grp_maxnc[jp.fold * G + grp]; a.hi = hi;
    a.GLb0 = a.HLb0 = a.GLs0 = a.HLs0 = 0;
    a.write_sib = has_sib && (level + 1 < jp.max_depth);
    a.goff = goff;
    a.nar_b = nar_b; a.nar_p = nar_p;
    a.nar_s = nodes[snode].narrow;
    a.src_b = reinterpret_cast<const char*>(hbuilt) + goff * (nar_b ? 8 : 16);
    a.src_p = has_sib ? reinterpret_cast<const char*>(hpar) + goff * (nar_p ? 8 : 16) : nullptr;
    a.hsib = node_hist(w, jp, level, snode);
    const float isg = (float)js.inv_sg, ish = (float)js.inv_sh;
    ChildTot tb{nodes[bnode].G, nodes[bnode].H, nodes[bnode].root_gain};
    const ChildTot ts{nodes[snode].G, nodes[snode].H, nodes[snode].root_gain};
    if (level == 0) tb.root_gain = calc_gain(jp, (double)tb.Gt * js.inv_sg, (double)tb.Ht * js.inv_sh);
    walk_dispatch(GainK{jp.alpha, jp.lambda, jp.mds}, tb, ts, hmin_q, isg, ish, a, ring, ring_s, rb, rs);
  }
  float v0 = rb.e1, v2 = rs.e1;
  for (int d = 16; d; d >>= 1) {
    v0 = fmaxf(v0, __shfl_xor_sync(0xffffffffu, v0, d)); v2 = fmaxf(v2, __shfl_xor_sync(0xffffffffu, v2, d));
  }
  // stage 2 prunes with this warp's own best estimate: a weaker bound than the pair-wide one (a few more exact
  // evaluations), but the warps of the pair meet only once, at the end
  // stage 2 and the node updates re-read the per-pair values they need: only the lane results live across the bin loop
  const double inv_sg = js.inv_sg, inv_sh = js.inv_sh;
  const float isg = (float)inv_sg, ish = (float)inv_sh;
  Cand bestc[2] = {Cand{0.0f, -1, -1, 0, 0}, Cand{0.0f, -1, -1, 0, 0}};
#pragma unroll
  for (int side = 0; side < 2; ++side) {
    if (side == 0 || has_sib) {
      const int node = side == 0 ? bnode : snode;
      ChildTot t{nodes[node].G, nodes[node].H, nodes[node].root_gain};
      if (level == 0) t.root_gain = calc_gain(jp, (double)t.Gt * inv_sg, (double)t.Ht * inv_sh);
      Cand best = lane_exact<true>(jp, t, side == 0 ? rb : rs, side, side == 0 ? v0 : v2, f, 0, hi, 0, 0,
                             hbuilt, nar_b, hpar, nar_p, goff, hmin_q, isg, ish, inv_sg, inv_sh);
      warp_merge<16>(best);
      bestc[side] = best;
    }
  }
  if (lane == 0) { s_c[grp][0] = bestc[0]; s_c[grp][1] = bestc[1]; }
  __syncthreads();
  // the first warp finishes both children: lanes 0,1 the built child's two grandchildren, lanes 2,3 the sibling's
  if (grp == 0 && lane < (has_sib ? 4 : 2)) {
    const int side = lane >> 1;
    Cand best = s_c[0][side];
    for (int k = 1; k < G; ++k) {
      const Cand o = s_c[k][side];
      if (better(o.chg, o.f, o.b, best.chg, best.f, best.b)) best = o;
    }
    const int node = side == 0 ? bnode : snode;
    ChildTot t{nodes[node].G, nodes[node].H, nodes[node].root_gain};
    if (level == 0) t.root_gain = calc_gain(jp, (double)t.Gt * inv_sg, (double)t.Ht * inv_sh);
    if (lane == 2) nodes[snode].absG = nodes[pr.parent].absG - nodes[bnode].absG;   // exact: both are sums over rows
    settle_node(jp, level, nodes, w.tree + (size_t)j * kNodeCap, node, t, best, inv_sg, inv_sh, hmin_q, lane & 1);
  }
}

// ------------------------------------------------------------------ X7: stable row partition
__global__ void __launch_bounds__(256) k_plan_part(Wave w, int level) {
  const int j = blockIdx.x;
  JobState& js = w.js[j];
  const JobParams& jp = w.jp[j];
  if (!js.active || level + 1 >= jp.max_depth) { if (threadIdx.x == 0) js.n_part_items = 0; return; }
  const Node* nodes = w.nodes + (size_t)j * kNodeCap;
  PartItem* items = w.pitems + (size_t)j * w.part_cap;
  __shared__ int s_scan[256];
  const int n_lvl = 1 << level, first = level_first(level);
  int nid[2], chunks[2];
  for (int s2 = 0; s2 < 2; ++s2) {
    const int li = 2 * threadIdx.x + s2;
    nid[s2] = -1; chunks[s2] = 0;
    if (li >= n_lvl) continue;
    const Node& nd = nodes[first + li];
    if (nd.state != kSplit) continue;
    if (nodes[2 * (first + li) + 1].state != kOpen && nodes[2 * (first + li) + 2].state != kOpen) continue;   // no child needs its rows
    nid[s2] = first + li;
    chunks[s2] = (int)((nd.cnt + kPartRows - 1) / kPartRows);
  }
  const int v = chunks[0] + chunks[1];
  s_scan[threadIdx.x] = v;
  __syncthreads();
  for (int d = 1; d < 256; d <<= 1) {
    const int t = threadIdx.x >= d ? s_scan[threadIdx.x - d] : 0;
    __syncthreads();
    s_scan[threadIdx.x] += t;
    __syncthreads();
  }
  int ni = s_scan[threadIdx.x] - v;
  for (int s2 = 0; s2 < 2; ++s2) {
    if (nid[s2] < 0) continue;
    const Node& nd = nodes[nid[s2]];
    const int first_item = ni;
    for (int c = 0; c < chunks[s2]; ++c, ++ni) {
      if (ni >= w.part_cap) { w.counters[7] = 2ull; break; }
      const uint32_t b = (uint32_t)c * kPartRows;
      const uint32_t cnt = nd.cnt - b < (uint32_t)kPartRows ? nd.cnt - b : (uint32_t)kPartRows;
      items[ni] = PartItem{nid[s2], nd.begin + b, cnt, first_item, chunks[s2], nd.begin, nd.cnt, (int)nd.buf};
    }
  }
  if (threadIdx.x == 255) {
    js.n_part_items = min(s_scan[255], w.part_cap);
    if (s_scan[255] > part_items_bound(w, level)) w.counters[7] = 2ull;
  }
}

// The split feature's bin of a row comes from the feature-major copy of the bin matrix: rows of a node
// are ascending, so neighbouring rows share DRAM lines (1 useful byte per 64-byte line otherwise).
// Two passes over a node's rows in items of kPartRows rows.  Pass 1 gathers the split feature's bin byte of every row,
// keeps the goes-left decision as one bit per row and counts the lefts of the item; pass 2 turns the bits into
// stable destinations (item-local ranks from the bits alone, item offsets from the counts of the node's earlier
// items) and moves the row ids.  Items are large so that a CTA lives long enough to amortise its start.
constexpr int kPartWords = kPartRows / 32;   // goes-left words per item
static_assert(kPartWords % 8 == 0, "one word per warp and iteration");

__global__ void __launch_bounds__(256) k_part_count(Wave w, int level) {
  const int j = blockIdx.y;
  const JobState& js = w.js[j];
  if ((int)blockIdx.x >= js.n_part_items) return;
  const JobParams& jp = w.jp[j];
  const PartItem it = w.pitems[(size_t)j * w.part_cap + blockIdx.x];
  const Node& nd = w.nodes[(size_t)j * kNodeCap + it.node];
  const uint32_t sbin = nd.bin;
  const uint8_t* col = w.folds[jp.fold].bins_col + (size_t)nd.feat * (size_t)w.n;
  const uint32_t* src = w.ridx + ((size_t)j * 2 + it.buf) * w.nt_stride + it.begin;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  uint32_t* flags = w.pflags + ((size_t)j * w.part_cap + blockIdx.x) * kPartWords;
  int cnt = 0;
  const int iters = (int)((it.cnt + 255u) / 256u);   // CTA-uniform
  for (int t0 = 0; t0 < iters; t0 += 8) {
    uint32_t rows[8];
    uint32_t bins[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      const uint32_t p = (uint32_t)(t0 + q) * 256u + threadIdx.x;
      rows[q] = p < it.cnt ? src[p] : 0xffffffffu;
    }
#pragma unroll
    for (int q = 0; q < 8; ++q) bins[q] = rows[q] != 0xffffffffu ? (uint32_t)col[rows[q]] : 0xffffffffu;
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      const uint32_t word = __ballot_sync(0xffffffffu, bins[q] <= sbin);   // rows past the end compare as 2^32 - 1: right
      if (t0 + q < iters) {
        if (lane == 0) flags[(t0 + q) * 8 + wid] = word;
        cnt += __popc(word);
      }
    }
  }
  __shared__ int s_cnt[8];
  if (lane == 0) s_cnt[wid] = cnt;
  __syncthreads();
  if (threadIdx.x == 0) {
    int tot = 0;
    for (int q = 0; q < 8; ++q) tot += s_cnt[q];
    w.pleft[(size_t)j * w.part_cap + blockIdx.x] = tot;
  }
}

__global__ void __launch_bounds__(256) k_part_scatter(Wave w, int level) {
  const int j = blockIdx.y;
  const JobState& js = w.js[j];
  if ((int)blockIdx.x >= js.n_part_items) return;
  const PartItem it = w.pitems[(size_t)j * w.part_cap + blockIdx.x];
  const uint32_t* flags = w.pflags + ((size_t)j * w.part_cap + blockIdx.x) * kPartWords;
  const int* pleft = w.pleft + (size_t)j * w.part_cap;
  const uint32_t* src = w.ridx + ((size_t)j * 2 + it.buf) * w.nt_stride + it.begin;
  uint32_t* dst = w.ridx + ((size_t)j * 2 + (it.buf ^ 1)) * w.nt_stride;
  __shared__ int s_pref[kPartWords];   // lefts of the item in front of each 32-row word
  __shared__ int s_warp[8];
  __shared__ int s_red[2][8];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int words = (int)((it.cnt + 31u) / 32u);
  // lefts of the node's earlier items, and of the whole node
  int prevL = 0, totL = 0;
  for (int i = it.first_item + threadIdx.x; i < it.first_item + it.n_items; i += blockDim.x) {
    const int v = pleft[i];
    totL += v;
    if (i < (int)blockIdx.x) prevL += v;
  }
  for (int d = 16; d; d >>= 1) { prevL += __shfl_xor_sync(0xffffffffu, prevL, d); totL += __shfl_xor_sync(0xffffffffu, totL, d); }
  if (lane == 0) { s_red[0][wid] = prevL; s_red[1][wid] = totL; }
  // exclusive scan of the words' left counts (word index = iteration * 8 + warp, i.e. row order); a thread takes
  // kWordsPerThread consecutive words
  {
    constexpr int kWordsPerThread = (kPartWords + 255) / 256;
    int c[kWordsPerThread], mine = 0;
#pragma unroll
    for (int q = 0; q < kWordsPerThread; ++q) {
      const int i = (int)threadIdx.x * kWordsPerThread + q;
      c[q] = i < words ? __popc(flags[i]) : 0;
      mine += c[q];
    }
    int incl = mine;
    for (int d = 1; d < 32; d <<= 1) { const int o = __shfl_up_sync(0xffffffffu, incl, d); if (lane >= d) incl += o; }
    if (lane == 31) s_warp[wid] = incl;
    __syncthreads();
    int run = incl - mine;
    for (int q = 0; q < wid; ++q) run += s_warp[q];
#pragma unroll
    for (int q = 0; q < kWordsPerThread; ++q) {
      const int i = (int)threadIdx.x * kWordsPerThread + q;
      if (i < kPartWords) s_pref[i] = run;
      run += c[q];
    }
  }
  __syncthreads();
  prevL = 0; totL = 0;
  for (int q = 0; q < 8; ++q) { prevL += s_red[0][q]; totL += s_red[1][q]; }
  const uint32_t rel0 = it.begin - it.node_begin;
  const int iters = (int)((it.cnt + 255u) / 256u);
  // eight iterations at a time: all their loads first (the stores below go to the other half of the same array, so
  // the compiler would not move a load above an earlier store by itself)
  for (int t0 = 0; t0 < iters; t0 += 8) {
    uint32_t rows[8], wordv[8];
    int pref[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      const int t = t0 + q;
      const uint32_t p = (uint32_t)t * 256u + threadIdx.x;
      rows[q] = p < it.cnt ? src[p] : 0u;
      wordv[q] = t < iters ? flags[t * 8 + wid] : 0u;
      pref[q] = t < iters ? s_pref[t * 8 + wid] : 0;
    }
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      const uint32_t p = (uint32_t)(t0 + q) * 256u + threadIdx.x;
      if (p < it.cnt) {
        const int lrank = pref[q] + __popc(wordv[q] & ((1u << lane) - 1u));
        const bool left = (wordv[q] >> lane) & 1u;
        const uint32_t rel = rel0 + p;
        const uint32_t d = left ? it.node_begin + (uint32_t)(prevL + lrank)
                                : it.node_begin + (uint32_t)totL + (rel - (uint32_t)(prevL + lrank));
        dst[d] = rows[q];
      }
    }
  }
  if ((int)blockIdx.x == it.first_item && threadIdx.x == 0) {
    Node* nodes = w.nodes + (size_t)j * kNodeCap;
    Node& L = nodes[2 * it.node + 1];
    Node& R = nodes[2 * it.node + 2];
    L.begin = it.node_begin; L.cnt = (uint32_t)totL; L.buf = (uint8_t)(it.buf ^ 1);
    R.begin = it.node_begin + (uint32_t)totL; R.cnt = it.node_cnt - (uint32_t)totL; R.buf = (uint8_t)(it.buf ^ 1);
    const Node& P = nodes[it.node];
    L.narrow = child_narrow(P, L.cnt, L.H); R.narrow = child_narrow(P, R.cnt, R.H);
  }
}

// ------------------------------------------------------------------ X8 + X9: prediction update and metric
// Every row (train slice and test slice) walks the new tree on its bin bytes (bin <= split bin  <=>
// raw value < threshold); the prediction is updated in place, test rows add their
// metric term, train rows record the gradient range of the next round.
//
// One CTA = (fold, chunk of kPredRows matrix rows).  All jobs of a fold read the same bin matrix, so the CTA stages
// the chunk's bin rows in shared memory ONCE and then takes the fold's jobs one after the other: tree into shared
// memory (double-buffered, the next job's nodes are fetched into registers while this job's rows are walked), each
// thread walks its rows out of shared memory.  With one CTA per (job, chunk) reading bytes from global memory the
// kernel was bound by L1 wavefronts (ncu: L1 99 % busy), so the shared layouts are chosen for one wavefront per access:
//  - bin rows transposed, word (4 features) major: s_rows[word][row] -- the 32 consecutive rows of a warp hit 32
//    different banks whatever feature each lane asks for;
//  - tree nodes split into a 4-byte split word (feature, bin, state) and a separate leaf value, heap order: the
//    nodes of a level are contiguous, so levels 0-5 are conflict-free and only the deep levels see random conflicts.
__device__ __forceinline__ uint32_t lds_u32(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
}
#ifndef GT_PRED_THREADS
#define GT_PRED_THREADS 1024
#endif
constexpr int kPredThreads = GT_PRED_THREADS;   // kPredRows / kPredThreads rows per thread and job
static_assert(sizeof(TreeNode) == 8 && kNodeCap % kPredThreads == 0, "k_predict splits TreeNode into two 32-bit words");
constexpr int kPredRowBytes = 128 * 1024;   // staged bin rows: kPredRows x 64 bytes fit
__host__ __device__ inline int pred_sub_rows(int fpad) {   // rows staged at a time (a chunk takes kPredRows / this many rounds)
  int r = kPredRows;
  while (r > 256 && (long long)r * fpad > kPredRowBytes) r >>= 1;
  return r;
}
constexpr int kPredTreeBytes = 2 * kNodeCap * 8;   // two buffers of (split word, leaf value) x kNodeCap
__host__ __device__ inline int pred_smem_bytes(int fpad, int n_inds) { return pred_sub_rows(fpad) * fpad + kPredTreeBytes + n_inds * 12; }

__global__ void __launch_bounds__(kPredThreads, 1) k_predict(Wave w) {
  extern __shared__ uint32_t s_pred[];
  __shared__ double s_m[2][32];
  __shared__ float s_g[2][32], s_h[2][32];
  __shared__ int s_nact;
  const int fold = blockIdx.x, chunk = blockIdx.y;
  const FoldDev fd = w.folds[fold];
  const int fpad = w.fpad, upr = fpad >> 4;   // 16-byte units per bin row
  const int SR = pred_sub_rows(fpad);
  uint32_t* s_split = s_pred;                                              // [2][kNodeCap] walk words, see store_tree
  float* s_leaf = reinterpret_cast<float*>(s_pred + 2 * kNodeCap);         // [2][kNodeCap]
  uint32_t* s_rows = s_pred + 4 * kNodeCap;                                // [fpad / 4][SR]
  int* s_list = reinterpret_cast<int*>(s_rows + (size_t)(fpad >> 2) * SR); // active individuals of this wave, in order
  float* s_spw = reinterpret_cast<float*>(s_list + w.n_inds);              // their scale_pos_weight ...
  int* s_nodes = reinterpret_cast<int*>(s_spw + w.n_inds);                 // ... and node counts, by list position
  const uint32_t split_base = (uint32_t)__cvta_generic_to_shared(s_split);
  const uint32_t rows_base = (uint32_t)__cvta_generic_to_shared(s_rows) + threadIdx.x * 4u;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const long long c0 = (long long)chunk * kPredRows;
  const long long chunk_rows = min((long long)kPredRows, w.n - c0);
  constexpr int kNodesPerThread = kNodeCap / kPredThreads;
  constexpr int R = kPredRows / kPredThreads;   // rows per thread and job, walked together so their latencies overlap

  // everything the job loop needs about the jobs comes from shared memory: the CTA is alone on its SM and moves in
  // lock step, so a global-memory round trip inside the loop would be paid in full every time
  if (wid == 0) {
    int cnt = 0;
    for (int base = 0; base < w.n_inds; base += 32) {
      const int q = base + lane;
      const bool act = q < w.n_inds && w.js[q * w.nfold + fold].active;
      const unsigned m = __ballot_sync(0xffffffffu, act);
      if (act) {
        const int pos = cnt + __popc(m & ((1u << lane) - 1u));
        const JobParams& jp = w.jp[q * w.nfold + fold];
        s_list[pos] = q; s_spw[pos] = jp.spw; s_nodes[pos] = (2 << jp.max_depth) - 1;
      }
      cnt += __popc(m);
    }
    if (lane == 0) s_nact = cnt;
  }
  __syncthreads();
  const int n_act = s_nact;

  // job at list position `k` of the previous loop step left its per-warp partial results in s_*[par]: fold and publish
  auto publish = [&](int k, int par, bool first_sub) {
    if (wid != 0) return;
    constexpr int kWarps = kPredThreads / 32;
    float gm = lane < kWarps ? s_g[par][lane] : 0.f, hm = lane < kWarps ? s_h[par][lane] : 0.f;
    double ms = lane < kWarps ? s_m[par][lane] : 0.0;
    for (int d = 16; d; d >>= 1) {
      gm = fmaxf(gm, __shfl_xor_sync(0xffffffffu, gm, d));
      hm = fmaxf(hm, __shfl_xor_sync(0xffffffffu, hm, d));
      ms += __shfl_xor_sync(0xffffffffu, ms, d);
    }
    if (lane == 0) {
      const int j = s_list[k] * w.nfold + fold;
      if (gm > 0.f) atomicMax(&w.js[j].gmax_bits, __float_as_uint(gm));
      if (hm > 0.f) atomicMax(&w.js[j].hmax_bits, __float_as_uint(hm));
      double* mp = w.mpart + (size_t)j * w.pred_chunks + chunk;
      *mp = first_sub ? ms : *mp + ms;
    }
  };
  uint2 tn[kNodesPerThread];   // next job's nodes on their way from global to shared memory
  auto fetch_tree = [&](int k) {
    const uint2* src = reinterpret_cast<const uint2*>(w.tree + (size_t)(s_list[k] * w.nfold + fold) * kNodeCap);
    const int n_nodes = s_nodes[k];
#pragma unroll
    for (int q = 0; q < kNodesPerThread; ++q) {
      const int i = threadIdx.x + q * kPredThreads;
      tn[q] = i < n_nodes ? src[i] : make_uint2(0u, 0u);
    }
  };
  auto store_tree = [&](int buf) {
#pragma unroll
    for (int q = 0; q < kNodesPerThread; ++q) {
      const int i = threadIdx.x + q * kPredThreads;
      s_leaf[buf * kNodeCap + i] = __uint_as_float(tn[q].x);
      // walk word: bits 24-31 split bin, bit 23 = split node, bits 2-22 byte offset of the feature's word row in
      // s_rows, bits 0-1 the feature's byte inside that word -- each field one instruction away from its use
      const uint32_t feat = tn[q].y & 0xffffu, sbin = (tn[q].y >> 16) & 0xffu;
      const bool is_split = (tn[q].y >> 24) == (uint32_t)kSplit;
      s_split[buf * kNodeCap + i] = is_split ? (sbin << 24) | 0x800000u | (((feat >> 2) * (uint32_t)SR * 4u) & 0x7ffffcu) | (feat & 3u) : 0u;
    }
  };

  for (long long s0 = 0; s0 < chunk_rows; s0 += SR) {
    const int sub_rows = (int)min((long long)SR, chunk_rows - s0);
    const bool first_sub = s0 == 0;
    __syncthreads();   // the previous sub-chunk's rows and trees are no longer read
    for (int part = 0; part < upr; ++part)
      for (int lr = threadIdx.x; lr < sub_rows; lr += kPredThreads) {
        const uint4 v = *(reinterpret_cast<const uint4*>(fd.bins + (size_t)(c0 + s0 + lr) * fpad) + part);
        uint32_t* dst = s_rows + (size_t)(part * 4) * SR + lr;
        dst[0] = v.x; dst[SR] = v.y; dst[2 * SR] = v.z; dst[3 * SR] = v.w;
      }
    // per-row values that do not depend on the job
    float yv[R];
    bool is_train[R], have[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const int i = threadIdx.x + r * kPredThreads;
      have[r] = i < sub_rows;
      const long long row = c0 + s0 + i;
      yv[r] = have[r] ? fd.y[row] : 0.f;
      is_train[r] = row < fd.ntrain;
    }
    const bool any_train = c0 + s0 < fd.ntrain, any_test = c0 + s0 + sub_rows > fd.ntrain;   // CTA-uniform
    float pn[R];   // the next job's current predictions, requested a job ahead
    auto fetch_pred = [&](int k) {
      const float* pred = w.pred + (size_t)(s_list[k] * w.nfold + fold) * w.n + c0 + s0;
#pragma unroll
      for (int r = 0; r < R; ++r) pn[r] = have[r] ? pred[threadIdx.x + r * kPredThreads] : 0.f;
    };
    // with few (fold, chunk) pairs the job list is dealt over gridDim.z CTAs, so that a small matrix still fills the GPU
    const int kz = blockIdx.z, Z = gridDim.z;
    if (kz < n_act) { fetch_tree(kz); store_tree(0); fetch_pred(kz); }
    int it = 0;
    for (int k = kz; k < n_act; k += Z, ++it) {
      const int par = it & 1;
      __syncthreads();   // tree of this job is in buffer par; everyone has left the previous job; the rows are staged
      float p0[R];
#pragma unroll
      for (int r = 0; r < R; ++r) p0[r] = pn[r];
      if (k + Z < n_act) { fetch_tree(k + Z); fetch_pred(k + Z); }
      if (it > 0) publish(k - Z, par ^ 1, first_sub);
      const float spw = s_spw[k];
      const float* leaf = s_leaf + par * kNodeCap;
      int nid[R];
      uint32_t t[R];
      bool go = false;
      const uint32_t split_s = split_base + (uint32_t)par * (kNodeCap * 4u);
#pragma unroll
      for (int r = 0; r < R; ++r) { nid[r] = 0; t[r] = lds_u32(split_s); go |= (t[r] & 0x800000u) != 0u; }
      while (go) {
        go = false;
#pragma unroll
        for (int r = 0; r < R; ++r) {
          if (t[r] & 0x800000u) {
            const uint32_t word = lds_u32(rows_base + (t[r] & 0x7ffffcu) + (uint32_t)r * (kPredThreads * 4u));
            const uint32_t bin = __byte_perm(word, 0u, 0x4440u | (t[r] & 3u));
            nid[r] = 2 * nid[r] + 1 + (bin > __byte_perm(t[r], 0u, 0x4443u) ? 1 : 0);
            t[r] = lds_u32(split_s + (uint32_t)nid[r] * 4u);
            go |= (t[r] & 0x800000u) != 0u;
          }
        }
      }
      float* pred = w.pred + (size_t)(s_list[k] * w.nfold + fold) * w.n + c0 + s0;
      float gm = 0.f, hm = 0.f;
      double msum = 0.0;
#pragma unroll
      for (int r = 0; r < R; ++r) {
        if (!have[r]) continue;
        const float p = p0[r] + leaf[nid[r]];
        pred[threadIdx.x + r * kPredThreads] = p;
        const float y = yv[r];
        if (is_train[r]) {
          float g, h;
          grad_pair(w.objective, p, y, spw, g, h);
          gm = fmaxf(gm, fabsf(g)); hm = fmaxf(hm, fabsf(h));
        } else {
          const float pv = w.objective == 1 ? sigmoid_x(p) : p;
          float e;
          if (w.metric == 1) {     // EvalRowLogLoss
            const float eps = 1e-16f, pneg = 1.0f - pv;
            if (pv < eps) e = -y * logf(eps) - (1.0f - y) * logf(1.0f - eps);
            else if (pneg < eps) e = -y * logf(1.0f - eps) - (1.0f - y) * logf(eps);
            else e = -y * logf(pv) - (1.0f - y) * logf(pneg);
          } else {                 // EvalRowRMSE
            const float d = y - pv;
            e = d * d;
          }
          msum += (double)e;
        }
      }
      if (k + Z < n_act) store_tree(par ^ 1);
      // a chunk lies in the train slice, in the test slice, or (one chunk per fold) across the boundary: only the
      // reductions it needs.  |g|, |h| >= 0, so their float order is the order of their bit patterns.
      if (any_train) {
        gm = __uint_as_float(__reduce_max_sync(0xffffffffu, __float_as_uint(gm)));
        hm = __uint_as_float(__reduce_max_sync(0xffffffffu, __float_as_uint(hm)));
      }
      if (any_test)
        for (int d = 16; d; d >>= 1) msum += __shfl_xor_sync(0xffffffffu, msum, d);
      if (lane == 0) { s_g[par][wid] = gm; s_h[par][wid] = hm; s_m[par][wid] = msum; }
    }
    __syncthreads();
    if (it > 0) publish(kz + (it - 1) * Z, (it - 1) & 1, first_sub);
  }
}

// Small-matrix variant: one CTA per (job, chunk), bytes read straight from global memory.  With only a handful of
// (fold, chunk) pairs -- winequality has 9 -- the kernel above cannot fill the GPU without re-staging the rows for every
// few jobs, and its start-up (job list, staging, first tree) is then most of its run time.
__global__ void __launch_bounds__(256) k_predict_small(Wave w) {
  const int j = blockIdx.x;
  const int chunk = blockIdx.y;
  const JobState& js = w.js[j];
  if (!js.active) return;
  const JobParams jp = w.jp[j];
  const FoldDev fd = w.folds[jp.fold];
  __shared__ TreeNode s_tree[kNodeCap];
  const int n_nodes = (2 << jp.max_depth) - 1;
  const TreeNode* tree = w.tree + (size_t)j * kNodeCap;
  for (int i = threadIdx.x; i < n_nodes; i += blockDim.x) s_tree[i] = tree[i];
  __syncthreads();
  float* pred = w.pred + (size_t)j * w.n;
  const long long r0 = (long long)chunk * kPredRows;
  float gm = 0.f, hm = 0.f;
  double msum = 0.0;
  for (int i = threadIdx.x; i < kPredRows; i += blockDim.x) {
    const long long r = r0 + i;
    if (r >= w.n) break;
    const uint8_t* row = fd.bins + (size_t)r * w.fpad;
    int nid = 0;
    TreeNode t = s_tree[0];
    while (t.state == kSplit) {
      nid = 2 * nid + 1 + (row[t.feat] > t.bin ? 1 : 0);
      t = s_tree[nid];
    }
    const float p = pred[r] + t.leaf;
    pred[r] = p;
    const float y = fd.y[r];
    if (r < fd.ntrain) {
      float g, h;
      grad_pair(w.objective, p, y, jp.spw, g, h);
      gm = fmaxf(gm, fabsf(g)); hm = fmaxf(hm, fabsf(h));
    } else {
      const float pv = w.objective == 1 ? sigmoid_x(p) : p;
      float e;
      if (w.metric == 1) {     // EvalRowLogLoss
        const float eps = 1e-16f, pneg = 1.0f - pv;
        if (pv < eps) e = -y * logf(eps) - (1.0f - y) * logf(1.0f - eps);
        else if (pneg < eps) e = -y * logf(1.0f - eps) - (1.0f - y) * logf(eps);
        else e = -y * logf(pv) - (1.0f - y) * logf(pneg);
      } else {                 // EvalRowRMSE
        const float d = y - pv;
        e = d * d;
      }
      msum += (double)e;
    }
  }
  for (int d = 16; d; d >>= 1) {
    gm = fmaxf(gm, __shfl_xor_sync(0xffffffffu, gm, d));
    hm = fmaxf(hm, __shfl_xor_sync(0xffffffffu, hm, d));
    msum += __shfl_xor_sync(0xffffffffu, msum, d);
  }
  __shared__ double s_m[8];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  if (lane == 0) {
    if (gm > 0.f) atomicMax(&w.js[j].gmax_bits, __float_as_uint(gm));
    if (hm > 0.f) atomicMax(&w.js[j].hmax_bits, __float_as_uint(hm));
    s_m[wid] = msum;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0;
    for (int q = 0; q < 8; ++q) t += s_m[q];
    w.mpart[(size_t)j * w.pred_chunks + chunk] = t;
  }
}

// aggcv mean over folds, EarlyStopping on the test mean (minimise, strict improvement), next-round set-up
__global__ void __launch_bounds__(32) k_round_end(Wave w, int first, double* fitness_out, int* rounds_out) {
  const int i = blockIdx.x;
  IndState& is = w.is[i];
  if (!is.active) return;
  const int lane = threadIdx.x;
  if (!first) {
    double fold_sum = 0.0;
    for (int k = 0; k < w.nfold; ++k) {
      const int j = i * w.nfold + k;
      const double* mp = w.mpart + (size_t)j * w.pred_chunks;
      double s = 0.0;
      for (int c = lane; c < w.pred_chunks; c += 32) s += mp[c];
      for (int d = 16; d; d >>= 1) s += __shfl_xor_sync(0xffffffffu, s, d);
      const double nt = (double)w.folds[w.jp[j].fold].ntest;
      fold_sum += (w.metric == 0) ? sqrt(s / nt) : s / nt;
    }
    const double mean = fold_sum / (double)w.nfold;
    bool done = false;
    if (lane == 0) {
      const int it = is.round;
      if (w.trace) w.trace[(size_t)i * w.num_boost_round + it] = mean;
      is.last_mean = mean;
      is.rounds_run = it + 1;
      if (w.early_stopping_rounds > 0) {
        if (it == 0) { is.best_score = mean; is.best_iter = 0; }
        else if (mean < is.best_score) { is.best_score = mean; is.best_iter = it; }
        else if (it - is.best_iter >= w.early_stopping_rounds) { is.stopped = 1; }
      }
      is.round = it + 1;
      if (is.stopped || is.round >= w.num_boost_round) {
        is.active = 0;
        is.fitness = is.stopped ? is.best_score : is.last_mean;   // cv() truncates to best_iteration + 1 rows
        fitness_out[is.out_index] = is.fitness;
        if (rounds_out) rounds_out[is.out_index] = is.rounds_run;
        for (int k = 0; k < w.nfold; ++k) w.js[i * w.nfold + k].active = 0;
        atomicAdd(&w.counters[2], (unsigned long long)(-1ll));
        done = true;
      }
    }
    done = __shfl_sync(0xffffffffu, (int)done, 0);
    if (done) return;
  }
  if (lane < w.nfold) prepare_next_round(w, i * w.nfold + lane);
}

// parity dumps
__global__ void k_dump_trees(Wave w, int round, int tree_rounds, gt_tree_node* out) {
  const int j = blockIdx.y;
  const Node* nodes = w.nodes + (size_t)j * kNodeCap;
  gt_tree_node* o = out + ((size_t)j * tree_rounds + round) * (kNodeCap - 1);
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < kNodeCap - 1; i += gridDim.x * blockDim.x) {
    const Node& nd = nodes[i];
    gt_tree_node r;
    if (nd.state == kSplit) { r.feat = nd.feat; r.bin = nd.bin; r.value = 0.f; r.gain = nd.loss_chg; }
    else if (nd.state == kLeaf) { r.feat = -1; r.bin = 0; r.value = nd.leaf; r.gain = 0.f; }
    else { r.feat = -2; r.bin = 0; r.value = 0.f; r.gain = 0.f; }
    o[i] = r;
  }
}

__global__ void k_dump_node_rows(Wave w, gt_node_rows* out) {
  const int j = blockIdx.y;
  const Node* nodes = w.nodes + (size_t)j * kNodeCap;
  gt_node_rows* o = out + (size_t)j * (kNodeCap - 1);
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < kNodeCap - 1; i += gridDim.x * blockDim.x) {
    const Node& nd = nodes[i];
    o[i] = gt_node_rows{nd.begin, nd.cnt, (int)nd.buf, (int)nd.state};
  }
}

}  // namespace

int hist_smem_bytes() { return kHistSmem; }

#ifdef GT_SCAN_STATS
extern "C" void dump_scan_stats() {
  unsigned long long h[4];
  cudaMemcpyFromSymbol(h, g_scan_stats, sizeof(h));
  fprintf(stderr, "scan stats: stage-2 warp calls %llu, alive lanes %llu, column walks %llu\n", h[0], h[1], h[2]);
}
#endif

int configure_kernels() {
  cudaError_t e = cudaFuncSetAttribute(k_hist<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, kHistSmem);
  if (e != cudaSuccess) return (int)e;
  e = cudaFuncSetAttribute(k_hist<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 2 * kHistSmem);
  if (e != cudaSuccess) return (int)e;
  e = cudaFuncSetAttribute(k_predict, cudaFuncAttributeMaxDynamicSharedMemorySize, kPredRowBytes + kPredTreeBytes + kMaxWaveInds * 12);
  if (e != cudaSuccess) return (int)e;
  return (int)cudaFuncSetAttribute(k_scan_deep, cudaFuncAttributeMaxDynamicSharedMemorySize, kDeepWarps * kRingWarpBytes);
}

void launch_init_wave(const Wave& w, cudaStream_t s) {
  k_init_wave<<<dim3(w.pred_chunks, w.n_jobs), 256, 0, s>>>(w);
}
void launch_grad(const Wave& w, cudaStream_t s) {
  int chunks = (int)((w.nt_stride + kPredRows - 1) / kPredRows);
  k_grad<<<dim3(chunks, w.n_jobs), 256, 0, s>>>(w);
}
// Per-level launches cover only the jobs whose trees reach that level (a prefix of the depth-ordered job array).
void launch_plan_hist(const Wave& w, int level, cudaStream_t s) { k_plan_hist<<<w.lvl_jobs[level], 256, 0, s>>>(w, level); }
void launch_hist_zero(const Wave& w, int level, cudaStream_t s) {
#ifdef GT_DBG_CLEAR_HIST
  k_dbg_clear_level<<<dim3(1 << level, w.lvl_jobs[level]), 256, 0, s>>>(w, level);
#endif
  k_hist_zero<<<dim3(32, w.lvl_jobs[level]), 256, 0, s>>>(w, level);
}
int launch_hist(const Wave& w, int level, cudaStream_t s) {   // returns the number of kernels launched
  int cap = hist_items_bound(w, level);
  const int jobs = w.lvl_jobs[level];
  const bool job_fastest = level <= kJobFastestLevels;
  // Two kernels: a 128-KiB CTA that takes a pair of feature groups (the whole 64-byte line of a 50-feature row; one CTA
  // per SM) and a 64-KiB CTA that takes one group (two CTAs per SM, whose fixed phases overlap).  Measured on B200
  // (profiles/r02_hist_experiments.md): pair CTAs win for 2 and 4 groups; for 7 groups seven single-group CTAs beat
  // three pairs plus a single, so an odd count goes to the single-group kernel throughout.
  int launched = 0;
  const int pairs = w.groups % 2 == 0 ? w.groups / 2 : 0;
  const int singles = w.groups - 2 * pairs;
  if (pairs) {
    k_hist<4><<<job_fastest ? dim3(jobs * pairs, cap) : dim3(cap * pairs, jobs), 512, 2 * kHistSmem, s>>>(w, level, 0, pairs);
    ++launched;
  }
  if (singles) {
    k_hist<2><<<job_fastest ? dim3(jobs * singles, cap) : dim3(cap * singles, jobs), 256, kHistSmem, s>>>(w, level, 2 * pairs, singles);
    ++launched;
  }
  return launched;
  return 1;
}
int launch_scan(const Wave& w, int level, cudaStream_t s) {   // returns the number of kernels launched
  const long long bound = (long long)w.lvl_jobs[level] * (level == 0 ? 1 : (1 << (level - 1)));
  const int grid = (int)(bound < 148 * 4 ? bound : 148 * 4);   // CTAs loop over the wave-wide pair list
  k_scan<<<grid, w.groups * scan_chunks(w.groups) * 32, 0, s>>>(w, level);
  if (bound < kDeepPairs) return 1;
  k_scan_deep<<<(unsigned)bound, w.groups * 32, w.groups * kRingWarpBytes, s>>>(w, level);
  return 2;
}
void launch_plan_part(const Wave& w, int level, cudaStream_t s) { k_plan_part<<<w.lvl_jobs[level + 1], 256, 0, s>>>(w, level); }
void launch_part_count(const Wave& w, int level, cudaStream_t s) {
  k_part_count<<<dim3(part_items_bound(w, level), w.lvl_jobs[level + 1]), 256, 0, s>>>(w, level);
}
void launch_part_scatter(const Wave& w, int level, cudaStream_t s) {
  k_part_scatter<<<dim3(part_items_bound(w, level), w.lvl_jobs[level + 1]), 256, 0, s>>>(w, level);
}
void launch_predict(const Wave& w, cudaStream_t s) {
  const int ctas = w.nfold * w.pred_chunks;
  if (ctas < 148) {   // fewer (fold, chunk) pairs than SMs
    k_predict_small<<<dim3(w.n_jobs, w.pred_chunks), 256, 0, s>>>(w);
    return;
  }
  int z = (2 * 148 + ctas - 1) / ctas;   // two CTAs' worth of work items per SM, if the wave has that many jobs
  z = z < 1 ? 1 : z > w.n_inds ? w.n_inds : z;
  k_predict<<<dim3(w.nfold, w.pred_chunks, z), kPredThreads, pred_smem_bytes(w.fpad, w.n_inds), s>>>(w);
}
void launch_round_end(const Wave& w, int first, double* fitness_out, int* rounds_out, cudaStream_t s) {
  k_round_end<<<w.n_inds, 32, 0, s>>>(w, first, fitness_out, rounds_out);
}
void launch_dump_trees(const Wave& w, int round, int tree_rounds, gt_tree_node* out, cudaStream_t s) {
  k_dump_trees<<<dim3(2, w.n_jobs), 256, 0, s>>>(w, round, tree_rounds, out);
}
void launch_dump_node_rows(const Wave& w, gt_node_rows* out, cudaStream_t s) {
  k_dump_node_rows<<<dim3(2, w.n_jobs), 256, 0, s>>>(w, out);
}

}  // namespace gt
