// The shared-memory-resident regime: data sets whose train slice has at most kSmallMaxTrain rows
// (winequality: 3 265 rows x 11 features per fold, configs C1 / C2 of BASELINE.json, reference scripts
// /root/reference/tests/test_wine-quality.py:17-25 and tests/test_grid_wine-quality.py:20-30).  There the streaming
// kernels of boost.cu spend their time on launches and empty CTAs -- up to 5000 rounds x 10 levels x ~10 kernels -- so
// this file runs the WHOLE of xgb.cv for one (individual, fold) job inside one persistent CTA: gradients, tree growth,
// prediction update, metric and early stopping loop on the device; one launch per gt_engine_eval.
//
// Same arithmetic as boost.cu, bit for bit: gradients quantised to the same fixed point, integer sums, the split
// arithmetic of split.cuh, the same two-stage (fp32 estimate, exact re-evaluation of the near-best) candidate search,
// settle_node(), the metric summed in the order of k_predict_small / k_round_end.  What differs is how the sums are
// formed.  A histogram per node costs 256 bins x F whatever the node holds, and deep levels have hundreds of nodes of
// a few rows each, so instead every feature keeps its rows SORTED BY BIN inside each node's segment (one stable
// partition per feature and level keeps them so): a prefix sum over that list gives, at the last row of every run of
// equal bins, exactly the (GL, HL) the histogram scan would have at that bin.  Cost per level: O(rows x features),
// independent of the number of nodes; no atomics on the sums, hence the same integers in any order.
//
// The folds of an individual must finish a round together (early stopping watches their mean): CTAs form teams of
// nfold, launched cooperatively so that a team's CTAs are co-resident, and meet at a per-round barrier in global memory.
#include <cooperative_groups.h>

#include "common.cuh"
#include "kernels.h"
#include "split.cuh"
#include <cstdio>

namespace gt {
namespace {

#ifdef GT_SMALL_PROF   // experiment builds: cycles of CTA 0 by phase
__device__ unsigned long long g_small_prof[16];
#define PROF_COUNT(i) atomicAdd(&g_small_prof[i], 1ull)
#define PROF_MARK(i) do { if (blockIdx.x == 0 && threadIdx.x == 0) { const long long now__ = clock64(); g_small_prof[i] += (unsigned long long)(now__ - prof_t); prof_t = now__; } } while (0)
#else
#define PROF_MARK(i)
#define PROF_COUNT(i)
#endif

constexpr int kST = kSmallThreads;
constexpr int kSlots = 1 << (kMaxDepth - 1);       // nodes of the deepest evaluated level
constexpr uint16_t kNoSlot = 0xffffu;
constexpr int kBatch = 32;                          // features whose sweeps run side by side
constexpr int kMaxChunks = (kSmallMaxTrain + 127) / 128;   // 128-element chunks of a list
static_assert(kSlots <= kST, "one thread per node slot");
static_assert(kSmallMaxTrain <= 8 * kST, "eight list elements per thread");

struct SlotArrays {        // one entry per node of the level being grown
  int beg[kSlots], end[kSlots];          // segment of the node in every feature's list
  int Gt[kSlots], Ht[kSlots];            // quantised totals (|sum| < 2^31: at most 4095 rows of 2^19)
  int baseG[kSlots], baseH[kSlots];      // prefix sums in front of the segment
  int bG[kSlots], bH[kSlots];            // prefix at the best candidate
  int baseL[kSlots], totL[kSlots];       // partition: lefts in front of the segment, lefts of the node
  float root_gain[kSlots];
  unsigned amax[kSlots];                 // best estimate so far (float bits; estimates are >= 0)
  unsigned long long key[kSlots];        // best exact candidate: gain bits << 32 | ~(feature << 8 | bin)
  uint16_t node[kSlots];                 // heap index
  int16_t sfeat[kSlots];                 // split feature (-1: not partitioned)
  uint8_t sbin[kSlots];
};

__device__ __forceinline__ int warp_incl_scan(int v) {
  const int lane = threadIdx.x & 31;
  for (int d = 1; d < 32; d <<= 1) { const int o = __shfl_up_sync(0xffffffffu, v, d); if (lane >= d) v += o; }
  return v;
}

// exclusive prefix of (a, b) over the block's threads; s_w: 2 * 32 ints of scratch.  Ends with a barrier.
__device__ __forceinline__ void block_excl_scan2(int a, int b, int& pa, int& pb, int* s_w) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int ia = warp_incl_scan(a), ib = warp_incl_scan(b);
  if (lane == 31) { s_w[wid] = ia; s_w[32 + wid] = ib; }
  __syncthreads();
  if (wid == 0) {
    const int va = lane < kST / 32 ? s_w[lane] : 0, vb = lane < kST / 32 ? s_w[32 + lane] : 0;
    const int sa = warp_incl_scan(va), sb = warp_incl_scan(vb);
    s_w[lane] = sa - va; s_w[32 + lane] = sb - vb;
  }
  __syncthreads();
  pa = ia - a + s_w[wid]; pb = ib - b + s_w[32 + wid];
  __syncthreads();   // scratch may be reused at once
}

// exact gain of one candidate as a sortable key (0: not valid or not positive).  Larger gain first, then the earlier
// (feature, bin).  One out-of-line copy: the double-precision code is long and only a few candidates get here.
__device__ __noinline__ unsigned long long exact_key(const JobParams& jp, int Gt, int Ht, float rg, int GL, int HL, int f, int bin,
                                                     double inv_sg, double inv_sh) {
  const ChildTot t{(long long)Gt, (long long)Ht, rg};
  Cand one{0.0f, -1, -1, 0, 0};
  exact_candidate(one, jp, t, (long long)GL, (long long)HL, f, bin, inv_sg, inv_sh);
  if (one.f < 0) return 0ull;
  return ((unsigned long long)__float_as_uint(one.chg) << 32) | (0xffffffffu - (((unsigned)f << 8) | (unsigned)bin));
}

__device__ __forceinline__ void team_barrier(unsigned* bar, unsigned& target, int nfold) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    atomicAdd(bar, 1u);
    target += (unsigned)nfold;
    while (*reinterpret_cast<volatile unsigned*>(bar) < target) __nanosleep(32);
    __threadfence();
  }
  __syncthreads();
}

}  // namespace

namespace {
constexpr size_t kTrainCap = (kSmallMaxTrain + 32) / 32 * 32;
constexpr size_t kOffPred = (sizeof(SlotArrays) + 15) / 16 * 16;
constexpr size_t kOffGq = kOffPred + (size_t)kSmallMaxRows * 4;
constexpr size_t kOffPos = kOffGq + kTrainCap * 8;
constexpr size_t kOffLeft = kOffPos + kTrainCap * 2;
constexpr size_t kOffWalk = kOffLeft + kTrainCap;
constexpr size_t kOffLeaf = kOffWalk + (size_t)kNodeCap * 4;
constexpr size_t kOffScr = kOffLeaf + (size_t)kNodeCap * 4;
constexpr size_t kOffStage = kOffScr + 1024;
constexpr size_t kSmallSmem = kOffStage + (size_t)(kSmallThreads / 32) * 128 * 16;
static_assert(kSmallSmem + 16 * 1024 <= 227 * 1024, "dynamic + static shared memory of k_small_cv must fit one SM");
}  // namespace

int small_smem_bytes(int n, int nt_stride) {
  (void)n; (void)nt_stride;
  return (int)kSmallSmem;
}

namespace {

__global__ void __launch_bounds__(kST, 1) k_small_cv(SmallWave w) {
  extern __shared__ __align__(16) unsigned char s_raw[];
  // fixed offsets (the maxima of the regime, not this data set's sizes): every array's address is a compile-time constant
  // off the one base, which the 64-register budget would otherwise make the compiler recompute at every use
  SlotArrays& S = *reinterpret_cast<SlotArrays*>(s_raw);
  float* s_pred = reinterpret_cast<float*>(s_raw + kOffPred);
  int2* s_gq = reinterpret_cast<int2*>(s_raw + kOffGq);
  uint16_t* s_pos = reinterpret_cast<uint16_t*>(s_raw + kOffPos);
  uint8_t* s_left = s_raw + kOffLeft;
  uint32_t* s_walk = reinterpret_cast<uint32_t*>(s_raw + kOffWalk);
  float* s_leaf = reinterpret_cast<float*>(s_raw + kOffLeaf);
  int* s_scr = reinterpret_cast<int*>(s_raw + kOffScr);         // 256 ints of scratch
  int4* s_stage = reinterpret_cast<int4*>(s_raw + kOffStage);   // per warp: the candidates of one chunk, dense
  __shared__ uint32_t s_tmask[kMaskWords], s_lmask[kMaskWords], s_in[kMaskWords];
  __shared__ double s_dm[kST / 256][8], s_mpart[8];
  __shared__ int2 s_cv[kBatch][kMaxChunks];            // look-back: running sums at the end of (feature in batch, chunk)
  __shared__ unsigned s_cf[kBatch][kMaxChunks];        // ... valid when it holds the sweep's epoch
  __shared__ uint16_t s_feats[256];                    // the sweep's features
  unsigned epoch = 0u;
  __shared__ float s_fm[2][kST / 32];
  __shared__ long long s_ll[2][kST / 32];
  __shared__ int s_i[4];
  __shared__ float s_f[4];
  __shared__ double s_d[2];

  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int team = blockIdx.x / w.nfold, fold = blockIdx.x % w.nfold;
  unsigned bar_target = 0u;
  unsigned* bar = w.team_bar + team;
  const FoldDev fd = w.folds[fold];
  const int ntrain = (int)fd.ntrain, n = w.n, fpad = w.fpad, F = w.f;
  Node* nodes = w.nodes + (size_t)blockIdx.x * kNodeCap;
  TreeNode* tree = w.tree + (size_t)blockIdx.x * kNodeCap;
  uint32_t* lists = w.lists + (size_t)blockIdx.x * 2 * F * w.nt_stride;
  const uint32_t* base_lists = w.base_lists + (size_t)fold * F * w.nt_stride;
  const int pred_chunks = (n + kPredRows - 1) / kPredRows;
  int parity = 0;
  for (int i = tid; i < kBatch * kMaxChunks; i += kST) (&s_cf[0][0])[i] = 0u;     // no entry carries an epoch yet (epochs start at 1)
  __syncthreads();
#ifdef GT_SMALL_PROF
  long long prof_t = clock64();
#endif

  for (;;) {
    // ---- next individual of the team
    if (fold == 0 && tid == 0) {
      const int at = atomicAdd(w.queue, 1);
      w.team_ind[team * 2 + parity] = at < w.n_inds ? w.order[at] : -1;
    }
    team_barrier(bar, bar_target, w.nfold);
    const int q = *reinterpret_cast<volatile int*>(&w.team_ind[team * 2 + parity]);
    parity ^= 1;
    if (q < 0) return;
    const JobParams jp = w.jp[q * w.nfold + fold];
    const GainK gk{jp.alpha, jp.lambda, jp.mds};
    const bool mds_on = jp.mds != 0.0f;
    const int n_nodes_max = (2 << jp.max_depth) - 1;

    // ---- k_init_wave: margins, gradient range of round 0
    float gm = 0.f, hm = 0.f;
    for (int r = tid; r < n; r += kST) {
      s_pred[r] = jp.base_margin;
      if (r < ntrain) {
        float g, h;
        grad_pair(w.objective, jp.base_margin, fd.y[r], jp.spw, g, h);
        gm = fmaxf(gm, fabsf(g)); hm = fmaxf(hm, fabsf(h));
      }
    }
    int best_iter = 0, stopped = 0, rounds_run = 0;
    double best_score = 0.0, last_mean = 0.0;

    for (int round = 0; round < w.num_boost_round; ++round) {
      // ---- prepare_next_round: power-of-two quantisation scales from the gradient range
      gm = __uint_as_float(__reduce_max_sync(0xffffffffu, __float_as_uint(gm)));
      hm = __uint_as_float(__reduce_max_sync(0xffffffffu, __float_as_uint(hm)));
      if (lane == 0) { s_fm[0][wid] = gm; s_fm[1][wid] = hm; }
      __syncthreads();
      if (tid == 0) {
        float gmax = 0.f, hmax = 0.f;
        for (int k = 0; k < kST / 32; ++k) { gmax = fmaxf(gmax, s_fm[0][k]); hmax = fmaxf(hmax, s_fm[1][k]); }
        int e;
        float sg = 1.0f, sh = 1.0f;
        if (gmax > 0.f && isfinite(gmax)) { frexpf(gmax, &e); sg = ldexpf(1.0f, kQuantBits - e); }
        if (hmax > 0.f && isfinite(hmax)) { frexpf(hmax, &e); sh = ldexpf(1.0f, kQuantBitsH - e); }
        s_f[0] = sg; s_f[1] = sh;
      }
      __syncthreads();
      const float sg = s_f[0], sh = s_f[1];
      const double inv_sg = 1.0 / (double)sg, inv_sh = 1.0 / (double)sh;
      const float isg = (float)inv_sg, ish = (float)inv_sh;
      const long long hmin_q = (long long)ceil((double)jp.mcw * (double)sh);
      const int hmin32 = hmin_q > 0x7fffffffll ? 0x7fffffff : (int)hmin_q;   // node sums are below 2^31: a larger bound just means "never"

      PROF_MARK(0);
      // ---- k_grad: objective gradient, row subsampling, quantisation, root sums
      {
        const uint64_t skey = stream_key(1u, (uint32_t)fold, (uint32_t)round, 0u);
        const bool sub = jp.subsample < 1.0f;
        long long sG = 0, sH = 0;
        for (int r = tid; r < ntrain; r += kST) {
          float g, h;
          grad_pair(w.objective, s_pred[r], fd.y[r], jp.spw, g, h);
          if (sub && !(u01(draw64(skey, (uint64_t)r)) < jp.subsample)) { g = 0.f; h = 0.f; }
          const int qg = __float2int_rn(g * sg), qh = __float2int_rn(h * sh);
          s_gq[r] = make_int2(qg, qh);
          s_pos[r] = 0;
          sG += qg; sH += qh;
        }
        for (int d = 16; d; d >>= 1) { sG += __shfl_xor_sync(0xffffffffu, sG, d); sH += __shfl_xor_sync(0xffffffffu, sH, d); }
        if (lane == 0) { s_ll[0][wid] = sG; s_ll[1][wid] = sH; }
        for (int i = tid; i < n_nodes_max; i += kST) { nodes[i].state = kAbsent; tree[i] = TreeNode{0.f, (int16_t)-1, (uint8_t)0, (uint8_t)kAbsent}; }
        __syncthreads();
        if (tid == 0) {
          long long a = 0, b = 0;
          for (int k = 0; k < kST / 32; ++k) { a += s_ll[0][k]; b += s_ll[1][k]; }
          Node& r = nodes[0];
          r.G = a; r.H = b; r.begin = 0; r.cnt = (uint32_t)ntrain; r.leaf = 0.f; r.root_gain = 0.f; r.loss_chg = 0.f;
          r.feat = -1; r.bin = 0; r.state = kOpen; r.buf = 0; r.narrow = 1; r.absG = 0ull;
          tree[0] = TreeNode{0.f, (int16_t)-1, (uint8_t)0, (uint8_t)kOpen};
          S.node[0] = 0; S.beg[0] = 0; S.end[0] = ntrain; S.Gt[0] = (int)a; S.Ht[0] = (int)b;
          S.root_gain[0] = calc_gain(jp, (double)a * inv_sg, (double)b * inv_sh);
        }
        __syncthreads();
      }

      PROF_MARK(1);
      // ---- tree growth, level by level.  Inside a level the warps work independently, one feature each: a warp sweeps
      // its feature's list in tiles of 32 consecutive elements (warp scans, no block barriers), so a level costs a handful
      // of barriers whatever the number of features.
      int m = 1;                                  // nodes of the level (slots)
      const uint32_t* cur = base_lists;           // level 0 reads the fold's base lists; later levels ping-pong
      int pp = 0;
      const int n_chunks = (ntrain + 127) / 128;
      for (int level = 0; level < jp.max_depth && m > 0; ++level) {
        // feature sets (k_plan_hist): colsample_bytree once per tree, colsample_bylevel at every level
        if (level == 0) {
          for (int k = tid; k < kMaskWords; k += kST) {
            uint32_t mk = 0u;
            for (int b = 0; b < 32; ++b) if (k * 32 + b < F) mk |= 1u << b;
            s_in[k] = mk;
          }
          __syncthreads();
          int n_tree = F;
          if (jp.coltree != 1.0f) {
            n_tree = (int)(jp.coltree * (float)F); if (n_tree < 1) n_tree = 1;
            sample_mask(stream_key(2u, (uint32_t)fold, (uint32_t)round, 0u), s_in, F, n_tree, s_tmask);
          } else {
            for (int k = tid; k < kMaskWords; k += kST) s_tmask[k] = s_in[k];
          }
          if (tid == 0) s_i[0] = n_tree;
          __syncthreads();
        }
        if (jp.collevel != 1.0f) {
          int n_lvl = (int)(jp.collevel * (float)s_i[0]); if (n_lvl < 1) n_lvl = 1;
          sample_mask(stream_key(3u, (uint32_t)fold, (uint32_t)round, (uint32_t)level), s_tmask, F, n_lvl, s_lmask);
        } else {
          for (int k = tid; k < kMaskWords; k += kST) s_lmask[k] = s_tmask[k];
        }
        for (int s = tid; s < m; s += kST) { S.amax[s] = 0u; S.key[s] = 0ull; S.bG[s] = 0; S.bH[s] = 0; S.totL[s] = 0; }
        __syncthreads();

        PROF_MARK(2);
        // ---- split search.  The prefix sums of a list run over ALL its rows; what lies in front of a node's segment
        // (the same rows for every feature, hence the same sums) is noted when the sweep enters the segment.
        // prefix sums in front of every node's segment: rows of finished nodes count as zero in the sweeps, and the nodes of a
        // level lie in the lists in slot order, so it is the sum of the earlier nodes' totals
        {
          int og, oh;
          const int gt = tid < m ? S.Gt[tid] : 0, ht = tid < m ? S.Ht[tid] : 0;
          block_excl_scan2(gt, ht, og, oh, s_scr);
          if (tid < m) { S.baseG[tid] = og; S.baseH[tid] = oh; }
        }
        // the level's features, compacted
        if (wid == 0) {
          int cnt = 0;
          for (int f0 = 0; f0 < F; f0 += 32) {
            const int f = f0 + lane;
            const bool on = f < F && ((s_lmask[f >> 5] >> (f & 31)) & 1u);
            const unsigned bal = __ballot_sync(0xffffffffu, on);
            if (on) s_feats[cnt + __popc(bal & ((1u << lane) - 1u))] = (uint16_t)f;
            cnt += __popc(bal);
          }
          if (lane == 0) s_i[2] = cnt;
        }
        __syncthreads();
        const int n_feats = s_i[2];
        // A list is cut into chunks of 128 elements; (chunk, feature) tasks are dealt to the warps chunk-major, so every warp
        // of the CTA works whatever the number of features, and a task's running sums come from its predecessor chunk through
        // shared memory (look-back: the predecessor always started earlier, possibly in this very warp).
        for (int fb = 0; fb < n_feats; fb += kBatch) {
          const int nfb = min(kBatch, n_feats - fb);
          const unsigned inv_nfb = 0xffffffffu / (unsigned)nfb + 1u;
          ++epoch;
          for (int t = wid; t < n_chunks * nfb; t += kST / 32) {
            const int chunk = nfb == 1 ? t : (int)__umulhi((unsigned)t, inv_nfb), fi = t - chunk * nfb;     // t < 2^11: exact (2^32 / 1 does not fit the reciprocal)
            const int f = s_feats[fb + fi];
            const uint32_t* lst = cur + (size_t)f * w.nt_stride;
            const int i0 = chunk * 128 + lane * 4;
            const uint4 q = i0 < ntrain ? reinterpret_cast<const uint4*>(lst)[chunk * 32 + lane] : make_uint4(0u, 0u, 0u, 0u);
            const uint32_t first_next = (lane == 31 && (chunk + 1) * 128 < ntrain) ? lst[(chunk + 1) * 128] : 0u;
            const uint32_t e[4] = {q.x, q.y, q.z, q.w};
            int pg[4], ph[4];
            uint16_t sl[4];
            int tg = 0, th = 0;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              const int row = (int)(e[k] & 0xffffu);
              sl[k] = (i0 + k < ntrain) ? s_pos[row] : kNoSlot;
              const int2 gh = sl[k] != kNoSlot ? s_gq[row] : make_int2(0, 0);
              tg += gh.x; th += gh.y;
              pg[k] = tg; ph[k] = th;
            }
            const int ig = warp_incl_scan(tg), ih = warp_incl_scan(th);
            int cg = 0, ch = 0;
            if (lane == 31) {
              if (chunk > 0) {
                while (*reinterpret_cast<volatile unsigned*>(&s_cf[fi][chunk - 1]) != epoch) {}
                cg = *reinterpret_cast<volatile int*>(&s_cv[fi][chunk - 1].x);
                ch = *reinterpret_cast<volatile int*>(&s_cv[fi][chunk - 1].y);
              }
              *reinterpret_cast<volatile int*>(&s_cv[fi][chunk].x) = cg + ig;
              *reinterpret_cast<volatile int*>(&s_cv[fi][chunk].y) = ch + ih;
              __threadfence_block();
              *reinterpret_cast<volatile unsigned*>(&s_cf[fi][chunk]) = epoch;
            }
            cg = __shfl_sync(0xffffffffu, cg, 31); ch = __shfl_sync(0xffffffffu, ch, 31);
            const int og = cg + ig - tg, oh = ch + ih - th;
            uint32_t nb_lane = __shfl_down_sync(0xffffffffu, e[0] >> 16, 1);
            if (lane == 31) nb_lane = first_next >> 16;
            // Candidates (the last row of a run of equal bins that is not its node's last row) are a minority of the
            // elements and scattered over the lanes, so they are first gathered into a dense per-warp list in shared
            // memory; the arithmetic below then runs with every lane busy.
            int4* stage = s_stage + wid * 128;
            int n_cand = 0;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              const uint32_t bin = e[k] >> 16;
              const uint32_t nb = k < 3 ? (e[(k + 1) & 3] >> 16) : nb_lane;
              const uint16_t slk = sl[k];
              const bool cand = slk != kNoSlot && i0 + k + 1 < S.end[slk] && nb != bin;
              const unsigned bal = __ballot_sync(0xffffffffu, cand);
              if (cand) stage[n_cand + __popc(bal & ((1u << lane) - 1u))] = make_int4(og + pg[k], oh + ph[k], (int)slk, (int)bin);
              n_cand += __popc(bal);
            }
            __syncwarp();
            for (int c = lane; c < n_cand; c += 32) {
              const int4 rec = stage[c];
              const int slk = rec.z;
              const int GL = rec.x - S.baseG[slk], HL = rec.y - S.baseH[slk];
              const int Ht = S.Ht[slk];
              if (HL >= hmin32 && Ht - HL >= hmin32) {
                // stage 1: fp32 estimate; stage 2: exact gain (reference expression order) if the estimate is within
                // its own error of the best estimate seen so far for the node -- a bound that only rises, so nothing
                // that could win is skipped
                const float Gt_f = (float)S.Gt[slk] * isg, Ht_f = (float)Ht * ish;
                const float glf = (float)GL * isg, hlf = (float)HL * ish;
                const float v = mds_on ? split_est<true>(gk, glf, hlf, Gt_f - glf, Ht_f - hlf)
                                       : split_est<false>(gk, glf, hlf, Gt_f - glf, Ht_f - hlf);
                PROF_COUNT(12);
                const float est = fmaxf(v, 0.0f);
                const float seen = __uint_as_float(S.amax[slk]);             // may be stale: any lower bound of the final maximum will do
                if (est > seen) atomicMax(&S.amax[slk], __float_as_uint(est));
                const float amax = fmaxf(seen, est), rg = S.root_gain[slk];
                const float margin = 3.2e-5f * (amax + fabsf(rg)) + 1e-12f;
                if (est >= fmaxf(amax, rg) - 2.0f * margin) {
                  PROF_COUNT(13);
                  const unsigned long long key = exact_key(jp, S.Gt[slk], Ht, rg, GL, HL, f, rec.w, inv_sg, inv_sh);
                  if (key) atomicMax(&S.key[slk], key);
                }
              }
            }
            __syncwarp();     // the list is rewritten by the warp's next task
          }
          __syncthreads();
        }

        PROF_MARK(3);
        // ---- the winning candidate of every node: which rows go left, and their sums (GL, HL, count)
        {
          uint8_t bv[8];
          uint16_t slv[8];
          unsigned lov[8];
#pragma unroll
          for (int k = 0; k < 8; ++k) {        // all the gathers of this thread's rows first, then their use
            const int r = k * kST + tid;
            slv[k] = kNoSlot; lov[k] = 0u; bv[k] = 255;
            if (r < ntrain) {
              slv[k] = s_pos[r];
              if (slv[k] != kNoSlot && S.key[slv[k]] != 0ull) {
                lov[k] = 0xffffffffu - (unsigned)(S.key[slv[k]] & 0xffffffffull);
                bv[k] = fd.bins[(size_t)r * fpad + (lov[k] >> 8)];
              } else {
                slv[k] = kNoSlot;
              }
            }
          }
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            const int r = k * kST + tid;
            if (k * kST >= ntrain) break;                                  // block-uniform
            const uint16_t sl = slv[k];
            const int l = (sl != kNoSlot && (unsigned)bv[k] <= (lov[k] & 255u)) ? 1 : 0;
            int2 gh = make_int2(0, 0);
            if (r < ntrain) { s_left[r] = (uint8_t)l; if (l) gh = s_gq[r]; }
            // rows of a warp mostly belong to one node at the top levels: one atomic per warp then, one per row otherwise
            const uint16_t sl0 = (uint16_t)__shfl_sync(0xffffffffu, (unsigned)sl, 0);
            if (__all_sync(0xffffffffu, sl == sl0)) {
              const int sg2 = __reduce_add_sync(0xffffffffu, gh.x), sh2 = __reduce_add_sync(0xffffffffu, gh.y);
              const int sc = __reduce_add_sync(0xffffffffu, l);
              if (lane == 0 && sl0 != kNoSlot && sc) { atomicAdd(&S.bG[sl0], sg2); atomicAdd(&S.bH[sl0], sh2); atomicAdd(&S.totL[sl0], sc); }
            } else if (l) {
              atomicAdd(&S.bG[sl], gh.x); atomicAdd(&S.bH[sl], gh.y); atomicAdd(&S.totL[sl], 1);
            }
          }
        }
        __syncthreads();

        PROF_MARK(4);
        // ---- settle the level's nodes (CPUExpandEntry::IsValid), create children
        for (int s = tid; s < m; s += kST) {
          const int node = S.node[s];
          const ChildTot t{(long long)S.Gt[s], (long long)S.Ht[s], S.root_gain[s]};
          Cand best{0.0f, -1, -1, 0, 0};
          if (S.key[s] != 0ull) {
            const unsigned lo = 0xffffffffu - (unsigned)(S.key[s] & 0xffffffffull);
            best.chg = __uint_as_float((unsigned)(S.key[s] >> 32)); best.f = (int)(lo >> 8); best.b = (int)(lo & 255u);
            best.GL = S.bG[s]; best.HL = S.bH[s];
          }
          settle_node(jp, level, nodes, tree, node, t, best, inv_sg, inv_sh, hmin_q, 0);
          settle_node(jp, level, nodes, tree, node, t, best, inv_sg, inv_sh, hmin_q, 1);
          const bool split = nodes[node].state == kSplit;
          const bool part = split && (nodes[2 * node + 1].state == kOpen || nodes[2 * node + 2].state == kOpen);
          S.sfeat[s] = part ? nodes[node].feat : (int16_t)-1;
          S.sbin[s] = nodes[node].bin;
        }
        __syncthreads();
        int any = 0;
        for (int s = tid; s < m; s += kST) any |= S.sfeat[s] >= 0;
        any = __syncthreads_or(any);
        if (!any || level + 1 >= jp.max_depth) { m = 0; break; }

        PROF_MARK(5);
        // ---- stable partition of every live feature's list (partition_builder.h), node by node in place; a warp per
        // feature again.  Lefts in front of a segment: noted on entering it, like the sums above.
        uint32_t* nxt = lists + (size_t)pp * F * w.nt_stride;
        // per row: 0 = stays where it is (finished node, or a node that is not partitioned), 1 = goes right, 2 = goes left;
        // the sweeps below then touch the node tables only for rows that move
        for (int r = tid; r < ntrain; r += kST) {
          const uint16_t sl = s_pos[r];
          s_left[r] = (sl != kNoSlot && S.sfeat[sl] >= 0) ? (uint8_t)(1 + s_left[r]) : (uint8_t)0;
        }
        __syncthreads();
        {   // lefts in front of every partitioned node's segment: the earlier partitioned nodes' left counts
          int ol, dummy;
          const int tl = (tid < m && S.sfeat[tid] >= 0) ? S.totL[tid] : 0;
          block_excl_scan2(tl, 0, ol, dummy, s_scr);
          if (tid < m) S.baseL[tid] = ol;
        }
        if (wid == 0) {     // the tree's features (features outside its set are never read again)
          int cnt = 0;
          for (int f0 = 0; f0 < F; f0 += 32) {
            const int f = f0 + lane;
            const bool on = f < F && ((s_tmask[f >> 5] >> (f & 31)) & 1u);
            const unsigned bal = __ballot_sync(0xffffffffu, on);
            if (on) s_feats[cnt + __popc(bal & ((1u << lane) - 1u))] = (uint16_t)f;
            cnt += __popc(bal);
          }
          if (lane == 0) s_i[2] = cnt;
        }
        __syncthreads();
        const int n_pfeats = s_i[2];
        for (int fb = 0; fb < n_pfeats; fb += kBatch) {
          const int nfb = min(kBatch, n_pfeats - fb);
          const unsigned inv_nfb = 0xffffffffu / (unsigned)nfb + 1u;
          ++epoch;
          for (int t = wid; t < n_chunks * nfb; t += kST / 32) {
            const int chunk = nfb == 1 ? t : (int)__umulhi((unsigned)t, inv_nfb), fi = t - chunk * nfb;     // t < 2^11: exact (2^32 / 1 does not fit the reciprocal)
            const int f = s_feats[fb + fi];
            const uint32_t* lst = cur + (size_t)f * w.nt_stride;
            uint32_t* out = nxt + (size_t)f * w.nt_stride;
            const int i0 = chunk * 128 + lane * 4;
            const uint4 q = i0 < ntrain ? reinterpret_cast<const uint4*>(lst)[chunk * 32 + lane] : make_uint4(0u, 0u, 0u, 0u);
            const uint32_t e[4] = {q.x, q.y, q.z, q.w};
            int pl[4], lk[4];
            uint16_t sl[4];
            int tl = 0;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              const int row = (int)(e[k] & 0xffffu);
              const int code = (i0 + k < ntrain) ? (int)s_left[row] : 0;
              sl[k] = code ? s_pos[row] : kNoSlot;
              lk[k] = code == 2;
              pl[k] = tl;                                                      // exclusive, lane-local
              tl += lk[k];
            }
            const int il = warp_incl_scan(tl);
            int cl = 0;
            if (lane == 31) {
              if (chunk > 0) {
                while (*reinterpret_cast<volatile unsigned*>(&s_cf[fi][chunk - 1]) != epoch) {}
                cl = *reinterpret_cast<volatile int*>(&s_cv[fi][chunk - 1].x);
              }
              *reinterpret_cast<volatile int*>(&s_cv[fi][chunk].x) = cl + il;
              __threadfence_block();
              *reinterpret_cast<volatile unsigned*>(&s_cf[fi][chunk]) = epoch;
            }
            cl = __shfl_sync(0xffffffffu, cl, 31);
            const int ol = cl + il - tl;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              if (i0 + k < ntrain) {
                int dest = i0 + k;
                if (sl[k] != kNoSlot) {
                  const int beg = S.beg[sl[k]], rankL = ol + pl[k] - S.baseL[sl[k]];
                  dest = lk[k] ? beg + rankL : beg + S.totL[sl[k]] + (i0 + k - beg - rankL);
                }
                out[dest] = e[k];
              }
            }
          }
          __syncthreads();
        }
        PROF_MARK(6);
        // ---- next level's nodes: the open children, in node order
        {
          int mine = 0;       // open children of the slot this thread owns (slot = tid: kSlots <= kST)
          int lo = 0, ro = 0;
          if (tid < m && S.sfeat[tid] >= 0) {
            const int node = S.node[tid];
            lo = nodes[2 * node + 1].state == kOpen; ro = nodes[2 * node + 2].state == kOpen;
            mine = lo + ro;
          }
          int off, dummy;
          block_excl_scan2(mine, 0, off, dummy, s_scr);
          // the children's records need this slot's values before the slot arrays are overwritten: read, barrier, write
          int node = 0, beg = 0, end = 0, totL = 0;
          if (tid < m) { node = S.node[tid]; beg = S.beg[tid]; end = S.end[tid]; totL = S.totL[tid]; }
          const bool part = tid < m && S.sfeat[tid] >= 0;
          if (tid == kST - 1) s_i[1] = off + mine;
          // row -> next slot, through the old slot: the two children's new slots are stashed in baseG / baseH
          if (tid < m) { S.baseG[tid] = part && lo ? off : -1; S.baseH[tid] = part && ro ? off + lo : -1; }
          __syncthreads();
          for (int r = tid; r < ntrain; r += kST) {
            const uint16_t sl = s_pos[r];
            if (sl == kNoSlot) continue;
            const int code = s_left[r];
            const int ns = code == 2 ? S.baseG[sl] : S.baseH[sl];
            s_pos[r] = (code && ns >= 0) ? (uint16_t)ns : kNoSlot;
          }
          __syncthreads();
          if (part) {
            Node& L = nodes[2 * node + 1];
            Node& R = nodes[2 * node + 2];
            L.begin = (uint32_t)beg; L.cnt = (uint32_t)totL; R.begin = (uint32_t)(beg + totL); R.cnt = (uint32_t)(end - beg - totL);
            L.narrow = 1; R.narrow = 1;
            int at = off;
            if (lo) {
              S.node[at] = (uint16_t)(2 * node + 1); S.beg[at] = beg; S.end[at] = beg + totL;
              S.Gt[at] = (int)L.G; S.Ht[at] = (int)L.H; S.root_gain[at] = L.root_gain; ++at;
            }
            if (ro) {
              S.node[at] = (uint16_t)(2 * node + 2); S.beg[at] = beg + totL; S.end[at] = end;
              S.Gt[at] = (int)R.G; S.Ht[at] = (int)R.H; S.root_gain[at] = R.root_gain;
            }
          }
          __syncthreads();
          m = s_i[1];
        }
        cur = nxt;
        pp ^= 1;
        PROF_MARK(7);
      }
      __syncthreads();
      PROF_MARK(8);

      // ---- k_predict_small: every row walks the new tree on its bin bytes; metric terms in that kernel's order
      for (int i = tid; i < kNodeCap; i += kST) {
        uint32_t wd = 0u; float lf = 0.f;
        if (i < n_nodes_max) {
          const TreeNode t = tree[i];
          lf = t.leaf;
          if (t.state == kSplit) wd = 0x80000000u | ((uint32_t)(uint16_t)t.feat << 8) | t.bin;
        }
        s_walk[i] = wd; s_leaf[i] = lf;
      }
      __syncthreads();
      gm = 0.f; hm = 0.f;
      for (int c0 = 0; c0 < pred_chunks; c0 += kST / 256) {
        const int chunk = c0 + (tid >> 8), vt = tid & 255;      // a virtual 256-thread CTA per chunk, like k_predict_small
        double msum = 0.0;
        if (chunk < pred_chunks) {
          for (int i = vt; i < kPredRows; i += 256) {
            const int r = chunk * kPredRows + i;
            if (r >= n) break;
            const uint8_t* row = fd.bins + (size_t)r * fpad;
            int nid = 0;
            uint32_t wd = s_walk[0];
            while (wd & 0x80000000u) {
              nid = 2 * nid + 1 + (row[(wd >> 8) & 0xffffu] > (wd & 0xffu) ? 1 : 0);
              wd = s_walk[nid];
            }
            const float p = s_pred[r] + s_leaf[nid];
            s_pred[r] = p;
            const float y = fd.y[r];
            if (r < ntrain) {
              float g, h;
              grad_pair(w.objective, p, y, jp.spw, g, h);
              gm = fmaxf(gm, fabsf(g)); hm = fmaxf(hm, fabsf(h));
            } else {
              const float pv = w.objective == 1 ? sigmoid_x(p) : p;
              float ev;
              if (w.metric == 1) {     // EvalRowLogLoss
                const float eps = 1e-16f, pneg = 1.0f - pv;
                if (pv < eps) ev = -y * logf(eps) - (1.0f - y) * logf(1.0f - eps);
                else if (pneg < eps) ev = -y * logf(1.0f - eps) - (1.0f - y) * logf(eps);
                else ev = -y * logf(pv) - (1.0f - y) * logf(pneg);
              } else {                 // EvalRowRMSE
                const float d = y - pv;
                ev = d * d;
              }
              msum += (double)ev;
            }
          }
        }
        for (int d = 16; d; d >>= 1) msum += __shfl_xor_sync(0xffffffffu, msum, d);
        if (lane == 0) s_dm[tid >> 8][(tid & 255) >> 5] = msum;
        __syncthreads();
        if (vt == 0 && chunk < pred_chunks) {
          double t = 0;
          for (int k = 0; k < 8; ++k) t += s_dm[tid >> 8][k];
          s_mpart[chunk] = t;
        }
        __syncthreads();
      }
      PROF_MARK(9);
      // ---- k_round_end: this fold's sum, then the mean over the team's folds
      if (wid == 0) {
        double s = 0.0;
        for (int c = lane; c < pred_chunks; c += 32) s += s_mpart[c];
        for (int d = 16; d; d >>= 1) s += __shfl_xor_sync(0xffffffffu, s, d);
        if (lane == 0) w.fold_s[((size_t)team * 2 + (round & 1)) * w.nfold + fold] = s;
      }
      team_barrier(bar, bar_target, w.nfold);
      if (tid == 0) {
        double fold_sum = 0.0;
        const volatile double* fs = w.fold_s + ((size_t)team * 2 + (round & 1)) * w.nfold;
        for (int k = 0; k < w.nfold; ++k) {
          const double nt = (double)w.folds[k].ntest;
          const double s = fs[k];
          fold_sum += (w.metric == 0) ? sqrt(s / nt) : s / nt;
        }
        s_d[0] = fold_sum / (double)w.nfold;
      }
      __syncthreads();
      const double mean = s_d[0];
      PROF_MARK(10);
      last_mean = mean; rounds_run = round + 1;
      if (w.early_stopping_rounds > 0) {
        if (round == 0) { best_score = mean; best_iter = 0; }
        else if (mean < best_score) { best_score = mean; best_iter = round; }
        else if (round - best_iter >= w.early_stopping_rounds) stopped = 1;
      }
      if (fold == 0 && tid == 0 && w.trace) w.trace[(size_t)q * w.num_boost_round + round] = mean;
      if (stopped) break;
    }
    if (fold == 0 && tid == 0) {
      const double fitness = stopped ? best_score : last_mean;     // cv() truncates to best_iteration + 1 rows
      const IndState& is = w.is[q];
      w.fitness_out[is.out_index] = fitness;
      if (w.rounds_out) w.rounds_out[is.out_index] = rounds_run;
    }
    __syncthreads();
  }
}

}  // namespace

#ifdef GT_SMALL_PROF
extern "C" void dump_small_prof() {
  unsigned long long h[16];
  cudaMemcpyFromSymbol(h, g_small_prof, sizeof(h));
  const char* names[] = {"round top", "grad", "masks", "search", "left pass", "settle", "partition", "frontier", "tree end", "predict", "round end"};
  unsigned long long tot = 0; for (int i = 0; i < 11; ++i) tot += h[i];
  for (int i = 0; i < 11; ++i) fprintf(stderr, "small prof %-10s %6.2f%%  %llu cycles\n", names[i], 100.0 * h[i] / (double)tot, h[i]);
  fprintf(stderr, "small prof candidates estimated %llu, evaluated exactly %llu (all CTAs)\n", h[12], h[13]);
}
#endif

int small_max_teams(int nfold, int smem_bytes) {
  int dev = 0, sms = 0, per_sm = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 0;
  if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return 0;
  if (cudaFuncSetAttribute(k_small_cv, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes) != cudaSuccess) return 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_small_cv, kST, smem_bytes) != cudaSuccess) return 0;
  return sms * per_sm / nfold;
}

int launch_small_cv(const SmallWave& w, int n_teams, int smem_bytes, cudaStream_t s) {
  SmallWave arg = w;
  void* args[] = {&arg};
  return (int)cudaLaunchCooperativeKernel((const void*)k_small_cv, dim3(n_teams * w.nfold), dim3(kST), args, (size_t)smem_bytes, s);
}

}  // namespace gt
