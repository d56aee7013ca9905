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

namespace gt {
namespace {

constexpr int kST = kSmallThreads;
constexpr int kSlots = 1 << (kMaxDepth - 1);       // nodes of the deepest evaluated level
constexpr uint16_t kNoSlot = 0xffffu;
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

int small_smem_bytes(int n, int nt_stride) {
  size_t b = sizeof(SlotArrays);
  b += (size_t)((n + 3) & ~3) * 4;    // predictions
  b += (size_t)nt_stride * 8;         // quantised gradient pairs
  b += (size_t)nt_stride * 2;         // node slot of every train row
  b += (size_t)nt_stride;             // goes-left flags
  b += (size_t)kNodeCap * 8;          // tree: walk words + leaf values
  b += 1024;                          // scratch
  return (int)((b + 15) / 16 * 16);
}

namespace {

__global__ void __launch_bounds__(kST, 2) k_small_cv(SmallWave w) {
  extern __shared__ __align__(16) unsigned char s_raw[];
  SlotArrays& S = *reinterpret_cast<SlotArrays*>(s_raw);
  float* s_pred = reinterpret_cast<float*>(s_raw + sizeof(SlotArrays));
  int2* s_gq = reinterpret_cast<int2*>(s_pred + ((w.n + 3) & ~3));   // 16-byte aligned whatever n is
  uint16_t* s_pos = reinterpret_cast<uint16_t*>(s_gq + w.nt_stride);
  uint8_t* s_left = reinterpret_cast<uint8_t*>(s_pos + w.nt_stride);
  uint32_t* s_walk = reinterpret_cast<uint32_t*>(s_left + w.nt_stride);
  float* s_leaf = reinterpret_cast<float*>(s_walk + kNodeCap);
  int* s_scr = reinterpret_cast<int*>(s_leaf + kNodeCap);     // 256 ints of scratch
  __shared__ uint32_t s_tmask[kMaskWords], s_lmask[kMaskWords], s_in[kMaskWords];
  __shared__ uint16_t s_firstbin[kST + 1];
  __shared__ double s_dm[2][8], s_mpart[8];
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
  const int C = (ntrain + kST - 1) / kST;                  // list elements per thread (contiguous chunk)
  Node* nodes = w.nodes + (size_t)blockIdx.x * kNodeCap;
  TreeNode* tree = w.tree + (size_t)blockIdx.x * kNodeCap;
  uint32_t* lists = w.lists + (size_t)blockIdx.x * 2 * F * w.nt_stride;
  const uint32_t* base_lists = w.base_lists + (size_t)fold * F * w.nt_stride;
  const int pred_chunks = (n + kPredRows - 1) / kPredRows;
  int parity = 0;

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

      // ---- tree growth, level by level
      int m = 1;                                  // nodes of the level (slots)
      const uint32_t* cur = base_lists;           // level 0 reads the fold's base lists; later levels ping-pong
      int pp = 0;
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
          __syncthreads();
        }
        for (int s = tid; s < m; s += kST) { S.amax[s] = 0u; S.key[s] = 0ull; S.bG[s] = 0; S.bH[s] = 0; }
        __syncthreads();

        // ---- split search: one pass per feature of the level's set
        for (int f = 0; f < F; ++f) {
          if (!((s_lmask[f >> 5] >> (f & 31)) & 1u)) continue;            // block-uniform
          const uint32_t* lst = cur + (size_t)f * w.nt_stride;
          const int i0 = tid * C;
          uint32_t e[8];
          int pg[8], ph[8];
          int tg = 0, th = 0;
#pragma unroll
          for (int c = 0; c < 8; ++c) {
            e[c] = 0u; pg[c] = 0; ph[c] = 0;
            if (c < C && i0 + c < ntrain) {
              e[c] = lst[i0 + c];
              const int2 gh = s_gq[e[c] & 0xffffu];
              tg += gh.x; th += gh.y;
              pg[c] = tg; ph[c] = th;                                      // inclusive, chunk-local
            }
          }
          s_firstbin[tid] = (uint16_t)(e[0] >> 16);
          if (tid == 0) s_firstbin[kST] = 0;
          int og, oh;
          block_excl_scan2(tg, th, og, oh, s_scr);
          // prefix in front of every node's segment (the same for every feature: the same rows precede it)
#pragma unroll
          for (int c = 0; c < 8; ++c) {
            if (c < C && i0 + c < ntrain) {
              const uint16_t sl = s_pos[e[c] & 0xffffu];
              if (sl != kNoSlot && i0 + c == S.beg[sl]) {
                const int2 gh = s_gq[e[c] & 0xffffu];
                S.baseG[sl] = og + pg[c] - gh.x; S.baseH[sl] = oh + ph[c] - gh.y;
              }
            }
          }
          __syncthreads();
          // stage 1: fp32 estimate of every candidate (last row of a run of equal bins, not the node's last row)
          float est[8];
          int cGL[8], cHL[8];
#pragma unroll
          for (int c = 0; c < 8; ++c) {
            est[c] = -1.0f;
            if (c < C && i0 + c < ntrain) {
              const uint16_t sl = s_pos[e[c] & 0xffffu];
              if (sl != kNoSlot && i0 + c + 1 < S.end[sl]) {
                const uint32_t nb = (c + 1 < C) ? (e[(c + 1) & 7] >> 16) : (uint32_t)s_firstbin[tid + 1];
                if (nb != (e[c] >> 16)) {
                  const int GL = og + pg[c] - S.baseG[sl], HL = oh + ph[c] - S.baseH[sl];
                  const int Ht = S.Ht[sl];
                  cGL[c] = GL; cHL[c] = HL;
                  if ((long long)HL >= hmin_q && (long long)(Ht - HL) >= hmin_q) {
                    const float Gt_f = (float)S.Gt[sl] * isg, Ht_f = (float)Ht * ish;
                    const float glf = (float)GL * isg, hlf = (float)HL * ish;
                    const float v = mds_on ? split_est<true>(gk, glf, hlf, Gt_f - glf, Ht_f - hlf)
                                           : split_est<false>(gk, glf, hlf, Gt_f - glf, Ht_f - hlf);
                    est[c] = fmaxf(v, 0.0f);
                    atomicMax(&S.amax[sl], __float_as_uint(est[c]));
                  }
                }
              }
            }
          }
          __syncthreads();
          // stage 2: exact gain (reference expression order) of the candidates within the estimate's error of the best
          unsigned long long mykey[8];
#pragma unroll
          for (int c = 0; c < 8; ++c) {
            mykey[c] = 0ull;
            if (est[c] >= 0.0f) {
              const uint16_t sl = s_pos[e[c] & 0xffffu];
              const float amax = __uint_as_float(S.amax[sl]), rg = S.root_gain[sl];
              const float margin = 3.2e-5f * (fmaxf(amax, 0.f) + fabsf(rg)) + 1e-12f;
              const float cut = fmaxf(amax, rg) - 2.0f * margin;
              if (est[c] >= cut) {
                const ChildTot t{(long long)S.Gt[sl], (long long)S.Ht[sl], rg};
                Cand one{0.0f, -1, -1, 0, 0};
                exact_candidate(one, jp, t, (long long)cGL[c], (long long)cHL[c], f, (int)(e[c] >> 16), inv_sg, inv_sh);
                if (one.f >= 0) {     // finite and > 0
                  mykey[c] = ((unsigned long long)__float_as_uint(one.chg) << 32) | (0xffffffffu - (((unsigned)f << 8) | (e[c] >> 16)));
                  atomicMax(&S.key[sl], mykey[c]);
                }
              }
            }
          }
          __syncthreads();
#pragma unroll
          for (int c = 0; c < 8; ++c) {
            if (mykey[c] != 0ull) {
              const uint16_t sl = s_pos[e[c] & 0xffffu];
              if (S.key[sl] == mykey[c]) { S.bG[sl] = cGL[c]; S.bH[sl] = cHL[c]; }
            }
          }
          // (the next feature's first barrier orders these writes before anything reads them)
        }
        __syncthreads();

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

        // ---- stable partition of every live feature's list (partition_builder.h), node by node in place
        for (int r = tid; r < ntrain; r += kST) {
          const uint16_t sl = s_pos[r];
          uint8_t l = 0;
          if (sl != kNoSlot && S.sfeat[sl] >= 0) l = fd.bins[(size_t)r * fpad + S.sfeat[sl]] <= S.sbin[sl] ? 1 : 0;
          s_left[r] = l;
        }
        __syncthreads();
        uint32_t* nxt = lists + (size_t)pp * F * w.nt_stride;
        bool first = true;
        for (int f = 0; f < F; ++f) {
          if (!((s_tmask[f >> 5] >> (f & 31)) & 1u)) continue;            // features outside the tree's set are never read again
          const uint32_t* lst = cur + (size_t)f * w.nt_stride;
          uint32_t* out = nxt + (size_t)f * w.nt_stride;
          const int i0 = tid * C;
          uint32_t e[8];
          int pl[8];
          int tl = 0;
#pragma unroll
          for (int c = 0; c < 8; ++c) {
            e[c] = 0u; pl[c] = 0;
            if (c < C && i0 + c < ntrain) {
              e[c] = lst[i0 + c];
              pl[c] = tl;                                                  // exclusive, chunk-local
              tl += s_left[e[c] & 0xffffu];
            }
          }
          int ol, dummy;
          block_excl_scan2(tl, 0, ol, dummy, s_scr);
          if (first) {     // lefts in front of each segment and inside it: the same for every feature
#pragma unroll
            for (int c = 0; c < 8; ++c) {
              if (c < C && i0 + c < ntrain) {
                const uint16_t sl = s_pos[e[c] & 0xffffu];
                if (sl != kNoSlot && S.sfeat[sl] >= 0) {
                  if (i0 + c == S.beg[sl]) S.baseL[sl] = ol + pl[c];
                  if (i0 + c == S.end[sl] - 1) S.totL[sl] = ol + pl[c] + s_left[e[c] & 0xffffu];     // inclusive count; base subtracted below
                }
              }
            }
            __syncthreads();
            for (int s = tid; s < m; s += kST) if (S.sfeat[s] >= 0) S.totL[s] -= S.baseL[s];
            __syncthreads();
            first = false;
          }
#pragma unroll
          for (int c = 0; c < 8; ++c) {
            if (c < C && i0 + c < ntrain) {
              const int row = e[c] & 0xffffu;
              const uint16_t sl = s_pos[row];
              int dest = i0 + c;
              if (sl != kNoSlot && S.sfeat[sl] >= 0) {
                const int beg = S.beg[sl], rankL = ol + pl[c] - S.baseL[sl];
                dest = s_left[row] ? beg + rankL : beg + S.totL[sl] + (i0 + c - beg - rankL);
              }
              out[dest] = e[c];
            }
          }
        }
        __syncthreads();
        // ---- next level's nodes: the open children, in node order
        {
          int mine = 0;       // open children of the slots this thread owns (slot = tid, one per thread: kSlots <= kST)
          int lo = 0, ro = 0;
          if (tid < m && S.sfeat[tid] >= 0) {
            const int node = S.node[tid];
            lo = nodes[2 * node + 1].state == kOpen; ro = nodes[2 * node + 2].state == kOpen;
            mine = lo + ro;
          }
          int off, dummy;
          block_excl_scan2(mine, 0, off, dummy, s_scr);
          // children records need this slot's values before the slot arrays are overwritten: read, barrier, write
          int node = 0, beg = 0, end = 0, totL = 0;
          if (tid < m) { node = S.node[tid]; beg = S.beg[tid]; end = S.end[tid]; totL = S.totL[tid]; }
          const bool part = tid < m && S.sfeat[tid] >= 0;
          if (tid == kST - 1) s_i[1] = off + mine;
          // row -> next slot, through the old slot: stash the two children's new slots in baseG / baseH
          if (tid < m) { S.baseG[tid] = part && lo ? off : -1; S.baseH[tid] = part && ro ? off + lo : -1; }
          __syncthreads();
          for (int r = tid; r < ntrain; r += kST) {
            const uint16_t sl = s_pos[r];
            if (sl == kNoSlot) continue;
            const int ns = s_left[r] ? S.baseG[sl] : S.baseH[sl];
            s_pos[r] = (S.sfeat[sl] >= 0 && ns >= 0) ? (uint16_t)ns : kNoSlot;
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
      }
      __syncthreads();

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
