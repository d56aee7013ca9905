// Split arithmetic and node bookkeeping shared by the streaming kernels (boost.cu) and the shared-memory-resident
// small-data kernel (small.cu): xgboost src/tree/param.h (CalcWeight, CalcGainGivenWeight, ThresholdL1),
// src/tree/split_evaluator.h, src/objective/regression_obj.cu, hist/evaluate_splits.h (SplitEntry::NeedReplace,
// CPUExpandEntry::IsValid).  One definition, so both regimes make the same decisions on the same sums.
#pragma once
#include "common.cuh"

namespace gt {
namespace {

constexpr float kRtEps = 1e-6f;

// ------------------------------------------------------------------ split arithmetic
__device__ __forceinline__ double thr_l1(double w, float alpha) {
  if (w > +alpha) return w - alpha;
  if (w < -alpha) return w + alpha;
  return 0.0;
}

__device__ __forceinline__ float calc_weight(const JobParams& p, double G, double H) {
  if (H < p.mcw || H <= 0.0) return 0.0f;
  double dw = -thr_l1(G, p.alpha) / (H + p.lambda);
  if (p.mds != 0.0f && fabs(dw) > p.mds) dw = copysign((double)p.mds, dw);
  return (float)dw;
}

__device__ __forceinline__ float gain_given_weight(const JobParams& p, double G, double H, float w) {
  if (H <= 0) return 0.0f;
  if (p.mds == 0.0f) {
    double t = thr_l1(G, p.alpha);
    return (float)((t * t) / (H + p.lambda));
  }
  float g = (float)G, h = (float)H;
  return -(2.0f * g * w + (h + p.lambda) * (w * w));
}

__device__ __forceinline__ float calc_gain(const JobParams& p, double G, double H) {
  return gain_given_weight(p, G, H, calc_weight(p, G, H));
}

__device__ __forceinline__ float split_gain(const JobParams& p, double GL, double HL, double GR, double HR) {
  float wl = calc_weight(p, GL, HL), wr = calc_weight(p, GR, HR);
  return gain_given_weight(p, GL, HL, wl) + gain_given_weight(p, GR, HR, wr);
}

__device__ __forceinline__ float sigmoid_x(float x) {   // upstream common::Sigmoid
  x = fminf(-x, 88.7f);
  float denom = expf(x) + 1.0f + 1e-16f;
  return 1.0f / denom;
}

__device__ __forceinline__ void grad_pair(int objective, float pred, float y, float spw, float& g, float& h) {
  float wgt = (y == 1.0f) ? spw : 1.0f;
  if (objective == 1) {
    float p = sigmoid_x(pred);
    g = (p - y) * wgt;
    h = fmaxf(p * (1.0f - p), 1e-16f) * wgt;
  } else {
    g = (pred - y) * wgt;
    h = 1.0f * wgt;
  }
}

__device__ __forceinline__ int level_first(int level) { return (1 << level) - 1; }

// ------------------------------------------------------------------ per-level planning
// ColumnSampler stand-in: keep the n candidates with the smallest (draw, index).
__device__ void sample_mask(uint64_t key, const uint32_t* in_mask, int f, int n_keep, uint32_t* out_mask /* smem */) {
  for (int q = threadIdx.x; q < kMaskWords; q += blockDim.x) out_mask[q] = 0u;
  __syncthreads();
  for (int jf = threadIdx.x; jf < f; jf += blockDim.x) {
    if (!((in_mask[jf >> 5] >> (jf & 31)) & 1u)) continue;
    uint64_t kj = draw64(key, (uint64_t)jf);
    int rank = 0;
    for (int i = 0; i < f; ++i) {
      if (!((in_mask[i >> 5] >> (i & 31)) & 1u)) continue;
      uint64_t ki = draw64(key, (uint64_t)i);
      rank += (ki < kj) || (ki == kj && i < jf);
    }
    if (rank < n_keep) atomicOr(&out_mask[jf >> 5], 1u << (jf & 31));
  }
  __syncthreads();
}

struct Cand { float chg; int f; int b; long long GL, HL; };

__device__ __forceinline__ bool better(float chg, int f, int b, float ochg, int of, int ob) {
  // SplitEntry::NeedReplace under an ascending (feature, bin) scan: strictly greater gain wins,
  // equal gain keeps the earlier (feature, bin)
  if (chg > ochg) return true;
  if (chg < ochg || f < 0) return false;
  if (of < 0) return true;
  return f < of || (f == of && b < ob);
}

// The scan runs in two stages.  Stage 1 walks the bins with an fp32 ESTIMATE of every candidate's gain and keeps, per
// lane, the best and second-best estimate and the prefix sums at the best.  Stage 2 re-evaluates, in the
// reference's double/float expression order, only the candidates whose estimate is within the estimate's error
// margin of the best one.  Stage 1 is the instruction-bound part, so it works on register copies of the three
// regularisation constants, uses a bare reciprocal, is compiled separately for max_delta_step == 0 / != 0, and
// runs in 32-bit integers when every histogram of the pair is in the narrow (int32) format.
struct GainK { float alpha, lambda, mds; };

struct ChildTot { long long Gt, Ht; float root_gain; };

__device__ __forceinline__ float rcp_fast(float x) {
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}

// fp32 estimate of one side's CalcGain (same formulas as gain_given_weight(calc_weight)) from the side's gradient sum
// G and its denominator d = H + lambda, branch-free: ThresholdL1 as G - clamp(G, -alpha, alpha).  Never negative.
template <bool MDS>
__device__ __forceinline__ float gain_term_k(const GainK& k, float G, float d) {
  const float t = G - fminf(fmaxf(G, -k.alpha), k.alpha);
  const float rd = rcp_fast(d);
  if (!MDS) return (t * t) * rd;
  const float wgt = fminf(fmaxf(-(t * rd), -k.mds), k.mds);
  return -(2.0f * G * wgt + d * (wgt * wgt));
}

// estimate of a candidate from real-valued prefix sums (stage 2's re-walks); an empty side (H <= 0) contributes nothing
template <bool MDS>
__device__ __forceinline__ float split_est(const GainK& k, float gl, float hl, float gr, float hr) {
  const float a = hl > 0.f ? gain_term_k<MDS>(k, gl, hl + k.lambda) : 0.f;
  const float r = hr > 0.f ? gain_term_k<MDS>(k, gr, hr + k.lambda) : 0.f;
  return a + r;
}

__device__ __forceinline__ void exact_candidate(Cand& best, const JobParams& jp, const ChildTot& t, long long GL, long long HL, int f, int b,
                                                double inv_sg, double inv_sh) {
  const double gl = (double)GL * inv_sg, hl = (double)HL * inv_sh;
  const double gr = (double)(t.Gt - GL) * inv_sg, hr = (double)(t.Ht - HL) * inv_sh;
  if (hl >= jp.mcw && hr >= jp.mcw) {
    const float chg = split_gain(jp, gl, hl, gr, hr) - t.root_gain;
    if (!isinf(chg) && chg > best.chg) { best.chg = chg; best.f = f; best.b = b; best.GL = GL; best.HL = HL; }
  }
}

// The evaluated node becomes a split or a leaf (CPUExpandEntry::IsValid), and a split's two children get their
// totals, weight and gain.  Called by two lanes: sd = 0 sets up the left child and writes the node itself, sd = 1
// the right child.
__device__ __forceinline__ void settle_node(const JobParams& jp, int level, Node* nodes, TreeNode* tree, int node, const ChildTot& t,
                                            const Cand& best, double inv_sg, double inv_sh, long long hmin_q, int sd) {
  Node& nd = nodes[node];
  const long long Gt = t.Gt, Ht = t.Ht;
  bool valid = best.f >= 0 && best.chg > kRtEps;
  if (valid) {
    const double hl = (double)best.HL * inv_sh, hr = (double)(Ht - best.HL) * inv_sh;
    if (hl == 0.0 || hr == 0.0) valid = false;
    if (best.chg < jp.gamma) valid = false;
  }
  float own_leaf = nd.leaf;
  if (level == 0) own_leaf = jp.eta * calc_weight(jp, (double)Gt * inv_sg, (double)Ht * inv_sh);
  if (sd == 0) {
    if (level == 0) { nd.root_gain = t.root_gain; nd.leaf = own_leaf; }
    nd.loss_chg = best.chg;
    if (!valid) {
      nd.state = kLeaf;
      tree[node] = TreeNode{own_leaf, (int16_t)-1, (uint8_t)0, (uint8_t)kLeaf};
    } else {
      nd.state = kSplit; nd.feat = (int16_t)best.f; nd.bin = (uint8_t)best.b;
      tree[node] = TreeNode{0.f, (int16_t)best.f, (uint8_t)best.b, (uint8_t)kSplit};
    }
  }
  if (!valid) return;
  const long long cgq = sd == 0 ? best.GL : Gt - best.GL, chq = sd == 0 ? best.HL : Ht - best.HL;
  // Driver::IsChildValid; in addition a child whose hessian sum is below 2 * min_child_weight has no valid
  // candidate (both sides need H >= min_child_weight), so evaluating it could only make it a leaf: settle that now
  // and spare its histogram, scan and partition
#ifdef GT_DBG_NO_EARLY_LEAF
  const bool can_split = true;
#else
  const bool can_split = hmin_q <= 0 || chq >= 2 * hmin_q;
#endif
  const uint8_t child_state = (level + 1 < jp.max_depth && can_split) ? kOpen : kLeaf;
  Node& chn = nodes[2 * node + 1 + sd];
  const double g = (double)cgq * inv_sg, h = (double)chq * inv_sh;
  chn.G = cgq; chn.H = chq;
  chn.leaf = jp.eta * calc_weight(jp, g, h);
  chn.root_gain = calc_gain(jp, g, h);
  chn.loss_chg = 0.f; chn.feat = -1; chn.bin = 0; chn.state = child_state;
  chn.begin = 0; chn.cnt = 0; chn.buf = 0; chn.narrow = 0; chn.absG = 0ull;
  tree[2 * node + 1 + sd] = TreeNode{chn.leaf, (int16_t)-1, (uint8_t)0, child_state};
}

}  // namespace
}  // namespace gt
