// Shared device-side types of the gentun_b200 engine (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace gt {

constexpr int kBins = 256;          // max_bin (xgboost default, never overridden by gentun)
constexpr int kLanes = 32;          // features per histogram group == smem banks
constexpr int kMaxDepth = 10;       // max_depth gene range 3..10 (gentun/individuals.py:167)
constexpr int kNodeCap = 2048;      // heap slots per tree (2^(10+1)-1 used)
#ifndef GT_HIST_ROWS
#define GT_HIST_ROWS 16384
#endif
constexpr int kHistRows = GT_HIST_ROWS;   // rows per histogram work item
constexpr int kEpochRows = 2048;    // rows between int32-overflow checks of the shared partial sums
constexpr int kNarrowRows = 4095;   // nodes up to this many rows keep int32 histogram cells whatever their gradients: 4095 * 2^19 < 2^31
#ifndef GT_PART_ROWS
#define GT_PART_ROWS 8192
#endif
constexpr int kPartRows = GT_PART_ROWS;   // rows per partition work item
constexpr int kQuantBits = 19;      // |quantised gradient|, quantised hessian <= 2^19 (one epoch adds < 2^30)
constexpr int kQuantBitsH = 17;     // hessians are near-constant (1 for squared error): fewer bits, fewer flushes
constexpr int kPredRows = 2048;     // rows per predict CTA
constexpr int kMaskWords = 8;       // feature masks: up to 256 features
constexpr int kMaxWaveInds = 4096;  // individuals per wave (k_predict keeps their job list in shared memory)
#ifndef GT_SMALL_THREADS
#define GT_SMALL_THREADS 1024
#endif
constexpr int kSmallThreads = GT_SMALL_THREADS;  // the shared-memory-resident regime (small.cu): threads per job (one CTA per SM)
constexpr int kSmallMaxTrain = 4095; // train rows per fold it takes: 4095 * 2^19 < 2^31 keeps every prefix sum in int32, 8 list elements per thread
constexpr int kSmallMaxRows = 8192;  // rows (train + test) it takes: predictions live in shared memory

enum NodeState : uint8_t { kAbsent = 0, kOpen = 1, kSplit = 2, kLeaf = 3 };

struct HistEntry { long long g, h; };   // fixed-point sums, one (feature, bin) cell, 16 B

struct Node {          // working record of one heap slot
  long long G, H;      // quantised gradient / hessian sums of the node
  uint32_t begin, cnt; // segment of row-set buffer `buf`
  float leaf;          // eta * weight
  float root_gain;     // CalcGain(G, H)
  float loss_chg;
  int16_t feat;
  uint8_t bin;
  uint8_t state;
  uint8_t buf;
  uint8_t narrow;      // histogram cells are int32 pairs (decided when the node gets its rows)
  uint8_t pad[2];
  unsigned long long absG;   // sum of |g_q| over the node's rows: bounds every cell of its histogram and of its children's
};

struct TreeNode {      // compact record read by the prediction kernel, 8 B
  float leaf;
  int16_t feat;
  uint8_t bin;
  uint8_t state;
};

struct HistItem { int node; uint32_t begin; uint32_t cnt; int flags; };  // flags bit0: shared (atomic flush), bit1: buf
struct PairRec { int parent; int build_right; int single; int job; };   // single: only the built child is evaluated (its sibling is already a leaf)
struct PartItem { int node; uint32_t begin; uint32_t cnt; int first_item; int n_items; uint32_t node_begin; uint32_t node_cnt; int buf; };

struct JobParams {     // constant over one evaluation
  float eta, mcw, gamma, mds, subsample, coltree, collevel, lambda, alpha, spw;
  float base_margin;
  int max_depth;
  int fold;
  int ind;             // index inside the wave
  int uniform_h;       // every train row of the fold has the same hessian under this job's genes (see k_hist)
  unsigned long long hist_off[2];   // element offsets of the even / odd level histogram buffers
};

struct JobState {
  int active;
  int n_hist_items, n_pairs, n_part_items, n_zero;
  unsigned gmax_bits, hmax_bits;    // float bit patterns, atomicMax'ed
  float sg, sh;                     // quantisation scales (powers of two)
  double inv_sg, inv_sh;
  uint32_t tmask[kMaskWords];       // colsample_bytree set of the current tree
  uint32_t lmask[kMaskWords];       // colsample_bylevel set of the current level
  int n_tree_feats;
};

struct IndState {
  int active;
  int round;          // next round to train
  int best_iter;
  int stopped;
  double best_score;
  double last_mean;
  double fitness;
  int rounds_run;
  int out_index;      // position in the caller's arrays
};

struct FoldDev {
  const uint8_t* bins;   // [n][fpad] fold order (train slice rows then test slice rows)
  const uint8_t* bins_col;  // [f][n] feature-major copy, read by the partition kernels
  const float* y;        // [n] fold order
  const float* cuts;     // [f][256]
  const int* ncuts;      // [f]
  const int* count_hist; // [groups][256][32] train rows per (feature, bin), laid out like a node histogram
  long long ntrain, ntest;
};

struct Wave {            // passed by value to kernels
  int n_jobs, n_inds, nfold;
  long long n;           // rows
  long long nt_stride;   // row-set stride (>= max ntrain, multiple of 32)
  int f, fpad, groups;
  int objective, metric, num_boost_round, early_stopping_rounds;
  int hist_cap, pair_cap, part_cap;  // per-job record capacities
  int pred_chunks;
  int lvl_jobs[kMaxDepth + 2];   // jobs whose max_depth exceeds the level; jobs are ordered deepest first, so they form a prefix
  const FoldDev* folds;
  const int* grp_maxnc;  // [nfold][groups] max ncuts inside the group
  JobParams* jp;
  JobState* js;
  IndState* is;
  Node* nodes;           // [n_jobs][kNodeCap]
  TreeNode* tree;        // [n_jobs][kNodeCap]
  float* pred;           // [n_jobs][n]
  int2* gq;              // [n_jobs][nt_stride]
  uint32_t* ridx;        // [n_jobs][2][nt_stride]
  HistEntry* hist;       // arena
  HistItem* hitems;      // [n_jobs][hist_cap]
  int* zlist;            // [n_jobs][hist_cap] nodes whose histogram must be cleared
  PairRec* pairs;        // [n_jobs * pair_cap] pairs of the current level, all jobs, appended by the planners
  PartItem* pitems;      // [n_jobs][part_cap]
  int* pleft;            // [n_jobs][part_cap]
  uint32_t* pflags;      // [n_jobs][part_cap][kPartRows / 32] goes-left bit of every row of a partition item
  double* mpart;         // [n_jobs][pred_chunks] per-CTA metric partial sums
  double* trace;         // [n_inds][num_boost_round] or null
  unsigned long long* counters;  // [0] rows through hist, [1] node histograms built, [2] active individuals, [5 + level parity] pairs of the level, [7] work-list overflow (an error)
};

struct SmallWave {       // arguments of the shared-memory-resident kernel (small.cu), passed by value
  int n_inds, nfold;
  int n, f, fpad, nt_stride;
  int objective, metric, num_boost_round, early_stopping_rounds;
  const FoldDev* folds;
  const uint32_t* base_lists;   // [nfold][f][nt_stride] (bin << 16 | row) of the train slice, sorted by bin, ties by row
  const JobParams* jp;          // [n_inds][nfold]
  const IndState* is;           // [n_inds] (out_index)
  Node* nodes;                  // [ctas][kNodeCap]
  TreeNode* tree;               // [ctas][kNodeCap]
  uint32_t* lists;              // [ctas][2][f][nt_stride] ping-pong copies of the per-feature lists
  double* fold_s;               // [teams][2][nfold] per-round metric sums of a team's folds
  unsigned* team_bar;           // [teams] barrier counters (zeroed before the launch)
  int* team_ind;                // [teams][2]
  int* queue;                   // [1] next position in `order` (zeroed before the launch)
  const int* order;             // [n_inds] individuals, most expensive first
  double* trace;                // [n_inds][num_boost_round] or null
  double* fitness_out;
  int* rounds_out;
};

// counter-based generator (definition shared with the oracle by specification)
__host__ __device__ inline uint64_t mix64(uint64_t z) {
  z += 0x9e3779b97f4a7c15ull;
  z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
  z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
  return z ^ (z >> 31);
}
__host__ __device__ inline uint64_t stream_key(uint32_t tag, uint32_t fold, uint32_t round, uint32_t level) {
  return mix64(((uint64_t)tag << 60) ^ ((uint64_t)fold << 48) ^ ((uint64_t)level << 40) ^ (uint64_t)round);
}
__host__ __device__ inline uint64_t draw64(uint64_t key, uint64_t index) { return mix64(key ^ (index * 0xd1342543de82ef95ull)); }
__host__ __device__ inline float u01(uint64_t h) { return (float)(h >> 40) * (1.0f / 16777216.0f); }

}  // namespace gt
