// Host-callable launchers of the engine's kernels (one translation unit per kernel family).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "common.cuh"
#include "../../include/gentun_b200.h"

namespace gt {

struct SketchPlan {
  int buf_cap;               // entries per summary buffer (>= 2 * limit_size + 2)
  int nlevel_alloc;          // summary levels reserved per stream
  int keys_per_stream;       // power of two >= 2 * limit_size
  long long entries_per_stream;
};

// sketch.cu
void launch_sketch(const float* X, long long ldx, int f, int nfold, const uint32_t* const* fold_rows,
                   const long long* fold_ntrain, const SketchPlan& plan, const int* fold_limit,
                   unsigned long long* key_arena, float4* entry_arena, float* cuts, int* ncuts, int stream0,
                   int count, cudaStream_t s);
int configure_sketch();   // cudaFuncSetAttribute for k_sketch_cuts' dynamic shared memory (0 on success); per device, before launch_sketch
void launch_bin(const float* X, long long ldx, int f, int fpad, const uint32_t* rows, long long n,
                const float* cuts, const int* ncuts, uint8_t* bins, cudaStream_t s);
void launch_bins_to_columns(const uint8_t* bins, int f, int fpad, long long n, uint8_t* out, cudaStream_t s);
void launch_ingest_tile(const void* tile, int dtype, int colmajor, long long ld, int rows, int f, float* X, long long ldx,
                        long long r0, cudaStream_t s);
void launch_count_hist(const uint8_t* bins, int f, int fpad, long long ntrain, int* out /* zeroed */, cudaStream_t s);
void launch_gather_y(const float* y, const uint32_t* rows, long long n, float* out, cudaStream_t s);
void launch_count_nonfinite(const float* X, long long n, int f, long long ldx, const float* y, unsigned long long* out /* [2], zeroed */, cudaStream_t s);

// engine.cu: message returned by gt_last_error() for the calling thread
void set_error_message(const char* msg);

// boost.cu
void launch_init_wave(const Wave& w, cudaStream_t s);
void launch_grad(const Wave& w, cudaStream_t s);
void launch_plan_hist(const Wave& w, int level, cudaStream_t s);
void launch_hist_zero(const Wave& w, int level, cudaStream_t s);
int launch_hist(const Wave& w, int level, cudaStream_t s);    // these two return how many kernels they launched
int launch_scan(const Wave& w, int level, cudaStream_t s);
void launch_plan_part(const Wave& w, int level, cudaStream_t s);
void launch_part_count(const Wave& w, int level, cudaStream_t s);
void launch_part_scatter(const Wave& w, int level, cudaStream_t s);
void launch_predict(const Wave& w, cudaStream_t s);
void launch_round_end(const Wave& w, int first, double* fitness_out, int* rounds_out, cudaStream_t s);
void launch_dump_trees(const Wave& w, int round, int tree_rounds, gt_tree_node* out, cudaStream_t s);
void launch_dump_node_rows(const Wave& w, gt_node_rows* out, cudaStream_t s);

// host-known upper bounds of the per-job work-item counts (grids are sized from these; the planners check them on the
// device and raise the wave's error flag if a count ever exceeds its bound).  A level builds one child per split
// parent -- usually the smaller one, but the LARGER one when its sibling was settled as a leaf at once -- so the rows
// built in a level can approach all nt rows: sum of ceil(cnt / kHistRows) <= nt / kHistRows + number of parents.
__host__ __device__ inline int hist_items_bound(const Wave& w, int level) {
  long long nt = w.nt_stride;
  long long b = level == 0 ? (nt + kHistRows - 1) / kHistRows : nt / kHistRows + (1ll << (level - 1)) + 1;
  return (int)(b < w.hist_cap ? b : w.hist_cap);
}
__host__ __device__ inline int part_items_bound(const Wave& w, int level) {
  long long b = (w.nt_stride + kPartRows - 1) / kPartRows + (1ll << level);
  return (int)(b < w.part_cap ? b : w.part_cap);
}
// small.cu
int small_smem_bytes(int n, int nt_stride);
int small_max_teams(int nfold, int smem_bytes);   // teams of nfold CTAs that can be co-resident (0 on failure)
int launch_small_cv(const SmallWave& w, int n_teams, int smem_bytes, cudaStream_t s);   // returns cudaError_t
int hist_smem_bytes();
int configure_kernels();   // cudaFuncSetAttribute for the dynamic-smem kernels; returns cudaError_t

}  // namespace gt
