/*
 * gentun_b200.h -- C ABI of the B200 fitness-evaluation engine.
 *
 * The reference (gentun) has no FFI of its own: its plugin boundary is the Python protocol
 *   GentunModel(x_train, y_train).cross_validate() -> float   (gentun/models/generic_models.py:7-18)
 * implemented for xgboost by XgboostModel (gentun/models/xgboost_models.py:11-37), whose two calls
 * into the third-party library -- xgb.DMatrix (:32) and xgb.cv (:33-36) -- are the boundary this
 * library replaces.  Each entry point below names the reference call it stands in for.
 *
 * Conventions: plain pointers and sizes; the caller owns every host buffer; the engine owns all
 * device memory; every function returns 0 on success or a negative GT_ERR_* code and never throws;
 * gt_last_error() gives the message for the calling thread.
 */
#ifndef GENTUN_B200_H_
#define GENTUN_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GT_OK 0
#define GT_ERR_ARG (-1)    /* bad argument / out-of-domain gene */
#define GT_ERR_CUDA (-2)   /* CUDA runtime failure (message has the CUDA error string) */
#define GT_ERR_NOMEM (-3)  /* device memory budget cannot hold even one individual */
#define GT_ERR_STATE (-4)  /* call order violated */

#define GT_NUM_GENES 11    /* order of XgboostIndividual's genome, gentun/individuals.py:162-175:
                              eta, min_child_weight, max_depth, gamma, max_delta_step, subsample,
                              colsample_bytree, colsample_bylevel, lambda, alpha, scale_pos_weight */
#define GT_MAX_BIN 256
#define GT_MAX_NODES 2047  /* heap-indexed tree of depth <= 10 (max_depth gene range 3..10) */

#define GT_OBJ_SQUAREDERROR 0 /* 'reg:linear' (individuals.py:159) / 'reg:squarederror' */
#define GT_OBJ_LOGISTIC 1     /* 'binary:logistic' */
#define GT_METRIC_RMSE 0
#define GT_METRIC_LOGLOSS 1

typedef struct gt_engine gt_engine;

typedef struct {
  int32_t device;           /* CUDA ordinal */
  int32_t x_on_device;      /* 0: X, y are host pointers (copied H2D); 1: already device pointers */
  int64_t mem_budget_bytes; /* 0 = 85% of free device memory at eval time */
  int32_t reserved[4];      /* [0] != 0: time the histogram kernel with CUDA events (gt_stats.hist_ms);
                               [1] != 0: always use the streaming kernels, even for data small enough for the
                               shared-memory-resident kernel (tests compare the two regimes); rest 0 */
} gt_data_opts;

/* The non-tuned parameters XgboostIndividual forwards (individuals.py:158-160, 208-216). */
typedef struct {
  int32_t objective;
  int32_t metric;
  int32_t nfold;
  int32_t num_boost_round;
  int32_t early_stopping_rounds;
  int32_t reserved[3];
} gt_fit_opts;

/* heap-indexed tree node as dumped for parity tests: feat -2 absent, -1 leaf, >= 0 split feature */
typedef struct {
  int32_t feat;
  int32_t bin;
  float value; /* leaf value (eta * weight) */
  float gain;  /* loss_chg of the split */
} gt_tree_node;

typedef struct {
  uint32_t begin, count; /* segment of the row-set buffer `buf` */
  int32_t buf;
  int32_t state;         /* 0 absent, 1 open, 2 split, 3 leaf */
} gt_node_rows;

/* optional outputs of gt_engine_eval_ex (any pointer may be NULL) */
typedef struct {
  int32_t tree_rounds;      /* dump trees of rounds [0, tree_rounds) */
  gt_tree_node* trees;      /* [n_ind][nfold][tree_rounds][GT_MAX_NODES] */
  double* trace_means;      /* [n_ind][num_boost_round] test-metric mean per round (NaN where not run) */
  int32_t rows_round;       /* dump row sets as they are at the end of this round's tree (-1 = none) */
  uint32_t* row_sets;       /* [n_ind][nfold][2][row_stride] both ping-pong row-set buffers */
  gt_node_rows* node_rows;  /* [n_ind][nfold][GT_MAX_NODES] */
  int64_t row_stride;       /* elements per row-set buffer (>= max train rows) */
} gt_dump;

typedef struct {
  double hist_ms;           /* summed device time of the histogram kernel (CUDA events on the engine stream) */
  double hist_alg_bytes;    /* algorithmic bytes those launches moved (DESIGN.md "histogram") */
  double hist_row_features; /* sum over tiles of rows x features accumulated */
  int64_t hist_launches;
  int64_t kernel_launches;  /* all kernels of this library launched by the last eval */
  double eval_ms;           /* device time of the last gt_engine_eval */
  double prepare_ms;        /* device time of sketch + binning for the last prepared nfold */
  int64_t waves;
  int64_t jobs;             /* individual x fold jobs of the last eval */
} gt_stats;

/* Replaces xgb.DMatrix(x_train, label=y_train) (xgboost_models.py:32), once per dataset instead of
 * once per individual.  X is row-major float32 [n][ldx] (first f columns used), y float32 [n]. */
int gt_engine_create(const float* X, int64_t n, int32_t f, int64_t ldx, const float* y,
                     const gt_data_opts* opts, gt_engine** out);

/* The same call for the arrays gentun's scripts really pass (tests/test_wine-quality.py:17-19 hands xgb.DMatrix a
 * pandas DataFrame of float64, which is column-major in memory, and an int64 Series): host X in its own element
 * type and layout, no host-side copy or cast by the caller.  Strides are in elements; either x_col_stride == 1
 * (row-major, x_row_stride >= f) or x_row_stride == 1 (column-major, x_col_stride >= n).  The matrix is staged
 * tile by tile through pinned buffers the library keeps, cast to float on the device (the cast DMatrix applies)
 * and laid out row-major there.  y is cast on the host.  opts->x_on_device must be 0.  While the matrix moves, a
 * helper thread of the engine draws xgb.cv's row permutation (it depends on n alone); gt_engine_prepare and
 * gt_engine_destroy join it.  With GENTUN_B200_TRACE set in the environment, create and prepare print the
 * wall-clock time of their phases on stderr. */
#define GT_DTYPE_F32 0
#define GT_DTYPE_F64 1
#define GT_DTYPE_I64 2   /* labels only */
int gt_engine_create_from(const void* X, int32_t x_dtype, int64_t n, int32_t f, int64_t x_row_stride,
                          int64_t x_col_stride, const void* y, int32_t y_dtype, const gt_data_opts* opts,
                          gt_engine** out);

/* Replaces what xgb.cv derives from (data, nfold) alone (training.py mknfold; per-slice quantile
 * sketch and GHistIndexMatrix built lazily by each Booster): folds, cut points, uint8 bin matrices.
 * Cached per nfold; gt_engine_eval calls it on demand. */
int gt_engine_prepare(gt_engine* e, int32_t nfold);

/* Replaces xgb.cv(params, d_train, num_boost_round, early_stopping_rounds, nfold) followed by
 * cv_result['test-<metric>-mean'][last] (xgboost_models.py:33-37) for n_ind gene sets at once --
 * the batch that Population.get_fittest (gentun/populations.py:55-58) evaluates serially.
 * genes: [n_ind][GT_NUM_GENES] doubles; fitness_out: [n_ind]; rounds_out: [n_ind] rounds trained (may be NULL). */
int gt_engine_eval(gt_engine* e, int32_t n_ind, const double* genes, const gt_fit_opts* opts,
                   double* fitness_out, int32_t* rounds_out);

/* Same, with parity-test dumps. */
int gt_engine_eval_ex(gt_engine* e, int32_t n_ind, const double* genes, const gt_fit_opts* opts,
                      double* fitness_out, int32_t* rounds_out, const gt_dump* dump);

void gt_engine_destroy(gt_engine* e);

/* Parity exports (X1-X3 of SURVEY.md 8a).  k is the fold number; rows are original row ids in
 * DMatrix.slice order. train_rows [ntrain], test_rows [ntest]; cuts [f][GT_MAX_BIN]; ncuts [f];
 * bins [n][f] row-major in fold order (train slice rows, then test slice rows). */
int gt_engine_fold_sizes(gt_engine* e, int32_t nfold, int32_t k, int64_t* ntrain, int64_t* ntest);
int gt_engine_get_fold_rows(gt_engine* e, int32_t nfold, int32_t k, int64_t* train_rows, int64_t* test_rows);
int gt_engine_get_cuts(gt_engine* e, int32_t nfold, int32_t k, float* cuts, int32_t* ncuts);
int gt_engine_get_bins(gt_engine* e, int32_t nfold, int32_t k, uint8_t* bins);

int gt_engine_get_stats(gt_engine* e, gt_stats* out);

/* Host-only: np.random.seed(seed); np.random.permutation(n); np.array_split(., nfold)
 * (xgboost training.py mknfold as called from xgboost_models.py:33-36). perm [n]; fold_begin [nfold+1]. */
int gt_make_folds(int64_t n, int32_t nfold, uint32_t seed, int64_t* perm, int64_t* fold_begin);

/* No counterpart in the reference (gentun never manages device memory).  Cut points of ONE column for
 * callers that may not allocate: a serving loop re-binning a feature while other streams run (cudaMalloc
 * and cudaFree stall the whole device), a process under a fixed memory reservation, a hook that may only
 * enqueue work.  Same sketch kernel as gt_engine_prepare, but every byte of scratch comes from a fixed
 * __device__ array inside the library (about 2.2 MB per GPU, held until exit) -- no cudaMalloc, no
 * cudaFree, no allocating library call.  d_values [n] (device, the column in DMatrix row order),
 * d_cuts [GT_MAX_BIN] and d_ncuts [1] (device) are the caller's; work is enqueued on `stream`
 * (a cudaStream_t, NULL = default stream) and the call returns without synchronising.  Only d_cuts[0 .. *d_ncuts)
 * is written; the rest of the caller's GT_MAX_BIN floats is left as it was.
 * Limits: 1 <= n <= gt_column_cuts_noalloc_max_rows() (65536), else GT_ERR_ARG.  All calls share the
 * one scratch array: issue them on one stream, or otherwise keep them from overlapping on the GPU --
 * overlapping calls race.  The first call on each device looks the array's address up
 * (cudaGetSymbolAddress); make that first call before any stream capture begins. */
int gt_column_cuts_noalloc(const float* d_values, int64_t n, float* d_cuts, int32_t* d_ncuts, void* stream);
int64_t gt_column_cuts_noalloc_max_rows(void);

const char* gt_last_error(void);
int gt_version(void);

#ifdef __cplusplus
}
#endif
#endif /* GENTUN_B200_H_ */
