/*
 * svmb200.h -- C ABI of the B200-native RBF-kernel SMO trainer (libsvmb200.so).
 *
 * This is the drop-in boundary for the reference's solver path.  The reference has
 * no FFI; the interface it does have is the free function
 *
 *     void SMO_train(const vector<double>& x_train, const vector<int>& y_train, size_t n_features,
 *                    vector<double>& alpha, double& b, double C = 10.0 [, bool init = true]);
 *                                            (code/main3.cpp:162, code/mpi_svm_main2.cpp:154)
 *
 * plus find_min_max / scale_features (code/main3.cpp:57,74), get_SV_indices (:297) and the
 * prediction loop of main() (:391-405).  Each entry point below names the reference lines it
 * replaces.  All pointers are HOST pointers unless the name says "dev"; every buffer is owned by
 * the caller.  Every function returns 0 on success and a negative SVMB200_E* code on failure (the
 * message is available from svmb200_last_error).  As in the reference, abnormal solver stops are
 * not errors: they are reported in svmb200_stats.stop_reason (code/main3.cpp:205-209,246-257,
 * 283-287 print to cerr, break, and still return b).
 *
 * Not thread-safe per context; independent contexts may be used from different threads.
 * There is no CPU fallback: without a CUDA device every call fails with SVMB200_ENODEVICE.
 */
#ifndef SVMB200_H
#define SVMB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SVMB200_OK          0
#define SVMB200_EINVAL     -1   /* bad argument                                   */
#define SVMB200_ECUDA      -2   /* a CUDA runtime call failed                     */
#define SVMB200_ENODEVICE  -3   /* no usable CUDA device                          */
#define SVMB200_ESTATE     -4   /* call order violated (e.g. train before upload) */
#define SVMB200_ELIMIT     -5   /* size beyond a documented limit                 */

/* stop_reason values: the five ways code/main3.cpp's loop ends */
#define SVMB200_STOP_CONVERGED   0  /* b_low <= b_high + 2 tau              main3.cpp:213 */
#define SVMB200_STOP_NOT_FOUND   1  /* i_high or i_low not found            main3.cpp:205 */
#define SVMB200_STOP_INFEASIBLE  2  /* infeasible U and V                   main3.cpp:246 */
#define SVMB200_STOP_ETA         3  /* non positive eta                     main3.cpp:253 */
#define SVMB200_STOP_MAX_ITER    4  /* num_iter > max_iter                  main3.cpp:283 */
#define SVMB200_STOP_PEER_LOST   5  /* a record of the exchange never arrived (~3 s): a peer GPU is gone (row sharding), or part of
                                      this GPU's own grid never ran (the GPU is shared with other work).  No counterpart in the reference */

typedef struct svmb200_ctx svmb200_ctx;

/* The reference's literals as parameters; svmb200_default_params fills in those literals:
 * C=10 (main3.cpp:163,342) gamma=0.00125 (:95) eps=1e-12 (:109,128,165) tau=1e-5 (:196)
 * max_iter=100000 (:198) sv_tol=1e-8 (:297,367). */
typedef struct {
    double  C;
    double  gamma;
    double  eps;
    double  tau;
    int32_t max_iter;
    double  sv_tol;
    int32_t wss;          /* working-set selection.  1 = first-order maximal violating pair, the only rule the reference has
                             (main3.cpp:107-142).  2 = second-order choice of i_low (Fan, Chen, Lin 2005, the rule of LIBSVM; NOT
                             in the reference, parity unpinned): i_high and the stop test as in 1, i_low = the row of I_low with
                             f_t > b_high that maximises (f_t - b_high)^2 / (2 - 2 K(i_high, t)), lowest index on ties, rows with
                             a curvature <= eps passed over; needs K(i_high, .) before i_low is known, i.e. one more pass (or
                             row-cache read) and one more exchange per iteration, for about a third fewer iterations.
                             Other values: EINVAL. */
    int32_t warm_start;   /* 0: alpha := 0, f := -y (main3.cpp:167-172).  1: alpha is in/out and f is
                             rebuilt from it (mpi_svm_main2.cpp:167-183, "init = false"). */
    int32_t cache_rows;   /* HBM kernel-row cache capacity in rows; 0 = off (the reference's behaviour:
                             only the previous row is reused, main3.cpp:216-232); -1 = auto. */
    int32_t accum_chunk;  /* fp32 products summed before widening to FP64: 1, 2 or 4 (DESIGN.md
                             "Numerics"); 0 = library default. */
    int32_t trace_cap;    /* record the first trace_cap (i_high, i_low) pairs for svmb200_get_trace */
    int32_t l2_persist;   /* 0 = stream X with evict-first; 1 = keep a leading fraction of every
                             block's rows resident in L2 (evict-last); -1 = auto */
    int32_t keep_cache;   /* 1 = start from the kernel rows the previous svmb200_smo_train on this X left in the
                             row cache (kernel values do not depend on the labels or on C): one-vs-rest over
                             several digits, C sweeps.  Needs the same cache_rows, accum_chunk and gamma. */
} svmb200_params;

typedef struct {
    int32_t num_iter;        /* the reference's printed counter: starts at 1, +1 per update (main3.cpp:197,281) */
    int32_t stop_reason;     /* SVMB200_STOP_* */
    int64_t i_high, i_low;   /* last selected pair (n if not found) */
    double  b_high, b_low;
    double  smo_ms;          /* device time of the SMO loop alone (CUDA events) */
    double  init_ms;         /* device time of state init (+ warm-start f rebuild) */
    int64_t x_passes;        /* iterations that streamed X (both kernel rows recomputed) */
    int64_t cache_hits;      /* iterations served entirely from the row cache */
    int64_t x_bytes_streamed;/* x_passes * n * row_pitch_bytes */
    int32_t n_sv;            /* count of alpha > sv_tol */
    int32_t grid_blocks;     /* persistent grid actually launched */
    int32_t launches;        /* kernels launched by this call */
    int32_t cluster_size;    /* CTAs per thread-block cluster of the persistent grid (2; 1 when the device refused clusters) */
} svmb200_stats;

void svmb200_default_params(svmb200_params *p);

/* Context bound to one CUDA device ordinal. */
int  svmb200_create(svmb200_ctx **out, int device);
void svmb200_destroy(svmb200_ctx *ctx);
const char *svmb200_last_error(const svmb200_ctx *ctx);   /* ctx may be NULL: last create error */
const char *svmb200_version(void);

/* ---- ingest + preprocessing: replaces find_min_max + scale_features (main3.cpp:57-89,338-339)
 * and the reference GPU's four scratch copies of X (gpu_svm_main3.cu:538-598).  X is row-major
 * [n x d]; y is +-1 (main3.cpp:49-52 maps "label != 1 -> -1" before the solver sees it).  The
 * device computes per-feature min/max in FP64, scales (x-min)/range with range<1e-12 -> 1 in FP64
 * and stores fp32 rows padded to a multiple of 4 floats.  min_out/max_out (optional, d doubles)
 * receive the statistics; they are also kept in the context for svmb200_predict_*. */
int svmb200_set_train_u8 (svmb200_ctx *ctx, const uint8_t *X, const int32_t *y, int64_t n, int32_t d,
                          double *min_out, double *max_out);
int svmb200_set_train_f64(svmb200_ctx *ctx, const double  *X, const int32_t *y, int64_t n, int32_t d,
                          double *min_out, double *max_out);
/* New +-1 labels for the rows already uploaded (the reference handles one digit, main3.cpp:49-52; one-vs-rest over all
 * ten digits is ten label vectors on the same X).  The kernel-row cache survives (see svmb200_params.keep_cache). */
int svmb200_set_labels(svmb200_ctx *ctx, const int32_t *y);

/* Storage precision of X for the NEXT svmb200_set_train_* call: 0 = the default (SVMB200_STORAGE_AUTO, below);
 * 1 = FP64 rows ("exact"): twice the bytes of fp32 rows per iteration, K formed from FP64 squared
 * differences as in main3.cpp:97-100, so that the selected (i_high, i_low) sequence is the reference's own
 * (DESIGN.md "Numerics").  Exact storage takes d <= 832. */
int svmb200_set_exact(svmb200_ctx *ctx, int32_t enable);

/* The same choice with a third option.  SVMB200_STORAGE_U8 (svmb200_set_train_u8 only; other uploads then fail with
 * SVMB200_ESTATE) keeps X as the bytes it arrived as -- a quarter of the fp32 bytes, so 60 000 x 784 rows stay in L2 from one
 * iteration to the next.  For integer-valued input min and range are integers and the reference's scaled squared difference
 * (main3.cpp:74-89, :97-100) is (u_k - p_k)^2 / range_k^2: the features that share the most frequent range are summed as exact
 * integers (four products per IDP4A), the rest in FP64 with their weights (DESIGN.md "Numerics").  K agrees with the reference's
 * to ~1e-16 and the solver walks the reference's own (i_high, i_low) sequence, like FP64 storage, at a fraction of the bytes.
 * The decision function (svmb200_predict_*, warm start) then works on scaled FP64 rows.  d <= 832.
 * SVMB200_STORAGE_AUTO, the default of a new context: uint8 storage for svmb200_set_train_u8 with d <= 832, fp32 rows for
 * every other upload. */
#define SVMB200_STORAGE_AUTO (-1)
#define SVMB200_STORAGE_F32 0
#define SVMB200_STORAGE_F64 1
#define SVMB200_STORAGE_U8  2
int svmb200_set_storage(svmb200_ctx *ctx, int32_t storage);
/* layout of the packed uint8 rows of the current training set (introspection): the range shared by the main class
 * (0 = none), the features in it, the features in the weighted tail, bytes per packed row */
int svmb200_get_u8_layout(svmb200_ctx *ctx, int32_t *main_range, int32_t *main_features, int32_t *tail_features, int32_t *row_bytes);

/* X already scaled by the caller -- exactly what the reference's SMO_train receives
 * (main3.cpp:347).  No statistics are computed; svmb200_predict_scaled_f64 is the matching call. */
int svmb200_set_train_scaled_f64(svmb200_ctx *ctx, const double *Xs, const int32_t *y, int64_t n, int32_t d);

/* The same two steps on their own, for hosts that keep the reference's call sequence
 * find_min_max(x, min, max, n, d); scale_features(x, min, max, n, d) on HOST vectors (main3.cpp:338-339,
 * :355 for the test set): min/max of X [n x d], and in-place FP64 scaling (x - min) / range with
 * range < 1e-12 -> 1.  Both run on the device in FP64 and return the reference's values bit for bit. */
int svmb200_minmax_f64(svmb200_ctx *ctx, const double *X, int64_t n, int32_t d, double *min_out, double *max_out);
int svmb200_scale_f64(svmb200_ctx *ctx, double *X_inout, const double *min_val, const double *max_val, int64_t n, int32_t d);

/* ---- the solver: replaces SMO_train (main3.cpp:162-294; warm start mpi_svm_main2.cpp:154-288).
 * alpha: n doubles, written on return; read first when params->warm_start != 0.  b: (b_high+b_low)/2
 * (main3.cpp:291).  stats may be NULL. */
int svmb200_smo_train(svmb200_ctx *ctx, const svmb200_params *params, double *alpha, double *b,
                      svmb200_stats *stats);

/* ---- replaces get_SV_indices (main3.cpp:297-304) on the last trained alpha: ascending row
 * indices with alpha > sv_tol.  ids may be NULL to query the count. */
int svmb200_sv_ids(svmb200_ctx *ctx, int32_t *ids, int32_t cap, int32_t *count);

/* ---- decision function + accuracy: replaces main3.cpp:391-405 (ge_zero = 0: "curr > 0") and
 * mpi_svm_main2.cpp:721-733 (ge_zero = 1: "s >= 0") and the reference GPU's predict/reduce_sum
 * (gpu_svm_main3.cu:277-315).  Test rows are scaled with the TRAIN statistics (main3.cpp:355).
 * Uses the support vectors and b of the last svmb200_smo_train.  y_test may be NULL (then *correct
 * is 0); scores (optional) receives m decision values sum_k alpha_k y_k K(x, x_k) - b. */
int svmb200_predict_u8 (svmb200_ctx *ctx, const uint8_t *X, const int32_t *y_test, int64_t m,
                        int32_t ge_zero, int32_t *correct, double *scores);
int svmb200_predict_f64(svmb200_ctx *ctx, const double  *X, const int32_t *y_test, int64_t m,
                        int32_t ge_zero, int32_t *correct, double *scores);
int svmb200_predict_scaled_f64(svmb200_ctx *ctx, const double *Xs, const int32_t *y_test, int64_t m,
                               int32_t ge_zero, int32_t *correct, double *scores);

/* ---- model persistence: the dump the reference left commented out (mpi_svm_main2.cpp:686-697: SV ids,
 * labels, alphas, b) as one binary file that also holds the scaled SV rows and the train statistics.
 * svmb200_load_model turns the context into a predict-only one (its training set, if any, is dropped):
 * svmb200_predict_* and svmb200_sv_ids work, svmb200_smo_train needs a new svmb200_set_train_*. */
int svmb200_save_model(svmb200_ctx *ctx, const char *path);
int svmb200_load_model(svmb200_ctx *ctx, const char *path);

/* ---- row-sharded SMO over the GPUs of one box (BASELINE.json configs[3]; the reference scales N only
 * through Cascade partitioning).  One process per GPU, each with its own context and the FULL training
 * set uploaded (X is replicated so that both pivot rows are always local); GPU `rank` owns rows
 * [n*rank/world, n*(rank+1)/world): their f, alpha, cache columns, and the X traffic of each iteration.
 * Inside the persistent kernel every GPU stores one 128-byte record per iteration into every peer's
 * mailbox over NVLink (peer-mapped memory) and reads the others' from its own -- no host, no NCCL on
 * the path.  Set-up: every rank calls svmb200_shard_setup (gets a 64-byte CUDA IPC handle), the handles
 * are all-gathered by the host (torch.distributed, MPI, ...), every rank calls svmb200_shard_connect
 * with the `world` handles in rank order.  Afterwards all ranks call svmb200_smo_train in lockstep with
 * the same parameters; on return b and stats are identical everywhere and alpha holds the rank's own
 * rows (the others are 0 / untouched): all-gather alpha and install it with svmb200_set_model before
 * svmb200_sv_ids / svmb200_predict_*.  Results are bit-identical to the single-GPU run. */
int svmb200_shard_setup(svmb200_ctx *ctx, int32_t rank, int32_t world, void *handle_out /* 64 bytes */);
int svmb200_shard_connect(svmb200_ctx *ctx, int32_t world, const void *handles /* world x 64 bytes */);
int svmb200_set_model(svmb200_ctx *ctx, const double *alpha, double b);

/* ---- Cascade SVM over the GPUs of one box (BASELINE.json configs[4]): replaces the round loops of the reference's MPI
 * programs, the modified two-layer ("star") cascade code/mpi_svm_main2.cpp:439-769 and the classical binary-tree cascade
 * code/mpi_svm_main3.cpp:565-828.  One rank per GPU (a process or a thread each with its own context); the support-vector
 * tuples the reference ships as blocking host messages (mpi_svm_main2.cpp:445-458,578-607,751-757; mpi_svm_main3.cpp:587-601,
 * 689-716) travel GPU to GPU with NCCL: ncclAllGather (star gather), ncclSend / ncclRecv (tree levels), ncclBroadcast (feedback
 * of the global set).  A rank's partition is uploaded once as the uint8 bytes it arrives as and stays on the device; every
 * stage's training set is assembled, packed and solved there (svmb200_smo_train with warm start, mpi_svm_main2.cpp:167-183);
 * only ids, labels and alphas are looked at on the host.  NCCL is bound at run time (libnccl.so.2).
 *   svmb200_cascade_unique_id   on one rank: 128 bytes to hand to every rank (MPI_Bcast, torch.distributed, a pipe ...)
 *   svmb200_cascade_connect     on every rank: joins the communicator (ncclCommInitRank)
 *   svmb200_cascade_set_partition_u8  this rank's rows [n_part x d] with global ids first_id .. first_id + n_part - 1
 *                               (contiguous chunks of ceil(n / P) rows, mpi_svm_main2.cpp:346-402) and labels; computes the
 *                               global min / max over all ranks' rows (the reference: rank 0 over all data, :404-419)
 *   svmb200_cascade_train       all ranks together: rounds until the support-vector id set repeats (:646-660) or max_rounds
 *                               (the reference's 50, :428); afterwards EVERY rank's context holds the final model:
 *                               svmb200_predict_u8 (ge_zero = 1 is the MPI programs' rule, :729), svmb200_cascade_get_model
 *   svmb200_cascade_get_model   global ids and alphas of the final support vectors
 * The tree needs a power-of-two number of ranks (mpi_svm_main3.cpp:420-428).  d <= 832. */
#define SVMB200_CASCADE_STAR 0
#define SVMB200_CASCADE_TREE 1
typedef struct {
    int32_t rounds;          /* rounds run (the reference prints "Cascade finished in N rounds") */
    int32_t converged;       /* 1: the id set repeated */
    int32_t n_sv;
    double  b;
    double  train_ms;        /* wall time of the call */
    int64_t smo_iterations;  /* SMO iterations of this rank's stages */
    int32_t stages;          /* SMO_train calls of this rank */
    double  solve_ms;        /* wall time of this rank's stages: labels up, packing, SMO, extraction of the support vectors */
    double  smo_ms;          /* of which inside the SMO kernels (CUDA events); train_ms - solve_ms = exchange of tuples, merging,
                              * and waiting for other ranks' stages */
} svmb200_cascade_stats;
int svmb200_cascade_unique_id(void *id128);
int svmb200_cascade_connect(svmb200_ctx *ctx, int32_t rank, int32_t world, const void *id128);
int svmb200_cascade_set_partition_u8(svmb200_ctx *ctx, const uint8_t *X_part, const int32_t *y_part, int64_t n_part, int32_t d,
                                     int64_t first_id, int64_t n_total);
int svmb200_cascade_train(svmb200_ctx *ctx, int32_t mode, const svmb200_params *params, int32_t max_rounds,
                          svmb200_cascade_stats *stats);
int svmb200_cascade_get_model(svmb200_ctx *ctx, int32_t *ids, double *alpha, int32_t cap, int32_t *count);

/* ---- the decision function on the tensor cores (tcgen05 / TMEM, TF32x3): svmb200_predict_* on fp32 storage first
 * forms all scores with the tensor cores (error up to ~1e-3: fp32 accumulation of three tf32 products), then
 * re-evaluates every row with |score| < recheck_threshold by the exact kernel, so predicted labels and accuracy are
 * those of the exact path; returned scores are exact for the re-evaluated rows and approximate for the others.
 * Off by default.  A threshold of 0.05 is ample for C = 10. */
int svmb200_set_tensor_cores(svmb200_ctx *ctx, int32_t enable, double recheck_threshold);

/* ---- introspection used by the parity tests ---- */
/* raw tensor-core dot products of the first min(n,128) training rows against themselves (128 x 128 floats, dots may
 * be NULL), and the number of rows the last svmb200_predict_* re-evaluated exactly */
int svmb200_debug_tc_dots(svmb200_ctx *ctx, float *dots, int64_t *rechecked);
/* (i_high, i_low) of the first min(cap, recorded) iterations that reached the update */
int svmb200_get_trace(svmb200_ctx *ctx, int32_t *pairs, int32_t cap, int32_t *count);
/* final error vector f (n doubles) */
int svmb200_get_f(svmb200_ctx *ctx, double *f);
/* one kernel row K(i, .) as the solver computes it: the replacement of main3.cpp:221-222 */
int svmb200_kernel_row(svmb200_ctx *ctx, int64_t i, double gamma, int32_t accum_chunk, double *row);
/* one working-set selection on caller-supplied state: replaces calc_i_high / calc_i_low
 * (main3.cpp:107-142): lowest index wins ties; n when the set is empty */
int svmb200_select_pair(svmb200_ctx *ctx, const double *alpha, const double *f, const int32_t *y,
                        int64_t n, double C, double eps, int64_t *i_high, int64_t *i_low);
/* the scalar part of one iteration exactly as the solver runs it: calculate_U_V (main3.cpp:145-159) and the pair
 * update (:235-270) for a pair with labels y_high / y_low, alphas, b_high = f[i_high], b_low = f[i_low] and
 * K12 = K(i_high, i_low).  out7 = {U, V, eta, alpha_high', alpha_low', delta_alpha_high * y_high, delta_alpha_low * y_low};
 * *stop = 0, SVMB200_STOP_INFEASIBLE (:246) or SVMB200_STOP_ETA (:253), in which case the alphas are returned unchanged */
int svmb200_pair_update(svmb200_ctx *ctx, int32_t y_high, int32_t y_low, double alpha_high, double alpha_low,
                        double b_high, double b_low, double K12, double C, double eps, double *out7, int32_t *stop);
/* the scaled fp32 training matrix as stored in HBM (n x d, padding stripped) */
int svmb200_get_x_scaled(svmb200_ctx *ctx, float *Xs);

/* ---- min/max for callers that cannot allocate (no counterpart in the reference) ----
 * Per-feature minimum and maximum of a uint8 matrix that is ALREADY in device memory, written to
 * device memory, for callers that must not allocate on the device: plugins running inside another
 * framework's stream-ordered or pool-managed region, hosts whose allocator owns the whole GPU, loops
 * that cannot afford the device-wide synchronisation an allocation may cause.  The reduction's
 * scratch is a fixed 1 MiB table inside the library image (held by every GPU that loads the library,
 * until the process exits), so this call performs no cudaMalloc / cudaFree and calls nothing that
 * allocates.  Limits and rules:
 *   - d <= SVMB200_NOALLOC_MAX_FEATURES, otherwise SVMB200_ELIMIT (message on stderr); n is unlimited;
 *   - dev_X [n x d] row-major uint8, dev_min / dev_max [d] doubles: device pointers on `device`;
 *   - the work is enqueued on `stream` (a cudaStream_t, NULL = default stream) and the call returns
 *     without waiting: the results are ready when the stream reaches that point;
 *   - the first call on each device looks up the table's address (one-time set-up): make that first
 *     call before starting any stream capture on that device;
 *   - that one-time set-up is not locked: two host threads making the first call on the same device at the same
 *     time both look the address up and both store the same value, which is harmless; later calls only read it;
 *   - all calls on one device share the table, so two calls whose kernels may overlap on the GPU
 *     (different streams, no ordering between them) race; calls on one stream are safe.  The table
 *     keeps the previous call's values, which is harmless: each call writes every entry it reads.
 * Gives the same values as the min_out / max_out of svmb200_set_train_u8. */
#define SVMB200_NOALLOC_MAX_FEATURES 1024
int svmb200_minmax_u8_noalloc(int device, const uint8_t *dev_X, int64_t n, int32_t d, double *dev_min,
                              double *dev_max, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* SVMB200_H */
