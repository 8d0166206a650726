/*
 * oracle/smo_oracle.h -- CPU restatement of the reference's RBF-kernel SMO path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product: only
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load this library, and there only as the checker or the timed CPU arm.
 *
 * Every function cites the reference file:line it follows (paths relative to the
 * upstream repository root, e.g. code/main3.cpp:92).  Parity of this restatement
 * is pinned by tests/test_oracle.py against oracle/_ref (the unmodified
 * code/main3.cpp compiled through oracle/ref_wrapper.cpp) and against the golden
 * fixtures under tests/golden/ that were generated from that build.
 */
#ifndef SMO_ORACLE_H
#define SMO_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Literals of code/main3.cpp turned into parameters (defaults = those literals):
 * C=10 (:163,:342) gamma=0.00125 (:95) eps=1e-12 (:109,:128,:165) tau=1e-5 (:196)
 * max_iter=100000 (:198) sv_tol=1e-8 (:297,:367). */
typedef struct {
    double C;
    double gamma;
    double eps;
    double tau;
    int    max_iter;
    double sv_tol;
    int    wss;        /* 1: the reference's first-order maximal violating pair (code/main3.cpp:107-142).
                          2: second-order choice of i_low (Fan, Chen, Lin 2005, "working set selection using second order
                          information", the rule LIBSVM uses); NOT in the reference -- see oracle_smo_train */
} oracle_params;

/* stop_reason mirrors the reference's loop exits (code/main3.cpp:205,213,246,253,283) */
enum {
    ORACLE_STOP_CONVERGED  = 0, /* b_low <= b_high + 2 tau            :213 */
    ORACLE_STOP_NOT_FOUND  = 1, /* i_high or i_low not found          :205 */
    ORACLE_STOP_INFEASIBLE = 2, /* infeasible U and V                 :246 */
    ORACLE_STOP_ETA        = 3, /* non positive eta                   :253 */
    ORACLE_STOP_MAX_ITER   = 4  /* num_iter > max_iter                :283 */
};

/* how kernel rows are evaluated */
enum {
    ORACLE_ARITH_F64      = 0, /* the reference's arithmetic: sequential FP64 sum, libm exp */
    ORACLE_ARITH_DEVICE   = 1, /* the CUDA path's arithmetic restated op for op (fp32 X,
                                  squared norms + fp32-chunk dot products, FP64 lane sums,
                                  fixed butterfly order, svm_exp): used to check the GPU bit for bit */
    ORACLE_ARITH_DEVICE_F64 = 2, /* the CUDA exact path (svmb200_set_exact: FP64 rows, squared
                                  differences, per-lane FP64 sums, same butterfly, svm_exp) */
    ORACLE_ARITH_DEVICE_U8 = 3   /* the CUDA path for uint8 input kept as uint8 (exact integer sums over the
                                  features of the most frequent range, weighted FP64 sums over the rest; see
                                  smo_oracle.c).  Needs oracle_u8_bind() on the raw training rows first; the
                                  warm-start rebuild of f and the decision function use ORACLE_ARITH_DEVICE_F64,
                                  as the device does */
};

typedef struct {
    int      num_iter;     /* the reference's printed counter: starts at 1, +1 per update (:197,:281) */
    int      stop_reason;
    int64_t  i_high, i_low;   /* last selected pair */
    double   b_high, b_low;
    int64_t  rows_computed;   /* kernel rows evaluated (1-row reuse counted, :216-232) */
} oracle_result;

void   oracle_default_params(oracle_params *p);

/* code/main3.cpp:92-104 with gamma as a parameter */
double oracle_kernel(const double *v1, const double *v2, size_t d, double gamma);
/* code/main3.cpp:57-71 and :74-89 */
void   oracle_find_min_max(const double *x, double *min_val, double *max_val, size_t n, size_t d);
void   oracle_scale_features(double *x, const double *min_val, const double *max_val, size_t n, size_t d);
/* code/main3.cpp:107-123, :126-142 -- return n when the set is empty */
size_t oracle_calc_i_high(const double *alpha, const int *y, const double *f, size_t n, double C, double eps);
size_t oracle_calc_i_low(const double *alpha, const int *y, const double *f, size_t n, double C, double eps);
/* code/main3.cpp:145-159 */
int    oracle_calculate_U_V(int s, double a_high, double a_low, double C, double *U, double *V);

/* code/main3.cpp:162-294 (init != 0) and code/mpi_svm_main2.cpp:154-288 (init == 0: alpha is
 * in/out and f is rebuilt from it, :167-183).  trace (optional) receives (i_high, i_low) of every
 * iteration that reached the update, trace_cap pairs at most.  f_out (optional) receives the final
 * error vector.  arith selects ORACLE_ARITH_*.  Returns 0.
 *
 * p->wss == 2 (no counterpart in the reference: PARITY UNPINNED, checked only through the solution it reaches):
 * i_high, the stop test and every other line stay as above; i_low becomes the row of I_low with f_t > b_high that
 * maximises (f_t - b_high)^2 / eta_t, eta_t = K11 + K_tt - 2 K(i_high, t) = 2 - 2 K(i_high, t) (the gain of the
 * two-variable sub-problem, Fan et al. 2005 eq. WSS 2 with a_it = eta_t, b_it = f_t - b_high), lowest index among equal
 * gains; rows with eta_t <= eps are passed over (the reference's own step would stop on them, :253); if no row qualifies
 * the first-order i_low stands. */
int    oracle_smo_train(const double *x, const int *y, size_t n, size_t d,
                        double *alpha, double *b, const oracle_params *p, int init,
                        int arith, oracle_result *res,
                        int32_t *trace, size_t trace_cap, double *f_out);

/* code/main3.cpp:297-304: ascending indices with alpha > tol; returns the count */
size_t oracle_sv_indices(const double *alpha, size_t n, double tol, int32_t *idx);

/* code/main3.cpp:391-405 (ge_zero == 0: curr > 0) and code/mpi_svm_main2.cpp:721-733
 * (ge_zero != 0: s >= 0).  SV rows given by sv_idx into x_train.  scores (optional) gets the
 * decision values.  Returns the number of correct predictions. */
int    oracle_predict(const double *x_test, const int *y_test, size_t m,
                      const double *x_train, const int *y_train, const double *alpha,
                      const int32_t *sv_idx, size_t n_sv, size_t d, double gamma, double b,
                      int ge_zero, int arith, double *scores);

/* ---- the CUDA path's arithmetic, restated (ORACLE_ARITH_DEVICE) ---- */
/* exp() restated with IEEE +,*,fma only, so host and device agree bit for bit */
double oracle_device_exp(double x);
/* fp32 squares summed before widening to FP64 (svmb200_params.accum_chunk): 1, 2 or 4 (default, as the library) */
void   oracle_set_device_chunk(int ch);
int    oracle_get_device_chunk(void);
/* dot product of two fp32 rows: products of `chunk` consecutive elements summed in fp32 (fmaf),
 * widened and added to the FP64 sum of lane l = q mod 32 (q = float4 group), lane sums combined by
 * an xor-butterfly 16,8,4,2,1 */
double oracle_device_dot(const float *a, const float *b, size_t d);
/* squared distance max(0, (|a|^2 + |b|^2) - 2 a.b) from those dot products */
double oracle_device_sqdist_n(double na, double nb, double ab);
double oracle_device_sqdist(const float *a, const float *b, size_t d);
double oracle_device_kernel(const float *a, const float *b, size_t d, double gamma);
double oracle_device_kernel_f64(const double *a, const double *b, size_t d, double gamma);
/* ORACLE_ARITH_DEVICE_U8: bind the raw uint8 training rows [n x d] (min / range are taken from them), then
 * oracle_smo_train(..., ORACLE_ARITH_DEVICE_U8, ...) on the same rows scaled.  row / piv index the bound rows. */
void   oracle_u8_bind(const uint8_t *rows, size_t n, size_t d);
void   oracle_u8_unbind(void);
void   oracle_u8_layout(int *R, int64_t *M, int64_t *T);
double oracle_u8_sqdist(size_t row, size_t piv);
double oracle_u8_kernel(size_t row, size_t piv, double gamma);

#ifdef __cplusplus
}
#endif
#endif
