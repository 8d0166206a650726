/*
 * oracle/smo_oracle.c -- CPU restatement of the reference's serial SMO (see smo_oracle.h).
 * TEST INFRASTRUCTURE ONLY: never linked into, loaded by, or called from the product path.
 *
 * Build: gcc -O2 -ffp-contract=off -fPIC -shared (oracle/Makefile).  -ffp-contract=off keeps
 * every FP64 operation separately rounded, as the reference's x86-64 build (g++ -O3, no -mfma,
 * README.md:139) does.
 */
#include "smo_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

void oracle_default_params(oracle_params *p)
{
    p->C = 10.0;          /* code/main3.cpp:163,342 */
    p->gamma = 0.00125;   /* code/main3.cpp:95      */
    p->eps = 1e-12;       /* code/main3.cpp:165     */
    p->tau = 1e-5;        /* code/main3.cpp:196     */
    p->max_iter = 100000; /* code/main3.cpp:198     */
    p->sv_tol = 1e-8;     /* code/main3.cpp:367     */
    p->wss = 1;           /* the reference's only rule, code/main3.cpp:107-142 */
}

/* code/main3.cpp:92-104 */
double oracle_kernel(const double *v1, const double *v2, size_t d, double gamma)
{
    double res = 0.0;
    for (size_t i = 0; i < d; ++i)
        res += (v1[i] - v2[i]) * (v1[i] - v2[i]);
    return exp(-gamma * res);
}

/* code/main3.cpp:57-71 */
void oracle_find_min_max(const double *x, double *min_val, double *max_val, size_t n, size_t d)
{
    for (size_t j = 0; j < d; ++j) {
        double lo = x[j], hi = x[j];
        for (size_t i = 0; i < n; ++i) {
            double v = x[i * d + j];
            lo = v < lo ? v : lo; /* std::min(a,b) = (b<a)?b:a */
            hi = hi < v ? v : hi; /* std::max(a,b) = (a<b)?b:a */
        }
        min_val[j] = lo;
        max_val[j] = hi;
    }
}

/* code/main3.cpp:74-89 */
void oracle_scale_features(double *x, const double *min_val, const double *max_val, size_t n, size_t d)
{
    for (size_t j = 0; j < d; ++j) {
        double range = max_val[j] - min_val[j];
        if (range < 1e-12)
            range = 1.0;
        for (size_t i = 0; i < n; ++i)
            x[i * d + j] = (x[i * d + j] - min_val[j]) / range;
    }
}

/* code/main3.cpp:107-123 */
size_t oracle_calc_i_high(const double *alpha, const int *y, const double *f, size_t n, double C, double eps)
{
    size_t ans = n;
    double best = INFINITY;
    for (size_t i = 0; i < n; ++i) {
        int in_set = (y[i] == 1 && alpha[i] < C - eps) || (y[i] == -1 && alpha[i] > 0.0 + eps);
        if (in_set && f[i] < best) {
            best = f[i];
            ans = i;
        }
    }
    return ans;
}

/* code/main3.cpp:126-142 */
size_t oracle_calc_i_low(const double *alpha, const int *y, const double *f, size_t n, double C, double eps)
{
    size_t ans = n;
    double best = -INFINITY;
    for (size_t i = 0; i < n; ++i) {
        int in_set = (y[i] == 1 && alpha[i] > 0.0 + eps) || (y[i] == -1 && alpha[i] < C - eps);
        if (in_set && f[i] > best) {
            best = f[i];
            ans = i;
        }
    }
    return ans;
}

/* code/main3.cpp:145-159 */
int oracle_calculate_U_V(int s, double a_high, double a_low, double C, double *U, double *V)
{
    if (s == -1) {
        double u = a_low - a_high, v = C + a_low - a_high;
        *U = (0.0 < u) ? u : 0.0; /* std::max(0.0, u) */
        *V = (v < C) ? v : C;     /* std::min(C, v)   */
    } else {
        double u = a_low + a_high - C, v = a_low + a_high;
        *U = (0.0 < u) ? u : 0.0;
        *V = (v < C) ? v : C;
    }
    return *U <= *V + 1e-12;
}

/* ------------------------------------------------------------------------------------------
 * The CUDA path's arithmetic, restated op for op (see DESIGN.md "Numerics").
 * ------------------------------------------------------------------------------------------ */

/* exp(x) from IEEE-exact primitives only: k = rint(x*log2 e); r = x - k ln2 (two-part, fma);
 * degree-13 Taylor polynomial in Horner form with fma; scaled by 2^k in two exact-power steps. */
double oracle_device_exp(double x)
{
    if (x != x) return x;
    if (x > 709.782712893384) return INFINITY;
    if (x < -745.2) return 0.0;
    const double LOG2E = 1.4426950408889634;
    const double LN2_HI = 6.93147180369123816490e-01;
    const double LN2_LO = 1.90821492927058770002e-10;
    double kf = rint(x * LOG2E);
    double r = fma(-kf, LN2_HI, x);
    r = fma(-kf, LN2_LO, r);
    double p = 1.0 / 6227020800.0;            /* 1/13! */
    p = fma(p, r, 1.0 / 479001600.0);         /* 1/12! */
    p = fma(p, r, 1.0 / 39916800.0);          /* 1/11! */
    p = fma(p, r, 1.0 / 3628800.0);           /* 1/10! */
    p = fma(p, r, 1.0 / 362880.0);            /* 1/9!  */
    p = fma(p, r, 1.0 / 40320.0);             /* 1/8!  */
    p = fma(p, r, 1.0 / 5040.0);              /* 1/7!  */
    p = fma(p, r, 1.0 / 720.0);               /* 1/6!  */
    p = fma(p, r, 1.0 / 120.0);               /* 1/5!  */
    p = fma(p, r, 1.0 / 24.0);                /* 1/4!  */
    p = fma(p, r, 1.0 / 6.0);                 /* 1/3!  */
    p = fma(p, r, 0.5);                       /* 1/2!  */
    p = fma(p, r, 1.0);
    p = fma(p, r, 1.0);
    int k = (int)kf;
    int k1 = k / 2, k2 = k - k1;
    union { uint64_t u; double f; } s1, s2;
    s1.u = (uint64_t)(k1 + 1023) << 52;
    s2.u = (uint64_t)(k2 + 1023) << 52;
    return (p * s1.f) * s2.f;
}

static int g_device_chunk = 4; /* fp32 squares summed before widening to FP64: 1, 2 or 4 */
void oracle_set_device_chunk(int ch) { g_device_chunk = (ch == 1 || ch == 2) ? ch : 4; }
int oracle_get_device_chunk(void) { return g_device_chunk; }

/* dot product of two fp32 rows as the CUDA path forms it: lane l = q mod 32 owns the float4 groups
 * q = l, l+32, ...; the products of `chunk` consecutive elements are summed in fp32 (fmaf), widened
 * and added to the lane's FP64 sum; the 32 lane sums are combined by an xor-butterfly 16,8,4,2,1. */
double oracle_device_dot(const float *a, const float *b, size_t d)
{
    double lane[32];
    for (int l = 0; l < 32; ++l) lane[l] = 0.0;
    size_t d4 = (d + 3) / 4;
    for (size_t q = 0; q < d4; ++q) {
        float av[4], bv[4];
        for (int c = 0; c < 4; ++c) {
            size_t k = 4 * q + c;
            av[c] = k < d ? a[k] : 0.0f; /* rows are zero-padded */
            bv[c] = k < d ? b[k] : 0.0f;
        }
        double acc = lane[q & 31];
        if (g_device_chunk == 4) {
            float p = av[0] * bv[0];
            p = fmaf(av[1], bv[1], p);
            p = fmaf(av[2], bv[2], p);
            p = fmaf(av[3], bv[3], p);
            acc = acc + (double)p;
        } else if (g_device_chunk == 2) {
            float p = av[0] * bv[0];
            p = fmaf(av[1], bv[1], p);
            acc = acc + (double)p;
            float t = av[2] * bv[2];
            t = fmaf(av[3], bv[3], t);
            acc = acc + (double)t;
        } else {
            for (int c = 0; c < 4; ++c) acc = fma((double)av[c], (double)bv[c], acc);
        }
        lane[q & 31] = acc;
    }
    for (int off = 16; off >= 1; off >>= 1) {
        double t[32];
        for (int l = 0; l < 32; ++l) t[l] = lane[l] + lane[l ^ off];
        for (int l = 0; l < 32; ++l) lane[l] = t[l];
    }
    return lane[0];
}

/* squared distance from the squared norms: (|a|^2 + |b|^2) - 2 a.b, clamped at 0 */
double oracle_device_sqdist_n(double na, double nb, double ab)
{
    double s = (na + nb) - 2.0 * ab;
    return s < 0.0 ? 0.0 : s;
}

double oracle_device_sqdist(const float *a, const float *b, size_t d)
{
    return oracle_device_sqdist_n(oracle_device_dot(a, a, d), oracle_device_dot(b, b, d), oracle_device_dot(a, b, d));
}

double oracle_device_kernel(const float *a, const float *b, size_t d, double gamma)
{
    return oracle_device_exp(-gamma * oracle_device_sqdist(a, b, d));
}

/* the exact path (svmb200_set_exact): FP64 rows, squared differences as in code/main3.cpp:97-100 but summed
 * per lane (lane l = g mod 32 over the 16-byte groups g of two values) and combined by the same xor-butterfly */
double oracle_device_kernel_f64(const double *a, const double *b, size_t d, double gamma)
{
    double lane[32];
    for (int l = 0; l < 32; ++l) lane[l] = 0.0;
    size_t groups = (d + 1) / 2;
    for (size_t g = 0; g < groups; ++g) {
        double acc = lane[g & 31];
        for (int c = 0; c < 2; ++c) {
            size_t k = 2 * g + c;
            double av = k < d ? a[k] : 0.0, bv = k < d ? b[k] : 0.0;
            double df = bv - av;
            acc = fma(df, df, acc);
        }
        lane[g & 31] = acc;
    }
    for (int off = 16; off >= 1; off >>= 1) {
        double t[32];
        for (int l = 0; l < 32; ++l) t[l] = lane[l] + lane[l ^ off];
        for (int l = 0; l < 32; ++l) lane[l] = t[l];
    }
    return oracle_device_exp(-gamma * lane[0]);
}


/* ------------------------------------------------------------------------------------------
 * The CUDA path for uint8 input kept as uint8 in memory (svmb200 storage "u8", ORACLE_ARITH_DEVICE_U8).
 * The reference scales every feature to (u - min_k) / range_k in FP64 (code/main3.cpp:74-89) and sums
 * squared differences (:97-100).  For integer-valued input min_k and range_k are integers and
 * (a_k - b_k)^2 = (u_k - p_k)^2 / range_k^2, so the device keeps the raw bytes and forms
 *     d2 = S_main * w_R + tail
 *   S_main = sum over the features whose range is the most frequent one, R, of (u_k - p_k)^2 : exact integer
 *   tail   = (N(u) + N(p)) - 2 D(u, p) over the other non-constant features, clamped at 0, with
 *            D(u, p) = sum_k u_k * (w_k p_k), N(u) = D(u, u), w_k = (1/range_k)^2, in FP64 as (D_0 + D_1) + (D_2 + D_3)
 *            where D_c adds the features of the 4-byte words w = c (mod 4) of the tail in ascending order with fma (the device walks only the
 *            positions where a pivot byte is non-zero: a zero byte adds +0).  D's first argument is the streamed row,
 *            the second the pivot row.
 * Constant features (range 0) contribute (u_k - p_k) = 0 and are left out.  K = exp(-gamma d2) (oracle_device_exp).
 * ------------------------------------------------------------------------------------------ */
static struct {
    size_t n, d;
    int R;                 /* range of the main class (0 = no main class) */
    size_t M, T;           /* features in the main class / in the tail */
    int *main_idx, *tail_idx;
    double wR, *wt;        /* weight of the main class, weights of the tail features */
    uint8_t *mainb, *tailb;   /* [n][M], [n][T] */
    int32_t *nmain;        /* [n] integer squared norms over the main class */
    double *ntail;         /* [n] N(u) over the tail */
} g_u8;

static double u8_tail_dot(const uint8_t *row, const uint8_t *piv)
{
    double acc[4] = {0.0, 0.0, 0.0, 0.0};
    for (size_t k = 0; k < g_u8.T; ++k) {
        double pw = g_u8.wt[k] * (double)piv[k];
        acc[(k >> 2) & 3] = fma((double)row[k], pw, acc[(k >> 2) & 3]);
    }
    return (acc[0] + acc[1]) + (acc[2] + acc[3]);
}

void oracle_u8_unbind(void)
{
    free(g_u8.main_idx); free(g_u8.tail_idx); free(g_u8.wt); free(g_u8.mainb); free(g_u8.tailb);
    free(g_u8.nmain); free(g_u8.ntail);
    memset(&g_u8, 0, sizeof g_u8);
}

void oracle_u8_bind(const uint8_t *rows, size_t n, size_t d)
{
    oracle_u8_unbind();
    int *range = (int *)calloc(d ? d : 1, sizeof(int));
    int count[256];
    memset(count, 0, sizeof count);
    for (size_t k = 0; k < d; ++k) {
        int lo = 255, hi = 0;
        for (size_t i = 0; i < n; ++i) {
            int v = rows[i * d + k];
            lo = v < lo ? v : lo;
            hi = v > hi ? v : hi;
        }
        range[k] = n ? hi - lo : 0;
        count[range[k]]++;
    }
    int R = 0;                                  /* most frequent non-zero range; the larger value wins a tie */
    for (int r = 1; r < 256; ++r)
        if (count[r] > 0 && (R == 0 || count[r] >= count[R])) R = r;
    if (R && count[R] < 32) R = 0;              /* too few to be worth an integer section */
    g_u8.n = n; g_u8.d = d; g_u8.R = R;
    g_u8.main_idx = (int *)malloc(sizeof(int) * (d ? d : 1));
    g_u8.tail_idx = (int *)malloc(sizeof(int) * (d ? d : 1));
    g_u8.wt = (double *)malloc(sizeof(double) * (d ? d : 1));
    for (size_t k = 0; k < d; ++k) {
        if (range[k] == 0) continue;
        if (range[k] == R) g_u8.main_idx[g_u8.M++] = (int)k;
        else {
            double s = 1.0 / (double)range[k];
            g_u8.wt[g_u8.T] = s * s;
            g_u8.tail_idx[g_u8.T++] = (int)k;
        }
    }
    { double s = R ? 1.0 / (double)R : 0.0; g_u8.wR = s * s; }
    g_u8.mainb = (uint8_t *)malloc(n * g_u8.M + 1);
    g_u8.tailb = (uint8_t *)malloc(n * g_u8.T + 1);
    g_u8.nmain = (int32_t *)malloc(sizeof(int32_t) * (n ? n : 1));
    g_u8.ntail = (double *)malloc(sizeof(double) * (n ? n : 1));
    for (size_t i = 0; i < n; ++i) {
        int32_t nm = 0;
        for (size_t k = 0; k < g_u8.M; ++k) {
            uint8_t v = rows[i * d + g_u8.main_idx[k]];
            g_u8.mainb[i * g_u8.M + k] = v;
            nm += (int32_t)v * v;
        }
        for (size_t k = 0; k < g_u8.T; ++k) g_u8.tailb[i * g_u8.T + k] = rows[i * d + g_u8.tail_idx[k]];
        g_u8.nmain[i] = nm;
        g_u8.ntail[i] = u8_tail_dot(g_u8.tailb + i * g_u8.T, g_u8.tailb + i * g_u8.T);
    }
    free(range);
}

void oracle_u8_layout(int *R, int64_t *M, int64_t *T)
{
    if (R) *R = g_u8.R;
    if (M) *M = (int64_t)g_u8.M;
    if (T) *T = (int64_t)g_u8.T;
}

/* squared distance of the scaled rows: streamed row `row`, pivot row `piv` (indices into the bound rows) */
double oracle_u8_sqdist(size_t row, size_t piv)
{
    int32_t dotm = 0;
    const uint8_t *a = g_u8.mainb + row * g_u8.M, *b = g_u8.mainb + piv * g_u8.M;
    for (size_t k = 0; k < g_u8.M; ++k) dotm += (int32_t)a[k] * b[k];
    int32_t S = g_u8.nmain[row] + g_u8.nmain[piv] - 2 * dotm;
    double dt = u8_tail_dot(g_u8.tailb + row * g_u8.T, g_u8.tailb + piv * g_u8.T);
    double tail = (g_u8.ntail[row] + g_u8.ntail[piv]) - 2.0 * dt;
    if (tail < 0.0) tail = 0.0;
    return fma((double)S, g_u8.wR, tail);
}

double oracle_u8_kernel(size_t row, size_t piv, double gamma)
{
    return oracle_device_exp(-gamma * oracle_u8_sqdist(row, piv));
}

static double device_kernel_n(const float *a, const float *b, double na, double nb, size_t d, double gamma)
{
    return oracle_device_exp(-gamma * oracle_device_sqdist_n(na, nb, oracle_device_dot(a, b, d)));
}

/* one kernel row K(i, .) in the chosen arithmetic */
static void kernel_row(const double *x, const float *x32, const double *sqn, size_t n, size_t d, size_t i,
                       double gamma, int arith, double *row)
{
    if (arith == ORACLE_ARITH_F64) {
        for (size_t j = 0; j < n; ++j)
            row[j] = oracle_kernel(x + i * d, x + j * d, d, gamma);
    } else if (arith == ORACLE_ARITH_DEVICE_F64) {
        for (size_t j = 0; j < n; ++j)
            row[j] = oracle_device_kernel_f64(x + i * d, x + j * d, d, gamma);
    } else if (arith == ORACLE_ARITH_DEVICE_U8) {
        for (size_t j = 0; j < n; ++j)
            row[j] = oracle_u8_kernel(j, i, gamma);
    } else {
        for (size_t j = 0; j < n; ++j)
            row[j] = device_kernel_n(x32 + i * d, x32 + j * d, sqn[i], sqn[j], d, gamma);
    }
}

/* code/main3.cpp:162-294 ; warm start code/mpi_svm_main2.cpp:161-184 */
int oracle_smo_train(const double *x, const int *y, size_t n, size_t d,
                     double *alpha, double *b, const oracle_params *p, int init,
                     int arith, oracle_result *res,
                     int32_t *trace, size_t trace_cap, double *f_out)
{
    const double eps = p->eps, C = p->C, tau = p->tau, gamma = p->gamma;
    double *f = (double *)malloc(sizeof(double) * (n ? n : 1));
    double *K_h = (double *)calloc(n ? n : 1, sizeof(double));
    double *K_l = (double *)calloc(n ? n : 1, sizeof(double));
    float *x32 = NULL;
    double *sqn = NULL;
    if (arith == ORACLE_ARITH_DEVICE) {
        x32 = (float *)malloc(sizeof(float) * (n * d ? n * d : 1));
        sqn = (double *)malloc(sizeof(double) * (n ? n : 1));
        for (size_t k = 0; k < n * d; ++k) x32[k] = (float)x[k];
        for (size_t i = 0; i < n; ++i) sqn[i] = oracle_device_dot(x32 + i * d, x32 + i * d, d);
    }

    if (init) {
        for (size_t i = 0; i < n; ++i) {
            alpha[i] = 0.0;                 /* :167 */
            f[i] = -(double)y[i];           /* :171-172 */
        }
    } else {
        /* code/mpi_svm_main2.cpp:169-183: f_i = sum_{alpha_j != 0} alpha_j y_j K(x_j, x_i) - y_i */
        for (size_t i = 0; i < n; ++i) {
            double sum = 0.0;
            for (size_t j = 0; j < n; ++j) {
                if (alpha[j] == 0.0) continue;
                double kij = (arith == ORACLE_ARITH_F64)          ? oracle_kernel(x + j * d, x + i * d, d, gamma)
                             : (arith == ORACLE_ARITH_DEVICE_F64 || arith == ORACLE_ARITH_DEVICE_U8) ? oracle_device_kernel_f64(x + j * d, x + i * d, d, gamma)
                                                                  : device_kernel_n(x32 + j * d, x32 + i * d, sqn[j], sqn[i], d, gamma);
                sum += alpha[j] * y[j] * kij;
            }
            f[i] = sum - (double)y[i];
        }
    }

    size_t i_high = n, i_low = n;
    double b_high = 0.0, b_low = 0.0; /* the reference leaves these uninitialised (:188-189) */
    size_t i_high_prev = n, i_low_prev = n;
    int num_iter = 1;
    int stop = ORACLE_STOP_CONVERGED;
    int64_t rows = 0;
    size_t ntrace = 0;

    for (;;) {
        i_high = oracle_calc_i_high(alpha, y, f, n, C, eps); /* :203 */
        i_low = oracle_calc_i_low(alpha, y, f, n, C, eps);   /* :204 */
        if (i_high >= n || i_low >= n) { stop = ORACLE_STOP_NOT_FOUND; break; } /* :205 */
        b_high = f[i_high];
        b_low = f[i_low];
        if (b_low <= b_high + 2.0 * tau) { stop = ORACLE_STOP_CONVERGED; break; } /* :213 */

        if (i_high != i_high_prev) { /* :216-223 */
            i_high_prev = i_high;
            kernel_row(x, x32, sqn, n, d, i_high, gamma, arith, K_h);
            ++rows;
        }
        if (p->wss == 2) {
            /* second-order choice of i_low (not in the reference; see smo_oracle.h) */
            size_t best = n;
            double best_gain = -INFINITY;
            for (size_t t = 0; t < n; ++t) {
                int in_low = (y[t] == 1 && alpha[t] > 0.0 + eps) || (y[t] == -1 && alpha[t] < C - eps);
                if (!in_low || !(f[t] > b_high)) continue;
                double eta_t = 1.0 + 1.0 - 2.0 * K_h[t];
                if (eta_t <= eps) continue;
                double diff = f[t] - b_high;
                double gain = diff * diff / eta_t;
                if (gain > best_gain) { best_gain = gain; best = t; }
            }
            if (best < n) { i_low = best; b_low = f[i_low]; }
        }
        if (i_low != i_low_prev) { /* :225-232 */
            i_low_prev = i_low;
            kernel_row(x, x32, sqn, n, d, i_low, gamma, arith, K_l);
            ++rows;
        }
        if (trace && ntrace < trace_cap) {
            trace[2 * ntrace] = (int32_t)i_high;
            trace[2 * ntrace + 1] = (int32_t)i_low;
        }
        ++ntrace;

        int s = y[i_high] * y[i_low];                    /* :235 */
        double K11 = K_h[i_high], K22 = K_l[i_low], K12 = K_h[i_low]; /* :239-241 */
        double eta = K11 + K22 - 2.0 * K12;              /* :243 */
        double U, V;
        if (!oracle_calculate_U_V(s, alpha[i_high], alpha[i_low], C, &U, &V)) { /* :246 */
            stop = ORACLE_STOP_INFEASIBLE; break;
        }
        if (eta <= eps) { stop = ORACLE_STOP_ETA; break; } /* :253 */

        double na_low = alpha[i_low] + y[i_low] * (b_high - b_low) / eta; /* :259 */
        if (na_low > V) na_low = V;
        if (na_low < U) na_low = U;
        double na_high = alpha[i_high] + s * (alpha[i_low] - na_low);      /* :266 */

        double d_high = na_high - alpha[i_high];
        double d_low = na_low - alpha[i_low];
        for (size_t i = 0; i < n; ++i) /* :271-275 */
            f[i] += d_high * y[i_high] * K_h[i] + d_low * y[i_low] * K_l[i];

        alpha[i_high] = na_high; /* :278-279 */
        alpha[i_low] = na_low;

        ++num_iter;              /* :281-287 */
        if (num_iter > p->max_iter) { stop = ORACLE_STOP_MAX_ITER; break; }
    }

    *b = (b_high + b_low) / 2; /* :291 */
    if (res) {
        res->num_iter = num_iter;
        res->stop_reason = stop;
        res->i_high = (int64_t)i_high;
        res->i_low = (int64_t)i_low;
        res->b_high = b_high;
        res->b_low = b_low;
        res->rows_computed = rows;
    }
    if (f_out) memcpy(f_out, f, sizeof(double) * n);
    free(f); free(K_h); free(K_l); free(x32); free(sqn);
    return 0;
}

/* code/main3.cpp:297-304 */
size_t oracle_sv_indices(const double *alpha, size_t n, double tol, int32_t *idx)
{
    size_t c = 0;
    for (size_t i = 0; i < n; ++i)
        if (alpha[i] > tol) {
            if (idx) idx[c] = (int32_t)i;
            ++c;
        }
    return c;
}

/* code/main3.cpp:391-405 ; code/mpi_svm_main2.cpp:721-733 */
int oracle_predict(const double *x_test, const int *y_test, size_t m,
                   const double *x_train, const int *y_train, const double *alpha,
                   const int32_t *sv_idx, size_t n_sv, size_t d, double gamma, double b,
                   int ge_zero, int arith, double *scores)
{
    int correct = 0;
    float *t32 = NULL, *s32 = NULL;
    if (arith == ORACLE_ARITH_DEVICE) {
        t32 = (float *)malloc(sizeof(float) * (m * d ? m * d : 1));
        s32 = (float *)malloc(sizeof(float) * (n_sv * d ? n_sv * d : 1));
        for (size_t k = 0; k < m * d; ++k) t32[k] = (float)x_test[k];
        for (size_t k = 0; k < n_sv; ++k)
            for (size_t c = 0; c < d; ++c) s32[k * d + c] = (float)x_train[(size_t)sv_idx[k] * d + c];
    }
    for (size_t i = 0; i < m; ++i) {
        double curr = -b;
        for (size_t k = 0; k < n_sv; ++k) {
            size_t j = (size_t)sv_idx[k];
            double kv = (arith == ORACLE_ARITH_F64)          ? oracle_kernel(x_test + i * d, x_train + j * d, d, gamma)
                        : (arith == ORACLE_ARITH_DEVICE_F64 || arith == ORACLE_ARITH_DEVICE_U8) ? oracle_device_kernel_f64(x_test + i * d, x_train + j * d, d, gamma)
                                                             : oracle_device_kernel(t32 + i * d, s32 + k * d, d, gamma);
            curr += alpha[j] * y_train[j] * kv;
        }
        if (scores) scores[i] = curr;
        int pred = ge_zero ? (curr >= 0 ? 1 : -1) : (curr > 0 ? 1 : -1);
        if (pred == y_test[i]) ++correct;
    }
    free(t32); free(s32);
    return correct;
}
