/*
 * oracle/ref_wrapper.cpp -- builds the UNMODIFIED reference serial trainer as a shared library.
 * TEST INFRASTRUCTURE ONLY.  No reference source is copied: the reference file is #included from
 * where it lies (REF_MAIN3 is set by oracle/Makefile to /root/reference/code/main3.cpp) and its
 * main() is renamed so that its free functions (read_CSV, find_min_max, scale_features, kernel,
 * calc_i_high, calc_i_low, calculate_U_V, SMO_train, get_SV_indices) can be called directly.
 * <iomanip> is force-included because code/main3.cpp:292 uses std::setprecision without it.
 * The output goes to oracle/_ref/ (git-ignored, travels to the GPU box with the snapshot).
 */
#include <iomanip>
#include <cstdint>
#include <cstring>
#define main ref_main3_main
#include REF_MAIN3
#undef main

namespace {
struct CoutSilencer { // SMO_train prints three lines to cout (code/main3.cpp:290-293); capture them
    std::streambuf *old;
    std::ostringstream cap;
    CoutSilencer() : old(std::cout.rdbuf(cap.rdbuf())) {}
    ~CoutSilencer() { std::cout.rdbuf(old); }
};
}

extern "C" {

double ref_kernel(const double *a, const double *b, size_t d) { return kernel(a, b, d); }

void ref_find_min_max(const double *x, double *mn, double *mx, size_t n, size_t d)
{
    vector<double> X(x, x + n * d), lo(d, 0.0), hi(d, 0.0);
    find_min_max(X, lo, hi, n, d);
    memcpy(mn, lo.data(), d * sizeof(double));
    memcpy(mx, hi.data(), d * sizeof(double));
}

void ref_scale_features(double *x, const double *mn, const double *mx, size_t n, size_t d)
{
    vector<double> X(x, x + n * d), lo(mn, mn + d), hi(mx, mx + d);
    scale_features(X, lo, hi, n, d);
    memcpy(x, X.data(), n * d * sizeof(double));
}

/* the reference's SMO_train (code/main3.cpp:162) run to its own stopping rule.
 * num_iter_out = the "number of iterations" line it prints (:290). */
void ref_smo_train(const double *x, const int *y, size_t n, size_t d, double *alpha, double *b,
                   double C, int *num_iter_out)
{
    vector<double> X(x, x + n * d), A;
    vector<int> Y(y, y + n);
    double bb = 0.0;
    CoutSilencer quiet;
    SMO_train(X, Y, d, A, bb, C);
    memcpy(alpha, A.data(), n * sizeof(double));
    *b = bb;
    if (num_iter_out) {
        std::string s = quiet.cap.str();
        size_t pos = s.find("number of iterations: ");
        *num_iter_out = pos == std::string::npos ? -1 : atoi(s.c_str() + pos + 22);
    }
}

size_t ref_sv_indices(const double *alpha, size_t n, double tol, int32_t *idx)
{
    vector<double> A(alpha, alpha + n);
    vector<int> v = get_SV_indices(A, tol);
    for (size_t i = 0; i < v.size(); ++i) idx[i] = v[i];
    return v.size();
}

/* the reference's prediction loop (code/main3.cpp:391-402) over its own kernel() */
int ref_predict(const double *x_test, const int *y_test, size_t m, const double *x_train,
                const int *y_train, const double *alpha, size_t n, size_t d, double b, double tol)
{
    vector<double> A(alpha, alpha + n);
    vector<int> sv_idx = get_SV_indices(A, tol);
    int correct = 0;
    for (size_t i = 0; i < m; ++i) {
        double curr = -b;
        for (size_t k = 0; k < sv_idx.size(); ++k) {
            size_t j = sv_idx[k];
            curr += alpha[j] * y_train[j] * kernel(x_test + i * d, x_train + j * d, d);
        }
        if ((curr > 0 ? 1 : -1) == y_test[i]) correct++;
    }
    return correct;
}

/* Timed CPU arm for bench.py --impl reference: exactly `iters` SMO iterations driven through the
 * reference's own calc_i_high / calc_i_low / kernel / calculate_U_V (its SMO_train cannot be
 * stopped early: max_iter is a local literal, code/main3.cpp:198).  The loop body follows
 * code/main3.cpp:200-288.  Returns the number of iterations done (fewer if it converged). */
int ref_smo_iterations(const double *x, const int *y, size_t n, size_t d, double C, int iters,
                       double *alpha_out, int64_t *rows_out)
{
    vector<double> alpha(n, 0.0), f(n), K_h(n, 0.0), K_l(n, 0.0);
    vector<int> Y(y, y + n);
    for (size_t i = 0; i < n; ++i) f[i] = -static_cast<double>(y[i]);
    size_t hp = n, lp = n;
    int done = 0;
    int64_t rows = 0;
    while (done < iters) {
        size_t ih = calc_i_high(alpha, Y, f, C), il = calc_i_low(alpha, Y, f, C);
        if (ih >= n || il >= n) break;
        double bh = f[ih], bl = f[il];
        if (bl <= bh + 2.0 * 1e-5) break;
        if (ih != hp) { hp = ih; ++rows; for (size_t j = 0; j < n; ++j) K_h[j] = kernel(x + ih * d, x + j * d, d); }
        if (il != lp) { lp = il; ++rows; for (size_t j = 0; j < n; ++j) K_l[j] = kernel(x + il * d, x + j * d, d); }
        int s = Y[ih] * Y[il];
        double eta = K_h[ih] + K_l[il] - 2.0 * K_h[il], U, V;
        if (!calculate_U_V(s, alpha[ih], alpha[il], C, U, V) || eta <= 1e-12) break;
        double nl = alpha[il] + Y[il] * (bh - bl) / eta;
        if (nl > V) nl = V;
        if (nl < U) nl = U;
        double nh = alpha[ih] + s * (alpha[il] - nl);
        double dh = nh - alpha[ih], dl = nl - alpha[il];
        for (size_t i = 0; i < n; ++i) f[i] += dh * Y[ih] * K_h[i] + dl * Y[il] * K_l[i];
        alpha[ih] = nh;
        alpha[il] = nl;
        ++done;
    }
    if (alpha_out) memcpy(alpha_out, alpha.data(), n * sizeof(double));
    if (rows_out) *rows_out = rows;
    return done;
}

} // extern "C"
