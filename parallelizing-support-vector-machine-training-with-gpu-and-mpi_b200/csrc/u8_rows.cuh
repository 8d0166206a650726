// u8_rows.cuh -- kernels around the uint8 storage of X (svmb200_set_storage "u8"): packing the raw rows into the
// [main class | tail] layout the SMO kernel walks, the per-row norms, scaled FP64 copies of chosen rows for the
// decision function, and one kernel row for the parity tests.  The arithmetic is described in smo_kernel.cuh.
//
// Replaces, for integer-valued input, scale_features + the FP64 copy of X the reference keeps (code/main3.cpp:74-89,
// :317): the statistics are still find_min_max's (:57-71), but X stays as the bytes it arrived as.
#pragma once
#include "smo_kernel.cuh"

namespace svm {

// rows of `src` ([n][d] raw bytes) -> [n][row_bytes]: byte b of a packed row is feature src_col[b] (or 0 where src_col[b] < 0)
__global__ void u8_pack_kernel(const unsigned char *__restrict__ src, long long n, int d, int row_bytes, const int *__restrict__ src_col,
                               unsigned char *__restrict__ out)
{
    const long long total = n * (row_bytes / 4);
    for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < total; g += (long long)gridDim.x * blockDim.x) {
        const long long r = g / (row_bytes / 4);
        const int w = (int)(g - r * (row_bytes / 4));
        unsigned v = 0;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const int col = src_col[4 * w + c];
            if (col >= 0) v |= (unsigned)src[r * d + col] << (8 * c);
        }
        reinterpret_cast<unsigned *>(out)[g] = v;
    }
}

// K-independent pieces of the uint8 arithmetic for one (row, pivot) pair, by ONE thread straight from global memory:
// the integer dot product over the main class and D(row, pivot) over the tail in the solver's order -- (D_0 + D_1) +
// (D_2 + D_3), D_c over the words w = c (mod 4) ascending (a zero byte of either row adds +0, so the dense walk here and
// the solver's walk over the pivots' non-zero positions give the same bits).
__device__ __forceinline__ void u8_pair_sums(const unsigned char *__restrict__ X, int row_bytes, int mw, int tw, const double *__restrict__ wt,
                                             long long row, long long piv, int &dot, double &D)
{
    const unsigned *xr = reinterpret_cast<const unsigned *>(X + (size_t)row * row_bytes);
    const unsigned *xp = reinterpret_cast<const unsigned *>(X + (size_t)piv * row_bytes);
    int s = 0;
    for (int w = 0; w < mw; ++w) s = (int)__dp4a(__ldg(xr + w), __ldg(xp + w), (unsigned)s);
    double acc[4] = {0.0, 0.0, 0.0, 0.0};
    for (int w = 0; w < tw; ++w) {
        const unsigned ur = __ldg(xr + mw + w), up = __ldg(xp + mw + w);
#pragma unroll
        for (int c = 0; c < 4; ++c)
            acc[w & 3] = fma((double)((ur >> (8 * c)) & 255u), wt[4 * w + c] * (double)((up >> (8 * c)) & 255u), acc[w & 3]);
    }
    dot = s;
    D = (acc[0] + acc[1]) + (acc[2] + acc[3]);
}

// per-row norms: the integer squared norm over the main class and the weighted norm N(u) = D(u, u) over the tail
__global__ void u8_norms_kernel(const unsigned char *__restrict__ X, long long n, int row_bytes, int mw, int tw, const double *__restrict__ wt,
                                int *__restrict__ nmain, double *__restrict__ ntail)
{
    const long long j = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    int dot;
    double D;
    u8_pair_sums(X, row_bytes, mw, tw, wt, j, j, dot, D);
    nmain[j] = dot;
    ntail[j] = D;
}

// scaled FP64 copies of chosen rows (ids == nullptr: rows 0 .. count-1) in the layout of the FP64-row kernels
// ([count][pitch2] 16-byte groups of two values, zero padded): (u - min) / range with range < 1e-12 -> 1, in FP64, as
// scale_features (code/main3.cpp:80-86).  pos[k] = byte position of feature k in a packed row, or -1 for a constant
// feature (every training row then holds min_k there, i.e. the scaled value 0).
__global__ void u8_gather_scale_kernel(const unsigned char *__restrict__ X, int row_bytes, const int *__restrict__ pos, const double *__restrict__ minv,
                                       const double *__restrict__ maxv, int d, const int *__restrict__ ids, long long count, int pitch2,
                                       double *__restrict__ out)
{
    const long long total = count * pitch2;
    for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < total; g += (long long)gridDim.x * blockDim.x) {
        const long long k = g / pitch2;
        const int q = (int)(g - k * pitch2);
        const long long row = ids ? ids[k] : k;
#pragma unroll
        for (int c = 0; c < 2; ++c) {
            const int col = 2 * q + c;
            double x = 0.0;
            if (col < d && pos[col] >= 0) {
                double range = maxv[col] - minv[col];
                if (range < 1e-12) range = 1.0;
                x = ((double)X[(size_t)row * row_bytes + pos[col]] - minv[col]) / range;
            }
            out[g * 2 + c] = x;
        }
    }
}

// introspection: one kernel row K(., pivot) exactly as the solver forms it (a thread per row)
__global__ void u8_kernel_row_kernel(const unsigned char *__restrict__ X, long long n, int row_bytes, int mw, int tw, const double *__restrict__ wt,
                                     double wR, const int *__restrict__ nmain, const double *__restrict__ ntail, long long pivot, double gamma,
                                     double *__restrict__ out)
{
    const long long j = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    int dot;
    double D;
    u8_pair_sums(X, row_bytes, mw, tw, wt, j, pivot, dot, D);
    const int S = nmain[j] + nmain[pivot] - 2 * dot;
    double tail = (ntail[j] + ntail[pivot]) - 2.0 * D;
    if (tail < 0.0) tail = 0.0;
    out[j] = svm_exp(-gamma * fma((double)S, wR, tail));
}

} // namespace svm
