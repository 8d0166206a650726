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

// per-row norms: the integer squared norm over the main class and the weighted norm N(u) = D(u, u) over the tail, formed
// by a warp exactly as the solver forms D(u, p) (lane l owns the words l, l + 32, ...; bytes in order; butterfly)
template <int QM, int QT>
__global__ void u8_norms_kernel(const unsigned char *__restrict__ X, long long n, int row_bytes, int mw, int tw, const double *__restrict__ wt,
                                int *__restrict__ nmain, double *__restrict__ ntail)
{
    const int lane = threadIdx.x & 31;
    const long long j = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (j >= n) return;
    const unsigned *x = reinterpret_cast<const unsigned *>(X + (size_t)j * row_bytes);
    int dot = 0;
#pragma unroll
    for (int i = 0; i < QM; ++i) {
        const int w = i * 32 + lane;
        if (w < mw) { const unsigned v = __ldg(x + w); dot = (int)__dp4a(v, v, (unsigned)dot); }
    }
    dot = __reduce_add_sync(FULL, dot);
    double acc = 0.0;
#pragma unroll
    for (int i = 0; i < QT; ++i) {
        const int w = i * 32 + lane;
        if (w < tw) {
            const unsigned v = __ldg(x + mw + w);
#pragma unroll
            for (int c = 0; c < 4; ++c) { const double ud = u8_to_f64(v, c); acc = fma(ud, wt[4 * w + c] * ud, acc); }
        }
    }
    acc = warp_sum_butterfly(acc);
    if (lane == 0) { nmain[j] = dot; ntail[j] = acc; }
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

// introspection: one kernel row K(., pivot) exactly as the solver forms it (a warp per row)
template <int QM, int QT>
__global__ void u8_kernel_row_kernel(const unsigned char *__restrict__ X, long long n, int row_bytes, int mw, int tw, const double *__restrict__ wt,
                                     double wR, const int *__restrict__ nmain, const double *__restrict__ ntail, long long pivot, double gamma,
                                     double *__restrict__ out)
{
    const int lane = threadIdx.x & 31;
    const long long j = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (j >= n) return;
    const double k = u8_kernel_value<QM, QT>(X, row_bytes, mw, tw, wt, wR, nmain, ntail, j, pivot, gamma, lane);
    if (lane == 0) out[j] = k;
}

} // namespace svm
