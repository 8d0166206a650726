// u8_rows.cuh -- kernels around the uint8 storage of X (svmb200_set_storage "u8"): packing the raw rows into the
// [main class | tail] layout the SMO kernel walks, the per-row norms, scaled FP64 copies of chosen rows for the
// decision function, and one kernel row for the parity tests.  The arithmetic is described in smo_kernel.cuh.
//
// Replaces, for integer-valued input, scale_features + the FP64 copy of X the reference keeps (code/main3.cpp:74-89,
// :317): the statistics are still find_min_max's (:57-71), but X stays as the bytes it arrived as.
#pragma once
#include "smo_kernel.cuh"

namespace svm {

// The packed rows: two arrays in one allocation -- [n][main_bytes] bytes of the main class (main_bytes = 4 * mw, whole
// 128-byte slots), then [n][tail_pitch] bytes of the tail (4 * tw bytes + 16 of padding; see SmoArgs).  A "position" of a
// byte is its offset in the virtual row [main class | tail incl. padding].
struct U8View {
    const unsigned char *main, *tail;
    int main_bytes, tail_pitch;
    __host__ __device__ U8View(const void *X, long long n, int mw, int tail_pitch_)
        : main(static_cast<const unsigned char *>(X)), tail(static_cast<const unsigned char *>(X) + (size_t)n * mw * 4), main_bytes(mw * 4),
          tail_pitch(tail_pitch_) {}
    __device__ __forceinline__ const unsigned *main_row(long long r) const { return reinterpret_cast<const unsigned *>(main + (size_t)r * main_bytes); }
    __device__ __forceinline__ const unsigned *tail_row(long long r) const { return reinterpret_cast<const unsigned *>(tail + (size_t)r * tail_pitch); }
    __device__ __forceinline__ unsigned char byte_at(long long r, int pos) const
    {
        return pos < main_bytes ? main[(size_t)r * main_bytes + pos] : tail[(size_t)r * tail_pitch + (pos - main_bytes)];
    }
};

// rows of `src` ([n][d] raw bytes) -> the packed arrays: byte b of a virtual row is feature src_col[b] (or 0 where src_col[b] < 0)
__global__ void u8_pack_kernel(const unsigned char *__restrict__ src, long long n, int d, int mw, int tail_pitch, const int *__restrict__ src_col,
                               unsigned char *__restrict__ out)
{
    const int vw = mw + tail_pitch / 4;                         // words of a virtual row
    const long long total = n * vw;
    unsigned *out_main = reinterpret_cast<unsigned *>(out);
    unsigned *out_tail = reinterpret_cast<unsigned *>(out + (size_t)n * mw * 4);
    for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < total; g += (long long)gridDim.x * blockDim.x) {
        const long long r = g / vw;
        const int w = (int)(g - r * vw);
        unsigned v = 0;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            const int col = src_col[4 * w + c];
            if (col >= 0) v |= (unsigned)src[r * d + col] << (8 * c);
        }
        if (w < mw) out_main[r * mw + w] = v;
        else out_tail[r * (tail_pitch / 4) + (w - mw)] = v;
    }
}

// K-independent pieces of the uint8 arithmetic for one (row, pivot) pair, by ONE thread straight from global memory:
// the integer dot product over the main class and D(row, pivot) over the tail in the solver's order -- (D_0 + D_1) +
// (D_2 + D_3), D_c over the words w = c (mod 4) ascending (a zero byte of either row adds +0, so the dense walk here and
// the solver's walk over the pivots' non-zero positions give the same bits).
__device__ __forceinline__ void u8_pair_sums(const U8View &X, int mw, int tw, const double *__restrict__ wt, long long row, long long piv, int &dot,
                                             double &D)
{
    const unsigned *xr = X.main_row(row), *xp = X.main_row(piv);
    int s = 0;
    for (int w = 0; w < mw; ++w) s = (int)__dp4a(__ldg(xr + w), __ldg(xp + w), (unsigned)s);
    const unsigned *tr = X.tail_row(row), *tp = X.tail_row(piv);
    double acc[4] = {0.0, 0.0, 0.0, 0.0};
    for (int w = 0; w < tw; ++w) {
        const unsigned ur = __ldg(tr + w), up = __ldg(tp + w);
#pragma unroll
        for (int c = 0; c < 4; ++c)
            acc[w & 3] = fma((double)((ur >> (8 * c)) & 255u), wt[4 * w + c] * (double)((up >> (8 * c)) & 255u), acc[w & 3]);
    }
    dot = s;
    D = (acc[0] + acc[1]) + (acc[2] + acc[3]);
}

// per-row norms: the integer squared norm over the main class and the weighted norm N(u) = D(u, u) over the tail
__global__ void u8_norms_kernel(const U8View X, long long n, int mw, int tw, const double *__restrict__ wt, int *__restrict__ nmain,
                                double *__restrict__ ntail)
{
    const long long j = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    int dot;
    double D;
    u8_pair_sums(X, mw, tw, wt, j, j, dot, D);
    nmain[j] = dot;
    ntail[j] = D;
}

// scaled FP64 copies of chosen rows (ids == nullptr: rows 0 .. count-1) in the layout of the FP64-row kernels
// ([count][pitch2] 16-byte groups of two values, zero padded): (u - min) / range with range < 1e-12 -> 1, in FP64, as
// scale_features (code/main3.cpp:80-86).  pos[k] = position of feature k in a packed row, or -1 for a constant
// feature (every training row then holds min_k there, i.e. the scaled value 0).
__global__ void u8_gather_scale_kernel(const U8View X, const int *__restrict__ pos, const double *__restrict__ minv, const double *__restrict__ maxv,
                                       int d, const int *__restrict__ ids, long long count, int pitch2, double *__restrict__ out)
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
                x = ((double)X.byte_at(row, pos[col]) - minv[col]) / range;
            }
            out[g * 2 + c] = x;
        }
    }
}

// introspection: one kernel row K(., pivot) exactly as the solver forms it (a thread per row)
__global__ void u8_kernel_row_kernel(const U8View X, long long n, int mw, int tw, const double *__restrict__ wt, double wR,
                                     const int *__restrict__ nmain, const double *__restrict__ ntail, long long pivot, double gamma,
                                     double *__restrict__ out)
{
    const long long j = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    int dot;
    double D;
    u8_pair_sums(X, mw, tw, wt, j, pivot, dot, D);
    const int S = nmain[j] + nmain[pivot] - 2 * dot;
    double tail = (ntail[j] + ntail[pivot]) - 2.0 * D;
    if (tail < 0.0) tail = 0.0;
    out[j] = svm_exp(-gamma * fma((double)S, wR, tail));
}

} // namespace svm
