// aux_kernels.cuh -- the kernels either side of the SMO loop: ingest/scale, state init, the batched
// decision function (prediction and warm-start f), and the small introspection kernels.
#pragma once
#include "device_math.cuh"
#include "smo_kernel.cuh"

namespace svm {

// ---- per-feature min/max: replaces find_min_max (code/main3.cpp:57-71) and the reference GPU's
// four full-size scratch copies + 4 passes (code/gpu_svm_main3.cu:540-572).  One read of X:
// CTA b scans rows [b*rows_per_cta, ...) with thread c on column c, then a second tiny kernel
// folds the per-CTA partials.  Scratch: 2 * gridDim.x * d doubles.
template <typename T>
__global__ void colminmax_partial_kernel(const T *__restrict__ X, long long n, int d, int rows_per_cta,
                                         double *__restrict__ pmin, double *__restrict__ pmax)
{
    const long long r0 = (long long)blockIdx.x * rows_per_cta;
    const long long r1 = min(n, r0 + rows_per_cta);
    for (int c = threadIdx.x; c < d; c += blockDim.x) {
        double lo = __longlong_as_double(0x7ff0000000000000LL), hi = __longlong_as_double(0xfff0000000000000LL);
        for (long long r = r0; r < r1; ++r) {
            const double v = (double)X[r * d + c];
            lo = v < lo ? v : lo;
            hi = hi < v ? v : hi;
        }
        pmin[(size_t)blockIdx.x * d + c] = lo;
        pmax[(size_t)blockIdx.x * d + c] = hi;
    }
}

__global__ void colminmax_final_kernel(const double *__restrict__ pmin, const double *__restrict__ pmax, int parts, int d,
                                       double *__restrict__ minv, double *__restrict__ maxv)
{
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= d) return;
    double lo = pmin[c], hi = pmax[c];
    for (int p = 1; p < parts; ++p) {
        const double a = pmin[(size_t)p * d + c], b = pmax[(size_t)p * d + c];
        lo = a < lo ? a : lo;
        hi = hi < b ? b : hi;
    }
    minv[c] = lo;
    maxv[c] = hi;
}

// ---- scale + narrow + pad: replaces scale_features (code/main3.cpp:74-89): (x - min) / range in
// FP64 with range < 1e-12 -> 1, then rounded once to fp32.  minv == nullptr: X is already scaled.
template <typename T, typename XT>
__global__ void scale_rows_kernel(const T *__restrict__ X, long long n, int d, int pitch4,
                                  const double *__restrict__ minv, const double *__restrict__ maxv,
                                  XT *__restrict__ out)
{
    constexpr int PER = 16 / (int)sizeof(XT);            // values per 16-byte group: 4 fp32 or 2 FP64
    const long long total = n * pitch4;
    for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < total; g += (long long)gridDim.x * blockDim.x) {
        const long long r = g / pitch4;
        const int q = (int)(g - r * pitch4);
#pragma unroll
        for (int c = 0; c < PER; ++c) {
            const int col = PER * q + c;
            double x = 0.0;
            if (col < d) {
                x = (double)X[r * d + col];
                if (minv) {
                    double range = maxv[col] - minv[col];
                    if (range < 1e-12) range = 1.0;
                    x = (x - minv[col]) / range;
                }
            }
            out[g * PER + c] = (XT)x;                      // fp32: the one rounding of the fast path; FP64: the reference's value
        }
    }
}

// ---- FP64 in-place scaling for hosts that call scale_features on their own vectors (main3.cpp:74-89)
__global__ void scale_f64_kernel(double *__restrict__ X, long long n, int d, const double *__restrict__ minv,
                                 const double *__restrict__ maxv)
{
    const long long total = n * d;
    for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < total; g += (long long)gridDim.x * blockDim.x) {
        const int col = (int)(g % d);
        double range = maxv[col] - minv[col];
        if (range < 1e-12) range = 1.0;
        X[g] = (X[g] - minv[col]) / range;
    }
}

// ---- squared norms |x_j|^2: the dot product of every row with itself, formed exactly as the solver
// forms x_i . x_j (a warp per row), so that duplicates are at distance exactly 0
template <int Q, int CH>
__global__ void row_sqnorm_kernel(const float *__restrict__ X, long long n, int pitch4, double *__restrict__ sqn)
{
    const int lane = threadIdx.x & 31;
    const long long j = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (j >= n) return;
    const float4 *xj = reinterpret_cast<const float4 *>(X) + (size_t)j * pitch4;
    double acc = 0.0;
#pragma unroll
    for (int i = 0; i < Q; ++i) {
        const int q = i * 32 + lane;
        if (q < pitch4) { const float4 v = __ldg(xj + q); accum_dot<CH>(v, v, acc); }   // fp32 rows only: FP64 rows need no norms
    }
    acc = warp_sum_butterfly(acc);
    if (lane == 0) sqn[j] = acc;
}

// ---- solver state init: alpha = 0, f = -y (code/main3.cpp:167-172; gpu_svm_main3.cu:152-161)
__global__ void init_state_kernel(const int *__restrict__ y, long long n, double *__restrict__ f, double *__restrict__ alpha,
                                  long long *__restrict__ cacheq, int zero_alpha, int set_f)
{
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        if (zero_alpha) alpha[i] = 0.0;
        if (set_f) f[i] = -(double)y[i];
        if (cacheq) cacheq[i] = -1;
    }
}

// ---- gather rows (support vectors) and their squared norms into a dense block
__global__ void gather_rows_kernel(const void *__restrict__ X, const double *__restrict__ sqn, int pitch4,
                                   const int *__restrict__ ids, int count, void *__restrict__ out, double *__restrict__ out_n)
{
    for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < count; k += gridDim.x * blockDim.x) out_n[k] = sqn ? sqn[ids[k]] : 0.0;
    const long long total = (long long)count * pitch4;
    for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < total; g += (long long)gridDim.x * blockDim.x) {
        const long long k = g / pitch4;
        const int q = (int)(g - k * pitch4);
        reinterpret_cast<float4 *>(out)[g] = reinterpret_cast<const float4 *>(X)[(size_t)ids[k] * pitch4 + q];
    }
}

// ---- batched decision function: replaces the prediction loop (code/main3.cpp:391-402), the
// reference GPU's predict kernel (gpu_svm_main3.cu:277-293, which loops over ALL n rows) and the
// warm-start f rebuild (code/mpi_svm_main2.cpp:169-183).
//   out[i] = init + sum_k coef[k] * K(q_i, sv_k)   (k ascending, sequential FP64, as the reference)
//   then, if ysub != nullptr, out[i] -= ysub[i]     (f_i = sum - y_i)
// A warp owns one or two query rows (registers: two when the rows are fp32, so that every support-vector row read
// from shared memory serves two dot products); the CTA streams the support-vector block through shared memory in
// tiles of DEC_TILE rows (bulk copies, double-buffered), every warp reading each tile.
constexpr int DEC_THREADS = 384;     // 12 warps: up to 168 registers each, room to keep the loads of several rows in flight
constexpr int DEC_WARPS = DEC_THREADS / 32;
constexpr int DEC_TILE = 16;
template <typename XT, int Q> constexpr int dec_rows_per_warp() { return (sizeof(XT) == 4 && Q <= 8) ? 2 : 1; }

// FULL: every one of the Q register groups reaches into the row ((Q - 1) * 32 < pitch4), so the fp32 loop needs no test at all
template <typename XT, int Q, int CH, bool FULL>
__global__ void __launch_bounds__(DEC_THREADS, 1)
decision_kernel(const void *__restrict__ Xq, long long m, const void *__restrict__ SV, const double *__restrict__ svn,
                const double *__restrict__ coef, int nsv, int pitch4, double gamma, double init, const int *__restrict__ ysub, double *__restrict__ out)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    constexpr int QR = dec_rows_per_warp<XT, Q>();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int row_bytes = pitch4 * 16;
    unsigned char *tiles = smem_raw;                                   // [2][DEC_TILE][row_bytes]
    uint64_t *bars = reinterpret_cast<uint64_t *>(tiles + (size_t)2 * DEC_TILE * row_bytes);

    if (threadIdx.x == 0) {
        mbar_init(&bars[0], 1);
        mbar_init(&bars[1], 1);
        mbar_fence_init();
    }
    __syncthreads();

    using Vec = typename RowVec<XT>::type;
    const long long qi0 = ((long long)blockIdx.x * (blockDim.x >> 5) + warp) * QR;     // the host picks the CTA size (<= DEC_THREADS)
    Vec qv[QR][Q];
    double qn[QR];                                                     // |q|^2, the same way the solver forms norms (fp32 rows)
#pragma unroll
    for (int u = 0; u < QR; ++u) {
        const bool active = qi0 + u < m;
#pragma unroll
        for (int i = 0; i < Q; ++i) {
            const int q = i * 32 + lane;
            qv[u][i] = (active && q < pitch4) ? __ldg(reinterpret_cast<const Vec *>(Xq) + (size_t)(qi0 + u) * pitch4 + q) : vec_zero<Vec>();
        }
        qn[u] = 0.0;
        if (sizeof(XT) == 4) {
#pragma unroll
            for (int i = 0; i < Q; ++i)
                if (i * 32 + lane < pitch4) accum_pair<CH>(qv[u][i], qv[u][i], qn[u]);
            qn[u] = warp_sum_butterfly(qn[u]);
        }
    }

    const int ntiles = (nsv + DEC_TILE - 1) / DEC_TILE;
    auto issue_tile = [&](int t) {      // thread 0
        const int rows = min(DEC_TILE, nsv - t * DEC_TILE);
        mbar_expect_tx(&bars[t & 1], (unsigned)(rows * row_bytes));
        bulk_g2s(tiles + (size_t)(t & 1) * DEC_TILE * row_bytes, static_cast<const unsigned char *>(SV) + (size_t)t * DEC_TILE * row_bytes,
                 (unsigned)(rows * row_bytes), &bars[t & 1], L2_EVICT_LAST);
    };
    if (threadIdx.x == 0 && ntiles > 0) issue_tile(0);

    if constexpr (QR == 2) {
        // Two query rows x 8 support-vector rows = 16 dot products at a time.  Every lane first forms its partial sum of
        // each (the lane's 16-byte groups, in order, exactly as everywhere else); the 16 warp sums are then formed by a
        // halving exchange: at offset 16 a lane keeps the 8 sums of "its" query row and sends the other 8, at offsets
        // 8, 4, 2 it keeps half of what it has, at offset 1 it adds.  Each addition pairs the same two operands as the
        // plain xor butterfly (which leaves the same bits in every lane), so lane l ends up holding the butterfly's
        // value for query row l >> 4 and support-vector row (l >> 1) & 7 -- with 16 shuffles per 16 sums instead of 80.
        double *prodbuf = reinterpret_cast<double *>(bars + 2) + warp * 16;      // [query row][8] products of the half tile
        const int my_u = lane >> 4, my_r = (lane >> 1) & 7;
        int qoff[Q];                                                            // this lane's 16-byte group of each register group, clamped into the row
#pragma unroll
        for (int i = 0; i < Q; ++i) qoff[i] = min(i * 32 + lane, pitch4 - 1);
        const double my_qn = my_u ? qn[QR - 1] : qn[0];
        double sum = init;                                                      // lanes 0-15: query row 0, lanes 16-31: query row 1
        for (int t = 0; t < ntiles; ++t) {
            if (threadIdx.x == 0 && t + 1 < ntiles) issue_tile(t + 1);   // buffer (t+1)&1 was released by the barrier below
            mbar_wait(&bars[t & 1], (t >> 1) & 1);
            const int rows = min(DEC_TILE, nsv - t * DEC_TILE);
            const unsigned char *tile = tiles + (size_t)(t & 1) * DEC_TILE * row_bytes;
#pragma unroll 1
            for (int half = 0; half < 2; ++half) {
                const int rows_here = min(8, rows - half * 8);
                if (rows_here <= 0) break;
                double P[16];                                                   // [u * 8 + r]
#pragma unroll
                for (int j = 0; j < 16; ++j) P[j] = 0.0;
#pragma unroll
                for (int r = 0; r < 8; ++r) {
                    const Vec *row = reinterpret_cast<const Vec *>(tile + (size_t)(half * 8 + r) * row_bytes);
#pragma unroll
                    for (int i = 0; i < Q; ++i) {
                        // no per-lane test: the query registers are zero beyond the row, so a (clamped, hence finite) element
                        // read there adds +-0 to the lane's sum; only groups wholly beyond the row are skipped (warp-uniform)
                        if (FULL || i * 32 < pitch4) {
                            const Vec sv = (FULL && i < Q - 1) ? row[i * 32 + lane] : row[qoff[i]];   // constant offsets from one address
                            accum_pair<CH>(sv, qv[0][i], P[r]);
                            accum_pair<CH>(sv, qv[QR - 1][i], P[8 + r]);
                        }
                    }
                }
                double v8[8], v4[4], v2[2], v1;
                {
                    const bool up = (lane & 16) != 0;
#pragma unroll
                    for (int j = 0; j < 8; ++j) v8[j] = (up ? P[8 + j] : P[j]) + __shfl_xor_sync(FULL, up ? P[j] : P[8 + j], 16);
                }
                {
                    const bool up = (lane & 8) != 0;
#pragma unroll
                    for (int j = 0; j < 4; ++j) v4[j] = (up ? v8[4 + j] : v8[j]) + __shfl_xor_sync(FULL, up ? v8[j] : v8[4 + j], 8);
                }
                {
                    const bool up = (lane & 4) != 0;
#pragma unroll
                    for (int j = 0; j < 2; ++j) v2[j] = (up ? v4[2 + j] : v4[j]) + __shfl_xor_sync(FULL, up ? v4[j] : v4[2 + j], 4);
                }
                {
                    const bool up = (lane & 2) != 0;
                    v1 = (up ? v2[1] : v2[0]) + __shfl_xor_sync(FULL, up ? v2[0] : v2[1], 2);
                }
                v1 = v1 + __shfl_xor_sync(FULL, v1, 1);
                const int k_mine = t * DEC_TILE + half * 8 + my_r;
                const bool valid = my_r < rows_here;
                const double kv = svm_exp(-gamma * dist2<XT>(my_qn, valid ? svn[k_mine] : 0.0, v1));
                __syncwarp();                                                   // the previous half's products have been read
                if (valid && !(lane & 1)) prodbuf[my_u * 8 + my_r] = coef[k_mine] * kv;
                __syncwarp();
                for (int r = 0; r < rows_here; ++r) sum = sum + prodbuf[my_u * 8 + r];   // curr += alpha[j]*y[j] * kernel(...), k ascending
            }
            fence_proxy_async();
            __syncthreads();                                              // everyone is done with buffer t&1
        }
        if ((lane & 15) == 0 && qi0 + my_u < m) {
            if (ysub) sum = sum - (double)ysub[qi0 + my_u];
            out[qi0 + my_u] = sum;
        }
    } else {
        double sum = init;
        for (int t = 0; t < ntiles; ++t) {
            if (threadIdx.x == 0 && t + 1 < ntiles) issue_tile(t + 1);   // buffer (t+1)&1 was released by the barrier below
            mbar_wait(&bars[t & 1], (t >> 1) & 1);
            const int rows = min(DEC_TILE, nsv - t * DEC_TILE);
            const unsigned char *tile = tiles + (size_t)(t & 1) * DEC_TILE * row_bytes;
            double myk = 0.0;
            for (int r = 0; r < rows; ++r) {
                const Vec *row = reinterpret_cast<const Vec *>(tile + (size_t)r * row_bytes);
                double acc = 0.0;
#pragma unroll
                for (int i = 0; i < Q; ++i) {
                    const int q = i * 32 + lane;
                    if (q < pitch4) accum_pair<CH>(row[q], qv[0][i], acc);
                }
                acc = warp_sum_butterfly(acc);
                if (lane == r) myk = acc;
            }
            const double myn = (sizeof(XT) == 4 && lane < rows) ? svn[t * DEC_TILE + lane] : 0.0;
            myk = svm_exp(-gamma * dist2<XT>(qn[0], myn, myk));        // lanes >= rows compute a value nobody reads
            const double mycoef = lane < rows ? coef[t * DEC_TILE + lane] : 0.0;
            for (int r = 0; r < rows; ++r) {
                const double k = __shfl_sync(FULL, myk, r);
                const double c = __shfl_sync(FULL, mycoef, r);
                sum = sum + c * k;                                        // curr += alpha[j]*y[j] * kernel(...)
            }
            fence_proxy_async();
            __syncthreads();                                              // everyone is done with buffer t&1
        }
        if (lane == 0 && qi0 < m) {
            if (ysub) sum = sum - (double)ysub[qi0];
            out[qi0] = sum;
        }
    }
}

// ---- the same decision values for a FEW query rows (the tensor-core path's re-evaluation list): `parts` CTAs per
// query row share out the support vectors (rows read straight from L2, two in flight per warp), the products
// coef_k K(q, sv_k) go to shared memory (parts == 1) or to `kprod` in global memory, and one thread of the row's last
// CTA to finish adds them up in k order -- the same operations in the same order as decision_kernel, hence the same
// bits, at a small fraction of the latency of one warp walking all support vectors.
// kprod: [rows][nsv] doubles, done: [rows] ints zeroed before the launch (both unused when parts == 1).
constexpr int DROWS_THREADS = 512;
template <typename XT, int Q, int CH>
__global__ void __launch_bounds__(DROWS_THREADS, 1)
decision_rows_kernel(const void *__restrict__ Xq, const void *__restrict__ SV, const double *__restrict__ svn,
                     const double *__restrict__ coef, int nsv, int pitch4, double gamma, double init, double *__restrict__ out,
                     int parts, double *__restrict__ kprod, int *__restrict__ done)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    double *kbuf = reinterpret_cast<double *>(smem_raw);
    __shared__ int is_last;
    using Vec = typename RowVec<XT>::type;
    constexpr int W = DROWS_THREADS / 32;
    constexpr bool TWO = sizeof(XT) == 4 && Q <= 8;                            // registers for a second row in flight
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int qrow = blockIdx.x / parts, part = blockIdx.x - qrow * parts;
    const int k_begin = (int)((long long)nsv * part / parts), k_end = (int)((long long)nsv * (part + 1) / parts);
    const Vec *xq = reinterpret_cast<const Vec *>(Xq) + (size_t)qrow * pitch4;
    double *dst = parts == 1 ? kbuf : kprod + (size_t)qrow * nsv;
    Vec qv[Q];
#pragma unroll
    for (int i = 0; i < Q; ++i) {
        const int q = i * 32 + lane;
        qv[i] = q < pitch4 ? __ldg(xq + q) : vec_zero<Vec>();
    }
    double qn = 0.0;
    if (sizeof(XT) == 4) {
#pragma unroll
        for (int i = 0; i < Q; ++i)
            if (i * 32 + lane < pitch4) accum_pair<CH>(qv[i], qv[i], qn);
        qn = warp_sum_butterfly(qn);
    }
    for (int k = k_begin + warp; k < k_end; k += (TWO ? 2 : 1) * W) {
        const int k2 = k + W;
        const bool second = TWO && k2 < k_end;
        const Vec *row = reinterpret_cast<const Vec *>(SV) + (size_t)k * pitch4;
        const Vec *row2 = reinterpret_cast<const Vec *>(SV) + (size_t)(second ? k2 : k) * pitch4;
        Vec r1[Q], r2[TWO ? Q : 1];
#pragma unroll
        for (int i = 0; i < Q; ++i) {
            const int q = i * 32 + lane;
            r1[i] = q < pitch4 ? __ldg(row + q) : vec_zero<Vec>();
            if (TWO) r2[i] = q < pitch4 ? __ldg(row2 + q) : vec_zero<Vec>();
        }
        double acc = 0.0, acc2 = 0.0;
#pragma unroll
        for (int i = 0; i < Q; ++i)
            if (i * 32 + lane < pitch4) accum_pair<CH>(r1[i], qv[i], acc);
        if (TWO) {
#pragma unroll
            for (int i = 0; i < Q; ++i)
                if (i * 32 + lane < pitch4) accum_pair<CH>(r2[i], qv[i], acc2);
            warp_sum_butterfly2(acc, acc2, lane);
        } else {
            acc = warp_sum_butterfly(acc);
        }
        if (lane == 0) {
            dst[k] = coef[k] * svm_exp(-gamma * dist2<XT>(qn, sizeof(XT) == 4 ? svn[k] : 0.0, acc));
            if (second) dst[k2] = coef[k2] * svm_exp(-gamma * dist2<XT>(qn, svn[k2], acc2));
        }
    }
    if (parts > 1) {
        __threadfence();
        __syncthreads();
        if (threadIdx.x == 0) is_last = atomicAdd(done + qrow, 1) == parts - 1;
        __syncthreads();
        if (!is_last) return;
        __threadfence();
        for (int k = threadIdx.x; k < nsv; k += DROWS_THREADS) kbuf[k] = __ldcg(dst + k);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        double sum = init;
        for (int k = 0; k < nsv; ++k) sum = sum + kbuf[k];
        out[qrow] = sum;
    }
}

inline size_t decision_smem_bytes(int pitch4) { return (size_t)2 * DEC_TILE * pitch4 * 16 + 16 + (size_t)DEC_WARPS * 16 * sizeof(double); }

// ---- accuracy: replaces the predict/compare/reduce_sum tail (code/main3.cpp:399-401;
// gpu_svm_main3.cu:291,299-315): "> 0" (serial/GPU) or ">= 0" (MPI programs)
__global__ void count_correct_kernel(const double *__restrict__ scores, const int *__restrict__ y, long long m, int ge_zero,
                                     int *__restrict__ correct)
{
    int c = 0;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < m; i += (long long)gridDim.x * blockDim.x) {
        const int pred = ge_zero ? (scores[i] >= 0 ? 1 : -1) : (scores[i] > 0 ? 1 : -1);
        c += pred == y[i];
    }
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) c += __shfl_xor_sync(FULL, c, off);
    if ((threadIdx.x & 31) == 0 && c) atomicAdd(correct, c);
}

// ---- introspection: one kernel row exactly as the solver forms it (a warp per row, no staging)
template <typename XT, int Q, int CH>
__global__ void kernel_row_kernel(const void *__restrict__ X, const double *__restrict__ sqn, long long n, int pitch4,
                                  long long pivot, double gamma, double *__restrict__ out)
{
    const int lane = threadIdx.x & 31;
    const long long j = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (j >= n) return;
    using Vec = typename RowVec<XT>::type;
    const Vec *xp = reinterpret_cast<const Vec *>(X) + (size_t)pivot * pitch4;
    const Vec *xj = reinterpret_cast<const Vec *>(X) + (size_t)j * pitch4;
    double acc = 0.0;
#pragma unroll
    for (int i = 0; i < Q; ++i) {
        const int q = i * 32 + lane;
        if (q < pitch4) accum_pair<CH>(__ldg(xj + q), __ldg(xp + q), acc);
    }
    acc = warp_sum_butterfly(acc);
    if (lane == 0) out[j] = svm_exp(-gamma * dist2<XT>(sizeof(XT) == 4 ? sqn[pivot] : 0.0, sizeof(XT) == 4 ? sqn[j] : 0.0, acc));
}

// ---- introspection: one working-set selection with the solver's own compare (single CTA)
__global__ void select_pair_kernel(const double *__restrict__ alpha, const double *__restrict__ f, const int *__restrict__ y,
                                   long long n, double C, double eps, long long *__restrict__ out)
{
    __shared__ Best slo[32], shi[32];
    LocalBest b;
    best_init(b);
    for (long long j = threadIdx.x; j < n; j += blockDim.x) best_consider(b, j, f[j], alpha[j], y[j], -1, C, eps);
    best_warp_reduce(b);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) { slo[warp] = b.lo; shi[warp] = b.hi; }
    __syncthreads();
    if (warp == 0) {
        best_init(b);
        if (lane < (int)(blockDim.x >> 5)) { b.lo = slo[lane]; b.hi = shi[lane]; }
        best_warp_reduce(b);
        if (lane == 0) {
            out[0] = b.lo.idx == IDX_NONE ? n : b.lo.idx;
            out[1] = b.hi.idx == IDX_NONE ? n : b.hi.idx;
        }
    }
}

// ---- introspection: the scalar part of one iteration (calculate_U_V + the pair update) exactly as the solver runs it
__global__ void pair_update_kernel(int y_h, int y_l, double a_h, double a_l, double b_high, double b_low, double K12, double C, double eps,
                                   double *__restrict__ out, int *__restrict__ stop)
{
    const PairStep r = pair_update(y_h, y_l, a_h, a_l, b_high, b_low, K12, C, eps);
    out[0] = r.U; out[1] = r.V; out[2] = r.eta; out[3] = r.na_h; out[4] = r.na_l; out[5] = r.c_h; out[6] = r.c_l;
    *stop = r.stop;
}

} // namespace svm
