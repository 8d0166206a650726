// tc_decision.cuh -- the batched decision function on the 5th-generation tensor cores (tcgen05 + TMEM).
//
// The one GEMM-shaped piece of the path: scores = Q [m x d] . SV^T [d x n_sv] followed by exp and a weighted row
// sum (code/main3.cpp:391-402 over all test rows at once; the reference GPU's predict kernel,
// code/gpu_svm_main3.cu:277-293, walks ALL n training rows per test row in scalar FP64).
//
// Work is cut into pairs (128 query rows = the M = 128 lanes of a TMEM accumulator) x (128 support vectors = N = 128
// accumulator columns).  Operands are split and laid out ONCE in global memory (tc_pack_kernel) in the canonical K-major,
// no-swizzle core-matrix layout: a tf32 operand keeps 10 mantissa bits, so every fp32 value x is split into hi = x with
// the low 13 mantissa bits cleared and lo = x - hi (exact), and each K step issues three tcgen05.mma kind::tf32:
// hi.hi + hi.lo + lo.hi ("TF32x3", fp32 accumulation in TMEM).  One K step (32 columns) of a tile is 32 KB contiguous
// (hi tile, lo tile) and reaches shared memory by one bulk copy (TMA) per operand.  The kernel is warp-specialised:
// warp 0 = producer (bulk copies into a 3-stage ring), warp 1 = MMA issuer (into one of two 128-column TMEM
// accumulators), warps 2-5 = epilogue (tcgen05.ld of the other accumulator, K = exp(-gamma (|q|^2 + |s|^2 - 2 q.s)),
// coef_k K added to the row's sum) -- copies, tensor core and epilogue all overlap.  Scores from this kernel carry an
// error of ~1e-3 at most (fp32 accumulation over d = 784 and 3 tf32 products, fp32 exp); the caller re-evaluates every
// row whose |score| is below a threshold with the exact kernel (decision_rows_kernel), so the predicted labels -- and
// the accuracy -- are those of the exact path.
#pragma once
#include "device_math.cuh"

namespace svm {

constexpr int TC_M = 128;          // query rows per CTA = TMEM lanes
constexpr int TC_N = 128;          // support vectors per tile = TMEM columns (fp32)
constexpr int TC_KB = 32;          // K elements staged per step (8 core-matrix columns of 4 tf32 each)
constexpr int TC_TILE_BYTES = TC_M * TC_KB * 4;   // 16 KB per operand tile

// K-major, no swizzle: core matrix = 8 rows x 16 bytes (128 B, rows 16 B apart); core matrices adjacent in K are
// 128 B apart (LBO), core matrices adjacent in M/N are 8 x 128 B = 1024 B apart (SBO)
__device__ __forceinline__ uint64_t umma_desc_kmajor_noswizzle(uint32_t smem_addr)
{
    const uint64_t start = (smem_addr & 0x3FFFF) >> 4;
    const uint64_t lbo = 128 >> 4, sbo = 1024 >> 4;
    return start | (lbo << 16) | (sbo << 32) | (1ull << 46);   // bits 46-47: descriptor version 1 (sm_100); layout_type 0
}
// kind::tf32, fp32 accumulate, both operands K-major, M x N
__device__ __forceinline__ uint32_t umma_idesc_tf32(int M, int N)
{
    return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate)
{
    asm volatile(
        "{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n}\n" ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t *bar)
{
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

__device__ __forceinline__ void split_store(const float4 v, unsigned char *hi, unsigned char *lo, int off)
{
    float4 t, l;
    t.x = __uint_as_float(__float_as_uint(v.x) & 0xffffe000u);
    t.y = __uint_as_float(__float_as_uint(v.y) & 0xffffe000u);
    t.z = __uint_as_float(__float_as_uint(v.z) & 0xffffe000u);
    t.w = __uint_as_float(__float_as_uint(v.w) & 0xffffe000u);
    l.x = v.x - t.x; l.y = v.y - t.y; l.z = v.z - t.z; l.w = v.w - t.w;
    *reinterpret_cast<float4 *>(hi + off) = t;      // already a tf32 value: nothing left for the tensor core to drop or round
    *reinterpret_cast<float4 *>(lo + off) = l;
}

constexpr int TCP_THREADS = 192;
constexpr int TCP_STAGES = 3;
constexpr int TCP_PANEL_BYTES = 2 * TC_TILE_BYTES;     // hi + lo of one operand for one K step: 32 KB

// dst[tile][kb][hi|lo][16 KB core-matrix layout] from row-major fp32 rows (zero beyond nrows / d)
__global__ void tc_pack_kernel(const float *__restrict__ src, long long nrows, int pitch, int d, int kblocks, unsigned char *__restrict__ dst)
{
    const long long ntiles = (nrows + TC_M - 1) / TC_M;
    const long long total = ntiles * kblocks * (TC_M * (TC_KB / 4));          // 16-byte chunks
    for (long long id = (long long)blockIdx.x * blockDim.x + threadIdx.x; id < total; id += (long long)gridDim.x * blockDim.x) {
        const int c = (int)(id & 7);
        const int r = (int)((id >> 3) & (TC_M - 1));
        const long long tk = id >> 10;                                         // tile * kblocks + kb
        const int kb = (int)(tk % kblocks);
        const long long tile = tk / kblocks;
        const long long row = tile * TC_M + r;
        const int k = kb * TC_KB + 4 * c;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (row < nrows && k < d) v = __ldg(reinterpret_cast<const float4 *>(src + (size_t)row * pitch + k));
        unsigned char *panel = dst + (size_t)tk * TCP_PANEL_BYTES;
        split_store(v, panel, panel + TC_TILE_BYTES, (r >> 3) * 1024 + c * 128 + (r & 7) * 16);
    }
}

__device__ __forceinline__ void mbar_arrive(uint64_t *bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// The (query tile, SV tile) pairs are numbered q * ntiles + n and dealt to the CTAs in equal contiguous ranges, so every SM
// has the same amount of work whatever m / 128 is.  A query tile whose SV tiles are spread over several CTAs gets one row
// of partial sums per CTA ("slot" = position of the CTA among those sharing the tile); tc_sum_partials_kernel adds the
// slots in order, so the result does not depend on scheduling.
__host__ __device__ inline long long tcp_first_pair(long long cta, long long pairs, int ctas) { return cta * pairs / ctas; }

__global__ void __launch_bounds__(TCP_THREADS, 1)
decision_tc_pipelined_kernel(const unsigned char *__restrict__ qP, const double *__restrict__ qn, long long m,
                             const unsigned char *__restrict__ svP, const double *__restrict__ svn, const double *__restrict__ coef,
                             int nsv, int kblocks, double gamma, double *__restrict__ partial, long long mpad, float *__restrict__ dots)
{
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    unsigned char *stages = smem_raw;                                                    // [3][A panel 32 KB | B panel 32 KB]
    uint64_t *full = reinterpret_cast<uint64_t *>(stages + TCP_STAGES * 2 * TCP_PANEL_BYTES);
    uint64_t *empty = full + TCP_STAGES, *tfull = empty + TCP_STAGES, *tempty = tfull + 2;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(tempty + 2);
    float *tile_coef = reinterpret_cast<float *>(tmem_slot + 4), *tile_svn = tile_coef + 2 * TC_N;   // per accumulator: coef and |s|^2 of the tile
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int ntiles = (nsv + TC_N - 1) / TC_N;
    const long long pairs = mpad / TC_M * ntiles;
    const long long p0 = tcp_first_pair(blockIdx.x, pairs, gridDim.x), p1 = tcp_first_pair(blockIdx.x + 1, pairs, gridDim.x);

    if (threadIdx.x == 0) {
        for (int s = 0; s < TCP_STAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
        for (int a = 0; a < 2; ++a) { mbar_init(&tfull[a], 1); mbar_init(&tempty[a], 4); }
        mbar_fence_init();
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(2 * TC_N) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ---------------- producer: one bulk copy per operand panel
        if (lane == 0) {
            unsigned it = 0;
            for (long long p = p0; p < p1; ++p) {
                const long long q = p / ntiles;
                const int n = (int)(p - q * ntiles);
                for (int kb = 0; kb < kblocks; ++kb, ++it) {
                    const int s = it % TCP_STAGES;
                    const unsigned u = it / TCP_STAGES;
                    if (u >= 1) mbar_wait(&empty[s], (u - 1) & 1);
                    unsigned char *dstA = stages + (size_t)s * 2 * TCP_PANEL_BYTES, *dstB = dstA + TCP_PANEL_BYTES;
                    mbar_expect_tx(&full[s], 2u * TCP_PANEL_BYTES);
                    bulk_g2s(dstA, qP + ((size_t)q * kblocks + kb) * TCP_PANEL_BYTES, TCP_PANEL_BYTES, &full[s], L2_EVICT_LAST);
                    bulk_g2s(dstB, svP + ((size_t)n * kblocks + kb) * TCP_PANEL_BYTES, TCP_PANEL_BYTES, &full[s], L2_EVICT_LAST);
                }
            }
        }
    } else if (warp == 1) {
        // ---------------- MMA issuer
        const uint32_t idesc = umma_idesc_tf32(TC_M, TC_N);
        unsigned it = 0;
        for (unsigned t = 0; t < (unsigned)(p1 - p0); ++t) {
            const int acc = t & 1;
            const unsigned v = t >> 1;
            if (v >= 1) mbar_wait(&tempty[acc], (v - 1) & 1);                 // the epilogue has drained this accumulator
            tc_fence_after();
            for (int kb = 0; kb < kblocks; ++kb, ++it) {
                const int s = it % TCP_STAGES;
                mbar_wait(&full[s], (it / TCP_STAGES) & 1);
                tc_fence_after();
                if (lane == 0) {
                    const uint32_t a_hi = smem_u32(stages + (size_t)s * 2 * TCP_PANEL_BYTES), a_lo = a_hi + TC_TILE_BYTES;
                    const uint32_t b_hi = a_hi + TCP_PANEL_BYTES, b_lo = b_hi + TC_TILE_BYTES;
                    const uint32_t d_tmem = tmem_base + (uint32_t)(acc * TC_N);
#pragma unroll
                    for (int ks = 0; ks < TC_KB / 8; ++ks) {
                        const uint32_t koff = ks * 2 * 128;
                        const uint64_t ah = umma_desc_kmajor_noswizzle(a_hi + koff), al = umma_desc_kmajor_noswizzle(a_lo + koff);
                        const uint64_t bh = umma_desc_kmajor_noswizzle(b_hi + koff), bl = umma_desc_kmajor_noswizzle(b_lo + koff);
                        umma_tf32(d_tmem, ah, bh, idesc, (kb | ks) ? 1u : 0u);
                        umma_tf32(d_tmem, ah, bl, idesc, 1u);
                        umma_tf32(d_tmem, al, bh, idesc, 1u);
                    }
                    umma_commit(&empty[s]);                                   // the stage may be refilled once these MMAs have read it
                    if (kb == kblocks - 1) umma_commit(&tfull[acc]);          // and the accumulator is complete
                }
                __syncwarp();
            }
        }
    } else {
        // ---------------- epilogue: warps 2..5 own TMEM lane quarters 2,3,0,1
        const int quarter = warp & 3;
        const int row_in_tile = quarter * 32 + lane;
        const float fgamma = (float)gamma;
        const int et = threadIdx.x - 64;                                       // 0..127 among the epilogue threads
        double sum = 0.0;
        float my_qn = 0.f;
        long long my_row = 0, cur_q = -1;
        int slot = 0;
        unsigned t = 0;
        for (long long p = p0; p < p1; ++p, ++t) {
            const long long q = p / ntiles;
            const int n = (int)(p - q * ntiles);
            if (q != cur_q) {                                                  // a new query tile: which of its partial-sum rows is mine
                cur_q = q;
                my_row = q * TC_M + row_in_tile;
                my_qn = my_row < m ? (float)qn[my_row] : 0.f;
                sum = 0.0;
                long long c = (q * ntiles) * (long long)gridDim.x / pairs;     // the CTA holding the tile's first pair
                while (c > 0 && tcp_first_pair(c, pairs, gridDim.x) > q * ntiles) --c;
                while (tcp_first_pair(c + 1, pairs, gridDim.x) <= q * ntiles) ++c;
                slot = (int)(blockIdx.x - c);
            }
            const int acc = t & 1;
            const int n0 = n * TC_N;
            // this tile's coefficients and norms (the accumulator-specific half of the arrays: the MMA of the next pair is already running)
            const int k_mine = n0 + et;
            tile_coef[acc * TC_N + et] = k_mine < nsv ? (float)coef[k_mine] : 0.f;
            tile_svn[acc * TC_N + et] = k_mine < nsv ? (float)svn[k_mine] : 0.f;
            asm volatile("bar.sync 1, 128;" ::: "memory");                   // the four epilogue warps only
            mbar_wait(&tfull[acc], (t >> 1) & 1);
            tc_fence_after();
#pragma unroll 1
            for (int c0 = 0; c0 < TC_N; c0 += 32) {
                uint32_t v[32];
                const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(acc * TC_N + c0);
                asm volatile(
                    "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, %19, "
                    "%20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                    : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
                      "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]),
                      "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]),
                      "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
                    : "r"(taddr)
                    : "memory");
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                float part = 0.f;                                              // 32 columns in fp32, then into the FP64 row sum
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    const float dot = __uint_as_float(v[j]);
                    if (dots && p == 0) dots[(size_t)row_in_tile * TC_N + c0 + j] = dot;
                    const float d2 = fmaxf(fmaf(-2.f, dot, my_qn + tile_svn[acc * TC_N + c0 + j]), 0.f);
                    part = fmaf(tile_coef[acc * TC_N + c0 + j], __expf(-fgamma * d2), part);
                }
                sum = sum + (double)part;
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tempty[acc]);                          // this warp is done with the accumulator
            if (p + 1 == p1 || (p + 1) / ntiles != q) partial[(size_t)slot * mpad + my_row] = sum;
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(2 * TC_N) : "memory");
}

// out[r] = init + the partial sums of row r, slot by slot (unused slots hold 0)
__global__ void tc_sum_partials_kernel(const double *__restrict__ partial, int slots, long long mpad, long long m, double init, double *__restrict__ out)
{
    const long long r = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= m) return;
    double s = init;
    for (int k = 0; k < slots; ++k) s = s + partial[(size_t)k * mpad + r];
    out[r] = s;
}

// how many CTAs, and how many partial-sum rows per query row, a launch over qtiles x svtiles pairs uses
inline void tcp_plan(long long qtiles, int svtiles, int num_sms, int *ctas, int *slots)
{
    const long long pairs = qtiles * svtiles;
    *ctas = (int)(pairs < num_sms ? pairs : num_sms);
    const long long per = pairs / *ctas;                                       // every CTA gets per or per + 1 pairs
    *slots = (int)(svtiles / per) + 2;
}

inline size_t decision_tc_pipelined_smem_bytes() { return (size_t)TCP_STAGES * 2 * TCP_PANEL_BYTES + 256 + 4 * TC_N * sizeof(float) + 1024; }


// |score| below the threshold: the row goes to the exact kernel
__global__ void flag_uncertain_kernel(const double *__restrict__ scores, long long m, double threshold, int *__restrict__ list, int *__restrict__ count)
{
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < m; i += (long long)gridDim.x * blockDim.x)
        if (!(fabs(scores[i]) >= threshold)) list[atomicAdd(count, 1)] = (int)i;
}
__global__ void scatter_scores_kernel(const double *__restrict__ exact, const int *__restrict__ list, int count, double *__restrict__ scores)
{
    for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < count; k += gridDim.x * blockDim.x) scores[list[k]] = exact[k];
}

} // namespace svm
