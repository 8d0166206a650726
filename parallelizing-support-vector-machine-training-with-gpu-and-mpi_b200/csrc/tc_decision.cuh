// tc_decision.cuh -- the batched decision function on the 5th-generation tensor cores (tcgen05 + TMEM).
//
// The one GEMM-shaped piece of the path: scores = Q [m x d] . SV^T [d x n_sv] followed by exp and a weighted row
// sum (code/main3.cpp:391-402 over all test rows at once; the reference GPU's predict kernel,
// code/gpu_svm_main3.cu:277-293, walks ALL n training rows per test row in scalar FP64).
//
// One CTA owns 128 query rows (the M = 128 lanes of its TMEM accumulator) and walks the support vectors in tiles of
// 128 (N = 128 accumulator columns).  Operands are staged in shared memory in the canonical K-major, no-swizzle
// core-matrix layout and multiplied with tcgen05.mma kind::tf32; a tf32 operand keeps 10 mantissa bits, so every
// fp32 value x is split into hi = x with the low 13 mantissa bits cleared and lo = x - hi (exact), and each
// K step issues three MMAs: hi.hi + hi.lo + lo.hi ("TF32x3", fp32 accumulation in TMEM).  The epilogue reads the
// accumulator with tcgen05.ld, forms K = exp(-gamma (|q|^2 + |s|^2 - 2 q.s)) and adds coef_k K to the row's FP64
// sum.  Scores from this kernel carry an error of ~1e-3 at most (fp32 accumulation over d = 784 and 3 tf32
// products); the caller re-evaluates every row whose |score| is below a threshold with the exact kernel
// (decision_kernel), so the predicted labels -- and the accuracy -- are those of the exact path.
#pragma once
#include "device_math.cuh"

namespace svm {

constexpr int TC_M = 128;          // query rows per CTA = TMEM lanes
constexpr int TC_N = 128;          // support vectors per tile = TMEM columns (fp32)
constexpr int TC_KB = 32;          // K elements staged per step (8 core-matrix columns of 4 tf32 each)
constexpr int TC_THREADS = 256;    // 8 warps stage operands; warps 0-3 also drain TMEM (warp w reads lanes 32w .. 32w+31)
constexpr int TC_TILE_BYTES = TC_M * TC_KB * 4;   // 16 KB per operand tile
constexpr int TC_STAGE_BYTES = 4 * TC_TILE_BYTES; // a_hi, a_lo, b_hi, b_lo

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

// stage one operand tile: rows [row0, row0+128) x K [k0, k0+32) of a row-major fp32 matrix with `pitch` floats per row,
// as hi (tf32-exact part) and lo (the rest) in the core-matrix layout; rows >= nrows and columns >= kmax are zero.
// A lane takes one row and one 32-byte sector of it (two 16-byte core-matrix rows), a warp 32 consecutive rows: the
// eight lanes of every quarter-warp hit eight different rows of a core matrix, so the 16-byte stores are conflict-free.
__device__ __forceinline__ void stage_tile(const float *__restrict__ src, long long row0, long long nrows, int pitch, int k0, int kmax,
                                           unsigned char *hi, unsigned char *lo)
{
#pragma unroll
    for (int i = 0; i < (TC_M * TC_KB / 8) / TC_THREADS; ++i) {       // 512 sectors per tile
        const int id = threadIdx.x + TC_THREADS * i;
        const int lane = id & 31, combo = id >> 5;
        const int r = (combo >> 2) * 32 + lane, c = (combo & 3) * 2;
        const int k = k0 + 4 * c;
        float4 v0 = make_float4(0.f, 0.f, 0.f, 0.f), v1 = v0;
        if (row0 + r < nrows) {
            const float4 *p = reinterpret_cast<const float4 *>(src + (size_t)(row0 + r) * pitch + k);
            if (k < kmax) v0 = __ldg(p);
            if (k + 4 < kmax) v1 = __ldg(p + 1);
        }
        const int off = (r >> 3) * 1024 + c * 128 + (r & 7) * 16;
        split_store(v0, hi, lo, off);
        split_store(v1, hi, lo, off + 128);
    }
}

// out[i] = init + sum_k coef[k] * exp(-gamma * max(0, qn[i] + svn[k] - 2 q_i . sv_k))   (approximate: see above)
// dots (optional, testing): raw q.s for the first tile of CTA 0, 128 x 128 row-major
// Two operand buffers: while the tensor core works through the MMAs of one K step, all warps stage the next one.
__global__ void __launch_bounds__(TC_THREADS, 1)
decision_tc_kernel(const float *__restrict__ Xq, const double *__restrict__ qn, long long m, const float *__restrict__ SV,
                   const double *__restrict__ svn, const double *__restrict__ coef, int nsv, int pitch, int d, double gamma,
                   double init, double *__restrict__ out, float *__restrict__ dots)
{
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    uint64_t *done = reinterpret_cast<uint64_t *>(smem_raw + 2 * TC_STAGE_BYTES);   // done[b]: the MMAs that read buffer b have finished
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(done + 2);
    const int warp = threadIdx.x >> 5;
    const long long row0 = (long long)blockIdx.x * TC_M;

    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(TC_N) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    if (threadIdx.x == 0) {
        mbar_init(&done[0], 1);
        mbar_init(&done[1], 1);
        mbar_fence_init();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t idesc = umma_idesc_tf32(TC_M, TC_N);

    const long long my_row = row0 + threadIdx.x;                     // thread t < 128 owns TMEM lane t = query row t of the tile
    const double my_qn = (threadIdx.x < TC_M && my_row < m) ? qn[my_row] : 0.0;
    double sum = init;
    const int kblocks = (d + TC_KB - 1) / TC_KB;
    unsigned it = 0;                                                 // K steps issued so far; step `it` uses buffer it & 1

    for (int n0 = 0; n0 < nsv; n0 += TC_N) {
        for (int kb = 0; kb < kblocks; ++kb, ++it) {
            unsigned char *buf = smem_raw + (it & 1) * TC_STAGE_BYTES;
            if (it >= 2) mbar_wait(&done[it & 1], ((it >> 1) - 1) & 1);   // the step that last used this buffer has been consumed
            stage_tile(Xq, row0, m, pitch, kb * TC_KB, d, buf, buf + TC_TILE_BYTES);
            stage_tile(SV, n0, nsv, pitch, kb * TC_KB, d, buf + 2 * TC_TILE_BYTES, buf + 3 * TC_TILE_BYTES);
            fence_proxy_async();                                     // generic-proxy stores -> visible to the tensor core
            __syncthreads();
            if (threadIdx.x == 0) {
                tc_fence_after();
                const uint32_t a_hi = smem_u32(buf), a_lo = a_hi + TC_TILE_BYTES, b_hi = a_lo + TC_TILE_BYTES, b_lo = b_hi + TC_TILE_BYTES;
#pragma unroll
                for (int ks = 0; ks < TC_KB / 8; ++ks) {             // one tf32 MMA covers K = 8 = two 16-byte core-matrix columns
                    const uint32_t koff = ks * 2 * 128;
                    const uint64_t ah = umma_desc_kmajor_noswizzle(a_hi + koff), al = umma_desc_kmajor_noswizzle(a_lo + koff);
                    const uint64_t bh = umma_desc_kmajor_noswizzle(b_hi + koff), bl = umma_desc_kmajor_noswizzle(b_lo + koff);
                    umma_tf32(tmem_base, ah, bh, idesc, (kb | ks) ? 1u : 0u);
                    umma_tf32(tmem_base, ah, bl, idesc, 1u);
                    umma_tf32(tmem_base, al, bh, idesc, 1u);
                }
                umma_commit(&done[it & 1]);                          // arrives when the MMAs above have read shared memory and written TMEM
            }
        }
        // the tile's last K step (and, the MMAs being in order, every earlier one) has landed in TMEM
        mbar_wait(&done[(it - 1) & 1], ((it - 1) >> 1) & 1);
        tc_fence_after();
        // epilogue: thread t reads its row of the 128 x 128 accumulator, 32 columns at a time
        if (threadIdx.x < TC_M) {
#pragma unroll 1
            for (int c0 = 0; c0 < TC_N; c0 += 32) {
                uint32_t v[32];
                const uint32_t taddr = tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0;
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
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    const int k = n0 + c0 + j;
                    const float dot = __uint_as_float(v[j]);
                    if (dots && blockIdx.x == 0 && n0 == 0) dots[(size_t)threadIdx.x * TC_N + c0 + j] = dot;
                    if (k < nsv && my_row < m) sum = sum + coef[k] * svm_exp(-gamma * sqdist_from_dot(my_qn, svn[k], (double)dot));
                }
            }
        }
        tc_fence_before();
        __syncthreads();                                             // the accumulator is drained before the next tile overwrites it
        tc_fence_after();
    }
    if (threadIdx.x < TC_M && my_row < m) out[my_row] = sum;
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TC_N) : "memory");
}

// ------------------------------------------------------------------------------------------------------------------
// The pipelined form: operands are split and laid out ONCE in global memory (tc_pack_kernel), so that a K step of a
// tile is 32 KB contiguous (hi tile then lo tile, already in the core-matrix layout) and reaches shared memory by one
// bulk copy (TMA) per operand.  Warp-specialised: warp 0 = producer (bulk copies into a 3-stage ring), warp 1 = MMA issuer
// (tcgen05.mma into one of two 128-column TMEM accumulators), warps 2-5 = epilogue (tcgen05.ld of the other accumulator,
// exp, weighted row sum) -- copies, tensor core and epilogue all overlap.
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

__global__ void __launch_bounds__(TCP_THREADS, 1)
decision_tc_pipelined_kernel(const unsigned char *__restrict__ qP, const double *__restrict__ qn, long long m,
                             const unsigned char *__restrict__ svP, const double *__restrict__ svn, const double *__restrict__ coef,
                             int nsv, int kblocks, double gamma, double init, double *__restrict__ out, float *__restrict__ dots)
{
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    unsigned char *stages = smem_raw;                                                    // [3][A panel 32 KB | B panel 32 KB]
    uint64_t *full = reinterpret_cast<uint64_t *>(stages + TCP_STAGES * 2 * TCP_PANEL_BYTES);
    uint64_t *empty = full + TCP_STAGES, *tfull = empty + TCP_STAGES, *tempty = tfull + 2;
    uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(tempty + 2);
    float *tile_coef = reinterpret_cast<float *>(tmem_slot + 4), *tile_svn = tile_coef + 2 * TC_N;   // per accumulator: coef and |s|^2 of the tile
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int ntiles = (nsv + TC_N - 1) / TC_N;

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
            for (int n = 0; n < ntiles; ++n)
                for (int kb = 0; kb < kblocks; ++kb, ++it) {
                    const int s = it % TCP_STAGES;
                    const unsigned u = it / TCP_STAGES;
                    if (u >= 1) mbar_wait(&empty[s], (u - 1) & 1);
                    unsigned char *dstA = stages + (size_t)s * 2 * TCP_PANEL_BYTES, *dstB = dstA + TCP_PANEL_BYTES;
                    mbar_expect_tx(&full[s], 2u * TCP_PANEL_BYTES);
                    bulk_g2s(dstA, qP + ((size_t)blockIdx.x * kblocks + kb) * TCP_PANEL_BYTES, TCP_PANEL_BYTES, &full[s], L2_EVICT_LAST);
                    bulk_g2s(dstB, svP + ((size_t)n * kblocks + kb) * TCP_PANEL_BYTES, TCP_PANEL_BYTES, &full[s], L2_EVICT_LAST);
                }
        }
    } else if (warp == 1) {
        // ---------------- MMA issuer
        const uint32_t idesc = umma_idesc_tf32(TC_M, TC_N);
        unsigned it = 0;
        for (int n = 0; n < ntiles; ++n) {
            const int acc = n & 1;
            const unsigned v = (unsigned)n >> 1;
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
        const long long my_row = (long long)blockIdx.x * TC_M + row_in_tile;
        const float my_qn = my_row < m ? (float)qn[my_row] : 0.f;
        const float fgamma = (float)gamma;
        const int et = threadIdx.x - 64;                                       // 0..127 among the epilogue threads
        double sum = init;
        for (int n = 0; n < ntiles; ++n) {
            const int acc = n & 1;
            const int n0 = n * TC_N;
            // this tile's coefficients and norms (the accumulator-specific half of the arrays: the MMA of tile n+1 is already running)
            const int k_mine = n0 + et;
            tile_coef[acc * TC_N + et] = k_mine < nsv ? (float)coef[k_mine] : 0.f;
            tile_svn[acc * TC_N + et] = k_mine < nsv ? (float)svn[k_mine] : 0.f;
            asm volatile("bar.sync 1, 128;" ::: "memory");                   // the four epilogue warps only
            mbar_wait(&tfull[acc], ((unsigned)n >> 1) & 1);
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
                    if (dots && blockIdx.x == 0 && n == 0) dots[(size_t)row_in_tile * TC_N + c0 + j] = dot;
                    const float d2 = fmaxf(fmaf(-2.f, dot, my_qn + tile_svn[acc * TC_N + c0 + j]), 0.f);
                    part = fmaf(tile_coef[acc * TC_N + c0 + j], __expf(-fgamma * d2), part);
                }
                sum = sum + (double)part;
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tempty[acc]);                          // this warp is done with the accumulator
        }
        if (my_row < m) out[my_row] = sum;
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(2 * TC_N) : "memory");
}

inline size_t decision_tc_pipelined_smem_bytes() { return (size_t)TCP_STAGES * 2 * TCP_PANEL_BYTES + 256 + 4 * TC_N * sizeof(float) + 1024; }

inline size_t decision_tc_smem_bytes() { return (size_t)2 * TC_STAGE_BYTES + 64 + 1024; }

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
