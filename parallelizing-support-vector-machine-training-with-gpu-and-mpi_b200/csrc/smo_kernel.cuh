// smo_kernel.cuh -- the persistent SMO kernel: one launch runs the whole training loop.
//
// Replaces the reference's SMO_train loop (code/main3.cpp:200-288) and, on the reference GPU, the
// <=9 launches + 11 blocking 8-byte copies + 2 device syncs it spends per iteration
// (code/gpu_svm_main3.cu:363-467).  One cooperative grid (one CTA per SM) stays resident:
//
//   per iteration, every CTA
//     1. knows the working pair (i_high, i_low) and the pair's (f, alpha, y) from the previous
//        record exchange, and redoes the scalar update redundantly (main3.cpp:235-270): K12, eta,
//        U/V, clipped alpha_low', alpha_high', the two f coefficients -- identical bits everywhere,
//        so no broadcast is needed;
//     2. X pass (a warp per row): the row arrives in shared memory by bulk copy (TMA, one 3136-byte
//        copy per row into a warp-private ring that keeps running across iterations), both squared
//        distances to the pivot rows held in registers are formed in one read of the row
//        (main3.cpp:221-222,230-231), f[j] is updated in the same pass (:271-275), and the row
//        joins the warp's running (value, lowest-index) argmin over I_high / argmax over I_low
//        (:107-142) with the pair's new alphas already applied (:278-279);
//        -- or, when both kernel rows are in the HBM row cache, a thread-per-row pass over the two
//        cached rows instead of X;
//     3. reduces its candidates, publishes one 64-byte record with a release store, and acquires
//        all other CTAs' records (an all-gather through L2: the only grid-wide synchronisation).
//
// All mutable state (f, alpha, cache slots, cache bookkeeping) of row j is touched only by the CTA
// that owns row j, so the record exchange is the only cross-CTA communication.
#pragma once
#include "device_math.cuh"

namespace svm {

constexpr int SMO_THREADS = 512;          // 16 warps
constexpr int SMO_WARPS = SMO_THREADS / 32;
constexpr int SMO_KSLOTS_MAX = 4;         // ring depth per warp: 16 x 4 x 3136 B = 196 KB in flight per SM

struct __align__(16) Cand {               // one side of a record: 32 bytes
    double f;
    double alpha;
    long long q;                          // row-cache allocation number of the row, -1 = never cached
    int idx;
    int y;
};

struct __align__(128) Rec {               // one CTA's (or one GPU's) candidates for one exchange
    Cand lo;                              // argmin f over I_high  -> i_high
    Cand hi;                              // argmax f over I_low   -> i_low
    int seq;                              // exchange number + 1, written last with release
    int pad[15];
};

struct SmoResult {
    int num_iter;
    int stop_reason;
    long long i_high, i_low;
    double b_high, b_low;
    long long x_passes;
    long long cache_hits;
    long long cache_counter;
};

struct SmoArgs {
    const float *X;        // [n][pitch4*4] fp32, zero padded
    const int *y;          // +-1
    double *f;
    double *alpha;
    long long *cacheq;     // per row: allocation number of its cached kernel row, or -1
    double *cacheK;        // [cache_cap][n] FP64 kernel rows
    Rec *rec;              // [2][grid]
    int *trace;            // [trace_cap][2]
    SmoResult *result;
    long long n;
    int pitch4;            // float4 groups per row
    int kslots;            // ring depth per warp (2..SMO_KSLOTS_MAX), chosen so the ring fits in shared memory
    int cache_cap;
    int trace_cap;
    int max_iter;
    double C, gamma, eps, tau;
    unsigned long long policy_stream;  // L2 policy for the streamed part of every CTA's rows
    unsigned long long policy_keep;    // L2 policy for the leading keep_frac of every warp's rows
    int keep_per_1024;                 // rows (per 1024) of each warp's range loaded with policy_keep
    // ---- multi-GPU row sharding (peers == 1: single GPU) ----
    int peers;                          // GPUs sharing the rows
    int rank;                           // this GPU
    long long row_begin, row_end;       // rows owned by this GPU (X is replicated, state is global-sized)
    Rec *mailbox;                       // [2][peers] local mailbox written by peers
    Rec *peer_mailbox[8];               // peers' mailboxes (peer-mapped device pointers)
};

// ------------------------------------------------------------------------------------------
template <int Q>
struct Pivots {
    float4 h[Q];
    float4 l[Q];
};

template <int Q, int CH>
__device__ __forceinline__ void row_sqdist2(const float4 *__restrict__ row, const Pivots<Q> &pv, int pitch4, int lane,
                                            double &dh, double &dl)
{
    double ah = 0.0, al = 0.0;
#pragma unroll
    for (int i = 0; i < Q; ++i) {
        const int q = i * 32 + lane;
        if (q < pitch4) {
            const float4 v = row[q];
            accum_sq<CH>(v, pv.h[i], ah);
            accum_sq<CH>(v, pv.l[i], al);
        }
    }
    dh = warp_sum_butterfly(ah);
    dl = warp_sum_butterfly(al);
}

// membership tests of calc_i_high / calc_i_low (code/main3.cpp:115,134)
__device__ __forceinline__ bool in_I_high(int y, double a, double C, double eps)
{
    return (y == 1 && a < C - eps) || (y == -1 && a > 0.0 + eps);
}
__device__ __forceinline__ bool in_I_low(int y, double a, double C, double eps)
{
    return (y == 1 && a > 0.0 + eps) || (y == -1 && a < C - eps);
}

struct LocalBest {
    double fmin; int imin;
    double fmax; int imax;
};
constexpr int IDX_NONE = 0x7fffffff;

__device__ __forceinline__ void best_init(LocalBest &b)
{
    b.fmin = __longlong_as_double(0x7ff0000000000000LL);
    b.imin = IDX_NONE;
    b.fmax = __longlong_as_double(0xfff0000000000000LL);
    b.imax = IDX_NONE;
}
__device__ __forceinline__ void best_merge(LocalBest &b, double fmin, int imin, double fmax, int imax)
{
    if (better_min(fmin, imin, b.fmin, b.imin)) { b.fmin = fmin; b.imin = imin; }
    if (better_max(fmax, imax, b.fmax, b.imax)) { b.fmax = fmax; b.imax = imax; }
}
__device__ __forceinline__ void best_warp_reduce(LocalBest &b)
{
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) {
        const double f1 = __shfl_xor_sync(FULL, b.fmin, off);
        const int i1 = __shfl_xor_sync(FULL, b.imin, off);
        const double f2 = __shfl_xor_sync(FULL, b.fmax, off);
        const int i2 = __shfl_xor_sync(FULL, b.imax, off);
        best_merge(b, f1, i1, f2, i2);
    }
}

struct __align__(16) Shared {
    Cand sel_lo, sel_hi;                  // the grid-wide winners of the last exchange
    double wfmin[SMO_WARPS], wfmax[SMO_WARPS];
    int wimin[SMO_WARPS], wimax[SMO_WARPS];
};

// One exchange: CTA-level reduce of the per-thread candidates, publish, gather, broadcast in smem.
// Returns with sh.sel_lo / sh.sel_hi valid for every thread of the CTA.
__device__ __forceinline__ void exchange(const SmoArgs &a, Shared &sh, LocalBest &mine, int gather_no)
{
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int G = gridDim.x;
    best_warp_reduce(mine);
    if (lane == 0) {
        sh.wfmin[warp] = mine.fmin; sh.wimin[warp] = mine.imin;
        sh.wfmax[warp] = mine.fmax; sh.wimax[warp] = mine.imax;
    }
    __syncthreads();
    if (warp == 0) {
        LocalBest b;
        best_init(b);
        if (lane < SMO_WARPS) { b.fmin = sh.wfmin[lane]; b.imin = sh.wimin[lane]; b.fmax = sh.wfmax[lane]; b.imax = sh.wimax[lane]; }
        best_warp_reduce(b);
        Rec *mine_rec = a.rec + (size_t)(gather_no & 1) * G + blockIdx.x;
        if (lane == 0) {
            Cand lo, hi;
            lo.f = b.fmin; lo.idx = b.imin; lo.alpha = 0.0; lo.y = 0; lo.q = -1;
            hi.f = b.fmax; hi.idx = b.imax; hi.alpha = 0.0; hi.y = 0; hi.q = -1;
            if (b.imin != IDX_NONE) { lo.alpha = a.alpha[b.imin]; lo.y = a.y[b.imin]; lo.q = a.cacheq ? a.cacheq[b.imin] : -1; }
            if (b.imax != IDX_NONE) { hi.alpha = a.alpha[b.imax]; hi.y = a.y[b.imax]; hi.q = a.cacheq ? a.cacheq[b.imax] : -1; }
            mine_rec->lo = lo;
            mine_rec->hi = hi;
            st_release_gpu(&mine_rec->seq, gather_no + 1);
        }
        // gather: lane l acquires the records of CTAs l, l+32, ...
        LocalBest g;
        best_init(g);
        int bmin = -1, bmax = -1;
        for (int c = lane; c < G; c += 32) {
            const Rec *r = a.rec + (size_t)(gather_no & 1) * G + c;
            while (ld_acquire_gpu(&r->seq) != gather_no + 1) { }
            const double f1 = r->lo.f; const int i1 = r->lo.idx;
            const double f2 = r->hi.f; const int i2 = r->hi.idx;
            if (better_min(f1, i1, g.fmin, g.imin)) { g.fmin = f1; g.imin = i1; bmin = c; }
            if (better_max(f2, i2, g.fmax, g.imax)) { g.fmax = f2; g.imax = i2; bmax = c; }
        }
#pragma unroll
        for (int off = 16; off >= 1; off >>= 1) {
            const double f1 = __shfl_xor_sync(FULL, g.fmin, off);
            const int i1 = __shfl_xor_sync(FULL, g.imin, off);
            const int c1 = __shfl_xor_sync(FULL, bmin, off);
            const double f2 = __shfl_xor_sync(FULL, g.fmax, off);
            const int i2 = __shfl_xor_sync(FULL, g.imax, off);
            const int c2 = __shfl_xor_sync(FULL, bmax, off);
            if (better_min(f1, i1, g.fmin, g.imin)) { g.fmin = f1; g.imin = i1; bmin = c1; }
            if (better_max(f2, i2, g.fmax, g.imax)) { g.fmax = f2; g.imax = i2; bmax = c2; }
        }
        if (lane == 0) {
            Cand lo, hi;
            lo.f = g.fmin; lo.idx = IDX_NONE; lo.alpha = 0.0; lo.y = 0; lo.q = -1;
            hi.f = g.fmax; hi.idx = IDX_NONE; hi.alpha = 0.0; hi.y = 0; hi.q = -1;
            if (bmin >= 0 && g.imin != IDX_NONE) lo = a.rec[(size_t)(gather_no & 1) * G + bmin].lo;
            if (bmax >= 0 && g.imax != IDX_NONE) hi = a.rec[(size_t)(gather_no & 1) * G + bmax].hi;
            sh.sel_lo = lo;
            sh.sel_hi = hi;
        }
    }
    __syncthreads();
}

// ------------------------------------------------------------------------------------------
template <int Q, int CH>
__global__ void __launch_bounds__(SMO_THREADS, 1) smo_persistent_kernel(const __grid_constant__ SmoArgs a)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int G = gridDim.x;
    const int row_bytes = a.pitch4 * 16;
    // shared-memory carve-up: [warps][slots] rows | mbarriers | Shared
    unsigned char *slots = smem_raw;
    const int KS = a.kslots;
    uint64_t *bars = reinterpret_cast<uint64_t *>(slots + (size_t)SMO_WARPS * KS * row_bytes);
    Shared &sh = *reinterpret_cast<Shared *>(bars + SMO_WARPS * SMO_KSLOTS_MAX);

    // rows owned by this CTA and, inside it, by this warp (contiguous ranges)
    const long long n_own = a.row_end - a.row_begin;
    const long long r0 = a.row_begin + (long long)blockIdx.x * n_own / G;
    const long long r1 = a.row_begin + (long long)(blockIdx.x + 1) * n_own / G;
    const long long R = r1 - r0;
    const long long w0 = r0 + (long long)warp * R / SMO_WARPS;
    const long long w1 = r0 + (long long)(warp + 1) * R / SMO_WARPS;
    const int m = (int)(w1 - w0);
    const int keep_rows = (int)(((long long)m * a.keep_per_1024) >> 10);

    unsigned char *my_slots = slots + (size_t)warp * KS * row_bytes;
    uint64_t *my_bars = bars + warp * SMO_KSLOTS_MAX;
    if (lane == 0) {
        for (int s = 0; s < KS; ++s) mbar_init(&my_bars[s], 1);
        mbar_fence_init();
    }
    __syncthreads();

    // the ring: rows of this warp's range are requested in order, cyclically, for ever
    unsigned issued = 0, consumed = 0;   // row copies requested / used so far (mod 2^32 is fine)
    int pf_row = 0;                      // next row of the range to request
    auto issue_one = [&]() {             // called by the whole warp (counters stay warp-uniform); lane 0 issues
        if (lane == 0) {
            const int s = issued % KS;
            mbar_expect_tx(&my_bars[s], (unsigned)row_bytes);
            bulk_g2s(my_slots + (size_t)s * row_bytes, a.X + (size_t)(w0 + pf_row) * a.pitch4 * 4, (unsigned)row_bytes,
                     &my_bars[s], pf_row < keep_rows ? a.policy_keep : a.policy_stream);
        }
        ++issued;
        if (++pf_row == m) pf_row = 0;
    };
    if (m > 0)
        for (int s = 0; s < KS; ++s) issue_one();

    // ---- exchange 0: selection over the initial state (main3.cpp:203-204 in the first iteration)
    LocalBest mine;
    best_init(mine);
    for (long long j = r0 + threadIdx.x; j < r1; j += SMO_THREADS) {
        const double fj = a.f[j], aj = a.alpha[j];
        const int yj = a.y[j];
        if (in_I_high(yj, aj, a.C, a.eps) && fj < mine.fmin) { mine.fmin = fj; mine.imin = (int)j; }
        if (in_I_low(yj, aj, a.C, a.eps) && fj > mine.fmax) { mine.fmax = fj; mine.imax = (int)j; }
    }
    int gather_no = 0;
    exchange(a, sh, mine, gather_no);

    int num_iter = 1;                    // main3.cpp:197
    int stop = 0;
    double b_high = 0.0, b_low = 0.0;    // uninitialised in the reference (main3.cpp:188-189)
    long long i_high = a.n, i_low = a.n;
    long long cache_counter = 0;         // replicated: every CTA sees the same hit/miss sequence
    long long x_passes = 0, cache_hits = 0;

    for (;;) {
        const Cand lo = sh.sel_lo, hi = sh.sel_hi;
        i_high = lo.idx == IDX_NONE ? a.n : lo.idx;
        i_low = hi.idx == IDX_NONE ? a.n : hi.idx;
        if (lo.idx == IDX_NONE || hi.idx == IDX_NONE) { stop = 1; break; }          // main3.cpp:205
        b_high = lo.f;
        b_low = hi.f;
        if (b_low <= b_high + 2.0 * a.tau) { stop = 0; break; }                     // main3.cpp:213

        // row-cache lookup (FIFO by allocation number; validity needs no notification)
        bool hit_h = false, hit_l = false;
        long long q_h = lo.q, q_l = hi.q;
        if (a.cache_cap > 0) {
            hit_h = q_h >= 0 && cache_counter - q_h < a.cache_cap;
            hit_l = q_l >= 0 && cache_counter - q_l < a.cache_cap;
        }
        const bool use_cache = hit_h && hit_l;

        // K12 = K(i_high, i_low)  (main3.cpp:241)
        Pivots<Q> pv;
        double K12;
        if (!use_cache) {
            const float4 *xh = reinterpret_cast<const float4 *>(a.X) + (size_t)i_high * a.pitch4;
            const float4 *xl = reinterpret_cast<const float4 *>(a.X) + (size_t)i_low * a.pitch4;
#pragma unroll
            for (int i = 0; i < Q; ++i) {
                const int q = i * 32 + lane;
                pv.h[i] = q < a.pitch4 ? __ldg(xh + q) : make_float4(0.f, 0.f, 0.f, 0.f);
                pv.l[i] = q < a.pitch4 ? __ldg(xl + q) : make_float4(0.f, 0.f, 0.f, 0.f);
            }
            double acc = 0.0;
#pragma unroll
            for (int i = 0; i < Q; ++i)
                if (i * 32 + lane < a.pitch4) accum_sq<CH>(pv.l[i], pv.h[i], acc);
            acc = warp_sum_butterfly(acc);
            K12 = svm_exp(-a.gamma * acc);
        } else {
            K12 = a.cacheK[(size_t)(q_h % a.cache_cap) * a.n + i_low];
        }

        // scalar update, main3.cpp:235-270, operation for operation
        const int s = lo.y * hi.y;
        const double K11 = 1.0, K22 = 1.0;                    // exp(-gamma * 0)
        const double eta = K11 + K22 - 2.0 * K12;
        const double a_h = lo.alpha, a_l = hi.alpha;
        double U, V;
        if (s == -1) {
            const double u = a_l - a_h, v = a.C + a_l - a_h;
            U = (0.0 < u) ? u : 0.0;
            V = (v < a.C) ? v : a.C;
        } else {
            const double u = a_l + a_h - a.C, v = a_l + a_h;
            U = (0.0 < u) ? u : 0.0;
            V = (v < a.C) ? v : a.C;
        }
        if (!(U <= V + 1e-12)) { stop = 2; break; }            // main3.cpp:246
        if (eta <= a.eps) { stop = 3; break; }                  // main3.cpp:253
        double na_l = a_l + hi.y * (b_high - b_low) / eta;      // main3.cpp:259
        if (na_l > V) na_l = V;
        if (na_l < U) na_l = U;
        const double na_h = a_h + s * (a_l - na_l);             // main3.cpp:266
        const double c_h = (na_h - a_h) * lo.y;                 // delta_alpha_i_high * y[i_high]
        const double c_l = (na_l - a_l) * hi.y;

        if (blockIdx.x == 0 && threadIdx.x == 0 && a.rank == 0 && num_iter - 1 < a.trace_cap) {
            a.trace[2 * (num_iter - 1)] = (int)i_high;
            a.trace[2 * (num_iter - 1) + 1] = (int)i_low;
        }

        // allocate cache slots for the rows about to be computed
        int slot_h = -1, slot_l = -1;
        if (a.cache_cap > 0) {
            if (!use_cache) {
                if (!hit_h) q_h = cache_counter++;
                if (!hit_l) q_l = cache_counter++;
            }
            slot_h = (int)(q_h % a.cache_cap);
            slot_l = (int)(q_l % a.cache_cap);
        }

        best_init(mine);
        if (!use_cache) {
            // ---------------- X pass: a warp per row ----------------
            ++x_passes;
            for (int base = 0; base < m; base += 32) {
                const int cnt = min(32, m - base);
                // lane t finishes row base+t: its state is requested now and used after the group
                const long long j = w0 + base + lane;
                double fj = 0.0, aj = 0.0;
                int yj = 0;
                if (lane < cnt) { fj = a.f[j]; aj = a.alpha[j]; yj = a.y[j]; }
                double my_dh = 0.0, my_dl = 0.0;
                for (int t = 0; t < cnt; ++t) {
                    const int sidx = consumed % KS;
                    mbar_wait(&my_bars[sidx], (consumed / KS) & 1);
                    double dh, dl;
                    row_sqdist2<Q, CH>(reinterpret_cast<const float4 *>(my_slots + (size_t)sidx * row_bytes), pv, a.pitch4, lane, dh, dl);
                    ++consumed;
                    if (lane == t) { my_dh = dh; my_dl = dl; }
                    // the slot is free again: request the next row of the ring into it
                    __syncwarp();
                    fence_proxy_async();
                    issue_one();
                }
                if (lane < cnt) {
                    const double kh = svm_exp(-a.gamma * my_dh);
                    const double kl = svm_exp(-a.gamma * my_dl);
                    fj = fj + (c_h * kh + c_l * kl);            // main3.cpp:274
                    a.f[j] = fj;
                    if (slot_h >= 0 && !hit_h) a.cacheK[(size_t)slot_h * a.n + j] = kh;
                    if (slot_l >= 0 && !hit_l) a.cacheK[(size_t)slot_l * a.n + j] = kl;
                    if (j == i_high) { aj = na_h; a.alpha[j] = aj; if (a.cacheq && !hit_h) a.cacheq[j] = q_h; }
                    if (j == i_low) { aj = na_l; a.alpha[j] = aj; if (a.cacheq && !hit_l) a.cacheq[j] = q_l; }
                    if (in_I_high(yj, aj, a.C, a.eps) && fj < mine.fmin) { mine.fmin = fj; mine.imin = (int)j; }
                    if (in_I_low(yj, aj, a.C, a.eps) && fj > mine.fmax) { mine.fmax = fj; mine.imax = (int)j; }
                }
            }
        } else {
            // ---------------- cached pass: a thread per row over two cached kernel rows ----------------
            ++cache_hits;
            const double *kh_row = a.cacheK + (size_t)slot_h * a.n;
            const double *kl_row = a.cacheK + (size_t)slot_l * a.n;
            for (long long j = r0 + threadIdx.x; j < r1; j += SMO_THREADS) {
                double fj = a.f[j], aj = a.alpha[j];
                const int yj = a.y[j];
                fj = fj + (c_h * kh_row[j] + c_l * kl_row[j]);
                a.f[j] = fj;
                if (j == i_high) { aj = na_h; a.alpha[j] = aj; }
                if (j == i_low) { aj = na_l; a.alpha[j] = aj; }
                if (in_I_high(yj, aj, a.C, a.eps) && fj < mine.fmin) { mine.fmin = fj; mine.imin = (int)j; }
                if (in_I_low(yj, aj, a.C, a.eps) && fj > mine.fmax) { mine.fmax = fj; mine.imax = (int)j; }
            }
        }

        ++gather_no;
        exchange(a, sh, mine, gather_no);

        ++num_iter;                                              // main3.cpp:281-287
        if (num_iter > a.max_iter) { stop = 4; break; }
    }

    // drain the ring: every requested copy must land before the CTA may exit
    while (consumed != issued) {
        mbar_wait(&my_bars[consumed % KS], (consumed / KS) & 1);
        ++consumed;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        SmoResult r;
        r.num_iter = num_iter; r.stop_reason = stop;
        r.i_high = i_high; r.i_low = i_low;
        r.b_high = b_high; r.b_low = b_low;
        r.x_passes = x_passes; r.cache_hits = cache_hits; r.cache_counter = cache_counter;
        *a.result = r;
    }
}

inline size_t smo_smem_bytes(int pitch4, int kslots)
{
    return (size_t)SMO_WARPS * kslots * pitch4 * 16 + sizeof(uint64_t) * SMO_WARPS * SMO_KSLOTS_MAX + sizeof(Shared) + 128;
}

} // namespace svm
