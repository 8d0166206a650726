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
//     2. X pass (a warp per group of four consecutive rows; two for the longest rows, one in the one-row variants): the
//        group arrives in shared memory by ONE bulk copy (TMA) into a warp-private ring that keeps running across
//        iterations, both dot products (FP64 rows: squared distances) with the pivot rows held in registers are formed
//        in one read of each row (main3.cpp:221-222,230-231), the 2 x 4 warp sums by one halving exchange, f[j] is
//        updated in the same pass (:271-275), and the row joins the warp's running (value, lowest-index) argmin over
//        I_high / argmax over I_low (:107-142) with the pair's new alphas already applied (:278-279).  A share of X is
//        loaded with an evict-last policy and stays in L2 from one pass to the next;
//        -- or, when both kernel rows are in the HBM row cache, a thread-per-row pass over the two
//        cached rows instead of X;
//     3. reduces its candidates, publishes one 128-byte self-validating record, adds 1 to an arrival counter, and
//        reads all other CTAs' records once the count is complete (an all-gather through L2: the only grid-wide
//        synchronisation).
//
// All mutable state (f, alpha, cache slots, cache bookkeeping) of row j is touched only by the CTA
// that owns row j, so the record exchange is the only cross-CTA communication.
#pragma once
#include <type_traits>
#include "device_math.cuh"

namespace svm {

constexpr int SMO_MAX_WARPS = 16;         // the CTA size is a template parameter (256 / 384 / 512 threads)
constexpr int SMO_KSLOTS_MAX = 8;         // ring depth per warp, sized so that warps x depth x row fills shared memory

struct Rec;   // one CTA's candidates for one exchange; defined with the exchange below

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
    const void *X;         // [n][pitch4] 16-byte groups: 4 fp32 or 2 FP64 values each, zero padded
    const double *sqn;     // [n] squared norms |x_j|^2 (dot product of the row with itself)
    const int *y;          // +-1
    double *f;
    double *alpha;
    long long *cacheq;     // per row: allocation number of its cached kernel row, or -1
    double *cacheK;        // [cache_cap][n] FP64 kernel rows
    Rec *rec;              // [2][grid]
    unsigned *arrive;      // arrival counter of the record exchange (zeroed before the launch)
    int *trace;            // [trace_cap][2]
    SmoResult *result;
    long long n;
    int pitch4;            // 16-byte groups per row
    int kslots;            // ring depth per warp (2..SMO_KSLOTS_MAX), chosen so the ring fits in shared memory
    int cache_cap;
    long long cache_counter0;  // allocation number to start from (> 0 when the cache is kept from an earlier run on the same X)
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
    Rec *peer_mailbox[8];               // peers' mailboxes (peer-mapped device pointers; [rank] is the local one)
    int epoch;                          // launch number, folded into the mailbox words so that stale records never match
    long long *timing;                  // optional [iterations][8] clock64 stamps of CTA 0 (profiling builds only)
    int timing_cap;
};

// Rows a warp takes per step of the X pass.  1: a row at a time (two sums per butterfly).  4 (fp32 rows, the 8-warp
// CTA; 2 for the longest rows, Q = 8): the ring holds groups of four consecutive rows (one bulk copy, one mbarrier per group); four rows x two pivots =
// eight sums per halving exchange, eight independent accumulation chains.  2: the same with FP64 rows (Q = 13), whose
// rows are twice as long.
template <typename XT, int Q, int THREADS>
constexpr int smo_rows_per_step()
{
    return (sizeof(XT) == 4 && THREADS <= 256) ? (Q == 8 ? 2 : 4) : (sizeof(XT) == 8 && Q == 13 && THREADS == 256) ? 2 : 1;
}

// ------------------------------------------------------------------------------------------
template <typename Vec, int Q>
struct Pivots {
    Vec h[Q];
    Vec l[Q];
};

template <typename Vec, int Q, int CH>
__device__ __forceinline__ void row_dot2(const Vec *__restrict__ row, const Pivots<Vec, Q> &pv, int pitch4, int lane,
                                         double &dh, double &dl)
{
    double ah = 0.0, al = 0.0;
#pragma unroll
    for (int i = 0; i < Q; ++i) {
        const int q = i * 32 + lane;
        if (q < pitch4) {
            const Vec v = row[q];
            accum_pair<CH>(v, pv.h[i], ah);
            accum_pair<CH>(v, pv.l[i], al);
        }
    }
    warp_sum_butterfly2(ah, al, lane);
    dh = ah;
    dl = al;
}

// One full group of RPS rows against both pivots, NG = the number of 32-lane register groups the row really has
// (the kernel is instantiated for a class of row lengths, Q >= NG): groups 0..NG-2 are whole and sit at constant
// offsets from one address per row; the last is read at a clamped offset (fp32: the pivot registers are zero beyond the
// row, so such an element adds +-0) or skipped per lane (FP64 rows accumulate squared differences).
template <typename Vec, int Q, int CH, int RPS, int NG>
__device__ __forceinline__ void group_dot2(const unsigned char *grp, int row_bytes, int lane, int q_last, bool last_valid,
                                           const Pivots<Vec, Q> &pv, double (&P)[2 * RPS])
{
    constexpr bool NEUTRAL_PAD = sizeof(Vec) == sizeof(float4) && std::is_same<Vec, float4>::value;
#pragma unroll
    for (int r = 0; r < RPS; ++r) {
        const unsigned char *rowp = grp + (size_t)r * row_bytes;
        const Vec *lanep = reinterpret_cast<const Vec *>(rowp) + lane;
#pragma unroll
        for (int i = 0; i < NG - 1; ++i) {
            const Vec v = lanep[i * 32];
            accum_pair<CH>(v, pv.h[i], P[r]);
            accum_pair<CH>(v, pv.l[i], P[RPS + r]);
        }
        if (NEUTRAL_PAD || last_valid) {
            const Vec v = *reinterpret_cast<const Vec *>(rowp + q_last);
            accum_pair<CH>(v, pv.h[NG - 1], P[r]);
            accum_pair<CH>(v, pv.l[NG - 1], P[RPS + r]);
        }
    }
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

constexpr int IDX_NONE = 0x7fffffff;

// A candidate for one side of the working pair, carried with everything the scalar update needs
// from that row (main3.cpp:235,246,259 read y and alpha of the pair), so that no CTA ever has to
// read another CTA's rows of alpha / y / cache bookkeeping.
struct Best {
    double f;
    double alpha;
    long long q;      // row-cache allocation number of the row, -1 = never cached
    int idx;
    int y;
};
struct LocalBest {
    Best lo;          // argmin f over I_high  -> i_high
    Best hi;          // argmax f over I_low   -> i_low
};

__device__ __forceinline__ void best_init(LocalBest &b)
{
    b.lo.f = __longlong_as_double(0x7ff0000000000000LL);
    b.hi.f = __longlong_as_double(0xfff0000000000000LL);
    b.lo.idx = b.hi.idx = IDX_NONE;
    b.lo.alpha = b.hi.alpha = 0.0;
    b.lo.y = b.hi.y = 0;
    b.lo.q = b.hi.q = -1;
}
// a row joins the running argmin / argmax (ascending j per thread + strict compare = lowest index)
__device__ __forceinline__ void best_consider(LocalBest &b, long long j, double fj, double aj, int yj, long long qj,
                                              double C, double eps)
{
    if (in_I_high(yj, aj, C, eps) && fj < b.lo.f) { b.lo.f = fj; b.lo.idx = (int)j; b.lo.alpha = aj; b.lo.y = yj; b.lo.q = qj; }
    if (in_I_low(yj, aj, C, eps) && fj > b.hi.f) { b.hi.f = fj; b.hi.idx = (int)j; b.hi.alpha = aj; b.hi.y = yj; b.hi.q = qj; }
}
// Warp-wide (value, lowest index) winner with the hardware integer reduction (REDUX): the double is
// mapped to a 64-bit key whose unsigned order is the double's order, its two halves are reduced one
// after the other, and the row index breaks ties.  f + 0.0 folds -0.0 into +0.0 so that, as in the
// reference's "<" / ">" scans, the two zeros tie.  Every lane returns the winner with its payload.
__device__ __forceinline__ unsigned long long order_key(double f)
{
    const unsigned long long u = (unsigned long long)__double_as_longlong(f + 0.0);
    return (u >> 63) ? ~u : (u | 0x8000000000000000ull);
}
template <bool IS_MIN>
__device__ __forceinline__ void best_warp_one(Best &b)
{
    unsigned long long key = order_key(b.f);
    if (!IS_MIN) key = ~key;
    const unsigned hi = (unsigned)(key >> 32), lo = (unsigned)key;
    const unsigned mhi = __reduce_min_sync(FULL, hi);
    const unsigned mlo = __reduce_min_sync(FULL, hi == mhi ? lo : 0xffffffffu);
    const bool tie = hi == mhi && lo == mlo;
    const unsigned midx = __reduce_min_sync(FULL, tie ? (unsigned)b.idx : 0xffffffffu);
    const unsigned holders = __ballot_sync(FULL, tie && (unsigned)b.idx == midx);   // a row index lives in exactly one lane
    const int src = __ffs(holders) - 1;
    b.f = __shfl_sync(FULL, b.f, src);
    b.alpha = __shfl_sync(FULL, b.alpha, src);
    b.q = __shfl_sync(FULL, b.q, src);
    b.idx = __shfl_sync(FULL, b.idx, src);
    b.y = __shfl_sync(FULL, b.y, src);
}
__device__ __forceinline__ void best_warp_reduce(LocalBest &b)
{
    best_warp_one<true>(b.lo);
    best_warp_one<false>(b.hi);
}
__device__ __forceinline__ void best_merge(LocalBest &b, const Best &lo, const Best &hi)
{
    if (better_min(lo.f, lo.idx, b.lo.f, b.lo.idx)) b.lo = lo;
    if (better_max(hi.f, hi.idx, b.hi.f, b.hi.idx)) b.hi = hi;
}

// ---- the record exchange ---------------------------------------------------------------------
// Every CTA stores its record, then adds 1 to an arrival counter.  ONE thread per CTA spins on the
// counter (148 pollers of one 4-byte word; having all 148 x 148 record readers spin instead measured
// ~10 us per exchange in L2 queueing); when the count is complete thread t reads CTA t's record.
// No fence orders the record against the counter.  Instead a record is 16 words of 8 bytes, each
// {32 data bits, 32-bit exchange number}: an aligned 8-byte access is single-copy atomic, so a word
// that shows the right number also shows its data, and a reader that is too early simply re-reads.
constexpr int REC_WORDS = 16;
struct __align__(128) Rec { unsigned long long w[REC_WORDS]; };

__device__ __forceinline__ unsigned rec_word(const LocalBest &b, int k)
{
    switch (k) {
    case 0: return (unsigned)__double2loint(b.lo.f);
    case 1: return (unsigned)__double2hiint(b.lo.f);
    case 2: return (unsigned)b.lo.idx;
    case 3: return (unsigned)__double2loint(b.hi.f);
    case 4: return (unsigned)__double2hiint(b.hi.f);
    case 5: return (unsigned)b.hi.idx;
    case 6: return (unsigned)__double2loint(b.lo.alpha);
    case 7: return (unsigned)__double2hiint(b.lo.alpha);
    case 8: return (unsigned)__double2loint(b.hi.alpha);
    case 9: return (unsigned)__double2hiint(b.hi.alpha);
    case 10: return (unsigned)b.lo.y;
    case 11: return (unsigned)b.hi.y;
    case 12: return (unsigned)(b.lo.q & 0xffffffffLL);
    case 13: return (unsigned)((unsigned long long)b.lo.q >> 32);
    case 14: return (unsigned)(b.hi.q & 0xffffffffLL);
    default: return (unsigned)((unsigned long long)b.hi.q >> 32);
    }
}
__device__ __forceinline__ void rec_unpack(const unsigned long long (&w)[REC_WORDS], LocalBest &b)
{
    auto d = [&](int k) { return (unsigned)(w[k] & 0xffffffffull); };
    b.lo.f = __hiloint2double((int)d(1), (int)d(0));
    b.lo.idx = (int)d(2);
    b.hi.f = __hiloint2double((int)d(4), (int)d(3));
    b.hi.idx = (int)d(5);
    b.lo.alpha = __hiloint2double((int)d(7), (int)d(6));
    b.hi.alpha = __hiloint2double((int)d(9), (int)d(8));
    b.lo.y = (int)d(10);
    b.hi.y = (int)d(11);
    b.lo.q = (long long)(((unsigned long long)d(13) << 32) | d(12));
    b.hi.q = (long long)(((unsigned long long)d(15) << 32) | d(14));
}
template <bool SYS>
__device__ __forceinline__ void rec_store(Rec *r, const LocalBest &b, unsigned seq)
{
#pragma unroll
    for (int k = 0; k < REC_WORDS; k += 2) {
        const unsigned long long v0 = ((unsigned long long)seq << 32) | rec_word(b, k);
        const unsigned long long v1 = ((unsigned long long)seq << 32) | rec_word(b, k + 1);
        if (SYS) asm volatile("st.relaxed.sys.global.v2.u64 [%0], {%1, %2};" ::"l"(&r->w[k]), "l"(v0), "l"(v1) : "memory");
        else asm volatile("st.relaxed.gpu.global.v2.u64 [%0], {%1, %2};" ::"l"(&r->w[k]), "l"(v0), "l"(v1) : "memory");
    }
}
// read a record; re-read until all 16 words carry `seq`.  Returns false when `budget` polls pass
// without that (only the multi-GPU mailbox reads pass a finite budget: a peer process may have died).
template <bool SYS>
__device__ __forceinline__ bool rec_read(const Rec *r, unsigned seq, LocalBest &out, long long deadline_clock = 0)
{
    unsigned long long w[REC_WORDS];
    for (;;) {
#pragma unroll
        for (int k = 0; k < REC_WORDS; k += 2) {
            if (SYS) asm volatile("ld.relaxed.sys.global.v2.u64 {%0, %1}, [%2];" : "=l"(w[k]), "=l"(w[k + 1]) : "l"(&r->w[k]) : "memory");
            else asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];" : "=l"(w[k]), "=l"(w[k + 1]) : "l"(&r->w[k]) : "memory");
        }
        bool ok = true;
#pragma unroll
        for (int k = 0; k < REC_WORDS; ++k) ok = ok && ((unsigned)(w[k] >> 32) == seq);
        if (ok) break;
        if (deadline_clock && clock64() > deadline_clock) return false;
    }
    rec_unpack(w, out);
    return true;
}
__device__ __forceinline__ void red_relaxed_gpu_add(unsigned *p, unsigned v)
{
    asm volatile("red.relaxed.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned ld_relaxed_gpu_u32(const unsigned *p)
{
    unsigned v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

struct __align__(16) Shared {
    Best wlo[SMO_MAX_WARPS], whi[SMO_MAX_WARPS];  // per-warp winners of this CTA's rows
    Best glo[SMO_MAX_WARPS], ghi[SMO_MAX_WARPS];  // per-warp winners over the gathered records
    Best xlo, xhi;                                // winners over the GPUs' records (row sharding)
    int xok, xpad_[3];
};

#ifdef SVM_TIMING
#define SVM_STAMP(k) do { if (a.timing && blockIdx.x == 0 && threadIdx.x == 0 && gather_no < a.timing_cap) a.timing[(size_t)gather_no * 16 + (k)] = clock64(); } while (0)
__device__ __forceinline__ unsigned long long gtimer() { unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }
#define SVM_GSTAMP(k) do { if (a.timing && threadIdx.x == 0 && gather_no >= 100 && gather_no < 108) a.timing[(size_t)a.timing_cap * 16 + ((size_t)(gather_no - 100) * gridDim.x + blockIdx.x) * 4 + (k)] = (long long)gtimer(); } while (0)
#else
#define SVM_STAMP(k) do { } while (0)
#define SVM_GSTAMP(k) do { } while (0)
#endif

constexpr long long PEER_TIMEOUT_CLOCKS = 6000000000LL;   // ~3 s at 1.97 GHz: a peer GPU's process is gone

// One exchange: CTA-level reduce of the per-thread candidates, publish, gather; with row sharding over
// several GPUs a second level follows in which every GPU stores its record into every peer's mailbox
// over NVLink (peer-mapped memory, same self-validating words) and reads the others' from its own.
// On return every thread of every CTA on every GPU holds the same winners in `mine`.  Returns false
// if a peer never showed up.
__device__ __forceinline__ bool exchange(const SmoArgs &a, Shared &sh, LocalBest &mine, int gather_no)
{
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int G = gridDim.x;
    const int nwarps = (int)(blockDim.x >> 5);
    const unsigned seq = (unsigned)gather_no + 1u;
    best_warp_reduce(mine);
    if (lane == 0) { sh.wlo[warp] = mine.lo; sh.whi[warp] = mine.hi; }
    SVM_STAMP(3);
    __syncthreads();
    SVM_STAMP(4);
    Rec *recs = a.rec + (size_t)(gather_no & 1) * G;
    if (warp == 0) {
        LocalBest b;
        best_init(b);
        if (lane < nwarps) { b.lo = sh.wlo[lane]; b.hi = sh.whi[lane]; }
        SVM_STAMP(8);
        best_warp_reduce(b);
        SVM_STAMP(9);
        if (lane == 0) {
            rec_store<false>(recs + blockIdx.x, b, seq);
            SVM_STAMP(10);
            red_relaxed_gpu_add(a.arrive, 1u);
            SVM_STAMP(5);
            const unsigned target = (unsigned)G * seq;
            while ((int)(ld_relaxed_gpu_u32(a.arrive) - target) < 0) { }
            __threadfence();      // orders the cached-row reads after the arrivals observed here (pairs with the writers' fence)
            SVM_STAMP(6);
        }
    }
    __syncthreads();
    // thread t reads the record of CTA t
    const int gather_warps = (G + 31) >> 5;
    if (warp < gather_warps) {
        LocalBest g;
        best_init(g);
        SVM_STAMP(11);
        if ((int)threadIdx.x < G) rec_read<false>(recs + threadIdx.x, seq, g);
        SVM_STAMP(12);
        best_warp_reduce(g);
        if (lane == 0) { sh.glo[warp] = g.lo; sh.ghi[warp] = g.hi; }
        SVM_STAMP(13);
    }
    __syncthreads();
    SVM_STAMP(14);
    best_init(mine);
    if (lane < gather_warps) { mine.lo = sh.glo[lane]; mine.hi = sh.ghi[lane]; }
    best_warp_reduce(mine);
    SVM_STAMP(7);
    if (a.peers == 1) return true;

    // ---- second level: one record per GPU, all-gathered through the peers' mailboxes
    const unsigned seq2 = ((unsigned)(a.epoch & 0x7ff) << 21) | seq;
    const int slot = (gather_no & 1) * a.peers;
    if (blockIdx.x == 0 && threadIdx.x < a.peers)        // thread p stores this GPU's record at peer p (p == rank: local)
        rec_store<true>(a.peer_mailbox[threadIdx.x] + slot + a.rank, mine, seq2);
    bool ok = true;
    if (warp == 0) {
        LocalBest g;
        best_init(g);
        if (lane < a.peers) ok = rec_read<true>(a.mailbox + slot + lane, seq2, g, clock64() + PEER_TIMEOUT_CLOCKS);
        ok = __all_sync(FULL, ok);
        best_warp_reduce(g);
        if (lane == 0) { sh.xlo = g.lo; sh.xhi = g.hi; sh.xok = ok ? 1 : 0; }
    }
    __syncthreads();
    mine.lo = sh.xlo;
    mine.hi = sh.xhi;
    ok = sh.xok != 0;
    return ok;                                            // xlo/xhi/xok are next written two CTA barriers from here
}

// ------------------------------------------------------------------------------------------
template <typename XT, int Q, int CH, int THREADS>
__global__ void __launch_bounds__(THREADS, 1) smo_persistent_kernel(const __grid_constant__ SmoArgs a)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int G = gridDim.x;
    constexpr int WARPS = THREADS / 32;
    using Vec = typename RowVec<XT>::type;
    const unsigned char *Xb = static_cast<const unsigned char *>(a.X);
    const int row_bytes = a.pitch4 * 16;
    // shared-memory carve-up: [warps][slots] rows | mbarriers | Shared
    unsigned char *slots = smem_raw;
    const int KS = a.kslots;
    uint64_t *bars = reinterpret_cast<uint64_t *>(slots + (size_t)WARPS * KS * smo_rows_per_step<XT, Q, THREADS>() * row_bytes);
    Shared &sh = *reinterpret_cast<Shared *>(bars + WARPS * SMO_KSLOTS_MAX);
    constexpr int RPS = smo_rows_per_step<XT, Q, THREADS>();
    // (RPS == 4) per warp: the two sums of each of the 32 rows of a block, parked until their lane finishes the rows
    double2 *stage = reinterpret_cast<double2 *>(reinterpret_cast<unsigned char *>(&sh) + ((sizeof(Shared) + 15) & ~(size_t)15)) + warp * 32;


    // rows owned by this CTA and, inside it, by this warp (contiguous ranges)
    const long long n_own = a.row_end - a.row_begin;
    const long long r0 = a.row_begin + (long long)blockIdx.x * n_own / G;
    const long long r1 = a.row_begin + (long long)(blockIdx.x + 1) * n_own / G;
    const long long R = r1 - r0;
    // warps take contiguous ranges, cut at multiples of RPS rows from r0 (so only the CTA's last rows form a short group)
    const long long units = (R + RPS - 1) / RPS;
    const long long w0 = min(r1, r0 + RPS * ((long long)warp * units / WARPS));
    const long long w1 = min(r1, r0 + RPS * ((long long)(warp + 1) * units / WARPS));
    const int m = (int)(w1 - w0);
    const int keep_rows = (int)(((long long)m * a.keep_per_1024) >> 10);

    unsigned char *my_slots = slots + (size_t)warp * KS * RPS * row_bytes;
    uint64_t *my_bars = bars + warp * SMO_KSLOTS_MAX;
    if (lane == 0) {
        for (int s = 0; s < KS; ++s) mbar_init(&my_bars[s], 1);
        mbar_fence_init();
    }
    __syncthreads();

    // the ring: rows of this warp's range are requested in order, cyclically, for ever
    // Each slot is refilled right after it is used, so "the slot to use next" is the whole state.
    int cslot = 0;                       // slot holding the next row to use
    unsigned cphase = 0;                 // mbarrier phase parity of that slot
    int pf_row = 0;                      // next row of the range to request
    auto issue_into = [&](int s) {       // called by the whole warp (state stays warp-uniform); lane 0 issues
        const int nrows = RPS == 1 ? 1 : min(RPS, m - pf_row);       // a slot holds RPS consecutive rows (fewer at the end of the range)
        if (lane == 0) {
            // which rows stay in L2 (policy_keep): with one row per slot the leading share of the warp's range; with groups
            // an evenly interleaved share, so that L2 hits and HBM reads alternate through the whole pass
            bool keep = pf_row < keep_rows;
            if (RPS > 1) {
                const int g = pf_row / RPS;
                keep = (((g + 1) * a.keep_per_1024) >> 10) != ((g * a.keep_per_1024) >> 10);
            }
            mbar_expect_tx(&my_bars[s], (unsigned)(nrows * row_bytes));
            bulk_g2s(my_slots + (size_t)s * RPS * row_bytes, Xb + (size_t)(w0 + pf_row) * row_bytes, (unsigned)(nrows * row_bytes),
                     &my_bars[s], keep ? a.policy_keep : a.policy_stream);
        }
        pf_row += nrows;
        if (pf_row >= m) pf_row = 0;
    };
    if (m > 0)
        for (int s = 0; s < KS; ++s) issue_into(s);

    // ---- exchange 0: selection over the initial state (main3.cpp:203-204 in the first iteration)
    LocalBest mine;
    best_init(mine);
    for (long long j = r0 + threadIdx.x; j < r1; j += THREADS)
        best_consider(mine, j, a.f[j], a.alpha[j], a.y[j], a.cacheq ? a.cacheq[j] : -1, a.C, a.eps);
    int gather_no = 0;
    int num_iter = 1;                    // main3.cpp:197
    int stop = 0;
    double b_high = 0.0, b_low = 0.0;    // uninitialised in the reference (main3.cpp:188-189)
    long long i_high = a.n, i_low = a.n;
    long long cache_counter = a.cache_counter0;   // replicated: every CTA sees the same hit/miss sequence
    long long x_passes = 0, cache_hits = 0;

    for (;;) {
        if (!exchange(a, sh, mine, gather_no)) { stop = 5; break; }                 // the one grid-wide synchronisation
        if (num_iter > a.max_iter) { stop = 4; break; }                             // main3.cpp:283 (b keeps this iteration's b_high/b_low)
        const Best lo = mine.lo, hi = mine.hi;
        SVM_STAMP(0);
        SVM_GSTAMP(0);
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

        // cached iteration: this thread's first row of the pass is requested now, so that its loads and the K12 load
        // below travel together instead of one L2 round trip after the other
        const long long pj = r0 + threadIdx.x, pj2 = pj + THREADS;
        double p_f = 0.0, p_a = 0.0, p_kh = 0.0, p_kl = 0.0, p2_f = 0.0, p2_a = 0.0, p2_kh = 0.0, p2_kl = 0.0;
        int p_y = 0, p2_y = 0;
        long long p_q = -1, p2_q = -1;
        if (use_cache && pj < r1) {
            p_f = a.f[pj]; p_a = a.alpha[pj]; p_y = a.y[pj]; p_q = a.cacheq[pj];
            p_kh = a.cacheK[(size_t)(q_h % a.cache_cap) * a.n + pj];
            p_kl = a.cacheK[(size_t)(q_l % a.cache_cap) * a.n + pj];
            if (pj2 < r1) {              // a CTA owns more rows than it has threads (405 against 256 at 60k): the second row's loads travel now too
                p2_f = a.f[pj2]; p2_a = a.alpha[pj2]; p2_y = a.y[pj2]; p2_q = a.cacheq[pj2];
                p2_kh = a.cacheK[(size_t)(q_h % a.cache_cap) * a.n + pj2];
                p2_kl = a.cacheK[(size_t)(q_l % a.cache_cap) * a.n + pj2];
            }
        }

        // K12 = K(i_high, i_low)  (main3.cpp:241)
        Pivots<Vec, Q> pv;
        double K12, n_h = 0.0, n_l = 0.0;
        if (!use_cache) {
            const Vec *xh = reinterpret_cast<const Vec *>(a.X) + (size_t)i_high * a.pitch4;
            const Vec *xl = reinterpret_cast<const Vec *>(a.X) + (size_t)i_low * a.pitch4;
#pragma unroll
            for (int i = 0; i < Q; ++i) {
                const int q = i * 32 + lane;
                pv.h[i] = q < a.pitch4 ? __ldg(xh + q) : vec_zero<Vec>();
                pv.l[i] = q < a.pitch4 ? __ldg(xl + q) : vec_zero<Vec>();
            }
            if (sizeof(XT) == 4) { n_h = a.sqn[i_high]; n_l = a.sqn[i_low]; }
            SVM_STAMP(15);
            double acc = 0.0;
#pragma unroll
            for (int i = 0; i < Q; ++i)
                if (i * 32 + lane < a.pitch4) accum_pair<CH>(pv.l[i], pv.h[i], acc);
            acc = warp_sum_butterfly(acc);
            K12 = svm_exp(-a.gamma * dist2<XT>(n_h, n_l, acc));
        } else if (i_low >= a.row_begin && i_low < a.row_end) {
            // the one read of a value another CTA wrote: column i_low of the cached row was stored by the CTA that
            // owns row i_low, at least one exchange ago (its __threadfence below, this CTA's fence after the counter
            // spin).  Read it from L2: this SM's L1 may still hold the line from when the slot cached another row.
            K12 = __ldcg(&a.cacheK[(size_t)(q_h % a.cache_cap) * a.n + i_low]);
        } else {
            // row sharding: column i_low of the cached rows lives on the GPU that owns row i_low; X is
            // replicated, so form K12 here the way that GPU formed it (same code, same bits)
            const Vec *xh = reinterpret_cast<const Vec *>(a.X) + (size_t)i_high * a.pitch4;
            const Vec *xl = reinterpret_cast<const Vec *>(a.X) + (size_t)i_low * a.pitch4;
            double acc = 0.0;
#pragma unroll
            for (int i = 0; i < Q; ++i) {
                const int q = i * 32 + lane;
                if (q < a.pitch4) accum_pair<CH>(__ldg(xl + q), __ldg(xh + q), acc);
            }
            acc = warp_sum_butterfly(acc);
            K12 = svm_exp(-a.gamma * dist2<XT>(sizeof(XT) == 4 ? a.sqn[i_high] : 0.0, sizeof(XT) == 4 ? a.sqn[i_low] : 0.0, acc));
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

        LocalBest acc_best;
        best_init(acc_best);
        SVM_STAMP(1);
        SVM_GSTAMP(1);
        if (!use_cache) {
            // ---------------- X pass: a warp per group of RPS rows (or per row, below) ----------------
            ++x_passes;
            if constexpr (RPS > 1) {
              constexpr bool NEUTRAL_PAD = sizeof(XT) == 4;   // fp32 rows form dot products: an element times a zero pivot register adds +-0.
                                                              // FP64 rows form squared differences: elements beyond the row must be skipped
              const int ng = (a.pitch4 + 31) >> 5;                                 // register groups the row really has (<= Q)
              const int q_last = min((ng - 1) * 32 + lane, a.pitch4 - 1) * 16;     // byte offset of the lane's last 16-byte group, clamped into the row
              const bool last_valid = (ng - 1) * 32 + lane < a.pitch4;
              for (int base = 0; base < m; base += 32) {
                const int cnt = min(32, m - base);
                const long long j = w0 + base + lane;
                double fj = 0.0, aj = 0.0;
                int yj = 0;
                long long qj = -1;
                double nj = 0.0;
                if (lane < cnt) { fj = a.f[j]; aj = a.alpha[j]; yj = a.y[j]; if (sizeof(XT) == 4) nj = a.sqn[j]; if (a.cacheq) qj = a.cacheq[j]; }
                for (int t = 0; t < cnt; t += RPS) {
                    const int rows_here = min(RPS, cnt - t);
                    mbar_wait(&my_bars[cslot], cphase);
                    const unsigned char *grp = my_slots + (size_t)cslot * RPS * row_bytes;
                    // the lane's partial sums of rows t..t+RPS-1 against both pivots: [pivot * RPS + row]
                    double P[2 * RPS];
#pragma unroll
                    for (int k = 0; k < 2 * RPS; ++k) P[k] = 0.0;
                    if (rows_here == RPS) {
                        // Q is the largest group count of its class: 1, 2, 3-4, 5-7, 8 (fp32) / 13 (FP64); ng is warp-uniform
                        if (ng == Q) group_dot2<Vec, Q, CH, RPS, Q>(grp, row_bytes, lane, q_last, last_valid, pv, P);
                        else if (Q >= 2 && ng == Q - 1) group_dot2<Vec, Q, CH, RPS, (Q >= 2 ? Q - 1 : 1)>(grp, row_bytes, lane, q_last, last_valid, pv, P);
                        else if (Q >= 3 && ng == Q - 2) group_dot2<Vec, Q, CH, RPS, (Q >= 3 ? Q - 2 : 1)>(grp, row_bytes, lane, q_last, last_valid, pv, P);
                        else {
#pragma unroll
                            for (int r = 0; r < RPS; ++r) {
                                const Vec *row = reinterpret_cast<const Vec *>(grp + (size_t)r * row_bytes);
#pragma unroll
                                for (int i = 0; i < Q; ++i) {
                                    const int q = i * 32 + lane;
                                    if (NEUTRAL_PAD ? i * 32 < a.pitch4 : q < a.pitch4) {
                                        const Vec v = row[min(q, a.pitch4 - 1)];
                                        accum_pair<CH>(v, pv.h[i], P[r]);
                                        accum_pair<CH>(v, pv.l[i], P[RPS + r]);
                                    }
                                }
                            }
                        }
                    } else {
#pragma unroll
                        for (int r = 0; r < RPS; ++r) {
                            if (r < rows_here) {
                                const Vec *row = reinterpret_cast<const Vec *>(grp + (size_t)r * row_bytes);
#pragma unroll
                                for (int i = 0; i < Q; ++i) {
                                    const int q = i * 32 + lane;
                                    if (NEUTRAL_PAD ? i * 32 < a.pitch4 : q < a.pitch4) {
                                        const Vec v = row[min(q, a.pitch4 - 1)];
                                        accum_pair<CH>(v, pv.h[i], P[r]);
                                        accum_pair<CH>(v, pv.l[i], P[RPS + r]);
                                    }
                                }
                            }
                        }
                    }
                    // the group's slot is free again: request the next group of the ring into it
                    __syncwarp();
                    fence_proxy_async();
                    issue_into(cslot);
                    if (++cslot == KS) { cslot = 0; cphase ^= 1u; }
                    // halving exchange: offset 16 splits by pivot, the next offsets by row (8 and 4 for four rows, 8 for two), the
                    // rest add.  Each addition pairs the same operands as the xor butterfly, so a lane ends with the butterfly's
                    // value for pivot lane >> 4 and row (lane >> 2) & 3 (four rows) / (lane >> 3) & 1 (two rows)
                    double v1;
                    if constexpr (RPS == 4) {
                        double v4[4], v2[2];
                        {
                            const bool up = (lane & 16) != 0;
#pragma unroll
                            for (int k = 0; k < 4; ++k) v4[k] = (up ? P[4 + k] : P[k]) + __shfl_xor_sync(FULL, up ? P[k] : P[4 + k], 16);
                        }
                        {
                            const bool up = (lane & 8) != 0;
#pragma unroll
                            for (int k = 0; k < 2; ++k) v2[k] = (up ? v4[2 + k] : v4[k]) + __shfl_xor_sync(FULL, up ? v4[k] : v4[2 + k], 8);
                        }
                        {
                            const bool up = (lane & 4) != 0;
                            v1 = (up ? v2[1] : v2[0]) + __shfl_xor_sync(FULL, up ? v2[0] : v2[1], 4);
                        }
                    } else {
                        double v2[2];
                        {
                            const bool up = (lane & 16) != 0;
#pragma unroll
                            for (int k = 0; k < 2; ++k) v2[k] = (up ? P[2 + k] : P[k]) + __shfl_xor_sync(FULL, up ? P[k] : P[2 + k], 16);
                        }
                        {
                            const bool up = (lane & 8) != 0;
                            v1 = (up ? v2[1] : v2[0]) + __shfl_xor_sync(FULL, up ? v2[0] : v2[1], 8);
                        }
                        v1 = v1 + __shfl_xor_sync(FULL, v1, 4);
                    }
                    v1 = v1 + __shfl_xor_sync(FULL, v1, 2);
                    v1 = v1 + __shfl_xor_sync(FULL, v1, 1);
                    {
                        const int r = RPS == 4 ? (lane >> 2) & 3 : (lane >> 3) & 1;
                        const bool writer = RPS == 4 ? (lane & 3) == 0 : (lane & 7) == 0;
                        if (writer && r < rows_here) *(reinterpret_cast<double *>(&stage[t + r]) + (lane >> 4)) = v1;
                    }
                }
                __syncwarp();
                if (lane < cnt) {
                    const double2 dd = stage[lane];
                    const double kh = svm_exp(-a.gamma * dist2<XT>(n_h, nj, dd.x));
                    const double kl = svm_exp(-a.gamma * dist2<XT>(n_l, nj, dd.y));
                    fj = fj + (c_h * kh + c_l * kl);            // main3.cpp:274
                    a.f[j] = fj;
                    if (slot_h >= 0 && !hit_h) a.cacheK[(size_t)slot_h * a.n + j] = kh;
                    if (slot_l >= 0 && !hit_l) a.cacheK[(size_t)slot_l * a.n + j] = kl;
                    if (slot_h >= 0 && !(hit_h && hit_l)) __threadfence();   // K12 of a later cached iteration is read by other CTAs
                    if (j == i_high) { aj = na_h; a.alpha[j] = aj; if (a.cacheq && !hit_h) { qj = q_h; a.cacheq[j] = qj; } }
                    if (j == i_low) { aj = na_l; a.alpha[j] = aj; if (a.cacheq && !hit_l) { qj = q_l; a.cacheq[j] = qj; } }
                    best_consider(acc_best, j, fj, aj, yj, qj, a.C, a.eps);
                }
                __syncwarp();                                  // the parked sums have been read: the next block may overwrite them
              }
            } else
            for (int base = 0; base < m; base += 32) {
                const int cnt = min(32, m - base);
                // lane t finishes row base+t: its state is requested now and used after the group
                const long long j = w0 + base + lane;
                double fj = 0.0, aj = 0.0;
                int yj = 0;
                long long qj = -1;
                double nj = 0.0;
                if (lane < cnt) { fj = a.f[j]; aj = a.alpha[j]; yj = a.y[j]; if (sizeof(XT) == 4) nj = a.sqn[j]; if (a.cacheq) qj = a.cacheq[j]; }
                double my_dh = 0.0, my_dl = 0.0;
                for (int t = 0; t < cnt; ++t) {
                    mbar_wait(&my_bars[cslot], cphase);
                    double dh, dl;
                    row_dot2<Vec, Q, CH>(reinterpret_cast<const Vec *>(my_slots + (size_t)cslot * row_bytes), pv, a.pitch4, lane, dh, dl);
                    if (lane == t) { my_dh = dh; my_dl = dl; }
                    // the slot is free again: request the next row of the ring into it
                    __syncwarp();
                    fence_proxy_async();
                    issue_into(cslot);
                    if (++cslot == KS) { cslot = 0; cphase ^= 1u; }
                }
                if (lane < cnt) {
                    const double kh = svm_exp(-a.gamma * dist2<XT>(n_h, nj, my_dh));
                    const double kl = svm_exp(-a.gamma * dist2<XT>(n_l, nj, my_dl));
                    fj = fj + (c_h * kh + c_l * kl);            // main3.cpp:274
                    a.f[j] = fj;
                    if (slot_h >= 0 && !hit_h) a.cacheK[(size_t)slot_h * a.n + j] = kh;
                    if (slot_l >= 0 && !hit_l) a.cacheK[(size_t)slot_l * a.n + j] = kl;
                    if (slot_h >= 0 && !(hit_h && hit_l)) __threadfence();   // K12 of a later cached iteration is read by other CTAs
                    if (j == i_high) { aj = na_h; a.alpha[j] = aj; if (a.cacheq && !hit_h) { qj = q_h; a.cacheq[j] = qj; } }
                    if (j == i_low) { aj = na_l; a.alpha[j] = aj; if (a.cacheq && !hit_l) { qj = q_l; a.cacheq[j] = qj; } }
                    best_consider(acc_best, j, fj, aj, yj, qj, a.C, a.eps);
                }
            }
        } else {
            // ---------------- cached pass: a thread per row over two cached kernel rows ----------------
            ++cache_hits;
            const double *kh_row = a.cacheK + (size_t)slot_h * a.n;
            const double *kl_row = a.cacheK + (size_t)slot_l * a.n;
            for (long long j = pj; j < r1; j += THREADS) {
                const bool first = j == pj, second = j == pj2;
                double fj = first ? p_f : second ? p2_f : a.f[j], aj = first ? p_a : second ? p2_a : a.alpha[j];
                const int yj = first ? p_y : second ? p2_y : a.y[j];
                const long long qj = first ? p_q : second ? p2_q : a.cacheq[j];
                const double khj = first ? p_kh : second ? p2_kh : kh_row[j], klj = first ? p_kl : second ? p2_kl : kl_row[j];
                fj = fj + (c_h * khj + c_l * klj);
                a.f[j] = fj;
                if (j == i_high) { aj = na_h; a.alpha[j] = aj; }
                if (j == i_low) { aj = na_l; a.alpha[j] = aj; }
                best_consider(acc_best, j, fj, aj, yj, qj, a.C, a.eps);
            }
        }

        SVM_STAMP(2);
        SVM_GSTAMP(2);
        ++gather_no;
        mine = acc_best;
        ++num_iter;                                              // main3.cpp:281
    }

    // drain the ring: every requested copy must land before the CTA may exit
    if (m > 0)
        for (int k = 0; k < KS; ++k) {
            mbar_wait(&my_bars[cslot], cphase);
            if (++cslot == KS) { cslot = 0; cphase ^= 1u; }
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

// kslots: ring depth per warp in slots; a slot holds rows_per_step rows
inline size_t smo_smem_bytes(int pitch4, int kslots, int warps, int rows_per_step = 1)
{
    return (size_t)warps * kslots * rows_per_step * pitch4 * 16 + sizeof(uint64_t) * warps * SMO_KSLOTS_MAX + sizeof(Shared) + 128 +
           (size_t)warps * 32 * sizeof(double2);                 // + the parked sums of the four-rows-per-step pass
}

} // namespace svm
