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
//     3. reduces its candidates inside its thread-block cluster (distributed shared memory), whose leader publishes one
//        128-byte self-validating record and polls the other clusters' records directly (an all-gather through L2: the
//        only grid-wide synchronisation; see "the record exchange" below).
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
    Rec *rec;              // [2][SMO_MAX_RECS] records of the exchange, zeroed before the launch
    unsigned *abort_word;  // set when a wait was given up (a peer GPU is gone): ends every other wait; zeroed before the launch
    int *trace;            // [trace_cap][2]
    SmoResult *result;
    long long n;
    int pitch4;            // 16-byte groups per row
    int kslots;            // ring depth per warp (2..SMO_KSLOTS_MAX), chosen so the ring fits in shared memory
    int cache_cap;
    long long cache_counter0;  // allocation number to start from (> 0 when the cache is kept from an earlier run on the same X)
    int trace_cap;
    int max_iter;
    int wss;               // 1: first-order pair (the reference); 2: second-order choice of i_low
    double C, gamma, eps, tau;
    unsigned long long policy_stream;  // L2 policy for the streamed part of every CTA's rows
    unsigned long long policy_keep;    // L2 policy for the leading keep_frac of every warp's rows
    int keep_per_1024;                 // rows (per 1024) of each warp's range loaded with policy_keep
    // ---- multi-GPU row sharding (peers == 1: single GPU) ----
    int peers;                          // GPUs sharing the rows
    int rank;                           // this GPU
    long long row_begin, row_end;       // rows owned by this GPU (X is replicated, state is global-sized)
    Rec *mailbox;                       // [2][SMO_MB_COPIES][8] local mailbox written by peers
    int mailbox_copies;                 // copies in use: min(SMO_MB_COPIES, clusters of the smallest grid among the GPUs)
    Rec *peer_mailbox[8];               // peers' mailboxes (peer-mapped device pointers; [rank] is the local one)
    int epoch;                          // launch number, folded into the mailbox words so that stale records never match
    // ---- uint8 rows (storage "u8"): X is two arrays, [n][4 * mw] bytes of the main class, then [n][tail_pitch] bytes of the
    //      tail (the sections of a row are copied by different bulk copies: the main class in groups of eight rows into the
    //      ring, the tails of a block of tail_rows rows straight into the buffer they are walked from)
    const int *nmain;                   // [n] integer squared norms over the main class
    const double *ntail;                // [n] weighted squared norms over the tail
    const double *wt;                   // [tw * 4] weights (1 / range)^2 of the tail features (0 for padding bytes)
    int mw, tw;                         // 4-byte words of a row in the main class / in the tail, each padded to a multiple of 32
    double wR;                          // weight of the main class
    int tail_pitch;                     // bytes per row of the tail array and of the per-warp buffer that receives a block of it: 4 * tw + 16
                                        // (the odd pitch in 16-byte units spreads the rows over the banks); 0 when there is no tail
    int tail_rows;                      // rows of such a block: 32 (a lane per row of the warp's 32-row block), 16 or 8 when shared memory is short
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
    return sizeof(XT) == 1 ? 8 : (sizeof(XT) == 4 && THREADS <= 256) ? (Q == 8 ? 2 : 4) : (sizeof(XT) == 8 && Q == 13 && THREADS == 256) ? 2 : 1;
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

// ---- uint8 rows.  For integer-valued input the reference's scaled difference (code/main3.cpp:74-89, :97-100) is
// (a_k - b_k)^2 = (u_k - p_k)^2 / range_k^2 with integer u, p, range, so the raw bytes are kept and
//     d2 = S_main * w_R + tail
//   S_main = sum over the features whose range is the most frequent one (R) of (u_k - p_k)^2, an exact integer formed as
//            |u|^2 + |p|^2 - 2 u.p with four products per IDP4A;
//   tail   = (N(u) + N(p)) - 2 D(u, p) over the other non-constant features, D(u, p) = sum_k u_k * (w_k p_k) in FP64,
//            as (D_0 + D_1) + (D_2 + D_3) where D_c adds, in ascending order with fma, the features of the 4-byte words
//            w = c (mod 4) of the tail;
//            N(u) = D(u, u), so duplicates are at distance exactly 0.
// A zero byte of the pivot adds u_k * 0 = +0 to its partial sum, i.e. nothing: the pass therefore walks only the tail
// positions where one of the two pivots is non-zero (MNIST's border pixels: a handful out of 233), from four sub-lists
// (one per c) that every warp builds at the top of the iteration.
// a byte (any integer below 2^32) as a double without the conversion unit: it sits in the mantissa of 2^52, one DADD takes 2^52
// away again (exact)
__device__ __forceinline__ double u32_to_double(unsigned v)
{
    return __hiloint2double(0x43300000, (int)v) - 4503599627370496.0;
}
struct U8Lists {
    double2 *pp;            // [4][cap] weighted pivot bytes {w_k p_high_k, w_k p_low_k}
    unsigned short *kk;     // [4][cap] tail position k
    unsigned char *xlt;     // [4 * tw] tail bytes of the low pivot (the "row" of K12)
    int tw;
    int cap;                // tw + 4: a sub-list holds at most tw entries, followed by four zero-weight ones
};
// Every warp calls this with the same arguments and writes the same values (no CTA barrier needed: a warp reads what it
// wrote itself).  Sub-list c holds the positions of the words w = c (mod 4), i.e. of the lanes l = c (mod 4), in ascending
// order; returns the length of the sub-list of this lane's class, lane & 3 (the sub-list the lane walks later).
__device__ __forceinline__ int u8_build_lists(const unsigned *__restrict__ xh_tail, const unsigned *__restrict__ xl_tail, const double *sh_wt,
                                              const U8Lists &L, int lane)
{
    const int c = lane & 3;
    const unsigned before = (0x11111111u << c) & ((1u << lane) - 1u);      // lower lanes of this lane's class
    const unsigned cls = 0x11111111u << c;
    int base = 0;                                                          // entries of this class so far
    for (int w0 = 0; w0 < L.tw; w0 += 64) {                                // two slots at a time: their four loads travel together
        const bool two = w0 + 32 < L.tw;                                   // tw is a multiple of 32
        unsigned thv[2], tlv[2];
        thv[0] = __ldg(xh_tail + w0 + lane);
        tlv[0] = __ldg(xl_tail + w0 + lane);
        thv[1] = two ? __ldg(xh_tail + w0 + 32 + lane) : 0u;
        tlv[1] = two ? __ldg(xl_tail + w0 + 32 + lane) : 0u;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            if (h == 1 && !two) break;
            const int w = w0 + 32 * h + lane;
            const unsigned th = thv[h], tl = tlv[h];
            reinterpret_cast<unsigned *>(L.xlt)[w] = tl;
            int e = c * L.cap + base, total = 0;
            bool nz[4];
#pragma unroll
            for (int b = 0; b < 4; ++b) {
                nz[b] = (((th | tl) >> (8 * b)) & 255u) != 0u;
                const unsigned ballot = __ballot_sync(FULL, nz[b]);
                e += __popc(ballot & before);                              // all bytes of the lower lanes come first
                total += __popc(ballot & cls);
            }
#pragma unroll
            for (int b = 0; b < 4; ++b)
                if (nz[b]) {
                    const double wk = sh_wt[4 * w + b];
                    L.kk[e] = (unsigned short)(4 * w + b);
                    L.pp[e] = make_double2(wk * u32_to_double((th >> (8 * b)) & 255u), wk * u32_to_double((tl >> (8 * b)) & 255u));
                    ++e;
                }
            base += total;
        }
    }
    // four zero-weight entries behind every sub-list: the walkers take four entries at a time, the first four without looking
    // at the length (u * 0 adds +0)
    if (lane < 16) {
        const int cc = lane & 3;                                           // lanes 0..15: class cc, pad entry lane >> 2
        const int n_cc = __shfl_sync(0xffffu, base, cc);
        const int e = cc * L.cap + n_cc + (lane >> 2);
        L.kk[e] = 0;
        L.pp[e] = make_double2(0.0, 0.0);
    }
    __syncwarp();
    return base;
}
// D(row, pivots) for the row whose tail bytes start at `rowtail` (shared memory): the lane walks sub-list c = lane & 3
// (`n` entries), the four lanes of a row then form (D_0 + D_1) + (D_2 + D_3).  Whole warp; n = 0 for a lane whose row does
// not exist.
// four entries of a sub-list against one row: independent loads first, then the two fma chains (entry order kept)
__device__ __forceinline__ void u8_walk4(const unsigned char *rowtail, const double2 *pp, const unsigned short *kk, int e, double &ah, double &al)
{
    const uint2 k4 = *reinterpret_cast<const uint2 *>(kk + e);             // e is a multiple of 4: 8-byte aligned
    const double2 p0 = pp[e], p1 = pp[e + 1], p2 = pp[e + 2], p3 = pp[e + 3];
    const double u0 = u32_to_double(rowtail[k4.x & 0xffffu]), u1 = u32_to_double(rowtail[k4.x >> 16]);
    const double u2 = u32_to_double(rowtail[k4.y & 0xffffu]), u3 = u32_to_double(rowtail[k4.y >> 16]);
    ah = fma(u0, p0.x, ah); al = fma(u0, p0.y, al);
    ah = fma(u1, p1.x, ah); al = fma(u1, p1.y, al);
    ah = fma(u2, p2.x, ah); al = fma(u2, p2.y, al);
    ah = fma(u3, p3.x, ah); al = fma(u3, p3.y, al);
}
__device__ __forceinline__ void u8_tail_dots(const unsigned char *rowtail, int n, const U8Lists &L, int lane, double &dh, double &dl)
{
    const int c = lane & 3;
    const double2 *pp = L.pp + c * L.cap;
    const unsigned short *kk = L.kk + c * L.cap;
    double ah = 0.0, al = 0.0;
    for (int e = 0; e < n; e += 4) u8_walk4(rowtail, pp, kk, e, ah, al);
    __syncwarp();
    ah = ah + __shfl_xor_sync(FULL, ah, 1);
    al = al + __shfl_xor_sync(FULL, al, 1);
    dh = ah + __shfl_xor_sync(FULL, ah, 2);
    dl = al + __shfl_xor_sync(FULL, al, 2);
}
// SLOTS = the number of 128-byte slots of the main class (a template parameter of the kernel: no branch between the slots,
// so the loads of all slots are in flight together).
template <int QM, int SLOTS, int RPS, int NR>
__device__ __forceinline__ void u8_main_sums(const unsigned char *lanep, int row_bytes, const unsigned (&pm_h)[QM], const unsigned (&pm_l)[QM],
                                             int (&SI)[2 * RPS])
{
#pragma unroll
    for (int i = 0; i < SLOTS; ++i) {
        unsigned w[NR];
#pragma unroll
        for (int r = 0; r < NR; ++r) w[r] = *reinterpret_cast<const unsigned *>(lanep + r * row_bytes + i * 128);
#pragma unroll
        for (int r = 0; r < NR; ++r) {
            SI[r] = (int)__dp4a(w[r], pm_h[i], (unsigned)SI[r]);
            SI[RPS + r] = (int)__dp4a(w[r], pm_l[i], (unsigned)SI[RPS + r]);
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
// Second-order choice of i_low (svmb200_params.wss = 2; Fan, Chen, Lin 2005 -- not in the reference): among the rows of
// I_low with f_t > b_high, the one that maximises the gain (f_t - b_high)^2 / eta_t of the two-variable sub-problem,
// eta_t = 2 - 2 K(i_high, t); rows with eta_t <= eps are passed over; lowest index among equal gains.  Both sides of the
// candidate elect the same row (min of -gain, max of gain), so that the record also carries f_t (in lo.alpha) next to
// alpha_t, y_t and the cache number.
__device__ __forceinline__ void gain_consider(LocalBest &b, long long j, double fj, double aj, int yj, long long qj, double kh, double b_high,
                                              double C, double eps)
{
    if (!in_I_low(yj, aj, C, eps) || !(fj > b_high)) return;
    const double eta = 1.0 + 1.0 - 2.0 * kh;
    if (eta <= eps) return;
    const double diff = fj - b_high;
    const double gain = diff * diff / eta;
    if (gain > b.hi.f) {
        b.hi.f = gain; b.hi.idx = (int)j; b.hi.alpha = aj; b.hi.y = yj; b.hi.q = qj;
        b.lo.f = -gain; b.lo.idx = (int)j; b.lo.alpha = fj; b.lo.y = yj; b.lo.q = qj;
    }
}

// What an iteration's pass does with a row once both kernel values are known.
struct IterState {
    double c_h, c_l, na_h, na_l, b_high;
    long long i_high, i_low, q_h, q_l;
    int slot_h, slot_l;
    bool hit_h, hit_l;
    bool gain;            // the pass only chooses i_low by gain (wss = 2): no f update; kl is a copy of kh
};
__device__ __forceinline__ void finish_row(const SmoArgs &a, const IterState &it, long long j, double fj, double aj, int yj, long long qj,
                                           double kh, double kl, LocalBest &acc)
{
    if (it.gain) {
        if (it.slot_h >= 0 && !it.hit_h) {
            a.cacheK[(size_t)it.slot_h * a.n + j] = kh;
            __threadfence();                                     // K12 of the update that follows is read by other CTAs
            if (j == it.i_high && a.cacheq) { qj = it.q_h; a.cacheq[j] = qj; }
        }
        gain_consider(acc, j, fj, aj, yj, qj, kh, it.b_high, a.C, a.eps);
        return;
    }
    fj = fj + (it.c_h * kh + it.c_l * kl);                       // main3.cpp:274
    a.f[j] = fj;
    if (it.slot_h >= 0 && !it.hit_h) a.cacheK[(size_t)it.slot_h * a.n + j] = kh;
    if (it.slot_l >= 0 && !it.hit_l) a.cacheK[(size_t)it.slot_l * a.n + j] = kl;
    if (it.slot_h >= 0 && !(it.hit_h && it.hit_l)) __threadfence();   // K12 of a later cached iteration is read by other CTAs
    if (j == it.i_high) { aj = it.na_h; a.alpha[j] = aj; if (a.cacheq && !it.hit_h) { qj = it.q_h; a.cacheq[j] = qj; } }
    if (j == it.i_low) { aj = it.na_l; a.alpha[j] = aj; if (a.cacheq && !it.hit_l) { qj = it.q_l; a.cacheq[j] = qj; } }
    best_consider(acc, j, fj, aj, yj, qj, a.C, a.eps);
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
// The lanes that hold the (value, lowest index) minimum of (flo, ilo) and maximum of (fhi, ihi).  The rounds of the two
// sides are issued side by side (a REDUX takes tens of cycles: the sides hide each other's latency); the round on the
// index runs only when more than one lane holds the winning value.
__device__ __forceinline__ void warp_arg_best2(double flo, int ilo, double fhi, int ihi, int &src_lo, int &src_hi)
{
    const unsigned long long ka = order_key(flo), kb = ~order_key(fhi);
    const unsigned ah = (unsigned)(ka >> 32), al = (unsigned)ka, bh = (unsigned)(kb >> 32), bl = (unsigned)kb;
    const unsigned mah = __reduce_min_sync(FULL, ah), mbh = __reduce_min_sync(FULL, bh);
    const unsigned mal = __reduce_min_sync(FULL, ah == mah ? al : 0xffffffffu), mbl = __reduce_min_sync(FULL, bh == mbh ? bl : 0xffffffffu);
    const bool ta = ah == mah && al == mal, tb = bh == mbh && bl == mbl;
    unsigned ba = __ballot_sync(FULL, ta), bb = __ballot_sync(FULL, tb);
    if (ba & (ba - 1)) {                                   // warp-uniform: several lanes hold the minimum
        const unsigned midx = __reduce_min_sync(FULL, ta ? (unsigned)ilo : 0xffffffffu);
        ba = __ballot_sync(FULL, ta && (unsigned)ilo == midx);
    }
    if (bb & (bb - 1)) {
        const unsigned midx = __reduce_min_sync(FULL, tb ? (unsigned)ihi : 0xffffffffu);
        bb = __ballot_sync(FULL, tb && (unsigned)ihi == midx);
    }
    src_lo = __ffs(ba) - 1;                                // lanes that tie on the index too hold the same row (or no row): any of them
    src_hi = __ffs(bb) - 1;
}
__device__ __forceinline__ void best_take(Best &b, int src)
{
    b.f = __shfl_sync(FULL, b.f, src);
    b.alpha = __shfl_sync(FULL, b.alpha, src);
    b.q = __shfl_sync(FULL, b.q, src);
    b.idx = __shfl_sync(FULL, b.idx, src);
    b.y = __shfl_sync(FULL, b.y, src);
}
__device__ __forceinline__ void best_warp_reduce(LocalBest &b)
{
    int src_lo, src_hi;
    warp_arg_best2(b.lo.f, b.lo.idx, b.hi.f, b.hi.idx, src_lo, src_hi);
    best_take(b.lo, src_lo);
    best_take(b.hi, src_hi);
}
__device__ __forceinline__ void best_merge(LocalBest &b, const Best &lo, const Best &hi)
{
    if (better_min(lo.f, lo.idx, b.lo.f, b.lo.idx)) b.lo = lo;
    if (better_max(hi.f, hi.idx, b.hi.f, b.hi.idx)) b.hi = hi;
}

// The scalar part of an iteration, code/main3.cpp:235-270 operation for operation: s, eta, calculate_U_V (:145-159),
// the clipped alpha_low', alpha_high', and the two coefficients of the f update.  Every CTA runs it redundantly on the
// same inputs (identical bits everywhere, so nothing is broadcast).  stop: 0 = go on, 2 = infeasible U and V (:246),
// 3 = non positive eta (:253).
struct PairStep { double U, V, eta, na_h, na_l, c_h, c_l; int stop; };
__device__ __forceinline__ PairStep pair_update(int y_h, int y_l, double a_h, double a_l, double b_high, double b_low, double K12,
                                                double C, double eps)
{
    PairStep r;
    const int s = y_h * y_l;
    const double K11 = 1.0, K22 = 1.0;                    // exp(-gamma * 0)
    r.eta = K11 + K22 - 2.0 * K12;
    if (s == -1) {
        const double u = a_l - a_h, v = C + a_l - a_h;
        r.U = (0.0 < u) ? u : 0.0;
        r.V = (v < C) ? v : C;
    } else {
        const double u = a_l + a_h - C, v = a_l + a_h;
        r.U = (0.0 < u) ? u : 0.0;
        r.V = (v < C) ? v : C;
    }
    r.na_h = a_h; r.na_l = a_l; r.c_h = 0.0; r.c_l = 0.0;
    r.stop = 0;
    if (!(r.U <= r.V + 1e-12)) { r.stop = 2; return r; }  // main3.cpp:246
    if (r.eta <= eps) { r.stop = 3; return r; }            // main3.cpp:253
    double na_l = a_l + y_l * (b_high - b_low) / r.eta;    // main3.cpp:259
    if (na_l > r.V) na_l = r.V;
    if (na_l < r.U) na_l = r.U;
    r.na_l = na_l;
    r.na_h = a_h + s * (a_l - na_l);                       // main3.cpp:266
    r.c_h = (r.na_h - a_h) * y_h;                          // delta_alpha_i_high * y[i_high]
    r.c_l = (r.na_l - a_l) * y_l;
    return r;
}

// ---- the record exchange ---------------------------------------------------------------------
// The one grid-wide synchronisation of an iteration, in four hops:
//   1. warp: winners by REDUX (both sides side by side) -> shared memory of the cluster's LEADER CTA: the leader's warps write
//      locally, the warps of the other CTA (two CTAs on neighbouring SMs form a cluster) through distributed shared memory
//      (st.async: the bytes are counted on the leader's mbarrier, no fence);
//   2. cluster: the leader's warp 0 folds the up to 32 entries in one REDUX round;
//   3. grid: the leader stores ONE 128-byte record per cluster (74 on a B200), in SMO_REC_COPIES copies, and polls the others'
//      records directly -- no arrival counter, one L2 round trip after the last record lands.  A record is 16 words of 8 bytes,
//      each {32 data bits, 32-bit exchange number}: an aligned 8-byte access is single-copy atomic, so a word that shows the
//      right number also shows its data and a reader that is too early simply re-reads.  The poll is warp-wide and
//      coalesced (8 lanes x 16 bytes = one record, four records per load), which keeps the number of L2 requests per
//      record line at one per polling leader and round;
//   4. cluster: the leader writes the grid's winners into its partner's shared memory and arrives on its mbarrier.
// With a cluster size of 1 (tiny grids, or a device that cannot co-schedule the clusters) hops 2 and 4 fall away.
constexpr int REC_WORDS = 16;
constexpr int SMO_MAX_RECS = 160;         // records per exchange: clusters of the grid (74 on a B200; 148 with a cluster size of 1)
// Loads that poll one line from many SMs serialise in L2 (measured: +29 ns per further reader of a line and round -- 74 readers
// of every record cost 2.6 us per exchange, 18 cost 1.0 us).  So every record is stored in SMO_REC_COPIES copies (lines) and
// leader c polls copy c mod SMO_REC_COPIES: a line then has at most five readers.  The peers' mailboxes likewise.
constexpr int SMO_REC_COPIES = 16;
constexpr int SMO_MB_COPIES = 16;                          // copies of a GPU's mailbox
constexpr int SMO_MB_RECS = 2 * SMO_MB_COPIES * 8;         // [generation][copy][sending GPU]; the rendezvous flags follow
struct __align__(128) Rec { unsigned long long w[REC_WORDS]; };

// word layout: the two 32-bit halves of one value share a 16-byte pair of words
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
// the 16 data words of a record (as staged in shared memory) -> winners
__device__ __forceinline__ void rec_unpack32(const unsigned *d, LocalBest &b)
{
    b.lo.f = __hiloint2double((int)d[1], (int)d[0]);
    b.lo.idx = (int)d[2];
    b.hi.f = __hiloint2double((int)d[4], (int)d[3]);
    b.hi.idx = (int)d[5];
    b.lo.alpha = __hiloint2double((int)d[7], (int)d[6]);
    b.hi.alpha = __hiloint2double((int)d[9], (int)d[8]);
    b.lo.y = (int)d[10];
    b.hi.y = (int)d[11];
    b.lo.q = (long long)(((unsigned long long)d[13] << 32) | d[12]);
    b.hi.q = (long long)(((unsigned long long)d[15] << 32) | d[14]);
}
// the word pair `pair` (0..7) of a record
__device__ __forceinline__ void rec_word_pair(const LocalBest &b, int pair, unsigned &d0, unsigned &d1)
{
    switch (pair) {
    case 0: d0 = rec_word(b, 0); d1 = rec_word(b, 1); break;
    case 1: d0 = rec_word(b, 2); d1 = rec_word(b, 3); break;
    case 2: d0 = rec_word(b, 4); d1 = rec_word(b, 5); break;
    case 3: d0 = rec_word(b, 6); d1 = rec_word(b, 7); break;
    case 4: d0 = rec_word(b, 8); d1 = rec_word(b, 9); break;
    case 5: d0 = rec_word(b, 10); d1 = rec_word(b, 11); break;
    case 6: d0 = rec_word(b, 12); d1 = rec_word(b, 13); break;
    default: d0 = rec_word(b, 14); d1 = rec_word(b, 15); break;
    }
}
// the whole warp publishes record `cid` in all copies, four copies per store instruction: lane l writes the word pair l & 7
// of the copies l >> 3, 4 + (l >> 3), ...
__device__ __forceinline__ void rec_store_copies(Rec *recs, int cid, const LocalBest &b, unsigned seq, int lane)
{
    static_assert(SMO_REC_COPIES % 4 == 0, "32 lanes = 4 copies x 8 word pairs");
    unsigned d0, d1;
    rec_word_pair(b, lane & 7, d0, d1);
    const unsigned long long v0 = ((unsigned long long)seq << 32) | d0, v1 = ((unsigned long long)seq << 32) | d1;
#pragma unroll
    for (int c = lane >> 3; c < SMO_REC_COPIES; c += 4) {
        Rec *r = recs + (size_t)c * SMO_MAX_RECS + cid;
        asm volatile("st.relaxed.gpu.global.v2.u64 [%0], {%1, %2};" ::"l"(&r->w[2 * (lane & 7)]), "l"(v0), "l"(v1) : "memory");
    }
}
// one copy of this GPU's record into every peer's mailbox (NVLink, system scope): lane l writes the word pair l & 7 for the
// peers l >> 3 and 4 + (l >> 3)
__device__ __forceinline__ void mailbox_store(Rec *const (&peer_mailbox)[8], int peers, int rank, int slot, const LocalBest &b, unsigned seq, int lane)
{
    unsigned d0, d1;
    rec_word_pair(b, lane & 7, d0, d1);
    const unsigned long long v0 = ((unsigned long long)seq << 32) | d0, v1 = ((unsigned long long)seq << 32) | d1;
#pragma unroll
    for (int p = lane >> 3; p < 8; p += 4)
        if (p < peers) {
            Rec *r = peer_mailbox[p] + slot + rank;
            asm volatile("st.relaxed.sys.global.v2.u64 [%0], {%1, %2};" ::"l"(&r->w[2 * (lane & 7)]), "l"(v0), "l"(v1) : "memory");
        }
}
template <bool SYS>
__device__ __forceinline__ void ld_pair(const unsigned long long *p, unsigned long long &w0, unsigned long long &w1)
{
    if (SYS) asm volatile("ld.relaxed.sys.global.v2.u64 {%0, %1}, [%2];" : "=l"(w0), "=l"(w1) : "l"(p) : "memory");
    else asm volatile("ld.relaxed.gpu.global.v2.u64 {%0, %1}, [%2];" : "=l"(w0), "=l"(w1) : "l"(p) : "memory");
}
__device__ __forceinline__ unsigned ld_relaxed_gpu_u32(const unsigned *p)
{
    unsigned v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed_gpu_u32(unsigned *p, unsigned v)
{
    asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// ---- thread-block cluster primitives (distributed shared memory + cluster-scope mbarrier operations)
__device__ __forceinline__ unsigned cluster_ctarank() { unsigned r; asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ unsigned cluster_nctarank() { unsigned r; asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(r)); return r; }
__device__ __forceinline__ void cluster_sync_all()
{
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cluster address of `p` (an address in this CTA's shared memory) in CTA `rank` of the cluster
__device__ __forceinline__ uint32_t dsmem_addr(const void *p, unsigned rank)
{
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(smem_u32(p)), "r"(rank));
    return r;
}
// 16 bytes into another CTA's shared memory, completing 16 bytes of the transaction count of ITS mbarrier: data and
// notification travel together, so the sender needs no fence and the receiver's wait on the barrier publishes the data
__device__ __forceinline__ void dsmem_st_async2(uint32_t addr, unsigned long long a, unsigned long long b, uint32_t bar_addr)
{
    asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v2.b64 [%0], {%1, %2}, [%3];"
                 ::"r"(addr), "l"(a), "l"(b), "r"(bar_addr) : "memory");
}
// wait for the phase with the given parity; false when `abort_word` becomes non-zero first (checked now and then)
__device__ __forceinline__ bool mbar_wait_abortable(uint64_t *bar, unsigned parity, const unsigned *abort_word)
{
    unsigned done = 0, spins = 0;
    for (;;) {
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
            "selp.u32 %0, 1, 0, p;\n"
            "}\n" : "=r"(done) : "r"(smem_u32(bar)), "r"(parity) : "memory");
        if (done) return true;
        if ((++spins & 63u) == 0 && ld_relaxed_gpu_u32(abort_word)) return false;
    }
}
constexpr unsigned BEST_PAIR_BYTES = 64;   // two winners of 32 bytes
// one side's winner (32 bytes) into CTA-remote shared memory at `addr`, counted on the mbarrier at `bar_addr` there
__device__ __forceinline__ void dsmem_put_best(uint32_t addr, const Best &b, uint32_t bar_addr)
{
    dsmem_st_async2(addr, (unsigned long long)__double_as_longlong(b.f), (unsigned long long)__double_as_longlong(b.alpha), bar_addr);
    dsmem_st_async2(addr + 16, (unsigned long long)b.q, ((unsigned long long)(unsigned)b.y << 32) | (unsigned)b.idx, bar_addr);
}

constexpr int POLL_GROUPS_MAX = 5;        // groups of four records a warp polls: 8 warps x 5 x 4 = 160
constexpr int REC_STRIDE = REC_WORDS + 1;

struct __align__(16) Shared {
    Best wlo[2 * SMO_MAX_WARPS], whi[2 * SMO_MAX_WARPS];  // per-warp winners of this CTA's rows; cluster leader: behind them (from
                                                  // [warps of a CTA] on) the partner CTA's, written through DSMEM
    Best rlo, rhi;                                // partner CTA: the grid's winners (written through DSMEM by the leader)
    uint64_t bar_in, bar_out;                     // mbarriers of the two cluster hops
    double k12;                                   // K12 of a cached iteration, read from L2 by one thread of the CTA
    unsigned recdata[SMO_MAX_RECS][REC_STRIDE];   // data words of the polled records (odd stride: lane l reads record l without bank conflicts)
};

#ifdef SVM_TIMING
#define SVM_STAMP_T(k, tid) do { if (a.timing && blockIdx.x == 0 && threadIdx.x == (tid) && gather_no < a.timing_cap) a.timing[(size_t)gather_no * 16 + (k)] = clock64(); } while (0)
#define SVM_STAMP(k) SVM_STAMP_T(k, 0)
__device__ __forceinline__ unsigned long long gtimer() { unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }
#define SVM_GSTAMP(k) do { if (a.timing && threadIdx.x == 0 && gather_no >= 100 && gather_no < 108) a.timing[(size_t)a.timing_cap * 16 + ((size_t)(gather_no - 100) * gridDim.x + blockIdx.x) * 4 + (k)] = (long long)gtimer(); } while (0)
// cycle sums of the pass phases of CTA 0's warp 0 (profiles/timeline_pass.py)
#define SVM_PH_DECL long long ph_t = clock64(), ph_acc[8] = {0, 0, 0, 0, 0, 0, 0, 0}
#define SVM_PH(k) do { const long long now_ = clock64(); ph_acc[k] += now_ - ph_t; ph_t = now_; } while (0)
#define SVM_PH_FLUSH do { if (a.timing && blockIdx.x == 0 && threadIdx.x == 0 && gather_no < a.timing_cap) for (int k_ = 0; k_ < 8; ++k_) a.timing[(size_t)a.timing_cap * 16 + 8 * 160 * 4 + (size_t)gather_no * 8 + k_] = ph_acc[k_]; } while (0)
#else
#define SVM_PH_DECL do { } while (0)
#define SVM_PH(k) do { } while (0)
#define SVM_PH_FLUSH do { } while (0)
#define SVM_STAMP_T(k, tid) do { } while (0)
#define SVM_STAMP(k) do { } while (0)
#define SVM_GSTAMP(k) do { } while (0)
#endif

constexpr long long PEER_TIMEOUT_CLOCKS = 6000000000LL;   // ~3 s at 1.97 GHz: a peer GPU's process is gone

// Poll `count` records starting at `recs` until every word carries `seq`, and stage their data words in shared memory.
// Called by all warps of a leader CTA; warp w takes the groups of four records w, w + nwarps, ... (8 lanes x 16 bytes
// cover a record, so a warp's load covers four records in one coalesced request); a group that is complete is staged
// and not asked for again.  (Two sets of loads in flight half a round trip apart were tried and were slower: loads that
// poll one line from many SMs serialise in L2, see SMO_REC_COPIES.)
// Returns false if the wait was given up: `abort_word` set by another CTA, or (deadline != 0) the clock ran out, in which
// case this CTA sets the abort word so that every other wait on this GPU ends too.
template <bool SYS, int G>
__device__ __forceinline__ bool poll_records_impl(const Rec *recs, int count, unsigned seq, unsigned (*stage)[REC_STRIDE], unsigned *abort_word,
                                                  long long deadline)
{
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int nwarps = (int)(blockDim.x >> 5);
    const int sub = lane >> 3, pair = lane & 7;
    unsigned long long w0[G], w1[G];
    unsigned pending = 0;                                  // warp-uniform: groups of this warp that are not complete yet
#pragma unroll
    for (int u = 0; u < G; ++u)
        if ((warp + u * nwarps) * 4 < count) pending |= 1u << u;
    unsigned spins = 0;
    while (pending) {
#pragma unroll
        for (int u = 0; u < G; ++u) {
            const int r = (warp + u * nwarps) * 4 + sub;
            if (((pending >> u) & 1u) && r < count) ld_pair<SYS>(&recs[r].w[2 * pair], w0[u], w1[u]);
        }
#pragma unroll
        for (int u = 0; u < G; ++u)
            if ((pending >> u) & 1u) {
                const int r = (warp + u * nwarps) * 4 + sub;
                const bool ok = r >= count || ((unsigned)(w0[u] >> 32) == seq && (unsigned)(w1[u] >> 32) == seq);
                if (__all_sync(FULL, ok)) {
                    if (r < count) { stage[r][2 * pair] = (unsigned)w0[u]; stage[r][2 * pair + 1] = (unsigned)w1[u]; }
                    pending &= ~(1u << u);
                }
            }
        if ((++spins & 127u) == 0) {
            if (ld_relaxed_gpu_u32(abort_word)) return false;
            if (deadline && clock64() > deadline) { st_relaxed_gpu_u32(abort_word, 1u); return false; }
        }
    }
    return true;
}
template <bool SYS>
__device__ __forceinline__ bool poll_records(const Rec *recs, int count, unsigned seq, unsigned (*stage)[REC_STRIDE], unsigned *abort_word,
                                             long long deadline)
{
    // up to three groups per warp with clusters of two CTAs (74 records = 19 groups over 8 warps), one for the peers' mailbox
    // (8 records), five without clusters (148 records)
    const int nwarps = (int)(blockDim.x >> 5);
    if (SYS) return poll_records_impl<SYS, 1>(recs, count, seq, stage, abort_word, deadline);
    if (count <= 12 * nwarps) return poll_records_impl<SYS, 3>(recs, count, seq, stage, abort_word, deadline);
    return poll_records_impl<SYS, POLL_GROUPS_MAX>(recs, count, seq, stage, abort_word, deadline);
}
// every warp folds the staged records by itself: lane l compares (f, idx) of records l, l + 32, ..., one REDUX round
// per side finds the winning record, and every lane reads that record's payload
__device__ __forceinline__ void fold_staged(unsigned (*stage)[REC_STRIDE], int count, LocalBest &g)
{
    const int lane = threadIdx.x & 31;
    double flo = __longlong_as_double(0x7ff0000000000000LL), fhi = __longlong_as_double(0xfff0000000000000LL);
    int ilo = IDX_NONE, ihi = IDX_NONE, rlo = 0, rhi = 0;
    for (int r = lane; r < count; r += 32) {
        const unsigned *d = stage[r];
        const double f0 = __hiloint2double((int)d[1], (int)d[0]), f1 = __hiloint2double((int)d[4], (int)d[3]);
        const int i0 = (int)d[2], i1 = (int)d[5];
        if (better_min(f0, i0, flo, ilo)) { flo = f0; ilo = i0; rlo = r; }
        if (better_max(f1, i1, fhi, ihi)) { fhi = f1; ihi = i1; rhi = r; }
    }
    int src_lo, src_hi;
    warp_arg_best2(flo, ilo, fhi, ihi, src_lo, src_hi);
    rlo = __shfl_sync(FULL, rlo, src_lo);
    rhi = __shfl_sync(FULL, rhi, src_hi);
    const unsigned *a = stage[rlo], *b = stage[rhi];
    g.lo.f = __hiloint2double((int)a[1], (int)a[0]);
    g.lo.idx = (int)a[2];
    g.lo.alpha = __hiloint2double((int)a[7], (int)a[6]);
    g.lo.y = (int)a[10];
    g.lo.q = (long long)(((unsigned long long)a[13] << 32) | a[12]);
    g.hi.f = __hiloint2double((int)b[4], (int)b[3]);
    g.hi.idx = (int)b[5];
    g.hi.alpha = __hiloint2double((int)b[9], (int)b[8]);
    g.hi.y = (int)b[11];
    g.hi.q = (long long)(((unsigned long long)b[15] << 32) | b[14]);
}

// One exchange.  On return every thread of every CTA on every GPU holds the same winners in `mine`.  Returns false
// if the exchange was abandoned (a peer GPU never showed up): all CTAs of this GPU then return false.
__device__ __forceinline__ bool exchange(const SmoArgs &a, Shared &sh, LocalBest &mine, int gather_no)
{
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int nwarps = (int)(blockDim.x >> 5);
    const unsigned cl = cluster_nctarank(), crank = cluster_ctarank();
    const int NC = (int)(gridDim.x / cl);                 // records per exchange
    const int cid = (int)(blockIdx.x / cl);
    const unsigned seq = (unsigned)gather_no + 1u;
    const unsigned parity = (unsigned)gather_no & 1u;
    const bool leader = crank == 0;
    Rec *recs = a.rec + (size_t)(gather_no & 1) * SMO_REC_COPIES * SMO_MAX_RECS;      // [copy][record] of this generation

    // hop 1: every warp's winners.  Hop 2 rides along: the warps of the non-leader CTA write theirs straight into the leader's
    // shared memory (behind the leader's own warps' entries), counted on the leader's mbarrier, so one fold by the leader's warp 0
    // covers the cluster
    best_warp_reduce(mine);
    if (leader) {
        if (lane == 0) { sh.wlo[warp] = mine.lo; sh.whi[warp] = mine.hi; }
    } else if (lane == 0) {
        const uint32_t bar = dsmem_addr(&sh.bar_in, 0);
        dsmem_put_best(dsmem_addr(&sh.wlo[nwarps + warp], 0), mine.lo, bar);
        dsmem_put_best(dsmem_addr(&sh.whi[nwarps + warp], 0), mine.hi, bar);
        if (warp == 0) mbar_expect_tx(&sh.bar_out, BEST_PAIR_BYTES);       // arms this CTA's barrier for hop 4
    }
    SVM_STAMP(3);
    if (leader) {
        __syncthreads();
        SVM_STAMP(4);
        if (warp == 0) {
            bool got = true;
            if (cl > 1) {                                  // the partner's entries
                if (lane == 0) mbar_expect_tx(&sh.bar_in, (unsigned)nwarps * BEST_PAIR_BYTES);
                __syncwarp();
                got = mbar_wait_abortable(&sh.bar_in, parity, a.abort_word);
            }
            SVM_STAMP(8);
            LocalBest b;
            best_init(b);
            if (lane < nwarps || (got && lane < (int)cl * nwarps)) { b.lo = sh.wlo[lane]; b.hi = sh.whi[lane]; }
            best_warp_reduce(b);
            SVM_STAMP(5);
            rec_store_copies(recs, cid, b, seq, lane);                     // hop 3: publish
            SVM_STAMP(6);
        }
    }
    bool ok = true;
    if (leader) {
        // (the deadline: on a GPU shared with other work part of the grid may never run)
        ok = poll_records<false>(recs + (size_t)(cid % SMO_REC_COPIES) * SMO_MAX_RECS, NC, seq, sh.recdata, a.abort_word,
                                 clock64() + PEER_TIMEOUT_CLOCKS);
        SVM_STAMP(11);
        ok = __syncthreads_and(ok ? 1 : 0) != 0;          // every warp's share is staged
        SVM_STAMP(12);
        fold_staged(sh.recdata, NC, mine);
        SVM_STAMP(13);
        if (a.peers > 1) {
            // second level (row sharding): one record per GPU, stored into every peer's mailbox over NVLink (peer-mapped
            // memory, same self-validating words at system scope), polled locally by every leader.  The mailbox is kept in
            // copies like the records: leader c < copies stores copy c everywhere, leader c polls copy c mod copies.
            // A peer that never shows up ends the wait after ~3 s; the abort word then ends every other wait on this GPU.
            const unsigned seq2 = ((unsigned)(a.epoch & 0x7ff) << 21) | seq;
            const int gen = (gather_no & 1) * SMO_MB_COPIES;
            if (cid < a.mailbox_copies && warp == 0) mailbox_store(a.peer_mailbox, a.peers, a.rank, (gen + cid) * 8, mine, seq2, lane);
            const int slot = (gen + cid % a.mailbox_copies) * 8;
            // the peers' records are staged behind the first level's (at most 148 of the 160 places), which slower warps may
            // still be folding: no CTA barrier in between
            unsigned (*stage2)[REC_STRIDE] = sh.recdata + (SMO_MAX_RECS - 8);
            const bool ok2 = poll_records<true>(a.mailbox + slot, a.peers, seq2, stage2, a.abort_word, clock64() + PEER_TIMEOUT_CLOCKS);
            ok = __syncthreads_and((ok && ok2) ? 1 : 0) != 0;
            fold_staged(stage2, a.peers, mine);
        }
        if (cl > 1 && warp == 0 && lane == 0) {           // hop 4, sender
            Best lo = mine.lo, hi = mine.hi;
            if (!ok) lo.y = hi.y = 99;                     // tells the partner that the exchange was abandoned
            const uint32_t bar = dsmem_addr(&sh.bar_out, 1);
            dsmem_put_best(dsmem_addr(&sh.rlo, 1), lo, bar);
            dsmem_put_best(dsmem_addr(&sh.rhi, 1), hi, bar);
        }
    } else {
        // hop 4, receiver: every thread waits for the leader's 64 bytes, then reads the winners
        ok = mbar_wait_abortable(&sh.bar_out, parity, a.abort_word);
        mine.lo = sh.rlo;
        mine.hi = sh.rhi;
        if (mine.lo.y == 99) ok = false;
        if (a.peers > 1) ok = __syncthreads_and(ok ? 1 : 0) != 0;
    }
    SVM_STAMP(7);
    return ok;
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
    // bytes of a row in the ring (uint8 rows: the main class only; the tails have their own array and buffer)
    const int row_bytes = sizeof(XT) == 1 ? a.mw * 4 : a.pitch4 * 16;
    // shared-memory carve-up: [warps][slots] rows | mbarriers | Shared
    unsigned char *slots = smem_raw;
    const int KS = a.kslots;
    uint64_t *bars = reinterpret_cast<uint64_t *>(slots + (size_t)WARPS * KS * smo_rows_per_step<XT, Q, THREADS>() * row_bytes);
    Shared &sh = *reinterpret_cast<Shared *>(bars + WARPS * SMO_KSLOTS_MAX);
    constexpr int RPS = smo_rows_per_step<XT, Q, THREADS>();
    // (RPS == 4) per warp: the two sums of each of the 32 rows of a block, parked until their lane finishes the rows
    double2 *stage_all = reinterpret_cast<double2 *>(reinterpret_cast<unsigned char *>(&sh) + ((sizeof(Shared) + 15) & ~(size_t)15));
    double2 *stage = stage_all + warp * 32;
    // (uint8 rows) per warp: the two integer sums of each of the 32 rows of a block; then the tail weights, shared by the CTA
    constexpr bool IS_U8 = sizeof(XT) == 1;
    constexpr int QM = IS_U8 ? (Q > 0 ? Q : 1) : 1;               // uint8 rows: Q = 128-byte slots of the main class (0 ... 8), exactly
    constexpr int MSLOTS = IS_U8 ? Q : 0;
    double *sh_wt = reinterpret_cast<double *>(reinterpret_cast<int2 *>(stage_all + WARPS * 32) + WARPS * 32);
    U8Lists lists;
    lists.tw = a.tw;
    lists.cap = a.tw + 4;
    lists.pp = reinterpret_cast<double2 *>(sh_wt + a.tw * 4);
    lists.kk = reinterpret_cast<unsigned short *>(lists.pp + lists.cap * 4);
    lists.xlt = reinterpret_cast<unsigned char *>(lists.kk + lists.cap * 4);
    // (uint8 rows) per warp: the tail bytes of a block of a.tail_rows rows, brought by one bulk copy from the tail array and
    // walked a lane per row; the odd pitch (in 16-byte units) spreads the rows over the banks
    unsigned char *tailbuf = reinterpret_cast<unsigned char *>((reinterpret_cast<uintptr_t>(lists.xlt + a.tw * 4) + 127) & ~(uintptr_t)127) +
                             (size_t)warp * a.tail_rows * a.tail_pitch;
    const unsigned char *Xt = Xb + (size_t)a.n * row_bytes;        // (uint8 rows) the tail array follows the main-class array
    if constexpr (IS_U8)
        for (int k = threadIdx.x; k < a.tw * 4; k += THREADS) sh_wt[k] = a.wt[k];

    // rows owned by this CTA and, inside it, by this warp (contiguous ranges)
    const long long n_own = a.row_end - a.row_begin;
    const long long r0 = a.row_begin + (long long)blockIdx.x * n_own / G;
    const long long r1 = a.row_begin + (long long)(blockIdx.x + 1) * n_own / G;
    const long long R = r1 - r0;
    // warps take contiguous ranges, cut at multiples of RPS rows from r0 (so only the CTA's last rows form a short group)
    // (uint8 rows: cut at any row -- a warp's last, short group costs half a group when it has up to four rows, so equal row
    // counts mean equal work)
    constexpr int CUT = sizeof(XT) == 1 ? 1 : RPS;
    const long long units = (R + CUT - 1) / CUT;
    const long long w0 = min(r1, r0 + CUT * ((long long)warp * units / WARPS));
    const long long w1 = min(r1, r0 + CUT * ((long long)(warp + 1) * units / WARPS));
    const int m = (int)(w1 - w0);
    const int keep_rows = (int)(((long long)m * a.keep_per_1024) >> 10);

    unsigned char *my_slots = slots + (size_t)warp * KS * RPS * row_bytes;
    uint64_t *my_bars = bars + warp * SMO_KSLOTS_MAX;
    if (lane == 0) {
        for (int s = 0; s < KS; ++s) mbar_init(&my_bars[s], 1);
        if (IS_U8) mbar_init(&my_bars[SMO_KSLOTS_MAX - 1], 1);       // the tail blocks' barrier (the ring uses at most four of the eight)
        if (warp == 0) { mbar_init(&sh.bar_in, 1); mbar_init(&sh.bar_out, 1); }
        mbar_fence_init();
    }
    __syncthreads();
    cluster_sync_all();                  // the partner CTA's mbarriers exist before anything is sent to them

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
    if (m > 0 && row_bytes > 0)
        for (int s = 0; s < KS; ++s) issue_into(s);
    // (uint8 rows) the tails: blocks of a.tail_rows consecutive rows of the warp's range, requested in order, cyclically, into
    // the one buffer -- the next block is requested when the block at hand has been walked
    uint64_t *tail_bar = &my_bars[SMO_KSLOTS_MAX - 1];
    unsigned tphase = 0;
    int tf_row = 0;                      // next row of the range whose tail to request
    auto issue_tail = [&]() {
        const int nrows = min(a.tail_rows, m - tf_row);
        if (lane == 0) {
            const int c = tf_row / a.tail_rows;
            const bool keep = (((c + 1) * a.keep_per_1024) >> 10) != ((c * a.keep_per_1024) >> 10);
            mbar_expect_tx(tail_bar, (unsigned)(nrows * a.tail_pitch));
            bulk_g2s(tailbuf, Xt + (size_t)(w0 + tf_row) * a.tail_pitch, (unsigned)(nrows * a.tail_pitch), tail_bar,
                     keep ? a.policy_keep : a.policy_stream);
        }
        tf_row += nrows;
        if (tf_row >= m) tf_row = 0;
    };
    if constexpr (IS_U8)
        if (m > 0 && a.tw > 0) issue_tail();

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

    Best lo, hi;                         // the working pair with what the scalar update needs of its rows
    lo.f = hi.f = lo.alpha = hi.alpha = 0.0; lo.q = hi.q = -1; lo.idx = hi.idx = IDX_NONE; lo.y = hi.y = 0;
    bool gain = false;                   // wss = 2: the pass at hand only chooses i_low (see gain_consider)
    for (;;) {
        if (!exchange(a, sh, mine, gather_no)) { stop = 5; break; }                 // the one grid-wide synchronisation
        if (!gain) {
            // the exchange that followed an update (or the start): i_high, the first-order i_low, the stop test
            if (num_iter > a.max_iter) { stop = 4; break; }                         // main3.cpp:283 (b keeps this iteration's b_high/b_low)
            lo = mine.lo; hi = mine.hi;
            SVM_STAMP(0);
            SVM_GSTAMP(0);
            i_high = lo.idx == IDX_NONE ? a.n : lo.idx;
            i_low = hi.idx == IDX_NONE ? a.n : hi.idx;
            if (lo.idx == IDX_NONE || hi.idx == IDX_NONE) { stop = 1; break; }      // main3.cpp:205
            b_high = lo.f;
            b_low = hi.f;
            if (b_low <= b_high + 2.0 * a.tau) { stop = 0; break; }                 // main3.cpp:213
            gain = a.wss == 2;                                                       // second-order: first a pass that chooses i_low
        } else {
            // the exchange that followed a gain pass: the row it elected replaces the first-order i_low (if any row qualified)
            gain = false;
            if (mine.hi.idx != IDX_NONE) {
                hi = mine.hi;
                hi.f = mine.lo.alpha;                                                // f of that row travelled in the other side
                i_low = hi.idx;
                b_low = hi.f;
            }
        }
        // in a gain pass both pivots are row i_high
        const long long piv_l = gain ? i_high : i_low;

        // row-cache lookup (FIFO by allocation number; validity needs no notification)
        bool hit_h = false, hit_l = false;
        long long q_h = lo.q, q_l = gain ? lo.q : hi.q;
        if (a.cache_cap > 0) {
            hit_h = q_h >= 0 && cache_counter - q_h < a.cache_cap;
            hit_l = q_l >= 0 && cache_counter - q_l < a.cache_cap;
        }
        const bool use_cache = hit_h && hit_l;

        // cached iteration: K12 is column i_low of the cached row of i_high -- the one read of a value another CTA wrote: it was
        // stored by the CTA that owns row i_low, at least one exchange ago (its __threadfence in finish_row, this thread's fence
        // here after the records it observed).  Read from L2: this SM's L1 may still hold the line from when the slot cached
        // another row.  ONE thread of the CTA reads it (loads of one line from many SMs and warps serialise in L2: with every
        // warp reading, this step cost 2 us of a 6 us iteration) and hands it to the others through shared memory, behind the
        // CTA barrier below; its fence comes first, while it has nothing in flight, and the K12 load and every thread's first
        // rows of the pass are requested together instead of one L2 round trip after the other.
        const bool local_low = i_low >= a.row_begin && i_low < a.row_end;
        const bool k12_cached = use_cache && !gain && local_low;            // uniform over the CTA (replicated state)
        double K12c = 1.0;
        if (k12_cached && threadIdx.x == 0) {
            __threadfence();
            sh.k12 = __ldcg(&a.cacheK[(size_t)(q_h % a.cache_cap) * a.n + i_low]);
        }
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

        if (k12_cached) {
            __syncthreads();                       // the next write of sh.k12 is at least one exchange (a CTA barrier) away
            K12c = sh.k12;
        }

        // K12 = K(i_high, i_low)  (main3.cpp:241)
        Pivots<Vec, (Q > 0 ? Q : 1)> pv;
        double K12, n_h = 0.0, n_l = 0.0;
        // uint8 rows: the pivots' main-class words in registers, the non-zero positions of their tails in the sub-lists
        unsigned pm_h[QM], pm_l[QM];
        int lcnt = 0;                                           // length of the sub-list this lane walks
        int nm_h = 0, nm_l = 0;
        if constexpr (IS_U8) {
          // pivots are needed for an X pass, and for K12 of a cached iteration whose row i_low lives on another GPU (row sharding:
          // X is replicated, so K12 is formed here the way that GPU formed it)
          if (!use_cache || (!gain && !local_low)) {
            const unsigned *xh = reinterpret_cast<const unsigned *>(Xb + (size_t)i_high * row_bytes);
            const unsigned *xl = reinterpret_cast<const unsigned *>(Xb + (size_t)piv_l * row_bytes);
            const unsigned *th = reinterpret_cast<const unsigned *>(Xt + (size_t)i_high * a.tail_pitch);
            const unsigned *tl = reinterpret_cast<const unsigned *>(Xt + (size_t)piv_l * a.tail_pitch);
#pragma unroll
            for (int i = 0; i < MSLOTS; ++i) {                  // a.mw == 32 * MSLOTS
                pm_h[i] = __ldg(xh + i * 32 + lane);
                pm_l[i] = __ldg(xl + i * 32 + lane);
            }
            nm_h = a.nmain[i_high]; nm_l = a.nmain[piv_l];
            n_h = a.ntail[i_high]; n_l = a.ntail[piv_l];
            lcnt = u8_build_lists(th, tl, sh_wt, lists, lane);
            SVM_STAMP(15);
            int dot = 0;
#pragma unroll
            for (int i = 0; i < MSLOTS; ++i) dot = __dp4a(pm_l[i], pm_h[i], (unsigned)dot);
            dot = __reduce_add_sync(FULL, dot);
            double dh, dl;
            u8_tail_dots(lists.xlt, lcnt, lists, lane, dh, dl);             // row = x_low, pivot = x_high
            const int S = nm_l + nm_h - 2 * dot;
            double tail = (n_l + n_h) - 2.0 * dh;
            if (tail < 0.0) tail = 0.0;
            K12 = svm_exp(-a.gamma * fma((double)S, a.wR, tail));
          } else {
            K12 = K12c;                                         // a gain pass makes no step: its K12 (1.0) is not used
          }
        } else
        if (!use_cache) {
            const Vec *xh = reinterpret_cast<const Vec *>(a.X) + (size_t)i_high * a.pitch4;
            const Vec *xl = reinterpret_cast<const Vec *>(a.X) + (size_t)piv_l * a.pitch4;
#pragma unroll
            for (int i = 0; i < Q; ++i) {
                const int q = i * 32 + lane;
                pv.h[i] = q < a.pitch4 ? __ldg(xh + q) : vec_zero<Vec>();
                pv.l[i] = q < a.pitch4 ? __ldg(xl + q) : vec_zero<Vec>();
            }
            if (sizeof(XT) == 4) { n_h = a.sqn[i_high]; n_l = a.sqn[piv_l]; }
            SVM_STAMP(15);
            double acc = 0.0;
#pragma unroll
            for (int i = 0; i < Q; ++i)
                if (i * 32 + lane < a.pitch4) accum_pair<CH>(pv.l[i], pv.h[i], acc);
            acc = warp_sum_butterfly(acc);
            K12 = svm_exp(-a.gamma * dist2<XT>(n_h, n_l, acc));
        } else if (gain || local_low) {
            K12 = K12c;                                             // a gain pass makes no step: its K12 (1.0) is not used
        } else {
            // row sharding: column i_low of the cached rows lives on the GPU that owns row i_low; X is
            // replicated, so form K12 here the way that GPU formed it (same code, same bits)
            const Vec *xh = reinterpret_cast<const Vec *>(a.X) + (size_t)i_high * a.pitch4;
            const Vec *xl = reinterpret_cast<const Vec *>(a.X) + (size_t)piv_l * a.pitch4;
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
        IterState it;
        it.gain = gain; it.b_high = b_high; it.i_high = i_high; it.i_low = i_low;
        it.c_h = it.c_l = 0.0; it.na_h = lo.alpha; it.na_l = hi.alpha;
        if (!gain) {
            const PairStep ps = pair_update(lo.y, hi.y, lo.alpha, hi.alpha, b_high, b_low, K12, a.C, a.eps);
            if (ps.stop) { stop = ps.stop; break; }             // main3.cpp:246 / :253
            it.na_h = ps.na_h; it.na_l = ps.na_l; it.c_h = ps.c_h; it.c_l = ps.c_l;
        }

        if (!gain && blockIdx.x == 0 && threadIdx.x == 0 && a.rank == 0 && num_iter - 1 < a.trace_cap) {
            a.trace[2 * (num_iter - 1)] = (int)i_high;
            a.trace[2 * (num_iter - 1) + 1] = (int)i_low;
        }

        // allocate cache slots for the rows about to be computed
        int slot_h = -1, slot_l = -1;
        if (a.cache_cap > 0) {
            if (!use_cache) {
                if (!hit_h) q_h = cache_counter++;
                if (gain) q_l = q_h;                          // one row, computed twice over
                else if (!hit_l) q_l = cache_counter++;
            }
            slot_h = (int)(q_h % a.cache_cap);
            slot_l = (int)(q_l % a.cache_cap);
            if (gain) { hit_l = true; lo.q = q_h; }           // nothing is stored for the copy; the update that follows finds row i_high cached
        }
        it.q_h = q_h; it.q_l = q_l; it.slot_h = slot_h; it.slot_l = slot_l; it.hit_h = hit_h; it.hit_l = hit_l;

        LocalBest acc_best;
        best_init(acc_best);
        SVM_STAMP(1);
        SVM_GSTAMP(1);
        if (!use_cache) {
            // ---------------- X pass: a warp per group of RPS rows (or per row, below) ----------------
            ++x_passes;
            if constexpr (IS_U8) {
              // ---------------- uint8 rows: a warp per group of eight rows ----------------
              // both sections of a row are padded with zero bytes to whole 128-byte slots (32 lanes x one word), so a lane's
              // words sit at constant offsets from one address per row and need no bounds test
              const bool has_tail = a.tw > 0;
              const int TB = a.tail_rows;
              const int n0 = __shfl_sync(FULL, lcnt, 0), n1 = __shfl_sync(FULL, lcnt, 1), n2 = __shfl_sync(FULL, lcnt, 2), n3 = __shfl_sync(FULL, lcnt, 3);
              SVM_PH_DECL;
              for (int base = 0; base < m; base += 32) {
                const int cnt = min(32, m - base);
                const long long j = w0 + base + lane;
                double fj = 0.0, aj = 0.0, ntj = 0.0;
                int yj = 0, nmj = 0;
                long long qj = -1;
                if (lane < cnt) { fj = a.f[j]; aj = a.alpha[j]; yj = a.y[j]; nmj = a.nmain[j]; ntj = a.ntail[j]; if (a.cacheq) qj = a.cacheq[j]; }
                int my_sh = 0, my_sl = 0;                  // the integer dot products of this lane's row of the block
                double2 dd = make_double2(0.0, 0.0);       // D(row, pivots) over the tail of this lane's row
                for (int t = 0; t < cnt; t += RPS) {
                    const int rows_here = min(RPS, cnt - t);
                    if constexpr (MSLOTS > 0) {
                        SVM_PH(5);
                        mbar_wait(&my_bars[cslot], cphase);
                        SVM_PH(0);
                        const unsigned char *grp = my_slots + (size_t)cslot * RPS * row_bytes;
                        int SI[2 * RPS];                   // [pivot * RPS + row] integer dot products over the main class
#pragma unroll
                        for (int k = 0; k < 2 * RPS; ++k) SI[k] = 0;
                        // main class: eight rows side by side (sixteen independent sums), or four when the group is the short one that
                        // ends the warp's range: rows beyond the group's last hold bytes of an earlier group, their sums are never read
                        const unsigned char *lanep = grp + lane * 4;
                        if (rows_here > 4) u8_main_sums<QM, MSLOTS, RPS, 8>(lanep, row_bytes, pm_h, pm_l, SI);
                        else u8_main_sums<QM, MSLOTS, RPS, 4>(lanep, row_bytes, pm_h, pm_l, SI);
                        SVM_PH(1);
                        __syncwarp();
                        fence_proxy_async();
                        issue_into(cslot);
                        if (++cslot == KS) { cslot = 0; cphase ^= 1u; }
                        SVM_PH(3);
                        // the integer sums are exact in any order: the hardware warp reduction, the row's own lane keeps the result
#pragma unroll
                        for (int r = 0; r < RPS; ++r) {
                            const int sh_ = __reduce_add_sync(FULL, SI[r]), sl_ = __reduce_add_sync(FULL, SI[RPS + r]);
                            if (lane == t + r) { my_sh = sh_; my_sl = sl_; }
                        }
                        SVM_PH(4);
                    }
                    // tail: when the rows of a tail block are through, a lane per row walks the four sub-lists one after the other
                    // (uniform trip counts, broadcast reads of the entries), then the buffer is handed back for the next block
                    const int done = t + rows_here;
                    if (has_tail && ((done & (TB - 1)) == 0 || done == cnt)) {
                        const int c0 = (done - 1) & ~(TB - 1);         // first row of the tail block within the 32-row block
                        mbar_wait(tail_bar, tphase);
                        tphase ^= 1u;
                        SVM_PH(2);
                        if (lane < done - c0) {
                            const unsigned char *myrow = tailbuf + (size_t)lane * a.tail_pitch;
                            double sh_[4] = {0.0, 0.0, 0.0, 0.0}, sl_[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll
                            for (int c = 0; c < 4; ++c)                 // the first four entries of every sub-list, no branch between them
                                u8_walk4(myrow, lists.pp + c * lists.cap, lists.kk + c * lists.cap, 0, sh_[c], sl_[c]);
#pragma unroll
                            for (int c = 0; c < 4; ++c) {
                                const int n = c == 0 ? n0 : c == 1 ? n1 : c == 2 ? n2 : n3;
                                for (int e = 4; e < n; e += 4) u8_walk4(myrow, lists.pp + c * lists.cap, lists.kk + c * lists.cap, e, sh_[c], sl_[c]);
                            }
                            dd = make_double2((sh_[0] + sh_[1]) + (sh_[2] + sh_[3]), (sl_[0] + sl_[1]) + (sl_[2] + sl_[3]));
                            if (TB < 32) stage[c0 + lane] = dd;        // parked until the row's own lane finishes the row
                        }
                        __syncwarp();
                        fence_proxy_async();
                        issue_tail();
                        SVM_PH(6);
                    }
                }
                __syncwarp();
                if (has_tail && TB < 32 && lane < cnt) dd = stage[lane];
                if (lane < cnt) {
                    double t_h = (ntj + n_h) - 2.0 * dd.x, t_l = (ntj + n_l) - 2.0 * dd.y;
                    if (t_h < 0.0) t_h = 0.0;
                    if (t_l < 0.0) t_l = 0.0;
                    const double kh = svm_exp(-a.gamma * fma((double)(nmj + nm_h - 2 * my_sh), a.wR, t_h));
                    const double kl = svm_exp(-a.gamma * fma((double)(nmj + nm_l - 2 * my_sl), a.wR, t_l));
                    finish_row(a, it, j, fj, aj, yj, qj, kh, kl, acc_best);
                }
                __syncwarp();                                  // the parked sums have been read: the next block may overwrite them
                SVM_PH(7);
              }
              SVM_PH_FLUSH;
            } else if constexpr (RPS > 1) {
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
                    finish_row(a, it, j, fj, aj, yj, qj, kh, kl, acc_best);
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
                    finish_row(a, it, j, fj, aj, yj, qj, kh, kl, acc_best);
                }
            }
        } else {
            // ---------------- cached pass: a thread per row over two cached kernel rows ----------------
            if (!gain) ++cache_hits;
            const double *kh_row = a.cacheK + (size_t)slot_h * a.n;
            const double *kl_row = a.cacheK + (size_t)slot_l * a.n;
            for (long long j = pj; j < r1; j += THREADS) {
                const bool first = j == pj, second = j == pj2;
                double fj = first ? p_f : second ? p2_f : a.f[j], aj = first ? p_a : second ? p2_a : a.alpha[j];
                const int yj = first ? p_y : second ? p2_y : a.y[j];
                const long long qj = first ? p_q : second ? p2_q : a.cacheq[j];
                const double khj = first ? p_kh : second ? p2_kh : kh_row[j], klj = first ? p_kl : second ? p2_kl : kl_row[j];
                finish_row(a, it, j, fj, aj, yj, qj, khj, klj, acc_best);
            }
        }

        SVM_STAMP(2);
        SVM_GSTAMP(2);
        ++gather_no;
        mine = acc_best;
        if (!gain) ++num_iter;                                   // main3.cpp:281
    }

    // drain the ring: every requested copy must land before the CTA may exit
    if (m > 0 && row_bytes > 0)
        for (int k = 0; k < KS; ++k) {
            mbar_wait(&my_bars[cslot], cphase);
            if (++cslot == KS) { cslot = 0; cphase ^= 1u; }
        }
    if constexpr (IS_U8)
        if (m > 0 && a.tw > 0) mbar_wait(tail_bar, tphase);      // and the tail block requested last
    cluster_sync_all();                  // no CTA leaves while its partner may still write into its shared memory
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        SmoResult r;
        r.num_iter = num_iter; r.stop_reason = stop;
        r.i_high = i_high; r.i_low = i_low;
        r.b_high = b_high; r.b_low = b_low;
        r.x_passes = x_passes; r.cache_hits = cache_hits; r.cache_counter = cache_counter;
        *a.result = r;
    }
}

// ---- row sharding: the GPUs meet before the timed launch.  Thread p stores this launch's number into GPU p's flag
// [rank] (peer-mapped memory) and waits until GPU p's number shows in its own flag [p]; gives up after ~3 s (the SMO
// kernel that follows then ends with PEER_LOST by itself).
struct PeerFlags { unsigned long long *peer[8]; };
__global__ void peer_rendezvous_kernel(PeerFlags pf, int rank, int peers, unsigned long long epoch)
{
    const int p = threadIdx.x;
    if (p >= peers || !pf.peer[p]) return;
    asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(pf.peer[p] + rank), "l"(epoch) : "memory");
    const long long deadline = clock64() + PEER_TIMEOUT_CLOCKS;
    for (;;) {
        unsigned long long v;
        asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(pf.peer[rank] + p) : "memory");
        if (v >= epoch || clock64() > deadline) break;
    }
}

// kslots: ring depth per warp in slots; a slot holds rows_per_step rows of pitch4 16-byte groups (uint8 rows: of the main class)
inline size_t smo_smem_bytes(int pitch4, int kslots, int warps, int rows_per_step = 1, int tail_words = 0, int tail_pitch = 0, int tail_rows = 32)
{
    return (size_t)warps * kslots * rows_per_step * pitch4 * 16 + sizeof(uint64_t) * warps * SMO_KSLOTS_MAX + sizeof(Shared) + 128 +
           (size_t)warps * 32 * (sizeof(double2) + sizeof(int2)) +   // + the parked sums of the rows-in-groups passes
           (size_t)tail_words * 4 * (sizeof(double) + 1) + (size_t)(tail_words + 4) * 4 * (sizeof(double2) + sizeof(unsigned short)) + 32 +  // + the tail weights and the
                                                                     //   sub-lists of the pivots' non-zero tail positions (uint8 rows)
           (tail_pitch ? 128 + (size_t)warps * tail_rows * tail_pitch : 0);   // + the buffers of the tail blocks
}

} // namespace svm
