// device_math.cuh -- arithmetic shared by every kernel of the path.
//
// The RBF kernel value replaces the reference's kernel() (code/main3.cpp:92-104).  The reference
// sums 784 FP64 squared differences sequentially; here K(i,j) = exp(-gamma * max(0, |xi|^2 + |xj|^2
// - 2 xi.xj)) with the dot product formed by a warp: lane l takes the float4 groups q = l, l+32, ...
// of the fp32 rows, sums the products of CH consecutive elements in fp32 (fmaf), widens to FP64 and
// accumulates per lane, then the 32 lane sums are combined with an xor-butterfly (16,8,4,2,1); the
// squared norms are dot products of a row with itself.  That order is fixed, so K(i,j) is
// bit-reproducible and the CPU checker (oracle/smo_oracle.c, ORACLE_ARITH_DEVICE) restates it op for op.
//
// Built with -fmad=false: every contraction in this library is an explicit fma()/fmaf(), so the
// FP64 scalar update rounds exactly like the reference's x86-64 build (no FMA).
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace svm {

constexpr unsigned FULL = 0xffffffffu;

// exp(x) from IEEE-exact primitives only (rint, fma, mul): identical bits on host and device.
__device__ __forceinline__ double svm_exp(double x)
{
    if (x != x) return x;
    if (x > 709.782712893384) return __longlong_as_double(0x7ff0000000000000LL);
    if (x < -745.2) return 0.0;
    const double LOG2E = 1.4426950408889634;
    const double LN2_HI = 6.93147180369123816490e-01;
    const double LN2_LO = 1.90821492927058770002e-10;
    double kf = rint(x * LOG2E);
    double r = fma(-kf, LN2_HI, x);
    r = fma(-kf, LN2_LO, r);
    double p = 1.0 / 6227020800.0;
    p = fma(p, r, 1.0 / 479001600.0);
    p = fma(p, r, 1.0 / 39916800.0);
    p = fma(p, r, 1.0 / 3628800.0);
    p = fma(p, r, 1.0 / 362880.0);
    p = fma(p, r, 1.0 / 40320.0);
    p = fma(p, r, 1.0 / 5040.0);
    p = fma(p, r, 1.0 / 720.0);
    p = fma(p, r, 1.0 / 120.0);
    p = fma(p, r, 1.0 / 24.0);
    p = fma(p, r, 1.0 / 6.0);
    p = fma(p, r, 0.5);
    p = fma(p, r, 1.0);
    p = fma(p, r, 1.0);
    int k = (int)kf;
    int k1 = k / 2, k2 = k - k1;
    double s1 = __longlong_as_double((long long)(k1 + 1023) << 52);
    double s2 = __longlong_as_double((long long)(k2 + 1023) << 52);
    return (p * s1) * s2;
}

// products of one float4 group of a row with the same group of a pivot row, added to the lane's
// FP64 sum.  CH = how many fp32 products are summed in fp32 before widening (CH = 1: every product
// is exact in FP64).
template <int CH>
__device__ __forceinline__ void accum_dot(const float4 &v, const float4 &p, double &acc)
{
    if (CH == 4) {
        float s = v.x * p.x;
        s = fmaf(v.y, p.y, s);
        s = fmaf(v.z, p.z, s);
        s = fmaf(v.w, p.w, s);
        acc = acc + (double)s;
    } else if (CH == 2) {
        float s = v.x * p.x;
        s = fmaf(v.y, p.y, s);
        acc = acc + (double)s;
        float t = v.z * p.z;
        t = fmaf(v.w, p.w, t);
        acc = acc + (double)t;
    } else {
        acc = fma((double)v.x, (double)p.x, acc);
        acc = fma((double)v.y, (double)p.y, acc);
        acc = fma((double)v.z, (double)p.z, acc);
        acc = fma((double)v.w, (double)p.w, acc);
    }
}

__device__ __forceinline__ double warp_sum_butterfly(double v)
{
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) v = v + __shfl_xor_sync(FULL, v, off);
    return v;
}

// Two butterflies for the price of one and a bit: after the first exchange the lower half-warp
// carries value a and the upper half value b; each half then finishes its own butterfly.  Every lane
// forms exactly the sums the plain xor-butterfly forms (x + y == y + x), so the results are the same
// bits.  On return a and b hold the two totals in every lane.
__device__ __forceinline__ void warp_sum_butterfly2(double &a, double &b, int lane)
{
    const bool upper = (lane & 16) != 0;
    const double send = upper ? a : b;
    const double keep = upper ? b : a;
    double v = keep + __shfl_xor_sync(FULL, send, 16);
#pragma unroll
    for (int off = 8; off >= 1; off >>= 1) v = v + __shfl_xor_sync(FULL, v, off);
    const double other = __shfl_xor_sync(FULL, v, 16);
    a = upper ? other : v;
    b = upper ? v : other;
}

// squared distance from squared norms and the dot product, clamped at 0 (duplicates give exactly 0:
// the norm of a row is its dot product with itself, formed by the same code)
__device__ __forceinline__ double sqdist_from_dot(double na, double nb, double ab)
{
    const double s = (na + nb) - 2.0 * ab;
    return s < 0.0 ? 0.0 : s;
}

// ---- the two storage precisions of X.  A row is a sequence of 16-byte groups either way: four fp32 values
// (the fast path: norm + dot product) or two FP64 values (the exact path: squared differences in FP64, the
// reference's own formula, main3.cpp:97-100, summed per lane instead of sequentially).
template <typename XT> struct RowVec;
template <> struct RowVec<float> { using type = float4; };
template <> struct RowVec<double> { using type = double2; };
template <> struct RowVec<unsigned char> { using type = uint4; };       // uint8 rows are walked in 4-byte words by code of their own
template <typename V> __device__ __forceinline__ V vec_zero();
template <> __device__ __forceinline__ float4 vec_zero<float4>() { return make_float4(0.f, 0.f, 0.f, 0.f); }
template <> __device__ __forceinline__ double2 vec_zero<double2>() { return make_double2(0.0, 0.0); }
template <> __device__ __forceinline__ uint4 vec_zero<uint4>() { return make_uint4(0u, 0u, 0u, 0u); }

template <int CH>
__device__ __forceinline__ void accum_pair(const float4 &v, const float4 &p, double &acc) { accum_dot<CH>(v, p, acc); }
template <int CH>
__device__ __forceinline__ void accum_pair(const double2 &v, const double2 &p, double &acc)
{
    const double d0 = v.x - p.x;
    acc = fma(d0, d0, acc);
    const double d1 = v.y - p.y;
    acc = fma(d1, d1, acc);
}
// squared distance from what accum_pair summed: fp32 rows give the dot product (norms needed), FP64 rows the distance itself
template <typename XT>
__device__ __forceinline__ double dist2(double na, double nb, double s)
{
    if (sizeof(XT) == 4) return sqdist_from_dot(na, nb, s);
    return s;
}

// (value, lowest index) ordering of calc_i_high / calc_i_low (code/main3.cpp:113-120,132-139):
// the reference scans ascending with a strict compare, so among equal f the lowest index wins.
__device__ __forceinline__ bool better_min(double fa, int ia, double fb, int ib)
{
    return fa < fb || (fa == fb && ia < ib);
}
__device__ __forceinline__ bool better_max(double fa, int ia, double fb, int ib)
{
    return fa > fb || (fa == fb && ia < ib);
}

// ---- mbarrier + bulk-copy (TMA, UBLKCP) primitives ----
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t *bar, unsigned count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, unsigned bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, unsigned parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra WAIT_DONE;\n"
        "bra WAIT_LOOP;\n"
        "WAIT_DONE:\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
// global -> shared bulk copy completing on an mbarrier; bytes and both addresses are multiples of 16
__device__ __forceinline__ void bulk_g2s(void *dst_smem, const void *src_gmem, unsigned bytes, uint64_t *bar, uint64_t policy)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
                 ::"r"(smem_u32(dst_smem)), "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar)), "l"(policy) : "memory");
}
constexpr uint64_t L2_EVICT_NORMAL = 0x1000000000000000ull;
constexpr uint64_t L2_EVICT_FIRST = 0x12F0000000000000ull;
constexpr uint64_t L2_EVICT_LAST = 0x14F0000000000000ull;

// release/acquire at GPU scope for the per-iteration record exchange
__device__ __forceinline__ void st_release_gpu(int *p, int v) { asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ int ld_acquire_gpu(const int *p)
{
    int v;
    asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys(int *p, int v) { asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ int ld_acquire_sys(const int *p)
{
    int v;
    asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

} // namespace svm
