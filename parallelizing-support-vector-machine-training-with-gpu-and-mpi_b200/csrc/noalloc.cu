// noalloc.cu -- per-feature min/max of a uint8 training matrix for callers that must not allocate.
//
// Who this is for: code that runs where device allocation is off limits -- a plugin called from
// inside another framework's stream-ordered region, a host whose memory pool owns every byte of the
// GPU and forbids cudaMalloc behind its back, a real-time loop where an allocation (which may
// synchronise the whole device) is not acceptable.  Such a caller already holds device buffers for
// its data and its results; what it cannot provide on the spot is the scratch this reduction needs
// between its two kernels.  So the scratch lives in this library's image instead: one fixed table in
// device memory, present on every GPU that loads libsvmb200.so from load until the process exits.
//
// The regular path (svmb200_set_train_u8 in capi.cu) sizes its scratch to the problem with
// cudaMalloc; this variant runs the same two-step reduction over a table of fixed size and therefore
// has a hard limit on the number of features.  Nothing here calls cudaMalloc, cudaFree, or any
// library routine that allocates.
#include "../../include/svmb200.h"

#include <cstdio>
#include <cuda_runtime.h>

namespace {

constexpr int NOALLOC_MAX_D = SVMB200_NOALLOC_MAX_FEATURES;   // widest matrix accepted
constexpr int NOALLOC_PARTS = 64;                              // row slabs reduced independently
constexpr int MAX_DEVICES = 64;

// [2][NOALLOC_PARTS][NOALLOC_MAX_D] doubles = 1 MiB: slab minima, then slab maxima.
static __device__ __align__(256) double g_minmax_scratch[2 * NOALLOC_PARTS * NOALLOC_MAX_D];

// In host code the name above is only a handle; the address the kernels need is looked up once per
// device (with that device current) and remembered here.
double *g_scratch_addr[MAX_DEVICES] = {};

// slab p scans rows [p * rows_per_part, ...); thread c takes column c
__global__ void minmax_slabs_kernel(const uint8_t *__restrict__ X, long long n, int d, int rows_per_part,
                                    double *__restrict__ scratch)
{
    double *pmin = scratch;
    double *pmax = scratch + (size_t)NOALLOC_PARTS * NOALLOC_MAX_D;
    const long long r0 = (long long)blockIdx.x * rows_per_part;
    const long long r1 = min(n, r0 + rows_per_part);
    for (int c = threadIdx.x; c < d; c += blockDim.x) {
        double lo = __longlong_as_double(0x7ff0000000000000LL), hi = __longlong_as_double(0xfff0000000000000LL);
        for (long long r = r0; r < r1; ++r) {
            const double v = (double)X[r * d + c];
            lo = v < lo ? v : lo;
            hi = hi < v ? v : hi;
        }
        pmin[(size_t)blockIdx.x * NOALLOC_MAX_D + c] = lo;
        pmax[(size_t)blockIdx.x * NOALLOC_MAX_D + c] = hi;
    }
}

// folds the `parts` slabs this call wrote; slabs beyond `parts` may hold an earlier call's values and are never read
__global__ void minmax_fold_kernel(const double *__restrict__ scratch, int parts, int d, double *__restrict__ minv,
                                   double *__restrict__ maxv)
{
    const double *pmin = scratch;
    const double *pmax = scratch + (size_t)NOALLOC_PARTS * NOALLOC_MAX_D;
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= d) return;
    double lo = pmin[c], hi = pmax[c];
    for (int p = 1; p < parts; ++p) {
        const double a = pmin[(size_t)p * NOALLOC_MAX_D + c], b = pmax[(size_t)p * NOALLOC_MAX_D + c];
        lo = a < lo ? a : lo;
        hi = hi < b ? b : hi;
    }
    minv[c] = lo;
    maxv[c] = hi;
}

int fail(int code, const char *what, cudaError_t e)
{
    if (e != cudaSuccess) fprintf(stderr, "svmb200_minmax_u8_noalloc: %s: %s\n", what, cudaGetErrorString(e));
    else fprintf(stderr, "svmb200_minmax_u8_noalloc: %s\n", what);
    return code;
}

} // namespace

extern "C" int svmb200_minmax_u8_noalloc(int device, const uint8_t *dev_X, int64_t n, int32_t d, double *dev_min,
                                         double *dev_max, void *stream)
{
    if (!dev_X || !dev_min || !dev_max || n <= 0 || d <= 0) return fail(SVMB200_EINVAL, "NULL pointer or non-positive size", cudaSuccess);
    if (d > NOALLOC_MAX_D) {
        fprintf(stderr, "svmb200_minmax_u8_noalloc: d=%d exceeds the fixed scratch table (at most %d features); "
                        "use svmb200_set_train_u8, which allocates scratch to size\n", (int)d, NOALLOC_MAX_D);
        return SVMB200_ELIMIT;
    }
    if (device < 0 || device >= MAX_DEVICES) return fail(SVMB200_EINVAL, "device ordinal out of range", cudaSuccess);

    int prev = -1;
    cudaError_t e = cudaGetDevice(&prev);
    if (e != cudaSuccess) return fail(SVMB200_ENODEVICE, "cudaGetDevice", e);
    if ((e = cudaSetDevice(device)) != cudaSuccess) return fail(SVMB200_ECUDA, "cudaSetDevice", e);

    int rc = SVMB200_OK;
    if (!g_scratch_addr[device]) {
        // set-up, once per device: resolve the table's device address while that device is current
        double *p = nullptr;
        e = cudaGetSymbolAddress((void **)&p, g_minmax_scratch);
        if (e != cudaSuccess || !p) rc = fail(SVMB200_ECUDA, "cudaGetSymbolAddress", e);
        else g_scratch_addr[device] = p;
    }
    if (rc == SVMB200_OK) {
        double *scratch = g_scratch_addr[device];
        cudaStream_t s = static_cast<cudaStream_t>(stream);
        const int parts = (int)((n + 63) / 64 < NOALLOC_PARTS ? (n + 63) / 64 : NOALLOC_PARTS);
        const int rows_per_part = (int)((n + parts - 1) / parts);
        minmax_slabs_kernel<<<parts, 256, 0, s>>>(dev_X, n, d, rows_per_part, scratch);
        minmax_fold_kernel<<<(d + 127) / 128, 128, 0, s>>>(scratch, parts, d, dev_min, dev_max);
        e = cudaGetLastError();
        if (e != cudaSuccess) rc = fail(SVMB200_ECUDA, "kernel launch", e);
    }
    cudaSetDevice(prev);
    return rc;
}
