// capi.cu -- the C ABI of include/svmb200.h over the CUDA kernels (sm_100a only).
//
// Host-side control only: buffer management, launches, the D2H copies of results.  There is no
// CPU implementation of any step here; if CUDA is unavailable every entry point fails loudly.
#include "../../include/svmb200.h"

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <string>
#include <unordered_map>
#include <vector>

#include "aux_kernels.cuh"
#include "smo_kernel.cuh"
#include "tc_decision.cuh"
#include "u8_rows.cuh"

using namespace svm;

static thread_local std::string g_create_error;
#ifdef SVM_TIMING
static long long *g_timing_dev = nullptr;
extern "C" int svmb200_debug_timing(long long *out, int rows)
{
    if (!g_timing_dev) return -1;
    return cudaMemcpy(out, g_timing_dev, (size_t)rows * 16 * sizeof(long long), cudaMemcpyDeviceToHost) == cudaSuccess ? 0 : -2;
}
// per-iteration cycle sums of warp 0's pass phases (uint8 rows): wait for data, main sums, tail, refill, integer reduce, finish
extern "C" int svmb200_debug_timing_pass(long long *out, int rows)
{
    if (!g_timing_dev) return -1;
    return cudaMemcpy(out, g_timing_dev + 4096 * 16 + 8 * 160 * 4, (size_t)rows * 8 * sizeof(long long), cudaMemcpyDeviceToHost) == cudaSuccess ? 0 : -2;
}
extern "C" int svmb200_debug_timing_grid(long long *out)
{
    if (!g_timing_dev) return -1;
    return cudaMemcpy(out, g_timing_dev + 4096 * 16, (size_t)8 * 160 * 4 * sizeof(long long), cudaMemcpyDeviceToHost) == cudaSuccess ? 0 : -2;
}
#endif

struct svmb200_cascade_state;
struct svmb200_ctx;
static void cascade_destroy(svmb200_ctx *ctx);

struct svmb200_ctx {
    int device = 0;
    int num_sms = 0;
    int max_smem_optin = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev2 = nullptr;
    std::string err;
    // device buffers of the training set are recycled inside the context instead of going back to the driver: a context that is
    // given one training set after the other (the cascade's stages: ~30 buffers per stage) would otherwise spend its time in
    // cudaFree / cudaMalloc, which serialise in the driver when several processes share a box.  All work of a context is on one
    // stream, so a recycled buffer is handed out in stream order.
    std::unordered_map<void *, size_t> pool_live;     // buffers handed out by dev_alloc: their sizes
    std::multimap<size_t, void *> pool_idle;          // buffers given back by dev_free, by size
    size_t pool_idle_bytes = 0;

    // training set
    long long n = 0;
    int d = 0, pitch4 = 0;
    void *X = nullptr;          // fp32 rows, FP64 rows when x_f64, packed uint8 rows when x_u8
    int storage_requested = SVMB200_STORAGE_AUTO;  // for the next svmb200_set_train_*
    int storage_resolved = SVMB200_STORAGE_F32;    // what AUTO became for the upload in progress
    bool x_f64 = false, x_u8 = false;
    // the rows the decision function works on (support-vector block, query rows): fp32 with fp32 storage, FP64 otherwise
    bool dec_f64 = false;
    int dec_pitch4 = 0;
    // uint8 storage: layout of a packed row and its per-row norms (smo_kernel.cuh "uint8 rows")
    int u8_mw = 0, u8_tw = 0, u8_tail_pitch = 0, u8_R = 0, u8_M = 0, u8_T = 0;
    double u8_wR = 0.0;
    int *u8_nmain = nullptr, *u8_pos = nullptr, *u8_src = nullptr;
    double *u8_ntail = nullptr, *u8_wt = nullptr;
    void *Xdec = nullptr;       // uint8 storage: scaled FP64 copy of all rows, built when a warm start needs it
    bool xdec_valid = false;
    int *y = nullptr;
    double *sqn = nullptr;      // |x_j|^2 formed with accumulation chunk sqn_chunk (0 = not computed)
    int sqn_chunk = 0;
    double *minv = nullptr, *maxv = nullptr;
    bool have_stats = false, have_train = false;

    // solver state
    double *f = nullptr, *alpha = nullptr;
    long long *cacheq = nullptr;
    double *cacheK = nullptr;
    size_t cacheK_elems = 0;
    // a cache kept across runs on the same X (kernel values do not depend on labels or on C): valid for (cap, chunk, gamma)
    bool cache_valid = false;
    int cache_valid_cap = 0, cache_valid_chunk = 0;
    double cache_valid_gamma = 0.0;
    long long cache_counter = 0;
    Rec *rec = nullptr;                  // [2][SMO_REC_COPIES][SMO_MAX_RECS] exchange records + the abort word
    SmoResult *result = nullptr;
    int *trace = nullptr;
    int trace_alloc = 0, trace_count = 0;

    // model
    bool trained = false;
    svmb200_params params{};
    double b = 0.0;
    std::vector<double> alpha_host;
    std::vector<int> sv_ids;
    void *svX = nullptr;
    double *svcoef = nullptr;
    double *svn = nullptr;
    bool model_ready = false;   // svX / svcoef / svn hold the support vectors of the current model
    int model_nsv = 0;
    std::vector<int> model_y;   // labels of the support vectors (kept for svmb200_save_model)
    int sv_alloc = 0;
    size_t svX_bytes = 0;          // capacity of svX (the row pitch may change between training sets)
    std::vector<double> model_alpha;   // alpha of the support vectors (kept for svmb200_save_model)
    bool have_params = false;      // params holds the values of the last svmb200_smo_train / svmb200_load_model
    int cluster_size = 0;          // cluster size of the last SMO launch (2, or 1 when clusters could not be used)

    // tensor-core decision function (svmb200_set_tensor_cores)
    bool tc_predict = false;
    double tc_threshold = 0.05;
    long long tc_rechecked = 0;

    // row sharding over several GPUs (one process per GPU; see svmb200_shard_*)
    int shard_rank = 0, shard_world = 1, shard_epoch = 0;
    Rec *mailbox = nullptr;              // [2][world], written by the peers over NVLink
    Rec *peer_mailbox[8] = {};           // peer-mapped mailboxes ([rank] = mailbox)

    // generic scratch
    void *scratch = nullptr;
    size_t scratch_bytes = 0;
    // grow-only work areas of the decision function (query rows, scores, packed panels; [1] = the rows sent to the exact kernel):
    // svmb200_predict_* allocates only when a call needs more than any earlier one did
    void *work[2] = {nullptr, nullptr};
    size_t work_bytes[2] = {0, 0};
    unsigned char *svP = nullptr;        // the model's support vectors packed for the tensor-core kernel
    size_t svP_bytes = 0;
    bool svP_valid = false;
    int launches = 0;
    svmb200_cascade_state *cascade = nullptr;     // Cascade SVM over several GPUs (cascade.cuh)
};

#define CK(call)                                                                                        \
    do {                                                                                                \
        cudaError_t e_ = (call);                                                                        \
        if (e_ != cudaSuccess) {                                                                        \
            char buf_[512];                                                                             \
            snprintf(buf_, sizeof buf_, "%s failed at %s:%d: %s", #call, __FILE__, __LINE__, cudaGetErrorString(e_)); \
            ctx->err = buf_;                                                                            \
            return SVMB200_ECUDA;                                                                       \
        }                                                                                               \
    } while (0)

#define FAIL(code, ...)                                  \
    do {                                                 \
        char buf_[512];                                  \
        snprintf(buf_, sizeof buf_, __VA_ARGS__);        \
        ctx->err = buf_;                                 \
        return (code);                                   \
    } while (0)

static int ensure_scratch(svmb200_ctx *ctx, size_t bytes)
{
    if (bytes <= ctx->scratch_bytes) return 0;
    if (ctx->scratch) CK(cudaFree(ctx->scratch));
    ctx->scratch = nullptr;
    ctx->scratch_bytes = 0;
    CK(cudaMalloc(&ctx->scratch, bytes));
    ctx->scratch_bytes = bytes;
    return 0;
}

static int ensure_work(svmb200_ctx *ctx, int which, size_t bytes)
{
    if (bytes <= ctx->work_bytes[which]) return 0;
    if (ctx->work[which]) CK(cudaFree(ctx->work[which]));
    ctx->work[which] = nullptr;
    ctx->work_bytes[which] = 0;
    bytes += bytes / 8;                                        // headroom: slightly larger batches do not reallocate
    CK(cudaMalloc(&ctx->work[which], bytes));
    ctx->work_bytes[which] = bytes;
    return 0;
}

static const size_t POOL_IDLE_MAX = (size_t)2 << 30;       // idle buffers kept: at most 2 GiB

static void pool_trim(svmb200_ctx *ctx, size_t keep_bytes)
{
    while (ctx->pool_idle_bytes > keep_bytes && !ctx->pool_idle.empty()) {
        auto it = std::prev(ctx->pool_idle.end());              // the largest first
        cudaFree(it->second);
        ctx->pool_idle_bytes -= it->first;
        ctx->pool_idle.erase(it);
    }
}
// a buffer of at least `bytes`: an idle one that is not more than twice as large, else a new one
template <typename T>
static int dev_alloc(svmb200_ctx *ctx, T **p, size_t bytes)
{
    bytes = (std::max<size_t>(bytes, 1) + 255) & ~(size_t)255;
    auto it = ctx->pool_idle.lower_bound(bytes);
    if (it != ctx->pool_idle.end() && it->first <= 2 * bytes + ((size_t)1 << 16)) {
        *p = static_cast<T *>(it->second);
        ctx->pool_live[it->second] = it->first;
        ctx->pool_idle_bytes -= it->first;
        ctx->pool_idle.erase(it);
        return 0;
    }
    void *q = nullptr;
    cudaError_t e = cudaMalloc(&q, bytes);
    if (e != cudaSuccess) {                                     // out of memory: give the idle buffers back and try once more
        (void)cudaGetLastError();
        pool_trim(ctx, 0);
        e = cudaMalloc(&q, bytes);
    }
    CK(e);
    ctx->pool_live[q] = bytes;
    *p = static_cast<T *>(q);
    return 0;
}
template <typename T>
static int dev_free(svmb200_ctx *ctx, T *&p)
{
    if (!p) return 0;
    auto it = ctx->pool_live.find((void *)p);
    if (it == ctx->pool_live.end()) CK(cudaFree(p));           // not from dev_alloc (the row cache, work areas)
    else {
        ctx->pool_idle.emplace(it->second, (void *)p);
        ctx->pool_idle_bytes += it->second;
        ctx->pool_live.erase(it);
        pool_trim(ctx, POOL_IDLE_MAX);
    }
    p = nullptr;
    return 0;
}

extern "C" {

const char *svmb200_version(void) { return "svmb200 0.1 (sm_100a)"; }

void svmb200_default_params(svmb200_params *p)
{
    p->C = 10.0;            // main3.cpp:163,342
    p->gamma = 0.00125;     // main3.cpp:95
    p->eps = 1e-12;         // main3.cpp:165
    p->tau = 1e-5;          // main3.cpp:196
    p->max_iter = 100000;   // main3.cpp:198
    p->sv_tol = 1e-8;       // main3.cpp:367
    p->wss = 1;
    p->warm_start = 0;
    p->cache_rows = 0;
    p->accum_chunk = 0;
    p->trace_cap = 0;
    p->l2_persist = -1;
    p->keep_cache = 0;
}

const char *svmb200_last_error(const svmb200_ctx *ctx) { return ctx ? ctx->err.c_str() : g_create_error.c_str(); }

int svmb200_create(svmb200_ctx **out, int device)
{
    if (!out) return SVMB200_EINVAL;
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) {
        g_create_error = std::string("no CUDA device: ") + cudaGetErrorString(e);
        return SVMB200_ENODEVICE;
    }
    if (device < 0 || device >= count) {
        g_create_error = "device ordinal out of range";
        return SVMB200_EINVAL;
    }
    svmb200_ctx *ctx = new svmb200_ctx();
    ctx->device = device;
    auto fail = [&](const char *what, cudaError_t err) {
        g_create_error = std::string(what) + ": " + cudaGetErrorString(err);
        delete ctx;
        return SVMB200_ECUDA;
    };
    if ((e = cudaSetDevice(device)) != cudaSuccess) return fail("cudaSetDevice", e);
    cudaDeviceProp prop;
    if ((e = cudaGetDeviceProperties(&prop, device)) != cudaSuccess) return fail("cudaGetDeviceProperties", e);
    if (prop.major != 10) {
        g_create_error = "svmb200 is built for sm_100a only; device is sm_" + std::to_string(prop.major) + std::to_string(prop.minor);
        delete ctx;
        return SVMB200_ENODEVICE;
    }
    ctx->num_sms = prop.multiProcessorCount;
    ctx->max_smem_optin = (int)prop.sharedMemPerBlockOptin;
    if ((e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking)) != cudaSuccess) return fail("cudaStreamCreate", e);
    if ((e = cudaEventCreate(&ctx->ev0)) != cudaSuccess) return fail("cudaEventCreate", e);
    if ((e = cudaEventCreate(&ctx->ev1)) != cudaSuccess) return fail("cudaEventCreate", e);
    if ((e = cudaEventCreate(&ctx->ev2)) != cudaSuccess) return fail("cudaEventCreate", e);
    if ((e = cudaMalloc(&ctx->result, sizeof(SmoResult))) != cudaSuccess) return fail("cudaMalloc", e);
    *out = ctx;
    return SVMB200_OK;
}

void svmb200_destroy(svmb200_ctx *ctx)
{
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    cascade_destroy(ctx);
    cudaFree(ctx->X); cudaFree(ctx->y); cudaFree(ctx->sqn); cudaFree(ctx->svn); cudaFree(ctx->minv); cudaFree(ctx->maxv);
    cudaFree(ctx->f); cudaFree(ctx->alpha); cudaFree(ctx->cacheq); cudaFree(ctx->cacheK);
    cudaFree(ctx->rec); cudaFree(ctx->result); cudaFree(ctx->trace);
    for (int p = 0; p < ctx->shard_world; ++p)
        if (p != ctx->shard_rank && ctx->peer_mailbox[p]) cudaIpcCloseMemHandle(ctx->peer_mailbox[p]);
    cudaFree(ctx->mailbox);
    cudaFree(ctx->svX); cudaFree(ctx->svcoef); cudaFree(ctx->scratch);
    cudaFree(ctx->work[0]); cudaFree(ctx->work[1]); cudaFree(ctx->svP);
    cudaFree(ctx->u8_nmain); cudaFree(ctx->u8_pos); cudaFree(ctx->u8_src); cudaFree(ctx->u8_ntail); cudaFree(ctx->u8_wt); cudaFree(ctx->Xdec);
    pool_trim(ctx, 0);
    cudaEventDestroy(ctx->ev0); cudaEventDestroy(ctx->ev1); cudaEventDestroy(ctx->ev2);
    cudaStreamDestroy(ctx->stream);
    delete ctx;
}

} // extern "C"

static int max_pitch(bool f64) { return f64 ? 13 * 32 : 8 * 32; }   // 16-byte groups per row the pivot registers can hold

// ------------------------------------------------------------------------------------------
static int alloc_train(svmb200_ctx *ctx, long long n, int d)
{
    if (n <= 0 || d <= 0) FAIL(SVMB200_EINVAL, "n and d must be positive (n=%lld d=%d)", n, d);
    if (n >= 0x7fffffffLL) FAIL(SVMB200_ELIMIT, "n=%lld exceeds the 2^31-2 row limit", n);
    const bool f64 = ctx->storage_resolved == SVMB200_STORAGE_F64, u8 = ctx->storage_resolved == SVMB200_STORAGE_U8;
    const bool dec_f64 = f64 || u8;
    const int dec_pitch4 = dec_f64 ? (d + 1) / 2 : (d + 3) / 4;  // 16-byte groups per row of the decision function's rows
    if (dec_pitch4 > max_pitch(dec_f64))
        FAIL(SVMB200_ELIMIT, "d=%d exceeds the %d-feature limit of the in-register pivot rows", d, dec_f64 ? 832 : 1024);
    CK(cudaSetDevice(ctx->device));
    if (n != ctx->n || d != ctx->d || f64 != ctx->x_f64 || u8 != ctx->x_u8 || u8) {
        ctx->n = 0;                 // nothing is valid until every buffer below exists
        ctx->d = 0;
        ctx->have_train = false;
        if (dev_free(ctx, ctx->X) || dev_free(ctx, ctx->y) || dev_free(ctx, ctx->sqn) || dev_free(ctx, ctx->f) || dev_free(ctx, ctx->alpha) ||
            dev_free(ctx, ctx->cacheq) || dev_free(ctx, ctx->minv) || dev_free(ctx, ctx->maxv) || dev_free(ctx, ctx->svX) ||
            dev_free(ctx, ctx->svcoef) || dev_free(ctx, ctx->svn) || dev_free(ctx, ctx->u8_nmain) || dev_free(ctx, ctx->u8_ntail) ||
            dev_free(ctx, ctx->Xdec))
            return SVMB200_ECUDA;
        ctx->sv_alloc = 0;          // the support-vector block was sized for the old row pitch
        ctx->svX_bytes = 0;
        const int pitch4 = u8 ? 0 : dec_pitch4;                  // uint8 rows: known (and X allocated) once the statistics are
        if (!u8) if (int rc = dev_alloc(ctx, &ctx->X, (size_t)n * pitch4 * 16)) return rc;
        if (int rc = dev_alloc(ctx, &ctx->y, (size_t)n * sizeof(int))) return rc;
        if (int rc = dev_alloc(ctx, &ctx->sqn, (size_t)n * sizeof(double))) return rc;
        if (int rc = dev_alloc(ctx, &ctx->f, (size_t)n * sizeof(double))) return rc;
        if (int rc = dev_alloc(ctx, &ctx->alpha, (size_t)n * sizeof(double))) return rc;
        if (int rc = dev_alloc(ctx, &ctx->cacheq, (size_t)n * sizeof(long long))) return rc;
        if (int rc = dev_alloc(ctx, &ctx->minv, (size_t)d * sizeof(double))) return rc;
        if (int rc = dev_alloc(ctx, &ctx->maxv, (size_t)d * sizeof(double))) return rc;
        if (u8) {
            if (int rc = dev_alloc(ctx, &ctx->u8_nmain, (size_t)n * sizeof(int))) return rc;
            if (int rc = dev_alloc(ctx, &ctx->u8_ntail, (size_t)n * sizeof(double))) return rc;
        }
        ctx->n = n;
        ctx->d = d;
        ctx->pitch4 = pitch4;
        ctx->x_f64 = f64;
        ctx->x_u8 = u8;
        ctx->dec_f64 = dec_f64;
        ctx->dec_pitch4 = dec_pitch4;
    }
    ctx->have_train = false;
    ctx->trained = false;
    ctx->have_stats = false;
    ctx->sqn_chunk = 0;
    ctx->model_ready = false;
    ctx->cache_valid = false;
    ctx->xdec_valid = false;
    return 0;
}

// stage raw host rows in device scratch, compute statistics (optional), scale into dst
template <typename T>
static int ingest(svmb200_ctx *ctx, const T *X_host, long long rows, bool compute_stats, bool apply_stats, void *dst, bool dst_f64, int pitch4)
{
    const int d = ctx->d;
    const size_t raw_bytes = (size_t)rows * d * sizeof(T);
    const int parts = (int)std::min<long long>(ctx->num_sms, (rows + 63) / 64);
    const size_t part_bytes = compute_stats ? (size_t)2 * parts * d * sizeof(double) : 0;
    const size_t raw_off = (part_bytes + 255) & ~(size_t)255;
    if (int rc = ensure_scratch(ctx, raw_off + raw_bytes)) return rc;
    T *raw = reinterpret_cast<T *>(static_cast<char *>(ctx->scratch) + raw_off);
    CK(cudaMemcpyAsync(raw, X_host, raw_bytes, cudaMemcpyHostToDevice, ctx->stream));
    if (compute_stats) {
        double *pmin = static_cast<double *>(ctx->scratch), *pmax = pmin + (size_t)parts * d;
        const int rows_per_cta = (int)((rows + parts - 1) / parts);
        colminmax_partial_kernel<T><<<parts, 256, 0, ctx->stream>>>(raw, rows, d, rows_per_cta, pmin, pmax);
        colminmax_final_kernel<<<(d + 127) / 128, 128, 0, ctx->stream>>>(pmin, pmax, parts, d, ctx->minv, ctx->maxv);
        ctx->launches += 2;
        CK(cudaGetLastError());
    }
    const long long total = rows * pitch4;
    const int blocks = (int)std::min<long long>((total + 255) / 256, (long long)ctx->num_sms * 16);
    if (!dst) return 0;                         // statistics only (uint8 storage packs the raw rows itself)
    if (dst_f64)
        scale_rows_kernel<T, double><<<blocks, 256, 0, ctx->stream>>>(raw, rows, d, pitch4, apply_stats ? ctx->minv : nullptr,
                                                                      apply_stats ? ctx->maxv : nullptr, static_cast<double *>(dst));
    else
        scale_rows_kernel<T, float><<<blocks, 256, 0, ctx->stream>>>(raw, rows, d, pitch4, apply_stats ? ctx->minv : nullptr,
                                                                     apply_stats ? ctx->maxv : nullptr, static_cast<float *>(dst));
    ctx->launches += 1;
    CK(cudaGetLastError());
    return 0;
}

// uint8 storage: the statistics are on the device (ctx->minv / maxv) and the raw rows in scratch at `raw`.  Lays the row
// out as [main class | tail] (smo_kernel.cuh "uint8 rows"), packs every row that way and forms the per-row norms.
static int pack_u8_rows(svmb200_ctx *ctx, const unsigned char *raw, long long n)
{
    const int d = ctx->d;
    std::vector<double> mn(d), mx(d);
    CK(cudaMemcpyAsync(mn.data(), ctx->minv, (size_t)d * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaMemcpyAsync(mx.data(), ctx->maxv, (size_t)d * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    // the main class: the most frequent non-zero range (the larger value wins a tie), if at least 32 features share it
    std::vector<int> range(d);
    int count[256] = {0};
    for (int k = 0; k < d; ++k) { range[k] = (int)(mx[k] - mn[k]); count[range[k]]++; }
    int R = 0;
    for (int r = 1; r < 256; ++r)
        if (count[r] > 0 && (R == 0 || count[r] >= count[R])) R = r;
    if (R && count[R] < 32) R = 0;
    std::vector<int> main_idx, tail_idx;
    for (int k = 0; k < d; ++k) {
        if (range[k] == 0) continue;              // a constant feature adds (u - p) = 0 to every distance
        (range[k] == R ? main_idx : tail_idx).push_back(k);
    }
    const int M = (int)main_idx.size(), T = (int)tail_idx.size();
    // both sections are padded with zero bytes (zero weights) to whole 128-byte slots: 32 lanes x one 4-byte word
    const int mw = ((M + 3) / 4 + 31) & ~31, tw = ((T + 3) / 4 + 31) & ~31;
    // two arrays (u8_rows.cuh: U8View): [n][4 * mw] bytes of the main class, then [n][tail_pitch] bytes of the tail
    const int tail_pitch = tw ? 4 * tw + 16 : 0;
    const int row_bytes = 4 * mw + tail_pitch;
    std::vector<int> src(std::max(row_bytes, 4), -1), pos(d, -1);
    std::vector<double> wt((size_t)std::max(tw, 1) * 4, 0.0);
    for (int k = 0; k < M; ++k) { src[k] = main_idx[k]; pos[main_idx[k]] = k; }
    for (int k = 0; k < T; ++k) {
        src[4 * mw + k] = tail_idx[k];
        pos[tail_idx[k]] = 4 * mw + k;
        const double sk = 1.0 / (double)range[tail_idx[k]];
        wt[k] = sk * sk;
    }
    const double sR = R ? 1.0 / (double)R : 0.0;
    ctx->u8_R = R; ctx->u8_M = M; ctx->u8_T = T; ctx->u8_mw = mw; ctx->u8_tw = tw; ctx->u8_tail_pitch = tail_pitch; ctx->u8_wR = sR * sR;
    ctx->pitch4 = row_bytes / 16;                   // bytes of a row over both arrays (accounting only)
    if (dev_free(ctx, ctx->u8_src) || dev_free(ctx, ctx->u8_pos) || dev_free(ctx, ctx->u8_wt) || dev_free(ctx, ctx->X)) return SVMB200_ECUDA;
    if (int rc = dev_alloc(ctx, &ctx->u8_src, src.size() * sizeof(int))) return rc;
    if (int rc = dev_alloc(ctx, &ctx->u8_pos, pos.size() * sizeof(int))) return rc;
    if (int rc = dev_alloc(ctx, &ctx->u8_wt, wt.size() * sizeof(double))) return rc;
    if (int rc = dev_alloc(ctx, &ctx->X, std::max<size_t>((size_t)n * row_bytes, 256))) return rc;
    CK(cudaMemcpyAsync(ctx->u8_src, src.data(), src.size() * sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemcpyAsync(ctx->u8_pos, pos.data(), pos.size() * sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemcpyAsync(ctx->u8_wt, wt.data(), wt.size() * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
    const long long words = n * (row_bytes / 4);
    if (words > 0) {
        u8_pack_kernel<<<(int)std::min<long long>((words + 255) / 256, (long long)ctx->num_sms * 16), 256, 0, ctx->stream>>>(
            raw, n, d, mw, tail_pitch, ctx->u8_src, static_cast<unsigned char *>(ctx->X));
        CK(cudaGetLastError());
    }
    u8_norms_kernel<<<(unsigned)((n + 127) / 128), 128, 0, ctx->stream>>>(U8View(ctx->X, n, mw, tail_pitch), n, mw, tw, ctx->u8_wt, ctx->u8_nmain,
                                                                             ctx->u8_ntail);
    CK(cudaGetLastError());
    ctx->launches += 2;
    CK(cudaStreamSynchronize(ctx->stream));      // src / pos / wt are host temporaries
    return 0;
}

template <typename T>
static int set_train(svmb200_ctx *ctx, const T *X, const int32_t *y, int64_t n, int32_t d, bool scale, double *min_out,
                     double *max_out)
{
    if (!ctx) return SVMB200_EINVAL;
    if (!X || !y) FAIL(SVMB200_EINVAL, "X and y must not be NULL");
    // AUTO: bytes stay bytes (svmb200_set_train_u8 with up to 832 features), everything else becomes fp32 rows
    const bool bytes_in = std::is_same<T, uint8_t>::value && scale;
    ctx->storage_resolved = ctx->storage_requested != SVMB200_STORAGE_AUTO ? ctx->storage_requested
                            : (bytes_in && d <= 832) ? SVMB200_STORAGE_U8 : SVMB200_STORAGE_F32;
    const bool u8 = ctx->storage_resolved == SVMB200_STORAGE_U8;
    if (u8 && !bytes_in)
        FAIL(SVMB200_ESTATE, "uint8 storage keeps the bytes of svmb200_set_train_u8: choose fp32 or FP64 storage for FP64 input");
    if (int rc = alloc_train(ctx, n, d)) return rc;
    if (int rc = ingest<T>(ctx, X, n, scale, scale, u8 ? nullptr : ctx->X, ctx->x_f64, ctx->pitch4)) return rc;
    if (u8) {
        const int parts = (int)std::min<long long>(ctx->num_sms, (n + 63) / 64);           // where ingest() staged the raw rows
        const size_t raw_off = ((size_t)2 * parts * d * sizeof(double) + 255) & ~(size_t)255;
        if (int rc = pack_u8_rows(ctx, static_cast<const unsigned char *>(ctx->scratch) + raw_off, n)) return rc;
    }
    CK(cudaMemcpyAsync(ctx->y, y, (size_t)n * sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
    if (scale) {
        if (min_out) CK(cudaMemcpyAsync(min_out, ctx->minv, (size_t)d * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
        if (max_out) CK(cudaMemcpyAsync(max_out, ctx->maxv, (size_t)d * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    }
    CK(cudaStreamSynchronize(ctx->stream));
    ctx->have_stats = scale;
    ctx->have_train = true;
    return SVMB200_OK;
}

extern "C" {

int svmb200_set_train_u8(svmb200_ctx *ctx, const uint8_t *X, const int32_t *y, int64_t n, int32_t d, double *min_out, double *max_out)
{
    return set_train<uint8_t>(ctx, X, y, n, d, true, min_out, max_out);
}
int svmb200_set_train_f64(svmb200_ctx *ctx, const double *X, const int32_t *y, int64_t n, int32_t d, double *min_out, double *max_out)
{
    return set_train<double>(ctx, X, y, n, d, true, min_out, max_out);
}
int svmb200_set_train_scaled_f64(svmb200_ctx *ctx, const double *Xs, const int32_t *y, int64_t n, int32_t d)
{
    return set_train<double>(ctx, Xs, y, n, d, false, nullptr, nullptr);
}

} // extern "C"

// new labels for the rows already uploaded (one-vs-rest over several digits on one X): the kernel-row cache stays valid
extern "C" int svmb200_set_labels(svmb200_ctx *ctx, const int32_t *y)
{
    if (!ctx || !y) return SVMB200_EINVAL;
    if (!ctx->have_train) FAIL(SVMB200_ESTATE, "svmb200_set_labels called before svmb200_set_train_*");
    CK(cudaSetDevice(ctx->device));
    CK(cudaMemcpyAsync(ctx->y, y, (size_t)ctx->n * sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    ctx->trained = false;
    ctx->model_ready = false;
    return SVMB200_OK;
}

// storage precision of X for the NEXT svmb200_set_train_*: 0 = fp32 (default), 1 = FP64 ("exact")
extern "C" int svmb200_set_exact(svmb200_ctx *ctx, int32_t enable)
{
    if (!ctx) return SVMB200_EINVAL;
    ctx->storage_requested = enable ? SVMB200_STORAGE_F64 : SVMB200_STORAGE_AUTO;
    return SVMB200_OK;
}

// storage of X for the NEXT svmb200_set_train_*: fp32 rows, FP64 rows, or (svmb200_set_train_u8 only) the bytes themselves
extern "C" int svmb200_set_storage(svmb200_ctx *ctx, int32_t storage)
{
    if (!ctx) return SVMB200_EINVAL;
    if (storage != SVMB200_STORAGE_AUTO && storage != SVMB200_STORAGE_F32 && storage != SVMB200_STORAGE_F64 && storage != SVMB200_STORAGE_U8)
        FAIL(SVMB200_EINVAL, "storage must be SVMB200_STORAGE_AUTO, _F32, _F64 or _U8");
    ctx->storage_requested = storage;
    return SVMB200_OK;
}

extern "C" int svmb200_get_u8_layout(svmb200_ctx *ctx, int32_t *main_range, int32_t *main_features, int32_t *tail_features, int32_t *row_bytes)
{
    if (!ctx) return SVMB200_EINVAL;
    if (!ctx->have_train || !ctx->x_u8) FAIL(SVMB200_ESTATE, "no training set in uint8 storage");
    if (main_range) *main_range = ctx->u8_R;
    if (main_features) *main_features = ctx->u8_M;
    if (tail_features) *tail_features = ctx->u8_T;
    if (row_bytes) *row_bytes = ctx->pitch4 * 16;
    return SVMB200_OK;
}

extern "C" int svmb200_minmax_f64(svmb200_ctx *ctx, const double *X, int64_t n, int32_t d, double *min_out, double *max_out)
{
    if (!ctx) return SVMB200_EINVAL;
    if (!X || !min_out || !max_out || n <= 0 || d <= 0) FAIL(SVMB200_EINVAL, "bad arguments");
    CK(cudaSetDevice(ctx->device));
    const int parts = (int)std::min<long long>(ctx->num_sms * 4, (n + 63) / 64);
    const size_t stat_bytes = ((size_t)(2 * parts + 2) * d * sizeof(double) + 255) & ~(size_t)255;
    if (int rc = ensure_scratch(ctx, stat_bytes + (size_t)n * d * sizeof(double))) return rc;
    double *pmin = static_cast<double *>(ctx->scratch), *pmax = pmin + (size_t)parts * d, *mn = pmax + (size_t)parts * d, *mx = mn + d;
    double *raw = reinterpret_cast<double *>(static_cast<char *>(ctx->scratch) + stat_bytes);
    CK(cudaMemcpyAsync(raw, X, (size_t)n * d * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
    colminmax_partial_kernel<double><<<parts, 256, 0, ctx->stream>>>(raw, n, d, (int)((n + parts - 1) / parts), pmin, pmax);
    colminmax_final_kernel<<<(d + 127) / 128, 128, 0, ctx->stream>>>(pmin, pmax, parts, d, mn, mx);
    ctx->launches += 2;
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(min_out, mn, (size_t)d * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaMemcpyAsync(max_out, mx, (size_t)d * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return SVMB200_OK;
}

extern "C" int svmb200_scale_f64(svmb200_ctx *ctx, double *X, const double *min_val, const double *max_val, int64_t n, int32_t d)
{
    if (!ctx) return SVMB200_EINVAL;
    if (!X || !min_val || !max_val || n <= 0 || d <= 0) FAIL(SVMB200_EINVAL, "bad arguments");
    CK(cudaSetDevice(ctx->device));
    const size_t stat_bytes = ((size_t)2 * d * sizeof(double) + 255) & ~(size_t)255;
    if (int rc = ensure_scratch(ctx, stat_bytes + (size_t)n * d * sizeof(double))) return rc;
    double *mn = static_cast<double *>(ctx->scratch), *mx = mn + d;
    double *raw = reinterpret_cast<double *>(static_cast<char *>(ctx->scratch) + stat_bytes);
    CK(cudaMemcpyAsync(mn, min_val, (size_t)d * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemcpyAsync(mx, max_val, (size_t)d * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemcpyAsync(raw, X, (size_t)n * d * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
    const long long total = (long long)n * d;
    scale_f64_kernel<<<(int)std::min<long long>((total + 255) / 256, (long long)ctx->num_sms * 16), 256, 0, ctx->stream>>>(raw, n, d, mn, mx);
    ctx->launches += 1;
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(X, raw, (size_t)n * d * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return SVMB200_OK;
}

// ------------------------------------------------------------------------------------------
// Kernel dispatch.  Q = 16-byte groups per lane per row (the pivot rows live in registers, so Q is a template
// parameter); fp32 rows: Q in {1,2,4,7,8} (d <= 1024), FP64 rows ("exact" storage): Q in {1,2,4,7,13} (d <= 832).
static int q_for_pitch(int pitch4, bool f64)
{
    const int need = (pitch4 + 31) / 32;
    if (need <= 1) return 1;
    if (need <= 2) return 2;
    if (need <= 4) return 4;
    if (need <= 7) return 7;
    return f64 ? 13 : 8;
}

struct KernelKey { bool f64; int Q, CH, threads; };

// calls fn.template operator()<XT, Q, CH, THREADS>() for the instantiation the key names
template <typename F>
static cudaError_t dispatch(const KernelKey &k, F &&fn)
{
    if (k.f64) {                      // FP64 rows: 384 threads (pivot rows take 104 registers at Q = 13), no accumulation chunk
        switch (k.Q) {
        case 1: return fn.template operator()<double, 1, 1, 384>();
        case 2: return fn.template operator()<double, 2, 1, 384>();
        case 4: return fn.template operator()<double, 4, 1, 384>();
        case 7: return fn.template operator()<double, 7, 1, 384>();
        default:
            if (k.threads == 256) return fn.template operator()<double, 13, 1, 256>();   // the SMO kernel's two-rows-per-step form
            return fn.template operator()<double, 13, 1, 384>();
        }
    }
#define SVM_F32_CASE(QQ, TT)                                                    \
    switch (k.CH) {                                                              \
    case 1: return fn.template operator()<float, QQ, 1, TT>();                   \
    case 2: return fn.template operator()<float, QQ, 2, TT>();                   \
    default: return fn.template operator()<float, QQ, 4, TT>();                  \
    }
    // 256 threads: the SMO kernel's rows-in-groups form; 512 (384 for Q = 7): a row per step, and every other kernel
    switch (k.Q) {
    case 1:
        if (k.threads == 256) { SVM_F32_CASE(1, 256) }
        SVM_F32_CASE(1, 512)
    case 2:
        if (k.threads == 256) { SVM_F32_CASE(2, 256) }
        SVM_F32_CASE(2, 512)
    case 4:
        if (k.threads == 256) { SVM_F32_CASE(4, 256) }
        SVM_F32_CASE(4, 512)
    case 7:
        if (k.threads == 256) { SVM_F32_CASE(7, 256) }
        if (k.threads == 384) { SVM_F32_CASE(7, 384) }
        SVM_F32_CASE(7, 512)
    default:
        if (k.threads == 256) { SVM_F32_CASE(8, 256) }
        SVM_F32_CASE(8, 512)
    }
#undef SVM_F32_CASE
}

// cooperative launch (all CTAs co-resident: they spin on each other's records) in clusters of `cluster` CTAs
template <typename K>
static cudaError_t launch_persistent(K kernel, const SmoArgs &args, int grid, int threads, int cluster, size_t smem, cudaStream_t stream)
{
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(threads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attrs[2];
    attrs[0].id = cudaLaunchAttributeCooperative;
    attrs[0].val.cooperative = 1;
    attrs[1].id = cudaLaunchAttributeClusterDimension;
    attrs[1].val.clusterDim.x = cluster;
    attrs[1].val.clusterDim.y = 1;
    attrs[1].val.clusterDim.z = 1;
    cfg.attrs = attrs;
    cfg.numAttrs = 2;
    if (cluster > 1) {
        int max_clusters = 0;
        e = cudaOccupancyMaxActiveClusters(&max_clusters, kernel, &cfg);
#ifdef SVM_TIMING
        fprintf(stderr, "svmb200: cudaOccupancyMaxActiveClusters -> %d clusters of %d (%s) for a grid of %d\n", max_clusters, cluster, cudaGetErrorString(e), grid);
#endif
        if (e != cudaSuccess) return e;
        if (max_clusters * cluster < grid) return cudaErrorCooperativeLaunchTooLarge;
    }
    return cudaLaunchKernelEx(&cfg, kernel, args);
}

// The SMO kernel's instantiations: fp32 rows in groups (256 threads) and FP64 rows; the one-row-per-step fp32 forms
// (384 / 512 threads) exist in the profiling build only.
static cudaError_t launch_smo_any(const KernelKey &k, const SmoArgs &args, int grid, int cluster, size_t smem, cudaStream_t stream)
{
    auto go = [&]<typename XT, int Q, int CH, int THREADS>() -> cudaError_t {
        return launch_persistent(smo_persistent_kernel<XT, Q, CH, THREADS>, args, grid, THREADS, cluster, smem, stream);
    };
#ifdef SVM_TIMING
    return dispatch(k, go);
#else
    if (k.f64) {
        switch (k.Q) {
        case 1: return go.template operator()<double, 1, 1, 384>();
        case 2: return go.template operator()<double, 2, 1, 384>();
        case 4: return go.template operator()<double, 4, 1, 384>();
        case 7: return go.template operator()<double, 7, 1, 384>();
        default:
            if (k.threads == 256) return go.template operator()<double, 13, 1, 256>();
            return go.template operator()<double, 13, 1, 384>();
        }
    }
#define SVM_F32_SMO(QQ)                                                          \
    switch (k.CH) {                                                              \
    case 1: return go.template operator()<float, QQ, 1, 256>();                  \
    case 2: return go.template operator()<float, QQ, 2, 256>();                  \
    default: return go.template operator()<float, QQ, 4, 256>();                 \
    }
    switch (k.Q) {
    case 1: SVM_F32_SMO(1)
    case 2: SVM_F32_SMO(2)
    case 4: SVM_F32_SMO(4)
    case 7: SVM_F32_SMO(7)
    default: SVM_F32_SMO(8)
    }
#undef SVM_F32_SMO
#endif
}

template <typename XT, int Q, int CH, bool FULL>
static cudaError_t launch_decision(const void *Xq, long long m, const void *SV, const double *svn, const double *coef, int nsv, int pitch4,
                                   double gamma, double init, const int *ysub, double *out, int num_sms, cudaStream_t stream)
{
    const size_t smem = decision_smem_bytes(pitch4);
    cudaError_t e = cudaFuncSetAttribute(decision_kernel<XT, Q, CH, FULL>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    // every CTA streams the whole support-vector block, so a launch costs (waves of CTAs) x (warps per CTA): pick the
    // CTA size between 8 and DEC_WARPS warps that wastes least of the last wave
    constexpr int per_warp = dec_rows_per_warp<XT, Q>();
    int warps = DEC_WARPS;
    long long best_cost = -1;
    for (int w = DEC_WARPS; w >= 8; --w) {
        const long long ctas = (m + (long long)w * per_warp - 1) / ((long long)w * per_warp);
        const long long cost = ((ctas + num_sms - 1) / num_sms) * w;
        if (best_cost < 0 || cost < best_cost) { best_cost = cost; warps = w; }
    }
    const int grid = (int)((m + (long long)warps * per_warp - 1) / ((long long)warps * per_warp));
    decision_kernel<XT, Q, CH, FULL><<<grid, warps * 32, smem, stream>>>(Xq, m, SV, svn, coef, nsv, pitch4, gamma, init, ysub, out);
    return cudaGetLastError();
}

static cudaError_t launch_decision_any(KernelKey k, const void *Xq, long long m, const void *SV, const double *svn, const double *coef,
                                       int nsv, int pitch4, double gamma, double init, const int *ysub, double *out, cudaStream_t stream)
{
    int dev = 0, num_sms = 148;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev);
    k.threads = k.f64 ? 384 : 512;      // one instantiation per (XT, Q, CH): the CTA size of this kernel is fixed
    return dispatch(k, [&]<typename XT, int Q, int CH, int THREADS>() -> cudaError_t {
        if (sizeof(XT) == 4 && (Q - 1) * 32 < pitch4)
            return launch_decision<XT, Q, CH, true>(Xq, m, SV, svn, coef, nsv, pitch4, gamma, init, ysub, out, num_sms, stream);
        return launch_decision<XT, Q, CH, false>(Xq, m, SV, svn, coef, nsv, pitch4, gamma, init, ysub, out, num_sms, stream);
    });
}

// few query rows: a handful of CTAs per row (falls back to the streaming kernel when the kernel values do not fit in
// shared memory).  rows_work: decision_rows_work_bytes(rows, nsv) of device memory, or NULL = one CTA per row.
static int decision_rows_parts(int rows, int num_sms) { return std::max(1, std::min(8, num_sms / std::max(rows, 1))); }
static size_t decision_rows_work_bytes(int rows, int nsv) { return ((size_t)rows * nsv * sizeof(double) + 255 & ~(size_t)255) + (size_t)rows * sizeof(int); }

static cudaError_t launch_decision_rows_any(KernelKey k, const void *Xq, int rows, const void *SV, const double *svn, const double *coef,
                                            int nsv, int pitch4, double gamma, double init, double *out, int max_smem, int num_sms,
                                            void *rows_work, cudaStream_t stream)
{
    const size_t smem = (size_t)nsv * sizeof(double) + 16;
    if (smem > (size_t)max_smem) return launch_decision_any(k, Xq, rows, SV, svn, coef, nsv, pitch4, gamma, init, nullptr, out, stream);
    const int parts = rows_work ? decision_rows_parts(rows, num_sms) : 1;
    double *kprod = static_cast<double *>(rows_work);
    int *done = rows_work ? reinterpret_cast<int *>(static_cast<char *>(rows_work) + ((size_t)rows * nsv * sizeof(double) + 255 & ~(size_t)255)) : nullptr;
    if (parts > 1) {
        cudaError_t e = cudaMemsetAsync(done, 0, (size_t)rows * sizeof(int), stream);
        if (e != cudaSuccess) return e;
    }
    k.threads = k.f64 ? 384 : 512;
    return dispatch(k, [&]<typename XT, int Q, int CH, int THREADS>() -> cudaError_t {
        cudaError_t e = cudaFuncSetAttribute(decision_rows_kernel<XT, Q, CH>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        decision_rows_kernel<XT, Q, CH><<<rows * parts, DROWS_THREADS, smem, stream>>>(Xq, SV, svn, coef, nsv, pitch4, gamma, init, out, parts, kprod, done);
        return cudaGetLastError();
    });
}

static cudaError_t launch_krow_any(KernelKey k, const void *X, const double *sqn, long long n, int pitch4, long long pivot, double gamma,
                                   double *out, cudaStream_t s)
{
    k.threads = k.f64 ? 384 : 512;
    return dispatch(k, [&]<typename XT, int Q, int CH, int THREADS>() -> cudaError_t {
        const long long threads = n * 32;
        kernel_row_kernel<XT, Q, CH><<<(unsigned)((threads + 255) / 256), 256, 0, s>>>(X, sqn, n, pitch4, pivot, gamma, out);
        return cudaGetLastError();
    });
}

static cudaError_t launch_sqnorm_any(KernelKey k, const float *X, long long n, int pitch4, double *sqn, cudaStream_t s)
{
    k.threads = 512;
    k.f64 = false;
    return dispatch(k, [&]<typename XT, int Q, int CH, int THREADS>() -> cudaError_t {
        const long long threads = n * 32;
        row_sqnorm_kernel<Q, CH><<<(unsigned)((threads + 255) / 256), 256, 0, s>>>(X, n, pitch4, sqn);
        return cudaGetLastError();
    });
}

// squared norms of the fp32 training rows for accumulation chunk CH (recomputed only when CH changes)
static int ensure_norms(svmb200_ctx *ctx, int CH)
{
    if (ctx->x_f64 || ctx->x_u8 || ctx->sqn_chunk == CH) return 0;
    KernelKey k{false, q_for_pitch(ctx->pitch4, false), CH, 512};
    cudaError_t e = launch_sqnorm_any(k, static_cast<const float *>(ctx->X), ctx->n, ctx->pitch4, ctx->sqn, ctx->stream);
    ctx->launches += 1;
    CK(e);
    ctx->sqn_chunk = CH;
    return 0;
}

// 8 warps x four rows per step (242 registers, no spills) + L2-resident share of X: 25.9 us/iteration at 60k; 12 warps x one row per
// step (165 registers): 33.4; 16 warps (128 registers, spills): 35.4
static const int SMO_DEFAULT_THREADS = 256;
static const int SMO_U8_THREADS = 256;      // uint8 rows: 8 warps x groups of eight rows
static const int DEFAULT_CHUNK = 4;   // fp32 storage of X dominates the error of K for every chunk (DESIGN.md "Numerics")

static int resolve_chunk(int c)
{
    if (c == 0) return DEFAULT_CHUNK;
    return (c == 1 || c == 2 || c == 4) ? c : -1;
}

// coefficients alpha_j * y_j and the dense SV block for the decision kernel
static int build_sv_block(svmb200_ctx *ctx, const std::vector<int> &ids, const double *alpha_host, const std::vector<int> &y_host_sv)
{
    const int nsv = (int)ids.size();
    ctx->svP_valid = false;
    if (nsv > ctx->sv_alloc || (size_t)nsv * ctx->dec_pitch4 * 16 > ctx->svX_bytes) {
        if (dev_free(ctx, ctx->svX) || dev_free(ctx, ctx->svcoef) || dev_free(ctx, ctx->svn)) return SVMB200_ECUDA;
        ctx->sv_alloc = 0;
        ctx->svX_bytes = 0;
        const int cap = std::max(nsv, 1024);
        if (int rc = dev_alloc(ctx, &ctx->svX, (size_t)cap * ctx->dec_pitch4 * 16)) return rc;
        ctx->svX_bytes = (size_t)cap * ctx->dec_pitch4 * 16;
        if (int rc = dev_alloc(ctx, &ctx->svcoef, (size_t)cap * sizeof(double))) return rc;
        if (int rc = dev_alloc(ctx, &ctx->svn, (size_t)cap * sizeof(double))) return rc;
        ctx->sv_alloc = cap;
    }
    if (nsv == 0) return 0;
    std::vector<double> coef(nsv);
    for (int k = 0; k < nsv; ++k) coef[k] = alpha_host[ids[k]] * y_host_sv[k];   // alpha[j] * y_train[j]
    if (int rc = ensure_scratch(ctx, (size_t)nsv * sizeof(int))) return rc;
    CK(cudaMemcpyAsync(ctx->scratch, ids.data(), (size_t)nsv * sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemcpyAsync(ctx->svcoef, coef.data(), (size_t)nsv * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
    const long long total = (long long)nsv * ctx->dec_pitch4;
    const int gblocks = (int)std::min<long long>((total + 255) / 256, ctx->num_sms * 16);
    if (ctx->x_u8) {            // the support vectors' bytes become scaled FP64 rows
        u8_gather_scale_kernel<<<gblocks, 256, 0, ctx->stream>>>(U8View(ctx->X, ctx->n, ctx->u8_mw, ctx->u8_tail_pitch), ctx->u8_pos, ctx->minv,
                                                                   ctx->maxv, ctx->d, static_cast<const int *>(ctx->scratch), nsv, ctx->dec_pitch4,
                                                                   static_cast<double *>(ctx->svX));
        CK(cudaMemsetAsync(ctx->svn, 0, (size_t)nsv * sizeof(double), ctx->stream));
    } else
    gather_rows_kernel<<<gblocks, 256, 0, ctx->stream>>>(
        ctx->X, ctx->x_f64 ? nullptr : ctx->sqn, ctx->pitch4, static_cast<const int *>(ctx->scratch), nsv, ctx->svX, ctx->svn);
    ctx->launches += 1;
    CK(cudaGetLastError());
    CK(cudaStreamSynchronize(ctx->stream));   // coef / ids are host temporaries
    return 0;
}

extern "C" int svmb200_smo_train(svmb200_ctx *ctx, const svmb200_params *params, double *alpha, double *b, svmb200_stats *stats)
{
    if (!ctx) return SVMB200_EINVAL;
    if (!params || !alpha || !b) FAIL(SVMB200_EINVAL, "params, alpha and b must not be NULL");
    if (!ctx->have_train) FAIL(SVMB200_ESTATE, "svmb200_smo_train called before svmb200_set_train_*");
    if (params->wss != 1 && params->wss != 2) FAIL(SVMB200_EINVAL, "wss=%d: 1 (the reference's first-order rule) or 2 (second-order choice of i_low)", params->wss);
    const int CH = resolve_chunk(params->accum_chunk);
    if (CH < 0) FAIL(SVMB200_EINVAL, "accum_chunk must be 0, 1, 2 or 4");
    if (params->max_iter < 1) FAIL(SVMB200_EINVAL, "max_iter must be >= 1");
    if (ctx->shard_world > 1 && params->max_iter >= (1 << 21))
        FAIL(SVMB200_ELIMIT, "row sharding numbers its exchanges with 21 bits: max_iter must be below %d", 1 << 21);
    CK(cudaSetDevice(ctx->device));
    const long long n = ctx->n;
    const int launches0 = ctx->launches;

    // persistent grid: one CTA per SM, fewer for tiny problems (the last 8 record places of the exchange stage the peer GPUs' records)
    const bool u8 = ctx->x_u8;
    const int Q = u8 ? 8 : q_for_pitch(ctx->pitch4, ctx->x_f64);
    const int QDEC = q_for_pitch(ctx->dec_pitch4, ctx->dec_f64);      // the decision kernels' row class (warm start)
    // fp32 rows: 8 warps taking rows in groups (SMO_DEFAULT_THREADS); FP64 rows: the same for the longest rows (Q = 13), else 12 warps, a row per step
    int threads = u8 ? SMO_U8_THREADS : ctx->x_f64 ? (Q == 13 ? 256 : 384) : SMO_DEFAULT_THREADS;
#ifdef SVM_TIMING      // measurement knobs exist in the profiling build only (libsvmb200_timing.so): the product reads no environment
    if (ctx->x_f64 && Q == 13 && getenv("SVMB200_F64_ONE_ROW")) threads = 384;
    if (!ctx->x_f64)
        if (const char *e = getenv("SVMB200_THREADS")) {   // 512 (384 at Q = 7) = a row per step
            const int t = atoi(e);
            if (t == 256 || t == 512 || (t == 384 && Q == 7)) threads = t;
        }
#endif
    const int warps = threads / 32;
    const int rps = u8 ? 8 : !ctx->x_f64 ? (threads <= 256 ? (Q == 8 ? 2 : 4) : 1) : (Q == 13 && threads == 256) ? 2 : 1;   // = smo_rows_per_step<XT, Q, THREADS>()
    const int tailw = u8 ? ctx->u8_tw : 0;
    // ring depth: as many groups as fit, at most 4 (two at the MNIST shape: one in use, one in flight); a row per step: 4 rows at Q = 7, else 2
    int kslots = u8 ? 3 : rps > 1 ? 4 : (Q == 7 && !ctx->x_f64) ? 4 : 2;          // uint8 rows: groups of the main class only, three deep
#ifdef SVM_TIMING
    if (const char *e = getenv("SVMB200_KSLOTS")) kslots = std::max(2, std::min(SMO_KSLOTS_MAX, atoi(e)));
#endif
    // uint8 rows: the ring carries the main class; a per-warp buffer receives the tails of a block of 32 rows (walked a lane per
    // row), of 16 or 8 rows when 32 do not fit beside a ring of two groups
    const int ring_pitch4 = u8 ? ctx->u8_mw / 4 : ctx->pitch4;
    const int tail_pitch = u8 ? ctx->u8_tail_pitch : 0;
    int tail_rows = 32;
    while (tail_rows > 8 && smo_smem_bytes(ring_pitch4, 2, warps, rps, tailw, tail_pitch, tail_rows) > (size_t)ctx->max_smem_optin) tail_rows /= 2;
    while (kslots > 2 && smo_smem_bytes(ring_pitch4, kslots, warps, rps, tailw, tail_pitch, tail_rows) > (size_t)ctx->max_smem_optin) --kslots;
    const size_t smem = smo_smem_bytes(ring_pitch4, kslots, warps, rps, tailw, tail_pitch, tail_rows);
    if (smem > (size_t)ctx->max_smem_optin) FAIL(SVMB200_ELIMIT, "row ring of %zu bytes does not fit in shared memory", smem);
    const long long n_own = n * (ctx->shard_rank + 1) / ctx->shard_world - n * ctx->shard_rank / ctx->shard_world;
    int grid = (int)std::max<long long>(1, std::min<long long>(std::min(ctx->num_sms, SMO_MAX_RECS - 8), (n_own + 7) / 8));
    // thread-block clusters of two CTAs (neighbouring SMs) halve the records of the exchange; the grid is then even
    int cluster = grid >= 2 ? 2 : 1;
#ifdef SVM_TIMING
    if (getenv("SVMB200_NO_CLUSTER")) cluster = 1;
#endif
    if (cluster == 2) grid &= ~1;

    // row cache
    int cache_cap = params->cache_rows;
    if (cache_cap < 0) {
        const size_t budget = (size_t)8 << 30;
        cache_cap = (int)std::min<long long>(n, (long long)(budget / ((size_t)n * sizeof(double))));
    }
    if (cache_cap == 1) cache_cap = 2;
    if (cache_cap > 0 && (size_t)cache_cap * n > ctx->cacheK_elems) {     // sized in elements: the same capacity needs more room when n grows
        // a context that is trained again and again on sets of changing size (the cascade's stages) should not free and
        // allocate gigabytes per call: grow by at least a half, up to the automatic budget
        size_t want = (size_t)cache_cap * n;
        if (ctx->cacheK_elems) want = std::max(want, std::min(ctx->cacheK_elems + ctx->cacheK_elems / 2, ((size_t)8 << 30) / sizeof(double)));
        if (dev_free(ctx, ctx->cacheK)) return SVMB200_ECUDA;
        ctx->cacheK_elems = 0;
        CK(cudaMalloc(&ctx->cacheK, want * sizeof(double)));
        ctx->cacheK_elems = want;
    }

    const size_t rec_bytes = (size_t)2 * SMO_REC_COPIES * SMO_MAX_RECS * sizeof(Rec) + 256;     // two generations of records (in copies) + the abort word
    if (!ctx->rec) CK(cudaMalloc(&ctx->rec, rec_bytes));
    const int trace_cap = std::max(0, params->trace_cap);
    if (trace_cap > ctx->trace_alloc) {
        if (dev_free(ctx, ctx->trace)) return SVMB200_ECUDA;
        CK(cudaMalloc(&ctx->trace, (size_t)trace_cap * 2 * sizeof(int)));
        ctx->trace_alloc = trace_cap;
    }

    // keep_cache: reuse the kernel rows cached by the previous run on this X (same capacity, arithmetic and gamma)
    const bool keep = params->keep_cache && cache_cap > 0 && ctx->cache_valid && ctx->cache_valid_cap == cache_cap &&
                      ctx->cache_valid_chunk == CH && ctx->cache_valid_gamma == params->gamma && ctx->shard_world == 1;
    CK(cudaEventRecord(ctx->ev0, ctx->stream));
    if (int rc = ensure_norms(ctx, CH)) return rc;
    CK(cudaMemsetAsync(ctx->rec, 0, rec_bytes, ctx->stream));
    if (params->warm_start) {
        // f_i = sum_{alpha_j != 0} alpha_j y_j K(x_j, x_i) - y_i   (mpi_svm_main2.cpp:169-183)
        CK(cudaMemcpyAsync(ctx->alpha, alpha, (size_t)n * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
        std::vector<int> y_host(n);
        CK(cudaMemcpyAsync(y_host.data(), ctx->y, (size_t)n * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
        CK(cudaStreamSynchronize(ctx->stream));
        std::vector<int> nz, nzy;
        for (long long j = 0; j < n; ++j)
            if (alpha[j] != 0.0) { nz.push_back((int)j); nzy.push_back(y_host[j]); }
        if (int rc = build_sv_block(ctx, nz, alpha, nzy)) return rc;
        ctx->model_ready = false;
        init_state_kernel<<<ctx->num_sms * 2, 256, 0, ctx->stream>>>(ctx->y, n, ctx->f, ctx->alpha, keep ? nullptr : ctx->cacheq, 0, 0);
        ctx->launches += 1;
        CK(cudaGetLastError());
        const void *Xd = ctx->X;
        if (u8) {               // the decision kernel walks scaled FP64 rows: one copy of all rows, kept until the next upload
            if (!ctx->Xdec) if (int rc = dev_alloc(ctx, &ctx->Xdec, (size_t)n * ctx->dec_pitch4 * 16)) return rc;
            if (!ctx->xdec_valid) {
                const long long total = n * ctx->dec_pitch4;
                u8_gather_scale_kernel<<<(int)std::min<long long>((total + 255) / 256, (long long)ctx->num_sms * 16), 256, 0, ctx->stream>>>(
                    U8View(ctx->X, n, ctx->u8_mw, ctx->u8_tail_pitch), ctx->u8_pos, ctx->minv, ctx->maxv, ctx->d, nullptr, n, ctx->dec_pitch4,
                    static_cast<double *>(ctx->Xdec));
                ctx->launches += 1;
                CK(cudaGetLastError());
                ctx->xdec_valid = true;
            }
            Xd = ctx->Xdec;
        }
        CK(launch_decision_any(KernelKey{ctx->dec_f64, QDEC, CH, 0}, Xd, n, ctx->svX, ctx->svn, ctx->svcoef, (int)nz.size(), ctx->dec_pitch4, params->gamma, 0.0,
                               ctx->y, ctx->f, ctx->stream));
        ctx->launches += 1;
    } else {
        init_state_kernel<<<ctx->num_sms * 2, 256, 0, ctx->stream>>>(ctx->y, n, ctx->f, ctx->alpha, keep ? nullptr : ctx->cacheq, 1, 1);
        ctx->launches += 1;
        CK(cudaGetLastError());
    }

    SmoArgs a{};
    a.X = ctx->X; a.sqn = (ctx->x_f64 || u8) ? nullptr : ctx->sqn; a.y = ctx->y; a.f = ctx->f; a.alpha = ctx->alpha;
    a.nmain = ctx->u8_nmain; a.ntail = ctx->u8_ntail; a.wt = ctx->u8_wt; a.mw = ctx->u8_mw; a.tw = ctx->u8_tw; a.wR = ctx->u8_wR;
    a.tail_pitch = tail_pitch; a.tail_rows = tail_rows;
    a.cacheq = cache_cap > 0 ? ctx->cacheq : nullptr;
    a.cacheK = ctx->cacheK; a.rec = ctx->rec;
    a.abort_word = reinterpret_cast<unsigned *>(reinterpret_cast<char *>(ctx->rec) + (size_t)2 * SMO_REC_COPIES * SMO_MAX_RECS * sizeof(Rec) + 128); a.trace = ctx->trace; a.result = ctx->result;
    a.n = n; a.pitch4 = ctx->pitch4; a.kslots = kslots; a.cache_cap = cache_cap; a.trace_cap = trace_cap;
    a.max_iter = params->max_iter;
    a.wss = params->wss;
    a.cache_counter0 = keep ? ctx->cache_counter : 0;
    a.C = params->C; a.gamma = params->gamma; a.eps = params->eps; a.tau = params->tau;
    a.policy_stream = L2_EVICT_FIRST; a.policy_keep = L2_EVICT_LAST; a.keep_per_1024 = 0;
    {
        // L2 residency: when this GPU's rows do not fit in L2, a share of them (evenly interleaved groups) is loaded with
        // evict-last and stays resident from one pass to the next; the rest streams with evict-first.  80 MB of the 126 MB
        // measured best at 60 000 x 784 (25.9 us per iteration; 64 MB 27.3, 88 MB 26.9, 96 MB 28.6, none 35.0).  It pays
        // only with the four-rows-per-step pass: with a row per step the SM side, not HBM, bounds the pass.
        const double x_bytes = (double)n_own * ctx->pitch4 * 16;
        double l2_keep_budget = 80.0 * 1024 * 1024;
#ifdef SVM_TIMING
        if (const char *e = getenv("SVMB200_L2KEEP_MB")) l2_keep_budget = atof(e) * 1024 * 1024;
#endif
        int persist = params->l2_persist;
        if (persist < 0) persist = rps > 1;
        if (x_bytes <= 96.0 * 1024 * 1024) { a.policy_stream = L2_EVICT_NORMAL; a.policy_keep = L2_EVICT_NORMAL; }
        else if (persist) a.keep_per_1024 = (int)std::min(1024.0, 1024.0 * l2_keep_budget / x_bytes);
    }
    a.peers = ctx->shard_world; a.rank = ctx->shard_rank;
    a.row_begin = n * ctx->shard_rank / ctx->shard_world;
    a.row_end = n * (ctx->shard_rank + 1) / ctx->shard_world;
    a.mailbox = ctx->mailbox;
    a.mailbox_copies = 1;
    for (int r = 0, first = 1; r < ctx->shard_world; ++r) {       // the clusters of the smallest grid among the GPUs (same rule as `grid` above)
        const long long own_r = n * (r + 1) / ctx->shard_world - n * r / ctx->shard_world;
        int g = (int)std::max<long long>(1, std::min<long long>(std::min(ctx->num_sms, SMO_MAX_RECS - 8), (own_r + 7) / 8));
        if (g >= 2) g /= 2;                                        // clusters of two (a launch that falls back to no clusters has more records)
        a.mailbox_copies = first ? std::min(SMO_MB_COPIES, g) : std::min(a.mailbox_copies, g);
        first = 0;
    }
    for (int p = 0; p < 8; ++p) a.peer_mailbox[p] = ctx->peer_mailbox[p];
    a.epoch = ++ctx->shard_epoch;
    a.timing = nullptr; a.timing_cap = 0;
#ifdef SVM_TIMING
    {
        static long long *tbuf = nullptr;
        if (!tbuf) CK(cudaMalloc(&tbuf, (4096 * 16 + 8 * 160 * 4 + 4096 * 8) * sizeof(long long)));
        CK(cudaMemsetAsync(tbuf, 0, (4096 * 16 + 8 * 160 * 4 + 4096 * 8) * sizeof(long long), ctx->stream));
        a.timing = tbuf; a.timing_cap = 4096;
        g_timing_dev = tbuf;
    }
#endif

    if (ctx->shard_world > 1) {
        // the GPUs of a row-sharded run meet on the device before the timed launch: every process reaches this point at
        // its own time (host skew of a millisecond or more), and the persistent kernels run in lock step once all are there
        PeerFlags pf{};
        for (int p = 0; p < 8; ++p) pf.peer[p] = ctx->peer_mailbox[p] ? reinterpret_cast<unsigned long long *>(ctx->peer_mailbox[p] + SMO_MB_RECS) : nullptr;
        peer_rendezvous_kernel<<<1, 32, 0, ctx->stream>>>(pf, ctx->shard_rank, ctx->shard_world, (unsigned long long)a.epoch);
        ctx->launches += 1;
        CK(cudaGetLastError());
    }
    CK(cudaEventRecord(ctx->ev1, ctx->stream));
    auto launch = [&](int cl) -> cudaError_t {
        if (u8) {               // instantiated on the exact number of 128-byte slots of the main class
            switch (ctx->u8_mw / 32) {
#define SVM_U8_CASE(S) case S: return launch_persistent(smo_persistent_kernel<unsigned char, S, 1, SMO_U8_THREADS>, a, grid, SMO_U8_THREADS, cl, smem, ctx->stream);
                SVM_U8_CASE(0) SVM_U8_CASE(1) SVM_U8_CASE(2) SVM_U8_CASE(3) SVM_U8_CASE(4) SVM_U8_CASE(5) SVM_U8_CASE(6) SVM_U8_CASE(7)
                default: SVM_U8_CASE(8)
#undef SVM_U8_CASE
            }
        }
        return launch_smo_any(KernelKey{ctx->x_f64, Q, CH, threads}, a, grid, cl, smem, ctx->stream);
    };
    cudaError_t le = launch(cluster);
    if (le != cudaSuccess && cluster > 1) {      // this device cannot co-schedule the clusters: same kernel, cluster size 1
#ifdef SVM_TIMING
        fprintf(stderr, "svmb200: cluster launch refused (%s), falling back to cluster size 1\n", cudaGetErrorString(le));
#endif
        (void)cudaGetLastError();
        cluster = 1;
        le = launch(cluster);
    }
    CK(le);
    ctx->cluster_size = cluster;
    ctx->launches += 1;
    CK(cudaEventRecord(ctx->ev2, ctx->stream));

    SmoResult r;
    CK(cudaMemcpyAsync(&r, ctx->result, sizeof r, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaMemcpyAsync(alpha, ctx->alpha, (size_t)n * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));

    *b = (r.b_high + r.b_low) / 2;   // main3.cpp:291
    ctx->b = *b;
    ctx->params = *params;
    ctx->params.accum_chunk = CH;
    ctx->have_params = true;
    ctx->alpha_host.assign(alpha, alpha + n);
    ctx->sv_ids.clear();
    for (long long i = 0; i < n; ++i)            // get_SV_indices, main3.cpp:297-304
        if (alpha[i] > params->sv_tol) ctx->sv_ids.push_back((int)i);
    ctx->cache_valid = cache_cap > 0 && ctx->shard_world == 1;
    ctx->cache_valid_cap = cache_cap; ctx->cache_valid_chunk = CH; ctx->cache_valid_gamma = params->gamma;
    ctx->cache_counter = r.cache_counter;
    ctx->trace_count = std::min(trace_cap, r.num_iter - 1);
    ctx->trained = true;
    ctx->model_ready = false;

    if (stats) {
        float ms_init = 0.f, ms_smo = 0.f;
        CK(cudaEventElapsedTime(&ms_init, ctx->ev0, ctx->ev1));
        CK(cudaEventElapsedTime(&ms_smo, ctx->ev1, ctx->ev2));
        stats->num_iter = r.num_iter;
        stats->stop_reason = r.stop_reason;
        stats->i_high = r.i_high;
        stats->i_low = r.i_low;
        stats->b_high = r.b_high;
        stats->b_low = r.b_low;
        stats->smo_ms = ms_smo;
        stats->init_ms = ms_init;
        stats->x_passes = r.x_passes;
        stats->cache_hits = r.cache_hits;
        stats->x_bytes_streamed = r.x_passes * n * (long long)ctx->pitch4 * 16;
        stats->n_sv = (int)ctx->sv_ids.size();
        stats->grid_blocks = grid;
        stats->launches = ctx->launches - launches0;
        stats->cluster_size = ctx->cluster_size;
    }
    return SVMB200_OK;
}

// ---- row sharding across GPUs: one process per GPU, X replicated, rows [n*rank/world, n*(rank+1)/world) owned
extern "C" int svmb200_shard_setup(svmb200_ctx *ctx, int32_t rank, int32_t world, void *handle_out)
{
    if (!ctx) return SVMB200_EINVAL;
    if (world < 1 || world > 8 || rank < 0 || rank >= world || !handle_out) FAIL(SVMB200_EINVAL, "rank/world out of range (world <= 8)");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    CK(cudaSetDevice(ctx->device));
    if (ctx->mailbox) FAIL(SVMB200_ESTATE, "svmb200_shard_setup called twice");
    CK(cudaMalloc(&ctx->mailbox, (size_t)(SMO_MB_RECS + 1) * sizeof(Rec)));      // [2][copies][8] records + 8 rendezvous flags
    CK(cudaMemset(ctx->mailbox, 0, (size_t)(SMO_MB_RECS + 1) * sizeof(Rec)));
    cudaIpcMemHandle_t h;
    CK(cudaIpcGetMemHandle(&h, ctx->mailbox));
    memcpy(handle_out, &h, sizeof h);
    ctx->shard_rank = rank;
    ctx->shard_world = 1;          // becomes `world` in svmb200_shard_connect
    ctx->peer_mailbox[rank] = ctx->mailbox;
    return SVMB200_OK;
}

extern "C" int svmb200_shard_connect(svmb200_ctx *ctx, int32_t world, const void *handles)
{
    if (!ctx) return SVMB200_EINVAL;
    if (!ctx->mailbox) FAIL(SVMB200_ESTATE, "svmb200_shard_connect before svmb200_shard_setup");
    if (world < 1 || world > 8 || ctx->shard_rank >= world || !handles) FAIL(SVMB200_EINVAL, "bad world / handles");
    CK(cudaSetDevice(ctx->device));
    for (int p = 0; p < world; ++p) {
        if (p == ctx->shard_rank) continue;
        cudaIpcMemHandle_t h;
        memcpy(&h, static_cast<const char *>(handles) + (size_t)p * 64, 64);
        void *ptr = nullptr;
        CK(cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess));
        ctx->peer_mailbox[p] = static_cast<Rec *>(ptr);
    }
    ctx->shard_world = world;
    ctx->shard_epoch = 0;
    return SVMB200_OK;
}

// install a model (alpha of all rows, b) -- used after a sharded run, when every GPU holds only its rows' alpha
extern "C" int svmb200_set_model(svmb200_ctx *ctx, const double *alpha, double b)
{
    if (!ctx || !alpha) return SVMB200_EINVAL;
    if (!ctx->have_train) FAIL(SVMB200_ESTATE, "no training set");
    CK(cudaSetDevice(ctx->device));
    // gamma, sv_tol and the arithmetic are those of the last svmb200_smo_train on this context (a new upload does not
    // forget them); a context that never trained takes the defaults
    if (!ctx->have_params) { svmb200_default_params(&ctx->params); ctx->params.accum_chunk = DEFAULT_CHUNK; ctx->have_params = true; }
    ctx->alpha_host.assign(alpha, alpha + ctx->n);
    ctx->b = b;
    ctx->sv_ids.clear();
    for (long long i = 0; i < ctx->n; ++i)
        if (alpha[i] > ctx->params.sv_tol) ctx->sv_ids.push_back((int)i);
    CK(cudaMemcpy(ctx->alpha, alpha, (size_t)ctx->n * sizeof(double), cudaMemcpyHostToDevice));
    ctx->trained = true;
    ctx->model_ready = false;
    return SVMB200_OK;
}

static int ensure_model(svmb200_ctx *ctx);

// ---- model persistence: the dump the reference left commented out (mpi_svm_main2.cpp:686-697: SV ids, labels,
// alphas, b), as one binary file that also carries the scaled SV rows and the scaling statistics, so that a
// predict-only process needs no training set.
namespace {
struct ModelHeader {
    char magic[8];          // "SVMB200M"
    uint32_t version;
    int32_t d, pitch4, n_sv, accum_chunk, have_stats, x_f64, reserved;
    double gamma, b, sv_tol;
};
}

extern "C" int svmb200_save_model(svmb200_ctx *ctx, const char *path)
{
    if (!ctx || !path) return SVMB200_EINVAL;
    if (!ctx->trained) FAIL(SVMB200_ESTATE, "no model to save");
    CK(cudaSetDevice(ctx->device));
    if (int rc = ensure_model(ctx)) return rc;
    const int nsv = ctx->model_nsv;
    ModelHeader h{};
    memcpy(h.magic, "SVMB200M", 8);
    h.version = 2; h.x_f64 = ctx->dec_f64 ? 1 : 0; h.d = ctx->d; h.pitch4 = ctx->dec_pitch4; h.n_sv = nsv; h.accum_chunk = ctx->params.accum_chunk;
    h.have_stats = ctx->have_stats ? 1 : 0; h.gamma = ctx->params.gamma; h.b = ctx->b; h.sv_tol = ctx->params.sv_tol;
    std::vector<double> mn(ctx->d), mx(ctx->d), al(nsv);
    std::vector<float> rows((size_t)nsv * ctx->dec_pitch4 * 4);
    if (ctx->have_stats) {
        CK(cudaMemcpy(mn.data(), ctx->minv, (size_t)ctx->d * sizeof(double), cudaMemcpyDeviceToHost));
        CK(cudaMemcpy(mx.data(), ctx->maxv, (size_t)ctx->d * sizeof(double), cudaMemcpyDeviceToHost));
    }
    if (nsv) CK(cudaMemcpy(rows.data(), ctx->svX, rows.size() * sizeof(float), cudaMemcpyDeviceToHost));
    for (int k = 0; k < nsv; ++k) al[k] = ctx->model_alpha[k];
    FILE *f = fopen(path, "wb");
    if (!f) FAIL(SVMB200_EINVAL, "cannot open %s for writing", path);
    bool ok = fwrite(&h, sizeof h, 1, f) == 1;
    if (ctx->have_stats) ok = ok && fwrite(mn.data(), sizeof(double), ctx->d, f) == (size_t)ctx->d && fwrite(mx.data(), sizeof(double), ctx->d, f) == (size_t)ctx->d;
    ok = ok && fwrite(ctx->sv_ids.data(), sizeof(int), nsv, f) == (size_t)nsv && fwrite(ctx->model_y.data(), sizeof(int), nsv, f) == (size_t)nsv &&
         fwrite(al.data(), sizeof(double), nsv, f) == (size_t)nsv && fwrite(rows.data(), sizeof(float), rows.size(), f) == rows.size();
    ok = (fclose(f) == 0) && ok;
    if (!ok) FAIL(SVMB200_EINVAL, "short write to %s", path);
    return SVMB200_OK;
}

extern "C" int svmb200_load_model(svmb200_ctx *ctx, const char *path)
{
    if (!ctx || !path) return SVMB200_EINVAL;
    CK(cudaSetDevice(ctx->device));
    FILE *f = fopen(path, "rb");
    if (!f) FAIL(SVMB200_EINVAL, "cannot open %s", path);
    ModelHeader h{};
    bool ok = fread(&h, sizeof h, 1, f) == 1 && memcmp(h.magic, "SVMB200M", 8) == 0 && h.version == 2 && h.d > 0 && h.d <= 1024 &&
              h.pitch4 == (h.x_f64 ? (h.d + 1) / 2 : (h.d + 3) / 4) && h.pitch4 <= max_pitch(h.x_f64 != 0) && h.n_sv >= 0 && resolve_chunk(h.accum_chunk) > 0;
    if (!ok) { fclose(f); FAIL(SVMB200_EINVAL, "%s is not a svmb200 model file (version 2)", path); }
    const int nsv = h.n_sv, d = h.d;
    std::vector<double> mn(d), mx(d), al(nsv), coef(nsv);
    std::vector<int> ids(nsv), ys(nsv);
    std::vector<float> rows((size_t)nsv * h.pitch4 * 4);
    if (h.have_stats) ok = fread(mn.data(), sizeof(double), d, f) == (size_t)d && fread(mx.data(), sizeof(double), d, f) == (size_t)d;
    ok = ok && fread(ids.data(), sizeof(int), nsv, f) == (size_t)nsv && fread(ys.data(), sizeof(int), nsv, f) == (size_t)nsv &&
         fread(al.data(), sizeof(double), nsv, f) == (size_t)nsv && fread(rows.data(), sizeof(float), rows.size(), f) == rows.size();
    fclose(f);
    if (!ok) FAIL(SVMB200_EINVAL, "%s is truncated", path);
    // the context now holds a model only: any training set is dropped
    if (dev_free(ctx, ctx->X) || dev_free(ctx, ctx->y) || dev_free(ctx, ctx->sqn) || dev_free(ctx, ctx->f) || dev_free(ctx, ctx->alpha) ||
        dev_free(ctx, ctx->cacheq) || dev_free(ctx, ctx->minv) || dev_free(ctx, ctx->maxv) || dev_free(ctx, ctx->svX) ||
        dev_free(ctx, ctx->svcoef) || dev_free(ctx, ctx->svn) || dev_free(ctx, ctx->u8_nmain) || dev_free(ctx, ctx->u8_ntail) || dev_free(ctx, ctx->Xdec))
        return SVMB200_ECUDA;
    ctx->n = 0; ctx->d = d; ctx->pitch4 = h.pitch4; ctx->x_f64 = h.x_f64 != 0; ctx->x_u8 = false; ctx->dec_f64 = ctx->x_f64; ctx->dec_pitch4 = h.pitch4; ctx->have_train = false; ctx->sqn_chunk = 0; ctx->sv_alloc = std::max(nsv, 1);
    ctx->svX_bytes = (size_t)ctx->sv_alloc * h.pitch4 * 16;
    CK(cudaMalloc(&ctx->minv, (size_t)d * sizeof(double)));
    CK(cudaMalloc(&ctx->maxv, (size_t)d * sizeof(double)));
    ctx->svP_valid = false;
    CK(cudaMalloc(&ctx->svX, (size_t)ctx->sv_alloc * h.pitch4 * 16));
    CK(cudaMalloc(&ctx->svcoef, (size_t)ctx->sv_alloc * sizeof(double)));
    CK(cudaMalloc(&ctx->svn, (size_t)ctx->sv_alloc * sizeof(double)));
    if (h.have_stats) {
        CK(cudaMemcpy(ctx->minv, mn.data(), (size_t)d * sizeof(double), cudaMemcpyHostToDevice));
        CK(cudaMemcpy(ctx->maxv, mx.data(), (size_t)d * sizeof(double), cudaMemcpyHostToDevice));
    }
    for (int k = 0; k < nsv; ++k) coef[k] = al[k] * ys[k];
    if (nsv) {
        CK(cudaMemcpy(ctx->svX, rows.data(), rows.size() * sizeof(float), cudaMemcpyHostToDevice));
        CK(cudaMemcpy(ctx->svcoef, coef.data(), (size_t)nsv * sizeof(double), cudaMemcpyHostToDevice));
        // norms of the SV rows (fp32 storage), formed on the device by the same code as everywhere else
        if (!h.x_f64) {
            cudaError_t e = launch_sqnorm_any(KernelKey{false, q_for_pitch(h.pitch4, false), h.accum_chunk, 512},
                                              static_cast<const float *>(ctx->svX), nsv, h.pitch4, ctx->svn, ctx->stream);
            ctx->launches += 1;
            CK(e);
        }
        CK(cudaStreamSynchronize(ctx->stream));
    }
    svmb200_default_params(&ctx->params);
    ctx->params.gamma = h.gamma; ctx->params.sv_tol = h.sv_tol; ctx->params.accum_chunk = h.accum_chunk;
    ctx->b = h.b; ctx->have_stats = h.have_stats != 0;
    ctx->sv_ids = ids; ctx->model_y = ys; ctx->model_alpha = al; ctx->alpha_host.clear(); ctx->have_params = true;
    ctx->model_nsv = nsv; ctx->model_ready = true; ctx->trained = true;
    return SVMB200_OK;
}

// ---- tensor-core decision function
extern "C" int svmb200_set_tensor_cores(svmb200_ctx *ctx, int32_t enable, double recheck_threshold)
{
    if (!ctx) return SVMB200_EINVAL;
    if (enable && !(recheck_threshold > 0.0)) FAIL(SVMB200_EINVAL, "recheck_threshold must be positive");
    ctx->tc_predict = enable != 0;
    if (enable) ctx->tc_threshold = recheck_threshold;
    return SVMB200_OK;
}

// tensor-core scores of m query rows against a support-vector block: operands packed once (tc_pack_kernel) and streamed
// by bulk copies through the warp-specialised kernel
static size_t tc_panel_bytes(const svmb200_ctx *ctx, long long rows)
{
    return (size_t)((rows + TC_M - 1) / TC_M) * ((ctx->d + TC_KB - 1) / TC_KB) * TCP_PANEL_BYTES;
}

static size_t tc_partial_bytes(const svmb200_ctx *ctx, long long m, int nsv)
{
    int ctas = 0, slots = 0;
    const long long qtiles = (m + TC_M - 1) / TC_M;
    tcp_plan(qtiles, (nsv + TC_N - 1) / TC_N, ctx->num_sms, &ctas, &slots);
    return (size_t)slots * qtiles * TC_M * sizeof(double);
}

// qP: tc_panel_bytes(m) of device memory for the packed query rows; partial: tc_partial_bytes(m, nsv).  svP: packed support
// vectors, or NULL = the current model's (ctx->svX), packed once per model and kept.
static int launch_tc_scores(svmb200_ctx *ctx, const float *Xq, const double *qn, long long m, unsigned char *qP, double *partial,
                            const float *svX, const double *svn, const double *svcoef, int nsv, unsigned char *svP, double init,
                            double *dscores, float *dots)
{
    const int pitch = ctx->dec_pitch4 * 4;
    const int kblocks = (ctx->d + TC_KB - 1) / TC_KB;
    const int blocks = ctx->num_sms * 8;
    if (!svP) {
        if (!ctx->svP_valid) {
            const size_t need = tc_panel_bytes(ctx, nsv);
            if (need > ctx->svP_bytes) {
                if (dev_free(ctx, ctx->svP)) return SVMB200_ECUDA;
                ctx->svP_bytes = 0;
                CK(cudaMalloc(&ctx->svP, need + need / 8));
                ctx->svP_bytes = need + need / 8;
            }
            tc_pack_kernel<<<blocks, 256, 0, ctx->stream>>>(svX, nsv, pitch, ctx->d, kblocks, ctx->svP);
            ctx->launches += 1;
            CK(cudaGetLastError());
            ctx->svP_valid = true;
        }
        svP = ctx->svP;
    } else {
        tc_pack_kernel<<<blocks, 256, 0, ctx->stream>>>(svX, nsv, pitch, ctx->d, kblocks, svP);
        ctx->launches += 1;
    }
    tc_pack_kernel<<<blocks, 256, 0, ctx->stream>>>(Xq, m, pitch, ctx->d, kblocks, qP);
    CK(cudaGetLastError());
    int ctas = 0, slots = 0;
    const long long qtiles = (m + TC_M - 1) / TC_M, mpad = qtiles * TC_M;
    tcp_plan(qtiles, (nsv + TC_N - 1) / TC_N, ctx->num_sms, &ctas, &slots);
    CK(cudaMemsetAsync(partial, 0, (size_t)slots * mpad * sizeof(double), ctx->stream));
    CK(cudaFuncSetAttribute(decision_tc_pipelined_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)decision_tc_pipelined_smem_bytes()));
    decision_tc_pipelined_kernel<<<ctas, TCP_THREADS, decision_tc_pipelined_smem_bytes(), ctx->stream>>>(
        qP, qn, m, svP, svn, svcoef, nsv, kblocks, ctx->params.gamma, partial, mpad, dots);
    tc_sum_partials_kernel<<<(unsigned)((m + 255) / 256), 256, 0, ctx->stream>>>(partial, slots, mpad, m, init, dscores);
    ctx->launches += 1;
    ctx->launches += 2;
    CK(cudaGetLastError());
    return 0;
}

// introspection: the raw tensor-core dot products of the first 128 training rows against themselves (128 x 128),
// and how many rows the last svmb200_predict_* sent to the exact kernel
extern "C" int svmb200_debug_tc_dots(svmb200_ctx *ctx, float *dots, int64_t *rechecked)
{
    if (!ctx) return SVMB200_EINVAL;
    if (rechecked) *rechecked = ctx->tc_rechecked;
    if (!dots) return SVMB200_OK;
    if (!ctx->have_train || ctx->x_f64 || ctx->x_u8) FAIL(SVMB200_ESTATE, "needs an fp32 training set");
    CK(cudaSetDevice(ctx->device));
    const long long rows = std::min<long long>(ctx->n, TC_M);
    if (int rc = ensure_norms(ctx, DEFAULT_CHUNK)) return rc;
    const size_t head = (size_t)TC_M * TC_N * sizeof(float) + (size_t)TC_M * 2 * sizeof(double), panel = tc_panel_bytes(ctx, rows);
    const size_t part = tc_partial_bytes(ctx, rows, (int)rows);
    if (int rc = ensure_work(ctx, 0, head + 2 * panel + part)) return rc;
    float *d_dots = static_cast<float *>(ctx->work[0]);
    double *coef = reinterpret_cast<double *>(d_dots + TC_M * TC_N), *out = coef + TC_M;
    unsigned char *qP = static_cast<unsigned char *>(ctx->work[0]) + head;
    CK(cudaMemsetAsync(ctx->work[0], 0, head, ctx->stream));
    if (int rc = launch_tc_scores(ctx, static_cast<const float *>(ctx->X), ctx->sqn, rows, qP, reinterpret_cast<double *>(qP + 2 * panel),
                                  static_cast<const float *>(ctx->X), ctx->sqn, coef, (int)rows, qP + panel, 0.0, out, d_dots))
        return rc;
    CK(cudaMemcpyAsync(dots, d_dots, (size_t)TC_M * TC_N * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return SVMB200_OK;
}

extern "C" int svmb200_sv_ids(svmb200_ctx *ctx, int32_t *ids, int32_t cap, int32_t *count)
{
    if (!ctx) return SVMB200_EINVAL;
    if (!ctx->trained) FAIL(SVMB200_ESTATE, "svmb200_sv_ids called before svmb200_smo_train");
    const int c = (int)ctx->sv_ids.size();
    if (count) *count = c;
    if (ids) {
        if (cap < c) FAIL(SVMB200_EINVAL, "ids capacity %d < %d support vectors", cap, c);
        memcpy(ids, ctx->sv_ids.data(), (size_t)c * sizeof(int));
    }
    return SVMB200_OK;
}

// dense support-vector block (rows, norms, alpha*y) of the current model, built once per model
static int ensure_model(svmb200_ctx *ctx)
{
    if (ctx->model_ready) return 0;
    if (!ctx->have_train) FAIL(SVMB200_ESTATE, "no training set to take the support vectors from");
    std::vector<int> y_host(ctx->n);
    CK(cudaMemcpyAsync(y_host.data(), ctx->y, (size_t)ctx->n * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    ctx->model_y.resize(ctx->sv_ids.size());
    ctx->model_alpha.resize(ctx->sv_ids.size());
    for (size_t k = 0; k < ctx->model_y.size(); ++k) {
        ctx->model_y[k] = y_host[ctx->sv_ids[k]];
        ctx->model_alpha[k] = ctx->alpha_host[ctx->sv_ids[k]];
    }
    if (int rc = ensure_norms(ctx, ctx->params.accum_chunk)) return rc;
    if (int rc = build_sv_block(ctx, ctx->sv_ids, ctx->alpha_host.data(), ctx->model_y)) return rc;
    ctx->model_nsv = (int)ctx->sv_ids.size();
    ctx->model_ready = true;
    return 0;
}

template <typename T>
static int predict_impl(svmb200_ctx *ctx, const T *X, const int32_t *y_test, int64_t m, bool scale, int32_t ge_zero,
                        int32_t *correct, double *scores)
{
    if (!ctx) return SVMB200_EINVAL;
    if (!ctx->trained) FAIL(SVMB200_ESTATE, "svmb200_predict_* called before svmb200_smo_train");
    if (!X || m <= 0) FAIL(SVMB200_EINVAL, "X must not be NULL and m must be positive");
    if (scale && !ctx->have_stats) FAIL(SVMB200_ESTATE, "the training set was uploaded pre-scaled: use svmb200_predict_scaled_f64");
    CK(cudaSetDevice(ctx->device));
    const int pitch4 = ctx->dec_pitch4;
    if (int rc = ensure_model(ctx)) return rc;

    // one grow-only work area: query rows | scores | labels + counter | (tensor cores: norms | list + counter | packed rows)
    auto up = [](size_t v) { return (v + 255) & ~(size_t)255; };
    const bool use_tc = ctx->tc_predict && !ctx->dec_f64 && ctx->model_nsv > 0;
    const size_t o_x = 0, o_scores = o_x + up((size_t)m * pitch4 * 16), o_y = o_scores + up((size_t)m * sizeof(double)),
                 o_qn = o_y + up(((size_t)m + 1) * sizeof(int)), o_list = o_qn + (use_tc ? up((size_t)m * sizeof(double)) : 0),
                 o_panel = o_list + (use_tc ? up(((size_t)m + 1) * sizeof(int)) : 0), o_part = o_panel + (use_tc ? tc_panel_bytes(ctx, m) : 0),
                 o_end = o_part + (use_tc ? tc_partial_bytes(ctx, m, ctx->model_nsv) : 0);
    if (int rc = ensure_work(ctx, 0, o_end)) return rc;
    char *w = static_cast<char *>(ctx->work[0]);
    void *Xq = w + o_x;
    double *dscores = reinterpret_cast<double *>(w + o_scores);
    int *dy = reinterpret_cast<int *>(w + o_y), *dcorrect = dy + m;
    if (int rc = ingest<T>(ctx, X, m, false, scale, Xq, ctx->dec_f64, pitch4)) return rc;
    const KernelKey key{ctx->dec_f64, q_for_pitch(pitch4, ctx->dec_f64), ctx->params.accum_chunk, 0};
    if (use_tc) {
        // tensor cores first (TF32x3, approximate), then the exact kernel for every row whose |score| is under the threshold
        const int nsv = ctx->model_nsv;
        double *qn = reinterpret_cast<double *>(w + o_qn);
        int *list = reinterpret_cast<int *>(w + o_list);
        CK(launch_sqnorm_any(key, static_cast<const float *>(Xq), m, pitch4, qn, ctx->stream));
        if (int rc = launch_tc_scores(ctx, static_cast<const float *>(Xq), qn, m, reinterpret_cast<unsigned char *>(w + o_panel),
                                      reinterpret_cast<double *>(w + o_part), static_cast<const float *>(ctx->svX), ctx->svn, ctx->svcoef, nsv, nullptr, -ctx->b, dscores, nullptr))
            return rc;
        int n_unc = 0;
        CK(cudaMemsetAsync(list + m, 0, sizeof(int), ctx->stream));
        flag_uncertain_kernel<<<std::min<long long>((m + 255) / 256, ctx->num_sms * 4), 256, 0, ctx->stream>>>(dscores, m, ctx->tc_threshold, list, list + m);
        CK(cudaGetLastError());
        CK(cudaMemcpyAsync(&n_unc, list + m, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
        CK(cudaStreamSynchronize(ctx->stream));
        ctx->launches += 2;
        if (n_unc > 0) {
            const size_t xu_bytes = up((size_t)n_unc * pitch4 * 16), ex_bytes = up((size_t)n_unc * sizeof(double));
            if (int rc = ensure_work(ctx, 1, xu_bytes + ex_bytes + decision_rows_work_bytes(n_unc, nsv))) return rc;
            void *Xu = ctx->work[1];
            double *exact = reinterpret_cast<double *>(static_cast<char *>(ctx->work[1]) + xu_bytes);
            void *rows_work = static_cast<char *>(ctx->work[1]) + xu_bytes + ex_bytes;
            const long long total = (long long)n_unc * pitch4;
            gather_rows_kernel<<<(int)std::min<long long>((total + 255) / 256, ctx->num_sms * 16), 256, 0, ctx->stream>>>(
                Xq, nullptr, pitch4, list, n_unc, Xu, qn /* norms unused by the exact kernel: scratch */);
            CK(cudaGetLastError());
            CK(launch_decision_rows_any(key, Xu, n_unc, ctx->svX, ctx->svn, ctx->svcoef, nsv, pitch4, ctx->params.gamma, -ctx->b, exact,
                                        ctx->max_smem_optin, ctx->num_sms, rows_work, ctx->stream));
            scatter_scores_kernel<<<(n_unc + 255) / 256, 256, 0, ctx->stream>>>(exact, list, n_unc, dscores);
            CK(cudaGetLastError());
            ctx->launches += 3;
        }
        ctx->tc_rechecked = n_unc;
    } else {
        CK(launch_decision_any(key, Xq, m, ctx->svX, ctx->svn, ctx->svcoef, ctx->model_nsv, pitch4, ctx->params.gamma, -ctx->b, nullptr, dscores,
                               ctx->stream));
        ctx->launches += 1;
    }
    int host_correct = 0;
    if (y_test) {
        CK(cudaMemcpyAsync(dy, y_test, (size_t)m * sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
        CK(cudaMemsetAsync(dcorrect, 0, sizeof(int), ctx->stream));
        count_correct_kernel<<<std::min<long long>((m + 255) / 256, ctx->num_sms * 4), 256, 0, ctx->stream>>>(dscores, dy, m, ge_zero, dcorrect);
        ctx->launches += 1;
        CK(cudaGetLastError());
        CK(cudaMemcpyAsync(&host_correct, dcorrect, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    }
    if (scores) CK(cudaMemcpyAsync(scores, dscores, (size_t)m * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    if (correct) *correct = host_correct;
    return SVMB200_OK;
}

extern "C" {

int svmb200_predict_u8(svmb200_ctx *ctx, const uint8_t *X, const int32_t *y_test, int64_t m, int32_t ge_zero, int32_t *correct, double *scores)
{
    return predict_impl<uint8_t>(ctx, X, y_test, m, true, ge_zero, correct, scores);
}
int svmb200_predict_f64(svmb200_ctx *ctx, const double *X, const int32_t *y_test, int64_t m, int32_t ge_zero, int32_t *correct, double *scores)
{
    return predict_impl<double>(ctx, X, y_test, m, true, ge_zero, correct, scores);
}
int svmb200_predict_scaled_f64(svmb200_ctx *ctx, const double *Xs, const int32_t *y_test, int64_t m, int32_t ge_zero, int32_t *correct, double *scores)
{
    return predict_impl<double>(ctx, Xs, y_test, m, false, ge_zero, correct, scores);
}

int svmb200_get_trace(svmb200_ctx *ctx, int32_t *pairs, int32_t cap, int32_t *count)
{
    if (!ctx) return SVMB200_EINVAL;
    if (!ctx->trained) FAIL(SVMB200_ESTATE, "no training run to trace");
    const int c = std::min(cap, ctx->trace_count);
    if (count) *count = c;
    if (pairs && c > 0) {
        CK(cudaSetDevice(ctx->device));
        CK(cudaMemcpy(pairs, ctx->trace, (size_t)c * 2 * sizeof(int), cudaMemcpyDeviceToHost));
    }
    return SVMB200_OK;
}

int svmb200_get_f(svmb200_ctx *ctx, double *f)
{
    if (!ctx || !f) return SVMB200_EINVAL;
    if (!ctx->trained) FAIL(SVMB200_ESTATE, "no training run");
    CK(cudaSetDevice(ctx->device));
    CK(cudaMemcpy(f, ctx->f, (size_t)ctx->n * sizeof(double), cudaMemcpyDeviceToHost));
    return SVMB200_OK;
}

int svmb200_kernel_row(svmb200_ctx *ctx, int64_t i, double gamma, int32_t accum_chunk, double *row)
{
    if (!ctx || !row) return SVMB200_EINVAL;
    if (!ctx->have_train) FAIL(SVMB200_ESTATE, "no training set");
    if (i < 0 || i >= ctx->n) FAIL(SVMB200_EINVAL, "row %lld out of range", (long long)i);
    const int CH = resolve_chunk(accum_chunk);
    if (CH < 0) FAIL(SVMB200_EINVAL, "accum_chunk must be 0, 1, 2 or 4");
    CK(cudaSetDevice(ctx->device));
    if (int rc = ensure_norms(ctx, CH)) return rc;
    if (int rc = ensure_scratch(ctx, (size_t)ctx->n * sizeof(double))) return rc;
    double *out = static_cast<double *>(ctx->scratch);
    cudaError_t e;
    if (ctx->x_u8) {
        u8_kernel_row_kernel<<<(unsigned)((ctx->n + 127) / 128), 128, 0, ctx->stream>>>(
            U8View(ctx->X, ctx->n, ctx->u8_mw, ctx->u8_tail_pitch), ctx->n, ctx->u8_mw, ctx->u8_tw, ctx->u8_wt, ctx->u8_wR, ctx->u8_nmain, ctx->u8_ntail,
            i, gamma, out);
        e = cudaGetLastError();
    } else
        e = launch_krow_any(KernelKey{ctx->x_f64, q_for_pitch(ctx->pitch4, ctx->x_f64), CH, 0}, ctx->X, ctx->x_f64 ? nullptr : ctx->sqn,
                            ctx->n, ctx->pitch4, i, gamma, out, ctx->stream);
    ctx->launches += 1;
    CK(e);
    CK(cudaMemcpyAsync(row, out, (size_t)ctx->n * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return SVMB200_OK;
}

int svmb200_select_pair(svmb200_ctx *ctx, const double *alpha, const double *f, const int32_t *y, int64_t n, double C,
                        double eps, int64_t *i_high, int64_t *i_low)
{
    if (!ctx || !alpha || !f || !y || !i_high || !i_low || n <= 0) return SVMB200_EINVAL;
    CK(cudaSetDevice(ctx->device));
    const size_t need = (size_t)n * (2 * sizeof(double) + sizeof(int)) + 64;
    if (int rc = ensure_scratch(ctx, need)) return rc;
    double *da = static_cast<double *>(ctx->scratch), *df = da + n;
    long long *dout = reinterpret_cast<long long *>(df + n);
    int *dy = reinterpret_cast<int *>(dout + 2);
    CK(cudaMemcpyAsync(da, alpha, (size_t)n * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemcpyAsync(df, f, (size_t)n * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemcpyAsync(dy, y, (size_t)n * sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
    select_pair_kernel<<<1, 1024, 0, ctx->stream>>>(da, df, dy, n, C, eps, dout);
    ctx->launches += 1;
    CK(cudaGetLastError());
    long long out[2];
    CK(cudaMemcpyAsync(out, dout, sizeof out, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    *i_high = out[0];
    *i_low = out[1];
    return SVMB200_OK;
}

int svmb200_pair_update(svmb200_ctx *ctx, int32_t y_high, int32_t y_low, double alpha_high, double alpha_low, double b_high,
                        double b_low, double K12, double C, double eps, double *out7, int32_t *stop)
{
    if (!ctx || !out7 || !stop) return SVMB200_EINVAL;
    CK(cudaSetDevice(ctx->device));
    if (int rc = ensure_scratch(ctx, 8 * sizeof(double))) return rc;
    double *dout = static_cast<double *>(ctx->scratch);
    int *dstop = reinterpret_cast<int *>(dout + 7);
    pair_update_kernel<<<1, 1, 0, ctx->stream>>>(y_high, y_low, alpha_high, alpha_low, b_high, b_low, K12, C, eps, dout, dstop);
    ctx->launches += 1;
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(out7, dout, 7 * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaMemcpyAsync(stop, dstop, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return SVMB200_OK;
}

int svmb200_get_x_scaled(svmb200_ctx *ctx, float *Xs)
{
    if (!ctx || !Xs) return SVMB200_EINVAL;
    if (!ctx->have_train) FAIL(SVMB200_ESTATE, "no training set");
    CK(cudaSetDevice(ctx->device));
    if (ctx->x_u8) {        // uint8 storage: the scaled FP64 rows, rounded to fp32
        const size_t bytes = (size_t)ctx->n * ctx->dec_pitch4 * 16;
        if (int rc = ensure_work(ctx, 1, bytes)) return rc;
        const long long total = ctx->n * ctx->dec_pitch4;
        u8_gather_scale_kernel<<<(int)std::min<long long>((total + 255) / 256, (long long)ctx->num_sms * 16), 256, 0, ctx->stream>>>(
            U8View(ctx->X, ctx->n, ctx->u8_mw, ctx->u8_tail_pitch), ctx->u8_pos, ctx->minv, ctx->maxv, ctx->d, nullptr, ctx->n, ctx->dec_pitch4,
            static_cast<double *>(ctx->work[1]));
        CK(cudaGetLastError());
        std::vector<double> tmp((size_t)ctx->n * ctx->d);
        CK(cudaMemcpy2DAsync(tmp.data(), (size_t)ctx->d * sizeof(double), ctx->work[1], (size_t)ctx->dec_pitch4 * 16, (size_t)ctx->d * sizeof(double),
                             (size_t)ctx->n, cudaMemcpyDeviceToHost, ctx->stream));
        CK(cudaStreamSynchronize(ctx->stream));
        for (size_t k = 0; k < tmp.size(); ++k) Xs[k] = (float)tmp[k];
        return SVMB200_OK;
    }
    if (ctx->x_f64) {       // exact storage: hand back the rows rounded to fp32
        std::vector<double> tmp((size_t)ctx->n * ctx->d);
        CK(cudaMemcpy2D(tmp.data(), (size_t)ctx->d * sizeof(double), ctx->X, (size_t)ctx->pitch4 * 16, (size_t)ctx->d * sizeof(double),
                        (size_t)ctx->n, cudaMemcpyDeviceToHost));
        for (size_t k = 0; k < tmp.size(); ++k) Xs[k] = (float)tmp[k];
        return SVMB200_OK;
    }
    CK(cudaMemcpy2D(Xs, (size_t)ctx->d * sizeof(float), ctx->X, (size_t)ctx->pitch4 * 16, (size_t)ctx->d * sizeof(float),
                    (size_t)ctx->n, cudaMemcpyDeviceToHost));
    return SVMB200_OK;
}

} // extern "C"

#include "cascade.cuh"
