// cascade.cuh -- Cascade SVM over the GPUs of one box, device resident (included by capi.cu: it uses the context's internals).
//
// Replaces the round loops of the reference's MPI programs: the modified two-layer ("star") cascade
// (code/mpi_svm_main2.cpp:439-769) and the classical binary-tree cascade (code/mpi_svm_main3.cpp:565-828).  One rank per
// GPU (a process or a thread), each with its own context.  What the reference ships between ranks as five blocking host
// messages per tuple (count; rows as FP64; y; alpha; id -- mpi_svm_main2.cpp:445-458,751-757) travels here GPU to GPU with
// NCCL: the rows as the uint8 bytes they arrived as (784 B instead of 6 272 B per row) and one 16-byte record
// {alpha, id, y} per row.
//   star: ncclAllGather of the counts, then of the padded row blocks and records (the rank-ordered MPI_Recv loop,
//         mpi_svm_main2.cpp:578-607); rank 0 solves the merged problem (:615) and ncclBroadcast returns the new global
//         support-vector set (:444-458);
//   tree: ncclSend / ncclRecv between the pairs of a level (mpi_svm_main3.cpp:689-716), ncclBroadcast of rank 0's set at
//         the start of a round (:574-613).
// A rank's partition is uploaded once and stays on the device as raw bytes; a stage's training set is assembled on the
// device (support-vector rows first, then the partition rows whose id is not among them: mpi_svm_main2.cpp:474-502),
// packed into the SMO kernel's uint8 layout with the GLOBAL statistics (min / max over all ranks' rows, which the
// reference computes on rank 0 over all data, mpi_svm_main2.cpp:404-419) and solved warm-started (init = false,
// :167-183).  Only ids, labels and alphas -- a few bytes per row -- are looked at on the host ("first id wins",
// convergence = the same id set twice, :646-660).
#pragma once
#include <dlfcn.h>
#include <set>

namespace svm {

// ---- NCCL, bound at run time: a process that already holds an NCCL (PyTorch's) shares it, and the library has no link
// dependency for the callers that never run a cascade
typedef struct ncclComm *ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
enum { ncclInt8 = 0, ncclUint8 = 1, ncclInt32 = 2, ncclFloat64 = 8 };
enum { ncclSum = 0, ncclMax = 2, ncclMin = 3 };
struct NcclApi {
    int (*GetUniqueId)(ncclUniqueId *);
    int (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int);
    int (*CommDestroy)(ncclComm_t);
    const char *(*GetErrorString)(int);
    int (*AllGather)(const void *, void *, size_t, int, ncclComm_t, cudaStream_t);
    int (*AllReduce)(const void *, void *, size_t, int, int, ncclComm_t, cudaStream_t);
    int (*Broadcast)(const void *, void *, size_t, int, int, ncclComm_t, cudaStream_t);
    int (*Send)(const void *, size_t, int, int, ncclComm_t, cudaStream_t);
    int (*Recv)(void *, size_t, int, int, ncclComm_t, cudaStream_t);
    int (*GroupStart)();
    int (*GroupEnd)();
    bool ok = false;
};
static NcclApi &nccl_api(std::string *err)
{
    static NcclApi api;
    static bool tried = false;
    if (tried) return api;
    tried = true;
    void *h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) { if (err) *err = std::string("cannot load libnccl.so.2: ") + dlerror(); return api; }
    bool all = true;
    auto sym = [&](const char *name) { void *p = dlsym(h, name); if (!p) all = false; return p; };
    api.GetUniqueId = (decltype(api.GetUniqueId))sym("ncclGetUniqueId");
    api.CommInitRank = (decltype(api.CommInitRank))sym("ncclCommInitRank");
    api.CommDestroy = (decltype(api.CommDestroy))sym("ncclCommDestroy");
    api.GetErrorString = (decltype(api.GetErrorString))sym("ncclGetErrorString");
    api.AllGather = (decltype(api.AllGather))sym("ncclAllGather");
    api.AllReduce = (decltype(api.AllReduce))sym("ncclAllReduce");
    api.Broadcast = (decltype(api.Broadcast))sym("ncclBroadcast");
    api.Send = (decltype(api.Send))sym("ncclSend");
    api.Recv = (decltype(api.Recv))sym("ncclRecv");
    api.GroupStart = (decltype(api.GroupStart))sym("ncclGroupStart");
    api.GroupEnd = (decltype(api.GroupEnd))sym("ncclGroupEnd");
    api.ok = all;
    if (!all && err) *err = "libnccl.so.2 lacks a needed entry point";
    return api;
}

// one 16-byte record per row of a support-vector tuple
struct __align__(16) SvMeta { double alpha; int id; int y; };

// rows of uint8 on the device with their (alpha, id, y) on the host
struct RowSet {
    unsigned char *rows = nullptr;     // [count][d], device
    size_t cap_rows = 0;
    std::vector<SvMeta> meta;          // host
    int count() const { return (int)meta.size(); }
};

struct CascadeState {
    ncclComm_t comm = nullptr;
    int rank = 0, world = 1;
    int d = 0;
    long long n_total = 0;
    unsigned char *part_rows = nullptr;          // this rank's partition, raw bytes, resident for the whole run
    std::vector<SvMeta> part_meta;               // ids = first_id + i, labels, alpha = 0
    double *gmin = nullptr, *gmax = nullptr;     // global statistics (device)
    unsigned char *train_raw = nullptr;          // a stage's training set, raw bytes
    size_t train_cap = 0;
    int *sel = nullptr;                          // gather list
    size_t sel_cap = 0;
    unsigned char *xfer = nullptr;               // all-gather / receive area for rows
    size_t xfer_cap = 0;
    SvMeta *meta_dev = nullptr;                  // the same for the records
    size_t meta_cap = 0;
    int *count_dev = nullptr;                    // [world] counts
    RowSet model;                                // the final global support-vector set
    double model_b = 0.0;
    long long smo_iterations = 0;
    int stages = 0;
    double solve_ms = 0.0, smo_ms = 0.0;      // wall time of this rank's stages (upload of labels, packing, SMO, SV extraction); SMO kernels alone
};

// out[k] = src[sel[k]] (rows of d bytes, d a multiple of 4 or not: byte-wise in words where aligned)
__global__ void gather_raw_rows_kernel(const unsigned char *__restrict__ src, const int *__restrict__ sel, long long count, int d,
                                       unsigned char *__restrict__ out)
{
    const long long total = count * d;
    for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < total; g += (long long)gridDim.x * blockDim.x) {
        const long long k = g / d;
        const int c = (int)(g - k * d);
        out[g] = src[(size_t)sel[k] * d + c];
    }
}
__global__ void minmax_u8_to_f64_kernel(const unsigned char *__restrict__ X, long long n, int d, double *__restrict__ mn, double *__restrict__ mx)
{
    // one thread per column, rows strided over blockIdx.y partitions folded with atomics on the integer image of the doubles
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= d) return;
    int lo = 255, hi = 0;
    for (long long r = blockIdx.y; r < n; r += gridDim.y) {
        const int v = X[r * d + c];
        lo = min(lo, v);
        hi = max(hi, v);
    }
    // values are small non-negative integers: their FP64 images order like their integer bit patterns
    atomicMin(reinterpret_cast<unsigned long long *>(mn) + c, (unsigned long long)__double_as_longlong((double)lo));
    atomicMax(reinterpret_cast<unsigned long long *>(mx) + c, (unsigned long long)__double_as_longlong((double)hi));
}

} // namespace svm

struct svmb200_cascade_state : svm::CascadeState {};

#define NCK(call)                                                                                       \
    do {                                                                                                \
        int r_ = (call);                                                                                \
        if (r_ != 0) {                                                                                  \
            char buf_[512];                                                                             \
            snprintf(buf_, sizeof buf_, "%s failed at %s:%d: %s", #call, __FILE__, __LINE__, N.GetErrorString(r_)); \
            ctx->err = buf_;                                                                            \
            return SVMB200_ECUDA;                                                                       \
        }                                                                                               \
    } while (0)

static int cas_reserve(svmb200_ctx *ctx, void **p, size_t *cap, size_t bytes)
{
    if (bytes <= *cap) return 0;
    if (*p) CK(cudaFree(*p));
    *p = nullptr;
    *cap = 0;
    bytes += bytes / 4 + 4096;
    CK(cudaMalloc(p, bytes));
    *cap = bytes;
    return 0;
}
static int rowset_reserve(svmb200_ctx *ctx, svm::RowSet &s, size_t rows, int d)
{
    void *p = s.rows;
    size_t cap = s.cap_rows * d;
    if (int rc = cas_reserve(ctx, &p, &cap, std::max<size_t>(rows, 1) * d)) return rc;
    s.rows = static_cast<unsigned char *>(p);
    s.cap_rows = cap / d;
    return 0;
}

// training set = `first` (its alphas kept: the warm start) followed by the rows of `second` whose id is not among them
// (alpha 0), duplicates inside `second` dropped too: mpi_svm_main2.cpp:474-502 / mpi_svm_main3.cpp:629-655.
// The rows are assembled in cas->train_raw; `meta` receives the (alpha, id, y) of the assembled rows.
static int cas_build_train_set(svmb200_ctx *ctx, svm::CascadeState *cas, const svm::RowSet &first, const unsigned char *second_rows,
                               const std::vector<svm::SvMeta> &second_meta, std::vector<svm::SvMeta> &meta)
{
    const int d = cas->d;
    meta.assign(first.meta.begin(), first.meta.end());
    std::set<int> seen;
    for (const auto &m : first.meta) seen.insert(m.id);
    std::vector<int> keep;
    for (size_t i = 0; i < second_meta.size(); ++i)
        if (seen.insert(second_meta[i].id).second) {
            keep.push_back((int)i);
            svm::SvMeta m = second_meta[i];
            m.alpha = 0.0;
            meta.push_back(m);
        }
    const size_t n = meta.size();
    void *p = cas->train_raw;
    if (int rc = cas_reserve(ctx, &p, &cas->train_cap, std::max<size_t>(n, 1) * d)) return rc;
    cas->train_raw = static_cast<unsigned char *>(p);
    if (first.count())
        CK(cudaMemcpyAsync(cas->train_raw, first.rows, (size_t)first.count() * d, cudaMemcpyDeviceToDevice, ctx->stream));
    if (!keep.empty()) {
        void *sp = cas->sel;
        size_t cap = cas->sel_cap;
        if (int rc = cas_reserve(ctx, &sp, &cap, keep.size() * sizeof(int))) return rc;
        cas->sel = static_cast<int *>(sp);
        cas->sel_cap = cap;
        CK(cudaMemcpyAsync(cas->sel, keep.data(), keep.size() * sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
        const long long total = (long long)keep.size() * d;
        svm::gather_raw_rows_kernel<<<(int)std::min<long long>((total + 255) / 256, (long long)ctx->num_sms * 16), 256, 0, ctx->stream>>>(
            second_rows, cas->sel, (long long)keep.size(), d, cas->train_raw + (size_t)first.count() * d);
        ctx->launches += 1;
        CK(cudaGetLastError());
        CK(cudaStreamSynchronize(ctx->stream));        // `keep` is a host temporary
    }
    return 0;
}

// One SMO_train(..., init = false) on the rows in cas->train_raw (mpi_svm_main2.cpp:507,615 / mpi_svm_main3.cpp:660), then
// the extraction of its support vectors (alpha > sv_tol, in row order: :516-538) into `out`.
static int cas_solve_stage(svmb200_ctx *ctx, svm::CascadeState *cas, const std::vector<svm::SvMeta> &meta, const svmb200_params *params,
                           svm::RowSet &out, double *b_out)
{
    const int d = cas->d;
    const long long n = (long long)meta.size();
    out.meta.clear();
    *b_out = 0.0;
    if (n == 0) return 0;
    struct StageClock {
        double *acc;
        std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
        ~StageClock() { *acc += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count(); }
    } stage_clock{&cas->solve_ms};
    ctx->storage_resolved = SVMB200_STORAGE_U8;
    if (int rc = alloc_train(ctx, n, d)) return rc;
    CK(cudaMemcpyAsync(ctx->minv, cas->gmin, (size_t)d * sizeof(double), cudaMemcpyDeviceToDevice, ctx->stream));
    CK(cudaMemcpyAsync(ctx->maxv, cas->gmax, (size_t)d * sizeof(double), cudaMemcpyDeviceToDevice, ctx->stream));
    if (int rc = pack_u8_rows(ctx, cas->train_raw, n)) return rc;
    std::vector<int> y(n);
    std::vector<double> alpha(n);
    bool any = false;
    for (long long i = 0; i < n; ++i) { y[i] = meta[i].y; alpha[i] = meta[i].alpha; any = any || alpha[i] != 0.0; }
    CK(cudaMemcpyAsync(ctx->y, y.data(), (size_t)n * sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    ctx->have_stats = true;
    ctx->have_train = true;
    svmb200_params p = *params;
    p.warm_start = any ? 1 : 0;             // all alphas zero: f = -y either way (mpi_svm_main2.cpp:167-183 with an empty sum)
    double b = 0.0;
    svmb200_stats st;
    if (int rc = svmb200_smo_train(ctx, &p, alpha.data(), &b, &st)) return rc;
    cas->smo_iterations += st.num_iter - 1;
    cas->stages += 1;
    cas->smo_ms += st.smo_ms;
    *b_out = b;
    std::vector<int> idx;
    for (long long i = 0; i < n; ++i)
        if (alpha[i] > params->sv_tol) {
            idx.push_back((int)i);
            out.meta.push_back(svm::SvMeta{alpha[i], meta[i].id, meta[i].y});
        }
    if (int rc = rowset_reserve(ctx, out, idx.size(), d)) return rc;
    if (!idx.empty()) {
        void *sp = cas->sel;
        size_t cap = cas->sel_cap;
        if (int rc = cas_reserve(ctx, &sp, &cap, idx.size() * sizeof(int))) return rc;
        cas->sel = static_cast<int *>(sp);
        cas->sel_cap = cap;
        CK(cudaMemcpyAsync(cas->sel, idx.data(), idx.size() * sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
        const long long total = (long long)idx.size() * d;
        svm::gather_raw_rows_kernel<<<(int)std::min<long long>((total + 255) / 256, (long long)ctx->num_sms * 16), 256, 0, ctx->stream>>>(
            cas->train_raw, cas->sel, (long long)idx.size(), d, out.rows);
        ctx->launches += 1;
        CK(cudaGetLastError());
        CK(cudaStreamSynchronize(ctx->stream));
    }
    return 0;
}

// rank `root`'s tuple to every rank (count, rows, records): mpi_svm_main2.cpp:444-458 / mpi_svm_main3.cpp:574-613
static int cas_broadcast(svmb200_ctx *ctx, svm::CascadeState *cas, svm::RowSet &s, int root)
{
    svm::NcclApi &N = svm::nccl_api(nullptr);
    const int d = cas->d;
    int k = cas->rank == root ? s.count() : 0;
    CK(cudaMemcpyAsync(cas->count_dev, &k, sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
    NCK(N.Broadcast(cas->count_dev, cas->count_dev, 1, svm::ncclInt32, root, cas->comm, ctx->stream));
    CK(cudaMemcpyAsync(&k, cas->count_dev, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    if (cas->rank != root) s.meta.resize(k);
    if (k == 0) return 0;
    if (int rc = rowset_reserve(ctx, s, k, d)) return rc;
    void *mp = cas->meta_dev;
    size_t cap = cas->meta_cap;
    if (int rc = cas_reserve(ctx, &mp, &cap, (size_t)k * sizeof(svm::SvMeta))) return rc;
    cas->meta_dev = static_cast<svm::SvMeta *>(mp);
    cas->meta_cap = cap;
    if (cas->rank == root) CK(cudaMemcpyAsync(cas->meta_dev, s.meta.data(), (size_t)k * sizeof(svm::SvMeta), cudaMemcpyHostToDevice, ctx->stream));
    NCK(N.GroupStart());
    NCK(N.Broadcast(s.rows, s.rows, (size_t)k * d, svm::ncclUint8, root, cas->comm, ctx->stream));
    NCK(N.Broadcast(cas->meta_dev, cas->meta_dev, (size_t)k * sizeof(svm::SvMeta), svm::ncclUint8, root, cas->comm, ctx->stream));
    NCK(N.GroupEnd());
    if (cas->rank != root) CK(cudaMemcpyAsync(s.meta.data(), cas->meta_dev, (size_t)k * sizeof(svm::SvMeta), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return 0;
}

// point-to-point tuple transfer of the tree (mpi_svm_main3.cpp:689-716): exactly one of dst / src is >= 0
static int cas_sendrecv(svmb200_ctx *ctx, svm::CascadeState *cas, svm::RowSet &s, int dst, int src)
{
    svm::NcclApi &N = svm::nccl_api(nullptr);
    const int d = cas->d;
    void *mp = cas->meta_dev;
    size_t cap = cas->meta_cap;
    if (dst >= 0) {
        int k = s.count();
        CK(cudaMemcpyAsync(cas->count_dev, &k, sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
        NCK(N.Send(cas->count_dev, 1, svm::ncclInt32, dst, cas->comm, ctx->stream));
        if (k) {
            if (int rc = cas_reserve(ctx, &mp, &cap, (size_t)k * sizeof(svm::SvMeta))) return rc;
            cas->meta_dev = static_cast<svm::SvMeta *>(mp);
            cas->meta_cap = cap;
            CK(cudaMemcpyAsync(cas->meta_dev, s.meta.data(), (size_t)k * sizeof(svm::SvMeta), cudaMemcpyHostToDevice, ctx->stream));
            NCK(N.GroupStart());
            NCK(N.Send(s.rows, (size_t)k * d, svm::ncclUint8, dst, cas->comm, ctx->stream));
            NCK(N.Send(cas->meta_dev, (size_t)k * sizeof(svm::SvMeta), svm::ncclUint8, dst, cas->comm, ctx->stream));
            NCK(N.GroupEnd());
        }
        CK(cudaStreamSynchronize(ctx->stream));
        return 0;
    }
    int k = 0;
    NCK(N.Recv(cas->count_dev, 1, svm::ncclInt32, src, cas->comm, ctx->stream));
    CK(cudaMemcpyAsync(&k, cas->count_dev, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    s.meta.resize(k);
    if (k == 0) return 0;
    if (int rc = rowset_reserve(ctx, s, k, d)) return rc;
    if (int rc = cas_reserve(ctx, &mp, &cap, (size_t)k * sizeof(svm::SvMeta))) return rc;
    cas->meta_dev = static_cast<svm::SvMeta *>(mp);
    cas->meta_cap = cap;
    NCK(N.GroupStart());
    NCK(N.Recv(s.rows, (size_t)k * d, svm::ncclUint8, src, cas->comm, ctx->stream));
    NCK(N.Recv(cas->meta_dev, (size_t)k * sizeof(svm::SvMeta), svm::ncclUint8, src, cas->comm, ctx->stream));
    NCK(N.GroupEnd());
    CK(cudaMemcpyAsync(s.meta.data(), cas->meta_dev, (size_t)k * sizeof(svm::SvMeta), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return 0;
}

static bool same_id_set(const std::vector<svm::SvMeta> &a, const std::vector<svm::SvMeta> &b)
{
    if (a.size() != b.size()) return false;                    // mpi_svm_main2.cpp:646-660: size, then membership
    std::set<int> s;
    for (const auto &m : a) s.insert(m.id);
    for (const auto &m : b)
        if (!s.count(m.id)) return false;
    return true;
}

// modified two-layer cascade, mpi_svm_main2.cpp:439-769
static int cas_run_star(svmb200_ctx *ctx, svm::CascadeState *cas, const svmb200_params *params, int max_rounds, int *rounds, int *converged)
{
    svm::NcclApi &N = svm::nccl_api(nullptr);
    const int d = cas->d, W = cas->world;
    svm::RowSet global, local, merged_sv;            // global: SV_global of the reference; starts empty (:436-438)
    std::vector<svm::SvMeta> meta;
    double b = 0.0;
    int rnd = 0;
    bool conv = false;
    while (rnd < max_rounds && !conv) {
        // layer 1: every rank trains on SV_global (alpha kept) + its partition (:474-507), extracts its SVs (:516-538)
        if (int rc = cas_build_train_set(ctx, cas, global, cas->part_rows, cas->part_meta, meta)) return rc;
        double b_local = 0.0;
        if (int rc = cas_solve_stage(ctx, cas, meta, params, local, &b_local)) return rc;
        // gather: counts, then padded row blocks and records from every rank
        int k = local.count();
        std::vector<int> counts(W);
        CK(cudaMemcpyAsync(cas->count_dev + W, &k, sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
        NCK(N.AllGather(cas->count_dev + W, cas->count_dev, 1, svm::ncclInt32, cas->comm, ctx->stream));
        CK(cudaMemcpyAsync(counts.data(), cas->count_dev, (size_t)W * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
        CK(cudaStreamSynchronize(ctx->stream));
        const int cap = std::max(1, *std::max_element(counts.begin(), counts.end()));
        void *xp = cas->xfer, *mp = cas->meta_dev;
        size_t xcap = cas->xfer_cap, mcap = cas->meta_cap;
        if (int rc = cas_reserve(ctx, &xp, &xcap, (size_t)(W + 1) * cap * d)) return rc;
        if (int rc = cas_reserve(ctx, &mp, &mcap, (size_t)(W + 1) * cap * sizeof(svm::SvMeta))) return rc;
        cas->xfer = static_cast<unsigned char *>(xp); cas->xfer_cap = xcap;
        cas->meta_dev = static_cast<svm::SvMeta *>(mp); cas->meta_cap = mcap;
        unsigned char *send_rows = cas->xfer + (size_t)W * cap * d;
        svm::SvMeta *send_meta = cas->meta_dev + (size_t)W * cap;
        if (k) {
            CK(cudaMemcpyAsync(send_rows, local.rows, (size_t)k * d, cudaMemcpyDeviceToDevice, ctx->stream));
            CK(cudaMemcpyAsync(send_meta, local.meta.data(), (size_t)k * sizeof(svm::SvMeta), cudaMemcpyHostToDevice, ctx->stream));
        }
        NCK(N.GroupStart());
        NCK(N.AllGather(send_rows, cas->xfer, (size_t)cap * d, svm::ncclUint8, cas->comm, ctx->stream));
        NCK(N.AllGather(send_meta, cas->meta_dev, (size_t)cap * sizeof(svm::SvMeta), svm::ncclUint8, cas->comm, ctx->stream));
        NCK(N.GroupEnd());
        std::vector<svm::SvMeta> all((size_t)W * cap);
        CK(cudaMemcpyAsync(all.data(), cas->meta_dev, all.size() * sizeof(svm::SvMeta), cudaMemcpyDeviceToHost, ctx->stream));
        CK(cudaStreamSynchronize(ctx->stream));
        // merge on rank 0: its own SVs keep their alpha, the others join with alpha 0, first id wins (:552-607); layer 2 (:615)
        svm::RowSet next;
        if (cas->rank == 0) {
            std::vector<svm::SvMeta> mm;
            std::vector<int> pick;
            std::set<int> seen;
            for (int r = 0; r < W; ++r)
                for (int i = 0; i < counts[r]; ++i) {
                    svm::SvMeta m = all[(size_t)r * cap + i];
                    if (!seen.insert(m.id).second) continue;
                    if (r != 0) m.alpha = 0.0;
                    mm.push_back(m);
                    pick.push_back(r * cap + i);
                }
            void *p = cas->train_raw;
            if (int rc = cas_reserve(ctx, &p, &cas->train_cap, std::max<size_t>(mm.size(), 1) * d)) return rc;
            cas->train_raw = static_cast<unsigned char *>(p);
            if (!pick.empty()) {
                void *sp = cas->sel;
                size_t scap = cas->sel_cap;
                if (int rc = cas_reserve(ctx, &sp, &scap, pick.size() * sizeof(int))) return rc;
                cas->sel = static_cast<int *>(sp); cas->sel_cap = scap;
                CK(cudaMemcpyAsync(cas->sel, pick.data(), pick.size() * sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
                const long long total = (long long)pick.size() * d;
                svm::gather_raw_rows_kernel<<<(int)std::min<long long>((total + 255) / 256, (long long)ctx->num_sms * 16), 256, 0, ctx->stream>>>(
                    cas->xfer, cas->sel, (long long)pick.size(), d, cas->train_raw);
                ctx->launches += 1;
                CK(cudaGetLastError());
                CK(cudaStreamSynchronize(ctx->stream));
            }
            if (int rc = cas_solve_stage(ctx, cas, mm, params, next, &b)) return rc;
        }
        // the new global set goes to every rank (:444-458 of the next round); convergence = the same id set twice (:646-660)
        if (int rc = cas_broadcast(ctx, cas, next, 0)) return rc;
        conv = same_id_set(next.meta, global.meta);
        std::swap(global.rows, next.rows); std::swap(global.cap_rows, next.cap_rows); global.meta.swap(next.meta);
        if (next.rows) CK(cudaFree(next.rows));
        ++rnd;
    }
    // b travels with the model
    CK(cudaMemcpyAsync(cas->gmin + d, &b, sizeof(double), cudaMemcpyHostToDevice, ctx->stream));      // scratch double behind the statistics
    NCK(N.Broadcast(cas->gmin + d, cas->gmin + d, 1, svm::ncclFloat64, 0, cas->comm, ctx->stream));
    CK(cudaMemcpyAsync(&cas->model_b, cas->gmin + d, sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    if (cas->model.rows) CK(cudaFree(cas->model.rows));
    cas->model = global;
    if (local.rows) CK(cudaFree(local.rows));
    if (merged_sv.rows) CK(cudaFree(merged_sv.rows));
    *rounds = rnd;
    *converged = conv ? 1 : 0;
    return 0;
}

// classical binary-tree cascade, mpi_svm_main3.cpp:565-828 (the number of ranks is a power of two, :420-428)
static int cas_run_tree(svmb200_ctx *ctx, svm::CascadeState *cas, const svmb200_params *params, int max_rounds, int *rounds, int *converged)
{
    svm::NcclApi &N = svm::nccl_api(nullptr);
    const int d = cas->d, W = cas->world, rank = cas->rank;
    svm::RowSet mine, recv, cur;                     // mine: X_sv .. ID_sv of the reference after a round; recv: recv_*
    std::vector<svm::SvMeta> meta, prev_ids;         // prev_ids: rank 0's global_ID_sv (:560)
    double b = 0.0;
    int rnd = 0;
    bool conv = false;
    while (rnd < max_rounds) {
        ++rnd;
        // rank 0 feeds its last result back to everyone (:574-613); round 1: nothing
        recv.meta = rank == 0 ? mine.meta : std::vector<svm::SvMeta>();
        if (rank == 0 && mine.count()) {
            if (int rc = rowset_reserve(ctx, recv, mine.count(), d)) return rc;
            CK(cudaMemcpyAsync(recv.rows, mine.rows, (size_t)mine.count() * d, cudaMemcpyDeviceToDevice, ctx->stream));
        }
        if (int rc = cas_broadcast(ctx, cas, recv, 0)) return rc;
        const unsigned char *cur_rows = cas->part_rows;           // X_sv := own partition (:616-619)
        const std::vector<svm::SvMeta> *cur_meta = &cas->part_meta;
        for (int step = 1; step <= W; step *= 2) {
            if (rank % step == 0) {                                // :623-685
                if (int rc = cas_build_train_set(ctx, cas, recv, cur_rows, *cur_meta, meta)) return rc;
                if (int rc = cas_solve_stage(ctx, cas, meta, params, cur, &b)) return rc;
                cur_rows = cur.rows;
                cur_meta = &cur.meta;
            }
            if (step < W) {                                        // :687-717
                if (rank % (2 * step) == step) {
                    if (int rc = cas_sendrecv(ctx, cas, cur, rank - step, -1)) return rc;
                } else if (rank % (2 * step) == 0) {
                    if (int rc = cas_sendrecv(ctx, cas, recv, -1, rank + step)) return rc;
                }
            }
        }
        // mine := cur
        std::swap(mine.rows, cur.rows); std::swap(mine.cap_rows, cur.cap_rows); mine.meta.swap(cur.meta);
        int flag = 0;
        if (rank == 0) {                                           // :720-743
            flag = same_id_set(prev_ids, mine.meta) ? 1 : 0;
            prev_ids = mine.meta;
        }
        CK(cudaMemcpyAsync(cas->count_dev, &flag, sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
        NCK(N.Broadcast(cas->count_dev, cas->count_dev, 1, svm::ncclInt32, 0, cas->comm, ctx->stream));     // :823-825
        CK(cudaMemcpyAsync(&flag, cas->count_dev, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
        CK(cudaStreamSynchronize(ctx->stream));
        conv = flag != 0;
        if (conv) break;
    }
    // the final model lives on rank 0 (as in the reference); every rank gets it for prediction
    if (int rc = cas_broadcast(ctx, cas, mine, 0)) return rc;
    CK(cudaMemcpyAsync(cas->gmin + d, &b, sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
    NCK(N.Broadcast(cas->gmin + d, cas->gmin + d, 1, svm::ncclFloat64, 0, cas->comm, ctx->stream));
    CK(cudaMemcpyAsync(&cas->model_b, cas->gmin + d, sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    if (cas->model.rows) CK(cudaFree(cas->model.rows));
    cas->model = mine;
    if (recv.rows) CK(cudaFree(recv.rows));
    if (cur.rows) CK(cudaFree(cur.rows));
    *rounds = rnd;
    *converged = conv ? 1 : 0;
    return 0;
}

extern "C" {

int svmb200_cascade_unique_id(void *id128)
{
    if (!id128) return SVMB200_EINVAL;
    std::string err;
    svm::NcclApi &N = svm::nccl_api(&err);
    if (!N.ok) { g_create_error = err; return SVMB200_ECUDA; }
    svm::ncclUniqueId id;
    if (N.GetUniqueId(&id) != 0) { g_create_error = "ncclGetUniqueId failed"; return SVMB200_ECUDA; }
    memcpy(id128, &id, sizeof id);
    return SVMB200_OK;
}

int svmb200_cascade_connect(svmb200_ctx *ctx, int32_t rank, int32_t world, const void *id128)
{
    if (!ctx) return SVMB200_EINVAL;
    if (!id128 || world < 1 || rank < 0 || rank >= world) FAIL(SVMB200_EINVAL, "bad rank / world / id");
    if (ctx->cascade) FAIL(SVMB200_ESTATE, "svmb200_cascade_connect called twice");
    std::string err;
    svm::NcclApi &N = svm::nccl_api(&err);
    if (!N.ok) FAIL(SVMB200_ECUDA, "%s", err.c_str());
    CK(cudaSetDevice(ctx->device));
    svm::ncclUniqueId id;
    memcpy(&id, id128, sizeof id);
    svmb200_cascade_state *cas = new svmb200_cascade_state();
    cas->rank = rank;
    cas->world = world;
    ctx->cascade = cas;
    NCK(N.CommInitRank(&cas->comm, world, id, rank));
    CK(cudaMalloc(&cas->count_dev, (size_t)(2 * world + 2) * sizeof(int)));
    CK(cudaMemsetAsync(cas->count_dev, 0, (size_t)(2 * world + 2) * sizeof(int), ctx->stream));
    // NCCL builds a collective's rings and a pair's point-to-point channels on first use (hundreds of milliseconds): one tiny
    // call of each kind the round loops make, and one message along every edge of the tree, happen here, like MPI_Init, not
    // inside the first round
    NCK(N.AllGather(cas->count_dev + world, cas->count_dev, 1, svm::ncclInt32, cas->comm, ctx->stream));
    NCK(N.Broadcast(cas->count_dev, cas->count_dev, 1, svm::ncclInt32, 0, cas->comm, ctx->stream));
    NCK(N.AllReduce(cas->count_dev + world, cas->count_dev + world, 1, svm::ncclInt32, svm::ncclMax, cas->comm, ctx->stream));
    NCK(N.GroupStart());
    for (int step = 1; step < world; step *= 2) {          // mpi_svm_main3.cpp:689-716: rank r receives from r + step, r + step sends to r
        if (rank % (2 * step) == 0 && rank + step < world) NCK(N.Recv(cas->count_dev + world + 1, 1, svm::ncclInt32, rank + step, cas->comm, ctx->stream));
        else if (rank % (2 * step) == step) NCK(N.Send(cas->count_dev + world, 1, svm::ncclInt32, rank - step, cas->comm, ctx->stream));
    }
    NCK(N.GroupEnd());
    CK(cudaStreamSynchronize(ctx->stream));
    return SVMB200_OK;
}

int svmb200_cascade_set_partition_u8(svmb200_ctx *ctx, const uint8_t *X_part, const int32_t *y_part, int64_t n_part, int32_t d, int64_t first_id,
                                     int64_t n_total)
{
    if (!ctx) return SVMB200_EINVAL;
    svm::CascadeState *cas = ctx->cascade;
    if (!cas) FAIL(SVMB200_ESTATE, "svmb200_cascade_set_partition_u8 before svmb200_cascade_connect");
    if (n_part < 0 || d <= 0 || d > 832 || (n_part > 0 && (!X_part || !y_part)) || n_total <= 0 || n_total >= 0x7fffffffLL)
        FAIL(SVMB200_EINVAL, "bad partition (n_part=%lld d=%d n_total=%lld; d <= 832)", (long long)n_part, d, (long long)n_total);
    svm::NcclApi &N = svm::nccl_api(nullptr);
    CK(cudaSetDevice(ctx->device));
    if (dev_free(ctx, cas->part_rows) || dev_free(ctx, cas->gmin) || dev_free(ctx, cas->gmax)) return SVMB200_ECUDA;
    cas->d = d;
    cas->n_total = n_total;
    CK(cudaMalloc(&cas->part_rows, std::max<size_t>((size_t)n_part * d, 1)));
    CK(cudaMalloc(&cas->gmin, (size_t)(d + 2) * sizeof(double)));
    CK(cudaMalloc(&cas->gmax, (size_t)(d + 2) * sizeof(double)));
    cas->part_meta.resize(n_part);
    for (int64_t i = 0; i < n_part; ++i) cas->part_meta[i] = svm::SvMeta{0.0, (int)(first_id + i), y_part[i]};
    if (n_part) CK(cudaMemcpyAsync(cas->part_rows, X_part, (size_t)n_part * d, cudaMemcpyHostToDevice, ctx->stream));
    // global statistics: min / max over every rank's rows (the reference: rank 0 over all data, mpi_svm_main2.cpp:404-419)
    std::vector<double> init_lo(d, 255.0), init_hi(d, 0.0);
    CK(cudaMemcpyAsync(cas->gmin, init_lo.data(), (size_t)d * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaMemcpyAsync(cas->gmax, init_hi.data(), (size_t)d * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
    if (n_part) {
        svm::minmax_u8_to_f64_kernel<<<dim3((d + 127) / 128, (unsigned)std::min<int64_t>(n_part, 256)), 128, 0, ctx->stream>>>(
            cas->part_rows, n_part, d, cas->gmin, cas->gmax);
        ctx->launches += 1;
        CK(cudaGetLastError());
    }
    NCK(N.GroupStart());
    NCK(N.AllReduce(cas->gmin, cas->gmin, (size_t)d, svm::ncclFloat64, svm::ncclMin, cas->comm, ctx->stream));
    NCK(N.AllReduce(cas->gmax, cas->gmax, (size_t)d, svm::ncclFloat64, svm::ncclMax, cas->comm, ctx->stream));
    NCK(N.GroupEnd());
    CK(cudaStreamSynchronize(ctx->stream));
    return SVMB200_OK;
}

int svmb200_cascade_train(svmb200_ctx *ctx, int32_t mode, const svmb200_params *params, int32_t max_rounds, svmb200_cascade_stats *stats)
{
    if (!ctx) return SVMB200_EINVAL;
    svm::CascadeState *cas = ctx->cascade;
    if (!cas || !cas->part_rows) FAIL(SVMB200_ESTATE, "svmb200_cascade_train before svmb200_cascade_set_partition_u8");
    if (!params || max_rounds < 1) FAIL(SVMB200_EINVAL, "params must not be NULL and max_rounds must be >= 1");
    if (mode != SVMB200_CASCADE_STAR && mode != SVMB200_CASCADE_TREE) FAIL(SVMB200_EINVAL, "mode must be SVMB200_CASCADE_STAR or _TREE");
    if (mode == SVMB200_CASCADE_TREE && (cas->world & (cas->world - 1)))
        FAIL(SVMB200_EINVAL, "This cascade expects number of ranks to be a power of two.");      // mpi_svm_main3.cpp:425
    CK(cudaSetDevice(ctx->device));
    cas->smo_iterations = 0;
    cas->stages = 0;
    cas->solve_ms = cas->smo_ms = 0.0;
    int rounds = 0, converged = 0;
    const auto t0 = std::chrono::steady_clock::now();
    const int rc = mode == SVMB200_CASCADE_STAR ? cas_run_star(ctx, cas, params, max_rounds, &rounds, &converged)
                                                 : cas_run_tree(ctx, cas, params, max_rounds, &rounds, &converged);
    if (rc) return rc;
    // the final model becomes this context's model: its support vectors are the training set, svmb200_predict_u8 scales
    // test rows with the global statistics (mpi_svm_main2.cpp:705-734)
    const int nsv = cas->model.count();
    if (nsv > 0) {
        ctx->storage_resolved = SVMB200_STORAGE_U8;
        if (int r2 = alloc_train(ctx, nsv, cas->d)) return r2;
        CK(cudaMemcpyAsync(ctx->minv, cas->gmin, (size_t)cas->d * sizeof(double), cudaMemcpyDeviceToDevice, ctx->stream));
        CK(cudaMemcpyAsync(ctx->maxv, cas->gmax, (size_t)cas->d * sizeof(double), cudaMemcpyDeviceToDevice, ctx->stream));
        if (int r2 = pack_u8_rows(ctx, cas->model.rows, nsv)) return r2;
        std::vector<int> y(nsv);
        std::vector<double> alpha(nsv);
        for (int i = 0; i < nsv; ++i) { y[i] = cas->model.meta[i].y; alpha[i] = cas->model.meta[i].alpha; }
        CK(cudaMemcpyAsync(ctx->y, y.data(), (size_t)nsv * sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
        CK(cudaStreamSynchronize(ctx->stream));
        ctx->have_stats = true;
        ctx->have_train = true;
        ctx->params = *params;
        ctx->params.accum_chunk = DEFAULT_CHUNK;
        ctx->have_params = true;
        if (int r2 = svmb200_set_model(ctx, alpha.data(), cas->model_b)) return r2;
    }
    if (stats) {
        stats->rounds = rounds;
        stats->converged = converged;
        stats->n_sv = nsv;
        stats->b = cas->model_b;
        stats->train_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
        stats->smo_iterations = cas->smo_iterations;
        stats->stages = cas->stages;
        stats->solve_ms = cas->solve_ms;
        stats->smo_ms = cas->smo_ms;
    }
    return SVMB200_OK;
}

int svmb200_cascade_get_model(svmb200_ctx *ctx, int32_t *ids, double *alpha, int32_t cap, int32_t *count)
{
    if (!ctx) return SVMB200_EINVAL;
    svm::CascadeState *cas = ctx->cascade;
    if (!cas) FAIL(SVMB200_ESTATE, "no cascade");
    const int c = cas->model.count();
    if (count) *count = c;
    if (ids || alpha) {
        if (cap < c) FAIL(SVMB200_EINVAL, "capacity %d < %d support vectors", cap, c);
        for (int i = 0; i < c; ++i) {
            if (ids) ids[i] = cas->model.meta[i].id;
            if (alpha) alpha[i] = cas->model.meta[i].alpha;
        }
    }
    return SVMB200_OK;
}

} // extern "C"

static void cascade_destroy(svmb200_ctx *ctx)
{
    svmb200_cascade_state *cas = ctx->cascade;
    if (!cas) return;
    svm::NcclApi &N = svm::nccl_api(nullptr);
    if (cas->comm && N.ok) N.CommDestroy(cas->comm);
    cudaFree(cas->part_rows); cudaFree(cas->gmin); cudaFree(cas->gmax); cudaFree(cas->train_raw); cudaFree(cas->sel);
    cudaFree(cas->xfer); cudaFree(cas->meta_dev); cudaFree(cas->count_dev); cudaFree(cas->model.rows);
    delete cas;
    ctx->cascade = nullptr;
}
