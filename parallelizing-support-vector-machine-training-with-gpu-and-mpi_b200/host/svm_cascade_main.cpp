// svm_cascade -- the reference's Cascade SVM programs over the GPUs of one box: mpi_svm2 (modified two-layer "star" cascade,
// code/mpi_svm_main2.cpp:300-786) and mpi_svm3 (classical binary-tree cascade, code/mpi_svm_main3.cpp:414-845).
// One rank per GPU, here one host thread per rank; each rank owns a context of the C ABI (include/svmb200.h) and the
// support-vector tuples travel GPU to GPU with NCCL inside svmb200_cascade_train.  What stays on the host is what the
// reference's rank 0 does before and after the rounds: read the files, deal the rows in contiguous chunks of ceil(n / P)
// (mpi_svm_main2.cpp:346-402), print the report lines (:681,700,733,777-779).
//
//   svm_cascade --mode star|tree --gpus P (--train-images F --train-labels F --test-images F --test-labels F |
//                                            --train-csv F --test-csv F)
//   options: --digit D  --C v  --gamma v  --tau v  --max-rounds k (50: mpi_svm_main2.cpp:428)  --limit n
#include <chrono>
#include <cstring>
#include <iomanip>
#include <iostream>
#include <thread>

#include "svmb200_host.hpp"

using namespace svmb200;
using namespace std::chrono;

int main(int argc, char **argv)
{
    std::string tri, trl, tei, tel, trc, tec, mode = "star";
    int digit = 1, gpus = 1, max_rounds = 50;
    long long limit = -1;
    svmb200_params params;
    svmb200_default_params(&params);
    params.cache_rows = -1;
    for (int i = 1; i < argc; ++i) {
        auto need = [&](const char *name) -> const char * {
            if (i + 1 >= argc) { std::cerr << name << " needs a value\n"; std::exit(1); }
            return argv[++i];
        };
        const std::string a = argv[i];
        if (a == "--train-images") tri = need("--train-images");
        else if (a == "--train-labels") trl = need("--train-labels");
        else if (a == "--test-images") tei = need("--test-images");
        else if (a == "--test-labels") tel = need("--test-labels");
        else if (a == "--train-csv") trc = need("--train-csv");
        else if (a == "--test-csv") tec = need("--test-csv");
        else if (a == "--mode") mode = need("--mode");
        else if (a == "--gpus") gpus = std::atoi(need("--gpus"));
        else if (a == "--digit") digit = std::atoi(need("--digit"));
        else if (a == "--C") params.C = std::atof(need("--C"));
        else if (a == "--gamma") params.gamma = std::atof(need("--gamma"));
        else if (a == "--tau") params.tau = std::atof(need("--tau"));
        else if (a == "--max-rounds") max_rounds = std::atoi(need("--max-rounds"));
        else if (a == "--limit") limit = std::atoll(need("--limit"));
        else { std::cerr << "unknown option " << a << "\n"; return 1; }
    }
    if (mode != "star" && mode != "tree") { std::cerr << "--mode must be star or tree\n"; return 1; }
    if (mode == "tree" && (gpus & (gpus - 1))) {                                 // mpi_svm_main3.cpp:420-428
        std::cerr << "This cascade expects number of ranks to be a power of two.\n";
        return 1;
    }

    // ---- rank 0 of the reference reads everything (mpi_svm_main2.cpp:316-340)
    std::vector<uint8_t> X, Xt, dg, dgt;
    std::vector<int> Y, Yt;
    size_t d = 0, dt = 0;
    if (!tri.empty()) {
        if (!read_idx(tri, trl, X, dg, d) || !read_idx(tei, tel, Xt, dgt, dt)) return 1;
        Y = binary_labels(dg, digit);
        Yt = binary_labels(dgt, digit);
    } else if (!trc.empty()) {
        std::vector<double> Xd, Xtd;                                             // the reference's CSV holds the same integers as text
        read_CSV(trc, Xd, Y, d, digit);
        read_CSV(tec, Xtd, Yt, dt, digit);
        X.assign(Xd.begin(), Xd.end());
        Xt.assign(Xtd.begin(), Xtd.end());
    } else {
        std::cerr << "give --train-images/--train-labels/--test-images/--test-labels or --train-csv/--test-csv\n";
        return 1;
    }
    if (d == 0 || d != dt || Y.empty()) { std::cerr << "No data loaded.\n"; return 1; }      // mpi_svm_main2.cpp:333-337
    size_t n = Y.size();
    if (limit > 0 && (size_t)limit < n) n = (size_t)limit;
    const size_t m = Yt.size();
    std::cout << "n = " << n << "\nn_features = " << d << "\n";

    const auto t_start = steady_clock::now();
    unsigned char id[128];
    if (svmb200_cascade_unique_id(id) != SVMB200_OK) { std::cerr << svmb200_last_error(nullptr) << "\n"; return 1; }
    std::vector<svmb200_cascade_stats> stats(gpus);
    std::vector<int> rc(gpus, 0), correct(gpus, 0);
    std::vector<std::string> errs(gpus);
    std::vector<std::thread> ranks;
    for (int r = 0; r < gpus; ++r)
        ranks.emplace_back([&, r] {
            svmb200_ctx *ctx = nullptr;
            auto fail = [&](int code) { rc[r] = code; errs[r] = svmb200_last_error(ctx); if (ctx) svmb200_destroy(ctx); };
            if (int e = svmb200_create(&ctx, r)) return fail(e);
            if (int e = svmb200_cascade_connect(ctx, r, gpus, id)) return fail(e);
            const size_t chunk = (n + gpus - 1) / gpus;                          // mpi_svm_main2.cpp:348
            const size_t start = std::min(n, r * chunk), end = std::min(n, (r + 1) * chunk);
            if (int e = svmb200_cascade_set_partition_u8(ctx, X.data() + start * d, Y.data() + start, (int64_t)(end - start), (int32_t)d,
                                                         (int64_t)start, (int64_t)n)) return fail(e);
            if (int e = svmb200_cascade_train(ctx, mode == "star" ? SVMB200_CASCADE_STAR : SVMB200_CASCADE_TREE, &params, max_rounds, &stats[r]))
                return fail(e);
            if (r == 0 && m > 0 && stats[r].n_sv > 0) {                          // rank 0 predicts with "s >= 0" (mpi_svm_main2.cpp:705-734)
                int32_t c = 0;
                if (int e = svmb200_predict_u8(ctx, Xt.data(), Yt.data(), (int64_t)m, 1, &c, nullptr)) return fail(e);
                correct[r] = c;
            }
            svmb200_destroy(ctx);
        });
    for (auto &t : ranks) t.join();
    for (int r = 0; r < gpus; ++r)
        if (rc[r]) { std::cerr << "[rank " << r << "] " << errs[r] << "\n"; return 1; }
    const auto t_end = steady_clock::now();

    const svmb200_cascade_stats &s0 = stats[0];
    std::cout << std::setprecision(15);
    if (s0.converged)      // the star program counts rounds from 0 (mpi_svm_main2.cpp:436,681,768), the tree program from 1 (mpi_svm_main3.cpp:562,568,750)
        std::cout << "[rank 0] Converged at round " << (mode == "star" ? s0.rounds - 1 : s0.rounds) << ", SV count = " << s0.n_sv << "\n";
    else
        std::cout << "[rank 0] Not converged after " << s0.rounds << " rounds, SV count = " << s0.n_sv << "\n";
    std::cout << "[rank 0] Final b = " << s0.b << "\n";
    if (m > 0 && s0.n_sv > 0)
        std::cout << "[rank 0] Test accuracy (final model) = " << (double)correct[0] / (double)m << " (" << correct[0] << "/" << m << ")\n";
    std::cout << "[rank 0] Final global SV count = " << s0.n_sv << "\n";
    std::cout << "[rank 0] Cascade finished in " << s0.rounds << " rounds\n";
    std::cout << "[rank 0] training time = " << s0.train_ms << " ms\n";
    std::cout << "[rank 0] elapsed time = " << duration<double, std::milli>(t_end - t_start).count() << " ms\n";
    return 0;
}
