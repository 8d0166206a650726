// svm_train -- the reference's main() (code/main3.cpp:306-417) over svmb200: read, scale, SMO_train, SV count,
// predict, the same printed lines.  Inputs: MNIST idx files (uint8 goes to the GPU as is) or the reference's CSV.
//
//   svm_train --train-images F --train-labels F [--test-images F --test-labels F]
//   svm_train --train-csv F [--test-csv F]
//   options: --digit D (positive class, default 1: main3.cpp:49-52)  --C v  --gamma v  --eps v  --tau v
//            --max-iter k  --sv-tol v  --cache-rows r (-1 auto, 0 off)  --limit n (first n rows: gpu_svm_main4.cu:34-38)
//            --device i  --exact (FP64 rows: the reference's own pair sequence)  --tensor-cores (predict on tcgen05)
//            --save-model F
#include <chrono>
#include <cstring>
#include <iomanip>
#include <iostream>

#include "svmb200_host.hpp"

using namespace svmb200;
using namespace std::chrono;

static void usage()
{
    std::cerr << "usage: svm_train (--train-images F --train-labels F [--test-images F --test-labels F] | --train-csv F [--test-csv F])\n"
                 "       [--digit D] [--C v] [--gamma v] [--eps v] [--tau v] [--max-iter k] [--sv-tol v] [--cache-rows r] [--limit n] [--device i] [--exact] [--tensor-cores] [--save-model F]\n";
}

int main(int argc, char **argv)
{
    std::string tri, trl, tei, tel, trc, tec;
    int digit = 1;
    bool exact = false, tensor_cores = false;
    std::string model_path;
    long long limit = -1;
    Options &opt = options();
    for (int i = 1; i < argc; ++i) {
        auto need = [&](const char *name) -> const char * {
            if (i + 1 >= argc) { std::cerr << name << " needs a value\n"; usage(); std::exit(1); }
            return argv[++i];
        };
        const std::string a = argv[i];
        if (a == "--train-images") tri = need("--train-images");
        else if (a == "--train-labels") trl = need("--train-labels");
        else if (a == "--test-images") tei = need("--test-images");
        else if (a == "--test-labels") tel = need("--test-labels");
        else if (a == "--train-csv") trc = need("--train-csv");
        else if (a == "--test-csv") tec = need("--test-csv");
        else if (a == "--digit") digit = std::atoi(need("--digit"));
        else if (a == "--C") opt.params.C = std::atof(need("--C"));
        else if (a == "--gamma") opt.params.gamma = std::atof(need("--gamma"));
        else if (a == "--eps") opt.params.eps = std::atof(need("--eps"));
        else if (a == "--tau") opt.params.tau = std::atof(need("--tau"));
        else if (a == "--max-iter") opt.params.max_iter = std::atoi(need("--max-iter"));
        else if (a == "--sv-tol") opt.params.sv_tol = std::atof(need("--sv-tol"));
        else if (a == "--cache-rows") opt.params.cache_rows = std::atoi(need("--cache-rows"));
        else if (a == "--limit") limit = std::atoll(need("--limit"));
        else if (a == "--device") opt.device = std::atoi(need("--device"));
        else if (a == "--exact") exact = true;
        else if (a == "--tensor-cores") tensor_cores = true;
        else if (a == "--save-model") model_path = need("--save-model");
        else if (a == "--help" || a == "-h") { usage(); return 0; }
        else { std::cerr << "unknown option " << a << "\n"; usage(); return 1; }
    }
    const bool use_idx = !tri.empty();
    if (use_idx == !trc.empty() || (use_idx && trl.empty())) { usage(); return 1; }

    svmb200_ctx *ctx = context();
    svmb200_set_exact(ctx, exact ? 1 : 0);
    if (tensor_cores) svmb200_set_tensor_cores(ctx, 1, 0.05);
    size_t n_features = 0, n = 0;
    std::vector<uint8_t> x8, digits;
    std::vector<double> x_train;
    std::vector<int> y_train;
    if (use_idx) {
        if (!read_idx(tri, trl, x8, digits, n_features)) return 1;
        y_train = binary_labels(digits, digit);
    } else {
        read_CSV(trc, x_train, y_train, n_features, digit);
    }
    n = y_train.size();
    if (limit >= 0 && (size_t)limit < n) n = (size_t)limit;          // gpu_svm_main4.cu:34-38
    y_train.resize(n);
    std::cout << "n = " << n << std::endl;
    std::cout << "n_features = " << n_features << std::endl;
    if (n == 0) {
        std::cerr << "Error: No data read from file." << std::endl;
        return 1;
    }

    auto start = high_resolution_clock::now();                        // main3.cpp:335
    std::vector<double> min_val(n_features), max_val(n_features);
    int rc = use_idx ? svmb200_set_train_u8(ctx, x8.data(), y_train.data(), (int64_t)n, (int32_t)n_features, min_val.data(), max_val.data())
                     : svmb200_set_train_f64(ctx, x_train.data(), y_train.data(), (int64_t)n, (int32_t)n_features, min_val.data(), max_val.data());
    if (rc != SVMB200_OK) { std::cerr << svmb200_last_error(ctx) << std::endl; return 2; }

    std::vector<double> alpha(n, 0.0);
    double b = 0.0;
    svmb200_stats st;
    if (svmb200_smo_train(ctx, &opt.params, alpha.data(), &b, &st) != SVMB200_OK) { std::cerr << svmb200_last_error(ctx) << std::endl; return 2; }
    static const char *stops[] = {nullptr, "i_high or i_low not found; iteration stops", "warning: infeasible U and V; iteration stops",
                                  "warning: non positive eta; iteration stops", "Too many iterations; stopping"};
    if (st.stop_reason > 0) std::cerr << stops[st.stop_reason] << std::endl;
    std::cout << "number of iterations: " << st.num_iter << std::endl;                                   // main3.cpp:290-293
    std::cout << "b = " << std::fixed << std::setprecision(15) << b << std::endl;
    std::cout << "(b_high - b_low)/2*1e10 = " << std::fixed << std::setprecision(15) << (st.b_high - st.b_low) / 2 * 1e10 << std::endl;
    std::cout << "Final SV count = " << st.n_sv << "\n";                                                 // main3.cpp:370
    auto end1 = high_resolution_clock::now();
    if (!model_path.empty() && svmb200_save_model(ctx, model_path.c_str()) != SVMB200_OK) { std::cerr << svmb200_last_error(ctx) << std::endl; return 2; }

    if (!tei.empty() || !tec.empty()) {
        std::vector<uint8_t> t8, tdig;
        std::vector<double> x_test;
        std::vector<int> y_test;
        size_t nf = 0;
        int32_t correct = 0;
        if (!tei.empty()) {
            if (!read_idx(tei, tel, t8, tdig, nf)) return 1;
            y_test = binary_labels(tdig, digit);
            rc = svmb200_predict_u8(ctx, t8.data(), y_test.data(), (int64_t)y_test.size(), 0, &correct, nullptr);
        } else {
            read_CSV(tec, x_test, y_test, nf, digit);
            rc = svmb200_predict_f64(ctx, x_test.data(), y_test.data(), (int64_t)y_test.size(), 0, &correct, nullptr);
        }
        if (rc != SVMB200_OK) { std::cerr << svmb200_last_error(ctx) << std::endl; return 2; }
        const size_t m = y_test.size();
        std::cout << "Test accuracy = " << std::fixed << std::setprecision(15) << double(correct) / double(m) << " (" << correct << "/" << m << ")\n";
    }
    auto end2 = high_resolution_clock::now();
    std::cout << "Training time: " << duration_cast<milliseconds>(end1 - start).count() << " ms" << std::endl;   // main3.cpp:412-414
    std::cout << "Prediction time: " << duration_cast<milliseconds>(end2 - end1).count() << " ms" << std::endl;
    std::cout << "Total Runtime: " << duration_cast<milliseconds>(end2 - start).count() << " ms" << std::endl;
    std::cout << "SMO loop on device: " << std::setprecision(3) << st.smo_ms << " ms, " << st.x_passes << " X passes, " << st.cache_hits
              << " cached iterations" << std::endl;
    return 0;
}
