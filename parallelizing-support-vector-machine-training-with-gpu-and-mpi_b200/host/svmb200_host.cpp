// svmb200_host.cpp -- see svmb200_host.hpp.  Host-side I/O and the reference-named wrappers over the C ABI.
#include "svmb200_host.hpp"

#include <cstdlib>
#include <fstream>
#include <iomanip>
#include <iostream>
#include <sstream>

namespace svmb200 {

namespace {
svmb200_ctx *g_ctx = nullptr;
svmb200_stats g_stats{};
Options g_options;

[[noreturn]] void die(const char *what, svmb200_ctx *ctx)
{
    std::cerr << "svmb200: " << what << ": " << svmb200_last_error(ctx) << std::endl;
    std::exit(2);
}
uint32_t be32(const unsigned char *p) { return (uint32_t)p[0] << 24 | (uint32_t)p[1] << 16 | (uint32_t)p[2] << 8 | p[3]; }
} // namespace

Options &options() { return g_options; }
const svmb200_stats &last_stats() { return g_stats; }

svmb200_ctx *context()
{
    if (!g_ctx && svmb200_create(&g_ctx, g_options.device) != SVMB200_OK) die("no usable GPU (there is no CPU path)", nullptr);
    return g_ctx;
}

// code/main3.cpp:13-54
void read_CSV(const std::string &filename, std::vector<double> &X, std::vector<int> &Y, size_t &n_features, int positive_label)
{
    std::ifstream file(filename);
    if (!file.is_open()) {
        std::cerr << "Error opening file: " << filename << std::endl;
        return;
    }
    std::string line, cell;
    std::getline(file, line);   // header
    {
        std::stringstream ss(line);
        size_t fields = 0;
        while (std::getline(ss, cell, ',')) ++fields;
        n_features = fields ? fields - 1 : 0;
    }
    std::vector<std::string> fields;
    while (std::getline(file, line)) {
        std::stringstream ss(line);
        fields.clear();
        while (std::getline(ss, cell, ',')) fields.push_back(cell);
        if (fields.size() < 2) continue;
        for (size_t i = 0; i + 1 < fields.size(); ++i) X.push_back(std::stod(fields[i]));
        const int label = std::stoi(fields.back());
        Y.push_back(label == positive_label ? 1 : -1);
    }
}

bool read_idx(const std::string &images_path, const std::string &labels_path, std::vector<uint8_t> &X,
              std::vector<uint8_t> &digits, size_t &n_features)
{
    std::ifstream fi(images_path, std::ios::binary), fl(labels_path, std::ios::binary);
    if (!fi.is_open() || !fl.is_open()) {
        std::cerr << "Error opening file: " << (fi.is_open() ? labels_path : images_path) << std::endl;
        return false;
    }
    unsigned char h[16];
    fi.read(reinterpret_cast<char *>(h), 16);
    if (!fi || be32(h) != 0x00000803u) { std::cerr << images_path << ": not an idx3-ubyte file" << std::endl; return false; }
    const size_t n = be32(h + 4), rows = be32(h + 8), cols = be32(h + 12);
    unsigned char g[8];
    fl.read(reinterpret_cast<char *>(g), 8);
    if (!fl || be32(g) != 0x00000801u || be32(g + 4) != n) { std::cerr << labels_path << ": idx1-ubyte header does not match" << std::endl; return false; }
    n_features = rows * cols;
    X.resize(n * n_features);
    digits.resize(n);
    fi.read(reinterpret_cast<char *>(X.data()), (std::streamsize)X.size());
    fl.read(reinterpret_cast<char *>(digits.data()), (std::streamsize)n);
    if (!fi || !fl) { std::cerr << "idx files are truncated" << std::endl; return false; }
    return true;
}

std::vector<int> binary_labels(const std::vector<uint8_t> &digits, int positive_digit)
{
    std::vector<int> y(digits.size());
    for (size_t i = 0; i < digits.size(); ++i) y[i] = digits[i] == positive_digit ? 1 : -1;
    return y;
}

void find_min_max(const std::vector<double> &X, std::vector<double> &min_val, std::vector<double> &max_val,
                  const size_t n_samples, const size_t n_features)
{
    min_val.resize(n_features);
    max_val.resize(n_features);
    if (svmb200_minmax_f64(context(), X.data(), (int64_t)n_samples, (int32_t)n_features, min_val.data(), max_val.data()) != SVMB200_OK)
        die("find_min_max", g_ctx);
}

void scale_features(std::vector<double> &X, const std::vector<double> &min_val, const std::vector<double> &max_val,
                    const size_t n_samples, const size_t n_features)
{
    if (svmb200_scale_f64(context(), X.data(), min_val.data(), max_val.data(), (int64_t)n_samples, (int32_t)n_features) != SVMB200_OK)
        die("scale_features", g_ctx);
}

void SMO_train(const std::vector<double> &x_train, const std::vector<int> &y_train, const size_t n_features,
               std::vector<double> &alpha, double &b, double C, bool init)
{
    svmb200_ctx *ctx = context();
    const int64_t n = (int64_t)y_train.size();
    if (init) alpha.assign((size_t)n, 0.0);                       // main3.cpp:167
    if (svmb200_set_train_scaled_f64(ctx, x_train.data(), y_train.data(), n, (int32_t)n_features) != SVMB200_OK) die("SMO_train (upload)", ctx);
    svmb200_params p = g_options.params;
    p.C = C;
    p.warm_start = init ? 0 : 1;
    if (svmb200_smo_train(ctx, &p, alpha.data(), &b, &g_stats) != SVMB200_OK) die("SMO_train", ctx);
    switch (g_stats.stop_reason) {                                 // the reference's cerr lines, main3.cpp:207,248,255,285
    case SVMB200_STOP_NOT_FOUND: std::cerr << "i_high or i_low not found; iteration stops" << std::endl; break;
    case SVMB200_STOP_INFEASIBLE: std::cerr << "warning: infeasible U and V; iteration stops" << std::endl; break;
    case SVMB200_STOP_ETA: std::cerr << "warning: non positive eta; iteration stops" << std::endl; break;
    case SVMB200_STOP_MAX_ITER: std::cerr << "Too many iterations; stopping" << std::endl; break;
    default: break;
    }
    if (!g_options.quiet) {                                        // main3.cpp:290-293
        std::cout << "number of iterations: " << g_stats.num_iter << std::endl;
        std::cout << "b = " << std::fixed << std::setprecision(15) << b << std::endl;
        std::cout << "(b_high - b_low)/2*1e10 = " << std::fixed << std::setprecision(15)
                  << (g_stats.b_high - g_stats.b_low) / 2 * 1e10 << std::endl;
    }
}

std::vector<int> get_SV_indices(const std::vector<double> &alpha, double tol)
{
    std::vector<int> idx;
    for (size_t i = 0; i < alpha.size(); ++i)
        if (alpha[i] > tol) idx.push_back((int)i);
    return idx;
}

int predict(const std::vector<double> &x_test_scaled, const std::vector<int> &y_test, const size_t n_features, bool ge_zero,
            std::vector<double> *scores)
{
    const int64_t m = (int64_t)(x_test_scaled.size() / n_features);
    int32_t correct = 0;
    if (scores) scores->resize((size_t)m);
    if (svmb200_predict_scaled_f64(context(), x_test_scaled.data(), y_test.empty() ? nullptr : y_test.data(), m, ge_zero ? 1 : 0,
                                   &correct, scores ? scores->data() : nullptr) != SVMB200_OK)
        die("predict", g_ctx);
    return correct;
}

} // namespace svmb200
