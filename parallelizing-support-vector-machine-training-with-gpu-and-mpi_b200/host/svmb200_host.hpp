// svmb200_host.hpp -- the C++ host layer: the reference's own function names and argument meaning,
// implemented over the C ABI (include/svmb200.h), so that the reference's main() programs compile
// against it unchanged.  Nothing here computes on the CPU: every numeric step is a
// call into libsvmb200.so; what stays on the host is I/O (CSV / idx readers) and bookkeeping.
#pragma once
#include <cstddef>
#include <cstdint>
#include <string>
#include <vector>

#include "../../include/svmb200.h"

namespace svmb200 {

// ---- I/O --------------------------------------------------------------------------------------
// code/main3.cpp:13-54: first line discarded, last column = label, label != 1 -> -1, rows with fewer
// than two fields skipped.  positive_label generalises the literal 1.
void read_CSV(const std::string &filename, std::vector<double> &X, std::vector<int> &Y, size_t &n_features,
              int positive_label = 1);
// MNIST idx3-ubyte images + idx1-ubyte labels (the input format BASELINE.json names; the reference has
// no idx reader).  Returns false (and prints to cerr, as read_CSV does) when a file cannot be read.
bool read_idx(const std::string &images_path, const std::string &labels_path, std::vector<uint8_t> &X,
              std::vector<uint8_t> &digits, size_t &n_features);
// code/main3.cpp:49-52 applied to idx digits
std::vector<int> binary_labels(const std::vector<uint8_t> &digits, int positive_digit = 1);

// ---- the reference's solver interface --------------------------------------------------------------
// code/main3.cpp:57-71 / :74-89 on host vectors (run on the device in FP64, same values bit for bit)
void find_min_max(const std::vector<double> &X, std::vector<double> &min_val, std::vector<double> &max_val,
                  const size_t n_samples, const size_t n_features);
void scale_features(std::vector<double> &X, const std::vector<double> &min_val, const std::vector<double> &max_val,
                    const size_t n_samples, const size_t n_features);
// code/main3.cpp:162 (init = true) and code/mpi_svm_main2.cpp:154 (init = false: alpha in/out).
// Prints what the reference prints (main3.cpp:290-293) and the same cerr lines on abnormal stops.
void SMO_train(const std::vector<double> &x_train, const std::vector<int> &y_train, const size_t n_features,
               std::vector<double> &alpha, double &b, double C = 10.0, bool init = true);
// code/main3.cpp:297-304
std::vector<int> get_SV_indices(const std::vector<double> &alpha, double tol = 1e-8);
// the prediction loop of code/main3.cpp:391-405 (ge_zero = false) / mpi_svm_main2.cpp:721-733 (true) on
// already-scaled test rows, with the model of the last SMO_train; returns the number of correct rows
int predict(const std::vector<double> &x_test_scaled, const std::vector<int> &y_test, const size_t n_features,
            bool ge_zero = false, std::vector<double> *scores = nullptr);

// ---- beyond the literals: everything the reference hard-codes, as a struct ------------------------------
struct Options {
    svmb200_params params;
    int device = 0;
    bool quiet = false;
    Options() { svmb200_default_params(&params); params.cache_rows = -1; }
};
Options &options();                    // process-wide options used by the functions above
svmb200_ctx *context();                // the process-wide context (created on first use)
const svmb200_stats &last_stats();     // stats of the last SMO_train

} // namespace svmb200
