// Drop-in check for the C++ host layer: the call sequence of the reference's serial program (read CSV, per-feature
// min/max, scale, SMO_train, support-vector indices, scale the test set with the train statistics, predict;
// code/main3.cpp:306-405) written against svmb200's reference-named functions.  argv[1] is the dataset prefix.
#include <cstdio>
#include <iostream>
#include "svmb200_host.hpp"

int main(int argc, char **argv)
{
    namespace sv = svmb200;
    const std::string prefix = argc > 1 ? argv[1] : "mnist3";
    std::vector<double> xtr, xte;
    std::vector<int> ytr, yte;
    size_t d = 0;
    sv::read_CSV(prefix + "_train_data.csv", xtr, ytr, d);
    const size_t n = d ? xtr.size() / d : 0;
    std::printf("n = %zu\nn_features = %zu\n", n, d);
    if (!n) return 1;

    std::vector<double> lo, hi;
    sv::find_min_max(xtr, lo, hi, n, d);
    sv::scale_features(xtr, lo, hi, n, d);

    std::vector<double> alpha;
    double b = 0.0;
    sv::SMO_train(xtr, ytr, d, alpha, b, 10.0);                 // prints iterations and b as the reference does
    const std::vector<int> svs = sv::get_SV_indices(alpha, 1e-8);
    std::printf("Final SV count = %zu\n", svs.size());

    sv::read_CSV(prefix + "_test_data.csv", xte, yte, d);
    const size_t m = xte.size() / d;
    sv::scale_features(xte, lo, hi, m, d);                       // train statistics, main3.cpp:355
    const int correct = sv::predict(xte, yte, d);
    std::printf("Test accuracy = %.15f (%d/%zu)\n", double(correct) / double(m), correct, m);
    std::printf("SV ids:");
    for (int i : svs) std::printf(" %d", i);
    std::printf("\n");

    double b2 = 0.0;                                             // the Cascade stages restart from alpha: init = false
    sv::SMO_train(xtr, ytr, d, alpha, b2, 10.0, false);
    std::printf("warm restart b = %.15f\n", b2);
    return 0;
}
