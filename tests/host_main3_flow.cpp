// The reference's main() (code/main3.cpp:306-417), statement for statement, compiled against the svmb200
// host layer instead of the reference's own function bodies.  The dataset name is argv[1] instead of the
// hard-coded "mnist3" (main3.cpp:311); everything else is the reference's sequence.  Prints SV ids too.
#include <chrono>
#include <iomanip>
#include <iostream>
#include "svmb200_host.hpp"
using namespace std;
using namespace std::chrono;
using namespace svmb200;

int main(int argc, char **argv)
{
    string dataset = argc > 1 ? argv[1] : "mnist3";
    string train_file = dataset + "_train_data.csv";
    vector<double> x_train;
    vector<int> y_train;
    size_t n_features;
    read_CSV(train_file, x_train, y_train, n_features);
    size_t n = x_train.size() / n_features;
    cout << "n = " << n << endl;
    cout << "n_features = " << n_features << endl;
    if (n == 0) { cerr << "Error: No data read from file." << endl; return 1; }
    vector<double> min_val(n_features, 0.0);
    vector<double> max_val(n_features, 0.0);
    find_min_max(x_train, min_val, max_val, n, n_features);
    scale_features(x_train, min_val, max_val, n, n_features);
    double C = 10.0;
    vector<double> alpha(n, 0.0);
    double b;
    SMO_train(x_train, y_train, n_features, alpha, b, C);
    string test_file = dataset + "_test_data.csv";
    vector<double> x_test;
    vector<int> y_test;
    read_CSV(test_file, x_test, y_test, n_features);
    size_t m = x_test.size() / n_features;
    scale_features(x_test, min_val, max_val, m, n_features);
    double tol = 1e-8;
    vector<int> sv_idx = get_SV_indices(alpha, tol);
    int sv_count = (int)sv_idx.size();
    cout << "Final SV count = " << sv_count << "\n";
    int correct = predict(x_test, y_test, n_features);
    double accuracy = double(correct) / double(m);
    cout << "Test accuracy = " << std::fixed << std::setprecision(15) << accuracy << " (" << correct << "/" << m << ")\n";
    cout << "SV ids:";
    for (int i : sv_idx) cout << " " << i;
    cout << "\n";
    // warm restart as the Cascade stages do (mpi_svm_main2.cpp:507): init = false keeps alpha
    double b2;
    SMO_train(x_train, y_train, n_features, alpha, b2, C, false);
    cout << "warm restart b = " << std::fixed << std::setprecision(15) << b2 << endl;
    return 0;
}
