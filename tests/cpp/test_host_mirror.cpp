// Drives the C++ host-side mirror (include/photon_ml_b200.hpp): reads a LibSVM file, packs it, checks the
// require() behaviour, and runs TRON on a logistic objective with the L2 mix-in.  Prints a few numbers that
// tests/test_cpp_host.py compares with the Python driver on the same data.
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <sstream>
#include <string>
#include <vector>

#include "photon_ml_b200.hpp"

using namespace photon_ml_b200;

int main(int argc, char **argv)
{
    if (argc < 4) {
        std::fprintf(stderr, "usage: %s file.libsvm d l2\n", argv[0]);
        return 2;
    }
    const int d_raw = std::atoi(argv[2]);
    const double l2 = std::atof(argv[3]);
    // LibSVM rows -> CSR, labels > 0 -> 1 else 0, one-based indices, intercept column appended
    // (client/io/deprecated/LibSVMInputDataFormat.scala:51-82)
    std::vector<int64_t> row_ptr{ 0 };
    std::vector<int32_t> col;
    std::vector<double> val, label;
    std::ifstream in(argv[1]);
    std::string line;
    while (std::getline(in, line)) {
        std::istringstream ss(line);
        std::string tok;
        if (!(ss >> tok)) continue;
        label.push_back(std::atof(tok.c_str()) > 0 ? 1.0 : 0.0);
        while (ss >> tok) {
            const size_t c = tok.find(':');
            col.push_back(std::atoi(tok.substr(0, c).c_str()) - 1);
            val.push_back(std::atof(tok.substr(c + 1).c_str()));
        }
        col.push_back(d_raw);
        val.push_back(1.0);
        row_ptr.push_back((int64_t)col.size());
    }
    const int64_t n = (int64_t)label.size();
    const int32_t d = d_raw + 1;

    Context ctx(0);
    DataSet data(ctx, n, d, row_ptr.data(), col.data(), val.data(), label.data());
    NormalizationContext none;

    // require(): coefficient size (ValueAndGradientAggregator.scala:141-143) -> IllegalArgumentException
    GLMLossFunction plain(PML_LOSS_LOGISTIC);
    bool threw = false;
    try {
        plain.calculate(data, Vector((size_t)d + 1, 0.0), none);
    } catch (const IllegalArgumentException &) {
        threw = true;
    }
    std::printf("require_size_throws %d\n", threw ? 1 : 0);

    L2RegularizationTwiceDiff<DistributedGLMLossFunction> objective(l2, PML_LOSS_LOGISTIC);
    const auto at_zero = objective.calculate(data, Vector((size_t)d, 0.0), none);
    std::printf("value_at_zero %.17g\n", at_zero.first);

    TRON tron(none);
    const auto result = tron.optimize(objective, Vector((size_t)d, 0.0), data);
    std::printf("iterations %d\n", tron.iterations);
    std::printf("reason %d\n", (int)tron.convergenceReason);
    std::printf("evaluations %d\n", tron.evaluations);
    std::printf("hv_evaluations %d\n", tron.hessianVectorEvaluations);
    std::printf("value %.17g\n", result.second);
    double checksum = 0.0;
    for (size_t j = 0; j < result.first.size(); ++j) checksum += (double)(j + 1) * result.first[j];
    std::printf("checksum %.17g\n", checksum);
    const Vector diag = objective.hessianDiagonal(data, result.first);
    std::printf("diag0 %.17g\n", diag[0]);
    return 0;
}
