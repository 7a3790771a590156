// Drives the C++ host-side mirror (include/photon_ml_b200.hpp): reads a LibSVM file, packs it, checks the
// require() behaviour, and runs TRON on a logistic objective with the L2 mix-in.  Prints a few numbers that
// tests/test_cpp_host.py compares with the Python driver on the same data.
#include <algorithm>
#include <cmath>
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

    // Batch: the first 64 rows as 8 entities of 8 rows each, every entity in the full d-dimensional space; the sum of
    // the per-entity values / the concatenated gradients must be what the same rows give as 8 separate data sets
    {
        const int64_t E = 8, per = 8;
        std::vector<int64_t> ent_row_ptr;
        for (int64_t e = 0; e <= E; ++e) ent_row_ptr.push_back(e * per);
        std::vector<int32_t> ent_dim((size_t)E, d);
        Batch batch(ctx, ent_row_ptr, ent_dim, row_ptr.data(), col.data(), val.data(), label.data());
        Vector coefs((size_t)batch.totalDimension());
        for (size_t j = 0; j < coefs.size(); ++j) coefs[j] = 0.01 * (double)((j * 7) % 13) - 0.05;
        Vector grads;
        const Vector values = batch.calculate(PML_LOSS_LOGISTIC, coefs, &grads);
        double worst = 0.0;
        for (int64_t e = 0; e < E; ++e) {
            const int64_t lo = row_ptr[(size_t)(e * per)];
            std::vector<int64_t> rp;
            for (int64_t i = e * per; i <= (e + 1) * per; ++i) rp.push_back(row_ptr[(size_t)i] - lo);
            DataSet one(ctx, per, d, rp.data(), col.data() + lo, val.data() + lo, label.data() + e * per);
            const Vector c(coefs.begin() + batch.coefPtr()[(size_t)e], coefs.begin() + batch.coefPtr()[(size_t)e + 1]);
            const auto ref = plain.calculate(one, c, none);
            worst = std::max(worst, std::fabs(ref.first - values[(size_t)e]) / std::fabs(ref.first));
            for (int32_t j = 0; j < d; ++j)
                worst = std::max(worst, std::fabs(ref.second[(size_t)j] - grads[(size_t)(batch.coefPtr()[(size_t)e] + j)]));
        }
        std::printf("batch_worst_diff %.3g\n", worst);
        bool batch_threw = false;
        try {
            batch.calculate(PML_LOSS_LOGISTIC, Vector(3, 0.0), nullptr);
        } catch (const IllegalArgumentException &) {
            batch_threw = true;
        }
        std::printf("batch_require_throws %d\n", batch_threw ? 1 : 0);
    }
    // Peer with a single rank: the merge kernel degenerates to getVector; the handle exchange is the identity
    {
        Peer peer(ctx, 0, 1, d, [](const std::vector<char> &mine) { return mine; });
        std::printf("peer_created %d\n", peer.handle() != nullptr ? 1 : 0);
        // host vectors through the one-rank exchange = the plain host-buffer call on the same data
        Vector w(d, 0.05), g_peer, g_plain(d, 0.0);
        const double f_peer = peer.calculate(data, PML_LOSS_LOGISTIC, w, g_peer);
        double f_plain = 0.0;
        check(pml_value_and_gradient(data.handle(), PML_LOSS_LOGISTIC, w.data(), &f_plain, g_plain.data()));
        double worst = std::fabs(f_peer - f_plain);
        for (int j = 0; j < d; ++j) worst = std::fmax(worst, std::fabs(g_peer[j] - g_plain[j]));
        const auto sl = peer.slice();
        std::printf("peer_host_diff %.3e\n", worst);
        std::printf("peer_slice_ok %d\n", (sl.first == 0 && sl.second == d && !peer.timedOut()) ? 1 : 0);
    }
    return 0;
}
