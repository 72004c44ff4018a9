// Replays the reference's driver loop (LinearRegression.scala:28-32,53-66,86) through the C++ host
// facade: model with zero weights, numIterations 50, stepSize 0.005, miniBatchFraction 1.0; per
// interval first the stats op (pre-training weights), then trainOn.  Batches come from binary files
// written by the test (n, d as int64, then X row-major, then labels); results go to stdout as JSON.
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

#include "twtml.hpp"

static bool read_batch(const char* path, int64_t& n, int64_t& d, std::vector<double>& X, std::vector<double>& y)
{
    FILE* f = std::fopen(path, "rb");
    if (!f) return false;
    bool ok = std::fread(&n, 8, 1, f) == 1 && std::fread(&d, 8, 1, f) == 1;
    if (ok) {
        X.resize((size_t)(n * d));
        y.resize((size_t)n);
        ok = (n == 0) || (std::fread(X.data(), 8, X.size(), f) == X.size() && std::fread(y.data(), 8, y.size(), f) == y.size());
    }
    std::fclose(f);
    return ok;
}

int main(int argc, char** argv)
{
    try {
        // "--devices 0,1,..": drive those GPUs through one group handle (twtml_create_all); default: GPU 0 alone
        std::vector<int> devices{0};
        if (argc > 2 && std::string(argv[1]) == "--devices") {
            devices.clear();
            for (const char* p = argv[2]; *p;) {
                devices.push_back((int)std::strtol(p, const_cast<char**>(&p), 10));
                if (*p == ',') ++p;
            }
            argv += 2;
            argc -= 2;
        }
        twtml::Context ctx(devices);
        twtml::StreamingLinearRegressionWithSGD model(ctx);
        try {
            model.latestModel();
            std::printf("{\"error\": \"expected IllegalArgumentException\"}\n");
            return 2;
        } catch (const twtml::IllegalArgumentException&) {
        }
        bool first = true;
        std::printf("{\"batches\": [");
        for (int a = 1; a < argc; ++a) {
            int64_t n, d;
            std::vector<double> X, y;
            if (!read_batch(argv[a], n, d, X, y)) return 3;
            if (first) model.setNumIterations(50).setStepSize(0.005).setMiniBatchFraction(1.0).setInitialWeights(std::vector<double>((size_t)d, 0.0));
            twtml::Batch rdd = twtml::Batch::fromDense(ctx, X.data(), d, y.data(), n, d);
            std::printf("%s{\"count\": %lld", first ? "" : ", ", (long long)rdd.count());
            if (!rdd.isEmpty()) {
                std::vector<double> pred;
                twtml_stats s = model.stats(rdd, nullptr, &pred);
                model.trainOn(rdd);
                std::printf(", \"mse\": %.17g, \"real_stdev\": %.17g, \"pred_stdev\": %.17g, \"iters\": %zu, \"pred0\": %.17g", s.mse,
                            s.real_stdev, s.pred_stdev, model.lossHistory.size(), pred.empty() ? 0.0 : pred[0]);
            }
            std::printf("}");
            first = false;
        }
        std::printf("], \"weights\": [");
        twtml::LinearRegressionModel lm = model.latestModel();
        for (size_t j = 0; j < lm.weights.size(); ++j) std::printf("%s%.17g", j ? ", " : "", lm.weights[j]);
        std::printf("]}\n");
        return 0;
    } catch (const std::exception& e) {
        std::printf("{\"error\": \"%s\"}\n", e.what());
        return 1;
    }
}
