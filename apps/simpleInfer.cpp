// simpleInfer -- drop-in for the reference's inference self-check (src/simpleInfer.cpp).
// batchTest (:112-147): B copies of one point go through nn.forward on the GPU; all outputs must be equal.
// singleTest (:81-110): prints tanh(nn.forward((0.1, 0.2, 0.3))).
// Differences: the model path may be given as argv[1] (default "model.h5", as in the reference, :83,114);
// argv[2] = batch size (default 1 000 000), argv[3] = NSDF_PREC_* rung; timing uses a steady clock and
// reports a warm second run too (the reference's std::clock figure includes first-call allocations).
#include <chrono>
#include <cmath>
#include <iomanip>
#include <iostream>
#include "neuralNetwork.hh"

static bool batchTest(NeuralNetwork& nn, int batchSize, bool doVerify) {
    Matrix X(Shape(3, batchSize));
    X.allocateMemory();
    for (int i = 0; i < batchSize; ++i) { X[3 * i] = 0.f; X[3 * i + 1] = 0.f; X[3 * i + 2] = 0.f; }
    X.copyHostToDevice();
    Matrix Y;
    for (int run = 0; run < 2; ++run) {
        const auto t0 = std::chrono::steady_clock::now();
        Y = nn.forward(X);
        nsdfCheck(nsdf_device_synchronize());
        const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
        std::cout << (run ? "Warm: " : "Took: ") << ms << " ms for " << batchSize << " inferences" << std::endl;
    }
    if (!doVerify) return true;
    Y.allocateMemory();   // host side of the returned alias
    Y.copyDeviceToHost();
    for (int i = 0; i < batchSize; ++i)
        if (Y[i] != Y[0]) {
            std::cout << "FAILED: " << Y[i] << " != " << Y[0] << " at " << i << std::endl;
            return false;
        }
    std::cout << "All " << batchSize << " outputs equal: raw " << Y[0] << ", tanh " << std::tanh(Y[0]) << std::endl;
    return true;
}

static void singleTest(NeuralNetwork& nn) {
    Matrix X(Shape(3, 1));
    X.allocateMemory();
    X[0] = 0.1f; X[1] = 0.2f; X[2] = 0.3f;
    X.copyHostToDevice();
    Matrix Y = nn.forward(X);
    Y.allocateMemory();
    Y.copyDeviceToHost();
    std::cout << "sdf(0.1, 0.2, 0.3) = " << std::tanh(Y[0]) << " (raw " << Y[0] << ")" << std::endl;
}

int main(int argc, char** argv) {
    const std::string path = argc > 1 ? argv[1] : "model.h5";
    const int batch = argc > 2 ? atoi(argv[2]) : 1000000;
    NeuralNetwork nn;
    bool ok = false;
    try {
        ok = nn.load(path);
    } catch (const std::exception& e) {
        std::cerr << e.what() << std::endl;
    }
    if (!ok) { std::cerr << "Failed to load model at " << path << std::endl; return 1; }
    if (argc > 3) nn.setPrecision(atoi(argv[3]));
    std::cout << std::setprecision(9);
    singleTest(nn);
    return batchTest(nn, batch, true) ? 0 : 1;
}
