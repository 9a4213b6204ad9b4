// Drop-in for the reference's DenseLayer (src/layers/denseLayer.hh:6-41, denseLayer.cu:180-278).
// W is stored [out][in] (W[y*in + x] = weights[x][y], denseLayer.cu:217-227), b is [out].  forward()
// is the per-layer operator: Z = act(W A + b) for every column (= point) of A, through
// nsdf_dense_layer_forward (same arithmetic as the CUTLASS SIMT kernels the reference instantiates).
// NeuralNetwork::forward does not go through here when the whole model fits the fused kernels.
#pragma once
#include <cassert>
#include <vector>
#include "layer.hh"

enum Activation {
    ReLU,
    Tanh   // selects the LINEAR epilogue, as in the reference (denseLayer.cu:93-113); callers apply tanh
};

class DenseLayer : public Layer {
private:
    Matrix W;
    Matrix b;
    Matrix Z;
    int activation;
    bool hostOnly = false;
    size_t zColumns = 0;   // columns Z was allocated with

public:
    DenseLayer(std::string name, std::vector<std::vector<float>> weights, std::vector<float> biases, int activation,
               bool hostOnly = false)
        : W(weights.size(), weights.empty() ? 0 : weights[0].size(), hostOnly), b(biases.size(), 1, hostOnly),
          activation(activation), hostOnly(hostOnly) {
        this->name = name;
        this->type = eDense;
        W.allocateMemory();
        b.allocateMemory();
        const size_t in = W.shape.x, out = W.shape.y;
        for (size_t x = 0; x < in; ++x)
            for (size_t y = 0; y < out; ++y) W[(int)(y * in + x)] = weights[x][y];
        for (size_t y = 0; y < biases.size(); ++y) b[(int)y] = biases[y];
        this->numWeightParams = (int)(in * out);
        this->numBiasParams = (int)biases.size();
        if (!hostOnly) {
            W.copyHostToDevice();
            b.copyHostToDevice();
        }
    }
    ~DenseLayer() {}

    Matrix& forward(Matrix& A, int maxBatchSize = -1) {
        const int batch = (int)A.shape.y;
        if (maxBatchSize < batch) maxBatchSize = batch;
        if (!zColumns) {                                                  // allocated once (matrix.cu:40-45)
            Z.maybeAllocateMemory(Shape(W.shape.y, (size_t)maxBatchSize));
            zColumns = (size_t)maxBatchSize;
        }
        if ((size_t)batch > zColumns) {
            std::printf("DenseLayer::forward: batch %d exceeds the %d columns Z was first allocated with\n", batch, (int)zColumns);
            std::exit(EXIT_FAILURE);
        }
        nsdfCheck(nsdf_dense_layer_forward(W.deviceData.get(), A.deviceData.get(), Z.deviceData.get(), b.deviceData.get(),
                                           (int)W.shape.y, 1, (int)W.shape.x, batch,
                                           activation == ReLU ? NSDF_ACT_RELU : NSDF_ACT_TANH_LINEAR, nullptr));
        Z.shape.y = (size_t)batch;
        return Z;
    }
    int getXDim() const { return (int)W.shape.x; }
    int getYDim() const { return (int)W.shape.y; }
    int getActivation() const { return activation; }
    Matrix getWeightsMatrix() const { return W; }
    Matrix getBiasVector() const { return b; }
};
