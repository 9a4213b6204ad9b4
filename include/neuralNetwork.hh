// Drop-in for the reference's NeuralNetwork (src/neuralNetwork.hh:9-30, neuralNetwork.cpp:36-151).
//
//   load(path)        walks the Keras save_weights .h5 like neuralNetwork.cpp:85-151 (name-sorted groups,
//                     /L/L/{bias:0,kernel:0}, last layer "Tanh" = linear, the others ReLU) through the
//                     HighFive-subset shim in include/highfive/ (own HDF5 reader, no libhdf5);
//   forward(X, max)   returns Matrix(1, B) of PRE-tanh values, device resident, valid until the next
//                     forward (the reference returns an alias of the last layer's Z, neuralNetwork.cpp:54-63).
//                     When the model is 3|4 -> 32 x k -> 1 the whole network runs as ONE fused kernel
//                     (nsdf_mlp_forward); other shapes go layer by layer through DenseLayer::forward.
//   setPrecision(p)   NSDF_PREC_* rung of the fused kernel (default fp32, bit-identical to the reference's
//                     CUTLASS SIMT order); also read once from the environment variable NSDF_PRECISION.
//
// Define NSDF_REFERENCE_NEURALNETWORK_CPP before including this header to get declarations only, so that the
// reference's own src/neuralNetwork.cpp can be compiled against these headers unmodified.
#pragma once
#include <string>
#include <vector>
#include "layers/layer.hh"

class NeuralNetwork {
private:
    std::vector<Layer*> layers;
    Matrix Y;
    // fused-path state (not in the reference)
    nsdf_model* model = nullptr;
    size_t modelLayers = 0;
    size_t yColumns = 0;
    int precision = -1;

public:
    NeuralNetwork();
    NeuralNetwork(std::string geomPath);
    ~NeuralNetwork();

    Matrix forward(Matrix X, int maxBatchSize = -1);
    void addLayer(Layer* layer);
    std::vector<Layer*> getLayers() const;
    int getNumWeightParams() const;
    int getNumBiasParams() const;
    bool load(std::string fp, bool hostOnly = false);

    // extensions
    void setPrecision(int nsdfPrecision) { precision = nsdfPrecision; }
    int getPrecision() const { return precision; }
    nsdf_model* handle();   // fused C-ABI model of the current layers (nullptr if they are not DenseLayers)
};

#ifndef NSDF_REFERENCE_NEURALNETWORK_CPP
#include "neuralNetwork.inl"
#endif
