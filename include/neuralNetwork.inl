// Inline implementation of include/neuralNetwork.hh (see there).
#pragma once
#include <cstdlib>
#include <iostream>
#include <highfive/H5DataSet.hpp>
#include <highfive/H5DataSpace.hpp>
#include <highfive/H5File.hpp>
#include "layers/denseLayer.hh"

inline NeuralNetwork::NeuralNetwork() {}

// neuralNetwork.cpp:36-42: a path that does not load is fatal.
inline NeuralNetwork::NeuralNetwork(std::string geomPath) {
    if (!load(geomPath)) {
        std::cout << "Failed to load model at " << geomPath << std::endl;
        std::exit(1);
    }
}

// neuralNetwork.cpp:44-48: the network owns its layers.
inline NeuralNetwork::~NeuralNetwork() {
    if (model) nsdf_model_destroy(model);
    for (Layer* l : layers) delete l;
}

inline void NeuralNetwork::addLayer(Layer* layer) { layers.push_back(layer); }
inline std::vector<Layer*> NeuralNetwork::getLayers() const { return layers; }
inline int NeuralNetwork::getNumWeightParams() const {
    int n = 0;
    for (Layer* l : layers) n += l->getNumWeightParams();
    return n;
}
inline int NeuralNetwork::getNumBiasParams() const {
    int n = 0;
    for (Layer* l : layers) n += l->getNumBiasParams();
    return n;
}

inline nsdf_model* NeuralNetwork::handle() {
    if (model && modelLayers == layers.size()) return model;
    if (model) { nsdf_model_destroy(model); model = nullptr; }
    std::vector<Matrix> keepW, keepB;
    std::vector<const float*> W, B;
    std::vector<int> in, out, act;
    for (Layer* l : layers) {
        DenseLayer* d = dynamic_cast<DenseLayer*>(l);
        if (!d) return nullptr;
        keepW.push_back(d->getWeightsMatrix());
        keepB.push_back(d->getBiasVector());
        W.push_back(keepW.back().hostData.get());
        B.push_back(keepB.back().hostData.get());
        in.push_back(d->getXDim());
        out.push_back(d->getYDim());
        act.push_back(d->getActivation() == ReLU ? NSDF_ACT_RELU : NSDF_ACT_TANH_LINEAR);
    }
    if (layers.empty()) return nullptr;
    nsdfCheck(nsdf_model_create(W.data(), B.data(), in.data(), out.data(), act.data(), (int)layers.size(), &model));
    modelLayers = layers.size();
    return model;
}

inline Matrix NeuralNetwork::forward(Matrix X, int maxBatchSize) {
    if (precision < 0) {
        const char* e = std::getenv("NSDF_PRECISION");
        precision = (e && *e) ? std::atoi(e) : NSDF_PREC_FP32;
    }
    nsdf_model* m = handle();
    const int in0 = layers.empty() ? 0 : static_cast<DenseLayer*>(layers[0])->getXDim();
    // The render path of the reference hands forward() a Matrix whose shape.x is 12 (4 points of 3
    // coordinates, src/volumeRender_kernel.cu:591,661) and a maxBatchSize of 4*W*H: what matters is the
    // flat array, read as consecutive points of in0 floats.
    const size_t batch = X.shape.x == (size_t)in0 ? X.shape.y : (in0 ? X.shape.x * X.shape.y / (size_t)in0 : 0);
    if (m && nsdf_model_is_fused(m)) {
        size_t cols = maxBatchSize > 0 && (size_t)maxBatchSize > batch ? (size_t)maxBatchSize : batch;
        if (cols > yColumns) {
            Y = Matrix(Shape(1, cols));
            Y.allocateMemory();
            yColumns = cols;
        }
        nsdfCheck(nsdf_mlp_forward(m, X.deviceData.get(), Y.deviceData.get(), (int64_t)batch, precision, nullptr));
        Y.shape = Shape(1, batch);
        return Y;
    }
    Matrix Z = X;
    for (Layer* l : layers) Z = l->forward(Z, maxBatchSize);
    Y = Z;
    return Y;
}

// neuralNetwork.cpp:85-151.
inline bool NeuralNetwork::load(std::string fp, bool hostOnly) {
    HighFive::File file(fp, HighFive::File::ReadOnly);
    // listObjectNames() order up to ten layers; Keras' own layer_names order beyond
    std::vector<std::string> kerasLayers = file.kerasLayerOrder();
    int layerCount = 0;
    for (size_t i = 0; i < kerasLayers.size(); ++i) {
        const std::string& lname = kerasLayers[i];
        if (file.getObjectType(lname) != HighFive::ObjectType::Group) {
            std::cout << "Unsupported Layer\n";
            return false;
        }
        HighFive::Group group = file.getGroup(lname);
        if (group.getNumberObjects() != 1) {
            std::cout << "Unsupported Layer\n";
            return false;
        }
        group = group.getGroup(lname);
        std::vector<std::vector<float>> weights;
        std::vector<float> biases;
        for (const std::string& mat : group.listObjectNames()) {
            if (group.getObjectType(mat) != HighFive::ObjectType::Dataset) {
                std::cout << "Unsupported Layer\n";
                return false;
            }
            HighFive::DataSet dataset = group.getDataSet(mat);
            const std::vector<size_t> dim = dataset.getDimensions();
            if (dim.size() == 1) dataset.read(biases);
            else if (dim.size() == 2) dataset.read(weights);
            else {
                std::cout << "Unsupported layer, to many dims!\n";
                return false;
            }
        }
        const int activation = (i + 1 == kerasLayers.size()) ? Tanh : ReLU;   // :136-139
        addLayer(new DenseLayer(std::string("Dense_") + std::to_string(layerCount), weights, biases, activation, hostOnly));
        ++layerCount;
    }
    return true;
}
