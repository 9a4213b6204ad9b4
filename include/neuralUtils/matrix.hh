// Drop-in for the reference's Matrix (src/neuralUtils/matrix.hh:9-37, matrix.cu:12-90): a host float
// array and a device float array, shared_ptr-owned (copies alias the storage), explicit copies, flat
// host operator[].  Element (x, y) of a Matrix(Shape(x_dim, y_dim)) is [y * shape.x + x]: for the network
// input Matrix(3, B) that is point b, coordinate c at [b*3 + c] (src/simpleInfer.cpp:116-124).
// Device memory goes through the C ABI (nsdf_dev_*), so this header needs no CUDA toolchain.
#pragma once
#include <memory>
#include "nsdf_host_common.hh"
#include "shape.hh"

class Matrix {
private:
    bool deviceAllocated = false;
    bool hostAllocated = false;
    bool hostOnly = false;
    void allocateDeviceMemory() {
        if (deviceAllocated || hostOnly) return;      // hostOnly skips the device side (matrix.cu:31-38)
        void* p = nullptr;
        nsdfCheck(nsdf_dev_malloc(&p, shape.x * shape.y * sizeof(float)));
        deviceData = std::shared_ptr<float>(static_cast<float*>(p), [](float* q) { nsdf_dev_free(q); });
        deviceAllocated = true;
    }
    void allocateHostMemory() {
        if (hostAllocated) return;
        hostData = std::shared_ptr<float>(new float[shape.x * shape.y](), [](float* q) { delete[] q; });
        hostAllocated = true;
    }

public:
    Shape shape;
    std::shared_ptr<float> deviceData;
    std::shared_ptr<float> hostData;

    Matrix(size_t x_dim = 1, size_t y_dim = 1, bool hostOnly = false) : hostOnly(hostOnly), shape(x_dim, y_dim) {}
    Matrix(Shape shape, bool hostOnly = false) : hostOnly(hostOnly), shape(shape) {}

    void allocateMemory() {
        allocateDeviceMemory();
        allocateHostMemory();
    }
    // Allocates once with the given shape and never re-allocates (matrix.cu:40-45).
    void maybeAllocateMemory(Shape s) {
        if (!deviceAllocated && !hostAllocated) {
            shape = s;
            allocateMemory();
        }
    }
    void copyHostToDevice() {
        if (deviceAllocated && hostAllocated) nsdfCheck(nsdf_memcpy_h2d(deviceData.get(), hostData.get(), shape.x * shape.y * sizeof(float)));
        else std::printf("Cannot copy host data to not allocated memory on device.\n");   // printf only, as matrix.cu:52-54
    }
    void copyDeviceToHost() {
        if (deviceAllocated && hostAllocated) nsdfCheck(nsdf_memcpy_d2h(hostData.get(), deviceData.get(), shape.x * shape.y * sizeof(float)));
        else std::printf("Cannot copy device data to not allocated memory on host.\n");
    }
    int size() { return (int)(shape.x * shape.y); }
    bool isDeviceAllocated() const { return deviceAllocated; }
    bool isHostOnly() const { return hostOnly; }
    float& operator[](const int index) { return hostData.get()[index]; }
    const float& operator[](const int index) const { return hostData.get()[index]; }
};
