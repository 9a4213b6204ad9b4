// Drop-in for the reference's Image (src/neuralUtils/image.hh:12-43, image.cu:12-140): like Matrix but
// packed RGBA8 texels (a<<24 | b<<16 | g<<8 | r, image.cu:58), with PNG load (matcaps) and save (frames).
// The PNG codec is the library's own (nsdf_png_load / nsdf_png_save), not lodepng.
#pragma once
#include <memory>
#include <string>
#include "nsdf_host_common.hh"
#include "shape.hh"

class Image {
private:
    bool deviceAllocated = false;
    bool hostAllocated = false;
    bool hostOnly = false;
    void allocateDeviceMemory() {
        if (deviceAllocated || hostOnly) return;
        void* p = nullptr;
        nsdfCheck(nsdf_dev_malloc(&p, shape.x * shape.y * sizeof(uint)));
        deviceData = std::shared_ptr<uint>(static_cast<uint*>(p), [](uint* q) { nsdf_dev_free(q); });
        deviceAllocated = true;
    }
    void allocateHostMemory() {
        if (hostAllocated) return;
        hostData = std::shared_ptr<uint>(new uint[shape.x * shape.y](), [](uint* q) { delete[] q; });
        hostAllocated = true;
    }

public:
    Shape shape;
    std::shared_ptr<uint> deviceData;
    std::shared_ptr<uint> hostData;

    Image(size_t x_dim = 1, size_t y_dim = 1, bool hostOnly = false) : hostOnly(hostOnly), shape(x_dim, y_dim) {}
    Image(Shape shape, bool hostOnly = false) : hostOnly(hostOnly), shape(shape) {}

    void allocateMemory() {
        allocateDeviceMemory();
        allocateHostMemory();
    }
    void maybeAllocateMemory(Shape s) {
        if (!deviceAllocated && !hostAllocated) {
            shape = s;
            allocateMemory();
        }
    }
    // image.cu:36-65: decode, pack, upload.  Returns false (after printing) when the file cannot be decoded.
    bool loadPNG(std::string filename) {
        int w = 0, h = 0;
        if (nsdf_png_load(filename.c_str(), nullptr, 0, &w, &h) != NSDF_OK || w <= 0 || h <= 0) {
            std::printf("decoder error: cannot read %s\n", filename.c_str());
            return false;
        }
        deviceAllocated = hostAllocated = false;
        deviceData.reset();
        hostData.reset();
        shape = Shape((size_t)w, (size_t)h);
        allocateMemory();
        if (nsdf_png_load(filename.c_str(), hostData.get(), (size_t)w * h, &w, &h) != NSDF_OK) return false;
        copyHostToDevice();
        return true;
    }
    // image.cu:67-110: doFlip reverses the whole RGBA byte stream (a 180 degree turn of the picture);
    // doMirror is accepted and ignored, exactly as in the reference (its branch is empty, :99-101).
    bool savePNG(std::string filename, bool doFlip = true, bool doMirror = true) {
        (void)doMirror;
        if (!hostAllocated) {
            std::printf("[ERROR] no data to save...\n");
            return false;
        }
        return nsdf_png_save(filename.c_str(), hostData.get(), (int)shape.x, (int)shape.y, doFlip ? 1 : 0) == NSDF_OK;
    }
    bool savePPM(std::string filename) { return hostAllocated && nsdf_ppm_save(filename.c_str(), hostData.get(), (int)shape.x, (int)shape.y) == NSDF_OK; }
    void copyHostToDevice() {
        if (deviceAllocated && hostAllocated) nsdfCheck(nsdf_memcpy_h2d(deviceData.get(), hostData.get(), shape.x * shape.y * sizeof(uint)));
        else std::printf("Cannot copy host data to not allocated memory on device.\n");
    }
    void copyDeviceToHost() {
        if (deviceAllocated && hostAllocated) nsdfCheck(nsdf_memcpy_d2h(hostData.get(), deviceData.get(), shape.x * shape.y * sizeof(uint)));
        else std::printf("Cannot copy device data to not allocated memory on host.\n");
    }
    int size() { return (int)(shape.x * shape.y); }
    uint& operator[](const int index) { return hostData.get()[index]; }
    const uint& operator[](const int index) const { return hostData.get()[index]; }
};
