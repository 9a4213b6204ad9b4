// =====================================================================================
// oracle/ref_driver.cu -- headless driver around the REFERENCE's own, unmodified CUDA code.
//
// TEST INFRASTRUCTURE ONLY (second parity witness + same-box CUDA baseline).
// oracle/build_ref.sh compiles the reference's translation units from where they lie under
// /root/reference (denseLayer.cu, matrix.cu, shape.cu, image.cu, volumeRender_kernel.cu,
// neuralNetwork.cpp, common/lodepng.cpp) together with this file into oracle/_ref/ref_driver.
// This file replaces only src/main.cpp (GLUT/OpenGL/Eigen: no graphics stack on the box) and calls
// the reference through its public interface: NeuralNetwork::load / forward, Image::loadPNG,
// copyStaticSettings, copyViewMatrices, render_kernel.
//
//   ref_driver mlp    MODEL.h5 POINTS.f32 OUT.f32 [repeat]
//         POINTS: n*3 floats.  nn.forward in chunks that avoid multiples of 65 536 (the reference's
//         GemmBatched grid.z bug).  OUT: n floats (pre-tanh).
//   ref_driver render MODEL.h5 W H MATCAP.png|- CAMERA.f32 FRAME OUT.u32 [repeat]
//         CAMERA: 12 floats c_invViewMatrix + 16 floats c_normalMatrix (src/main.cpp:207-222).
//         OUT: W*H packed texels as the reference leaves them in d_output.
// Timings (CUDA events around the reference calls, after one warm-up call) go to stdout as JSON.
// =====================================================================================
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include <cuda_runtime.h>
#include <helper_cuda.h>

#include "neuralNetwork.hh"
#include "layers/denseLayer.hh"
#include "neuralUtils/image.hh"

extern "C" void render_kernel(dim3 gridSize, dim3 blockSize, uint* d_output, uint imageW, uint imageH, uint numInputs,
                              NeuralNetwork& nn, Image matcap);
extern "C" void copyViewMatrices(float* invViewMatrix, size_t sizeofInvModelViewMat, float* normalMatrix,
                                 size_t sizeofNormalMatrix, int frameNumber, int numInputs);
extern "C" void copyStaticSettings(int colorType, int numInputs);

static std::vector<float> read_floats(const char* path) {
    FILE* f = std::fopen(path, "rb");
    if (!f) { std::fprintf(stderr, "cannot open %s\n", path); std::exit(2); }
    std::fseek(f, 0, SEEK_END);
    const long n = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    std::vector<float> v((size_t)n / 4);
    if (std::fread(v.data(), 4, v.size(), f) != v.size()) { std::fprintf(stderr, "short read %s\n", path); std::exit(2); }
    std::fclose(f);
    return v;
}
static void write_bytes(const char* path, const void* p, size_t bytes) {
    FILE* f = std::fopen(path, "wb");
    if (!f || std::fwrite(p, 1, bytes, f) != bytes) { std::fprintf(stderr, "cannot write %s\n", path); std::exit(2); }
    std::fclose(f);
}

static int run_mlp(int argc, char** argv) {
    if (argc < 5) return 1;
    const int repeat = argc > 5 ? atoi(argv[5]) : 1;
    NeuralNetwork nn;
    if (!nn.load(argv[2])) return 3;
    const std::vector<float> pts = read_floats(argv[3]);
    const size_t n = pts.size() / 3;
    std::vector<float> out(n);
    size_t chunk = 1000000;
    if (chunk > n) chunk = n;
    if (chunk % 65536 == 0 && chunk > 1) --chunk;
    // the first forward sizes every layer's Z once and for all (matrix.cu:40-45): give it the largest chunk
    const int maxBatch = (int)chunk;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    float best_ms = 1e30f, first_ms = 0.f;
    for (int rep = 0; rep < repeat + 1; ++rep) {      // rep 0 = warm-up (includes the first-call cudaMallocs)
        float total = 0.f;
        for (size_t base = 0; base < n; base += chunk) {
            size_t m = n - base < chunk ? n - base : chunk;
            if (m % 65536 == 0 && m > 1) {            // split a forbidden size into two legal ones
                --m;
            }
            Matrix X(Shape(3, m));
            X.allocateMemory();
            std::memcpy(X.hostData.get(), pts.data() + 3 * base, m * 12);
            X.copyHostToDevice();
            cudaEventRecord(e0);
            Matrix Y = nn.forward(X, maxBatch);
            cudaEventRecord(e1);
            checkCudaErrors(cudaEventSynchronize(e1));
            float ms = 0.f;
            cudaEventElapsedTime(&ms, e0, e1);
            total += ms;
            Y.copyDeviceToHost();
            std::memcpy(out.data() + base, Y.hostData.get(), m * 4);
            if (m != (n - base < chunk ? n - base : chunk)) {   // the one point left over by the split
                base += m;
                Matrix X1(Shape(3, 1));
                X1.allocateMemory();
                std::memcpy(X1.hostData.get(), pts.data() + 3 * base, 12);
                X1.copyHostToDevice();
                Matrix Y1 = nn.forward(X1, maxBatch);
                Y1.copyDeviceToHost();
                out[base] = Y1.hostData.get()[0];
                base += 1;
                base -= chunk;                        // the loop adds chunk back
            }
        }
        if (rep == 0) first_ms = total;
        else if (total < best_ms) best_ms = total;
    }
    write_bytes(argv[4], out.data(), n * 4);
    std::printf("{\"mode\": \"mlp\", \"n\": %zu, \"first_call_ms\": %.4f, \"best_ms\": %.4f, \"queries_per_s\": %.6g}\n", n, first_ms,
                repeat ? best_ms : first_ms, n / ((repeat ? best_ms : first_ms) * 1e-3));
    return 0;
}

static int run_render(int argc, char** argv) {
    if (argc < 10) return 1;
    const int W = atoi(argv[3]), H = atoi(argv[4]);
    const bool useMatcap = std::strcmp(argv[5], "-") != 0;
    std::vector<float> cam = read_floats(argv[6]);
    if (cam.size() < 28) return 2;
    const int frame = atoi(argv[7]);
    const int repeat = argc > 9 ? atoi(argv[9]) : 1;
    NeuralNetwork nn;
    if (!nn.load(argv[2])) return 3;
    Image matcap;
    if (useMatcap && !matcap.loadPNG(argv[5])) return 4;
    copyStaticSettings(useMatcap ? 1 : 0, 3);
    copyViewMatrices(cam.data(), 48, cam.data() + 12, 64, frame, 3);
    uint* d_output = nullptr;
    checkCudaErrors(cudaMalloc((void**)&d_output, (size_t)W * H * sizeof(uint)));
    const dim3 block(8, 8), grid((W + 7) / 8, (H + 7) / 8);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    float best_ms = 1e30f, first_ms = 0.f, sum_ms = 0.f;
    for (int rep = 0; rep < repeat + 1; ++rep) {
        checkCudaErrors(cudaMemset(d_output, 0, (size_t)W * H * sizeof(uint)));
        checkCudaErrors(cudaDeviceSynchronize());      // generateSingleImage: sync, timer, render_kernel, sync (main.cpp:414-431)
        cudaEventRecord(e0);
        render_kernel(grid, block, d_output, (uint)W, (uint)H, 3u, nn, matcap);
        cudaEventRecord(e1);
        checkCudaErrors(cudaDeviceSynchronize());
        float ms = 0.f;
        cudaEventElapsedTime(&ms, e0, e1);
        if (rep == 0) first_ms = ms;
        else {
            if (ms < best_ms) best_ms = ms;
            sum_ms += ms;
        }
    }
    std::vector<uint> out((size_t)W * H);
    checkCudaErrors(cudaMemcpy(out.data(), d_output, out.size() * 4, cudaMemcpyDeviceToHost));
    write_bytes(argv[8], out.data(), out.size() * 4);
    std::printf("{\"mode\": \"render\", \"W\": %d, \"H\": %d, \"first_call_ms\": %.4f, \"best_ms\": %.4f, \"mean_ms\": %.4f, \"frames\": %d}\n",
                W, H, first_ms, repeat ? best_ms : first_ms, repeat ? sum_ms / repeat : first_ms, repeat);
    return 0;
}

int main(int argc, char** argv) {
    if (argc >= 2 && std::strcmp(argv[1], "mlp") == 0) return run_mlp(argc, argv);
    if (argc >= 2 && std::strcmp(argv[1], "render") == 0) return run_render(argc, argv);
    std::fprintf(stderr, "usage: ref_driver mlp MODEL POINTS OUT [repeat] | ref_driver render MODEL W H MATCAP|- CAMERA FRAME OUT [repeat]\n");
    return 1;
}
