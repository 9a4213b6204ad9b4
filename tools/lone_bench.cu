// Development microbenchmarks (not product):
//   1. latency of ONE dependent network evaluation per warp: mlp32_eval_warp<1> (weights loaded after the barrier) against
//      mlp32_eval_lone (next layer's weights prefetched into registers), with 1..4 such warps per SM sub-partition;
//   2. issue rate of FFMA against the packed fma.rn.f32x2 (SASS FFMA2) on independent accumulators.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -fmad=false -o tools/lone_bench tools/lone_bench.cu
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_runtime.h>
#include "../cudaneuralrender_b200/csrc/mlp_fp32.cuh"

using namespace nsdf;

template <int VARIANT>
__global__ void __launch_bounds__(512, 1) lone_kernel(const float* __restrict__ blob, uint32_t blob_bytes, int n_sq, int iters,
                                                      long long* __restrict__ cycles, float* __restrict__ sink) {
    extern __shared__ __align__(16) float s_w[];
    __shared__ uint64_t s_bar;
    stage_weights(s_w, blob, blob_bytes, &s_bar);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float* abuf = s_w + (blob_bytes >> 2) + warp * (4 * WARP_ACT_STRIDE);
    float px = 0.1f + 0.01f * warp, py = 0.2f, pz = 0.3f;
    __syncthreads();
    const long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
        float y;
        if (VARIANT == 0) {
            const float x[1][3] = {{px, py, pz}};
            float yy[1];
            mlp32_eval_warp<1>(s_w, n_sq, abuf, x, yy, lane);
            y = yy[0];
        } else {
            y = mlp32_eval_lone(s_w, n_sq, abuf, px, py, pz, lane);
        }
        px = fmaf(y, 1e-3f, px);   // the next point depends on this result, like a march step
    }
    const long long t1 = clock64();
    if (lane == 0) {
        cycles[blockIdx.x * (blockDim.x >> 5) + warp] = t1 - t0;
        sink[blockIdx.x * (blockDim.x >> 5) + warp] = px;
    }
}

__global__ void __launch_bounds__(256) ffma_kernel(int iters, float* sink, long long* cycles) {
    float a[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) a[i] = threadIdx.x * 1e-3f + i;
    const float w0 = 1.0001f, w1 = 0.9999f, c = 1e-7f;
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 32; i += 2) {
            a[i] = fmaf(a[i], w0, c);
            a[i + 1] = fmaf(a[i + 1], w1, c);
        }
    }
    const long long t1 = clock64();
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 32; ++i) s += a[i];
    sink[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

__device__ __forceinline__ void ffma2(float& d0, float& d1, float a0, float a1, float b0, float b1) {
    unsigned long long d, a, b;
    asm("mov.b64 %0, {%1, %2};" : "=l"(d) : "f"(d0), "f"(d1));
    asm("mov.b64 %0, {%1, %2};" : "=l"(a) : "f"(a0), "f"(a1));
    asm("mov.b64 %0, {%1, %2};" : "=l"(b) : "f"(b0), "f"(b1));
    asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(d) : "l"(a), "l"(b));
    asm("mov.b64 {%0, %1}, %2;" : "=f"(d0), "=f"(d1) : "l"(d));
}

__global__ void __launch_bounds__(256) ffma2_kernel(int iters, float* sink, long long* cycles) {
    unsigned long long a[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        const float lo = threadIdx.x * 1e-3f + 2 * i, hi = lo + 1.0f;
        asm("mov.b64 %0, {%1, %2};" : "=l"(a[i]) : "f"(lo), "f"(hi));
    }
    unsigned long long w, c;
    {
        const float w0 = 1.0001f, w1 = 0.9999f, cc = 1e-7f;
        asm("mov.b64 %0, {%1, %2};" : "=l"(w) : "f"(w0), "f"(w1));
        asm("mov.b64 %0, {%1, %2};" : "=l"(c) : "f"(cc), "f"(cc));
    }
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int i = 0; i < 16; ++i) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(a[i]) : "l"(w), "l"(c));
    }
    const long long t1 = clock64();
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        float lo, hi;
        asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(a[i]));
        s += lo + hi;
    }
    sink[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

int main() {
    const int n_sq = 7;
    const int floats = fused_blob_floats(n_sq);
    std::vector<float> h(floats);
    srand(1);
    for (auto& v : h) v = (rand() / (float)RAND_MAX - 0.5f) * 0.5f;
    float* d_blob;
    cudaMalloc(&d_blob, floats * 4);
    cudaMemcpy(d_blob, h.data(), floats * 4, cudaMemcpyHostToDevice);
    long long* d_cyc;
    float* d_sink;
    cudaMalloc(&d_cyc, 148 * 16 * 8);
    cudaMalloc(&d_sink, 148 * 256 * 4 * 4);
    const uint32_t blob_bytes = floats * 4;
    const size_t smem = blob_bytes + 16 * 4 * WARP_ACT_STRIDE * 4;
    cudaFuncSetAttribute(lone_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaFuncSetAttribute(lone_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    const int iters = 2000;
    for (int warps : {1, 4, 8, 16}) {
        for (int variant = 0; variant < 2; ++variant) {
            for (int rep = 0; rep < 2; ++rep) {
                if (variant == 0) lone_kernel<0><<<148, warps * 32, smem>>>(d_blob, blob_bytes, n_sq, iters, d_cyc, d_sink);
                else lone_kernel<1><<<148, warps * 32, smem>>>(d_blob, blob_bytes, n_sq, iters, d_cyc, d_sink);
                cudaDeviceSynchronize();
            }
            std::vector<long long> c(148 * warps);
            cudaMemcpy(c.data(), d_cyc, c.size() * 8, cudaMemcpyDeviceToHost);
            double sum = 0, mx = 0;
            for (auto v : c) { sum += v; if (v > mx) mx = v; }
            printf("{\"bench\": \"lone_eval\", \"variant\": \"%s\", \"warps_per_sm\": %d, \"clk_per_eval_mean\": %.1f, \"clk_per_eval_max\": %.1f}\n",
                   variant ? "eval_lone(prefetch)" : "eval_warp<1>", warps, sum / c.size() / iters, mx / iters);
        }
    }
    const int fi = 4096;
    for (int which = 0; which < 2; ++which) {
        cudaEvent_t e0, e1;
        cudaEventCreate(&e0); cudaEventCreate(&e1);
        for (int rep = 0; rep < 2; ++rep) {
            cudaEventRecord(e0);
            if (which == 0) ffma_kernel<<<148 * 8, 256>>>(fi, d_sink, d_cyc);
            else ffma2_kernel<<<148 * 8, 256>>>(fi, d_sink, d_cyc);
            cudaEventRecord(e1);
            cudaDeviceSynchronize();
        }
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        const double fma = 148.0 * 8 * 256 * 32.0 * fi;
        printf("{\"bench\": \"%s\", \"ms\": %.4f, \"tflops\": %.2f, \"err\": \"%s\"}\n", which ? "ffma2" : "ffma", ms, 2 * fma / (ms * 1e-3) / 1e12,
               cudaGetErrorString(cudaGetLastError()));
    }
    return 0;
}
