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

namespace nsdf {
// Prototype measured here and NOT adopted (1 warp per SM: 2319 clk per evaluation against 1821 for mlp32_eval_warp<1>; the 32
// prefetch loads are hoisted in front of the activation loads and delay the start of the chain more than they save).
// The serial tail of a frame is ONE ray's chain of dependent evaluations, so what counts here is the
// latency of a single evaluation, whose floor is the dependent FFMA chain 3 + n_sq*32 + 32 (4 clocks each).
// mlp32_eval_warp<1> pays ~140 clocks per layer on top of the 128-clock chain: its 32 weight loads are issued after the
// __syncwarp() that publishes the previous layer (a barrier orders memory operations, so the compiler cannot hoist them),
// then wait out the shared-memory latency before the chain can start.  Weights do not depend on activations: here lane o
// loads the column W[o][0..31] of layer l+1 into registers BEFORE layer l's chain (two register buffers, layers unrolled
// in pairs), so that after the barrier only the eight activation loads stand between one chain and the next; the last
// layer's weights are fetched the same way.  Same arithmetic, same bits (ascending-k fmaf chain from 0, + bias, relu).
// x is warp-uniform, the result is returned in every lane.  abuf: 2 * WARP_ACT_STRIDE floats of per-warp scratch.
__device__ __forceinline__ void lone_load_col(const float* __restrict__ Wt, int lane, float (&w)[32], float& bias) {
#pragma unroll
    for (int k = 0; k < 32; ++k) w[k] = Wt[k * 32 + lane];
    bias = Wt[1024 + lane];
}
__device__ __forceinline__ void lone_load_last(const float* __restrict__ LL, float (&w)[32], float& bias) {
#pragma unroll
    for (int kc = 0; kc < 8; ++kc) {
        const float4 v = *reinterpret_cast<const float4*>(LL + 4 * kc);
        w[4 * kc + 0] = v.x; w[4 * kc + 1] = v.y; w[4 * kc + 2] = v.z; w[4 * kc + 3] = v.w;
    }
    bias = LL[32];
}
__device__ __forceinline__ float lone_chain(const float (&w)[32], const float* __restrict__ act) {
    float acc = 0.0f;
#pragma unroll
    for (int kc = 0; kc < 8; ++kc) {
        const float4 av = *reinterpret_cast<const float4*>(act + 4 * kc);
        acc = fmaf(w[4 * kc + 0], av.x, acc);
        acc = fmaf(w[4 * kc + 1], av.y, acc);
        acc = fmaf(w[4 * kc + 2], av.z, acc);
        acc = fmaf(w[4 * kc + 3], av.w, acc);
    }
    return acc;
}
__device__ __forceinline__ float mlp32_eval_lone(const float* __restrict__ sw, int n_sq, float* __restrict__ abuf, float x0,
                                                 float x1, float x2, int lane, float t4 = 0.0f, bool in4 = false) {
    float* cur = abuf;
    float* nxt = abuf + WARP_ACT_STRIDE;
    float wA[32], wB[32], bA, bB;
    const float* H = sw + FUSED_L0_FLOATS;
    const float* LL = H + n_sq * FUSED_SQ_FLOATS;
    {
        const float w0 = sw[lane], w1 = sw[32 + lane], w2 = sw[64 + lane], bb = sw[128 + lane];
        if (n_sq > 0) lone_load_col(H, lane, wA, bA); else lone_load_last(LL, wA, bA);
        float c = fmaf(w2, x2, fmaf(w1, x1, fmaf(w0, x0, 0.0f)));
        if (in4) c = fmaf(sw[96 + lane], t4, c);
        cur[lane] = relu_det(c + bb);
    }
    __syncwarp();
    int l = 0;
#pragma unroll 1
    for (; l + 1 < n_sq; l += 2) {
        lone_load_col(H + (l + 1) * FUSED_SQ_FLOATS, lane, wB, bB);
        nxt[lane] = relu_det(lone_chain(wA, cur) + bA);
        __syncwarp();
        if (l + 2 < n_sq) lone_load_col(H + (l + 2) * FUSED_SQ_FLOATS, lane, wA, bA); else lone_load_last(LL, wA, bA);
        cur[lane] = relu_det(lone_chain(wB, nxt) + bB);
        __syncwarp();
    }
    float out;
    if (l < n_sq) {   // odd number of hidden layers: wA holds the last hidden one
        lone_load_last(LL, wB, bB);
        nxt[lane] = relu_det(lone_chain(wA, cur) + bA);
        __syncwarp();
        out = lone_chain(wB, nxt) + bB;
    } else {
        out = lone_chain(wA, cur) + bA;
    }
    __syncwarp();   // abuf may be rewritten by the caller's next evaluation
    return out;
}

}  // namespace nsdf

using namespace nsdf;

template <int VARIANT>
__global__ void __launch_bounds__(512, 1) lone_kernel(const float* __restrict__ blob, uint32_t blob_bytes, int n_sq, int iters,
                                                      long long* __restrict__ cycles, float* __restrict__ sink) {
    extern __shared__ __align__(16) float s_w[];
    __shared__ uint64_t s_bar;
    stage_weights(s_w, blob, blob_bytes, &s_bar);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float* abuf = s_w + (blob_bytes >> 2) + warp * (16 * WARP_ACT_STRIDE);
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
        } else if (VARIANT >= 2) {   // VARIANT = R independent rays marched together by one warp (lane = neuron)
            constexpr int R = VARIANT >= 2 ? VARIANT : 1;
            float x[R][3], yy[R];
#pragma unroll
            for (int r = 0; r < R; ++r) { x[r][0] = px + 0.01f * r; x[r][1] = py; x[r][2] = pz; }
            mlp32_eval_warp<R>(s_w, n_sq, abuf, x, yy, lane);
            y = yy[0];
#pragma unroll
            for (int r = 1; r < R; ++r) y += yy[r];
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
    const size_t smem = blob_bytes + 16 * 16 * WARP_ACT_STRIDE * 4;
    cudaFuncSetAttribute(lone_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaFuncSetAttribute(lone_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaFuncSetAttribute(lone_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaFuncSetAttribute(lone_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaFuncSetAttribute(lone_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    const int iters = 2000;
    for (int warps : {4, 8, 16}) {
        for (int variant : {0, 1, 2, 4, 8}) {
            for (int rep = 0; rep < 2; ++rep) {
                if (variant == 0) lone_kernel<0><<<148, warps * 32, smem>>>(d_blob, blob_bytes, n_sq, iters, d_cyc, d_sink);
                else if (variant == 1) lone_kernel<1><<<148, warps * 32, smem>>>(d_blob, blob_bytes, n_sq, iters, d_cyc, d_sink);
                else if (variant == 2) lone_kernel<2><<<148, warps * 32, smem>>>(d_blob, blob_bytes, n_sq, iters, d_cyc, d_sink);
                else if (variant == 4) lone_kernel<4><<<148, warps * 32, smem>>>(d_blob, blob_bytes, n_sq, iters, d_cyc, d_sink);
                else lone_kernel<8><<<148, warps * 32, smem>>>(d_blob, blob_bytes, n_sq, iters, d_cyc, d_sink);
                cudaDeviceSynchronize();
            }
            std::vector<long long> c(148 * warps);
            cudaMemcpy(c.data(), d_cyc, c.size() * 8, cudaMemcpyDeviceToHost);
            double sum = 0, mx = 0;
            for (auto v : c) { sum += v; if (v > mx) mx = v; }
            printf("{\"bench\": \"lone_eval\", \"variant\": \"%s\", \"warps_per_sm\": %d, \"clk_per_eval_mean\": %.1f, \"clk_per_eval_max\": %.1f}\n",
                   variant == 0 ? "eval_warp<1>" : variant == 1 ? "eval_lone(prefetch)" : variant == 2 ? "eval_warp<2>" : variant == 4 ? "eval_warp<4>" : "eval_warp<8>",
                   warps, sum / c.size() / iters, mx / iters);
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
