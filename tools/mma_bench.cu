// Development microbenchmark: issue rate of tcgen05.mma.kind::f16 (A from TMEM, B from smem) for small N.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mma_bench mma_bench.cu ; ./mma_bench
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../cudaneuralrender_b200/csrc/mlp_fp32.cuh"
#include "../cudaneuralrender_b200/csrc/tc_common.cuh"
using namespace nsdf;

// ctas_per_sm CTAs of 128 threads; each allocates 64 TMEM columns and issues `iters` batches of `per` MMAs + commit + wait.
__device__ __forceinline__ void wait_poll(uint64_t* bar, uint32_t parity) {
    const uint32_t a = (uint32_t)__cvta_generic_to_shared(bar);
    uint32_t done;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done) : "r"(a), "r"(parity) : "memory");
    } while (!done);
}
__global__ void __launch_bounds__(128) bench(int N, int per, int iters, int lbo, long long* out, int poll) {
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ uint64_t bar;
    __shared__ uint32_t s_tmem;
    for (int i = threadIdx.x; i < 8192 / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
    if (threadIdx.x == 0) { mbar_init(&bar, 1); mbar_fence_init(); }
    __syncthreads();
    if (threadIdx.x < 32) tc::tmem_alloc(&s_tmem, 64);
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem = s_tmem;
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    long long t0 = 0, t1 = 0;
    if (threadIdx.x == 0) {
        const uint32_t idesc = tc::make_idesc(0, 128, N);
        const uint64_t d = tc::make_bdesc(smem_u32(smem), lbo);
        uint32_t phase = 0;
        t0 = clock64();
        for (int it = 0; it < iters; ++it) {
            for (int j = 0; j < per; ++j) tc::mma_ts(tmem, tmem + 32 + 8 * (j & 1), d, idesc, j ? 1u : 0u);
            tc::mma_commit(&bar);
            if (poll) wait_poll(&bar, phase); else tc::mbar_wait_guard(&bar, phase);
            phase ^= 1;
        }
        t1 = clock64();
        out[blockIdx.x] = t1 - t0;
    }
    __syncthreads();
    if (threadIdx.x < 32) tc::tmem_dealloc(tmem, 64);
}


// ONE CTA per SM, `chains` warps; warp w loops {`per` UMMAs into its own 32 accumulator columns, commit, wait} like a tile-layer of
// the dense kernel without any epilogue.  Shows whether the 75 clk per tile-layer of eight chains is the chains' latency (then more
// chains in flight lower it towards the tensor pipe's 3 x 17 clk) or a rate limit of the issue / commit path (then it stays).
__global__ void __launch_bounds__(512) bench_chains(int per, int iters, long long* out, int poll) {
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ uint64_t bar[16];
    __shared__ uint32_t s_tmem;
    for (int i = threadIdx.x; i < 8192 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
    if (threadIdx.x == 0) { for (int i = 0; i < 16; ++i) mbar_init(&bar[i], 1); mbar_fence_init(); }
    __syncthreads();
    if (threadIdx.x < 32) tc::tmem_alloc(&s_tmem, 512);
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem = s_tmem;
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    const int warp = threadIdx.x >> 5;
    const uint32_t idesc = tc::make_idesc(0, 128, 32);
    const uint64_t d = tc::make_bdesc(smem_u32(smem), 32 * 16);
    const uint32_t acc = tmem + 32u * (uint32_t)warp;      // chains 0..14: columns 0..479; the A operand: columns 480..511
    uint32_t phase = 0;
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
        if (tc::elect_one()) {
            for (int j = 0; j < per; ++j) tc::mma_ts(acc, tmem + 480 + 8 * (j & 1), d, idesc, j ? 1u : 0u);
            tc::mma_commit(&bar[warp]);
        }
        __syncwarp();
        if (poll) wait_poll(&bar[warp], phase); else tc::mbar_wait_guard(&bar[warp], phase);
        phase ^= 1;
    }
    const long long t1 = clock64();
    if ((threadIdx.x & 31) == 0) atomicMax((unsigned long long*)&out[blockIdx.x], (unsigned long long)(t1 - t0));
    __syncthreads();
    if (threadIdx.x < 32) tc::tmem_dealloc(tmem, 512);
}

static void run_chains(long long* d_out) {
    cudaFuncSetAttribute(bench_chains, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    for (int poll = 0; poll < 2; ++poll)
    for (int per : {3, 2}) {
        for (int chains : {1, 2, 4, 5, 6, 7, 8, 9, 10, 12, 15}) {
            const int iters = 512;
            cudaMemset(d_out, 0, 148 * 8);
            bench_chains<<<148, chains * 32, 24 * 1024>>>(per, iters, d_out, poll);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("chains: error %s\n", cudaGetErrorString(e)); return; }
            long long h[148];
            cudaMemcpy(h, d_out, sizeof(h), cudaMemcpyDeviceToHost);
            double avg = 0;
            for (int i = 0; i < 148; ++i) avg += h[i];
            avg /= 148;
            printf("%s chains/SM %2d, %d UMMAs (M128 N32 K16) + commit + wait per chain step: %.1f clk per chain step, %.1f clk per tile-layer SM-wide\n",
                   poll ? "test_wait spin" : "try_wait      ", chains, per, avg / iters, avg / iters / chains);
        }
    }
}

int main() {
    long long* d_out;
    cudaMalloc(&d_out, 148 * 8 * 8);
    cudaFuncSetAttribute(bench, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
    const int Ns[] = {32};
    for (int poll = 0; poll < 1; ++poll)
    for (int ctas = 1; ctas <= 8; ctas *= 2)
        for (int N : Ns) {
            if (N > 32 && false) continue;
            for (int per : {1, 2, 3, 4, 8, 16}) {
                const int iters = 256 / per > 8 ? 256 / per : 8;
                const int smem = 24 * 1024;   // <= 8 CTAs/SM (TMEM: 8 x 64 columns)
                bench<<<148 * ctas, 128, smem>>>(N, per, iters, N * 16, d_out, poll);
                cudaError_t e = cudaDeviceSynchronize();
                if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
                long long h[148 * 8];
                cudaMemcpy(h, d_out, sizeof(long long) * 148 * ctas, cudaMemcpyDeviceToHost);
                double avg = 0;
                for (int i = 0; i < 148 * ctas; ++i) avg += h[i];
                avg /= 148 * ctas;
                printf("poll %d ctas/SM %d N %3d per-batch %2d : %.1f cycles/batch  %.1f cycles/MMA (SM-level %.1f cycles/MMA)\n", poll, ctas, N, per,
                       avg / iters, avg / iters / per, avg / iters / per / ctas);
            }
        }
    run_chains(d_out);
    return 0;
}
