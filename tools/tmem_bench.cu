// Development microbenchmark: tcgen05.ld / tcgen05.st throughput per SM (32x32b.x32, one 4 KB slab per warp instruction).
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tmem_bench tmem_bench.cu ; ./tmem_bench
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../cudaneuralrender_b200/csrc/mlp_fp32.cuh"
#include "../cudaneuralrender_b200/csrc/tc_common.cuh"
#include "../cudaneuralrender_b200/csrc/tc_mlp4.cuh"
using namespace nsdf;

__device__ __forceinline__ void tmem_ld16_pack(uint32_t taddr, uint32_t (&r)[16]) {   // 32 columns of 16-bit cells -> 16 registers
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.pack::16b.b32 "
        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr)
        : "memory");
}

template <int MODE>   // 4 = ld x16.pack::16b (32 columns -> 16 regs); 0 = ld x32, 1 = st x32, 2 = ld x16, 3 = ld x32 + st x16 (the epilogue's mix)
__global__ void __launch_bounds__(1024) bench(int iters, int depth, long long* out, uint32_t* sink) {
    __shared__ uint32_t s_tmem;
    if (threadIdx.x < 32) tc::tmem_alloc(&s_tmem, 512);
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const int warp = threadIdx.x >> 5;
    const uint32_t t = s_tmem + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)(warp >> 2) * 64u;
    uint32_t v[32];
    for (int i = 0; i < 32; ++i) v[i] = threadIdx.x + i;
    uint32_t acc = 0;
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
        for (int d = 0; d < depth; ++d) {
            if (MODE == 0) tc::tmem_ld32(t + 32 * (d & 1), v);
            if (MODE == 1) tc::tmem_st32(t + 32 * (d & 1), v);
            if (MODE == 2) tc::tmem_ld16(t + 16 * (d & 3), reinterpret_cast<uint32_t(&)[16]>(v));
            if (MODE == 4) tmem_ld16_pack(t + 32 * (d & 1), reinterpret_cast<uint32_t(&)[16]>(v));
            if (MODE == 3) { tc::tmem_ld32(t, v); tc::wait_ld(); tc::tmem_st16(t + 32, reinterpret_cast<uint32_t(&)[16]>(v)); }
        }
        if (MODE == 1 || MODE == 3) tc::wait_st();
        if (MODE != 1) tc::wait_ld();
        acc += v[it & 31];
    }
    const long long t1 = clock64();
    __syncthreads();
    if (threadIdx.x == 0) out[blockIdx.x] = t1 - t0;
    if (acc == 0x12345u) sink[0] = acc;
    if (threadIdx.x < 32) tc::tmem_dealloc(s_tmem, 512);
}

template <int MODE>
void run(const char* name, int bytes_per_op, long long* d_out, uint32_t* d_sink) {
    for (int warps : {4, 16}) {
        for (int depth : {4}) {
            const int iters = 2000;
            bench<MODE><<<148, warps * 32>>>(iters, depth, d_out, d_sink);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return; }
            long long h[148];
            cudaMemcpy(h, d_out, sizeof(h), cudaMemcpyDeviceToHost);
            double avg = 0;
            for (int i = 0; i < 148; ++i) avg += h[i];
            avg /= 148;
            const double bytes = (double)iters * depth * warps * bytes_per_op;
            printf("%-14s warps/SM %2d ops-in-flight %d : %.1f cycles/iter  %.1f B/clk/SM  (%.1f clk per warp-op)\n", name, warps, depth,
                   avg / iters, bytes / avg, avg / iters / depth);
        }
    }
}

int main() {
    long long* d_out;
    uint32_t* d_sink;
    cudaMalloc(&d_out, 148 * 8);
    cudaMalloc(&d_sink, 4);
    run<0>("ld.x32", 4096, d_out, d_sink);
    run<2>("ld.x16", 2048, d_out, d_sink);
    run<1>("st.x32", 4096, d_out, d_sink);
    run<4>("ld.x16.pack16", 2048, d_out, d_sink);
    return 0;
}
