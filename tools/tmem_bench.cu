// Development microbenchmark: tcgen05.ld / tcgen05.st throughput per SM (32x32b.x32, one 4 KB slab per warp instruction).
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tmem_bench tmem_bench.cu ; ./tmem_bench
#include <cstdio>
#include <cstdlib>
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

// The other load shapes of tcgen05.ld, each delivering 32 registers per thread (4 KB per warp instruction) like 32x32b.x32.
__device__ __forceinline__ void tmem_ld_16x64b_x32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile("tcgen05.ld.sync.aligned.16x64b.x32.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
                 : "r"(taddr)
                 : "memory");
}
__device__ __forceinline__ void tmem_ld_16x128b_x16(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile("tcgen05.ld.sync.aligned.16x128b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
                 : "r"(taddr)
                 : "memory");
}
__device__ __forceinline__ void tmem_ld_16x256b_x8(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile("tcgen05.ld.sync.aligned.16x256b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
                 : "r"(taddr)
                 : "memory");
}

// MODE 10..13: 32x32b.x32 / 16x64b.x32 / 16x128b.x16 / 16x256b.x8 with FOUR independent destination register sets (no WAW
// dependence between the loads in flight); MODE 5..7: the 16-lane shapes into one register set, like MODE 0.
template <int MODE>
__global__ void __launch_bounds__(1024) bench_shapes(int iters, long long* out, uint32_t* sink) {
    __shared__ uint32_t s_tmem;
    if (threadIdx.x < 32) tc::tmem_alloc(&s_tmem, 512);
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const int warp = threadIdx.x >> 5;
    const uint32_t t = s_tmem + ((uint32_t)((warp & 3) * 32) << 16) + (uint32_t)(warp >> 2) * 64u;
    uint32_t a[32], b[32], c[32], d[32];
    for (int i = 0; i < 32; ++i) a[i] = b[i] = c[i] = d[i] = threadIdx.x + i;
    uint32_t acc = 0;
    __syncthreads();
    const long long t0 = clock64();
    for (int it = 0; it < iters; ++it) {
        // a 16-lane shape with 32 registers spans 16 lanes x 64 columns: the warp's own 64-column slab, lower lanes (t) / upper lanes (u)
        const uint32_t u = t + (16u << 16);
        if (MODE == 5) { tmem_ld_16x64b_x32(t, a); tmem_ld_16x64b_x32(t, a); tmem_ld_16x64b_x32(t, a); tmem_ld_16x64b_x32(t, a); }
        if (MODE == 6) { tmem_ld_16x128b_x16(t, a); tmem_ld_16x128b_x16(t, a); tmem_ld_16x128b_x16(t, a); tmem_ld_16x128b_x16(t, a); }
        if (MODE == 7) { tmem_ld_16x256b_x8(t, a); tmem_ld_16x256b_x8(t, a); tmem_ld_16x256b_x8(t, a); tmem_ld_16x256b_x8(t, a); }
        if (MODE == 10) { tc::tmem_ld32(t, a); tc::tmem_ld32(t + 32, b); tc::tmem_ld32(t, c); tc::tmem_ld32(t + 32, d); }
        if (MODE == 11) { tmem_ld_16x64b_x32(t, a); tmem_ld_16x64b_x32(u, b); tmem_ld_16x64b_x32(t, c); tmem_ld_16x64b_x32(u, d); }
        if (MODE == 12) { tmem_ld_16x128b_x16(t, a); tmem_ld_16x128b_x16(u, b); tmem_ld_16x128b_x16(t, c); tmem_ld_16x128b_x16(u, d); }
        if (MODE == 13) { tmem_ld_16x256b_x8(t, a); tmem_ld_16x256b_x8(u, b); tmem_ld_16x256b_x8(t, c); tmem_ld_16x256b_x8(u, d); }
        if (MODE == 14) { tmem_ld_16x64b_x32(t, a); tmem_ld_16x64b_x32(t, b); tmem_ld_16x64b_x32(t, c); tmem_ld_16x64b_x32(t, d); }
        tc::wait_ld();
        acc += a[0] + a[31] + b[1] + b[30] + c[2] + c[29] + d[3] + d[28];      // static indices: the arrays stay in registers
    }
    const long long t1 = clock64();
    __syncthreads();
    if (threadIdx.x == 0) out[blockIdx.x] = t1 - t0;
    if (acc == 0x12345u) sink[0] = acc;
    if (threadIdx.x < 32) tc::tmem_dealloc(s_tmem, 512);
}

template <int MODE>
void run_shapes(const char* name, long long* d_out, uint32_t* d_sink) {
    for (int warps : {4, 8, 16, 32}) {
        const int iters = 2000, depth = 4;
        bench_shapes<MODE><<<148, warps * 32>>>(iters, d_out, d_sink);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("%s: error %s\n", name, cudaGetErrorString(e)); return; }
        long long h[148];
        cudaMemcpy(h, d_out, sizeof(h), cudaMemcpyDeviceToHost);
        double avg = 0;
        for (int i = 0; i < 148; ++i) avg += h[i];
        avg /= 148;
        printf("%-26s warps/SM %2d loads-in-flight %d : %.1f B/clk/SM  (%.1f clk per warp-op)\n", name, warps, depth,
               (double)iters * depth * warps * 4096 / avg, avg / iters / depth);
    }
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
        acc += v[0] + v[15];      // static indices: v stays in registers (a dynamic index sends it to local memory)
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

int main(int argc, char** argv) {
    // one mode per process (an illegal access in one shape must not take the others down): tmem_bench [mode]
    const int only = argc > 1 ? atoi(argv[1]) : -1;
    long long* d_out;
    uint32_t* d_sink;
    cudaMalloc(&d_out, 148 * 8);
    cudaMalloc(&d_sink, 4);
    if (only < 0 || only == 0) run<0>("ld.x32", 4096, d_out, d_sink);
    if (only < 0 || only == 2) run<2>("ld.x16", 2048, d_out, d_sink);
    if (only < 0 || only == 1) run<1>("st.x32", 4096, d_out, d_sink);
    if (only < 0 || only == 4) run<4>("ld.x16.pack16", 2048, d_out, d_sink);
    if (only == 5) run_shapes<5>("ld.16x64b.x32  one set", d_out, d_sink);
    if (only == 6) run_shapes<6>("ld.16x128b.x16 one set", d_out, d_sink);
    if (only == 7) run_shapes<7>("ld.16x256b.x8  one set", d_out, d_sink);
    if (only == 10) run_shapes<10>("ld.32x32b.x32  four sets", d_out, d_sink);
    if (only == 11) run_shapes<11>("ld.16x64b.x32  four sets, both lane halves", d_out, d_sink);
    if (only == 12) run_shapes<12>("ld.16x128b.x16 four sets, both lane halves", d_out, d_sink);
    if (only == 13) run_shapes<13>("ld.16x256b.x8  four sets, both lane halves", d_out, d_sink);
    if (only == 14) run_shapes<14>("ld.16x64b.x32  four sets, lower lanes", d_out, d_sink);
    return 0;
}
