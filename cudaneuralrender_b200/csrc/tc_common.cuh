// =====================================================================================
// tc_common.cuh -- tcgen05 / TMEM building blocks shared by the tensor-core kernels
// (mlp_tc.cu: dense queries; render_tc: the mixed-precision tracer).  sm_100a only.
//
// Operand conventions used everywhere in this library (one 128-query tile per CTA slot):
//   D  (accumulator, fp32)  TMEM lanes 0..127 = tile rows (queries), 32 columns = neurons
//   A  (activations)        TMEM, same lanes; 16-bit values packed two per 32-bit column
//                           (k even in the low half), so K = 32 occupies 16 columns
//   B  (layer weights)      shared memory, K-major no-swizzle canonical layout: 8x8 core
//                           matrices of 128 contiguous bytes (row n%8 at +16 B), 8-row groups
//                           SBO = 128 B apart, 8-element k-chunks LBO = 512 B apart.
//                           B[n][k] = W[out = n][in = k], i.e. DenseLayer::W of the reference
//                           (src/layers/denseLayer.cu:217-227), so D = A * W^T.
// =====================================================================================
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace nsdf {
namespace tc {

constexpr uint32_t B_LAYER_BYTES = 2048;   // 32 x 32 16-bit weights
constexpr uint32_t B_LBO = 512;            // k-chunk (8 elements) stride
constexpr uint32_t B_SBO = 128;            // 8-row group stride
constexpr uint32_t B_MMA_BYTES = 1024;     // one K = 16 slice (two k-chunks)

// byte offset of element (n, k) inside a B image of `rows` rows (k-chunk stride = rows * 16 bytes)
__host__ __device__ constexpr uint32_t b_elem_offset(int n, int k, int rows = 32) {
    return (uint32_t)(k >> 3) * (uint32_t)(rows * 16) + (uint32_t)(n >> 3) * B_SBO + (uint32_t)(n & 7) * 16u +
           (uint32_t)(k & 7) * 2u;
}

// tcgen05 instruction descriptor, kind::f16, fp32 accumulate, A and B K-major.
// fmt: 0 = fp16, 1 = bf16 (bits [7,10) and [10,13)); N>>3 at [17,23); M>>4 at [24,29).
// d_f32: accumulator format, bits [4,6): 1 = fp32, 0 = fp16 (one 16-bit value per 32-bit TMEM column).
__host__ __device__ constexpr uint32_t make_idesc(int fmt, int M, int N, bool d_f32 = true) {
    return (d_f32 ? (1u << 4) : 0u) | ((uint32_t)fmt << 7) | ((uint32_t)fmt << 10) | ((uint32_t)(N >> 3) << 17) |
           ((uint32_t)(M >> 4) << 24);
}

// shared-memory matrix descriptor (no swizzle, K-major): start>>4 [0,14), LBO>>4 [16,30),
// SBO>>4 [32,46), version = 1 at [46,48).
__device__ __forceinline__ uint64_t make_bdesc(uint32_t smem_addr, uint32_t lbo = B_LBO) {
    return (uint64_t)((smem_addr & 0x3FFFFu) >> 4) | ((uint64_t)(lbo >> 4) << 16) | ((uint64_t)(B_SBO >> 4) << 32) |
           (1ull << 46);
}

__device__ __forceinline__ void tmem_alloc(uint32_t* smem_dst, uint32_t cols) {   // whole warp
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                     (uint32_t)__cvta_generic_to_shared(smem_dst)),
                 "r"(cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t cols) {     // whole warp
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
}
__device__ __forceinline__ void fence_before_sync() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_after_sync() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// One lane of a CONVERGED warp.  Issuing tcgen05.mma under `if (elect_one())` lets ptxas keep the operands on the
// uniform datapath and emit back-to-back UTCHMMA; under `if (threadIdx.x == 0)` it wraps every MMA in an
// ELECT / BRA.U.ANY loop that costs ~50 cycles per instruction (measured with tools/mma_bench.cu).
__device__ __forceinline__ uint32_t elect_one() {
    uint32_t pred;
    asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
    return pred;
}

// D[tmem_d] (+)= A[tmem_a] * B[desc]   (one elected thread)
__device__ __forceinline__ void mma_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(tmem_d),
        "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// Arrives on the mbarrier when every MMA issued so far by this thread has completed.
__device__ __forceinline__ void mma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(
                     (uint32_t)__cvta_generic_to_shared(bar))
                 : "memory");
}

// mbarrier wait that traps instead of hanging the device if the arrival never comes.
__device__ __forceinline__ void mbar_wait_guard(uint64_t* bar, uint32_t parity) {
    const uint32_t a = (uint32_t)__cvta_generic_to_shared(bar);
    uint32_t done;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(done)
                 : "r"(a), "r"(parity)
                 : "memory");
    if (done) return;
    const long long t0 = clock64();
    for (;;) {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done)
                     : "r"(a), "r"(parity)
                     : "memory");
        if (done) return;
        if (clock64() - t0 > 4000000000LL) __trap();   // ~2 s: a lost arrival, not a slow one
    }
}

// 32 lanes x 32-bit, N consecutive columns; each thread owns TMEM lane 32*(warp%4)+lane.
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
          "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr)
        : "memory");
}
// 32 columns holding one 16-bit value each (fp16 accumulators) -> 16 registers of packed pairs.  Costs the TMEM read
// port half of what tmem_ld32 does (tools/tmem_bench.cu: the port is bound by delivered register bytes).
__device__ __forceinline__ void tmem_ld16_pack(uint32_t taddr, uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.pack::16b.b32 "
        "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr)
        : "memory");
}
// relu on a packed pair of fp16 (HMNMX2)
__device__ __forceinline__ uint32_t relu_f16x2(uint32_t v) {
    uint32_t d;
    asm("max.f16x2 %0, %1, %2;" : "=r"(d) : "r"(v), "r"(0u));
    return d;
}
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t (&r)[16]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
        "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};" ::"r"(taddr),
        "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
        "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
        : "memory");
}
__device__ __forceinline__ void tmem_st8(uint32_t taddr, const uint32_t (&r)[8]) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"r"(taddr), "r"(r[0]), "r"(r[1]),
                 "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7])
                 : "memory");
}

__device__ __forceinline__ uint32_t tmem_ld1(uint32_t taddr) {
    uint32_t r;
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x1.b32 {%0}, [%1];" : "=r"(r) : "r"(taddr) : "memory");
    return r;
}

// relu + round-to-nearest pack: low half = lo, high half = hi.
__device__ __forceinline__ uint32_t pack_relu_f16(float lo, float hi) {
    uint32_t d;
    asm("cvt.rn.relu.f16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(hi), "f"(lo));
    return d;
}
__device__ __forceinline__ uint32_t pack_relu_bf16(float lo, float hi) {
    uint32_t d;
    asm("cvt.rn.relu.bf16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(hi), "f"(lo));
    return d;
}
__device__ __forceinline__ uint32_t pack_f16(float lo, float hi) {
    uint32_t d;
    asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(hi), "f"(lo));
    return d;
}
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
    uint32_t d;
    asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(hi), "f"(lo));
    return d;
}
// relu + round-toward-zero pack: for v > 0 the packed value never exceeds v, so v - packed >= 0
__device__ __forceinline__ uint32_t pack_relu_rz_f16(float lo, float hi) {
    uint32_t d;
    asm("cvt.rz.relu.f16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(hi), "f"(lo));
    return d;
}
__device__ __forceinline__ float f16lo_to_f32(uint32_t p) {
    float f;
    asm("{\n\t.reg .b16 l, h;\n\tmov.b32 {l, h}, %1;\n\tcvt.f32.f16 %0, l;\n\t}" : "=f"(f) : "r"(p));
    return f;
}
__device__ __forceinline__ float f16hi_to_f32(uint32_t p) {
    float f;
    asm("{\n\t.reg .b16 l, h;\n\tmov.b32 {l, h}, %1;\n\tcvt.f32.f16 %0, h;\n\t}" : "=f"(f) : "r"(p));
    return f;
}

}  // namespace tc
}  // namespace nsdf
