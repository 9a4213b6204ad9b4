// =====================================================================================
// mlp_fp32.cuh -- fused all-layer SDF-MLP evaluators, fp32, reference arithmetic order.
//
// Replaces the nine CUTLASS SIMT GemmBatched launches of one NeuralNetwork::forward
// (src/neuralNetwork.cpp:54-63 -> src/layers/denseLayer.cu:126-176) by ONE kernel: the
// whole network's weights are staged once per CTA into shared memory with a TMA bulk copy
// (cp.async.bulk + mbarrier) and activations never leave registers / shared memory.
//
// Arithmetic (identical to the CUTLASS 2.0 SIMT kernels the reference instantiates, so the
// results are the reference's bit for bit): per output, one accumulator starting at 0,
// acc = fma(W[o][k], a[k], acc) for k ascending; z = acc + bias; ReLU on hidden layers.
//
// Two evaluators over the same packed weights:
//   * mlp32_eval_thread<P>  one thread = P points, weights read as warp-uniform LDS.128
//                           (throughput form: 64 FFMA per LDS.128 at P = 2)
//   * mlp32_eval_warp<R>    one warp = R points, lane = neuron (latency form: one layer is a
//                           32-deep FFMA chain instead of 1024 FFMA issue slots)
// =====================================================================================
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace nsdf {

// ---- packed "fused blob" (floats) ---------------------------------------------------------
//   [0,128)    Wt0[k][o], k < 4 (rows >= in0 are zero), o < 32
//   [128,160)  b0[o]
//   160 + l*1056 : hidden layer l (l < n_sq):  Wt[k][o] (1024) then b[o] (32)
//   160 + n_sq*1056 : last layer w[k] (32), b (1), 3 pad
constexpr int FUSED_L0_FLOATS = 160;
constexpr int FUSED_SQ_FLOATS = 1056;
constexpr int FUSED_LAST_FLOATS = 36;
constexpr int FUSED_MAX_SQ = 15;
__host__ __device__ constexpr int fused_blob_floats(int n_sq) {
    return FUSED_L0_FLOATS + n_sq * FUSED_SQ_FLOATS + FUSED_LAST_FLOATS;
}

// ---- TMA bulk staging -----------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
// 1-D TMA: global -> shared::cta, completion on an mbarrier (SASS: UBLKCP).
__device__ __forceinline__ void tma_bulk_g2s(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(smem_dst)),
                 "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

// Stages `bytes` (multiple of 16) of the fused blob into shared memory.  Must be called by all
// threads of the CTA; returns after the data is visible to every thread.
__device__ __forceinline__ void stage_weights(float* s_dst, const float* __restrict__ g_blob, uint32_t bytes,
                                              uint64_t* s_bar) {
    if (threadIdx.x == 0) {
        mbar_init(s_bar, 1);
        mbar_fence_init();
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        mbar_arrive_expect_tx(s_bar, bytes);
        // chunks of <= 32 KB keep each bulk request small enough for any smem carve-out
        uint32_t off = 0;
        while (off < bytes) {
            const uint32_t n = (bytes - off) < 32768u ? (bytes - off) : 32768u;
            tma_bulk_g2s((char*)s_dst + off, (const char*)g_blob + off, n, s_bar);
            off += n;
        }
    }
    mbar_wait(s_bar, 0);
}

__device__ __forceinline__ float relu_det(float v) { return (v > 0.0f) ? v : 0.0f; }

// ---- thread-per-point evaluator ---------------------------------------------------------------
// x[p][0..2] -> y[p].  sw = fused blob in shared memory.  All weight reads are warp-uniform
// 16-byte loads; with -fmad=false the only FMAs are the explicit ones, in reference order.
//
// 4-input (x, y, z, t) animation nets (src/volumeRender_kernel.cu:524-545, `--animation`): in4 == true adds the fourth
// product, k = 3, to the ascending-k chain with the call-uniform t4 (the frame number); in4 is uniform and false by
// default, so 3-input callers compile to what they were.
template <int P>
__device__ __forceinline__ void mlp32_eval_thread(const float* __restrict__ sw, int n_sq, const float (&x)[P][3],
                                                  float (&y)[P], float t4 = 0.0f, bool in4 = false) {
    float a[P][32];
    if (in4) {
        const float4* L0 = reinterpret_cast<const float4*>(sw);
#pragma unroll
        for (int oc = 0; oc < 8; ++oc) {
            const float4 w0 = L0[oc], w1 = L0[8 + oc], w2 = L0[16 + oc], w3 = L0[24 + oc], bb = L0[32 + oc];
#pragma unroll
            for (int p = 0; p < P; ++p) {
                a[p][4 * oc + 0] = relu_det(fmaf(w3.x, t4, fmaf(w2.x, x[p][2], fmaf(w1.x, x[p][1], fmaf(w0.x, x[p][0], 0.0f)))) + bb.x);
                a[p][4 * oc + 1] = relu_det(fmaf(w3.y, t4, fmaf(w2.y, x[p][2], fmaf(w1.y, x[p][1], fmaf(w0.y, x[p][0], 0.0f)))) + bb.y);
                a[p][4 * oc + 2] = relu_det(fmaf(w3.z, t4, fmaf(w2.z, x[p][2], fmaf(w1.z, x[p][1], fmaf(w0.z, x[p][0], 0.0f)))) + bb.z);
                a[p][4 * oc + 3] = relu_det(fmaf(w3.w, t4, fmaf(w2.w, x[p][2], fmaf(w1.w, x[p][1], fmaf(w0.w, x[p][0], 0.0f)))) + bb.w);
            }
        }
    } else {
        const float4* L0 = reinterpret_cast<const float4*>(sw);
#pragma unroll
        for (int oc = 0; oc < 8; ++oc) {
            const float4 w0 = L0[oc], w1 = L0[8 + oc], w2 = L0[16 + oc], bb = L0[32 + oc];
#pragma unroll
            for (int p = 0; p < P; ++p) {
                a[p][4 * oc + 0] = relu_det(fmaf(w2.x, x[p][2], fmaf(w1.x, x[p][1], fmaf(w0.x, x[p][0], 0.0f))) + bb.x);
                a[p][4 * oc + 1] = relu_det(fmaf(w2.y, x[p][2], fmaf(w1.y, x[p][1], fmaf(w0.y, x[p][0], 0.0f))) + bb.y);
                a[p][4 * oc + 2] = relu_det(fmaf(w2.z, x[p][2], fmaf(w1.z, x[p][1], fmaf(w0.z, x[p][0], 0.0f))) + bb.z);
                a[p][4 * oc + 3] = relu_det(fmaf(w2.w, x[p][2], fmaf(w1.w, x[p][1], fmaf(w0.w, x[p][0], 0.0f))) + bb.w);
            }
        }
    }
#pragma unroll 1
    for (int l = 0; l < n_sq; ++l) {
        const float4* Wt = reinterpret_cast<const float4*>(sw + FUSED_L0_FLOATS + l * FUSED_SQ_FLOATS);
        float z[P][32];
#pragma unroll
        for (int p = 0; p < P; ++p)
#pragma unroll
            for (int o = 0; o < 32; ++o) z[p][o] = 0.0f;
#pragma unroll
        for (int k = 0; k < 32; ++k) {
#pragma unroll
            for (int oc = 0; oc < 8; ++oc) {
                const float4 w = Wt[k * 8 + oc];
#pragma unroll
                for (int p = 0; p < P; ++p) {
                    z[p][4 * oc + 0] = fmaf(w.x, a[p][k], z[p][4 * oc + 0]);
                    z[p][4 * oc + 1] = fmaf(w.y, a[p][k], z[p][4 * oc + 1]);
                    z[p][4 * oc + 2] = fmaf(w.z, a[p][k], z[p][4 * oc + 2]);
                    z[p][4 * oc + 3] = fmaf(w.w, a[p][k], z[p][4 * oc + 3]);
                }
            }
        }
#pragma unroll
        for (int oc = 0; oc < 8; ++oc) {
            const float4 bb = Wt[256 + oc];
#pragma unroll
            for (int p = 0; p < P; ++p) {
                a[p][4 * oc + 0] = relu_det(z[p][4 * oc + 0] + bb.x);
                a[p][4 * oc + 1] = relu_det(z[p][4 * oc + 1] + bb.y);
                a[p][4 * oc + 2] = relu_det(z[p][4 * oc + 2] + bb.z);
                a[p][4 * oc + 3] = relu_det(z[p][4 * oc + 3] + bb.w);
            }
        }
    }
    {
        const float4* LL = reinterpret_cast<const float4*>(sw + FUSED_L0_FLOATS + n_sq * FUSED_SQ_FLOATS);
        float acc[P];
#pragma unroll
        for (int p = 0; p < P; ++p) acc[p] = 0.0f;
#pragma unroll
        for (int kc = 0; kc < 8; ++kc) {
            const float4 w = LL[kc];
#pragma unroll
            for (int p = 0; p < P; ++p) {
                acc[p] = fmaf(w.x, a[p][4 * kc + 0], acc[p]);
                acc[p] = fmaf(w.y, a[p][4 * kc + 1], acc[p]);
                acc[p] = fmaf(w.z, a[p][4 * kc + 2], acc[p]);
                acc[p] = fmaf(w.w, a[p][4 * kc + 3], acc[p]);
            }
        }
        const float bl = sw[FUSED_L0_FLOATS + n_sq * FUSED_SQ_FLOATS + 32];
#pragma unroll
        for (int p = 0; p < P; ++p) y[p] = acc[p] + bl;
    }
}

// ---- thread-per-4-points evaluator, activations in shared memory -------------------------------------
// Measured on a B200 (profiles/r01_smem_delivery_notes.md): shared memory delivers ~128 B/clk/SM into registers
// (a warp-uniform LDS.128 costs two wavefronts, a quarter-warp-replicated one four), so an evaluator is bound by that
// pipe unless every loaded weight feeds >= 4 FFMAs of its lane.  mlp32_eval_thread<2> sits at 8 FFMA per LDS.128
// (pipe-balanced, 53 % of the FMA peak); this form keeps the 4 x 32 accumulators of FOUR points in registers and
// the points' activations in a thread-private shared-memory column (act[k * stride + thread] = float4 of the four
// points' a[k]): per k one LDS.128 of activations and eight warp-uniform LDS.128 of weights feed 128 FFMAs.
// P = 2 is the same idea inside a 128-register budget (64 accumulators): 8 FFMA per weight load like
// mlp32_eval_thread<2>, at half its registers.  Same chains, same bits.
template <int P> struct ActVec;
template <> struct ActVec<2> {
    typedef float2 type;
    static __device__ __forceinline__ float2 make(const float (&z)[2]) { return make_float2(z[0], z[1]); }
    static __device__ __forceinline__ void get(const float2& v, float (&a)[2]) { a[0] = v.x; a[1] = v.y; }
};
template <> struct ActVec<4> {
    typedef float4 type;
    static __device__ __forceinline__ float4 make(const float (&z)[4]) { return make_float4(z[0], z[1], z[2], z[3]); }
    static __device__ __forceinline__ void get(const float4& v, float (&a)[4]) { a[0] = v.x; a[1] = v.y; a[2] = v.z; a[3] = v.w; }
};
// act: this thread's column (element k at act[k * stride]); P = 2 or 4 points per thread.
template <int P, int KU = 4, int PEEL = 1>
__device__ __forceinline__ void mlp32_eval_thread_s(const float* __restrict__ sw, int n_sq,
                                                    typename ActVec<P>::type* __restrict__ act, int stride,
                                                    const float (&x)[P][3], float (&y)[P], float t4 = 0.0f,
                                                    bool in4 = false) {
    float z[32][P];
    {
        const float4* L0 = reinterpret_cast<const float4*>(sw);
#pragma unroll
        for (int oc = 0; oc < 8; ++oc) {
            const float4 w0 = L0[oc], w1 = L0[8 + oc], w2 = L0[16 + oc], bb = L0[32 + oc];
            float4 w3 = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
            if (in4) w3 = L0[24 + oc];
#pragma unroll
            for (int p = 0; p < P; ++p) {
                float c0 = fmaf(w2.x, x[p][2], fmaf(w1.x, x[p][1], fmaf(w0.x, x[p][0], 0.0f)));
                float c1 = fmaf(w2.y, x[p][2], fmaf(w1.y, x[p][1], fmaf(w0.y, x[p][0], 0.0f)));
                float c2 = fmaf(w2.z, x[p][2], fmaf(w1.z, x[p][1], fmaf(w0.z, x[p][0], 0.0f)));
                float c3 = fmaf(w2.w, x[p][2], fmaf(w1.w, x[p][1], fmaf(w0.w, x[p][0], 0.0f)));
                if (in4) { c0 = fmaf(w3.x, t4, c0); c1 = fmaf(w3.y, t4, c1); c2 = fmaf(w3.z, t4, c2); c3 = fmaf(w3.w, t4, c3); }
                z[4 * oc + 0][p] = relu_det(c0 + bb.x);
                z[4 * oc + 1][p] = relu_det(c1 + bb.y);
                z[4 * oc + 2][p] = relu_det(c2 + bb.z);
                z[4 * oc + 3][p] = relu_det(c3 + bb.w);
            }
        }
    }
#pragma unroll 1
    for (int l = 0; l < n_sq; ++l) {
        const float4* Wt = reinterpret_cast<const float4*>(sw + FUSED_L0_FLOATS + l * FUSED_SQ_FLOATS);
        // PEEL: k = 0 apart -- its activation is still in registers, and fma(w, a, +0) starts every chain without 32 x P clears
        // (dense 256^3 sweep 3.44 -> 3.63 Gq/s); callers that need the smallest instruction footprint keep the plain loop
        if (PEEL > 0) {      // PEEL = 2 takes k = 1 from registers as well (an even trip count is left for a k-unroll of 2)
            float a0[P], a1[P];
#pragma unroll
            for (int p = 0; p < P; ++p) { a0[p] = z[0][p]; a1[p] = z[1][p]; }
#pragma unroll
            for (int o = PEEL; o < 32; ++o) act[o * stride] = ActVec<P>::make(z[o]);
#pragma unroll
            for (int oc = 0; oc < 8; ++oc) {
                const float4 w = Wt[oc];
#pragma unroll
                for (int p = 0; p < P; ++p) {
                    z[4 * oc + 0][p] = fmaf(w.x, a0[p], 0.0f);
                    z[4 * oc + 1][p] = fmaf(w.y, a0[p], 0.0f);
                    z[4 * oc + 2][p] = fmaf(w.z, a0[p], 0.0f);
                    z[4 * oc + 3][p] = fmaf(w.w, a0[p], 0.0f);
                }
            }
            if (PEEL > 1) {
#pragma unroll
                for (int oc = 0; oc < 8; ++oc) {
                    const float4 w = Wt[8 + oc];
#pragma unroll
                    for (int p = 0; p < P; ++p) {
                        z[4 * oc + 0][p] = fmaf(w.x, a1[p], z[4 * oc + 0][p]);
                        z[4 * oc + 1][p] = fmaf(w.y, a1[p], z[4 * oc + 1][p]);
                        z[4 * oc + 2][p] = fmaf(w.z, a1[p], z[4 * oc + 2][p]);
                        z[4 * oc + 3][p] = fmaf(w.w, a1[p], z[4 * oc + 3][p]);
                    }
                }
            }
        } else {
#pragma unroll
            for (int o = 0; o < 32; ++o) {
                act[o * stride] = ActVec<P>::make(z[o]);
#pragma unroll
                for (int p = 0; p < P; ++p) z[o][p] = 0.0f;
            }
        }
#pragma unroll KU
        for (int k = PEEL; k < 32; ++k) {
            float av[P];
            ActVec<P>::get(act[k * stride], av);
#pragma unroll
            for (int oc = 0; oc < 8; ++oc) {
                const float4 w = Wt[k * 8 + oc];
#pragma unroll
                for (int p = 0; p < P; ++p) {
                    z[4 * oc + 0][p] = fmaf(w.x, av[p], z[4 * oc + 0][p]);
                    z[4 * oc + 1][p] = fmaf(w.y, av[p], z[4 * oc + 1][p]);
                    z[4 * oc + 2][p] = fmaf(w.z, av[p], z[4 * oc + 2][p]);
                    z[4 * oc + 3][p] = fmaf(w.w, av[p], z[4 * oc + 3][p]);
                }
            }
        }
#pragma unroll
        for (int oc = 0; oc < 8; ++oc) {
            const float4 bb = Wt[256 + oc];
#pragma unroll
            for (int p = 0; p < P; ++p) {
                z[4 * oc + 0][p] = relu_det(z[4 * oc + 0][p] + bb.x);
                z[4 * oc + 1][p] = relu_det(z[4 * oc + 1][p] + bb.y);
                z[4 * oc + 2][p] = relu_det(z[4 * oc + 2][p] + bb.z);
                z[4 * oc + 3][p] = relu_det(z[4 * oc + 3][p] + bb.w);
            }
        }
    }
    {
        const float4* LL = reinterpret_cast<const float4*>(sw + FUSED_L0_FLOATS + n_sq * FUSED_SQ_FLOATS);
        float acc[P];
#pragma unroll
        for (int p = 0; p < P; ++p) acc[p] = 0.0f;
#pragma unroll
        for (int kc = 0; kc < 8; ++kc) {
            const float4 w = LL[kc];
#pragma unroll
            for (int p = 0; p < P; ++p) {
                acc[p] = fmaf(w.x, z[4 * kc + 0][p], acc[p]);
                acc[p] = fmaf(w.y, z[4 * kc + 1][p], acc[p]);
                acc[p] = fmaf(w.z, z[4 * kc + 2][p], acc[p]);
                acc[p] = fmaf(w.w, z[4 * kc + 3][p], acc[p]);
            }
        }
        const float bl = sw[FUSED_L0_FLOATS + n_sq * FUSED_SQ_FLOATS + 32];
#pragma unroll
        for (int p = 0; p < P; ++p) y[p] = acc[p] + bl;
    }
}

// ---- warp-cooperative evaluator ------------------------------------------------------------------
// One warp evaluates R points; lane = neuron.  x must be warp-uniform.  abuf: per-warp scratch of
// 2*R*WARP_ACT_STRIDE floats in shared memory.  y is returned warp-uniform.  Same bits as above.
constexpr int WARP_ACT_STRIDE = 36;  // 32 + 4 pad: rows start 4 banks apart (last-layer reads)
template <int R>
__device__ __forceinline__ void mlp32_eval_warp(const float* __restrict__ sw, int n_sq, float* __restrict__ abuf,
                                                const float (&x)[R][3], float (&y)[R], int lane, float t4 = 0.0f,
                                                bool in4 = false) {
    float* cur = abuf;
    float* nxt = abuf + R * WARP_ACT_STRIDE;
    {
        const float w0 = sw[lane], w1 = sw[32 + lane], w2 = sw[64 + lane], bb = sw[128 + lane];
        if (in4) {
            const float w3 = sw[96 + lane];
#pragma unroll
            for (int r = 0; r < R; ++r)
                cur[r * WARP_ACT_STRIDE + lane] =
                    relu_det(fmaf(w3, t4, fmaf(w2, x[r][2], fmaf(w1, x[r][1], fmaf(w0, x[r][0], 0.0f)))) + bb);
        } else {
#pragma unroll
            for (int r = 0; r < R; ++r)
                cur[r * WARP_ACT_STRIDE + lane] =
                    relu_det(fmaf(w2, x[r][2], fmaf(w1, x[r][1], fmaf(w0, x[r][0], 0.0f))) + bb);
        }
    }
    __syncwarp();
#pragma unroll 1
    for (int l = 0; l < n_sq; ++l) {
        const float* Wt = sw + FUSED_L0_FLOATS + l * FUSED_SQ_FLOATS;
        float z[R];
#pragma unroll
        for (int r = 0; r < R; ++r) z[r] = 0.0f;
#pragma unroll
        for (int kc = 0; kc < 8; ++kc) {
            const float wa = Wt[(4 * kc + 0) * 32 + lane], wb = Wt[(4 * kc + 1) * 32 + lane],
                        wc = Wt[(4 * kc + 2) * 32 + lane], wd = Wt[(4 * kc + 3) * 32 + lane];
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const float4 av = *reinterpret_cast<const float4*>(cur + r * WARP_ACT_STRIDE + 4 * kc);
                z[r] = fmaf(wa, av.x, z[r]);
                z[r] = fmaf(wb, av.y, z[r]);
                z[r] = fmaf(wc, av.z, z[r]);
                z[r] = fmaf(wd, av.w, z[r]);
            }
        }
        const float bb = Wt[1024 + lane];
#pragma unroll
        for (int r = 0; r < R; ++r) nxt[r * WARP_ACT_STRIDE + lane] = relu_det(z[r] + bb);
        __syncwarp();
        float* t = cur; cur = nxt; nxt = t;
    }
    float out = 0.0f;
    {
        const float* LL = sw + FUSED_L0_FLOATS + n_sq * FUSED_SQ_FLOATS;
        const int r = lane < R ? lane : 0;
        float acc = 0.0f;
#pragma unroll
        for (int kc = 0; kc < 8; ++kc) {
            const float4 w = *reinterpret_cast<const float4*>(LL + 4 * kc);
            const float4 av = *reinterpret_cast<const float4*>(cur + r * WARP_ACT_STRIDE + 4 * kc);
            acc = fmaf(w.x, av.x, acc);
            acc = fmaf(w.y, av.y, acc);
            acc = fmaf(w.z, av.z, acc);
            acc = fmaf(w.w, av.w, acc);
        }
        out = acc + LL[32];
    }
#pragma unroll
    for (int r = 0; r < R; ++r) y[r] = __shfl_sync(0xffffffffu, out, r);
    __syncwarp();   // abuf may be rewritten by the caller's next evaluation
}

// ---- warp-cooperative evaluator, 16 points, register-tiled ------------------------------------------
// One warp evaluates 16 points; lane = (g = lane & 7, rg = lane >> 3) owns the 4 x 4 tile
// {neurons 4g..4g+3} x {points rg, rg+4, rg+8, rg+12}.  A hidden layer is still a 32-deep FFMA chain per output
// (the latency of mlp32_eval_warp), but every LDS.128 now feeds 16 FFMA instead of 4: per warp and layer
// 32 weight + 32 activation wavefronts + 16 store wavefronts against 512 FFMA instructions, so the FMA pipe and
// not the shared-memory pipe bounds it (mlp32_eval_warp<8>: 104 wavefronts per 256 FFMA, LSU 94 % busy in ncu).
// Bank layout: weights Wt[k][4g..4g+3] are 8 consecutive float4 (one wavefront); activations live point-major with
// the WARP_ACT_STRIDE = 36 row stride, so the four points rg (+4j) of one load start 4 banks apart (one wavefront),
// and the float4 stores of a quarter warp (g = 0..7) cover all 32 banks.
// xin: abuf + 16*WARP_ACT_STRIDE (the second activation buffer) holds the 16 points as float4 {x, y, z, -} written
// by the caller before the call (a __syncwarp() is issued here).  Returns point (lane & 15)'s value in every lane
// of that residue.  Same bits as the evaluators above (ascending-k fmaf chain from 0, + bias, relu).
__device__ __forceinline__ float mlp32_eval_warp16(const float* __restrict__ sw, int n_sq, float* __restrict__ abuf,
                                                   int lane, float t4 = 0.0f, bool in4 = false) {
    float* cur = abuf;
    float* nxt = abuf + 16 * WARP_ACT_STRIDE;
    const int g = lane & 7, rg = lane >> 3;
    __syncwarp();
    {
        const float4* S4 = reinterpret_cast<const float4*>(sw);
        const float4 w0 = S4[g], w1 = S4[8 + g], w2 = S4[16 + g], w3 = S4[24 + g], bb = S4[32 + g];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int r = rg + 4 * j;
            const float4 x = *reinterpret_cast<const float4*>(nxt + 4 * r);
            float4 z;
            z.x = fmaf(w2.x, x.z, fmaf(w1.x, x.y, fmaf(w0.x, x.x, 0.0f)));
            z.y = fmaf(w2.y, x.z, fmaf(w1.y, x.y, fmaf(w0.y, x.x, 0.0f)));
            z.z = fmaf(w2.z, x.z, fmaf(w1.z, x.y, fmaf(w0.z, x.x, 0.0f)));
            z.w = fmaf(w2.w, x.z, fmaf(w1.w, x.y, fmaf(w0.w, x.x, 0.0f)));
            if (in4) {
                z.x = fmaf(w3.x, t4, z.x); z.y = fmaf(w3.y, t4, z.y);
                z.z = fmaf(w3.z, t4, z.z); z.w = fmaf(w3.w, t4, z.w);
            }
            z.x = relu_det(z.x + bb.x); z.y = relu_det(z.y + bb.y);
            z.z = relu_det(z.z + bb.z); z.w = relu_det(z.w + bb.w);
            *reinterpret_cast<float4*>(cur + r * WARP_ACT_STRIDE + 4 * g) = z;
        }
    }
    __syncwarp();
#pragma unroll 1
    for (int l = 0; l < n_sq; ++l) {
        const float4* Wt = reinterpret_cast<const float4*>(sw + FUSED_L0_FLOATS + l * FUSED_SQ_FLOATS) + g;
        const float* arow = cur + rg * WARP_ACT_STRIDE;
        float z[4][4];   // [neuron i][point j]
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) z[i][j] = 0.0f;
#pragma unroll
        for (int kc = 0; kc < 8; ++kc) {
            float a[4][4];   // [point j][k within the chunk]
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float4 av = *reinterpret_cast<const float4*>(arow + 4 * j * WARP_ACT_STRIDE + 4 * kc);
                a[j][0] = av.x; a[j][1] = av.y; a[j][2] = av.z; a[j][3] = av.w;
            }
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) {
                const float4 w = Wt[(4 * kc + kk) * 8];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    z[0][j] = fmaf(w.x, a[j][kk], z[0][j]);
                    z[1][j] = fmaf(w.y, a[j][kk], z[1][j]);
                    z[2][j] = fmaf(w.z, a[j][kk], z[2][j]);
                    z[3][j] = fmaf(w.w, a[j][kk], z[3][j]);
                }
            }
        }
        const float4 bb = Wt[256];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            float4 o;
            o.x = relu_det(z[0][j] + bb.x); o.y = relu_det(z[1][j] + bb.y);
            o.z = relu_det(z[2][j] + bb.z); o.w = relu_det(z[3][j] + bb.w);
            *reinterpret_cast<float4*>(nxt + (rg + 4 * j) * WARP_ACT_STRIDE + 4 * g) = o;
        }
        __syncwarp();
        float* t = cur; cur = nxt; nxt = t;
    }
    float out;
    {
        const float* LL = sw + FUSED_L0_FLOATS + n_sq * FUSED_SQ_FLOATS;
        const float* arow = cur + (lane & 15) * WARP_ACT_STRIDE;
        float acc = 0.0f;
#pragma unroll
        for (int kc = 0; kc < 8; ++kc) {
            const float4 w = *reinterpret_cast<const float4*>(LL + 4 * kc);
            const float4 av = *reinterpret_cast<const float4*>(arow + 4 * kc);
            acc = fmaf(w.x, av.x, acc);
            acc = fmaf(w.y, av.y, acc);
            acc = fmaf(w.z, av.z, acc);
            acc = fmaf(w.w, av.w, acc);
        }
        out = acc + LL[32];
    }
    __syncwarp();   // abuf may be rewritten by the caller's next evaluation
    return out;
}

// ---- warp-cooperative evaluator, 8 neurons x RJ points per lane ---------------------------------------------
// One warp evaluates 8*RJ points (RJ = 4: 32 points, RJ = 8: 64 points); lane (g = lane & 3, rg = lane >> 2) owns the tile
// {neurons 8g..8g+7} x {points rg + 8j, j < RJ}.  Why: shared memory fills registers at 128 B per clock per SM whatever the
// access pattern (only a warp-uniform load is cheaper), so an evaluator is bound by that pipe unless a loaded byte feeds at
// least one FFMA of its lane.  Per k the tile loads 8 weights + RJ activations for 8*RJ FFMA: 1.5 B per FFMA at RJ = 4 and
// 1 B at RJ = 8, against 2 B for the 4 x 4 tile of mlp32_eval_warp16 (ncu: LSU wavefronts 88 % busy, FMA pipe 53 %).
// Bank layout: weights Wt[k][8g..8g+7] are two float4 per lane, four distinct addresses per warp, conflict-free; activations
// live point-major with the WARP_ACT_STRIDE = 36 row stride, the eight rows rg + 8j of one load start four banks apart, and
// the two float4 stores per point of a quarter warp cover all 32 banks.
// xin: abuf + 8*RJ*WARP_ACT_STRIDE (the second activation buffer) holds the points as float4 {x, y, z, -}, written by the
// caller (a __syncwarp() is issued here).  Returns the value of point `lane` in y[0] (and of point lane + 32 in y[1] for
// RJ = 8).  Same bits as every other evaluator (ascending-k fmaf chain from 0, + bias, relu).
template <int RJ>
__device__ __forceinline__ void mlp32_eval_tile(const float* __restrict__ sw, int n_sq, float* __restrict__ abuf, int lane,
                                                float (&y)[RJ / 4], float t4 = 0.0f, bool in4 = false) {
    constexpr int NP = 8 * RJ;
    float* cur = abuf;
    float* nxt = abuf + NP * WARP_ACT_STRIDE;
    const int g = lane & 3, rg = lane >> 2;
    __syncwarp();
    {
        const float4* S4 = reinterpret_cast<const float4*>(sw) + 2 * g;
        float4 w[4][2], bb[2];
#pragma unroll
        for (int k = 0; k < 4; ++k) { w[k][0] = S4[8 * k]; w[k][1] = S4[8 * k + 1]; }
        bb[0] = S4[32]; bb[1] = S4[33];
#pragma unroll
        for (int j = 0; j < RJ; ++j) {
            const int r = rg + 8 * j;
            const float4 x = *reinterpret_cast<const float4*>(nxt + 4 * r);
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                float4 z;
                z.x = fmaf(w[2][h].x, x.z, fmaf(w[1][h].x, x.y, fmaf(w[0][h].x, x.x, 0.0f)));
                z.y = fmaf(w[2][h].y, x.z, fmaf(w[1][h].y, x.y, fmaf(w[0][h].y, x.x, 0.0f)));
                z.z = fmaf(w[2][h].z, x.z, fmaf(w[1][h].z, x.y, fmaf(w[0][h].z, x.x, 0.0f)));
                z.w = fmaf(w[2][h].w, x.z, fmaf(w[1][h].w, x.y, fmaf(w[0][h].w, x.x, 0.0f)));
                if (in4) {
                    z.x = fmaf(w[3][h].x, t4, z.x); z.y = fmaf(w[3][h].y, t4, z.y);
                    z.z = fmaf(w[3][h].z, t4, z.z); z.w = fmaf(w[3][h].w, t4, z.w);
                }
                z.x = relu_det(z.x + bb[h].x); z.y = relu_det(z.y + bb[h].y);
                z.z = relu_det(z.z + bb[h].z); z.w = relu_det(z.w + bb[h].w);
                *reinterpret_cast<float4*>(cur + r * WARP_ACT_STRIDE + 8 * g + 4 * h) = z;
            }
        }
    }
    __syncwarp();
#pragma unroll 1
    for (int l = 0; l < n_sq; ++l) {
        const float4* Wt = reinterpret_cast<const float4*>(sw + FUSED_L0_FLOATS + l * FUSED_SQ_FLOATS) + 2 * g;
        const float* arow = cur + rg * WARP_ACT_STRIDE;
        float z[8][RJ];   // [neuron i][point j]
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
            for (int j = 0; j < RJ; ++j) z[i][j] = 0.0f;
#pragma unroll
        for (int kc = 0; kc < 8; ++kc) {
            float a[RJ][4];   // [point j][k within the chunk]
#pragma unroll
            for (int j = 0; j < RJ; ++j) {
                const float4 av = *reinterpret_cast<const float4*>(arow + 8 * j * WARP_ACT_STRIDE + 4 * kc);
                a[j][0] = av.x; a[j][1] = av.y; a[j][2] = av.z; a[j][3] = av.w;
            }
#pragma unroll
            for (int kk = 0; kk < 4; ++kk) {
                const float4 w0 = Wt[(4 * kc + kk) * 8], w1 = Wt[(4 * kc + kk) * 8 + 1];
#pragma unroll
                for (int j = 0; j < RJ; ++j) {
                    z[0][j] = fmaf(w0.x, a[j][kk], z[0][j]);
                    z[1][j] = fmaf(w0.y, a[j][kk], z[1][j]);
                    z[2][j] = fmaf(w0.z, a[j][kk], z[2][j]);
                    z[3][j] = fmaf(w0.w, a[j][kk], z[3][j]);
                    z[4][j] = fmaf(w1.x, a[j][kk], z[4][j]);
                    z[5][j] = fmaf(w1.y, a[j][kk], z[5][j]);
                    z[6][j] = fmaf(w1.z, a[j][kk], z[6][j]);
                    z[7][j] = fmaf(w1.w, a[j][kk], z[7][j]);
                }
            }
        }
        const float4 b0 = Wt[256], b1 = Wt[257];
#pragma unroll
        for (int j = 0; j < RJ; ++j) {
            float4 o0, o1;
            o0.x = relu_det(z[0][j] + b0.x); o0.y = relu_det(z[1][j] + b0.y);
            o0.z = relu_det(z[2][j] + b0.z); o0.w = relu_det(z[3][j] + b0.w);
            o1.x = relu_det(z[4][j] + b1.x); o1.y = relu_det(z[5][j] + b1.y);
            o1.z = relu_det(z[6][j] + b1.z); o1.w = relu_det(z[7][j] + b1.w);
            float* dst = nxt + (rg + 8 * j) * WARP_ACT_STRIDE + 8 * g;
            *reinterpret_cast<float4*>(dst) = o0;
            *reinterpret_cast<float4*>(dst + 4) = o1;
        }
        __syncwarp();
        float* t = cur; cur = nxt; nxt = t;
    }
    {
        const float* LL = sw + FUSED_L0_FLOATS + n_sq * FUSED_SQ_FLOATS;
        float acc[RJ / 4];
#pragma unroll
        for (int q = 0; q < RJ / 4; ++q) acc[q] = 0.0f;
#pragma unroll
        for (int kc = 0; kc < 8; ++kc) {
            const float4 w = *reinterpret_cast<const float4*>(LL + 4 * kc);
#pragma unroll
            for (int q = 0; q < RJ / 4; ++q) {
                const float4 av = *reinterpret_cast<const float4*>(cur + (lane + 32 * q) * WARP_ACT_STRIDE + 4 * kc);
                acc[q] = fmaf(w.x, av.x, acc[q]);
                acc[q] = fmaf(w.y, av.y, acc[q]);
                acc[q] = fmaf(w.z, av.z, acc[q]);
                acc[q] = fmaf(w.w, av.w, acc[q]);
            }
        }
#pragma unroll
        for (int q = 0; q < RJ / 4; ++q) y[q] = acc[q] + LL[32];
    }
    __syncwarp();   // abuf may be rewritten by the caller's next evaluation
}

}  // namespace nsdf
