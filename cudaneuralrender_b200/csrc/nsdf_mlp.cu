// =====================================================================================
// nsdf_mlp.cu -- dense-query kernels (fp32 rungs) and the per-layer operators.
// Compiled with -fmad=false: the explicit fmaf() calls are the only fused operations.
// =====================================================================================
#include "lattice.cuh"
#include "mlp_fp32.cuh"
#include "mlp_generic.cuh"
#include "nsdf_internal.h"

namespace nsdf {

#ifdef NSDF_DEV_SCHEDULES   // development rungs (make dev); the product runs mlp_fp32_thread_s_kernel and mlp_fp32_warp_kernel<4>
// Thread-per-point fused MLP.  Persistent: chunks of blockDim*P points are strided over the grid.
template <int P>
__global__ void __launch_bounds__(256, (P >= 2) ? 1 : 2)
mlp_fp32_thread_kernel(const float* __restrict__ blob, uint32_t blob_bytes, int n_sq, const float* __restrict__ xin,
                       float* __restrict__ yout, long long n, const LatticeParams lp) {
    extern __shared__ __align__(16) float s_w[];
    __shared__ uint64_t s_bar;
    stage_weights(s_w, blob, blob_bytes, &s_bar);
    const long long chunk = (long long)blockDim.x * P;
    for (long long base = (long long)blockIdx.x * chunk; base < n; base += (long long)gridDim.x * chunk) {
        float x[P][3], y[P];
#pragma unroll
        for (int p = 0; p < P; ++p) {
            const long long i = base + (long long)p * blockDim.x + threadIdx.x;
            x[p][0] = x[p][1] = x[p][2] = 0.0f;
            if (i < n) {
                if (lp.enabled) lattice_point(lp, lp.first + i, x[p]);
                else { x[p][0] = __ldg(xin + 3 * i); x[p][1] = __ldg(xin + 3 * i + 1); x[p][2] = __ldg(xin + 3 * i + 2); }
            }
        }
        mlp32_eval_thread<P>(s_w, n_sq, x, y);
#pragma unroll
        for (int p = 0; p < P; ++p) {
            const long long i = base + (long long)p * blockDim.x + threadIdx.x;
            if (i < n) yout[i] = y[p];
        }
    }
}
#endif  // NSDF_DEV_SCHEDULES

// Thread-per-4-points fused MLP (mlp32_eval_thread4s): accumulators in registers, activations in a thread-private
// shared-memory column behind the weights.  One CTA per SM.
template <int P, int THREADS, int MINB>
__global__ void __launch_bounds__(THREADS, MINB)
mlp_fp32_thread_s_kernel(const float* __restrict__ blob, uint32_t blob_bytes, int n_sq, const float* __restrict__ xin,
                        float* __restrict__ yout, long long n, const LatticeParams lp) {
    extern __shared__ __align__(16) float s_w[];
    __shared__ uint64_t s_bar;
    stage_weights(s_w, blob, blob_bytes, &s_bar);
    typename ActVec<P>::type* act = reinterpret_cast<typename ActVec<P>::type*>(s_w + (blob_bytes >> 2)) + threadIdx.x;
    const long long chunk = (long long)THREADS * P;
    for (long long base = (long long)blockIdx.x * chunk; base < n; base += (long long)gridDim.x * chunk) {
        float x[P][3], y[P];
#pragma unroll
        for (int p = 0; p < P; ++p) {
            const long long i = base + (long long)p * THREADS + threadIdx.x;
            x[p][0] = x[p][1] = x[p][2] = 0.0f;
            if (i < n) {
                if (lp.enabled) lattice_point(lp, lp.first + i, x[p]);
                else { x[p][0] = __ldg(xin + 3 * i); x[p][1] = __ldg(xin + 3 * i + 1); x[p][2] = __ldg(xin + 3 * i + 2); }
            }
        }
        mlp32_eval_thread_s<P>(s_w, n_sq, act, THREADS, x, y);
#pragma unroll
        for (int p = 0; p < P; ++p) {
            const long long i = base + (long long)p * THREADS + threadIdx.x;
            if (i < n) yout[i] = y[p];
        }
    }
}

// Warp-cooperative fused MLP on a dense batch (exercises the latency evaluator of the tracer).
template <int R>
__global__ void __launch_bounds__(1024, 1)
mlp_fp32_warp_kernel(const float* __restrict__ blob, uint32_t blob_bytes, int n_sq, const float* __restrict__ xin,
                     float* __restrict__ yout, long long n, const LatticeParams lp) {
    extern __shared__ __align__(16) float s_w[];
    __shared__ uint64_t s_bar;
    stage_weights(s_w, blob, blob_bytes, &s_bar);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, wpb = blockDim.x >> 5;
    float* abuf = s_w + (blob_bytes >> 2) + warp * (2 * R * WARP_ACT_STRIDE);
    const long long gw = (long long)blockIdx.x * wpb + warp, nw = (long long)gridDim.x * wpb;
    for (long long base = gw * R; base < n; base += nw * R) {
        float x[R][3], y[R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const long long i = base + r;
            x[r][0] = x[r][1] = x[r][2] = 0.0f;
            if (i < n) {
                if (lp.enabled) lattice_point(lp, lp.first + i, x[r]);
                else { x[r][0] = __ldg(xin + 3 * i); x[r][1] = __ldg(xin + 3 * i + 1); x[r][2] = __ldg(xin + 3 * i + 2); }
            }
        }
        mlp32_eval_warp<R>(s_w, n_sq, abuf, x, y, lane);
#pragma unroll
        for (int r = 0; r < R; ++r)
            if (lane == r && base + r < n) yout[base + r] = y[r];
    }
}

#ifdef NSDF_DEV_SCHEDULES   // development rungs: the tracer's tile evaluators on dense batches, constant-memory weights, the warp pipeline
// The 16-point register-tiled form of the same evaluator (mlp32_eval_warp16): what the tracer's warps run while
// more than 8 of their 16 ray slots are busy.
__global__ void __launch_bounds__(256, 2)
mlp_fp32_warp16_kernel(const float* __restrict__ blob, uint32_t blob_bytes, int n_sq, const float* __restrict__ xin,
                       float* __restrict__ yout, long long n, const LatticeParams lp) {
    extern __shared__ __align__(16) float s_w[];
    __shared__ uint64_t s_bar;
    stage_weights(s_w, blob, blob_bytes, &s_bar);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, wpb = blockDim.x >> 5;
    float* abuf = s_w + (blob_bytes >> 2) + warp * (2 * 16 * WARP_ACT_STRIDE);
    const long long gw = (long long)blockIdx.x * wpb + warp, nw = (long long)gridDim.x * wpb;
    for (long long base = gw * 16; base < n; base += nw * 16) {
        const long long i = base + lane;
        if (lane < 16) {
            float x[3] = {0.0f, 0.0f, 0.0f};
            if (i < n) {
                if (lp.enabled) lattice_point(lp, lp.first + i, x);
                else { x[0] = __ldg(xin + 3 * i); x[1] = __ldg(xin + 3 * i + 1); x[2] = __ldg(xin + 3 * i + 2); }
            }
            *reinterpret_cast<float4*>(abuf + 16 * WARP_ACT_STRIDE + 4 * lane) = make_float4(x[0], x[1], x[2], 0.0f);
        }
        const float y = mlp32_eval_warp16(s_w, n_sq, abuf, lane);
        if (lane < 16 && i < n) yout[i] = y;
    }
}

// The 8-neuron x RJ-point register tiles (mlp32_eval_tile): 32 or 64 points per warp.
template <int RJ>
__global__ void __launch_bounds__(RJ == 4 ? 512 : 256, 1)
mlp_fp32_tile_kernel(const float* __restrict__ blob, uint32_t blob_bytes, int n_sq, const float* __restrict__ xin,
                     float* __restrict__ yout, long long n, const LatticeParams lp) {
    constexpr int NP = 8 * RJ;
    extern __shared__ __align__(16) float s_w[];
    __shared__ uint64_t s_bar;
    stage_weights(s_w, blob, blob_bytes, &s_bar);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, wpb = blockDim.x >> 5;
    float* abuf = s_w + (blob_bytes >> 2) + warp * (2 * NP * WARP_ACT_STRIDE);
    const long long gw = (long long)blockIdx.x * wpb + warp, nw = (long long)gridDim.x * wpb;
    for (long long base = gw * NP; base < n; base += nw * NP) {
#pragma unroll
        for (int q = 0; q < RJ / 4; ++q) {
            const long long i = base + lane + 32 * q;
            float x[3] = {0.0f, 0.0f, 0.0f};
            if (i < n) {
                if (lp.enabled) lattice_point(lp, lp.first + i, x);
                else { x[0] = __ldg(xin + 3 * i); x[1] = __ldg(xin + 3 * i + 1); x[2] = __ldg(xin + 3 * i + 2); }
            }
            *reinterpret_cast<float4*>(abuf + NP * WARP_ACT_STRIDE + 4 * (lane + 32 * q)) = make_float4(x[0], x[1], x[2], 0.0f);
        }
        float y[RJ / 4];
        mlp32_eval_tile<RJ>(s_w, n_sq, abuf, lane, y);
#pragma unroll
        for (int q = 0; q < RJ / 4; ++q) {
            const long long i = base + lane + 32 * q;
            if (i < n) yout[i] = y[q];
        }
    }
}

// Development rung (NSDF_FP32_P=40): the thread-per-point evaluator with the weights in __constant__ memory, so that they reach
// the FFMAs through the uniform datapath / constant-bank operands instead of the shared-memory pipe (ncu: every LDS.128 costs four
// wavefronts, which is what bounds the shared-memory forms).  One symbol per device: callers on different streams would overwrite
// each other's weights, hence not a default.  Same arithmetic order, same bits.
constexpr int CONST_MAX_SQ = 8;
__constant__ float c_blob[FUSED_L0_FLOATS + CONST_MAX_SQ * FUSED_SQ_FLOATS + FUSED_LAST_FLOATS];

template <int P>
__global__ void __launch_bounds__(256, 2)
mlp_fp32_const_kernel(int n_sq, const float* __restrict__ xin, float* __restrict__ yout, long long n, const LatticeParams lp) {
    const long long stride = (long long)gridDim.x * 256 * P;
    for (long long base = (long long)blockIdx.x * 256 * P; base < n; base += stride) {
        float a[P][32];
#pragma unroll
        for (int p = 0; p < P; ++p) {
            const long long i = base + (long long)p * 256 + threadIdx.x;
            float x[3] = {0.0f, 0.0f, 0.0f};
            if (i < n) {
                if (lp.enabled) lattice_point(lp, lp.first + i, x);
                else { x[0] = __ldg(xin + 3 * i); x[1] = __ldg(xin + 3 * i + 1); x[2] = __ldg(xin + 3 * i + 2); }
            }
#pragma unroll
            for (int o = 0; o < 32; ++o)
                a[p][o] = relu_det(fmaf(c_blob[64 + o], x[2], fmaf(c_blob[32 + o], x[1], fmaf(c_blob[o], x[0], 0.0f))) + c_blob[128 + o]);
        }
#pragma unroll 1
        for (int l = 0; l < n_sq; ++l) {
            const int wb = FUSED_L0_FLOATS + l * FUSED_SQ_FLOATS;
            float z[P][32];
#pragma unroll
            for (int p = 0; p < P; ++p)
#pragma unroll
                for (int o = 0; o < 32; ++o) z[p][o] = 0.0f;
#pragma unroll
            for (int k = 0; k < 32; ++k)
#pragma unroll
                for (int o = 0; o < 32; ++o)
#pragma unroll
                    for (int p = 0; p < P; ++p) z[p][o] = fmaf(c_blob[wb + k * 32 + o], a[p][k], z[p][o]);
#pragma unroll
            for (int p = 0; p < P; ++p)
#pragma unroll
                for (int o = 0; o < 32; ++o) a[p][o] = relu_det(z[p][o] + c_blob[wb + 1024 + o]);
        }
        const int lb = FUSED_L0_FLOATS + n_sq * FUSED_SQ_FLOATS;
#pragma unroll
        for (int p = 0; p < P; ++p) {
            float acc = 0.0f;
#pragma unroll
            for (int k = 0; k < 32; ++k) acc = fmaf(c_blob[lb + k], a[p][k], acc);
            const long long i = base + (long long)p * 256 + threadIdx.x;
            if (i < n) yout[i] = acc + c_blob[lb + 32];
        }
    }
}

// Development rung (NSDF_FP32_WARP_R=77, networks with 7 hidden layers): weight-stationary warp pipeline.  Warp l < 7 keeps the
// 4 neurons x 32 k weights of hidden layer l in registers for the life of the kernel; 16-point tiles stream through the warps via
// shared memory, one layer per tick, one __syncthreads per tick; warp 7 feeds (layer 0) and drains (last layer).  Per tile-layer the
// shared-memory traffic is activations only (32 LDS.128 + 4 STS.128) -- the tiled evaluator re-reads the weights for every 16 points
// and is bound by that (profiles/r01_ncu_full_trace_coop_s26.txt).  Same k-ascending chains, same bits.
template <int NT>   // NT 16-point tiles per tick (NT * 4 points per lane)
__global__ void __launch_bounds__(256, 1)
mlp_fp32_pipe_kernel(const float* __restrict__ blob, const float* __restrict__ xin, float* __restrict__ yout, long long n,
                     const LatticeParams lp) {
    constexpr int RAYS = 16 * NT, NJ = 4 * NT;
    extern __shared__ __align__(16) float pipe_smem[];
    float* const buf = pipe_smem;                                   // [2][8][RAYS * WARP_ACT_STRIDE]
    float4* const xs = reinterpret_cast<float4*>(pipe_smem + 2 * 8 * RAYS * WARP_ACT_STRIDE);   // [RAYS]
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int g = lane & 7, rg = lane >> 3;
    const long long n_tiles = (n + RAYS - 1) / RAYS;
    const long long T = (n_tiles + gridDim.x - 1) / gridDim.x;
    float4 wreg[32];
    float4 breg;
    if (warp < 7) {
        const float4* Wt = reinterpret_cast<const float4*>(blob + FUSED_L0_FLOATS + warp * FUSED_SQ_FLOATS) + g;
#pragma unroll
        for (int k = 0; k < 32; ++k) wreg[k] = __ldg(Wt + k * 8);
        breg = __ldg(Wt + 256);
    } else {
        const float4* S4 = reinterpret_cast<const float4*>(blob);
        wreg[0] = __ldg(S4 + g); wreg[1] = __ldg(S4 + 8 + g); wreg[2] = __ldg(S4 + 16 + g); breg = __ldg(S4 + 32 + g);
        const float4* LL = reinterpret_cast<const float4*>(blob + FUSED_L0_FLOATS + 7 * FUSED_SQ_FLOATS);
#pragma unroll
        for (int k = 0; k < 8; ++k) wreg[8 + k] = __ldg(LL + k);
        wreg[16] = __ldg(LL + 8);   // .x = bias of the last layer
    }
    auto stage = [&](int par, int st) -> float* { return buf + ((par * 8 + st) * RAYS) * WARP_ACT_STRIDE; };
    for (long long t = 0; t < T + 9; ++t) {
        const int par = (int)(t & 1);
        if (warp < 7) {
            const long long tile = t - 1 - warp;
            if (tile >= 0 && tile < T) {
                const float* in = stage(par, warp) + rg * WARP_ACT_STRIDE;
                float z[4][NJ];
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int j = 0; j < NJ; ++j) z[i][j] = 0.0f;
#pragma unroll
                for (int kc = 0; kc < 8; ++kc) {
                    float a[NJ][4];
#pragma unroll
                    for (int j = 0; j < NJ; ++j) {
                        const float4 av = *reinterpret_cast<const float4*>(in + 4 * j * WARP_ACT_STRIDE + 4 * kc);
                        a[j][0] = av.x; a[j][1] = av.y; a[j][2] = av.z; a[j][3] = av.w;
                    }
#pragma unroll
                    for (int kk = 0; kk < 4; ++kk) {
                        const float4 w = wreg[4 * kc + kk];
#pragma unroll
                        for (int j = 0; j < NJ; ++j) {
                            z[0][j] = fmaf(w.x, a[j][kk], z[0][j]);
                            z[1][j] = fmaf(w.y, a[j][kk], z[1][j]);
                            z[2][j] = fmaf(w.z, a[j][kk], z[2][j]);
                            z[3][j] = fmaf(w.w, a[j][kk], z[3][j]);
                        }
                    }
                }
                float* out = stage(par ^ 1, warp + 1);
#pragma unroll
                for (int j = 0; j < NJ; ++j) {
                    float4 o;
                    o.x = relu_det(z[0][j] + breg.x); o.y = relu_det(z[1][j] + breg.y);
                    o.z = relu_det(z[2][j] + breg.z); o.w = relu_det(z[3][j] + breg.w);
                    *reinterpret_cast<float4*>(out + (rg + 4 * j) * WARP_ACT_STRIDE + 4 * g) = o;
                }
            }
        } else {
            if (t < T) {                                   // feed: layer 0 of tile t
                for (int r = lane; r < RAYS; r += 32) {
                    const long long i = ((long long)blockIdx.x + t * gridDim.x) * RAYS + r;
                    float x[3] = {0.0f, 0.0f, 0.0f};
                    if (i < n) {
                        if (lp.enabled) lattice_point(lp, lp.first + i, x);
                        else { x[0] = __ldg(xin + 3 * i); x[1] = __ldg(xin + 3 * i + 1); x[2] = __ldg(xin + 3 * i + 2); }
                    }
                    xs[r] = make_float4(x[0], x[1], x[2], 0.0f);
                }
                __syncwarp();
                float* out = stage(par ^ 1, 0);
#pragma unroll
                for (int j = 0; j < NJ; ++j) {
                    const float4 x = xs[rg + 4 * j];
                    float4 o;
                    o.x = relu_det(fmaf(wreg[2].x, x.z, fmaf(wreg[1].x, x.y, fmaf(wreg[0].x, x.x, 0.0f))) + breg.x);
                    o.y = relu_det(fmaf(wreg[2].y, x.z, fmaf(wreg[1].y, x.y, fmaf(wreg[0].y, x.x, 0.0f))) + breg.y);
                    o.z = relu_det(fmaf(wreg[2].z, x.z, fmaf(wreg[1].z, x.y, fmaf(wreg[0].z, x.x, 0.0f))) + breg.z);
                    o.w = relu_det(fmaf(wreg[2].w, x.z, fmaf(wreg[1].w, x.y, fmaf(wreg[0].w, x.x, 0.0f))) + breg.w);
                    *reinterpret_cast<float4*>(out + (rg + 4 * j) * WARP_ACT_STRIDE + 4 * g) = o;
                }
                __syncwarp();
            }
            const long long to = t - 8;                    // drain: last layer of tile t - 8
            if (to >= 0 && to < T) {
              for (int r = lane; r < RAYS; r += 32) {
                const float* in = stage(par, 7) + r * WARP_ACT_STRIDE;
                float acc = 0.0f;
#pragma unroll
                for (int kc = 0; kc < 8; ++kc) {
                    const float4 w = wreg[8 + kc];
                    const float4 av = *reinterpret_cast<const float4*>(in + 4 * kc);
                    acc = fmaf(w.x, av.x, acc);
                    acc = fmaf(w.y, av.y, acc);
                    acc = fmaf(w.z, av.z, acc);
                    acc = fmaf(w.w, av.w, acc);
                }
                const long long i = ((long long)blockIdx.x + to * gridDim.x) * RAYS + r;
                if (i < n) yout[i] = acc + wreg[16].x;
              }
            }
        }
        __syncthreads();
    }
}

#endif  // NSDF_DEV_SCHEDULES

// Generic dense layer (any M, K; N = 1 per batch): replaces denseLayerForward /
// batchedDenseLayerForward (src/layers/denseLayer.cu:86-176) with the same arithmetic.
__global__ void __launch_bounds__(256)
dense_layer_kernel(const float* __restrict__ W, const float* __restrict__ A, float* __restrict__ Z,
                   const float* __restrict__ b, int M, int K, long long num_batches, int activation) {
    const long long total = num_batches * M;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total;
         t += (long long)gridDim.x * blockDim.x) {
        const long long j = t / M;
        const int m = (int)(t % M);
        const float* w = W + (size_t)m * K;
        const float* a = A + (size_t)j * K;
        float acc = 0.0f;
        for (int k = 0; k < K; ++k) acc = fmaf(__ldg(w + k), __ldg(a + k), acc);
        float v = acc + __ldg(b + m);
        if (activation == NSDF_ACT_RELU) v = relu_det(v);
        Z[t] = v;
    }
}

#ifndef NSDF_DEV_SCHEDULES
// Product dispatch: the thread-per-4-points evaluator (3.44 Gq/s on the 256^3 sweep, 68 % of the FFMA pipe) and, for
// NSDF_PREC_FP32_WARP, the lane = neuron evaluator the tracer's thinning warps run.  No environment reads.
int fp32_forward(nsdf_model* m, const float* d_x, float* d_y, int64_t n, int precision, bool lattice, int res,
                 float lo, float hi, int64_t first, cudaStream_t stream) {
    if (n == 0) return NSDF_OK;
    LatticeParams lp{lattice ? 1 : 0, res, lo, hi, (long long)first};
    const int sms = m->sm_count > 0 ? m->sm_count : 148;
    if (precision == NSDF_PREC_FP32_WARP) {
        constexpr int R = 4, threads = 512;
        const size_t smem = m->blob_bytes + (threads / 32) * 2 * R * WARP_ACT_STRIDE * sizeof(float);
        cudaError_t e = cudaFuncSetAttribute(mlp_fp32_warp_kernel<R>, cudaFuncAttributeMaxDynamicSharedMemorySize, 110 * 1024);
        if (e != cudaSuccess) return (int)e;
        const long long want = (n + (long long)(threads / 32) * R - 1) / ((long long)(threads / 32) * R);
        const int grid = (int)(want < (long long)sms ? want : (long long)sms);
        mlp_fp32_warp_kernel<R><<<grid, threads, smem, stream>>>(m->d_blob, m->blob_bytes, m->n_sq, d_x, d_y, n, lp);
        return (int)cudaGetLastError();
    }
    // 384 threads x 512 B of activations + the weights must fit the 227 KB a CTA may own (deep nets: 256 threads)
    const bool wide = m->blob_bytes + 384u * 512u <= 227u * 1024u;
    const int th = wide ? 384 : 256;
    const size_t smem4 = m->blob_bytes + (size_t)th * 32 * 4 * sizeof(float);
    cudaError_t e = wide ? cudaFuncSetAttribute(mlp_fp32_thread_s_kernel<4, 384, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem4)
                         : cudaFuncSetAttribute(mlp_fp32_thread_s_kernel<4, 256, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem4);
    if (e != cudaSuccess) return (int)e;
    const long long chunk = (long long)th * 4;
    const long long want = (n + chunk - 1) / chunk;
    const int grid = (int)(want < (long long)sms ? want : (long long)sms);
    if (wide) mlp_fp32_thread_s_kernel<4, 384, 1><<<grid, 384, smem4, stream>>>(m->d_blob, m->blob_bytes, m->n_sq, d_x, d_y, n, lp);
    else mlp_fp32_thread_s_kernel<4, 256, 1><<<grid, 256, smem4, stream>>>(m->d_blob, m->blob_bytes, m->n_sq, d_x, d_y, n, lp);
    return (int)cudaGetLastError();
}
#else  // development build: every rung, selected through the environment
static int env_int(const char* name, int dflt) {
    const char* s = getenv(name);
    return (s && *s) ? atoi(s) : dflt;
}

int fp32_forward(nsdf_model* m, const float* d_x, float* d_y, int64_t n, int precision, bool lattice, int res,
                 float lo, float hi, int64_t first, cudaStream_t stream) {
    if (n == 0) return NSDF_OK;
    LatticeParams lp{lattice ? 1 : 0, res, lo, hi, (long long)first};
    const int sms = m->sm_count > 0 ? m->sm_count : 148;
    cudaError_t e;
    if (precision == NSDF_PREC_FP32_WARP) {
        const int R = env_int("NSDF_FP32_WARP_R", 4);
        const int threads = env_int("NSDF_FP32_WARP_THREADS", 512);
        const int ctas = env_int("NSDF_FP32_WARP_CTAS", 1);
        auto go = [&](auto kern, int r) -> int {
            const size_t smem = m->blob_bytes + (threads / 32) * 2 * r * WARP_ACT_STRIDE * sizeof(float);
            cudaError_t e2 = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 110 * 1024);
            if (e2 != cudaSuccess) return (int)e2;
            long long want = (n + (long long)(threads / 32) * r - 1) / ((long long)(threads / 32) * r);
            const long long cap = (long long)sms * ctas;
            const int grid = (int)(want < cap ? want : cap);
            kern<<<grid, threads, smem, stream>>>(m->d_blob, m->blob_bytes, m->n_sq, d_x, d_y, n, lp);
            return (int)cudaGetLastError();
        };
        if ((R == 77 || R == 78 || R == 79) && m->n_sq == 7 && m->in[0] == 3) {   // development: weight-stationary warp pipeline (1 / 2 tiles per tick)
            const int nt = R - 76;
            const long long n_tiles = (n + 16 * nt - 1) / (16 * nt);
            const int grid = (int)(n_tiles < (long long)sms ? n_tiles : (long long)sms);
            const size_t smem = (size_t)(2 * 8 * 16 * nt * WARP_ACT_STRIDE) * sizeof(float) + 16 * nt * sizeof(float4);
            cudaError_t e2 = cudaFuncSetAttribute(mlp_fp32_pipe_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
            if (e2 != cudaSuccess) return (int)e2;
            e2 = cudaFuncSetAttribute(mlp_fp32_pipe_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, 120 * 1024);
            if (e2 != cudaSuccess) return (int)e2;
            if (nt == 3) mlp_fp32_pipe_kernel<3><<<grid, 256, smem, stream>>>(m->d_blob, d_x, d_y, n, lp);
            else if (nt == 1) mlp_fp32_pipe_kernel<1><<<grid, 256, smem, stream>>>(m->d_blob, d_x, d_y, n, lp);
            else mlp_fp32_pipe_kernel<2><<<grid, 256, smem, stream>>>(m->d_blob, d_x, d_y, n, lp);
            return (int)cudaGetLastError();
        }
        if (R == 32 || R == 64) {   // 8 x 4 / 8 x 8 register tiles, one CTA per SM
            const int np = R, thr = env_int("NSDF_FP32_WARP_THREADS", R == 32 ? 512 : 256);
            const size_t smem = m->blob_bytes + (size_t)(thr / 32) * 2 * np * WARP_ACT_STRIDE * sizeof(float);
            cudaError_t e2 = cudaFuncSetAttribute(mlp_fp32_tile_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
            if (e2 != cudaSuccess) return (int)e2;
            e2 = cudaFuncSetAttribute(mlp_fp32_tile_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
            if (e2 != cudaSuccess) return (int)e2;
            if (smem > 220 * 1024) return NSDF_EINVAL;
            const long long per_cta = (long long)(thr / 32) * np;
            const long long want = (n + per_cta - 1) / per_cta;
            const int grid = (int)(want < (long long)sms ? want : (long long)sms);
            if (R == 32) mlp_fp32_tile_kernel<4><<<grid, thr, smem, stream>>>(m->d_blob, m->blob_bytes, m->n_sq, d_x, d_y, n, lp);
            else mlp_fp32_tile_kernel<8><<<grid, thr, smem, stream>>>(m->d_blob, m->blob_bytes, m->n_sq, d_x, d_y, n, lp);
            return (int)cudaGetLastError();
        }
        if (R == 16) {
            const size_t smem = m->blob_bytes + 8 * 2 * 16 * WARP_ACT_STRIDE * sizeof(float);
            cudaError_t e2 = cudaFuncSetAttribute(mlp_fp32_warp16_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 110 * 1024);
            if (e2 != cudaSuccess) return (int)e2;
            long long want = (n + 127) / 128;
            const int grid = (int)(want < 2LL * sms ? want : 2LL * sms);
            mlp_fp32_warp16_kernel<<<grid, 256, smem, stream>>>(m->d_blob, m->blob_bytes, m->n_sq, d_x, d_y, n, lp);
            return (int)cudaGetLastError();
        }
        if (R == 8) return go(mlp_fp32_warp_kernel<8>, 8);
        if (R == 2) return go(mlp_fp32_warp_kernel<2>, 2);
        if (R == 1) return go(mlp_fp32_warp_kernel<1>, 1);
        return go(mlp_fp32_warp_kernel<4>, 4);
    }
    // measured on 16.8 M points (tools/fp32_p4_probe.py): P = 2 2.72 Gq/s, P = 4 with 128/256/384 threads 2.90/3.34/3.46 Gq/s
    const int P = env_int("NSDF_FP32_P", 4);
    const size_t smem = m->blob_bytes;
    NSDF_PER_DEVICE_FLAG(attr1);
    NSDF_PER_DEVICE_FLAG(attr2);
    if (P == 4) {
        // 384 threads x 512 B of activations + the weights must fit the 227 KB a CTA may own (deep nets: 256 threads)
        const int threads = env_int("NSDF_FP32_P4_THREADS", m->blob_bytes + 384u * 512u <= 227u * 1024u ? 384 : 256);
        auto go4 = [&](auto kern, int th, int pp = 4, int minb = 1) -> int {
            const size_t smem4 = m->blob_bytes + (size_t)th * 32 * pp * sizeof(float);
            cudaError_t e2 = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem4);
            if (e2 != cudaSuccess) return (int)e2;
            const long long chunk = (long long)th * pp;
            long long want = (n + chunk - 1) / chunk;
            const int grid = (int)(want < (long long)sms * minb ? want : (long long)sms * minb);
            kern<<<grid, th, smem4, stream>>>(m->d_blob, m->blob_bytes, m->n_sq, d_x, d_y, n, lp);
            return (int)cudaGetLastError();
        };
        if (threads == 384) return go4(mlp_fp32_thread_s_kernel<4, 384, 1>, 384);
        if (threads == 128) return go4(mlp_fp32_thread_s_kernel<4, 128, 1>, 128);
        return go4(mlp_fp32_thread_s_kernel<4, 256, 1>, 256);
    }
    if ((P == 40 || P == 41) && m->n_sq <= CONST_MAX_SQ && m->in[0] == 3) {   // development: weights through the constant path
        e = cudaMemcpyToSymbolAsync(c_blob, m->d_blob, m->blob_bytes, 0, cudaMemcpyDeviceToDevice, stream);
        if (e != cudaSuccess) return (int)e;
        const int pp = P == 40 ? 2 : 1;
        const long long chunk = 256LL * pp;
        long long want = (n + chunk - 1) / chunk;
        const int grid = (int)(want < 2LL * sms ? want : 2LL * sms);
        if (pp == 2) mlp_fp32_const_kernel<2><<<grid, 256, 0, stream>>>(m->n_sq, d_x, d_y, n, lp);
        else mlp_fp32_const_kernel<1><<<grid, 256, 0, stream>>>(m->n_sq, d_x, d_y, n, lp);
        return (int)cudaGetLastError();
    }
    if (P == 22) {   // development: two points per thread, activations in shared memory, the tracer's register budget
        auto go2 = [&](auto kern) -> int {
            const size_t smem2 = m->blob_bytes + (size_t)256 * 32 * 2 * sizeof(float);
            cudaError_t e2 = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2);
            if (e2 != cudaSuccess) return (int)e2;
            long long want = (n + 511) / 512;
            const int grid = (int)(want < 2LL * sms ? want : 2LL * sms);
            kern<<<grid, 256, smem2, stream>>>(m->d_blob, m->blob_bytes, m->n_sq, d_x, d_y, n, lp);
            return (int)cudaGetLastError();
        };
        return go2(mlp_fp32_thread_s_kernel<2, 256, 2>);
    }
    if (P == 1) {
        if (!attr1) {
            e = cudaFuncSetAttribute(mlp_fp32_thread_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
            if (e != cudaSuccess) return (int)e;
            attr1 = true;
        }
        const long long chunk = 256;
        long long want = (n + chunk - 1) / chunk;
        const int grid = (int)(want < 2LL * sms ? want : 2LL * sms);
        mlp_fp32_thread_kernel<1><<<grid, 256, smem, stream>>>(m->d_blob, m->blob_bytes, m->n_sq, d_x, d_y, n, lp);
    } else {
        if (!attr2) {
            e = cudaFuncSetAttribute(mlp_fp32_thread_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
            if (e != cudaSuccess) return (int)e;
            attr2 = true;
        }
        const long long chunk = 512;
        long long want = (n + chunk - 1) / chunk;
        const int grid = (int)(want < sms ? want : sms);
        mlp_fp32_thread_kernel<2><<<grid, 256, smem, stream>>>(m->d_blob, m->blob_bytes, m->n_sq, d_x, d_y, n, lp);
    }
    return (int)cudaGetLastError();
}

#endif  // NSDF_DEV_SCHEDULES

// ---- any other dense stack, one fused launch (mlp_generic.cuh) -------------------------------------------------------
__global__ void __launch_bounds__(128) mlp_generic_kernel(const __grid_constant__ GenericNet g, const float* __restrict__ xin,
                                                          float* __restrict__ yout, long long n) {
    extern __shared__ __align__(16) float s_params[];
    const float* P = generic_stage(g, s_params);
    const int in0 = g.dims[0], outL = g.dims[g.n_layers];
    float a[GENERIC_MAX_WIDTH], z[GENERIC_MAX_WIDTH];
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        for (int k = 0; k < in0; ++k) a[k] = __ldg(xin + i * in0 + k);
        mlp_generic_eval(g, P, a, z);
        for (int o = 0; o < outL; ++o) yout[i * outL + o] = a[o];
    }
}

void generic_net_of(const nsdf_model* m, GenericNet* g) {
    g->params = m->d_params;
    g->n_layers = m->n_layers;
    g->dims[0] = m->in[0];
    for (int l = 0; l < m->n_layers; ++l) {
        g->dims[l + 1] = m->out[l];
        g->w_off[l] = (unsigned int)m->w_off[l];
        g->b_off[l] = (unsigned int)m->b_off[l];
    }
    g->n_params = (unsigned int)m->h_params.size();
    g->staged = (size_t)g->n_params * sizeof(float) <= GENERIC_SMEM_LIMIT ? 1 : 0;
}

int generic_forward(nsdf_model* m, const float* d_x, float* d_y, int64_t n, cudaStream_t stream) {
    if (!m->generic) return NSDF_EUNSUPPORTED;
    if (n == 0) return NSDF_OK;
    GenericNet g;
    generic_net_of(m, &g);
    const size_t smem = g.staged ? (size_t)g.n_params * sizeof(float) : 0;
    cudaError_t e = cudaFuncSetAttribute(mlp_generic_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)GENERIC_SMEM_LIMIT);
    if (e != cudaSuccess) return (int)e;
    const int sms = m->sm_count > 0 ? m->sm_count : 148;
    const long long want = (n + 127) / 128;
    const long long cap = (long long)sms * (smem > 100 * 1024 ? 1 : 4);
    mlp_generic_kernel<<<(unsigned int)(want < cap ? want : cap), 128, smem, stream>>>(g, d_x, d_y, (long long)n);
    return (int)cudaGetLastError();
}

int dense_layer_launch(const float* d_W, const float* d_A, float* d_Z, const float* d_b, int M, int K,
                       int64_t num_batches, int activation, cudaStream_t stream) {
    if (num_batches == 0) return NSDF_OK;
    const long long total = (long long)num_batches * M;
    long long blocks = (total + 255) / 256;
    if (blocks > 148LL * 32) blocks = 148LL * 32;
    dense_layer_kernel<<<(int)blocks, 256, 0, stream>>>(d_W, d_A, d_Z, d_b, M, K, (long long)num_batches, activation);
    return (int)cudaGetLastError();
}

}  // namespace nsdf
