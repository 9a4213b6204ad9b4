// =====================================================================================
// mlp_tc.cu -- tensor-core rungs of the fused SDF-MLP (tcgen05 + TMEM, sm_100a).
//
// Replaces the same nine CUTLASS SIMT GemmBatched launches of NeuralNetwork::forward
// (src/neuralNetwork.cpp:54-63 -> src/layers/denseLayer.cu:126-176) as mlp_fp32.cuh, but the
// seven 32x32 hidden contractions run on the 5th-generation tensor cores:
//
//   per CTA (128 threads = 128 queries = the M dimension of one UMMA tile)
//     * all weights are staged ONCE into shared memory by one TMA bulk copy (UBLKCP + mbarrier):
//       fp32 first/last layer + biases, and the hidden layers as 16-bit K-major B operands;
//     * layer 0 (K = 3) and the last layer (N = 1) run in fp32 on the ALU, so the query
//       coordinates and the output never see 16-bit rounding;
//     * hidden layer l:  D[128x32] (TMEM, fp32) = A[128x32] (TMEM, 16-bit) * W_l^T (smem) by two
//       tcgen05.mma.kind::f16 (K = 16 each) issued by one thread; the epilogue threads read
//       their row of D with tcgen05.ld, add the bias, apply ReLU + convert with one
//       cvt.rn.relu.f16x2 per pair and write the next layer's A operand straight back into
//       TMEM with tcgen05.st -- activations never touch shared or global memory.
//   Several CTAs are resident per SM (64 TMEM columns each), so one CTA's epilogue overlaps the
//   others' MMAs.
//
// Rungs (include/nsdf.h): TC_FP16, TC_BF16 (single pass), TC_FP16X3 (activations and weights
// split hi + lo, three MMA passes: hi*hi + lo*hi + hi*lo, fp32-grade products).
// Compiled with -fmad=false like the rest of the library; every FMA here is explicit.
// =====================================================================================
#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include <cstdlib>
#include <cstring>
#include <vector>

#include "lattice.cuh"
#include "mlp_fp32.cuh"
#include "nsdf_internal.h"
#include "tc_common.cuh"

namespace nsdf {

namespace {

enum { MODE_FP16 = 0, MODE_BF16 = 1, MODE_FP16X3 = 2, N_MODES = 3 };

// ---- weight image ------------------------------------------------------------------------
//   floats: [0,160) layer 0 (Wt0[4][32], b0[32]) | hidden biases 32*n_sq | last w[32], b, 3 pad
//   then (128-byte aligned) B_hi: n_sq * 2048 bytes, and for FP16X3 B_lo: n_sq * 2048 bytes
struct ImageLayout {
    uint32_t bias_off_f;    // floats
    uint32_t last_off_f;    // floats
    uint32_t b_off;         // bytes
    uint32_t bytes;         // total, multiple of 128
};
__host__ __device__ inline ImageLayout image_layout(int n_sq, int mode) {
    ImageLayout L;
    L.bias_off_f = FUSED_L0_FLOATS;
    L.last_off_f = L.bias_off_f + 32u * (uint32_t)n_sq;
    const uint32_t fbytes = (L.last_off_f + FUSED_LAST_FLOATS) * 4u;
    L.b_off = (fbytes + 127u) & ~127u;
    L.bytes = L.b_off + (uint32_t)n_sq * tc::B_LAYER_BYTES * (mode == MODE_FP16X3 ? 2u : 1u);
    return L;
}

struct TcDeviceImages {
    uint8_t* d[N_MODES];
    uint32_t bytes[N_MODES];
};

struct TcParams {
    const uint8_t* image;
    uint32_t image_bytes, b_off, bias_off_f, last_off_f;
    int n_sq;
    const float* x;
    float* y;
    long long n;
    LatticeParams lp;
};

template <int MODE>
__device__ __forceinline__ uint32_t pack_act(float lo, float hi) {
    if (MODE == MODE_BF16) return tc::pack_relu_bf16(lo, hi);
    return tc::pack_relu_f16(lo, hi);
}

// One CTA = one 128-query tile at a time; thread t owns row t = TMEM lane t.
template <int MODE>
__global__ void __launch_bounds__(128) mlp_tc_kernel(const __grid_constant__ TcParams p) {
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ uint64_t s_bar_w, s_bar_mma;
    __shared__ uint32_t s_tmem;
    constexpr uint32_t TMEM_COLS = 64;
    constexpr uint32_t COL_D = 0, COL_AHI = 32, COL_ALO = 48;
    constexpr uint32_t IDESC = tc::make_idesc(MODE == MODE_BF16 ? 1 : 0, 128, 32);

    const int tid = threadIdx.x, warp = tid >> 5;
    if (tid == 0) {
        mbar_init(&s_bar_w, 1);
        mbar_init(&s_bar_mma, 1);
        mbar_fence_init();
    }
    __syncthreads();
    if (tid == 0) {
        mbar_arrive_expect_tx(&s_bar_w, p.image_bytes);
        uint32_t off = 0;
        while (off < p.image_bytes) {
            const uint32_t nb = (p.image_bytes - off) < 32768u ? (p.image_bytes - off) : 32768u;
            tma_bulk_g2s(smem + off, p.image + off, nb, &s_bar_w);
            off += nb;
        }
    }
    __syncwarp();
    if (warp == 0) tc::tmem_alloc(&s_tmem, TMEM_COLS);
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    const uint32_t tmem = s_tmem;
    const uint32_t t_row = tmem + ((uint32_t)(warp * 32) << 16);   // this warp's lane quadrant
    tc::mbar_wait_guard(&s_bar_w, 0);

    const float* sf = reinterpret_cast<const float*>(smem);
    const uint32_t b_base = smem_u32(smem + p.b_off);
    const uint32_t blo_base = b_base + (uint32_t)p.n_sq * tc::B_LAYER_BYTES;
    uint32_t phase = 0;

    const long long n_tiles = (p.n + 127) >> 7;
    for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const long long row = (tile << 7) + tid;
        float x[3] = {0.0f, 0.0f, 0.0f};
        if (row < p.n) {
            if (p.lp.enabled) lattice_point(p.lp, p.lp.first + row, x);
            else { x[0] = __ldg(p.x + 3 * row); x[1] = __ldg(p.x + 3 * row + 1); x[2] = __ldg(p.x + 3 * row + 2); }
        }
        // ---- layer 0 in fp32 (reference order), packed straight into the A operand
        {
            const float4* L0 = reinterpret_cast<const float4*>(sf);
            uint32_t ahi[16], alo[16];
#pragma unroll
            for (int oc = 0; oc < 8; ++oc) {
                const float4 w0 = L0[oc], w1 = L0[8 + oc], w2 = L0[16 + oc], bb = L0[32 + oc];
                const float h0 = fmaf(w2.x, x[2], fmaf(w1.x, x[1], fmaf(w0.x, x[0], 0.0f))) + bb.x;
                const float h1 = fmaf(w2.y, x[2], fmaf(w1.y, x[1], fmaf(w0.y, x[0], 0.0f))) + bb.y;
                const float h2 = fmaf(w2.z, x[2], fmaf(w1.z, x[1], fmaf(w0.z, x[0], 0.0f))) + bb.z;
                const float h3 = fmaf(w2.w, x[2], fmaf(w1.w, x[1], fmaf(w0.w, x[0], 0.0f))) + bb.w;
                ahi[2 * oc] = pack_act<MODE>(h0, h1);
                ahi[2 * oc + 1] = pack_act<MODE>(h2, h3);
                if (MODE == MODE_FP16X3) {
                    alo[2 * oc] = tc::pack_f16(fmaxf(h0, 0.0f) - tc::f16lo_to_f32(ahi[2 * oc]),
                                               fmaxf(h1, 0.0f) - tc::f16hi_to_f32(ahi[2 * oc]));
                    alo[2 * oc + 1] = tc::pack_f16(fmaxf(h2, 0.0f) - tc::f16lo_to_f32(ahi[2 * oc + 1]),
                                                   fmaxf(h3, 0.0f) - tc::f16hi_to_f32(ahi[2 * oc + 1]));
                }
            }
            tc::tmem_st16(t_row + COL_AHI, ahi);
            if (MODE == MODE_FP16X3) tc::tmem_st16(t_row + COL_ALO, alo);
            tc::wait_st();
        }
        float y = 0.0f;
#pragma unroll 1
        for (int l = 0; l < p.n_sq; ++l) {
            tc::fence_before_sync();
            __syncthreads();
            if (tid == 0) {
                tc::fence_after_sync();
                const uint64_t d0 = tc::make_bdesc(b_base + (uint32_t)l * tc::B_LAYER_BYTES);
                const uint64_t d1 = tc::make_bdesc(b_base + (uint32_t)l * tc::B_LAYER_BYTES + tc::B_MMA_BYTES);
                tc::mma_ts(tmem + COL_D, tmem + COL_AHI, d0, IDESC, 0u);
                tc::mma_ts(tmem + COL_D, tmem + COL_AHI + 8, d1, IDESC, 1u);
                if (MODE == MODE_FP16X3) {
                    const uint64_t e0 = tc::make_bdesc(blo_base + (uint32_t)l * tc::B_LAYER_BYTES);
                    const uint64_t e1 = tc::make_bdesc(blo_base + (uint32_t)l * tc::B_LAYER_BYTES + tc::B_MMA_BYTES);
                    tc::mma_ts(tmem + COL_D, tmem + COL_ALO, d0, IDESC, 1u);
                    tc::mma_ts(tmem + COL_D, tmem + COL_ALO + 8, d1, IDESC, 1u);
                    tc::mma_ts(tmem + COL_D, tmem + COL_AHI, e0, IDESC, 1u);
                    tc::mma_ts(tmem + COL_D, tmem + COL_AHI + 8, e1, IDESC, 1u);
                }
                tc::mma_commit(&s_bar_mma);
            }
            tc::mbar_wait_guard(&s_bar_mma, phase);
            phase ^= 1u;
            tc::fence_after_sync();

            uint32_t v[32];
            tc::tmem_ld32(t_row + COL_D, v);
            tc::wait_ld();
            const float4* bias = reinterpret_cast<const float4*>(sf + p.bias_off_f + 32 * l);
            if (l + 1 < p.n_sq) {
                uint32_t ahi[16], alo[16];
#pragma unroll
                for (int oc = 0; oc < 8; ++oc) {
                    const float4 bb = bias[oc];
                    const float h0 = __uint_as_float(v[4 * oc + 0]) + bb.x, h1 = __uint_as_float(v[4 * oc + 1]) + bb.y;
                    const float h2 = __uint_as_float(v[4 * oc + 2]) + bb.z, h3 = __uint_as_float(v[4 * oc + 3]) + bb.w;
                    ahi[2 * oc] = pack_act<MODE>(h0, h1);
                    ahi[2 * oc + 1] = pack_act<MODE>(h2, h3);
                    if (MODE == MODE_FP16X3) {
                        alo[2 * oc] = tc::pack_f16(fmaxf(h0, 0.0f) - tc::f16lo_to_f32(ahi[2 * oc]),
                                                   fmaxf(h1, 0.0f) - tc::f16hi_to_f32(ahi[2 * oc]));
                        alo[2 * oc + 1] = tc::pack_f16(fmaxf(h2, 0.0f) - tc::f16lo_to_f32(ahi[2 * oc + 1]),
                                                       fmaxf(h3, 0.0f) - tc::f16hi_to_f32(ahi[2 * oc + 1]));
                    }
                }
                tc::tmem_st16(t_row + COL_AHI, ahi);
                if (MODE == MODE_FP16X3) tc::tmem_st16(t_row + COL_ALO, alo);
                tc::wait_st();
            } else {
                // last hidden layer: keep fp32, and contract with the output layer on the ALU
                const float4* LL = reinterpret_cast<const float4*>(sf + p.last_off_f);
                float acc = 0.0f;
#pragma unroll
                for (int oc = 0; oc < 8; ++oc) {
                    const float4 bb = bias[oc], w = LL[oc];
                    acc = fmaf(w.x, fmaxf(__uint_as_float(v[4 * oc + 0]) + bb.x, 0.0f), acc);
                    acc = fmaf(w.y, fmaxf(__uint_as_float(v[4 * oc + 1]) + bb.y, 0.0f), acc);
                    acc = fmaf(w.z, fmaxf(__uint_as_float(v[4 * oc + 2]) + bb.z, 0.0f), acc);
                    acc = fmaf(w.w, fmaxf(__uint_as_float(v[4 * oc + 3]) + bb.w, 0.0f), acc);
                }
                y = acc + sf[p.last_off_f + 32];
            }
        }
        if (row < p.n) p.y[row] = y;
    }

    tc::fence_before_sync();
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc(tmem, TMEM_COLS);
}

static int env_int(const char* name, int dflt) {
    const char* s = getenv(name);
    return (s && *s) ? atoi(s) : dflt;
}

void put16(std::vector<uint8_t>& img, size_t off, uint16_t v) { std::memcpy(img.data() + off, &v, 2); }

std::vector<uint8_t> build_image(const nsdf_model* m, int mode) {
    const int n_sq = m->n_sq;
    const ImageLayout L = image_layout(n_sq, mode);
    std::vector<uint8_t> img(L.bytes, 0);
    float* f = reinterpret_cast<float*>(img.data());
    const float* P = m->h_params.data();
    // layer 0: Wt0[k][o] (k < in0), b0
    for (int k = 0; k < m->in[0]; ++k)
        for (int o = 0; o < 32; ++o) f[k * 32 + o] = P[m->w_off[0] + (size_t)o * m->in[0] + k];
    for (int o = 0; o < 32; ++o) f[128 + o] = P[m->b_off[0] + o];
    for (int l = 0; l < n_sq; ++l) {
        const float* W = P + m->w_off[1 + l];
        for (int o = 0; o < 32; ++o) f[L.bias_off_f + 32 * l + o] = P[m->b_off[1 + l] + o];
        for (int n = 0; n < 32; ++n)
            for (int k = 0; k < 32; ++k) {
                const float w = W[n * 32 + k];
                const size_t off = L.b_off + (size_t)l * tc::B_LAYER_BYTES + tc::b_elem_offset(n, k);
                if (mode == MODE_BF16) {
                    const __nv_bfloat16 h = __float2bfloat16_rn(w);
                    uint16_t bits;
                    std::memcpy(&bits, &h, 2);
                    put16(img, off, bits);
                } else {
                    const __half h = __float2half_rn(w);
                    uint16_t bits;
                    std::memcpy(&bits, &h, 2);
                    put16(img, off, bits);
                    if (mode == MODE_FP16X3) {
                        const __half lo = __float2half_rn(w - __half2float(h));
                        std::memcpy(&bits, &lo, 2);
                        put16(img, off + (size_t)n_sq * tc::B_LAYER_BYTES, bits);
                    }
                }
            }
    }
    const int last = m->n_layers - 1;
    for (int k = 0; k < 32; ++k) f[L.last_off_f + k] = P[m->w_off[last] + k];
    f[L.last_off_f + 32] = P[m->b_off[last]];
    return img;
}

int ensure_images(nsdf_model* m) {
    if (m->d_tc) return NSDF_OK;
    TcDeviceImages* t = new TcDeviceImages();
    size_t total = 0;
    std::vector<std::vector<uint8_t>> imgs;
    for (int mode = 0; mode < N_MODES; ++mode) {
        imgs.push_back(build_image(m, mode));
        t->bytes[mode] = (uint32_t)imgs.back().size();
        total += imgs.back().size();
    }
    uint8_t* d = nullptr;
    cudaError_t e = cudaMalloc(&d, total);
    if (e != cudaSuccess) { delete t; return (int)e; }
    size_t off = 0;
    for (int mode = 0; mode < N_MODES; ++mode) {
        t->d[mode] = d + off;
        e = cudaMemcpy(d + off, imgs[mode].data(), imgs[mode].size(), cudaMemcpyHostToDevice);
        if (e != cudaSuccess) { cudaFree(d); delete t; return (int)e; }
        off += imgs[mode].size();
    }
    m->d_tc = t;
    m->tc_bytes = total;
    return NSDF_OK;
}

template <int MODE>
int launch(nsdf_model* m, const TcDeviceImages* t, const float* d_x, float* d_y, int64_t n, const LatticeParams& lp,
           cudaStream_t stream) {
    const ImageLayout L = image_layout(m->n_sq, MODE);
    TcParams p;
    p.image = t->d[MODE];
    p.image_bytes = L.bytes;
    p.b_off = L.b_off;
    p.bias_off_f = L.bias_off_f;
    p.last_off_f = L.last_off_f;
    p.n_sq = m->n_sq;
    p.x = d_x;
    p.y = d_y;
    p.n = n;
    p.lp = lp;
    static bool attr_done = false;
    if (!attr_done) {
        cudaError_t e = cudaFuncSetAttribute(mlp_tc_kernel<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024);
        if (e != cudaSuccess) return (int)e;
        attr_done = true;
    }
    const int sms = m->sm_count > 0 ? m->sm_count : 148;
    // resident CTAs per SM: 8 x 64 TMEM columns = the whole TMEM; shared memory is padded so that
    // no more CTAs than that are ever co-resident (tcgen05.alloc would otherwise spin)
    const int per_sm = env_int("NSDF_TC_CTAS_PER_SM", 8);
    size_t smem = L.bytes;
    const size_t min_smem = (size_t)(227 * 1024) / (size_t)(per_sm + 1) + 1;
    if (smem < min_smem) smem = min_smem;
    const long long n_tiles = (n + 127) / 128;
    const long long cap = (long long)sms * per_sm;
    const int grid = (int)(n_tiles < cap ? n_tiles : cap);
    mlp_tc_kernel<MODE><<<grid, 128, smem, stream>>>(p);
    return (int)cudaGetLastError();
}

}  // namespace

int tc_forward(nsdf_model* m, const float* d_x, float* d_y, int64_t n, int precision, bool lattice, int res, float lo,
               float hi, int64_t first, cudaStream_t stream) {
    if (n == 0) return NSDF_OK;
    if (m->n_sq < 1) return NSDF_EUNSUPPORTED;
    int rc = ensure_images(m);
    if (rc) return rc;
    const TcDeviceImages* t = static_cast<const TcDeviceImages*>(m->d_tc);
    LatticeParams lp{lattice ? 1 : 0, res, lo, hi, (long long)first};
    switch (precision) {
        case NSDF_PREC_TC_FP16: return launch<MODE_FP16>(m, t, d_x, d_y, n, lp, stream);
        case NSDF_PREC_TC_BF16: return launch<MODE_BF16>(m, t, d_x, d_y, n, lp, stream);
        case NSDF_PREC_TC_FP16X3: return launch<MODE_FP16X3>(m, t, d_x, d_y, n, lp, stream);
        default: return NSDF_EINVAL;
    }
}

void tc_release(nsdf_model* m) {
    if (m->d_tc) {
        TcDeviceImages* t = static_cast<TcDeviceImages*>(m->d_tc);
        if (t->d[0]) cudaFree(t->d[0]);
        delete t;
    }
    m->d_tc = nullptr;
    m->tc_bytes = 0;
}

}  // namespace nsdf
