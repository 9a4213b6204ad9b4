// =====================================================================================
// tc_mlp4.cuh -- dense-query kernel of the tensor-core rungs, schedule v4.
//
// What the measurements said (tools/mma_bench.cu, tools/tc3_trace.py, profiles/r01_tc_schedule_notes.md):
//   * every tcgen05.mma / tcgen05.commit costs the ISSUING WARP ~50 cycles, whatever N is, so a
//     warp that issues the three UMMAs + commit of one layer is busy ~200 cycles; one dedicated
//     issuer warp per SM (v3) caps the SM at one tile-layer per ~400 cycles;
//   * the per-layer chain of a tile (sync -> issue -> UMMA latency ~105 -> tcgen05.ld/cvt/st ~206)
//     is ~550 cycles against 32-48 cycles of tensor-pipe work, so throughput = tiles in flight / chain.
// Hence v4:
//   * G tile groups of 4 warps per CTA (named barrier per group, no __syncthreads in the steady
//     state), G * 48 TMEM columns, 2 CTAs per SM -> up to 10 tiles of 128 queries in flight per SM;
//   * the bias is no longer a third K = 16 UMMA against a ones block: the epilogue of layer l
//     pre-loads the accumulator with the fp32 bias of layer l+1 (tcgen05.st), every UMMA accumulates,
//     so the UMMAs of a layer are order-free and are issued IN PARALLEL by different warps of the
//     group (one UMMA + one commit each): ~100 cycles of issue in the chain instead of ~200, and
//     32 instead of 48 tensor-pipe cycles per layer (floor 16 + 7*32 + 16 = 256 cycles per tile);
//   * bias is exact fp32 (was fp16 hi+lo).
// Arithmetic otherwise as tc_mlp.cuh.  Replaces NeuralNetwork::forward (src/neuralNetwork.cpp:54-63).
// =====================================================================================
#pragma once
#include "tc_mlp3.cuh"

namespace nsdf {
namespace tc {

struct Tc4Params {
    const uint8_t* image;
    uint32_t image_bytes;
    const float* blob;     // fused fp32 blob (mlp_fp32.cuh): source of the fp32 biases
    int n_sq;
    const float* x;
    float* y;
    long long n;
    LatticeParams lp;
};

__device__ __forceinline__ void group_sync(int g) { asm volatile("bar.sync %0, 128;" ::"r"(g + 1) : "memory"); }

__device__ __forceinline__ void tmem_st32(uint32_t taddr, const uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
        "{%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31,%32};" ::"r"(taddr),
        "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]), "r"(r[9]),
        "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15]), "r"(r[16]), "r"(r[17]), "r"(r[18]),
        "r"(r[19]), "r"(r[20]), "r"(r[21]), "r"(r[22]), "r"(r[23]), "r"(r[24]), "r"(r[25]), "r"(r[26]), "r"(r[27]),
        "r"(r[28]), "r"(r[29]), "r"(r[30]), "r"(r[31])
        : "memory");
}

// Epilogue of step `step` (layer output in D): A(next) = relu(D) as 16-bit, then D := bias of the next layer.
// bias: shared memory, 32 floats of the next layer (warp-uniform address -> broadcast LDS.128).
template <int MODE>
__device__ __forceinline__ void epilogue4(uint32_t t_slot, const float* bias) {
    uint32_t v[32];
    tmem_ld32(t_slot + Cols3<MODE>::D, v);
    wait_ld();
    uint32_t ahi[16], alo[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) {
        const float h0 = __uint_as_float(v[2 * j]), h1 = __uint_as_float(v[2 * j + 1]);
        if (MODE != MODE_FP16X3) {
            ahi[j] = pack2_relu<MODE>(h0, h1);
        } else {
            ahi[j] = pack_relu_rz_f16(h0, h1);
            alo[j] = pack_relu_f16(h0 - f16lo_to_f32(ahi[j]), h1 - f16hi_to_f32(ahi[j]));
        }
    }
    tmem_st16(t_slot + Cols3<MODE>::A_HI, ahi);
    if (MODE == MODE_FP16X3) tmem_st16(t_slot + Cols3<MODE>::A_LO, alo);
    const float4* b4 = reinterpret_cast<const float4*>(bias);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const float4 b = b4[j];
        v[4 * j] = __float_as_uint(b.x);
        v[4 * j + 1] = __float_as_uint(b.y);
        v[4 * j + 2] = __float_as_uint(b.z);
        v[4 * j + 3] = __float_as_uint(b.w);
    }
    tmem_st32(t_slot + Cols3<MODE>::D, v);
}

// Same in two 16-column passes: fits the 48-register budget of 1280 resident threads per SM.
template <int MODE>
__device__ __forceinline__ void epilogue4_halves(uint32_t t_slot, const float* bias) {
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        uint32_t v[16];
        tmem_ld16(t_slot + Cols3<MODE>::D + 16 * h, v);
        wait_ld();
        uint32_t ahi[8], alo[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float h0 = __uint_as_float(v[2 * j]), h1 = __uint_as_float(v[2 * j + 1]);
            if (MODE != MODE_FP16X3) {
                ahi[j] = pack2_relu<MODE>(h0, h1);
            } else {
                ahi[j] = pack_relu_rz_f16(h0, h1);
                alo[j] = pack_relu_f16(h0 - f16lo_to_f32(ahi[j]), h1 - f16hi_to_f32(ahi[j]));
            }
        }
        tmem_st8(t_slot + Cols3<MODE>::A_HI + 8 * h, ahi);
        if (MODE == MODE_FP16X3) tmem_st8(t_slot + Cols3<MODE>::A_LO + 8 * h, alo);
    }
    const float4* b4 = reinterpret_cast<const float4*>(bias);
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        uint32_t v[16];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float4 b = b4[4 * h + j];
            v[4 * j] = __float_as_uint(b.x);
            v[4 * j + 1] = __float_as_uint(b.y);
            v[4 * j + 2] = __float_as_uint(b.z);
            v[4 * j + 3] = __float_as_uint(b.w);
        }
        tmem_st16(t_slot + Cols3<MODE>::D + 16 * h, v);
    }
}

template <int MODE, int G, int CPS, bool SINGLE_POLLER = false>
__global__ void __launch_bounds__(G * 128, CPS) mlp_tc4_kernel(const __grid_constant__ Tc4Params p) {
    constexpr uint32_t SLOT = Cols3<MODE>::SLOT;
    constexpr uint32_t TMEM_COLS = pow2_cols(G * SLOT);
    static_assert(TMEM_COLS * CPS <= 512, "TMEM over-subscribed");
    constexpr int NTHREADS = G * 128;
    constexpr int NISS = (MODE == MODE_FP16X3) ? 4 : 2;   // warps of a group that issue UMMAs

    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ uint64_t s_bar_w, s_bar_d[G];
    __shared__ uint32_t s_tmem;
    __shared__ __align__(16) uint64_t s_desc[(FUSED_MAX_SQ + 2) * 8];
    __shared__ __align__(16) float s_bias[(FUSED_MAX_SQ + 1) * 32];   // [l]: bias pre-loaded before step l+1

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int n_sq = p.n_sq;
    if (tid == 0) {
        mbar_init(&s_bar_w, 1);
        for (int s = 0; s < G; ++s) mbar_init(&s_bar_d[s], NISS);
        mbar_fence_init();
    }
    tile_build_desc<MODE>(s_desc, smem_u32(smem), n_sq, tid, NTHREADS);
    for (int i = tid; i < (n_sq + 1) * 32; i += NTHREADS) {
        const int l = i >> 5, o = i & 31;
        float b;
        if (l < n_sq) b = __ldg(p.blob + FUSED_L0_FLOATS + l * FUSED_SQ_FLOATS + 1024 + o);
        else b = (o == 0) ? __ldg(p.blob + FUSED_L0_FLOATS + n_sq * FUSED_SQ_FLOATS + 32) : 0.0f;
        s_bias[i] = b;
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
    if (warp == 0) tmem_alloc(&s_tmem, TMEM_COLS);
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    const uint32_t tmem = s_tmem;
    mbar_wait_guard(&s_bar_w, 0);

    const int g = warp >> 2, q = warp & 3;
    const uint32_t t_grp = tmem + (uint32_t)g * SLOT;                      // lane 0 of this group's slot (UMMA addresses)
    const uint32_t t_slot = t_grp + ((uint32_t)(q * 32) << 16);           // this warp's lane quadrant
    const uint32_t bar_d = smem_u32(&s_bar_d[g]);
    const int row_in_tile = q * 32 + lane;
    constexpr uint32_t ID32 = make_idesc(MODE == MODE_BF16 ? 1 : 0, 128, 32);
    constexpr uint32_t ID16 = make_idesc(MODE == MODE_BF16 ? 1 : 0, 128, 16);
    uint32_t phase = 0;

    const long long n_tiles = (p.n + 127) >> 7;
    const long long stride = (long long)gridDim.x * G;
    for (long long tile = (long long)blockIdx.x * G + g; tile < n_tiles; tile += stride) {
        const long long row = (tile << 7) + row_in_tile;
        float x[3] = {0.0f, 0.0f, 0.0f};
        if (row < p.n) {
            if (p.lp.enabled) lattice_point_fast(p.lp, p.lp.first + row, x);
            else { x[0] = __ldg(p.x + 3 * row); x[1] = __ldg(p.x + 3 * row + 1); x[2] = __ldg(p.x + 3 * row + 2); }
        }
        store_query3<MODE>(t_slot, x);
        wait_st();
        fence_before_sync();
        group_sync(g);
#pragma unroll 1
        for (int step = 0; step <= n_sq + 1; ++step) {
            // ---- issue: the UMMAs of this layer, spread over the first NISS warps of the group
            if (q < NISS) {
                if (elect_one()) {
                    fence_after_sync();
                    const uint64_t* d = s_desc + step * 8;
                    const uint32_t idesc = step > n_sq ? ID16 : ID32;
                    const uint32_t tD = t_grp + Cols3<MODE>::D;
                    bool issued = false;
                    if (step == 0) {
                        if (q == 0) { mma_ts(tD, t_grp + Cols3<MODE>::A_HI, d[0], idesc, 0u); issued = true; }
                    } else if (MODE != MODE_FP16X3) {
                        mma_ts(tD, t_grp + Cols3<MODE>::A_HI + 8u * (uint32_t)q, d[q], idesc, 1u);
                        issued = true;
                    } else {
                        if (q == 0) {
                            mma_ts(tD, t_grp + Cols3<MODE>::A_HI, d[0], idesc, 1u);
                            mma_ts(tD, t_grp + Cols3<MODE>::A_HI + 8, d[1], idesc, 1u);
                        } else if (q == 1) {
                            mma_ts(tD, t_grp + Cols3<MODE>::A_LO, d[3], idesc, 1u);
                            mma_ts(tD, t_grp + Cols3<MODE>::A_LO + 8, d[4], idesc, 1u);
                        } else {
                            mma_ts(tD, t_grp + Cols3<MODE>::A_HI + 8u * (uint32_t)(q - 2), d[3 + q], idesc, 1u);
                        }
                        issued = true;
                    }
                    if (issued) mma_commit_addr(bar_d);
                    else mbar_arrive(bar_d);
                }
                __syncwarp();
            }
            if (SINGLE_POLLER) {          // one warp of the group polls, the group barrier releases the other three (tc_mlp5.cuh)
                if (q == 0) mbar_wait_spin(bar_d, phase);
                group_sync(g);
            } else {
                mbar_wait_spin(bar_d, phase);
            }
            phase ^= 1u;
            fence_after_sync();
            if (step <= n_sq) {
                if (G * 128 * CPS > 1024) epilogue4_halves<MODE>(t_slot, s_bias + step * 32);
                else epilogue4<MODE>(t_slot, s_bias + step * 32);
                wait_st();
                fence_before_sync();
                group_sync(g);
            } else {
                const float y = __uint_as_float(tmem_ld1(t_slot + Cols3<MODE>::D));
                wait_ld();
                if (row < p.n) p.y[row] = y;
            }
        }
    }

    fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem, TMEM_COLS);
}

}  // namespace tc
}  // namespace nsdf
