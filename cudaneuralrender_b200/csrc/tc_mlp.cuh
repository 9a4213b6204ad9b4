// =====================================================================================
// tc_mlp.cuh -- the whole SDF-MLP on tcgen05 tensor cores, one 128-query tile per call.
//
// Replaces NeuralNetwork::forward (src/neuralNetwork.cpp:54-63; nine CUTLASS SIMT GemmBatched
// launches, src/layers/denseLayer.cu:126-176).  Every layer is a UMMA with the tile's queries
// as M = 128:
//
//   layer 0   K = 16 : A0 = [xh yh zh | xl yl zl | xh yh zh | 1 1 | 0..]  (x = xh + xl, 16-bit halves)
//                      B0 = [wh wh wh | wh wh wh | wl wl wl | bh bl | 0..] -> x*w + b to ~2^-22
//   hidden l  K = 48 : A  = [a_0..a_31 | 1 1 0..]   (the ones block is written to TMEM once)
//                      B  = [W_l[n][0..31] | bh bl 0..]   -> bias rides in the MMA, the epilogue
//                      is ONE cvt.rn.relu.f16x2 per pair of neurons
//   last      K = 48, N = 16 (column 0 used): same, no ReLU; the thread reads its scalar.
//   FP16X3 rung: activations and weights are split hi + lo and each layer adds the passes
//                      A_lo * W_hi and A_hi * W_lo (K = 32 each): fp32-grade products.
//
// Activations live in TMEM from the first layer to the last (tcgen05.ld -> cvt -> tcgen05.st);
// nothing but the query point and the result crosses the register file / HBM.
// =====================================================================================
#pragma once
#include "mlp_fp32.cuh"
#include "tc_common.cuh"

namespace nsdf {
namespace tc {

enum { MODE_FP16 = 0, MODE_BF16 = 1, MODE_FP16X3 = 2, N_MODES = 3 };

// ---- weight image (bytes; every section a multiple of 128) -----------------------------------
//   B0      : 32 rows x K16          1024
//   HID_HI  : n_sq x (32 rows x K48) 3072 each
//   LAST_HI : 16 rows x K48          1536
//   HID_LO  : n_sq x (32 rows x K32) 2048 each   (FP16X3 only)
//   LAST_LO : 16 rows x K32          1024        (FP16X3 only)
constexpr uint32_t IMG_B0_BYTES = 1024, IMG_HID_HI_BYTES = 3072, IMG_LAST_HI_BYTES = 1536, IMG_HID_LO_BYTES = 2048,
                   IMG_LAST_LO_BYTES = 1024;
constexpr uint32_t LAST_LBO = 256;   // 16 rows x 16 bytes

struct ImageLayout {
    uint32_t b0, hid_hi, last_hi, hid_lo, last_lo, bytes;
};
__host__ __device__ inline ImageLayout image_layout(int n_sq, int mode) {
    ImageLayout L;
    L.b0 = 0;
    L.hid_hi = L.b0 + IMG_B0_BYTES;
    L.last_hi = L.hid_hi + (uint32_t)n_sq * IMG_HID_HI_BYTES;
    L.hid_lo = L.last_hi + IMG_LAST_HI_BYTES;
    L.last_lo = L.hid_lo;
    L.bytes = L.hid_lo;
    if (mode == MODE_FP16X3) {
        L.last_lo = L.hid_lo + (uint32_t)n_sq * IMG_HID_LO_BYTES;
        L.bytes = L.last_lo + IMG_LAST_LO_BYTES;
    }
    L.bytes = (L.bytes + 127u) & ~127u;
    return L;
}

// TMEM columns of one tile slot
template <int MODE>
struct Cols {
    static constexpr uint32_t D = 0;        // 32 fp32 accumulator columns (last layer: 16)
    static constexpr uint32_t A_HI = 32;    // 16 columns (K = 32); layer 0 uses the first 8
    static constexpr uint32_t ONES = 48;    // 8 columns (K = 16): [1 1 0 ...], written once
    static constexpr uint32_t A_LO = 64;    // 16 columns, FP16X3 only
    static constexpr uint32_t TOTAL = (MODE == MODE_FP16X3) ? 128 : 64;   // power of two >= 32
};

template <int MODE> __device__ __forceinline__ uint32_t pack2(float lo, float hi) {
    return MODE == MODE_BF16 ? pack_bf16(lo, hi) : pack_f16(lo, hi);
}
template <int MODE> __device__ __forceinline__ uint32_t pack2_relu(float lo, float hi) {
    return MODE == MODE_BF16 ? pack_relu_bf16(lo, hi) : pack_relu_f16(lo, hi);
}
template <int MODE> __device__ __forceinline__ float unpack_lo(uint32_t p) {
    return MODE == MODE_BF16 ? __uint_as_float(p << 16) : f16lo_to_f32(p);
}
template <int MODE> __device__ __forceinline__ float unpack_hi(uint32_t p) {
    return MODE == MODE_BF16 ? __uint_as_float(p & 0xffff0000u) : f16hi_to_f32(p);
}
template <int MODE> __device__ __forceinline__ uint32_t ones_word() { return MODE == MODE_BF16 ? 0x3F803F80u : 0x3C003C00u; }

// Per-CTA-slot state of the evaluator.  All threads of the slot hold the same values except t_row.
struct TileCtx {
    uint32_t tmem;        // TMEM base of this slot (lane 0)
    uint32_t t_row;       // tmem + (lane quadrant of this warp << 16)
    const uint64_t* desc; // shared-memory table of B descriptors: [step][8], step = layer + 1
    uint64_t* bar_d;      // "accumulator ready" mbarrier (tcgen05.commit arrives)
    uint64_t* bar_a;      // "A operand ready" mbarrier (one arrival per epilogue warp)
    uint32_t phase;       // running parity of the barrier this role waits on
    int n_sq;
};

// Writes the constant ones block of this thread's row.  Call once after TMEM allocation.
template <int MODE>
__device__ __forceinline__ void tile_init_ones(const TileCtx& c) {
    const uint32_t r[8] = {ones_word<MODE>(), 0u, 0u, 0u, 0u, 0u, 0u, 0u};
    tmem_st8(c.t_row + Cols<MODE>::ONES, r);
    wait_st();
}

// MMAs of one step: count, and for MMA j the A operand column and the B descriptor.
//   step 0 = layer 0 (1 MMA), steps 1..n_sq = hidden, step n_sq+1 = last.
template <int MODE> __host__ __device__ constexpr int mmas_in_step(int step) {
    return step == 0 ? 1 : (MODE == MODE_FP16X3 ? 7 : 3);
}
template <int MODE> __host__ __device__ constexpr uint32_t mma_a_col(int j) {
    // hi k0-15, hi k16-31, ones | lo k0-15, lo k16-31 (x W_hi) | hi k0-15, hi k16-31 (x W_lo)
    return j == 0 ? Cols<MODE>::A_HI : j == 1 ? Cols<MODE>::A_HI + 8 : j == 2 ? Cols<MODE>::ONES
         : j == 3 ? Cols<MODE>::A_LO : j == 4 ? Cols<MODE>::A_LO + 8 : j == 5 ? Cols<MODE>::A_HI : Cols<MODE>::A_HI + 8;
}
// Fills the descriptor table (any number of threads; entry e = step * 8 + j).
template <int MODE>
__device__ __forceinline__ void tile_build_desc(uint64_t* table, uint32_t img, int n_sq, int tid, int nthreads) {
    const ImageLayout L = image_layout(n_sq, MODE);
    for (int e = tid; e < (n_sq + 2) * 8; e += nthreads) {
        const int step = e >> 3, j = e & 7;
        uint64_t d = 0;
        if (j < mmas_in_step<MODE>(step)) {
            if (step == 0) {
                d = make_bdesc(img + L.b0);
            } else if (step <= n_sq) {
                const uint32_t bh = img + L.hid_hi + (uint32_t)(step - 1) * IMG_HID_HI_BYTES;
                const uint32_t bl = img + L.hid_lo + (uint32_t)(step - 1) * IMG_HID_LO_BYTES;
                const uint32_t a = j < 3 ? bh + 1024u * j : j < 5 ? bh + 1024u * (j - 3) : bl + 1024u * (j - 5);
                d = make_bdesc(a);
            } else {
                const uint32_t bh = img + L.last_hi, bl = img + L.last_lo;
                const uint32_t a = j < 3 ? bh + 512u * j : j < 5 ? bh + 512u * (j - 3) : bl + 512u * (j - 5);
                d = make_bdesc(a, LAST_LBO);
            }
        }
        table[e] = d;
    }
}

// Issues the MMAs of one step and commits them to bar_d.  ONE thread (the MMA warp's lane 0).
// ACC16: layer 0 and the hidden layers accumulate in fp16 (the last layer always in fp32).
template <int MODE, bool ACC16 = false>
__device__ __forceinline__ void tile_issue_step(const TileCtx& c, int step) {
    constexpr uint32_t ID32 = make_idesc(MODE == MODE_BF16 ? 1 : 0, 128, 32, !ACC16);
    constexpr uint32_t ID16 = make_idesc(MODE == MODE_BF16 ? 1 : 0, 128, 16);
    const uint32_t idesc = step > c.n_sq ? ID16 : ID32;
    const uint64_t* d = c.desc + step * 8;
    const uint32_t tD = c.tmem + Cols<MODE>::D;
    fence_after_sync();
    mma_ts(tD, c.tmem + mma_a_col<MODE>(0), d[0], idesc, 0u);
    if (step > 0) {
#pragma unroll
        for (int j = 1; j < mmas_in_step<MODE>(1); ++j) mma_ts(tD, c.tmem + mma_a_col<MODE>(j), d[j], idesc, 1u);
    }
    mma_commit(c.bar_d);
}

// One elected arrival per epilogue warp on bar_a, after the warp's tcgen05.st have completed.
__device__ __forceinline__ void tile_signal_a(const TileCtx& c) {
    fence_before_sync();
    __syncwarp();
    if ((threadIdx.x & 31) == 0)
        asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"((uint32_t)__cvta_generic_to_shared(c.bar_a)) : "memory");
}

// This thread's query -> the layer-0 A operand (8 packed columns).
template <int MODE>
__device__ __forceinline__ void tile_store_query(const TileCtx& c, const float (&x)[3]) {
    const uint32_t h01 = pack2<MODE>(x[0], x[1]);
    const uint32_t h2_ = pack2<MODE>(x[2], 0.0f);
    const float l0 = x[0] - unpack_lo<MODE>(h01), l1 = x[1] - unpack_hi<MODE>(h01), l2 = x[2] - unpack_lo<MODE>(h2_);
    const uint32_t one = ones_word<MODE>();
    uint32_t r[8];
    r[0] = h01;                                                  // k0 k1  = xh yh
    r[1] = (h2_ & 0xffffu) | (pack2<MODE>(0.0f, l0) & 0xffff0000u);  // k2 k3  = zh xl
    r[2] = pack2<MODE>(l1, l2);                                  // k4 k5  = yl zl
    r[3] = h01;                                                  // k6 k7  = xh yh
    r[4] = (h2_ & 0xffffu) | (one & 0xffff0000u);                // k8 k9  = zh 1
    r[5] = one & 0xffffu;                                        // k10    = 1
    r[6] = 0u;
    r[7] = 0u;
    tmem_st8(c.t_row + Cols<MODE>::A_HI, r);
    wait_st();
}

// Epilogue of layer 0 / a hidden layer: D -> relu -> 16-bit A operand of the next layer.
// HALVES = 2 processes 16 accumulator columns at a time (smaller live register set).
template <int MODE, int HALVES = 1, bool ACC16 = false>
__device__ __forceinline__ void tile_epilogue(const TileCtx& c) {
    if (ACC16) {   // fp16 accumulators: packed read (half the TMEM read traffic), relu on pairs, no conversion
        uint32_t v[16];
        tmem_ld16_pack(c.t_row + Cols<MODE>::D, v);
        wait_ld();
#pragma unroll
        for (int j = 0; j < 16; ++j) v[j] = relu_f16x2(v[j]);
        tmem_st16(c.t_row + Cols<MODE>::A_HI, v);
        wait_st();
        return;
    }
    constexpr int NC = 32 / HALVES;          // accumulator columns per pass
#pragma unroll
    for (int half = 0; half < HALVES; ++half) {
        uint32_t v[NC];
        if (HALVES == 1) tmem_ld32(c.t_row + Cols<MODE>::D, reinterpret_cast<uint32_t(&)[32]>(v));
        else tmem_ld16(c.t_row + Cols<MODE>::D + 16 * half, reinterpret_cast<uint32_t(&)[16]>(v));
        wait_ld();
        uint32_t ahi[NC / 2], alo[NC / 2];
#pragma unroll
        for (int j = 0; j < NC / 2; ++j) {
            const float h0 = __uint_as_float(v[2 * j]), h1 = __uint_as_float(v[2 * j + 1]);
            if (MODE != MODE_FP16X3) {
                ahi[j] = pack2_relu<MODE>(h0, h1);
            } else {
                ahi[j] = pack_relu_rz_f16(h0, h1);                 // hi <= h for h > 0, hi = 0 for h <= 0
                alo[j] = pack_relu_f16(h0 - f16lo_to_f32(ahi[j]), h1 - f16hi_to_f32(ahi[j]));   // >= 0 where h > 0
            }
        }
        if (HALVES == 1) {
            tmem_st16(c.t_row + Cols<MODE>::A_HI, reinterpret_cast<uint32_t(&)[16]>(ahi));
            if (MODE == MODE_FP16X3) tmem_st16(c.t_row + Cols<MODE>::A_LO, reinterpret_cast<uint32_t(&)[16]>(alo));
        } else {
            tmem_st8(c.t_row + Cols<MODE>::A_HI + 8 * half, reinterpret_cast<uint32_t(&)[8]>(ahi));
            if (MODE == MODE_FP16X3) tmem_st8(c.t_row + Cols<MODE>::A_LO + 8 * half, reinterpret_cast<uint32_t(&)[8]>(alo));
        }
    }
    wait_st();
}

}  // namespace tc
}  // namespace nsdf
