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

// Per-CTA-slot state of the evaluator.  All 128 threads of the slot hold the same values except t_row.
struct TileCtx {
    uint32_t tmem;        // TMEM base of this slot (lane 0)
    uint32_t t_row;       // tmem + (lane quadrant of this warp << 16)
    uint32_t img;         // shared-memory address of the weight image
    uint64_t* bar;        // MMA-complete mbarrier of this slot
    uint32_t phase;       // its running parity
    int n_sq;
    int dbg;              // development knobs (NSDF_TC_DBG): 1 no LDTM, 2 no STTM, 4 no MMA, 8 one MMA per layer
};

// Writes the constant ones block of this thread's row.  Call once after TMEM allocation.
template <int MODE>
__device__ __forceinline__ void tile_init_ones(const TileCtx& c) {
    const uint32_t r[8] = {ones_word<MODE>(), 0u, 0u, 0u, 0u, 0u, 0u, 0u};
    tmem_st8(c.t_row + Cols<MODE>::ONES, r);
    wait_st();
}

// Issues the MMAs of one layer and commits them to the slot barrier.  ONE thread.
//   layer: -1 = layer 0, 0..n_sq-1 = hidden, n_sq = last.
template <int MODE>
__device__ __forceinline__ void tile_issue_layer(const TileCtx& c, int layer) {
    constexpr uint32_t ID32 = make_idesc(MODE == MODE_BF16 ? 1 : 0, 128, 32);
    constexpr uint32_t ID16 = make_idesc(MODE == MODE_BF16 ? 1 : 0, 128, 16);
    const ImageLayout L = image_layout(c.n_sq, MODE);
    const uint32_t tD = c.tmem + Cols<MODE>::D, tA = c.tmem + Cols<MODE>::A_HI, tL = c.tmem + Cols<MODE>::A_LO;
    fence_after_sync();
    if (c.dbg & 4) { mma_commit(c.bar); return; }
    if (layer < 0) {
        mma_ts(tD, tA, make_bdesc(c.img + L.b0), ID32, 0u);
    } else if (layer < c.n_sq) {
        const uint32_t bh = c.img + L.hid_hi + (uint32_t)layer * IMG_HID_HI_BYTES;
        mma_ts(tD, tA, make_bdesc(bh), ID32, 0u);
        if (c.dbg & 8) { mma_commit(c.bar); return; }
        mma_ts(tD, tA + 8, make_bdesc(bh + 1024), ID32, 1u);
        mma_ts(tD, tA + 16, make_bdesc(bh + 2048), ID32, 1u);           // ones block x bias rows
        if (MODE == MODE_FP16X3) {
            const uint32_t bl = c.img + L.hid_lo + (uint32_t)layer * IMG_HID_LO_BYTES;
            mma_ts(tD, tL, make_bdesc(bh), ID32, 1u);
            mma_ts(tD, tL + 8, make_bdesc(bh + 1024), ID32, 1u);
            mma_ts(tD, tA, make_bdesc(bl), ID32, 1u);
            mma_ts(tD, tA + 8, make_bdesc(bl + 1024), ID32, 1u);
        }
    } else {
        const uint32_t bh = c.img + L.last_hi;
        mma_ts(tD, tA, make_bdesc(bh, LAST_LBO), ID16, 0u);
        mma_ts(tD, tA + 8, make_bdesc(bh + 512, LAST_LBO), ID16, 1u);
        mma_ts(tD, tA + 16, make_bdesc(bh + 1024, LAST_LBO), ID16, 1u);
        if (MODE == MODE_FP16X3) {
            const uint32_t bl = c.img + L.last_lo;
            mma_ts(tD, tL, make_bdesc(bh, LAST_LBO), ID16, 1u);
            mma_ts(tD, tL + 8, make_bdesc(bh + 512, LAST_LBO), ID16, 1u);
            mma_ts(tD, tA, make_bdesc(bl, LAST_LBO), ID16, 1u);
            mma_ts(tD, tA + 8, make_bdesc(bl + 512, LAST_LBO), ID16, 1u);
        }
    }
    mma_commit(c.bar);
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
template <int MODE>
__device__ __forceinline__ void tile_epilogue(const TileCtx& c) {
    uint32_t v[32];
    if (c.dbg & 1) {
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = c.phase + j;
    } else {
        tmem_ld32(c.t_row + Cols<MODE>::D, v);
        wait_ld();
    }
    uint32_t ahi[16];
    if (MODE != MODE_FP16X3) {
#pragma unroll
        for (int j = 0; j < 16; ++j) ahi[j] = pack2_relu<MODE>(__uint_as_float(v[2 * j]), __uint_as_float(v[2 * j + 1]));
        if (!(c.dbg & 2)) tmem_st16(c.t_row + Cols<MODE>::A_HI, ahi);
        else if (ahi[0] == 0x12345u && ahi[7] == 0x777u && ahi[15] == 3u) __trap();
    } else {
        uint32_t alo[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) {
            const float h0 = __uint_as_float(v[2 * j]), h1 = __uint_as_float(v[2 * j + 1]);
            ahi[j] = pack_relu_rz_f16(h0, h1);                 // hi <= h for h > 0, hi = 0 for h <= 0
            alo[j] = pack_relu_f16(h0 - f16lo_to_f32(ahi[j]), h1 - f16hi_to_f32(ahi[j]));   // >= 0 where h > 0
        }
        tmem_st16(c.t_row + Cols<MODE>::A_HI, ahi);
        tmem_st16(c.t_row + Cols<MODE>::A_LO, alo);
    }
    wait_st();
}

}  // namespace tc
}  // namespace nsdf
