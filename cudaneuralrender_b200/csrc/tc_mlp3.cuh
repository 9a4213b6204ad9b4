// =====================================================================================
// tc_mlp3.cuh -- warp-specialised dense-query kernel for the tensor-core rungs (v3).
//
// Same operands and arithmetic as tc_mlp.cuh (every layer a tcgen05 UMMA with the bias folded
// into K, activations resident in TMEM), different schedule.  v2 ran 8 independent 128-thread
// CTAs per SM, each a strictly serial chain  sync -> issue -> wait MMA -> epilogue  per layer;
// ncu showed the tensor pipe 33 % active with a third of all stall samples on the MMA-completion
// wait (profiles/r01_ncu_full_mlp_tc_fp16_v2.txt).  v3 is ONE persistent CTA per SM:
//
//   * NM MMA-issuer warps: one elected lane waits for "A operand ready" (mbarrier bar_a[slot]),
//     issues the UMMAs of that slot's next layer and commits them to bar_d[slot].
//   * GW epilogue warp-groups of 4 warps (one per TMEM lane quadrant).  A group owns T tile slots
//     and walks them in turn: wait bar_d[slot] -> tcgen05.ld D -> cvt.rn.relu.f16x2 -> tcgen05.st A
//     -> arrive bar_a[slot].  While it converts slot t, the UMMAs of its other slots are in flight.
//   * No __syncthreads in the steady state; GW*T (up to 10) tiles of 128 queries in flight per SM in
//     GW*T*SLOT + 8 TMEM columns (the constant ones block is shared by all slots).
//
// Replaces NeuralNetwork::forward (src/neuralNetwork.cpp:54-63) for the tensor-core rungs.
// =====================================================================================
#pragma once
#include "lattice.cuh"
#include "tc_mlp.cuh"

namespace nsdf {
namespace tc {

template <int MODE>
struct Cols3 {
    static constexpr uint32_t D = 0;       // 32 fp32 accumulator columns
    static constexpr uint32_t A_HI = 32;   // 16 columns (K = 32 packed 16-bit)
    static constexpr uint32_t A_LO = 48;   // 16 columns, FP16X3 only
    static constexpr uint32_t SLOT = (MODE == MODE_FP16X3) ? 64 : 48;
};

__host__ __device__ constexpr uint32_t pow2_cols(uint32_t need) { return need <= 32 ? 32 : need <= 64 ? 64 : need <= 128 ? 128 : need <= 256 ? 256 : 512; }

struct Tc3Params {
    const uint8_t* image;
    uint32_t image_bytes;
    int n_sq;
    const float* x;
    float* y;
    long long n;
    LatticeParams lp;
    long long* trace;   // development (NSDF_TC_TRACE): clock64 stamps of CTA 0, issuer 0 at [0,512), epilogue warp 0 at [512,1024)
};

// bounded mbarrier wait: the hardware try_wait suspends the thread for a while by itself, so the loop
// body is short; a lost arrival traps instead of hanging the device.
__device__ __forceinline__ void mbar_wait_spin(uint32_t bar_addr, uint32_t parity) {
    uint32_t done, spins = 0;
    long long t0 = 0;
    for (;;) {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done)
                     : "r"(bar_addr), "r"(parity)
                     : "memory");
        if (done) return;
        if ((++spins & 255u) == 0u) {            // look at the clock rarely: the loop body stays 3 instructions
            const long long now = clock64();
            if (t0 == 0) t0 = now;
            else if (now - t0 > 4000000000LL) __trap();   // ~2 s: a lost arrival, not a slow one
        }
    }
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar_addr) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar_addr) : "memory");
}
__device__ __forceinline__ void mma_commit_addr(uint32_t bar_addr) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar_addr) : "memory");
}

__device__ __forceinline__ void lattice_point_fast(const LatticeParams& lp, long long idx, float (&x)[3]) {
    if ((lp.res & (lp.res - 1)) == 0) {   // power of two: no 64-bit divisions
        const int sh = __ffs(lp.res) - 1;
        const unsigned long long u = (unsigned long long)idx;
        const int ix = (int)(u & (unsigned)(lp.res - 1)), iy = (int)((u >> sh) & (unsigned)(lp.res - 1)), iz = (int)(u >> (2 * sh));
        const float s = __fdiv_rn(lp.hi - lp.lo, (float)lp.res);
        x[0] = fmaf((float)ix + 0.5f, s, lp.lo);
        x[1] = fmaf((float)iy + 0.5f, s, lp.lo);
        x[2] = fmaf((float)iz + 0.5f, s, lp.lo);
    } else {
        lattice_point(lp, idx, x);
    }
}

template <int MODE>
__device__ __forceinline__ void store_query3(uint32_t t_slot, const float (&x)[3]) {
    const uint32_t h01 = pack2<MODE>(x[0], x[1]);
    const uint32_t h2_ = pack2<MODE>(x[2], 0.0f);
    const float l0 = x[0] - unpack_lo<MODE>(h01), l1 = x[1] - unpack_hi<MODE>(h01), l2 = x[2] - unpack_lo<MODE>(h2_);
    const uint32_t one = ones_word<MODE>();
    uint32_t r[8];
    r[0] = h01;                                                      // k0 k1  = xh yh
    r[1] = (h2_ & 0xffffu) | (pack2<MODE>(0.0f, l0) & 0xffff0000u);  // k2 k3  = zh xl
    r[2] = pack2<MODE>(l1, l2);                                      // k4 k5  = yl zl
    r[3] = h01;                                                      // k6 k7  = xh yh
    r[4] = (h2_ & 0xffffu) | (one & 0xffff0000u);                    // k8 k9  = zh 1
    r[5] = one & 0xffffu;                                            // k10    = 1
    r[6] = 0u;
    r[7] = 0u;
    tmem_st8(t_slot + Cols3<MODE>::A_HI, r);
}

// D -> relu -> 16-bit A operand of the next layer (this thread's TMEM lane of one slot)
template <int MODE>
__device__ __forceinline__ void epilogue3(uint32_t t_slot) {
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
}

template <int MODE, int GW, int T, int NM>
__global__ void __launch_bounds__((GW * 4 + NM) * 32, 1) mlp_tc3_kernel(const __grid_constant__ Tc3Params p) {
    constexpr int NSLOT = GW * T;
    constexpr uint32_t SLOT = Cols3<MODE>::SLOT;
    constexpr uint32_t ONES_COL = NSLOT * SLOT;
    constexpr uint32_t TMEM_COLS = pow2_cols(ONES_COL + 8);
    static_assert(ONES_COL + 8 <= 512, "too many tile slots for TMEM");
    static_assert(NSLOT <= 16, "phase bits");
    constexpr int NTHREADS = (GW * 4 + NM) * 32;
    constexpr int NMMA = (MODE == MODE_FP16X3) ? 7 : 3;

    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ uint64_t s_bar_w, s_bar_d[NSLOT], s_bar_a[NSLOT];
    __shared__ uint32_t s_tmem;
    __shared__ __align__(16) uint64_t s_desc[(FUSED_MAX_SQ + 2) * 8];

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (tid == 0) {
        mbar_init(&s_bar_w, 1);
        for (int s = 0; s < NSLOT; ++s) {
            mbar_init(&s_bar_d[s], 1);
            mbar_init(&s_bar_a[s], 4);
        }
        mbar_fence_init();
    }
    tile_build_desc<MODE>(s_desc, smem_u32(smem), p.n_sq, tid, NTHREADS);
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
    if (warp < 4) {   // the shared ones block, one row per TMEM lane
        const uint32_t r[8] = {ones_word<MODE>(), 0u, 0u, 0u, 0u, 0u, 0u, 0u};
        tmem_st8(tmem + ((uint32_t)(warp * 32) << 16) + ONES_COL, r);
        wait_st();
    }
    mbar_wait_guard(&s_bar_w, 0);
    fence_before_sync();
    __syncthreads();
    fence_after_sync();

#ifdef NSDF_TC_TRACE
    long long* trace = nullptr;
    int ti = 0;
    if (p.trace && blockIdx.x == 0 && lane == 0) {
        if (warp == 0) trace = p.trace + 512;
        if (warp == GW * 4) trace = p.trace;
    }
#define TC3_STAMP() do { if (trace && ti < 500) trace[ti++] = clock64(); } while (0)
#else
#define TC3_STAMP() do { } while (0)
#endif
    const long long n_tiles = (p.n + 127) >> 7;
    const long long stride = (long long)gridDim.x * NSLOT;
    const long long tile0 = (long long)blockIdx.x * NSLOT;   // tile of (round r, slot s) = tile0 + s + r * stride
    const int n_sq = p.n_sq;
    const uint32_t bar_d0 = smem_u32(&s_bar_d[0]), bar_a0 = smem_u32(&s_bar_a[0]);

    if (warp < GW * 4) {
        // ------------------------------------------------------------------ epilogue warps
        const int gw = warp >> 2, quad = warp & 3;
        const uint32_t t_lane = tmem + ((uint32_t)(quad * 32) << 16);
        const int row_in_tile = quad * 32 + lane;
        uint32_t phase = 0;
        float xq[T][3];

        auto load_query = [&](long long tile, float (&x)[3]) {
            const long long row = (tile << 7) + row_in_tile;
            x[0] = x[1] = x[2] = 0.0f;
            if (row < p.n) {
                if (p.lp.enabled) lattice_point_fast(p.lp, p.lp.first + row, x);
                else { x[0] = __ldg(p.x + 3 * row); x[1] = __ldg(p.x + 3 * row + 1); x[2] = __ldg(p.x + 3 * row + 2); }
            }
        };
        auto signal_a = [&](int slot) {
            wait_st();
            fence_before_sync();
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_a0 + 8u * (uint32_t)slot);
        };

#pragma unroll
        for (int t = 0; t < T; ++t) {
            const int slot = gw * T + t;
            const long long tile = tile0 + slot;
            if (tile < n_tiles) {
                load_query(tile, xq[t]);
                store_query3<MODE>(t_lane + slot * SLOT, xq[t]);
                signal_a(slot);
            }
        }
        for (long long base = tile0; base < n_tiles; base += stride) {
#pragma unroll 1
            for (int step = 0; step <= n_sq + 1; ++step) {
                const bool last = step == n_sq + 1;
#pragma unroll
                for (int t = 0; t < T; ++t) {
                    const int slot = gw * T + t;
                    const long long tile = base + slot;
                    if (tile >= n_tiles) continue;
                    const uint32_t t_slot = t_lane + slot * SLOT;
                    const bool more = tile + stride < n_tiles;
                    if (last && more) load_query(tile + stride, xq[t]);   // global loads fly during the wait
                    mbar_wait_spin(bar_d0 + 8u * (uint32_t)slot, (phase >> t) & 1u);
                    phase ^= 1u << t;
                    fence_after_sync();
                    TC3_STAMP();
                    if (!last) {
                        epilogue3<MODE>(t_slot);
                        signal_a(slot);
                        TC3_STAMP();
                    } else {
                        const float y = __uint_as_float(tmem_ld1(t_slot + Cols3<MODE>::D));
                        wait_ld();
                        const long long row = (tile << 7) + row_in_tile;
                        if (row < p.n) p.y[row] = y;
                        if (more) {
                            store_query3<MODE>(t_slot, xq[t]);
                            signal_a(slot);
                        }
                        TC3_STAMP();
                    }
                }
            }
        }
    } else {
        // ------------------------------------------------------------------ MMA issuer warps (converged; one elected lane issues)
        const int m = warp - GW * 4;
        constexpr uint32_t ID32 = make_idesc(MODE == MODE_BF16 ? 1 : 0, 128, 32);
        constexpr uint32_t ID16 = make_idesc(MODE == MODE_BF16 ? 1 : 0, 128, 16);
        uint32_t phase = 0;
        for (long long base = tile0; base < n_tiles; base += stride) {
#pragma unroll 1
            for (int step = 0; step <= n_sq + 1; ++step) {
                const uint32_t idesc = step > n_sq ? ID16 : ID32;
                uint64_t d[NMMA];
#pragma unroll
                for (int j = 0; j < NMMA; ++j) d[j] = s_desc[step * 8 + j];
#pragma unroll
                for (int t = 0; t < T; ++t) {
#pragma unroll
                    for (int g = 0; g < GW; ++g) {
                        const int slot = g * T + t;
                        if ((slot % NM) != m) continue;
                        if (base + slot >= n_tiles) continue;
                        mbar_wait_spin(bar_a0 + 8u * (uint32_t)slot, (phase >> slot) & 1u);
                        phase ^= 1u << slot;
                        TC3_STAMP();
                        if (elect_one()) {
                            fence_after_sync();
                            const uint32_t ts = tmem + slot * SLOT;
                            const uint32_t tD = ts + Cols3<MODE>::D;
                            mma_ts(tD, ts + Cols3<MODE>::A_HI, d[0], idesc, 0u);
                            if (step > 0) {
                                mma_ts(tD, ts + Cols3<MODE>::A_HI + 8, d[1], idesc, 1u);
                                mma_ts(tD, tmem + ONES_COL, d[2], idesc, 1u);
                                if constexpr (MODE == MODE_FP16X3) {
                                    mma_ts(tD, ts + Cols3<MODE>::A_LO, d[3], idesc, 1u);
                                    mma_ts(tD, ts + Cols3<MODE>::A_LO + 8, d[4], idesc, 1u);
                                    mma_ts(tD, ts + Cols3<MODE>::A_HI, d[NMMA - 2], idesc, 1u);
                                    mma_ts(tD, ts + Cols3<MODE>::A_HI + 8, d[NMMA - 1], idesc, 1u);
                                }
                            }
                            mma_commit_addr(bar_d0 + 8u * (uint32_t)slot);
                        }
                        __syncwarp();
                        TC3_STAMP();
                    }
                }
            }
        }
    }

    fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem, TMEM_COLS);
}

}  // namespace tc
}  // namespace nsdf
