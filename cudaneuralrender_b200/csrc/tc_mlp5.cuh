// =====================================================================================
// tc_mlp5.cuh -- dense-query kernel of the fp16 / bf16 tensor-core rungs, schedule v5.
//
// What round 2's microbenchmarks said (profiles/r02_tmem_mma_bench.txt): the TMEM read port
// (1.3-2.0 KB/clk/SM) and the UMMA issue path (48 clk per tile-layer once >= 9 chains are in flight) are far from
// binding; the rate is (tiles in flight per SM) / (serial chain of a tile-layer), and TMEM capacity allows ten
// 48-column tile slots per SM where v2 (tc_mlp.cuh + mlp_tc_kernel) holds eight 64-column ones.  v4 (tc_mlp4.cuh) had the
// ten slots but a longer chain (a 32-column bias store per layer, two arrivals per step).  v5 = v4's structure with v2's
// step:
//   * G tile groups of 4 warps per CTA (one named barrier per group), slots of 48 columns (32 accumulator + 16 A-operand),
//     CPS CTAs per SM: 5 x 2 = ten tiles of 128 queries in flight per SM;
//   * the bias rides in the MMA as in v2 (third UMMA, K = 16, against a ones block) -- ONE ones block per CTA, shared by
//     its slots (the UMMA reads it as the A operand; nothing ever writes it again);
//   * one issuing warp per group (3 UMMAs + 1 commit, one arrival); the same warp alone polls the accumulator mbarrier and a
//     group barrier releases the other three; epilogue in two 16-column halves (48-register budget of 1280 resident threads).
// Measured (256^3 lattice, fp16): v2 35.5-35.9 Gq/s; v5 with 2 groups x 4 CTAs 41.0, 4 x 2 38.6, 5 x 2 (ten tiles) 38.8 -- more tiles
// in flight do not help; compiled-out experiments (-DNSDF_TC5_DIAG, tools/tc5_diag.py, profiles/r02_tc5_diag.txt): without
// conversion 38.8 (of 39.0), without tcgen05.ld 38.5, without tcgen05.st 42.7, without any epilogue 48.8, and with a single
// poller, no fences and two UMMAs 57.7: the step's synchronisation skeleton, not a pipe, is what bounds the rung.
// Arithmetic exactly as tc_mlp.cuh (same weight image, same conversions): results equal v2's bit for bit.
// Replaces NeuralNetwork::forward (src/neuralNetwork.cpp:54-63).
// =====================================================================================
#pragma once
#include "tc_mlp4.cuh"

namespace nsdf {
namespace tc {

template <int MODE, int G, int CPS>
__global__ void __launch_bounds__(G * 128, CPS) mlp_tc5_kernel(const __grid_constant__ Tc4Params p) {
    static_assert(MODE != MODE_FP16X3, "v5 carries the 16-bit rungs only");
    constexpr uint32_t SLOT = 48;
    constexpr uint32_t ONES_COL = G * SLOT;
    constexpr uint32_t TMEM_COLS = pow2_cols(ONES_COL + 8);
    static_assert(TMEM_COLS * CPS <= 512, "TMEM over-subscribed");
    constexpr int NTHREADS = G * 128;

    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ uint64_t s_bar_w, s_bar_d[G];
    __shared__ uint32_t s_tmem;
    __shared__ __align__(16) uint64_t s_desc[(FUSED_MAX_SQ + 2) * 8];

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int n_sq = p.n_sq;
#ifdef NSDF_TC5_DIAG     // development: timing experiments that leave parts of the step out (results are wrong under it)
    constexpr int diag = NSDF_TC5_DIAG;
#else
    constexpr int diag = 0;
#endif
    if (tid == 0) {
        mbar_init(&s_bar_w, 1);
        for (int s = 0; s < G; ++s) mbar_init(&s_bar_d[s], 1);
        mbar_fence_init();
    }
    tile_build_desc<MODE>(s_desc, smem_u32(smem), n_sq, tid, NTHREADS);
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

    const int g = warp >> 2, q = warp & 3;
    const uint32_t t_grp = tmem + (uint32_t)g * SLOT;                      // lane 0 of this group's slot (UMMA addresses)
    const uint32_t t_slot = t_grp + ((uint32_t)(q * 32) << 16);           // this warp's lane quadrant
    const uint32_t t_ones = tmem + ONES_COL;
    if (g == 0) {          // the CTA's ones block: [1 1 0 ...] in every row, written once
        const uint32_t r[8] = {ones_word<MODE>(), 0u, 0u, 0u, 0u, 0u, 0u, 0u};
        tmem_st8(t_ones + ((uint32_t)(q * 32) << 16), r);
        wait_st();
    }
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    mbar_wait_guard(&s_bar_w, 0);

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
#pragma unroll 1
        for (int step = 0; step <= n_sq + 1; ++step) {
            if (!(diag & 16)) fence_before_sync();
            if (!(diag & 32)) group_sync(g);
            if (q == 0) {
                if (elect_one()) {
                    if (!(diag & 16)) fence_after_sync();
                    const uint64_t* d = s_desc + step * 8;
                    const uint32_t idesc = step > n_sq ? ID16 : ID32;
                    const uint32_t tD = t_grp + Cols3<MODE>::D;
                    mma_ts(tD, t_grp + Cols3<MODE>::A_HI, d[0], idesc, 0u);
                    if (step > 0) {
                        mma_ts(tD, t_grp + Cols3<MODE>::A_HI + 8, d[1], idesc, 1u);
                        if (!(diag & 8)) mma_ts(tD, t_ones, d[2], idesc, 1u);
                    }
                    mma_commit_addr(bar_d);
                }
                __syncwarp();
            }
            // ONE polling warp per tile; the group barrier releases the other three (40 warps polling mbarriers cost the
            // step's skeleton 9 %: 48.8 -> 53.2 Gq/s with the epilogue compiled out, 39.0 -> 41.0 with it)
            if ((diag & 64) || q == 0) mbar_wait_spin(bar_d, phase);
            group_sync(g);
            phase ^= 1u;
            if (!(diag & 16)) fence_after_sync();
            if (step <= n_sq) {
                if (G * 128 * CPS <= 1024 && diag == 0) {      // 64 registers per thread: one pass over the 32 accumulator columns
                    uint32_t v[32];
                    tmem_ld32(t_slot + Cols3<MODE>::D, v);
                    wait_ld();
                    uint32_t a[16];
#pragma unroll
                    for (int j = 0; j < 16; ++j) a[j] = pack2_relu<MODE>(__uint_as_float(v[2 * j]), __uint_as_float(v[2 * j + 1]));
                    tmem_st16(t_slot + Cols3<MODE>::A_HI, a);
                    wait_st();
                } else {
    #pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        uint32_t v[16];
                        if (!(diag & 4)) {
                            tmem_ld16(t_slot + Cols3<MODE>::D + 16 * h, v);
                            wait_ld();
                        } else {
    #pragma unroll
                            for (int j = 0; j < 16; ++j) v[j] = (uint32_t)(step + j);
                        }
                        uint32_t a[8];
                        if (!(diag & 1)) {
    #pragma unroll
                            for (int j = 0; j < 8; ++j) a[j] = pack2_relu<MODE>(__uint_as_float(v[2 * j]), __uint_as_float(v[2 * j + 1]));
                        } else {
    #pragma unroll
                            for (int j = 0; j < 8; ++j) a[j] = v[2 * j] ^ v[2 * j + 1];
                        }
                        if (!(diag & 2)) tmem_st8(t_slot + Cols3<MODE>::A_HI + 8 * h, a);
                        else if (a[0] == 0x12345u && a[7] == 0x54321u) p.y[0] = 1.0f;
                    }
                    if (!(diag & 2)) wait_st();
                }
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

#ifdef NSDF_DEV_SCHEDULES
// ---- schedule v6: v5 with TWO tile slots per warpgroup (ping-pong); development builds only: measured SLOWER than v5 ----------------
// (fp16, 256^3: ten tiles with 640 threads 34.3 Gq/s, eight tiles with 512 threads 29.5-31.0; v5 41.6): the chain of a tile-layer is
// mostly the warpgroup's own serial work -- barrier, tcgen05.ld, convert, tcgen05.st, fence, barrier, issue: ~560 clk -- not the UMMA
// round trip, so halving the threads per tile in flight costs more than the overlap gains.
// A tile needs four warps (one per TMEM lane quadrant), so ten tiles in flight cost v5 1280 threads and the two-pass epilogue.  Here a
// warpgroup owns two 48-column slots and alternates between them: while the UMMAs of one slot are in flight (~400 clk round trip) it
// runs the other slot's epilogue and issue.  G warpgroups x CPS CTAs per SM hold 2 * G * CPS tiles: 5 x 1 = ten tiles with 640 threads.
// Arithmetic, weight image and results as v5 / v2.
template <int MODE, int G, int CPS>
__global__ void __launch_bounds__(G * 128, CPS) mlp_tc6_kernel(const __grid_constant__ Tc4Params p) {
    static_assert(MODE != MODE_FP16X3, "v6 carries the 16-bit rungs only");
    constexpr uint32_t SLOT = 48;
    constexpr uint32_t ONES_COL = 2 * G * SLOT;
    constexpr uint32_t TMEM_COLS = pow2_cols(ONES_COL + 8);
    static_assert(TMEM_COLS * CPS <= 512, "TMEM over-subscribed");
    constexpr int NTHREADS = G * 128;

    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ uint64_t s_bar_w, s_bar_d[2 * G];
    __shared__ uint32_t s_tmem;
    __shared__ __align__(16) uint64_t s_desc[(FUSED_MAX_SQ + 2) * 8];

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int n_sq = p.n_sq;
    if (tid == 0) {
        mbar_init(&s_bar_w, 1);
        for (int s = 0; s < 2 * G; ++s) mbar_init(&s_bar_d[s], 1);
        mbar_fence_init();
    }
    tile_build_desc<MODE>(s_desc, smem_u32(smem), n_sq, tid, NTHREADS);
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

    const int g = warp >> 2, q = warp & 3;
    const uint32_t lane_off = (uint32_t)(q * 32) << 16;
    const uint32_t t_ones = tmem + ONES_COL;
    if (g == 0) {          // the CTA's ones block: [1 1 0 ...] in every row, written once
        const uint32_t r[8] = {ones_word<MODE>(), 0u, 0u, 0u, 0u, 0u, 0u, 0u};
        tmem_st8(t_ones + lane_off, r);
        wait_st();
    }
    fence_before_sync();
    __syncthreads();
    fence_after_sync();
    mbar_wait_guard(&s_bar_w, 0);

    const int row_in_tile = q * 32 + lane;
    constexpr uint32_t ID32 = make_idesc(MODE == MODE_BF16 ? 1 : 0, 128, 32);
    constexpr uint32_t ID16 = make_idesc(MODE == MODE_BF16 ? 1 : 0, 128, 16);
    uint32_t phase = 0;                 // bit s = parity of slot s's accumulator barrier

    // issue of step `step` of slot s: the slot's issuing warp is warp s of the group, so the two slots' issues overlap
    auto issue = [&](int s, int step) {
        if (q == s) {
            if (elect_one()) {
                fence_after_sync();
                const uint32_t t_grp = tmem + (uint32_t)(2 * g + s) * SLOT;
                const uint64_t* d = s_desc + step * 8;
                const uint32_t idesc = step > n_sq ? ID16 : ID32;
                const uint32_t tD = t_grp + Cols3<MODE>::D;
                mma_ts(tD, t_grp + Cols3<MODE>::A_HI, d[0], idesc, 0u);
                if (step > 0) {
                    mma_ts(tD, t_grp + Cols3<MODE>::A_HI + 8, d[1], idesc, 1u);
                    mma_ts(tD, t_ones, d[2], idesc, 1u);
                }
                mma_commit_addr(smem_u32(&s_bar_d[2 * g + s]));
            }
            __syncwarp();
        }
    };

    const long long n_tiles = (p.n + 127) >> 7;
    const long long stride = (long long)gridDim.x * (2 * G);
    for (long long tile0 = ((long long)blockIdx.x * G + g) * 2; tile0 < n_tiles; tile0 += stride) {
        long long row[2];
#pragma unroll
        for (int s = 0; s < 2; ++s) {
            row[s] = ((tile0 + s) << 7) + row_in_tile;       // a tile past the end computes on zeros and stores nothing
            float x[3] = {0.0f, 0.0f, 0.0f};
            if (row[s] < p.n) {
                if (p.lp.enabled) lattice_point_fast(p.lp, p.lp.first + row[s], x);
                else { x[0] = __ldg(p.x + 3 * row[s]); x[1] = __ldg(p.x + 3 * row[s] + 1); x[2] = __ldg(p.x + 3 * row[s] + 2); }
            }
            store_query3<MODE>(tmem + (uint32_t)(2 * g + s) * SLOT + lane_off, x);
        }
        wait_st();
        fence_before_sync();
        group_sync(g);
        issue(0, 0);
        issue(1, 0);
#pragma unroll 1
        for (int step = 0; step <= n_sq + 1; ++step) {
#pragma unroll
            for (int s = 0; s < 2; ++s) {
                const uint32_t t_slot = tmem + (uint32_t)(2 * g + s) * SLOT + lane_off;
                if (q == s) mbar_wait_spin(smem_u32(&s_bar_d[2 * g + s]), (phase >> s) & 1u);     // the slot's issuing warp polls
                group_sync(g);
                phase ^= 1u << s;
                fence_after_sync();
                if (step <= n_sq) {
                    uint32_t v[32];
                    tmem_ld32(t_slot + Cols3<MODE>::D, v);
                    wait_ld();
                    uint32_t a[16];
#pragma unroll
                    for (int j = 0; j < 16; ++j) a[j] = pack2_relu<MODE>(__uint_as_float(v[2 * j]), __uint_as_float(v[2 * j + 1]));
                    tmem_st16(t_slot + Cols3<MODE>::A_HI, a);
                    wait_st();
                    fence_before_sync();
                    group_sync(g);
                    issue(s, step + 1);
                } else {
                    const float y = __uint_as_float(tmem_ld1(t_slot + Cols3<MODE>::D));
                    wait_ld();
                    if (row[s] < p.n) p.y[row[s]] = y;
                }
            }
        }
    }

    fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem, TMEM_COLS);
}

#endif  // NSDF_DEV_SCHEDULES (v6)

}  // namespace tc
}  // namespace nsdf
