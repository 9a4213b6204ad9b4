// =====================================================================================
// mlp_tc.cu -- dense-query kernel and host side of the tensor-core rungs (tcgen05 + TMEM).
//
// The device-side evaluator is tc_mlp.cuh (all nine layers as UMMAs, bias folded into K,
// activations resident in TMEM).  This file holds
//   * the weight-image builder (16-bit K-major B operands, hi/lo split for FP16X3),
//   * mlp_tc_kernel: one CTA = 128 threads = one 128-query tile at a time, persistent over
//     tiles; several CTAs are resident per SM (64 or 128 TMEM columns each) so one CTA's
//     epilogue overlaps the others' MMAs,
//   * nsdf::tc_forward, the rung dispatcher behind nsdf_mlp_forward (include/nsdf.h).
// =====================================================================================
#include <cuda_bf16.h>
#include <cuda_fp16.h>

#include <cstdlib>
#include <cstring>
#include <vector>

#include "lattice.cuh"
#include "nsdf_internal.h"
#include "tc_mlp.cuh"
#include "tc_mlp3.cuh"
#include "tc_mlp4.cuh"
#include "tc_mlp5.cuh"

namespace nsdf {

namespace {

using tc::ImageLayout;
using tc::MODE_BF16;
using tc::MODE_FP16;
using tc::MODE_FP16X3;
using tc::N_MODES;

struct TcDeviceImages {
    uint8_t* d[N_MODES];
    uint32_t bytes[N_MODES];
};

struct TcParams {
    const uint8_t* image;
    uint32_t image_bytes;
    int n_sq;
    const float* x;
    float* y;
    long long n;
    LatticeParams lp;
    long long* trace;   // development: clock64 stamps of CTA 0 (threads 0 and 96), 2 x 512 entries, or null
};

constexpr int TC_THREADS = 128;   // thread t = query row t of the tile = TMEM lane t; thread 0 also issues the MMAs

template <int MODE, int HALVES, bool ACC16 = false>
__global__ void __launch_bounds__(TC_THREADS, 512 / tc::Cols<MODE>::TOTAL) mlp_tc_kernel(const __grid_constant__ TcParams p) {
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ uint64_t s_bar_w, s_bar_d;
    __shared__ uint32_t s_tmem;
    __shared__ __align__(16) uint64_t s_desc[(FUSED_MAX_SQ + 2) * 8];
    constexpr uint32_t TMEM_COLS = tc::Cols<MODE>::TOTAL;

    const int tid = threadIdx.x, warp = tid >> 5;
    if (tid == 0) {
        mbar_init(&s_bar_w, 1);
        mbar_init(&s_bar_d, 1);
        mbar_fence_init();
    }
    tc::tile_build_desc<MODE>(s_desc, smem_u32(smem), p.n_sq, tid, TC_THREADS);
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

    tc::TileCtx c;
    c.tmem = s_tmem;
    c.t_row = c.tmem + ((uint32_t)(warp * 32) << 16);
    c.desc = s_desc;
    c.bar_d = &s_bar_d;
    c.bar_a = nullptr;
    c.phase = 0;
    c.n_sq = p.n_sq;
    tc::tile_init_ones<MODE>(c);
    tc::mbar_wait_guard(&s_bar_w, 0);

#ifdef NSDF_TC_TRACE
    long long* trace = (p.trace && blockIdx.x == 0 && (tid == 0 || tid == 96)) ? p.trace + (tid ? 512 : 0) : nullptr;
    int ti = 0;
#define TC_STAMP() do { if (trace && ti < 500) trace[ti++] = clock64(); } while (0)
#else
#define TC_STAMP() do { } while (0)
#endif

    const long long n_tiles = (p.n + 127) >> 7;
    for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const long long row = (tile << 7) + tid;
        float x[3] = {0.0f, 0.0f, 0.0f};
        if (row < p.n) {
            if (p.lp.enabled) tc::lattice_point_fast(p.lp, p.lp.first + row, x);
            else { x[0] = __ldg(p.x + 3 * row); x[1] = __ldg(p.x + 3 * row + 1); x[2] = __ldg(p.x + 3 * row + 2); }
        }
        TC_STAMP();
        tc::tile_store_query<MODE>(c, x);
#pragma unroll 1
        for (int step = 0; step <= p.n_sq + 1; ++step) {
            TC_STAMP();
            tc::fence_before_sync();
            __syncthreads();
            TC_STAMP();
            if (warp == 0) {
                if (tc::elect_one()) tc::tile_issue_step<MODE, ACC16>(c, step);
                __syncwarp();
            }
            TC_STAMP();
            tc::mbar_wait_guard(c.bar_d, c.phase);
            c.phase ^= 1u;
            tc::fence_after_sync();
            TC_STAMP();
            if (step <= p.n_sq) tc::tile_epilogue<MODE, HALVES, ACC16>(c);
        }
        const float y = __uint_as_float(tc::tmem_ld1(c.t_row + tc::Cols<MODE>::D));
        tc::wait_ld();
        if (row < p.n) p.y[row] = y;
    }
#undef TC_STAMP

    tc::fence_before_sync();
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc(c.tmem, TMEM_COLS);
}

}  // namespace
}  // namespace nsdf

#include "mlp_tc_host.inl"
