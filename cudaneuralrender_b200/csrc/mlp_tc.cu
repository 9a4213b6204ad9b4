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
    int dbg;
};

template <int MODE>
__global__ void __launch_bounds__(128) mlp_tc_kernel(const __grid_constant__ TcParams p) {
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ uint64_t s_bar_w, s_bar_mma;
    __shared__ uint32_t s_tmem;
    constexpr uint32_t TMEM_COLS = tc::Cols<MODE>::TOTAL;

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

    tc::TileCtx c;
    c.tmem = s_tmem;
    c.t_row = c.tmem + ((uint32_t)(warp * 32) << 16);
    c.img = smem_u32(smem);
    c.bar = &s_bar_mma;
    c.phase = 0;
    c.n_sq = p.n_sq;
    c.dbg = p.dbg;
    tc::tile_init_ones<MODE>(c);
    tc::mbar_wait_guard(&s_bar_w, 0);

    const long long n_tiles = (p.n + 127) >> 7;
    for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const long long row = (tile << 7) + tid;
        float x[3] = {0.0f, 0.0f, 0.0f};
        if (row < p.n) {
            if (p.lp.enabled) lattice_point(p.lp, p.lp.first + row, x);
            else { x[0] = __ldg(p.x + 3 * row); x[1] = __ldg(p.x + 3 * row + 1); x[2] = __ldg(p.x + 3 * row + 2); }
        }
        tc::tile_store_query<MODE>(c, x);
#pragma unroll 1
        for (int layer = -1; layer <= p.n_sq; ++layer) {
            tc::fence_before_sync();
            __syncthreads();
            if (tid == 0) tc::tile_issue_layer<MODE>(c, layer);
            tc::mbar_wait_guard(c.bar, c.phase);
            c.phase ^= 1u;
            tc::fence_after_sync();
            if (layer < p.n_sq) tc::tile_epilogue<MODE>(c);
        }
        const float y = __uint_as_float(tc::tmem_ld1(c.t_row + tc::Cols<MODE>::D));
        tc::wait_ld();
        if (row < p.n) p.y[row] = y;
    }

    tc::fence_before_sync();
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc(c.tmem, TMEM_COLS);
}

}  // namespace
}  // namespace nsdf

#include "mlp_tc_host.inl"
