// =====================================================================================
// trace_tc.cuh -- far-field pass of the mixed-precision tracer (NSDF_RENDER_TC_FARFIELD).
//
// Measured on the BASELINE frames, 75-81 % of all network evaluations of a frame are march steps
// whose step length t = tanh(nSDF) is >= 1e-3 .. 1e-4 (profiles/r01_farfield_fraction.txt): the ray
// is far from the surface and only needs a safe step, not an exact one.  This kernel marches every
// ray through that regime on the tensor cores:
//
//   CTA = 128 threads = one 128-ray tile = the 128 TMEM lanes of one tcgen05 accumulator; thread t
//   owns ray slot t for its whole life (RaySlot in registers, no HBM ray state).  Per march step the
//   tile goes through the nine UMMAs of tc_mlp.cuh (activations resident in TMEM).  Idle slots are
//   refilled from the global pixel counter by warp-ballot compaction, so tiles stay full.
//
//   t >= thr : the ray advances by t - margin (margin >= the rung's stated error bound, so it never
//              oversteps the fp32 path) -- singleMarch (src/volumeRender_kernel.cu:416-477) otherwise.
//   t <  thr : the ray is handed, at its CURRENT point, to the fp32 path through the straggler
//              queue (trace_fused.cuh); the consumer warps of trace_coop_kernel re-evaluate it in
//              fp32 and finish marching, accept test (t < 1e-6), normals and shading exactly as the
//              bit-exact tracer does.
//
// Frames of this mode are NOT bit-identical to the reference order of operations (the ray reaches the
// surface band through different points); tests/test_gpu_parity.py::test_tc_farfield_frame measures
// hit-mask mismatches and the fraction of pixels within 1/255.  Default rendering stays bit-exact.
// =====================================================================================
#pragma once
#include "render_kernels.cuh"
#include "tc_mlp.cuh"
#include "trace_fused.cuh"

namespace nsdf {

struct TraceTcParams {
    const uint8_t* image;   // tensor-core weight image of the rung (tc_mlp.cuh)
    uint32_t image_bytes;
    float thr;              // hand-over threshold on t
    float margin;           // subtracted from every tensor-core step
    int max_tc_steps;       // a ray that has marched this many steps here is handed over too: a tensor-core step has ~5x the
                            // latency of a cooperative fp32 step, long rays belong on the latency path
    int drain_below;        // once the pixels are exhausted, a warp with fewer active rays than this hands all of them over
};

template <int MODE>
__global__ void __launch_bounds__(128, 512 / tc::Cols<MODE>::TOTAL)
trace_tc_kernel(const __grid_constant__ RenderParams rp, const TraceTcParams tp) {
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ uint64_t s_bar_w, s_bar_d;
    __shared__ uint32_t s_tmem;
    __shared__ __align__(16) uint64_t s_desc[(FUSED_MAX_SQ + 2) * 8];
    constexpr uint32_t TMEM_COLS = tc::Cols<MODE>::TOTAL;

    const int tid = threadIdx.x, warp = tid >> 5;
    const unsigned int lane = tid & 31;
    if (tid == 0) {
        mbar_init(&s_bar_w, 1);
        mbar_init(&s_bar_d, 1);
        mbar_fence_init();
    }
    tc::tile_build_desc<MODE>(s_desc, smem_u32(smem), rp.n_sq, tid, 128);
    __syncthreads();
    if (tid == 0) {
        mbar_arrive_expect_tx(&s_bar_w, tp.image_bytes);
        uint32_t off = 0;
        while (off < tp.image_bytes) {
            const uint32_t nb = (tp.image_bytes - off) < 32768u ? (tp.image_bytes - off) : 32768u;
            tma_bulk_g2s(smem + off, tp.image + off, nb, &s_bar_w);
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
    c.n_sq = rp.n_sq;
    tc::tile_init_ones<MODE>(c);
    tc::mbar_wait_guard(&s_bar_w, 0);

    RaySlot slot;
    slot.mode = -1;
    LocalStats st = {0, 0, 0, 0, 0, 0};
    bool exhausted = false;   // warp-uniform

    for (;;) {
        // ---- refill idle slots (ballot compaction, one atomic per warp)
        while (!exhausted) {
            const bool need = slot.mode < 0;
            const unsigned int m = __ballot_sync(0xffffffffu, need);
            if (!m) break;
            const int leader = __ffs(m) - 1;
            const unsigned int cnt = __popc(m);
            unsigned int base = 0;
            if ((int)lane == leader) base = atomicAdd(rp.pixel_counter, cnt);
            base = __shfl_sync(0xffffffffu, base, leader);
            if (need) {
                const unsigned int idx = base + __popc(m & lanemask_lt());
                if (idx < rp.total_idx) slot_start(slot, idx, rp, st);
            }
            if (base + cnt >= rp.total_idx) exhausted = true;
        }
        const bool active = slot.mode >= 0;
        tc::fence_before_sync();
        if (!__syncthreads_or(active ? 1 : 0)) break;   // also orders the previous step's TMEM reads before the next UMMA

        // ---- one network evaluation of the tile on the tensor cores
        float x[3] = {0.0f, 0.0f, 0.0f};
        if (active) { x[0] = slot.p.x; x[1] = slot.p.y; x[2] = slot.p.z; }
        tc::tile_store_query<MODE>(c, x);
#pragma unroll 1
        for (int step = 0; step <= rp.n_sq + 1; ++step) {
            tc::fence_before_sync();
            __syncthreads();
            if (warp == 0) {
                if (tc::elect_one()) tc::tile_issue_step<MODE>(c, step);
                __syncwarp();
            }
            tc::mbar_wait_guard(c.bar_d, c.phase);
            c.phase ^= 1u;
            tc::fence_after_sync();
            if (step <= rp.n_sq) tc::tile_epilogue<MODE, 1>(c);
        }
        const float nsdf = __uint_as_float(tc::tmem_ld1(c.t_row + tc::Cols<MODE>::D));
        tc::wait_ld();

        // ---- far-field march step or hand-over
        bool hand_over = false;
        if (active) {
            ++st.queries;
            float rem_unused = -1.0f;
            const float t = scene_eval(rp, 0, slot.p, nsdf, rem_unused);
            if (t >= tp.thr) {
                const float ts = t - tp.margin;
                slot.tfar = slot.tfar - ts;
                ++slot.steps;
                if (slot.tfar <= 0.0f) {
                    rp.out[slot.pix] = 0u;
                    slot.mode = -1;
                } else {
                    slot.p = f3{fmaf(slot.d.x, ts, slot.p.x), fmaf(slot.d.y, ts, slot.p.y), fmaf(slot.d.z, ts, slot.p.z)};
                    if (slot.steps >= rp.max_steps) {
                        rp.out[slot.pix] = 0u;
                        ++st.unfinished;
                        slot.mode = -1;
                    }
                }
                if (slot.mode == 0 && slot.steps >= tp.max_tc_steps) hand_over = true;
            } else {
                hand_over = true;   // NaN lands here too: the fp32 path decides
            }
        }
        if (exhausted) {
            const int n_act = __popc(__ballot_sync(0xffffffffu, slot.mode == 0));
            if (n_act < tp.drain_below && slot.mode == 0) hand_over = true;
        }
        fused_push(slot, hand_over, rp, st);
    }
    flush_stats(st, rp.stats);

    tc::fence_before_sync();
    __syncthreads();
    if (warp == 0) tc::tmem_dealloc(c.tmem, TMEM_COLS);
}

}  // namespace nsdf
