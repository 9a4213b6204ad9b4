// =====================================================================================
// tile_priority.cuh -- start the frame's longest rays first.
//
// A frame's critical path is a handful of grazing rays that need thousands of DEPENDENT network
// evaluations (one plane_1 ray runs all 6000 steps = 7 ms at 1.17 us per step), while
// the rest of the frame is ~11 ms of throughput work.  If those rays start when the frame starts, their
// chain hides behind the bulk; if their tile comes up late, the chain is a tail.  Pixels are independent
// and every pixel is written exactly once, so the ORDER in which tiles are traced is free -- frames stay
// bit-identical (tests/test_gpu_parity.py::test_frame_schedules_agree).
//
//   tile_probe_kernel : marches every 8th pixel of every 8th row of this call's tiles (the same rays the
//                       frame will trace: u = x/W*2-1 does not depend on the stride) for at most
//                       `probe_steps` steps with the exact fp32 evaluator and counts, per tile, the probes
//                       that are still marching.  Nothing is written to the frame or to the statistics.
//   tile_sort_kernel  : one block; counting sort of the tiles by that count, descending -> tile_order[].
//
// Cost: W*H/64 rays x <= 48 steps, two tiny launches (~0.25 ms at 1080p).  The tracers read
// RenderParams::tile_order in decode_pixel.  No reference counterpart: render_kernel
// (src/volumeRender_kernel.cu:608-692) advances all rays in lock step.
// =====================================================================================
#pragma once
#include "render_kernels.cuh"

namespace nsdf {

constexpr int PROBE_STRIDE = 8;                                   // pixels between probes
constexpr int PROBES_PER_TILE = (TILE / PROBE_STRIDE) * (TILE / PROBE_STRIDE);   // 64
constexpr int PROBE_MAX_SCORE = PROBES_PER_TILE;

__global__ void __launch_bounds__(128) tile_probe_kernel(const __grid_constant__ RenderParams rp, int n_jobs, int probe_steps,
                                                          int* __restrict__ tile_score) {
    extern __shared__ __align__(16) float s_w[];
    __shared__ uint64_t s_bar;
    stage_weights(s_w, rp.blob, rp.blob_bytes, &s_bar);
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    const int j = g / PROBES_PER_TILE, s = g % PROBES_PER_TILE;   // tile slot of this call, probe inside the tile
    bool marching = false;
    f3 p = {0.f, 0.f, 0.f}, d = {0.f, 0.f, 0.f};
    float tfar = 0.f, rem = -1.0f;
    const int view = j < n_jobs ? j / rp.my_tiles : 0;
    if (j < n_jobs) {
        const int t = rp.tile_rank + (j - view * rp.my_tiles) * rp.tile_world;
        const int tx = t % rp.tiles_x, ty = t / rp.tiles_x;
        const int x = tx * TILE + (s % (TILE / PROBE_STRIDE)) * PROBE_STRIDE + PROBE_STRIDE / 2 - 1;
        const int y = ty * TILE + (s / (TILE / PROBE_STRIDE)) * PROBE_STRIDE + PROBE_STRIDE / 2 - 1;
        if (x < rp.W && y < rp.H) {
            const RayInit r = init_ray(view_cam(rp, view), x, y, rp.W, rp.H, rp.refmath != 0);
            if (r.hit) { marching = true; p = r.p; d = r.d; tfar = r.tfar; }
        }
    }
    for (int k = 0; k < probe_steps; ++k) {
        if (!__ballot_sync(0xffffffffu, marching)) break;
        float x[1][3] = {{p.x, p.y, p.z}}, y[1];
        mlp32_eval_thread<1>(s_w, rp.n_sq, x, y, rp.t4, rp.n_in == 4);
        if (marching) {
            const float t = scene_eval(rp, view, p, y[0], rem);
            tfar = tfar - t;
            rem = rem - clearance_step(t);
            if (tfar <= 0.0f || t < NSDF_MARCHING_EPSILON) marching = false;
            else p = f3{fmaf(d.x, t, p.x), fmaf(d.y, t, p.y), fmaf(d.z, t, p.z)};
        }
    }
    if (marching) atomicAdd(tile_score + j, 1);
}

// tile_order[rank among this call's jobs by descending score] = job (view * my_tiles + tile slot).  One block of 256 threads.
__global__ void __launch_bounds__(256) tile_sort_kernel(const int* __restrict__ tile_score, int n_my_tiles,
                                                         int* __restrict__ tile_order) {
    __shared__ int s_count[PROBE_MAX_SCORE + 1], s_start[PROBE_MAX_SCORE + 1];
    for (int i = threadIdx.x; i <= PROBE_MAX_SCORE; i += blockDim.x) s_count[i] = 0;
    __syncthreads();
    for (int j = threadIdx.x; j < n_my_tiles; j += blockDim.x) {
        int sc = tile_score[j];
        sc = sc < 0 ? 0 : (sc > PROBE_MAX_SCORE ? PROBE_MAX_SCORE : sc);
        atomicAdd(&s_count[sc], 1);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        int acc = 0;
        for (int sc = PROBE_MAX_SCORE; sc >= 0; --sc) { s_start[sc] = acc; acc += s_count[sc]; }
    }
    __syncthreads();
    for (int j = threadIdx.x; j < n_my_tiles; j += blockDim.x) {
        int sc = tile_score[j];
        sc = sc < 0 ? 0 : (sc > PROBE_MAX_SCORE ? PROBE_MAX_SCORE : sc);
        const int pos = atomicAdd(&s_start[sc], 1);
        tile_order[pos] = j;
    }
}

}  // namespace nsdf
