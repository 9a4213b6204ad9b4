// =====================================================================================
// render_kernels.cuh -- persistent sphere tracer with fused normals + matcap shading.
//
// Replaces the reference's host-driven wavefront loop (render_kernel,
// src/volumeRender_kernel.cu:608-692: up to 6000 x {9 GEMM launches, singleMarch, thrust
// scan, blocking 4-byte D2H, createBatch}) by two launches with no host round trip:
//
//   trace_throughput_kernel<P>  persistent; one lane = P ray slots; the MLP is evaluated
//       in registers (mlp32_eval_thread).  Finished slots are refilled from a global pixel
//       counter with warp-ballot compaction (replaces thrust::exclusive_scan + createBatch),
//       so lanes stay full.  Rays that are still marching after `evict_steps` steps (and
//       the last few rays of a draining warp) are appended to the straggler queue.
//   trace_latency_kernel<R>     persistent; one warp = R ray slots, lane = neuron
//       (mlp32_eval_warp): one march step costs ~1.5k cycles instead of ~15k, which bounds
//       the serial tail of rays that need thousands of steps.
//
// Per ray both kernels perform exactly the per-pixel projection of the reference loop
// (initMarcher :293-358, createBatch :504-547, singleMarch :416-477), in the same
// arithmetic, so frames are bit-identical to the oracle whatever the schedule.
// =====================================================================================
#pragma once
#include "device_math.cuh"
#include "scene_program.cuh"
#include "mlp_fp32.cuh"
#include "mlp_generic.cuh"
#include "../../include/nsdf.h"

namespace nsdf {

constexpr int TILE = 64;           // multi-GPU partition granule (pixels)
constexpr int TILE_PIXELS = TILE * TILE;

struct StragglerEntry {            // 48 bytes, 16-byte aligned
    int pix, steps;
    float px, py;
    float pz, dx, dy, dz;
    float tfar, pad0, pad1;
    unsigned int ready;         // == RenderParams::frame_seq once the entry is complete (fused kernel)
};

struct RenderParams {
    Camera cam;                    // the view of a single-view call (n_views == 1); batched calls read views[]
    int n_views;                   // views rendered by this launch (nsdf_render_views): their tails overlap each other's bulk
    const Camera* views;           // device array [n_views] (global memory), null when n_views == 1
    unsigned int view_pixels;      // W * H: view v's frame starts at out + v * view_pixels
    int my_tiles;                  // tiles of ONE view this call renders; job g = view * my_tiles + tile slot
    uint32_t* stage;               // staged frame stores (tile-partitioned calls): [job][64 x 64] finished pixels, flushed to `out`
    unsigned int* tile_done;       //   as 16-byte row segments by the warp that completes a tile ([job] pixel counters); null = direct
    const float* blob;
    int n_sq;
    uint32_t blob_bytes;
    const uint32_t* matcap;
    int mw, mh;
    int color_type, scene;
    uint32_t* out;
    int W, H;
    int tiles_x, n_tiles;
    int tile_rank, tile_world;
    const int* tile_order;         // [j] = job traced j-th by this call (tile_priority.cuh), or null: job j
    unsigned int total_idx;        // pixel indices this call enumerates (jobs * 4096)
    int max_steps, evict_steps;
    int evict_steps_hi;            // fused kernel: unconditional hand-over (evict_steps applies while the backlog is short)
    unsigned int queue_room;       // fused kernel: backlog (tail - head) below which evict_steps applies
    unsigned int* pixel_counter;   // [0] next pixel index, [1] straggler tail, [2] straggler head,
                                   // [3] producer warps finished (fused kernel)
    unsigned int frame_seq;        // publication stamp of straggler entries (fused kernel), never 0
    unsigned int consumer_period;  // fused kernel: global warp w is a consumer from the start when
                                   // w % consumer_period == consumer_period - 1 (0 = none)
    StragglerEntry* queue;
    unsigned int queue_capacity;
    int isolate_steps;             // trace_coop_kernel: a warp holding a ray that has marched this many steps stops taking new
                                   // rays, so the long ray soon runs alone at the 1-ray latency (0 = off)
    int marathon_steps;            // trace_coop_kernel: a consumer warp holding a ray with this many steps stops popping (0 = off)
    int consumer_warps;            // trace_coop_kernel: the last this-many warps of every CTA consume from the start
    int pop_max;                   // trace_coop_kernel: a consumer warp holds at most this many popped rays (0 = all slots)
    StragglerEntry* xqueue;        // trace_coop_kernel: express queue (counters pixel_counter[4] tail, [5] head, [6] express rays in flight)
    unsigned int xqueue_soft_cap;  //   nothing is pushed beyond this index (the allocation has room for every warp's last push behind it)
    int express_steps;             //   a consumer's ray with this many steps moves to an express lane when one is free (0 = off)
    unsigned int express_period;   //   the warps 0..3 of every CTA on SMs with smid % express_period == express_period - 1 are express
    unsigned int express_lanes;    //   lanes: they hold ONE ray each, so that it marches at the 1-ray latency on a quiet scheduler
    int express_warps;             //   express lanes per CTA of an express SM (warps 0..express_warps-1)
    int tail_lanes;                //   once the pixels are gone, warps 0..tail_lanes-1 of every CTA are lanes for the express queue (0 = off)
    int tail_push_max;             //   and a warp outside them that is down to this many marching rays hands them over
    int express_rays;              //   rays an express lane marches together (1..8)
    int express_solo_steps;        //   a lane holding a ray with this many steps takes no further rays (0 = never alone by rule)
    int express_steps_old, express_check_old;   //   rays past express_steps_old steps ask for a lane whenever steps & express_check_old == 0
    int express_from_bulk;         //   producing warps hand rays over to free express lanes too (else only the consumers do)
    int timeline;                  // development: record the timeline of long rays (trace_coop.cuh g_timeline)
    int queue_static;              // trace_coop_kernel: the queue was filled by an earlier launch and IS the ray source
                                   // (second pass of the mixed-precision tracer): slots refill from it, nothing is pushed
    SceneProgram prog;             // scene == NSDF_SCENE_PROGRAM: the op list (scene_program.cuh)
    int n_in;                      // network inputs: 3 (x, y, z) or 4 (x, y, z, t) -- animation nets, volumeRender_kernel.cu:524-545
    float t4;                      // the fourth input of every query of this call: (float)frame number (c_frameNumber)
    int refmath;                   // nsdf_render_opts.math == NSDF_MATH_REFERENCE: rsqrtf / libdevice tanhf / __saturatef as the
                                   // reference's device code calls them (device_math.cuh); 0 = the IEEE-only forms the oracle restates
    int hybrid;                    // trace_coop_kernel: producing warps march one ray per LANE (mlp32_eval_thread, 32 rays per warp,
                                   // a third of the shared-memory wavefronts per query) and hand rays that reach evict_steps_hi
                                   // steps to the cooperative consumers on the quiet SMs; out of pixels they become consumers
    int b4_refill_min;             // trace_bulk4_kernel: a producer warp asks for pixels once this many of its 128 slots are idle
    nsdf_render_stats* stats;
};

// sceneSDF (volumeRender_kernel.cu:217-230) of this call: one of the two compiled scenes, or the op list
// The camera of view v: kernel parameters for a single-view call, a (cached) global array for a batch.
__device__ __forceinline__ const Camera& view_cam(const RenderParams& rp, int view) {
    return rp.n_views > 1 ? rp.views[view] : rp.cam;
}
// RaySlot::pix packs the view into the top byte: (view << 24) | (y * W + x); batched calls require W * H <= 2^24, views <= 128.
__device__ __forceinline__ int slot_view(int packed) { return (int)((unsigned int)packed >> 24); }
__device__ __forceinline__ int slot_pixel(int packed) { return packed & 0x00ffffff; }

// rem: the ray's clearance state (device_math.cuh, many_sphere_union); pass a local -1 where no state is carried.
__device__ __forceinline__ float scene_eval(const RenderParams& rp, int view, f3 p, float nsdf, float& rem) {
    const double zoff = view_cam(rp, view).zoff;
    if (rp.scene == NSDF_SCENE_PROGRAM) return scene_program_eval(rp.prog, p, nsdf, zoff, rp.refmath != 0);
    return scene_sdf(rp.scene, p, nsdf, zoff, rp.refmath != 0, rem);
}

struct RaySlot {
    int pix;      // linear pixel id y*W+x
    int steps;    // march steps taken
    int mode;     // -1 idle, 0 marching, 1..4 shading: next tetrahedron vertex = mode-1
    f3 p, d;
    float tfar;
    float s0, s1, s2;
    float rem;    // clearance to the scene's analytic parts at p (many_sphere_union), < 0 = unknown
};

struct LocalStats { unsigned int queries, hits, rays, evicted, unfinished, pixels; };

__device__ __forceinline__ bool decode_pixel(const RenderParams& rp, unsigned int idx, int& x, int& y, int& view) {
    const unsigned int j = idx / TILE_PIXELS, within = idx % TILE_PIXELS;
    const int g = rp.tile_order ? __ldg(rp.tile_order + j) : (int)j;
    view = g / rp.my_tiles;
    const int t = rp.tile_rank + (g - view * rp.my_tiles) * rp.tile_world;
    const int tx = t % rp.tiles_x, ty = t / rp.tiles_x;
    const unsigned int sb = within >> 5, l = within & 31u;        // 8x4-pixel sub-blocks: one warp's first fill
    x = tx * TILE + (int)(sb & 7u) * 8 + (int)(l & 7u);
    y = ty * TILE + (int)(sb >> 3) * 4 + (int)(l >> 3);
    return x < rp.W && y < rp.H;
}

// ---- frame stores ---------------------------------------------------------------------------------------------------
// Direct: one 4-byte store per finished pixel (the frame stays in L2 until the launch ends; ncu: DRAM write traffic below the
// frame size).  Staged (rp.stage != null; tile-partitioned calls, whose `out` is a peer GPU's memory reached over NVLink): the
// pixel goes to this GPU's staging tile and counts the tile down; the warp whose pixel completes the tile copies its 64 rows
// to `out` as coalesced 16-byte stores (256 B per row), so that the frame gather is a few thousand 256-byte row writes
// instead of millions of 4-byte ones, still inside the tracing kernel.
// emit_pixel returns the job whose tile this pixel completed (the caller's WARP must then call flush_tile), or -1.
__device__ __forceinline__ int emit_pixel(const RenderParams& rp, int packed, uint32_t color) {
    const int view = slot_view(packed), pix = slot_pixel(packed);
    if (!rp.stage) {
        rp.out[(size_t)view * rp.view_pixels + pix] = color;
        return -1;
    }
    const int x = pix % rp.W, y = pix / rp.W;
    const int tx = x / TILE, ty = y / TILE;
    const int t = ty * rp.tiles_x + tx;
    const int g = view * rp.my_tiles + (t - rp.tile_rank) / rp.tile_world;
    rp.stage[(size_t)g * TILE_PIXELS + (y - ty * TILE) * TILE + (x - tx * TILE)] = color;
    __threadfence();
    const int tw = min(TILE, rp.W - tx * TILE), th = min(TILE, rp.H - ty * TILE);
    const unsigned int old = atomicAdd(rp.tile_done + g, 1u);
    return (old + 1u == (unsigned int)(tw * th)) ? g : -1;
}
// All 32 lanes: copy the finished staging tile of job g to the frame.
__device__ __forceinline__ void flush_tile(const RenderParams& rp, int g, int lane) {
    __threadfence();
    const int view = g / rp.my_tiles;
    const int t = rp.tile_rank + (g - view * rp.my_tiles) * rp.tile_world;
    const int x0 = (t % rp.tiles_x) * TILE, y0 = (t / rp.tiles_x) * TILE;
    const int tw = min(TILE, rp.W - x0), th = min(TILE, rp.H - y0);
    const uint32_t* src = rp.stage + (size_t)g * TILE_PIXELS;
    uint32_t* dst = rp.out + (size_t)view * rp.view_pixels + (size_t)y0 * rp.W + x0;
    if ((rp.W & 3) == 0 && (tw & 3) == 0 && (((size_t)rp.out) & 15) == 0) {
        const int qpr = tw >> 2;                       // 16-byte segments per row
        for (int i = lane; i < th * 16; i += 32) {
            const int row = i >> 4, q = i & 15;
            if (q < qpr) {
                const uint4 v = __ldcg(reinterpret_cast<const uint4*>(src + row * TILE) + q);
                reinterpret_cast<uint4*>(dst + (size_t)row * rp.W)[q] = v;
            }
        }
    } else {
        for (int i = lane; i < th * TILE; i += 32) {
            const int row = i >> 6, c = i & 63;
            if (c < tw) dst[(size_t)row * rp.W + c] = __ldcg(src + row * TILE + c);
        }
    }
}
// Warp-level epilogue of a step: flush every tile one of the lanes completed (flush_g = emit_pixel's result, -1 = none).
__device__ __forceinline__ void flush_completed(const RenderParams& rp, int flush_g, int lane) {
    if (!rp.stage) return;
    unsigned int fm = __ballot_sync(0xffffffffu, flush_g >= 0);
    while (fm) {
        const int src = __ffs(fm) - 1;
        flush_tile(rp, __shfl_sync(0xffffffffu, flush_g, src), lane);
        fm &= fm - 1u;
    }
}

// Point the network is asked about (createBatch, volumeRender_kernel.cu:526-545).
__device__ __forceinline__ f3 slot_query_point(const RaySlot& s) {
    if (s.mode <= 0) return s.p;
    const f3 t = tet_vertex(s.mode - 1);
    return f3{fmaf(t.x, NSDF_NORMAL_EPSILON, s.p.x), fmaf(t.y, NSDF_NORMAL_EPSILON, s.p.y),
              fmaf(t.z, NSDF_NORMAL_EPSILON, s.p.z)};
}

// singleMarch (volumeRender_kernel.cu:416-477) for one slot whose request returned `nsdf`.
// flush_g receives the job whose staged tile this update completed (emit_pixel), else it is left alone.
__device__ __forceinline__ void slot_update(RaySlot& s, float nsdf, f3 q, const RenderParams& rp, LocalStats& st, int& flush_g) {
    ++st.queries;
    const int view = slot_view(s.pix);
    // q == s.p while marching; a tetrahedron vertex is 1.74e-5 away from it
    float rem = s.mode > 0 ? s.rem - 2.0e-5f : s.rem;
#ifdef NSDF_DEBUG_TSCALE   // development: every ray crawls and runs to max_steps (per-step latency probe)
    const float sv = s.mode > 0 ? scene_eval(rp, view, q, nsdf, rem) : scene_eval(rp, view, q, nsdf, rem) * NSDF_DEBUG_TSCALE + 2e-6f;
#else
    const float sv = scene_eval(rp, view, q, nsdf, rem);
#endif
    if (s.mode <= 0) s.rem = rem;
    if (s.mode > 0) {
        if (s.mode == 1) s.s0 = sv;
        else if (s.mode == 2) s.s1 = sv;
        else if (s.mode == 3) s.s2 = sv;
        if (s.mode < 4) { ++s.mode; return; }
        const float sarr[4] = {s.s0, s.s1, s.s2, sv};
        const f3 n = tet_normal(sarr, rp.refmath != 0);
        const uint32_t color = (rp.color_type == NSDF_COLOR_FACING)
                                   ? facing_color(n, s.d, rp.refmath != 0)
                                   : matcap_color(view_cam(rp, view), n, rp.matcap, rp.mw, rp.mh, rp.refmath != 0);
        flush_g = emit_pixel(rp, s.pix, color);
        ++st.hits;
        s.mode = -1;
        return;
    }
    const float t = sv;
    s.tfar = s.tfar - t;
    ++s.steps;
    if (s.tfar <= 0.0f) {             // left the march budget: background (:463-467)
        flush_g = emit_pixel(rp, s.pix, 0u);
        s.mode = -1;
        return;
    }
    s.p = f3{fmaf(s.d.x, t, s.p.x), fmaf(s.d.y, t, s.p.y), fmaf(s.d.z, t, s.p.z)};
    s.rem = s.rem - clearance_step(t);
    if (t < NSDF_MARCHING_EPSILON) s.mode = 1;
    if (s.steps >= rp.max_steps) {    // loop bound of render_kernel (:652): never shaded, stays 0
        flush_g = emit_pixel(rp, s.pix, 0u);
        ++st.unfinished;
        s.mode = -1;
    }
}

__device__ __forceinline__ void slot_update(RaySlot& s, float nsdf, f3 q, const RenderParams& rp, LocalStats& st) {
    int unused = -1;
    slot_update(s, nsdf, q, rp, st, unused);   // direct stores only (the host never stages for the kernels that call this form)
}

// Starts the ray of enumeration index idx in slot s (initMarcher).  Returns with s.mode == -1 when
// the pixel is outside the frame or misses the bounding sphere.
__device__ __forceinline__ void slot_start(RaySlot& s, unsigned int idx, const RenderParams& rp, LocalStats& st, int& flush_g) {
    int x, y, view;
    s.mode = -1;
    if (!decode_pixel(rp, idx, x, y, view)) return;
    ++st.pixels;
    const RayInit r = init_ray(view_cam(rp, view), x, y, rp.W, rp.H, rp.refmath != 0);
    const int pix = (y * rp.W + x) | (view << 24);
    if (!r.hit || rp.max_steps <= 0) { flush_g = emit_pixel(rp, pix, 0u); return; }
    ++st.rays;
    s.pix = pix; s.steps = 0; s.mode = 0; s.p = r.p; s.d = r.d; s.tfar = r.tfar; s.rem = -1.0f;
}

__device__ __forceinline__ void flush_stats(const LocalStats& st, nsdf_render_stats* g) {
    if (!g) return;
    unsigned int v[6] = {st.queries, st.hits, st.rays, st.evicted, st.unfinished, st.pixels};
#pragma unroll
    for (int i = 0; i < 6; ++i) {
        unsigned int x = v[i];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
        v[i] = x;
    }
    if ((threadIdx.x & 31) == 0) {
        unsigned long long* p = reinterpret_cast<unsigned long long*>(g);
#pragma unroll
        for (int i = 0; i < 6; ++i)
            if (v[i]) atomicAdd(p + i, (unsigned long long)v[i]);
    }
}

__device__ __forceinline__ unsigned int lanemask_lt() {
    unsigned int m;
    asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
    return m;
}

__device__ __forceinline__ void slot_start(RaySlot& s, unsigned int idx, const RenderParams& rp, LocalStats& st) {
    int unused = -1;
    slot_start(s, idx, rp, st, unused);
}

// Warp-aggregated append of marching rays to the straggler queue.
__device__ __forceinline__ void evict_slot(RaySlot& s, bool do_evict, const RenderParams& rp, LocalStats& st) {
    const unsigned int m = __ballot_sync(0xffffffffu, do_evict);
    if (!m) return;
    const int leader = __ffs(m) - 1;
    unsigned int base = 0;
    if ((int)(threadIdx.x & 31) == leader) base = atomicAdd(rp.pixel_counter + 1, (unsigned int)__popc(m));
    base = __shfl_sync(0xffffffffu, base, leader);
    if (do_evict) {
        const unsigned int slot = base + __popc(m & lanemask_lt());
        if (slot < rp.queue_capacity) {
            StragglerEntry e;
            e.pix = s.pix; e.steps = s.steps; e.px = s.p.x; e.py = s.p.y; e.pz = s.p.z;
            e.dx = s.d.x; e.dy = s.d.y; e.dz = s.d.z; e.tfar = s.tfar; e.pad0 = e.pad1 = 0.0f; e.ready = 0u;
            float4* dst = reinterpret_cast<float4*>(rp.queue + slot);
            const float4* src = reinterpret_cast<const float4*>(&e);
            dst[0] = src[0]; dst[1] = src[1]; dst[2] = src[2];
            ++st.evicted;
            s.mode = -1;
        }
    }
}

#ifdef NSDF_DEV_SCHEDULES   // superseded A/B schedules: development builds only (make dev)
// ---------------------------------------------------------------------------------------------
template <int P>
__global__ void __launch_bounds__(256, (P >= 2) ? 1 : 2) trace_throughput_kernel(const __grid_constant__ RenderParams rp) {
    extern __shared__ __align__(16) float s_w[];
    __shared__ uint64_t s_bar;
    stage_weights(s_w, rp.blob, rp.blob_bytes, &s_bar);

    const unsigned int lane = threadIdx.x & 31u;
    RaySlot slot[P];
#pragma unroll
    for (int p = 0; p < P; ++p) slot[p].mode = -1;
    LocalStats st = {0, 0, 0, 0, 0, 0};
    bool exhausted = false;   // warp-uniform

    for (;;) {
        // ---- refill idle slots from the global pixel counter (ballot compaction)
#pragma unroll
        for (int p = 0; p < P; ++p) {
            while (!exhausted) {
                const bool need = slot[p].mode < 0;
                const unsigned int m = __ballot_sync(0xffffffffu, need);
                if (!m) break;
                const int leader = __ffs(m) - 1;
                const unsigned int cnt = __popc(m);
                unsigned int base = 0;
                if ((int)lane == leader) base = atomicAdd(rp.pixel_counter, cnt);
                base = __shfl_sync(0xffffffffu, base, leader);
                if (need) {
                    const unsigned int idx = base + __popc(m & lanemask_lt());
                    if (idx < rp.total_idx) slot_start(slot[p], idx, rp, st);
                }
                if (base + cnt >= rp.total_idx) exhausted = true;
            }
        }
        bool any = false;
#pragma unroll
        for (int p = 0; p < P; ++p) any |= slot[p].mode >= 0;
        const unsigned int act = __ballot_sync(0xffffffffu, any);
        if (!act) break;

        // ---- one network evaluation per slot
        float x[P][3], y[P];
        f3 q[P];
#pragma unroll
        for (int p = 0; p < P; ++p) {
            q[p] = (slot[p].mode >= 0) ? slot_query_point(slot[p]) : f3{0.0f, 0.0f, 0.0f};
            x[p][0] = q[p].x; x[p][1] = q[p].y; x[p][2] = q[p].z;
        }
        mlp32_eval_thread<P>(s_w, rp.n_sq, x, y, rp.t4, rp.n_in == 4);
#pragma unroll
        for (int p = 0; p < P; ++p)
            if (slot[p].mode >= 0) slot_update(slot[p], y[p], q[p], rp, st);

        // ---- hand long rays (and the remains of a draining warp) to the latency kernel
        if (rp.queue) {
            int n_act = 0;
#pragma unroll
            for (int p = 0; p < P; ++p) n_act += __popc(__ballot_sync(0xffffffffu, slot[p].mode >= 0));
            const bool drain = exhausted && n_act <= 8 * P;
#pragma unroll
            for (int p = 0; p < P; ++p)
                evict_slot(slot[p], slot[p].mode == 0 && (drain || slot[p].steps >= rp.evict_steps), rp, st);
        }
    }
    flush_stats(st, rp.stats);
}

// ---------------------------------------------------------------------------------------------
// Latency path: lane r < R owns ray slot r; every lane is a neuron.
template <int R>
__global__ void __launch_bounds__(512, 1) trace_latency_kernel(const __grid_constant__ RenderParams rp) {
    extern __shared__ __align__(16) float s_w[];
    __shared__ uint64_t s_bar;
    stage_weights(s_w, rp.blob, rp.blob_bytes, &s_bar);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float* abuf = s_w + (rp.blob_bytes >> 2) + warp * (2 * R * WARP_ACT_STRIDE);

    unsigned int n_queue = rp.pixel_counter[1];
    if (n_queue > rp.queue_capacity) n_queue = rp.queue_capacity;
    RaySlot slot;
    slot.mode = -1;
    LocalStats st = {0, 0, 0, 0, 0, 0};
    bool exhausted = false;

    for (;;) {
        if (!exhausted) {
            const bool need = lane < R && slot.mode < 0;
            const unsigned int m = __ballot_sync(0xffffffffu, need);
            if (m) {
                const int leader = __ffs(m) - 1;
                const unsigned int cnt = __popc(m);
                unsigned int base = 0;
                if (lane == leader) base = atomicAdd(rp.pixel_counter + 2, cnt);
                base = __shfl_sync(0xffffffffu, base, leader);
                if (need) {
                    const unsigned int idx = base + __popc(m & lanemask_lt());
                    if (idx < n_queue) {
                        const float4* src = reinterpret_cast<const float4*>(rp.queue + idx);
                        const float4 e0 = src[0], e1 = src[1], e2 = src[2];
                        slot.pix = __float_as_int(e0.x); slot.steps = __float_as_int(e0.y);
                        slot.p = f3{e0.z, e0.w, e1.x}; slot.d = f3{e1.y, e1.z, e1.w};
                        slot.tfar = e2.x; slot.mode = 0; slot.rem = -1.0f;
                    }
                }
                if (base + cnt >= n_queue) exhausted = true;
            }
        }
        const unsigned int act = __ballot_sync(0xffffffffu, lane < R && slot.mode >= 0);
        if (!act) { if (exhausted) break; else continue; }

        const f3 q = (lane < R && slot.mode >= 0) ? slot_query_point(slot) : f3{0.0f, 0.0f, 0.0f};
        float x[R][3], y[R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            x[r][0] = __shfl_sync(0xffffffffu, q.x, r);
            x[r][1] = __shfl_sync(0xffffffffu, q.y, r);
            x[r][2] = __shfl_sync(0xffffffffu, q.z, r);
        }
        mlp32_eval_warp<R>(s_w, rp.n_sq, abuf, x, y, lane, rp.t4, rp.n_in == 4);
        float mine = y[0];
#pragma unroll
        for (int r = 1; r < R; ++r) mine = (lane == r) ? y[r] : mine;
        if (lane < R && slot.mode >= 0) slot_update(slot, mine, q, rp, st);
    }
    flush_stats(st, rp.stats);
}

#endif  // NSDF_DEV_SCHEDULES

// ---------------------------------------------------------------------------------------------
// Debug / A-B kernel: one thread per enumeration index, no refill, no eviction.
__global__ void __launch_bounds__(128) trace_simple_kernel(const __grid_constant__ RenderParams rp) {
    extern __shared__ __align__(16) float s_w[];
    __shared__ uint64_t s_bar;
    stage_weights(s_w, rp.blob, rp.blob_bytes, &s_bar);
    const unsigned int idx = blockIdx.x * blockDim.x + threadIdx.x;
    RaySlot slot;
    slot.mode = -1;
    LocalStats st = {0, 0, 0, 0, 0, 0};
    if (idx < rp.total_idx) slot_start(slot, idx, rp, st);
    while (__ballot_sync(0xffffffffu, slot.mode >= 0)) {
        const f3 q = (slot.mode >= 0) ? slot_query_point(slot) : f3{0.0f, 0.0f, 0.0f};
        float x[1][3] = {{q.x, q.y, q.z}}, y[1];
        mlp32_eval_thread<1>(s_w, rp.n_sq, x, y, rp.t4, rp.n_in == 4);
        if (slot.mode >= 0) slot_update(slot, y[0], q, rp, st);
    }
    flush_stats(st, rp.stats);
}

// ---------------------------------------------------------------------------------------------
// Any other network the reference's renderer would run (3 or 4 inputs, one output, any widths up to 128): one thread per
// enumeration index with the generic fused evaluator (mlp_generic.cuh).  Same per-ray arithmetic (slot_start / slot_update).
__global__ void __launch_bounds__(128) trace_generic_kernel(const __grid_constant__ RenderParams rp, const __grid_constant__ GenericNet g) {
    extern __shared__ __align__(16) float s_params[];
    const float* P = generic_stage(g, s_params);
    const unsigned int idx = blockIdx.x * blockDim.x + threadIdx.x;
    RaySlot slot;
    slot.mode = -1;
    LocalStats st = {0, 0, 0, 0, 0, 0};
    if (idx < rp.total_idx) slot_start(slot, idx, rp, st);
    float a[GENERIC_MAX_WIDTH], z[GENERIC_MAX_WIDTH];
    while (__ballot_sync(0xffffffffu, slot.mode >= 0)) {
        const f3 q = (slot.mode >= 0) ? slot_query_point(slot) : f3{0.0f, 0.0f, 0.0f};
        a[0] = q.x; a[1] = q.y; a[2] = q.z;
        if (rp.n_in == 4) a[3] = rp.t4;
        mlp_generic_eval(g, P, a, z);
        if (slot.mode >= 0) slot_update(slot, a[0], q, rp, st);
    }
    flush_stats(st, rp.stats);
}

}  // namespace nsdf
