// =====================================================================================
// nsdf_render.cu -- host side of the persistent sphere tracer (launches of
// render_kernels.cuh).  Replaces render_kernel / allocateBuffers
// (src/volumeRender_kernel.cu:587-692).  Compiled with -fmad=false.
// =====================================================================================
#include <cstdlib>
#include <cstring>
#include <vector>

#include "nsdf_internal.h"
#include "render_kernels.cuh"
#include "trace_coop.cuh"
#include "trace_fused.cuh"
#ifdef NSDF_DEV_SCHEDULES
#include "trace_tc.cuh"
#endif
#include "tile_priority.cuh"
#include "trace_bulk4.cuh"

namespace nsdf {

#ifdef NSDF_DEV_SCHEDULES
// Development build (make dev -> libnsdf_dev.so): every knob of the tracer can be overridden through the environment, read per
// call, and the superseded A/B schedules are compiled in.  The product build reads NO environment variable: its knobs are
// nsdf_render_opts.tuning (include/nsdf.h nsdf_tuning).
static int env_int(const char* name, int dflt) {
    const char* s = getenv(name);
    return (s && *s) ? atoi(s) : dflt;
}
#endif

#ifdef NSDF_DEV_SCHEDULES
constexpr int LAT_R = 4;            // ray slots per warp in the latency kernel
constexpr int LAT_THREADS = 512;

static size_t latency_smem(const nsdf_model* m) {
    return m->blob_bytes + (LAT_THREADS / 32) * 2 * LAT_R * WARP_ACT_STRIDE * sizeof(float);
}

static size_t fused_smem(const nsdf_model* m) {
    return m->blob_bytes + (FUSED_THREADS / 32) * 2 * FUSED_R * WARP_ACT_STRIDE * sizeof(float);
}

#endif  // NSDF_DEV_SCHEDULES

static size_t coop_smem(const nsdf_model* m) {
    const size_t need = m->blob_bytes + (COOP_THREADS / 32) * 2 * COOP_R * WARP_ACT_STRIDE * sizeof(float);
    return need > COOP_MIN_SMEM ? need : COOP_MIN_SMEM;   // more than half an SM: one CTA per SM
}

// The cooperative tracer's warps wait on one another (queues, termination counters): every CTA of the grid must be resident.
// cudaLaunchCooperativeKernel guarantees that or fails; the grid never exceeds one CTA per SM.
static cudaError_t launch_coop(const RenderParams& rp, unsigned int grid, size_t smem, cudaStream_t stream) {
    void* args[] = {const_cast<RenderParams*>(&rp)};
    return cudaLaunchCooperativeKernel((const void*)trace_coop_kernel, dim3(grid), dim3(COOP_THREADS), args, smem, stream);
}

// express queue of the cooperative tracer: pushes stop at half of it; the other half is room for every warp's last push
constexpr size_t XQUEUE_ENTRIES = 1u << 17;        // per view of a call

static int ensure_scratch(nsdf_model* m, size_t queue_entries, size_t xqueue_entries) {
    cudaError_t e;
    if (!m->d_counters) {
        e = cudaMalloc(&m->d_counters, 8 * sizeof(unsigned int));
        if (e != cudaSuccess) return (int)e;
    }
    if (xqueue_entries > m->xqueue_capacity) {
        if (m->d_xqueue) cudaFree(m->d_xqueue);
        m->d_xqueue = nullptr;
        m->xqueue_capacity = 0;
        e = cudaMalloc(&m->d_xqueue, xqueue_entries * sizeof(StragglerEntry));
        if (e != cudaSuccess) return (int)e;
        e = cudaMemset(m->d_xqueue, 0, xqueue_entries * sizeof(StragglerEntry));
        if (e != cudaSuccess) return (int)e;
        m->xqueue_capacity = xqueue_entries;
    }
    if (queue_entries > m->queue_capacity) {
        if (m->d_queue) cudaFree(m->d_queue);
        m->d_queue = nullptr;
        m->queue_capacity = 0;
        e = cudaMalloc(&m->d_queue, queue_entries * sizeof(StragglerEntry));
        if (e != cudaSuccess) return (int)e;
        e = cudaMemset(m->d_queue, 0, queue_entries * sizeof(StragglerEntry));   // no entry carries a live frame stamp
        if (e != cudaSuccess) return (int)e;
        m->queue_capacity = queue_entries;
    }
    return NSDF_OK;
}

void render_release(nsdf_model* m) {
    if (m->d_counters) cudaFree(m->d_counters);
    if (m->d_queue) cudaFree(m->d_queue);
    if (m->d_xqueue) cudaFree(m->d_xqueue);
    m->d_xqueue = nullptr;
    m->xqueue_capacity = 0;
    if (m->d_views) cudaFree(m->d_views);
    m->d_views = nullptr;
    m->views_capacity = 0;
    if (m->d_frame_stage) cudaFree(m->d_frame_stage);
    m->d_frame_stage = nullptr;
    m->frame_stage_jobs = 0;
    if (m->d_stats_scratch) cudaFree(m->d_stats_scratch);
    if (m->d_tile_scratch) cudaFree(m->d_tile_scratch);
    m->d_tile_scratch = nullptr;
    m->tile_scratch_capacity = 0;
    m->d_counters = nullptr;
    m->d_queue = nullptr;
    m->d_stats_scratch = nullptr;
    m->queue_capacity = 0;
}

// nsdf_scene_op list (host) -> kernel-parameter form; rejects unknown opcodes so that a typo is an error, not a scene
int scene_program_pack(const nsdf_scene_op* program, int n_ops, SceneProgram* out) {
    if (!program || n_ops < 1 || n_ops > SCENE_MAX_OPS) return NSDF_EINVAL;
    static_assert(sizeof(SceneOp) == sizeof(nsdf_scene_op) && SCENE_MAX_OPS == NSDF_SCENE_MAX_OPS, "scene op layout");
    std::memset(out, 0, sizeof(*out));
    out->n_ops = n_ops;
    for (int i = 0; i < n_ops; ++i) {
        const int pr = program[i].prim, cb = program[i].combine;
        const bool prim_ok = (pr >= NSDF_PRIM_NSDF && pr <= NSDF_PRIM_CYLINDER);
        const bool mod_ok = (pr >= NSDF_MOD_DISPLACE && pr <= NSDF_MOD_MANY_CYLINDER_CUT);
        if (!prim_ok && !mod_ok) return NSDF_EINVAL;
        if (prim_ok && (cb < NSDF_COMB_SET || cb > NSDF_COMB_SMOOTH_SUBTRACT_FROM)) return NSDF_EINVAL;
        std::memcpy(&out->ops[i], &program[i], sizeof(SceneOp));
    }
    return NSDF_OK;
}

__global__ void __launch_bounds__(256) scene_eval_kernel(const __grid_constant__ SceneProgram prog, const float* __restrict__ xyz,
                                                         const float* __restrict__ nsdf, double zoff, int refmath,
                                                         float* __restrict__ out, int64_t n) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    out[i] = scene_program_eval(prog, f3{xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]}, nsdf[i], zoff, refmath != 0);
}

int scene_eval_launch(const nsdf_scene_op* program, int n_ops, const float* d_xyz, const float* d_nsdf, int frame, int math,
                      float* d_out, int64_t n, cudaStream_t stream) {
    SceneProgram prog;
    const int rc = scene_program_pack(program, n_ops, &prog);
    if (rc) return rc;
    if (n <= 0) return NSDF_OK;
    scene_eval_kernel<<<(unsigned int)((n + 255) / 256), 256, 0, stream>>>(prog, d_xyz, d_nsdf, many_sphere_zoff(frame),
                                                                           math == NSDF_MATH_REFERENCE ? 1 : 0, d_out, n);
    return (int)cudaGetLastError();
}

#ifdef NSDF_DEV_SCHEDULES
// The development build's launcher: all schedules, all knobs from the environment (defaults = the product's).
int render_launch(nsdf_model* m, const nsdf_camera* cams, int n_views, const uint32_t* d_matcap, int mw, int mh,
                  const nsdf_render_opts* opts, uint32_t* d_out, int W, int H, nsdf_render_stats* d_stats,
                  cudaStream_t stream) {
    if (!m->fused || (m->in[0] != 3 && m->in[0] != 4)) return NSDF_EUNSUPPORTED;
    if (n_views < 1 || n_views > 128) return NSDF_EINVAL;
    if (n_views > 1 && (long long)W * H > (1LL << 24)) return NSDF_EINVAL;   // RaySlot::pix carries the view in its top byte
    const nsdf_camera* cam = cams;
    // 4-input (x, y, z, t) nets: fp32 tracers only; t = the frame number (createBatch, volumeRender_kernel.cu:532-545)
    if (m->in[0] == 4 && (opts->flags & NSDF_RENDER_TC_FARFIELD)) return NSDF_EUNSUPPORTED;
    if (opts->tile_world < 1 || opts->tile_rank < 0 || opts->tile_rank >= opts->tile_world) return NSDF_EINVAL;
    if (opts->color_type == NSDF_COLOR_MATCAP && (!d_matcap || mw <= 0 || mh <= 0)) return NSDF_EINVAL;

    RenderParams rp;
    std::memset(&rp, 0, sizeof(rp));
    std::memcpy(rp.cam.inv_view, cam->inv_view, sizeof(rp.cam.inv_view));
    std::memcpy(rp.cam.normal, cam->normal, sizeof(rp.cam.normal));
    rp.cam.frame = cam->frame;
    rp.cam.zoff = many_sphere_zoff(cam->frame);
    rp.n_views = n_views;
    rp.views = nullptr;
    rp.view_pixels = (unsigned int)((long long)W * H);
    rp.blob = m->d_blob;
    rp.n_sq = m->n_sq;
    rp.blob_bytes = m->blob_bytes;
    rp.matcap = d_matcap;
    rp.mw = mw;
    rp.mh = mh;
    rp.color_type = opts->color_type;
    rp.scene = opts->scene;
    if (opts->scene == NSDF_SCENE_PROGRAM) {
        const int rc_prog = scene_program_pack(opts->program, opts->n_ops, &rp.prog);
        if (rc_prog) return rc_prog;
    } else if (opts->scene != NSDF_SCENE_TANH && opts->scene != NSDF_SCENE_MANYSPHERE) {
        return NSDF_EINVAL;
    }
    rp.out = d_out;
    rp.W = W;
    rp.H = H;
    rp.tiles_x = (W + TILE - 1) / TILE;
    const int tiles_y = (H + TILE - 1) / TILE;
    rp.n_tiles = rp.tiles_x * tiles_y;
    rp.tile_rank = opts->tile_rank;
    rp.tile_world = opts->tile_world;
    const int my_tiles = (rp.n_tiles - opts->tile_rank + opts->tile_world - 1) / opts->tile_world;
    if (my_tiles <= 0) return NSDF_OK;
    const long long n_jobs_ll = (long long)my_tiles * n_views;       // job = (view, tile of this call)
    if (n_jobs_ll * TILE_PIXELS >= (1LL << 32)) return NSDF_EINVAL;
    const int n_jobs = (int)n_jobs_ll;
    rp.my_tiles = my_tiles;
    rp.total_idx = (unsigned int)n_jobs * TILE_PIXELS;
    rp.max_steps = opts->max_steps;
    rp.isolate_steps = env_int("NSDF_COOP_ISOLATE", 0);
    rp.marathon_steps = env_int("NSDF_COOP_MARATHON", 0);
    rp.timeline = env_int("NSDF_COOP_TIMELINE", 0);
    rp.xqueue = nullptr;
    rp.xqueue_soft_cap = 0u;
    rp.express_steps = 0;
    rp.express_period = 0u;
    rp.express_lanes = 0u;
    rp.express_from_bulk = 0;
    rp.express_warps = env_int("NSDF_COOP_EXPRESS_WARPS", 4);
    rp.express_rays = env_int("NSDF_COOP_EXPRESS_RAYS", 1);   // measured: several rays per lane slow the longest chains more than the extra lanes help
    if (rp.express_rays < 1) rp.express_rays = 1;
    if (rp.express_rays > 8) rp.express_rays = 8;
    rp.express_solo_steps = env_int("NSDF_COOP_EXPRESS_SOLO", 768);
    rp.tail_lanes = env_int("NSDF_COOP_TAIL_LANES", 0);   // measured neutral at 1080p and harmful for small frames: off
    rp.tail_push_max = env_int("NSDF_COOP_TAIL_PUSH", 2);
    rp.express_steps_old = env_int("NSDF_COOP_EXPRESS_OLD", 192);      // older rays ask for a lane every 2nd step instead of every 16th:
    rp.express_check_old = env_int("NSDF_COOP_EXPRESS_OLD_MASK", 1);   // the lottery for a free lane favours them (median 11.9 -> 11.3 ms)
    rp.consumer_warps = env_int("NSDF_COOP_CONSUMER_WARPS", 0);
    rp.pop_max = env_int("NSDF_COOP_POP_MAX", 0);
    rp.n_in = m->in[0];
    rp.t4 = (float)cam->frame;
    if (opts->math != NSDF_MATH_REFERENCE && opts->math != NSDF_MATH_DETERMINISTIC) return NSDF_EINVAL;
    rp.refmath = opts->math == NSDF_MATH_REFERENCE ? 1 : 0;
    rp.tile_order = nullptr;
    rp.stats = d_stats;

    const bool farfield = (opts->flags & NSDF_RENDER_TC_FARFIELD) != 0;
    // two-pass schedule: asked for, or chosen for large calls when the caller expressed no schedule preference
    // large calls (the bulk dominates, not the longest ray): ray-per-lane producers.  NSDF_LARGE_SCHEDULE: 1 = hybrid single
    // launch, 2 = two-pass, 0 = the cooperative tracer at every size (default since its 16-ray tiled evaluator: 4K 49.5 -> 41.3 ms)
    const bool large = !farfield && opts->flags == 0 && opts->evict_steps == 0 &&
                       rp.total_idx >= (unsigned int)env_int("NSDF_TWO_PASS_MIN_PIXELS", 3000000);
    const int large_schedule = large ? env_int("NSDF_LARGE_SCHEDULE", 0) : 0;
    const bool two_pass = !farfield && ((opts->flags & NSDF_RENDER_TWO_PASS) != 0 || large_schedule == 2);
    const bool hybrid = !farfield && !two_pass && ((opts->flags & NSDF_RENDER_HYBRID) != 0 || large_schedule == 1 ||
                                                   env_int("NSDF_COOP_HYBRID", 0) != 0);
    const bool simple = !farfield && (opts->flags & NSDF_RENDER_SIMPLE) != 0;
    const bool coop = !simple && !two_pass && !(opts->flags & (NSDF_RENDER_NO_STRAGGLER | NSDF_RENDER_PHASE_THROUGHPUT | NSDF_RENDER_PHASE_LATENCY |
                                                   NSDF_RENDER_TWO_KERNELS | NSDF_RENDER_QUEUE));
    const bool coop_queue = coop && !(opts->flags & NSDF_RENDER_NO_QUIET_SMS) && env_int("NSDF_COOP_QUEUE", 1) != 0;
    const bool straggler = farfield || (!simple && (coop ? coop_queue : !(opts->flags & NSDF_RENDER_NO_STRAGGLER)));
    const bool two_kernel = !farfield && straggler && (two_pass || (opts->flags & (NSDF_RENDER_PHASE_THROUGHPUT | NSDF_RENDER_PHASE_LATENCY |
                                                         NSDF_RENDER_TWO_KERNELS)) != 0);
    // fused kernel: only the long tail leaves the throughput lanes early; two-kernel path: its old default
    rp.evict_steps = opts->evict_steps > 0 ? opts->evict_steps
                                           : (two_pass ? env_int("NSDF_TWO_PASS_EVICT", 128)
                                              : two_kernel ? env_int("NSDF_EVICT_STEPS", 48) : env_int("NSDF_FUSED_EVICT_STEPS", 32));
    rp.evict_steps_hi = opts->evict_steps > 0 ? opts->evict_steps : env_int("NSDF_FUSED_EVICT_HI", 256);
    size_t qcap = 0;
    if (straggler) {
        qcap = rp.total_idx;   // every ray is queued at most once
        if (two_kernel) {
            const size_t cap_max = (size_t)env_int("NSDF_QUEUE_MAX", 4 << 20);
            if (qcap > cap_max) qcap = cap_max;
        }
    }
    int rc = ensure_scratch(m, qcap, XQUEUE_ENTRIES * (size_t)n_views);
    if (rc) return rc;
    const unsigned int xq_soft_cap = (unsigned int)(XQUEUE_ENTRIES * (size_t)n_views / 2);
    if (n_views > 1) {
        // batched views: only the cooperative single-launch schedule enumerates (view, tile) jobs
        if (simple || farfield || !coop || hybrid) return NSDF_EUNSUPPORTED;
        if ((size_t)n_views > m->views_capacity) {
            if (m->d_views) cudaFree(m->d_views);
            m->d_views = nullptr;
            m->views_capacity = 0;
            if (cudaMalloc(&m->d_views, (size_t)n_views * sizeof(Camera)) != cudaSuccess) return NSDF_ENOMEM;
            m->views_capacity = (size_t)n_views;
        }
        std::vector<Camera> hv((size_t)n_views);
        for (int v = 0; v < n_views; ++v) {
            std::memcpy(hv[v].inv_view, cams[v].inv_view, sizeof(hv[v].inv_view));
            std::memcpy(hv[v].normal, cams[v].normal, sizeof(hv[v].normal));
            hv[v].frame = cams[v].frame;
            hv[v].zoff = many_sphere_zoff(cams[v].frame);
        }
        cudaError_t ev = cudaMemcpyAsync(m->d_views, hv.data(), hv.size() * sizeof(Camera), cudaMemcpyHostToDevice, stream);
        if (ev != cudaSuccess) return (int)ev;
        ev = cudaStreamSynchronize(stream);   // hv is a local: the copy must have left it (a few KB, once per batch)
        if (ev != cudaSuccess) return (int)ev;
        rp.views = reinterpret_cast<const Camera*>(m->d_views);
    }
    // staged frame stores (render_kernels.cuh emit_pixel / flush_tile): for tile-partitioned calls, whose d_out is a peer
    // GPU's memory in the multi-GPU gather, or on request
    rp.stage = nullptr;
    rp.tile_done = nullptr;
    if (coop && !hybrid && !farfield && (opts->tile_world > 1 || (opts->flags & NSDF_RENDER_STAGED_STORES))) {
        if ((size_t)n_jobs > m->frame_stage_jobs) {
            if (m->d_frame_stage) cudaFree(m->d_frame_stage);
            m->d_frame_stage = nullptr;
            m->frame_stage_jobs = 0;
            if (cudaMalloc(&m->d_frame_stage, (size_t)n_jobs * (TILE_PIXELS + 1) * sizeof(uint32_t)) != cudaSuccess) return NSDF_ENOMEM;
            m->frame_stage_jobs = (size_t)n_jobs;
        }
        rp.stage = reinterpret_cast<uint32_t*>(m->d_frame_stage);
        rp.tile_done = reinterpret_cast<unsigned int*>(m->d_frame_stage) + m->frame_stage_jobs * TILE_PIXELS;
        cudaError_t es = cudaMemsetAsync(rp.tile_done, 0, (size_t)n_jobs * sizeof(unsigned int), stream);
        if (es != cudaSuccess) return (int)es;
    }
    rp.pixel_counter = m->d_counters;
    rp.queue = straggler ? reinterpret_cast<StragglerEntry*>(m->d_queue) : nullptr;
    rp.queue_capacity = straggler ? (unsigned int)qcap : 0u;

    const bool only_latency = straggler && (opts->flags & NSDF_RENDER_PHASE_LATENCY) != 0;
    const bool only_throughput = straggler && (opts->flags & NSDF_RENDER_PHASE_THROUGHPUT) != 0;
    cudaError_t e = cudaSuccess;
    if (!only_latency) {
        e = cudaMemsetAsync(m->d_counters, 0, 8 * sizeof(unsigned int), stream);
        if (e != cudaSuccess) return (int)e;
    }

    const int sms = m->sm_count > 0 ? m->sm_count : 148;
    NSDF_PER_DEVICE_FLAG(attr_done);
    if (!attr_done) {
        e = cudaFuncSetAttribute(trace_throughput_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
        if (e != cudaSuccess) return (int)e;
        e = cudaFuncSetAttribute(trace_throughput_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
        if (e != cudaSuccess) return (int)e;
        e = cudaFuncSetAttribute(trace_simple_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
        if (e != cudaSuccess) return (int)e;
        e = cudaFuncSetAttribute(trace_latency_kernel<LAT_R>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        if (e != cudaSuccess) return (int)e;
        e = cudaFuncSetAttribute(trace_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
        if (e != cudaSuccess) return (int)e;
        e = cudaFuncSetAttribute(trace_coop_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        if (e != cudaSuccess) return (int)e;
        attr_done = true;
    }

    // ---- tile priority: probe every 8th pixel for a few steps and trace the tiles with grazing rays first, so that the
    // frame's longest dependent chains start at t = 0 and hide behind the bulk (tile_priority.cuh).  Order-free: same bytes.
    // Measured (tools/priority_probe.py): pays when the tail would outlast the bulk -- the single-launch tracer on 0.5-3 M
    // pixels (1080p: 18.5 -> 15.2 ms); neutral-to-negative for small frames (probe cost), for the two-pass schedule (its
    // second pass only starts after the bulk) and for the mixed-precision mode (already at one ray's chain).
    const int probe_steps = env_int("NSDF_TILE_PRIORITY_STEPS", 32);
    const bool priority_default = coop && n_jobs >= env_int("NSDF_TILE_PRIORITY_MIN_TILES", 128);
    if (probe_steps > 0 && !simple && !only_latency && (priority_default || env_int("NSDF_TILE_PRIORITY_FORCE", 0))) {
        if ((size_t)n_jobs > m->tile_scratch_capacity) {
            if (m->d_tile_scratch) cudaFree(m->d_tile_scratch);
            m->d_tile_scratch = nullptr;
            m->tile_scratch_capacity = 0;
            e = cudaMalloc(&m->d_tile_scratch, 2 * (size_t)n_jobs * sizeof(int));
            if (e != cudaSuccess) return (int)e;
            m->tile_scratch_capacity = (size_t)n_jobs;
        }
        int* d_score = m->d_tile_scratch;
        int* d_order = m->d_tile_scratch + m->tile_scratch_capacity;
        e = cudaMemsetAsync(d_score, 0, (size_t)n_jobs * sizeof(int), stream);
        if (e != cudaSuccess) return (int)e;
        NSDF_PER_DEVICE_FLAG(probe_attr);
        if (!probe_attr) {
            e = cudaFuncSetAttribute(tile_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
            if (e != cudaSuccess) return (int)e;
            probe_attr = true;
        }
        const unsigned int probes = (unsigned int)n_jobs * PROBES_PER_TILE;
        tile_probe_kernel<<<(probes + 127u) / 128u, 128, m->blob_bytes, stream>>>(rp, n_jobs, probe_steps, d_score);
        tile_sort_kernel<<<1, 256, 0, stream>>>(d_score, n_jobs, d_order);
        e = cudaGetLastError();
        if (e != cudaSuccess) return (int)e;
        rp.tile_order = d_order;
    }

    if (simple) {
        const unsigned int blocks = (rp.total_idx + 127u) / 128u;
        trace_simple_kernel<<<blocks, 128, m->blob_bytes, stream>>>(rp);
        return (int)cudaGetLastError();
    }

    if (opts->flags & NSDF_RENDER_TC_FARFIELD) {
        // ---- mixed precision: far-field marching on the tensor cores (trace_tc.cuh), then the fp32 consumers
        // of trace_coop_kernel finish every ray that got close to the surface (bit-exact arithmetic from there on)
        if (opts->scene != NSDF_SCENE_TANH) return NSDF_EUNSUPPORTED;
        const bool x3 = env_int("NSDF_FARFIELD_FP16", 0) == 0;
        const int mode = x3 ? tc::MODE_FP16X3 : tc::MODE_FP16;
        TraceTcParams tp;
        tp.image = tc_image(m, mode, &tp.image_bytes);
        if (!tp.image) return NSDF_ENOMEM;
        const char* e_thr = getenv("NSDF_FARFIELD_THR");
        const char* e_mar = getenv("NSDF_FARFIELD_MARGIN");
        tp.thr = (e_thr && *e_thr) ? (float)atof(e_thr) : (x3 ? 1e-4f : 1e-2f);
        tp.margin = (e_mar && *e_mar) ? (float)atof(e_mar) : (x3 ? 8e-6f : 2.5e-3f);
        tp.max_tc_steps = env_int("NSDF_FARFIELD_MAX_STEPS", opts->max_steps);   // no cap measured best on the 1080p / 4K frames
        tp.drain_below = env_int("NSDF_FARFIELD_DRAIN", 16);
        m->frame_seq = m->frame_seq + 1u ? m->frame_seq + 1u : 1u;
        rp.frame_seq = m->frame_seq;
        const int per_sm = 512 / (int)(x3 ? tc::Cols<tc::MODE_FP16X3>::TOTAL : tc::Cols<tc::MODE_FP16>::TOTAL);
        size_t smem = tp.image_bytes;
        const size_t min_smem = (size_t)(227 * 1024) / (size_t)(per_sm + 1) + 1;   // never more CTAs than TMEM columns allow
        if (smem < min_smem) smem = min_smem;
        NSDF_PER_DEVICE_FLAG(tc_attr_done);
        if (!tc_attr_done) {
            e = cudaFuncSetAttribute(trace_tc_kernel<tc::MODE_FP16X3>, cudaFuncAttributeMaxDynamicSharedMemorySize, 120 * 1024);
            if (e != cudaSuccess) return (int)e;
            e = cudaFuncSetAttribute(trace_tc_kernel<tc::MODE_FP16>, cudaFuncAttributeMaxDynamicSharedMemorySize, 120 * 1024);
            if (e != cudaSuccess) return (int)e;
            tc_attr_done = true;
        }
        const unsigned int want_tc = (rp.total_idx + 127u) / 128u;
        const unsigned int cap_tc = (unsigned int)(sms * per_sm);
        const unsigned int grid_tc = want_tc < cap_tc ? want_tc : cap_tc;
        if (x3) trace_tc_kernel<tc::MODE_FP16X3><<<grid_tc, 128, smem, stream>>>(rp, tp);
        else trace_tc_kernel<tc::MODE_FP16><<<grid_tc, 128, smem, stream>>>(rp, tp);
        e = cudaGetLastError();
        if (e != cudaSuccess) return (int)e;
        int occ = 0;
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, trace_coop_kernel, COOP_THREADS, coop_smem(m));
        if (e != cudaSuccess) return (int)e;
        if (occ < 1) return NSDF_EUNSUPPORTED;
        if (env_int("NSDF_FARFIELD_SKIP_PASS2", 0)) return NSDF_OK;   // development: time the first pass alone
        RenderParams rp2 = rp;
        rp2.total_idx = 0;          // no pixels of its own: ray slots refill from the queue the first pass filled
        rp2.consumer_period = 0;
        rp2.queue_static = 1;
        rp2.evict_steps_hi = 1 << 30;
        rp2.express_period = (unsigned int)env_int("NSDF_COOP_EXPRESS_PERIOD", 12);   // the serial tail of this pass: express lanes
        rp2.express_steps = env_int("NSDF_COOP_EXPRESS_STEPS", 160);
        rp2.express_from_bulk = 1;
        if (rp2.express_period > 0 && rp2.express_steps > 0) {
            rp2.xqueue = reinterpret_cast<StragglerEntry*>(m->d_xqueue);
            rp2.xqueue_soft_cap = xq_soft_cap;
            rp2.express_lanes = ((unsigned int)(sms * occ) / rp2.express_period) * (unsigned int)(rp.express_warps * rp.express_rays);
        }
        return (int)launch_coop(rp2, (unsigned int)(sms * occ), coop_smem(m), stream);
    }

    if (coop) {
        // ---- default: one persistent launch, every warp marches COOP_R rays cooperatively (trace_coop.cuh)
        int occ = 0;
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, trace_coop_kernel, COOP_THREADS, coop_smem(m));
        if (e != cudaSuccess) return (int)e;
        if (occ < 1) return NSDF_EUNSUPPORTED;
        { const int occ_max = env_int("NSDF_COOP_OCC", 0); if (occ_max > 0 && occ > occ_max) occ = occ_max; }
        const unsigned int per_cta = (COOP_THREADS / 32) * COOP_R;
        const unsigned int cap = (unsigned int)(sms * occ);
        const unsigned int want = (rp.total_idx + per_cta - 1) / per_cta;
        m->frame_seq = m->frame_seq + 1u ? m->frame_seq + 1u : 1u;
        rp.frame_seq = m->frame_seq;
        // Serial tail.  Measured with the timeline hook (tools/timeline_probe.py): a ray inside a 16-ray warp steps
        // every ~13 us, a lone ray beside crowded warps every 3.5-5.5 us, a lone ray on an SM whose schedulers run nothing but lone
        // rays every 1.15-1.6 us.  So every express_period-th SM is an express SM (4 warps per CTA hold ONE ray each) and any warp
        // hands a ray that has marched express_steps steps to a free express lane; rays that find no lane stay where they are.
        // That replaces the quiet SMs + unconditional hand-over at 256 steps (kept for callers that set evict_steps).
        // Only for full grids: a grid smaller than the machine could land on express SMs alone and trace nothing.
        const bool full_grid = want >= cap;
        const bool caller_evicts = opts->evict_steps > 0;
        rp.consumer_period = full_grid ? (unsigned int)env_int("NSDF_COOP_QUIET_PERIOD", caller_evicts ? 16 : 0) : 0u;
        rp.evict_steps_hi = caller_evicts ? opts->evict_steps : env_int("NSDF_COOP_EVICT", 1 << 30);
        rp.express_period = full_grid && !caller_evicts && !(hybrid && coop_queue) ? (unsigned int)env_int("NSDF_COOP_EXPRESS_PERIOD", rp.total_idx >= 3000000u ? 24 : 12) : 0u;   // large calls are bound by the bulk
        rp.express_steps = env_int("NSDF_COOP_EXPRESS_STEPS", 160);
        rp.express_from_bulk = env_int("NSDF_COOP_EXPRESS_FROM_BULK", 1);
        if (coop_queue && rp.express_period > 0 && rp.express_steps > 0) {
            rp.xqueue = reinterpret_cast<StragglerEntry*>(m->d_xqueue);
            rp.xqueue_soft_cap = xq_soft_cap;
            rp.express_lanes = (cap / rp.express_period) * (unsigned int)(rp.express_warps * rp.express_rays);
        }
        if (hybrid && coop_queue) {
            rp.hybrid = env_int("NSDF_COOP_HYBRID", 8);   // a dry producer warp with <= this many rays hands them all over
            if (rp.hybrid < 1) rp.hybrid = 1;
            rp.consumer_period = (unsigned int)env_int("NSDF_COOP_QUIET_PERIOD", 32);
            rp.evict_steps_hi = opts->evict_steps > 0 ? opts->evict_steps : env_int("NSDF_COOP_EVICT", 192);
            const unsigned int quiet_sms = rp.consumer_period ? (unsigned int)sms / rp.consumer_period : 0u;
            rp.queue_room = opts->evict_steps > 0 ? 0u : (unsigned int)env_int("NSDF_COOP_ROOM", (int)(quiet_sms * 2u * per_cta));
        }
        return (int)launch_coop(rp, want < cap ? want : cap, coop_smem(m), stream);
    }

    if (!two_kernel) {
        // ---- A/B: producer lanes + consumer warps through a queue (trace_fused.cuh); grid = co-resident CTAs
        int occ = 0;
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, trace_fused_kernel, FUSED_THREADS, fused_smem(m));
        if (e != cudaSuccess) return (int)e;
        if (occ < 1) return NSDF_EUNSUPPORTED;
        const unsigned int cap = (unsigned int)(sms * occ);
        const unsigned int want = (rp.total_idx + FUSED_THREADS - 1) / FUSED_THREADS;
        const unsigned int grid = want < cap ? want : cap;
        m->frame_seq = m->frame_seq + 1u ? m->frame_seq + 1u : 1u;
        rp.frame_seq = m->frame_seq;
        const unsigned int total_warps = grid * (FUSED_THREADS / 32);
        unsigned int period = (unsigned int)env_int("NSDF_FUSED_CONSUMER_PERIOD", 16);
        if (period > total_warps) period = total_warps;   // at least one consumer from the start
        if (period < 2) period = total_warps >= 2 ? 2 : 0; // a lone warp produces first, then consumes
        rp.consumer_period = straggler ? period : 0u;
        const unsigned int consumers0 = period ? total_warps / period : 0u;
        rp.queue_room = (unsigned int)env_int("NSDF_FUSED_ROOM", (int)(2u * FUSED_R * (consumers0 ? consumers0 : 1u)));
        trace_fused_kernel<<<grid, FUSED_THREADS, fused_smem(m), stream>>>(rp);
        return (int)cudaGetLastError();
    }

    if (only_latency) {
        trace_latency_kernel<LAT_R><<<sms, LAT_THREADS, latency_smem(m), stream>>>(rp);
        return (int)cudaGetLastError();
    }
    // throughput kernel: persistent, grid = SMs x resident CTAs
    const int P = env_int("NSDF_TRACE_P", 1);
    const unsigned int per_cta = 256u * (unsigned)P;
    unsigned int want = (rp.total_idx + per_cta - 1) / per_cta;
    if (P == 2) {
        const unsigned int grid = want < (unsigned)sms ? want : (unsigned)sms;
        trace_throughput_kernel<2><<<grid, 256, m->blob_bytes, stream>>>(rp);
    } else {
        const unsigned int cap = 2u * (unsigned)sms;
        const unsigned int grid = want < cap ? want : cap;
        trace_throughput_kernel<1><<<grid, 256, m->blob_bytes, stream>>>(rp);
    }
    e = cudaGetLastError();
    if (e != cudaSuccess) return (int)e;
    if (straggler && !only_throughput) {
        if (two_pass || env_int("NSDF_TWO_PASS2_COOP", 0)) {   // second pass = cooperative tracer refilling from the static queue
            int occ = 0;
            e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, trace_coop_kernel, COOP_THREADS, coop_smem(m));
            if (e != cudaSuccess) return (int)e;
            RenderParams rp2 = rp;
            rp2.total_idx = 0;
            rp2.consumer_period = 0;
            rp2.queue_static = 1;
            e = launch_coop(rp2, (unsigned int)(sms * occ), coop_smem(m), stream);
            if (e != cudaSuccess) return (int)e;
        } else {
            trace_latency_kernel<LAT_R><<<sms, LAT_THREADS, latency_smem(m), stream>>>(rp);
        }
        e = cudaGetLastError();
    }
    return (int)e;
}

#else  // ---- product build --------------------------------------------------------------------------------------------------

static void resolve_tuning(const nsdf_render_opts* opts, unsigned int total_idx, int n_views, nsdf_tuning* T) {
    nsdf_tuning_default(T);
    if (opts->tuning) *T = *opts->tuning;
    // Automatic choices (measured, profiles/r02_tuning_sweeps.txt).  A call of many views is pure throughput: the tails of its
    // views hide behind each other's bulk, and once the pixels are gone the last long rays run alone in thinning warps anyway,
    // so express SMs and the priority probe only cost (64 views: 10.39 -> 9.70 ms per frame without them; 8 views: equal).
    const bool many_views = n_views >= 12;
    if (T->express_period < 0) T->express_period = many_views ? 0 : (total_idx >= 3000000u ? 24 : 8);
    if (T->tile_priority_steps < 0) T->tile_priority_steps = many_views ? 0 : 32;
    if (T->express_rays < 1) T->express_rays = 1;
    if (T->express_rays > 8) T->express_rays = 8;
    if (T->express_warps < 1) T->express_warps = 1;
    if (T->express_warps > COOP_THREADS / 32) T->express_warps = COOP_THREADS / 32;
    // ... and their bulk runs best on the four-rays-per-lane evaluator (trace_bulk4.cuh; measured per 1080p frame, cooperative
    // tracer -> bulk4: 4 views 14.03 -> 13.58 ms, 8: 13.48 -> 12.29, 16: 11.90 -> 10.77, 64: 9.47 -> 8.55; one view 17.8 -> 24.3)
    // Small calls stay on the cooperative tracer: a bulk4 warp steps its 128 slots in ~40 us whatever their occupancy, so the tail
    // of a call must hide behind enough bulk (64 views of 512^2 = 16.8 M pixels: 1.377 -> 1.323 ms per frame; 16 views = 4.2 M and
    // 8 views = 2.1 M pixels per call: 0.86 x and 0.39 x the cooperative tracer's rate, profiles/r02_scaling_summary.txt)
    if (T->bulk4 < 0) T->bulk4 = (n_views >= 4 && total_idx >= 12000000u) ? 1 : 0;
    if (T->bulk4_evict_steps <= 0) T->bulk4_evict_steps = 1 << 30;
    if (T->bulk4_period == 1) T->bulk4_period = 2;     // somebody has to take pixels
    if (T->bulk4_drain < 0) T->bulk4_drain = 0;
    if (T->bulk4_drain > 128) T->bulk4_drain = 128;
    if (T->bulk4_refill_min < 1) T->bulk4_refill_min = 1;
    if (T->bulk4_refill_min > 64) T->bulk4_refill_min = 64;
}

// The product launcher: one schedule (the cooperative persistent tracer, trace_coop.cuh) plus the one-thread-per-pixel
// debug kernel; knobs from nsdf_render_opts.tuning, no environment reads.
int render_launch(nsdf_model* m, const nsdf_camera* cams, int n_views, const uint32_t* d_matcap, int mw, int mh,
                  const nsdf_render_opts* opts, uint32_t* d_out, int W, int H, nsdf_render_stats* d_stats,
                  cudaStream_t stream) {
    if (m->in[0] != 3 && m->in[0] != 4) return NSDF_EUNSUPPORTED;
    // shapes other than (3|4) -> 32 x k -> 1: the generic fused evaluator, one thread per pixel (any widths up to 128)
    const bool generic = !m->fused;
    if (generic && !(m->generic && m->out[m->n_layers - 1] == 1)) return NSDF_EUNSUPPORTED;
    if (generic && n_views > 1) return NSDF_EUNSUPPORTED;
    if (n_views < 1 || n_views > 128) return NSDF_EINVAL;
    if (n_views > 1 && (long long)W * H > (1LL << 24)) return NSDF_EINVAL;   // RaySlot::pix carries the view in its top byte
    // the A/B schedules, the mixed-precision far field and the caller-set hand-over threshold exist in development builds only
    const unsigned int product_flags = NSDF_RENDER_SIMPLE | NSDF_RENDER_NO_QUIET_SMS | NSDF_RENDER_STAGED_STORES;
    if (((unsigned int)opts->flags & ~product_flags) != 0u || opts->evict_steps != 0) return NSDF_EUNSUPPORTED;
    if (opts->tile_world < 1 || opts->tile_rank < 0 || opts->tile_rank >= opts->tile_world) return NSDF_EINVAL;
    if (opts->color_type == NSDF_COLOR_MATCAP && (!d_matcap || mw <= 0 || mh <= 0)) return NSDF_EINVAL;
    if (opts->math != NSDF_MATH_REFERENCE && opts->math != NSDF_MATH_DETERMINISTIC) return NSDF_EINVAL;
    const bool simple = (opts->flags & NSDF_RENDER_SIMPLE) != 0;
    if (simple && n_views > 1) return NSDF_EUNSUPPORTED;

    RenderParams rp;
    std::memset(&rp, 0, sizeof(rp));
    std::memcpy(rp.cam.inv_view, cams->inv_view, sizeof(rp.cam.inv_view));
    std::memcpy(rp.cam.normal, cams->normal, sizeof(rp.cam.normal));
    rp.cam.frame = cams->frame;
    rp.cam.zoff = many_sphere_zoff(cams->frame);
    rp.n_views = n_views;
    rp.view_pixels = (unsigned int)((long long)W * H);
    rp.blob = m->d_blob;
    rp.n_sq = m->n_sq;
    rp.blob_bytes = m->blob_bytes;
    rp.matcap = d_matcap;
    rp.mw = mw;
    rp.mh = mh;
    rp.color_type = opts->color_type;
    rp.scene = opts->scene;
    if (opts->scene == NSDF_SCENE_PROGRAM) {
        const int rc_prog = scene_program_pack(opts->program, opts->n_ops, &rp.prog);
        if (rc_prog) return rc_prog;
    } else if (opts->scene != NSDF_SCENE_TANH && opts->scene != NSDF_SCENE_MANYSPHERE) {
        return NSDF_EINVAL;
    }
    rp.out = d_out;
    rp.W = W;
    rp.H = H;
    rp.tiles_x = (W + TILE - 1) / TILE;
    rp.n_tiles = rp.tiles_x * ((H + TILE - 1) / TILE);
    rp.tile_rank = opts->tile_rank;
    rp.tile_world = opts->tile_world;
    const int my_tiles = (rp.n_tiles - opts->tile_rank + opts->tile_world - 1) / opts->tile_world;
    if (my_tiles <= 0) return NSDF_OK;
    const long long n_jobs_ll = (long long)my_tiles * n_views;       // job = (view, tile of this call)
    if (n_jobs_ll * TILE_PIXELS >= (1LL << 32)) return NSDF_EINVAL;
    const int n_jobs = (int)n_jobs_ll;
    rp.my_tiles = my_tiles;
    rp.total_idx = (unsigned int)n_jobs * TILE_PIXELS;
    rp.max_steps = opts->max_steps;
    rp.n_in = m->in[0];
    rp.t4 = (float)cams->frame;
    rp.refmath = opts->math == NSDF_MATH_REFERENCE ? 1 : 0;
    rp.stats = d_stats;
    rp.evict_steps = 1 << 30;
    rp.evict_steps_hi = 1 << 30;
    nsdf_tuning T;
    resolve_tuning(opts, rp.total_idx, n_views, &T);

    int rc = ensure_scratch(m, 0, XQUEUE_ENTRIES * (size_t)n_views);
    if (rc) return rc;
    cudaError_t e = cudaSuccess;
    if (n_views > 1) {
        if ((size_t)n_views > m->views_capacity) {
            if (m->d_views) cudaFree(m->d_views);
            m->d_views = nullptr;
            m->views_capacity = 0;
            if (cudaMalloc(&m->d_views, (size_t)n_views * sizeof(Camera)) != cudaSuccess) return NSDF_ENOMEM;
            m->views_capacity = (size_t)n_views;
        }
        std::vector<Camera> hv((size_t)n_views);
        for (int v = 0; v < n_views; ++v) {
            std::memcpy(hv[v].inv_view, cams[v].inv_view, sizeof(hv[v].inv_view));
            std::memcpy(hv[v].normal, cams[v].normal, sizeof(hv[v].normal));
            hv[v].frame = cams[v].frame;
            hv[v].zoff = many_sphere_zoff(cams[v].frame);
        }
        e = cudaMemcpyAsync(m->d_views, hv.data(), hv.size() * sizeof(Camera), cudaMemcpyHostToDevice, stream);
        if (e != cudaSuccess) return (int)e;
        e = cudaStreamSynchronize(stream);   // hv is a local: the copy must have left it (a few KB, once per batch)
        if (e != cudaSuccess) return (int)e;
        rp.views = reinterpret_cast<const Camera*>(m->d_views);
    }
    // staged frame stores (render_kernels.cuh emit_pixel / flush_tile): for tile-partitioned calls, whose d_out is a peer
    // GPU's memory in the multi-GPU gather, or on request
    if (!simple && !generic && (opts->tile_world > 1 || (opts->flags & NSDF_RENDER_STAGED_STORES))) {
        if ((size_t)n_jobs > m->frame_stage_jobs) {
            if (m->d_frame_stage) cudaFree(m->d_frame_stage);
            m->d_frame_stage = nullptr;
            m->frame_stage_jobs = 0;
            if (cudaMalloc(&m->d_frame_stage, (size_t)n_jobs * (TILE_PIXELS + 1) * sizeof(uint32_t)) != cudaSuccess) return NSDF_ENOMEM;
            m->frame_stage_jobs = (size_t)n_jobs;
        }
        rp.stage = reinterpret_cast<uint32_t*>(m->d_frame_stage);
        rp.tile_done = reinterpret_cast<unsigned int*>(m->d_frame_stage) + m->frame_stage_jobs * TILE_PIXELS;
        e = cudaMemsetAsync(rp.tile_done, 0, (size_t)n_jobs * sizeof(unsigned int), stream);
        if (e != cudaSuccess) return (int)e;
    }
    rp.pixel_counter = m->d_counters;
    e = cudaMemsetAsync(m->d_counters, 0, 8 * sizeof(unsigned int), stream);
    if (e != cudaSuccess) return (int)e;

    const int sms = m->sm_count > 0 ? m->sm_count : 148;
    NSDF_PER_DEVICE_FLAG(attr_done);
    if (!attr_done) {
        e = cudaFuncSetAttribute(trace_simple_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
        if (e != cudaSuccess) return (int)e;
        e = cudaFuncSetAttribute(trace_coop_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        if (e != cudaSuccess) return (int)e;
        e = cudaFuncSetAttribute(tile_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
        if (e != cudaSuccess) return (int)e;
        attr_done = true;
    }
    if (generic) {
        GenericNet g;
        generic_net_of(m, &g);
        NSDF_PER_DEVICE_FLAG(gen_attr);
        if (!gen_attr) {
            e = cudaFuncSetAttribute(trace_generic_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)GENERIC_SMEM_LIMIT);
            if (e != cudaSuccess) return (int)e;
            gen_attr = true;
        }
        const unsigned int blocks = (rp.total_idx + 127u) / 128u;
        trace_generic_kernel<<<blocks, 128, g.staged ? (size_t)g.n_params * sizeof(float) : 0, stream>>>(rp, g);
        return (int)cudaGetLastError();
    }
    if (simple) {
        const unsigned int blocks = (rp.total_idx + 127u) / 128u;
        trace_simple_kernel<<<blocks, 128, m->blob_bytes, stream>>>(rp);
        return (int)cudaGetLastError();
    }

    // ---- tile priority: probe every 8th pixel for a few steps and trace the jobs (view, tile) with grazing rays first, so that
    // the call's longest dependent chains start at t = 0 and hide behind the bulk (tile_priority.cuh).  Order-free: same bytes.
    if (T.tile_priority_steps > 0 && n_jobs >= T.tile_priority_min_jobs) {
        if ((size_t)n_jobs > m->tile_scratch_capacity) {
            if (m->d_tile_scratch) cudaFree(m->d_tile_scratch);
            m->d_tile_scratch = nullptr;
            m->tile_scratch_capacity = 0;
            e = cudaMalloc(&m->d_tile_scratch, 2 * (size_t)n_jobs * sizeof(int));
            if (e != cudaSuccess) return (int)e;
            m->tile_scratch_capacity = (size_t)n_jobs;
        }
        int* d_score = m->d_tile_scratch;
        int* d_order = m->d_tile_scratch + m->tile_scratch_capacity;
        e = cudaMemsetAsync(d_score, 0, (size_t)n_jobs * sizeof(int), stream);
        if (e != cudaSuccess) return (int)e;
        const unsigned int probes = (unsigned int)n_jobs * PROBES_PER_TILE;
        tile_probe_kernel<<<(probes + 127u) / 128u, 128, m->blob_bytes, stream>>>(rp, n_jobs, T.tile_priority_steps, d_score);
        tile_sort_kernel<<<1, 256, 0, stream>>>(d_score, n_jobs, d_order);
        e = cudaGetLastError();
        if (e != cudaSuccess) return (int)e;
        rp.tile_order = d_order;
    }

    // ---- calls of many views: the four-rays-per-lane throughput tracer (trace_bulk4.cuh), one CTA of 256 threads per SM
    if (T.bulk4 > 0 && (size_t)rp.total_idx >= (size_t)sms * (B4_THREADS / 32) * 128u * 4u) {
        NSDF_PER_DEVICE_FLAG(b4_attr);
        const size_t smem = bulk4_smem(m->blob_bytes);
        const void* b4_kernel = T.bulk4_unroll == 1 ? (const void*)trace_bulk4_kernel<1>
                                : T.bulk4_unroll == 4 ? (const void*)trace_bulk4_kernel<4> : (const void*)trace_bulk4_kernel<2>;
        if (!b4_attr) {
            for (const void* k : {(const void*)trace_bulk4_kernel<1>, (const void*)trace_bulk4_kernel<2>, (const void*)trace_bulk4_kernel<4>}) {
                e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
                if (e != cudaSuccess) return (int)e;
            }
            b4_attr = true;
        }
        rc = ensure_scratch(m, B4_RING_ENTRIES, 0);
        if (rc) return rc;
        e = cudaMemsetAsync(m->d_queue, 0, (size_t)B4_RING_ENTRIES * sizeof(StragglerEntry), stream);   // no stale `ready` stamps
        if (e != cudaSuccess) return (int)e;
        rp.queue = reinterpret_cast<StragglerEntry*>(m->d_queue);
        rp.queue_capacity = B4_RING_ENTRIES;
        rp.evict_steps_hi = T.bulk4_evict_steps;
        rp.consumer_period = T.bulk4_period > 0 ? (unsigned int)T.bulk4_period : 0u;
        rp.hybrid = T.bulk4_drain;
        rp.b4_refill_min = T.bulk4_refill_min;
        void* args[] = {&rp};
        const unsigned int b4_grid = (T.max_ctas > 0 && T.max_ctas < sms) ? (unsigned int)T.max_ctas : (unsigned int)sms;
        return (int)cudaLaunchCooperativeKernel(b4_kernel, dim3(b4_grid), dim3(B4_THREADS), args, smem, stream);
    }

    // ---- one persistent, cooperative launch: one CTA per SM (trace_coop.cuh)
    int occ = 0;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, trace_coop_kernel, COOP_THREADS, coop_smem(m));
    if (e != cudaSuccess) return (int)e;
    if (occ < 1) return NSDF_EUNSUPPORTED;
    const unsigned int per_cta = (COOP_THREADS / 32) * COOP_R;
    unsigned int cap = (unsigned int)(sms * occ);
    if (T.max_ctas > 0 && (unsigned int)T.max_ctas < cap) cap = (unsigned int)T.max_ctas;
    const unsigned int want = (rp.total_idx + per_cta - 1) / per_cta;
    m->frame_seq = m->frame_seq + 1u ? m->frame_seq + 1u : 1u;
    rp.frame_seq = m->frame_seq;
    // Serial tail: every express_period-th CTA (= SM) is an express SM whose lanes hold ONE ray each; any warp
    // hands a ray that has marched express_steps steps to a free lane.  Only for grids that cover the machine (a smaller grid
    // could consist of express CTAs alone).
    const bool full_grid = want >= cap;
    const bool express = full_grid && !(opts->flags & NSDF_RENDER_NO_QUIET_SMS) && T.express_period > 0 && T.express_steps > 0;
    if (express) {
        rp.queue = reinterpret_cast<StragglerEntry*>(m->d_xqueue);   // non-null: enables the consumer machinery (nothing is pushed to it)
        rp.queue_capacity = 0u;
        rp.xqueue = reinterpret_cast<StragglerEntry*>(m->d_xqueue);
        rp.xqueue_soft_cap = (unsigned int)(XQUEUE_ENTRIES * (size_t)n_views / 2);
        rp.express_period = (unsigned int)T.express_period;
        rp.express_steps = T.express_steps;
        rp.express_warps = T.express_warps;
        rp.express_rays = T.express_rays;
        rp.express_solo_steps = T.express_solo_steps;
        rp.express_steps_old = T.express_old_steps;
        rp.express_check_old = T.express_old_mask;
        rp.express_from_bulk = T.express_from_bulk;
        rp.express_lanes = (cap / rp.express_period) * (unsigned int)(rp.express_warps * rp.express_rays);
        rp.tail_lanes = T.tail_lanes;
        rp.tail_push_max = T.tail_push_max;
    }
    return (int)launch_coop(rp, want < cap ? want : cap, coop_smem(m), stream);
}
#endif  // NSDF_DEV_SCHEDULES

// =====================================================================================
// Allocation-free variant (nsdf_render_noalloc).  No counterpart in the reference: added
// for callers that cannot allocate device memory (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
// The cooperative tracer's scratch -- its job counters and the tile-priority scores / order -- lives in ONE file-scope device
// array of fixed size instead of the model handle's cudaMalloc'ed buffers, which bounds the frame at NSDF_NOALLOC_MAX_TILES
// 64x64 tiles (3840x2160 fits).  No express lanes (their queue is 6 MB per view): the frame's bytes are the same -- the schedule
// never changes them -- but the serial tail of a frame is not shortened.  Every call on a device shares the array: the counters
// and scores are cleared on the stream at the start of each call; two calls that overlap on the GPU race.
// =====================================================================================
constexpr int NOALLOC_COUNTER_WORDS = 64;         // 8 used; the tile arrays start 256 bytes in
static __device__ __align__(256) unsigned int g_noalloc_scratch[NOALLOC_COUNTER_WORDS + 2 * NSDF_NOALLOC_MAX_TILES];

struct NoallocDevice {        // per device: set up on the first call with that device current
    unsigned int* scratch = nullptr;     // DEVICE address of g_noalloc_scratch (cudaGetSymbolAddress; the symbol itself is a host-side stand-in)
    int sms = 0;
    int occ = 0;
};
static NoallocDevice g_noalloc_dev[64];

int render_noalloc_launch(const nsdf_model* m, const nsdf_camera* cam, const uint32_t* d_matcap, int mw, int mh,
                          const nsdf_render_opts* opts, uint32_t* d_out, int W, int H, nsdf_render_stats* d_stats,
                          cudaStream_t stream) {
    if (!m->fused || (m->in[0] != 3 && m->in[0] != 4)) return NSDF_EUNSUPPORTED;
    if (opts->flags != 0 || opts->evict_steps != 0 || opts->tile_world != 1 || opts->tile_rank != 0) return NSDF_EUNSUPPORTED;
    if (opts->color_type == NSDF_COLOR_MATCAP && (!d_matcap || mw <= 0 || mh <= 0)) return NSDF_EINVAL;
    if (opts->math != NSDF_MATH_REFERENCE && opts->math != NSDF_MATH_DETERMINISTIC) return NSDF_EINVAL;
    const long long tiles_x = ((long long)W + TILE - 1) / TILE, tiles_y = ((long long)H + TILE - 1) / TILE;
    if (tiles_x * tiles_y > NSDF_NOALLOC_MAX_TILES) return NSDF_ETOOLARGE;

    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return (int)e;
    if (dev < 0 || dev >= 64) return NSDF_EUNSUPPORTED;
    NoallocDevice& D = g_noalloc_dev[dev];
    if (!D.scratch) {
        unsigned int* ptr = nullptr;
        e = cudaGetSymbolAddress((void**)&ptr, g_noalloc_scratch);
        if (e != cudaSuccess) return (int)e;
        if (!ptr) return NSDF_EUNSUPPORTED;
        e = cudaFuncSetAttribute(trace_coop_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        if (e != cudaSuccess) return (int)e;
        e = cudaFuncSetAttribute(tile_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
        if (e != cudaSuccess) return (int)e;
        int sms = 0, occ = 0;
        e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        if (e != cudaSuccess) return (int)e;
        e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, trace_coop_kernel, COOP_THREADS, coop_smem(m));
        if (e != cudaSuccess) return (int)e;
        if (sms < 1 || occ < 1) return NSDF_EUNSUPPORTED;
        D.sms = sms;
        D.occ = occ;
        D.scratch = ptr;
    }

    RenderParams rp;
    std::memset(&rp, 0, sizeof(rp));
    std::memcpy(rp.cam.inv_view, cam->inv_view, sizeof(rp.cam.inv_view));
    std::memcpy(rp.cam.normal, cam->normal, sizeof(rp.cam.normal));
    rp.cam.frame = cam->frame;
    rp.cam.zoff = many_sphere_zoff(cam->frame);
    rp.n_views = 1;
    rp.view_pixels = (unsigned int)((long long)W * H);
    rp.blob = m->d_blob;
    rp.n_sq = m->n_sq;
    rp.blob_bytes = m->blob_bytes;
    rp.matcap = d_matcap;
    rp.mw = mw;
    rp.mh = mh;
    rp.color_type = opts->color_type;
    rp.scene = opts->scene;
    if (opts->scene == NSDF_SCENE_PROGRAM) {
        const int rc_prog = scene_program_pack(opts->program, opts->n_ops, &rp.prog);
        if (rc_prog) return rc_prog;
    } else if (opts->scene != NSDF_SCENE_TANH && opts->scene != NSDF_SCENE_MANYSPHERE) {
        return NSDF_EINVAL;
    }
    rp.out = d_out;
    rp.W = W;
    rp.H = H;
    rp.tiles_x = (int)tiles_x;
    rp.n_tiles = (int)(tiles_x * tiles_y);
    rp.tile_rank = 0;
    rp.tile_world = 1;
    rp.my_tiles = rp.n_tiles;
    rp.total_idx = (unsigned int)rp.n_tiles * TILE_PIXELS;
    rp.max_steps = opts->max_steps;
    rp.n_in = m->in[0];
    rp.t4 = (float)cam->frame;
    rp.refmath = opts->math == NSDF_MATH_REFERENCE ? 1 : 0;
    rp.stats = d_stats;
    rp.evict_steps = 1 << 30;
    rp.evict_steps_hi = 1 << 30;
    rp.frame_seq = 1u;
    rp.pixel_counter = D.scratch;
    int* const d_score = reinterpret_cast<int*>(D.scratch + NOALLOC_COUNTER_WORDS);
    int* const d_order = d_score + NSDF_NOALLOC_MAX_TILES;

    nsdf_tuning T;
    nsdf_tuning_default(&T);
    if (opts->tuning) T = *opts->tuning;
    if (T.tile_priority_steps < 0) T.tile_priority_steps = 32;
    const int n_jobs = rp.n_tiles;
    const bool priority = T.tile_priority_steps > 0 && n_jobs >= T.tile_priority_min_jobs;
    // the array keeps the last call's contents: counters, and the scores the probe adds to, start at zero
    e = cudaMemsetAsync(D.scratch, 0, (size_t)(NOALLOC_COUNTER_WORDS + (priority ? n_jobs : 0)) * sizeof(unsigned int), stream);
    if (e != cudaSuccess) return (int)e;
    if (priority) {
        const unsigned int probes = (unsigned int)n_jobs * PROBES_PER_TILE;
        tile_probe_kernel<<<(probes + 127u) / 128u, 128, m->blob_bytes, stream>>>(rp, n_jobs, T.tile_priority_steps, d_score);
        tile_sort_kernel<<<1, 256, 0, stream>>>(d_score, n_jobs, d_order);
        e = cudaGetLastError();
        if (e != cudaSuccess) return (int)e;
        rp.tile_order = d_order;
    }
    const unsigned int per_cta = (COOP_THREADS / 32) * COOP_R;
    const unsigned int cap = (unsigned int)(D.sms * D.occ);
    const unsigned int want = (rp.total_idx + per_cta - 1) / per_cta;
    return (int)launch_coop(rp, want < cap ? want : cap, coop_smem(m), stream);
}

}  // namespace nsdf

extern "C" int nsdf_build_has_dev_schedules(void) {
#ifdef NSDF_DEV_SCHEDULES
    return 1;
#else
    return 0;
#endif
}

#ifdef NSDF_DEV_SCHEDULES
// Development hook (not part of include/nsdf.h): copies the tracer's timeline log to the host and clears it.
extern "C" int nsdf_dev_timeline_read(unsigned long long* out, int n_words) {
    const size_t total = sizeof(unsigned long long) * (4 + 4 * nsdf::TIMELINE_EVENTS);
    size_t bytes = sizeof(unsigned long long) * (size_t)n_words;
    if (bytes > total) bytes = total;
    cudaError_t e = cudaDeviceSynchronize();
    if (e == cudaSuccess) e = cudaMemcpyFromSymbol(out, nsdf::g_timeline, bytes);
    if (e == cudaSuccess) {
        static unsigned long long zeros[4] = {0, 0, 0, 0};
        e = cudaMemcpyToSymbol(nsdf::g_timeline, zeros, sizeof(zeros));
    }
    return (int)e;
}
#endif  // NSDF_DEV_SCHEDULES
