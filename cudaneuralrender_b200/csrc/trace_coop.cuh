// =====================================================================================
// trace_coop.cuh -- the default tracer: ONE persistent launch, every warp marches up to
// COOP_R = 16 rays with the warp-cooperative evaluators of mlp_fp32.cuh.
//
// Replaces the host loop of render_kernel (src/volumeRender_kernel.cu:608-692) and its
// per-iteration thrust::exclusive_scan + createBatch (:504-563): a warp's idle ray slots are
// refilled from the global pixel counter by ballot compaction, so there is no batch to build.
//
// Why a warp per ray group rather than a ray per lane: one march step of a ray is a chain of nine
// dependent layers.  With a ray per lane a step costs the whole warp ~9k instructions (~18 us on a
// busy SM), which turns the few rays that need thousands of steps into tens of
// milliseconds; cooperatively a layer is a 32-deep FFMA chain per output.  With 9..16 busy slots the
// warp runs the register-tiled 16-ray evaluator (mlp32_eval_warp16: 4 neurons x 4 rays per lane, ~13 us
// per step on a full SM); with fewer it drops to the lane = neuron evaluators for 8/4/2/1 rays, and a
// warp left with one ray runs lone_ray_march (1.15 us per step when its SM is otherwise quiet).
//
// Serial tail: "express lanes".  Every express_period-th CTA (= SM: one CTA per SM, cooperative launch) takes no pixels; its
// warps 0..3 hold ONE ray each, popped from the express queue.  Any other warp hands a ray that has marched
// express_steps steps to that queue when a lane is free (three L2 counters), and keeps it otherwise.
// The older scheme (quiet SMs consuming a main queue fed at evict_steps_hi steps) is still here for
// callers that set evict_steps, for the hybrid schedule and for the second pass of the mixed-precision
// tracer, whose source IS that queue.  When all pixels are taken every idle warp serves both queues.
// NSDF_COOP_TIMELINE=1 records when long rays are pushed/popped and how fast they step (g_timeline).
//
// Per ray the arithmetic is that of render_kernels.cuh (slot_start / slot_update): frames are
// bit-identical to the oracle whatever the schedule.
// =====================================================================================
#pragma once
#include "render_kernels.cuh"
#include "trace_fused.cuh"   // fused_push / fused_pop (the straggler queue)

namespace nsdf {

// Development timeline (NSDF_COOP_TIMELINE=1): [0] kernel start, [1] last warp out of pixels, [2] kernel end (globaltimer ns),
// [3] number of events, then events {time, pix, steps, kind}: kind 100 = pushed, 101 = popped, else rays in the warp.
#ifdef NSDF_DEV_SCHEDULES
#define NSDF_TIMELINE(rp) ((rp).timeline)
#else
#define NSDF_TIMELINE(rp) false      // product build: the timeline hooks and the A/B schedules are compiled out
#endif
constexpr int TIMELINE_EVENTS = 32768;
__device__ unsigned long long g_timeline[4 + 4 * TIMELINE_EVENTS];
__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ void timeline_event(int pix, int steps, int kind) {
    const unsigned long long i = atomicAdd(&g_timeline[3], 1ull);
    if (i < TIMELINE_EVENTS) {
        g_timeline[4 + 4 * i] = global_ns(); g_timeline[5 + 4 * i] = (unsigned long long)pix;
        g_timeline[6 + 4 * i] = (unsigned long long)steps; g_timeline[7 + 4 * i] = (unsigned long long)kind;
    }
}

#ifndef NSDF_COOP_THREADS
#define NSDF_COOP_THREADS 512
#endif
constexpr int COOP_THREADS = NSDF_COOP_THREADS;   // ONE CTA per SM (the launch pads its shared memory past half an SM): 16 warps, 4 per scheduler
#ifndef NSDF_COOP_R
#define NSDF_COOP_R 16
#endif
constexpr int COOP_R = NSDF_COOP_R;  // ray slots per warp (lanes 0..COOP_R-1 own them).  32 (every lane owns a ray, 8 x 4 register tile) was measured
                                     // SLOWER in the tracer: 1080p 19.2 -> 24.0 ms, batched turntable 10.9 -> 11.5 ms per frame (step latency doubles)
static_assert(COOP_R == 16 || COOP_R == 32, "COOP_R");
// Shared memory a CTA asks for at least, so that two can never share an SM: roles are per CTA (= per SM), and an express
// CTA's lone rays must have their schedulers to themselves.
constexpr size_t COOP_MIN_SMEM = 116 * 1024;

// Evaluates the K lowest-ranked active rays of the warp.  src[j] = lane of the j-th active ray.
template <int K>
__device__ __forceinline__ float coop_eval(const float* s_w, int n_sq, float* abuf, const f3& q, const int (&src)[8],
                                           int rank, int lane, float t4, bool in4) {
    float x[K][3], y[K];
#pragma unroll
    for (int j = 0; j < K; ++j) {
        x[j][0] = __shfl_sync(0xffffffffu, q.x, src[j]);
        x[j][1] = __shfl_sync(0xffffffffu, q.y, src[j]);
        x[j][2] = __shfl_sync(0xffffffffu, q.z, src[j]);
    }
    mlp32_eval_warp<K>(s_w, n_sq, abuf, x, y, lane, t4, in4);
    float mine = y[0];
#pragma unroll
    for (int j = 1; j < K; ++j) mine = (rank == j) ? y[j] : mine;
    return mine;
}

// A warp left with ONE marching ray and nothing to refill from: the serial tail of the frame (one
// plane_1 ray runs all 6000 steps).  Every lane keeps an identical copy of the ray state, so a march step is the
// network evaluation plus a handful of scalar instructions -- no ballots, shuffles, refill or queue traffic per step
// (measured: 1.62 us per step in the general loop).  Arithmetic is slot_update's, bit for bit; the owner lane takes the
// state back when the ray converges, leaves, hits the step cap, or after max_iter steps (so that queues get polled).
__device__ __forceinline__ void lone_ray_march(RaySlot& slot, int owner, const float* s_w, int n_sq, float* abuf,
                                               const RenderParams& rp, LocalStats& st, int lane, int max_iter) {
    f3 p, d;
    p.x = __shfl_sync(0xffffffffu, slot.p.x, owner); p.y = __shfl_sync(0xffffffffu, slot.p.y, owner);
    p.z = __shfl_sync(0xffffffffu, slot.p.z, owner);
    d.x = __shfl_sync(0xffffffffu, slot.d.x, owner); d.y = __shfl_sync(0xffffffffu, slot.d.y, owner);
    d.z = __shfl_sync(0xffffffffu, slot.d.z, owner);
    float tfar = __shfl_sync(0xffffffffu, slot.tfar, owner);
    int steps = __shfl_sync(0xffffffffu, slot.steps, owner);
    float rem = __shfl_sync(0xffffffffu, slot.rem, owner);
    const int view = slot_view(__shfl_sync(0xffffffffu, slot.pix, owner));
    int mode = 0;
    bool background = false, unfinished = false;
    unsigned int q = 0;
#pragma unroll 1
    for (int it = 0; it < max_iter; ++it) {
        const float x[1][3] = {{p.x, p.y, p.z}};
        float y[1];
        mlp32_eval_warp<1>(s_w, n_sq, abuf, x, y, lane, rp.t4, rp.n_in == 4);
        ++q;
#ifdef NSDF_DEBUG_TSCALE
        const float t = scene_eval(rp, view, p, y[0], rem) * NSDF_DEBUG_TSCALE + 2e-6f;
#else
        const float t = scene_eval(rp, view, p, y[0], rem);
#endif
        tfar = tfar - t;
        ++steps;
        if (tfar <= 0.0f) { background = true; mode = -1; break; }
        p = f3{fmaf(d.x, t, p.x), fmaf(d.y, t, p.y), fmaf(d.z, t, p.z)};
        rem = rem - clearance_step(t);
        if (t < NSDF_MARCHING_EPSILON) mode = 1;
        if (steps >= rp.max_steps) { unfinished = true; mode = -1; break; }
        if (mode != 0) break;
    }
    int flush_g = -1;
    if (lane == owner) {
        slot.p = p; slot.tfar = tfar; slot.steps = steps; slot.mode = mode; slot.rem = rem;
        st.queries += q;
        if (background || unfinished) flush_g = emit_pixel(rp, slot.pix, 0u);
        if (unfinished) ++st.unfinished;
    }
    flush_completed(rp, flush_g, lane);
}

__device__ __forceinline__ unsigned int sm_id() {
    unsigned int v;
    asm("mov.u32 %0, %%smid;" : "=r"(v));
    return v;
}

// rp.queue != nullptr enables the "quiet SM" scheme for the serial tail: rays that have marched
// rp.evict_steps_hi steps are pushed to the straggler queue and continued by the warps of every
// rp.consumer_period-th SM, which take no pixels of their own -- a lone long ray there is not slowed
// by busy neighbours.  Warps that run out of pixels join the consumers.  Termination: every warp has
// stopped producing (pixel_counter[3] == all warps) and the queue is empty.
__global__ void __launch_bounds__(COOP_THREADS, 1) trace_coop_kernel(const __grid_constant__ RenderParams rp) {
    extern __shared__ __align__(16) float s_w[];
    __shared__ uint64_t s_bar;
    stage_weights(s_w, rp.blob, rp.blob_bytes, &s_bar);

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (NSDF_TIMELINE(rp) && blockIdx.x == 0 && threadIdx.x == 0) g_timeline[0] = global_ns();
    float* abuf = s_w + (rp.blob_bytes >> 2) + warp * (2 * COOP_R * WARP_ACT_STRIDE);
    const unsigned int total_warps = gridDim.x * (COOP_THREADS / 32);
    const bool from_queue = rp.queue_static != 0;
    // second pass over a static queue: the consumer machinery runs only for the express lanes (nothing is pushed to the main queue)
    const bool use_queue = rp.queue != nullptr && (!from_queue || rp.xqueue != nullptr);
    unsigned int n_src = rp.total_idx;
    if (from_queue) {
        n_src = rp.pixel_counter[1];
        if (n_src > rp.queue_capacity) n_src = rp.queue_capacity;
    }
    unsigned int* const src_counter = from_queue ? rp.pixel_counter + 2 : rp.pixel_counter;

    RaySlot slot;
    slot.mode = -1;
    LocalStats st = {0, 0, 0, 0, 0, 0};
    // warp-uniform role state
    // consumers from the start: every warp of the quiet SMs, and the last rp.consumer_warps warps of every CTA ("fast lanes":
    // a warp that marches one or two long rays is latency-bound and takes few issue slots from the 16-ray warps beside it)
    // express SMs (trace timeline: a lone ray steps in 1.14 us on an SM where nothing else runs, in 3.5 us beside the crowded
    // warps of a quiet SM, in 5.5 us beside 16-ray warps): warps 0..3 of their CTAs hold ONE ray each, taken from the express queue
    // Roles go by CTA: the launch is cooperative with one CTA per SM, so CTA b has SM-wide resources to itself whichever SM it got.
    // An express CTA's lanes are its warps 0..express_warps-1: with 4 of them each sits alone on its scheduler (measured,
    // profiles/r02_lone_bench.txt: 1836 clocks per dependent evaluation with one lone warp per scheduler, 3097 with two).
    const bool express_sm = use_queue && rp.xqueue && rp.express_period && (blockIdx.x % rp.express_period) == rp.express_period - 1;
    const bool express_lane = express_sm && warp < rp.express_warps;
    bool producing = !(use_queue && (express_sm || (rp.consumer_period && (blockIdx.x % rp.consumer_period) == rp.consumer_period - 1) ||
                                     warp >= COOP_THREADS / 32 - rp.consumer_warps));
    bool bulk_over = false;   // warp-uniform
    int xcount = 0;           // express rays this lane holds (counted in pc[6]), warp-uniform
    unsigned int* const pc = rp.pixel_counter;
    const int pop_max = rp.pop_max > 0 && rp.pop_max < COOP_R ? rp.pop_max : COOP_R;
    bool exhausted = !producing;
    if (use_queue && !producing && lane == 0) atomicAdd(rp.pixel_counter + 3, 1u);
    int poll = 0;
    unsigned int idle = 0;

#ifdef NSDF_DEV_SCHEDULES
    // ---- hybrid schedule: the bulk of the frame on the ray-per-lane evaluator, long rays to the quiet SMs
    if (rp.hybrid && use_queue && producing) {
        RaySlot ps;
        ps.mode = -1;
        bool dry = false;   // warp-uniform
        for (;;) {
            while (!dry) {
                const bool need = ps.mode < 0;
                const unsigned int m = __ballot_sync(0xffffffffu, need);
                if (!m) break;
                const int leader = __ffs(m) - 1;
                const unsigned int cnt = __popc(m);
                unsigned int base = 0;
                if (lane == leader) base = atomicAdd(rp.pixel_counter, cnt);
                base = __shfl_sync(0xffffffffu, base, leader);
                if (need) {
                    const unsigned int idx = base + __popc(m & lanemask_lt());
                    if (idx < rp.total_idx) slot_start(ps, idx, rp, st);
                }
                if (base + cnt >= rp.total_idx) dry = true;
            }
            if (!__ballot_sync(0xffffffffu, ps.mode >= 0)) break;
            const f3 q = ps.mode >= 0 ? slot_query_point(ps) : f3{0.0f, 0.0f, 0.0f};
            float x[1][3] = {{q.x, q.y, q.z}}, y[1];
            mlp32_eval_thread<1>(s_w, rp.n_sq, x, y, rp.t4, rp.n_in == 4);
            if (ps.mode >= 0) slot_update(ps, y[0], q, rp, st);
            const int n_act = __popc(__ballot_sync(0xffffffffu, ps.mode >= 0));
            const bool drain = dry && n_act <= rp.hybrid;
            // a long ray leaves for the quiet SMs only while their backlog is short: marching on in a full warp costs nothing
            // extra, waiting in the queue does (looked at every 8th step: the counters live in L2)
            bool longray = ps.mode == 0 && ps.steps >= rp.evict_steps_hi && (rp.queue_room == 0u || (ps.steps & 7) == 0);
            if (rp.queue_room && !drain && __ballot_sync(0xffffffffu, longray)) {
                unsigned int backlog = 0;
                if (lane == 0) backlog = ld_volatile_u32(rp.pixel_counter + 1) - ld_volatile_u32(rp.pixel_counter + 2);
                backlog = __shfl_sync(0xffffffffu, backlog, 0);
                if ((int)backlog >= (int)rp.queue_room) longray = false;
            }
            fused_push(ps, ps.mode == 0 && (drain || longray), rp, st);
        }
        producing = false;
        exhausted = true;
        if (lane == 0) { __threadfence(); atomicAdd(rp.pixel_counter + 3, 1u); }
    }
#endif

    for (;;) {
        // ---- refill idle ray slots (ballot compaction over the slot lanes); not while a long ray is on board
        const bool hold = rp.isolate_steps > 0 &&
                          __ballot_sync(0xffffffffu, lane < COOP_R && slot.mode == 0 && slot.steps >= rp.isolate_steps) != 0u;
        int flush_g = -1;    // staged stores: job whose tile this lane completed in this iteration
        while (!exhausted && !hold) {
            flush_completed(rp, flush_g, lane);
            flush_g = -1;
            const bool need = lane < COOP_R && slot.mode < 0;
            const unsigned int m = __ballot_sync(0xffffffffu, need);
            if (!m) break;
            const int leader = __ffs(m) - 1;
            const unsigned int cnt = __popc(m);
            unsigned int base = 0;
            if (lane == leader) base = atomicAdd(src_counter, cnt);
            base = __shfl_sync(0xffffffffu, base, leader);
            if (need) {
                const unsigned int idx = base + __popc(m & lanemask_lt());
                if (idx < n_src) {
                    if (!from_queue) {
                        slot_start(slot, idx, rp, st, flush_g);
                    } else {   // a ray the far-field pass handed over: continue it from its current point
                        const float4* src = reinterpret_cast<const float4*>(rp.queue + idx);
                        const float4 e0 = __ldcg(src), e1 = __ldcg(src + 1);
                        const float tf = __ldcg(reinterpret_cast<const float*>(src + 2));
                        slot.pix = __float_as_int(e0.x); slot.steps = __float_as_int(e0.y);
                        slot.p = f3{e0.z, e0.w, e1.x}; slot.d = f3{e1.y, e1.z, e1.w};
                        slot.tfar = tf; slot.mode = 0; slot.rem = -1.0f;
                    }
                }
            }
            if (base + cnt >= n_src) {
                exhausted = true;
                if (NSDF_TIMELINE(rp) && lane == 0) atomicMax(&g_timeline[1], global_ns());
            }
        }
        flush_completed(rp, flush_g, lane);
        flush_g = -1;
        unsigned int act = __ballot_sync(0xffffffffu, lane < COOP_R && slot.mode >= 0);
        if (rp.xqueue && rp.express_steps > 0 && (!producing || rp.express_from_bulk) && !express_sm && !bulk_over) {
            // rays that have come this far move to an express lane, if one is free
            // (looked at when a ray's step count is a multiple of 16: the counters live in L2; a lone ray comes back every 64 steps)
            const bool go = lane < COOP_R && slot.mode == 0 && slot.steps >= rp.express_steps &&
                            (slot.steps & (slot.steps >= rp.express_steps_old ? rp.express_check_old : 15)) == 0;
            const unsigned int gm = __ballot_sync(0xffffffffu, go);
            if (gm) {
                unsigned int ok = 0, over = 0;
                if (lane == 0) {
                    const unsigned int tail2 = ld_volatile_u32(pc + 4), head2 = ld_volatile_u32(pc + 5);
                    const unsigned int infl = ld_volatile_u32(pc + 6);
                    over = ld_volatile_u32(pc + 3) >= total_warps ? 1u : 0u;
                    ok = (!over && (tail2 - head2) + infl + __popc(gm) <= rp.express_lanes &&
                          tail2 + 32u <= rp.xqueue_soft_cap) ? 1u : 0u;
                }
                ok = __shfl_sync(0xffffffffu, ok, 0);
                bulk_over = __shfl_sync(0xffffffffu, over, 0) != 0u;
                if (ok) {
                    if (NSDF_TIMELINE(rp) && go) timeline_event(slot.pix, slot.steps, 102);
                    queue_push(slot, go, rp.xqueue, pc + 4, rp, st);
                    if (!exhausted) continue;      // a producing warp refills the slots it has just emptied
                    act = __ballot_sync(0xffffffffu, lane < COOP_R && slot.mode >= 0);
                }
            }
        }
        // ---- tail hand-over: when the pixels are gone, a frame's last milliseconds are a few hundred long rays scattered over all
        // warps, 16 lone warps per SM stepping at a third of the speed four would (lone_bench: 6081 vs 1836 clocks per step).
        // A warp that is not a lane (warp >= tail_lanes) and is down to tail_push_max marching rays hands them to the express
        // queue; the lanes of all CTAs take them one at a time.
        if (rp.xqueue && rp.tail_lanes > 0 && exhausted && !express_lane && warp >= rp.tail_lanes && act &&
            __popc(act) <= rp.tail_push_max) {
            const bool go = lane < COOP_R && slot.mode == 0;
            const unsigned int gm = __ballot_sync(0xffffffffu, go);
            if (gm) {
                unsigned int ok = 0;
                if (lane == 0) ok = (ld_volatile_u32(pc + 4) + 32u <= rp.xqueue_soft_cap) ? 1u : 0u;
                if (__shfl_sync(0xffffffffu, ok, 0)) {
                    if (NSDF_TIMELINE(rp) && go) timeline_event(slot.pix, slot.steps, 104);
                    queue_push(slot, go, rp.xqueue, pc + 4, rp, st);
                    act = __ballot_sync(0xffffffffu, lane < COOP_R && slot.mode >= 0);
                }
            }
        }
        if (use_queue && exhausted) {
            if (producing && !act) {          // out of pixels and rays: from now on a consumer
                producing = false;
                if (lane == 0) { __threadfence(); atomicAdd(rp.pixel_counter + 3, 1u); }
            }
            // a consumer warp that holds a marathon ray (>= rp.marathon_steps steps) takes no more rays: its other rays
            // finish within a few hundred steps and the marathon ray then marches alone at the 1-ray latency instead of
            // sharing every step with up to 15 newcomers for as long as the queue has a backlog
            const bool marathon = rp.marathon_steps > 0 &&
                __ballot_sync(0xffffffffu, lane < COOP_R && slot.mode == 0 && slot.steps >= rp.marathon_steps) != 0u;
            if (!producing && !marathon && (!act || --poll <= 0)) {
                poll = 8;
                if (express_sm) {
                    // an express lane marches up to express_rays rays together (lone_bench: 1836 clocks per step for one ray, 2594
                    // for two, 3204 for four: four rays in a lane advance 2.3 x as many steps per second as one), but takes no
                    // company once one of its rays has passed express_solo_steps: the frame's longest chains run alone
                    const int now = __popc(act);
                    if (xcount > now) {                        // express rays of this lane that have finished
                        if (lane == 0) atomicSub(pc + 6, (unsigned int)(xcount - now));
                        xcount = now;
                    }
                    const bool solo = rp.express_solo_steps > 0 &&
                        __ballot_sync(0xffffffffu, lane < COOP_R && slot.mode == 0 && slot.steps >= rp.express_solo_steps) != 0u;
                    if (express_lane && now < rp.express_rays && !solo) {
                        const bool popped = queue_pop(slot, lane < rp.express_rays && slot.mode < 0, rp.xqueue, pc + 4, pc + 5, rp);
                        if (NSDF_TIMELINE(rp) && popped) timeline_event(slot.pix, slot.steps, 103);
                        const int got = __popc(__ballot_sync(0xffffffffu, popped));
                        if (got) {
                            xcount += got;
                            if (lane == 0) atomicAdd(pc + 6, (unsigned int)got);
                        }
                    }
                    // once every warp is out of pixels an express SM consumes like any other
                    if (bulk_over && !__ballot_sync(0xffffffffu, lane < COOP_R && slot.mode >= 0)) {
                        fused_pop(slot, lane < pop_max && slot.mode < 0, rp);
                        if (!express_lane && warp < rp.tail_lanes && !__ballot_sync(0xffffffffu, lane < COOP_R && slot.mode >= 0))
                            queue_pop(slot, lane == 0, rp.xqueue, pc + 4, pc + 5, rp);
                    }
                } else {
                    const bool popped = fused_pop(slot, lane < pop_max && slot.mode < 0, rp);
                    if (NSDF_TIMELINE(rp) && popped) timeline_event(slot.pix, slot.steps, 101);
                    // tail of the frame: once every warp is out of pixels, warps 0..tail_lanes-1 of EVERY CTA are lanes for the
                    // express queue (one lone ray each: up to four per SM step at full speed, lone_bench), fed by tail_handover below
                    if (rp.xqueue && bulk_over && warp < rp.tail_lanes && !__ballot_sync(0xffffffffu, lane < COOP_R && slot.mode >= 0))
                        queue_pop(slot, lane == 0, rp.xqueue, pc + 4, pc + 5, rp);
                }
                act = __ballot_sync(0xffffffffu, lane < COOP_R && slot.mode >= 0);
            }
            if (!act) {
                unsigned int done = 0, over = 0;
                if (lane == 0) {
                    const unsigned int fin = ld_volatile_u32(pc + 3);
                    const unsigned int tail = ld_volatile_u32(pc + 1);
                    const unsigned int head = ld_volatile_u32(pc + 2);
                    over = fin >= total_warps ? 1u : 0u;
                    done = (over && head >= tail) ? 1u : 0u;
                    if (done && rp.xqueue) done = ld_volatile_u32(pc + 5) >= ld_volatile_u32(pc + 4) ? 1u : 0u;
                }
                bulk_over = __shfl_sync(0xffffffffu, over, 0) != 0u;
                if (__shfl_sync(0xffffffffu, done, 0)) break;
                if (bulk_over) poll = 0;                   // look at the queues again right away
                __nanosleep(express_sm && !express_lane && !bulk_over ? 2000 : 200);
                if (++idle > 50000000u) __trap();     // > 10 s without work: never hang the device
                continue;
            }
            idle = 0;
        } else if (!act) {
            break;
        }
        if (__popc(act) == 1 && (exhausted || hold)) {
            const int owner = __ffs(act) - 1;
            if (__shfl_sync(0xffffffffu, slot.mode, owner) == 0) {
                // chunks end on multiples of 64 steps (the queues and the express hand-over are looked at in between)
                const int chunk = 64 - (__shfl_sync(0xffffffffu, slot.steps, owner) & 63);
                lone_ray_march(slot, owner, s_w, rp.n_sq, abuf, rp, st, lane, use_queue ? chunk : (1 << 30));
                if (NSDF_TIMELINE(rp) && lane == owner && (slot.steps & 255) < 64 && slot.steps >= 256) timeline_event(slot.pix, slot.steps, 1);
                continue;
            }
        }
        const bool active = lane < COOP_R && slot.mode >= 0;
        const int na = __popc(act);
        const int rank = __popc(act & lanemask_lt());
        const f3 q = active ? slot_query_point(slot) : f3{0.0f, 0.0f, 0.0f};
        float mine;
#if NSDF_COOP_R == 32
        if (na > 16) {
            // 17..32 busy slots: the 8-neuron x 4-ray register tile over 32 points (mlp32_eval_tile<4>): 1.5 B of shared memory
            // per FFMA instead of 2 (dense 3.19 against 2.84 Gq/s), and every lane owns a ray, so the scene function and the
            // ray update run at full lane efficiency
            float4* xin = reinterpret_cast<float4*>(abuf + 32 * WARP_ACT_STRIDE);
            if (active) xin[rank] = make_float4(q.x, q.y, q.z, 0.0f);
            else xin[na + __popc(~act & lanemask_lt())] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);   // idle lanes zero the unused ranks
            float y[1];
            mlp32_eval_tile<4>(s_w, rp.n_sq, abuf, lane, y, rp.t4, rp.n_in == 4);
            mine = __shfl_sync(0xffffffffu, y[0], rank);
        } else
#endif
        if (na > 8) {
            // 9..16 busy slots: the register-tiled 16-point evaluator.  Owners write their query point at their rank,
            // unused ranks are zeroed (their results are computed and dropped).
            float4* xin = reinterpret_cast<float4*>(abuf + 16 * WARP_ACT_STRIDE);
            if (active) xin[rank] = make_float4(q.x, q.y, q.z, 0.0f);
            else if (lane < 16 && lane >= na) xin[lane] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
            const float out = mlp32_eval_warp16(s_w, rp.n_sq, abuf, lane, rp.t4, rp.n_in == 4);
            mine = __shfl_sync(0xffffffffu, out, rank & 15);
        } else {
            int src[8];
            unsigned int mm = act;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                src[j] = mm ? (__ffs(mm) - 1) : 0;
                mm &= mm - 1u;
            }
            if (na > 4) mine = coop_eval<8>(s_w, rp.n_sq, abuf, q, src, rank, lane, rp.t4, rp.n_in == 4);
            else if (na > 2) mine = coop_eval<4>(s_w, rp.n_sq, abuf, q, src, rank, lane, rp.t4, rp.n_in == 4);
            else if (na == 2) mine = coop_eval<2>(s_w, rp.n_sq, abuf, q, src, rank, lane, rp.t4, rp.n_in == 4);
            else mine = coop_eval<1>(s_w, rp.n_sq, abuf, q, src, rank, lane, rp.t4, rp.n_in == 4);
        }
        if (active) slot_update(slot, mine, q, rp, st, flush_g);
        flush_completed(rp, flush_g, lane);
        if (NSDF_TIMELINE(rp) && active && slot.mode == 0 && slot.steps >= 256 && (slot.steps & 255) == 0) timeline_event(slot.pix, slot.steps, na);
        // ---- long rays leave for the quiet SMs
        if (use_queue && producing)
        {
            const bool leave = lane < COOP_R && slot.mode == 0 && slot.steps >= rp.evict_steps_hi;
            if (NSDF_TIMELINE(rp) && leave) timeline_event(slot.pix, slot.steps, 100);
            fused_push(slot, leave, rp, st);
        }
    }
    if (NSDF_TIMELINE(rp) && lane == 0) atomicMax(&g_timeline[2], global_ns());
    flush_stats(st, rp.stats);
}

}  // namespace nsdf
