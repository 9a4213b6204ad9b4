// =====================================================================================
// trace_fused.cuh -- ONE persistent launch per frame: throughput lanes and latency warps
// run side by side, so the serial tail of long rays overlaps the bulk.
//
// Replaces the host loop of render_kernel (src/volumeRender_kernel.cu:608-692).  Every warp is
// first a PRODUCER (one ray per lane, mlp32_eval_thread, lanes refilled from the global pixel
// counter by ballot compaction) unless it is one of the few warps that are CONSUMERS from the
// start; a producer whose pixels are exhausted hands its last rays over and becomes a consumer.
// CONSUMERS pop rays from the straggler queue and march R of them per warp with the
// warp-cooperative evaluator (lane = neuron, mlp32_eval_warp): ~10x lower latency per step.
//
//   producer -> queue : a ray that has marched `evict_steps` steps, or the remains of a
//                       draining warp.  Slots are reserved with one atomicAdd per warp, filled,
//                       then published by writing frame_seq into the entry's `ready` word.
//   queue -> consumer : head is advanced with atomicCAS up to the reserved tail; the consumer
//                       lane then waits for its entry's `ready` stamp.
//   termination       : consumers leave when every producer warp has finished
//                       (pixel_counter[3] == total producer warps) and head == tail.
// The grid never exceeds the number of co-resident CTAs, so the waiting consumers cannot starve
// a producer that has not started.
//
// Per ray the arithmetic is that of render_kernels.cuh (slot_start / slot_update), whatever warp
// evaluates it: frames are bit-identical to the oracle for every schedule.
// =====================================================================================
#pragma once
#include "render_kernels.cuh"

namespace nsdf {

constexpr int FUSED_THREADS = 256;
constexpr int FUSED_R = 4;   // rays per consumer warp

__device__ __forceinline__ unsigned int ld_volatile_u32(const unsigned int* p) {
    unsigned int v;
    asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// Producer side: append marching rays to the queue (warp-aggregated) and publish them.
__device__ __forceinline__ void queue_push(RaySlot& s, bool do_evict, StragglerEntry* queue, unsigned int* tail_counter,
                                           const RenderParams& rp, LocalStats& st) {
    const unsigned int m = __ballot_sync(0xffffffffu, do_evict);
    if (!m) return;
    const int leader = __ffs(m) - 1;
    unsigned int base = 0;
    if ((int)(threadIdx.x & 31) == leader) base = atomicAdd(tail_counter, (unsigned int)__popc(m));
    base = __shfl_sync(0xffffffffu, base, leader);
    if (do_evict) {
        const unsigned int slot = base + __popc(m & lanemask_lt());
        // main queue: capacity == number of pixels of this call, every ray is pushed at most once
        float4* dst = reinterpret_cast<float4*>(queue + slot);
        dst[0] = make_float4(__int_as_float(s.pix), __int_as_float(s.steps), s.p.x, s.p.y);
        dst[1] = make_float4(s.p.z, s.d.x, s.d.y, s.d.z);
        reinterpret_cast<float2*>(dst + 2)[0] = make_float2(s.tfar, 0.0f);
        __threadfence();
        asm volatile("st.volatile.global.u32 [%0], %1;" ::"l"(&queue[slot].ready), "r"(rp.frame_seq) : "memory");
        if (queue == rp.queue) ++st.evicted;
        s.mode = -1;
    }
}
__device__ __forceinline__ void fused_push(RaySlot& s, bool do_evict, const RenderParams& rp, LocalStats& st) {
    queue_push(s, do_evict, rp.queue, rp.pixel_counter + 1, rp, st);
}

// Consumer side: lanes with want == true try to pop one entry each.  Returns true in the lanes that got one.
__device__ __forceinline__ bool queue_pop(RaySlot& s, bool want, const StragglerEntry* queue, unsigned int* tail_counter,
                                          unsigned int* head_counter, const RenderParams& rp) {
    const unsigned int m = __ballot_sync(0xffffffffu, want);
    if (!m) return false;
    const int leader = __ffs(m) - 1;
    unsigned int base = 0, got = 0;
    if ((int)(threadIdx.x & 31) == leader) {
        const unsigned int cnt = __popc(m);
        unsigned int head = ld_volatile_u32(head_counter);
        for (;;) {
            const unsigned int tail = ld_volatile_u32(tail_counter);
            if (head >= tail) { got = 0; break; }
            const unsigned int take = (tail - head) < cnt ? (tail - head) : cnt;
            const unsigned int old = atomicCAS(head_counter, head, head + take);
            if (old == head) { base = head; got = take; break; }
            head = old;
        }
    }
    base = __shfl_sync(0xffffffffu, base, leader);
    got = __shfl_sync(0xffffffffu, got, leader);
    const unsigned int rank = __popc(m & lanemask_lt());
    const bool mine = want && rank < got;
    if (mine) {
        const StragglerEntry* e = queue + base + rank;
        unsigned int spins = 0;
        while (ld_volatile_u32(&e->ready) != rp.frame_seq) {
            __nanosleep(20);
            if (++spins > 100000000u) __trap();   // a reserved entry is published within microseconds
        }
        __threadfence();
        const float4* src = reinterpret_cast<const float4*>(e);
        const float4 e0 = __ldcg(src), e1 = __ldcg(src + 1);
        const float tf = __ldcg(reinterpret_cast<const float*>(src + 2));
        s.pix = __float_as_int(e0.x); s.steps = __float_as_int(e0.y);
        s.p = f3{e0.z, e0.w, e1.x}; s.d = f3{e1.y, e1.z, e1.w};
        s.tfar = tf; s.mode = 0; s.rem = -1.0f;
    }
    return mine;
}
__device__ __forceinline__ bool fused_pop(RaySlot& s, bool want, const RenderParams& rp) {
    return queue_pop(s, want, rp.queue, rp.pixel_counter + 1, rp.pixel_counter + 2, rp);
}

#ifdef NSDF_DEV_SCHEDULES   // the producer/consumer A/B schedule; the queue helpers above serve the default tracer
__global__ void __launch_bounds__(FUSED_THREADS, 2) trace_fused_kernel(const __grid_constant__ RenderParams rp) {
    extern __shared__ __align__(16) float s_w[];
    __shared__ uint64_t s_bar;
    stage_weights(s_w, rp.blob, rp.blob_bytes, &s_bar);

    const unsigned int lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const unsigned int gwarp = blockIdx.x * (FUSED_THREADS / 32) + warp;
    const unsigned int total_warps = gridDim.x * (FUSED_THREADS / 32);
    const unsigned int cp = rp.consumer_period;
    const bool start_as_consumer = rp.queue && cp && (gwarp % cp == cp - 1);
    const unsigned int total_producers = (rp.queue && cp) ? total_warps - total_warps / cp : total_warps;
    float* abuf = s_w + (rp.blob_bytes >> 2) + warp * (2 * FUSED_R * WARP_ACT_STRIDE);

    RaySlot slot;
    slot.mode = -1;
    LocalStats st = {0, 0, 0, 0, 0, 0};

    if (!start_as_consumer) {
        // ------------------------------------------------------------------ producer
        bool exhausted = false;   // warp-uniform
        bool room = true;         // warp-uniform: the consumers' backlog is short (refreshed every 4 steps)
        unsigned int it = 0;
        for (;;) {
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
            const unsigned int act = __ballot_sync(0xffffffffu, slot.mode >= 0);
            if (!act) break;
            const f3 q = (slot.mode >= 0) ? slot_query_point(slot) : f3{0.0f, 0.0f, 0.0f};
            float x[1][3] = {{q.x, q.y, q.z}}, y[1];
            mlp32_eval_thread<1>(s_w, rp.n_sq, x, y, rp.t4, rp.n_in == 4);
            if (slot.mode >= 0) slot_update(slot, y[0], q, rp, st);
            if (rp.queue) {
                if ((++it & 3u) == 0u) {
                    unsigned int r = 0;
                    if (lane == 0)
                        r = (ld_volatile_u32(rp.pixel_counter + 1) - ld_volatile_u32(rp.pixel_counter + 2)) < rp.queue_room;
                    room = __shfl_sync(0xffffffffu, r, 0) != 0u;
                }
                const int n_act = __popc(__ballot_sync(0xffffffffu, slot.mode >= 0));
                const bool drain = exhausted && n_act <= 8;
                const bool longray = slot.steps >= rp.evict_steps_hi || (room && slot.steps >= rp.evict_steps);
                fused_push(slot, slot.mode == 0 && (drain || longray), rp, st);
            }
        }
        if (rp.queue && lane == 0) {
            __threadfence();
            atomicAdd(rp.pixel_counter + 3, 1u);
        }
    }
    if (rp.queue) {
        // ------------------------------------------------------------------ consumer
        slot.mode = -1;
        unsigned int idle = 0;
        int poll = 0;
        for (;;) {
            unsigned int act = __ballot_sync(0xffffffffu, lane < FUSED_R && slot.mode >= 0);
            // the queue lives in L2 (~700 cycles per look): look every step only while idle
            if (!act || --poll <= 0) {
                fused_pop(slot, lane < FUSED_R && slot.mode < 0, rp);
                poll = 8;
                act = __ballot_sync(0xffffffffu, lane < FUSED_R && slot.mode >= 0);
            }
            if (!act) {
                unsigned int done = 0;
                if (lane == 0) {
                    const unsigned int fin = ld_volatile_u32(rp.pixel_counter + 3);
                    const unsigned int tail = ld_volatile_u32(rp.pixel_counter + 1);
                    const unsigned int head = ld_volatile_u32(rp.pixel_counter + 2);
                    done = (fin >= total_producers && head >= tail) ? 1u : 0u;
                }
                done = __shfl_sync(0xffffffffu, done, 0);
                if (done) break;
                __nanosleep(200);
                if (++idle > 50000000u) __trap();     // > 10 s without work: never hang the device
                continue;
            }
            idle = 0;
            const bool mine_active = lane < FUSED_R && slot.mode >= 0;
            const f3 q = mine_active ? slot_query_point(slot) : f3{0.0f, 0.0f, 0.0f};
            float mine;
            if (__popc(act) == 1) {
                // lone ray (the serial tail): one chain per lane instead of FUSED_R
                const int src = __ffs(act) - 1;
                float x[1][3], y[1];
                x[0][0] = __shfl_sync(0xffffffffu, q.x, src);
                x[0][1] = __shfl_sync(0xffffffffu, q.y, src);
                x[0][2] = __shfl_sync(0xffffffffu, q.z, src);
                mlp32_eval_warp<1>(s_w, rp.n_sq, abuf, x, y, (int)lane, rp.t4, rp.n_in == 4);
                mine = y[0];
            } else {
                float x[FUSED_R][3], y[FUSED_R];
#pragma unroll
                for (int r = 0; r < FUSED_R; ++r) {
                    x[r][0] = __shfl_sync(0xffffffffu, q.x, r);
                    x[r][1] = __shfl_sync(0xffffffffu, q.y, r);
                    x[r][2] = __shfl_sync(0xffffffffu, q.z, r);
                }
                mlp32_eval_warp<FUSED_R>(s_w, rp.n_sq, abuf, x, y, (int)lane, rp.t4, rp.n_in == 4);
                mine = y[0];
#pragma unroll
                for (int r = 1; r < FUSED_R; ++r) mine = ((int)lane == r) ? y[r] : mine;
            }
            if (mine_active) slot_update(slot, mine, q, rp, st);
        }
    }
    flush_stats(st, rp.stats);
}

#endif  // NSDF_DEV_SCHEDULES

}  // namespace nsdf
