// =====================================================================================
// trace_bulk4.cuh -- the THROUGHPUT tracer for calls of many views (nsdf_render_views: >= 4 views and >= 12 M pixels).
//
// Replaces, like trace_coop.cuh, the host loop of render_kernel (src/volumeRender_kernel.cu:608-692) and its
// per-iteration scan + createBatch (:504-563).  A call of many views is a pure throughput problem: the serial tails of
// its views hide behind each other's bulk, so what counts is FFMA-pipe utilisation of the bulk, and the 16-ray warp tiles
// of trace_coop.cuh are bound by shared-memory bandwidth (2 B per FFMA; ncu: LSU 81 %, FMA 52 %).  Here the bulk runs on
// the evaluator that is FMA-bound on dense batches (mlp32_eval_thread_s<4>: one lane = FOUR rays x all 32 neurons, the
// weights as warp-uniform LDS.128, 0.6 B per FFMA, 67.5 % of the FFMA pipe on the 256^3 sweep):
//
//   producer warps   128 ray slots per warp (4 per lane, their state in a warp-private shared-memory block).  Idle slots are refilled
//                    from the global job counter with ONE atomic per warp (prefix sum over the lanes' idle counts) once
//                    `b4_refill_min` (16) of them are free.  Slot work runs in rolled loops over the four slot rows (one copy of
//                    slot_start / slot_update in the instruction stream).  A warp step costs ~40 us whatever the occupancy, so a
//                    warp that is out of pixels and down to `rp.hybrid` (16) rays hands them to a ring queue in L2; rays may also
//                    leave after `evict_steps_hi` steps (off by default: with many views the tails hide anyway);
//   consumer warps   (every producer warp once it has drained; optionally all warps of every `consumer_period`-th CTA from the
//                    start) pop rays from the ring and march them 16 / 8 / 4 / 2 / 1 at a time with the cooperative evaluators of
//                    trace_coop.cuh (a lone ray: ~1.3 us per dependent step).
//
// Per ray the arithmetic is slot_start / slot_update of render_kernels.cuh, whoever evaluates it: frames are bit-identical
// to trace_coop_kernel's, to the oracle (deterministic math) and to the reference binary's (reference math).
// One CTA of 256 threads per SM (cooperative launch): 214 KB of shared memory, up to 255 registers per thread.
// =====================================================================================
#pragma once
#include "render_kernels.cuh"
#include "trace_coop.cuh"

namespace nsdf {

constexpr int B4_THREADS = 256;
constexpr int B4_P = 4;                                   // ray slots per lane
constexpr int B4_WORDS = 14;                              // words of a RaySlot
constexpr size_t B4_ACT_BYTES = 32 * 32 * sizeof(float4); // per warp: 32 k x 32 lanes x float4 (mlp32_eval_thread_s<4>)
constexpr size_t B4_STATE_BYTES = B4_P * B4_WORDS * 32 * sizeof(uint32_t);
constexpr unsigned int B4_RING_ENTRIES = 1u << 19;        // 48-byte entries; pushes keep B4_RING_MARGIN free
constexpr unsigned int B4_RING_MARGIN = 148u * 8u * 128u + 4096u;

__host__ __device__ constexpr size_t bulk4_smem(uint32_t blob_bytes) {
    return ((size_t)blob_bytes + 15u) / 16u * 16u + (B4_THREADS / 32) * (B4_ACT_BYTES + B4_STATE_BYTES);
}

// ---- ray state of slot p of this lane in the warp's state block: word w at [(p * B4_WORDS + w) * 32 + lane] -----------
__device__ __forceinline__ void b4_store(uint32_t* sst, int p, const RaySlot& s) {
    uint32_t* b = sst + p * B4_WORDS * 32;
    b[0 * 32] = (uint32_t)s.pix; b[1 * 32] = (uint32_t)s.steps; b[2 * 32] = (uint32_t)s.mode;
    b[3 * 32] = __float_as_uint(s.p.x); b[4 * 32] = __float_as_uint(s.p.y); b[5 * 32] = __float_as_uint(s.p.z);
    b[6 * 32] = __float_as_uint(s.d.x); b[7 * 32] = __float_as_uint(s.d.y); b[8 * 32] = __float_as_uint(s.d.z);
    b[9 * 32] = __float_as_uint(s.tfar); b[10 * 32] = __float_as_uint(s.s0); b[11 * 32] = __float_as_uint(s.s1);
    b[12 * 32] = __float_as_uint(s.s2); b[13 * 32] = __float_as_uint(s.rem);
}
__device__ __forceinline__ void b4_load(const uint32_t* sst, int p, RaySlot& s) {
    const uint32_t* b = sst + p * B4_WORDS * 32;
    s.pix = (int)b[0 * 32]; s.steps = (int)b[1 * 32]; s.mode = (int)b[2 * 32];
    s.p = f3{__uint_as_float(b[3 * 32]), __uint_as_float(b[4 * 32]), __uint_as_float(b[5 * 32])};
    s.d = f3{__uint_as_float(b[6 * 32]), __uint_as_float(b[7 * 32]), __uint_as_float(b[8 * 32])};
    s.tfar = __uint_as_float(b[9 * 32]); s.s0 = __uint_as_float(b[10 * 32]); s.s1 = __uint_as_float(b[11 * 32]);
    s.s2 = __uint_as_float(b[12 * 32]); s.rem = __uint_as_float(b[13 * 32]);
}

// ---- ring queue in L2: counters pc[1] = tail, pc[2] = head (absolute indices), entry = queue[abs & (cap - 1)],
// published by writing abs + 1 into its `ready` word (the host zeroes the ring before every launch) ----------------------
__device__ __forceinline__ void ring_push(RaySlot& s, bool want, const RenderParams& rp, LocalStats& st) {
    const unsigned int m = __ballot_sync(0xffffffffu, want);
    if (!m) return;
    const int lane = threadIdx.x & 31, leader = __ffs(m) - 1;
    unsigned int base = 0, ok = 0;
    if (lane == leader) {
        const unsigned int cnt = __popc(m);
        const unsigned int tail = ld_volatile_u32(rp.pixel_counter + 1), head = ld_volatile_u32(rp.pixel_counter + 2);
        if (tail - head + cnt + B4_RING_MARGIN <= rp.queue_capacity) {     // head only grows: the backlog is over-estimated
            base = atomicAdd(rp.pixel_counter + 1, cnt);
            ok = 1;
        }
    }
    ok = __shfl_sync(0xffffffffu, ok, leader);
    if (!ok) return;                                                       // ring full: the rays stay where they are
    base = __shfl_sync(0xffffffffu, base, leader);
    if (want) {
        const unsigned int abs = base + __popc(m & lanemask_lt());
        StragglerEntry* e = rp.queue + (abs & (rp.queue_capacity - 1u));
        float4* dst = reinterpret_cast<float4*>(e);
        dst[0] = make_float4(__int_as_float(s.pix), __int_as_float(s.steps), s.p.x, s.p.y);
        dst[1] = make_float4(s.p.z, s.d.x, s.d.y, s.d.z);
        reinterpret_cast<float2*>(dst + 2)[0] = make_float2(s.tfar, s.rem);
        __threadfence();
        asm volatile("st.volatile.global.u32 [%0], %1;" ::"l"(&e->ready), "r"(abs + 1u) : "memory");
        ++st.evicted;
        s.mode = -1;
    }
}
// lanes with want == true try to pop one entry each; returns true in the lanes that got one
__device__ __forceinline__ bool ring_pop(RaySlot& s, bool want, const RenderParams& rp) {
    const unsigned int m = __ballot_sync(0xffffffffu, want);
    if (!m) return false;
    const int lane = threadIdx.x & 31, leader = __ffs(m) - 1;
    unsigned int base = 0, got = 0;
    if (lane == leader) {
        const unsigned int cnt = __popc(m);
        unsigned int head = ld_volatile_u32(rp.pixel_counter + 2);
        for (;;) {
            const unsigned int tail = ld_volatile_u32(rp.pixel_counter + 1);
            if ((int)(tail - head) <= 0) { got = 0; break; }
            const unsigned int take = (tail - head) < cnt ? (tail - head) : cnt;
            const unsigned int old = atomicCAS(rp.pixel_counter + 2, head, head + take);
            if (old == head) { base = head; got = take; break; }
            head = old;
        }
    }
    base = __shfl_sync(0xffffffffu, base, leader);
    got = __shfl_sync(0xffffffffu, got, leader);
    const unsigned int rank = __popc(m & lanemask_lt());
    const bool mine = want && rank < got;
    if (mine) {
        const unsigned int abs = base + rank;
        const StragglerEntry* e = rp.queue + (abs & (rp.queue_capacity - 1u));
        unsigned int spins = 0;
        while (ld_volatile_u32(&e->ready) != abs + 1u) {
            __nanosleep(20);
            if (++spins > 100000000u) __trap();   // a reserved entry is published within microseconds
        }
        __threadfence();
        const float4* src = reinterpret_cast<const float4*>(e);
        const float4 e0 = __ldcg(src), e1 = __ldcg(src + 1);
        const float2 e2 = __ldcg(reinterpret_cast<const float2*>(src + 2));
        s.pix = __float_as_int(e0.x); s.steps = __float_as_int(e0.y);
        s.p = f3{e0.z, e0.w, e1.x}; s.d = f3{e1.y, e1.z, e1.w};
        s.tfar = e2.x; s.rem = e2.y; s.mode = 0;
    }
    return mine;
}

// One evaluation step of up to 16 rays held by lanes 0..15 of a consumer warp (the evaluator ladder of trace_coop.cuh).
__device__ __forceinline__ float b4_consumer_eval(const float* s_w, float* abuf, const RenderParams& rp, unsigned int act,
                                                  bool active, const f3& q, int lane) {
    const int na = __popc(act);
    const int rank = __popc(act & lanemask_lt());
    if (na > 8) {
        float4* xin = reinterpret_cast<float4*>(abuf + 16 * WARP_ACT_STRIDE);
        if (active) xin[rank] = make_float4(q.x, q.y, q.z, 0.0f);
        else if (lane < 16 && lane >= na) xin[lane] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        const float out = mlp32_eval_warp16(s_w, rp.n_sq, abuf, lane, rp.t4, rp.n_in == 4);
        return __shfl_sync(0xffffffffu, out, rank & 15);
    }
    int src[8];
    unsigned int mm = act;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        src[j] = mm ? (__ffs(mm) - 1) : 0;
        mm &= mm - 1u;
    }
    if (na > 4) return coop_eval<8>(s_w, rp.n_sq, abuf, q, src, rank, lane, rp.t4, rp.n_in == 4);
    if (na > 2) return coop_eval<4>(s_w, rp.n_sq, abuf, q, src, rank, lane, rp.t4, rp.n_in == 4);
    if (na == 2) return coop_eval<2>(s_w, rp.n_sq, abuf, q, src, rank, lane, rp.t4, rp.n_in == 4);
    return coop_eval<1>(s_w, rp.n_sq, abuf, q, src, rank, lane, rp.t4, rp.n_in == 4);
}

template <int KU, int PEEL = 0>
__global__ void __launch_bounds__(B4_THREADS, 1) trace_bulk4_kernel(const __grid_constant__ RenderParams rp) {
    extern __shared__ __align__(16) float s_w[];
    __shared__ uint64_t s_bar;
    stage_weights(s_w, rp.blob, rp.blob_bytes, &s_bar);

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned char* const warp_base = reinterpret_cast<unsigned char*>(s_w) + ((rp.blob_bytes + 15u) & ~15u) +
                                     (size_t)warp * (B4_ACT_BYTES + B4_STATE_BYTES);
    float4* const act4 = reinterpret_cast<float4*>(warp_base) + lane;
    float* const abuf = reinterpret_cast<float*>(warp_base);            // the consumer's activation rows live in the same block
    uint32_t* const sst = reinterpret_cast<uint32_t*>(warp_base + B4_ACT_BYTES) + lane;
    unsigned int* const pc = rp.pixel_counter;

    const int period = (int)rp.consumer_period;
    const unsigned int n_cons_cta = period > 0 ? gridDim.x / (unsigned int)period : 0u;
    const unsigned int total_producers = (gridDim.x - n_cons_cta) * (B4_THREADS / 32);
    const bool producer = !(period > 0 && (int)(blockIdx.x % (unsigned int)period) == period - 1);
    LocalStats st = {0, 0, 0, 0, 0, 0};

    if (producer) {
        // Per-slot work runs in ROLLED loops over the four slot rows (one copy of slot_start / slot_update in the instruction
        // stream: four inlined copies made the kernel 275 KB of SASS and the warps waited on instruction fetch); the query points
        // and the network outputs pass through the warp's activation block, which the evaluator only overwrites after reading x.
        float* const xq = reinterpret_cast<float*>(act4);      // xq[(j * 32) * 4 + p] = coordinate j of slot p (this lane's float4s)
#pragma unroll 1
        for (int p = 0; p < B4_P; ++p) {
            sst[(p * B4_WORDS + 2) * 32] = (uint32_t)-1;
            xq[0 * 128 + p] = 0.0f; xq[1 * 128 + p] = 0.0f; xq[2 * 128 + p] = 0.0f;
        }
        bool exhausted = false;     // warp-uniform
        const int refill_min = rp.b4_refill_min;
        for (;;) {
            // ---- refill: one atomic for all idle slots of the warp
            int n_busy = 0;
            for (;;) {
                int idle_mine = 0;
#pragma unroll
                for (int p = 0; p < B4_P; ++p) idle_mine += (int)sst[(p * B4_WORDS + 2) * 32] < 0 ? 1 : 0;
                int incl = idle_mine;          // inclusive prefix sum over the lanes
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) {
                    const int v = __shfl_up_sync(0xffffffffu, incl, d);
                    if (lane >= d) incl += v;
                }
                const int total = __shfl_sync(0xffffffffu, incl, 31);
                n_busy = 32 * B4_P - total;
                if (exhausted || total == 0 || (total < refill_min && n_busy > 0)) break;
                unsigned int base = 0;
                if (lane == 31) base = atomicAdd(pc, (unsigned int)total);
                base = __shfl_sync(0xffffffffu, base, 31);
                unsigned int idx = base + (unsigned int)(incl - idle_mine);
#pragma unroll 1
                for (int p = 0; p < B4_P; ++p) {
                    int flush_g = -1;
                    if ((int)sst[(p * B4_WORDS + 2) * 32] < 0) {
                        if (idx < rp.total_idx) {
                            RaySlot s;
                            slot_start(s, idx, rp, st, flush_g);
                            if (s.mode >= 0) {
                                b4_store(sst, p, s);
                                xq[0 * 128 + p] = s.p.x; xq[1 * 128 + p] = s.p.y; xq[2 * 128 + p] = s.p.z;
                            }
                        }
                        ++idx;
                    }
                    flush_completed(rp, flush_g, lane);
                }
                if (base + (unsigned int)total >= rp.total_idx) exhausted = true;
            }
            if (n_busy == 0) break;
            // ---- one network evaluation for the four slots of every lane (their query points are in place: written when a ray
            // started and by the previous update; zeros for idle slots)
            float x[B4_P][3], y[B4_P];
            {
                const float4 x0 = act4[0], x1 = act4[32], x2 = act4[64];
                x[0][0] = x0.x; x[1][0] = x0.y; x[2][0] = x0.z; x[3][0] = x0.w;
                x[0][1] = x1.x; x[1][1] = x1.y; x[2][1] = x1.z; x[3][1] = x1.w;
                x[0][2] = x2.x; x[1][2] = x2.y; x[2][2] = x2.z; x[3][2] = x2.w;
            }
            mlp32_eval_thread_s<B4_P, KU, PEEL>(s_w, rp.n_sq, act4, 32, x, y, rp.t4, rp.n_in == 4);
            act4[0] = make_float4(y[0], y[1], y[2], y[3]);
            // ---- ray updates and the next query points; long rays, and the remnants of a warp that has nothing to refill from,
            // go to the consumers
            const bool drain = exhausted && n_busy <= rp.hybrid;
#pragma unroll 1
            for (int p = 0; p < B4_P; ++p) {
                int flush_g = -1;
                RaySlot s;
                s.mode = (int)sst[(p * B4_WORDS + 2) * 32];
                const bool was_busy = s.mode >= 0;
                bool go = false;
                if (was_busy) {
                    b4_load(sst, p, s);
                    const f3 q = slot_query_point(s);
                    slot_update(s, xq[p], q, rp, st, flush_g);
                    go = s.mode == 0 && (drain || s.steps >= rp.evict_steps_hi);
                }
                flush_completed(rp, flush_g, lane);
                ring_push(s, go, rp, st);
                f3 qn = f3{0.0f, 0.0f, 0.0f};
                if (was_busy) {
                    if (s.mode >= 0) {
                        b4_store(sst, p, s);
                        qn = slot_query_point(s);
                    } else {
                        sst[(p * B4_WORDS + 2) * 32] = (uint32_t)-1;
                    }
                }
                xq[0 * 128 + p] = qn.x; xq[1 * 128 + p] = qn.y; xq[2 * 128 + p] = qn.z;
            }
        }
        if (lane == 0) { __threadfence(); atomicAdd(pc + 3, 1u); }
    }

    // ---- consumer: 16 ray slots in lanes 0..15, fed from the ring
    RaySlot slot;
    slot.mode = -1;
    int poll = 0;
    unsigned int idle = 0;
    bool bulk_over = false;
    for (;;) {
        unsigned int act = __ballot_sync(0xffffffffu, lane < 16 && slot.mode >= 0);
        if (!act || --poll <= 0) {
            poll = 8;
            ring_pop(slot, lane < 16 && slot.mode < 0, rp);
            act = __ballot_sync(0xffffffffu, lane < 16 && slot.mode >= 0);
        }
        if (!act) {
            unsigned int done = 0, over = 0;
            if (lane == 0) {
                const unsigned int fin = ld_volatile_u32(pc + 3);
                const unsigned int tail = ld_volatile_u32(pc + 1);
                const unsigned int head = ld_volatile_u32(pc + 2);
                over = fin >= total_producers ? 1u : 0u;
                done = (over && (int)(tail - head) <= 0) ? 1u : 0u;
            }
            bulk_over = __shfl_sync(0xffffffffu, over, 0) != 0u;
            if (__shfl_sync(0xffffffffu, done, 0)) break;
            __nanosleep(bulk_over ? 100 : 500);
            if (++idle > 50000000u) __trap();     // never hang the device
            poll = 0;
            continue;
        }
        idle = 0;
        if (__popc(act) == 1) {
            const int owner = __ffs(act) - 1;
            if (__shfl_sync(0xffffffffu, slot.mode, owner) == 0) {
                const int chunk = 64 - (__shfl_sync(0xffffffffu, slot.steps, owner) & 63);
                lone_ray_march(slot, owner, s_w, rp.n_sq, abuf, rp, st, lane, chunk);
                poll = 0;
                continue;
            }
        }
        const bool active = lane < 16 && slot.mode >= 0;
        const f3 q = active ? slot_query_point(slot) : f3{0.0f, 0.0f, 0.0f};
        const float mine = b4_consumer_eval(s_w, abuf, rp, act, active, q, lane);
        int flush_g = -1;
        if (active) slot_update(slot, mine, q, rp, st, flush_g);
        flush_completed(rp, flush_g, lane);
    }
    flush_stats(st, rp.stats);
}

}  // namespace nsdf
