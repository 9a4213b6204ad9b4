/* =====================================================================================
 * nsdf.h -- C ABI of the B200-native neural-SDF hot path (libnsdf.so).
 *
 * This is the drop-in boundary: plain pointers and sizes, opaque handles, int status
 * (0 = success, otherwise a cudaError_t value or one of the NSDF_E* codes below).
 * Nothing here depends on torch, on C++ types, or on the CPU oracle.  Every entry
 * point names the interface of the reference (daviesthomas/cudaNeuralRender, paths
 * relative to its root) that it replaces.  INTEGRATION.md shows the reference-side
 * binding (the patch a maintainer would apply to src/layers/denseLayer.cu,
 * src/neuralNetwork.cpp and src/volumeRender_kernel.cu).
 *
 * Conventions
 *   - "d_" pointers are device pointers on the CURRENT device, "h_" pointers are host.
 *   - `stream` is a cudaStream_t passed as void* (NULL = default stream).  Calls that take a
 *     stream are asynchronous with respect to the host unless stated otherwise.
 *   - A model handle belongs to the device that was current when it was created.
 *   - ONE render at a time per model handle: the tracer's queues, counters and staging buffers are scratch of the handle
 *     (the reference keeps file-scope globals, src/volumeRender_kernel.cu:578-585).  Use one handle per stream / thread;
 *     nsdf_multi_* holds one per device.  The tracer is launched cooperatively (cudaLaunchCooperativeKernel, one CTA per SM):
 *     its warps wait on one another, so the launch either gets the whole grid resident or fails with a CUDA error -- it
 *     never spins on CTAs that cannot run.
 * ===================================================================================== */
#ifndef NSDF_H_
#define NSDF_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NSDF_OK 0
#define NSDF_EINVAL 10001       /* bad argument                                           */
#define NSDF_EUNSUPPORTED 10002 /* model shape not handled by the requested kernel        */
#define NSDF_EIO 10003          /* file could not be read / parsed                        */
#define NSDF_ENOMEM 10004
#define NSDF_ETOOLARGE 10005   /* size beyond the fixed maximum of an allocation-free entry point */

/* Activation ids: the reference's `enum Activation { ReLU, Tanh }` (src/layers/denseLayer.hh:6-9).
 * "Tanh" selects the LINEAR epilogue in the reference (src/layers/denseLayer.cu:93-113,150-166);
 * the caller applies tanh itself (src/simpleInfer.cpp:109, src/volumeRender_kernel.cu:229). */
#define NSDF_ACT_RELU 0
#define NSDF_ACT_TANH_LINEAR 1

/* Arithmetic rungs of the fused MLP kernel (DESIGN.md "Precision ladder"). */
#define NSDF_PREC_FP32 0       /* fp32 FFMA, ascending-k single accumulator: bit-identical to the reference order */
#define NSDF_PREC_TC_FP16 1    /* tcgen05 kind::f16, fp16 operands, fp32 accumulate (bound 4e-3 abs)              */
#define NSDF_PREC_TC_BF16 2    /* tcgen05 kind::f16, bf16 operands, fp32 accumulate (bound 5e-2 abs)              */
#define NSDF_PREC_TC_FP16X3 3  /* tcgen05, split-fp16 (hi+lo) operands, 3 MMAs per layer (bound 1e-4 abs)        */
#define NSDF_PREC_TC_FP16_ACC16 5 /* tcgen05, fp16 operands AND fp16 accumulators in layers 0..n-1 (last layer fp32): halves
                                    the TMEM read-back that bounds the other rungs (bound 2e-2 abs)              */
#define NSDF_PREC_FP32_WARP 4  /* fp32, warp-cooperative evaluator (lanes = neurons); same bits as NSDF_PREC_FP32 */

/* Scene functions: sceneSDF, src/volumeRender_kernel.cu:217-230. */
#define NSDF_SCENE_TANH 0       /* tanh(nSDF)  (:229, the shipped renders)                 */
#define NSDF_SCENE_MANYSPHERE 1 /* manySphere(p, nSDF, true)  (:221, what HEAD evaluates)  */
#define NSDF_SCENE_PROGRAM 2    /* nsdf_render_opts.program: the analytic-SDF library of :67-196 composed at run time */

/* ---- scene programs: the reference's analytic SDF primitives and operators (src/volumeRender_kernel.cu:67-196), which it
 * composes by editing sceneSDF (:217-230) and recompiling, as data.  A program is a chain over one accumulator s:
 *     for each op:  d = primitive(p, nSDF);  s = combine(s, d)        (modifiers act on s alone)
 * and the scene's distance is the final s.  Every scene the reference's sceneSDF holds (live or commented out) is such a
 * chain.  Arithmetic: fp32 as written in the reference, its double-promoting literals kept, tanh / saturate / sin in the
 * deterministic forms of DESIGN.md section 4; device and oracle agree bit for bit.                                          */
typedef struct nsdf_scene_op {
    int prim;    /* NSDF_PRIM_* or NSDF_MOD_*                                   */
    int combine; /* NSDF_COMB_* (ignored for modifiers)                         */
    float a[8];  /* parameters, see below                                       */
} nsdf_scene_op;
#define NSDF_SCENE_MAX_OPS 16
/* primitives: d = ...                                                                                          */
#define NSDF_PRIM_NSDF 0     /* the network output, raw                                                          */
#define NSDF_PRIM_TANH 1     /* tanh(nSDF) (:229)                                                                */
#define NSDF_PRIM_SPHERE 2   /* sdfSphere(p - a[0..2], a[3])                                       (:67-71)      */
#define NSDF_PRIM_BOX 3      /* sdfBox(p - a[0..2], b = a[3..5], r = a[6])                          (:81-89)      */
#define NSDF_PRIM_PLANE 4    /* sdfPlane: p.y - a[0]  (the reference hard-codes 0.5)                (:91-94)      */
#define NSDF_PRIM_CYLINDER 5 /* sdfCylinder(p - a[0..2], c = a[3..5]): length(p.xy - (c.x, c.z)) - c.y (:96-101) */
/* modifiers: s = ...                                                                                           */
#define NSDF_MOD_DISPLACE 16 /* sdfOpDisplace: s + sin(a[0] x) sin(a[0] y) sin(a[0] z) * a[1]  (5, 0.05 in :103-110) */
#define NSDF_MOD_ROUND 17    /* sdfOpRound: s - a[0]                                                (:112-115)    */
#define NSDF_MOD_ONION 18    /* sdfOpOnion: |s| - a[0]                                              (:117-120)    */
#define NSDF_MOD_MANY_SPHERE 19        /* manySphere(p, s, a[0] != 0)   (:176-196, animated by the frame number)  */
#define NSDF_MOD_MANY_CYLINDER_CUT 20  /* manyCylinderCut(p, s)         (:156-174)                                */
/* combiners: s = ...                                                                                           */
#define NSDF_COMB_SET 0              /* d                                                                        */
#define NSDF_COMB_UNION 1            /* min(s, d)                                                   (:128-131)    */
#define NSDF_COMB_INTERSECT 2        /* max(s, d)                                                   (:123-126)    */
#define NSDF_COMB_SUBTRACT 3         /* max(s, -d): d carved out of s                               (:133-136)    */
#define NSDF_COMB_SUBTRACT_FROM 4    /* max(d, -s): s carved out of d                                            */
#define NSDF_COMB_SMOOTH_UNION 5     /* sdfOpSmoothUnion(s, d, k = a[7])                            (:144-149)    */
#define NSDF_COMB_SMOOTH_SUBTRACT 6  /* sdfOpSmoothSubtraction(s, d, k = a[7])                      (:138-142)    */
#define NSDF_COMB_SMOOTH_SUBTRACT_FROM 7 /* sdfOpSmoothSubtraction(d, s, k = a[7])  (the commented scene of :225-227) */

/* Arithmetic of the three operations the reference's device code takes from the SFU / libdevice: normalize's rsqrtf
 * (src/common/helper_math.h:1309-1313), tanh (src/volumeRender_kernel.cu:229) and __saturatef (:140,146,268-271).
 *   NSDF_MATH_REFERENCE      the same instructions the reference compiles to (rsqrt.approx.f32, libdevice tanhf, .sat): on the
 *                            GPU family the reference binary runs on, frames are the reference's frames.  The drop-in default.
 *   NSDF_MATH_DETERMINISTIC  IEEE-only forms (1/sqrt correctly rounded, a polynomial tanh, compare/select) that a CPU reproduces
 *                            bit for bit: the mode the oracle tests run in.  Frames differ from the reference's in the last bits
 *                            of ray directions and normals (0 hit-mask mismatches, 97.7-99.98 % of hit pixels within 1/255).   */
#define NSDF_MATH_REFERENCE 0
#define NSDF_MATH_DETERMINISTIC 1

/* c_coloringType, src/volumeRender_kernel.cu:33 */
#define NSDF_COLOR_FACING 0
#define NSDF_COLOR_MATCAP 1

typedef struct nsdf_model nsdf_model;

/* ---- library --------------------------------------------------------------------------- */
const char* nsdf_version(void);
/* Human-readable text for a status returned by this library. */
const char* nsdf_error_string(int status);
/* Number of SMs of the current device (0 if there is no usable device). */
int nsdf_device_sm_count(void);

/* ---- device buffers: what Matrix / Image (src/neuralUtils/matrix.cu:12-72, image.cu:12-35,112-140)
 * do with cudaMalloc / cudaMemcpy / cudaFree, so that the C++ host layer needs no CUDA toolchain.
 * Synchronous; pointers are plain device pointers on the current device. */
int nsdf_dev_malloc(void** d_ptr_out, size_t bytes);
int nsdf_dev_free(void* d_ptr);
int nsdf_dev_memset(void* d_ptr, int value, size_t bytes);
int nsdf_memcpy_h2d(void* d_dst, const void* h_src, size_t bytes);
int nsdf_memcpy_d2h(void* h_dst, const void* d_src, size_t bytes);
int nsdf_device_synchronize(void);
int nsdf_set_device(int ordinal);
int nsdf_device_count(void);

/* ---- model: replaces NeuralNetwork::load / addLayer and the DenseLayer constructor ------
 * (src/neuralNetwork.cpp:50-52,85-151; src/layers/denseLayer.cu:180-227).
 * W[l] is HOST memory, [out][in] row-major, exactly DenseLayer::W after initializeWeights
 * (W[y*in + x] = weights[x][y], denseLayer.cu:220); b[l] is host, [out].  Weights are copied to
 * the device and re-packed for the fused kernels; the caller keeps ownership of its buffers. */
int nsdf_model_create(const float* const* W, const float* const* b, const int* in_dim, const int* out_dim,
                      const int* activation, int n_layers, nsdf_model** out);
/* Same, from one packed host array: per layer W[out][in] then b[out]; dims has n_layers+1 entries.
 * Activations follow src/neuralNetwork.cpp:136-139 (last layer "Tanh", the others ReLU). */
int nsdf_model_create_packed(const float* params, const int* dims, int n_layers, nsdf_model** out);
/* Keras `save_weights` .h5 file, walked like NeuralNetwork::load (name-sorted groups,
 * /L/L/{bias:0,kernel:0}; src/neuralNetwork.cpp:85-151) by the library's own HDF5 reader. */
int nsdf_model_load_h5(const char* path, nsdf_model** out);
/* Host-only part of the same walk (no device needed): returns the layer count, dims (n_layers+1
 * ints, capacity dims_capacity) and the packed parameters in the layout of nsdf_model_create_packed.
 * Any output pointer may be NULL; n_params_out always receives the float count. */
int nsdf_h5_read_keras(const char* path, int* n_layers_out, int* dims_out, int dims_capacity, float* params_out,
                       size_t params_capacity, size_t* n_params_out);
/* Writes a Keras `save_weights`-style .h5 (the subset the shipped files use: superblock v0, symbol-table groups
 * /L/L/{bias:0,kernel:0}, contiguous float32, root attribute `layer_names`) from packed parameters in the layout of
 * nsdf_model_create_packed; groups are named dense, dense_1, ...  The reader orders layers by `layer_names` when it names
 * exactly the groups present, so stacks of more than ten layers load in network order (the reference's name sort,
 * src/neuralNetwork.cpp:88-91, puts dense_10 before dense_2).  No device needed.  SURVEY 8 f1. */
int nsdf_h5_write_keras(const char* path, const float* params, const int* dims, int n_layers);
/* Same with a storage layout for the datasets.  The reader (nsdf_model_load_h5, nsdf_h5_read_keras, the HighFive shim) takes
 * whatever libhdf5 would hand the reference for float32/float64 data: contiguous, compact, and chunked datasets with the
 * shuffle, deflate (gzip) and fletcher32 filters (what h5py writes with compression="gzip"). */
#define NSDF_H5_CONTIGUOUS 0
#define NSDF_H5_CHUNKED 1
#define NSDF_H5_CHUNKED_DEFLATE 2 /* chunked + shuffle + deflate filter pipeline; the streams are stored (uncompressed) blocks */
int nsdf_h5_write_keras_ex(const char* path, const float* params, const int* dims, int n_layers, int storage);
void nsdf_model_destroy(nsdf_model* m);
int nsdf_model_num_layers(const nsdf_model* m);
/* dims_out receives n_layers+1 ints. */
int nsdf_model_dims(const nsdf_model* m, int* dims_out);
/* 1 if the fused kernels apply (3|4 -> 32 x k -> 1, ReLU hidden, linear last), else 0.  4-input nets are the reference's
 * --animation mode (x, y, z, t; src/volumeRender_kernel.cu:524-545): nsdf_render feeds t = (float)cam->frame to every query
 * (fp32 schedules only; NSDF_RENDER_TC_FARFIELD returns NSDF_EUNSUPPORTED); nsdf_mlp_forward takes B x 4 inputs through the
 * layer-by-layer kernels. */
int nsdf_model_is_fused(const nsdf_model* m);
/* Copies the packed host parameters (layout of nsdf_model_create_packed) into params_out
 * (capacity in floats); returns the number of floats through n_out. */
int nsdf_model_params(const nsdf_model* m, float* params_out, size_t capacity, size_t* n_out);

/* ---- MLP: replaces NeuralNetwork::forward (src/neuralNetwork.cpp:54-63) -------------------
 * d_x: n points, element [b*in0 + c] -- the layout of Matrix(Shape(3, B)) (src/simpleInfer.cpp:116);
 * d_y: n values, pre-tanh, the layout of the returned Matrix(1, B).  Any n >= 0 (the reference
 * fails when n is a multiple of 65536). */
int nsdf_mlp_forward(const nsdf_model* m, const float* d_x, float* d_y, int64_t n, int precision, void* stream);
/* Host-buffer variant (what a plugin caller with host Matrix data uses): H2D, kernel, D2H,
 * synchronous.  h_x / h_y may be pageable or pinned. */
int nsdf_mlp_forward_host(const nsdf_model* m, const float* h_x, float* h_y, int64_t n, int precision);
/* Dense lattice sweep (BASELINE config 3): evaluates the res^3 lattice p = lo + (hi-lo)*(i+0.5)/res,
 * x fastest, for linear indices [first, first+count); points are generated in-kernel. */
int nsdf_mlp_forward_lattice(const nsdf_model* m, float* d_y, int res, float lo, float hi, int64_t first,
                             int64_t count, int precision, void* stream);

/* ---- per-layer operators: replace the free functions of src/layers/denseLayer.cu ----------
 * denseLayerForward (:86-124) and batchedDenseLayerForward (:126-176): Z[:,j] = act(W A[:,j] + b)
 * for j < num_batches; W is [M][K] row-major, A has column stride K, Z column stride M
 * (N must be 1, as in every call the reference makes).  All device pointers, caller-owned. */
int nsdf_dense_layer_forward(const float* d_W, const float* d_A, float* d_Z, const float* d_b, int M, int N, int K,
                             int64_t num_batches, int activation, void* stream);

/* ---- renderer: replaces render_kernel / copyViewMatrices / copyStaticSettings ---------------
 * (src/volumeRender_kernel.cu:608-706) and the kernels it drives (initMarcher :293,
 * createBatch :504, singleMarch :416, thrust scan :553). */
typedef struct nsdf_camera {
    float inv_view[12]; /* c_invViewMatrix: rows 0..2 of modelView (src/main.cpp:218-221)            */
    float normal[16];   /* c_normalMatrix: inverse(modelView), row-major 4x4 (src/main.cpp:214-217) */
    int frame;          /* c_frameNumber                                                           */
} nsdf_camera;

/* Knobs of the persistent tracer (csrc/trace_coop.cuh); nsdf_tuning_default fills the measured best for 1080p..4K frames on a
 * B200.  The product library reads no environment variable: this struct is the only way in.  The serial tail of a frame is a
 * handful of rays that need thousands of DEPENDENT network evaluations: "express" SMs take no pixels and hold
 * one such ray per lane (warp), stepping it at ~1.3 us instead of the ~14 us a ray pays inside a full 16-ray warp.            */
typedef struct nsdf_tuning {
    int express_period;      /* every express_period-th SM is an express SM; 0 = none; -1 = automatic: 8, 24 for calls of >= 3 M
                                pixels, none for calls of >= 12 views (pure throughput: their tails overlap)                    */
    int express_steps;       /* a ray that has marched this many steps moves to a free express lane (160)                      */
    int express_warps;       /* lanes per express SM (4: one per scheduler steps fastest)                                       */
    int express_old_steps;   /* rays older than this (192) ask for a lane whenever steps & express_old_mask == 0 ...            */
    int express_old_mask;    /* ... (1: every 2nd step) instead of every 16th, so that a freed lane goes to the older ray      */
    int express_from_bulk;   /* 1: warps that still take pixels hand rays over too                                             */
    int express_rays;        /* rays a lane marches together (1; more measured slower for the longest chains)                  */
    int express_solo_steps;  /* with express_rays > 1: a lane holding a ray this old takes no company (768)                    */
    int tile_priority_steps; /* probe every 8th pixel for this many steps and trace jobs with grazing rays first (0 = off;
                                -1 = automatic: 32, off for calls of >= 12 views)                                                 */
    int tile_priority_min_jobs; /* ... for calls of at least this many (view, tile) jobs (128)                                 */
    int tail_lanes;          /* once the pixels are gone, warps 0..tail_lanes-1 of every SM are lanes too (0 = off)             */
    int tail_push_max;       /* ... fed by warps that are down to this many marching rays (2)                                   */
    int bulk4;               /* calls of several views on the four-rays-per-lane throughput tracer (csrc/trace_bulk4.cuh):
                                1 = on, 0 = off, -1 = automatic (calls of >= 4 views and >= 12 M pixels)                                            */
    int bulk4_evict_steps;   /* ... a ray still marching after this many steps moves to the consumer warps (0 = never: measured
                                best; consumers then only finish what thinned-out warps hand over at the end of the call)       */
    int bulk4_period;        /* ... every bulk4_period-th SM consumes from the start (0 = none)                                 */
    int bulk4_drain;         /* ... a warp that is out of pixels hands its rays over once this few are left of its 128 (16)     */
    int bulk4_refill_min;    /* ... a warp asks for new pixels once this many of its 128 slots are idle (16)                    */
    int bulk4_unroll;        /* ... k-unroll of the evaluator's inner loop: 1, 2 or 4 (2: the loop body must stay in the
                                instruction cache next to the ray code)                                                          */
    int max_ctas;            /* the persistent tracers launch at most this many CTAs (= SMs); 0 = all.  Several calls with a share of
                                the SMs each, on different streams and model handles, run side by side: their tails overlap            */
} nsdf_tuning;
void nsdf_tuning_default(nsdf_tuning* t);
/* 1 if this library was built with -DNSDF_DEV_SCHEDULES (libnsdf_dev.so: superseded A/B schedules, mixed-precision far field,
 * knobs from the environment); the product build returns 0 and answers NSDF_EUNSUPPORTED to their flags. */
int nsdf_build_has_dev_schedules(void);

typedef struct nsdf_render_opts {
    int color_type;   /* NSDF_COLOR_*                                                      */
    int scene;        /* NSDF_SCENE_*                                                      */
    int max_steps;    /* MAX_STEPS of the reference = 6000 (volumeRender_kernel.cu:61)     */
    int tile_rank;    /* this call renders the 64x64 tiles t with t % tile_world == tile_rank */
    int tile_world;   /* 1 = whole frame                                                   */
    int evict_steps;  /* rays still marching after this many steps move to the latency path; 0 = library default */
    int flags;        /* NSDF_RENDER_* bits                                                */
    const nsdf_scene_op* program; /* HOST array, read during the call; used when scene == NSDF_SCENE_PROGRAM */
    int n_ops;        /* 1..NSDF_SCENE_MAX_OPS                                             */
    int math;         /* NSDF_MATH_*                                                       */
    const nsdf_tuning* tuning; /* NULL = nsdf_tuning_default                                       */
} nsdf_render_opts;

#define NSDF_RENDER_SIMPLE 1u      /* one-thread-per-pixel kernel without refill (debug / A-B testing) */
#define NSDF_RENDER_NO_STRAGGLER 2u /* ray-per-lane kernel with refill, nothing handed to cooperative warps (A/B) */
#define NSDF_RENDER_PHASE_THROUGHPUT 4u /* launch only the throughput kernel (stragglers stay queued)       */
#define NSDF_RENDER_NO_QUIET_SMS 64u    /* default kernel without the quiet-SM queue for long rays (A/B)            */
#define NSDF_RENDER_QUEUE 32u           /* ray-per-lane producers + cooperative consumers through a queue, one launch (A/B) */
#define NSDF_RENDER_TWO_KERNELS 16u     /* throughput kernel then latency kernel instead of the fused launch (A/B) */
#define NSDF_RENDER_TWO_PASS 256u       /* ray-per-lane throughput kernel hands rays over after evict_steps (default 128) march steps,
                                           then the cooperative tracer refills its ray slots from that queue.  Bit-identical to
                                           the single launch (4K: 64.0 -> 55.4 ms; superseded as the large-call default by NSDF_RENDER_HYBRID)                                        */
#define NSDF_RENDER_HYBRID 512u         /* ONE launch: producing warps march one ray per lane (a third of the shared-memory traffic
                                           per query) and hand rays that reach evict_steps (default 192) march steps to cooperative
                                           consumer warps on every 32nd SM; out of pixels they become consumers themselves.
                                           Bit-identical; chosen automatically for calls covering >= 3 M pixels (4K: 54.9 -> 49.4 ms
                                           against NSDF_RENDER_TWO_PASS, the previous choice there)                                */
#define NSDF_RENDER_TC_FARFIELD 128u    /* mixed precision: march on the tensor cores while t >= threshold (fp16x3 rung,
                                           |error| <= 1e-4), hand every ray that gets close to the surface to the fp32 path
                                           (accept test, normals, shading unchanged).  Not bit-identical to the default. */
#define NSDF_RENDER_STAGED_STORES 1024u /* finished pixels go to a staging tile and every completed 64x64 tile is copied to d_out as 16-byte
                                           row segments by the warp that completes it; always on for tile-partitioned calls
                                           (tile_world > 1, where d_out is usually a peer GPU's memory), optional otherwise          */
#define NSDF_RENDER_PHASE_LATENCY 8u    /* launch only the latency kernel on the queue a PHASE_THROUGHPUT call
                                           of the same model handle left behind (per-kernel timing)            */

typedef struct nsdf_render_stats {
    unsigned long long queries;    /* MLP evaluations = sum over rays (march steps + 4 per shaded hit) */
    unsigned long long hit_pixels; /* shaded pixels                                                     */
    unsigned long long rays;       /* rays that entered the bounding sphere                             */
    unsigned long long evicted;    /* rays handed to the latency path                                   */
    unsigned long long unfinished; /* rays stopped by max_steps                                         */
    unsigned long long pixels;     /* pixels this call covered                                          */
} nsdf_render_stats;

/* Fills opts with what the reference's HEAD build does: sceneSDF = manySphere(p, nSDF, true) (src/volumeRender_kernel.cu:221),
 * MAX_STEPS 6000, facing-ratio colouring until copyStaticSettings says otherwise (src/main.cpp:655-662), reference math,
 * whole frame. */
void nsdf_render_opts_default(nsdf_render_opts* opts);
/* sceneSDF for n points on the device: d_out[i] = program(d_xyz[3i..3i+2], d_nsdf[i]) with c_frameNumber = frame
 * (what singleMarch :459 and surfaceNormal :361-377 evaluate per query).  program is a HOST array. */
int nsdf_scene_eval(const nsdf_scene_op* program, int n_ops, const float* d_xyz, const float* d_nsdf, int frame, int math,
                    float* d_out, int64_t n, void* stream);

/* updateViewMatrices (src/main.cpp:207-222): modelView from viewRotation (degrees) and
 * viewTranslation, without Eigen.  Host-side helper. */
void nsdf_camera_from_view(float rx_deg, float ry_deg, float tx, float ty, float tz, int frame, nsdf_camera* cam);

/* One frame into d_out (W*H packed a<<24|b<<16|g<<8|r, volumeRender_kernel.cu:273).  Every pixel
 * of the selected tiles is written (0 where the reference leaves the caller's memset).
 * d_matcap: mw*mh texels packed like Image::loadPNG (src/neuralUtils/image.cu:58); may be NULL
 * when color_type == NSDF_COLOR_FACING.  d_stats (device, may be NULL) is ACCUMULATED into.
 * d_out may be a peer-mapped pointer (tile partitioning across GPUs). */
int nsdf_render(const nsdf_model* m, const nsdf_camera* cam, const uint32_t* d_matcap, int mw, int mh,
                const nsdf_render_opts* opts, uint32_t* d_out, int W, int H, nsdf_render_stats* d_stats,
                void* stream);
/* n_views frames in ONE persistent launch (the multi-view loop of src/main.cpp:470-478, `--spin`): view v is rendered with
 * cams[v] (HOST array) into d_out + v * W * H.  The (view, tile) jobs of all views share the launch, so the serial tail of one
 * view (a handful of rays with thousands of dependent steps) hides behind the bulk of the others instead of idling the GPU
 * once per frame.  1 <= n_views <= 128, W * H <= 2^24 for n_views > 1; frames are bit-identical to n_views calls of
 * nsdf_render.  Only the default schedule (opts->flags without the A/B bits). */
int nsdf_render_views(const nsdf_model* m, const nsdf_camera* cams, int n_views, const uint32_t* d_matcap, int mw, int mh,
                      const nsdf_render_opts* opts, uint32_t* d_out, int W, int H, nsdf_render_stats* d_stats, void* stream);
/* nsdf_render for callers that cannot allocate device memory.  NO COUNTERPART IN THE REFERENCE: added
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  Same arguments and
 * the same frame, byte for byte, as nsdf_render; the call performs no cudaMalloc / cudaFree and calls no allocating library
 * routine: the tracer's scratch (job counters, tile-priority scores and order, 16.6 KB) is a file-scope device array that
 * every GPU which loads the library holds until exit.  Limits: the frame has at most NSDF_NOALLOC_MAX_TILES 64x64 tiles
 * (3840x2160 fits; beyond that NSDF_ETOOLARGE); (3|4) -> 32 x k -> 1 models, whole frames (tile_world 1), opts->flags 0; no
 * express lanes, so a single frame's serial tail is not shortened (the bytes do not depend on that).  Every call on a device
 * shares the array: it is cleared on `stream` at the start of each call, so calls on one stream may follow each other
 * freely, but two calls that OVERLAP on the GPU (different streams, or a different process using the same device through
 * this library instance) race.  The model handle m, d_matcap, d_out and d_stats are the caller's, as for nsdf_render. */
#define NSDF_NOALLOC_MAX_TILES 2048
int nsdf_render_noalloc(const nsdf_model* m, const nsdf_camera* cam, const uint32_t* d_matcap, int mw, int mh,
                        const nsdf_render_opts* opts, uint32_t* d_out, int W, int H, nsdf_render_stats* d_stats,
                        void* stream);
/* Host-buffer variant: uploads the matcap, renders, copies the frame back; synchronous.
 * h_stats may be NULL. */
int nsdf_render_host(const nsdf_model* m, const nsdf_camera* cam, const uint32_t* h_matcap, int mw, int mh,
                     const nsdf_render_opts* opts, uint32_t* h_out, int W, int H, nsdf_render_stats* h_stats);
/* Host-buffer variant of nsdf_render_views: h_out receives n_views * W * H texels. */
int nsdf_render_views_host(const nsdf_model* m, const nsdf_camera* cams, int n_views, const uint32_t* h_matcap, int mw, int mh,
                           const nsdf_render_opts* opts, uint32_t* h_out, int W, int H, nsdf_render_stats* h_stats);

/* ---- multi-GPU frame gather over NVLink (one process per GPU) --------------------------------
 * The presenting rank exports its frame buffer; the other ranks open it and pass the mapped
 * pointer as d_out of nsdf_render, so finished tiles are stored straight into the root's HBM. */
#define NSDF_IPC_HANDLE_BYTES 64
int nsdf_ipc_export(void* d_ptr, unsigned char handle_out[NSDF_IPC_HANDLE_BYTES]);
int nsdf_ipc_open(const unsigned char handle[NSDF_IPC_HANDLE_BYTES], void** d_ptr_out);
int nsdf_ipc_close(void* d_ptr);

/* ---- one process, several GPUs ------------------------------------------------------------------------------
 * The reference drives one device from one host thread.  nsdf_multi clones a network onto the listed devices
 * (devices[0] presents: d_out of nsdf_multi_render lives there; every other device needs peer access to it, else
 * NSDF_EUNSUPPORTED).
 *   n_views == 1 : the frame is partitioned by 64x64 tiles, tile t -> device t % n; each device's tracer stores its finished
 *                  tiles straight into d_out over NVLink (256-byte row segments from inside the kernel).
 *   n_views  > 1 : view v -> device v % n (the `--spin` loop of src/main.cpp:470-478 spread over the devices); each device
 *                  renders its views in one persistent launch and copies them to d_out + v * W * H on the presenting device.
 * h_matcap is a HOST array (uploaded to every device); opts->tile_rank / tile_world must be 0 / 1.  Synchronous.  Frames are
 * bit-identical to nsdf_render on one device.  h_stats (may be NULL) receives the totals over all devices. */
typedef struct nsdf_multi nsdf_multi;
int nsdf_multi_create(const nsdf_model* m, const int* devices, int n_devices, nsdf_multi** out);
void nsdf_multi_destroy(nsdf_multi* mm);
int nsdf_multi_device_count(const nsdf_multi* mm);
int nsdf_multi_render(nsdf_multi* mm, const nsdf_camera* cams, int n_views, const uint32_t* h_matcap, int mw, int mh,
                      const nsdf_render_opts* opts, uint32_t* d_out, int W, int H, nsdf_render_stats* h_stats);

/* ---- image I/O used by the drop-in executables (src/neuralUtils/image.cu:36-110,
 * src/common/helper_image.h:252-329) ------------------------------------------------------------ */
/* Decodes an 8-bit RGB/RGBA/grey non-interlaced PNG into packed texels (a<<24|b<<16|g<<8|r).
 * Call with texels == NULL to query w/h. */
int nsdf_png_load(const char* path, uint32_t* texels, size_t capacity, int* w, int* h);
/* Image::savePNG: flip != 0 reverses the texel stream (the reference's default 180-degree turn). */
int nsdf_png_save(const char* path, const uint32_t* texels, int w, int h, int flip);
/* The zlib / DEFLATE decoder the PNG reader and the HDF5 reader (deflate-filtered chunks) share (include/nsdf_inflate.hpp). */
int nsdf_zlib_inflate(const unsigned char* in, size_t n, unsigned char* out, size_t capacity, size_t* n_out);
/* sdkSavePPM4ub / __savePPM: "P6\n<W>\n<H>\n255\n" + RGB. */
int nsdf_ppm_save(const char* path, const uint32_t* texels, int w, int h);

#ifdef __cplusplus
}
#endif
#endif /* NSDF_H_ */
