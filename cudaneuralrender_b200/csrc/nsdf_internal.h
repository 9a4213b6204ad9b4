// Internal declarations shared by the translation units of libnsdf.so (not part of the ABI).
#pragma once
#include <cstdint>
#include <cstddef>
#include <string>
#include <vector>
#include <cuda_runtime.h>
#include "../../include/nsdf.h"

struct nsdf_model {
    int device = 0;
    int n_layers = 0;
    std::vector<int> in, out, act;
    std::vector<float> h_params;          // packed: per layer W[out][in] then b[out]
    std::vector<size_t> w_off, b_off;     // offsets (floats) into the packed array
    float* d_params = nullptr;            // same layout on the device (per-layer operators)
    bool fused = false;                   // 3 -> 32 x k -> 1, ReLU hidden, linear last
    bool generic = false;                 // any other stack mlp_generic.cuh covers: <= 16 layers of <= 128 neurons, ReLU hidden, linear last
    int n_sq = 0;                         // number of 32x32 layers
    float* d_blob = nullptr;              // fused fp32 blob (mlp_fp32.cuh)
    uint32_t blob_bytes = 0;
    void* d_tc = nullptr;                 // tensor-core operand images (mlp_tc.cu), lazily built
    size_t tc_bytes = 0;
    // render scratch, grown on demand (the reference keeps file-scope globals,
    // src/volumeRender_kernel.cu:578-585); one render at a time per model handle
    unsigned int* d_counters = nullptr;   // [0] pixel counter [1] queue tail [2] queue head [3] pad
    void* d_queue = nullptr;
    void* d_xqueue = nullptr;             // express queue of the cooperative tracer (trace_coop.cuh), grown with the views of a call
    size_t xqueue_capacity = 0;           // entries
    void* d_views = nullptr;              // per-view cameras of a batched call (nsdf_render_views)
    size_t views_capacity = 0;
    void* d_frame_stage = nullptr;        // staged frame stores: [jobs][64 x 64] pixels, then [jobs] tile counters
    size_t frame_stage_jobs = 0;
    size_t queue_capacity = 0;
    unsigned int frame_seq = 0;           // publication stamp of the fused tracer's queue entries
    nsdf_render_stats* d_stats_scratch = nullptr;
    int* d_tile_scratch = nullptr;        // [0,cap) probe scores, [cap,2cap) tile order (tile_priority.cuh)
    size_t tile_scratch_capacity = 0;     // tiles
    float* d_act[2] = {nullptr, nullptr}; // ping-pong activations for the per-layer path
    size_t act_capacity = 0;              // floats per buffer
    int sm_count = 0;
    // device staging of the host-buffer entry points (nsdf_render_host / nsdf_mlp_forward_host), grown on demand
    void* d_stage[3] = {nullptr, nullptr, nullptr};   // [0] frame / y, [1] matcap / x, [2] stats
    size_t stage_bytes[3] = {0, 0, 0};
};

// cudaFuncSetAttribute is per DEVICE: a once-flag must be too (a process may drive several devices through nsdf_set_device).
#define NSDF_PER_DEVICE_FLAG(name)                  \
    static bool name##_by_device[64] = {};          \
    int name##_dev = 0;                             \
    cudaGetDevice(&name##_dev);                     \
    bool& name = name##_by_device[name##_dev & 63]

namespace nsdf {

// fp32 rungs + per-layer operator (nsdf_mlp.cu)
int fp32_forward(nsdf_model* m, const float* d_x, float* d_y, int64_t n, int precision, bool lattice, int res,
                 float lo, float hi, int64_t first, cudaStream_t stream);
int generic_forward(nsdf_model* m, const float* d_x, float* d_y, int64_t n, cudaStream_t stream);
struct GenericNet;
void generic_net_of(const nsdf_model* m, GenericNet* g);
int dense_layer_launch(const float* d_W, const float* d_A, float* d_Z, const float* d_b, int M, int K,
                       int64_t num_batches, int activation, cudaStream_t stream);

// renderer (nsdf_render.cu)
int scene_eval_launch(const nsdf_scene_op* program, int n_ops, const float* d_xyz, const float* d_nsdf, int frame, int math,
                      float* d_out, int64_t n, cudaStream_t stream);
int render_launch(nsdf_model* m, const nsdf_camera* cams, int n_views, const uint32_t* d_matcap, int mw, int mh,
                  const nsdf_render_opts* opts, uint32_t* d_out, int W, int H, nsdf_render_stats* d_stats,
                  cudaStream_t stream);
void render_release(nsdf_model* m);
// allocation-free single-view render on the file-scope scratch array of nsdf_render.cu (nsdf_render_noalloc)
int render_noalloc_launch(const nsdf_model* m, const nsdf_camera* cam, const uint32_t* d_matcap, int mw, int mh,
                          const nsdf_render_opts* opts, uint32_t* d_out, int W, int H, nsdf_render_stats* d_stats,
                          cudaStream_t stream);

// tensor-core path (mlp_tc.cu)
int tc_forward(nsdf_model* m, const float* d_x, float* d_y, int64_t n, int precision, bool lattice, int res, float lo,
               float hi, int64_t first, cudaStream_t stream);
void tc_release(nsdf_model* m);
void tc_set_trace(long long* d_buf);
// device pointer + size of the tensor-core weight image of one rung (tc_mlp.cuh MODE_*), built lazily
const uint8_t* tc_image(nsdf_model* m, int mode, uint32_t* bytes);

}  // namespace nsdf
