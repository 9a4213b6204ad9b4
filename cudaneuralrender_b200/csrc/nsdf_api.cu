// =====================================================================================
// nsdf_api.cu -- the C ABI of libnsdf.so (include/nsdf.h): model handles, dispatch to the
// fused MLP kernels and the tracer, host-buffer variants, IPC helpers, camera helper.
// No CPU fallback: every compute entry point runs CUDA kernels or returns an error.
// =====================================================================================
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>

#include "../../include/nsdf_h5.hpp"
#include "mlp_fp32.cuh"
#include "nsdf_internal.h"

namespace {

#define NSDF_TRY(expr)                        \
    do {                                      \
        cudaError_t e__ = (expr);             \
        if (e__ != cudaSuccess) return (int)e__; \
    } while (0)

// Packs the reference's per-layer [out][in] weights into the fused blob of mlp_fp32.cuh
// (transposed to [k][o] so that one warp-uniform 16-byte read feeds four outputs).
void build_fused_blob(const nsdf_model& m, std::vector<float>& blob) {
    const int n_sq = m.n_sq;
    blob.assign((size_t)nsdf::fused_blob_floats(n_sq), 0.0f);
    const float* P = m.h_params.data();
    {   // layer 0: in0 (3 or 4) -> 32
        const float* W = P + m.w_off[0];
        const float* b = P + m.b_off[0];
        const int in0 = m.in[0];
        for (int k = 0; k < in0; ++k)
            for (int o = 0; o < 32; ++o) blob[(size_t)k * 32 + o] = W[(size_t)o * in0 + k];
        for (int o = 0; o < 32; ++o) blob[128 + o] = b[o];
    }
    for (int l = 0; l < n_sq; ++l) {
        const float* W = P + m.w_off[l + 1];
        const float* b = P + m.b_off[l + 1];
        float* dst = blob.data() + nsdf::FUSED_L0_FLOATS + (size_t)l * nsdf::FUSED_SQ_FLOATS;
        for (int k = 0; k < 32; ++k)
            for (int o = 0; o < 32; ++o) dst[k * 32 + o] = W[o * 32 + k];
        for (int o = 0; o < 32; ++o) dst[1024 + o] = b[o];
    }
    {
        const int L = m.n_layers - 1;
        const float* W = P + m.w_off[L];
        const float* b = P + m.b_off[L];
        float* dst = blob.data() + nsdf::FUSED_L0_FLOATS + (size_t)n_sq * nsdf::FUSED_SQ_FLOATS;
        for (int k = 0; k < 32; ++k) dst[k] = W[k];
        dst[32] = b[0];
    }
}

int finish_model(nsdf_model* m) {
    const int L = m->n_layers;
    // fused shape: (3|4) -> 32, k x (32 -> 32), 32 -> 1; ReLU hidden, linear last
    bool fused = L >= 2 && (m->in[0] == 3 || m->in[0] == 4) && m->out[L - 1] == 1 && m->act[L - 1] == NSDF_ACT_TANH_LINEAR;
    for (int l = 0; l < L && fused; ++l) {
        if (l < L - 1 && (m->out[l] != 32 || m->act[l] != NSDF_ACT_RELU)) fused = false;
        if (l > 0 && m->in[l] != 32) fused = false;
    }
    if (fused && L - 2 > nsdf::FUSED_MAX_SQ) fused = false;
    m->fused = fused;
    m->n_sq = fused ? L - 2 : 0;
    bool generic = L >= 1 && L <= 16 && m->in[0] >= 1 && m->in[0] <= 128 && m->act[L - 1] == NSDF_ACT_TANH_LINEAR;
    for (int l = 0; l < L && generic; ++l) {
        if (m->out[l] < 1 || m->out[l] > 128) generic = false;
        if (l < L - 1 && m->act[l] != NSDF_ACT_RELU) generic = false;
        if (l > 0 && m->in[l] != m->out[l - 1]) generic = false;
    }
    m->generic = generic;
    NSDF_TRY(cudaGetDevice(&m->device));
    NSDF_TRY(cudaDeviceGetAttribute(&m->sm_count, cudaDevAttrMultiProcessorCount, m->device));
    NSDF_TRY(cudaMalloc(&m->d_params, m->h_params.size() * sizeof(float)));
    NSDF_TRY(cudaMemcpy(m->d_params, m->h_params.data(), m->h_params.size() * sizeof(float), cudaMemcpyHostToDevice));
    if (fused) {
        std::vector<float> blob;
        build_fused_blob(*m, blob);
        m->blob_bytes = (uint32_t)(blob.size() * sizeof(float));
        NSDF_TRY(cudaMalloc(&m->d_blob, m->blob_bytes));
        NSDF_TRY(cudaMemcpy(m->d_blob, blob.data(), m->blob_bytes, cudaMemcpyHostToDevice));
    }
    return NSDF_OK;
}

int ensure_act(nsdf_model* m, size_t floats) {
    if (floats <= m->act_capacity) return NSDF_OK;
    for (int i = 0; i < 2; ++i) {
        if (m->d_act[i]) cudaFree(m->d_act[i]);
        m->d_act[i] = nullptr;
    }
    m->act_capacity = 0;
    NSDF_TRY(cudaMalloc(&m->d_act[0], floats * sizeof(float)));
    NSDF_TRY(cudaMalloc(&m->d_act[1], floats * sizeof(float)));
    m->act_capacity = floats;
    return NSDF_OK;
}

int ensure_stage(nsdf_model* m, int slot, size_t bytes) {
    if (bytes <= m->stage_bytes[slot]) return NSDF_OK;
    if (m->d_stage[slot]) cudaFree(m->d_stage[slot]);
    m->d_stage[slot] = nullptr;
    m->stage_bytes[slot] = 0;
    NSDF_TRY(cudaMalloc(&m->d_stage[slot], bytes));
    m->stage_bytes[slot] = bytes;
    return NSDF_OK;
}

// Layer-by-layer path for shapes the fused kernels do not cover (same arithmetic).
int layered_forward(nsdf_model* m, const float* d_x, float* d_y, int64_t n, cudaStream_t stream) {
    int maxw = 0;
    for (int l = 0; l < m->n_layers; ++l) maxw = m->out[l] > maxw ? m->out[l] : maxw;
    int rc = ensure_act(m, (size_t)maxw * (size_t)n);
    if (rc) return rc;
    const float* a = d_x;
    for (int l = 0; l < m->n_layers; ++l) {
        float* z = (l == m->n_layers - 1) ? d_y : m->d_act[l & 1];
        rc = nsdf::dense_layer_launch(m->d_params + m->w_off[l], a, z, m->d_params + m->b_off[l], m->out[l], m->in[l], n,
                                      m->act[l], stream);
        if (rc) return rc;
        a = z;
    }
    return NSDF_OK;
}

int forward_dispatch(nsdf_model* m, const float* d_x, float* d_y, int64_t n, int precision, bool lattice, int res,
                     float lo, float hi, int64_t first, cudaStream_t stream) {
    if (n < 0) return NSDF_EINVAL;
    if (n == 0) return NSDF_OK;
    switch (precision) {
        case NSDF_PREC_FP32:
        case NSDF_PREC_FP32_WARP:
            if (!m->fused || m->in[0] != 3) {
                if (lattice) return NSDF_EUNSUPPORTED;
                if (m->generic) return nsdf::generic_forward(m, d_x, d_y, n, stream);   // other widths / depths: still ONE fused launch
                return layered_forward(m, d_x, d_y, n, stream);
            }
            return nsdf::fp32_forward(m, d_x, d_y, n, precision, lattice, res, lo, hi, first, stream);
        case NSDF_PREC_TC_FP16:
        case NSDF_PREC_TC_BF16:
        case NSDF_PREC_TC_FP16X3:
        case NSDF_PREC_TC_FP16_ACC16:
            if (!m->fused || m->in[0] != 3) return NSDF_EUNSUPPORTED;
            return nsdf::tc_forward(m, d_x, d_y, n, precision, lattice, res, lo, hi, first, stream);
        default:
            return NSDF_EINVAL;
    }
}

}  // namespace

extern "C" {

/* development hook (not in include/nsdf.h): clock64 trace buffer of the dense TC kernel, 1024 int64 */
void nsdf_debug_tc_trace(long long* d_buf) { nsdf::tc_set_trace(d_buf); }

const char* nsdf_version(void) { return "nsdf-b200 0.1 (sm_100a)"; }

const char* nsdf_error_string(int status) {
    switch (status) {
        case NSDF_OK: return "success";
        case NSDF_EINVAL: return "nsdf: invalid argument";
        case NSDF_EUNSUPPORTED: return "nsdf: model shape not supported by the requested kernel";
        case NSDF_EIO: return "nsdf: file could not be read or parsed";
        case NSDF_ENOMEM: return "nsdf: out of host memory";
        case NSDF_ETOOLARGE: return "nsdf: size beyond the fixed maximum of an allocation-free entry point (NSDF_NOALLOC_MAX_TILES)";
        default: return cudaGetErrorString((cudaError_t)status);
    }
}

int nsdf_device_sm_count(void) {
    int dev = 0, sms = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 0;
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return 0;
    return sms;
}

int nsdf_dev_malloc(void** p, size_t bytes) {
    if (!p) return NSDF_EINVAL;
    *p = nullptr;
    return bytes ? (int)cudaMalloc(p, bytes) : NSDF_OK;
}
int nsdf_dev_free(void* p) { return p ? (int)cudaFree(p) : NSDF_OK; }
int nsdf_dev_memset(void* p, int v, size_t bytes) { return bytes ? (int)cudaMemset(p, v, bytes) : NSDF_OK; }
int nsdf_memcpy_h2d(void* d, const void* h, size_t bytes) { return bytes ? (int)cudaMemcpy(d, h, bytes, cudaMemcpyHostToDevice) : NSDF_OK; }
int nsdf_memcpy_d2h(void* h, const void* d, size_t bytes) { return bytes ? (int)cudaMemcpy(h, d, bytes, cudaMemcpyDeviceToHost) : NSDF_OK; }
int nsdf_device_synchronize(void) { return (int)cudaDeviceSynchronize(); }
int nsdf_set_device(int ordinal) { return (int)cudaSetDevice(ordinal); }
int nsdf_device_count(void) {
    int n = 0;
    return cudaGetDeviceCount(&n) == cudaSuccess ? n : 0;
}

int nsdf_model_create(const float* const* W, const float* const* b, const int* in_dim, const int* out_dim,
                      const int* activation, int n_layers, nsdf_model** out) {
    if (!W || !b || !in_dim || !out_dim || !activation || n_layers <= 0 || !out) return NSDF_EINVAL;
    for (int l = 0; l < n_layers; ++l) {
        if (in_dim[l] <= 0 || out_dim[l] <= 0 || !W[l] || !b[l]) return NSDF_EINVAL;
        if (l > 0 && in_dim[l] != out_dim[l - 1]) return NSDF_EINVAL;
        if (activation[l] != NSDF_ACT_RELU && activation[l] != NSDF_ACT_TANH_LINEAR) return NSDF_EINVAL;
    }
    nsdf_model* m = new (std::nothrow) nsdf_model();
    if (!m) return NSDF_ENOMEM;
    m->n_layers = n_layers;
    for (int l = 0; l < n_layers; ++l) {
        m->in.push_back(in_dim[l]);
        m->out.push_back(out_dim[l]);
        m->act.push_back(activation[l]);
        m->w_off.push_back(m->h_params.size());
        m->h_params.insert(m->h_params.end(), W[l], W[l] + (size_t)in_dim[l] * out_dim[l]);
        m->b_off.push_back(m->h_params.size());
        m->h_params.insert(m->h_params.end(), b[l], b[l] + out_dim[l]);
    }
    const int rc = finish_model(m);
    if (rc) { nsdf_model_destroy(m); return rc; }
    *out = m;
    return NSDF_OK;
}

int nsdf_model_create_packed(const float* params, const int* dims, int n_layers, nsdf_model** out) {
    if (!params || !dims || n_layers <= 0 || n_layers > 4096) return NSDF_EINVAL;
    std::vector<const float*> W(n_layers), b(n_layers);
    std::vector<int> in(n_layers), od(n_layers), act(n_layers);
    const float* p = params;
    for (int l = 0; l < n_layers; ++l) {
        if (dims[l] <= 0 || dims[l + 1] <= 0) return NSDF_EINVAL;
        in[l] = dims[l];
        od[l] = dims[l + 1];
        act[l] = (l == n_layers - 1) ? NSDF_ACT_TANH_LINEAR : NSDF_ACT_RELU;   // src/neuralNetwork.cpp:136-139
        W[l] = p; p += (size_t)in[l] * od[l];
        b[l] = p; p += od[l];
    }
    return nsdf_model_create(W.data(), b.data(), in.data(), od.data(), act.data(), n_layers, out);
}

// Walks the file like NeuralNetwork::load and packs W[out][in] + b[out] per layer.
static int read_keras_packed(const char* path, std::vector<int>& dims, std::vector<float>& params) {
    std::vector<nsdf::h5::KerasLayer> layers;
    try {
        nsdf::h5::File f(path);
        std::string err;
        if (!nsdf::h5::read_keras_layers(f, layers, err)) {
            std::printf("%s\n", err.c_str());   // NeuralNetwork::load prints and returns false
            return NSDF_EIO;
        }
    } catch (const std::exception& e) {
        std::fprintf(stderr, "nsdf: reading %s: %s\n", path, e.what());
        return NSDF_EIO;
    }
    dims.clear();
    params.clear();
    for (size_t l = 0; l < layers.size(); ++l) {
        const int in = (int)layers[l].kernel_dims[0], od = (int)layers[l].kernel_dims[1];
        if (l == 0) dims.push_back(in);
        else if (in != dims.back()) return NSDF_EIO;
        dims.push_back(od);
        // DenseLayer::initializeWeights, src/layers/denseLayer.cu:217-227: W[y*in + x] = weights[x][y]
        const size_t base = params.size();
        params.resize(base + (size_t)in * od);
        for (int x = 0; x < in; ++x)
            for (int y = 0; y < od; ++y) params[base + (size_t)y * in + x] = layers[l].kernel[(size_t)x * od + y];
        params.insert(params.end(), layers[l].bias.begin(), layers[l].bias.end());
    }
    return NSDF_OK;
}

int nsdf_h5_read_keras(const char* path, int* n_layers_out, int* dims_out, int dims_capacity, float* params_out,
                       size_t params_capacity, size_t* n_params_out) {
    if (!path) return NSDF_EINVAL;
    std::vector<int> dims;
    std::vector<float> params;
    const int rc = read_keras_packed(path, dims, params);
    if (rc) return rc;
    if (n_layers_out) *n_layers_out = (int)dims.size() - 1;
    if (n_params_out) *n_params_out = params.size();
    if (dims_out) {
        if (dims_capacity < (int)dims.size()) return NSDF_EINVAL;
        std::memcpy(dims_out, dims.data(), dims.size() * sizeof(int));
    }
    if (params_out) {
        if (params_capacity < params.size()) return NSDF_EINVAL;
        std::memcpy(params_out, params.data(), params.size() * sizeof(float));
    }
    return NSDF_OK;
}

int nsdf_h5_write_keras(const char* path, const float* params, const int* dims, int n_layers) {
    return nsdf_h5_write_keras_ex(path, params, dims, n_layers, NSDF_H5_CONTIGUOUS);
}

int nsdf_h5_write_keras_ex(const char* path, const float* params, const int* dims, int n_layers, int storage) {
    if (!path || !params || !dims || n_layers <= 0 || n_layers > 256) return NSDF_EINVAL;
    if (storage < NSDF_H5_CONTIGUOUS || storage > NSDF_H5_CHUNKED_DEFLATE) return NSDF_EINVAL;
    std::vector<nsdf::h5::KerasLayer> layers((size_t)n_layers);
    const float* p = params;
    for (int l = 0; l < n_layers; ++l) {
        const int in = dims[l], od = dims[l + 1];
        if (in <= 0 || od <= 0) return NSDF_EINVAL;
        nsdf::h5::KerasLayer& L = layers[(size_t)l];
        L.name = l ? "dense_" + std::to_string(l) : std::string("dense");
        L.kernel_dims = {(size_t)in, (size_t)od};
        L.kernel.resize((size_t)in * od);
        for (int x = 0; x < in; ++x)   // inverse of DenseLayer::initializeWeights (denseLayer.cu:217-227)
            for (int y = 0; y < od; ++y) L.kernel[(size_t)x * od + y] = p[(size_t)y * in + x];
        p += (size_t)in * od;
        L.bias.assign(p, p + od);
        p += od;
    }
    try {
        nsdf::h5::write_keras_weights(path, layers, storage);
    } catch (const std::exception& e) {
        std::fprintf(stderr, "nsdf: writing %s: %s\n", path, e.what());
        return NSDF_EIO;
    }
    return NSDF_OK;
}

int nsdf_model_load_h5(const char* path, nsdf_model** out) {
    if (!path || !out) return NSDF_EINVAL;
    std::vector<int> dims;
    std::vector<float> params;
    const int rc = read_keras_packed(path, dims, params);
    if (rc) return rc;
    return nsdf_model_create_packed(params.data(), dims.data(), (int)dims.size() - 1, out);
}

void nsdf_model_destroy(nsdf_model* m) {
    if (!m) return;
    int prev = -1;
    cudaGetDevice(&prev);
    if (prev != m->device) cudaSetDevice(m->device);
    nsdf::tc_release(m);
    nsdf::render_release(m);
    if (m->d_params) cudaFree(m->d_params);
    if (m->d_blob) cudaFree(m->d_blob);
    for (int i = 0; i < 2; ++i)
        if (m->d_act[i]) cudaFree(m->d_act[i]);
    for (int i = 0; i < 3; ++i)
        if (m->d_stage[i]) cudaFree(m->d_stage[i]);
    if (prev >= 0 && prev != m->device) cudaSetDevice(prev);
    delete m;
}

int nsdf_model_num_layers(const nsdf_model* m) { return m ? m->n_layers : 0; }

int nsdf_model_dims(const nsdf_model* m, int* dims_out) {
    if (!m || !dims_out) return NSDF_EINVAL;
    dims_out[0] = m->in[0];
    for (int l = 0; l < m->n_layers; ++l) dims_out[l + 1] = m->out[l];
    return NSDF_OK;
}

int nsdf_model_is_fused(const nsdf_model* m) { return (m && m->fused) ? 1 : 0; }

int nsdf_model_params(const nsdf_model* m, float* params_out, size_t capacity, size_t* n_out) {
    if (!m) return NSDF_EINVAL;
    if (n_out) *n_out = m->h_params.size();
    if (!params_out) return NSDF_OK;
    if (capacity < m->h_params.size()) return NSDF_EINVAL;
    std::memcpy(params_out, m->h_params.data(), m->h_params.size() * sizeof(float));
    return NSDF_OK;
}

int nsdf_mlp_forward(const nsdf_model* m, const float* d_x, float* d_y, int64_t n, int precision, void* stream) {
    if (!m || (n > 0 && (!d_x || !d_y))) return NSDF_EINVAL;
    return forward_dispatch(const_cast<nsdf_model*>(m), d_x, d_y, n, precision, false, 0, 0.f, 0.f, 0, (cudaStream_t)stream);
}

int nsdf_mlp_forward_lattice(const nsdf_model* m, float* d_y, int res, float lo, float hi, int64_t first, int64_t count,
                             int precision, void* stream) {
    if (!m || res <= 0 || first < 0 || count < 0 || (count > 0 && !d_y)) return NSDF_EINVAL;
    if (first + count > (int64_t)res * res * res) return NSDF_EINVAL;
    return forward_dispatch(const_cast<nsdf_model*>(m), nullptr, d_y, count, precision, true, res, lo, hi, first,
                            (cudaStream_t)stream);
}

int nsdf_mlp_forward_host(const nsdf_model* cm, const float* h_x, float* h_y, int64_t n, int precision) {
    nsdf_model* m = const_cast<nsdf_model*>(cm);
    if (!m || n < 0 || (n > 0 && (!h_x || !h_y))) return NSDF_EINVAL;
    if (n == 0) return NSDF_OK;
    const size_t xb = (size_t)n * m->in[0] * sizeof(float), yb = (size_t)n * m->out[m->n_layers - 1] * sizeof(float);
    int rc = ensure_stage(m, 1, xb);
    if (rc) return rc;
    rc = ensure_stage(m, 0, yb);
    if (rc) return rc;
    float* d_x = (float*)m->d_stage[1];
    float* d_y = (float*)m->d_stage[0];
    NSDF_TRY(cudaMemcpyAsync(d_x, h_x, xb, cudaMemcpyHostToDevice, 0));
    rc = nsdf_mlp_forward(m, d_x, d_y, n, precision, nullptr);
    if (rc) return rc;
    NSDF_TRY(cudaMemcpyAsync(h_y, d_y, yb, cudaMemcpyDeviceToHost, 0));
    NSDF_TRY(cudaStreamSynchronize(0));
    return NSDF_OK;
}

int nsdf_dense_layer_forward(const float* d_W, const float* d_A, float* d_Z, const float* d_b, int M, int N, int K,
                             int64_t num_batches, int activation, void* stream) {
    if (!d_W || !d_A || !d_Z || !d_b || M <= 0 || K <= 0 || num_batches < 0) return NSDF_EINVAL;
    if (N != 1) return NSDF_EUNSUPPORTED;
    if (activation != NSDF_ACT_RELU && activation != NSDF_ACT_TANH_LINEAR) return NSDF_EINVAL;
    return nsdf::dense_layer_launch(d_W, d_A, d_Z, d_b, M, K, num_batches, activation, (cudaStream_t)stream);
}

void nsdf_render_opts_default(nsdf_render_opts* o) {
    if (!o) return;
    o->color_type = NSDF_COLOR_FACING;
    o->scene = NSDF_SCENE_MANYSPHERE;   // what sceneSDF returns at the reference's HEAD (src/volumeRender_kernel.cu:221)
    o->max_steps = 6000;
    o->tile_rank = 0;
    o->tile_world = 1;
    o->evict_steps = 0;
    o->flags = 0;
    o->program = nullptr;
    o->n_ops = 0;
    o->math = NSDF_MATH_REFERENCE;
    o->tuning = nullptr;
}

void nsdf_tuning_default(nsdf_tuning* t) {
    if (!t) return;
    t->express_period = -1;
    t->express_steps = 160;
    t->express_warps = 4;
    t->express_old_steps = 192;
    t->express_old_mask = 1;
    t->express_from_bulk = 1;
    t->express_rays = 1;
    t->express_solo_steps = 768;
    t->tile_priority_steps = -1;
    t->tile_priority_min_jobs = 128;
    t->tail_lanes = 0;
    t->tail_push_max = 2;
    t->bulk4 = -1;
    t->bulk4_evict_steps = 0;
    t->bulk4_period = 0;
    t->bulk4_drain = 16;
    t->bulk4_refill_min = 16;
    t->bulk4_unroll = 2;
    t->max_ctas = 0;
}

int nsdf_scene_eval(const nsdf_scene_op* program, int n_ops, const float* d_xyz, const float* d_nsdf, int frame, int math,
                    float* d_out, int64_t n, void* stream) {
    if (n < 0 || (n > 0 && (!d_xyz || !d_nsdf || !d_out))) return NSDF_EINVAL;
    if (math != NSDF_MATH_REFERENCE && math != NSDF_MATH_DETERMINISTIC) return NSDF_EINVAL;
    return nsdf::scene_eval_launch(program, n_ops, d_xyz, d_nsdf, frame, math, d_out, n, (cudaStream_t)stream);
}

// updateViewMatrices, src/main.cpp:207-222.  Eigen composes AngleAxis(-rx, X) * AngleAxis(-ry, Y)
// through quaternions and converts to a matrix; modelView = [m | m * (-t)]; normalMatrix = inverse.
// The inverse is taken analytically (rigid transform): [m^T | -m^T (m v)].
void nsdf_camera_from_view(float rx_deg, float ry_deg, float tx, float ty, float tz, int frame, nsdf_camera* cam) {
    if (!cam) return;
    const float ax = (float)(-(double)rx_deg * M_PI / 180), ay = (float)(-(double)ry_deg * M_PI / 180);
    const float cx = cosf(0.5f * ax), sx = sinf(0.5f * ax), cy = cosf(0.5f * ay), sy = sinf(0.5f * ay);
    const float qw = cx * cy, qx = sx * cy, qy = cx * sy, qz = sx * sy;
    const float t2x = 2 * qx, t2y = 2 * qy, t2z = 2 * qz;
    const float twx = t2x * qw, twy = t2y * qw, twz = t2z * qw;
    const float txx = t2x * qx, txy = t2y * qx, txz = t2z * qx;
    const float tyy = t2y * qy, tyz = t2z * qy, tzz = t2z * qz;
    float m[3][3];
    m[0][0] = 1 - (tyy + tzz); m[0][1] = txy - twz;       m[0][2] = txz + twy;
    m[1][0] = txy + twz;       m[1][1] = 1 - (txx + tzz); m[1][2] = tyz - twx;
    m[2][0] = txz - twy;       m[2][1] = tyz + twx;       m[2][2] = 1 - (txx + tyy);
    const float v[3] = {-tx, -ty, -tz};
    float mv[3];
    for (int i = 0; i < 3; ++i) {
        mv[i] = m[i][0] * v[0] + m[i][1] * v[1] + m[i][2] * v[2];
        for (int j = 0; j < 3; ++j) cam->inv_view[4 * i + j] = m[i][j];
        cam->inv_view[4 * i + 3] = mv[i];
    }
    for (int i = 0; i < 3; ++i) {
        for (int j = 0; j < 3; ++j) cam->normal[4 * i + j] = m[j][i];
        cam->normal[4 * i + 3] = -(m[0][i] * mv[0] + m[1][i] * mv[1] + m[2][i] * mv[2]);
    }
    cam->normal[12] = cam->normal[13] = cam->normal[14] = 0.0f;
    cam->normal[15] = 1.0f;
    cam->frame = frame;
}

int nsdf_render_views(const nsdf_model* m, const nsdf_camera* cams, int n_views, const uint32_t* d_matcap, int mw, int mh,
                      const nsdf_render_opts* opts, uint32_t* d_out, int W, int H, nsdf_render_stats* d_stats, void* stream) {
    if (!m || !cams || !d_out || W <= 0 || H <= 0 || n_views < 1) return NSDF_EINVAL;
    nsdf_render_opts o;
    if (opts) o = *opts;
    else nsdf_render_opts_default(&o);
    return nsdf::render_launch(const_cast<nsdf_model*>(m), cams, n_views, d_matcap, mw, mh, &o, d_out, W, H, d_stats,
                               (cudaStream_t)stream);
}

int nsdf_render(const nsdf_model* m, const nsdf_camera* cam, const uint32_t* d_matcap, int mw, int mh,
                const nsdf_render_opts* opts, uint32_t* d_out, int W, int H, nsdf_render_stats* d_stats, void* stream) {
    return nsdf_render_views(m, cam, 1, d_matcap, mw, mh, opts, d_out, W, H, d_stats, stream);
}

int nsdf_render_noalloc(const nsdf_model* m, const nsdf_camera* cam, const uint32_t* d_matcap, int mw, int mh,
                        const nsdf_render_opts* opts, uint32_t* d_out, int W, int H, nsdf_render_stats* d_stats, void* stream) {
    if (!m || !cam || !d_out || W <= 0 || H <= 0) return NSDF_EINVAL;
    nsdf_render_opts o;
    if (opts) o = *opts;
    else nsdf_render_opts_default(&o);
    return nsdf::render_noalloc_launch(m, cam, d_matcap, mw, mh, &o, d_out, W, H, d_stats, (cudaStream_t)stream);
}

int nsdf_render_views_host(const nsdf_model* cm, const nsdf_camera* cams, int n_views, const uint32_t* h_matcap, int mw, int mh,
                           const nsdf_render_opts* opts, uint32_t* h_out, int W, int H, nsdf_render_stats* h_stats) {
    nsdf_model* m = const_cast<nsdf_model*>(cm);
    if (!m || !cams || !h_out || W <= 0 || H <= 0 || n_views < 1) return NSDF_EINVAL;
    const size_t npx = (size_t)W * H * (size_t)n_views;
    const bool have_mat = h_matcap && mw > 0 && mh > 0;
    int rc = ensure_stage(m, 0, npx * sizeof(uint32_t));
    if (rc) return rc;
    if (have_mat && (rc = ensure_stage(m, 1, (size_t)mw * mh * sizeof(uint32_t)))) return rc;
    if (h_stats && (rc = ensure_stage(m, 2, sizeof(nsdf_render_stats)))) return rc;
    uint32_t* d_out = (uint32_t*)m->d_stage[0];
    uint32_t* d_mat = have_mat ? (uint32_t*)m->d_stage[1] : nullptr;
    nsdf_render_stats* d_stats = h_stats ? (nsdf_render_stats*)m->d_stage[2] : nullptr;
    // every pixel of the frame is written by the tracer, so the caller's memset (src/main.cpp:407-408)
    // needs no device-side counterpart
    if (have_mat) NSDF_TRY(cudaMemcpyAsync(d_mat, h_matcap, (size_t)mw * mh * sizeof(uint32_t), cudaMemcpyHostToDevice, 0));
    if (d_stats) NSDF_TRY(cudaMemsetAsync(d_stats, 0, sizeof(nsdf_render_stats), 0));
    rc = nsdf_render_views(m, cams, n_views, d_mat, mw, mh, opts, d_out, W, H, d_stats, nullptr);
    if (rc) return rc;
    NSDF_TRY(cudaMemcpyAsync(h_out, d_out, npx * sizeof(uint32_t), cudaMemcpyDeviceToHost, 0));
    if (d_stats) NSDF_TRY(cudaMemcpyAsync(h_stats, d_stats, sizeof(nsdf_render_stats), cudaMemcpyDeviceToHost, 0));
    NSDF_TRY(cudaStreamSynchronize(0));
    return NSDF_OK;
}

int nsdf_render_host(const nsdf_model* cm, const nsdf_camera* cam, const uint32_t* h_matcap, int mw, int mh,
                     const nsdf_render_opts* opts, uint32_t* h_out, int W, int H, nsdf_render_stats* h_stats) {
    return nsdf_render_views_host(cm, cam, 1, h_matcap, mw, mh, opts, h_out, W, H, h_stats);
}

int nsdf_ipc_export(void* d_ptr, unsigned char handle_out[NSDF_IPC_HANDLE_BYTES]) {
    static_assert(sizeof(cudaIpcMemHandle_t) == NSDF_IPC_HANDLE_BYTES, "IPC handle size");
    if (!d_ptr || !handle_out) return NSDF_EINVAL;
    cudaIpcMemHandle_t h;
    NSDF_TRY(cudaIpcGetMemHandle(&h, d_ptr));
    std::memcpy(handle_out, &h, sizeof(h));
    return NSDF_OK;
}

int nsdf_ipc_open(const unsigned char handle[NSDF_IPC_HANDLE_BYTES], void** d_ptr_out) {
    if (!handle || !d_ptr_out) return NSDF_EINVAL;
    cudaIpcMemHandle_t h;
    std::memcpy(&h, handle, sizeof(h));
    NSDF_TRY(cudaIpcOpenMemHandle(d_ptr_out, h, cudaIpcMemLazyEnablePeerAccess));
    return NSDF_OK;
}

int nsdf_ipc_close(void* d_ptr) {
    if (!d_ptr) return NSDF_EINVAL;
    NSDF_TRY(cudaIpcCloseMemHandle(d_ptr));
    return NSDF_OK;
}

}  // extern "C"
