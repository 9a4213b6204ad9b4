// =====================================================================================
// mlp_generic.cuh -- fused evaluator for ANY dense stack the reference can load.
//
// The reference's forward is a loop over whatever DenseLayers NeuralNetwork::load built
// (src/neuralNetwork.cpp:54-63, src/layers/denseLayer.cu:180-278): any widths, any depth; its renderer runs any such
// network with 3 or 4 inputs and one output (src/volumeRender_kernel.cu:652-689).  The shipped shapes are all 3 -> 32 x 8 -> 1
// and take the tuned evaluators of mlp_fp32.cuh; every OTHER shape up to 128 neurons per layer and 16 layers takes this
// one: one thread = one point, all layers in ONE launch, activations in two thread-private arrays, weights read with
// warp-uniform loads from shared memory (when the whole network fits) or through the read-only cache.
// Arithmetic = the reference's CUTLASS SIMT order (one accumulator from 0, acc = fma(W[o][k], a[k], acc) for k ascending,
// z = acc + b, ReLU on hidden layers): bit-identical to the oracle and to the per-layer operators.
// =====================================================================================
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace nsdf {

constexpr int GENERIC_MAX_LAYERS = 16;
constexpr int GENERIC_MAX_WIDTH = 128;
constexpr size_t GENERIC_SMEM_LIMIT = 200 * 1024;   // stage the packed parameters in shared memory up to this size

struct GenericNet {
    const float* params;                  // device: per layer W[out][in] then b[out] (nsdf_model::d_params)
    int n_layers;
    int dims[GENERIC_MAX_LAYERS + 1];     // dims[0] = inputs, dims[l + 1] = outputs of layer l
    unsigned int w_off[GENERIC_MAX_LAYERS], b_off[GENERIC_MAX_LAYERS];   // offsets in floats
    unsigned int n_params;                // floats in total
    int staged;                           // 1: the launch provides n_params floats of dynamic shared memory
};

// All threads of the CTA: copies the parameters into shared memory when the launch staged them; returns the base to read from.
__device__ __forceinline__ const float* generic_stage(const GenericNet& g, float* s_params) {
    if (!g.staged) return g.params;
    for (unsigned int i = threadIdx.x; i < g.n_params; i += blockDim.x) s_params[i] = __ldg(g.params + i);
    __syncthreads();
    return s_params;
}

// x: dims[0] inputs in a[0..); on return the outputs of the last layer are in a[0..dims[n_layers]).
__device__ __forceinline__ void mlp_generic_eval(const GenericNet& g, const float* __restrict__ P, float (&a)[GENERIC_MAX_WIDTH],
                                                 float (&z)[GENERIC_MAX_WIDTH]) {
#pragma unroll 1
    for (int l = 0; l < g.n_layers; ++l) {
        const int in = g.dims[l], out = g.dims[l + 1];
        const float* W = P + g.w_off[l];
        const float* b = P + g.b_off[l];
        const bool hidden = l + 1 < g.n_layers;
#pragma unroll 1
        for (int o = 0; o < out; ++o) {
            const float* w = W + (size_t)o * in;
            float acc = 0.0f;
#pragma unroll 4
            for (int k = 0; k < in; ++k) acc = fmaf(w[k], a[k], acc);
            const float v = acc + b[o];
            z[o] = hidden ? ((v > 0.0f) ? v : 0.0f) : v;
        }
#pragma unroll 1
        for (int o = 0; o < out; ++o) a[o] = z[o];
    }
}

}  // namespace nsdf
