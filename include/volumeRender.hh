// The three C-linkage entry points main.cpp calls in the reference (src/volumeRender_kernel.cu:608-706,
// declared in src/main.cpp:97-109), kept with their signatures, over the C ABI (nsdf_render).
//   render_kernel      gridSize / blockSize are accepted and ignored: the persistent tracer sizes its own grid.
//                      d_output is the caller's device buffer of imageW*imageH packed texels; every pixel is
//                      written (0 where the reference leaves the caller's memset).  The network is BORROWED
//                      (the reference's by-value declaration destroys the caller's layers).
//   copyViewMatrices   c_invViewMatrix (3x4) and c_normalMatrix (4x4) + frame number, as cudaMemcpyToSymbol did.
//   copyStaticSettings c_coloringType / c_numInputs.
// Defaults are those of the reference's HEAD build (sceneSDF = manySphere, reference math; nsdf_render_opts_default).
// Extensions: setRenderOptions (scene function, MAX_STEPS, NSDF_RENDER_* flags), setMathMode, setSceneProgram, getLastRenderStats,
// updateViewMatrices (src/main.cpp:207-222 without Eigen).
#pragma once
#include <cstddef>
#if defined(__has_include)
#if __has_include(<vector_types.h>)
#include <vector_types.h>
#define NSDF_HAVE_DIM3 1
#endif
#endif
#ifndef NSDF_HAVE_DIM3
struct dim3 {
    unsigned int x, y, z;
    dim3(unsigned int x = 1, unsigned int y = 1, unsigned int z = 1) : x(x), y(y), z(z) {}
};
#endif
#include "neuralNetwork.hh"
#include "neuralUtils/image.hh"

extern "C" void render_kernel(dim3 gridSize, dim3 blockSize, uint* d_output, uint imageW, uint imageH, uint numInputs,
                              NeuralNetwork& nn, Image matcap);
extern "C" void copyViewMatrices(float* invViewMatrix, size_t sizeofInvModelViewMat, float* normalMatrix,
                                 size_t sizeofNormalMatrix, int frameNumber, int numInputs);
extern "C" void copyStaticSettings(int colorType, int numInputs);

extern "C" void setRenderOptions(int scene, int maxSteps, int flags);
// NSDF_MATH_REFERENCE (default: the reference binary's frames) or NSDF_MATH_DETERMINISTIC (the CPU oracle's frames)
extern "C" void setMathMode(int math);
// sceneSDF (volumeRender_kernel.cu:217-230) as data: the op list is copied; selects NSDF_SCENE_PROGRAM (include/nsdf.h)
extern "C" void setSceneProgram(const nsdf_scene_op* ops, int nOps);
// the options render_kernel would use now (for callers of nsdf_render_views / nsdf_multi_render)
extern "C" void getRenderOptions(nsdf_render_opts* out);
extern "C" void getLastRenderStats(nsdf_render_stats* out);
extern "C" void updateViewMatrices(float rxDeg, float ryDeg, float tx, float ty, float tz, int frameNumber);
