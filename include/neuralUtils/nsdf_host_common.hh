// Shared helpers of the C++ host layer (the reference-compatible classes over include/nsdf.h).
// Error convention of the reference: CUDA errors print and exit(EXIT_FAILURE)
// (src/common/helper_cuda.h:577-591 checkCudaErrors); kept here as nsdfCheck().
#pragma once
#include <cstdio>
#include <cstdlib>
#include <ctime>
#include "../nsdf.h"

// The reference's matrix.hh pulls in <cuda_runtime.h> (src/neuralUtils/matrix.hh:4), and its callers use a few
// runtime calls directly (src/simpleInfer.cpp:103,130).  With a CUDA toolkit on the include path the real
// header is used; without one the handful of names those callers need map onto the C ABI.
#if defined(__has_include)
#if __has_include(<cuda_runtime.h>)
#include <cuda_runtime.h>
#define NSDF_HAVE_CUDA_RUNTIME_H 1
#endif
#endif
#ifndef NSDF_HAVE_CUDA_RUNTIME_H
typedef int cudaError_t;
#ifndef cudaSuccess
#define cudaSuccess 0
#endif
inline cudaError_t cudaDeviceSynchronize() { return nsdf_device_synchronize(); }
inline cudaError_t cudaGetLastError() { return 0; }
inline const char* cudaGetErrorString(cudaError_t e) { return nsdf_error_string(e); }
#endif

#ifndef NSDF_HOST_UINT_DEFINED
#define NSDF_HOST_UINT_DEFINED
typedef unsigned int uint;   // the reference uses glibc's `uint` (src/neuralUtils/image.hh:25)
#endif

inline void nsdfCheckImpl(int status, const char* what, const char* file, int line) {
    if (status != NSDF_OK) {
        std::fprintf(stderr, "NSDF error at %s:%d code=%d(%s) \"%s\"\n", file, line, status, nsdf_error_string(status), what);
        std::exit(EXIT_FAILURE);
    }
}
#define nsdfCheck(call) nsdfCheckImpl((call), #call, __FILE__, __LINE__)
