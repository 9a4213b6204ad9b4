// lattice.cuh -- in-kernel generation of the dense query lattice (BASELINE config 3).
#pragma once
#include <cuda_runtime.h>

namespace nsdf {

struct LatticeParams { int enabled; int res; float lo, hi; long long first; };

// p = lo + (hi-lo)*(i+0.5)/res per axis, x fastest (BASELINE config 3 / SURVEY 8(d) C3).
__device__ __forceinline__ void lattice_point(const LatticeParams& lp, long long idx, float (&x)[3]) {
    const long long r = lp.res;
    const int ix = (int)(idx % r), iy = (int)((idx / r) % r), iz = (int)(idx / (r * r));
    const float s = __fdiv_rn(lp.hi - lp.lo, (float)lp.res);
    x[0] = fmaf((float)ix + 0.5f, s, lp.lo);
    x[1] = fmaf((float)iy + 0.5f, s, lp.lo);
    x[2] = fmaf((float)iz + 0.5f, s, lp.lo);
}

}  // namespace nsdf
