// placeholder until the tcgen05 kernel lands (next commit)
#include "nsdf_internal.h"
namespace nsdf {
int tc_forward(nsdf_model*, const float*, float*, int64_t, int, bool, int, float, float, int64_t, cudaStream_t) {
    return NSDF_EUNSUPPORTED;
}
void tc_release(nsdf_model* m) {
    if (m->d_tc) cudaFree(m->d_tc);
    m->d_tc = nullptr;
}
}  // namespace nsdf
