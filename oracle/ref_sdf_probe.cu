// =====================================================================================
// oracle/ref_sdf_probe.cu -- evaluates the REFERENCE's own analytic-SDF device functions on given points.
//
// TEST INFRASTRUCTURE ONLY (SURVEY 8 f4 pin).  src/volumeRender_kernel.cu:67-196 holds sphere / box / plane /
// cylinder, displace / round / onion, the hard and smooth booleans, manyCylinderCut and manySphere; at HEAD only
// manySphere(..., true) is reachable from sceneSDF, so the reference binary never executes the others.  This file
// #includes the reference's translation unit UNMODIFIED (from where it lies under /root/reference; -I on the command
// line, oracle/build_ref.sh) and adds one kernel that calls those functions, so that their arithmetic -- as nvcc
// compiles the reference's own source text with the reference's flags -- can be recorded on a B200
// (tools/make_ref_golden.py -> tests/golden/ref_cuda_sdf.npz) and compared with the scene programs of the product
// (csrc/scene_program.cuh) and of the oracle.
//
//   ref_sdf_probe POINTS.f32 NSDF.f32 FRAME OUT.f32
//       POINTS: n*3 floats, NSDF: n floats (the "network output" fed to the functions), OUT: n*REF_SDF_FUNCS floats,
//       function-major ([f*n + i]).
// =====================================================================================
#include "volumeRender_kernel.cu"   // the reference's file, resolved through -I/root/reference/src

#define REF_SDF_FUNCS 16

__global__ void refSdfProbe(const float* __restrict__ xyz, const float* __restrict__ nsdf, int n, float* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float3 p = make_float3(xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]);
    const float s = nsdf[i];
    // the commented scene of sceneSDF (:224-228); its offsets as float literals, which is what a float-parameter scene program
    // can express (the source's double literals 0.2 / 0.4 would round p.y + 0.2 differently in rare cases)
    const float3 boxp = make_float3(p.x + 0.5f, p.y + 0.2f, p.z - 0.4f);
    const float3 boxb = make_float3(0.1, 0.1, 0.1);
    out[0 * n + i] = sdfSphere(p, 0.9);                                              // :219
    out[1 * n + i] = sdfBox(boxp, boxb);
    out[2 * n + i] = sdfBox(boxp, boxb, 0.04);
    out[3 * n + i] = sdfPlane(p);
    out[4 * n + i] = sdfCylinder(boxp, boxb);
    out[5 * n + i] = displacementPattern(p, s);                                      // :218
    out[6 * n + i] = sdfOpRound(tanh(s), 0.04);                                      // :220
    out[7 * n + i] = sdfOpOnion(s, 0.02);
    out[8 * n + i] = sdfOpSmoothSubtraction(sdfCylinder(boxp, boxb), tanh(s), 0.05); // :227
    out[9 * n + i] = manyCylinderCut(p, s);                                          // :222
    out[10 * n + i] = manySphere(p, s, false);
    out[11 * n + i] = manySphere(p, s, true);                                        // :221 (HEAD)
    out[12 * n + i] = sdfOpSmoothUnion(s, sdfSphere(p, 0.5), 0.05);
    out[13 * n + i] = tanh(s);                                                       // :229
    out[14 * n + i] = sdfOpSubtraction(s, sdfSphere(p, 0.5));
    out[15 * n + i] = sdfOpIntersect(sdfOpUnion(s, sdfPlane(p)), sdfSphere(p, 0.9));
}

static std::vector<float> probe_read(const char* path) {
    FILE* f = std::fopen(path, "rb");
    if (!f) { std::fprintf(stderr, "cannot open %s\n", path); std::exit(2); }
    std::fseek(f, 0, SEEK_END);
    const long bytes = std::ftell(f);
    std::fseek(f, 0, SEEK_SET);
    std::vector<float> v((size_t)bytes / 4);
    if (std::fread(v.data(), 4, v.size(), f) != v.size()) { std::fprintf(stderr, "short read %s\n", path); std::exit(2); }
    std::fclose(f);
    return v;
}

int main(int argc, char** argv) {
    if (argc < 5) {
        std::fprintf(stderr, "usage: ref_sdf_probe POINTS.f32 NSDF.f32 FRAME OUT.f32\n");
        return 1;
    }
    const std::vector<float> pts = probe_read(argv[1]), sd = probe_read(argv[2]);
    const int n = (int)sd.size();
    if (pts.size() != (size_t)n * 3) { std::fprintf(stderr, "size mismatch\n"); return 2; }
    float identity[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
    copyViewMatrices(identity, 48, identity, 64, atoi(argv[3]), 3);   // sets c_frameNumber (manySphere :182)
    float *d_p = nullptr, *d_s = nullptr, *d_o = nullptr;
    checkCudaErrors(cudaMalloc(&d_p, pts.size() * 4));
    checkCudaErrors(cudaMalloc(&d_s, sd.size() * 4));
    checkCudaErrors(cudaMalloc(&d_o, (size_t)n * REF_SDF_FUNCS * 4));
    checkCudaErrors(cudaMemcpy(d_p, pts.data(), pts.size() * 4, cudaMemcpyHostToDevice));
    checkCudaErrors(cudaMemcpy(d_s, sd.data(), sd.size() * 4, cudaMemcpyHostToDevice));
    refSdfProbe<<<(n + 127) / 128, 128>>>(d_p, d_s, n, d_o);
    checkCudaErrors(cudaDeviceSynchronize());
    std::vector<float> out((size_t)n * REF_SDF_FUNCS);
    checkCudaErrors(cudaMemcpy(out.data(), d_o, out.size() * 4, cudaMemcpyDeviceToHost));
    FILE* f = std::fopen(argv[4], "wb");
    if (!f || std::fwrite(out.data(), 4, out.size(), f) != out.size()) { std::fprintf(stderr, "cannot write %s\n", argv[4]); return 2; }
    std::fclose(f);
    std::printf("{\"mode\": \"sdf_probe\", \"n\": %d, \"funcs\": %d}\n", n, REF_SDF_FUNCS);
    return 0;
}
