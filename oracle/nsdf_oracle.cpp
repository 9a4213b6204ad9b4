// =====================================================================================
// oracle/nsdf_oracle.cpp -- CPU restatement of the reference's neural-SDF hot path.
//
// TEST INFRASTRUCTURE ONLY.  Nothing under cudaneuralrender_b200/, include/ or apps/ may
// include, link or call this file.  Only tests/, __graft_entry__.smoke() and the
// cpu_baseline / --impl reference legs of bench.py use it (through oracle/oracle.py).
//
// The reference (daviesthomas/cudaNeuralRender) has NO CPU implementation of this path
// (src/simpleInfer.cpp runs nn.forward on the GPU), so this file restates, in plain
// C++ with IEEE-754 single/double arithmetic only, what the reference's CUDA code
// computes.  Every function cites the reference lines it follows (paths relative to
// /root/reference).
//
// Parity status: PINNED against the reference itself.
//   * MLP (oracle_mlp_forward): arithmetic order taken from the vendored CUTLASS 2.0 sources; checked
//     against the known-answer table of SURVEY.md App. B (tests/golden/kat_sdf.json) and BIT-EXACT
//     against nn.forward of the reference's own CUDA kernels run on a B200 (tests/golden/
//     ref_cuda_mlp.npz: 5 models x 20 000 points, made by tools/make_ref_golden.py through
//     oracle/_ref/ref_driver; tests/test_ref_cuda_golden.py).
//   * Renderer (oracle_render): the reference's tests hold no golden vectors, and its device code uses
//     MUFU.RSQ (`rsqrtf`) and libdevice `tanhf`, which are not reproducible on a CPU.  The oracle
//     therefore fixes a *deterministic* arithmetic (1/sqrtf, own tanhf, explicit fmaf where nvcc
//     contracts -- pattern read from the SASS of the reference build, DESIGN.md "Arithmetic contract")
//     that the product CUDA kernels reproduce bit for bit.  Against frames of the real reference
//     (tests/golden/ref_cuda_frames.npz, render_kernel at HEAD) parity is statistical and measured:
//     0 hit-mask mismatches, 97.7-99.98 % of hit pixels within 1/255, 94-98 % identical.
//
// Build: see oracle/Makefile (g++ -O2 -ffp-contract=off -mavx2 -mfma; NO -ffast-math).
// =====================================================================================
#include <cmath>
#include <cstdint>
#include <cstring>
#include <cstdlib>
#include <thread>
#include <vector>
#include <atomic>
#include <algorithm>

namespace {

// ------------------------------------------------------------------ model
struct Model {
    int n_layers = 0;
    std::vector<int> in, out, act;          // act: 0 = ReLU, 1 = "Tanh" == linear epilogue
    std::vector<const float*> W, b;         // W[l] is [out][in] row-major, b[l] is [out]
    int max_width = 0;
};

// params: per layer W[out][in] then b[out]; dims: n_layers+1 ints.
// Activation rule of src/neuralNetwork.cpp:136-139: last layer "Tanh" (linear), others ReLU.
Model make_model(const float* params, const int* dims, int n_layers) {
    Model m;
    m.n_layers = n_layers;
    const float* p = params;
    for (int l = 0; l < n_layers; ++l) {
        int i = dims[l], o = dims[l + 1];
        m.in.push_back(i);
        m.out.push_back(o);
        m.act.push_back(l == n_layers - 1 ? 1 : 0);
        m.W.push_back(p); p += (size_t)i * o;
        m.b.push_back(p); p += o;
        m.max_width = std::max(m.max_width, std::max(i, o));
    }
    return m;
}

// One dense layer on L points held SoA (a[k*L + j]).
// Follows src/layers/denseLayer.cu:126-176 (GemmBatched, N = 1 per point) with the
// arithmetic of CUTLASS 2.0 SIMT:
//   submodules/cutlass/include/cutlass/gemm/thread/mma_sm50.h:136-145  (k outermost, acc carried)
//   submodules/cutlass/include/cutlass/arch/mma_sm50.h:64              (d = a*b + c  -> FFMA)
//   submodules/cutlass/include/cutlass/epilogue/thread/linear_combination_relu.h:171-174
//        x = beta*bias (beta = 1, exact); x = alpha*acc + x (alpha = 1 -> one rounding of acc + bias);
//        ReLU = fmaxf(x, 0)  (functional.h:244-248)
template <int L>
inline void dense_layer(const float* __restrict W, const float* __restrict b, int in, int out, int act,
                        const float* __restrict a, float* __restrict z) {
    for (int o = 0; o < out; ++o) {
        float acc[L];
        for (int j = 0; j < L; ++j) acc[j] = 0.0f;
        const float* w = W + (size_t)o * in;
        for (int k = 0; k < in; ++k) {
            const float wk = w[k];
            const float* ak = a + (size_t)k * L;
            for (int j = 0; j < L; ++j) acc[j] = fmaf(wk, ak[j], acc[j]);
        }
        const float bo = b[o];
        for (int j = 0; j < L; ++j) {
            float v = acc[j] + bo;
            if (act == 0) v = (v > 0.0f) ? v : 0.0f;   // fmaxf(v, 0): NaN -> 0, -0 -> +0
            z[(size_t)o * L + j] = v;
        }
    }
}

// Whole network on L points.  x: SoA [in0][L]; y: SoA [out_last][L].
// Follows src/neuralNetwork.cpp:54-63 (sequential layers).
template <int L>
inline void mlp_block(const Model& m, const float* x, float* y, float* s0, float* s1) {
    const float* a = x;
    float* bufs[2] = {s0, s1};
    for (int l = 0; l < m.n_layers; ++l) {
        float* z = (l == m.n_layers - 1) ? y : bufs[l & 1];
        dense_layer<L>(m.W[l], m.b[l], m.in[l], m.out[l], m.act[l], a, z);
        a = z;
    }
}

// ------------------------------------------------------------------ scalar helpers
struct f3 { float x, y, z; };

// helper_math.h:1253 `a.x*b.x + a.y*b.y + a.z*b.z` as nvcc -fmad=true compiles it
// (SASS of the reference build: FMUL y*y ; FFMA x*x+ ; FFMA z*z+).
inline float dot3(f3 a, f3 b) { return fmaf(a.z, b.z, fmaf(a.x, b.x, a.y * b.y)); }

// helper_math.h:1309-1313 normalize = v * rsqrtf(dot(v,v)).  Device rsqrtf is MUFU.RSQ
// (not CPU-reproducible); the deterministic contract uses 1/sqrtf (the host fallback of
// helper_math.h:62-65).
inline float rsqrt_det(float x) { return 1.0f / sqrtf(x); }
inline f3 normalize3(f3 v) {
    float r = rsqrt_det(dot3(v, v));
    return f3{v.x * r, v.y * r, v.z * r};
}

inline float saturatef(float x) {            // __saturatef: clamp to [0,1], NaN -> 0
    if (!(x > 0.0f)) return 0.0f;
    return x > 1.0f ? 1.0f : x;
}

// Deterministic tanhf (|err| < 1.5 ulp).  The reference calls libdevice tanhf
// (src/volumeRender_kernel.cu:229), whose large-argument branch uses MUFU.EX2/RCP.
// Contract: |x| < 0.55 -> x + x^3 * P(x^2) (degree-4 minimax fit, oracle/tanhfit.py);
// otherwise 1 - 2/(exp(2|x|)+1) with exp by Cody-Waite reduction + degree-7 Taylor.
inline float tanhf_det(float x) {
    const float ax = fabsf(x);
    if (ax < 0.55f) {
        const float x2 = x * x;
        float p = -0x1.9b14e4p-8f;
        p = fmaf(p, x2, 0x1.59370cp-6f);
        p = fmaf(p, x2, -0x1.b9278cp-5f);
        p = fmaf(p, x2, 0x1.110d1ep-3f);
        p = fmaf(p, x2, -0x1.55554ap-2f);
        return fmaf(x * x2, p, x);
    }
    float r;
    if (ax > 9.125f) {
        r = 1.0f;                                   // also catches +-inf; NaN handled below
    } else {
        const float y = ax + ax;
        const float n = rintf(y * 0x1.715476p+0f);   // y * log2(e)
        float t = fmaf(n, -0x1.62e400p-1f, y);       // ln2 hi (exact product for |n| < 2^9)
        t = fmaf(n, -0x1.7f7d1cp-20f, t);            // ln2 lo
        float e = 0x1.a01a02p-13f;                   // 1/5040
        e = fmaf(e, t, 0x1.6c16c2p-10f);             // 1/720
        e = fmaf(e, t, 0x1.111112p-7f);              // 1/120
        e = fmaf(e, t, 0x1.555556p-5f);              // 1/24
        e = fmaf(e, t, 0x1.555556p-3f);              // 1/6
        e = fmaf(e, t, 0.5f);
        e = fmaf(e, t, 1.0f);
        e = fmaf(e, t, 1.0f);
        uint32_t sb = (uint32_t)((int)n + 127) << 23;
        float s; std::memcpy(&s, &sb, 4);
        e = e * s;                                   // exact scaling by 2^n
        r = 1.0f - 2.0f / (e + 1.0f);
    }
    if (ax != ax) return x + x;                      // NaN
    return copysignf(r, x);
}

// ------------------------------------------------------------------ scene functions
// sdfOpSmoothUnion, src/volumeRender_kernel.cu:144-149, with its double promotions and the
// contraction nvcc/ptxas produce (PTX fma.rn.f64 for the mix; SASS DFMA for the last line).
inline float smooth_union(float d1, float d2, float k) {
    const float diff = d2 - d1;
    const double hd = ((double)diff * 0.5) / (double)k + 0.5;
    const float h = saturatef((float)hd);
    const double omh = 1.0 - (double)h;
    const float d1h = d1 * h;
    const float mix = (float)fma(omh, (double)d2, (double)d1h);
    const float kh = k * h;                           // k*h in float, then promoted
    return (float)fma(-omh, (double)kh, (double)mix);
}

// manySphere(p, nSDF, true), src/volumeRender_kernel.cu:176-196 -- the scene function the
// reference's HEAD actually evaluates (sceneSDF :217-222).
inline float many_sphere_union(f3 p, float nsdf, int frame) {
    float s = nsdf;
    float cx = p.x;
    float cy = (float)((double)p.y - 0.6);
    const double zoff = -0.7 + ((double)(frame * 2) * 0.7) / 360.0;
    const float cz = (float)((double)p.z + zoff);
    const float zz = cz * cz;
    for (int i = 0; i < 9; ++i) {
        if (i % 3 == 0) {
            cy = (float)((double)cy + 0.4);
            cx = p.x + 0.5f;                          // float(double(p.x)+0.5) == p.x + 0.5f
        }
        // sdfSphere: length(cP) - 0.1f ; SASS: zz + fma(x,x, y*y) (zz hoisted, plain FADD)
        const float len2 = zz + fmaf(cx, cx, cy * cy);
        const float d = sqrtf(len2) + (-0.1f);
        s = smooth_union(s, d, 0.01f);
        cx = (float)((double)cx - 0.4);
    }
    return s;
}

// ---- the rest of the analytic library, src/volumeRender_kernel.cu:67-174.  Dead code in the reference build
// (sceneSDF returns before reaching it), so there is no SASS to read contraction from: fp32 as written, the source's
// double-promoting literals as double operations, dot products the way nvcc contracts helper_math's dot
// (fma(z,z, fma(x,x, y*y))), min/max with explicit NaN rules, sin by the deterministic form below.
inline float min_det(float a, float b) { return (a != a) ? b : ((b != b) ? a : ((b < a) ? b : a)); }
inline float max_det(float a, float b) { return (a != a) ? b : ((b != b) ? a : ((b > a) ? b : a)); }

// sin (sdfOpDisplace :103-110 calls sinf): Cody-Waite by pi/2 in three parts, Cephes sinf/cosf kernels on |r| <= pi/4
inline float sinf_det(float x) {
    if (!(fabsf(x) < 1.0e6f)) return x - x;
    const float q = rintf(x * 0.636619772f);
    float r = fmaf(q, -1.5703125f, x);
    r = fmaf(q, -4.837512969970703125e-4f, r);
    r = fmaf(q, -7.54978995489188216e-8f, r);
    const int n = (int)q;
    const float z = r * r;
    const float sn = fmaf(fmaf(fmaf(-1.9515295891e-4f, z, 8.3321608736e-3f), z, -1.6666654611e-1f) * z, r, r);
    const float cs = fmaf(fmaf(fmaf(2.443315711809948e-5f, z, -1.388731625493765e-3f), z, 4.166664568298827e-2f) * z, z,
                          fmaf(-0.5f, z, 1.0f));
    const float v = (n & 1) ? cs : sn;
    return (n & 2) ? -v : v;
}

// sdfOpSmoothSubtraction :138-142: h = saturate(0.5 - 0.5*(d1+d2)/k); mix = d1*(1.0-h) - d2*h; mix + k*h*(1.0-h)
inline float smooth_subtract(float d1, float d2, float k) {
    const float sum = d1 + d2;
    const double hd = 0.5 - ((double)sum * 0.5) / (double)k;
    const float h = saturatef((float)hd);
    const double omh = 1.0 - (double)h;
    const float d2h = d2 * h;
    const float mix = (float)fma(omh, (double)d1, -(double)d2h);
    const float kh = k * h;
    return (float)fma(omh, (double)kh, (double)mix);
}

inline float length3(f3 v) { return sqrtf(dot3(v, v)); }

// sdfBox :81-89
inline float sdf_box(f3 p, f3 b, float r) {
    const f3 q = f3{fabsf(p.x) - b.x, fabsf(p.y) - b.y, fabsf(p.z) - b.z};
    const f3 m = f3{max_det(q.x, 0.0f), max_det(q.y, 0.0f), max_det(q.z, 0.0f)};
    const float inner = max_det(q.x, max_det(q.y, q.z));
    const double in0 = (double)inner < 0.0 ? (double)inner : 0.0;           // min(float, 0.0) is a double min
    return (float)(((double)length3(m) + in0) - (double)r);
}

// sdfCylinder :96-101
inline float sdf_cylinder(f3 p, f3 c) {
    const float dx = p.x - c.x, dy = p.y - c.z;
    return sqrtf(fmaf(dx, dx, dy * dy)) - c.y;
}

// manySphere(p, nSDF, false) :176-196 (the subtraction form)
inline float many_sphere_subtract(f3 p, float nsdf, int frame) {
    float s = nsdf;
    float cx = p.x;
    float cy = (float)((double)p.y - 0.6);
    const double zoff = -0.7 + ((double)(frame * 2) * 0.7) / 360.0;
    const float cz = (float)((double)p.z + zoff);
    const float zz = cz * cz;
    for (int i = 0; i < 9; ++i) {
        if (i % 3 == 0) {
            cy = (float)((double)cy + 0.4);
            cx = p.x + 0.5f;
        }
        const float len2 = zz + fmaf(cx, cx, cy * cy);
        const float d = sqrtf(len2) + (-0.1f);
        s = smooth_subtract(s, d, 0.01f);
        cx = (float)((double)cx - 0.4);
    }
    return s;
}

// manyCylinderCut :156-174
inline float many_cylinder_cut(f3 p, float nsdf) {
    float s = nsdf;
    const f3 c = f3{0.02f, 0.02f, 0.02f};
    f3 cp = p;
    cp.y = (float)((double)cp.y - 0.5);
    for (int i = 0; i < 300; ++i) {
        if (i % 20 == 0) {
            cp.y = (float)((double)cp.y + 0.1);
            cp.x = (float)((double)p.x + 0.9);
        }
        // sdfCylinder (:96-101) as nvcc contracts it INSIDE this loop: dy is loop-invariant per row, and the compiled reference
        // (oracle/_ref/ref_sdf_probe SASS: FMUL dx,dx ; FFMA dy,dy,.) forms fma(dy, dy, dx*dx) here, not the stand-alone
        // function's fma(dx, dx, dy*dy)
        const float dx = cp.x - c.x, dy = cp.y - c.z;
        const float cyl = sqrtf(fmaf(dy, dy, dx * dx)) - c.y;
        s = smooth_subtract(s, cyl, 0.01f);
        cp.x = (float)((double)cp.x - 0.1);
    }
    return s;
}

// One accumulator chain over the library: the run-time form of editing sceneSDF (:217-230).  Layout = nsdf_scene_op.
struct SceneOp { int prim; int combine; float a[8]; };
static std::vector<SceneOp> g_program;   // set by oracle_set_scene_program; scene == 2 evaluates it

inline float scene_program_eval(const SceneOp* ops, int n, f3 p, float nsdf, int frame) {
    float s = 0.0f;
    for (int i = 0; i < n; ++i) {
        const SceneOp& op = ops[i];
        const float* a = op.a;
        if (op.prim >= 16) {
            switch (op.prim) {
                case 16: {   // sdfOpDisplace :103-110 (frequency 5 and amplitude 0.05 there; float parameters here)
                    const float w = (sinf_det(a[0] * p.x) * sinf_det(a[0] * p.y)) * sinf_det(a[0] * p.z);
                    // :107 multiplies by the DOUBLE literal 0.05; a float parameter equal to 0.05f stands for that literal
                    const double amp = a[1] == 0.05f ? 0.05 : (double)a[1];
                    s = (float)fma((double)w, amp, (double)s);
                    break;
                }
                case 17: s = s - a[0]; break;                              // sdfOpRound :112-115
                case 18: s = fabsf(s) - a[0]; break;                       // sdfOpOnion :117-120
                case 19: s = (a[0] != 0.0f) ? many_sphere_union(p, s, frame) : many_sphere_subtract(p, s, frame); break;
                case 20: s = many_cylinder_cut(p, s); break;
                default: break;
            }
            continue;
        }
        float d;
        const f3 q = f3{p.x - a[0], p.y - a[1], p.z - a[2]};
        switch (op.prim) {
            case 0: d = nsdf; break;
            case 1: d = tanhf_det(nsdf); break;                            // :229
            case 2: d = length3(q) - a[3]; break;                          // sdfSphere :67-71
            case 3: d = sdf_box(q, f3{a[3], a[4], a[5]}, a[6]); break;
            case 4: d = p.y - a[0]; break;                                 // sdfPlane :91-94 (0.5 there)
            case 5: d = sdf_cylinder(q, f3{a[3], a[4], a[5]}); break;
            default: d = nsdf; break;
        }
        switch (op.combine) {
            case 1: s = min_det(s, d); break;                              // sdfOpUnion :128-131
            case 2: s = max_det(s, d); break;                              // sdfOpIntersect :123-126
            case 3: s = max_det(s, -d); break;                             // sdfOpSubtraction :133-136
            case 4: s = max_det(d, -s); break;
            case 5: s = smooth_union(s, d, a[7]); break;
            case 6: s = smooth_subtract(s, d, a[7]); break;
            case 7: s = smooth_subtract(d, s, a[7]); break;
            default: s = d; break;
        }
    }
    return s;
}

enum Scene { SCENE_TANH = 0, SCENE_MANYSPHERE = 1, SCENE_PROGRAM = 2 };

// sceneSDF, src/volumeRender_kernel.cu:217-230.  scene 1 = HEAD (manySphere union);
// scene 0 = the unreachable-but-intended `tanh(nSDF)` (:229) used for the shipped renders;
// scene 2 = the op list last given to oracle_set_scene_program.
inline float scene_sdf(int scene, f3 p, float nsdf, int frame) {
    if (scene == SCENE_MANYSPHERE) return many_sphere_union(p, nsdf, frame);
    if (scene == SCENE_PROGRAM) return scene_program_eval(g_program.data(), (int)g_program.size(), p, nsdf, frame);
    return tanhf_det(nsdf);
}

const float TET[4][3] = {{1, -1, -1}, {-1, -1, 1}, {-1, 1, -1}, {1, 1, 1}};   // :38-43
const float NORMAL_EPSILON = 0.00001f;     // :59
const float MARCHING_EPSILON = 0.000001f;  // :60

struct Camera {
    float inv_view[12];   // c_invViewMatrix, 3 rows x 4   (:31)
    float normal[16];     // c_normalMatrix, 4 rows x 4    (:32)
    int frame;
};

struct Ray { f3 o, d, p; float tfar; bool hit; };

// initMarcher, src/volumeRender_kernel.cu:293-358 (+ intersectSphere :200-215).
inline Ray init_ray(const Camera& c, int x, int y, int W, int H) {
    Ray r;
    const float u = fmaf((float)x / (float)W, 2.0f, -1.0f);
    const float v = fmaf((float)y / (float)H, 2.0f, -1.0f);
    const float* M = c.inv_view;
    // o = M * (0,0,0,1): 0*m.y ; fma(0,m.x) ; fma(0,m.z) ; + m.w
    r.o = f3{M[3] + fmaf(M[2], 0.0f, fmaf(M[0], 0.0f, M[1] * 0.0f)),
             M[7] + fmaf(M[6], 0.0f, fmaf(M[4], 0.0f, M[5] * 0.0f)),
             M[11] + fmaf(M[10], 0.0f, fmaf(M[8], 0.0f, M[9] * 0.0f))};
    // normalize(u, v, -2): dot = fma(u,u, v*v) + 4 ; components v*rs, with z = rs * -2
    const float rs = rsqrt_det(fmaf(u, u, v * v) + 4.0f);
    const f3 d0{u * rs, v * rs, rs * -2.0f};
    r.d = f3{dot3(d0, f3{M[0], M[1], M[2]}), dot3(d0, f3{M[4], M[5], M[6]}), dot3(d0, f3{M[8], M[9], M[10]})};
    // intersectSphere with c = 0, r = 1.2f
    const float a = dot3(r.d, r.d);
    const float qd = dot3(r.o, r.d);
    const float b = qd + qd;                                   // 2.0 * dot, exact
    const float cc = dot3(r.o, r.o) + (-1.4400000572204589844f);   // r*r folded by the compiler
    const float disc = fmaf(b, b, -(cc * (a * 4.0f)));
    r.hit = disc > 0.0f;
    r.tfar = 0.0f;
    r.p = r.o;
    if (!r.hit) return r;
    const float sq = sqrtf(disc);
    const double two_a = (double)a + (double)a;
    float tnear = (float)((double)(-b - sq) / two_a);
    r.tfar = (float)((double)(sq - b) / two_a);
    if (tnear < 0.0f) tnear = 0.0f;
    r.p = f3{fmaf(r.d.x, tnear, r.o.x), fmaf(r.d.y, tnear, r.o.y), fmaf(r.d.z, tnear, r.o.z)};
    return r;
}

inline uint32_t rgba_float_to_int(float r, float g, float b, float a) {   // :266-274
    r = saturatef(r); g = saturatef(g); b = saturatef(b); a = saturatef(a);
    return ((uint32_t)(a * 255.0f) << 24) | ((uint32_t)(b * 255.0f) << 16) | ((uint32_t)(g * 255.0f) << 8) |
           (uint32_t)(r * 255.0f);
}

// facingColor, :380-384.  SASS: t = rx*nx ; t = fma(-ry, ny, -t) ; t = fma(-rz, nz, t)
inline uint32_t facing_color(f3 n, f3 ray) {
    float t = ray.x * n.x;
    t = fmaf(-ray.y, n.y, -t);
    t = fmaf(-ray.z, n.z, t);
    double rd = (double)t;
    rd = (rd > 0.0) ? rd : 0.0;                                // max(0.0, x), NaN -> 0
    const float ratio = (float)rd;
    return rgba_float_to_int(ratio, ratio, ratio, 1.0f);
}

// matCapColor, :387-413.
inline uint32_t matcap_color(const Camera& c, f3 n, const uint32_t* matcap, int mw, int mh) {
    const float* N = c.normal;
    float e[3];
    for (int i = 0; i < 3; ++i) {
        const float* row = N + 4 * i;
        float t = n.y * row[1];
        t = fmaf(n.x, row[0], t);
        t = fmaf(n.z, row[2], t);
        t = fmaf(row[3], 0.0f, t);                             // w = 0
        e[i] = t;
    }
    const f3 ne = normalize3(f3{e[0], e[1], e[2]});
    const float fu = (float)fma((double)ne.x, 0.5, 0.5);
    const float fv = (float)fma((double)ne.y, 0.5, 0.5);
    const int ux = (int)(fu * (float)(mw - 1));
    const int uy = (int)(fv * (float)(mh - 1));
    const int index = uy * mw + ux;
    if (index < 0) return 0u;                                  // rgbaFloatToInt(0)
    return matcap[index];
}

// ------------------------------------------------------------------ renderer
// One block of L pixels marched in lock step.  Per pixel this is exactly the per-pixel
// projection of render_kernel's wavefront loop (src/volumeRender_kernel.cu:608-692):
// iteration i runs nn.forward on the requests formed after iteration i-1 and then
// singleMarch (:416-477).  mask 1 = step request, mask 4 = shade request (4 tetrahedron
// points, createBatch :536-545), mask 0 = retired.  A shade consumes one iteration; a ray
// whose converging step is iteration MAX_STEPS-1 is never shaded and keeps the caller's
// memset 0.  E-6 (stale SDF slot of the very last pixel) is waived.
// Here a shade request is served by four consecutive single-point passes (tet 0..3); that
// changes nothing observable because pixels never interact.
template <int L>
void render_block(const Model& m, const Camera& cam, const uint32_t* matcap, int mw, int mh, int color_type,
                  int scene, int W, int H, int max_steps, const int* px, const int* py, int n, uint32_t* out,
                  int32_t* steps_out, uint8_t* hit_out, int64_t& queries, float* scratch) {
    const int in0 = m.in[0];                    // 3, or 4 for (x, y, z, t) animation nets
    float* xin = scratch;                       // [in0][L]
    float* yout = xin + in0 * L;                // [L]
    float* s0 = yout + L;                       // [max_width][L]
    float* s1 = s0 + (size_t)m.max_width * L;
    Ray ray[L];
    int mask[L], iter[L], tet[L];
    float sv[L][4];
    int active = 0;
    for (int j = 0; j < L; ++j) {
        mask[j] = 0; iter[j] = 0; tet[j] = 0;
        if (j >= n) continue;
        ray[j] = init_ray(cam, px[j], py[j], W, H);
        const size_t id = (size_t)py[j] * W + px[j];
        out[id] = 0u;                                          // caller's memset == BACKGROUND_COLOR
        if (steps_out) steps_out[id] = 0;
        if (hit_out) hit_out[id] = 0;
        if (ray[j].hit && max_steps > 0) { mask[j] = 1; ++active; }
    }
    while (active > 0) {
        // createBatch (:504-547): one request point per lane per pass
        for (int j = 0; j < L; ++j) {
            f3 q{0, 0, 0};
            if (mask[j] == 1) q = ray[j].p;
            else if (mask[j] == 4) {
                const float* t = TET[tet[j]];
                q = f3{fmaf(t[0], NORMAL_EPSILON, ray[j].p.x), fmaf(t[1], NORMAL_EPSILON, ray[j].p.y),
                       fmaf(t[2], NORMAL_EPSILON, ray[j].p.z)};
            }
            xin[0 * L + j] = q.x; xin[1 * L + j] = q.y; xin[2 * L + j] = q.z;
            // createBatch :532-545: the fourth input of every request is c_frameNumber.  (For the four normal requests the
            // reference strides its writes by 3 instead of 4 and scrambles them; this is the intended layout)
            if (in0 == 4) xin[3 * L + j] = (float)cam.frame;
        }
        mlp_block<L>(m, xin, yout, s0, s1);
        for (int j = 0; j < L; ++j) {
            if (mask[j] == 0) continue;
            ++queries;
            const size_t id = (size_t)py[j] * W + px[j];
            if (mask[j] == 4) {
                // surfaceNormal (:361-377): sceneSDF(p + tet_i*eps, sdf_i)
                const f3 q{xin[0 * L + j], xin[1 * L + j], xin[2 * L + j]};
                sv[j][tet[j]] = scene_sdf(scene, q, yout[j], cam.frame);
                if (++tet[j] < 4) continue;
                // p0+p1+p2+p3 as compiled: fma(t3,s3, fma(t2,s2, fma(t0,s0, t1*s1)))
                const float* s = sv[j];
                f3 nn;
                nn.x = fmaf(TET[3][0], s[3], fmaf(TET[2][0], s[2], fmaf(TET[0][0], s[0], TET[1][0] * s[1])));
                nn.y = fmaf(TET[3][1], s[3], fmaf(TET[2][1], s[2], fmaf(TET[0][1], s[0], TET[1][1] * s[1])));
                nn.z = fmaf(TET[3][2], s[3], fmaf(TET[2][2], s[2], fmaf(TET[0][2], s[0], TET[1][2] * s[1])));
                const f3 nrm = normalize3(nn);
                out[id] = (color_type == 0) ? facing_color(nrm, ray[j].d) : matcap_color(cam, nrm, matcap, mw, mh);
                if (hit_out) hit_out[id] = 1;
                mask[j] = 0; --active;
                continue;
            }
            // march step (:459-476)
            const float t = scene_sdf(scene, ray[j].p, yout[j], cam.frame);
            ray[j].tfar = ray[j].tfar - t;
            ++iter[j];
            if (steps_out) steps_out[id] = iter[j];
            if (ray[j].tfar <= 0.0f) {                          // NaN keeps marching (PTX: setp.gtu)
                mask[j] = 0; --active; out[id] = 0u;
                continue;
            }
            ray[j].p = f3{fmaf(ray[j].d.x, t, ray[j].p.x), fmaf(ray[j].d.y, t, ray[j].p.y),
                          fmaf(ray[j].d.z, t, ray[j].p.z)};
            if (t < MARCHING_EPSILON) { mask[j] = 4; tet[j] = 0; }
            if (iter[j] >= max_steps) { mask[j] = 0; --active; }   // loop bound, render_kernel :652
        }
    }
}

constexpr int LANES = 16;

template <class F>
void parallel_for(int64_t n_items, int nthreads, F fn) {
    if (nthreads <= 0) nthreads = (int)std::thread::hardware_concurrency();
    if (nthreads < 1) nthreads = 1;
    if (nthreads == 1 || n_items <= 1) { fn(0, n_items, 0); return; }
    std::atomic<int64_t> next(0);
    const int64_t chunk = std::max<int64_t>(1, n_items / ((int64_t)nthreads * 16));
    std::vector<std::thread> th;
    for (int t = 0; t < nthreads; ++t)
        th.emplace_back([&, t] {
            for (;;) {
                int64_t b = next.fetch_add(chunk);
                if (b >= n_items) break;
                fn(b, std::min(n_items, b + chunk), t);
            }
        });
    for (auto& x : th) x.join();
}

}  // namespace

extern "C" {

int oracle_hardware_threads() { return (int)std::thread::hardware_concurrency(); }

float oracle_tanhf(float x) { return tanhf_det(x); }

void oracle_tanhf_array(const float* x, float* y, int64_t n) {
    for (int64_t i = 0; i < n; ++i) y[i] = tanhf_det(x[i]);
}

float oracle_scene_sdf(int scene, float x, float y, float z, float nsdf, int frame) {
    return scene_sdf(scene, f3{x, y, z}, nsdf, frame);
}

// ops: n records of {int prim, int combine, float a[8]} (include/nsdf.h, nsdf_scene_op).  Not thread-safe against a
// render in flight; tests set it between calls.
void oracle_set_scene_program(const void* ops, int n) {
    g_program.clear();
    if (ops && n > 0) g_program.assign((const SceneOp*)ops, (const SceneOp*)ops + n);
}

void oracle_scene_program_eval(const float* xyz, const float* nsdf, int frame, float* out, int64_t n) {
    for (int64_t i = 0; i < n; ++i)
        out[i] = scene_program_eval(g_program.data(), (int)g_program.size(), f3{xyz[3 * i], xyz[3 * i + 1], xyz[3 * i + 2]},
                                    nsdf[i], frame);
}

float oracle_sinf(float x) { return sinf_det(x); }

// NeuralNetwork::forward (src/neuralNetwork.cpp:54-63) on n points.
// x: [n][in0] (the reference's Matrix(Shape(3, B)) layout, element b*3+c); y: [n][out_last].
int oracle_mlp_forward(const float* params, const int* dims, int n_layers, const float* x, int64_t n, float* y,
                       int nthreads) {
    if (n_layers <= 0 || n < 0) return 1;
    const Model m = make_model(params, dims, n_layers);
    const int in0 = m.in[0], outl = m.out[n_layers - 1];
    const int64_t nblocks = (n + LANES - 1) / LANES;
    parallel_for(nblocks, nthreads, [&](int64_t b0, int64_t b1, int) {
        std::vector<float> xin((size_t)in0 * LANES), yo((size_t)outl * LANES), s0((size_t)m.max_width * LANES),
            s1((size_t)m.max_width * LANES);
        for (int64_t blk = b0; blk < b1; ++blk) {
            const int64_t base = blk * LANES;
            const int cnt = (int)std::min<int64_t>(LANES, n - base);
            for (int j = 0; j < LANES; ++j)
                for (int c = 0; c < in0; ++c) xin[(size_t)c * LANES + j] = (j < cnt) ? x[(base + j) * in0 + c] : 0.0f;
            mlp_block<LANES>(m, xin.data(), yo.data(), s0.data(), s1.data());
            for (int j = 0; j < cnt; ++j)
                for (int c = 0; c < outl; ++c) y[(base + j) * outl + c] = yo[(size_t)c * LANES + j];
        }
    });
    return 0;
}

// updateViewMatrices (src/main.cpp:207-222) without Eigen: quaternion(angle-axis X) *
// quaternion(angle-axis Y) -> rotation matrix m; modelView = [m | m*v], v = -viewTranslation;
// inv_view = rows 0..2; normal = inverse(modelView) (general cofactor inverse, float).
// Eigen (un-vendored, absent here) may differ from this in the last ulp for rx,ry != 0; for the
// default camera (rx = ry = 0) every entry is exact.
void oracle_view_matrices(float rx_deg, float ry_deg, float tx, float ty, float tz, float* inv_view12,
                          float* normal16) {
    const float ax = (float)(-(double)rx_deg * M_PI / 180);
    const float ay = (float)(-(double)ry_deg * M_PI / 180);
    const float hx = 0.5f * ax, hy = 0.5f * ay;
    const float cx = cosf(hx), sx = sinf(hx), cy = cosf(hy), sy = sinf(hy);
    // (cx, sx,0,0) * (cy, 0,sy,0)
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
    float M[4][4] = {{0}};
    for (int i = 0; i < 3; ++i) {
        for (int j = 0; j < 3; ++j) M[i][j] = m[i][j];
        M[i][3] = m[i][0] * v[0] + m[i][1] * v[1] + m[i][2] * v[2];
    }
    M[3][3] = 1.0f;
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 4; ++j) inv_view12[4 * i + j] = M[i][j];
    // general 4x4 inverse by cofactors
    const float* a = &M[0][0];
    float inv[16];
    inv[0] = a[5] * a[10] * a[15] - a[5] * a[11] * a[14] - a[9] * a[6] * a[15] + a[9] * a[7] * a[14] + a[13] * a[6] * a[11] - a[13] * a[7] * a[10];
    inv[4] = -a[4] * a[10] * a[15] + a[4] * a[11] * a[14] + a[8] * a[6] * a[15] - a[8] * a[7] * a[14] - a[12] * a[6] * a[11] + a[12] * a[7] * a[10];
    inv[8] = a[4] * a[9] * a[15] - a[4] * a[11] * a[13] - a[8] * a[5] * a[15] + a[8] * a[7] * a[13] + a[12] * a[5] * a[11] - a[12] * a[7] * a[9];
    inv[12] = -a[4] * a[9] * a[14] + a[4] * a[10] * a[13] + a[8] * a[5] * a[14] - a[8] * a[6] * a[13] - a[12] * a[5] * a[10] + a[12] * a[6] * a[9];
    inv[1] = -a[1] * a[10] * a[15] + a[1] * a[11] * a[14] + a[9] * a[2] * a[15] - a[9] * a[3] * a[14] - a[13] * a[2] * a[11] + a[13] * a[3] * a[10];
    inv[5] = a[0] * a[10] * a[15] - a[0] * a[11] * a[14] - a[8] * a[2] * a[15] + a[8] * a[3] * a[14] + a[12] * a[2] * a[11] - a[12] * a[3] * a[10];
    inv[9] = -a[0] * a[9] * a[15] + a[0] * a[11] * a[13] + a[8] * a[1] * a[15] - a[8] * a[3] * a[13] - a[12] * a[1] * a[11] + a[12] * a[3] * a[9];
    inv[13] = a[0] * a[9] * a[14] - a[0] * a[10] * a[13] - a[8] * a[1] * a[14] + a[8] * a[2] * a[13] + a[12] * a[1] * a[10] - a[12] * a[2] * a[9];
    inv[2] = a[1] * a[6] * a[15] - a[1] * a[7] * a[14] - a[5] * a[2] * a[15] + a[5] * a[3] * a[14] + a[13] * a[2] * a[7] - a[13] * a[3] * a[6];
    inv[6] = -a[0] * a[6] * a[15] + a[0] * a[7] * a[14] + a[4] * a[2] * a[15] - a[4] * a[3] * a[14] - a[12] * a[2] * a[7] + a[12] * a[3] * a[6];
    inv[10] = a[0] * a[5] * a[15] - a[0] * a[7] * a[13] - a[4] * a[1] * a[15] + a[4] * a[3] * a[13] + a[12] * a[1] * a[7] - a[12] * a[3] * a[5];
    inv[14] = -a[0] * a[5] * a[14] + a[0] * a[6] * a[13] + a[4] * a[1] * a[14] - a[4] * a[2] * a[13] - a[12] * a[1] * a[6] + a[12] * a[2] * a[5];
    inv[3] = -a[1] * a[6] * a[11] + a[1] * a[7] * a[10] + a[5] * a[2] * a[11] - a[5] * a[3] * a[10] - a[9] * a[2] * a[7] + a[9] * a[3] * a[6];
    inv[7] = a[0] * a[6] * a[11] - a[0] * a[7] * a[10] - a[4] * a[2] * a[11] + a[4] * a[3] * a[10] + a[8] * a[2] * a[7] - a[8] * a[3] * a[6];
    inv[11] = -a[0] * a[5] * a[11] + a[0] * a[7] * a[9] + a[4] * a[1] * a[11] - a[4] * a[3] * a[9] - a[8] * a[1] * a[7] + a[8] * a[3] * a[5];
    inv[15] = a[0] * a[5] * a[10] - a[0] * a[6] * a[9] - a[4] * a[1] * a[10] + a[4] * a[2] * a[9] + a[8] * a[1] * a[6] - a[8] * a[2] * a[5];
    const float det = a[0] * inv[0] + a[1] * inv[4] + a[2] * inv[8] + a[3] * inv[12];
    const float idet = 1.0f / det;
    for (int i = 0; i < 16; ++i) normal16[i] = inv[i] * idet;
}

// initMarcher only: per pixel [ox,oy,oz, dx,dy,dz, px,py,pz, tfar] and hit flag (ray-level tests).
int oracle_init_rays(const float* inv_view12, int W, int H, float* rays10, uint8_t* hit) {
    Camera cam{};
    std::memcpy(cam.inv_view, inv_view12, sizeof(cam.inv_view));
    for (int y = 0; y < H; ++y)
        for (int x = 0; x < W; ++x) {
            const Ray r = init_ray(cam, x, y, W, H);
            float* o = rays10 + ((size_t)y * W + x) * 10;
            o[0] = r.o.x; o[1] = r.o.y; o[2] = r.o.z; o[3] = r.d.x; o[4] = r.d.y; o[5] = r.d.z;
            o[6] = r.p.x; o[7] = r.p.y; o[8] = r.p.z; o[9] = r.tfar;
            hit[(size_t)y * W + x] = r.hit ? 1 : 0;
        }
    return 0;
}

// render_kernel (src/volumeRender_kernel.cu:608-692) for one frame, or for the rows
// y = y0, y0+ystride, ... < y1 of it (bounded CPU-baseline samples; other rows are not touched).
//   color_type: 0 facing ratio, 1 matcap (c_coloringType :33);  scene: 0 tanh, 1 manySphere (HEAD)
//   out: W*H packed ABGR uint32 (:273);  steps_out/hit_out optional;  queries_out = Q_frame.
int oracle_render_rows(const float* params, const int* dims, int n_layers, const float* inv_view12,
                       const float* normal16, int frame, const uint32_t* matcap, int mw, int mh, int color_type,
                       int scene, int W, int H, int max_steps, int y0, int y1, int ystride, uint32_t* out,
                       int32_t* steps_out, uint8_t* hit_out, int64_t* queries_out, int nthreads) {
    if (n_layers <= 0 || W <= 0 || H <= 0 || ystride <= 0 || y0 < 0 || y1 > H) return 1;
    if (color_type != 0 && (!matcap || mw <= 0 || mh <= 0)) return 2;
    const Model m = make_model(params, dims, n_layers);
    if ((m.in[0] != 3 && m.in[0] != 4) || m.out[n_layers - 1] != 1) return 3;
    Camera cam{};
    std::memcpy(cam.inv_view, inv_view12, sizeof(cam.inv_view));
    std::memcpy(cam.normal, normal16, sizeof(cam.normal));
    cam.frame = frame;
    const int64_t nrows = y1 > y0 ? (y1 - y0 + ystride - 1) / ystride : 0;
    const int64_t N = nrows * W;
    const int64_t nblocks = (N + LANES - 1) / LANES;
    std::atomic<int64_t> total_q(0);
    parallel_for(nblocks, nthreads, [&](int64_t b0, int64_t b1, int) {
        std::vector<float> scratch((size_t)(5 + 2 * m.max_width) * LANES);
        int64_t q = 0;
        int px[LANES], py[LANES];
        for (int64_t blk = b0; blk < b1; ++blk) {
            const int64_t base = blk * LANES;
            const int cnt = (int)std::min<int64_t>(LANES, N - base);
            for (int j = 0; j < cnt; ++j) {
                px[j] = (int)((base + j) % W);
                py[j] = y0 + (int)((base + j) / W) * ystride;
            }
            render_block<LANES>(m, cam, matcap, mw, mh, color_type, scene, W, H, max_steps, px, py, cnt, out,
                                steps_out, hit_out, q, scratch.data());
        }
        total_q += q;
    });
    if (queries_out) *queries_out = total_q.load();
    return 0;
}

int oracle_render(const float* params, const int* dims, int n_layers, const float* inv_view12, const float* normal16,
                  int frame, const uint32_t* matcap, int mw, int mh, int color_type, int scene, int W, int H,
                  int max_steps, uint32_t* out, int32_t* steps_out, uint8_t* hit_out, int64_t* queries_out,
                  int nthreads) {
    return oracle_render_rows(params, dims, n_layers, inv_view12, normal16, frame, matcap, mw, mh, color_type, scene, W,
                              H, max_steps, 0, H, 1, out, steps_out, hit_out, queries_out, nthreads);
}

}  // extern "C"
