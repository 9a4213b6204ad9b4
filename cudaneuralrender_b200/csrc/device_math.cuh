// =====================================================================================
// device_math.cuh -- scalar arithmetic of the neural-SDF renderer, sm_100a device side.
//
// Every function states the reference lines it replaces (paths relative to the
// reference repo root).  The translation unit is compiled with -fmad=false, so the ONLY
// fused multiply-adds are the explicit fmaf()/fma() below; their placement is the
// "arithmetic contract" of DESIGN.md (the contraction pattern nvcc -fmad=true produces
// for the reference's expressions), which makes every result reproducible bit for bit.
//
// Two math modes (nsdf_render_opts.math), selected by the warp-uniform `ref` argument:
//   ref == true   NSDF_MATH_REFERENCE: the three operations the reference's device code takes from
//                 the SFU / libdevice are taken from the same place -- rsqrtf (rsqrt.approx.f32 =
//                 MUFU.RSQ with the denormal pre-scale, what helper_math.h:1309-1313 compiles to),
//                 libdevice tanhf (:229) and __saturatef (:140,146,268-271).  On the GPU family the
//                 reference binary runs on, frames equal the reference's frames.
//   ref == false  NSDF_MATH_DETERMINISTIC: IEEE-only forms of the same three, which a CPU can
//                 reproduce bit for bit -- the mode the oracle tests run in.
// libdevice's tanhf / sinf / rsqrtf compile to the same SASS under -fmad=false and -fmad=true
// (checked with cuobjdump), so calling them from this translation unit is exact.
// =====================================================================================
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace nsdf {

struct f3 { float x, y, z; };

// helper_math.h:1248-1253 dot(float3,float3): y*y first, then fma x, then fma z.
__device__ __forceinline__ float dot3(f3 a, f3 b) { return fmaf(a.z, b.z, fmaf(a.x, b.x, a.y * b.y)); }

// helper_math.h:1309-1313 normalize = v * rsqrtf(dot(v,v)).  MUFU.RSQ is replaced by the
// correctly rounded 1/sqrt (the header's own host fallback, helper_math.h:62-65).
__device__ __forceinline__ float rsqrt_det(float x) { return __fdiv_rn(1.0f, __fsqrt_rn(x)); }
__device__ __forceinline__ float rsqrt_sel(float x, bool ref) { return ref ? rsqrtf(x) : rsqrt_det(x); }
__device__ __forceinline__ f3 normalize3(f3 v, bool ref) {
    const float r = rsqrt_sel(dot3(v, v), ref);
    return f3{v.x * r, v.y * r, v.z * r};
}

// __saturatef with NaN -> 0 (volumeRender_kernel.cu:140,146,268-271)
__device__ __forceinline__ float saturate_det(float x) {
    if (!(x > 0.0f)) return 0.0f;
    return x > 1.0f ? 1.0f : x;
}
__device__ __forceinline__ float saturate_sel(float x, bool ref) { return ref ? __saturatef(x) : saturate_det(x); }

// tanh of the network output (volumeRender_kernel.cu:229; simpleInfer.cpp:109).  libdevice
// tanhf goes through MUFU.EX2/RCP; this polynomial/Cody-Waite form is < 1.5 ulp and uses only
// IEEE operations.
__device__ __forceinline__ float tanhf_det(float x) {
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
        r = 1.0f;
    } else {
        const float y = ax + ax;
        const float n = rintf(y * 0x1.715476p+0f);
        float t = fmaf(n, -0x1.62e400p-1f, y);
        t = fmaf(n, -0x1.7f7d1cp-20f, t);
        float e = 0x1.a01a02p-13f;
        e = fmaf(e, t, 0x1.6c16c2p-10f);
        e = fmaf(e, t, 0x1.111112p-7f);
        e = fmaf(e, t, 0x1.555556p-5f);
        e = fmaf(e, t, 0x1.555556p-3f);
        e = fmaf(e, t, 0.5f);
        e = fmaf(e, t, 1.0f);
        e = fmaf(e, t, 1.0f);
        const float s = __uint_as_float((uint32_t)((int)n + 127) << 23);
        e = e * s;
        r = 1.0f - __fdiv_rn(2.0f, e + 1.0f);
    }
    if (ax != ax) return x + x;
    return copysignf(r, x);
}
__device__ __forceinline__ float tanh_sel(float x, bool ref) { return ref ? tanhf(x) : tanhf_det(x); }

// sdfOpSmoothUnion (volumeRender_kernel.cu:144-149) with its double promotions.
// Outside the blend band the double arithmetic collapses exactly: d2 - d1 >= k gives h == 1 and the value d1, d2 - d1 <= -k
// gives h == 0 and the value d2 (every product with the zero weight is a signed zero that the final roundings absorb, which
// is why zero operands and non-finite differences take the long way).  Same bits, no division on the march's critical path.
__device__ __forceinline__ float smooth_union(float d1, float d2, float k, bool ref) {
    const float diff = d2 - d1;
    if (diff >= k && diff <= 1.0e30f && d1 != 0.0f) return d1;
    if (diff <= -k && diff >= -1.0e30f && d2 != 0.0f) return d2;
    const double hd = __dadd_rn(__ddiv_rn(__dmul_rn((double)diff, 0.5), (double)k), 0.5);
    const float h = saturate_sel((float)hd, ref);
    const double omh = __dsub_rn(1.0, (double)h);
    const float d1h = d1 * h;
    const float mix = (float)fma(omh, (double)d2, (double)d1h);
    const float kh = k * h;
    return (float)fma(-omh, (double)kh, (double)mix);
}

// manySphere(p, nSDF, true) (volumeRender_kernel.cu:176-196): what sceneSDF (:217-222) evaluates at the reference's HEAD.
//
// The nine sphere centres depend on the point alone, and almost every query is nowhere near a sphere: then every one of the
// nine smooth unions takes smooth_union's d2 - d1 >= k exit and the function is the identity on nSDF.  That case is decided
// WITHOUT the nine square roots and without the chain: with m = min_i |p - c_i|^2 (computed exactly as the chain computes
// each len2), sqrt(len2_i) >= nSDF + 0.11001 for every i  =>  d_i = fl(fl(sqrt) - 0.1f) >= nSDF + 0.0100097 > nSDF + k
// =>  fl(d_i - s) >= k (rounding is monotonic; the 1e-5 in the constant and the 1e-6 in the factor are far above the
// roundings involved), so m >= (nSDF + 0.11001)^2 proves it.  m is the "pre" value: a marching warp takes it before the
// network answers, off the dependent chain (~160 clocks of double adds instead of ~2000 clocks of square roots with their
// slow-path branches on every march step: measured 2.35 us per dependent step against 1.35 us for the tanh scene).
// zoff = -0.7 + (c_frameNumber*2*0.7/360) (:182) depends on the frame number alone: many_sphere_zoff computes it once per
// call on the host in the same IEEE double operations (a device-side double division of 0 by 360 -- frame 0 -- takes the
// slow path of the division routine on every step of every ray).
__host__ __device__ inline double many_sphere_zoff(int frame) {
    volatile double num = (double)(frame * 2) * 0.7;
    volatile double q = num / 360.0;
    return -0.7 + q;
}
// The reference walks the 3 x 3 grid with a running cx / cy (:183-195): cx restarts at p.x + 0.5f in every row and loses 0.4 (a
// double literal: the subtraction is done in double and rounded back to float) per sphere, cy gains 0.4 per row.  The nine centres
// therefore have only three distinct x and three distinct y offsets, each the same chain of roundings whatever the row: computed
// once here (a dependent chain of 2 + 4 conversions instead of 13), same bits.
__device__ __forceinline__ void many_sphere_len2(f3 p, double zoff, float (&len2)[9]) {
    float xs[3], ys[3];
    xs[0] = p.x + 0.5f;
    xs[1] = (float)__dsub_rn((double)xs[0], 0.4);
    xs[2] = (float)__dsub_rn((double)xs[1], 0.4);
    const float y_init = (float)__dsub_rn((double)p.y, 0.6);
    ys[0] = (float)__dadd_rn((double)y_init, 0.4);
    ys[1] = (float)__dadd_rn((double)ys[0], 0.4);
    ys[2] = (float)__dadd_rn((double)ys[1], 0.4);
    const float cz = (float)__dadd_rn((double)p.z, zoff);
    const float zz = cz * cz;
#pragma unroll
    for (int r = 0; r < 3; ++r) {
        const float yy = ys[r] * ys[r];
#pragma unroll
        for (int c = 0; c < 3; ++c) len2[3 * r + c] = zz + fmaf(xs[c], xs[c], yy);
    }
}
// The chain itself for a point that IS near a sphere (rays grazing a sphere stay in this case for thousands of steps, so it
// is the frame's critical path: measured 2.7 us per dependent step with the reference's loop against 1.3 us elsewhere).
// Same smooth unions in the same order, but a sphere whose squared distance proves d_i - s >= k for every value s can take
// along the chain (the running s never exceeds nSDF: a smooth union is <= min(d1, d2) up to an ulp, covered by the margin) is
// the identity and is skipped without its square root or its double-precision blend; typically one or two spheres are left.
// len2 are the nine squared centre distances many_sphere_len2 has just produced; the loop walks the set bits of the
// near-sphere mask (lanes of a warp work on different spheres in the same iteration), picking its len2 with selects.
__device__ __forceinline__ float many_sphere_union_chain(const float (&len2)[9], float nsdf, bool ref) {
    const bool plain = nsdf != 0.0f && fabsf(nsdf) <= 1.0e30f;
    const float thr = nsdf + 0.11002f;
    const float thr2 = (thr * thr) * 1.000001f;
    unsigned int near = 0u;
#pragma unroll
    for (int i = 0; i < 9; ++i) {
        const bool far = plain && len2[i] <= 1.0e30f && (thr <= 0.0f || len2[i] >= thr2);
        near |= far ? 0u : (1u << i);
    }
    float s = nsdf;
#pragma unroll 1
    while (near) {
        const int i = __ffs(near) - 1;
        near &= near - 1u;
        float l = len2[0];
#pragma unroll
        for (int j = 1; j < 9; ++j) l = (i == j) ? len2[j] : l;
        s = smooth_union(s, __fsqrt_rn(l) + (-0.1f), 0.01f, ref);
    }
    return s;
}
// Clearance tracking.  `rem` is a lower bound of min_i sqrt(len2_i) at the ray's current point (negative = unknown), carried
// from step to step: a march step moves the point by at most |t| (|d| = 1 up to rounding), so the clearance shrinks by at most
// that; the caller subtracts clearance_step(t) after every step.  While rem >= nSDF + 0.11002 the union is the identity (see
// above) and a march step costs one compare -- no double arithmetic, no square root; when the bound is exhausted it is
// refreshed from the exact minimum at the current point.  Slack: the computed centre offsets are within ~4e-7 of the true
// ones and a step's rounding moves the point by < 3e-7, hence the 3e-6 at refresh and the 5e-7 per step.
__device__ __forceinline__ float clearance_step(float t) { return fmaf(fabsf(t), 1.000002f, 5.0e-7f); }
__device__ __forceinline__ float many_sphere_union(f3 p, float nsdf, double zoff, bool ref, float& rem) {
    const bool plain = nsdf != 0.0f && fabsf(nsdf) <= 1.0e30f;
    const float thr = nsdf + 0.11002f;
    if (plain && rem >= 0.0f && (thr <= 0.0f || rem >= thr)) return nsdf;
    float l[9];
    many_sphere_len2(p, zoff, l);
    float m = l[0];
    bool ok = true;       // a NaN anywhere (a NaN point) sends the query the long way
#pragma unroll
    for (int i = 1; i < 9; ++i) m = (l[i] < m) ? l[i] : m;
#pragma unroll
    for (int i = 0; i < 9; ++i) ok = ok && (l[i] == l[i]);
    const bool valid = ok && m <= 1.0e30f;
    rem = valid ? __fsqrt_rd(m) - 3.0e-6f : -1.0f;
    const float thr2 = nsdf + 0.11001f;
    if (plain && valid && (thr2 <= 0.0f || m >= (thr2 * thr2) * 1.000001f)) return nsdf;
    return many_sphere_union_chain(l, nsdf, ref);
}

enum { SCENE_TANH = 0, SCENE_MANYSPHERE = 1 };

// sceneSDF (volumeRender_kernel.cu:217-230): scene 1 = HEAD, scene 0 = `tanh(nSDF)` (:229).  rem: clearance state of the ray
// (many_sphere_union), untouched by the tanh scene.
__device__ __forceinline__ float scene_sdf(int scene, f3 p, float nsdf, double zoff, bool ref, float& rem) {
    if (scene == SCENE_MANYSPHERE) return many_sphere_union(p, nsdf, zoff, ref, rem);
    return tanh_sel(nsdf, ref);
}

#define NSDF_NORMAL_EPSILON 0.00001f     // volumeRender_kernel.cu:59
#define NSDF_MARCHING_EPSILON 0.000001f  // volumeRender_kernel.cu:60

// tetrahedronVerts (volumeRender_kernel.cu:38-43) as sign bits: bit i of TET_NEG_{X,Y,Z}
// set <=> component of vertex i is -1.
#define NSDF_TET_NEG_X 0x6u  // (+,-,-,+)
#define NSDF_TET_NEG_Y 0x3u  // (-,-,+,+)
#define NSDF_TET_NEG_Z 0x5u  // (-,+,-,+)
__device__ __forceinline__ f3 tet_vertex(int i) {
    return f3{((NSDF_TET_NEG_X >> i) & 1u) ? -1.0f : 1.0f, ((NSDF_TET_NEG_Y >> i) & 1u) ? -1.0f : 1.0f,
              ((NSDF_TET_NEG_Z >> i) & 1u) ? -1.0f : 1.0f};
}

// float -> int/uint conversion with the semantics of cvt.rzi (NaN -> 0, saturating); this is
// what the reference's device code gets for int(f) / uint(f) (:273, :398-399).
__device__ __forceinline__ int f2i_rz(float f) { return __float2int_rz(f); }
__device__ __forceinline__ uint32_t f2u_rz(float f) { return __float2uint_rz(f); }

struct Camera {
    float inv_view[12];  // c_invViewMatrix rows 0..2 (volumeRender_kernel.cu:31)
    float normal[16];    // c_normalMatrix (:32)
    int frame;           // c_frameNumber (:35)
    double zoff;         // many_sphere_zoff(frame), filled by the host side of the launch
};

struct RayInit { f3 o, d, p; float tfar; bool hit; };

// initMarcher (volumeRender_kernel.cu:293-358) + intersectSphere (:200-215), sphere (0, 1.2).
__device__ __forceinline__ RayInit init_ray(const Camera& c, int x, int y, int W, int H, bool ref) {
    RayInit r;
    const float u = fmaf(__fdiv_rn((float)x, (float)W), 2.0f, -1.0f);
    const float v = fmaf(__fdiv_rn((float)y, (float)H), 2.0f, -1.0f);
    const float* M = c.inv_view;
    r.o = f3{M[3] + fmaf(M[2], 0.0f, fmaf(M[0], 0.0f, M[1] * 0.0f)),
             M[7] + fmaf(M[6], 0.0f, fmaf(M[4], 0.0f, M[5] * 0.0f)),
             M[11] + fmaf(M[10], 0.0f, fmaf(M[8], 0.0f, M[9] * 0.0f))};
    const float rs = rsqrt_sel(fmaf(u, u, v * v) + 4.0f, ref);
    const f3 d0{u * rs, v * rs, rs * -2.0f};
    r.d = f3{dot3(d0, f3{M[0], M[1], M[2]}), dot3(d0, f3{M[4], M[5], M[6]}), dot3(d0, f3{M[8], M[9], M[10]})};
    const float a = dot3(r.d, r.d);
    const float qd = dot3(r.o, r.d);
    const float b = qd + qd;
    const float cc = dot3(r.o, r.o) + (-1.4400000572204589844f);
    const float disc = fmaf(b, b, -(cc * (a * 4.0f)));
    r.hit = disc > 0.0f;
    r.tfar = 0.0f;
    r.p = r.o;
    if (!r.hit) return r;
    const float sq = __fsqrt_rn(disc);
    const double two_a = __dadd_rn((double)a, (double)a);
    float tnear = (float)__ddiv_rn((double)(-b - sq), two_a);
    r.tfar = (float)__ddiv_rn((double)(sq - b), two_a);
    if (tnear < 0.0f) tnear = 0.0f;
    r.p = f3{fmaf(r.d.x, tnear, r.o.x), fmaf(r.d.y, tnear, r.o.y), fmaf(r.d.z, tnear, r.o.z)};
    return r;
}

// rgbaFloatToInt (volumeRender_kernel.cu:266-274)
__device__ __forceinline__ uint32_t rgba_float_to_int(float r, float g, float b, float a, bool ref) {
    r = saturate_sel(r, ref); g = saturate_sel(g, ref); b = saturate_sel(b, ref); a = saturate_sel(a, ref);
    return (f2u_rz(a * 255.0f) << 24) | (f2u_rz(b * 255.0f) << 16) | (f2u_rz(g * 255.0f) << 8) | f2u_rz(r * 255.0f);
}

// facingColor (volumeRender_kernel.cu:380-384)
__device__ __forceinline__ uint32_t facing_color(f3 n, f3 ray, bool ref) {
    float t = ray.x * n.x;
    t = fmaf(-ray.y, n.y, -t);
    t = fmaf(-ray.z, n.z, t);
    double rd = (double)t;
    rd = (rd > 0.0) ? rd : 0.0;
    const float ratio = (float)rd;
    return rgba_float_to_int(ratio, ratio, ratio, 1.0f, ref);
}

// matCapColor (volumeRender_kernel.cu:387-413): nearest texel, truncating, texel returned as is.
__device__ __forceinline__ uint32_t matcap_color(const Camera& c, f3 n, const uint32_t* __restrict__ matcap, int mw,
                                                 int mh, bool ref) {
    const float* N = c.normal;
    float e[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        const float* row = N + 4 * i;
        float t = n.y * row[1];
        t = fmaf(n.x, row[0], t);
        t = fmaf(n.z, row[2], t);
        t = fmaf(row[3], 0.0f, t);
        e[i] = t;
    }
    const f3 ne = normalize3(f3{e[0], e[1], e[2]}, ref);
    const float fu = (float)fma((double)ne.x, 0.5, 0.5);
    const float fv = (float)fma((double)ne.y, 0.5, 0.5);
    const int ux = f2i_rz(fu * (float)(mw - 1));
    const int uy = f2i_rz(fv * (float)(mh - 1));
    const int index = (int)((uint32_t)uy * (uint32_t)mw + (uint32_t)ux);
    if (index < 0 || index >= mw * mh) return 0u;  // :403-407 (the upper bound is our guard; the reference would fault)
    return __ldg(matcap + index);
}

// surfaceNormal (volumeRender_kernel.cu:361-377): p0+p1+p2+p3 as contracted by nvcc.
__device__ __forceinline__ f3 tet_normal(const float (&s)[4], bool ref) {
    f3 nn;
    const f3 t0 = tet_vertex(0), t1 = tet_vertex(1), t2 = tet_vertex(2), t3 = tet_vertex(3);
    nn.x = fmaf(t3.x, s[3], fmaf(t2.x, s[2], fmaf(t0.x, s[0], t1.x * s[1])));
    nn.y = fmaf(t3.y, s[3], fmaf(t2.y, s[2], fmaf(t0.y, s[0], t1.y * s[1])));
    nn.z = fmaf(t3.z, s[3], fmaf(t2.z, s[2], fmaf(t0.z, s[0], t1.z * s[1])));
    return normalize3(nn, ref);
}

}  // namespace nsdf
