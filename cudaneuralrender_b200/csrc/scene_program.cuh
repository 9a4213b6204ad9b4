// =====================================================================================
// scene_program.cuh -- the reference's analytic SDF library as a run-time program.
//
// src/volumeRender_kernel.cu:67-196 holds sphere / box / plane / cylinder primitives, displace /
// round / onion modifiers, hard and smooth boolean operators and two composite scenes
// (manyCylinderCut, manySphere); the scene is chosen by editing sceneSDF (:217-230) and
// recompiling.  Here the same functions are interpreted from a short op list that travels in
// the kernel parameters (include/nsdf.h, nsdf_scene_op): one accumulator, each op computes a
// primitive and combines it, or modifies the accumulator.  Every scene sceneSDF holds (live or
// commented out) is such a chain.
//
// Arithmetic contract (the oracle's scene_program_eval follows the same lines): fp32 as written in
// the reference, its double-promoting literals kept as double operations, dot products contracted
// the way nvcc contracts helper_math's (fma(z,z, fma(x,x, y*y))), min/max with explicit NaN rules,
// sin through sinf_det (Cody-Waite by pi/2 in three parts + the Cephes single-precision kernels,
// only explicit fmaf).  Most of these functions are dead code in the reference build, so there is no
// SASS to pin contraction against: parity for this file is oracle <-> device, bit for bit.
// =====================================================================================
#pragma once
#include "device_math.cuh"

namespace nsdf {

constexpr int SCENE_MAX_OPS = 16;
struct SceneOp { int prim; int combine; float a[8]; };
struct SceneProgram { int n_ops; int pad; SceneOp ops[SCENE_MAX_OPS]; };

enum { PRIM_NSDF = 0, PRIM_TANH = 1, PRIM_SPHERE = 2, PRIM_BOX = 3, PRIM_PLANE = 4, PRIM_CYLINDER = 5,
       MOD_DISPLACE = 16, MOD_ROUND = 17, MOD_ONION = 18, MOD_MANY_SPHERE = 19, MOD_MANY_CYLINDER_CUT = 20 };
enum { COMB_SET = 0, COMB_UNION = 1, COMB_INTERSECT = 2, COMB_SUBTRACT = 3, COMB_SUBTRACT_FROM = 4, COMB_SMOOTH_UNION = 5,
       COMB_SMOOTH_SUBTRACT = 6, COMB_SMOOTH_SUBTRACT_FROM = 7 };

__device__ __forceinline__ float min_det(float a, float b) { return (a != a) ? b : ((b != b) ? a : ((b < a) ? b : a)); }
__device__ __forceinline__ float max_det(float a, float b) { return (a != a) ? b : ((b != b) ? a : ((b > a) ? b : a)); }

// sin for |x| < 1e6 (else NaN): q = rint(x * 2/pi), r = x - q*pi/2 in three parts, Cephes sinf/cosf kernels on |r| <= pi/4.
__device__ __forceinline__ float sinf_det(float x) {
    if (!(fabsf(x) < 1.0e6f)) return x - x;
    const float q = rintf(x * 0.636619772f);
    float r = fmaf(q, -1.5703125f, x);
    r = fmaf(q, -4.837512969970703125e-4f, r);
    r = fmaf(q, -7.54978995489188216e-8f, r);
    const int n = (int)q;
    const float z = r * r;
    const float s = fmaf(fmaf(fmaf(-1.9515295891e-4f, z, 8.3321608736e-3f), z, -1.6666654611e-1f) * z, r, r);
    const float c = fmaf(fmaf(fmaf(2.443315711809948e-5f, z, -1.388731625493765e-3f), z, 4.166664568298827e-2f) * z, z,
                         fmaf(-0.5f, z, 1.0f));
    const float v = (n & 1) ? c : s;
    return (n & 2) ? -v : v;
}

// sdfOpSmoothSubtraction (volumeRender_kernel.cu:138-142) with its double promotions.
// Outside the blend band it collapses exactly like smooth_union: d1 + d2 >= k gives h == 0 and the value d1, d1 + d2 <= -k gives
// h == 1 and the value -d2 (manyCylinderCut applies 300 of these per query; almost all take one of the two exits).
__device__ __forceinline__ float smooth_subtract(float d1, float d2, float k, bool ref) {
    const float sum = d1 + d2;
    if (sum >= k && sum <= 1.0e30f && d1 != 0.0f && fabsf(d2) <= 1.0e30f) return d1;
    if (sum <= -k && sum >= -1.0e30f && d2 != 0.0f && fabsf(d1) <= 1.0e30f) return -d2;
    const double hd = __dsub_rn(0.5, __ddiv_rn(__dmul_rn((double)sum, 0.5), (double)k));
    const float h = saturate_sel((float)hd, ref);
    const double omh = __dsub_rn(1.0, (double)h);
    const float d2h = d2 * h;
    const float mix = (float)fma(omh, (double)d1, -(double)d2h);
    const float kh = k * h;
    return (float)fma(omh, (double)kh, (double)mix);
}

__device__ __forceinline__ float length3_det(f3 v) { return __fsqrt_rn(dot3(v, v)); }

// sdfBox (:81-89): length(max(q,0)) + min(max(q.x,max(q.y,q.z)), 0.0) - r, the tail in double like the source
__device__ __forceinline__ float sdf_box(f3 p, f3 b, float r) {
    const f3 q = f3{fabsf(p.x) - b.x, fabsf(p.y) - b.y, fabsf(p.z) - b.z};
    const f3 m = f3{max_det(q.x, 0.0f), max_det(q.y, 0.0f), max_det(q.z, 0.0f)};
    const float inner = max_det(q.x, max_det(q.y, q.z));
    const double in0 = (double)inner < 0.0 ? (double)inner : 0.0;
    return (float)__dsub_rn(__dadd_rn((double)length3_det(m), in0), (double)r);
}

// sdfCylinder (:96-101): length(p.xy - (c.x, c.z)) - c.y
__device__ __forceinline__ float sdf_cylinder(f3 p, f3 c) {
    const float dx = p.x - c.x, dy = p.y - c.z;
    return __fsqrt_rn(fmaf(dx, dx, dy * dy)) - c.y;
}

// manySphere(p, nSDF, doUnion) (:176-196); the union form is many_sphere_union of device_math.cuh, bit for bit.
__device__ __forceinline__ float many_sphere(f3 p, float nsdf, double zoff, bool do_union, bool ref) {
    if (do_union) { float rem = -1.0f; return many_sphere_union(p, nsdf, zoff, ref, rem); }
    float s = nsdf;
    float cx = p.x;
    float cy = (float)__dsub_rn((double)p.y, 0.6);
    const float cz = (float)__dadd_rn((double)p.z, zoff);
    const float zz = cz * cz;
#pragma unroll 1
    for (int i = 0; i < 9; ++i) {
        if (i % 3 == 0) {
            cy = (float)__dadd_rn((double)cy, 0.4);
            cx = p.x + 0.5f;
        }
        const float len2 = zz + fmaf(cx, cx, cy * cy);
        const float d = __fsqrt_rn(len2) + (-0.1f);
        s = smooth_subtract(s, d, 0.01f, ref);
        cx = (float)__dsub_rn((double)cx, 0.4);
    }
    return s;
}

// manyCylinderCut (:156-174): 300 cylinders of radius 0.02 carved out smoothly, 15 rows of 20
__device__ __forceinline__ float many_cylinder_cut(f3 p, float nsdf, bool ref) {
    float s = nsdf;
    const f3 c = f3{0.02f, 0.02f, 0.02f};
    f3 cp = p;
    cp.y = (float)__dsub_rn((double)cp.y, 0.5);
#pragma unroll 1
    for (int i = 0; i < 300; ++i) {
        if (i % 20 == 0) {
            cp.y = (float)__dadd_rn((double)cp.y, 0.1);
            cp.x = (float)__dadd_rn((double)p.x, 0.9);
        }
        // sdfCylinder as nvcc contracts it inside this loop of the reference (ref_sdf_probe SASS: FMUL dx,dx ; FFMA dy,dy,.):
        // fma(dy, dy, dx*dx), not the stand-alone function's fma(dx, dx, dy*dy)
        const float dx = cp.x - c.x, dy = cp.y - c.z;
        const float cyl = __fsqrt_rn(fmaf(dy, dy, dx * dx)) - c.y;
        s = smooth_subtract(s, cyl, 0.01f, ref);
        cp.x = (float)__dsub_rn((double)cp.x, 0.1);
    }
    return s;
}

__device__ __noinline__ float scene_program_eval(const SceneProgram& prog, f3 p, float nsdf, double zoff, bool ref) {
    float s = 0.0f;
    const int n = prog.n_ops < SCENE_MAX_OPS ? prog.n_ops : SCENE_MAX_OPS;
#pragma unroll 1
    for (int i = 0; i < n; ++i) {
        const SceneOp& op = prog.ops[i];
        const float* a = op.a;
        if (op.prim >= MOD_DISPLACE) {
            switch (op.prim) {
                case MOD_DISPLACE: {
                    const float sx = ref ? sinf(a[0] * p.x) : sinf_det(a[0] * p.x), sy = ref ? sinf(a[0] * p.y) : sinf_det(a[0] * p.y),
                                sz = ref ? sinf(a[0] * p.z) : sinf_det(a[0] * p.z);
                    const float w = (sx * sy) * sz;
                    // the reference's amplitude is the DOUBLE literal 0.05 (:107); a float parameter equal to 0.05f means that literal
                    const double amp = a[1] == 0.05f ? 0.05 : (double)a[1];
                    s = (float)fma((double)w, amp, (double)s);
                    break;
                }
                case MOD_ROUND: s = s - a[0]; break;
                case MOD_ONION: s = fabsf(s) - a[0]; break;
                case MOD_MANY_SPHERE: s = many_sphere(p, s, zoff, a[0] != 0.0f, ref); break;
                case MOD_MANY_CYLINDER_CUT: s = many_cylinder_cut(p, s, ref); break;
                default: break;
            }
            continue;
        }
        float d;
        const f3 q = f3{p.x - a[0], p.y - a[1], p.z - a[2]};
        switch (op.prim) {
            case PRIM_NSDF: d = nsdf; break;
            case PRIM_TANH: d = tanh_sel(nsdf, ref); break;
            case PRIM_SPHERE: d = length3_det(q) - a[3]; break;
            case PRIM_BOX: d = sdf_box(q, f3{a[3], a[4], a[5]}, a[6]); break;
            case PRIM_PLANE: d = p.y - a[0]; break;
            case PRIM_CYLINDER: d = sdf_cylinder(q, f3{a[3], a[4], a[5]}); break;
            default: d = nsdf; break;
        }
        switch (op.combine) {
            case COMB_UNION: s = min_det(s, d); break;
            case COMB_INTERSECT: s = max_det(s, d); break;
            case COMB_SUBTRACT: s = max_det(s, -d); break;
            case COMB_SUBTRACT_FROM: s = max_det(d, -s); break;
            case COMB_SMOOTH_UNION: s = smooth_union(s, d, a[7], ref); break;
            case COMB_SMOOTH_SUBTRACT: s = smooth_subtract(s, d, a[7], ref); break;
            case COMB_SMOOTH_SUBTRACT_FROM: s = smooth_subtract(d, s, a[7], ref); break;
            default: s = d; break;
        }
    }
    return s;
}

}  // namespace nsdf
