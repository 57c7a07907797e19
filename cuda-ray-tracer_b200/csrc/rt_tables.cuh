// rt_tables.cuh — FP32 sphere-rejection tables specialised per ray family (queue kernel, rt_queue.cu).
//
// Almost every ray of a frame belongs to a family that shares a point or a direction:
//   primary rays            all start at the camera origin                      (update-cpu.cpp:84-89,123)
//   point-light shadow rays all END (within the shadow bias) at the light       (light_impl.h:20-22)
//   directional shadow rays all have the light's direction                      (light_impl.h:24-25)
// For such a family everything in the ray/sphere discriminant that does not depend on the individual ray is
// hoisted into a per-(family, sphere) float4, so the hot loop costs 5 (pivot families) or 4 (direction
// families) FP32 instructions per sphere instead of the 14 of the generic filter (rt_filter.cuh), which stays
// in use for reflection rays.  Like the generic filter, a table can only REJECT a sphere when the infinite
// line provably misses it (with a margin covering every FP32 rounding and, for shadow rays, the distance
// between the true segment and the family's idealised line); everything else goes to the exact FP64 test
// (intersect_sphere, reference operation order), so results are bit-identical to testing every sphere exactly.
//
// Pivot family, pivot X, filter direction e (FP32), A = e.e, oc = X - C:
//   real discriminant/4 of the line X + s e against sphere (C, r):  disc = (oc.e)^2 - A (|oc|^2 - r^2)
//   entry {oc (rounded to FP32), c' = |oc|^2 - r^2 - m},  m = 2 u (18 |oc|^2 + 7 |r^2|)  (u = 2^-24)
//   slack = fma(b, b, -(A c')) with b = oc.e in FP32.  First-order error of the FP32 evaluation:
//     |b_f - b| <= 5 u Q D, |b_f^2 - b^2| <= 10 u Q^2 D^2, |A_f c_f - A c| <= 6 u D^2 (Q^2 + r^2), final fma
//     rounding <= u D^2 (2 Q^2 + r^2)  =>  |disc_f - disc| <= u D^2 (18 Q^2 + 7 r^2); m is twice that per unit A.
//   A shadow segment runs from o' = P + 1e-2 n to o' + float(L - P): it stays within
//   eps = 1e-2 + |rounding of the direction| of the idealised segment [L, o'], so it can only touch sphere (C, r)
//   if the idealised line touches (C, r + eps):  disc >= -A (2 r eps + eps^2) =: -kappa.  The kernel therefore
//   rejects only when slack < -kappa (kappa = 0 for primary rays, whose rays pass through X exactly).
//
// Direction family, unit direction s with an orthonormal pair (u, v) spanning the plane normal to s:
//   squared distance from C to the line through o' = ((o'-C).u)^2 + ((o'-C).v)^2
//   entry {C.u, C.v, thr = r^2 (1 + 1e-6) rounded up, 0};  slack = thr - (du^2 + dv^2), du = o'.u - C.u (FP32)
//   with e1 = 2 u (max(|o'.u|, |o'.v|) + max|C|) bounding the error of du, dv:
//     g <= r^2  =>  f <= r^2 (1 + 3u) + 3 r e1 + 2.1 e1^2, so reject only when slack < -kappa,
//     kappa = 2 (3 r_max e1 + 2.1 e1^2).
#ifndef RT_TABLES_CUH
#define RT_TABLES_CUH

#include "rt_filter.cuh"

namespace rt {

enum FilterKind { FK_GENERIC = 0, FK_PIVOT = 1, FK_DIRECTION = 2 };

// Per-light filter family (built at upload).
struct LightFilter {
    double ux, uy, uz;   // FK_DIRECTION: orthonormal pair normal to the light direction
    double vx, vy, vz;
    double sx, sy, sz;   // directional lights: the shadow direction exactly as shadow_ray() returns it
                         // (rounded to float and widened, light_impl.h:24-25) — the same for every hit point
    int kind;            // FilterKind; FK_GENERIC when the light's data is unusable for a table (zero / huge vectors)
    int pad;
};

#define RT_NEG_INF_F (-CUDART_INF_F)

// ---- table builders (device code shared by the setup kernels in rt_queue.cu) -------------------------------------
__device__ __forceinline__ float4 pivot_entry(const double4 k, double X, double Y, double Z)
{
    // |p|^2 + l.p + c = 0  ->  centre -l/2, r^2 = |centre|^2 - c   (surface.cpp:4-15)
    const double cx = -0.5 * k.x, cy = -0.5 * k.y, cz = -0.5 * k.z;
    const double r2 = cx * cx + cy * cy + cz * cz - k.w;
    const double ox = X - cx, oy = Y - cy, oz = Z - cz;
    const double q2 = ox * ox + oy * oy + oz * oz;
    const double m = 2.0 * 5.9604645e-8 * (18.0 * q2 + 7.0 * fabs(r2)) * 1.000001;
    return make_float4(__double2float_rn(ox), __double2float_rn(oy), __double2float_rn(oz),
                       __double2float_rd(q2 - r2 - m));
}

__device__ __forceinline__ float4 direction_entry(const double4 k, const LightFilter &lf)
{
    const double cx = -0.5 * k.x, cy = -0.5 * k.y, cz = -0.5 * k.z;
    const double r2 = cx * cx + cy * cy + cz * cz - k.w;
    return make_float4(__double2float_rn(cx * lf.ux + cy * lf.uy + cz * lf.uz),
                       __double2float_rn(cx * lf.vx + cy * lf.vy + cz * lf.vz),
                       __double2float_ru(r2 + 1e-6 * fabs(r2)), 0.0f);
}

// Entry no ray can ever select (padding up to a multiple of four spheres).
__device__ __forceinline__ float4 never_entry(int kind)
{
    if (kind == FK_DIRECTION) return make_float4(0.f, 0.f, RT_NEG_INF_F, 0.f);
    if (kind == FK_PIVOT) return make_float4(0.f, 0.f, 0.f, CUDART_INF_F);
    return make_float4(0.f, 0.f, 0.f, RT_NEG_INF_F);   // generic {centre, r^2}: c = q2 + inf -> disc = -inf
}

// ---- hot-loop halves: slack of one sphere, larger = closer to (or inside) the sphere -------------------------------
struct PivotRay {
    float ex, ey, ez, A;
};
struct DirRay {
    float ou, ov;
};

__device__ __forceinline__ float pivot_slack(const PivotRay &r, const float4 s)
{
    const float b = fmaf(s.x, r.ex, fmaf(s.y, r.ey, s.z * r.ez));
    return fmaf(b, b, -(r.A * s.w));
}

__device__ __forceinline__ float dir_slack(const DirRay &r, const float4 s)
{
    const float du = r.ou - s.x, dv = r.ov - s.y;
    return fmaf(-dv, dv, fmaf(-du, du, s.z));
}

__device__ __forceinline__ float max4(float a, float b, float c, float d) { return fmaxf(fmaxf(a, b), fmaxf(c, d)); }

} // namespace rt
#endif
