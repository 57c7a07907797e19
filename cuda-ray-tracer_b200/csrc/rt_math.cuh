// rt_math.cuh — exact (reference-order) FP64 ray/surface arithmetic for sm_100a.
//
// Everything here rounds where the reference's expressions round: this translation unit is compiled
// with -fmad=false, IEEE division and square root, so the only differences from the CPU oracle are the
// last-ulp differences between libdevice's and glibc's cbrt / acos / cos.
//   include/surface_impl.h:21-155   intersect_ray  -> expand_* (per structural class) + solve_poly
//   include/surface_impl.h:157-172  normal_vector  -> surface_normal
//   include/light_impl.h:17-49      shadow_ray / surface_color / reflect_ray -> shadow_dir, shade_light, reflect
#ifndef RT_MATH_CUH
#define RT_MATH_CUH

#include <math_constants.h>

#include "rt_scene.h"

namespace rt {

// include/surface_impl.h:16-19
#define RT_EPS 1e-7
#define RT_TWO_THIRD_PI (3.14159265358979323846 * 2.0 / 3.0)
#define RT_SHADOW_BIAS 1e-2
#define RT_MAX_T 1e6
#define RT_PI_F 3.14159274101257324f /* M_PIf32 */

struct d3 {
    double x, y, z;
};

__device__ __forceinline__ double dot3(const d3 &a, const d3 &b) { return a.x * b.x + a.y * b.y + a.z * b.z; }

// GLM normalize: v * (1 / sqrt(dot(v, v)))
__device__ __forceinline__ d3 normalize3(const d3 &v)
{
    double inv = 1.0 / sqrt(dot3(v, v));
    return d3{v.x * inv, v.y * inv, v.z * inv};
}

// Coefficients of F(o + t d) = t3 t^3 + t2 t^2 + t1 t + t0.
struct Poly {
    double t3, t2, t1, t0;
};

// ---- closed-form root selection, surface_impl.h:105-154 ----------------------------------------
__device__ __forceinline__ double solve_quadratic_or_linear(double t2, double t1, double t0)
{
    if (fabs(t2) > RT_EPS) {
        double delta = t1 * t1 - 4.0 * t2 * t0;
        if (delta < 0) return -1.0;
        delta = sqrt(delta);
        double x = (-t1 - delta) / (2.0 * t2);
        if (x >= RT_EPS) return x;
        return (-t1 + delta) / (2.0 * t2);
    }
    if (fabs(t1) > RT_EPS) return -t0 / t1;
    return -1.0;
}

// x / b for a finite non-zero divisor, with numerators that are exactly zero short-cut: 0 / b is the zero with the
// quotient's sign, but on the GPU it sends the FP64 division sequence to its slow path (a subroutine of ~60
// instructions) — and exact zeros are common here: with the camera at the origin of a surface's frame whole
// coefficients of F(o + t d) vanish (Fermat cubic: t2 = t1 = 0 for every primary ray).
// (Measured too: the three divisions by t3 and the divisions by 9, 54 and 3 as FMA corrections of a product with the
// correctly rounded reciprocal — same bits, fewer instructions, but not faster over the five cubic scenes.)
__device__ __forceinline__ double div_nz(double x, double b)
{
    if (x == 0.0) return x * copysign(1.0, b);
    return x / b;
}

__device__ __forceinline__ double solve_cubic_inline(double t3, double t2, double t1, double t0)
{
    t2 = div_nz(t2, t3);   // |t3| > 1e-7 here (surface_impl.h:106)
    t1 = div_nz(t1, t3);
    t0 = div_nz(t0, t3);
    double q = div_nz(3.0 * t1 - t2 * t2, 9.0);
    double r = div_nz(9.0 * t2 * t1 - 27.0 * t0 - 2.0 * t2 * t2 * t2, 54.0);
    const double t2_3 = div_nz(t2, 3.0);   // t2 / 3.0, the same value in every root below
    double delta = q * q * q + r * r;
    if (delta > 0) {
        delta = sqrt(delta);
        q = cbrt(r + delta);
        r = cbrt(r - delta);
        return q + r - t2_3;
    }
    double theta = acos(r / sqrt(-q * q * q)) / 3.0;
    double c = 2.0 * sqrt(-q);
    double x = c * cos(theta) - t2_3;
    double x1 = c * cos(theta + RT_TWO_THIRD_PI) - t2_3;
    if (x1 >= RT_EPS && x1 < x) x = x1;
    x1 = c * cos(theta + 2.0 * RT_TWO_THIRD_PI) - t2_3;
    if (x1 >= RT_EPS && x1 < x) x = x1;
    return x;
}

// out-of-line copy for kernels that call the solver from many places (exact and wavefront kernels)
static __device__ __noinline__ double solve_cubic(double t3, double t2, double t1, double t0)
{
    return solve_cubic_inline(t3, t2, t1, t0);
}

__device__ __forceinline__ double solve_poly(const Poly &p)
{
    if (fabs(p.t3) > RT_EPS) return solve_cubic(p.t3, p.t2, p.t1, p.t0);
    return solve_quadratic_or_linear(p.t2, p.t1, p.t0);
}

// ---- per-class expansions (non-zero terms of surface_impl.h:46-103, same left-to-right order) ---
__device__ __forceinline__ double intersect_plane(const double4 k, const d3 &o, const d3 &d)
{
    double t1 = k.x * d.x + k.y * d.y + k.z * d.z;
    double t0 = k.x * o.x + k.y * o.y + k.z * o.z + k.w;
    if (fabs(t1) > RT_EPS) return -t0 / t1;
    return -1.0;
}

// x2 = y2 = z2 = 1 (multiplying by 1.0 is exact, so the products with the unit coefficients vanish)
__device__ __forceinline__ double intersect_sphere(const double4 k, const d3 &o, const d3 &d)
{
    double t2 = d.x * d.x + d.y * d.y + d.z * d.z;
    double t1 = 2.0 * o.x * d.x + 2.0 * o.y * d.y + 2.0 * o.z * d.z + k.x * d.x + k.y * d.y + k.z * d.z;
    double t0 = o.x * o.x + o.y * o.y + o.z * o.z + k.x * o.x + k.y * o.y + k.z * o.z + k.w;
    return solve_quadratic_or_linear(t2, t1, t0);
}

// q = {x2,y2,z2,xy,xz,yz,x,y,z,c}
__device__ __forceinline__ Poly expand_quadric(const double *__restrict__ q, const d3 &o, const d3 &d)
{
    Poly p;
    p.t3 = 0.0;
    p.t2 = q[0] * (d.x * d.x) + q[1] * (d.y * d.y) + q[2] * (d.z * d.z)
           + q[3] * (d.x * d.y) + q[4] * (d.x * d.z) + q[5] * (d.y * d.z);
    p.t1 = q[0] * (2.0 * o.x * d.x) + q[1] * (2.0 * o.y * d.y) + q[2] * (2.0 * o.z * d.z)
           + q[3] * (o.x * d.y + d.x * o.y) + q[4] * (o.x * d.z + d.x * o.z) + q[5] * (o.y * d.z + d.y * o.z)
           + q[6] * d.x + q[7] * d.y + q[8] * d.z;
    p.t0 = q[0] * (o.x * o.x) + q[1] * (o.y * o.y) + q[2] * (o.z * o.z)
           + q[3] * (o.x * o.y) + q[4] * (o.x * o.z) + q[5] * (o.y * o.z)
           + q[6] * o.x + q[7] * o.y + q[8] * o.z + q[9];
    return p;
}

__device__ __forceinline__ double intersect_quadric(const double *__restrict__ q, const d3 &o, const d3 &d)
{
    double t2 = q[0] * (d.x * d.x) + q[1] * (d.y * d.y) + q[2] * (d.z * d.z)
                + q[3] * (d.x * d.y) + q[4] * (d.x * d.z) + q[5] * (d.y * d.z);
    double t1 = q[0] * (2.0 * o.x * d.x) + q[1] * (2.0 * o.y * d.y) + q[2] * (2.0 * o.z * d.z)
                + q[3] * (o.x * d.y + d.x * o.y) + q[4] * (o.x * d.z + d.x * o.z) + q[5] * (o.y * d.z + d.y * o.z)
                + q[6] * d.x + q[7] * d.y + q[8] * d.z;
    double t0 = q[0] * (o.x * o.x) + q[1] * (o.y * o.y) + q[2] * (o.z * o.z)
                + q[3] * (o.x * o.y) + q[4] * (o.x * o.z) + q[5] * (o.y * o.z)
                + q[6] * o.x + q[7] * o.y + q[8] * o.z + q[9];
    return solve_quadratic_or_linear(t2, t1, t0);
}

// (a0 + t a)^2 (b0 + t b): t^2 and t^1 parts, surface_impl.h:33-34
__device__ __forceinline__ double sq_lin_t2(double oa, double da, double ob, double db)
{
    return da * (da * ob + 2.0 * oa * db);
}
__device__ __forceinline__ double sq_lin_t1(double oa, double da, double ob, double db)
{
    return oa * (oa * db + 2.0 * da * ob);
}

__device__ __forceinline__ Poly expand_cubic(const double *__restrict__ k, const d3 &o, const d3 &d)
{
    Poly p;
    p.t3 = k[X3] * (d.x * d.x * d.x) + k[Y3] * (d.y * d.y * d.y) + k[Z3] * (d.z * d.z * d.z)
           + k[X2Y] * (d.x * d.x * d.y) + k[XY2] * (d.x * d.y * d.y)
           + k[X2Z] * (d.x * d.x * d.z) + k[XZ2] * (d.x * d.z * d.z)
           + k[Y2Z] * (d.y * d.y * d.z) + k[YZ2] * (d.y * d.z * d.z)
           + k[XYZ] * (d.x * d.y * d.z);
    p.t2 = k[X3] * (3.0 * o.x * d.x * d.x) + k[Y3] * (3.0 * o.y * d.y * d.y) + k[Z3] * (3.0 * o.z * d.z * d.z)
           + k[X2Y] * sq_lin_t2(o.x, d.x, o.y, d.y) + k[XY2] * sq_lin_t2(o.y, d.y, o.x, d.x)
           + k[X2Z] * sq_lin_t2(o.x, d.x, o.z, d.z) + k[XZ2] * sq_lin_t2(o.z, d.z, o.x, d.x)
           + k[Y2Z] * sq_lin_t2(o.y, d.y, o.z, d.z) + k[YZ2] * sq_lin_t2(o.z, d.z, o.y, d.y)
           + k[XYZ] * (d.x * d.y * o.z + d.x * o.y * d.z + o.x * d.y * d.z)
           + k[X2] * (d.x * d.x) + k[Y2] * (d.y * d.y) + k[Z2] * (d.z * d.z)
           + k[XY] * (d.x * d.y) + k[XZ] * (d.x * d.z) + k[YZ] * (d.y * d.z);
    p.t1 = k[X3] * (3.0 * o.x * o.x * d.x) + k[Y3] * (3.0 * o.y * o.y * d.y) + k[Z3] * (3.0 * o.z * o.z * d.z)
           + k[X2Y] * sq_lin_t1(o.x, d.x, o.y, d.y) + k[XY2] * sq_lin_t1(o.y, d.y, o.x, d.x)
           + k[X2Z] * sq_lin_t1(o.x, d.x, o.z, d.z) + k[XZ2] * sq_lin_t1(o.z, d.z, o.x, d.x)
           + k[Y2Z] * sq_lin_t1(o.y, d.y, o.z, d.z) + k[YZ2] * sq_lin_t1(o.z, d.z, o.y, d.y)
           + k[XYZ] * (d.x * o.y * o.z + o.x * d.y * o.z + o.x * o.y * d.z)
           + k[X2] * (2.0 * o.x * d.x) + k[Y2] * (2.0 * o.y * d.y) + k[Z2] * (2.0 * o.z * d.z)
           + k[XY] * (o.x * d.y + d.x * o.y) + k[XZ] * (o.x * d.z + d.x * o.z) + k[YZ] * (o.y * d.z + d.y * o.z)
           + k[LX] * d.x + k[LY] * d.y + k[LZ] * d.z;
    p.t0 = k[X3] * (o.x * o.x * o.x) + k[Y3] * (o.y * o.y * o.y) + k[Z3] * (o.z * o.z * o.z)
           + k[X2Y] * (o.x * o.x * o.y) + k[XY2] * (o.x * o.y * o.y)
           + k[X2Z] * (o.x * o.x * o.z) + k[XZ2] * (o.x * o.z * o.z)
           + k[Y2Z] * (o.y * o.y * o.z) + k[YZ2] * (o.y * o.z * o.z)
           + k[XYZ] * (o.x * o.y * o.z)
           + k[X2] * (o.x * o.x) + k[Y2] * (o.y * o.y) + k[Z2] * (o.z * o.z)
           + k[XY] * (o.x * o.y) + k[XZ] * (o.x * o.z) + k[YZ] * (o.y * o.z)
           + k[LX] * o.x + k[LY] * o.y + k[LZ] * o.z + k[CC];
    return p;
}

__device__ __forceinline__ double intersect_cubic(const double *__restrict__ k, const d3 &o, const d3 &d)
{
    return solve_poly(expand_cubic(k, o, d));
}

// surface_impl.h:157-172 on the full 20 coefficients (once per hit)
__device__ __forceinline__ d3 surface_normal(const double *__restrict__ k, const d3 &p)
{
    d3 g;
    g.x = 3.0 * k[X3] * p.x * p.x + 2.0 * k[X2] * p.x + k[LX];
    g.y = 3.0 * k[Y3] * p.y * p.y + 2.0 * k[Y2] * p.y + k[LY];
    g.z = 3.0 * k[Z3] * p.z * p.z + 2.0 * k[Z2] * p.z + k[LZ];
    g.x += 2.0 * p.x * (k[X2Y] * p.y + k[X2Z] * p.z) + p.y * (k[XY2] * p.y + k[XYZ] * p.z + k[XY])
           + p.z * (k[XZ2] * p.z + k[XZ]);
    g.y += 2.0 * p.y * (k[XY2] * p.x + k[Y2Z] * p.z) + p.x * (k[X2Y] * p.x + k[XYZ] * p.z + k[XY])
           + p.z * (k[YZ2] * p.z + k[YZ]);
    g.z += 2.0 * p.z * (k[XZ2] * p.x + k[YZ2] * p.y) + p.x * (k[X2Z] * p.x + k[XYZ] * p.y + k[XZ])
           + p.y * (k[Y2Z] * p.y + k[YZ]);
    return normalize3(g);
}

// Per-class gradients: surface_normal() restricted to the coefficients a class can have.  The terms left out are
// products with coefficients that are exactly 0, i.e. +-0 added to the sums: the gradient — and with it the normal —
// is the same number; only the sign of a component that is exactly zero may differ, which nothing downstream can see
// (x + (+-0) = x, (+-0) * y adds +-0 to a dot product, and the shading factor compares with 0).
//   plane   (only x, y, z, c):  grad F = (x, y, z), the same at every point: normalised once per CTA
//   sphere  (x2 = y2 = z2 = 1): grad F = 2 p + (x, y, z)       — 2.0 * 1.0 * p.x == 2.0 * p.x
//   quadric (degree <= 2)
__device__ __forceinline__ d3 sphere_gradient(const double4 k, const d3 &p)
{
    return d3{2.0 * p.x + k.x, 2.0 * p.y + k.y, 2.0 * p.z + k.z};
}

// q = {x2,y2,z2,xy,xz,yz,x,y,z,c}; association as in surface_normal: (2 k2 p + l) + (p_b k_ab + p_c k_ac)
__device__ __forceinline__ d3 quadric_gradient(const double *__restrict__ q, const d3 &p)
{
    d3 g;
    g.x = 2.0 * q[0] * p.x + q[6];
    g.y = 2.0 * q[1] * p.y + q[7];
    g.z = 2.0 * q[2] * p.z + q[8];
    g.x += p.y * q[3] + p.z * q[4];
    g.y += p.x * q[3] + p.z * q[5];
    g.z += p.x * q[4] + p.y * q[5];
    return g;
}

// light_impl.h:17-27: direction rounded to float and widened again; max_t = 1 (point) or 1e6
__device__ __forceinline__ d3 shadow_dir(const LightDev &l, const d3 &p, double &max_t)
{
    if (l.is_spherical) {
        max_t = 1.0;
        return d3{(double) (float) (l.px - p.x), (double) (float) (l.py - p.y), (double) (float) (l.pz - p.z)};
    }
    max_t = 1e6;
    return d3{(double) (float) l.px, (double) (float) l.py, (double) (float) l.pz};
}

// light_impl.h:29-44, accumulated into acc (update-cpu.cpp:74)
__device__ __forceinline__ void shade_light(const LightDev &l, const d3 &p, const d3 &n, const float3 &obj, float3 &acc)
{
    d3 dir;
    float cr, cg, cb;
    if (l.is_spherical) {
        dir = d3{l.px - p.x, l.py - p.y, l.pz - p.z};
        float denom = 4.0f * RT_PI_F * (float) dot3(dir, dir);
        cr = l.r / denom;
        cg = l.g / denom;
        cb = l.b / denom;
        dir = normalize3(dir);
    }
    else {
        dir = d3{l.px, l.py, l.pz};
        cr = l.r;
        cg = l.g;
        cb = l.b;
    }
    float cosv = (float) dot3(n, dir);
    float lambert = (0.0f < cosv) ? cosv : 0.0f;
    acc.x += obj.x / RT_PI_F * cr * lambert;
    acc.y += obj.y / RT_PI_F * cg * lambert;
    acc.z += obj.z / RT_PI_F * cb * lambert;
}

// The same with the object's colour already divided by pi (SceneView::material_pi, divided once at upload with the
// same IEEE FP32 division): three divisions fewer per (hit, light).  A colour channel of exactly 0 skips the division
// of the point-light term too: 0 / denom takes the division's slow path on the GPU, and for 0 < denom < inf the
// quotient is the signed zero itself.
__device__ __forceinline__ void shade_light_pi(const LightDev &l, const d3 &p, const d3 &n, const float3 &obj_pi, float3 &acc)
{
    d3 dir;
    float cr, cg, cb;
    if (l.is_spherical) {
        dir = d3{l.px - p.x, l.py - p.y, l.pz - p.z};
        float denom = 4.0f * RT_PI_F * (float) dot3(dir, dir);
        const bool plain = denom > 0.0f && denom < CUDART_INF_F;
        cr = (plain && l.r == 0.0f) ? l.r : l.r / denom;
        cg = (plain && l.g == 0.0f) ? l.g : l.g / denom;
        cb = (plain && l.b == 0.0f) ? l.b : l.b / denom;
        dir = normalize3(dir);
    }
    else {
        dir = d3{l.px, l.py, l.pz};
        cr = l.r;
        cg = l.g;
        cb = l.b;
    }
    float cosv = (float) dot3(n, dir);
    float lambert = (0.0f < cosv) ? cosv : 0.0f;
    acc.x += obj_pi.x * cr * lambert;
    acc.y += obj_pi.y * cg * lambert;
    acc.z += obj_pi.z * cb * lambert;
}

// The factor surface_color() multiplies a light's colour with (light_impl.h:43): max(0, (float) dot(n, dir)).  When it
// is 0 and the colour terms are finite the light adds exactly +-0 to the pixel whether or not it is occluded, so its
// shadow ray cannot influence the frame.  Returns true only when that is CERTAIN (false = trace the ray as usual):
//   directional: the reference's own (float) dot(n, p) is not positive;
//   point light: the reference normalises dir = p - P first (a positive factor), so the sign is that of
//       u = dot(n, dir) unless u is within rounding of 0 — required: u below -1e-14 times the sum of |terms|, and
//       |dir| > 1e-7 so that colour / (4 pi |dir|^2) stays finite.
// The caller has checked once per scene that every light and object colour is finite and moderate (cull_safe).
__device__ __forceinline__ bool light_cannot_contribute(const LightDev &l, const d3 &p, const d3 &n)
{
    if (l.is_spherical) {
        const double dx = l.px - p.x, dy = l.py - p.y, dz = l.pz - p.z;
        const double a = n.x * dx, b = n.y * dy, c = n.z * dz;
        const double u = (a + b) + c;
        return u < -1e-14 * ((fabs(a) + fabs(b)) + fabs(c)) && (fabs(dx) + fabs(dy)) + fabs(dz) > 1e-7;
    }
    const float cosv = (float) dot3(n, d3{l.px, l.py, l.pz});
    return !(0.0f < cosv);
}

// light_impl.h:46-49
__device__ __forceinline__ d3 reflect(const d3 &dir, const d3 &n)
{
    double twice = 2.0 * dot3(dir, n);
    return d3{dir.x - twice * n.x, dir.y - twice * n.y, dir.z - twice * n.z};
}

// GLM mat4 * vec4 = (m0*x + m1*y) + (m2*z + m3*w), xyz rows only
__device__ __forceinline__ d3 mat_point(const double *__restrict__ m, double x, double y, double z, double w)
{
    return d3{(m[0] * x + m[4] * y) + (m[8] * z + m[12] * w),
              (m[1] * x + m[5] * y) + (m[9] * z + m[13] * w),
              (m[2] * x + m[6] * y) + (m[10] * z + m[14] * w)};
}

// update-cpu.cpp:84-89
__device__ __forceinline__ d3 primary_dir(const FrameParams &f, const d3 &origin, int px, int py)
{
    double ndc_x = (px + 0.5) / f.width;
    double ndc_y = (py + 0.5) / f.height;
    double cam_x = (2.0 * ndc_x - 1.0) * f.aspect * f.tan_half_fov;
    double cam_y = (2.0 * ndc_y - 1.0) * f.tan_half_fov;
    d3 v = mat_point(f.m, cam_x, cam_y, 1.0, 1.0);
    return normalize3(d3{v.x - origin.x, v.y - origin.y, v.z - origin.z});
}

// The same in two steps, for kernels that compute the per-column / per-row factors once per tile: cam_x depends only
// on the pixel column, cam_y only on the image row (identical expressions, hence identical bits).
__device__ __forceinline__ double primary_cam_x(const FrameParams &f, int px)
{
    double ndc_x = (px + 0.5) / f.width;
    return (2.0 * ndc_x - 1.0) * f.aspect * f.tan_half_fov;
}
__device__ __forceinline__ double primary_cam_y(const FrameParams &f, int py)
{
    double ndc_y = (py + 0.5) / f.height;
    return (2.0 * ndc_y - 1.0) * f.tan_half_fov;
}
__device__ __forceinline__ d3 primary_dir_cam(const FrameParams &f, const d3 &origin, double cam_x, double cam_y)
{
    d3 v = mat_point(f.m, cam_x, cam_y, 1.0, 1.0);
    return normalize3(d3{v.x - origin.x, v.y - origin.y, v.z - origin.z});
}

// update-cuda.cu:149-156: (int)(c*255 + 0.5) per channel, alpha 255
__device__ __forceinline__ uchar4 pack_rgba8(float r, float g, float b)
{
    uchar4 px;
    px.x = (unsigned char) (int) (r * 255.0f + 0.5f);
    px.y = (unsigned char) (int) (g * 255.0f + 0.5f);
    px.z = (unsigned char) (int) (b * 255.0f + 0.5f);
    px.w = 255;
    return px;
}

} // namespace rt
#endif
