// rt_cubic.cuh — mask-driven exact expansion of F(o + t d) for cubic-class objects.
//
// The reference evaluates all 20 monomial groups for every object (include/surface_impl.h:46-103) although
// the bundled cubics use 4 (Fermat, monkey saddle), 8 (ding-dong), 9 (Cayley) or 19 (Clebsch) of them.
// rt_b200_upload_scene records which coefficients are exactly zero in a 20-bit mask; all lanes of a warp
// test the same object, so the `mask & bit` branches are warp-uniform.  A skipped group would have added
// exact zeros, so t3..t0 round exactly as in the reference (sums start from +0 and run left to right in
// coefficient order, the order of the reference's expressions).
#ifndef RT_CUBIC_CUH
#define RT_CUBIC_CUH

#include "rt_math.cuh"

namespace rt {

#define RT_BIT(c) (1u << (c))

__device__ __forceinline__ Poly expand_cubic_masked(const double *__restrict__ k, unsigned mask, const d3 &o, const d3 &d)
{
    double t3 = 0.0, t2 = 0.0, t1 = 0.0, t0 = 0.0;
    // ---- pure cubes: c (a0 + t a)^3 ----------------------------------------------------------------
    if (mask & RT_BIT(X3)) {
        const double c = k[X3];
        t3 += c * (d.x * d.x * d.x);
        t2 += c * (3.0 * o.x * d.x * d.x);
        t1 += c * (3.0 * o.x * o.x * d.x);
        t0 += c * (o.x * o.x * o.x);
    }
    if (mask & RT_BIT(Y3)) {
        const double c = k[Y3];
        t3 += c * (d.y * d.y * d.y);
        t2 += c * (3.0 * o.y * d.y * d.y);
        t1 += c * (3.0 * o.y * o.y * d.y);
        t0 += c * (o.y * o.y * o.y);
    }
    if (mask & RT_BIT(Z3)) {
        const double c = k[Z3];
        t3 += c * (d.z * d.z * d.z);
        t2 += c * (3.0 * o.z * d.z * d.z);
        t1 += c * (3.0 * o.z * o.z * d.z);
        t0 += c * (o.z * o.z * o.z);
    }
    // ---- c (a0 + t a)^2 (b0 + t b) -------------------------------------------------------------------
#define RT_SQ_LIN(BIT, A, B, T3EXPR, T0EXPR)                                   \
    if (mask & RT_BIT(BIT)) {                                                  \
        const double c = k[BIT];                                               \
        t3 += c * (T3EXPR);                                                    \
        t2 += c * sq_lin_t2(o.A, d.A, o.B, d.B);                               \
        t1 += c * sq_lin_t1(o.A, d.A, o.B, d.B);                               \
        t0 += c * (T0EXPR);                                                    \
    }
    RT_SQ_LIN(X2Y, x, y, d.x * d.x * d.y, o.x * o.x * o.y)
    RT_SQ_LIN(XY2, y, x, d.x * d.y * d.y, o.x * o.y * o.y)
    RT_SQ_LIN(X2Z, x, z, d.x * d.x * d.z, o.x * o.x * o.z)
    RT_SQ_LIN(XZ2, z, x, d.x * d.z * d.z, o.x * o.z * o.z)
    RT_SQ_LIN(Y2Z, y, z, d.y * d.y * d.z, o.y * o.y * o.z)
    RT_SQ_LIN(YZ2, z, y, d.y * d.z * d.z, o.y * o.z * o.z)
#undef RT_SQ_LIN
    if (mask & RT_BIT(XYZ)) {
        const double c = k[XYZ];
        t3 += c * (d.x * d.y * d.z);
        t2 += c * (d.x * d.y * o.z + d.x * o.y * d.z + o.x * d.y * d.z);
        t1 += c * (d.x * o.y * o.z + o.x * d.y * o.z + o.x * o.y * d.z);
        t0 += c * (o.x * o.y * o.z);
    }
    // ---- squares and cross terms ---------------------------------------------------------------------
#define RT_SQUARE(BIT, A)                                                      \
    if (mask & RT_BIT(BIT)) {                                                  \
        const double c = k[BIT];                                               \
        t2 += c * (d.A * d.A);                                                 \
        t1 += c * (2.0 * o.A * d.A);                                           \
        t0 += c * (o.A * o.A);                                                 \
    }
    RT_SQUARE(X2, x)
    RT_SQUARE(Y2, y)
    RT_SQUARE(Z2, z)
#undef RT_SQUARE
#define RT_CROSS(BIT, A, B)                                                    \
    if (mask & RT_BIT(BIT)) {                                                  \
        const double c = k[BIT];                                               \
        t2 += c * (d.A * d.B);                                                 \
        t1 += c * (o.A * d.B + d.A * o.B);                                     \
        t0 += c * (o.A * o.B);                                                 \
    }
    RT_CROSS(XY, x, y)
    RT_CROSS(XZ, x, z)
    RT_CROSS(YZ, y, z)
#undef RT_CROSS
    // ---- linear and constant ---------------------------------------------------------------------------
    if (mask & RT_BIT(LX)) { const double c = k[LX]; t1 += c * d.x; t0 += c * o.x; }
    if (mask & RT_BIT(LY)) { const double c = k[LY]; t1 += c * d.y; t0 += c * o.y; }
    if (mask & RT_BIT(LZ)) { const double c = k[LZ]; t1 += c * d.z; t0 += c * o.z; }
    if (mask & RT_BIT(CC)) t0 += k[CC];
    Poly p;
    p.t3 = t3; p.t2 = t2; p.t1 = t1; p.t0 = t0;
    return p;
}

__device__ __forceinline__ double intersect_cubic_masked(const double *__restrict__ k, unsigned mask, const d3 &o, const d3 &d)
{
    return solve_poly(expand_cubic_masked(k, mask, o, d));
}

} // namespace rt
#endif
