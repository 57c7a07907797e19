// rt_cubic_filter.cuh — conservative FP32 decision "does this ray hit this cubic-class surface inside (EPS, hi)?"
// in front of the exact FP64 test (rt_cubic.cuh + solve_poly).
//
// The reference answers the question by expanding F(o + t d) = t3 t^3 + t2 t^2 + t1 t + t0 and solving the cubic in
// closed form (include/surface_impl.h:46-137): ~290 FP64 operations for the expansion plus cbrt/acos/cos.  Whatever
// branch it takes, the root it returns lies in [EPS, hi) exactly when the cubic has a real root there: Cardano
// (:117-121) returns the only real root, the trigonometric branch (:123-136) the smallest root >= EPS (else the
// largest root, which is then < EPS).  So a test that PROVES "no real root in a slightly wider interval" (MISS) or
// "a sign change strictly inside a slightly narrower interval" (HIT) predicts the reference's decision without
// solving anything; only when neither can be proven (CV_UNKNOWN) does the exact FP64 path run.  Results are therefore
// identical to running the exact test everywhere; the filter only moves work from the FP64 pipe to FP32 FMAs.
//
// How the proof goes (u = 2^-24):
//   1. T_i = t_i evaluated in FP32 with FMAs from an FP32 table of the coefficients and the FP32-rounded origin and
//      direction (expand_cubic_f32<false>).  |T_i - t_i| <= E_i = 32 u B_i, where B_i is the SAME evaluation scheme
//      run over absolute values (expand_cubic_f32<true>): it bounds the sum of the absolute values of everything
//      that is added up, so (running-error argument) each of the <= 16 roundings on the longest path and the <= 4
//      rounded inputs of a term contribute at most u B_i.  Exact zeros stay exact: a direction along an axis makes
//      T3 = B3 = 0, which is what lets the degree switch below be decided.
//   2. Degree switches (surface_impl.h:106,138,150): a branch is certain when |T_i| - E_i > 1e-7 or |T_i| + E_i < 1e-7;
//      the lower-degree branches drop the higher coefficients exactly as the reference does.  For the quadratic
//      branch with t2 < 0 the reference tries the LARGER root first, so only a downward crossing proves a hit.
//   3. The reference's own root error is ~64 u64 M with M = 4 max(|t2/t3|, sqrt|t1/t3|, cbrt|t0/t3|) (more in the
//      trigonometric branch near delta = 0, which 3b excludes); a per-ray bound of
//      that error (3c in the code: cancellation in Cardano's r - sqrt(delta) when q is small, acos near +-1 in the
//      trigonometric branch) must stay below EPS / 4 = 2.5e-8, inside the slack the interval ends get below (factor 2
//      around EPS, 2e-6 relative around hi).
//   3b. The reference's choice between Cardano and the trigonometric formula must not hinge on FP64 rounding: the
//      sign of the normalised discriminant has to be certifiable in FP32 (see the code).
//   4. With s = sign(T0), for all t >= 0:  s P(t) >= g(t) and s P(t) <= h(t), where g, h are the cubics with
//      coefficients s T_i -/+ E_i — EXACTLY known polynomials, so their extrema over an interval sit at its ends
//      and at their own critical points, which a compensated discriminant (Kahan) locates to a few ulps; a
//      critical point misplaced by eta changes the sampled extremum only by O(eta^2).  Every sample carries its
//      FP32 Horner error 32 u A(x), A = Horner on absolute values.
//      MISS: g > 0 on [EPS/2, 2 EPS] (uniform bound), on [hi (1 -/+ 2e-6)] (value at hi minus the variation bound)
//            and at every critical point of g in between  =>  P has no root in [EPS/2, hi (1 + 2e-6)].
//      HIT:  g > 0 on [EPS/2, 2 EPS] and h < 0 on the hi range or at some sampled point of (2 EPS, hi (1 - 2e-6))
//            =>  P changes sign strictly inside, so a real root of odd multiplicity lies in (2 EPS, hi (1 - 2e-6)).
//   Anything non-finite makes a comparison false and ends in UNKNOWN.
// Not covered (measure-zero): a reference-side NaN from acos of 1 + ulp exactly at a double root (delta == 0-).
//
// The header is plain C++ when compiled without nvcc: tests/test_cubic_filter.py builds it for the host and checks
// the verdicts against the oracle on the bundled scenes' primary and shadow rays.
#ifndef RT_CUBIC_FILTER_CUH
#define RT_CUBIC_FILTER_CUH

#include <math.h>

#ifdef __CUDACC__
#define RT_HD __host__ __device__ __forceinline__
#else
#define RT_HD static inline
#endif

namespace rt {

enum CubicVerdict { CV_UNKNOWN = 0, CV_MISS = 1, CV_HIT = 2 };
#ifdef RT_CF_DEBUG
#define RT_CF_UNK(n) (-(n))   /* host debugging: which test gave up */
#else
#define RT_CF_UNK(n) CV_UNKNOWN
#endif

#define RT_CF_FLOATS 28          /* per-object FP32 table, see make_cubicf */
#define RT_CF_U 5.9604645e-8f    /* 2^-24 */
#define RT_CF_G (32.0f * RT_CF_U * 1.001f)

// Per-object FP32 table.  With C, Q, L the homogeneous cubic, quadratic and linear parts of F:
//   F(o + t d) = [C(o) + Q(o) + L(o) + c] + t [grad C(o).d + grad Q(o).d + L(d)] + t^2 [grad C(d).o + Q(d)] + t^3 C(d)
// and C(v) = grad C(v).v / 3 (Euler), so everything follows from grad C at o and at d:
//   [0..5]   d/dx C : coefficients of xx, xy, yy, xz, zz, yz   = 3 x3, 2 x2y, xy2, 2 x2z, xz2, xyz
//   [6..11]  d/dy C :                                            x2y, 2 xy2, 3 y3, xyz, yz2, 2 y2z
//   [12..17] d/dz C :                                            x2z, xyz, y2z, 2 xz2, 3 z3, 2 yz2
//   [18..23] Q      : x2, y2, z2, xy, xz, yz      [24..26] L : x, y, z      [27] c
RT_HD void make_cubicf(const double *k, float *out)
{
    const double x3 = k[0], y3 = k[1], z3 = k[2], x2y = k[3], xy2 = k[4], x2z = k[5], xz2 = k[6], y2z = k[7], yz2 = k[8], xyz = k[9];
    out[0] = (float) (3.0 * x3); out[1] = (float) (2.0 * x2y); out[2] = (float) xy2;
    out[3] = (float) (2.0 * x2z); out[4] = (float) xz2; out[5] = (float) xyz;
    out[6] = (float) x2y; out[7] = (float) (2.0 * xy2); out[8] = (float) (3.0 * y3);
    out[9] = (float) xyz; out[10] = (float) yz2; out[11] = (float) (2.0 * y2z);
    out[12] = (float) x2z; out[13] = (float) xyz; out[14] = (float) y2z;
    out[15] = (float) (2.0 * xz2); out[16] = (float) (3.0 * z3); out[17] = (float) (2.0 * yz2);
    for (int i = 10; i < 20; i++) out[18 + i - 10] = (float) k[i];
}

struct RayC {
    float ox, oy, oz, dx, dy, dz;   // origin and direction rounded to FP32
};

RT_HD RayC make_rayc(double ox, double oy, double oz, double dx, double dy, double dz)
{
    RayC r;
    r.ox = (float) ox; r.oy = (float) oy; r.oz = (float) oz;
    r.dx = (float) dx; r.dy = (float) dy; r.dz = (float) dz;
    return r;
}

// FP32 expansion of F(o + t d) from the table above, in two halves that can be cached separately: everything that
// depends on the ORIGIN (shared by all shadow rays of one hit point, and by all primary rays of a frame) and
// everything that depends on the DIRECTION (shared by all shadow rays towards one directional light).
// The a* / b* members are the same sums over |table| and |components|: upper bounds of the sums of the terms'
// absolute values, hence (times the rounding count) of the rounding errors of the signed values.
#if defined(__CUDACC__)
#define RT_CF_ALIGN __align__(16)   /* shared-memory copies move as 128-bit words */
#else
#define RT_CF_ALIGN
#endif
struct RT_CF_ALIGN CubicO {
    float ox, oy, oz;       // origin rounded to FP32
    float hx, hy, hz;       // grad C(o) + grad Q(o) + l : t1 = h . d
    float t0;               // C(o) + Q(o) + L(o) + c
    float ahx, ahy, ahz, b0;
    float pad0;
};
struct RT_CF_ALIGN CubicD {
    float dx, dy, dz;       // direction rounded to FP32
    float gx, gy, gz;       // grad C(d) : t2 = g . o + qd
    float qd, t3;           // Q(d), C(d)
    float agx, agy, agz, aqd, b3;
    float pad0, pad1, pad2;
};

#define RT_F_GRAD(A, B, xx, xy, yy, xz, zz, yz)                                                                        \
    fmaf(A(k[B]), xx, fmaf(A(k[B + 1]), xy, fmaf(A(k[B + 2]), yy, fmaf(A(k[B + 3]), xz, fmaf(A(k[B + 4]), zz, A(k[B + 5]) * yz)))))
#define RT_F_ID(x) (x)
#define RT_CF_THIRD 0.33333334f /* rounded up: keeps the absolute sums upper bounds */

RT_HD CubicO make_cubic_o(const float *k, float ox, float oy, float oz)
{
    CubicO r;
    r.pad0 = 0.0f;
    r.ox = ox; r.oy = oy; r.oz = oz;
    const float xx = ox * ox, xy = ox * oy, yy = oy * oy, xz = ox * oz, zz = oz * oz, yz = oy * oz;
    const float gx = RT_F_GRAD(RT_F_ID, 0, xx, xy, yy, xz, zz, yz);   // grad C(o)
    const float gy = RT_F_GRAD(RT_F_ID, 6, xx, xy, yy, xz, zz, yz);
    const float gz = RT_F_GRAD(RT_F_ID, 12, xx, xy, yy, xz, zz, yz);
    const float q0 = k[18], q1 = k[19], q2 = k[20], q3 = k[21], q4 = k[22], q5 = k[23];
    // grad Q(o) = (2 q0 ox + q3 oy + q4 oz, q3 ox + 2 q1 oy + q5 oz, q4 ox + q5 oy + 2 q2 oz)
    r.hx = fmaf(2.0f * q0, ox, fmaf(q3, oy, fmaf(q4, oz, gx + k[24])));
    r.hy = fmaf(q3, ox, fmaf(2.0f * q1, oy, fmaf(q5, oz, gy + k[25])));
    r.hz = fmaf(q4, ox, fmaf(q5, oy, fmaf(2.0f * q2, oz, gz + k[26])));
    const float qo = fmaf(q0, xx, fmaf(q1, yy, fmaf(q2, zz, fmaf(q3, xy, fmaf(q4, xz, q5 * yz)))));
    const float co = RT_CF_THIRD * fmaf(gx, ox, fmaf(gy, oy, gz * oz));   // Euler: C(o) = grad C(o) . o / 3
    r.t0 = fmaf(k[24], ox, fmaf(k[25], oy, fmaf(k[26], oz, (co + qo) + k[27])));
    // the same over absolute values (|xy| = |x||y|: the products are reused)
    const float ax = fabsf(ox), ay = fabsf(oy), az = fabsf(oz);
    const float axy = fabsf(xy), axz = fabsf(xz), ayz = fabsf(yz);
    const float agx = RT_F_GRAD(fabsf, 0, xx, axy, yy, axz, zz, ayz);
    const float agy = RT_F_GRAD(fabsf, 6, xx, axy, yy, axz, zz, ayz);
    const float agz = RT_F_GRAD(fabsf, 12, xx, axy, yy, axz, zz, ayz);
    const float a0 = fabsf(q0), a1 = fabsf(q1), a2 = fabsf(q2), a3 = fabsf(q3), a4 = fabsf(q4), a5 = fabsf(q5);
    r.ahx = fmaf(2.0f * a0, ax, fmaf(a3, ay, fmaf(a4, az, agx + fabsf(k[24]))));
    r.ahy = fmaf(a3, ax, fmaf(2.0f * a1, ay, fmaf(a5, az, agy + fabsf(k[25]))));
    r.ahz = fmaf(a4, ax, fmaf(a5, ay, fmaf(2.0f * a2, az, agz + fabsf(k[26]))));
    const float aqo = fmaf(a0, xx, fmaf(a1, yy, fmaf(a2, zz, fmaf(a3, axy, fmaf(a4, axz, a5 * ayz)))));
    const float aco = RT_CF_THIRD * fmaf(agx, ax, fmaf(agy, ay, agz * az));
    r.b0 = fmaf(fabsf(k[24]), ax, fmaf(fabsf(k[25]), ay, fmaf(fabsf(k[26]), az, (aco + aqo) + fabsf(k[27]))));
    return r;
}

RT_HD CubicD make_cubic_d(const float *k, float dx, float dy, float dz)
{
    CubicD r;
    r.pad0 = r.pad1 = r.pad2 = 0.0f;
    r.dx = dx; r.dy = dy; r.dz = dz;
    const float xx = dx * dx, xy = dx * dy, yy = dy * dy, xz = dx * dz, zz = dz * dz, yz = dy * dz;
    r.gx = RT_F_GRAD(RT_F_ID, 0, xx, xy, yy, xz, zz, yz);   // grad C(d)
    r.gy = RT_F_GRAD(RT_F_ID, 6, xx, xy, yy, xz, zz, yz);
    r.gz = RT_F_GRAD(RT_F_ID, 12, xx, xy, yy, xz, zz, yz);
    r.qd = fmaf(k[18], xx, fmaf(k[19], yy, fmaf(k[20], zz, fmaf(k[21], xy, fmaf(k[22], xz, k[23] * yz)))));
    r.t3 = RT_CF_THIRD * fmaf(r.gx, dx, fmaf(r.gy, dy, r.gz * dz));
    const float axy = fabsf(xy), axz = fabsf(xz), ayz = fabsf(yz);
    r.agx = RT_F_GRAD(fabsf, 0, xx, axy, yy, axz, zz, ayz);
    r.agy = RT_F_GRAD(fabsf, 6, xx, axy, yy, axz, zz, ayz);
    r.agz = RT_F_GRAD(fabsf, 12, xx, axy, yy, axz, zz, ayz);
    r.aqd = fmaf(fabsf(k[18]), xx, fmaf(fabsf(k[19]), yy, fmaf(fabsf(k[20]), zz,
            fmaf(fabsf(k[21]), axy, fmaf(fabsf(k[22]), axz, fabsf(k[23]) * ayz)))));
    r.b3 = RT_CF_THIRD * fmaf(r.agx, fabsf(dx), fmaf(r.agy, fabsf(dy), r.agz * fabsf(dz)));
    return r;
}
#undef RT_F_GRAD
#undef RT_F_ID

// Value, error bound and absolute Horner sum of the cubic at one x >= 0:
//   lower bound of s P(x) = s T(x) - E(x) - ev (|T|(x) + E(x)),   upper bound = s T(x) + E(x) + ev (|T|(x) + E(x))
struct CfSample {
    float lo, up;
};
RT_HD CfSample cf_sample(float s, const float T[4], const float E[4], float x)
{
    const float v = fmaf(fmaf(fmaf(T[3], x, T[2]), x, T[1]), x, T[0]);
    const float e = fmaf(fmaf(fmaf(E[3], x, E[2]), x, E[1]), x, E[0]);
    const float a = fmaf(fmaf(fmaf(fabsf(T[3]), x, fabsf(T[2])), x, fabsf(T[1])), x, fabsf(T[0]));
    const float m = fmaf(32.0f * RT_CF_U, a + e, e);
    CfSample r;
    r.lo = s * v - m;
    r.up = s * v + m;
    return r;
}

// Fast approximate quotient / square root for LOCATING critical points (their position enters the proof only to
// second order); plain operators on the host.
#if defined(__CUDA_ARCH__)
#define RT_CF_DIV(x, y) __fdividef((x), (y))
RT_HD float cf_sqrt(float x)
{
    float r;
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
RT_HD float cf_cbrt(float x) { return exp2f(__log2f(x) * 0.33333334f); }   // x > 0; ~1e-6 relative, used in bounds only
#else
#define RT_CF_DIV(x, y) ((x) / (y))
RT_HD float cf_sqrt(float x) { return sqrtf(x); }
RT_HD float cf_cbrt(float x) { return cbrtf(x); }
#endif

// The decision from the FP32 coefficients T[i] (index = power of t) and the sums of absolute values B[i].
// hi = the caller's max_t (1 for point lights, 1e6 for directional lights and for nearest-hit queries).
// Written with selects instead of early exits: on the GPU almost every lane runs the whole function anyway.
RT_HD int cubic_verdict_tb(float T[4], const float B[4], float hi, bool want_hit)
{
    float E[4] = {RT_CF_G * B[0], RT_CF_G * B[1], RT_CF_G * B[2], RT_CF_G * B[3]};
    // 2. which branch of surface_impl.h:106-154 the reference takes (its switches compare |t_i| with 1e-7):
    //    c_i = "|t_i| > 1e-7 for sure", z_i = "|t_i| <= 1e-7 for sure"
    const bool c3 = fabsf(T[3]) - E[3] > 1.0001e-7f, z3 = fabsf(T[3]) + E[3] < 0.9999e-7f;
    const bool c2 = fabsf(T[2]) - E[2] > 1.0001e-7f, z2 = fabsf(T[2]) + E[2] < 0.9999e-7f;
    const bool c1 = fabsf(T[1]) - E[1] > 1.0001e-7f, z1 = fabsf(T[1]) + E[1] < 0.9999e-7f;
    bool ok = c3 || (z3 && (c2 || (z2 && (c1 || z1))));
    if (z3 && z2 && z1) return CV_MISS;                 // :153 returns -1
    const int deg = c3 ? 3 : (c2 ? 2 : 1);
    if (deg < 3) { T[3] = 0.0f; E[3] = 0.0f; }          // the reference drops the higher coefficients altogether
    if (deg < 2) { T[2] = 0.0f; E[2] = 0.0f; }
    // 3. quadratic branch (:138-149): |t1| < 1e4 |t2| and |t0| < 1e8 |t2| keep the reference's own root error (u64 times the
    //    intermediates of the quadratic formula) far below EPS / 2
    if (deg == 2) {
        const float lead = fabsf(T[2]) - E[2];
        ok = ok && fabsf(T[1]) + E[1] < 1e4f * lead && fabsf(T[0]) + E[0] < 1e8f * lead;
    }
    // 3b. cubic branch: the reference chooses between Cardano and the trigonometric formula by the sign of
    //    delta = q^3 + r^2 of the NORMALISED cubic (surface_impl.h:112-116).  Two roots that are close relative to the
    //    largest root magnitude (e.g. a nearly quadratic ray: third root far away at -t2/t3) make that sign a matter
    //    of FP64 rounding — the reference may then return the far root and ignore a crossing that P really has, or the
    //    other way round.  Nothing can be predicted there, so a verdict needs the sign of delta CERTIFIED in FP32:
    //    |delta| > 2 e_delta + 1e-12 (Qs^3 + Rs^2), with first-order error propagation through a_i = T_i / T_3.
    float eroot = 1e-9f;   // bound of |reference's computed root - exact root of its polynomial|; degree <= 2: see 3.
    if (deg == 3) {
        const float inv = RT_CF_DIV(1.0f, T[3]);            // a few ulps: covered by the 4 u terms below
        const float x3 = E[3] * fabsf(inv);                 // relative error of T3; 1 / (1 - x) <= 1 + 1.34 x for x <= 1/4
        const float ainv = fabsf(inv) * fmaf(1.34f, x3, 1.0002f);
        const float a2 = T[2] * inv, a1 = T[1] * inv, a0 = T[0] * inv;
        const float e2a = fmaf(fmaf(fabsf(a2), E[3], E[2]), ainv, 4.0f * RT_CF_U * fabsf(a2));
        const float e1a = fmaf(fmaf(fabsf(a1), E[3], E[1]), ainv, 4.0f * RT_CF_U * fabsf(a1));
        const float e0a = fmaf(fmaf(fabsf(a0), E[3], E[0]), ainv, 4.0f * RT_CF_U * fabsf(a0));
        const float A2 = fabsf(a2) + e2a, A1 = fabsf(a1) + e1a;   // upper bounds of the true |a_i|
        const float a22 = a2 * a2;
        const float q = fmaf(3.0f, a1, -a22) * (1.0f / 9.0f);
        const float qs = fmaf(3.0f, fabsf(a1), a22) * (1.0f / 9.0f);
        const float eq = fmaf(fmaf(2.0f * A2, e2a, 3.0f * e1a), 1.0f / 9.0f, 8.0f * RT_CF_U * qs);
        const float a21 = a2 * a1;
        const float r = fmaf(-2.0f * a22, a2, fmaf(9.0f, a21, -27.0f * a0)) * (1.0f / 54.0f);
        const float rs = fmaf(2.0f * a22, fabsf(a2), fmaf(9.0f, fabsf(a21), 27.0f * fabsf(a0))) * (1.0f / 54.0f);
        const float er = fmaf(fmaf(6.0f * A2 * A2, e2a, fmaf(9.0f, fmaf(A2, e1a, A1 * e2a), 27.0f * e0a)), 1.0f / 54.0f,
                              8.0f * RT_CF_U * rs);
        const float delta = fmaf(q * q, q, r * r);
        const float scale = fmaf(qs * qs, qs, rs * rs);
        const float Q = fabsf(q) + eq, R = fabsf(r) + er;
        const float ed = fmaf(3.0f * Q * Q, eq, fmaf(2.0f * R, er, 8.0f * RT_CF_U * scale));
        ok = ok && x3 < 0.25f && fabsf(delta) > fmaf(2.0f, ed, 1e-12f * scale) && scale < 1e37f;
        // 3c. how far the reference's computed root can be from the exact root of its own cubic (KU = 32 u64):
        //   Cardano: t = cbrt(r + sd) + cbrt(r - sd) - a2/3, sd = sqrt(delta).  r +- sd carry an absolute error
        //     dx = KU (rs + scale / sd); the smaller one, xm = |q|^3 / (|r| + sd), suffers cancellation when q is small,
        //     and cbrt amplifies by 1 / (3 x^(2/3)) — or, when dx is not small against xm, by nothing better than
        //     2 cbrt(xm + dx).
        //   trigonometric: t = c cos(theta + 2 k pi / 3) - a2/3, c = 2 sqrt(-q), theta = acos(v) / 3, v = r / sqrt(-q^3):
        //     dv = KU (rs / |q|^1.5 + 2 qs / |q|), dtheta = 2 dv / (3 sqrt(1 - v^2)), 1 - v^2 = |delta| / |q|^3
        //     (first order; required: dv < (1 - v^2) / 4).
        //   Both plus KU times the magnitudes added up.  The bound becomes the slack of the interval ends (4.); a
        //   verdict needs it below 1e-3.
        const float KU = 32.0f * 1.1102230e-16f;
        const float aq = fabsf(q), ar = fabsf(r), a2third = fabsf(a2) * (1.0f / 3.0f);
        const float sd = cf_sqrt(fabsf(delta));
        if (delta > 0.0f) {
            const float xM = ar + sd;
            const float xm = RT_CF_DIV(aq * aq * aq, xM);
            const float dx = KU * (rs + RT_CF_DIV(scale, sd));
            const float cM = cf_cbrt(xM);
            const float eS = RT_CF_DIV(dx, 3.0f * cM * cM);
            float eT, cm;
            if (dx < 0.5f * xm) {
                cm = cf_cbrt(0.5f * xm);
                eT = RT_CF_DIV(dx, 3.0f * cm * cm);
            }
            else {
                cm = cf_cbrt(xm + dx);
                eT = 2.0f * cm;
            }
            eroot = 1.01f * (eS + eT) + KU * (cM + 1.3f * cm + a2third);
        }
        else {
            const float sq = cf_sqrt(aq), c = 2.0f * sq, q15 = aq * sq;
            const float dv = KU * (RT_CF_DIV(rs, q15) + 2.0f * RT_CF_DIV(qs, aq));
            const float one_v2 = RT_CF_DIV(fabsf(delta), q15 * q15);
            ok = ok && dv < 0.25f * one_v2;
            eroot = 1.01f * c * RT_CF_DIV(2.0f * dv, 3.0f * cf_sqrt(one_v2)) + KU * (c + a2third);
        }
        ok = ok && eroot < 1e-3f;
    }
    // 4. sign-normalised lower polynomial g = s T - E <= s P
    const float s = T[0] > 0.0f ? 1.0f : -1.0f;
    // quadratic branch with t2 < 0 (:138-149) returns the LARGER root first: only a downward crossing proves a hit
    const bool hit_ok = want_hit && !(deg == 2 && T[2] < 0.0f && s < 0.0f);
    const float g0 = s * T[0] - E[0], g1 = s * T[1] - E[1], g2 = s * T[2] - E[2], g3 = s * T[3] - E[3];
    // uniform lower bound of s P over |t| <= tau, tau = EPS + eroot (at least 2 EPS): no root of P within eroot of EPS on
    // either side — valid for negative t too, hence with |T_i| + E_i
    const float e2 = fmaxf(2e-7f, 1.01f * (1e-7f + eroot));
    const float la = g0 - 1.01f * e2 * fmaf(fmaf(fabsf(T[3]) + E[3], e2, fabsf(T[2]) + E[2]), e2, fabsf(T[1]) + E[1]) -
                     4.0f * RT_CF_U * fabsf(g0);
    ok = ok && la > 0.0f;
    // critical points of g: 3 g3 x^2 + 2 g2 x + g1 = 0 (degree 3), 2 g2 x + g1 = 0 (degree 2)
    float x1 = -1.0f, x2 = -1.0f;   // "not in range"
    {
        // discriminant / 4 = g2^2 - (3 g3) g1, compensated (Kahan); degree 2 has g3 = 0: disc = g2^2, q = -2 g2, x2 = -g1 / (2 g2)
        const float a = 3.0f * g3;
        const float pp = g2 * g2, qq = a * g1;
        const float disc = (pp - qq) + (fmaf(g2, g2, -pp) - fmaf(a, g1, -qq));
        const float sd = cf_sqrt(fmaxf(disc, 0.0f));
        const float q = -(g2 + copysignf(sd, g2));
        const bool two = disc > 0.0f && deg >= 2;
        ok = ok && (disc == disc) && fabsf(q) < 1e30f && fabsf(a) < 1e30f;   // NaN / range of the fast quotient
        if (two) {
            x2 = RT_CF_DIV(g1, q);
            if (deg == 3) x1 = RT_CF_DIV(q, a);
        }
        ok = ok && fabsf(x1) < 3e38f && fabsf(x2) < 3e38f;
    }
    if (!ok) return RT_CF_UNK(1);
    const float b_hi = fmaf(hi, 1.0f + 2e-6f, eroot), b_lo = fmaf(hi, 1.0f - 2e-6f, -eroot);
    // bounds over [b_lo, b_hi]: sample at hi, widened by the variation over the range
    const float var = fmaf(2.1e-6f, hi, 1.01f * eroot) * fmaf(fmaf(3.0f * (fabsf(T[3]) + E[3]), hi, 2.0f * (fabsf(T[2]) + E[2])), hi, fabsf(T[1]) + E[1]);
    const CfSample sb = cf_sample(s, T, E, hi);
    const bool in1 = x1 > e2 && x1 < b_hi, in2 = x2 > e2 && x2 < b_hi;
    const CfSample s1 = cf_sample(s, T, E, in1 ? x1 : hi);   // out-of-range points re-sample hi: no effect
    const CfSample s2 = cf_sample(s, T, E, in2 ? x2 : hi);
    if (sb.lo - var > 0.0f && (!in1 || s1.lo > 0.0f) && (!in2 || s2.lo > 0.0f)) return CV_MISS;
    // a certainly negative value strictly inside proves a sign change
    if (hit_ok && (sb.up + var < 0.0f || (in1 && x1 < b_lo && s1.up < 0.0f) || (in2 && x2 < b_lo && s2.up < 0.0f))) return CV_HIT;
    return RT_CF_UNK(2);
}

RT_HD int cubic_verdict_od(const CubicO &o, const CubicD &d, float hi, bool want_hit)
{
    float T[4], B[4];
    T[3] = d.t3;
    B[3] = d.b3;
    T[2] = fmaf(d.gx, o.ox, fmaf(d.gy, o.oy, fmaf(d.gz, o.oz, d.qd)));
    B[2] = fmaf(d.agx, fabsf(o.ox), fmaf(d.agy, fabsf(o.oy), fmaf(d.agz, fabsf(o.oz), d.aqd)));
    T[1] = fmaf(o.hx, d.dx, fmaf(o.hy, d.dy, o.hz * d.dz));
    B[1] = fmaf(o.ahx, fabsf(d.dx), fmaf(o.ahy, fabsf(d.dy), o.ahz * fabsf(d.dz)));
    T[0] = o.t0;
    B[0] = o.b0;
    return cubic_verdict_tb(T, B, hi, want_hit);
}

RT_HD int cubic_verdict(const float *kf, const RayC &r, float hi, bool want_hit)
{
    const CubicO o = make_cubic_o(kf, r.ox, r.oy, r.oz);
    const CubicD d = make_cubic_d(kf, r.dx, r.dy, r.dz);
    return cubic_verdict_od(o, d, hi, want_hit);
}

} // namespace rt
#endif
