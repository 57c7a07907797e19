// rt_filter.cuh — conservative FP32 ray/sphere rejection test in front of the exact FP64 test.
//
// For a sphere-class object the reference's decision depends on delta = t1^2 - 4 t2 t0
// (include/surface_impl.h:140-148).  With oc = o - C, A = d.d, b = oc.d, c = oc.oc - r^2 the same quantity is
// delta = 4 (b^2 - A c).  The filter evaluates disc = b^2 - A c in FP32 (FMA) from centre-relative data and
// rejects the sphere only when disc < -margin, where margin bounds |disc_fp32 - disc_exact| (derivation below).
// Everything not rejected goes to the exact FP64 test, so the result is bit-identical to running the exact
// test on every sphere; only the amount of FP64 work changes.
//
// Error bound (u = 2^-24, D = |d|, Q = |o - C|, e1 >= |oc_fp32 - oc_exact|_inf):
//   e1   <= u (|o|_inf + |C|_inf + |oc|_inf) <= 2 u (|o|_inf + Cmax)            (rounding o, C and the subtraction)
//   |db| <= sqrt3 e1 D + sqrt3 u Q D + 3 u Q D            <= D (2 e1 + 5 u Q)
//   |d(b^2)|  <= 2 Q D |db| + |db|^2 + u b^2               <= D^2 (4 e1 Q + 11 u Q^2) + ...
//   |d(q2)|   <= 2 sqrt3 Q e1 + 3 e1^2 + 3 u Q^2           <= 4 e1 Q + 3 u Q^2 + 3 e1^2
//   |d(A c)|  <= D^2 (4 e1 Q + 3 u Q^2 + 2 u r^2 + 6 u (Q^2 + r^2))
//   |d disc|  <= D^2 (8 e1 Q + 21 u Q^2 + 8 u r^2 + 6 e1^2)
// and with Q <= (Q^2 / S + S) / 2 for the scene scale S:
//   |d disc|  <= D^2 ( (4 e1 / S + 21 u) Q^2 + 4 e1 S + 8 u r2max + 6 e1^2 )
// The kernel uses twice that (k1, k0 below), evaluated with the rounded-up FP32 A and the FP32 q2.
#ifndef RT_FILTER_CUH
#define RT_FILTER_CUH

#include "rt_math.cuh"

namespace rt {

#define RT_U32F 5.9604645e-8f /* 2^-24 */

// Per-ray constants of the sphere filter.
struct RayF {
    float ox, oy, oz;   // origin rounded to FP32
    float dx, dy, dz;   // direction rounded to FP32
    float A;            // d.d (FP32)
    float kA1, kA0;     // margin = kA1 * q2 + kA0   (already multiplied by A, rounded up)
    float kb1, kb0;     // |b_fp32 - b_exact| <= kb1 * q2 + kb0
    float kc1, kc0;     // |c_fp32 - c_exact| <= kc1 * q2 + kc0
    bool safe;          // false: magnitudes could overflow FP32 (inf - inf = NaN) -> skip the filter for this ray
};

__device__ __forceinline__ RayF make_rayf(const d3 &o, const d3 &d, float center_max, float r2_max)
{
    RayF r;
    r.ox = (float) o.x; r.oy = (float) o.y; r.oz = (float) o.z;
    r.dx = (float) d.x; r.dy = (float) d.y; r.dz = (float) d.z;
    r.A = fmaf(r.dx, r.dx, fmaf(r.dy, r.dy, r.dz * r.dz));
    const float omax = fmaxf(fabsf(r.ox), fmaxf(fabsf(r.oy), fabsf(r.oz)));
    const float e1 = 2.0f * RT_U32F * (omax + center_max) * 1.001f;
    const float S = fmaxf(1.0f, center_max);
    const float Aup = r.A * 1.0001f;
    const float Dup = sqrtf(Aup) * 1.0001f;
    const float k1 = 2.0f * (4.0f * e1 / S + 21.0f * RT_U32F);
    const float k0 = 2.0f * (4.0f * e1 * S + 8.0f * RT_U32F * r2_max + 6.0f * e1 * e1);
    r.kA1 = Aup * k1;
    r.kA0 = Aup * k0;
    // b: D (2 e1 + 5 u Q), Q <= (Q^2 + 1) / 2
    r.kb1 = Dup * 3.0f * RT_U32F * 2.0f;
    r.kb0 = Dup * (2.0f * e1 + 3.0f * RT_U32F) * 2.0f;
    // c: 4 e1 Q + 3 u Q^2 + 3 e1^2 + u r^2 + u |c|, Q <= (Q^2 / S + S) / 2
    r.kc1 = 2.0f * (2.0f * e1 / S + 5.0f * RT_U32F);
    r.kc0 = 2.0f * (2.0f * e1 * S + 3.0f * e1 * e1 + 2.0f * RT_U32F * r2_max);
    // q2 <= 3 (omax + Cmax)^2 and A q2 must stay finite in FP32 (max 3.4e38): keep every factor below 1e9
    // ... and the reference must really be in its quadratic branch: t2 = d.d > 1e-7 (surface_impl.h:138); a direction
    // shorter than that (a point light within 0.3 mm of the hit) sends it to the linear formula, which the geometry
    // above says nothing about
    r.safe = omax < 1e9f && center_max < 1e9f && r.A < 1e18f && r2_max < 1e18f && r.A > 2e-7f;
    return r;
}

// Hot-loop half of the filter: disc + margin for one sphere (negative = the line provably misses it).
// 14 FP32 instructions, no branch; four of these are evaluated back to back for instruction-level parallelism.
__device__ __forceinline__ float sphere_slack(const RayF &r, const float4 s)
{
    const float ocx = r.ox - s.x, ocy = r.oy - s.y, ocz = r.oz - s.z;
    const float b = fmaf(ocx, r.dx, fmaf(ocy, r.dy, ocz * r.dz));
    const float q2 = fmaf(ocx, ocx, fmaf(ocy, ocy, ocz * ocz));
    const float c = q2 - s.w;
    const float disc = fmaf(b, b, -(r.A * c));
    return disc + fmaf(q2, r.kA1, r.kA0);
}

// Outcome of the filter for one sphere {cx, cy, cz, r^2}.
//   true  -> provably no intersection usable by the caller: no real root, or both roots behind the origin,
//            or (when segment01: the caller's max_t is 1, a point light's shadow ray) both roots beyond t = 1 — the exact test would not have produced a hit/occluder
//   false -> undecided: run the exact FP64 test
__device__ __forceinline__ bool sphere_rejected(const RayF &r, const float4 s, bool segment01)
{
    const float ocx = r.ox - s.x, ocy = r.oy - s.y, ocz = r.oz - s.z;
    const float b = fmaf(ocx, r.dx, fmaf(ocy, r.dy, ocz * r.dz));
    const float q2 = fmaf(ocx, ocx, fmaf(ocy, ocy, ocz * ocz));
    const float c = q2 - s.w;
    const float disc = fmaf(b, b, -(r.A * c));
    const float margin = fmaf(q2, r.kA1, r.kA0);
    if (disc < -margin) return true;           // the common case: the line misses the sphere
    // rare path: the line (nearly) touches the sphere; cheap sign arguments on t = (-b +- sqrt(disc)) / A
    const float eb = fmaf(q2, r.kb1, r.kb0);
    const float ec = fmaf(q2, r.kc1, r.kc0);
    if (b > eb && c > ec) return true;         // origin outside, centre behind: both roots negative
    if (segment01) {
        // both roots > 1  <=>  p(1) = A + 2b + c > 0 and the vertex -b/A > 1, with margins
        const float p1 = r.A + 2.0f * b + c;
        const float e1 = 2.0f * eb + ec + 4.0f * RT_U32F * (r.A + 2.0f * fabsf(b) + fabsf(c)) + 4.0f * RT_U32F * r.A;
        if (p1 > e1 && -b > r.A * 1.0001f + eb) return true;
    }
    return false;
}

} // namespace rt
#endif
