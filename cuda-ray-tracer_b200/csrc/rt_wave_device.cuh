// rt_wave_device.cuh — device functions of the wavefront kernel (phases A and C), shared by the two translation
// units that instantiate it: rt_wave.cu (scenes with cubic-class objects) and rt_wave_plain.cu (scenes without).
#ifndef RT_WAVE_DEVICE_CUH
#define RT_WAVE_DEVICE_CUH

#include <atomic>

#include "rt_kernels.h"
#include "rt_wave.cuh"

namespace rt {

namespace {

// ---- sphere class through the uniform grid (rt_grid.cuh; option accel) -----------------------------------
// The grid only proposes candidates: each one goes through the same FP32 filter and exact FP64 test as in the
// brute-force loops below, so the frame is identical.  Out of line on purpose: the walker's registers must not
// weigh on the scenes that never call it.
struct GridNearestResult {
    double t;
    int slot;            // class-sorted slot, -1 = none so far
    unsigned confirms, tests;
};

struct GridNearestVisitor {
    const SceneView &g;
    const RayF &rf;
    const d3 &o, &d;
    GridNearestResult r;
    __device__ __forceinline__ bool test(int j)
    {
        r.tests++;
        if (!rf.safe || !sphere_rejected(rf, __ldg(&g.sphere_f[j]), false)) {
            r.confirms++;
            consider_lazy_t(r.t, r.slot, intersect_sphere(g.sphere[j], o, d), g.n_plane + j, g.orig_index);
        }
        return false;
    }
    __device__ __forceinline__ double bound() const { return r.t; }
};

struct GridOccluderVisitor {
    const SceneView &g;
    const RayF &rf;
    const d3 &o, &d;
    double max_t;
    bool segment01;
    int found;
    unsigned confirms, tests;
    __device__ __forceinline__ bool test(int j)
    {
        tests++;
        if (!rf.safe || !sphere_rejected(rf, __ldg(&g.sphere_f[j]), segment01)) {
            confirms++;
            if (blocks(intersect_sphere(g.sphere[j], o, d), max_t)) {
                found = g.n_plane + j;
                return true;
            }
        }
        return false;
    }
    __device__ __forceinline__ double bound() const { return max_t; }
};

// best_t / best_slot: the nearest hit found so far among the classes visited earlier (it bounds the walk)
static __device__ __noinline__ GridNearestResult nearest_spheres_grid(const SceneView &g, float sphere_r2_max, d3 o, d3 d,
                                                                      double best_t, int best_slot)
{
    const RayF rf = make_rayf(o, d, g.sphere_center_max, sphere_r2_max);
    GridNearestVisitor v{g, rf, o, d, GridNearestResult{best_t, best_slot, 0u, 0u}};
    grid_walk(g.grid, o.x, o.y, o.z, d.x, d.y, d.z, RT_MAX_T, v);
    return v.r;
}

struct GridOccluderResult {
    int found;           // occluding slot or -1
    unsigned confirms, tests;
};

static __device__ __noinline__ GridOccluderResult occluder_spheres_grid(const SceneView &g, float sphere_r2_max, d3 o, d3 d,
                                                                        double max_t, bool segment01)
{
    const RayF rf = make_rayf(o, d, g.sphere_center_max, sphere_r2_max);
    GridOccluderVisitor v{g, rf, o, d, max_t, segment01, -1, 0u, 0u};
    grid_walk(g.grid, o.x, o.y, o.z, d.x, d.y, d.z, max_t, v);
    return GridOccluderResult{v.found, v.confirms, v.tests};
}

// ---- phase A: nearest hit (update-cpu.cpp:48-56) ---------------------------------------------------------
// CUBIC = false compiles the cubic class out (scenes without cubics run a leaner kernel).
// GRID = true compiles the grid walk in (its own kernel instantiations, so the brute-force kernels carry none of it).
// WIDE = true (the 512-thread instantiations, which the launcher picks when the sphere array is large): the brute-force
// sphere loops run eight wide first — the loop's own instructions (address, maximum, branch) are a fifth of a four-wide
// iteration; C5 brute force -9 %.  Not in the 256-thread kernels: 20spheres (five iterations of four) is 2 % slower with it,
// and merely having the extra loop in the code shifts the small scenes' kernels by 1 %.
template <bool CUBIC, bool GRID, bool WIDE>
__device__ __forceinline__ Hit nearest_hit_wave(const SceneView &g, const SceneSmem &s, const WaveConfig &cfg,
                                                const d3 &o, const d3 &d, bool primary, bool filter, Tally &tl)
{
    Hit h;
    h.t = CUDART_INF;
    h.slot = -1;
    h.orig = 0x7fffffff;
    int slot = 0;
    for (int i = 0; i < g.n_plane; i++, slot++) consider_lazy(h, intersect_plane(s.plane[i], o, d), slot, g.orig_index);
    if (GRID && g.n_sphere > 0 && cfg.use_grid && grid_ray_ok(g.grid, o.x, o.y, o.z, d.x, d.y, d.z)) {
        const GridNearestResult r = nearest_spheres_grid(g, cfg.sphere_r2_max, o, d, h.t, h.slot);
        h.t = r.t;
        h.slot = r.slot;
        slot += g.n_sphere;
        tl.n[CNT_CONFIRM] += r.confirms;
        tl.n[CNT_T_SPHERE] += r.tests;
    }
    else if (g.n_sphere > 0) {
        const RayF rf = make_rayf(o, d, g.sphere_center_max, cfg.sphere_r2_max);
        unsigned confirms = 0;
        const int n4 = rf.safe ? (g.n_sphere & ~3) : 0;
        int i = 0;
        if (WIDE) {
            const int n8 = rf.safe ? (g.n_sphere & ~7) : 0;
            for (; i < n8; i += 8) {
                const float s0 = sphere_slack(rf, s.sphere_f[i]), s1 = sphere_slack(rf, s.sphere_f[i + 1]);
                const float s2 = sphere_slack(rf, s.sphere_f[i + 2]), s3 = sphere_slack(rf, s.sphere_f[i + 3]);
                const float s4 = sphere_slack(rf, s.sphere_f[i + 4]), s5 = sphere_slack(rf, s.sphere_f[i + 5]);
                const float s6 = sphere_slack(rf, s.sphere_f[i + 6]), s7 = sphere_slack(rf, s.sphere_f[i + 7]);
                if (!(fmaxf(fmaxf(fmaxf(s0, s1), fmaxf(s2, s3)), fmaxf(fmaxf(s4, s5), fmaxf(s6, s7))) < 0.0f)) {
#pragma unroll 1
                    for (int j = i; j < i + 8; j++) {
                        if (!sphere_rejected(rf, s.sphere_f[j], false)) {
                            confirms++;
                            consider_lazy(h, intersect_sphere(g.sphere[j], o, d), slot + j, g.orig_index);
                        }
                    }
                }
            }
        }
        for (; i < n4; i += 4) {
            // four independent filter evaluations, one branch: candidates are rare
            const float s0 = sphere_slack(rf, s.sphere_f[i]), s1 = sphere_slack(rf, s.sphere_f[i + 1]);
            const float s2 = sphere_slack(rf, s.sphere_f[i + 2]), s3 = sphere_slack(rf, s.sphere_f[i + 3]);
            if (!(fmaxf(fmaxf(s0, s1), fmaxf(s2, s3)) < 0.0f)) {
#pragma unroll 1
                for (int j = i; j < i + 4; j++) {
                    if (!sphere_rejected(rf, s.sphere_f[j], false)) {
                        confirms++;
                        consider_lazy(h, intersect_sphere(g.sphere[j], o, d), slot + j, g.orig_index);
                    }
                }
            }
        }
        for (; i < g.n_sphere; i++) {
            if (!rf.safe || !sphere_rejected(rf, s.sphere_f[i], false)) {
                confirms++;
                consider_lazy(h, intersect_sphere(g.sphere[i], o, d), slot + i, g.orig_index);
            }
        }
        slot += g.n_sphere;
        tl.n[CNT_CONFIRM] += confirms;
        tl.n[CNT_T_SPHERE] += g.n_sphere;
    }
    for (int i = 0; i < g.n_quadric; i++, slot++)
        consider_lazy(h, intersect_quadric(s.quadric + 10 * i, o, d), slot, g.orig_index);
    if (CUBIC && g.n_cubic > 0) {
        // FP32 filter first: a proven miss needs no FP64 expansion and no cbrt/acos/cos (rt_cubic_filter.cuh).
        // Primary rays share the origin, so their origin half comes from the per-CTA cache.
        const float ofx = (float) o.x, ofy = (float) o.y, ofz = (float) o.z;
        const float dfx = (float) d.x, dfy = (float) d.y, dfz = (float) d.z;
        unsigned exact = 0;
        for (int i = 0; i < g.n_cubic; i++, slot++) {
            if (cfg.use_cubic_filter && filter) {
                const float *kf = s.cubic_f + RT_CF_FLOATS * i;
                const CubicD cd = make_cubic_d(kf, dfx, dfy, dfz);
                const CubicO co = primary ? s.cam_o[i] : make_cubic_o(kf, ofx, ofy, ofz);
                int v;
                if (g.n_cubic > 1) {
                    // Several cubic surfaces: a candidate matters only if its t is below the best t so far (or ties with
                    // it), so the filter is asked about [EPS, best_t] — rounded up to FP32, and the proven-empty interval
                    // extends 2e-6 (relative) past its end — and only the few surfaces that can still improve the hit
                    // reach the exact test.
                    v = cubic_verdict_od(co, cd, fminf((float) RT_MAX_T, __double2float_ru(h.t)), false);
                }
                else {
                    v = cubic_verdict_od(co, cd, (float) RT_MAX_T, false);   // constant end: the compiler folds it
                }
                if (v == CV_MISS) continue;
            }
            exact++;
            consider_lazy(h, intersect_cubic_masked(s.cubic + 20 * i, s.cubic_mask[i], o, d), slot, g.orig_index);
        }
        tl.n[CNT_CUBIC_EXACT] += exact;
    }
    if (h.slot >= 0) h.orig = __ldg(&g.orig_index[h.slot]);
    tl.n[CNT_T_PLANE] += g.n_plane;
    tl.n[CNT_T_QUADRIC] += g.n_quadric;
    if (CUBIC) tl.n[CNT_T_CUBIC] += g.n_cubic;
    return h;
}

// exact test of one class-sorted slot (shadow cache probe; planes, spheres and quadrics)
__device__ __forceinline__ double intersect_slot(const SceneView &g, const SceneSmem &s, int slot, const d3 &o,
                                                 const d3 &d, Tally &tl)
{
    if (slot < g.n_plane) {
        tl.n[CNT_T_PLANE]++;
        return intersect_plane(s.plane[slot], o, d);
    }
    slot -= g.n_plane;
    if (slot < g.n_sphere) {
        tl.n[CNT_T_SPHERE]++;
        tl.n[CNT_CONFIRM]++;
        return intersect_sphere(g.sphere[slot], o, d);
    }
    slot -= g.n_sphere;
    if (slot < g.n_quadric) {
        tl.n[CNT_T_QUADRIC]++;
        return intersect_quadric(s.quadric + 10 * slot, o, d);
    }
    return -1.0;   // cached cubic slots are not probed here: they become the first cubic of the class search (occluder_cubics)
}

// ---- phase C: any occluder with EPS < t < max_t (update-cpu.cpp:66-72); returns the occluding slot or -1 ----
// The boolean is order-independent, so classes are visited cheapest-first, except that the cubic class moves
// in front of the sphere class while cubics are the more productive occluders per unit of work (block-local
// statistics in WaveCtl::occ_found, updated as occluders are found).
// Code size matters here (the kernel's instruction footprint was 27 400 SASS instructions in round 1, two thirds of it
// copies of these searches, and "no instruction" the top stall on dingdong): every class search below exists ONCE —
// the point-light / directional distinction (segment01) is a run-time flag and the sphere / cubic order a two-pass loop.
template <bool GRID, bool WIDE>
__device__ __forceinline__ int occluder_spheres(const SceneView &g, const SceneSmem &s, const WaveConfig &cfg,
                                                const d3 &o, const d3 &d, double max_t, bool segment01, Tally &tl)
{
    if (GRID && cfg.use_grid && grid_ray_ok(g.grid, o.x, o.y, o.z, d.x, d.y, d.z)) {
        const GridOccluderResult r = occluder_spheres_grid(g, cfg.sphere_r2_max, o, d, max_t, segment01);
        tl.n[CNT_T_SPHERE] += r.tests;
        tl.n[CNT_CONFIRM] += r.confirms;
        return r.found;
    }
    const int slot = g.n_plane;
    const RayF rf = make_rayf(o, d, g.sphere_center_max, cfg.sphere_r2_max);
    int found = -1;
    const int n4 = rf.safe ? (g.n_sphere & ~3) : 0;
    int i = 0;
    if (WIDE) {   // see nearest_hit_wave
        const int n8 = rf.safe ? (g.n_sphere & ~7) : 0;
        for (; i < n8 && found < 0; i += 8) {
            const float s0 = sphere_slack(rf, s.sphere_f[i]), s1 = sphere_slack(rf, s.sphere_f[i + 1]);
            const float s2 = sphere_slack(rf, s.sphere_f[i + 2]), s3 = sphere_slack(rf, s.sphere_f[i + 3]);
            const float s4 = sphere_slack(rf, s.sphere_f[i + 4]), s5 = sphere_slack(rf, s.sphere_f[i + 5]);
            const float s6 = sphere_slack(rf, s.sphere_f[i + 6]), s7 = sphere_slack(rf, s.sphere_f[i + 7]);
            if (!(fmaxf(fmaxf(fmaxf(s0, s1), fmaxf(s2, s3)), fmaxf(fmaxf(s4, s5), fmaxf(s6, s7))) < 0.0f)) {
#pragma unroll 1
                for (int j = i; j < i + 8; j++) {
                    if (!sphere_rejected(rf, s.sphere_f[j], segment01)) {
                        tl.n[CNT_CONFIRM]++;
                        if (blocks(intersect_sphere(g.sphere[j], o, d), max_t)) {
                            found = slot + j;
                            break;
                        }
                    }
                }
            }
        }
    }
    for (; i < n4 && found < 0; i += 4) {
        const float s0 = sphere_slack(rf, s.sphere_f[i]), s1 = sphere_slack(rf, s.sphere_f[i + 1]);
        const float s2 = sphere_slack(rf, s.sphere_f[i + 2]), s3 = sphere_slack(rf, s.sphere_f[i + 3]);
        if (!(fmaxf(fmaxf(s0, s1), fmaxf(s2, s3)) < 0.0f)) {
#pragma unroll 1
            for (int j = i; j < i + 4; j++) {
                if (!sphere_rejected(rf, s.sphere_f[j], segment01)) {
                    tl.n[CNT_CONFIRM]++;
                    if (blocks(intersect_sphere(g.sphere[j], o, d), max_t)) {
                        found = slot + j;
                        break;
                    }
                }
            }
        }
    }
#pragma unroll 1
    for (; i < g.n_sphere && found < 0; i++) {
        if (!rf.safe || !sphere_rejected(rf, s.sphere_f[i], segment01)) {
            tl.n[CNT_CONFIRM]++;
            if (blocks(intersect_sphere(g.sphere[i], o, d), max_t)) found = slot + i;
        }
    }
    tl.n[CNT_T_SPHERE] += i;
    return found;
}

// Cubic class.  hit_slot / light: wave slot of the hit this shadow ray starts from and index of the light it goes to,
// the keys of the filter caches (SceneSmem::hit_o, light_d) that exist when the scene has exactly one cubic: the origin
// half of the FP32 expansion per hit, the direction half per directional light (a point light's direction differs per
// hit).  One verdict call site and one exact test for both the cached and the generic form.
// first: the cubic to test first (the per-light shadow cache's last occluder when that was a cubic), the others follow
// cyclically — the cached occluder costs no code of its own.
__device__ __forceinline__ int occluder_cubics(const SceneView &g, const SceneSmem &s, const WaveConfig &cfg, const d3 &o,
                                               const d3 &d, double max_t, bool segment01, int hit_slot, int light, int first, Tally &tl)
{
    const int base = g.n_plane + g.n_sphere + g.n_quadric;
    const float hi = segment01 ? 1.0f : (float) RT_MAX_T;   // = max_t, exact in FP32
    const float ofx = (float) o.x, ofy = (float) o.y, ofz = (float) o.z;
    const float dfx = (float) d.x, dfy = (float) d.y, dfz = (float) d.z;
    const bool cached = s.hit_o != nullptr;
#pragma unroll 1
    for (int n = 0; n < g.n_cubic; n++) {
        int i = n + first;
        if (i >= g.n_cubic) i -= g.n_cubic;
        tl.n[CNT_T_CUBIC]++;
        if (cfg.use_cubic_filter) {
            const float *kf = s.cubic_f + RT_CF_FLOATS * i;
            const CubicO co = cached ? s.hit_o[hit_slot] : make_cubic_o(kf, ofx, ofy, ofz);
            const CubicD cd = (cached && !segment01) ? s.light_d[light] : make_cubic_d(kf, dfx, dfy, dfz);
            const int v = cubic_verdict_od(co, cd, hi, true);
            if (v == CV_MISS) continue;
            if (v == CV_HIT) return base + i;
        }
        tl.n[CNT_CUBIC_EXACT]++;
        if (blocks(intersect_cubic_masked(s.cubic + 20 * i, s.cubic_mask[i], o, d), max_t)) return base + i;
    }
    return -1;
}

template <bool CUBIC, bool GRID, bool WIDE>
__device__ __forceinline__ int occluder_wave(const SceneView &g, const SceneSmem &s, const WaveConfig &cfg, WaveCtl &ctl,
                                             const d3 &o, const d3 &d, double max_t, bool segment01, int hit_slot, int light,
                                             int cached_cubic, Tally &tl)
{
    for (int i = 0; i < g.n_plane; i++) {
        tl.n[CNT_T_PLANE]++;
        if (blocks(intersect_plane(s.plane[i], o, d), max_t)) return i;
    }
    const int quad_base = g.n_plane + g.n_sphere;
    for (int i = 0; i < g.n_quadric; i++) {
        tl.n[CNT_T_QUADRIC]++;
        if (blocks(intersect_quadric(s.quadric + 10 * i, o, d), max_t)) return quad_base + i;
    }
    const bool have_s = g.n_sphere > 0, have_c = CUBIC && g.n_cubic > 0;
    bool cubic_first = !have_s;
    if (have_s && have_c) {
        // relative cost of a full pass: ~4 issue slots x 4 per sphere filter (a flat ~2000 for a grid walk), ~600 per
        // exact cubic test
        const volatile int *found = ctl.occ_found;
        const long long sphere_pass = (GRID && cfg.use_grid) ? 2000LL : 16LL * g.n_sphere;
        cubic_first = (long long) found[1] * sphere_pass > (long long) found[0] * (600LL * g.n_cubic);
        if (cached_cubic >= 0) cubic_first = true;   // this light's last occluder was a cubic: try it before anything else
    }
    // straight-line: the sphere search (small code, the hot loop of sphere-heavy scenes) is instantiated twice, the cubic
    // search (large code) once.  (Both inside one two-pass loop cost the sphere loop its unrolling: C5 brute force +45 %.)
    const bool both = have_s && have_c;
    if (have_s && !cubic_first) {
        const int by = occluder_spheres<GRID, WIDE>(g, s, cfg, o, d, max_t, segment01, tl);
        if (by >= 0) {
            if (both) atomicAdd(&ctl.occ_found[0], 1);
            return by;
        }
    }
    if (have_c) {
        const int by = occluder_cubics(g, s, cfg, o, d, max_t, segment01, hit_slot, light, cached_cubic > 0 ? cached_cubic : 0, tl);
        if (by >= 0) {
            if (both) atomicAdd(&ctl.occ_found[1], 1);
            return by;
        }
    }
    if (have_s && cubic_first) {
        const int by = occluder_spheres<GRID, WIDE>(g, s, cfg, o, d, max_t, segment01, tl);
        if (by >= 0) {
            if (both) atomicAdd(&ctl.occ_found[0], 1);
            return by;
        }
    }
    return -1;
}

} // namespace

} // namespace rt

#endif
