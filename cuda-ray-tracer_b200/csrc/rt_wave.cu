// rt_wave.cu — block-local wavefront kernel (option kernel=1, the default).
//
// One persistent CTA per SM slot pulls 32x8 pixel tiles from a global counter.  Rays are processed in WAVES of
// 256 (one per thread); every wave runs the same four phases, separated by __syncthreads():
//   A  nearest hit          thread <-> ray.   planes/quadrics exact FP64, spheres FP32 filter + exact confirm,
//                           cubics exact FP64 with the zero-coefficient mask (rt_cubic.cuh)
//   B  hit records          P, n into shared memory; ballot/popc prefix sums build the compact hit list
//   C  shadow rays          (hit, light) pairs spread evenly over ALL 256 threads, light-major so neighbouring
//                           lanes carry neighbouring pixels' rays to the same light; result = one bit per light
//   D  shading + bounce     lights summed in index order (FP32, as the reference), blend, then reflective hits
//                           append their bounce ray to the block's POOL with a warp-ballot compaction
// A wave is either the primary rays of the next tile or, as soon as the pool holds a full wave (or tiles ran
// out), 256 pooled bounce rays from any earlier tile — so terminated and non-reflective rays never idle lanes
// in later bounces.  Ray state lives in registers/shared memory only; HBM sees the scene once per CTA and each
// pixel once (128-bit stores where four neighbouring lanes finish together).
//
// Arithmetic is the reference CPU path's exact FP64 order (rt_math.cuh); see rt_filter.cuh for why the FP32
// sphere filter cannot change a result.  Reference: src/update-cpu.cpp:45-139 (what is computed),
// src/update-cuda.cu:104-158 (the kernel this replaces; not a port of it).
#include "rt_kernels.h"
#include "rt_wave.cuh"

namespace rt {

namespace {

// ---- phase A: nearest hit (update-cpu.cpp:48-56) ---------------------------------------------------------
__device__ __forceinline__ Hit nearest_hit_wave(const SceneView &g, const SceneSmem &s, const WaveConfig &cfg,
                                                const d3 &o, const d3 &d, Tally &tl)
{
    Hit h;
    h.t = CUDART_INF;
    h.slot = -1;
    h.orig = 0x7fffffff;
    int slot = 0;
    for (int i = 0; i < g.n_plane; i++, slot++)
        consider(h, intersect_plane(s.plane[i], o, d), slot, __ldg(&g.orig_index[slot]));
    if (g.n_sphere > 0) {
        const RayF rf = make_rayf(o, d, g.sphere_center_max, cfg.sphere_r2_max);
        unsigned confirms = 0;
#pragma unroll 4
        for (int i = 0; i < g.n_sphere; i++) {
            if (!sphere_rejected<false>(rf, s.sphere_f[i])) {
                confirms++;
                consider(h, intersect_sphere(g.sphere[i], o, d), slot + i, __ldg(&g.orig_index[slot + i]));
            }
        }
        slot += g.n_sphere;
        tl.n[CNT_CONFIRM] += confirms;
    }
    for (int i = 0; i < g.n_quadric; i++, slot++)
        consider(h, intersect_quadric(s.quadric + 10 * i, o, d), slot, __ldg(&g.orig_index[slot]));
    for (int i = 0; i < g.n_cubic; i++, slot++)
        consider(h, intersect_cubic_masked(s.cubic + 20 * i, s.cubic_mask[i], o, d), slot, __ldg(&g.orig_index[slot]));
    tl.n[CNT_T_PLANE] += g.n_plane;
    tl.n[CNT_T_SPHERE] += g.n_sphere;
    tl.n[CNT_T_QUADRIC] += g.n_quadric;
    tl.n[CNT_T_CUBIC] += g.n_cubic;
    return h;
}

// exact test of one class-sorted slot (shadow cache probe)
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
    slot -= g.n_quadric;
    tl.n[CNT_T_CUBIC]++;
    return intersect_cubic_masked(s.cubic + 20 * slot, s.cubic_mask[slot], o, d);
}

// ---- phase C: any occluder with EPS < t < max_t (update-cpu.cpp:66-72); returns the occluding slot or -1 ----
template <bool SEGMENT01>
__device__ __forceinline__ int occluder_wave(const SceneView &g, const SceneSmem &s, const WaveConfig &cfg,
                                             const d3 &o, const d3 &d, double max_t, Tally &tl)
{
    int slot = 0;
    for (int i = 0; i < g.n_plane; i++, slot++) {
        tl.n[CNT_T_PLANE]++;
        if (blocks(intersect_plane(s.plane[i], o, d), max_t)) return slot;
    }
    const int quad_base = g.n_plane + g.n_sphere;
    for (int i = 0; i < g.n_quadric; i++) {
        tl.n[CNT_T_QUADRIC]++;
        if (blocks(intersect_quadric(s.quadric + 10 * i, o, d), max_t)) return quad_base + i;
    }
    if (g.n_sphere > 0) {
        const RayF rf = make_rayf(o, d, g.sphere_center_max, cfg.sphere_r2_max);
        int found = -1;
        int i = 0;
#pragma unroll 4
        for (; i < g.n_sphere; i++) {
            if (!sphere_rejected<SEGMENT01>(rf, s.sphere_f[i])) {
                tl.n[CNT_CONFIRM]++;
                if (blocks(intersect_sphere(g.sphere[i], o, d), max_t)) {
                    found = slot + i;
                    i++;
                    break;
                }
            }
        }
        tl.n[CNT_T_SPHERE] += i;
        if (found >= 0) return found;
    }
    const int cubic_base = quad_base + g.n_quadric;
    for (int i = 0; i < g.n_cubic; i++) {
        tl.n[CNT_T_CUBIC]++;
        if (blocks(intersect_cubic_masked(s.cubic + 20 * i, s.cubic_mask[i], o, d), max_t)) return cubic_base + i;
    }
    return -1;
}

} // namespace

} // namespace rt

#include "rt_wave_kernel.inl"
