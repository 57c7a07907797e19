// rt_trace.cuh — nearest-hit and any-hit searches over the class-sorted scene, exact FP64.
//
//   nearest_hit  = the loop at src/update-cpu.cpp:48-56: argmin_i t_i with EPS <= t < MAX_T, ties -> lowest
//                  original index (classes are visited in turn, so the tie rule is applied explicitly)
//   occluded     = the inner loop at src/update-cpu.cpp:66-72: any object with EPS < t < max_t.  The
//                  reference stops at the first occluder in index order; the boolean is order-independent.
#ifndef RT_TRACE_CUH
#define RT_TRACE_CUH

#include "rt_math.cuh"

namespace rt {

struct Tally {
    unsigned int n[CNT_COUNT];
    __device__ __forceinline__ void clear()
    {
#pragma unroll
        for (int i = 0; i < CNT_COUNT; i++) n[i] = 0;
    }
};

struct Hit {
    double t;
    int slot;   // class-sorted slot, -1 = none
    int orig;   // original object index
};

__device__ __forceinline__ void consider(Hit &h, double t, int slot, int orig)
{
    if (t >= RT_EPS && t < RT_MAX_T && (t < h.t || (t == h.t && orig < h.orig))) {
        h.t = t;
        h.slot = slot;
        h.orig = orig;
    }
}

// Same rule with the slot -> original-index lookup deferred: it is only needed on an exact tie in t.
__device__ __forceinline__ void consider_lazy(Hit &h, double t, int slot, const int *__restrict__ orig_index)
{
    if (t >= RT_EPS && t < RT_MAX_T) {
        if (t < h.t) {
            h.t = t;
            h.slot = slot;
        }
        else if (t == h.t && h.slot >= 0 && __ldg(&orig_index[slot]) < __ldg(&orig_index[h.slot])) {
            h.slot = slot;
        }
    }
}

// The same rule on two scalars (grid walk visitors).
__device__ __forceinline__ void consider_lazy_t(double &best_t, int &best_slot, double t, int slot,
                                                const int *__restrict__ orig_index)
{
    if (t >= RT_EPS && t < RT_MAX_T) {
        if (t < best_t) {
            best_t = t;
            best_slot = slot;
        }
        else if (t == best_t && best_slot >= 0 && __ldg(&orig_index[slot]) < __ldg(&orig_index[best_slot])) {
            best_slot = slot;
        }
    }
}

__device__ __forceinline__ Hit nearest_hit_exact(const SceneView &sc, const d3 &o, const d3 &d, Tally &tl)
{
    Hit h;
    h.t = CUDART_INF;
    h.slot = -1;
    h.orig = 0x7fffffff;
    int slot = 0;
    for (int i = 0; i < sc.n_plane; i++, slot++) consider(h, intersect_plane(sc.plane[i], o, d), slot, __ldg(&sc.orig_index[slot]));
    for (int i = 0; i < sc.n_sphere; i++, slot++) consider(h, intersect_sphere(sc.sphere[i], o, d), slot, __ldg(&sc.orig_index[slot]));
    for (int i = 0; i < sc.n_quadric; i++, slot++) consider(h, intersect_quadric(sc.quadric + 10 * i, o, d), slot, __ldg(&sc.orig_index[slot]));
    for (int i = 0; i < sc.n_cubic; i++, slot++) consider(h, intersect_cubic(sc.cubic + 20 * i, o, d), slot, __ldg(&sc.orig_index[slot]));
    tl.n[CNT_T_PLANE] += sc.n_plane;
    tl.n[CNT_T_SPHERE] += sc.n_sphere;
    tl.n[CNT_T_QUADRIC] += sc.n_quadric;
    tl.n[CNT_T_CUBIC] += sc.n_cubic;
    return h;
}

__device__ __forceinline__ bool blocks(double t, double max_t) { return t > RT_EPS && t < max_t; }

__device__ __forceinline__ bool occluded_exact(const SceneView &sc, const d3 &o, const d3 &d, double max_t, Tally &tl)
{
    for (int i = 0; i < sc.n_plane; i++) {
        tl.n[CNT_T_PLANE]++;
        if (blocks(intersect_plane(sc.plane[i], o, d), max_t)) return true;
    }
    for (int i = 0; i < sc.n_sphere; i++) {
        tl.n[CNT_T_SPHERE]++;
        if (blocks(intersect_sphere(sc.sphere[i], o, d), max_t)) return true;
    }
    for (int i = 0; i < sc.n_quadric; i++) {
        tl.n[CNT_T_QUADRIC]++;
        if (blocks(intersect_quadric(sc.quadric + 10 * i, o, d), max_t)) return true;
    }
    for (int i = 0; i < sc.n_cubic; i++) {
        tl.n[CNT_T_CUBIC]++;
        if (blocks(intersect_cubic(sc.cubic + 20 * i, o, d), max_t)) return true;
    }
    return false;
}

// src/update-cpu.cpp:45-80 — returns original object index or -1; on a hit fills colour / point / normal.
__device__ __forceinline__ int trace_and_shade_exact(const SceneView &sc, const d3 &o, const d3 &d, float3 &color,
                                                     d3 &point, d3 &normal, Tally &tl)
{
    Hit h = nearest_hit_exact(sc, o, d, tl);
    if (h.slot < 0) return -1;
    tl.n[CNT_HITS]++;
    d3 p = d3{o.x + h.t * d.x, o.y + h.t * d.y, o.z + h.t * d.z};
    d3 n = surface_normal(sc.coef_full + 20 * h.orig, p);
    float4 mat = __ldg(&sc.material[h.orig]);
    float3 obj = make_float3(mat.y, mat.z, mat.w);
    float3 acc = make_float3(0.0f, 0.0f, 0.0f);
    d3 so = d3{p.x + RT_SHADOW_BIAS * n.x, p.y + RT_SHADOW_BIAS * n.y, p.z + RT_SHADOW_BIAS * n.z};
    for (int j = 0; j < sc.n_lights; j++) {
        const LightDev l = sc.lights[j];
        double max_t;
        d3 sd = shadow_dir(l, p, max_t);
        tl.n[CNT_SHADOW]++;
        if (!occluded_exact(sc, so, sd, max_t, tl)) {
            shade_light(l, p, n, obj, acc);
            tl.n[CNT_LIT]++;
        }
    }
    color.x = (acc.x < 1.0f) ? acc.x : 1.0f;   // glm::min(vec3(1), c), update-cpu.cpp:77
    color.y = (acc.y < 1.0f) ? acc.y : 1.0f;
    color.z = (acc.z < 1.0f) ? acc.z : 1.0f;
    point = p;
    normal = n;
    return h.orig;
}

// Add a thread's tallies into the global 64-bit counters: warp shuffle reduce, one atomic per warp per slot.
__device__ __forceinline__ void flush_tally(const Tally &tl, unsigned long long *counters)
{
    if (counters == nullptr) return;
#pragma unroll
    for (int i = 0; i < CNT_COUNT; i++) {
        unsigned long long v = tl.n[i];
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
        if ((threadIdx.x & 31) == 0 && v != 0) atomicAdd(&counters[i], v);
    }
}

} // namespace rt
#endif
