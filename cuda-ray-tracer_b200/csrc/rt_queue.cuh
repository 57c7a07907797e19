// rt_queue.cuh — shared-memory layout and launch configuration of the queue kernel (rt_queue.cu).
#ifndef RT_QUEUE_CUH
#define RT_QUEUE_CUH

#include "rt_cubic.cuh"
#include "rt_tables.cuh"
#include "rt_trace.cuh"

namespace rt {

constexpr int QK_MAX_MASKS = 64;   // cubic zero-coefficient masks carried in the kernel parameters (uniform branches)
constexpr int QK_CAND = 8;         // per-thread queue of sphere groups waiting for the exact test
constexpr int QK_CACHE = 64;       // lights with a "last occluder" slot

// Kernel parameters beyond SceneView / FrameParams.
struct QueueConfig {
    int tiles_x, tiles_y;              // 32 x (W/32) pixel tiles over [local_rows][width]
    unsigned long long *tile_counter;
    int cap;                           // path records per CTA (>= 2 W)
    int streaming;                     // 0: every filter table resident in shared memory; 1: tables streamed in chunks
    int chunk;                         // spheres per chunk (multiple of 4); resident mode: n_sphere_pad
    int n_sphere_pad;                  // n_sphere rounded up to a multiple of 4 (table stride)
    int tables_safe, primary_safe;     // magnitudes small enough for the FP32 tables (else every sphere is tested exactly)
    int use_shadow_cache;
    int direct;                        // 1: pixels are written at their image position (peer-memory frame), 0: packed rows
    float r_max, center_max, center_norm_max, r2_max;
    const float4 *tables;              // global: [2 + n_lights][n_sphere_pad]  (0 generic, 1 primary, 2+j light j)
    const LightFilter *light_filter;   // global: [n_lights]
    unsigned cubic_mask[QK_MAX_MASKS];
};

// One path record = the state of one pixel's path between two kernel phases.  It is created by the wave that finds
// a hit, shaded in place, and — if the surface reflects — traced again in place, until the path ends.
// SoA over `cap` slots; pointers into dynamic shared memory.
struct Records {
    double *px, *py, *pz;      // hit point
    double *nx, *ny, *nz;      // unit normal at the hit point (not oriented, surface_impl.h:157-172)
    double *dx, *dy, *dz;      // direction of the ray that produced the hit (reflect_ray needs it)
    float *out_r, *out_g, *out_b, *ratio;   // colour accumulated so far and cumulative reflection ratio
    int *pixel, *obj, *bounces;
};

struct QueueCtl {
    int n_free, n_shade, n_trace;   // stack heights
    int mode, take, base, tile;     // decision of the scheduler for the current wave
    int tiles_left;
    int cache[QK_CACHE];            // last sphere that occluded light j (any occluder will do), -1 = none
};

__host__ __device__ inline size_t qk_record_bytes(int cap) { return (size_t) cap * (9 * 8 + 4 * 4 + 3 * 4); }

__host__ __device__ inline size_t qk_fixed_smem(int W, int cap)
{
    return align16(qk_record_bytes(cap)) + align16((size_t) 3 * cap * sizeof(unsigned short)) +
           align16((size_t) W * QK_CAND * sizeof(unsigned short)) + align16(sizeof(QueueCtl));
}

__host__ __device__ inline size_t qk_scene_smem(const SceneView &sc)
{
    return align16(sizeof(double4) * sc.n_plane) + align16(sizeof(double) * 10 * sc.n_quadric) +
           align16(sizeof(double) * 20 * sc.n_cubic) + align16(sizeof(LightDev) * sc.n_lights) +
           align16(sizeof(LightFilter) * sc.n_lights);
}

} // namespace rt
#endif
