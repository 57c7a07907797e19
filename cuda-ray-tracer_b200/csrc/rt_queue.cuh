// rt_queue.cuh — parameter blocks and shared-memory layout of the queue kernel (rt_queue.cu).
#ifndef RT_QUEUE_CUH
#define RT_QUEUE_CUH

#include "rt_cubic.cuh"
#include "rt_tables.cuh"
#include "rt_trace.cuh"

namespace rt {

// Capacity of the scene block carried in the kernel parameters.  Everything except the spheres lives there:
// constant-bank operands cost no load instruction and no register, and a loop index that is uniform across
// the warp turns the cubic zero-coefficient masks into uniform branches.  Scenes beyond these limits are
// rendered by the wavefront kernel (rt_wave.cu) instead.
constexpr int QK_MAX_PLANE = 32;
constexpr int QK_MAX_QUADRIC = 16;
constexpr int QK_MAX_CUBIC = 64;
constexpr int QK_MAX_LIGHT = 64;
constexpr int QK_CAND = 8;         // per-thread queue of sphere groups waiting for the exact test
constexpr int QK_CAP_MULT = 3;     // path records per CTA = 3 x threads: a full wave is always available (rt_queue.cu)

struct QueueScene {
    double4 plane[QK_MAX_PLANE];            // {x, y, z, c}
    double quadric[QK_MAX_QUADRIC][10];     // {x2,y2,z2,xy,xz,yz,x,y,z,c}
    double cubic[QK_MAX_CUBIC][20];         // reference coefficient order
    unsigned cubic_mask[QK_MAX_CUBIC];      // bit i: coefficient i is non-zero
    LightDev light[QK_MAX_LIGHT];
    LightFilter lf[QK_MAX_LIGHT];
};

// Kernel parameters beyond SceneView / FrameParams / QueueScene.
struct QueueConfig {
    int tiles_x, tiles_y;              // 32 x (W/32) pixel tiles over [local_rows][width]
    unsigned long long *tile_counter;
    int streaming;                     // 0: every filter table resident in shared memory; 1: tables streamed in chunks
    int chunk;                         // spheres per chunk (multiple of 4); resident mode: n_sphere_pad
    int n_sphere_pad;                  // n_sphere rounded up to a multiple of 4 (table stride)
    int tables_safe, primary_safe;     // magnitudes small enough for the FP32 tables (else every sphere is tested exactly)
    int use_shadow_cache;
    int poly_first;                    // any-hit rays without a cached occluder: planes/quadrics/cubics before spheres
    float r_max, center_max, center_norm_max, r2_max;
    const float4 *tables;              // global: [2 + n_lights][n_sphere_pad]  (0 generic, 1 primary, 2+j light j)
};

// Shared memory of one CTA (W threads, CAP = CAPM * W path records; CAPM = 3 guarantees full waves, 2 fits a third
// CTA per SM).  Fixed layout, so that every access is
// [register + immediate].  One path record = the state of one pixel's path between two phases: created by the
// wave that finds a hit, shaded in place, and — if the surface reflects — traced again in place.
template <int W, int CAPM>
struct QueueSmem {
    static constexpr int CAP = CAPM * W;
    double px[CAP], py[CAP], pz[CAP];       // hit point
    double nx[CAP], ny[CAP], nz[CAP];       // unit normal at the hit point (not oriented, surface_impl.h:157-172)
    double dx[CAP], dy[CAP], dz[CAP];       // direction of the ray that produced the hit (reflect_ray needs it)
    float out_r[CAP], out_g[CAP], out_b[CAP], ratio[CAP];   // colour so far, cumulative reflection ratio
    int pixel[CAP], obj[CAP], bounces[CAP];
    unsigned short stk_free[CAP], stk_shade[CAP], stk_trace[CAP];   // stacks of record indices
    unsigned short cand[QK_CAND][W];        // per-thread queue of candidate sphere groups
    int n_free, n_shade, n_trace;           // stack heights
    int mode, take, base, tile;             // decision of the scheduler for the current wave
    int tiles_left;
    int cache[QK_MAX_LIGHT];                // last sphere that occluded light j (any occluder will do), -1 = none
};

template <int W, int CAPM>
__host__ __device__ inline size_t qk_fixed_smem() { return align16(sizeof(QueueSmem<W, CAPM>)); }

} // namespace rt
#endif
