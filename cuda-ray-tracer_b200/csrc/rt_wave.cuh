// rt_wave.cuh — shared-memory layout and scene access of the block-local wavefront kernel (rt_wave.cu).
#ifndef RT_WAVE_CUH
#define RT_WAVE_CUH

#include <cstdio>

#include "rt_cubic.cuh"
#include "rt_cubic_filter.cuh"
#include "rt_filter.cuh"
#include "rt_trace.cuh"

namespace rt {

// Checked build (build.py --checked -> librt_b200_checked.so): every index into the ray pool, the hit records, the slot
// table and the frame is asserted in range; a violation prints its source line and traps, which the host sees as a
// failed launch.  The GPU tests run against that library too (tools/gpu_checked.sh).  compute-sanitizer is closed on
// this pool (profiles/r2_compute_sanitizer_closed.log), so these are the bounds checks there are.
#ifdef RT_B200_CHECKED
#define RT_CHECK(cond)                                                                                    \
    do {                                                                                                  \
        if (!(cond)) {                                                                                    \
            printf("RT_CHECK failed: %s  (%s:%d, block %d thread %d)\n", #cond, __FILE__, __LINE__, (int) blockIdx.x, (int) threadIdx.x); \
            __trap();                                                                                     \
        }                                                                                                 \
    } while (0)
#else
#define RT_CHECK(cond) ((void) 0)
#endif

constexpr int LIGHT_CHUNK = 32;  // lights handled per visibility word
constexpr int MAX_WARPS = 16;    // largest CTA: 512 threads

// Pixel tile of a W-thread CTA: every warp owns an 8x4 sub-tile (coherent rays; each warp writes four fully used
// 32-byte sectors); up to four warps sit side by side, further warps stack upwards.
template <int W>
struct TileGeom {
    static constexpr int WARPS = W / 32;
    static constexpr int WX = WARPS >= 4 ? 4 : WARPS;
    static constexpr int WY = WARPS / WX;
    static constexpr int TILE_W = 8 * WX;
    static constexpr int TILE_ROWS = 4 * WY;
};

// Scene arrays staged once per CTA into shared memory (SoA by class, rt_scene.h).
struct SceneSmem {
    const double4 *plane;
    const d3 *plane_n;           // [n_plane] unit normals: a plane's gradient is the same at every point
    const float4 *sphere_f;      // shared or global, see WaveConfig::spheres_in_smem
    const double *quadric;
    const double *cubic;
    const unsigned *cubic_mask;
    const float *cubic_f;        // [n_cubic][RT_CF_FLOATS]: FP32 tables of the cubic filter (rt_cubic_filter.cuh)
    const CubicO *cam_o;         // [n_cubic]: origin half of the filter for the camera origin (all primary rays)
    const CubicD *light_d;       // [n_lights]: direction half per directional light   } only when the scene has
    CubicO *hit_o;               // [W]: origin half per hit of the current wave        } exactly one cubic, else null
    const LightDev *lights;
};

// Ray pool: reflection rays waiting for a full wave (block-local, never leaves the SM).  W = threads per CTA
// = rays per wave; capacity 2 W (a wave is taken as soon as W rays wait, and adds at most W).
template <int W>
struct Pool {
    double ox[2 * W], oy[2 * W], oz[2 * W];
    double dx[2 * W], dy[2 * W], dz[2 * W];
    float out_r[2 * W], out_g[2 * W], out_b[2 * W], ratio[2 * W];
    int pixel[2 * W];
    int bounces[2 * W];
};

// Per-wave hit records (indexed by wave slot, not compacted: only the index list is).
template <int W>
struct WaveHits {
    double px[W], py[W], pz[W];
    double nx[W], ny[W], nz[W];
    unsigned occluded[W];        // bit j: light (chunk base + j) is shadowed (or cannot contribute)
    unsigned short hitlist[W];
};

struct WaveCtl {
    int pool_count;
    int n_hits[2];               // hits of the current wave ([wave & 1]); the other word is zeroed for the next wave
    int next_tile[2];            // this CTA's next tile (n_tiles = none left), double-buffered: while a wave traces the tile
                                 // of buffer b, warp 0 pulls the following one into buffer b ^ 1 — with its cam_x / cam_y —
                                 // so a primary wave starts without a barrier of its own
    int last_cta;                // epilogue: this CTA was the last one of the launch to finish
    int shadow_cache[LIGHT_CHUNK];
    int occ_found[2];            // occluders found so far by [0] the sphere pass, [1] the cubic pass
    int prim_rays, prim_hits;    // recent primary rays / hits of this CTA (decayed): the cubic filter runs on primary
    int prim_filter;             // rays only while enough of them miss for a proven miss to pay for the filter
    int dyn_slot, dyn_group;     // dynamic ownership: the last slot -> group lookup (thread 0 only)
    double cam_x[2][32], cam_y[2][MAX_WARPS];   // camera-plane coordinates of the buffered tile's columns / rows
};

// Frame synchronisation of the fused multi-GPU gather, done by the render kernel itself (no extra launches):
//   prologue  rank 0 publishes go = seq; the other ranks wait for go >= seq - 1 before their first store into rank 0's
//             frame (rank 0 may still be reading the frame that used this buffer two frames ago)
//   epilogue  the last CTA to finish publishes flags[rank] = seq after a system-scope fence; on rank 0 it then
//             waits until every rank's flag has reached seq, so the kernel's completion means "frame assembled"
//   dyn        dynamic ownership of row groups across GPUs, see WaveDyn below
struct WaveSync {
    unsigned *go;                // in rank 0's memory (null: no fused gather, nothing of this runs)
    unsigned *flags;             // [world], in rank 0's memory
    unsigned *timed_out;         // mapped host word
    unsigned seq;
    int rank, world;
};

// Dynamic ownership of the image across GPUs (fused gather only: any GPU may store any pixel into rank 0's frame).
// The image is cut into GROUPS of group_rows rows.  A job-wide counter in rank 0's memory hands out groups; each GPU
// keeps a table slot -> group and its CTAs pull TILES from the GPU's own counter as before: local tile t is tile
// t % tiles_per_group of slot t / tiles_per_group.  The CTA that takes the first tile of slot k fetches the group of
// slot k + ahead from the job's counter (one system-scope atomic over NVLink, ~3 us, for a whole group of tiles) and
// publishes it in the table, so by the time anybody needs slot k + ahead its group is known; the first `ahead` slots
// of every GPU are assigned statically (group = slot * world + rank).  Remote traffic: one atomic per group.  A GPU that
// is done early simply takes more groups — no GPU waits for another while there is work left.
//   counters[2]  in rank 0's memory, used alternately: frame seq counts in counters[seq & 1]; rank 0 zeroes the other
//                before it publishes go = seq (nobody counts there: rank 0 finished frame seq - 1 only after every
//                rank had, and nobody starts frame seq + 1 before go >= seq)
//   slot_group   this GPU's table; -1 = not fetched yet, -2 = no groups left; re-armed by the last CTA of each launch
struct WaveDyn {
    unsigned long long *counters;   // null: static strips (each rank renders its own strips)
    int *slot_group;
    int n_groups, group_rows, max_slots, ahead;
};

// q = n / d for n * d < 2^32 with inv = ceil(2^32 / d) (d >= 2; d == 1 is passed as inv = 0)
__host__ __device__ inline unsigned magic_inv(unsigned d) { return d > 1u ? 0xffffffffu / d + 1u : 0u; }
__device__ __forceinline__ unsigned magic_div(unsigned n, unsigned inv) { return inv ? __umulhi(n, inv) : n; }

struct WaveConfig {
    int tiles_x, tiles_y;        // 32 x (W/32) pixel tiles over [local_rows][width]
    unsigned inv_tiles_x;        // magic_inv(tiles_x): tile -> (column, row) without an integer division per thread
    unsigned tiles_per_group, inv_tiles_per_group;   // dynamic ownership: tiles in one row group
    int spheres_in_smem;
    int use_shadow_cache;
    int use_cubic_filter;
    int use_grid;                // sphere class through the uniform grid (rt_grid.cuh) instead of the brute-force loop
    int cull_backfacing;         // skip shadow rays of lights whose Lambert factor is 0 at the hit (they add exactly 0)
    int lone_kernel;             // host side: use the lone-cubic instantiation for scenes of exactly one cubic object
    unsigned long long *tile_counter;   // [0] next tile, [1] CTAs finished; the last CTA leaves both at 0
    WaveSync sync;
    WaveDyn dyn;
    const unsigned *cubic_mask;  // global copy
    float sphere_r2_max;
};

// dynamic shared memory: [Pool][WaveHits][WaveCtl][scene arrays...]
template <int W>
__host__ __device__ inline size_t wave_fixed_smem()
{
    return align16(sizeof(Pool<W>)) + align16(sizeof(WaveHits<W>)) + align16(sizeof(WaveCtl));
}

__host__ __device__ inline size_t wave_scene_smem(const SceneView &sc, bool spheres)
{
    return align16(sizeof(double4) * sc.n_plane) + align16(sizeof(d3) * sc.n_plane) + (spheres ? align16(sizeof(float4) * sc.n_sphere) : 0) +
           align16(sizeof(double) * 10 * sc.n_quadric) + align16(sizeof(double) * 20 * sc.n_cubic) +
           align16(sizeof(unsigned) * sc.n_cubic) + align16(sizeof(float) * RT_CF_FLOATS * sc.n_cubic) +
           align16(sizeof(LightDev) * sc.n_lights);
}

// filter caches behind the scene arrays: per cubic the camera-origin half; for single-cubic scenes also the per-light
// direction halves and one origin half per wave slot (the shadow origin of that slot's hit)
__host__ __device__ inline bool wave_single_cubic(const SceneView &sc) { return sc.n_cubic == 1 && sc.n_lights <= 256; }

__host__ __device__ inline size_t wave_cubic_smem(const SceneView &sc, int wave)
{
    size_t n = align16(sizeof(CubicO) * sc.n_cubic);
    if (wave_single_cubic(sc)) n += align16(sizeof(CubicD) * sc.n_lights) + align16(sizeof(CubicO) * wave);
    return n;
}

} // namespace rt
#endif
