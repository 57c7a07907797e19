// rt_wave.cuh — shared-memory layout and scene access of the block-local wavefront kernel (rt_wave.cu).
#ifndef RT_WAVE_CUH
#define RT_WAVE_CUH

#include "rt_cubic.cuh"
#include "rt_filter.cuh"
#include "rt_trace.cuh"

namespace rt {

constexpr int WAVE = 256;        // threads per CTA == rays per wave
constexpr int POOL_CAP = 2 * WAVE;
constexpr int LIGHT_CHUNK = 32;  // lights handled per visibility word

// Scene arrays staged once per CTA into shared memory (SoA by class, rt_scene.h).
struct SceneSmem {
    const double4 *plane;
    const float4 *sphere_f;      // shared or global, see WaveConfig::spheres_in_smem
    const double *quadric;
    const double *cubic;
    const unsigned *cubic_mask;
    const LightDev *lights;
};

// Ray pool: reflection rays waiting for a full wave (block-local, never leaves the SM).
struct Pool {
    double ox[POOL_CAP], oy[POOL_CAP], oz[POOL_CAP];
    double dx[POOL_CAP], dy[POOL_CAP], dz[POOL_CAP];
    float out_r[POOL_CAP], out_g[POOL_CAP], out_b[POOL_CAP], ratio[POOL_CAP];
    int pixel[POOL_CAP];
    int bounces[POOL_CAP];
};

// Per-wave hit records (indexed by wave slot, not compacted: only the index list is).
struct WaveHits {
    double px[WAVE], py[WAVE], pz[WAVE];
    double nx[WAVE], ny[WAVE], nz[WAVE];
    unsigned occluded[WAVE];     // bit j: light (chunk base + j) is shadowed
    unsigned short hitlist[WAVE];
};

struct WaveCtl {
    int pool_count;
    int n_hits;
    int tile;
    int warp_base[WAVE / 32];
    int shadow_cache[LIGHT_CHUNK];
    int occ_found[2];            // occluders found so far by [0] the sphere pass, [1] the cubic pass
};

struct WaveConfig {
    int tiles_x, tiles_y;        // 32x8 pixel tiles over [local_rows][width]
    int spheres_in_smem;
    int use_shadow_cache;
    unsigned long long *tile_counter;
    const unsigned *cubic_mask;  // global copy
    float sphere_r2_max;
};

__host__ __device__ inline size_t align16(size_t v) { return (v + 15) & ~(size_t) 15; }

// dynamic shared memory: [Pool][WaveHits][WaveCtl][scene arrays...]
__host__ __device__ inline size_t wave_fixed_smem() { return align16(sizeof(Pool)) + align16(sizeof(WaveHits)) + align16(sizeof(WaveCtl)); }

__host__ __device__ inline size_t wave_scene_smem(const SceneView &sc, bool spheres)
{
    return align16(sizeof(double4) * sc.n_plane) + (spheres ? align16(sizeof(float4) * sc.n_sphere) : 0) +
           align16(sizeof(double) * 10 * sc.n_quadric) + align16(sizeof(double) * 20 * sc.n_cubic) +
           align16(sizeof(unsigned) * sc.n_cubic) + align16(sizeof(LightDev) * sc.n_lights);
}

} // namespace rt
#endif
