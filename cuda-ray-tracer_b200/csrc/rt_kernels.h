// rt_kernels.h — host-callable launchers of the device kernels (implemented in rt_kernels.cu).
#ifndef RT_KERNELS_H
#define RT_KERNELS_H

#include "rt_scene.h"

namespace rt {

// Exact FP64 kernel, one thread per pixel (the known-good device path every faster kernel is diffed against).
cudaError_t launch_render_exact(const SceneView &scene, const FrameParams &frame, cudaStream_t stream);

// Block-local wavefront kernel (rt_wave.cu).
struct WaveLaunch {
    unsigned long long *tile_counter;   // device, two 64-bit slots, zero between launches
    const unsigned *cubic_mask;         // device, [n_cubic] non-zero coefficient masks
    float sphere_r2_max;                // max r^2 over sphere-class objects (filter error bound)
    int shadow_cache;                   // 1 / 0 / -1 = automatic (on when the scene has >= 8 objects)
    int variant;                        // 0 = automatic; 2 / 3 = 256-thread CTAs, two / three per SM; 5 = one 512-thread CTA
    int cubic_filter;                   // 1 = FP32 filter in front of the exact cubic test (rt_cubic_filter.cuh)
    int cull_backfacing;                // 1 = no shadow ray towards a light whose Lambert factor is 0 at the hit
    int accel;                          // 1 = sphere class through the uniform grid (SceneView::grid), 0 = brute force
    // fused multi-GPU gather: sequence words in rank 0's memory (go == nullptr: none), see WaveSync in rt_wave.cuh
    unsigned *sync_go, *sync_flags, *sync_timed_out;
    unsigned sync_seq;
    int sync_rank, sync_world;          // this GPU's rank in the job (FrameParams may describe the whole image as one rank's)
    // dynamic ownership of row groups across GPUs (WaveDyn in rt_wave.cuh); dyn_counters == nullptr: static strips
    unsigned long long *dyn_counters;   // [2] in rank 0's memory
    int *dyn_slot_group;                // this GPU's slot -> group table
    int dyn_n_groups, dyn_group_rows, dyn_max_slots, dyn_ahead;
    int lone_kernel = 1;                // 1 = scenes of exactly one cubic object run the lone-cubic instantiation (option "lone_kernel")
    int spheres_smem = -1;              // -1 = automatic, 0 = never stage the FP32 sphere records in shared memory (option "spheres_smem")
};
// (re)initialise a slot -> group table: the first `ahead` slots statically, the rest "not fetched yet"
cudaError_t launch_dyn_table_init(int *slot_group, int max_slots, int ahead, int n_groups, int rank, int world, cudaStream_t stream);
cudaError_t launch_render_wave(const SceneView &scene, const FrameParams &frame, const WaveLaunch &wl, cudaStream_t stream);
bool wave_scene_fits(const SceneView &scene);   // false: too many lights / quadrics / cubics for the per-CTA scene copy
int wave_tile_rows(const SceneView &scene, const WaveLaunch &wl);   // 8 or 16: tile height of the variant that launch would use

// Undo the strip interleave after the all-gather: gathered = [world][local_rows][width] -> frame[height][width].
cudaError_t launch_unpack_strips(const uchar4 *gathered, uchar4 *frame, int width, int height, int strip_rows,
                                 int world, int local_rows, cudaStream_t stream);

// Fused-gather signalling through sequence words in peer-mapped memory (rt_kernels.cu).
cudaError_t launch_sync_store(unsigned *word, unsigned value, cudaStream_t stream);
cudaError_t launch_sync_wait(const unsigned *words, int first, int n, unsigned need, unsigned *timed_out, cudaStream_t stream);

} // namespace rt
#endif
