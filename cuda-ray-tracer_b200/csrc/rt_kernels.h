// rt_kernels.h — host-callable launchers of the device kernels (implemented in rt_kernels.cu).
#ifndef RT_KERNELS_H
#define RT_KERNELS_H

#include "rt_scene.h"

namespace rt {

// Exact FP64 kernel, one thread per pixel (the known-good device path every faster kernel is diffed against).
cudaError_t launch_render_exact(const SceneView &scene, const FrameParams &frame, cudaStream_t stream);

// Block-local wavefront kernel (rt_wave.cu).
struct WaveLaunch {
    unsigned long long *tile_counter;   // device, one 64-bit slot
    const unsigned *cubic_mask;         // device, [n_cubic] non-zero coefficient masks
    float sphere_r2_max;                // max r^2 over sphere-class objects (filter error bound)
    int shadow_cache;                   // 1 / 0 / -1 = automatic (on when the scene has >= 8 objects)
    int variant;                        // 0 = automatic; 1/2/3 = 256-thread CTAs, min 1/2/3 per SM; 5 = 512-thread CTA
};
cudaError_t launch_render_wave(const SceneView &scene, const FrameParams &frame, const WaveLaunch &wl, cudaStream_t stream);

// Undo the strip interleave after the all-gather: gathered = [world][local_rows][width] -> frame[height][width].
cudaError_t launch_unpack_strips(const uchar4 *gathered, uchar4 *frame, int width, int height, int strip_rows,
                                 int world, int local_rows, cudaStream_t stream);

// Number of kernels a call to launch_render_* enqueues (for bench.py's gpu_launches accounting).
int render_launch_count(int kernel_kind);

} // namespace rt
#endif
