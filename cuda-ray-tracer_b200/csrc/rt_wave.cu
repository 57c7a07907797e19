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
// Two translation units instantiate the kernel so that they compile in parallel: this one with the cubic class
// (CUBIC = true) plus the dispatcher, rt_wave_plain.cu without it.
#include "rt_wave_device.cuh"
#include "rt_wave_kernel.inl"

namespace rt {

cudaError_t launch_render_wave_plain(const SceneView &scene, const FrameParams &frame, const WaveLaunch &wl, cudaStream_t stream);

cudaError_t launch_render_wave(const SceneView &scene, const FrameParams &frame, const WaveLaunch &wl, cudaStream_t stream)
{
    return scene.n_cubic > 0 ? launch_render_wave_t<true>(scene, frame, wl, stream)
                             : launch_render_wave_plain(scene, frame, wl, stream);
}

// Does the wavefront kernel's per-CTA copy of the scene (planes, quadrics, cubics + filter tables, lights; the sphere
// filter array falls back to global memory on its own) fit the device's shared memory?  The limits are a few thousand
// lights or quadrics, a few hundred cubics.  When it does not, rt_api.cu renders with the exact per-pixel kernel, which
// reads the scene from global memory and has no such limit — the reference accepts any object and light count.
bool wave_scene_fits(const SceneView &scene)
{
    int dev = 0, max_smem = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return false;
    cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    return wave_fixed_smem<256>() + wave_scene_smem(scene, false) + wave_cubic_smem(scene, 256) <= (size_t) max_smem;
}

// Tile height (rows) of the variant launch_render_wave will use: the dynamic split hands out row groups that are a
// multiple of it.
int wave_tile_rows(const SceneView &scene, const WaveLaunch &wl)
{
    int dev = 0, max_smem = 0, sm_smem = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 16;
    cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    cudaDeviceGetAttribute(&sm_smem, cudaDevAttrMaxSharedMemoryPerMultiprocessor, dev);
    return wave_pick_variant(scene, wl, max_smem, sm_smem) == 5 ? 16 : 8;
}

} // namespace rt
