// rt_wave_plain.cu — the wavefront kernel compiled without the cubic class (CUBIC = false): a leaner kernel for
// scenes of planes, spheres and quadrics.  See rt_wave.cu.
#include "rt_wave_device.cuh"
#include "rt_wave_kernel.inl"

namespace rt {

cudaError_t launch_render_wave_plain(const SceneView &scene, const FrameParams &frame, const WaveLaunch &wl, cudaStream_t stream)
{
    return launch_render_wave_t<false>(scene, frame, wl, stream);
}

} // namespace rt
