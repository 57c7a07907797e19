// rt_wave_grid.cu — the wavefront kernel with the sphere-grid walk compiled in (GRID = true), scenes with cubic-class
// objects.  See rt_wave.cu for the kernel and rt_grid.cuh for the walk.
#define RT_WAVE_GRID_TU launch_wave_grid_cubic
#define RT_WAVE_GRID_CUBIC true
#include "rt_wave_device.cuh"
#include "rt_wave_kernel.inl"
