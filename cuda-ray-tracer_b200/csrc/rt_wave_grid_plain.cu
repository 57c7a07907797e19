// rt_wave_grid_plain.cu — the wavefront kernel with the sphere-grid walk compiled in (GRID = true), scenes without
// cubic-class objects.  See rt_wave.cu for the kernel and rt_grid.cuh for the walk.
#define RT_WAVE_GRID_TU launch_wave_grid_plain
#define RT_WAVE_GRID_CUBIC false
#include "rt_wave_device.cuh"
#include "rt_wave_kernel.inl"
