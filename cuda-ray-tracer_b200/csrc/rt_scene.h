// rt_scene.h — device-side scene layout (SoA by structural class) and per-frame parameters.
//
// The reference keeps the scene as AoS Object[] / LightSource[] in global memory and evaluates the full
// 20-coefficient expansion for every (ray, object) pair (src/update-cuda.cu:20-22,71-77,
// include/surface_impl.h:46-103).  Here rt_b200_upload_scene classifies every object once by its
// STRUCTURALLY non-zero coefficients and repacks each class into its own dense array:
//
//   plane    only x,y,z,c                      double4 {x, y, z, c}
//   sphere   x2 = y2 = z2 = 1, x,y,z,c         double4 {x, y, z, c}   (+ float4 {cx,cy,cz,r^2} FP32 filter)
//   quadric  any other degree-2 polynomial     10 doubles {x2,y2,z2,xy,xz,yz,x,y,z,c}
//   cubic    any third-order term              20 doubles (reference order)
//
// Dropping terms whose coefficient is exactly 0 does not change any rounded intermediate of the
// reference's left-to-right sums (x + 0 == x), so per-class expansion stays bit-faithful to
// intersect_ray() while doing 13 / 38 / 81 / 320 flops instead of 320 for everything.
#ifndef RT_SCENE_H
#define RT_SCENE_H

#include <cuda_runtime.h>
#include <stdint.h>

#include "rt_grid.cuh"

namespace rt {

enum ObjClass { CLS_PLANE = 0, CLS_SPHERE = 1, CLS_QUADRIC = 2, CLS_CUBIC = 3 };

// Index of each coefficient inside the reference's SurfaceCoefs (include/surface.h:10-14).
enum Coef { X3, Y3, Z3, X2Y, XY2, X2Z, XZ2, Y2Z, YZ2, XYZ, X2, Y2, Z2, XY, XZ, YZ, LX, LY, LZ, CC };

struct LightDev {
    double px, py, pz;   // -normalised direction (directional) or position (spherical), light.cpp:12,24
    float r, g, b;       // intensity * colour
    int is_spherical;
};

struct SceneView {
    int n_plane, n_sphere, n_quadric, n_cubic;
    int n_objects, n_lights;
    const double4 *plane;     // [n_plane]
    const double4 *sphere;    // [n_sphere]   exact coefficients {x, y, z, c}
    const float4 *sphere_f;   // [n_sphere]   FP32 filter {cx, cy, cz, r^2 (rounded up)}
    const double *quadric;    // [n_quadric][10]
    const double *cubic;      // [n_cubic][20]
    const int *orig_index;    // [n_objects]  class-sorted slot -> original object index
    const double *coef_full;  // [n_objects][20] by ORIGINAL index (normal_vector at the hit)
    const float4 *material;   // [n_objects] by ORIGINAL index {reflection_ratio, r, g, b}
    const float4 *material_pi;   // [n_objects] {reflection_ratio, r / pi, g / pi, b / pi}: light_impl.h:43 divides the
                                 // object colour by pi for every light; the wavefront kernel reads the quotients
    const LightDev *lights;   // [n_lights]
    float sphere_center_max;  // max |coordinate| over all sphere centres (FP32 filter error bound)
    SphereGrid grid;          // uniform grid over the sphere class (rt_grid.cuh); cell_start == nullptr: none
};

// Work counters, one 64-bit slot each; order matches rt_b200_counters in include/rt_b200.h.
enum Counter {
    CNT_PRIMARY, CNT_REFLECTION, CNT_SHADOW, CNT_HITS, CNT_LIT,
    CNT_T_PLANE, CNT_T_SPHERE, CNT_T_QUADRIC, CNT_T_CUBIC, CNT_CONFIRM, CNT_CUBIC_EXACT, CNT_COUNT
};

struct FrameParams {
    double m[16];          // camera matrix, GLM column-major
    double origin[3];      // m * (0,0,0,1), update-cpu.cpp:123
    double aspect;         // W / H
    double tan_half_fov;   // tan(fov / 2)
    float bg[3];
    int max_reflections;
    int width, height;
    // multi-GPU interleave: this rank renders strips s with s % world == rank, packed contiguously
    int strip_rows, rank, world;
    int local_rows;        // rows in this rank's packed buffer (including padding rows past the image)
    int row0, rows;        // this launch renders the packed rows [row0, row0 + rows) (a band; whole frame: 0, local_rows)
    int direct;            // 0: out is this rank's packed buffer [local_rows][width]
                           // 1: out is the assembled frame [height][width] (possibly in a peer GPU's memory): pixels
                           //    are stored at their image position and padding rows are not written
    uchar4 *out;
    unsigned long long *counters;  // [CNT_COUNT] or nullptr
    unsigned long long *timeline;  // diagnostic build only (-DRT_B200_TIMELINE, option "timeline"): [CTA][RT_TL_WORDS] or nullptr
};

#define RT_TL_WORDS 8      /* per CTA: global-timer ns at entry, after the prologue, when no tile was left, at loop exit, at kernel
                              exit; waves run; SM id; tiles rendered */
#define RT_TL_MAX_CTAS 2048

// local (packed) row -> image row; may be >= height for padding rows
__host__ __device__ inline int local_to_image_row(int local_row, int strip_rows, int rank, int world)
{
    int s_local = local_row / strip_rows;
    int r = local_row - s_local * strip_rows;
    return (s_local * world + rank) * strip_rows + r;
}

// index into FrameParams::out of pixel x of packed row lrow (image row y)
__host__ __device__ inline size_t out_index(const FrameParams &f, int lrow, int y, int x)
{
    return (size_t) (f.direct ? y : lrow) * f.width + x;
}

__host__ __device__ inline size_t align16(size_t v) { return (v + 15) & ~(size_t) 15; }

} // namespace rt
#endif
