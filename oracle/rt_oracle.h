/* TEST INFRASTRUCTURE — plain-C restatement of the reference's per-pixel frame update (CPU).
 *
 * This is the parity ORACLE ("port" kind).  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline / --impl reference legs may load it; the product library never links or calls it.
 * It follows, operation for operation (IEEE double / float, no FMA contraction):
 *   /root/reference/include/surface_impl.h:16-172   intersect_ray, normal_vector, constants
 *   /root/reference/include/light_impl.h:17-49      shadow_ray, surface_color, reflect_ray
 *   /root/reference/src/update-cpu.cpp:45-139       get_color_and_object, render_pixel, update
 * and the GLM operation orders listed in SURVEY.md §8(c) (GLM itself is not under /root/reference).
 * Parity pinning: the reference has no tests or golden vectors, so this restatement is
 * pinned against the reference itself compiled unmodified (oracle/_ref/libref_cpu.so) — bit-exact on
 * every bundled scene and on random rays (tests/test_oracle.py) — and against the fixtures under
 * tests/golden/ that were generated from oracle/_ref by tests/golden/make_golden.py.
 */
#ifndef RT_ORACLE_H
#define RT_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Reference in-memory layouts (include/scene.h:8-15, include/light.h:6-13, include/surface.h:10-14). */
typedef struct {
    double coef[20]; /* x3,y3,z3,x2y,xy2,x2z,xz2,y2z,yz2,xyz, x2,y2,z2,xy,xz,yz, x,y,z,c */
    float reflection_ratio;
    float color[3];
} rto_object; /* 176 bytes */

typedef struct {
    uint8_t is_spherical;
    uint8_t pad0[7];
    double p[3];
    float light_color[3];
    uint32_t pad1;
} rto_light; /* 48 bytes */

/* What the reference would have executed, for cross-checking the GPU counters. */
typedef struct {
    uint64_t primary_rays;
    uint64_t reflection_rays;
    uint64_t shadow_rays;
    uint64_t nearest_tests;  /* (ray, object) tests in nearest-hit loops */
    uint64_t shadow_tests;   /* (ray, object) tests in shadow loops, up to and including the first occluder */
    uint64_t hits;           /* nearest-hit loops that found an object */
    uint64_t lit;            /* unshadowed light evaluations */
} rto_counters;

typedef struct rto_scene rto_scene;

rto_scene *rto_scene_create(uint32_t width, uint32_t height, double vertical_fov_rad, const float bg_color[3],
                            uint32_t max_reflections, const void *objects, size_t n_objects,
                            const void *lights, size_t n_lights);
void rto_scene_destroy(rto_scene *scene);

double rto_intersect_ray(const double coef[20], const double origin[3], const double dir[3]);
void rto_normal_vector(const double coef[20], const double pos[3], double out[3]);

/* One pixel, float RGB as the reference's render_pixel returns it. counters may be NULL. */
void rto_render_pixel(const rto_scene *scene, const double camera_matrix[16], int x, int y, float rgb[3],
                      rto_counters *counters);

/* Rows y0, y0+ystep, ... < y1 split over n_threads pthreads; rgb_out = packed W*3 float rows.
 * Returns elapsed milliseconds (pixel loop only, like the reference's update()). */
float rto_render_rows(const rto_scene *scene, const double camera_matrix[16], int y0, int y1, int ystep,
                      int n_threads, float *rgb_out, rto_counters *counters);

/* Arbitrary pixels (xy = n pairs). */
void rto_render_pixels(const rto_scene *scene, const double camera_matrix[16], const int *xy, size_t n,
                       float *rgb_out);

/* float RGB -> RGBA8 with the reference CUDA back-end's rounding, (int)(c*255.0f + 0.5f), alpha 255
 * (src/update-cuda.cu:149-156; GL's float->unorm8 conversion of the CPU back-end gives the same value). */
void rto_quantize_rgba8(const float *rgb, size_t n_pixels, uint8_t *rgba_out);

#ifdef __cplusplus
}
#endif
#endif
