/* TEST INFRASTRUCTURE — see rt_oracle.h.  Plain-C restatement of the reference CPU frame update.
 * Build with -O2 -ffp-contract=off (oracle/Makefile): every double/float operation below is meant to
 * round exactly where the reference's (GLM-expanded) expression rounds. */
#include "rt_oracle.h"

#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

/* include/surface_impl.h:16-19 */
static const double K_EPS = 1e-7;
static const double K_TWO_THIRD_PI = M_PI * 2.0 / 3.0;
static const double K_SHADOW_BIAS = 1e-2;
static const double K_MAX_T = 1e6;
static const float K_PI_F = (float) M_PI; /* == M_PIf32 */

enum { X3, Y3, Z3, X2Y, XY2, X2Z, XZ2, Y2Z, YZ2, XYZ, X2, Y2, Z2, XY, XZ, YZ, LX, LY, LZ, CC };

struct rto_scene {
    int width, height;
    double aspect_ratio, tan_half_fov;
    float bg[3];
    int max_reflections;
    size_t n_objects, n_lights;
    rto_object *objects;
    rto_light *lights;
};

typedef struct { double x, y, z; } d3;

/* ---- monomial pieces of F(o + t d), grouped by the power of t they multiply --------------------
 * surface_impl.h:25-44.  (a,b,c) name components of dir (d*) and origin (o*); products associate
 * left to right exactly as written. */
static inline double t3_aaa(double da, double db, double dc) { return da * db * dc; }
static inline double t0_aaa(double oa, double ob, double oc) { return oa * ob * oc; }
static inline double t2_cube(double oa, double da) { return 3.0 * oa * da * da; }
static inline double t1_cube(double oa, double da) { return 3.0 * oa * oa * da; }
/* (a0 + t a)^2 (b0 + t b) */
static inline double t2_sq_lin(double oa, double da, double ob, double db) { return da * (da * ob + 2.0 * oa * db); }
static inline double t1_sq_lin(double oa, double da, double ob, double db) { return oa * (oa * db + 2.0 * da * ob); }

/* surface_impl.h:21-155 */
double rto_intersect_ray(const double *k, const double *origin, const double *dir)
{
    const double ox = origin[0], oy = origin[1], oz = origin[2];
    const double dx = dir[0], dy = dir[1], dz = dir[2];

    double t3 = k[X3] * t3_aaa(dx, dx, dx)
                + k[Y3] * t3_aaa(dy, dy, dy)
                + k[Z3] * t3_aaa(dz, dz, dz)
                + k[X2Y] * t3_aaa(dx, dx, dy)
                + k[XY2] * t3_aaa(dx, dy, dy)
                + k[X2Z] * t3_aaa(dx, dx, dz)
                + k[XZ2] * t3_aaa(dx, dz, dz)
                + k[Y2Z] * t3_aaa(dy, dy, dz)
                + k[YZ2] * t3_aaa(dy, dz, dz)
                + k[XYZ] * t3_aaa(dx, dy, dz);
    double t2 = k[X3] * t2_cube(ox, dx)
                + k[Y3] * t2_cube(oy, dy)
                + k[Z3] * t2_cube(oz, dz)
                + k[X2Y] * t2_sq_lin(ox, dx, oy, dy)
                + k[XY2] * t2_sq_lin(oy, dy, ox, dx)
                + k[X2Z] * t2_sq_lin(ox, dx, oz, dz)
                + k[XZ2] * t2_sq_lin(oz, dz, ox, dx)
                + k[Y2Z] * t2_sq_lin(oy, dy, oz, dz)
                + k[YZ2] * t2_sq_lin(oz, dz, oy, dy)
                + k[XYZ] * (dx * dy * oz + dx * oy * dz + ox * dy * dz)
                + k[X2] * (dx * dx)
                + k[Y2] * (dy * dy)
                + k[Z2] * (dz * dz)
                + k[XY] * (dx * dy)
                + k[XZ] * (dx * dz)
                + k[YZ] * (dy * dz);
    double t1 = k[X3] * t1_cube(ox, dx)
                + k[Y3] * t1_cube(oy, dy)
                + k[Z3] * t1_cube(oz, dz)
                + k[X2Y] * t1_sq_lin(ox, dx, oy, dy)
                + k[XY2] * t1_sq_lin(oy, dy, ox, dx)
                + k[X2Z] * t1_sq_lin(ox, dx, oz, dz)
                + k[XZ2] * t1_sq_lin(oz, dz, ox, dx)
                + k[Y2Z] * t1_sq_lin(oy, dy, oz, dz)
                + k[YZ2] * t1_sq_lin(oz, dz, oy, dy)
                + k[XYZ] * (dx * oy * oz + ox * dy * oz + ox * oy * dz)
                + k[X2] * (2.0 * ox * dx)
                + k[Y2] * (2.0 * oy * dy)
                + k[Z2] * (2.0 * oz * dz)
                + k[XY] * (ox * dy + dx * oy)
                + k[XZ] * (ox * dz + dx * oz)
                + k[YZ] * (oy * dz + dy * oz)
                + k[LX] * dx + k[LY] * dy + k[LZ] * dz;
    double t0 = k[X3] * t0_aaa(ox, ox, ox)
                + k[Y3] * t0_aaa(oy, oy, oy)
                + k[Z3] * t0_aaa(oz, oz, oz)
                + k[X2Y] * t0_aaa(ox, ox, oy)
                + k[XY2] * t0_aaa(ox, oy, oy)
                + k[X2Z] * t0_aaa(ox, ox, oz)
                + k[XZ2] * t0_aaa(ox, oz, oz)
                + k[Y2Z] * t0_aaa(oy, oy, oz)
                + k[YZ2] * t0_aaa(oy, oz, oz)
                + k[XYZ] * t0_aaa(ox, oy, oz)
                + k[X2] * (ox * ox)
                + k[Y2] * (oy * oy)
                + k[Z2] * (oz * oz)
                + k[XY] * (ox * oy)
                + k[XZ] * (ox * oz)
                + k[YZ] * (oy * oz)
                + k[LX] * ox + k[LY] * oy + k[LZ] * oz + k[CC];

    if (fabs(t3) > K_EPS) {
        /* cubic: normalise, depressed-cubic invariants, Cardano or trigonometric (surface_impl.h:106-137) */
        t2 /= t3;
        t1 /= t3;
        t0 /= t3;
        double q = (3.0 * t1 - t2 * t2) / 9.0;
        double r = (9.0 * t2 * t1 - 27.0 * t0 - 2.0 * t2 * t2 * t2) / 54.0;
        double delta = q * q * q + r * r;
        if (delta > 0) {
            delta = sqrt(delta);
            q = cbrt(r + delta);
            r = cbrt(r - delta);
            return q + r - t2 / 3.0; /* single real root, returned unfiltered */
        }
        double theta = acos(r / sqrt(-q * q * q)) / 3.0;
        double c = 2.0 * sqrt(-q);
        double x = c * cos(theta) - t2 / 3.0;
        double x1 = c * cos(theta + K_TWO_THIRD_PI) - t2 / 3.0;
        if (x1 >= K_EPS && x1 < x) x = x1;
        x1 = c * cos(theta + 2.0 * K_TWO_THIRD_PI) - t2 / 3.0;
        if (x1 >= K_EPS && x1 < x) x = x1;
        return x;
    }
    if (fabs(t2) > K_EPS) {
        /* quadratic (surface_impl.h:138-149) */
        double delta = t1 * t1 - 4.0 * t2 * t0;
        if (delta < 0) return -1.0;
        delta = sqrt(delta);
        double x = (-t1 - delta) / (2.0 * t2);
        if (x >= K_EPS) return x;
        return (-t1 + delta) / (2.0 * t2);
    }
    if (fabs(t1) > K_EPS) return -t0 / t1; /* linear (surface_impl.h:150-153) */
    return -1.0;
}

static inline double dot3(d3 a, d3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }

static inline d3 normalize3(d3 v)
{
    double inv = 1.0 / sqrt(dot3(v, v));
    d3 r = {v.x * inv, v.y * inv, v.z * inv};
    return r;
}

/* surface_impl.h:157-172 */
static d3 normal_at(const double *k, d3 p)
{
    d3 g;
    g.x = 3.0 * k[X3] * p.x * p.x + 2.0 * k[X2] * p.x + k[LX];
    g.y = 3.0 * k[Y3] * p.y * p.y + 2.0 * k[Y2] * p.y + k[LY];
    g.z = 3.0 * k[Z3] * p.z * p.z + 2.0 * k[Z2] * p.z + k[LZ];
    g.x += 2.0 * p.x * (k[X2Y] * p.y + k[X2Z] * p.z)
           + p.y * (k[XY2] * p.y + k[XYZ] * p.z + k[XY])
           + p.z * (k[XZ2] * p.z + k[XZ]);
    g.y += 2.0 * p.y * (k[XY2] * p.x + k[Y2Z] * p.z)
           + p.x * (k[X2Y] * p.x + k[XYZ] * p.z + k[XY])
           + p.z * (k[YZ2] * p.z + k[YZ]);
    g.z += 2.0 * p.z * (k[XZ2] * p.x + k[YZ2] * p.y)
           + p.x * (k[X2Z] * p.x + k[XYZ] * p.y + k[XZ])
           + p.y * (k[Y2Z] * p.y + k[YZ]);
    return normalize3(g);
}

void rto_normal_vector(const double *coef, const double *pos, double *out)
{
    d3 p = {pos[0], pos[1], pos[2]};
    d3 n = normal_at(coef, p);
    out[0] = n.x;
    out[1] = n.y;
    out[2] = n.z;
}

/* light_impl.h:29-44 */
static void shade_light(const rto_light *l, d3 p, d3 n, const float *obj_color, float *acc)
{
    d3 dir;
    float lc[3];
    if (l->is_spherical) {
        dir.x = l->p[0] - p.x;
        dir.y = l->p[1] - p.y;
        dir.z = l->p[2] - p.z;
        float denom = 4.0f * K_PI_F * (float) dot3(dir, dir);
        lc[0] = l->light_color[0] / denom;
        lc[1] = l->light_color[1] / denom;
        lc[2] = l->light_color[2] / denom;
        dir = normalize3(dir);
    }
    else {
        dir.x = l->p[0];
        dir.y = l->p[1];
        dir.z = l->p[2];
        lc[0] = l->light_color[0];
        lc[1] = l->light_color[1];
        lc[2] = l->light_color[2];
    }
    float cosv = (float) dot3(n, dir);
    float lambert = (0.0f < cosv) ? cosv : 0.0f;
    for (int c = 0; c < 3; c++) acc[c] += obj_color[c] / K_PI_F * lc[c] * lambert;
}

/* update-cpu.cpp:45-80.  Returns the hit object index or -1; on a hit fills colour, point, normal. */
static int trace_and_shade(const rto_scene *s, d3 o, d3 d, float *color, d3 *point, d3 *normal, rto_counters *cnt)
{
    int best = -1;
    double best_t = INFINITY;
    const double ov[3] = {o.x, o.y, o.z}, dv[3] = {d.x, d.y, d.z};
    for (size_t i = 0; i < s->n_objects; i++) {
        double t = rto_intersect_ray(s->objects[i].coef, ov, dv);
        if (t >= K_EPS && t < K_MAX_T && t < best_t) {
            best_t = t;
            best = (int) i;
        }
    }
    if (cnt) cnt->nearest_tests += s->n_objects;
    if (best < 0) return best;
    if (cnt) cnt->hits++;

    color[0] = color[1] = color[2] = 0.0f;
    d3 p = {o.x + best_t * d.x, o.y + best_t * d.y, o.z + best_t * d.z};
    d3 n = normal_at(s->objects[best].coef, p);
    const float *obj_color = s->objects[best].color;
    for (size_t j = 0; j < s->n_lights; j++) {
        const rto_light *l = &s->lights[j];
        /* light_impl.h:17-27 — the shadow direction is rounded to float, then widened again */
        double max_t;
        float sdf[3];
        if (l->is_spherical) {
            max_t = 1.0f;
            sdf[0] = (float) (l->p[0] - p.x);
            sdf[1] = (float) (l->p[1] - p.y);
            sdf[2] = (float) (l->p[2] - p.z);
        }
        else {
            max_t = 1e6;
            sdf[0] = (float) l->p[0];
            sdf[1] = (float) l->p[1];
            sdf[2] = (float) l->p[2];
        }
        const double sd[3] = {sdf[0], sdf[1], sdf[2]};
        const double so[3] = {p.x + K_SHADOW_BIAS * n.x, p.y + K_SHADOW_BIAS * n.y, p.z + K_SHADOW_BIAS * n.z};
        if (cnt) cnt->shadow_rays++;
        int occluded = 0;
        for (size_t k = 0; k < s->n_objects; k++) {
            double t = rto_intersect_ray(s->objects[k].coef, so, sd);
            if (cnt) cnt->shadow_tests++;
            if (t > K_EPS && t < max_t) {
                occluded = 1;
                break;
            }
        }
        if (!occluded) {
            shade_light(l, p, n, obj_color, color);
            if (cnt) cnt->lit++;
        }
    }
    for (int c = 0; c < 3; c++) color[c] = (color[c] < 1.0f) ? color[c] : 1.0f;
    *point = p;
    *normal = n;
    return best;
}

static void mat_mul_point(const double *m, double x, double y, double z, double w, double out[4])
{
    /* GLM: (m[0]*x + m[1]*y) + (m[2]*z + m[3]*w), column-major m[col*4+row] */
    for (int r = 0; r < 4; r++) out[r] = (m[r] * x + m[4 + r] * y) + (m[8 + r] * z + m[12 + r] * w);
}

/* update-cpu.cpp:82-119 */
void rto_render_pixel(const rto_scene *s, const double *m, int px, int py, float *rgb, rto_counters *cnt)
{
    double o4[4], v4[4];
    mat_mul_point(m, 0.0, 0.0, 0.0, 1.0, o4); /* update-cpu.cpp:123 */
    d3 origin = {o4[0], o4[1], o4[2]};

    double ndc_x = (px + 0.5) / s->width;
    double ndc_y = (py + 0.5) / s->height;
    double cam_x = (2.0 * ndc_x - 1.0) * s->aspect_ratio * s->tan_half_fov;
    double cam_y = (2.0 * ndc_y - 1.0) * s->tan_half_fov;
    mat_mul_point(m, cam_x, cam_y, 1.0, 1.0, v4);
    d3 dir = {v4[0] - origin.x, v4[1] - origin.y, v4[2] - origin.z};
    dir = normalize3(dir);
    if (cnt) cnt->primary_rays++;

    float hit_color[3];
    d3 p, n;
    int idx = trace_and_shade(s, origin, dir, hit_color, &p, &n, cnt);
    if (idx < 0) {
        rgb[0] = s->bg[0];
        rgb[1] = s->bg[1];
        rgb[2] = s->bg[2];
        return;
    }
    float out[3] = {hit_color[0], hit_color[1], hit_color[2]};
    float ratio = 1.0f;
    int bounces = 0;
    while (s->objects[idx].reflection_ratio > K_EPS) {
        ratio *= s->objects[idx].reflection_ratio;
        const float *blend = s->bg;
        int stop = 1;
        if (bounces != s->max_reflections) {
            bounces++;
            double twice = 2.0 * dot3(dir, n); /* light_impl.h:46-49 */
            d3 rd = {dir.x - twice * n.x, dir.y - twice * n.y, dir.z - twice * n.z};
            d3 ro = {p.x + K_SHADOW_BIAS * n.x, p.y + K_SHADOW_BIAS * n.y, p.z + K_SHADOW_BIAS * n.z};
            dir = rd;
            if (cnt) cnt->reflection_rays++;
            idx = trace_and_shade(s, ro, dir, hit_color, &p, &n, cnt);
            if (idx >= 0) {
                blend = hit_color;
                stop = 0;
            }
        }
        for (int c = 0; c < 3; c++) out[c] = (1.0f - ratio) * out[c] + ratio * blend[c];
        if (stop) break;
    }
    rgb[0] = out[0];
    rgb[1] = out[1];
    rgb[2] = out[2];
}

rto_scene *rto_scene_create(uint32_t width, uint32_t height, double vertical_fov_rad, const float *bg_color,
                            uint32_t max_reflections, const void *objects, size_t n_objects, const void *lights,
                            size_t n_lights)
{
    rto_scene *s = (rto_scene *) calloc(1, sizeof(rto_scene));
    if (!s) return NULL;
    /* update-cpu.cpp:22-32 */
    s->width = (int) width;
    s->height = (int) height;
    s->aspect_ratio = (double) width / height;
    s->tan_half_fov = tan(0.5 * vertical_fov_rad);
    memcpy(s->bg, bg_color, sizeof(s->bg));
    s->max_reflections = (int) max_reflections;
    s->n_objects = n_objects;
    s->n_lights = n_lights;
    s->objects = (rto_object *) malloc(sizeof(rto_object) * (n_objects ? n_objects : 1));
    s->lights = (rto_light *) malloc(sizeof(rto_light) * (n_lights ? n_lights : 1));
    if (n_objects) memcpy(s->objects, objects, sizeof(rto_object) * n_objects);
    if (n_lights) memcpy(s->lights, lights, sizeof(rto_light) * n_lights);
    return s;
}

void rto_scene_destroy(rto_scene *s)
{
    if (!s) return;
    free(s->objects);
    free(s->lights);
    free(s);
}

typedef struct {
    const rto_scene *scene;
    const double *m;
    const int *rows;
    size_t n_rows;
    int tid, n_threads;
    float *out;
    rto_counters cnt;
    int count;
} row_job;

static void *row_worker(void *arg)
{
    row_job *j = (row_job *) arg;
    const int w = j->scene->width;
    for (size_t k = (size_t) j->tid; k < j->n_rows; k += (size_t) j->n_threads) {
        for (int x = 0; x < w; x++) {
            rto_render_pixel(j->scene, j->m, x, j->rows[k], j->out + 3 * (k * (size_t) w + (size_t) x),
                             j->count ? &j->cnt : NULL);
        }
    }
    return NULL;
}

float rto_render_rows(const rto_scene *s, const double *m, int y0, int y1, int ystep, int n_threads,
                      float *rgb_out, rto_counters *counters)
{
    if (ystep < 1) ystep = 1;
    if (n_threads < 1) n_threads = 1;
    size_t n_rows = 0;
    for (int y = y0; y < y1; y += ystep) n_rows++;
    int *rows = (int *) malloc(sizeof(int) * (n_rows ? n_rows : 1));
    size_t k = 0;
    for (int y = y0; y < y1; y += ystep) rows[k++] = y;

    row_job *jobs = (row_job *) calloc((size_t) n_threads, sizeof(row_job));
    pthread_t *threads = (pthread_t *) calloc((size_t) n_threads, sizeof(pthread_t));
    struct timespec a, b;
    clock_gettime(CLOCK_MONOTONIC, &a);
    for (int t = 0; t < n_threads; t++) {
        jobs[t].scene = s;
        jobs[t].m = m;
        jobs[t].rows = rows;
        jobs[t].n_rows = n_rows;
        jobs[t].tid = t;
        jobs[t].n_threads = n_threads;
        jobs[t].out = rgb_out;
        jobs[t].count = counters != NULL;
        if (t > 0) pthread_create(&threads[t], NULL, row_worker, &jobs[t]);
    }
    row_worker(&jobs[0]);
    for (int t = 1; t < n_threads; t++) pthread_join(threads[t], NULL);
    clock_gettime(CLOCK_MONOTONIC, &b);
    if (counters) {
        memset(counters, 0, sizeof(*counters));
        for (int t = 0; t < n_threads; t++) {
            counters->primary_rays += jobs[t].cnt.primary_rays;
            counters->reflection_rays += jobs[t].cnt.reflection_rays;
            counters->shadow_rays += jobs[t].cnt.shadow_rays;
            counters->nearest_tests += jobs[t].cnt.nearest_tests;
            counters->shadow_tests += jobs[t].cnt.shadow_tests;
            counters->hits += jobs[t].cnt.hits;
            counters->lit += jobs[t].cnt.lit;
        }
    }
    free(jobs);
    free(threads);
    free(rows);
    return (float) ((b.tv_sec - a.tv_sec) * 1e3 + (b.tv_nsec - a.tv_nsec) * 1e-6);
}

void rto_render_pixels(const rto_scene *s, const double *m, const int *xy, size_t n, float *rgb_out)
{
    for (size_t i = 0; i < n; i++) rto_render_pixel(s, m, xy[2 * i], xy[2 * i + 1], rgb_out + 3 * i, NULL);
}

void rto_quantize_rgba8(const float *rgb, size_t n_pixels, uint8_t *rgba_out)
{
    for (size_t i = 0; i < n_pixels; i++) {
        for (int c = 0; c < 3; c++) rgba_out[4 * i + c] = (uint8_t) (int) (rgb[3 * i + c] * 255.0f + 0.5f);
        rgba_out[4 * i + 3] = 255;
    }
}
