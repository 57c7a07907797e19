// TEST INFRASTRUCTURE — C-callable harness around the UNMODIFIED reference CPU back-end.
//
// oracle/Makefile compiles this file together with /root/reference/src/{update-cpu,surface,light,
// scene-exception}.cpp (where they lie; nothing is copied) against the header shims in oracle/shim/
// into oracle/_ref/libref_cpu.so.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
// --impl reference legs may load that library.  The product (librt_b200.so) never does.
//
// The harness only adds what the reference keeps in files that need yaml-cpp (src/scene.cpp:9-22, the
// Object constructor) and a way in and out of the reference's file-scope globals
// (src/update-cpu.cpp:10-19), which have external linkage.
#include <cstdint>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include <GLFW/glfw3.h>

#include "scene-exception.h"
#include "update.h"

rt_shim_texture_upload g_rt_shim_last_upload = {0, 0, 0, nullptr};

// src/scene.cpp:9-14 is not compiled (it needs yaml-cpp); same semantics: validate, then store.
Object::Object(SurfaceCoefs surface, float reflection_ratio, const glm::vec3 &color)
        : surface{surface}, reflection_ratio{reflection_ratio}, color{color}
{
    validate_positive("object reflection ratio", reflection_ratio);
    validate_color(color);
}

// Globals and functions of src/update-cpu.cpp with external linkage.
extern int g_width, g_height;
extern std::vector<float> g_data;
extern glm::dvec3 g_ray_origin;
glm::vec3 render_pixel(const glm::dmat4 &camera_matrix, int pixel_x, int pixel_y);
// include/surface_impl.h defines these as non-inline functions, so they are linked from update-cpu.o
// rather than by including the header a second time.
double intersect_ray(const SurfaceCoefs &coef, const glm::dvec3 &origin, const glm::dvec3 &dir);
glm::dvec3 normal_vector(const SurfaceCoefs &coef, const glm::dvec3 &pos);

static_assert(sizeof(SurfaceCoefs) == 160, "SurfaceCoefs layout");
static_assert(sizeof(Object) == 176, "Object layout");
static_assert(offsetof(Object, reflection_ratio) == 160, "Object layout");
static_assert(offsetof(Object, color) == 164, "Object layout");
static_assert(sizeof(LightSource) == 48, "LightSource layout");
static_assert(offsetof(LightSource, p) == 8, "LightSource layout");
static_assert(offsetof(LightSource, light_color) == 32, "LightSource layout");

static std::string g_error;

static glm::dmat4 to_mat(const double *m)
{
    glm::dmat4 r;
    for (int c = 0; c < 4; c++) {
        r[c] = glm::dvec4(m[4 * c], m[4 * c + 1], m[4 * c + 2], m[4 * c + 3]);
    }
    return r;
}

extern "C" {

const char *ref_last_error() { return g_error.c_str(); }

// objects: n_objects records of 176 bytes, lights: n_lights records of 48 bytes, the reference layouts.
int ref_init(uint32_t width, uint32_t height, double vertical_fov_rad, const float *bg_color,
             uint32_t max_reflections, const void *objects, size_t n_objects, const void *lights,
             size_t n_lights)
{
    try {
        Scene scene;
        scene.px_width = width;
        scene.px_height = height;
        scene.vertical_fov = vertical_fov_rad;
        scene.bg_color = glm::vec3(bg_color[0], bg_color[1], bg_color[2]);
        scene.max_reflections = max_reflections;
        for (size_t i = 0; i < n_objects; i++) {
            SurfaceCoefs coefs;
            float refl;
            float color[3];
            const char *rec = static_cast<const char *>(objects) + 176 * i;
            std::memcpy(&coefs, rec, 160);
            std::memcpy(&refl, rec + 160, 4);
            std::memcpy(color, rec + 164, 12);
            scene.objects.push_back(Object(coefs, refl, glm::vec3(color[0], color[1], color[2])));
        }
        for (size_t i = 0; i < n_lights; i++) {
            LightSource light{};
            std::memcpy(&light, static_cast<const char *>(lights) + 48 * i, 48);
            scene.lights.push_back(light);
        }
        init_update(0, scene);
    }
    catch (std::exception &e) {
        g_error = e.what();
        return 1;
    }
    return 0;
}

// One full frame through the reference's own update(); rgb_out (may be null) gets W*H*3 floats in
// the reference's g_data order (row 0 = bottom).  Returns the reference's own elapsed ms.
float ref_update(const double *camera_matrix, float *rgb_out)
{
    float ms = update(to_mat(camera_matrix));
    if (rgb_out != nullptr) {
        std::memcpy(rgb_out, g_rt_shim_last_upload.pixels, sizeof(float) * 3 * (size_t) g_width * g_height);
    }
    return ms;
}

// Rows y0, y0+ystep, ... < y1 through the reference's render_pixel(), split over n_threads host
// threads (render_pixel only reads the globals once g_ray_origin is set, update-cpu.cpp:82-119).
// rgb_out holds one W*3 float row per rendered row, packed in order.  Returns elapsed ms.
float ref_render_rows(const double *camera_matrix, int y0, int y1, int ystep, int n_threads, float *rgb_out)
{
    const glm::dmat4 m = to_mat(camera_matrix);
    g_ray_origin = m * glm::dvec4(0.0, 0.0, 0.0, 1.0);
    std::vector<int> rows;
    for (int y = y0; y < y1; y += ystep) rows.push_back(y);
    if (n_threads < 1) n_threads = 1;
    const int width = g_width;
    auto work = [&](int tid) {
        for (size_t k = tid; k < rows.size(); k += n_threads) {
            for (int x = 0; x < width; x++) {
                glm::vec3 c = render_pixel(m, x, rows[k]);
                float *o = rgb_out + 3 * (k * width + x);
                o[0] = c.r;
                o[1] = c.g;
                o[2] = c.b;
            }
        }
    };
    const double start = glfwGetTime();
    std::vector<std::thread> pool;
    for (int t = 1; t < n_threads; t++) pool.emplace_back(work, t);
    work(0);
    for (auto &t: pool) t.join();
    return (float) ((glfwGetTime() - start) * 1000.0);
}

// Arbitrary pixel subset (xy = n pairs) — the sparse oracle for frames too big to render on the CPU.
void ref_render_pixels(const double *camera_matrix, const int *xy, size_t n, float *rgb_out)
{
    const glm::dmat4 m = to_mat(camera_matrix);
    g_ray_origin = m * glm::dvec4(0.0, 0.0, 0.0, 1.0);
    for (size_t i = 0; i < n; i++) {
        glm::vec3 c = render_pixel(m, xy[2 * i], xy[2 * i + 1]);
        rgb_out[3 * i] = c.r;
        rgb_out[3 * i + 1] = c.g;
        rgb_out[3 * i + 2] = c.b;
    }
}

// Per-function entry points for known-answer tests.
double ref_intersect_ray(const double *coefs, const double *origin, const double *dir)
{
    SurfaceCoefs c;
    std::memcpy(&c, coefs, 160);
    return intersect_ray(c, glm::dvec3(origin[0], origin[1], origin[2]), glm::dvec3(dir[0], dir[1], dir[2]));
}

void ref_normal_vector(const double *coefs, const double *pos, double *out)
{
    SurfaceCoefs c;
    std::memcpy(&c, coefs, 160);
    glm::dvec3 n = normal_vector(c, glm::dvec3(pos[0], pos[1], pos[2]));
    out[0] = n.x;
    out[1] = n.y;
    out[2] = n.z;
}

// Factories (src/surface.cpp:4-60, src/light.cpp:4-26).  kind: 0 sphere(a=center, s=radius), 1 plane(a=origin,
// b=normal), 2 dingDong(a=origin), 3 clebsch, 4 cayley.  Returns 0, or 1 with ref_last_error() set.
int ref_make_surface(int kind, const double *a, const double *b, double s, double *coefs_out)
{
    try {
        SurfaceCoefs c{};
        glm::dvec3 va(0.0), vb(0.0);
        if (a) va = glm::dvec3(a[0], a[1], a[2]);
        if (b) vb = glm::dvec3(b[0], b[1], b[2]);
        switch (kind) {
            case 0: c = SurfaceCoefs::sphere(va, s); break;
            case 1: c = SurfaceCoefs::plane(va, vb); break;
            case 2: c = SurfaceCoefs::dingDong(va); break;
            case 3: c = SurfaceCoefs::clebsch(); break;
            case 4: c = SurfaceCoefs::cayley(); break;
            default: g_error = "unknown surface kind"; return 1;
        }
        std::memcpy(coefs_out, &c, 160);
    }
    catch (std::exception &e) {
        g_error = e.what();
        return 1;
    }
    return 0;
}

int ref_make_object(const double *coefs, float reflection_ratio, const float *color, void *object_out)
{
    try {
        SurfaceCoefs c;
        std::memcpy(&c, coefs, 160);
        Object o(c, reflection_ratio, glm::vec3(color[0], color[1], color[2]));
        std::memset(object_out, 0, 176);
        std::memcpy(object_out, &o.surface, 160);
        std::memcpy((char *) object_out + 160, &o.reflection_ratio, 4);
        std::memcpy((char *) object_out + 164, &o.color, 12);
    }
    catch (std::exception &e) {
        g_error = e.what();
        return 1;
    }
    return 0;
}

// spherical != 0: LightSource::spherical(intensity, v = position, color); else directional(intensity, v = direction, color).
int ref_make_light(int spherical, float intensity, const double *v, const float *color, void *light_out)
{
    try {
        glm::dvec3 dv(v[0], v[1], v[2]);
        glm::vec3 col(color[0], color[1], color[2]);
        LightSource l = spherical ? LightSource::spherical(intensity, dv, col)
                                  : LightSource::directional(intensity, dv, col);
        std::memset(light_out, 0, 48);
        unsigned char flag = l.is_spherical ? 1 : 0;
        std::memcpy(light_out, &flag, 1);
        std::memcpy((char *) light_out + 8, &l.p, 24);
        std::memcpy((char *) light_out + 32, &l.light_color, 12);
    }
    catch (std::exception &e) {
        g_error = e.what();
        return 1;
    }
    return 0;
}

} // extern "C"
