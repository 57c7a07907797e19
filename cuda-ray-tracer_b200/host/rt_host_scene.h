// rt_host_scene.h — host-side scene model of the B200 back-end's C++ host (loader + headless driver).
//
// Mirrors the reference's host types with the SAME in-memory layouts, so that `scene.objects.data()` and
// `scene.lights.data()` can be handed to rt_b200_upload_scene exactly as the reference would hand
// `&scene.objects[0]`:
//   SurfaceCoefs  include/surface.h:10-14   20 doubles, order x3,y3,z3,x2y,xy2,x2z,xz2,y2z,yz2,xyz,x2,y2,z2,xy,xz,yz,x,y,z,c
//   Object        include/scene.h:8-15      176 bytes: SurfaceCoefs, float reflection_ratio @160, float color[3] @164
//   LightSource   include/light.h:6-13      48 bytes: bool is_spherical @0, double p[3] @8, float light_color[3] @32
//   Scene         include/scene.h:17-36     px_width, px_height, vertical_fov (radians), bg_color, max_reflections
// glm is not on this image; the three-component vectors are plain arrays.
#ifndef RT_HOST_SCENE_H
#define RT_HOST_SCENE_H

#include <cstddef>
#include <exception>
#include <string>
#include <vector>

namespace rt_host {

// include/scene-exception.h:10-25
class SceneException : public std::exception {
public:
    explicit SceneException(std::string error) : error_(std::move(error)) {}
    const char *what() const noexcept override { return error_.c_str(); }

private:
    std::string error_;
};

struct SurfaceCoefs {
    double x3, y3, z3, x2y, xy2, x2z, xz2, y2z, yz2, xyz, x2, y2, z2, xy, xz, yz, x, y, z, c;

    // src/surface.cpp:4-60
    static SurfaceCoefs sphere(const double center[3], double radius);
    static SurfaceCoefs plane(const double origin[3], const double normal[3]);
    static SurfaceCoefs dingDong(const double origin[3]);
    static SurfaceCoefs clebsch();
    static SurfaceCoefs cayley();
};

struct Object {
    SurfaceCoefs surface;
    float reflection_ratio;
    float color[3];
    Object(const SurfaceCoefs &surface, float reflection_ratio, const float color[3]);   // validates, scene.cpp:9-14
};

struct LightSource {
    bool is_spherical;
    double p[3];            // -normalised direction (directional) or position (spherical)
    float light_color[3];   // intensity * colour
    static LightSource directional(float intensity, const double direction[3], const float color[3]);   // light.cpp:4-15
    static LightSource spherical(float intensity, const double position[3], const float color[3]);      // light.cpp:17-26
};

static_assert(sizeof(SurfaceCoefs) == 160, "SurfaceCoefs layout");
static_assert(sizeof(Object) == 176 && offsetof(Object, reflection_ratio) == 160 && offsetof(Object, color) == 164, "Object layout");
static_assert(sizeof(LightSource) == 48 && offsetof(LightSource, p) == 8 && offsetof(LightSource, light_color) == 32,
              "LightSource layout");

struct Scene {
    unsigned int px_width, px_height;
    double vertical_fov;   // radians (scene.cpp:20)
    float bg_color[3];
    unsigned int max_reflections;
    std::vector<Object> objects;
    std::vector<LightSource> lights;

    Scene(unsigned int px_width, unsigned int px_height, double vertical_fov_deg, unsigned int max_reflections,
          const float bg_color[3]);                       // scene.cpp:16-22
    double aspect_ratio() const { return (double) px_width / px_height; }   // scene.h:32-33

    // Scene::load_from_file, src/scene.cpp:154-203: same keys, defaults, validation and error messages
    static Scene load_from_file(const char *path);
    static Scene load_from_string(const std::string &yaml);
};

void validate_color(const float color[3]);   // src/scene-exception.cpp:3-11

} // namespace rt_host
#endif
