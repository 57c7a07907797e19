// rt_host_scene.cpp — scene factories, validation and the YAML loader of the C++ host.
// Follows the reference's src/surface.cpp:4-60, src/light.cpp:4-26, src/scene-exception.cpp:3-11 and
// src/scene.cpp:6-22,24-203 (defaults, required keys, validation order, error messages incl. line/column).
#include "rt_host_scene.h"

#include <cmath>
#include <cstring>
#include <sstream>

#include "rt_yaml.h"

namespace rt_host {

namespace {

// scene.cpp:6-7
const unsigned int MAX_REFLECTIONS = 5;
const float BG_COLOR[3] = {1.0f, 1.0f, 1.0f};

double dot3(const double a[3], const double b[3]) { return (a[0] * b[0] + a[1] * b[1]) + a[2] * b[2]; }   // glm::dot

template <typename T>
void validate_positive(const char *what, const T &value)   // scene-exception.h:27-35: rejects only value < 0
{
    if (value < 0) {
        std::stringstream error;
        error << "Negative value for " << what << ": " << value;
        throw SceneException(error.str());
    }
}

std::string mark_to_string(rt_yaml::Mark mark)   // scene.cpp:24-29
{
    std::stringstream ss;
    ss << "line: " << mark.line + 1 << " column: " << mark.column + 1;
    return ss.str();
}

SceneException undefined_value(rt_yaml::Mark parent, const char *key)
{
    return SceneException(std::string("Value '") + key + "' undefined, " + mark_to_string(parent));
}

SceneException invalid_type(rt_yaml::Mark mark, const char *key)
{
    return SceneException(std::string("Value '") + key + "' is invalid, " + mark_to_string(mark));
}

using rt_yaml::Node;

// convert<glm::vec<3, T>>::decode, scene.cpp:78-94
template <typename T>
bool decode_vec3(const Node &node, T out[3])
{
    if (!node.IsSequence() || node.size() != 3) return false;
    for (int i = 0; i < 3; i++) out[i] = (T) node[(size_t) i].as_double();
    return true;
}

// get_value<T>(node, key), scene.cpp:41-54: required key, conversion errors carry the value's mark
const Node &required(const Node &node, const char *key)
{
    const Node &v = node[key];
    if (!v.IsDefined()) throw undefined_value(node.mark, key);
    return v;
}

template <typename T>
void get_vec3(const Node &node, const char *key, T out[3])
{
    const Node &v = required(node, key);
    try {
        if (!decode_vec3(v, out)) throw rt_yaml::BadConversion(v.mark);
    }
    catch (rt_yaml::BadConversion &e) {
        throw invalid_type(e.mark, key);
    }
}

// node[key].as<vec3>(fallback): the fallback when the key is absent or is not a 3-element sequence
template <typename T>
void vec3_or(const Node &node, const char *key, const T fallback[3], T out[3])
{
    const Node &v = node[key];
    try {
        if (v.IsDefined() && decode_vec3(v, out)) return;
    }
    catch (rt_yaml::BadConversion &e) {
        throw invalid_type(e.mark, key);   // a non-numeric component (yaml-cpp lets this escape as BadConversion)
    }
    for (int i = 0; i < 3; i++) out[i] = fallback[i];
}

double double_or(const Node &node, const char *key, double fallback)
{
    const Node &v = node[key];
    if (!v.IsDefined()) return fallback;
    try {
        return v.as_double();
    }
    catch (rt_yaml::BadConversion &) {
        return fallback;   // as<T>(fallback) swallows a failed conversion
    }
}

unsigned unsigned_or(const Node &node, const char *key, unsigned fallback)
{
    const Node &v = node[key];
    if (!v.IsDefined()) return fallback;
    try {
        return v.as_unsigned();
    }
    catch (rt_yaml::BadConversion &) {
        return fallback;
    }
}

void check_sequence(const Node &node, const char *key)   // scene.cpp:56-65
{
    const Node &v = node[key];
    if (!v.IsDefined()) throw undefined_value(node.mark, key);
    if (!v.IsSequence())
        throw SceneException(std::string("Value '") + key + "' must be a sequence, " + mark_to_string(v.mark));
}

void check_map(const Node &node, const char *key)   // scene.cpp:67-76
{
    const Node &v = node[key];
    if (!v.IsDefined()) throw undefined_value(node.mark, key);
    if (!v.IsMap()) throw SceneException(std::string("Value '") + key + "' must be a mapping, " + mark_to_string(v.mark));
}

std::string get_string(const Node &node, const char *key)
{
    const Node &v = required(node, key);
    try {
        return v.as_string();
    }
    catch (rt_yaml::BadConversion &e) {
        throw invalid_type(e.mark, key);
    }
}

SurfaceCoefs parse_surface(const Node &node)   // scene.cpp:97-151
{
    const std::string type = get_string(node, "type");
    const double zero[3] = {0.0, 0.0, 0.0}, up[3] = {0.0, 1.0, 0.0};
    double a[3], b[3];
    if (type == "sphere") {
        vec3_or(node, "center", zero, a);
        return SurfaceCoefs::sphere(a, double_or(node, "radius", 1.0));
    }
    if (type == "plane") {
        vec3_or(node, "origin", zero, a);
        vec3_or(node, "normal", up, b);
        return SurfaceCoefs::plane(a, b);
    }
    if (type == "dingDong") {
        vec3_or(node, "origin", zero, a);
        return SurfaceCoefs::dingDong(a);
    }
    if (type == "clebsch") return SurfaceCoefs::clebsch();
    if (type == "cayley") return SurfaceCoefs::cayley();
    if (type == "polynomial") {
        check_map(node, "coefficients");
        const Node &k = node["coefficients"];
        SurfaceCoefs s{};
#define RT_LOAD_COEF(name) s.name = double_or(k, #name, 0.0)
        RT_LOAD_COEF(x3); RT_LOAD_COEF(y3); RT_LOAD_COEF(z3); RT_LOAD_COEF(x2y); RT_LOAD_COEF(xy2);
        RT_LOAD_COEF(x2z); RT_LOAD_COEF(xz2); RT_LOAD_COEF(y2z); RT_LOAD_COEF(yz2); RT_LOAD_COEF(xyz);
        RT_LOAD_COEF(x2); RT_LOAD_COEF(y2); RT_LOAD_COEF(z2); RT_LOAD_COEF(xy); RT_LOAD_COEF(xz);
        RT_LOAD_COEF(yz); RT_LOAD_COEF(x); RT_LOAD_COEF(y); RT_LOAD_COEF(z); RT_LOAD_COEF(c);
#undef RT_LOAD_COEF
        return s;
    }
    throw SceneException(std::string("Unknown surface type: '") + type + "', " + mark_to_string(node["type"].mark));
}

Scene scene_from_document(const Node &doc)   // scene.cpp:164-202
{
    unsigned width, height;
    double fov;
    try {
        width = required(doc, "width").as_unsigned();
    }
    catch (rt_yaml::BadConversion &e) { throw invalid_type(e.mark, "width"); }
    try {
        height = required(doc, "height").as_unsigned();
    }
    catch (rt_yaml::BadConversion &e) { throw invalid_type(e.mark, "height"); }
    try {
        fov = required(doc, "fov").as_double();
    }
    catch (rt_yaml::BadConversion &e) { throw invalid_type(e.mark, "fov"); }
    float bg[3];
    vec3_or(doc, "bg_color", BG_COLOR, bg);
    Scene scene(width, height, fov, unsigned_or(doc, "max_reflections", MAX_REFLECTIONS), bg);
    check_sequence(doc, "objects");
    check_sequence(doc, "light_sources");
    for (const Node &node : doc["objects"].items) {
        const SurfaceCoefs surface = parse_surface(node);
        const float refl = (float) double_or(node, "reflection_ratio", 0.0);
        float color[3];
        get_vec3(node, "color", color);
        scene.objects.push_back(Object(surface, refl, color));
    }
    const float white[3] = {1.0f, 1.0f, 1.0f};
    for (const Node &node : doc["light_sources"].items) {
        const std::string type = get_string(node, "type");
        const float intensity = (float) double_or(node, "intensity", 1.0);
        double v[3];
        float color[3];
        if (type == "directional") {
            get_vec3(node, "direction", v);
            vec3_or(node, "color", white, color);
            scene.lights.push_back(LightSource::directional(intensity, v, color));
        }
        else if (type == "spherical") {
            get_vec3(node, "position", v);
            vec3_or(node, "color", white, color);
            scene.lights.push_back(LightSource::spherical(intensity, v, color));
        }
        else {
            throw SceneException(std::string("Light source type must be 'spherical' or 'directional', ") +
                                 mark_to_string(node["type"].mark));
        }
    }
    return scene;
}

} // namespace

void validate_color(const float color[3])   // scene-exception.cpp:3-11
{
    for (int i = 0; i < 3; i++) {
        if (color[i] < 0.0f || color[i] > 1.0f) {
            std::stringstream error;
            error << "Invalid color: (" << color[0] << ", " << color[1] << ", " << color[2] << ")";
            throw SceneException(error.str());
        }
    }
}

// ---- surface factories, src/surface.cpp ---------------------------------------------------------------------------
SurfaceCoefs SurfaceCoefs::sphere(const double center[3], double radius)
{
    validate_positive("sphere radius", radius);
    SurfaceCoefs s{};
    s.x2 = s.y2 = s.z2 = 1.0;
    s.x = -2.0 * center[0];
    s.y = -2.0 * center[1];
    s.z = -2.0 * center[2];
    s.c = dot3(center, center) - radius * radius;
    return s;
}

SurfaceCoefs SurfaceCoefs::plane(const double origin[3], const double normal[3])
{
    SurfaceCoefs s{};
    s.x = normal[0];
    s.y = normal[1];
    s.z = normal[2];
    s.c = -dot3(origin, normal);
    return s;
}

SurfaceCoefs SurfaceCoefs::dingDong(const double o[3])
{
    SurfaceCoefs s{};
    s.x2 = s.y3 = s.z2 = 1.0;
    s.y2 = -1.0 - 3.0 * o[1];
    s.x = -2.0 * o[0];
    s.z = -2.0 * o[2];
    s.y = (2.0 + 3.0 * o[1]) * o[1];
    s.c = std::pow(o[0], 2) + std::pow(o[2], 2) - std::pow(o[1], 2) * (1.0 + o[1]);
    return s;
}

SurfaceCoefs SurfaceCoefs::clebsch()
{
    SurfaceCoefs s{};
    // surface.cpp:44 assigns x3 twice and never z3: z3 stays 0.  Part of the oracle; not "fixed" here.
    s.x3 = s.y3 = 81.0;
    s.x2y = s.x2z = s.xy2 = s.y2z = s.xz2 = s.yz2 = -189.0;
    s.xyz = 54.0;
    s.xy = s.yz = s.xz = 126.0;
    s.x2 = s.y2 = s.z2 = -9.0;
    s.x = s.y = s.z = 9.0;
    s.c = 1.0;
    return s;
}

SurfaceCoefs SurfaceCoefs::cayley()
{
    SurfaceCoefs s{};
    s.x2y = s.x2z = s.xy2 = s.y2z = s.xz2 = s.yz2 = -5.0;
    s.xy = s.yz = s.xz = 2.0;
    return s;
}

// ---- lights, src/light.cpp ------------------------------------------------------------------------------------------
static LightSource make_light(bool spherical, float intensity, const double v[3], const float color[3])
{
    validate_positive("light intensity", intensity);
    validate_color(color);
    LightSource l;
    std::memset(&l, 0, sizeof(l));
    l.is_spherical = spherical;
    for (int i = 0; i < 3; i++) {
        l.p[i] = v[i];
        l.light_color[i] = intensity * color[i];
    }
    return l;
}

LightSource LightSource::directional(float intensity, const double direction[3], const float color[3])
{
    const double inv = 1.0 / std::sqrt(dot3(direction, direction));   // glm::normalize: v * (1 / sqrt(dot(v, v)))
    const double p[3] = {-(direction[0] * inv), -(direction[1] * inv), -(direction[2] * inv)};
    return make_light(false, intensity, p, color);
}

LightSource LightSource::spherical(float intensity, const double position[3], const float color[3])
{
    return make_light(true, intensity, position, color);
}

Object::Object(const SurfaceCoefs &surface_, float reflection_ratio_, const float color_[3])
    : surface(surface_), reflection_ratio(reflection_ratio_)
{
    for (int i = 0; i < 3; i++) color[i] = color_[i];
    validate_positive("object reflection ratio", reflection_ratio);
    validate_color(color);
}

Scene::Scene(unsigned int w, unsigned int h, double vertical_fov_deg, unsigned int max_refl, const float bg[3])
    : px_width(w), px_height(h), max_reflections(max_refl)
{
    vertical_fov = vertical_fov_deg * 0.01745329251994329576923690768489;   // glm::radians
    for (int i = 0; i < 3; i++) bg_color[i] = bg[i];
    validate_color(bg_color);
}

Scene Scene::load_from_file(const char *path)
{
    Node doc;
    try {
        doc = rt_yaml::LoadFile(path);
    }
    catch (rt_yaml::BadFile &) {
        throw SceneException(std::string("Cannot read the file ") + path);
    }
    catch (rt_yaml::ParseError &e) {
        throw SceneException(std::string("YAML parser error: ") + e.what());
    }
    return scene_from_document(doc);
}

Scene Scene::load_from_string(const std::string &yaml)
{
    Node doc;
    try {
        std::istringstream in(yaml);
        doc = rt_yaml::Load(in);
    }
    catch (rt_yaml::ParseError &e) {
        throw SceneException(std::string("YAML parser error: ") + e.what());
    }
    return scene_from_document(doc);
}

} // namespace rt_host
