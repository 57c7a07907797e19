// rt_headless.cpp — headless driver: the reference's main() (src/ray-tracer.cpp:136-248) without the window.
//
//   rt_headless <scene.yml> [--width W] [--height H] [--max-reflections N]      (override the YAML values)
//               [--frames K] [--position x,y,z] [--yaw deg] [--pitch deg]       (camera: ray-tracer.cpp:25-32)
//               [--orbit]                                                       (scripted fly-through: yaw sweeps 60..120)
//               [--out frame.ppm] [--raw frame.rgba] [--kernel 0|1]
//               [--gpus N]                                                      (one context + host thread per GPU, one process)
//               [--dump-camera]                                                 (print the 16 doubles handed to update())
//               [--option name=value ...]                                       (any rt_b200_set_option, e.g. accel=1)
//               [--dump-scene]                                                  (print the parsed scene; needs no GPU)
//
// What stays of the reference's main: argv[1] is the scene path and a SceneException ends the program with its
// message (ray-tracer.cpp:145-158); init_update once, update(camera_matrix()) per frame, cleanup_update at exit
// (ray-tracer.cpp:215,226,245); once per second — and after the last frame — the same line the reference prints,
// "FPS: %.4lf, last render time: %.4f ms" (ray-tracer.cpp:239).  What is gone: GLFW/GLEW/OpenGL (no graphics
// stack on this box); the frame is written to a file instead (PPM, top row first, i.e. rows flipped, because
// row 0 of the frame is the bottom of the picture — ray-tracer.cpp:12-17).
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <utility>
#include <vector>

#include "rt_camera.h"
#include "rt_host_scene.h"
#include "update_b200.h"

namespace {

void usage(const char *argv0)
{
    std::fprintf(stderr,
                 "Usage: %s <scene.yml> [--width W] [--height H] [--max-reflections N] [--frames K]\n"
                 "          [--position x,y,z] [--yaw deg] [--pitch deg] [--orbit] [--out f.ppm] [--raw f.rgba]\n"
                 "          [--kernel 0|1] [--gpus N] [--option name=value ...] [--dump-scene] [--dump-camera]\n", argv0);
}

void dump_scene(const rt_host::Scene &s)
{
    std::printf("scene %u %u %.17g %.9g %.9g %.9g %u %zu %zu\n", s.px_width, s.px_height, s.vertical_fov, s.bg_color[0],
                s.bg_color[1], s.bg_color[2], s.max_reflections, s.objects.size(), s.lights.size());
    for (const auto &o : s.objects) {
        const unsigned char *b = reinterpret_cast<const unsigned char *>(&o);
        std::printf("object ");
        for (size_t i = 0; i < sizeof(o); i++) std::printf("%02x", b[i]);
        std::printf("\n");
    }
    for (const auto &l : s.lights) {
        // padding bytes are not part of the contract: print the fields
        unsigned char b[48];
        std::memset(b, 0, sizeof(b));
        b[0] = l.is_spherical ? 1 : 0;
        std::memcpy(b + 8, l.p, 24);
        std::memcpy(b + 32, l.light_color, 12);
        std::printf("light ");
        for (size_t i = 0; i < sizeof(b); i++) std::printf("%02x", b[i]);
        std::printf("\n");
    }
}

bool write_ppm(const char *path, const unsigned char *rgba, unsigned w, unsigned h)
{
    FILE *f = std::fopen(path, "wb");
    if (!f) return false;
    std::fprintf(f, "P6\n%u %u\n255\n", w, h);
    std::string row(3 * (size_t) w, '\0');
    for (unsigned y = 0; y < h; y++) {
        const unsigned char *src = rgba + 4ull * w * (h - 1 - y);   // flip: frame row 0 is the bottom of the picture
        for (unsigned x = 0; x < w; x++) {
            row[3 * x] = (char) src[4 * x];
            row[3 * x + 1] = (char) src[4 * x + 1];
            row[3 * x + 2] = (char) src[4 * x + 2];
        }
        std::fwrite(row.data(), 1, row.size(), f);
    }
    return std::fclose(f) == 0;
}

} // namespace

int main(int argc, char **argv)
{
    if (argc < 2) {   // ray-tracer.cpp:145-148
        usage(argv[0]);
        return EXIT_FAILURE;
    }
    const char *scene_path = argv[1];
    long width = 0, height = 0, max_refl = -1, frames = 1, kernel = -1;
    bool orbit = false, dump = false, dump_camera = false;
    long gpus = 1;
    const char *out_ppm = nullptr, *out_raw = nullptr;
    std::vector<std::pair<std::string, long long>> options;
    rt_host::Camera cam;
    for (int i = 2; i < argc; i++) {
        const std::string a = argv[i];
        auto next = [&]() -> const char * {
            if (i + 1 >= argc) {
                usage(argv[0]);
                std::exit(EXIT_FAILURE);
            }
            return argv[++i];
        };
        if (a == "--width") width = std::atol(next());
        else if (a == "--height") height = std::atol(next());
        else if (a == "--max-reflections") max_refl = std::atol(next());
        else if (a == "--frames") frames = std::atol(next());
        else if (a == "--kernel") kernel = std::atol(next());
        else if (a == "--yaw") cam.yaw = std::atof(next());
        else if (a == "--pitch") cam.pitch = std::atof(next());
        else if (a == "--position") {
            if (std::sscanf(next(), "%lf,%lf,%lf", &cam.position[0], &cam.position[1], &cam.position[2]) != 3) {
                usage(argv[0]);
                return EXIT_FAILURE;
            }
        }
        else if (a == "--orbit") orbit = true;
        else if (a == "--out") out_ppm = next();
        else if (a == "--raw") out_raw = next();
        else if (a == "--dump-scene") dump = true;
        else if (a == "--dump-camera") dump_camera = true;
        else if (a == "--gpus") gpus = std::atol(next());
        else if (a == "--option") {
            const std::string kv = next();
            const size_t eq = kv.find('=');
            if (eq == std::string::npos || eq == 0) {
                usage(argv[0]);
                return EXIT_FAILURE;
            }
            options.emplace_back(kv.substr(0, eq), std::atoll(kv.c_str() + eq + 1));
        }
        else {
            usage(argv[0]);
            return EXIT_FAILURE;
        }
    }

    try {   // ray-tracer.cpp:150-158
        rt_host::Scene scene = rt_host::Scene::load_from_file(scene_path);
        if (width > 0) scene.px_width = (unsigned) width;
        if (height > 0) scene.px_height = (unsigned) height;
        if (max_refl >= 0) scene.max_reflections = (unsigned) max_refl;
        if (dump) {
            dump_scene(scene);
            return EXIT_SUCCESS;
        }
        if (scene.px_width < 1 || scene.px_height < 1 || frames < 1) {
            std::fprintf(stderr, "Empty frame or no frames requested\n");
            return EXIT_FAILURE;
        }
        if (gpus < 1 || gpus > 64) {
            std::fprintf(stderr, "--gpus must be 1..64\n");
            return EXIT_FAILURE;
        }
        update_set_option("gpus", gpus);
        if (kernel >= 0) update_set_option("kernel", kernel);
        for (const auto &o : options) update_set_option(o.first.c_str(), o.second);
        init_update(0, scene);   // ray-tracer.cpp:215

        double m[16];
        float ms = 0.f;
        auto t_last = std::chrono::steady_clock::now();
        long n_since = 0;
        for (long k = 0; k < frames; k++) {   // ray-tracer.cpp:222-243
            if (orbit && frames > 1) cam.yaw = 60.0 + 60.0 * (double) k / (double) (frames - 1);
            cam.matrix(m);
            if (dump_camera) {   // what update() receives: glm::dmat4 in column-major order (ray-tracer.cpp:54-58,226)
                std::printf("camera");
                for (int q = 0; q < 16; q++) std::printf(" %.17g", m[q]);
                std::printf("\n");
            }
            ms = update(m);
            n_since++;
            const auto now = std::chrono::steady_clock::now();
            const double dt = std::chrono::duration<double>(now - t_last).count();
            if (dt >= 1.0 || k + 1 == frames) {
                std::printf("FPS: %.4lf, last render time: %.4f ms\n", n_since / dt, ms);   // ray-tracer.cpp:239
                t_last = now;
                n_since = 0;
            }
        }
        size_t bytes = 0;
        const unsigned char *frame = update_frame(&bytes);
        int rc = EXIT_SUCCESS;
        if (out_ppm && !write_ppm(out_ppm, frame, scene.px_width, scene.px_height)) {
            std::fprintf(stderr, "Cannot write %s\n", out_ppm);
            rc = EXIT_FAILURE;
        }
        if (out_raw) {
            FILE *f = std::fopen(out_raw, "wb");
            if (!f || std::fwrite(frame, 1, bytes, f) != bytes || std::fclose(f) != 0) {
                std::fprintf(stderr, "Cannot write %s\n", out_raw);
                rc = EXIT_FAILURE;
            }
        }
        cleanup_update();   // ray-tracer.cpp:245
        return rc;
    }
    catch (rt_host::SceneException &e) {
        std::fprintf(stderr, "%s\n", e.what());
        return EXIT_FAILURE;
    }
}
