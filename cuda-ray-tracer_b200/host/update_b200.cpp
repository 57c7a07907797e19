// update_b200.cpp — third back-end of the reference's update.h (next to update-cpu.cpp and update-cuda.cu).
#include "update_b200.h"

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <utility>
#include <vector>

#include "../../include/rt_b200.h"

namespace {

rt_b200_ctx *g_ctx = nullptr;               // file-scope state, like update-cuda.cu:15-27
unsigned char *g_frame = nullptr;           // page-locked, so that the streamed device->host copy overlaps rendering
size_t g_frame_bytes = 0;
std::vector<std::pair<std::string, long long>> g_options;

void check(int rc)   // same policy as helper_cuda_opengl.h:13-44: print and exit
{
    if (rc != RT_B200_OK) {
        std::fprintf(stderr, "rt_b200: %s\n", rt_b200_last_error());
        std::exit(EXIT_FAILURE);
    }
}

} // namespace

void update_set_option(const char *name, long long value) { g_options.emplace_back(name, value); }

void init_update(unsigned int /*texture*/, const rt_host::Scene &scene)
{
    cleanup_update();
    check(rt_b200_upload_scene(&g_ctx, scene.px_width, scene.px_height, scene.vertical_fov, scene.bg_color,
                               scene.max_reflections, scene.objects.data(), scene.objects.size(), scene.lights.data(),
                               scene.lights.size(), /*device*/ 0, /*rank*/ 0, /*world_size*/ 1));
    for (const auto &o : g_options) check(rt_b200_set_option(g_ctx, o.first.c_str(), o.second));
    g_frame_bytes = 4ull * scene.px_width * scene.px_height;
    void *p = nullptr;
    check(rt_b200_host_alloc(&p, g_frame_bytes));
    g_frame = static_cast<unsigned char *>(p);
    std::memset(g_frame, 0, g_frame_bytes);
}

float update(const double camera_matrix[16])
{
    if (!g_ctx) {
        std::fprintf(stderr, "rt_b200: update() before init_update()\n");
        std::exit(EXIT_FAILURE);
    }
    float ms = 0.f;
    check(rt_b200_update_to_host(g_ctx, camera_matrix, g_frame, g_frame_bytes, &ms));
    return ms;   // what ray-tracer.cpp:239 prints as "last render time"
}

void cleanup_update()
{
    if (g_ctx) rt_b200_destroy(g_ctx);
    g_ctx = nullptr;
    if (g_frame) rt_b200_host_free(g_frame);
    g_frame = nullptr;
    g_frame_bytes = 0;
}

const unsigned char *update_frame(size_t *bytes)
{
    if (bytes) *bytes = g_frame_bytes;
    return g_frame;
}
