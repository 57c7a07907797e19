// update_b200.cpp — third back-end of the reference's update.h (next to update-cpu.cpp and update-cuda.cu).
//
// One GPU: a single context, the frame streamed to a page-locked host buffer (rt_b200_update_to_host).
// Several GPUs (update_set_option("gpus", N) before init_update): one context per device in THIS process, rank r on
// device r, each driven by its own host thread — rank 0 by the caller's thread, so that update() stays the synchronous
// call of update.h:7.  The frame is assembled on GPU 0 by the render kernels themselves (fused gather over peer access,
// rt_b200_frame_attach_peer); rank 0 then copies it to the host frame.
#include "update_b200.h"

#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <utility>
#include <vector>

#include "../../include/rt_b200.h"

namespace {

std::vector<rt_b200_ctx *> g_ctx;           // file-scope state, like update-cuda.cu:15-27; [rank]
unsigned char *g_frame = nullptr;           // page-locked, so that the streamed device->host copy overlaps rendering
size_t g_frame_bytes = 0;
std::vector<std::pair<std::string, long long>> g_options;
int g_gpus = 1;
bool g_same_device = false;                 // test rig: all ranks on device 0 (a box with one GPU)

// worker threads of ranks 1..N-1: wait for a frame number, render it, report
std::vector<std::thread> g_workers;
std::mutex g_mu;
std::condition_variable g_cv;
unsigned long g_frame_no = 0;               // frame the workers should render (monotonic)
unsigned g_done = 0;                        // workers that finished frame g_frame_no
bool g_quit = false;
double g_matrix[16];
std::string g_worker_error;

void check(int rc)   // same policy as helper_cuda_opengl.h:13-44: print and exit
{
    if (rc != RT_B200_OK) {
        std::fprintf(stderr, "rt_b200: %s\n", rt_b200_last_error());
        std::exit(EXIT_FAILURE);
    }
}

void worker(int rank)
{
    unsigned long seen = 0;
    for (;;) {
        double m[16];
        {
            std::unique_lock<std::mutex> lock(g_mu);
            g_cv.wait(lock, [&] { return g_quit || g_frame_no != seen; });
            if (g_quit) return;
            seen = g_frame_no;
            std::memcpy(m, g_matrix, sizeof(m));
        }
        const int rc = rt_b200_update(g_ctx[(size_t) rank], m, nullptr);
        {
            std::lock_guard<std::mutex> lock(g_mu);
            if (rc != RT_B200_OK && g_worker_error.empty()) g_worker_error = rt_b200_last_error();
            g_done++;
        }
        g_cv.notify_all();
    }
}

} // namespace

void update_set_option(const char *name, long long value)
{
    if (!std::strcmp(name, "gpus")) g_gpus = value < 1 ? 1 : (int) value;
    else if (!std::strcmp(name, "same_device")) g_same_device = value != 0;
    else g_options.emplace_back(name, value);
}

void init_update(unsigned int /*texture*/, const rt_host::Scene &scene)
{
    cleanup_update();
    g_ctx.assign((size_t) g_gpus, nullptr);
    for (int r = 0; r < g_gpus; r++) {
        check(rt_b200_upload_scene(&g_ctx[(size_t) r], scene.px_width, scene.px_height, scene.vertical_fov, scene.bg_color,
                                   scene.max_reflections, scene.objects.data(), scene.objects.size(), scene.lights.data(),
                                   scene.lights.size(), /*device*/ g_same_device ? 0 : r, /*rank*/ r, /*world_size*/ g_gpus));
        for (const auto &o : g_options) check(rt_b200_set_option(g_ctx[(size_t) r], o.first.c_str(), o.second));
    }
    if (g_gpus > 1) {
        unsigned char handle[RT_B200_IPC_HANDLE_BYTES];   // unused here: the ranks live in one process
        check(rt_b200_frame_export(g_ctx[0], handle));
        for (int r = 0; r < g_gpus; r++) check(rt_b200_frame_attach_peer(g_ctx[(size_t) r], g_ctx[0]));
        g_quit = false;
        g_frame_no = 0;
        for (int r = 1; r < g_gpus; r++) g_workers.emplace_back(worker, r);
    }
    g_frame_bytes = 4ull * scene.px_width * scene.px_height;
    void *p = nullptr;
    check(rt_b200_host_alloc(&p, g_frame_bytes));
    g_frame = static_cast<unsigned char *>(p);
    std::memset(g_frame, 0, g_frame_bytes);
}

float update(const double camera_matrix[16])
{
    if (g_ctx.empty()) {
        std::fprintf(stderr, "rt_b200: update() before init_update()\n");
        std::exit(EXIT_FAILURE);
    }
    float ms = 0.f;
    if (g_gpus == 1) {
        check(rt_b200_update_to_host(g_ctx[0], camera_matrix, g_frame, g_frame_bytes, &ms));
        return ms;   // what ray-tracer.cpp:239 prints as "last render time"
    }
    {
        std::lock_guard<std::mutex> lock(g_mu);
        std::memcpy(g_matrix, camera_matrix, sizeof(g_matrix));
        g_done = 0;
        g_frame_no++;
    }
    g_cv.notify_all();
    check(rt_b200_update(g_ctx[0], camera_matrix, &ms));   // returns when every rank's pixels are in GPU 0's frame
    check(rt_b200_gather(g_ctx[0], g_frame, g_frame_bytes, nullptr));
    {
        std::unique_lock<std::mutex> lock(g_mu);
        g_cv.wait(lock, [&] { return g_done == (unsigned) (g_gpus - 1); });
        if (!g_worker_error.empty()) {
            std::fprintf(stderr, "rt_b200: %s\n", g_worker_error.c_str());
            std::exit(EXIT_FAILURE);
        }
    }
    return ms;
}

void cleanup_update()
{
    if (!g_workers.empty()) {
        {
            std::lock_guard<std::mutex> lock(g_mu);
            g_quit = true;
        }
        g_cv.notify_all();
        for (auto &t : g_workers) t.join();
        g_workers.clear();
    }
    for (size_t r = g_ctx.size(); r-- > 0;)      // rank 0 owns the shared frame: destroy it last
        if (g_ctx[r]) rt_b200_destroy(g_ctx[r]);
    g_ctx.clear();
    if (g_frame) rt_b200_host_free(g_frame);
    g_frame = nullptr;
    g_frame_bytes = 0;
}

const unsigned char *update_frame(size_t *bytes)
{
    if (bytes) *bytes = g_frame_bytes;
    return g_frame;
}
