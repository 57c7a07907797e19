// rt_api.cu — implementation of the C ABI declared in include/rt_b200.h.
//
// Host-side work per call:
//   rt_b200_upload_scene  classify + repack the reference's AoS Object[] / LightSource[] into the class-sorted
//                         SoA of rt_scene.h, upload once, allocate the frame buffers (≙ update-cuda.cu:34-63)
//   rt_b200_update        fill FrameParams from the camera matrix, launch, optionally all-gather the strips,
//                         event-time it (≙ update-cuda.cu:160-190)
//   rt_b200_gather        hand out the assembled RGBA8 frame (replaces the GL texture)
// No CPU rendering path exists in this library: if CUDA is unavailable the calls fail with RT_B200_ERR_CUDA.
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <fcntl.h>
#include <nccl.h>
#include <sched.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <time.h>
#include <unistd.h>

#include <cerrno>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <algorithm>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/rt_b200.h"
#include "rt_grid_build.h"
#include "rt_kernels.h"

namespace {

thread_local std::string g_last_error;

int fail(int code, const char *fmt, ...)
{
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    g_last_error = buf;
    return code;
}

#define CUDA_TRY(expr)                                                                              \
    do {                                                                                            \
        cudaError_t e_ = (expr);                                                                    \
        if (e_ != cudaSuccess) return fail(RT_B200_ERR_CUDA, "%s: %s", #expr, cudaGetErrorString(e_)); \
    } while (0)

// ---- NCCL, loaded on first use (the Python host already has torch's libnccl.so.2 mapped) -----------
struct NcclApi {
    void *handle = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId *) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllReduce)(const void *, void *, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    const char *(*GetErrorString)(ncclResult_t) = nullptr;
};
NcclApi g_nccl;

int load_nccl()
{
    if (g_nccl.handle) return RT_B200_OK;
    void *h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) return fail(RT_B200_ERR_NCCL, "cannot load libnccl.so.2: %s", dlerror());
    g_nccl.GetUniqueId = (decltype(g_nccl.GetUniqueId)) dlsym(h, "ncclGetUniqueId");
    g_nccl.CommInitRank = (decltype(g_nccl.CommInitRank)) dlsym(h, "ncclCommInitRank");
    g_nccl.CommDestroy = (decltype(g_nccl.CommDestroy)) dlsym(h, "ncclCommDestroy");
    g_nccl.AllGather = (decltype(g_nccl.AllGather)) dlsym(h, "ncclAllGather");
    g_nccl.GetErrorString = (decltype(g_nccl.GetErrorString)) dlsym(h, "ncclGetErrorString");
    g_nccl.AllReduce = (decltype(g_nccl.AllReduce)) dlsym(h, "ncclAllReduce");
    if (!g_nccl.AllReduce || !g_nccl.GetUniqueId || !g_nccl.CommInitRank || !g_nccl.CommDestroy || !g_nccl.AllGather ||
        !g_nccl.GetErrorString) {
        dlclose(h);
        return fail(RT_B200_ERR_NCCL, "libnccl.so.2 lacks a required symbol");
    }
    g_nccl.handle = h;
    return RT_B200_OK;
}

#define NCCL_TRY(expr)                                                                                 \
    do {                                                                                               \
        ncclResult_t r_ = (expr);                                                                      \
        if (r_ != ncclSuccess) return fail(RT_B200_ERR_NCCL, "%s: %s", #expr, g_nccl.GetErrorString(r_)); \
    } while (0)

// ---- host frames shared between the processes of a multi-GPU job -------------------------------------
// A file-backed MAP_SHARED mapping (the caller picks the path, normally under /dev/shm), page-locked in every
// process that maps it, so each rank's copy engine writes its own strips of the frame straight into the one host
// buffer the consumer reads: G PCIe links fill one frame.  Behind the frame sits one page of sequence words:
//   go        last frame rank 0 (the consumer) has entered — rank r may overwrite the buffer for frame k once go >= k
//   done[r]   last frame whose strips rank r has finished copying
struct SharedTrailer {
    unsigned go;
    unsigned pad[15];
    unsigned done[1000];
};
static_assert(sizeof(SharedTrailer) <= 4096, "one page of sync words");

struct SharedBlock {
    void *base = nullptr;
    size_t bytes = 0, mapped = 0;
    std::string path;
    bool creator = false;
    unsigned seq = 0;   // frames this process has sent through the block
    SharedTrailer *trailer() const { return reinterpret_cast<SharedTrailer *>(static_cast<char *>(base) + mapped - 4096); }
};
std::mutex g_blocks_mutex;
std::vector<SharedBlock *> g_blocks;

SharedBlock *find_block(const void *p)
{
    std::lock_guard<std::mutex> lock(g_blocks_mutex);
    for (SharedBlock *b : g_blocks)
        if (b->base == p) return b;
    return nullptr;
}

double now_seconds()
{
    timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (double) ts.tv_sec + 1e-9 * (double) ts.tv_nsec;
}

// spin until *word >= need (acquire); false after `timeout_s`
bool host_wait_word(const unsigned *word, unsigned need, double timeout_s)
{
    const double t0 = now_seconds();
    for (unsigned spins = 0;; spins++) {
        if ((int) (__atomic_load_n(word, __ATOMIC_ACQUIRE) - need) >= 0) return true;
        if ((spins & 1023) == 1023) {
            if (now_seconds() - t0 > timeout_s) return false;
            sched_yield();
        }
    }
}

template <typename T>
cudaError_t upload(const std::vector<T> &host, T **dev)
{
    *dev = nullptr;
    size_t bytes = sizeof(T) * (host.empty() ? 1 : host.size());
    cudaError_t e = cudaMalloc((void **) dev, bytes);
    if (e != cudaSuccess) return e;
    if (!host.empty()) e = cudaMemcpy(*dev, host.data(), sizeof(T) * host.size(), cudaMemcpyHostToDevice);
    return e;
}

} // namespace

#define RT_MAX_BANDS 16
#define RT_DYN_AHEAD 2         /* slots whose group is fetched in advance (and assigned statically at the start): a group of
                                  8 x 3840 pixels is ~2 us of one GPU's work, the fetch ~3 us; what a GPU has reserved when
                                  the groups run out is the tail imbalance (ahead 4 x 16 rows cost 17 us per frame) */
#define RT_GRID_AUTO_SPHERES 64

struct rt_b200_ctx {
    int device = 0, rank = 0, world = 1;
    int width = 0, height = 0;
    double vertical_fov = 0, aspect = 0, tan_half_fov = 0;
    float bg[3] = {0, 0, 0};
    int max_reflections = 0;
    uint32_t class_counts[4] = {0, 0, 0, 0};

    // device scene
    rt::SceneView view{};
    double4 *d_plane = nullptr, *d_sphere = nullptr;
    float4 *d_sphere_f = nullptr, *d_material = nullptr, *d_material_pi = nullptr;
    double *d_quadric = nullptr, *d_cubic = nullptr, *d_coef_full = nullptr;
    int *d_orig = nullptr;
    rt::LightDev *d_lights = nullptr;

    // frame buffers
    int strip_rows = 8;
    int local_rows = 0;             // rows of the packed per-rank buffer (padded to whole strips per rank)
    uchar4 *d_local = nullptr;      // [local_rows][width]; for world == 1 this IS the frame
    uchar4 *d_gathered = nullptr;   // [world][local_rows][width] (world > 1)
    uchar4 *d_frame = nullptr;      // [height][width] (world > 1)
    unsigned long long *d_counters = nullptr;
    unsigned long long *d_tile_counter = nullptr;
    unsigned *d_cubic_mask = nullptr;
    float sphere_r2_max = 0.0f;
    int shadow_cache = -1;
    int wave_variant = 0;
    int cubic_filter = 1;
    int cull_backfacing = 1;
    int lone_kernel = 1;             // option: scenes of exactly one cubic object use the kernel specialised for them
    int spheres_smem = -1;           // option: -1 automatic, 0 = sphere filter records through L1 instead of shared memory
    int timeline = 0;                // option (diagnostic build only): per-CTA time stamps of the last launch, rt_b200_debug_timeline
    // sphere grid (rt_grid.cuh): built at upload; accel = -1 automatic (on from RT_GRID_AUTO_SPHERES
    // sphere-class objects), 0 brute force, 1 on.  Frames are identical either way.
    int accel = -1;
    int *d_grid_cells = nullptr, *d_grid_items = nullptr;
    std::vector<double> sphere_centres, sphere_r2;   // kept for rebuilding the grid at another density
    uint32_t grid_info[6] = {0, 0, 0, 0, 0, 0};      // nx, ny, nz, gridded spheres, loose spheres, registrations
    bool sync_in_kernel = false;       // this frame's go / flag words are handled inside the wavefront kernel
    bool tile_counter_dirty = false;   // another kernel used the tile counter: zero it before the wavefront kernel
    double trace[5] = {0, 0, 0, 0, 0};   // RT_B200_TRACE=1: seconds spent per stage of update_to_host (shared block)
    unsigned trace_frames = 0;
    bool cull_safe = true;           // every light / object colour is finite and moderate: 0 * colour is exactly 0
    bool buffers_dirty = true;

    int kernel_kind = 1;             // option "kernel"
    bool wave_fits = true;           // the wavefront kernel's per-CTA scene copy fits shared memory (else kernel 0 renders)
    int kernel_now() const { return (kernel_kind == 1 && !wave_fits) ? 0 : kernel_kind; }
    int launches = 0;                // render launches of the last update
    bool count = true;
    bool rendered = false;

    // fused gather (world > 1): the assembled frame lives on rank 0, double-buffered; the other ranks map it through
    // CUDA IPC and their render kernels store pixels straight into it over NVLink
    uchar4 *d_shared_base = nullptr;   // rank 0: cudaMalloc; other ranks: cudaIpcOpenMemHandle
    uchar4 *d_shared[2] = {nullptr, nullptr};
    bool shared_owner = false, shared_ready = false;
    bool shared_local = false;         // the frame belongs to another context of THIS process (peer access, no IPC, no NCCL)
    int parity = 0;
    int *d_barrier = nullptr;
    // sequence words behind the two frames in rank 0's allocation: flags[rank] = last frame this rank finished,
    // go = last frame rank 0 started (back-pressure: rank r starts frame k only when go >= k - 1)
    unsigned *d_flags = nullptr, *d_go = nullptr;
    // dynamic ownership of row groups across GPUs (WaveDyn in rt_wave.cuh): fused gather + kernel 1
    unsigned long long *d_tiles = nullptr;   // [2] behind go in rank 0's allocation: the job's group counters
    int *d_slot_group = nullptr;     // this GPU's slot -> group table
    int dyn_max_slots = 0, dyn_n_groups = 0, dyn_group_rows = 0;
    int dynamic_tiles = 0;           // option: 1 = on, 0 (default) = every rank renders its own strips — measured faster on the
                                     // bundled frames at 2, 4 and 8 GPUs (profiles/r2_scaling_8gpu.json)
    bool dynamic_now = false;        // ... in effect for the frame being issued
    unsigned *h_timed_out = nullptr, *d_timed_out = nullptr;   // page-locked, mapped: the wait kernel's timeout report
    unsigned seq = 0;
    int flag_barrier = 1;            // 1: flag words (default), 0: 4-byte NCCL all-reduce
    const uchar4 *last_frame = nullptr;   // what rt_b200_gather hands out

    cudaStream_t stream = nullptr, copy_stream = nullptr, stream2 = nullptr;   // stream2: odd bands of a streamed frame
    cudaEvent_t ev_join = nullptr;
    int band_overlap = 1;                 // streamed frames: alternate the bands between two streams (tail overlap)
    cudaEvent_t ev_start = nullptr, ev_render = nullptr, ev_end = nullptr, ev_copy = nullptr;
    cudaEvent_t ev_band[RT_MAX_BANDS] = {};
    int host_bands = 0;                   // 0 = automatic
    int band_taper = -1;                  // 0 = equal bands; r in 1..99: band k holds (r/100)^k of the rows of band 0
                                          // (shrinking); 100 + r: the same sizes in reverse order (growing);
                                          // -1 = automatic: whichever of 0 / 65 / 165 delivered this scene's recent frames fastest
    float taper_ms[3] = {0.f, 0.f, 0.f};  // automatic mode: smoothed frame time with [0] equal, [1] shrinking, [2] growing bands
    unsigned taper_frames = 0;
    float ms_render = 0, ms_gather = 0;

    ncclComm_t comm = nullptr;
};

namespace {

// rows of one rank's packed buffer
int packed_rows(int height, int strip_rows, int world)
{
    if (world == 1) return height;
    const int n_strips = (height + strip_rows - 1) / strip_rows;
    return ((n_strips + world - 1) / world) * strip_rows;
}

int free_buffers(rt_b200_ctx *c)
{
    if (c->d_local) cudaFree(c->d_local);
    if (c->d_gathered) cudaFree(c->d_gathered);
    if (c->d_frame) cudaFree(c->d_frame);
    c->d_local = c->d_gathered = c->d_frame = nullptr;
    return RT_B200_OK;
}

int ensure_buffers(rt_b200_ctx *c)
{
    if (!c->buffers_dirty) return RT_B200_OK;
    free_buffers(c);
    c->local_rows = packed_rows(c->height, c->strip_rows, c->world);
    const size_t row_bytes = sizeof(uchar4) * (size_t) c->width;
    CUDA_TRY(cudaMalloc((void **) &c->d_local, row_bytes * c->local_rows));
    if (c->world > 1) {
        CUDA_TRY(cudaMalloc((void **) &c->d_gathered, row_bytes * c->local_rows * c->world));
        CUDA_TRY(cudaMalloc((void **) &c->d_frame, row_bytes * c->height));
        CUDA_TRY(cudaMemsetAsync(c->d_gathered, 0, row_bytes * c->local_rows * c->world, c->stream));
        CUDA_TRY(cudaMemsetAsync(c->d_frame, 0, row_bytes * c->height, c->stream));
    }
    c->buffers_dirty = false;
    c->rendered = false;
    return RT_B200_OK;
}

// Structural class of one object: which coefficients are exactly zero.
// (Re)build the sphere grid at `density` cells per sphere (<= 0: the builder's default) and point the view at it.
int build_grid(rt_b200_ctx *c, double density)
{
    if (c->stream) cudaStreamSynchronize(c->stream);
    cudaFree(c->d_grid_cells);
    cudaFree(c->d_grid_items);
    c->d_grid_cells = c->d_grid_items = nullptr;
    c->view.grid = rt::SphereGrid{};
    std::memset(c->grid_info, 0, sizeof(c->grid_info));
    const int n = (int) c->sphere_r2.size();
    if (n == 0) return RT_B200_OK;
    rt::SphereGridHost H = rt::build_sphere_grid(c->sphere_centres.data(), c->sphere_r2.data(), n, density);
    cudaError_t e = upload(H.cell_start, &c->d_grid_cells);
    if (e == cudaSuccess) e = upload(H.items, &c->d_grid_items);
    if (e != cudaSuccess) return fail(RT_B200_ERR_CUDA, "upload sphere grid: %s", cudaGetErrorString(e));
    H.g.cell_start = c->d_grid_cells;
    H.g.items = c->d_grid_items;
    c->view.grid = H.g;
    c->grid_info[0] = (uint32_t) H.g.n[0];
    c->grid_info[1] = (uint32_t) H.g.n[1];
    c->grid_info[2] = (uint32_t) H.g.n[2];
    c->grid_info[3] = (uint32_t) H.g.n_gridded;
    c->grid_info[4] = (uint32_t) (H.g.loose_end - H.g.loose_begin);
    c->grid_info[5] = (uint32_t) H.g.loose_begin;
    return RT_B200_OK;
}

// Row-band boundaries of a streamed frame (rt_b200_update_to_host): n_bands bands over `rows` rows, every boundary a
// multiple of `unit` rows.  taper 0: equal bands; 1..99: band k holds (taper/100)^k of the rows of band 0 (shrinking);
// 101..199: the same sizes in reverse order (growing).  Every band gets at least one unit (n_bands <= ceil(rows / unit)).
void band_layout(int rows, int n_bands, int unit, int taper, int *start /* [n_bands + 1] */)
{
    const int band_rows = ((rows + n_bands - 1) / n_bands + unit - 1) / unit * unit;
    for (int b = 0; b <= n_bands; b++) start[b] = std::min(b * band_rows, rows);
    if (taper <= 0 || n_bands < 2) return;
    const int units = (rows + unit - 1) / unit;
    const bool growing = taper > 100;
    const double r = (growing ? taper - 100 : taper) / 100.0;
    double w[RT_MAX_BANDS], sum = 0.0;
    for (int b = 0; b < n_bands; b++) sum += (w[b] = std::pow(r, growing ? n_bands - 1 - b : b));
    double acc = 0.0;
    int prev = 0;
    for (int b = 1; b < n_bands; b++) {
        acc += w[b - 1];
        int u = (int) std::llround(acc / sum * units);
        u = std::max(prev + 1, std::min(u, units - (n_bands - b)));
        start[b] = std::min(u * unit, rows);
        prev = u;
    }
    start[n_bands] = rows;
}

rt::ObjClass classify(const double *k)
{
    for (int i = rt::X3; i <= rt::XYZ; i++)
        if (k[i] != 0.0) return rt::CLS_CUBIC;
    const bool cross = k[rt::XY] != 0.0 || k[rt::XZ] != 0.0 || k[rt::YZ] != 0.0;
    const bool squares = k[rt::X2] != 0.0 || k[rt::Y2] != 0.0 || k[rt::Z2] != 0.0;
    if (!cross && !squares) return rt::CLS_PLANE;
    if (!cross && k[rt::X2] == 1.0 && k[rt::Y2] == 1.0 && k[rt::Z2] == 1.0) return rt::CLS_SPHERE;
    return rt::CLS_QUADRIC;
}

} // namespace

extern "C" {

const char *rt_b200_last_error(void) { return g_last_error.c_str(); }

const char *rt_b200_version(void) { return "rt_b200 0.1 sm_100a"; }

int rt_b200_upload_scene(rt_b200_ctx **ctx_out, uint32_t px_width, uint32_t px_height, double vertical_fov_rad,
                         const float bg_color[3], uint32_t max_reflections, const void *objects, size_t n_objects,
                         const void *lights, size_t n_lights, int device, int rank, int world_size)
{
    if (!ctx_out) return fail(RT_B200_ERR_ARG, "ctx_out is null");
    *ctx_out = nullptr;
    if (px_width == 0 || px_height == 0) return fail(RT_B200_ERR_ARG, "empty frame %ux%u", px_width, px_height);
    if (!bg_color) return fail(RT_B200_ERR_ARG, "bg_color is null");
    if ((n_objects && !objects) || (n_lights && !lights)) return fail(RT_B200_ERR_ARG, "null scene array");
    if (world_size < 1 || rank < 0 || rank >= world_size)
        return fail(RT_B200_ERR_ARG, "bad rank %d of %d", rank, world_size);
    if (n_objects > (size_t) 1 << 24 || n_lights > (size_t) 1 << 20) return fail(RT_B200_ERR_ARG, "scene too large");

    int n_dev = 0;
    CUDA_TRY(cudaGetDeviceCount(&n_dev));
    if (device < 0 || device >= n_dev) return fail(RT_B200_ERR_CUDA, "no CUDA device %d (found %d)", device, n_dev);
    CUDA_TRY(cudaSetDevice(device));

    rt_b200_ctx *c = new rt_b200_ctx();
    c->device = device;
    c->rank = rank;
    c->world = world_size;
    c->width = (int) px_width;
    c->height = (int) px_height;
    c->vertical_fov = vertical_fov_rad;
    c->aspect = (double) px_width / px_height;          // Scene::aspect_ratio, scene.h:32-33
    c->tan_half_fov = std::tan(0.5 * vertical_fov_rad); // update-cpu.cpp:27
    std::memcpy(c->bg, bg_color, sizeof(c->bg));
    c->max_reflections = (int) max_reflections;

    // ---- classify and repack --------------------------------------------------------------------
    std::vector<double4> plane, sphere;
    std::vector<float4> sphere_f, material(n_objects), material_pi(n_objects);
    std::vector<double> quadric, cubic, coef_full(20 * n_objects);
    std::vector<int> orig[4];
    std::vector<unsigned> cubic_mask;
    float center_max = 0.0f, r2_max = 0.0f;
    const char *obj_bytes = static_cast<const char *>(objects);
    for (size_t i = 0; i < n_objects; i++) {
        double k[20];
        float refl, color[3];
        std::memcpy(k, obj_bytes + RT_B200_OBJECT_BYTES * i, 160);
        std::memcpy(&refl, obj_bytes + RT_B200_OBJECT_BYTES * i + 160, 4);
        std::memcpy(color, obj_bytes + RT_B200_OBJECT_BYTES * i + 164, 12);
        std::memcpy(&coef_full[20 * i], k, 160);
        material[i] = make_float4(refl, color[0], color[1], color[2]);
        const float pi_f = 3.14159274101257324f;   // glm::pi<float>(); IEEE FP32 division here as on the device
        material_pi[i] = make_float4(refl, color[0] / pi_f, color[1] / pi_f, color[2] / pi_f);
        if (!(fabsf(color[0]) < 1e5f && fabsf(color[1]) < 1e5f && fabsf(color[2]) < 1e5f)) c->cull_safe = false;
        rt::ObjClass cls = classify(k);
        orig[cls].push_back((int) i);
        switch (cls) {
            case rt::CLS_PLANE:
                plane.push_back(make_double4(k[rt::LX], k[rt::LY], k[rt::LZ], k[rt::CC]));
                break;
            case rt::CLS_SPHERE: {
                sphere.push_back(make_double4(k[rt::LX], k[rt::LY], k[rt::LZ], k[rt::CC]));
                // |p|^2 + l.p + c = 0  ->  centre -l/2, r^2 = |centre|^2 - c
                double cx = -0.5 * k[rt::LX], cy = -0.5 * k[rt::LY], cz = -0.5 * k[rt::LZ];
                double r2 = cx * cx + cy * cy + cz * cz - k[rt::CC];
                sphere_f.push_back(make_float4((float) cx, (float) cy, (float) cz, (float) r2));
                c->sphere_centres.insert(c->sphere_centres.end(), {cx, cy, cz});
                c->sphere_r2.push_back(r2);
                r2_max = fmaxf(r2_max, nextafterf((float) fabs(r2), INFINITY));
                center_max = fmaxf(center_max, fmaxf(fabsf((float) cx), fmaxf(fabsf((float) cy), fabsf((float) cz))));
                break;
            }
            case rt::CLS_QUADRIC:
                for (int j = rt::X2; j <= rt::CC; j++) quadric.push_back(k[j]);
                break;
            case rt::CLS_CUBIC: {
                unsigned mask = 0;
                for (int j = 0; j < 20; j++) {
                    cubic.push_back(k[j]);
                    if (k[j] != 0.0) mask |= 1u << j;
                }
                cubic_mask.push_back(mask);
                break;
            }
        }
    }
    std::vector<int> orig_index;
    for (int cls = 0; cls < 4; cls++) {
        c->class_counts[cls] = (uint32_t) orig[cls].size();
        orig_index.insert(orig_index.end(), orig[cls].begin(), orig[cls].end());
    }
    std::vector<rt::LightDev> ldev(n_lights);
    const char *light_bytes = static_cast<const char *>(lights);
    for (size_t j = 0; j < n_lights; j++) {
        const char *rec = light_bytes + RT_B200_LIGHT_BYTES * j;
        double p[3];
        float col[3];
        std::memcpy(p, rec + 8, 24);
        std::memcpy(col, rec + 32, 12);
        ldev[j] = rt::LightDev{p[0], p[1], p[2], col[0], col[1], col[2], rec[0] != 0 ? 1 : 0};
        if (!(fabsf(col[0]) < 1e15f && fabsf(col[1]) < 1e15f && fabsf(col[2]) < 1e15f)) c->cull_safe = false;
    }

    int rc = RT_B200_OK;
    auto up = [&](cudaError_t e, const char *what) {
        if (e != cudaSuccess && rc == RT_B200_OK) rc = fail(RT_B200_ERR_CUDA, "upload %s: %s", what, cudaGetErrorString(e));
    };
    up(upload(plane, &c->d_plane), "planes");
    up(upload(sphere, &c->d_sphere), "spheres");
    up(upload(sphere_f, &c->d_sphere_f), "sphere filter");
    up(upload(quadric, &c->d_quadric), "quadrics");
    up(upload(cubic, &c->d_cubic), "cubics");
    up(upload(coef_full, &c->d_coef_full), "coefficients");
    up(upload(material, &c->d_material), "materials");
    up(upload(material_pi, &c->d_material_pi), "materials / pi");
    up(upload(orig_index, &c->d_orig), "index map");
    up(upload(ldev, &c->d_lights), "lights");
    up(upload(cubic_mask, &c->d_cubic_mask), "cubic masks");
    // one {next tile, CTAs finished} pair per band, so that the bands of a streamed frame may overlap; pair 0 otherwise
    up(cudaMalloc((void **) &c->d_tile_counter, 2 * RT_MAX_BANDS * sizeof(unsigned long long)), "tile counter");
    if (c->d_tile_counter) up(cudaMemset(c->d_tile_counter, 0, 2 * RT_MAX_BANDS * sizeof(unsigned long long)), "tile counter");
    c->sphere_r2_max = r2_max;
    // work counters, and behind them the per-CTA timeline of the diagnostic build (option "timeline"; 128 KB)
    up(cudaMalloc((void **) &c->d_counters, sizeof(unsigned long long) * (rt::CNT_COUNT + RT_TL_MAX_CTAS * RT_TL_WORDS)), "counters");
    up(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking), "stream");
    up(cudaEventCreate(&c->ev_start), "event");
    up(cudaEventCreate(&c->ev_render), "event");
    up(cudaEventCreate(&c->ev_end), "event");
    up(cudaEventCreate(&c->ev_copy), "event");
    up(cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking), "copy stream");
    up(cudaStreamCreateWithFlags(&c->stream2, cudaStreamNonBlocking), "second render stream");
    up(cudaEventCreateWithFlags(&c->ev_join, cudaEventDisableTiming), "event");
    for (int b = 0; b < RT_MAX_BANDS; b++) up(cudaEventCreateWithFlags(&c->ev_band[b], cudaEventDisableTiming), "event");
    up(cudaMalloc((void **) &c->d_barrier, sizeof(int)), "barrier word");
    up(cudaHostAlloc((void **) &c->h_timed_out, sizeof(unsigned), cudaHostAllocMapped), "timeout word");
    if (c->h_timed_out) {
        *c->h_timed_out = 0;
        up(cudaHostGetDevicePointer((void **) &c->d_timed_out, c->h_timed_out, 0), "timeout word");
    }
    if (rc != RT_B200_OK) {
        rt_b200_destroy(c);
        return rc;
    }
    cudaMemset(c->d_counters, 0, sizeof(unsigned long long) * rt::CNT_COUNT);

    rt::SceneView &v = c->view;
    v.n_plane = (int) plane.size();
    v.n_sphere = (int) sphere.size();
    v.n_quadric = (int) quadric.size() / 10;
    v.n_cubic = (int) cubic.size() / 20;
    v.n_objects = (int) n_objects;
    v.n_lights = (int) n_lights;
    v.plane = c->d_plane;
    v.sphere = c->d_sphere;
    v.sphere_f = c->d_sphere_f;
    v.quadric = c->d_quadric;
    v.cubic = c->d_cubic;
    v.orig_index = c->d_orig;
    v.coef_full = c->d_coef_full;
    v.material = c->d_material;
    v.material_pi = c->d_material_pi;
    v.lights = c->d_lights;
    v.sphere_center_max = center_max;

    c->wave_fits = rt::wave_scene_fits(v);
    rc = build_grid(c, 0.0);
    if (rc == RT_B200_OK) rc = ensure_buffers(c);
    if (rc != RT_B200_OK) {
        rt_b200_destroy(c);
        return rc;
    }
    *ctx_out = c;
    return RT_B200_OK;
}

} // extern "C"

namespace {

// What the wavefront kernel's launcher needs from the context (the frame hand-shake words are added per frame).
rt::WaveLaunch wave_launch_of(const rt_b200_ctx *c, int counter_pair)
{
    return rt::WaveLaunch{c->d_tile_counter + 2 * counter_pair, c->d_cubic_mask, c->sphere_r2_max, c->shadow_cache, c->wave_variant,
                          c->cubic_filter, (c->cull_backfacing && c->cull_safe) ? 1 : 0,
                          (c->accel > 0 || (c->accel < 0 && c->view.n_sphere >= RT_GRID_AUTO_SPHERES)) ? 1 : 0,
                          nullptr, nullptr, nullptr, 0u, c->rank, c->world, nullptr, nullptr, 0, 0, 0, 0, c->lone_kernel, c->spheres_smem};
}

// One render launch of the selected kernel over the packed rows [f.row0, f.row0 + f.rows).
// stream / counter_pair: the stream to launch on and which tile-counter pair the wavefront kernel uses (bands that may
// run concurrently need their own).
int launch_kernel(rt_b200_ctx *c, const rt::FrameParams &f, cudaStream_t stream, int counter_pair)
{
    if (c->kernel_now() == 0) {
        CUDA_TRY(rt::launch_render_exact(c->view, f, stream));
    }
    else {
        if (c->tile_counter_dirty) {
            CUDA_TRY(cudaMemsetAsync(c->d_tile_counter, 0, 2 * RT_MAX_BANDS * sizeof(unsigned long long), stream));
            c->tile_counter_dirty = false;
        }
        rt::WaveLaunch wl = wave_launch_of(c, counter_pair);
        if (c->sync_in_kernel) {   // fused gather: the render kernel does the frame hand-shake itself
            wl.sync_go = c->d_go;
            wl.sync_flags = c->d_flags;
            wl.sync_timed_out = c->d_timed_out;
            wl.sync_seq = c->seq;
            if (c->dynamic_now) {
                wl.dyn_counters = c->d_tiles;
                wl.dyn_slot_group = c->d_slot_group;
                wl.dyn_n_groups = c->dyn_n_groups;
                wl.dyn_group_rows = c->dyn_group_rows;
                wl.dyn_max_slots = c->dyn_max_slots;
                wl.dyn_ahead = RT_DYN_AHEAD;
            }
        }
        const cudaError_t e = rt::launch_render_wave(c->view, f, wl, stream);
        if (e != cudaSuccess) {
            c->tile_counter_dirty = true;
            return fail(RT_B200_ERR_CUDA, "launch_render_wave: %s", cudaGetErrorString(e));
        }
    }
    return RT_B200_OK;
}

// update() proper.  host_sink != nullptr (single GPU): the frame is rendered in row bands and every finished band is
// copied to the host on a second stream while the next band renders, so the device->host transfer of the frame
// overlaps the rendering instead of following it.
int render_frame(rt_b200_ctx *c, const double camera_matrix[16], void *host_sink, size_t host_bytes, float *elapsed_ms)
{
    if (!c || !camera_matrix) return fail(RT_B200_ERR_ARG, "null argument");
    CUDA_TRY(cudaSetDevice(c->device));
    int rc = ensure_buffers(c);
    if (rc != RT_B200_OK) return rc;
    const size_t frame_bytes = sizeof(uchar4) * (size_t) c->width * c->height;
    if (host_sink && host_bytes < frame_bytes)
        return fail(RT_B200_ERR_ARG, "host buffer holds %zu bytes, frame needs %zu", host_bytes, frame_bytes);

    rt::FrameParams f;
    std::memcpy(f.m, camera_matrix, sizeof(f.m));
    // origin = M * (0,0,0,1) in GLM's association: (m0*0 + m1*0) + (m2*0 + m3*1), update-cpu.cpp:123
    for (int r = 0; r < 3; r++)
        f.origin[r] = (f.m[r] * 0.0 + f.m[4 + r] * 0.0) + (f.m[8 + r] * 0.0 + f.m[12 + r] * 1.0);
    f.aspect = c->aspect;
    f.tan_half_fov = c->tan_half_fov;
    std::memcpy(f.bg, c->bg, sizeof(f.bg));
    f.max_reflections = c->max_reflections;
    f.width = c->width;
    f.height = c->height;
    f.strip_rows = (c->world == 1) ? c->height : c->strip_rows;
    f.rank = c->rank;
    f.world = c->world;
    f.local_rows = c->local_rows;
    f.row0 = 0;
    f.rows = c->local_rows;
    f.counters = c->count ? c->d_counters : nullptr;
    f.timeline = c->timeline ? c->d_counters + rt::CNT_COUNT : nullptr;
    if (c->timeline) CUDA_TRY(cudaMemsetAsync(c->d_counters + rt::CNT_COUNT, 0, sizeof(unsigned long long) * RT_TL_MAX_CTAS * RT_TL_WORDS, c->stream));
    // host frame shared by all ranks (rt_b200_host_alloc_shared): every rank copies its own strips to the host
    SharedBlock *blk = (host_sink && c->world > 1) ? find_block(host_sink) : nullptr;
    if (blk && blk->bytes < frame_bytes)
        return fail(RT_B200_ERR_ARG, "shared host block holds %zu bytes, frame needs %zu", blk->bytes, frame_bytes);
    // fused gather: needs the frame attached, and either a communicator (flag_barrier = 0 uses an all-reduce) or the
    // flag words alone (contexts of one process, rt_b200_frame_attach_peer)
    const bool shared = c->world > 1 && c->shared_ready && !blk && (c->comm || (c->shared_local && c->flag_barrier));
    if (shared) {
        c->parity ^= 1;                 // rank 0 may still be reading the previous frame: write the other buffer
        f.out = c->d_shared[c->parity];
        f.direct = 1;
    }
    else {
        f.out = c->d_local;
        f.direct = 0;
    }

    // row bands of the streamed host copy: ~4 MB each, whole 16-row tiles (and whole strips when ranks share the
    // host frame)
    int n_bands = 1;
    const bool streamed = host_sink != nullptr && (c->world == 1 || blk);
    int band_unit = 16;
    if (blk) {
        int a = 16, b = c->strip_rows;
        while (b) { const int t = a % b; a = b; b = t; }
        band_unit = 16 / a * c->strip_rows;   // lcm(16, strip_rows)
    }
    if (streamed) {
        const size_t local_bytes = sizeof(uchar4) * (size_t) c->width * c->local_rows;
        n_bands = c->host_bands > 0 ? c->host_bands : (int) (local_bytes / (4u << 20));   // measured: smaller bands do not pay, at 1 or 8 GPUs
        if (n_bands > RT_MAX_BANDS) n_bands = RT_MAX_BANDS;
        if (n_bands > (c->local_rows + band_unit - 1) / band_unit) n_bands = (c->local_rows + band_unit - 1) / band_unit;
        if (n_bands < 1) n_bands = 1;
    }
    // Band boundaries (band_layout).  Equal bands unless a taper is chosen below.
    int band_start[RT_MAX_BANDS + 1];
    // Which is better depends on what bounds the frame.  Rendering slower than the PCIe copy: shrinking bands hide all but
    // the last, small copy.  Copy slower than rendering: the frame takes (first band's render) + (whole copy), so growing
    // bands — a tiny first one — start the copy engine earliest.  Automatic mode tries the three shapes on the first
    // streamed frames of a scene, keeps the fastest and re-tries another every 32nd frame (the camera moves).
    int taper = c->band_taper;
    int taper_mode = -1;
    if (taper < 0) {
        taper = 0;
        if (streamed && !blk && n_bands > 1) {
            const unsigned k = c->taper_frames++;
            if (k < 3) taper_mode = (int) k;
            else {
                taper_mode = 0;
                for (int mde = 1; mde < 3; mde++)
                    if (c->taper_ms[mde] < c->taper_ms[taper_mode]) taper_mode = mde;
                if (k % 32 == 0) taper_mode = (taper_mode + 1 + (int) ((k / 32) & 1)) % 3;
            }
            taper = taper_mode == 0 ? 0 : (taper_mode == 1 ? 65 : 165);
        }
    }
    band_layout(c->local_rows, n_bands, band_unit, (streamed && !blk) ? taper : 0, band_start);
    unsigned host_seq = 0;
    if (blk) {
        host_seq = ++blk->seq;
        SharedTrailer *t = blk->trailer();
        if (c->rank == 0) __atomic_store_n(&t->go, host_seq, __ATOMIC_RELEASE);
    }
    static const bool tracing = std::getenv("RT_B200_TRACE") != nullptr;
    const double t_enter = tracing ? now_seconds() : 0.0;
    double t_go = 0.0;
    bool go_seen = (blk == nullptr) || c->rank == 0;

    if (c->count) CUDA_TRY(cudaMemsetAsync(c->d_counters, 0, sizeof(unsigned long long) * rt::CNT_COUNT, c->stream));
    CUDA_TRY(cudaEventRecord(c->ev_start, c->stream));
    // kernel 1 does the go / flag hand-shake of the fused gather in its own prologue / epilogue: no extra launches
    c->sync_in_kernel = shared && c->flag_barrier && c->kernel_now() == 1;
    // Dynamic tile ownership: the frame is one tile space for all GPUs; this rank's launch describes
    // the WHOLE image (as a single-rank frame) and its CTAs pull tiles from the job's counter in rank 0's memory.
    c->dynamic_now = c->sync_in_kernel && c->dynamic_tiles && c->d_tiles != nullptr;
    if (c->dynamic_now) {
        // groups of one tile height of rows (8; 16 for the 512-thread variant): re-made when the variant changes
        const int group_rows = rt::wave_tile_rows(c->view, wave_launch_of(c, 0));
        if (!c->d_slot_group || group_rows != c->dyn_group_rows) {
            CUDA_TRY(cudaStreamSynchronize(c->stream));
            cudaFree(c->d_slot_group);
            c->d_slot_group = nullptr;
            c->dyn_group_rows = group_rows;
            c->dyn_n_groups = (c->height + group_rows - 1) / group_rows;
            c->dyn_max_slots = c->dyn_n_groups + RT_DYN_AHEAD + 1;
            CUDA_TRY(cudaMalloc((void **) &c->d_slot_group, sizeof(int) * (size_t) c->dyn_max_slots));
            CUDA_TRY(rt::launch_dyn_table_init(c->d_slot_group, c->dyn_max_slots, RT_DYN_AHEAD, c->dyn_n_groups, c->rank, c->world, c->stream));
        }
        f.strip_rows = c->height;
        f.rank = 0;
        f.world = 1;
        f.local_rows = c->height;
        band_start[0] = 0;
        band_start[1] = c->dyn_n_groups * c->dyn_group_rows;   // whole groups (rows past the image are masked by local_rows);
                                                               // n_bands == 1: the fused gather never streams
    }
    if (shared && c->flag_barrier) {
        c->seq++;
        if (!c->sync_in_kernel) {
            if (c->rank == 0) CUDA_TRY(rt::launch_sync_store(c->d_go, c->seq, c->stream));
            else CUDA_TRY(rt::launch_sync_wait(c->d_go, 0, 1, c->seq - 1, c->d_timed_out, c->stream));
        }
    }
    c->launches = 0;
    const bool overlap = streamed && n_bands > 1 && c->band_overlap && c->kernel_now() == 1 && !c->tile_counter_dirty;
    for (int b = 0; b < n_bands; b++) {
        f.row0 = band_start[b];
        f.rows = band_start[b + 1] - band_start[b];
        if (f.rows <= 0) break;
        // Overlap: even bands on the render stream, odd bands on a second one, each band with its own tile counter.  The
        // persistent CTAs of band b fill the GPU, so band b + 1 only gets the SM slots band b's tail leaves empty —
        // the bands still finish in order, but the tail of one no longer idles the machine before the next starts.
        cudaStream_t bs = (overlap && (b & 1)) ? c->stream2 : c->stream;
        if (overlap && b == 1) CUDA_TRY(cudaStreamWaitEvent(c->stream2, c->ev_start, 0));
        rc = launch_kernel(c, f, bs, overlap ? b : 0);
        if (rc != RT_B200_OK) return rc;
        c->launches++;
        if (streamed) {
            CUDA_TRY(cudaEventRecord(c->ev_band[b], bs));
            CUDA_TRY(cudaStreamWaitEvent(c->copy_stream, c->ev_band[b], 0));
            const size_t off = (size_t) f.row0 * c->width;
            if (!blk) {
                CUDA_TRY(cudaMemcpyAsync(static_cast<uchar4 *>(host_sink) + off, c->d_local + off,
                                         sizeof(uchar4) * (size_t) f.rows * c->width, cudaMemcpyDeviceToHost, c->copy_stream));
            }
            else {
                if (!go_seen) {
                    // the block may still hold the frame rank 0's caller is reading: wait until rank 0 has entered
                    // this frame too (rendering is already queued, only the copies wait)
                    const double t0 = tracing ? now_seconds() : 0.0;
                    if (!host_wait_word(&blk->trailer()->go, host_seq, 20.0)) {
                        // nothing of this frame may still be in flight when the caller gets the error
                        cudaStreamSynchronize(c->stream);
                        cudaStreamSynchronize(c->stream2);
                        cudaStreamSynchronize(c->copy_stream);
                        return fail(RT_B200_ERR_STATE, "rank 0 did not enter frame %u of the shared host block (timeout)", host_seq);
                    }
                    go_seen = true;
                    if (tracing) t_go = now_seconds() - t0;
                }
                // local strip s of this rank is image strip s * world + rank: one pitched copy scatters the band's
                // strips to their image rows; a ragged last strip goes separately; padding strips are not copied
                const int n_strips = (c->height + c->strip_rows - 1) / c->strip_rows;
                const int s0 = f.row0 / c->strip_rows, s1 = (f.row0 + f.rows) / c->strip_rows;
                const size_t strip_bytes = sizeof(uchar4) * (size_t) c->strip_rows * c->width;
                int full = 0, ragged = -1;
                for (int sl = s0; sl < s1; sl++) {
                    const int g = sl * c->world + c->rank;
                    if (g >= n_strips) break;
                    if ((g + 1) * c->strip_rows <= c->height) full++;
                    else ragged = sl;
                }
                char *dst = static_cast<char *>(host_sink) + ((size_t) s0 * c->world + c->rank) * strip_bytes;
                const char *src = reinterpret_cast<const char *>(c->d_local) + (size_t) s0 * strip_bytes;
                if (full > 0)
                    CUDA_TRY(cudaMemcpy2DAsync(dst, strip_bytes * c->world, src, strip_bytes, strip_bytes, full,
                                               cudaMemcpyDeviceToHost, c->copy_stream));
                if (ragged >= 0) {
                    const int g = ragged * c->world + c->rank;
                    const size_t rows = (size_t) c->height - (size_t) g * c->strip_rows;
                    CUDA_TRY(cudaMemcpyAsync(static_cast<char *>(host_sink) + (size_t) g * strip_bytes,
                                             reinterpret_cast<const char *>(c->d_local) + (size_t) ragged * strip_bytes,
                                             sizeof(uchar4) * rows * c->width, cudaMemcpyDeviceToHost, c->copy_stream));
                }
            }
        }
    }
    if (overlap) {
        CUDA_TRY(cudaEventRecord(c->ev_join, c->stream2));
        CUDA_TRY(cudaStreamWaitEvent(c->stream, c->ev_join, 0));
    }
    CUDA_TRY(cudaEventRecord(c->ev_render, c->stream));
    c->last_frame = c->d_local;
    if (blk) {
        c->last_frame = nullptr;   // the frame is assembled in host memory; rt_b200_gather would see this rank's rows only
    }
    else if (c->world > 1 && (c->comm || shared)) {
        if (shared && c->flag_barrier) {
            // every rank's pixels are already in rank 0's frame: each rank publishes "frame seq done" in rank 0's
            // memory, and only rank 0 — the consumer of the frame — waits for all of them
            if (!c->sync_in_kernel) {
                CUDA_TRY(rt::launch_sync_store(c->d_flags + c->rank, c->seq, c->stream));
                if (c->rank == 0) CUDA_TRY(rt::launch_sync_wait(c->d_flags, 1, c->world - 1, c->seq, c->d_timed_out, c->stream));
            }
            c->last_frame = c->d_shared[c->parity];
        }
        else if (shared) {
            // same, with a 4-byte all-reduce as the frame-complete barrier (every rank waits for every rank)
            NCCL_TRY(g_nccl.AllReduce(c->d_barrier, c->d_barrier, 1, ncclInt, ncclSum, c->comm, c->stream));
            c->last_frame = c->d_shared[c->parity];
        }
        else {
            const size_t bytes = sizeof(uchar4) * (size_t) c->width * c->local_rows;
            NCCL_TRY(g_nccl.AllGather(c->d_local, c->d_gathered, bytes, ncclUint8, c->comm, c->stream));
            CUDA_TRY(rt::launch_unpack_strips(c->d_gathered, c->d_frame, c->width, c->height, c->strip_rows, c->world,
                                              c->local_rows, c->stream));
            c->last_frame = c->d_frame;
        }
    }
    else if (c->world > 1) {
        c->last_frame = nullptr;   // assembled on demand by rt_b200_gather (this rank's rows only)
    }
    CUDA_TRY(cudaEventRecord(c->ev_end, c->stream));
    const double t_issued = tracing ? now_seconds() : 0.0;
    if (streamed) {
        CUDA_TRY(cudaEventRecord(c->ev_copy, c->copy_stream));
        CUDA_TRY(cudaStreamWaitEvent(c->stream, c->ev_copy, 0));   // keep the next frame from overwriting rows in flight
        CUDA_TRY(cudaEventRecord(c->ev_copy, c->stream));
        CUDA_TRY(cudaEventSynchronize(c->ev_copy));
    }
    CUDA_TRY(cudaEventSynchronize(c->ev_end));
    CUDA_TRY(cudaGetLastError());
    if (shared && c->flag_barrier) {
        if (*static_cast<volatile unsigned *>(c->h_timed_out)) {
            // report once and re-arm: the peers may recover; nothing of this frame is still in flight (events waited above)
            *static_cast<volatile unsigned *>(c->h_timed_out) = 0u;
            cudaStreamSynchronize(c->stream);
            return fail(RT_B200_ERR_STATE, "a peer rank did not finish its part of the frame (sync timeout)");
        }
    }
    CUDA_TRY(cudaEventElapsedTime(&c->ms_render, c->ev_start, c->ev_render));
    CUDA_TRY(cudaEventElapsedTime(&c->ms_gather, c->ev_render, c->ev_end));
    float total = c->ms_render + c->ms_gather;
    if (streamed) CUDA_TRY(cudaEventElapsedTime(&total, c->ev_start, c->ev_copy));
    if (taper_mode >= 0) c->taper_ms[taper_mode] = c->taper_ms[taper_mode] > 0.f ? 0.5f * (c->taper_ms[taper_mode] + total) : total;
    const double t_synced = tracing ? now_seconds() : 0.0;
    if (blk) {
        // this rank's strips are in host memory (the event wait above); publish that, and rank 0 — whose caller
        // reads the frame — waits for every rank
        SharedTrailer *t = blk->trailer();
        __atomic_store_n(&t->done[c->rank], host_seq, __ATOMIC_RELEASE);
        if (c->rank == 0)
            for (int r = 1; r < c->world; r++)
                if (!host_wait_word(&t->done[r], host_seq, 20.0))
                    return fail(RT_B200_ERR_STATE, "rank %d did not deliver its strips of frame %u (timeout)", r, host_seq);
    }
    if (tracing && blk) {
        const double t_done = now_seconds();
        c->trace[0] += t_issued - t_enter - t_go;   // enqueue kernels and copies
        c->trace[1] += t_go;                        // waiting for rank 0 to enter the frame
        c->trace[2] += t_synced - t_issued;         // waiting for the GPU (render + own copies)
        c->trace[3] += t_done - t_synced;           // rank 0: waiting for the other ranks' strips
        c->trace[4] += c->ms_render * 1e-3;
        c->trace_frames++;
    }
    c->rendered = true;
    if (elapsed_ms) *elapsed_ms = total;
    if (host_sink && !streamed) return rt_b200_gather(c, host_sink, host_bytes, nullptr);
    return RT_B200_OK;
}

} // namespace

extern "C" {

int rt_b200_update(rt_b200_ctx *c, const double camera_matrix[16], float *elapsed_ms)
{
    return render_frame(c, camera_matrix, nullptr, 0, elapsed_ms);
}

int rt_b200_update_to_host(rt_b200_ctx *c, const double camera_matrix[16], void *host_rgba8, size_t host_bytes,
                           float *elapsed_ms)
{
    if (!host_rgba8) return fail(RT_B200_ERR_ARG, "host buffer is null");
    return render_frame(c, camera_matrix, host_rgba8, host_bytes, elapsed_ms);
}

int rt_b200_host_alloc(void **ptr_out, size_t bytes)
{
    if (!ptr_out || bytes == 0) return fail(RT_B200_ERR_ARG, "bad host allocation request");
    *ptr_out = nullptr;
    CUDA_TRY(cudaHostAlloc(ptr_out, bytes, cudaHostAllocPortable));
    return RT_B200_OK;
}

int rt_b200_host_free(void *ptr)
{
    if (ptr) CUDA_TRY(cudaFreeHost(ptr));
    return RT_B200_OK;
}

int rt_b200_host_alloc_shared(const char *path, size_t bytes, int create, void **ptr_out)
{
    if (!path || !ptr_out || bytes == 0) return fail(RT_B200_ERR_ARG, "bad shared host allocation request");
    *ptr_out = nullptr;
    const size_t mapped = ((bytes + 4095) & ~(size_t) 4095) + 4096;
    const int fd = open(path, create ? (O_RDWR | O_CREAT | O_EXCL) : O_RDWR, 0600);
    if (fd < 0) return fail(RT_B200_ERR_ARG, "cannot %s %s: %s", create ? "create" : "open", path, std::strerror(errno));
    if (create && ftruncate(fd, (off_t) mapped) != 0) {   // zero-filled: all sequence words start at 0
        const int e = errno;
        close(fd);
        unlink(path);
        return fail(RT_B200_ERR_ARG, "cannot size %s to %zu bytes: %s", path, mapped, std::strerror(e));
    }
    struct stat st;
    if (fstat(fd, &st) != 0 || (size_t) st.st_size < mapped) {
        close(fd);
        return fail(RT_B200_ERR_ARG, "%s is smaller than the %zu bytes this frame needs", path, mapped);
    }
    void *base = mmap(nullptr, mapped, PROT_READ | PROT_WRITE, MAP_SHARED | MAP_POPULATE, fd, 0);
    close(fd);
    if (base == MAP_FAILED) {
        if (create) unlink(path);
        return fail(RT_B200_ERR_ARG, "cannot map %s: %s", path, std::strerror(errno));
    }
    const cudaError_t e = cudaHostRegister(base, mapped, cudaHostRegisterPortable);
    if (e != cudaSuccess) {
        munmap(base, mapped);
        if (create) unlink(path);
        return fail(RT_B200_ERR_CUDA, "cudaHostRegister(%s): %s", path, cudaGetErrorString(e));
    }
    SharedBlock *b = new SharedBlock;
    b->base = base;
    b->bytes = bytes;
    b->mapped = mapped;
    b->path = path;
    b->creator = create != 0;
    {
        std::lock_guard<std::mutex> lock(g_blocks_mutex);
        g_blocks.push_back(b);
    }
    *ptr_out = base;
    return RT_B200_OK;
}

int rt_b200_host_free_shared(void *ptr)
{
    if (!ptr) return RT_B200_OK;
    SharedBlock *b = nullptr;
    {
        std::lock_guard<std::mutex> lock(g_blocks_mutex);
        for (size_t i = 0; i < g_blocks.size(); i++)
            if (g_blocks[i]->base == ptr) {
                b = g_blocks[i];
                g_blocks.erase(g_blocks.begin() + (long) i);
                break;
            }
    }
    if (!b) return fail(RT_B200_ERR_ARG, "not a block of rt_b200_host_alloc_shared");
    cudaHostUnregister(b->base);
    munmap(b->base, b->mapped);
    if (b->creator) unlink(b->path.c_str());
    delete b;
    return RT_B200_OK;
}

int rt_b200_frame_export(rt_b200_ctx *c, void *handle_out)
{
    if (!c || !handle_out) return fail(RT_B200_ERR_ARG, "null argument");
    if (c->world < 2) return fail(RT_B200_ERR_ARG, "a shared frame needs world_size > 1");
    CUDA_TRY(cudaSetDevice(c->device));
    static_assert(sizeof(cudaIpcMemHandle_t) == RT_B200_IPC_HANDLE_BYTES, "cudaIpcMemHandle_t size");
    const size_t px = (size_t) c->width * c->height;
    if (!c->d_shared_base) {
        CUDA_TRY(cudaMalloc((void **) &c->d_shared_base, 2 * sizeof(uchar4) * px + 4096));
        CUDA_TRY(cudaMemset(c->d_shared_base, 0, 2 * sizeof(uchar4) * px + 4096));
        c->shared_owner = true;
    }
    if (!c->shared_owner) return fail(RT_B200_ERR_STATE, "this context maps another rank's frame");
    cudaIpcMemHandle_t h;
    CUDA_TRY(cudaIpcGetMemHandle(&h, c->d_shared_base));
    std::memcpy(handle_out, &h, sizeof(h));
    return RT_B200_OK;
}

int rt_b200_frame_attach(rt_b200_ctx *c, const void *handle)
{
    if (!c || !handle) return fail(RT_B200_ERR_ARG, "null argument");
    if (c->world < 2) return fail(RT_B200_ERR_ARG, "a shared frame needs world_size > 1");
    CUDA_TRY(cudaSetDevice(c->device));
    const size_t px = (size_t) c->width * c->height;
    if (!c->shared_owner) {
        if (c->d_shared_base) return fail(RT_B200_ERR_STATE, "a shared frame is already attached");
        cudaIpcMemHandle_t h;
        std::memcpy(&h, handle, sizeof(h));
        void *base = nullptr;
        CUDA_TRY(cudaIpcOpenMemHandle(&base, h, cudaIpcMemLazyEnablePeerAccess));
        c->d_shared_base = static_cast<uchar4 *>(base);
    }
    c->d_shared[0] = c->d_shared_base;
    c->d_shared[1] = c->d_shared_base + px;
    if (c->world > 1000) return fail(RT_B200_ERR_ARG, "too many ranks for the sync words");
    c->d_flags = reinterpret_cast<unsigned *>(c->d_shared_base + 2 * px);   // [world] + go
    c->d_go = c->d_flags + 1000;
    c->d_tiles = reinterpret_cast<unsigned long long *>(c->d_flags + 1008);
    c->seq = 0;
    c->shared_ready = true;
    CUDA_TRY(cudaMemset(c->d_barrier, 0, sizeof(int)));
    return RT_B200_OK;
}

int rt_b200_frame_attach_peer(rt_b200_ctx *c, rt_b200_ctx *owner)
{
    if (!c || !owner) return fail(RT_B200_ERR_ARG, "null argument");
    if (c->world < 2 || owner->world != c->world) return fail(RT_B200_ERR_ARG, "contexts of one multi-GPU job expected");
    if (c->width != owner->width || c->height != owner->height) return fail(RT_B200_ERR_ARG, "frame sizes differ");
    if (c->world > 1000) return fail(RT_B200_ERR_ARG, "too many ranks for the sync words");
    const size_t px = (size_t) c->width * c->height;
    if (c != owner) {
        if (c->d_shared_base) return fail(RT_B200_ERR_STATE, "a shared frame is already attached");
        if (!owner->d_shared_base || !owner->shared_owner) return fail(RT_B200_ERR_STATE, "the owner has not exported a frame (rt_b200_frame_export)");
        CUDA_TRY(cudaSetDevice(c->device));
        if (c->device != owner->device) {
            int can = 0;
            CUDA_TRY(cudaDeviceCanAccessPeer(&can, c->device, owner->device));
            if (!can) return fail(RT_B200_ERR_CUDA, "device %d cannot access device %d's memory", c->device, owner->device);
            const cudaError_t e = cudaDeviceEnablePeerAccess(owner->device, 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled)
                return fail(RT_B200_ERR_CUDA, "cudaDeviceEnablePeerAccess: %s", cudaGetErrorString(e));
            (void) cudaGetLastError();
        }
        c->d_shared_base = owner->d_shared_base;
        c->shared_local = true;
    }
    else {
        if (!c->shared_owner) return fail(RT_B200_ERR_STATE, "the owner has not exported a frame (rt_b200_frame_export)");
        CUDA_TRY(cudaSetDevice(c->device));
        c->shared_local = true;
    }
    c->d_shared[0] = c->d_shared_base;
    c->d_shared[1] = c->d_shared_base + px;
    c->d_flags = reinterpret_cast<unsigned *>(c->d_shared_base + 2 * px);
    c->d_go = c->d_flags + 1000;
    c->d_tiles = reinterpret_cast<unsigned long long *>(c->d_flags + 1008);
    c->seq = 0;
    c->shared_ready = true;
    return RT_B200_OK;
}

int rt_b200_gather(rt_b200_ctx *c, void *host_rgba8, size_t host_bytes, const void **device_rgba8)
{
    if (!c) return fail(RT_B200_ERR_ARG, "null context");
    if (!c->rendered) return fail(RT_B200_ERR_STATE, "rt_b200_gather before the first rt_b200_update");
    CUDA_TRY(cudaSetDevice(c->device));
    const size_t frame_bytes = sizeof(uchar4) * (size_t) c->width * c->height;
    const uchar4 *frame = c->last_frame;
    if (!frame) {
        // world > 1 without a communicator: expose this rank's rows in place, the others stay zero
        CUDA_TRY(cudaMemsetAsync(c->d_gathered, 0, sizeof(uchar4) * (size_t) c->width * c->local_rows * c->world, c->stream));
        CUDA_TRY(cudaMemcpyAsync(c->d_gathered + (size_t) c->rank * c->local_rows * c->width, c->d_local,
                                 sizeof(uchar4) * (size_t) c->width * c->local_rows, cudaMemcpyDeviceToDevice, c->stream));
        CUDA_TRY(rt::launch_unpack_strips(c->d_gathered, c->d_frame, c->width, c->height, c->strip_rows, c->world,
                                          c->local_rows, c->stream));
        CUDA_TRY(cudaStreamSynchronize(c->stream));
        frame = c->d_frame;
    }
    if (host_rgba8) {
        if (host_bytes < frame_bytes)
            return fail(RT_B200_ERR_ARG, "host buffer holds %zu bytes, frame needs %zu", host_bytes, frame_bytes);
        CUDA_TRY(cudaMemcpyAsync(host_rgba8, frame, frame_bytes, cudaMemcpyDeviceToHost, c->stream));
        CUDA_TRY(cudaStreamSynchronize(c->stream));
    }
    if (device_rgba8) *device_rgba8 = frame;
    return RT_B200_OK;
}

int rt_b200_destroy(rt_b200_ctx *c)
{
    if (c && c->trace_frames) {
        const double k = 1e6 / c->trace_frames;
        std::fprintf(stderr, "[rt_b200 trace] rank %d: %u shared-host frames, per frame: enqueue %.1f us, wait go %.1f us, "
                     "wait gpu %.1f us (render %.1f us), wait ranks %.1f us\n", c->rank, c->trace_frames,
                     c->trace[0] * k, c->trace[1] * k, c->trace[2] * k, c->trace[4] * k, c->trace[3] * k);
    }
    if (!c) return RT_B200_OK;
    cudaSetDevice(c->device);
    if (c->comm && g_nccl.CommDestroy) g_nccl.CommDestroy(c->comm);
    if (c->stream) cudaStreamSynchronize(c->stream);
    free_buffers(c);
    cudaFree(c->d_plane);
    cudaFree(c->d_sphere);
    cudaFree(c->d_sphere_f);
    cudaFree(c->d_quadric);
    cudaFree(c->d_cubic);
    cudaFree(c->d_coef_full);
    cudaFree(c->d_material);
    cudaFree(c->d_material_pi);
    cudaFree(c->d_orig);
    cudaFree(c->d_lights);
    cudaFree(c->d_counters);
    cudaFree(c->d_tile_counter);
    cudaFree(c->d_cubic_mask);
    cudaFree(c->d_grid_cells);
    cudaFree(c->d_grid_items);
    cudaFree(c->d_slot_group);
    if (c->ev_start) cudaEventDestroy(c->ev_start);
    if (c->ev_render) cudaEventDestroy(c->ev_render);
    if (c->ev_end) cudaEventDestroy(c->ev_end);
    if (c->ev_copy) cudaEventDestroy(c->ev_copy);
    for (int b = 0; b < RT_MAX_BANDS; b++)
        if (c->ev_band[b]) cudaEventDestroy(c->ev_band[b]);
    if (c->copy_stream) cudaStreamDestroy(c->copy_stream);
    if (c->stream2) cudaStreamDestroy(c->stream2);
    if (c->ev_join) cudaEventDestroy(c->ev_join);
    cudaFree(c->d_barrier);
    if (c->h_timed_out) cudaFreeHost(c->h_timed_out);
    if (c->d_shared_base) {
        if (c->shared_owner) cudaFree(c->d_shared_base);
        else if (!c->shared_local) cudaIpcCloseMemHandle(c->d_shared_base);   // a peer pointer of this process is not ours to free
    }
    if (c->stream) cudaStreamDestroy(c->stream);
    delete c;
    return RT_B200_OK;
}

int rt_b200_comm_unique_id(void *id_out)
{
    if (!id_out) return fail(RT_B200_ERR_ARG, "id_out is null");
    int rc = load_nccl();
    if (rc != RT_B200_OK) return rc;
    static_assert(sizeof(ncclUniqueId) == RT_B200_UNIQUE_ID_BYTES, "ncclUniqueId size");
    ncclUniqueId id;
    NCCL_TRY(g_nccl.GetUniqueId(&id));
    std::memcpy(id_out, &id, sizeof(id));
    return RT_B200_OK;
}

int rt_b200_comm_init(rt_b200_ctx *c, const void *id_bytes)
{
    if (!c || !id_bytes) return fail(RT_B200_ERR_ARG, "null argument");
    if (c->world < 2) return fail(RT_B200_ERR_ARG, "communicator needs world_size > 1");
    int rc = load_nccl();
    if (rc != RT_B200_OK) return rc;
    CUDA_TRY(cudaSetDevice(c->device));
    ncclUniqueId id;
    std::memcpy(&id, id_bytes, sizeof(id));
    NCCL_TRY(g_nccl.CommInitRank(&c->comm, c->world, id, c->rank));
    return RT_B200_OK;
}

int rt_b200_get_counters(rt_b200_ctx *c, rt_b200_counters *out)
{
    if (!c || !out) return fail(RT_B200_ERR_ARG, "null argument");
    static_assert(sizeof(rt_b200_counters) == sizeof(unsigned long long) * rt::CNT_COUNT, "counter layout");
    CUDA_TRY(cudaSetDevice(c->device));
    CUDA_TRY(cudaMemcpy(out, c->d_counters, sizeof(*out), cudaMemcpyDeviceToHost));
    return RT_B200_OK;
}

int rt_b200_set_option(rt_b200_ctx *c, const char *name, int64_t value)
{
    if (!c || !name) return fail(RT_B200_ERR_ARG, "null argument");
    if (!std::strcmp(name, "kernel")) {
        if (value < 0 || value > 1) return fail(RT_B200_ERR_ARG, "kernel must be 0 or 1");
        c->kernel_kind = (int) value;
    }
    else if (!std::strcmp(name, "strip_rows")) {
        if (value < 1 || value > 4096) return fail(RT_B200_ERR_ARG, "strip_rows out of range");
        if (c->strip_rows != (int) value) {
            c->strip_rows = (int) value;
            c->buffers_dirty = true;
        }
    }
    else if (!std::strcmp(name, "variant")) {
        if (value < 0 || value > 64) return fail(RT_B200_ERR_ARG, "variant out of range (see rt_wave_kernel.inl: RT_WAVE_VARIANTS)");
        c->wave_variant = (int) value;
    }
    else if (!std::strcmp(name, "shadow_cache")) {
        if (value < -1 || value > 1) return fail(RT_B200_ERR_ARG, "shadow_cache must be -1, 0 or 1");
        c->shadow_cache = (int) value;
    }
    else if (!std::strcmp(name, "flag_barrier")) {
        c->flag_barrier = value != 0;
    }
    else if (!std::strcmp(name, "spheres_smem")) {
        c->spheres_smem = value ? -1 : 0;
    }
    else if (!std::strcmp(name, "timeline")) {
        c->timeline = value ? 1 : 0;
    }
    else if (!std::strcmp(name, "lone_kernel")) {
        c->lone_kernel = value ? 1 : 0;
    }
    else if (!std::strcmp(name, "dynamic_tiles")) {
        c->dynamic_tiles = value ? 1 : 0;
    }
    else if (!std::strcmp(name, "cull_backfacing")) {
        c->cull_backfacing = value ? 1 : 0;
    }
    else if (!std::strcmp(name, "cubic_filter")) {
        c->cubic_filter = value ? 1 : 0;
    }
    else if (!std::strcmp(name, "accel")) {
        if (value < -1 || value > 1) return fail(RT_B200_ERR_ARG, "accel must be -1, 0 or 1");
        c->accel = (int) value;
    }
    else if (!std::strcmp(name, "grid_density")) {   // cells per sphere x 100 (0 = default)
        if (value < 0 || value > 100000) return fail(RT_B200_ERR_ARG, "grid_density must be 0..100000");
        return build_grid(c, (double) value / 100.0);
    }
    else if (!std::strcmp(name, "host_bands")) {
        if (value < 0 || value > RT_MAX_BANDS) return fail(RT_B200_ERR_ARG, "host_bands must be 0..%d", RT_MAX_BANDS);
        c->host_bands = (int) value;
    }
    else if (!std::strcmp(name, "band_overlap")) {
        c->band_overlap = value ? 1 : 0;
    }
    else if (!std::strcmp(name, "band_taper")) {
        if (value < -1 || value > 199 || value == 100) return fail(RT_B200_ERR_ARG, "band_taper must be -1..99 or 101..199");
        c->band_taper = (int) value;
        c->taper_frames = 0;
        c->taper_ms[0] = c->taper_ms[1] = c->taper_ms[2] = 0.f;
    }
    else if (!std::strcmp(name, "count")) {
        c->count = value != 0;
    }
    else {
        return fail(RT_B200_ERR_ARG, "unknown option '%s'", name);
    }
    return RT_B200_OK;
}

int rt_b200_strip_layout(uint32_t px_height, uint32_t strip_rows, int rank, int world_size, int32_t *image_rows,
                         size_t capacity, uint32_t *local_rows)
{
    if (px_height == 0 || strip_rows == 0 || world_size < 1 || rank < 0 || rank >= world_size || !local_rows)
        return fail(RT_B200_ERR_ARG, "bad strip layout query");
    const int eff_strip = (world_size == 1) ? (int) px_height : (int) strip_rows;
    const int rows = packed_rows((int) px_height, (int) strip_rows, world_size);
    *local_rows = (uint32_t) rows;
    if (image_rows) {
        if (capacity < (size_t) rows) return fail(RT_B200_ERR_ARG, "image_rows holds %zu entries, need %d", capacity, rows);
        for (int i = 0; i < rows; i++) {
            int y = rt::local_to_image_row(i, eff_strip, rank, world_size);
            image_rows[i] = y < (int) px_height ? y : -1;
        }
    }
    return RT_B200_OK;
}

int rt_b200_get_timing(rt_b200_ctx *c, float out_ms[2])
{
    if (!c || !out_ms) return fail(RT_B200_ERR_ARG, "null argument");
    out_ms[0] = c->ms_render;
    out_ms[1] = c->ms_gather;
    return RT_B200_OK;
}

int rt_b200_get_launches(rt_b200_ctx *c, uint32_t *launches)
{
    if (!c || !launches) return fail(RT_B200_ERR_ARG, "null argument");
    uint32_t n = (uint32_t) c->launches;
    if (c->world > 1 && c->comm && !c->shared_ready) n += 1;                 // strip un-interleave (the all-gather itself is NCCL's)
    if (c->world > 1 && c->shared_ready && c->flag_barrier && !c->sync_in_kernel) n += c->rank == 0 ? 3 : 2;   // sequence-word kernels (kernel 0 only: kernel 1 does the hand-shake itself)
    *launches = n;
    return RT_B200_OK;
}

int rt_b200_debug_timeline(rt_b200_ctx *c, uint64_t *out, size_t capacity_words, uint32_t *words_per_cta)
{
    if (!c || !out || !words_per_cta) return fail(RT_B200_ERR_ARG, "null argument");
    *words_per_cta = RT_TL_WORDS;
    const size_t n = std::min(capacity_words, (size_t) RT_TL_MAX_CTAS * RT_TL_WORDS);
    CUDA_TRY(cudaSetDevice(c->device));
    CUDA_TRY(cudaMemcpy(out, c->d_counters + rt::CNT_COUNT, sizeof(uint64_t) * n, cudaMemcpyDeviceToHost));
    return RT_B200_OK;
}

int rt_b200_get_class_counts(rt_b200_ctx *c, uint32_t out[4])
{
    if (!c || !out) return fail(RT_B200_ERR_ARG, "null argument");
    std::memcpy(out, c->class_counts, sizeof(c->class_counts));
    return RT_B200_OK;
}

int rt_b200_get_grid_info(rt_b200_ctx *c, uint32_t out[6])
{
    if (!c || !out) return fail(RT_B200_ERR_ARG, "null argument");
    std::memcpy(out, c->grid_info, sizeof(c->grid_info));
    return RT_B200_OK;
}

int rt_b200_band_layout(uint32_t rows, uint32_t n_bands, uint32_t unit, int taper, int32_t *starts)
{
    if (!starts || rows == 0 || unit == 0 || n_bands == 0 || n_bands > RT_MAX_BANDS || taper < 0 || taper > 199 || taper == 100)
        return fail(RT_B200_ERR_ARG, "bad band layout query");
    if (n_bands > (rows + unit - 1) / unit) return fail(RT_B200_ERR_ARG, "%u bands do not fit %u rows in units of %u", n_bands, rows, unit);
    int tmp[RT_MAX_BANDS + 1];
    band_layout((int) rows, (int) n_bands, (int) unit, taper, tmp);
    for (uint32_t b = 0; b <= n_bands; b++) starts[b] = tmp[b];
    return RT_B200_OK;
}

} // extern "C"

// =====================================================================================================================
// rt_b200_render_no_alloc — one frame without any allocation (include/rt_b200.h; no counterpart in the reference).
// Nothing above this line is changed by it.
//
// The wavefront kernel's device scratch (tile counters + the class-sorted scene arrays it reads) lives in the static
// array below instead of cudaMalloc'ed buffers.  In HOST code the array's name is only a host-side stand-in: its device
// address comes from cudaGetSymbolAddress, once per device, with that device current (no_alloc_device_scratch).
// =====================================================================================================================
namespace {

static __device__ __align__(256) unsigned char g_no_alloc_scratch[RT_B200_NO_ALLOC_SCRATCH_BYTES];

// Host image of the scratch, built here and copied to the device in one transfer (static: no host allocation either).
alignas(256) unsigned char g_no_alloc_host[RT_B200_NO_ALLOC_SCRATCH_BYTES];
std::mutex g_no_alloc_mutex;                    // one call at a time: the staging area and the device array are shared
constexpr int NO_ALLOC_MAX_DEVICES = 64;
unsigned char *g_no_alloc_dev_ptr[NO_ALLOC_MAX_DEVICES] = {};   // device address of g_no_alloc_scratch per device ordinal

// Offsets of the arrays inside the scratch (every array 256-byte aligned).
struct NoAllocLayout {
    size_t tile_counter, plane, sphere, sphere_f, quadric, cubic, cubic_mask, orig, coef_full, material, material_pi, lights;
    size_t total;
    size_t n_class[4];
};

size_t bump(size_t &at, size_t bytes)
{
    const size_t here = at;
    at = (at + (bytes ? bytes : 1) + 255) & ~(size_t) 255;
    return here;
}

NoAllocLayout no_alloc_layout(const void *objects, size_t n_objects, size_t n_lights)
{
    NoAllocLayout L{};
    const char *obj_bytes = static_cast<const char *>(objects);
    for (size_t i = 0; i < n_objects; i++) {
        double k[20];
        std::memcpy(k, obj_bytes + RT_B200_OBJECT_BYTES * i, 160);
        L.n_class[classify(k)]++;
    }
    size_t at = 0;
    L.tile_counter = bump(at, 2 * sizeof(unsigned long long));   // must start at zero: the host image carries the zeros
    L.plane = bump(at, sizeof(double4) * L.n_class[rt::CLS_PLANE]);
    L.sphere = bump(at, sizeof(double4) * L.n_class[rt::CLS_SPHERE]);
    L.sphere_f = bump(at, sizeof(float4) * L.n_class[rt::CLS_SPHERE]);
    L.quadric = bump(at, sizeof(double) * 10 * L.n_class[rt::CLS_QUADRIC]);
    L.cubic = bump(at, sizeof(double) * 20 * L.n_class[rt::CLS_CUBIC]);
    L.cubic_mask = bump(at, sizeof(unsigned) * L.n_class[rt::CLS_CUBIC]);
    L.orig = bump(at, sizeof(int) * n_objects);
    L.coef_full = bump(at, sizeof(double) * 20 * n_objects);
    L.material = bump(at, sizeof(float4) * n_objects);
    L.material_pi = bump(at, sizeof(float4) * n_objects);
    L.lights = bump(at, sizeof(rt::LightDev) * n_lights);
    L.total = at;
    return L;
}

// Device address of the static scratch on `device` (current).  cudaGetSymbolAddress once per device.
int no_alloc_device_scratch(int device, unsigned char **out)
{
    if (device >= 0 && device < NO_ALLOC_MAX_DEVICES && g_no_alloc_dev_ptr[device]) {
        *out = g_no_alloc_dev_ptr[device];
        return RT_B200_OK;
    }
    unsigned char *ptr = nullptr;
    const cudaError_t e = cudaGetSymbolAddress((void **) &ptr, g_no_alloc_scratch);
    if (e != cudaSuccess || ptr == nullptr)
        return fail(RT_B200_ERR_CUDA, "cudaGetSymbolAddress(no-alloc scratch) on device %d: %s", device, cudaGetErrorString(e));
    if (device >= 0 && device < NO_ALLOC_MAX_DEVICES) g_no_alloc_dev_ptr[device] = ptr;
    *out = ptr;
    return RT_B200_OK;
}

} // namespace

extern "C" {

int rt_b200_no_alloc_scratch_needed(const void *objects, size_t n_objects, size_t n_lights, size_t *bytes)
{
    if (!bytes || (n_objects && !objects)) return fail(RT_B200_ERR_ARG, "null argument");
    if (n_objects > (size_t) 1 << 24 || n_lights > (size_t) 1 << 20) return fail(RT_B200_ERR_ARG, "scene too large");
    *bytes = no_alloc_layout(objects, n_objects, n_lights).total;
    return RT_B200_OK;
}

int rt_b200_render_no_alloc(uint32_t px_width, uint32_t px_height, double vertical_fov_rad, const float bg_color[3],
                            uint32_t max_reflections, const void *objects, size_t n_objects, const void *lights,
                            size_t n_lights, const double camera_matrix[16], int device, void *device_rgba8,
                            size_t device_bytes, float *elapsed_ms)
{
    if (px_width == 0 || px_height == 0) return fail(RT_B200_ERR_ARG, "empty frame %ux%u", px_width, px_height);
    if (!bg_color || !camera_matrix || !device_rgba8) return fail(RT_B200_ERR_ARG, "null argument");
    if ((n_objects && !objects) || (n_lights && !lights)) return fail(RT_B200_ERR_ARG, "null scene array");
    if (n_objects > (size_t) 1 << 24 || n_lights > (size_t) 1 << 20) return fail(RT_B200_ERR_ARG, "scene too large");
    const size_t frame_bytes = sizeof(uchar4) * (size_t) px_width * px_height;
    if (device_bytes < frame_bytes)
        return fail(RT_B200_ERR_ARG, "device buffer holds %zu bytes, frame needs %zu", device_bytes, frame_bytes);
    if (reinterpret_cast<uintptr_t>(device_rgba8) & 15u) return fail(RT_B200_ERR_ARG, "device buffer must be 16-byte aligned");
    if ((unsigned long long) px_width * px_height >= (1ULL << 31)) return fail(RT_B200_ERR_ARG, "frame too large");

    const NoAllocLayout L = no_alloc_layout(objects, n_objects, n_lights);
    if (L.total > (size_t) RT_B200_NO_ALLOC_SCRATCH_BYTES)
        return fail(RT_B200_ERR_ARG, "scene (%zu objects, %zu lights) needs %zu bytes of device scratch; rt_b200_render_no_alloc "
                    "holds %d (use rt_b200_upload_scene / rt_b200_update for larger scenes)", n_objects, n_lights, L.total,
                    RT_B200_NO_ALLOC_SCRATCH_BYTES);

    int n_dev = 0;
    CUDA_TRY(cudaGetDeviceCount(&n_dev));
    if (device < 0 || device >= n_dev) return fail(RT_B200_ERR_CUDA, "no CUDA device %d (found %d)", device, n_dev);
    CUDA_TRY(cudaSetDevice(device));
    cudaPointerAttributes pa{};
    if (cudaPointerGetAttributes(&pa, device_rgba8) != cudaSuccess ||
        (pa.type != cudaMemoryTypeDevice && pa.type != cudaMemoryTypeManaged)) {
        (void) cudaGetLastError();
        return fail(RT_B200_ERR_ARG, "device_rgba8 is not a device pointer");
    }
    if (pa.type == cudaMemoryTypeDevice && pa.device != device)
        return fail(RT_B200_ERR_ARG, "device_rgba8 lives on device %d, the frame is rendered on device %d", pa.device, device);

    std::lock_guard<std::mutex> lock(g_no_alloc_mutex);
    unsigned char *d_scratch = nullptr;
    int rc = no_alloc_device_scratch(device, &d_scratch);
    if (rc != RT_B200_OK) return rc;

    // ---- classify and repack into the host image (same rules as rt_b200_upload_scene) ----------------------------
    unsigned char *H = g_no_alloc_host;
    std::memset(H, 0, L.total);                       // tile counters = 0; no stale bytes of the previous call
    double4 *plane = reinterpret_cast<double4 *>(H + L.plane), *sphere = reinterpret_cast<double4 *>(H + L.sphere);
    float4 *sphere_f = reinterpret_cast<float4 *>(H + L.sphere_f);
    double *quadric = reinterpret_cast<double *>(H + L.quadric), *cubic = reinterpret_cast<double *>(H + L.cubic);
    unsigned *cubic_mask = reinterpret_cast<unsigned *>(H + L.cubic_mask);
    int *orig = reinterpret_cast<int *>(H + L.orig);
    double *coef_full = reinterpret_cast<double *>(H + L.coef_full);
    float4 *material = reinterpret_cast<float4 *>(H + L.material), *material_pi = reinterpret_cast<float4 *>(H + L.material_pi);
    rt::LightDev *ldev = reinterpret_cast<rt::LightDev *>(H + L.lights);
    size_t filled[4] = {0, 0, 0, 0};
    size_t class_base[4] = {0, L.n_class[0], L.n_class[0] + L.n_class[1], L.n_class[0] + L.n_class[1] + L.n_class[2]};
    float center_max = 0.0f, r2_max = 0.0f;
    bool cull_safe = true;
    const char *obj_bytes = static_cast<const char *>(objects);
    for (size_t i = 0; i < n_objects; i++) {
        double k[20];
        float refl, color[3];
        std::memcpy(k, obj_bytes + RT_B200_OBJECT_BYTES * i, 160);
        std::memcpy(&refl, obj_bytes + RT_B200_OBJECT_BYTES * i + 160, 4);
        std::memcpy(color, obj_bytes + RT_B200_OBJECT_BYTES * i + 164, 12);
        std::memcpy(coef_full + 20 * i, k, 160);
        const float pi_f = 3.14159274101257324f;
        material[i] = make_float4(refl, color[0], color[1], color[2]);
        material_pi[i] = make_float4(refl, color[0] / pi_f, color[1] / pi_f, color[2] / pi_f);
        if (!(fabsf(color[0]) < 1e5f && fabsf(color[1]) < 1e5f && fabsf(color[2]) < 1e5f)) cull_safe = false;
        const rt::ObjClass cls = classify(k);
        const size_t j = filled[cls]++;
        orig[class_base[cls] + j] = (int) i;          // class-sorted slot -> original index (order kept inside a class)
        switch (cls) {
            case rt::CLS_PLANE:
                plane[j] = make_double4(k[rt::LX], k[rt::LY], k[rt::LZ], k[rt::CC]);
                break;
            case rt::CLS_SPHERE: {
                sphere[j] = make_double4(k[rt::LX], k[rt::LY], k[rt::LZ], k[rt::CC]);
                const double cx = -0.5 * k[rt::LX], cy = -0.5 * k[rt::LY], cz = -0.5 * k[rt::LZ];
                const double r2 = cx * cx + cy * cy + cz * cz - k[rt::CC];
                sphere_f[j] = make_float4((float) cx, (float) cy, (float) cz, (float) r2);
                r2_max = fmaxf(r2_max, nextafterf((float) fabs(r2), INFINITY));
                center_max = fmaxf(center_max, fmaxf(fabsf((float) cx), fmaxf(fabsf((float) cy), fabsf((float) cz))));
                break;
            }
            case rt::CLS_QUADRIC:
                for (int q = rt::X2; q <= rt::CC; q++) quadric[10 * j + (q - rt::X2)] = k[q];
                break;
            case rt::CLS_CUBIC: {
                unsigned mask = 0;
                for (int q = 0; q < 20; q++) {
                    cubic[20 * j + q] = k[q];
                    if (k[q] != 0.0) mask |= 1u << q;
                }
                cubic_mask[j] = mask;
                break;
            }
        }
    }
    const char *light_bytes = static_cast<const char *>(lights);
    for (size_t j = 0; j < n_lights; j++) {
        const char *rec = light_bytes + RT_B200_LIGHT_BYTES * j;
        double p[3];
        float col[3];
        std::memcpy(p, rec + 8, 24);
        std::memcpy(col, rec + 32, 12);
        ldev[j] = rt::LightDev{p[0], p[1], p[2], col[0], col[1], col[2], rec[0] != 0 ? 1 : 0};
        if (!(fabsf(col[0]) < 1e15f && fabsf(col[1]) < 1e15f && fabsf(col[2]) < 1e15f)) cull_safe = false;
    }

    // ---- scene view and frame parameters over the DEVICE copy of the image -----------------------------------------
    rt::SceneView v{};                                 // grid.cell_start == nullptr: brute force
    v.n_plane = (int) L.n_class[rt::CLS_PLANE];
    v.n_sphere = (int) L.n_class[rt::CLS_SPHERE];
    v.n_quadric = (int) L.n_class[rt::CLS_QUADRIC];
    v.n_cubic = (int) L.n_class[rt::CLS_CUBIC];
    v.n_objects = (int) n_objects;
    v.n_lights = (int) n_lights;
    v.plane = reinterpret_cast<const double4 *>(d_scratch + L.plane);
    v.sphere = reinterpret_cast<const double4 *>(d_scratch + L.sphere);
    v.sphere_f = reinterpret_cast<const float4 *>(d_scratch + L.sphere_f);
    v.quadric = reinterpret_cast<const double *>(d_scratch + L.quadric);
    v.cubic = reinterpret_cast<const double *>(d_scratch + L.cubic);
    v.orig_index = reinterpret_cast<const int *>(d_scratch + L.orig);
    v.coef_full = reinterpret_cast<const double *>(d_scratch + L.coef_full);
    v.material = reinterpret_cast<const float4 *>(d_scratch + L.material);
    v.material_pi = reinterpret_cast<const float4 *>(d_scratch + L.material_pi);
    v.lights = reinterpret_cast<const rt::LightDev *>(d_scratch + L.lights);
    v.sphere_center_max = center_max;

    rt::FrameParams f;
    std::memcpy(f.m, camera_matrix, sizeof(f.m));
    for (int r = 0; r < 3; r++)   // origin = M * (0,0,0,1) in GLM's association, update-cpu.cpp:123
        f.origin[r] = (f.m[r] * 0.0 + f.m[4 + r] * 0.0) + (f.m[8 + r] * 0.0 + f.m[12 + r] * 1.0);
    f.aspect = (double) px_width / px_height;
    f.tan_half_fov = std::tan(0.5 * vertical_fov_rad);
    std::memcpy(f.bg, bg_color, sizeof(f.bg));
    f.max_reflections = (int) max_reflections;
    f.width = (int) px_width;
    f.height = (int) px_height;
    f.strip_rows = (int) px_height;
    f.rank = 0;
    f.world = 1;
    f.local_rows = (int) px_height;
    f.row0 = 0;
    f.rows = (int) px_height;
    f.direct = 0;
    f.out = static_cast<uchar4 *>(device_rgba8);
    f.counters = nullptr;
    f.timeline = nullptr;

    cudaStream_t stream = nullptr;   // legacy default stream: nothing to create
    CUDA_TRY(cudaMemcpyAsync(d_scratch, H, L.total, cudaMemcpyHostToDevice, stream));
    CUDA_TRY(cudaStreamSynchronize(stream));          // the staging area is pageable: done with it before the clock starts
    const double t0 = now_seconds();
    if (rt::wave_scene_fits(v)) {
        const rt::WaveLaunch wl{reinterpret_cast<unsigned long long *>(d_scratch + L.tile_counter),
                                reinterpret_cast<const unsigned *>(d_scratch + L.cubic_mask), r2_max, -1, 0, 1, cull_safe ? 1 : 0, 0,
                                nullptr, nullptr, nullptr, 0u, 0, 1, nullptr, nullptr, 0, 0, 0, 0};
        const cudaError_t e = rt::launch_render_wave(v, f, wl, stream);
        if (e != cudaSuccess) return fail(RT_B200_ERR_CUDA, "launch_render_wave: %s", cudaGetErrorString(e));
    }
    else {
        CUDA_TRY(rt::launch_render_exact(v, f, stream));
    }
    CUDA_TRY(cudaStreamSynchronize(stream));
    if (elapsed_ms) *elapsed_ms = (float) ((now_seconds() - t0) * 1e3);
    return RT_B200_OK;
}

} // extern "C"
