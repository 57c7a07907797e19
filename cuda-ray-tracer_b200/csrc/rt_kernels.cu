// rt_kernels.cu — sm_100a kernels of the per-pixel frame update.
//
// Replaces the reference's single update_kernel (src/update-cuda.cu:104-158) — not a port of it: the scene is
// class-sorted SoA (rt_scene.h), the arithmetic is the reference CPU path's exact FP64 order (rt_math.cuh),
// the frame is a plain RGBA8 device buffer instead of a GL surface, and rows are strip-interleaved over ranks.
#include "rt_kernels.h"
#include "rt_trace.cuh"

namespace rt {

// ---------------------------------------------------------------------------------------------------
// Exact kernel: 256 threads = a 32x8 pixel tile, each warp an 8x4 sub-tile (coherent rays, and each
// warp writes four fully used 32-byte sectors).
// ---------------------------------------------------------------------------------------------------
constexpr int TILE_W = 32;
constexpr int TILE_H = 8;

__global__ void __launch_bounds__(TILE_W *TILE_H)
render_exact_kernel(const __grid_constant__ SceneView sc, const __grid_constant__ FrameParams f)
{
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int x = blockIdx.x * TILE_W + (warp & 3) * 8 + (lane & 7);
    const int lrow = f.row0 + blockIdx.y * TILE_H + (warp >> 2) * 4 + (lane >> 3);
    const int y = local_to_image_row(lrow, f.strip_rows, f.rank, f.world);
    const bool in_band = x < f.width && lrow < f.row0 + f.rows && lrow < f.local_rows;
    const bool live = in_band && y < f.height;

    Tally tl;
    tl.clear();
    if (live) {
        const d3 origin = d3{f.origin[0], f.origin[1], f.origin[2]};
        d3 dir = primary_dir(f, origin, x, y);
        tl.n[CNT_PRIMARY]++;

        float3 hit_color;
        d3 p, n;
        int idx = trace_and_shade_exact(sc, origin, dir, hit_color, p, n, tl);
        float3 out;
        if (idx < 0) {
            out = make_float3(f.bg[0], f.bg[1], f.bg[2]);
        }
        else {
            // reflection loop, update-cpu.cpp:97-117; blending uses the cumulative ratio
            out = hit_color;
            float ratio = 1.0f;
            int bounces = 0;
            float refl = __ldg(&sc.material[idx]).x;
            while (refl > RT_EPS) {
                ratio *= refl;
                float3 blend = make_float3(f.bg[0], f.bg[1], f.bg[2]);
                bool stop = true;
                if (bounces != f.max_reflections) {
                    bounces++;
                    d3 rd = reflect(dir, n);
                    d3 ro = d3{p.x + RT_SHADOW_BIAS * n.x, p.y + RT_SHADOW_BIAS * n.y, p.z + RT_SHADOW_BIAS * n.z};
                    dir = rd;
                    tl.n[CNT_REFLECTION]++;
                    idx = trace_and_shade_exact(sc, ro, dir, hit_color, p, n, tl);
                    if (idx >= 0) {
                        blend = hit_color;
                        stop = false;
                        refl = __ldg(&sc.material[idx]).x;
                    }
                }
                out.x = (1.0f - ratio) * out.x + ratio * blend.x;
                out.y = (1.0f - ratio) * out.y + ratio * blend.y;
                out.z = (1.0f - ratio) * out.z + ratio * blend.z;
                if (stop) break;
            }
        }
        f.out[out_index(f, lrow, y, x)] = pack_rgba8(out.x, out.y, out.z);
    }
    else if (in_band && !f.direct) {
        f.out[(size_t) lrow * f.width + x] = make_uchar4(0, 0, 0, 0);   // padding row past the image
    }
    flush_tally(tl, f.counters);
}

cudaError_t launch_render_exact(const SceneView &scene, const FrameParams &frame, cudaStream_t stream)
{
    dim3 block(TILE_W * TILE_H);
    dim3 grid((frame.width + TILE_W - 1) / TILE_W, (frame.rows + TILE_H - 1) / TILE_H);
    render_exact_kernel<<<grid, block, 0, stream>>>(scene, frame);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------------
// Strip un-interleave after the all-gather.  16 bytes (4 pixels) per thread when the width allows.
// ---------------------------------------------------------------------------------------------------
__global__ void unpack_strips_kernel(const uchar4 *__restrict__ gathered, uchar4 *__restrict__ frame, int width,
                                     int height, int strip_rows, int world, int local_rows)
{
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = blockIdx.y;
    if (x >= width || y >= height) return;
    const int strip = y / strip_rows;
    const int owner = strip % world;
    const int lrow = (strip / world) * strip_rows + (y - strip * strip_rows);
    frame[(size_t) y * width + x] = gathered[((size_t) owner * local_rows + lrow) * width + x];
}

cudaError_t launch_unpack_strips(const uchar4 *gathered, uchar4 *frame, int width, int height, int strip_rows,
                                 int world, int local_rows, cudaStream_t stream)
{
    dim3 block(256);
    dim3 grid((width + 255) / 256, height);
    unpack_strips_kernel<<<grid, block, 0, stream>>>(gathered, frame, width, height, strip_rows, world, local_rows);
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------------
// Frame-complete signalling for the fused gather (one process per GPU, flags in rank 0's memory, mapped by
// every rank through CUDA IPC).  A kernel boundary orders a rank's pixel stores before its flag store.
// ---------------------------------------------------------------------------------------------------
__global__ void sync_store_kernel(unsigned *word, unsigned value)
{
    __threadfence_system();
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(word), "r"(value) : "memory");
}

// Spin until every word[i], i in [first, first + n), has reached `need` (sequence numbers only grow); gives up
// after ~20 s of GPU clock and reports through *timed_out, so a dead peer cannot hang the box.
__global__ void sync_wait_kernel(const unsigned *words, int first, int n, unsigned need, unsigned *timed_out)
{
    const long long t0 = clock64();
    for (int i = first + threadIdx.x; i < first + n; i += blockDim.x) {
        unsigned v;
        for (;;) {
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(words + i) : "memory");
            if ((int) (v - need) >= 0) break;
            if (clock64() - t0 > 40000000000LL) {
                *timed_out = 1u;
                break;
            }
            __nanosleep(200);
        }
    }
}

__global__ void dyn_table_init_kernel(int *slot_group, int max_slots, int ahead, int n_groups, int rank, int world)
{
    for (int k = blockIdx.x * blockDim.x + threadIdx.x; k < max_slots; k += gridDim.x * blockDim.x) {
        int g = -1;
        if (k < ahead) {
            g = k * world + rank;
            if (g >= n_groups) g = -2;
        }
        slot_group[k] = g;
    }
}

cudaError_t launch_dyn_table_init(int *slot_group, int max_slots, int ahead, int n_groups, int rank, int world, cudaStream_t stream)
{
    dyn_table_init_kernel<<<(max_slots + 255) / 256, 256, 0, stream>>>(slot_group, max_slots, ahead, n_groups, rank, world);
    return cudaGetLastError();
}

cudaError_t launch_sync_store(unsigned *word, unsigned value, cudaStream_t stream)
{
    sync_store_kernel<<<1, 1, 0, stream>>>(word, value);
    return cudaGetLastError();
}

cudaError_t launch_sync_wait(const unsigned *words, int first, int n, unsigned need, unsigned *timed_out, cudaStream_t stream)
{
    if (n <= 0) return cudaSuccess;
    sync_wait_kernel<<<1, 32, 0, stream>>>(words, first, n, need, timed_out);
    return cudaGetLastError();
}

} // namespace rt
