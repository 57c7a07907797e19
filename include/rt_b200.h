/* rt_b200.h — C ABI of the B200-native frame-update library (librt_b200.so).
 *
 * This is the drop-in boundary for the reference's back-end call surface,
 * /root/reference/include/update.h:6-8:
 *
 *     void  init_update(unsigned int texture, const Scene& scene);     -> rt_b200_upload_scene
 *     float update(const glm::dmat4& camera_matrix);                   -> rt_b200_update
 *     void  cleanup_update();                                          -> rt_b200_destroy
 *
 * plus rt_b200_gather, which replaces the GL texture the reference writes into
 * (src/update-cuda.cu:59,166-174) as the result sink: there is no graphics stack on the target box, so
 * the frame lives in a plain device buffer and is copied to the caller on request.
 *
 * Conventions: plain C types only; every call returns an int status (RT_B200_OK == 0) instead of the
 * reference's print-and-exit(1) (include/helper_cuda_opengl.h:13-44); rt_b200_last_error() describes
 * the last failure on the calling thread; state lives in an opaque context instead of file-scope
 * globals (src/update-cuda.cu:15-27); one caller thread per context; every call is synchronous.
 * There is no CPU fallback: without a CUDA device every compute entry point fails with
 * RT_B200_ERR_CUDA.
 */
#ifndef RT_B200_H
#define RT_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RT_B200_OK 0
#define RT_B200_ERR_ARG 1      /* bad argument (null pointer, zero size, unknown option, small buffer) */
#define RT_B200_ERR_CUDA 2     /* a CUDA runtime call failed (including: no device)                  */
#define RT_B200_ERR_NCCL 3     /* NCCL missing or an NCCL call failed                                */
#define RT_B200_ERR_STATE 4    /* call out of order (e.g. gather before the first update)            */

#define RT_B200_OBJECT_BYTES 176 /* sizeof(Object),      include/scene.h:8-15  */
#define RT_B200_LIGHT_BYTES 48   /* sizeof(LightSource), include/light.h:6-13  */
#define RT_B200_UNIQUE_ID_BYTES 128
#define RT_B200_IPC_HANDLE_BYTES 64

typedef struct rt_b200_ctx rt_b200_ctx;

/* Work actually executed by the last rt_b200_update on THIS rank (64-bit, exact); with option cull_backfacing = 0
 * this is the reference's own ray census for the frame.
 * rays = primary + reflection + shadow, the reference's ray definition. */
typedef struct {
    uint64_t primary_rays;    /* one per pixel rendered by this rank                                   */
    uint64_t reflection_rays; /* bounce rays cast (update-cpu.cpp:104-111)                             */
    uint64_t shadow_rays;     /* one per light per hit (update-cpu.cpp:62-64)                          */
    uint64_t hits;            /* nearest-hit searches that found an object                             */
    uint64_t lit;             /* unshadowed light evaluations (surface_color calls)                    */
    uint64_t tests_plane;     /* (ray, object) tests this kernel performed, by structural class ...    */
    uint64_t tests_sphere;
    uint64_t tests_quadric;
    uint64_t tests_cubic;
    uint64_t fp64_confirms;   /* sphere tests that fell through the FP32 filter to the exact FP64 test */
    uint64_t cubic_exact;     /* cubic-class tests that needed the exact FP64 expansion + solve (the others were
                               * decided by the FP32 filter); kernel 1 only, kernel 0 tests exactly and leaves 0 */
} rt_b200_counters;

/* == init_update(texture, scene), update.h:6.
 * Copies everything it needs; the caller keeps ownership of all inputs.
 *   px_width, px_height, vertical_fov_rad, bg_color, max_reflections : Scene scalars, include/scene.h:19-22
 *       (vertical_fov in RADIANS, as Scene stores it after scene.cpp:20)
 *   objects : n_objects records in the reference's Object layout (176 B: 20 doubles in SurfaceCoefs order
 *             x3,y3,z3,x2y,xy2,x2z,xz2,y2z,yz2,xyz,x2,y2,z2,xy,xz,yz,x,y,z,c; float reflection_ratio @160;
 *             float color[3] @164) — i.e. &scene.objects[0]
 *   lights  : n_lights records in the reference's LightSource layout (48 B: bool is_spherical @0,
 *             double p[3] @8, float light_color[3] @32) — i.e. &scene.lights[0]
 *   device  : CUDA device ordinal this context renders on
 *   rank, world_size : this context renders the image strips s with s % world_size == rank
 *             (1-GPU use: rank 0, world_size 1).  For world_size > 1 call rt_b200_comm_init next. */
int rt_b200_upload_scene(rt_b200_ctx **ctx_out, uint32_t px_width, uint32_t px_height, double vertical_fov_rad,
                         const float bg_color[3], uint32_t max_reflections, const void *objects, size_t n_objects,
                         const void *lights, size_t n_lights, int device, int rank, int world_size);

/* == update(camera_matrix), update.h:7.
 * camera_matrix: 16 doubles, GLM column-major (&m[0][0]).  Renders this rank's strips (world_size 1: the frame) and
 * assembles the frame inside the call, in one of two ways:
 *   - a shared frame is attached (rt_b200_frame_attach / _attach_peer, the fused gather below): the render kernel itself
 *     stores every pixel at its place in the frame on rank 0's GPU; rank 0's call returns when the frame is complete;
 *   - only a communicator is attached (rt_b200_comm_init): the packed strips are all-gathered over NCCL and un-interleaved,
 *     so that every rank holds the full frame.
 * With world_size > 1 and neither, only this rank's rows are rendered (rt_b200_gather exposes them in place).
 * *elapsed_ms (may be NULL) = CUDA-event time from the first launch to the end of the assembly, the value
 * the reference's update() returns (src/update-cuda.cu:178-189). */
int rt_b200_update(rt_b200_ctx *ctx, const double camera_matrix[16], float *elapsed_ms);

/* Result sink (replaces the GL texture).  Frame = px_width*px_height RGBA8 pixels, row 0 = bottom row of the
 * picture (as in the reference: ray-tracer.cpp:12-17, update-cpu.cpp:85-87), alpha = 255.
 *   host_rgba8 != NULL : copy the frame into it (host_bytes must be >= px_width*px_height*4)
 *   device_rgba8 != NULL : receives the device pointer of the assembled frame on this rank's GPU
 * With world_size > 1 and no communicator attached only this rank's rows are valid. */
int rt_b200_gather(rt_b200_ctx *ctx, void *host_rgba8, size_t host_bytes, const void **device_rgba8);

/* update() and the result sink in one call, for hosts that want the frame in host memory every frame (what the
 * reference's CPU back-end hands to glTexImage2D, update-cpu.cpp:136-137).  On one GPU the frame is rendered in row
 * bands and each finished band is copied to host_rgba8 on a second stream while the next band renders, so the
 * device->host transfer overlaps the rendering; *elapsed_ms then covers launch -> last byte on the host.
 * host_rgba8 should be page-locked for the overlap.  With world_size > 1 there are two forms:
 *   - host_rgba8 from rt_b200_host_alloc_shared (the same block mapped by every rank): every rank renders its strips
 *     and its own copy engine writes them to their image rows of the one host frame, so G PCIe links fill it in
 *     parallel and nothing crosses NVLink; all ranks call it for every frame, the frame is complete when RANK 0's
 *     call returns (the other ranks return when their own strips are out), and a rank does not start overwriting
 *     the block for frame k+1 before rank 0 has entered its call for frame k+1;
 *   - any other buffer: rt_b200_update followed by rt_b200_gather. */
int rt_b200_update_to_host(rt_b200_ctx *ctx, const double camera_matrix[16], void *host_rgba8, size_t host_bytes,
                           float *elapsed_ms);

/* Page-locked host memory for frames (so that the streamed copy above really overlaps); plain C, no CUDA headers
 * needed by the host.  rt_b200_host_free(NULL) is a no-op. */
int rt_b200_host_alloc(void **ptr_out, size_t bytes);
int rt_b200_host_free(void *ptr);

/* A host frame shared by the processes of a multi-GPU job: a MAP_SHARED mapping of the file `path` (normally under
 * /dev/shm), page-locked in the calling process.  Exactly one rank passes create != 0 (the file must not exist; it
 * is sized to `bytes` + one page of synchronisation words and unlinked again when that rank frees the block); the
 * others open it afterwards with create == 0 and the same `bytes`.  New with the multi-GPU path — the reference is
 * one process, one GPU. */
int rt_b200_host_alloc_shared(const char *path, size_t bytes, int create, void **ptr_out);
int rt_b200_host_free_shared(void *ptr);

/* == cleanup_update(), update.h:8. */
int rt_b200_destroy(rt_b200_ctx *ctx);

const char *rt_b200_last_error(void);

/* ---- multi-GPU: one process (or thread) per GPU, NCCL over NVLink --------------------------------
 * Rank 0 obtains an id, the host distributes the 128 bytes to all ranks by any means (the Python host
 * uses torch.distributed), then every rank attaches its context.  libnccl.so.2 is loaded on first use. */
int rt_b200_comm_unique_id(void *id_out /* RT_B200_UNIQUE_ID_BYTES */);
int rt_b200_comm_init(rt_b200_ctx *ctx, const void *id /* RT_B200_UNIQUE_ID_BYTES */);

/* Fused gather (optional): instead of all-gathering packed strips, every rank's render kernel stores its pixels straight
 * into ONE assembled frame that lives on rank 0 (two of them, used alternately) — a CUDA IPC mapping over NVLink for
 * one process per GPU, plain peer access for contexts of one process — so the transfer overlaps the tracing tile by
 * tile.  Rank 0 exports, the host hands the 64 bytes to every rank, and every rank (rank 0 included) attaches after
 * rt_b200_comm_init; contexts of ONE process attach with rt_b200_frame_attach_peer instead (no IPC handle, no NCCL).
 * Frame-complete is signalled through sequence words behind the frames in rank 0's memory (option "flag_barrier" 1, the
 * default): every rank publishes flags[rank] = frame number after a system-scope fence, and RANK 0 — the consumer —
 * waits for all of them inside its own rt_b200_update.  Guarantees:
 *   - when rank 0's rt_b200_update returns, the frame is complete; rt_b200_gather on rank 0 hands it out;
 *   - a rank > 0 returns when its OWN pixels are stored; rt_b200_gather there returns a peer pointer to rank 0's
 *     frame, which other ranks may still be writing — a host that wants to read it on rank r > 0 adds its own
 *     barrier after rank 0 has returned (tools/multi_gpu_check.py does);
 *   - no rank overwrites a frame rank 0 may still be reading: rank r starts frame k only after rank 0 has entered
 *     frame k - 1 (two buffers).
 * With "flag_barrier" 0 a 4-byte NCCL all-reduce is the barrier instead (every rank waits for every rank; needs the
 * communicator).  All ranks must call rt_b200_update for every frame, each from its own host thread or process: rank
 * 0's call does not return before the others have rendered. */
int rt_b200_frame_export(rt_b200_ctx *ctx, void *handle_out /* RT_B200_IPC_HANDLE_BYTES */);
int rt_b200_frame_attach(rt_b200_ctx *ctx, const void *handle /* RT_B200_IPC_HANDLE_BYTES */);
/* Same-process form: `owner` is rank 0's context (after rt_b200_frame_export) living in THIS process, possibly on
 * another device; rank 0 passes its own context.  New with the multi-GPU path (the reference is one GPU). */
int rt_b200_frame_attach_peer(rt_b200_ctx *ctx, rt_b200_ctx *owner);

/* ---- introspection and tuning -------------------------------------------------------------------- */
int rt_b200_get_counters(rt_b200_ctx *ctx, rt_b200_counters *out);

/* Options: "kernel"     0 = exact FP64 kernel (thread per pixel; the on-device cross-check), 1 = wavefront kernel (default)
 *          "host_bands" row bands of rt_b200_update_to_host (0 = automatic, about 4 MB each)
 *          "strip_rows" image rows per strip of the multi-GPU interleave (default 8)
 *          "count"      1 = maintain rt_b200_counters (default), 0 = skip them
 *          "cubic_filter" 1 = FP32 filter in front of the exact cubic test in kernel 1 (default), 0 = off
 *          "cull_backfacing" 1 (default, kernel 1): no shadow ray is traced towards a light whose Lambert factor
 *                       max(0, n.dir) is 0 at the hit — such a light adds exactly 0 whether occluded or not, so
 *                       frames are identical; rt_b200_counters then report the rays actually traced.  0 = trace every
 *                       shadow ray the reference casts (the counters equal the reference's ray census)
 *           "accel"     sphere-class acceleration (wavefront kernel): 1 = rays look only at the
 *                       spheres registered in the uniform-grid cells they pass (csrc/rt_grid.cuh), each candidate still
 *                       going through the exact test, so frames are identical to the brute-force loops of
 *                       update-cpu.cpp:48-56,66-72; 0 = brute force (what the roofline figures are quoted on);
 *                       -1 (default) = on for scenes with at least 64 sphere-class objects.  With the grid on, the
 *                       tests_sphere / confirms counters report the candidates actually examined
 *           "band_taper"  rt_b200_update_to_host: 0 = equal row bands; r in 1..99 = band k holds (r/100)^k of the rows
 *                       of band 0, so the device-to-host copy that remains after the last band has rendered is small
 *                       (pays when rendering is slower than the copy); 100 + r = the same sizes in reverse order, a tiny
 *                       band first (pays when the copy is slower: it starts earliest); -1 (default) = try 0, 65 and 165
 *                       on a scene's first streamed frames, keep the fastest, re-try another every 32nd frame
 *           "band_overlap"  rt_b200_update_to_host, wavefront kernel: 1 (default) = consecutive row bands alternate between
 *                       two streams, each with its own tile counter, so the tail of one band overlaps the start of the next
 *           "grid_density"  rebuild the grid at value / 100 cells per sphere (0 = default 1.5)
 *           "dynamic_tiles" fused gather, wavefront kernel: 1 = the image is cut into groups of one
 *                       tile height of rows that a counter in rank 0's memory hands out to the GPUs on demand (each GPU
 *                       fetches two groups ahead over NVLink), so a GPU that finishes early takes more; 0 (default) =
 *                       every rank renders its own interleaved strips.  Frames are identical; on the bundled scenes the
 *                       static interleave is already balanced and measured 2-5 % faster (profiles/r2_scaling_8gpu.json)
 *           "variant"   wavefront kernel shape: 0 = automatic, 2 / 3 = 256-thread CTAs, two / three per SM, 5 = 512 x 1 */
int rt_b200_set_option(rt_b200_ctx *ctx, const char *name, int64_t value);

/* Split timing of the last update, milliseconds: [0] render kernels, [1] gather (all-gather + unpack). */
int rt_b200_get_timing(rt_b200_ctx *ctx, float out_ms[2]);

/* Kernels of THIS library launched by the last rt_b200_update / rt_b200_update_to_host on this rank: render launches
 * (one per row band), the strip un-interleave after an NCCL all-gather, and — kernel 0 with the fused gather only — the
 * sequence-word kernels (kernel 1 does that hand-shake inside the render kernel). */
int rt_b200_get_launches(rt_b200_ctx *ctx, uint32_t *launches);

/* Diagnostic (meaningful only with the timeline build of the library, `build.py --timeline`, and option "timeline" = 1; the
 * product build leaves zeros): per CTA of the last wavefront-kernel launch *words_per_cta 64-bit words — global-timer ns at kernel
 * entry, after the prologue, when the CTA first found no tile left, at loop exit, at kernel exit; waves run; SM id; tiles rendered.
 * tools/gpu_timeline.py prints where the time of a launch goes. */
int rt_b200_debug_timeline(rt_b200_ctx *ctx, uint64_t *out, size_t capacity_words, uint32_t *words_per_cta);

/* Structural class counts of the uploaded scene: planes, spheres, quadrics, cubics. */
int rt_b200_get_class_counts(rt_b200_ctx *ctx, uint32_t out[4]);

/* Host-only helper (needs no GPU): the row bands rt_b200_update_to_host cuts a frame of `rows` rows into — n_bands
 * (1..16) bands, boundaries at multiples of `unit` rows, shape as option "band_taper" (0 equal, 1..99 shrinking,
 * 101..199 growing).  starts receives n_bands + 1 row indices, starts[0] = 0, starts[n_bands] = rows. */
int rt_b200_band_layout(uint32_t rows, uint32_t n_bands, uint32_t unit, int taper, int32_t *starts);

/* The sphere grid built at upload (option "accel"): cells along x, y, z; spheres registered in cells; spheres kept on
 * the always-tested list (too large, too small or non-finite for the grid); cell registrations in total. */
int rt_b200_get_grid_info(rt_b200_ctx *ctx, uint32_t out[6]);

/* Host-only helper (needs no GPU): the strip interleave of the multi-GPU split.  Rank r of world_size owns
 * the strips s (of strip_rows image rows each) with s % world_size == r and stores them packed, padded to
 * ceil(n_strips / world_size) strips so that every rank contributes the same byte count to the all-gather.
 * *local_rows receives the packed row count; image_rows (may be NULL; capacity entries) receives, for each
 * packed row, the image row it holds or -1 for a padding row.  world_size == 1: the identity. */
int rt_b200_strip_layout(uint32_t px_height, uint32_t strip_rows, int rank, int world_size, int32_t *image_rows,
                         size_t capacity, uint32_t *local_rows);

/* ---- one frame without any allocation -------------------------------------------------------------------------------
 * No counterpart in the reference; added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 *
 * Renders ONE frame of a scene given in the reference's layouts (the same arguments as rt_b200_upload_scene, plus the
 * camera matrix of rt_b200_update) into a DEVICE buffer the caller owns — no context, no rt_b200_upload_scene.  The
 * wavefront kernel's device scratch — its tile counters and the class-sorted copy of the scene it reads, which
 * rt_b200_upload_scene would cudaMalloc — lives in one file-scope `static __device__ __align__(256)` array of
 * RT_B200_NO_ALLOC_SCRATCH_BYTES (1 MiB) inside the library; every GPU that loads the library holds that array until the
 * process exits.  The call makes no cudaMalloc / cudaFree / cudaHostAlloc, creates no stream or event, and calls no
 * allocating library routine: one host-to-device copy of the repacked scene (from a static host staging area), one kernel
 * launch on the legacy default stream, one stream synchronise.
 *   device_rgba8 : device pointer on `device`, 16-byte aligned, device_bytes >= px_width * px_height * 4; receives the
 *                  frame exactly as rt_b200_update + rt_b200_gather deliver it (RGBA8, row 0 = bottom, alpha 255)
 *   *elapsed_ms  : (may be NULL) host clock around launch -> kernel complete (no events are created)
 * Limits (RT_B200_ERR_ARG with a message naming the sizes): the repacked scene must fit the scratch array — about 250
 * bytes per sphere or plane, 280 per quadric, 360 per cubic, 40 per light: the synthetic 4096 + 64 + 16 scene of
 * BASELINE.json needs 1 024 000 bytes and fits, 6 000 spheres do not; rt_b200_no_alloc_scratch_needed tells exactly.  Scenes
 * whose lights / quadrics / cubics exceed the kernel's shared memory render through the exact per-pixel kernel, as in
 * rt_b200_update.  The sphere grid (option "accel") is not available here: brute force, like the reference.
 * Sharing: every call uses the same array (and the same host staging area), which keeps the previous call's contents;
 * each call rewrites everything it reads, including the tile counters that must start at zero.  Calls are serialised by
 * a mutex and are synchronous, so calls through this entry point never overlap on the GPU; anything ELSE that touched the
 * array concurrently would race — there is nothing else: the array is private to this entry point. */
#define RT_B200_NO_ALLOC_SCRATCH_BYTES 1048576
int rt_b200_render_no_alloc(uint32_t px_width, uint32_t px_height, double vertical_fov_rad, const float bg_color[3],
                            uint32_t max_reflections, const void *objects, size_t n_objects, const void *lights,
                            size_t n_lights, const double camera_matrix[16], int device, void *device_rgba8,
                            size_t device_bytes, float *elapsed_ms);
/* Host-only helper (needs no GPU): scratch bytes rt_b200_render_no_alloc needs for a scene with these objects / lights. */
int rt_b200_no_alloc_scratch_needed(const void *objects, size_t n_objects, size_t n_lights, size_t *bytes);

/* Library build tag, e.g. "rt_b200 0.1 sm_100a". */
const char *rt_b200_version(void);

#ifdef __cplusplus
}
#endif
#endif
