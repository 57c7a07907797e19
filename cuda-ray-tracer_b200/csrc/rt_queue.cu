// rt_queue.cu — queue kernel (option kernel=2): block-local wavefront over PATH RECORDS.
//
// One persistent CTA per SM slot.  A pixel whose primary ray hits something gets a path record in shared
// memory; the record moves between three stacks of record indices:
//     free  ->  (hit found)  ->  shade  ->  (surface reflects)  ->  trace  ->  (bounce ray hits)  ->  shade ...
// and every wave takes W records from ONE stack (or W pixels of the next image tile), so that all W threads
// always run the same phase:
//   TILE   primary rays of the next 32 x (W/32) tile; misses are final, hits allocate a record
//   POOL   reflection rays of W waiting records (update-cpu.cpp:104-111)
//   SHADE  W hit records: one thread per hit loops over the lights — shadow ray (any-hit), Lambert term in
//          light order (update-cpu.cpp:62-77), blend, and either finish the pixel or queue the bounce
// A full wave is taken as soon as a stack holds W records, tiles only when W records are free; with
// cap >= 3 W this never forces a partial wave before the image runs out.  Terminated and non-reflective
// paths therefore never idle lanes in later bounces, and shading waves stay full even where only a few
// pixels per tile hit anything.
//
// All three phases run their rays through ONE inlined query (single instantiation of every intersection
// routine — the kernel stays small enough for the instruction cache):
//   spheres   FP32 rejection tables specialised per ray family (rt_tables.cuh) + exact FP64 confirmation of
//             the few candidates, batched per lane so that the exact tests of a warp run together
//   planes, quadrics, cubics   exact FP64 in the reference's operation order (rt_math.cuh, rt_cubic.cuh).  Their
//             coefficients, the cubic zero-coefficient masks and the lights are KERNEL PARAMETERS (QueueScene):
//             constant-bank operands indexed by a warp-uniform loop counter, so a coefficient costs neither a
//             load nor a register and the mask tests are uniform branches
// Any-hit queries leave the loops through warp votes, never through per-lane branches.
//
// Reference: src/update-cpu.cpp:45-139 (what is computed); src/update-cuda.cu:104-158 is the kernel this
// replaces (one thread per pixel, AoS scene in global memory) — not a port of it.
#include <cstring>

#include "rt_kernels.h"
#include "rt_queue.cuh"

namespace rt {

namespace {

constexpr unsigned FULL = 0xffffffffu;
enum Mode { M_EXIT = 0, M_TILE = 1, M_POOL = 2, M_SHADE = 3 };

// Finished pixels leave the SM.  Lanes 4i..4i+3 that hold four horizontally adjacent, 16-byte aligned pixels
// store them as one 128-bit uchar4x4; everything else is stored as single pixels.
__device__ __forceinline__ void store_pixels(uchar4 *out, bool final_px, int index, uchar4 px)
{
    const unsigned lane = threadIdx.x & 31;
    const int lead = __shfl_sync(FULL, index, lane & ~3u);
    const bool fits = final_px && index == lead + (int) (lane & 3u) && (lead & 3) == 0;
    const unsigned ok = __ballot_sync(FULL, fits);
    const unsigned group = (ok >> (lane & ~3u)) & 0xfu;
    const unsigned v = *reinterpret_cast<unsigned *>(&px);
    const unsigned v1 = __shfl_down_sync(FULL, v, 1);
    const unsigned v2 = __shfl_down_sync(FULL, v, 2);
    const unsigned v3 = __shfl_down_sync(FULL, v, 3);
    if (group == 0xfu) {
        if ((lane & 3u) == 0) *reinterpret_cast<uint4 *>(out + index) = make_uint4(v, v1, v2, v3);
    }
    else if (final_px) {
        out[index] = px;
    }
}

__device__ __forceinline__ float blend(float out, float ratio, float c) { return (1.0f - ratio) * out + ratio * c; }

// Push `idx` on a stack for every lane with `want`: warp ballot + prefix popc, one shared atomic per warp.
__device__ __forceinline__ void push_stack(unsigned short *stack, int *count, bool want, int idx)
{
    const unsigned lane = threadIdx.x & 31;
    const unsigned m = __ballot_sync(FULL, want);
    int base = 0;
    if (lane == 0 && m) base = atomicAdd(count, __popc(m));
    base = __shfl_sync(FULL, base, 0);
    if (want) stack[base + __popc(m & ((1u << lane) - 1u))] = (unsigned short) idx;
}

// Four groups of four spheres against one ray; lanes whose ray might touch a group queue the group index.
template <int KIND>
__device__ __forceinline__ void scan_block(const float4 *__restrict__ tab, int g0, int n_groups, const PivotRay &pr,
                                           const DirRay &dr, const RayF &rf, float thr, unsigned short *cq, int stride,
                                           int &cn)
{
#pragma unroll
    for (int u = 0; u < 4; u++) {
        if (u < n_groups) {
            const int gi = g0 + u;
            const float4 a = tab[4 * gi], b = tab[4 * gi + 1], c = tab[4 * gi + 2], d = tab[4 * gi + 3];
            float m;
            if (KIND == FK_PIVOT) m = max4(pivot_slack(pr, a), pivot_slack(pr, b), pivot_slack(pr, c), pivot_slack(pr, d));
            else if (KIND == FK_DIRECTION) m = max4(dir_slack(dr, a), dir_slack(dr, b), dir_slack(dr, c), dir_slack(dr, d));
            else m = max4(sphere_slack(rf, a), sphere_slack(rf, b), sphere_slack(rf, c), sphere_slack(rf, d));
            if (!(m < thr)) {
                cq[cn * stride] = (unsigned short) gi;
                cn++;
            }
        }
    }
}

} // namespace

// ---- setup kernel ------------------------------------------------------------------------------------------------
// Rows [row_begin, row_end) of the table set: 0 generic {centre, r^2}, 1 pivot = (X, Y, Z) (primary rays, rebuilt
// every frame), 2 + j light j (once per scene).
__global__ void build_table_rows_kernel(const double4 *sphere, const float4 *sphere_f, int n_sphere, int n_pad,
                                        const __grid_constant__ QueueScene qs, int row_begin, int row_end,
                                        double X, double Y, double Z, float4 *tables)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int row = row_begin + blockIdx.y;
    if (i >= n_pad || row >= row_end) return;
    float4 e;
    if (row == 0) {
        e = i < n_sphere ? sphere_f[i] : never_entry(FK_GENERIC);
    }
    else if (row == 1) {
        e = i < n_sphere ? pivot_entry(sphere[i], X, Y, Z) : never_entry(FK_PIVOT);
    }
    else {
        const LightFilter f = qs.lf[row - 2];
        const LightDev l = qs.light[row - 2];
        if (f.kind == FK_PIVOT) e = i < n_sphere ? pivot_entry(sphere[i], l.px, l.py, l.pz) : never_entry(FK_PIVOT);
        else if (f.kind == FK_DIRECTION) e = i < n_sphere ? direction_entry(sphere[i], f) : never_entry(FK_DIRECTION);
        else e = i < n_sphere ? sphere_f[i] : never_entry(FK_GENERIC);
    }
    tables[(size_t) row * n_pad + i] = e;
}

// ---- the render kernel ---------------------------------------------------------------------------------------------
template <int W, int MINB, int CAPM, bool COUNT>
__global__ void __launch_bounds__(W, MINB)
render_queue_kernel(const __grid_constant__ SceneView g, const __grid_constant__ FrameParams f,
                    const __grid_constant__ QueueConfig cfg, const __grid_constant__ QueueScene qs)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    constexpr int TILE_ROWS = W / 32;
    using Smem = QueueSmem<W, CAPM>;
    Smem &S = *reinterpret_cast<Smem *>(smem_raw);
    float4 *s_tab = reinterpret_cast<float4 *>(smem_raw + qk_fixed_smem<W, CAPM>());
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    unsigned short *cq = &S.cand[0][tid];   // this thread's candidate queue, stride W

    // ---- once per CTA: resident tables, empty stacks ---------------------------------------------------------------
    if (!cfg.streaming) {
        const int n = (2 + g.n_lights) * cfg.n_sphere_pad;
        for (int i = tid; i < n; i += W) s_tab[i] = cfg.tables[i];
    }
    for (int i = tid; i < Smem::CAP; i += W) S.stk_free[i] = (unsigned short) i;
    if (tid < QK_MAX_LIGHT) S.cache[tid] = -1;
    if (tid == 0) {
        S.n_free = Smem::CAP;
        S.n_shade = S.n_trace = 0;
        S.tiles_left = 1;
    }
    __syncthreads();

    Tally tl;
    tl.clear();
    const d3 cam_origin = d3{f.origin[0], f.origin[1], f.origin[2]};
    const int n_tiles = cfg.tiles_x * cfg.tiles_y;
    const int n_poly = g.n_plane + g.n_quadric + g.n_cubic;
    unsigned waves = 0;

    for (;;) {
        if (COUNT && (++waves & 63u) == 0) {   // keep the 32-bit per-thread tallies far from overflow
            flush_tally(tl, f.counters);
            tl.clear();
        }
        // ---- scheduler: which stack feeds this wave ------------------------------------------------------------------
        if (tid == 0) {
            const int ns = S.n_shade, nt = S.n_trace, nf = S.n_free;
            int mode = M_EXIT, take = 0, tile = -1;
            if (ns >= W) mode = M_SHADE;
            else if (nt >= W) mode = M_POOL;
            else {
                if (S.tiles_left && nf >= W) {
                    tile = (int) atomicAdd(cfg.tile_counter, 1ULL);
                    if (tile < n_tiles) mode = M_TILE;
                    else S.tiles_left = 0;
                }
                if (mode == M_EXIT) {   // nothing full: the larger leftover (only when the image has run out)
                    if (ns > 0 && ns >= nt) mode = M_SHADE;
                    else if (nt > 0) mode = M_POOL;
                }
            }
            if (mode == M_SHADE) {
                take = ns < W ? ns : W;
                S.n_shade = ns - take;
                S.base = ns - take;
            }
            else if (mode == M_POOL) {
                take = nt < W ? nt : W;
                S.n_trace = nt - take;
                S.base = nt - take;
            }
            S.mode = mode;
            S.take = take;
            S.tile = tile;
        }
        __syncthreads();
        const int mode = S.mode;
        if (mode == M_EXIT) break;

        // ---- this thread's work item ---------------------------------------------------------------------------------
        bool active = false;
        int rec = 0, pixel = 0;
        d3 qo = d3{0.0, 0.0, 0.0}, qd = d3{0.0, 0.0, 1.0};   // the ray of the current query
        d3 hp = qo;                                          // SHADE: hit point
        if (mode == M_TILE) {
            const int tile = S.tile;
            const int tx = tile % cfg.tiles_x, ty = tile / cfg.tiles_x;
            const int x = tx * 32 + (warp & 3) * 8 + (lane & 7);
            const int lrow = f.row0 + ty * TILE_ROWS + (warp >> 2) * 4 + (lane >> 3);
            const int y = local_to_image_row(lrow, f.strip_rows, f.rank, f.world);
            const bool in_buf = x < f.width && lrow < f.row0 + f.rows && lrow < f.local_rows;
            active = in_buf && y < f.height;
            pixel = (int) out_index(f, lrow, y, x);
            if (in_buf && !active && !f.direct) f.out[pixel] = make_uchar4(0, 0, 0, 0);   // padding row past the image
            if (active) {
                qo = cam_origin;
                qd = primary_dir(f, cam_origin, x, y);
                if (COUNT) tl.n[CNT_PRIMARY]++;
            }
        }
        else {
            active = tid < S.take;
            if (active) {
                rec = (mode == M_SHADE ? S.stk_shade : S.stk_trace)[S.base + tid];
                hp = d3{S.px[rec], S.py[rec], S.pz[rec]};
                const d3 n = d3{S.nx[rec], S.ny[rec], S.nz[rec]};
                // shadow rays and the reflection ray start at P + 1e-2 n (update-cpu.cpp:67,110)
                qo = d3{hp.x + RT_SHADOW_BIAS * n.x, hp.y + RT_SHADOW_BIAS * n.y, hp.z + RT_SHADOW_BIAS * n.z};
                if (mode == M_POOL) qd = reflect(d3{S.dx[rec], S.dy[rec], S.dz[rec]}, n);   // light_impl.h:46-49
            }
        }

        // ---- the rays of this wave: one nearest-hit ray (TILE, POOL) or one shadow ray per light (SHADE) ----------
        const bool any = (mode == M_SHADE);
        const int n_iter = any ? g.n_lights : 1;
        float acc_r = 0.f, acc_g = 0.f, acc_b = 0.f;
        double best_t = CUDART_INF;
        int best_slot = -1;
        for (int it = 0; it < n_iter; it++) {
            // -- ray and filter family (it, kind, row are uniform across the CTA)
            double tmax = RT_MAX_T;
            int kind = FK_GENERIC, row = 0;
            if (mode == M_TILE) {
                kind = FK_PIVOT;
                row = 1;
            }
            else if (any) {
                kind = qs.lf[it].kind;
                row = (kind == FK_GENERIC) ? 0 : 2 + it;
                if (qs.light[it].is_spherical) {   // light_impl.h:20-22: float(L - P), max_t = 1
                    tmax = 1.0;
                    qd = d3{(double) (float) (qs.light[it].px - hp.x), (double) (float) (qs.light[it].py - hp.y),
                            (double) (float) (qs.light[it].pz - hp.z)};
                }
                else {
                    qd = d3{qs.lf[it].sx, qs.lf[it].sy, qs.lf[it].sz};
                }
                if (COUNT && active) tl.n[CNT_SHADOW]++;
            }
            // -- two stages, cheaper or more promising one first: an any-hit ray starts with the class of the object
            //    that last occluded this light (shadow cache, any occluder will do — update-cpu.cpp:66-72)
            const int sphere_lo = g.n_plane, sphere_hi = g.n_plane + g.n_sphere;
            const int cache_slot = (any && cfg.use_shadow_cache) ? S.cache[it] : -1;
            const bool cache_is_sphere = cache_slot >= sphere_lo && cache_slot < sphere_hi;
            const bool poly_first = cache_slot >= 0 ? !cache_is_sphere : (cfg.poly_first != 0);
            bool done = !active;        // lanes without a ray, and any-hit lanes that found their occluder
            best_t = CUDART_INF;
            best_slot = -1;

            // -- planes, quadrics, cubics: exact FP64; the cached occluder goes first.  Inlined twice (before and after
            //    the sphere stage); only one of the two runs.
            auto poly_stage = [&]() {
                int k = 0;
                if (cache_slot >= 0 && !cache_is_sphere) k = cache_slot < sphere_lo ? cache_slot : cache_slot - g.n_sphere;
                for (int kk = 0; kk < n_poly; kk++, k = (k + 1 == n_poly) ? 0 : k + 1) {
                    if (!__all_sync(FULL, done)) {
                        Poly p;
                        int slot;
                        if (k < g.n_plane) {
                            p.t3 = 0.0;
                            p.t2 = 0.0;
                            p.t1 = qs.plane[k].x * qd.x + qs.plane[k].y * qd.y + qs.plane[k].z * qd.z;
                            p.t0 = qs.plane[k].x * qo.x + qs.plane[k].y * qo.y + qs.plane[k].z * qo.z + qs.plane[k].w;
                            slot = k;
                            if (COUNT && !done) tl.n[CNT_T_PLANE]++;
                        }
                        else if (k < g.n_plane + g.n_quadric) {
                            const int qi = k - g.n_plane;
                            p = expand_quadric(qs.quadric[qi], qo, qd);
                            slot = g.n_plane + g.n_sphere + qi;
                            if (COUNT && !done) tl.n[CNT_T_QUADRIC]++;
                        }
                        else {
                            const int ci = k - g.n_plane - g.n_quadric;
                            p = expand_cubic_masked(qs.cubic[ci], qs.cubic_mask[ci], qo, qd);
                            slot = g.n_plane + g.n_sphere + g.n_quadric + ci;
                            if (COUNT && !done) tl.n[CNT_T_CUBIC]++;
                        }
                        const double t = solve_poly_inline(p);
                        if (!done) {
                            if (any) {
                                if (blocks(t, tmax)) {
                                    best_slot = slot;
                                    done = true;
                                    if (cfg.use_shadow_cache) S.cache[it] = slot;   // benign race
                                }
                            }
                            else {
                                consider_lazy_t(best_t, best_slot, t, slot, g.orig_index);
                            }
                        }
                    }
                }
            };
            if (poly_first) poly_stage();

            // -- spheres: FP32 table scan, exact FP64 confirmation of the queued candidates
            if (g.n_sphere > 0) {
                PivotRay pr = {0.f, 0.f, 0.f, 0.f};
                DirRay dr = {0.f, 0.f};
                RayF rf;
                float thr = RT_NEG_INF_F;   // -inf: every group is a candidate (the exact test decides)
                if (kind == FK_PIVOT) {
                    if (mode == M_TILE) {
                        pr.ex = (float) qd.x; pr.ey = (float) qd.y; pr.ez = (float) qd.z;
                        pr.A = fmaf(pr.ex, pr.ex, fmaf(pr.ey, pr.ey, pr.ez * pr.ez));
                        if (cfg.tables_safe && cfg.primary_safe && pr.A < 1e18f && pr.A > 2e-7f) thr = 0.0f;
                    }
                    else {
                        pr.ex = (float) (qo.x - qs.light[it].px);
                        pr.ey = (float) (qo.y - qs.light[it].py);
                        pr.ez = (float) (qo.z - qs.light[it].pz);
                        pr.A = fmaf(pr.ex, pr.ex, fmaf(pr.ey, pr.ey, pr.ez * pr.ez));
                        const float De = sqrtf(pr.A) * 1.0001f;
                        const float eps = 0.010001f + 4.0f * RT_U32F * (De + 0.02f);
                        const float kappa = pr.A * 1.0002f * (2.0f * cfg.r_max * eps + eps * eps);
                        if (cfg.tables_safe && pr.A < 1e18f && pr.A > 2e-7f) thr = -kappa;   // > 2e-7: the reference is in its quadratic branch
                    }
                }
                else if (kind == FK_DIRECTION) {
                    dr.ou = (float) (qo.x * qs.lf[it].ux + qo.y * qs.lf[it].uy + qo.z * qs.lf[it].uz);
                    dr.ov = (float) (qo.x * qs.lf[it].vx + qo.y * qs.lf[it].vy + qo.z * qs.lf[it].vz);
                    const float mag = fmaxf(fabsf(dr.ou), fabsf(dr.ov));
                    const float e1 = 2.0f * RT_U32F * (mag + cfg.center_norm_max) * 1.001f;
                    const float kappa = 2.0f * (3.0f * cfg.r_max * e1 + 2.1f * e1 * e1);
                    if (cfg.tables_safe && mag < 1e9f) thr = -kappa * 1.0001f;
                }
                else {
                    rf = make_rayf(qo, qd, cfg.center_max, cfg.r2_max);
                    if (rf.safe) thr = 0.0f;
                }
                // shadow cache: the sphere that last occluded this light is tried first
                if (cache_is_sphere) {
                    const int cs = cache_slot - sphere_lo;
                    if (!done) {
                        if (COUNT) { tl.n[CNT_T_SPHERE]++; tl.n[CNT_CONFIRM]++; }
                        if (blocks(intersect_sphere(g.sphere[cs], qo, qd), tmax)) {
                            best_slot = g.n_plane + cs;
                            done = true;
                        }
                    }
                }
                int cn = 0;
                const float4 *gtab = cfg.tables + (size_t) row * cfg.n_sphere_pad;
                for (int c0 = 0; c0 < cfg.n_sphere_pad; c0 += cfg.chunk) {
                    const int n_in = min(cfg.chunk, cfg.n_sphere_pad - c0);
                    const float4 *tab;
                    if (cfg.streaming) {
                        // all threads of the CTA walk the chunks together and leave together once every ray is
                        // decided; this barrier also orders the previous readers of the chunk buffer before the refill
                        if (__syncthreads_and(done ? 1 : 0)) break;
                        for (int i = tid; i < n_in; i += W) s_tab[i] = gtab[c0 + i];
                        __syncthreads();
                        tab = s_tab;
                    }
                    else {
                        tab = s_tab + row * cfg.n_sphere_pad + c0;
                    }
                    const int groups = n_in >> 2;
                    bool skip = __all_sync(FULL, done);
                    // any-hit rays confirm their candidates after 32, 64, 128, ... groups (an occluder found early ends
                    // the scan early); nearest-hit rays only when a lane's queue fills up or the chunk ends
                    int next_drain = any ? 32 : groups;
                    for (int g0 = 0; g0 < groups && !skip; g0 += 4) {
                        const int ng = min(4, groups - g0);
                        if (COUNT && !done) tl.n[CNT_T_SPHERE] += min(4 * ng, g.n_sphere - c0 - 4 * g0);
                        if (kind == FK_PIVOT) scan_block<FK_PIVOT>(tab, g0, ng, pr, dr, rf, thr, cq, W, cn);
                        else if (kind == FK_DIRECTION) scan_block<FK_DIRECTION>(tab, g0, ng, pr, dr, rf, thr, cq, W, cn);
                        else scan_block<FK_GENERIC>(tab, g0, ng, pr, dr, rf, thr, cq, W, cn);
                        if (done) cn = 0;
                        const bool due = g0 + 4 >= groups || g0 + 4 >= next_drain;
                        if ((due && __any_sync(FULL, cn > 0)) || __any_sync(FULL, cn > QK_CAND - 4)) {
                            // drain: exact tests of the queued groups, all lanes of the warp together
                            while (__any_sync(FULL, cn > 0)) {
                                if (cn > 0) {
                                    cn--;
                                    const int first = c0 + 4 * (int) cq[cn * W];
#pragma unroll 1
                                    for (int idx = first; idx < first + 4; idx++) {
                                        // which of the four spheres made the group a candidate: its own slack again
                                        const float4 e = tab[idx - c0];
                                        const float sl = kind == FK_PIVOT ? pivot_slack(pr, e)
                                                         : (kind == FK_DIRECTION ? dir_slack(dr, e) : sphere_slack(rf, e));
                                        if (idx < g.n_sphere && !done && !(sl < thr)) {
                                            if (COUNT) tl.n[CNT_CONFIRM]++;
                                            const double t = intersect_sphere(g.sphere[idx], qo, qd);
                                            if (any) {
                                                if (blocks(t, tmax)) {
                                                    best_slot = g.n_plane + idx;
                                                    done = true;
                                                    if (cfg.use_shadow_cache) S.cache[it] = g.n_plane + idx;   // benign race
                                                }
                                            }
                                            else {
                                                consider_lazy_t(best_t, best_slot, t, g.n_plane + idx, g.orig_index);
                                            }
                                        }
                                    }
                                    if (done) cn = 0;
                                }
                            }
                            skip = __all_sync(FULL, done);
                        }
                        if (g0 + 4 >= next_drain) next_drain *= 2;
                    }
                }
            }

            if (!poly_first) poly_stage();

            // -- SHADE: the light contributes unless something blocks it (update-cpu.cpp:66-76)
            if (any && active && best_slot < 0) {
                const d3 n = d3{S.nx[rec], S.ny[rec], S.nz[rec]};
                const float4 mat = __ldg(&g.material[S.obj[rec]]);
                float3 acc = make_float3(acc_r, acc_g, acc_b);
                shade_light(qs.light[it], hp, n, make_float3(mat.y, mat.z, mat.w), acc);
                acc_r = acc.x; acc_g = acc.y; acc_b = acc.z;
                if (COUNT) tl.n[CNT_LIT]++;
            }
        }

        // ---- epilogue of the wave ------------------------------------------------------------------------------------
        bool final_px = false;
        float out_r = 0.f, out_g = 0.f, out_b = 0.f;
        bool to_shade = false, to_trace = false, to_free = false;
        if (any) {
            if (active) {
                const float cr = (acc_r < 1.0f) ? acc_r : 1.0f;   // glm::min(vec3(1), c), update-cpu.cpp:77
                const float cg = (acc_g < 1.0f) ? acc_g : 1.0f;
                const float cb = (acc_b < 1.0f) ? acc_b : 1.0f;
                float ratio = S.ratio[rec];
                // a primary hit starts from out = 0, ratio = 1: (1 - 1) * 0 + 1 * c == c exactly (update-cpu.cpp:93)
                out_r = blend(S.out_r[rec], ratio, cr);
                out_g = blend(S.out_g[rec], ratio, cg);
                out_b = blend(S.out_b[rec], ratio, cb);
                pixel = S.pixel[rec];
                const float refl = __ldg(&g.material[S.obj[rec]]).x;
                final_px = true;
                if (refl > RT_EPS) {   // update-cpu.cpp:97-111
                    ratio *= refl;
                    const int bounces = S.bounces[rec];
                    if (bounces == f.max_reflections) {
                        out_r = blend(out_r, ratio, f.bg[0]); out_g = blend(out_g, ratio, f.bg[1]); out_b = blend(out_b, ratio, f.bg[2]);
                    }
                    else {
                        S.bounces[rec] = bounces + 1;
                        S.ratio[rec] = ratio;
                        S.out_r[rec] = out_r; S.out_g[rec] = out_g; S.out_b[rec] = out_b;
                        final_px = false;
                        to_trace = true;
                        if (COUNT) tl.n[CNT_REFLECTION]++;
                    }
                }
                to_free = final_px;
            }
        }
        else {
            const bool hit = active && best_slot >= 0;
            if (active && !hit) {
                // miss: background (primary) or blend towards it (update-cpu.cpp:94-96,112-115)
                if (mode == M_TILE) { out_r = f.bg[0]; out_g = f.bg[1]; out_b = f.bg[2]; }
                else {
                    const float ratio = S.ratio[rec];
                    out_r = blend(S.out_r[rec], ratio, f.bg[0]);
                    out_g = blend(S.out_g[rec], ratio, f.bg[1]);
                    out_b = blend(S.out_b[rec], ratio, f.bg[2]);
                    pixel = S.pixel[rec];
                    to_free = true;
                }
                final_px = true;
            }
            if (mode == M_TILE) {   // records for the new hits: pop from the free stack, one shared atomic per warp
                const unsigned m = __ballot_sync(FULL, hit);
                int base = 0;
                if (lane == 0 && m) base = atomicSub(&S.n_free, __popc(m)) - __popc(m);
                base = __shfl_sync(FULL, base, 0);
                if (hit) {
                    rec = S.stk_free[base + __popc(m & ((1u << lane) - 1u))];
                    S.out_r[rec] = 0.f; S.out_g[rec] = 0.f; S.out_b[rec] = 0.f;
                    S.ratio[rec] = 1.0f;
                    S.bounces[rec] = 0;
                    S.pixel[rec] = pixel;
                }
            }
            if (hit) {
                const int obj = __ldg(&g.orig_index[best_slot]);
                const d3 p = d3{qo.x + best_t * qd.x, qo.y + best_t * qd.y, qo.z + best_t * qd.z};   // update-cpu.cpp:59
                const d3 n = surface_normal(g.coef_full + 20 * obj, p);
                S.px[rec] = p.x; S.py[rec] = p.y; S.pz[rec] = p.z;
                S.nx[rec] = n.x; S.ny[rec] = n.y; S.nz[rec] = n.z;
                S.dx[rec] = qd.x; S.dy[rec] = qd.y; S.dz[rec] = qd.z;
                S.obj[rec] = obj;
                to_shade = true;
                if (COUNT) tl.n[CNT_HITS]++;
            }
        }
        if (!any) push_stack(S.stk_shade, &S.n_shade, to_shade, rec);
        if (any) push_stack(S.stk_trace, &S.n_trace, to_trace, rec);
        if (mode != M_TILE) push_stack(S.stk_free, &S.n_free, to_free, rec);
        store_pixels(f.out, final_px, pixel, pack_rgba8(out_r, out_g, out_b));
        __syncthreads();
    }
    if (COUNT) flush_tally(tl, f.counters);
}

// ---- host side ----------------------------------------------------------------------------------------------------
namespace {

template <int W, int MINB, int CAPM>
cudaError_t launch_queue_variant(const SceneView &scene, const FrameParams &frame, QueueConfig cfg,
                                 const QueueScene &qs, size_t smem, int n_sm, int max_smem, cudaStream_t stream)
{
    static bool attr_set = false;
    auto kern = frame.counters ? render_queue_kernel<W, MINB, CAPM, true> : render_queue_kernel<W, MINB, CAPM, false>;
    cudaError_t e;
    if (!attr_set) {
        e = cudaFuncSetAttribute(render_queue_kernel<W, MINB, CAPM, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem);
        if (e != cudaSuccess) return e;
        e = cudaFuncSetAttribute(render_queue_kernel<W, MINB, CAPM, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem);
        if (e != cudaSuccess) return e;
        attr_set = true;
    }
    cfg.tiles_x = (frame.width + 31) / 32;
    cfg.tiles_y = (frame.rows + W / 32 - 1) / (W / 32);
    int per_sm = 1;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, W, smem);
    if (e != cudaSuccess) return e;
    if (per_sm < 1) return cudaErrorInvalidConfiguration;
    const int n_tiles = cfg.tiles_x * cfg.tiles_y;
    int grid = n_sm * per_sm;
    if (grid > n_tiles) grid = n_tiles;
    e = cudaMemsetAsync(cfg.tile_counter, 0, sizeof(unsigned long long), stream);
    if (e != cudaSuccess) return e;
    kern<<<grid, W, smem, stream>>>(scene, frame, cfg, qs);
    return cudaGetLastError();
}

} // namespace

bool queue_scene_fits(const SceneView &scene)
{
    return scene.n_plane <= QK_MAX_PLANE && scene.n_quadric <= QK_MAX_QUADRIC && scene.n_cubic <= QK_MAX_CUBIC &&
           scene.n_lights <= QK_MAX_LIGHT;
}

// Filter family of one light (host, once per scene): pivot for point lights, an orthonormal pair normal to the
// direction for directional lights, generic when the light's vector is unusable for an FP32 table.
LightFilter make_light_filter(const LightDev &l)
{
    LightFilter f;
    memset(&f, 0, sizeof(f));
    // the direction exactly as shadow_ray() hands it to intersect_ray(): rounded to float, widened (light_impl.h:24-25)
    f.sx = (double) (float) l.px;
    f.sy = (double) (float) l.py;
    f.sz = (double) (float) l.pz;
    const bool finite = fabs(l.px) < 1e9 && fabs(l.py) < 1e9 && fabs(l.pz) < 1e9;
    f.kind = FK_GENERIC;
    if (!finite) return f;
    if (l.is_spherical) {
        f.kind = FK_PIVOT;
        return f;
    }
    const double len = sqrt(f.sx * f.sx + f.sy * f.sy + f.sz * f.sz);
    if (!(len > 1e-30)) return f;
    const double nx = f.sx / len, ny = f.sy / len, nz = f.sz / len;
    // u = normalize(n x axis) for the axis least aligned with n, v = n x u
    const double ax = fabs(nx), ay = fabs(ny), az = fabs(nz);
    double ex = 0.0, ey = 0.0, ez = 0.0;
    if (ax <= ay && ax <= az) ex = 1.0;
    else if (ay <= az) ey = 1.0;
    else ez = 1.0;
    double ux = ny * ez - nz * ey, uy = nz * ex - nx * ez, uz = nx * ey - ny * ex;
    const double ul = sqrt(ux * ux + uy * uy + uz * uz);
    ux /= ul; uy /= ul; uz /= ul;
    f.ux = ux; f.uy = uy; f.uz = uz;
    f.vx = ny * uz - nz * uy;
    f.vy = nz * ux - nx * uz;
    f.vz = nx * uy - ny * ux;
    f.kind = FK_DIRECTION;
    return f;
}

cudaError_t build_queue_tables(const SceneView &scene, const QueueTables &qt, const QueueScene &qs, int row_begin,
                               int row_end, const double pivot[3], cudaStream_t stream)
{
    if (scene.n_sphere == 0 || row_end <= row_begin) return cudaSuccess;
    dim3 grid((qt.n_sphere_pad + 127) / 128, row_end - row_begin);
    build_table_rows_kernel<<<grid, 128, 0, stream>>>(scene.sphere, scene.sphere_f, scene.n_sphere, qt.n_sphere_pad, qs,
                                                      row_begin, row_end, pivot[0], pivot[1], pivot[2], qt.tables);
    return cudaGetLastError();
}

// Launch configuration: 256-thread CTAs, two per SM, with every table resident while that fits; otherwise one
// 512-thread CTA per SM, resident if the tables fit beside its records, else streamed through a chunk buffer.
cudaError_t launch_render_queue(const SceneView &scene, const FrameParams &frame, const QueueLaunch &ql, cudaStream_t stream)
{
    if (!ql.scene || !queue_scene_fits(scene)) return cudaErrorInvalidValue;
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    int n_sm = 0, max_smem = 0, sm_smem = 0;
    cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev);
    cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    cudaDeviceGetAttribute(&sm_smem, cudaDevAttrMaxSharedMemoryPerMultiprocessor, dev);

    QueueConfig cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.tile_counter = ql.tile_counter;
    cfg.n_sphere_pad = ql.tables.n_sphere_pad;
    cfg.tables_safe = ql.tables.tables_safe;
    cfg.primary_safe = ql.primary_safe;
    cfg.use_shadow_cache = ql.shadow_cache >= 0 ? ql.shadow_cache : (scene.n_objects >= 4 ? 1 : 0);
    // rough issue-slot cost of one full pass over each stage: cheaper stage first when no occluder is cached
    cfg.poly_first = (15LL * scene.n_plane + 150LL * scene.n_quadric + 700LL * scene.n_cubic < 7LL * scene.n_sphere) ? 1 : 0;
    cfg.r_max = ql.tables.r_max;
    cfg.center_max = scene.sphere_center_max;
    cfg.center_norm_max = ql.tables.center_norm_max;
    cfg.r2_max = ql.tables.r2_max;
    cfg.tables = ql.tables.tables;

    const size_t resident = sizeof(float4) * (size_t) (2 + scene.n_lights) * cfg.n_sphere_pad;
    // variants: 1: 256 threads x 2 CTAs/SM, 3 W records, tables resident     2: 512 threads, streamed
    //           3: 512 threads, resident     4: 256 threads x 2, streamed     5: 256 threads x 3 CTAs/SM, 2 W records, resident
    int variant = ql.variant;
    const size_t fixed[6] = {0, qk_fixed_smem<256, 3>(), qk_fixed_smem<512, 3>(), qk_fixed_smem<512, 3>(),
                             qk_fixed_smem<256, 3>(), qk_fixed_smem<256, 2>()};
    if (variant == 0) {
        if (2 * (fixed[1] + resident + 1024) <= (size_t) sm_smem) variant = 1;
        else if (fixed[3] + resident <= (size_t) max_smem) variant = 3;
        else variant = 2;
    }
    const bool streaming = (variant == 2 || variant == 4);
    cfg.streaming = streaming ? 1 : 0;
    size_t smem = fixed[variant];
    if (streaming) {
        const size_t room = (size_t) max_smem > smem ? (size_t) max_smem - smem : 0;
        int chunk = (int) (room / sizeof(float4)) & ~63;
        if (chunk > 4096) chunk = 4096;
        if (chunk > cfg.n_sphere_pad) chunk = cfg.n_sphere_pad;
        if (chunk < 64 && chunk < cfg.n_sphere_pad) return cudaErrorInvalidConfiguration;
        cfg.chunk = chunk > 0 ? chunk : 4;
        smem += sizeof(float4) * (size_t) cfg.chunk;
    }
    else {
        cfg.chunk = cfg.n_sphere_pad > 0 ? cfg.n_sphere_pad : 4;
        smem += resident;
    }
    if (smem > (size_t) max_smem) return cudaErrorInvalidConfiguration;
    switch (variant) {
        case 1: case 4: return launch_queue_variant<256, 2, 3>(scene, frame, cfg, *ql.scene, smem, n_sm, max_smem, stream);
        case 2: case 3: return launch_queue_variant<512, 1, 3>(scene, frame, cfg, *ql.scene, smem, n_sm, max_smem, stream);
        case 5: return launch_queue_variant<256, 3, 2>(scene, frame, cfg, *ql.scene, smem, n_sm, max_smem, stream);
        default: return cudaErrorInvalidValue;
    }
}

} // namespace rt
