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
//   planes, quadrics, cubics   exact FP64 in the reference's operation order (rt_math.cuh, rt_cubic.cuh); the
//             cubic zero-coefficient masks come from the kernel parameters, so their branches are uniform
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

// ---- setup kernels -------------------------------------------------------------------------------------------------
// Light families (once per scene): basis of every directional light, then one table row per light.
__global__ void build_light_filters_kernel(const LightDev *lights, int n_lights, LightFilter *lf)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n_lights) return;
    const LightDev l = lights[j];
    LightFilter f;
    f.ux = f.uy = f.uz = f.vx = f.vy = f.vz = 0.0;
    f.pad = 0;
    const bool finite = fabs(l.px) < 1e9 && fabs(l.py) < 1e9 && fabs(l.pz) < 1e9;
    if (!finite) {
        f.kind = FK_GENERIC;
    }
    else if (l.is_spherical) {
        f.kind = FK_PIVOT;
    }
    else {
        // direction exactly as shadow_ray() hands it to intersect_ray(): rounded to float, widened (light_impl.h:24-25)
        const double sx = (double) (float) l.px, sy = (double) (float) l.py, sz = (double) (float) l.pz;
        const double len = sqrt(sx * sx + sy * sy + sz * sz);
        if (!(len > 1e-30)) {
            f.kind = FK_GENERIC;
        }
        else {
            const double nx = sx / len, ny = sy / len, nz = sz / len;
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
        }
    }
    lf[j] = f;
}

// Row `row` of the table set: 0 generic {centre, r^2}, 1 pivot = (X, Y, Z) (primary rays), 2 + j light j.
__global__ void build_table_rows_kernel(const double4 *sphere, const float4 *sphere_f, int n_sphere, int n_pad,
                                        const LightDev *lights, const LightFilter *lf, int row_begin, int row_end,
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
        const LightFilter f = lf[row - 2];
        const LightDev l = lights[row - 2];
        if (f.kind == FK_PIVOT) e = i < n_sphere ? pivot_entry(sphere[i], l.px, l.py, l.pz) : never_entry(FK_PIVOT);
        else if (f.kind == FK_DIRECTION) e = i < n_sphere ? direction_entry(sphere[i], f) : never_entry(FK_DIRECTION);
        else e = i < n_sphere ? sphere_f[i] : never_entry(FK_GENERIC);
    }
    tables[(size_t) row * n_pad + i] = e;
}

// ---- the render kernel ---------------------------------------------------------------------------------------------
template <int W, bool COUNT>
__global__ void __launch_bounds__(W, (W <= 256 ? 2 : 1))
render_queue_kernel(const __grid_constant__ SceneView g, const __grid_constant__ FrameParams f,
                    const __grid_constant__ QueueConfig cfg)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    constexpr int TILE_ROWS = W / 32;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int cap = cfg.cap;

    // ---- carve shared memory -------------------------------------------------------------------------------------
    unsigned char *sp = smem_raw;
    Records R;
    {
        double *dp = reinterpret_cast<double *>(sp);
        R.px = dp; R.py = dp + cap; R.pz = dp + 2 * cap;
        R.nx = dp + 3 * cap; R.ny = dp + 4 * cap; R.nz = dp + 5 * cap;
        R.dx = dp + 6 * cap; R.dy = dp + 7 * cap; R.dz = dp + 8 * cap;
        float *fp = reinterpret_cast<float *>(dp + 9 * cap);
        R.out_r = fp; R.out_g = fp + cap; R.out_b = fp + 2 * cap; R.ratio = fp + 3 * cap;
        int *ip = reinterpret_cast<int *>(fp + 4 * cap);
        R.pixel = ip; R.obj = ip + cap; R.bounces = ip + 2 * cap;
        sp += align16(qk_record_bytes(cap));
    }
    unsigned short *stk_free = reinterpret_cast<unsigned short *>(sp);
    unsigned short *stk_shade = stk_free + cap;
    unsigned short *stk_trace = stk_shade + cap;
    sp += align16((size_t) 3 * cap * sizeof(unsigned short));
    unsigned short *cq = reinterpret_cast<unsigned short *>(sp) + tid;   // this thread's candidate queue, stride W
    sp += align16((size_t) W * QK_CAND * sizeof(unsigned short));
    QueueCtl &ctl = *reinterpret_cast<QueueCtl *>(sp);
    sp += align16(sizeof(QueueCtl));

    // ---- stage the scene once per CTA ------------------------------------------------------------------------------
    const double4 *s_plane;
    const double *s_quadric, *s_cubic;
    const LightDev *s_lights;
    const LightFilter *s_lf;
    float4 *s_tab;
    {
        double4 *pl = reinterpret_cast<double4 *>(sp);
        sp += align16(sizeof(double4) * g.n_plane);
        for (int i = tid; i < g.n_plane; i += W) pl[i] = g.plane[i];
        s_plane = pl;
        double *qd = reinterpret_cast<double *>(sp);
        sp += align16(sizeof(double) * 10 * g.n_quadric);
        for (int i = tid; i < 10 * g.n_quadric; i += W) qd[i] = g.quadric[i];
        s_quadric = qd;
        double *cb = reinterpret_cast<double *>(sp);
        sp += align16(sizeof(double) * 20 * g.n_cubic);
        for (int i = tid; i < 20 * g.n_cubic; i += W) cb[i] = g.cubic[i];
        s_cubic = cb;
        LightDev *lt = reinterpret_cast<LightDev *>(sp);
        sp += align16(sizeof(LightDev) * g.n_lights);
        for (int i = tid; i < g.n_lights; i += W) lt[i] = g.lights[i];
        s_lights = lt;
        LightFilter *lf = reinterpret_cast<LightFilter *>(sp);
        sp += align16(sizeof(LightFilter) * g.n_lights);
        for (int i = tid; i < g.n_lights; i += W) lf[i] = cfg.light_filter[i];
        s_lf = lf;
        s_tab = reinterpret_cast<float4 *>(sp);
        if (!cfg.streaming) {
            const int n = (2 + g.n_lights) * cfg.n_sphere_pad;
            for (int i = tid; i < n; i += W) s_tab[i] = cfg.tables[i];
        }
        for (int i = tid; i < cap; i += W) stk_free[i] = (unsigned short) i;
        if (tid < QK_CACHE) ctl.cache[tid] = -1;
        if (tid == 0) {
            ctl.n_free = cap;
            ctl.n_shade = ctl.n_trace = 0;
            ctl.tiles_left = 1;
        }
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
            const int ns = ctl.n_shade, nt = ctl.n_trace, nf = ctl.n_free;
            int mode = M_EXIT, take = 0, tile = -1;
            if (ns >= W) mode = M_SHADE;
            else if (nt >= W) mode = M_POOL;
            else {
                if (ctl.tiles_left && nf >= W) {
                    tile = (int) atomicAdd(cfg.tile_counter, 1ULL);
                    if (tile < n_tiles) mode = M_TILE;
                    else ctl.tiles_left = 0;
                }
                if (mode == M_EXIT) {   // nothing full: the larger leftover (end of the image, or records running out)
                    if (ns > 0 && ns >= nt) mode = M_SHADE;
                    else if (nt > 0) mode = M_POOL;
                }
            }
            if (mode == M_SHADE) {
                take = ns < W ? ns : W;
                ctl.n_shade = ns - take;
                ctl.base = ns - take;
            }
            else if (mode == M_POOL) {
                take = nt < W ? nt : W;
                ctl.n_trace = nt - take;
                ctl.base = nt - take;
            }
            ctl.mode = mode;
            ctl.take = take;
            ctl.tile = tile;
        }
        __syncthreads();
        const int mode = ctl.mode;
        if (mode == M_EXIT) break;
        const int take = ctl.take;

        // ---- this thread's work item ---------------------------------------------------------------------------------
        bool active = false;
        int rec = -1, pixel = 0;
        d3 o = d3{0.0, 0.0, 0.0}, d = d3{0.0, 0.0, 1.0};
        if (mode == M_TILE) {
            const int tile = ctl.tile;
            const int tx = tile % cfg.tiles_x, ty = tile / cfg.tiles_x;
            const int x = tx * 32 + (warp & 3) * 8 + (lane & 7);
            const int lrow = ty * TILE_ROWS + (warp >> 2) * 4 + (lane >> 3);
            const int y = local_to_image_row(lrow, f.strip_rows, f.rank, f.world);
            const bool in_buf = x < f.width && lrow < f.local_rows;
            active = in_buf && y < f.height;
            pixel = cfg.direct ? y * f.width + x : lrow * f.width + x;
            if (in_buf && !active && !cfg.direct) f.out[pixel] = make_uchar4(0, 0, 0, 0);   // padding row past the image
            if (active) {
                o = cam_origin;
                d = primary_dir(f, cam_origin, x, y);
                if (COUNT) tl.n[CNT_PRIMARY]++;
            }
        }
        else {
            active = tid < take;
            if (active) {
                rec = (mode == M_SHADE ? stk_shade : stk_trace)[ctl.base + tid];
                if (mode == M_POOL) {   // reflect_ray + biased origin, update-cpu.cpp:109-110
                    const d3 p = d3{R.px[rec], R.py[rec], R.pz[rec]};
                    const d3 n = d3{R.nx[rec], R.ny[rec], R.nz[rec]};
                    d = reflect(d3{R.dx[rec], R.dy[rec], R.dz[rec]}, n);
                    o = d3{p.x + RT_SHADOW_BIAS * n.x, p.y + RT_SHADOW_BIAS * n.y, p.z + RT_SHADOW_BIAS * n.z};
                }
            }
        }

        // ---- the rays of this wave: one nearest-hit ray (TILE, POOL) or one shadow ray per light (SHADE) ----------
        const int n_iter = (mode == M_SHADE) ? g.n_lights : 1;
        float acc_r = 0.f, acc_g = 0.f, acc_b = 0.f;
        double best_t = CUDART_INF;
        int best_slot = -1;
        for (int it = 0; it < n_iter; it++) {
            // -- ray and filter family
            d3 qo = o, qd = d;
            double tmax = RT_MAX_T;
            const bool any = (mode == M_SHADE);
            int kind = FK_GENERIC, row = 0;
            if (mode == M_TILE) {
                kind = FK_PIVOT;
                row = 1;
            }
            else if (mode == M_SHADE) {
                kind = s_lf[it].kind;
                row = (kind == FK_GENERIC) ? 0 : 2 + it;
                if (active) {
                    const d3 p = d3{R.px[rec], R.py[rec], R.pz[rec]};
                    const d3 n = d3{R.nx[rec], R.ny[rec], R.nz[rec]};
                    qo = d3{p.x + RT_SHADOW_BIAS * n.x, p.y + RT_SHADOW_BIAS * n.y, p.z + RT_SHADOW_BIAS * n.z};
                    qd = shadow_dir(s_lights[it], p, tmax);
                    if (COUNT) tl.n[CNT_SHADOW]++;
                }
            }
            bool done = !active;        // lanes without a ray, and any-hit lanes that found their occluder
            best_t = CUDART_INF;
            best_slot = -1;

            // -- spheres: FP32 table scan, exact FP64 confirmation of the queued candidates
            if (g.n_sphere > 0) {
                PivotRay pr = {0.f, 0.f, 0.f, 0.f};
                DirRay dr = {0.f, 0.f};
                RayF rf;
                float thr = RT_NEG_INF_F;   // -inf: every group is a candidate (exact test decides)
                if (kind == FK_PIVOT) {
                    if (mode == M_TILE) {
                        pr.ex = (float) qd.x; pr.ey = (float) qd.y; pr.ez = (float) qd.z;
                        pr.A = fmaf(pr.ex, pr.ex, fmaf(pr.ey, pr.ey, pr.ez * pr.ez));
                        if (cfg.tables_safe && cfg.primary_safe && pr.A < 1e18f) thr = 0.0f;
                    }
                    else {
                        const LightDev &l = s_lights[it];
                        pr.ex = (float) (qo.x - l.px); pr.ey = (float) (qo.y - l.py); pr.ez = (float) (qo.z - l.pz);
                        pr.A = fmaf(pr.ex, pr.ex, fmaf(pr.ey, pr.ey, pr.ez * pr.ez));
                        const float De = sqrtf(pr.A) * 1.0001f;
                        const float eps = 0.010001f + 4.0f * RT_U32F * (De + 0.02f);
                        const float kappa = pr.A * 1.0002f * (2.0f * cfg.r_max * eps + eps * eps);
                        if (cfg.tables_safe && pr.A < 1e18f) thr = -kappa;
                    }
                }
                else if (kind == FK_DIRECTION) {
                    const LightFilter &lf = s_lf[it];
                    dr.ou = (float) (qo.x * lf.ux + qo.y * lf.uy + qo.z * lf.uz);
                    dr.ov = (float) (qo.x * lf.vx + qo.y * lf.vy + qo.z * lf.vz);
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
                if (any && cfg.use_shadow_cache && it < QK_CACHE) {
                    const int cs = ctl.cache[it];
                    if (cs >= 0 && !done) {
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
                    for (int g0 = 0; g0 < groups && !skip; g0 += 4) {
                        const int ng = min(4, groups - g0);
                        if (COUNT && !done) tl.n[CNT_T_SPHERE] += min(4 * ng, g.n_sphere - c0 - 4 * g0);
                        if (kind == FK_PIVOT) scan_block<FK_PIVOT>(tab, g0, ng, pr, dr, rf, thr, cq, W, cn);
                        else if (kind == FK_DIRECTION) scan_block<FK_DIRECTION>(tab, g0, ng, pr, dr, rf, thr, cq, W, cn);
                        else scan_block<FK_GENERIC>(tab, g0, ng, pr, dr, rf, thr, cq, W, cn);
                        if (done) cn = 0;
                        const bool last = g0 + 4 >= groups;
                        if (last || __any_sync(FULL, cn > QK_CAND - 4)) {
                            // drain: exact tests of the queued groups, all lanes of the warp together
                            while (__any_sync(FULL, cn > 0)) {
                                if (cn > 0) {
                                    cn--;
                                    const int first = c0 + 4 * (int) cq[cn * W];
#pragma unroll 1
                                    for (int idx = first; idx < first + 4; idx++) {
                                        if (idx < g.n_sphere && !done) {
                                            if (COUNT) tl.n[CNT_CONFIRM]++;
                                            const double t = intersect_sphere(g.sphere[idx], qo, qd);
                                            if (any) {
                                                if (blocks(t, tmax)) {
                                                    best_slot = g.n_plane + idx;
                                                    done = true;
                                                    if (cfg.use_shadow_cache && it < QK_CACHE) ctl.cache[it] = idx;   // benign race
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
                    }
                    // (the barrier at the top of the next chunk — or of the next query — protects the chunk buffer)
                }
            }

            // -- planes, quadrics, cubics: exact FP64, one solver instantiation
            for (int k = 0; k < n_poly; k++) {
                if (__all_sync(FULL, done)) break;
                Poly p;
                int slot;
                if (k < g.n_plane) {
                    const double4 c = s_plane[k];
                    p.t3 = 0.0;
                    p.t2 = 0.0;
                    p.t1 = c.x * qd.x + c.y * qd.y + c.z * qd.z;
                    p.t0 = c.x * qo.x + c.y * qo.y + c.z * qo.z + c.w;
                    slot = k;
                    if (COUNT && !done) tl.n[CNT_T_PLANE]++;
                }
                else if (k < g.n_plane + g.n_quadric) {
                    const int qi = k - g.n_plane;
                    p = expand_quadric(s_quadric + 10 * qi, qo, qd);
                    slot = g.n_plane + g.n_sphere + qi;
                    if (COUNT && !done) tl.n[CNT_T_QUADRIC]++;
                }
                else {
                    const int ci = k - g.n_plane - g.n_quadric;
                    const unsigned mask = ci < QK_MAX_MASKS ? cfg.cubic_mask[ci] : 0xfffffu;
                    p = expand_cubic_masked(s_cubic + 20 * ci, mask, qo, qd);
                    slot = g.n_plane + g.n_sphere + g.n_quadric + ci;
                    if (COUNT && !done) tl.n[CNT_T_CUBIC]++;
                }
                const double t = solve_poly_inline(p);
                if (!done) {
                    if (any) {
                        if (blocks(t, tmax)) {
                            best_slot = slot;
                            done = true;
                        }
                    }
                    else {
                        consider_lazy_t(best_t, best_slot, t, slot, g.orig_index);
                    }
                }
            }

            // -- SHADE: the light contributes unless something blocks it (update-cpu.cpp:66-76)
            if (mode == M_SHADE && active && best_slot < 0) {
                const d3 p = d3{R.px[rec], R.py[rec], R.pz[rec]};
                const d3 n = d3{R.nx[rec], R.ny[rec], R.nz[rec]};
                const float4 mat = __ldg(&g.material[R.obj[rec]]);
                float3 acc = make_float3(acc_r, acc_g, acc_b);
                shade_light(s_lights[it], p, n, make_float3(mat.y, mat.z, mat.w), acc);
                acc_r = acc.x; acc_g = acc.y; acc_b = acc.z;
                if (COUNT) tl.n[CNT_LIT]++;
            }
        }

        // ---- epilogue of the wave ------------------------------------------------------------------------------------
        bool final_px = false;
        float out_r = 0.f, out_g = 0.f, out_b = 0.f;
        bool to_shade = false, to_trace = false, to_free = false;
        if (mode == M_SHADE) {
            if (active) {
                const float cr = (acc_r < 1.0f) ? acc_r : 1.0f;   // glm::min(vec3(1), c), update-cpu.cpp:77
                const float cg = (acc_g < 1.0f) ? acc_g : 1.0f;
                const float cb = (acc_b < 1.0f) ? acc_b : 1.0f;
                float ratio = R.ratio[rec];
                // a primary hit starts from out = 0, ratio = 1: (1 - 1) * 0 + 1 * c == c exactly (update-cpu.cpp:93)
                out_r = blend(R.out_r[rec], ratio, cr);
                out_g = blend(R.out_g[rec], ratio, cg);
                out_b = blend(R.out_b[rec], ratio, cb);
                pixel = R.pixel[rec];
                const float refl = __ldg(&g.material[R.obj[rec]]).x;
                final_px = true;
                if (refl > RT_EPS) {   // update-cpu.cpp:97-111
                    ratio *= refl;
                    const int bounces = R.bounces[rec];
                    if (bounces == f.max_reflections) {
                        out_r = blend(out_r, ratio, f.bg[0]); out_g = blend(out_g, ratio, f.bg[1]); out_b = blend(out_b, ratio, f.bg[2]);
                    }
                    else {
                        R.bounces[rec] = bounces + 1;
                        R.ratio[rec] = ratio;
                        R.out_r[rec] = out_r; R.out_g[rec] = out_g; R.out_b[rec] = out_b;
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
                    const float ratio = R.ratio[rec];
                    out_r = blend(R.out_r[rec], ratio, f.bg[0]);
                    out_g = blend(R.out_g[rec], ratio, f.bg[1]);
                    out_b = blend(R.out_b[rec], ratio, f.bg[2]);
                    pixel = R.pixel[rec];
                    to_free = true;
                }
                final_px = true;
            }
            if (mode == M_TILE) {   // records for the new hits: pop from the free stack, one shared atomic per warp
                const unsigned m = __ballot_sync(FULL, hit);
                int base = 0;
                if (lane == 0 && m) base = atomicSub(&ctl.n_free, __popc(m)) - __popc(m);
                base = __shfl_sync(FULL, base, 0);
                if (hit) {
                    rec = stk_free[base + __popc(m & ((1u << lane) - 1u))];
                    R.out_r[rec] = 0.f; R.out_g[rec] = 0.f; R.out_b[rec] = 0.f;
                    R.ratio[rec] = 1.0f;
                    R.bounces[rec] = 0;
                    R.pixel[rec] = pixel;
                }
            }
            if (hit) {
                const int obj = __ldg(&g.orig_index[best_slot]);
                const d3 p = d3{o.x + best_t * d.x, o.y + best_t * d.y, o.z + best_t * d.z};   // update-cpu.cpp:59
                const d3 n = surface_normal(g.coef_full + 20 * obj, p);
                R.px[rec] = p.x; R.py[rec] = p.y; R.pz[rec] = p.z;
                R.nx[rec] = n.x; R.ny[rec] = n.y; R.nz[rec] = n.z;
                R.dx[rec] = d.x; R.dy[rec] = d.y; R.dz[rec] = d.z;
                R.obj[rec] = obj;
                to_shade = true;
                if (COUNT) tl.n[CNT_HITS]++;
            }
        }
        if (mode != M_SHADE) push_stack(stk_shade, &ctl.n_shade, to_shade, rec);
        if (mode == M_SHADE) push_stack(stk_trace, &ctl.n_trace, to_trace, rec);
        if (mode != M_TILE) push_stack(stk_free, &ctl.n_free, to_free, rec);
        store_pixels(f.out, final_px, pixel, pack_rgba8(out_r, out_g, out_b));
        __syncthreads();
    }
    if (COUNT) flush_tally(tl, f.counters);
}

// ---- host side ----------------------------------------------------------------------------------------------------
namespace {

template <int W>
cudaError_t launch_queue_variant(const SceneView &scene, const FrameParams &frame, QueueConfig cfg, size_t smem,
                                 int n_sm, int max_smem, cudaStream_t stream)
{
    static bool attr_set = false;
    auto kern = frame.counters ? render_queue_kernel<W, true> : render_queue_kernel<W, false>;
    cudaError_t e;
    if (!attr_set) {
        e = cudaFuncSetAttribute(render_queue_kernel<W, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem);
        if (e != cudaSuccess) return e;
        e = cudaFuncSetAttribute(render_queue_kernel<W, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem);
        if (e != cudaSuccess) return e;
        attr_set = true;
    }
    cfg.tiles_x = (frame.width + 31) / 32;
    cfg.tiles_y = (frame.local_rows + W / 32 - 1) / (W / 32);
    int per_sm = 1;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, W, smem);
    if (e != cudaSuccess) return e;
    if (per_sm < 1) return cudaErrorInvalidConfiguration;
    const int n_tiles = cfg.tiles_x * cfg.tiles_y;
    int grid = n_sm * per_sm;
    if (grid > n_tiles) grid = n_tiles;
    e = cudaMemsetAsync(cfg.tile_counter, 0, sizeof(unsigned long long), stream);
    if (e != cudaSuccess) return e;
    kern<<<grid, W, smem, stream>>>(scene, frame, cfg);
    return cudaGetLastError();
}

} // namespace

cudaError_t build_queue_tables(const SceneView &scene, const QueueTables &qt, int row_begin, int row_end,
                               const double pivot[3], cudaStream_t stream)
{
    if (scene.n_sphere == 0 || row_end <= row_begin) return cudaSuccess;
    dim3 grid((qt.n_sphere_pad + 127) / 128, row_end - row_begin);
    build_table_rows_kernel<<<grid, 128, 0, stream>>>(scene.sphere, scene.sphere_f, scene.n_sphere, qt.n_sphere_pad,
                                                      scene.lights, qt.light_filter, row_begin, row_end, pivot[0],
                                                      pivot[1], pivot[2], qt.tables);
    return cudaGetLastError();
}

cudaError_t build_light_filters(const SceneView &scene, const QueueTables &qt, cudaStream_t stream)
{
    if (scene.n_lights == 0) return cudaSuccess;
    build_light_filters_kernel<<<(scene.n_lights + 127) / 128, 128, 0, stream>>>(scene.lights, scene.n_lights,
                                                                                 qt.light_filter);
    return cudaGetLastError();
}

// Launch configuration: 256-thread CTAs, two per SM, with every table resident while that fits; otherwise one
// 512-thread CTA per SM that streams the tables through a chunk buffer.
cudaError_t launch_render_queue(const SceneView &scene, const FrameParams &frame, const QueueLaunch &ql, cudaStream_t stream)
{
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
    cfg.use_shadow_cache = ql.shadow_cache >= 0 ? ql.shadow_cache : (scene.n_sphere >= 4 ? 1 : 0);
    cfg.direct = ql.direct;
    cfg.r_max = ql.tables.r_max;
    cfg.center_max = scene.sphere_center_max;
    cfg.center_norm_max = ql.tables.center_norm_max;
    cfg.r2_max = ql.tables.r2_max;
    cfg.tables = ql.tables.tables;
    cfg.light_filter = ql.tables.light_filter;
    for (int i = 0; i < QK_MAX_MASKS; i++) cfg.cubic_mask[i] = i < ql.n_masks ? ql.cubic_mask_host[i] : 0xfffffu;

    const size_t scene_smem = qk_scene_smem(scene);
    const size_t resident = sizeof(float4) * (size_t) (2 + scene.n_lights) * cfg.n_sphere_pad;
    const int cap_mult = ql.cap_mult >= 2 ? ql.cap_mult : 3;
    int variant = ql.variant;   // 0 auto, 1: W=256 resident, 2: W=512 streaming, 3: W=512 resident, 4: W=256 streaming
    const size_t need256 = qk_fixed_smem(256, cap_mult * 256) + scene_smem + resident;
    const size_t need512 = qk_fixed_smem(512, cap_mult * 512) + scene_smem;
    if (variant == 0) {
        if (2 * (need256 + 1024) <= (size_t) sm_smem && need256 <= (size_t) max_smem) variant = 1;
        else if (need512 + resident <= (size_t) max_smem && resident <= 32 * 1024) variant = 3;
        else variant = 2;
    }
    const bool streaming = (variant == 2 || variant == 4);
    const int W = (variant == 1 || variant == 4) ? 256 : 512;
    cfg.cap = cap_mult * W;
    cfg.streaming = streaming ? 1 : 0;
    size_t smem = qk_fixed_smem(W, cfg.cap) + scene_smem;
    if (streaming) {
        size_t room = (size_t) max_smem > smem ? (size_t) max_smem - smem : 0;
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
    if (cfg.cap > 65535) return cudaErrorInvalidConfiguration;
    if (W == 256) return launch_queue_variant<256>(scene, frame, cfg, smem, n_sm, max_smem, stream);
    return launch_queue_variant<512>(scene, frame, cfg, smem, n_sm, max_smem, stream);
}

} // namespace rt
