// rt_wave_kernel.inl — body of the wavefront kernel (included by rt_wave.cu).
namespace rt {

namespace {

// Write finished pixels of a PRIMARY wave.  Lanes 4i..4i+3 of a warp hold four horizontally adjacent pixels
// (8x4 sub-tile per warp); when all four are final the group leader stores them as one 128-bit uchar4x4.
__device__ __forceinline__ void store_primary(uchar4 *out, size_t index, bool final_px, uchar4 px, bool vec_ok)
{
    const unsigned lane = threadIdx.x & 31;
    const unsigned done = __ballot_sync(0xffffffffu, final_px);
    const unsigned group = (done >> (lane & ~3u)) & 0xfu;
    unsigned v = *reinterpret_cast<unsigned *>(&px);
    const unsigned v1 = __shfl_down_sync(0xffffffffu, v, 1);
    const unsigned v2 = __shfl_down_sync(0xffffffffu, v, 2);
    const unsigned v3 = __shfl_down_sync(0xffffffffu, v, 3);
    if (vec_ok && group == 0xfu) {
        if ((lane & 3u) == 0) *reinterpret_cast<uint4 *>(out + index) = make_uint4(v, v1, v2, v3);
    }
    else if (final_px) {
        out[index] = px;
    }
}

// Next tile for this CTA (thread 0 only): index into the tile space of the launch's frame, n_tiles = none left.
// Static split: this GPU's counter counts the tiles of its own strips.  Dynamic ownership (WaveDyn): the counter
// counts tiles of SLOTS; the slot's row group comes from the table, which is filled `ahead` slots in advance from the
// job's counter in rank 0's memory.
// (Measured in round 2 and dropped: pulling several tiles per atomic, pulling the next tile one wave ahead — no gain on
// one GPU, the other CTAs of the SM cover the round trip, and a coarser tail on several — and a warp-collective
// variant that pre-reads the slot table while the atomic is in flight: 2 us per frame better with the dynamic split,
// 1 % worse without.)
// Split in two so that the counter's round trip overlaps a wave of tracing: pull_tile() issues the atomic, resolve_tile()
// — called after phase A — turns its result into a tile index.
__device__ __forceinline__ unsigned pull_tile(const WaveConfig &cfg)
{
    return atomicAdd(reinterpret_cast<unsigned *>(cfg.tile_counter), 1u);
}

__device__ __forceinline__ int resolve_tile(const WaveConfig &cfg, WaveCtl &ctl, unsigned t, int n_tiles)
{
    if (cfg.dyn.counters == nullptr) return t < (unsigned) n_tiles ? (int) t : n_tiles;
    const unsigned tiles_per_group = cfg.tiles_per_group;
    const unsigned k = magic_div(t, cfg.inv_tiles_per_group), r = t - k * tiles_per_group;   // t * tiles_per_group < 2^32
    // past the table: the CTAs' last pulls overshoot the counter (one each); every slot that can hold a group lies below
    // max_slots = n_groups + ahead + 1 and has been handed out completely by then
    if (k >= (unsigned) cfg.dyn.max_slots) return n_tiles;
    if (r == 0u && k + cfg.dyn.ahead < (unsigned) cfg.dyn.max_slots) {
        // first tile of slot k: fetch the group of slot k + ahead (groups 0 .. world * ahead - 1 were assigned statically)
        const unsigned got = atomicAdd_system(reinterpret_cast<unsigned *>(cfg.dyn.counters + (cfg.sync.seq & 1u)), 1u);
        const long long grp = (long long) got + (long long) cfg.sync.world * cfg.dyn.ahead;
        const int val = grp < cfg.dyn.n_groups ? (int) grp : -2;
        // relaxed: the group number is the whole message (an acquire on the reading side would flush the SM's L1 per tile)
        asm volatile("st.relaxed.gpu.global.s32 [%0], %1;" ::"l"(cfg.dyn.slot_group + k + cfg.dyn.ahead), "r"(val) : "memory");
    }
    int grp = ctl.dyn_group;
    if (ctl.dyn_slot != (int) k) {      // one table look-up per slot and CTA, not per tile
        const long long t0 = clock64();
        for (;;) {
            asm volatile("ld.relaxed.gpu.global.s32 %0, [%1];" : "=r"(grp) : "l"(cfg.dyn.slot_group + k) : "memory");
            if (grp != -1) break;
            if (clock64() - t0 > 40000000000LL) {   // ~20 s: report instead of hanging the box
                *cfg.sync.timed_out = 1u;
                grp = -2;
                break;
            }
            __nanosleep(40);
        }
        ctl.dyn_slot = (int) k;
        ctl.dyn_group = grp;
    }
    if (grp < 0) return n_tiles;
    RT_CHECK(grp < cfg.dyn.n_groups && (unsigned) grp * tiles_per_group + r < (unsigned) n_tiles);
    // the group's tiles in the frame's tile space: group g covers tile rows [g * h, (g + 1) * h), h = group_rows / tile_rows
    return (int) ((unsigned) grp * tiles_per_group + r);
}

// Warp 0: publish tile t (n_tiles = none left) in buffer b together with the camera-plane coordinates of its columns and
// rows — the two divisions of primary ray generation (update-cpu.cpp:84-87), once per column / row of the tile.
template <int W>
__device__ __forceinline__ void publish_tile(WaveCtl &ctl, int b, int t, int n_tiles, const FrameParams &f, const WaveConfig &cfg, int lane)
{
    if (lane == 0) ctl.next_tile[b] = t;
    if (t < n_tiles) {
        const int ty = (int) magic_div((unsigned) t, cfg.inv_tiles_x), tx = t - ty * cfg.tiles_x;
        if (lane < TileGeom<W>::TILE_W) ctl.cam_x[b][lane] = primary_cam_x(f, tx * TileGeom<W>::TILE_W + lane);
        if (lane < TileGeom<W>::TILE_ROWS) {
            const int lrow = f.row0 + ty * TileGeom<W>::TILE_ROWS + lane;
            ctl.cam_y[b][lane] = primary_cam_y(f, local_to_image_row(lrow, f.strip_rows, f.rank, f.world));
        }
    }
}

#ifdef RT_B200_TIMELINE
__device__ __forceinline__ unsigned long long tl_now()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
#define RT_TL(k, v) do { if (f.timeline != nullptr && threadIdx.x == 0 && blockIdx.x < RT_TL_MAX_CTAS) f.timeline[blockIdx.x * RT_TL_WORDS + (k)] = (v); } while (0)
#else
#define RT_TL(k, v) ((void) 0)
#endif

__device__ __forceinline__ float blend(float out, float ratio, float c) { return (1.0f - ratio) * out + ratio * c; }

} // namespace

// LONE: the scene is exactly one cubic-class object (four of the eight bundled scenes).  The class counts become
// compile-time constants, so the plane / sphere / quadric code and the class loops disappear from that instantiation:
// round 2's ncu captures showed only ~3 700 of the generic kernel's 11 800 instructions executing on such a scene, but
// scattered over its whole 185 KB, with 10-20 % of the stall samples waiting for instruction fetch at the far jumps.
template <int W, int MINB, bool COUNT, bool CUBIC, bool GRID, bool LONE = false>
__global__ void __launch_bounds__(W, MINB)
render_wave_kernel(const __grid_constant__ SceneView g_param, const __grid_constant__ FrameParams f,
                   const __grid_constant__ WaveConfig cfg)
{
    SceneView g_lone;
    if (LONE) {
        g_lone = g_param;
        g_lone.n_plane = g_lone.n_sphere = g_lone.n_quadric = 0;
        g_lone.n_cubic = 1;
        g_lone.n_objects = 1;
    }
    const SceneView &g = LONE ? g_lone : g_param;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    constexpr int WAVE = W;
    constexpr int WX = TileGeom<W>::WX;              // warps side by side in a tile, an 8x4 sub-tile per warp
    constexpr int TILE_W = TileGeom<W>::TILE_W;      // tile = TILE_W x TILE_ROWS pixels
    constexpr int TILE_ROWS = TileGeom<W>::TILE_ROWS;
    Pool<W> &pool = *reinterpret_cast<Pool<W> *>(smem_raw);
    WaveHits<W> &wh = *reinterpret_cast<WaveHits<W> *>(smem_raw + align16(sizeof(Pool<W>)));
    WaveCtl &ctl = *reinterpret_cast<WaveCtl *>(smem_raw + align16(sizeof(Pool<W>)) + align16(sizeof(WaveHits<W>)));
    unsigned char *sp = smem_raw + wave_fixed_smem<W>();

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    RT_TL(0, tl_now());
#ifdef RT_B200_TIMELINE
    unsigned tl_tiles = 0;
    bool tl_dry = false;
#endif

    // ---- stage the scene once per CTA ----------------------------------------------------------------
    SceneSmem s;
    {
        double4 *pl = reinterpret_cast<double4 *>(sp);
        sp += align16(sizeof(double4) * g.n_plane);
        for (int i = tid; i < g.n_plane; i += WAVE) pl[i] = g.plane[i];
        s.plane = pl;
        d3 *pn = reinterpret_cast<d3 *>(sp);
        sp += align16(sizeof(d3) * g.n_plane);
        for (int i = tid; i < g.n_plane; i += WAVE) {
            const double4 k = g.plane[i];
            pn[i] = normalize3(d3{k.x, k.y, k.z});
        }
        s.plane_n = pn;
        if (cfg.spheres_in_smem) {
            float4 *sf = reinterpret_cast<float4 *>(sp);
            sp += align16(sizeof(float4) * g.n_sphere);
            for (int i = tid; i < g.n_sphere; i += WAVE) sf[i] = g.sphere_f[i];
            s.sphere_f = sf;
        }
        else {
            s.sphere_f = g.sphere_f;
        }
        double *qd = reinterpret_cast<double *>(sp);
        sp += align16(sizeof(double) * 10 * g.n_quadric);
        for (int i = tid; i < 10 * g.n_quadric; i += WAVE) qd[i] = g.quadric[i];
        s.quadric = qd;
        double *cb = reinterpret_cast<double *>(sp);
        sp += align16(sizeof(double) * 20 * g.n_cubic);
        for (int i = tid; i < 20 * g.n_cubic; i += WAVE) cb[i] = g.cubic[i];
        s.cubic = cb;
        unsigned *cm = reinterpret_cast<unsigned *>(sp);
        sp += align16(sizeof(unsigned) * g.n_cubic);
        for (int i = tid; i < g.n_cubic; i += WAVE) cm[i] = cfg.cubic_mask[i];
        s.cubic_mask = cm;
        float *cf = reinterpret_cast<float *>(sp);
        sp += align16(sizeof(float) * RT_CF_FLOATS * g.n_cubic);
        for (int i = tid; i < g.n_cubic; i += WAVE) make_cubicf(g.cubic + 20 * i, cf + RT_CF_FLOATS * i);
        s.cubic_f = cf;
        LightDev *lt = reinterpret_cast<LightDev *>(sp);
        for (int i = tid; i < g.n_lights; i += WAVE) lt[i] = g.lights[i];
        s.lights = lt;
        if (tid == 0) {
            ctl.pool_count = 0;
            ctl.n_hits[0] = ctl.n_hits[1] = 0;
            ctl.occ_found[0] = ctl.occ_found[1] = 0;
            ctl.prim_rays = ctl.prim_hits = 0;
            ctl.prim_filter = 1;
            ctl.dyn_slot = -1;
            ctl.dyn_group = -1;
        }
        if (tid < LIGHT_CHUNK) ctl.shadow_cache[tid] = -1;
        __syncthreads();
        // cubic filter caches (rt_cubic_filter.cuh): what depends only on the camera origin / on a directional light
        sp += align16(sizeof(LightDev) * g.n_lights);
        CubicO *co = reinterpret_cast<CubicO *>(sp);
        sp += align16(sizeof(CubicO) * g.n_cubic);
        for (int i = tid; i < g.n_cubic; i += WAVE)
            co[i] = make_cubic_o(s.cubic_f + RT_CF_FLOATS * i, (float) f.origin[0], (float) f.origin[1], (float) f.origin[2]);
        s.cam_o = co;
        s.light_d = nullptr;
        s.hit_o = nullptr;
        if (CUBIC && wave_single_cubic(g)) {
            CubicD *ld = reinterpret_cast<CubicD *>(sp);
            sp += align16(sizeof(CubicD) * g.n_lights);
            for (int j = tid; j < g.n_lights; j += WAVE) {
                const LightDev l = s.lights[j];      // directional: shadow_dir() is the float-rounded light vector
                ld[j] = make_cubic_d(s.cubic_f, (float) l.px, (float) l.py, (float) l.pz);
            }
            s.light_d = ld;
            s.hit_o = reinterpret_cast<CubicO *>(sp);
        }
        // fused gather: frame hand-shake with rank 0 (WaveSync) — before the first pixel leaves this GPU
        if (cfg.sync.go != nullptr && tid == 0) {
            if (cfg.sync.rank == 0) {
                if (blockIdx.x == 0) {
                    if (cfg.dyn.counters != nullptr) cfg.dyn.counters[(cfg.sync.seq + 1u) & 1u] = 0ULL;   // the next frame's group counter (WaveDyn)
                    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(cfg.sync.go), "r"(cfg.sync.seq) : "memory");
                }
            }
            else {
                // a rank may run one frame ahead of rank 0 (two frame buffers).  With dynamic ownership a rank that starts
                // a frame early simply takes groups until the others join: work-conserving, nobody waits.  (Making
                // everybody wait for rank 0 here would deadlock ranks that share one GPU: the waiting CTAs would hold
                // the SM slots rank 0's kernel needs.)
                const unsigned need = cfg.sync.seq - 1u;
                const long long t0 = clock64();
                for (;;) {
                    unsigned v;
                    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(cfg.sync.go) : "memory");
                    if ((int) (v - need) >= 0) break;
                    if (clock64() - t0 > 40000000000LL) {   // ~20 s: report instead of hanging the box
                        *cfg.sync.timed_out = 1u;
                        break;
                    }
                    __nanosleep(100);
                }
            }
        }
    }
    __syncthreads();

    Tally tl;
    tl.clear();
    const d3 cam_origin = d3{f.origin[0], f.origin[1], f.origin[2]};
    const int n_tiles = cfg.tiles_x * cfg.tiles_y;
    const bool vec_ok = (f.width & 3) == 0;   // rows start 16-byte aligned (packed and assembled layout alike)
    unsigned waves = 0;
    // Barriers per wave: one after the hit list is built, one after the shadow rays (per chunk of 32 lights), one at the
    // end.  A primary wave needs none of its own: its tile and the tile's camera terms were put into ctl.next_tile /
    // cam_x / cam_y [par] by warp 0 during an earlier wave (round 1 spent three barriers on picking a tile, publishing it
    // and publishing the camera terms; `barrier` was 10-18 % of the stall samples).
    if (warp == 0) {
        int t = 0;
        if (lane == 0) t = resolve_tile(cfg, ctl, pull_tile(cfg), n_tiles);
        t = __shfl_sync(0xffffffffu, t, 0);
        publish_tile<W>(ctl, 0, t, n_tiles, f, cfg, lane);
    }
    __syncthreads();
    RT_TL(1, tl_now());
    int par = 0;   // the buffer that holds this CTA's next tile (uniform over the CTA)
    for (;;) {
        if (COUNT && (waves & 63u) == 63u) {   // keep the 32-bit per-thread tallies far from overflow
            flush_tally(tl, f.counters);
            tl.clear();
        }
        // ---- choose the wave: pooled bounce rays when a full wave is waiting (or no tile is left), else the next tile ----
        // (every thread decides for itself from words that are stable since the last barrier)
        const int pooled = ctl.pool_count;
        const int tile_next = ctl.next_tile[par];
        const bool prim_filter = ctl.prim_filter != 0;
        const bool primary = pooled < WAVE && tile_next < n_tiles;
#ifdef RT_B200_TIMELINE
        if (!tl_dry && tile_next >= n_tiles) { tl_dry = true; RT_TL(2, tl_now()); }
        if (primary) tl_tiles++;
#endif
        if (!primary && pooled == 0) break;
        const int tile = primary ? tile_next : -1;
        const int hp = (int) (waves & 1u);
        if (tid == 0) ctl.n_hits[hp ^ 1] = 0;   // last read before the previous wave's final barrier, next used after this wave's
        unsigned pulled = 0u;
        if (primary && tid == 0) pulled = pull_tile(cfg);   // the tile after this one: the result is consumed after phase A
        const int take = pooled < WAVE ? pooled : WAVE;
        const int pool_base = pooled - take;                // bounce wave: what stays in the pool

        // ---- ray of this thread -----------------------------------------------------------------------
        bool active;
        d3 o, d;
        float out_r = 0.f, out_g = 0.f, out_b = 0.f, ratio = 1.0f;
        int bounces = 0, pixel = 0;
        if (primary) {
            const int ty = (int) magic_div((unsigned) tile, cfg.inv_tiles_x), tx = tile - ty * cfg.tiles_x;
            const int x = tx * TILE_W + (warp % WX) * 8 + (lane & 7);
            const int lrow = f.row0 + ty * TILE_ROWS + (warp / WX) * 4 + (lane >> 3);
            const int y = local_to_image_row(lrow, f.strip_rows, f.rank, f.world);
            const bool in_buf = x < f.width && lrow < f.row0 + f.rows && lrow < f.local_rows;
            active = in_buf && y < f.height;
            pixel = (int) out_index(f, lrow, y, x);
            if (in_buf && !active && !f.direct) f.out[pixel] = make_uchar4(0, 0, 0, 0);   // padding row past the image
            if (active) {
                o = cam_origin;
                d = primary_dir_cam(f, cam_origin, ctl.cam_x[par][(warp % WX) * 8 + (lane & 7)], ctl.cam_y[par][(warp / WX) * 4 + (lane >> 3)]);
                if (COUNT) tl.n[CNT_PRIMARY]++;
            }
        }
        else {
            active = tid < take;
            if (active) {
                const int r = pool_base + tid;
                RT_CHECK(r >= 0 && r < 2 * WAVE);
                o = d3{pool.ox[r], pool.oy[r], pool.oz[r]};
                d = d3{pool.dx[r], pool.dy[r], pool.dz[r]};
                out_r = pool.out_r[r]; out_g = pool.out_g[r]; out_b = pool.out_b[r];
                ratio = pool.ratio[r];
                pixel = pool.pixel[r];
                bounces = pool.bounces[r];
            }
        }

        // ---- phase A + B: nearest hit, hit record, compact hit list ---------------------------------------
        int obj = -1;
        if (active) {
            Hit h = nearest_hit_wave<CUBIC, GRID, (W >= 512)>(g, s, cfg, o, d, primary, !primary || prim_filter, tl);
            if (h.slot >= 0) {
                obj = h.orig;
                const d3 p = d3{o.x + h.t * d.x, o.y + h.t * d.y, o.z + h.t * d.z};
                // normal_vector() (surface_impl.h:157-172).  Kernels without the cubic class evaluate only the coefficients
                // the hit object's class can have (rt_math.cuh: *_gradient).  The kernels with it keep the full formula for
                // every class: the class branches cost them more than they save (measured: +1 % on the four single-cubic
                // scenes; a mask-driven gradient for the cubic class itself +3 % on clebsch).
                d3 n;
                if (!CUBIC) {
                    int sl = h.slot;
                    if (sl < g.n_plane) n = s.plane_n[sl];
                    else if ((sl -= g.n_plane) < g.n_sphere) n = normalize3(sphere_gradient(g.sphere[sl], p));
                    else n = normalize3(quadric_gradient(s.quadric + 10 * (sl - g.n_sphere), p));
                }
                else {
                    n = surface_normal(g.coef_full + 20 * obj, p);
                }
                wh.px[tid] = p.x; wh.py[tid] = p.y; wh.pz[tid] = p.z;
                wh.nx[tid] = n.x; wh.ny[tid] = n.y; wh.nz[tid] = n.z;
                wh.occluded[tid] = 0u;
                if (CUBIC && s.hit_o && cfg.use_cubic_filter)   // origin half of the filter, shared by this hit's shadow rays
                    s.hit_o[tid] = make_cubic_o(s.cubic_f, (float) (p.x + RT_SHADOW_BIAS * n.x), (float) (p.y + RT_SHADOW_BIAS * n.y),
                                                (float) (p.z + RT_SHADOW_BIAS * n.z));
                if (COUNT) tl.n[CNT_HITS]++;
            }
        }
        const bool hit = obj >= 0;
        bool final_px = false;
        if (active && !hit) {
            // miss: background (primary) or blend towards it (update-cpu.cpp:94-96,112-115)
            if (primary) { out_r = f.bg[0]; out_g = f.bg[1]; out_b = f.bg[2]; }
            else { out_r = blend(out_r, ratio, f.bg[0]); out_g = blend(out_g, ratio, f.bg[1]); out_b = blend(out_b, ratio, f.bg[2]); }
            final_px = true;
        }
        if (primary && warp == 0) {
            // the tile after this one, pulled before phase A: into the other buffer, nobody reads that one during this wave.
            // (Pulling it after the shadow phase instead — so that a CTA grinding through a heavy wave does not sit on a
            // reserved tile — was measured: 20spheres -15 % at an eighth of a frame, the cheap scenes +3 %, the step unchanged.)
            int t = 0;
            if (lane == 0) t = resolve_tile(cfg, ctl, pulled, n_tiles);
            t = __shfl_sync(0xffffffffu, t, 0);
            publish_tile<W>(ctl, par ^ 1, t, n_tiles, f, cfg, lane);
        }
        {
            // hit list: one shared atomic per warp (the order of the list is immaterial: it only spreads the shadow tasks)
            const unsigned m = __ballot_sync(0xffffffffu, hit);
            int wbase = 0;
            if (lane == 0 && m) wbase = atomicAdd(&ctl.n_hits[hp], __popc(m));
            wbase = __shfl_sync(0xffffffffu, wbase, 0);
            RT_CHECK(!hit || wbase + __popc(m & ((1u << lane) - 1u)) < WAVE);
            if (hit) wh.hitlist[wbase + __popc(m & ((1u << lane) - 1u))] = (unsigned short) tid;
        }
        __syncthreads();
        const int n_hits = ctl.n_hits[hp];
        if (!primary && tid == 0) ctl.pool_count = pool_base;   // everybody has read it; appends follow the next barrier
        if (CUBIC && primary && tid == 0 && g.n_cubic > 0) {
            // the filter costs about a third of an exact test and saves one only for (ray, cubic) pairs that cannot give
            // the nearest hit: a ray has one nearest hit, so with several cubics most pairs are of that kind
            int rays = ctl.prim_rays + WAVE, hits = ctl.prim_hits + n_hits;
            if (rays >= 16 * WAVE) { rays >>= 1; hits >>= 1; }
            ctl.prim_rays = rays;
            ctl.prim_hits = hits;
            ctl.prim_filter = (5LL * hits < 3LL * rays * g.n_cubic) ? 1 : 0;   // one cubic: on while fewer than 60 % of the primary rays hit
        }

        // ---- phases C + D per chunk of 32 lights ---------------------------------------------------------------
        float acc_r = 0.f, acc_g = 0.f, acc_b = 0.f;
        float4 mat = make_float4(0.f, 0.f, 0.f, 0.f);
        if (hit) mat = __ldg(&g.material_pi[obj]);   // {reflection_ratio, colour / pi}
        for (int l0 = 0; l0 < g.n_lights; l0 += LIGHT_CHUNK) {
            const int lc = min(LIGHT_CHUNK, g.n_lights - l0);
            if (l0 > 0) {   // first chunk: cleared with the hit record, before the barrier that published the records
                if (hit) wh.occluded[tid] = 0u;
                __syncthreads();
            }
            // C: task = (light j, hit i), light-major
            const unsigned inv_hits = n_hits > 1 ? 0xffffffffu / (unsigned) n_hits + 1u : 0u;   // ceil(2^32 / n), 32-bit division
            const bool lone_cubic = CUBIC && s.hit_o != nullptr && cfg.use_cubic_filter && g.n_objects == 1;
            for (int task = tid; task < n_hits * lc; task += WAVE) {
                const int j = n_hits > 1 ? (int) __umulhi((unsigned) task, inv_hits) : task;   // task / n_hits (task * n_hits < 2^32)
                RT_CHECK(j >= 0 && j < lc && task - j * n_hits >= 0 && task - j * n_hits < n_hits);
                const int k = wh.hitlist[task - j * n_hits];
                RT_CHECK(k >= 0 && k < WAVE);
                const LightDev l = s.lights[l0 + j];
                const d3 p = d3{wh.px[k], wh.py[k], wh.pz[k]};
                const d3 n = d3{wh.nx[k], wh.ny[k], wh.nz[k]};
                if (cfg.cull_backfacing && light_cannot_contribute(l, p, n)) {
                    atomicOr(&wh.occluded[k], 1u << j);     // phase D skips it, exactly as if it were shadowed
                    continue;
                }
                if (COUNT) tl.n[CNT_SHADOW]++;
                bool occ = false;
                int verdict = CV_UNKNOWN;
                if (lone_cubic) {
                    // the only object is one cubic (four of the five bundled cubic scenes): the cached filter halves decide
                    // most shadow rays without touching FP64 (rt_cubic_filter.cuh), not even for the ray's origin and
                    // direction; the exact test below runs for the undecided ones
                    if (COUNT) tl.n[CNT_T_CUBIC]++;
                    const CubicD cd = l.is_spherical ? make_cubic_d(s.cubic_f, (float) (l.px - p.x), (float) (l.py - p.y), (float) (l.pz - p.z))
                                                     : s.light_d[l0 + j];
                    verdict = cubic_verdict_od(s.hit_o[k], cd, l.is_spherical ? 1.0f : (float) RT_MAX_T, true);
                    occ = verdict == CV_HIT;
                }
                if (verdict == CV_UNKNOWN) {
                    const d3 so = d3{p.x + RT_SHADOW_BIAS * n.x, p.y + RT_SHADOW_BIAS * n.y, p.z + RT_SHADOW_BIAS * n.z};
                    double max_t;
                    const d3 sd = shadow_dir(l, p, max_t);
                    if (lone_cubic) {
                        if (COUNT) tl.n[CNT_CUBIC_EXACT]++;
                        occ = blocks(intersect_cubic_masked(s.cubic, s.cubic_mask[0], so, sd), max_t);
                    }
                    else {
                        // per-light shadow cache: the slot that occluded this light last.  A plane / sphere / quadric is
                        // probed on its own; a cubic becomes the first cubic of the class search
                        int cached_cubic = -1;
                        if (cfg.use_shadow_cache) {
                            const int cs = ctl.shadow_cache[j];
                            const int n_other = g.n_plane + g.n_sphere + g.n_quadric;
                            if (cs >= n_other) cached_cubic = cs - n_other;
                            else if (cs >= 0) occ = blocks(intersect_slot(g, s, cs, so, sd, tl), max_t);
                        }
                        if (!occ) {
                            const int by = occluder_wave<CUBIC, GRID, (W >= 512)>(g, s, cfg, ctl, so, sd, max_t, l.is_spherical != 0, k, l0 + j, cached_cubic, tl);
                            occ = by >= 0;
                            if (occ && cfg.use_shadow_cache) ctl.shadow_cache[j] = by;   // benign race: any occluder will do
                        }
                    }
                }
                if (occ) atomicOr(&wh.occluded[k], 1u << j);
            }
            __syncthreads();
            // D (part 1): sum the unshadowed lights of this chunk in index order
            if (hit) {
                const unsigned occ = wh.occluded[tid];
                const d3 p = d3{wh.px[tid], wh.py[tid], wh.pz[tid]};
                const d3 n = d3{wh.nx[tid], wh.ny[tid], wh.nz[tid]};
                float3 acc = make_float3(acc_r, acc_g, acc_b);
                const float3 oc = make_float3(mat.y, mat.z, mat.w);
                for (int j = 0; j < lc; j++) {
                    if (!((occ >> j) & 1u)) {
                        shade_light_pi(s.lights[l0 + j], p, n, oc, acc);
                        if (COUNT) tl.n[CNT_LIT]++;
                    }
                }
                acc_r = acc.x; acc_g = acc.y; acc_b = acc.z;
            }
            if (l0 + LIGHT_CHUNK < g.n_lights) __syncthreads();   // the next chunk reuses the occlusion words
        }
        if (g.n_lights == 0) __syncthreads();   // no shadow phase: the pool bookkeeping above still has to precede the appends below

        // ---- phase D (part 2): clamp, blend, bounce ---------------------------------------------------------------
        bool spawn = false;
        d3 no, nd;
        if (hit) {
            const float cr = (acc_r < 1.0f) ? acc_r : 1.0f;   // glm::min(vec3(1), c), update-cpu.cpp:77
            const float cg = (acc_g < 1.0f) ? acc_g : 1.0f;
            const float cb = (acc_b < 1.0f) ? acc_b : 1.0f;
            if (primary) { out_r = cr; out_g = cg; out_b = cb; }
            else { out_r = blend(out_r, ratio, cr); out_g = blend(out_g, ratio, cg); out_b = blend(out_b, ratio, cb); }
            final_px = true;
            if (mat.x > RT_EPS) {   // update-cpu.cpp:100-111
                ratio *= mat.x;
                if (bounces == f.max_reflections) {
                    out_r = blend(out_r, ratio, f.bg[0]); out_g = blend(out_g, ratio, f.bg[1]); out_b = blend(out_b, ratio, f.bg[2]);
                }
                else {
                    bounces++;
                    const d3 p = d3{wh.px[tid], wh.py[tid], wh.pz[tid]};
                    const d3 n = d3{wh.nx[tid], wh.ny[tid], wh.nz[tid]};
                    nd = reflect(d, n);
                    no = d3{p.x + RT_SHADOW_BIAS * n.x, p.y + RT_SHADOW_BIAS * n.y, p.z + RT_SHADOW_BIAS * n.z};
                    spawn = true;
                    final_px = false;
                    if (COUNT) tl.n[CNT_REFLECTION]++;
                }
            }
        }
        // append bounce rays to the pool: warp ballot + prefix popc, one shared atomic per warp
        {
            const unsigned m = __ballot_sync(0xffffffffu, spawn);
            int wbase = 0;
            if (lane == 0 && m) wbase = atomicAdd(&ctl.pool_count, __popc(m));
            wbase = __shfl_sync(0xffffffffu, wbase, 0);
            if (spawn) {
                const int r = wbase + __popc(m & ((1u << lane) - 1u));
                RT_CHECK(r >= 0 && r < 2 * WAVE);
                pool.ox[r] = no.x; pool.oy[r] = no.y; pool.oz[r] = no.z;
                pool.dx[r] = nd.x; pool.dy[r] = nd.y; pool.dz[r] = nd.z;
                pool.out_r[r] = out_r; pool.out_g[r] = out_g; pool.out_b[r] = out_b;
                pool.ratio[r] = ratio;
                pool.pixel[r] = pixel;
                pool.bounces[r] = bounces;
            }
        }
        // finished pixels leave the SM
        RT_CHECK(!final_px || (pixel >= 0 && (size_t) pixel < (size_t) f.width * (size_t) (f.direct ? f.height : f.local_rows)));
        const uchar4 px = pack_rgba8(out_r, out_g, out_b);
        if (primary) store_primary(f.out, (size_t) pixel, final_px, px, vec_ok);
        else if (final_px) f.out[pixel] = px;
        __syncthreads();
        if (primary) par ^= 1;   // the buffer filled during this wave holds the next tile
        waves++;
    }
    if (COUNT) flush_tally(tl, f.counters);
    RT_TL(3, tl_now());
#ifdef RT_B200_TIMELINE
    RT_TL(5, (unsigned long long) waves);
    { unsigned smid; asm volatile("mov.u32 %0, %%smid;" : "=r"(smid)); RT_TL(6, (unsigned long long) smid); }
    RT_TL(7, (unsigned long long) tl_tiles);
#endif

    // ---- end of launch: the last CTA re-arms the counters and, with the fused gather, signals "frame complete" ----
    if (cfg.sync.go != nullptr) __threadfence_system();   // this thread's pixel stores, visible to rank 0's consumer
    __syncthreads();
    if (tid == 0) {
        const unsigned long long done = atomicAdd(cfg.tile_counter + 1, 1ULL);
        const bool last = done == (unsigned long long) gridDim.x - 1ULL;
        if (last) {
            cfg.tile_counter[0] = 0ULL;     // the next launch on this stream starts from tile 0 without a memset
            cfg.tile_counter[1] = 0ULL;
            if (cfg.sync.go != nullptr) {
                __threadfence_system();
                asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(cfg.sync.flags + cfg.sync.rank), "r"(cfg.sync.seq) : "memory");
            }
        }
        ctl.last_cta = last ? 1 : 0;
    }
    if (cfg.dyn.counters != nullptr) {
        // the last CTA re-arms this GPU's slot table for the next launch (every other CTA is done with it)
        __syncthreads();
        if (ctl.last_cta) {
            for (int k = tid; k < cfg.dyn.max_slots; k += WAVE) {
                int grp = -1;
                if (k < cfg.dyn.ahead) {
                    grp = k * cfg.sync.world + cfg.sync.rank;
                    if (grp >= cfg.dyn.n_groups) grp = -2;
                }
                cfg.dyn.slot_group[k] = grp;
            }
        }
    }
    if (cfg.sync.go != nullptr && cfg.sync.rank == 0) {
        __syncthreads();
        if (ctl.last_cta && tid < cfg.sync.world) {   // rank 0 consumes the frame: wait for every rank's flag
            const long long t0 = clock64();
            for (;;) {
                unsigned v;
                asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(cfg.sync.flags + tid) : "memory");
                if ((int) (v - cfg.sync.seq) >= 0) break;
                if (clock64() - t0 > 40000000000LL) {
                    *cfg.sync.timed_out = 1u;
                    break;
                }
                __nanosleep(100);
            }
        }
    }
    RT_TL(4, tl_now());
}

namespace {

template <int W, int MINB, bool CUBIC, bool GRID = false, bool LONE = false>
cudaError_t launch_wave_variant(const SceneView &scene, const FrameParams &frame, WaveConfig cfg, size_t scene_smem,
                                int n_sm, int max_smem, cudaStream_t stream)
{
    static std::atomic<unsigned long long> attr_set{0ULL};   // bit d: the attribute is set on device d (it is per device)
    auto kern = frame.counters ? render_wave_kernel<W, MINB, true, CUBIC, GRID, LONE> : render_wave_kernel<W, MINB, false, CUBIC, GRID, LONE>;
    cudaError_t e;
    int dev = 0;
    e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    // Setting it twice from two threads is harmless; devices past 63 set it on every launch.
    if (dev >= 64 || !((attr_set.load(std::memory_order_acquire) >> dev) & 1ULL)) {
        e = cudaFuncSetAttribute(render_wave_kernel<W, MINB, true, CUBIC, GRID, LONE>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(render_wave_kernel<W, MINB, false, CUBIC, GRID, LONE>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem);
        if (e != cudaSuccess) return e;
        if (dev < 64) attr_set.fetch_or(1ULL << dev, std::memory_order_release);
    }
    const size_t smem = wave_fixed_smem<W>() + scene_smem + wave_cubic_smem(scene, W);
    cfg.tiles_x = (frame.width + TileGeom<W>::TILE_W - 1) / TileGeom<W>::TILE_W;
    cfg.tiles_y = (frame.rows + TileGeom<W>::TILE_ROWS - 1) / TileGeom<W>::TILE_ROWS;
    if ((unsigned long long) cfg.tiles_x * cfg.tiles_y * (unsigned long long) cfg.tiles_x >= (1ULL << 32)) return cudaErrorInvalidConfiguration;
    cfg.inv_tiles_x = magic_inv((unsigned) cfg.tiles_x);
    cfg.tiles_per_group = (unsigned) (cfg.tiles_x * (cfg.dyn.group_rows > 0 ? cfg.dyn.group_rows / TileGeom<W>::TILE_ROWS : 1));
    cfg.inv_tiles_per_group = magic_inv(cfg.tiles_per_group);
    int per_sm = 1;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, W, smem);
    if (e != cudaSuccess) return e;
    if (per_sm < 1) return cudaErrorInvalidConfiguration;
    const int n_tiles = cfg.tiles_x * cfg.tiles_y;
    int grid = n_sm * per_sm;
    if (grid > n_tiles) grid = n_tiles;
    kern<<<grid, W, smem, stream>>>(scene, frame, cfg);   // tile counter: left at 0 by the previous launch's last CTA
    return cudaGetLastError();
}

// Variants (option "variant"): code -> (threads per CTA, CTAs per SM the register budget is cut for).
//   2: 256 x 2 (128 registers)   3: 256 x 3 (80 registers)   5: 512 x 1 (128 registers)
// Smaller CTAs at the same 24 warps per SM (128 x 6, 64 x 12, 32 x 24; 128 x 4, 64 x 8, 32 x 16 at 128 registers) were
// measured in round 2 (profiles/r2_a_variants.jsonl): slower on every bundled scene but quadratic / reflection_test
// (128 x 6: -2 %), so barriers are not what limits the kernel; they are not compiled.
#define RT_WAVE_VARIANTS(X) X(2, 256, 2) X(3, 256, 3) X(5, 512, 1)
} // namespace

// The kernels with the grid walk are instantiated in their own translation units (rt_wave_grid.cu with the cubic class,
// rt_wave_grid_plain.cu without), so that the four units compile in parallel.
cudaError_t launch_wave_grid_cubic(const SceneView &scene, const FrameParams &frame, const WaveConfig &cfg, size_t scene_smem,
                                   int n_sm, int max_smem, int variant, cudaStream_t stream);
cudaError_t launch_wave_grid_plain(const SceneView &scene, const FrameParams &frame, const WaveConfig &cfg, size_t scene_smem,
                                   int n_sm, int max_smem, int variant, cudaStream_t stream);
#ifdef RT_WAVE_GRID_TU
// two CTAs per SM at 128 registers by default: the walk spills at the 80 registers of the three-CTA variant (C5 at 4K
// 15.2 vs 16.1 ms)
cudaError_t RT_WAVE_GRID_TU(const SceneView &scene, const FrameParams &frame, const WaveConfig &cfg, size_t scene_smem,
                            int n_sm, int max_smem, int variant, cudaStream_t stream)
{
    if (variant == 3) return launch_wave_variant<256, 3, RT_WAVE_GRID_CUBIC, true>(scene, frame, cfg, scene_smem, n_sm, max_smem, stream);
    return launch_wave_variant<256, 2, RT_WAVE_GRID_CUBIC, true>(scene, frame, cfg, scene_smem, n_sm, max_smem, stream);
}
#endif

namespace {

// The variant launch_render_wave_t will pick for this scene: the host needs the tile height before the launch.
inline int wave_pick_variant(const SceneView &scene, const WaveLaunch &wl, int max_smem, int sm_smem)
{
    const bool grid = wl.accel && scene.grid.cell_start != nullptr && scene.n_sphere > 0;
    if (grid) return wl.variant == 3 ? 3 : 2;
    if (wl.variant > 0) return wl.variant;
    size_t scene_smem = wave_scene_smem(scene, true);
    if (wave_fixed_smem<512>() + scene_smem + wave_cubic_smem(scene, 512) > (size_t) max_smem) scene_smem = wave_scene_smem(scene, false);
    const size_t cta256 = wave_fixed_smem<256>() + scene_smem + wave_cubic_smem(scene, 256) + 1024;
    return (3 * cta256 <= (size_t) sm_smem) ? 3 : 5;
}

template <bool CUBIC>
cudaError_t launch_render_wave_t(const SceneView &scene, const FrameParams &frame, const WaveLaunch &wl, cudaStream_t stream)
{
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    int n_sm = 0, max_smem = 0, sm_smem = 0;
    cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev);
    cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    cudaDeviceGetAttribute(&sm_smem, cudaDevAttrMaxSharedMemoryPerMultiprocessor, dev);

    WaveConfig cfg;
    cfg.tiles_x = cfg.tiles_y = 0;
    cfg.tile_counter = wl.tile_counter;
    cfg.sync = WaveSync{wl.sync_go, wl.sync_flags, wl.sync_timed_out, wl.sync_seq, wl.sync_rank, wl.sync_world};
    cfg.dyn = WaveDyn{wl.dyn_counters, wl.dyn_slot_group, wl.dyn_n_groups, wl.dyn_group_rows, wl.dyn_max_slots, wl.dyn_ahead};
    cfg.cubic_mask = wl.cubic_mask;
    cfg.sphere_r2_max = wl.sphere_r2_max;
    cfg.use_shadow_cache = wl.shadow_cache >= 0 ? wl.shadow_cache : (scene.n_objects >= 8 ? 1 : 0);
    cfg.use_cubic_filter = wl.cubic_filter;
    cfg.cull_backfacing = wl.cull_backfacing;
    cfg.lone_kernel = wl.lone_kernel;
    // with the grid the FP32 sphere array is read a few candidates at a time (L1/L2): it is not staged, which leaves
    // room for three CTAs per SM; rays the walker refuses (rt_grid.cuh: grid_ray_ok) run the brute-force loop from L2
    cfg.use_grid = (wl.accel && scene.grid.cell_start != nullptr && scene.n_sphere > 0) ? 1 : 0;
    cfg.spheres_in_smem = cfg.use_grid ? 0 : 1;
    if (wl.spheres_smem == 0) cfg.spheres_in_smem = 0;   // option "spheres_smem": read the FP32 sphere records through L1 instead
    size_t scene_smem = wave_scene_smem(scene, cfg.spheres_in_smem != 0);
    if (cfg.spheres_in_smem && wave_fixed_smem<512>() + scene_smem + wave_cubic_smem(scene, 512) > (size_t) max_smem) {
        cfg.spheres_in_smem = 0;
        scene_smem = wave_scene_smem(scene, false);
    }
    if (!cfg.spheres_in_smem && wave_fixed_smem<256>() + scene_smem + wave_cubic_smem(scene, 256) > (size_t) max_smem)
        return cudaErrorInvalidConfiguration;
    const size_t cta256 = wave_fixed_smem<256>() + scene_smem + wave_cubic_smem(scene, 256) + 1024;   // + per-CTA reservation
    // the grid walk is compiled into the 256-thread variants only (nothing is staged for the spheres, so three CTAs fit)
    if (cfg.use_grid)
        return CUBIC ? launch_wave_grid_cubic(scene, frame, cfg, scene_smem, n_sm, max_smem, wl.variant, stream)
                     : launch_wave_grid_plain(scene, frame, cfg, scene_smem, n_sm, max_smem, wl.variant, stream);
    int variant = wl.variant;
    if (variant <= 0) variant = (3 * cta256 <= (size_t) sm_smem) ? 3 : 5;
    if (CUBIC && cfg.lone_kernel && scene.n_objects == 1 && scene.n_cubic == 1) {   // the lone-cubic instantiations
        if (variant == 3) return launch_wave_variant<256, 3, CUBIC, false, CUBIC>(scene, frame, cfg, scene_smem, n_sm, max_smem, stream);
        if (variant == 2) return launch_wave_variant<256, 2, CUBIC, false, CUBIC>(scene, frame, cfg, scene_smem, n_sm, max_smem, stream);
    }
    switch (variant) {
#define RT_WAVE_CASE(CODE, THREADS, MINB) \
        case CODE: return launch_wave_variant<THREADS, MINB, CUBIC>(scene, frame, cfg, scene_smem, n_sm, max_smem, stream);
        RT_WAVE_VARIANTS(RT_WAVE_CASE)
#undef RT_WAVE_CASE
        default: return cudaErrorInvalidValue;
    }
}

} // namespace

} // namespace rt
