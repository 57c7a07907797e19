// rt_grid.cuh — uniform grid over the sphere class: the walker, host- and device-callable.
//
// The reference tests every object for every ray (src/update-cpu.cpp:48-56, 66-72).  For the sphere class the set of
// objects a ray can hit is bounded in space, so a ray only has to look at the spheres registered in the grid cells
// it passes through.  The grid never decides a hit: every sphere it proposes still goes through the FP32 filter and
// the exact FP64 test of the brute-force path (rt_filter.cuh, rt_math.cuh: intersect_sphere), and the nearest-hit
// rule (smallest t, ties to the lowest original index) is applied to whatever those tests return.  What has to be
// shown is only that no sphere the REFERENCE would report is skipped:
//
//   * A reference hit at parameter t_s (EPS <= t_s < limit, quadratic branch, i.e. d.d > 1e-7) lies within rho of the
//     sphere's surface, where rho covers the rounding of the reference's expanded-form discriminant
//     (|d delta| <~ 2e-15 A M^2 with M = |o| + |C|, so a line missing the sphere by up to ~5e-15 M^2 / (2 r) can still
//     "hit") and the displacement of a near-tangent root along the ray (~1e-7 M, i.e. ~1e-14 M^2 / (2 r) off the
//     surface).  The builder (rt_grid_build.h) registers every sphere in all cells touched by its bounding box padded
//     by pad = 1e-6 S + 1e-10 Mmax^2 / r (S = scene scale, Mmax = 30 S), two to three orders of magnitude above rho,
//     and the walker is only used for rays with |o|_inf <= 16 S (SphereGrid::o_limit), finite components and
//     2e-7 < d.d < 1e30; every other ray takes the brute-force loop.  Spheres whose pad would exceed half a cell, or
//     that are much larger than the typical sphere, are kept on a "loose" list that is tested for every ray.
//   * The walker is a 3-D DDA in FP64.  Its cell sequence can differ from the exact one only where the ray passes
//     within ~1e-13 (relative) of a cell face; a point that close to a neighbouring cell is inside the padded box of
//     its sphere in that neighbour too, so the sphere is registered there as well.
//   * Nearest hit: the walk stops after a cell whose exit parameter t_out exceeds the best t so far.  A sphere with
//     t_s <= best_t < t_out has its hit point in the current or an earlier cell, hence was proposed already (an exact
//     tie in t included, so the lowest-index rule sees both candidates).
//   * Any hit (shadow rays): the walk covers [0, max_t]; the boolean does not depend on the order of the tests.
//
// A sphere spanning several cells is proposed several times; the tests are pure functions of (sphere, ray), so
// repeats cannot change the outcome.  tests/test_grid.py checks walker + builder on the CPU against brute force over
// the oracle's intersect_ray (random and adversarial scenes, rays through cell corners and along faces);
// tests/test_parity_gpu.py checks frames with the grid on against the grid off, byte for byte.
#ifndef RT_GRID_CUH
#define RT_GRID_CUH

#include <math.h>

#if defined(__CUDACC__)
#define RT_GHD __host__ __device__ __forceinline__
#else
#define RT_GHD static inline
#endif
#if defined(__CUDA_ARCH__)
#define RT_GLD(p) __ldg(p)
#define RT_GUNROLL _Pragma("unroll")
#else
#define RT_GLD(p) (*(p))
#define RT_GUNROLL
#endif

namespace rt {

struct SphereGrid {
    const int *cell_start;   // [nx * ny * nz + 1] offsets into items (null: no grid)
    const int *items;        // sphere-class indices per cell, then the loose list
    double lo[3], hi[3];     // padded bounds of all gridded spheres
    double cell[3];          // cell edge per axis
    double inv_cell[3];
    double o_limit;          // the walker is used for rays with |origin|_inf <= o_limit only (pad derivation above)
    int n[3];
    int n_gridded;           // spheres registered in cells
    int loose_begin, loose_end;   // items[loose_begin .. loose_end): spheres tested for every ray
};

// Rays the walker may be used for (see the header); everything else takes the brute-force loop.
RT_GHD bool grid_ray_ok(const SphereGrid &G, double ox, double oy, double oz, double dx, double dy, double dz)
{
    const double omax = fmax(fabs(ox), fmax(fabs(oy), fabs(oz)));
    const double A = dx * dx + dy * dy + dz * dz;
    return omax <= G.o_limit && A > 2e-7 && A < 1e30;   // NaN fails every comparison
}

// Walks the cells along o + t d for t in [0, t_limit] and calls v.test(sphere_index) for every registered sphere
// (loose list first).  v.test returns true to stop (any-hit found); v.bound() is the parameter beyond which nothing
// can matter any more (nearest hit: the best t so far).
template <class Visitor>
RT_GHD void grid_walk(const SphereGrid &G, double ox, double oy, double oz, double dx, double dy, double dz, double t_limit,
                      Visitor &v)
{
    for (int k = G.loose_begin; k < G.loose_end; k++)
        if (v.test(RT_GLD(G.items + k))) return;
    if (G.n_gridded == 0) return;

    const double o[3] = {ox, oy, oz}, d[3] = {dx, dy, dz};
    const double dmax = fmax(fabs(dx), fmax(fabs(dy), fabs(dz)));
    double inv[3];
    bool moving[3];
    double t0 = 0.0, t1 = t_limit;
RT_GUNROLL
    for (int a = 0; a < 3; a++) {
        // a component 1e-30 of the largest moves the point by less than 1e-30 of the path inside the grid: parallel
        moving[a] = fabs(d[a]) > 1e-30 * dmax;
        if (moving[a]) {
            inv[a] = 1.0 / d[a];
            double ta = (G.lo[a] - o[a]) * inv[a], tb = (G.hi[a] - o[a]) * inv[a];
            if (ta > tb) {
                const double tmp = ta;
                ta = tb;
                tb = tmp;
            }
            t0 = fmax(t0, ta);
            t1 = fmin(t1, tb);
        }
        else {
            inv[a] = 0.0;
            if (o[a] < G.lo[a] || o[a] > G.hi[a]) return;
        }
    }
    if (!(t0 <= t1)) return;

    int idx[3], step[3];
    double tn[3], dt[3];
RT_GUNROLL
    for (int a = 0; a < 3; a++) {
        const double p = o[a] + t0 * d[a];
        int i = (int) floor((p - G.lo[a]) * G.inv_cell[a]);
        i = i < 0 ? 0 : (i >= G.n[a] ? G.n[a] - 1 : i);
        idx[a] = i;
        if (moving[a]) {
            step[a] = d[a] > 0.0 ? 1 : -1;
            const double face = G.lo[a] + (double) (i + (d[a] > 0.0 ? 1 : 0)) * G.cell[a];
            tn[a] = (face - o[a]) * inv[a];
            dt[a] = G.cell[a] * fabs(inv[a]);
        }
        else {
            step[a] = 0;
            tn[a] = INFINITY;
            dt[a] = 0.0;
        }
    }
    for (;;) {
        const int c = idx[0] + G.n[0] * (idx[1] + G.n[1] * idx[2]);
        const int b = RT_GLD(G.cell_start + c), e = RT_GLD(G.cell_start + c + 1);
        for (int k = b; k < e; k++)
            if (v.test(RT_GLD(G.items + k))) return;
        // leave through the nearest face
        int a = tn[0] <= tn[1] ? (tn[0] <= tn[2] ? 0 : 2) : (tn[1] <= tn[2] ? 1 : 2);
        const double t_out = a == 0 ? tn[0] : (a == 1 ? tn[1] : tn[2]);
        if (!(t_out <= t1) || v.bound() < t_out) return;
        if (a == 0) {
            idx[0] += step[0];
            if (idx[0] < 0 || idx[0] >= G.n[0]) return;
            tn[0] += dt[0];
        }
        else if (a == 1) {
            idx[1] += step[1];
            if (idx[1] < 0 || idx[1] >= G.n[1]) return;
            tn[1] += dt[1];
        }
        else {
            idx[2] += step[2];
            if (idx[2] < 0 || idx[2] >= G.n[2]) return;
            tn[2] += dt[2];
        }
    }
}

} // namespace rt
#endif
