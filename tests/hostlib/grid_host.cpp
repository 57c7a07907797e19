// Host build of csrc/rt_grid.cuh + rt_grid_build.h for tests/test_grid.py (g++, no CUDA): the walker and builder the
// kernels use, driven with the ORACLE's intersect_ray as the exact test, against brute force over the same function.
#include <cstddef>
#include <cstdint>
#include <vector>

#include "../../cuda-ray-tracer_b200/csrc/rt_grid_build.h"

typedef double (*intersect_fn)(const double coef[20], const double origin[3], const double dir[3]);

namespace {

const double EPS = 1e-7, MAX_T = 1e6;

struct Nearest {
    intersect_fn f;
    const double *coef, *o, *d;
    double best_t;
    int best_i;
    uint64_t tests;
    bool test(int i)
    {
        tests++;
        const double t = f(coef + 20 * i, o, d);
        if (t >= EPS && t < MAX_T && (t < best_t || (t == best_t && i < best_i))) {
            best_t = t;
            best_i = i;
        }
        return false;
    }
    double bound() const { return best_t; }
};

struct Any {
    intersect_fn f;
    const double *coef, *o, *d;
    double max_t;
    int found;
    uint64_t tests;
    bool test(int i)
    {
        tests++;
        const double t = f(coef + 20 * i, o, d);
        if (t > EPS && t < max_t) {
            found = i;
            return true;
        }
        return false;
    }
    double bound() const { return max_t; }
};

} // namespace

extern "C" {

// coef: n sphere-class objects (20 doubles each, x2 = y2 = z2 = 1).  Rays: o, d [m][3]; max_t[m] (any-hit limit).
// Outputs per ray: nearest index / t by brute force and through the grid, any-hit booleans both ways, whether the
// walker was allowed for the ray, and the number of sphere tests the grid path made.  info: nx, ny, nz, n_gridded,
// n_loose, n_items.
void gridtest_run(intersect_fn f, const double *coef, int n, const double *o, const double *d, const double *max_t, size_t m,
                  double cells_per_sphere, int *brute_i, double *brute_t, int *grid_i, double *grid_t, int *brute_any,
                  int *grid_any, int *walked, uint64_t *grid_tests, int *info)
{
    std::vector<double> centres(3 * (size_t) n), r2(n);
    for (int i = 0; i < n; i++) {
        const double *k = coef + 20 * i;
        const double cx = -0.5 * k[16], cy = -0.5 * k[17], cz = -0.5 * k[18];
        centres[3 * i] = cx;
        centres[3 * i + 1] = cy;
        centres[3 * i + 2] = cz;
        r2[i] = cx * cx + cy * cy + cz * cz - k[19];
    }
    rt::SphereGridHost H = rt::build_sphere_grid(centres.data(), r2.data(), n, cells_per_sphere);
    H.g.cell_start = H.cell_start.data();
    H.g.items = H.items.data();
    info[0] = H.g.n[0];
    info[1] = H.g.n[1];
    info[2] = H.g.n[2];
    info[3] = H.g.n_gridded;
    info[4] = H.g.loose_end - H.g.loose_begin;
    info[5] = (int) H.items.size();
    for (size_t r = 0; r < m; r++) {
        const double *ro = o + 3 * r, *rd = d + 3 * r;
        Nearest nb{f, coef, ro, rd, INFINITY, 0x7fffffff, 0};
        for (int i = 0; i < n; i++) nb.test(i);
        Any ab{f, coef, ro, rd, max_t[r], -1, 0};
        for (int i = 0; i < n && ab.found < 0; i++) ab.test(i);
        brute_i[r] = nb.best_t < INFINITY ? nb.best_i : -1;
        brute_t[r] = nb.best_t;
        brute_any[r] = ab.found >= 0;
        const bool ok = rt::grid_ray_ok(H.g, ro[0], ro[1], ro[2], rd[0], rd[1], rd[2]);
        walked[r] = ok;
        if (ok) {
            Nearest ng{f, coef, ro, rd, INFINITY, 0x7fffffff, 0};
            rt::grid_walk(H.g, ro[0], ro[1], ro[2], rd[0], rd[1], rd[2], MAX_T, ng);
            Any ag{f, coef, ro, rd, max_t[r], -1, 0};
            rt::grid_walk(H.g, ro[0], ro[1], ro[2], rd[0], rd[1], rd[2], max_t[r], ag);
            grid_i[r] = ng.best_t < INFINITY ? ng.best_i : -1;
            grid_t[r] = ng.best_t;
            grid_any[r] = ag.found >= 0;
            grid_tests[r] = ng.tests;
        }
        else {
            grid_i[r] = brute_i[r];
            grid_t[r] = brute_t[r];
            grid_any[r] = brute_any[r];
            grid_tests[r] = nb.tests;
        }
    }
}

} // extern "C"
