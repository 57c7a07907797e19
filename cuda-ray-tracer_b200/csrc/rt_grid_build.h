// rt_grid_build.h — host-side construction of the sphere grid (rt_grid.cuh holds the walker and the argument why
// the padding below cannot lose a hit).  Plain C++ so that rt_api.cu and the CPU test harness
// (tests/hostlib/grid_host.cpp) build the same structure.
#ifndef RT_GRID_BUILD_H
#define RT_GRID_BUILD_H

#include <algorithm>
#include <cmath>
#include <vector>

#include "rt_grid.cuh"

namespace rt {

struct SphereGridHost {
    SphereGrid g{};               // pointers left null: the caller points them at its copy of the two arrays
    std::vector<int> cell_start;  // [cells + 1]
    std::vector<int> items;       // cell lists, then the loose list
};

// centres[3 i .. 3 i + 2], r2[i] = squared radius (may be <= 0 or non-finite: such spheres go to the loose list).
// cells_per_sphere: target ratio cells / gridded spheres (<= 0: default 1.5).  max_dim bounds each axis.
inline SphereGridHost build_sphere_grid(const double *centres, const double *r2, int n, double cells_per_sphere = 0.0,
                                        int max_dim = 128)
{
    SphereGridHost H;
    SphereGrid &G = H.g;
    G.cell_start = nullptr;
    G.items = nullptr;
    G.n[0] = G.n[1] = G.n[2] = 0;
    G.n_gridded = 0;
    G.o_limit = 0.0;
    for (int a = 0; a < 3; a++) G.lo[a] = G.hi[a] = G.cell[a] = G.inv_cell[a] = 0.0;
    if (cells_per_sphere <= 0.0) cells_per_sphere = 1.5;

    std::vector<double> rad(n, 0.0);
    std::vector<char> loose(n, 0);
    std::vector<double> finite_r;
    for (int i = 0; i < n; i++) {
        const double *c = centres + 3 * i;
        const bool ok = std::isfinite(c[0]) && std::isfinite(c[1]) && std::isfinite(c[2]) && std::isfinite(r2[i]) &&
                        r2[i] > 0.0 && std::fabs(c[0]) < 1e9 && std::fabs(c[1]) < 1e9 && std::fabs(c[2]) < 1e9 && r2[i] < 1e18;
        if (!ok) {
            loose[i] = 1;
            continue;
        }
        rad[i] = std::sqrt(r2[i]);
        finite_r.push_back(rad[i]);
    }
    // spheres much larger than the typical one would each fill thousands of cells: keep them out of the grid
    if (!finite_r.empty()) {
        std::nth_element(finite_r.begin(), finite_r.begin() + finite_r.size() / 2, finite_r.end());
        const double r_med = finite_r[finite_r.size() / 2];
        for (int i = 0; i < n; i++)
            if (!loose[i] && rad[i] > 16.0 * r_med) loose[i] = 1;
    }

    // two passes: the padding depends on the scene scale S, which depends on which spheres are gridded
    double S = 1.0;
    double cell = 0.0;
    for (int pass = 0; pass < 2; pass++) {
        double lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
        int m = 0;
        double r_sum = 0.0;
        for (int i = 0; i < n; i++) {
            if (loose[i]) continue;
            m++;
            r_sum += rad[i];
            for (int a = 0; a < 3; a++) {
                lo[a] = std::min(lo[a], centres[3 * i + a] - rad[i]);
                hi[a] = std::max(hi[a], centres[3 * i + a] + rad[i]);
            }
        }
        if (m == 0) break;
        S = 1.0;
        for (int a = 0; a < 3; a++) S = std::max(S, std::max(std::fabs(lo[a]), std::fabs(hi[a])));
        S *= 1.01;
        double ext[3], vol = 1.0;
        const double r_mean = r_sum / m;
        for (int a = 0; a < 3; a++) {
            ext[a] = std::max(hi[a] - lo[a], 2.0 * r_mean);
            vol *= ext[a];
        }
        cell = std::cbrt(vol / (cells_per_sphere * m));
        cell = std::max(cell, r_mean);          // cells smaller than the spheres only multiply the registrations
        for (int a = 0; a < 3; a++) cell = std::max(cell, ext[a] / max_dim);
        const double Mmax = 30.0 * S;
        bool changed = false;
        for (int i = 0; i < n; i++) {
            if (loose[i]) continue;
            const double pad = 1e-6 * S + 1e-10 * Mmax * Mmax / rad[i];
            if (!(pad < 0.5 * cell)) {
                loose[i] = 1;
                changed = true;
            }
        }
        if (!changed) break;
    }

    std::vector<int> gridded;
    for (int i = 0; i < n; i++)
        if (!loose[i]) gridded.push_back(i);
    G.n_gridded = (int) gridded.size();
    if (!gridded.empty()) {
        const double Mmax = 30.0 * S;
        std::vector<double> blo(3 * gridded.size()), bhi(3 * gridded.size());
        double lo[3] = {INFINITY, INFINITY, INFINITY}, hi[3] = {-INFINITY, -INFINITY, -INFINITY};
        for (size_t k = 0; k < gridded.size(); k++) {
            const int i = gridded[k];
            const double pad = 1e-6 * S + 1e-10 * Mmax * Mmax / rad[i];
            for (int a = 0; a < 3; a++) {
                blo[3 * k + a] = centres[3 * i + a] - rad[i] - pad;
                bhi[3 * k + a] = centres[3 * i + a] + rad[i] + pad;
                lo[a] = std::min(lo[a], blo[3 * k + a]);
                hi[a] = std::max(hi[a], bhi[3 * k + a]);
            }
        }
        for (int a = 0; a < 3; a++) {
            lo[a] -= 1e-3 * cell;
            hi[a] += 1e-3 * cell;
            int na = (int) std::ceil((hi[a] - lo[a]) / cell);
            na = std::max(1, std::min(max_dim, na));
            G.n[a] = na;
            G.lo[a] = lo[a];
            G.hi[a] = hi[a];
            G.cell[a] = (hi[a] - lo[a]) / na;
            G.inv_cell[a] = 1.0 / G.cell[a];
        }
        G.o_limit = 16.0 * S;
        const size_t cells = (size_t) G.n[0] * G.n[1] * G.n[2];
        auto range = [&](size_t k, int a, int &i0, int &i1) {
            // one extra cell either side absorbs the rounding of this very computation
            i0 = (int) std::floor((blo[3 * k + a] - G.lo[a]) * G.inv_cell[a] - 1e-9);
            i1 = (int) std::floor((bhi[3 * k + a] - G.lo[a]) * G.inv_cell[a] + 1e-9);
            i0 = std::max(0, std::min(G.n[a] - 1, i0));
            i1 = std::max(0, std::min(G.n[a] - 1, i1));
        };
        H.cell_start.assign(cells + 1, 0);
        for (int fill = 0; fill < 2; fill++) {
            for (size_t k = 0; k < gridded.size(); k++) {
                int x0, x1, y0, y1, z0, z1;
                range(k, 0, x0, x1);
                range(k, 1, y0, y1);
                range(k, 2, z0, z1);
                for (int z = z0; z <= z1; z++)
                    for (int y = y0; y <= y1; y++)
                        for (int x = x0; x <= x1; x++) {
                            const size_t c = x + (size_t) G.n[0] * (y + (size_t) G.n[1] * z);
                            if (fill == 0) H.cell_start[c + 1]++;
                            else H.items[H.cell_start[c]++] = gridded[k];
                        }
            }
            if (fill == 0) {
                for (size_t c = 0; c < cells; c++) H.cell_start[c + 1] += H.cell_start[c];
                H.items.resize(H.cell_start[cells]);
            }
            else {   // the fill advanced every start to its end: shift back
                for (size_t c = cells; c > 0; c--) H.cell_start[c] = H.cell_start[c - 1];
                H.cell_start[0] = 0;
            }
        }
    }
    else {
        H.cell_start.assign(1, 0);
    }
    G.loose_begin = (int) H.items.size();
    for (int i = 0; i < n; i++)
        if (loose[i]) H.items.push_back(i);
    G.loose_end = (int) H.items.size();
    return H;
}

} // namespace rt
#endif
