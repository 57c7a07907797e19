// Host build of csrc/rt_cubic_filter.cuh for tests/test_cubic_filter.py (g++, no CUDA): the same source the
// kernels compile, called on arrays of rays.
#include <cstddef>

#include "../../cuda-ray-tracer_b200/csrc/rt_cubic_filter.cuh"

extern "C" void rtf_cubic_verdicts(const double *k, const double *o, const double *d, size_t n, double hi, int want_hit,
                                   int *out)
{
    float kf[RT_CF_FLOATS];
    rt::make_cubicf(k, kf);
    for (size_t i = 0; i < n; i++) {
        const rt::RayC r = rt::make_rayc(o[3 * i], o[3 * i + 1], o[3 * i + 2], d[3 * i], d[3 * i + 1], d[3 * i + 2]);
        out[i] = rt::cubic_verdict(kf, r, (float) hi, want_hit != 0);
    }
}

// Per-ray interval end (the nearest-hit search asks about [EPS, best_t so far], rt_wave_device.cuh).
extern "C" void rtf_cubic_verdicts_hi(const double *k, const double *o, const double *d, size_t n, const float *hi,
                                      int want_hit, int *out)
{
    float kf[RT_CF_FLOATS];
    rt::make_cubicf(k, kf);
    for (size_t i = 0; i < n; i++) {
        const rt::RayC r = rt::make_rayc(o[3 * i], o[3 * i + 1], o[3 * i + 2], d[3 * i], d[3 * i + 1], d[3 * i + 2]);
        out[i] = rt::cubic_verdict(kf, r, hi[i], want_hit != 0);
    }
}
