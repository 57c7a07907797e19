"""CPU: the FP32 cubic filter (csrc/rt_cubic_filter.cuh, built for the host) never contradicts the exact test.

For primary rays, shadow rays and random rays against every cubic-class surface of the bundled scenes and of the
synthetic scene: a MISS verdict implies the oracle's intersect_ray returns no t in [EPS, hi), a HIT verdict implies
EPS < t < hi; and the filter decides the bulk of the rays (otherwise it would be useless, not wrong)."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
EPS, MAX_T, BIAS = 1e-7, 1e6, 1e-2
CUBIC_SCENES = ["cubic", "cayley", "clebsch", "dingdong", "monkey_saddle"]


@pytest.fixture(scope="module")
def filt(tmp_path_factory):
    so = str(tmp_path_factory.mktemp("hostlib") / "libcubic_filter_host.so")
    subprocess.run(["g++", "-O2", "-std=c++17", "-shared", "-fPIC", "-o", so,
                    os.path.join(HERE, "hostlib", "cubic_filter_host.cpp")], check=True)
    lib = C.CDLL(so)
    lib.rtf_cubic_verdicts.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_double, C.c_int, C.c_void_p]

    lib.rtf_cubic_verdicts_hi.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p, C.c_int, C.c_void_p]

    def verdicts(k, o, d, hi, want_hit):
        """hi: a scalar, or one float32 per ray (the running best t of a nearest-hit search)."""
        k = np.ascontiguousarray(k, dtype=np.float64)
        o = np.ascontiguousarray(o, dtype=np.float64).reshape(-1, 3)
        d = np.ascontiguousarray(d, dtype=np.float64).reshape(-1, 3)
        out = np.zeros(len(o), dtype=np.int32)
        if np.ndim(hi) == 0:
            lib.rtf_cubic_verdicts(k.ctypes.data, o.ctypes.data, d.ctypes.data, len(o), hi, int(want_hit), out.ctypes.data)
        else:
            hi = np.ascontiguousarray(hi, dtype=np.float32)
            assert len(hi) == len(o)
            lib.rtf_cubic_verdicts_hi(k.ctypes.data, o.ctypes.data, d.ctypes.data, len(o), hi.ctypes.data, int(want_hit), out.ctypes.data)
        return out
    return verdicts


def exact_ts(oracle, k, o, d):
    return np.array([oracle.PortOracle.intersect_ray(k, o[i], d[i]) for i in range(len(o))])


def check(verd, t, hi, label, stats):
    inside = (t > EPS) & (t < hi)             # NaN -> False, as in the reference's comparisons
    wrong_miss = (verd == 1) & inside
    wrong_hit = (verd == 2) & ~inside
    assert not wrong_miss.any(), (label, "MISS but exact t =", t[wrong_miss][:5])
    assert not wrong_hit.any(), (label, "HIT but exact t =", t[wrong_hit][:5])
    # usefulness: share of the decidable rays that got a verdict (primary queries can only use MISS)
    decidable = ~inside if label[-1] == "primary" else np.ones_like(inside)
    stats.append((label, int((verd[decidable] != 0).sum()), int(decidable.sum())))


def primary_rays(sc, m, nx, ny):
    m = np.asarray(m, dtype=np.float64).reshape(4, 4).T          # column-major -> [row, col]
    origin = m[:3, 3].copy()
    aspect = sc.px_width / sc.px_height
    th = np.tan(0.5 * sc.vertical_fov)
    xs = (np.arange(nx) + 0.5) / nx
    ys = (np.arange(ny) + 0.5) / ny
    gx, gy = np.meshgrid(xs, ys)
    cam = np.stack([(2 * gx - 1) * aspect * th, (2 * gy - 1) * th, np.ones_like(gx), np.ones_like(gx)], axis=-1).reshape(-1, 4)
    p = cam @ m.T
    d = p[:, :3] - origin
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    return np.broadcast_to(origin, d.shape).copy(), d


def cubic_objects(sc):
    return [np.array(o["coef"], dtype=np.float64).reshape(20) for o in sc.objects
            if np.any(np.array(o["coef"], dtype=np.float64).reshape(20)[:10] != 0.0)]


@pytest.mark.parametrize("name", CUBIC_SCENES)
def test_filter_agrees_with_exact_on_scene_rays(pkg, oracle, scenes_dir, filt, name):
    from cuda_ray_tracer_b200 import camera
    sc = pkg.load_from_file(os.path.join(scenes_dir, name + ".yml"))
    stats = []
    for pose in ("initial", "pose_a", "pose_b"):
        o, d = primary_rays(sc, camera.TEST_POSES[pose], 64, 36)
        for k in cubic_objects(sc):
            t = exact_ts(oracle, k, o, d)
            check(filt(k, o, d, MAX_T, False), t, MAX_T, (name, pose, "primary"), stats)
            hit = (t >= EPS) & (t < MAX_T)
            if not hit.any():
                continue
            p = o[hit] + t[hit, None] * d[hit]
            n = np.array([oracle.PortOracle.normal_vector(k, q) for q in p])
            so = p + BIAS * n
            for l in sc.lights:
                if l["is_spherical"]:
                    sd = (l["p"][None, :] - p).astype(np.float32).astype(np.float64)   # light_impl.h:21: float-rounded
                    hi = 1.0
                else:
                    sd = np.broadcast_to(np.asarray(l["p"], dtype=np.float32).astype(np.float64), p.shape).copy()
                    hi = MAX_T
                ts = exact_ts(oracle, k, so, sd)
                check(filt(k, so, sd, hi, True), ts, hi, (name, pose, "shadow"), stats)
    for kind in ("primary", "shadow"):
        got = sum(a for l, a, n in stats if l[-1] == kind)
        of = sum(n for l, a, n in stats if l[-1] == kind)
        print(name, kind, "decided", got, "of", of)
        assert of == 0 or got > 0.6 * of, stats


def test_filter_on_random_rays_and_transformed_cubics(pkg, oracle, filt):
    from cuda_ray_tracer_b200 import synth
    rng = np.random.default_rng(20)
    sc = synth.synthetic_scene(width=64, height=36)
    ks = cubic_objects(sc)[:12]
    assert ks
    stats = []
    for k in ks:
        o = rng.uniform(-60, 60, size=(400, 3)) + np.array([0.0, 0.0, 150.0])
        d = rng.normal(size=(400, 3))
        d[:200] /= np.linalg.norm(d[:200], axis=1, keepdims=True)    # the rest stays unnormalised, like point-light rays
        t = exact_ts(oracle, k, o, d)
        for hi in (1.0, MAX_T):
            check(filt(k, o, d, hi, True), t, hi, ("random", hi), stats)
    # degenerate inputs end in UNKNOWN, never in a verdict
    k = ks[0]
    bad = np.array([[np.inf, 0, 0], [1e30, 1e30, 1e30], [np.nan, 0, 0], [0, 0, 0]], dtype=np.float64)
    dirs = np.array([[0, 0, 1], [1, 1, 1], [0, 0, 1], [0, 0, 0]], dtype=np.float64)
    assert not filt(k, bad, dirs, MAX_T, True)[:3].any()


def float32_round_up(x):
    """Smallest float32 >= x (what __double2float_ru gives the kernel)."""
    f = np.asarray(x, dtype=np.float64).astype(np.float32)
    low = f.astype(np.float64) < x
    f[low] = np.nextafter(f[low], np.float32(np.inf))
    return f


def test_filter_with_running_nearest_bound(pkg, oracle, filt, scenes_dir):
    """The nearest-hit search asks the filter about [EPS, best t so far] (rt_wave_device.cuh): a MISS verdict for an
    interval end hi must mean the exact t is not in [EPS, hi] — hi itself included, because an exact tie in t is
    resolved by object index and must reach the exact test.  Interval ends: random magnitudes, the exact root rounded
    up to FP32 (the tie), and values a few ulps either side of a root."""
    from cuda_ray_tracer_b200 import camera, synth
    rng = np.random.default_rng(77)
    surfaces = cubic_objects(synth.synthetic_scene(width=64, height=36))[:9]
    for name in CUBIC_SCENES:
        surfaces += cubic_objects(pkg.load_from_file(os.path.join(scenes_dir, name + ".yml")))
    pruned = total = 0
    for k in surfaces:
        o = np.concatenate([np.zeros((300, 3)), rng.uniform(-30, 30, size=(300, 3)) + np.array([0.0, 0.0, 60.0])])
        d = rng.normal(size=(600, 3)) * np.array([1.0, 0.6, 1.0]) + np.array([0.0, 0.0, 1.5])
        d /= np.linalg.norm(d, axis=1, keepdims=True)
        d[450:] *= (10.0 ** rng.uniform(-2, 2, 150))[:, None]
        t = exact_ts(oracle, k, o, d)
        has_t = np.isfinite(t) & (t >= EPS) & (t < MAX_T)
        tt = np.where(has_t, t, 1.0)
        his = [float32_round_up(np.minimum(MAX_T, 10.0 ** rng.uniform(-6.5, 6, 600))),      # anything
               float32_round_up(np.minimum(MAX_T, tt)),                                      # the tie: hi = ru(t)
               np.nextafter(float32_round_up(tt), np.float32(0)),                            # one FP32 ulp below ru(t)
               float32_round_up(np.minimum(MAX_T, tt * (1.0 - 10.0 ** rng.uniform(-7, -1, 600)))),   # just before the root
               float32_round_up(np.minimum(MAX_T, tt * (1.0 + 10.0 ** rng.uniform(-7, -1, 600))))]   # just past it
        for hi in his:
            hi = np.minimum(hi, np.float32(MAX_T))
            v = filt(k, o, d, hi, False)
            inside = (t >= EPS) & (t <= hi.astype(np.float64)) & (t < MAX_T)
            wrong = (v == 1) & inside
            assert not wrong.any(), ("MISS but exact t within the bound", t[wrong][:4], hi[wrong][:4])
            assert not (v == 2).any()                    # want_hit = False never answers HIT
            pruned += int(((v == 1) & ~inside).sum())
            total += int((~inside).sum())
    print("pairs outside the bound proven so:", pruned, "of", total)
    assert pruned > 0.5 * total
