"""CPU: the sphere grid (csrc/rt_grid.cuh walker + csrc/rt_grid_build.h builder, built for the host) proposes every
sphere the reference's brute-force loops would report.

The harness (tests/hostlib/grid_host.cpp) runs nearest-hit and any-hit searches twice per ray — brute force over all
spheres and through the grid — with the ORACLE's intersect_ray as the exact test in both, and the results must be
identical: same object index, same t bit for bit, same occlusion boolean.  Scenes: the synthetic C5 sphere cloud,
random clouds with touching / nested / duplicate / tiny / huge spheres, lattices whose sphere surfaces and ray paths
run exactly along cell faces and through cell corners."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
EPS, MAX_T = 1e-7, 1e6


def sphere_coef(c, r):
    """SurfaceCoefs of |p - c|^2 = r^2 in the reference's expanded form (surface.cpp:4-16)."""
    k = np.zeros(20)
    k[10:13] = 1.0
    k[16:19] = -2.0 * np.asarray(c, dtype=np.float64)
    k[19] = float(np.dot(c, c) - r * r)
    return k


@pytest.fixture(scope="module")
def grid(tmp_path_factory, oracle):
    so = str(tmp_path_factory.mktemp("hostlib") / "libgrid_host.so")
    subprocess.run(["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-shared", "-fPIC", "-o", so,
                    os.path.join(HERE, "hostlib", "grid_host.cpp")], check=True)
    lib = C.CDLL(so)
    olib = oracle.PortOracle.lib()
    fn = C.cast(olib.rto_intersect_ray, C.c_void_p)
    lib.gridtest_run.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_double] + \
        [C.c_void_p] * 9

    def run(coef, o, d, max_t, cells_per_sphere=0.0):
        coef = np.ascontiguousarray(coef, dtype=np.float64).reshape(-1, 20)
        o = np.ascontiguousarray(o, dtype=np.float64).reshape(-1, 3)
        d = np.ascontiguousarray(d, dtype=np.float64).reshape(-1, 3)
        m = len(o)
        max_t = np.ascontiguousarray(np.broadcast_to(np.asarray(max_t, dtype=np.float64), (m,)))
        out = {k: np.zeros(m, dtype=t) for k, t in (("brute_i", np.int32), ("brute_t", np.float64), ("grid_i", np.int32),
                                                    ("grid_t", np.float64), ("brute_any", np.int32), ("grid_any", np.int32),
                                                    ("walked", np.int32), ("tests", np.uint64))}
        info = np.zeros(6, dtype=np.int32)
        lib.gridtest_run(fn, coef.ctypes.data, len(coef), o.ctypes.data, d.ctypes.data, max_t.ctypes.data, m,
                         float(cells_per_sphere), *[out[k].ctypes.data for k in
                                                    ("brute_i", "brute_t", "grid_i", "grid_t", "brute_any", "grid_any", "walked", "tests")],
                         info.ctypes.data)
        out["info"] = dict(zip(("nx", "ny", "nz", "gridded", "loose", "items"), info.tolist()))
        return out
    return run


def assert_same(res, label):
    bad = res["brute_i"] != res["grid_i"]
    assert not bad.any(), (label, "nearest index differs", np.flatnonzero(bad)[:5], res["brute_i"][bad][:5], res["grid_i"][bad][:5])
    hit = res["brute_i"] >= 0
    assert (res["brute_t"][hit].view(np.uint64) == res["grid_t"][hit].view(np.uint64)).all(), (label, "t differs")
    bad = res["brute_any"] != res["grid_any"]
    assert not bad.any(), (label, "occlusion differs", np.flatnonzero(bad)[:5])


def random_dirs(rng, n):
    d = rng.normal(size=(n, 3))
    return d / np.linalg.norm(d, axis=1, keepdims=True)


def test_synthetic_cloud(pkg, grid):
    """C5's 4096 spheres: primary-like rays from the camera, shadow-like segments between random surface points and the
    point lights, random rays from inside the cloud."""
    from cuda_ray_tracer_b200 import synth
    sc = synth.synthetic_scene(width=64, height=36)
    ks = [np.array(o["coef"], dtype=np.float64).reshape(20) for o in sc.objects]
    spheres = np.array([k for k in ks if not k[:10].any() and (k[10:13] == 1).all() and not k[13:16].any()])
    assert len(spheres) == 4096
    rng = np.random.default_rng(1)
    n = 3000
    # camera rays (fov 60, looking down +z)
    d = np.stack([rng.uniform(-1.0, 1.0, n), rng.uniform(-0.6, 0.6, n), np.ones(n)], axis=1)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    res = grid(spheres, np.zeros((n, 3)), d, MAX_T)
    assert_same(res, "camera")
    assert res["walked"].all() and res["info"]["loose"] == 0
    assert (res["brute_i"] >= 0).mean() > 0.2
    assert res["tests"].mean() < 0.05 * len(spheres), res["tests"].mean()      # the point of the exercise
    # shadow segments: from a point on a random sphere (biased along the normal) to each point light, d = float(light - P)
    lights = np.array([l["p"] for l in sc.lights])
    i = rng.integers(0, len(spheres), n)
    c = -0.5 * spheres[i, 16:19]
    r = np.sqrt((c * c).sum(1) - spheres[i, 19])
    nrm = random_dirs(rng, n)
    p = c + r[:, None] * nrm
    so = p + 1e-2 * nrm
    sd = (lights[rng.integers(0, len(lights), n)] - p).astype(np.float32).astype(np.float64)
    res = grid(spheres, so, sd, 1.0)
    assert_same(res, "shadow")
    assert 0.02 < res["brute_any"].mean() < 0.98
    # rays from inside the cloud, un-normalised directions of very different lengths
    o = np.stack([rng.uniform(-60, 60, n), rng.uniform(-34, 34, n), rng.uniform(40, 160, n)], axis=1)
    d = random_dirs(rng, n) * (10.0 ** rng.uniform(-3, 3, n))[:, None]
    assert_same(grid(spheres, o, d, MAX_T), "inside")
    assert_same(grid(spheres, o, d, 1.0), "inside01")


@pytest.mark.parametrize("seed", range(6))
def test_random_clouds(grid, seed):
    """Random clouds with the awkward cases mixed in: duplicates (exact ties in t -> lowest index), nested and touching
    spheres, tiny and huge radii (loose list), spheres far from the cluster, rays starting inside spheres, rays that
    the walker must refuse (origin far away, direction shorter than the quadratic-branch threshold)."""
    rng = np.random.default_rng(100 + seed)
    n = int(rng.integers(1, 400))
    scale = 10.0 ** rng.uniform(-1, 3)
    c = rng.uniform(-1, 1, (n, 3)) * scale + rng.uniform(-1, 1, 3) * scale * rng.choice([0.0, 5.0])
    r = rng.uniform(0.01, 0.1, n) * scale
    if n > 8:
        c[1] = c[0]; r[1] = r[0]                       # duplicate: tie on every ray
        c[3] = c[2]; r[3] = 0.5 * r[2]                 # nested
        c[5] = c[4] + [r[4] + r[5], 0, 0]              # touching
        r[6] = 40 * scale                              # huge (contains everything): loose
        r[7] = 1e-7 * scale                            # tiny: its pad exceeds a cell -> loose, or never hit
        c[8] = c[8] + 50 * scale                       # outlier stretching the bounds
    coef = np.array([sphere_coef(c[i], r[i]) for i in range(n)])
    m = 1500
    o = rng.uniform(-1.5, 1.5, (m, 3)) * scale
    o[: m // 5] = c[rng.integers(0, n, m // 5)] + rng.normal(size=(m // 5, 3)) * 0.02 * scale    # inside / near spheres
    tgt = c[rng.integers(0, n, m)] + rng.normal(size=(m, 3)) * r[rng.integers(0, n, m), None]
    d = tgt - o
    d[::3] = random_dirs(rng, len(d[::3]))
    d[5::50] *= 1e-5                                   # below the reference's quadratic threshold -> brute force
    o[7::60] *= 1e4                                    # far origins -> brute force
    d[11::70, rng.integers(0, 3)] = 0.0                # axis-parallel
    for cps in (0.0, 0.2, 8.0):
        res = grid(coef, o, d, MAX_T, cps)
        assert_same(res, ("random", seed, cps))
        assert_same(grid(coef, o, d, 1.0, cps), ("random01", seed, cps))
    assert res["walked"].mean() > 0.8


def test_lattice_faces_and_corners(grid):
    """Unit spheres on an integer lattice; rays along lattice lines, through lattice points and along the planes where
    spheres touch: the walker's face / corner decisions are exercised with exact arithmetic ties."""
    g = np.arange(-4, 5, 2.0)
    c = np.array([[x, y, z] for x in g for y in g for z in g])
    coef = np.array([sphere_coef(ci, 1.0) for ci in c])           # neighbours touch
    rng = np.random.default_rng(7)
    o, d = [], []
    for _ in range(600):
        p = rng.integers(-5, 6, 3).astype(np.float64)
        q = rng.integers(-5, 6, 3).astype(np.float64)
        if (p == q).all():
            continue
        o.append(p - 3.0 * (q - p))
        d.append(q - p)
        o.append(p + [0.5, 0.0, 0.0])
        d.append((q - p) * [0.0, 1.0, 1.0] + [0.0, 0.0, 1e-30])
    o, d = np.array(o), np.array(d)
    ok = np.abs(d).sum(1) > 0
    for cps in (0.0, 1.0, 5.0):
        assert_same(grid(coef, o[ok], d[ok], MAX_T, cps), ("lattice", cps))
        assert_same(grid(coef, o[ok], d[ok], 1.0, cps), ("lattice01", cps))


def test_degenerate_inputs(grid):
    """No spheres in cells (all loose), a single sphere, non-positive r^2, NaN / inf rays."""
    one = np.array([sphere_coef([0.0, 0.0, 5.0], 1.0)])
    bad = np.array([sphere_coef([0.0, 0.0, 5.0], 1.0), sphere_coef([1.0, 0.0, 5.0], 0.0)])
    bad[1, 19] += 1.0                                              # r^2 = -1
    o = np.zeros((6, 3))
    d = np.array([[0, 0, 1.0], [0, 1.0, 0], [np.nan, 0, 1], [np.inf, 0, 1], [0, 0, 1e-9], [1e-3, 0, 1.0]])
    for coef in (one, bad):
        res = grid(coef, o, d, MAX_T)
        assert_same(res, "degenerate")
        assert res["brute_i"][0] == 0
    assert grid(bad, o, d, MAX_T)["info"]["loose"] == 1
