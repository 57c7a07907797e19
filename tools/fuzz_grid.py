"""CPU fuzzing of the sphere grid (csrc/rt_grid.cuh + rt_grid_build.h, host build) against brute force over the oracle's
intersect_ray: random clouds at scales 1e-2 .. 1e4, rays aimed tangentially at sphere surfaces (offset by 0, +-1e-9,
+-1e-6 of the radius), un-normalised directions, nearest-hit and any-hit (max_t 1e6 and 1), two grid densities.
    python tools/fuzz_grid.py [n_seeds] [first_seed]
Round 1: 120 seeds, 576 000 rays x 4 queries, no mismatch."""
import ctypes as C
import os
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from oracle import oracle_py

oracle_py.build()
so = os.path.join(tempfile.mkdtemp(), "libgrid_host.so")
subprocess.run(["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-shared", "-fPIC", "-o", so,
                os.path.join(ROOT, "tests", "hostlib", "grid_host.cpp")], check=True)
lib = C.CDLL(so)
fn = C.cast(oracle_py.PortOracle.lib().rto_intersect_ray, C.c_void_p)
lib.gridtest_run.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_double] + [C.c_void_p] * 9


def run(coef, o, d, max_t, cps):
    coef = np.ascontiguousarray(coef, dtype=np.float64).reshape(-1, 20)
    o, d = np.ascontiguousarray(o), np.ascontiguousarray(d)
    m = len(o)
    mt = np.ascontiguousarray(np.broadcast_to(np.asarray(max_t, dtype=np.float64), (m,)))
    outs = [np.zeros(m, dtype=t) for t in (np.int32, np.float64, np.int32, np.float64, np.int32, np.int32, np.int32, np.uint64)]
    info = np.zeros(6, dtype=np.int32)
    lib.gridtest_run(fn, coef.ctypes.data, len(coef), o.ctypes.data, d.ctypes.data, mt.ctypes.data, m, float(cps),
                     *[x.ctypes.data for x in outs], info.ctypes.data)
    return outs


def sphere_coef(c, r):
    k = np.zeros(20)
    k[10:13] = 1.0
    k[16:19] = -2.0 * np.asarray(c)
    k[19] = float(np.dot(c, c) - r * r)
    return k


n_seeds = int(sys.argv[1]) if len(sys.argv) > 1 else 120
first = int(sys.argv[2]) if len(sys.argv) > 2 else 1000
bad = rays = 0
for seed in range(first, first + n_seeds):
    rng = np.random.default_rng(seed)
    n = int(rng.integers(1, 600))
    scale = 10.0 ** rng.uniform(-2, 4)
    c = rng.uniform(-1, 1, (n, 3)) * scale * rng.uniform(0.1, 1, 3) + rng.uniform(-1, 1, 3) * scale * rng.choice([0.0, 3.0, 30.0])
    r = rng.uniform(0.005, 0.2, n) * scale * 10.0 ** rng.uniform(-1, 0)
    m = 1200
    i = rng.integers(0, n, m)
    nrm = rng.normal(size=(m, 3))
    nrm /= np.linalg.norm(nrm, axis=1, keepdims=True)
    p = c[i] + r[i, None] * nrm * (1.0 + rng.choice([0.0, 1e-9, -1e-9, 1e-6, -1e-6], m))[:, None]
    tang = np.cross(nrm, rng.normal(size=(m, 3)))
    tang /= np.linalg.norm(tang, axis=1, keepdims=True)
    o = p - tang * rng.uniform(0.1, 3, m)[:, None] * scale
    d = tang * (10.0 ** rng.uniform(-2, 2, m))[:, None]
    half = m // 2
    o[half:] = rng.uniform(-1.5, 1.5, (m - half, 3)) * scale
    d[half:] = rng.normal(size=(m - half, 3))
    coef = np.array([sphere_coef(c[j], r[j]) for j in range(n)])
    for cps in (0.0, 6.0):
        for mt in (1e6, 1.0):
            bi, bt, gi, gt, ba, ga, wk, ts = run(coef, o, d, mt, cps)
            hit = bi >= 0
            ok = (bi == gi).all() and (bt[hit].view(np.uint64) == gt[hit].view(np.uint64)).all() and (ba == ga).all()
            if not ok:
                bad += 1
                print("MISMATCH seed", seed, "density", cps, "max_t", mt, np.flatnonzero(bi != gi)[:5], np.flatnonzero(ba != ga)[:5], flush=True)
    rays += m
print("GRID FUZZ", "OK" if bad == 0 else "FAILED", n_seeds, "seeds,", rays, "rays x 4 queries,", bad, "mismatching cases")
sys.exit(1 if bad else 0)
