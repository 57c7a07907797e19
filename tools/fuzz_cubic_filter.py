"""CPU fuzzing of the FP32 cubic filter (csrc/rt_cubic_filter.cuh, host build) against the oracle's exact
intersect_ray: random and adversarial surfaces x rays; a verdict must never contradict the exact result.
    python tools/fuzz_cubic_filter.py [n_rounds] [seed]"""
import ctypes as C
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools"))
from loader import load_package
pkg = load_package()
from oracle import oracle_py
from cuda_ray_tracer_b200 import scene as S

EPS, MAX_T = 1e-7, 1e6
oracle_py.build()
so = os.path.join(tempfile.mkdtemp(), "libcf.so")
subprocess.run(["g++", "-O2", "-std=c++17", "-shared", "-fPIC", "-o", so, os.path.join(ROOT, "tests", "hostlib", "cubic_filter_host.cpp")], check=True)
lib = C.CDLL(so)
olib = oracle_py.PortOracle.lib()


def verdicts(k, o, d, hi):
    out = np.zeros(len(o), dtype=np.int32)
    lib.rtf_cubic_verdicts(C.c_void_p(k.ctypes.data), C.c_void_p(o.ctypes.data), C.c_void_p(d.ctypes.data), C.c_size_t(len(o)),
                           C.c_double(hi), C.c_int(1), C.c_void_p(out.ctypes.data))
    return out


def exact(k, o, d):
    kp = k.ctypes.data_as(C.POINTER(C.c_double))
    f = olib.rto_intersect_ray
    out = np.empty(len(o))
    for i in range(len(o)):
        out[i] = f(kp, o[i].ctypes.data_as(C.POINTER(C.c_double)), d[i].ctypes.data_as(C.POINTER(C.c_double)))
    return out


def surfaces(rng):
    out = [S.surface_cayley(), S.surface_clebsch(), S.surface_ding_dong([0.0, 0.0, 0.0]),
           S.surface_polynomial({"x3": 1, "y3": 1, "z3": 1, "c": -1}), S.surface_polynomial({"x3": 1, "xy2": -3, "z": -1})]
    for _ in range(6):      # dense random
        out.append(rng.normal(size=20) * 10.0 ** rng.uniform(-2, 2))
    for _ in range(6):      # sparse random, mixed magnitudes
        k = np.zeros(20)
        idx = rng.choice(20, size=rng.integers(2, 7), replace=False)
        k[idx] = rng.normal(size=len(idx)) * 10.0 ** rng.uniform(-3, 3, size=len(idx))
        if not k[:10].any():
            k[rng.integers(0, 10)] = rng.normal()
        out.append(k)
    for _ in range(4):      # tiny cubic part on top of a quadric: the degree switch is in play
        k = np.zeros(20)
        k[10:13] = 1.0
        k[19] = -rng.uniform(0.5, 4.0)
        k[rng.integers(0, 10)] = 10.0 ** rng.uniform(-9, -5)
        out.append(k)
    return [np.ascontiguousarray(k, dtype=np.float64) for k in out]


def rays(rng, k, n):
    scale = 10.0 ** rng.uniform(-1, 3)
    o = rng.uniform(-1, 1, size=(n, 3)) * scale
    d = rng.normal(size=(n, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    kind = rng.integers(0, 6, size=n)
    axis = np.eye(3)[rng.integers(0, 3, size=n)] * rng.choice([-1.0, 1.0], size=(n, 1))
    d[kind == 1] = axis[kind == 1]                                          # axis-aligned
    d[kind == 2] = axis[kind == 2] + 1e-9 * rng.normal(size=(int((kind == 2).sum()), 3))   # almost axis-aligned
    d[kind == 3] *= 100.0                                                    # long, like point-light rays
    d[kind == 4] *= 1e-3
    d = d.astype(np.float32).astype(np.float64) if rng.random() < 0.5 else d  # shadow directions are float-rounded
    # a third of the rays start just off the surface, like shadow rays
    t = exact(k, np.ascontiguousarray(o), np.ascontiguousarray(d))
    on = (t > EPS) & (t < MAX_T) & (rng.random(n) < 0.5)
    p = o[on] + t[on, None] * d[on]
    nrm = np.array([oracle_py.PortOracle.normal_vector(k, q) for q in p]) if on.any() else np.zeros((0, 3))
    o[on] = p + 1e-2 * nrm
    d[on] = rng.normal(size=(int(on.sum()), 3)) * 10.0 ** rng.uniform(-1, 1)
    good = np.isfinite(o).all(axis=1) & np.isfinite(d).all(axis=1)
    return np.ascontiguousarray(o[good]), np.ascontiguousarray(d[good])


def main():
    rounds = int(sys.argv[1]) if len(sys.argv) > 1 else 20
    rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 1)
    total = decided = 0
    for r in range(rounds):
        for k in surfaces(rng):
            o, d = rays(rng, k, 400)
            t = exact(k, o, d)
            for hi in (1.0, MAX_T):
                v = verdicts(k, o, d, hi)
                inside = (t > EPS) & (t < hi)
                bad = ((v == 1) & inside) | ((v == 2) & ~inside)
                if bad.any():
                    i = int(np.where(bad)[0][0])
                    print("WRONG VERDICT", v[i], "exact t", t[i], "hi", hi, "\nk =", repr(k), "\no =", repr(o[i]), "\nd =", repr(d[i]))
                    sys.exit(1)
                total += len(v)
                decided += int((v != 0).sum())
        print("round", r, "rays", total, "decided", round(decided / total, 4), flush=True)
    print("FUZZ OK", total, "verdicts, none wrong")


main()
