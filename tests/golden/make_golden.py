#!/usr/bin/env python
"""Mint the golden fixtures from the reference itself (oracle/_ref/libref_cpu.so = the reference's
update-cpu.cpp compiled unmodified).  Run in the build container, where /root/reference exists:

    make -C oracle && python tests/golden/make_golden.py

Writes (all small, committed):
  frames_160x120.npz   RGBA8 frames of the 8 bundled scenes at 160x120, max_reflections 4, 3 cameras
  golden.json          FNV-1a-64 of the native-resolution identity-camera frames + non-background fraction
  rays.npz             known-answer vectors for intersect_ray / normal_vector on random rays
The reference has no tests or golden vectors of its own; these pin the oracle port and the
GPU path to the reference's actual output.
"""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
from loader import load_package  # noqa: E402
from oracle import oracle_py  # noqa: E402

pkg = load_package()
from cuda_ray_tracer_b200 import camera  # noqa: E402

SCENES = ["20spheres", "quadratic", "cubic", "cayley", "clebsch", "dingdong", "monkey_saddle", "reflection_test"]
POSES = ["identity", "pose_a", "pose_b"]
SMALL = (160, 120)


def main():
    assert oracle_py.have_ref(), "build oracle/_ref first"
    frames = {}
    meta = {"scenes": {}, "small": {"width": SMALL[0], "height": SMALL[1], "max_reflections": 4, "poses": POSES}}
    for name in SCENES:
        sc = pkg.load_from_file(os.path.join(ROOT, "scenes", name + ".yml"))
        ref = oracle_py.RefCpu(sc)
        rgb, _ = ref.update(camera.IDENTITY)
        q = oracle_py.quantize_rgba8(rgb)
        bgq = oracle_py.quantize_rgba8(sc.bg_color.reshape(1, 3))[0]
        meta["scenes"][name] = {
            "width": sc.px_width, "height": sc.px_height,
            "fnv1a64_rgb_identity": oracle_py.fnv1a64(q[..., :3].tobytes()),
            "non_background_frac": float((q != bgq).any(axis=-1).mean()),
        }
        small = sc.with_overrides(width=SMALL[0], height=SMALL[1], max_reflections=4)
        ref = oracle_py.RefCpu(small)
        for pose in POSES:
            rgb, _ = ref.update(camera.TEST_POSES[pose])
            frames["{}/{}".format(name, pose)] = oracle_py.quantize_rgba8(rgb)
        print(name, meta["scenes"][name])
    np.savez_compressed(os.path.join(HERE, "frames_160x120.npz"), **frames)

    # known-answer rays: every bundled surface + random polynomials of each structural class
    rng = np.random.default_rng(0xB200)
    coefs = []
    for name in SCENES:
        sc = pkg.load_from_file(os.path.join(ROOT, "scenes", name + ".yml"))
        coefs.extend(sc.objects["coef"][:3])
    for _ in range(8):
        k = np.zeros(20); k[16:] = rng.normal(size=4); coefs.append(k)                       # planes
        k = np.zeros(20); k[10:13] = 1.0; k[16:] = rng.normal(size=4) * 3; coefs.append(k)   # spheres
        k = np.zeros(20); k[10:] = rng.normal(size=10); coefs.append(k)                      # quadrics
        k = rng.normal(size=20); coefs.append(k)                                             # cubics
    coefs = np.array(coefs)
    n_rays = 64
    origins = rng.normal(size=(n_rays, 3)) * 4.0
    dirs = rng.normal(size=(n_rays, 3))
    dirs[: n_rays // 2] /= np.linalg.norm(dirs[: n_rays // 2], axis=1, keepdims=True)  # half unit, half not
    t = np.empty((len(coefs), n_rays))
    normals = np.empty((len(coefs), n_rays, 3))
    for i, k in enumerate(coefs):
        for j in range(n_rays):
            t[i, j] = oracle_py.RefCpu.intersect_ray(k, origins[j], dirs[j])
            normals[i, j] = oracle_py.RefCpu.normal_vector(k, origins[j])
    np.savez_compressed(os.path.join(HERE, "rays.npz"), coefs=coefs, origins=origins, dirs=dirs, t=t, normals=normals)

    with open(os.path.join(HERE, "golden.json"), "w") as f:
        json.dump(meta, f, indent=1, sort_keys=True)
    print("wrote fixtures to", HERE)


if __name__ == "__main__":
    main()
