"""Synthetic scaled scene of BASELINE.json configs[4] / SURVEY.md §8(d) "C5":
4096 spheres + 64 cubic surfaces + 16 point lights at 7680x4320, fov 60, max_reflections 4.

Deterministic (SplitMix64, seed 0xB200).  Every number is rounded to 6 significant digits BEFORE it is
used, so the YAML text this module can emit (`to_yaml`) is the single source of truth: loading that YAML
with either loader gives the same records as `synthetic_scene()`.

Cubic surfaces are the ding-dong (x^2 + z^2 = y^2 (1 - y), cf. src/surface.cpp:27-39), Fermat
(x^3 + y^3 + z^3 = 1, scenes/cubic.yml) and monkey-saddle (x^3 - 3 x z^2 + 10 y + 5, scenes/monkey_saddle.yml)
polynomials under a translation + uniform scale, expanded into the 20 `type: polynomial` coefficients.
"""
from __future__ import annotations

import itertools

import numpy as np

from . import scene as S

SEED = 0xB200
MASK = (1 << 64) - 1


class SplitMix64:
    def __init__(self, seed: int):
        self.s = seed & MASK

    def next_u64(self) -> int:
        self.s = (self.s + 0x9E3779B97F4A7C15) & MASK
        z = self.s
        z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & MASK
        z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & MASK
        return z ^ (z >> 31)

    def uniform(self, lo: float, hi: float) -> float:
        return lo + (hi - lo) * ((self.next_u64() >> 11) * (1.0 / (1 << 53)))


def sig6(v: float) -> float:
    return float("{:.6g}".format(v))


# ---- trivariate polynomials of degree <= 3 as {(i, j, k): coefficient} -------------------------------
_MONO = {"x3": (3, 0, 0), "y3": (0, 3, 0), "z3": (0, 0, 3), "x2y": (2, 1, 0), "xy2": (1, 2, 0), "x2z": (2, 0, 1),
         "xz2": (1, 0, 2), "y2z": (0, 2, 1), "yz2": (0, 1, 2), "xyz": (1, 1, 1), "x2": (2, 0, 0), "y2": (0, 2, 0),
         "z2": (0, 0, 2), "xy": (1, 1, 0), "xz": (1, 0, 1), "yz": (0, 1, 1), "x": (1, 0, 0), "y": (0, 1, 0),
         "z": (0, 0, 1), "c": (0, 0, 0)}

BASE_CUBICS = {
    "dingdong": {"x2": 1.0, "z2": 1.0, "y3": 1.0, "y2": -1.0},
    "fermat": {"x3": 1.0, "y3": 1.0, "z3": 1.0, "c": -1.0},
    "monkey_saddle": {"x3": 1.0, "xz2": -3.0, "y": 10.0, "c": 5.0},
}


def _pmul(a: dict, b: dict) -> dict:
    out = {}
    for (ea, ca), (eb, cb) in itertools.product(a.items(), b.items()):
        e = (ea[0] + eb[0], ea[1] + eb[1], ea[2] + eb[2])
        out[e] = out.get(e, 0.0) + ca * cb
    return out


def _ppow(a: dict, n: int) -> dict:
    out = {(0, 0, 0): 1.0}
    for _ in range(n):
        out = _pmul(out, a)
    return out


def transform_polynomial(coefs: dict, center, scale: float) -> dict:
    """Coefficients of F((p - center) / scale) by name."""
    lin = []
    for axis in range(3):
        e = [0, 0, 0]
        e[axis] = 1
        lin.append({tuple(e): 1.0 / scale, (0, 0, 0): -center[axis] / scale})
    total = {}
    for name, c in coefs.items():
        i, j, k = _MONO[name]
        term = _pmul(_pmul(_ppow(lin[0], i), _ppow(lin[1], j)), _ppow(lin[2], k))
        for e, v in term.items():
            total[e] = total.get(e, 0.0) + c * v
    inv = {v: k for k, v in _MONO.items()}
    return {inv[e]: v for e, v in total.items() if v != 0.0}


def describe(n_spheres=4096, n_cubics=64, n_lights=16, width=7680, height=4320, seed=SEED) -> dict:
    """The scene as the plain dict a YAML loader would produce (all numbers already rounded to 6 digits)."""
    rng = SplitMix64(seed)
    u = lambda lo, hi: sig6(rng.uniform(lo, hi))
    spheres = []
    for i in range(n_spheres):
        spheres.append({"type": "sphere", "center": [u(-60, 60), u(-34, 34), u(40, 160)], "radius": u(0.3, 1.0),
                        "color": [u(0.1, 0.9), u(0.1, 0.9), u(0.1, 0.9)],
                        "reflection_ratio": 0.3 if i % 4 == 0 else 0.0})
    cubics = []
    kinds = list(BASE_CUBICS)
    for i in range(n_cubics):
        center = [u(-150, 150), u(-80, 80), u(200, 400)]
        scale = u(5, 20)
        poly = transform_polynomial(BASE_CUBICS[kinds[i % 3]], center, scale)
        cubics.append({"type": "polynomial", "coefficients": {k: sig6(v) for k, v in poly.items() if sig6(v) != 0.0},
                       "color": [u(0.1, 0.9), u(0.1, 0.9), u(0.1, 0.9)]})
    # cubics are spread through the object list (one after every n_spheres/n_cubics spheres) so object order != class order
    objects = []
    per = max(1, n_spheres // max(1, n_cubics))
    ci = 0
    for i, s in enumerate(spheres):
        objects.append(s)
        if (i + 1) % per == 0 and ci < n_cubics:
            objects.append(cubics[ci])
            ci += 1
    objects.extend(cubics[ci:])
    lights = [{"type": "spherical", "position": [u(-60, 60), u(-34, 34), u(40, 160)], "intensity": 20000.0,
               "color": [u(0.5, 1.0), u(0.5, 1.0), u(0.5, 1.0)]} for _ in range(n_lights)]
    return {"width": width, "height": height, "fov": 60, "max_reflections": 4, "bg_color": [0.0, 0.1, 0.2],
            "objects": objects, "light_sources": lights}


def synthetic_scene(n_spheres=4096, n_cubics=64, n_lights=16, width=7680, height=4320, seed=SEED) -> S.Scene:
    name = "synth_{}s_{}c_{}l".format(n_spheres, n_cubics, n_lights)
    return S.scene_from_dict(describe(n_spheres, n_cubics, n_lights, width, height, seed), name)


def to_yaml(desc: dict) -> str:
    """Emit the description in the scene-file schema (presentation/Instrukcja.md:17-32)."""
    vec = lambda v: "[" + ", ".join(repr(x) for x in v) + "]"
    out = ["width: {}".format(desc["width"]), "height: {}".format(desc["height"]), "fov: {}".format(desc["fov"]),
           "max_reflections: {}".format(desc["max_reflections"]), "bg_color: " + vec(desc["bg_color"]), "objects:"]
    for o in desc["objects"]:
        out.append("  - type: " + o["type"])
        if o["type"] == "sphere":
            out.append("    center: " + vec(o["center"]))
            out.append("    radius: " + repr(o["radius"]))
        else:
            out.append("    coefficients:")
            for k in S.COEF_NAMES:
                if k in o["coefficients"]:
                    out.append("      {}: {}".format(k, repr(o["coefficients"][k])))
        out.append("    color: " + vec(o["color"]))
        if o.get("reflection_ratio", 0.0):
            out.append("    reflection_ratio: " + repr(o["reflection_ratio"]))
    out.append("light_sources:")
    for l in desc["light_sources"]:
        out.append("  - type: " + l["type"])
        out.append("    position: " + vec(l["position"]))
        out.append("    intensity: " + repr(l["intensity"]))
        out.append("    color: " + vec(l["color"]))
    return "\n".join(out) + "\n"
