"""Scene model and YAML loader — Python mirror of the reference's host-side scene code.

Follows /root/reference/src/scene.cpp:6-22,97-203 (defaults, required keys, validation),
src/surface.cpp:4-60 (surface factories, including the clebsch z3 = 0 quirk at :44),
src/light.cpp:4-26 (light factories) and src/scene-exception.cpp:3-11 (colour validation).

Records use the reference's exact in-memory layouts (include/scene.h:8-15 Object = 176 bytes,
include/light.h:6-13 LightSource = 48 bytes) so `objects.ctypes.data` can be handed straight to
`rt_b200_upload_scene` — the same bytes `&scene.objects[0]` would be in the C++ host.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field

import numpy as np
import yaml

COEF_NAMES = ("x3", "y3", "z3", "x2y", "xy2", "x2z", "xz2", "y2z", "yz2", "xyz",
              "x2", "y2", "z2", "xy", "xz", "yz", "x", "y", "z", "c")
COEF_INDEX = {n: i for i, n in enumerate(COEF_NAMES)}

OBJECT_DTYPE = np.dtype([("coef", "<f8", (20,)), ("reflection_ratio", "<f4"), ("color", "<f4", (3,))])
LIGHT_DTYPE = np.dtype([("is_spherical", "u1"), ("pad0", "u1", (7,)), ("p", "<f8", (3,)),
                        ("light_color", "<f4", (3,)), ("pad1", "<u4")])
assert OBJECT_DTYPE.itemsize == 176 and LIGHT_DTYPE.itemsize == 48

# scene.cpp:6-7
DEFAULT_MAX_REFLECTIONS = 5
DEFAULT_BG_COLOR = (1.0, 1.0, 1.0)
_DEG_TO_RAD = 0.01745329251994329576923690768489  # glm::radians


class SceneException(Exception):
    """include/scene-exception.h:10-25."""


def _fmt(v) -> str:
    return "{:g}".format(float(v))


def validate_positive(what: str, value) -> None:
    """include/scene-exception.h:27-35 — rejects only values < 0."""
    if value < 0:
        raise SceneException("Negative value for {}: {}".format(what, _fmt(value)))


def validate_color(color) -> None:
    """src/scene-exception.cpp:3-11 — every channel must lie in [0, 1]."""
    c = [np.float32(v) for v in color]
    if any((v < 0.0) or (v > 1.0) for v in c):
        raise SceneException("Invalid color: ({}, {}, {})".format(*(_fmt(v) for v in c)))


def _dot(a, b) -> float:
    return (a[0] * b[0] + a[1] * b[1]) + a[2] * b[2]


# ---- surface factories (src/surface.cpp) --------------------------------------------------------
def surface_sphere(center, radius) -> np.ndarray:
    validate_positive("sphere radius", radius)
    cx, cy, cz = (float(v) for v in center)
    k = np.zeros(20)
    k[COEF_INDEX["x2"]] = k[COEF_INDEX["y2"]] = k[COEF_INDEX["z2"]] = 1.0
    k[COEF_INDEX["x"]] = -2.0 * cx
    k[COEF_INDEX["y"]] = -2.0 * cy
    k[COEF_INDEX["z"]] = -2.0 * cz
    k[COEF_INDEX["c"]] = _dot((cx, cy, cz), (cx, cy, cz)) - float(radius) * float(radius)
    return k


def surface_plane(origin, normal) -> np.ndarray:
    o = [float(v) for v in origin]
    n = [float(v) for v in normal]
    k = np.zeros(20)
    k[COEF_INDEX["x"]], k[COEF_INDEX["y"]], k[COEF_INDEX["z"]] = n
    k[COEF_INDEX["c"]] = -_dot(o, n)
    return k


def surface_ding_dong(origin) -> np.ndarray:
    ox, oy, oz = (float(v) for v in origin)
    k = np.zeros(20)
    k[COEF_INDEX["x2"]] = k[COEF_INDEX["y3"]] = k[COEF_INDEX["z2"]] = 1.0
    k[COEF_INDEX["y2"]] = -1.0 - 3.0 * oy
    k[COEF_INDEX["x"]] = -2.0 * ox
    k[COEF_INDEX["z"]] = -2.0 * oz
    k[COEF_INDEX["y"]] = (2.0 + 3.0 * oy) * oy
    k[COEF_INDEX["c"]] = math.pow(ox, 2) + math.pow(oz, 2) - math.pow(oy, 2) * (1.0 + oy)
    return k


def surface_clebsch() -> np.ndarray:
    k = np.zeros(20)
    # surface.cpp:44 assigns x3 twice and never z3: z3 stays 0.  Part of the oracle; do not "fix".
    k[COEF_INDEX["x3"]] = k[COEF_INDEX["y3"]] = 81.0
    for n in ("x2y", "x2z", "xy2", "y2z", "xz2", "yz2"):
        k[COEF_INDEX[n]] = -189.0
    k[COEF_INDEX["xyz"]] = 54.0
    for n in ("xy", "yz", "xz"):
        k[COEF_INDEX[n]] = 126.0
    for n in ("x2", "y2", "z2"):
        k[COEF_INDEX[n]] = -9.0
    for n in ("x", "y", "z"):
        k[COEF_INDEX[n]] = 9.0
    k[COEF_INDEX["c"]] = 1.0
    return k


def surface_cayley() -> np.ndarray:
    k = np.zeros(20)
    for n in ("x2y", "x2z", "xy2", "y2z", "xz2", "yz2"):
        k[COEF_INDEX[n]] = -5.0
    for n in ("xy", "yz", "xz"):
        k[COEF_INDEX[n]] = 2.0
    return k


def surface_polynomial(coefficients: dict) -> np.ndarray:
    k = np.zeros(20)
    for name, idx in COEF_INDEX.items():
        k[idx] = float(coefficients.get(name, 0.0))
    return k


# ---- light factories (src/light.cpp) -------------------------------------------------------------
def _light_record(spherical: bool, intensity, vec, color) -> np.ndarray:
    validate_positive("light intensity", intensity)
    validate_color(color)
    rec = np.zeros((), dtype=LIGHT_DTYPE)
    rec["is_spherical"] = 1 if spherical else 0
    rec["light_color"] = np.float32(intensity) * np.asarray(color, dtype=np.float32)
    rec["p"] = vec
    return rec


def light_directional(intensity, direction, color) -> np.ndarray:
    d = [float(v) for v in direction]
    inv = 1.0 / math.sqrt(_dot(d, d))
    return _light_record(False, intensity, [-(v * inv) for v in d], color)


def light_spherical(intensity, position, color) -> np.ndarray:
    return _light_record(True, intensity, [float(v) for v in position], color)


def make_object(coef, reflection_ratio, color) -> np.ndarray:
    """scene.cpp:9-14."""
    validate_positive("object reflection ratio", reflection_ratio)
    validate_color(color)
    rec = np.zeros((), dtype=OBJECT_DTYPE)
    rec["coef"] = coef
    rec["reflection_ratio"] = np.float32(reflection_ratio)
    rec["color"] = np.asarray(color, dtype=np.float32)
    return rec


@dataclass
class Scene:
    """include/scene.h:17-36.  vertical_fov is stored in radians (scene.cpp:20)."""
    px_width: int
    px_height: int
    vertical_fov: float
    bg_color: np.ndarray
    max_reflections: int
    objects: np.ndarray = field(default_factory=lambda: np.zeros(0, dtype=OBJECT_DTYPE))
    lights: np.ndarray = field(default_factory=lambda: np.zeros(0, dtype=LIGHT_DTYPE))
    name: str = ""

    def aspect_ratio(self) -> float:
        return float(self.px_width) / self.px_height

    def with_overrides(self, width=None, height=None, max_reflections=None) -> "Scene":
        """The benchmark configs render bundled scenes at other resolutions / bounce limits."""
        return Scene(int(width or self.px_width), int(height or self.px_height), self.vertical_fov,
                     self.bg_color.copy(),
                     self.max_reflections if max_reflections is None else int(max_reflections),
                     self.objects.copy(), self.lights.copy(), self.name)


def _require(node: dict, key: str, where: str):
    if not isinstance(node, dict) or key not in node or node[key] is None:
        raise SceneException("Value '{}' undefined, {}".format(key, where))
    return node[key]


def _vec3(value, key: str, where: str):
    if not isinstance(value, (list, tuple)) or len(value) != 3:
        raise SceneException("Value '{}' is invalid, {}".format(key, where))
    try:
        return [float(v) for v in value]
    except (TypeError, ValueError):
        raise SceneException("Value '{}' is invalid, {}".format(key, where))


def parse_surface(node: dict, where: str) -> np.ndarray:
    """scene.cpp:97-151."""
    kind = _require(node, "type", where)
    if kind == "sphere":
        return surface_sphere(_vec3(node.get("center", [0.0] * 3), "center", where), float(node.get("radius", 1.0)))
    if kind == "plane":
        return surface_plane(_vec3(node.get("origin", [0.0] * 3), "origin", where),
                             _vec3(node.get("normal", [0.0, 1.0, 0.0]), "normal", where))
    if kind == "dingDong":
        return surface_ding_dong(_vec3(node.get("origin", [0.0] * 3), "origin", where))
    if kind == "clebsch":
        return surface_clebsch()
    if kind == "cayley":
        return surface_cayley()
    if kind == "polynomial":
        coefs = _require(node, "coefficients", where)
        if not isinstance(coefs, dict):
            raise SceneException("Value 'coefficients' must be a mapping, {}".format(where))
        return surface_polynomial(coefs)
    raise SceneException("Unknown surface type: '{}', {}".format(kind, where))


def scene_from_dict(desc: dict, name: str = "") -> Scene:
    """scene.cpp:154-203 on an already parsed YAML document."""
    where = "document"
    try:
        width = int(_require(desc, "width", where))
        height = int(_require(desc, "height", where))
        fov_deg = float(_require(desc, "fov", where))
    except (TypeError, ValueError):
        raise SceneException("Value 'width/height/fov' is invalid, {}".format(where))
    if width < 0 or height < 0:
        raise SceneException("Value 'width/height' is invalid, {}".format(where))
    bg = _vec3(desc.get("bg_color", list(DEFAULT_BG_COLOR)), "bg_color", where)
    validate_color(bg)
    scene = Scene(width, height, fov_deg * _DEG_TO_RAD, np.asarray(bg, dtype=np.float32),
                  int(desc.get("max_reflections", DEFAULT_MAX_REFLECTIONS)), name=name)
    for key in ("objects", "light_sources"):
        if key not in desc or desc[key] is None:
            raise SceneException("Value '{}' undefined, {}".format(key, where))
        if not isinstance(desc[key], list):
            raise SceneException("Value '{}' must be a sequence, {}".format(key, where))
    objects = []
    for i, node in enumerate(desc["objects"]):
        w = "objects[{}]".format(i)
        coef = parse_surface(node, w)
        refl = float(node.get("reflection_ratio", 0.0))
        color = _vec3(_require(node, "color", w), "color", w)
        objects.append(make_object(coef, refl, color))
    lights = []
    for i, node in enumerate(desc["light_sources"]):
        w = "light_sources[{}]".format(i)
        kind = _require(node, "type", w)
        intensity = float(node.get("intensity", 1.0))
        color = _vec3(node.get("color", [1.0, 1.0, 1.0]), "color", w)
        if kind == "directional":
            lights.append(light_directional(intensity, _vec3(_require(node, "direction", w), "direction", w), color))
        elif kind == "spherical":
            lights.append(light_spherical(intensity, _vec3(_require(node, "position", w), "position", w), color))
        else:
            raise SceneException("Light source type must be 'spherical' or 'directional', {}".format(w))
    scene.objects = np.array(objects, dtype=OBJECT_DTYPE) if objects else np.zeros(0, dtype=OBJECT_DTYPE)
    scene.lights = np.array(lights, dtype=LIGHT_DTYPE) if lights else np.zeros(0, dtype=LIGHT_DTYPE)
    return scene


def load_from_file(path: str) -> Scene:
    """Scene::load_from_file, scene.cpp:154-203."""
    import os
    try:
        with open(path, "r") as f:
            text = f.read()
    except OSError:
        raise SceneException("Cannot read the file {}".format(path))
    try:
        desc = yaml.safe_load(text)
    except yaml.YAMLError as e:
        raise SceneException("YAML parser error: {}".format(e))
    if not isinstance(desc, dict):
        raise SceneException("Value 'width' undefined, document")
    return scene_from_dict(desc, os.path.splitext(os.path.basename(path))[0])
