"""TEST INFRASTRUCTURE — ctypes wrappers around the two CPU checkers.

  RefCpu      oracle/_ref/libref_cpu.so : the reference's own CPU back-end compiled unmodified
              (built by oracle/Makefile only where /root/reference exists; the .so travels to the GPU box)
  PortOracle  oracle/librt_oracle.so    : the plain-C restatement (oracle/rt_oracle.c)

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this
module.  Nothing under cuda-ray-tracer_b200/ (the product) may.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_LIB = os.path.join(HERE, "_ref", "libref_cpu.so")
REF_FMA_LIB = os.path.join(HERE, "_ref", "libref_cpu_fma.so")   # FMA-contracted build: for the ill-conditioned mask only
PORT_LIB = os.path.join(HERE, "librt_oracle.so")

IDENTITY = np.eye(4, dtype=np.float64).reshape(16)


def build(quiet: bool = True) -> None:
    """make -C oracle: always builds the port; builds _ref when /root/reference is present."""
    subprocess.run(["make", "-C", HERE] + (["-s"] if quiet else []), check=True)


def have_ref() -> bool:
    return os.path.exists(REF_LIB)


def _dp(a):
    return a.ctypes.data_as(C.c_void_p)


class _Counters(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in ("primary_rays", "reflection_rays", "shadow_rays", "nearest_tests",
                                          "shadow_tests", "hits", "lit")]


def quantize_rgba8(rgb: np.ndarray) -> np.ndarray:
    """float RGB -> RGBA8 with (int)(c*255.0f + 0.5f), alpha 255 (update-cuda.cu:149-156)."""
    rgb = np.ascontiguousarray(rgb, dtype=np.float32)
    q = (rgb * np.float32(255.0) + np.float32(0.5)).astype(np.int32).astype(np.uint8)
    out = np.full(rgb.shape[:-1] + (4,), 255, dtype=np.uint8)
    out[..., :3] = q
    return out


def fnv1a64(data: bytes) -> str:
    h = 0xcbf29ce484222325
    for b in data:
        h = ((h ^ b) * 0x100000001b3) & 0xFFFFFFFFFFFFFFFF
    return "{:016x}".format(h)


def fnv1a64_np(arr: np.ndarray) -> str:
    """Vector-friendly FNV-1a-64 is inherently serial; go through a tiny C loop via numpy bytes."""
    return fnv1a64(np.ascontiguousarray(arr).tobytes())


class PortOracle:
    """One scene in the plain-C restatement.  Re-entrant; any number may coexist."""

    _lib = None

    @classmethod
    def lib(cls):
        if cls._lib is None:
            if not os.path.exists(PORT_LIB):
                build()
            L = C.CDLL(PORT_LIB)
            L.rto_scene_create.restype = C.c_void_p
            L.rto_scene_create.argtypes = [C.c_uint32, C.c_uint32, C.c_double, C.c_void_p, C.c_uint32,
                                           C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t]
            L.rto_scene_destroy.argtypes = [C.c_void_p]
            L.rto_intersect_ray.restype = C.c_double
            L.rto_intersect_ray.argtypes = [C.c_void_p] * 3
            L.rto_normal_vector.argtypes = [C.c_void_p] * 3
            L.rto_render_rows.restype = C.c_float
            L.rto_render_rows.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int,
                                          C.c_void_p, C.c_void_p]
            L.rto_render_pixels.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
            cls._lib = L
        return cls._lib

    def __init__(self, scene):
        L = self.lib()
        self.scene = scene
        bg = np.ascontiguousarray(scene.bg_color, dtype=np.float32)
        objs = np.ascontiguousarray(scene.objects)
        lights = np.ascontiguousarray(scene.lights)
        self._h = L.rto_scene_create(scene.px_width, scene.px_height, float(scene.vertical_fov), _dp(bg),
                                     scene.max_reflections, _dp(objs), len(objs), _dp(lights), len(lights))
        if not self._h:
            raise MemoryError("rto_scene_create failed")

    def close(self):
        if self._h:
            self.lib().rto_scene_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def render_rows(self, camera=IDENTITY, y0=0, y1=None, ystep=1, threads=1, count=False):
        """Returns (rgb float32 [rows, W, 3], elapsed_ms[, counters dict])."""
        s = self.scene
        y1 = s.px_height if y1 is None else y1
        n_rows = len(range(y0, y1, ystep))
        out = np.empty((n_rows, s.px_width, 3), dtype=np.float32)
        m = np.ascontiguousarray(camera, dtype=np.float64).reshape(16)
        cnt = _Counters()
        ms = self.lib().rto_render_rows(self._h, _dp(m), y0, y1, ystep, threads, _dp(out),
                                        C.byref(cnt) if count else None)
        if count:
            return out, ms, {n: getattr(cnt, n) for n, _ in _Counters._fields_}
        return out, ms

    def render(self, camera=IDENTITY, threads=1):
        return self.render_rows(camera, threads=threads)[0]

    def render_pixels(self, camera, xy):
        xy = np.ascontiguousarray(xy, dtype=np.int32).reshape(-1, 2)
        out = np.empty((len(xy), 3), dtype=np.float32)
        m = np.ascontiguousarray(camera, dtype=np.float64).reshape(16)
        self.lib().rto_render_pixels(self._h, _dp(m), _dp(xy), len(xy), _dp(out))
        return out

    def render_pixels_mt(self, camera, xy, threads=None):
        """render_pixels over `threads` host threads (the restatement is re-entrant and ctypes drops the GIL) — the
        sparse oracle of frames whose full CPU render would take hours (C5 at 7680x4320: ~0.75 ms per pixel)."""
        from concurrent.futures import ThreadPoolExecutor
        xy = np.ascontiguousarray(xy, dtype=np.int32).reshape(-1, 2)
        threads = max(1, min(threads or (os.cpu_count() or 1), len(xy)))
        chunks = np.array_split(np.arange(len(xy)), threads * 4)
        out = np.empty((len(xy), 3), dtype=np.float32)
        with ThreadPoolExecutor(max_workers=threads) as ex:
            for idx, rgb in zip(chunks, ex.map(lambda i: self.render_pixels(camera, xy[i]), chunks)):
                out[idx] = rgb
        return out

    @classmethod
    def intersect_ray(cls, coef, origin, direction) -> float:
        a = np.ascontiguousarray(coef, dtype=np.float64)
        o = np.ascontiguousarray(origin, dtype=np.float64)
        d = np.ascontiguousarray(direction, dtype=np.float64)
        return cls.lib().rto_intersect_ray(_dp(a), _dp(o), _dp(d))

    @classmethod
    def normal_vector(cls, coef, pos):
        a = np.ascontiguousarray(coef, dtype=np.float64)
        p = np.ascontiguousarray(pos, dtype=np.float64)
        out = np.empty(3)
        cls.lib().rto_normal_vector(_dp(a), _dp(p), _dp(out))
        return out


class RefCpu:
    """The unmodified reference CPU back-end.  It keeps its state in file-scope globals
    (update-cpu.cpp:10-19), so exactly one scene is live per process (and library): init() replaces it."""

    _lib = None
    LIB_PATH = REF_LIB

    @classmethod
    def lib(cls):
        if cls._lib is None:
            if not os.path.exists(cls.LIB_PATH):
                raise FileNotFoundError(cls.LIB_PATH + " (build with `make -C oracle` where /root/reference exists)")
            L = C.CDLL(cls.LIB_PATH)
            L.ref_last_error.restype = C.c_char_p
            L.ref_init.argtypes = [C.c_uint32, C.c_uint32, C.c_double, C.c_void_p, C.c_uint32, C.c_void_p,
                                   C.c_size_t, C.c_void_p, C.c_size_t]
            L.ref_update.restype = C.c_float
            L.ref_update.argtypes = [C.c_void_p, C.c_void_p]
            L.ref_render_rows.restype = C.c_float
            L.ref_render_rows.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p]
            L.ref_render_pixels.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
            L.ref_intersect_ray.restype = C.c_double
            L.ref_intersect_ray.argtypes = [C.c_void_p] * 3
            L.ref_normal_vector.argtypes = [C.c_void_p] * 3
            L.ref_make_surface.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_double, C.c_void_p]
            L.ref_make_object.argtypes = [C.c_void_p, C.c_float, C.c_void_p, C.c_void_p]
            L.ref_make_light.argtypes = [C.c_int, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p]
            cls._lib = L
        return cls._lib

    def __init__(self, scene):
        L = self.lib()
        self.scene = scene
        bg = np.ascontiguousarray(scene.bg_color, dtype=np.float32)
        objs = np.ascontiguousarray(scene.objects)
        lights = np.ascontiguousarray(scene.lights)
        rc = L.ref_init(scene.px_width, scene.px_height, float(scene.vertical_fov), _dp(bg), scene.max_reflections,
                        _dp(objs), len(objs), _dp(lights), len(lights))
        if rc != 0:
            raise RuntimeError(L.ref_last_error().decode())

    def update(self, camera=IDENTITY):
        """The reference's own update(): returns (rgb float32 [H, W, 3], the reference's ms)."""
        s = self.scene
        out = np.empty((s.px_height, s.px_width, 3), dtype=np.float32)
        m = np.ascontiguousarray(camera, dtype=np.float64).reshape(16)
        ms = self.lib().ref_update(_dp(m), _dp(out))
        return out, ms

    def render_rows(self, camera=IDENTITY, y0=0, y1=None, ystep=1, threads=1):
        s = self.scene
        y1 = s.px_height if y1 is None else y1
        n_rows = len(range(y0, y1, ystep))
        out = np.empty((n_rows, s.px_width, 3), dtype=np.float32)
        m = np.ascontiguousarray(camera, dtype=np.float64).reshape(16)
        ms = self.lib().ref_render_rows(_dp(m), y0, y1, ystep, threads, _dp(out))
        return out, ms

    def render_pixels(self, camera, xy):
        xy = np.ascontiguousarray(xy, dtype=np.int32).reshape(-1, 2)
        out = np.empty((len(xy), 3), dtype=np.float32)
        m = np.ascontiguousarray(camera, dtype=np.float64).reshape(16)
        self.lib().ref_render_pixels(_dp(m), _dp(xy), len(xy), _dp(out))
        return out

    @classmethod
    def intersect_ray(cls, coef, origin, direction) -> float:
        a = np.ascontiguousarray(coef, dtype=np.float64)
        o = np.ascontiguousarray(origin, dtype=np.float64)
        d = np.ascontiguousarray(direction, dtype=np.float64)
        return cls.lib().ref_intersect_ray(_dp(a), _dp(o), _dp(d))

    @classmethod
    def normal_vector(cls, coef, pos):
        a = np.ascontiguousarray(coef, dtype=np.float64)
        p = np.ascontiguousarray(pos, dtype=np.float64)
        out = np.empty(3)
        cls.lib().ref_normal_vector(_dp(a), _dp(p), _dp(out))
        return out

    @classmethod
    def make_surface(cls, kind, a=None, b=None, s=0.0):
        kinds = {"sphere": 0, "plane": 1, "dingDong": 2, "clebsch": 3, "cayley": 4}
        out = np.empty(20)
        av = np.ascontiguousarray(a, dtype=np.float64) if a is not None else None
        bv = np.ascontiguousarray(b, dtype=np.float64) if b is not None else None
        rc = cls.lib().ref_make_surface(kinds[kind], _dp(av) if av is not None else None,
                                        _dp(bv) if bv is not None else None, float(s), _dp(out))
        if rc != 0:
            raise RuntimeError(cls.lib().ref_last_error().decode())
        return out

    @classmethod
    def make_object(cls, coef, reflection_ratio, color):
        out = np.zeros(176, dtype=np.uint8)
        k = np.ascontiguousarray(coef, dtype=np.float64)
        c = np.ascontiguousarray(color, dtype=np.float32)
        rc = cls.lib().ref_make_object(_dp(k), float(reflection_ratio), _dp(c), _dp(out))
        if rc != 0:
            raise RuntimeError(cls.lib().ref_last_error().decode())
        return out

    @classmethod
    def make_light(cls, spherical, intensity, vec, color):
        out = np.zeros(48, dtype=np.uint8)
        v = np.ascontiguousarray(vec, dtype=np.float64)
        c = np.ascontiguousarray(color, dtype=np.float32)
        rc = cls.lib().ref_make_light(1 if spherical else 0, float(intensity), _dp(v), _dp(c), _dp(out))
        if rc != 0:
            raise RuntimeError(cls.lib().ref_last_error().decode())
        return out


class RefCpuFma(RefCpu):
    """The same sources compiled with -mfma -ffp-contract=fast.  NOT an oracle: a perturbed run for the mask below."""
    _lib = None
    LIB_PATH = REF_FMA_LIB


def cpu_has_fma() -> bool:
    try:
        with open("/proc/cpuinfo") as f:
            return " fma " in f.read().replace("\n", " ")
    except OSError:
        return False


def have_ref_fma() -> bool:
    return os.path.exists(REF_FMA_LIB) and cpu_has_fma()


def _dilate(mask: np.ndarray) -> np.ndarray:
    """3 x 3 dilation (the mask's 1-pixel safety margin)."""
    out = mask.copy()
    h, w = mask.shape
    for dy in (-1, 0, 1):
        for dx in (-1, 0, 1):
            ys, yd = (slice(max(dy, 0), h + min(dy, 0)), slice(max(-dy, 0), h + min(-dy, 0)))
            xs, xd = (slice(max(dx, 0), w + min(dx, 0)), slice(max(-dx, 0), w + min(-dx, 0)))
            out[yd, xd] |= mask[ys, xs]
    return out


def ill_conditioned_mask(scene, camera=IDENTITY, threads=1) -> np.ndarray:
    """bool [H, W]: the pixels where the REFERENCE's own result is ill-conditioned (SURVEY.md §7.1 step 1, §8c acceptance
    rule "outliers within the mask").  Three runs of the reference CPU back-end on the same frame —
      A  the oracle build (-O2 -ffp-contract=off)
      B  the same sources with FMA contraction (-mfma -ffp-contract=fast; skipped on a CPU without FMA)
      C  build A with every entry of the camera matrix moved up by one ulp
    — a pixel is in the mask when any two of them differ by more than 1 LSB in some 8-bit channel, plus one pixel of
    dilation.  Root solving that survives a perturbation of one rounding is well-conditioned; what flips is not."""
    m = np.ascontiguousarray(camera, dtype=np.float64).reshape(16)
    runs = [quantize_rgba8(RefCpu(scene).render_rows(m, threads=threads)[0])]
    if have_ref_fma():
        runs.append(quantize_rgba8(RefCpuFma(scene).render_rows(m, threads=threads)[0]))
    m_up = np.nextafter(m, np.where(m >= 0, np.inf, -np.inf))
    m_up[m == 0.0] = 0.0      # structural zeros of the matrix stay zero
    runs.append(quantize_rgba8(RefCpu(scene).render_rows(m_up, threads=threads)[0]))
    mask = np.zeros(runs[0].shape[:2], dtype=bool)
    for i in range(len(runs)):
        for j in range(i + 1, len(runs)):
            mask |= np.abs(runs[i].astype(np.int16) - runs[j].astype(np.int16)).max(axis=-1) > 1
    return _dilate(mask)
