"""ctypes binding of librt_b200.so (include/rt_b200.h) and a host-side mirror of the reference's
back-end call surface, /root/reference/include/update.h:6-8:

    init_update(texture, scene)   -> rt_b200_upload_scene
    update(camera_matrix) -> ms   -> rt_b200_update
    cleanup_update()              -> rt_b200_destroy

`Renderer` is the re-entrant form (one opaque context per scene); the three module-level functions keep
one global context exactly like the reference's translation units do (update-cuda.cu:15-27).

There is NO CPU fallback here: if the CUDA library is missing or no device is present every call raises.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("RT_B200_LIB", os.path.join(HERE, "librt_b200.so"))   # (override: A/B experiments)

COUNTER_FIELDS = ("primary_rays", "reflection_rays", "shadow_rays", "hits", "lit",
                  "tests_plane", "tests_sphere", "tests_quadric", "tests_cubic", "fp64_confirms", "cubic_exact")

# every symbol include/rt_b200.h declares
EXPORTS = ("rt_b200_upload_scene", "rt_b200_update", "rt_b200_gather", "rt_b200_destroy", "rt_b200_last_error",
           "rt_b200_comm_unique_id", "rt_b200_comm_init", "rt_b200_get_counters", "rt_b200_set_option",
           "rt_b200_get_timing", "rt_b200_get_class_counts", "rt_b200_strip_layout", "rt_b200_version",
           "rt_b200_update_to_host", "rt_b200_frame_export", "rt_b200_frame_attach", "rt_b200_get_launches",
           "rt_b200_host_alloc", "rt_b200_host_free", "rt_b200_host_alloc_shared", "rt_b200_host_free_shared",
           "rt_b200_get_grid_info", "rt_b200_band_layout", "rt_b200_frame_attach_peer",
           "rt_b200_render_no_alloc", "rt_b200_no_alloc_scratch_needed", "rt_b200_debug_timeline")

UNIQUE_ID_BYTES = 128
IPC_HANDLE_BYTES = 64


class RtError(RuntimeError):
    pass


class _Counters(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in COUNTER_FIELDS]


_lib = None


def load_library():
    """dlopen librt_b200.so (build it first with build.py / __graft_entry__.build()).  Fails loudly."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RtError("{} is missing: run `python cuda-ray-tracer_b200/build.py` (there is no CPU fallback)".format(LIB_PATH))
    L = C.CDLL(LIB_PATH)
    vp = C.c_void_p
    L.rt_b200_upload_scene.argtypes = [C.POINTER(vp), C.c_uint32, C.c_uint32, C.c_double, vp, C.c_uint32, vp,
                                       C.c_size_t, vp, C.c_size_t, C.c_int, C.c_int, C.c_int]
    L.rt_b200_update.argtypes = [vp, vp, C.POINTER(C.c_float)]
    L.rt_b200_gather.argtypes = [vp, vp, C.c_size_t, C.POINTER(vp)]
    L.rt_b200_destroy.argtypes = [vp]
    L.rt_b200_last_error.restype = C.c_char_p
    L.rt_b200_comm_unique_id.argtypes = [vp]
    L.rt_b200_comm_init.argtypes = [vp, vp]
    L.rt_b200_get_counters.argtypes = [vp, C.POINTER(_Counters)]
    L.rt_b200_set_option.argtypes = [vp, C.c_char_p, C.c_int64]
    L.rt_b200_get_timing.argtypes = [vp, C.POINTER(C.c_float * 2)]
    L.rt_b200_get_class_counts.argtypes = [vp, C.POINTER(C.c_uint32 * 4)]
    L.rt_b200_get_grid_info.argtypes = [vp, C.POINTER(C.c_uint32 * 6)]
    L.rt_b200_band_layout.argtypes = [C.c_uint32, C.c_uint32, C.c_uint32, C.c_int, C.POINTER(C.c_int32)]
    L.rt_b200_version.restype = C.c_char_p
    L.rt_b200_update_to_host.argtypes = [vp, vp, vp, C.c_size_t, C.POINTER(C.c_float)]
    L.rt_b200_frame_export.argtypes = [vp, vp]
    L.rt_b200_frame_attach.argtypes = [vp, vp]
    L.rt_b200_frame_attach_peer.argtypes = [vp, vp]
    L.rt_b200_debug_timeline.argtypes = [vp, vp, C.c_size_t, C.POINTER(C.c_uint32)]
    L.rt_b200_render_no_alloc.argtypes = [C.c_uint32, C.c_uint32, C.c_double, vp, C.c_uint32, vp, C.c_size_t, vp, C.c_size_t,
                                          vp, C.c_int, vp, C.c_size_t, C.POINTER(C.c_float)]
    L.rt_b200_no_alloc_scratch_needed.argtypes = [vp, C.c_size_t, C.c_size_t, C.POINTER(C.c_size_t)]
    L.rt_b200_get_launches.argtypes = [vp, C.POINTER(C.c_uint32)]
    L.rt_b200_host_alloc.argtypes = [C.POINTER(vp), C.c_size_t]
    L.rt_b200_host_free.argtypes = [vp]
    L.rt_b200_host_alloc_shared.argtypes = [C.c_char_p, C.c_size_t, C.c_int, C.POINTER(vp)]
    L.rt_b200_host_free_shared.argtypes = [vp]
    L.rt_b200_strip_layout.argtypes = [C.c_uint32, C.c_uint32, C.c_int, C.c_int, vp, C.c_size_t, C.POINTER(C.c_uint32)]
    _lib = L
    return L


def _check(rc: int):
    if rc != 0:
        raise RtError("rt_b200 error {}: {}".format(rc, load_library().rt_b200_last_error().decode()))


def _dp(a):
    return a.ctypes.data_as(C.c_void_p)


def strip_layout(height: int, strip_rows: int, rank: int, world_size: int) -> np.ndarray:
    """Image row held by each packed row of `rank` (-1 = padding).  Host-only; needs no GPU."""
    L = load_library()
    n = C.c_uint32()
    _check(L.rt_b200_strip_layout(height, strip_rows, rank, world_size, None, 0, C.byref(n)))
    rows = np.empty(n.value, dtype=np.int32)
    _check(L.rt_b200_strip_layout(height, strip_rows, rank, world_size, _dp(rows), rows.size, C.byref(n)))
    return rows


def band_layout(rows: int, n_bands: int, unit: int = 16, taper: int = 0) -> np.ndarray:
    """Row-band boundaries of a streamed frame (n_bands + 1 row indices).  Host-only; needs no GPU."""
    starts = np.empty(n_bands + 1, dtype=np.int32)
    _check(load_library().rt_b200_band_layout(rows, n_bands, unit, taper, starts.ctypes.data_as(C.POINTER(C.c_int32))))
    return starts


def comm_unique_id() -> bytes:
    buf = (C.c_char * UNIQUE_ID_BYTES)()
    _check(load_library().rt_b200_comm_unique_id(buf))
    return bytes(buf)


def shared_frame_path(nbytes: int) -> str:
    """A fresh path for rt_b200_host_alloc_shared: /dev/shm when it has room, else the temp directory."""
    import shutil
    import tempfile
    import uuid
    root = "/dev/shm"
    try:
        if shutil.disk_usage(root).free < nbytes + (1 << 20):
            root = tempfile.gettempdir()
    except OSError:
        root = tempfile.gettempdir()
    return os.path.join(root, "rt_b200_{}_{}".format(os.getpid(), uuid.uuid4().hex[:8]))


def host_alloc_shared(path: str, nbytes: int, create: bool) -> int:
    """Host frame shared by all ranks of a multi-GPU job (one rank creates, the others open); returns the address."""
    p = C.c_void_p()
    _check(load_library().rt_b200_host_alloc_shared(path.encode(), nbytes, 1 if create else 0, C.byref(p)))
    return int(p.value)


def host_free_shared(ptr: int):
    _check(load_library().rt_b200_host_free_shared(C.c_void_p(ptr)))


def host_view(ptr: int, height: int, width: int) -> np.ndarray:
    """uint8 [H, W, 4] view of a host frame at `ptr` (no copy)."""
    buf = (C.c_uint8 * (height * width * 4)).from_address(ptr)
    return np.frombuffer(buf, dtype=np.uint8).reshape(height, width, 4)


NO_ALLOC_SCRATCH_BYTES = 1048576


def no_alloc_scratch_needed(scene) -> int:
    """Bytes of the library's static device scratch rt_b200_render_no_alloc needs for `scene` (host-only)."""
    objs = np.ascontiguousarray(scene.objects)
    n = C.c_size_t()
    _check(load_library().rt_b200_no_alloc_scratch_needed(_dp(objs), len(objs), len(scene.lights), C.byref(n)))
    return int(n.value)


def render_no_alloc(scene, camera_matrix, device_ptr: int, nbytes: int, device: int = 0) -> float:
    """rt_b200_render_no_alloc: one frame of `scene` into the caller's DEVICE buffer, no context and no allocation inside
    the library (its scratch is a static __device__ array).  Returns the host-clock render time in ms."""
    bg = np.ascontiguousarray(scene.bg_color, dtype=np.float32)
    objs = np.ascontiguousarray(scene.objects)
    lights = np.ascontiguousarray(scene.lights)
    m = np.ascontiguousarray(camera_matrix, dtype=np.float64).reshape(16)
    ms = C.c_float()
    _check(load_library().rt_b200_render_no_alloc(int(scene.px_width), int(scene.px_height), float(scene.vertical_fov), _dp(bg),
                                                  int(scene.max_reflections), _dp(objs), len(objs), _dp(lights), len(lights),
                                                  _dp(m), int(device), C.c_void_p(device_ptr), int(nbytes), C.byref(ms)))
    return ms.value


class Renderer:
    """One uploaded scene on one GPU (one rank of `world_size`)."""

    def __init__(self, scene, device: int = 0, rank: int = 0, world_size: int = 1):
        L = load_library()
        self.scene = scene
        self.width, self.height = int(scene.px_width), int(scene.px_height)
        self._h = C.c_void_p()
        bg = np.ascontiguousarray(scene.bg_color, dtype=np.float32)
        objs = np.ascontiguousarray(scene.objects)
        lights = np.ascontiguousarray(scene.lights)
        _check(L.rt_b200_upload_scene(C.byref(self._h), self.width, self.height, float(scene.vertical_fov), _dp(bg),
                                      int(scene.max_reflections), _dp(objs), len(objs), _dp(lights), len(lights),
                                      int(device), int(rank), int(world_size)))

    # -- update.h:7
    def update(self, camera_matrix) -> float:
        m = np.ascontiguousarray(camera_matrix, dtype=np.float64).reshape(16)
        ms = C.c_float()
        _check(load_library().rt_b200_update(self._h, _dp(m), C.byref(ms)))
        return ms.value

    def update_to_host(self, camera_matrix, host_ptr: int, nbytes: int) -> float:
        """update() + result sink in one call: the frame lands in caller-owned (ideally pinned) host memory, the
        device->host copy of finished row bands overlapping the rendering of the next ones."""
        m = np.ascontiguousarray(camera_matrix, dtype=np.float64).reshape(16)
        ms = C.c_float()
        _check(load_library().rt_b200_update_to_host(self._h, _dp(m), C.c_void_p(host_ptr), nbytes, C.byref(ms)))
        return ms.value

    def frame_export(self) -> bytes:
        """Rank 0: IPC handle of the assembled frame every rank's kernel will write into (fused gather)."""
        buf = (C.c_char * IPC_HANDLE_BYTES)()
        _check(load_library().rt_b200_frame_export(self._h, buf))
        return bytes(buf)

    def frame_attach(self, handle: bytes):
        buf = C.create_string_buffer(handle, IPC_HANDLE_BYTES)
        _check(load_library().rt_b200_frame_attach(self._h, buf))

    def frame_attach_peer(self, owner: "Renderer"):
        """Same-process fused gather: `owner` is rank 0's Renderer (after frame_export), possibly on another device."""
        _check(load_library().rt_b200_frame_attach_peer(self._h, owner._h))

    def debug_timeline(self) -> np.ndarray:
        """uint64 [2048, words]: per-CTA stamps of the last launch (timeline build of the library + option "timeline")."""
        out = np.zeros(2048 * 8, dtype=np.uint64)
        w = C.c_uint32()
        _check(load_library().rt_b200_debug_timeline(self._h, _dp(out), out.size, C.byref(w)))
        return out.reshape(2048, int(w.value))

    def launches(self) -> int:
        n = C.c_uint32()
        _check(load_library().rt_b200_get_launches(self._h, C.byref(n)))
        return int(n.value)

    def gather(self, out: np.ndarray | None = None) -> np.ndarray:
        """Frame as uint8 [H, W, 4], row 0 = bottom of the picture."""
        if out is None:
            out = np.empty((self.height, self.width, 4), dtype=np.uint8)
        assert out.dtype == np.uint8 and out.flags.c_contiguous
        _check(load_library().rt_b200_gather(self._h, _dp(out), out.nbytes, None))
        return out

    def gather_into(self, host_ptr: int, nbytes: int):
        """Copy the frame into caller-owned host memory (e.g. a pinned torch tensor's data_ptr)."""
        _check(load_library().rt_b200_gather(self._h, C.c_void_p(host_ptr), nbytes, None))

    def device_frame_ptr(self) -> int:
        p = C.c_void_p()
        _check(load_library().rt_b200_gather(self._h, None, 0, C.byref(p)))
        return p.value

    def comm_init(self, unique_id: bytes):
        buf = C.create_string_buffer(unique_id, UNIQUE_ID_BYTES)
        _check(load_library().rt_b200_comm_init(self._h, buf))

    def counters(self) -> dict:
        c = _Counters()
        _check(load_library().rt_b200_get_counters(self._h, C.byref(c)))
        d = {n: int(getattr(c, n)) for n in COUNTER_FIELDS}
        d["rays"] = d["primary_rays"] + d["reflection_rays"] + d["shadow_rays"]
        return d

    def set_option(self, name: str, value: int):
        _check(load_library().rt_b200_set_option(self._h, name.encode(), int(value)))

    def timing(self):
        t = (C.c_float * 2)()
        _check(load_library().rt_b200_get_timing(self._h, C.byref(t)))
        return {"render_ms": t[0], "gather_ms": t[1]}

    def class_counts(self):
        t = (C.c_uint32 * 4)()
        _check(load_library().rt_b200_get_class_counts(self._h, C.byref(t)))
        return {"plane": t[0], "sphere": t[1], "quadric": t[2], "cubic": t[3]}

    def grid_info(self):
        """The sphere grid built at upload (option accel, csrc/rt_grid.cuh)."""
        t = (C.c_uint32 * 6)()
        _check(load_library().rt_b200_get_grid_info(self._h, C.byref(t)))
        return {"nx": t[0], "ny": t[1], "nz": t[2], "gridded": t[3], "loose": t[4], "registrations": t[5]}

    # -- update.h:8
    def close(self):
        if self._h:
            load_library().rt_b200_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()


# ---- the reference's three free functions, one global context (update.h:6-8) ------------------------
_global: Renderer | None = None


def init_update(texture, scene, device: int = 0) -> None:
    """update.h:6.  `texture` (a GL name in the reference) is ignored: the frame is read with gather()."""
    global _global
    if _global is not None:
        _global.close()
    _global = Renderer(scene, device=device)


def update(camera_matrix) -> float:
    """update.h:7.  Returns the render time in milliseconds, like the reference."""
    if _global is None:
        raise RtError("update() before init_update()")
    return _global.update(camera_matrix)


def gather(out=None) -> np.ndarray:
    if _global is None:
        raise RtError("gather() before init_update()")
    return _global.gather(out)


def cleanup_update() -> None:
    """update.h:8."""
    global _global
    if _global is not None:
        _global.close()
        _global = None
