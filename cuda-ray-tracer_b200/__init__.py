"""cuda-ray-tracer_b200 — B200-native per-pixel frame update for the implicit-surface ray tracer.

The directory name is not a Python identifier; load it with `tools/loader.py` (`load_package()`),
which registers it as the module `cuda_ray_tracer_b200`.
"""
from . import camera, scene, synth  # noqa: F401
from .scene import Scene, SceneException, load_from_file  # noqa: F401
