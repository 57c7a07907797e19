"""Render a few frames of one workload scene — the short command ncu wraps.
    python tools/prof_scene.py SCENE [--frames 3] [--kernel 1] [--width 3840 --height 2160]"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
from loader import load_package

pkg = load_package()
from cuda_ray_tracer_b200 import binding, camera

ap = argparse.ArgumentParser()
ap.add_argument("scene")
ap.add_argument("--frames", type=int, default=3)
ap.add_argument("--kernel", type=int, default=1)
ap.add_argument("--width", type=int, default=3840)
ap.add_argument("--height", type=int, default=2160)
ap.add_argument("--bounces", type=int, default=4)
ap.add_argument("--pose", default="initial")
ap.add_argument("--accel", type=int, default=-1)
ap.add_argument("--spheres", type=int, default=4096)
ap.add_argument("--cubics", type=int, default=64)
ap.add_argument("--lights", type=int, default=16)
ap.add_argument("--count", type=int, default=0, help="1 = the counting instantiation of the kernel")
ap.add_argument("--option", action="append", default=[], help="name=value, passed to rt_b200_set_option")
a = ap.parse_args()
if a.scene == "synthetic":
    from cuda_ray_tracer_b200 import synth
    sc = synth.synthetic_scene(n_spheres=a.spheres, n_cubics=a.cubics, n_lights=a.lights, width=a.width, height=a.height)
else:
    sc = pkg.load_from_file(os.path.join(ROOT, "scenes", a.scene + ".yml")).with_overrides(
        width=a.width, height=a.height, max_reflections=a.bounces)
with binding.Renderer(sc) as r:
    r.set_option("kernel", a.kernel)
    r.set_option("accel", a.accel)
    r.set_option("count", a.count)
    for o in a.option:
        k, v = o.split("=")
        r.set_option(k, int(v))
    for _ in range(a.frames):
        ms = r.update(camera.TEST_POSES[a.pose])
    print(a.scene, "kernel", a.kernel, "last frame ms", ms, r.counters() if a.count else "")
