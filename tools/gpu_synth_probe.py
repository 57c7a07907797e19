"""Synthetic C5 scene on a small frame: kernel ms per variant, brute force and grid, with the test counters."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools"))
from loader import load_package
pkg = load_package()
from cuda_ray_tracer_b200 import binding, camera, synth
w, h = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (960, 540)
sc = synth.synthetic_scene(width=w, height=h)
m = camera.TEST_POSES["initial"]
with binding.Renderer(sc) as r:
    r.set_option("kernel", 0); r.update(m); base = r.gather(); r.set_option("kernel", 1)
    for accel, smem in ((0, 1), (0, 0), (1, 1)):
        r.set_option("spheres_smem", smem)
        for v in (0, 3, 2, 5):
            try:
                r.set_option("accel", accel); r.set_option("variant", v); r.set_option("count", 1)
                r.update(m); c = r.counters(); same = bool((r.gather() == base).all())
                r.set_option("count", 0)
                ms = min(r.update(m) for _ in range(5))
                print("accel", accel, "spheres_smem", smem, "variant", v, "ms %.4f" % ms, "same", same, "sphere tests", c["tests_sphere"], "cubic tests", c["tests_cubic"], "exact", c["cubic_exact"], flush=True)
            except binding.RtError as e:
                print("accel", accel, "spheres_smem", smem, "variant", v, "n/a", str(e)[:60], flush=True)
