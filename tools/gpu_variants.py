"""Kernel time of every wavefront-kernel variant (CTA size / occupancy) per scene, and byte equality among them."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools"))
import numpy as np
from loader import load_package
pkg = load_package()
from cuda_ray_tracer_b200 import binding, camera, synth
names = ["20spheres", "quadratic", "cubic", "cayley", "clebsch", "dingdong", "monkey_saddle", "reflection_test"]
todo = [(n, pkg.load_from_file(os.path.join(ROOT, "scenes", n + ".yml")).with_overrides(width=3840, height=2160, max_reflections=4)) for n in names]
todo.append(("synthetic_960x540", synth.synthetic_scene(width=960, height=540)))
variants = [int(v) for v in sys.argv[1].split(",")] if len(sys.argv) > 1 else [2, 3, 5]
opts = {}
for a in sys.argv[2:]:
    k, v = a.split("=")
    opts[k] = int(v)
m = camera.TEST_POSES["initial"]
for name, sc in todo:
    row = {"scene": name}
    base = None
    with binding.Renderer(sc) as r:
        for k, v in opts.items():
            r.set_option(k, v)
        r.set_option("kernel", 0)       # the exact per-pixel kernel: what every variant's frame must equal
        r.update(m)
        base = r.gather()
        r.set_option("kernel", 1)
        r.set_option("count", 0)        # time the product instantiation (counters off)
        for v in variants:
            r.set_option("variant", v)
            try:
                ms = min(r.update(m) for _ in range(6))
            except binding.RtError as e:
                row["v%d" % v] = "n/a"
                continue
            img = r.gather()
            if base is None:
                base = img
            row["v%d" % v] = round(ms, 4)
            row["same%d" % v] = bool(np.array_equal(img, base))
    print(json.dumps(row), flush=True)
