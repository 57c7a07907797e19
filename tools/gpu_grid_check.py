"""Sphere grid (option accel) on the GPU: frame equality with the brute-force loops and kernel time per grid density /
kernel variant, on the synthetic C5 scene (and any bundled scene with spheres).

    python tools/gpu_grid_check.py [width height] > gpurun_out/grid_check.jsonl
"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools"))
import numpy as np
from loader import load_package
pkg = load_package()
from cuda_ray_tracer_b200 import binding, camera, synth

w, h = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (3840, 2160)
todo = [("synthetic_{}x{}".format(w, h), synth.synthetic_scene(width=w, height=h)),
        ("20spheres_4k", pkg.load_from_file(os.path.join(ROOT, "scenes", "20spheres.yml")).with_overrides(width=3840, height=2160, max_reflections=4))]
m = camera.TEST_POSES["initial"]
for name, sc in todo:
    with binding.Renderer(sc) as r:
        r.set_option("count", 1)
        r.set_option("accel", 0)
        ms = min(r.update(m) for _ in range(2))
        base = r.gather()
        c = r.counters()
        print(json.dumps({"scene": name, "accel": 0, "ms": round(ms, 4), "rays": c["rays"], "tests_sphere": c["tests_sphere"],
                          "confirms": c["fp64_confirms"], "tests_cubic": c["tests_cubic"], "cubic_exact": c["cubic_exact"]}), flush=True)
        r.set_option("accel", 1)
        for variant, density in ((0, 0), (3, 0), (0, 50), (0, 400), (0, 1000), (3, 400)):
            r.set_option("variant", variant)
            r.set_option("grid_density", density)
            r.set_option("count", 1)
            r.update(m)
            c = r.counters()
            same = bool(np.array_equal(r.gather(), base))
            r.set_option("count", 0)
            ms = min(r.update(m) for _ in range(3))
            print(json.dumps({"scene": name, "accel": 1, "variant": variant, "density": density, "grid": r.grid_info(), "ms": round(ms, 4),
                              "same_frame": same, "rays": c["rays"], "tests_sphere": c["tests_sphere"], "confirms": c["fp64_confirms"],
                              "tests_cubic": c["tests_cubic"], "cubic_exact": c["cubic_exact"]}), flush=True)
