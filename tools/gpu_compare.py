"""Wavefront kernel (1) vs exact kernel (0): byte equality and kernel times, 4K bundled scenes + small synthetic."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools"))
import numpy as np
from loader import load_package
pkg = load_package()
from cuda_ray_tracer_b200 import binding, camera, synth

names = ["20spheres", "quadratic", "cubic", "cayley", "clebsch", "dingdong", "monkey_saddle", "reflection_test"]
todo = [(n, pkg.load_from_file(os.path.join(ROOT, "scenes", n + ".yml")).with_overrides(width=3840, height=2160, max_reflections=4)) for n in names]
sw, sh = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (640, 360)
todo.append(("synthetic_%dx%d" % (sw, sh), synth.synthetic_scene(width=sw, height=sh)))
skip_exact_synth = len(sys.argv) > 3
res = []
for name, sc in todo:
    for pose in ("initial", "pose_a"):
        m = camera.TEST_POSES[pose]
        row = {"scene": name, "pose": pose}
        frames = {}
        with binding.Renderer(sc) as r:
            for k in (1, 0):
                if k == 0 and name.startswith("synthetic") and skip_exact_synth:
                    continue
                r.set_option("kernel", k)
                ms = min(r.update(m) for _ in range(3 if not name.startswith("synthetic") else 1))
                frames[k] = r.gather()
                c = r.counters()
                row["ms_k%d" % k] = round(ms, 4)
                row["cnt_k%d" % k] = {x: c[x] for x in ("rays", "tests_sphere", "tests_cubic", "fp64_confirms", "hits", "lit")}
        if 0 in frames:
            d = np.abs(frames[0].astype(np.int16) - frames[1].astype(np.int16)).max(axis=-1)
            row["identical"] = bool((d == 0).all()); row["diff_px"] = int((d > 0).sum())
        print(json.dumps(row), flush=True)
        res.append(row)
json.dump(res, open(os.path.join(ROOT, "gpurun_out", "compare.json"), "w"), indent=1)
