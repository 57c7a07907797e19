"""Queue kernel (kernel=2) against the exact per-pixel kernel (kernel=0) on the device: byte equality of the frame,
equality of the ray census, and kernel time per launch variant.
    python tools/gpu_queue_check.py [--variants 0,1,2,3,4] [--scenes a,b] [--width W --height H] [--synth WxH]"""
import argparse, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools"))
import numpy as np
from loader import load_package
pkg = load_package()
from cuda_ray_tracer_b200 import binding, camera, synth
ap = argparse.ArgumentParser()
ap.add_argument("--variants", default="0,1,2,3,4")
ap.add_argument("--scenes", default="20spheres,quadratic,cubic,cayley,clebsch,dingdong,monkey_saddle,reflection_test")
ap.add_argument("--width", type=int, default=3840)
ap.add_argument("--height", type=int, default=2160)
ap.add_argument("--synth", default="960x540")
ap.add_argument("--poses", default="initial,pose_a")
ap.add_argument("--cap", type=int, default=0)  # unused
a = ap.parse_args()
todo = [(n, pkg.load_from_file(os.path.join(ROOT, "scenes", n + ".yml")).with_overrides(width=a.width, height=a.height, max_reflections=4))
        for n in a.scenes.split(",") if n]
if a.synth:
    w, h = (int(v) for v in a.synth.split("x"))
    todo.append(("synthetic_%dx%d" % (w, h), synth.synthetic_scene(width=w, height=h)))
CENSUS = ("primary_rays", "reflection_rays", "shadow_rays", "hits", "lit")
bad = 0
for name, sc in todo:
    for pose in a.poses.split(","):
        m = camera.TEST_POSES[pose]
        row = {"scene": name, "pose": pose}
        with binding.Renderer(sc) as r:
            r.set_option("kernel", 0)
            row["exact_ms"] = round(min(r.update(m) for _ in range(2)), 4)
            base, cb = r.gather(), r.counters()
            r.set_option("kernel", 1)
            row["wave_ms"] = round(min(r.update(m) for _ in range(3)), 4)
            r.set_option("kernel", 2)
            for v in (int(x) for x in a.variants.split(",")):
                r.set_option("queue_variant", v)
                try:
                    ms = min(r.update(m) for _ in range(3))
                except binding.RtError as e:
                    row["q%d" % v] = "n/a"
                    continue
                img, c = r.gather(), r.counters()
                same = bool(np.array_equal(img, base))
                census = all(c[k] == cb[k] for k in CENSUS)
                row["q%d" % v] = round(ms, 4)
                if not (same and census):
                    bad += 1
                    diff = np.abs(img.astype(int) - base.astype(int)).max(axis=2)
                    row["q%d_diff" % v] = {"pixels": int((diff > 0).sum()), "max": int(diff.max()),
                                           "census": {k: (c[k], cb[k]) for k in CENSUS if c[k] != cb[k]}}
                if v == 0:
                    row["tests"] = {k: c[k] for k in ("tests_plane", "tests_sphere", "tests_quadric", "tests_cubic", "fp64_confirms")}
            r.set_option("count", 0)
            r.set_option("queue_variant", 0)
            row["q0_nocount"] = round(min(r.update(m) for _ in range(3)), 4)
        print(json.dumps(row), flush=True)
print("MISMATCHES", bad)
sys.exit(1 if bad else 0)
