"""One GPU, no communicator: kernel time of ONE rank's share of an N-rank strip split (N = 1, 2, 4, 8), per bundled
scene at 3840x2160 — what the render kernel itself costs at 1/N of the work, without any multi-GPU hand-shake.
    python tools/gpu_share_probe.py [option=value ...]"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools"))
from loader import load_package
pkg = load_package()
from cuda_ray_tracer_b200 import binding, camera
names = ["20spheres", "quadratic", "cubic", "cayley", "clebsch", "dingdong", "monkey_saddle", "reflection_test"]
opts = dict((a.split("=")[0], int(a.split("=")[1])) for a in sys.argv[1:])
m = camera.TEST_POSES["initial"]
total = {}
for name in names:
    sc = pkg.load_from_file(os.path.join(ROOT, "scenes", name + ".yml")).with_overrides(width=3840, height=2160, max_reflections=4)
    row = {"scene": name}
    for world in (1, 2, 4, 8):
        worst = 0.0
        for rank in sorted({0, world - 1, world // 2}):
            with binding.Renderer(sc, rank=rank, world_size=world) as r:
                r.set_option("count", 0)
                for k, v in opts.items():
                    r.set_option(k, v)
                for _ in range(3):
                    r.update(m)
                worst = max(worst, min(r.update(m) for _ in range(8)))
        row["n%d" % world] = round(worst, 4)
        total[world] = total.get(world, 0.0) + worst
    print(json.dumps(row), flush=True)
print(json.dumps({"sum_ms": {k: round(v, 4) for k, v in total.items()},
                  "speedup_vs_n1": {k: round(total[1] / v, 3) for k, v in total.items()}}), flush=True)
