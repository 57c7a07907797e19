"""First GPU contact: render every bundled scene with the exact kernel and diff against the oracle."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
from loader import load_package
from oracle import oracle_py
from parity import compare_frames

pkg = load_package()
from cuda_ray_tracer_b200 import binding, camera

names = ["20spheres", "quadratic", "cubic", "cayley", "clebsch", "dingdong", "monkey_saddle", "reflection_test"]
out = []
for name in names:
    sc = pkg.load_from_file(os.path.join(ROOT, "scenes", name + ".yml"))
    port = oracle_py.PortOracle(sc)
    with binding.Renderer(sc) as r:
        for pose in ("identity", "pose_a"):
            m = camera.TEST_POSES[pose]
            ref_rgb, cpu_ms, cnt = port.render_rows(m, threads=os.cpu_count(), count=True)
            ref = oracle_py.quantize_rgba8(ref_rgb)
            ms = min(r.update(m) for _ in range(3))
            img = r.gather()
            res = compare_frames(img, ref)
            c = r.counters()
            res.update(scene=name, pose=pose, gpu_ms=ms, cpu_ms_mt=cpu_ms, rays=c["rays"],
                       rays_match=(c["primary_rays"] == cnt["primary_rays"] and c["shadow_rays"] == cnt["shadow_rays"]
                                   and c["reflection_rays"] == cnt["reflection_rays"] and c["hits"] == cnt["hits"]
                                   and c["lit"] == cnt["lit"]),
                       classes=r.class_counts())
            print(json.dumps(res), flush=True)
            out.append(res)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "first_light.json"), "w"), indent=1)
