"""End-to-end frame time (rt_b200_update_to_host into pinned memory) against the number of row bands and their taper."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools"))
import numpy as np
import torch
from loader import load_package
pkg = load_package()
from cuda_ray_tracer_b200 import binding, camera
names = sys.argv[1].split(",") if len(sys.argv) > 1 else ["clebsch", "20spheres", "quadratic"]
m = camera.TEST_POSES["initial"]
host = torch.empty(3840 * 2160 * 4, dtype=torch.uint8).pin_memory()
for name in names:
    sc = pkg.load_from_file(os.path.join(ROOT, "scenes", name + ".yml")).with_overrides(width=3840, height=2160, max_reflections=4)
    row = {"scene": name}
    with binding.Renderer(sc) as r:
        r.set_option("count", 0)
        row["device_ms"] = round(min(r.update(m) for _ in range(5)), 4)
        for overlap in (1,):
            r.set_option("band_overlap", overlap)
            for bands, taper in ((8, 0), (12, 0), (8, 65), (8, 165), (8, 150), (8, 180), (12, 165), (12, 175), (0, -1)):
                r.set_option("host_bands", bands)
                r.set_option("band_taper", taper)
                for _ in range(3):
                    r.update_to_host(m, host.data_ptr(), host.numel())
                t0 = time.perf_counter()
                for _ in range(10):
                    r.update_to_host(m, host.data_ptr(), host.numel())
                row["o%d_b%d_t%d" % (overlap, bands, taper)] = round((time.perf_counter() - t0) * 100, 4)
    print(json.dumps(row), flush=True)
