"""Where the time of ONE wavefront-kernel launch goes, per CTA (diagnostic build: python cuda-ray-tracer_b200/build.py --timeline).
Renders rank 0's share of an N-rank strip split on one GPU (N = 1: the whole frame) and prints, in microseconds from the first CTA's
entry: when CTAs enter (launch ramp), how long the prologue takes, when CTAs first find no tile left, when they leave the loop.
    RT_B200_LIB=$PWD/cuda-ray-tracer_b200/librt_b200_timeline.so python tools/gpu_timeline.py [scene ...] [--world N]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools"))
import numpy as np
from loader import load_package
pkg = load_package()
from cuda_ray_tracer_b200 import binding, camera
args = [a for i, a in enumerate(sys.argv[1:], 1) if not a.startswith("--") and sys.argv[i - 1] != "--world"]
world = int(sys.argv[sys.argv.index("--world") + 1]) if "--world" in sys.argv else 8
names = args or ["20spheres", "dingdong", "quadratic"]
m = camera.TEST_POSES["initial"]


def pct(a, q):
    return float(np.percentile(a, q))


for name in names:
    sc = pkg.load_from_file(os.path.join(ROOT, "scenes", name + ".yml")).with_overrides(width=3840, height=2160, max_reflections=4)
    with binding.Renderer(sc, rank=0, world_size=world) as r:
        r.set_option("count", 0)
        r.set_option("timeline", 1)
        for _ in range(4):
            ms = r.update(m)
        t = r.debug_timeline().astype(np.float64)
    t = t[t[:, 0] > 0]
    if len(t) == 0:
        print(name, "no stamps: not the timeline build of the library?")
        continue
    t0 = t[:, 0].min()
    entry, pro, dry, loop_end, kexit = [(t[:, k] - t0) / 1e3 for k in range(5)]
    waves, sm, tiles = t[:, 5], t[:, 6], t[:, 7]
    end = kexit.max()
    busy = (loop_end - entry)
    print("{} world {} (rank 0's share): event time {:.1f} us, {} CTAs on {} SMs, last exit at {:.1f} us".format(
        name, world, ms * 1e3, len(t), len(set(sm.tolist())), end))
    print("   entry          : median {:6.1f}  p90 {:6.1f}  max {:6.1f}".format(pct(entry, 50), pct(entry, 90), entry.max()))
    print("   prologue length: median {:6.1f}  p90 {:6.1f}  max {:6.1f}".format(pct(pro - entry, 50), pct(pro - entry, 90), (pro - entry).max()))
    print("   first 'no tile': min    {:6.1f}  median {:6.1f}  max {:6.1f}".format(dry.min(), pct(dry, 50), dry.max()))
    print("   loop exit      : min    {:6.1f}  median {:6.1f}  p90 {:6.1f}  max {:6.1f}".format(loop_end.min(), pct(loop_end, 50), pct(loop_end, 90), loop_end.max()))
    print("   tiles per CTA  : min {:.0f} median {:.0f} max {:.0f};  waves per CTA: median {:.0f} max {:.0f}".format(tiles.min(), pct(tiles, 50), tiles.max(), pct(waves, 50), waves.max()))
    idle_tail = float((end - loop_end).mean())
    print("   mean idle per CTA slot: before entry {:.1f} us, prologue {:.1f} us, after its loop exit until the last CTA is done {:.1f} us  (of {:.1f})".format(
        float(entry.mean()), float((pro - entry).mean()), idle_tail, end))
    late = np.argsort(loop_end)[-5:]
    print("   the five last CTAs: " + "; ".join("exit {:.1f} (tiles {:.0f}, waves {:.0f}, dry at {:.1f})".format(loop_end[i], tiles[i], waves[i], dry[i]) for i in late))
