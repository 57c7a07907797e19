"""torchrun -N ranks: per-rank render / gather milliseconds of every bundled 4K scene, for both frame-assembly paths."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools"))
import numpy as np
import torch
import torch.distributed as dist
from loader import load_package
pkg = load_package()
from cuda_ray_tracer_b200 import binding, camera

rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
names = ["20spheres", "quadratic", "cubic", "cayley", "clebsch", "dingdong", "monkey_saddle", "reflection_test"]
m = camera.TEST_POSES["initial"]
for mode in sys.argv[1:] or ["nccl", "p2p"]:
    rs = []
    for n in names:
        sc = pkg.load_from_file(os.path.join(ROOT, "scenes", n + ".yml")).with_overrides(width=3840, height=2160, max_reflections=4)
        r = binding.Renderer(sc, device=local, rank=rank, world_size=world)
        r.set_option("count", 0)
        ident = [binding.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ident, src=0)
        r.comm_init(ident[0])
        if mode == "p2p":
            h = [r.frame_export() if rank == 0 else None]
            dist.broadcast_object_list(h, src=0)
            r.frame_attach(h[0])
        rs.append(r)
    acc = np.zeros((len(names), 2))
    for it in range(13):
        for i, r in enumerate(rs):
            r.update(m)
            t = r.timing()
            if it >= 3:
                acc[i] += (t["render_ms"], t["gather_ms"])
    acc /= 10
    allacc = [None] * world
    dist.all_gather_object(allacc, acc)
    if rank == 0:
        for i, n in enumerate(names):
            print(mode, n, " ".join("r{}: {:.3f}+{:.3f}".format(k, a[i][0], a[i][1]) for k, a in enumerate(allacc)), flush=True)
    for r in rs:
        r.close()
    dist.barrier()
dist.destroy_process_group()
