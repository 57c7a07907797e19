"""torchrun -N: per-rank kernel time and pixel share of static strips vs dynamic row groups, frames started in lockstep
(host barrier before every frame) and free-running.  Prints one line per (scene, mode, sync)."""
import os, sys, time
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
names = sys.argv[1].split(",") if len(sys.argv) > 1 else ["quadratic", "clebsch", "reflection_test"]
variant = int(sys.argv[2]) if len(sys.argv) > 2 else 0
m = camera.TEST_POSES["initial"]
for name in names:
    sc = pkg.load_from_file(os.path.join(ROOT, "scenes", name + ".yml")).with_overrides(width=3840, height=2160, max_reflections=4)
    for dyn in (0, 1):
        r = binding.Renderer(sc, device=local, rank=rank, world_size=world)
        r.set_option("variant", variant)
        r.set_option("dynamic_tiles", dyn)
        ident = [binding.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ident, src=0)
        r.comm_init(ident[0])
        handle = [r.frame_export() if rank == 0 else None]
        dist.broadcast_object_list(handle, src=0)
        r.frame_attach(handle[0])
        r.set_option("count", 1)
        r.update(m)
        px = r.counters()["primary_rays"]
        r.set_option("count", 0)
        for lock in (1, 0):
            for _ in range(5):
                dist.barrier()
                r.update(m)
            ms = []
            dist.barrier()
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for _ in range(40):
                if lock:
                    dist.barrier()
                ms.append(r.update(m))
            dist.barrier()
            wall = (time.perf_counter() - t0) / 40 * 1e3
            t = torch.tensor([float(np.median(ms)), float(np.mean(ms)), float(px)], dtype=torch.float64, device="cuda")
            allt = [torch.zeros_like(t) for _ in range(world)]
            dist.all_gather(allt, t)
            if rank == 0:
                print("{} world {} dynamic {} lockstep {}: wall {:.4f} ms/frame; per rank median ms {} pixels {}".format(
                    name, world, dyn, lock, wall, [round(float(a[0]), 4) for a in allt], [int(a[2]) for a in allt]), flush=True)
        dist.barrier()
        r.close()
dist.destroy_process_group()
