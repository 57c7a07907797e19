"""torchrun -N ranks: every rank renders its strips and EVERY rank checks that the assembled frame is byte-identical to a
single-GPU render of the same frame on its own device — for both assembly paths: NCCL all-gather of packed strips, and
the fused path (kernels store straight into rank 0's frame through CUDA IPC, 4-byte all-reduce as barrier)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools"))
import numpy as np
import torch
import torch.distributed as dist
from loader import load_package
pkg = load_package()
from cuda_ray_tracer_b200 import binding, camera, synth

rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
ok = True
cases = [("reflection_test", pkg.load_from_file(os.path.join(ROOT, "scenes", "reflection_test.yml")).with_overrides(width=3840, height=2160, max_reflections=4), 8),
         ("dingdong", pkg.load_from_file(os.path.join(ROOT, "scenes", "dingdong.yml")).with_overrides(width=1000, height=562, max_reflections=4), 7),
         ("synthetic", synth.synthetic_scene(width=480, height=270), 8)]
for name, sc, strip in cases:
    m = camera.TEST_POSES["pose_a"]
    single = binding.Renderer(sc, device=local)
    ms1 = min(single.update(m) for _ in range(3))
    ref = single.gather()
    single.close()
    for mode in ("nccl", "p2p-allreduce", "p2p-static", "p2p"):
        r = binding.Renderer(sc, device=local, rank=rank, world_size=world)
        r.set_option("strip_rows", strip)
        ident = [binding.comm_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ident, src=0)
        r.comm_init(ident[0])
        if mode != "nccl":
            r.set_option("flag_barrier", 0 if mode == "p2p-allreduce" else 1)
            r.set_option("dynamic_tiles", 1 if mode == "p2p" else 0)   # p2p: row groups handed out by rank 0's counter (f3)
            handle = [r.frame_export() if rank == 0 else None]
            dist.broadcast_object_list(handle, src=0)
            r.frame_attach(handle[0])
        same = True
        for i in range(4):      # several frames: the shared frame is double-buffered
            ms = r.update(camera.TEST_POSES["pose_b" if i % 2 else "pose_a"])
        ms = r.update(m)
        if mode in ("p2p", "p2p-static"):
            dist.barrier()      # flag words: only rank 0 waits for every rank; the others need a host barrier to read
        multi = r.gather()
        dist.barrier()          # nobody starts the next frame while a peer still reads this one
        t = r.timing()
        same = bool(np.array_equal(multi, ref))
        ok &= same
        r.set_option("count", 1)
        r.update(m)
        px = r.counters()["primary_rays"]      # pixels THIS rank rendered: the dynamic split moves work between ranks
        r.set_option("count", 0)
        dist.barrier()
        print("rank {} {} world {} strips {} {}: identical={} multi {:.3f} ms (render {:.3f} gather {:.3f}) single {:.3f} ms, pixels here {}".format(
            rank, name, world, strip, mode, same, ms, t["render_ms"], t["gather_ms"], ms1, px), flush=True)
        r.close()
        dist.barrier()
    # host frame shared by all ranks: every rank's copy engine writes its own strips (no NVLink, no communicator)
    nbytes = sc.px_width * sc.px_height * 4
    path = [binding.shared_frame_path(nbytes) if rank == 0 else None]
    ptr = binding.host_alloc_shared(path[0], nbytes, True) if rank == 0 else 0
    dist.broadcast_object_list(path, src=0)
    if rank != 0:
        ptr = binding.host_alloc_shared(path[0], nbytes, False)
    for bands in (0, 1, 3):
        r = binding.Renderer(sc, device=local, rank=rank, world_size=world)
        r.set_option("strip_rows", strip)
        r.set_option("host_bands", bands)
        for i in range(4):
            r.update_to_host(camera.TEST_POSES["pose_b" if i % 2 else "pose_a"], ptr, nbytes)
        ms = r.update_to_host(m, ptr, nbytes)
        same = True
        if rank == 0:           # the frame is complete when rank 0's call returns
            same = bool(np.array_equal(binding.host_view(ptr, sc.px_height, sc.px_width), ref))
        ok &= same
        print("rank {} {} world {} strips {} host-shared bands {}: identical={} {:.3f} ms".format(
            rank, name, world, strip, bands, same, ms), flush=True)
        dist.barrier()          # rank 0 has compared before anybody renders into the block again
        r.close()
    dist.barrier()
    binding.host_free_shared(ptr)
flag = torch.tensor([1 if ok else 0], device="cuda")
dist.all_reduce(flag, op=dist.ReduceOp.MIN)
dist.destroy_process_group()
if rank == 0:
    print("MULTI_GPU_CHECK", "OK" if flag.item() == 1 else "FAILED")
sys.exit(0 if flag.item() == 1 else 1)
