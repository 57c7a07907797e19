"""CPU, world_size 2 over gloo: the host-side logic of the multi-GPU split (strip interleave, equal-sized
contributions, all-gather, un-interleave).  The per-rank "render" here is the oracle (test infrastructure)
standing in for the GPU kernel; the GPU version of the same check is tests/test_parity_gpu.py::test_strip_split."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, strip_rows, out_path):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    for p in (ROOT, os.path.join(ROOT, "tools")):
        sys.path.insert(0, p)
    from loader import load_package
    pkg = load_package()
    from cuda_ray_tracer_b200 import binding, camera
    from oracle import oracle_py
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        sc = pkg.load_from_file(os.path.join(ROOT, "scenes", "reflection_test.yml")).with_overrides(
            width=64, height=45, max_reflections=4)
        port_oracle = oracle_py.PortOracle(sc)
        rows = binding.strip_layout(sc.px_height, strip_rows, rank, world)
        local = np.zeros((len(rows), sc.px_width, 4), dtype=np.uint8)
        m = camera.TEST_POSES["pose_a"]
        for i, y in enumerate(rows):
            if y >= 0:
                rgb, _ = port_oracle.render_rows(m, int(y), int(y) + 1)
                local[i] = oracle_py.quantize_rgba8(rgb)[0]
        mine = torch.from_numpy(local)
        gathered = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(gathered, mine)                       # equal sizes on every rank, as NCCL all-gather needs
        frame = np.zeros((sc.px_height, sc.px_width, 4), dtype=np.uint8)
        for r in range(world):
            for i, y in enumerate(binding.strip_layout(sc.px_height, strip_rows, r, world)):
                if y >= 0:
                    frame[y] = gathered[r][i].numpy()
        full = oracle_py.quantize_rgba8(port_oracle.render(m))
        ok = bool(np.array_equal(frame, full))                # byte-identical to the unsplit frame
        flags = [None] * world
        dist.all_gather_object(flags, ok)
        if rank == 0:
            with open(out_path, "w") as f:
                f.write("ok" if all(flags) else "mismatch")
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("strip_rows", [8, 7])
def test_two_rank_strip_gather(tmp_path, strip_rows):
    from cuda_ray_tracer_b200 import build  # noqa: F401  (conftest put the package on the path)
    out = tmp_path / "result.txt"
    port = 29500 + (os.getpid() % 2000) + strip_rows
    mp.spawn(_worker, args=(2, port, strip_rows, str(out)), nprocs=2, join=True)
    assert out.read_text() == "ok"
