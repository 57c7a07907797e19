"""Device -> pinned host copy bandwidth of every GPU of the box, one at a time and all at once (one process, one stream
per GPU, 32 MiB copies) — the ceiling of any 'frame in host memory every frame' path.
    python tools/gpu_d2h_probe.py"""
import json, time
import torch
n = torch.cuda.device_count()
size = 32 << 20
src = [torch.empty(size, dtype=torch.uint8, device="cuda:%d" % i) for i in range(n)]
dst = [torch.empty(size, dtype=torch.uint8).pin_memory() for _ in range(n)]
streams = [torch.cuda.Stream(device=i) for i in range(n)]


def run(devs, reps=20):
    evs = {}
    for i in devs:
        with torch.cuda.device(i), torch.cuda.stream(streams[i]):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(reps):
                dst[i].copy_(src[i], non_blocking=True)
            b.record()
            evs[i] = (a, b)
    out = {}
    for i in devs:
        evs[i][1].synchronize()
        out[i] = round(size * reps / (evs[i][0].elapsed_time(evs[i][1]) * 1e-3) / 1e9, 1)
    return out


for i in range(n):
    run([i], 3)
alone = {}
for i in range(n):
    alone.update(run([i]))
together = run(list(range(n)))
half_a = run(list(range(n // 2))) if n > 1 else {}
half_b = run(list(range(n // 2, n))) if n > 1 else {}
print(json.dumps({"gpus": n, "GBps_alone": alone, "GBps_all_concurrent": together, "sum_all_concurrent": round(sum(together.values()), 1),
                  "GBps_first_half_concurrent": half_a, "GBps_second_half_concurrent": half_b}))
