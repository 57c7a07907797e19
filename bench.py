#!/usr/bin/env python
"""bench.py — headline benchmark of the B200-native frame update (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload bundled4k|synthetic]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W

Workload `bundled4k` (default, the configuration the metric is quoted on): one STEP = one 3840x2160,
4-bounce frame of each of the eight bundled scenes (the reference's initial camera pose), i.e. eight
rt_b200_update calls.  `value` = rays cast per second over the whole job (primary + reflection + shadow
rays), inputs resident on the device.  `e2e` = the same through the C ABI with HOST
buffers: camera matrix from host memory in, finished RGBA8 frame copied to pinned host memory out, every
frame.  With N > 1 ranks (one process per GPU) every rank renders its interleaved strips of every frame and
its render kernel stores the pixels straight into the frame assembled on rank 0 (CUDA IPC over NVLink, flag words
as the frame barrier; `--gather nccl` selects the NCCL all-gather of packed strips instead); time is the max over
ranks.  Before anything is timed every frame of the workload is CHECKED (config.frame_check): against the exact
per-pixel kernel, against the reference CPU build on a pixel sample, and — N > 1 — the assembled frame against a
1-GPU render; a mismatch ends the run with exit code 3.

`--impl reference` times the reference's own CPU implementation (oracle/_ref = update-cpu.cpp compiled
unmodified, else the C restatement) on the host cores, on a bounded row sample of the same workload.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "tools")):
    if p not in sys.path:
        sys.path.insert(0, p)

import numpy as np  # noqa: E402

BUNDLED = ["20spheres", "quadratic", "cubic", "cayley", "clebsch", "dingdong", "monkey_saddle", "reflection_test"]
WIDTH, HEIGHT, BOUNCES = 3840, 2160, 4
METRIC, UNIT = "Mrays/s", "Mrays/s"

# Canonical algorithmic flops: per (ray, object) test by structural class, per primary ray,
# per hit, per unshadowed light evaluation, per bounce.
FLOPS = {"tests_plane": 13, "tests_sphere": 38, "tests_quadric": 81, "tests_cubic": 320,
         "primary_rays": 50, "hits": 70, "lit": 35, "reflection_rays": 12}


def canonical_flops(c: dict) -> float:
    return float(sum(c[k] * w for k, w in FLOPS.items()))


def load_peaks() -> dict:
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            p = json.load(f)
        p["_source"] = "measured"
        return p
    return {"hbm_gbs": 6650.0, "sm_max_mhz": 1965.0, "_source": "fallback"}


def ncu_summaries(labels):
    """Pipe utilisation and DRAM traffic of the render kernel per scene, from the committed ncu --set full summaries
    (profiles/r2_final_wave_<scene>.txt, written by tools/ncu_summary.py from captures of this code)."""
    import re
    out = {}
    for label in labels:
        path = os.path.join(ROOT, "profiles", "r2_final_wave_{}.txt".format(label))
        try:
            t = open(path).read()
        except OSError:
            continue

        def g(key):
            m = re.search(re.escape(key) + r"\s+([0-9.]+)\s*(\w*)", t)
            return (float(m.group(1)), m.group(2)) if m else (None, "")
        scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        rd, ru = g("dram__bytes_read.sum")
        wr, wu = g("dram__bytes_write.sum")
        dur, du = g("gpu__time_duration.sum")
        out[label] = {
            "fma_pipe_pct": g("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active")[0],
            "fp64_pipe_pct": g("sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active")[0],
            "issue_active_pct": g("smsp__issue_active.avg.pct_of_peak_sustained_active")[0],
            "lanes_per_instruction": g("smsp__thread_inst_executed_per_inst_executed.ratio")[0],
            "dram_bytes": (rd * scale.get(ru, 1.0) + wr * scale.get(wu, 1.0)) if rd is not None and wr is not None else None,
            "ncu_ms": dur * (1.0 if du == "ms" else 1e-3) if dur is not None else None,
        }
    return out


def fp32_peak_tflops(n_sms: int, sm_mhz: float) -> float:
    """FP32 SIMT peak: SMs x 128 lanes x 2 flop (FMA) x clock."""
    return n_sms * 128 * 2 * sm_mhz * 1e6 / 1e12


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.lines:
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for n, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        # samples under load = the upper half (the sampler also sees idle gaps between steps)
        sm_sorted = sorted(sm)
        load = sm_sorted[len(sm_sorted) // 2:] if sm_sorted else []
        return {"sm_mhz": float(np.median(load)) if load else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def dist_env():
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    return rank, local, world


def build_workload(pkg, name: str):
    """Returns [(label, Scene)] and a description."""
    if name == "bundled4k":
        scenes = []
        for s in BUNDLED:
            sc = pkg.load_from_file(os.path.join(ROOT, "scenes", s + ".yml")).with_overrides(
                width=WIDTH, height=HEIGHT, max_reflections=BOUNCES)
            scenes.append((s, sc))
        return scenes, "8 bundled scenes @ {}x{}, max_reflections {}, initial camera".format(WIDTH, HEIGHT, BOUNCES)
    if name == "synthetic":
        from cuda_ray_tracer_b200 import synth
        sc = synth.synthetic_scene()
        return [("synth_4096s_64c_16l", sc)], "synthetic 4096 spheres + 64 cubics + 16 point lights @ 7680x4320, 4 bounces"
    raise SystemExit("unknown workload " + name)


# --------------------------------------------------------------------------------------------------
def run_reference(args):
    """The reference's own CPU implementation on the host cores, bounded sample of the same workload."""
    rank, _, world = dist_env()
    if rank != 0:
        return
    from loader import load_package
    pkg = load_package()
    from cuda_ray_tracer_b200 import camera
    from oracle import oracle_py
    oracle_py.build()
    scenes, desc = build_workload(pkg, args.workload)
    cores = os.cpu_count() or 1
    kind = "reference" if oracle_py.have_ref() else "port"
    m = camera.TEST_POSES["initial"]
    # each step a bounded sample: about 1.5 s of wall time on all host threads (every 2nd row for the bundled step)
    ystep = args.ref_row_step or max(2, pick_row_step(oracle_py, scenes, m, cores, 1.5))
    # ray census of the sampled rows (untimed, via the counting restatement)
    rays_sample = 0
    for _, sc in scenes:
        _, _, cnt = oracle_py.PortOracle(sc).render_rows(m, 0, sc.px_height, ystep, threads=cores, count=True)
        rays_sample += cnt["primary_rays"] + cnt["reflection_rays"] + cnt["shadow_rays"]

    def step() -> float:
        """Seconds inside the pixel loops only — the reference's own timing convention (update-cpu.cpp:124-138);
        its init_update (a 100 MB float buffer fill per 4K scene) is left out."""
        ms_total = 0.0
        for _, sc in scenes:
            if kind == "reference":
                ms_total += oracle_py.RefCpu(sc).render_rows(m, 0, sc.px_height, ystep, threads=cores)[1]
            else:
                ms_total += oracle_py.PortOracle(sc).render_rows(m, 0, sc.px_height, ystep, threads=cores)[1]
        return ms_total / 1e3

    for _ in range(args.warmup):
        step()
    dt = sum(step() for _ in range(args.steps)) / args.steps
    value = rays_sample / dt / 1e6
    h0 = scenes[0][1].px_height
    sample = "every {}th image row of each frame ({} rows per {}), {} host threads".format(
        ystep, len(range(0, h0, ystep)), h0, cores)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": args.workload, "description": desc, "sampled": sample},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------------------------------
def run_b200(args):
    import torch
    import torch.distributed as dist
    from loader import load_package
    pkg = load_package()
    from cuda_ray_tracer_b200 import binding, camera

    rank, local, world = dist_env()
    if world != args.gpus and world != 1:
        raise SystemExit("--gpus {} but WORLD_SIZE {}".format(args.gpus, world))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the B200 path has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    binding.load_library()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    scenes, desc = build_workload(pkg, args.workload)
    m = np.ascontiguousarray(camera.TEST_POSES["initial"])
    renderers = []
    fused = world > 1 and args.gather == "p2p"
    for label, sc in scenes:
        r = binding.Renderer(sc, device=local, rank=rank, world_size=world)
        r.set_option("kernel", args.kernel)
        r.set_option("accel", args.accel)
        r.set_option("variant", args.variant)
        if world > 1:
            r.set_option("strip_rows", args.strip_rows)
            r.set_option("dynamic_tiles", args.dynamic)
            ident = [binding.comm_unique_id() if rank == 0 else None]
            dist.broadcast_object_list(ident, src=0)
            r.comm_init(ident[0])
            if fused:   # every rank's kernel stores its pixels straight into rank 0's frame (CUDA IPC over NVLink)
                handle = [r.frame_export() if rank == 0 else None]
                dist.broadcast_object_list(handle, src=0)
                r.frame_attach(handle[0])
        renderers.append((label, sc, r))

    frame_bytes = max(sc.px_width * sc.px_height * 4 for _, sc in scenes)
    host_frame = torch.empty(frame_bytes, dtype=torch.uint8).pin_memory()
    host_cam = torch.from_numpy(m.copy()).pin_memory()
    shared_ptr = 0
    if world > 1 and args.e2e == "shared":
        # one host frame mapped (and page-locked) by every rank: each GPU's copy engine delivers its own strips
        path = [binding.shared_frame_path(frame_bytes) if rank == 0 else None]
        if rank == 0:
            shared_ptr = binding.host_alloc_shared(path[0], frame_bytes, True)
        dist.broadcast_object_list(path, src=0)
        if rank != 0:
            shared_ptr = binding.host_alloc_shared(path[0], frame_bytes, False)

    def step_e2e():
        """The user-facing call with HOST buffers: camera matrix in, finished frame in pinned host memory out."""
        for _, sc, r in renderers:
            nbytes = sc.px_width * sc.px_height * 4
            if world == 1:
                r.update_to_host(host_cam.numpy(), host_frame.data_ptr(), nbytes)
            elif shared_ptr:
                r.update_to_host(host_cam.numpy(), shared_ptr, nbytes)
            else:
                r.update(host_cam.numpy())
                if rank == 0:           # the frame is assembled on GPU 0 (north_star); rank 0 hands it to the host
                    r.gather_into(host_frame.data_ptr(), nbytes)

    # ---- frame check (untimed): every frame this run times is first compared with the exact per-pixel kernel, with
    # the reference CPU build on a pixel sample, and — at N > 1 — rank 0's ASSEMBLED frame with a 1-GPU render of the
    # same frame (fused gather, and once through the NCCL all-gather path).  A mismatch ends the run with exit code 3.
    frame_check = None
    if not args.no_frame_check:
        frame_check = check_frames(args, pkg, binding, renderers, m, rank, local, world, barrier, dist if world > 1 else None)

    # ---- census (untimed): rays and tests of one step, with the kernel's counters on -----------------
    # The metric counts the rays the REFERENCE casts for these frames: census with culling off.  The
    # timed runs use the product default (culling on): shadow rays that provably add 0 to the pixel are not traced;
    # that smaller number is reported beside it as rays_traced_per_step.
    local_counts, traced_counts = {}, {}
    for label, _, r in renderers:
        r.set_option("count", 1)
        r.set_option("cull_backfacing", 0)
        r.set_option("accel", 0)        # the census is the reference's brute-force loops, whatever the timed mode is
        r.update(m)
        local_counts[label] = r.counters()
        r.set_option("cull_backfacing", 1 if args.cull else 0)
        r.set_option("accel", args.accel)
        r.update(m)
        traced_counts[label] = r.counters()
        r.set_option("count", 0)        # counters are instrumentation, not part of the frame: off while timing

    # ---- device-resident timing ------------------------------------------------------------------
    for _ in range(max(args.warmup, 3)):
        for _, _, r in renderers:
            r.update(m)
    sampler = ClockSampler(local)
    per_scene = {label: {"render_ms": 0.0, "gather_ms": 0.0} for label, _, _ in renderers}
    launches = 0
    if rank == 0:
        sampler.start()                 # before the barrier: forking nvidia-smi must not delay rank 0's first frame
        time.sleep(0.05)
    barrier()
    t0 = time.perf_counter()
    ev_ms = 0.0
    for _ in range(args.steps):
        for label, _, r in renderers:
            ev_ms += r.update(m)        # CUDA events on the library's render stream: first launch -> frame complete
            t = r.timing()
            per_scene[label]["render_ms"] += t["render_ms"]
            per_scene[label]["gather_ms"] += t["gather_ms"]
            launches += r.launches()
    barrier()
    dt = time.perf_counter() - t0
    clocks = sampler.stop() if rank == 0 else None

    # ---- end-to-end timing (host buffers) ----------------------------------------------------------
    for _ in range(max(args.warmup, 3)):      # also lets update_to_host settle its band shape per scene (three frames)
        step_e2e()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step_e2e()
    barrier()
    dt_e2e = time.perf_counter() - t0

    keys = list(binding.COUNTER_FIELDS)
    mat = torch.tensor([[local_counts[label][k] for k in keys] for label, _, _ in renderers], dtype=torch.int64,
                       device="cuda")
    times = torch.tensor([dt, dt_e2e, ev_ms / 1e3], dtype=torch.float64, device="cuda")
    kern = torch.tensor([[per_scene[label]["render_ms"], per_scene[label]["gather_ms"]] for label, _, _ in renderers],
                        dtype=torch.float64, device="cuda")
    nl = torch.tensor([launches, sum(c["primary_rays"] + c["reflection_rays"] + c["shadow_rays"]
                                     for c in traced_counts.values())], dtype=torch.int64, device="cuda")
    if world > 1:
        dist.all_reduce(mat, op=dist.ReduceOp.SUM)
        dist.all_reduce(times, op=dist.ReduceOp.MAX)
        dist.all_reduce(kern, op=dist.ReduceOp.MAX)
        dist.all_reduce(nl, op=dist.ReduceOp.SUM)
    mat, times, kern = mat.cpu().numpy(), times.cpu().numpy(), kern.cpu().numpy() / args.steps
    launches, rays_traced = int(nl[0].item()), int(nl[1].item())

    if rank == 0:
        peaks = load_peaks()
        props = torch.cuda.get_device_properties(local)
        n_sms = props.multi_processor_count
        fp32_peak = fp32_peak_tflops(n_sms, float(peaks.get("sm_max_mhz", 1965.0)))
        totals = {k: int(mat[:, i].sum()) for i, k in enumerate(keys)}
        rays_step = totals["primary_rays"] + totals["reflection_rays"] + totals["shadow_rays"]
        flops_step = canonical_flops(totals)
        ms_step = times[2] / args.steps * 1e3          # CUDA events, max over ranks
        ms_wall = times[0] / args.steps * 1e3          # host clock around the same loop (barrier + sync on both sides)
        ms_e2e = times[1] / args.steps * 1e3
        render_ms_sum = float(kern[:, 0].sum())          # per step, max over ranks per scene
        scenes_out = {}
        for i, (label, sc, _) in enumerate(renderers):
            c = {k: int(mat[i, j]) for j, k in enumerate(keys)}
            rays = c["primary_rays"] + c["reflection_rays"] + c["shadow_rays"]
            fl = canonical_flops(c)
            rms = float(kern[i, 0])
            scenes_out[label] = {
                "ms_frame_kernel": rms, "ms_gather": float(kern[i, 1]), "mrays_s_kernel": rays / rms / 1e3,
                "rays": rays, "gflop": fl / 1e9, "tflops": fl / rms / 1e9, "fp32_frac": fl / rms / 1e9 / (fp32_peak * world),
                "fb_write_gbs": sc.px_width * sc.px_height * 4 / rms / 1e6,
            }
        achieved = flops_step / render_ms_sum / 1e9 / world    # TFLOP/s per GPU over the render kernels
        grid_mode = args.accel != 0 and any(len(sc.objects) >= 64 for _, sc, _ in renderers)
        ncu_labels = [label if args.workload == "bundled4k" else ("synthetic_a1" if grid_mode else "synthetic_a0") for label, _, _ in renderers]
        ncu = ncu_summaries(ncu_labels)
        weights = {nl: float(kern[i, 0]) for i, nl in enumerate(ncu_labels)}

        def ncu_w(key):
            have = [(ncu[k][key], weights[k]) for k in ncu if ncu[k].get(key) is not None]
            return sum(v * w for v, w in have) / sum(w for _, w in have) if have else None
        tr = [ncu[k]["dram_bytes"] for k in ncu if ncu[k].get("dram_bytes") is not None]
        traffic = int(sum(tr) / len(tr)) if tr else None     # dram read + write per launch, mean over the captured frames
        # the same flops split by the pipe that executes them here: sphere tests run through the FP32 filter, every
        # other class (and hit / shading geometry) is FP64 in the reference's operation order, at half the FP32 rate
        fl32 = float(totals["tests_sphere"] * FLOPS["tests_sphere"] + totals["lit"] * FLOPS["lit"])
        fl64 = flops_step - fl32
        t_min_ms = (fl32 / fp32_peak + fl64 / (fp32_peak / 2.0)) / 1e9 / world
        line = {
            "metric": METRIC, "value": rays_step / ms_step / 1e3, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_step, "ms_per_step_wall": ms_wall, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": args.workload, "description": desc, "frames_per_step": len(renderers),
                       "rays_per_step": rays_step, "rays_traced_per_step": rays_traced,
                       "ray_count": "rays_per_step = rays the reference casts (metric); rays_traced_per_step = rays this kernel "
                                    "traces: shadow rays towards lights with Lambert factor 0 at the hit add exactly 0 and are skipped", "kernel": args.kernel,
                       "accel": {"mode": args.accel, "sphere_tests_examined_rank0": int(sum(c["tests_sphere"] for c in traced_counts.values())),
                                 "what": "0 = brute force over all objects like the reference (the roofline is quoted on this); 1 = sphere "
                                         "class through the uniform grid (SURVEY §8 f4): identical frames, canonical flops still those "
                                         "of the brute-force census, so roofline.frac is reference-equivalent work and can exceed 1"},
                       "parallelism": "strips{}x{}rows,{}".format(world, args.strip_rows, ("p2p-store+flags,dynamic-groups" if args.dynamic else "p2p-store+flags,static-strips") if fused else
                                                                  "nccl-allgather") if world > 1 else "1gpu",
                       "timing": "CUDA events on the render stream per frame, summed; max over ranks",
                       "ms_per_step_wall": ms_wall, "frame_check": frame_check,
                       "value_traced": rays_traced / ms_step / 1e3,
                       "l2": "outputs ({} MB per step) exceed the 126 MB L2; the scene (KBs) is cache-resident by design"
                             .format(len(renderers) * frame_bytes // 1000000)},
            "e2e": {"value": rays_step / ms_e2e / 1e3, "unit": UNIT, "ms_per_step": ms_e2e,
                    "h2d_bytes_per_step": len(renderers) * 128,
                    "how": "rt_b200_update_to_host: banded render with the D2H copy of finished bands overlapped"
                           if world == 1 else
                           "rt_b200_update_to_host on every rank into one shared page-locked host frame: each GPU copies its own strips over its own PCIe link"
                           if args.e2e == "shared" else
                           "rt_b200_update on every rank + rt_b200_gather to pinned host memory on rank 0",
                    "d2h_bytes_per_step": int(sum(sc.px_width * sc.px_height * 4 for _, sc, _ in renderers))},
            "gpu_launches": launches,
            "clocks": clocks,
            "roofline": {"bound": "fp32", "achieved": achieved if not grid_mode else None, "peak": fp32_peak, "unit": "TFLOP/s",
                         "frac": achieved / fp32_peak if not grid_mode else None, "traffic": traffic,
                         "reference_equivalent_rate": {"tflops": achieved, "of_fp32_peak": achieved / fp32_peak,
                                                       "what": "canonical flops of the reference's brute-force loops / kernel time: a work rate, "
                                                               "not a pipe utilisation — with the sphere grid the kernel skips most of that work"} if grid_mode else None,
                         "ncu_fma_pipe_pct": ncu_w("fma_pipe_pct"), "ncu_fp64_pipe_pct": ncu_w("fp64_pipe_pct"),
                         "ncu_issue_active_pct": ncu_w("issue_active_pct"), "ncu_lanes_per_instruction": ncu_w("lanes_per_instruction"),
                         "ncu_per_scene": ncu, "ncu_what": "sm__inst_executed_pipe_{fma,fp64}, smsp__issue_active (% of peak sustained active) from the committed "
                                                           "captures profiles/r2_final_wave_*.txt, averaged over the step weighted by each frame's kernel time",
                         "peak_source": "148 SMs x 128 lanes x 2 x sm_max_mhz of MEASURED_PEAKS.json ({})".format(peaks["_source"]),
                         "algorithmic_gflop_per_step": flops_step / 1e9, "kernel_ms_per_step": render_ms_sum,
                         "blended": {"frac": t_min_ms / render_ms_sum, "fp32_gflop": fl32 / 1e9, "fp64_gflop": fl64 / 1e9,
                                     "what": "time at peak (fp32 flops / FP32 peak + fp64 flops / (FP32 peak / 2)) / kernel time"},
                         "note": "canonical flops of SURVEY.md §8d over the render kernels' CUDA-event time; "
                                 "the path is compute-bound (not hbm/tensor)",
                         "hbm": {"achieved": len(renderers) * frame_bytes / render_ms_sum / 1e6, "peak": peaks["hbm_gbs"],
                                 "unit": "GB/s", "what": "framebuffer write only"}},
            "per_scene": scenes_out,
        }
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(pkg, scenes, args)
        print(json.dumps(line), flush=True)

    for _, _, r in renderers:
        r.close()
    if shared_ptr:
        barrier()
        binding.host_free_shared(shared_ptr)
    if world > 1:
        dist.destroy_process_group()


def check_frames(args, pkg, binding, renderers, m, rank, local, world, barrier, dist) -> dict:
    """Untimed.  oracle/ is used here as the checker only (never timed, never shipped)."""
    import numpy as np
    out = {"scenes": len(renderers), "vs_exact_kernel": "identical", "vs_ref_pixels": 0, "outliers": 0, "max_diff": 0,
           "assembled_equals_1gpu": None if world == 1 else True, "nccl_allgather_equals_1gpu": None,
           "oracle": None, "mismatches": []}
    oracle_py = None
    if rank == 0:
        from oracle import oracle_py
        oracle_py.build()
        out["oracle"] = "reference build (oracle/_ref)" if oracle_py.have_ref() else "restatement (oracle/rt_oracle.c)"
    n_px = 1000 if args.workload == "bundled4k" else 1200
    for i, (label, sc, r) in enumerate(renderers):
        r.set_option("kernel", args.kernel)
        r.set_option("accel", args.accel)
        r.set_option("cull_backfacing", 1 if args.cull else 0)
        r.update(m)
        img = None
        if rank == 0:
            img = r.gather()            # world > 1: the frame assembled on rank 0 (rank 0's update returns when it is whole)
        barrier()
        nccl_img = None
        if world > 1 and i == 0:        # the all-gather assembly path, once: a second context without the fused frame
            r2 = binding.Renderer(sc, device=local, rank=rank, world_size=world)
            r2.set_option("strip_rows", args.strip_rows)
            ident = [binding.comm_unique_id() if rank == 0 else None]
            dist.broadcast_object_list(ident, src=0)
            r2.comm_init(ident[0])
            r2.update(m)
            if rank == 0:
                nccl_img = r2.gather()
            barrier()
            r2.close()
        if rank == 0:
            with binding.Renderer(sc, device=local) as r1:
                r1.set_option("kernel", 0)
                r1.update(m)
                exact = r1.gather()
                if world > 1:
                    r1.set_option("kernel", args.kernel)
                    r1.set_option("accel", args.accel)
                    r1.update(m)
                    single = r1.gather()
                    if not np.array_equal(img, single):
                        out["assembled_equals_1gpu"] = False
                        out["mismatches"].append(label + ": assembled frame differs from the 1-GPU frame")
                    if nccl_img is not None:
                        out["nccl_allgather_equals_1gpu"] = bool(np.array_equal(nccl_img, single))
                        if not out["nccl_allgather_equals_1gpu"]:
                            out["mismatches"].append(label + ": NCCL all-gather frame differs from the 1-GPU frame")
            if not np.array_equal(img, exact):
                out["vs_exact_kernel"] = "differs"
                out["mismatches"].append("{}: {} pixels differ from kernel 0".format(label, int((img != exact).any(axis=-1).sum())))
            rng = np.random.default_rng(1000 + i)
            xy = np.stack([rng.integers(0, sc.px_width, n_px), rng.integers(0, sc.px_height, n_px)], axis=1).astype(np.int32)
            o = oracle_py.RefCpu(sc) if oracle_py.have_ref() else oracle_py.PortOracle(sc)
            want = oracle_py.quantize_rgba8(o.render_pixels(m, xy))
            d = np.abs(img[xy[:, 1], xy[:, 0]].astype(np.int16) - want.astype(np.int16)).max(axis=-1)
            out["vs_ref_pixels"] += int(n_px)
            out["outliers"] += int((d > 1).sum())
            out["max_diff"] = max(out["max_diff"], int(d.max()))
            if (d > 1).sum() > max(1, n_px // 1000):
                out["mismatches"].append("{}: {} of {} sampled pixels beyond +-1 LSB of the CPU reference".format(label, int((d > 1).sum()), n_px))
    bad = [len(out["mismatches"])]
    if world > 1:
        dist.broadcast_object_list(bad, src=0)
    if bad[0]:
        if rank == 0:
            print(json.dumps({"frame_check": out}), file=sys.stderr, flush=True)
        raise SystemExit(3)
    del out["mismatches"]
    return out


def pick_row_step(oracle_py, scenes, m, threads: int, budget_s: float) -> int:
    """Row stride that bounds one CPU pass over the workload to about `budget_s` seconds on `threads` host threads:
    a 96-pixel probe per scene estimates the cost of a whole frame (C5 at 7680x4320 would take hours)."""
    import numpy as np
    est = 0.0
    for _, sc in scenes:
        rng = np.random.default_rng(5)
        xy = np.stack([rng.integers(0, sc.px_width, 96), rng.integers(0, sc.px_height, 96)], axis=1).astype(np.int32)
        t0 = time.perf_counter()
        oracle_py.PortOracle(sc).render_pixels(m, xy)
        est += (time.perf_counter() - t0) / 96 * sc.px_width * sc.px_height
    return max(1, int(np.ceil(est / threads / budget_s)))


def cpu_baseline(pkg, scenes, args) -> dict:
    """The reference CPU path on this box's host cores, bounded row sample (about 10-30 s of CPU work): all host
    threads (rows split over threads; the reference itself is a serial loop, update-cpu.cpp:125-133) and, beside it,
    one thread "as shipped" on a sparser sample."""
    from cuda_ray_tracer_b200 import camera
    from oracle import oracle_py
    oracle_py.build()
    cores = os.cpu_count() or 1
    kind = "reference" if oracle_py.have_ref() else "port"
    m = camera.TEST_POSES["initial"]

    def leg(ystep, threads):
        rays, secs = 0, 0.0
        for _, sc in scenes:
            _, _, cnt = oracle_py.PortOracle(sc).render_rows(m, 0, sc.px_height, ystep, threads=cores, count=True)
            rays += cnt["primary_rays"] + cnt["reflection_rays"] + cnt["shadow_rays"]
            if kind == "reference":
                _, ms = oracle_py.RefCpu(sc).render_rows(m, 0, sc.px_height, ystep, threads=threads)
            else:
                _, ms = oracle_py.PortOracle(sc).render_rows(m, 0, sc.px_height, ystep, threads=threads)
            secs += ms / 1e3
        what = "every image row" if ystep == 1 else "every {}th image row".format(ystep)
        return rays / secs / 1e6, "{} of each frame of the workload, {} host thread{}, {:.1f} s".format(
            what, threads, "" if threads == 1 else "s", secs)

    ystep = args.ref_row_step or pick_row_step(oracle_py, scenes, m, cores, 10.0)
    value, sample = leg(ystep, cores)
    ystep1 = pick_row_step(oracle_py, scenes, m, 1, 6.0)
    value1, sample1 = leg(ystep1, 1)
    return {"value": value, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample,
            "one_thread": {"value": value1, "unit": UNIT, "cores": 1, "sample": sample1,
                           "what": "the reference as shipped: a serial loop over the pixels (update-cpu.cpp:125-133)"}}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="bundled4k", choices=["bundled4k", "synthetic"])
    ap.add_argument("--kernel", type=int, default=1)
    ap.add_argument("--strip-rows", type=int, default=8)
    ap.add_argument("--ref-row-step", type=int, default=0, help="CPU legs render every n-th image row (0 = 1 for the "
                    "cpu_baseline leg, 2 for --impl reference steps)")
    ap.add_argument("--gather", default="p2p", choices=["p2p", "nccl"], help="multi-GPU frame assembly")
    ap.add_argument("--e2e", default="shared", choices=["shared", "gather"], help="multi-GPU end-to-end leg: every rank "
                    "copies its strips into one shared host frame, or rank 0 gathers on the device and copies alone")
    ap.add_argument("--accel", type=int, default=0, choices=[-1, 0, 1], help="sphere-class grid (separate mode, SURVEY §8 f4): "
                    "0 = brute force like the reference (default, what the roofline is quoted on), 1 = on, -1 = library default")
    ap.add_argument("--dynamic", type=int, default=0, choices=[0, 1], help="multi-GPU, fused gather: 1 = row groups of the image are "
                    "handed out to the GPUs on demand by a counter in rank 0's memory (SURVEY §8 f3), 0 = static 8-row strips (default: "
                    "measured faster at 2, 4 and 8 GPUs, profiles/r2_scaling_8gpu.json)")
    ap.add_argument("--variant", type=int, default=0, help="wavefront kernel shape (0 = automatic)")
    ap.add_argument("--no-cull", dest="cull", action="store_false", help="trace every shadow ray the reference casts")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-frame-check", action="store_true", help="skip the untimed frame verification before the timed loops")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
