#!/bin/bash
# Multi-GPU numbers on one 8-GPU box (gpurun --gpus 8): bundled4k at N = 8 (frame check + RT_B200_TRACE), 4, 2 and the
# synthetic C5 workload at N = 8.  Output: gpurun_out/scale/*.json, trace lines, one summary line per run.
cd "${GRAFT_REPO_ROOT:-/root/repo}"
O=gpurun_out/scale; mkdir -p $O; PORT=29700
bench() {  # name, nproc, args...
  local name=$1 n=$2; shift 2
  PORT=$((PORT+1))
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $n "$@" 2> $O/$name.err | grep '"metric"' > $O/$name.json
  python - "$O/$name.json" "$name" <<'PY'
import json, sys
try:
    d = json.load(open(sys.argv[1]))
    print(sys.argv[2], "ms/step", round(d["ms_per_step"], 4), "wall", round(d["config"]["ms_per_step_wall"], 4), "e2e", round(d["e2e"]["ms_per_step"], 4),
          "frame_check", d["config"].get("frame_check"), "clocks", d.get("clocks"), flush=True)
except Exception as e:
    print(sys.argv[2], "FAILED", e, flush=True)
PY
}
RT_B200_TRACE=1 bench b8 8 --steps 20 --warmup 3
grep "rt_b200 trace" $O/b8.err | sort > $O/b8_trace.txt; rm -f $O/b8.err
bench b4 4 --steps 20 --warmup 3 --no-frame-check
bench b2 2 --steps 20 --warmup 3 --no-frame-check
bench synth8_brute 8 --workload synthetic --steps 3 --warmup 3 --accel 0 --no-frame-check
bench synth8_grid 8 --workload synthetic --steps 5 --warmup 3 --accel 1 --no-frame-check
