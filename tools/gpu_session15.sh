#!/bin/bash
# Nearest-hit pruning of cubic tests by the running bound: parity suite, scene fuzzing with many cubics, timings.
cd "${GRAFT_REPO_ROOT:-/root/repo}"
O=gpurun_out
mkdir -p $O
timeout 300 python -m pytest tests -x -q -m gpu > $O/s15_pytest_gpu.log 2>&1; echo "pytest gpu exit $?" | tee -a $O/s15_status.txt
timeout 150 python tools/gpu_fuzz_scenes.py 1200 15 8 > $O/s15_fuzz8.log 2>&1; echo "fuzz8 exit $?" | tee -a $O/s15_status.txt
timeout 100 python tools/gpu_fuzz_scenes.py 800 16 3 > $O/s15_fuzz3.log 2>&1; echo "fuzz3 exit $?" | tee -a $O/s15_status.txt
timeout 240 python tools/gpu_grid_check.py 3840 2160 > $O/s15_grid_check.jsonl 2> $O/s15_grid_check.err; echo "grid check exit $?" | tee -a $O/s15_status.txt
timeout 300 python bench.py --no-cpu-baseline > $O/s15_bench.json 2> $O/s15_bench.err; echo "bench exit $?" | tee -a $O/s15_status.txt
timeout 200 python bench.py --workload synthetic --steps 3 --warmup 3 --no-cpu-baseline --accel 0 > $O/s15_bench_synth_brute.json 2> $O/s15_bench_synth_brute.err; echo "synth brute exit $?" | tee -a $O/s15_status.txt
timeout 200 python bench.py --workload synthetic --steps 10 --warmup 3 --no-cpu-baseline --accel 1 > $O/s15_bench_synth_grid.json 2> $O/s15_bench_synth_grid.err; echo "synth grid exit $?" | tee -a $O/s15_status.txt
timeout 200 ncu --set full --clock-control none --import-source on -k regex:render_wave -s 2 -c 1 -o $O/s15_synth_grid -f python tools/prof_scene.py synthetic --width 3840 --height 2160 --accel 1 > $O/s15_ncu.log 2>&1; echo "ncu exit $?" | tee -a $O/s15_status.txt
tail -3 $O/s15_pytest_gpu.log; tail -2 $O/s15_fuzz8.log $O/s15_fuzz3.log; head -3 $O/s15_grid_check.jsonl | cut -c1-300
