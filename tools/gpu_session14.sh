#!/bin/bash
# One GPU call: grid test first (new code), then measurements, then the whole GPU suite.  Every step writes its own
# file under gpurun_out/ so that a cut-off call still leaves what finished.
cd "${GRAFT_REPO_ROOT:-/root/repo}"
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm --format=csv > $O/s14_gpu.txt 2>&1
timeout 300 python -m pytest tests/test_parity_gpu.py -q -m gpu -k "sphere_grid" > $O/s14_pytest_grid.log 2>&1; echo "grid test exit $?" | tee -a $O/s14_status.txt
timeout 240 python tools/gpu_grid_check.py 3840 2160 > $O/s14_grid_check.jsonl 2> $O/s14_grid_check.err; echo "grid check exit $?" | tee -a $O/s14_status.txt
timeout 300 python bench.py > $O/s14_bench.json 2> $O/s14_bench.err; echo "bench exit $?" | tee -a $O/s14_status.txt
timeout 200 python bench.py --workload synthetic --steps 3 --warmup 3 --no-cpu-baseline --accel 0 > $O/s14_bench_synth_brute.json 2> $O/s14_bench_synth_brute.err; echo "synth brute exit $?" | tee -a $O/s14_status.txt
timeout 200 python bench.py --workload synthetic --steps 10 --warmup 3 --no-cpu-baseline --accel 1 > $O/s14_bench_synth_grid.json 2> $O/s14_bench_synth_grid.err; echo "synth grid exit $?" | tee -a $O/s14_status.txt
timeout 200 ncu --set full --clock-control none --import-source on -k regex:render_wave -s 2 -c 1 -o $O/s14_synth_grid -f python tools/prof_scene.py synthetic --width 3840 --height 2160 --accel 1 > $O/s14_ncu.log 2>&1; echo "ncu exit $?" | tee -a $O/s14_status.txt
timeout 600 python -m pytest tests -x -q -m gpu > $O/s14_pytest_gpu.log 2>&1; echo "pytest gpu exit $?" | tee -a $O/s14_status.txt
tail -3 $O/s14_pytest_grid.log; tail -3 $O/s14_pytest_gpu.log; cat $O/s14_grid_check.jsonl | cut -c1-300
