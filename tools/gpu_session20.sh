#!/bin/bash
cd "${GRAFT_REPO_ROOT:-/root/repo}"
O=gpurun_out
mkdir -p $O
timeout 300 python -m pytest tests -x -q -m gpu > $O/s20_pytest_gpu.log 2>&1; echo "pytest gpu exit $?" | tee -a $O/s20_status.txt
timeout 150 python tools/gpu_fuzz_scenes.py 1200 45 8 > $O/s20_fuzz8.log 2>&1; echo "fuzz8 exit $?" | tee -a $O/s20_status.txt
timeout 300 python bench.py --no-cpu-baseline > $O/s20_bench.json 2> $O/s20_bench.err; echo "bench exit $?" | tee -a $O/s20_status.txt
timeout 240 python tools/gpu_grid_check.py 3840 2160 > $O/s20_grid_check.jsonl 2> $O/s20_grid_check.err; echo "grid check exit $?" | tee -a $O/s20_status.txt
tail -3 $O/s20_pytest_gpu.log; tail -n 2 $O/s20_fuzz8.log; head -3 $O/s20_grid_check.jsonl | cut -c1-200
