#!/bin/bash
cd "${GRAFT_REPO_ROOT:-/root/repo}"
O=gpurun_out
mkdir -p $O
timeout 300 python -m pytest tests -x -q -m gpu > $O/s17_pytest_gpu.log 2>&1; echo "pytest gpu exit $?" | tee -a $O/s17_status.txt
timeout 150 python tools/gpu_fuzz_scenes.py 1200 35 8 > $O/s17_fuzz8.log 2>&1; echo "fuzz8 exit $?" | tee -a $O/s17_status.txt
timeout 300 python bench.py --no-cpu-baseline > $O/s17_bench.json 2> $O/s17_bench.err; echo "bench exit $?" | tee -a $O/s17_status.txt
timeout 200 python tools/gpu_variants.py 2,3,5 > $O/s17_variants.jsonl 2> $O/s17_variants.err; echo "variants exit $?" | tee -a $O/s17_status.txt
timeout 240 python tools/gpu_grid_check.py 3840 2160 > $O/s17_grid_check.jsonl 2> $O/s17_grid_check.err; echo "grid check exit $?" | tee -a $O/s17_status.txt
timeout 120 ncu --set full --clock-control none --import-source on -k regex:render_wave -s 2 -c 1 -o $O/s17_dingdong -f python tools/prof_scene.py dingdong > $O/s17_ncu_dingdong.log 2>&1; echo "ncu exit $?" | tee -a $O/s17_status.txt
tail -3 $O/s17_pytest_gpu.log; tail -n 2 $O/s17_fuzz8.log; cat $O/s17_variants.jsonl | cut -c1-200
