#!/bin/bash
cd "${GRAFT_REPO_ROOT:-/root/repo}"
O=gpurun_out
mkdir -p $O
timeout 300 python -m pytest tests -x -q -m gpu > $O/s16_pytest_gpu.log 2>&1; echo "pytest gpu exit $?" | tee -a $O/s16_status.txt
timeout 150 python tools/gpu_fuzz_scenes.py 1200 25 8 > $O/s16_fuzz8.log 2>&1; echo "fuzz8 exit $?" | tee -a $O/s16_status.txt
timeout 240 python tools/gpu_grid_check.py 3840 2160 > $O/s16_grid_check.jsonl 2> $O/s16_grid_check.err; echo "grid check exit $?" | tee -a $O/s16_status.txt
timeout 300 python bench.py --no-cpu-baseline > $O/s16_bench.json 2> $O/s16_bench.err; echo "bench exit $?" | tee -a $O/s16_status.txt
for sc in 20spheres clebsch dingdong; do
timeout 120 ncu --set full --clock-control none --import-source on -k regex:render_wave -s 2 -c 1 -o $O/s16_$sc -f python tools/prof_scene.py $sc > $O/s16_ncu_$sc.log 2>&1; echo "ncu $sc exit $?" | tee -a $O/s16_status.txt
done
tail -3 $O/s16_pytest_gpu.log; tail -n 2 $O/s16_fuzz8.log; head -3 $O/s16_grid_check.jsonl | cut -c1-300
