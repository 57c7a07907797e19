#!/bin/bash
# Final single-GPU evidence of the round: tests, default bench (with the CPU baseline), synthetic benches, launch list,
# ncu --set full captures of the top kernel on three scenes.
cd "${GRAFT_REPO_ROOT:-/root/repo}"
O=gpurun_out
mkdir -p $O
timeout 300 python -m pytest tests -x -q -m gpu > $O/final_pytest_gpu.log 2>&1; echo "pytest gpu exit $?" | tee -a $O/final_status.txt
timeout 100 python -c "import __graft_entry__ as g; g.smoke()" > $O/final_smoke.log 2>&1; echo "smoke exit $?" | tee -a $O/final_status.txt
timeout 400 python bench.py > $O/final_bench.json 2> $O/final_bench.err; echo "bench exit $?" | tee -a $O/final_status.txt
timeout 200 python bench.py --workload synthetic --steps 3 --warmup 3 --no-cpu-baseline --accel 0 > $O/final_bench_synth_brute.json 2> $O/final_bench_synth_brute.err; echo "synth brute exit $?" | tee -a $O/final_status.txt
timeout 200 python bench.py --workload synthetic --steps 10 --warmup 3 --no-cpu-baseline --accel 1 > $O/final_bench_synth_grid.json 2> $O/final_bench_synth_grid.err; echo "synth grid exit $?" | tee -a $O/final_status.txt
timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/final_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $O/final_ncu_launches.log 2>&1; echo "launch list exit $?" | tee -a $O/final_status.txt
for sc in 20spheres clebsch; do
timeout 120 ncu --set full --clock-control none --import-source on -k regex:render_wave -s 2 -c 1 -o $O/final_$sc -f python tools/prof_scene.py $sc > $O/final_ncu_$sc.log 2>&1; echo "ncu $sc exit $?" | tee -a $O/final_status.txt
done
timeout 200 ncu --set full --clock-control none --import-source on -k regex:render_wave -s 2 -c 1 -o $O/final_synth_grid -f python tools/prof_scene.py synthetic --width 3840 --height 2160 --accel 1 > $O/final_ncu_synth.log 2>&1; echo "ncu synth exit $?" | tee -a $O/final_status.txt
tail -3 $O/final_pytest_gpu.log; tail -2 $O/final_smoke.log
