#!/bin/bash
# Final single-GPU evidence of the round (run under gpurun, one GPU):
#   tests + smoke, ncu --set full of the render kernel on all 8 bundled scenes and on C5 (3840x2160, brute force and grid)
#   -> summaries profiles/r2_final_wave_*.txt (bench.py reads them), default bench with the CPU baseline, C5 benches,
#   ncu launch list of the default bench command.  Everything is also copied to gpurun_out/final/ for the way back.
cd "${GRAFT_REPO_ROOT:-/root/repo}"
O=gpurun_out/final; mkdir -p $O
S=$O/status.txt; : > $S
timeout 600 python -m pytest tests -x -q -m gpu > $O/pytest_gpu.log 2>&1; echo "pytest gpu exit $?: $(tail -1 $O/pytest_gpu.log)" | tee -a $S
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke exit $?" | tee -a $S
tools/gpu_ncu_scenes.sh final/ncu 20spheres quadratic cubic cayley clebsch dingdong monkey_saddle reflection_test 2>&1 | tee -a $S
timeout 400 python tools/prof_scene.py synthetic --width 3840 --height 2160 --accel 0 > $O/ncu_synthetic_a0.plain.log 2>&1 && \
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:render_wave -s 2 -c 1 -o $O/ncu_synthetic_a0 -f python tools/prof_scene.py synthetic --width 3840 --height 2160 --accel 0 > $O/ncu_synthetic_a0.log 2>&1; echo "ncu synthetic brute exit $?" | tee -a $S
timeout 400 python tools/prof_scene.py synthetic --width 3840 --height 2160 --accel 1 > $O/ncu_synthetic_a1.plain.log 2>&1 && \
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:render_wave -s 2 -c 1 -o $O/ncu_synthetic_a1 -f python tools/prof_scene.py synthetic --width 3840 --height 2160 --accel 1 > $O/ncu_synthetic_a1.log 2>&1; echo "ncu synthetic grid exit $?" | tee -a $S
for sc in 20spheres quadratic cubic cayley clebsch dingdong monkey_saddle reflection_test synthetic_a0 synthetic_a1; do
  python tools/ncu_summary.py $O/ncu_$sc.ncu-rep profiles/r2_final_wave_$sc.txt > /dev/null 2>&1 && cp profiles/r2_final_wave_$sc.txt $O/
done
ls profiles/r2_final_wave_*.txt | wc -l | sed 's/^/summaries written: /' | tee -a $S
# gpurun returns at most 64 MiB: keep three reports (per-line attribution here), drop the others
for sc in quadratic cubic cayley dingdong monkey_saddle reflection_test synthetic_a1; do rm -f $O/ncu_$sc.ncu-rep; done
rm -f gpurun_out/final/*.plain.log
timeout 600 python bench.py > $O/bench.json 2> $O/bench.err; echo "bench exit $?" | tee -a $S
timeout 400 python bench.py --workload synthetic --steps 3 --warmup 3 --accel 0 > $O/bench_synth_brute.json 2> $O/bench_synth_brute.err; echo "synth brute exit $?" | tee -a $S
timeout 400 python bench.py --workload synthetic --steps 10 --warmup 3 --accel 1 > $O/bench_synth_grid.json 2> $O/bench_synth_grid.err; echo "synth grid exit $?" | tee -a $S
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/bench_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-frame-check > $O/ncu_launches.log 2>&1; echo "launch list exit $?" | tee -a $S
python - <<'PY'
import json
for n in ("bench", "bench_synth_brute", "bench_synth_grid"):
    try:
        d = json.loads([l for l in open("gpurun_out/final/%s.json" % n) if l.startswith("{")][-1])
        r = d["roofline"]
        print(n, "ms/step", round(d["ms_per_step"], 4), "e2e", round(d["e2e"]["ms_per_step"], 4), "frac", r["frac"], "fma%", r["ncu_fma_pipe_pct"],
              "fp64%", r["ncu_fp64_pipe_pct"], "check", d["config"]["frame_check"], "cpu", (d.get("cpu_baseline") or {}).get("value"))
    except Exception as e:
        print(n, "FAILED", e)
PY
