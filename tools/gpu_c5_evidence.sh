cd "${GRAFT_REPO_ROOT:-/root/repo}"; O=gpurun_out/c5; mkdir -p $O
python -m pytest tests -m gpu -x -q 2>&1 | tail -1
timeout 400 python tools/prof_scene.py synthetic --width 3840 --height 2160 --accel 0 > $O/plain.log 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:render_wave -s 2 -c 1 -o $O/ncu_synthetic_a0 -f python tools/prof_scene.py synthetic --width 3840 --height 2160 --accel 0 > $O/ncu.log 2>&1; echo "ncu exit $?"
python tools/ncu_summary.py $O/ncu_synthetic_a0.ncu-rep profiles/r2_final_wave_synthetic_a0.txt > /dev/null 2>&1 && cp profiles/r2_final_wave_synthetic_a0.txt $O/
rm -f $O/ncu_synthetic_a0.ncu-rep
timeout 400 python bench.py --workload synthetic --steps 3 --warmup 3 --accel 0 > $O/bench_synth_brute.json 2> $O/bench_synth_brute.err; echo "synth brute exit $?"
timeout 400 python bench.py --workload synthetic --steps 10 --warmup 3 --accel 1 --no-cpu-baseline > $O/bench_synth_grid.json 2> $O/bench_synth_grid.err; echo "synth grid exit $?"
python - <<'PY'
import json
for n in ("bench_synth_brute", "bench_synth_grid"):
    d = json.loads([l for l in open("gpurun_out/c5/%s.json" % n) if l.startswith("{")][-1]); r = d["roofline"]
    print(n, "ms", round(d["ms_per_step"], 2), "Mrays/s", round(d["value"]), "frac", r["frac"], "fma%", r["ncu_fma_pipe_pct"], "issue%", r["ncu_issue_active_pct"], "check", d["config"]["frame_check"]["outliers"], d["config"]["frame_check"]["vs_exact_kernel"])
PY
