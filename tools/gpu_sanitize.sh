#!/bin/bash
# compute-sanitizer (memcheck, racecheck, synccheck, initcheck) on small frames of the wavefront kernel -> gpurun_out/<tag>_sanitize_*.log
#   tools/gpu_sanitize.sh TAG
cd "${GRAFT_REPO_ROOT:-/root/repo}"
O=gpurun_out; mkdir -p $O; TAG=${1:-r2}
run() {  # tool, label, args...
  local tool=$1 label=$2; shift 2
  timeout 600 compute-sanitizer --tool $tool --print-limit 20 python tools/prof_scene.py "$@" > $O/${TAG}_sanitize_${tool}_${label}.log 2>&1
  echo "$tool $label exit $? : $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY' $O/${TAG}_sanitize_${tool}_${label}.log | tail -1)"
}
for tool in memcheck racecheck synccheck; do
  run $tool dingdong dingdong --width 160 --height 90 --frames 2
  run $tool reflection_test reflection_test --width 160 --height 90 --frames 2
  run $tool 20spheres 20spheres --width 160 --height 90 --frames 2
  run $tool synthetic128_brute synthetic --spheres 128 --cubics 2 --lights 2 --width 160 --height 90 --frames 2 --accel 0
  run $tool synthetic128_grid synthetic --spheres 128 --cubics 2 --lights 2 --width 160 --height 90 --frames 2 --accel 1
done
