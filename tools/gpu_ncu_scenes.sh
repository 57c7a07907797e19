#!/bin/bash
# ncu --set full capture of the render kernel (third frame) for the given scenes -> gpurun_out/<tag>_<scene>.ncu-rep
#   tools/gpu_ncu_scenes.sh TAG scene [scene ...]       (scene "synthetic" = C5 at 3840x2160, brute force)
cd "${GRAFT_REPO_ROOT:-/root/repo}"
O=gpurun_out
mkdir -p $O
TAG=$1; shift
for sc in "$@"; do
  extra=""
  if [ "$sc" = "synthetic" ]; then extra="--accel 0"; fi
  if [ "$sc" = "synthetic_grid" ]; then sc=synthetic; extra="--accel 1"; fi
  timeout 60 python tools/prof_scene.py $sc $extra > $O/${TAG}_$sc.plain.log 2>&1 || { echo "plain run of $sc failed"; continue; }
  timeout 300 ncu --set full --clock-control none --import-source on -k regex:render_wave -s 2 -c 1 -o $O/${TAG}_$sc$( [ -n "$extra" ] && echo "_a${extra: -1}" ) -f \
      python tools/prof_scene.py $sc $extra > $O/${TAG}_$sc.ncu.log 2>&1
  echo "ncu $sc exit $?"
done
