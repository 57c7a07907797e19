#!/bin/bash
# The GPU test-suite and one bench step against the CHECKED build of the library (index assertions in the kernels).
#   python cuda-ray-tracer_b200/build.py --checked   (here, before gpurun)
cd "${GRAFT_REPO_ROOT:-/root/repo}"
O=gpurun_out; mkdir -p $O; TAG=${1:-r2}
export RT_B200_LIB=$PWD/cuda-ray-tracer_b200/librt_b200_checked.so
{ echo "library: $RT_B200_LIB"; timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
  timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu-baseline 2>&1 | grep -c '"frame_check"' | sed 's/^/bench lines with frame_check: /'
  echo "(a failed RT_CHECK traps: the launch fails and the test with it)"; } 2>&1 | tee $O/${TAG}_checked.log
