#!/bin/bash
cd "${GRAFT_REPO_ROOT:-/root/repo}"
O=gpurun_out
mkdir -p $O
timeout 300 python -m pytest tests -x -q -m gpu > $O/s22_pytest_gpu.log 2>&1; echo "pytest gpu exit $?" | tee -a $O/s22_status.txt
timeout 300 python tools/gpu_e2e_bands.py clebsch,20spheres,quadratic,reflection_test > $O/s22_bands.jsonl 2> $O/s22_bands.err; echo "bands exit $?" | tee -a $O/s22_status.txt
timeout 300 python bench.py --no-cpu-baseline > $O/s22_bench.json 2> $O/s22_bench.err; echo "bench exit $?" | tee -a $O/s22_status.txt
tail -3 $O/s22_pytest_gpu.log; cat $O/s22_bands.jsonl
