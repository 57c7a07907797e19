#!/bin/bash
# 2-GPU sanity of the final code: byte identity of every assembly path, one bench line.
cd "${GRAFT_REPO_ROOT:-/root/repo}"
O=gpurun_out
mkdir -p $O
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/multi_gpu_check.py > $O/g2_check.log 2>&1; echo "check exit $?" | tee -a $O/g2_status.txt
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 20 --warmup 3 > $O/g2_bench2.json 2> $O/g2_bench2.err; echo "bench2 exit $?" | tee -a $O/g2_status.txt
grep -c "identical=True" $O/g2_check.log; grep -c "identical=False" $O/g2_check.log; tail -3 $O/g2_check.log
