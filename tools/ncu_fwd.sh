#!/bin/bash
# ncu --set full capture of the forward score kernel (one launch; with the k-mer filter every step launches the chained flavour twice: -s 6 = the sub-window pass of the fourth step), tag = $1 (tools/gpu_session5.sh is the full session)
TAG=${1:-x}
CMD="python bench.py --steps 2 --warmup 3 --pairs 50000 --no-cpu-baseline"
timeout 300 $CMD > gpurun_out/plain_$TAG.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:sw_score_kernel -s 6 -c 1 -f -o gpurun_out/prof_fwd_$TAG $CMD > gpurun_out/ncu_full_$TAG.log 2>&1
echo "exit $?"
