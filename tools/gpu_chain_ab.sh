#!/bin/bash
# One gpurun call: step statistics of the current build, then an A/B of the forward pass over build flags.
O=gpurun_out; mkdir -p $O
bash tools/exp_stats.sh > $O/step_stats.txt 2>&1; cat $O/step_stats.txt
timeout 900 bash tools/exp_build_run.sh "$@" 2>&1 | tee $O/chain_ab.txt
