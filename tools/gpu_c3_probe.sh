#!/bin/bash
# One gpurun call: the two flavours of the forward kernels (CUDASW_CHAIN=1: quiet chains, =0: without) on C2, C3 and a
# shape sweep; for C3 also with the variant launches serialised (CUDASW_NO_SPREAD=1) and as an ncu launch list
# (per-kernel times in isolation). Default (no CUDASW_CHAIN): chains unless more than 4 variants run side by side.
O=gpurun_out; mkdir -p $O
for ch in "" 1 0; do
  echo "CUDASW_CHAIN=[$ch]"
  for env in "X=1" "CUDASW_NO_SPREAD=1"; do
    echo " env [$env]"
    env ${ch:+CUDASW_CHAIN=$ch} $env timeout 400 python tools/bench_configs.py 2:400000 3:1000000 2>&1 | grep '^{' | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print('  ',{k:d[k] for k in ('cfg','n','ms_total','ms_forward','ms_reverse','ms_traceback','GCUPS_fwd')})
"
  done
done 2>&1 | tee $O/c3_probe.txt
if [ "$1" = "ncu" ]; then
for ch in 1 0; do
  CUDASW_CHAIN=$ch timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:sw_score_kernel --csv --log-file $O/c3_launches_chain$ch.csv python tools/bench_configs.py 3:300000 > /dev/null 2>&1
done
fi
