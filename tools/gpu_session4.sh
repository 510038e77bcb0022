#!/bin/bash
# the three captures gpu_session3.sh got wrong: banded reverse (the G = 1 variant that takes C2), strip-mined forward, C5 reverse
TAG=${1:-r02}
O=gpurun_out
mkdir -p $O
CMD="python bench.py --steps 2 --warmup 3 --pairs 50000 --no-cpu-baseline --no-windows --no-homology --no-worst-case --no-int-peak"
C5="python tools/bench_configs.py 5:1000"
NCU="timeout 900 ncu --set full --clock-control none --import-source on --kernel-name-base mangled -c 1 -f"
$NCU -k regex:sw_rev_band_kernelILi1ELi16ELb1E -s 2 -o $O/prof_rev_$TAG $CMD > $O/ncu_full_rev_$TAG.log 2>&1; echo "rev exit $?"
$NCU -k regex:sw_score_kernelILi32ELi20ELb0ELb1E -s 1 -o $O/prof_fwdstrip_$TAG $C5 > $O/ncu_full_fwdstrip_$TAG.log 2>&1; echo "fwdstrip exit $?"
$NCU -k regex:sw_rev_band_kernelILi32E -s 1 -o $O/prof_revc5_$TAG $C5 > $O/ncu_full_revc5_$TAG.log 2>&1; echo "revc5 exit $?"
ls -la $O/ | grep "prof_.*$TAG"
