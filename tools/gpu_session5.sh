#!/bin/bash
# Evidence session for the k-mer-filtered pipeline (round 2, second half): plain bench of the profiled command, ncu launch
# list of the same command, full captures of the forward kernel's sub-window pass, the filter kernel and the banded reverse
# pass. Summaries: python tools/profile_summaries.py <tag> r02f
TAG=${1:-r02t}
O=gpurun_out
mkdir -p $O
CMD="python bench.py --steps 2 --warmup 3 --pairs 50000 --no-cpu-baseline --no-windows --no-homology --no-worst-case --no-int-peak"
timeout 300 $CMD > $O/plain_$TAG.log 2> $O/plain_$TAG.err; echo "plain exit $?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file $O/launches_$TAG.csv $CMD > $O/ncu_list_$TAG.log 2>&1; echo "list exit $?"
NCU="timeout 900 ncu --set full --clock-control none --import-source on --kernel-name-base mangled -c 1 -f"
# forward, chained query-profile flavour: two launches per step (sub-window pass, then the full-window pass of the unproven pairs)
$NCU -k regex:sw_score_kernelILi8ELi19ELb0ELb0ELb1ELb1E -s 6 -o $O/prof_fwd_$TAG $CMD > $O/ncu_full_fwd_$TAG.log 2>&1; echo "fwd exit $?"
$NCU -k regex:filter_kernel -s 3 -o $O/prof_filter_$TAG $CMD > $O/ncu_full_filter_$TAG.log 2>&1; echo "filter exit $?"
$NCU -k regex:sw_rev_band_kernelILi1ELi16ELb1E -s 3 -o $O/prof_rev_$TAG $CMD > $O/ncu_full_rev_$TAG.log 2>&1; echo "rev exit $?"
ls -la $O/ | grep "prof_.*$TAG"
