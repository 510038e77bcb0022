#!/bin/bash
# round 2 measurement session: bench of record (both arms), per-config timings, ncu launch list and --set full captures of
# the forward, banded reverse, cooperative traceback (warp and CTA flavours) and unpack kernels. Outputs -> gpurun_out/ (tag $1).
TAG=${1:-r02}
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/smi_$TAG.txt 2>&1
timeout 600 python bench.py > $O/bench_$TAG.json 2> $O/bench_$TAG.err; echo "bench exit $?"
tail -c 400 $O/bench_$TAG.err
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_ref_$TAG.json 2>> $O/bench_$TAG.err; echo "ref exit $?"
timeout 600 python tools/bench_configs.py 1:10000 1:400000 2:1000000 3:1000000 4:1000000 5:1000 > $O/configs_$TAG.txt 2>&1; echo "configs exit $?"
CMD="python bench.py --steps 2 --warmup 3 --pairs 50000 --no-cpu-baseline --no-windows --no-homology --no-worst-case --no-int-peak"
timeout 300 $CMD > $O/plain_$TAG.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file $O/launches_$TAG.csv $CMD > $O/ncu_launches_$TAG.log 2>&1
echo "launch list exit $?"
cap() {  # name, kernel regex, skip count, command
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:$2 -s $3 -c 1 -f -o $O/prof_$1_$TAG ${@:4} > $O/ncu_full_$1_$TAG.log 2>&1
  echo "full capture ($1) exit $?"
}
cap fwd sw_score_kernel 2 $CMD
cap rev sw_rev_band_kernel 3 $CMD
cap unpack unpack_kernel 3 $CMD
cap tb sw_traceback_kernel 12 $CMD
C3="python tools/bench_configs.py 3:200000"
C5="python tools/bench_configs.py 5:1000"
cap tbwarp 'sw_traceback_warp_kernel' 5 $C3
cap tbcta 'sw_traceback_block_kernel' 1 $C5
cap fwdstrip 'sw_score_kernel<32' 1 $C5
cap revc5 'sw_rev_band_kernel<32' 1 $C5
ls -la $O | tail -30
