#!/bin/bash
# One gpurun call: GPU parity tests, bench (both arms), per-config timings, ncu launch list and full captures of the
# forward score kernel and the one-thread traceback kernel. Outputs land in gpurun_out/ (tag = $1).
TAG=${1:-r01}
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/smi_$TAG.txt 2>&1
( timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_$TAG.log 2>&1; echo "pytest exit $?" >> $O/pytest_$TAG.log )
tail -3 $O/pytest_$TAG.log
timeout 600 python bench.py > $O/bench_$TAG.json 2> $O/bench_$TAG.err; echo "bench exit $?"
tail -c 600 $O/bench_$TAG.err
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_ref_$TAG.json 2>> $O/bench_$TAG.err; echo "ref exit $?"
timeout 600 python tools/bench_configs.py 1:10000 1:400000 2:1000000 3:1000000 4:1000000 5:1000 > $O/configs_$TAG.txt 2>&1; echo "configs exit $?"
CUDASW_CHAIN=1 timeout 200 python tests/fuzz_gpu.py --seconds 60 > $O/fuzz_$TAG.json 2> $O/fuzz_$TAG.err; echo "fuzz exit $?"
for p in int_peak; do
  [ -x tools/$p ] && timeout 120 tools/$p > $O/${p}_$TAG.txt 2>&1
done
if [ "$2" != "noncu" ]; then
CMD="python bench.py --steps 2 --warmup 3 --pairs 50000 --no-cpu-baseline --no-windows --no-homology"
timeout 300 $CMD > $O/plain_$TAG.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file $O/launches_$TAG.csv $CMD > $O/ncu_launches_$TAG.log 2>&1
echo "launch list exit $?"
timeout 300 $CMD > $O/plain2_$TAG.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:sw_score_kernel -s 2 -c 1 -f -o $O/prof_fwd_$TAG $CMD > $O/ncu_full_$TAG.log 2>&1
echo "full capture (forward) exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:sw_traceback_kernel -s 5 -c 1 -f -o $O/prof_tb_$TAG $CMD > $O/ncu_full_tb_$TAG.log 2>&1
echo "full capture (traceback) exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:encode_kernel -s 9 -c 1 -f -o $O/prof_enc_$TAG $CMD > $O/ncu_full_enc_$TAG.log 2>&1
echo "full capture (encode) exit $?"
HCMD="python bench.py --steps 1 --warmup 3 --pairs 50000 --no-cpu-baseline --no-windows --homology-breakpoints 100000"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:window_gather_kernel -s 4 -c 1 -f -o $O/prof_gather_$TAG $HCMD > $O/ncu_full_gather_$TAG.log 2>&1
echo "full capture (query/window gather of the homology path) exit $?"
fi
ls -la $O | tail -20
