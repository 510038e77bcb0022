#!/bin/bash
# One gpurun call: C2/C3 forward times per build flavour of the quiet chains, then parity tests + fuzz on the default build.
O=gpurun_out; mkdir -p $O
for flags in "" "-DCUSW_NO_CHAIN" "-DCUSW_CHAIN_MAXG=16" "-DCUSW_CHAIN_GATE=0"; do
  CUDASW_NVCC_EXTRA="$flags" python gridss_b200/build.py --force > /dev/null 2>&1
  echo "flags [$flags]"
  timeout 400 python tools/bench_configs.py 2:400000 3:1000000 5:1000 2>&1 | grep '^{' | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print({k:d[k] for k in ('cfg','n','ms_total','ms_forward','ms_reverse','ms_traceback','GCUPS_fwd')})
"
done | tee $O/chain_c3.txt
python gridss_b200/build.py --force > /dev/null 2>&1
( timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_chain2.log 2>&1; echo "pytest exit $?" >> $O/pytest_chain2.log ); tail -3 $O/pytest_chain2.log
timeout 120 python tests/fuzz_gpu.py --seconds 45 > $O/fuzz_chain2.json 2> $O/fuzz_chain2.err; echo "fuzz exit $?"; head -c 300 $O/fuzz_chain2.json
