#!/bin/bash
# round 2 evidence session: bounds-checked GPU suite, volume fuzz vs the reference binary, C5 volume, bench of record
TAG=${1:-r02}
O=gpurun_out
mkdir -p $O
( CUDASW_LIB=$PWD/gridss_b200/libcudasw_check.so timeout 1500 python -m pytest tests -m gpu -x -q > $O/bounds_check_$TAG.log 2>&1; echo "pytest (libcudasw_check.so, -DCUSW_BOUNDS_CHECK) exit $?" >> $O/bounds_check_$TAG.log )
tail -4 $O/bounds_check_$TAG.log
timeout 700 python tests/fuzz_gpu.py --seconds ${FUZZ_SECONDS:-420} > $O/fuzz_$TAG.json 2> $O/fuzz_$TAG.err; echo "fuzz exit $?"; cut -c1-600 $O/fuzz_$TAG.json
timeout 900 python tests/fuzz_gpu.py --c5 ${C5_PAIRS:-2000} > $O/fuzz_c5_$TAG.json 2> $O/fuzz_c5_$TAG.err; echo "fuzz c5 exit $?"; cut -c1-400 $O/fuzz_c5_$TAG.json
