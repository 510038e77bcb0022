#!/bin/bash
# One gpurun call: forward-kernel shape sweep per build flavour (quiet chains on / off / third-step variant).
O=gpurun_out; mkdir -p $O
SH="50x1000 150x2000 300x1000 300x4000 600x4000 1000x4000 1500x4000 2000x4000 2000x10000 1000x1000"
for flags in "" "-DCUSW_NO_CHAIN" "-DCUSW_CHAIN_STEP3"; do
  CUDASW_NVCC_EXTRA="$flags" python gridss_b200/build.py --force > /dev/null 2>&1
  echo "flags [$flags]"
  timeout 400 python tools/shape_sweep.py $SH 2>&1 | grep '^{' | cut -c1-200
done | tee $O/chain_sweep.txt
python gridss_b200/build.py --force > /dev/null 2>&1
