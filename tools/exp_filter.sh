#!/bin/bash
# experiment: rebuild libcudasw.so on the GPU box with extra nvcc flags, then time the phases of a device-resident C2 / C1 / C3
# batch with the k-mer filter (one chunk, so the phase times do not overlap)
for flags in "$@"; do
  CUDASW_NVCC_EXTRA="$flags" python gridss_b200/build.py --force > /dev/null 2>&1
  echo "flags [$flags]"
  python tools/filter_phases.py 2>&1 | grep -v "cudasw filter"
done
python gridss_b200/build.py --force > /dev/null 2>&1
