#!/bin/bash
# experiment: rebuild libcudasw.so on the GPU box with extra nvcc flags, then time the forward pass (results may be wrong under EXP flags)
for flags in "$@"; do
  CUDASW_NVCC_EXTRA="$flags" python gridss_b200/build.py --force > /dev/null 2>&1
  echo "flags [$flags]"
  python - <<'P' 2>&1 | tail -3
import sys; sys.path.insert(0, ".")
import numpy as np, torch
import gridss_b200 as g
from tools import synth as S
n = 400000
al = g.CudaAligner()
sb, so, rb, ro = S.generate(2, n)
d = [torch.from_numpy(x).cuda() for x in (sb, so, rb, ro)]
torch.cuda.synchronize()
for rep in range(3):
    try:
        al.align_blobs_device(d[0].data_ptr(), d[1].data_ptr(), d[2].data_ptr(), d[3].data_ptr(), so, ro)
    except Exception as e:
        print("  (call failed:", str(e)[:80], ")")
    st = al.stats()
print("  fwd ms", round(st["ms_forward"], 2), "TCUPS", round(n * 300 * 4000 / st["ms_forward"] / 1e9, 3), "rev", round(st["ms_reverse"], 2), "tb", round(st["ms_traceback"], 2))
P
done
python gridss_b200/build.py --force > /dev/null 2>&1
