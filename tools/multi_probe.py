"""GPU experiment: where a cudasw_multi_align_batch call spends its wall time (per-GPU device time vs call wall time)."""
import json
import sys
import time

import numpy as np

sys.path.insert(0, ".")
import gridss_b200 as g
from gridss_b200.aligner import PinnedBuffer
from tools import synth as S

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1000000
ma = g.MultiGpuAligner()
lens = S.lengths(4, n)
ps, pr = PinnedBuffer(int(lens[0][-1])), PinnedBuffer(int(lens[1][-1]))
sb, so, rb, ro = S.generate(4, n, out=(ps.array, pr.array))
pres, pcig = PinnedBuffer(n * g.RESULT_DTYPE.itemsize), PinnedBuffer(4 * 24 * n)
res, cig = pres.array.view(g.RESULT_DTYPE), pcig.array.view(np.uint32)
for rep in range(4):
    t = time.perf_counter()
    ma.align_blobs(sb, so, rb, ro, res=res, cig=cig)
    dt = (time.perf_counter() - t) * 1e3
st = ma.slice_stats()
print(json.dumps({"n": n, "gpus": ma.n_devices, "call_wall_ms": round(dt, 2), "per_gpu_ms_total": [round(s["ms_total"], 2) for s in st],
                  "per_gpu_pairs": [s["pairs"] for s in st], "per_gpu_phases": [{k: round(s[k], 2) for k in ("ms_h2d", "ms_pack", "ms_forward", "ms_reverse", "ms_traceback", "ms_d2h")} for s in st][:2]}))
