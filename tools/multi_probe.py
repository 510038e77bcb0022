"""GPU experiment: where a cudasw_multi_align_batch call spends its wall time (per-GPU device time vs call wall time),
ASCII and packed host inputs, against the one-engine rate on the same pairs."""
import json
import sys
import time

import numpy as np

sys.path.insert(0, ".")
import gridss_b200 as g
from gridss_b200.aligner import PinnedBuffer
from tools import synth as S

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1000000
cfg = int(sys.argv[2]) if len(sys.argv) > 2 else 4
ma = g.MultiGpuAligner()
lens = S.lengths(cfg, n)
ps, pr = PinnedBuffer(int(lens[0][-1])), PinnedBuffer(int(lens[1][-1]))
sb, so, rb, ro = S.generate(cfg, n, out=(ps.array, pr.array))
pres, pcig = PinnedBuffer(n * g.RESULT_DTYPE.itemsize), PinnedBuffer(4 * 24 * n)
res, cig = pres.array.view(g.RESULT_DTYPE), pcig.array.view(np.uint32)
w2q, wnq = g.packed_words(int(so[-1]))
w2w, wnw = g.packed_words(int(ro[-1]))
pq2, pqn, pw2, pwn = PinnedBuffer(4 * w2q), PinnedBuffer(4 * wnq), PinnedBuffer(4 * w2w), PinnedBuffer(4 * wnw)
q2, qn, w2, wn = (b.array.view(np.uint32) for b in (pq2, pqn, pw2, pwn))
_, _, aq = g.pack_bases(sb, q2, qn, n=int(so[-1]), threads=16)
_, _, aw = g.pack_bases(rb, w2, wn, n=int(ro[-1]), threads=16)
out = {"n": n, "cfg": cfg, "gpus": ma.n_devices}
for name, call in (("ascii", lambda: ma.align_blobs(sb, so, rb, ro, res=res, cig=cig)),
                   ("packed", lambda: ma.align_blobs_packed(q2, qn if aq else None, so, w2, wn if aw else None, ro, res=res, cig=cig))):
    for rep in range(4):
        t = time.perf_counter()
        call()
        dt = (time.perf_counter() - t) * 1e3
    st = ma.slice_stats()
    out[name] = {"call_wall_ms": round(dt, 2), "pairs_per_s": round(n / dt * 1e3), "per_gpu_ms_total": [round(s["ms_total"], 2) for s in st],
                 "per_gpu_pairs": [s["pairs"] for s in st],
                 "gpu0_phases": {k: round(st[0][k], 2) for k in ("ms_h2d", "ms_pack", "ms_forward", "ms_reverse", "ms_traceback", "ms_d2h")}}
# the same pairs on ONE engine in slices of n / gpus: the per-GPU rate a rank-per-GPU deployment would see
one = g.CudaAligner(device=0)
k = n // ma.n_devices
so1 = np.ascontiguousarray(so[: k + 1])
ro1 = np.ascontiguousarray(ro[: k + 1])
for rep in range(3):
    t = time.perf_counter()
    one.align_blobs_packed(q2, qn if aq else None, so1, w2, wn if aw else None, ro1, res=res[:k], cig=cig)
    dt1 = (time.perf_counter() - t) * 1e3
out["one_engine_slice"] = {"pairs": k, "call_wall_ms": round(dt1, 2), "pairs_per_s": round(k / dt1 * 1e3)}
out["multi_vs_ranks"] = round(out["packed"]["pairs_per_s"] / (out["one_engine_slice"]["pairs_per_s"] * ma.n_devices), 3)
print(json.dumps(out))
