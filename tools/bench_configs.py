"""GPU experiment: phase timings and GCUPS of every BASELINE.json config shape (not the bench of record)."""
import json
import sys
import time

import numpy as np

sys.path.insert(0, ".")
import gridss_b200 as g
from tools import synth as S
from gridss_b200.aligner import PinnedBuffer

al = g.AlignerFactory.create()
runs = [(1, 10000), (1, 400000), (2, 200000), (3, 100000), (4, 400000), (5, 64), (5, 1000)]
if len(sys.argv) > 1:
    runs = [tuple(int(x) for x in a.split(":")) for a in sys.argv[1:]]
for cfg, n in runs:
    lens = S.lengths(cfg, n)
    pin_s, pin_r = PinnedBuffer(int(lens[0][-1])), PinnedBuffer(int(lens[1][-1]))
    sb, so, rb, ro = S.generate(cfg, n, out=(pin_s.array, pin_r.array))
    for rep in range(2):
        t = time.time()
        res, cig = al.align_blobs(sb, so, rb, ro)
        dt = time.time() - t
    st = al.stats()
    cells = float((np.diff(so) * np.diff(ro)).sum())
    print(json.dumps({"cfg": cfg, "n": n, "wall_ms": round(dt * 1e3, 2), "pairs_per_s": round(n / dt), "GCUPS_wall": round(cells / dt / 1e9, 1),
                      "GCUPS_fwd": round(cells / max(st["ms_forward"], 1e-3) / 1e6, 1), "bad_status": int((res["status"] != 0).sum()),
                      **{k: round(v, 2) for k, v in st.items() if k.startswith("ms_")}}), flush=True)
