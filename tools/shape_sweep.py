"""GPU experiment: forward-kernel throughput as a function of the pair shape (uniform batches, query = mutated window slice)."""
import json
import sys

import numpy as np

sys.path.insert(0, ".")
import gridss_b200 as g

al = g.AlignerFactory.create()
rng = np.random.default_rng(5)
shapes = [(100, 2000), (300, 4000), (600, 4000), (1200, 4000), (2000, 4000), (2000, 10000), (1000, 1000)]
if len(sys.argv) > 1:
    shapes = [tuple(int(x) for x in a.split("x")) for a in sys.argv[1:]]
for q, w in shapes:
    n = max(2000, int(2.4e11 / (q * w)))
    win = rng.integers(0, 4, size=(n, w), dtype=np.uint8)
    st = rng.integers(0, w - q + 1, size=n)
    idx = st[:, None] + np.arange(q)[None, :]
    qs = np.take_along_axis(win, idx, axis=1)
    mut = rng.random((n, q)) < 0.02
    qs = np.where(mut, (qs + rng.integers(1, 4, size=(n, q), dtype=np.uint8)) & 3, qs).astype(np.uint8)
    lut = np.frombuffer(b"ACGT", dtype=np.uint8)
    sb, rb = lut[qs].reshape(-1), lut[win].reshape(-1)
    so, ro = np.arange(n + 1, dtype=np.int64) * q, np.arange(n + 1, dtype=np.int64) * w
    for _ in range(2):
        res, cig = al.align_blobs(sb, so, rb, ro)
    s = al.stats()
    print(json.dumps({"q": q, "w": w, "n": n, "variant": al.variant_for(q), "fwd_TCUPS": round(q * w * n / s["ms_forward"] / 1e9, 3),
                      "padded_TCUPS": round(s["forward_cells_padded"] / s["ms_forward"] / 1e9, 3),
                      **{k: round(s[k], 2) for k in ("ms_forward", "ms_reverse", "ms_traceback")}}), flush=True)
