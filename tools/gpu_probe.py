"""First-contact probe for a gpurun call: quick parity + phase timings at growing batch sizes."""
import json
import sys
import time

import numpy as np

sys.path.insert(0, ".")
import gridss_b200 as g
from oracle import oracle as O
from tools import synth as S

al = g.AlignerFactory.create()
FIELDS = ["score", "ref_begin", "ref_end", "read_begin", "read_end", "cigar_len", "status"]


def check(sb, so, rb, ro, label):
    m = O.model_align_blobs(sb, so, rb, ro, threads=32)
    res, cig = al.align_blobs(sb, so, rb, ro)
    bad = 0
    first = None
    for i in range(len(res)):
        a = tuple(int(res[i][f]) for f in FIELDS)
        b = tuple(int(m.res[i][f]) for f in FIELDS)
        ca = tuple(cig[res[i]["cigar_off"]: res[i]["cigar_off"] + res[i]["cigar_len"]])
        cb = tuple(m.cig[m.res[i]["cigar_off"]: m.res[i]["cigar_off"] + m.res[i]["cigar_len"]])
        if a != b or ca != cb:
            bad += 1
            if first is None:
                first = (i, a, b, g.cigar_string(ca), g.cigar_string(cb))
    print(label, "pairs", len(res), "bad", bad, first, flush=True)
    return bad


tot = 0
for kind in S.ADVERSARIAL_KINDS:
    seqs, refs = S.adversarial(kind, 500, seed=1)
    sb, so = O.pack(seqs)
    rb, ro = O.pack(refs)
    tot += check(sb, so, rb, ro, kind)
for cfg, n in ((1, 2000), (2, 2000), (3, 500)):
    tot += check(*S.generate(cfg, n), f"C{cfg}")
print("TOTAL BAD", tot, flush=True)

for cfg, n in ((2, 20000), (2, 200000), (1, 200000), (3, 50000)):
    sb, so, rb, ro = S.generate(cfg, n)
    for rep in range(2):
        t = time.time()
        res, cig = al.align_blobs(sb, so, rb, ro)
        dt = time.time() - t
    st = al.stats()
    cells = float((np.diff(so) * np.diff(ro)).sum())
    print(json.dumps({"cfg": cfg, "n": n, "wall_s": round(dt, 4), "GCUPS_wall": round(cells / dt / 1e9, 1),
                      "GCUPS_fwd_kernel": round(cells / st["ms_forward"] / 1e6, 1),
                      **{k: (round(v, 3) if isinstance(v, float) else v) for k, v in st.items()}}), flush=True)
