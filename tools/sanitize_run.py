"""Small end-to-end run for compute-sanitizer (memcheck): every kernel family on small, awkward inputs."""
import sys

import numpy as np

sys.path.insert(0, ".")
import gridss_b200 as g
from gridss_b200.aligner import pack
from tools import synth as S

al = g.CudaAligner()
for kind in ("mutated", "with_n", "indel_heavy", "tiny", "q_gt_w", "zero", "tandem"):
    seqs, refs = S.adversarial(kind, 150, seed=3, qmax=260, wmax=700)
    al.align_batch(seqs, refs)
for cfg, n in ((1, 300), (2, 200), (3, 120), (5, 2)):  # uniform batches (C1, C2) take the quiet-chain flavour; CUDASW_CHAIN=1 forces it everywhere
    sb, so, rb, ro = S.generate(cfg, n)
    al.align_blobs(sb, so, rb, ro)
# wide-band traceback paths
rng = np.random.default_rng(1)
a = rng.choice(list(b"ACGT"), size=500).astype(np.uint8).tobytes()
b = rng.choice(list(b"ACGT"), size=500).astype(np.uint8).tobytes()
gap = rng.choice(list(b"ACGT"), size=700).astype(np.uint8).tobytes()
al.align_batch([a + b, a + gap + b, a + b], [a + gap + b, a + b, a + gap[:40] + b])
# device-resident reference + windows, multi handle, dispatcher
contig = rng.choice(list(b"ACGTN"), size=50001, p=[.24, .24, .24, .24, .04]).astype(np.uint8).tobytes()
al.load_reference([contig, contig[:33]])
win = np.zeros(40, dtype=g.WINDOW_DTYPE)
seqs = []
for i in range(40):
    s = int(rng.integers(0, 48000)); L = int(rng.integers(30, 1500))
    win[i] = (0, i & 1, s, s + L)
    seqs.append(al.get_subsequence(0, s, s + L, bool(i & 1))[: max(10, L // 3)].replace(b"N", b"A"))
sb, so = pack(seqs)
al.align_blobs_windows(sb, so, win, np.arange(41, dtype=np.int64))
ma = g.MultiGpuAligner(devices=[0, 0])
sb, so, rb, ro = S.generate(4, 200)
ma.align_blobs(sb, so, rb, ro)
ma.close()
al.start_dispatcher(16, 0.5)
ts = [al.submit(seqs[i], contig[:2000]) for i in range(40)]
for t in ts:
    assert al.wait(t, 20000) is not None
al.stop_dispatcher()
al.close()
print("sanitize run done")
