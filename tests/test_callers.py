"""The reference's two callers of the hot path, batched (gridss_b200/callers.py):
known answers from BreakpointHomologyTest / SAMRecordUtilTest (tests/golden/golden_callers.json), the window-descriptor
formulation against a one-pair-at-a-time string restatement, on the CPU through the oracle and on the GPU through the
C ABI (cudasw_reference_load + cudasw_align_batch_windows)."""
import itertools
import json
import os

import numpy as np
import pytest

from oracle.callers_restatement import homology_strings, rc, realign_strings
from gridss_b200 import callers as C
from gridss_b200.aligner import RESULT_DTYPE, WINDOW_REVCOMP

GOLD = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "golden_callers.json")))


def contigs_of(case):
    named = {"@polyA10000": b"A" * 10000, "@random600": GOLD["random600"].encode()}
    return [named.get(c) or c.encode() for c in case["contigs"]]


def read_of(case):
    r = case["read"]
    return (r[0], r[1], r[2], r[3].encode())


class OracleWindowAligner:
    """align_blobs_windows of CudaAligner restated on the host for the CPU tests: windows are cut from the contigs
    (upper-cased, as the packed reference stores them) and the pairs go to the oracle."""

    def __init__(self, oracle, contigs):
        self.O, self.contigs = oracle, [bytes(c).upper() for c in contigs]
        self.calls = 0

    def align_blobs_windows(self, seq_blob, seq_off, win, win_off, res=None, cig=None):
        self.calls += 1
        n = len(seq_off) - 1
        seqs = [seq_blob[seq_off[i]:seq_off[i + 1]].tobytes() for i in range(n)]
        refs = []
        for i in range(n):
            parts = []
            for w in win[win_off[i]:win_off[i + 1]]:
                b = self.contigs[w["contig"]][int(w["start"]):int(w["end"])]
                parts.append(rc(b) if w["flags"] & WINDOW_REVCOMP else b)
            refs.append(b"".join(parts))
        fn = self.O.ref_align if self.O.have_reference() else self.O.model_align
        r = fn(seqs, refs)
        out = np.zeros(n, dtype=RESULT_DTYPE)
        for f in out.dtype.names:
            if f in r.res.dtype.names:
                out[f] = r.res[f]
        return out, r.cig


def model_align_one(oracle):
    def f(seq, ref):
        score, start, cigar = oracle.model_align([seq], [ref]).tuple(0)
        return start, cigar
    return f


def bp_of(t):
    return C.BreakpointSummary(*t)


def random_genome(rng, n_contigs=3, length=3000):
    """Contigs with planted repeats so that breakpoints between copies have real homology."""
    out = []
    unit = rng.integers(0, 4, size=60)
    for _ in range(n_contigs):
        a = rng.integers(0, 4, size=length)
        for p in rng.integers(0, length - 60, size=12):
            k = int(rng.integers(10, 60))
            a[p:p + k] = unit[:k]
        s = np.frombuffer(b"ACGT", dtype=np.uint8)[a].copy()
        s[rng.integers(0, length, size=5)] = ord("N")
        out.append(s.tobytes())
    return out


def random_breakpoints(rng, contigs, n):
    bps, ins = [], []
    for _ in range(n):
        r1, r2 = (int(x) for x in rng.integers(0, len(contigs), size=2))
        if rng.random() < 0.5:
            r2 = r1
        p1 = int(rng.integers(1, len(contigs[r1]) + 1))
        p2 = int(rng.integers(1, len(contigs[r2]) + 1))
        if r1 == r2 and rng.random() < 0.3:
            p2 = min(len(contigs[r1]), p1 + int(rng.integers(0, 40)))
        bps.append((r1, "fb"[int(rng.integers(0, 2))], p1, r2, "fb"[int(rng.integers(0, 2))], p2))
        k = int(rng.integers(0, 8)) if rng.random() < 0.4 else 0
        ins.append("".join("ACGT"[int(x)] for x in rng.integers(0, 4, size=k)))
    return bps, ins


# ---------------------------------------------------------------- CPU: host logic through the oracle


def test_restatement_reproduces_reference_expectations(oracle):
    f = model_align_one(oracle)
    for c in GOLD["homology"]:
        got = homology_strings(f, contigs_of(c), tuple(c["breakpoint"]), c["inserted"], c["maxBreakendLength"], c["margin"])
        assert got == (c["local"], c["remote"]), c["cite"]
    for c in GOLD["realign"]:
        assert realign_strings(f, contigs_of(c), read_of(c), c["windowSize"], c["extend"]) == (c["start"], c["cigar"]), c["cite"]


def test_batched_homology_known_answers_cpu(oracle):
    for c in GOLD["homology"]:
        contigs = contigs_of(c)
        al = OracleWindowAligner(oracle, contigs)
        got = C.breakpoint_homology_batch(al, contigs, [bp_of(c["breakpoint"])], [c["inserted"]], c["maxBreakendLength"], c["margin"])
        assert got == [(c["local"], c["remote"])], c["cite"]
        assert al.calls == 1


def test_batched_homology_matches_string_restatement_cpu(oracle):
    rng = np.random.default_rng(5)
    contigs = random_genome(rng)
    bps, ins = random_breakpoints(rng, contigs, 400)
    al = OracleWindowAligner(oracle, contigs)
    f = model_align_one(oracle)
    for mbl, margin in ((20, 10), (300, 0), (7, 50)):
        got = C.breakpoint_homology_batch(al, contigs, [bp_of(b) for b in bps], ins, mbl, margin)
        want = [homology_strings(f, [c.upper() for c in contigs], b, i, mbl, margin) for b, i in zip(bps, ins)]
        assert got == want
        assert any(x != (0, 0) for x in want)
    assert al.calls == 3   # one batch per list, not one call per pair


def bounds_cases(step):
    n = 11
    out = []
    for b1, b2, d1, d2 in itertools.product(range(1, n + 1), range(1, n + 1), "fb", "fb"):
        out.append((0, d1, b1, 1, d2, b2))
    return out[::step]


def test_homology_never_leaves_contig_bounds_cpu(oracle):
    """BreakpointHomologyTest.java:109-131 (subsampled here, in full on the GPU): no homology, no out-of-range window."""
    contigs = [b"C" * 11, b"A" * 11]
    al = OracleWindowAligner(oracle, contigs)
    bps = [bp_of(b) for b in bounds_cases(7)]
    for mbl, margin in ((1, 0), (5, 12), (12, 3)):
        got = C.breakpoint_homology_batch(al, contigs, bps, [""] * len(bps), mbl, margin)
        assert got == [(0, 0)] * len(bps)


def test_event_size_and_exactness():
    assert bp_of((0, "f", 100, 0, "b", 103)).getEventSize() == 2      # BreakpointSummary.java:226-236
    assert bp_of((0, "b", 100, 0, "f", 103)).getEventSize() == 4
    assert bp_of((0, "f", 100, 0, "f", 103)).getEventSize() == 3
    assert bp_of((0, "f", 100, 1, "b", 103)).getEventSize() is None
    b = bp_of((0, "f", 6, 1, "b", 7))
    assert b.remoteBreakpoint().remoteBreakpoint() == b
    assert b.local().advance(3).position == 9 and b.remoteBreakend().advance(3).position == 4
    assert C.BreakendSummary(0, "f", 2).anchor_segment([12], 5) == (0, 0, 2, False)
    assert C.BreakendSummary(0, "b", 11).anchor_segment([12], 5) == (0, 10, 12, True)
    assert C.BreakendSummary(0, "f", 20).advance(10).anchor_segment([12], 5) is None


def test_batched_realign_cpu(oracle):
    c = GOLD["realign"][0]
    contigs = contigs_of(c)
    al = OracleWindowAligner(oracle, contigs)
    read = read_of(c)
    rng = np.random.default_rng(9)
    g = GOLD["random600"][:400]
    reads = [read, (2, 101, "100M", b""), (2, 50, "10S80M10S", (g[30:80] + g[95:145]).encode()),
             (2, 380, "5S30M", g[360:395].encode()), (2, 1, "20M3S", ("TTT" + g[:20]).encode())]
    for _ in range(40):
        p = int(rng.integers(1, 300))
        k = int(rng.integers(20, 90))
        s = bytearray(g[p - 1:p - 1 + k].encode())
        if rng.random() < 0.5:
            s[k // 2] = ord("n")
        lead = int(rng.integers(0, 6))
        reads.append((2, p, (f"{lead}S" if lead else "") + f"{k}M", b"G" * lead + bytes(s)))
    f = model_align_one(oracle)
    for window, extend in ((0, True), (10, False), (500, True)):
        got = C.realign_batch(al, [len(x) for x in contigs], reads, window, extend)
        assert got == [realign_strings(f, contigs, r, window, extend) for r in reads]
    assert C.realign_batch(al, [len(x) for x in contigs], reads[1:2], 0, True) == [(101, "100M")]   # no bases: unchanged


def realign_known_answers(make_aligner):
    """Every realign case of the reference's tests, grouped by genome so each group is one batch."""
    groups = {}
    for c in GOLD["realign"]:
        groups.setdefault((tuple(c["contigs"]), c["windowSize"], c["extend"]), []).append(c)
    for (_, window, extend), cases in groups.items():
        contigs = contigs_of(cases[0])
        got = C.realign_batch(make_aligner(contigs), [len(x) for x in contigs], [read_of(c) for c in cases], window, extend)
        for c, g in zip(cases, got):
            assert g == (c["start"], c["cigar"]), c["cite"]


def test_batched_realign_known_answers_cpu(oracle):
    realign_known_answers(lambda contigs: OracleWindowAligner(oracle, contigs))


def streaming_case(make_aligner, oracle, n_reads=60, contig_len=600, batch=16):
    rng = np.random.default_rng(21)
    contig = GOLD["random600"][:contig_len].encode() if contig_len <= 600 else random_genome(rng, 1, contig_len)[0]
    contigs = [b"ACGT" * 5, b"A" * 30, contig]
    sw = C.SmithWatermanStreamingAligner(make_aligner(contigs), [len(x) for x in contigs], 2, batch=batch)
    reads = []
    for k in range(n_reads):
        p, q = int(rng.integers(0, contig_len - 120)), int(rng.integers(20, 120))
        b = bytearray(contig[p:p + q])
        if k % 3 == 0:
            b[q // 2:q // 2] = b"TTTT"
        if k % 4 == 0:
            b = bytearray(b"GGGGGG") + b
        reads.append((f"r{k}", bytes(b)))
    for name, b in reads:
        sw.asyncAlign(name, b)
    assert sw.processedAlignmentRecords() == n_reads // batch * batch and sw.outstandingAlignmentRecord() == n_reads % batch
    sw.flush()
    assert sw.processedAlignmentRecords() == n_reads and sw.outstandingAlignmentRecord() == 0
    want = oracle.model_align([b for _, b in reads], [contig.upper()] * n_reads)
    for i, (name, b) in enumerate(reads):
        score, start, cigar = want.tuple(i)
        assert sw.getAlignment() == (name, 2, start + 1, cigar)      # SmithWatermanStreamingAligner.java:36-38
    with pytest.raises(ValueError):
        sw.asyncAlign("empty", b"")


def test_streaming_whole_contig_aligner_cpu(oracle):
    streaming_case(lambda contigs: OracleWindowAligner(oracle, contigs), oracle)


# ---------------------------------------------------------------- GPU: through the C ABI


@pytest.mark.gpu
def test_streaming_whole_contig_aligner_gpu(aligner, oracle):
    def loaded(contigs):
        aligner.load_reference(contigs)
        return aligner
    streaming_case(loaded, oracle)
    streaming_case(loaded, oracle, n_reads=300, contig_len=12000, batch=128)


@pytest.mark.gpu
def test_batched_homology_known_answers_gpu(aligner):
    for c in GOLD["homology"]:
        contigs = contigs_of(c)
        aligner.load_reference(contigs)
        got = C.breakpoint_homology_batch(aligner, contigs, [bp_of(c["breakpoint"])], [c["inserted"]], c["maxBreakendLength"], c["margin"])
        assert got == [(c["local"], c["remote"])], c["cite"]


@pytest.mark.gpu
def test_homology_never_leaves_contig_bounds_gpu(aligner):
    """BreakpointHomologyTest.java:109-131 in full: 484 breakpoints x 12 lengths x 13 margins, one batch per setting."""
    contigs = [b"C" * 11, b"A" * 11]
    aligner.load_reference(contigs)
    bps = [bp_of(b) for b in bounds_cases(1)]
    for mbl in range(1, 13):
        for margin in range(0, 13):
            got = C.breakpoint_homology_batch(aligner, contigs, bps, [""] * len(bps), mbl, margin)
            assert got == [(0, 0)] * len(bps), (mbl, margin)


@pytest.mark.gpu
def test_batched_homology_gpu_matches_oracle(aligner, oracle):
    rng = np.random.default_rng(11)
    contigs = random_genome(rng, n_contigs=4, length=20000)
    bps, ins = random_breakpoints(rng, contigs, 20000)
    aligner.load_reference(contigs)
    ref = OracleWindowAligner(oracle, contigs)
    for mbl, margin in ((20, 10), (300, 0), (100, 100)):
        got = C.breakpoint_homology_batch(aligner, contigs, [bp_of(b) for b in bps], ins, mbl, margin)
        want = C.breakpoint_homology_batch(ref, contigs, [bp_of(b) for b in bps], ins, mbl, margin)
        assert got == want
        f = model_align_one(oracle)
        assert got[:200] == [homology_strings(f, [c.upper() for c in contigs], b, i, mbl, margin) for b, i in zip(bps[:200], ins[:200])]


@pytest.mark.gpu
def test_batched_realign_gpu(aligner, oracle):
    def loaded(contigs):
        aligner.load_reference(contigs)
        return aligner
    realign_known_answers(loaded)
    rng = np.random.default_rng(13)
    genome = random_genome(rng, n_contigs=2, length=50000)
    aligner.load_reference(genome)
    reads = []
    for _ in range(5000):
        ci = int(rng.integers(0, 2))
        p = int(rng.integers(1, 49000))
        k = int(rng.integers(30, 400))
        s = bytearray(genome[ci][p - 1:p - 1 + k])
        if rng.random() < 0.5 and k > 60:
            del s[k // 2:k // 2 + int(rng.integers(1, 9))]
        lead = int(rng.integers(0, 20))
        reads.append((ci, p, (f"{lead}S" if lead else "") + f"{len(s)}M", bytes(rng.integers(65, 70, size=lead, dtype=np.uint8)) + bytes(s)))
    ref = OracleWindowAligner(oracle, genome)
    lens = [len(x) for x in genome]
    for window, extend in ((0, True), (25, False)):
        assert C.realign_batch(aligner, lens, reads, window, extend) == C.realign_batch(ref, lens, reads, window, extend)


@pytest.mark.gpu
def test_device_homology_known_answers(aligner):
    """cudasw_breakpoint_homology: the reference tests' expected lengths, segments derived and gathered by the library."""
    for c in GOLD["homology"]:
        contigs = contigs_of(c)
        aligner.load_reference(contigs)
        got = C.breakpoint_homology_device(aligner, [bp_of(c["breakpoint"])], [c["inserted"]], c["maxBreakendLength"], c["margin"])
        assert got == [(c["local"], c["remote"])], c["cite"]
        st = aligner.stats()
        assert st["d2h_bytes"] == 32 * st["pairs"]      # clip lengths only: no traceback ran, no CIGAR came back


@pytest.mark.gpu
def test_device_homology_contig_bounds(aligner):
    contigs = [b"C" * 11, b"A" * 11]
    aligner.load_reference(contigs)
    bps = C.breakpoint_array([bp_of(b) for b in bounds_cases(1)])
    for mbl in range(1, 13):
        for margin in range(0, 13):
            local, remote = aligner.breakpoint_homology(bps, None, None, mbl, margin)
            assert not local.any() and not remote.any(), (mbl, margin)


@pytest.mark.gpu
def test_device_homology_matches_batched_and_restatement(aligner, oracle):
    rng = np.random.default_rng(17)
    contigs = random_genome(rng, n_contigs=4, length=20000)
    contigs[1] = contigs[1].lower()
    bps, ins = random_breakpoints(rng, contigs, 30000)
    ins = [i.lower() if k % 5 == 0 else i for k, i in enumerate(ins)]
    aligner.load_reference(contigs)
    f = model_align_one(oracle)
    for mbl, margin in ((20, 10), (300, 10), (100, 100), (0, 5)):
        got = C.breakpoint_homology_device(aligner, [bp_of(b) for b in bps], ins, mbl, margin)
        want = C.breakpoint_homology_batch(aligner, contigs, [bp_of(b) for b in bps], ins, mbl, margin)
        assert got == want
        assert got[:150] == [homology_strings(f, [c.upper() for c in contigs], b, i, mbl, margin) for b, i in zip(bps[:150], ins[:150])]
    # no inserted sequence at all: ins_off may be NULL
    a = C.breakpoint_homology_device(aligner, [bp_of(b) for b in bps[:2000]], None, 300, 10)
    b = C.breakpoint_homology_device(aligner, [bp_of(b) for b in bps[:2000]], [""] * 2000, 300, 10)
    assert a == b
    assert C.breakpoint_homology_device(aligner, [], None, 300, 10) == []


@pytest.mark.gpu
def test_device_homology_rejects_bad_breakpoints(aligner):
    import gridss_b200 as g
    aligner.load_reference([b"ACGT" * 10])
    bad = np.zeros(1, dtype=g.BREAKPOINT_DTYPE)
    bad[0] = (0, 0, 5, 3, 1, 6)
    with pytest.raises(ValueError):
        aligner.breakpoint_homology(bad, None, None, 10, 0)
    bad[0] = (0, 2, 5, 0, 1, 6)
    with pytest.raises(ValueError):
        aligner.breakpoint_homology(bad, None, None, 10, 0)
    fresh = g.CudaAligner()
    bad[0] = (0, 0, 5, 0, 1, 6)
    with pytest.raises(ValueError):     # no reference loaded
        fresh.breakpoint_homology(bad, None, None, 10, 0)


@pytest.mark.gpu
def test_multi_handle_homology_matches_single(aligner):
    """cudasw_multi_breakpoint_homology: the list cut over the GPUs of the box (twice the same GPU on a 1-GPU box)."""
    import torch

    import gridss_b200 as g
    n_dev = torch.cuda.device_count()
    ma = g.MultiGpuAligner(devices=list(range(n_dev)) if n_dev > 1 else [0, 0])
    rng = np.random.default_rng(23)
    contigs = random_genome(rng, n_contigs=3, length=15000)
    bps, ins = random_breakpoints(rng, contigs, 9001)
    aligner.load_reference(contigs)
    ma.load_reference(contigs)
    bp = C.breakpoint_array([bp_of(b) for b in bps])
    blob, off = g.aligner.pack([i.encode() for i in ins])
    for mbl, margin in ((300, 10), (25, 0)):
        l1, r1 = aligner.breakpoint_homology(bp, blob, off, mbl, margin)
        lm, rm = ma.breakpoint_homology(bp, blob, off, mbl, margin)
        assert np.array_equal(l1, lm) and np.array_equal(r1, rm)
    l0, r0 = ma.breakpoint_homology(bp, None, None, 300, 10)
    assert np.array_equal(l0, aligner.breakpoint_homology(bp, None, None, 300, 10)[0])
    ma.close()
