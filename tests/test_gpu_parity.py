"""GPU: the CUDA path (through the C ABI) against the oracle restatement, bit-exact
(score, start, end, read begin/end, CIGAR). The oracle itself is pinned in test_oracle.py."""
import json
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
FIELDS = ["score", "ref_begin", "ref_end", "read_begin", "read_end", "cigar_len", "status"]


def _cig(res, cig, i):
    r = res[i]
    return tuple(int(x) for x in cig[r["cigar_off"]: r["cigar_off"] + r["cigar_len"]])


def _compare_blobs(aligner_obj, oracle, sb, so, rb, ro, scoring=(1, -4, -4, 6, 1), threads=16):
    m = oracle.model_align_blobs(sb, so, rb, ro, scoring, threads=threads)
    res, cig = aligner_obj.align_blobs(sb, so, rb, ro)
    bad = []
    undef = 0
    for i in range(len(res)):
        if m.res[i]["status"] == -5 or res[i]["status"] == -5:
            undef += 1
            continue
        a = tuple(int(res[i][f]) for f in FIELDS)
        b = tuple(int(m.res[i][f]) for f in FIELDS)
        ca = _cig(res, cig, i)
        cb = tuple(int(x) for x in m.cig[m.res[i]["cigar_off"]: m.res[i]["cigar_off"] + m.res[i]["cigar_len"]])
        if a != b or ca != cb:
            bad.append((i, a, b, ca[:6], cb[:6]))
    return bad, undef


def _compare(aligner_obj, oracle, seqs, refs, scoring=(1, -4, -4, 6, 1)):
    from gridss_b200.aligner import pack

    sb, so = pack(seqs)
    rb, ro = pack(refs)
    return _compare_blobs(aligner_obj, oracle, sb, so, rb, ro, scoring)


def test_reference_golden_vectors(aligner):
    g = json.load(open(os.path.join(HERE, "golden", "golden_reference_tests.json")))
    for x in g:
        a = aligner.align_smith_waterman(x["seq"].encode("latin1"), x["ref"].encode("latin1"))
        assert (a.score, a.getStartPosition(), a.getCigar()) == (x["score"], x["start"], x["cigar"]), x["cite"]


def test_sswjni_golden_fuzz(aligner):
    import gridss_b200 as gb

    g = json.load(open(os.path.join(HERE, "golden", "golden_sswjni_fuzz.json")))
    by_sc = {}
    for x in g:
        by_sc.setdefault(tuple(x["scoring"]), []).append(x)
    for sc, xs in by_sc.items():
        al = aligner if sc == (1, -4, -4, 6, 1) else gb.AlignerFactory.create(*sc)
        out = al.align_batch([x["seq"].encode("latin1") for x in xs], [x["ref"].encode("latin1") for x in xs])
        for a, x in zip(out, xs):
            assert (a.score, a.getStartPosition(), a.getCigar()) == (x["score"], x["start"], x["cigar"]), (x["kind"], sc)


@pytest.mark.parametrize("kind", ["random", "mutated", "two_letter", "homopolymer", "dinuc", "tandem", "subst_block",
                                  "with_n", "zero", "tiny", "q_gt_w", "indel_heavy"])
def test_adversarial_vs_oracle(aligner, oracle, synth, kind):
    for seed, qmax, wmax in ((1, 200, 400), (2, 700, 1500)):
        seqs, refs = synth.adversarial(kind, 1500, seed=seed, qmax=qmax, wmax=wmax)
        bad, undef = _compare(aligner, oracle, seqs, refs)
        assert bad == [], bad[:5]
        assert undef == 0


@pytest.mark.parametrize("cfg,n", [(1, 10000), (2, 4000), (3, 1500), (4, 4000)])
def test_synthetic_configs_vs_oracle(aligner, oracle, synth, cfg, n):
    sb, so, rb, ro = synth.generate(cfg, n)
    bad, undef = _compare_blobs(aligner, oracle, sb, so, rb, ro)
    assert bad == [], bad[:5]
    assert undef == 0


@pytest.mark.parametrize("sc", [(2, -6, -1, 5, 3), (1, -1, -1, 2, 1), (3, -2, -5, 4, 0), (5, -4, -2, 10, 1)])
def test_other_scoring_vs_oracle(oracle, synth, sc):
    import gridss_b200 as gb

    al = gb.AlignerFactory.create(*sc)
    for kind in ("mutated", "with_n", "indel_heavy", "subst_block"):
        seqs, refs = synth.adversarial(kind, 800, seed=9)
        bad, undef = _compare(al, oracle, seqs, refs, sc)
        assert bad == [], (kind, bad[:5])
    al.close()


def test_every_variant_boundary(aligner, oracle):
    """query lengths straddling every (G, R) variant's row capacity"""
    rng = np.random.default_rng(3)
    seqs, refs = [], []
    for rows in (16, 24, 32, 40, 48, 64, 80, 96, 128, 160, 192, 256, 320, 384, 512, 640, 768, 1024, 1280, 1536, 2048):
        for q in (rows - 1, rows, rows + 1):
            if q > 2048:
                continue
            w = int(rng.integers(q, 3 * q + 50))
            r = bytes(rng.choice(list(b"ACGT"), size=w).tolist())
            st = int(rng.integers(0, w - q + 1))
            s = bytearray(r[st:st + q])
            for k in range(max(1, q // 30)):
                s[int(rng.integers(0, q))] = int(rng.choice(list(b"ACGT")))
            seqs.append(bytes(s))
            refs.append(r)
    bad, undef = _compare(aligner, oracle, seqs, refs)
    assert bad == [] and undef == 0, bad[:5]


def test_empty_and_oversize_inputs(aligner):
    import gridss_b200 as gb
    from gridss_b200.aligner import pack

    with pytest.raises(ValueError):
        aligner.align_smith_waterman(b"", b"ACGT")  # SswJniAligner.java:45-47
    with pytest.raises(ValueError):
        aligner.align_smith_waterman(b"ACGT", b"")  # SswJniAligner.java:48-50
    # per-pair status: one bad pair does not fail the batch
    sb, so = pack([b"ACGT", b"", b"A" * 40000, b"GGGG"])
    rb, ro = pack([b"ACGT", b"ACGT", b"A" * 50, b"CCGGGGCC"])
    res, cig = aligner.align_blobs(sb, so, rb, ro)
    assert [int(x) for x in res["status"]] == [0, gb.aligner.CUDASW_EINVAL, gb.aligner.CUDASW_ETOOLARGE, 0]
    assert int(res[3]["ref_begin"]) == 2
    # empty batch
    res, cig = aligner.align_blobs(np.zeros(1, np.uint8), np.zeros(1, np.int64), np.zeros(1, np.uint8), np.zeros(1, np.int64))
    assert len(res) == 0


def test_async_trio_matches_sync(aligner, synth):
    import gridss_b200 as gb

    seqs, refs = synth.adversarial("mutated", 300, seed=77)
    sync = aligner.align_batch(seqs, refs)
    b = gb.BatchingAligner(aligner, batch_size=128)
    for s, r in zip(seqs, refs):
        b.asyncAlign(s, r)
    b.flush()
    assert b.outstandingAlignmentRecord() == 0
    got = [b.getAlignment() for _ in seqs]
    assert got == sync
    with pytest.raises(RuntimeError):
        b.getAlignment()


def test_device_resident_entry_point(aligner, oracle, synth):
    torch = pytest.importorskip("torch")
    sb, so, rb, ro = synth.generate(2, 2000)
    d = [torch.from_numpy(x).cuda() for x in (sb, so, rb, ro)]
    torch.cuda.synchronize()
    res, cig = aligner.align_blobs_device(d[0].data_ptr(), d[1].data_ptr(), d[2].data_ptr(), d[3].data_ptr(), so, ro)
    res2, cig2 = aligner.align_blobs(sb, so, rb, ro)
    for f in FIELDS:
        assert np.array_equal(res[f], res2[f])
    assert all(_cig(res, cig, i) == _cig(res2, cig2, i) for i in range(len(res)))


def test_order_invariance_and_idempotence(aligner, synth):
    """size-independent properties: results do not depend on batch composition or order"""
    sb, so, rb, ro = synth.generate(4, 20000)
    res, cig = aligner.align_blobs(sb, so, rb, ro)
    res2, cig2 = aligner.align_blobs(sb, so, rb, ro)
    assert np.array_equal(res[FIELDS], res2[FIELDS])
    idx = np.random.default_rng(0).permutation(20000)[:3000]
    seqs = [bytes(sb[so[i]:so[i + 1]]) for i in idx]
    refs = [bytes(rb[ro[i]:ro[i + 1]]) for i in idx]
    out = aligner.align_batch(seqs, refs)
    for a, i in zip(out, idx):
        assert a.getStartPosition() == int(res[i]["ref_begin"])
        assert a.score == int(res[i]["score"])
        from gridss_b200 import cigar_string
        assert a.getCigar() == cigar_string(_cig(res, cig, i))


def test_long_queries_strip_mined(aligner, oracle, synth):
    """queries longer than one 2048-row strip go through the strip-mined warp kernel (C5 path)"""
    rng = np.random.default_rng(11)
    seqs, refs = [], []
    for q in (2049, 2500, 2561, 3000, 4096, 4097, 5121, 6500):
        w = int(q * rng.uniform(1.0, 2.0)) + 10
        r = rng.choice(list(b"ACGT"), size=w).astype(np.uint8)
        st = int(rng.integers(0, w - q + 1))
        s = r[st:st + q].copy()
        for k in range(q // 40):
            s[int(rng.integers(0, q))] = int(rng.choice(list(b"ACGT")))
        s = bytearray(s.tobytes())
        # a few indels
        for k in range(3):
            p = int(rng.integers(50, len(s) - 50))
            if rng.random() < 0.5:
                del s[p:p + int(rng.integers(1, 20))]
            else:
                s[p:p] = bytes(rng.choice(list(b"ACGT"), size=int(rng.integers(1, 20))).tolist())
        seqs.append(bytes(s))
        refs.append(r.tobytes())
    bad, undef = _compare(aligner, oracle, seqs, refs)
    assert bad == [] and undef == 0, bad[:3]


@pytest.mark.parametrize("flavour", ["cta_per_pair", "warp_per_pair"])
def test_c5_long_contigs_vs_oracle(aligner, oracle, synth, flavour, monkeypatch):
    """BASELINE.json C5 (10 kb contigs vs 20 kb windows: strip-mined score passes, band doubling, long CIGARs), 64 pairs"""
    if flavour == "warp_per_pair":
        monkeypatch.setenv("CUDASW_TB_NO_BLOCK", "1")
    sb, so, rb, ro = synth.generate(5, 64)
    bad, undef = _compare_blobs(aligner, oracle, sb, so, rb, ro)
    assert bad == [] and undef == 0, bad[:3]


def test_chunked_pipeline_matches_oracle(aligner, oracle, synth, monkeypatch):
    """host batches are split into chunks whose H2D overlaps the previous chunk's kernels; force many small chunks"""
    monkeypatch.setenv("CUDASW_CHUNK_MB", "1")
    sb, so, rb, ro = synth.generate(4, 4000)
    bad, undef = _compare_blobs(aligner, oracle, sb, so, rb, ro)
    assert bad == [] and undef == 0, bad[:5]
    res, cig = aligner.align_blobs(sb, so, rb, ro)
    monkeypatch.delenv("CUDASW_CHUNK_MB")
    res2, cig2 = aligner.align_blobs(sb, so, rb, ro)
    for f in ("score", "ref_begin", "ref_end", "read_begin", "read_end", "cigar_len", "status"):
        assert np.array_equal(res[f], res2[f])
    assert all(_cig(res, cig, i) == _cig(res2, cig2, i) for i in range(len(res)))


@pytest.mark.parametrize("kind", ["random", "mutated", "two_letter", "homopolymer", "dinuc", "tandem", "subst_block",
                                  "with_n", "zero", "tiny", "q_gt_w", "indel_heavy"])
@pytest.mark.parametrize("flavour", ["cta_per_pair", "warp_per_pair"])
def test_warp_traceback_forced(aligner, oracle, synth, kind, flavour, monkeypatch):
    """every pair through the cooperative traceback kernels (normally only long / wide-band pairs take them): the
    CTA-per-pair kernel small batches get, and the one-warp-per-pair kernels of large batches"""
    monkeypatch.setenv("CUDASW_TB_WIDE_CELLS", "0")
    if flavour == "warp_per_pair":
        monkeypatch.setenv("CUDASW_TB_NO_BLOCK", "1")
    for seed, qmax, wmax in ((5, 120, 300), (6, 700, 1500)):
        seqs, refs = synth.adversarial(kind, 800, seed=seed, qmax=qmax, wmax=wmax)
        bad, undef = _compare(aligner, oracle, seqs, refs)
        assert bad == [], bad[:5]
        assert undef == 0


@pytest.mark.parametrize("flavour", ["cta_per_pair", "warp_per_pair"])
def test_warp_traceback_wide_bands(aligner, oracle, flavour, monkeypatch):
    """long indels force band doubling through every register frame width, the hand-over from the CTA-per-pair kernel
    to the one-warp kernel (bands above 511) and the memory-backed state (bands above 1023)"""
    monkeypatch.setenv("CUDASW_TB_WIDE_CELLS", "0")
    if flavour == "warp_per_pair":
        monkeypatch.setenv("CUDASW_TB_NO_BLOCK", "1")
    rng = np.random.default_rng(21)
    seqs, refs = [], []
    for gap in (40, 90, 200, 300, 450, 700, 1100, 1500):
        for kind in ("del", "ins"):
            a = rng.choice(list(b"ACGT"), size=400).astype(np.uint8).tobytes()
            b = rng.choice(list(b"ACGT"), size=400).astype(np.uint8).tobytes()
            g = rng.choice(list(b"ACGT"), size=gap).astype(np.uint8).tobytes()
            flank = rng.choice(list(b"ACGT"), size=100).astype(np.uint8).tobytes()
            if kind == "del":   # reference has extra bases
                seqs.append(a + b)
                refs.append(flank + a + g + b + flank)
            else:               # query has extra bases
                seqs.append(a + g + b)
                refs.append(flank + a + b + flank)
    bad, undef = _compare(aligner, oracle, seqs, refs)
    assert bad == [] and undef == 0, bad[:3]


# ---- device-resident reference + window descriptors
_COMP = bytes.maketrans(b"ACGTN", b"TGCAN")


def _toy_genome(rng):
    """contigs with lower case, N runs and stray IUPAC codes (all of which clean() maps the way the packing does)"""
    contigs = []
    for L in (70001, 33, 12345, 1, 40000):
        c = rng.choice(list(b"ACGT"), size=L).astype(np.uint8)
        for _ in range(max(1, L // 4000)):
            p = int(rng.integers(0, L))
            c[p:p + int(rng.integers(1, 60))] = ord("N")
        low = rng.random(L) < 0.1
        c[low] |= 0x20
        odd = rng.random(L) < 0.001
        c[odd] = ord("R")
        contigs.append(c.tobytes())
    return contigs


def _clean(b):
    u = bytes(b).upper()
    return bytes(x if x in b"ACGT" else ord("N") for x in u)


def test_reference_fetch_matches_getSubsequenceAt():
    import gridss_b200 as gb

    rng = np.random.default_rng(11)
    al = gb.CudaAligner()
    contigs = _toy_genome(rng)
    al.load_reference(contigs)
    for _ in range(300):
        c = int(rng.integers(0, len(contigs)))
        L = len(contigs[c])
        s = int(rng.integers(0, L))
        e = int(rng.integers(s, min(L, s + 3000) + 1))
        want = _clean(contigs[c][s:e])
        assert al.get_subsequence(c, s, e) == want
        assert al.get_subsequence(c, s, e, revcomp=True) == want.translate(_COMP)[::-1]
    assert al.get_subsequence(1, 0, 33) == _clean(contigs[1])
    with pytest.raises(ValueError):
        al.get_subsequence(1, 0, 34)
    with pytest.raises(ValueError):
        al.get_subsequence(9, 0, 1)
    al.close()


def test_window_descriptors_match_shipped_bytes(oracle):
    """cudasw_align_batch_windows == cudasw_align_batch on the same windows fetched on the host == oracle"""
    import gridss_b200 as gb
    from gridss_b200.aligner import pack

    rng = np.random.default_rng(12)
    al = gb.CudaAligner()
    contigs = _toy_genome(rng)
    al.load_reference(contigs)
    n = 3000
    seqs, refs, segs, win_off = [], [], [], [0]
    for i in range(n):
        nseg = 1 if i % 3 else 2  # BreakpointHomology-style two-piece windows every third pair
        w = b""
        for _ in range(nseg):
            c = int(rng.choice([0, 2, 4]))
            L = len(contigs[c])
            wl = int(rng.integers(50, 2500))
            s = int(rng.integers(0, L - wl))
            rcf = int(rng.integers(0, 2))
            piece = _clean(contigs[c][s:s + wl])
            w += piece.translate(_COMP)[::-1] if rcf else piece
            segs.append((c, rcf, s, s + wl))
        win_off.append(len(segs))
        ql = int(rng.integers(20, min(400, len(w))))
        qs = int(rng.integers(0, len(w) - ql))
        q = bytearray(w[qs:qs + ql])
        for _ in range(ql // 25):
            q[int(rng.integers(0, ql))] = int(rng.choice(list(b"ACGT")))
        if i % 5 == 0:
            p = int(rng.integers(1, ql - 1))
            del q[p:p + int(rng.integers(1, 12))]
        seqs.append(bytes(q))
        refs.append(w)
    sb, so = pack(seqs)
    rb, ro = pack(refs)
    win = np.array(segs, dtype=gb.WINDOW_DTYPE)
    woff = np.array(win_off, dtype=np.int64)
    res_w, cig_w = al.align_blobs_windows(sb, so, win, woff)
    assert al.stats()["h2d_bytes"] < int(so[-1]) + 64 * n + 1024  # no window bytes crossed PCIe
    res_b, cig_b = al.align_blobs(sb, so, rb, ro)
    m = oracle.model_align_blobs(sb, so, rb, ro, threads=16)
    for f in FIELDS:
        assert np.array_equal(res_w[f], res_b[f]), f
        assert np.array_equal(res_w[f], m.res[f]), f
    for i in range(n):
        assert _cig(res_w, cig_w, i) == _cig(res_b, cig_b, i)
        assert _cig(res_w, cig_w, i) == tuple(int(x) for x in m.cig[m.res[i]["cigar_off"]: m.res[i]["cigar_off"] + m.res[i]["cigar_len"]])

    # a bad descriptor fails its own pair only
    win2 = win.copy()
    win2[woff[7]]["end"] = len(contigs[win2[woff[7]]["contig"]]) + 5
    win2[woff[8]]["contig"] = 99
    res2, cig2 = al.align_blobs_windows(sb, so, win2, woff)
    assert res2[7]["status"] == -1 and res2[8]["status"] == -1
    ok = np.ones(n, dtype=bool)
    ok[[7, 8]] = False
    for f in FIELDS:
        assert np.array_equal(res2[f][ok], res_b[f][ok]), f
    al.close()


def test_windows_without_reference_is_an_error(aligner):
    import gridss_b200 as gb
    from gridss_b200.aligner import pack

    al = gb.CudaAligner()
    sb, so = pack([b"ACGT"])
    win = np.zeros(1, dtype=gb.WINDOW_DTYPE)
    with pytest.raises(gb.CudaSwError):
        al.align_blobs_windows(sb, so, win, np.array([0, 1], dtype=np.int64))
    al.close()


# ---- several engines behind one handle (cudasw_multi_*); on a one-GPU box the slices share device 0
def _visible_devices():
    import gridss_b200 as gb

    n = gb.load_library().cudasw_device_count()
    return list(range(n)) if n > 1 else [0, 0, 0]


def test_multi_handle_matches_single(aligner, synth):
    import gridss_b200 as gb

    ma = gb.MultiGpuAligner(devices=_visible_devices())
    for cfg, n in ((4, 6000), (3, 900)):
        sb, so, rb, ro = synth.generate(cfg, n)
        res1, cig1 = aligner.align_blobs(sb, so, rb, ro)
        resm, cigm = ma.align_blobs(sb, so, rb, ro)
        for f in FIELDS:
            assert np.array_equal(res1[f], resm[f]), f
        assert all(_cig(res1, cig1, i) == _cig(resm, cigm, i) for i in range(n))
        sl = ma.slice_stats()
        assert sum(s["pairs"] for s in sl) == n and all(s["pairs"] > 0 for s in sl)
        cost = [float((np.diff(so)[s["first_pair"]: s["first_pair"] + s["pairs"]] * np.diff(ro)[s["first_pair"]: s["first_pair"] + s["pairs"]]).sum()) for s in sl]
        assert max(cost) < 1.25 * min(cost) + 3e7  # slices are cost-balanced
    # a CIGAR blob that is too small is reported with a capacity that works
    sb, so, rb, ro = synth.generate(4, 3000)
    small = np.zeros(3000, dtype=np.uint32)
    resm, cigm = ma.align_blobs(sb, so, rb, ro, cig=small)
    res1, cig1 = aligner.align_blobs(sb, so, rb, ro)
    assert all(_cig(res1, cig1, i) == _cig(resm, cigm, i) for i in range(3000))
    # fewer pairs than engines, and the single-pair interface
    a = ma.align_smith_waterman(b"AAACCCCCCCCCCCCCTTAATTAATTAATTTT", b"AAACCCCCCCCCCCCCGTTAATTAATTAATTTT")
    assert (a.getStartPosition(), a.getCigar()) == (0, "16M1D16M")
    ma.close()


def test_multi_handle_windows(aligner):
    import gridss_b200 as gb
    from gridss_b200.aligner import pack

    rng = np.random.default_rng(14)
    contigs = _toy_genome(rng)
    ma = gb.MultiGpuAligner(devices=_visible_devices())
    ma.load_reference(contigs)
    seqs, refs, segs = [], [], []
    for i in range(2000):
        c = int(rng.choice([0, 2, 4]))
        wl = int(rng.integers(100, 3000))
        s = int(rng.integers(0, len(contigs[c]) - wl))
        rcf = int(rng.integers(0, 2))
        w = _clean(contigs[c][s:s + wl])
        w = w.translate(_COMP)[::-1] if rcf else w
        ql = int(rng.integers(30, min(300, wl)))
        qs = int(rng.integers(0, wl - ql))
        seqs.append(w[qs:qs + ql])
        refs.append(w)
        segs.append((c, rcf, s, s + wl))
    sb, so = pack(seqs)
    rb, ro = pack(refs)
    win = np.array(segs, dtype=gb.WINDOW_DTYPE)
    woff = np.arange(len(segs) + 1, dtype=np.int64)
    resm, cigm = ma.align_blobs_windows(sb, so, win, woff)
    res1, cig1 = aligner.align_blobs(sb, so, rb, ro)
    for f in FIELDS:
        assert np.array_equal(res1[f], resm[f]), f
    assert all(_cig(res1, cig1, i) == _cig(resm, cigm, i) for i in range(len(segs)))
    ma.close()


def test_dispatcher_thread_batches_concurrent_submitters(oracle, synth):
    """worker threads only submit + wait; one native dispatcher thread turns the stream into GPU batches"""
    import threading

    import gridss_b200 as gb

    al = gb.CudaAligner()
    seqs, refs = synth.adversarial("mutated", 1200, seed=31, qmax=150, wmax=400)
    m = oracle.model_align(seqs, refs, threads=8)
    al.start_dispatcher(max_pairs=256, max_delay_ms=1.0)
    got = [None] * len(seqs)
    errs = []

    def worker(k, nthreads):
        try:
            for i in range(k, len(seqs), nthreads):
                t = al.submit(seqs[i], refs[i])
                if i % 3 == 0:          # some wait at once (age-triggered flush), some collect later
                    got[i] = al.wait(t, 20000)
                else:
                    got[i] = t
            for i in range(k, len(seqs), nthreads):
                if isinstance(got[i], int):
                    got[i] = al.wait(got[i], 20000)
        except Exception as e:  # pragma: no cover
            errs.append(e)

    ths = [threading.Thread(target=worker, args=(k, 6)) for k in range(6)]
    for t in ths:
        t.start()
    for t in ths:
        t.join()
    assert not errs, errs[:1]
    for i in range(len(seqs)):
        assert got[i] is not None, i
        assert (got[i].score, got[i].getStartPosition(), got[i].getCigar()) == m.tuple(i), i
    assert al.outstanding() == 0 and al.processed() == 0
    with pytest.raises(KeyError):
        al.wait(10**9, 10)
    # stop() flushes what is pending; the ticket can still be collected
    t = al.submit(b"ACGTACGTAC", b"TTACGTACGTACTT")
    al.stop_dispatcher()
    a = al.wait(t, 5000)
    assert a is not None and a.getCigar() == "10M"
    al.close()


def _low_score_pair(rng, q, w, plants):
    """Query over {G,T}, window over {A,C}: nothing matches except the planted k-mers, so the whole forward pass
    runs on quiet steps (every H <= gapOpen when k <= 6) and the end cell comes from the tie-breaks alone."""
    qs = rng.choice(np.frombuffer(b"GT", dtype=np.uint8), size=q)
    ws = rng.choice(np.frombuffer(b"AC", dtype=np.uint8), size=w)
    for k in plants:
        if k > q or k > w:
            continue
        i = int(rng.integers(0, q - k + 1))
        j = int(rng.integers(0, w - k + 1))
        ws[j:j + k] = qs[i:i + k]
    return qs.tobytes(), ws.tobytes()


@pytest.mark.parametrize("q", [30, 100, 150, 300, 304, 600, 1000])
def test_quiet_chain_edges_vs_oracle(aligner, oracle, q):
    """The chained quiet steps of the forward kernel (sw_kernels.cuh, qstep): windows around the 3 x rows gate, maxima
    that are found and tie-broken inside chains (scores <= gapOpen), chains broken at either ping-pong parity by
    planted 7..12-mers, windows of different length in one warp."""
    rng = np.random.default_rng(900 + q)
    rows = aligner.variant_for(q)[0] * aligner.variant_for(q)[1] * 2
    seqs, refs = [], []
    for w in (3 * rows - 2, 3 * rows - 1, 3 * rows, 3 * rows + 1, 3 * rows + 2, 3 * rows + 3, 3 * rows + 64, 4000, 4001):
        if w < q:
            continue
        for rep in range(24):
            kind = rep % 4
            if kind == 0:    # only short plants: everything is decided on quiet steps
                plants = [int(rng.integers(1, 7)) for _ in range(int(rng.integers(1, 12)))]
            elif kind == 1:  # equal plants: first column, then smallest row
                plants = [5] * int(rng.integers(2, 10))
            elif kind == 2:  # events that break chains, at arbitrary steps
                plants = [int(rng.integers(7, 13)) for _ in range(int(rng.integers(1, 30)))]
            else:            # a mix
                plants = [int(rng.integers(1, 13)) for _ in range(int(rng.integers(1, 40)))]
            s, r = _low_score_pair(rng, q, w, plants)
            seqs.append(s)
            refs.append(r)
    bad, undef = _compare(aligner, oracle, seqs, refs)
    assert bad == [], bad[:5]
    assert undef == 0


# ---- packed host inputs (cudasw_align_batch_packed: 2 bits per base + N mask, north_star (1))
def _same(res1, cig1, res2, cig2):
    for f in FIELDS:
        assert np.array_equal(res1[f], res2[f]), f
    assert all(_cig(res1, cig1, i) == _cig(res2, cig2, i) for i in range(len(res1)))


@pytest.mark.parametrize("cfg,n", [(1, 6000), (2, 3000), (3, 800), (4, 5000)])
def test_packed_inputs_match_ascii_and_oracle(aligner, oracle, synth, cfg, n):
    import gridss_b200 as gb

    sb, so, rb, ro = synth.generate(cfg, n)
    if cfg == 1:  # lower case and IUPAC codes must pack to what clean() makes of them
        rb = rb.copy()
        rb[::97] = ord("n")
        rb[5::211] = ord("R")
        sb = sb.copy()
        sb[3::301] = ord("y")
    bad, undef = _compare_blobs(aligner, oracle, sb, so, rb, ro)
    assert bad == [] and undef == 0, bad[:5]
    res1, cig1 = aligner.align_blobs(sb, so, rb, ro)
    s2, sn, samb = gb.pack_bases(sb, n=int(so[-1]))
    r2, rn, ramb = gb.pack_bases(rb, n=int(ro[-1]))
    res2, cig2 = aligner.align_blobs_packed(s2, sn, so, r2, rn, ro)
    _same(res1, cig1, res2, cig2)
    st = aligner.stats()
    assert st["h2d_bytes"] < 0.4 * (int(so[-1]) + int(ro[-1])) + 16 * (n + 1) + 64
    if samb == 0 and ramb == 0:  # a NULL mask = "no ambiguous base in this blob"
        res3, cig3 = aligner.align_blobs_packed(s2, None, so, r2, None, ro)
        _same(res1, cig1, res3, cig3)


def test_packed_inputs_chunked_and_sliced(aligner, synth, monkeypatch):
    """many small chunks (chunk borders inside a 2-bit word) and a batch that starts deep inside a packed blob"""
    import gridss_b200 as gb

    sb, so, rb, ro = synth.generate(4, 5000)
    rb = rb.copy()
    rb[7::1013] = ord("N")
    s2, sn, _ = gb.pack_bases(sb, n=int(so[-1]))
    r2, rn, _ = gb.pack_bases(rb, n=int(ro[-1]))
    res1, cig1 = aligner.align_blobs(sb, so, rb, ro)
    monkeypatch.setenv("CUDASW_CHUNK_MB", "1")
    res2, cig2 = aligner.align_blobs_packed(s2, sn, so, r2, rn, ro)
    monkeypatch.delenv("CUDASW_CHUNK_MB")
    _same(res1, cig1, res2, cig2)
    for first, cnt in ((1, 1), (1237, 2000), (4999, 1), (3, 4997)):
        so_v = np.ascontiguousarray(so[first: first + cnt + 1])
        ro_v = np.ascontiguousarray(ro[first: first + cnt + 1])
        res3, cig3 = aligner.align_blobs_packed(s2, sn, so_v, r2, rn, ro_v)
        _same(res1[first: first + cnt], cig1, res3, cig3)


def test_packed_inputs_multi_handle(aligner, synth):
    import gridss_b200 as gb

    ma = gb.MultiGpuAligner(devices=_visible_devices())
    sb, so, rb, ro = synth.generate(4, 6000)
    s2, sn, _ = gb.pack_bases(sb, n=int(so[-1]))
    r2, rn, _ = gb.pack_bases(rb, n=int(ro[-1]))
    res1, cig1 = aligner.align_blobs(sb, so, rb, ro)
    resm, cigm = ma.align_blobs_packed(s2, sn, so, r2, rn, ro)
    _same(res1, cig1, resm, cigm)
    ma.close()


# ---- JAligner-compatible tie-break mode (cudasw_create_ex(..., CUDASW_TIEBREAK_JALIGNER)). UNPINNED: the
# checker is oracle/jaligner_model.py, a restatement of jaligner 1.0's published algorithm (class files only, no JVM).
def _ja_expected(seqs, refs, scoring=(1, -4, -4, 6, 1)):
    from oracle import jaligner_model as J

    return [J.align(s, r, *scoring) for s, r in zip(seqs, refs)]


def _ja_check(al, seqs, refs, scoring=(1, -4, -4, 6, 1)):
    import gridss_b200 as gb
    from gridss_b200.aligner import pack

    sb, so = pack(seqs)
    rb, ro = pack(refs)
    res, cig = al.align_blobs(sb, so, rb, ro)
    want = _ja_expected(seqs, refs, scoring)
    bad = []
    for i, (sc, st, cg, ref_end, read_end) in enumerate(want):
        r = res[i]
        got = (int(r["status"]), int(r["score"]), int(r["ref_begin"]), gb.cigar_string(cig[r["cigar_off"]: r["cigar_off"] + r["cigar_len"]]))
        exp = (0, sc, st, cg)
        if got != exp or (sc > 0 and (int(r["ref_end"]), int(r["read_end"])) != (ref_end, read_end)):
            bad.append((i, got, exp, seqs[i][:40], refs[i][:40]))
    return bad


def test_jaligner_mode_contract_vectors():
    import gridss_b200 as gb

    al = gb.CudaAligner(tie_break="jaligner")
    # SmithWatermanAlignerTest.java:13-35, run by the reference against every backend (JAlignerAlignerTest)
    for seq, ref, start, cigar in ((b"AAACCCCCCCCCCCCCTTAATTAATTAATTTT", b"AAACCCCCCCCCCCCCGTTAATTAATTAATTTT", 0, "16M1D16M"),
                                   (b"AACCCTTTTTT", b"AAACCCTTTTTT", 1, "11M"), (b"GGGGGGTTTTTT", b"AAACCCTTTTTT", 6, "6S6M")):
        a = al.align_smith_waterman(seq, ref)
        assert (a.getStartPosition(), a.getCigar()) == (start, cigar)
    a = al.align_smith_waterman(b"AAAA", b"CCCCCC")  # no positive cell: alignmentToCigar still appends its (empty) last run
    assert (a.getStartPosition(), a.getCigar(), a.score) == (0, "0M4S", 0)
    al.close()


@pytest.mark.parametrize("kind", ["random", "mutated", "two_letter", "homopolymer", "dinuc", "tandem", "subst_block", "zero",
                                  "tiny", "q_gt_w", "indel_heavy"])
def test_jaligner_mode_vs_restatement(synth, kind):
    import gridss_b200 as gb

    al = gb.CudaAligner(tie_break="jaligner")
    for seed, qmax, wmax, n in ((31, 60, 140, 250), (32, 260, 420, 40)):
        seqs, refs = synth.adversarial(kind, n, seed=seed, qmax=qmax, wmax=wmax)
        bad = _ja_check(al, seqs, refs)
        assert bad == [], bad[:3]
    al.close()


def test_jaligner_mode_other_scoring_and_n(synth):
    """gap extension free / cheaper to open than to extend (the open-beats-extend tie rule), and N in both sequences"""
    import gridss_b200 as gb

    rng = np.random.default_rng(5)
    for sc in ((2, -6, -1, 5, 3), (3, -2, -5, 4, 0), (1, -1, -1, 2, 1), (1, -4, -4, 1, 3)):
        al = gb.CudaAligner(*sc, tie_break="jaligner")
        seqs, refs = [], []
        for kind in ("mutated", "indel_heavy", "tandem"):
            s, r = synth.adversarial(kind, 60, seed=41, qmax=70, wmax=150)
            seqs += s
            refs += r
        for _ in range(60):  # ACGTN only, either case: where clean() and jaligner's raw-byte matrix agree
            r = rng.choice(list(b"ACGTNacgtn"), size=int(rng.integers(20, 120))).astype(np.uint8).tobytes()
            st = int(rng.integers(0, len(r) // 2))
            seqs.append(r[st: st + int(rng.integers(5, 60))].swapcase())
            refs.append(r)
        bad = _ja_check(al, seqs, refs, sc)
        assert bad == [], (sc, bad[:3])
        al.close()


def test_jaligner_mode_shares_score_and_end_with_ssw_mode(aligner, synth):
    import gridss_b200 as gb

    al = gb.CudaAligner(tie_break="jaligner")
    sb, so, rb, ro = synth.generate(4, 3000)
    a, _ = aligner.align_blobs(sb, so, rb, ro)
    b, _ = al.align_blobs(sb, so, rb, ro)
    assert (b["status"] == 0).all()
    for f in ("score", "ref_end", "read_end"):
        assert np.array_equal(a[f], b[f]), f
    # SSW reports the start column nearest the end among all full-score alignments; jaligner's walk ends in one of them
    ok = a["score"] > 0
    assert (b["ref_begin"][ok] <= a["ref_begin"][ok]).all()
    assert (b["ref_begin"] == a["ref_begin"]).mean() > 0.9
    al.close()


def test_strip_mined_queries_across_many_chunks(aligner, oracle, monkeypatch):
    """queries above the tallest single-strip variant (strip-mined forward pass, boundary rows in a bump-allocated
    pool) spread over MANY host chunks with w ~ q, so that the reverse pass (which runs once over the whole batch) needs more
    boundary memory than any single chunk's forward pass did; both reverse-pass implementations."""
    rng = np.random.default_rng(77)
    seqs, refs = [], []
    for _ in range(160):
        w = int(rng.integers(2150, 3100))
        r = rng.choice(list(b"ACGT"), size=w).astype(np.uint8)
        s = r[int(rng.integers(0, 40)): w - int(rng.integers(0, 40))].copy()
        for k in rng.integers(0, len(s), size=len(s) // 40):
            s[k] = rng.choice(list(b"ACGT"))
        seqs.append(s.tobytes())
        refs.append(r.tobytes())
    monkeypatch.setenv("CUDASW_CHUNK_MB", "1")
    bad, undef = _compare(aligner, oracle, seqs, refs)
    assert bad == [] and undef == 0, bad[:3]
    monkeypatch.setenv("CUDASW_REV_FULL", "1")  # round 1's full-rectangle reverse kernels (strip pool sized for the whole batch)
    bad, undef = _compare(aligner, oracle, seqs, refs)
    assert bad == [] and undef == 0, bad[:3]


# ---- the k-mer filter in front of the forward pass (csrc/band_filter.cuh): identical results with and without it
def _digest_on_off(sb, so, rb, ro, monkeypatch, scoring=(1, -4, -4, 6, 1)):
    import gridss_b200 as gb
    from gridss_b200.shard import pair_digests

    monkeypatch.setenv("CUDASW_FILTER", "0")
    off = gb.AlignerFactory.create(*scoring)
    monkeypatch.setenv("CUDASW_FILTER", "2")  # always, also for batches that fill the machine less than twice
    on = gb.AlignerFactory.create(*scoring)
    r0, c0 = off.align_blobs(sb, so, rb, ro)
    r1, c1 = on.align_blobs(sb, so, rb, ro)
    d0, d1 = pair_digests(r0, c0), pair_digests(r1, c1)
    off.close()
    on.close()
    return np.flatnonzero(d0 != d1), r0, r1


@pytest.mark.parametrize("cfg,n", [(1, 60000), (2, 40000), (3, 6000), (4, 50000)])
def test_filter_on_off_identical_on_configs(synth, monkeypatch, cfg, n):
    sb, so, rb, ro = synth.generate(cfg, n)
    bad, r0, r1 = _digest_on_off(sb, so, rb, ro, monkeypatch)
    assert bad.size == 0, (bad[:5], r0[bad[:3]], r1[bad[:3]])


def _repeat_cases(rng, n):
    """Pairs built to mislead a seed filter: the query occurs several times in the window (exactly, or with a few edits, or with a
    gap), copies at equal score in different columns (tie-break: the FIRST column wins), tandem repeats, a strong decoy far away
    from a slightly better true hit, queries that are half noise."""
    A = np.frombuffer(b"ACGT", dtype=np.uint8)
    seqs, refs = [], []
    for t in range(n):
        q = int(rng.integers(40, 600))
        w = int(rng.integers(2 * q + 400, 7000))
        ref = rng.choice(A, size=w)
        core = rng.choice(A, size=q)
        kind = t % 6
        def put(at, s):
            at = max(0, min(w - len(s), at))
            ref[at: at + len(s)] = s
        def edit(s, k):
            s = s.copy()
            for x in rng.integers(0, len(s), size=k):
                s[x] = rng.choice(A)
            return s
        if kind == 0:    # two exact copies
            put(int(rng.integers(0, w // 2 - q)) if w // 2 > q else 0, core); put(int(rng.integers(w // 2, w - q)), core)
        elif kind == 1:  # a decoy with a few more edits, placed FIRST
            put(int(rng.integers(0, w // 2 - q)) if w // 2 > q else 0, edit(core, 3)); put(int(rng.integers(w // 2, w - q)), edit(core, 2))
        elif kind == 2:  # tandem repeat unit in both
            unit = rng.choice(A, size=int(rng.integers(2, 30)))
            core = np.resize(unit, q)
            put(int(rng.integers(0, w - 3 * q)) if w > 3 * q else 0, np.resize(unit, min(3 * q, w)))
        elif kind == 3:  # copy with a long gap in the window, and an ungapped shorter copy elsewhere
            g = int(rng.integers(5, 120))
            h = q // 2
            s = np.concatenate([core[:h], rng.choice(A, size=g), core[h:]])
            put(int(rng.integers(0, max(1, w // 2 - len(s)))), s); put(int(rng.integers(w // 2, w - q)), core[: int(q * 0.8)])
        elif kind == 4:  # half of the query is noise; two copies of the good half
            core[q // 2:] = rng.choice(A, size=q - q // 2)
            hq = core[: q // 2]
            put(int(rng.integers(0, w - q)), hq); put(int(rng.integers(0, w - q)), edit(hq, 1))
        else:            # many weak copies
            for _ in range(6):
                put(int(rng.integers(0, w - q)), edit(core, int(rng.integers(0, max(2, q // 10)))))
        seqs.append(core.tobytes())
        refs.append(ref.tobytes())
    return seqs, refs


def test_filter_repeats_decoys_and_ties_vs_oracle(aligner, oracle, monkeypatch):
    from gridss_b200.aligner import pack

    import gridss_b200 as gb

    rng = np.random.default_rng(20260)
    seqs, refs = _repeat_cases(rng, 1800)
    monkeypatch.setenv("CUDASW_FILTER", "2")
    forced = gb.CudaAligner(1, -4, -4, 6, 1)  # (AlignerFactory.create() without arguments is the shared default instance)
    bad, undef = _compare(forced, oracle, seqs, refs)
    forced.close()
    assert bad == [], bad[:3]
    sb, so = pack(seqs)
    rb, ro = pack(refs)
    bad2, r0, r1 = _digest_on_off(sb, so, rb, ro, monkeypatch)
    assert bad2.size == 0, (bad2[:5], r0[bad2[:3]], r1[bad2[:3]])


def test_filter_other_scorings_on_off(synth, monkeypatch):
    sb, so, rb, ro = synth.generate(3, 3000)
    for sc in ((2, -3, -1, 5, 2), (1, -1, -1, 2, 1), (3, -2, -2, 4, 0), (1, -4, -4, 0, 1), (5, -4, -4, 10, 1)):
        bad, r0, r1 = _digest_on_off(sb, so, rb, ro, monkeypatch, sc)
        assert bad.size == 0, (sc, bad[:5], r0[bad[:3]], r1[bad[:3]])


@pytest.mark.parametrize("cfg,n", [(2, 30000), (3, 5000)])
def test_filter_band_mode_identical(synth, monkeypatch, cfg, n):
    """CUDASW_BAND=1: certified pairs sweep their diagonal band with the strip kernels run forwards (not the default)."""
    monkeypatch.setenv("CUDASW_BAND", "1")
    sb, so, rb, ro = synth.generate(cfg, n)
    bad, r0, r1 = _digest_on_off(sb, so, rb, ro, monkeypatch)
    assert bad.size == 0, (bad[:5], r0[bad[:3]], r1[bad[:3]])
