"""CPU: the JAligner restatement (oracle/jaligner_model.py; SURVEY.md §8a A6, §8f N4 — groundwork, parity unpinned).
Anchors: the three contract vectors the reference runs against every Aligner backend
(SmithWatermanAlignerTest.java:13-35 via JAlignerAlignerTest.java), and the invariant that score and end cell equal the
SSW backend's (same recurrences, same first-strictly-greater scan order); start/CIGAR may differ on ties only."""
import numpy as np

from oracle import jaligner_model as J


def test_contract_vectors_shared_by_every_backend():
    # SmithWatermanAlignerTest.java:13-20, 21-26, 27-35
    s, st, cg, _, _ = J.align(b"AAACCCCCCCCCCCCCTTAATTAATTAATTTT", b"AAACCCCCCCCCCCCCGTTAATTAATTAATTTT")
    assert (st, cg) == (0, "16M1D16M")
    s, st, cg, _, _ = J.align(b"AACCCTTTTTT", b"AAACCCTTTTTT")
    assert (st, cg) == (1, "11M")
    s, st, cg, _, _ = J.align(b"GGGGGGTTTTTT", b"AAACCCTTTTTT")
    assert (st, cg) == (6, "6S6M")


def test_case_and_ambiguity_follow_the_matrix():
    # JAlignerAligner.createMatrix (JAlignerAligner.java:38-54): no clean(); case-insensitive equality is a match
    assert J.align(b"acgtacgt", b"TTACGTACGTTT")[1:3] == (2, "8M")
    assert J.align(b"ACGTNNACGT", b"ACGTNNACGT")[0] == 10


def test_score_and_end_cell_equal_the_ssw_backend(oracle, synth):
    rng = np.random.default_rng(77)
    seqs, refs = [], []
    for kind in ("mutated", "random", "indel_heavy", "tandem", "two_letter"):
        s, r = synth.adversarial(kind, 60, seed=5, qmax=60, wmax=120)
        seqs += s
        refs += r
    m = oracle.model_align(seqs, refs, threads=4)
    same_path = 0
    n = 0
    for i, (s, r) in enumerate(zip(seqs, refs)):
        if any(c not in b"ACGTacgt" for c in s + r):
            continue  # SswJniAligner.clean() rewrites other symbols; JAligner does not
        sc, st, cg, ref_end, read_end = J.align(s, r)
        ssw = m.res[i]
        if sc == 0:
            assert int(ssw["score"]) == 0
            continue
        n += 1
        assert (sc, ref_end, read_end) == (int(ssw["score"]), int(ssw["ref_end"]), int(ssw["read_end"])), (s, r)
        same_path += (st, cg) == (int(ssw["ref_begin"]), m.cigar(i))
    assert n > 150
    # the two backends differ on tie-breaks only (longest zero-prefix start, pointer priority): most paths coincide
    assert same_path >= 0.8 * n, (same_path, n)
