"""CPU: pins the oracle restatement (oracle/ssw_model.c) against the reference's golden vectors and,
when oracle/_ref/libsswjni.so is staged, against the reference's shipped binary itself."""
import json
import os

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
FIELDS = ["score", "ref_begin", "ref_end", "read_begin", "read_end", "cigar_len"]


def _load(name):
    return json.load(open(os.path.join(HERE, "golden", name)))


def test_model_reproduces_reference_test_vectors(oracle):
    g = _load("golden_reference_tests.json")
    m = oracle.model_align([x["seq"].encode("latin1") for x in g], [x["ref"].encode("latin1") for x in g])
    for i, x in enumerate(g):
        assert m.tuple(i) == (x["score"], x["start"], x["cigar"]), x["cite"]


def test_model_matches_sswjni_golden_fuzz(oracle):
    g = _load("golden_sswjni_fuzz.json")
    by_sc = {}
    for x in g:
        by_sc.setdefault(tuple(x["scoring"]), []).append(x)
    for sc, xs in by_sc.items():
        m = oracle.model_align([x["seq"].encode("latin1") for x in xs], [x["ref"].encode("latin1") for x in xs], sc, threads=4)
        for i, x in enumerate(xs):
            assert m.res[i]["status"] == 0
            assert m.tuple(i) == (x["score"], x["start"], x["cigar"]), (x["kind"], sc, i)


def _same(r, m):
    bad = []
    for i in range(len(r.res)):
        if m.res[i]["status"] != 0 or tuple(r.res[i][f] for f in FIELDS) != tuple(m.res[i][f] for f in FIELDS) or r.cigar(i) != m.cigar(i):
            bad.append(i)
    return bad


@pytest.mark.parametrize("kind", ["random", "mutated", "two_letter", "homopolymer", "dinuc", "tandem", "subst_block",
                                  "with_n", "zero", "tiny", "q_gt_w", "indel_heavy"])
def test_model_vs_reference_binary_adversarial(oracle, synth, kind):
    if not oracle.have_reference():
        pytest.skip("oracle/_ref/libsswjni.so not staged on this machine")
    seqs, refs = synth.adversarial(kind, 600, seed=2024)
    r = oracle.ref_align(seqs, refs, threads=4)
    m = oracle.model_align(seqs, refs, threads=4)
    assert _same(r, m) == []


@pytest.mark.parametrize("cfg,n", [(1, 1500), (2, 500), (3, 150), (4, 300), (5, 1)])
def test_model_vs_reference_binary_synthetic_configs(oracle, synth, cfg, n):
    if not oracle.have_reference():
        pytest.skip("oracle/_ref/libsswjni.so not staged on this machine")
    sb, so, rb, ro = synth.generate(cfg, n)
    r = oracle.ref_align_blobs(sb, so, rb, ro, threads=8)
    m = oracle.model_align_blobs(sb, so, rb, ro, threads=8)
    assert _same(r, m) == []


@pytest.mark.parametrize("sc", [(2, -6, -1, 5, 3), (1, -1, -1, 2, 1), (3, -2, -5, 4, 0), (5, -4, -2, 10, 1)])
def test_model_vs_reference_binary_other_scoring(oracle, synth, sc):
    if not oracle.have_reference():
        pytest.skip("oracle/_ref/libsswjni.so not staged on this machine")
    for kind in ("mutated", "with_n", "indel_heavy"):
        seqs, refs = synth.adversarial(kind, 300, seed=5)
        assert _same(oracle.ref_align(seqs, refs, sc, threads=4), oracle.model_align(seqs, refs, sc, threads=4)) == []


def test_synth_is_counter_based(synth):
    a = synth.generate(3, 64, first=100)
    b = synth.generate(3, 200, first=0)
    so, ro = b[1], b[3]
    assert bytes(a[0]) == bytes(b[0][so[100]:so[164]])
    assert bytes(a[2]) == bytes(b[2][ro[100]:ro[164]])
    assert np.array_equal(np.diff(a[1]), np.diff(so[100:165]))
