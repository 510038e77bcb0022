"""CPU: the certificate arithmetic of the k-mer filter (gridss_b200/csrc/band_filter.cuh) - the very functions the kernels
call, compiled for the host by tests/native/filter_math.cu - against brute force, against the Python model of
tools/filter_probe.py and against the optimal alignments the oracle reports. The GPU side (results identical with the
filter on and off) is tests/test_gpu_parity.py::test_filter_*."""
import ctypes
import os
import re
import subprocess

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
SCORINGS = [(1, -4, -4, 6, 1), (2, -3, -1, 5, 2), (1, -1, -1, 2, 1), (5, -4, -4, 10, 1), (1, -4, -4, 6, 3), (1, -6, -2, 7, 1), (3, -9, -9, 16, 2)]


@pytest.fixture(scope="module")
def fm():
    src = os.path.join(HERE, "native", "filter_math.cu")
    out_dir = os.path.join(HERE, "native", "_build")
    os.makedirs(out_dir, exist_ok=True)
    so = os.path.join(out_dir, "libfm.so")
    deps = [src, os.path.join(HERE, "..", "gridss_b200", "csrc", "band_filter.cuh"), os.path.join(HERE, "..", "gridss_b200", "csrc", "sw_kernels.cuh")]
    if not os.path.exists(so) or any(os.path.getmtime(d) > os.path.getmtime(so) for d in deps):
        subprocess.check_call([NVCC, "-O2", "-std=c++17", "-w", "--expt-relaxed-constexpr", "-Xcompiler", "-fPIC", "-shared", "-o", so, src])
    lib = ctypes.CDLL(so)
    i, ll = ctypes.c_int, ctypes.c_longlong
    lib.fm_hits_floor.restype = ll
    lib.fm_hits_floor.argtypes = [i, i, i, i, i, ll, ll]
    lib.fm_max_gap_bases.restype = ll
    lib.fm_max_gap_bases.argtypes = [ll, i, i, i, i]
    lib.fm_bins_of.argtypes = [ll]
    lib.fm_tight_bins.argtypes = [i, i, i, i, i, ll, ll, i, ll, ctypes.c_void_p, ctypes.POINTER(ll)]
    return lib


def test_kmer_length_matches_the_model(fm):
    from tools import filter_probe as F

    assert fm.fm_k() == F.K == 6


def test_hits_floor_never_exceeds_the_brute_force_minimum(fm):
    """T(S0) must be a LOWER bound of M - (k-1)(X + g + 1) over every (M, X, g) with a M - p X - go g >= S0 and M + X <= n."""
    k1 = fm.fm_k() - 1
    checked = 0
    for sc in SCORINGS:
        a, mm, amb, go, ge = sc
        if not fm.fm_ok(*sc):
            continue
        p = min(-mm, -amb)
        for n in (7, 12, 23, 40):
            for S0 in range(1, a * n + 1, max(1, a * n // 9)):
                best = None
                for M in range(n + 1):
                    for X in range(n - M + 1):
                        g_hi = (a * M - p * X - S0) // go if go > 0 else 0
                        if a * M - p * X < S0:
                            continue
                        # more gaps only lower the count: the most that the score allows
                        v = M - k1 * (X + max(g_hi, 0) + 1)
                        best = v if best is None else min(best, v)
                if best is None:
                    continue
                t = fm.fm_hits_floor(*sc, S0, n)
                assert t <= best, (sc, n, S0, t, best)
                assert t >= best - 2 - k1, (sc, n, S0, t, best)  # and it is not vacuous (LP relaxation + floor)
                checked += 1
    assert checked > 100


def test_native_functions_equal_the_python_model(fm):
    from tools import filter_probe as F

    rng = np.random.default_rng(5)
    for sc in SCORINGS:
        for _ in range(300):
            n = int(rng.integers(6, 3000))
            S0 = int(rng.integers(1, sc[0] * n + 1))
            if fm.fm_ok(*sc):
                assert fm.fm_hits_floor(*sc, S0, n) == F.hits_floor(S0, n, sc)
            assert fm.fm_max_gap_bases(n, sc[0], min(S0, 32767), sc[3], sc[4]) == F.max_gap_bases(n, sc[0], min(S0, 32767), sc[3], sc[4])
    for L in list(range(0, 200)) + [1000, 4095, 4096, 4097]:
        assert fm.fm_bins_of(L) == F.bins_of(L)
        # a window of L + 1 consecutive diagonals never touches more bins than that, and some window touches as many
        assert max((d + L) // F.BIN - d // F.BIN + 1 for d in range(2 * F.BIN)) == F.bins_of(L)


def _path_stats(seq, ref, r, cigar):
    """Walks the oracle's alignment: (score parts) matches, other diagonal steps, gaps, gap bases, 6-mer hits on the path."""
    code = {65: 0, 67: 1, 71: 2, 84: 3}
    i, j = int(r["read_begin"]), int(r["ref_begin"])
    M = X = g = Lg = 0
    runs = []
    run = 0
    for ln, op in re.findall(r"(\d+)([MIDS])", cigar):
        ln = int(ln)
        if op == "S":
            continue
        if op == "M":
            for _ in range(ln):
                a, b = code.get(seq[i] & 0xDF, 4), code.get(ref[j] & 0xDF, 4)
                if a == b and a < 4:
                    M += 1
                    run += 1
                else:
                    X += 1
                    runs.append(run)
                    run = 0
                i += 1
                j += 1
        else:
            g += 1
            Lg += ln
            runs.append(run)
            run = 0
            if op == "I":
                i += ln
            else:
                j += ln
    runs.append(run)
    return M, X, g, Lg, sum(max(0, x - 5) for x in runs)


def test_bounds_hold_on_the_oracle_s_optimal_alignments(fm, oracle, synth):
    """hits >= T(score), gap bases <= L(score), and the evidence bound with the path's own hits, on the optimal paths of
    C1-C3 samples and adversarial families (windows without N, as the filter requires)."""
    sc = (1, -4, -4, 6, 1)
    from gridss_b200.aligner import pack

    batches = [synth.generate(c, n) for c, n in ((1, 400), (2, 400), (3, 150))]
    for kind in ("mutated", "indel_heavy", "tandem", "subst_block"):
        s, r = synth.adversarial(kind, 150, seed=3)
        batches.append(pack(s) + pack(r))
    checked = 0
    for sb, so, rb, ro in batches:
        m = oracle.model_align_blobs(sb, so, rb, ro, sc, threads=8)
        for p in range(len(so) - 1):
            r = m.res[p]
            if r["status"] != 0 or r["score"] <= 0:
                continue
            seq, ref = sb[so[p]: so[p + 1]], rb[ro[p]: ro[p + 1]]
            if any((x & 0xDF) not in (65, 67, 71, 84) for x in ref.tolist()):
                continue
            M, X, g, Lg, hits = _path_stats(seq, ref, r, m.cigar(p))
            S, n = int(r["score"]), min(len(seq), len(ref))
            assert hits >= fm.fm_hits_floor(*sc, S, n), (p, S, n, hits)
            L = fm.fm_max_gap_bases(n, sc[0], S, sc[3], sc[4])
            assert L < 0 or Lg <= L, (p, S, n, Lg, L)
            # evidence: with hc(j) >= the path's own hits for every j, the bound must still cover the path's gap bases
            hc = (ctypes.c_uint32 * 16)(*([hits] * 16))
            gb = ctypes.c_longlong(0)
            m0 = fm.fm_bins_of(L)
            mt = fm.fm_tight_bins(*sc, S, n, m0, L, hc, ctypes.byref(gb))
            assert Lg <= gb.value and fm.fm_bins_of(Lg) <= mt <= m0, (p, S, n, Lg, gb.value, mt, m0, hits)
            checked += 1
    assert checked > 1000


@pytest.fixture(scope="module")
def fb():
    src = os.path.join(HERE, "native", "filter_bruteforce.cu")
    out_dir = os.path.join(HERE, "native", "_build")
    os.makedirs(out_dir, exist_ok=True)
    so = os.path.join(out_dir, "libfb.so")
    deps = [src, os.path.join(HERE, "..", "gridss_b200", "csrc", "band_filter.cuh"), os.path.join(HERE, "..", "gridss_b200", "csrc", "sw_kernels.cuh")]
    if not os.path.exists(so) or any(os.path.getmtime(d) > os.path.getmtime(so) for d in deps):
        subprocess.check_call([NVCC, "-O2", "-std=c++17", "-w", "--expt-relaxed-constexpr", "-Xcompiler", "-fPIC", "-shared", "-o", so, src])
    lib = ctypes.CDLL(so)
    lib.fb_run.argtypes = [ctypes.c_uint64] + [ctypes.c_int] * 8 + [ctypes.POINTER(ctypes.c_longlong)]
    return lib


@pytest.mark.parametrize("sc", [(1, -4, -4, 6, 1), (2, -3, -1, 5, 2), (1, -1, -1, 2, 1), (5, -4, -4, 10, 1), (1, -6, -2, 7, 1), (1, -4, -4, 6, 3)])
def test_certified_range_equals_the_full_sweep_by_brute_force(fb, sc):
    """The theorem itself on small instances (tests/native/filter_bruteforce.cu): whenever the certificate holds for the score
    found inside the column range, (score, first column, smallest row) of the range equal those of the whole matrix; and a
    range never scores above the whole. Exact and mutated copies, two copies at equal score (ties), decoys placed first,
    tandem repeats, indel-heavy copies, two-letter noise, half-noise queries, N in the query."""
    out = (ctypes.c_longlong * 5)()
    tried = certified = 0
    for seed, qmax, wmax in ((1, 60, 300), (2, 140, 700), (3, 300, 1500)):
        assert fb.fb_run(seed, 1600, *sc, qmax, wmax, out) == 0
        assert out[2] == 0 and out[3] == 0, (sc, seed, list(out))
        tried += out[0]
        certified += out[1]
    assert tried > 1000
    if sc == (1, -4, -4, 6, 1):
        assert certified > 500, (tried, certified)  # the default scoring certifies most planted copies
