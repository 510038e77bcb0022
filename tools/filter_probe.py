"""CPU model of the lossless k-mer filter in front of the forward pass: how often it certifies a
diagonal band, how many cells the banded DP then visits, on the synthetic configs. Tooling only (uses the oracle for the
true score); the device kernels in csrc/band_filter.cuh follow the same arithmetic."""
from __future__ import annotations

import sys

import numpy as np

K = 6
BIN = 16
CORE = 2  # bins either side of the densest pair that belong to the core
EVIDENCE = True
REV = []


def max_gap_bases(rows, match, target, go, ge):
    budget = rows * match - target
    if budget < 0:
        return 0
    if go >= ge:
        if budget < go:
            return 0
        if ge == 0:
            return -1
        return (budget - go) // ge + 1
    if go == 0:
        return -1
    return budget // go


def hits_floor(S0, n, sc, k=K):
    """Lower bound (floor) on the exact k-mer hits of any path of score >= S0 inside n = min(q, w) diagonal steps."""
    a, mm, amb, go, ge = sc
    pmin = min(-mm, -amb)
    gmin = go
    if a <= 0:
        return -1
    cands = [(0, 1)]
    c1, A1 = pmin - (k - 1) * a, pmin + a
    cands.append((c1, A1))
    c2, A2 = gmin - (k - 1) * a, gmin
    if A2 == 0:
        if c2 < 0:
            return -1
    else:
        cands.append((c2, A2))
    cn, cd = min(cands, key=lambda x: x[0] / x[1])
    if cn > 0:
        cn, cd = 0, 1
    bud = a * n - S0
    if bud < 0:
        bud = 0
    num = (S0 - (k - 1) * a) * cd + bud * cn
    return num // (a * cd)  # python floor division


def bins_of(Lg):
    return (Lg + BIN - 1) // BIN + 1


def code(b):
    t = np.full(256, 4, np.uint8)
    for i, ch in enumerate(b"ACGT"):
        t[ch] = i
        t[ch | 0x20] = i
    return t[b]


def kmers(c, k=K):
    n = len(c) - k + 1
    if n <= 0:
        return np.zeros(0, np.int64), np.zeros(0, bool)
    v = np.zeros(n, np.int64)
    bad = np.zeros(n, bool)
    for x in range(k):
        v = v * 4 + (c[x : x + n] & 3)
        bad |= c[x : x + n] > 3
    return v, bad


def strip_cells(q, wd, w):
    best = None
    for S in (32, 64, 128, 256, 1024):
        cells = ((q + S - 1) // S) * S * min(S + wd, w)
        if best is None or cells < best:
            best = cells
    return best


def probe(cfg, n, sc=(1, -4, -4, 6, 1), est=0.8, est_t=1.0):
    from oracle import oracle as O
    from tools import synth as S

    sb, so, rb, ro = S.generate(cfg, n)
    m = O.model_align_blobs(sb, so, rb, ro, sc, threads=16)
    tot_cells = 0
    band_cells = 0
    n_try = n_cert = 0
    why = {}
    for p in range(n):
        qs = code(sb[so[p] : so[p + 1]])
        ws = code(rb[ro[p] : ro[p + 1]])
        q, w = len(qs), len(ws)
        tot_cells += q * w
        full = q * w

        def fail(r):
            nonlocal band_cells
            why[r] = why.get(r, 0) + 1
            band_cells += full

        if (ws > 3).any() or q < K or w < K or q > 2048 or q + w > 32768:
            fail("ineligible")
            continue
        kq, bq = kmers(qs)
        kw, bw = kmers(ws)
        nb = ((q + w - 2) // BIN) + 1
        cnt = np.zeros(nb + 2, np.int64)
        order = np.argsort(kq, kind="stable")
        ks = kq[order]
        lo = np.searchsorted(ks, kw, "left")
        hi = np.searchsorted(ks, kw, "right")
        for j in np.flatnonzero((hi > lo) & ~bw):
            for i in order[lo[j] : hi[j]]:
                if not bq[i]:
                    cnt[(j - i + q - 1) // BIN] += 1
        pair = cnt[:-1] + cnt[1:]
        bstar = int(np.argmax(pair[:nb]))
        hstar = int(pair[bstar])
        S_est = int(est * sc[0] * hstar)
        nmin = min(q, w)
        Lg = max_gap_bases(nmin, sc[0], S_est, sc[3], sc[4])
        if Lg < 0 or S_est <= 0:
            fail("unbounded")
            continue
        m_r = bins_of(Lg)
        # noise: max m_r-run-sum over runs entirely in bins <= bstar-2 or >= bstar+3
        cs = np.concatenate([[0], np.cumsum(cnt[:nb])])
        noise = 0
        for bs in range(0, nb):
            be = min(bs + m_r, nb)
            if be - 1 <= bstar - CORE - 1 or bs >= bstar + CORE + 2:
                noise = max(noise, int(cs[be] - cs[bs]))
        # evidence-based gap bound: a path through the core sees at most hc(j) hits when it spans j extra bins either side
        def hc(j):
            lo, hi = max(0, bstar - CORE - j), min(nb - 1, bstar + 1 + CORE + j)
            return int(cs[hi + 1] - cs[lo])

        def m_tight(S, m0):
            a, mm, amb, go, ge = sc
            kap = K - 1
            pmin = min(-mm, -amb)
            if go - a * kap <= 0 or ge <= 0:
                return m0, None
            xmax = (a * nmin - S) // (pmin + a)
            alpha = max(0, a * kap - pmin)
            m = m0
            Lt = None
            for _ in range(8):
                R = a * hc(m - 1) + a * kap - S + alpha * xmax
                Lt = max(0, R // min(go - a * kap, ge))
                m2 = min(m, bins_of(Lt))
                if m2 == m:
                    break
                m = m2
            return m, Lt

        m_old = m_r
        if EVIDENCE:
            m_r, _ = m_tight(S_est, m_old)
        Dlo, Dhi = BIN * (bstar - CORE - m_r + 1), BIN * (bstar + 1 + CORE + m_r) - 1
        Dlo, Dhi = max(Dlo, 0), min(Dhi, q + w - 2)
        wd = Dhi - Dlo + 1
        c0, c1 = max(0, Dlo - (q - 1)), min(w - 1, Dhi)
        cells = q * (c1 - c0 + 1)  # v1: the rectangle of columns the band's rows can reach (existing forward kernels on a sub-window)
        if hits_floor(int(est_t * sc[0] * hstar), nmin, sc) <= noise:
            fail("hopeless")
            continue
        if cells > 0.7 * full:
            fail("no gain")
            continue
        n_try += 1
        S0 = int(m.res[p]["score"])  # (the band DP's score; equals the true one whenever the certificate holds)
        Lg0 = max_gap_bases(nmin, sc[0], S0, sc[3], sc[4])
        if EVIDENCE:
            ok = S0 > 0 and Lg0 >= 0 and bins_of(Lg0) <= m_old and hits_floor(S0, nmin, sc) > noise
            if ok:
                mt, Lt = m_tight(S0, bins_of(Lg0))
                ok = mt <= m_r
                if ok:
                    REV.append((Lg0, min(Lg0, Lt) if Lt is not None else Lg0))
        else:
            ok = S0 > 0 and Lg0 >= 0 and bins_of(Lg0) <= m_r and hits_floor(S0, nmin, sc) > noise
        if ok:
            n_cert += 1
            band_cells += cells
        else:
            why["cert failed"] = why.get("cert failed", 0) + 1
            band_cells += cells + full
    if REV:
        r = np.array(REV)
        print(f"   forward gap bound: worst-case {r[:, 0].mean():.1f} -> evidence {r[:, 1].mean():.1f} (mean over certified pairs)")
        REV.clear()
    print(f"C{cfg}: pairs {n} tried {n_try} certified {n_cert}; cells visited {band_cells / tot_cells:.3f} of full; other: {why}")


if __name__ == "__main__":
    cfgs = [int(x) for x in sys.argv[1].split(",")] if len(sys.argv) > 1 else [1, 2, 3]
    n = int(sys.argv[2]) if len(sys.argv) > 2 else 200
    est = float(sys.argv[3]) if len(sys.argv) > 3 else 0.8
    est_t = float(sys.argv[4]) if len(sys.argv) > 4 else 1.0
    for c in cfgs:
        probe(c, n, est=est, est_t=est_t)
