"""CPU analysis for the next round: how much of C2's forward matrix is "quiet" at different
granularities. A column of a pair is quiet when no H in it exceeds gapOpen and no E or F is alive, i.e. the column is
exactly the gap-free recurrence H' = max(0, Hdiag + s). Reports the quiet share per pair-column (what a per-pair
sparse/bit-sliced pre-pass could skip), per 19-row block of a column (the unit a virtual lane owns) and per warp-step
of the current kernel (4 pairs x 16 virtual lanes, skewed by one column per lane, anchors binned together or not).

  python tools/quiet_column_probe.py [pairs]
"""
import sys

import numpy as np

sys.path.insert(0, ".")
from tools import synth as S

GO, GE, MATCH, MIS = 6, 1, 1, -4


def column_flags(q, w):
    """Column-by-column Gotoh (F by prefix maximum, valid for go >= ge). Returns live[rows, cols]: cell has H > go or a
    live E / F (E, F > 0)."""
    nq, nw = len(q), len(w)
    H = np.zeros(nq + 1, dtype=np.int32)
    E = np.full(nq + 1, -10**6, dtype=np.int32)
    ramp = np.arange(nq + 1, dtype=np.int32) * GE
    live = np.zeros((nq, nw), dtype=bool)
    for j in range(nw):
        s = np.where(q == w[j], MATCH, MIS).astype(np.int32)
        E = np.maximum(E - GE, H - GO)                 # gap along the window: from column j-1 of the same row
        Ht = np.maximum(0, np.maximum(H[:-1] + s, E[1:]))  # rows 1..nq without F
        # F(i) = max_{k<i} Ht(k) - go - (i-k-1) ge
        t = np.maximum.accumulate(Ht + ramp[1:])       # max_k<=i (Ht(k) + k ge), 1-based rows
        F = np.empty(nq, dtype=np.int32)
        F[0] = -10**6
        F[1:] = t[:-1] - GO + GE - ramp[2:] + 0        # row i (1-based i+1): t[i-1] - go - (i - k - 1) ge with k ge folded in
        Hn = np.maximum(Ht, F)
        live[:, j] = (Hn > GO) | (E[1:] > 0) | (F > 0)
        H = np.concatenate(([0], Hn)).astype(np.int32)
    return live


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 64
    S.build()
    sb, so, rb, ro = S.generate(2, n)
    R, G = 19, 8
    col_quiet, blk_quiet = [], []
    lane_live = []  # per pair: [16 virtual lanes, cols] any live cell in the lane's rows
    for p in range(n):
        q = np.frombuffer(sb[so[p]:so[p + 1]].tobytes().upper(), dtype=np.uint8)
        w = np.frombuffer(rb[ro[p]:ro[p + 1]].tobytes().upper(), dtype=np.uint8)
        live = column_flags(q, w)
        col_quiet.append(1.0 - live.any(axis=0).mean())
        pad = np.zeros((2 * G * R, live.shape[1]), dtype=bool)
        pad[:live.shape[0]] = live
        ll = pad.reshape(2 * G, R, -1).any(axis=1)
        blk_quiet.append(1.0 - ll.mean())
        lane_live.append(ll)
    print(f"C2, {n} pairs (300 x 4000, default scoring)")
    print(f"quiet pair-columns (no live cell in the whole column of one pair): {np.mean(col_quiet):.3f}")
    print(f"quiet 19-row blocks (one virtual lane's share of a column):          {np.mean(blk_quiet):.3f}")
    # warp-steps of the wavefront kernel: lane v of a pair is at column t - v; a step is quiet when all 4 x 16 lanes are
    def warp_quiet(order):
        out = []
        for a in range(0, len(order) - 3, 4):
            W = lane_live[0].shape[1]
            steps = np.zeros(W + 2 * G, dtype=bool)
            for p in order[a:a + 4]:
                ll = lane_live[p]
                for v in range(2 * G):
                    steps[v:v + W] |= ll[v]
            # a live step also spoils the next one (E/F of the step before are part of the pre-check): already in `live`
            out.append(1.0 - steps.mean())
        return float(np.mean(out))
    first_live = [int(np.argmax(ll.any(axis=0))) for ll in lane_live]
    # anchor column = where the long live stretch starts: use the column of the maximum live density
    dens = [np.convolve(ll.any(axis=0).astype(np.int32), np.ones(64, dtype=np.int32), "same").argmax() for ll in lane_live]
    print(f"quiet warp-steps, 4 pairs per warp in input order:               {warp_quiet(list(range(n))):.3f}")
    print(f"quiet warp-steps, pairs binned by anchor column:                 {warp_quiet(list(np.argsort(dens))):.3f}")


if __name__ == "__main__":
    main()
