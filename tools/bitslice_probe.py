"""CPU check of the bit-sliced quiet-column update sketched in DESIGN.md §8a (next round; nothing here is product code).

At the default scoring (match 1, mismatch -4, gapOpen 6) a quiet cell holds H in 0..6; 7 is kept as the saturating
"event" value (some H reached gapOpen + 1: the dense kernel has to take over there). Per cell
    H' = match ? min(Hdiag + 1, 7) : max(0, Hdiag - 4)
and on three bit-planes (b2 b1 b0, 32 rows per word) this is
    sat = b2 & b1 & b0
    o0 = M ? (~b0 | sat)            : (b2 & b0)
    o1 = M ? ((b1 ^ b0) | sat)      : (b2 & b1)
    o2 = M ? ((b2 ^ (b1 & b0)) | sat) : 0
(H - 4 is H & 3 when b2 is set, else <= 0), i.e. one LOP3 for sat and two per plane, plus the diagonal shift of the
planes by one row (three funnel shifts with the carry of the word above), the match mask of the column's base (one
select among four precomputed query masks) and the event accumulation (o2 & o1 & o0 OR-ed into a flag): about 15
logic instructions per 32 cells. The script checks the circuit against the integer recurrence on random data.
"""
import sys

import numpy as np


def step_int(Hd, match):
    return np.where(match, np.minimum(Hd + 1, 7), np.maximum(0, Hd - 4))


def step_planes(b0, b1, b2, M):
    sat = b2 & b1 & b0
    o0 = (M & (~b0 | sat)) | (~M & (b2 & b0))
    o1 = (M & ((b1 ^ b0) | sat)) | (~M & (b2 & b1))
    o2 = M & ((b2 ^ (b1 & b0)) | sat)
    return o0, o1, o2


def shift_down(p):
    """diagonal move: row i of the new column reads row i-1 of the old one; bit r of word k is row 32k + r"""
    carry = np.concatenate(([np.uint32(0)], p[:-1] >> np.uint32(31)))
    return ((p << np.uint32(1)) | carry).astype(np.uint32)


def main():
    rng = np.random.default_rng(3)
    rows, cols = 320, 4000
    q = rng.integers(0, 4, size=rows)
    w = rng.integers(0, 4, size=cols)
    w[1000:1040] = q[100:140]  # an exact 40-mer: saturates
    words = rows // 32
    eq = np.zeros((4, words), dtype=np.uint32)  # per-base match masks of the query
    for i in range(rows):
        eq[q[i], i // 32] |= np.uint32(1) << np.uint32(i % 32)
    H = np.zeros(rows, dtype=np.int64)
    b0 = np.zeros(words, dtype=np.uint32)
    b1 = b0.copy()
    b2 = b0.copy()
    events = 0
    for j in range(cols):
        Hd = np.concatenate(([0], H[:-1]))
        H = step_int(Hd, q == w[j])
        b0, b1, b2 = step_planes(shift_down(b0), shift_down(b1), shift_down(b2), eq[w[j]])
        got = np.zeros(rows, dtype=np.int64)
        for i in range(rows):
            k, r = divmod(i, 32)
            got[i] = ((int(b0[k]) >> r) & 1) | (((int(b1[k]) >> r) & 1) << 1) | (((int(b2[k]) >> r) & 1) << 2)
        if not np.array_equal(got, H):
            print("MISMATCH at column", j)
            return 1
        events += int((H == 7).any())
    print(f"bit-sliced update == integer recurrence on {rows} x {cols} cells; columns holding an event value: {events}")
    return 0


if __name__ == "__main__":
    sys.exit(main())
