"""TEST INFRASTRUCTURE ONLY (never imported by the product path): CPU restatement of the reference's second, slower
`Aligner` backend, `JAlignerAligner`.

PARITY UNPINNED: the arithmetic lives in a third-party dependency that /root/reference ships as class files only
(`jaligner:jaligner:1.0`, pom.xml:119-123; `repo/jaligner/jaligner/1.0/jaligner-1.0.jar`), and this image has no JVM to
run it. What follows restates the published algorithm of that library (JAligner 1.0, `SmithWatermanGotoh.align /
construct / traceback`) as SURVEY.md §8a A6 describes it, plus the reference's own wrapper
(`alignment/JAlignerAligner.java:30-37` call, `:38-54` matrix, `:59-93` CIGAR). It is anchored on the three contract
vectors the reference runs against every backend (`SmithWatermanAlignerTest.java:13-35`, subclassed by
`JAlignerAlignerTest.java`) and on the invariant that score and end cell equal the SSW backend's (both are exact Gotoh
with the same gap convention and the same first-strictly-greater scan order). Start position and CIGAR may differ
from SSW on ties: JAligner walks stored pointers (STOP / DIAGONAL / UP / LEFT priority) until H == 0.

Pure-Python loops: small cases only.
"""
from __future__ import annotations

STOP, LEFT, DIAGONAL, UP = 0, 1, 2, 3  # jaligner.Directions
MATRIX_SIZE = 127                      # jaligner.matrix.Matrix.SIZE
_VALID = set(b"ACGTacgt")              # htsjdk SequenceUtil.isValidBase


def _score(a: int, b: int, match: int, mismatch: int, ambiguous: int) -> int:
    # JAlignerAligner.createMatrix (JAlignerAligner.java:38-54): no clean(), raw bytes index the matrix
    if chr(a).upper() == chr(b).upper():
        return match
    if a in _VALID and b in _VALID:
        return mismatch
    return ambiguous


def align(seq: bytes, ref: bytes, match: int = 1, mismatch: int = -4, ambiguous: int = -4, gap_open: int = 6,
          gap_extend: int = 1):
    """Returns (score, start_position, cigar, ref_end, read_end) of JAlignerAligner.align_smith_waterman(seq, ref).

    s1 = ref (rows i), s2 = seq (columns j) as in JAlignerAligner.java:33.
    """
    if any(c >= MATRIX_SIZE for c in seq) or any(c >= MATRIX_SIZE for c in ref):
        raise ValueError("byte outside the 127 x 127 matrix (the reference throws ArrayIndexOutOfBoundsException)")
    a1, a2 = ref, seq
    m, n = len(a1) + 1, len(a2) + 1
    pointers = bytearray(m * n)           # borders are STOP (= 0)
    size_v = [1] * (m * n)                # lengths of the vertical / horizontal gap ending in a cell
    size_h = [1] * (m * n)
    ninf = float("-inf")
    g = [ninf] * n                        # best score ending in a vertical gap (consumes ref only: D)
    v = [0] * n                           # best score per cell of the previous / current row
    best, bi, bj = 0, 0, 0
    o, e = gap_open, gap_extend
    for i in range(1, m):
        h = ninf                          # best score ending in a horizontal gap (consumes seq only: I)
        vdiag = v[0]
        k = i * n
        c1 = a1[i - 1]
        for j in range(1, n):
            l = k + j
            f = vdiag + _score(c1, a2[j - 1], match, mismatch, ambiguous)
            g1, g2 = g[j] - e, v[j] - o   # extend / open (the opening penalty is the cost of the first gapped base)
            if g1 > g2:
                g[j] = g1
                size_v[l] = size_v[l - n] + 1
            else:
                g[j] = g2
            h1, h2 = h - e, v[j - 1] - o
            if h1 > h2:
                h = h1
                size_h[l] = size_h[l - 1] + 1
            else:
                h = h2
            vdiag = v[j]
            x = max(f, g[j], h, 0)
            v[j] = x
            if x == 0:
                pointers[l] = STOP
            elif x == f:
                pointers[l] = DIAGONAL
            elif x == g[j]:
                pointers[l] = UP
            else:
                pointers[l] = LEFT
            if x > best:                  # strict: first cell in (ref, seq) order
                best, bi, bj = x, i, j
    # traceback: whole gaps are jumped by their stored sizes, until a STOP pointer (H == 0)
    i, j = bi, bj
    k = i * n + j
    ops = []                              # reversed M / I / D
    while True:
        p = pointers[k]
        if p == UP:
            for _ in range(size_v[k]):
                i -= 1
                k -= n
                ops.append("D")
        elif p == DIAGONAL:
            i -= 1
            j -= 1
            k -= n + 1
            ops.append("M")
        elif p == LEFT:
            for _ in range(size_h[k]):
                j -= 1
                k -= 1
                ops.append("I")
        else:
            break
    ops.reverse()
    # JAlignerAligner.alignmentToCigar (JAlignerAligner.java:59-93)
    cigar = []
    if j > 0:
        cigar.append((j, "S"))
    length, op = 0, "M"
    for c in ops:
        if c != op:
            if length > 0:
                cigar.append((length, op))
            op, length = c, 0
        length += 1
    cigar.append((length, op))
    consumed = sum(n_ for n_, o_ in cigar if o_ in "MIS")
    if consumed != len(seq):
        cigar.append((len(seq) - consumed, "S"))
    return int(best), i, "".join(f"{n_}{o_}" for n_, o_ in cigar), bi - 1, bj - 1
