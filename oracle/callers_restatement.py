"""TEST INFRASTRUCTURE (like everything under oracle/; never imported by gridss_b200/): one-pair-at-a-time string restatements of the reference's two callers of
Aligner.align_smith_waterman, written the way the reference does it (host strings, one aligner call each), to
check gridss_b200/callers.py's batched / window-descriptor formulation against.

  homology_strings : BreakpointHomology.calculate   (alignment/BreakpointHomology.java:38-103)
  realign_strings  : SAMRecordUtil.realign          (sam/SAMRecordUtil.java:584-661)
`align(seq: bytes, ref: bytes) -> (start, cigar string)` is any Aligner.
"""
import re

_COMP = bytes.maketrans(b"ACGTNacgtn", b"TGCANtgcan")
_CIG = re.compile(r"(\d+)([MIDNSHP=X])")


def rc(b: bytes) -> bytes:
    return b.translate(_COMP)[::-1]


def anchor(contigs, ref_index, direction, pos, length) -> bytes:
    """BreakendSummary.getAnchorSequence (BreakendSummary.java:73-106), exact position."""
    if direction == "f":
        end = pos
        start = end - length + 1
    else:
        start = pos
        end = start + length - 1
    start = max(1, start)
    end = min(len(contigs[ref_index]), end)
    if start > end:
        return b""
    s = bytes(contigs[ref_index][start - 1:end])
    return rc(s) if direction == "b" else s


def event_size(bp):
    r1, d1, p1, r2, d2, p2 = bp
    if r1 != r2:
        return None
    d = abs(p2 - p1)
    if d1 == d2:
        return d
    if (p1 < p2 and d1 == "f") or (p2 < p1 and d1 == "b"):
        return d - 1
    return d + 1


def soft_clips(cigar: str):
    ops = [(int(n), op) for n, op in _CIG.findall(cigar)]

    def first(seq):
        for n, op in seq:
            if op == "S":
                return n
            if op != "H":
                return 0
        return 0

    return first(ops), first(ops[::-1])


def homology_pairs(contigs, bp, inserted, max_breakend_length, margin):
    """The two aligner calls of BreakpointHomology.calculate as (seq, ref) byte pairs (None where the reference skips
    the call), plus the lengths of localBsSeq and remoteBsSeq."""
    r1, d1, p1, r2, d2, p2 = bp
    ins = inserted or b""
    ins = ins.encode() if isinstance(ins, str) else bytes(ins)
    if d1 == "b":
        ins = rc(ins)
    seq_len = max_breakend_length
    ref_len = max_breakend_length + len(ins) + margin
    ev = event_size(bp)
    if ev is not None:
        seq_len, ref_len = min(seq_len, ev), min(ref_len, ev)
    adv = lambda p, d: p + ref_len if d == "f" else p - ref_len  # noqa: E731
    local_seq = anchor(contigs, r1, d1, p1, ref_len)
    local_bs = anchor(contigs, r1, d1, p1, seq_len)
    local_ref = anchor(contigs, r1, d1, adv(p1, d1), ref_len)
    remote_seq = rc(anchor(contigs, r2, d2, p2, ref_len))
    remote_bs = rc(anchor(contigs, r2, d2, p2, seq_len))
    remote_ref = rc(anchor(contigs, r2, d2, adv(p2, d2), ref_len))
    breakend = local_bs + ins + remote_bs
    local = local_seq + local_ref
    remote = remote_ref + remote_seq
    pair_a = (breakend, local) if breakend and local else None
    pair_b = (rc(breakend), rc(remote)) if breakend and remote else None
    return len(local_bs), len(remote_bs), pair_a, pair_b


def homology_from_cigars(n_local_bs, n_remote_bs, cigar_a, cigar_b):
    """(localHomologyLength, remoteHomologyLength) from the CIGARs of the two calls (None = call skipped)."""
    local_h = remote_h = 0
    if cigar_a is not None:
        lead, tail = soft_clips(cigar_a)
        remote_h = 0 if lead > 0 else max(0, n_remote_bs - tail)
    if cigar_b is not None:
        lead, tail = soft_clips(cigar_b)
        local_h = 0 if lead > 0 else max(0, n_local_bs - tail)
    return local_h, remote_h


def homology_strings(align, contigs, bp, inserted, max_breakend_length, margin):
    nl, nr, pair_a, pair_b = homology_pairs(contigs, bp, inserted, max_breakend_length, margin)
    cig_a = align(*pair_a)[1] if pair_a else None
    cig_b = align(*pair_b)[1] if pair_b else None
    return homology_from_cigars(nl, nr, cig_a, cig_b)


def realign_strings(align, contigs, read, window, extend):
    ref_index, pos0, cigar, bases = read
    pos = start = end = pos0
    ops = [(int(n), op) for n, op in _CIG.findall(cigar)]
    for n, op in ops:
        if op in "MDN=X":
            pos += n
        start = min(start, pos)
        end = max(end, pos - 1)
    start -= window
    end += window
    if extend:
        lead, tail = soft_clips(cigar)
        start -= lead
        end += tail
    start = max(1, start)
    end = min(len(contigs[ref_index]), end)
    ref = bytes(contigs[ref_index][start - 1:end]) if end >= start else b""
    if not bases or not ref:
        return pos0, cigar
    a_start, a_cigar = align(bytes(bases), ref)
    return start + a_start, a_cigar
