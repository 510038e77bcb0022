"""Batched mirrors of the reference's two live callers of `Aligner.align_smith_waterman`:

  BreakpointHomology.calculate            alignment/BreakpointHomology.java:38-103  (two SW calls per breakpoint)
  SAMRecordUtil.realign                   sam/SAMRecordUtil.java:584-661            (one SW call per read)

The reference makes one synchronous aligner call per pair from a worker pool. Here a caller hands over a whole list
and ONE GPU batch is issued for it: the queries (short) are built on the host and shipped, the reference windows
(long) are only NAMED as segments of the device-resident 2-bit reference and gathered on the GPU
(cudasw_align_batch_windows), including the reverse complements BreakpointHomology applies on the host.
Coordinates follow the reference: 1-based inclusive positions in BreakendSummary / SAM records.
"""
from __future__ import annotations

import re
from dataclasses import dataclass
from typing import Sequence

import numpy as np

from .aligner import BREAKPOINT_DTYPE, CUDASW_EINVAL, WINDOW_DTYPE, WINDOW_REVCOMP, CudaAligner, CudaSwError, cigar_string, pack

FWD, BWD = "f", "b"
_COMP = bytes.maketrans(b"ACGTNacgtn", b"TGCANtgcan")


def reverse_complement(b: bytes) -> bytes:
    return bytes(b).translate(_COMP)[::-1]


@dataclass(frozen=True)
class BreakendSummary:
    """BreakendSummary.java (exact positions only: start == end == nominal)."""

    referenceIndex: int
    direction: str
    position: int  # 1-based

    def advance(self, bases: int) -> "BreakendSummary":  # BreakendSummary.java:211-217
        return BreakendSummary(self.referenceIndex, self.direction, self.position + (bases if self.direction == FWD else -bases))

    def anchor_segment(self, contig_len: Sequence[int], length: int):
        """getAnchorSequence (BreakendSummary.java:73-106) as a reference segment: (contig, start0, end0, revcomp),
        clamped to the contig like the reference (its N padding is built but never returned); None when empty."""
        if self.direction == FWD:
            a_end = self.position
            a_start = a_end - length + 1
        else:
            a_start = self.position
            a_end = a_start + length - 1
        a_start = max(1, a_start)
        a_end = min(contig_len[self.referenceIndex], a_end)
        if a_start > a_end:
            return None
        return (self.referenceIndex, a_start - 1, a_end, self.direction == BWD)


@dataclass(frozen=True)
class BreakpointSummary:
    """BreakpointSummary.java:39-43 (exact breakpoints)."""

    referenceIndex: int
    direction: str
    position: int
    referenceIndex2: int
    direction2: str
    position2: int

    def local(self) -> BreakendSummary:
        return BreakendSummary(self.referenceIndex, self.direction, self.position)

    def remoteBreakend(self) -> BreakendSummary:  # BreakpointSummary.java:69-71
        return BreakendSummary(self.referenceIndex2, self.direction2, self.position2)

    def remoteBreakpoint(self) -> "BreakpointSummary":
        return BreakpointSummary(self.referenceIndex2, self.direction2, self.position2, self.referenceIndex, self.direction, self.position)

    def getEventSize(self):  # BreakpointSummary.java:226-236
        if self.referenceIndex != self.referenceIndex2:
            return None
        d = abs(self.position2 - self.position)
        if self.direction == self.direction2:
            return d
        if (self.position < self.position2 and self.direction == FWD) or (self.position2 < self.position and self.direction == BWD):
            return d - 1
        return d + 1


def _seg_bytes(contigs, seg) -> bytes:
    if seg is None:
        return b""
    c, s, e, rc = seg
    b = bytes(contigs[c][s:e])
    return reverse_complement(b) if rc else b


def _flip(seg):
    return None if seg is None else (seg[0], seg[1], seg[2], not seg[3])


def _start_soft_clip(ops) -> int:
    """getStartSoftClipLength (SAMRecordUtil.java:91-103) over [(length, op char)]."""
    for n, op in ops:
        if op == "S":
            return n
        if op != "H":
            break
    return 0


def _end_soft_clip(ops) -> int:
    """getEndSoftClipLength (SAMRecordUtil.java:113-123)."""
    return _start_soft_clip(ops[::-1])


_OPS = "MIDNSHP=X"


def _clips(words) -> tuple[int, int]:
    """(start, end) soft clip lengths of packed CIGAR words (length << 4 | op)."""
    ops = [(int(w) >> 4, _OPS[int(w) & 15]) for w in words]
    return _start_soft_clip(ops), _end_soft_clip(ops)


def breakpoint_homology_batch(aligner: CudaAligner, contigs: Sequence[bytes], breakpoints: Sequence[BreakpointSummary],
                              inserted: Sequence[str | bytes | None], maxBreakendLength: int, margin: int):
    """BreakpointHomology.calculate for a list of breakpoints in one GPU batch -> [(localHomologyLength,
    remoteHomologyLength)]. `aligner` must hold `contigs` as its device-resident reference (load_reference)."""
    clen = [len(c) for c in contigs]
    seqs, segs, win_off, owner = [], [], [0], []
    meta = []
    for k, (bs, ins) in enumerate(zip(breakpoints, inserted)):
        ins = b"" if ins is None else (ins.encode("ascii") if isinstance(ins, str) else bytes(ins))
        if bs.direction == BWD:
            ins = reverse_complement(ins)  # BreakpointHomology.java:45-47
        seq_len = maxBreakendLength
        ref_len = maxBreakendLength + len(ins) + margin
        ev = bs.getEventSize()
        if ev is not None:
            seq_len, ref_len = min(seq_len, ev), min(ref_len, ev)
        loc, rem = bs.local(), bs.remoteBreakend()
        local_seq = loc.anchor_segment(clen, ref_len)
        local_bs = loc.anchor_segment(clen, seq_len)
        local_ref = loc.advance(ref_len).anchor_segment(clen, ref_len)
        remote_seq = _flip(rem.anchor_segment(clen, ref_len))      # reverseComplement(...) of :64-66
        remote_bs = _flip(rem.anchor_segment(clen, seq_len))
        remote_ref = _flip(rem.advance(ref_len).anchor_segment(clen, ref_len))
        local_bs_b, remote_bs_b = _seg_bytes(contigs, local_bs), _seg_bytes(contigs, remote_bs)
        breakend = local_bs_b + ins + remote_bs_b               # :67
        meta.append((len(local_bs_b), len(remote_bs_b)))
        if not breakend:
            continue
        local_win = [s for s in (local_seq, local_ref) if s is not None]       # strLocal = localSeq + localRef (:68)
        # strRemote = remoteRef + remoteSeq (:69), reverse-complemented together with the breakend (:87-89):
        # rc(remoteRef + remoteSeq) = rc(remoteSeq) + rc(remoteRef)
        remote_win = [_flip(s) for s in (remote_seq, remote_ref) if s is not None]
        if local_win:
            seqs.append(breakend)
            segs.extend(local_win)
            win_off.append(len(segs))
            owner.append((k, 0))
        if remote_win:
            seqs.append(reverse_complement(breakend))
            segs.extend(remote_win)
            win_off.append(len(segs))
            owner.append((k, 1))
    out = [[0, 0] for _ in breakpoints]
    if seqs:
        sb, so = pack(seqs)
        win = np.zeros(len(segs), dtype=WINDOW_DTYPE)
        for i, (c, s, e, rc) in enumerate(segs):
            win[i] = (c, WINDOW_REVCOMP if rc else 0, s, e)
        res, cig = aligner.align_blobs_windows(sb, so, win, np.asarray(win_off, dtype=np.int64))
        for i, (k, side) in enumerate(owner):
            r = res[i]
            if r["status"] == CUDASW_EINVAL:
                continue  # a window the reference clipped to nothing / outside the contig: it skips the call (:76, :86)
            if r["status"] != 0:  # ENOMEM / ETOOLARGE / EUNDEF are not "no alignment": the reference would have thrown or crashed
                raise CudaSwError(f"breakpoint {k}: alignment failed with status {int(r['status'])}")
            lead, tail = _clips(cig[r["cigar_off"]: r["cigar_off"] + r["cigar_len"]])
            n_local, n_remote = meta[k]
            if side == 0:   # local alignment -> remote homology (:79-85)
                out[k][1] = 0 if lead > 0 else max(0, n_remote - tail)
            else:           # remote alignment -> local homology (:91-100)
                out[k][0] = 0 if lead > 0 else max(0, n_local - tail)
    return [tuple(x) for x in out]


def breakpoint_array(breakpoints: Sequence[BreakpointSummary]) -> np.ndarray:
    a = np.zeros(len(breakpoints), dtype=BREAKPOINT_DTYPE)
    for k, b in enumerate(breakpoints):
        a[k] = (b.referenceIndex, int(b.direction == BWD), b.position, b.referenceIndex2, int(b.direction2 == BWD), b.position2)
    return a


def breakpoint_homology_device(aligner: CudaAligner, breakpoints, inserted: Sequence[str | bytes | None] | None,
                               maxBreakendLength: int, margin: int):
    """Same answers as breakpoint_homology_batch through cudasw_breakpoint_homology: the library derives every query
    and window segment itself, gathers both on the device and skips the traceback (only clip lengths are read).
    `breakpoints` is a list of BreakpointSummary or a BREAKPOINT_DTYPE array."""
    bp = breakpoints if isinstance(breakpoints, np.ndarray) else breakpoint_array(breakpoints)
    blob = off = None
    if inserted is not None:
        ins = [b"" if i is None else (i.encode("ascii") if isinstance(i, str) else bytes(i)) for i in inserted]
        blob, off = pack(ins)
    local, remote = aligner.breakpoint_homology(bp, blob, off, maxBreakendLength, margin)
    return list(zip(local.tolist(), remote.tolist()))


_CIG = re.compile(r"(\d+)([MIDNSHP=X])")


def realign_batch(aligner: CudaAligner, contig_len: Sequence[int], reads, windowSize: int, extendWindowForClippedBases: bool):
    """SAMRecordUtil.realign for a list of mapped reads in one GPU batch. A read is (referenceIndex, alignmentStart
    (1-based), cigar string, read bases); the result per read is (alignmentStart, cigar) — the input values when the
    reference would have returned the read unchanged."""
    seqs, win, keep = [], [], []
    starts = []
    for ref_idx, pos0, cigar, bases in reads:
        pos = start = end = pos0
        ops = [(int(n), op) for n, op in _CIG.findall(cigar)]
        for n, op in ops:                                   # SAMRecordUtil.java:595-601
            if op in "MDN=X":
                pos += n
            start, end = min(start, pos), max(end, pos - 1)
        start -= windowSize
        end += windowSize
        if extendWindowForClippedBases:                     # :605-608
            start -= _start_soft_clip(ops)
            end += _end_soft_clip(ops)
        start = max(1, start)
        end = min(contig_len[ref_idx], end)
        bases = bytes(bases)
        if not bases or end < start:
            keep.append(None)
            continue
        keep.append(len(seqs))
        seqs.append(bases)                                   # non-ACGT -> N happens in the engine (clean(), :618-627)
        win.append((ref_idx, 0, start - 1, end))
        starts.append(start)
    out = []
    res = cig = None
    if seqs:
        sb, so = pack(seqs)
        res, cig = aligner.align_blobs_windows(sb, so, np.array(win, dtype=WINDOW_DTYPE), np.arange(len(win) + 1, dtype=np.int64))
    for (ref_idx, pos0, cigar, bases), k in zip(reads, keep):
        if k is None or res[k]["status"] != 0:
            # alignment == null -> return read (:643-645); the reference swallows EVERY exception of the aligner here
            # (catch (Exception e), :632-644), so any per-pair failure keeps the original alignment; it is counted so
            # that a caller can tell "unchanged" from "could not be realigned"
            if k is not None:
                realign_batch.failed += 1
            out.append((pos0, cigar))
            continue
        r = res[k]
        new_cigar = cigar_string(cig[r["cigar_off"]: r["cigar_off"] + r["cigar_len"]])
        out.append((starts[k] + int(r["ref_begin"]), new_cigar))   # :647
    return out


realign_batch.failed = 0  # reads whose realignment failed in the engine (kept unchanged, as the reference does)


class SmithWatermanStreamingAligner:
    """SmithWatermanStreamingAligner.java:16-69 (and the loop of SmithWatermanFastqAligner.java:24-45): every read is
    aligned against one WHOLE contig. The reference aligns inside asyncAlign; here reads queue up and flush() (or
    reaching `batch` queued reads) issues one GPU batch whose windows all name the same contig-long segment of the
    device-resident reference. Records come back in submission order as (readName, referenceIndex,
    alignmentStart (1-based), cigar)."""

    def __init__(self, aligner: CudaAligner, contig_len: Sequence[int], referenceIndex: int, batch: int = 4096):
        self.aligner, self.referenceIndex, self.batch = aligner, referenceIndex, batch
        self._win = np.zeros(1, dtype=WINDOW_DTYPE)
        self._win[0] = (referenceIndex, 0, 0, contig_len[referenceIndex])
        self._pending: list[tuple[str, bytes]] = []
        self._done: list[tuple[str, int, int, str]] = []

    def asyncAlign(self, readName: str, bases: bytes) -> None:
        if not bases:
            raise ValueError("seq and ref must be non-zero size")   # SswJniAligner.java:45-50
        self._pending.append((readName, bytes(bases)))
        if len(self._pending) >= self.batch:
            self.flush()

    def flush(self) -> None:
        if not self._pending:
            return
        n = len(self._pending)
        sb, so = pack([b for _, b in self._pending])
        res, cig = self.aligner.align_blobs_windows(sb, so, np.repeat(self._win, n), np.arange(n + 1, dtype=np.int64))
        for (name, _), r in zip(self._pending, res):
            if r["status"] != 0:
                # e.g. CUDASW_ETOOLARGE: a contig beyond 2^24 - 2 columns or a read beyond 32 k rows. The reference would
                # have produced an alignment (or died in native code); silently emitting start 0 / empty CIGAR is neither
                self._pending.clear()
                raise CudaSwError(f"read {name}: alignment failed with status {int(r['status'])}")
            words = cig[r["cigar_off"]: r["cigar_off"] + r["cigar_len"]]
            self._done.append((name, self.referenceIndex, int(r["ref_begin"]) + 1, cigar_string(words)))   # :36-38
        self._pending.clear()

    def processedAlignmentRecords(self) -> int:
        return len(self._done)

    def outstandingAlignmentRecord(self) -> int:
        return len(self._pending)

    def getAlignment(self):
        if not self._done:
            self.flush()
        return self._done.pop(0)

    def close(self) -> None:
        self.flush()
