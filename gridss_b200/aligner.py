"""Python mirror of the reference's Aligner API over the C ABI of libcudasw.so (ctypes).

Mirrors, name for name where Python allows:
  Aligner.align_smith_waterman(seq, ref) -> Alignment           alignment/Aligner.java:10
  Alignment.getStartPosition() / getCigar()                      alignment/Alignment.java:23,30
  AlignerFactory.create() / create(match, mismatch, ambiguous, gapOpen, gapExtend)
                                                                 alignment/AlignerFactory.java:108-120
  StreamingAligner-style asyncAlign / flush / getAlignment       alignment/StreamingAligner.java:10-44
Error behaviour follows SswJniAligner.java:45-50 (empty seq/ref -> IllegalArgumentException, here
ValueError) and ssw.Aligner.align's argument checks (gap in [0,255], |score| <= 127).
"""
from __future__ import annotations

import ctypes
import os
import threading
from typing import Iterable, Sequence

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
# CUDASW_LIB selects another build of the same library (the -DCUSW_BOUNDS_CHECK flavour, gridss_b200/build.py --check)
LIB_PATH = os.environ.get("CUDASW_LIB") or os.path.join(HERE, "libcudasw.so")

CUDASW_OK = 0
CUDASW_EINVAL = -1
CUDASW_ENOMEM = -2
CUDASW_ECUDA = -3
CUDASW_ETOOLARGE = -4
CUDASW_EUNDEF = -5
TIEBREAK_SSW, TIEBREAK_JALIGNER = 0, 1

RESULT_DTYPE = np.dtype(
    [
        ("score", "<i4"),
        ("ref_begin", "<i4"),
        ("ref_end", "<i4"),
        ("read_begin", "<i4"),
        ("read_end", "<i4"),
        ("cigar_off", "<i4"),
        ("cigar_len", "<i4"),
        ("status", "<i4"),
    ]
)
# cudasw_window (include/cudasw.h): one window segment = bases [start, end) of a loaded contig, optionally
# reverse-complemented (flags & WINDOW_REVCOMP)
WINDOW_DTYPE = np.dtype([("contig", "<i4"), ("flags", "<i4"), ("start", "<i8"), ("end", "<i8")])
WINDOW_REVCOMP = 1
# cudasw_breakpoint (include/cudasw.h): exact 1-based positions, dir 0 = forward / 1 = backward
BREAKPOINT_DTYPE = np.dtype([("contig1", "<i4"), ("dir1", "<i4"), ("pos1", "<i8"), ("contig2", "<i4"), ("dir2", "<i4"), ("pos2", "<i8")])
_OPS = "MIDNSHP=X"


class CudaSwError(RuntimeError):
    pass


class _Scoring(ctypes.Structure):
    _fields_ = [(n, ctypes.c_int32) for n in ("match", "mismatch", "ambiguous", "gap_open", "gap_extend")]


class Stats(ctypes.Structure):
    _fields_ = [
        ("pairs", ctypes.c_uint64),
        ("cells", ctypes.c_uint64),
        ("kernel_launches", ctypes.c_uint64),
        ("ms_total", ctypes.c_float),
        ("ms_h2d", ctypes.c_float),
        ("ms_pack", ctypes.c_float),
        ("ms_forward", ctypes.c_float),
        ("ms_reverse", ctypes.c_float),
        ("ms_traceback", ctypes.c_float),
        ("ms_d2h", ctypes.c_float),
        ("h2d_bytes", ctypes.c_uint64),
        ("d2h_bytes", ctypes.c_uint64),
        ("forward_cells_padded", ctypes.c_uint64),
        ("ms_encode", ctypes.c_float),
        ("encode_bytes", ctypes.c_uint64),
        ("filter_banded", ctypes.c_uint64),
        ("filter_failed", ctypes.c_uint64),
    ]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


_lib = None
_lib_lock = threading.Lock()

# every symbol include/cudasw.h declares: (name, restype, argtypes)
_vp, _i64, _i32, _sz = ctypes.c_void_p, ctypes.c_int64, ctypes.c_int32, ctypes.c_size_t
ABI = [
    ("cudasw_create", ctypes.c_int, [ctypes.POINTER(_Scoring), ctypes.c_int, ctypes.POINTER(_vp)]),
    ("cudasw_create_ex", ctypes.c_int, [ctypes.POINTER(_Scoring), ctypes.c_int, ctypes.c_int, ctypes.POINTER(_vp)]),
    ("cudasw_destroy", None, [_vp]),
    ("cudasw_align_batch", ctypes.c_int, [_vp, _i64, _vp, _vp, _vp, _vp, _vp, _vp, _i64, ctypes.POINTER(_i64)]),
    ("cudasw_align_batch_packed", ctypes.c_int, [_vp, _i64, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i64, ctypes.POINTER(_i64)]),
    ("cudasw_pack_bases", _i64, [_vp, _i64, _i64, _vp, _vp]),
    ("cudasw_multi_align_batch_packed", ctypes.c_int, [_vp, _i64, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i64, ctypes.POINTER(_i64)]),
    ("cudasw_align_batch_device", ctypes.c_int, [_vp, _i64, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _i64, ctypes.POINTER(_i64)]),
    ("cudasw_reference_load", ctypes.c_int, [_vp, _i32, _vp, _vp]),
    ("cudasw_reference_fetch", ctypes.c_int, [_vp, _vp, _i32, _vp, _i64]),
    ("cudasw_align_batch_windows", ctypes.c_int, [_vp, _i64, _vp, _vp, _vp, _vp, _vp, _vp, _i64, ctypes.POINTER(_i64)]),
    ("cudasw_breakpoint_homology", ctypes.c_int, [_vp, _i64, _vp, _vp, _vp, _i32, _i32, _vp, _vp]),
    ("cudasw_multi_create", ctypes.c_int, [ctypes.POINTER(_Scoring), _vp, ctypes.c_int, ctypes.POINTER(_vp)]),
    ("cudasw_multi_destroy", None, [_vp]),
    ("cudasw_multi_device_count", ctypes.c_int, [_vp]),
    ("cudasw_multi_align_batch", ctypes.c_int, [_vp, _i64, _vp, _vp, _vp, _vp, _vp, _vp, _i64, ctypes.POINTER(_i64)]),
    ("cudasw_multi_reference_load", ctypes.c_int, [_vp, _i32, _vp, _vp]),
    ("cudasw_multi_align_batch_windows", ctypes.c_int, [_vp, _i64, _vp, _vp, _vp, _vp, _vp, _vp, _i64, ctypes.POINTER(_i64)]),
    ("cudasw_multi_breakpoint_homology", ctypes.c_int, [_vp, _i64, _vp, _vp, _vp, _i32, _i32, _vp, _vp]),
    ("cudasw_multi_get_stats", ctypes.c_int, [_vp, ctypes.c_int, ctypes.POINTER(Stats), ctypes.POINTER(_i64), ctypes.POINTER(_i64)]),
    ("cudasw_multi_last_error", ctypes.c_char_p, [_vp]),
    ("cudasw_device_alloc", ctypes.c_int, [_vp, _sz, ctypes.POINTER(_vp)]),
    ("cudasw_device_free", ctypes.c_int, [_vp, _vp]),
    ("cudasw_upload", ctypes.c_int, [_vp, _vp, _vp, _sz]),
    ("cudasw_host_alloc", ctypes.c_int, [_sz, ctypes.POINTER(_vp)]),
    ("cudasw_host_free", ctypes.c_int, [_vp]),
    ("cudasw_submit", ctypes.c_int, [_vp, _vp, _i64, _vp, _i64, ctypes.POINTER(_i64)]),
    ("cudasw_flush", ctypes.c_int, [_vp]),
    ("cudasw_outstanding", _i64, [_vp]),
    ("cudasw_processed", _i64, [_vp]),
    ("cudasw_poll", ctypes.c_int, [_vp, _i64, _vp, _vp, _i32, ctypes.POINTER(_i32)]),
    ("cudasw_wait", ctypes.c_int, [_vp, _i64, _i64, _vp, _vp, _i32, ctypes.POINTER(_i32)]),
    ("cudasw_dispatcher_start", ctypes.c_int, [_vp, _i64, _i64]),
    ("cudasw_dispatcher_stop", ctypes.c_int, [_vp]),
    ("cudasw_get_stats", ctypes.c_int, [_vp, ctypes.POINTER(Stats)]),
    ("cudasw_variant_for", ctypes.c_int, [_vp, _i32, ctypes.POINTER(_i32), ctypes.POINTER(_i32), ctypes.POINTER(_i32)]),
    ("cudasw_last_error", ctypes.c_char_p, [_vp]),
    ("cudasw_device_count", ctypes.c_int, []),
    ("cudasw_version", ctypes.c_char_p, []),
    ("cudasw_cigar_string", ctypes.c_int, [_vp, _i32, ctypes.c_char_p, _i32]),
    # no allocation, no handle: device pointers in and out (include/cudasw.h; tests/test_zz_no_alloc.py)
    ("cudasw_score_batch_noalloc", ctypes.c_int, [ctypes.POINTER(_Scoring), ctypes.c_int, _i32, _vp, _vp, _i64, _vp, _vp, _i64, _vp, _vp, ctypes.c_char_p, _sz]),
]


def load_library():
    """Load libcudasw.so (built in-tree by gridss_b200/build.py). Fails loudly if it is missing."""
    global _lib
    with _lib_lock:
        if _lib is None:
            if not os.path.exists(LIB_PATH):
                raise CudaSwError(
                    f"{LIB_PATH} not built (run `python -m gridss_b200.build`); gridss_b200 has no CPU fallback"
                )
            lib = ctypes.CDLL(LIB_PATH)
            for name, res, args in ABI:
                fn = getattr(lib, name)
                fn.restype = res
                fn.argtypes = args
            _lib = lib
    return _lib


def cigar_string(words: Iterable[int]) -> str:
    return "".join(f"{int(w) >> 4}{_OPS[int(w) & 15]}" for w in words)


def pack(seqs: Sequence[bytes]):
    off = np.zeros(len(seqs) + 1, dtype=np.int64)
    if len(seqs):
        off[1:] = np.cumsum([len(s) for s in seqs])
    blob = np.frombuffer(b"".join(bytes(s) for s in seqs), dtype=np.uint8)
    if blob.size == 0:
        blob = np.zeros(1, dtype=np.uint8)
    return np.ascontiguousarray(blob), off


def packed_words(n_bases: int) -> tuple[int, int]:
    """(two_bit words, nmask words) a packed blob of n_bases needs: whole 32-base groups."""
    g = (int(n_bases) + 31) // 32
    return 2 * max(g, 1), max(g, 1)


def pack_bases(blob: np.ndarray, two_bit: np.ndarray | None = None, nmask: np.ndarray | None = None, base0: int = 0,
               n: int | None = None, threads: int = 1):
    """ASCII bases -> the packed layout of cudasw_align_batch_packed (2 bits per base + N mask), through the library's
    own host packer cudasw_pack_bases. Returns (two_bit, nmask, n_ambiguous). `threads` > 1 packs disjoint 32-base
    aligned ranges in parallel (the C call releases the GIL)."""
    lib = load_library()
    n = int(len(blob) if n is None else n)
    w2, wn = packed_words(base0 + n)
    if two_bit is None:
        two_bit = np.zeros(w2, dtype=np.uint32)
    if nmask is None:
        nmask = np.zeros(wn, dtype=np.uint32)
    assert two_bit.dtype == np.uint32 and nmask.dtype == np.uint32 and len(two_bit) >= w2 and len(nmask) >= wn
    if threads <= 1 or n < (1 << 20) or base0 & 31:
        amb = lib.cudasw_pack_bases(blob.ctypes.data, n, base0, two_bit.ctypes.data, nmask.ctypes.data)
    else:
        from concurrent.futures import ThreadPoolExecutor
        step = ((n + threads - 1) // threads + 31) & ~31
        cuts = list(range(0, n, step))
        with ThreadPoolExecutor(max_workers=threads) as ex:
            parts = list(ex.map(lambda a: lib.cudasw_pack_bases(blob.ctypes.data + a, min(step, n - a), base0 + a,
                                                                two_bit.ctypes.data, nmask.ctypes.data), cuts))
        amb = -1 if any(x < 0 for x in parts) else sum(parts)
    if amb < 0:
        raise ValueError("cudasw_pack_bases: bad argument")
    return two_bit, nmask, int(amb)


class Alignment:
    """alignment/Alignment.java:8-34."""

    __slots__ = ("_start", "_cigar", "score")

    def __init__(self, startPosition: int, cigar: str, score: int | None = None):
        self._start = int(startPosition)
        self._cigar = cigar
        self.score = score

    def getStartPosition(self) -> int:
        return self._start

    def getCigar(self) -> str:
        return self._cigar

    def __repr__(self):
        return f"Alignment(startPosition={self._start}, cigar={self._cigar!r})"

    def __eq__(self, other):
        return isinstance(other, Alignment) and (self._start, self._cigar) == (other._start, other._cigar)


class Aligner:
    """alignment/Aligner.java:3-11."""

    def align_smith_waterman(self, seq: bytes, ref: bytes) -> Alignment:  # pragma: no cover - interface
        raise NotImplementedError


class PinnedBuffer:
    """Page-locked host memory from cudasw_host_alloc, exposed as a numpy uint8 array."""

    def __init__(self, nbytes: int):
        lib = load_library()
        p = ctypes.c_void_p()
        rc = lib.cudasw_host_alloc(max(1, nbytes), ctypes.byref(p))
        if rc != 0:
            raise CudaSwError("cudasw_host_alloc failed")
        self.ptr = p.value
        self.nbytes = nbytes
        self.array = np.ctypeslib.as_array((ctypes.c_uint8 * max(1, nbytes)).from_address(self.ptr))

    def close(self):
        if self.ptr:
            load_library().cudasw_host_free(self.ptr)
            self.ptr = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class CudaAligner(Aligner):
    """The drop-in for SswJniAligner (alignment/SswJniAligner.java:8-61) on one B200."""

    def __init__(self, match=1, mismatch=-4, ambiguous=-4, gapOpen=6, gapExtend=1, device: int = 0, tie_break: str = "ssw"):
        """tie_break: "ssw" = bit-compatible with SswJniAligner (the reference's default backend); "jaligner" = with
        JAlignerAligner, its --nojni fallback (unpinned: no JVM here; see include/cudasw.h)."""
        lib = load_library()
        self._lib = lib
        self.scoring = (match, mismatch, ambiguous, gapOpen, gapExtend)
        sc = _Scoring(match, mismatch, ambiguous, gapOpen, gapExtend)
        h = ctypes.c_void_p()
        rc = lib.cudasw_create_ex(ctypes.byref(sc), device, {"ssw": TIEBREAK_SSW, "jaligner": TIEBREAK_JALIGNER}[tie_break], ctypes.byref(h))
        if rc != 0:
            msg = (lib.cudasw_last_error(None) or b"").decode()
            if rc == CUDASW_EINVAL:
                raise ValueError(msg)  # IllegalArgumentException in the reference
            raise CudaSwError(f"cudasw_create failed ({rc}): {msg}")
        self._h = h
        self.device = device

    # -- lifecycle
    def close(self):
        if getattr(self, "_h", None):
            self._lib.cudasw_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _err(self) -> str:
        return (self._lib.cudasw_last_error(self._h) or b"").decode()

    # -- Aligner.align_smith_waterman (Aligner.java:10)
    def align_smith_waterman(self, seq: bytes, ref: bytes) -> Alignment:
        if seq is None or len(seq) == 0:
            raise ValueError("seq must be non-zero size")  # SswJniAligner.java:45-47
        if ref is None or len(ref) == 0:
            raise ValueError("ref must be non-zero size")  # SswJniAligner.java:48-50
        return self.align_batch([seq], [ref])[0]

    # -- batched entry points
    def align_blobs(self, seq_blob, seq_off, ref_blob, ref_off, cigar_cap: int | None = None, res=None, cig=None):
        """Host blobs (numpy uint8 + int64 offsets) -> (RESULT_DTYPE[n], uint32 cigar blob).

        `res` / `cig` may supply preallocated output arrays (e.g. views of pinned memory)."""
        n = len(seq_off) - 1
        if res is None:
            res = np.zeros(n, dtype=RESULT_DTYPE)
        assert len(res) >= n
        cap = int(cigar_cap if cigar_cap is not None else max(64, 24 * n))
        if cig is not None:
            cap = len(cig)
        while True:
            if cig is None or len(cig) < cap:
                cig = np.zeros(cap, dtype=np.uint32)
            used = ctypes.c_int64(0)
            rc = self._lib.cudasw_align_batch(
                self._h, n, seq_blob.ctypes.data, seq_off.ctypes.data, ref_blob.ctypes.data,
                ref_off.ctypes.data, res.ctypes.data, cig.ctypes.data, cap, ctypes.byref(used))
            if rc == CUDASW_ENOMEM and used.value > cap:
                cap = int(used.value) + 16
                continue
            if rc != 0:
                raise CudaSwError(f"cudasw_align_batch failed ({rc}): {self._err()}")
            return res, cig[: used.value]

    _packed_fn = "cudasw_align_batch_packed"

    def align_blobs_packed(self, seq2, seqN, seq_off, ref2, refN, ref_off, res=None, cig=None):
        """Host blobs packed by pack_bases (2 bits per base; seqN / refN may be None = no ambiguous base in that blob).
        Same outputs as align_blobs on the ASCII bytes; 3/8 of the bytes cross PCIe."""
        n = len(seq_off) - 1
        res = np.zeros(n, dtype=RESULT_DTYPE) if res is None else res
        cap = max(64, 24 * n) if cig is None else len(cig)
        while True:
            if cig is None or len(cig) < cap:
                cig = np.zeros(cap, dtype=np.uint32)
            used = ctypes.c_int64(0)
            rc = getattr(self._lib, self._packed_fn)(
                self._h, n, seq2.ctypes.data, None if seqN is None else seqN.ctypes.data, seq_off.ctypes.data,
                ref2.ctypes.data, None if refN is None else refN.ctypes.data, ref_off.ctypes.data,
                res.ctypes.data, cig.ctypes.data, cap, ctypes.byref(used))
            if rc == CUDASW_ENOMEM and used.value > cap:
                cap = int(used.value) + 16
                continue
            if rc != 0:
                raise CudaSwError(f"{self._packed_fn} failed ({rc}): {self._err()}")
            return res, (cig[: used.value] if self._packed_fn == "cudasw_align_batch_packed" else cig)

    def align_blobs_device(self, d_seq, d_seq_off, d_ref, d_ref_off, h_seq_off, h_ref_off, res=None, cig=None):
        """Inputs already resident in device memory (raw device pointers as ints)."""
        n = len(h_seq_off) - 1
        res = np.zeros(n, dtype=RESULT_DTYPE) if res is None else res
        cig = np.zeros(max(64, 24 * n), dtype=np.uint32) if cig is None else cig
        used = ctypes.c_int64(0)
        rc = self._lib.cudasw_align_batch_device(
            self._h, n, d_seq, d_seq_off, d_ref, d_ref_off, h_seq_off.ctypes.data, h_ref_off.ctypes.data,
            res.ctypes.data, cig.ctypes.data, cig.size, ctypes.byref(used))
        if rc != 0:
            raise CudaSwError(f"cudasw_align_batch_device failed ({rc}): {self._err()}")
        return res, cig[: used.value]

    def align_batch(self, seqs: Sequence[bytes], refs: Sequence[bytes]) -> list[Alignment]:
        """Alignment[] alignBatch(byte[][] seqs, byte[][] refs)."""
        if len(seqs) != len(refs):
            raise ValueError("seqs and refs differ in length")
        for s, r in zip(seqs, refs):
            if s is None or len(s) == 0:
                raise ValueError("seq must be non-zero size")
            if r is None or len(r) == 0:
                raise ValueError("ref must be non-zero size")
        sb, so = pack(seqs)
        rb, ro = pack(refs)
        res, cig = self.align_blobs(sb, so, rb, ro)
        out = []
        for r in res:
            if r["status"] != 0:
                raise CudaSwError(f"pair failed with status {int(r['status'])}")
            out.append(Alignment(int(r["ref_begin"]), cigar_string(cig[r["cigar_off"] : r["cigar_off"] + r["cigar_len"]]), int(r["score"])))
        return out

    # -- device-resident reference (TwoBitBufferedReferenceSequenceFile.java:97-125)
    def load_reference(self, contigs: Sequence[bytes]) -> None:
        """Pack the contigs into HBM (2 bits per base + an N mask). Replaces any earlier reference."""
        blob, off = pack(contigs)
        rc = self._lib.cudasw_reference_load(self._h, len(contigs), blob.ctypes.data, off.ctypes.data)
        if rc != 0:
            raise (ValueError if rc == CUDASW_EINVAL else CudaSwError)(f"cudasw_reference_load failed ({rc}): {self._err()}")
        self._contig_len = [len(c) for c in contigs]

    def load_reference_blob(self, blob: np.ndarray, contig_off: np.ndarray) -> None:
        """Same from one uint8 blob + int64 contig offsets (no copy on the host)."""
        rc = self._lib.cudasw_reference_load(self._h, len(contig_off) - 1, blob.ctypes.data, contig_off.ctypes.data)
        if rc != 0:
            raise (ValueError if rc == CUDASW_EINVAL else CudaSwError)(f"cudasw_reference_load failed ({rc}): {self._err()}")

    def get_subsequence(self, contig: int, start: int, end: int, revcomp: bool = False) -> bytes:
        """getSubsequenceAt(contig, start+1, end) of the packed reference: upper-case ACGT, anything else N."""
        win = np.zeros(1, dtype=WINDOW_DTYPE)
        win[0] = (contig, WINDOW_REVCOMP if revcomp else 0, start, end)
        out = np.zeros(max(1, end - start), dtype=np.uint8)
        rc = self._lib.cudasw_reference_fetch(self._h, win.ctypes.data, 1, out.ctypes.data, out.size)
        if rc < 0:
            raise (ValueError if rc == CUDASW_EINVAL else CudaSwError)(f"cudasw_reference_fetch failed ({rc}): {self._err()}")
        return out[:rc].tobytes()

    def align_blobs_windows(self, seq_blob, seq_off, win, win_off, res=None, cig=None):
        """Queries from host blobs, windows named by WINDOW_DTYPE segments of the loaded reference
        (pair i uses win[win_off[i]:win_off[i+1]] concatenated). Same outputs as align_blobs."""
        n = len(seq_off) - 1
        assert win.dtype == WINDOW_DTYPE and len(win_off) == n + 1
        res = np.zeros(n, dtype=RESULT_DTYPE) if res is None else res
        cap = max(64, 24 * n) if cig is None else len(cig)
        while True:
            if cig is None or len(cig) < cap:
                cig = np.zeros(cap, dtype=np.uint32)
            used = ctypes.c_int64(0)
            rc = self._lib.cudasw_align_batch_windows(
                self._h, n, seq_blob.ctypes.data, seq_off.ctypes.data, win.ctypes.data, win_off.ctypes.data,
                res.ctypes.data, cig.ctypes.data, cap, ctypes.byref(used))
            if rc == CUDASW_ENOMEM and used.value > cap:
                cap = int(used.value) + 16
                continue
            if rc != 0:
                raise CudaSwError(f"cudasw_align_batch_windows failed ({rc}): {self._err()}")
            return res, cig[: used.value]

    _homology_fn = "cudasw_breakpoint_homology"

    def breakpoint_homology(self, bp: np.ndarray, ins_blob=None, ins_off=None, max_breakend_length: int = 300, margin: int = 10):
        """BreakpointHomology.calculate (BreakpointHomology.java:38-103) for BREAKPOINT_DTYPE[n] against the loaded
        reference -> (localHomologyLength int32[n], remoteHomologyLength int32[n]). The defaults are
        variantcalling.maxBreakendHomologyLength / breakendHomologyAlignmentMargin of gridss.properties:88-89."""
        assert bp.dtype == BREAKPOINT_DTYPE
        n = len(bp)
        local = np.zeros(n, dtype=np.int32)
        remote = np.zeros(n, dtype=np.int32)
        if ins_off is not None:
            ins_off = np.ascontiguousarray(ins_off, dtype=np.int64)
            assert len(ins_off) == n + 1
            if ins_blob is None or len(ins_blob) == 0:
                ins_blob = np.zeros(1, dtype=np.uint8)
        bp = np.ascontiguousarray(bp)
        rc = getattr(self._lib, self._homology_fn)(
            self._h, n, bp.ctypes.data, None if ins_off is None else ins_blob.ctypes.data,
            None if ins_off is None else ins_off.ctypes.data, max_breakend_length, margin, local.ctypes.data, remote.ctypes.data)
        if rc != 0:
            raise (ValueError if rc == CUDASW_EINVAL else CudaSwError)(f"{self._homology_fn} failed ({rc}): {self._err()}")
        return local, remote

    def variant_for(self, query_len: int) -> tuple[int, int, int]:
        """(lanes per pair, rows per virtual lane, strips) of the score kernel a query length maps to."""
        g, r, s = ctypes.c_int32(), ctypes.c_int32(), ctypes.c_int32()
        self._lib.cudasw_variant_for(self._h, query_len, ctypes.byref(g), ctypes.byref(r), ctypes.byref(s))
        return g.value, r.value, s.value

    def stats(self) -> dict:
        st = Stats()
        self._lib.cudasw_get_stats(self._h, ctypes.byref(st))
        return st.as_dict()

    # -- async trio (StreamingAligner.java:10-44)
    def submit(self, seq: bytes, ref: bytes) -> int:
        if not seq or not ref:
            raise ValueError("seq and ref must be non-zero size")
        t = ctypes.c_int64(0)
        s = np.frombuffer(bytes(seq), dtype=np.uint8)
        r = np.frombuffer(bytes(ref), dtype=np.uint8)
        rc = self._lib.cudasw_submit(self._h, s.ctypes.data, len(s), r.ctypes.data, len(r), ctypes.byref(t))
        if rc != 0:
            raise CudaSwError(f"cudasw_submit failed ({rc})")
        return t.value

    def flush(self) -> None:
        rc = self._lib.cudasw_flush(self._h)
        if rc != 0:
            raise CudaSwError(f"cudasw_flush failed ({rc}): {self._err()}")

    def outstanding(self) -> int:
        return int(self._lib.cudasw_outstanding(self._h))

    def processed(self) -> int:
        return int(self._lib.cudasw_processed(self._h))

    def start_dispatcher(self, max_pairs: int = 65536, max_delay_ms: float = 2.0) -> None:
        """One native thread turns the submit stream into GPU batches (size- or age-triggered flush)."""
        rc = self._lib.cudasw_dispatcher_start(self._h, int(max_pairs), int(max_delay_ms * 1000))
        if rc != 0:
            raise CudaSwError(f"cudasw_dispatcher_start failed ({rc})")

    def stop_dispatcher(self) -> None:
        rc = self._lib.cudasw_dispatcher_stop(self._h)
        if rc != 0:
            raise CudaSwError(f"cudasw_dispatcher_stop failed ({rc}): {self._err()}")

    def wait(self, ticket: int, timeout_ms: int = -1) -> Alignment | None:
        """Blocking getAlignment for one ticket (None on timeout). Releases the GIL while waiting."""
        res = np.zeros(1, dtype=RESULT_DTYPE)
        cap = 256
        while True:
            cig = np.zeros(cap, dtype=np.uint32)
            need = ctypes.c_int32(0)
            rc = self._lib.cudasw_wait(self._h, ticket, timeout_ms, res.ctypes.data, cig.ctypes.data, cap, ctypes.byref(need))
            if rc == CUDASW_ENOMEM and need.value > cap:
                cap = need.value
                continue
            break
        if rc == 0:
            return None
        if rc < 0:
            raise KeyError(f"unknown ticket {ticket}")
        r = res[0]
        if r["status"] != 0:
            raise CudaSwError(f"pair failed with status {int(r['status'])}")
        return Alignment(int(r["ref_begin"]), cigar_string(cig[: need.value]), int(r["score"]))

    def poll(self, ticket: int) -> Alignment | None:
        res = np.zeros(1, dtype=RESULT_DTYPE)
        cap = 256
        while True:
            cig = np.zeros(cap, dtype=np.uint32)
            need = ctypes.c_int32(0)
            rc = self._lib.cudasw_poll(self._h, ticket, res.ctypes.data, cig.ctypes.data, cap, ctypes.byref(need))
            if rc == CUDASW_ENOMEM and need.value > cap:
                cap = need.value
                continue
            break
        if rc == 0:
            return None
        if rc < 0:
            raise KeyError(f"unknown ticket {ticket}")
        r = res[0]
        if r["status"] != 0:
            raise CudaSwError(f"pair failed with status {int(r['status'])}")
        return Alignment(int(r["ref_begin"]), cigar_string(cig[: need.value]), int(r["score"]))


class MultiGpuAligner(Aligner):
    """Several B200s behind one in-process handle (cudasw_multi_*): contiguous cost-balanced
    slices, one host thread and engine per GPU, results at the original ordinals, no collective."""

    def __init__(self, match=1, mismatch=-4, ambiguous=-4, gapOpen=6, gapExtend=1, devices: Sequence[int] | None = None):
        lib = load_library()
        self._lib = lib
        sc = _Scoring(match, mismatch, ambiguous, gapOpen, gapExtend)
        h = ctypes.c_void_p()
        devs = np.asarray(list(devices), dtype=np.int32) if devices is not None else None
        rc = lib.cudasw_multi_create(ctypes.byref(sc), devs.ctypes.data if devs is not None else None,
                                     len(devs) if devs is not None else 0, ctypes.byref(h))
        if rc != 0:
            msg = (lib.cudasw_multi_last_error(None) or b"").decode()
            if rc == CUDASW_EINVAL:
                raise ValueError(msg)
            raise CudaSwError(f"cudasw_multi_create failed ({rc}): {msg}")
        self._h = h
        self.n_devices = lib.cudasw_multi_device_count(h)

    def close(self):
        if getattr(self, "_h", None):
            self._lib.cudasw_multi_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _err(self) -> str:
        return (self._lib.cudasw_multi_last_error(self._h) or b"").decode()

    def _call(self, fn, name, n, args, res, cig):
        res = np.zeros(n, dtype=RESULT_DTYPE) if res is None else res
        cap = max(64, 24 * n) if cig is None else len(cig)
        while True:
            if cig is None or len(cig) < cap:
                cig = np.zeros(cap, dtype=np.uint32)
            used = ctypes.c_int64(0)
            rc = fn(self._h, n, *args, res.ctypes.data, cig.ctypes.data, cap, ctypes.byref(used))
            if rc == CUDASW_ENOMEM and used.value > cap:
                cap = int(used.value) + 16
                continue
            if rc != 0:
                raise CudaSwError(f"{name} failed ({rc}): {self._err()}")
            return res, cig  # not compact: cigar_off is absolute into `cig`

    def align_blobs(self, seq_blob, seq_off, ref_blob, ref_off, res=None, cig=None):
        return self._call(self._lib.cudasw_multi_align_batch, "cudasw_multi_align_batch", len(seq_off) - 1,
                          (seq_blob.ctypes.data, seq_off.ctypes.data, ref_blob.ctypes.data, ref_off.ctypes.data), res, cig)

    def load_reference(self, contigs: Sequence[bytes]) -> None:
        blob, off = pack(contigs)
        rc = self._lib.cudasw_multi_reference_load(self._h, len(contigs), blob.ctypes.data, off.ctypes.data)
        if rc != 0:
            raise (ValueError if rc == CUDASW_EINVAL else CudaSwError)(f"cudasw_multi_reference_load failed ({rc}): {self._err()}")

    def align_blobs_windows(self, seq_blob, seq_off, win, win_off, res=None, cig=None):
        assert win.dtype == WINDOW_DTYPE
        return self._call(self._lib.cudasw_multi_align_batch_windows, "cudasw_multi_align_batch_windows", len(seq_off) - 1,
                          (seq_blob.ctypes.data, seq_off.ctypes.data, win.ctypes.data, win_off.ctypes.data), res, cig)

    _packed_fn = "cudasw_multi_align_batch_packed"
    align_blobs_packed = CudaAligner.align_blobs_packed  # same marshalling; the CIGAR blob is not compact here

    _homology_fn = "cudasw_multi_breakpoint_homology"
    breakpoint_homology = CudaAligner.breakpoint_homology   # same marshalling, the multi-GPU entry point

    def align_batch(self, seqs: Sequence[bytes], refs: Sequence[bytes]) -> list[Alignment]:
        if len(seqs) != len(refs):
            raise ValueError("seqs and refs differ in length")
        for s, r in zip(seqs, refs):
            if not s:
                raise ValueError("seq must be non-zero size")
            if not r:
                raise ValueError("ref must be non-zero size")
        sb, so = pack(seqs)
        rb, ro = pack(refs)
        res, cig = self.align_blobs(sb, so, rb, ro)
        out = []
        for r in res:
            if r["status"] != 0:
                raise CudaSwError(f"pair failed with status {int(r['status'])}")
            out.append(Alignment(int(r["ref_begin"]), cigar_string(cig[r["cigar_off"]: r["cigar_off"] + r["cigar_len"]]), int(r["score"])))
        return out

    def align_smith_waterman(self, seq: bytes, ref: bytes) -> Alignment:
        return self.align_batch([seq], [ref])[0]

    def slice_stats(self) -> list[dict]:
        out = []
        for d in range(self.n_devices):
            st, f, c = Stats(), ctypes.c_int64(0), ctypes.c_int64(0)
            self._lib.cudasw_multi_get_stats(self._h, d, ctypes.byref(st), ctypes.byref(f), ctypes.byref(c))
            out.append({"first_pair": f.value, "pairs": c.value, **st.as_dict()})
        return out


class BatchingAligner(Aligner):
    """Caller-side batch-collect layer, modelled on BwaStreamingAligner's
    size-triggered flush (alignment/BwaStreamingAligner.java:79-127) and the StreamingAligner
    contract (alignment/StreamingAligner.java:10-44): asyncAlign queues, flush dispatches one GPU
    batch, getAlignment returns results in submission order."""

    def __init__(self, aligner: CudaAligner, batch_size: int = 65536):
        self.aligner = aligner
        self.batch_size = batch_size
        self._tickets: list[int] = []
        self._lock = threading.Lock()

    def asyncAlign(self, seq: bytes, ref: bytes) -> int:
        with self._lock:
            t = self.aligner.submit(seq, ref)
            self._tickets.append(t)
            if self.aligner.outstanding() >= self.batch_size:
                self.aligner.flush()
            return t

    def flush(self) -> None:
        with self._lock:
            self.aligner.flush()

    def outstandingAlignmentRecord(self) -> int:
        return self.aligner.outstanding()

    def processedAlignmentRecords(self) -> int:
        return self.aligner.processed()

    def getAlignment(self) -> Alignment:
        with self._lock:
            if not self._tickets:
                raise RuntimeError("no alignment record is available")  # IllegalStateException
            a = self.aligner.poll(self._tickets[0])
            if a is None:
                raise RuntimeError("no alignment record is available; call flush()")
            self._tickets.pop(0)
            return a

    def align_smith_waterman(self, seq: bytes, ref: bytes) -> Alignment:
        t = self.asyncAlign(seq, ref)
        with self._lock:
            self.aligner.flush()
            a = self.aligner.poll(t)
            self._tickets.remove(t)
        return a


class AlignerFactory:
    """alignment/AlignerFactory.java:108-120: one shared default instance (1,-4,-4,6,1 at :90);
    create(...) with explicit scoring builds a new aligner. No fallback chain here: the reference
    falls back SSW -> JAligner on the GRIDSS side, this package only provides the CUDA backend."""

    _default: CudaAligner | None = None
    _lock = threading.Lock()

    @classmethod
    def create(cls, match=None, mismatch=None, ambiguous=None, gapOpen=None, gapExtend=None, device: int = 0) -> CudaAligner:
        if match is None:
            with cls._lock:
                if cls._default is None:
                    cls._default = CudaAligner(1, -4, -4, 6, 1, device=device)
                return cls._default
        return CudaAligner(match, mismatch, ambiguous, gapOpen, gapExtend, device=device)
