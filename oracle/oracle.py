"""ctypes front-end to the CPU oracles.  TEST / BENCH INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py (cpu_baseline / --impl reference) may import this
module; the product package gridss_b200/ never does.

Two checkers with one batch interface:
  * ``reference``: the reference's own shipped libsswjni.so (staged into oracle/_ref/ by
    oracle/make_ref.sh) driven by oracle/ref_driver.c exactly as SswJniAligner drives it
    (SswJniAligner.java:42-61).
  * ``model``: oracle/ssw_model.c, our scalar restatement, pinned against the
    reference binary and the reference's golden vectors by tests/test_oracle.py.
"""
from __future__ import annotations

import ctypes
import os
import subprocess
from dataclasses import dataclass

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(HERE, "_ref")

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

DEFAULT_SCORING = (1, -4, -4, 6, 1)  # AlignerFactory.java:90
_OPS = "MIDNSHP=X"


def build(force: bool = False) -> None:
    """Compile the model + the dlopen driver and stage the reference binary (make_ref.sh)."""
    need = force or not os.path.exists(os.path.join(HERE, "libssw_model.so")) or not os.path.exists(
        os.path.join(REF_DIR, "libref_driver.so")
    )
    if need:
        subprocess.check_call(["sh", os.path.join(HERE, "make_ref.sh")], stdout=subprocess.DEVNULL)


def pack(seqs) -> tuple[np.ndarray, np.ndarray]:
    """list of bytes -> (uint8 blob, int64 offsets[n+1])."""
    off = np.zeros(len(seqs) + 1, dtype=np.int64)
    if len(seqs):
        off[1:] = np.cumsum([len(s) for s in seqs])
    blob = np.frombuffer(b"".join(bytes(s) for s in seqs), dtype=np.uint8).copy()
    if blob.size == 0:
        blob = np.zeros(1, dtype=np.uint8)
    return blob, off


def cigar_string(words) -> str:
    return "".join(f"{int(w) >> 4}{_OPS[int(w) & 15]}" for w in words)


@dataclass
class BatchResult:
    res: np.ndarray  # RESULT_DTYPE[n]
    cig: np.ndarray  # uint32 blob

    def cigar(self, i: int) -> str:
        r = self.res[i]
        return cigar_string(self.cig[r["cigar_off"] : r["cigar_off"] + r["cigar_len"]])

    def tuple(self, i: int):
        r = self.res[i]
        return int(r["score"]), int(r["ref_begin"]), self.cigar(i)


class _Lib:
    def __init__(self, path: str, fn: str):
        self.lib = ctypes.CDLL(path)
        self.fn = getattr(self.lib, fn)
        self.fn.restype = ctypes.c_int
        self.fn.argtypes = [
            ctypes.c_int64,
            ctypes.c_void_p,
            ctypes.c_void_p,
            ctypes.c_void_p,
            ctypes.c_void_p,
            ctypes.c_int,
            ctypes.c_int,
            ctypes.c_int,
            ctypes.c_int,
            ctypes.c_int,
            ctypes.c_void_p,
            ctypes.c_void_p,
            ctypes.c_int64,
            ctypes.c_int,
        ]

    def run(self, seq_blob, seq_off, ref_blob, ref_off, scoring, threads, stride):
        n = len(seq_off) - 1
        res = np.zeros(n, dtype=RESULT_DTYPE)
        cig = np.zeros(max(1, n * stride), dtype=np.uint32)
        rc = self.fn(
            n,
            seq_blob.ctypes.data,
            seq_off.ctypes.data,
            ref_blob.ctypes.data,
            ref_off.ctypes.data,
            *[int(x) for x in scoring],
            res.ctypes.data,
            cig.ctypes.data,
            stride,
            threads,
        )
        if rc != 0:
            raise RuntimeError(f"oracle batch call failed rc={rc}")
        return BatchResult(res, cig)


_model = None
_ref = None


def have_reference() -> bool:
    return os.path.exists(os.path.join(REF_DIR, "libsswjni.so")) and os.path.exists(
        os.path.join(REF_DIR, "libref_driver.so")
    )


def _get_model() -> _Lib:
    global _model
    if _model is None:
        build()
        _model = _Lib(os.path.join(HERE, "libssw_model.so"), "model_align_batch")
    return _model


def _get_ref() -> _Lib:
    global _ref
    if _ref is None:
        build()
        if not have_reference():
            raise RuntimeError("oracle/_ref/libsswjni.so not staged (run oracle/make_ref.sh)")
        lib = _Lib(os.path.join(REF_DIR, "libref_driver.so"), "ref_align_batch")
        lib.lib.ref_load.argtypes = [ctypes.c_char_p]
        lib.lib.ref_load.restype = ctypes.c_int
        rc = lib.lib.ref_load(os.path.join(REF_DIR, "libsswjni.so").encode())
        if rc != 0:
            raise RuntimeError(f"ref_load failed rc={rc}")
        _ref = lib
    return _ref


def _stride_for(seq_off, ref_off, stride):
    if stride:
        return stride
    q = int(np.max(np.diff(seq_off))) if len(seq_off) > 1 else 1
    return int(min(2 * q + 8, 4096))


def model_align_blobs(seq_blob, seq_off, ref_blob, ref_off, scoring=DEFAULT_SCORING, threads=1, stride=0):
    return _get_model().run(seq_blob, seq_off, ref_blob, ref_off, scoring, threads, _stride_for(seq_off, ref_off, stride))


def ref_align_blobs(seq_blob, seq_off, ref_blob, ref_off, scoring=DEFAULT_SCORING, threads=1, stride=0):
    return _get_ref().run(seq_blob, seq_off, ref_blob, ref_off, scoring, threads, _stride_for(seq_off, ref_off, stride))


def model_align(seqs, refs, scoring=DEFAULT_SCORING, threads=1) -> BatchResult:
    sb, so = pack(seqs)
    rb, ro = pack(refs)
    return model_align_blobs(sb, so, rb, ro, scoring, threads)


def ref_align(seqs, refs, scoring=DEFAULT_SCORING, threads=1) -> BatchResult:
    sb, so = pack(seqs)
    rb, ro = pack(refs)
    return ref_align_blobs(sb, so, rb, ro, scoring, threads)
