"""Synthetic workloads of SURVEY.md §8d (ctypes over tools/synth.c) + adversarial sets for fuzzing.

Test/bench tooling: not part of the product path.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SEEDS = {1: 1001, 2: 2002, 3: 3003, 4: 4004, 5: 5005}
_lib = None


def build(force: bool = False) -> None:
    so = os.path.join(HERE, "libsynth.so")
    src = os.path.join(HERE, "synth.c")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["gcc", "-O2", "-fPIC", "-shared", "-pthread", "-o", so, src, "-lm"])


def _get():
    global _lib
    if _lib is None:
        build()
        _lib = ctypes.CDLL(os.path.join(HERE, "libsynth.so"))
        _lib.synth_lengths.argtypes = [ctypes.c_int, ctypes.c_uint64, ctypes.c_int64, ctypes.c_int64, ctypes.c_void_p, ctypes.c_void_p]
        _lib.synth_lengths.restype = None
        _lib.synth_fill.argtypes = [ctypes.c_int, ctypes.c_uint64, ctypes.c_int64, ctypes.c_int64, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int]
        _lib.synth_fill.restype = None
    return _lib


def generate(cfg: int, n: int, first: int = 0, seed: int | None = None, threads: int | None = None,
             out=None):
    """-> (seq_blob u8, seq_off i64[n+1], ref_blob u8, ref_off i64[n+1]) for pairs [first, first+n).

    `out` may supply preallocated (seq_blob, ref_blob) arrays (e.g. views of pinned memory)."""
    lib = _get()
    seed = SEEDS[cfg] if seed is None else seed
    threads = threads or min(64, os.cpu_count() or 1)
    so = np.zeros(n + 1, dtype=np.int64)
    ro = np.zeros(n + 1, dtype=np.int64)
    lib.synth_lengths(cfg, seed, first, n, so.ctypes.data, ro.ctypes.data)
    if out is None:
        sb = np.empty(max(1, int(so[-1])), dtype=np.uint8)
        rb = np.empty(max(1, int(ro[-1])), dtype=np.uint8)
    else:
        sb, rb = out
        assert sb.size >= so[-1] and rb.size >= ro[-1]
    lib.synth_fill(cfg, seed, first, n, sb.ctypes.data, so.ctypes.data, rb.ctypes.data, ro.ctypes.data, threads)
    return sb, so, rb, ro


# ---------------------------------------------------------------------------------------------
# adversarial sets (SURVEY.md §8c "Volume parity plan"): python, small
# ---------------------------------------------------------------------------------------------
def lengths(cfg: int, n: int, first: int = 0, seed: int | None = None):
    """-> (seq_off i64[n+1], ref_off i64[n+1]) without generating the bases (to size pinned buffers)."""
    lib = _get()
    seed = SEEDS[cfg] if seed is None else seed
    so = np.zeros(n + 1, dtype=np.int64)
    ro = np.zeros(n + 1, dtype=np.int64)
    lib.synth_lengths(cfg, seed, first, n, so.ctypes.data, ro.ctypes.data)
    return so, ro


def _rand(rng, n, alphabet=b"ACGT"):
    return bytes(rng.choice(np.frombuffer(alphabet, dtype=np.uint8), size=n).tolist())


def adversarial(kind: str, n: int, seed: int = 7, qmax: int = 200, wmax: int = 400):
    """-> (seqs, refs) lists of bytes."""
    rng = np.random.default_rng(seed)
    seqs, refs = [], []
    for _ in range(n):
        q = int(rng.integers(1, qmax + 1))
        w = int(rng.integers(1, wmax + 1))
        if kind == "random":
            s, r = _rand(rng, q), _rand(rng, w)
        elif kind == "mutated":
            r = _rand(rng, w)
            st = int(rng.integers(0, w))
            frag = bytearray(r[st : st + q])
            out = bytearray()
            for b in frag:
                u = rng.random()
                if u < 0.05:
                    out.append(int(rng.choice(list(b"ACGT"))))
                elif u < 0.08:
                    out.extend(_rand(rng, int(rng.integers(1, 8))))
                    out.append(b)
                elif u < 0.11:
                    continue
                else:
                    out.append(b)
            if rng.random() < 0.5:
                out = bytearray(_rand(rng, int(rng.integers(0, 30)))) + out
            if rng.random() < 0.5:
                out = out + bytearray(_rand(rng, int(rng.integers(0, 30))))
            s = bytes(out) or b"A"
        elif kind == "two_letter":
            s, r = _rand(rng, q, b"AC"), _rand(rng, w, b"AC")
        elif kind == "homopolymer":
            s = b"A" * q
            r = bytearray(b"A" * w)
            for k in range(int(rng.integers(0, 4))):
                r[int(rng.integers(0, w))] = ord("C")
            r = bytes(r)
        elif kind == "dinuc":
            unit = _rand(rng, int(rng.integers(2, 5)))
            s = (unit * (q // len(unit) + 1))[:q]
            r = (unit * (w // len(unit) + 1))[:w]
            if rng.random() < 0.7 and w > 4:
                r = bytearray(r)
                for k in range(int(rng.integers(1, 4))):
                    r[int(rng.integers(0, w))] = int(rng.choice(list(b"ACGT")))
                r = bytes(r)
        elif kind == "tandem":
            unit = _rand(rng, int(rng.integers(5, 40)))
            r = _rand(rng, w // 3) + unit * int(rng.integers(1, 5)) + _rand(rng, w // 3)
            s = (_rand(rng, int(rng.integers(0, 20))) + unit * int(rng.integers(1, 5)) + _rand(rng, int(rng.integers(0, 20))))
        elif kind == "subst_block":
            r = _rand(rng, max(w, 60))
            st = int(rng.integers(0, len(r) - 40))
            frag = bytearray(r[st : st + max(40, q)])
            bl = int(rng.integers(3, 14))
            p = int(rng.integers(10, len(frag) - bl - 5)) if len(frag) > bl + 20 else 5
            frag[p : p + bl] = _rand(rng, bl)
            s = bytes(frag)
        elif kind == "with_n":
            r = bytearray(_rand(rng, w, b"ACGTN"))
            s = bytearray(_rand(rng, q, b"ACGTNacgtnRYx."))
            s, r = bytes(s), bytes(r)
        elif kind == "zero":
            s, r = b"A" * q, b"C" * w
        elif kind == "tiny":
            s, r = _rand(rng, int(rng.integers(1, 4))), _rand(rng, int(rng.integers(1, 4)))
        elif kind == "q_gt_w":
            r = _rand(rng, max(1, w // 4))
            s = _rand(rng, int(rng.integers(0, 50))) + r[: len(r)] + _rand(rng, int(rng.integers(0, 50)))
        elif kind == "indel_heavy":
            r = _rand(rng, max(w, 120))
            st = int(rng.integers(0, len(r) - 100))
            frag = bytearray(r[st : st + 100 + q])
            out = bytearray()
            i = 0
            while i < len(frag):
                u = rng.random()
                if u < 0.02:
                    out.extend(_rand(rng, int(rng.integers(1, 25))))
                elif u < 0.04:
                    i += int(rng.integers(1, 25))
                    continue
                out.append(frag[i])
                i += 1
            s = bytes(out) or b"A"
        else:
            raise ValueError(kind)
        seqs.append(bytes(s))
        refs.append(bytes(r))
    return seqs, refs


ADVERSARIAL_KINDS = ["random", "mutated", "two_letter", "homopolymer", "dinuc", "tandem", "subst_block",
                     "with_n", "zero", "tiny", "q_gt_w", "indel_heavy"]
