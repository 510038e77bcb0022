"""Host-side logic of the pair-partitioned multi-GPU path: which pairs a rank owns,
an order-independent checksum of results, and the host-side gather by original ordinal.

Pairs are independent (`Aligner` is stateless, SswJniAligner.java:42-61), so ranks never exchange
alignment data: the only collectives here are the checksum reduction and the optional result
gather, both over host memory (gloo), never on the timed device path. The reference's own
scale-out is the same idea one level up (chunk striding, AssemblyEvidenceSource.java:94-95).
"""
from __future__ import annotations

import numpy as np

_M1 = np.uint64(0xBF58476D1CE4E5B9)
_M2 = np.uint64(0x94D049BB133111EB)
_G = np.uint64(0x9E3779B97F4A7C15)


def rank_slice(total_pairs: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous slice [first, first + count) of a stream of `total_pairs` owned by `rank`.
    Slices differ by at most one pair and tile the stream exactly."""
    if world <= 0 or not (0 <= rank < world) or total_pairs < 0:
        raise ValueError("bad rank/world/total")
    base, extra = divmod(total_pairs, world)
    first = rank * base + min(rank, extra)
    return first, base + (1 if rank < extra else 0)


def balance_cuts(seq_off: np.ndarray, ref_off: np.ndarray, parts: int, per_pair: float = 20000.0) -> np.ndarray:
    """Cut points (parts + 1) of contiguous slices with balanced cost sum(q * w + per_pair): the same rule
    cudasw_multi_align_batch applies (gridss_b200/csrc/multi.cpp)."""
    n = len(seq_off) - 1
    cost = np.maximum(np.diff(seq_off), 0).astype(np.float64) * np.maximum(np.diff(ref_off), 0) + per_pair
    pre = np.concatenate([[0.0], np.cumsum(cost)])
    cuts = [0]
    for d in range(1, parts):
        cuts.append(max(cuts[-1], int(np.searchsorted(pre, pre[-1] * d / parts, side="left"))))
    cuts.append(n)
    return np.asarray(cuts, dtype=np.int64)


def _mix(x: np.ndarray) -> np.ndarray:
    x = x.astype(np.uint64, copy=True)
    x ^= x >> np.uint64(30)
    x *= _M1
    x ^= x >> np.uint64(27)
    x *= _M2
    x ^= x >> np.uint64(31)
    return x


def pair_digests(res: np.ndarray, cig: np.ndarray, first_ordinal: int = 0) -> np.ndarray:
    """One 64-bit digest per pair over (global ordinal, status, score, begin/end coordinates, CIGAR words).
    Independent of where a pair's words sit in the blob."""
    n = len(res)
    with np.errstate(over="ignore"):
        h = _mix(np.arange(first_ordinal, first_ordinal + n, dtype=np.uint64) + _G)
        for f in ("status", "score", "ref_begin", "ref_end", "read_begin", "read_end", "cigar_len"):
            h = _mix(h ^ (res[f].astype(np.int64).astype(np.uint64) * _G))
        lens = res["cigar_len"].astype(np.int64)
        lens = np.where(res["status"] == 0, lens, 0)
        tot = int(lens.sum())
        if tot:
            starts = np.concatenate([[0], np.cumsum(lens)[:-1]])
            pos = np.arange(tot, dtype=np.int64) - np.repeat(starts, lens)
            idx = np.repeat(res["cigar_off"].astype(np.int64), lens) + pos
            w = _mix(cig[idx].astype(np.uint64) + (pos.astype(np.uint64) + np.uint64(1)) * _G)
            nz = lens > 0
            sums = np.add.reduceat(w, starts[nz])
            c = np.zeros(n, dtype=np.uint64)
            c[nz] = sums
            h = _mix(h ^ c)
    return h


class Checksum:
    """Order-independent accumulator: (sum mod 2^64, xor, count) of pair digests."""

    def __init__(self, s: int = 0, x: int = 0, n: int = 0):
        self.sum, self.xor, self.count = int(s) & (2**64 - 1), int(x), int(n)

    def add(self, digests: np.ndarray) -> "Checksum":
        with np.errstate(over="ignore"):
            self.sum = (self.sum + int(np.add.reduce(digests.astype(np.uint64), dtype=np.uint64))) & (2**64 - 1)
        self.xor ^= int(np.bitwise_xor.reduce(digests.astype(np.uint64))) if len(digests) else 0
        self.count += len(digests)
        return self

    def merge(self, other: "Checksum") -> "Checksum":
        self.sum = (self.sum + other.sum) & (2**64 - 1)
        self.xor ^= other.xor
        self.count += other.count
        return self

    def as_tuple(self):
        return (self.sum, self.xor, self.count)

    def __eq__(self, other):
        return isinstance(other, Checksum) and self.as_tuple() == other.as_tuple()

    def __repr__(self):
        return f"Checksum(sum=0x{self.sum:016x}, xor=0x{self.xor:016x}, count={self.count})"


def allreduce_checksum(local: Checksum, group=None) -> Checksum:
    """Combine the ranks' checksums (host tensors; works on a gloo group). Every rank gets the total."""
    import torch
    import torch.distributed as dist

    if not (dist.is_available() and dist.is_initialized()):
        return Checksum(*local.as_tuple())
    world = dist.get_world_size(group)
    # 64-bit values travel as two 32-bit halves in int64 lanes (no signed overflow in the transport)
    mine = torch.tensor([local.sum & 0xFFFFFFFF, local.sum >> 32, local.xor & 0xFFFFFFFF, local.xor >> 32, local.count], dtype=torch.int64)
    parts = [torch.zeros_like(mine) for _ in range(world)]
    dist.all_gather(parts, mine, group=group)
    out = Checksum()
    for p in parts:
        v = [int(x) for x in p.tolist()]
        out.merge(Checksum(v[0] | (v[1] << 32), v[2] | (v[3] << 32), v[4]))
    return out


def gather_by_ordinal(res_local: np.ndarray, first: int, total: int, group=None, dst: int = 0):
    """Host-side gather: rank `dst` receives the result records of all ranks placed at their original
    ordinals (cigar_off stays an offset into the owning rank's CIGAR blob)."""
    import torch
    import torch.distributed as dist

    if not (dist.is_available() and dist.is_initialized()):
        assert first == 0 and len(res_local) == total
        return res_local
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    counts = [rank_slice(total, r, world)[1] for r in range(world)]
    assert rank_slice(total, rank, world) == (first, len(res_local))
    width = res_local.dtype.itemsize
    mx = max(counts)
    buf = np.zeros((mx, width), dtype=np.uint8)
    buf[: len(res_local)] = np.ascontiguousarray(res_local).view(np.uint8).reshape(len(res_local), width)
    mine = torch.from_numpy(buf)
    parts = [torch.zeros_like(mine) for _ in range(world)] if rank == dst else None
    dist.gather(mine, parts, dst=dst, group=group)
    if rank != dst:
        return None
    out = np.zeros(total, dtype=res_local.dtype)
    for r in range(world):
        f, c = rank_slice(total, r, world)
        out[f: f + c] = parts[r].numpy()[:c].reshape(-1).view(res_local.dtype)
    return out


def run_sharded(total_pairs: int, chunk_pairs: int, rank: int, world: int, generate, align, prefetch: bool = True):
    """Drive one rank's share of a counter-based pair stream (BASELINE.json C4: WGS-scale volume).

    generate(first, n) -> (seq_blob, seq_off, ref_blob, ref_off) for pairs [first, first + n) of the stream;
    align(seq_blob, seq_off, ref_blob, ref_off) -> (results, cigar_blob).
    The next chunk is generated on a host thread while the current one is aligned. Returns the rank's
    Checksum and a dict of counters (pairs, cells, seconds in align, seconds waiting for the generator)."""
    import threading
    import time

    first, count = rank_slice(total_pairs, rank, world)
    chunks = [(first + o, min(chunk_pairs, count - o)) for o in range(0, count, max(1, chunk_pairs))]
    cs = Checksum()
    st = {"pairs": 0, "cells": 0.0, "align_s": 0.0, "wait_s": 0.0, "chunks": len(chunks), "bad_status": 0}
    nxt = {}

    def produce(k):
        nxt[k] = generate(*chunks[k])

    th = None
    if chunks:
        produce(0)
    for k, (f, n) in enumerate(chunks):
        if th is not None:
            t0 = time.perf_counter()
            th.join()
            st["wait_s"] += time.perf_counter() - t0
            th = None
        sb, so, rb, ro = nxt.pop(k)
        if k + 1 < len(chunks):
            if prefetch:
                th = threading.Thread(target=produce, args=(k + 1,), daemon=True)
                th.start()
            else:
                t0 = time.perf_counter()
                produce(k + 1)
                st["wait_s"] += time.perf_counter() - t0
        t0 = time.perf_counter()
        res, cig = align(sb, so, rb, ro)
        st["align_s"] += time.perf_counter() - t0
        cs.add(pair_digests(res[:n], cig, first_ordinal=f))
        st["pairs"] += n
        st["cells"] += float((np.diff(so).astype(np.float64) * np.diff(ro)).sum())
        st["bad_status"] += int((res["status"][:n] != 0).sum())
    return cs, st
