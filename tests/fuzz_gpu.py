#!/usr/bin/env python
"""Differential fuzz of the CUDA path against the reference's own shipped libsswjni.so (SURVEY.md §8c "Volume parity
plan"). Test infrastructure: run by hand on a GPU box, not collected by pytest.

  python tests/fuzz_gpu.py --seconds 180          # about that much reference-binary time on all host cores

Streams fresh slices of the counter-based C1-C4 workloads (new seeds every round) and the adversarial families
through cudasw_align_batch and through the checker, compares (score, ref_begin, ref_end, read_begin, read_end,
CIGAR words, status) per pair and prints one JSON summary line; any mismatch is listed and the exit code is 1.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
FIELDS = ["score", "ref_begin", "ref_end", "read_begin", "read_end", "cigar_len", "status"]


def compare(res, cig, m):
    bad = 0
    first = None
    same = np.ones(len(res), dtype=bool)
    for f in FIELDS:
        same &= res[f] == m.res[f]
    # CIGAR words: compare through the per-pair digests of the product's own checksum helper, then exactly where cheap
    from gridss_b200.shard import pair_digests

    same &= pair_digests(res, cig, 0) == pair_digests(m.res, m.cig, 0)
    undef = (res["status"] == -5) | (m.res["status"] == -5)
    same |= undef
    bad = int((~same).sum())
    if bad:
        first = int(np.flatnonzero(~same)[0])
    return bad, int(undef.sum()), first


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--seconds", type=float, default=180.0)
    ap.add_argument("--checker", default="reference", choices=["reference", "model"])
    ap.add_argument("--c4-sample", type=int, default=0, help="instead: parity on the first N pairs of the C4 stream itself (seed 4004)")
    ap.add_argument("--c5", type=int, default=0, help="instead: parity on the first N pairs of C5 (10 kb contigs vs 20 kb windows, seed 5005)")
    ap.add_argument("--c5-first", type=int, default=0, help="with --c5: start at this pair of the stream (pairs [first, N))")
    args = ap.parse_args()
    import gridss_b200 as g
    from gridss_b200.aligner import pack
    from oracle import oracle as O
    from tools import synth as S

    O.build()
    S.build()
    cores = os.cpu_count() or 1
    use_ref = args.checker == "reference" and O.have_reference()
    check = O.ref_align_blobs if use_ref else O.model_align_blobs
    al = g.CudaAligner()
    calls = {"ascii": 0, "packed": 0}

    class _Alt:
        """every other batch goes through the packed host entry point (cudasw_align_batch_packed)"""

        def align_blobs(self, sb, so, rb, ro):
            if (calls["ascii"] + calls["packed"]) % 2 == 0:
                calls["ascii"] += 1
                return al_raw.align_blobs(sb, so, rb, ro)
            calls["packed"] += 1
            s2, sn, _ = g.pack_bases(sb, n=int(so[-1]), threads=8)
            r2, rn, _ = g.pack_bases(rb, n=int(ro[-1]), threads=8)
            return al_raw.align_blobs_packed(s2, sn, so, r2, rn, ro)

        def close(self):
            al_raw.close()

    al_raw, al = al, None
    al = _Alt()
    plan = [(1, 200000), (2, 60000), (4, 100000), (3, 30000)]
    kinds = ["random", "mutated", "two_letter", "homopolymer", "dinuc", "tandem", "subst_block", "with_n", "zero", "tiny", "q_gt_w", "indel_heavy"]
    tot = {"pairs": 0, "cells": 0.0, "mismatches": 0, "undefined_in_reference": 0, "by_source": {}}
    examples = []
    if args.c4_sample > 0:
        from gridss_b200.shard import Checksum, pair_digests

        cs_gpu, cs_ref = Checksum(), Checksum()
        step = 250000
        for first in range(0, args.c4_sample, step):
            n = min(step, args.c4_sample - first)
            sb, so, rb, ro = S.generate(4, n, first=first)
            res, cig = al.align_blobs(sb, so, rb, ro)
            m = check(sb, so, rb, ro, threads=cores)
            bad, undef, bad_first = compare(res, cig, m)
            cs_gpu.add(pair_digests(res, cig, first))
            cs_ref.add(pair_digests(m.res, m.cig, first))
            tot["pairs"] += n
            tot["cells"] += float((np.diff(so).astype(np.float64) * np.diff(ro)).sum())
            tot["mismatches"] += bad
            tot["undefined_in_reference"] += undef
        tot["by_source"] = {"C4 stream, seed 4004, pairs [0, %d)" % args.c4_sample: {"pairs": tot["pairs"], "mismatches": tot["mismatches"]}}
        tot["checksum_gpu"] = repr(cs_gpu)
        tot["checksum_reference"] = repr(cs_ref)
        tot["checker"] = ("libsswjni.so as shipped by the reference" if use_ref else "oracle/ssw_model.c") + f", {cores} host threads"
        print(json.dumps(tot), flush=True)
        al.close()
        return 1 if tot["mismatches"] or cs_gpu != cs_ref else 0
    if args.c5 > 0:
        # the long-contig case takes the reference's 16-bit rerun and its band doubling
        step = 250
        for first in range(args.c5_first, args.c5, step):
            n = min(step, args.c5 - first)
            sb, so, rb, ro = S.generate(5, n, first=first)
            res, cig = al.align_blobs(sb, so, rb, ro)
            m = check(sb, so, rb, ro, threads=cores)
            bad, undef, bad_first = compare(res, cig, m)
            tot["pairs"] += n
            tot["cells"] += float((np.diff(so).astype(np.float64) * np.diff(ro)).sum())
            tot["mismatches"] += bad
            tot["undefined_in_reference"] += undef
            if bad and len(examples) < 5:
                examples.append({"source": "C5", "pair": first + bad_first})
        tot["by_source"] = {"C5 stream, seed 5005, pairs [%d, %d)" % (args.c5_first, args.c5): {"pairs": tot["pairs"], "mismatches": tot["mismatches"]}}
        tot["checker"] = ("libsswjni.so as shipped by the reference" if use_ref else "oracle/ssw_model.c") + f", {cores} host threads"
        tot["examples"] = examples
        print(json.dumps(tot), flush=True)
        al.close()
        return 1 if tot["mismatches"] else 0
    t_end = time.time() + args.seconds
    rnd = 0
    while time.time() < t_end:
        rnd += 1
        for cfg, n in plan:
            if time.time() >= t_end:
                break
            seed = 100000 * cfg + rnd
            sb, so, rb, ro = S.generate(cfg, n, seed=seed)
            res, cig = al.align_blobs(sb, so, rb, ro)
            m = check(sb, so, rb, ro, threads=cores)
            bad, undef, first = compare(res, cig, m)
            key = f"C{cfg}"
            d = tot["by_source"].setdefault(key, {"pairs": 0, "mismatches": 0})
            d["pairs"] += n
            d["mismatches"] += bad
            tot["pairs"] += n
            tot["cells"] += float((np.diff(so).astype(np.float64) * np.diff(ro)).sum())
            tot["mismatches"] += bad
            tot["undefined_in_reference"] += undef
            if bad and len(examples) < 5:
                examples.append({"source": key, "seed": seed, "pair": first})
        for kind in kinds:
            if time.time() >= t_end:
                break
            seqs, refs = S.adversarial(kind, 1500, seed=1000 + rnd, qmax=300 + 100 * (rnd % 5), wmax=600 + 300 * (rnd % 4))
            sb, so = pack(seqs)
            rb, ro = pack(refs)
            res, cig = al.align_blobs(sb, so, rb, ro)
            m = check(sb, so, rb, ro, threads=cores)
            bad, undef, first = compare(res, cig, m)
            d = tot["by_source"].setdefault("adversarial:" + kind, {"pairs": 0, "mismatches": 0})
            d["pairs"] += len(seqs)
            d["mismatches"] += bad
            tot["pairs"] += len(seqs)
            tot["cells"] += float((np.diff(so).astype(np.float64) * np.diff(ro)).sum())
            tot["mismatches"] += bad
            tot["undefined_in_reference"] += undef
            if bad and len(examples) < 5:
                examples.append({"source": kind, "seed": 1000 + rnd, "pair": first})
    tot["checker"] = ("libsswjni.so as shipped by the reference" if use_ref else "oracle/ssw_model.c") + f", {cores} host threads"
    tot["rounds"] = rnd
    tot["entry_points"] = calls
    tot["examples"] = examples
    print(json.dumps(tot), flush=True)
    al.close()
    return 1 if tot["mismatches"] else 0


if __name__ == "__main__":
    sys.exit(main())
