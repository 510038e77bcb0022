#!/usr/bin/env python
"""C4 of BASELINE.json: the WGS-scale synthetic workload (~50 M pairs, 70 % C1-shape / 30 % C2-shape,
seed 4004), pair-sharded over the GPUs of one box with no collective on the data path.

  python tools/wgs_run.py --total 50000000                      one GPU
  torchrun --nproc-per-node N ... tools/wgs_run.py --total ...  one rank per GPU (rank r owns a contiguous slice)
  python tools/wgs_run.py --multi --total ...                   one process, cudasw_multi_* over every visible GPU

Every mode prints one JSON line with the same order-independent result checksum (gridss_b200/shard.py), so
1/2/4/8-GPU runs can be compared for equality. Pairs are generated on the host from the counter-based stream while
the previous chunk is on the GPU; alignments/s is reported both for the whole run and for the time spent inside the
aligner calls (host generation is not part of the hot path).
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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--total", type=int, default=50_000_000)
    ap.add_argument("--chunk", type=int, default=1_000_000, help="pairs per aligner call per rank")
    ap.add_argument("--cfg", type=int, default=4)
    ap.add_argument("--multi", action="store_true", help="one process, every visible GPU behind cudasw_multi_*")
    ap.add_argument("--packed", action="store_true", help="ship the pairs 2 bits per base (cudasw_[multi_]align_batch_packed); packing time is reported apart")
    args = ap.parse_args()
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))

    import gridss_b200 as g
    from gridss_b200 import shard
    from gridss_b200.aligner import PinnedBuffer
    from tools import synth as S

    S.build()
    group = None
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("gloo")  # host-side checksum reduction only; alignment data never crosses ranks
    if args.multi:
        al = g.MultiGpuAligner()
        n_gpus = al.n_devices
    else:
        al = g.CudaAligner(device=local)
        n_gpus = world
    threads = max(1, (os.cpu_count() or 1) // max(world, 1))

    # two pinned input buffer sets (the generator fills one while the other is on the GPU) + pinned outputs
    lens = S.lengths(args.cfg, min(args.chunk, max(1, args.total)))
    cap_q = int(lens[0][-1] * 1.3) + (1 << 20)
    cap_w = int(lens[1][-1] * 1.3) + (1 << 20)
    bufs = [(PinnedBuffer(cap_q), PinnedBuffer(cap_w)) for _ in range(2)]
    pin_res = PinnedBuffer(args.chunk * g.RESULT_DTYPE.itemsize)
    pin_cig = PinnedBuffer(4 * 24 * args.chunk)
    res_p, cig_p = pin_res.array.view(g.RESULT_DTYPE), pin_cig.array.view(np.uint32)
    flip = [0]

    def generate(first, n):
        b = bufs[flip[0]]
        flip[0] ^= 1
        return S.generate(args.cfg, n, first=first, threads=threads, out=(b[0].array, b[1].array))

    pack_s = [0.0]
    if args.packed:
        w2q, wnq = g.packed_words(cap_q)
        w2w, wnw = g.packed_words(cap_w)
        pk = [PinnedBuffer(4 * w2q), PinnedBuffer(4 * wnq), PinnedBuffer(4 * w2w), PinnedBuffer(4 * wnw)]
        q2, qn, w2, wn = (b.array.view(np.uint32) for b in pk)

    def align(sb, so, rb, ro):
        n = len(so) - 1
        if not args.packed:
            return al.align_blobs(sb, so, rb, ro, res=res_p[:n], cig=cig_p)
        t = time.perf_counter()
        _, _, aq = g.pack_bases(sb, q2, qn, n=int(so[-1]), threads=threads)
        _, _, aw = g.pack_bases(rb, w2, wn, n=int(ro[-1]), threads=threads)
        pack_s[0] += time.perf_counter() - t
        return al.align_blobs_packed(q2, qn if aq else None, so, w2, wn if aw else None, ro, res=res_p[:n], cig=cig_p)

    t0 = time.perf_counter()
    cs, st = shard.run_sharded(args.total, args.chunk, rank, world, generate, align)
    wall = time.perf_counter() - t0
    tot = shard.allreduce_checksum(cs, group)
    if world > 1:
        import torch
        import torch.distributed as dist

        t = torch.tensor([wall, st["align_s"], st["wait_s"]], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        wall, align_s, wait_s = [float(x) for x in t.tolist()]
        c = torch.tensor([st["pairs"], st["cells"], st["bad_status"]], dtype=torch.float64)
        dist.all_reduce(c, op=dist.ReduceOp.SUM)
        pairs, cells, bad = [float(x) for x in c.tolist()]
    else:
        align_s, wait_s, pairs, cells, bad = st["align_s"], st["wait_s"], st["pairs"], st["cells"], st["bad_status"]
    if rank == 0:
        print(json.dumps({
            "workload": f"C{args.cfg} (seed {S.SEEDS[args.cfg]}), {args.total} pairs, chunks of {args.chunk} per rank",
            "mode": "multi-handle (one process)" if args.multi else f"one rank per GPU x{world}",
            "n_gpus": n_gpus, "pairs": int(pairs), "bad_status": int(bad),
            "checksum": {"sum": f"0x{tot.sum:016x}", "xor": f"0x{tot.xor:016x}", "count": tot.count},
            "entry_point": "packed (2 bits per base)" if args.packed else "ASCII", "host_pack_s_rank0": round(pack_s[0], 3),
            "wall_s": round(wall, 3), "align_s_max_rank": round(align_s - (pack_s[0] if args.packed else 0.0), 3), "generator_wait_s_max_rank": round(wait_s, 3),
            "alignments_per_s_wall": pairs / wall, "alignments_per_s_in_aligner": pairs / max(align_s - (pack_s[0] if args.packed else 0.0), 1e-9),
            "gcups_in_aligner": cells / max(align_s - (pack_s[0] if args.packed else 0.0), 1e-9) / 1e9, "host_threads_per_rank": threads,
        }), flush=True)
    if world > 1:
        import torch.distributed as dist

        dist.barrier()
        dist.destroy_process_group()
    al.close()


if __name__ == "__main__":
    main()
