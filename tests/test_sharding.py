"""CPU: host-side logic of the pair-partitioned multi-GPU path (gridss_b200/shard.py), including a
world_size-2 gloo run in which the oracle stands in for the GPU engine (tests may use the oracle as the checker;
the product never does)."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_rank_slices_tile_the_stream():
    from gridss_b200.shard import rank_slice

    for total in (0, 1, 7, 8, 9, 1000003):
        for world in (1, 2, 3, 8):
            sl = [rank_slice(total, r, world) for r in range(world)]
            assert sl[0][0] == 0 and sum(c for _, c in sl) == total
            assert all(sl[r][0] + sl[r][1] == sl[r + 1][0] for r in range(world - 1))
            assert max(c for _, c in sl) - min(c for _, c in sl) <= 1
    with pytest.raises(ValueError):
        rank_slice(10, 2, 2)


def test_balance_cuts_are_cost_balanced(synth):
    from gridss_b200.shard import balance_cuts

    sb, so, rb, ro = synth.generate(3, 20000)
    cost = np.diff(so).astype(np.float64) * np.diff(ro) + 20000.0
    for parts in (2, 4, 8):
        cuts = balance_cuts(so, ro, parts)
        assert cuts[0] == 0 and cuts[-1] == 20000 and np.all(np.diff(cuts) > 0)
        per = [cost[cuts[d]:cuts[d + 1]].sum() for d in range(parts)]
        assert max(per) / min(per) < 1.05


def test_checksum_is_order_and_layout_independent(oracle, synth):
    from gridss_b200.shard import Checksum, pair_digests

    sb, so, rb, ro = synth.generate(4, 600)
    m = oracle.model_align_blobs(sb, so, rb, ro, threads=4)
    whole = Checksum().add(pair_digests(m.res, m.cig, 0))
    # two halves computed separately (different cigar blobs / offsets), merged in the other order
    a = oracle.model_align_blobs(sb[: so[250]], so[:251], rb[: ro[250]], ro[:251], threads=2)
    b = oracle.model_align_blobs(sb[so[250]:], so[250:] - so[250], rb[ro[250]:], ro[250:] - ro[250], threads=2)
    parts = Checksum().add(pair_digests(b.res, b.cig, 250)).merge(Checksum().add(pair_digests(a.res, a.cig, 0)))
    assert parts == whole and whole.count == 600
    # any change to a result, a CIGAR word or a pair's ordinal is seen
    r2 = m.res.copy()
    r2["ref_begin"][17] += 1
    assert Checksum().add(pair_digests(r2, m.cig, 0)) != whole
    c2 = m.cig.copy()
    c2[m.res["cigar_off"][5]] ^= 16
    assert Checksum().add(pair_digests(m.res, c2, 0)) != whole
    assert Checksum().add(pair_digests(m.res, m.cig, 1)) != whole
    r3 = m.res.copy()
    r3[[3, 4]] = r3[[4, 3]]
    assert Checksum().add(pair_digests(r3, m.cig, 0)) != whole


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _rank_main(rank, world, port, total, chunk, q):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    import torch.distributed as dist

    from gridss_b200 import shard
    from oracle import oracle as O
    from tools import synth as S

    dist.init_process_group("gloo", rank=rank, world_size=world)

    def generate(first, n):
        return S.generate(4, n, first=first, threads=2)

    kept = []

    def align(sb, so, rb, ro):
        m = O.model_align_blobs(sb, so, rb, ro, threads=2)  # the checker stands in for the engine on CPU
        kept.append(m.res.copy())
        return m.res, m.cig

    cs, st = shard.run_sharded(total, chunk, rank, world, generate, align)
    tot = shard.allreduce_checksum(cs)
    first, count = shard.rank_slice(total, rank, world)
    res_local = np.concatenate(kept) if kept else np.zeros(0, dtype=O.RESULT_DTYPE)
    gathered = shard.gather_by_ordinal(res_local, first, total)
    if rank == 0:
        q.put((tot.as_tuple(), st["pairs"], gathered["score"].tolist(), gathered["ref_begin"].tolist()))
    dist.barrier()
    dist.destroy_process_group()


def test_world_size_2_gloo_matches_single_process(oracle, synth):
    import torch.multiprocessing as mp

    from gridss_b200 import shard

    total, chunk = 901, 200
    sb, so, rb, ro = synth.generate(4, total)
    m = oracle.model_align_blobs(sb, so, rb, ro, threads=4)
    want = shard.Checksum().add(shard.pair_digests(m.res, m.cig, 0))
    # single process, chunked
    cs1, st1 = shard.run_sharded(total, chunk, 0, 1, lambda f, n: synth.generate(4, n, first=f),
                                 lambda a, b, c, d: (lambda r: (r.res, r.cig))(oracle.model_align_blobs(a, b, c, d, threads=4)))
    assert cs1 == want and st1["pairs"] == total and st1["chunks"] == 5
    # two ranks over gloo
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_rank_main, args=(r, 2, port, total, chunk, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = q.get(timeout=240)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert got[0] == want.as_tuple()
    assert got[1] == 451  # rank 0's share
    assert got[2] == m.res["score"].tolist() and got[3] == m.res["ref_begin"].tolist()
