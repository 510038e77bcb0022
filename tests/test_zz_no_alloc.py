"""GPU: cudasw_score_batch_noalloc (include/cudasw.h) - the forward score pass with all of its scratch in a static
__device__ array - against the same fields of the existing entry point cudasw_align_batch, bit for bit, on calls made
back to back (the array keeps the previous call's contents). Collected last on purpose (file name).
The test allocates the device inputs/outputs (torch); the entry point itself allocates nothing."""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX_PAIRS, MAX_SEQ, MAX_REF = 4096, 1 << 20, 1 << 20  # CUDASW_NOALLOC_MAX_* of include/cudasw.h


def _pack(seqs):
    off = np.zeros(len(seqs) + 1, dtype=np.int64)
    off[1:] = np.cumsum([len(s) for s in seqs])
    blob = np.frombuffer(b"".join(seqs), dtype=np.uint8).copy() if off[-1] else np.zeros(1, dtype=np.uint8)
    return blob, off


def _noalloc(lib, scoring, sb, so, rb, ro, device=0):
    """One call through the C ABI with device pointers; returns the (n, 4) int32 records."""
    import torch

    from gridss_b200.aligner import _Scoring

    n = len(so) - 1
    dev = torch.device("cuda", device)
    d_sb = torch.from_numpy(np.ascontiguousarray(sb)).to(dev)
    d_so = torch.from_numpy(np.ascontiguousarray(so)).to(dev)
    d_rb = torch.from_numpy(np.ascontiguousarray(rb)).to(dev)
    d_ro = torch.from_numpy(np.ascontiguousarray(ro)).to(dev)
    d_out = torch.full((max(n, 1), 4), -77, dtype=torch.int32, device=dev)
    err = ctypes.create_string_buffer(512)
    sc = _Scoring(*scoring)
    # NULL stream = the legacy default stream, which is torch's current stream here: ordered after the copies above
    rc = lib.cudasw_score_batch_noalloc(ctypes.byref(sc), device, n, d_sb.data_ptr(), d_so.data_ptr(), int(so[-1]),
                                        d_rb.data_ptr(), d_ro.data_ptr(), int(ro[-1]), d_out.data_ptr(), None, err, 512)
    assert rc == 0, err.value
    torch.cuda.synchronize(dev)
    return d_out.cpu().numpy()[:n]


def _expected(al, sb, so, rb, ro):
    res, _ = al.align_blobs(sb, so, rb, ro)
    return np.stack([res["score"], res["ref_end"], res["read_end"], res["status"]], axis=1).astype(np.int32)


def _batches(synth):
    # 1: mixed lengths (50-2000 bp contigs vs 1-10 kb windows: queries above 1280 rows take strip memory), cut to the limits
    sb, so, rb, ro = synth.generate(3, 400)
    n = int(min(np.searchsorted(so, MAX_SEQ, side="right"), np.searchsorted(ro, MAX_REF, side="right"))) - 1
    assert n >= 100 and int((np.diff(so[: n + 1]) > 1280).sum()) >= 5
    first = (sb[: so[n]], so[: n + 1], rb[: ro[n]], ro[: n + 1])
    # 2: a smaller, different batch: lower case, N, non-ACGT bytes, an empty query, an empty window, identical pair, q = 1
    seqs, refs = synth.adversarial("mutated", 60, seed=11)
    s2, r2 = synth.adversarial("random", 40, seed=12, qmax=1500, wmax=3000)
    seqs += s2 + [b"", b"ACGT", b"acgtnACGTN-x", b"A", b"ACGTACGTAC" * 200]
    refs += r2 + [b"ACGT", b"", b"ACGTNacgtn*y", b"C", b"ACGTACGTAC" * 200]
    second = _pack(seqs) + _pack(refs)
    return first, second


def test_noalloc_matches_align_batch_on_consecutive_calls(aligner, synth):
    import gridss_b200 as g

    lib = g.load_library()
    first, second = _batches(synth)
    default = (1, -4, -4, 6, 1)
    for k, b in enumerate((first, second, first)):  # the third call finds the second call's leftovers in the array
        got = _noalloc(lib, default, *b)
        want = _expected(aligner, *b)
        bad = np.flatnonzero((got != want).any(axis=1))
        assert bad.size == 0, (k, bad[:8], got[bad[:4]], want[bad[:4]])
    assert (want[:, 3] == 0).all() and (want[:, 0] > 0).any()
    st = _expected(aligner, *second)[:, 3]
    assert (st == -1).sum() == 2  # the empty query and the empty window are per-pair CUDASW_EINVAL on both entry points

    # another scoring on the same array, right after
    other = (2, -3, -1, 5, 2)
    al2 = g.AlignerFactory.create(*other)
    got = _noalloc(lib, other, *second)
    want = _expected(al2, *second)
    assert np.array_equal(got, want)
    al2.close()


def test_noalloc_rejects_sizes_beyond_the_static_scratch():
    import torch

    import gridss_b200 as g

    lib = g.load_library()
    d = torch.zeros(64, dtype=torch.int64, device="cuda:0")
    err = ctypes.create_string_buffer(512)
    p = d.data_ptr()
    for n, sbytes, rbytes in ((MAX_PAIRS + 1, 16, 16), (1, MAX_SEQ + 1, 16), (1, 16, MAX_REF + 1)):
        rc = lib.cudasw_score_batch_noalloc(None, 0, n, p, p, sbytes, p, p, rbytes, p, None, err, 512)
        assert rc == -4 and b"exceed the static scratch" in err.value, (rc, err.value)
    assert lib.cudasw_score_batch_noalloc(None, 0, 0, None, None, 0, None, None, 0, None, None, err, 512) == 0
