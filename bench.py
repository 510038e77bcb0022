#!/usr/bin/env python
"""bench.py - alignments/s and GCUPS of the batched Smith-Waterman hot path (BASELINE.json metric).

A step = one pass of the whole hot path (encode -> forward score pass -> reverse pass -> banded
traceback -> CIGARs on the host) over one batch of the C2 workload (BASELINE.json configs[1]:
synthetic 300-bp breakend contigs vs 4 kb windows, 1M pairs per GPU, SURVEY.md §8d seed 2002).

  value : whole-job alignments/s with the raw ASCII inputs already resident in HBM
          (cudasw_align_batch_device), device-timed with CUDA events on the library's stream.
  e2e   : the same through the reference-facing C-ABI call with HOST (pinned) buffers holding the
          shim's packed layout (cudasw_align_batch_packed: 2 bits per base + N mask, 3/8 of the ASCII
          bytes): H2D of every pair and D2H of every result inside the timed region. `e2e_ascii` is the
          same with raw ASCII host blobs (cudasw_align_batch), `e2e_windows` with windows named.
  roofline : the forward score-pass kernel against the INT32 issue peak.
  cpu_baseline : the reference's own shipped libsswjni.so on this box's host cores (bounded sample).

`--impl reference` times only that CPU reference arm. Multi-GPU: one rank per GPU under torchrun,
pairs sharded by rank (no collective on the data path), weak scaling.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

CFG = 2
WORKLOAD = "C2: synthetic 300-bp breakend assembly contigs vs 4 kb reference windows (seed 2002), default scoring 1/-4/-4/6/1"
Q_LEN, W_LEN = 300, 4000
INT_OPS_PER_CELL = 12


def env_int(name, default):
    try:
        return int(os.environ.get(name, default))
    except ValueError:
        return default


class ClockSampler(threading.Thread):
    """nvidia-smi clocks/throttle reasons during the timed region (B200_PROFILING.md clocks line)."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index = index
        self.samples = []
        self.stop_flag = threading.Event()

    def run(self):
        while not self.stop_flag.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.samples.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            self.stop_flag.wait(0.1)

    def summary(self):
        self.stop_flag.set()
        self.join(timeout=6)
        sm = [float(s[0]) for s in self.samples if s and s[0].replace(".", "").isdigit()]
        mx = [float(s[1]) for s in self.samples if len(s) > 1 and s[1].replace(".", "").isdigit()]
        reasons = set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for s in self.samples:
            for k, nm in enumerate(names):
                if len(s) > 3 + k and s[3 + k].lower().startswith("active"):
                    reasons.add(nm)
        pw = [float(s[2]) for s in self.samples if len(s) > 2 and s[2].replace(".", "").isdigit()]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "reasons": sorted(reasons), "samples": len(self.samples)}


def bind_to_gpu_numa_node(torch, local):
    """One process per GPU: run this rank's host threads (and first-touch its pinned staging buffers) on the NUMA node
    the GPU hangs off, so that H2D copies do not cross the socket interconnect. Best effort; returns the node or None."""
    try:
        p = torch.cuda.get_device_properties(local)
        dev = "%04x:%02x:%02x.0" % (p.pci_domain_id, p.pci_bus_id, p.pci_device_id)
        base = "/sys/bus/pci/devices/" + dev
        node = int(open(base + "/numa_node").read())
        cpus = set()
        for part in open(base + "/local_cpulist").read().strip().split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        cpus &= os.sched_getaffinity(0)
        if node < 0 or len(cpus) < 2:
            return None
        os.sched_setaffinity(0, cpus)
        return node
    except Exception:
        return None


def reference_arm(steps: int, warmup: int, n_gpus: int, budget_s: float = 150.0):
    """The reference's own CPU implementation of the path on this box's host cores."""
    from oracle import oracle as O
    from tools import synth as S

    rank = env_int("RANK", 0)
    if rank != 0:
        return None
    O.build()
    S.build()
    cores = os.cpu_count() or 1
    kind = "reference" if O.have_reference() else "port"
    run = O.ref_align_blobs if kind == "reference" else O.model_align_blobs
    # calibrate on a small slice, then size each step to fit the budget
    sb, so, rb, ro = S.generate(CFG, 4 * cores)
    t = time.perf_counter()
    run(sb, so, rb, ro, threads=cores)
    rate = 4 * cores / max(time.perf_counter() - t, 1e-6)
    per_step = max(cores, int(rate * budget_s / max(steps + warmup, 1)))
    per_step = min(per_step, 200000)
    sb, so, rb, ro = S.generate(CFG, per_step)
    for _ in range(warmup):
        run(sb, so, rb, ro, threads=cores)
    t0 = time.perf_counter()
    for _ in range(steps):
        r = run(sb, so, rb, ro, threads=cores)
    dt = time.perf_counter() - t0
    aps = per_step * steps / dt
    cells = float((np.diff(so) * np.diff(ro)).sum()) * steps
    sample = f"{per_step} pairs of the C2 workload per step, {cores} host threads, libsswjni.so as shipped (-O0 SSE2)" if kind == "reference" \
        else f"{per_step} pairs of the C2 workload per step, {cores} host threads, oracle/ssw_model.c -O3"
    line = {
        "impl": "reference", "metric": "alignments/sec", "value": aps, "unit": "alignments/s", "n_gpus": n_gpus,
        "steps": steps, "warmup": warmup, "ms_per_step": dt / steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8/i16 (SSE2 striped) + i32 traceback", "data": "synthetic",
        "config": {"workload": WORKLOAD, "pairs_per_step": per_step, "query_len": Q_LEN, "window_len": W_LEN},
        "gcups": cells / dt / 1e9,
        "cpu_baseline": {"value": aps, "unit": "alignments/s", "cores": cores, "kind": kind, "sample": sample, "gcups": cells / dt / 1e9},
        "e2e": {"value": aps, "unit": "alignments/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    return line


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="cudasw", choices=["cudasw", "reference"])
    ap.add_argument("--pairs", type=int, default=1_000_000, help="pairs per GPU per step (C2 is quoted at 1M)")
    ap.add_argument("--cpu-baseline-seconds", type=float, default=15.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-windows", action="store_true", help="skip the device-resident-reference e2e variant")
    ap.add_argument("--no-homology", action="store_true", help="skip the caller-level line (cudasw_breakpoint_homology)")
    ap.add_argument("--homology-breakpoints", type=int, default=500000)
    ap.add_argument("--no-worst-case", action="store_true", help="skip the zero-quiet-steps forward line")
    ap.add_argument("--no-int-peak", action="store_true", help="do not run tools/int_peak (keeps an ncu launch list to the library's own kernels)")
    ap.add_argument("--worst-case-pairs", type=int, default=1_000_000)
    args = ap.parse_args()
    rank, world, local = env_int("RANK", 0), env_int("WORLD_SIZE", 1), env_int("LOCAL_RANK", 0)

    if args.impl == "reference":
        line = reference_arm(args.steps, max(args.warmup, 0), args.gpus)
        if line is not None:
            print(json.dumps(line), flush=True)
        return 0

    warmup = max(args.warmup, 3)
    import torch
    import torch.distributed as dist

    import gridss_b200 as g
    from gridss_b200.aligner import PinnedBuffer
    from tools import synth as S

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (gridss_b200 has no CPU fallback); use --impl reference for the CPU arm")
    torch.cuda.set_device(local)
    saved_stdout = -1
    if world > 1:
        # stdout carries exactly one JSON line: whatever native libraries print there meanwhile (NCCL's
        # "NCCL version ..." banner at the first collective) is sent to stderr until the line is due
        sys.stdout.flush()
        saved_stdout = os.dup(1)
        os.dup2(2, 1)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    S.build()
    n = args.pairs
    numa = bind_to_gpu_numa_node(torch, local) if world > 1 and not os.environ.get("CUDASW_NO_NUMA_BIND") else None
    al = g.CudaAligner(1, -4, -4, 6, 1, device=local)

    # ---- inputs: this rank's shard of the counter-based C2 stream, in pinned host memory
    first = rank * n
    so = np.zeros(n + 1, dtype=np.int64)
    ro = np.zeros(n + 1, dtype=np.int64)
    so[:] = np.arange(n + 1, dtype=np.int64) * Q_LEN
    ro[:] = np.arange(n + 1, dtype=np.int64) * W_LEN
    pin_s, pin_r = PinnedBuffer(int(so[-1])), PinnedBuffer(int(ro[-1]))
    sb, so2, rb, ro2 = S.generate(CFG, n, first=first, out=(pin_s.array, pin_r.array))
    assert np.array_equal(so, so2) and np.array_equal(ro, ro2)
    cells_per_step = float(n) * Q_LEN * W_LEN
    d_seq = torch.from_numpy(sb[: so[-1]]).cuda()
    d_ref = torch.from_numpy(rb[: ro[-1]]).cuda()
    d_so = torch.from_numpy(so).cuda()
    d_ro = torch.from_numpy(ro).cuda()
    torch.cuda.synchronize()
    # result buffers of both arms are pinned, as the JNI shim's direct ByteBuffers are
    pin_res0, pin_cig0 = PinnedBuffer(n * g.RESULT_DTYPE.itemsize), PinnedBuffer(4 * 24 * n)
    res, cig = pin_res0.array.view(g.RESULT_DTYPE), pin_cig0.array.view(np.uint32)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step_device():
        r, c = al.align_blobs_device(d_seq.data_ptr(), d_so.data_ptr(), d_ref.data_ptr(), d_ro.data_ptr(), so, ro, res, cig)
        return al.stats()

    # e2e buffers: every host buffer the C ABI touches is pinned, as the JNI shim's direct ByteBuffers are
    pin_so, pin_ro = PinnedBuffer(8 * (n + 1)), PinnedBuffer(8 * (n + 1))
    so_p, ro_p = pin_so.array.view(np.int64), pin_ro.array.view(np.int64)
    so_p[:], ro_p[:] = so, ro
    pin_res, pin_cig = PinnedBuffer(n * g.RESULT_DTYPE.itemsize), PinnedBuffer(4 * 24 * n)
    res_p, cig_p = pin_res.array.view(g.RESULT_DTYPE), pin_cig.array.view(np.uint32)

    def step_host():
        r, c = al.align_blobs(sb, so_p, rb, ro_p, res=res_p, cig=cig_p)
        return al.stats(), r

    # ---- value: inputs resident in HBM
    for _ in range(warmup):
        step_device()
    sampler = ClockSampler(local)
    sampler.start()
    barrier()
    t0 = time.perf_counter()
    dev_ms = 0.0
    fwd_ms = 0.0
    launches = 0
    phase = {}
    visited_cells = 0.0
    flt_banded = flt_failed = 0
    for _ in range(args.steps):
        st = step_device()
        dev_ms += st["ms_total"]
        fwd_ms += st["ms_forward"]
        launches += st["kernel_launches"]
        for k in ("ms_pack", "ms_encode", "ms_forward", "ms_reverse", "ms_traceback", "ms_d2h"):
            phase[k] = phase.get(k, 0.0) + st[k] / args.steps
        enc_bytes = st["encode_bytes"]
        visited_cells += float(st["forward_cells_padded"])
        flt_banded, flt_failed = int(st["filter_banded"]), int(st["filter_failed"])
    barrier()
    wall_ms = (time.perf_counter() - t0) * 1e3
    clocks = sampler.summary()
    checksum = int(res["score"].astype(np.int64).sum()) ^ int(res["ref_begin"].astype(np.int64).sum() << 8)
    statuses_ok = bool((res["status"] == 0).all())

    from gridss_b200.shard import pair_digests
    dig_resident = pair_digests(res, cig, first_ordinal=first)  # ordinal, status, coordinates and every CIGAR word

    # ---- e2e (headline): the shim's packed host layout through cudasw_align_batch_packed (H2D + D2H inside).
    # Packing is the caller's job (it replaces the array copies + clean() the reference does per pair while it builds
    # the call, SswJniAligner.java:42-44); its host rate is reported beside the line, outside the timed region.
    w2q, wnq = g.packed_words(int(so[-1]))
    w2w, wnw = g.packed_words(int(ro[-1]))
    pin_q2, pin_qn, pin_w2, pin_wn = PinnedBuffer(4 * w2q), PinnedBuffer(4 * wnq), PinnedBuffer(4 * w2w), PinnedBuffer(4 * wnw)
    q2, qn, w2, wn = (b.array.view(np.uint32) for b in (pin_q2, pin_qn, pin_w2, pin_wn))
    pack_threads = max(1, min(16, (os.cpu_count() or 1) // max(1, world)))
    t0 = time.perf_counter()
    _, _, amb_q = g.pack_bases(sb, q2, qn, n=int(so[-1]), threads=pack_threads)
    _, _, amb_w = g.pack_bases(rb, w2, wn, n=int(ro[-1]), threads=pack_threads)
    host_pack_s = time.perf_counter() - t0

    def step_packed():
        r, c = al.align_blobs_packed(q2, qn if amb_q else None, so_p, w2, wn if amb_w else None, ro_p, res=res_p, cig=cig_p)
        return al.stats(), r, c

    for _ in range(2):
        step_packed()
    barrier()
    t0 = time.perf_counter()
    e2e_dev_ms = 0.0
    e2e_phase = {}
    for _ in range(args.steps):
        st_h, r_h, c_h = step_packed()
        e2e_dev_ms += st_h["ms_total"]
        for k in ("ms_h2d", "ms_pack", "ms_forward", "ms_reverse", "ms_traceback", "ms_d2h"):
            e2e_phase[k] = e2e_phase.get(k, 0.0) + st_h[k] / args.steps
    barrier()
    e2e_wall_ms = (time.perf_counter() - t0) * 1e3
    assert np.array_equal(pair_digests(r_h, c_h, first_ordinal=first), dig_resident), "packed e2e results differ from the resident path"

    # ---- e2e_ascii: raw ASCII host blobs through cudasw_align_batch (round 1's e2e)
    for _ in range(2):
        step_host()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        st_a, r_a = step_host()
    barrier()
    ascii_wall_ms = (time.perf_counter() - t0) * 1e3
    assert np.array_equal(pair_digests(r_a, cig_p, first_ordinal=first), dig_resident), "ASCII e2e results differ from the resident path"

    # ---- e2e with the windows named instead of shipped: the windows of this shard are loaded once
    # as one device-resident 2-bit "contig"; each step then copies queries + 24-byte descriptors only
    win_line = None
    if not args.no_windows:
        off1 = np.array([0, int(ro[-1])], dtype=np.int64)
        al.load_reference_blob(rb[: ro[-1]], off1)
        pin_win, pin_woff = PinnedBuffer(n * g.WINDOW_DTYPE.itemsize), PinnedBuffer(8 * (n + 1))
        win = pin_win.array.view(g.WINDOW_DTYPE)
        win["contig"], win["flags"], win["start"], win["end"] = 0, 0, ro[:-1], ro[1:]
        woff = pin_woff.array.view(np.int64)
        woff[:] = np.arange(n + 1, dtype=np.int64)
        for _ in range(2):
            al.align_blobs_windows(sb, so_p, win, woff, res=res_p, cig=cig_p)
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            r_w, _c = al.align_blobs_windows(sb, so_p, win, woff, res=res_p, cig=cig_p)
            st_w = al.stats()
        barrier()
        win_wall_ms = (time.perf_counter() - t0) * 1e3
        assert np.array_equal(pair_digests(r_w, _c, first_ordinal=first), dig_resident), "windows e2e results differ from the resident path"
        win_line = {"value": None, "unit": "alignments/s", "h2d_bytes_per_step": int(st_w["h2d_bytes"]),
                    "d2h_bytes_per_step": int(st_w["d2h_bytes"]), "ms_per_step_wall": win_wall_ms / args.steps,
                    "note": "cudasw_align_batch_windows: windows gathered on the GPU from the device-resident 2-bit reference "
                            "(loaded once, outside the timed region); result digests (coordinates + every CIGAR word) asserted identical to the resident path; "
                            "value = all ranks' pairs / max-over-ranks time"}

    # ---- caller level: BreakpointHomology.calculate for a list of breakpoints through
    # cudasw_breakpoint_homology, with GRIDSS's defaults (gridss.properties:88-89: 300 bp breakend, 10 bp margin). The
    # "genome" is this shard's windows as one contig; 32 bytes per breakpoint (+ inserted bases) cross PCIe per step.
    hom_line = None
    if not args.no_homology:
        glen = int(ro[-1])
        if args.no_windows:
            al.load_reference_blob(rb[:glen], np.array([0, glen], dtype=np.int64))
        nb = args.homology_breakpoints
        rng_h = np.random.default_rng(7007 + rank)
        bp = np.zeros(nb, dtype=g.BREAKPOINT_DTYPE)
        bp["pos1"] = rng_h.integers(1, glen + 1, size=nb)
        near = rng_h.random(nb) < 0.5     # half of the events are small (window sizes shrink to the event size)
        bp["pos2"] = np.where(near, np.minimum(glen, bp["pos1"] + rng_h.integers(0, 2000, size=nb)), rng_h.integers(1, glen + 1, size=nb))
        bp["dir1"] = rng_h.integers(0, 2, size=nb)
        bp["dir2"] = rng_h.integers(0, 2, size=nb)
        ins_len = np.where(rng_h.random(nb) < 0.3, rng_h.integers(1, 30, size=nb), 0)
        ins_off = np.zeros(nb + 1, dtype=np.int64)
        ins_off[1:] = np.cumsum(ins_len)
        ins_blob = np.frombuffer(b"ACGT", dtype=np.uint8)[rng_h.integers(0, 4, size=max(1, int(ins_off[-1])))].copy()
        for _ in range(2):
            hl, hr = al.breakpoint_homology(bp, ins_blob, ins_off, 300, 10)
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            hl, hr = al.breakpoint_homology(bp, ins_blob, ins_off, 300, 10)
            st_b = al.stats()
        barrier()
        hom_wall_ms = (time.perf_counter() - t0) * 1e3
        hom_line = {"metric": "breakpoints annotated/s per GPU", "value": float(nb) * args.steps / (hom_wall_ms / 1e3),
                    "alignments_per_s": float(st_b["pairs"]) * args.steps / (hom_wall_ms / 1e3), "breakpoints_per_step": nb,
                    "pairs_per_step": int(st_b["pairs"]), "ms_per_step_wall": hom_wall_ms / args.steps, "ms_per_step_device": st_b["ms_total"],
                    "h2d_bytes_per_step": int(st_b["h2d_bytes"]), "d2h_bytes_per_step": int(st_b["d2h_bytes"]),
                    "gcups": float(st_b["cells"]) * args.steps / (hom_wall_ms / 1e3) / 1e9,
                    "forward_tcups": float(st_b["cells"]) / (st_b["ms_forward"] / 1e3) / 1e12 if st_b["ms_forward"] else None,
                    "phases_ms": {k: round(float(st_b[k]), 3) for k in ("ms_h2d", "ms_pack", "ms_forward", "ms_reverse", "ms_traceback", "ms_d2h")},
                    "maxBreakendLength": 300, "margin": 10, "with_homology": int(((hl > 0) | (hr > 0)).sum()),
                    "note": "cudasw_breakpoint_homology: segment arithmetic on the host, queries and windows gathered from the "
                            "device-resident 2-bit reference, score passes only (clip lengths are all the caller reads)"}

    # ---- max over ranks
    t = torch.tensor([dev_ms, wall_ms, e2e_wall_ms, fwd_ms, e2e_dev_ms, ascii_wall_ms, win_wall_ms if win_line is not None else 0.0,
                      hom_wall_ms if hom_line is not None else 0.0], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms, wall_ms, e2e_wall_ms, fwd_ms, e2e_dev_ms, ascii_wall_ms, win_wall_ms, hom_wall_max_ms = [float(x) for x in t.tolist()]
    if win_line is not None:
        win_line["value"] = float(n) * world * args.steps / (win_wall_ms / 1e3)
        win_line["ms_per_step_wall"] = win_wall_ms / args.steps
    if hom_line is not None:
        hom_line["value_all_gpus"] = float(args.homology_breakpoints) * world * args.steps / (hom_wall_max_ms / 1e3)

    # ---- worst case for the forward kernel: query == a slice of a window of the same length, so the alignment (and the
    # E/F values it leaves behind) crosses every step of the sweep and no step is quiet. Rank 0 only; explains `value`.
    worst = None
    if rank == 0 and not args.no_worst_case:
        nw = args.worst_case_pairs
        rng_w = np.random.default_rng(9009)
        wq = np.frombuffer(b"ACGT", dtype=np.uint8)[rng_w.integers(0, 4, size=(nw, Q_LEN), dtype=np.uint8)]
        wso = np.arange(nw + 1, dtype=np.int64) * Q_LEN
        d_wq = torch.from_numpy(wq.reshape(-1)).cuda()
        d_wso = torch.from_numpy(wso).cuda()
        for _ in range(3):
            al.align_blobs_device(d_wq.data_ptr(), d_wso.data_ptr(), d_wq.data_ptr(), d_wso.data_ptr(), wso, wso)
        wms = wfwd = 0.0
        for _ in range(args.steps):
            r_wc, _ = al.align_blobs_device(d_wq.data_ptr(), d_wso.data_ptr(), d_wq.data_ptr(), d_wso.data_ptr(), wso, wso)
            st_wc = al.stats()
            wms += st_wc["ms_total"]
            wfwd += st_wc["ms_forward"]
        assert (r_wc["score"] == Q_LEN).all() and (r_wc["cigar_len"] == 1).all()
        worst = {"workload": f"{nw} pairs, query == window ({Q_LEN} x {Q_LEN}, identical sequences): zero quiet steps",
                 "alignments_per_s": nw * args.steps / (wms / 1e3), "forward_tcups": nw * float(Q_LEN) * Q_LEN * args.steps / (wfwd / 1e3) / 1e12,
                 "ms_per_step": wms / args.steps, "ms_forward": wfwd / args.steps, "inputs": "resident in HBM"}
        del d_wq, d_wso

    line = None
    if rank == 0:
        props = torch.cuda.get_device_properties(local)
        sms = props.multi_processor_count
        f_sm = (clocks["sm_mhz"] or 1965.0) * 1e6
        total_pairs = float(n) * world * args.steps
        total_cells = cells_per_step * world * args.steps
        value = total_pairs / (dev_ms / 1e3)
        # Cells: the k-mer filter (csrc/band_filter.cuh) lets the forward kernels sweep a proven column range of most
        # windows, so the cells they VISIT (counted on the device, padded rows and fill/drain columns included) are a fraction
        # of the algorithmic q x w. The roofline is about the kernel: it counts visited cells; the algorithmic rate is
        # reported beside it as gcups_effective.
        eff_cups = cells_per_step * args.steps / (fwd_ms / 1e3)  # q x w of every pair / forward time (rank max time)
        fwd_cups = visited_cells / (fwd_ms / 1e3)                # cells the forward kernels computed / forward time
        peak_laneops = sms * 128 * f_sm  # INT32 lane-ops/s at the clock seen during the run
        achieved_laneops = fwd_cups * INT_OPS_PER_CELL / 2.0  # packed s16x2: two cells per lane-op
        # the integer pipes measured on THIS box in THIS run (tools/int_peak.cu, dependent-free asm streams, < 1 s)
        measured = {}
        ip = os.path.join(ROOT, "tools", "int_peak")
        if os.path.exists(ip) and not args.no_int_peak:
            try:
                env = dict(os.environ, CUDA_VISIBLE_DEVICES=str(local) if "CUDA_VISIBLE_DEVICES" not in os.environ else os.environ["CUDA_VISIBLE_DEVICES"].split(",")[local])
                measured = json.loads(subprocess.run([ip], capture_output=True, text=True, timeout=120, env=env).stdout.strip().splitlines()[-1])
            except Exception as e:  # noqa: BLE001
                measured = {"error": repr(e)}
        # ALU-pipe instructions the forward kernel executes per packed cell pair: from the committed ncu capture of this
        # kernel on this workload (profiles/fwd_alu_mix.json, written by tools/profile_summaries.py), NOT from this run
        mix_file = os.path.join(ROOT, "profiles", "fwd_alu_mix.json")
        mix = json.load(open(mix_file)) if os.path.exists(mix_file) else {}
        alu_issue = measured.get("VIADDMNMX.S16x2_Tlaneops")
        alu_pipe_frac = None
        if mix.get("alu_lane_ops_per_visited_cell") and alu_issue:
            alu_pipe_frac = fwd_cups * mix["alu_lane_ops_per_visited_cell"] / (alu_issue * 1e12)
        line = {
            "metric": "alignments/sec", "value": value, "unit": "alignments/s", "n_gpus": world, "steps": args.steps,
            "warmup": warmup, "ms_per_step": dev_ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "s16x2", "data": "synthetic",
            "config": {"workload": WORKLOAD, "arithmetic": "packed int16 (s16x2) score passes, int32 traceback", "pairs_per_gpu_per_step": n, "query_len": Q_LEN, "window_len": W_LEN,
                       "sharding": f"pair-partition x{world}, no collective" + (f", ranks bound to their GPU's NUMA node {numa}" if numa is not None else ", no NUMA node exposed for the GPU in sysfs (no binding)" if world > 1 else ""),
                       "l2": "inputs (4.3 GB/step/GPU) exceed the 126 MB L2; no flush needed",
                       "bit_exact_checked": "tests/test_gpu_parity.py; bench asserts equal result digests (coordinates + every CIGAR word) across the resident, packed, ASCII and windows paths"},
            "gcups": total_cells / (dev_ms / 1e3) / 1e9,
            "wall_ms_per_step": wall_ms / args.steps,
            "phases_ms": {k: round(v, 3) for k, v in phase.items()},
            "e2e": {"value": float(n) * world * args.steps / (e2e_wall_ms / 1e3), "unit": "alignments/s",
                    "h2d_bytes_per_step": int(st_h["h2d_bytes"]), "d2h_bytes_per_step": int(st_h["d2h_bytes"]),
                    "gcups": total_cells / (e2e_wall_ms / 1e3) / 1e9, "ms_per_step_wall": e2e_wall_ms / args.steps,
                    "ms_per_step_device": e2e_dev_ms / args.steps,
                    "entry_point": "cudasw_align_batch_packed (2 bits per base + N mask: the shim's pinned-buffer layout)",
                    "host_buffers": "pinned (cudasw_host_alloc)",
                    "host_pack": {"seconds": host_pack_s, "threads": pack_threads, "gb_per_s_ascii_in": (int(so[-1]) + int(ro[-1])) / host_pack_s / 1e9,
                                  "note": "cudasw_pack_bases over this step's 4.3 GB, outside the timed region: it replaces the per-pair array copies + clean() "
                                          "the reference's wrapper does while building a call (SswJniAligner.java:42-44,71-79)"},
                    "phases_ms": {k: round(v, 3) for k, v in e2e_phase.items()},
                    "pipeline": "H2D in geometric chunks on a copy stream, overlapped with the previous chunk's kernels; ms_h2d = copy time not hidden"},
            "e2e_ascii": {"value": float(n) * world * args.steps / (ascii_wall_ms / 1e3), "unit": "alignments/s", "h2d_bytes_per_step": int(st_a["h2d_bytes"]),
                          "d2h_bytes_per_step": int(st_a["d2h_bytes"]), "ms_per_step_wall": ascii_wall_ms / args.steps,
                          "entry_point": "cudasw_align_batch (raw ASCII host blobs; round 1's e2e)"},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": {
                "bound": "int32_alu", "kernel": "sw_score_kernel<G=%d,R=%d,REV=false,PROF=true> (forward score pass, query-profile flavour)" % al.variant_for(Q_LEN)[:2],
                "achieved": achieved_laneops / 1e12, "peak": peak_laneops / 1e12, "unit": "T lane-op/s",
                "frac": achieved_laneops / peak_laneops,
                "traffic": mix.get("dram_bytes_per_pair", 0) * n if mix.get("dram_bytes_per_pair") else None,
                "traffic_source": ("%s (dram__bytes_read+write per pair x pairs per launch) - from that committed ncu capture, NOT measured by this run; algorithmic q+w = 4300 B/pair" % mix.get("source")) if mix else None,
                "gcups_kernel": fwd_cups / 1e9, "gcups_effective": eff_cups / 1e9, "int_ops_per_cell": INT_OPS_PER_CELL, "packing": 2,
                "cells_algorithmic_per_step": cells_per_step, "cells_visited_per_step": visited_cells / args.steps,
                "visited_frac": visited_cells / args.steps / cells_per_step,
                "filter": {"pairs_banded_last_step": flt_banded, "certificates_failed_last_step": flt_failed,
                           "what": "lossless k-mer filter + certificate in front of the forward pass (csrc/band_filter.cuh): banded pairs sweep only the columns a proven "
                                   "diagonal band can reach; a failed certificate sends the pair through the full window; results are bit-identical either way "
                                   "(tests/test_gpu_parity.py::test_filter_*, CUDASW_FILTER=0 switches it off)"},
                "peak_source": f"{sms} SMs x 128 lanes x {f_sm / 1e6:.0f} MHz (nominal INT32 lane rate at the median SM clock sampled during the run); "
                               "measured on this box in this run by tools/int_peak (dependent-free asm streams): see measured_issue_rates_Tlaneops",
                "measured_issue_rates_Tlaneops": {k.replace("_Tlaneops", ""): v for k, v in measured.items() if k.endswith("_Tlaneops")} or measured,
                "alu_pipe_frac": alu_pipe_frac,
                "alu_pipe_frac_how": "ALU-pipe lane-ops the kernel EXECUTES per visited cell (%s, from %s) x visited cells/s of this run / the measured VIADDMNMX.S16x2 issue rate; "
                                     "this is the fraction of the pipe the kernel actually lives on" % (mix.get("alu_lane_ops_per_visited_cell"), mix.get("source")),
                "note": "achieved = SURVEY.md 8d's 12 integer ops per cell x the cells the forward kernels VISITED per second / 2 (s16x2), against the nominal INT32 rate. "
                        "It can still exceed what the pipe can issue for this mix (one VIADDMNMX does three of those ops, quiet steps skip the E/F updates exactly): "
                        "alu_pipe_frac is the bounded yardstick. gcups_effective counts every q x w cell, visited or proven irrelevant.",
                "measured_cellmix_ceiling_TCUPS": measured.get("cellmix_TCUPS"),
                "frac_of_measured_cellmix_ceiling": (fwd_cups / 1e12 / measured["cellmix_TCUPS"]) if measured.get("cellmix_TCUPS") else None,
            },
            "result_checksum": checksum, "all_status_ok": statuses_ok,
        }
        if worst is not None:
            line["worst_case"] = worst
        peaks_file = os.path.join(ROOT, "MEASURED_PEAKS.json")
        hbm_peak = json.load(open(peaks_file)).get("hbm_gbs") if os.path.exists(peaks_file) else None
        if phase.get("ms_encode"):
            gbs = enc_bytes / (phase["ms_encode"] / 1e3) / 1e9
            line["roofline"]["load_phase"] = {"kernel": "encode_kernel (clean() + selector encoding, 1 B read + 1 B written per base)", "bound": "hbm",
                                              "achieved": gbs, "peak": hbm_peak or 6650.0, "unit": "GB/s", "frac": gbs / (hbm_peak or 6650.0),
                                              "peak_source": "MEASURED_PEAKS.json hbm_gbs" if hbm_peak else "B200_PROFILING.md fallback (6.65 TB/s)",
                                              "bytes_per_launch_pair": enc_bytes, "ms": phase["ms_encode"]}
        if st_h.get("ms_encode"):
            gbs_u = st_h["encode_bytes"] / (st_h["ms_encode"] / 1e3) / 1e9
            line["roofline"]["load_phase_packed"] = {"kernel": "unpack_kernel (2 bit + N mask words -> encoded bytes: 3/8 B read + 1 B written per base), e2e path",
                                                     "bound": "hbm", "achieved": gbs_u, "peak": hbm_peak or 6650.0, "unit": "GB/s", "frac": gbs_u / (hbm_peak or 6650.0),
                                                     "ms": st_h["ms_encode"], "note": "summed over the chunks of the last e2e step; later chunks unpack beside the previous chunk's forward kernels, so this is a lower bound of the kernel's own rate"}
        if win_line is not None:
            line["e2e_windows"] = win_line
        if hom_line is not None:
            line["callers"] = {"breakpoint_homology": hom_line}
        if not args.no_cpu_baseline and world >= 1:
            from oracle import oracle as O
            O.build()
            cores = os.cpu_count() or 1
            kind = "reference" if O.have_reference() else "port"
            run = O.ref_align_blobs if kind == "reference" else O.model_align_blobs
            k0 = 4 * cores
            t0 = time.perf_counter()
            run(sb[: so[k0]], so[: k0 + 1], rb[: ro[k0]], ro[: k0 + 1], threads=cores)
            rate = k0 / max(time.perf_counter() - t0, 1e-6)
            k = int(min(n, max(k0, rate * args.cpu_baseline_seconds)))
            t0 = time.perf_counter()
            cpu = run(sb[: so[k]], so[: k + 1], rb[: ro[k]], ro[: k + 1], threads=cores)
            dt = time.perf_counter() - t0
            # full result digests on both sides: ordinal, status, coordinates and every CIGAR word
            same = bool(np.array_equal(pair_digests(cpu.res, cpu.cig, first_ordinal=first), dig_resident[:k]))
            line["cpu_baseline"] = {"value": k / dt, "unit": "alignments/s", "cores": cores, "kind": kind,
                                    "sample": f"first {k} pairs of rank 0's C2 shard, {cores} host threads, "
                                              + ("libsswjni.so as shipped by the reference (-O0 SSE2 build)" if kind == "reference" else "oracle/ssw_model.c -O3"),
                                    "gcups": k * Q_LEN * W_LEN / dt / 1e9, "gpu_results_identical_on_sample": same,
                                    "identity_check": "pair digests over coordinates and every CIGAR word (gridss_b200.shard.pair_digests)"}
            if kind == "reference":
                # the same algorithm compiled -O3 (oracle/ssw_model.c, scalar), so the ratio is not flattered
                # by the reference's -O0 build flags
                k3 = int(min(n, max(k0, k // 2)))
                t0 = time.perf_counter()
                cpu3 = O.model_align_blobs(sb[: so[k3]], so[: k3 + 1], rb[: ro[k3]], ro[: k3 + 1], threads=cores)
                dt3 = time.perf_counter() - t0
                line["cpu_baseline_o3"] = {"value": k3 / dt3, "unit": "alignments/s", "cores": cores, "kind": "port",
                                           "sample": f"first {k3} pairs of the same shard, {cores} host threads, oracle/ssw_model.c (scalar restatement) gcc -O3",
                                           "gcups": k3 * Q_LEN * W_LEN / dt3 / 1e9,
                                           "gpu_results_identical_on_sample": bool(np.array_equal(pair_digests(cpu3.res, cpu3.cig, first_ordinal=first), dig_resident[:k3]))}
            if hom_line is not None:
                # the same breakpoints the way the reference does them: host strings, two libsswjni calls each
                from oracle import callers_restatement as R
                kb = min(nb, 4000)
                contig = [rb[:glen]]
                prs, owners = [], []
                for i in range(kb):
                    b = bp[i]
                    ins_b = ins_blob[ins_off[i]:ins_off[i + 1]].tobytes()
                    nl, nr, pa, pb = R.homology_pairs(contig, (0, "fb"[b["dir1"]], int(b["pos1"]), 0, "fb"[b["dir2"]], int(b["pos2"])), ins_b, 300, 10)
                    owners.append((nl, nr, len(prs) if pa else -1, len(prs) + (1 if pa else 0) if pb else -1))
                    prs += [p for p in (pa, pb) if p]
                t0 = time.perf_counter()
                cpu_h = (O.ref_align if kind == "reference" else O.model_align)([p[0] for p in prs], [p[1] for p in prs], threads=cores)
                dt_h = time.perf_counter() - t0
                want = [R.homology_from_cigars(nl, nr, cpu_h.cigar(a) if a >= 0 else None, cpu_h.cigar(c) if c >= 0 else None)
                        for nl, nr, a, c in owners]
                same_h = want == list(zip(hl[:kb].tolist(), hr[:kb].tolist()))
                hom_line["cpu_baseline"] = {"value": kb / dt_h, "unit": "breakpoints/s", "cores": cores, "kind": kind,
                                            "sample": f"first {kb} breakpoints ({len(prs)} aligner calls), {cores} host threads, aligner time only "
                                                      "(window strings prepared outside the timed region)",
                                            "gpu_results_identical_on_sample": bool(same_h)}
        if saved_stdout >= 0:
            sys.stdout.flush()
            os.dup2(saved_stdout, 1)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    al.close()
    return 0


if __name__ == "__main__":
    sys.exit(main())
