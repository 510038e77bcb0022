#!/usr/bin/env python
"""Turns the raw ncu outputs of tools/gpu_session.sh (gpurun_out/) into the tracked summaries under profiles/.

  python tools/profile_summaries.py r01z        # tag of the session
"""
import collections
import csv
import io
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
O = os.path.join(ROOT, "gpurun_out")
P = os.path.join(ROOT, "profiles")
KEYS = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__registers_per_thread", "launch__occupancy_limit_registers",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "smsp__inst_executed.sum", "sm__cycles_elapsed.avg", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__thread_inst_executed_per_inst_executed.ratio",
    "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
    "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
]


def launch_list(tag, cmd):
    rows = list(csv.reader(l for l in open(os.path.join(O, f"launches_{tag}.csv")) if l.startswith('"')))
    h = rows[0]
    ki, vi, ui = h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Unit")
    agg = collections.OrderedDict()
    for r in rows[1:]:
        k = re.sub(r"\(.*", "", r[ki])
        v = float(r[vi].replace(",", ""))
        v = v / 1e6 if r[ui] in ("ns", "nsecond") else v / 1e3 if r[ui] in ("us", "usecond") else v
        a = agg.setdefault(k, [0, 0.0])
        a[0] += 1
        a[1] += v
    tot = sum(a[1] for a in agg.values())
    out = [f"# ncu launch list ({tag})", "", f"`ncu --metrics gpu__time_duration.sum --clock-control none -c 900 {cmd}`",
           "(per-launch times are cold-cache and serialised: compare shares)", "", "| total ms | share | launches | kernel |", "|---|---|---|---|"]
    for k, a in sorted(agg.items(), key=lambda x: -x[1][1])[:28]:
        out.append(f"| {a[1]:.3f} | {100 * a[1] / tot:.2f}% | {a[0]} | `{k[:100]}` |")
    out.append(f"| {tot:.3f} | 100% | {sum(a[0] for a in agg.values())} | all {len(agg)} kernels |")
    open(os.path.join(P, f"launches_{tag}.md"), "w").write("\n".join(out) + "\n")
    import shutil
    shutil.copy(os.path.join(O, f"launches_{tag}.csv"), os.path.join(P, f"launches_{tag}.csv"))


def full_capture(rep, name, title, cmd, notes=""):
    raw = subprocess.run(["ncu", "-i", os.path.join(O, rep), "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    h, u, v = rows[0], rows[1], rows[2]
    out = [f"# ncu --set full, {title}", "", f"`{cmd}`", "", "| metric | unit | value |", "|---|---|---|"]
    for k in ["Kernel Name", "Grid Size"] + KEYS:
        if k in h:
            i = h.index(k)
            out.append(f"| {k} | {u[i]} | {v[i]} |")
    if notes:
        out += ["", notes]
    open(os.path.join(P, name), "w").write("\n".join(out) + "\n")
    UNITS.clear()
    UNITS.update({k: u[h.index(k)] for k in h if k in KEYS})
    return {k: v[h.index(k)] for k in h if k in KEYS}


UNITS = {}  # units of the last full_capture()


def mbytes(m, k):
    x = float(str(m[k]).replace(",", ""))
    return x * {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}.get(UNITS.get(k, "Mbyte"), 1.0)


def alu_mix(tag, m, cmd_pairs=50000, q=300, w=4000):
    """profiles/fwd_alu_mix.json: ALU-pipe lane-ops per DP cell and DRAM bytes per pair of the forward kernel, from the
    --set full capture (bench.py reads it for roofline.alu_pipe_frac / roofline.traffic)."""
    import json

    def f(k):
        return float(str(m[k]).replace(",", ""))

    cyc, alu = f("sm__cycles_elapsed.avg"), f("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active") / 100.0
    smsps = 148 * 4
    warp_inst = alu * cyc * smsps * 0.5  # the ALU pipe takes one warp instruction every other clock (16 lanes)
    cells = float(cmd_pairs) * q * w
    rd, wr = f("dram__bytes_read.sum"), f("dram__bytes_write.sum")  # Mbyte
    out = {"alu_lane_ops_per_cell": round(warp_inst * 32 / cells, 3), "dram_bytes_per_pair": round((rd + wr) * 1e6 / cmd_pairs, 1),
           "source": f"profiles/ncu_fwd_{tag}.md",
           "how": f"sm__inst_executed_pipe_alu {alu * 100:.3f} % x sm__cycles_elapsed {cyc:.0f} x {smsps} SMSPs x 0.5 warp-inst/clk (16-lane ALU pipe) x 32 lanes / "
                  f"({cmd_pairs} pairs x {q} x {w} cells); dram__bytes_read+write {rd + wr:.2f} MB / {cmd_pairs} pairs"}
    json.dump(out, open(os.path.join(P, "fwd_alu_mix.json"), "w"), indent=1)
    print("fwd_alu_mix", out)


if __name__ == "__main__" and len(sys.argv) > 4 and sys.argv[1] == "one":
    # python tools/profile_summaries.py one <rep in gpurun_out> <md under profiles> <title> [command]
    m = full_capture(sys.argv[2], sys.argv[3], sys.argv[4], sys.argv[5] if len(sys.argv) > 5 else "")
    print({k: m[k] for k in list(m)[:10]})
    sys.exit(0)

def alu_mix_visited(tag, m, plain_log, q=300, w=4000, rows_padded=304, lanes=8):
    """profiles/fwd_alu_mix.json for the k-mer-filtered pipeline: ALU-pipe lane-ops per VISITED cell of the forward kernel's
    sub-window pass. The visited cells of the captured launch = what the bench line of the same command reports per step,
    minus the second (full-window) pass of the pairs whose certificate failed."""
    import json

    def f(k):
        return float(str(m[k]).replace(",", ""))

    line = json.loads(open(plain_log).read().strip().splitlines()[-1])
    rf = line["roofline"]
    pairs = line["config"]["pairs_per_gpu_per_step"]
    failed = rf["filter"]["certificates_failed_last_step"]
    visited = rf["cells_visited_per_step"] - failed * rows_padded * (w + 2 * lanes - 1)
    cyc, alu = f("sm__cycles_elapsed.avg"), f("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active") / 100.0
    smsps = 148 * 4
    warp_inst = alu * cyc * smsps * 0.5  # the ALU pipe takes one warp instruction every other clock (16 lanes)
    rd, wr = mbytes(m, "dram__bytes_read.sum"), mbytes(m, "dram__bytes_write.sum")
    out = {"alu_lane_ops_per_visited_cell": round(warp_inst * 32 / visited, 3), "visited_cells_of_the_launch": visited,
           "dram_bytes_per_pair": round((rd + wr) * 1e6 / pairs, 1), "source": f"profiles/ncu_fwd_{tag}.md",
           "how": f"sm__inst_executed_pipe_alu {alu * 100:.3f} % x sm__cycles_elapsed {cyc:.0f} x {smsps} SMSPs x 0.5 warp-inst/clk (16-lane ALU pipe) x 32 lanes / "
                  f"{visited:.4g} visited cells (bench line of the same command: cells_visited_per_step {rf['cells_visited_per_step']:.4g} minus {failed} unproven pairs x "
                  f"{rows_padded} x {w + 2 * lanes - 1} of the second pass); dram__bytes_read+write {rd + wr:.2f} MB / {pairs} pairs"}
    json.dump(out, open(os.path.join(P, "fwd_alu_mix.json"), "w"), indent=1)
    print("fwd_alu_mix", out)


if __name__ == "__main__" and len(sys.argv) > 2 and sys.argv[2] == "r02f":
    # round 2, k-mer-filtered pipeline (tools/gpu_session5.sh)
    tag = sys.argv[1]
    cmd = "python bench.py --steps 2 --warmup 3 --pairs 50000 --no-cpu-baseline --no-windows --no-homology --no-worst-case --no-int-peak"
    launch_list(tag, cmd)
    caps = [("fwd", "forward score kernel, sub-window pass of the k-mer-filtered pipeline (C2, chained query-profile flavour)", "sw_score_kernelILi8ELi19ELb0ELb0ELb1ELb1E -s 6", cmd),
            ("filter", "k-mer filter kernel (hits per diagonal bin, column range, certificate data; C2)", "filter_kernel -s 3", cmd),
            ("rev", "banded reverse pass with the certified gap bound (C2)", "sw_rev_band_kernelILi1ELi16ELb1E -s 3", cmd)]
    for name, title, kre, c in caps:
        rep = f"prof_{name}_{tag}.ncu-rep"
        if os.path.exists(os.path.join(O, rep)):
            m = full_capture(rep, f"ncu_{name}_{tag}.md", title, f"ncu --set full --clock-control none --import-source on --kernel-name-base mangled -k regex:{kre} -c 1 {c}")
            print(name, {k: m[k] for k in list(m)[:8]})
            if name == "fwd":
                alu_mix_visited(tag, m, os.path.join(O, f"plain_{tag}.log"))
    import shutil
    shutil.copy(os.path.join(O, f"plain_{tag}.log"), os.path.join(P, f"bench_{tag}_50k_profiled_command.json"))
    sys.exit(0)

if __name__ == "__main__" and len(sys.argv) > 2 and sys.argv[2] == "r02":
    # round 2 session (tools/gpu_session3.sh)
    tag = sys.argv[1]
    cmd = "python bench.py --steps 2 --warmup 3 --pairs 50000 --no-cpu-baseline --no-windows --no-homology --no-worst-case"
    launch_list(tag, cmd)
    caps = [("fwd", "forward score kernel (C2, chained query-profile flavour)", "sw_score_kernel -s 2", cmd),
            ("rev", "banded reverse pass (C2)", "sw_rev_band_kernel -s 3", cmd),
            ("unpack", "unpack kernel (packed host inputs -> encoded bytes; load phase of the packed path)", "unpack_kernel -s 3", cmd),
            ("tb", "one-thread traceback kernel (C2)", "sw_traceback_kernel -s 12", cmd),
            ("tbwarp", "one-warp-per-pair cooperative traceback (C3, 200 k pairs)", "sw_traceback_warp_kernel -s 5", "python tools/bench_configs.py 3:200000"),
            ("tbcta", "CTA-per-pair cooperative traceback (C5, 1000 pairs)", "sw_traceback_block_kernel -s 1", "python tools/bench_configs.py 5:1000"),
            ("fwdstrip", "strip-mined forward kernel (C5: 10 kb contigs, one warp per pair)", "'sw_score_kernel<32' -s 1", "python tools/bench_configs.py 5:1000"),
            ("revc5", "banded reverse pass, one warp per pair (C5)", "'sw_rev_band_kernel<32' -s 1", "python tools/bench_configs.py 5:1000")]
    for name, title, kre, c in caps:
        rep = f"prof_{name}_{tag}.ncu-rep"
        if os.path.exists(os.path.join(O, rep)):
            m = full_capture(rep, f"ncu_{name}_{tag}.md", title, f"ncu --set full --clock-control none --import-source on -k regex:{kre} -c 1 {c}")
            print(name, {k: m[k] for k in list(m)[:8]})
            if name == "fwd":
                alu_mix(tag, m)
    sys.exit(0)

if __name__ == "__main__":
    tag = sys.argv[1]
    cmd = "python bench.py --steps 2 --warmup 3 --pairs 50000 --no-cpu-baseline --no-windows --no-homology"
    hcmd = "python bench.py --steps 1 --warmup 3 --pairs 50000 --no-cpu-baseline --no-windows --homology-breakpoints 100000"
    launch_list(tag, cmd)
    for rep, name, title, kre in ((f"prof_fwd_{tag}.ncu-rep", f"ncu_fwd_{tag}.md", "forward score kernel", "sw_score_kernel -s 2"),
                                  (f"prof_tb_{tag}.ncu-rep", f"ncu_tb_{tag}.md", "one-thread traceback kernel", "sw_traceback_kernel -s 5"),
                                  (f"prof_enc_{tag}.ncu-rep", f"ncu_enc_{tag}.md", "encode kernel (load phase)", "encode_kernel -s 9")):
        if os.path.exists(os.path.join(O, rep)):
            m = full_capture(rep, name, title, f"ncu --set full --clock-control none --import-source on -k regex:{kre} -c 1 {cmd}")
            print(name, {k: m[k] for k in list(m)[:6]})
    rep = f"prof_gather_{tag}.ncu-rep"
    if os.path.exists(os.path.join(O, rep)):
        m = full_capture(rep, f"ncu_gather_{tag}.md", "query gather of cudasw_breakpoint_homology (window_gather_kernel, 200 k pairs x 3 segment slots)",
                         f"ncu --set full --clock-control none --import-source on -k regex:window_gather_kernel -s 4 -c 1 {hcmd}")
        print("gather", {k: m[k] for k in list(m)[:6]})
