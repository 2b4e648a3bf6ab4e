#!/usr/bin/env python
"""Summaries of ncu output, read on the CPU box.
  python tools/ncu_summary.py launches <launches.csv> [first] [count]   per-launch durations
  python tools/ncu_summary.py raw <prof.ncu-rep>                        key metrics per captured launch
  python tools/ncu_summary.py source <prof.ncu-rep> <launch> [top]      per-source-line instruction share / active threads / stall samples
"""
import collections
import csv
import io
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum", "launch__registers_per_thread",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio", "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio", "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio", "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio",
        "sm__cycles_active.avg", "sm__cycles_elapsed.max", "l1tex__t_sectors_pipe_lsu_mem_local_op_ld.sum", "l1tex__t_sectors_pipe_lsu_mem_local_op_st.sum",
        "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "lts__t_sectors_op_read.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_active",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
        "sm__maximum_warps_per_active_cycle_pct", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smsp__inst_executed_pipe_lsu.sum",
        "smsp__inst_executed_pipe_fma.sum", "smsp__inst_executed_pipe_alu.sum", "smsp__inst_executed_pipe_xu.sum"]


def ncu(args):
    return subprocess.run(["ncu", *args], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout


def launches(path, first=0, count=10**9):
    rows = [r for r in csv.reader(open(path)) if len(r) > 10 and r[0].isdigit()]
    total = 0.0
    for r in rows[first:first + count]:
        name = r[4].split("(")[0].replace("void ", "").replace("crtdev::", "")
        ms = float(r[-1]) / 1e6
        total += ms
        print(f"{int(r[0]):4d} {name[:44]:44s} grid {r[8]:16s} {ms:10.4f} ms")
    print(f"sum {total:.4f} ms over {len(rows[first:first + count])} launches")


def raw(rep):
    rows = list(csv.reader(io.StringIO(ncu(["-i", rep, "--page", "raw", "--csv"]))))
    hdr, units = rows[0], rows[1]
    for k in KEYS:
        if k in hdr:
            i = hdr.index(k)
            print(f"{k:90s} {units[i]:10s} {[r[i][:12] for r in rows[2:]]}")


def source(rep, launch, top=40):
    text = ncu(["-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--launch-skip", str(launch), "--launch-count", "1"])
    cur, hdr, agg = None, None, collections.OrderedDict()
    for r in csv.reader(io.StringIO(text)):
        if not r:
            continue
        if r[0] == "File Path":
            cur = r[1].split("/")[-1]
        elif r[0] == "Line No":
            hdr = r
        elif hdr and len(r) == len(hdr) and r[0].isdigit():
            d = dict(zip(hdr, r))
            a = agg.setdefault((cur, int(r[0])), [0.0, 0.0, 0.0, collections.Counter()])
            a[0] += float(d["Instructions Executed"] or 0)
            a[1] += float(d["Predicated-On Thread Instructions Executed"] or 0)
            a[2] += float(d["# Samples"] or 0)
            for k, v in d.items():
                if k.startswith("stall_") and "Not Issued" not in k and v not in ("", "0"):
                    a[3][k[6:]] += float(v)
    ti = sum(a[0] for a in agg.values()) or 1
    ts = sum(a[2] for a in agg.values()) or 1
    print(f"warp instructions {ti:.4g}, thread instructions {sum(a[1] for a in agg.values()):.4g}, samples {ts:.0f}")
    for (f, ln), a in sorted(agg.items(), key=lambda kv: -(kv[1][0] / ti + kv[1][2] / ts))[:top]:
        stalls = ",".join(f"{k}:{v / ts * 100:.1f}" for k, v in a[3].most_common(3))
        print(f"{f[:20]:20s}:{ln:4d} inst {a[0] / ti * 100:5.1f}%  thr/inst {a[1] / max(a[0], 1):5.1f}  samples {a[2] / ts * 100:5.1f}%  {stalls}")


if __name__ == "__main__":
    cmd = sys.argv[1]
    if cmd == "launches":
        launches(sys.argv[2], *(int(v) for v in sys.argv[3:5]))
    elif cmd == "raw":
        raw(sys.argv[2])
    else:
        source(sys.argv[2], int(sys.argv[3]), *(int(v) for v in sys.argv[4:5]))
