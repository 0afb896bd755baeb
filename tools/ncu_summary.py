#!/usr/bin/env python
"""Summarise ncu artefacts brought back in gpurun_out/ into small text files for profiles/.

  python tools/ncu_summary.py launches <launches.csv>          # per-kernel share of the step
  python tools/ncu_summary.py raw <report.ncu-rep>              # key metrics per captured launch
  python tools/ncu_summary.py lines <report.ncu-rep> [top]      # hottest source lines (needs -lineinfo)
"""
import collections
import csv
import io
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "launch__registers_per_thread", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smsp__cycles_active.avg"]


def launches(path):
    rows = list(csv.reader(open(path)))
    hdr = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
    h = rows[hdr]
    ki, vi, ui = h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Unit")
    agg = collections.OrderedDict()
    for r in rows[hdr + 1:]:
        if len(r) <= vi:
            continue
        v = float(r[vi].replace(",", ""))
        v = v / 1e6 if r[ui] in ("ns", "nsecond") else (v / 1e3 if r[ui] in ("us", "usecond") else v)
        a = agg.setdefault(r[ki][:90], [0, 0.0])
        a[0] += 1
        a[1] += v
    tot = sum(a[1] for a in agg.values())
    print(f"# {path}: {sum(a[0] for a in agg.values())} launches, {tot:.3f} ms total (ncu: cold cache, serialised -> compare SHARES)")
    for k, a in sorted(agg.items(), key=lambda x: -x[1][1]):
        print(f"{a[1]:11.3f} ms {a[0]:5d}x {100 * a[1] / tot:6.2f}%  {k}")


def _ncu(rep, *args):
    return subprocess.run(["ncu", "-i", rep, *args], capture_output=True, text=True).stdout


def raw(rep):
    rows = list(csv.reader(io.StringIO(_ncu(rep, "--page", "raw", "--csv"))))
    h = rows[0]
    names = [r[h.index("Kernel Name")][:60] for r in rows[2:]]
    print("# " + rep)
    print("metric | " + " | ".join(names))
    for k in KEYS:
        if k in h:
            i = h.index(k)
            print(f"{k} [{rows[1][i]}] | " + " | ".join(r[i] for r in rows[2:]))


def lines(rep, top=28):
    rows = list(csv.reader(io.StringIO(_ncu(rep, "--page", "source", "--csv", "--print-source", "cuda,sass"))))
    cur_file = cur_fn = None
    out = {}
    for r in rows:
        if not r:
            continue
        if r[0] == "File Path":
            cur_file = r[1].split("/")[-1]
        elif r[0] == "Function Name":
            cur_fn = r[1]
        elif r[0].isdigit():
            try:
                out[(cur_fn, cur_file, int(r[0]))] = (int(r[4]), int(r[7]), int(r[8]), r[1].strip()[:100])
            except (ValueError, IndexError):
                pass
    for fn in sorted(set(k[0] for k in out)):
        sel = [(k, v) for k, v in out.items() if k[0] == fn]
        ts, ti = sum(v[0] for _, v in sel), sum(v[1] for _, v in sel)
        print(f"== {fn}: {ts} stall samples, {ti} warp instructions")
        print("  %samples %warp-inst thr/inst  file:line  source")
        for k, v in sorted(sel, key=lambda x: -x[1][0])[:top]:
            print(f"  {100 * v[0] / ts:6.1f}  {100 * v[1] / ti:6.1f}  {v[2] / max(v[1], 1):5.1f}  {k[1]}:{k[2]}  {v[3]}")


if __name__ == "__main__":
    {"launches": launches, "raw": raw, "lines": lambda p, t=28: lines(p, int(t))}[sys.argv[1]](*sys.argv[2:])
