#!/usr/bin/env python
"""Markdown table of the bench lines of one tag:  python tools/bench_table.py profiles/r01_h [profiles/r01_i ...]
(later prefixes override earlier ones per (config, mode); used to fill the tables of DESIGN.md)"""
import glob
import json
import re
import sys

rows = {}
for prefix in sys.argv[1:]:
    for f in sorted(glob.glob(prefix + "_bench_cfg*.json")):
        m = re.search(r"_bench_(cfg\d)_(analytic|numerical)(_\w+)?\.json$", f)
        if not m or m.group(3):
            continue
        for line in open(f):
            if line.startswith("{"):
                rows[(m.group(1), m.group(2))] = (json.loads(line), f)
print("| config | mode | forward / backward kernel | step | e2e step | Gaussian-bin evals/s | frac fwd / bwd | source |")
print("|---|---|---|---|---|---|---|---|")
for (cfg, mode), (d, f) in sorted(rows.items()):
    r = d["roofline"]
    print(f"| {cfg} | {mode} | {r['fwd']['launch_ms']:.2f} / {r['bwd']['launch_ms']:.2f} ms | {d['ms_per_step']:.1f} ms | "
          f"{d['e2e']['ms_per_step']:.1f} ms | {d['value']:.3g} | {r['fwd']['frac']:.2f} / {r['bwd']['frac']:.2f} ({r['fwd']['binding_pipe']}) | `{f}` |")
