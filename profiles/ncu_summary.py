#!/usr/bin/env python
"""Summarises an .ncu-rep (read here, no GPU needed) into the markdown tables committed under profiles/.

    python profiles/ncu_summary.py gpurun_out/prof.ncu-rep [--lines N] > profiles/rNN_....md

Raw page: the metrics the judge greps (DRAM bytes, pipe utilisation, occupancy, stall reasons).
Source page: instructions executed / stall samples aggregated per source line (needs -lineinfo).
"""
import csv
import io
import re
import subprocess
import sys
from collections import defaultdict

KEEP = re.compile(
    r"dram__bytes_(read|write)\.sum$|gpu__time_duration\.sum|gpu__dram_throughput.avg.pct|launch__(registers_per_thread|occupancy_limit|"
    r"shared_mem_per_block_dynamic|waves|grid_size|block_size)|sm__warps_active.avg.pct|sm__throughput.avg.pct|sm__cycles_elapsed.max|"
    r"sm__inst_executed_pipe_(fp64|alu|fma|lsu|adu|xu|uniform|cbu)\.avg\.pct|sm__pipe_fp64_cycles_active.avg.pct|smsp__inst_executed\.sum$|"
    r"smsp__issue_active.avg.pct|smsp__average_warps_issue_stalled_.*_per_issue_active|l1tex__data_bank_conflicts_pipe_lsu_mem_shared\.sum$|"
    r"l1tex__data_pipe_lsu_wavefronts_mem_shared\.sum$|l1tex__t_sector_hit_rate|lts__t_sector_hit_rate|sm__sass_thread_inst_executed_op_d(fma|mul|add)_pred_on\.sum$|"
    r"smsp__sass_inst_executed_op_(shared|global|local)_(ld|st)\.sum$|local_load|local_store|sm__inst_executed_pipe_fp64\.sum")


def ncu(args):
    return subprocess.run(["ncu"] + args, capture_output=True, text=True, check=True).stdout


def raw(rep):
    rows = list(csv.reader(io.StringIO(ncu(["-i", rep, "--page", "raw", "--csv"]))))
    head, units, data = rows[0], rows[1], rows[2:]
    for d in data:
        name = d[head.index("Kernel Name")]
        print(f"## {name}  grid {d[head.index('Grid Size')]} block {d[head.index('Block Size')]}\n")
        print("| metric | unit | value |\n|---|---|---|")
        for h, u, v in sorted(zip(head, units, d)):
            if KEEP.search(h):
                print(f"| {h} | {u} | {v} |")
        print()


def source(rep, top):
    out = ncu(["-i", rep, "--page", "source", "--csv", "--print-source", "sass,cuda" if False else "cuda,sass"])
    rows = list(csv.reader(io.StringIO(out)))
    # find header row
    hi = next(i for i, r in enumerate(rows) if "Source" in r and any("Instructions Executed" in c for c in r))
    head = rows[hi]
    col = {c: i for i, c in enumerate(head)}
    src_i = col["Source"]
    ie = next(i for c, i in col.items() if c == "Instructions Executed")
    ws = next((i for c, i in col.items() if c.startswith("Warp Stall Sampling (All")), None)
    agg = defaultdict(lambda: [0.0, 0.0, ""])
    cur = None
    for r in rows[hi + 1:]:
        if len(r) <= max(src_i, ie):
            continue
        s = r[src_i]
        key = s.strip()[:110]
        try:
            n = float(r[ie] or 0)
        except ValueError:
            continue
        w = float(r[ws] or 0) if ws is not None and r[ws] else 0.0
        a = agg[key]
        a[0] += n
        a[1] += w
    tot_i = sum(a[0] for a in agg.values()) or 1.0
    tot_w = sum(a[1] for a in agg.values()) or 1.0
    print(f"## source hot spots (top {top} by stall samples; totals: {tot_i:.4g} warp-insts, {tot_w:.4g} samples)\n")
    print("| % insts | % stall samples | source / sass |\n|---|---|---|")
    for key, a in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
        print(f"| {100 * a[0] / tot_i:.2f} | {100 * a[1] / tot_w:.2f} | `{key}` |")


if __name__ == "__main__":
    rep = sys.argv[1]
    top = int(sys.argv[sys.argv.index("--lines") + 1]) if "--lines" in sys.argv else 40
    raw(rep)
    if "--no-source" not in sys.argv:
        source(rep, top)
