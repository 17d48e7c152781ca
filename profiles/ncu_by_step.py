#!/usr/bin/env python
"""Attributes an ncu capture of the scene-specialised kernel to the STEPS of the walk (the lines of the generated
omo_jit_kernel.cu): warp-instructions executed, FP64 instructions, stall samples by reason, per step.

    ncu -i prof.ncu-rep --page source --csv --print-source sass > sass.csv
    nvdisasm --print-line-info-inline kernel.cubin > dis.txt        (cubin from OMO_JIT_DUMP / probe_specialised_kernel)
    python profiles/ncu_by_step.py sass.csv dis.txt kernel.cu [--inner]  > profiles/rNN_....md

--inner additionally aggregates by the innermost source line (file:line of the .cuh device code).
"""
import csv
import re
import sys
from collections import defaultdict


def line_map(dis_path, kernel="omo_jit_kernel"):
    """offset -> (outer line in the generated .cu, innermost 'file:line')"""
    out = {}
    inside = False
    pend = []
    for ln in open(dis_path):
        if ln.startswith(".text." + kernel):
            inside = True
            continue
        if inside and ln.startswith(".") and not ln.startswith(".L"):
            if out:
                break
        if not inside:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m:
            pend.append((m.group(1).split("/")[-1], int(m.group(2))))
            continue
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
        if m:
            off = int(m.group(1), 16)
            if pend:
                outer = next((l for f, l in reversed(pend) if f.endswith("omo_jit_kernel.cu")), -1)
                inner = f"{pend[0][0]}:{pend[0][1]}"
                cur = (outer, inner)
                pend = []
            out[off] = cur
    return out


def main():
    sass_csv, dis, gen = sys.argv[1:4]
    inner_too = "--inner" in sys.argv
    lm = line_map(dis)
    rows = list(csv.reader(open(sass_csv)))
    hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    head = rows[hi]
    col = {c: i for i, c in enumerate(head)}
    stall_cols = [c for c in head if c.startswith("stall_") and "Not Issued" not in c]
    base = None
    agg = defaultdict(lambda: defaultdict(float))
    agg_in = defaultdict(lambda: defaultdict(float))
    tot = defaultdict(float)
    for r in rows[hi + 1:]:
        if len(r) < len(head) or not r[0].startswith("0x"):
            continue
        addr = int(r[0], 16)
        if base is None:
            base = addr
        off = addr - base
        outer, inner = lm.get(off, (-1, "?"))
        sass = r[col["Source"]].strip()
        op = re.sub(r"^@!?U?P\d+\s+", "", sass).split()[0].split(".")[0]
        n = float(r[col["Instructions Executed"]] or 0)
        smp = float(r[col["# Samples"]] or 0)
        for a in (agg[outer], tot) + ((agg_in[inner],) if inner_too else ()):
            a["inst"] += n
            a["samples"] += smp
            if op in ("DFMA", "DMUL", "DADD", "DSETP"):
                a["fp64"] += n
            if op in ("LDG", "LDS", "STS", "LDL", "STL", "LDC", "LDCU"):
                a[op] += n
            for c in stall_cols:
                a[c] += float(r[col[c]] or 0)
    src = open(gen).read().split("\n")
    top = ["stall_wait", "stall_long_sb", "stall_short_sb", "stall_no_inst", "stall_math", "stall_not_selected", "stall_barrier", "stall_branch_resolving"]
    print(f"totals: {tot['inst']:.4g} warp-insts, {tot['fp64']:.4g} FP64 ({100 * tot['fp64'] / tot['inst']:.1f} %), {tot['samples']:.0f} samples\n")
    print("| line | step | % insts | % fp64 | % samples | " + " | ".join(s.replace("stall_", "") for s in top) + " |")
    print("|---|---|---|---|---|" + "---|" * len(top))
    for ln in sorted(agg):
        a = agg[ln]
        text = src[ln - 1].strip() if 0 < ln <= len(src) else "?"
        print(f"| {ln} | `{text[:60]}` | {100 * a['inst'] / tot['inst']:.2f} | {100 * a['fp64'] / max(tot['fp64'], 1):.2f} | {100 * a['samples'] / tot['samples']:.2f} | "
              + " | ".join(f"{100 * a[s] / tot['samples']:.2f}" for s in top) + " |")
    print("\n| all | | 100 | 100 | 100 | " + " | ".join(f"{100 * tot[s] / tot['samples']:.2f}" for s in top) + " |")
    if inner_too:
        print("\n## innermost source lines, top 40 by samples\n\n| file:line | % insts | % samples | wait | long_sb | short_sb | no_inst |\n|---|---|---|---|---|---|---|")
        for k, a in sorted(agg_in.items(), key=lambda kv: -kv[1]["samples"])[:40]:
            print(f"| {k} | {100 * a['inst'] / tot['inst']:.2f} | {100 * a['samples'] / tot['samples']:.2f} | {100 * a['stall_wait'] / tot['samples']:.2f} | "
                  f"{100 * a['stall_long_sb'] / tot['samples']:.2f} | {100 * a['stall_short_sb'] / tot['samples']:.2f} | {100 * a['stall_no_inst'] / tot['samples']:.2f} |")


if __name__ == "__main__":
    main()
