"""Static issue-cycle count of the loops of a cubin that contain a given opcode (default MUFU.RCP64H = the Newton solve):
2 cycles per FP64 instruction (DFMA / DMUL / DADD / DSETP hold the issue port for two, profiles/micro/icache_mix), 1 otherwise.
  python profiles/sass_loops.py file.cubin [opcode] [-v]"""
import re
import subprocess
import sys
from collections import Counter

cubin = sys.argv[1]
needle = sys.argv[2] if len(sys.argv) > 2 and not sys.argv[2].startswith("-") else "MUFU.RCP64H"
verbose = "-v" in sys.argv
out = subprocess.run(["cuobjdump", "-sass", cubin], capture_output=True, text=True).stdout
funcs, cur = {}, None
for line in out.splitlines():
    m = re.match(r"\s+Function : (\S+)", line)
    if m:
        cur = m.group(1); funcs[cur] = []
        continue
    m = re.match(r"\s+/\*([0-9a-f]{4,6})\*/\s+(.*?);", line)
    if m and cur:
        funcs[cur].append((int(m.group(1), 16), m.group(2).strip()))
for fn, ins in funcs.items():
    addr_idx = {a: i for i, (a, _) in enumerate(ins)}
    loops = []
    for i, (a, t) in enumerate(ins):
        m = re.search(r"\bBRA(?:\.U)?\b.*?(0x[0-9a-f]+)\s*$", t)
        if m:
            tgt = int(m.group(1), 16)
            if tgt <= a and tgt in addr_idx:
                loops.append((addr_idx[tgt], i))
    for lo, hi in loops:
        body = ins[lo:hi + 1]
        if not any(needle in t for _, t in body):
            continue
        if any(l2 > lo and h2 < hi and any(needle in t for _, t in ins[l2:h2 + 1]) for l2, h2 in loops):
            continue   # report the innermost loop holding the opcode
        ops = Counter()
        for _, t in body:
            op = t.split()[1] if t.startswith("@") else t.split()[0]
            ops[op.split(".")[0]] += 1
        fp64 = sum(v for k, v in ops.items() if k in ("DFMA", "DMUL", "DADD", "DSETP"))
        n = len(body)
        print(f"{fn[:40]}  loop {ins[lo][0]:#x}..{ins[hi][0]:#x}: {n} instructions, {fp64} FP64, issue cycles {2 * fp64 + (n - fp64)}  "
              + " ".join(f"{k}:{v}" for k, v in ops.most_common(14)))
        if verbose:
            for a, t in body:
                print(f"    {a:#07x}  {t}")
