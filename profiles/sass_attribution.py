"""Static attribution of a generated kernel's SASS to the source functions it was inlined from: `nvdisasm -g` on the NVRTC cubin
(compiled with -lineinfo) prints a `//## File "...", line N` marker in front of every run of instructions; the line is mapped to
the enclosing __device__ function of the embedded header. Needs no GPU.

  python profiles/sass_attribution.py <cubin> [function]   ->  markdown table on stdout
  (cubins: optimo_b200/csrc/jit_cache/*.cubin, or profiles/exp_prebuild.py <workload> <dump_prefix>)
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC_DIRS = [os.path.join(ROOT, "optimo_b200", "csrc")]
FUNC = re.compile(r"\s*(?:template\s*<.*>)?\s*(?:static\s+)?(?:__device__|__global__).*?\b([A-Za-z_0-9]+)\s*\(")
INSTR = re.compile(r"^\s+/\*[0-9a-f]{4,6}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)")


def enclosing(src, fname, line):
    lines = src.get(fname)
    if not lines:
        return fname
    for i in range(min(line, len(lines)) - 1, -1, -1):
        m = FUNC.match(lines[i])
        if m and not lines[i].strip().startswith("//"):
            return f"{fname}:{m.group(1)}"
    return fname


def main():
    cubin = sys.argv[1]
    kernel = sys.argv[2] if len(sys.argv) > 2 else "omo_jit_kernel"
    dis = subprocess.run(["nvdisasm", "-g", cubin], capture_output=True, text=True, check=True).stdout.splitlines()
    start = next(i for i, l in enumerate(dis) if l.startswith(f".text.{kernel}:"))
    per_line, fp64 = collections.Counter(), collections.Counter()
    cur = None
    for l in dis[start + 1:]:
        if l.startswith("\t.section"):
            break
        m = re.match(r'\s*//## File "([^"]+)", line (\d+)', l)
        if m:
            cur = (os.path.basename(m.group(1)), int(m.group(2)))
            continue
        m = INSTR.match(l)
        if m:
            per_line[cur] += 1
            if m.group(1) in ("DFMA", "DMUL", "DADD", "DSETP"):
                fp64[cur] += 1
    src = {}
    for f in {k[0] for k in per_line if k}:
        for d in SRC_DIRS:
            if os.path.exists(os.path.join(d, f)):
                src[f] = open(os.path.join(d, f)).read().splitlines()
    by_fn, by_fn64 = collections.Counter(), collections.Counter()
    for key, v in per_line.items():
        fn = enclosing(src, *key) if key else "(no line info)"
        by_fn[fn] += v
        by_fn64[fn] += fp64[key]
    total = sum(by_fn.values())
    print(f"`{kernel}` of `{os.path.basename(cubin)}`: {total} instructions, {sum(by_fn64.values())} of them FP64\n")
    print("| inlined from | instructions | share | FP64 |\n|---|---|---|---|")
    for fn, v in by_fn.most_common():
        if v * 200 >= total:
            print(f"| `{fn}` | {v} | {100.0 * v / total:.1f} % | {by_fn64[fn]} |")


if __name__ == "__main__":
    main()
