#!/usr/bin/env python
"""Static SASS size of one kernel by source line / file, from `nvdisasm --print-line-info-inline` of a cubin
(cuobjdump -xelf all librptx.so), optionally joined with the executed-instruction counts of an ncu source page
(`ncu -i rep --page source --csv --print-source sass,cuda`) to show which lines make up the HOT code bytes.
usage: sass_lines.py <cubin> <kernel-substring> [ncu_src.csv]"""
import collections
import csv
import re
import subprocess
import sys


def static_counts(cubin, kernel):
    text = subprocess.run(["nvdisasm", "--print-line-info-inline", cubin], capture_output=True, text=True).stdout.splitlines()
    cnt, cur, inside = collections.Counter(), None, False
    for line in text:
        if line.lstrip().startswith(".section"):
            inside = ".text." in line and kernel in line
            continue
        if not inside:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', line)
        if m:
            if "inlined at" not in line:  # innermost location of the instruction that follows
                cur = (m.group(1).split("/")[-1], int(m.group(2)))
            continue
        if re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+\S", line) and cur:
            cnt[cur] += 1
    return cnt


def executed(path):
    out, fname, hdr = {}, None, None
    for r in csv.reader(open(path)):
        if not r:
            continue
        if r[0] == "File Path":
            fname = r[1].split("/")[-1]
            continue
        if r[0] == "Line No":
            hdr = {h: i for i, h in enumerate(r)}
            continue
        if hdr is None or not r[0].isdigit():
            continue
        try:
            out[(fname, int(r[0]))] = float(r[hdr["Instructions Executed"]] or 0)
        except (ValueError, IndexError):
            pass
    return out


def main():
    cnt = static_counts(sys.argv[1], sys.argv[2])
    tot = sum(cnt.values())
    print(f"static instructions {tot} = {tot * 16 / 1024:.1f} KB")
    byfile = collections.Counter()
    for (f, _), c in cnt.items():
        byfile[f] += c
    print("by file:", [(f, f"{c * 16 / 1024:.1f} KB") for f, c in byfile.most_common()])
    ex = executed(sys.argv[3]) if len(sys.argv) > 3 else None
    if ex:
        hot = {k: c for k, c in cnt.items() if ex.get(k, 0) > 0}
        print(f"lines executed at least once: {sum(hot.values()) * 16 / 1024:.1f} KB")
        hf = collections.Counter()
        for (f, _), c in hot.items():
            hf[f] += c
        print("hot by file:", [(f, f"{c * 16 / 1024:.1f} KB") for f, c in hf.most_common()])
        print("largest hot lines (static instr, executed warp-inst):")
        for k, c in sorted(hot.items(), key=lambda kv: -kv[1])[:45]:
            print(f"  {c:5d} instr  {ex[k]:.3g} exec  {k[0]}:{k[1]}")
    else:
        for k, c in cnt.most_common(40):
            print(f"  {c:5d} instr  {k[0]}:{k[1]}")


if __name__ == "__main__":
    main()
