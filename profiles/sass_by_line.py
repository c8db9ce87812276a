#!/usr/bin/env python
"""Attribute an `ncu --page source --csv` (SASS view) dump to CUDA source lines.

ncu's CSV export of the SASS view carries no line column; `nvdisasm -g -c` of the same cubin
does.  Both list the kernel's instructions in address order, so they are matched by position.

usage: sass_by_line.py <ncu_sass.csv> <nvdisasm.txt> <kernel-name-substring> [top N]
"""
import csv
import re
import sys
from collections import defaultdict


def main():
    sass_csv, disasm, kname = sys.argv[1:4]
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    rows = list(csv.reader(open(sass_csv)))
    hdr = rows[1]
    col = {h: i for i, h in enumerate(hdr)}
    ins = rows[2:]
    # nvdisasm: find the kernel's .text section, collect (line, instruction)
    lines = open(disasm).read().split("\n")
    in_k = False
    cur = None
    seq = []
    for l in lines:
        if l.startswith(".section") or l.lstrip().startswith(".section"):
            in_k = (".text." in l) and (kname in l)
            continue
        if not in_k:
            continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', l)
        if m:
            inl = "inlined" in l
            cur = (m.group(1).split("/")[-1], int(m.group(2)))
            continue
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
        if m:
            seq.append((cur, m.group(2).strip()))
    print("ncu instructions: %d, nvdisasm instructions: %d" % (len(ins), len(seq)))
    n = min(len(ins), len(seq))
    by = defaultdict(lambda: [0, 0, 0])
    tot = [0, 0]
    for i in range(n):
        ex = int(ins[i][col["Instructions Executed"]] or 0)
        sm = int(ins[i][col["# Samples"]] or 0)
        k = seq[i][0]
        by[k][0] += ex
        by[k][1] += sm
        by[k][2] += 1
        tot[0] += ex
        tot[1] += sm
    print("total warp-instr %d, samples %d" % tuple(tot))
    print("%-28s %14s %7s %9s %7s %6s" % ("file:line", "warp-instr", "%", "samples", "%", "sass"))
    for k, v in sorted(by.items(), key=lambda kv: -kv[1][1])[:top]:
        print("%-28s %14d %6.2f%% %9d %6.2f%% %6d" % (
            "%s:%d" % k if k else "?", v[0], 100.0 * v[0] / max(1, tot[0]), v[1],
            100.0 * v[1] / max(1, tot[1]), v[2]))


if __name__ == "__main__":
    main()
