#!/usr/bin/env python3
"""Warp instructions and stall samples of an ncu report (-lineinfo) per named source region of render.cu.
usage: tools/ncu_regions.py report.ncu-rep kernel_regex "name=marker text" ...   (regions start at the first line containing
the marker text and end at the next region's start)"""
import csv
import os
import subprocess
import sys


def main():
    rep, kern = sys.argv[1], sys.argv[2]
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv", "--kernel-name",
                          "regex:" + kern], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    cur, ie, si = None, None, None
    per = {}
    for r in rows:
        if r and r[0] in ("File Name", "File Path"):
            cur = os.path.basename(r[1])
            continue
        if "Instructions Executed" in r:
            ie, si = r.index("Instructions Executed"), r.index("# Samples")
            continue
        if ie is None or len(r) <= ie or r[2] != "-":
            continue
        try:
            per[(cur, int(r[0]))] = (int(r[ie]), int(r[si]))
        except ValueError:
            pass
    src = open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "atray_b200", "csrc", "render.cu")).read().split("\n")
    marks = []
    for a in sys.argv[3:]:
        name, pat = a.split("=", 1)
        line = next(i + 1 for i, l in enumerate(src) if pat in l)
        marks.append((line, name))
    marks.sort()
    marks.append((len(src) + 1, "end"))
    tot = sum(v[0] for v in per.values()) or 1
    ts = sum(v[1] for v in per.values()) or 1
    print("attributed warp instructions %.1fM (inlined header lines count once per file), samples %d" % (tot / 1e6, ts))
    files = {}
    for (f, l), v in per.items():
        files[f] = tuple(a + b for a, b in zip(files.get(f, (0, 0)), v))
    for f, v in sorted(files.items(), key=lambda kv: -kv[1][0]):
        print("  file %-28s inst %5.1f%%  samples %5.1f%%" % (f, 100.0 * v[0] / tot, 100.0 * v[1] / ts))
    first = marks[0][0]
    i0 = sum(v[0] for (f, l), v in per.items() if f == "render.cu" and l < first)
    s0 = sum(v[1] for (f, l), v in per.items() if f == "render.cu" and l < first)
    print("  %-34s L1-%d inst %5.1f%% samples %5.1f%%" % ("(before the first region)", first, 100.0 * i0 / tot, 100.0 * s0 / ts))
    for (a, n), (b, _) in zip(marks, marks[1:]):
        i = sum(v[0] for (f, l), v in per.items() if f == "render.cu" and a <= l < b)
        s = sum(v[1] for (f, l), v in per.items() if f == "render.cu" and a <= l < b)
        print("  %-34s L%d-%d inst %5.1f%% (%.1fM) samples %5.1f%%" % (n, a, b, 100.0 * i / tot, i / 1e6, 100.0 * s / ts))


if __name__ == "__main__":
    main()
