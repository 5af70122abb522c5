#!/usr/bin/env python3
"""Per-source-line summary of an ncu report (needs -lineinfo): instructions executed and stall samples per line.
usage: tools/ncu_lines.py report.ncu-rep kernel_regex [top_n]"""
import csv
import subprocess
import sys


def main():
    rep, kern = sys.argv[1], sys.argv[2]
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv", "--kernel-name",
                          "regex:" + kern], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hi = [i for i, r in enumerate(rows) if "Instructions Executed" in r][0]
    h = rows[hi]
    ie, si = h.index("Instructions Executed"), h.index("# Samples")
    data = []
    for r in rows[hi + 1:]:
        if len(r) > ie and r[2] == "-":  # source-line rows (SASS rows carry an address)
            try:
                data.append((int(r[ie]), int(r[si]), r[0], r[1].strip()[:120]))
            except ValueError:
                pass
    tot, ts = sum(d[0] for d in data) or 1, sum(d[1] for d in data) or 1
    print("total warp instructions %d, samples %d" % (tot, ts))
    for d in sorted(data, key=lambda d: -d[1])[:top]:
        print("%5.1f%% inst %5.1f%% samples  L%-4s %s" % (100 * d[0] / tot, 100 * d[1] / ts, d[2], d[3]))


if __name__ == "__main__":
    main()
