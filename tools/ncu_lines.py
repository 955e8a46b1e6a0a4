#!/usr/bin/env python
"""Aggregates an ncu report's per-SASS-instruction counters by CUDA source
line (needs -lineinfo and --import-source on).  Usage:
    python tools/ncu_lines.py gpurun_out/prof.ncu-rep [top_n]
"""
import collections
import csv
import subprocess
import sys


def main():
    rep = sys.argv[1]
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    cur = None
    inst = collections.Counter()
    samp = collections.Counter()
    text = {}
    col_i = col_s = None
    for r in rows:
        if len(r) == 2 and r[0] == "File Path":
            cur = r[1].split("/")[-1]
            continue
        if len(r) > 7 and r[0] == "Line No":
            col_i = r.index("Instructions Executed")
            col_s = r.index("# Samples")
            continue
        if col_i and len(r) > col_i and r[0].isdigit():
            key = (cur, int(r[0]))
            text[key] = r[1].strip()
            try:
                inst[key] += int(r[col_i])
                samp[key] += int(r[col_s])
            except ValueError:
                pass
    ti, ts = sum(inst.values()), sum(samp.values())
    print("total warp-instructions %d, samples %d" % (ti, ts))
    for key, v in inst.most_common(top):
        print("%-16s:%4d %5.1f%% inst %5.1f%% samp  %s" % (key[0], key[1], 100.0 * v / ti, 100.0 * samp[key] / max(ts, 1), text[key][:88]))


if __name__ == "__main__":
    main()
