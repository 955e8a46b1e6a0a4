#!/usr/bin/env python
"""Per-source-line instruction and stall counts of one kernel from an ncu
report (captured with --set full --import-source on; library built with
-lineinfo).  The report's SASS page gives counters per instruction address;
nvdisasm -g on the cubin inside dash_b200/libdash_b200.so gives the source
line of every instruction offset; the two are joined on the offset.  The
library must be the build the report was taken from.

    python tools/ncu_lines.py gpurun_out/r2_ksign_src.ncu-rep k_signILb1ELi1 [top_n] [stall_column]

stall_column: e.g. stall_long_sb, stall_math, stall_wait (default: all samples)
"""
import collections
import csv
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "dash_b200", "libdash_b200.so")


def disassemble(tmp):
    subprocess.run(["cuobjdump", "-xelf", "all", LIB], cwd=tmp, check=True, capture_output=True)
    out = ""
    for f in sorted(os.listdir(tmp)):
        if f.endswith(".cubin"):
            out += subprocess.run(["nvdisasm", "-g", "-c", f], cwd=tmp, capture_output=True, text=True).stdout
    return out.splitlines()


def line_map(dis, pattern):
    """offset -> (file, line) for the first .text section whose name contains `pattern`"""
    start = [i for i, l in enumerate(dis) if l.startswith(".text.") and pattern in l and l.endswith(":")]
    if not start:
        sys.exit("no kernel matching %r in %s" % (pattern, LIB))
    m, cur = {}, None
    for l in dis[start[0] + 1:]:
        if l.startswith("\t.section"):
            break
        f = re.search(r'//## File "([^"]+)", line (\d+)', l)
        if f:
            cur = (os.path.basename(f.group(1)), int(f.group(2)))
            continue
        i = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*);", l)
        if i:
            m[int(i.group(1), 16)] = cur
    return m, dis[start[0]][6:-1]


def main():
    rep, pattern = sys.argv[1], sys.argv[2]
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
    stall = sys.argv[4] if len(sys.argv) > 4 else "# Samples"
    with tempfile.TemporaryDirectory() as tmp:
        lines, name = line_map(disassemble(tmp), pattern)
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    h = rows[1]
    ia, ie, isamp = h.index("Address"), h.index("Instructions Executed"), h.index(stall)
    base = int(rows[2][ia], 16)
    inst, samp = collections.Counter(), collections.Counter()
    for r in rows[2:]:
        ln = lines.get(int(r[ia], 16) - base)
        inst[ln] += int(r[ie])
        samp[ln] += int(r[isamp])
    ti, ts = sum(inst.values()), max(1, sum(samp.values()))
    src = {}
    print("%s\n%s\nwarp-instructions %d, %s %d" % (rows[0][1], name, ti, stall, ts))
    order = inst if stall == "# Samples" else samp
    for ln, _ in order.most_common(top):
        n = inst[ln]
        text = ""
        if ln:
            if ln[0] not in src:
                p = os.path.join(ROOT, "dash_b200", "csrc", ln[0])
                src[ln[0]] = open(p).read().splitlines() if os.path.exists(p) else []
            if 0 < ln[1] <= len(src[ln[0]]):
                text = src[ln[0]][ln[1] - 1].strip()[:90]
        print("%-16s:%5s %5.1f%% inst %5.1f%% %s  %s" % (ln[0] if ln else "?", ln[1] if ln else "", 100.0 * n / ti,
                                                          100.0 * samp[ln] / ts, "samp" if stall == "# Samples" else stall, text))


if __name__ == "__main__":
    main()
