#!/usr/bin/env python
"""Writes profiles/r2_traffic.json from one `ncu --set full` capture of
k_sign<relu> (read here with the ncu CLI): DRAM bytes, issue-slot utilisation
and warp instructions per (element, warp), tagged with the git blob hash of
the kernels.cu it was captured from -- bench.py withholds `roofline.traffic`
when that hash is not the tree's.

    python tools/make_traffic.py gpurun_out/r2_ksign_big.ncu-rep ELEMENTS INFERENCES
"""
import csv
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

rep, elements, inferences = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
rows = list(csv.reader(out.splitlines()))
h, v = rows[0], rows[2]
m = {n: v[i] for i, n in enumerate(h)}
units = {n: rows[1][i] for i, n in enumerate(h)}


def num(name):
    x = float(m[name].replace(",", ""))
    u = units[name].lower()
    scale = {"kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9, "byte": 1.0}.get(u, 1.0)
    return x * scale


grid = int(m["Grid Size"].strip("()").split(",")[0])
block = int(m["Block Size"].strip("()").split(",")[0])
assert grid * block >= elements * inferences > (grid - 1) * block, "the capture is not the launch you named"
warps = grid * block // 32
dram = num("dram__bytes_read.sum") + num("dram__bytes_write.sum")
j = {
    "kernels_cu_blob": bench.git_blob_hash(os.path.join(ROOT, "dash_b200", "csrc", "kernels.cu")),
    "source": os.path.basename(rep) + " (ncu --set full --clock-control none, one launch of the concurrent bench step)",
    "k_sign_relu": {
        "launch": "%d elements x %d inferences, grid %d x %d threads" % (elements, inferences, grid, block),
        "duration_us": num("gpu__time_duration.sum"),
        "dram_bytes_per_launch": dram,
        "dram_bytes_per_element_inference": dram / (elements * inferences),
        "algorithmic_bytes_per_element_inference_k9": 3120,
        "issue_slot_frac": num("smsp__issue_active.avg.pct_of_peak_sustained_active") / 100.0,
        "warp_instructions_per_element": num("smsp__inst_executed.sum") / warps,
        "registers_per_thread": int(float(m["launch__registers_per_thread"])),
        "warps_active_frac": num("sm__warps_active.avg.pct_of_peak_sustained_active") / 100.0,
        "pipe_alu_frac": num("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active") / 100.0,
        "pipe_fma_frac": num("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active") / 100.0,
        "pipe_lsu_frac": num("sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active") / 100.0,
    },
}
json.dump(j, open(os.path.join(ROOT, "profiles", "r2_traffic.json"), "w"), indent=1)
print(json.dumps(j, indent=1))
