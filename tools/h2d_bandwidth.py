#!/usr/bin/env python
"""Raw pinned-host -> device copy bandwidth of the box, at the sizes bench.py's
e2e leg moves per step (187.9 MB as int16 digits, 28.9 MB compressed): tells
whether `dashb_evaluate` from host buffers is bound by the link or by the
library.   python tools/h2d_bandwidth.py"""
import json

import torch

for mb in (5, 29, 47, 188, 1024):
    n = mb << 20
    h = torch.empty(n, dtype=torch.uint8).pin_memory()
    d = torch.empty(n, dtype=torch.uint8, device="cuda")
    for _ in range(3):
        d.copy_(h, non_blocking=True)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 10
    e0.record()
    for _ in range(reps):
        d.copy_(h, non_blocking=True)
    e1.record()
    e1.synchronize()
    ms = e0.elapsed_time(e1) / reps
    h2d = n / ms / 1e6
    e0.record()
    for _ in range(reps):
        h.copy_(d, non_blocking=True)
    e1.record()
    e1.synchronize()
    ms2 = e0.elapsed_time(e1) / reps
    print(json.dumps({"MiB": mb, "h2d_GBps": round(h2d, 2), "d2h_GBps": round(n / ms2 / 1e6, 2)}), flush=True)
