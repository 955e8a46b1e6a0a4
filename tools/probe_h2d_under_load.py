"""Pinned-host -> device copy rate while the evaluation kernels run: is the
pipelined e2e bound by the link, or does the link slow down under compute?
   python tools/probe_h2d_under_load.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import bench  # noqa: E402
from dash_b200 import api  # noqa: E402

B = 256
gcm, labels, _ = bench.build_material("MODEL_B_POOL_REPL", B)
api.init(0)
circ = api.Circuit.load(gcm, max_batch=B)
circ.move_inputs(labels)
n = 188 << 20
h = torch.empty(n, dtype=torch.uint8).pin_memory()
d = torch.empty(n, dtype=torch.uint8, device="cuda")
cs = torch.cuda.Stream()


def copies(reps, chunks):
    step = n // chunks
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(cs):
        e0.record()
        for _ in range(reps):
            for c in range(chunks):
                d[c * step:(c + 1) * step].copy_(h[c * step:(c + 1) * step], non_blocking=True)
        e1.record()
    return e0, e1


for chunks in (1, 36):
    for load in (False, True):
        torch.cuda.synchronize()
        if load:
            for _ in range(3):
                circ.evaluate_device_nosync() if hasattr(circ, "evaluate_device_nosync") else None
        e0, e1 = copies(10, chunks)
        if load:
            for _ in range(12):
                circ.evaluate_device()  # blocks the host per step; the copies were all enqueued before
        e1.synchronize()
        ms = e0.elapsed_time(e1) / 10
        print("chunks %2d, evaluation running: %-5s  %.2f ms per 188 MiB = %.1f GB/s" % (chunks, load, ms, n / ms / 1e6), flush=True)
