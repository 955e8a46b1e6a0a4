"""Probe: how long do the H2D copies take, and do they block the host?"""
import sys, time, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from dash_b200 import api

B = 256
gcm, labels, _ = bench.build_material("MODEL_B_POOL_REPL", B)
api.init(0)
circ = api.Circuit.load(gcm, max_batch=B)
pinned = []
for l in labels:
    t = torch.empty(l.shape, dtype=torch.int16).pin_memory()
    t.numpy()[...] = l
    pinned.append(t)
in_np = [t.numpy() for t in pinned]
in_ptrs = api._ptr_array(in_np)
stream = torch.cuda.ExternalStream(api.stream_handle())
L = circ.L
for name, arrs in (("pinned", in_np), ("pageable", [np.ascontiguousarray(l) for l in labels])):
    ptrs = api._ptr_array(arrs)
    for rep in range(3):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record(stream)
        api._check(L.dashb_move_inputs(circ.h, ptrs, B))
        t1 = time.perf_counter()
        e1.record(stream)
        e1.synchronize()
        t2 = time.perf_counter()
        print(name, "host enqueue %.2f ms, device %.2f ms, wall %.2f ms" % ((t1 - t0) * 1e3, e0.elapsed_time(e1), (t2 - t0) * 1e3))
torch.cuda.synchronize()
t0 = time.perf_counter()
api._check(L.dashb_evaluate_device(circ.h))
t1 = time.perf_counter()
stream.synchronize()
t2 = time.perf_counter()
print("evaluate_device host enqueue %.2f ms, wall %.2f ms" % ((t1 - t0) * 1e3, (t2 - t0) * 1e3))
