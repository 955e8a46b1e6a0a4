"""Pipelined end-to-end rate (dashb_evaluate_async, two steps in flight) of
MODEL_B batch 256 from pinned host buffers, next to the device-resident step:
   python tools/probe_async.py   (env: DASHB_LANES, DASHB_INGEST_CTAS, DASHB_DIRECT_INGEST, DASHB_LANE_SPLIT)"""
import ctypes as C
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import bench  # noqa: E402
from dash_b200 import api  # noqa: E402

B = 256
gcm, labels, _ = bench.build_material("MODEL_B_POOL_REPL", B)
api.init(0)
circ = api.Circuit.load(gcm, max_batch=B)
L = circ.L
L.dashb_evaluate_async.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p]
L.dashb_evaluate_wait.argtypes = [C.c_void_p]
pin = []
for l in labels:
    t = torch.empty(l.shape, dtype=torch.int16).pin_memory()
    t.numpy()[...] = l
    pin.append(t)
ip = api._ptr_array([t.numpy() for t in pin])
outs = [[torch.empty((B, circ.output_size, api.nr_comps(q)), dtype=torch.int16).pin_memory() for q in circ.out_moduli] for _ in range(2)]
ops = [api._ptr_array([t.numpy() for t in o]) for o in outs]
circ.move_inputs([t.numpy() for t in pin])
for _ in range(5):
    circ.evaluate_device()
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(30):
    circ.evaluate_device()
torch.cuda.synchronize()
dev = (time.perf_counter() - t0) / 30 * 1e3
for i in range(6):
    api._check(L.dashb_evaluate_async(circ.h, ip, B, ops[i & 1]))
api._check(L.dashb_evaluate_wait(circ.h))
t0 = time.perf_counter()
N = 40
tq = 0.0
for i in range(N):
    t1 = time.perf_counter()
    api._check(L.dashb_evaluate_async(circ.h, ip, B, ops[i & 1]))
    tq += time.perf_counter() - t1
api._check(L.dashb_evaluate_wait(circ.h))
e2e = (time.perf_counter() - t0) / N * 1e3
print("env %s: device-resident %.3f ms/step, pipelined e2e %.3f ms/step (ratio %.3f), host time inside the async call %.3f ms" % (
    {k: v for k, v in os.environ.items() if k.startswith("DASHB_") and k != "DASHB_CACHE"}, dev, e2e, dev / e2e, tq / N * 1e3), flush=True)
