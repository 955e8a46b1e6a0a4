import sys, os, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, bench
from dash_b200 import api
B = 256
gcm, labels, _ = bench.build_material("MODEL_B_POOL_REPL", B)
api.init(0)
circ = api.Circuit.load(gcm, max_batch=B)
pin = []
for l in labels:
    t = torch.empty(l.shape, dtype=torch.int16).pin_memory(); t.numpy()[...] = l; pin.append(t)
in_np = [t.numpy() for t in pin]
out_t = [torch.empty((B, circ.output_size, api.nr_comps(q)), dtype=torch.int16).pin_memory() for q in circ.out_moduli]
out_np = [t.numpy() for t in out_t]
ip, op = api._ptr_array(in_np), api._ptr_array(out_np)
for r in range(3):
    t0 = time.perf_counter()
    api._check(circ.L.dashb_evaluate(circ.h, ip, B, op))
    print("wall %.2f ms" % ((time.perf_counter() - t0) * 1e3), file=sys.stderr)
