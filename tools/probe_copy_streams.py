#!/usr/bin/env python
"""End-to-end dashb_evaluate (host int16 labels in, host labels out) of the
benchmark model against the number of H2D copy streams (DASHB_COPY_STREAMS,
read at finalize); checks that every setting returns the same labels.
    python tools/probe_copy_streams.py [1 2 3 4]"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402

import bench  # noqa: E402
from dash_b200 import api  # noqa: E402

B = 256
settings = [int(x) for x in sys.argv[1:]] or [1, 2, 3, 4]
gcm, labels, _ = bench.build_material("MODEL_B_POOL_REPL", B)
api.init(0)
pin = []
for l in labels:
    t = torch.empty(l.shape, dtype=torch.int16).pin_memory()
    t.numpy()[...] = l
    pin.append(t)
in_np = [t.numpy() for t in pin]
first = None
for n in settings:
    os.environ["DASHB_COPY_STREAMS"] = str(n)
    circ = api.Circuit.load(gcm, max_batch=B)
    out_t = [torch.empty((B, circ.output_size, api.nr_comps(q)), dtype=torch.int16).pin_memory() for q in circ.out_moduli]
    out_np = [t.numpy() for t in out_t]
    ip, op = api._ptr_array(in_np), api._ptr_array(out_np)
    for _ in range(3):
        api._check(circ.L.dashb_evaluate(circ.h, ip, B, op))
    steps = 30
    t0 = time.perf_counter()
    for _ in range(steps):
        api._check(circ.L.dashb_evaluate(circ.h, ip, B, op))
    ms = (time.perf_counter() - t0) * 1e3 / steps
    got = [o.copy() for o in out_np]
    same = True if first is None else all(np.array_equal(a, b) for a, b in zip(first, got))
    if first is None:
        first = got
    print(json.dumps({"copy_streams": n, "e2e_ms_per_step": round(ms, 3), "inferences_per_s": round(B / ms * 1e3),
                      "same_labels_as_first": bool(same)}), flush=True)
    circ.close()
