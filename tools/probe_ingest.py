"""End-to-end input path of dashb_evaluate on MODEL_B batch 256: how long the
labels take from pinned host memory into the digit planes -- staged
(cudaMemcpyAsync + k_ingest) against direct (k_ingest_direct reads the mapped
host buffers) -- alone and under the pipelined evaluation (DASHB_TRACE).
   python tools/probe_ingest.py [direct|staged]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
mode = sys.argv[1] if len(sys.argv) > 1 else "direct"
os.environ["DASHB_DIRECT_INGEST"] = "1" if mode == "direct" else "0"
os.environ["DASHB_TRACE"] = "1"
import torch  # noqa: E402

import bench  # noqa: E402
from dash_b200 import api  # noqa: E402

B = 256
gcm, labels, _ = bench.build_material("MODEL_B_POOL_REPL", B)
api.init(0)
circ = api.Circuit.load(gcm, max_batch=B)
pin = []
for l in labels:
    t = torch.empty(l.shape, dtype=torch.int16).pin_memory()
    t.numpy()[...] = l
    pin.append(t)
in_np = [t.numpy() for t in pin]
nbytes = sum(a.nbytes for a in in_np)
out_t = [torch.empty((B, circ.output_size, api.nr_comps(q)), dtype=torch.int16).pin_memory() for q in circ.out_moduli]
ip, op = api._ptr_array(in_np), api._ptr_array([t.numpy() for t in out_t])
stream = torch.cuda.ExternalStream(api.stream_handle())
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
# the input path alone
for r in range(6):
    flush.zero_()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    api._check(circ.L.dashb_move_inputs(circ.h, ip, B))
    e1.record(stream)
    e1.synchronize()
    ms = e0.elapsed_time(e1)
    if r >= 3:
        print("%s: move_inputs alone %.3f ms = %.1f GB/s" % (mode, ms, nbytes / ms / 1e6), file=sys.stderr)
for r in range(5):
    flush.zero_()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    api._check(circ.L.dashb_evaluate(circ.h, ip, B, op))
    print("%s: dashb_evaluate wall %.2f ms" % (mode, (time.perf_counter() - t0) * 1e3), file=sys.stderr)
