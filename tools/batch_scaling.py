#!/usr/bin/env python
"""Throughput against batch size for one model garbled on the GPU
(B in {1, 16, 256, 4096}).  One JSON line per batch.
    python tools/batch_scaling.py MODEL_B_POOL_REPL 1 16 256 1024 4096"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

import workloads as W  # noqa: E402
from dash_b200 import api, garble  # noqa: E402


def main():
    name = sys.argv[1]
    batches = [int(x) for x in sys.argv[2:]] or [1, 16, 256, 4096]
    api.init(0)
    spec = W.MODELS[name]
    plain, n_in = W.plain_layers(spec)
    stream = torch.cuda.ExternalStream(api.stream_handle())
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for B in batches:
        c, _ = garble.garble_to_circuit(spec["crt"], spec.get("mrs", []), n_in, plain, B)
        c.garble_inputs(W.synthetic_inputs(spec, B, seed=1))
        for _ in range(3):
            c.evaluate_device()
        steps, ms = 10, 0.0
        for _ in range(steps):
            flush.zero_()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            c.evaluate_device()
            e1.record(stream)
            e1.synchronize()
            ms += e0.elapsed_time(e1)
        ms /= steps
        _, ok = c.decode_outputs()
        print(json.dumps({"model": name, "batch": B, "ms_per_step": ms, "inferences_per_s": B / (ms / 1e3),
                          "all_outputs_decode": bool(ok.all())}), flush=True)
        c.close()


if __name__ == "__main__":
    main()
