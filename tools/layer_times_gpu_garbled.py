#!/usr/bin/env python
"""Per-layer device times of a model garbled on the GPU (no reference needed):
    python tools/layer_times_gpu_garbled.py MODEL_F_GNNP_POOL_REPL 64"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import workloads as W  # noqa: E402
from dash_b200 import api, garble  # noqa: E402


def main():
    name, B = sys.argv[1], int(sys.argv[2])
    api.init(0)
    spec = W.MODELS[name]
    plain, n_in = W.plain_layers(spec)
    c, _ = garble.garble_to_circuit(spec["crt"], spec.get("mrs", []), n_in, plain, B)
    c.garble_inputs(W.synthetic_inputs(spec, B, seed=1))
    for _ in range(3):
        c.evaluate_device()
    c.set_layer_timing(True)
    acc, n = None, 3
    for _ in range(n):
        c.evaluate_device()
        t = c.layer_times_ms()
        acc = t if acc is None else [(a[0], a[1] + b[1]) for a, b in zip(acc, t)]
    by = {}
    for a, v in acc:
        by[a] = by.get(a, 0.0) + v / n
    print(json.dumps({"model": name, "batch": B, "by_type_ms": {k: round(v, 3) for k, v in by.items()},
                      "layers": [(a, round(v / n, 3)) for a, v in acc], "total_ms": round(sum(v for _, v in acc) / n, 3)}))


if __name__ == "__main__":
    main()
