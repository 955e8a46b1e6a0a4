#!/usr/bin/env python
"""Per-layer device times (CUDA events, lanes serialised) of one model.

    python tools/layer_times.py [MODEL] [BATCH] [LIB.so ...]

With extra arguments, each is a build of libdash_b200.so to time in turn (the
library is copied over dash_b200/libdash_b200.so in a fresh process)."""
import json
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def one(model, batch):
    import numpy as np
    import workloads as W
    from dash_b200 import api

    api.init(0)
    spec = W.MODELS[model]
    ref = W.build_reference_circuit(spec, seed=42, weight_seed=42)
    path = "/tmp/lt_%s.gcm" % model
    ref.export(path)
    X = W.synthetic_inputs(spec, min(batch, 32), seed=42)
    per = [ref.garble_inputs(x) for x in X]
    reps = -(-batch // len(X))
    labels = [np.concatenate([np.stack([p[j] for p in per])] * reps)[:batch].copy() for j in range(ref.k)]
    c = api.Circuit.load(path, max_batch=batch)
    c.move_inputs(labels)
    for _ in range(3):
        c.evaluate_device()
    c.set_layer_timing(True)
    acc = None
    n = 5
    for _ in range(n):
        c.evaluate_device()
        t = c.layer_times_ms()
        acc = t if acc is None else [(a[0], a[1] + b[1]) for a, b in zip(acc, t)]
    print(json.dumps({"model": model, "batch": batch, "layers": [(a, round(v / n, 4)) for a, v in acc],
                      "total_ms": round(sum(v for _, v in acc) / n, 4)}))


if __name__ == "__main__":
    model = sys.argv[1] if len(sys.argv) > 1 else "MODEL_B_POOL_REPL"
    batch = int(sys.argv[2]) if len(sys.argv) > 2 else 256
    libs = sys.argv[3:]
    if not libs:
        one(model, batch)
    else:
        for lib in libs:
            shutil.copy(lib, os.path.join(ROOT, "dash_b200", "libdash_b200.so"))
            print("#", lib, flush=True)
            subprocess.check_call([sys.executable, os.path.abspath(__file__), model, str(batch)])
