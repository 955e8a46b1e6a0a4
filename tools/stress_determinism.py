"""Repeats one evaluation many times and reports runs whose output labels
differ from the first (a race shows up as a non-zero count)."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import workloads as W
from dash_b200 import api, garble
api.init(0)
spec = dict(input=(12, 12, 2),
            layers=[("conv", 16, 3, 1), ("relu",), ("conv", 32, 2, 2), ("relu",), ("flatten",), ("dense", 64), ("relu",), ("dense", 10)],
            quant=("scaleplus", 32), crt=W.SCALE_CRT, mrs=W.SCALE_MRS, max_modulus=32)
plain, n_in = W.plain_layers(spec, weight_seed=1)
batch = int(sys.argv[1]) if len(sys.argv) > 1 else 80
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 200
X = W.synthetic_inputs(spec, batch, seed=2)
circ, _ = garble.garble_to_circuit(spec["crt"], spec["mrs"], n_in, plain, batch, seed=b"smoke")
n_layers = circ.n_layers
first = None
bad = 0
where = {}
for r in range(reps):
    circ.garble_inputs(X)
    circ.evaluate_device()
    got = circ.move_outputs()
    if first is None:
        first = got
        continue
    if not all(np.array_equal(a, b) for a, b in zip(got, first)):
        bad += 1
        for j, (a, b) in enumerate(zip(got, first)):
            d = np.argwhere((a != b).any(axis=2))
            if len(d):
                where.setdefault(j, []).append((int(d[0][0]), int(d[0][1]), len(d)))
print("batch", batch, "reps", reps, "runs differing from the first:", bad, {k: v[:3] for k, v in where.items()})
