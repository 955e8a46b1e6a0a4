#!/usr/bin/env python
"""One rescale layer (factor 32, CIFAR base) garbled on the GPU and evaluated a few times: a profiler target."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import workloads as W
from dash_b200 import api, garble
api.init(0)
n, B = 16384, 64
plain = [{"type": "rescale", "factors": (32,)}]
c, _ = garble.garble_to_circuit(W.SCALE_CRT, W.SCALE_MRS, n, plain, B)
c.garble_inputs(np.random.default_rng(0).integers(0, 256, size=(B, n), dtype=np.int64))
for _ in range(3):
    c.evaluate_device()
c.sync()
c.set_layer_timing(True)
c.evaluate_device()
print(c.layer_times_ms())
