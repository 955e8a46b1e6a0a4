"""The smoke sequence (garble -> file -> load -> garble_inputs -> evaluate ->
compare with the oracle) repeated; prints which side deviates if any."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import workloads as W
from dash_b200 import api, garble, gcm
from oracle import oraclebind
api.init(0)
spec = dict(input=(12, 12, 2),
            layers=[("conv", 16, 3, 1), ("relu",), ("conv", 32, 2, 2), ("relu",), ("flatten",), ("dense", 64), ("relu",), ("dense", 10)],
            quant=("scaleplus", 32), crt=W.SCALE_CRT, mrs=W.SCALE_MRS, max_modulus=32)
plain, n_in = W.plain_layers(spec, weight_seed=1)
batch = 80
X = W.synthetic_inputs(spec, batch, seed=2)
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 30
files = []
for r in range(reps):
    path = "/tmp/stress_%d.gcm" % (r % 2)
    garble.garble_to_file(spec["crt"], spec["mrs"], n_in, plain, path, seed=b"smoke")
    blob = open(path, "rb").read()
    files.append(hash(blob))
    circ = api.Circuit.load(path, max_batch=batch)
    circ.garble_inputs(X)
    circ.evaluate_device()
    got = circ.move_outputs()
    g = gcm.read_gcm(path)
    oc = oraclebind.OracleCircuit(path)
    lab = gcm.encode_inputs(g, X[0])
    w1 = oc.evaluate(lab, threads=1)
    w2 = oc.evaluate(lab, threads=2)
    e_gpu = all(np.array_equal(got[j][0], w1[j]) for j in range(7))
    e_or = all(np.array_equal(w1[j], w2[j]) for j in range(7))
    if not (e_gpu and e_or):
        print("rep", r, "gpu==oracle1", e_gpu, "oracle1==oracle2", e_or, "file same as first", files[r] == files[0])
    circ.close()
print("done", reps, "distinct files", len(set(files)))
