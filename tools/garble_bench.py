#!/usr/bin/env python
"""Offline phase on the GPU: garbles a model with dash_b200.garble (synthetic
quantised weights of the model's shape), evaluates a batch on the GPU
evaluator and checks one inference label for label against the pinned CPU
oracle evaluating the same GPU-garbled material.  One JSON line per model.

    python tools/garble_bench.py MODEL_B_POOL_REPL 256
    python tools/garble_bench.py MODEL_F_GNNP_POOL_REPL 16 [--ref-garble-s 172.3]
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

import workloads as W  # noqa: E402
from dash_b200 import api, garble, gcm  # noqa: E402


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "MODEL_B_POOL_REPL"
    B = int(sys.argv[2]) if len(sys.argv) > 2 else 256
    check = "--no-oracle" not in sys.argv
    ref_s = float(sys.argv[sys.argv.index("--ref-garble-s") + 1]) if "--ref-garble-s" in sys.argv else None
    spec = W.MODELS[name]
    api.init(0)
    plain, input_size = W.plain_layers(spec)
    path = "/tmp/garbled_%s.gcm" % name
    garble.garble_to_file(spec["crt"], spec.get("mrs", []), input_size, plain[:1], "/tmp/warm.gcm")  # context / module warm-up
    t0 = time.time()
    st = garble.garble_to_file(spec["crt"], spec.get("mrs", []), input_size, plain, path, max_modulus=spec.get("max_modulus", 0))
    wall = time.time() - t0
    t0 = time.time()
    c = api.Circuit.load(path, max_batch=B)
    load_s = time.time() - t0
    c.close()
    # the same circuit, garbled straight into the evaluator (tables stay on the device)
    torch.cuda.synchronize()
    t0 = time.time()
    c, st2 = garble.garble_to_circuit(spec["crt"], spec.get("mrs", []), input_size, plain, B, max_modulus=spec.get("max_modulus", 0))
    torch.cuda.synchronize()
    direct_s = time.time() - t0
    X = W.synthetic_inputs(spec, B, seed=1)
    c.garble_inputs(X)
    for _ in range(2):
        c.evaluate_device()
    stream = torch.cuda.ExternalStream(api.stream_handle())
    steps = 5
    ms = 0.0
    for _ in range(steps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        c.evaluate_device()
        e1.record(stream)
        e1.synchronize()
        ms += e0.elapsed_time(e1)
    ms /= steps
    got = c.move_outputs()
    vals, ok = c.decode_outputs()
    exact = None
    oracle_s = None
    if check:
        from oracle import oraclebind as O

        g = gcm.read_gcm(path)
        oc = O.OracleCircuit(path)
        t0 = time.time()
        out = oc.evaluate(gcm.encode_inputs(g, X[0]), threads=len(os.sched_getaffinity(0)))
        oracle_s = time.time() - t0
        exact = all(np.array_equal(a[0], b) for a, b in zip(got, out))
        v, missing = oc.decode(out)
        exact = bool(exact and missing == 0 and np.array_equal(v, vals[0]))
    relu = sum(1 for l in plain if l["type"] == "relu")
    print(json.dumps({
        "model": name, "k": len(spec["crt"]), "layers": len(plain), "relu_layers": relu,
        "garble_kernel_ms": st["kernel_ms"], "garble_total_s": st["total_ms"] / 1e3, "garble_wall_s": wall,
        "gates": st["gates"], "ciphertext_rows": st["rows"], "table_GB": st["table_bytes"] / 1e9,
        "rows_per_s_in_kernels": st["rows"] / (st["kernel_ms"] / 1e3) if st["kernel_ms"] else None,
        "reference_cpu_garble_s": ref_s, "load_s": load_s, "garble_to_ready_circuit_s": direct_s,
        "eval_batch": B, "eval_ms_per_step": ms, "eval_inferences_per_s": B / (ms / 1e3),
        "all_outputs_decode": bool(ok.all()), "oracle_check_bit_exact": exact, "oracle_eval_s": oracle_s,
    }), flush=True)
    c.close()


if __name__ == "__main__":
    main()
