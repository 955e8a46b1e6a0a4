#!/usr/bin/env python
"""Config sweep: per-model and per-layer
online-evaluation throughput on one B200, next to the reference's CPU
evaluator on the box's host cores.  Prints one JSON line per case.

    python tools/sweep.py models          # C1..C3-class models
    python tools/sweep.py micro           # C5: dense / conv / relu / sign / rescale
    python tools/sweep.py gnnp            # C3 at full size (minutes of garbling)

Material comes from the reference's garbler (oracle/_ref, test
infrastructure), outside every timed region.
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
from dash_b200 import api  # noqa: E402
from oracle import refbind  # noqa: E402

PEAK = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(
    os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0


def run_case(name, spec, batches, steps=10, cpu_reps=3, seed=42, check=2):
    t0 = time.time()
    ref = W.build_reference_circuit(spec, seed=seed, weight_seed=seed)
    garble_s = time.time() - t0
    path = "/tmp/sweep_%s.gcm" % name
    ref.export(path)
    maxb = max(batches)
    X = W.synthetic_inputs(spec, min(maxb, 64), seed=seed)
    per = [ref.garble_inputs(x) for x in X]
    base = [np.stack([p[j] for p in per]) for j in range(ref.k)]
    threads = len(os.sched_getaffinity(0))
    # reference CPU evaluator: one inference at a time, all host threads
    ref.cpu_evaluate([l[0] for l in base], threads)
    cpu_ms = []
    for r in range(cpu_reps):
        _, ms = ref.cpu_evaluate([l[r % len(X)] for l in base], threads)
        cpu_ms.append(ms)
    cpu_ms = float(np.median(cpu_ms))
    stream = torch.cuda.ExternalStream(api.stream_handle())
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for B in batches:
        reps = -(-B // len(X))
        labels = [np.concatenate([l] * reps)[:B].copy() for l in base]
        c = api.Circuit.load(path, max_batch=B)
        got = c.evaluate(labels)
        ok = True
        for b in range(min(check, B)):
            want, _ = ref.cpu_evaluate([l[b] for l in labels], threads)
            ok = ok and all(np.array_equal(g[b], w) for g, w in zip(got, want))
        c.move_inputs(labels)
        for _ in range(3):
            c.evaluate_device()
        ms = 0.0
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
        c.set_layer_timing(True)
        c.evaluate_device()
        by = {}
        for n, v in c.layer_times_ms():
            by[n] = by.get(n, 0.0) + v
        c.set_layer_timing(False)
        alg = c.algorithmic_bytes(-1) * B
        line = {
            "case": name, "batch": B, "k": ref.k, "ms_per_step": ms, "inferences_per_s": B / (ms / 1e3),
            "algorithmic_GBps": alg / (ms / 1e3) / 1e9, "frac_of_hbm_peak": alg / (ms / 1e3) / 1e9 / PEAK,
            "aes_blocks_per_s": c.aes_blocks(-1) * B / (ms / 1e3), "reference_aes_per_inference": c.reference_aes_blocks(-1),
            "layer_ms_serialised": by, "bit_exact_vs_reference": bool(ok),
            "cpu_reference_ms_per_inference": cpu_ms, "cpu_threads": threads, "cpu_inferences_per_s": 1e3 / cpu_ms,
            "garble_s": garble_s,
        }
        print(json.dumps(line), flush=True)
        c.close()
    ref.close()


def relu_spec(n, k, acc, kind="relu"):
    return dict(input=(n, 1, 1), layers=[("flatten",), (kind,)], quant=("simple", 1.0), **W.cpm(k, acc))


def main():
    what = sys.argv[1] if len(sys.argv) > 1 else "models"
    api.init(0)
    if what == "models":
        for name, bs in (("MODEL_A", [1, 16, 256]), ("MODEL_B_POOL_REPL", [1, 16, 256]), ("MODEL_C", [256]),
                         ("MODEL_D_POOL_REPL", [256]), ("MODEL_E_30", [256]), ("CIFAR_TINY_SCALE", [256])):
            run_case(name, W.MODELS[name], bs)
    elif what == "gnnp":
        run_case("MODEL_F_GNNP_POOL_REPL", W.MODELS["MODEL_F_GNNP_POOL_REPL"], [16], steps=3, cpu_reps=1, check=1)
    elif what == "dense1":  # one case, for a profiler
        run_case("dense_2048", dict(input=(2048, 1, 1), layers=[("flatten",), ("dense", 2048)],
                                    quant=("simple", 0.02), crt=W.PRIMES[:8], mrs=[], max_modulus=0), [256], steps=2, cpu_reps=1, check=1)
    elif what == "dense":
        for dim in (128, 512, 2048):
            run_case("dense_%d" % dim, dict(input=(dim, 1, 1), layers=[("flatten",), ("dense", dim)],
                                            quant=("simple", 0.02), crt=W.PRIMES[:8], mrs=[], max_modulus=0), [1, 16, 256])
        run_case("MODEL_A", W.MODELS["MODEL_A"], [256])
    elif what == "conv":
        # the convolution shapes of MODEL_F_GNNP / MINIONN (32x32 CIFAR models), one layer each
        for name, (w, h, ch, f, fs, st) in (("conv_30x30x32_32f3", (30, 30, 32, 32, 3, 1)),
                                            ("conv_28x28x32_32f2s2", (28, 28, 32, 32, 2, 2)),
                                            ("conv_14x14x32_64f3", (14, 14, 32, 64, 3, 1)),
                                            ("conv_12x12x64_64f3", (12, 12, 64, 64, 3, 1)),
                                            ("conv_5x5x64_128f3", (5, 5, 64, 128, 3, 1))):
            run_case(name, dict(input=(w, h, ch), layers=[("conv", f, fs, st)], quant=("simple", 0.02),
                                crt=[32, 3, 5, 7, 11, 13, 17], mrs=[], max_modulus=32), [16, 64])
        run_case("MODEL_D_POOL_REPL", W.MODELS["MODEL_D_POOL_REPL"], [256])
    elif what == "rescale1":  # one case, for a profiler
        run_case("rescale32_n16384", dict(input=(16384, 1, 1), layers=[("flatten",), ("rescale", (32,))], quant=("simple", 1.0),
                                          crt=W.SCALE_CRT, mrs=W.SCALE_MRS, max_modulus=32), [64], steps=2, cpu_reps=1, check=1)
    elif what == "rescale":
        # C3/C4 scaling gadget: factor 32 on the {32,3,5,7,11,13,17} base, and the k=8 prime base with factor 2
        for n in (1024, 16384):
            run_case("rescale32_n%d" % n, dict(input=(n, 1, 1), layers=[("flatten",), ("rescale", (32,))], quant=("simple", 1.0),
                                               crt=W.SCALE_CRT, mrs=W.SCALE_MRS, max_modulus=32), [16, 256] if n == 1024 else [64])
        run_case("rescale2_n4096", dict(input=(4096, 1, 1), layers=[("flatten",), ("rescale", (2,))],
                                        quant=("simple", 1.0), **W.cpm(8)), [256])
    elif what == "micro":
        # benchmarks/micro_benchmarks/non_sgx/main.cpp:577-683 grids (subset that garbles in seconds)
        for n in (128, 1024, 4096):
            for acc in (99.0, 100.0):
                run_case("relu_n%d_acc%g" % (n, acc), relu_spec(n, 8, acc), [1, 16, 256])
        for n in (128, 1024):
            run_case("sign_n%d" % n, relu_spec(n, 8, 100.0, "sign"), [1, 256])
            run_case("rescale2_n%d" % n, dict(input=(n, 1, 1), layers=[("flatten",), ("rescale", (2,))],
                                              quant=("simple", 1.0), **W.cpm(8)), [1, 256])
            run_case("rescale_legacy_n%d" % n, dict(input=(n, 1, 1), layers=[("flatten",), ("rescale_legacy", 1)],
                                                    quant=("simple", 1.0), **W.cpm(8)), [1, 256])
        for dim in (128, 512, 2048):
            run_case("dense_%d" % dim, dict(input=(dim, 1, 1), layers=[("flatten",), ("dense", dim)],
                                            quant=("simple", 0.02), crt=W.PRIMES[:8], mrs=[], max_modulus=0), [1, 16, 256])
        run_case("conv_64x64x3_16f4s2", dict(input=(64, 64, 3), layers=[("conv", 16, 4, 2)], quant=("simple", 0.02),
                                             crt=W.PRIMES[:8], mrs=[], max_modulus=0), [1, 16, 256])


if __name__ == "__main__":
    main()
