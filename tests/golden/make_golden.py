"""Generates the golden fixtures in this directory with the UNMODIFIED
reference (oracle/_ref/libdash_ref.so, built from /root/reference by
oracle/Makefile).  Run in the build container only:

    python tests/golden/make_golden.py

For every unit circuit: <name>.gcm (garbled material, seeded garbling) and
<name>.npz with integer inputs `x`, the reference's input labels `in<j>`,
the output labels of every garbled layer `l<layer>_<j>` as computed by the
reference's own cpu_evaluate, and the decoded outputs `dec`.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))

import workloads as W  # noqa: E402
from oracle import refbind as R  # noqa: E402

GOLDEN = ["SIGN_TINY", "RELU_99", "RESCALE_LEGACY", "RESCALE_2", "RESCALE_2_3", "RESCALE_OPT", "PROJECTION",
          "MLP_TINY", "CONV_UNIT", "RELU_EDGE", "PROJ_UP_BACK", "PROJ_DOWN_BACK", "PROJ_2X2",
          "MAXPOOL_2X2", "MAXPOOL_P9", "MAXPOOL_1X1", "MAXPOOL_4X4C2_DENSE", "BASE_EXT_1", "BASE_EXT_2"]

# extra unit circuits only used as fixtures
W.MODELS["CONV_UNIT"] = dict(  # stride 2, two channels, three filters (cf. test_conv2d.h:133-261)
    input=(5, 5, 2), layers=[("conv", 3, 3, 2), ("relu",), ("flatten",), ("dense", 2)],
    quant=("simple", 0.05), crt=[2, 3, 5, 7, 11], mrs=[54, 4, 3], max_modulus=0)
W.MODELS["RELU_EDGE"] = dict(  # test_relu.h:144-225: crt {2,3,5}, mrs {26,6,3,2}
    input=(14, 1, 1), layers=[("flatten",), ("relu",)], quant=("simple", 1.0), crt=[2, 3, 5], mrs=[26, 6, 3, 2],
    max_modulus=0)

# dash/test/test_projection.h:18-188, the three circuits with the test's own functions
W.MODELS["PROJ_UP_BACK"] = dict(   # :18-50  19 -x*2-> 91 -x/2-10-> 19, max_modulus 100; 15 -> 5
    input=(1, 1, 1), layers=[("flatten",), ("projection", [19], [91], 2), ("projection", [91], [19], 4)],
    quant=("simple", 1.0), crt=[19], mrs=[], max_modulus=100)
W.MODELS["PROJ_DOWN_BACK"] = dict(  # :52-83  19 -x/4-> 10 -x*2-> 19; 12 -> 6
    input=(1, 1, 1), layers=[("flatten",), ("projection", [19], [10], 5), ("projection", [10], [19], 2)],
    quant=("simple", 1.0), crt=[19], mrs=[], max_modulus=0)
W.MODELS["PROJ_2X2"] = dict(        # :85-120 one projection x*2 over a 2x2 tensor; {3,2,1,-1} -> {6,4,2,-2}
    input=(2, 2, 1), layers=[("flatten",), ("projection", [19], [19], 2)],
    quant=("simple", 1.0), crt=[19], mrs=[], max_modulus=100)

# dash/test/test_maxpool2d.h: a 2x2 window over a 2x2x1 input (stride 1), with the test's bases and values
W.MODELS["MAXPOOL_2X2"] = dict(    # :93-170 (crt {2,3,5}, mrs {26,6,3,2}); mixed_high {3,-11,5,10} -> 10
    input=(2, 2, 1), layers=[("maxpool", 2, 1)], quant=("simple", 1.0), crt=[2, 3, 5], mrs=[26, 6, 3, 2], max_modulus=0)
W.MODELS["MAXPOOL_P9"] = dict(     # :173-212 (nine primes, mrs {76,7,7,7,7,7,5,5}); mid_p9 / neg_high_p9 / low_p9
    input=(2, 2, 1), layers=[("maxpool", 2, 1)], quant=("simple", 1.0), **W.cpm(9))
W.MODELS["MAXPOOL_1X1"] = dict(    # :74-84 TestGarbledMaxPool2DMultKernel: 1x1 kernel, outputs = inputs
    input=(2, 2, 1), layers=[("maxpool", 1, 1)], quant=("simple", 1.0), crt=[2, 3, 5], mrs=[26, 6, 3, 2], max_modulus=0)
W.MODELS["MAXPOOL_4X4C2_DENSE"] = dict(  # two channels, 2x2 / stride 2 windows, then a dense layer on the pooled labels
    input=(4, 4, 2), layers=[("maxpool", 2, 2), ("flatten",), ("dense", 3)], quant=("simple", 0.1), crt=[2, 3, 5, 7],
    mrs=[26, 3], max_modulus=0)

# the standalone base-extension layer (layer/garbled_base_extension.h; no test in the reference): one and two
# extra moduli; the inputs are zero because the layer starts the extra residues from the circuit's input BASE
# labels (:74-80), i.e. from the labels of the value 0
W.MODELS["BASE_EXT_1"] = dict(input=(5, 1, 1), layers=[("flatten",), ("base_extension", (7,))], quant=("simple", 1.0),
                              crt=[2, 3, 5, 7], mrs=[], max_modulus=0)
W.MODELS["BASE_EXT_2"] = dict(input=(4, 1, 1), layers=[("flatten",), ("base_extension", (3, 11))], quant=("simple", 1.0),
                              crt=[2, 3, 5, 7, 11, 13], mrs=[], max_modulus=0)

BATCH = 3


def inputs_for(name, spec, n):
    if name == "RELU_EDGE":
        P = 30  # 0, +-1, +-mid, +-extremes of the signed range (-P/2, P/2)
        row = np.array([0, 1, -1, 7, -7, 14, -14, 2, -2, 13, -13, 5, -5, 3], dtype=np.int64)
        return np.stack([row, -row, np.roll(row, 3)])[:, :n]
    if name.startswith("BASE_EXT"):  # row 0: zeros (decodes); rows 1, 2: other labels on the non-extra residues
        return np.array([[0] * n, list(range(1, n + 1)), [-(i + 1) for i in range(n)]], dtype=np.int64)
    if name in ("MAXPOOL_2X2", "MAXPOOL_1X1"):   # first row: mixed_high of the reference test
        return np.array([[3, -11, 5, 10], [0, 0, 0, 0], [-14, 1, 0, 7]], dtype=np.int64)
    if name == "MAXPOOL_P9":     # mid_p9, neg_high_p9, low_p9
        return np.array([[55773217] * 4, [-111546435] * 4, [1, 1, 1, 1]], dtype=np.int64)
    if name == "MAXPOOL_4X4C2_DENSE":
        return np.random.default_rng(5).integers(-2, 3, size=(BATCH, n), dtype=np.int64)
    if name == "PROJ_UP_BACK":   # first row = the reference test's input
        return np.array([[15], [4], [9]], dtype=np.int64)
    if name == "PROJ_DOWN_BACK":
        return np.array([[12], [8], [17]], dtype=np.int64)
    if name == "PROJ_2X2":
        return np.array([[3, 2, 1, -1], [0, 4, -4, 9], [-9, 7, -2, 5]], dtype=np.int64)
    if name in ("RESCALE_LEGACY", "RESCALE_2", "RESCALE_2_3"):
        rng = np.random.default_rng(3)
        return rng.integers(-1000, 1000, size=(BATCH, n), dtype=np.int64)
    return W.synthetic_inputs(spec, BATCH, seed=11)


def main():
    regenerate_all = "--all" in sys.argv
    for name in GOLDEN:
        if not regenerate_all and os.path.exists(os.path.join(HERE, name + ".npz")):
            continue  # existing fixtures stay byte-identical; pass --all to regenerate everything
        spec = W.MODELS[name]
        c = W.build_reference_circuit(spec, seed=1234, weight_seed=7)
        c.export(os.path.join(HERE, name + ".gcm"))
        X = inputs_for(name, spec, c.input_size)
        out = {"x": X}
        ins = [[] for _ in range(c.k)]
        lay = {}
        dec = []
        for x in X:
            lab = c.garble_inputs(x)
            for j in range(c.k):
                ins[j].append(lab[j])
            for l in range(c.n_layers):
                o = c.cpu_evaluate_layer(lab, l, 1)
                for j in range(c.k):
                    lay.setdefault((l, j), []).append(o[j])
            final, _ = c.cpu_evaluate(lab, 1)
            if all(q == p for q, p in zip(c.out_moduli, c.crt)):
                dec.append(c.decode_outputs(final))
        for j in range(c.k):
            out["in%d" % j] = np.stack(ins[j])
        for (l, j), v in lay.items():
            out["l%d_%d" % (l, j)] = np.stack(v)
        if dec:
            out["dec"] = np.stack(dec)
        np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
        print(name, "layers", c.n_layers, "gcm KB", os.path.getsize(os.path.join(HERE, name + ".gcm")) >> 10)


if __name__ == "__main__":
    main()
