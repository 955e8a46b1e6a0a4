"""Workload definitions: the circuits of BASELINE.json's configs and of the
reference's test suite, as layer specs.

A spec is turned into a plaintext circuit + garbled material by the
*reference's own* classes through oracle/refbind.py (test infrastructure);
the B200 evaluator only ever sees the exported DGCM material.  Nothing here
computes on labels.

Architectures: models/train_pytorch.ipynb cell 8 of the reference;
bases: benchmarks/model_benchmarks/non_sgx/main.cpp:36-86,227-335 and
GarbledCircuitInterface::get_mrs_base (garbling/garbled_circuit_interface.h:
888-930); weight init: benchmarks/micro_benchmarks/non_sgx/main.cpp:20-55.
"""
import numpy as np

PRIMES = [2, 3, 5, 7, 11, 13, 17, 19, 23, 29, 31]

# (k, accuracy) -> mrs base; subset of garbled_circuit_interface.h:888-930
MRS_BASE = {
    (4, 100.0): [26, 3],
    (5, 100.0): [54, 4, 3],
    (6, 100.0): [60, 5, 5, 5],
    (7, 100.0): [86, 7, 6, 6, 5],
    (8, 100.0): [98, 9, 8, 8, 7, 5],
    (8, 99.999): [102, 7, 6, 5, 5],
    (8, 99.99): [78, 7, 5, 4],
    (8, 99.9): [78, 5, 3],
    (8, 99.0): [126],
    (9, 100.0): [76, 7, 7, 7, 7, 7, 5, 5],
    (9, 99.0): [138],
}

SCALE_CRT = [32, 3, 5, 7, 11, 13, 17]      # model_benchmarks/non_sgx/main.cpp:62-68
SCALE_MRS = [10, 9, 9, 8, 7, 7, 6]


def cpm(k, acc=100.0):
    return dict(crt=PRIMES[:k], mrs=MRS_BASE[(k, acc)], max_modulus=0)


def _mlp(dims):
    layers = [("flatten",)]
    for i, d in enumerate(dims):
        layers.append(("dense", d))
        if i + 1 < len(dims):
            layers.append(("relu",))
    return layers


MODELS = {
    # config C1 (BASELINE.json configs[0])
    "MODEL_A": dict(input=(28, 28, 1), layers=_mlp([128, 128, 10]), quant=("simple", 0.02), **cpm(8)),
    # config C2 (configs[1]): MNIST LeNet-5-style CNN, pooling replaced by strided conv
    "MODEL_B_POOL_REPL": dict(
        input=(28, 28, 1),
        layers=[("conv", 5, 5, 1), ("relu",), ("conv", 5, 3, 3), ("relu",), ("conv", 10, 3, 1), ("relu",),
                ("conv", 10, 3, 3), ("flatten",), ("dense", 100), ("relu",), ("dense", 10)],
        quant=("simple", 0.02), **cpm(9)),
    "MODEL_C": dict(
        input=(28, 28, 1),
        layers=[("conv", 5, 4, 2), ("relu",), ("flatten",), ("dense", 100), ("relu",), ("dense", 10)],
        quant=("simple", 0.02), **cpm(9)),
    "MODEL_D_POOL_REPL": dict(
        input=(28, 28, 1),
        layers=[("conv", 16, 6, 2), ("relu",), ("conv", 16, 6, 2), ("relu",), ("flatten",), ("dense", 100), ("relu",),
                ("dense", 10)],
        quant=("simple", 0.02), **cpm(8)),
    "MODEL_E_30": dict(input=(28, 28, 1), layers=[("flatten",), ("dense", 30), ("sign",), ("dense", 10)],
                       quant=("simple", 0.02), **cpm(5)),
    # config C3 (configs[2]): CIFAR-10 small CNN, ScaleQuantPlus s=32, rescale after every linear layer
    "MODEL_F_GNNP_POOL_REPL": dict(
        input=(32, 32, 3),
        layers=[("conv", 32, 3, 1), ("relu",), ("conv", 32, 3, 1), ("relu",), ("conv", 32, 2, 2),
                ("conv", 64, 3, 1), ("relu",), ("conv", 64, 3, 1), ("relu",), ("conv", 64, 2, 2),
                ("conv", 128, 3, 1), ("relu",), ("conv", 128, 3, 1), ("relu",), ("flatten",), ("dense", 10)],
        quant=("scaleplus", 32), crt=SCALE_CRT, mrs=SCALE_MRS, max_modulus=32),
    # config C4 (configs[3])
    "MODEL_F_MINIONN_POOL_REPL": dict(
        input=(32, 32, 3),
        layers=[("conv", 64, 3, 1), ("relu",), ("conv", 64, 3, 1), ("relu",), ("conv", 64, 2, 2),
                ("conv", 64, 3, 1), ("relu",), ("conv", 64, 3, 1), ("relu",), ("conv", 64, 2, 2),
                ("conv", 64, 3, 1), ("relu",), ("conv", 64, 1, 1), ("relu",), ("conv", 16, 1, 1), ("relu",),
                ("flatten",), ("dense", 10)],
        quant=("scaleplus", 32), crt=SCALE_CRT, mrs=SCALE_MRS, max_modulus=32),
    # a reduced CIFAR CNN with the same structure (scaling gadget after each
    # linear layer) that the reference garbles in seconds; parity-test case
    "CIFAR_TINY_SCALE": dict(
        input=(8, 8, 3),
        layers=[("conv", 4, 3, 1), ("relu",), ("conv", 4, 2, 2), ("relu",), ("flatten",), ("dense", 10)],
        quant=("scaleplus", 32), crt=SCALE_CRT, mrs=SCALE_MRS, max_modulus=32),
    "MNIST_TINY_CNN": dict(
        input=(10, 10, 1),
        layers=[("conv", 3, 3, 1), ("relu",), ("conv", 4, 3, 3), ("relu",), ("flatten",), ("dense", 12), ("relu",),
                ("dense", 10)],
        quant=("simple", 0.05), **cpm(9)),
    # unit circuits modelled on the reference's tests (dash/test/test_*.h)
    "SIGN_TINY": dict(input=(6, 1, 1), layers=[("flatten",), ("dense", 5), ("sign",), ("dense", 3)],
                      quant=("simple", 0.05), **cpm(5)),
    "RELU_99": dict(input=(6, 1, 1), layers=[("flatten",), ("relu",)], quant=("simple", 1.0), **cpm(8, 99.0)),
    "RESCALE_LEGACY": dict(input=(5, 1, 1), layers=[("flatten",), ("rescale_legacy", 2)], quant=("simple", 1.0),
                           **cpm(9)),                                   # test_rescale.h:28-57
    "RESCALE_2": dict(input=(5, 1, 1), layers=[("flatten",), ("rescale", (2,))], quant=("simple", 1.0), **cpm(8)),
    "RESCALE_2_3": dict(input=(5, 1, 1), layers=[("flatten",), ("rescale", (2, 3))], quant=("simple", 1.0),
                        **cpm(6)),
    "RESCALE_OPT": dict(input=(4, 1, 1), layers=[("flatten",), ("dense", 3), ("rescale", (32,)), ("relu",)],
                        quant=("simple", 0.05), crt=[32, 97, 107], mrs=[22, 19, 15, 13], max_modulus=107),
    # a modulus large enough for k_rescale's 32-bit accumulators ((k+1)(q-1)^2 >= 2^16)
    "RESCALE_WIDE": dict(input=(6, 1, 1), layers=[("flatten",), ("rescale", (2,))], quant=("simple", 1.0),
                         crt=[2, 3, 251], mrs=[], max_modulus=255),
    "PROJECTION": dict(input=(4, 1, 1), layers=[("flatten",), ("projection", [19], [91], 0),
                                                ("projection", [91], [19], 0)],
                       quant=("simple", 1.0), crt=[19], mrs=[], max_modulus=100),   # test_projection.h:18-50
    "MLP_TINY": dict(input=(4, 4, 1), layers=_mlp([8, 8, 4]), quant=("simple", 0.05), **cpm(8)),
}


def _uniform(rng, bound, shape):
    return rng.uniform(-bound, bound, size=shape).astype(np.float32)


def build_reference_circuit(spec, seed=42, weight_seed=42):
    """Instantiates `spec` with the reference's classes and garbles it.

    Returns the oracle.refbind.RefCircuit (test infrastructure)."""
    from oracle import refbind

    rng = np.random.default_rng(weight_seed)
    qname, qval = spec["quant"]
    q_method = {"simple": 0, "scale": 1, "scaleplus": 2}[qname]
    q_const = float(qval) if qname == "simple" else 10.0
    q_param = -1 if qname == "simple" else int(qval)
    c = refbind.RefCircuit()
    w, h, ch = spec["input"]
    flat = None
    for layer in spec["layers"]:
        kind = layer[0]
        dims = [flat] if flat is not None else [w, h, ch]
        if kind == "flatten":
            c.add_flatten([w, h, ch])
            flat = w * h * ch
        elif kind in ("dense", "dense_tf"):
            out = layer[1]
            assert flat is not None, "dense needs a flatten first"
            bound = np.sqrt(1.0 / flat)
            # dense_tf: the tf2onnx channel-last input permutation (Dense::get_channel_tf, label_tensor.h:820-858),
            # channel count = filters of the convolution in front of the flatten
            c.add_dense(_uniform(rng, bound, (out, flat)), _uniform(rng, bound, (out,)), q_method, q_const, q_param,
                        ch if kind == "dense_tf" else 0)
            flat = out
            if qname == "scale":
                c.add_rescale_legacy(q_param, [flat])
            elif qname == "scaleplus":
                c.add_rescale([q_param], [flat])
        elif kind == "conv":
            _, filt, ksz, stride = layer
            bound = np.sqrt(1.0 / (ch * ksz * ksz))
            c.add_conv2d(_uniform(rng, bound, (filt, ch, ksz, ksz)), _uniform(rng, bound, (filt,)), w, h, stride,
                         stride, q_method, q_const, q_param)
            w, h, ch = (w - ksz) // stride + 1, (h - ksz) // stride + 1, filt
            if qname == "scale":
                c.add_rescale_legacy(q_param, [w, h, ch])
            elif qname == "scaleplus":
                c.add_rescale([q_param], [w, h, ch])
        elif kind == "base_extension":
            c.add_base_extension(list(layer[1]), dims)
        elif kind == "maxpool":
            _, ksz, stride = layer
            c.add_maxpool2d(w, h, ch, ksz, ksz, stride, stride)
            w, h = (w - ksz) // stride + 1, (h - ksz) // stride + 1
        elif kind == "relu":
            c.add_relu(dims)
        elif kind == "sign":
            c.add_sign(dims)
        elif kind == "rescale_legacy":
            c.add_rescale_legacy(layer[1], dims)
        elif kind == "rescale":
            c.add_rescale(list(layer[1]), dims)
        elif kind == "projection":
            c.add_projection(flat, layer[1], layer[2], layer[3] if len(layer) > 3 else 0)
        else:
            raise ValueError(kind)
    c.garble(spec["crt"], spec.get("mrs", []), spec.get("max_modulus", 0), seed)
    return c


def plain_layers(spec, weight_seed=42, wmax=7):
    """The circuit of `spec` as plaintext layer dicts for the GPU garbler
    (dash_b200.garble.garble_to_file), with synthetic quantised weights
    U[-wmax, wmax]: same layer structure build_reference_circuit gives the
    reference (a rescale after every linear layer for the scaling quantisers),
    no reference code involved.  Returns (layers, input_size)."""
    rng = np.random.default_rng(weight_seed)
    qname, qval = spec["quant"]
    w, h, ch = spec["input"]
    flat = None
    out = []

    def after_linear():
        if qname == "scaleplus":
            out.append({"type": "rescale", "factors": (int(qval),)})
        elif qname == "scale":
            raise ValueError("legacy rescale is not supported by the GPU garbler")

    for layer in spec["layers"]:
        kind = layer[0]
        if kind == "flatten":
            flat = w * h * ch
        elif kind == "dense":
            m = layer[1]
            out.append({"type": "dense", "M": m, "K": flat, "channel_tf": 0,
                        "weights": rng.integers(-wmax, wmax + 1, size=(m, flat), dtype=np.int64),
                        "biases": rng.integers(-wmax, wmax + 1, size=(m,), dtype=np.int64)})
            flat = m
            after_linear()
        elif kind == "conv":
            _, filt, ksz, stride = layer
            out.append({"type": "conv2d", "geom": (w, h, ch, filt, ksz, ksz, stride, stride),
                        "weights": rng.integers(-wmax, wmax + 1, size=(filt * ch * ksz * ksz,), dtype=np.int64),
                        "biases": rng.integers(-wmax, wmax + 1, size=(filt,), dtype=np.int64)})
            w, h, ch = (w - ksz) // stride + 1, (h - ksz) // stride + 1, filt
            after_linear()
        elif kind in ("relu", "sign"):
            out.append({"type": kind})
        elif kind == "rescale":
            out.append({"type": "rescale", "factors": tuple(layer[1])})
        else:
            raise ValueError("layer %r is not supported by the GPU garbler" % (kind,))
    return out, spec["input"][0] * spec["input"][1] * spec["input"][2]


def synthetic_inputs(spec, batch, seed=42):
    """Integer inputs U[0,255] (micro_benchmarks/non_sgx/main.cpp:57-65) of
    the circuit's input shape, one row per inference."""
    rng = np.random.default_rng(seed)
    w, h, ch = spec["input"]
    return rng.integers(0, 256, size=(batch, w * h * ch), dtype=np.int64)


def relu_elements(spec):
    """(#ReLU, #sign, #rescale) elements of one inference; for reporting."""
    w, h, ch = spec["input"]
    flat = None
    relu = sign = resc = 0
    scale = spec["quant"][0] in ("scale", "scaleplus")
    for layer in spec["layers"]:
        kind = layer[0]
        n = flat if flat is not None else w * h * ch
        if kind == "flatten":
            flat = w * h * ch
        elif kind == "dense":
            flat = layer[1]
            resc += flat if scale else 0
        elif kind == "conv":
            _, filt, ksz, stride = layer
            w, h, ch = (w - ksz) // stride + 1, (h - ksz) // stride + 1, filt
            resc += w * h * ch if scale else 0
        elif kind == "relu":
            relu += n
        elif kind == "sign":
            sign += n
        elif kind in ("rescale", "rescale_legacy"):
            resc += n
    return relu, sign, resc
