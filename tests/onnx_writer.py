"""A minimal ONNX writer for the tests (protobuf wire format by hand; there is
no onnx package in this environment): just the messages and fields
circuit/onnx_modelloader.h reads, in the two flavours it distinguishes."""
import struct

import numpy as np


def _varint(v):
    out = bytearray()
    v &= (1 << 64) - 1
    while True:
        b = v & 0x7F
        v >>= 7
        if v:
            out.append(b | 0x80)
        else:
            out.append(b)
            return bytes(out)


def _key(field, wire):
    return _varint((field << 3) | wire)


def _ld(field, payload):
    return _key(field, 2) + _varint(len(payload)) + payload


def _int(field, v):
    return _key(field, 0) + _varint(v)


def _str(field, s):
    return _ld(field, s.encode())


def tensor(name, array, raw=True, packed_dims=False):
    a = np.ascontiguousarray(array, dtype="<f4")
    if packed_dims:
        out = _ld(1, b"".join(_varint(d) for d in a.shape))
    else:
        out = b"".join(_int(1, d) for d in a.shape)
    out += _int(2, 1)  # FLOAT
    out += _str(8, name)
    if raw:
        out += _ld(9, a.tobytes())
    else:
        out += _ld(4, a.tobytes())  # packed float_data
    return out


def attribute(name, ints):
    return _str(1, name) + b"".join(_int(8, i) for i in ints) + _int(20, 7)  # type INTS


def node(op_type, inputs, outputs, attributes=()):
    out = b"".join(_str(1, i) for i in inputs) + b"".join(_str(2, o) for o in outputs)
    out += _str(4, op_type)
    out += b"".join(_ld(5, a) for a in attributes)
    return out


def value_info(name, dims):
    shape = b"".join(_ld(1, _int(1, d)) for d in dims)      # TensorShapeProto.dim.dim_value
    ttype = _int(1, 1) + _ld(2, shape)                      # Tensor: elem_type, shape
    return _str(1, name) + _ld(2, _ld(1, ttype))            # ValueInfoProto: name, type.tensor_type


def model(producer, input_dims, nodes, initializers):
    graph = b"".join(_ld(1, n) for n in nodes) + _str(2, "g") + b"".join(_ld(5, t) for t in initializers)
    graph += _ld(11, value_info("input", input_dims)) + _ld(12, value_info("output", [1, 1]))
    return _int(1, 8) + _str(2, producer) + _str(3, "test") + _ld(7, graph) + _ld(8, _int(2, 13))


def cnn(flavour, in_w, in_h, channels, layers, seed=0, raw=True):
    """layers: ("conv", filters, k, stride) | ("relu",) | ("tanh",) | ("flatten",) | ("dense", out).
    Returns (model bytes, [(kind, float weights, float biases), ...]) with the
    weights exactly as the file stores them."""
    rng = np.random.default_rng(seed)
    nodes, inits, params = [], [], []
    w, h, c = in_w, in_h, channels
    flat = None
    idx = 0
    for l in layers:
        kind = l[0]
        if kind == "conv":
            _, filt, k, stride = l
            wt = rng.uniform(-0.5, 0.5, size=(filt, c, k, k)).astype(np.float32)
            bs = rng.uniform(-0.5, 0.5, size=(filt,)).astype(np.float32)
            if flavour == "pytorch":
                attrs = [attribute("dilations", [1, 1]), attribute("group", [1]), attribute("kernel_shape", [k, k]),
                         attribute("pads", [0, 0, 0, 0]), attribute("strides", [stride, stride])]
            else:
                attrs = [attribute("dilations", [1, 1]), attribute("strides", [stride, stride]), attribute("kernel_shape", [k, k])]
            nodes.append(node("Conv", ["x%d" % idx, "w%d" % idx, "b%d" % idx], ["x%d" % (idx + 1)], attrs))
            inits.append((tensor("w%d" % idx, wt, raw), tensor("b%d" % idx, bs, raw)))
            params.append(("conv", wt, bs, (w, h, c, filt, k, stride)))
            w, h, c = (w - k) // stride + 1, (h - k) // stride + 1, filt
        elif kind == "dense":
            out = l[1]
            wt = rng.uniform(-0.5, 0.5, size=(out, flat)).astype(np.float32)
            bs = rng.uniform(-0.5, 0.5, size=(out,)).astype(np.float32)
            stored = wt if flavour == "pytorch" else np.ascontiguousarray(wt.T)  # tf2onnx: [in][out]
            nodes.append(node("Gemm", ["x%d" % idx, "w%d" % idx, "b%d" % idx], ["x%d" % (idx + 1)],
                              [attribute("transB", [1])]))
            inits.append((tensor("w%d" % idx, stored, raw, packed_dims=True), tensor("b%d" % idx, bs, raw)))
            params.append(("dense", wt, bs, (out, flat)))
            flat = out
        elif kind == "relu":
            nodes.append(node("Relu", ["x%d" % idx], ["x%d" % (idx + 1)]))
            params.append(("relu",))
        elif kind == "tanh":
            nodes.append(node("Tanh", ["x%d" % idx], ["x%d" % (idx + 1)]))
            params.append(("sign",))
        elif kind == "flatten":
            if flavour == "pytorch":
                nodes.append(node("Flatten", ["x%d" % idx], ["x%d" % (idx + 1)], [attribute("axis", [1])]))
            else:
                nodes.append(node("Reshape", ["x%d" % idx, "shape"], ["sequential/flatten/Reshape:0"]))
            flat = w * h * c
            params.append(("flatten", (w, h, c)))
        idx += 1
    if flavour == "pytorch":
        initializers = [t for pair in inits for t in pair]
        dims = [1, channels, in_h, in_w]
    else:  # read from the back: bias, then weights, first layer last
        initializers = [t for pair in reversed(inits) for t in pair]
        dims = [1, in_h, in_w, channels]
    return model("pytorch" if flavour == "pytorch" else "tf2onnx", dims, nodes, initializers), params
