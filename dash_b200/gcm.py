"""Reader of DGCM material files (include/dash_gcm_format.h) and the pieces
of host logic that only need the file: the plaintext circuit inside it and
the garbler-side input encoding."""
import math
import struct

import numpy as np


def nr_comps(q):
    return int(math.floor(128 / math.log2(q)))


def read_gcm(path):
    data = open(path, "rb").read()
    if data[:8] != b"DASHGCM1":
        raise ValueError("not a DGCM file: " + str(path))
    pos = 8
    g = {"layers": [], "dec": {}, "in_base": {}, "in_offset": {}, "zero": None, "upshift": None}
    while pos < len(data):
        tag, aux, n = struct.unpack_from("<4sIQ", data, pos)
        pos += 16
        payload = data[pos:pos + n]
        pos += (n + 15) & ~15
        if tag == b"HEAD":
            h = np.frombuffer(payload, dtype=np.int64)
            g.update(k=int(h[0]), n_mrs=int(h[1]), max_modulus=int(h[2]), input_size=int(h[3]), n_layers=int(h[4]))
            g["crt"] = [int(x) for x in h[8:8 + g["k"]]]
            g["mrs"] = [int(x) for x in h[40:40 + g["n_mrs"]]]
        elif tag == b"ZERO":
            g["zero"] = np.frombuffer(payload, dtype=np.int16)
        elif tag == b"UPSH":
            g["upshift"] = np.frombuffer(payload, dtype=np.int16)
        elif tag == b"LAYR":
            g["layers"].append({"p": np.frombuffer(payload, dtype=np.int64), "arrays": []})
        elif tag == b"ARRY":
            g["layers"][-1]["arrays"].append(payload)
        elif tag == b"DECI":
            g["dec"][aux] = payload
        elif tag == b"GINB":
            g["in_base"][aux] = np.frombuffer(payload, dtype=np.int16)
        elif tag == b"GINO":
            g["in_offset"][aux] = np.frombuffer(payload, dtype=np.int16)
        elif tag == b"ENDF":
            break
    return g


def zero_label(g, q):
    off = sum(nr_comps(m) for m in range(2, q))
    return g["zero"][off:off + nr_comps(q)]


def _crt_combine(residues, moduli):
    """Smallest non-negative x with x = residues[j] mod moduli[j]."""
    P = 1
    for m in moduli:
        P *= m
    x = 0
    for r, m in zip(residues, moduli):
        Pi = P // m
        x += int(r) * Pi * pow(Pi, -1, m)
    return x % P


def plain_circuit(g):
    """The plaintext circuit inside a DGCM file, as the layer dicts
    dash_b200.garble.garble_to_file takes.  Biases are recovered from the bias
    labels: label = b * R + Z with R[0] = 1, so b = label[0] - Z[0] (mod p)."""
    crt, k = g["crt"], g["k"]
    out = []
    for L in g["layers"]:
        p, a = L["p"], L["arrays"]
        t = int(p[0])
        if t in (1, 2):
            M = int(p[1]) if t == 1 else int(p[4])
            res = []
            for j in range(k):
                nc = nr_comps(crt[j])
                lab = np.frombuffer(a[1 + j], dtype=np.int16).reshape(M, nc)
                res.append((lab[:, 0].astype(np.int64) - int(zero_label(g, crt[j])[0])) % crt[j])
            biases = np.array([_crt_combine([res[j][r] for j in range(k)], crt) for r in range(M)], dtype=np.int64)
            w = np.frombuffer(a[0], dtype=np.int64)
            if t == 1:
                out.append({"type": "dense", "M": M, "K": int(p[2]), "channel_tf": int(p[3]), "weights": w, "biases": biases})
            else:
                out.append({"type": "conv2d", "geom": tuple(int(x) for x in p[1:9]), "weights": w, "biases": biases})
        elif t == 3:
            out.append({"type": "relu"})
        elif t == 4:
            out.append({"type": "sign"})
        elif t == 5:
            out.append({"type": "rescale", "factors": tuple(int(x) for x in p[3:3 + int(p[2])])})
        elif t == 6:
            out.append({"type": "rescale_legacy", "l": int(p[2])})
        else:
            raise ValueError("layer type %d has no plaintext form here" % t)
    return out


def encode_inputs(g, x):
    """garble_inputs (garbled_circuit_interface.h:325-343) from the file's
    GINB/GINO sections: per CRT modulus int16 [input_size][nc]."""
    x = np.asarray(x, dtype=np.int64).ravel()
    labels = []
    for j, q in enumerate(g["crt"]):
        nc = nr_comps(q)
        base = g["in_base"][j].reshape(-1, nc).astype(np.int64)
        off = g["in_offset"][j].astype(np.int64)
        labels.append(((base + (x % q)[:, None] * off[None, :]) % q).astype(np.int16))
    return labels


def compress_labels(q, labels):
    """LabelTensor::compress (garbling/label_tensor.h:715-727) for a whole
    array: int16 digits [..., nc] -> uint64 [..., 2] (low, high half of the
    128-bit value sum_i d_i q^i), vectorised on 32-bit limbs."""
    labels = np.asarray(labels)
    nc = labels.shape[-1]
    lab = labels.reshape(-1, nc).astype(np.uint64)
    x = [np.zeros(lab.shape[0], dtype=np.uint64) for _ in range(4)]
    qq = np.uint64(q)
    mask, s32 = np.uint64(0xFFFFFFFF), np.uint64(32)
    for d in range(nc - 1, -1, -1):
        carry = lab[:, d]
        for i in range(4):
            t = x[i] * qq + carry
            x[i] = t & mask
            carry = t >> s32
    out = np.stack([x[0] | (x[1] << s32), x[2] | (x[3] << s32)], axis=-1)
    return np.ascontiguousarray(out.reshape(labels.shape[:-1] + (2,)))


def decompress_labels(q, values):
    """LabelTensor::decompress: uint64 [..., 2] -> int16 digits [..., nc(q)]
    (exact base-q digits; valid labels are below q^nc)."""
    values = np.asarray(values, dtype=np.uint64)
    nc = nr_comps(q)
    v = values.reshape(-1, 2)
    mask, s32 = np.uint64(0xFFFFFFFF), np.uint64(32)
    x = [v[:, 0] & mask, v[:, 0] >> s32, v[:, 1] & mask, v[:, 1] >> s32]
    qq = np.uint64(q)
    out = np.zeros((v.shape[0], nc), dtype=np.int16)
    for d in range(nc):
        r = np.zeros(v.shape[0], dtype=np.uint64)
        for i in range(3, -1, -1):
            cur = (r << s32) | x[i]
            x[i] = cur // qq
            r = cur % qq
        out[:, d] = r.astype(np.int16)
    return out.reshape(values.shape[:-1] + (nc,))
