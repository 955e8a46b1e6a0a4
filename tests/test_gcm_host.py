"""Host logic around the DGCM file (dash_b200/gcm.py), CPU tier: the input
encoding read back from a reference-garbled file reproduces the reference's
own garble_inputs, and the plaintext circuit recovered from it has the weights
the reference garbled."""
import numpy as np
import pytest


def _ref(name, tmp_path):
    import workloads as W
    from oracle import refbind as R

    if not R.available():
        pytest.skip("oracle/_ref not built")
    spec = W.MODELS[name]
    ref = W.build_reference_circuit(spec, seed=21, weight_seed=21)
    path = str(tmp_path / "m.gcm")
    ref.export(path)
    return spec, ref, path


@pytest.mark.parametrize("name", ["MLP_TINY", "CIFAR_TINY_SCALE"])
def test_encode_inputs_equals_reference_garble_inputs(tmp_path, name):
    import workloads as W
    from dash_b200 import gcm

    spec, ref, path = _ref(name, tmp_path)
    g = gcm.read_gcm(path)
    for x in W.synthetic_inputs(spec, 3, seed=8):
        want = ref.garble_inputs(x)
        got = gcm.encode_inputs(g, x)
        assert len(got) == ref.k
        for a, b in zip(got, want):
            assert np.array_equal(a, np.asarray(b).reshape(a.shape))


def test_plain_circuit_recovers_layers_and_bias_residues(tmp_path):
    from dash_b200 import gcm

    spec, ref, path = _ref("MNIST_TINY_CNN", tmp_path)
    g = gcm.read_gcm(path)
    plain = gcm.plain_circuit(g)
    assert [l["type"] for l in plain] == ["conv2d", "relu", "conv2d", "relu", "dense", "relu", "dense"]
    P = 1
    for p in g["crt"]:
        P *= p
    for l, L in zip(plain, g["layers"]):
        if l["type"] not in ("dense", "conv2d"):
            continue
        assert ((0 <= l["biases"]) & (l["biases"] < P)).all()
        # the recovered bias reproduces every modulus' bias label digit 0: label = b*R + Z, R[0] = 1
        for j, q in enumerate(g["crt"]):
            nc = gcm.nr_comps(q)
            lab = np.frombuffer(L["arrays"][1 + j], dtype=np.int16).reshape(-1, nc)
            z0 = int(gcm.zero_label(g, q)[0])
            assert np.array_equal((l["biases"] % q + z0) % q, lab[:, 0])


def test_malformed_material_files_are_rejected(tmp_path):
    """A truncated, inconsistent or hostile DGCM file must be refused by the
    loader before any pointer reaches dashb_add_* (dashb_material_check is the
    host half of dashb_circuit_load)."""
    import ctypes as C
    import struct

    from dash_b200 import api
    from helpers import GOLDEN_DIR

    L = api.lib()
    L.dashb_material_check.argtypes = [C.c_char_p]
    import os

    for name in ("MLP_TINY", "RESCALE_2_3", "RESCALE_LEGACY", "PROJECTION", "SIGN_TINY", "CONV_UNIT"):
        assert L.dashb_material_check(os.path.join(GOLDEN_DIR, name + ".gcm").encode()) == 0, name
    good = open(os.path.join(GOLDEN_DIR, "MLP_TINY.gcm"), "rb").read()

    def check(blob):
        p = str(tmp_path / "bad.gcm")
        open(p, "wb").write(blob)
        return L.dashb_material_check(p.encode())

    assert check(good) == 0
    assert check(good[: len(good) // 2]) == -5           # truncated
    assert check(good[:8] + good[8:8 + 16 + 72 * 8] * 2 + good[8 + 16 + 72 * 8:]) == -5   # HEAD twice
    # records: tag(4) aux(4) n(8) payload padded to 16
    recs, pos = [], 8
    while pos < len(good):
        tag, aux, n = struct.unpack_from("<4sIQ", good, pos)
        size = 16 + ((n + 15) & ~15)
        recs.append((tag, pos, n, size))
        pos += size
    arr = [r for r in recs if r[0] == b"ARRY"]
    tag, pos, n, size = arr[1]                            # first bias array of the first dense layer
    # length near 2^64: padded size wraps to 0
    assert check(good[:pos + 8] + struct.pack("<Q", 2**64 - 3) + good[pos + 16:]) == -5
    # an array 16 bytes shorter than its layer implies
    short = good[:pos + 8] + struct.pack("<Q", n - 16) + good[pos + 16:pos + size - 16] + good[pos + size:]
    assert check(short) == -1
    # the same array twice
    assert check(good[:pos + size] + good[pos:pos + size] + good[pos + size:]) == -5
    # a dense layer whose K does not match the labels arriving
    lay = [r for r in recs if r[0] == b"LAYR"][0]
    p2 = bytearray(good)
    struct.pack_into("<q", p2, lay[1] + 16 + 2 * 8, 17)  # p[2] = K
    assert check(bytes(p2)) == -1


def test_wrong_digit_count_is_rejected_before_the_library_reads():
    """Circuit._in_arrays checks the last dimension against nc(q): a wrong nc
    would make the library read past the host buffer."""
    from dash_b200 import api

    class Fake(api.Circuit):
        def __init__(self):
            self.k, self.input_size, self.in_moduli = 2, 3, [3, 5]

    f = Fake()
    f.h = None
    ok = [np.zeros((4, 3, api.nr_comps(3)), dtype=np.int16), np.zeros((4, 3, api.nr_comps(5)), dtype=np.int16)]
    assert f._in_arrays(ok)[1] == 4
    bad = [ok[0], np.zeros((4, 3, api.nr_comps(5) - 1), dtype=np.int16)]
    with pytest.raises(ValueError):
        f._in_arrays(bad)
