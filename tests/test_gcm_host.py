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
