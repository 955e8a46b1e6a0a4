"""The GPU garbler (include/dash_b200_garble.h): material
garbled on the B200 must evaluate -- on the pinned CPU oracle and on the GPU
evaluator -- to exactly the values the reference-garbled circuit decodes to
for the same plaintext inputs, and must have the reference's structure."""
import numpy as np
import pytest

from helpers import assert_labels_equal

pytestmark = pytest.mark.gpu

CASES = [("MLP_TINY", 9), ("MNIST_TINY_CNN", 7), ("CIFAR_TINY_SCALE", 6), ("SIGN_TINY", 11), ("RESCALE_2_3", 5),
         ("RESCALE_OPT", 4), ("RELU_99", 8), ("RESCALE_2", 3), ("RESCALE_LEGACY", 10)]


@pytest.mark.parametrize("name,batch", CASES)
def test_gpu_garbled_circuit_evaluates_like_the_reference(gpu, tmp_path, name, batch):
    import workloads as W
    from dash_b200 import garble, gcm
    from oracle import oraclebind as O
    from oracle import refbind as R

    if not R.available():
        pytest.skip("oracle/_ref not present on this box")
    spec = W.MODELS[name]
    ref = W.build_reference_circuit(spec, seed=5, weight_seed=5)
    ref_path = str(tmp_path / "ref.gcm")
    ref.export(ref_path)
    gref = gcm.read_gcm(ref_path)
    plain = gcm.plain_circuit(gref)

    gpu_path = str(tmp_path / "gpu.gcm")
    stats = garble.garble_to_file(gref["crt"], gref["mrs"], gref["input_size"], plain, gpu_path, seed=b"seed-%s" % name.encode(),
                                  max_modulus=gref["max_modulus"])
    assert stats["gates"] > 0 or all(l["type"] in ("dense", "conv2d") for l in plain)
    ggpu = gcm.read_gcm(gpu_path)
    # same structure as the reference's material: layer types, parameters, size of every table
    assert len(ggpu["layers"]) == len(gref["layers"])
    for a, b in zip(ggpu["layers"], gref["layers"]):
        assert list(a["p"][:12]) == list(b["p"][:12])
        assert [len(x) for x in a["arrays"]] == [len(x) for x in b["arrays"]]
    assert {j: len(v) for j, v in ggpu["dec"].items()} == {j: len(v) for j, v in gref["dec"].items()}

    X = W.synthetic_inputs(spec, batch, seed=9)
    want = []
    for x in X:
        lab = ref.garble_inputs(x)
        out, _ = ref.cpu_evaluate(lab, 2)
        want.append(ref.decode_outputs(out))
    want = np.array(want)

    c = gpu.Circuit.load(gpu_path, max_batch=batch)
    c.garble_inputs(np.asarray(X, dtype=np.int64).reshape(batch, -1))
    c.evaluate_device()
    got_labels = c.move_outputs()
    vals, ok = c.decode_outputs()
    assert ok.all(), "outputs of the GPU-garbled circuit do not decode"
    assert np.array_equal(vals, want.reshape(vals.shape))

    oc = O.OracleCircuit(gpu_path)
    for b in range(batch):
        lab = gcm.encode_inputs(ggpu, X[b])
        out = oc.evaluate(lab, threads=2)
        assert_labels_equal([g[b] for g in got_labels], out, "%s: oracle vs GPU evaluator on GPU-garbled material, inference %d" % (name, b))
        v, missing = oc.decode(out)
        assert missing == 0 and np.array_equal(v, want[b].ravel())
    c.close()


def test_same_seed_same_circuit_other_seed_other_labels(gpu, tmp_path):
    import workloads as W
    from dash_b200 import garble, gcm
    from oracle import refbind as R

    if not R.available():
        pytest.skip("oracle/_ref not present on this box")
    spec = W.MODELS["MLP_TINY"]
    ref = W.build_reference_circuit(spec, seed=5, weight_seed=5)
    ref.export(str(tmp_path / "ref.gcm"))
    gref = gcm.read_gcm(str(tmp_path / "ref.gcm"))
    plain = gcm.plain_circuit(gref)
    paths = []
    for i, seed in enumerate((b"one", b"one", b"two")):
        p = str(tmp_path / ("g%d.gcm" % i))
        garble.garble_to_file(gref["crt"], gref["mrs"], gref["input_size"], plain, p, seed=seed, max_modulus=gref["max_modulus"])
        paths.append(open(p, "rb").read())
    assert paths[0] == paths[1]
    assert paths[0] != paths[2] and len(paths[0]) == len(paths[2])


def test_garble_straight_into_an_evaluator_circuit(gpu, tmp_path):
    """dashb_garble_to_circuit: same seed as the file path -> the same labels,
    without the tables ever leaving the device."""
    import workloads as W
    from dash_b200 import garble

    spec = W.MODELS["CIFAR_TINY_SCALE"]
    plain, n_in = W.plain_layers(spec, weight_seed=3)
    path = str(tmp_path / "g.gcm")
    garble.garble_to_file(spec["crt"], spec["mrs"], n_in, plain, path, seed=b"abc")
    X = W.synthetic_inputs(spec, 12, seed=2)
    a = gpu.Circuit.load(path, max_batch=12)
    a.garble_inputs(X)
    a.evaluate_device()
    want = a.move_outputs()
    want_vals, ok = a.decode_outputs()
    assert ok.all()
    b, stats = garble.garble_to_circuit(spec["crt"], spec["mrs"], n_in, plain, 12, seed=b"abc")
    b.garble_inputs(X)
    b.evaluate_device()
    assert_labels_equal(b.move_outputs(), want, "file path vs in-memory path")
    vals, ok = b.decode_outputs()
    assert ok.all() and np.array_equal(vals, want_vals)
    assert stats["rows"] > 0
    a.close()
    b.close()


def _plain_eval(layers, x, P):
    """Integer semantics of the plaintext circuit (circuit/layer/*.h
    plain_eval): values are residues mod P read as signed numbers."""
    def signed(v):
        v = np.mod(v, P)
        return np.where(v >= P // 2, v - P, v)

    v = np.asarray(x, dtype=object)
    for l in layers:
        if l["type"] == "dense":
            w = np.asarray(l["weights"], dtype=object).reshape(l["M"], l["K"])
            v = signed(w.dot(v) + np.asarray(l["biases"], dtype=object))
        elif l["type"] == "conv2d":
            in_w, in_h, ch, filt, fw, fh, sw, sh = l["geom"]
            ow, oh = (in_w - fw) // sw + 1, (in_h - fh) // sh + 1
            w = np.asarray(l["weights"], dtype=object)
            out = np.zeros(filt * ow * oh, dtype=object)
            fs = fw * fh
            for f in range(filt):
                for y in range(oh):
                    for xx in range(ow):
                        acc = int(l["biases"][f])
                        for c in range(ch):
                            for i in range(fh):
                                for j in range(fw):
                                    # weight index as the reference reads it (label_tensor.h:967-969)
                                    acc += int(w[f * fs * ch + c * fs + i * fh + j]) * int(v[c * in_w * in_h + (y * sh + i) * in_w + xx * sw + j])
                        out[f * ow * oh + y * ow + xx] = acc
            v = signed(out)
        elif l["type"] == "relu":
            v = np.where(v > 0, v, 0)
        else:
            raise ValueError(l["type"])
    return np.array([int(t) for t in v], dtype=np.int64)


@pytest.mark.parametrize("name", ["MLP_TINY", "MNIST_TINY_CNN"])
def test_gpu_garbled_circuit_decodes_to_the_plain_evaluation(gpu, name):
    """Independent of the reference: with an exact sign base (accuracy 100 %)
    decode(evaluate(garble(C), encode(x))) must equal C(x) computed with plain
    integers, for circuits whose values stay inside (-P/2, P/2)."""
    import workloads as W
    from dash_b200 import garble

    spec = W.MODELS[name]
    plain, n_in = W.plain_layers(spec, weight_seed=11, wmax=2)
    P = 1
    for p in spec["crt"]:
        P *= p
    batch = 6
    X = W.synthetic_inputs(spec, batch, seed=4) % 8  # small inputs keep every intermediate value far from P/2
    c, _ = garble.garble_to_circuit(spec["crt"], spec["mrs"], n_in, plain, batch, seed=b"plain")
    c.garble_inputs(X)
    c.evaluate_device()
    vals, ok = c.decode_outputs()
    assert ok.all()
    for b in range(batch):
        assert np.array_equal(vals[b], _plain_eval(plain, X[b], P)), "inference %d" % b
    c.close()


def test_onnx_model_to_gpu_garbled_inference(gpu, tmp_path):
    """The whole pipeline without protobuf and without the reference's
    garbler: ONNX file -> integration/dash_onnx_lite.h (reference layer
    classes quantise) -> plaintext circuit -> GPU garbler -> GPU evaluator;
    decoded values equal the reference's own garble + evaluate + decode of the
    same model."""
    import sys

    from helpers import ROOT
    sys.path.insert(0, ROOT + "/tests")
    from onnx_writer import cnn

    from dash_b200 import garble, gcm
    from oracle import refbind as R

    if not R.available():
        pytest.skip("oracle/_ref not present on this box")
    layers = [("conv", 4, 3, 1), ("relu",), ("conv", 4, 2, 2), ("relu",), ("flatten",), ("dense", 8), ("relu",), ("dense", 4)]
    blob, _ = cnn("pytorch", 10, 10, 1, layers, seed=12)
    path = str(tmp_path / "m.onnx")
    open(path, "wb").write(blob)
    ref = R.RefCircuit()
    ref.load_onnx(path, 2, 0.02, 32)  # ScaleQuantPlus, s = 32: a Rescale after every linear layer
    crt, mrs = [32, 3, 5, 7, 11, 13, 17], [10, 9, 9, 8, 7, 7, 6]
    ref.garble(crt, mrs, 32, 3)
    ref.export(str(tmp_path / "ref.gcm"))
    g = gcm.read_gcm(str(tmp_path / "ref.gcm"))
    plain = gcm.plain_circuit(g)
    assert [l["type"] for l in plain].count("rescale") == 4
    batch = 5
    X = np.random.default_rng(1).integers(0, 256, size=(batch, 100), dtype=np.int64)
    want = np.array([ref.decode_outputs(ref.cpu_evaluate(ref.garble_inputs(x), 2)[0]) for x in X])
    c, _ = garble.garble_to_circuit(crt, mrs, 100, plain, batch, seed=b"onnx", max_modulus=g["max_modulus"])
    c.garble_inputs(X)
    c.evaluate_device()
    vals, ok = c.decode_outputs()
    assert ok.all() and np.array_equal(vals, want.reshape(vals.shape))
    c.close()


def test_garbling_is_deterministic_under_repetition(gpu, tmp_path):
    """Regression: uploads and table clears on the NULL stream used to race
    the gate kernel on the library's non-blocking stream (about one run in
    twenty produced different, wrong material).  Thirty garblings of the same
    circuit with the same seed must give one file."""
    import hashlib

    import workloads as W
    from dash_b200 import garble

    spec = dict(input=(12, 12, 2),
                layers=[("conv", 16, 3, 1), ("relu",), ("conv", 32, 2, 2), ("relu",), ("flatten",), ("dense", 64), ("relu",),
                        ("dense", 10)],
                quant=("scaleplus", 32), crt=W.SCALE_CRT, mrs=W.SCALE_MRS, max_modulus=32)
    plain, n_in = W.plain_layers(spec, weight_seed=1)
    digests = set()
    for r in range(30):
        path = str(tmp_path / ("g%d.gcm" % (r % 2)))
        garble.garble_to_file(spec["crt"], spec["mrs"], n_in, plain, path, seed=b"smoke")
        digests.add(hashlib.sha256(open(path, "rb").read()).hexdigest())
    assert len(digests) == 1
