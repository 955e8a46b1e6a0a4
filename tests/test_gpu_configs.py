"""BASELINE.json's configurations at their full sizes and the linear-layer
variant the round-1 suite did not reach, through the C ABI on the B200:

  * dense layers with the tf2onnx channel-last input permutation
    (LabelTensor::matvecmul_tf, garbling/label_tensor.h:820-858; kernel
    MatMulModAddZeroTF, misc/cuda_util.h:116-138) behind a convolution;
  * configs[0] MODEL_A and configs[1] MODEL_B_POOL_REPL at batch 256 against
    the reference's own evaluator (oracle/_ref);
  * configs[2] MODEL_F_GNNP_POOL_REPL and configs[3] MODEL_F_MINIONN_POOL_REPL
    at full size on GPU-garbled material against the pinned C oracle.
"""
import os

import numpy as np
import pytest

from helpers import assert_labels_equal

pytestmark = pytest.mark.gpu


def _ref_or_skip():
    from oracle import refbind as R

    if not R.available():
        pytest.skip("oracle/_ref not present on this box")
    return R


def _encode(ref, X):
    per = [ref.garble_inputs(x) for x in X]
    return [np.stack([p[j] for p in per]) for j in range(ref.k)]


# ------------------------------------------------------------- (a) TF dense
@pytest.mark.parametrize("batch", [1, 17, 40])
@pytest.mark.parametrize("geom", [
    # (width, height, channels, filters, filter, stride, dense rows)
    (6, 6, 1, 4, 3, 1, 12),     # 4x4x4 = 64 taps, channel_tf = 4
    (7, 5, 2, 3, 2, 1, 20),     # odd channel count: 6x4x3 = 72 taps, channel_tf = 3
    (9, 9, 1, 16, 2, 1, 130),   # 1024 taps x 130 rows: would take the tcgen05 path without the permutation
])
def test_dense_channel_tf_behind_convolution(gpu, tmp_path, geom, batch):
    import workloads as W

    _ref_or_skip()
    w, h, ch, filt, f, stride, rows = geom
    spec = dict(input=(w, h, ch), layers=[("conv", filt, f, stride), ("relu",), ("flatten",), ("dense_tf", rows), ("relu",)],
                quant=("simple", 0.02), **W.cpm(8, 99.0))
    ref = W.build_reference_circuit(spec, seed=w * 100 + rows, weight_seed=filt)
    path = str(tmp_path / "tf.gcm")
    ref.export(path)
    from dash_b200 import gcm

    dense = [l for l in gcm.read_gcm(path)["layers"] if int(l["p"][0]) == 1]
    assert len(dense) == 1 and int(dense[0]["p"][3]) == filt, "the exported dense layer must carry channel_tf"
    labels = _encode(ref, W.synthetic_inputs(spec, batch, seed=8))
    c = gpu.Circuit.load(path, max_batch=batch)
    got = c.evaluate(labels)
    for b in sorted({0, batch // 2, batch - 1}):
        want, _ = ref.cpu_evaluate([l[b] for l in labels], 4)
        assert_labels_equal([g[b] for g in got], want, "TF dense %s inference %d" % (geom, b))
    c.close()


def test_tf2onnx_model_end_to_end(gpu, tmp_path):
    """A tf2onnx-flavoured ONNX file (weights stored last-to-first, dense
    weights for channel-last activations) -> integration/dash_onnx_lite.h ->
    reference layer objects and garbler -> GPU evaluator: output labels and
    decoded values equal the reference's own evaluation."""
    import sys

    from helpers import ROOT

    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from onnx_writer import cnn

    R = _ref_or_skip()
    layers = [("conv", 3, 3, 1), ("relu",), ("conv", 4, 2, 2), ("relu",), ("flatten",), ("dense", 6), ("relu",), ("dense", 5)]
    blob, _ = cnn("tf2onnx", 8, 8, 2, layers, seed=21)
    onnx_path = str(tmp_path / "tf.onnx")
    open(onnx_path, "wb").write(blob)
    ref = R.RefCircuit()
    ref.load_onnx(onnx_path, 0, 0.02, -1)
    ref.garble([2, 3, 5, 7, 11, 13, 17, 19], [98, 9, 8, 8, 7, 5], 0, 5)
    path = str(tmp_path / "tf.gcm")
    ref.export(path)
    from dash_b200 import gcm

    assert any(int(l["p"][0]) == 1 and int(l["p"][3]) > 0 for l in gcm.read_gcm(path)["layers"]), "no TF-permuted dense layer in the model"
    batch = 23
    X = np.random.default_rng(6).integers(0, 256, size=(batch, ref.input_size), dtype=np.int64)
    labels = _encode(ref, X)
    c = gpu.Circuit.load(path, max_batch=batch)
    got = c.evaluate(labels)
    vals, ok = c.decode_outputs()
    assert ok.all()
    for b in range(batch):
        want, _ = ref.cpu_evaluate([l[b] for l in labels], 2)
        assert_labels_equal([g[b] for g in got], want, "tf2onnx model inference %d" % b)
        assert np.array_equal(vals[b], ref.decode_outputs(want))
    c.close()


# --------------------------------------- (b), (c) configs[0] / configs[1]
def _full_batch_vs_reference(gpu, model, label_sample):
    import bench

    _ref_or_skip()
    B = 256
    gcm_path, labels, ref = bench.build_material(model, B, want_ref=True)
    c = gpu.Circuit.load(gcm_path, max_batch=B)
    got = c.evaluate(labels)
    vals, ok = c.decode_outputs()
    assert ok.all()
    threads = len(os.sched_getaffinity(0))
    for b in range(B):
        want, _ = ref.cpu_evaluate([l[b] for l in labels], threads)
        if b in label_sample:
            assert_labels_equal([g[b] for g in got], want, "%s inference %d" % (model, b))
        assert np.array_equal(vals[b], ref.decode_outputs(want)), "%s: decoded values of inference %d" % (model, b)
    c.close()


def test_full_size_model_a_batch_256(gpu):
    """BASELINE.json configs[0] (MNIST MLP 784-128-128-10, k = 8): all 256
    decoded outputs equal the reference's, 16 sampled inferences label-exact."""
    _full_batch_vs_reference(gpu, "MODEL_A", set(range(0, 256, 17)) | {255})


def test_full_size_model_b_all_256_decoded(gpu):
    """BASELINE.json configs[1]: all 256 decoded outputs equal the
    reference's, 40 sampled inferences label-exact (every lane boundary
    included)."""
    sample = set(range(0, 256, 8)) | {63, 64, 127, 128, 191, 192, 255, 1}
    _full_batch_vs_reference(gpu, "MODEL_B_POOL_REPL", sample)


# ---------------------------------------------- (d) configs[2] / configs[3]
def _rescale_plain(v, s, P):
    """What the garbled rescaling computes on plain integers: upshift by P/2
    (garbled_circuit_interface.h:948-954), exact division of the value minus
    its residue mod s (rescale_gadget.h:271-307), downshift by P/(2s), both
    shifts in integer arithmetic (:969-980).  For the reference's prime bases
    (P/2 odd, s = 2) this is ScalarTensor::rescale's ceil(v / s)
    (circuit/scalar_tensor.h:477-493); for the base {32,3,...,17} P/(2*32) is
    not an integer and the result is v/32 rounded to nearest."""
    return (v + P // 2) // s - P // (2 * s)


@pytest.mark.parametrize("model,batch", [("MODEL_F_GNNP_POOL_REPL", 16), ("MODEL_F_MINIONN_POOL_REPL", 16)])
def test_full_size_cifar_models_vs_oracle(gpu, tmp_path, model, batch):
    """configs[2] / configs[3] at full size (78 k / 125 k ReLU elements and as
    many rescale elements per inference), material garbled on the GPU:
    two inferences of the batch label-exact against oracle/dash_oracle.c (the
    pinned restatement), every inference decodes, and decoding agrees with
    the oracle's decode."""
    import workloads as W
    from dash_b200 import garble, gcm
    from oracle import oraclebind as O

    spec = W.MODELS[model]
    plain, n_in = W.plain_layers(spec)
    path = str(tmp_path / "cifar.gcm")
    garble.garble_to_file(spec["crt"], spec["mrs"], n_in, plain, path, seed=b"cifar-full-size")
    g = gcm.read_gcm(path)
    X = W.synthetic_inputs(spec, batch, seed=31)
    c = gpu.Circuit.load(path, max_batch=batch)
    c.garble_inputs(X)
    c.evaluate_device()
    got = c.move_outputs()
    vals, ok = c.decode_outputs()
    assert ok.all()
    oc = O.OracleCircuit(path)
    threads = len(os.sched_getaffinity(0))
    for b in (0, batch - 1):
        want = oc.evaluate(gcm.encode_inputs(g, X[b]), threads=threads)
        assert_labels_equal([o[b] for o in got], want, "%s inference %d vs oracle" % (model, b))
        v, missing = oc.decode(want)
        assert missing == 0 and np.array_equal(v, vals[b].ravel())
    c.close()
    oc.close()
    os.remove(path)


@pytest.mark.parametrize("model", ["MODEL_F_GNNP_POOL_REPL", "MODEL_F_MINIONN_POOL_REPL"])
def test_full_size_first_scaled_layer_decodes_to_plain_integers(gpu, model):
    """The first convolution + Rescale{32} of configs[2] / configs[3] at full
    size as a circuit of its own (no sign approximation involved): every
    decoded output equals the plain integer semantics, convolution
    (Conv2d::plain_q_eval) then _rescale_plain, for the whole batch."""
    import workloads as W
    from dash_b200 import garble

    spec = W.MODELS[model]
    plain, n_in = W.plain_layers(spec)
    plain = plain[:2]
    assert [l["type"] for l in plain] == ["conv2d", "rescale"]
    batch = 8
    X = W.synthetic_inputs(spec, batch, seed=5)
    c, _ = garble.garble_to_circuit(spec["crt"], spec["mrs"], n_in, plain, batch, seed=b"first-layer")
    c.garble_inputs(X)
    c.evaluate_device()
    vals, ok = c.decode_outputs()
    assert ok.all()
    in_w, in_h, ch, filt, fw, fh, sw, sh = plain[0]["geom"]
    ow, oh = (in_w - fw) // sw + 1, (in_h - fh) // sh + 1
    w = np.asarray(plain[0]["weights"], dtype=np.int64)
    fs = fw * fh
    # weight of tap (filter v, channel c, row i, column j) as the reference reads it (label_tensor.h:967-969)
    wt = np.zeros((filt, ch, fh, fw), dtype=np.int64)
    for ci in range(ch):
        for i in range(fh):
            for j in range(fw):
                wt[:, ci, i, j] = w[np.arange(filt) * fs * ch + ci * fs + i * fh + j]
    x = X.reshape(batch, ch, in_h, in_w)
    out = np.zeros((batch, filt, oh, ow), dtype=np.int64)
    for i in range(fh):
        for j in range(fw):
            patch = x[:, :, i:i + sh * (oh - 1) + 1:sh, j:j + sw * (ow - 1) + 1:sw]
            out += np.einsum("bchw,fc->bfhw", patch, wt[:, :, i, j])
    out += np.asarray(plain[0]["biases"], dtype=np.int64)[None, :, None, None]
    P = int(np.prod(np.asarray(spec["crt"], dtype=np.int64)))
    want = _rescale_plain(out, int(plain[1]["factors"][0]), P).reshape(batch, -1)
    assert np.array_equal(vals, want)
    c.close()


def test_large_convolution_takes_several_launches(gpu, tmp_path):
    """A convolution whose (pixel group, row group) count passes the 65535
    limit of grid.y on the dp4a kernel (363 x 363 outputs, 15 filters:
    65 886 pairs) is split over several launches; every output label equals
    the reference's."""
    import workloads as W

    _ref_or_skip()
    spec = dict(input=(364, 364, 1), layers=[("conv", 15, 2, 1)], quant=("simple", 0.05), crt=[7, 11], mrs=[], max_modulus=0)
    ref = W.build_reference_circuit(spec, seed=2, weight_seed=2)
    path = str(tmp_path / "bigconv.gcm")
    ref.export(path)
    X = W.synthetic_inputs(spec, 1, seed=4)
    labels = _encode(ref, X)
    c = gpu.Circuit.load(path, max_batch=1)
    got = c.evaluate(labels)
    want, _ = ref.cpu_evaluate([l[0] for l in labels], len(os.sched_getaffinity(0)))
    assert_labels_equal([g[0] for g in got], want, "363x363x15 convolution")
    c.close()


def test_garbler_writes_secrets_to_a_separate_file(gpu, tmp_path):
    """dashb_garble_to_files: the evaluator's file carries no input encoding
    and no decoding information; the garbler's file carries exactly those and
    no layer; together they evaluate like the single-file form of the same
    seed."""
    import workloads as W
    from dash_b200 import api, garble, gcm

    spec = W.MODELS["MNIST_TINY_CNN"]
    plain, n_in = W.plain_layers(spec, weight_seed=3)
    one = str(tmp_path / "one.gcm")
    ev, sec = str(tmp_path / "evaluator.gcm"), str(tmp_path / "garbler.gcm")
    garble.garble_to_file(spec["crt"], spec["mrs"], n_in, plain, one, seed=b"split")
    garble.garble_to_file(spec["crt"], spec["mrs"], n_in, plain, ev, seed=b"split", secrets_path=sec)
    g_ev, g_sec, g_one = gcm.read_gcm(ev), gcm.read_gcm(sec), gcm.read_gcm(one)
    assert not g_ev["dec"] and not g_ev["in_base"] and not g_ev["in_offset"]
    assert len(g_ev["layers"]) == len(g_one["layers"]) and not g_sec["layers"]
    assert set(g_sec["dec"]) == set(g_sec["in_base"]) == set(g_sec["in_offset"]) == set(range(len(spec["crt"])))
    batch = 5
    X = W.synthetic_inputs(spec, batch, seed=1)
    a = gpu.Circuit.load(one, max_batch=batch)
    a.garble_inputs(X)
    a.evaluate_device()
    want = a.move_outputs()
    b = gpu.Circuit.load(ev, max_batch=batch)
    with pytest.raises(api.DashB200Error):
        b.garble_inputs(X)  # the evaluator's circuit cannot encode inputs ...
    labels = [np.stack([gcm.encode_inputs(g_sec, x)[j] for x in X]) for j in range(len(spec["crt"]))]
    got = b.evaluate(labels)  # ... it evaluates the labels the garbler sends
    assert_labels_equal(got, want, "split files vs single file")
    with pytest.raises(api.DashB200Error):
        b.decode_outputs()  # ... and cannot decode
    a.close()
    b.close()
    # a garbling without a seed draws a fresh one: two runs, two different circuits
    r1, r2 = str(tmp_path / "r1.gcm"), str(tmp_path / "r2.gcm")
    garble.garble_to_file(spec["crt"], spec["mrs"], n_in, plain, r1)
    garble.garble_to_file(spec["crt"], spec["mrs"], n_in, plain, r2)
    assert open(r1, "rb").read() != open(r2, "rb").read()
