"""integration/dash_onnx_lite.h: an ONNX file read without
protobuf gives the same Dash circuit as handing the same float weights to the
reference's layer constructors directly -- checked on the garbled material
(same garbling seed -> byte-identical DGCM export), for a pytorch-flavoured
and a tf2onnx-flavoured model and all three quantisers."""
import numpy as np
import pytest

from onnx_writer import cnn

LAYERS = [("conv", 3, 3, 1), ("relu",), ("conv", 4, 2, 2), ("relu",), ("flatten",), ("dense", 6), ("tanh",), ("dense", 5)]
CRT = [2, 3, 5, 7, 11, 13, 17, 19]
MRS = [98, 9, 8, 8, 7, 5]


def _programmatic(R, params, flavour, q_method, q_const, q_param, channel_first_dense):
    c = R.RefCircuit()
    first_dense = True
    conv_filters = 0
    for p in params:
        if p[0] == "conv":
            _, w, b, (in_w, in_h, ch, filt, k, stride) = p
            c.add_conv2d(w, b, in_w, in_h, stride, stride, q_method, q_const, q_param)
            conv_filters = filt
            dims = [(in_w - k) // stride + 1, (in_h - k) // stride + 1, filt]
            if q_method == 1:
                c.add_rescale_legacy(q_param, dims)
            elif q_method == 2:
                c.add_rescale([q_param], dims)
            cur = dims
        elif p[0] == "dense":
            _, w, b, (out, flat) = p
            tf = conv_filters if (flavour == "tf2onnx" and first_dense and channel_first_dense) else 0
            first_dense = False
            c.add_dense(w, b, q_method, q_const, q_param, tf)
            if q_method == 1:
                c.add_rescale_legacy(q_param, [out])
            elif q_method == 2:
                c.add_rescale([q_param], [out])
            cur = [out]
        elif p[0] == "relu":
            c.add_relu(cur)
        elif p[0] == "sign":
            c.add_sign(cur)
        elif p[0] == "flatten":
            c.add_flatten(list(p[1]))
            cur = [p[1][0] * p[1][1] * p[1][2]]
    return c


def _assert_same_material(pa, pb):
    """Every section equal, except the bytes the reference's garbler never
    writes: a mini gate only fills the first two int16 of its 16-byte row
    (projection_gate_mini.h:28-44), the rest is uninitialised heap."""
    from dash_b200 import gcm

    ga, gb = gcm.read_gcm(pa), gcm.read_gcm(pb)
    assert (ga["crt"], ga["mrs"], ga["max_modulus"], ga["input_size"]) == (gb["crt"], gb["mrs"], gb["max_modulus"], gb["input_size"])
    assert np.array_equal(ga["zero"], gb["zero"]) and len(ga["layers"]) == len(gb["layers"])
    for la, lb in zip(ga["layers"], gb["layers"]):
        assert list(la["p"]) == list(lb["p"])
        for i, (x, y) in enumerate(zip(la["arrays"], lb["arrays"])):
            if int(la["p"][0]) == 3 and i == 5:  # ReLU evaluator table: [element][modulus][3] rows
                x = np.frombuffer(x, dtype=np.uint8).reshape(-1, 3, 16).copy()
                y = np.frombuffer(y, dtype=np.uint8).reshape(-1, 3, 16).copy()
                x[:, 2, 4:] = 0
                y[:, 2, 4:] = 0
                assert np.array_equal(x, y)
            else:
                assert x == y
    assert ga["dec"] == gb["dec"]
    for j in ga["in_base"]:
        assert np.array_equal(ga["in_base"][j], gb["in_base"][j])


@pytest.mark.parametrize("flavour", ["pytorch", "tf2onnx"])
@pytest.mark.parametrize("q_method,q_param", [(0, -1), (2, 2)])
def test_onnx_file_gives_the_same_garbled_circuit(tmp_path, flavour, q_method, q_param):
    from oracle import refbind as R

    if not R.available():
        pytest.skip("oracle/_ref not built")
    blob, params = cnn(flavour, 8, 8, 2, LAYERS, seed=3, raw=(flavour == "pytorch"))
    path = str(tmp_path / "m.onnx")
    open(path, "wb").write(blob)
    a = R.RefCircuit()
    n = a.load_onnx(path, q_method, 0.02, q_param)
    assert n == len(LAYERS) + (4 if q_method == 2 else 0)
    a.garble(CRT, MRS, 0, 7)
    a.export(str(tmp_path / "a.gcm"))
    b = _programmatic(R, params, flavour, q_method, 0.02, q_param, True)
    b.garble(CRT, MRS, 0, 7)
    b.export(str(tmp_path / "b.gcm"))
    _assert_same_material(str(tmp_path / "a.gcm"), str(tmp_path / "b.gcm"))
    # and it evaluates: one inference decodes to the plain quantised evaluation
    x = np.arange(8 * 8 * 2, dtype=np.int64) % 7
    out, _ = a.cpu_evaluate(a.garble_inputs(x), 2)
    assert np.array_equal(a.decode_outputs(out), b.decode_outputs(b.cpu_evaluate(b.garble_inputs(x), 2)[0]))


def test_malformed_model_is_rejected(tmp_path):
    from oracle import refbind as R

    if not R.available():
        pytest.skip("oracle/_ref not built")
    blob, _ = cnn("pytorch", 8, 8, 2, LAYERS, seed=3)
    path = str(tmp_path / "bad.onnx")
    open(path, "wb").write(blob[: len(blob) // 2])
    with pytest.raises(RuntimeError):
        R.RefCircuit().load_onnx(path)


def test_reference_fixture_model_dense():
    """dash/test/test_onnx_modelloader.h:10-43 restated: the reference's own
    fixture (dash/test/fixtures/model_dense.onnx, committed unchanged as test
    data under tests/golden/) read through integration/dash_onnx_lite.h gives
    the input/output dims and the six float values the reference's
    protobuf-based loader is tested for."""
    import os

    from helpers import GOLDEN_DIR
    from oracle import refbind as R

    if not R.available():
        pytest.skip("oracle/_ref not built")
    c = R.RefCircuit()
    n = c.load_onnx(os.path.join(GOLDEN_DIR, "model_dense.onnx"), 0, 10.0, -1)  # SimpleQuant, q_const 10
    assert n == 2                                    # Flatten + Dense
    assert c.plain_layer_dims(0) == [28, 28, 1]
    assert c.plain_layer_dims(1, output=True)[0] == 10
    w = c.plain_dense_floats(1)
    b = c.plain_dense_floats(1, biases=True)
    assert w.shape == (7840,) and b.shape == (10,)
    f32 = np.float32
    assert w[0] == f32(-0.02056167833507061004638671875)
    assert w[9] == f32(0.00375685538165271282196044921875)
    assert w[-1] == f32(0.018187098205089569091796875)
    assert b[0] == f32(-0.02799860946834087371826171875)
    assert b[4] == f32(-0.0069468836300075054168701171875)
    assert b[-1] == f32(-0.02649968676269054412841796875)
