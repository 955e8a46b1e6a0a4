"""Parity of the CUDA evaluator (through the C ABI) with the reference:
bit-exact output labels, per-layer labels and decoded values.  Run on the
B200 box: `pytest -m gpu`."""
import numpy as np
import pytest

from helpers import GOLDEN, Golden, assert_labels_equal, tile_batch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", GOLDEN)
@pytest.mark.parametrize("batch", [1, 3, 17, 40])
def test_golden_final_labels(gpu, name, batch):
    g = Golden(name)
    c = gpu.Circuit.load(g.gcm, max_batch=batch)
    got = c.evaluate(tile_batch(g.inputs, batch))
    assert_labels_equal(got, tile_batch(g.layers[-1], batch), "%s final" % name)
    if g.dec is not None:
        vals, ok = c.decode_outputs()
        assert ok.all()
        assert np.array_equal(vals, tile_batch([g.dec], batch)[0])
    c.close()


@pytest.mark.parametrize("name", GOLDEN)
def test_golden_every_layer(gpu, name):
    g = Golden(name)
    c = gpu.Circuit.load(g.gcm, max_batch=g.batch)
    c.move_inputs(g.inputs)
    for l in range(g.n_layers):
        want = g.layers[l]
        mods_nc = [w.shape[2] for w in want]
        out = [np.zeros((g.batch, want[0].shape[1], nc), dtype=np.int16) for nc in mods_nc]
        from dash_b200.api import _check, _ptr_array
        _check(c.L.dashb_layer_outputs(c.h, l, _ptr_array(out)))
        assert_labels_equal(out, want, "%s layer %d" % (name, l))
    c.close()


@pytest.mark.parametrize("name", ["MLP_TINY", "RELU_EDGE", "SIGN_TINY", "RESCALE_2_3", "RESCALE_LEGACY", "RESCALE_OPT"])
def test_division_fallback(gpu, name, monkeypatch):
    """The gadget kernels take a label's digits by multiplication where the
    modulus has a plan and by long division otherwise (a modulus cut into more
    than 10 chunks); no circuit of the configurations has such a modulus, so
    DASHB_NO_FRAC switches the plans off when the circuit is created."""
    if name not in GOLDEN:
        pytest.skip("no such golden circuit")
    monkeypatch.setenv("DASHB_NO_FRAC", "1")
    g = Golden(name)
    batch = 19
    c = gpu.Circuit.load(g.gcm, max_batch=batch)
    got = c.evaluate(tile_batch(g.inputs, batch))
    assert_labels_equal(got, tile_batch(g.layers[-1], batch), "%s final, division path" % name)
    c.close()


# ---- circuits garbled on the box by the reference build (oracle/_ref) -----
def _live(name, gpu, tmp_path, batch, seed=77):
    import workloads as W
    from oracle import oraclebind as O
    from oracle import refbind as R

    if not R.available():
        pytest.skip("oracle/_ref not present on this box")
    spec = W.MODELS[name]
    ref = W.build_reference_circuit(spec, seed=seed, weight_seed=seed)
    path = str(tmp_path / (name + ".gcm"))
    ref.export(path)
    X = W.synthetic_inputs(spec, batch, seed=seed)
    per = [ref.garble_inputs(x) for x in X]
    labels = [np.stack([p[j] for p in per]) for j in range(ref.k)]
    c = gpu.Circuit.load(path, max_batch=batch)
    got = c.evaluate(labels)
    return ref, O.OracleCircuit(path), labels, c, got


@pytest.mark.parametrize("name,batch", [("MNIST_TINY_CNN", 33), ("CIFAR_TINY_SCALE", 20), ("SIGN_TINY", 64),
                                        ("MODEL_E_30", 5), ("RESCALE_LEGACY", 48), ("RESCALE_2_3", 37),
                                        ("RESCALE_OPT", 9)])
def test_live_reference_circuits(gpu, tmp_path, name, batch):
    ref, oc, labels, c, got = _live(name, gpu, tmp_path, batch)
    vals, ok = c.decode_outputs()
    assert ok.all()
    for b in range(0, batch, max(1, batch // 6)):
        lab = [l[b] for l in labels]
        want, _ = ref.cpu_evaluate(lab, 4)              # the reference itself
        assert_labels_equal([g[b] for g in got], want, "%s inference %d vs reference" % (name, b))
        assert_labels_equal(oc.evaluate(lab, threads=4), want, "%s oracle vs reference" % name)
        assert np.array_equal(vals[b], ref.decode_outputs(want))
    c.close()


def test_rescale_wide_accumulators(gpu, tmp_path):
    """CRT base {2,3,251}: (k+1)(q-1)^2 >= 2^16, so k_rescale runs with its
    32-bit accumulators.  Labels only: this base is outside the range where
    the reference's own decode finds a match."""
    import workloads as W
    from oracle import refbind as R

    if not R.available():
        pytest.skip("oracle/_ref not present on this box")
    spec = W.MODELS["RESCALE_WIDE"]
    ref = W.build_reference_circuit(spec, seed=77, weight_seed=77)
    path = str(tmp_path / "rw.gcm")
    ref.export(path)
    batch = 21
    X = W.synthetic_inputs(spec, batch, seed=77)
    per = [ref.garble_inputs(x) for x in X]
    labels = [np.stack([p[j] for p in per]) for j in range(ref.k)]
    c = gpu.Circuit.load(path, max_batch=batch)
    got = c.evaluate(labels)
    for b in range(batch):
        want, _ = ref.cpu_evaluate([l[b] for l in labels], 2)
        assert_labels_equal([g[b] for g in got], want, "RESCALE_WIDE inference %d" % b)
    c.close()


def test_full_size_model_b_batch_256(gpu, tmp_path):
    """BASELINE.json configs[1] at its full size: MODEL_B_POOL_REPL, batch 256.
    Bit-exact against the reference on a sample of the batch, plus size-
    independent properties: an inference's labels do not depend on its
    position or neighbours in the batch, and every output decodes."""
    import bench
    from oracle import refbind as R

    if not R.available():
        pytest.skip("oracle/_ref not present on this box")
    B = 256
    gcm, labels, ref = bench.build_material("MODEL_B_POOL_REPL", B, want_ref=True)
    c = gpu.Circuit.load(gcm, max_batch=B)
    got = c.evaluate(labels)
    vals, ok = c.decode_outputs()
    assert ok.all()
    for b in (0, 1, 31, 32, 100, 255):
        want, _ = ref.cpu_evaluate([l[b] for l in labels], 8)
        assert_labels_equal([g[b] for g in got], want, "MODEL_B inference %d" % b)
        assert np.array_equal(vals[b], ref.decode_outputs(want))
    # permutation property: reverse the batch, outputs reverse with it
    rev = c.evaluate([l[::-1].copy() for l in labels])
    assert_labels_equal([r[::-1] for r in rev], got, "batch permutation")
    # a smaller, ragged batch reproduces the same labels
    sub = c.evaluate([l[:37].copy() for l in labels])
    assert_labels_equal(sub, [g[:37] for g in got], "ragged batch 37")
    c.close()


@pytest.mark.parametrize("name", ["MLP_TINY", "CONV_UNIT", "RESCALE_2_3", "RELU_EDGE"])
@pytest.mark.parametrize("batch", [3, 70])
def test_device_garble_inputs(gpu, name, batch):
    """garble_inputs on the device (garbled_circuit_interface.h:325-343):
    plaintext inputs -> labels -> evaluation gives the same output labels and
    decoded values as feeding the reference's own input labels."""
    g = Golden(name)
    c = gpu.Circuit.load(g.gcm, max_batch=batch)
    x = tile_batch([g.x], batch)[0]
    c.garble_inputs(x)
    c.evaluate_device()
    got = c.move_outputs()
    assert_labels_equal(got, tile_batch(g.layers[-1], batch), "%s via device garble_inputs" % name)
    if g.dec is not None:
        vals, ok = c.decode_outputs()
        assert ok.all() and np.array_equal(vals, tile_batch([g.dec], batch)[0])
    c.close()


def test_inprocess_cpp_example(gpu):
    """integration/example_inprocess.cpp: a Dash program (reference classes,
    compiled from the unmodified headers) that hands its garbled circuit to
    the B200 evaluator through integration/dash_adapter.h + the C ABI and
    decodes the result with the reference's own decode_outputs."""
    import os
    import subprocess

    from helpers import ROOT

    exe = os.path.join(ROOT, "oracle", "_ref", "example_inprocess")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/example_inprocess not built")
    res = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert res.returncode == 0, res.stdout + res.stderr
    line = [l for l in res.stdout.splitlines() if l.startswith("b200:")][0]
    b200, cpu = line.split("cpu:")
    assert b200.replace("b200:", "").split() == cpu.split() == ["3", "3"]


@pytest.mark.parametrize("M,K", [(16, 33), (50, 100), (130, 257), (128, 128), (200, 640)])
@pytest.mark.parametrize("batch", [1, 17, 40])
def test_dense_tensor_core_path(gpu, tmp_path, M, K, batch):
    """Dense layers large enough for the tcgen05 kind::i8 kernel (dense_tc.cu):
    odd shapes (rows/taps not multiples of the 128-wide tiles, ragged batch),
    followed by a ReLU so that a wrong digit anywhere would also break the
    hashes.  Bit-exact against the reference."""
    import workloads as W
    from oracle import refbind as R

    if not R.available():
        pytest.skip("oracle/_ref not present on this box")
    spec = dict(input=(K, 1, 1), layers=[("flatten",), ("dense", M), ("relu",)], quant=("simple", 0.02), **W.cpm(8, 99.0))
    ref = W.build_reference_circuit(spec, seed=K, weight_seed=M)
    path = str(tmp_path / "dense.gcm")
    ref.export(path)
    X = W.synthetic_inputs(spec, batch, seed=3)
    per = [ref.garble_inputs(x) for x in X]
    labels = [np.stack([p[j] for p in per]) for j in range(ref.k)]
    c = gpu.Circuit.load(path, max_batch=batch)
    got = c.evaluate(labels)
    for b in sorted({0, batch // 2, batch - 1}):
        want, _ = ref.cpu_evaluate([l[b] for l in labels], 4)
        assert_labels_equal([g[b] for g in got], want, "dense %dx%d inference %d" % (M, K, b))
    c.close()


@pytest.mark.parametrize("geom", [
    # (width, height, channels, filters, filter size, stride)
    (8, 8, 8, 16, 3, 1),     # K = 72: one K block, 14 channels would fit
    (10, 9, 20, 40, 3, 1),   # 20 channels = 14 + 6: second box partly out of bounds
    (12, 12, 16, 32, 2, 2),  # 2x2/2 "pool replacement", K = 64
    (9, 9, 33, 130, 2, 1),   # two filter blocks (130 > 128), 33 channels = 32 + 1
    (7, 7, 3, 16, 5, 2),     # 5x5 filter: 5 channels per block, only 3 present
    (9, 9, 20, 64, 3, 1),    # 64 filters: filters-on-N kernel, two K blocks
    (6, 6, 40, 128, 3, 1),   # 128 filters on N, three K blocks
    (8, 8, 8, 96, 2, 2),     # 96 filters on N (TMEM allocation rounds up to 128 columns)
    # stride 1, 32 or 64 filters: the strip kernel (k_conv_tc_strip, several output pixels per CTA)
    (10, 10, 32, 32, 3, 1),  # one channel block of 32, strips of 4 over 8 output columns
    (11, 9, 64, 64, 3, 1),   # two channel blocks, 9 output columns: last strip partly outside
    (9, 9, 64, 64, 1, 1),    # 1x1 filter: one column per pixel
    (7, 7, 33, 32, 3, 1),    # 33 channels: second block holds one channel, the rest zero-filled
    (12, 12, 16, 32, 2, 1),  # 2x2 filter, 16 channels: 32 rows per column
    (30, 30, 32, 32, 3, 1),  # the GNNP 32 -> 32 convolution at full width
])
@pytest.mark.parametrize("batch", [1, 19])
def test_conv_tensor_core_path(gpu, tmp_path, geom, batch, monkeypatch):
    """Convolutions wide enough for the tcgen05 kernel (im2col by a 4-D TMA
    box, dense_tc.cu), followed by a ReLU; bit-exact against the reference
    and against the dp4a kernel's result for the same circuit."""
    import workloads as W
    from oracle import refbind as R

    if not R.available():
        pytest.skip("oracle/_ref not present on this box")
    w, h, ch, filters, f, stride = geom
    monkeypatch.setenv("DASHB_CONV_N_MAX", "128")  # read once per process: also exercises the filters-on-N kernel at 64..128 filters
    spec = dict(input=(w, h, ch), layers=[("conv", filters, f, stride), ("relu",)], quant=("simple", 0.02), **W.cpm(8, 99.0))
    ref = W.build_reference_circuit(spec, seed=w * h, weight_seed=filters)
    path = str(tmp_path / "conv.gcm")
    ref.export(path)
    X = W.synthetic_inputs(spec, batch, seed=5)
    per = [ref.garble_inputs(x) for x in X]
    labels = [np.stack([p[j] for p in per]) for j in range(ref.k)]
    c = gpu.Circuit.load(path, max_batch=batch)
    got = c.evaluate(labels)
    for b in sorted({0, batch - 1}):
        want, _ = ref.cpu_evaluate([l[b] for l in labels], 4)
        assert_labels_equal([g[b] for g in got], want, "conv %s inference %d" % (geom, b))
    c.close()


@pytest.mark.parametrize("name", ["MLP_TINY", "CONV_UNIT", "RESCALE_2_3"])
@pytest.mark.parametrize("batch", [1, 37, 130])
def test_compressed_wire_labels(gpu, name, batch):
    """dashb_evaluate_compressed / move_*_compressed: labels cross the
    boundary as LabelTensor::compress leaves them (16 bytes each); the result,
    decompressed, is the int16-path result bit for bit."""
    from dash_b200 import gcm

    g = Golden(name)
    c = gpu.Circuit.load(g.gcm, max_batch=batch)
    labels = tile_batch(g.inputs, batch)
    want = c.evaluate(labels)
    info = gcm.read_gcm(g.gcm)
    comp = [gcm.compress_labels(q, l) for q, l in zip(info["crt"], labels)]
    got = c.evaluate_compressed(comp)
    for j, q in enumerate(c.out_moduli):
        assert np.array_equal(gcm.decompress_labels(q, got[j]), want[j]), "modulus %d" % q
    # the separate calls
    c.move_inputs_compressed(comp)
    c.evaluate_device()
    again = c.move_outputs_compressed()
    for a, b in zip(again, got):
        assert np.array_equal(a, b)
    assert_labels_equal(c.move_outputs(), want, "int16 outputs after compressed inputs")
    c.close()


@pytest.mark.parametrize("name", ["MLP_TINY", "CONV_UNIT", "RESCALE_2_3"])
@pytest.mark.parametrize("batch,skew", [(1, 0), (37, 1), (130, 3), (256, 5)])
def test_direct_ingest_from_pinned_host_memory(gpu, name, batch, skew):
    """Pinned (mapped) host buffers are read by k_ingest_direct itself, without
    the staging copy; the labels it lays out must be those of the staged path
    (pageable numpy buffers).  `skew` shifts every buffer by that many int16
    from its allocation, so the 16-byte windows of the kernel start at every
    possible misalignment, and the buffers end exactly at the last digit."""
    import ctypes as C

    import torch

    from dash_b200.api import _check, _ptr_array

    g = Golden(name)
    c = gpu.Circuit.load(g.gcm, max_batch=batch)
    labels = tile_batch(g.inputs, batch)
    want = c.evaluate(labels)  # pageable numpy arrays: staged path
    pinned, views = [], []
    for l in labels:
        t = torch.empty(l.size + skew, dtype=torch.int16).pin_memory()
        v = t[skew:].view(l.shape)
        v.numpy()[...] = l
        pinned.append(t)
        views.append(v)
    out = [torch.empty(w.shape, dtype=torch.int16).pin_memory() for w in want]
    in_ptrs = (C.c_void_p * len(views))(*[v.data_ptr() for v in views])
    out_ptrs = (C.c_void_p * len(out))(*[o.data_ptr() for o in out])
    _check(c.L.dashb_evaluate(c.h, in_ptrs, batch, out_ptrs))
    assert_labels_equal([o.numpy() for o in out], want, "%s pinned inputs (direct ingest)" % name)
    # the separate calls: move_inputs from pinned memory, evaluate on the device, outputs back
    _check(c.L.dashb_move_inputs(c.h, in_ptrs, batch))
    c.batch = batch
    c.evaluate_device()
    assert_labels_equal(c.move_outputs(), want, "%s pinned inputs, separate calls" % name)
    c.close()


@pytest.mark.parametrize("name", ["MLP_TINY", "CONV_UNIT"])
def test_async_evaluation_two_steps_in_flight(gpu, name):
    """dashb_evaluate_async / dashb_evaluate_wait: a stream of batches with two
    steps in flight gives, batch for batch, the labels of the synchronous call;
    batches differ in size and content so that a mixed-up lane set or buffer
    would show."""
    import ctypes as C

    import torch

    from dash_b200.api import _check

    g = Golden(name)
    c = gpu.Circuit.load(g.gcm, max_batch=160)
    L = c.L
    L.dashb_evaluate_async.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p]
    L.dashb_evaluate_wait.argtypes = [C.c_void_p]
    batches = [160, 37, 130, 1, 96, 160, 64]
    keep, want = [], []
    for r, b in enumerate(batches):
        labels = [np.roll(l, r, axis=0) for l in tile_batch(g.inputs, b)]
        want.append(c.evaluate(labels))
        ins = []
        for l in labels:
            t = torch.empty(l.shape, dtype=torch.int16).pin_memory()
            t.numpy()[...] = l
            ins.append(t)
        outs = [torch.full(w.shape, -7, dtype=torch.int16).pin_memory() for w in want[-1]]
        keep.append((ins, outs))
    for (ins, outs), b in zip(keep, batches):
        ip = (C.c_void_p * len(ins))(*[t.data_ptr() for t in ins])
        op = (C.c_void_p * len(outs))(*[t.data_ptr() for t in outs])
        _check(L.dashb_evaluate_async(c.h, ip, b, op))
    _check(L.dashb_evaluate_wait(c.h))
    for r, ((ins, outs), w) in enumerate(zip(keep, want)):
        assert_labels_equal([o.numpy() for o in outs], w, "%s async step %d" % (name, r))
    # a synchronous call after the asynchronous ones still works
    labels = tile_batch(g.inputs, 5)
    assert_labels_equal(c.evaluate(labels), tile_batch(g.layers[-1], 5), "%s sync after async" % name)
    c.close()


@pytest.mark.parametrize("name", ["MLP_TINY", "CONV_UNIT"])
def test_async_evaluation_compressed_labels(gpu, name):
    """dashb_evaluate_compressed_async: the streaming call with 128-bit labels
    on both sides returns, batch for batch, what the synchronous compressed
    call returns."""
    import ctypes as C

    import torch

    from dash_b200 import gcm as gcm_io
    from dash_b200.api import _check

    g = Golden(name)
    crt = gcm_io.read_gcm(g.gcm)["crt"]
    c = gpu.Circuit.load(g.gcm, max_batch=96)
    L = c.L
    L.dashb_evaluate_compressed_async.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p]
    L.dashb_evaluate_wait.argtypes = [C.c_void_p]
    batches = [96, 5, 64, 33, 96]
    keep, want = [], []
    for r, b in enumerate(batches):
        labels = [np.roll(l, r, axis=0) for l in tile_batch(g.inputs, b)]
        comp = [np.ascontiguousarray(gcm_io.compress_labels(q, l)) for q, l in zip(crt, labels)]
        want.append(c.evaluate_compressed(comp))
        ins = []
        for l in comp:
            t = torch.empty(l.shape, dtype=torch.uint64).pin_memory()
            t.numpy()[...] = l
            ins.append(t)
        outs = [torch.zeros(w.shape, dtype=torch.uint64).pin_memory() for w in want[-1]]
        keep.append((ins, outs))
    for (ins, outs), b in zip(keep, batches):
        ip = (C.c_void_p * len(ins))(*[t.data_ptr() for t in ins])
        op = (C.c_void_p * len(outs))(*[t.data_ptr() for t in outs])
        _check(L.dashb_evaluate_compressed_async(c.h, ip, b, op))
    _check(L.dashb_evaluate_wait(c.h))
    for r, ((ins, outs), w) in enumerate(zip(keep, want)):
        for j, (o, x) in enumerate(zip(outs, w)):
            assert np.array_equal(o.numpy(), x), "%s compressed async step %d modulus %d" % (name, r, j)
    c.close()


@pytest.mark.parametrize("name", ["MLP_TINY", "CONV_UNIT", "SIGN_TINY", "RELU_EDGE"])
@pytest.mark.parametrize("batch,replicas", [(37, 37), (130, 130), (40, 7)])
def test_table_replicas_give_the_same_labels(gpu, name, batch, replicas):
    """dashb_set_table_replicas: the sign / ReLU tables stored once per
    inference (rows interleaved by inference) -- the memory layout of one
    garbling per inference -- must evaluate to the same labels as the shared
    tables, whatever lane an inference falls into (replica = global inference
    index mod replicas)."""
    g = Golden(name)
    c = gpu.Circuit.load(g.gcm, max_batch=batch, table_replicas=replicas)
    got = c.evaluate(tile_batch(g.inputs, batch))
    assert_labels_equal(got, tile_batch(g.layers[-1], batch), "%s with %d table replicas" % (name, replicas))
    c.close()
