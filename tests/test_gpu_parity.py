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
