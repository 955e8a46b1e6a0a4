"""dashb_projection_static: the entry point that allocates nothing.

Collected last on purpose (file name): a run that stops at the first failure
has then already run every other test.

Who needs it: a host that may not allocate while it evaluates -- its device
memory was handed to its own arena up front, or every cudaMalloc would be an
ocall out of an SGX enclave, or a malloc/free would stall a latency-critical
loop.  Such a caller owns the label and table buffers (here: torch tensors
stand in for that arena) and the library keeps the digit planes it needs in
between in a fixed 8 MiB array inside its own image.

Checked here, through the C ABI, on two calls in a row (the second one reuses
the array the first one dirtied, with another batch size): the output labels
equal, digit for digit, those of the allocating path (Circuit.load +
dashb_layer_outputs) on the same inputs and tables, and the labels the
reference itself computed for that layer (the golden fixture).
"""
import ctypes as C

import numpy as np
import pytest

from helpers import Golden, read_gcm_layers, tile_batch

pytestmark = pytest.mark.gpu


def _static_call(api, torch, labels_in, ciphers_bytes, n, q_in, q_out):
    """labels_in: int16 [batch][n][nc(q_in)] on the host."""
    L = api.lib()
    L.dashb_projection_static.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64, C.c_int, C.c_int, C.c_void_p]
    batch = labels_in.shape[0]
    # the caller's own device memory (its "arena")
    d_in = torch.from_numpy(np.ascontiguousarray(labels_in)).cuda()
    d_ci = torch.from_numpy(np.frombuffer(ciphers_bytes, dtype=np.uint8).copy()).cuda()
    d_out = torch.full((batch, n, api.nr_comps(q_out)), -1, dtype=torch.int16, device="cuda")
    torch.cuda.synchronize()
    rc = L.dashb_projection_static(d_in.data_ptr(), d_ci.data_ptr(), d_out.data_ptr(), n, batch, q_in, q_out, None)
    assert rc == 0, api.lib().dashb_last_error().decode()
    stream = torch.cuda.ExternalStream(api.stream_handle())
    stream.synchronize()
    return d_out.cpu().numpy()


def test_projection_static_matches_allocating_path_twice(gpu):
    import torch

    api = gpu
    g = Golden("PROJECTION")  # 19 -> 91 -> 19 (dash/test/test_projection.h:18-50), k = 1
    head, layers = read_gcm_layers(g.gcm)
    params, arrays = layers[0]
    n, q_in, q_out = int(params[1]), int(params[2]), int(params[3])
    assert (q_in, q_out) == (19, 91)

    circ = api.Circuit.load(g.gcm, max_batch=64)
    for batch in (3, 37):  # second call: other size, same (now dirty) static array
        labels = tile_batch(g.inputs, batch)
        if batch == 37:
            labels = [l[::-1].copy() for l in labels]
        circ.move_inputs(labels)
        want = circ.layer_outputs(0, n, [q_out])[0]
        got = _static_call(api, torch, labels[0], arrays[0], n, q_in, q_out)
        assert got.shape == want.shape
        assert np.array_equal(got, want), "static-scratch projection differs from the allocating path (batch %d)" % batch
        # and directly against the reference's own labels for this layer
        ref_want = tile_batch(g.layers[0], batch)[0]
        if batch == 37:
            ref_want = ref_want[::-1]
        assert np.array_equal(got, ref_want), "static-scratch projection differs from the reference's labels (batch %d)" % batch
    circ.close()


def test_projection_static_rejects_oversize(gpu):
    import torch

    api = gpu
    L = api.lib()
    L.dashb_projection_static.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64, C.c_int, C.c_int, C.c_void_p]
    dummy = torch.zeros(16, dtype=torch.int16, device="cuda")
    # 4096 gates x 256 inferences of modulus 2 (128 digits) need far more than 8 MiB of planes
    rc = L.dashb_projection_static(dummy.data_ptr(), dummy.data_ptr(), dummy.data_ptr(), 4096, 256, 2, 2, None)
    assert rc == -1
    msg = L.dashb_last_error().decode()
    assert "static scratch" in msg and "split" in msg
