"""Output-channel sharding on one GPU: W circuits
with shards (r, W) stand in for W ranks, the exchange of activation slices is
done with device-to-device copies instead of NCCL.  Results must be
bit-identical to the unsharded evaluation / the reference.  The real
multi-process NCCL path is tools/run_channel_shard.py (torchrun)."""
import numpy as np
import pytest

from helpers import Golden, assert_labels_equal, tile_batch

pytestmark = pytest.mark.gpu


def _planes(torch, c):
    from dash_b200.scheduler import _DeviceBytes

    out = []
    for j in range(c.k):
        ptr, bpe, n = c.current_plane(j)
        out.append((torch.as_tensor(_DeviceBytes(ptr, bpe * n), device="cuda"), bpe))
    return out


@pytest.mark.parametrize("name", ["MLP_TINY", "CONV_UNIT", "SIGN_TINY", "RESCALE_OPT", "RESCALE_LEGACY", "PROJECTION"])
@pytest.mark.parametrize("world", [2, 3])
def test_channel_sharded_equals_reference(gpu, name, world):
    import torch

    from dash_b200.scheduler import layer_groups

    g = Golden(name)
    batch = 5
    labels = tile_batch(g.inputs, batch)
    ranks = [gpu.Circuit.load(g.gcm, max_batch=batch, shard=(r, world)) for r in range(world)]
    groups = layer_groups([ranks[0].layer_type(i) for i in range(ranks[0].n_layers)])
    for c in ranks:
        c.move_inputs(labels)
    for first, last in groups:
        for c in ranks:
            c.evaluate_layers(first, last)
        for c in ranks:
            c.sync()
        planes = [_planes(torch, c) for c in ranks]
        for src in range(world):
            e0, e1 = ranks[0].layer_slice(last, src)
            for dst in range(world):
                if dst == src:
                    continue
                for j in range(ranks[0].k):
                    p_src, bpe = planes[src][j]
                    planes[dst][j][0][e0 * bpe:e1 * bpe].copy_(p_src[e0 * bpe:e1 * bpe])
        torch.cuda.synchronize()
    want = tile_batch(g.layers[-1], batch)
    for r, c in enumerate(ranks):
        assert_labels_equal(c.move_outputs(), want, "%s rank %d/%d" % (name, r, world))
        c.close()


def test_sharded_circuit_refuses_unsharded_call(gpu):
    g = Golden("MLP_TINY")
    c = gpu.Circuit.load(g.gcm, max_batch=2, shard=(0, 2))
    c.move_inputs(tile_batch(g.inputs, 2))
    with pytest.raises(gpu.DashB200Error):
        c.evaluate_device()
    c.close()
