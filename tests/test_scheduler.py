"""Batch sharding across ranks: host logic, exercised with world_size 2 on
the gloo backend (CPU)."""
import os
import socket

import numpy as np
import pytest
import torch.multiprocessing as mp

from dash_b200.scheduler import shard_range


def test_shard_range_partitions():
    for total in [0, 1, 5, 256, 257]:
        for world in [1, 2, 3, 8]:
            spans = [shard_range(total, world, r) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == total
            for (a, b), (c, d) in zip(spans, spans[1:]):
                assert b == c
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard_range(4, 2, 2)


def _worker(rank, world, port, total, q):
    import torch.distributed as dist

    from dash_b200.scheduler import BatchShardedEvaluator

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rng = np.random.default_rng(0)
    labels = [rng.integers(0, 7, size=(total, 6, 5), dtype=np.int16), rng.integers(0, 3, size=(total, 6, 9), dtype=np.int16)]

    def fake_eval(ls):  # stands in for Circuit.evaluate: any per-inference function
        return [(ls[0][:, :2, :] * 3 + 1).astype(np.int16), (ls[1][:, :2, :] + 2).astype(np.int16)]

    ev = BatchShardedEvaluator(fake_eval, [(2, 5), (2, 9)])
    out = ev.evaluate(labels)
    want = fake_eval(labels)
    ok = all(np.array_equal(a, b) for a, b in zip(out, want))
    q.put((rank, ok))
    dist.destroy_process_group()


@pytest.mark.parametrize("total", [7, 8, 1])
def test_batch_sharded_gather_world2(total):
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, total, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert sorted(res) == [(0, True), (1, True)]


def test_layer_groups():
    from dash_b200.scheduler import layer_groups

    D, C2, R, S, RS, P = 1, 2, 3, 4, 5, 7
    assert layer_groups([C2, R, C2, R, D, R, D]) == [(0, 1), (2, 3), (4, 5), (6, 6)]
    assert layer_groups([R, D]) == [(0, 0), (1, 1)]
    assert layer_groups([C2, RS, R, C2, RS]) == [(0, 2), (3, 4)]
    assert layer_groups([D]) == [(0, 0)]
    assert layer_groups([]) == []
