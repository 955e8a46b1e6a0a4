"""One rank of tests/test_gpu_p2p_shard.py: `world` processes on the SAME GPU
stand in for `world` GPUs.  The exchange is the real one -- CUDA IPC mappings
of the other processes' gather buffers, k_shard_push / signal / wait -- only
the wires are the GPU's own memory instead of NVLink."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    rank, world, port, name, batch, lanes = int(sys.argv[1]), int(sys.argv[2]), sys.argv[3], sys.argv[4], int(sys.argv[5]), int(sys.argv[6])
    import torch.distributed as dist

    from dash_b200 import api
    from dash_b200.scheduler import ChannelShardedEvaluator
    from helpers import Golden, assert_labels_equal, tile_batch

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = port
    dist.init_process_group("gloo", rank=rank, world_size=world)
    api.init(0)
    g = Golden(name)
    c = api.Circuit.load(g.gcm, max_batch=batch, shard=(rank, world), shard_lanes=lanes)
    ev = ChannelShardedEvaluator(c, rank, world, dist=dist, p2p=True)
    for rep in range(3):  # epochs keep counting across evaluations; other inputs each time
        labels = [np.roll(l, rep, axis=0) for l in tile_batch(g.inputs, batch)]
        want = [np.roll(l, rep, axis=0) for l in tile_batch(g.layers[-1], batch)]
        got = ev.evaluate(labels)
        assert_labels_equal(got, want, "%s rank %d/%d repetition %d" % (name, rank, world, rep))
    dist.barrier()
    c.close()
    dist.destroy_process_group()
    print("rank %d ok" % rank)


if __name__ == "__main__":
    main()
