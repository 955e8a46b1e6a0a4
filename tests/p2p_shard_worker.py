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
    if name.startswith("GARBLED:"):
        # a circuit garbled on the GPU (same seed on every rank): sharded evaluation against the unsharded one
        # of the same material (itself pinned against the oracle in tests/test_gpu_garbler.py)
        import workloads as W
        from dash_b200 import garble

        spec = W.MODELS[name[8:]]
        plain, n_in = W.plain_layers(spec, weight_seed=3)
        X = W.synthetic_inputs(spec, batch, seed=2)
        whole, _ = garble.garble_to_circuit(spec["crt"], spec["mrs"], n_in, plain, batch, seed=b"p2p")
        whole.garble_inputs(X)
        whole.evaluate_device()
        want = whole.move_outputs()
        whole.close()
        c, _ = garble.garble_to_circuit(spec["crt"], spec["mrs"], n_in, plain, batch, seed=b"p2p", shard=(rank, world),
                                        shard_lanes=lanes)
        ev = ChannelShardedEvaluator(c, rank, world, dist=dist, p2p=True)
        for rep in range(2):
            ev.load_values(X)
            ev.run()
            got = c.move_outputs()
            assert_labels_equal(got, want, "%s rank %d/%d repetition %d" % (name, rank, world, rep))
            assert not any(c.shard_timed_out(l) for l in range(c.lanes_used()))
        dist.barrier()
        c.close()
        dist.destroy_process_group()
        print("rank %d ok" % rank)
        return
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
