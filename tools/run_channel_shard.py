#!/usr/bin/env python
"""Output-channel-sharded evaluation across the GPUs of one box with NCCL
(one process per GPU):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N \
        --master-addr 127.0.0.1 --master-port 29533 tools/run_channel_shard.py [MODEL] [BATCH]

Every rank garbles/loads the same material, evaluates its channel slices,
exchanges activation slices over NVLink, and rank 0 checks the result against
the reference and prints one JSON line with timings."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import bench  # noqa: E402
from dash_b200 import api  # noqa: E402
from dash_b200.scheduler import ChannelShardedEvaluator  # noqa: E402


def main():
    model = sys.argv[1] if len(sys.argv) > 1 else "MODEL_B_POOL_REPL"
    B = int(sys.argv[2]) if len(sys.argv) > 2 else 64
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    if rank == 0:
        gcm, labels, ref = bench.build_material(model, B, want_ref=True)
    dist.barrier()
    if rank != 0:
        gcm, labels, ref = bench.build_material(model, B)
    api.init(local)
    c = api.Circuit.load(gcm, max_batch=B, shard=(rank, world))
    ev = ChannelShardedEvaluator(c, rank, world, dist=dist)
    out = ev.evaluate(labels)  # warm-up + result
    ev.bytes_exchanged = 0
    dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    reps = 5
    for _ in range(reps):
        for first, last in ev.groups:
            c.evaluate_layers(first, last)
            ev._gather(last)
    torch.cuda.synchronize()
    dist.barrier()
    ms = (time.perf_counter() - t0) * 1e3 / reps
    t = torch.tensor([ms], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        ok = True
        for b in (0, B // 2, B - 1):
            want, _ = ref.cpu_evaluate([l[b] for l in labels], 8)
            ok = ok and all(np.array_equal(o[b], w) for o, w in zip(out, want))
        print(json.dumps({"case": "channel_sharded", "model": model, "batch": B, "n_gpus": world,
                          "ms_per_batch": float(t.item()), "inferences_per_s": B / (float(t.item()) / 1e3),
                          "groups": ev.groups, "nvlink_bytes_received_per_batch_per_rank": ev.bytes_exchanged // reps,
                          "bit_exact_vs_reference": bool(ok)}), flush=True)
    outs = [None] * world
    dist.all_gather_object(outs, [o.tobytes() for o in out][:1])
    if rank == 0:
        assert all(o == outs[0] for o in outs), "ranks disagree"
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
