#!/usr/bin/env python
"""BASELINE.json configs[2] and configs[3] on the GPUs of one box, with the
material garbled on every rank's own GPU (same seed -> identical circuits):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29544 tools/run_multi_gpu_cifar.py batch   MODEL_F_GNNP_POOL_REPL 16
    ... tools/run_multi_gpu_cifar.py channel MODEL_F_MINIONN_POOL_REPL 16

batch:   every rank evaluates its own BATCH inferences (weak scaling, no
         data-path collective), value = N * BATCH / max-over-ranks step time.
channel: output-channel sharding: one batch, every rank owns
         a slice of each layer, activation slices exchanged over NCCL.
Rank 0 checks one inference label for label against the CPU oracle."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import workloads as W  # noqa: E402
from dash_b200 import api, garble, gcm  # noqa: E402
from dash_b200.scheduler import ChannelShardedEvaluator  # noqa: E402


def main():
    mode, model, B = sys.argv[1], sys.argv[2], int(sys.argv[3])
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    api.init(local)
    spec = W.MODELS[model]
    plain, input_size = W.plain_layers(spec)
    path = "/tmp/mg_%s_%d.gcm" % (model, rank)
    t0 = time.time()
    st = garble.garble_to_file(spec["crt"], spec.get("mrs", []), input_size, plain, path, seed=b"multi-gpu-cifar")  # every rank: same seed, same circuit
    garble_s = time.time() - t0
    g = gcm.read_gcm(path) if (rank == 0 or mode == "channel") else None
    X = W.synthetic_inputs(spec, B, seed=100 + (rank if mode == "batch" else 0))
    stream_of = lambda: torch.cuda.ExternalStream(api.stream_handle())  # noqa: E731
    steps = 5
    if mode == "batch":
        c = api.Circuit.load(path, max_batch=B)
        c.garble_inputs(X)
        for _ in range(2):
            c.evaluate_device()
        s = stream_of()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(s)
        for _ in range(steps):
            c.evaluate_device()
        e1.record(s)
        e1.synchronize()
        ms = e0.elapsed_time(e1) / steps
        out = c.move_outputs()
        nbytes = 0
    else:
        c = api.Circuit.load(path, max_batch=B, shard=(rank, world))
        ev = ChannelShardedEvaluator(c, rank, world, dist=dist if world > 1 else None, compressed="--compressed" in sys.argv)
        labels = [np.stack([gcm.encode_inputs(g, x)[j] for x in X]) for j in range(len(spec["crt"]))]
        out = ev.evaluate(labels)
        ev.bytes_exchanged = 0
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(steps):
            for first, last in ev.groups:
                c.evaluate_layers(first, last)
                ev._gather(last)
        torch.cuda.synchronize()
        ms = (time.perf_counter() - t0) * 1e3 / steps
        nbytes = ev.bytes_exchanged // steps
    t = torch.tensor([ms], device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())
    if rank == 0:
        from oracle import oraclebind as O

        oc = O.OracleCircuit(path)
        want = oc.evaluate(gcm.encode_inputs(g, X[0]), threads=len(os.sched_getaffinity(0)))
        exact = all(np.array_equal(o[0], w) for o, w in zip(out, want))
        total = B * (world if mode == "batch" else 1)
        print(json.dumps({"case": mode + "_sharded", "model": model, "n_gpus": world, "batch_per_step": total,
                          "ms_per_step": ms, "inferences_per_s": total / (ms / 1e3), "garble_s_per_rank": garble_s,
                          "garble_kernel_ms": st["kernel_ms"], "table_GB": st["table_bytes"] / 1e9,
                          "nvlink_bytes_received_per_step_per_rank": nbytes, "bit_exact_vs_oracle": bool(exact),
                          "scaling": "weak" if mode == "batch" else "strong",
                          "exchange": None if mode == "batch" else ("compressed labels" if "--compressed" in sys.argv else "digit planes")}), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
