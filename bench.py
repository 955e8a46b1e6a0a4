#!/usr/bin/env python
"""bench.py -- online garbled-evaluation throughput of the B200 evaluator.

    python bench.py --gpus N --steps K --warmup W [--impl reference]

One "step" = one pass of the hot path (every garbled layer of the model)
over one batch of synthetic inferences.  Workload at N=1: BASELINE.json
configs[1], the MNIST LeNet-5-style CNN (MODEL_B_POOL_REPL), batch 256 per
GPU; N>1 shards independent inferences (weak scaling, no data-path
collective).  Prints ONE JSON line on rank 0; see DESIGN.md "Measurement".

The garbled material and the input labels are *inputs* of the path: they are
produced outside the timed region by the reference's own garbler
(oracle/_ref, test infrastructure) with a fixed seed and cached under
$DASHB_CACHE.  The reference arm (--impl reference) and the cpu_baseline leg
time the reference's own CPU evaluator on the same material.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import workloads as W  # noqa: E402

METRIC = "online garbled-eval inferences/s"
CACHE = os.environ.get("DASHB_CACHE", "/tmp/dashb_cache")
SEED = 42


def git_blob_hash(path):
    """What `git hash-object` prints for the file (the .git directory does not travel to the GPU box)."""
    import hashlib

    data = open(path, "rb").read()
    return hashlib.sha1(b"blob %d\0" % len(data) + data).hexdigest()


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def material_paths(model):
    os.makedirs(CACHE, exist_ok=True)
    return os.path.join(CACHE, "%s_s%d.gcm" % (model, SEED)), os.path.join(CACHE, "%s_s%d_in.npz" % (model, SEED))


def build_material(model, n_inputs, want_ref=False):
    """Garbles `model` with the reference (seeded) and encodes n_inputs
    synthetic inputs.  Returns (gcm path, labels [k][n_inputs][n][nc], ref)."""
    from oracle import refbind

    gcm, npz = material_paths(model)
    spec = W.MODELS[model]
    ref = None
    have = os.path.exists(gcm) and os.path.exists(npz)
    if have:
        z = np.load(npz)
        labels = [z["in%d" % j] for j in range(len(spec["crt"]))]
        have = labels[0].shape[0] >= n_inputs
    if not have or want_ref:
        if not refbind.available():
            raise RuntimeError("oracle/_ref/libdash_ref.so missing: build it where /root/reference exists (make -C oracle ref)")
        ref = W.build_reference_circuit(spec, seed=SEED, weight_seed=SEED)
    if not have:
        ref.export(gcm + ".tmp")
        os.replace(gcm + ".tmp", gcm)
        X = W.synthetic_inputs(spec, n_inputs, seed=SEED)
        per = [ref.garble_inputs(x) for x in X]
        labels = [np.stack([p[j] for p in per]) for j in range(ref.k)]
        np.savez(npz + ".tmp.npz", x=X, **{"in%d" % j: labels[j] for j in range(ref.k)})
        os.replace(npz + ".tmp.npz", npz)
    build_material.x = np.load(npz)["x"][:n_inputs]
    return gcm, [l[:n_inputs] for l in labels], ref


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons during the timed region
    (one background `nvidia-smi -lms 50` process)."""

    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.samples = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return
        self.proc.terminate()
        try:
            out, _ = self.proc.communicate(timeout=5)
        except Exception:
            self.proc.kill()
            out = ""
        for line in out.splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) >= 6 and f[0].replace(".", "").isdigit():
                self.samples.append(f)

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unsampled"]}
        sm = sorted(float(s[0]) for s in self.samples)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(s[2 + i] == "Active" for s in self.samples)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(self.samples[0][1]), "reasons": reasons,
                "samples": len(self.samples)}


def run_reference(args, rank):
    """Reference arm: the reference's own CPU evaluator (oracle/_ref) with all
    host threads on the same model; a step is a bounded sample of inferences."""
    if rank != 0:
        return
    threads = host_threads()
    per_step = args.ref_inferences
    n = per_step * (args.steps + args.warmup)
    gcm, labels, ref = build_material(args.model, max(n, 1), want_ref=True)
    times = []
    idx = 0
    for s in range(args.steps + args.warmup):
        t0 = time.perf_counter()
        for _ in range(per_step):
            ref.cpu_evaluate([l[idx % labels[0].shape[0]] for l in labels], threads, want_output=False)
            idx += 1
        if s >= args.warmup:
            times.append(time.perf_counter() - t0)
    total = sum(times)
    value = per_step * args.steps / total
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "inferences/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": "%s garbled online evaluation (BASELINE.json configs[1]), reference CPU evaluator" % args.model,
                   "batch": per_step},
        "cpu_baseline": {"value": value, "unit": "inferences/s", "cores": threads, "kind": "reference",
                         "sample": "%d inferences per step, GarbledCircuitInterface::cpu_evaluate, %d threads" % (per_step, threads)},
        "e2e": {"value": value, "unit": "inferences/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def union_ms(iv):
    """Length of the union of the intervals [(begin, end)]."""
    tot, end = 0.0, None
    for a, b in sorted(iv):
        if end is None or a > end:
            tot += b - a
            end = b
        elif b > end:
            tot += b - end
            end = b
    return tot


def oracle_parity(path, X, got, rows, threads):
    """Output labels of inferences `rows` against oracle/dash_oracle.c (the pinned restatement) on the material file."""
    from dash_b200 import gcm as gcm_io
    from oracle import oraclebind as O

    g = gcm_io.read_gcm(path)
    oc = O.OracleCircuit(path)
    ok = True
    for b in rows:
        want = oc.evaluate(gcm_io.encode_inputs(g, X[b]), threads=threads)
        ok = ok and all(np.array_equal(o[b], w) for o, w in zip(got, want))
    oc.close()
    return bool(ok)


def run_c3(torch, dist, rank, world, local_rank, barrier, steps=5, batch=64):
    """BASELINE.json configs[2]: MODEL_F_GNNP_POOL_REPL (CIFAR-10 CNN, Rescale{32} after every linear layer),
    material garbled on every rank's own GPU, `batch` inferences per GPU, batch-sharded (no data-path collective)."""
    from dash_b200 import garble

    model = "MODEL_F_GNNP_POOL_REPL"
    spec = W.MODELS[model]
    plain, n_in = W.plain_layers(spec)
    X = W.synthetic_inputs(spec, batch, seed=100 + rank)
    seed = b"bench-c3-gnnp"
    c, gst = garble.garble_to_circuit(spec["crt"], spec["mrs"], n_in, plain, batch, seed=seed)
    c.garble_inputs(X)
    stream = torch.cuda.ExternalStream(__import__("dash_b200").api.stream_handle())
    for _ in range(2):
        c.evaluate_device()
    torch.cuda.synchronize()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(steps):
        c.evaluate_device()
    e1.record(stream)
    e1.synchronize()
    ms = e0.elapsed_time(e1) / steps
    if world > 1:
        t = torch.tensor([ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    # share of the scaling gadget in the step (events on the lane streams inside the concurrent step)
    c.set_layer_timing(2)
    c.evaluate_device()
    iv = {}
    last = 0.0
    for name, _li, _lane, a, b in c.layer_intervals_ms():
        iv.setdefault(name, []).append((a, b))
        last = max(last, b)
    c.set_layer_timing(0)
    got = c.move_outputs()
    vals, ok = c.decode_outputs()
    out = {"model": model, "config": "BASELINE.json configs[2]", "batch_per_gpu": batch, "n_gpus": world, "scaling": "weak",
           "ms_per_step": ms, "inferences_per_s": batch * world / (ms / 1e3), "all_outputs_decode": bool(ok.all()),
           "k_rescale_share_of_step": union_ms(iv.get("rescale", [])) / last if last else None,
           "busy_ms_by_layer_type": {n: union_ms(v) for n, v in iv.items()},
           "material": "GPU garbler, %.0f ms of kernels, %.2f GB of tables per GPU" % (gst["kernel_ms"], gst["table_bytes"] / 1e9)}
    c.close()
    if rank == 0:
        os.makedirs(CACHE, exist_ok=True)
        path = os.path.join(CACHE, "c3_gnnp.gcm")
        garble.garble_to_file(spec["crt"], spec["mrs"], n_in, plain, path, seed=seed)
        out["bit_exact_vs_oracle"] = oracle_parity(path, X, got, (0, batch - 1), host_threads())
        out["parity_inferences_checked"] = 2
        os.remove(path)
    return out


def run_c4(torch, dist, rank, world, local_rank, barrier, steps=5, batch=64):
    """BASELINE.json configs[3]: MODEL_F_MINIONN_POOL_REPL, ONE global batch, output-channel sharding across the
    ranks (strong scaling): every rank evaluates its block of each layer's output channels, the activation slices
    are exchanged over NVLink after every layer group."""
    from dash_b200 import garble, gcm as gcm_io
    from dash_b200.scheduler import ChannelShardedEvaluator

    model = "MODEL_F_MINIONN_POOL_REPL"
    spec = W.MODELS[model]
    plain, n_in = W.plain_layers(spec)
    X = W.synthetic_inputs(spec, batch, seed=100)
    seed = b"bench-c4-minionn"
    mode = os.environ.get("DASHB_C4_EXCHANGE", "p2p")
    if world == 1:  # the strong-scaling baseline: the same batch on one GPU, nothing to exchange
        c, gst = garble.garble_to_circuit(spec["crt"], spec["mrs"], n_in, plain, batch, seed=seed)
        c.evaluate_layers = lambda first, last: c.evaluate_device() if first == 0 else None
    else:
        c, gst = garble.garble_to_circuit(spec["crt"], spec["mrs"], n_in, plain, batch, seed=seed, shard=(rank, world),
                                          shard_lanes=int(os.environ.get("DASHB_C4_LANES", "2")) if mode == "p2p" else 0)
    ev = ChannelShardedEvaluator(c, rank, world, dist=dist if world > 1 else None, p2p=(mode == "p2p"))
    ev.load_values(X)
    ev.run()  # warm-up (peer mappings / NCCL channels, lazy module loading)
    ev.run()
    torch.cuda.synchronize()
    barrier()
    ev.bytes_exchanged = 0
    t0 = time.perf_counter()
    for _ in range(steps):
        ev.run()
    torch.cuda.synchronize()
    ms = (time.perf_counter() - t0) * 1e3 / steps
    if world > 1:
        t = torch.tensor([ms], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    trace = None
    if world > 1 and mode == "p2p" and os.environ.get("DASHB_C4_TRACE"):
        trace = ev.trace()
    got = c.move_outputs()
    out = {"model": model, "config": "BASELINE.json configs[3]", "global_batch": batch, "n_gpus": world, "scaling": "strong",
           "ms_per_step": ms, "inferences_per_s": batch / (ms / 1e3), "exchange": ev.describe(),
           "nvlink_bytes_received_per_rank_per_step": ev.bytes_exchanged // steps,
           "timing": "wall clock around %d steps between device synchronisations, max over ranks" % steps}
    if trace is not None:
        out["trace_ms_rank0_[group][lane](layers_done, exchange_done)"] = trace
    c.close()
    if rank == 0:
        os.makedirs(CACHE, exist_ok=True)
        path = os.path.join(CACHE, "c4_minionn.gcm")
        garble.garble_to_file(spec["crt"], spec["mrs"], n_in, plain, path, seed=seed)
        out["bit_exact_vs_oracle"] = oracle_parity(path, X, got, (0,), host_threads())
        out["parity_inferences_checked"] = 1
        os.remove(path)
    return out


def bind_to_gpu_numa_node(local_rank):
    """Multi-rank runs: keep this process (and the pinned host buffers it first-touches) on the CPUs of the NUMA node
    the GPU hangs off (sysfs local_cpulist of the GPU's PCI function), so that the step's host->device traffic does not
    cross the socket interconnect.  Returns a description for the JSON line, None when nothing was done."""
    try:
        import pynvml

        pynvml.nvmlInit()
        bus = pynvml.nvmlDeviceGetPciInfo(pynvml.nvmlDeviceGetHandleByIndex(local_rank)).busId
        bus = bus.decode() if isinstance(bus, bytes) else bus
        dom, rest = bus.split(":", 1)
        path = "/sys/bus/pci/devices/%s:%s/local_cpulist" % (dom[-4:].lower(), rest.lower())
        cpus = set()
        for part in open(path).read().strip().split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return "%d cpus local to %s" % (len(cpus), bus)
    except Exception:  # no NVML, no sysfs entry, a container without the right: run unbound
        return None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--model", default="MODEL_B_POOL_REPL")
    ap.add_argument("--batch", type=int, default=256, help="inferences per GPU per step")
    ap.add_argument("--ref-inferences", type=int, default=2, help="reference arm: inferences per step")
    ap.add_argument("--cpu-seconds", type=float, default=15.0, help="cpu_baseline leg budget")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the c3_batch_sharded / c4_channel_sharded legs")
    ap.add_argument("--only", default="", choices=["", "c3", "c4"], help="run one extra leg alone and print its JSON (experiments)")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank)
        return
    if args.warmup < 3:
        args.warmup = 3
    numa = bind_to_gpu_numa_node(local_rank) if world > 1 else None

    import torch
    import torch.distributed as dist

    from dash_b200 import api

    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        if world > 1:
            dist.barrier()

    if args.only:
        api.init(local_rank)
        res = (run_c3 if args.only == "c3" else run_c4)(torch, dist, rank, world, local_rank, barrier)
        if rank == 0:
            print(json.dumps(res), flush=True)
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- inputs of the path (outside the timed region) -------------------
    B = args.batch
    if rank == 0:
        gcm, labels, _ = build_material(args.model, B)
    barrier()
    if rank != 0:
        gcm, labels, _ = build_material(args.model, B)
    api.init(local_rank)
    circ = api.Circuit.load(gcm, max_batch=B)
    stream = torch.cuda.ExternalStream(api.stream_handle())

    # pinned host buffers: what a caller hands to the C ABI
    pinned_in = []
    for l in labels:
        t = torch.empty(l.shape, dtype=torch.int16).pin_memory()
        t.numpy()[...] = l
        pinned_in.append(t)
    in_np = [t.numpy() for t in pinned_in]
    h2d = sum(l.nbytes for l in in_np)
    out_t = [torch.empty((B, circ.output_size, api.nr_comps(q)), dtype=torch.int16).pin_memory() for q in circ.out_moduli]
    out_np = [t.numpy() for t in out_t]
    d2h = sum(o.nbytes for o in out_np)
    in_ptrs = api._ptr_array(in_np)
    out_ptrs = api._ptr_array(out_np)
    L = circ.L

    circ.move_inputs(in_np)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    def step_device():
        api._check(L.dashb_evaluate_device(circ.h))

    def step_e2e():
        api._check(L.dashb_evaluate(circ.h, in_ptrs, B, out_ptrs))

    # the streaming form of the same call: dashb_evaluate_async keeps two steps in flight (two sets of lanes), so
    # step n+1's labels cross PCIe while step n computes; outputs alternate between two host buffer sets
    out_t2 = [torch.empty_like(t).pin_memory() for t in out_t]
    out_sets = [out_ptrs, api._ptr_array([t.numpy() for t in out_t2])]
    L.dashb_evaluate_async.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p]
    L.dashb_evaluate_wait.argtypes = [C.c_void_p]

    def timed_pipelined(steps, warmup, call=None, ins=None, outs=None):
        call = call or L.dashb_evaluate_async
        ins = ins if ins is not None else in_ptrs
        outs = outs or out_sets
        for i in range(warmup):
            api._check(call(circ.h, ins, B, outs[i & 1]))
        api._check(L.dashb_evaluate_wait(circ.h))
        torch.cuda.synchronize()
        barrier()
        flush.zero_()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for i in range(steps):
            api._check(call(circ.h, ins, B, outs[i & 1]))
        api._check(L.dashb_evaluate_wait(circ.h))  # host waits for the last step; e1 then closes the region
        e1.record(stream)
        e1.synchronize()
        ms = e0.elapsed_time(e1)
        torch.cuda.synchronize()
        barrier()
        if world > 1:
            t = torch.tensor([ms], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    def timed(fn, steps, warmup):
        for _ in range(warmup):
            fn()
        torch.cuda.synchronize()
        barrier()
        torch.cuda.synchronize()
        ms = 0.0
        for _ in range(steps):
            flush.zero_()  # L2 flush between timed iterations (not timed)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            fn()
            e1.record(stream)
            e1.synchronize()
            ms += e0.elapsed_time(e1)
        torch.cuda.synchronize()
        barrier()
        if world > 1:
            t = torch.tensor([ms], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    # plaintext values in -> decoded values out (garble_inputs + evaluate +
    # decode_outputs all on the device): the reference user's whole call chain
    xs = torch.empty((B, circ.input_size), dtype=torch.int64).pin_memory()
    xs.numpy()[...] = build_material.x[:B]
    vals_t = torch.empty((B, circ.output_size), dtype=torch.int64).pin_memory()
    ok_t = torch.empty((B,), dtype=torch.uint8).pin_memory()

    def step_values():
        api._check(L.dashb_garble_inputs(circ.h, xs.data_ptr(), B))
        api._check(L.dashb_evaluate_device(circ.h))
        api._check(L.dashb_decode_outputs(circ.h, vals_t.data_ptr(), ok_t.data_ptr()))

    # the same call with compressed 128-bit labels on the host side
    # (LabelTensor::compress form, dashb_evaluate_compressed)
    from dash_b200 import gcm as gcm_io

    crt_list = gcm_io.read_gcm(gcm)["crt"]

    comp_in = [torch.from_numpy(gcm_io.compress_labels(q, l)).pin_memory() for q, l in zip(crt_list, in_np)]
    comp_out = [torch.empty((B, circ.output_size, 2), dtype=torch.uint64).pin_memory() for _ in range(circ.k)]
    comp_in_ptrs = (C.c_void_p * circ.k)(*[t.data_ptr() for t in comp_in])
    comp_out_ptrs = (C.c_void_p * circ.k)(*[t.data_ptr() for t in comp_out])
    L.dashb_evaluate_compressed.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p]

    def step_compressed():
        api._check(L.dashb_evaluate_compressed(circ.h, comp_in_ptrs, B, comp_out_ptrs))

    comp_out2 = [torch.empty_like(t).pin_memory() for t in comp_out]
    comp_out_sets = [comp_out_ptrs, (C.c_void_p * circ.k)(*[t.data_ptr() for t in comp_out2])]
    L.dashb_evaluate_compressed_async.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p]

    sampler = ClockSampler(local_rank)
    sampler.start()
    time.sleep(0.2)
    ms_dev = timed(step_device, args.steps, args.warmup)
    launches = circ.launch_count()
    ms_e2e_sync = timed(step_e2e, args.steps, args.warmup)
    ms_e2e = timed_pipelined(args.steps, args.warmup)
    pipelined_equal = all(bool(torch.equal(a, b)) for a, b in zip(out_t, out_t2))
    ms_val = timed(step_values, args.steps, args.warmup)
    ms_comp = timed(step_compressed, args.steps, args.warmup)
    ms_comp_pipe = timed_pipelined(args.steps, args.warmup, L.dashb_evaluate_compressed_async, comp_in_ptrs, comp_out_sets)
    comp_pipelined_equal = all(bool(torch.equal(a, b)) for a, b in zip(comp_out, comp_out2))
    values_ok = bool(ok_t.numpy().all())
    circ.move_inputs(in_np)  # restore the label inputs for the profiling pass
    # Per-kernel durations for the roofline, taken in the step as it runs (sub-batch lanes concurrent on their own
    # streams): an event on the lane's stream before and after every layer (dashb_set_layer_timing(2)).  A layer
    # type's busy time is the union of its (lane, layer) intervals; it can never exceed the step it is part of.
    circ.set_layer_timing(2)
    prof_steps = min(args.steps, 20)
    busy = {}       # layer type -> union of its intervals, ms summed over the profiled steps
    launch_ms = {}  # layer type -> [duration of every (lane, layer) launch]
    prof_step_ms = 0.0

    for it in range(3 + prof_steps):
        flush.zero_()
        torch.cuda.synchronize()
        step_device()
        if it < 3:
            continue
        by = {}
        last = 0.0
        for name, _li, _lane, a, b in circ.layer_intervals_ms():
            by.setdefault(name, []).append((a, b))
            launch_ms.setdefault(name, []).append(b - a)
            last = max(last, b)
        prof_step_ms += last
        for name, iv in by.items():
            busy[name] = busy.get(name, 0.0) + union_ms(iv)
    circ.set_layer_timing(0)
    time.sleep(0.1)
    sampler.stop()
    busy = {n: v / prof_steps for n, v in busy.items()}
    prof_step_ms /= prof_steps

    total_inf = B * world * args.steps
    value = total_inf / (ms_dev / 1e3)
    e2e_value = total_inf / (ms_e2e / 1e3)
    ms_step = ms_dev / args.steps

    # ---- roofline of the dominant kernel (the fused sign/ReLU gadget) -----
    peaks_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(peaks_path):
        peak, peak_src = json.load(open(peaks_path))["hbm_gbs"], "measured (MEASURED_PEAKS.json)"
    else:
        peak, peak_src = 6650.0, "fallback (B200_PROFILING.md)"
    relu_ms = busy.get("approx_relu", 0.0)
    if relu_ms > 1.05 * max(ms_step, prof_step_ms):
        raise RuntimeError("roofline: ReLU busy time %.3f ms exceeds the step it is part of (%.3f ms)" % (relu_ms, ms_step))
    relu_layers = [i for i in range(circ.n_layers) if circ.layer_type(i) == 3]
    relu_bytes = sum(circ.algorithmic_bytes(i) for i in relu_layers) * B
    relu_aes = sum(circ.aes_blocks(i) for i in relu_layers) * B
    relu_units = sum(circ.layer_size(i) for i in relu_layers) * B  # (element, inference) pairs per step
    achieved = relu_bytes / (relu_ms / 1e3) / 1e9 if relu_ms > 0 else 0.0
    n_lanes = circ.L.dashb_lanes_used(circ.h)
    n_launch = max(1, len(relu_layers) * n_lanes)
    relu_launch = launch_ms.get("approx_relu", [0.0])
    # DRAM traffic / issue statistics of the same kernel: one `ncu --set full` capture, valid only for the
    # kernels.cu it was taken from (git blob hash stored with it; tools/make_traffic.py writes the file)
    traffic, capture = None, None
    tpath = os.path.join(ROOT, "profiles", "r2_traffic.json")
    if os.path.exists(tpath) and args.model == "MODEL_B_POOL_REPL" and B == 256:
        tj = json.load(open(tpath))
        if tj.get("kernels_cu_blob") == git_blob_hash(os.path.join(ROOT, "dash_b200", "csrc", "kernels.cu")):
            capture = tj["k_sign_relu"]
            # per launch like `achieved`: the captured launch's DRAM bytes per (element, inference) times the
            # average launch's (element, inference) count
            traffic = capture["dram_bytes_per_element_inference"] * relu_units / n_launch
        else:
            capture = {"stale": "profiles/r2_traffic.json was captured from another kernels.cu; traffic withheld"}
    aes_peaks = json.load(open(os.path.join(ROOT, "profiles", "aes_peaks.json")))
    aes_rate = relu_aes / (relu_ms / 1e3) if relu_ms > 0 else 0.0
    roofline = {
        "bound": "hbm", "kernel": "k_sign<relu> (fused sign gadget + half gates)", "achieved": achieved, "peak": peak,
        "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
        "algorithmic_bytes_per_launch": relu_bytes / n_launch,
        "launches_per_step": n_launch,
        "element_inferences_per_launch": relu_units / n_launch,
        "busy_ms_per_step": relu_ms,
        "avg_launch_ms": sum(relu_launch) / len(relu_launch),
        "min_max_launch_ms": [min(relu_launch), max(relu_launch)],
        "timing": "CUDA events on the lane streams around every layer inside the concurrent step (%d sub-batch lanes), "
                  "%d steps after the timed region; busy time = union of the lanes' intervals" % (n_lanes, prof_steps),
        "step_ms_in_timing_pass": prof_step_ms,
        "share_of_step": relu_ms / prof_step_ms if prof_step_ms else None,
        "aes_blocks_per_s": aes_rate,
        "aes_blocks_per_s_in_step": circ.aes_blocks(-1) * B * args.steps / (ms_dev / 1e3),
        "aes_peak_4kb_tables": aes_peaks["aes_peak_4kb_tables"],
        "aes_peak_conflict_free": aes_peaks["aes_peak_conflict_free"],
        "aes_frac_of_4kb_tables_peak": aes_rate / aes_peaks["aes_peak_4kb_tables"]["blocks_per_s"],
        "aes_frac_of_conflict_free_peak": aes_rate / aes_peaks["aes_peak_conflict_free"]["blocks_per_s"],
        "ncu_capture": capture,
        "issue_slot_frac": capture.get("issue_slot_frac") if capture else None,
        "warp_instructions_per_element": capture.get("warp_instructions_per_element") if capture else None,
        "note": "issue-bound on AES T-table lookups + base-q digit arithmetic, not on HBM (see DESIGN.md)",
        "whole_step_algorithmic_GBps": circ.algorithmic_bytes(-1) * B / ms_step / 1e6,
        "busy_ms_by_layer_type": busy,
    }

    line = {
        "metric": METRIC, "value": value, "unit": "inferences/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_dev / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": "%s garbled online evaluation (BASELINE.json configs[1]): conv2d + ReLU projections + dense, k=%d CRT moduli" % (args.model, circ.k),
                   "batch_per_gpu": B, "global_batch": B * world, "parallelism": "batch-sharded x%d" % world,
                   "cache": "L2 flushed (256 MiB write) between timed iterations; per-step working set also exceeds L2",
                   "material": "reference garbler, seed %d, tables shared by the batch" % SEED},
        "e2e": {"value": e2e_value, "unit": "inferences/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": ms_e2e / args.steps,
                "api": "dashb_evaluate_async x steps + dashb_evaluate_wait: int16 LabelTensor layout in pinned host memory -> output "
                       "labels in host memory, two steps in flight (step n+1's labels cross PCIe while step n computes); "
                       "every step moves its own inputs and outputs inside the timed region",
                "cache": "no flush between pipelined steps: a step's working set (94 MB of digit planes + 111 MB of tables + "
                         "activations) exceeds the 126 MB L2",
                "outputs_equal_across_steps": pipelined_equal},
        "e2e_sync": {"value": total_inf / (ms_e2e_sync / 1e3), "unit": "inferences/s", "ms_per_step": ms_e2e_sync / args.steps,
                     "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                     "api": "dashb_evaluate, one synchronous call per step (the reference's cuda_move_inputs + cuda_evaluate + "
                            "cuda_move_outputs), L2 flushed between steps"},
        "e2e_values": {"value": total_inf / (ms_val / 1e3), "unit": "inferences/s", "ms_per_step": ms_val / args.steps,
                       "h2d_bytes_per_step": int(xs.numel() * 8), "d2h_bytes_per_step": int(vals_t.numel() * 8 + B),
                       "all_outputs_decoded": values_ok,
                       "api": "dashb_garble_inputs + dashb_evaluate_device + dashb_decode_outputs (int64 values in, int64 values out)"},
        "e2e_compressed": {"value": total_inf / (ms_comp / 1e3), "unit": "inferences/s", "ms_per_step": ms_comp / args.steps,
                           "h2d_bytes_per_step": int(sum(t.numel() * 8 for t in comp_in)),
                           "d2h_bytes_per_step": int(sum(t.numel() * 8 for t in comp_out)),
                           "api": "dashb_evaluate_compressed (labels as 16-byte LabelTensor::compress values in pinned host memory)"},
        "e2e_compressed_pipelined": {"value": total_inf / (ms_comp_pipe / 1e3), "unit": "inferences/s",
                                     "ms_per_step": ms_comp_pipe / args.steps,
                                     "h2d_bytes_per_step": int(sum(t.numel() * 8 for t in comp_in)),
                                     "d2h_bytes_per_step": int(sum(t.numel() * 8 for t in comp_out)),
                                     "both_output_sets_equal": comp_pipelined_equal,
                                     "api": "dashb_evaluate_compressed_async x steps + dashb_evaluate_wait: the compressed wire "
                                            "form with two steps in flight (13 % of the int16 layout's PCIe bytes)"},
        "gpu_launches": int(launches * args.steps),
        "clocks": sampler.summary(),
        "host_affinity": numa,
        "roofline": roofline,
    }

    # ---- cpu_baseline + parity on a bounded sample (rank 0, N=1) ----------
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        from oracle import refbind

        if refbind.available():
            _, _, ref = build_material(args.model, B, want_ref=True)
            threads = host_threads()
            gpu_out = [o.copy() for o in out_np]
            t_used, n_done, equal = 0.0, 0, True
            ref.cpu_evaluate([l[0] for l in labels], threads)  # warm-up (table init)
            while n_done < B and (n_done < 2 or t_used < args.cpu_seconds):
                t0 = time.perf_counter()
                ro, _ = ref.cpu_evaluate([l[n_done] for l in labels], threads)
                t_used += time.perf_counter() - t0
                equal = equal and all(np.array_equal(ro[j], gpu_out[j][n_done]) for j in range(ref.k))
                n_done += 1
            line["cpu_baseline"] = {"value": n_done / t_used, "unit": "inferences/s", "cores": threads, "kind": "reference",
                                    "sample": "%d inferences of the same batch, GarbledCircuitInterface::cpu_evaluate (oracle/_ref), %d threads" % (n_done, threads)}
            line["parity"] = {"inferences_checked": n_done, "bit_exact_output_labels": bool(equal)}
        else:
            line["cpu_baseline"] = {"value": None, "unit": "inferences/s", "cores": 0, "kind": "reference", "sample": "oracle/_ref not built"}
    # ---- offline phase of the same model on the GPU (extra key; rank 0, N=1) ----
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            from dash_b200 import garble as gb
            gref = gcm_io.read_gcm(gcm)
            plain = gcm_io.plain_circuit(gref)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            gc2, gst = gb.garble_to_circuit(gref["crt"], gref["mrs"], gref["input_size"], plain, B, max_modulus=gref["max_modulus"])
            torch.cuda.synchronize()
            ready_s = time.perf_counter() - t0
            gc2.garble_inputs(np.asarray(build_material.x, dtype=np.int64).reshape(B, -1))
            gc2.evaluate_device()
            v2, ok2 = gc2.decode_outputs()
            gc2.close()
            line["offline_phase"] = {
                "what": "the same plaintext circuit garbled on the GPU (dashb_garble_to_circuit), then evaluated and decoded",
                "kernel_ms": gst["kernel_ms"], "ready_circuit_s": ready_s, "gates": gst["gates"], "ciphertext_rows": gst["rows"],
                "rows_per_s_in_kernels": gst["rows"] / (gst["kernel_ms"] / 1e3) if gst["kernel_ms"] else None,
                "decodes_to_the_reference_garbled_circuits_values": bool(ok2.all() and values_ok and np.array_equal(v2, vals_t.numpy())),
            }
        except Exception as ex:  # the extra key must never break the contract line
            line["offline_phase"] = {"error": str(ex)[:200]}
    # ---- one set of ciphertext tables per inference (extra key; rank 0, N=1) ----
    # The headline batch shares ONE garbled circuit (test / benchmark use, see include/dash_b200.h); a deployment
    # garbles once per inference.  Same kernels, same labels, but every sign / ReLU table stored B times with the
    # rows interleaved by inference (dashb_set_table_replicas): each inference streams its own rows from HBM.
    if rank == 0 and world == 1 and not args.no_extras:
        try:
            want_out = [o.copy() for o in out_np]
            circ_r = api.Circuit.load(gcm, max_batch=B, table_replicas=B)
            circ_r.move_inputs(in_np)

            def step_rep():
                api._check(L.dashb_evaluate_device(circ_r.h))

            ms_rep = timed(step_rep, min(args.steps, 20), args.warmup)
            got_r = circ_r.move_outputs()
            line["per_inference_tables"] = {
                "value": B * min(args.steps, 20) / (ms_rep / 1e3), "unit": "inferences/s",
                "ms_per_step": ms_rep / min(args.steps, 20), "replicas": B,
                "table_bytes_on_device": int(sum(circ_r.aes_blocks(i) for i in range(circ_r.n_layers)) * 0) or None,
                "same_output_labels_as_shared_tables": bool(all(np.array_equal(a, b) for a, b in zip(got_r, want_out))),
                "what": "dashb_evaluate_device with every sign/ReLU table stored once per inference of the batch, rows "
                        "interleaved by inference (the layout B independent garblings give a batch); labels and "
                        "linear-layer constants shared, so the outputs stay checkable",
            }
            line["per_inference_tables"].pop("table_bytes_on_device")
            circ_r.close()
            del circ_r
        except Exception as ex:
            line["per_inference_tables"] = {"error": "%s: %s" % (type(ex).__name__, str(ex)[:300])}
    # ---- BASELINE.json configs[2] and configs[3] on the same run (extra keys; every rank takes part) ----
    circ.close()
    del flush
    torch.cuda.empty_cache()
    if not args.no_extras:
        for key, fn in (("c3_batch_sharded", run_c3), ("c4_channel_sharded", run_c4)):
            t0 = time.perf_counter()
            try:
                res = fn(torch, dist, rank, world, local_rank, barrier)
                res["leg_seconds"] = time.perf_counter() - t0
            except Exception as ex:  # an extra key must never break the contract line
                res = {"error": "%s: %s" % (type(ex).__name__, str(ex)[:300])}
            line[key] = res
            barrier()
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
