"""Batch scheduler: shards independent inferences across the GPUs of one box.

One process per GPU (torch.distributed, NCCL over NVLink on the GPU box,
gloo in the CPU tests).  Inferences are independent, weights and ciphertext
tables are replicated once per GPU, so evaluation needs no collective at
all; the only exchange is the final gather of the output labels
(k * output_size * nc int16 per inference).  The reference has no
counterpart: it evaluates one inference per call on one GPU
(garbling/garbled_circuit_interface.h:603-612).
"""
import numpy as np


def shard_range(total, world_size, rank):
    """[begin, end) of the inferences rank `rank` evaluates: contiguous,
    sizes differing by at most one, empty shards allowed."""
    if world_size < 1 or not (0 <= rank < world_size):
        raise ValueError("bad world_size/rank")
    base, extra = divmod(total, world_size)
    begin = rank * base + min(rank, extra)
    return begin, begin + base + (1 if rank < extra else 0)


class BatchShardedEvaluator:
    """Evaluates a global batch: every rank runs `evaluate_fn` on its shard
    and the output labels of all ranks are gathered on every rank."""

    def __init__(self, evaluate_fn, out_shapes, device=None):
        """evaluate_fn(list of k int16 arrays [b][n_in][nc]) -> list of k
        int16 arrays [b][n_out][nc_out]; out_shapes = [(n_out, nc_out_j)]."""
        import torch.distributed as dist

        self.dist = dist
        self.evaluate_fn = evaluate_fn
        self.out_shapes = out_shapes
        self.world = dist.get_world_size() if dist.is_initialized() else 1
        self.rank = dist.get_rank() if dist.is_initialized() else 0
        self.device = device

    def evaluate(self, labels):
        import torch

        total = labels[0].shape[0]
        b, e = shard_range(total, self.world, self.rank)
        local = self.evaluate_fn([l[b:e] for l in labels]) if e > b else [
            np.zeros((0, n, nc), dtype=np.int16) for (n, nc) in self.out_shapes]
        if self.world == 1:
            return local
        out = []
        for j, (n, nc) in enumerate(self.out_shapes):
            sizes = [shard_range(total, self.world, r) for r in range(self.world)]
            mx = max(hi - lo for lo, hi in sizes)
            pad = np.zeros((mx, n, nc), dtype=np.int16)
            pad[: e - b] = local[j]
            # gathered as raw bytes: int16 is not a collective dtype on every backend
            t = torch.from_numpy(pad.view(np.uint8).reshape(-1))
            if self.device is not None:
                t = t.to(self.device)
            parts = [torch.empty_like(t) for _ in range(self.world)]
            self.dist.all_gather(parts, t)
            out.append(np.concatenate(
                [p.cpu().numpy().view(np.int16).reshape(mx, n, nc)[: hi - lo] for p, (lo, hi) in zip(parts, sizes)],
                axis=0))
        return out


LINEAR_TYPES = (1, 2)  # DASHB_LAYER_DENSE, DASHB_LAYER_CONV2D


def layer_groups(layer_types):
    """Cuts a layer list into groups that can run on a rank's slice without
    communication: a linear layer (dense / conv2d) needs every input element,
    so each one starts a new group; the element-wise gadget layers after it
    stay with it.  Returns [(first, last)]."""
    groups = []
    start = 0
    for i, t in enumerate(layer_types):
        if t in LINEAR_TYPES and i > start:
            groups.append((start, i - 1))
            start = i
    if layer_types:
        groups.append((start, len(layer_types) - 1))
    return groups


class _DeviceBytes:
    """Wraps raw device memory for torch.as_tensor (CUDA array interface)."""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (int(nbytes),), "typestr": "|u1", "data": (int(ptr), False), "version": 2}


class ChannelShardedEvaluator:
    """Output-channel sharding of ONE batch across the GPUs of a box (config
    C4): every rank owns a block of each linear layer's output rows and the
    same elements of the gadget layers behind it; after each group of layers
    the ranks' slices of the activation planes are exchanged.  Weights and
    tables are replicated.

    Two exchanges.  p2p=True (the circuit was prepared with shard_lanes > 0):
    the library's own all-gather, plain stores into every peer's gather
    buffer over NVLink (CUDA IPC mappings exchanged once through
    torch.distributed), per sub-batch lane, so that one lane's exchange
    overlaps the other lane's layer group -- no NCCL in the data path.
    Otherwise: NCCL all-gathers / broadcasts straight on the device planes
    (one per CRT modulus), one lane."""

    def __init__(self, circuit, rank, world, dist=None, as_tensor=None, compressed=False, p2p=False):
        self.c = circuit
        self.rank = rank
        self.world = world
        self.dist = dist
        self.as_tensor = as_tensor
        self.p2p = bool(p2p) and world > 1
        if self.p2p:
            if dist is None:
                raise ValueError("the peer-memory exchange needs torch.distributed to hand the IPC handles round")
            n_lanes = circuit.L.dashb_shard_lanes_used(circuit.h) or 1
            # lanes exist from dashb_finalize on; export every lane the circuit has
            handles = []
            lane = 0
            while True:
                try:
                    handles.append(circuit.shard_export(lane))
                except Exception:
                    break
                lane += 1
            everyone = [None] * world
            dist.all_gather_object(everyone, handles)
            for r in range(world):
                if r != rank:
                    for lane, h in enumerate(everyone[r]):
                        circuit.shard_import(lane, r, h)
            dist.barrier()
        self.groups = layer_groups([circuit.layer_type(i) for i in range(circuit.n_layers)])
        self.bytes_exchanged = 0
        self._ext = None
        # exchange 16-byte compressed labels instead of digit planes: 2.9x fewer NVLink bytes, but measured
        # slower on NVSwitch (pack/unpack cost more than the transfer saves, profiles/r1_multi_gpu_cifar.jsonl)
        self.compressed = compressed
        self._wire = {}

    def _gather(self, layer):
        """Exchanges the ranks' slices of the planes the layer group just
        produced.  The collectives are enqueued behind the library's stream
        (no host synchronisation): NCCL orders its work after the current
        torch stream, which is set to the library stream here.  Equal slices
        in rank order go as ONE in-place all-gather per CRT modulus; uneven
        ones (a last dense layer with 10 rows on 8 ranks) as broadcasts."""
        import torch

        if self.world == 1 or self.dist is None:
            return
        if self._ext is None and self.as_tensor is None:
            from . import api

            self._ext = torch.cuda.ExternalStream(api.stream_handle())
        slices = [self.c.layer_slice(layer, r) for r in range(self.world)]
        sizes = {e1 - e0 for e0, e1 in slices}
        even = len(sizes) == 1 and slices[0][0] == 0 and all(slices[r][0] == slices[r - 1][1] for r in range(1, self.world))

        def exchange_compressed():
            """even slices: every rank packs its slice to 16-byte labels,
            one all-gather per modulus, everybody unpacks the other slices"""
            n_slice = slices[0][1] - slices[0][0]
            per = n_slice * self.c.batch * 16
            for j in range(self.c.k):
                key = (j, per)
                if key not in self._wire:
                    self._wire[key] = torch.empty(per * self.world, dtype=torch.uint8, device="cuda")
                buf = self._wire[key]
                e0, e1 = slices[self.rank]
                self.c.slice_compress(j, e0, e1, buf.data_ptr() + self.rank * per)
                self.dist.all_gather_into_tensor(buf, buf[self.rank * per:(self.rank + 1) * per])
                for r in range(self.world):
                    if r != self.rank:
                        self.c.slice_decompress(j, slices[r][0], slices[r][1], buf.data_ptr() + r * per)
                self.bytes_exchanged += per * (self.world - 1)

        def exchange():
            if even and self.compressed and self.as_tensor is None and hasattr(self.dist, "all_gather_into_tensor"):
                return exchange_compressed()
            for j in range(self.c.k):
                ptr, bpe, n = self.c.current_plane(j)
                if self.as_tensor is not None:
                    plane = self.as_tensor(ptr, bpe * n)
                else:
                    plane = torch.as_tensor(_DeviceBytes(ptr, bpe * n), device="cuda")
                if even and hasattr(self.dist, "all_gather_into_tensor") and self.as_tensor is None:
                    e0, e1 = slices[self.rank]
                    self.dist.all_gather_into_tensor(plane[:slices[-1][1] * bpe], plane[e0 * bpe:e1 * bpe])
                    self.bytes_exchanged += (slices[-1][1] - (e1 - e0)) * bpe
                    continue
                for r in range(self.world):
                    e0, e1 = slices[r]
                    if e1 > e0:
                        self.dist.broadcast(plane[e0 * bpe:e1 * bpe], src=r)
                        if r != self.rank:
                            self.bytes_exchanged += (e1 - e0) * bpe

        if self._ext is not None:
            with torch.cuda.stream(self._ext):
                exchange()
        else:  # emulated ranks / CPU tests: the caller's tensors, synchronous
            self.c.sync()
            exchange()

    def describe(self):
        if self.world == 1:
            return "none (one rank)"
        if self.p2p:
            return ("stores into peer memory over NVLink (k_shard_push + epoch flags), %d sub-batch lanes, "
                    "one exchange per layer group and lane" % self.c.lanes_used())
        return "compressed labels over NCCL" if self.compressed else "NCCL all-gather of the digit planes, one per CRT modulus and layer group"

    def load_labels(self, labels):
        self.c.move_inputs(labels)

    def load_values(self, values):
        """Plaintext inputs, encoded on the device (needs the input encoding in the material)."""
        self.c.garble_inputs(values)

    def trace(self):
        """One run with CUDA events on the lane streams: per layer group and
        lane (ms since the start) when its layers finished and when its
        exchange finished.  p2p mode only."""
        import torch

        lanes = list(range(self.c.lanes_used()))
        streams = [torch.cuda.ExternalStream(self.c.L.dashb_lane_stream(self.c.h, l)) for l in lanes]
        lib_stream = torch.cuda.ExternalStream(self.c.L.dashb_stream())
        t0 = torch.cuda.Event(enable_timing=True)
        t0.record(lib_stream)
        ev = []
        for first, last in self.groups:
            row = []
            for lane in lanes:
                self.c.evaluate_layers_lane(lane, first, last)
                a = torch.cuda.Event(enable_timing=True)
                a.record(streams[lane])
                row.append([a, None])
            for lane in lanes:
                self.c.shard_exchange(lane, last)
                b = torch.cuda.Event(enable_timing=True)
                b.record(streams[lane])
                row[lane][1] = b
            ev.append(row)
        torch.cuda.synchronize()
        return [[(t0.elapsed_time(a), t0.elapsed_time(b)) for a, b in row] for row in ev]

    def run(self):
        """All layer groups on the inputs loaded last, exchanges included; asynchronous."""
        if self.p2p:
            lanes = range(self.c.lanes_used())
            for first, last in self.groups:
                for lane in lanes:
                    self.c.evaluate_layers_lane(lane, first, last)
                for lane in lanes:
                    self.c.shard_exchange(lane, last)
                    self.bytes_exchanged += self.c.shard_exchange_bytes(lane, last)
            return
        for first, last in self.groups:
            self.c.evaluate_layers(first, last)
            if self.world > 1:
                self._gather(last)

    def evaluate(self, labels):
        """labels: the whole batch on every rank.  Returns the output labels
        (identical on every rank)."""
        self.load_labels(labels)
        self.run()
        out = self.c.move_outputs()
        if self.p2p:
            for lane in range(self.c.lanes_used()):
                if self.c.shard_timed_out(lane):
                    raise RuntimeError("rank %d: a peer never delivered its slice (lane %d)" % (self.rank, lane))
        return out
