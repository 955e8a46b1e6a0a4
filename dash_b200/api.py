"""ctypes front end of libdash_b200.so -- host-side mirror of the reference's
GarbledCircuitInterface GPU calls (garbling/garbled_circuit_interface.h:
477-612,733-736):

    reference                      here
    ---------                      ----
    init_cuda() + cuda_move()      init(device); Circuit.load(path, max_batch)
    cuda_move_inputs(g_inputs)     Circuit.move_inputs(labels)
    cuda_evaluate(dev_in)          Circuit.evaluate_device()
    cuda_move_outputs()            Circuit.move_outputs()
    decode_outputs(g_out)          Circuit.decode_outputs()

Labels cross this boundary exactly as LabelTensor::get_components() lays
them out: per CRT modulus an int16 array [batch][labels][nc].  All compute
is in the CUDA library; this module only marshals pointers and raises
DashB200Error when the library reports a failure (including "no CUDA
device": there is no CPU path).
"""
import ctypes as C
import os

import numpy as np

from . import build as _build

_lib = None


class DashB200Error(RuntimeError):
    pass


def lib():
    """Loads (building if necessary) the shared library."""
    global _lib
    if _lib is None:
        path = _build.LIB
        if not os.path.exists(path):
            _build.build()
        L = C.CDLL(path)
        L.dashb_last_error.restype = C.c_char_p
        L.dashb_stream.restype = C.c_void_p
        L.dashb_circuit_load.argtypes = [C.POINTER(C.c_void_p), C.c_char_p]
        L.dashb_circuit_load_replicated.argtypes = [C.POINTER(C.c_void_p), C.c_char_p, C.c_int]
        L.dashb_finalize.argtypes = [C.c_void_p, C.c_uint64]
        L.dashb_move_inputs.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64]
        L.dashb_evaluate_device.argtypes = [C.c_void_p]
        L.dashb_garble_inputs.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64]
        L.dashb_move_outputs.argtypes = [C.c_void_p, C.c_void_p]
        L.dashb_evaluate.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p]
        L.dashb_layer_outputs.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        L.dashb_move_inputs_compressed.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64]
        L.dashb_move_outputs_compressed.argtypes = [C.c_void_p, C.c_void_p]
        L.dashb_evaluate_compressed.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p]
        L.dashb_decode_outputs.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        for f in (L.dashb_input_size, L.dashb_output_size, L.dashb_last_launch_count):
            f.restype = C.c_uint64
            f.argtypes = [C.c_void_p]
        for f in (L.dashb_num_layers, L.dashb_crt_base_size):
            f.argtypes = [C.c_void_p]
        L.dashb_output_modulus.argtypes = [C.c_void_p, C.c_int]
        L.dashb_crt_modulus.argtypes = [C.c_void_p, C.c_int]
        L.dashb_layer_type.argtypes = [C.c_void_p, C.c_int]
        L.dashb_layer_size.restype = C.c_uint64
        L.dashb_layer_size.argtypes = [C.c_void_p, C.c_int]
        for f in (L.dashb_algorithmic_bytes, L.dashb_aes_blocks, L.dashb_reference_aes_blocks):
            f.restype = C.c_double
            f.argtypes = [C.c_void_p, C.c_int]
        L.dashb_set_shard.argtypes = [C.c_void_p, C.c_int, C.c_int]
        L.dashb_evaluate_layers.argtypes = [C.c_void_p, C.c_int, C.c_int]
        L.dashb_current_plane.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
        L.dashb_current_slice_compress.argtypes = [C.c_void_p, C.c_int, C.c_uint64, C.c_uint64, C.c_void_p]
        L.dashb_current_slice_decompress.argtypes = [C.c_void_p, C.c_int, C.c_uint64, C.c_uint64, C.c_void_p]
        L.dashb_layer_slice.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
        L.dashb_set_shard_lanes.argtypes = [C.c_void_p, C.c_int]
        L.dashb_shard_lanes_used.argtypes = [C.c_void_p]
        L.dashb_lane_stream.restype = C.c_void_p
        L.dashb_lane_stream.argtypes = [C.c_void_p, C.c_int]
        L.dashb_shard_export.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
        L.dashb_shard_import.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.dashb_evaluate_layers_lane.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int]
        L.dashb_shard_exchange.argtypes = [C.c_void_p, C.c_int, C.c_int]
        L.dashb_evaluate_sharded.argtypes = [C.c_void_p]
        L.dashb_shard_exchange_bytes.restype = C.c_uint64
        L.dashb_shard_exchange_bytes.argtypes = [C.c_void_p, C.c_int, C.c_int]
        L.dashb_shard_status.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_uint32)]
        L.dashb_set_layer_timing.argtypes = [C.c_void_p, C.c_int]
        L.dashb_layer_time_ms.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_float)]
        L.dashb_lanes_used.argtypes = [C.c_void_p]
        L.dashb_layer_interval_ms.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_float), C.POINTER(C.c_float)]
        L.dashb_circuit_destroy.argtypes = [C.c_void_p]
        L.dashb_hostcheck_compress.argtypes = [C.c_int, C.c_void_p, C.c_void_p]
        L.dashb_hostcheck_digits.argtypes = [C.c_int, C.c_void_p, C.c_void_p]
        L.dashb_hostcheck_colour.argtypes = [C.c_int, C.c_void_p]
        L.dashb_hostcheck_digits_frac.argtypes = [C.c_int, C.c_void_p, C.c_void_p]
        L.dashb_hostcheck_frac_sweep.argtypes = [C.c_int, C.c_int, C.c_int, C.c_uint64, C.c_void_p]
        L.dashb_hostcheck_hash.argtypes = [C.c_void_p, C.c_void_p]
        L.dashb_hostcheck_ladder_sweep.argtypes = [C.c_int, C.c_int, C.c_int, C.c_uint64, C.c_int, C.c_void_p]
        L.dashb_hostcheck_compress_words.argtypes = [C.c_int, C.c_void_p, C.c_void_p]
        _lib = L
    return _lib


def _check(rc):
    if rc != 0:
        raise DashB200Error("dash_b200 error %d: %s" % (rc, lib().dashb_last_error().decode()))


def init(device=0):
    """init_cuda() + CUDAAESEngine setup of the reference, for one device."""
    _check(lib().dashb_init(int(device)))


def nr_comps(q):
    return lib().dashb_nr_comps(int(q))


def stream_handle():
    return lib().dashb_stream()


LAYER_NAMES = {1: "dense", 2: "conv2d", 3: "approx_relu", 4: "sign", 5: "rescale", 6: "rescale", 7: "projection",
               8: "max_pool", 9: "base_extension"}


def _ptr_array(arrs):
    return (C.c_void_p * len(arrs))(*[a.ctypes.data for a in arrs])


class Circuit:
    """A garbled circuit resident on the GPU (the result of cuda_move())."""

    def __init__(self, handle, max_batch):
        self.L = lib()
        self.h = handle
        self.max_batch = max_batch
        self.k = self.L.dashb_crt_base_size(self.h)
        self.input_size = self.L.dashb_input_size(self.h)
        self.output_size = self.L.dashb_output_size(self.h)
        self.n_layers = self.L.dashb_num_layers(self.h)
        self.out_moduli = [self.L.dashb_output_modulus(self.h, j) for j in range(self.k)]
        self.batch = 0
        self.in_moduli = [self.L.dashb_crt_modulus(self.h, j) for j in range(self.k)]

    @classmethod
    def load(cls, path, max_batch=1, shard=None, shard_lanes=0, table_replicas=1):
        """Reads a DGCM material file and moves the circuit to the device.
        shard=(rank, world) makes it one rank of an output-channel-sharded
        evaluation (see dash_b200.scheduler.ChannelShardedEvaluator);
        shard_lanes > 0 prepares the peer-memory exchange with that many
        sub-batch lanes (otherwise the exchange goes through NCCL)."""
        L = lib()
        h = C.c_void_p()
        if table_replicas > 1:  # one copy of the sign / ReLU tables per inference (measurement; dashb_set_table_replicas)
            _check(L.dashb_circuit_load_replicated(C.byref(h), os.fsencode(path), int(table_replicas)))
        else:
            _check(L.dashb_circuit_load(C.byref(h), os.fsencode(path)))
        if shard is not None:
            _check(L.dashb_set_shard(h, int(shard[0]), int(shard[1])))
            if shard_lanes:
                _check(L.dashb_set_shard_lanes(h, int(shard_lanes)))
        _check(L.dashb_finalize(h, int(max_batch)))
        return cls(h, max_batch)

    # ---- channel sharding -------------------------------------------------
    def evaluate_layers(self, first, last):
        _check(self.L.dashb_evaluate_layers(self.h, int(first), int(last)))

    # peer-memory exchange (one process per GPU, CUDA IPC)
    def shard_export(self, lane):
        buf = C.create_string_buffer(192)
        _check(self.L.dashb_shard_export(self.h, lane, buf))
        return bytes(buf.raw)

    def shard_import(self, lane, peer_rank, handles):
        _check(self.L.dashb_shard_import(self.h, lane, int(peer_rank), C.create_string_buffer(bytes(handles), 192)))

    def evaluate_layers_lane(self, lane, first, last):
        _check(self.L.dashb_evaluate_layers_lane(self.h, int(lane), int(first), int(last)))

    def shard_exchange(self, lane, layer):
        _check(self.L.dashb_shard_exchange(self.h, int(lane), int(layer)))

    def shard_exchange_bytes(self, lane, layer):
        return self.L.dashb_shard_exchange_bytes(self.h, int(lane), int(layer))

    def shard_timed_out(self, lane):
        v = C.c_uint32(0)
        _check(self.L.dashb_shard_status(self.h, int(lane), C.byref(v)))
        return v.value

    def lanes_used(self):
        return self.L.dashb_lanes_used(self.h)

    def current_plane(self, j):
        """(device pointer, bytes per element, elements) of plane j of the
        tensor the last evaluated layer produced."""
        ptr, bpe, n = C.c_void_p(), C.c_uint64(), C.c_uint64()
        _check(self.L.dashb_current_plane(self.h, j, C.byref(ptr), C.byref(bpe), C.byref(n)))
        return ptr.value, bpe.value, n.value

    def slice_compress(self, j, e0, e1, dev_ptr):
        _check(self.L.dashb_current_slice_compress(self.h, j, e0, e1, dev_ptr))

    def slice_decompress(self, j, e0, e1, dev_ptr):
        _check(self.L.dashb_current_slice_decompress(self.h, j, e0, e1, dev_ptr))

    def layer_slice(self, layer, rank):
        e0, e1 = C.c_uint64(), C.c_uint64()
        _check(self.L.dashb_layer_slice(self.h, layer, rank, C.byref(e0), C.byref(e1)))
        return e0.value, e1.value

    def sync(self):
        import torch

        torch.cuda.ExternalStream(stream_handle()).synchronize()

    def close(self):
        if self.h:
            self.L.dashb_circuit_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _in_arrays(self, labels):
        labels = [np.ascontiguousarray(l, dtype=np.int16) for l in labels]
        if len(labels) != self.k:
            raise ValueError("expected %d label tensors" % self.k)
        batch = labels[0].shape[0]
        for q, l in zip(self.in_moduli, labels):
            # the library reads batch * input_size * nc(q) digits from each buffer: a wrong last dimension
            # would make it read past the array
            if l.ndim != 3 or l.shape[0] != batch or l.shape[1] != self.input_size or l.shape[2] != nr_comps(q):
                raise ValueError("labels of modulus %d must be [batch][input_size][%d], got %s" % (q, nr_comps(q), l.shape))
        return labels, batch

    def move_inputs(self, labels):
        labels, batch = self._in_arrays(labels)
        _check(self.L.dashb_move_inputs(self.h, _ptr_array(labels), batch))
        self.batch = batch
        self._keep = labels

    def garble_inputs(self, values):
        """garble_inputs on the device: values int64 [batch][input_size]
        (needs a material file that carries the garbler's input encoding)."""
        v = np.ascontiguousarray(values, dtype=np.int64)
        if v.ndim != 2 or v.shape[1] != self.input_size:
            raise ValueError("values must be [batch][input_size]")
        _check(self.L.dashb_garble_inputs(self.h, v.ctypes.data, v.shape[0]))
        self.batch = v.shape[0]
        self._keep = v

    def evaluate_device(self):
        _check(self.L.dashb_evaluate_device(self.h))

    def _out_arrays(self, n, moduli):
        return [np.zeros((self.batch, n, nr_comps(q)), dtype=np.int16) for q in moduli]

    def move_outputs(self):
        out = self._out_arrays(self.output_size, self.out_moduli)
        _check(self.L.dashb_move_outputs(self.h, _ptr_array(out)))
        return out

    def evaluate(self, labels):
        """cuda_move_inputs + cuda_evaluate + cuda_move_outputs."""
        labels, batch = self._in_arrays(labels)
        self.batch = batch
        out = self._out_arrays(self.output_size, self.out_moduli)
        _check(self.L.dashb_evaluate(self.h, _ptr_array(labels), batch, _ptr_array(out)))
        return out

    # ---- compressed 128-bit labels on the host side (uint64 [batch][label][2]) ----
    def _in_compressed(self, labels):
        labels = [np.ascontiguousarray(l, dtype=np.uint64) for l in labels]
        if len(labels) != self.k:
            raise ValueError("expected %d label tensors" % self.k)
        batch = labels[0].shape[0]
        for l in labels:
            if l.shape != (batch, self.input_size, 2):
                raise ValueError("compressed labels must be uint64 [batch][input_size][2]")
        return labels, batch

    def move_inputs_compressed(self, labels):
        labels, batch = self._in_compressed(labels)
        _check(self.L.dashb_move_inputs_compressed(self.h, _ptr_array(labels), batch))
        self.batch = batch
        self._keep = labels

    def move_outputs_compressed(self):
        out = [np.zeros((self.batch, self.output_size, 2), dtype=np.uint64) for _ in range(self.k)]
        _check(self.L.dashb_move_outputs_compressed(self.h, _ptr_array(out)))
        return out

    def evaluate_compressed(self, labels):
        labels, batch = self._in_compressed(labels)
        self.batch = batch
        out = [np.zeros((batch, self.output_size, 2), dtype=np.uint64) for _ in range(self.k)]
        _check(self.L.dashb_evaluate_compressed(self.h, _ptr_array(labels), batch, _ptr_array(out)))
        return out

    def layer_outputs(self, layer, n, moduli):
        out = self._out_arrays(n, moduli)
        _check(self.L.dashb_layer_outputs(self.h, layer, _ptr_array(out)))
        return out

    def decode_outputs(self):
        vals = np.zeros((self.batch, self.output_size), dtype=np.int64)
        ok = np.zeros(self.batch, dtype=np.uint8)
        _check(self.L.dashb_decode_outputs(self.h, vals.ctypes.data, ok.ctypes.data))
        return vals, ok.astype(bool)

    def launch_count(self):
        return self.L.dashb_last_launch_count(self.h)

    def algorithmic_bytes(self, layer=-1):
        return self.L.dashb_algorithmic_bytes(self.h, layer)

    def aes_blocks(self, layer=-1):
        return self.L.dashb_aes_blocks(self.h, layer)

    def reference_aes_blocks(self, layer=-1):
        return self.L.dashb_reference_aes_blocks(self.h, layer)

    def layer_type(self, layer):
        return self.L.dashb_layer_type(self.h, layer)

    def layer_size(self, layer):
        return self.L.dashb_layer_size(self.h, layer)

    def set_layer_timing(self, enabled):
        """False/0 off, True/1 per-layer events with the lanes serialised, 2 per-(lane, layer) intervals in the
        concurrent step (layer_intervals_ms)."""
        _check(self.L.dashb_set_layer_timing(self.h, int(enabled)))

    def layer_intervals_ms(self):
        """[(layer name, layer index, lane, begin ms, end ms)] of the last evaluation in timing mode 2."""
        out = []
        for lane in range(self.L.dashb_lanes_used(self.h)):
            for i in range(self.n_layers):
                a, b = C.c_float(0), C.c_float(0)
                _check(self.L.dashb_layer_interval_ms(self.h, lane, i, C.byref(a), C.byref(b)))
                out.append((LAYER_NAMES.get(self.layer_type(i), "?"), i, lane, a.value, b.value))
        return out

    def layer_times_ms(self):
        """Per layer-type device time of the last timed evaluation, keyed like
        the reference's BENCHMARK map (garbled_circuit_interface.h:40-57)."""
        per_layer = []
        for i in range(self.n_layers):
            ms = C.c_float(0)
            _check(self.L.dashb_layer_time_ms(self.h, i, C.byref(ms)))
            per_layer.append((LAYER_NAMES.get(self.layer_type(i), "?"), ms.value))
        return per_layer


# ---- host self-checks of the device primitives (CPU-only test tier) ----
def hostcheck_compress(q, digits):
    d = np.ascontiguousarray(digits, dtype=np.int16)
    o = np.zeros(2, dtype=np.uint64)
    lib().dashb_hostcheck_compress(int(q), d.ctypes.data, o.ctypes.data)
    return int(o[0]) | (int(o[1]) << 64)


def hostcheck_digits(q, x):
    c = np.array([x & (2**64 - 1), x >> 64], dtype=np.uint64)
    d = np.zeros(nr_comps(q), dtype=np.int16)
    lib().dashb_hostcheck_digits(int(q), c.ctypes.data, d.ctypes.data)
    return d


def hostcheck_digits_frac(q, x):
    """Digits of a valid payload (x < q^nc) through the multiplicative
    extraction of the gadget kernels; None if the modulus has no plan."""
    c = np.array([x & (2**64 - 1), x >> 64], dtype=np.uint64)
    d = np.zeros(nr_comps(q), dtype=np.int16)
    if not lib().dashb_hostcheck_digits_frac(int(q), c.ctypes.data, d.ctypes.data):
        return None
    return d


def hostcheck_frac_sweep(q_lo, q_hi, samples, seed=1):
    """(moduli on which the two digit extractions differ, moduli with a plan)
    over [q_lo, q_hi], `samples` random valid payloads + edge payloads each."""
    n = C.c_int(0)
    bad = lib().dashb_hostcheck_frac_sweep(int(q_lo), int(q_hi), int(samples), int(seed), C.byref(n))
    return bad, n.value


def hostcheck_ladder_sweep(q_lo, q_hi, samples, seed=1, aligned=False):
    """Same sweep through the width-ladder walk the sign kernel uses, for plain
    or word-aligned plans: (moduli that differ or whose plan bookkeeping is
    inconsistent, moduli with a plan)."""
    n = C.c_int(0)
    bad = lib().dashb_hostcheck_ladder_sweep(int(q_lo), int(q_hi), int(samples), int(seed), int(bool(aligned)), C.byref(n))
    return bad, n.value


def hostcheck_compress_words(q, digits):
    """Word-wise compress of a u8 digit-plane label (gc::WordCompressor); None for q = 256."""
    d = np.ascontiguousarray(digits, dtype=np.int16)
    o = np.zeros(2, dtype=np.uint64)
    if not lib().dashb_hostcheck_compress_words(int(q), d.ctypes.data, o.ctypes.data):
        return None
    return int(o[0]) | (int(o[1]) << 64)


def hostcheck_colour(q, x):
    c = np.array([x & (2**64 - 1), x >> 64], dtype=np.uint64)
    return lib().dashb_hostcheck_colour(int(q), c.ctypes.data)


def hostcheck_hash(x):
    c = np.array([x & (2**64 - 1), x >> 64], dtype=np.uint64)
    o = np.zeros(2, dtype=np.uint64)
    lib().dashb_hostcheck_hash(c.ctypes.data, o.ctypes.data)
    return int(o[0]) | (int(o[1]) << 64)
