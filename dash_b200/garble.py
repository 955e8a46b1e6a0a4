"""ctypes front end of dashb_garble_to_file (include/dash_b200_garble.h): the
offline phase on the GPU.  A plaintext circuit is a list of layer dicts:

    {"type": "dense", "M": m, "K": k, "channel_tf": 0, "weights": int64 [M][K], "biases": int64 [M]}
    {"type": "conv2d", "geom": (in_w, in_h, channels, filters, f_w, f_h, s_w, s_h),
     "weights": int64 (reference order), "biases": int64 [filters]}
    {"type": "relu"} | {"type": "sign"} | {"type": "rescale", "factors": (32,)} | {"type": "rescale_legacy", "l": 2}
"""
import ctypes as C
import os

import numpy as np

from . import api

TYPES = {"dense": 1, "conv2d": 2, "relu": 3, "sign": 4, "rescale": 5, "rescale_legacy": 6}


class PlainLayer(C.Structure):
    _fields_ = [("p", C.c_int64 * 16), ("weights", C.c_void_p), ("biases", C.c_void_p)]


class GarbleStats(C.Structure):
    _fields_ = [("kernel_ms", C.c_double), ("total_ms", C.c_double), ("gates", C.c_uint64), ("rows", C.c_uint64),
                ("table_bytes", C.c_uint64)]


def _marshal(layers):
    arr = (PlainLayer * len(layers))()
    keep = []
    for i, l in enumerate(layers):
        t = TYPES[l["type"]]
        arr[i].p[0] = t
        if t == 1:
            arr[i].p[1], arr[i].p[2], arr[i].p[3] = int(l["M"]), int(l["K"]), int(l.get("channel_tf", 0))
        elif t == 2:
            for x, v in enumerate(l["geom"]):
                arr[i].p[1 + x] = int(v)
        elif t == 6:
            arr[i].p[2] = int(l["l"])
        elif t == 5:
            arr[i].p[2] = len(l["factors"])
            for x, v in enumerate(l["factors"]):
                arr[i].p[3 + x] = int(v)
        if t in (1, 2):
            w = np.ascontiguousarray(l["weights"], dtype=np.int64).ravel()
            b = np.ascontiguousarray(l["biases"], dtype=np.int64).ravel()
            keep += [w, b]
            arr[i].weights = w.ctypes.data
            arr[i].biases = b.ctypes.data
    return arr, keep


def _stats(st):
    return {f: getattr(st, f) for f, _ in GarbleStats._fields_}


def _seed16(seed):
    """The PRF key of a garbling.  None = 16 fresh random bytes from the OS
    (every label, the offsets R_q included, derives from it, so a deployment
    must never reuse or publish it); tests and benchmarks that need
    reproducible material pass their own constant."""
    if seed is None:
        return os.urandom(16)
    return (bytes(seed) + b"\0" * 16)[:16]


def garble_to_file(crt, mrs, input_size, layers, path, seed=None, max_modulus=0, secrets_path=None):
    """Garbles `layers` on the current device and writes the DGCM material to
    `path`.  With `secrets_path` the input encoding (base labels, offsets R_q)
    and the decoding information go to that second, garbler-only file and
    `path` holds only what the evaluating party may see; without it one file
    holds everything (single-party tests and benchmarks).  Returns the
    statistics as a dict."""
    L = api.lib()
    arr, keep = _marshal(layers)
    crt_a = np.ascontiguousarray(crt, dtype=np.int32)
    mrs_a = np.ascontiguousarray(mrs if len(mrs) else [0], dtype=np.int32)
    seed = _seed16(seed)
    st = GarbleStats()
    if secrets_path is None:
        L.dashb_garble_to_file.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_uint64, C.c_void_p, C.c_int,
                                           C.c_char_p, C.c_char_p, C.c_void_p]
        api._check(L.dashb_garble_to_file(crt_a.ctypes.data, len(crt), mrs_a.ctypes.data, len(mrs), int(max_modulus),
                                          int(input_size), C.byref(arr), len(layers), seed, os.fsencode(path), C.byref(st)))
    else:
        L.dashb_garble_to_files.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_uint64, C.c_void_p, C.c_int,
                                            C.c_char_p, C.c_char_p, C.c_char_p, C.c_void_p]
        api._check(L.dashb_garble_to_files(crt_a.ctypes.data, len(crt), mrs_a.ctypes.data, len(mrs), int(max_modulus),
                                           int(input_size), C.byref(arr), len(layers), seed, os.fsencode(path),
                                           os.fsencode(secrets_path), C.byref(st)))
    del keep
    return _stats(st)


def garble_to_circuit(crt, mrs, input_size, layers, max_batch, seed=None, max_modulus=0, shard=None, shard_lanes=0):
    """Garbles `layers` straight into an evaluator circuit on the current
    device (the tables never leave the GPU).  shard=(rank, world) makes it one
    rank of an output-channel-sharded evaluation (every rank garbles the same
    seed).  Returns (api.Circuit, stats)."""
    L = api.lib()
    L.dashb_garble_to_circuit.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_uint64, C.c_void_p, C.c_int,
                                          C.c_char_p, C.POINTER(C.c_void_p), C.c_void_p]
    arr, keep = _marshal(layers)
    crt_a = np.ascontiguousarray(crt, dtype=np.int32)
    mrs_a = np.ascontiguousarray(mrs if len(mrs) else [0], dtype=np.int32)
    seed = _seed16(seed)
    st = GarbleStats()
    h = C.c_void_p()
    api._check(L.dashb_garble_to_circuit(crt_a.ctypes.data, len(crt), mrs_a.ctypes.data, len(mrs), int(max_modulus),
                                         int(input_size), C.byref(arr), len(layers), seed, C.byref(h), C.byref(st)))
    del keep
    if shard is not None:
        api._check(L.dashb_set_shard(h, int(shard[0]), int(shard[1])))
        if shard_lanes:
            api._check(L.dashb_set_shard_lanes(h, int(shard_lanes)))
    api._check(L.dashb_finalize(h, int(max_batch)))
    return api.Circuit(h, max_batch), _stats(st)
