"""ctypes bindings for oracle/libdash_oracle.so -- TEST INFRASTRUCTURE.

The plain-C restatement of the reference's evaluator (oracle/dash_oracle.c),
used only as the parity checker by tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline leg.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_SO = os.path.join(_HERE, "libdash_oracle.so")
_lib = None


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE, "libdash_oracle.so"])


def lib():
    global _lib
    if _lib is None:
        try:  # make is a no-op when the library is newer than its sources
            build()
        except Exception:
            if not os.path.exists(ORACLE_SO):
                raise
        L = C.CDLL(ORACLE_SO)
        L.oracle_load.restype = C.c_void_p
        L.oracle_load.argtypes = [C.c_char_p]
        L.oracle_free.argtypes = [C.c_void_p]
        for f in (L.oracle_k, L.oracle_num_layers):
            f.argtypes = [C.c_void_p]
        L.oracle_input_size.restype = C.c_uint64
        L.oracle_input_size.argtypes = [C.c_void_p]
        L.oracle_layer_out_size.restype = C.c_uint64
        L.oracle_layer_out_size.argtypes = [C.c_void_p, C.c_int]
        L.oracle_layer_out_modulus.argtypes = [C.c_void_p, C.c_int, C.c_int]
        L.oracle_evaluate.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int]
        L.oracle_decode.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.oracle_aes_blocks.restype = C.c_double
        L.oracle_aes_blocks.argtypes = [C.c_void_p]
        L.oracle_compress.argtypes = [C.c_int, C.c_void_p, C.c_void_p]
        L.oracle_decompress.argtypes = [C.c_int, C.c_void_p, C.c_void_p]
        L.oracle_hash.argtypes = [C.c_void_p, C.c_void_p]
        L.oracle_aes128_fixed_key.argtypes = [C.c_void_p, C.c_void_p]
        L.oracle_mul_inv.restype = C.c_int64
        L.oracle_mul_inv.argtypes = [C.c_uint64, C.c_int]
        _lib = L
    return _lib


def _ptr_array(arrs):
    return (C.c_void_p * len(arrs))(*[a.ctypes.data for a in arrs])


def nr_comps(q):
    return lib().oracle_nr_comps(int(q))


def aes128_fixed_key(block16):
    b = np.frombuffer(bytes(block16), dtype=np.uint8).copy()
    o = np.zeros(16, dtype=np.uint8)
    lib().oracle_aes128_fixed_key(b.ctypes.data, o.ctypes.data)
    return bytes(o)


def compress(q, digits):
    d = np.ascontiguousarray(digits, dtype=np.int16)
    o = np.zeros(2, dtype=np.uint64)
    lib().oracle_compress(int(q), d.ctypes.data, o.ctypes.data)
    return int(o[0]) | (int(o[1]) << 64)


def decompress(q, x):
    c = np.array([x & (2**64 - 1), x >> 64], dtype=np.uint64)
    d = np.zeros(nr_comps(q), dtype=np.int16)
    lib().oracle_decompress(int(q), c.ctypes.data, d.ctypes.data)
    return d


def hash128(x):
    c = np.array([x & (2**64 - 1), x >> 64], dtype=np.uint64)
    o = np.zeros(2, dtype=np.uint64)
    lib().oracle_hash(c.ctypes.data, o.ctypes.data)
    return int(o[0]) | (int(o[1]) << 64)


class OracleCircuit:
    def __init__(self, path):
        self.L = lib()
        self.h = C.c_void_p(self.L.oracle_load(os.fsencode(path)))
        if not self.h:
            raise RuntimeError("oracle: cannot load " + str(path))
        self.k = self.L.oracle_k(self.h)
        self.n_layers = self.L.oracle_num_layers(self.h)
        self.input_size = self.L.oracle_input_size(self.h)
        self.crt = [self.L.oracle_layer_out_modulus(self.h, -1, j) for j in range(self.k)] if self.n_layers == 0 else None

    def close(self):
        if self.h:
            self.L.oracle_free(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def out_shape(self, layer=-1):
        n = self.L.oracle_layer_out_size(self.h, layer)
        mods = [self.L.oracle_layer_out_modulus(self.h, layer, j) for j in range(self.k)]
        return n, mods

    def evaluate(self, labels, upto=-1, threads=1):
        """labels: list of k int16 arrays [input_size][nc]; one inference."""
        labels = [np.ascontiguousarray(l, dtype=np.int16) for l in labels]
        n, mods = self.out_shape(upto)
        out = [np.zeros((n, nr_comps(q)), dtype=np.int16) for q in mods]
        rc = self.L.oracle_evaluate(self.h, _ptr_array(labels), upto, _ptr_array(out), threads)
        if rc != 0:
            raise RuntimeError("oracle_evaluate failed")
        return out

    def decode(self, labels):
        labels = [np.ascontiguousarray(l, dtype=np.int16) for l in labels]
        n, _ = self.out_shape(-1)
        vals = np.zeros(n, dtype=np.int64)
        missing = self.L.oracle_decode(self.h, _ptr_array(labels), vals.ctypes.data)
        return vals, missing

    def aes_blocks(self):
        return self.L.oracle_aes_blocks(self.h)
