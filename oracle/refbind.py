"""ctypes bindings for oracle/_ref/libdash_ref.so -- TEST INFRASTRUCTURE.

The shared object is the UNMODIFIED reference (UzL-ITS/dash) compiled from
/root/reference by oracle/Makefile; see oracle/ref_driver.cu.  Only tests/,
__graft_entry__.smoke() and bench.py's reference / cpu_baseline legs may
import this module.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
REF_SO = os.path.join(_HERE, "_ref", "libdash_ref.so")

_lib = None


def available():
    return os.path.exists(REF_SO)


def lib():
    global _lib
    if _lib is None:
        if not available():
            raise RuntimeError("oracle/_ref/libdash_ref.so not built (run `make -C oracle ref` where /root/reference exists)")
        L = C.CDLL(REF_SO)
        L.ref_new.restype = C.c_void_p
        L.ref_free.argtypes = [C.c_void_p]
        L.ref_error.restype = C.c_char_p
        L.ref_error.argtypes = [C.c_void_p]
        L.ref_add_flatten.argtypes = [C.c_void_p] + [C.c_uint64] * 3
        L.ref_load_onnx.argtypes = [C.c_void_p, C.c_char_p, C.c_int, C.c_float, C.c_int]
        L.ref_add_dense.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_void_p, C.c_void_p, C.c_int, C.c_float, C.c_int, C.c_int]
        L.ref_add_conv2d.argtypes = [C.c_void_p] + [C.c_uint64] * 8 + [C.c_void_p, C.c_void_p, C.c_int, C.c_float, C.c_int]
        for f in (L.ref_add_relu, L.ref_add_sign):
            f.argtypes = [C.c_void_p] + [C.c_uint64] * 3
        L.ref_add_rescale_legacy.argtypes = [C.c_void_p, C.c_int] + [C.c_uint64] * 3
        L.ref_add_rescale.argtypes = [C.c_void_p, C.c_void_p, C.c_int] + [C.c_uint64] * 3
        L.ref_add_maxpool2d.argtypes = [C.c_void_p] + [C.c_uint64] * 7
        L.ref_add_base_extension.argtypes = [C.c_void_p, C.c_void_p, C.c_int] + [C.c_uint64] * 3
        L.ref_add_projection.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_int, C.c_int]
        L.ref_garble.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_uint64]
        for f in (L.ref_input_size, L.ref_output_size):
            f.restype = C.c_uint64
            f.argtypes = [C.c_void_p]
        for f in (L.ref_k, L.ref_num_garbled_layers, L.ref_max_modulus):
            f.argtypes = [C.c_void_p]
        L.ref_layer_out_modulus.argtypes = [C.c_void_p, C.c_int, C.c_int]
        L.ref_layer_out_size.restype = C.c_uint64
        L.ref_layer_out_size.argtypes = [C.c_void_p, C.c_int]
        L.ref_export.argtypes = [C.c_void_p, C.c_char_p]
        L.ref_garble_inputs.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.ref_cpu_evaluate.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.POINTER(C.c_double)]
        L.ref_cpu_evaluate_upto.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p]
        L.ref_decode_outputs.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.ref_plain_q_eval.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.ref_compress_hash.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        L.ref_decompress.argtypes = [C.c_int, C.c_void_p, C.c_void_p]
        L.ref_plain_num_layers.argtypes = [C.c_void_p]
        L.ref_plain_layer_type.argtypes = [C.c_void_p, C.c_int]
        L.ref_plain_layer_dims.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int]
        L.ref_plain_dense_floats.restype = C.c_long
        L.ref_plain_dense_floats.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_long]
        _lib = L
    return _lib


def _ptr_array(arrs):
    return (C.c_void_p * len(arrs))(*[a.ctypes.data for a in arrs])


def nr_comps(q):
    return lib().ref_nr_comps(int(q))


class RefCircuit:
    """Builds, garbles and evaluates one circuit with the reference."""

    def __init__(self):
        self.L = lib()
        self.h = C.c_void_p(self.L.ref_new())
        self.crt = None

    def close(self):
        if self.h:
            self.L.ref_free(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _chk(self, rc):
        if rc != 0:
            raise RuntimeError("reference: " + self.L.ref_error(self.h).decode())

    @staticmethod
    def _d3(dims):
        d = list(dims) + [0, 0]
        return d[0], d[1], d[2]

    def load_onnx(self, path, q_method=0, q_const=0.02, q_param=-1):
        """Appends the layers of an ONNX model through integration/dash_onnx_lite.h
        (the protobuf-free restatement of circuit/onnx_modelloader.h)."""
        n = self.L.ref_load_onnx(self.h, os.fsencode(path), q_method, q_const, q_param)
        if n < 0:
            raise RuntimeError("ref_load_onnx: " + self.L.ref_error(self.h).decode())
        return n

    def plain_layer_dims(self, layer, output=False):
        d = np.zeros(8, dtype=np.uint64)
        n = self.L.ref_plain_layer_dims(self.h, layer, int(output), d.ctypes.data, 8)
        return [int(x) for x in d[:n]]

    def plain_dense_floats(self, layer, biases=False):
        """float32 weights (or biases) of plaintext Dense layer `layer`, as stored by the loader."""
        n = self.L.ref_plain_dense_floats(self.h, layer, int(biases), None, 0)
        if n < 0:
            raise RuntimeError("plaintext layer %d is not a Dense layer" % layer)
        out = np.zeros(n, dtype=np.float32)
        self.L.ref_plain_dense_floats(self.h, layer, int(biases), out.ctypes.data, n)
        return out

    def add_flatten(self, dims):
        self._chk(self.L.ref_add_flatten(self.h, *self._d3(dims)))

    def add_dense(self, w, b, q_method=0, q_const=10.0, q_param=-1, channel_tf=0):
        w = np.ascontiguousarray(w, dtype=np.float32)
        b = np.ascontiguousarray(b, dtype=np.float32)
        self._chk(self.L.ref_add_dense(self.h, w.shape[0], w.shape[1], w.ctypes.data, b.ctypes.data, q_method, q_const, q_param, channel_tf))

    def add_conv2d(self, w, b, in_w, in_h, stride_w, stride_h, q_method=0, q_const=10.0, q_param=-1):
        """w: [filters][channels][f_h][f_w] float (OIHW)."""
        w = np.ascontiguousarray(w, dtype=np.float32)
        b = np.ascontiguousarray(b, dtype=np.float32)
        F, Cc, fh, fw = w.shape
        self._chk(self.L.ref_add_conv2d(self.h, in_w, in_h, Cc, F, fw, fh, stride_w, stride_h, w.ctypes.data, b.ctypes.data, q_method, q_const, q_param))

    def add_relu(self, dims):
        self._chk(self.L.ref_add_relu(self.h, *self._d3(dims)))

    def add_sign(self, dims):
        self._chk(self.L.ref_add_sign(self.h, *self._d3(dims)))

    def add_rescale_legacy(self, l, dims):
        self._chk(self.L.ref_add_rescale_legacy(self.h, l, *self._d3(dims)))

    def add_rescale(self, factors, dims):
        f = np.asarray(factors, dtype=np.int32)
        self._chk(self.L.ref_add_rescale(self.h, f.ctypes.data, len(f), *self._d3(dims)))

    def add_base_extension(self, extra_moduli, dims):
        e = np.asarray(extra_moduli, dtype=np.int32)
        self._chk(self.L.ref_add_base_extension(self.h, e.ctypes.data, len(e), *self._d3(dims)))

    def add_maxpool2d(self, w, h, c, kw, kh, sw=1, sh=1):
        self._chk(self.L.ref_add_maxpool2d(self.h, w, h, c, kw, kh, sw, sh))

    def add_projection(self, n, in_moduli, out_moduli, func=0):
        im = np.asarray(in_moduli, dtype=np.int32)
        om = np.asarray(out_moduli, dtype=np.int32)
        self._chk(self.L.ref_add_projection(self.h, n, im.ctypes.data, om.ctypes.data, len(im), func))

    def garble(self, crt_base, mrs_base=(), max_modulus=0, seed=1):
        crt = np.asarray(crt_base, dtype=np.int32)
        mrs = np.asarray(mrs_base, dtype=np.int32)
        self._chk(self.L.ref_garble(self.h, crt.ctypes.data, len(crt), mrs.ctypes.data if len(mrs) else None, len(mrs), max_modulus, seed))
        self.crt = [int(x) for x in crt]
        self.k = len(self.crt)
        self.input_size = self.L.ref_input_size(self.h)
        self.output_size = self.L.ref_output_size(self.h)
        self.n_layers = self.L.ref_num_garbled_layers(self.h)
        self.out_moduli = [self.L.ref_layer_out_modulus(self.h, -1, j) for j in range(self.k)] if self.n_layers else list(self.crt)

    def export(self, path):
        self._chk(self.L.ref_export(self.h, os.fsencode(path)))

    def garble_inputs(self, x):
        x = np.ascontiguousarray(x, dtype=np.int64).reshape(-1)
        assert x.size == self.input_size
        out = [np.zeros((self.input_size, nr_comps(q)), dtype=np.int16) for q in self.crt]
        self._chk(self.L.ref_garble_inputs(self.h, x.ctypes.data, _ptr_array(out)))
        return out

    def cpu_evaluate(self, labels, threads=1, want_output=True):
        """Returns (output labels per modulus, milliseconds inside gc->cpu_evaluate)."""
        labels = [np.ascontiguousarray(l, dtype=np.int16) for l in labels]
        out = [np.zeros((self.output_size, nr_comps(q)), dtype=np.int16) for q in self.out_moduli]
        ms = C.c_double(0)
        self._chk(self.L.ref_cpu_evaluate(self.h, _ptr_array(labels), threads, _ptr_array(out) if want_output else None, C.byref(ms)))
        return out, ms.value

    def cpu_evaluate_layer(self, labels, layer, threads=1):
        labels = [np.ascontiguousarray(l, dtype=np.int16) for l in labels]
        n = self.L.ref_layer_out_size(self.h, layer)
        mods = [self.L.ref_layer_out_modulus(self.h, layer, j) for j in range(self.k)]
        out = [np.zeros((n, nr_comps(q)), dtype=np.int16) for q in mods]
        self._chk(self.L.ref_cpu_evaluate_upto(self.h, _ptr_array(labels), threads, layer, _ptr_array(out)))
        return out

    def decode_outputs(self, labels):
        labels = [np.ascontiguousarray(l, dtype=np.int16) for l in labels]
        vals = np.zeros(self.output_size, dtype=np.int64)
        rc = self.L.ref_decode_outputs(self.h, _ptr_array(labels), vals.ctypes.data)
        if rc != 0:
            raise RuntimeError("reference decode_outputs: no matching label")
        return vals

    def plain_q_eval(self, x):
        x = np.ascontiguousarray(x, dtype=np.int64).reshape(-1)
        y = np.zeros(self.output_size, dtype=np.int64)
        self._chk(self.L.ref_plain_q_eval(self.h, x.ctypes.data, y.ctypes.data))
        return y


def compress_hash(q, digits):
    d = np.ascontiguousarray(digits, dtype=np.int16)
    c = np.zeros(2, dtype=np.uint64)
    h = np.zeros(2, dtype=np.uint64)
    lib().ref_compress_hash(int(q), d.ctypes.data, c.ctypes.data, h.ctypes.data)
    return (int(c[0]) | (int(c[1]) << 64)), (int(h[0]) | (int(h[1]) << 64))


def decompress(q, x):
    c = np.array([x & (2**64 - 1), x >> 64], dtype=np.uint64)
    d = np.zeros(nr_comps(q), dtype=np.int16)
    lib().ref_decompress(int(q), c.ctypes.data, d.ctypes.data)
    return d
