"""The C-ABI library loads, exports every symbol include/dash_b200.h
declares, fails loudly without a GPU, and its device primitives -- run on
the host through the dashb_hostcheck_* hooks -- agree with the oracle.
CPU only: no compute call needs a device."""
import ctypes
import os
import random
import re

import pytest

from helpers import ROOT
from oracle import oraclebind as O


@pytest.fixture(scope="module")
def api():
    from dash_b200 import api as a

    a.lib()
    return a


def declared_symbols():
    syms = set()
    for header in ("dash_b200.h", "dash_b200_garble.h"):
        text = open(os.path.join(ROOT, "include", header)).read()
        text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
        syms |= set(re.findall(r"\b(dashb_[a-z0-9_]+)\s*\(", text))
    return sorted(syms)


def test_library_exports_every_declared_symbol(api):
    L = api.lib()
    syms = declared_symbols()
    assert len(syms) >= 25
    for s in syms:
        assert hasattr(L, s), "libdash_b200.so does not export " + s


def test_garbler_fails_loudly_without_a_device(api):
    """The offline phase has no CPU path either."""
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from dash_b200 import garble

    layers = [{"type": "dense", "M": 2, "K": 3, "weights": [[1, 2, 3], [4, 5, 6]], "biases": [0, 1]}]
    with pytest.raises(api.DashB200Error):
        garble.garble_to_file([2, 3, 5], [], 3, layers, "/tmp/never.gcm")
    assert not os.path.exists("/tmp/never.gcm")


def test_no_gpu_fails_loudly(api):
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(api.DashB200Error) as e:
        api.init(0)
    assert "no CUDA device" in str(e.value) or "CPU" in str(e.value)
    # nothing can be built or evaluated either
    h = ctypes.c_void_p()
    rc = api.lib().dashb_circuit_load(ctypes.byref(h), os.fsencode(os.path.join(ROOT, "tests", "golden", "MLP_TINY.gcm")))
    assert rc == -2 and not h  # DASHB_ENODEV


def test_product_does_not_link_the_oracle():
    import subprocess

    from dash_b200 import build

    out = subprocess.run(["nm", "-D", build.LIB], capture_output=True, text=True).stdout
    assert "oracle_" not in out and "ref_" not in out.replace("deref", "")


def test_device_primitives_on_host_match_oracle(api):
    rng = random.Random(11)
    pt = int.from_bytes(bytes.fromhex("00112233445566778899aabbccddeeff"), "little")
    assert api.hostcheck_hash(pt).to_bytes(16, "little").hex() == "69c4e0d86a7b0430d8cdb78070b4c55a"
    for q in [2, 3, 5, 7, 8, 9, 11, 13, 17, 19, 23, 32, 45, 63, 72, 81, 91, 97, 98, 107, 126, 138, 167, 173, 255, 256, 882]:
        nc = O.nr_comps(q)
        assert api.nr_comps(q) == nc
        for t in range(60):
            d = [rng.randrange(q) for _ in range(nc)]
            c = O.compress(q, d)
            assert api.hostcheck_compress(q, d) == c
            assert list(api.hostcheck_digits(q, c)) == d
            assert api.hostcheck_colour(q, c) == d[0]
            x = rng.getrandbits(128)
            assert api.hostcheck_hash(x) == O.hash128(x)
            assert api.hostcheck_colour(q, x) == x % q
            if q <= 181:  # out-of-range payloads: identical digits as long as the top quotient fits int16
                assert list(api.hostcheck_digits(q, x)) == list(O.decompress(q, x))


def test_multiplicative_digit_extraction_matches_division(api):
    """gc::for_each_chunk_msb (the gadget kernels' digit extraction for valid
    payloads, x < q^nc) against the oracle's decompress (LabelTensor::
    decompress, garbling/label_tensor.h:726-742) and the long-division
    extraction, for every modulus a circuit can use: CRT moduli 2..256, MRS
    digits and cast moduli (k+1)*m_i; edge payloads included."""
    rng = random.Random(5)
    planned = 0
    for q in list(range(2, 257)) + [9 * 98, 10 * 76, 9 * 9, 10 * 7, 8 * 10, 9 * 126, 13 * 126, 1000, 4096, 1638, 32767]:
        nc = O.nr_comps(q)
        D = q**nc
        xs = [0, 1, D - 1, D - 2, D // 2, D // 3] + [q**i for i in range(nc)] + [q**i - 1 for i in range(1, nc + 1)]
        xs += [rng.randrange(D) for _ in range(40)]
        for x in xs:
            got = api.hostcheck_digits_frac(q, x)
            if got is None:  # no plan: the kernels fall back to the division
                break
            want = [(x // q**i) % q for i in range(nc)]
            assert list(got) == want, (q, x)
            if q <= 256:
                assert list(O.decompress(q, x)) == want
            assert list(api.hostcheck_digits(q, x)) == want
        else:
            planned += 1
    assert planned >= 260


def test_multiplicative_digit_extraction_every_modulus(api):
    """Every modulus a circuit can name (CRT moduli, MRS digits, cast moduli
    (k+1)*m_i: 2..32767): the plan's extraction against the long division on
    random valid payloads and the edge payloads, in C++ (the Python loop above
    covers 2..256 against the oracle)."""
    bad, planned = api.hostcheck_frac_sweep(2, 32767, 200, seed=3)
    assert bad == 0
    assert planned > 1500  # every modulus cut into at most 10 chunks has a plan


def test_ladder_walk_every_modulus(api):
    """gc::for_each_chunk_ladder (stage width known at compile time, trip
    counts from the plan) against the long division for every modulus
    2..32767, and the word-aligned plans the ReLU half gates use for the CRT
    moduli: each stage after the first covers whole 32-bit words of a digit
    plane."""
    bad, planned = api.hostcheck_ladder_sweep(2, 32767, 100, seed=5)
    assert bad == 0 and planned > 1500
    bad, planned = api.hostcheck_ladder_sweep(2, 256, 300, seed=6, aligned=True)
    assert bad == 0
    assert planned >= 70  # every modulus whose chunks hold at least four digits (q <= 76 and a few more)
    for q in (2, 3, 5, 7, 11, 13, 17, 19, 23, 29, 31, 32):  # the CRT moduli of every configuration
        assert api.hostcheck_ladder_sweep(q, q, 1, seed=1, aligned=True) == (0, 1), q


def test_word_wise_compress_matches_oracle(api):
    """gc::WordCompressor (four digits per step, dp4a) == LabelTensor::compress
    for every u8 modulus; q = 256 has no word-wise form and keeps the
    digit-wise path."""
    rng = random.Random(11)
    for q in range(2, 257):
        nc = O.nr_comps(q)
        for t in range(12):
            d = [rng.randrange(q) for _ in range(nc)] if t else [q - 1] * nc
            got = api.hostcheck_compress_words(q, d)
            if q == 256:
                assert got is None
                break
            assert got == O.compress(q, d), (q, t)
