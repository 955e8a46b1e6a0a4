"""The inline-PTX carry chains of gc_common.cuh (frac_init, frac_stage) checked
instruction by instruction against big-integer arithmetic on the CPU.

The multiplicative digit extraction (gc::for_each_chunk_msb) has a plain C
statement that the host tests pin against the oracle
(test_cabi.py::test_multiplicative_digit_extraction_matches_division) and a
device statement written as PTX with mad.lo.cc / madc.hi.cc chains, which no
CPU test can execute.  This test reads the asm blocks out of the source file
and interprets them (mul.wide, mov.b64, mul.lo/hi, add.cc/addc, mad.lo.cc,
madc.lo/hi.cc with the carry flag), so a wrong operand number or a missing
carry shows up here and not only on the GPU."""
import os
import random
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = open(os.path.join(ROOT, "dash_b200", "csrc", "gc_common.cuh")).read()
M = 1 << 32


def asm_blocks(start, end):
    seg = SRC[SRC.index(start):SRC.index(end)]
    out = []
    for m in re.finditer(r"asm\((.*?)\);\n", seg, re.S):
        body = m.group(1)
        cut = re.search(r"\n\s+:", body)
        strs = re.findall(r'"((?:[^"\\]|\\.)*)"', body[:cut.start()])
        out.append("".join(strs).replace("\\n", "\n").replace("\\t", ""))
    return out


def run(text, regs):
    cc = 0
    loc = {}

    def get(t):
        t = t.strip()
        if t.startswith("%"):
            return regs[t]
        return int(t) if t.isdigit() else loc[t]

    def put(t, v):
        t = t.strip()
        if t.startswith("%"):
            regs[t] = v
        else:
            loc[t] = v

    for line in text.split("\n"):
        line = re.sub(r"/\*.*?\*/", "", line).strip().rstrip(";").strip()
        if not line or line in "{}" or line.startswith(".reg"):
            continue
        op, rest = line.split(None, 1)
        if op == "mov.b64":
            m = re.match(r"\{(.*?),(.*?)\},\s*(\w+)", rest)
            v = loc[m.group(3)]
            put(m.group(1), v % M)
            put(m.group(2), v >> 32)
            continue
        a = [x.strip() for x in rest.split(",")]
        if op == "mul.wide.u32":
            loc[a[0]] = get(a[1]) * get(a[2])
        elif op == "mov.u32":
            put(a[0], get(a[1]))
        elif op == "mul.lo.u32":
            put(a[0], (get(a[1]) * get(a[2])) % M)
        elif op == "mul.hi.u32":
            put(a[0], (get(a[1]) * get(a[2])) >> 32)
        elif op in ("add.cc.u32", "addc.cc.u32", "addc.u32"):
            v = get(a[1]) + get(a[2]) + (cc if op.startswith("addc") else 0)
            put(a[0], v % M)
            if ".cc" in op:
                cc = v >> 32
        elif op in ("mad.lo.cc.u32", "madc.lo.cc.u32", "madc.hi.cc.u32", "mad.hi.cc.u32"):
            p = get(a[1]) * get(a[2])
            p = p % M if ".lo" in op else p >> 32
            v = p + get(a[3]) + (cc if op.startswith("madc") else 0)
            put(a[0], v % M)
            cc = v >> 32
        else:
            raise AssertionError("unknown PTX instruction " + op)
    return regs


def rnd(rng, bits):
    c = rng.random()
    if c < 0.2:
        return (1 << bits) - 1
    if c < 0.3:
        return 0
    if c < 0.5:  # long runs of ones from the top: carries that ripple
        return rng.getrandbits(bits) | (((1 << bits) - 1) ^ ((1 << rng.randrange(bits)) - 1))
    return rng.getrandbits(bits)


def test_frac_init_chain():
    (text,) = asm_blocks("GC_HD void frac_init(", "/* fraction *= mult")
    rng = random.Random(7)
    n = 0
    while n < 4000:
        x, mu, addc = rnd(rng, 128), rnd(rng, 160), rnd(rng, 31)
        if x * mu >= 1 << 288:  # the kernels never see this: mu < 2^160 and x < q^nc keep x*mu below 2^288
            continue
        n += 1
        xl = [(x >> (32 * i)) % M for i in range(4)]
        ml = [(mu >> (32 * i)) % M for i in range(5)]
        regs = {"%5": xl[3], "%6": xl[2], "%7": xl[1], "%8": xl[0], "%14": addc}
        for j in range(5):
            regs["%%%d" % (9 + j)] = ml[j]
        run(text, regs)
        got = sum(regs["%%%d" % i] << (32 * i) for i in range(5))
        t = sum((xl[i] * ml[j]) << (32 * (i + j)) for i in range(4) for j in range(5) if i + j >= 3)
        assert got == ((t >> 128) + addc) % (1 << 160), (hex(x), hex(mu), addc)


def test_frac_stage_chains():
    blocks = asm_blocks("GC_HD uint32_t frac_stage(", "template <typename F>\nGC_HD void for_each_chunk_msb")
    assert len(blocks) == 4  # W = 5, 4, 3, 2 (W = 1 is plain C)
    rng = random.Random(9)
    for W, text in zip([5, 4, 3, 2], blocks):
        for _ in range(1500):
            limbs = [rnd(rng, 32) for _ in range(W)]
            mult = rnd(rng, 32)
            regs = {"%%%d" % (i + 1): v for i, v in enumerate(limbs)}
            regs["%%%d" % (W + 1)] = mult
            run(text, regs)
            p = sum(v << (32 * i) for i, v in enumerate(limbs)) * mult
            assert regs["%0"] == p >> (32 * W), (W, limbs, mult)
            assert sum(regs["%%%d" % (i + 1)] << (32 * i) for i in range(W)) == p % (1 << (32 * W))
