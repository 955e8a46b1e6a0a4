/*
 * gc_common.cuh -- label primitives shared by every kernel of the B200
 * evaluator: 128-bit arithmetic, base-q compress / digit extraction, the
 * fixed-key AES-128 hash.  Everything that is pure arithmetic is
 * __host__ __device__ so that the CPU-only test tier can run the very same
 * code (dashb_hostcheck_* in dashb_api.cu) against the oracle.
 *
 * Reference semantics reproduced here (paths relative to dash/include/):
 *   compress   LabelTensor::compress      garbling/label_tensor.h:715-724
 *   digits     LabelTensor::decompress    garbling/label_tensor.h:726-742
 *   hash       CPUAESEngine::cipher       crypto/cpu_aes_engine.h:34-40
 *              (== cuda_cipher, crypto/cuda_aes_engine.h:352-522)
 */
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define GC_HD __host__ __device__ __forceinline__
#define GC_D __device__ __forceinline__
#else
#define GC_HD inline
#define GC_D inline
#endif

namespace gc {

struct u128 {
    uint64_t lo, hi;
};

GC_HD u128 make_u128(uint64_t lo, uint64_t hi) {
    u128 r;
    r.lo = lo;
    r.hi = hi;
    return r;
}

GC_HD u128 sub128(u128 a, u128 b) {
    u128 r;
    r.lo = a.lo - b.lo;
    r.hi = a.hi - b.hi - (a.lo < b.lo ? 1u : 0u);
    return r;
}

GC_HD u128 add128(u128 a, u128 b) {
    u128 r;
    r.lo = a.lo + b.lo;
    r.hi = a.hi + b.hi + (r.lo < a.lo ? 1u : 0u);
    return r;
}

GC_HD uint64_t mulhi64(uint64_t a, uint64_t b) {
#if defined(__CUDA_ARCH__)
    return __umul64hi(a, b);
#else
    return (uint64_t)(((unsigned __int128)a * b) >> 64);
#endif
}

GC_HD uint32_t mulhi32(uint32_t a, uint32_t b) {
#if defined(__CUDA_ARCH__)
    return __umulhi(a, b);
#else
    return (uint32_t)(((uint64_t)a * b) >> 32);
#endif
}

/* x * m + a (mod 2^128), m < 2^32: four 32x32->64 multiply-adds on 32-bit
 * limbs (one IMAD.WIDE each) instead of 64-bit products */
GC_HD u128 muladd_small(u128 x, uint32_t m, uint32_t a) {
    uint64_t t = (uint64_t)(uint32_t)x.lo * m + a;
    const uint32_t r0 = (uint32_t)t;
    t = (uint64_t)(uint32_t)(x.lo >> 32) * m + (t >> 32);
    const uint32_t r1 = (uint32_t)t;
    t = (uint64_t)(uint32_t)x.hi * m + (t >> 32);
    const uint32_t r2 = (uint32_t)t;
    const uint32_t r3 = (uint32_t)(x.hi >> 32) * m + (uint32_t)(t >> 32);
    u128 r;
    r.lo = (uint64_t)r0 | ((uint64_t)r1 << 32);
    r.hi = (uint64_t)r2 | ((uint64_t)r3 << 32);
    return r;
}

/*
 * Everything a kernel needs to know about one label modulus q.
 * A label's nc digits are cut into chunks of t digits from the least
 * significant end; Q = q^t with q^(t+1) <= 2^32, so that a chunk value fits
 * 32 bits and r/q for r < Q is exact with the 32-bit reciprocal mq.
 * The 128-bit value is divided by Q limb by limb (32-bit limbs) with the
 * Moeller-Granlund 2-by-1 step on the normalised divisor dn = Q << s.
 */
struct ModInfo {
    uint32_t q;        /* modulus */
    uint32_t nc;       /* digits per label, floor(128/log2 q) */
    uint32_t t;        /* digits per chunk */
    uint32_t nchunks;  /* ceil(nc / t) */
    uint32_t Q;        /* q^t */
    uint32_t mq;       /* 2^32/q for powers of two, else floor(2^32 / q) + 1 */
    uint32_t lim;      /* r / q == mulhi(r, mq) for every r <= lim */
    uint32_t top_len;  /* digits in the most significant chunk */
    uint32_t s;        /* clz(Q) */
    uint32_t dn;       /* Q << s */
    uint32_t v;        /* floor((2^64 - 1) / dn) - 2^32 */
    uint32_t limbs;    /* 2 bits per chunk c: (significant 32-bit limbs before division c) - 1 */
    uint32_t pow2s;    /* 1 << s */
    /* word-wise compression of u8 digit planes (four digits per 32-bit word):
     * a chunk is cw whole words, Qw = q^(4 cw) < 2^32; cw = 0 (q^4 >= 2^32,
     * i.e. q = 256) keeps the digit-wise Compressor */
    uint32_t cw;       /* words per chunk */
    uint32_t cwf;      /* words of the most significant chunk: ceil(nc/4) - (chunks - 1) cw, in 1..cw */
    uint32_t q4;       /* q^4 */
    uint32_t Qw;       /* q^(4 cw) */
};

/* r / q for r <= m.lim (true for every chunk value and its quotients). */
GC_HD uint32_t div_q(uint32_t r, const ModInfo& m) { return mulhi32(r, m.mq); }

/* r / q for any 32-bit r */
GC_HD uint32_t div_q_any(uint32_t r, const ModInfo& m) { return r <= m.lim ? mulhi32(r, m.mq) : r / m.q; }

/* a mod q for small a (a <= m.lim): digit sums before reduction. */
GC_HD uint32_t mod_q(uint32_t a, const ModInfo& m) { return a - div_q(a, m) * m.q; }

GC_HD uint32_t funnel_hi(uint32_t lo, uint32_t sh) {
    /* (lo >> (32 - sh)), 0 for sh == 0 */
#if defined(__CUDA_ARCH__)
    return __funnelshift_l(lo, 0u, sh);
#else
    return sh ? (lo >> (32 - sh)) : 0u;
#endif
}

/* One step of the long division by Q: (rs:w) / dn with rs the running
 * remainder in the shifted domain (rem << s).  Returns the quotient limb,
 * updates rs.  Moeller & Granlund, "Improved division by invariant
 * integers", Alg. 4. */
GC_HD uint32_t div_step(uint32_t& rs, uint32_t w, const ModInfo& m) {
    /* (rs:0) + (w << s) as two words; written with multiplies by 2^s so that
     * the shifts run on the FMA pipe (the ALU pipe is the busy one, ncu) */
    const uint32_t u1 = mulhi32(w, m.pow2s) + rs;
    const uint32_t u0 = w * m.pow2s;
    const uint64_t qq = (uint64_t)m.v * u1 + (((uint64_t)u1 << 32) | u0);
    uint32_t q1 = (uint32_t)(qq >> 32) + 1u;
    const uint32_t q0 = (uint32_t)qq;
    uint32_t r = u0 - q1 * m.dn;
    /* the two corrections, branch-free */
    const uint32_t dec = r > q0 ? 1u : 0u;
    q1 -= dec;
    r += dec ? m.dn : 0u;
    const uint32_t inc = r >= m.dn ? 1u : 0u;
    q1 += inc;
    r -= inc ? m.dn : 0u;
    rs = r;
    return q1;
}

/* x = x / Q, returns x mod Q; `limbs` = number of 32-bit limbs of x that can
 * be non-zero (1..4). */
GC_HD uint32_t divmod_Q(uint32_t (&w)[4], const ModInfo& m, uint32_t limbs) {
    uint32_t rs = 0;
    if (limbs > 3) w[3] = div_step(rs, w[3], m);
    if (limbs > 2) w[2] = div_step(rs, w[2], m);
    if (limbs > 1) w[1] = div_step(rs, w[1], m);
    w[0] = div_step(rs, w[0], m);
    return rs >> m.s;
}

GC_HD uint32_t limbs_before(const ModInfo& m, uint32_t c) { return ((m.limbs >> (2 * c)) & 3u) + 1u; }

/*
 * Calls f(d, digit) for d = 0..nc-1 with the base-q digits of x, least
 * significant first -- LabelTensor::decompress.  Digits 0..nc-2 are the
 * exact base-q digits of x for every 128-bit x; the top digit is
 * floor(x / q^(nc-1)) mod q, which equals the reference's int16 arithmetic
 * whenever that quotient is below 2^15 (always for q <= 181, and for every
 * payload of a valid evaluation, where x < q^nc).
 */
template <typename F>
GC_HD void for_each_digit(u128 x, const ModInfo& m, F&& f) {
    uint32_t w[4] = {(uint32_t)x.lo, (uint32_t)(x.lo >> 32), (uint32_t)x.hi, (uint32_t)(x.hi >> 32)};
    uint32_t d = 0;
    for (uint32_t c = 0; c + 1 < m.nchunks; ++c) {
        uint32_t r = divmod_Q(w, m, limbs_before(m, c));
        for (uint32_t i = 0; i < m.t; ++i) {
            const uint32_t qd = div_q(r, m);
            f(d++, r - qd * m.q);
            r = qd;
        }
    }
    /* most significant chunk: what is left of x (fits 32 bits, see
     * DESIGN.md); an out-of-range payload can push it past the reciprocal's
     * exact range, hence div_q_any */
    uint32_t r = w[0];
    for (uint32_t i = 0; i < m.top_len; ++i) {
        const uint32_t qd = div_q_any(r, m);
        f(d++, r - qd * m.q);
        r = qd;
    }
}

/* Same, for two labels of the same modulus in lockstep: f(d, dx, dy). */
template <typename F>
GC_HD void for_each_digit2(u128 x, u128 y, const ModInfo& m, F&& f) {
    uint32_t wx[4] = {(uint32_t)x.lo, (uint32_t)(x.lo >> 32), (uint32_t)x.hi, (uint32_t)(x.hi >> 32)};
    uint32_t wy[4] = {(uint32_t)y.lo, (uint32_t)(y.lo >> 32), (uint32_t)y.hi, (uint32_t)(y.hi >> 32)};
    uint32_t d = 0;
    for (uint32_t c = 0; c + 1 < m.nchunks; ++c) {
        const uint32_t L = limbs_before(m, c);
        uint32_t rx = divmod_Q(wx, m, L);
        uint32_t ry = divmod_Q(wy, m, L);
        for (uint32_t i = 0; i < m.t; ++i) {
            const uint32_t qx = div_q(rx, m), qy = div_q(ry, m);
            f(d++, rx - qx * m.q, ry - qy * m.q);
            rx = qx;
            ry = qy;
        }
    }
    uint32_t rx = wx[0], ry = wy[0];
    for (uint32_t i = 0; i < m.top_len; ++i) {
        const uint32_t qx = div_q_any(rx, m), qy = div_q_any(ry, m);
        f(d++, rx - qx * m.q, ry - qy * m.q);
        rx = qx;
        ry = qy;
    }
}

/*
 * Digit extraction by multiplication, most significant chunk first, for
 * VALID payloads (x < q^nc, which every label of a correct evaluation is).
 * x / q^nc is held as a fixed-point fraction N / 2^160 (five 32-bit limbs,
 * n[4] the most significant); multiplying the fraction by q^len pushes the
 * next `len` digits out of the top as one chunk value.  With D = q^nc,
 * sh = bit length of q and mu = ceil(2^(288-sh) / D) < 2^160,
 *     N0 = floor(x * mu / 2^128) + addc       (fraction N0 / 2^(160-sh)),
 * the first multiplier carries the extra 2^sh that aligns the fraction to
 * 2^160.  The result is exact as long as the fraction over-estimates x/D by
 * less than 1/D: every later value x / q^r has a fractional part that is a
 * multiple of 1/q^r, so an excess below 1/q^r cannot reach the next integer.
 * Sources of error, all accounted for by the host when it builds the plan
 * (make_fracinfo, dashb_api.cu): mu is rounded up (< 1 unit), the columns
 * 0..2 of the product x * mu are not formed (< 4 units), the product is
 * truncated (< 1 unit) and the limbs a stage no longer needs are dropped
 * before its multiplication (stage i keeps the top W_i limbs).  addc makes up
 * for every truncation, so the fraction never under-estimates, and
 * (addc + 1) * D <= 2^(160-sh) bounds the excess.  A long division by Q costs
 * ~13 Moeller-Granlund steps of ~16 instructions per label; this costs 14
 * wide multiplies up front and W_i per stage (15 in total for q = 70), on the
 * FMA pipe instead of the ALU pipe.  tests/test_cabi.py pins it against
 * for_each_digit on the host.
 */
constexpr int FRAC_MAXST = 10;
struct FracInfo {
    uint32_t mu[5];             /* ceil(2^(288-sh) / q^nc), little-endian limbs */
    uint32_t addc;              /* added to N0 (see above) */
    uint32_t nst;               /* stages (chunks); 0 = no plan, use for_each_digit */
    uint32_t mult[FRAC_MAXST];  /* stage s multiplies by q^len_s (stage 0: << sh) */
    uint32_t desc[FRAC_MAXST];  /* len_s | W_s << 8: digits of the chunk, limbs of the fraction the stage keeps (1..5) */
    uint32_t ladder;            /* stages per width, 4 bits each: bits 4(W-1)..4(W-1)+3 = stages that keep W limbs */
    uint32_t words;             /* sum over the stages of ceil(len_s / 2): accumulator words in stage-local pair layout */
    uint32_t t4;                /* 1: five or six stages of four digits (the last one of the six may hold fewer) with
                                 * widths 5,[5,]4,3,2,1 -- the shape of every cast modulus (k+1) m_i in 41..76; the sign
                                 * kernel runs it as straight-line code */
    uint32_t aligned;           /* 1: every stage but the first delivers a multiple of four digits and the first one's range
                                 * starts on a multiple of four (make_fracinfo(aligned4)): stages cover whole plane words */
    uint32_t pad[1];            /* 128 bytes: indexing an array of plans is a shift */
};

GC_HD void frac_init(u128 x, const FracInfo& fi, uint32_t (&n)[5]) {
    const uint32_t x0 = (uint32_t)x.lo, x1 = (uint32_t)(x.lo >> 32), x2 = (uint32_t)x.hi, x3 = (uint32_t)(x.hi >> 32);
#if defined(__CUDA_ARCH__)
    /* Columns 3..8 of x * mu, products of even and odd columns kept in two
     * sets of 64-bit accumulators (e4:e5, e6:e7, e8 and o3:o4, o5:o6, o7:o8)
     * so that every product is one wide multiply-add with the carry chained
     * into the next pair; the sets are added once at the end.  addc rides in
     * on the first product of column 4. */
    const uint32_t m0 = fi.mu[0], m1 = fi.mu[1], m2 = fi.mu[2], m3 = fi.mu[3], m4 = fi.mu[4];
    uint32_t e4, e5, e6, e7, e8;
    asm("{\n\t"
        ".reg .u64 t;\n\t"
        ".reg .u32 o3, o4, o5, o6, o7, o8;\n\t"
        "mul.wide.u32 t, %5, %9;\n\t" /* x3*m0 -> o3:o4 */
        "mov.b64 {o3, o4}, t;\n\t"
        "mul.wide.u32 t, %5, %11;\n\t" /* x3*m2 -> o5:o6 */
        "mov.b64 {o5, o6}, t;\n\t"
        "mul.wide.u32 t, %5, %13;\n\t" /* x3*m4 -> o7:o8 */
        "mov.b64 {o7, o8}, t;\n\t"
        "mul.wide.u32 t, %5, %10;\n\t" /* x3*m1 (+ addc) -> e4:e5 */
        "mov.b64 {%0, %1}, t;\n\t"
        "add.cc.u32 %0, %0, %14;\n\t"
        "addc.u32 %1, %1, 0;\n\t"
        "mul.wide.u32 t, %5, %12;\n\t" /* x3*m3 -> e6:e7 */
        "mov.b64 {%2, %3}, t;\n\t"
        "mov.u32 %4, 0;\n\t"
        /* row x2: m1 (col 3), m3 (col 5) odd; m2 (col 4), m4 (col 6) even */
        "mad.lo.cc.u32 o3, %6, %10, o3;\n\t"
        "madc.hi.cc.u32 o4, %6, %10, o4;\n\t"
        "madc.lo.cc.u32 o5, %6, %12, o5;\n\t"
        "madc.hi.cc.u32 o6, %6, %12, o6;\n\t"
        "addc.cc.u32 o7, o7, 0;\n\t"
        "addc.u32 o8, o8, 0;\n\t"
        "mad.lo.cc.u32 %0, %6, %11, %0;\n\t"
        "madc.hi.cc.u32 %1, %6, %11, %1;\n\t"
        "madc.lo.cc.u32 %2, %6, %13, %2;\n\t"
        "madc.hi.cc.u32 %3, %6, %13, %3;\n\t"
        "addc.u32 %4, %4, 0;\n\t"
        /* row x1: m2 (col 3), m4 (col 5) odd; m3 (col 4) even */
        "mad.lo.cc.u32 o3, %7, %11, o3;\n\t"
        "madc.hi.cc.u32 o4, %7, %11, o4;\n\t"
        "madc.lo.cc.u32 o5, %7, %13, o5;\n\t"
        "madc.hi.cc.u32 o6, %7, %13, o6;\n\t"
        "addc.cc.u32 o7, o7, 0;\n\t"
        "addc.u32 o8, o8, 0;\n\t"
        "mad.lo.cc.u32 %0, %7, %12, %0;\n\t"
        "madc.hi.cc.u32 %1, %7, %12, %1;\n\t"
        "addc.cc.u32 %2, %2, 0;\n\t"
        "addc.cc.u32 %3, %3, 0;\n\t"
        "addc.u32 %4, %4, 0;\n\t"
        /* row x0: m3 (col 3) odd; m4 (col 4) even */
        "mad.lo.cc.u32 o3, %8, %12, o3;\n\t"
        "madc.hi.cc.u32 o4, %8, %12, o4;\n\t"
        "addc.cc.u32 o5, o5, 0;\n\t"
        "addc.cc.u32 o6, o6, 0;\n\t"
        "addc.cc.u32 o7, o7, 0;\n\t"
        "addc.u32 o8, o8, 0;\n\t"
        "mad.lo.cc.u32 %0, %8, %13, %0;\n\t"
        "madc.hi.cc.u32 %1, %8, %13, %1;\n\t"
        "addc.cc.u32 %2, %2, 0;\n\t"
        "addc.cc.u32 %3, %3, 0;\n\t"
        "addc.u32 %4, %4, 0;\n\t"
        /* columns 4..8: even + odd */
        "add.cc.u32 %0, %0, o4;\n\t"
        "addc.cc.u32 %1, %1, o5;\n\t"
        "addc.cc.u32 %2, %2, o6;\n\t"
        "addc.cc.u32 %3, %3, o7;\n\t"
        "addc.u32 %4, %4, o8;\n\t"
        "}"
        : "=&r"(e4), "=&r"(e5), "=&r"(e6), "=&r"(e7), "=&r"(e8)
        : "r"(x3), "r"(x2), "r"(x1), "r"(x0), "r"(m0), "r"(m1), "r"(m2), "r"(m3), "r"(m4), "r"(fi.addc));
    n[0] = e4, n[1] = e5, n[2] = e6, n[3] = e7, n[4] = e8;
#else
    const uint32_t xl[4] = {x0, x1, x2, x3};
    uint32_t r[6] = {0, 0, 0, 0, 0, 0}; /* columns 3..8 of x * mu */
    for (int i = 0; i < 4; ++i) {
        uint32_t carry = 0;
        for (int j = 0; j < 5; ++j) {
            if (i + j < 3) continue;
            const uint64_t t = (uint64_t)xl[i] * fi.mu[j] + r[i + j - 3] + carry;
            r[i + j - 3] = (uint32_t)t;
            carry = (uint32_t)(t >> 32);
        }
        r[i + 2] = carry;
    }
    uint64_t t = (uint64_t)r[1] + fi.addc;
    n[0] = (uint32_t)t;
    t = (uint64_t)r[2] + (t >> 32);
    n[1] = (uint32_t)t;
    t = (uint64_t)r[3] + (t >> 32);
    n[2] = (uint32_t)t;
    t = (uint64_t)r[4] + (t >> 32);
    n[3] = (uint32_t)t;
    n[4] = r[5] + (uint32_t)(t >> 32);
#endif
}

/* fraction *= mult, keeping the top W limbs; returns the integer part */
GC_HD uint32_t frac_stage(uint32_t (&n)[5], uint32_t mult, uint32_t W) {
#if defined(__CUDA_ARCH__)
    /* n'[i] = lo(n[i]*mult) + hi(n[i-1]*mult) + carry: one wide multiply and
     * one add per limb */
    uint32_t chunk;
    if (W > 4) {
        asm("{\n\t"
            ".reg .u32 h0, h1, h2, h3;\n\t"
            "mul.hi.u32 h0, %1, %6;\n\t"
            "mul.lo.u32 %1, %1, %6;\n\t"
            "mul.hi.u32 h1, %2, %6;\n\t"
            "mul.lo.u32 %2, %2, %6;\n\t"
            "mul.hi.u32 h2, %3, %6;\n\t"
            "mul.lo.u32 %3, %3, %6;\n\t"
            "mul.hi.u32 h3, %4, %6;\n\t"
            "mul.lo.u32 %4, %4, %6;\n\t"
            "mul.hi.u32 %0, %5, %6;\n\t"
            "mul.lo.u32 %5, %5, %6;\n\t"
            "add.cc.u32 %2, %2, h0;\n\t"
            "addc.cc.u32 %3, %3, h1;\n\t"
            "addc.cc.u32 %4, %4, h2;\n\t"
            "addc.cc.u32 %5, %5, h3;\n\t"
            "addc.u32 %0, %0, 0;\n\t"
            "}"
            : "=&r"(chunk), "+r"(n[0]), "+r"(n[1]), "+r"(n[2]), "+r"(n[3]), "+r"(n[4])
            : "r"(mult));
    } else if (W > 3) {
        asm("{\n\t"
            ".reg .u32 h1, h2, h3;\n\t"
            "mul.hi.u32 h1, %1, %5;\n\t"
            "mul.lo.u32 %1, %1, %5;\n\t"
            "mul.hi.u32 h2, %2, %5;\n\t"
            "mul.lo.u32 %2, %2, %5;\n\t"
            "mul.hi.u32 h3, %3, %5;\n\t"
            "mul.lo.u32 %3, %3, %5;\n\t"
            "mul.hi.u32 %0, %4, %5;\n\t"
            "mul.lo.u32 %4, %4, %5;\n\t"
            "add.cc.u32 %2, %2, h1;\n\t"
            "addc.cc.u32 %3, %3, h2;\n\t"
            "addc.cc.u32 %4, %4, h3;\n\t"
            "addc.u32 %0, %0, 0;\n\t"
            "}"
            : "=&r"(chunk), "+r"(n[1]), "+r"(n[2]), "+r"(n[3]), "+r"(n[4])
            : "r"(mult));
    } else if (W > 2) {
        asm("{\n\t"
            ".reg .u32 h2, h3;\n\t"
            "mul.hi.u32 h2, %1, %4;\n\t"
            "mul.lo.u32 %1, %1, %4;\n\t"
            "mul.hi.u32 h3, %2, %4;\n\t"
            "mul.lo.u32 %2, %2, %4;\n\t"
            "mul.hi.u32 %0, %3, %4;\n\t"
            "mul.lo.u32 %3, %3, %4;\n\t"
            "add.cc.u32 %2, %2, h2;\n\t"
            "addc.cc.u32 %3, %3, h3;\n\t"
            "addc.u32 %0, %0, 0;\n\t"
            "}"
            : "=&r"(chunk), "+r"(n[2]), "+r"(n[3]), "+r"(n[4])
            : "r"(mult));
    } else if (W > 1) {
        asm("{\n\t"
            ".reg .u32 h3;\n\t"
            "mul.hi.u32 h3, %1, %3;\n\t"
            "mul.lo.u32 %1, %1, %3;\n\t"
            "mul.hi.u32 %0, %2, %3;\n\t"
            "mul.lo.u32 %2, %2, %3;\n\t"
            "add.cc.u32 %2, %2, h3;\n\t"
            "addc.u32 %0, %0, 0;\n\t"
            "}"
            : "=&r"(chunk), "+r"(n[3]), "+r"(n[4])
            : "r"(mult));
    } else {
        chunk = __umulhi(n[4], mult);
        n[4] *= mult;
    }
    return chunk;
#else
    uint64_t t = 0;
    if (W > 4) {
        t = (uint64_t)n[0] * mult;
        n[0] = (uint32_t)t;
        t >>= 32;
    }
    if (W > 3) {
        t += (uint64_t)n[1] * mult;
        n[1] = (uint32_t)t;
        t >>= 32;
    }
    if (W > 2) {
        t += (uint64_t)n[2] * mult;
        n[2] = (uint32_t)t;
        t >>= 32;
    }
    if (W > 1) {
        t += (uint64_t)n[3] * mult;
        n[3] = (uint32_t)t;
        t >>= 32;
    }
    t += (uint64_t)n[4] * mult;
    n[4] = (uint32_t)t;
    return (uint32_t)(t >> 32);
#endif
}

/* Calls f(d0, len, r) once per chunk, most significant chunk first: r holds
 * the digits d0 .. d0+len-1 of x, r = sum_i digit_{d0+i} q^i.  x < q^nc. */
template <typename F>
GC_HD void for_each_chunk_msb(u128 x, const ModInfo& m, const FracInfo& fi, F&& f) {
    uint32_t n[5];
    frac_init(x, fi, n);
    uint32_t d0 = m.nc;
    /* one loop, one copy of f; multiplier, length and width come from the
     * plan's per-stage tables (two loads from the parameter bank) */
#if defined(__CUDA_ARCH__)
#pragma unroll 1
#endif
    for (uint32_t s = 0; s < fi.nst; ++s) {
        const uint32_t desc = fi.desc[s];
        const uint32_t len = desc & 0xffu;
        d0 -= len;
        f(d0, len, frac_stage(n, fi.mult[s], desc >> 8));
    }
}

/* The same walk with the stage width known at compile time: the stages of a
 * plan keep 5, then 4, ... then 1 limbs (never more than the stage before), so
 * five short loops -- one per width, trip counts from fi.ladder -- replace the
 * per-stage width dispatch; every loop body is the multiply chain of its width
 * plus one copy of f.  f(len, r): r holds the next `len` digits below those
 * already delivered (most significant chunk first), r = sum_i digit_i q^i. */
#if defined(__CUDA_ARCH__)
#define GC_LADDER_STEP(W)                                                            \
    _Pragma("unroll 1") for (uint32_t c = (cnt >> (4 * ((W)-1))) & 15u; c; --c, ++s) \
        f(fi.desc[s] & 0xffu, frac_stage(n, fi.mult[s], (W)));
#else
#define GC_LADDER_STEP(W) \
    for (uint32_t c = (cnt >> (4 * ((W)-1))) & 15u; c; --c, ++s) f(fi.desc[s] & 0xffu, frac_stage(n, fi.mult[s], (W)));
#endif
template <typename F>
GC_HD void for_each_chunk_ladder(u128 x, const FracInfo& fi, F&& f) {
    uint32_t n[5];
    frac_init(x, fi, n);
    const uint32_t cnt = fi.ladder;
    uint32_t s = 0;
    GC_LADDER_STEP(5)
    GC_LADDER_STEP(4)
    GC_LADDER_STEP(3)
    GC_LADDER_STEP(2)
    GC_LADDER_STEP(1)
}

/* Colour digit of a compressed label: x mod q, from the residues of the
 * limb weights 2^32, 2^64, 2^96 mod q (a dozen multiply-adds instead of a
 * four-limb long division).  The constants depend on the modulus only. */
struct ColourConsts {
    uint32_t c16, c32, c64, c96; /* 2^16, 2^32, 2^64, 2^96 mod q */
};
GC_HD ColourConsts colour_consts(const ModInfo& m) {
    ColourConsts c;
    c.c16 = mod_q(65536u, m);
    c.c32 = mod_q(c.c16 * c.c16, m);
    c.c64 = mod_q(c.c32 * c.c32, m);
    c.c96 = mod_q(c.c64 * c.c32, m);
    return c;
}
GC_HD uint32_t colour_of(u128 x, const ModInfo& m, const ColourConsts& c) {
    const uint64_t s = (uint64_t)(uint32_t)x.lo + (uint64_t)(uint32_t)(x.lo >> 32) * c.c32 + (uint64_t)(uint32_t)x.hi * c.c64 +
                       (uint64_t)(uint32_t)(x.hi >> 32) * c.c96; /* < 2^42 */
    const uint32_t sl = (uint32_t)s, sh = (uint32_t)(s >> 32);
    const uint32_t r1 = mod_q((sl >> 16) * c.c16 + (sl & 0xffffu), m); /* < 2^24 + 2^16 <= lim */
    return mod_q(sh * c.c32 + r1, m);
}
GC_HD uint32_t colour_of(u128 x, const ModInfo& m) { return colour_of(x, m, colour_consts(m)); }

/*
 * Horner accumulation of digits fed most significant first, chunk-wise:
 * LabelTensor::compress.  Usage: Compressor c(m); push the nc digits from
 * digit nc-1 down to digit 0; x = c.value().
 */
struct Compressor {
    u128 x;
    uint32_t chunk;
    uint32_t left; /* digits still missing in the current chunk */
    const ModInfo& m;
    GC_HD explicit Compressor(const ModInfo& mi) : chunk(0), left(mi.top_len), m(mi) { x = make_u128(0, 0); }
    GC_HD void push(uint32_t digit) {
        chunk = chunk * m.q + digit;
        if (--left == 0) { /* chunk complete */
            x = muladd_small(x, m.Q, chunk);
            chunk = 0;
            left = m.t;
        }
    }
    GC_HD u128 value() const { return x; }
};

/*
 * LabelTensor::compress for a label held as u8 digits, four to a 32-bit word
 * (digit 4w+i in byte i of word w, padding digits zero): words are pushed
 * from the most significant one down; a word's value b0 + q b1 + q^2 b2 +
 * q^3 b3 is two byte dot products (dp4a) and one multiply-add, a chunk of cw
 * words is folded into the 128-bit value at once.  Needs m.cw != 0.  Leading
 * zero digits (the padding of the top word) do not change the value.
 */
GC_HD uint32_t dot4_u8(uint32_t a, uint32_t b) {
#if defined(__CUDA_ARCH__)
    return __dp4a(a, b, 0u);
#else
    return (a & 0xff) * (b & 0xff) + ((a >> 8) & 0xff) * ((b >> 8) & 0xff) + ((a >> 16) & 0xff) * ((b >> 16) & 0xff) +
           (a >> 24) * (b >> 24);
#endif
}
struct WordCompressor {
    u128 x;
    uint32_t chunk, left;
    uint32_t pat, q2;
    const ModInfo& m;
    GC_HD explicit WordCompressor(const ModInfo& mi) : chunk(0), left(mi.cwf), pat(1u | (mi.q << 8)), q2(mi.q * mi.q), m(mi) {
        x = make_u128(0, 0);
    }
    GC_HD void push(uint32_t word) {
        const uint32_t v = dot4_u8(word, pat << 16) * q2 + dot4_u8(word, pat);
        chunk = chunk * m.q4 + v;
        if (--left == 0) {
            x = muladd_small(x, m.Qw, chunk);
            chunk = 0;
            left = m.cw;
        }
    }
    GC_HD u128 value() const { return x; }
};

/* ------------------------------------------------------------------ AES */
/*
 * AES-128 encryption of one block with expanded key rk[44] (big-endian
 * words, FIPS-197), T-table formulation.  te0 has 256 entries
 * (02*S, S, S, 03*S) MSB first; te1..te3 are its byte rotations; sbox the
 * S-box.  All four tables are passed in so the same code runs on the host
 * (plain arrays) and on the device (shared memory).  The block is the
 * compressed label as 16 little-endian bytes in, 16 little-endian bytes out.
 */
GC_HD uint32_t bswap32(uint32_t x) {
#if defined(__CUDA_ARCH__)
    return __byte_perm(x, 0, 0x0123);
#else
    return __builtin_bswap32(x);
#endif
}

/* byte i (0 = least significant) of s, times 4: the T-table byte offset.
 * (Writing the shifts as mul.hi by a power of two moves them to the FMA pipe
 * but measured slower on B200: IMAD.HI is not full rate.) */
GC_HD uint32_t tt_off3(uint32_t s) {
    return (s >> 22) & 0x3fcu;
}
GC_HD uint32_t tt_off2(uint32_t s) {
    return (s >> 14) & 0x3fcu;
}
GC_HD uint32_t tt_off1(uint32_t s) {
    return (s >> 6) & 0x3fcu;
}
GC_HD uint32_t tt_off0(uint32_t s) { return (s << 2) & 0x3fcu; }

template <typename T32, typename T8, typename K32>
GC_HD u128 aes128_encrypt(u128 in, const T32* te0, const T32* te1, const T32* te2, const T32* te3, const T8* sbox,
                          const K32* rk) {
    uint32_t s0 = bswap32((uint32_t)in.lo) ^ rk[0];
    uint32_t s1 = bswap32((uint32_t)(in.lo >> 32)) ^ rk[1];
    uint32_t s2 = bswap32((uint32_t)in.hi) ^ rk[2];
    uint32_t s3 = bswap32((uint32_t)(in.hi >> 32)) ^ rk[3];
    /* rolled on purpose: the nine middle rounds are one ~1 KB loop body that
     * stays in the SM's L0 instruction cache; unrolled (9.6 KB) every call
     * streamed through it (ncu: stall_no_inst 24 %) */
#pragma unroll 1
    for (int r = 1; r < 10; ++r) {
        uint32_t t0 = te0[tt_off3(s0) >> 2] ^ te1[tt_off2(s1) >> 2] ^ te2[tt_off1(s2) >> 2] ^ te3[tt_off0(s3) >> 2] ^ rk[4 * r];
        uint32_t t1 = te0[tt_off3(s1) >> 2] ^ te1[tt_off2(s2) >> 2] ^ te2[tt_off1(s3) >> 2] ^ te3[tt_off0(s0) >> 2] ^ rk[4 * r + 1];
        uint32_t t2 = te0[tt_off3(s2) >> 2] ^ te1[tt_off2(s3) >> 2] ^ te2[tt_off1(s0) >> 2] ^ te3[tt_off0(s1) >> 2] ^ rk[4 * r + 2];
        uint32_t t3 = te0[tt_off3(s3) >> 2] ^ te1[tt_off2(s0) >> 2] ^ te2[tt_off1(s1) >> 2] ^ te3[tt_off0(s2) >> 2] ^ rk[4 * r + 3];
        s0 = t0;
        s1 = t1;
        s2 = t2;
        s3 = t3;
    }
    uint32_t o0 = ((uint32_t)sbox[s0 >> 24] << 24) ^ ((uint32_t)sbox[(s1 >> 16) & 0xff] << 16) ^
                  ((uint32_t)sbox[(s2 >> 8) & 0xff] << 8) ^ (uint32_t)sbox[s3 & 0xff] ^ rk[40];
    uint32_t o1 = ((uint32_t)sbox[s1 >> 24] << 24) ^ ((uint32_t)sbox[(s2 >> 16) & 0xff] << 16) ^
                  ((uint32_t)sbox[(s3 >> 8) & 0xff] << 8) ^ (uint32_t)sbox[s0 & 0xff] ^ rk[41];
    uint32_t o2 = ((uint32_t)sbox[s2 >> 24] << 24) ^ ((uint32_t)sbox[(s3 >> 16) & 0xff] << 16) ^
                  ((uint32_t)sbox[(s0 >> 8) & 0xff] << 8) ^ (uint32_t)sbox[s1 & 0xff] ^ rk[42];
    uint32_t o3 = ((uint32_t)sbox[s3 >> 24] << 24) ^ ((uint32_t)sbox[(s0 >> 16) & 0xff] << 16) ^
                  ((uint32_t)sbox[(s1 >> 8) & 0xff] << 8) ^ (uint32_t)sbox[s2 & 0xff] ^ rk[43];
    u128 out;
    out.lo = (uint64_t)bswap32(o0) | ((uint64_t)bswap32(o1) << 32);
    out.hi = (uint64_t)bswap32(o2) | ((uint64_t)bswap32(o3) << 32);
    return out;
}

}  // namespace gc
