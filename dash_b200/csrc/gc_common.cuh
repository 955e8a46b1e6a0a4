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
