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

/* x * m + a (mod 2^128), m < 2^32 */
GC_HD u128 muladd_small(u128 x, uint32_t m, uint32_t a) {
    u128 r;
    uint64_t carry = mulhi64(x.lo, m);
    r.lo = x.lo * m;
    r.hi = x.hi * m + carry;
    uint64_t lo2 = r.lo + a;
    r.hi += (lo2 < r.lo) ? 1u : 0u;
    r.lo = lo2;
    return r;
}

/*
 * Everything a kernel needs to know about one label modulus q.
 * A label's nc digits are cut into chunks of t digits from the least
 * significant end; Q = q^t with q^(t+1) <= 2^32, so that a chunk value fits
 * 32 bits and r/q for r < Q is exact with the 32-bit reciprocal mq.
 */
struct ModInfo {
    uint32_t q;        /* modulus */
    uint32_t nc;       /* digits per label, floor(128/log2 q) */
    uint32_t t;        /* digits per chunk */
    uint32_t nchunks;  /* ceil(nc / t) */
    uint32_t Q;        /* q^t */
    uint32_t mq;       /* floor(2^32 / q) + 1 (0 when q is a power of two) */
    uint32_t shift;    /* log2 q when q is a power of two, else 0 */
    uint32_t top_len;  /* digits in the most significant chunk */
    uint64_t MQ;       /* floor((2^64 - 1) / Q) */
};

/* r / q for r < 2^32 / q (true for every chunk value and its quotients). */
GC_HD uint32_t div_q(uint32_t r, const ModInfo& m) {
    return m.shift ? (r >> m.shift) : mulhi32(r, m.mq);
}

/* a mod q for small a (a < 2^32 / q): digit sums before reduction. */
GC_HD uint32_t mod_q(uint32_t a, const ModInfo& m) {
    return a - div_q(a, m) * m.q;
}

/* x = x / Q, returns x mod Q.  Long division over four 32-bit limbs with the
 * 64-bit reciprocal MQ; limbs known to be zero are skipped by the caller's
 * `limbs` (number of significant 32-bit limbs, 1..4). */
GC_HD uint32_t divmod_Q(u128& x, const ModInfo& m, int limbs = 4) {
    uint32_t w[4] = {(uint32_t)x.lo, (uint32_t)(x.lo >> 32), (uint32_t)x.hi, (uint32_t)(x.hi >> 32)};
    uint64_t rem = 0;
#pragma unroll
    for (int i = 3; i >= 0; --i) {
        if (i >= limbs) continue;
        uint64_t cur = (rem << 32) | w[i];
        uint64_t qd = mulhi64(cur, m.MQ);
        uint64_t r = cur - qd * m.Q;
        while (r >= m.Q) {
            r -= m.Q;
            ++qd;
        }
        w[i] = (uint32_t)qd;
        rem = r;
    }
    x.lo = (uint64_t)w[0] | ((uint64_t)w[1] << 32);
    x.hi = (uint64_t)w[2] | ((uint64_t)w[3] << 32);
    return (uint32_t)rem;
}

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
    uint32_t d = 0;
    for (uint32_t c = 0; c + 1 < m.nchunks; ++c) {
        uint32_t r = divmod_Q(x, m);
        for (uint32_t i = 0; i < m.t; ++i) {
            uint32_t qd = div_q(r, m);
            f(d++, r - qd * m.q);
            r = qd;
        }
    }
    /* most significant chunk: what is left of x (fits 32 bits, see
     * DESIGN.md); plain division because an out-of-range payload can push
     * it past the reciprocal's exact range */
    uint32_t r = (uint32_t)x.lo;
    for (uint32_t i = 0; i + 1 < m.top_len; ++i) {
        uint32_t qd = r / m.q;
        f(d++, r - qd * m.q);
        r = qd;
    }
    f(d, r % m.q);
}

/* Same, for two labels of the same modulus in lockstep: f(d, dx, dy). */
template <typename F>
GC_HD void for_each_digit2(u128 x, u128 y, const ModInfo& m, F&& f) {
    uint32_t d = 0;
    for (uint32_t c = 0; c + 1 < m.nchunks; ++c) {
        uint32_t rx = divmod_Q(x, m);
        uint32_t ry = divmod_Q(y, m);
        for (uint32_t i = 0; i < m.t; ++i) {
            uint32_t qx = div_q(rx, m), qy = div_q(ry, m);
            f(d++, rx - qx * m.q, ry - qy * m.q);
            rx = qx;
            ry = qy;
        }
    }
    uint32_t rx = (uint32_t)x.lo, ry = (uint32_t)y.lo;
    for (uint32_t i = 0; i + 1 < m.top_len; ++i) {
        uint32_t qx = rx / m.q, qy = ry / m.q;
        f(d++, rx - qx * m.q, ry - qy * m.q);
        rx = qx;
        ry = qy;
    }
    f(d, rx % m.q, ry % m.q);
}

/* Colour digit of a compressed label: x mod q (q divides Q). */
GC_HD uint32_t colour_of(u128 x, const ModInfo& m) {
    if (m.shift) return (uint32_t)x.lo & (m.q - 1);
    uint32_t w[4] = {(uint32_t)x.lo, (uint32_t)(x.lo >> 32), (uint32_t)x.hi, (uint32_t)(x.hi >> 32)};
    uint64_t rem = 0;
#pragma unroll
    for (int i = 3; i >= 0; --i) {
        uint64_t cur = (rem << 32) | w[i];
        uint64_t qd = mulhi64(cur, m.MQ);
        uint64_t r = cur - qd * m.Q;
        while (r >= m.Q) r -= m.Q;
        rem = r;
    }
    return mod_q((uint32_t)rem, m);
}

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

template <typename T32, typename T8, typename K32>
GC_HD u128 aes128_encrypt(u128 in, const T32* te0, const T32* te1, const T32* te2, const T32* te3, const T8* sbox,
                          const K32* rk) {
    uint32_t s0 = bswap32((uint32_t)in.lo) ^ rk[0];
    uint32_t s1 = bswap32((uint32_t)(in.lo >> 32)) ^ rk[1];
    uint32_t s2 = bswap32((uint32_t)in.hi) ^ rk[2];
    uint32_t s3 = bswap32((uint32_t)(in.hi >> 32)) ^ rk[3];
#pragma unroll
    for (int r = 1; r < 10; ++r) {
        uint32_t t0 = te0[s0 >> 24] ^ te1[(s1 >> 16) & 0xff] ^ te2[(s2 >> 8) & 0xff] ^ te3[s3 & 0xff] ^ rk[4 * r];
        uint32_t t1 = te0[s1 >> 24] ^ te1[(s2 >> 16) & 0xff] ^ te2[(s3 >> 8) & 0xff] ^ te3[s0 & 0xff] ^ rk[4 * r + 1];
        uint32_t t2 = te0[s2 >> 24] ^ te1[(s3 >> 16) & 0xff] ^ te2[(s0 >> 8) & 0xff] ^ te3[s1 & 0xff] ^ rk[4 * r + 2];
        uint32_t t3 = te0[s3 >> 24] ^ te1[(s0 >> 16) & 0xff] ^ te2[(s1 >> 8) & 0xff] ^ te3[s2 & 0xff] ^ rk[4 * r + 3];
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
