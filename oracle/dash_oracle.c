/*
 * dash_oracle.c -- TEST INFRASTRUCTURE (the parity checker), NOT product.
 *
 * A plain-C, single-file restatement of the online (evaluator-side) half of
 * UzL-ITS/dash, written from the reference's algorithm, operating on DGCM
 * material files (include/dash_gcm_format.h).  Every function cites the
 * reference lines it follows (paths relative to dash/include/).  Nothing in
 * the product path may call into this file; only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline leg do.
 *
 * Parity status: PINNED.  The restatement is checked (tests/test_oracle.py)
 *  - against FIPS-197 Appendix C.1 for the hash's block cipher,
 *  - against the reference itself run here (oracle/_ref, built from
 *    /root/reference by oracle/Makefile) label for label, and
 *  - against the golden vectors under tests/golden/ that the same reference
 *    build generated (tests/golden/make_golden.py).
 *
 * Arithmetic notes that matter for bit-exactness:
 *  - a label of modulus q has nc(q) = floor(128/log2 q) int16 digits, digit 0
 *    is the colour (label_tensor.h:1323-1326,1353-1357);
 *  - compress is Horner in 128 bits, decompress the reference's int16
 *    wrap-around digit extraction (label_tensor.h:715-742), restated
 *    literally so that even out-of-range payloads give identical digits;
 *  - the hash is AES-128 with the fixed key 00..0f over the 16 little-endian
 *    bytes of the compressed label (crypto/cpu_aes_engine.h:23-40).
 */
#include "dash_oracle.h"

#include <omp.h>

#include "dash_gcm_format.h"

typedef unsigned __int128 u128;

/* ------------------------------------------------------------------ AES */
/* FIPS-197 AES-128 encryption, byte oriented.  The reference uses AES-NI
 * (crypto/simple_aes_ni.h: AES_128_Key_Expansion, AES_ECB_encrypt) with the
 * key of cpu_aes_engine.h:23-25; any FIPS-197 implementation is equal. */
static const uint8_t SBOX[256] = {
    0x63, 0x7c, 0x77, 0x7b, 0xf2, 0x6b, 0x6f, 0xc5, 0x30, 0x01, 0x67, 0x2b, 0xfe, 0xd7, 0xab, 0x76, 0xca, 0x82, 0xc9,
    0x7d, 0xfa, 0x59, 0x47, 0xf0, 0xad, 0xd4, 0xa2, 0xaf, 0x9c, 0xa4, 0x72, 0xc0, 0xb7, 0xfd, 0x93, 0x26, 0x36, 0x3f,
    0xf7, 0xcc, 0x34, 0xa5, 0xe5, 0xf1, 0x71, 0xd8, 0x31, 0x15, 0x04, 0xc7, 0x23, 0xc3, 0x18, 0x96, 0x05, 0x9a, 0x07,
    0x12, 0x80, 0xe2, 0xeb, 0x27, 0xb2, 0x75, 0x09, 0x83, 0x2c, 0x1a, 0x1b, 0x6e, 0x5a, 0xa0, 0x52, 0x3b, 0xd6, 0xb3,
    0x29, 0xe3, 0x2f, 0x84, 0x53, 0xd1, 0x00, 0xed, 0x20, 0xfc, 0xb1, 0x5b, 0x6a, 0xcb, 0xbe, 0x39, 0x4a, 0x4c, 0x58,
    0xcf, 0xd0, 0xef, 0xaa, 0xfb, 0x43, 0x4d, 0x33, 0x85, 0x45, 0xf9, 0x02, 0x7f, 0x50, 0x3c, 0x9f, 0xa8, 0x51, 0xa3,
    0x40, 0x8f, 0x92, 0x9d, 0x38, 0xf5, 0xbc, 0xb6, 0xda, 0x21, 0x10, 0xff, 0xf3, 0xd2, 0xcd, 0x0c, 0x13, 0xec, 0x5f,
    0x97, 0x44, 0x17, 0xc4, 0xa7, 0x7e, 0x3d, 0x64, 0x5d, 0x19, 0x73, 0x60, 0x81, 0x4f, 0xdc, 0x22, 0x2a, 0x90, 0x88,
    0x46, 0xee, 0xb8, 0x14, 0xde, 0x5e, 0x0b, 0xdb, 0xe0, 0x32, 0x3a, 0x0a, 0x49, 0x06, 0x24, 0x5c, 0xc2, 0xd3, 0xac,
    0x62, 0x91, 0x95, 0xe4, 0x79, 0xe7, 0xc8, 0x37, 0x6d, 0x8d, 0xd5, 0x4e, 0xa9, 0x6c, 0x56, 0xf4, 0xea, 0x65, 0x7a,
    0xae, 0x08, 0xba, 0x78, 0x25, 0x2e, 0x1c, 0xa6, 0xb4, 0xc6, 0xe8, 0xdd, 0x74, 0x1f, 0x4b, 0xbd, 0x8b, 0x8a, 0x70,
    0x3e, 0xb5, 0x66, 0x48, 0x03, 0xf6, 0x0e, 0x61, 0x35, 0x57, 0xb9, 0x86, 0xc1, 0x1d, 0x9e, 0xe1, 0xf8, 0x98, 0x11,
    0x69, 0xd9, 0x8e, 0x94, 0x9b, 0x1e, 0x87, 0xe9, 0xce, 0x55, 0x28, 0xdf, 0x8c, 0xa1, 0x89, 0x0d, 0xbf, 0xe6, 0x42,
    0x68, 0x41, 0x99, 0x2d, 0x0f, 0xb0, 0x54, 0xbb, 0x16};

static uint8_t g_rk[176];
static int g_rk_ready = 0;

static uint8_t xtime(uint8_t x) { return (uint8_t)((x << 1) ^ ((x >> 7) * 0x1b)); }

static void aes_key_expand(void) {
    static const uint8_t rcon[10] = {0x01, 0x02, 0x04, 0x08, 0x10, 0x20, 0x40, 0x80, 0x1b, 0x36};
    for (int i = 0; i < 16; ++i) g_rk[i] = (uint8_t)i; /* key 00 01 .. 0f */
    for (int i = 4; i < 44; ++i) {
        uint8_t t[4];
        memcpy(t, g_rk + 4 * (i - 1), 4);
        if (i % 4 == 0) {
            uint8_t u = t[0];
            t[0] = SBOX[t[1]] ^ rcon[i / 4 - 1];
            t[1] = SBOX[t[2]];
            t[2] = SBOX[t[3]];
            t[3] = SBOX[u];
        }
        for (int b = 0; b < 4; ++b) g_rk[4 * i + b] = g_rk[4 * (i - 4) + b] ^ t[b];
    }
    g_rk_ready = 1;
}

void oracle_aes128_fixed_key(const uint8_t in[16], uint8_t out[16]) {
    if (!g_rk_ready) {
#pragma omp critical(oracle_rk)
        if (!g_rk_ready) aes_key_expand();
    }
    uint8_t s[16], t[16];
    for (int i = 0; i < 16; ++i) s[i] = in[i] ^ g_rk[i];
    for (int r = 1; r <= 10; ++r) {
        /* SubBytes + ShiftRows: state byte (row,col) lives at s[4*col+row] */
        for (int c = 0; c < 4; ++c)
            for (int rw = 0; rw < 4; ++rw) t[4 * c + rw] = SBOX[s[4 * ((c + rw) & 3) + rw]];
        if (r < 10) {
            for (int c = 0; c < 4; ++c) {
                uint8_t a0 = t[4 * c], a1 = t[4 * c + 1], a2 = t[4 * c + 2], a3 = t[4 * c + 3];
                s[4 * c + 0] = (uint8_t)(xtime(a0) ^ (xtime(a1) ^ a1) ^ a2 ^ a3);
                s[4 * c + 1] = (uint8_t)(a0 ^ xtime(a1) ^ (xtime(a2) ^ a2) ^ a3);
                s[4 * c + 2] = (uint8_t)(a0 ^ a1 ^ xtime(a2) ^ (xtime(a3) ^ a3));
                s[4 * c + 3] = (uint8_t)((xtime(a0) ^ a0) ^ a1 ^ a2 ^ xtime(a3));
            }
        } else
            memcpy(s, t, 16);
        for (int i = 0; i < 16; ++i) s[i] ^= g_rk[16 * r + i];
    }
    memcpy(out, s, 16);
}

/* CPUAESEngine::cipher (crypto/cpu_aes_engine.h:34-40): the __uint128_t is
 * read and written as 16 bytes in memory order (little endian). */
static u128 hash128(u128 x) {
    uint8_t in[16], out[16];
    memcpy(in, &x, 16);
    oracle_aes128_fixed_key(in, out);
    u128 y;
    memcpy(&y, out, 16);
    return y;
}

/* ------------------------------------------------------- label primitives */
int oracle_nr_comps(int q) { return gcm_nr_comps(q); }

/* LabelTensor::compress (label_tensor.h:715-724) */
static u128 compress(const int16_t* d, int nc, int q) {
    u128 x = (u128)(__int128)d[nc - 1];
    for (int i = nc - 1; i > 0; --i) {
        x *= (u128)(__int128)(int16_t)q;
        x += (u128)(__int128)d[i - 1];
    }
    return x;
}

/* LabelTensor::decompress (label_tensor.h:726-742), literally: truncating
 * int16 stores, then the running `sub` correction, then `%=` on the top
 * component. */
static void decompress(u128 x, int q, int nc, int16_t* d) {
    int16_t m = (int16_t)q;
    for (int i = 0; i < nc; ++i) {
        d[i] = (int16_t)x;
        x /= (u128)(__int128)m;
    }
    int16_t sub = 0;
    for (int i = nc - 1; i > 0; --i) {
        sub = (int16_t)(sub - d[i] * m);
        d[i - 1] = (int16_t)(d[i - 1] + sub);
        sub = (int16_t)(sub * m);
    }
    d[nc - 1] = (int16_t)(d[nc - 1] % m);
}

void oracle_compress(int q, const int16_t* digits, uint64_t out2[2]) {
    u128 x = compress(digits, gcm_nr_comps(q), q);
    out2[0] = (uint64_t)x;
    out2[1] = (uint64_t)(x >> 64);
}
void oracle_decompress(int q, const uint64_t in2[2], int16_t* digits) {
    u128 x = ((u128)in2[1] << 64) | in2[0];
    decompress(x, q, gcm_nr_comps(q), digits);
}
void oracle_hash(const uint64_t in2[2], uint64_t out2[2]) {
    u128 y = hash128(((u128)in2[1] << 64) | in2[0]);
    out2[0] = (uint64_t)y;
    out2[1] = (uint64_t)(y >> 64);
}

/* util::modulo (misc/util.h:43-62): always non-negative */
static inline int64_t modulo(int64_t a, int64_t m) { return (a % m + m) % m; }

/* util::mul_inv (misc/util.h:92-103), literally, including the uint16_t
 * divisor variable. */
static uint64_t mul_inv(u128 a, uint16_t b) {
    long long b0 = b, t, q;
    long long x0 = 0, x1 = 1;
    if (b == 1) return 1;
    while (a > 1) {
        q = (long long)(a / b);
        t = b, b = (uint16_t)(a % b), a = (u128)(__int128)t;
        t = x0, x0 = x1 - q * x0, x1 = t;
    }
    if (x1 < 0) x1 += b0;
    return (uint64_t)x1;
}
int64_t oracle_mul_inv(uint64_t a, int b) { return (int64_t)mul_inv(a, (uint16_t)b); }

/* label += / -= label, *= const (label_tensor.h:577-702) */
static void lab_add(int16_t* a, const int16_t* b, int nc, int q) {
    for (int i = 0; i < nc; ++i) a[i] = (int16_t)modulo((int64_t)a[i] + b[i], q);
}
static void lab_sub(int16_t* a, const int16_t* b, int nc, int q) {
    for (int i = 0; i < nc; ++i) a[i] = (int16_t)modulo((int64_t)a[i] - b[i], q);
}
static void lab_mul(int16_t* a, int64_t c, int nc, int q) {
    for (int i = 0; i < nc; ++i) a[i] = (int16_t)modulo((int64_t)a[i] * c, q);
}

/* ProjectionGate::cpu_evaluate (gates/projection_gate.h:89-102) */
static void proj_eval(const u128* ciphers, const int16_t* in, int q_in, int q_out, int16_t* out) {
    u128 key = compress(in, gcm_nr_comps(q_in), q_in);
    u128 h = hash128(key);
    int colour = in[0];
    u128 payload = ciphers[colour] - h;
    decompress(payload, q_out, gcm_nr_comps(q_out), out);
}

/* ProjectionGateMini::cpu_evaluate (gates/projection_gate_mini.h:45-56) */
static int16_t proj_mini_eval(const u128* cipher, const int16_t* in, int q_in) {
    u128 key = compress(in, gcm_nr_comps(q_in), q_in);
    u128 h = hash128(key);
    int colour = in[0];
    int16_t row;
    memcpy(&row, (const char*)cipher + 2 * colour, 2);
    return (int16_t)(row - (int16_t)h);
}

/* ----------------------------------------------------------- the circuit */
struct oracle_circuit {
    gcm_circuit g;
    const int16_t* zero[4096]; /* by modulus */
    int nc[32];
};

typedef struct labels {
    int q[32];
    int nc[32];
    uint64_t n;
    int16_t* d[32];
} labels;

static void labels_alloc(labels* l, int k, const int* q, uint64_t n) {
    l->n = n;
    for (int j = 0; j < k; ++j) {
        l->q[j] = q[j];
        l->nc[j] = gcm_nr_comps(q[j]);
        l->d[j] = (int16_t*)calloc((size_t)(n ? n : 1) * (size_t)l->nc[j], sizeof(int16_t));
    }
}
static void labels_free(labels* l, int k) {
    for (int j = 0; j < k; ++j) free(l->d[j]);
}

oracle_circuit* oracle_load(const char* path) {
    oracle_circuit* c = (oracle_circuit*)calloc(1, sizeof(*c));
    if (!c) return NULL;
    if (gcm_load(&c->g, path) != 0 || c->g.max_modulus >= 4096) {
        free(c);
        return NULL;
    }
    const int16_t* z = (const int16_t*)c->g.zero.data;
    uint64_t off = 0;
    for (int q = 2; q <= c->g.max_modulus; ++q) {
        c->zero[q] = z + off;
        off += (uint64_t)gcm_nr_comps(q);
    }
    for (int j = 0; j < c->g.k; ++j) c->nc[j] = gcm_nr_comps(c->g.crt_base[j]);
    return c;
}
void oracle_free(oracle_circuit* c) {
    if (!c) return;
    gcm_free(&c->g);
    free(c);
}
int oracle_k(const oracle_circuit* c) { return c->g.k; }
int oracle_num_layers(const oracle_circuit* c) { return c->g.n_layers; }
uint64_t oracle_input_size(const oracle_circuit* c) { return c->g.input_size; }

static uint64_t layer_out_size(const gcm_layer* L) {
    switch ((int)L->p[0]) {
        case GCM_DENSE: return (uint64_t)L->p[1];
        case GCM_CONV2D: {
            uint64_t ow = ((uint64_t)L->p[1] - (uint64_t)L->p[5]) / (uint64_t)L->p[7] + 1;
            uint64_t oh = ((uint64_t)L->p[2] - (uint64_t)L->p[6]) / (uint64_t)L->p[8] + 1;
            return ow * oh * (uint64_t)L->p[4];
        }
        case GCM_MAXPOOL2D: {
            uint64_t ow = ((uint64_t)L->p[1] - (uint64_t)L->p[4]) / (uint64_t)L->p[6] + 1;
            uint64_t oh = ((uint64_t)L->p[2] - (uint64_t)L->p[5]) / (uint64_t)L->p[7] + 1;
            return ow * oh * (uint64_t)L->p[3];
        }
        default: return (uint64_t)L->p[1];
    }
}
uint64_t oracle_layer_out_size(const oracle_circuit* c, int layer) {
    if (c->g.n_layers == 0) return c->g.input_size;
    if (layer < 0) layer = c->g.n_layers - 1;
    return layer_out_size(&c->g.layers[layer]);
}
int oracle_layer_out_modulus(const oracle_circuit* c, int layer, int j) {
    if (layer < 0) layer = c->g.n_layers - 1;
    int q = c->g.crt_base[j];
    for (int l = 0; l <= layer; ++l)
        if (c->g.layers[l].p[0] == GCM_PROJECTION) q = (int)c->g.layers[l].p[2 + c->g.k + j];
    return q;
}

/* LabelTensor::matvecmul / matvecmul_tf (label_tensor.h:783-858) followed by
 * `+= bias_label` (layer/garbled_dense.h:136-141). */
static void eval_dense(const oracle_circuit* c, const gcm_layer* L, const labels* in, labels* out, int thr) {
    const int64_t M = L->p[1], K = L->p[2], ch = L->p[3];
    const int64_t* W = (const int64_t*)L->a[0].data;
    for (int j = 0; j < c->g.k; ++j) {
        const int q = in->q[j], nc = in->nc[j];
        const int16_t* z = c->zero[q];
        const int16_t* bias = (const int16_t*)L->a[1 + j].data;
#pragma omp parallel for num_threads(thr)
        for (int64_t r = 0; r < M; ++r) {
            int16_t* o = out->d[j] + r * nc;
            const int64_t* a = W + r * K;
            for (int64_t kk = 0; kk < K; ++kk) {
                if (modulo(a[kk], q) == 0) {
                    for (int i = 0; i < nc; ++i) o[i] = (int16_t)((o[i] + z[i]) % q);
                } else {
                    int64_t idx = kk;
                    if (ch != 0) idx = kk / ch + (kk % ch) * (K / ch);
                    const int16_t* b = in->d[j] + idx * nc;
                    for (int i = 0; i < nc; ++i) o[i] = (int16_t)(((int64_t)o[i] + (int64_t)b[i] * a[kk]) % q);
                }
            }
            lab_add(o, bias + r * nc, nc, q);
        }
    }
}

/* LabelTensor::conv2d (label_tensor.h:920-1002), including the weight index
 * `i * filter_height + j` of line 969. */
static void eval_conv2d(const oracle_circuit* c, const gcm_layer* L, const labels* in, labels* out, int thr) {
    const int64_t iw = L->p[1], ih = L->p[2], ch = L->p[3], fl = L->p[4], fw = L->p[5], fh = L->p[6], sw = L->p[7],
                  sh = L->p[8];
    const int64_t ow = (iw - fw) / sw + 1, oh = (ih - fh) / sh + 1;
    const int64_t osz = ow * oh, fsz = fw * fh, isz = iw * ih;
    const int64_t* we = (const int64_t*)L->a[0].data;
    for (int jm = 0; jm < c->g.k; ++jm) {
        const int q = in->q[jm], nc = in->nc[jm];
        const int16_t* z = c->zero[q];
        const int16_t* bi = (const int16_t*)L->a[1 + jm].data;
        const int16_t* input = in->d[jm];
        int16_t* output = out->d[jm];
#pragma omp parallel for collapse(2) num_threads(thr)
        for (int64_t v = 0; v < fl; ++v)
            for (int64_t l = 0; l < oh; ++l)
                for (int64_t kx = 0; kx < ow; ++kx)
                    for (int m = 0; m < nc; ++m) {
                        int64_t oidx = v * osz * nc + l * ow * nc + kx * nc + m;
                        int64_t acc = output[oidx];
                        for (int64_t w = 0; w < ch; ++w)
                            for (int64_t i = 0; i < fh; ++i)
                                for (int64_t j = 0; j < fw; ++j) {
                                    int64_t wei = we[v * fsz * ch + w * fsz + i * fh + j];
                                    if (modulo(wei, q) == 0)
                                        acc = (acc + z[m]) % q;
                                    else
                                        acc = (acc + wei * input[w * isz * nc + i * iw * nc + j * nc + kx * sw * nc +
                                                                 l * sh * iw * nc + m]) %
                                              q;
                                }
                        output[oidx] = (int16_t)((acc + bi[v * nc + m]) % q);
                    }
    }
}

/* SignGadget::cpu_evaluate (gadgets/sign_gadget.h:583-671) with
 * ApproxResGadget::cpu_evaluate (gadgets/approx_res_gadget.h:94-102).
 * tables: approx, cast1, cast2, sign.  out has n_out moduli. */
static void eval_sign_gadget_at(const oracle_circuit* c, const gcm_array* t, uint64_t g0, uint64_t n, const labels* in,
                                labels* out, int n_out, int thr);
static void eval_sign_gadget(const oracle_circuit* c, const gcm_array* t, uint64_t n, const labels* in, labels* out,
                             int n_out, int thr) {
    eval_sign_gadget_at(c, t, 0, n, in, out, n_out, thr);
}
/* element e of `in` uses the tables of gadget instance g0 + e */
static void eval_sign_gadget_at(const oracle_circuit* c, const gcm_array* t, uint64_t g0, uint64_t n, const labels* in,
                                labels* out, int n_out, int thr) {
    const int k = c->g.k, nm = c->g.n_mrs;
    const int32_t* crt = c->g.crt_base;
    const int32_t* mrs = c->g.mrs_base;
    int S = 0, Mp = 0;
    for (int j = 0; j < k; ++j) S += crt[j];
    for (int i = 1; i < nm; ++i) Mp += mrs[i];
    const u128* approx = (const u128*)t[0].data;
    const u128* cast1 = (const u128*)t[1].data;
    const u128* cast2 = (const u128*)t[2].data;
    const u128* sgn = (const u128*)t[3].data;
#pragma omp parallel for num_threads(thr)
    for (uint64_t e = 0; e < n; ++e) {
        int16_t mrs_label[32][32][128]; /* [crt j][mrs i][digits] */
        int16_t carry[128], sum[128], sum2[128], cast[128];
        /* Step 1: approximate residues in mixed radix (sign_gadget.h:594-601) */
        int prefix = 0;
        for (int j = 0; j < k; ++j) {
            const u128* cj = approx + (g0 + e) * (uint64_t)nm * S + (uint64_t)nm * prefix;
            for (int i = 0; i < nm; ++i)
                proj_eval(cj + (uint64_t)i * crt[j], in->d[j] + e * in->nc[j], crt[j], mrs[i], mrs_label[j][i]);
            prefix += crt[j];
        }
        /* Step 2: mixed-radix addition, least to most significant (:603-654) */
        uint64_t cnt1 = (g0 + e) * (uint64_t)(k + 1) * Mp, cnt2 = cnt1;
        int q_carry = mrs[nm - 1];
        memcpy(carry, c->zero[q_carry], sizeof(int16_t) * gcm_nr_comps(q_carry));
        for (int i = nm - 1; i >= 0; --i) {
            const int m = mrs[i], ncm = gcm_nr_comps(m);
            memcpy(sum, carry, sizeof(int16_t) * ncm);
            for (int j = 0; j < k; ++j) lab_add(sum, mrs_label[j][i], ncm, m);
            if (i == 0) break;
            const int om = (k + 1) * m, nco = gcm_nr_comps(om);
            for (int j = 0; j <= k; ++j) {
                proj_eval(cast1 + cnt1, j < k ? mrs_label[j][i] : carry, m, om, cast);
                cnt1 += m;
                if (j == 0)
                    memcpy(sum2, cast, sizeof(int16_t) * nco);
                else
                    lab_add(sum2, cast, nco, om);
            }
            proj_eval(cast2 + cnt2, sum2, om, mrs[i - 1], carry);
            cnt2 += om;
        }
        /* Step 3: sign of the most significant sum (:657-662) */
        for (int j = 0; j < n_out; ++j)
            proj_eval(sgn + (g0 + e) * (uint64_t)n_out * mrs[0] + (uint64_t)j * mrs[0], sum, mrs[0], out->q[j],
                      out->d[j] + e * out->nc[j]);
    }
}

/* GarbledRelu::cpu_evaluate (layer/garbled_relu.h:181-223) and
 * MixedModHalfGate::cpu_evaluate (gates/mixed_mod_half_gate.h:91-104). */
static void eval_relu_at(const oracle_circuit* c, const gcm_layer* L, uint64_t g0, uint64_t n, const labels* in, labels* out,
                         int thr);
static void eval_relu(const oracle_circuit* c, const gcm_layer* L, const labels* in, labels* out, int thr) {
    eval_relu_at(c, L, 0, (uint64_t)L->p[1], in, out, thr);
}
/* n elements of `in`; element e uses the tables of gadget instance g0 + e
 * (ReluGadget::cpu_evaluate, gadgets/relu_gadget.h, is the same evaluation) */
static void eval_relu_at(const oracle_circuit* c, const gcm_layer* L, uint64_t g0, uint64_t n, const labels* in, labels* out,
                         int thr) {
    const int k = c->g.k;
    labels s;
    int two = 2;
    labels_alloc(&s, 1, &two, n);
    eval_sign_gadget_at(c, &L->a[0], g0, n, in, &s, 1, thr);
    const u128* gar = (const u128*)L->a[4].data;
    const u128* eva = (const u128*)L->a[5].data;
    int S = 0;
    for (int j = 0; j < k; ++j) S += c->g.crt_base[j];
#pragma omp parallel for num_threads(thr)
    for (uint64_t e = 0; e < n; ++e) {
        int prefix = 0;
        for (int j = 0; j < k; ++j) {
            const int p = c->g.crt_base[j], nc = in->nc[j];
            const u128* g_c = gar + (g0 + e) * (uint64_t)S + prefix;
            const u128* e_c = eva + (g0 + e) * (uint64_t)k * 3 + (uint64_t)j * 3;
            const int16_t* x = in->d[j] + e * nc;
            const int16_t* sl = s.d[0] + e * s.nc[0];
            int16_t gl[128], el[128], xs[128];
            proj_eval(g_c, x, p, p, gl);
            proj_eval(e_c, sl, 2, p, el);
            int16_t y_plus_r = proj_mini_eval(e_c + 2, sl, 2);
            memcpy(xs, x, sizeof(int16_t) * nc);
            lab_mul(xs, y_plus_r, nc, p);
            lab_add(el, xs, nc, p);
            lab_sub(el, gl, nc, p);
            memcpy(out->d[j] + e * nc, el, sizeof(int16_t) * nc);
            prefix += p;
        }
    }
    labels_free(&s, 1);
}

/* GarbledMaxPool2d::cpu_evaluate (layer/garbled_maxpool2d.h:143-224) with
 * MaxGadget::cpu_evaluate (gadgets/max_gadget.h:116-160): per output
 * (i, j, channel) the running maximum starts as the window's first input and
 * every kernel position (l, m) -- l outer, (0, 0) included -- applies
 * cur = cur + relu(in(x, y, z) - cur) with its own ReLU gadget, instance
 * o * k_w k_h + l k_h + m of the layer's tables.  This layer addresses
 * labels with the LAST index fastest on both sides: input (x, y, z) of the
 * W x H x C tensor at (x H + y) C + z, output (i, j, channel) at
 * o = (i oh + j) C + channel. */
static void eval_maxpool(const oracle_circuit* c, const gcm_layer* L, const labels* in, labels* out, int thr) {
    const int k = c->g.k;
    const uint64_t W = (uint64_t)L->p[1], H = (uint64_t)L->p[2], C = (uint64_t)L->p[3], kw = (uint64_t)L->p[4],
                   kh = (uint64_t)L->p[5], sw = (uint64_t)L->p[6], sh = (uint64_t)L->p[7];
    const uint64_t ow = (W - kw) / sw + 1, oh = (H - kh) / sh + 1, T = kw * kh;
    (void)thr;
    for (uint64_t o = 0; o < ow * oh * C; ++o) {
        /* both the window's slices (LabelSlice::get_first_label_id_from_slice_dims, label_tensor.h:1574-1579)
         * and set_label(label, {i, j, channel}) (:1290-1301) count the LAST index fastest */
        const uint64_t z = o % C, j = (o / C) % oh, i = o / (C * oh);
        labels cur, diff, res;
        labels_alloc(&cur, k, in->q, 1);
        labels_alloc(&diff, k, in->q, 1);
        labels_alloc(&res, k, in->q, 1);
        const uint64_t first = ((i * sw) * H + j * sh) * C + z;
        for (int m = 0; m < k; ++m) memcpy(cur.d[m], in->d[m] + first * in->nc[m], sizeof(int16_t) * in->nc[m]);
        for (uint64_t l = 0; l < kw; ++l)
            for (uint64_t mm = 0; mm < kh; ++mm) {
                const uint64_t pos = ((i * sw + l) * H + (j * sh + mm)) * C + z;
                for (int m = 0; m < k; ++m) {
                    memcpy(diff.d[m], in->d[m] + pos * in->nc[m], sizeof(int16_t) * in->nc[m]);
                    lab_sub(diff.d[m], cur.d[m], in->nc[m], in->q[m]);
                }
                eval_relu_at(c, L, o * T + l * kh + mm, 1, &diff, &res, 1);
                for (int m = 0; m < k; ++m) lab_add(cur.d[m], res.d[m], in->nc[m], in->q[m]);
            }
        for (int m = 0; m < k; ++m) memcpy(out->d[m] + o * out->nc[m], cur.d[m], sizeof(int16_t) * in->nc[m]);
        labels_free(&cur, k);
        labels_free(&diff, k);
        labels_free(&res, k);
    }
}

/* BaseExtensionGadget::cpu_evaluate (gadgets/base_extension_gadget.h:
 * 197-295) on the k residue labels of all n elements, in place in `lab`
 * (the rescale gadget calls it with in == out). */
static void eval_base_extension(const oracle_circuit* c, const u128* ciph, uint64_t n, const int* extra, int n_extra,
                                labels* lab, int thr) {
    const int k = c->g.k;
    int swapped[32], perm[32], extra_idx[32];
    for (int j = 0; j < k; ++j) perm[j] = j;
    for (int x = 0; x < n_extra; ++x) {
        extra_idx[x] = -1;
        for (int j = 0; j < k; ++j)
            if (c->g.crt_base[j] == extra[x] && extra_idx[x] < 0) extra_idx[x] = j;
    }
    /* swap the i-th last extra modulus to the i-th last position (:206-213,
     * get_out_moduli_swapped :334-345) */
    for (int i = 1; i <= n_extra; ++i) {
        int a = extra_idx[n_extra - i], b = k - i, t = perm[a];
        perm[a] = perm[b];
        perm[b] = t;
    }
    for (int j = 0; j < k; ++j) swapped[j] = c->g.crt_base[perm[j]];
    const int non_ext = k - n_extra;
    /* inverse tables (:347-377) */
    int16_t inv_partial[32][32];
    for (int i = 0; i < k - 1; ++i)
        for (int j = i + 1; j < k; ++j) inv_partial[i][j - i - 1] = (int16_t)mul_inv((u128)swapped[i], (uint16_t)swapped[j]);
    int16_t invv[32];
    for (int x = 0; x < n_extra; ++x) {
        /* std::accumulate with an `int` initial value (:351): the running
         * product is narrowed to int after every step */
        int acc_i = 1;
        for (int j = 0; j < non_ext; ++j) acc_i = (int)((unsigned long)(long)acc_i * (unsigned long)swapped[j]);
        const unsigned long acc = (unsigned long)(long)acc_i;
        int16_t inv_total = (int16_t)mul_inv((u128)acc, (uint16_t)extra[x]);
        invv[x] = (int16_t)mul_inv((u128)(__int128)inv_total, (uint16_t)extra[x]);
    }
    /* cipher offsets of gate (i,j): sequential [i][j][element][m_i] (:120-152) */
    uint64_t base_off[32][32];
    uint64_t cnt = 0;
    for (int i = 0; i < non_ext; ++i)
        for (int j = 0; j < k - i - 1; ++j) {
            base_off[i][j] = cnt;
            cnt += n * (uint64_t)swapped[i];
        }
#pragma omp parallel for num_threads(thr)
    for (uint64_t e = 0; e < n; ++e) {
        int16_t lw[32][128];
        for (int i = 0; i < k; ++i)
            memcpy(lw[i], lab->d[perm[i]] + e * lab->nc[perm[i]], sizeof(int16_t) * lab->nc[perm[i]]);
        for (int i = 0; i < non_ext; ++i)
            for (int j = 0; j < k - i - 1; ++j) {
                const int t = j + i + 1, qt = swapped[t], nct = gcm_nr_comps(qt);
                int16_t pr[128];
                proj_eval(ciph + base_off[i][j] + e * (uint64_t)swapped[i], lw[i], swapped[i], qt, pr);
                lab_sub(lw[t], pr, nct, qt);
                lab_mul(lw[t], inv_partial[i][j], nct, qt);
            }
        /* write back: only the extra residues change (:253-294) */
        for (int x = 0; x < n_extra; ++x) {
            const int ci = extra_idx[x];
            int pos = -1;
            for (int j = 0; j < k; ++j)
                if (swapped[j] == c->g.crt_base[ci] && pos < 0) pos = j;
            int16_t* o = lab->d[ci] + e * lab->nc[ci];
            memcpy(o, lw[pos], sizeof(int16_t) * lab->nc[ci]);
            lab_mul(o, -(int64_t)invv[x], lab->nc[ci], lab->q[ci]);
        }
    }
}

/* RescaleGadget::cpu_evaluate (gadgets/rescale_gadget.h:244-362); `it`
 * selects the iteration's tables for the legacy variant. */
static void eval_rescale_gadget(const oracle_circuit* c, const gcm_layer* L, int legacy, int it, const labels* in,
                                labels* out, int thr) {
    const int k = c->g.k;
    const uint64_t n = (uint64_t)L->p[1];
    int factors[16], nf;
    if (legacy) {
        nf = 1;
        factors[0] = 2;
    } else {
        nf = (int)L->p[2];
        for (int i = 0; i < nf; ++i) factors[i] = (int)L->p[3 + i];
    }
    const int16_t* up = (const int16_t*)c->g.upshift.data;
    /* Step 1: upshift (:261-273) */
    uint64_t uoff = 0;
    for (int j = 0; j < k; ++j) {
        const int nc = in->nc[j];
        for (uint64_t e = 0; e < n; ++e) {
            int16_t* o = out->d[j] + e * nc;
            if (o != in->d[j] + e * nc) memcpy(o, in->d[j] + e * nc, sizeof(int16_t) * nc);
            lab_add(o, up + uoff, nc, in->q[j]);
        }
        uoff += (uint64_t)nc;
    }
    /* Step 2: per factor, project its residue into every other modulus,
     * subtract, multiply by the factor's inverse (:276-313) */
    const u128* tm = (const u128*)L->a[legacy ? k + 5 * it : k].data;
    int ignored[16], n_ign = 0;
    uint64_t cnt = 0;
    for (int i = 0; i < nf; ++i) {
        int fidx = -1;
        for (int j = 0; j < k; ++j)
            if (c->g.crt_base[j] == factors[i] && fidx < 0) fidx = j;
        ignored[n_ign++] = fidx;
        for (int j = 0; j < k; ++j) {
            int skip = 0;
            for (int g = 0; g < n_ign; ++g)
                if (ignored[g] == j) skip = 1;
            if (skip) continue;
            const int q = out->q[j], nc = out->nc[j];
            const int16_t inv = (int16_t)mul_inv((u128)factors[i], (uint16_t)q);
#pragma omp parallel for num_threads(thr)
            for (uint64_t e = 0; e < n; ++e) {
                int16_t pr[128];
                proj_eval(tm + cnt + e * (uint64_t)factors[i], out->d[fidx] + e * out->nc[fidx], factors[i], q, pr);
                int16_t* o = out->d[j] + e * nc;
                lab_sub(o, pr, nc, q);
                lab_mul(o, inv, nc, q);
            }
            cnt += n * (uint64_t)factors[i];
        }
    }
    /* Step 2.3: zero label on the ignored residues (:315-326) */
    for (int g = 0; g < n_ign; ++g) {
        const int j = ignored[g], nc = out->nc[j];
        for (uint64_t e = 0; e < n; ++e) memcpy(out->d[j] + e * nc, c->zero[out->q[j]], sizeof(int16_t) * nc);
    }
    /* Step 3: base extension (:328-346) */
    if (legacy) {
        labels s;
        int two = 2;
        labels_alloc(&s, 1, &two, n);
        eval_sign_gadget(c, &L->a[k + 5 * it + 1], n, out, &s, 1, thr);
        memcpy(out->d[0], s.d[0], sizeof(int16_t) * n * (uint64_t)s.nc[0]);
        labels_free(&s, 1);
    } else {
        eval_base_extension(c, (const u128*)L->a[k + 1].data, n, factors, nf, out, thr);
    }
    /* Step 4: downshift (:348-361) */
    for (int j = 0; j < k; ++j) {
        const int nc = out->nc[j];
        const int16_t* dn = (const int16_t*)L->a[j].data;
        for (uint64_t e = 0; e < n; ++e) lab_sub(out->d[j] + e * nc, dn, nc, out->q[j]);
    }
}

/* GarbledProjection::cpu_evaluate (layer/garbled_projection.h:118-134) */
static void eval_projection(const oracle_circuit* c, const gcm_layer* L, const labels* in, labels* out, int thr) {
    const int k = c->g.k;
    const uint64_t n = (uint64_t)L->p[1];
    for (int j = 0; j < k; ++j) {
        const int qi = (int)L->p[2 + j], qo = (int)L->p[2 + k + j];
        const u128* ciph = (const u128*)L->a[j].data;
#pragma omp parallel for num_threads(thr)
        for (uint64_t e = 0; e < n; ++e)
            proj_eval(ciph + e * (uint64_t)qi, in->d[j] + e * in->nc[j], qi, qo, out->d[j] + e * out->nc[j]);
    }
}

/* GarbledCircuitInterface::cpu_evaluate (garbled_circuit_interface.h:
 * 345-380): the layer loop.  Flatten only changes dims there and has no
 * entry in the material. */
int oracle_evaluate(const oracle_circuit* c, const int16_t* const* in, int upto, int16_t* const* out,
                    int nr_threads) {
    const int k = c->g.k;
    if (nr_threads < 1) nr_threads = 1;
    if (upto < 0) upto = c->g.n_layers - 1;
    labels cur;
    int q0[32];
    for (int j = 0; j < k; ++j) q0[j] = c->g.crt_base[j];
    labels_alloc(&cur, k, q0, c->g.input_size);
    for (int j = 0; j < k; ++j) memcpy(cur.d[j], in[j], sizeof(int16_t) * c->g.input_size * (uint64_t)cur.nc[j]);
    int rc = 0;
    for (int l = 0; l <= upto && l < c->g.n_layers && !rc; ++l) {
        const gcm_layer* L = &c->g.layers[l];
        labels nxt;
        int qn[32];
        for (int j = 0; j < k; ++j) qn[j] = cur.q[j];
        if (L->p[0] == GCM_PROJECTION)
            for (int j = 0; j < k; ++j) qn[j] = (int)L->p[2 + k + j];
        labels_alloc(&nxt, k, qn, layer_out_size(L));
        switch ((int)L->p[0]) {
            case GCM_DENSE: eval_dense(c, L, &cur, &nxt, nr_threads); break;
            case GCM_CONV2D: eval_conv2d(c, L, &cur, &nxt, nr_threads); break;
            case GCM_RELU: eval_relu(c, L, &cur, &nxt, nr_threads); break;
            case GCM_SIGN: eval_sign_gadget(c, &L->a[0], (uint64_t)L->p[1], &cur, &nxt, k, nr_threads); break;
            case GCM_RESCALE: eval_rescale_gadget(c, L, 0, 0, &cur, &nxt, nr_threads); break;
            case GCM_RESCALE_LEGACY:
                /* garbled_rescale.h:88-100: gadget 0 from the inputs, the
                 * following ones in place */
                eval_rescale_gadget(c, L, 1, 0, &cur, &nxt, nr_threads);
                for (int it = 1; it < (int)L->p[2]; ++it) eval_rescale_gadget(c, L, 1, it, &nxt, &nxt, nr_threads);
                break;
            case GCM_PROJECTION: eval_projection(c, L, &cur, &nxt, nr_threads); break;
            case GCM_MAXPOOL2D: eval_maxpool(c, L, &cur, &nxt, nr_threads); break;
            case GCM_BASE_EXTENSION: {
                /* GarbledBaseExtension::cpu_evaluate (layer/garbled_base_extension.h:67-97): the extra residues'
                 * labels are replaced by the layer's start labels (the circuit's input base labels, :74-80),
                 * then the gadget runs on all residues; the others pass through */
                const int nx = (int)L->p[2];
                int extra[16];
                for (int j = 0; j < k; ++j) memcpy(nxt.d[j], cur.d[j], sizeof(int16_t) * cur.n * (uint64_t)cur.nc[j]);
                for (int x = 0; x < nx; ++x) {
                    extra[x] = (int)L->p[3 + x];
                    for (int j = 0; j < k; ++j)
                        if (c->g.crt_base[j] == extra[x]) {
                            memcpy(nxt.d[j], L->a[1 + x].data, sizeof(int16_t) * cur.n * (uint64_t)cur.nc[j]);
                            break;
                        }
                }
                eval_base_extension(c, (const u128*)L->a[0].data, cur.n, extra, nx, &nxt, nr_threads);
                break;
            }
            default: rc = -1;
        }
        labels_free(&cur, k);
        cur = nxt;
    }
    if (!rc)
        for (int j = 0; j < k; ++j) memcpy(out[j], cur.d[j], sizeof(int16_t) * cur.n * (uint64_t)cur.nc[j]);
    labels_free(&cur, k);
    return rc;
}

/* decode_outputs (garbled_circuit_interface.h:422-467) with
 * util::chinese_remainder (misc/util.h:116-141); `prod` there is
 * accumulated in an int (std::accumulate with an int initial value). */
int oracle_decode(const oracle_circuit* c, const int16_t* const* labels_, int64_t* values) {
    const int k = c->g.k;
    if (!c->g.has_dec) return -1;
    const uint64_t n = oracle_layer_out_size(c, -1);
    int16_t* res = (int16_t*)calloc((size_t)k * n, sizeof(int16_t));
    int missing = 0;
    for (int j = 0; j < k; ++j) {
        const int q = c->g.crt_base[j], nc = gcm_nr_comps(q);
        const u128* dec = (const u128*)c->g.dec[j].data;
        for (uint64_t l = 0; l < n; ++l) {
            u128 h = hash128(compress(labels_[j] + l * nc, nc, q));
            int found = 0;
            for (int r = 0; r < q; ++r)
                if (dec[(uint64_t)r * n + l] == h) {
                    res[(uint64_t)j * n + l] = (int16_t)r;
                    found = 1;
                }
            if (!found) ++missing;
        }
    }
    int prod_i = 1;
    long long crt_modulus = 1;
    for (int j = 0; j < k; ++j) {
        prod_i = (int)((u128)(__int128)prod_i * (u128)c->g.crt_base[j]);
        crt_modulus *= c->g.crt_base[j];
    }
    u128 prod = (u128)(__int128)prod_i;
    for (uint64_t l = 0; l < n; ++l) {
        u128 sum = 0;
        for (int j = 0; j < k; ++j) {
            u128 p = prod / (u128)c->g.crt_base[j];
            sum += (((u128)(__int128)res[(uint64_t)j * n + l] * ((u128)mul_inv(p, (uint16_t)c->g.crt_base[j]) % prod)) %
                    prod * p) %
                   prod;
        }
        long long v = (long long)(sum % prod);
        if (v >= crt_modulus / 2) v -= crt_modulus;
        values[l] = v;
    }
    free(res);
    return missing;
}

/* AES blocks the reference evaluator computes per inference (one per
 * projection gate it evaluates). */
double oracle_aes_blocks(const oracle_circuit* c) {
    const int k = c->g.k, nm = c->g.n_mrs;
    double sign_in = (double)k * nm + (double)(k + 1) * (nm > 0 ? nm - 1 : 0) + (nm > 0 ? nm - 1 : 0);
    double total = 0;
    for (int l = 0; l < c->g.n_layers; ++l) {
        const gcm_layer* L = &c->g.layers[l];
        double n = (double)L->p[1];
        switch ((int)L->p[0]) {
            case GCM_RELU: total += n * (sign_in + 1 + 3.0 * k); break;
            case GCM_MAXPOOL2D: total += (double)layer_out_size(L) * (double)(L->p[4] * L->p[5]) * (sign_in + 1 + 3.0 * k); break;
            case GCM_SIGN: total += n * (sign_in + k); break;
            case GCM_RESCALE: {
                int nf = (int)L->p[2];
                double a = 0;
                for (int i = 0; i < nf; ++i) a += k - 1 - i;
                for (int i = 0; i < k - nf; ++i) a += k - i - 1;
                total += n * a;
                break;
            }
            case GCM_RESCALE_LEGACY: total += n * L->p[2] * ((k - 1) + sign_in + 1); break;
            case GCM_PROJECTION: total += n * k; break;
            case GCM_BASE_EXTENSION: {
                const int nx = (int)L->p[2];
                for (int i = 0; i < k - nx; ++i) total += n * (k - i - 1);
                break;
            }
            default: break;
        }
    }
    return total;
}
