/*
 * dashb_api.cu -- host side of the B200 evaluator: the C ABI declared in
 * include/dash_b200.h, the circuit object (device-resident weights,
 * ciphertext tables, activation planes) and the per-layer launch plan.
 * No label arithmetic happens on the host except the one-off preparation of
 * constants at circuit-build time (weights mod q, constant labels, modular
 * inverses) and the dashb_hostcheck_* self-test hooks.
 */
#include <cuda_runtime.h>

#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <string>
#include <chrono>
#include <vector>

#include "../../include/dash_b200.h"
#include "../../include/dash_b200_gcm.h"
#include "kernels.h"

using gc::ModInfo;
using gc::u128;
using namespace dashb;

namespace {

thread_local std::string g_err;
int g_device = -1;
cudaStream_t g_stream = nullptr;
int g_load_replicas = 1; /* consumed by dashb_circuit_create when dashb_circuit_load_replicated drives it */

int fail(int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    g_err = buf;
    return code;
}

#define CU(expr)                                                                              \
    do {                                                                                      \
        cudaError_t _e = (expr);                                                              \
        if (_e != cudaSuccess)                                                                \
            return fail(_e == cudaErrorMemoryAllocation ? DASHB_ENOMEM : DASHB_ECUDA, "%s: %s", #expr, \
                        cudaGetErrorString(_e));                                              \
    } while (0)

int nr_comps(int q) { return (int)floor(128 / log2((double)q)); }

ModInfo make_modinfo(uint32_t q) {
    ModInfo m;
    memset(&m, 0, sizeof(m));
    m.q = q;
    m.nc = (uint32_t)nr_comps((int)q);
    /* t = max digits per chunk with q^(t+1) <= 2^32 */
    uint64_t pw = q;
    uint32_t t = 0;
    while (pw * q <= (1ull << 32) && t < m.nc) {
        pw *= q;
        ++t;
    }
    if (t == 0) t = 1;
    if (t > m.nc) t = m.nc;
    m.t = t;
    uint64_t Q = 1;
    for (uint32_t i = 0; i < t; ++i) Q *= q;
    m.Q = (uint32_t)Q;
    m.nchunks = (m.nc + t - 1) / t;
    m.top_len = m.nc - (m.nchunks - 1) * t;
    if ((q & (q - 1)) == 0) {
        m.mq = (uint32_t)((1ull << 32) / q);
        m.lim = 0xffffffffu;
    } else {
        m.mq = (uint32_t)((1ull << 32) / q) + 1;
        m.lim = 0xffffffffu / q;
    }
    m.s = (uint32_t)__builtin_clz(m.Q);
    m.dn = m.Q << m.s;
    m.pow2s = 1u << m.s;
    m.v = (uint32_t)(~0ull / m.dn - (1ull << 32));
    /* significant limbs of the running quotient before division c: it is
     * below 2^128 / Q^c for every 128-bit payload */
    unsigned __int128 bound = ~(unsigned __int128)0;
    m.limbs = 0;
    for (uint32_t c = 0; c + 1 < m.nchunks && c < 16; ++c) {
        uint32_t l = 1;
        for (int i = 3; i >= 1; --i)
            if ((uint32_t)(bound >> (32 * i)) != 0) {
                l = (uint32_t)i + 1;
                break;
            }
        m.limbs |= (l - 1) << (2 * c);
        bound /= Q;
    }
    /* word-wise compression: whole words of four digits per chunk, q^(4 cw) < 2^32 */
    if ((uint64_t)q * q * q * q <= 0xffffffffull) {
        m.q4 = q * q * q * q;
        uint64_t pw = m.q4;
        m.cw = 1;
        while (pw * m.q4 <= 0xffffffffull) {
            pw *= m.q4;
            ++m.cw;
        }
        m.Qw = (uint32_t)pw;
        const uint32_t nw = (m.nc + 3) / 4;
        m.cwf = nw - ((nw - 1) / m.cw) * m.cw;
    }
    return m;
}

/* ---- plan of the multiplicative digit extraction (gc::FracInfo) ---- */
namespace {
/* little-endian 32-bit limbs, enough for 2^288 */
struct Big {
    uint32_t w[10];
    Big() { memset(w, 0, sizeof(w)); }
    void mul_small(uint32_t m) {
        uint64_t c = 0;
        for (int i = 0; i < 10; ++i) {
            c += (uint64_t)w[i] * m;
            w[i] = (uint32_t)c;
            c >>= 32;
        }
    }
    void shl1(uint32_t bit) {
        for (int i = 9; i > 0; --i) w[i] = (w[i] << 1) | (w[i - 1] >> 31);
        w[0] = (w[0] << 1) | bit;
    }
    void set_bit(int b) { w[b >> 5] |= 1u << (b & 31); }
    bool is_zero() const {
        for (int i = 0; i < 10; ++i)
            if (w[i]) return false;
        return true;
    }
    static int cmp(const Big& a, const Big& b) {
        for (int i = 9; i >= 0; --i)
            if (a.w[i] != b.w[i]) return a.w[i] < b.w[i] ? -1 : 1;
        return 0;
    }
    void sub(const Big& b) {
        uint64_t br = 0;
        for (int i = 0; i < 10; ++i) {
            const uint64_t d = (uint64_t)w[i] - b.w[i] - br;
            w[i] = (uint32_t)d;
            br = (d >> 32) & 1u;
        }
    }
};
}  // namespace

/* aligned4: every stage but the first delivers a multiple of four digits and
 * the first one nc mod 4 more than that, so that the stages' digit ranges
 * start on the 32-bit words of a u8 digit plane (the half gates of k_sign
 * write their output labels word by word). */
gc::FracInfo make_fracinfo(const ModInfo& m, bool aligned4 = false) {
    gc::FracInfo fi;
    memset(&fi, 0, sizeof(fi));
    const uint32_t q = m.q, nc = m.nc;
    uint32_t t = m.t;
    if (aligned4) {
        t &= ~3u;
        if (t == 0) return fi;
    }
    if (q < 2 || nc < 2) return fi;
    /* DASHB_NO_FRAC: no plans, every kernel takes its long-division path
     * (the one moduli with more than FRAC_MAXST chunks take; tests/
     * test_gpu_parity.py::test_division_fallback keeps it covered) */
    if (getenv("DASHB_NO_FRAC")) return fi;
    uint32_t sh = 0;
    while ((q >> sh) != 0) ++sh; /* bit length: q < 2^sh */
    const int F = 160 - (int)sh;
    Big D;
    D.w[0] = 1;
    for (uint32_t i = 0; i < nc; ++i) D.mul_small(q);
    /* mu = ceil(2^(288-sh) / D), bit by bit */
    Big rem, mu;
    const int E = 288 - (int)sh;
    for (int b = E; b >= 0; --b) {
        rem.shl1(b == E ? 1u : 0u);
        if (Big::cmp(rem, D) >= 0) {
            rem.sub(D);
            mu.set_bit(b);
        }
    }
    if (!rem.is_zero()) {
        /* mu += 1 */
        uint64_t c = 1;
        for (int i = 0; i < 10 && c; ++i) {
            c += mu.w[i];
            mu.w[i] = (uint32_t)c;
            c >>= 32;
        }
    }
    for (int i = 5; i < 10; ++i)
        if (mu.w[i]) return fi; /* cannot happen: mu < q * 2^(160-sh) + 1 <= 2^160 */
    /* first chunk: as many digits as fit next to the 2^sh alignment factor */
    uint32_t a = 0;
    uint64_t pa = 1;
    while (a < t && a < nc && ((pa * q) << sh) <= 0xffffffffull) {
        pa *= q;
        ++a;
    }
    if (aligned4) { /* largest a' <= a with a' = nc (mod 4) */
        const uint32_t want = nc & 3u;
        while (a > 0 && (a & 3u) != want) {
            pa /= q;
            --a;
        }
    }
    if (a == 0 || a >= nc) return fi;
    const uint32_t rest = nc - a;
    const uint32_t nmid = rest / t, rl = rest % t;
    const uint32_t nst = 1 + nmid + (rl ? 1 : 0);
    if (nst < 2 || nst > (uint32_t)gc::FRAC_MAXST) return fi;
    uint64_t plast = 1, qt = 1;
    for (uint32_t i = 0; i < (rl ? rl : t); ++i) plast *= q;
    for (uint32_t i = 0; i < t; ++i) qt *= q;
    uint32_t lens[10];
    long double mults[10];
    lens[0] = a;
    mults[0] = (long double)(pa << sh);
    for (uint32_t s = 1; s < nst; ++s) {
        lens[s] = t;
        mults[s] = (long double)qt;
    }
    lens[nst - 1] = rl ? rl : t;
    mults[nst - 1] = (long double)plast;
    const long double lq = log2l((long double)q);
    /* the dropped limbs' loss for a width ladder, the compensation addc and the exact excess bound
     * (addc + 1) * D <= 2^F; fills the plan when the ladder passes */
    auto try_ladder = [&](const uint32_t* W) -> bool {
        long double tot = 0, amp = 1;
        for (uint32_t s = 0; s < nst; ++s) {
            if (W[s] < 5) tot += ldexpl(1.0L, 32 * (5 - (int)W[s])) / amp;
            amp *= mults[s];
        }
        const long double addc_ld = 5 + ceill(tot * (1 + 1e-12L)) + 2;
        if (addc_ld >= 2147483648.0L) return false;
        const uint32_t addc = (uint32_t)addc_ld;
        Big lhs = D, rhs;
        lhs.mul_small(addc + 1);
        rhs.set_bit(F);
        if (Big::cmp(lhs, rhs) > 0) return false;
        for (int i = 0; i < 5; ++i) fi.mu[i] = mu.w[i];
        fi.addc = addc;
        fi.nst = nst;
        fi.aligned = aligned4 ? 1u : 0u;
        fi.ladder = fi.words = 0;
        for (uint32_t s = 0; s < nst; ++s) {
            fi.mult[s] = (uint32_t)mults[s];
            fi.desc[s] = lens[s] | W[s] << 8;
            fi.ladder += 1u << (4 * (W[s] - 1));
            fi.words += (lens[s] + 1) / 2;
        }
        return true;
    };
    /* the straight-line shape of k_sign (FracInfo::t4): stages of four digits, widths 5,[5,]4,3,2,1 */
    if (!aligned4 && (nst == 5 || nst == 6) && lens[nst - 1] <= 4) {
        bool four = true;
        for (uint32_t s = 0; s + 1 < nst; ++s) four = four && lens[s] == 4;
        const uint32_t W5[5] = {5, 4, 3, 2, 1}, W6[6] = {5, 5, 4, 3, 2, 1};
        if (four && try_ladder(nst == 5 ? W5 : W6)) {
            fi.t4 = 1;
            return fi;
        }
    }
    for (int guard = 6; guard < 40; ++guard) {
        uint32_t W[10];
        uint32_t remd = nc;
        for (uint32_t s = 0; s < nst; ++s) {
            const long double need = lq * remd + guard;
            uint32_t w = (uint32_t)ceill(need / 32);
            if (w < 1) w = 1;
            if (w > 5) w = 5;
            if (s > 0 && w > W[s - 1]) w = W[s - 1];
            W[s] = w;
            remd -= lens[s];
        }
        if (try_ladder(W)) return fi;
    }
    memset(&fi, 0, sizeof(fi));
    return fi;
}

/* util::mul_inv (misc/util.h:92-103) */
uint64_t mul_inv(unsigned __int128 a, uint16_t b) {
    long long b0 = b, t, q;
    long long x0 = 0, x1 = 1;
    if (b == 1) return 1;
    while (a > 1) {
        if (b == 0) return 0; /* not invertible: rejected by the callers */
        q = (long long)(a / b);
        t = b, b = (uint16_t)(a % b), a = (unsigned __int128)(__int128)t;
        t = x0, x0 = x1 - q * x0, x1 = t;
    }
    if (x1 < 0) x1 += b0;
    return (uint64_t)x1;
}

struct DevMem {
    void* p{nullptr};
    size_t bytes{0};
    DevMem() = default;
    DevMem(const DevMem&) = delete;
    DevMem& operator=(const DevMem&) = delete;
    ~DevMem() {
        if (p) cudaFree(p);
    }
    cudaError_t alloc(size_t n) {
        if (p) cudaFree(p);
        p = nullptr;
        bytes = n;
        return cudaMalloc(&p, n ? n : 16);
    }
    cudaError_t upload(const void* src, size_t n) {
        cudaError_t e = alloc(n);
        if (e != cudaSuccess) return e;
        /* cudaMemcpyDefault: the source may also be device memory (tables
         * that the GPU garbler left on the device) */
        if (n) e = cudaMemcpy(p, src, n, cudaMemcpyDefault);
        /* The library's streams are non-blocking: nothing orders them behind
         * the NULL stream this copy ran on, and cudaMemcpy may return while a
         * pageable staging DMA or a device-to-device copy is still in flight. */
        if (e == cudaSuccess && n) e = cudaStreamSynchronize(cudaStreamLegacy);
        return e;
    }
};

/* int16 label digits -> u8, zero padded to a multiple of 4 digits */
std::vector<uint8_t> to_u8(const int16_t* d, size_t n) {
    std::vector<uint8_t> v((n + 3) & ~(size_t)3, 0);
    for (size_t i = 0; i < n; ++i) v[i] = (uint8_t)d[i];
    return v;
}

u128 compress_host(const int16_t* d, const ModInfo& m) {
    gc::Compressor c(m);
    for (int i = (int)m.nc - 1; i >= 0; --i) c.push((uint32_t)d[i]);
    return c.value();
}

struct Tensor {
    uint64_t n{0};
    int k{0};
    int q[MAXK];
    uint8_t* plane[MAXK];
};

size_t tensor_bytes(uint64_t n, int k, const int* q, uint32_t Bp) {
    size_t total = 0;
    for (int j = 0; j < k; ++j) total += ((size_t)n * ((nr_comps(q[j]) + 3) & ~3) * Bp + 255) & ~(size_t)255;
    return total;
}

void carve(Tensor& t, uint8_t* base, uint64_t n, int k, const int* q, uint32_t Bp) {
    t.n = n;
    t.k = k;
    size_t off = 0;
    for (int j = 0; j < k; ++j) {
        t.q[j] = q[j];
        t.plane[j] = base + off;
        off += ((size_t)n * ((nr_comps(q[j]) + 3) & ~3) * Bp + 255) & ~(size_t)255;
    }
}

struct Layer {
    int type{0};
    uint64_t in_size{0}, out_size{0};
    int out_q[MAXK];
    /* linear */
    uint64_t M{0}, K{0};
    uint32_t Kp{0}, npix{1}, ow{1}, step_x{0}, step_y{0};
    DevMem w4[MAXK], cst[MAXK], koff;
    /* tensor-core dense path: row-major u8 weights [M][Kp16] and their TMA maps */
    DevMem wrm[MAXK];
    TensorMapBlob map_a[MAXK];
    bool use_tc{false};
    /* tensor-core convolution: geometry for the im2col TMA box, cb channels per 128-column K block */
    uint32_t cv_w{0}, cv_h{0}, cv_c{0}, cv_fw{0}, cv_fh{0}, cv_sw{0}, cv_sh{0}, cv_cb{0}, tc_nkb{0};
    /* strip convolution (k_conv_tc_strip): weights in (channel block, filter column, filter)-major rows of 128 bytes */
    uint32_t st_strip{0}, st_rows{0}, st_cblk{0}, st_ncb{0}, st_wboxes{0};
    DevMem wst[MAXK];
    TensorMapBlob map_st[MAXK];
    /* gadget tables */
    DevMem approx, cast1, cast2, sign, garbler, evaluator;
    /* rescale */
    int nf{0}, l_iter{0};
    DevMem down[MAXK], tm, be;
    DevMem start[MAXK]; /* base extension layer: per-element start labels of the extra residues (u8 [n][4*nc4]) */
    std::vector<std::unique_ptr<DevMem>> it_tables; /* legacy: 5 per iteration */
    RescaleParams rp;
    /* projection */
    int in_q[MAXK];
    DevMem proj[MAXK];
    /* max pooling */
    uint32_t pl_w{0}, pl_h{0}, pl_c{0}, pl_kw{0}, pl_kh{0}, pl_sw{0}, pl_sh{0}, pl_ow{0}, pl_oh{0};
    /* sharding: output elements [e0, e1) / linear rows [v0, v1) of this rank; or, a convolution sharded by
     * output pixels [p0, p1) (peer-memory exchange): the elements are seg = p1 - p0 pixels of each of the
     * nseg filters, run s starting at element e0 + s * stride (stride = pixels per filter) */
    uint64_t e0{0}, e1{0};
    uint32_t v0{0}, v1{0};
    uint32_t p0{0}, p1{0};
    uint64_t seg{0}, stride{0}, nseg{0};
    uint64_t slice_count() const { return seg ? seg * nseg : e1 - e0; }
    /* accounting */
    double label_bytes{0}, weight_bytes{0}, aes{0}, ref_aes{0};
    float ms{0};
};

/* One sub-batch pipeline: its own stream, staging and activation buffers.
 * A batch is cut into contiguous slices, one per lane; the lanes' copies and
 * kernels overlap (H2D of lane i+1 under the kernels of lane i, the small
 * late layers of one lane under the big early layers of another). */
struct Lane {
    cudaStream_t s{nullptr};
    cudaEvent_t done{nullptr}, in_done{nullptr}, ing_done{nullptr};
    cudaEvent_t in_done_x[3] = {nullptr, nullptr, nullptr}; /* the lane's copies on the further copy streams */
    DevMem buf_in, buf_a, buf_b, stage_in, stage_out, xvals;
    DevMem buf_in2; /* dashb_evaluate_async, in-order mode: the input planes of the step after the one computing */
    DevMem pool; /* max pooling: the ReLU gadget's input (y - x) and output, two tensors of the layer's output size */
    /* output-channel sharding over peer memory: two gather buffers, the epoch flags, and every rank's copies of them */
    DevMem gat[2], flags;
    uint8_t* peer_gat[2][MAXR] = {};
    uint32_t* peer_flags[MAXR] = {};
    int gcur{0};
    uint32_t epoch{0};
    Tensor t_in, t_cur;
    int which{0};
    uint32_t b0{0}, batch{0}, Bp{0};
};

}  // namespace

struct dashb_circuit {
    int k{0}, n_mrs{0}, max_modulus{0};
    int crt[MAXK];
    int mrs[MAXM];
    ModInfo crt_mi[MAXK], mrs_mi[MAXM], cast_mi[MAXM];
    gc::FracInfo crt_fi[MAXK], cast_fi[MAXM]; /* multiplicative digit extraction plans */
    gc::FracInfo crt_fa[MAXK];                /* ... with word-aligned stages (half gates of k_sign) */
    uint64_t input_size{0};
    std::vector<std::vector<int16_t>> zero; /* by modulus-2 */
    DevMem d_zero[MAXK], d_up[MAXK];
    DevMem d_nolabel; /* 512 zero bytes: "no upshift" for the standalone base extension */
    bool has_up{false};
    u128 carry0;
    std::vector<std::unique_ptr<Layer>> layers;
    int cur_q[MAXK];      /* moduli of the labels leaving the last added layer */
    uint64_t cur_size{0}; /* their count */
    bool finalized{false};
    uint64_t max_batch{0};
    uint32_t batch{0};
    std::vector<std::unique_ptr<Lane>> lanes;
    int n_lanes{0};
    int lanes_used{0};
    /* dashb_evaluate_async: a second set of lanes so that two steps are in
     * flight (step n+1's inputs cross PCIe while step n computes) */
    std::vector<std::unique_ptr<Lane>> lanes_alt;
    cudaEvent_t async_done[2] = {nullptr, nullptr};
    cudaStream_t join_s{nullptr}; /* waits for every lane's end of step without holding a lane back */
    /* DASHB_TRACE_ASYNC: per step the host time of the call and, per lane, events after inputs / layers /
     * outputs; printed relative to the first step's enqueue at dashb_evaluate_wait (experiments only) */
    struct AsyncTrace {
        int step, set, lanes;
        double host_in_ms, host_out_ms;
        cudaEvent_t ev[1 + 4 * 8];
    };
    std::vector<AsyncTrace> atrace;
    cudaEvent_t atrace_base{nullptr};
    std::chrono::steady_clock::time_point atrace_t0;
    int atrace_steps{0};
    bool async_busy[2] = {false, false};
    int async_next{0};
    int shard_rank{0}, shard_world{1};
    int shard_lanes{0}; /* > 0: peer-memory exchange with that many sub-batch lanes (dashb_set_shard_lanes) */
    int table_replicas{1}; /* dashb_set_table_replicas: the sign/ReLU tables exist this many times, interleaved by row */
    std::vector<void*> ipc_opened;
    uint32_t lane_cap{0};
    std::vector<uint32_t> lane_weights; /* DASHB_LANE_SPLIT */
    cudaEvent_t ev_fork{nullptr};
    cudaStream_t copy_s{nullptr};   /* H2D copies of the pipelined dashb_evaluate */
    static constexpr int MAX_COPY = 4;
    cudaStream_t copy_x[MAX_COPY - 1] = {nullptr, nullptr, nullptr}; /* further copy streams (DASHB_COPY_STREAMS) */
    int n_copy{1};
    cudaStream_t ingest_s{nullptr}; /* its ingest kernels (highest priority) */
    DevMem dec[MAXK];
    bool has_dec{false};
    DevMem in_base[MAXK], in_off[MAXK];
    bool has_ginp{false};
    DevMem d_values, d_ok;
    uint64_t out_n{0};   /* labels per inference in the lanes' current tensors */
    int out_q[MAXK];     /* their moduli */
    uint64_t launches{0};
    int timing{0}; /* 0 off, 1 per-layer events with the lanes serialised, 2 per-(lane, layer) events in the concurrent step */
    std::vector<cudaEvent_t> ev;
    std::vector<cudaEvent_t> iv;  /* mode 2: [lane][layer][begin, end] */
    std::vector<float> iv_ms;     /* ... as milliseconds since the step began */
    cudaEvent_t iv_t0{nullptr};
    ~dashb_circuit() {
        for (auto e : ev) cudaEventDestroy(e);
        for (auto e : iv) cudaEventDestroy(e);
        if (iv_t0) cudaEventDestroy(iv_t0);
        if (ev_fork) cudaEventDestroy(ev_fork);
        if (copy_s) cudaStreamDestroy(copy_s);
        for (auto x : copy_x)
            if (x) cudaStreamDestroy(x);
        if (ingest_s) cudaStreamDestroy(ingest_s);
        if (join_s) cudaStreamDestroy(join_s);
        for (auto e : async_done)
            if (e) cudaEventDestroy(e);
        for (void* p : ipc_opened) cudaIpcCloseMemHandle(p);
        for (auto& l : lanes_alt) lanes.push_back(std::move(l));
        for (auto& l : lanes) {
            if (l->done) cudaEventDestroy(l->done);
            if (l->in_done) cudaEventDestroy(l->in_done);
            for (auto x : l->in_done_x)
                if (x) cudaEventDestroy(x);
            if (l->ing_done) cudaEventDestroy(l->ing_done);
            if (l->s) cudaStreamDestroy(l->s);
        }
    }
};

namespace {

int check_building(dashb_circuit* c) {
    if (!c) return fail(DASHB_EINVAL, "null circuit");
    if (c->finalized) return fail(DASHB_ESTATE, "circuit already finalized");
    return DASHB_OK;
}

int check_all_crt(dashb_circuit* c, const char* what) {
    for (int j = 0; j < c->k; ++j)
        if (c->cur_q[j] != c->crt[j]) return fail(DASHB_EINVAL, "%s needs its input labels in the CRT base", what);
    return DASHB_OK;
}

/* Shared preparation of dense / conv2d: packs W mod q_j into the dp4a
 * layout, builds the constant label (bias + zero-weight rule) per row. */
int prep_linear(dashb_circuit* c, Layer& L, const std::vector<int64_t>& wrow /* [M][K] */, uint64_t M, uint64_t K,
                const int16_t* const* bias_labels) {
    if (K > 60000) return fail(DASHB_EINVAL, "linear layer with K=%llu taps exceeds the exact u32 range", (unsigned long long)K);
    L.M = M;
    L.K = K;
    L.Kp = (uint32_t)((K + 3) & ~3ull);
    const uint32_t k4 = L.Kp / 4, nvg = (uint32_t)((M + 7) / 8);
    for (int j = 0; j < c->k; ++j) {
        const uint32_t q = (uint32_t)c->crt[j], nc = c->crt_mi[j].nc;
        std::vector<uint32_t> w4((size_t)nvg * k4 * 8, 0);
        const uint32_t ncp = (nc + 3) & ~3u; /* rows padded to whole words */
        std::vector<uint8_t> cst((size_t)M * ncp, 0);
        const int16_t* z = c->zero[q - 2].data();
        for (uint64_t r = 0; r < M; ++r) {
            uint64_t nz = 0;
            for (uint64_t kk = 0; kk < K; ++kk) {
                int64_t w = wrow[r * K + kk] % (int64_t)q;
                if (w < 0) w += q;
                if (w == 0) ++nz;
                w4[((size_t)(r / 8) * k4 + kk / 4) * 8 + (r % 8)] |= (uint32_t)w << (8 * (kk % 4));
            }
            for (uint32_t d = 0; d < nc; ++d)
                cst[r * ncp + d] = (uint8_t)(((uint64_t)bias_labels[j][r * nc + d] + (nz % q) * (uint64_t)z[d]) % q);
        }
        CU(L.w4[j].upload(w4.data(), w4.size() * 4));
        CU(L.cst[j].upload(cst.data(), cst.size()));
    }
    L.label_bytes = 16.0 * c->k * (double)(L.in_size + L.out_size);
    L.weight_bytes = (double)c->k * (double)M * (double)K;
    return DASHB_OK;
}

int upload_tables(DevMem& dst, const dashb_u128* src, size_t rows, int replicas = 1) {
    if (!src && rows) return fail(DASHB_EINVAL, "null ciphertext table");
    if (replicas <= 1) {
        CU(dst.upload(src, rows * 16));
        return DASHB_OK;
    }
    DevMem once;
    CU(once.upload(src, rows * 16));
    CU(dst.alloc(rows * 16 * (size_t)replicas));
    CU(launch_replicate_rows(once.p, dst.p, rows, (uint32_t)replicas, cudaStreamLegacy));
    CU(cudaStreamSynchronize(cudaStreamLegacy));
    return DASHB_OK;
}

int sign_table_rows(dashb_circuit* c, uint64_t n, int n_out, size_t rows[4]) {
    if (c->n_mrs < 1) return fail(DASHB_EINVAL, "sign gadget needs an MRS base");
    size_t S = 0, Mp = 0;
    for (int j = 0; j < c->k; ++j) S += (size_t)c->crt[j];
    for (int i = 1; i < c->n_mrs; ++i) Mp += (size_t)c->mrs[i];
    rows[0] = n * c->n_mrs * S;
    rows[1] = n * (c->k + 1) * Mp;
    rows[2] = rows[1];
    rows[3] = n * (size_t)n_out * c->mrs[0];
    return DASHB_OK;
}

void fill_sign_params(dashb_circuit* c, SignParams& p, uint64_t n, int n_out) {
    memset(&p, 0, sizeof(p));
    p.k = c->k;
    p.n_mrs = c->n_mrs;
    p.n_out = n_out;
    p.lanes = c->lanes_used;
    p.n = n;
    uint32_t S = 0, Mp = 0;
    for (int j = 0; j < c->k; ++j) {
        p.crt[j] = c->crt_mi[j];
        p.crtf[j] = c->crt_fa[j];
        p.prefix[j] = S;
        S += (uint32_t)c->crt[j];
    }
    for (int i = 0; i < c->n_mrs; ++i) {
        p.mrs[i] = c->mrs_mi[i];
        p.cast[i] = c->cast_mi[i];
        p.castf[i] = c->cast_fi[i];
        if (i) Mp += (uint32_t)c->mrs[i];
    }
    p.S = S;
    p.Mp = Mp;
    p.carry0 = c->carry0;
    p.tab_mul = 1;
    p.tab_add = 0;
    p.rep = c->table_replicas > 1 ? (uint32_t)c->table_replicas : 1u;
    p.b0 = 0;
}

double sign_inner_aes_ref(const dashb_circuit* c) {
    const int k = c->k, n = c->n_mrs;
    return (double)k * n + (double)(k + 1) * (n - 1) + (n - 1);
}
double sign_inner_aes(const dashb_circuit* c) {
    const int k = c->k, n = c->n_mrs;
    return (double)k + (double)(k + 1) * (n - 1) + (n - 1) + 1;
}

void push_layer(dashb_circuit* c, std::unique_ptr<Layer> L) {
    for (int j = 0; j < c->k; ++j) c->cur_q[j] = L->out_q[j];
    c->cur_size = L->out_size;
    c->layers.push_back(std::move(L));
}

}  // namespace

namespace dashb {
gc::ModInfo modinfo_for(uint32_t q) { return make_modinfo(q); }
int report_error(int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof(buf), fmt, ap);
    va_end(ap);
    g_err = buf;
    return code;
}
cudaStream_t library_stream() { return g_stream; }
int initialised_device() { return g_device; }
}  // namespace dashb

extern "C" {

const char* dashb_last_error(void) { return g_err.c_str(); }
int dashb_nr_comps(int modulus) { return modulus >= 2 ? nr_comps(modulus) : 0; }
void* dashb_stream(void) { return (void*)g_stream; }

int dashb_init(int device) {
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0)
        return fail(DASHB_ENODEV, "no CUDA device (%s); this library has no CPU path",
                    e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
    if (device < 0 || device >= count) return fail(DASHB_EINVAL, "device %d out of range (%d devices)", device, count);
    CU(cudaSetDevice(device));
    if (g_device != device) {
        if (g_stream) cudaStreamDestroy(g_stream);
        g_stream = nullptr;
        CU(cudaStreamCreateWithFlags(&g_stream, cudaStreamNonBlocking));
    }
    CU(upload_aes_tables());
    CU(cudaDeviceSynchronize()); /* the tables went up on the NULL stream; the library's streams are non-blocking */
    g_device = device;
    return DASHB_OK;
}

int dashb_circuit_create(dashb_circuit** out, const int32_t* crt_base, int k, const int32_t* mrs_base, int n_mrs,
                         int max_modulus, const int16_t* const* zero_labels, const int16_t* const* upshift_labels,
                         uint64_t input_size) {
    if (!out) return fail(DASHB_EINVAL, "null out pointer");
    *out = nullptr;
    if (g_device < 0) return fail(DASHB_ENODEV, "dashb_init() has not succeeded; no device, no evaluation");
    if (!crt_base || k < 1 || k > MAXK) return fail(DASHB_EINVAL, "crt base size %d unsupported (1..%d)", k, MAXK);
    if (n_mrs < 0 || n_mrs > MAXM || (n_mrs && !mrs_base)) return fail(DASHB_EINVAL, "mrs base size %d unsupported", n_mrs);
    if (!zero_labels || max_modulus < 2) return fail(DASHB_EINVAL, "zero labels missing");
    if (input_size == 0) return fail(DASHB_EINVAL, "empty circuit input");
    auto c = std::make_unique<dashb_circuit>();
    c->k = k;
    c->n_mrs = n_mrs;
    c->max_modulus = max_modulus;
    c->input_size = input_size;
    for (int j = 0; j < k; ++j) {
        if (crt_base[j] < 2 || crt_base[j] > 256)
            return fail(DASHB_EINVAL, "CRT modulus %d outside the u8 digit range [2,256]", crt_base[j]);
        if (crt_base[j] > max_modulus) return fail(DASHB_EINVAL, "no zero label for CRT modulus %d", crt_base[j]);
        c->crt[j] = crt_base[j];
        c->crt_mi[j] = make_modinfo((uint32_t)crt_base[j]);
        c->crt_fi[j] = make_fracinfo(c->crt_mi[j]);
        c->crt_fa[j] = make_fracinfo(c->crt_mi[j], true);
        c->cur_q[j] = crt_base[j];
    }
    c->cur_size = input_size;
    c->zero.resize((size_t)max_modulus - 1);
    for (int q = 2; q <= max_modulus; ++q) {
        if (!zero_labels[q - 2]) return fail(DASHB_EINVAL, "zero label of modulus %d missing", q);
        c->zero[q - 2].assign(zero_labels[q - 2], zero_labels[q - 2] + nr_comps(q));
    }
    for (int i = 0; i < n_mrs; ++i) {
        if (mrs_base[i] < 2 || mrs_base[i] > 32767) return fail(DASHB_EINVAL, "bad MRS modulus %d", mrs_base[i]);
        c->mrs[i] = mrs_base[i];
        c->mrs_mi[i] = make_modinfo((uint32_t)mrs_base[i]);
        const int64_t cm = (int64_t)(k + 1) * mrs_base[i];
        if (i > 0 && cm > 32767) return fail(DASHB_EINVAL, "cast modulus %lld too large", (long long)cm);
        /* the sign kernels sum k+1 digits of this modulus in 16-bit fields */
        if ((int64_t)(k + 1) * ((i > 0 ? cm : (int64_t)mrs_base[i]) - 1) > 65535)
            return fail(DASHB_EINVAL, "MRS modulus %d: the sum of %d digits does not fit 16 bits", mrs_base[i], k + 1);
        c->cast_mi[i] = make_modinfo((uint32_t)(i > 0 ? cm : mrs_base[i]));
        c->cast_fi[i] = make_fracinfo(c->cast_mi[i]);
    }
    if (n_mrs) {
        const int ql = mrs_base[n_mrs - 1];
        if (ql > max_modulus) return fail(DASHB_EINVAL, "no zero label for MRS modulus %d", ql);
        c->carry0 = compress_host(c->zero[ql - 2].data(), c->mrs_mi[n_mrs - 1]);
    }
    for (int j = 0; j < k; ++j) {
        auto z = to_u8(c->zero[c->crt[j] - 2].data(), c->crt_mi[j].nc);
        CU(c->d_zero[j].upload(z.data(), z.size()));
        if (upshift_labels) {
            if (!upshift_labels[j]) return fail(DASHB_EINVAL, "upshift label %d missing", j);
            auto u = to_u8(upshift_labels[j], c->crt_mi[j].nc);
            CU(c->d_up[j].upload(u.data(), u.size()));
        }
    }
    c->has_up = upshift_labels != nullptr;
    c->table_replicas = g_load_replicas;
    *out = c.release();
    return DASHB_OK;
}

int dashb_add_dense(dashb_circuit* c, uint64_t M, uint64_t K, int channel_tf, const int64_t* weights,
                    const int16_t* const* bias_labels) {
    int rc = check_building(c);
    if (rc) return rc;
    if ((rc = check_all_crt(c, "dense"))) return rc;
    if (!weights || !bias_labels || M == 0 || K == 0) return fail(DASHB_EINVAL, "dense: bad arguments");
    if (K != c->cur_size) return fail(DASHB_EINVAL, "dense: K=%llu but previous layer has %llu labels", (unsigned long long)K, (unsigned long long)c->cur_size);
    if (channel_tf < 0 || (channel_tf > 0 && K % (uint64_t)channel_tf != 0)) return fail(DASHB_EINVAL, "dense: bad channel_tf");
    auto L = std::make_unique<Layer>();
    L->type = DASHB_LAYER_DENSE;
    L->in_size = K;
    L->out_size = M;
    for (int j = 0; j < c->k; ++j) L->out_q[j] = c->crt[j];
    std::vector<int64_t> w(weights, weights + M * K);
    if ((rc = prep_linear(c, *L, w, M, K, bias_labels))) return rc;
    std::vector<int32_t> koff(L->Kp, 0);
    for (uint64_t kk = 0; kk < K; ++kk) {
        /* matvecmul_tf input permutation (label_tensor.h:848) */
        koff[kk] = channel_tf ? (int32_t)(kk / channel_tf + (kk % channel_tf) * (K / channel_tf)) : (int32_t)kk;
    }
    CU(L->koff.upload(koff.data(), koff.size() * 4));
    /* tcgen05 path: plain input order, exact s32 accumulation, worth a 128-row tile */
    static const bool tc_off = getenv("DASHB_NO_TC") != nullptr;
    if (!tc_off && channel_tf == 0 && K * 65025ull < (1ull << 31) && M >= 16 && K >= 32) {
        const uint64_t Kp16 = (K + 15) & ~15ull;
        for (int j = 0; j < c->k; ++j) {
            const int64_t q = c->crt[j];
            std::vector<uint8_t> wr((size_t)M * Kp16, 0);
            for (uint64_t r = 0; r < M; ++r)
                for (uint64_t kk = 0; kk < K; ++kk) {
                    int64_t wv = weights[r * K + kk] % q;
                    if (wv < 0) wv += q;
                    wr[r * Kp16 + kk] = (uint8_t)wv;
                }
            CU(L->wrm[j].upload(wr.data(), wr.size()));
            CU(dense_tc_make_map(&L->map_a[j], L->wrm[j].p, Kp16, M));
        }
        L->use_tc = true;
        L->tc_nkb = (uint32_t)((K + 127) / 128);
    }
    push_layer(c, std::move(L));
    return DASHB_OK;
}

int dashb_add_conv2d(dashb_circuit* c, uint64_t in_w, uint64_t in_h, uint64_t channels, uint64_t filters,
                     uint64_t f_w, uint64_t f_h, uint64_t s_w, uint64_t s_h, const int64_t* weights,
                     const int16_t* const* bias_labels) {
    int rc = check_building(c);
    if (rc) return rc;
    if ((rc = check_all_crt(c, "conv2d"))) return rc;
    if (!weights || !bias_labels || !in_w || !in_h || !channels || !filters || !f_w || !f_h || !s_w || !s_h ||
        f_w > in_w || f_h > in_h)
        return fail(DASHB_EINVAL, "conv2d: bad geometry");
    if (in_w * in_h * channels != c->cur_size) return fail(DASHB_EINVAL, "conv2d: input size mismatch");
    const uint64_t ow = (in_w - f_w) / s_w + 1, oh = (in_h - f_h) / s_h + 1;
    const uint64_t fs = f_w * f_h, K = channels * fs, nw = filters * channels * fs;
    auto L = std::make_unique<Layer>();
    L->type = DASHB_LAYER_CONV2D;
    L->in_size = c->cur_size;
    L->out_size = ow * oh * filters;
    for (int j = 0; j < c->k; ++j) L->out_q[j] = c->crt[j];
    /* taps ordered (channel w, filter row i, filter column j); the weight of
     * a tap is read at the reference's index v*fs*C + w*fs + i*f_h + j
     * (label_tensor.h:967-969) */
    std::vector<int64_t> wrow(filters * K);
    std::vector<int32_t> koff((K + 3) & ~3ull, 0);
    for (uint64_t v = 0; v < filters; ++v)
        for (uint64_t w = 0; w < channels; ++w)
            for (uint64_t i = 0; i < f_h; ++i)
                for (uint64_t jx = 0; jx < f_w; ++jx) {
                    const uint64_t widx = v * fs * channels + w * fs + i * f_h + jx;
                    if (widx >= nw) return fail(DASHB_EINVAL, "conv2d: the reference's weight index leaves the array for this filter shape");
                    const uint64_t kk = (w * f_h + i) * f_w + jx;
                    wrow[v * K + kk] = weights[widx];
                    koff[kk] = (int32_t)(w * in_w * in_h + i * in_w + jx);
                }
    if ((rc = prep_linear(c, *L, wrow, filters, K, bias_labels))) return rc;
    L->npix = (uint32_t)(ow * oh);
    L->ow = (uint32_t)ow;
    L->step_x = (uint32_t)s_w;
    L->step_y = (uint32_t)(s_h * in_w);
    if (ow * oh > 0x7fffffffull || ((ow * oh + 3) / 4) * ((filters + 7) / 8) > 0xffffffffull)
        return fail(DASHB_EINVAL, "conv2d: output grid too large");
    CU(L->koff.upload(koff.data(), koff.size() * 4));
    /* tcgen05 path (dense_tc.cu): one GEMM per output pixel, im2col by a 4-D
     * TMA box of cb channels; weights regrouped into 128-column K blocks of
     * cb*fs taps each, zero padded */
    static const bool tc_off = getenv("DASHB_NO_TC") != nullptr;
    if (!tc_off && filters >= 16 && K >= 64 && fs <= 128 && f_w <= 256 && f_h <= 256 && K * 65025ull < (1ull << 31) &&
        ow * oh * ((filters + 127) / 128) <= 65535) {
        const uint64_t cb = std::min<uint64_t>(128 / fs, channels);
        const uint64_t nkb = (channels + cb - 1) / cb, Kpad = nkb * 128;
        for (int j = 0; j < c->k; ++j) {
            const int64_t q = c->crt[j];
            std::vector<uint8_t> wr((size_t)filters * Kpad, 0);
            for (uint64_t v = 0; v < filters; ++v)
                for (uint64_t w = 0; w < channels; ++w)
                    for (uint64_t t = 0; t < fs; ++t) {
                        int64_t wv = wrow[v * K + w * fs + t] % q;
                        if (wv < 0) wv += q;
                        wr[v * Kpad + (w / cb) * 128 + (w % cb) * fs + t] = (uint8_t)wv;
                    }
            CU(L->wrm[j].upload(wr.data(), wr.size()));
            CU(dense_tc_make_map(&L->map_a[j], L->wrm[j].p, Kpad, filters));
        }
        L->use_tc = true;
        L->tc_nkb = (uint32_t)nkb;
        L->cv_w = (uint32_t)in_w, L->cv_h = (uint32_t)in_h, L->cv_c = (uint32_t)channels;
        L->cv_fw = (uint32_t)f_w, L->cv_fh = (uint32_t)f_h, L->cv_sw = (uint32_t)s_w, L->cv_sh = (uint32_t)s_h;
        L->cv_cb = (uint32_t)cb;
        /* strip kernel: stride 1 in x, 32 or 64 filters, cblk channels per block with cblk * f_h a multiple of
         * the MMA's K = 32 and at most 128 rows per input column */
        static const int strip_env = getenv("DASHB_CONV_STRIP") ? atoi(getenv("DASHB_CONV_STRIP")) : 4;
        uint64_t cblk = 0;
        for (uint64_t cand = 128 / f_h; cand >= 1; --cand)
            if ((cand * f_h) % 32 == 0) {
                cblk = cand;
                break;
            }
        /* fewer channels than a block: the smallest block that still holds them */
        for (uint64_t cand = 1; cblk && cand < cblk; ++cand)
            if ((cand * f_h) % 32 == 0 && cand >= channels) {
                cblk = cand;
                break;
            }
        if (strip_env >= 1 && s_w == 1 && (filters == 32 || filters == 64) && cblk && f_w >= 1) {
            uint32_t P = (uint32_t)strip_env;
            const uint32_t R = (uint32_t)(cblk * f_h), wboxes = (uint32_t)((f_w * filters + 127) / 128);
            while (P > 1 && (P + f_w - 1 > 8 || (P + f_w - 1) * R * 128ull + wboxes * 16384ull > 108 * 1024 || P * filters > 256)) --P;
            if (P + f_w - 1 <= 8 && (P + f_w - 1) * R * 128ull + wboxes * 16384ull <= 196 * 1024 && P * filters <= 256) {
                const uint64_t ncb = (channels + cblk - 1) / cblk;
                const uint64_t rows = ncb * f_w * filters;
                for (int j = 0; j < c->k; ++j) {
                    const int64_t q = c->crt[j];
                    std::vector<uint8_t> wt((size_t)rows * 128, 0);
                    for (uint64_t v = 0; v < filters; ++v)
                        for (uint64_t w = 0; w < channels; ++w)
                            for (uint64_t i = 0; i < f_h; ++i)
                                for (uint64_t jx = 0; jx < f_w; ++jx) {
                                    int64_t wv = wrow[v * K + (w * f_h + i) * f_w + jx] % q;
                                    if (wv < 0) wv += q;
                                    wt[(((w / cblk) * f_w + jx) * filters + v) * 128 + (w % cblk) * f_h + i] = (uint8_t)wv;
                                }
                    CU(L->wst[j].upload(wt.data(), wt.size()));
                    CU(dense_tc_make_map(&L->map_st[j], L->wst[j].p, 128, rows));
                }
                L->st_strip = P;
                L->st_rows = R;
                L->st_cblk = (uint32_t)cblk;
                L->st_ncb = (uint32_t)ncb;
                L->st_wboxes = wboxes;
            }
        }
    }
    push_layer(c, std::move(L));
    return DASHB_OK;
}

static int add_sign_common(dashb_circuit* c, Layer& L, uint64_t n, int n_out, const dashb_u128* approx,
                           const dashb_u128* cast1, const dashb_u128* cast2, const dashb_u128* sign) {
    size_t rows[4];
    int rc = sign_table_rows(c, n, n_out, rows);
    if (rc) return rc;
    const int R = c->table_replicas;
    if ((rc = upload_tables(L.approx, approx, rows[0], R))) return rc;
    if ((rc = upload_tables(L.cast1, cast1, rows[1], R))) return rc;
    if ((rc = upload_tables(L.cast2, cast2, rows[2], R))) return rc;
    return upload_tables(L.sign, sign, rows[3], R);
}

int dashb_add_relu(dashb_circuit* c, uint64_t n, const dashb_u128* approx, const dashb_u128* cast1,
                   const dashb_u128* cast2, const dashb_u128* sign, const dashb_u128* garbler,
                   const dashb_u128* evaluator) {
    int rc = check_building(c);
    if (rc) return rc;
    if ((rc = check_all_crt(c, "relu"))) return rc;
    if (n != c->cur_size) return fail(DASHB_EINVAL, "relu: size mismatch");
    if (2 > c->max_modulus) return fail(DASHB_EINVAL, "relu: modulus 2 unavailable");
    auto L = std::make_unique<Layer>();
    L->type = DASHB_LAYER_RELU;
    L->in_size = L->out_size = n;
    for (int j = 0; j < c->k; ++j) L->out_q[j] = c->crt[j];
    if ((rc = add_sign_common(c, *L, n, 1, approx, cast1, cast2, sign))) return rc;
    size_t S = 0;
    for (int j = 0; j < c->k; ++j) S += (size_t)c->crt[j];
    if ((rc = upload_tables(L->garbler, garbler, n * S, c->table_replicas))) return rc;
    if ((rc = upload_tables(L->evaluator, evaluator, n * (size_t)c->k * 3, c->table_replicas))) return rc;
    L->ref_aes = (double)n * (sign_inner_aes_ref(c) + 1 + 3.0 * c->k);
    L->aes = (double)n * (sign_inner_aes(c) + 1);
    L->label_bytes = (double)n * (32.0 * c->k + 16.0 * (sign_inner_aes_ref(c) + 1 + 3.0 * c->k));
    push_layer(c, std::move(L));
    return DASHB_OK;
}

int dashb_add_maxpool2d(dashb_circuit* c, uint64_t in_w, uint64_t in_h, uint64_t channels, uint64_t k_w, uint64_t k_h,
                        uint64_t s_w, uint64_t s_h, const dashb_u128* approx, const dashb_u128* cast1, const dashb_u128* cast2,
                        const dashb_u128* sign, const dashb_u128* garbler, const dashb_u128* evaluator) {
    int rc = check_building(c);
    if (rc) return rc;
    if ((rc = check_all_crt(c, "maxpool2d"))) return rc;
    if (!in_w || !in_h || !channels || !k_w || !k_h || !s_w || !s_h || k_w > in_w || k_h > in_h || in_w > 65535 || in_h > 65535)
        return fail(DASHB_EINVAL, "maxpool2d: bad geometry");
    if (in_w * in_h * channels != c->cur_size) return fail(DASHB_EINVAL, "maxpool2d: input size mismatch");
    const uint64_t ow = (in_w - k_w) / s_w + 1, oh = (in_h - k_h) / s_h + 1;
    /* the reference sizes the output tensor as (W / k_w, H / k_h, C) (circuit/layer/max_pool2d.h:26-29) and walks
     * ((W - k_w) / s_w + 1, ...) windows (layer/garbled_maxpool2d.h:150-153): only shapes where the two agree are defined */
    if (ow != in_w / k_w || oh != in_h / k_h)
        return fail(DASHB_EINVAL, "maxpool2d: %llux%llu windows do not match the reference's %llux%llu output tensor",
                    (unsigned long long)ow, (unsigned long long)oh, (unsigned long long)(in_w / k_w), (unsigned long long)(in_h / k_h));
    if (c->shard_world > 1) return fail(DASHB_EINVAL, "maxpool2d: not available in channel-sharded circuits");
    if (c->table_replicas > 1) return fail(DASHB_EINVAL, "maxpool2d: table replicas are not supported");
    const uint64_t n_out = ow * oh * channels, T = k_w * k_h, n = n_out * T;
    auto L = std::make_unique<Layer>();
    L->type = DASHB_LAYER_MAXPOOL2D;
    L->in_size = c->cur_size;
    L->out_size = n_out;
    for (int j = 0; j < c->k; ++j) L->out_q[j] = c->crt[j];
    if ((rc = add_sign_common(c, *L, n, 1, approx, cast1, cast2, sign))) return rc;
    size_t S = 0;
    for (int j = 0; j < c->k; ++j) S += (size_t)c->crt[j];
    if ((rc = upload_tables(L->garbler, garbler, n * S))) return rc;
    if ((rc = upload_tables(L->evaluator, evaluator, n * (size_t)c->k * 3))) return rc;
    L->pl_w = (uint32_t)in_w, L->pl_h = (uint32_t)in_h, L->pl_c = (uint32_t)channels;
    L->pl_kw = (uint32_t)k_w, L->pl_kh = (uint32_t)k_h, L->pl_sw = (uint32_t)s_w, L->pl_sh = (uint32_t)s_h;
    L->pl_ow = (uint32_t)ow, L->pl_oh = (uint32_t)oh;
    L->ref_aes = (double)n * (sign_inner_aes_ref(c) + 1 + 3.0 * c->k);
    L->aes = (double)n * (sign_inner_aes(c) + 1);
    L->label_bytes = (double)n * (32.0 * c->k + 16.0 * (sign_inner_aes_ref(c) + 1 + 3.0 * c->k));
    push_layer(c, std::move(L));
    return DASHB_OK;
}

int dashb_add_sign(dashb_circuit* c, uint64_t n, const dashb_u128* approx, const dashb_u128* cast1,
                   const dashb_u128* cast2, const dashb_u128* sign) {
    int rc = check_building(c);
    if (rc) return rc;
    if ((rc = check_all_crt(c, "sign"))) return rc;
    if (n != c->cur_size) return fail(DASHB_EINVAL, "sign: size mismatch");
    auto L = std::make_unique<Layer>();
    L->type = DASHB_LAYER_SIGN;
    L->in_size = L->out_size = n;
    for (int j = 0; j < c->k; ++j) L->out_q[j] = c->crt[j];
    if ((rc = add_sign_common(c, *L, n, c->k, approx, cast1, cast2, sign))) return rc;
    L->ref_aes = (double)n * (sign_inner_aes_ref(c) + c->k);
    L->aes = (double)n * sign_inner_aes(c);
    L->label_bytes = (double)n * (32.0 * c->k + 16.0 * (sign_inner_aes_ref(c) + c->k));
    push_layer(c, std::move(L));
    return DASHB_OK;
}

static int upload_downshift(dashb_circuit* c, Layer& L, const int16_t* const* downshift_labels) {
    if (!downshift_labels) return fail(DASHB_EINVAL, "rescale: downshift labels missing");
    if (!c->has_up) return fail(DASHB_EINVAL, "rescale: circuit was created without upshift labels");
    for (int j = 0; j < c->k; ++j) {
        if (!downshift_labels[j]) return fail(DASHB_EINVAL, "rescale: downshift label %d missing", j);
        auto d = to_u8(downshift_labels[j], c->crt_mi[j].nc);
        CU(L.down[j].upload(d.data(), d.size()));
    }
    return DASHB_OK;
}

/* fills the parts of RescaleParams that do not depend on the batch */
static int fill_rescale_static(dashb_circuit* c, Layer& L, uint64_t n, const int32_t* factors, int nf, int legacy) {
    RescaleParams& p = L.rp;
    memset(&p, 0, sizeof(p));
    const int k = c->k;
    p.k = k;
    p.nf = nf;
    p.n = n;
    p.legacy = legacy;
    uint32_t off = 0;
    for (int j = 0; j < k; ++j) {
        p.crt[j] = c->crt_mi[j];
        p.crtf[j] = c->crt_fa[j].nst ? c->crt_fa[j] : c->crt_fi[j]; /* word-aligned stages where the modulus has them */
        p.loff[j] = off;
        off += (c->crt_mi[j].nc + 3) & ~3u;
    }
    /* Step 2 bookkeeping: sequential cipher counter over (factor, modulus) */
    bool ignored[MAXK] = {false};
    uint64_t cnt = 0;
    for (int i = 0; i < nf; ++i) {
        int f = -1;
        for (int j = 0; j < k; ++j)
            if (c->crt[j] == factors[i] && f < 0) f = j;
        if (f < 0) return fail(DASHB_EINVAL, "rescale: factor %d is not a CRT modulus", factors[i]);
        if (ignored[f]) return fail(DASHB_EINVAL, "rescale: duplicate factor %d", factors[i]);
        ignored[f] = true;
        p.fidx[i] = f;
        p.factor[i] = (uint32_t)factors[i];
        for (int j = 0; j < k; ++j) {
            if (ignored[j]) {
                p.tm_off[i][j] = ~0ull;
                continue;
            }
            p.tm_off[i][j] = cnt;
            cnt += n * (uint64_t)factors[i];
            const uint64_t inv = mul_inv((unsigned __int128)factors[i], (uint16_t)c->crt[j]);
            if ((inv * (uint64_t)factors[i]) % (uint64_t)c->crt[j] != 1)
                return fail(DASHB_EINVAL, "rescale: factor %d not invertible mod %d", factors[i], c->crt[j]);
            p.inv_f[i][j] = (uint16_t)inv;
        }
    }
    /* lazy-reduction form of step 2 (k_rescale): for residue j the updates
     * w <- (w - p_i) * inv_i over its factors i collapse to w0*a1 + sum p_i*b1 */
    for (int j = 0; j < k; ++j) {
        const uint64_t q = (uint64_t)c->crt[j];
        if ((uint64_t)(k + 1) * (q - 1) * (q - 1) >= 65536) p.wide = 1;
        uint64_t suffix = 1; /* product of inv_f[i'][j] over the valid i' >= i */
        for (int i = nf - 1; i >= 0; --i) {
            if (p.tm_off[i][j] == ~0ull) continue;
            suffix = suffix * p.inv_f[i][j] % q;
            p.b1[i][j] = (uint16_t)((q - suffix) % q);
        }
        p.a1[j] = (uint16_t)suffix;
    }
    for (int i = 0; i < nf; ++i) { /* factor i's own residue: only the factors before it */
        const int j = p.fidx[i];
        const uint64_t q = (uint64_t)c->crt[j];
        uint64_t suffix = 1;
        for (int i2 = i - 1; i2 >= 0; --i2) {
            if (p.tm_off[i2][j] == ~0ull) continue;
            suffix = suffix * p.inv_f[i2][j] % q;
            p.b1f[i][i2] = (uint16_t)((q - suffix) % q);
        }
        p.a1f[i] = (uint16_t)suffix;
    }
    if (legacy) return DASHB_OK;
    /* base extension constants (base_extension_gadget.h:334-377) */
    int perm[MAXK], swapped[MAXK];
    for (int j = 0; j < k; ++j) perm[j] = j;
    for (int i = 1; i <= nf; ++i) {
        const int a = p.fidx[nf - i], b = k - i;
        std::swap(perm[a], perm[b]);
    }
    for (int j = 0; j < k; ++j) {
        swapped[j] = c->crt[perm[j]];
        p.perm[j] = perm[j];
    }
    p.non_ext = k - nf;
    uint64_t bcnt = 0;
    for (int i = 0; i < p.non_ext; ++i)
        for (int t = i + 1; t < k; ++t) {
            p.be_off[i][t] = bcnt;
            bcnt += n * (uint64_t)swapped[i];
            p.inv_partial[i][t] = (uint16_t)(int16_t)mul_inv((unsigned __int128)swapped[i], (uint16_t)swapped[t]);
        }
    for (int t = 0; t < k; ++t) { /* lazy-reduction form of the eliminations into target t */
        const uint64_t q = (uint64_t)swapped[t];
        const int ni = t < p.non_ext ? t : p.non_ext;
        uint64_t suffix = 1;
        for (int i = ni - 1; i >= 0; --i) {
            suffix = suffix * (p.inv_partial[i][t] % q) % q;
            p.b2[i][t] = (uint16_t)((q - suffix) % q);
        }
        p.a2[t] = (uint16_t)suffix;
    }
    for (int x = 0; x < nf; ++x) {
        int acc_i = 1; /* std::accumulate with an int accumulator (:351) */
        for (int j = 0; j < p.non_ext; ++j) acc_i = (int)((unsigned long)(long)acc_i * (unsigned long)swapped[j]);
        const unsigned long acc = (unsigned long)(long)acc_i;
        const int16_t inv_total = (int16_t)mul_inv((unsigned __int128)acc, (uint16_t)factors[x]);
        const int16_t invv = (int16_t)mul_inv((unsigned __int128)(__int128)inv_total, (uint16_t)factors[x]);
        int64_t neg = (-(int64_t)invv) % (int64_t)factors[x];
        if (neg < 0) neg += factors[x];
        p.neg_invv[x] = (uint16_t)neg;
        p.extra_idx[x] = p.fidx[x];
        for (int j = 0; j < k; ++j)
            if (swapped[j] == factors[x]) {
                p.extra_pos[x] = j;
                break;
            }
    }
    return DASHB_OK;
}

int dashb_add_rescale(dashb_circuit* c, uint64_t n, const int32_t* factors, int nf,
                      const int16_t* const* downshift_labels, const dashb_u128* trans_mod,
                      const dashb_u128* be_ciphers) {
    int rc = check_building(c);
    if (rc) return rc;
    if ((rc = check_all_crt(c, "rescale"))) return rc;
    if (n != c->cur_size) return fail(DASHB_EINVAL, "rescale: size mismatch");
    if (!factors || nf < 1 || nf > MAXF || nf >= c->k) return fail(DASHB_EINVAL, "rescale: bad factor list");
    auto L = std::make_unique<Layer>();
    L->type = DASHB_LAYER_RESCALE;
    L->in_size = L->out_size = n;
    L->nf = nf;
    for (int j = 0; j < c->k; ++j) L->out_q[j] = c->crt[j];
    if ((rc = upload_downshift(c, *L, downshift_labels))) return rc;
    if ((rc = fill_rescale_static(c, *L, n, factors, nf, 0))) return rc;
    const int k = c->k;
    size_t tm_rows = 0, be_rows = 0;
    for (int i = 0; i < nf; ++i) tm_rows += (size_t)(k - 1 - i) * n * (size_t)factors[i];
    for (int i = 0; i < L->rp.non_ext; ++i) be_rows += (size_t)(k - i - 1) * n * (size_t)c->crt[L->rp.perm[i]];
    if ((rc = upload_tables(L->tm, trans_mod, tm_rows))) return rc;
    if ((rc = upload_tables(L->be, be_ciphers, be_rows))) return rc;
    double gates = 0, hashes = nf + L->rp.non_ext;
    for (int i = 0; i < nf; ++i) gates += k - 1 - i;
    for (int i = 0; i < L->rp.non_ext; ++i) gates += k - i - 1;
    L->ref_aes = (double)n * gates;
    L->aes = (double)n * hashes;
    L->label_bytes = (double)n * (32.0 * k + 16.0 * gates);
    push_layer(c, std::move(L));
    return DASHB_OK;
}

int dashb_add_base_extension(dashb_circuit* c, uint64_t n, const int32_t* extra_moduli, int n_extra,
                             const int16_t* const* start_labels, const dashb_u128* be_ciphers) {
    int rc = check_building(c);
    if (rc) return rc;
    if ((rc = check_all_crt(c, "base extension"))) return rc;
    if (n != c->cur_size) return fail(DASHB_EINVAL, "base extension: size mismatch");
    if (!extra_moduli || !start_labels || n_extra < 1 || n_extra > MAXF || n_extra >= c->k)
        return fail(DASHB_EINVAL, "base extension: bad list of extra moduli");
    auto L = std::make_unique<Layer>();
    L->type = DASHB_LAYER_BASE_EXTENSION;
    L->in_size = L->out_size = n;
    L->nf = n_extra;
    for (int j = 0; j < c->k; ++j) L->out_q[j] = c->crt[j];
    /* the gadget's constants are those of the rescale layer's step 3 with the extra moduli as factors */
    if ((rc = fill_rescale_static(c, *L, n, extra_moduli, n_extra, 0))) return rc;
    RescaleParams& p = L->rp;
    for (int i = 0; i < MAXF; ++i)
        for (int j = 0; j < MAXK; ++j) p.tm_off[i][j] = ~0ull; /* no division step: nothing is projected from the extra residues */
    for (int j = 0; j < c->k; ++j) p.a1[j] = 1;
    for (int i = 0; i < MAXF; ++i) p.a1f[i] = 1;
    for (int x = 0; x < n_extra; ++x) {
        const int j = p.fidx[x];
        if (!start_labels[x]) return fail(DASHB_EINVAL, "base extension: start labels %d missing", x);
        const uint32_t nc = c->crt_mi[j].nc, ncp = (nc + 3) & ~3u;
        std::vector<uint8_t> b((size_t)n * ncp, 0);
        for (uint64_t e = 0; e < n; ++e)
            for (uint32_t d = 0; d < nc; ++d) b[e * ncp + d] = (uint8_t)start_labels[x][e * nc + d];
        CU(L->start[j].upload(b.data(), b.size()));
    }
    if (!c->d_nolabel.p) {
        std::vector<uint8_t> z(512, 0);
        CU(c->d_nolabel.upload(z.data(), z.size()));
    }
    size_t be_rows = 0;
    for (int i = 0; i < p.non_ext; ++i) be_rows += (size_t)(c->k - i - 1) * n * (size_t)c->crt[p.perm[i]];
    if ((rc = upload_tables(L->be, be_ciphers, be_rows))) return rc;
    double gates = 0;
    for (int i = 0; i < p.non_ext; ++i) gates += c->k - i - 1;
    L->ref_aes = (double)n * gates;
    L->aes = (double)n * p.non_ext;
    L->label_bytes = (double)n * (32.0 * c->k + 16.0 * gates);
    push_layer(c, std::move(L));
    return DASHB_OK;
}

int dashb_add_rescale_legacy(dashb_circuit* c, uint64_t n, int l, const int16_t* const* downshift_labels,
                             const dashb_u128* const* trans_mod, const dashb_u128* const* approx,
                             const dashb_u128* const* cast1, const dashb_u128* const* cast2,
                             const dashb_u128* const* sign) {
    int rc = check_building(c);
    if (rc) return rc;
    if ((rc = check_all_crt(c, "rescale"))) return rc;
    if (n != c->cur_size) return fail(DASHB_EINVAL, "rescale: size mismatch");
    if (l < 1 || l > 16 || !trans_mod || !approx || !cast1 || !cast2 || !sign) return fail(DASHB_EINVAL, "rescale: bad arguments");
    if (c->crt[0] != 2 || c->k < 2) return fail(DASHB_EINVAL, "legacy rescale needs crt_base[0] == 2");
    if (c->table_replicas > 1) return fail(DASHB_EINVAL, "legacy rescale: table replicas are not supported");
    auto L = std::make_unique<Layer>();
    L->type = DASHB_LAYER_RESCALE_LEGACY;
    L->in_size = L->out_size = n;
    L->l_iter = l;
    for (int j = 0; j < c->k; ++j) L->out_q[j] = c->crt[j];
    if ((rc = upload_downshift(c, *L, downshift_labels))) return rc;
    const int32_t two = 2;
    if ((rc = fill_rescale_static(c, *L, n, &two, 1, 1))) return rc;
    size_t rows[4];
    if ((rc = sign_table_rows(c, n, 1, rows))) return rc;
    for (int it = 0; it < l; ++it) {
        const dashb_u128* src[5] = {trans_mod[it], approx[it], cast1[it], cast2[it], sign[it]};
        const size_t nrows[5] = {n * (size_t)(c->k - 1) * 2, rows[0], rows[1], rows[2], rows[3]};
        for (int a = 0; a < 5; ++a) {
            auto m = std::make_unique<DevMem>();
            if ((rc = upload_tables(*m, src[a], nrows[a]))) return rc;
            L->it_tables.push_back(std::move(m));
        }
    }
    const double gates = (c->k - 1) + sign_inner_aes_ref(c) + 1;
    L->ref_aes = (double)n * l * gates;
    L->aes = (double)n * l * (1 + sign_inner_aes(c));
    L->label_bytes = (double)n * l * (32.0 * c->k + 16.0 * gates);
    push_layer(c, std::move(L));
    return DASHB_OK;
}

int dashb_add_projection(dashb_circuit* c, uint64_t n, const int32_t* in_moduli, const int32_t* out_moduli,
                         const dashb_u128* const* ciphers) {
    int rc = check_building(c);
    if (rc) return rc;
    if (!in_moduli || !out_moduli || !ciphers) return fail(DASHB_EINVAL, "projection: bad arguments");
    if (n != c->cur_size) return fail(DASHB_EINVAL, "projection: size mismatch");
    auto L = std::make_unique<Layer>();
    L->type = DASHB_LAYER_PROJECTION;
    L->in_size = L->out_size = n;
    for (int j = 0; j < c->k; ++j) {
        if (in_moduli[j] != c->cur_q[j]) return fail(DASHB_EINVAL, "projection: input modulus %d does not match the incoming labels (%d)", in_moduli[j], c->cur_q[j]);
        if (out_moduli[j] < 2 || out_moduli[j] > 256) return fail(DASHB_EINVAL, "projection: output modulus %d outside the u8 digit range", out_moduli[j]);
        L->in_q[j] = in_moduli[j];
        L->out_q[j] = out_moduli[j];
        if ((rc = upload_tables(L->proj[j], ciphers[j], n * (size_t)in_moduli[j]))) return rc;
    }
    L->ref_aes = L->aes = (double)n * c->k;
    L->label_bytes = 48.0 * n * c->k;
    push_layer(c, std::move(L));
    return DASHB_OK;
}

int dashb_set_decoding_info(dashb_circuit* c, const dashb_u128* const* dec) {
    if (!c || !dec) return fail(DASHB_EINVAL, "null argument");
    if (c->finalized && c->has_dec) return fail(DASHB_ESTATE, "decoding information already set");
    for (int j = 0; j < c->k; ++j) {
        if (!dec[j]) return fail(DASHB_EINVAL, "decoding information %d missing", j);
        CU(c->dec[j].upload(dec[j], (size_t)c->crt[j] * c->cur_size * 16));
    }
    c->has_dec = true;
    return DASHB_OK;
}

int dashb_set_input_encoding(dashb_circuit* c, const int16_t* const* base_labels,
                             const int16_t* const* offset_labels) {
    int rc = check_building(c);
    if (rc) return rc;
    if (!base_labels || !offset_labels) return fail(DASHB_EINVAL, "null argument");
    for (int j = 0; j < c->k; ++j) {
        if (!base_labels[j] || !offset_labels[j]) return fail(DASHB_EINVAL, "input encoding %d missing", j);
        const uint32_t nc = c->crt_mi[j].nc, ncp = (nc + 3) & ~3u;
        std::vector<uint8_t> b((size_t)c->input_size * ncp, 0);
        for (uint64_t e = 0; e < c->input_size; ++e)
            for (uint32_t d = 0; d < nc; ++d) b[e * ncp + d] = (uint8_t)base_labels[j][e * nc + d];
        CU(c->in_base[j].upload(b.data(), b.size()));
        auto o = to_u8(offset_labels[j], nc);
        CU(c->in_off[j].upload(o.data(), o.size()));
    }
    c->has_ginp = true;
    return DASHB_OK;
}

/* Output slice of `rank` for every layer: a linear layer splits its rows
 * (dense rows / conv filters) evenly, the element-wise layers after it keep
 * the same elements. */
static void layer_slice(const dashb_circuit* c, int layer, int rank, uint64_t* e0, uint64_t* e1, uint32_t* v0, uint32_t* v1) {
    const int W = c->shard_world;
    uint64_t a = 0, b = 0;
    uint32_t r0 = 0, r1 = 0;
    for (int li = 0; li <= layer; ++li) {
        const Layer& L = *c->layers[li];
        if (L.type == DASHB_LAYER_DENSE || L.type == DASHB_LAYER_CONV2D) {
            r0 = (uint32_t)(L.M * (uint64_t)rank / W);
            r1 = (uint32_t)(L.M * (uint64_t)(rank + 1) / W);
            a = (uint64_t)r0 * L.npix;
            b = (uint64_t)r1 * L.npix;
        } else if (li == 0) {
            a = L.in_size * (uint64_t)rank / W;
            b = L.in_size * (uint64_t)(rank + 1) / W;
        }
    }
    *e0 = a;
    *e1 = b;
    *v0 = r0;
    *v1 = r1;
}

/* pixel sharding of the convolutions: with the peer-memory exchange (which moves strided slices) and without
 * legacy rescale layers (k_add_label works on contiguous slices) */
static bool pixel_sharding(const dashb_circuit* c) {
    if (c->shard_world <= 1 || c->shard_lanes < 1) return false;
    for (auto& L : c->layers)
        if (L->type == DASHB_LAYER_RESCALE_LEGACY) return false;
    return true;
}

/* this rank's slice of every layer, strided form included (see Layer) */
static void assign_slices(dashb_circuit* c) {
    const bool by_pixel = pixel_sharding(c);
    const int W = c->shard_world, rank = c->shard_rank;
    uint64_t seg = 0, stride = 0, nseg = 0;
    uint32_t p0 = 0, p1 = 0;
    for (int li = 0; li < (int)c->layers.size(); ++li) {
        Layer& L = *c->layers[li];
        layer_slice(c, li, rank, &L.e0, &L.e1, &L.v0, &L.v1);
        if (L.type == DASHB_LAYER_DENSE) {
            seg = stride = nseg = 0;
            p0 = 0, p1 = 1;
        } else if (L.type == DASHB_LAYER_CONV2D) {
            if (by_pixel) {
                p0 = (uint32_t)((uint64_t)L.npix * rank / W);
                p1 = (uint32_t)((uint64_t)L.npix * (rank + 1) / W);
                seg = p1 - p0;
                stride = L.npix;
                nseg = L.M;
                L.v0 = 0;
                L.v1 = (uint32_t)L.M;
                L.e0 = p0;
                L.e1 = p0; /* not a contiguous range: slice_count() counts the runs */
            } else {
                seg = stride = nseg = 0;
                p0 = 0, p1 = L.npix;
            }
        }
        if (seg) {
            L.e0 = p0;
            L.e1 = p0;
        }
        L.seg = seg, L.stride = stride, L.nseg = nseg;
        L.p0 = p0, L.p1 = p1;
    }
}

int dashb_set_shard(dashb_circuit* c, int rank, int world) {
    int rc = check_building(c);
    if (rc) return rc;
    if (world < 1 || rank < 0 || rank >= world) return fail(DASHB_EINVAL, "bad shard %d of %d", rank, world);
    if (world > 1)
        for (auto& L : c->layers)
            if (L->type == DASHB_LAYER_MAXPOOL2D) return fail(DASHB_EINVAL, "maxpool2d: not available in channel-sharded circuits");
    c->shard_rank = rank;
    c->shard_world = world;
    return DASHB_OK;
}

/* One set of sub-batch lanes (streams, events, input / activation / staging
 * buffers) sized for c->lane_cap inferences per lane. */
static int make_lanes(dashb_circuit* c, std::vector<std::unique_ptr<Lane>>& lanes) {
    const uint32_t Bp = c->lane_cap;
    size_t in_bytes = tensor_bytes(c->input_size, c->k, c->crt, Bp);
    size_t act_bytes = 256;
    size_t stage_in = 0, stage_out = 0;
    for (int j = 0; j < c->k; ++j) stage_in += (size_t)Bp * c->input_size * c->crt_mi[j].nc * 2;
    for (auto& L : c->layers) {
        size_t b = tensor_bytes(L->out_size, c->k, L->out_q, Bp);
        if (b > act_bytes) act_bytes = b;
        size_t so = 0;
        for (int j = 0; j < c->k; ++j) so += (size_t)Bp * L->out_size * nr_comps(L->out_q[j]) * 2;
        if (so > stage_out) stage_out = so;
    }
    if (stage_in > stage_out) stage_out = stage_in; /* zero-layer circuits echo the input */
    size_t pool_bytes = 0;
    for (auto& L : c->layers)
        if (L->type == DASHB_LAYER_MAXPOOL2D) pool_bytes = std::max(pool_bytes, 2 * tensor_bytes(L->out_size, c->k, L->out_q, Bp));
    for (int i = 0; i < c->n_lanes; ++i) {
        auto ln = std::make_unique<Lane>();
        CU(cudaStreamCreateWithFlags(&ln->s, cudaStreamNonBlocking));
        CU(cudaEventCreateWithFlags(&ln->done, cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&ln->in_done, cudaEventDisableTiming));
        for (auto& x : ln->in_done_x) CU(cudaEventCreateWithFlags(&x, cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&ln->ing_done, cudaEventDisableTiming));
        CU(ln->buf_in.alloc(in_bytes));
        CU(ln->buf_a.alloc(act_bytes));
        CU(ln->buf_b.alloc(act_bytes));
        CU(ln->stage_in.alloc(stage_in));
        CU(ln->stage_out.alloc(stage_out));
        if (c->has_ginp) CU(ln->xvals.alloc((size_t)Bp * c->input_size * 8));
        if (pool_bytes) {
            CU(ln->pool.alloc(pool_bytes));
            CU(cudaMemset(ln->pool.p, 0, pool_bytes));
        }
        if (c->shard_world > 1 && c->shard_lanes > 0) {
            for (auto& g : ln->gat) {
                CU(g.alloc(act_bytes));
                CU(cudaMemset(g.p, 0, act_bytes));
            }
            CU(ln->flags.alloc(4 * (MAXR + 1)));
            CU(cudaMemset(ln->flags.p, 0, 4 * (MAXR + 1)));
            ln->peer_gat[0][c->shard_rank] = (uint8_t*)ln->gat[0].p;
            ln->peer_gat[1][c->shard_rank] = (uint8_t*)ln->gat[1].p;
            ln->peer_flags[c->shard_rank] = (uint32_t*)ln->flags.p;
        }
        CU(cudaMemset(ln->buf_in.p, 0, in_bytes));
        CU(cudaMemset(ln->buf_a.p, 0, act_bytes));
        CU(cudaMemset(ln->buf_b.p, 0, act_bytes));
        /* (asynchronous on the NULL stream: the callers end with a device synchronisation) */
        lanes.push_back(std::move(ln));
    }
    return DASHB_OK;
}

int dashb_finalize(dashb_circuit* c, uint64_t max_batch) {
    int rc = check_building(c);
    if (rc) return rc;
    if (max_batch == 0 || max_batch > (1u << 24)) return fail(DASHB_EINVAL, "bad max_batch");
    c->max_batch = max_batch;
    assign_slices(c);
    int n_lanes = max_batch >= 128 ? 4 : (max_batch >= 64 ? 2 : 1);
    if (c->shard_world > 1) n_lanes = c->shard_lanes > 0 ? c->shard_lanes : 1; /* NCCL path: whole planes, one lane */
    if (const char* env = getenv("DASHB_LANES")) {
        int v = atoi(env);
        if (v >= 1 && v <= 16 && c->shard_world == 1) n_lanes = v;
    }
    c->lane_cap = (uint32_t)((((max_batch + n_lanes - 1) / n_lanes) + 15) & ~15ull);
    if (const char* env = getenv("DASHB_LANE_SPLIT")) {
        if (c->shard_world == 1) {
            uint64_t wsum = 0, wmax = 0;
            for (const char* p = env; *p;) {
                char* endp = nullptr;
                const long w = strtol(p, &endp, 10);
                if (endp == p) break;
                if (w > 0 && c->lane_weights.size() < 16) {
                    c->lane_weights.push_back((uint32_t)w);
                    wsum += (uint64_t)w;
                    wmax = std::max<uint64_t>(wmax, (uint64_t)w);
                }
                p = *endp ? endp + 1 : endp;
            }
            if (c->lane_weights.size() >= 2) {
                n_lanes = (int)c->lane_weights.size();
                c->lane_cap = (uint32_t)(((max_batch * wmax + wsum - 1) / wsum + 31) & ~15ull);
            } else {
                c->lane_weights.clear();
            }
        }
    }
    c->n_lanes = n_lanes;
    if ((rc = make_lanes(c, c->lanes))) return rc;
    CU(cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming));
    CU(cudaStreamCreateWithFlags(&c->copy_s, cudaStreamNonBlocking));
    /* A host-to-device copy of a few MB (one modulus of one lane) runs well
     * below the link rate on its own (tools/h2d_bandwidth.py: 22 GB/s at
     * 5 MiB, 55 GB/s from 188 MiB).  Spreading the copies over several
     * streams does not recover that: measured 7.38 / 8.83 / 8.91 / 9.34 ms per
     * end-to-end step with 1 / 2 / 3 / 4 copy streams
     * (tools/probe_copy_streams.py), so one stream stays the default;
     * DASHB_COPY_STREAMS keeps the experiment reproducible. */
    c->n_copy = 1;
    if (const char* env = getenv("DASHB_COPY_STREAMS")) {
        const int v = atoi(env);
        if (v >= 1 && v <= dashb_circuit::MAX_COPY) c->n_copy = v;
    }
    for (int i = 0; i + 1 < c->n_copy; ++i) CU(cudaStreamCreateWithFlags(&c->copy_x[i], cudaStreamNonBlocking));
    {
        int lo = 0, hi = 0;
        CU(cudaDeviceGetStreamPriorityRange(&lo, &hi));
        CU(cudaStreamCreateWithPriority(&c->ingest_s, cudaStreamNonBlocking, hi));
    }
    CU(c->d_values.alloc((size_t)max_batch * c->cur_size * 8));
    CU(c->d_ok.alloc(max_batch));
    c->ev.resize(c->layers.size() + 1);
    for (auto& e : c->ev) CU(cudaEventCreate(&e));
    /* the buffer clears above and every upload made while building ran on
     * the NULL stream; the evaluation streams are non-blocking, so wait here
     * once instead of racing the first evaluation */
    CU(cudaDeviceSynchronize());
    c->finalized = true;
    return DASHB_OK;
}

/* lane streams start after everything already enqueued on the library stream */
static int wait_async(dashb_circuit* c); /* drains dashb_evaluate_async steps in flight */

static int fork_lanes(dashb_circuit* c) {
    CU(cudaEventRecord(c->ev_fork, g_stream));
    for (int i = 0; i < c->lanes_used; ++i) CU(cudaStreamWaitEvent(c->lanes[i]->s, c->ev_fork, 0));
    CU(cudaStreamWaitEvent(c->copy_s, c->ev_fork, 0));
    for (int i = 0; i + 1 < c->n_copy; ++i) CU(cudaStreamWaitEvent(c->copy_x[i], c->ev_fork, 0));
    CU(cudaStreamWaitEvent(c->ingest_s, c->ev_fork, 0));
    return DASHB_OK;
}
/* ... and the library stream continues after all of them */
static int join_lanes(dashb_circuit* c) {
    for (int i = 0; i < c->lanes_used; ++i) {
        CU(cudaEventRecord(c->lanes[i]->done, c->lanes[i]->s));
        CU(cudaStreamWaitEvent(g_stream, c->lanes[i]->done, 0));
    }
    return DASHB_OK;
}

/* Sub-batch lanes of a batch.  Equal slices by default.  DASHB_LANE_SPLIT =
 * "w0,w1,..." cuts the batch in proportion to the weights instead (one lane
 * per weight, at most the lanes the circuit was finalized with): a small first
 * lane starts computing early in the pipelined dashb_evaluate, a small last
 * one shortens the tail after the last input has arrived. */
static int split_batch(dashb_circuit* c, uint64_t batch) {
    if (batch == 0 || batch > c->max_batch)
        return fail(DASHB_EINVAL, "batch %llu outside 1..%llu", (unsigned long long)batch, (unsigned long long)c->max_batch);
    const int n_lanes = c->n_lanes;
    c->batch = (uint32_t)batch;
    std::vector<uint32_t> sizes;
    if (!c->lane_weights.empty() && batch >= 16ull * c->lane_weights.size()) {
        uint64_t wsum = 0, given = 0;
        for (uint32_t w : c->lane_weights) wsum += w;
        for (size_t i = 0; i < c->lane_weights.size(); ++i) {
            uint64_t sz = i + 1 == c->lane_weights.size() ? batch - given : ((batch * c->lane_weights[i] / wsum) + 15) & ~15ull;
            sz = std::min<uint64_t>(std::min<uint64_t>(sz, batch - given), c->lane_cap);
            if (sz) sizes.push_back((uint32_t)sz);
            given += sz;
        }
        if (given != batch) sizes.clear(); /* the caps did not allow this split: equal slices */
    }
    if (sizes.empty()) {
        uint32_t per = (uint32_t)((((batch + n_lanes - 1) / n_lanes) + 15) & ~15ull);
        if (per > c->lane_cap) per = c->lane_cap;
        for (uint64_t b0 = 0; b0 < batch; b0 += per) sizes.push_back((uint32_t)std::min<uint64_t>(per, batch - b0));
    }
    c->lanes_used = (int)sizes.size();
    uint32_t b0 = 0;
    for (int i = 0; i < c->lanes_used; ++i) {
        Lane& ln = *c->lanes[i];
        ln.b0 = b0;
        ln.batch = sizes[i];
        b0 += sizes[i];
        ln.Bp = (ln.batch + 15) & ~15u;
        carve(ln.t_in, (uint8_t*)ln.buf_in.p, c->input_size, c->k, c->crt, ln.Bp);
        ln.t_cur = ln.t_in;
        ln.which = 0;
    }
    c->out_n = c->input_size;
    for (int j = 0; j < c->k; ++j) c->out_q[j] = c->crt[j];
    return DASHB_OK;
}

/* H2D copies of one lane's slice into its staging area (stream cs), then the
 * layout-changing ingest kernels (stream s).  With cs != s the copies of
 * all lanes run back to back on the copy engine while earlier lanes compute;
 * s waits for the lane's last copy. */
/* bytes one label of modulus index j takes on the host side of the boundary */
static size_t wire_label_bytes(const dashb_circuit* c, int j, bool compressed) {
    return compressed ? 16 : (size_t)c->crt_mi[j].nc * 2;
}

/* Device-visible addresses of the caller's label buffers when every one of
 * them is pinned host memory mapped into this device's address space
 * (cudaHostAlloc / cudaHostRegister under unified addressing): k_ingest_direct
 * then reads them over PCIe itself.  Pageable buffers (or DASHB_DIRECT_INGEST=0)
 * take the staged path: cudaMemcpyAsync into stage_in, then k_ingest. */
static bool mapped_host_inputs(const dashb_circuit* c, const void* const* in, const void** dev) {
    static const bool off = getenv("DASHB_DIRECT_INGEST") != nullptr && atoi(getenv("DASHB_DIRECT_INGEST")) == 0;
    if (off) return false;
    for (int j = 0; j < c->k; ++j) {
        cudaPointerAttributes a;
        if (cudaPointerGetAttributes(&a, in[j]) != cudaSuccess) {
            cudaGetLastError();
            return false;
        }
        if (a.type != cudaMemoryTypeHost || !a.devicePointer) return false;
        dev[j] = a.devicePointer;
    }
    return true;
}

/* `in`: per CRT modulus int16 [batch][label][nc], or (compressed) dashb_u128 [batch][label] */
static int enqueue_inputs(dashb_circuit* c, Lane& ln, const void* const* in, bool compressed, cudaStream_t cs, cudaStream_t s,
                          const void* const* mapped = nullptr, cudaEvent_t copied = nullptr) {
    /* timing experiments only (tools/probe_async.py): evaluate whatever the planes hold, move no input */
    static const bool skip = getenv("DASHB_DEBUG_SKIP_INPUTS") != nullptr;
    if (skip) {
        ln.t_cur = ln.t_in;
        ln.which = 0;
        return DASHB_OK;
    }
    if (mapped && !compressed) {
        /* no copy, no staging: one kernel per lane reads the host buffers; in the pipelined path it runs on
         * the high-priority stream so that its CTAs take the SM slots the computing lanes free up */
        cudaStream_t is = (cs != s) ? c->ingest_s : s;
        IngestParams p;
        memset(&p, 0, sizeof(p));
        p.k = c->k;
        p.n = c->input_size;
        p.batch = ln.batch;
        p.Bp = ln.Bp;
        for (int j = 0; j < c->k; ++j) {
            const size_t per = c->input_size * (size_t)c->crt_mi[j].nc;
            p.src[j] = (const int16_t*)mapped[j] + (size_t)ln.b0 * per;
            p.end[j] = (const int16_t*)mapped[j] + (size_t)c->batch * per;
            p.plane[j] = ln.t_in.plane[j];
            p.nc[j] = c->crt_mi[j].nc;
        }
        CU(launch_ingest_direct(p, is));
        c->launches += 1;
        if (is != s) {
            CU(cudaEventRecord(ln.ing_done, is));
            CU(cudaStreamWaitEvent(s, ln.ing_done, 0));
        }
        ln.t_cur = ln.t_in;
        ln.which = 0;
        return DASHB_OK;
    }
    size_t off = 0;
    for (int j = 0; j < c->k; ++j) {
        const size_t per = c->input_size * wire_label_bytes(c, j, compressed);
        const size_t bytes = (size_t)ln.batch * per;
        /* pipelined path: the moduli's copies alternate between the copy streams */
        const int which = (cs == c->copy_s) ? j % c->n_copy : 0;
        cudaStream_t cj = which ? c->copy_x[which - 1] : cs;
        CU(cudaMemcpyAsync((char*)ln.stage_in.p + off, (const char*)in[j] + (size_t)ln.b0 * per, bytes, cudaMemcpyHostToDevice, cj));
        off += bytes;
    }
    /* the ingest kernels are short; in the pipelined path they run on a
     * high-priority stream so that they are not queued behind the thousands
     * of CTAs of the lanes that already compute */
    cudaStream_t is = (cs != s) ? c->ingest_s : s;
    if (copied) CU(cudaEventRecord(copied, cs)); /* DASHB_TRACE_ASYNC */
    if (cs != s) {
        CU(cudaEventRecord(ln.in_done, cs));
        CU(cudaStreamWaitEvent(is, ln.in_done, 0));
        if (cs == c->copy_s)
            for (int i = 0; i + 1 < c->n_copy; ++i) {
                CU(cudaEventRecord(ln.in_done_x[i], c->copy_x[i]));
                CU(cudaStreamWaitEvent(is, ln.in_done_x[i], 0));
            }
    }
    off = 0;
    for (int j = 0; j < c->k; ++j) {
        const void* st = (const char*)ln.stage_in.p + off;
        if (compressed)
            CU(launch_ingest_compressed(st, ln.t_in.plane[j], c->input_size, c->crt_mi[j], ln.batch, ln.Bp, is));
        else
            CU(launch_ingest((const int16_t*)st, ln.t_in.plane[j], c->input_size, c->crt_mi[j].nc, ln.batch, ln.Bp, is));
        off += (size_t)ln.batch * c->input_size * wire_label_bytes(c, j, compressed);
    }
    if (is != s) {
        CU(cudaEventRecord(ln.ing_done, is));
        CU(cudaStreamWaitEvent(s, ln.ing_done, 0));
    }
    ln.t_cur = ln.t_in;
    ln.which = 0;
    return DASHB_OK;
}

static int move_inputs_any(dashb_circuit* c, const void* const* in, uint64_t batch, bool compressed) {
    if (!c || !in) return fail(DASHB_EINVAL, "null argument");
    if (!c->finalized) return fail(DASHB_ESTATE, "circuit not finalized");
    for (int j = 0; j < c->k; ++j)
        if (!in[j]) return fail(DASHB_EINVAL, "input labels %d missing", j);
    int rc = wait_async(c); /* the lanes and their input planes may still belong to a streamed step */
    if (rc) return rc;
    if ((rc = split_batch(c, batch))) return rc;
    const void* dev[MAXK];
    const bool direct = !compressed && mapped_host_inputs(c, in, dev);
    if ((rc = fork_lanes(c))) return rc;
    for (int i = 0; i < c->lanes_used; ++i)
        if ((rc = enqueue_inputs(c, *c->lanes[i], in, compressed, c->lanes[i]->s, c->lanes[i]->s, direct ? dev : nullptr))) return rc;
    return join_lanes(c);
}

int dashb_move_inputs(dashb_circuit* c, const int16_t* const* in, uint64_t batch) {
    return move_inputs_any(c, reinterpret_cast<const void* const*>(in), batch, false);
}

int dashb_move_inputs_compressed(dashb_circuit* c, const dashb_u128* const* in, uint64_t batch) {
    return move_inputs_any(c, reinterpret_cast<const void* const*>(in), batch, true);
}

int dashb_garble_inputs(dashb_circuit* c, const int64_t* values, uint64_t batch) {
    if (!c || !values) return fail(DASHB_EINVAL, "null argument");
    if (!c->finalized) return fail(DASHB_ESTATE, "circuit not finalized");
    if (!c->has_ginp) return fail(DASHB_ESTATE, "no input encoding (dashb_set_input_encoding)");
    int rc = wait_async(c);
    if (rc) return rc;
    if ((rc = split_batch(c, batch))) return rc;
    if ((rc = fork_lanes(c))) return rc;
    for (int i = 0; i < c->lanes_used; ++i) {
        Lane& ln = *c->lanes[i];
        CU(cudaMemcpyAsync(ln.xvals.p, values + (size_t)ln.b0 * c->input_size, (size_t)ln.batch * c->input_size * 8,
                           cudaMemcpyHostToDevice, ln.s));
        GarbleInParams p;
        memset(&p, 0, sizeof(p));
        p.x = (const int64_t*)ln.xvals.p;
        p.n = c->input_size;
        p.batch = ln.batch;
        p.Bp = ln.Bp;
        p.k = c->k;
        for (int j = 0; j < c->k; ++j) {
            p.out[j] = ln.t_in.plane[j];
            p.base[j] = (const uint8_t*)c->in_base[j].p;
            p.off[j] = (const uint8_t*)c->in_off[j].p;
            p.crt[j] = c->crt_mi[j];
        }
        CU(launch_garble_inputs(p, ln.s));
    }
    return join_lanes(c);
}

/* enqueues layer li of one lane on stream s; input = the lane's current tensor */
static int run_layer(dashb_circuit* c, Lane& ln, int li, cudaStream_t s) {
    Layer& L = *c->layers[li];
    const Tensor cur = ln.t_cur;
    uint8_t* bufs[2] = {(uint8_t*)ln.buf_a.p, (uint8_t*)ln.buf_b.p};
    const uint32_t B = ln.batch, Bp = ln.Bp;
    Tensor out;
    carve(out, bufs[ln.which], L.out_size, c->k, L.out_q, Bp);
    ln.which ^= 1;
    switch (L.type) {
        case DASHB_LAYER_DENSE:
        case DASHB_LAYER_CONV2D: {
            if (L.use_tc) {
                DenseTcMaps maps;
                DenseTcParams tp;
                memset(&tp, 0, sizeof(tp));
                tp.k = c->k;
                tp.M = (uint32_t)L.M;
                tp.K = (uint32_t)L.K;
                tp.Bp = Bp;
                tp.v0 = L.v0;
                tp.v1 = L.v1;
                tp.p0 = L.p0;
                tp.p1 = L.p1;
                tp.nkb = L.tc_nkb;
                const bool conv = L.type == DASHB_LAYER_CONV2D;
                tp.conv = conv;
                tp.npix = L.npix;
                tp.ow = L.ow;
                tp.sx = L.cv_sw;
                tp.sy = L.cv_sh;
                tp.cb = L.cv_cb;
                tp.b_box_bytes = conv ? 128u * L.cv_fw * L.cv_fh * L.cv_cb : 128u * 128u;
                /* narrow convolutions: filters on the N dimension of the MMA (dense_tc.cu: k_conv_tc_n) */
                static const bool conv_n_off = getenv("DASHB_CONV_N") != nullptr && atoi(getenv("DASHB_CONV_N")) == 0;
                /* measured equal or better up to 32 filters, slower above (both are L2->SM bound); the
                 * kernel itself handles 32 | filters <= 128 (DASHB_CONV_N_MAX for experiments and tests) */
                static const int conv_n_max = getenv("DASHB_CONV_N_MAX") ? atoi(getenv("DASHB_CONV_N_MAX")) : 32;
                if (conv && !conv_n_off && L.M % 32 == 0 && L.M <= (uint64_t)conv_n_max) tp.npad = (uint32_t)L.M;
                const bool strip = conv && L.st_strip != 0;
                if (strip) { /* k_conv_tc_strip: a strip of output pixels per CTA, filters on N */
                    tp.npad = (uint32_t)L.M;
                    tp.strip = L.st_strip;
                    tp.st_rows = L.st_rows;
                    tp.st_cblk = L.st_cblk;
                    tp.st_ncb = L.st_ncb;
                    tp.st_fw = L.cv_fw;
                    tp.st_wboxes = L.st_wboxes;
                }
                for (int j = 0; j < c->k; ++j) {
                    maps.a[j] = strip ? L.map_st[j] : L.map_a[j];
                    const uint64_t Nj = (uint64_t)((c->crt_mi[j].nc + 3) & ~3u) * Bp;
                    if (strip)
                        CU(conv_tc_make_map(&maps.b[j], cur.plane[j], Nj, L.cv_w, L.cv_h, L.cv_c, 1, L.cv_fh, L.st_cblk));
                    else if (conv)
                        CU(conv_tc_make_map(&maps.b[j], cur.plane[j], Nj, L.cv_w, L.cv_h, L.cv_c, L.cv_fw, L.cv_fh, L.cv_cb));
                    else
                        CU(dense_tc_make_map(&maps.b[j], cur.plane[j], Nj, L.K));
                    tp.out[j] = out.plane[j];
                    tp.cst[j] = (const uint8_t*)L.cst[j].p;
                    tp.mod[j] = c->crt_mi[j];
                }
                CU(launch_dense_tc(maps, tp, s));
                c->launches += 1;
                break;
            }
            LinearParams p;
            memset(&p, 0, sizeof(p));
            p.k = c->k;
            for (int j = 0; j < c->k; ++j) {
                p.in[j] = cur.plane[j];
                p.out[j] = out.plane[j];
                p.w4[j] = (const uint32_t*)L.w4[j].p;
                p.cst[j] = (const uint8_t*)L.cst[j].p;
                p.mod[j] = c->crt_mi[j];
            }
            p.koff = (const int32_t*)L.koff.p;
            p.M = (uint32_t)L.M;
            p.v0 = L.v0;
            p.v1 = L.v1;
            p.p0 = L.p0;
            p.p1 = L.p1;
            p.K = (uint32_t)L.K;
            p.Kp = L.Kp;
            p.npix = L.npix;
            p.ow = L.ow;
            p.step_x = L.step_x;
            p.step_y = L.step_y;
            p.Bp = Bp;
            CU(launch_linear(p, s));
            c->launches += 1;
            break;
        }
        case DASHB_LAYER_RELU:
        case DASHB_LAYER_SIGN: {
            const bool relu = L.type == DASHB_LAYER_RELU;
            SignParams p;
            fill_sign_params(c, p, L.slice_count(), relu ? 1 : c->k);
            p.e0 = L.e0;
            p.seg = L.seg;
            p.stride = L.stride;
            for (int j = 0; j < c->k; ++j) {
                p.in[j] = cur.plane[j];
                p.out[j] = out.plane[j];
                p.outm[j] = c->crt_mi[j];
            }
            p.approx = (const u128*)L.approx.p;
            p.cast1 = (const u128*)L.cast1.p;
            p.cast2 = (const u128*)L.cast2.p;
            p.sign = (const u128*)L.sign.p;
            p.garbler = (const u128*)L.garbler.p;
            p.evaluator = (const u128*)L.evaluator.p;
            p.batch = B;
            p.Bp = Bp;
            p.b0 = ln.b0;
            CU(launch_sign(p, relu, s));
            c->launches += 1;
            break;
        }
        case DASHB_LAYER_RESCALE:
        case DASHB_LAYER_RESCALE_LEGACY: {
            const bool legacy = L.type == DASHB_LAYER_RESCALE_LEGACY;
            const int iters = legacy ? L.l_iter : 1;
            for (int it = 0; it < iters; ++it) {
                RescaleParams p = L.rp;
                for (int j = 0; j < c->k; ++j) {
                    p.in[j] = it == 0 ? cur.plane[j] : out.plane[j];
                    p.out[j] = out.plane[j];
                    p.up[j] = (const uint8_t*)c->d_up[j].p;
                    p.down[j] = (const uint8_t*)L.down[j].p;
                    p.zero[j] = (const uint8_t*)c->d_zero[j].p;
                }
                p.tm = (const u128*)(legacy ? L.it_tables[5 * it]->p : L.tm.p);
                p.be = (const u128*)L.be.p;
                p.e0 = L.e0;
                p.n = L.slice_count();
                p.seg = L.seg;
                p.stride = L.stride;
                p.batch = B;
                p.Bp = Bp;
                CU(launch_rescale(p, s));
                c->launches += 1;
                if (legacy) {
                    /* base extension by sign gadget: regenerates residue 0 in place */
                    SignParams sp;
                    fill_sign_params(c, sp, L.e1 - L.e0, 1);
                    sp.e0 = L.e0;
                    for (int j = 0; j < c->k; ++j) sp.in[j] = out.plane[j];
                    sp.out[0] = out.plane[0];
                    sp.outm[0] = c->crt_mi[0];
                    sp.approx = (const u128*)L.it_tables[5 * it + 1]->p;
                    sp.cast1 = (const u128*)L.it_tables[5 * it + 2]->p;
                    sp.cast2 = (const u128*)L.it_tables[5 * it + 3]->p;
                    sp.sign = (const u128*)L.it_tables[5 * it + 4]->p;
                    sp.batch = B;
                    sp.Bp = Bp;
                    CU(launch_sign(sp, false, s));
                    AddLabelParams ap;
                    memset(&ap, 0, sizeof(ap));
                    ap.k = c->k;
                    ap.n = L.e1 - L.e0;
                    ap.e0 = L.e0;
                    ap.Bp = Bp;
                    ap.subtract = 1;
                    for (int j = 0; j < c->k; ++j) {
                        ap.plane[j] = out.plane[j];
                        ap.label[j] = (const uint8_t*)L.down[j].p;
                        ap.crt[j] = c->crt_mi[j];
                    }
                    CU(launch_add_label(ap, s));
                    c->launches += 2;
                }
            }
            break;
        }
        case DASHB_LAYER_BASE_EXTENSION: {
            /* k_rescale without its division step and shifts: the other residues pass through (and are hashed
             * as the sources of the mixed-radix conversion), the extra residues start from the layer's labels */
            RescaleParams p = L.rp;
            for (int j = 0; j < c->k; ++j) {
                p.in[j] = cur.plane[j];
                p.out[j] = out.plane[j];
                p.up[j] = (const uint8_t*)c->d_nolabel.p;
                p.down[j] = nullptr;
                p.zero[j] = (const uint8_t*)(L.start[j].p ? L.start[j].p : c->d_zero[j].p);
                p.zstride[j] = L.start[j].p ? ((c->crt_mi[j].nc + 3) & ~3u) : 0u;
            }
            p.tm = nullptr;
            p.be = (const u128*)L.be.p;
            p.e0 = L.e0;
            p.n = L.slice_count();
            p.seg = L.seg;
            p.stride = L.stride;
            p.batch = B;
            p.Bp = Bp;
            CU(launch_rescale(p, s));
            c->launches += 1;
            break;
        }
        case DASHB_LAYER_MAXPOOL2D: {
            /* k_w k_h MaxGadget steps per output, one after the other (the running maximum feeds the next):
             * k_pool_step forms y - x, k_sign<relu> evaluates that step's ReLU gadgets, the next
             * k_pool_step adds the result */
            Tensor diff, relu;
            const size_t half = tensor_bytes(L.out_size, c->k, L.out_q, Bp);
            carve(diff, (uint8_t*)ln.pool.p, L.out_size, c->k, L.out_q, Bp);
            carve(relu, (uint8_t*)ln.pool.p + half, L.out_size, c->k, L.out_q, Bp);
            PoolParams pp;
            memset(&pp, 0, sizeof(pp));
            pp.k = c->k;
            pp.n_out = L.out_size;
            pp.Bp = Bp;
            pp.in_w = L.pl_w, pp.in_h = L.pl_h, pp.ch = L.pl_c, pp.ow = L.pl_ow, pp.oh = L.pl_oh, pp.s_w = L.pl_sw, pp.s_h = L.pl_sh;
            for (int j = 0; j < c->k; ++j) {
                pp.in[j] = cur.plane[j];
                pp.cur[j] = out.plane[j];
                pp.diff[j] = diff.plane[j];
                pp.relu[j] = relu.plane[j];
                pp.crt[j] = c->crt_mi[j];
            }
            const uint32_t T = L.pl_kw * L.pl_kh;
            for (uint32_t t = 0; t < T; ++t) {
                pp.l = t / L.pl_kh;
                pp.m = t % L.pl_kh;
                pp.mode = t == 0 ? 0 : 1;
                CU(launch_pool_step(pp, s));
                SignParams sp;
                fill_sign_params(c, sp, L.out_size, 1);
                sp.tab_mul = T;
                sp.tab_add = t;
                for (int j = 0; j < c->k; ++j) {
                    sp.in[j] = diff.plane[j];
                    sp.out[j] = relu.plane[j];
                    sp.outm[j] = c->crt_mi[j];
                }
                sp.approx = (const u128*)L.approx.p;
                sp.cast1 = (const u128*)L.cast1.p;
                sp.cast2 = (const u128*)L.cast2.p;
                sp.sign = (const u128*)L.sign.p;
                sp.garbler = (const u128*)L.garbler.p;
                sp.evaluator = (const u128*)L.evaluator.p;
                sp.batch = B;
                sp.Bp = Bp;
                CU(launch_sign(sp, true, s));
                c->launches += 2;
            }
            pp.mode = 2;
            CU(launch_pool_step(pp, s));
            c->launches += 1;
            break;
        }
        case DASHB_LAYER_PROJECTION: {
            ProjParams p;
            memset(&p, 0, sizeof(p));
            p.k = c->k;
            p.n = L.slice_count();
            p.e0 = L.e0;
            p.seg = L.seg;
            p.stride = L.stride;
            p.batch = B;
            p.Bp = Bp;
            for (int j = 0; j < c->k; ++j) {
                p.in[j] = cur.plane[j];
                p.out[j] = out.plane[j];
                p.min[j] = make_modinfo((uint32_t)L.in_q[j]);
                p.mout[j] = make_modinfo((uint32_t)L.out_q[j]);
                p.ciphers[j] = (const u128*)L.proj[j].p;
            }
            CU(launch_projection(p, s));
            c->launches += 1;
            break;
        }
        default: return fail(DASHB_EINVAL, "unknown layer type");
    }
    ln.t_cur = out;
    return DASHB_OK;
}

/* Layers 0..upto for every lane, layer-major so that the lanes' launches of
 * one layer sit next to each other.  With per-layer timing everything runs
 * on the library stream (serialised) so that event pairs delimit a layer. */
static int run_layers(dashb_circuit* c, int upto) {
    int rc;
    c->launches = 0;
    for (int i = 0; i < c->lanes_used; ++i) {
        c->lanes[i]->t_cur = c->lanes[i]->t_in;
        c->lanes[i]->which = 0;
    }
    if (c->timing == 1) {
        CU(cudaEventRecord(c->ev[0], g_stream));
        for (int li = 0; li <= upto; ++li) {
            for (int i = 0; i < c->lanes_used; ++i)
                if ((rc = run_layer(c, *c->lanes[i], li, g_stream))) return rc;
            CU(cudaEventRecord(c->ev[li + 1], g_stream));
        }
    } else {
        /* mode 2: the step as it always runs (lanes concurrent), with an event
         * on the lane's stream before and after every layer */
        const bool trace = c->timing == 2;
        const size_t nl = c->layers.size();
        if (trace) {
            const size_t need = (size_t)c->n_lanes * nl * 2;
            while (c->iv.size() < need) {
                cudaEvent_t e;
                CU(cudaEventCreate(&e));
                c->iv.push_back(e);
            }
            if (!c->iv_t0) CU(cudaEventCreate(&c->iv_t0));
            CU(cudaEventRecord(c->iv_t0, g_stream));
        }
        if ((rc = fork_lanes(c))) return rc;
        for (int li = 0; li <= upto; ++li)
            for (int i = 0; i < c->lanes_used; ++i) {
                if (trace) CU(cudaEventRecord(c->iv[(i * nl + li) * 2], c->lanes[i]->s));
                if ((rc = run_layer(c, *c->lanes[i], li, c->lanes[i]->s))) return rc;
                if (trace) CU(cudaEventRecord(c->iv[(i * nl + li) * 2 + 1], c->lanes[i]->s));
            }
        if ((rc = join_lanes(c))) return rc;
    }
    if (upto >= 0) {
        c->out_n = c->layers[upto]->out_size;
        for (int j = 0; j < c->k; ++j) c->out_q[j] = c->layers[upto]->out_q[j];
    }
    return DASHB_OK;
}

/* Channel-sharded evaluation: layers first..last on this rank's slices,
 * continuing from the current tensors (first == 0 restarts from the inputs).
 * The caller gathers the slices of all ranks (dashb_current_plane) before a
 * layer that needs every input element. */
int dashb_evaluate_layers(dashb_circuit* c, int first, int last) {
    if (!c) return fail(DASHB_EINVAL, "null circuit");
    if (!c->finalized || c->batch == 0) return fail(DASHB_ESTATE, "no inputs moved");
    if (first < 0 || last < first || last >= (int)c->layers.size()) return fail(DASHB_EINVAL, "bad layer range");
    if (c->lanes_used != 1) return fail(DASHB_ESTATE, "dashb_evaluate_layers needs a single lane (sharded circuit or DASHB_LANES=1)");
    Lane& ln = *c->lanes[0];
    if (first == 0) {
        ln.t_cur = ln.t_in;
        ln.which = 0;
        c->launches = 0;
    }
    for (int li = first; li <= last; ++li) {
        int rc = run_layer(c, ln, li, g_stream);
        if (rc) return rc;
    }
    c->out_n = c->layers[last]->out_size;
    for (int j = 0; j < c->k; ++j) c->out_q[j] = c->layers[last]->out_q[j];
    return DASHB_OK;
}

int dashb_set_table_replicas(dashb_circuit* c, int replicas) {
    int rc = check_building(c);
    if (rc) return rc;
    if (!c->layers.empty()) return fail(DASHB_ESTATE, "set the table replicas before the first layer is added");
    if (replicas < 1 || replicas > 65536) return fail(DASHB_EINVAL, "replicas %d outside 1..65536", replicas);
    c->table_replicas = replicas;
    return DASHB_OK;
}

int dashb_set_shard_lanes(dashb_circuit* c, int lanes) {
    int rc = check_building(c);
    if (rc) return rc;
    if (lanes < 1 || lanes > 8) return fail(DASHB_EINVAL, "shard lanes %d outside 1..8", lanes);
    if (c->shard_world > MAXR) return fail(DASHB_EINVAL, "peer exchange supports up to %d ranks", MAXR);
    c->shard_lanes = lanes;
    return DASHB_OK;
}

int dashb_shard_lanes_used(const dashb_circuit* c) { return c ? c->lanes_used : 0; }
void* dashb_lane_stream(const dashb_circuit* c, int lane) {
    return (c && lane >= 0 && lane < (int)c->lanes.size()) ? (void*)c->lanes[lane]->s : nullptr;
}

/* 3 x 64 bytes: the IPC handles of lane `lane`'s two gather buffers and its flag array */
int dashb_shard_export(dashb_circuit* c, int lane, void* handles) {
    if (!c || !handles) return fail(DASHB_EINVAL, "null argument");
    if (!c->finalized || c->shard_lanes < 1 || lane < 0 || lane >= c->n_lanes) return fail(DASHB_ESTATE, "no such sharded lane");
    Lane& ln = *c->lanes[lane];
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "handle size");
    cudaIpcMemHandle_t* h = (cudaIpcMemHandle_t*)handles;
    CU(cudaIpcGetMemHandle(&h[0], ln.gat[0].p));
    CU(cudaIpcGetMemHandle(&h[1], ln.gat[1].p));
    CU(cudaIpcGetMemHandle(&h[2], ln.flags.p));
    return DASHB_OK;
}

int dashb_shard_import(dashb_circuit* c, int lane, int peer_rank, const void* handles) {
    if (!c || !handles) return fail(DASHB_EINVAL, "null argument");
    if (!c->finalized || c->shard_lanes < 1 || lane < 0 || lane >= c->n_lanes) return fail(DASHB_ESTATE, "no such sharded lane");
    if (peer_rank < 0 || peer_rank >= c->shard_world || peer_rank == c->shard_rank) return fail(DASHB_EINVAL, "bad peer rank %d", peer_rank);
    Lane& ln = *c->lanes[lane];
    const cudaIpcMemHandle_t* h = (const cudaIpcMemHandle_t*)handles;
    void* p[3] = {nullptr, nullptr, nullptr};
    for (int i = 0; i < 3; ++i) {
        CU(cudaIpcOpenMemHandle(&p[i], h[i], cudaIpcMemLazyEnablePeerAccess));
        c->ipc_opened.push_back(p[i]);
    }
    ln.peer_gat[0][peer_rank] = (uint8_t*)p[0];
    ln.peer_gat[1][peer_rank] = (uint8_t*)p[1];
    ln.peer_flags[peer_rank] = (uint32_t*)p[2];
    return DASHB_OK;
}

/* layers first..last of one lane on that lane's stream */
int dashb_evaluate_layers_lane(dashb_circuit* c, int lane, int first, int last) {
    if (!c) return fail(DASHB_EINVAL, "null circuit");
    if (!c->finalized || c->batch == 0) return fail(DASHB_ESTATE, "no inputs moved");
    if (first < 0 || last < first || last >= (int)c->layers.size()) return fail(DASHB_EINVAL, "bad layer range");
    if (lane < 0 || lane >= c->lanes_used) return fail(DASHB_EINVAL, "bad lane %d", lane);
    Lane& ln = *c->lanes[lane];
    if (first == 0) {
        /* behind the inputs (dashb_move_inputs / dashb_garble_inputs joined the library stream) */
        CU(cudaEventRecord(ln.in_done, g_stream));
        CU(cudaStreamWaitEvent(ln.s, ln.in_done, 0));
        ln.t_cur = ln.t_in;
        ln.which = 0;
        if (lane == 0) c->launches = 0;
    }
    for (int li = first; li <= last; ++li) {
        int rc = run_layer(c, ln, li, ln.s);
        if (rc) return rc;
    }
    c->out_n = c->layers[last]->out_size;
    for (int j = 0; j < c->k; ++j) c->out_q[j] = c->layers[last]->out_q[j];
    return DASHB_OK;
}

/* DASHB_SHARD_COMPRESS=0: the exchange moves digit planes as they are instead of compressed labels */
static bool shard_compressed() {
    static const bool on = !(getenv("DASHB_SHARD_COMPRESS") != nullptr && atoi(getenv("DASHB_SHARD_COMPRESS")) == 0);
    return on;
}

/* After layer `layer`: this rank's slice of the lane's current tensor goes to
 * every rank's gather buffer; the lane's stream continues once every rank's
 * slice has arrived here.  The lane's current tensor is then the gathered one. */
int dashb_shard_exchange(dashb_circuit* c, int lane, int layer) {
    if (!c) return fail(DASHB_EINVAL, "null circuit");
    if (!c->finalized || c->shard_lanes < 1 || c->batch == 0) return fail(DASHB_ESTATE, "no sharded evaluation in flight");
    if (lane < 0 || lane >= c->lanes_used || layer < 0 || layer >= (int)c->layers.size()) return fail(DASHB_EINVAL, "bad lane or layer");
    Lane& ln = *c->lanes[lane];
    const int W = c->shard_world, x = ln.gcur;
    for (int r = 0; r < W; ++r)
        if (!ln.peer_gat[x][r] || !ln.peer_flags[r]) return fail(DASHB_ESTATE, "rank %d's buffers were not imported (dashb_shard_import)", r);
    const Tensor cur = ln.t_cur;
    const Layer& SL = *c->layers[layer];
    if (shard_compressed()) {
        /* labels cross NVLink as 16 bytes each; the gathered tensor is expanded back into the lane's own planes */
        ShardCompParams q;
        memset(&q, 0, sizeof(q));
        q.k = cur.k;
        q.world = W;
        q.n = cur.n;
        q.e0 = SL.e0;
        q.seg = SL.seg;
        q.stride = SL.stride;
        q.count = SL.slice_count();
        q.Bp = ln.Bp;
        q.local = (const uint8_t*)ln.gat[x].p;
        for (int j = 0; j < cur.k; ++j) {
            q.plane[j] = cur.plane[j];
            q.mod[j] = make_modinfo((uint32_t)cur.q[j]);
        }
        for (int r = 0; r < W; ++r) q.peer[r] = ln.peer_gat[x][r];
        const uint32_t epoch = ++ln.epoch;
        CU(launch_shard_push_comp(q, ln.s));
        CU(launch_shard_signal(ln.peer_flags, W, c->shard_rank, epoch, ln.s));
        CU(launch_shard_wait((uint32_t*)ln.flags.p, W, epoch, ln.s));
        CU(launch_shard_unpack(q, ln.s));
        c->launches += 4;
        ln.gcur ^= 1;
        return DASHB_OK;
    }
    Tensor dst;
    carve(dst, (uint8_t*)ln.gat[x].p, cur.n, cur.k, cur.q, ln.Bp);
    ShardPushParams p;
    memset(&p, 0, sizeof(p));
    p.k = cur.k;
    p.world = W;
    p.e0 = SL.e0;
    p.e1 = SL.e1;
    p.seg = SL.seg;
    p.stride = SL.stride;
    p.nseg = SL.nseg;
    for (int j = 0; j < cur.k; ++j) {
        p.src[j] = cur.plane[j];
        p.dst_off[j] = (uint64_t)(dst.plane[j] - (uint8_t*)ln.gat[x].p);
        p.row[j] = (uint64_t)((nr_comps(cur.q[j]) + 3) & ~3) * ln.Bp;
    }
    for (int r = 0; r < W; ++r) p.peer[r] = ln.peer_gat[x][r];
    const uint32_t epoch = ++ln.epoch;
    CU(launch_shard_push(p, ln.s));
    CU(launch_shard_signal(ln.peer_flags, W, c->shard_rank, epoch, ln.s));
    CU(launch_shard_wait((uint32_t*)ln.flags.p, W, epoch, ln.s));
    c->launches += 3;
    ln.t_cur = dst;
    ln.gcur ^= 1;
    return DASHB_OK;
}

/* Every layer group of every lane with its exchange, in the order all ranks
 * must follow: a linear layer (dense / conv2d) needs all of its inputs, so
 * each one starts a group; the element-wise layers behind it stay with it.
 * Asynchronous; dashb_move_outputs / dashb_decode_outputs order themselves
 * behind it. */
int dashb_evaluate_sharded(dashb_circuit* c) {
    if (!c) return fail(DASHB_EINVAL, "null circuit");
    if (!c->finalized || c->shard_lanes < 1 || c->batch == 0) return fail(DASHB_ESTATE, "no inputs moved, or no peer exchange prepared");
    const int nl = (int)c->layers.size();
    int first = 0;
    for (int li = 0; li <= nl; ++li) {
        const bool linear = li < nl && (c->layers[li]->type == DASHB_LAYER_DENSE || c->layers[li]->type == DASHB_LAYER_CONV2D);
        if ((linear && li > first) || li == nl) {
            if (li == first) break;
            for (int lane = 0; lane < c->lanes_used; ++lane) {
                int rc = dashb_evaluate_layers_lane(c, lane, first, li - 1);
                if (rc) return rc;
            }
            if (c->shard_world > 1)
                for (int lane = 0; lane < c->lanes_used; ++lane) {
                    int rc = dashb_shard_exchange(c, lane, li - 1);
                    if (rc) return rc;
                }
            first = li;
        }
    }
    return DASHB_OK;
}

/* bytes this rank receives from its peers in dashb_shard_exchange(lane, layer) */
uint64_t dashb_shard_exchange_bytes(dashb_circuit* c, int lane, int layer) {
    if (!c || !c->finalized || lane < 0 || lane >= c->lanes_used || layer < 0 || layer >= (int)c->layers.size()) return 0;
    Lane& ln = *c->lanes[lane];
    const Layer& L = *c->layers[layer];
    uint64_t per = 0;
    for (int j = 0; j < c->k; ++j) per += (shard_compressed() ? 16ull : (uint64_t)((nr_comps(L.out_q[j]) + 3) & ~3)) * ln.Bp;
    return (L.out_size - L.slice_count()) * per;
}

/* 0, or the epoch at which a lane's wait for its peers timed out */
int dashb_shard_status(dashb_circuit* c, int lane, uint32_t* timed_out_epoch) {
    if (!c || !timed_out_epoch || lane < 0 || lane >= c->n_lanes || !c->lanes[lane]->flags.p) return fail(DASHB_EINVAL, "bad argument");
    CU(cudaMemcpy(timed_out_epoch, (uint32_t*)c->lanes[lane]->flags.p + MAXR, 4, cudaMemcpyDeviceToHost));
    return DASHB_OK;
}

int dashb_current_plane(dashb_circuit* c, int j, void** dev_ptr, uint64_t* bytes_per_element, uint64_t* n_elements) {
    if (!c || !dev_ptr || !bytes_per_element || !n_elements) return fail(DASHB_EINVAL, "null argument");
    if (!c->finalized || c->batch == 0 || c->lanes_used != 1) return fail(DASHB_ESTATE, "no single-lane evaluation in flight");
    if (j < 0 || j >= c->k) return fail(DASHB_EINVAL, "bad modulus index");
    Lane& ln = *c->lanes[0];
    *dev_ptr = ln.t_cur.plane[j];
    *bytes_per_element = (uint64_t)((nr_comps(ln.t_cur.q[j]) + 3) & ~3) * ln.Bp;
    *n_elements = ln.t_cur.n;
    return DASHB_OK;
}

/* Elements [e0, e1) of plane j of the current tensor <-> compressed labels
 * in device memory, dashb_u128 [batch][e1 - e0]; enqueued on the library
 * stream.  The exchange of output-channel sharding sends these 16 bytes per
 * label instead of the digit planes. */
static int slice_args(dashb_circuit* c, int j, uint64_t e0, uint64_t e1, const void* dev) {
    if (!c || !dev) return fail(DASHB_EINVAL, "null argument");
    if (!c->finalized || c->batch == 0 || c->lanes_used != 1) return fail(DASHB_ESTATE, "no single-lane evaluation in flight");
    if (j < 0 || j >= c->k || e1 < e0 || e1 > c->lanes[0]->t_cur.n) return fail(DASHB_EINVAL, "bad modulus index or element range");
    return DASHB_OK;
}

int dashb_current_slice_compress(dashb_circuit* c, int j, uint64_t e0, uint64_t e1, void* dev_dst) {
    int rc = slice_args(c, j, e0, e1, dev_dst);
    if (rc) return rc;
    Lane& ln = *c->lanes[0];
    const ModInfo m = make_modinfo((uint32_t)ln.t_cur.q[j]);
    const uint64_t row = (uint64_t)((m.nc + 3) & ~3u) * ln.Bp;
    CU(launch_egress_compressed(ln.t_cur.plane[j] + e0 * row, dev_dst, e1 - e0, m, ln.batch, ln.Bp, g_stream));
    return DASHB_OK;
}

int dashb_current_slice_decompress(dashb_circuit* c, int j, uint64_t e0, uint64_t e1, const void* dev_src) {
    int rc = slice_args(c, j, e0, e1, dev_src);
    if (rc) return rc;
    Lane& ln = *c->lanes[0];
    const ModInfo m = make_modinfo((uint32_t)ln.t_cur.q[j]);
    const uint64_t row = (uint64_t)((m.nc + 3) & ~3u) * ln.Bp;
    CU(launch_ingest_compressed(dev_src, ln.t_cur.plane[j] + e0 * row, e1 - e0, m, ln.batch, ln.Bp, g_stream));
    return DASHB_OK;
}

int dashb_layer_slice(const dashb_circuit* c, int layer, int rank, uint64_t* e0, uint64_t* e1) {
    if (!c || !e0 || !e1 || layer < 0 || layer >= (int)c->layers.size() || rank < 0 || rank >= c->shard_world)
        return fail(DASHB_EINVAL, "bad argument");
    if (pixel_sharding(c) && c->layers[layer]->seg)
        return fail(DASHB_ESTATE, "layer %d is sharded by output pixels (peer-memory exchange): its slice is not one contiguous range", layer);
    uint32_t v0, v1;
    layer_slice(c, layer, rank, e0, e1, &v0, &v1);
    return DASHB_OK;
}

int dashb_evaluate_device(dashb_circuit* c) {
    if (!c) return fail(DASHB_EINVAL, "null circuit");
    if (!c->finalized || c->batch == 0) return fail(DASHB_ESTATE, "no inputs moved");
    if (c->shard_world > 1)
        return fail(DASHB_ESTATE, "channel-sharded circuit: use dashb_evaluate_layers and gather between layer groups");
    int rc = wait_async(c);
    if (rc) return rc;
    if ((rc = run_layers(c, (int)c->layers.size() - 1))) return rc;
    if (c->timing == 1) {
        CU(cudaStreamSynchronize(g_stream));
        for (size_t i = 0; i < c->layers.size(); ++i) CU(cudaEventElapsedTime(&c->layers[i]->ms, c->ev[i], c->ev[i + 1]));
    } else if (c->timing == 2) {
        CU(cudaStreamSynchronize(g_stream));
        const size_t nl = c->layers.size();
        c->iv_ms.assign((size_t)c->n_lanes * nl * 2, 0.f);
        for (int i = 0; i < c->lanes_used; ++i)
            for (size_t li = 0; li < nl; ++li)
                for (int x = 0; x < 2; ++x)
                    CU(cudaEventElapsedTime(&c->iv_ms[(i * nl + li) * 2 + x], c->iv_t0, c->iv[(i * nl + li) * 2 + x]));
    }
    return DASHB_OK;
}

/* Device-visible addresses of the caller's output buffers when all of them are mapped pinned host memory
 * (see mapped_host_inputs): the egress kernels then store into them directly, no staging copy.
 * DASHB_DIRECT_EGRESS=0 keeps the staged path. */
static bool mapped_host_outputs(const dashb_circuit* c, void* const* out, void** dev) {
    static const bool off = getenv("DASHB_DIRECT_EGRESS") != nullptr && atoi(getenv("DASHB_DIRECT_EGRESS")) == 0;
    if (off) return false;
    for (int j = 0; j < c->k; ++j) {
        cudaPointerAttributes a;
        if (cudaPointerGetAttributes(&a, out[j]) != cudaSuccess) {
            cudaGetLastError();
            return false;
        }
        if (a.type != cudaMemoryTypeHost || !a.devicePointer) return false;
        dev[j] = a.devicePointer;
    }
    return true;
}

static int enqueue_outputs(dashb_circuit* c, Lane& ln, void* const* out, bool compressed, cudaStream_t s) {
    void* dev[MAXK];
    const bool direct = mapped_host_outputs(c, out, dev);
    size_t off = 0;
    EgressParams ep;
    memset(&ep, 0, sizeof(ep));
    ep.k = c->k;
    ep.n = ln.t_cur.n;
    ep.batch = ln.batch;
    ep.Bp = ln.Bp;
    size_t pers[MAXK], offs[MAXK];
    for (int j = 0; j < c->k; ++j) {
        const ModInfo m = make_modinfo((uint32_t)ln.t_cur.q[j]);
        const size_t per = ln.t_cur.n * (compressed ? 16 : (size_t)m.nc * 2);
        pers[j] = per;
        offs[j] = off;
        void* dst = direct ? (void*)((char*)dev[j] + (size_t)ln.b0 * per) : (void*)((char*)ln.stage_out.p + off);
        if (compressed) {
            CU(launch_egress_compressed(ln.t_cur.plane[j], dst, ln.t_cur.n, m, ln.batch, ln.Bp, s));
            c->launches += 1;
        } else {
            ep.plane[j] = ln.t_cur.plane[j];
            ep.dst[j] = (int16_t*)dst;
            ep.nc[j] = m.nc;
        }
        off += (size_t)ln.batch * per;
    }
    if (!compressed) {
        CU(launch_egress_all(ep, s));
        c->launches += 1;
    }
    if (!direct)
        for (int j = 0; j < c->k; ++j)
            CU(cudaMemcpyAsync((char*)out[j] + (size_t)ln.b0 * pers[j], (char*)ln.stage_out.p + offs[j], (size_t)ln.batch * pers[j],
                               cudaMemcpyDeviceToHost, s));
    return DASHB_OK;
}

static int copy_out(dashb_circuit* c, void* const* out, bool compressed) {
    for (int j = 0; j < c->k; ++j)
        if (!out[j]) return fail(DASHB_EINVAL, "output buffer %d missing", j);
    int rc = fork_lanes(c);
    if (rc) return rc;
    for (int i = 0; i < c->lanes_used; ++i)
        if ((rc = enqueue_outputs(c, *c->lanes[i], out, compressed, c->lanes[i]->s))) return rc;
    if ((rc = join_lanes(c))) return rc;
    CU(cudaStreamSynchronize(g_stream));
    return DASHB_OK;
}

int dashb_move_outputs_compressed(dashb_circuit* c, dashb_u128* const* out) {
    if (!c || !out) return fail(DASHB_EINVAL, "null argument");
    if (!c->finalized || c->batch == 0) return fail(DASHB_ESTATE, "nothing evaluated");
    return copy_out(c, reinterpret_cast<void* const*>(out), true);
}

int dashb_move_outputs(dashb_circuit* c, int16_t* const* out) {
    if (!c || !out) return fail(DASHB_EINVAL, "null argument");
    if (!c->finalized || c->batch == 0) return fail(DASHB_ESTATE, "nothing evaluated");
    return copy_out(c, reinterpret_cast<void* const*>(out), false);
}

/* cuda_move_inputs + cuda_evaluate + cuda_move_outputs as one pipelined
 * pass: every lane copies its slice in, runs all layers and copies its
 * outputs back on its own stream. */
static int wait_async(dashb_circuit* c) {
    for (int set = 0; set < 2; ++set)
        if (c->async_busy[set]) {
            c->async_busy[set] = false;
            CU(cudaEventSynchronize(c->async_done[set]));
        }
    if (!c->atrace.empty()) {
        for (auto& t : c->atrace) {
            float t0 = 0;
            cudaEventElapsedTime(&t0, c->atrace_base, t.ev[0]);
            fprintf(stderr, "[dashb async] step %d set %d: call %.2f..%.2f ms (host), enqueue mark %.2f ms;", t.step, t.set, t.host_in_ms,
                    t.host_out_ms, t0);
            for (int i = 0; i < t.lanes; ++i) {
                float a = 0, b = 0, d = 0;
                cudaEventElapsedTime(&a, c->atrace_base, t.ev[1 + 3 * i]);
                cudaEventElapsedTime(&b, c->atrace_base, t.ev[2 + 3 * i]);
                cudaEventElapsedTime(&d, c->atrace_base, t.ev[3 + 3 * i]);
                float cp = 0;
                if (cudaEventElapsedTime(&cp, c->atrace_base, t.ev[1 + 3 * t.lanes + i]) != cudaSuccess) {
                    cudaGetLastError(); /* not recorded: the zero-copy path has no copies */
                    cp = 0;
                }
                fprintf(stderr, " lane %d copied %.2f in %.2f layers %.2f out %.2f;", i, cp, a, b, d);
            }
            fprintf(stderr, "\n");
            for (int i = 0; i < 1 + 4 * t.lanes; ++i) cudaEventDestroy(t.ev[i]);
        }
        c->atrace.clear();
    }
    return DASHB_OK;
}

static int evaluate_any(dashb_circuit* c, const void* const* in, uint64_t batch, void* const* out, bool compressed,
                        bool async = false) {
    if (!c || !in || !out) return fail(DASHB_EINVAL, "null argument");
    if (!c->finalized) return fail(DASHB_ESTATE, "circuit not finalized");
    if (async) {
        const auto t_call = std::chrono::steady_clock::now();
        if (c->shard_world > 1) return fail(DASHB_ESTATE, "channel-sharded circuits evaluate synchronously");
        for (int j = 0; j < c->k; ++j)
            if (!in[j] || !out[j]) return fail(DASHB_EINVAL, "label buffer %d missing", j);
        const int set = c->async_next;
        if (!c->async_done[set]) CU(cudaEventCreateWithFlags(&c->async_done[set], cudaEventDisableTiming));
        if (c->async_busy[set]) { /* this set's previous step: its buffers and the caller's are free once it is done */
            c->async_busy[set] = false;
            CU(cudaEventSynchronize(c->async_done[set]));
        }
        /* Two ways to keep two steps in flight (DASHB_ASYNC_MODE):
         * "inorder" (default): ONE set of lanes; a lane's stream runs step after step in order, only the input
         *   planes are double-buffered, so step n+1's labels are copied and re-laid-out while step n computes
         *   and lane i starts step n+1 the moment it has finished step n -- no second set of lanes competing
         *   for the SMs, the small late layers of step n run under the big first layers of step n+1.
         * "sets": a second set of lanes (streams + all buffers) computes step n+1 beside step n. */
        static const bool two_sets = getenv("DASHB_ASYNC_MODE") != nullptr && !strcmp(getenv("DASHB_ASYNC_MODE"), "sets");
        if (two_sets && set == 1 && c->lanes_alt.empty()) {
            int rc = make_lanes(c, c->lanes_alt);
            if (rc) return rc;
            CU(cudaStreamSynchronize(cudaStreamLegacy)); /* the clears of the new buffers */
        }
        if (two_sets && set == 1) std::swap(c->lanes, c->lanes_alt);
        if (!two_sets && !c->join_s) {
            CU(cudaStreamCreateWithFlags(&c->join_s, cudaStreamNonBlocking));
            const size_t in_bytes = tensor_bytes(c->input_size, c->k, c->crt, c->lane_cap);
            for (auto& l : c->lanes) {
                CU(l->buf_in2.alloc(in_bytes));
                CU(cudaMemsetAsync(l->buf_in2.p, 0, in_bytes, cudaStreamLegacy));
            }
            CU(cudaStreamSynchronize(cudaStreamLegacy));
        }
        /* nothing here is ordered behind dashb_stream(): the lanes' own streams
         * carry the order (the host waited for the step before the last above) */
        int rc = split_batch(c, batch);
        if (!rc && !two_sets) {
            for (int i = 0; i < c->lanes_used; ++i) {
                Lane& ln = *c->lanes[i];
                if (set == 1) {
                    carve(ln.t_in, (uint8_t*)ln.buf_in2.p, c->input_size, c->k, c->crt, ln.Bp);
                    ln.t_cur = ln.t_in;
                }
                /* the lane's staging area is read by the previous step's ingest kernels */
                CU(cudaStreamWaitEvent(c->copy_s, ln.ing_done, 0));
            }
        }
        const void* dev[MAXK];
        /* in-order mode: the copy engine by default (52-55 GB/s next to the kernels; the zero-copy ingest kernel
         * reads the host at 36-38 GB/s, which is below the compute rate of the benchmark model) */
        static const bool async_direct = getenv("DASHB_ASYNC_DIRECT") != nullptr && atoi(getenv("DASHB_ASYNC_DIRECT")) != 0;
        const bool direct = !rc && !compressed && (two_sets || async_direct) && mapped_host_inputs(c, in, dev);
        c->launches = 0;
        static const bool atrace = getenv("DASHB_TRACE_ASYNC") != nullptr;
        dashb_circuit::AsyncTrace* tr = nullptr;
        if (atrace && !rc && c->lanes_used <= 8 && c->atrace.size() < 64) {
            if (!c->atrace_base) {
                cudaEventCreate(&c->atrace_base);
                cudaEventRecord(c->atrace_base, c->ingest_s);
                c->atrace_t0 = t_call;
            }
            c->atrace.emplace_back();
            tr = &c->atrace.back();
            tr->step = c->atrace_steps++;
            tr->set = set;
            tr->lanes = c->lanes_used;
            tr->host_in_ms = std::chrono::duration<double, std::milli>(t_call - c->atrace_t0).count();
            for (int i = 0; i < 1 + 4 * tr->lanes; ++i) cudaEventCreate(&tr->ev[i]);
            cudaEventRecord(tr->ev[0], c->ingest_s);
        }
        for (int i = 0; !rc && i < c->lanes_used; ++i) {
            rc = enqueue_inputs(c, *c->lanes[i], in, compressed, c->copy_s, c->lanes[i]->s, direct ? dev : nullptr,
                                tr ? tr->ev[1 + 3 * tr->lanes + i] : nullptr);
            if (tr) cudaEventRecord(tr->ev[1 + 3 * i], c->lanes[i]->s);
        }
        /* experiment: no lane computes before the whole step's inputs are in */
        static const bool gather = getenv("DASHB_ASYNC_GATHER") != nullptr && atoi(getenv("DASHB_ASYNC_GATHER")) != 0;
        if (gather && !rc && c->lanes_used > 1)
            for (int i = 0; i + 1 < c->lanes_used; ++i) cudaStreamWaitEvent(c->lanes[i]->s, c->lanes[c->lanes_used - 1]->ing_done, 0);
        const int nl = (int)c->layers.size();
        for (int li = 0; !rc && li < nl; ++li)
            for (int i = 0; !rc && i < c->lanes_used; ++i) rc = run_layer(c, *c->lanes[i], li, c->lanes[i]->s);
        if (tr)
            for (int i = 0; i < c->lanes_used; ++i) cudaEventRecord(tr->ev[2 + 3 * i], c->lanes[i]->s);
        if (!rc && nl > 0) {
            c->out_n = c->layers[nl - 1]->out_size;
            for (int j = 0; j < c->k; ++j) c->out_q[j] = c->layers[nl - 1]->out_q[j];
        }
        for (int i = 0; !rc && i < c->lanes_used; ++i) {
            rc = enqueue_outputs(c, *c->lanes[i], out, compressed, c->lanes[i]->s);
            if (tr) cudaEventRecord(tr->ev[3 + 3 * i], c->lanes[i]->s);
        }
        if (tr) tr->host_out_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - c->atrace_t0).count();
        cudaError_t e = cudaSuccess;
        cudaStream_t js = two_sets ? c->lanes[0]->s : c->join_s;
        for (int i = two_sets ? 1 : 0; e == cudaSuccess && i < c->lanes_used; ++i) {
            e = cudaEventRecord(c->lanes[i]->done, c->lanes[i]->s);
            if (e == cudaSuccess) e = cudaStreamWaitEvent(js, c->lanes[i]->done, 0);
        }
        if (e == cudaSuccess) e = cudaEventRecord(c->async_done[set], js);
        if (two_sets && set == 1) std::swap(c->lanes, c->lanes_alt);
        if (e != cudaSuccess) return fail(DASHB_ECUDA, "dashb_evaluate_async: %s", cudaGetErrorString(e));
        c->async_busy[set] = true;
        c->async_next = set ^ 1;
        return rc;
    }
    {
        int rc = wait_async(c); /* a synchronous call behind asynchronous ones: drain them first */
        if (rc) return rc;
    }
    if (c->timing == 1) { /* serialised, per-layer events */
        int rc = move_inputs_any(c, in, batch, compressed);
        if (rc) return rc;
        if ((rc = dashb_evaluate_device(c))) return rc;
        return copy_out(c, out, compressed);
    }
    for (int j = 0; j < c->k; ++j)
        if (!in[j] || !out[j]) return fail(DASHB_EINVAL, "label buffer %d missing", j);
    int rc = split_batch(c, batch);
    if (rc) return rc;
    c->launches = 0;
    static const bool trace = getenv("DASHB_TRACE") != nullptr;
    cudaEvent_t tr[1 + 3 * 16];
    if (trace) {
        for (auto& e : tr) cudaEventCreate(&e);
        cudaEventRecord(tr[0], g_stream);
    }
    /* a failure part-way leaves the lane streams forked: join them again and
     * drop the trace events before reporting it (the message survives: the
     * clean-up calls do not touch the error text unless they fail themselves) */
    auto bail = [&](int code) {
        const std::string msg = g_err;
        join_lanes(c);
        cudaStreamSynchronize(g_stream);
        if (trace)
            for (auto& e : tr) cudaEventDestroy(e);
        g_err = msg;
        return code;
    };
    const void* dev[MAXK];
    const bool direct = !compressed && mapped_host_inputs(c, in, dev);
    if ((rc = fork_lanes(c))) return bail(rc);
    for (int i = 0; i < c->lanes_used; ++i) {
        /* all H2D copies go through one copy stream, lane after lane, so that
         * lane i-1 computes while lane i still transfers (pinned, mapped
         * buffers: no copies at all, the lane's ingest kernel reads the host) */
        if ((rc = enqueue_inputs(c, *c->lanes[i], in, compressed, c->copy_s, c->lanes[i]->s, direct ? dev : nullptr))) return bail(rc);
        if (trace) cudaEventRecord(tr[1 + 3 * i], c->lanes[i]->s);
    }
    const int nl = (int)c->layers.size();
    for (int li = 0; li < nl; ++li)
        for (int i = 0; i < c->lanes_used; ++i) {
            if ((rc = run_layer(c, *c->lanes[i], li, c->lanes[i]->s))) return bail(rc);
            if (trace && li == nl - 1) cudaEventRecord(tr[2 + 3 * i], c->lanes[i]->s);
        }
    if (nl > 0) {
        c->out_n = c->layers[nl - 1]->out_size;
        for (int j = 0; j < c->k; ++j) c->out_q[j] = c->layers[nl - 1]->out_q[j];
    }
    for (int i = 0; i < c->lanes_used; ++i) {
        if ((rc = enqueue_outputs(c, *c->lanes[i], out, compressed, c->lanes[i]->s))) return bail(rc);
        if (trace) cudaEventRecord(tr[3 + 3 * i], c->lanes[i]->s);
    }
    if ((rc = join_lanes(c))) return bail(rc);
    CU(cudaStreamSynchronize(g_stream));
    if (trace) {
        for (int i = 0; i < c->lanes_used; ++i) {
            float a = 0, b = 0, d = 0;
            cudaEventElapsedTime(&a, tr[0], tr[1 + 3 * i]);
            cudaEventElapsedTime(&b, tr[0], tr[2 + 3 * i]);
            cudaEventElapsedTime(&d, tr[0], tr[3 + 3 * i]);
            fprintf(stderr, "[dashb trace] lane %d: inputs done %.2f ms, layers done %.2f ms, outputs done %.2f ms\n", i, a, b, d);
        }
        for (auto& e : tr) cudaEventDestroy(e);
    }
    return DASHB_OK;
}

int dashb_evaluate(dashb_circuit* c, const int16_t* const* in, uint64_t batch, int16_t* const* out) {
    return evaluate_any(c, reinterpret_cast<const void* const*>(in), batch, reinterpret_cast<void* const*>(out), false);
}

int dashb_evaluate_compressed(dashb_circuit* c, const dashb_u128* const* in, uint64_t batch, dashb_u128* const* out) {
    return evaluate_any(c, reinterpret_cast<const void* const*>(in), batch, reinterpret_cast<void* const*>(out), true);
}

int dashb_evaluate_async(dashb_circuit* c, const int16_t* const* in, uint64_t batch, int16_t* const* out) {
    return evaluate_any(c, reinterpret_cast<const void* const*>(in), batch, reinterpret_cast<void* const*>(out), false, true);
}

int dashb_evaluate_compressed_async(dashb_circuit* c, const dashb_u128* const* in, uint64_t batch, dashb_u128* const* out) {
    return evaluate_any(c, reinterpret_cast<const void* const*>(in), batch, reinterpret_cast<void* const*>(out), true, true);
}

int dashb_evaluate_wait(dashb_circuit* c) {
    if (!c) return fail(DASHB_EINVAL, "null circuit");
    return wait_async(c);
}

int dashb_layer_outputs(dashb_circuit* c, int layer, int16_t* const* out) {
    if (!c || !out) return fail(DASHB_EINVAL, "null argument");
    if (!c->finalized || c->batch == 0) return fail(DASHB_ESTATE, "no inputs moved");
    if (layer < 0 || layer >= (int)c->layers.size()) return fail(DASHB_EINVAL, "layer index out of range");
    const int t = c->timing;
    c->timing = 0;
    int rc = run_layers(c, layer);
    c->timing = t;
    if (rc) return rc;
    return copy_out(c, reinterpret_cast<void* const*>(out), false);
}

int dashb_decode_outputs(dashb_circuit* c, int64_t* values, uint8_t* ok) {
    if (!c || !values) return fail(DASHB_EINVAL, "null argument");
    if (!c->finalized || c->batch == 0) return fail(DASHB_ESTATE, "nothing evaluated");
    if (!c->has_dec) return fail(DASHB_ESTATE, "no decoding information");
    for (int j = 0; j < c->k; ++j)
        if (c->out_q[j] != c->crt[j]) return fail(DASHB_EINVAL, "outputs are not in the CRT base");
    DecodeParams p;
    memset(&p, 0, sizeof(p));
    p.k = c->k;
    p.n = c->out_n;
    /* util::chinese_remainder accumulates `prod` in an int (misc/util.h:123) */
    int prod_i = 1;
    long long crt_modulus = 1;
    for (int j = 0; j < c->k; ++j) {
        prod_i = (int)((unsigned __int128)(__int128)prod_i * (unsigned __int128)c->crt[j]);
        crt_modulus *= c->crt[j];
    }
    if (prod_i <= 0) return fail(DASHB_EINVAL, "CRT modulus overflows the reference's int accumulator");
    p.prod = prod_i;
    p.crt_modulus = crt_modulus;
    for (int j = 0; j < c->k; ++j) {
        p.crt[j] = c->crt_mi[j];
        p.dec[j] = (const u128*)c->dec[j].p;
        p.pdiv[j] = prod_i / c->crt[j];
        p.inv[j] = (int64_t)(mul_inv((unsigned __int128)p.pdiv[j], (uint16_t)c->crt[j]) % (uint64_t)prod_i);
    }
    CU(cudaMemsetAsync(c->d_ok.p, 1, c->batch, g_stream));
    for (int i = 0; i < c->lanes_used; ++i) {
        Lane& ln = *c->lanes[i];
        for (int j = 0; j < c->k; ++j) p.in[j] = ln.t_cur.plane[j];
        p.batch = ln.batch;
        p.Bp = ln.Bp;
        p.values = (int64_t*)c->d_values.p + (size_t)ln.b0 * p.n;
        p.ok = (uint8_t*)c->d_ok.p + ln.b0;
        CU(launch_decode(p, g_stream));
    }
    CU(cudaMemcpyAsync(values, c->d_values.p, (size_t)c->batch * p.n * 8, cudaMemcpyDeviceToHost, g_stream));
    if (ok) CU(cudaMemcpyAsync(ok, c->d_ok.p, c->batch, cudaMemcpyDeviceToHost, g_stream));
    CU(cudaStreamSynchronize(g_stream));
    return DASHB_OK;
}

uint64_t dashb_input_size(const dashb_circuit* c) { return c ? c->input_size : 0; }
uint64_t dashb_output_size(const dashb_circuit* c) { return c ? c->cur_size : 0; }
int dashb_num_layers(const dashb_circuit* c) { return c ? (int)c->layers.size() : 0; }
int dashb_crt_base_size(const dashb_circuit* c) { return c ? c->k : 0; }
int dashb_crt_modulus(const dashb_circuit* c, int j) { return (c && j >= 0 && j < c->k) ? c->crt[j] : 0; }
int dashb_output_modulus(const dashb_circuit* c, int j) { return (c && j >= 0 && j < c->k) ? c->cur_q[j] : 0; }
uint64_t dashb_last_launch_count(const dashb_circuit* c) { return c ? c->launches : 0; }
uint64_t dashb_layer_size(const dashb_circuit* c, int layer) {
    return (c && layer >= 0 && layer < (int)c->layers.size()) ? c->layers[layer]->out_size : 0;
}
int dashb_layer_type(const dashb_circuit* c, int layer) {
    return (c && layer >= 0 && layer < (int)c->layers.size()) ? c->layers[layer]->type : 0;
}

static double layer_sum(const dashb_circuit* c, int layer, int what) {
    if (!c) return 0;
    double s = 0;
    for (int i = 0; i < (int)c->layers.size(); ++i) {
        if (layer >= 0 && i != layer) continue;
        const Layer& L = *c->layers[i];
        if (what == 0) s += L.label_bytes + L.weight_bytes / (c->batch ? c->batch : 1);
        if (what == 1) s += L.aes;
        if (what == 2) s += L.ref_aes;
    }
    return s;
}
double dashb_algorithmic_bytes(const dashb_circuit* c, int layer) { return layer_sum(c, layer, 0); }
double dashb_aes_blocks(const dashb_circuit* c, int layer) { return layer_sum(c, layer, 1); }
double dashb_reference_aes_blocks(const dashb_circuit* c, int layer) { return layer_sum(c, layer, 2); }

int dashb_set_layer_timing(dashb_circuit* c, int enabled) {
    if (!c) return fail(DASHB_EINVAL, "null circuit");
    if (enabled < 0 || enabled > 2) return fail(DASHB_EINVAL, "timing mode %d (0 off, 1 serialised, 2 concurrent intervals)", enabled);
    c->timing = enabled;
    return DASHB_OK;
}
int dashb_lanes_used(const dashb_circuit* c) { return c ? c->lanes_used : 0; }
int dashb_layer_interval_ms(dashb_circuit* c, int lane, int layer, float* begin_ms, float* end_ms) {
    if (!c || !begin_ms || !end_ms || layer < 0 || layer >= (int)c->layers.size() || lane < 0 || lane >= c->lanes_used)
        return fail(DASHB_EINVAL, "bad argument");
    const size_t at = ((size_t)lane * c->layers.size() + (size_t)layer) * 2;
    if (at + 1 >= c->iv_ms.size()) return fail(DASHB_ESTATE, "no evaluation in timing mode 2 yet");
    *begin_ms = c->iv_ms[at];
    *end_ms = c->iv_ms[at + 1];
    return DASHB_OK;
}
int dashb_layer_time_ms(dashb_circuit* c, int layer, float* ms) {
    if (!c || !ms || layer < 0 || layer >= (int)c->layers.size()) return fail(DASHB_EINVAL, "bad argument");
    *ms = c->layers[layer]->ms;
    return DASHB_OK;
}

int dashb_circuit_load_replicated(dashb_circuit** out, const char* path, int table_replicas) {
    if (table_replicas < 1 || table_replicas > 65536) return fail(DASHB_EINVAL, "replicas %d outside 1..65536", table_replicas);
    g_load_replicas = table_replicas;
    const int rc = dashb_circuit_load(out, path);
    g_load_replicas = 1;
    return rc;
}

int dashb_circuit_load(dashb_circuit** out, const char* path) {
    if (!out || !path) return fail(DASHB_EINVAL, "null argument");
    *out = nullptr;
    if (g_device < 0) return fail(DASHB_ENODEV, "dashb_init() has not succeeded; no device, no evaluation");
    gcm_circuit g;
    if (gcm_load(&g, path) != 0) return fail(DASHB_EIO, "cannot read DGCM file %s (missing, truncated or malformed)", path);
    if (!gcm_sizes_ok(&g)) {
        gcm_free(&g);
        return fail(DASHB_EINVAL, "DGCM file %s: an array's size does not match its layer's parameters", path);
    }
    int rc = dashb_circuit_from_gcm(&g, out);
    gcm_free(&g);
    return rc;
}

int dashb_material_check(const char* path) {
    if (!path) return fail(DASHB_EINVAL, "null argument");
    gcm_circuit g;
    if (gcm_load(&g, path) != 0) return fail(DASHB_EIO, "cannot read DGCM file %s (missing, truncated or malformed)", path);
    const int ok = gcm_sizes_ok(&g);
    gcm_free(&g);
    if (!ok) return fail(DASHB_EINVAL, "DGCM file %s: an array's size does not match its layer's parameters", path);
    return DASHB_OK;
}

void dashb_circuit_destroy(dashb_circuit* c) { delete c; }

/* ---- host self-checks of the device primitives (CPU-only test tier) ----
 * These run the __host__ __device__ label primitives of gc_common.cuh on the
 * host so that `pytest -m "not gpu"` can pin them against the oracle without
 * a GPU.  They are not an evaluation path: no circuit can be run with them. */
void dashb_hostcheck_compress(int q, const int16_t* digits, uint64_t out2[2]) {
    ModInfo m = make_modinfo((uint32_t)q);
    u128 x = compress_host(digits, m);
    out2[0] = x.lo;
    out2[1] = x.hi;
}
void dashb_hostcheck_digits(int q, const uint64_t in2[2], int16_t* digits) {
    ModInfo m = make_modinfo((uint32_t)q);
    gc::for_each_digit(gc::make_u128(in2[0], in2[1]), m, [&](uint32_t d, uint32_t v) { digits[d] = (int16_t)v; });
}
/* same digits through the multiplicative extraction (valid payloads only);
 * returns 0 if the modulus has no plan */
int dashb_hostcheck_digits_frac(int q, const uint64_t in2[2], int16_t* digits) {
    ModInfo m = make_modinfo((uint32_t)q);
    gc::FracInfo fi = make_fracinfo(m);
    if (fi.nst == 0) return 0;
    gc::for_each_chunk_msb(gc::make_u128(in2[0], in2[1]), m, fi, [&](uint32_t d0, uint32_t len, uint32_t r) {
        for (uint32_t i = 0; i < len; ++i) {
            digits[d0 + i] = (int16_t)(r % m.q);
            r /= m.q;
        }
    });
    return 1;
}
/* For every modulus in [q_lo, q_hi] that has a plan: `samples` random valid
 * payloads plus the edge payloads 0, 1, q^nc - 1, q^i and q^i - 1 through both
 * extractions; returns the number of moduli on which they differ (0 = all
 * agree) and, in *planned, how many moduli had a plan. */
int dashb_hostcheck_frac_sweep(int q_lo, int q_hi, int samples, uint64_t seed, int* planned) {
    int bad = 0, np = 0;
    uint64_t st = seed * 0x9e3779b97f4a7c15ull + 1;
    auto next = [&]() { /* splitmix64 */
        uint64_t z = (st += 0x9e3779b97f4a7c15ull);
        z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
        z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
        return z ^ (z >> 31);
    };
    for (int q = q_lo; q <= q_hi; ++q) {
        if (q < 2) continue;
        const ModInfo m = make_modinfo((uint32_t)q);
        const gc::FracInfo fi = make_fracinfo(m);
        if (fi.nst == 0) continue;
        ++np;
        /* D = q^nc as 128 bits (2^128 itself for the powers of two that reach it: then every payload is valid) */
        unsigned __int128 D = 1;
        bool full = false;
        for (uint32_t i = 0; i < m.nc; ++i) {
            const unsigned __int128 nd = D * (unsigned)q;
            if (nd / (unsigned)q != D) full = true;
            D = nd;
        }
        bool differs = false;
        auto check = [&](unsigned __int128 x) {
            if (!full && x >= D) return;
            uint32_t a[128] = {0}, b[128] = {0};
            const gc::u128 v = gc::make_u128((uint64_t)x, (uint64_t)(x >> 64));
            gc::for_each_digit(v, m, [&](uint32_t d, uint32_t dg) { a[d] = dg; });
            gc::for_each_chunk_msb(v, m, fi, [&](uint32_t d0, uint32_t len, uint32_t r) {
                for (uint32_t i = 0; i < len; ++i) {
                    b[d0 + i] = r % m.q;
                    r /= m.q;
                }
                if (r) differs = true; /* a chunk must hold exactly len digits */
            });
            for (uint32_t d = 0; d < m.nc; ++d)
                if (a[d] != b[d]) differs = true;
        };
        check(0);
        check(1);
        check(full ? ~(unsigned __int128)0 : D - 1);
        check(full ? ~(unsigned __int128)0 - 1 : D - 2);
        unsigned __int128 pw = 1;
        for (uint32_t i = 0; i < m.nc; ++i) {
            check(pw);
            check(pw - 1);
            pw *= (unsigned)q;
        }
        for (int t = 0; t < samples; ++t) {
            unsigned __int128 x = ((unsigned __int128)next() << 64) | next();
            if (!full) x %= D;
            check(x);
        }
        if (differs) ++bad;
    }
    if (planned) *planned = np;
    return bad;
}
/* The ladder walk (gc::for_each_chunk_ladder) over plain (aligned = 0) or
 * word-aligned (aligned = 1) plans against the division path, same payload
 * set as dashb_hostcheck_frac_sweep; also checks the plan's bookkeeping
 * (ladder counts, accumulator words, word alignment of the stages). */
int dashb_hostcheck_ladder_sweep(int q_lo, int q_hi, int samples, uint64_t seed, int aligned, int* planned) {
    int bad = 0, np = 0;
    uint64_t st = seed * 0x9e3779b97f4a7c15ull + 1;
    auto next = [&]() {
        uint64_t z = (st += 0x9e3779b97f4a7c15ull);
        z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
        z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
        return z ^ (z >> 31);
    };
    for (int q = q_lo; q <= q_hi; ++q) {
        if (q < 2) continue;
        const ModInfo m = make_modinfo((uint32_t)q);
        const gc::FracInfo fi = make_fracinfo(m, aligned != 0);
        if (fi.nst == 0) continue;
        ++np;
        bool differs = false;
        /* bookkeeping */
        uint32_t stages = 0, words = 0, digits = 0, lastW = 5;
        for (int W = 1; W <= 5; ++W) stages += (fi.ladder >> (4 * (W - 1))) & 15u;
        for (uint32_t s = 0; s < fi.nst; ++s) {
            const uint32_t len = fi.desc[s] & 0xffu, W = fi.desc[s] >> 8;
            if (W > lastW || W < 1) differs = true;
            lastW = W;
            words += (len + 1) / 2;
            digits += len;
            if (aligned && s > 0 && (len & 3u)) differs = true;
        }
        if (stages != fi.nst || words != fi.words || digits != m.nc) differs = true;
        if (aligned && ((fi.desc[0] & 0xffu) & 3u) != (m.nc & 3u)) differs = true;
        unsigned __int128 D = 1;
        bool full = false;
        for (uint32_t i = 0; i < m.nc; ++i) {
            const unsigned __int128 nd = D * (unsigned)q;
            if (nd / (unsigned)q != D) full = true;
            D = nd;
        }
        auto check = [&](unsigned __int128 x) {
            if (!full && x >= D) return;
            uint32_t a[128] = {0}, b[128] = {0};
            const gc::u128 v = gc::make_u128((uint64_t)x, (uint64_t)(x >> 64));
            gc::for_each_digit(v, m, [&](uint32_t d, uint32_t dg) { a[d] = dg; });
            uint32_t d0 = m.nc;
            gc::for_each_chunk_ladder(v, fi, [&](uint32_t len, uint32_t r) {
                d0 -= len;
                for (uint32_t i = 0; i < len; ++i) {
                    b[d0 + i] = r % m.q;
                    r /= m.q;
                }
                if (r) differs = true;
            });
            if (d0 != 0) differs = true;
            for (uint32_t d = 0; d < m.nc; ++d)
                if (a[d] != b[d]) differs = true;
        };
        check(0);
        check(1);
        check(full ? ~(unsigned __int128)0 : D - 1);
        check(full ? ~(unsigned __int128)0 - 1 : D - 2);
        unsigned __int128 pw = 1;
        for (uint32_t i = 0; i < m.nc; ++i) {
            check(pw);
            check(pw - 1);
            pw *= (unsigned)q;
        }
        for (int t = 0; t < samples; ++t) {
            unsigned __int128 x = ((unsigned __int128)next() << 64) | next();
            if (!full) x %= D;
            check(x);
        }
        if (differs) ++bad;
    }
    if (planned) *planned = np;
    return bad;
}
/* gc::WordCompressor (word-wise compress of a u8 digit plane label) on the
 * host: digits int16[nc] -> 128-bit value; returns 0 when the modulus has no
 * word-wise form (q = 256). */
int dashb_hostcheck_compress_words(int q, const int16_t* digits, uint64_t out2[2]) {
    const ModInfo m = make_modinfo((uint32_t)q);
    if (!m.cw) return 0;
    const uint32_t nw = (m.nc + 3) / 4;
    gc::WordCompressor c(m);
    for (int w = (int)nw - 1; w >= 0; --w) {
        uint32_t word = 0;
        for (uint32_t i = 0; i < 4 && 4 * w + i < m.nc; ++i) word |= (uint32_t)(uint8_t)digits[4 * w + i] << (8 * i);
        c.push(word);
    }
    out2[0] = c.value().lo;
    out2[1] = c.value().hi;
    return 1;
}
int dashb_hostcheck_colour(int q, const uint64_t in2[2]) {
    ModInfo m = make_modinfo((uint32_t)q);
    return (int)gc::colour_of(gc::make_u128(in2[0], in2[1]), m);
}
void dashb_hostcheck_hash(const uint64_t in2[2], uint64_t out2[2]) {
    const uint32_t *t0, *t1, *t2, *t3, *rk;
    const uint8_t* sb;
    host_aes_tables(&t0, &t1, &t2, &t3, &sb, &rk);
    u128 y = gc::aes128_encrypt(gc::make_u128(in2[0], in2[1]), t0, t1, t2, t3, sb, rk);
    out2[0] = y.lo;
    out2[1] = y.hi;
}

}  // extern "C"
