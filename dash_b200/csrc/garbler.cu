/*
 * garbler.cu -- host side of the GPU garbler (include/dash_b200_garble.h).
 *
 * Walks the plaintext circuit layer by layer the way GarbledCircuit::garble
 * does (garbling/garbled_circuit.h:122-176), carrying the wires' BASE labels
 * (the label of value 0):
 *   - linear layers run the evaluator's own linear kernels on the base
 *     labels as a batch of one (garbling a dense/conv layer IS evaluating it
 *     on the base labels with the zero label as bias,
 *     layer/garbled_dense.h:89-118, layer/garbled_conv2d.h:86-112);
 *   - gadget layers are turned into gate descriptors (kernels.h: GateDesc)
 *     for k_garble_gates, one descriptor per projection gate of the gadget,
 *     all elements of the layer in one launch.
 * The result is assembled as a gcm_circuit and written with gcm_save.
 */
#include <cuda_runtime.h>

#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <memory>
#include <numeric>
#include <vector>

#include "../../include/dash_b200.h"
#include "../../include/dash_b200_garble.h"
#include "../../include/dash_gcm_format.h"
#include "../../include/dash_b200_gcm.h"
#include "kernels.h"

using namespace dashb;
using gc::ModInfo;
using gc::u128;

namespace {

#define GCU(x)                                                                                   \
    do {                                                                                         \
        cudaError_t e_ = (x);                                                                    \
        if (e_ != cudaSuccess) return report_error(DASHB_ECUDA, "%s: %s", #x, cudaGetErrorString(e_)); \
    } while (0)
#define GRC(x)                   \
    do {                         \
        int rc_ = (x);           \
        if (rc_ != DASHB_OK) return rc_; \
    } while (0)

typedef std::vector<int16_t> Label; /* nc digits */

struct DevBuf {
    void* p{nullptr};
    size_t n{0};
    ~DevBuf() { cudaFree(p); }
    cudaError_t alloc(size_t bytes) {
        cudaFree(p);
        p = nullptr;
        n = bytes;
        return bytes ? cudaMalloc(&p, bytes) : cudaSuccess;
    }
};

int64_t pmod(int64_t a, int64_t q) {
    int64_t r = a % q;
    return r < 0 ? r + q : r;
}

/* util::mul_inv (misc/util.h:92-103) */
int64_t mul_inv_ll(long long a, long long b) {
    long long b0 = b, t, q;
    long long x0 = 0, x1 = 1;
    if (b == 1) return 1;
    while (a > 1) {
        if (b == 0) return 0;
        q = a / b;
        t = b, b = a % b, a = t;
        t = x0, x0 = x1 - q * x0, x1 = t;
    }
    if (x1 < 0) x1 += b0;
    return x1;
}

uint32_t nc4x4(uint32_t nc) { return (nc + 3) & ~3u; }

/* AES-128 key expansion with the S-box of the hash tables */
void expand_key(const uint8_t key[16], const uint8_t* sbox, uint32_t rk[44]) {
    static const uint8_t rcon[10] = {0x01, 0x02, 0x04, 0x08, 0x10, 0x20, 0x40, 0x80, 0x1b, 0x36};
    for (int i = 0; i < 4; ++i)
        rk[i] = ((uint32_t)key[4 * i] << 24) | ((uint32_t)key[4 * i + 1] << 16) | ((uint32_t)key[4 * i + 2] << 8) | key[4 * i + 3];
    for (int i = 4; i < 44; ++i) {
        uint32_t t = rk[i - 1];
        if (i % 4 == 0) {
            t = (t << 8) | (t >> 24);
            t = ((uint32_t)sbox[t >> 24] << 24) | ((uint32_t)sbox[(t >> 16) & 0xff] << 16) | ((uint32_t)sbox[(t >> 8) & 0xff] << 8) |
                sbox[t & 0xff];
            t ^= (uint32_t)rcon[i / 4 - 1] << 24;
        }
        rk[i] = rk[i - 4] ^ t;
    }
}

/* LookupApproxSign::gen_lookup (gadgets/lookup_approx_sign.h:73-98): for CRT
 * residue j of modulus i, the mixed-radix digits (most significant first) of
 * round(alpha_i * j / P, to 1/prod(mrs)) * prod(mrs). */
std::vector<uint8_t> approx_lookup(const std::vector<int>& crt, const std::vector<int>& mrs, std::vector<uint32_t>& lut_off) {
    const int t = (int)mrs.size();
    long level = 1;
    for (int m : mrs) level *= m;
    long pk = 1;
    for (int p : crt) pk *= p;
    std::vector<uint8_t> lut;
    lut_off.clear();
    for (size_t i = 0; i < crt.size(); ++i) {
        lut_off.push_back((uint32_t)lut.size());
        long double A = pk / crt[i];
        long double alpha = A * mul_inv_ll((long long)A, (uint16_t)crt[i]);
        for (int j = 0; j < crt[i]; ++j) {
            long double d = alpha * j / pk;
            double res = (double)d * level; /* round_nearest_frac(double, long) */
            res = std::round(res);
            d = res / level;
            d = d * level;
            uint32_t dec = static_cast<uint32_t>(std::round(d));
            std::vector<uint8_t> mr(t);
            for (int x = t - 1; x >= 0; --x) {
                mr[x] = (uint8_t)(dec % (uint32_t)mrs[x]);
                dec = dec / (uint32_t)mrs[x];
            }
            lut.insert(lut.end(), mr.begin(), mr.end());
        }
    }
    return lut;
}

struct Garbler {
    int k{0}, n_mrs{0}, max_modulus{0};
    std::vector<int> crt, mrs;
    int64_t P{1}; /* CRT modulus */
    std::vector<ModInfo> mods; /* by modulus */
    /* host AES for constant labels */
    const uint32_t *te0, *te1, *te2, *te3;
    const uint8_t* sbox;
    uint32_t rk_prf[44];
    uint64_t host_ctr{0};
    /* constant labels as u8 digits (device copy: d_labels) */
    std::vector<uint8_t> labels;
    std::vector<uint32_t> r_off, z_off; /* by modulus */
    std::vector<Label> R, Z;            /* by modulus */
    std::vector<Label> U;               /* upshift base labels, per CRT modulus */
    std::vector<uint32_t> u_off;
    DevBuf d_labels, d_r_off, d_mods, d_lut;
    std::vector<uint32_t> lut_off;
    /* current base labels, per CRT modulus int16 [n][nc] */
    uint64_t cur_n{0};
    std::vector<std::vector<int16_t>> cur;
    /* output */
    gcm_circuit g;
    std::vector<std::unique_ptr<std::vector<uint8_t>>> store; /* owns every host array of g */
    std::vector<std::unique_ptr<DevBuf>> dev_store;           /* ciphertext tables stay on the device */
    std::vector<gcm_layer> glayers;
    dashb_garble_stats st;
    cudaEvent_t ev0{nullptr}, ev1{nullptr};
    uint64_t layer_tag{0};

    ~Garbler() {
        if (ev0) cudaEventDestroy(ev0);
        if (ev1) cudaEventDestroy(ev1);
    }

    uint32_t nc(int q) const { return mods[q].nc; }

    gcm_array keep(const void* data, size_t nbytes) {
        store.emplace_back(new std::vector<uint8_t>((const uint8_t*)data, (const uint8_t*)data + nbytes));
        gcm_array a;
        a.data = store.back()->data();
        a.nbytes = nbytes;
        return a;
    }

    gcm_array keep_dev(std::unique_ptr<DevBuf> d) {
        gcm_array a;
        a.data = d->p;
        a.nbytes = d->n;
        dev_store.push_back(std::move(d));
        return a;
    }
    bool on_device(const void* ptr) const {
        for (const auto& d : dev_store)
            if (d->p == ptr) return true;
        return false;
    }

    /* a fresh label of modulus q from the host-side PRF stream */
    Label random_label(int q) {
        const u128 x = gc::aes128_encrypt(gc::make_u128(host_ctr++, 0xffffffffull << 32), te0, te1, te2, te3, sbox, rk_prf);
        Label l(nc(q));
        gc::for_each_digit(x, mods[q], [&](uint32_t d, uint32_t v) { l[d] = (int16_t)v; });
        return l;
    }
    Label lin(const Label& a, int64_t ca, const Label& b, int64_t cb, int q) const { /* ca*a + cb*b mod q */
        Label r(a.size());
        for (size_t d = 0; d < a.size(); ++d) r[d] = (int16_t)pmod(ca * a[d] + cb * b[d], q);
        return r;
    }
    uint32_t add_const_label(const Label& l) { /* returns its byte offset in `labels` */
        const uint32_t off = (uint32_t)labels.size();
        labels.resize(off + nc4x4((uint32_t)l.size()), 0);
        for (size_t d = 0; d < l.size(); ++d) labels[off + d] = (uint8_t)l[d];
        return off;
    }
    int upload_labels() {
        GCU(d_labels.alloc(labels.size()));
        GCU(cudaMemcpy(d_labels.p, labels.data(), labels.size(), cudaMemcpyHostToDevice));
        return DASHB_OK;
    }
};

GateTerm term_plane(uint32_t j, int64_t coef) { return GateTerm{GB_TERM_PLANE, (uint32_t)coef, j}; }
GateTerm term_const(uint32_t off, int64_t coef) { return GateTerm{GB_TERM_CONST, (uint32_t)coef, off}; }
GateTerm term_prf(uint64_t tag, int64_t coef) { return GateTerm{GB_TERM_PRF, (uint32_t)coef, tag}; }

GateDesc blank_desc() {
    GateDesc d;
    memset(&d, 0, sizeof(d));
    d.off_plane = GB_NONE;
    d.r_plane = GB_NONE;
    d.out_plane = -1;
    return d;
}

/* PRF tags inside one layer */
enum { TAG_A = 1, TAG_C1, TAG_C2, TAG_S, TAG_G, TAG_E, TAG_T, TAG_B };
uint64_t tag(int kind, int a, int b) { return ((uint64_t)kind << 24) | ((uint64_t)a << 12) | (uint64_t)b; }

}  // namespace

namespace {

/* u8 [n][4*nc4] device copy of the current base labels of CRT modulus j */
int upload_plane(const std::vector<int16_t>& lab, uint64_t n, uint32_t nc, DevBuf& d) {
    const uint32_t w = nc4x4(nc);
    std::vector<uint8_t> h((size_t)n * w, 0);
    for (uint64_t e = 0; e < n; ++e)
        for (uint32_t x = 0; x < nc; ++x) h[e * w + x] = (uint8_t)lab[e * nc + x];
    GCU(d.alloc(h.size()));
    GCU(cudaMemcpy(d.p, h.data(), h.size(), cudaMemcpyHostToDevice));
    return DASHB_OK;
}

int download_plane(const DevBuf& d, uint64_t n, uint32_t nc, std::vector<int16_t>& lab) {
    const uint32_t w = nc4x4(nc);
    std::vector<uint8_t> h((size_t)n * w);
    GCU(cudaMemcpy(h.data(), d.p, h.size(), cudaMemcpyDeviceToHost));
    lab.assign((size_t)n * nc, 0);
    for (uint64_t e = 0; e < n; ++e)
        for (uint32_t x = 0; x < nc; ++x) lab[e * nc + x] = h[e * w + x];
    return DASHB_OK;
}

/*
 * Runs one batch of gate descriptors over the n elements of the current
 * tensor.  table_rows[t] = rows of output table t (zero-initialised on the
 * device); out_q[j] != 0 asks for an output plane of that modulus.  Tables
 * come back in `tables`, output planes replace G.cur when replace_cur.
 */
int run_gates(Garbler& G, std::vector<GateDesc>& descs, const std::vector<uint64_t>& table_rows,
              std::vector<gcm_array>& tables, const std::vector<int>& out_q, bool replace_cur) {
    const uint64_t n = G.cur_n;
    GarbleParams P;
    memset(&P, 0, sizeof(P));
    std::vector<DevBuf> d_in(G.k), d_out(out_q.size());
    std::vector<std::unique_ptr<DevBuf>> d_tab;
    for (size_t t = 0; t < table_rows.size(); ++t) d_tab.emplace_back(new DevBuf);
    uint32_t nc_max = 1;
    for (int j = 0; j < G.k; ++j) {
        GRC(upload_plane(G.cur[j], n, G.nc(G.crt[j]), d_in[j]));
        P.in[j] = (const uint8_t*)d_in[j].p;
        P.in_nc[j] = G.nc(G.crt[j]);
    }
    for (size_t j = 0; j < out_q.size(); ++j) {
        if (!out_q[j]) continue;
        GCU(d_out[j].alloc((size_t)n * nc4x4(G.nc(out_q[j]))));
        P.out[j] = (uint8_t*)d_out[j].p;
    }
    if (table_rows.size() > sizeof(P.tables) / sizeof(P.tables[0])) return report_error(DASHB_EINVAL, "garbler: too many tables in one launch");
    for (size_t t = 0; t < table_rows.size(); ++t) {
        GCU(d_tab[t]->alloc(table_rows[t] * 16));
        if (table_rows[t]) GCU(cudaMemset(d_tab[t]->p, 0, table_rows[t] * 16));
        P.tables[t] = (u128*)d_tab[t]->p;
    }
    for (const GateDesc& d : descs) {
        if (d.q_in) nc_max = std::max(nc_max, G.nc((int)d.q_in));
        if (d.q_out) nc_max = std::max(nc_max, G.nc((int)d.q_out));
        if (d.q_in) {
            G.st.gates += n;
            G.st.rows += n * d.q_in;
        }
    }
    DevBuf d_desc;
    GCU(d_desc.alloc(descs.size() * sizeof(GateDesc)));
    GCU(cudaMemcpy(d_desc.p, descs.data(), descs.size() * sizeof(GateDesc), cudaMemcpyHostToDevice));
    P.desc = (const GateDesc*)d_desc.p;
    P.n_desc = (uint32_t)descs.size();
    P.nc_max = nc_max;
    P.n = n;
    P.mods = (const ModInfo*)G.d_mods.p;
    P.labels = (const uint8_t*)G.d_labels.p;
    P.r_off = (const uint32_t*)G.d_r_off.p;
    P.layer_tag = G.layer_tag;
    cudaStream_t s = library_stream();
    /* every upload and table clear above ran on the NULL stream (and a
     * pageable cudaMemcpy may return before its DMA lands); the library
     * stream is non-blocking, so nothing orders the kernel behind them */
    GCU(cudaDeviceSynchronize());
    GCU(cudaEventRecord(G.ev0, s));
    GCU(launch_garble_gates(P, s));
    GCU(cudaEventRecord(G.ev1, s));
    GCU(cudaStreamSynchronize(s));
    float ms = 0;
    GCU(cudaEventElapsedTime(&ms, G.ev0, G.ev1));
    G.st.kernel_ms += ms;
    tables.clear();
    for (size_t t = 0; t < table_rows.size(); ++t) {
        G.st.table_bytes += table_rows[t] * 16;
        tables.push_back(G.keep_dev(std::move(d_tab[t])));
    }
    if (replace_cur) {
        std::vector<std::vector<int16_t>> next(out_q.size());
        for (size_t j = 0; j < out_q.size(); ++j)
            if (out_q[j]) GRC(download_plane(d_out[j], n, G.nc(out_q[j]), next[j]));
        G.cur.swap(next);
    }
    return DASHB_OK;
}

/*
 * Descriptors of the approximate sign gadget (SignGadget::garble,
 * gadgets/sign_gadget.h:420-581) on the current tensor.  Tables 0..3 =
 * approx, cast1, cast2, sign in the reference's layout (include/dash_b200.h).
 * Output o gets base label PRF(TAG_S, o) of modulus out_mod[o]; with
 * to_plane it is also the layer's output label of plane o.
 */
void sign_gadget_descs(const Garbler& G, std::vector<GateDesc>& D, const std::vector<int>& out_mod, int lower, int upper,
                       bool to_plane, uint64_t rows[4]) {
    const int k = G.k, nm = G.n_mrs;
    uint32_t S = 0, Mp = 0;
    std::vector<uint32_t> prefix(k);
    for (int j = 0; j < k; ++j) {
        prefix[j] = S;
        S += (uint32_t)G.crt[j];
    }
    for (int i = 1; i < nm; ++i) Mp += (uint32_t)G.mrs[i];
    const uint64_t n = G.cur_n;
    rows[0] = n * nm * S;
    rows[1] = n * (uint64_t)(k + 1) * Mp;
    rows[2] = rows[1];
    rows[3] = n * out_mod.size() * (uint64_t)G.mrs[0];
    /* step 1: approximate residues (approx_res_gadget.h:59-83) */
    for (int j = 0; j < k; ++j)
        for (int i = 0; i < nm; ++i) {
            GateDesc d = blank_desc();
            d.q_in = (uint32_t)G.crt[j];
            d.in[d.n_in++] = term_plane((uint32_t)j, 1);
            d.q_out = (uint32_t)G.mrs[i];
            d.out[d.n_out++] = term_prf(tag(TAG_A, j, i), 1);
            d.f = GB_F_LOOKUP;
            d.flut = (uint64_t)((const uint8_t*)G.d_lut.p + G.lut_off[j] + i);
            d.flut_stride = (uint32_t)nm;
            d.table = 0;
            d.table_off = (uint64_t)nm * prefix[j] + (uint64_t)i * G.crt[j];
            d.table_stride = (uint64_t)nm * S;
            D.push_back(d);
        }
    /* step 2: mixed-radix addition, least significant digit first (:441-544) */
    uint64_t cnt1 = 0, cnt2 = 0;
    for (int i = nm - 1; i >= 1; --i) {
        const uint32_t mi = (uint32_t)G.mrs[i], C = (uint32_t)(k + 1) * mi;
        for (int j = 0; j <= k; ++j) { /* 2.2: cast the k digits and the carry to modulus (k+1)*m_i */
            GateDesc d = blank_desc();
            d.q_in = mi;
            if (j < k)
                d.in[d.n_in++] = term_prf(tag(TAG_A, j, i), 1);
            else if (i == nm - 1)
                d.in[d.n_in++] = term_const(G.z_off[mi], 1); /* initial carry: the zero label (:438) */
            else
                d.in[d.n_in++] = term_prf(tag(TAG_C2, i + 1, 0), 1);
            d.q_out = C;
            d.out[d.n_out++] = term_prf(tag(TAG_C1, i, j), 1);
            d.f = GB_F_ID;
            d.table = 1;
            d.table_off = cnt1 + (uint64_t)j * mi;
            d.table_stride = (uint64_t)(k + 1) * Mp;
            D.push_back(d);
        }
        cnt1 += (uint64_t)(k + 1) * mi;
        { /* 2.3 + 2.4: sum of the casts -> carry into digit i-1, f(x) = x / m_i */
            GateDesc d = blank_desc();
            d.q_in = C;
            for (int j = 0; j <= k; ++j) d.in[d.n_in++] = term_prf(tag(TAG_C1, i, j), 1);
            d.q_out = (uint32_t)G.mrs[i - 1];
            d.out[d.n_out++] = term_prf(tag(TAG_C2, i, 0), 1);
            d.f = GB_F_DIV;
            d.fa = (int32_t)mi;
            d.table = 2;
            d.table_off = cnt2;
            d.table_stride = (uint64_t)(k + 1) * Mp;
            D.push_back(d);
        }
        cnt2 += C;
    }
    /* step 3: sign of the most significant sum (:547-571) */
    for (size_t o = 0; o < out_mod.size(); ++o) {
        GateDesc d = blank_desc();
        d.q_in = (uint32_t)G.mrs[0];
        if (nm == 1)
            d.in[d.n_in++] = term_const(G.z_off[G.mrs[0]], 1);
        else
            d.in[d.n_in++] = term_prf(tag(TAG_C2, 1, 0), 1);
        for (int j = 0; j < k; ++j) d.in[d.n_in++] = term_prf(tag(TAG_A, j, 0), 1);
        d.q_out = (uint32_t)out_mod[o];
        d.out[d.n_out++] = term_prf(tag(TAG_S, (int)o, 0), 1);
        d.f = GB_F_SIGN;
        d.fa = G.mrs[0] / 2;
        d.fb = upper;
        d.fc = lower;
        d.table = 3;
        d.table_off = (uint64_t)o * G.mrs[0];
        d.table_stride = (uint64_t)out_mod.size() * G.mrs[0];
        if (to_plane) d.out_plane = (int32_t)o;
        D.push_back(d);
    }
}

/* GarbledRelu::garble (layer/garbled_relu.h:118-176) / GarbledSign::garble
 * (layer/garbled_sign.h:16-40) */
int garble_sign_layer(Garbler& G, bool relu) {
    if (G.n_mrs < 1) return report_error(DASHB_EINVAL, "garbler: ReLU/sign need an MRS base");
    if (G.P % 2 != 0 || G.mrs[0] % 2 != 0) return report_error(DASHB_EINVAL, "garbler: the sign gadget needs an even CRT modulus and an even first MRS modulus");
    const int k = G.k;
    std::vector<GateDesc> D;
    uint64_t rows[4];
    std::vector<uint64_t> table_rows;
    std::vector<int> out_q(k);
    for (int j = 0; j < k; ++j) out_q[j] = G.crt[j];
    if (relu) {
        /* sign bit in modulus 2, bounds (0, 1) (garbled_relu.h:65), then x * sign per modulus */
        sign_gadget_descs(G, D, std::vector<int>{2}, 0, 1, false, rows);
        uint32_t S = 0;
        for (int j = 0; j < k; ++j) {
            const uint32_t p = (uint32_t)G.crt[j];
            /* MixedModHalfGate::garble (gates/mixed_mod_half_gate.h:51-89): r = colour of x's base label */
            GateDesc g = blank_desc(); /* garbler half gate: f(x) = x * r */
            g.q_in = p;
            g.in[g.n_in++] = term_plane((uint32_t)j, 1);
            g.q_out = p;
            g.out[g.n_out++] = term_prf(tag(TAG_G, j, 0), 1);
            g.f = GB_F_MULR;
            g.r_plane = (uint32_t)j;
            g.table = 4;
            g.table_off = S;
            D.push_back(g);
            GateDesc e = blank_desc(); /* evaluator half gate on the sign bit: payload = -(y + r) * x + sk04 */
            e.q_in = 2;
            e.in[e.n_in++] = term_prf(tag(TAG_S, 0, 0), 1);
            e.q_out = p;
            e.out[e.n_out++] = term_prf(tag(TAG_E, j, 0), 1);
            e.off_plane = (uint32_t)j;
            e.f = GB_F_NEGSUM;
            e.r_plane = (uint32_t)j;
            e.table = 5;
            e.table_off = (uint64_t)j * 3;
            e.table_stride = (uint64_t)k * 3;
            D.push_back(e);
            GateDesc m = blank_desc(); /* mini gate: (y + r) mod p in 16-bit slots of row 2 */
            m.q_in = 2;
            m.in[m.n_in++] = term_prf(tag(TAG_S, 0, 0), 1);
            m.mode = GB_MODE_MINI;
            m.f = GB_F_ADDR;
            m.fa = (int32_t)p;
            m.r_plane = (uint32_t)j;
            m.table = 5;
            m.table_off = (uint64_t)j * 3 + 2;
            m.table_stride = (uint64_t)k * 3;
            D.push_back(m);
            GateDesc o = blank_desc(); /* output base label sk04 - sk03 */
            o.q_out = p;
            o.out[o.n_out++] = term_prf(tag(TAG_E, j, 0), 1);
            o.out[o.n_out++] = term_prf(tag(TAG_G, j, 0), p - 1);
            o.out_plane = j;
            D.push_back(o);
            S += p;
        }
        for (GateDesc& d : D)
            if (d.table == 4 && d.q_in) d.table_stride = S;
        table_rows = {rows[0], rows[1], rows[2], rows[3], G.cur_n * S, G.cur_n * (uint64_t)k * 3};
    } else {
        /* sign layer: outputs on every CRT modulus, bounds (-1, 1) (garbled_sign.h:22-23) */
        sign_gadget_descs(G, D, G.crt, -1, 1, true, rows);
        table_rows = {rows[0], rows[1], rows[2], rows[3]};
    }
    std::vector<gcm_array> tables;
    const uint64_t n = G.cur_n;
    GRC(run_gates(G, D, table_rows, tables, out_q, true));
    gcm_layer L;
    memset(&L, 0, sizeof(L));
    L.p[0] = relu ? GCM_RELU : GCM_SIGN;
    L.p[1] = (int64_t)n;
    L.n_arrays = (int)tables.size();
    for (size_t t = 0; t < tables.size(); ++t) L.a[t] = tables[t];
    G.glayers.push_back(L);
    return DASHB_OK;
}

}  // namespace

namespace {

/* RescaleGadget::garble with the base-extension variant
 * (gadgets/rescale_gadget.h:114-243, base_extension_gadget.h:63-196). */
int garble_rescale(Garbler& G, const int64_t* factors_in, int nf) {
    const int k = G.k;
    if (nf < 1 || nf > MAXF || nf >= k) return report_error(DASHB_EINVAL, "garbler: bad number of rescale factors");
    const uint64_t n = G.cur_n;
    std::vector<int> factors(nf), fidx(nf);
    std::vector<bool> ignored(k, false);
    /* step 2 bookkeeping: which (factor, modulus) pairs have a trans-modulus gate, in cipher order */
    std::vector<std::vector<int64_t>> tm_off(nf, std::vector<int64_t>(k, -1));
    std::vector<std::vector<int64_t>> inv_f(nf, std::vector<int64_t>(k, 0));
    uint64_t tm_rows = 0;
    int64_t prod_f = 1;
    for (int i = 0; i < nf; ++i) {
        factors[i] = (int)factors_in[i];
        int f = -1;
        for (int j = 0; j < k; ++j)
            if (G.crt[j] == factors[i] && f < 0) f = j;
        if (f < 0 || ignored[f]) return report_error(DASHB_EINVAL, "garbler: rescale factor %d is not a (distinct) CRT modulus", factors[i]);
        ignored[f] = true;
        fidx[i] = f;
        prod_f *= factors[i];
        for (int j = 0; j < k; ++j) {
            if (ignored[j]) continue;
            tm_off[i][j] = (int64_t)tm_rows;
            tm_rows += n * (uint64_t)factors[i];
            inv_f[i][j] = mul_inv_ll(factors[i], G.crt[j]);
            if (inv_f[i][j] * factors[i] % G.crt[j] != 1) return report_error(DASHB_EINVAL, "garbler: factor %d not invertible mod %d", factors[i], G.crt[j]);
        }
    }
    /* swapped order of the base extension (base_extension_gadget.h:334-345) */
    std::vector<int> perm(k);
    for (int j = 0; j < k; ++j) perm[j] = j;
    for (int i = 1; i <= nf; ++i) std::swap(perm[fidx[nf - i]], perm[k - i]);
    const int non_ext = k - nf;
    std::vector<int> swapped(k);
    for (int j = 0; j < k; ++j) swapped[j] = G.crt[perm[j]];
    std::vector<std::vector<int64_t>> be_off(k, std::vector<int64_t>(k, -1)), inv_p(k, std::vector<int64_t>(k, 0));
    uint64_t be_rows = 0;
    for (int i = 0; i < non_ext; ++i)
        for (int t = i + 1; t < k; ++t) {
            be_off[i][t] = (int64_t)be_rows;
            be_rows += n * (uint64_t)swapped[i];
            inv_p[i][t] = mul_inv_ll(swapped[i], swapped[t]);
        }
    /* fresh downshift base labels of this layer, and the labels the evaluator subtracts
     * (garbled_circuit_interface.h:969-979): (P / (2 * prod factors)) * R + base */
    std::vector<uint32_t> d_off(k);
    std::vector<Label> down(k);
    for (int j = 0; j < k; ++j) {
        const int q = G.crt[j];
        const Label base = G.random_label(q);
        d_off[j] = G.add_const_label(base);
        down[j] = G.lin(G.R[q], pmod(G.P / (2 * prod_f), q), base, 1, q);
    }
    GRC(G.upload_labels());

    /* a label as a list of terms; scaling multiplies every coefficient */
    typedef std::vector<GateTerm> Terms;
    auto scaled = [](const Terms& t, int64_t c, int q) {
        Terms r = t;
        for (GateTerm& x : r) x.coef = (uint32_t)pmod((int64_t)x.coef * c, q);
        return r;
    };
    /* w_j after step 1 (upshift) and the divisions by factors 0..upto-1 (steps 2.1-2.2) */
    auto after_factors = [&](int j, int upto) {
        const int q = G.crt[j];
        Terms w = {term_plane((uint32_t)j, 1), term_const(G.u_off[j], 1)};
        for (int i = 0; i < upto; ++i) {
            if (tm_off[i][j] < 0) continue;
            w.push_back(term_prf(tag(TAG_T, i, j), q - 1)); /* minus the projected residue */
            w = scaled(w, inv_f[i][j], q);                  /* times factor^-1 */
        }
        return w;
    };
    std::vector<GateDesc> D;
    auto put_terms = [&](GateTerm* dst, uint32_t& cnt, const Terms& t) -> bool {
        if (t.size() > (size_t)GB_MAXTERMS) return false;
        for (const GateTerm& x : t) dst[cnt++] = x;
        return true;
    };
    /* step 2: trans-modulus projections of the factor's residue into every other modulus */
    for (int i = 0; i < nf; ++i)
        for (int j = 0; j < k; ++j) {
            if (tm_off[i][j] < 0) continue;
            GateDesc d = blank_desc();
            d.q_in = (uint32_t)factors[i];
            if (!put_terms(d.in, d.n_in, after_factors(fidx[i], i))) return report_error(DASHB_EINVAL, "garbler: too many label terms");
            d.q_out = (uint32_t)G.crt[j];
            d.out[d.n_out++] = term_prf(tag(TAG_T, i, j), 1);
            d.f = GB_F_ID;
            d.table = 0;
            d.table_off = (uint64_t)tm_off[i][j];
            d.table_stride = (uint64_t)factors[i];
            D.push_back(d);
        }
    /* step 3: base extension.  lw[t] = the label the elimination starts from, in swapped order */
    std::vector<Terms> lw(k);
    for (int t = 0; t < k; ++t) {
        const int j = perm[t];
        if (t < non_ext)
            lw[t] = after_factors(j, nf);
        else
            lw[t] = {term_const(G.z_off[G.crt[j]], 1)}; /* step 2.3: zero label on the factors' residues */
    }
    for (int i = 0; i < non_ext; ++i) {
        for (int t = i + 1; t < k; ++t) {
            GateDesc d = blank_desc();
            d.q_in = (uint32_t)swapped[i];
            if (!put_terms(d.in, d.n_in, lw[i])) return report_error(DASHB_EINVAL, "garbler: too many label terms (CRT base too long for the base extension)");
            d.q_out = (uint32_t)swapped[t];
            d.out[d.n_out++] = term_prf(tag(TAG_B, i, t), 1);
            d.f = GB_F_ID;
            d.table = 1;
            d.table_off = (uint64_t)be_off[i][t];
            d.table_stride = (uint64_t)swapped[i];
            D.push_back(d);
            /* l_w[t] = (l_w[t] - projected l_w[i]) * inv (base_extension_gadget.h:143-151) */
            lw[t].push_back(term_prf(tag(TAG_B, i, t), swapped[t] - 1));
            lw[t] = scaled(lw[t], inv_p[i][t], swapped[t]);
        }
    }
    /* step 4 + outputs: untouched residues leave after step 2, the factors'
     * residues are the extended digit times -invv (base_extension_gadget.h:
     * 186-195); both minus the downshift base */
    std::vector<int> out_q(k);
    for (int j = 0; j < k; ++j) out_q[j] = G.crt[j];
    for (int t = 0; t < k; ++t) {
        const int j = perm[t], q = G.crt[j];
        Terms o;
        if (t < non_ext) {
            o = after_factors(j, nf);
        } else {
            int x = -1;
            for (int i = 0; i < nf; ++i)
                if (fidx[i] == j) x = i;
            int acc_i = 1; /* std::accumulate with an int accumulator (base_extension_gadget.h:351) */
            for (int i = 0; i < non_ext; ++i) acc_i = (int)((unsigned long)(long)acc_i * (unsigned long)swapped[i]);
            const int16_t inv_total = (int16_t)mul_inv_ll((long long)(unsigned long)(long)acc_i, (uint16_t)factors[x]);
            const int16_t invv = (int16_t)mul_inv_ll((long long)inv_total, (uint16_t)factors[x]);
            o = scaled(lw[t], pmod(-(int64_t)invv, q), q);
        }
        o.push_back(term_const(d_off[j], q - 1));
        GateDesc d = blank_desc();
        d.q_out = (uint32_t)q;
        if (!put_terms(d.out, d.n_out, o)) return report_error(DASHB_EINVAL, "garbler: too many label terms");
        d.out_plane = j;
        D.push_back(d);
    }
    std::vector<gcm_array> tables;
    GRC(run_gates(G, D, std::vector<uint64_t>{tm_rows, be_rows}, tables, out_q, true));
    gcm_layer L;
    memset(&L, 0, sizeof(L));
    L.p[0] = GCM_RESCALE;
    L.p[1] = (int64_t)n;
    L.p[2] = nf;
    for (int i = 0; i < nf; ++i) L.p[3 + i] = factors[i];
    L.n_arrays = k + 2;
    for (int j = 0; j < k; ++j) L.a[j] = G.keep(down[j].data(), down[j].size() * 2);
    L.a[k] = tables[0];
    L.a[k + 1] = tables[1];
    G.glayers.push_back(L);
    return DASHB_OK;
}

/* Legacy rescale (layer/garbled_rescale.h:27-35,52-75; RescaleGadget with
 * use_sign_base_extension, rescale_gadget.h:114-243): l iterations of
 * "divide by 2, regenerate the mod-2 residue with a full sign gadget".  Each
 * iteration is two launches: the trans-modulus gates (their outputs become the
 * current tensor), then the sign gadget on that tensor. */
int garble_rescale_legacy(Garbler& G, int l_iter) {
    const int k = G.k;
    if (l_iter < 1 || l_iter > 16) return report_error(DASHB_EINVAL, "garbler: bad number of legacy rescale iterations");
    if (G.crt[0] != 2) return report_error(DASHB_EINVAL, "garbler: legacy rescale needs modulus 2 first in the CRT base (the reference writes the regenerated residue to index 0, rescale_gadget.h:219-222)");
    if (G.n_mrs < 1) return report_error(DASHB_EINVAL, "garbler: legacy rescale needs an MRS base");
    const uint64_t n = G.cur_n;
    std::vector<uint32_t> d_off(k);
    std::vector<Label> down(k);
    for (int j = 0; j < k; ++j) {
        const int q = G.crt[j];
        const Label base = G.random_label(q);
        d_off[j] = G.add_const_label(base);
        down[j] = G.lin(G.R[q], pmod(G.P / 4, q), base, 1, q);
    }
    GRC(G.upload_labels());
    gcm_layer L;
    memset(&L, 0, sizeof(L));
    L.p[0] = GCM_RESCALE_LEGACY;
    L.p[1] = (int64_t)n;
    L.p[2] = l_iter;
    L.n_arrays = k + 5 * l_iter;
    for (int j = 0; j < k; ++j) L.a[j] = G.keep(down[j].data(), down[j].size() * 2);
    std::vector<int> out_q(k);
    for (int j = 0; j < k; ++j) out_q[j] = G.crt[j];
    const uint64_t tag0 = G.layer_tag;
    for (int it = 0; it < l_iter; ++it) {
        /* steps 1-2.3: upshift, project residue mod 2 into every other modulus, subtract, halve */
        G.layer_tag = tag0 | ((uint64_t)(2 * it + 1) << 16);
        std::vector<GateDesc> D;
        uint64_t tm_rows = 0;
        for (int j = 1; j < k; ++j) {
            const int q = G.crt[j];
            GateDesc d = blank_desc();
            d.q_in = 2;
            d.in[d.n_in++] = term_plane(0, 1);
            d.in[d.n_in++] = term_const(G.u_off[0], 1);
            d.q_out = (uint32_t)q;
            d.out[d.n_out++] = term_prf(tag(TAG_T, 0, j), 1);
            d.f = GB_F_ID;
            d.table = 0;
            d.table_off = tm_rows;
            d.table_stride = 2;
            tm_rows += n * 2;
            D.push_back(d);
            const int64_t inv2 = mul_inv_ll(2, q);
            GateDesc o = blank_desc();
            o.q_out = (uint32_t)q;
            o.out[o.n_out++] = term_plane((uint32_t)j, inv2);
            o.out[o.n_out++] = term_const(G.u_off[j], inv2);
            o.out[o.n_out++] = term_prf(tag(TAG_T, 0, j), pmod(-inv2, q));
            o.out_plane = j;
            D.push_back(o);
        }
        GateDesc z = blank_desc(); /* step 2.3 */
        z.q_out = 2;
        z.out[z.n_out++] = term_const(G.z_off[2], 1);
        z.out_plane = 0;
        D.push_back(z);
        std::vector<gcm_array> t1;
        GRC(run_gates(G, D, std::vector<uint64_t>{tm_rows}, t1, out_q, true));
        L.a[k + 5 * it] = t1[0];
        /* step 3: sign gadget with bounds (1, 0) regenerates residue 2; step 4: downshift */
        G.layer_tag = tag0 | ((uint64_t)(2 * it + 2) << 16);
        D.clear();
        uint64_t rows[4];
        sign_gadget_descs(G, D, std::vector<int>{2}, 1, 0, false, rows);
        GateDesc s0 = blank_desc();
        s0.q_out = 2;
        s0.out[s0.n_out++] = term_prf(tag(TAG_S, 0, 0), 1);
        s0.out[s0.n_out++] = term_const(d_off[0], 1); /* minus the downshift base, mod 2 */
        s0.out_plane = 0;
        D.push_back(s0);
        for (int j = 1; j < k; ++j) {
            GateDesc o = blank_desc();
            o.q_out = (uint32_t)G.crt[j];
            o.out[o.n_out++] = term_plane((uint32_t)j, 1);
            o.out[o.n_out++] = term_const(d_off[j], G.crt[j] - 1);
            o.out_plane = j;
            D.push_back(o);
        }
        std::vector<gcm_array> t2;
        GRC(run_gates(G, D, std::vector<uint64_t>{rows[0], rows[1], rows[2], rows[3]}, t2, out_q, true));
        for (int a = 0; a < 4; ++a) L.a[k + 5 * it + 1 + a] = t2[a];
    }
    G.layer_tag = tag0;
    G.glayers.push_back(L);
    return DASHB_OK;
}

/* Dense / conv2d: the output base labels are the layer evaluated on the
 * input base labels with the zero label as bias (layer/garbled_dense.h:
 * 89-118, garbled_conv2d.h:86-112); done by the evaluator itself, batch 1. */
int garble_linear(Garbler& G, const dashb_plain_layer& pl) {
    const int k = G.k;
    const bool conv = pl.p[0] == DASHB_LAYER_CONV2D;
    uint64_t M, nw;
    if (conv) {
        const uint64_t in_w = (uint64_t)pl.p[1], in_h = (uint64_t)pl.p[2], ch = (uint64_t)pl.p[3], fl = (uint64_t)pl.p[4],
                       fw = (uint64_t)pl.p[5], fh = (uint64_t)pl.p[6];
        if (in_w * in_h * ch != G.cur_n) return report_error(DASHB_EINVAL, "garbler: conv2d input size mismatch");
        M = fl;
        nw = fl * ch * fw * fh;
    } else {
        M = (uint64_t)pl.p[1];
        if ((uint64_t)pl.p[2] != G.cur_n) return report_error(DASHB_EINVAL, "garbler: dense K=%lld but the previous layer has %llu labels", (long long)pl.p[2], (unsigned long long)G.cur_n);
        nw = M * G.cur_n;
    }
    if (!pl.weights || !pl.biases) return report_error(DASHB_EINVAL, "garbler: linear layer without weights or biases");
    /* weights reduced to [0, P) as the reference stores them (garbled_dense.h:93-94) */
    std::vector<int64_t> W(nw);
    for (uint64_t i = 0; i < nw; ++i) W[i] = pmod(pl.weights[i], G.P);
    /* bias labels: b * R + Z (garble_values, garbled_circuit_interface.h:301-318) */
    std::vector<std::vector<int16_t>> bias(k), zbias(k);
    std::vector<const int16_t*> bias_p(k), zbias_p(k), zero_p(G.max_modulus - 1), in_p(k);
    for (int j = 0; j < k; ++j) {
        const int q = G.crt[j];
        const uint32_t nc = G.nc(q);
        bias[j].resize(M * nc);
        zbias[j].resize(M * nc);
        for (uint64_t r = 0; r < M; ++r) {
            const Label b = G.lin(G.R[q], pmod(pl.biases[r], q), G.Z[q], 1, q);
            memcpy(&bias[j][r * nc], b.data(), nc * 2);
            memcpy(&zbias[j][r * nc], G.Z[q].data(), nc * 2);
        }
        bias_p[j] = bias[j].data();
        zbias_p[j] = zbias[j].data();
        in_p[j] = G.cur[j].data();
    }
    for (int q = 2; q <= G.max_modulus; ++q) zero_p[q - 2] = G.Z[q].data();
    dashb_circuit* c = nullptr;
    GRC(dashb_circuit_create(&c, G.crt.data(), k, G.mrs.data(), G.n_mrs, G.max_modulus, zero_p.data(), nullptr, G.cur_n));
    int rc = conv ? dashb_add_conv2d(c, (uint64_t)pl.p[1], (uint64_t)pl.p[2], (uint64_t)pl.p[3], (uint64_t)pl.p[4], (uint64_t)pl.p[5],
                                     (uint64_t)pl.p[6], (uint64_t)pl.p[7], (uint64_t)pl.p[8], W.data(), zbias_p.data())
                  : dashb_add_dense(c, M, G.cur_n, (int)pl.p[3], W.data(), zbias_p.data());
    uint64_t out_n = 0;
    std::vector<std::vector<int16_t>> next(k);
    std::vector<int16_t*> out_p(k);
    if (rc == DASHB_OK) rc = dashb_finalize(c, 1);
    if (rc == DASHB_OK) {
        out_n = dashb_output_size(c);
        for (int j = 0; j < k; ++j) {
            next[j].resize(out_n * G.nc(G.crt[j]));
            out_p[j] = next[j].data();
        }
        rc = dashb_evaluate(c, in_p.data(), 1, out_p.data());
    }
    dashb_circuit_destroy(c);
    if (rc != DASHB_OK) return rc;
    gcm_layer L;
    memset(&L, 0, sizeof(L));
    for (int i = 0; i < 9; ++i) L.p[i] = pl.p[i];
    if (!conv) L.p[2] = (int64_t)G.cur_n;
    L.n_arrays = 1 + k;
    L.a[0] = G.keep(W.data(), W.size() * 8);
    for (int j = 0; j < k; ++j) L.a[1 + j] = G.keep(bias[j].data(), bias[j].size() * 2);
    G.glayers.push_back(L);
    G.cur.swap(next);
    G.cur_n = out_n;
    return DASHB_OK;
}

/* gen_decoding_information (garbled_circuit_interface.h:386-415): for every
 * output and every value v of its modulus, H(compress(base + v * R)). */
int garble_decoding_info(Garbler& G) {
    const int k = G.k;
    std::vector<GateDesc> D;
    std::vector<uint64_t> rows(k);
    for (int j = 0; j < k; ++j) {
        GateDesc d = blank_desc();
        d.q_in = (uint32_t)G.crt[j];
        d.in[d.n_in++] = term_plane((uint32_t)j, 1);
        d.mode = GB_MODE_DEC;
        d.table = (uint32_t)j;
        D.push_back(d);
        rows[j] = (uint64_t)G.crt[j] * G.cur_n;
    }
    std::vector<gcm_array> tables;
    GRC(run_gates(G, D, rows, tables, std::vector<int>(), false));
    G.g.has_dec = 1;
    for (int j = 0; j < k; ++j) G.g.dec[j] = tables[j];
    return DASHB_OK;
}

}  // namespace

/* downloads every device-resident array of the circuit and writes the file */
/* Writes the material.  With `secrets_path` the file at `path` is what the
 * evaluating party may hold (tables, weights, bias / zero / shift labels)
 * and everything that opens inputs or outputs -- the input base labels, the
 * offset labels R_q and the decoding information -- goes to `secrets_path`
 * alone (a DGCM file without layers). */
static int save_material(Garbler& G, const char* path, const char* secrets_path) {
    gcm_circuit g = G.g;
    std::vector<gcm_layer> layers = G.glayers;
    std::vector<std::unique_ptr<std::vector<uint8_t>>> host;
    auto to_host = [&](gcm_array& a) -> int {
        if (!G.on_device(a.data)) return DASHB_OK;
        host.emplace_back(new std::vector<uint8_t>(a.nbytes));
        if (a.nbytes) GCU(cudaMemcpy(host.back()->data(), a.data, a.nbytes, cudaMemcpyDeviceToHost));
        a.data = host.back()->data();
        return DASHB_OK;
    };
    for (gcm_layer& L : layers)
        for (int a = 0; a < L.n_arrays; ++a) GRC(to_host(L.a[a]));
    for (int j = 0; j < g.k; ++j) GRC(to_host(g.dec[j]));
    g.layers = layers.data();
    if (secrets_path) {
        gcm_circuit pub = g;
        pub.has_dec = 0;
        pub.has_ginp = 0;
        if (gcm_save(&pub, path) != 0) return report_error(DASHB_EIO, "garbler: cannot write %s", path);
        gcm_circuit sec = g;
        sec.n_layers = 0;
        sec.layers = nullptr;
        if (gcm_save(&sec, secrets_path) != 0) return report_error(DASHB_EIO, "garbler: cannot write %s", secrets_path);
        return DASHB_OK;
    }
    if (gcm_save(&g, path) != 0) return report_error(DASHB_EIO, "garbler: cannot write %s", path);
    return DASHB_OK;
}

static int garble_impl(const int32_t* crt_base, int k, const int32_t* mrs_base, int n_mrs, int max_modulus,
                       uint64_t input_size, const dashb_plain_layer* layers, int n_layers, const uint8_t seed[16],
                       const char* path, const char* secrets_path, dashb_circuit** circuit, dashb_garble_stats* stats) {
    const auto t_start = std::chrono::steady_clock::now();
    if (!crt_base || k < 1 || k > MAXK || n_mrs < 0 || n_mrs > MAXM || (n_mrs && !mrs_base) || !layers || n_layers < 1 || !seed ||
        (!path && !circuit) || input_size == 0)
        return report_error(DASHB_EINVAL, "dashb_garble: bad arguments");
    if (initialised_device() < 0) return report_error(DASHB_ENODEV, "dashb_garble: call dashb_init() first");
    Garbler G;
    memset(&G.g, 0, sizeof(G.g));
    memset(&G.st, 0, sizeof(G.st));
    G.k = k;
    G.n_mrs = n_mrs;
    G.crt.assign(crt_base, crt_base + k);
    G.mrs.assign(mrs_base, mrs_base + n_mrs);
    int need = 2;
    for (int j = 0; j < k; ++j) {
        if (G.crt[j] < 2 || G.crt[j] > 256) return report_error(DASHB_EINVAL, "garbler: CRT modulus %d outside [2,256]", G.crt[j]);
        need = std::max(need, G.crt[j]);
        G.P *= G.crt[j];
    }
    for (int i = 0; i < n_mrs; ++i) need = std::max(need, i == 0 ? G.mrs[0] : (k + 1) * G.mrs[i]);
    /* a given max_modulus only ever raises the computed one (garbled_circuit_interface.h:128-135) */
    if (max_modulus < need) max_modulus = need;
    if (max_modulus > 1024) return report_error(DASHB_EINVAL, "garbler: max_modulus %d too large", max_modulus);
    if (need > 256) return report_error(DASHB_EINVAL, "garbler: gadget modulus %d above 256", need);
    G.max_modulus = max_modulus;
    G.mods.resize(max_modulus + 1);
    for (int q = 2; q <= max_modulus; ++q) G.mods[q] = modinfo_for((uint32_t)q);
    GCU(cudaEventCreate(&G.ev0));
    GCU(cudaEventCreate(&G.ev1));
    /* PRF key = the seed; host-side stream for the circuit-wide labels */
    const uint32_t* rk_fixed;
    host_aes_tables(&G.te0, &G.te1, &G.te2, &G.te3, &G.sbox, &rk_fixed);
    expand_key(seed, G.sbox, G.rk_prf);
    GCU(garbler_set_prf_key(G.rk_prf));
    /* zero labels and offset labels of every modulus (init_garbling, :933-936; R_q[0] = 1, label_tensor.h:1303-1307) */
    G.R.resize(max_modulus + 1);
    G.Z.resize(max_modulus + 1);
    G.r_off.assign(max_modulus + 1, 0);
    G.z_off.assign(max_modulus + 1, 0);
    std::vector<int16_t> zero_all;
    for (int q = 2; q <= max_modulus; ++q) {
        G.Z[q] = G.random_label(q);
        G.R[q] = G.random_label(q);
        G.R[q][0] = 1;
        zero_all.insert(zero_all.end(), G.Z[q].begin(), G.Z[q].end());
        if (q <= 256) {
            G.z_off[q] = G.add_const_label(G.Z[q]);
            G.r_off[q] = G.add_const_label(G.R[q]);
        }
    }
    /* upshift base labels and upshift labels (:938-954) */
    G.U.resize(k);
    G.u_off.resize(k);
    std::vector<int16_t> up_all;
    for (int j = 0; j < k; ++j) {
        const int q = G.crt[j];
        G.U[j] = G.random_label(q);
        G.u_off[j] = G.add_const_label(G.U[j]);
        const Label up = G.lin(G.R[q], pmod(G.P / 2, q), G.U[j], 1, q);
        up_all.insert(up_all.end(), up.begin(), up.end());
    }
    GRC(G.upload_labels());
    GCU(G.d_r_off.alloc(G.r_off.size() * 4));
    GCU(cudaMemcpy(G.d_r_off.p, G.r_off.data(), G.r_off.size() * 4, cudaMemcpyHostToDevice));
    GCU(G.d_mods.alloc(G.mods.size() * sizeof(ModInfo)));
    GCU(cudaMemcpy(G.d_mods.p, G.mods.data(), G.mods.size() * sizeof(ModInfo), cudaMemcpyHostToDevice));
    if (n_mrs) {
        const std::vector<uint8_t> lut = approx_lookup(G.crt, G.mrs, G.lut_off);
        GCU(G.d_lut.alloc(lut.size()));
        GCU(cudaMemcpy(G.d_lut.p, lut.data(), lut.size(), cudaMemcpyHostToDevice));
    }
    /* input base labels (init_base_label, :833-839) */
    G.cur_n = input_size;
    G.cur.resize(k);
    for (int j = 0; j < k; ++j) {
        const int q = G.crt[j];
        const uint32_t nc = G.nc(q);
        G.cur[j].resize(input_size * nc);
        for (uint64_t e = 0; e < input_size; ++e) {
            const Label l = G.random_label(q);
            memcpy(&G.cur[j][e * nc], l.data(), nc * 2);
        }
        G.g.in_base[j] = G.keep(G.cur[j].data(), G.cur[j].size() * 2);
        G.g.in_offset[j] = G.keep(G.R[q].data(), G.R[q].size() * 2);
    }
    G.g.has_ginp = 1;
    for (int l = 0; l < n_layers; ++l) {
        G.layer_tag = (uint64_t)l + 1;
        const dashb_plain_layer& pl = layers[l];
        switch ((int)pl.p[0]) {
            case DASHB_LAYER_DENSE:
            case DASHB_LAYER_CONV2D: GRC(garble_linear(G, pl)); break;
            case DASHB_LAYER_RELU: GRC(garble_sign_layer(G, true)); break;
            case DASHB_LAYER_SIGN: GRC(garble_sign_layer(G, false)); break;
            case DASHB_LAYER_RESCALE: GRC(garble_rescale(G, &pl.p[3], (int)pl.p[2])); break;
            case DASHB_LAYER_RESCALE_LEGACY: GRC(garble_rescale_legacy(G, (int)pl.p[2])); break;
            default: return report_error(DASHB_EINVAL, "garbler: layer type %d is not supported", (int)pl.p[0]);
        }
    }
    G.layer_tag = (uint64_t)n_layers + 1;
    GRC(garble_decoding_info(G));
    G.g.k = k;
    G.g.n_mrs = n_mrs;
    G.g.max_modulus = max_modulus;
    G.g.input_size = input_size;
    for (int j = 0; j < k; ++j) G.g.crt_base[j] = G.crt[j];
    for (int i = 0; i < n_mrs; ++i) G.g.mrs_base[i] = G.mrs[i];
    G.g.zero = G.keep(zero_all.data(), zero_all.size() * 2);
    G.g.upshift = G.keep(up_all.data(), up_all.size() * 2);
    G.g.n_layers = (int)G.glayers.size();
    if (path) GRC(save_material(G, path, secrets_path));
    if (circuit) {
        /* straight into an evaluator circuit: the tables are copied device to device */
        G.g.layers = G.glayers.data();
        const int rc = dashb_circuit_from_gcm(&G.g, circuit);
        G.g.layers = nullptr;
        if (rc != DASHB_OK) return rc;
    }
    G.st.total_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_start).count();
    if (stats) *stats = G.st;
    return DASHB_OK;
}

extern "C" int dashb_garble_to_file(const int32_t* crt_base, int k, const int32_t* mrs_base, int n_mrs, int max_modulus,
                                    uint64_t input_size, const dashb_plain_layer* layers, int n_layers, const uint8_t seed[16],
                                    const char* path, dashb_garble_stats* stats) {
    if (!path) return report_error(DASHB_EINVAL, "dashb_garble_to_file: no path");
    return garble_impl(crt_base, k, mrs_base, n_mrs, max_modulus, input_size, layers, n_layers, seed, path, nullptr, nullptr, stats);
}

extern "C" int dashb_garble_to_files(const int32_t* crt_base, int k, const int32_t* mrs_base, int n_mrs, int max_modulus,
                                     uint64_t input_size, const dashb_plain_layer* layers, int n_layers, const uint8_t seed[16],
                                     const char* evaluator_path, const char* garbler_path, dashb_garble_stats* stats) {
    if (!evaluator_path || !garbler_path) return report_error(DASHB_EINVAL, "dashb_garble_to_files: two paths are needed");
    return garble_impl(crt_base, k, mrs_base, n_mrs, max_modulus, input_size, layers, n_layers, seed, evaluator_path, garbler_path,
                       nullptr, stats);
}

extern "C" int dashb_garble_to_circuit(const int32_t* crt_base, int k, const int32_t* mrs_base, int n_mrs, int max_modulus,
                                       uint64_t input_size, const dashb_plain_layer* layers, int n_layers,
                                       const uint8_t seed[16], dashb_circuit** circuit, dashb_garble_stats* stats) {
    if (!circuit) return report_error(DASHB_EINVAL, "dashb_garble_to_circuit: no circuit pointer");
    return garble_impl(crt_base, k, mrs_base, n_mrs, max_modulus, input_size, layers, n_layers, seed, nullptr, nullptr, circuit, stats);
}
