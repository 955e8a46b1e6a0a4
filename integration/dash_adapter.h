/*
 * dash_adapter.h -- the glue a Dash maintainer adds to drive the B200
 * evaluator from an existing, already garbled `GarbledCircuit`.
 *
 * It includes the *unmodified* reference headers, walks
 * gc->get_garbled_layer() (garbling/garbled_circuit_interface.h:760) and
 * describes every layer's garbled material as a flat `gcm_circuit` view
 * (include/dash_gcm_format.h) that points straight into the reference's own
 * arrays -- nothing is copied except the small concatenated label blocks.
 * From the view one can
 *   - write a DGCM file            (dash_view_save), or
 *   - build a device circuit       (dash_b200_move, needs libdash_b200.so).
 *
 * The ciphertext arrays of the gadgets are private members without getters
 * (layer/garbled_relu.h:50-52, gadgets/sign_gadget.h:234-253,
 * gadgets/rescale_gadget.h:33-36, gadgets/base_extension_gadget.h:379-388,
 * layer/garbled_projection.h:25).  They are reached with explicit template
 * instantiations naming the members (access checks do not apply to the
 * arguments of an explicit instantiation, [temp.spec]/6), so no compiler
 * flag and no edit of the reference is needed.  A maintainer would instead
 * add one `friend` line or a getter per class.
 *
 * Requires: garbling/garbled_circuit.h already included by the TU.
 */
#ifndef DASH_ADAPTER_H
#define DASH_ADAPTER_H

#include <algorithm>
#include <map>
#include <memory>
#include <numeric>
#include <stdexcept>
#include <vector>

#include "dash_gcm_format.h"

struct DashMaterialView {
    gcm_circuit c{};
    std::vector<gcm_layer> layers;
    std::vector<int16_t> zero_buf, upshift_buf;
    // keeps the concatenated per-layer label blocks alive
    std::vector<std::unique_ptr<std::vector<int16_t>>> owned;

    DashMaterialView() = default;
    DashMaterialView(const DashMaterialView&) = delete;
    DashMaterialView& operator=(const DashMaterialView&) = delete;
};

namespace dash_adapter_detail {

template <typename Tag, typename Tag::type M>
struct Rob {
    friend typename Tag::type get(Tag) { return M; }
};
#define DASH_ADAPTER_MEMBER(Tag, Class, Type, Member) \
    struct Tag {                                      \
        typedef Type Class::*type;                    \
        friend type get(Tag);                         \
    };                                                \
    template struct Rob<Tag, &Class::Member>;

DASH_ADAPTER_MEMBER(SG_n, SignGadget, size_t, m_nr_inputs)
DASH_ADAPTER_MEMBER(SG_out, SignGadget, std::vector<crt_val_t>, m_out_moduli)
DASH_ADAPTER_MEMBER(SG_approx, SignGadget, __uint128_t*, m_approx_res_ciphers)
DASH_ADAPTER_MEMBER(SG_c1, SignGadget, __uint128_t*, m_1_cast_ciphers)
DASH_ADAPTER_MEMBER(SG_c2, SignGadget, __uint128_t*, m_2_cast_ciphers)
DASH_ADAPTER_MEMBER(SG_sign, SignGadget, __uint128_t*, m_sign_ciphers)
DASH_ADAPTER_MEMBER(GD_w, GarbledDense, ScalarTensor<q_val_t>, m_qe_weights)
DASH_ADAPTER_MEMBER(GC_w, GarbledConv2d, ScalarTensor<q_val_t>, m_qe_weights)
DASH_ADAPTER_MEMBER(GR_sg, GarbledRelu, SignGadget*, m_sign_gadget)
DASH_ADAPTER_MEMBER(GR_g, GarbledRelu, __uint128_t*, m_garbler_ciphers)
DASH_ADAPTER_MEMBER(GR_e, GarbledRelu, __uint128_t*, m_evaluator_ciphers)
DASH_ADAPTER_MEMBER(GS_sg, GarbledSign, SignGadget*, m_sign_gadget)
DASH_ADAPTER_MEMBER(GRS_g, GarbledRescale, std::vector<RescaleGadget*>, m_rescale_gadgets)
DASH_ADAPTER_MEMBER(RG_tm, RescaleGadget, __uint128_t*, m_trans_mod_ciphers)
DASH_ADAPTER_MEMBER(RG_sg, RescaleGadget, SignGadget*, m_sign_gadget)
DASH_ADAPTER_MEMBER(RG_be, RescaleGadget, BaseExtensionGadget*, m_be_gadget)
DASH_ADAPTER_MEMBER(BE_c, BaseExtensionGadget, __uint128_t*, m_proj_gates_base_change_loop_ciphers)
DASH_ADAPTER_MEMBER(GP_c, GarbledProjection, std::vector<__uint128_t*>, m_ciphers)
DASH_ADAPTER_MEMBER(GBE_g, GarbledBaseExtension, BaseExtensionGadget*, m_base_extension_gadget)
typedef std::map<std::vector<size_t>, MaxGadget*> MaxGadgetMap;
DASH_ADAPTER_MEMBER(GMP_g, GarbledMaxPool2d, MaxGadgetMap, m_max_gadgets)
DASH_ADAPTER_MEMBER(MG_r, MaxGadget, ReluGadget*, m_relu_gadget)
DASH_ADAPTER_MEMBER(RLG_sg, ReluGadget, SignGadget*, m_sign_gadget)
DASH_ADAPTER_MEMBER(RLG_g, ReluGadget, __uint128_t*, M_GARBLER_CIPHERS)
DASH_ADAPTER_MEMBER(RLG_e, ReluGadget, __uint128_t*, M_EVALUATOR_CIPHERS)
#undef DASH_ADAPTER_MEMBER

inline gcm_array arr(const void* p, uint64_t nbytes) { return gcm_array{p, nbytes}; }

inline void push(gcm_layer& L, const void* p, uint64_t nbytes) {
    if (L.n_arrays >= GCM_MAX_ARRAYS) throw std::runtime_error("dash_adapter: too many arrays");
    L.a[L.n_arrays++] = arr(p, nbytes);
}

inline void push_labels(gcm_layer& L, std::vector<LabelTensor*>* labels) {
    for (auto* l : *labels) push(L, l->get_components(), l->size() * sizeof(crt_val_t));
}

// SignGadget tables, in the order approx, cast1, cast2, sign.
inline void push_sign_gadget(gcm_layer& L, SignGadget* sg, GarbledCircuitInterface* gc) {
    const size_t k = gc->get_crt_base().size();
    const size_t n = sg->*get(SG_n());
    const size_t S = std::accumulate(gc->get_crt_base().begin(), gc->get_crt_base().end(), size_t{0});
    const size_t n_mrs = gc->get_mrs_base().size();
    const size_t Mp = std::accumulate(gc->get_mrs_base().begin() + 1, gc->get_mrs_base().end(), size_t{0});
    push(L, sg->*get(SG_approx()), n * n_mrs * S * 16);
    push(L, sg->*get(SG_c1()), n * (k + 1) * Mp * 16);
    // allocated (n_mrs-1)x larger than used; element stride is (k+1)*M'
    // (gadgets/sign_gadget.h:443-445)
    push(L, sg->*get(SG_c2()), n * (k + 1) * Mp * 16);
    push(L, sg->*get(SG_sign()), n * (sg->*get(SG_out())).size() * gc->get_mrs_base().at(0) * 16);
}

}  // namespace dash_adapter_detail

/* Fills `v` from a garbled circuit.  The view borrows the circuit's arrays:
 * keep `gc` alive (and do not re-garble it) while the view is in use. */
inline void dash_view_from_circuit(GarbledCircuit* gc, DashMaterialView& v, bool with_decoding = true) {
    using namespace dash_adapter_detail;
    auto& crt = gc->get_crt_base();
    auto& mrs = gc->get_mrs_base();
    const size_t k = crt.size();
    gcm_circuit& c = v.c;
    memset(&c, 0, sizeof(c));
    c.k = (int)k;
    c.n_mrs = (int)mrs.size();
    c.max_modulus = gc->get_max_modulus();
    c.input_size = gc->get_circuit()->get_input_size();
    for (size_t j = 0; j < k; ++j) c.crt_base[j] = crt[j];
    for (size_t i = 0; i < mrs.size(); ++i) c.mrs_base[i] = mrs[i];

    v.zero_buf.clear();
    for (auto* z : *gc->get_zero_labels()) {
        auto comps = z->get_components();
        v.zero_buf.insert(v.zero_buf.end(), comps, comps + z->size());
    }
    c.zero = arr(v.zero_buf.data(), v.zero_buf.size() * 2);

    bool has_rescale = false;
    for (auto* gl : gc->get_garbled_layer())
        if (gl->get_layer()->get_type() == Layer::rescale) has_rescale = true;
    v.upshift_buf.clear();
    if (has_rescale) {
        for (size_t j = 0; j < k; ++j) {
            auto* u = gc->get_upshift_label(j);
            v.upshift_buf.insert(v.upshift_buf.end(), u->get_components(), u->get_components() + u->size());
        }
    }
    c.upshift = arr(v.upshift_buf.data(), v.upshift_buf.size() * 2);

    v.layers.clear();
    for (auto* gl : gc->get_garbled_layer()) {
        gcm_layer L;
        memset(&L, 0, sizeof(L));
        Layer* pl = gl->get_layer();
        switch (pl->get_type()) {
            case Layer::dense: {
                auto* g = dynamic_cast<GarbledDense*>(gl);
                L.p[0] = GCM_DENSE;
                L.p[1] = (int64_t)pl->get_output_size();
                L.p[2] = (int64_t)pl->get_input_size();
                L.p[3] = dynamic_cast<Dense*>(pl)->get_channel_tf();
                push(L, (g->*get(GD_w())).data(), (g->*get(GD_w())).size() * sizeof(q_val_t));
                push_labels(L, g->get_qe_bias_label());
                break;
            }
            case Layer::conv2d: {
                auto* g = dynamic_cast<GarbledConv2d*>(gl);
                Conv2d* cv = dynamic_cast<Conv2d*>(pl);
                L.p[0] = GCM_CONV2D;
                L.p[1] = cv->get_input_width();
                L.p[2] = cv->get_input_height();
                L.p[3] = cv->get_channel();
                L.p[4] = cv->get_filter();
                L.p[5] = cv->get_filter_width();
                L.p[6] = cv->get_filter_height();
                L.p[7] = cv->get_stride_width();
                L.p[8] = cv->get_stride_height();
                push(L, (g->*get(GC_w())).data(), (g->*get(GC_w())).size() * sizeof(q_val_t));
                push_labels(L, g->get_qe_bias_label());
                break;
            }
            case Layer::approx_relu: {
                auto* g = dynamic_cast<GarbledRelu*>(gl);
                const size_t n = pl->get_input_size();
                const size_t S = std::accumulate(crt.begin(), crt.end(), size_t{0});
                L.p[0] = GCM_RELU;
                L.p[1] = (int64_t)n;
                push_sign_gadget(L, g->*get(GR_sg()), gc);
                push(L, g->*get(GR_g()), n * S * 16);
                push(L, g->*get(GR_e()), n * k * 3 * 16);
                break;
            }
            case Layer::sign: {
                auto* g = dynamic_cast<GarbledSign*>(gl);
                L.p[0] = GCM_SIGN;
                L.p[1] = (int64_t)pl->get_input_size();
                push_sign_gadget(L, g->*get(GS_sg()), gc);
                break;
            }
            case Layer::rescale: {
                auto* g = dynamic_cast<GarbledRescale*>(gl);
                const size_t n = pl->get_input_size();
                Rescale* rs = dynamic_cast<Rescale*>(pl);
                std::vector<crt_val_t> factors =
                    rs->get_use_sign_base_extension() ? std::vector<crt_val_t>{2} : rs->get_s();
                auto ds = std::make_unique<std::vector<int16_t>>();
                std::vector<size_t> ds_off;
                for (size_t j = 0; j < k; ++j) {
                    auto* d = gc->get_downshift_label(factors, j);
                    ds_off.push_back(ds->size());
                    ds->insert(ds->end(), d->get_components(), d->get_components() + d->size());
                }
                L.p[1] = (int64_t)n;
                for (size_t j = 0; j < k; ++j)
                    push(L, ds->data() + ds_off[j], (uint64_t)LabelTensor::get_nr_comps(crt[j]) * 2);
                v.owned.push_back(std::move(ds));
                if (rs->get_use_sign_base_extension()) {
                    L.p[0] = GCM_RESCALE_LEGACY;
                    L.p[2] = rs->get_l();
                    for (auto* rg : g->*get(GRS_g())) {
                        push(L, rg->*get(RG_tm()), n * (k - 1) * 2 * 16);
                        push_sign_gadget(L, rg->*get(RG_sg()), gc);
                    }
                } else {
                    L.p[0] = GCM_RESCALE;
                    L.p[2] = (int64_t)factors.size();
                    size_t tm_rows = 0;
                    for (size_t i = 0; i < factors.size(); ++i) {
                        L.p[3 + i] = factors[i];
                        tm_rows += (k - 1 - i) * n * (size_t)factors[i];
                    }
                    auto* rg = (g->*get(GRS_g())).at(0);
                    push(L, rg->*get(RG_tm()), tm_rows * 16);
                    auto* be = rg->*get(RG_be());
                    // rows per element: sum_i (k-i-1) * m_i over the swapped
                    // base (base_extension_gadget.h:77-85)
                    auto swapped = be->get_out_moduli();
                    size_t be_rows = 0;
                    for (size_t i = 0; i + factors.size() < k; ++i) be_rows += (k - i - 1) * (size_t)swapped[i];
                    push(L, be->*get(BE_c()), be_rows * n * 16);
                }
                break;
            }
            case Layer::base_extension: {
                auto* g = dynamic_cast<GarbledBaseExtension*>(gl);
                const size_t n = pl->get_input_size();
                const auto extra = dynamic_cast<BaseExtension*>(pl)->get_extra_moduli();
                L.p[0] = GCM_BASE_EXTENSION;
                L.p[1] = (int64_t)n;
                L.p[2] = (int64_t)extra.size();
                auto* be = g->*get(GBE_g());
                auto swapped = be->get_out_moduli();
                size_t be_rows = 0;
                for (size_t i = 0; i + extra.size() < k; ++i) be_rows += (k - i - 1) * (size_t)swapped[i];
                push(L, be->*get(BE_c()), be_rows * n * 16);
                for (size_t x = 0; x < extra.size(); ++x) {
                    L.p[3 + x] = extra[x];
                    // the labels the layer starts the extra residue from: the circuit's input base labels (garbled_base_extension.h:74-80)
                    const size_t idx = std::find(crt.begin(), crt.end(), extra[x]) - crt.begin();
                    auto* base = gc->get_base_label().at(idx);
                    push(L, base->get_components(), n * (size_t)LabelTensor::get_nr_comps(extra[x]) * sizeof(crt_val_t));
                }
                break;
            }
            case Layer::max_pool: {
                /* one MaxGadget -> ReluGadget -> SignGadget (one element each) per (output, kernel position):
                 * their arrays are concatenated into the six tables of a ReLU layer with outputs * k_w k_h
                 * elements, instance (output label o, l, m) at o * k_w k_h + l * k_h + m (the only copy this adapter makes) */
                auto* g = dynamic_cast<GarbledMaxPool2d*>(gl);
                MaxPool2d* mp = dynamic_cast<MaxPool2d*>(pl);
                // the garbled layer takes its geometry from the circuit's input labels (garbled_maxpool2d.h:66-68)
                const auto dims = gc->get_base_label().at(0)->get_dims();
                const size_t W = dims.at(0), H = dims.at(1), C = dims.at(2);
                const size_t kw = mp->get_kernel_width(), kh = mp->get_kernel_height();
                const size_t sw = mp->get_stride_width(), sh = mp->get_stride_height();
                const size_t ow = (W - kw) / sw + 1, oh = (H - kh) / sh + 1, T = kw * kh;
                const size_t S = std::accumulate(crt.begin(), crt.end(), size_t{0});
                const size_t n_mrs = mrs.size();
                const size_t Mp = std::accumulate(mrs.begin() + 1, mrs.end(), size_t{0});
                const size_t per[6] = {n_mrs * S, (k + 1) * Mp, (k + 1) * Mp, (size_t)mrs.at(0), S, k * 3};
                std::unique_ptr<std::vector<int16_t>> blk[6];
                for (int a = 0; a < 6; ++a) blk[a] = std::make_unique<std::vector<int16_t>>(ow * oh * C * T * per[a] * 8);
                auto& gadgets = g->*get(GMP_g());
                for (size_t i = 0; i < ow; ++i)
                    for (size_t j = 0; j < oh; ++j)
                        for (size_t z = 0; z < C; ++z)
                            for (size_t l = 0; l < kw; ++l)
                                for (size_t m = 0; m < kh; ++m) {
                                    // output (i, j, z) lives at label index (i * oh + j) * C + z (set_label, label_tensor.h:1290-1301)
                                    const size_t inst = ((i * oh + j) * C + z) * T + l * kh + m;
                                    ReluGadget* rg = gadgets.at({i, j, z, l, m})->*get(MG_r());
                                    SignGadget* sg = rg->*get(RLG_sg());
                                    const __uint128_t* src[6] = {sg->*get(SG_approx()), sg->*get(SG_c1()), sg->*get(SG_c2()),
                                                                 sg->*get(SG_sign()), rg->*get(RLG_g()), rg->*get(RLG_e())};
                                    for (int a = 0; a < 6; ++a)
                                        memcpy(reinterpret_cast<__uint128_t*>(blk[a]->data()) + inst * per[a], src[a], per[a] * 16);
                                }
                L.p[0] = GCM_MAXPOOL2D;
                L.p[1] = (int64_t)W, L.p[2] = (int64_t)H, L.p[3] = (int64_t)C;
                L.p[4] = (int64_t)kw, L.p[5] = (int64_t)kh, L.p[6] = (int64_t)sw, L.p[7] = (int64_t)sh;
                for (int a = 0; a < 6; ++a) {
                    push(L, blk[a]->data(), (uint64_t)ow * oh * C * T * per[a] * 16);
                    v.owned.push_back(std::move(blk[a]));
                }
                break;
            }
            case Layer::projection: {
                auto* g = dynamic_cast<GarbledProjection*>(gl);
                const size_t n = pl->get_input_size();
                Projection* pr = dynamic_cast<Projection*>(pl);
                L.p[0] = GCM_PROJECTION;
                L.p[1] = (int64_t)n;
                for (size_t j = 0; j < k; ++j) {
                    L.p[2 + j] = pr->get_in_moduli().at(j);
                    L.p[2 + k + j] = pr->get_out_moduli().at(j);
                    push(L, (g->*get(GP_c())).at(j), n * (size_t)pr->get_in_moduli().at(j) * 16);
                }
                break;
            }
            default:
                throw std::runtime_error(
                    "dash_adapter: layer type outside the B200 hot path (mult / mixed-mod mult / max layer)");
        }
        v.layers.push_back(L);
    }
    c.n_layers = (int)v.layers.size();
    c.layers = v.layers.data();

    c.has_dec = 0;
    if (with_decoding && !gc->get_decoding_information().empty()) {
        c.has_dec = 1;
        const size_t out = gc->get_circuit()->get_output_size();
        for (size_t j = 0; j < k; ++j) {
            auto blk = std::make_unique<std::vector<int16_t>>();
            blk->resize((size_t)crt[j] * out * 8);
            auto* dst = reinterpret_cast<__uint128_t*>(blk->data());
            for (int r = 0; r < crt[j]; ++r)
                memcpy(dst + (size_t)r * out, gc->get_decoding_information().at(j).at(r).data(), out * 16);
            c.dec[j] = arr(blk->data(), (uint64_t)crt[j] * out * 16);
            v.owned.push_back(std::move(blk));
        }
    }
}

/* Adds the garbler-side input encoding (base labels and offset labels) to the
 * view, for deployments where the party that evaluates also encodes inputs
 * (benchmarks, tests).  The arrays are borrowed from the circuit. */
inline void dash_view_add_input_encoding(GarbledCircuit* gc, DashMaterialView& v) {
    const size_t k = gc->get_crt_base().size();
    v.c.has_ginp = 1;
    for (size_t j = 0; j < k; ++j) {
        auto* base = gc->get_base_label().at(j);
        v.c.in_base[j] = dash_adapter_detail::arr(base->get_components(), base->size() * sizeof(crt_val_t));
        LabelSlice off = gc->get_label_offset(gc->get_crt_base().at(j));
        v.c.in_offset[j] = dash_adapter_detail::arr(off.get_components(), off.get_nr_comps() * sizeof(crt_val_t));
    }
}

inline int dash_view_save(const DashMaterialView& v, const char* path) {
    gcm_circuit c = v.c;
    c.layers = const_cast<gcm_layer*>(v.layers.data());
    return gcm_save(&c, path);
}

#ifdef DASH_B200_GCM_H
/* In-process route, the drop-in for gc->cuda_move(): builds the device
 * circuit through the C ABI straight from the view (no file).  Include
 * dash_b200_gcm.h before this header to get it; see INTEGRATION.md. */
inline int dash_b200_move(GarbledCircuit* gc, dashb_circuit** out, uint64_t max_batch) {
    DashMaterialView v;
    dash_view_from_circuit(gc, v);
    int rc = dashb_circuit_from_gcm(&v.c, out);
    if (rc != DASHB_OK) return rc;
    return dashb_finalize(*out, max_batch);
}
#endif

#endif /* DASH_ADAPTER_H */
