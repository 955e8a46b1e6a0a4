/*
 * ref_driver.cu -- TEST INFRASTRUCTURE, not part of the product.
 *
 * A thin C-callable harness around the UNMODIFIED reference (UzL-ITS/dash,
 * headers included from /root/reference/dash/include at build time; nothing
 * is copied into this repository).  It builds plaintext circuits layer by
 * layer, garbles them with the reference's own garbler made deterministic
 * by a seeded stand-in for _rdrand16_step (garbling/label_tensor.h:1405),
 * runs the reference's own CPU evaluator (GarbledCircuitInterface::
 * cpu_evaluate, garbling/garbled_circuit_interface.h:345-380) and exports
 * the garbled material as a DGCM file through integration/dash_adapter.h.
 *
 * Built by oracle/Makefile into oracle/_ref/libdash_ref.so (git-ignored).
 * Only tests/, __graft_entry__.smoke() and bench.py's reference/cpu_baseline
 * legs may load it.  nvcc (-x cu) is needed only because the reference
 * headers pull in __device__ declarations; no CUDA call is made.
 */
#include <immintrin.h>
#include <omp.h>

#include <chrono>
#include <cstdint>
#include <cstring>
#include <iostream>
#include <random>
#include <string>

namespace {
std::mt19937_64 g_rng{0};
}
static inline int dash_ref_seeded_rdrand16(unsigned short* p) {
    *p = (unsigned short)(g_rng() & 0xffff);
    return 1;
}
#define _rdrand16_step dash_ref_seeded_rdrand16

#include "circuit/circuit.h"
#include "circuit/layer/conv2d.h"
#include "circuit/layer/dense.h"
#include "circuit/layer/base_extension.h"
#include "circuit/layer/flatten.h"
#include "circuit/layer/max_pool2d.h"
#include "circuit/layer/projection.h"
#include "circuit/layer/relu.h"
#include "circuit/layer/rescale.h"
#include "circuit/layer/sign.h"
#include "circuit/scalar_tensor.h"
#include "garbling/garbled_circuit.h"
#include "misc/datatypes.h"

#include "../integration/dash_adapter.h"
#define ONNX_LITE_BUILD_CIRCUIT
#include "../integration/dash_onnx_lite.h"

namespace {

crt_val_t proj_identity(crt_val_t x, void*) { return x; }
crt_val_t proj_plus_one(crt_val_t x, void*) { return x + 1; }
crt_val_t proj_times_two(crt_val_t x, void*) { return 2 * x; }
crt_val_t proj_half(crt_val_t x, void*) { return x / 2; }
crt_val_t proj_half_minus_ten(crt_val_t x, void*) { return x / 2 - 10; } /* test_projection.h:15 */
crt_val_t proj_quarter(crt_val_t x, void*) { return x / 4; }             /* test_projection.h:16 */

struct RefHandle {
    std::vector<Layer*> layers;  // owned by circuit once built
    Circuit* circuit{nullptr};
    GarbledCircuit* gc{nullptr};
    std::string err;
    ~RefHandle() {
        delete gc;
        if (circuit)
            delete circuit;
        else
            for (auto* l : layers) delete l;
    }
};

QuantizationMethod qm(int m) {
    return m == 0 ? QuantizationMethod::SimpleQuant
                  : (m == 1 ? QuantizationMethod::ScaleQuant : QuantizationMethod::ScaleQuantPlus);
}

std::vector<LabelTensor*>* labels_from_raw(GarbledCircuit* gc, const std::vector<crt_val_t>& moduli,
                                           const int16_t* const* raw, dim_t dims) {
    auto* v = new std::vector<LabelTensor*>();
    for (size_t j = 0; j < moduli.size(); ++j) {
        size_t nc = LabelTensor::get_nr_comps(moduli[j]);
        v->push_back(new LabelTensor(moduli[j], const_cast<crt_val_t*>(raw[j]), nc, dims));
    }
    return v;
}

void free_labels(std::vector<LabelTensor*>* v) {
    for (auto* l : *v) delete l;
    delete v;
}

}  // namespace

extern "C" {

void* ref_new(void) { return new RefHandle(); }
void ref_free(void* h) { delete static_cast<RefHandle*>(h); }
const char* ref_error(void* h) { return static_cast<RefHandle*>(h)->err.c_str(); }

/* layers of an ONNX model (integration/dash_onnx_lite.h), appended like the ref_add_* calls */
int ref_load_onnx(void* hv, const char* path, int q_method, float q_const, int q_param) {
    auto* h = static_cast<RefHandle*>(hv);
    dash_onnx::Model m;
    if (!dash_onnx::read_model(path, m)) {
        h->err = "cannot parse ONNX model";
        return -1;
    }
    std::vector<Layer*> layers = dash_onnx::build_layers(m, qm(q_method), q_const, q_param);
    if (layers.empty()) {
        h->err = "ONNX model has no supported layers or is malformed";
        return -1;
    }
    for (auto* l : layers) h->layers.push_back(l);
    return (int)layers.size();
}

/* Plaintext layer introspection for the ONNX-reader test (what
 * dash/test/test_onnx_modelloader.h:10-43 asserts on): kind of plaintext
 * layer `layer` (Layer::LayerType as int), dims of the circuit input, and the
 * float weights / biases of a Dense layer as the loader stored them. */
int ref_plain_num_layers(void* hv) { return (int)static_cast<RefHandle*>(hv)->layers.size(); }
int ref_plain_layer_type(void* hv, int layer) { return (int)static_cast<RefHandle*>(hv)->layers.at(layer)->get_type(); }
int ref_plain_layer_dims(void* hv, int layer, int output, uint64_t* dims, int cap) {
    auto* l = static_cast<RefHandle*>(hv)->layers.at(layer);
    dim_t d = output ? l->get_output_dims() : l->get_input_dims();
    for (int i = 0; i < cap && i < (int)d.size(); ++i) dims[i] = d[i];
    return (int)d.size();
}
/* which: 0 weights, 1 biases; returns the element count (copies at most cap floats) */
long ref_plain_dense_floats(void* hv, int layer, int which, float* out, long cap) {
    auto* d = dynamic_cast<Dense*>(static_cast<RefHandle*>(hv)->layers.at(layer));
    if (!d) return -1;
    auto& t = which ? d->get_biases() : d->get_weights();
    const long n = (long)t.size();
    for (long i = 0; i < n && i < cap; ++i) out[i] = (float)t.data()[i];
    return n;
}

int ref_add_flatten(void* hv, uint64_t d0, uint64_t d1, uint64_t d2) {
    auto* h = static_cast<RefHandle*>(hv);
    h->layers.push_back(new Flatten(dim_t{d0, d1, d2}));
    return 0;
}

/* weights float [out][in] row-major, biases float [out] */
int ref_add_dense(void* hv, uint64_t out, uint64_t in, const float* w, const float* b, int q_method,
                  float q_const, int q_param, int channel_tf) {
    auto* h = static_cast<RefHandle*>(hv);
    std::vector<wandb_t> wv(w, w + out * in), bv(b, b + out);
    ScalarTensor<wandb_t> wt(wv, dim_t{out, in});
    ScalarTensor<wandb_t> bt(bv, dim_t{out});
    h->layers.push_back(new Dense(wt, bt, q_param, qm(q_method), q_const, channel_tf));
    return 0;
}

/* weights float [filter][channel][fh][fw] (PyTorch OIHW flat order, as the
 * ONNX loader passes them, circuit/onnx_modelloader.h:316-322) */
int ref_add_conv2d(void* hv, uint64_t in_w, uint64_t in_h, uint64_t ch, uint64_t filt, uint64_t fw,
                   uint64_t fh, uint64_t sw, uint64_t sh, const float* w, const float* b, int q_method,
                   float q_const, int q_param) {
    auto* h = static_cast<RefHandle*>(hv);
    std::vector<wandb_t> wv(w, w + filt * ch * fw * fh), bv(b, b + filt);
    ScalarTensor<wandb_t> wt(wv, dim_t{fw, fh, ch, filt});
    ScalarTensor<wandb_t> bt(bv, dim_t{filt});
    h->layers.push_back(new Conv2d(wt, bt, in_w, in_h, ch, filt, fw, fh, sw, sh, q_param, qm(q_method), q_const));
    return 0;
}

static dim_t dims3(uint64_t d0, uint64_t d1, uint64_t d2) {
    if (d1 == 0) return dim_t{d0};
    return dim_t{d0, d1, d2};
}

int ref_add_relu(void* hv, uint64_t d0, uint64_t d1, uint64_t d2) {
    static_cast<RefHandle*>(hv)->layers.push_back(new Relu(dims3(d0, d1, d2)));
    return 0;
}
int ref_add_sign(void* hv, uint64_t d0, uint64_t d1, uint64_t d2) {
    static_cast<RefHandle*>(hv)->layers.push_back(new Sign(dims3(d0, d1, d2)));
    return 0;
}
int ref_add_rescale_legacy(void* hv, int l, uint64_t d0, uint64_t d1, uint64_t d2) {
    static_cast<RefHandle*>(hv)->layers.push_back(new Rescale(l, dims3(d0, d1, d2)));
    return 0;
}
int ref_add_rescale(void* hv, const int* factors, int nf, uint64_t d0, uint64_t d1, uint64_t d2) {
    std::vector<crt_val_t> s(factors, factors + nf);
    static_cast<RefHandle*>(hv)->layers.push_back(new Rescale(s, dims3(d0, d1, d2)));
    return 0;
}
/* BaseExtension(input_dims, extra_moduli) (circuit/layer/base_extension.h:20) */
int ref_add_base_extension(void* hv, const int* extra, int nx, uint64_t d0, uint64_t d1, uint64_t d2) {
    std::vector<crt_val_t> e(extra, extra + nx);
    static_cast<RefHandle*>(hv)->layers.push_back(new BaseExtension(dims3(d0, d1, d2), e));
    return 0;
}
/* MaxPool2d(input_width, input_height, channel, kernel_width, kernel_height, stride_width, stride_height)
 * (circuit/layer/max_pool2d.h:24-37) */
int ref_add_maxpool2d(void* hv, uint64_t w, uint64_t h_, uint64_t c, uint64_t kw, uint64_t kh, uint64_t sw, uint64_t sh) {
    static_cast<RefHandle*>(hv)->layers.push_back(new MaxPool2d(w, h_, c, kw, kh, sw, sh));
    return 0;
}
/* func: 0 identity, 1 x+1, 2 2x, 3 x/2, 4 x/2-10, 5 x/4 (cf. dash/test/test_projection.h:14-16) */
int ref_add_projection(void* hv, uint64_t n, const int* in_mod, const int* out_mod, int k, int func) {
    std::vector<crt_val_t> im(in_mod, in_mod + k), om(out_mod, out_mod + k);
    crt_val_t (*f)(crt_val_t, void*) =
        func == 0   ? proj_identity
        : func == 1 ? proj_plus_one
        : func == 2 ? proj_times_two
        : func == 3 ? proj_half
        : func == 4 ? proj_half_minus_ten
                    : proj_quarter;
    static_cast<RefHandle*>(hv)->layers.push_back(new Projection(dim_t{n}, im, om, f));
    return 0;
}

/* Garbles with GarbledCircuit(circuit, crt_base, [mrs_base,] max_modulus)
 * (garbling/garbled_circuit.h:85-109), single-threaded so that the seeded
 * random stream is consumed in a fixed order. */
int ref_garble(void* hv, const int* crt_base, int k, const int* mrs_base, int n_mrs, int max_modulus,
               uint64_t seed) {
    auto* h = static_cast<RefHandle*>(hv);
    try {
        if (h->gc) {
            delete h->gc;
            h->gc = nullptr;
        }
        if (!h->circuit) h->circuit = new Circuit(h->layers);
        std::vector<crt_val_t> crt(crt_base, crt_base + k);
        std::vector<mrs_val_t> mrs(mrs_base, mrs_base + n_mrs);
        g_rng.seed(seed);
        int prev = omp_get_max_threads();
        omp_set_num_threads(1);
        if (n_mrs > 0)
            h->gc = new GarbledCircuit(h->circuit, crt, mrs, (crt_val_t)max_modulus);
        else
            h->gc = new GarbledCircuit(h->circuit, crt, (crt_val_t)max_modulus);
        omp_set_num_threads(prev);
    } catch (const std::exception& e) {
        h->err = e.what();
        return -1;
    } catch (const char* e) {
        h->err = e;
        return -1;
    }
    return 0;
}

uint64_t ref_input_size(void* hv) { return static_cast<RefHandle*>(hv)->circuit->get_input_size(); }
uint64_t ref_output_size(void* hv) { return static_cast<RefHandle*>(hv)->circuit->get_output_size(); }
int ref_k(void* hv) { return (int)static_cast<RefHandle*>(hv)->gc->get_crt_base().size(); }
int ref_num_garbled_layers(void* hv) { return (int)static_cast<RefHandle*>(hv)->gc->get_garbled_layer().size(); }
int ref_max_modulus(void* hv) { return static_cast<RefHandle*>(hv)->gc->get_max_modulus(); }
int ref_nr_comps(int q) { return (int)LabelTensor::get_nr_comps(q); }

/* Modulus j / label count of garbled layer `layer`'s output (-1 = last). */
int ref_layer_out_modulus(void* hv, int layer, int j) {
    auto* h = static_cast<RefHandle*>(hv);
    auto& gl = h->gc->get_garbled_layer();
    if (layer < 0) layer = (int)gl.size() - 1;
    return gl.at(layer)->get_out_label()->at(j)->get_modulus();
}
uint64_t ref_layer_out_size(void* hv, int layer) {
    auto* h = static_cast<RefHandle*>(hv);
    auto& gl = h->gc->get_garbled_layer();
    if (layer < 0) layer = (int)gl.size() - 1;
    return gl.at(layer)->get_layer()->get_output_size();
}

int ref_export(void* hv, const char* path) {
    auto* h = static_cast<RefHandle*>(hv);
    try {
        DashMaterialView v;
        dash_view_from_circuit(h->gc, v);
        dash_view_add_input_encoding(h->gc, v);
        return dash_view_save(v, path);
    } catch (const std::exception& e) {
        h->err = e.what();
        return -1;
    }
}

/* garble_inputs (garbled_circuit_interface.h:325-343): out[j] receives
 * int16 [input_size][nc(p_j)]. */
int ref_garble_inputs(void* hv, const int64_t* x, int16_t* const* out) {
    auto* h = static_cast<RefHandle*>(hv);
    auto dims = h->circuit->get_input_dims();
    std::vector<q_val_t> xv(x, x + h->circuit->get_input_size());
    ScalarTensor<q_val_t> in(xv, dims);
    auto* g = h->gc->garble_inputs(in);
    for (size_t j = 0; j < g->size(); ++j)
        memcpy(out[j], g->at(j)->get_components(), g->at(j)->size() * sizeof(crt_val_t));
    free_labels(g);
    return 0;
}

/* The reference's own evaluator, one inference.  in[j] = int16
 * [input_size][nc(p_j)], out[j] = int16 [output_size][nc(q_j)].
 * Returns the wall time of gc->cpu_evaluate alone in *ms (may be NULL). */
int ref_cpu_evaluate(void* hv, const int16_t* const* in, int nr_threads, int16_t* const* out, double* ms) {
    auto* h = static_cast<RefHandle*>(hv);
    try {
        auto* g_in = labels_from_raw(h->gc, h->gc->get_crt_base(), in, h->circuit->get_input_dims());
        auto t1 = std::chrono::high_resolution_clock::now();
        auto* g_out = h->gc->cpu_evaluate(g_in, nr_threads);
        auto t2 = std::chrono::high_resolution_clock::now();
        if (ms) *ms = std::chrono::duration<double, std::milli>(t2 - t1).count();
        if (out)
            for (size_t j = 0; j < g_out->size(); ++j)
                memcpy(out[j], g_out->at(j)->get_components(), g_out->at(j)->size() * sizeof(crt_val_t));
        // with zero garbled layers cpu_evaluate returns its argument
        free_labels(g_in);
    } catch (const std::exception& e) {
        h->err = e.what();
        return -1;
    } catch (const char* e) {
        h->err = e;
        return -1;
    }
    return 0;
}

/* Same walk as GarbledCircuitInterface::cpu_evaluate
 * (garbled_circuit_interface.h:345-380) but copying the output labels of
 * garbled layer `layer` as well: out_layer[j] = int16 [size][nc]. */
int ref_cpu_evaluate_upto(void* hv, const int16_t* const* in, int nr_threads, int layer,
                          int16_t* const* out_layer) {
    auto* h = static_cast<RefHandle*>(hv);
    try {
        auto* g_in = labels_from_raw(h->gc, h->gc->get_crt_base(), in, h->circuit->get_input_dims());
        auto* cur = g_in;
        auto lay = h->circuit->get_layer();
        auto& gl = h->gc->get_garbled_layer();
        for (size_t i = 0; i < lay.size(); ++i) {
            if (lay.at(i)->get_type() == Layer::flatten)
                for (auto* l : *cur) l->flatten();
            if (i == gl.size()) break;
            cur = gl.at(i)->cpu_evaluate(cur, nr_threads);
            if ((int)i == layer) {
                for (size_t j = 0; j < cur->size(); ++j)
                    memcpy(out_layer[j], cur->at(j)->get_components(), cur->at(j)->size() * sizeof(crt_val_t));
                break;
            }
        }
        free_labels(g_in);
    } catch (const std::exception& e) {
        h->err = e.what();
        return -1;
    } catch (const char* e) {
        h->err = e;
        return -1;
    }
    return 0;
}

/* decode_outputs (garbled_circuit_interface.h:422-467); labels[j] = int16
 * [output_size][nc(p_j)].  Returns -1 when a residue does not decode. */
int ref_decode_outputs(void* hv, const int16_t* const* labels, int64_t* values) {
    auto* h = static_cast<RefHandle*>(hv);
    auto* g = labels_from_raw(h->gc, h->gc->get_crt_base(), labels, h->circuit->get_output_dims());
    int rc = 0;
    // the reference only reports a failed match through assert(); re-check
    // here so that a release (-DNDEBUG) build still reports it
    for (size_t j = 0; j < g->size() && !rc; ++j) {
        LabelTensor dec{*g->at(j)};
        dec.compress();
        dec.hash();
        for (size_t l = 0; l < h->circuit->get_output_size(); ++l) {
            bool found = false;
            for (int r = 0; r < h->gc->get_crt_base().at(j); ++r)
                if (dec.get_hashed()[l] == h->gc->get_decoding_information().at(j).at(r).data()[l]) found = true;
            if (!found) rc = -1;
        }
    }
    if (!rc) {
        auto res = h->gc->decode_outputs(g);
        for (size_t i = 0; i < res.size(); ++i) values[i] = res.data()[i];
    }
    free_labels(g);
    return rc;
}

/* Circuit::plain_q_eval: the plaintext quantised semantics. */
int ref_plain_q_eval(void* hv, const int64_t* x, int64_t* y) {
    auto* h = static_cast<RefHandle*>(hv);
    std::vector<q_val_t> xv(x, x + h->circuit->get_input_size());
    ScalarTensor<q_val_t> in(xv, h->circuit->get_input_dims());
    auto out = h->circuit->plain_q_eval(in);
    for (size_t i = 0; i < out.size(); ++i) y[i] = out.data()[i];
    return 0;
}

/* Single-label primitives for pinning the restatement:
 * LabelTensor::compress / hash / decompress (label_tensor.h:715-742,775-780). */
void ref_compress_hash(int modulus, const int16_t* digits, uint64_t* compressed2, uint64_t* hashed2) {
    size_t nc = LabelTensor::get_nr_comps(modulus);
    LabelTensor l((crt_val_t)modulus, const_cast<crt_val_t*>(digits), nc, dim_t{1});
    l.compress();
    l.hash();
    memcpy(compressed2, l.get_compressed(), 16);
    memcpy(hashed2, l.get_hashed(), 16);
}
void ref_decompress(int modulus, const uint64_t* compressed2, int16_t* digits) {
    LabelTensor l((crt_val_t)modulus, dim_t{1});
    __uint128_t x;
    memcpy(&x, compressed2, 16);
    l.set_compressed(x, 0);
    l.decompress();
    memcpy(digits, l.get_components(), l.size() * sizeof(crt_val_t));
}

}  // extern "C"
