/*
 * dash_onnx_lite.h -- ONNX models into Dash circuits without protobuf
 * ("next" row f1).
 *
 * The reference's loader (circuit/onnx_modelloader.h:156-369) needs
 * libprotobuf and a generated onnx.proto3.pb.h that the repository does not
 * ship (README.md:37-44).  This header reads the handful of ONNX messages the
 * loader looks at straight from the protobuf wire format and then builds the
 * SAME reference layer objects (Dense, Conv2d, Relu, Sign, Flatten,
 * MaxPool2d, Rescale) with the reference's own constructors, so quantisation
 * is the reference's code, not a restatement.  Include it after the Dash
 * headers:
 *
 *     #include "circuit/circuit.h"            // Dash, unmodified
 *     #include "circuit/layer/dense.h" ...     // (the layer headers)
 *     #include "dash_onnx_lite.h"
 *     Circuit* c = dash_onnx::load_circuit("model.onnx", QuantizationMethod::SimpleQuant, 0.02f, -1);
 *
 * Loader semantics kept on purpose, quirks included: initializers are taken
 * in file order (pytorch) or reverse order (tf2onnx), two per Gemm/Conv;
 * Conv attributes are read by POSITION (kernel_shape = attribute 2, strides =
 * attribute 4 for pytorch / 1 for tf2onnx); a Reshape counts as Flatten when
 * its output name contains "flatten"; tf2onnx Gemm weights are transposed and
 * the first dense layer after a convolution gets the channel count for the
 * channel-first/-last permutation; ScaleQuant / ScaleQuantPlus append a
 * Rescale after every linear layer.
 */
#ifndef DASH_ONNX_LITE_H
#define DASH_ONNX_LITE_H

#include <cstdint>
#include <cstring>
#include <fstream>
#include <string>
#include <vector>

namespace dash_onnx {

/* ---- protobuf wire format, just enough of it ---- */
struct Reader {
    const uint8_t* p;
    const uint8_t* end;
    bool ok;
    Reader(const uint8_t* b, size_t n) : p(b), end(b + n), ok(true) {}
    bool more() const { return ok && p < end; }
    uint64_t varint() {
        uint64_t v = 0;
        for (int shift = 0; shift < 64 && p < end; shift += 7) {
            const uint8_t b = *p++;
            v |= (uint64_t)(b & 0x7f) << shift;
            if (!(b & 0x80)) return v;
        }
        ok = false;
        return 0;
    }
    /* next field: number and wire type; for length-delimited fields `sub` spans the payload */
    bool field(uint32_t& number, uint32_t& wire, uint64_t& value, Reader& sub) {
        if (!more()) return false;
        const uint64_t key = varint();
        number = (uint32_t)(key >> 3);
        wire = (uint32_t)(key & 7);
        value = 0;
        if (wire == 0) {
            value = varint();
        } else if (wire == 1) {
            if (end - p < 8) return ok = false;
            memcpy(&value, p, 8);
            p += 8;
        } else if (wire == 5) {
            if (end - p < 4) return ok = false;
            uint32_t v32;
            memcpy(&v32, p, 4);
            value = v32;
            p += 4;
        } else if (wire == 2) {
            const uint64_t n = varint();
            if (!ok || (uint64_t)(end - p) < n) return ok = false;
            sub = Reader(p, (size_t)n);
            p += n;
        } else {
            return ok = false;
        }
        return ok;
    }
    std::string str() const { return std::string((const char*)p, (size_t)(end - p)); }
};

/* repeated int64: packed (wire 2) or one element per field (wire 0) */
inline void add_ints(std::vector<int64_t>& v, uint32_t wire, uint64_t value, Reader sub) {
    if (wire == 0) {
        v.push_back((int64_t)value);
    } else if (wire == 2) {
        while (sub.more()) v.push_back((int64_t)sub.varint());
    }
}

struct Tensor { /* TensorProto: dims = 1, data_type = 2, float_data = 4, name = 8, raw_data = 9 */
    std::vector<int64_t> dims;
    std::vector<float> data;
    std::string name;
};
struct Attribute { /* AttributeProto: name = 1, ints = 8 */
    std::string name;
    std::vector<int64_t> ints;
};
struct Node { /* NodeProto: input = 1, output = 2, op_type = 4, attribute = 5 */
    std::string op_type;
    std::vector<std::string> inputs, outputs;
    std::vector<Attribute> attributes;
};
struct Model { /* ModelProto: producer_name = 2, graph = 7; GraphProto: node = 1, initializer = 5, input = 11 */
    std::string producer;
    std::vector<Node> nodes;
    std::vector<Tensor> initializers;
    std::vector<int64_t> input_dims; /* shape of graph.input(0) */
};

inline bool parse_tensor(Reader r, Tensor& t) {
    uint32_t f, w;
    uint64_t v;
    Reader sub(nullptr, 0);
    while (r.field(f, w, v, sub)) {
        if (f == 1) {
            add_ints(t.dims, w, v, sub);
        } else if (f == 4) { /* float_data, packed or single */
            if (w == 2) {
                const size_t n = (size_t)(sub.end - sub.p) / 4;
                const size_t at = t.data.size();
                t.data.resize(at + n);
                memcpy(t.data.data() + at, sub.p, n * 4);
            } else if (w == 5) {
                float x;
                const uint32_t bits = (uint32_t)v;
                memcpy(&x, &bits, 4);
                t.data.push_back(x);
            }
        } else if (f == 8 && w == 2) {
            t.name = sub.str();
        } else if (f == 9 && w == 2) { /* raw_data: little-endian float32 (parse_raw_data, onnx_modelloader.h:89-122) */
            const size_t n = (size_t)(sub.end - sub.p) / 4;
            t.data.resize(n);
            memcpy(t.data.data(), sub.p, n * 4);
        }
    }
    return r.ok;
}

inline bool parse_node(Reader r, Node& n) {
    uint32_t f, w;
    uint64_t v;
    Reader sub(nullptr, 0);
    while (r.field(f, w, v, sub)) {
        if (w != 2) continue;
        if (f == 1) n.inputs.push_back(sub.str());
        else if (f == 2) n.outputs.push_back(sub.str());
        else if (f == 4) n.op_type = sub.str();
        else if (f == 5) {
            Attribute a;
            Reader ar = sub, as(nullptr, 0);
            uint32_t af, aw;
            uint64_t av;
            while (ar.field(af, aw, av, as)) {
                if (af == 1 && aw == 2) a.name = as.str();
                else if (af == 8) add_ints(a.ints, aw, av, as);
            }
            if (!ar.ok) return false;
            n.attributes.push_back(a);
        }
    }
    return r.ok;
}

/* ValueInfoProto.type(2).tensor_type(1).shape(2).dim(1).dim_value(1) */
inline bool parse_input_dims(Reader r, std::vector<int64_t>& dims) {
    uint32_t f, w;
    uint64_t v;
    Reader type(nullptr, 0);
    while (r.field(f, w, v, type)) {
        if (f != 2 || w != 2) continue;
        Reader tt(nullptr, 0), tr = type;
        while (tr.field(f, w, v, tt)) {
            if (f != 1 || w != 2) continue;
            Reader sh(nullptr, 0), ttr = tt;
            while (ttr.field(f, w, v, sh)) {
                if (f != 2 || w != 2) continue;
                Reader dim(nullptr, 0), shr = sh;
                while (shr.field(f, w, v, dim)) {
                    if (f != 1 || w != 2) continue;
                    int64_t value = 0;
                    Reader x(nullptr, 0), dr = dim;
                    uint32_t df, dw;
                    uint64_t dv;
                    while (dr.field(df, dw, dv, x))
                        if (df == 1 && dw == 0) value = (int64_t)dv;
                    dims.push_back(value);
                }
            }
        }
    }
    return r.ok;
}

inline bool parse_model(const uint8_t* bytes, size_t n, Model& m) {
    Reader r(bytes, n), sub(nullptr, 0);
    uint32_t f, w;
    uint64_t v;
    while (r.field(f, w, v, sub)) {
        if (f == 2 && w == 2) {
            m.producer = sub.str();
        } else if (f == 7 && w == 2) {
            Reader g = sub, gs(nullptr, 0);
            bool first_input = true;
            while (g.field(f, w, v, gs)) {
                if (w != 2) continue;
                if (f == 1) {
                    m.nodes.emplace_back();
                    if (!parse_node(gs, m.nodes.back())) return false;
                } else if (f == 5) {
                    m.initializers.emplace_back();
                    if (!parse_tensor(gs, m.initializers.back())) return false;
                } else if (f == 11 && first_input) {
                    first_input = false;
                    if (!parse_input_dims(gs, m.input_dims)) return false;
                }
            }
            if (!g.ok) return false;
        }
    }
    return r.ok;
}

inline bool read_model(const std::string& path, Model& m) {
    std::ifstream in(path, std::ios::binary | std::ios::ate);
    if (!in) return false;
    const std::streamsize size = in.tellg();
    in.seekg(0, std::ios::beg);
    std::vector<uint8_t> buf((size_t)size);
    if (!in.read((char*)buf.data(), size)) return false;
    return parse_model(buf.data(), buf.size(), m);
}

}  // namespace dash_onnx

/* ---- the circuit: needs the Dash headers included first ---- */
#ifdef ONNX_LITE_BUILD_CIRCUIT
namespace dash_onnx {

/* create_circuit_from_onnx_model (circuit/onnx_modelloader.h:156-369) on the
 * parsed model; returns the layer list (empty on a malformed model). */
inline std::vector<Layer*> build_layers(const Model& m, QuantizationMethod q_method, wandb_t q_const, int q_parameter) {
    std::vector<Layer*> layers;
    const bool tf = m.producer == "tf2onnx";
    long cursor = 0; /* next initializer: counts up for pytorch, down for tf2onnx (:169-171) */
    if (tf) {
        for (const Node& n : m.nodes)
            if (n.op_type == "Gemm" || n.op_type == "Conv") cursor += 2;
        cursor -= 1;
    }
    const std::vector<int64_t>& d = m.input_dims;
    size_t channel, in_h, in_w;
    if (tf && d.size() == 3) {
        channel = (size_t)d[0], in_h = (size_t)d[1], in_w = (size_t)d[2];
    } else if (tf && d.size() == 4) {
        channel = (size_t)d[3], in_h = (size_t)d[1], in_w = (size_t)d[2];
    } else if (d.size() == 4) {
        channel = (size_t)d[1], in_h = (size_t)d[2], in_w = (size_t)d[3];
    } else {
        return layers;
    }
    dim_t next_dim{in_w, in_h, channel};
    bool first_dense = true;
    auto tensor_at = [&](long i) -> const Tensor* { return i >= 0 && i < (long)m.initializers.size() ? &m.initializers[(size_t)i] : nullptr; };
    auto fail = [&]() {
        for (Layer* l : layers) delete l;
        layers.clear();
        return layers;
    };
    auto after_linear = [&](const dim_t& out_dims) {
        if (q_method == QuantizationMethod::ScaleQuant) {
            layers.push_back(new Rescale(q_parameter, out_dims));
        } else if (q_method == QuantizationMethod::ScaleQuantPlus) {
            std::vector<crt_val_t> qs = {static_cast<crt_val_t>(q_parameter)};
            layers.push_back(new Rescale(qs, out_dims));
        }
    };
    for (const Node& node : m.nodes) {
        const std::string& op = node.op_type;
        if (op == "Flatten" || (op == "Reshape" && !node.outputs.empty() && node.outputs[0].find("flatten") != std::string::npos)) {
            layers.push_back(new Flatten(next_dim));
        } else if (op == "Gemm") {
            const Tensor *wt, *bt;
            size_t in_dim, out_dim;
            if (tf) { /* biases first, walking backwards; weights stored [in][out] */
                bt = tensor_at(cursor), wt = tensor_at(cursor - 1);
                if (!wt || !bt || wt->dims.size() < 2) return fail();
                in_dim = (size_t)wt->dims[0], out_dim = (size_t)wt->dims[1];
                cursor -= 2;
            } else {
                wt = tensor_at(cursor), bt = tensor_at(cursor + 1);
                if (!wt || !bt || wt->dims.size() < 2) return fail();
                in_dim = (size_t)wt->dims[1], out_dim = (size_t)wt->dims[0];
                cursor += 2;
            }
            if (wt->data.size() != in_dim * out_dim || bt->data.size() != out_dim) return fail();
            std::vector<wandb_t> wv(wt->data.begin(), wt->data.end()), bv(bt->data.begin(), bt->data.end());
            ScalarTensor<wandb_t> biases(bv, dim_t{out_dim});
            Dense* dense;
            if (tf) {
                ScalarTensor<wandb_t> weights(wv, dim_t{in_dim, out_dim});
                weights.transpose();
                /* channel count of the convolution in front of the flatten, first dense layer only (:225-236) */
                int ch = 0;
                for (Layer* l : layers) {
                    if (l->get_type() == Layer::LayerType::conv2d) ch = (int)((Conv2d*)l)->get_filter();
                    if (l->get_type() == Layer::LayerType::flatten) break;
                }
                bool has_flatten = false;
                for (Layer* l : layers) has_flatten = has_flatten || l->get_type() == Layer::LayerType::flatten;
                if (!has_flatten) ch = 0;
                if (ch > 0 && first_dense) {
                    first_dense = false;
                    dense = new Dense(weights, biases, q_parameter, q_method, q_const, ch);
                } else {
                    dense = new Dense(weights, biases, q_parameter, q_method, q_const);
                }
            } else {
                ScalarTensor<wandb_t> weights(wv, dim_t{out_dim, in_dim});
                dense = new Dense(weights, biases, q_parameter, q_method, q_const);
            }
            layers.push_back(dense);
            after_linear(dense->get_output_dims());
            next_dim = dense->get_output_dims();
        } else if (op == "Conv") {
            /* the loader's own naming: dims are {width, height, channel} but it reads height first (:274-277) */
            const size_t input_height = next_dim.at(0), input_width = next_dim.at(1), ch = next_dim.at(2);
            const size_t stride_attr = tf ? 1 : 4;
            if (node.attributes.size() <= 2 || node.attributes.size() <= stride_attr) return fail();
            const std::vector<int64_t>& ks = node.attributes[2].ints;
            const std::vector<int64_t>& st = node.attributes[stride_attr].ints;
            if (ks.size() < 2 || st.size() < 2) return fail();
            const size_t fh = (size_t)ks[0], fw = (size_t)ks[1], sh = (size_t)st[0], sw = (size_t)st[1];
            const Tensor *wt, *bt;
            if (tf) {
                bt = tensor_at(cursor), wt = tensor_at(cursor - 1);
            } else {
                wt = tensor_at(cursor), bt = tensor_at(cursor + 1);
            }
            /* the filter count comes from the tensor at the cursor (:299): the bias for tf2onnx, the weights for pytorch */
            const Tensor* at_cursor = tensor_at(cursor);
            if (!wt || !bt || !at_cursor || at_cursor->dims.empty()) return fail();
            const size_t filter = (size_t)at_cursor->dims[0];
            cursor += tf ? -2 : 2;
            if (wt->data.size() != fw * fh * ch * filter || bt->data.size() != filter) return fail();
            std::vector<wandb_t> wv(wt->data.begin(), wt->data.end()), bv(bt->data.begin(), bt->data.end());
            ScalarTensor<wandb_t> weights(wv, dim_t{fw, fh, ch, filter});
            ScalarTensor<wandb_t> biases(bv, dim_t{filter});
            auto conv = new Conv2d{weights, biases, input_width, input_height, ch, filter, fw, fh, sw, sh, q_parameter, q_method, q_const};
            layers.push_back(conv);
            after_linear(conv->get_output_dims());
            next_dim = conv->get_output_dims();
        } else if (op == "MaxPool") {
            if (node.attributes.size() <= 3 || node.attributes[1].ints.size() < 2 || node.attributes[3].ints.size() < 2) return fail();
            const size_t input_height = next_dim.at(0), input_width = next_dim.at(1), ch = next_dim.at(2);
            auto pool = new MaxPool2d(input_width, input_height, ch, (size_t)node.attributes[1].ints[1], (size_t)node.attributes[1].ints[0],
                                      (size_t)node.attributes[3].ints[1], (size_t)node.attributes[3].ints[0]);
            layers.push_back(pool);
            next_dim = pool->get_output_dims();
        } else if (op == "Relu") {
            layers.push_back(new Relu(next_dim));
        } else if (op == "Tanh" || op == "SignTanh") {
            layers.push_back(new Sign(next_dim));
        }
    }
    return layers;
}

inline Circuit* load_circuit(const std::string& path, QuantizationMethod q_method, wandb_t q_const, int q_parameter) {
    Model m;
    if (!read_model(path, m)) return nullptr;
    std::vector<Layer*> layers = build_layers(m, q_method, q_const, q_parameter);
    if (layers.empty()) return nullptr;
    return new Circuit(layers, q_parameter);
}

}  // namespace dash_onnx
#endif /* ONNX_LITE_BUILD_CIRCUIT */
#endif /* DASH_ONNX_LITE_H */
