// See netdef.h.  Host-only C++ (no CUDA).
#include "netdef.h"

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>
#include <stdexcept>

namespace fcnod {

// ------------------------------------------------------------------ prototxt (protobuf text format)
namespace {

struct Tok {
    enum Kind { ATOM, STRING, LBRACE, RBRACE, COLON, END } kind;
    std::string text;
};

class Lexer {
public:
    explicit Lexer(const std::string& s) : s_(s), i_(0) {}
    Tok next() {
        for (;;) {
            while (i_ < s_.size() && isspace((unsigned char)s_[i_])) ++i_;
            if (i_ < s_.size() && s_[i_] == '#') {
                while (i_ < s_.size() && s_[i_] != '\n') ++i_;
                continue;
            }
            break;
        }
        Tok t;
        if (i_ >= s_.size()) { t.kind = Tok::END; return t; }
        char c = s_[i_];
        if (c == '{') { ++i_; t.kind = Tok::LBRACE; return t; }
        if (c == '}') { ++i_; t.kind = Tok::RBRACE; return t; }
        if (c == ':') { ++i_; t.kind = Tok::COLON; return t; }
        if (c == '"' || c == '\'') {
            char q = c;
            ++i_;
            t.kind = Tok::STRING;
            while (i_ < s_.size() && s_[i_] != q) {
                if (s_[i_] == '\\' && i_ + 1 < s_.size()) ++i_;
                t.text.push_back(s_[i_++]);
            }
            if (i_ >= s_.size()) throw std::runtime_error("prototxt: unterminated string");
            ++i_;
            return t;
        }
        t.kind = Tok::ATOM;
        while (i_ < s_.size() && !isspace((unsigned char)s_[i_]) && s_[i_] != '{' && s_[i_] != '}' &&
               s_[i_] != ':' && s_[i_] != '#' && s_[i_] != '"' && s_[i_] != '\'')
            t.text.push_back(s_[i_++]);
        return t;
    }

private:
    const std::string& s_;
    size_t i_;
};

struct Msg {
    // field name -> scalars (as text) / sub-messages, both in order of appearance
    std::multimap<std::string, std::string> scalars;
    std::vector<std::pair<std::string, Msg> > subs;
    std::vector<std::pair<std::string, std::string> > ordered;  // scalars in file order

    std::vector<std::string> all(const std::string& k) const {
        std::vector<std::string> out;
        for (size_t i = 0; i < ordered.size(); ++i)
            if (ordered[i].first == k) out.push_back(ordered[i].second);
        return out;
    }
    bool has(const std::string& k) const { return scalars.find(k) != scalars.end(); }
    std::string str(const std::string& k, const std::string& d = "") const {
        std::multimap<std::string, std::string>::const_iterator it = scalars.find(k);
        return it == scalars.end() ? d : it->second;
    }
    double num(const std::string& k, double d) const {
        std::multimap<std::string, std::string>::const_iterator it = scalars.find(k);
        if (it == scalars.end()) return d;
        char* end = 0;
        double v = strtod(it->second.c_str(), &end);
        if (end == it->second.c_str()) throw std::runtime_error("prototxt: field '" + k + "' is not a number");
        return v;
    }
    const Msg* sub(const std::string& k) const {
        for (size_t i = 0; i < subs.size(); ++i)
            if (subs[i].first == k) return &subs[i].second;
        return 0;
    }
};

Msg parse_msg(Lexer& lx, bool top) {
    Msg m;
    for (;;) {
        Tok t = lx.next();
        if (t.kind == Tok::END) {
            if (!top) throw std::runtime_error("prototxt: missing '}'");
            return m;
        }
        if (t.kind == Tok::RBRACE) {
            if (top) throw std::runtime_error("prototxt: unbalanced '}'");
            return m;
        }
        if (t.kind != Tok::ATOM) throw std::runtime_error("prototxt: expected a field name");
        std::string name = t.text;
        Tok v = lx.next();
        if (v.kind == Tok::COLON) v = lx.next();
        if (v.kind == Tok::LBRACE) {
            m.subs.push_back(std::make_pair(name, parse_msg(lx, false)));
        } else if (v.kind == Tok::ATOM || v.kind == Tok::STRING) {
            m.scalars.insert(std::make_pair(name, v.text));
            m.ordered.push_back(std::make_pair(name, v.text));
        } else {
            throw std::runtime_error("prototxt: bad value for field '" + name + "'");
        }
    }
}

LayerType type_of(const std::string& s) {
    if (s == "Convolution") return L_CONV;
    if (s == "ReLU") return L_RELU;
    if (s == "Pooling") return L_POOL;
    if (s == "LRN") return L_LRN;
    if (s == "Concat") return L_CONCAT;
    if (s == "Power") return L_POWER;
    if (s == "Dropout") return L_DROPOUT;
    if (s == "Sigmoid") return L_SIGMOID;
    return L_UNKNOWN;
}

}  // namespace

NetDef parse_prototxt_text(const std::string& text, bool train) {
    Lexer lx(text);
    Msg root = parse_msg(lx, true);
    NetDef net;
    net.input_name = root.str("input");
    std::vector<std::string> dims;
    if (const Msg* shp = root.sub("input_shape")) dims = shp->all("dim");
    else dims = root.all("input_dim");
    if (net.input_name.empty() || dims.size() != 4) {
        if (!train) throw std::runtime_error("prototxt: need `input:` and a 4-d input_shape / input_dim");
    } else {
        for (int i = 0; i < 4; ++i) net.input_dims[i] = atoi(dims[i].c_str());
    }
    for (size_t i = 0; i < root.subs.size(); ++i) {
        if (root.subs[i].first != "layer") continue;
        const Msg& L = root.subs[i].second;
        LayerDef d;
        d.name = L.str("name");
        d.type_name = L.str("type");
        d.type = type_of(d.type_name);
        d.bottom = L.all("bottom");
        d.top = L.all("top");
        if (train && (d.type_name == "Data" || d.type_name == "Slice" || d.type_name == "Eltwise" || d.type_name == "L1Loss" ||
                      d.type_name == "EuclideanLoss")) {
            TrainSideLayer t;
            t.name = d.name; t.type_name = d.type_name; t.bottom = d.bottom; t.top = d.top;
            if (const Msg* sp = L.sub("slice_param")) {
                std::vector<std::string> pts = sp->all("slice_point");
                for (size_t k = 0; k < pts.size(); ++k) t.slice_points.push_back(atoi(pts[k].c_str()));
                if ((int)sp->num("slice_dim", sp->num("axis", 1)) != 1)
                    throw std::runtime_error("prototxt: '" + d.name + "': only channel slicing is supported");
            }
            if (const Msg* ep = L.sub("eltwise_param")) t.eltwise_op = ep->str("operation", "SUM");
            t.loss_weight = (float)L.num("loss_weight", 1.0);
            net.train_side.push_back(t);
            continue;
        }
        if (d.type == L_UNKNOWN)
            throw std::runtime_error("prototxt: layer '" + d.name + "' has unsupported type '" + d.type_name + "'");
        if (d.top.size() != 1) throw std::runtime_error("prototxt: layer '" + d.name + "' must have exactly one top");
        if (d.type == L_CONV) {
            const Msg* p = L.sub("convolution_param");
            if (!p) throw std::runtime_error("prototxt: '" + d.name + "' lacks convolution_param");
            d.num_output = (int)p->num("num_output", 0);
            d.kernel = (int)p->num("kernel_size", 1);
            d.stride = (int)p->num("stride", 1);
            d.pad = (int)p->num("pad", 0);
            d.bias_term = p->str("bias_term", "true") != "false";
            {
                int np = 0;
                for (size_t k = 0; k < L.subs.size() && np < 2; ++k)
                    if (L.subs[k].first == "param") {
                        d.lr_mult[np] = (float)L.subs[k].second.num("lr_mult", 1.0);
                        d.decay_mult[np] = (float)L.subs[k].second.num("decay_mult", 1.0);
                        ++np;
                    }
            }
            if (d.num_output <= 0) throw std::runtime_error("prototxt: '" + d.name + "' num_output missing");
            // square kernels, one stride / pad, dense, ungrouped -- everything models/deploy.prototxt uses.  Anything else
            // would silently compute a different convolution: refuse it.
            static const char* const unsupported[] = {"kernel_h", "kernel_w", "pad_h", "pad_w", "stride_h", "stride_w"};
            for (size_t u = 0; u < sizeof(unsupported) / sizeof(unsupported[0]); ++u)
                if (p->has(unsupported[u]))
                    throw std::runtime_error("prototxt: '" + d.name + "': convolution_param." + unsupported[u] + " is not supported (square kernels only)");
            if (p->all("kernel_size").size() > 1 || p->all("stride").size() > 1 || p->all("pad").size() > 1)
                throw std::runtime_error("prototxt: '" + d.name + "': per-axis kernel_size / stride / pad lists are not supported");
            if ((int)p->num("dilation", 1) != 1) throw std::runtime_error("prototxt: '" + d.name + "': dilation != 1 is not supported");
            if ((int)p->num("group", 1) != 1) throw std::runtime_error("prototxt: '" + d.name + "': group != 1 is not supported");
            if ((int)p->num("axis", 1) != 1) throw std::runtime_error("prototxt: '" + d.name + "': convolution axis != 1 is not supported");
            if (d.kernel <= 0 || d.stride <= 0 || d.pad < 0) throw std::runtime_error("prototxt: '" + d.name + "': bad kernel_size / stride / pad");
        } else if (d.type == L_POOL) {
            const Msg* p = L.sub("pooling_param");
            if (!p) throw std::runtime_error("prototxt: '" + d.name + "' lacks pooling_param");
            d.pool_max = p->str("pool", "MAX") == "MAX";
            d.kernel = (int)p->num("kernel_size", 1);
            d.stride = (int)p->num("stride", 1);
            d.pad = (int)p->num("pad", 0);
            if (!d.pool_max) throw std::runtime_error("prototxt: '" + d.name + "': only MAX pooling is supported");
            static const char* const unsupported[] = {"kernel_h", "kernel_w", "pad_h", "pad_w", "stride_h", "stride_w"};
            for (size_t u = 0; u < sizeof(unsupported) / sizeof(unsupported[0]); ++u)
                if (p->has(unsupported[u]))
                    throw std::runtime_error("prototxt: '" + d.name + "': pooling_param." + unsupported[u] + " is not supported (square windows only)");
            if (p->str("global_pooling", "false") == "true") throw std::runtime_error("prototxt: '" + d.name + "': global_pooling is not supported");
            if (d.kernel <= 0 || d.stride <= 0 || d.pad < 0) throw std::runtime_error("prototxt: '" + d.name + "': bad kernel_size / stride / pad");
        } else if (d.type == L_LRN) {
            const Msg* p = L.sub("lrn_param");
            if (p) {
                d.local_size = (int)p->num("local_size", 5);
                d.alpha = (float)p->num("alpha", 1.0);
                d.beta = (float)p->num("beta", 0.75);
                d.lrn_k = (float)p->num("k", 1.0);
                if (p->str("norm_region", "ACROSS_CHANNELS") != "ACROSS_CHANNELS")
                    throw std::runtime_error("prototxt: '" + d.name + "': only ACROSS_CHANNELS LRN is supported");
            }
        } else if (d.type == L_DROPOUT) {
            if (const Msg* p = L.sub("dropout_param")) d.dropout_ratio = (float)p->num("dropout_ratio", 0.5);
        } else if (d.type == L_POWER) {
            const Msg* p = L.sub("power_param");
            if (p) {
                d.power = (float)p->num("power", 1.0);
                d.scale = (float)p->num("scale", 1.0);
                d.shift = (float)p->num("shift", 0.0);
            }
        }
        net.layers.push_back(d);
    }
    if (net.input_name.empty()) {
        // train form: the network input is the bottom of the first body layer (`data`, train_val.prototxt:27-35)
        if (net.layers.empty() || net.layers[0].bottom.size() != 1) throw std::runtime_error("prototxt: no body layers");
        net.input_name = net.layers[0].bottom[0];
    }
    return net;
}

NetDef parse_prototxt_file(const std::string& path, bool train) {
    std::ifstream f(path.c_str());
    if (!f) throw std::runtime_error("cannot open prototxt '" + path + "'");
    std::stringstream ss;
    ss << f.rdbuf();
    return parse_prototxt_text(ss.str(), train);
}

// ------------------------------------------------------------------ caffemodel (protobuf wire format)
// caffe.proto field numbers: NetParameter{layers(V1)=2, layer=100};
// LayerParameter{name=1, blobs=7}; V1LayerParameter{name=4, blobs=6};
// BlobProto{num=1, channels=2, height=3, width=4, data=5, shape=7, double_data=8};
// BlobShape{dim=1}.
// NVCaffe (the fork the reference trained with: train/train.sh:19) stores weights, unless store_blobs_in_old_format is
// set, as BlobProto{raw_data_type=10 (enum Type: DOUBLE=0, FLOAT=1, FLOAT16=2), raw_data=12 (bytes)} [ext: field and
// enum numbers from NVCaffe 0.16/0.17's caffe.proto, not vendored upstream and not verifiable here].  The element width
// is taken from len(raw_data) / count (4, 2 or 8 bytes: float32, IEEE half, float64), which does not depend on the enum
// numbering; a declared type that contradicts it is an error.
// Datum{channels=1, height=2, width=3, data=4 (bytes), label=5, float_data=6, encoded=7}.
namespace {

struct Reader {
    const uint8_t* p;
    const uint8_t* end;
    bool done() const { return p >= end; }
    uint64_t varint() {
        uint64_t v = 0;
        int shift = 0;
        for (;;) {
            if (p >= end) throw std::runtime_error("caffemodel: truncated varint");
            uint8_t b = *p++;
            v |= (uint64_t)(b & 0x7F) << shift;
            if (!(b & 0x80)) return v;
            shift += 7;
            if (shift > 63) throw std::runtime_error("caffemodel: varint too long");
        }
    }
    // returns field number; sets wire type and, for length-delimited, the sub-range
    int field(int* wt, Reader* sub, uint64_t* val) {
        uint64_t key = varint();
        *wt = (int)(key & 7);
        switch (*wt) {
            case 0: *val = varint(); break;
            case 1:
                if (end - p < 8) throw std::runtime_error("caffemodel: truncated fixed64");
                memcpy(val, p, 8); p += 8; break;
            case 5: {
                if (end - p < 4) throw std::runtime_error("caffemodel: truncated fixed32");
                uint32_t v32; memcpy(&v32, p, 4); *val = v32; p += 4; break;
            }
            case 2: {
                uint64_t n = varint();
                if ((uint64_t)(end - p) < n) throw std::runtime_error("caffemodel: truncated field");
                sub->p = p; sub->end = p + n; p += n; break;
            }
            default: throw std::runtime_error("caffemodel: unsupported wire type");
        }
        return (int)(key >> 3);
    }
};

float half_to_float(uint16_t h) {
    const uint32_t sign = (uint32_t)(h & 0x8000u) << 16;
    uint32_t exp = (h >> 10) & 0x1Fu, man = h & 0x3FFu, bits;
    if (exp == 0) {
        if (man == 0) bits = sign;
        else {                                  // subnormal: renormalise
            int e = -1;
            do { ++e; man <<= 1; } while (!(man & 0x400u));
            bits = sign | ((uint32_t)(127 - 15 - e) << 23) | ((man & 0x3FFu) << 13);
        }
    } else if (exp == 31) bits = sign | 0x7F800000u | (man << 13);
    else bits = sign | ((exp + 112u) << 23) | (man << 13);
    float f;
    memcpy(&f, &bits, 4);
    return f;
}

HostBlob read_blob(Reader r) {
    HostBlob b;
    int legacy[4] = {1, 1, 1, 1};
    bool has_legacy = false, has_shape = false;
    std::vector<double> dd;
    const uint8_t* raw = 0;
    size_t raw_len = 0;
    int raw_type = -1;
    while (!r.done()) {
        int wt; Reader s; uint64_t v = 0;
        int f = r.field(&wt, &s, &v);
        if (f >= 1 && f <= 4 && wt == 0) { legacy[f - 1] = (int)v; has_legacy = true; }
        else if (f == 5 && wt == 2) {
            size_t n = (size_t)(s.end - s.p) / 4, old = b.data.size();
            b.data.resize(old + n);
            memcpy(b.data.data() + old, s.p, n * 4);
        } else if (f == 5 && wt == 5) {
            uint32_t u = (uint32_t)v; float x; memcpy(&x, &u, 4); b.data.push_back(x);
        } else if (f == 8 && wt == 2) {
            size_t n = (size_t)(s.end - s.p) / 8;
            for (size_t i = 0; i < n; ++i) { double x; memcpy(&x, s.p + 8 * i, 8); dd.push_back(x); }
        } else if (f == 10 && wt == 0) {
            raw_type = (int)v;
        } else if (f == 12 && wt == 2) {
            raw = s.p; raw_len = (size_t)(s.end - s.p);
        } else if (f == 7 && wt == 2) {
            has_shape = true;
            while (!s.done()) {
                int wt2; Reader s2; uint64_t v2 = 0;
                int f2 = s.field(&wt2, &s2, &v2);
                if (f2 != 1) continue;
                if (wt2 == 2) while (!s2.done()) b.shape.push_back((int)s2.varint());
                else b.shape.push_back((int)v2);
            }
        }
    }
    if (b.data.empty() && !dd.empty()) { b.data.resize(dd.size()); for (size_t i = 0; i < dd.size(); ++i) b.data[i] = (float)dd[i]; }
    if (!has_shape) { (void)has_legacy; b.shape.assign(legacy, legacy + 4); }
    size_t n = 1;
    for (size_t i = 0; i < b.shape.size(); ++i) n *= (size_t)b.shape[i];
    if (b.data.empty() && raw && n > 0) {       // NVCaffe raw_data
        if (raw_len % n) throw std::runtime_error("caffemodel: raw_data length is not a multiple of the blob's element count");
        const size_t es = raw_len / n;
        const int expect = es == 8 ? 0 : (es == 4 ? 1 : (es == 2 ? 2 : -2));
        if (expect == -2) throw std::runtime_error("caffemodel: raw_data with " + std::to_string(es) + "-byte elements is not supported");
        if (raw_type >= 0 && raw_type != expect)
            throw std::runtime_error("caffemodel: raw_data_type " + std::to_string(raw_type) + " contradicts the element width of raw_data (" +
                                     std::to_string(es) + " bytes)");
        b.data.resize(n);
        for (size_t i = 0; i < n; ++i) {
            if (es == 4) memcpy(&b.data[i], raw + 4 * i, 4);
            else if (es == 8) { double x; memcpy(&x, raw + 8 * i, 8); b.data[i] = (float)x; }
            else { uint16_t h; memcpy(&h, raw + 2 * i, 2); b.data[i] = half_to_float(h); }
        }
    }
    if (n != b.data.size()) throw std::runtime_error("caffemodel: blob shape does not match its data length");
    return b;
}

}  // namespace

Datum decode_datum(const uint8_t* buf, size_t len) {
    Datum d;
    Reader r; r.p = buf; r.end = buf + len;
    while (!r.done()) {
        int wt; Reader s; uint64_t v = 0;
        const int f = r.field(&wt, &s, &v);
        if (f == 1 && wt == 0) d.channels = (int)v;
        else if (f == 2 && wt == 0) d.height = (int)v;
        else if (f == 3 && wt == 0) d.width = (int)v;
        else if (f == 4 && wt == 2) d.bytes.assign(s.p, s.end);
        else if (f == 5 && wt == 0) { d.label = (int)v; d.has_label = true; }
        else if (f == 6 && wt == 5) { uint32_t u = (uint32_t)v; float x; memcpy(&x, &u, 4); d.floats.push_back(x); }   // unpacked (proto2 default)
        else if (f == 6 && wt == 2) {                                                                              // packed
            const size_t n = (size_t)(s.end - s.p) / 4, old = d.floats.size();
            d.floats.resize(old + n);
            memcpy(d.floats.data() + old, s.p, n * 4);
        } else if (f == 7 && wt == 0) d.encoded = v != 0;
    }
    if (d.encoded) throw std::runtime_error("Datum: encoded (compressed) images are not supported");
    const size_t n = (size_t)d.channels * d.height * d.width;
    if (d.channels <= 0 || d.height <= 0 || d.width <= 0) throw std::runtime_error("Datum: missing channels / height / width");
    if (!d.bytes.empty() && !d.floats.empty()) throw std::runtime_error("Datum: both data and float_data are set");
    if ((d.bytes.empty() ? d.floats.size() : d.bytes.size()) != n) throw std::runtime_error("Datum: data length does not match channels x height x width");
    return d;
}

WeightMap read_caffemodel(const std::string& path) {
    std::ifstream f(path.c_str(), std::ios::binary);
    if (!f) throw std::runtime_error("cannot open caffemodel '" + path + "'");
    std::vector<uint8_t> buf((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
    Reader r; r.p = buf.data(); r.end = buf.data() + buf.size();
    WeightMap out;
    while (!r.done()) {
        int wt; Reader s; uint64_t v = 0;
        int f0 = r.field(&wt, &s, &v);
        if (wt != 2 || (f0 != 100 && f0 != 2)) continue;
        const int name_field = f0 == 100 ? 1 : 4, blob_field = f0 == 100 ? 7 : 6;
        std::string name;
        std::vector<HostBlob> blobs;
        while (!s.done()) {
            int wt2; Reader s2; uint64_t v2 = 0;
            int f2 = s.field(&wt2, &s2, &v2);
            if (f2 == name_field && wt2 == 2) name.assign((const char*)s2.p, (size_t)(s2.end - s2.p));
            else if (f2 == blob_field && wt2 == 2) blobs.push_back(read_blob(s2));
        }
        if (!blobs.empty()) out[name] = blobs;
    }
    return out;
}

}  // namespace fcnod
