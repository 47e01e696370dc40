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

NetDef parse_prototxt_text(const std::string& text) {
    Lexer lx(text);
    Msg root = parse_msg(lx, true);
    NetDef net;
    net.input_name = root.str("input");
    std::vector<std::string> dims;
    if (const Msg* shp = root.sub("input_shape")) dims = shp->all("dim");
    else dims = root.all("input_dim");
    if (net.input_name.empty() || dims.size() != 4)
        throw std::runtime_error("prototxt: need `input:` and a 4-d input_shape / input_dim");
    for (int i = 0; i < 4; ++i) net.input_dims[i] = atoi(dims[i].c_str());
    for (size_t i = 0; i < root.subs.size(); ++i) {
        if (root.subs[i].first != "layer") continue;
        const Msg& L = root.subs[i].second;
        LayerDef d;
        d.name = L.str("name");
        d.type_name = L.str("type");
        d.type = type_of(d.type_name);
        d.bottom = L.all("bottom");
        d.top = L.all("top");
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
            if (d.num_output <= 0) throw std::runtime_error("prototxt: '" + d.name + "' num_output missing");
        } else if (d.type == L_POOL) {
            const Msg* p = L.sub("pooling_param");
            if (!p) throw std::runtime_error("prototxt: '" + d.name + "' lacks pooling_param");
            d.pool_max = p->str("pool", "MAX") == "MAX";
            d.kernel = (int)p->num("kernel_size", 1);
            d.stride = (int)p->num("stride", 1);
            d.pad = (int)p->num("pad", 0);
            if (!d.pool_max) throw std::runtime_error("prototxt: '" + d.name + "': only MAX pooling is supported");
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
    return net;
}

NetDef parse_prototxt_file(const std::string& path) {
    std::ifstream f(path.c_str());
    if (!f) throw std::runtime_error("cannot open prototxt '" + path + "'");
    std::stringstream ss;
    ss << f.rdbuf();
    return parse_prototxt_text(ss.str());
}

// ------------------------------------------------------------------ caffemodel (protobuf wire format)
// caffe.proto field numbers: NetParameter{layers(V1)=2, layer=100};
// LayerParameter{name=1, blobs=7}; V1LayerParameter{name=4, blobs=6};
// BlobProto{num=1, channels=2, height=3, width=4, data=5, shape=7, double_data=8};
// BlobShape{dim=1}.
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

HostBlob read_blob(Reader r) {
    HostBlob b;
    int legacy[4] = {1, 1, 1, 1};
    bool has_legacy = false, has_shape = false;
    std::vector<double> dd;
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
    if (n != b.data.size()) throw std::runtime_error("caffemodel: blob shape does not match its data length");
    return b;
}

}  // namespace

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
