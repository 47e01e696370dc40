// Network plan + executor + the B2 half of the C ABI (see include/fcnod_b200.h).
//
// Replaces caffe.Net for /root/reference/scripts/fcn_object_detector.py:317,82,87,89-90:
// parses the prototxt, loads the caffemodel, plans one kernel launch per fused
// layer group on NHWC activations, and runs them on a stream.
//
// Plan-time fusions (none of them exist in the reference, which runs one Caffe
// layer per prototxt entry):
//   * Convolution + in-place ReLU              -> one conv launch (epilogue)
//   * Concat                                    -> eliminated: producers write
//                                                  straight into channel slices
//   * Dropout (TEST)                            -> alias
//   * Power on the input + NCHW->NHWC           -> one transform launch
#include <cuda.h>

#include <algorithm>
#include <cmath>
#include <cstring>
#include <map>
#include <memory>
#include <set>
#include <sstream>

#include "common.cuh"
#include "conv_simt.cuh"
#include "conv_tc.cuh"
#include "elementwise.cuh"
#include "kernels.h"
#include "netdef.h"

namespace fcnod {

static thread_local std::string g_last_error;
void set_last_error(const std::string& msg) { g_last_error = msg; }

namespace {

const int kNumSMs = 148;

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode_tiled() {
    static EncodeTiledFn fn = 0;
    if (!fn) {
        void* p = 0;
        cudaDriverEntryPointQueryResult qres;
        FCNOD_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres));
        FCNOD_CHECK(p && qres == cudaDriverEntryPointSuccess, "cuTensorMapEncodeTiled is not available in this driver");
        fn = (EncodeTiledFn)p;
    }
    return fn;
}

enum DType { DT_F32 = 0, DT_BF16 = 1 };
inline size_t dsize(int dt) { return dt == DT_F32 ? 4 : 2; }

struct Blob {
    std::string name;
    int C = 0, H = 0, W = 0;
    int dtype = DT_F32;
    int buf = -1;     // buffer index
    int c_off = 0;    // channel offset inside the buffer
    int ldc = 0;      // channels of the buffer
    bool is_output = false;
    int out_index = -1;
};

struct Buffer {
    int C = 0, H = 0, W = 0, dtype = DT_F32;
    void* ptr = 0;
};

enum StepKind { S_TRANSFORM, S_CONV_SIMT, S_CONV_TC, S_POOL, S_LRN, S_UNARY, S_COPY, S_EXPORT };

struct ConvWeights {
    int Cout = 0, Cin = 0, k = 0;
    std::vector<float> w, b;   // host copies: [Cout][Cin][k][k], [Cout]
    // SIMT
    float* d_w_simt = 0;
    int ldw = 0;
    // TC
    bf16* d_w_tc = 0;
    float* d_bias = 0;         // padded
    int CK = 0, n_cchunks = 0, BN = 0, n_tiles = 0;
};

struct Step {
    StepKind kind;
    std::string desc;
    int in = -1, out = -1;     // blob indices
    int layer = -1;
    int conv = -1;             // index into convs
    bool relu = false;
    int unary_op = 0;
    // TC state (rebuilt on reshape)
    CUtensorMap tmap_a, tmap_b;
    ConvTcParams tc;
    size_t tc_smem = 0;
    dim3 tc_grid;
};

}  // namespace
}  // namespace fcnod

using namespace fcnod;

struct fcnod_net {
    int device = 0, precision = 0, batch = 0;
    NetDef def;
    std::vector<Blob> blobs;
    std::map<std::string, int> blob_index;
    std::vector<Buffer> buffers;
    std::vector<ConvWeights> convs;
    std::vector<Step> steps;
    std::vector<int> outputs;              // blob indices of network outputs
    std::vector<float*> out_dev;           // dense NCHW fp32 per output
    float* data_dev = 0;                   // staging for host-fed forward
    float* scratch_export = 0;
    size_t scratch_export_elems = 0;
    // detection state
    void* post_scratch = 0;
    int32_t *det_dev = 0, *label_dev = 0, *weight_dev = 0, *count_dev = 0, *srchw_dev = 0;
    uint8_t* img_dev = 0;
    size_t img_dev_bytes = 0;
    float* pre_scratch = 0;
    cudaGraphExec_t graph = 0;
    cudaStream_t stream = 0;
    float* pinned_in = 0;
    size_t pinned_in_bytes = 0;
    int32_t* pinned_out = 0;
};

namespace fcnod {
namespace {

void require_device(const fcnod_net* net) {
    FCNOD_CHECK(net->device >= 0, "plan-only handle (device < 0): no compute is possible; there is no CPU fallback");
    FCNOD_CUDA(cudaSetDevice(net->device));
}

int blob_id(fcnod_net* n, const std::string& name) {
    std::map<std::string, int>::iterator it = n->blob_index.find(name);
    FCNOD_CHECK(it != n->blob_index.end(), "unknown blob '" + name + "'");
    return it->second;
}

int caffe_pool_out(int h, int k, int s, int p) {
    int o = (int)std::ceil((h + 2 * p - k) / (double)s) + 1;
    if (p > 0 && (o - 1) * s >= h + p) --o;
    return o;
}

void* view_ptr(const fcnod_net* n, const Blob& b) {
    const Buffer& bf = n->buffers[b.buf];
    return (char*)bf.ptr + (size_t)b.c_off * dsize(bf.dtype);
}

int grid_for(long work_items, int threads) {
    long blocks = (work_items + threads - 1) / threads;
    long cap = (long)kNumSMs * 16;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    return (int)blocks;
}

void choose_tc_tiling(ConvWeights& cw);

// ------------------------------------------------------------------------------------ planning
void build_plan(fcnod_net* net, const WeightMap& weights) {
    const NetDef& def = net->def;
    const bool bf = net->precision == FCNOD_PRECISION_BF16;
    FCNOD_CHECK(def.input_dims[1] > 0 && def.input_dims[2] > 0 && def.input_dims[3] > 0, "bad input dims");

    // ---- pass 1: shapes, consumers
    struct Shape { int C, H, W; };
    std::map<std::string, Shape> shp;
    std::map<std::string, int> producer;        // blob -> first layer writing it (not in-place)
    std::map<std::string, std::vector<int> > consumers;
    shp[def.input_name] = Shape{def.input_dims[1], def.input_dims[2], def.input_dims[3]};
    for (size_t i = 0; i < def.layers.size(); ++i) {
        const LayerDef& L = def.layers[i];
        FCNOD_CHECK(!L.bottom.empty(), "layer '" + L.name + "' has no bottom");
        for (size_t b = 0; b < L.bottom.size(); ++b) {
            FCNOD_CHECK(shp.count(L.bottom[b]), "layer '" + L.name + "': unknown bottom '" + L.bottom[b] + "'");
            if (!(L.top[0] == L.bottom[b])) consumers[L.bottom[b]].push_back((int)i);
        }
        Shape s = shp[L.bottom[0]], o = s;
        switch (L.type) {
            case L_CONV:
                o.C = L.num_output;
                o.H = (s.H + 2 * L.pad - L.kernel) / L.stride + 1;
                o.W = (s.W + 2 * L.pad - L.kernel) / L.stride + 1;
                break;
            case L_POOL:
                o.H = caffe_pool_out(s.H, L.kernel, L.stride, L.pad);
                o.W = caffe_pool_out(s.W, L.kernel, L.stride, L.pad);
                break;
            case L_CONCAT:
                o.C = 0;
                for (size_t b = 0; b < L.bottom.size(); ++b) {
                    Shape t = shp[L.bottom[b]];
                    FCNOD_CHECK(t.H == s.H && t.W == s.W, "Concat '" + L.name + "': spatial size mismatch");
                    o.C += t.C;
                }
                break;
            default: break;
        }
        FCNOD_CHECK(o.H > 0 && o.W > 0, "layer '" + L.name + "' produces an empty blob");
        if (shp.count(L.top[0])) {
            Shape e = shp[L.top[0]];
            FCNOD_CHECK(e.C == o.C && e.H == o.H && e.W == o.W, "in-place layer '" + L.name + "' changes shape");
        } else {
            shp[L.top[0]] = o;
            producer[L.top[0]] = (int)i;
        }
    }

    // ---- pass 2: aliasing (Dropout = same view; Concat bottoms = slices of the top)
    struct Alias { std::string root; int off; };
    std::map<std::string, Alias> alias;
    std::set<std::string> needs_copy;   // "<concat layer>|<bottom>" that could not be aliased
    for (size_t i = 0; i < def.layers.size(); ++i) {
        const LayerDef& L = def.layers[i];
        if (L.type == L_CONCAT) {
            int off = 0;
            for (size_t b = 0; b < L.bottom.size(); ++b) {
                const std::string& bn = L.bottom[b];
                bool ok = producer.count(bn) && !alias.count(bn);
                if (ok) {
                    LayerType pt = def.layers[producer[bn]].type;
                    ok = (pt == L_CONV || pt == L_POOL || pt == L_LRN);
                }
                if (ok) alias[bn] = Alias{L.top[0], off};
                else needs_copy.insert(L.name + "|" + bn);
                off += shp[bn].C;
            }
        }
    }
    // Dropout: the top aliases the bottom (processed when creating blobs, in order)

    // ---- network outputs = blobs nobody consumes
    std::set<std::string> outputs;
    for (std::map<std::string, Shape>::iterator it = shp.begin(); it != shp.end(); ++it)
        if (!consumers.count(it->first) && it->first != def.input_name) outputs.insert(it->first);

    // ---- dtype per blob
    std::map<std::string, int> dt;
    for (std::map<std::string, Shape>::iterator it = shp.begin(); it != shp.end(); ++it) {
        int d = bf ? DT_BF16 : DT_F32;
        const std::string& nm = it->first;
        if (nm == def.input_name) d = DT_F32;
        if (outputs.count(nm)) d = DT_F32;
        if (producer.count(nm) && def.layers[producer[nm]].type == L_POWER &&
            def.layers[producer[nm]].bottom[0] == def.input_name) d = DT_F32;
        if (bf && consumers.count(nm)) {   // feeds only Sigmoid layers that produce outputs
            bool all = true;
            for (size_t c = 0; c < consumers[nm].size(); ++c) {
                const LayerDef& L = def.layers[consumers[nm][c]];
                if (!(L.type == L_SIGMOID && outputs.count(L.top[0]))) all = false;
            }
            if (all) d = DT_F32;
        }
        dt[nm] = d;
    }

    // ---- create blobs + buffers in definition order
    net->blobs.clear(); net->blob_index.clear(); net->buffers.clear();
    struct Maker {
        fcnod_net* net; std::map<std::string, Shape>& shp; std::map<std::string, int>& dt;
        std::map<std::string, Alias>& alias; std::set<std::string>& outputs;
        int make(const std::string& nm) {
            if (net->blob_index.count(nm)) return net->blob_index[nm];
            Blob b; b.name = nm; Shape s = shp[nm]; b.C = s.C; b.H = s.H; b.W = s.W; b.dtype = dt[nm];
            if (alias.count(nm)) {
                int r = make(alias[nm].root);
                const Blob& rb = net->blobs[r];
                FCNOD_CHECK(rb.dtype == b.dtype, "concat slice dtype mismatch for '" + nm + "'");
                b.buf = rb.buf; b.c_off = rb.c_off + alias[nm].off; b.ldc = rb.ldc;
            } else {
                Buffer bu; bu.C = s.C; bu.H = s.H; bu.W = s.W; bu.dtype = b.dtype;
                net->buffers.push_back(bu);
                b.buf = (int)net->buffers.size() - 1; b.c_off = 0; b.ldc = s.C;
            }
            b.is_output = outputs.count(nm) > 0;
            net->blobs.push_back(b);
            net->blob_index[nm] = (int)net->blobs.size() - 1;
            return net->blob_index[nm];
        }
    } mk = {net, shp, dt, alias, outputs};

    // the input blob has no NHWC buffer of its own (it arrives NCHW); give it a descriptor only
    {
        Blob b; b.name = def.input_name; Shape s = shp[def.input_name]; b.C = s.C; b.H = s.H; b.W = s.W; b.dtype = DT_F32;
        b.buf = -1; b.ldc = s.C;
        net->blobs.push_back(b); net->blob_index[b.name] = 0;
    }

    net->steps.clear(); net->convs.clear(); net->outputs.clear();
    for (size_t i = 0; i < def.layers.size(); ++i) {
        const LayerDef& L = def.layers[i];
        const std::string& top = L.top[0];
        const bool inplace = (L.bottom[0] == top);
        if (L.type == L_DROPOUT) {
            if (!inplace) {   // alias top -> bottom view
                int bi = blob_id(net, L.bottom[0]);
                Blob b = net->blobs[bi];
                b.name = top; b.is_output = outputs.count(top) > 0;
                net->blobs.push_back(b); net->blob_index[top] = (int)net->blobs.size() - 1;
            }
            continue;
        }
        Step st; st.layer = (int)i;
        switch (L.type) {
            case L_POWER: {
                FCNOD_CHECK(L.bottom[0] == def.input_name, "Power is only supported directly on the network input ('" + L.name + "')");
                st.kind = S_TRANSFORM; st.in = 0; st.out = mk.make(top);
                st.desc = "transform " + L.name;
                net->steps.push_back(st);
                break;
            }
            case L_CONV: {
                st.in = blob_id(net, L.bottom[0]);
                FCNOD_CHECK(net->blobs[st.in].buf >= 0, "Convolution directly on the NCHW input is not supported; insert a Power layer ('" + L.name + "')");
                st.out = mk.make(top);
                // fuse a directly following in-place ReLU
                if (i + 1 < def.layers.size() && def.layers[i + 1].type == L_RELU && def.layers[i + 1].bottom[0] == top &&
                    def.layers[i + 1].top[0] == top)
                    st.relu = true;
                WeightMap::const_iterator wi = weights.find(L.name);
                FCNOD_CHECK(wi != weights.end(), "caffemodel has no blobs for layer '" + L.name + "'");
                const std::vector<HostBlob>& wb = wi->second;
                const Blob& ib = net->blobs[st.in];
                ConvWeights cw; cw.Cout = L.num_output; cw.Cin = ib.C; cw.k = L.kernel;
                size_t wn = (size_t)cw.Cout * cw.Cin * cw.k * cw.k;
                FCNOD_CHECK(wb.size() >= 1 && wb[0].data.size() == wn, "weight blob of '" + L.name + "' has the wrong size");
                cw.w = wb[0].data;
                if (L.bias_term) {
                    FCNOD_CHECK(wb.size() >= 2 && wb[1].data.size() == (size_t)cw.Cout, "bias blob of '" + L.name + "' has the wrong size");
                    cw.b = wb[1].data;
                } else cw.b.assign(cw.Cout, 0.f);
                const Blob& ob = net->blobs[st.out];
                bool tc_ok = bf && L.stride == 1 && 2 * L.pad == L.kernel - 1 && ib.dtype == DT_BF16 && ib.C % 8 == 0 &&
                             ib.ldc % 8 == 0 && ib.c_off % 8 == 0 &&
                             (ob.dtype == DT_F32 || (ob.ldc % 8 == 0 && ob.c_off % 8 == 0));
                st.kind = tc_ok ? S_CONV_TC : S_CONV_SIMT;
                if (tc_ok) choose_tc_tiling(cw);
                net->convs.push_back(cw);
                st.conv = (int)net->convs.size() - 1;
                std::ostringstream d;
                d << (tc_ok ? "conv_tc " : "conv_simt ") << L.name << " " << L.kernel << "x" << L.kernel << "/" << L.stride
                  << " Cin=" << cw.Cin << " Cout=" << cw.Cout << (st.relu ? " +relu" : "");
                st.desc = d.str();
                net->steps.push_back(st);
                break;
            }
            case L_RELU: {
                // skip when fused into the preceding convolution
                if (i > 0 && def.layers[i - 1].type == L_CONV && def.layers[i - 1].top[0] == top && inplace) break;
                st.kind = S_UNARY; st.unary_op = 0; st.in = blob_id(net, L.bottom[0]); st.out = mk.make(top);
                st.desc = "relu " + L.name;
                net->steps.push_back(st);
                break;
            }
            case L_SIGMOID: {
                st.kind = S_UNARY; st.unary_op = 1; st.in = blob_id(net, L.bottom[0]); st.out = mk.make(top);
                st.desc = "sigmoid " + L.name;
                net->steps.push_back(st);
                break;
            }
            case L_POOL: {
                st.kind = S_POOL; st.in = blob_id(net, L.bottom[0]); st.out = mk.make(top);
                FCNOD_CHECK(net->blobs[st.in].C % 8 == 0, "Pooling '" + L.name + "' needs channels % 8 == 0");
                st.desc = "pool " + L.name;
                net->steps.push_back(st);
                break;
            }
            case L_LRN: {
                st.kind = S_LRN; st.in = blob_id(net, L.bottom[0]); st.out = mk.make(top);
                FCNOD_CHECK(net->blobs[st.in].C % 8 == 0, "LRN '" + L.name + "' needs channels % 8 == 0");
                FCNOD_CHECK(L.local_size == 3 || L.local_size == 5 || L.local_size == 7 || L.local_size == 9,
                            "LRN '" + L.name + "': local_size must be 3, 5, 7 or 9");
                st.desc = "lrn " + L.name;
                net->steps.push_back(st);
                break;
            }
            case L_CONCAT: {
                int t = mk.make(top);
                int off = 0;
                for (size_t b = 0; b < L.bottom.size(); ++b) {
                    int bi = blob_id(net, L.bottom[b]);
                    if (needs_copy.count(L.name + "|" + L.bottom[b])) {
                        Step cs; cs.layer = (int)i; cs.kind = S_COPY; cs.in = bi; cs.out = t; cs.unary_op = off;
                        cs.desc = "concat-copy " + L.bottom[b];
                        FCNOD_CHECK(net->blobs[bi].dtype == net->blobs[t].dtype, "concat copy dtype mismatch");
                        net->steps.push_back(cs);
                    }
                    off += net->blobs[bi].C;
                }
                break;
            }
            default: FCNOD_CHECK(false, "unsupported layer type in '" + L.name + "'");
        }
    }
    // exports
    for (size_t i = 0; i < net->blobs.size(); ++i)
        if (net->blobs[i].is_output) {
            net->blobs[i].out_index = (int)net->outputs.size();
            net->outputs.push_back((int)i);
            Step st; st.kind = S_EXPORT; st.in = (int)i; st.desc = "export " + net->blobs[i].name;
            net->steps.push_back(st);
        }
}

// ------------------------------------------------------------------------------------ tcgen05 tiling (host only)
void choose_tc_tiling(ConvWeights& cw) {
    // K-block width: minimise padded K with a small per-block cost
    int best = 64; long best_cost = -1;
    const int cand[3] = {64, 32, 16};
    for (int i = 0; i < 3; ++i) {
        int ck = cand[i];
        long cost = (long)((cw.Cin + ck - 1) / ck) * (ck + 8);
        if (best_cost < 0 || cost < best_cost) { best_cost = cost; best = ck; }
    }
    cw.CK = best;
    cw.n_cchunks = (cw.Cin + cw.CK - 1) / cw.CK;
    cw.n_tiles = (cw.Cout + 127) / 128;
    cw.BN = ((cw.Cout + cw.n_tiles - 1) / cw.n_tiles + 15) / 16 * 16;
}

void fill_tc_geometry(const fcnod_net* net, Step& st, int batch) {
    const LayerDef& L = net->def.layers[st.layer];
    const ConvWeights& cw = net->convs[st.conv];
    const Blob& ib = net->blobs[st.in];
    const Blob& ob = net->blobs[st.out];
    ConvTcParams& p = st.tc;
    p.out = 0; p.bias = 0;
    p.H = ib.H; p.W = ib.W;
    p.Cout = cw.Cout; p.ldo = ob.ldc;
    p.kh = p.kw = cw.k; p.pad = L.pad;
    p.n_cchunks = cw.n_cchunks; p.CK = cw.CK;
    p.BW = std::min(ib.W, 128);
    p.BH = std::max(1, std::min(ib.H, 128 / p.BW));
    p.tiles_x = ceil_div(ib.W, p.BW);
    p.tiles_y = ceil_div(ib.H, p.BH);
    p.BN = cw.BN;
    p.relu = st.relu ? 1 : 0;
    p.out_fp32 = ob.dtype == DT_F32;
    uint32_t cols = 32;
    while ((int)cols < p.BN) cols <<= 1;
    p.tmem_cols = cols;
    const uint32_t span = cw.CK * 2;
    const size_t stage_bytes = 128 * span + align_up((size_t)p.BN * span, 1024);
    const int KB = cw.k * cw.k * cw.n_cchunks;
    int stages = (int)(96 * 1024 / stage_bytes);
    stages = std::max(2, std::min(stages, 8));
    stages = std::min(stages, std::max(2, KB));
    p.stages = stages;
    st.tc_smem = stages * stage_bytes + 16 * stages + 16 + 1024;
    st.tc_grid = dim3(p.tiles_x * p.tiles_y * batch, cw.n_tiles, 1);
}

// ------------------------------------------------------------------------------------ weights
void upload_weights(fcnod_net* net) {
    for (size_t si = 0; si < net->steps.size(); ++si) {
        Step& st = net->steps[si];
        if (st.kind != S_CONV_SIMT && st.kind != S_CONV_TC) continue;
        ConvWeights& cw = net->convs[st.conv];
        const int k = cw.k, Cin = cw.Cin, Cout = cw.Cout;
        if (st.kind == S_CONV_SIMT) {
            cw.ldw = (Cout + 63) / 64 * 64;
            const int K = k * k * Cin;
            std::vector<float> w((size_t)K * cw.ldw, 0.f);
            for (int o = 0; o < Cout; ++o)
                for (int c = 0; c < Cin; ++c)
                    for (int r = 0; r < k; ++r)
                        for (int s = 0; s < k; ++s)
                            w[(size_t)((r * k + s) * Cin + c) * cw.ldw + o] = cw.w[(((size_t)o * Cin + c) * k + r) * k + s];
            FCNOD_CUDA(cudaMalloc(&cw.d_w_simt, w.size() * 4));
            FCNOD_CUDA(cudaMemcpy(cw.d_w_simt, w.data(), w.size() * 4, cudaMemcpyHostToDevice));
            FCNOD_CUDA(cudaMalloc(&cw.d_bias, (size_t)cw.ldw * 4));
            std::vector<float> b(cw.ldw, 0.f);
            std::copy(cw.b.begin(), cw.b.end(), b.begin());
            FCNOD_CUDA(cudaMemcpy(cw.d_bias, b.data(), b.size() * 4, cudaMemcpyHostToDevice));
        } else {
            const int Kp = k * k * cw.n_cchunks * cw.CK, Np = cw.n_tiles * cw.BN;
            std::vector<bf16> w((size_t)Np * Kp, __float2bfloat16(0.f));
            for (int o = 0; o < Cout; ++o)
                for (int c = 0; c < Cin; ++c)
                    for (int r = 0; r < k; ++r)
                        for (int s = 0; s < k; ++s) {
                            size_t kk = ((size_t)(r * k + s) * cw.n_cchunks + c / cw.CK) * cw.CK + c % cw.CK;
                            w[(size_t)o * Kp + kk] = __float2bfloat16(cw.w[(((size_t)o * Cin + c) * k + r) * k + s]);
                        }
            FCNOD_CUDA(cudaMalloc(&cw.d_w_tc, w.size() * 2));
            FCNOD_CUDA(cudaMemcpy(cw.d_w_tc, w.data(), w.size() * 2, cudaMemcpyHostToDevice));
            std::vector<float> b(Np, 0.f);
            std::copy(cw.b.begin(), cw.b.end(), b.begin());
            FCNOD_CUDA(cudaMalloc(&cw.d_bias, b.size() * 4));
            FCNOD_CUDA(cudaMemcpy(cw.d_bias, b.data(), b.size() * 4, cudaMemcpyHostToDevice));
        }
    }
}

// ------------------------------------------------------------------------------------ (re)allocation
void free_activations(fcnod_net* net) {
    for (size_t i = 0; i < net->buffers.size(); ++i)
        if (net->buffers[i].ptr) { cudaFree(net->buffers[i].ptr); net->buffers[i].ptr = 0; }
    for (size_t i = 0; i < net->out_dev.size(); ++i) if (net->out_dev[i]) cudaFree(net->out_dev[i]);
    net->out_dev.clear();
    if (net->data_dev) { cudaFree(net->data_dev); net->data_dev = 0; }
    if (net->scratch_export) { cudaFree(net->scratch_export); net->scratch_export = 0; net->scratch_export_elems = 0; }
    if (net->post_scratch) { cudaFree(net->post_scratch); net->post_scratch = 0; }
    if (net->det_dev) { cudaFree(net->det_dev); net->det_dev = 0; }
    if (net->label_dev) { cudaFree(net->label_dev); net->label_dev = 0; }
    if (net->weight_dev) { cudaFree(net->weight_dev); net->weight_dev = 0; }
    if (net->count_dev) { cudaFree(net->count_dev); net->count_dev = 0; }
    if (net->srchw_dev) { cudaFree(net->srchw_dev); net->srchw_dev = 0; }
    if (net->img_dev) { cudaFree(net->img_dev); net->img_dev = 0; net->img_dev_bytes = 0; }
    if (net->pre_scratch) { cudaFree(net->pre_scratch); net->pre_scratch = 0; }
    if (net->graph) { cudaGraphExecDestroy(net->graph); net->graph = 0; }
    if (net->pinned_in) { cudaFreeHost(net->pinned_in); net->pinned_in = 0; net->pinned_in_bytes = 0; }
    if (net->pinned_out) { cudaFreeHost(net->pinned_out); net->pinned_out = 0; }
}

void setup_tc_step(fcnod_net* net, Step& st) {
    const LayerDef& L = net->def.layers[st.layer];
    ConvWeights& cw = net->convs[st.conv];
    const Blob& ib = net->blobs[st.in];
    const Blob& ob = net->blobs[st.out];
    const int B = net->batch;
    fill_tc_geometry(net, st, B);
    ConvTcParams& p = st.tc;
    p.out = view_ptr(net, ob);
    p.bias = cw.d_bias;
    const uint32_t span = cw.CK * 2;

    EncodeTiledFn enc = get_encode_tiled();
    const CUtensorMapSwizzle sw = span == 128 ? CU_TENSOR_MAP_SWIZZLE_128B
                                              : (span == 64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_32B);
    {
        cuuint64_t dims[4] = {(cuuint64_t)ib.C, (cuuint64_t)ib.W, (cuuint64_t)ib.H, (cuuint64_t)B};
        cuuint64_t strides[3] = {(cuuint64_t)ib.ldc * 2, (cuuint64_t)ib.W * ib.ldc * 2, (cuuint64_t)ib.H * ib.W * ib.ldc * 2};
        cuuint32_t box[4] = {(cuuint32_t)cw.CK, (cuuint32_t)p.BW, (cuuint32_t)p.BH, 1};
        cuuint32_t es[4] = {1, 1, 1, 1};
        CUresult r = enc(&st.tmap_a, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, view_ptr(net, ib), dims, strides, box, es,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        FCNOD_CHECK(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled(A) failed for '" + L.name + "' (" + std::to_string((int)r) + ")");
    }
    {
        const cuuint64_t Kp = (cuuint64_t)cw.k * cw.k * cw.n_cchunks * cw.CK;
        cuuint64_t dims[2] = {Kp, (cuuint64_t)cw.n_tiles * cw.BN};
        cuuint64_t strides[1] = {Kp * 2};
        cuuint32_t box[2] = {(cuuint32_t)cw.CK, (cuuint32_t)cw.BN};
        cuuint32_t es[2] = {1, 1};
        CUresult r = enc(&st.tmap_b, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, cw.d_w_tc, dims, strides, box, es,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        FCNOD_CHECK(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled(B) failed for '" + L.name + "' (" + std::to_string((int)r) + ")");
    }
}

void reshape(fcnod_net* net, int batch) {
    FCNOD_CHECK(batch > 0, "batch must be positive");
    FCNOD_CHECK(net->device >= 0, "plan-only handle (device < 0): no compute is possible; there is no CPU fallback");
    require_device(net);
    if (batch == net->batch && !net->buffers.empty() && net->buffers[0].ptr) return;
    FCNOD_CUDA(cudaDeviceSynchronize());
    free_activations(net);
    net->batch = batch;
    for (size_t i = 0; i < net->buffers.size(); ++i) {
        Buffer& b = net->buffers[i];
        size_t bytes = (size_t)batch * b.C * b.H * b.W * dsize(b.dtype);
        FCNOD_CUDA(cudaMalloc(&b.ptr, bytes));
        FCNOD_CUDA(cudaMemset(b.ptr, 0, bytes));
    }
    net->out_dev.assign(net->outputs.size(), (float*)0);
    for (size_t i = 0; i < net->outputs.size(); ++i) {
        const Blob& b = net->blobs[net->outputs[i]];
        FCNOD_CUDA(cudaMalloc(&net->out_dev[i], (size_t)batch * b.C * b.H * b.W * 4));
    }
    const Blob& in = net->blobs[0];
    FCNOD_CUDA(cudaMalloc(&net->data_dev, (size_t)batch * in.C * in.H * in.W * 4));
    for (size_t i = 0; i < net->steps.size(); ++i)
        if (net->steps[i].kind == S_CONV_TC) setup_tc_step(net, net->steps[i]);
}

// ------------------------------------------------------------------------------------ execution
template <typename T>
void launch_pool(const fcnod_net* net, const Step& st, const LayerDef& L, cudaStream_t s) {
    const Blob& ib = net->blobs[st.in];
    const Blob& ob = net->blobs[st.out];
    long items = (long)net->batch * ob.H * ob.W * (ib.C / 8);
    pool_max_kernel<T><<<grid_for(items, 256), 256, 0, s>>>((const T*)view_ptr(net, ib), (T*)view_ptr(net, ob), net->batch,
                                                             ib.C, ib.H, ib.W, ib.ldc, ob.H, ob.W, ob.ldc, L.kernel,
                                                             L.stride, L.pad);
}

template <typename T>
void launch_lrn(const fcnod_net* net, const Step& st, const LayerDef& L, cudaStream_t s) {
    const Blob& ib = net->blobs[st.in];
    const Blob& ob = net->blobs[st.out];
    long npix = (long)net->batch * ib.H * ib.W;
    long items = npix * (ib.C / 8);
    int g = grid_for(items, 256);
    const T* in = (const T*)view_ptr(net, ib);
    T* out = (T*)view_ptr(net, ob);
    float aon = L.alpha / L.local_size;
    switch (L.local_size) {
        case 3: lrn_kernel<T, 1><<<g, 256, 0, s>>>(in, out, npix, ib.C, ib.ldc, ob.ldc, aon, L.beta, L.lrn_k); break;
        case 5: lrn_kernel<T, 2><<<g, 256, 0, s>>>(in, out, npix, ib.C, ib.ldc, ob.ldc, aon, L.beta, L.lrn_k); break;
        case 7: lrn_kernel<T, 3><<<g, 256, 0, s>>>(in, out, npix, ib.C, ib.ldc, ob.ldc, aon, L.beta, L.lrn_k); break;
        default: lrn_kernel<T, 4><<<g, 256, 0, s>>>(in, out, npix, ib.C, ib.ldc, ob.ldc, aon, L.beta, L.lrn_k); break;
    }
}

template <typename T>
void launch_unary(const fcnod_net* net, const Step& st, cudaStream_t s) {
    const Blob& ib = net->blobs[st.in];
    const Blob& ob = net->blobs[st.out];
    long npix = (long)net->batch * ib.H * ib.W;
    int g = grid_for(npix * ib.C, 256);
    if (st.unary_op == 0)
        unary_kernel<T, 0><<<g, 256, 0, s>>>((const T*)view_ptr(net, ib), (T*)view_ptr(net, ob), npix, ib.C, ib.ldc, ob.ldc);
    else
        unary_kernel<T, 1><<<g, 256, 0, s>>>((const T*)view_ptr(net, ib), (T*)view_ptr(net, ob), npix, ib.C, ib.ldc, ob.ldc);
}

void launch_conv_simt(const fcnod_net* net, const Step& st, const LayerDef& L, cudaStream_t s) {
    const Blob& ib = net->blobs[st.in];
    const Blob& ob = net->blobs[st.out];
    const ConvWeights& cw = net->convs[st.conv];
    ConvSimtParams p;
    p.in = view_ptr(net, ib); p.w = cw.d_w_simt; p.bias = cw.d_bias; p.out = view_ptr(net, ob);
    p.N = net->batch; p.H = ib.H; p.W = ib.W; p.Cin = ib.C; p.ldi = ib.ldc;
    p.Ho = ob.H; p.Wo = ob.W; p.Cout = cw.Cout; p.ldo = ob.ldc; p.ldw = cw.ldw;
    p.kh = p.kw = cw.k; p.stride = L.stride; p.pad = L.pad; p.relu = st.relu ? 1 : 0;
    long M = (long)net->batch * ob.H * ob.W;
    dim3 grid((unsigned)((M + 63) / 64), (unsigned)(cw.ldw / 64));
    if (ib.dtype == DT_F32 && ob.dtype == DT_F32) conv_simt_kernel<float, float><<<grid, 256, 0, s>>>(p);
    else if (ib.dtype == DT_F32 && ob.dtype == DT_BF16) conv_simt_kernel<float, bf16><<<grid, 256, 0, s>>>(p);
    else if (ib.dtype == DT_BF16 && ob.dtype == DT_BF16) conv_simt_kernel<bf16, bf16><<<grid, 256, 0, s>>>(p);
    else conv_simt_kernel<bf16, float><<<grid, 256, 0, s>>>(p);
}

void launch_export(const fcnod_net* net, const Blob& b, float* dst, cudaStream_t s) {
    long total = (long)net->batch * b.C * b.H * b.W;
    int g = grid_for(total, 256);
    if (b.dtype == DT_F32) export_nchw_kernel<float><<<g, 256, 0, s>>>((const float*)view_ptr(net, b), dst, net->batch, b.C, b.H * b.W, b.ldc);
    else export_nchw_kernel<bf16><<<g, 256, 0, s>>>((const bf16*)view_ptr(net, b), dst, net->batch, b.C, b.H * b.W, b.ldc);
}

void run_step(fcnod_net* net, const Step& st, const float* data_dev, cudaStream_t s) {
    const LayerDef* L = st.layer >= 0 ? &net->def.layers[st.layer] : 0;
    switch (st.kind) {
        case S_TRANSFORM: {
            const Blob& ob = net->blobs[st.out];
            long npix = (long)net->batch * ob.H * ob.W;
            input_transform_kernel<<<grid_for(npix, 256), 256, 0, s>>>(data_dev, (float*)view_ptr(net, ob), net->batch, ob.C,
                                                                        ob.H * ob.W, L->shift, L->scale, L->power);
            break;
        }
        case S_CONV_SIMT: launch_conv_simt(net, st, *L, s); break;
        case S_CONV_TC:
            conv_tc_kernel<<<st.tc_grid, kConvTcThreads, st.tc_smem, s>>>(st.tmap_a, st.tmap_b, st.tc);
            break;
        case S_POOL:
            if (net->blobs[st.in].dtype == DT_F32) launch_pool<float>(net, st, *L, s);
            else launch_pool<bf16>(net, st, *L, s);
            break;
        case S_LRN:
            if (net->blobs[st.in].dtype == DT_F32) launch_lrn<float>(net, st, *L, s);
            else launch_lrn<bf16>(net, st, *L, s);
            break;
        case S_UNARY:
            FCNOD_CHECK(net->blobs[st.in].dtype == net->blobs[st.out].dtype, "unary dtype mismatch");
            if (net->blobs[st.in].dtype == DT_F32) launch_unary<float>(net, st, s);
            else launch_unary<bf16>(net, st, s);
            break;
        case S_COPY: {
            const Blob& ib = net->blobs[st.in];
            const Blob& ob = net->blobs[st.out];
            long npix = (long)net->batch * ib.H * ib.W;
            int g = grid_for(npix * ib.C, 256);
            if (ib.dtype == DT_F32)
                copy_channels_kernel<float><<<g, 256, 0, s>>>((const float*)view_ptr(net, ib), (float*)view_ptr(net, ob) + st.unary_op, npix, ib.C, ib.ldc, ob.ldc);
            else
                copy_channels_kernel<bf16><<<g, 256, 0, s>>>((const bf16*)view_ptr(net, ib), (bf16*)view_ptr(net, ob) + st.unary_op, npix, ib.C, ib.ldc, ob.ldc);
            break;
        }
        case S_EXPORT: {
            const Blob& b = net->blobs[st.in];
            launch_export(net, b, net->out_dev[b.out_index], s);
            break;
        }
    }
    FCNOD_CUDA(cudaGetLastError());
}

void forward_device(fcnod_net* net, const float* data_dev, cudaStream_t s) {
    FCNOD_CHECK(net->device >= 0, "plan-only handle (device < 0): no compute is possible; there is no CPU fallback");
    FCNOD_CHECK(net->batch > 0, "net not reshaped");
    for (size_t i = 0; i < net->steps.size(); ++i) run_step(net, net->steps[i], data_dev, s);
}

void ensure_detect_buffers(fcnod_net* net) {
    const int B = net->batch;
    const int c_cvg = net->blobs[blob_id(net, "coverage")].C;
    if (!net->post_scratch) FCNOD_CUDA(cudaMalloc(&net->post_scratch, postnet_scratch_bytes(B, c_cvg)));
    if (!net->det_dev) FCNOD_CUDA(cudaMalloc(&net->det_dev, (size_t)B * FCNOD_MAX_DET * 5 * 4));
    if (!net->label_dev) FCNOD_CUDA(cudaMalloc(&net->label_dev, (size_t)B * FCNOD_MAX_DET * 4));
    if (!net->weight_dev) FCNOD_CUDA(cudaMalloc(&net->weight_dev, (size_t)B * FCNOD_MAX_DET * 4));
    if (!net->count_dev) FCNOD_CUDA(cudaMalloc(&net->count_dev, (size_t)B * 4));
    if (!net->srchw_dev) FCNOD_CUDA(cudaMalloc(&net->srchw_dev, (size_t)B * 2 * 4));
    if (!net->pinned_out) FCNOD_CUDA(cudaMallocHost(&net->pinned_out, (size_t)B * (FCNOD_MAX_DET * 7 + 1) * 4));
}

void detect_device(fcnod_net* net, const float* data_dev, const int32_t* srchw_dev, const fcnod_postnet_params& p,
                   cudaStream_t s) {
    forward_device(net, data_dev, s);
    int ic = blob_id(net, "coverage"), ib = blob_id(net, "bboxes");
    const Blob& bc = net->blobs[ic];
    const Blob& bb = net->blobs[ib];
    FCNOD_CHECK(bc.is_output && bb.is_output, "blobs 'coverage' and 'bboxes' must be network outputs");
    FCNOD_CHECK(bc.H == bb.H && bc.W == bb.W, "coverage / bboxes grid mismatch");
    ensure_detect_buffers(net);
    postnet_launch(net->out_dev[bc.out_index], net->out_dev[bb.out_index], net->batch, bc.C, bb.C, bc.H, bc.W, srchw_dev, p,
                   net->post_scratch, net->det_dev, net->label_dev, net->weight_dev, net->count_dev, s);
}

void copy_detections_out(fcnod_net* net, int B, int32_t* out_det, int32_t* out_label, int32_t* out_weight,
                         int32_t* out_count, cudaStream_t s) {
    int32_t* po = net->pinned_out;
    int32_t* p_det = po;
    int32_t* p_lab = p_det + (size_t)B * FCNOD_MAX_DET * 5;
    int32_t* p_w = p_lab + (size_t)B * FCNOD_MAX_DET;
    int32_t* p_cnt = p_w + (size_t)B * FCNOD_MAX_DET;
    FCNOD_CUDA(cudaMemcpyAsync(p_det, net->det_dev, (size_t)B * FCNOD_MAX_DET * 5 * 4, cudaMemcpyDeviceToHost, s));
    FCNOD_CUDA(cudaMemcpyAsync(p_lab, net->label_dev, (size_t)B * FCNOD_MAX_DET * 4, cudaMemcpyDeviceToHost, s));
    FCNOD_CUDA(cudaMemcpyAsync(p_w, net->weight_dev, (size_t)B * FCNOD_MAX_DET * 4, cudaMemcpyDeviceToHost, s));
    FCNOD_CUDA(cudaMemcpyAsync(p_cnt, net->count_dev, (size_t)B * 4, cudaMemcpyDeviceToHost, s));
    FCNOD_CUDA(cudaStreamSynchronize(s));
    memcpy(out_det, p_det, (size_t)B * FCNOD_MAX_DET * 5 * 4);
    memcpy(out_label, p_lab, (size_t)B * FCNOD_MAX_DET * 4);
    memcpy(out_weight, p_w, (size_t)B * FCNOD_MAX_DET * 4);
    memcpy(out_count, p_cnt, (size_t)B * 4);
}

}  // namespace
}  // namespace fcnod

// =========================================================================================== C ABI
extern "C" {

int fcnod_version(void) { return 100; }
const char* fcnod_last_error(void) { return g_last_error.c_str(); }

int fcnod_device_count(int* count) {
    FCNOD_API_BEGIN
    FCNOD_CUDA(cudaGetDeviceCount(count));
    FCNOD_API_END
}

int fcnod_net_create(const char* prototxt_path, const char* caffemodel_path, int device, int precision, fcnod_net** out) {
    FCNOD_API_BEGIN
    FCNOD_CHECK(out, "out is null");
    *out = 0;
    FCNOD_CHECK(prototxt_path && caffemodel_path, "prototxt / caffemodel path is null");
    FCNOD_CHECK(precision == FCNOD_PRECISION_FP32 || precision == FCNOD_PRECISION_BF16, "bad precision");
    std::unique_ptr<fcnod_net> net(new fcnod_net());
    net->device = device;
    net->precision = precision;
    net->def = parse_prototxt_file(prototxt_path);
    WeightMap w = read_caffemodel(caffemodel_path);
    build_plan(net.get(), w);
    if (device < 0) {   // plan-only handle: host-side introspection (describe / dims), every compute call fails
        net->batch = net->def.input_dims[0] > 0 ? net->def.input_dims[0] : 1;
        for (size_t i = 0; i < net->steps.size(); ++i)
            if (net->steps[i].kind == S_CONV_TC) fill_tc_geometry(net.get(), net->steps[i], net->batch);
        *out = net.release();
        return 0;
    }
    // CUDA from here on: fail loudly when there is no device (no CPU fallback exists).
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    FCNOD_CHECK(e == cudaSuccess && ndev > 0, std::string("no usable CUDA device: ") + cudaGetErrorString(e));
    FCNOD_CHECK(device >= 0 && device < ndev, "device index out of range");
    FCNOD_CUDA(cudaSetDevice(device));
    cudaDeviceProp prop;
    FCNOD_CUDA(cudaGetDeviceProperties(&prop, device));
    FCNOD_CHECK(prop.major == 10, "fcnod_b200 is built for sm_100a (Blackwell B200) only; found sm_" +
                                      std::to_string(prop.major) + std::to_string(prop.minor));
    FCNOD_CUDA(cudaFuncSetAttribute(conv_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    FCNOD_CUDA(cudaStreamCreateWithFlags(&net->stream, cudaStreamNonBlocking));
    upload_weights(net.get());
    reshape(net.get(), net->def.input_dims[0] > 0 ? net->def.input_dims[0] : 1);
    *out = net.release();
    FCNOD_API_END
}

int fcnod_net_destroy(fcnod_net* net) {
    FCNOD_API_BEGIN
    if (net && net->device < 0) { delete net; net = 0; }
    if (net) {
        cudaSetDevice(net->device);
        cudaDeviceSynchronize();
        free_activations(net);
        for (size_t i = 0; i < net->convs.size(); ++i) {
            if (net->convs[i].d_w_simt) cudaFree(net->convs[i].d_w_simt);
            if (net->convs[i].d_w_tc) cudaFree(net->convs[i].d_w_tc);
            if (net->convs[i].d_bias) cudaFree(net->convs[i].d_bias);
        }
        if (net->stream) cudaStreamDestroy(net->stream);
        delete net;
    }
    FCNOD_API_END
}

int fcnod_net_reshape(fcnod_net* net, int batch) {
    FCNOD_API_BEGIN
    FCNOD_CHECK(net, "net is null");
    reshape(net, batch);
    FCNOD_API_END
}

int fcnod_net_input_dims(const fcnod_net* net, int dims[4]) {
    FCNOD_API_BEGIN
    FCNOD_CHECK(net && dims, "null argument");
    const Blob& b = net->blobs[0];
    dims[0] = net->batch; dims[1] = b.C; dims[2] = b.H; dims[3] = b.W;
    FCNOD_API_END
}

int fcnod_net_blob_dims(const fcnod_net* net, const char* blob, int dims[4]) {
    FCNOD_API_BEGIN
    FCNOD_CHECK(net && blob && dims, "null argument");
    const Blob& b = net->blobs[blob_id(const_cast<fcnod_net*>(net), blob)];
    dims[0] = net->batch; dims[1] = b.C; dims[2] = b.H; dims[3] = b.W;
    FCNOD_API_END
}

int fcnod_net_num_blobs(const fcnod_net* net) { return net ? (int)net->blobs.size() : 0; }
const char* fcnod_net_blob_name(const fcnod_net* net, int index) {
    if (!net || index < 0 || index >= (int)net->blobs.size()) return 0;
    return net->blobs[index].name.c_str();
}

int fcnod_net_launches_per_forward(const fcnod_net* net) { return net ? (int)net->steps.size() : 0; }

int fcnod_net_describe(const fcnod_net* net, char* buf, size_t buflen) {
    FCNOD_API_BEGIN
    FCNOD_CHECK(net && buf && buflen > 0, "null argument");
    std::ostringstream o;
    for (size_t i = 0; i < net->steps.size(); ++i) {
        const Step& st = net->steps[i];
        o << i << ": " << st.desc;
        if (st.kind == S_CONV_TC) {
            const ConvWeights& cw = net->convs[st.conv];
            o << " [CK=" << cw.CK << " BN=" << cw.BN << "x" << cw.n_tiles << " stages=" << st.tc.stages << " grid=" << st.tc_grid.x
              << "x" << st.tc_grid.y << " smem=" << st.tc_smem << "]";
        }
        if (st.out >= 0) {
            const Blob& b = net->blobs[st.out];
            o << " -> " << b.name << " (" << b.C << "," << b.H << "," << b.W << ") " << (b.dtype == DT_F32 ? "f32" : "bf16")
              << " buf" << b.buf << "+" << b.c_off << "/" << b.ldc;
        }
        o << "\n";
    }
    std::string s = o.str();
    FCNOD_CHECK(s.size() + 1 <= buflen, "describe buffer too small (need " + std::to_string(s.size() + 1) + ")");
    memcpy(buf, s.c_str(), s.size() + 1);
    FCNOD_API_END
}

int fcnod_net_forward_device(fcnod_net* net, const float* data_dev, void* stream) {
    FCNOD_API_BEGIN
    FCNOD_CHECK(net && data_dev, "null argument");
    require_device(net);
    forward_device(net, data_dev, (cudaStream_t)stream);
    FCNOD_API_END
}

int fcnod_net_forward(fcnod_net* net, const float* data_host) {
    FCNOD_API_BEGIN
    FCNOD_CHECK(net && data_host, "null argument");
    require_device(net);
    const Blob& in = net->blobs[0];
    size_t bytes = (size_t)net->batch * in.C * in.H * in.W * 4;
    FCNOD_CUDA(cudaMemcpyAsync(net->data_dev, data_host, bytes, cudaMemcpyHostToDevice, net->stream));
    forward_device(net, net->data_dev, net->stream);
    FCNOD_CUDA(cudaStreamSynchronize(net->stream));
    FCNOD_API_END
}

int fcnod_net_blob_read(fcnod_net* net, const char* blob, float* host_out, size_t count) {
    FCNOD_API_BEGIN
    FCNOD_CHECK(net && blob && host_out, "null argument");
    require_device(net);
    const Blob& b = net->blobs[blob_id(net, blob)];
    size_t n = (size_t)net->batch * b.C * b.H * b.W;
    FCNOD_CHECK(count == n, "blob_read: count mismatch (blob has " + std::to_string(n) + " elements)");
    if (b.buf < 0) {
        FCNOD_CUDA(cudaMemcpy(host_out, net->data_dev, n * 4, cudaMemcpyDeviceToHost));
    } else if (b.is_output) {
        FCNOD_CUDA(cudaStreamSynchronize(net->stream));
        FCNOD_CUDA(cudaMemcpy(host_out, net->out_dev[b.out_index], n * 4, cudaMemcpyDeviceToHost));
    } else {
        if (net->scratch_export_elems < n) {
            if (net->scratch_export) cudaFree(net->scratch_export);
            FCNOD_CUDA(cudaMalloc(&net->scratch_export, n * 4));
            net->scratch_export_elems = n;
        }
        launch_export(net, b, net->scratch_export, net->stream);
        FCNOD_CUDA(cudaGetLastError());
        FCNOD_CUDA(cudaStreamSynchronize(net->stream));
        FCNOD_CUDA(cudaMemcpy(host_out, net->scratch_export, n * 4, cudaMemcpyDeviceToHost));
    }
    FCNOD_API_END
}

int fcnod_net_output_device(fcnod_net* net, const char* blob, const float** dev_ptr) {
    FCNOD_API_BEGIN
    FCNOD_CHECK(net && blob && dev_ptr, "null argument");
    const Blob& b = net->blobs[blob_id(net, blob)];
    FCNOD_CHECK(b.is_output, std::string("'") + blob + "' is not a network output");
    *dev_ptr = net->out_dev[b.out_index];
    FCNOD_API_END
}

int fcnod_net_run_step(fcnod_net* net, int step, int reps, void* stream) {
    FCNOD_API_BEGIN
    FCNOD_CHECK(net && step >= 0 && step < (int)net->steps.size(), "bad step index");
    require_device(net);
    for (int i = 0; i < reps; ++i) run_step(net, net->steps[step], net->data_dev, (cudaStream_t)stream);
    FCNOD_API_END
}

int fcnod_detect_device(fcnod_net* net, const float* data_dev, int B, const fcnod_postnet_params* p,
                        const int32_t** det_dev, const int32_t** label_dev, const int32_t** weight_dev,
                        const int32_t** count_dev, void* stream) {
    FCNOD_API_BEGIN
    FCNOD_CHECK(net && data_dev && p, "null argument");
    require_device(net);
    reshape(net, B);
    detect_device(net, data_dev, 0, *p, (cudaStream_t)stream);
    if (det_dev) *det_dev = net->det_dev;
    if (label_dev) *label_dev = net->label_dev;
    if (weight_dev) *weight_dev = net->weight_dev;
    if (count_dev) *count_dev = net->count_dev;
    FCNOD_API_END
}

int fcnod_detect_graph_build(fcnod_net* net, const float* data_dev, int B, const fcnod_postnet_params* p, void* stream) {
    FCNOD_API_BEGIN
    FCNOD_CHECK(net && data_dev && p, "null argument");
    require_device(net);
    reshape(net, B);
    cudaStream_t s = (cudaStream_t)stream;
    // warm run outside capture: allocates the detection buffers
    detect_device(net, data_dev, 0, *p, s);
    FCNOD_CUDA(cudaStreamSynchronize(s));
    if (net->graph) { cudaGraphExecDestroy(net->graph); net->graph = 0; }
    cudaGraph_t g = 0;
    FCNOD_CUDA(cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal));
    try {
        detect_device(net, data_dev, 0, *p, s);
    } catch (...) {
        cudaStreamEndCapture(s, &g);
        if (g) cudaGraphDestroy(g);
        throw;
    }
    FCNOD_CUDA(cudaStreamEndCapture(s, &g));
    cudaError_t e = cudaGraphInstantiate(&net->graph, g, 0);
    cudaGraphDestroy(g);
    FCNOD_CUDA(e);
    FCNOD_API_END
}

int fcnod_detect_graph_launch(fcnod_net* net, void* stream) {
    FCNOD_API_BEGIN
    FCNOD_CHECK(net && net->graph, "no graph built");
    FCNOD_CUDA(cudaGraphLaunch(net->graph, (cudaStream_t)stream));
    FCNOD_API_END
}

int fcnod_detect_from_data(fcnod_net* net, const float* data_host, int B, const fcnod_postnet_params* p,
                           int32_t* out_det, int32_t* out_label, int32_t* out_weight, int32_t* out_count) {
    FCNOD_API_BEGIN
    FCNOD_CHECK(net && data_host && p && out_det && out_label && out_weight && out_count, "null argument");
    require_device(net);
    reshape(net, B);
    const Blob& in = net->blobs[0];
    size_t bytes = (size_t)B * in.C * in.H * in.W * 4;
    cudaStream_t s = net->stream;
    FCNOD_CUDA(cudaMemcpyAsync(net->data_dev, data_host, bytes, cudaMemcpyHostToDevice, s));
    detect_device(net, net->data_dev, 0, *p, s);
    copy_detections_out(net, B, out_det, out_label, out_weight, out_count, s);
    FCNOD_API_END
}

int fcnod_detect(fcnod_net* net, const uint8_t* images, int B, int src_h, int src_w, const fcnod_postnet_params* p,
                 int32_t* out_det, int32_t* out_label, int32_t* out_weight, int32_t* out_count) {
    FCNOD_API_BEGIN
    FCNOD_CHECK(net && images && p && out_det && out_label && out_weight && out_count, "null argument");
    FCNOD_CHECK(src_h > 0 && src_w > 0, "bad source size");
    require_device(net);
    reshape(net, B);
    const Blob& in = net->blobs[0];
    FCNOD_CHECK(in.C == 3, "detect needs a 3-channel input blob");
    cudaStream_t s = net->stream;
    size_t ibytes = (size_t)B * src_h * src_w * 3;
    if (net->img_dev_bytes < ibytes) {
        if (net->img_dev) cudaFree(net->img_dev);
        FCNOD_CUDA(cudaMalloc(&net->img_dev, ibytes));
        net->img_dev_bytes = ibytes;
    }
    if (!net->pre_scratch) FCNOD_CUDA(cudaMalloc(&net->pre_scratch, (size_t)B * 6 * 4));
    ensure_detect_buffers(net);
    FCNOD_CUDA(cudaMemcpyAsync(net->img_dev, images, ibytes, cudaMemcpyHostToDevice, s));
    preprocess_launch(net->img_dev, B, src_h, src_w, in.H, in.W, net->data_dev, net->pre_scratch, s);
    std::vector<int32_t> hw((size_t)B * 2);
    for (int i = 0; i < B; ++i) { hw[2 * i] = src_h; hw[2 * i + 1] = src_w; }
    FCNOD_CUDA(cudaMemcpyAsync(net->srchw_dev, hw.data(), hw.size() * 4, cudaMemcpyHostToDevice, s));
    FCNOD_CUDA(cudaStreamSynchronize(s));   // hw is a stack-lifetime host buffer
    detect_device(net, net->data_dev, net->srchw_dev, *p, s);
    copy_detections_out(net, B, out_det, out_label, out_weight, out_count, s);
    FCNOD_API_END
}

}  // extern "C"
