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
// bf16 ("deploy") precision only -- the fp32 precision is the layer-by-layer parity mode and keeps
// every blob materialised (set FCNOD_NO_FUSION=1 to get that plan in bf16 too):
//   * Power + first stride-2 convolution        -> 2x2 space-to-depth transform + ONE stride-1 tcgen05
//                                                  convolution; scale folded into the weights, shift into a
//                                                  border-aware bias table (zero padding happens AFTER the shift)
//   * Pooling -> LRN, LRN -> Pooling            -> one launch each (fused_pool.cuh)
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
#include "conv_tc2.cuh"
#include "elementwise.cuh"
#include "fused_pool.cuh"
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
    int buf = -1;     // buffer index; -1 = the NCHW input; -2 = fused away (never materialised)
    int c_off = 0;    // channel offset inside the buffer
    int ldc = 0;      // channels of the buffer
    bool is_output = false;
    int out_index = -1;
};

struct Buffer {
    int C = 0, H = 0, W = 0, dtype = DT_F32;
    void* ptr = 0;
};

enum StepKind { S_TRANSFORM, S_S2D, S_CONV_SIMT, S_CONV_TC, S_POOL, S_LRN, S_POOL_LRN, S_LRN_POOL, S_UNARY, S_COPY, S_EXPORT };

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
    bool halo = false;         // k x k taps served from one shared-memory patch per channel chunk (conv_tc2.cuh)
    bool pool = false;         // a 3x3 / stride 1 / pad 1 max-pool of the input runs on the A-operand path (POOL mode)
    bool lrn = false;          // the LRN that follows (local_size 5, beta 0.75) runs in the epilogue; needs one N tile
    float lrn_aon = 0.f, lrn_k = 1.f;
    // space-to-depth form of a stride-2 convolution on the (Power-transformed) network input
    bool s2d = false;
    int ka = 0;                // taps per axis after space-to-depth (padded to 1/2/4)
    int q = 0;                 // even padding: ih = 2*(oh - q/2 + a) + dy
    int stride = 1, pad = 0;
    float in_scale = 1.f, in_shift = 0.f;   // the folded Power layer: y = in_shift + in_scale * x
    float* d_bias_tab = 0;
    int bt_lo_h = 0, bt_hi_h = 0, bt_lo_w = 0, bt_hi_w = 0;
};

const int kS2dCpad = 16;       // channels per space-to-depth pixel (4*Cin padded)

// Geometry of the padded, pre-swizzled space-to-depth image that feeds the first convolution with plain bulk copies
// (cp.async.bulk, one per patch row): box widths are multiples of 8 pixels so that the 32-byte-swizzle phase of a pixel
// in shared memory, (x >> 2) & 1, is the same as that of its physical column in global memory.
struct S2dGeom { int BW, PW, BH, tiles_x, tiles_y, Wp, Hp; };
static S2dGeom s2d_geom(int H2, int W2, int ka) {
    S2dGeom g;
    for (int ntx = (W2 + 127) / 128;; ++ntx) {
        g.BW = ((W2 + ntx - 1) / ntx + 7) / 8 * 8;
        g.PW = (g.BW + ka - 1 + 7) / 8 * 8;
        if (g.PW <= 128) break;
    }
    g.BH = std::max(1, std::min(H2, 128 / g.PW));
    g.tiles_x = (W2 + g.BW - 1) / g.BW;
    g.tiles_y = (H2 + g.BH - 1) / g.BH;
    g.Wp = (g.tiles_x - 1) * g.BW + g.PW;
    g.Hp = g.tiles_y * g.BH + ka - 1;
    return g;
}

struct Step {
    StepKind kind;
    std::string desc;
    int in = -1, out = -1;     // blob indices
    int layer = -1;
    int layer2 = -1;           // second layer of a fused pair (Pool->LRN, LRN->Pool)
    int conv = -1;             // index into convs
    int s2d_buf = -1;          // S_S2D / s2d convolution: buffer index of the space-to-depth image
    int tc_ncand = 0, tc_pick = -1;      // tensor-core tiling candidates (ranked by the model) and the one in use
    int s2d_wp = 0, s2d_hp = 0, s2d_padl = 0;    // padded space-to-depth image: width, height, zero columns / rows before (0, 0)
    bool s2d_overlap = false;  // one K-block per kernel row (overlapping-window tensor map) vs one per tap
    bool relu = false;
    int unary_op = 0;
    // tensor-core convolution: output segments (one per fused convolution) + launch state (rebuilt on reshape)
    struct Seg { int blob; int direct_blob; int col0, ncols, flags; std::string name; };
    std::vector<Seg> segs;
    CUtensorMap tmap_a, tmap_b;
    ConvTc2Params tc;
    size_t tc_smem = 0;
    int tc_grid = 0;
    // lanes (plan_lanes): independent branches of the layer graph run on different streams
    int lane = 0;
    std::vector<int> waits;    // steps on OTHER lanes whose completion event this step waits for
    bool signal = false;       // some later step on another lane waits for this one
    std::vector<int> deps;     // every earlier step this one depends on (RAW / WAR / WAW on a buffer region)
    int grid_cap = 0;          // > 0: this launch may use only that many CTAs (its share of the SMs while independent launches of
                               //      other lanes run beside it, see partition_lane_groups)
    float solo_ms = 0.f;       // launch time alone on the whole device (autotuner)
    cudaEvent_t ev = 0;
};

}  // namespace
}  // namespace fcnod

using namespace fcnod;

struct fcnod_net {
    int device = 0, precision = 0, batch = 0;
    int capacity = 0;                      // frames the activation buffers were allocated for (>= batch)
    std::map<int, std::vector<int> > tc_picks;   // batch -> tiling pick of every step (autotuned once per batch size)
    std::map<int, std::vector<int> > tc_caps;    // batch -> grid cap of every step (partition_lane_groups)
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
    // the forward's lanes: internal streams with descending priority (lane 0 carries the critical chain); they fork from
    // the caller's stream at the start of a forward and join it again at the end
    cudaStream_t lane_stream[3] = {0, 0, 0};
    cudaEvent_t fork_ev = 0, join_ev[3] = {0, 0, 0};
    int n_lanes = 1;
    // detect paths on the fused bf16 plan: the prelude writes the first convolution's space-to-depth operand directly and
    // the fp32 blob `data` is only materialised when somebody reads it (from the frames still on the device)
    bool fused_prelude = false;
    bool data_valid = true;
    int srchw_b = 0, srchw_h = 0, srchw_w = 0;      // what srchw_dev holds (streaming detect)
    const uint8_t* data_img = 0;
    int data_img_b = 0, data_img_h = 0, data_img_w = 0;
    cudaStream_t copy_stream = 0;        // uploads of fcnod_detect's second half-batch, overlapped with the first one's compute
    cudaEvent_t copy_ev[2] = {0, 0};
    int32_t* pinned_big = 0;             // detections of a whole fcnod_detect call (all chunks)
    size_t pinned_big_bytes = 0;
    float* pinned_in = 0;
    size_t pinned_in_bytes = 0;
    int32_t* pinned_out = 0;
    // streaming detect (fcnod_detect_submit / _collect): two batches in flight
    struct DetectSlot {
        uint8_t* img = 0;                // device copy of the batch's frames
        size_t img_bytes = 0;
        int32_t* pinned = 0;             // [B*2 src_hw | det | label | weight | count]
        size_t pinned_bytes = 0;
        cudaEvent_t up = 0, done = 0;    // upload (+ prelude) finished (copy stream) / detections on the host (compute stream)
        // fused-prelude plan: the slot's own space-to-depth image, min/max scratch and graph (conv1 reads THIS image), so that
        // the prelude of batch i+1 runs on the copy stream while batch i is still computing
        void* s2d = 0;
        void* mm = 0;
        cudaGraphExec_t graph = 0;
        fcnod_postnet_params graph_p;
        int B = 0;
        int ticket = -1;                 // -1 = free
    } slots[2];
    int next_ticket = 0, oldest_ticket = 0;
    cudaGraphExec_t sgraph = 0;          // forward + post-net from data_dev / srchw_dev, replayed by every submit
    fcnod_postnet_params sgraph_p;       // the thresholds baked into it
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
void plan_lanes(fcnod_net* net);

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
    std::map<std::string, int> inplace_ops;     // blob -> number of in-place layers rewriting it
    std::map<std::string, bool> nonneg;         // blob is >= 0 everywhere (ReLU / Sigmoid outputs and what pools, LRNs, concats make of them)
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
        switch (L.type) {
            case L_RELU: case L_SIGMOID: nonneg[L.top[0]] = true; break;
            case L_POOL: case L_LRN: case L_DROPOUT: nonneg[L.top[0]] = nonneg[L.bottom[0]] && (L.type != L_LRN || L.lrn_k > 0.f); break;
            case L_CONCAT: {
                bool all = true;
                for (size_t b = 0; b < L.bottom.size(); ++b) all = all && nonneg[L.bottom[b]];
                nonneg[L.top[0]] = all;
                break;
            }
            default: nonneg[L.top[0]] = false; break;
        }
        if (shp.count(L.top[0])) {
            Shape e = shp[L.top[0]];
            FCNOD_CHECK(e.C == o.C && e.H == o.H && e.W == o.W, "in-place layer '" + L.name + "' changes shape");
            inplace_ops[L.top[0]]++;
        } else {
            shp[L.top[0]] = o;
            producer[L.top[0]] = (int)i;
        }
    }

    // ---- network outputs = blobs nobody consumes
    std::set<std::string> outputs;
    for (std::map<std::string, Shape>::iterator it = shp.begin(); it != shp.end(); ++it)
        if (!consumers.count(it->first) && it->first != def.input_name) outputs.insert(it->first);

    // ---- pass 1b: fusions of the bf16 plan (see the header comment)
    const bool fuse = bf && !getenv("FCNOD_NO_FUSION");
    std::set<std::string> virt;            // blobs that are never materialised
    std::map<int, int> s2d_conv;           // conv layer -> Power layer folded into it
    std::map<int, int> s2d_power;          // Power layer -> its conv
    std::map<int, int> fuse_first;         // first layer of a fused pair -> second
    std::map<int, int> fuse_second;        // second -> first
    if (fuse) {
        for (size_t i = 0; i < def.layers.size(); ++i) {
            const LayerDef& L = def.layers[i];
            const std::string& top = L.top[0];
            if (L.bottom[0] == top || inplace_ops.count(top) || outputs.count(top)) continue;
            if (!consumers.count(top) || consumers[top].size() != 1) continue;
            const int ci = consumers[top][0];
            const LayerDef& Cn = def.layers[ci];
            if (Cn.bottom.size() != 1 || Cn.top[0] == top) continue;
            const Shape si = shp[L.bottom[0]], so = shp[Cn.top[0]];
            if (L.type == L_POWER && L.bottom[0] == def.input_name && L.power == 1.f && Cn.type == L_CONV && Cn.stride == 2 &&
                4 * si.C <= kS2dCpad && si.H % 2 == 0 && si.W % 2 == 0 && so.H == si.H / 2 && so.W == si.W / 2 &&
                so.W >= 8 && Cn.num_output % 8 == 0) {
                const int q = Cn.pad + (Cn.pad & 1);
                const int ka = (q - Cn.pad + Cn.kernel + 1) / 2;
                if (ka <= 4) { s2d_conv[ci] = (int)i; s2d_power[(int)i] = ci; virt.insert(top); }
            } else if (L.type == L_POOL && Cn.type == L_LRN && L.pool_max && L.kernel == 3 && L.stride == 2 &&
                       Cn.local_size == 5 && si.C % 8 == 0 && si.C / 8 <= 256 && !fuse_second.count((int)i)) {
                fuse_first[(int)i] = ci; fuse_second[ci] = (int)i; virt.insert(top);
            } else if (L.type == L_LRN && Cn.type == L_POOL && Cn.pool_max && Cn.kernel == 3 && Cn.stride == 2 &&
                       L.local_size == 5 && si.C % 8 == 0 && (size_t)9 * 17 * si.C * 2 <= 200 * 1024 &&
                       !fuse_second.count((int)i)) {
                fuse_first[(int)i] = ci; fuse_second[ci] = (int)i; virt.insert(top);
            }
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

    // ---- pass 2b: tensor-core eligibility, horizontal groups, directly written heads (bf16 plan)
    std::map<std::string, std::string> drop_alias;
    for (size_t i = 0; i < def.layers.size(); ++i)
        if (def.layers[i].type == L_DROPOUT && def.layers[i].bottom[0] != def.layers[i].top[0])
            drop_alias[def.layers[i].top[0]] = def.layers[i].bottom[0];
    struct SliceOf {
        std::map<std::string, Alias>& alias; std::map<std::string, std::string>& drop; std::map<std::string, Shape>& shp;
        void operator()(std::string nm, int& off, int& ldc, std::string& root) const {
            off = 0;
            for (;;) {
                if (drop.count(nm)) nm = drop[nm];
                else if (alias.count(nm)) { off += alias[nm].off; nm = alias[nm].root; }
                else break;
            }
            ldc = shp[nm].C; root = nm;
        }
    } slice_of = {alias, drop_alias, shp};
    std::map<int, int> sig_of;                   // conv layer -> Sigmoid layer folded into its epilogue
    std::map<int, std::string> direct_blob;      // conv layer -> network output it writes as NCHW float32
    std::set<int> sig_fused;                     // Sigmoid layers folded away
    std::set<std::string> direct_outputs;        // outputs that need no export step
    std::map<int, std::vector<int> > group;      // leader conv layer -> all members (leader first)
    std::map<int, int> group_leader;             // non-leader member -> leader
    std::set<int> tc_layers;                     // conv layers that run on tcgen05
    if (bf) {
        for (size_t i = 0; i < def.layers.size(); ++i) {
            const LayerDef& L = def.layers[i];
            if (L.type != L_CONV) continue;
            const std::string& top = L.top[0];
            if (s2d_conv.count((int)i)) { tc_layers.insert((int)i); continue; }
            if (L.stride != 1 || 2 * L.pad != L.kernel - 1) continue;
            if (L.bottom[0] == def.input_name || virt.count(L.bottom[0]) || dt[L.bottom[0]] != DT_BF16) continue;
            int off, ldc; std::string root;
            slice_of(L.bottom[0], off, ldc, root);
            if (shp[L.bottom[0]].C % 8 || off % 8 || ldc % 8) continue;
            // heads: a convolution that is a network output, or feeds only a Sigmoid that is one
            const bool relu_only = !inplace_ops.count(top) ||
                                   (inplace_ops[top] == 1 && i + 1 < def.layers.size() && def.layers[i + 1].type == L_RELU &&
                                    def.layers[i + 1].top[0] == top);
            if (fuse && relu_only && outputs.count(top)) {
                direct_blob[(int)i] = top;
            } else if (fuse && !inplace_ops.count(top) && consumers.count(top) && consumers[top].size() == 1) {
                const LayerDef& S = def.layers[consumers[top][0]];
                if (S.type == L_SIGMOID && S.top[0] != top && outputs.count(S.top[0]) && !inplace_ops.count(S.top[0])) {
                    sig_of[(int)i] = consumers[top][0];
                    direct_blob[(int)i] = S.top[0];
                }
            }
            if (!direct_blob.count((int)i)) {
                if (dt[top] != DT_BF16) continue;            // fp32 NHWC outputs stay on the SIMT kernel (unfused plan only)
                slice_of(top, off, ldc, root);
                if (shp[top].C % 8 || off % 8 || ldc % 8) continue;
            }
            tc_layers.insert((int)i);
        }
        for (std::map<int, int>::iterator it = sig_of.begin(); it != sig_of.end(); ++it) {
            sig_fused.insert(it->second);
            virt.insert(def.layers[it->first].top[0]);
        }
        for (std::map<int, std::string>::iterator it = direct_blob.begin(); it != direct_blob.end(); ++it)
            direct_outputs.insert(it->second);
        if (fuse) {
            for (std::set<int>::iterator it = tc_layers.begin(); it != tc_layers.end(); ++it) {
                const int i = *it;
                if (group_leader.count(i) || s2d_conv.count(i)) continue;
                const LayerDef& L = def.layers[i];
                std::vector<int> mem(1, i);
                for (std::set<int>::iterator jt = tc_layers.upper_bound(i); jt != tc_layers.end() && mem.size() < 4; ++jt) {
                    const LayerDef& M = def.layers[*jt];
                    if (group_leader.count(*jt) || group.count(*jt) || s2d_conv.count(*jt)) continue;
                    if (M.bottom[0] != L.bottom[0] || M.kernel != L.kernel || M.pad != L.pad) continue;
                    if (direct_blob.count(*jt) != direct_blob.count(i)) continue;
                    {   // the epilogue applies one ReLU flag to the whole launch
                        const bool ri = (size_t)i + 1 < def.layers.size() && def.layers[i + 1].type == L_RELU && def.layers[i + 1].top[0] == L.top[0];
                        const bool rj = (size_t)*jt + 1 < def.layers.size() && def.layers[*jt + 1].type == L_RELU && def.layers[*jt + 1].top[0] == M.top[0];
                        if (ri != rj) continue;
                    }
                    mem.push_back(*jt);
                }
                if (mem.size() > 1) {
                    group[i] = mem;
                    for (size_t m = 1; m < mem.size(); ++m) group_leader[mem[m]] = i;
                }
            }
        }
    }

    // ---- pass 2c: max-pool 3x3 / stride 1 / pad 1 feeding exactly one 1x1 tensor-core convolution (the `pool ->
    // pool_proj` branch of every inception module, deploy.prototxt:331-377): the pool runs on the convolution's A-operand
    // path (conv_tc2.cuh POOL mode) and its blob is never materialised.  Zero fill replaces the clipped windows, which
    // is only the same thing when the input is non-negative.
    std::map<int, int> pool_of;                  // conv layer -> pool layer folded into it
    std::set<int> pool_fused;                    // pool layers folded away
    if (bf && fuse && !getenv("FCNOD_NO_POOLPROJ")) {
        for (size_t i = 0; i < def.layers.size(); ++i) {
            const LayerDef& L = def.layers[i];
            if (L.type != L_POOL || !L.pool_max || L.kernel != 3 || L.stride != 1 || L.pad != 1) continue;
            const std::string& top = L.top[0];
            if (L.bottom[0] == top || inplace_ops.count(top) || outputs.count(top) || virt.count(top) || alias.count(top)) continue;
            if (!consumers.count(top) || consumers[top].size() != 1) continue;
            const int ci = consumers[top][0];
            const LayerDef& Cn = def.layers[ci];
            if (Cn.type != L_CONV || Cn.kernel != 1 || Cn.stride != 1 || Cn.pad != 0 || Cn.bottom.size() != 1) continue;
            if (!tc_layers.count(ci) || group.count(ci) || group_leader.count(ci) || direct_blob.count(ci)) continue;
            const std::string& src = L.bottom[0];
            if (!nonneg[src] || virt.count(src) || dt[src] != DT_BF16 || src == def.input_name) continue;
            int off, ldc; std::string root;
            slice_of(src, off, ldc, root);
            if (shp[src].C % 8 || off % 8 || ldc % 8) continue;
            pool_of[ci] = (int)i; pool_fused.insert((int)i); virt.insert(top);
        }
    }

    // ---- pass 2d: Convolution (+ReLU) -> LRN(5, beta 0.75) with the whole channel extent in one N tile (conv2/3x3 ->
    // conv2/norm2, deploy.prototxt:108-151): the LRN runs in the convolution's epilogue, the un-normalised blob is never
    // materialised.  Takes precedence over the LRN -> Pool fusion, which then reverts to a plain pooling launch.
    std::map<int, int> lrn_of;                   // conv layer -> LRN layer folded into its epilogue
    std::set<int> lrn_fused;
    if (bf && fuse && !getenv("FCNOD_NO_CONV_LRN")) {
        for (size_t i = 0; i < def.layers.size(); ++i) {
            const LayerDef& R = def.layers[i];
            if (R.type != L_LRN || R.local_size != 5 || R.beta != 0.75f || R.bottom[0] == R.top[0]) continue;
            const std::string& src = R.bottom[0];
            if (!producer.count(src) || outputs.count(src) || virt.count(src) || alias.count(src)) continue;
            const int ci = producer[src];
            const LayerDef& Cn = def.layers[ci];
            if (Cn.type != L_CONV || !tc_layers.count(ci) || group.count(ci) || group_leader.count(ci) || direct_blob.count(ci) ||
                s2d_conv.count(ci) || pool_of.count(ci)) continue;
            if (Cn.num_output > 192 || Cn.num_output % 16) continue;          // <= 48 columns per epilogue warp
            if ((Cn.kernel != 1 && Cn.kernel != 3) || shp[Cn.bottom[0]].C <= 32) continue;      // the instantiations that exist: CK = 64
            if (!consumers.count(src) || consumers[src].size() != 1 || consumers[src][0] != (int)i) continue;
            // exactly one in-place layer on the convolution's top, and it is the ReLU right behind it
            if (!inplace_ops.count(src) || inplace_ops[src] != 1 || (size_t)ci + 1 >= def.layers.size() ||
                def.layers[ci + 1].type != L_RELU || def.layers[ci + 1].top[0] != src) continue;
            if (inplace_ops.count(R.top[0]) || outputs.count(R.top[0]) || alias.count(R.top[0]) || dt[R.top[0]] != DT_BF16) continue;
            // undo an LRN -> Pool fusion of this LRN: the pool becomes a plain launch again
            if (fuse_first.count((int)i)) {
                const int pl = fuse_first[(int)i];
                fuse_second.erase(pl); fuse_first.erase((int)i);
                virt.erase(R.top[0]);
            }
            if (fuse_second.count((int)i)) continue;        // Pool -> LRN pairs are left alone
            lrn_of[ci] = (int)i; lrn_fused.insert((int)i); virt.insert(src);
        }
    }

    // ---- create blobs + buffers in definition order
    net->blobs.clear(); net->blob_index.clear(); net->buffers.clear();
    struct Maker {
        fcnod_net* net; std::map<std::string, Shape>& shp; std::map<std::string, int>& dt;
        std::map<std::string, Alias>& alias; std::set<std::string>& outputs; std::set<std::string>& virt;
        int make(const std::string& nm) {
            if (net->blob_index.count(nm)) return net->blob_index[nm];
            Blob b; b.name = nm; Shape s = shp[nm]; b.C = s.C; b.H = s.H; b.W = s.W; b.dtype = dt[nm];
            if (virt.count(nm)) {
                b.buf = -2; b.c_off = 0; b.ldc = s.C;
            } else if (alias.count(nm)) {
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
    } mk = {net, shp, dt, alias, outputs, virt};

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
                if (s2d_power.count((int)i)) {
                    // folded into the following stride-2 convolution: emit the space-to-depth transform only
                    const LayerDef& Cn = def.layers[s2d_power[(int)i]];
                    const Shape si = shp[def.input_name];
                    const int q = Cn.pad + (Cn.pad & 1);
                    int ka = (q - Cn.pad + Cn.kernel + 1) / 2;
                    ka = ka <= 1 ? 1 : (ka <= 2 ? 2 : 4);
                    mk.make(top);   // virtual
                    const S2dGeom sg = s2d_geom(si.H / 2, si.W / 2, ka);
                    Buffer bu; bu.C = kS2dCpad; bu.H = sg.Hp; bu.W = sg.Wp; bu.dtype = DT_BF16;
                    net->buffers.push_back(bu);
                    st.kind = S_S2D; st.in = 0; st.out = -1;
                    st.s2d_buf = (int)net->buffers.size() - 1; st.s2d_wp = sg.Wp; st.s2d_hp = sg.Hp; st.s2d_padl = q / 2;
                    st.desc = "s2d " + L.name + " (Power folded into " + Cn.name + ")";
                    net->steps.push_back(st);
                    break;
                }
                st.kind = S_TRANSFORM; st.in = 0; st.out = mk.make(top);
                st.desc = "transform " + L.name;
                net->steps.push_back(st);
                break;
            }
            case L_CONV: {
                if (group_leader.count((int)i)) break;      // emitted with its group leader
                const bool pooled = pool_of.count((int)i) > 0;
                st.in = blob_id(net, pooled ? def.layers[pool_of[(int)i]].bottom[0] : L.bottom[0]);
                const bool is_s2d = s2d_conv.count((int)i) > 0;
                const bool tc_ok = tc_layers.count((int)i) > 0;
                FCNOD_CHECK(is_s2d || net->blobs[st.in].buf >= 0, "Convolution directly on the NCHW input is not supported; insert a Power layer ('" + L.name + "')");
                std::vector<int> members(1, (int)i);
                if (group.count((int)i)) members = group[(int)i];
                const int Cin = net->blobs[st.in].C;
                ConvWeights cw; cw.Cin = Cin; cw.k = L.kernel; cw.stride = L.stride; cw.pad = L.pad; cw.Cout = 0;
                cw.pool = pooled;
                std::string names;
                for (size_t m = 0; m < members.size(); ++m) {
                    const int li = members[m];
                    const LayerDef& M = def.layers[li];
                    const std::string& mtop = M.top[0];
                    const bool relu = (size_t)li + 1 < def.layers.size() && def.layers[li + 1].type == L_RELU &&
                                      def.layers[li + 1].bottom[0] == mtop && def.layers[li + 1].top[0] == mtop;
                    WeightMap::const_iterator wi = weights.find(M.name);
                    FCNOD_CHECK(wi != weights.end(), "caffemodel has no blobs for layer '" + M.name + "'");
                    const std::vector<HostBlob>& wb = wi->second;
                    const size_t per = (size_t)Cin * cw.k * cw.k, wn = (size_t)M.num_output * per;
                    FCNOD_CHECK(wb.size() >= 1 && wb[0].data.size() == wn, "weight blob of '" + M.name + "' has the wrong size");
                    if (M.bias_term)
                        FCNOD_CHECK(wb.size() >= 2 && wb[1].data.size() == (size_t)M.num_output, "bias blob of '" + M.name + "' has the wrong size");
                    Step::Seg sg;
                    sg.name = M.name;
                    sg.blob = mk.make(mtop);
                    if (lrn_of.count(li)) {       // the epilogue normalises: what is written is the LRN's top
                        const LayerDef& R = def.layers[lrn_of[li]];
                        sg.blob = mk.make(R.top[0]);
                        cw.lrn = true; cw.lrn_aon = R.alpha / R.local_size; cw.lrn_k = R.lrn_k;
                    }
                    sg.direct_blob = -1;
                    sg.col0 = (cw.Cout + 7) / 8 * 8;
                    sg.ncols = M.num_output;
                    sg.flags = relu ? SEG_RELU : 0;
                    if (direct_blob.count(li)) {
                        sg.direct_blob = mk.make(direct_blob[li]);
                        sg.flags |= SEG_DIRECT;
                        if (sig_of.count(li)) sg.flags |= SEG_SIGMOID;
                    }
                    cw.w.resize((size_t)sg.col0 * per, 0.f);
                    cw.b.resize(sg.col0, 0.f);
                    cw.w.insert(cw.w.end(), wb[0].data.begin(), wb[0].data.end());
                    if (M.bias_term) cw.b.insert(cw.b.end(), wb[1].data.begin(), wb[1].data.end());
                    else cw.b.resize(sg.col0 + M.num_output, 0.f);
                    cw.Cout = sg.col0 + M.num_output;
                    st.segs.push_back(sg);
                    names += (m ? " + " : "") + M.name;
                }
                st.out = st.segs[0].blob;
                st.relu = (st.segs[0].flags & SEG_RELU) != 0;
                if (is_s2d) {
                    const LayerDef& P = def.layers[s2d_conv[(int)i]];
                    cw.s2d = true; cw.in_scale = P.scale; cw.in_shift = P.shift;
                    cw.q = L.pad + (L.pad & 1);
                    int ka = (cw.q - L.pad + L.kernel + 1) / 2;
                    cw.ka = ka <= 1 ? 1 : (ka <= 2 ? 2 : 4);
                    for (size_t sj = 0; sj < net->steps.size(); ++sj)
                        if (net->steps[sj].kind == S_S2D && net->steps[sj].layer == s2d_conv[(int)i]) {
                            st.s2d_buf = net->steps[sj].s2d_buf; st.s2d_wp = net->steps[sj].s2d_wp;
                            st.s2d_padl = net->steps[sj].s2d_padl; st.s2d_hp = net->steps[sj].s2d_hp;
                        }
                    FCNOD_CHECK(st.s2d_buf >= 0, "internal: no space-to-depth buffer for '" + L.name + "'");
                    st.s2d_overlap = true;
                }
                st.kind = tc_ok ? S_CONV_TC : S_CONV_SIMT;
                if (tc_ok) {
                    choose_tc_tiling(cw);
                    for (size_t m = 0; m < st.segs.size(); ++m) {
                        if (st.segs[m].flags & SEG_DIRECT) continue;
                        const Blob& sb = net->blobs[st.segs[m].blob];
                        FCNOD_CHECK(sb.dtype == DT_BF16 && sb.C % 8 == 0 && sb.ldc % 8 == 0 && sb.c_off % 8 == 0,
                                    "tensor-core convolution '" + st.segs[m].name + "': output slice is not bf16 / 16-byte aligned");
                    }
                }
                net->convs.push_back(cw);
                st.conv = (int)net->convs.size() - 1;
                std::ostringstream d;
                d << (tc_ok ? "conv_tc " : "conv_simt ") << names << " " << L.kernel << "x" << L.kernel << "/" << L.stride
                  << " Cin=" << cw.Cin << " Cout=" << cw.Cout << (st.relu ? " +relu" : "") << (is_s2d ? " +s2d+power" : "");
                if (pooled) d << " +maxpool3x3(" << def.layers[pool_of[(int)i]].name << ")";
                if (lrn_of.count((int)i)) d << " +lrn(" << def.layers[lrn_of[(int)i]].name << ")";
                for (size_t m = 0; m < st.segs.size(); ++m)
                    if (st.segs[m].flags & SEG_DIRECT) d << ((st.segs[m].flags & SEG_SIGMOID) ? " +sigmoid" : "") << " +nchw";
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
                if (sig_fused.count((int)i)) { mk.make(top); break; }     // folded into the producing convolution
                st.kind = S_UNARY; st.unary_op = 1; st.in = blob_id(net, L.bottom[0]); st.out = mk.make(top);
                st.desc = "sigmoid " + L.name;
                net->steps.push_back(st);
                break;
            }
            case L_POOL: {
                if (pool_fused.count((int)i)) { mk.make(top); break; }      // runs inside its 1x1 consumer (POOL mode)
                if (fuse_first.count((int)i)) { mk.make(top); break; }      // Pool -> LRN: emitted at the LRN
                if (fuse_second.count((int)i)) {                            // LRN -> Pool
                    const LayerDef& F = def.layers[fuse_second[(int)i]];
                    st.kind = S_LRN_POOL; st.layer = fuse_second[(int)i]; st.layer2 = (int)i;
                    st.in = blob_id(net, F.bottom[0]); st.out = mk.make(top);
                    st.desc = "lrn+pool " + F.name + " + " + L.name;
                    net->steps.push_back(st);
                    break;
                }
                st.kind = S_POOL; st.in = blob_id(net, L.bottom[0]); st.out = mk.make(top);
                FCNOD_CHECK(net->blobs[st.in].C % 8 == 0, "Pooling '" + L.name + "' needs channels % 8 == 0");
                st.desc = "pool " + L.name;
                net->steps.push_back(st);
                break;
            }
            case L_LRN: {
                if (lrn_fused.count((int)i)) { mk.make(top); break; }       // runs in its convolution's epilogue
                if (fuse_first.count((int)i)) { mk.make(top); break; }      // LRN -> Pool: emitted at the Pool
                if (fuse_second.count((int)i)) {                            // Pool -> LRN
                    const LayerDef& F = def.layers[fuse_second[(int)i]];
                    st.kind = S_POOL_LRN; st.layer = fuse_second[(int)i]; st.layer2 = (int)i;
                    st.in = blob_id(net, F.bottom[0]); st.out = mk.make(top);
                    st.desc = "pool+lrn " + F.name + " + " + L.name;
                    net->steps.push_back(st);
                    break;
                }
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
            if (direct_outputs.count(net->blobs[i].name)) continue;   // written as NCHW float32 by its convolution
            Step st; st.kind = S_EXPORT; st.in = (int)i; st.desc = "export " + net->blobs[i].name;
            net->steps.push_back(st);
        }
    plan_lanes(net);
    net->fused_prelude = !net->steps.empty() && net->steps[0].kind == S_S2D && net->blobs[0].C == 3 &&
                         net->blobs[0].H % 2 == 0 && net->blobs[0].W % 2 == 0 && !getenv("FCNOD_NO_PRELUDE_FUSION");
}

// ------------------------------------------------------------------------------------ lanes
// The layer graph is not a chain: inside an inception module (deploy.prototxt:163-377) the 3x3 and the 5x5 convolution
// only need their own reduce outputs, and `pool -> pool_proj` only needs the module's input.  Every launch is a
// persistent grid whose last round leaves most SMs idle (448 tiles on 148 SMs = 3.03 rounds); launches that do not
// depend on each other go to different streams ("lanes") so that the block scheduler fills one kernel's tail -- and its
// prologue -- with the CTAs of the next.  Dependencies are derived from the buffer regions every step reads and writes
// (RAW, WAR and WAW); cross-lane edges become event waits, which stream capture turns into graph edges.
struct Region { int buf, c0, c1; };
bool overlaps(const Region& a, const Region& b) { return a.buf == b.buf && a.c0 < b.c1 && b.c0 < a.c1; }

void plan_lanes(fcnod_net* net) {
    const int kLanes = 3, kOutBase = 1 << 20;
    const size_t n = net->steps.size();
    static const bool off = getenv("FCNOD_LANES") && atoi(getenv("FCNOD_LANES")) <= 1;
    std::vector<std::vector<Region> > rd(n), wr(n);
    auto blob_region = [&](int bi) {
        const Blob& b = net->blobs[bi];
        Region r = {b.buf, b.c_off, b.c_off + b.C};      // buf -1 = the NCHW input
        return r;
    };
    for (size_t i = 0; i < n; ++i) {
        const Step& st = net->steps[i];
        switch (st.kind) {
            case S_S2D: rd[i].push_back(blob_region(0)); wr[i].push_back(Region{st.s2d_buf, 0, 1 << 30}); break;
            case S_CONV_SIMT: case S_CONV_TC:
                if (net->convs[st.conv].s2d) rd[i].push_back(Region{st.s2d_buf, 0, 1 << 30});
                else rd[i].push_back(blob_region(st.in));
                for (size_t k = 0; k < st.segs.size(); ++k) {
                    if (st.segs[k].flags & SEG_DIRECT) wr[i].push_back(Region{kOutBase + st.segs[k].direct_blob, 0, 1 << 30});
                    else wr[i].push_back(blob_region(st.segs[k].blob));
                }
                break;
            case S_COPY: {
                rd[i].push_back(blob_region(st.in));
                const Blob& ob = net->blobs[st.out];
                wr[i].push_back(Region{ob.buf, ob.c_off + st.unary_op, ob.c_off + st.unary_op + net->blobs[st.in].C});
                break;
            }
            case S_EXPORT: rd[i].push_back(blob_region(st.in)); wr[i].push_back(Region{kOutBase + st.in, 0, 1 << 30}); break;
            default: rd[i].push_back(blob_region(st.in)); wr[i].push_back(blob_region(st.out)); break;
        }
    }
    int last_on[kLanes];
    for (int l = 0; l < kLanes; ++l) last_on[l] = -1;
    net->n_lanes = 1;
    for (size_t j = 0; j < n; ++j) {
        Step& st = net->steps[j];
        st.lane = 0; st.waits.clear(); st.signal = false; st.deps.clear();
        std::vector<int> deps;
        for (size_t i = 0; i < j; ++i) {
            bool dep = false;
            for (size_t a = 0; a < wr[i].size() && !dep; ++a) {
                for (size_t b = 0; b < rd[j].size() && !dep; ++b) dep = overlaps(wr[i][a], rd[j][b]);
                for (size_t b = 0; b < wr[j].size() && !dep; ++b) dep = overlaps(wr[i][a], wr[j][b]);
            }
            for (size_t a = 0; a < rd[i].size() && !dep; ++a)
                for (size_t b = 0; b < wr[j].size() && !dep; ++b) dep = overlaps(rd[i][a], wr[j][b]);
            if (dep) deps.push_back((int)i);
        }
        st.deps = deps;
        if (off) continue;
        int lane = 0;
        if (!deps.empty()) {
            // continue a chain: among the dependencies that are still the last step of their lane, the one on the lowest
            // (= highest-priority) lane.  Lane 0 thus carries fused-1x1 -> 3x3 -> next module's fused-1x1 -> ... (the
            // critical chain); the 5x5 and pool_proj branches hang off it on lanes 1 and 2.
            lane = -1;
            for (size_t d = 0; d < deps.size(); ++d) {
                const int l = net->steps[deps[d]].lane;
                if (last_on[l] == deps[d] && (lane < 0 || l < lane)) lane = l;
            }
            if (lane < 0) {                                    // a branch: the lane that has been idle the longest
                lane = -1;
                for (int l = 0; l < kLanes; ++l)
                    if (l != net->steps[j - 1].lane && (lane < 0 || last_on[l] < last_on[lane])) lane = l;
            }
        }
        st.lane = lane;
        // one wait per other lane: its newest dependency (events of a lane complete in order)
        for (int l = 0; l < kLanes; ++l) {
            if (l == lane) continue;
            int newest = -1;
            for (size_t d = 0; d < deps.size(); ++d)
                if (net->steps[deps[d]].lane == l) newest = deps[d];
            if (newest < 0) continue;
            // already implied by an earlier step of this lane that waited for it (or for something newer on lane l)?
            bool implied = false;
            for (size_t i = 0; i < j && !implied; ++i)
                if (net->steps[i].lane == lane)
                    for (size_t w = 0; w < net->steps[i].waits.size(); ++w)
                        if (net->steps[net->steps[i].waits[w]].lane == l && net->steps[i].waits[w] >= newest) implied = true;
            if (!implied) { st.waits.push_back(newest); net->steps[newest].signal = true; }
        }
        last_on[lane] = (int)j;
        net->n_lanes = std::max(net->n_lanes, lane + 1);
    }
}

// ------------------------------------------------------------------------------------ tcgen05 tiling (host only)
const size_t kTcSmemBudget = 225 * 1024;

// K layout of the packed weights (independent of the batch): channel-chunk width, HALO mode.
void choose_tc_tiling(ConvWeights& cw) {
    if (cw.s2d) {   // K-block = one tap of the ka x ka convolution on the space-to-depth image: 16 channels = 32-byte rows
        cw.CK = kS2dCpad;
        cw.n_cchunks = 1;
        cw.halo = true;
        return;
    }
    if (cw.pool) {  // POOL mode pools 64-channel chunks (128-byte swizzled rows on both sides of the pooling warps)
        cw.halo = false;
        cw.CK = 64;
        cw.n_cchunks = (cw.Cin + 63) / 64;
        return;
    }
    const char* nh = getenv("FCNOD_NO_HALO");
    cw.halo = cw.k > 1 && !(nh && nh[0] == '1');
    if (cw.halo) {
        // HALO mode shifts the A descriptor by whole patch pixels of CK channels (128-, 64- or 32-byte swizzled rows):
        // narrow inputs (the 5x5 branches: 16 .. 32 channels) take the narrowest block that holds them in one chunk
        cw.CK = cw.Cin <= 16 ? 16 : (cw.Cin <= 32 ? 32 : 64);
    } else {
        // K-block width: minimise padded K with a small per-block cost
        int best = 64; long best_cost = -1;
        const int cand[3] = {64, 32, 16};
        for (int i = 0; i < 3; ++i) {
            int ck = cand[i];
            long cost = (long)((cw.Cin + ck - 1) / ck) * (ck + 8);
            if (best_cost < 0 || cost < best_cost) { best_cost = cost; best = ck; }
        }
        cw.CK = best;
    }
    cw.n_cchunks = (cw.Cin + cw.CK - 1) / cw.CK;
}

// Everything of ConvTc2Params that does not need device pointers: pixel box, N tiling, pipeline depth.
// Pixel tilings of one launch: the spatial box (BW x BH pixels of one image, halo patches in HALO mode) and, for plain
// 1x1 convolutions, the FLAT tiling: the pixel index runs over batch * H * W and a tile is any 128 consecutive pixels.
// A 28x28 image holds 7 boxes of 28x4 = 112 pixels (448 tiles per 64 frames = 3.03 rounds over 148 SMs); flat, the same
// 50,176 pixels are 392 full 128-row tiles (2.65 rounds, 3 instead of 4).
struct TcGeo { int flat, H, W, BW, PW, BH, tiles_x, tiles_y, m_tiles; uint32_t a_tx, a_stage; };

void fill_tc_geometry(const fcnod_net* net, Step& st, int batch, int pick = -1) {
    const ConvWeights& cw = net->convs[st.conv];
    const Blob& ob = net->blobs[st.out];
    ConvTc2Params& p = st.tc;
    memset(&p, 0, sizeof(p));
    p.kh = p.kw = cw.k; p.pad_h = p.pad_w = cw.pad;
    p.n_cchunks = cw.n_cchunks; p.CK = cw.CK;
    p.halo = cw.halo ? 1 : 0;
    p.pool = cw.pool ? 1 : 0;
    if (cw.pool) p.pad_h = p.pad_w = 1;    // the raw patch starts one pixel up / left of the tile
    if (cw.s2d) {
        p.kh = p.kw = cw.ka; p.CK = kS2dCpad;                // HALO mode on the space-to-depth image, one K step per tap
        p.n_cchunks = 1;
        p.pad_h = 0; p.pad_w = 0;          // the padding is physical (zero rows / columns of the space-to-depth image)
        p.a_bulk = 1;                      // A patch rows arrive by cp.async.bulk from the pre-swizzled image
        p.bt_lo_h = cw.bt_lo_h; p.bt_hi_h = cw.bt_hi_h; p.bt_lo_w = cw.bt_lo_w; p.bt_hi_w = cw.bt_hi_w;
    }
    p.ksteps_last = p.CK / 16;
    if (!cw.s2d) p.ksteps_last = std::min(p.CK / 16, ceil_div(cw.Cin - (p.n_cchunks - 1) * p.CK, 16));
    const uint32_t span = p.CK * 2;
    const int halo_w = p.halo ? p.kw - 1 : 0, halo_h = p.halo ? p.kh - 1 : 0;
    std::vector<TcGeo> geos;
    {   // spatial box: the fewest boxes per image row, as wide as they can be made equal (224 -> 2 x 112, not 128 + 96)
        TcGeo g;
        g.flat = 0; g.H = ob.H; g.W = ob.W;   // == input size for the stride-1 "same" convolutions; == s2d image size for s2d
        // POOL mode: boxes at most 30 pixels wide -- one shuffle pass of the pooling warps (32 lanes load 32 patch
        // columns, 30 come out) and a squarer box, i.e. less halo: 28x4 needs a 30x6 patch (1.6x), 56x2 a 58x4 one (2.1x)
        const int bw_max = cw.pool ? 30 : 128 - halo_w;
        g.BW = ceil_div(ob.W, ceil_div(ob.W, bw_max));
        g.PW = g.BW + halo_w;
        g.BH = std::max(1, std::min(ob.H, 128 / g.PW));
        g.tiles_x = ceil_div(ob.W, g.BW);
        g.tiles_y = ceil_div(ob.H, g.BH);
        if (cw.s2d) {
            const S2dGeom sg = s2d_geom(ob.H, ob.W, cw.ka);
            g.BW = sg.BW; g.PW = sg.PW; g.BH = sg.BH; g.tiles_x = sg.tiles_x; g.tiles_y = sg.tiles_y;
        }
        g.m_tiles = g.tiles_x * g.tiles_y * batch;
        if (p.halo) {
            g.a_tx = (uint32_t)(g.PW * (g.BH + halo_h)) * span;
            const int rows = std::max(128 + halo_h * g.PW + halo_w, g.PW * (g.BH + halo_h));
            g.a_stage = (uint32_t)align_up((size_t)rows * span, 1024);
        } else {
            g.a_tx = (uint32_t)(g.BW * g.BH) * span;
            g.a_stage = 128 * span;
        }
        geos.push_back(g);
    }
    static const bool flat_off = getenv("FCNOD_FLAT") && getenv("FCNOD_FLAT")[0] == '0';
    const long npix = (long)batch * ob.H * ob.W;
    if (!flat_off && !p.halo && !cw.s2d && !cw.pool && cw.k == 1 && cw.pad == 0 && !(st.segs[0].flags & SEG_DIRECT) &&
        npix < (1L << 30) && geos[0].BW * geos[0].BH < 128) {
        TcGeo g;
        g.flat = 1; g.H = 1; g.W = (int)npix;
        g.BW = g.PW = 128; g.BH = 1;
        g.tiles_x = ceil_div((int)npix, 128); g.tiles_y = 1; g.m_tiles = g.tiles_x;
        g.a_tx = 128 * span; g.a_stage = 128 * span;
        geos.push_back(g);
    }
    const int taps = p.kh * p.kw;
    const int Kp = taps * p.n_cchunks * p.CK;
    const int sms = st.grid_cap > 0 ? st.grid_cap : kNumSMs;       // CTAs this launch may occupy
    // N tiling / pipeline shape: every candidate that fits in shared memory, ranked by a per-tile model
    //   tile time = max(MMA issue, smem bandwidth, L2->smem bytes, epilogue) + fixed;  total = rounds over the persistent CTAs
    // (constants measured with tools/micro/mma_rate.cu and the conv timelines).  `pick` selects the pick-th best; the
    // device autotuner (autotune_tc_steps) times the first few and keeps the fastest.
    const int N8 = (cw.Cout + 7) / 8 * 8;
    const char* nres = getenv("FCNOD_NO_RESIDENT");
    const bool no_resident = nres && nres[0] == '1';
    const char* de = getenv("FCNOD_DUAL");
    const int dual_bn = de ? (de[0] == '0' ? 0 : (de[0] == '1' ? 256 : atoi(de))) : 256;
    const double ksteps_tile = (double)taps * ((p.n_cchunks - 1) * (p.CK / 16) + p.ksteps_last);   // real 16-wide K steps
    const size_t fixed = (size_t)kTcEpiWarps * 256 + 1024 + 1024;     // bias staging, barriers, alignment slack
    struct Cand { int geo, nt, BN, a_st, b_st, bt, resident, dual, r_st; uint32_t b_stage; double cost; };
    std::vector<Cand> cands;
    // POOL mode: raw halo patches of (BW+2) x (BH+2) pixels x 64 channels in a ring of their own (2..4 stages); the eight
    // remaining epilogue warps drain at most 128 columns; one issuing thread
    const uint32_t r_tx = cw.pool ? (uint32_t)((geos[0].BW + 2) * (geos[0].BH + 2)) * span : 0;
    const uint32_t r_stage = (uint32_t)align_up(r_tx, 1024);
    for (int r_st = cw.pool ? 2 : 0; r_st <= (cw.pool ? 6 : 0); ++r_st)
    for (size_t gi = 0; gi < geos.size(); ++gi) {
      const TcGeo& G = geos[gi];
      if ((size_t)r_st * r_stage + 2 * (size_t)G.a_stage > kTcSmemBudget - fixed) continue;
      const size_t avail = kTcSmemBudget - fixed - (size_t)r_st * r_stage;
      for (int nt = ceil_div(N8, cw.pool ? 128 : 256); nt <= (cw.lrn ? 1 : ceil_div(N8, 16)); ++nt) {
        const int BN = ceil_div(ceil_div(N8, nt), 16) * 16;
        if (BN > (cw.pool ? 128 : 256) || nt * BN - N8 >= BN) continue;
        // a CTA owns one N tile: grid = a multiple of nt; pixel tiles are dealt to the grid / nt CTAs of each N tile
        // (ranked for the whole device whatever the launch's grid cap: a pick index must always mean the same tiling)
        const int mslots = std::max(1, std::min(G.m_tiles, kNumSMs / nt));
        const long rounds = ((long)G.m_tiles + mslots - 1) / mslots;
        for (int bt = (p.halo ? p.kw : 1); bt >= 1; bt = (bt > 1 ? 1 : 0)) {       // B stage = kw taps (HALO) or 1
            const uint32_t b_stage = (uint32_t)align_up((size_t)BN * span, 1024) * bt;
            const int rounds_b = taps * p.n_cchunks / bt;                               // B stages consumed per tile
            for (int resident = 0; resident <= (no_resident ? 0 : 1); ++resident)
                for (int dual = 0; dual <= ((BN <= dual_bn && G.m_tiles > 1 && !cw.pool) ? 1 : 0); ++dual) {
                    const int nr = dual ? 2 : 1;
                    int a_st = 0, b_st = 0;
                    if (resident) {
                        // the whole K extent of the N tile stays in shared memory (loaded once per CTA, shared by both rings)
                        const size_t bres = (size_t)rounds_b * b_stage;
                        const int a_min = p.halo ? 1 : 2;
                        if (rounds_b > 24 || bres + (size_t)nr * a_min * G.a_stage > avail) continue;
                        b_st = rounds_b;
                        // halo patches stream from HBM / L2 with ~1 us of latency and a tile consumes one in well under that:
                        // as many patch stages as shared memory allows (FCNOD_HALO_STAGES caps it; 3 / 4 before round 2)
                        static const int halo_cap = getenv("FCNOD_HALO_STAGES") ? atoi(getenv("FCNOD_HALO_STAGES")) : 0;
                        const size_t a_cap = p.halo ? (halo_cap > 0 ? (size_t)halo_cap : (dual ? 6 : 8)) : 8;
                        a_st = (int)std::min<size_t>(a_cap, (avail - bres) / ((size_t)nr * G.a_stage));
                        if (nr * a_st + b_st > 30) a_st = (30 - b_st) / nr;
                        if (a_st < a_min) continue;
                    } else if (p.halo) {
                        for (a_st = 2; a_st >= 1; --a_st) {
                            if (avail < (size_t)nr * a_st * G.a_stage + (size_t)nr * 2 * b_stage) continue;
                            b_st = (int)std::min<size_t>(bt > 1 ? 6 : 8, (avail - (size_t)nr * a_st * G.a_stage) / ((size_t)nr * b_stage));
                            break;
                        }
                        if (a_st < 1) continue;
                    } else {
                        const size_t pair = (size_t)G.a_stage + b_stage;
                        a_st = b_st = (int)std::min<size_t>(8, avail / (nr * pair));
                        if (a_st < 2) continue;
                    }
                    if (!resident && !dual && b_st >= 3) b_st &= ~1;     // two warps feed the B ring round-robin: even count
                    // one M=128 K=16 instruction: max(93, 43 + N/2) cycles from one issuing thread, ~max(68, 22 + N/2) from two
                    // (measured before the MMA warps issued converged through uniform registers, see conv_tc2.cuh; the issue
                    // floor is lower now, the N/2 tensor term is not -- the ranking only orders candidates for the autotuner);
                    // the one issuing thread also spends ~300 cycles per barrier round (wait + commit)
                    const double mma = ksteps_tile * (dual ? std::max(68.0, 22.0 + BN / 2.0) : std::max(93.0, 43.0 + BN / 2.0));
                    const double brounds = resident ? (double)(p.halo ? p.n_cchunks : taps * p.n_cchunks) : (double)rounds_b;
                    const double issue = (brounds * 300.0 + ksteps_tile * 25.0) / (dual ? 2.0 : 1.0);
                    const double a_bytes = p.halo ? (double)p.n_cchunks * G.a_tx : (double)taps * p.n_cchunks * G.a_tx;
                    const double b_bytes = resident ? 0.0 : (double)BN * Kp * 2;
                    const double load = (a_bytes + b_bytes) / 32.0;                            // ~32 B/clk/SM seen from L2
                    const double smem = (ksteps_tile * (4096.0 + 32.0 * BN) + a_bytes + b_bytes) / 128.0;
                    const double epi = 300.0 + 10.0 * BN;
                    double tile = std::max(std::max(std::max(mma, smem), issue), std::max(load, epi)) + 400.0;
                    if (a_st < 2 && !dual) tile += 0.5 * a_bytes / 32.0;                       // no A prefetch across tiles
                    if (cw.pool) {   // per 64-channel chunk: ~4.5 + 1 shared-memory 128-bit accesses per pooled 16-byte piece
                        const double pool = p.n_cchunks * ((double)r_tx + 5.5 * G.a_tx) / 128.0 + 300.0;
                        tile = std::max(tile, pool) + (r_st < 4 ? 0.1 * (4 - r_st) * pool : 0.0);
                    }
                    Cand c = {(int)gi, nt, BN, a_st, b_st, bt, resident, dual, r_st, b_stage, 0.0};
                    c.cost = rounds * tile + (resident ? (double)BN * Kp * 2 / 32.0 : 0.0);
                    cands.push_back(c);
                }
        }
      }
    }
    FCNOD_CHECK(!cands.empty(), "no tensor-core tiling fits in shared memory for '" + st.desc + "'");
    std::stable_sort(cands.begin(), cands.end(), [](const Cand& x, const Cand& y) { return x.cost < y.cost; });
    st.tc_ncand = (int)cands.size();
    {
        const Cand& c = cands[std::min<size_t>(pick < 0 ? 0 : pick, cands.size() - 1)];
        p.n_tiles = c.nt; p.BN = c.BN; p.a_stages = c.a_st; p.b_stages = c.b_st; p.b_taps = c.bt;
        p.b_stage_bytes = c.b_stage;
        p.b_resident = c.resident; p.dual = c.dual;
        p.r_stages = c.r_st; p.r_stage_bytes = r_stage; p.r_tx_bytes = r_tx;
        const TcGeo& G = geos[c.geo];
        p.flat = G.flat; p.H = G.H; p.W = G.W; p.BW = G.BW; p.PW = G.PW; p.BH = G.BH;
        p.tiles_x = G.tiles_x; p.tiles_y = G.tiles_y; p.m_tiles = G.m_tiles;
        p.a_tx_bytes = G.a_tx; p.a_stage_bytes = G.a_stage;
    }
    uint32_t acc = 32;
    while ((int)acc < p.BN) acc <<= 1;
    p.acc_stride = acc;
    // four accumulators when they fit the 512 TMEM columns; narrow N tiles are drained by 4 (or 8) warps each, alternate
    // tiles by alternate warp groups, instead of by all 16 (ConvTc2Params::egroups)
    p.acc_lg = 4 * acc <= 512 ? 2 : 1;
    static const int egroups_env = getenv("FCNOD_EGROUPS") ? atoi(getenv("FCNOD_EGROUPS")) : 0;
    p.egroups = p.BN <= 64 ? 4 : (p.BN <= 128 ? 2 : 1);
    if (egroups_env == 1 || egroups_env == 2) p.egroups = std::min<uint32_t>(p.egroups, (uint32_t)egroups_env);
    if (p.acc_lg == 1 && p.egroups > 2) p.egroups = 2;
    const int epi_per_quadrant = cw.pool ? 2 : 4;                    // POOL mode: eight of the sixteen warps pool
    if (cw.pool) p.egroups = p.BN <= 64 ? 2 : 1;
    if (cw.lrn) p.egroups = p.BN <= 48 ? 4 : (p.BN <= 96 ? 2 : 1);   // the LRN epilogue keeps <= 48 columns per warp in registers
    if (p.acc_lg == 1 && p.egroups > 2) p.egroups = 2;
    p.cols_half0 = ceil_div(p.BN, 16 * (epi_per_quadrant / (int)p.egroups)) * 16;   // columns per epilogue warp (<= 64)
    p.tmem_cols = std::max<uint32_t>((1u << p.acc_lg) * acc, 32);
    p.bt_ld = (int)align_up((size_t)N8, 256) + 256;
    st.tc_smem = (size_t)(p.dual ? 2 : 1) * p.a_stages * p.a_stage_bytes +
                 (size_t)((p.dual && !p.b_resident) ? 2 : 1) * p.b_stages * p.b_stage_bytes + (size_t)p.r_stages * p.r_stage_bytes +
                 (size_t)kTcEpiWarps * 256 + 1024 + 1024;
    p.mstride = std::max(1, std::min(p.m_tiles, std::max(sms, p.n_tiles) / p.n_tiles));
    st.tc_grid = p.mstride * p.n_tiles;
    {   // the pixel-tile stride in the mixed radix (img, ty, tx) of the tile index (TileIter::next)
        int g = p.mstride;
        p.step_tx = g % p.tiles_x; g /= p.tiles_x;
        p.step_ty = g % p.tiles_y; g /= p.tiles_y;
        p.step_img = g;
    }
    p.lrn = cw.lrn ? 1 : 0; p.lrn_aon = cw.lrn_aon; p.lrn_k = cw.lrn_k;
    p.nseg = (int)st.segs.size();
    p.flags = st.segs[0].flags & (SEG_RELU | SEG_DIRECT);
    p.bt_interior = cw.bt_lo_h * (cw.bt_lo_w + cw.bt_hi_w + 1) + cw.bt_lo_w;
    for (int i = 0; i < p.nseg; ++i) {
        FCNOD_CHECK((st.segs[i].flags & (SEG_RELU | SEG_DIRECT)) == p.flags, "internal: fused convolutions disagree on ReLU / direct output");
        p.seg[i].col0 = st.segs[i].col0; p.seg[i].ncols = st.segs[i].ncols; p.seg[i].flags = st.segs[i].flags;
    }
}

// ------------------------------------------------------------------------------------ weights
void upload_weights(fcnod_net* net) {
    for (size_t si = 0; si < net->steps.size(); ++si) {
        Step& st = net->steps[si];
        if (st.kind != S_CONV_SIMT && st.kind != S_CONV_TC) continue;
        ConvWeights& cw = net->convs[st.conv];
        const int k = cw.k, Cin = cw.Cin, Cout = cw.Cout;
        if (st.kind == S_CONV_SIMT) {
            FCNOD_CHECK(st.segs.size() == 1, "internal: grouped convolution on the SIMT path");
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
        } else if (cw.s2d) {
            // stride-2 convolution on the Power-transformed input as a stride-1 ka x ka convolution on the
            // 2x2 space-to-depth image:  ih = 2*oh - pad + r,  r = 2*a + dy - (q - pad),  q = pad rounded up to even.
            // GEMM weights = in_scale * w (bf16); the Power shift goes into the border-aware bias table
            // bias[class(oh)][class(ow)][o] = b[o] + in_shift * sum over the taps that fall INSIDE the image.
            const int ka = cw.ka, off = cw.q - cw.pad, Kp = ka * ka * kS2dCpad;
            const int Nrows = (Cout + 15) / 16 * 16, Np = (int)align_up((size_t)(Cout + 7) / 8 * 8, 256) + 256;
            // packed [K-block = tap ah * ka + aw][N][(dy*2+dx)*Cin + c]  (HALO order: the taps of a kernel row adjacent)
            std::vector<bf16> w((size_t)Nrows * Kp, __float2bfloat16(0.f));
            const size_t kbs = (size_t)Nrows * kS2dCpad;
            for (int o = 0; o < Cout; ++o)
                for (int ah = 0; ah < ka; ++ah)
                    for (int aw = 0; aw < ka; ++aw)
                        for (int dy = 0; dy < 2; ++dy)
                            for (int dx = 0; dx < 2; ++dx) {
                                const int r = 2 * ah + dy - off, sx = 2 * aw + dx - off;
                                if (r < 0 || r >= k || sx < 0 || sx >= k) continue;
                                for (int c = 0; c < Cin; ++c)
                                    w[(size_t)(ah * ka + aw) * kbs + (size_t)o * kS2dCpad + (dy * 2 + dx) * Cin + c] =
                                        __float2bfloat16(cw.in_scale * cw.w[(((size_t)o * Cin + c) * k + r) * k + sx]);
                            }
            FCNOD_CUDA(cudaMalloc(&cw.d_w_tc, w.size() * 2));
            FCNOD_CUDA(cudaMemcpy(cw.d_w_tc, w.data(), w.size() * 2, cudaMemcpyHostToDevice));
            std::vector<float> b(Np, 0.f);
            std::copy(cw.b.begin(), cw.b.end(), b.begin());
            FCNOD_CUDA(cudaMalloc(&cw.d_bias, b.size() * 4));
            FCNOD_CUDA(cudaMemcpy(cw.d_bias, b.data(), b.size() * 4, cudaMemcpyHostToDevice));
            const Blob& ibv = net->blobs[st.in];
            const Blob& obv = net->blobs[st.out];
            struct Cls { int lo, hi; };
            Cls ch, cwd;
            const int dimsI[2] = {ibv.H, ibv.W}, dimsO[2] = {obv.H, obv.W};
            Cls* cl[2] = {&ch, &cwd};
            for (int d = 0; d < 2; ++d) {
                cl[d]->lo = (cw.pad + 1) / 2;
                const int first_bad = (dimsI[d] + cw.pad - k + 1 + 1) / 2;     // first o with 2o - pad + k - 1 >= size
                cl[d]->hi = std::max(0, dimsO[d] - first_bad);
                FCNOD_CHECK(cl[d]->lo + cl[d]->hi < dimsO[d], "space-to-depth convolution: image too small for the border classes");
            }
            cw.bt_lo_h = ch.lo; cw.bt_hi_h = ch.hi; cw.bt_lo_w = cwd.lo; cw.bt_hi_w = cwd.hi;
            const int nrc = ch.lo + ch.hi + 1, ncc = cwd.lo + cwd.hi + 1;
            std::vector<float> tab((size_t)nrc * ncc * Np, 0.f);
            for (int rc = 0; rc < nrc; ++rc) {
                const int oh = rc < ch.lo ? rc : (rc == ch.lo ? ch.lo : obv.H - ch.hi + (rc - ch.lo - 1));
                for (int cc = 0; cc < ncc; ++cc) {
                    const int ow = cc < cwd.lo ? cc : (cc == cwd.lo ? cwd.lo : obv.W - cwd.hi + (cc - cwd.lo - 1));
                    for (int o = 0; o < Cout; ++o) {
                        double sum = 0.0;
                        for (int r = 0; r < k; ++r) {
                            const int ih = 2 * oh - cw.pad + r;
                            if (ih < 0 || ih >= ibv.H) continue;
                            for (int sx = 0; sx < k; ++sx) {
                                const int iw = 2 * ow - cw.pad + sx;
                                if (iw < 0 || iw >= ibv.W) continue;
                                for (int c = 0; c < Cin; ++c) sum += (double)cw.w[(((size_t)o * Cin + c) * k + r) * k + sx];
                            }
                        }
                        tab[((size_t)rc * ncc + cc) * Np + o] = (float)((double)cw.b[o] + (double)cw.in_shift * sum);
                    }
                }
            }
            FCNOD_CUDA(cudaMalloc(&cw.d_bias_tab, tab.size() * 4));
            FCNOD_CUDA(cudaMemcpy(cw.d_bias_tab, tab.data(), tab.size() * 4, cudaMemcpyHostToDevice));
        } else {
            const int Kp = k * k * cw.n_cchunks * cw.CK;
            const int Nrows = (Cout + 15) / 16 * 16, Np = (int)align_up((size_t)(Cout + 7) / 8 * 8, 256) + 256;
            // packed [K-block][N][CK]; K-block order = the kernel's consumption order:
            //   TAP  (r * k + s) * n_cchunks + cc        HALO  (r * n_cchunks + cc) * k + s   (taps of a kernel row adjacent)
            std::vector<bf16> w((size_t)Nrows * Kp, __float2bfloat16(0.f));
            for (int o = 0; o < Cout; ++o)
                for (int c = 0; c < Cin; ++c)
                    for (int r = 0; r < k; ++r)
                        for (int s = 0; s < k; ++s) {
                            const int cc = c / cw.CK;
                            const size_t kb = cw.halo ? (size_t)(r * cw.n_cchunks + cc) * k + s : (size_t)(r * k + s) * cw.n_cchunks + cc;
                            w[(kb * Nrows + o) * cw.CK + c % cw.CK] = __float2bfloat16(cw.w[(((size_t)o * Cin + c) * k + r) * k + s]);
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
// captured graphs hold kernel parameters (tensor maps, tilings, batch): any change of those invalidates them
void drop_graphs(fcnod_net* net) {
    if (net->graph) { cudaGraphExecDestroy(net->graph); net->graph = 0; }
    if (net->sgraph) { cudaGraphExecDestroy(net->sgraph); net->sgraph = 0; }
    for (int i = 0; i < 2; ++i)
        if (net->slots[i].graph) { cudaGraphExecDestroy(net->slots[i].graph); net->slots[i].graph = 0; }
}

void free_activations(fcnod_net* net) {
    for (size_t i = 0; i < net->buffers.size(); ++i)
        if (net->buffers[i].ptr) { cudaFree(net->buffers[i].ptr); net->buffers[i].ptr = 0; }
    for (size_t i = 0; i < net->out_dev.size(); ++i) if (net->out_dev[i]) cudaFree(net->out_dev[i]);
    net->out_dev.clear();
    if (net->data_dev) { cudaFree(net->data_dev); net->data_dev = 0; }
    if (net->scratch_export) { cudaFree(net->scratch_export); net->scratch_export = 0; net->scratch_export_elems = 0; }
    if (net->post_scratch) { cudaFree(net->post_scratch); net->post_scratch = 0; }
    if (net->det_dev) { cudaFree(net->det_dev); net->det_dev = 0; }      // one block: det | label | weight | count
    net->label_dev = net->weight_dev = net->count_dev = 0;
    if (net->srchw_dev) { cudaFree(net->srchw_dev); net->srchw_dev = 0; }
    net->srchw_b = 0;
    if (net->img_dev) { cudaFree(net->img_dev); net->img_dev = 0; net->img_dev_bytes = 0; }
    if (net->pre_scratch) { cudaFree(net->pre_scratch); net->pre_scratch = 0; }
    for (int i = 0; i < 2; ++i) {
        if (net->slots[i].s2d) { cudaFree(net->slots[i].s2d); net->slots[i].s2d = 0; }
        if (net->slots[i].mm) { cudaFree(net->slots[i].mm); net->slots[i].mm = 0; }
    }
    drop_graphs(net);
    if (net->pinned_in) { cudaFreeHost(net->pinned_in); net->pinned_in = 0; net->pinned_in_bytes = 0; }
    if (net->pinned_out) { cudaFreeHost(net->pinned_out); net->pinned_out = 0; }
}

CUresult encode_map(CUtensorMap* m, void* base, int rank, const cuuint64_t* dims, const cuuint64_t* strides,
                    const cuuint32_t* box, uint32_t span, CUtensorMapL2promotion l2) {
    cuuint32_t es[4] = {1, 1, 1, 1};
    const CUtensorMapSwizzle sw = span == 128 ? CU_TENSOR_MAP_SWIZZLE_128B
                                              : (span == 64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_32B);
    return get_encode_tiled()(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, rank, base, dims, strides, box, es,
                              CU_TENSOR_MAP_INTERLEAVE_NONE, sw, l2, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
}

void bind_tc_pointers(fcnod_net* net, Step& st) {
    ConvWeights& cw = net->convs[st.conv];
    ConvTc2Params& p = st.tc;
    p.bias = cw.d_bias_tab ? cw.d_bias_tab + (size_t)p.bt_interior * p.bt_ld : cw.d_bias;
    p.bias_tab = cw.d_bias_tab;
    for (size_t i = 0; i < st.segs.size(); ++i) {
        const Step::Seg& sg = st.segs[i];
        if (sg.flags & SEG_DIRECT) {
            const Blob& ob = net->blobs[sg.direct_blob];
            FCNOD_CHECK(ob.out_index >= 0, "internal: direct segment without an output slot");
            p.seg[i].out = net->out_dev[ob.out_index];
            p.seg[i].ldo = ob.C;
        } else {
            const Blob& ob = net->blobs[sg.blob];
            p.seg[i].out = view_ptr(net, ob);
            p.seg[i].ldo = ob.ldc;
        }
    }
}

// B = packed weights [K-block][Nrows][CK]; box = b_taps consecutive K-blocks x BN rows x CK
void encode_weight_map(Step& st, const ConvWeights& cw, int Nrows, int n_kblocks, const std::string& name) {
    const ConvTc2Params& p = st.tc;
    const uint32_t span = p.CK * 2;
    cuuint64_t dims[3] = {(cuuint64_t)p.CK, (cuuint64_t)Nrows, (cuuint64_t)n_kblocks};
    cuuint64_t strides[2] = {(cuuint64_t)span, (cuuint64_t)Nrows * span};
    cuuint32_t box[3] = {(cuuint32_t)p.CK, (cuuint32_t)p.BN, (cuuint32_t)p.b_taps};
    CUresult r = encode_map(&st.tmap_b, cw.d_w_tc, 3, dims, strides, box, span, CU_TENSOR_MAP_L2_PROMOTION_L2_256B);
    FCNOD_CHECK(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled(B) failed for '" + name + "' (" + std::to_string((int)r) + ")");
}

void setup_tc_step(fcnod_net* net, Step& st) {
    const LayerDef& L = net->def.layers[st.layer];
    ConvWeights& cw = net->convs[st.conv];
    const Blob& ib = net->blobs[st.in];
    const Blob& ob = net->blobs[st.out];
    const int B = net->batch;
    const int Nrows = (cw.Cout + 15) / 16 * 16;
    if (cw.s2d) {
        // A = the W-padded space-to-depth image [B][H2][Wp][16] bf16.  ONE K-block per kernel row: its A rows are the
        // 128-byte windows of ka consecutive pixels, fetched through an overlapping-stride tensor map
        // (dim0 = ka*16 contiguous elements, dim1 = window start, one pixel = 32 bytes apart).
        const int H2 = ob.H, W2 = ob.W, Wp = st.s2d_wp;
        void* base = net->buffers[st.s2d_buf].ptr;
        const cuuint64_t pix = kS2dCpad * 2;
        fill_tc_geometry(net, st, B, st.tc_pick);
        ConvTc2Params& p = st.tc;
        const uint32_t span = p.CK * 2;
        (void)W2;
        (void)H2;       // the tensor map is only a valid placeholder in bulk mode (the kernel prefetches it)
        cuuint64_t dims[4] = {(cuuint64_t)p.CK, (cuuint64_t)Wp, (cuuint64_t)st.s2d_hp, (cuuint64_t)B};
        cuuint64_t strides[3] = {pix, (cuuint64_t)Wp * pix, (cuuint64_t)st.s2d_hp * Wp * pix};
        cuuint32_t box[4] = {(cuuint32_t)p.CK, (cuuint32_t)p.PW, (cuuint32_t)(p.BH + p.kh - 1), 1};
        CUresult r = encode_map(&st.tmap_a, base, 4, dims, strides, box, span, CU_TENSOR_MAP_L2_PROMOTION_L2_128B);
        FCNOD_CHECK(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled(A, s2d) failed for '" + L.name + "' (" + std::to_string((int)r) + ")");
        encode_weight_map(st, cw, Nrows, cw.ka * cw.ka, L.name);
        bind_tc_pointers(net, st);
        FCNOD_CHECK(st.s2d_hp >= p.tiles_y * p.BH + p.kh - 1 && Wp >= (p.tiles_x - 1) * p.BW + p.PW,
                    "internal: space-to-depth image smaller than the bulk-copy patches of '" + L.name + "'");
        p.a_src = base;
        p.a_row_pitch = (unsigned)(Wp * pix);
        p.a_img_pitch = (unsigned long long)st.s2d_hp * Wp * pix;
        return;
    }
    fill_tc_geometry(net, st, B, st.tc_pick);
    ConvTc2Params& p = st.tc;
    const uint32_t span = cw.CK * 2;
    {
        cuuint64_t dims[4] = {(cuuint64_t)ib.C, (cuuint64_t)ib.W, (cuuint64_t)ib.H, (cuuint64_t)B};
        cuuint64_t strides[3] = {(cuuint64_t)ib.ldc * 2, (cuuint64_t)ib.W * ib.ldc * 2, (cuuint64_t)ib.H * ib.W * ib.ldc * 2};
        cuuint32_t box[4] = {(cuuint32_t)cw.CK, (cuuint32_t)p.PW, (cuuint32_t)(p.halo ? p.BH + p.kh - 1 : p.BH), 1};
        if (p.pool) { box[1] = (cuuint32_t)(p.BW + 2); box[2] = (cuuint32_t)(p.BH + 2); }     // raw halo patch of the 3x3 pool
        if (p.flat) {   // all pixels of the batch as one row: [C][B*H*W][1][1]
            dims[1] = (cuuint64_t)B * ib.H * ib.W; dims[2] = 1; dims[3] = 1;
            strides[1] = dims[1] * ib.ldc * 2; strides[2] = strides[1];
        }
        CUresult r = encode_map(&st.tmap_a, view_ptr(net, ib), 4, dims, strides, box, span, CU_TENSOR_MAP_L2_PROMOTION_L2_128B);
        FCNOD_CHECK(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled(A) failed for '" + L.name + "' (" + std::to_string((int)r) + ")");
    }
    encode_weight_map(st, cw, Nrows, cw.k * cw.k * cw.n_cchunks, L.name);
    bind_tc_pointers(net, st);
}

void launch_conv_tc2(const Step& st, cudaStream_t s, bool prof);

// The tiling model ranks the candidates; the device has the last word: the first few are timed on the real buffers
// (contents are irrelevant, every tiling computes the same sums in the same K order) and the fastest one is kept.
// FCNOD_AUTOTUNE=0 keeps the model's first choice; FCNOD_AUTOTUNE=n times n candidates (default 12).
const int kAutotuneDefault = 12;
void autotune_tc_steps(fcnod_net* net) {
    const char* e = getenv("FCNOD_AUTOTUNE");
    const int ntry = e ? atoi(e) : kAutotuneDefault;
    if (ntry <= 1) return;
    cudaEvent_t e0, e1;
    FCNOD_CUDA(cudaEventCreate(&e0));
    FCNOD_CUDA(cudaEventCreate(&e1));
    cudaStream_t s = net->stream;
    for (size_t i = 0; i < net->steps.size(); ++i) {
        Step& st = net->steps[i];
        if (st.kind != S_CONV_TC) continue;
        int best = 0;
        float best_ms = 0.f;
        const int n = std::min(ntry, st.tc_ncand);
        for (int c = 0; c < n; ++c) {
            st.tc_pick = c;
            setup_tc_step(net, st);
            launch_conv_tc2(st, s, false);                       // warm-up (instruction cache, L2)
            FCNOD_CUDA(cudaEventRecord(e0, s));
            for (int r = 0; r < 3; ++r) launch_conv_tc2(st, s, false);
            FCNOD_CUDA(cudaEventRecord(e1, s));
            FCNOD_CUDA(cudaEventSynchronize(e1));
            FCNOD_CUDA(cudaGetLastError());
            float ms = 0.f;
            FCNOD_CUDA(cudaEventElapsedTime(&ms, e0, e1));
            if (c == 0 || ms < best_ms) { best_ms = ms; best = c; }
        }
        st.tc_pick = best;
        st.solo_ms = best_ms / 3.f;
        setup_tc_step(net, st);
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
}

void forward_device(fcnod_net* net, const float* data_dev, cudaStream_t s, bool skip_input_step = false);

// Independent launches of different lanes (the 3x3, the 5x5 and the pool_proj convolution of one inception module) only
// overlap at their tails while each of them asks for all 148 SMs: the first one takes the device, the others start as its
// CTAs retire.  Here every such group gets a PARTITION of the SMs instead -- grids that add up to 148, sized from the
// launches' solo times so that all members finish together -- so that they run side by side from the start: no member
// pays its own last, mostly empty round (448 tiles on 148 SMs = 3.03 rounds), and three prologues overlap.  The device
// has the last word: the whole forward is timed with and without each group's partition and the faster form is kept.
void partition_lane_groups(fcnod_net* net) {
    // Opt-in (FCNOD_PARTITION=1).  Measured on the bench (batch 64): the device check keeps the partition for 4 of the 9
    // inception modules and the step goes from 1.840 to 1.832 ms -- inside the run-to-run noise -- while every partitioned
    // launch, timed alone for the roofline table, looks slower than it is in the step.  Not worth it by default.
    static const bool on = getenv("FCNOD_PARTITION") && getenv("FCNOD_PARTITION")[0] == '1';
    if (!on || net->n_lanes <= 1) return;
    const size_t n = net->steps.size();
    std::vector<std::vector<int> > groups;
    std::vector<int> cur;
    for (size_t j = 0; j < n; ++j) {
        bool dep = false;
        for (size_t g = 0; g < cur.size() && !dep; ++g)
            dep = std::find(net->steps[j].deps.begin(), net->steps[j].deps.end(), cur[g]) != net->steps[j].deps.end();
        if (dep) { groups.push_back(cur); cur.clear(); }
        cur.push_back((int)j);
    }
    groups.push_back(cur);
    const double F = 0.004;     // ms of a launch that do not scale with its rounds (prologue, pipeline fill, drain)
    std::vector<std::vector<int> > planned;
    for (size_t gi = 0; gi < groups.size(); ++gi) {
        const std::vector<int>& G = groups[gi];
        if (G.size() < 2 || G.size() > 3) continue;
        bool ok = true;
        std::set<int> lanes;
        for (size_t k = 0; k < G.size(); ++k) {
            const Step& st = net->steps[G[k]];
            ok = ok && st.kind == S_CONV_TC && st.solo_ms > 0.f;
            lanes.insert(st.lane);
        }
        if (!ok || lanes.size() != G.size()) continue;
        // per-round time of every member from its solo launch
        double tau[3]; int tiles[3], nt[3];
        for (size_t k = 0; k < G.size(); ++k) {
            const Step& st = net->steps[G[k]];
            nt[k] = st.tc.n_tiles; tiles[k] = st.tc.m_tiles;
            const int slots = std::max(1, std::min(tiles[k], kNumSMs / nt[k]));
            const int rounds = (tiles[k] + slots - 1) / slots;
            tau[k] = std::max((double)st.solo_ms - F, 0.3 * st.solo_ms) / rounds;
        }
        auto cost = [&](int k, int g) {
            const int slots = std::max(1, g / nt[k]);
            return F + (double)((tiles[k] + slots - 1) / slots) * tau[k];
        };
        double best = 1e30; int bg[3] = {0, 0, 0};
        if (G.size() == 2) {
            for (int g0 = nt[0]; g0 + nt[1] <= kNumSMs; g0 += nt[0]) {
                const int g1 = (kNumSMs - g0) / nt[1] * nt[1];
                const double t = std::max(cost(0, g0), cost(1, g1));
                if (t < best) { best = t; bg[0] = g0; bg[1] = g1; }
            }
        } else {
            for (int g0 = nt[0]; g0 + nt[1] + nt[2] <= kNumSMs; g0 += nt[0])
                for (int g1 = nt[1]; g0 + g1 + nt[2] <= kNumSMs; g1 += nt[1]) {
                    const int g2 = (kNumSMs - g0 - g1) / nt[2] * nt[2];
                    const double t = std::max(std::max(cost(0, g0), cost(1, g1)), cost(2, g2));
                    if (t < best) { best = t; bg[0] = g0; bg[1] = g1; bg[2] = g2; }
                }
        }
        if (best >= 1e30) continue;
        for (size_t k = 0; k < G.size(); ++k) { net->steps[G[k]].grid_cap = bg[k]; setup_tc_step(net, net->steps[G[k]]); }
        planned.push_back(G);
    }
    if (planned.empty()) return;
    // device check: the forward with all partitions, then each group reverted in turn; keep what is faster
    cudaEvent_t e0, e1;
    FCNOD_CUDA(cudaEventCreate(&e0));
    FCNOD_CUDA(cudaEventCreate(&e1));
    cudaStream_t s = net->stream;
    auto time_forward = [&]() {
        float best_ms = 1e30f;
        forward_device(net, net->data_dev, s, false);
        for (int rep = 0; rep < 3; ++rep) {
            FCNOD_CUDA(cudaEventRecord(e0, s));
            for (int r = 0; r < 2; ++r) forward_device(net, net->data_dev, s, false);
            FCNOD_CUDA(cudaEventRecord(e1, s));
            FCNOD_CUDA(cudaEventSynchronize(e1));
            float ms = 0.f;
            FCNOD_CUDA(cudaEventElapsedTime(&ms, e0, e1));
            best_ms = std::min(best_ms, ms);
        }
        FCNOD_CUDA(cudaGetLastError());
        return best_ms;
    };
    float t_cur = time_forward();
    for (size_t gi = 0; gi < planned.size(); ++gi) {
        std::vector<int> saved;
        for (size_t k = 0; k < planned[gi].size(); ++k) {
            Step& st = net->steps[planned[gi][k]];
            saved.push_back(st.grid_cap);
            st.grid_cap = 0; setup_tc_step(net, st);
        }
        const float t_off = time_forward();
        if (t_off < t_cur) { t_cur = t_off; continue; }          // this group is better off with full grids
        for (size_t k = 0; k < planned[gi].size(); ++k) {
            Step& st = net->steps[planned[gi][k]];
            st.grid_cap = saved[k]; setup_tc_step(net, st);
        }
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
}

void reshape(fcnod_net* net, int batch) {
    FCNOD_CHECK(batch > 0, "batch must be positive");
    if (net->device < 0) {   // plan-only handle: re-plan the tilings for the new batch (host only)
        net->batch = batch;
        for (size_t i = 0; i < net->steps.size(); ++i)
            if (net->steps[i].kind == S_CONV_TC) fill_tc_geometry(net, net->steps[i], batch);
        return;
    }
    require_device(net);
    if (batch == net->batch && !net->buffers.empty() && net->buffers[0].ptr) return;
    FCNOD_CUDA(cudaDeviceSynchronize());
    drop_graphs(net);
    // Buffers are [batch][H][W][C] with the batch outermost: a smaller batch runs in a prefix of a larger allocation
    // (the zero padding of the space-to-depth image sits at the same offsets), so alternating batch sizes -- the
    // node's batch-1 run_detector next to batched calls -- neither reallocates nor re-times the tilings.
    if (batch > net->capacity) {
        free_activations(net);
        net->capacity = batch;
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
    }
    net->batch = batch;
    std::map<int, std::vector<int> >::const_iterator pk = net->tc_picks.find(batch);
    for (size_t i = 0; i < net->steps.size(); ++i)
        if (net->steps[i].kind == S_CONV_TC) {
            net->steps[i].tc_pick = pk != net->tc_picks.end() ? pk->second[i] : -1;
            net->steps[i].grid_cap = pk != net->tc_picks.end() ? net->tc_caps[batch][i] : 0;
            setup_tc_step(net, net->steps[i]);
        }
    if (pk == net->tc_picks.end()) {
        autotune_tc_steps(net);
        partition_lane_groups(net);
        std::vector<int> picks(net->steps.size(), -1), caps(net->steps.size(), 0);
        for (size_t i = 0; i < net->steps.size(); ++i) { picks[i] = net->steps[i].tc_pick; caps[i] = net->steps[i].grid_cap; }
        net->tc_picks[batch] = picks;
        net->tc_caps[batch] = caps;
    }
}

// ------------------------------------------------------------------------------------ execution
template <typename T>
void launch_pool(const fcnod_net* net, const Step& st, const LayerDef& L, cudaStream_t s) {
    const Blob& ib = net->blobs[st.in];
    const Blob& ob = net->blobs[st.out];
    long items = (long)net->batch * ob.H * ob.W * (ib.C / 8);
    if (L.kernel == 3 && (L.stride == 1 || L.stride == 2) && L.pad <= 1 && ib.C % 8 == 0 && !getenv("FCNOD_POOL_V1")) {
        // register-marching kernel: CTA = 8 channel groups x 32 output columns x a strip of output rows
        const int cg_chunks = ceil_div(ib.C / 8, 8), col_chunks = ceil_div(ob.W, 32);
        static const int rs_env = getenv("FCNOD_POOL_RS") ? atoi(getenv("FCNOD_POOL_RS")) : 0;
        const int RS = rs_env > 0 ? std::min(rs_env, ob.H) : (ob.H <= 14 ? ob.H : 14), strips = ceil_div(ob.H, RS);
        const long blocks = (long)net->batch * strips * col_chunks * cg_chunks;
        if (L.stride == 1)
            pool3_march_kernel<T, 1><<<(unsigned)blocks, 256, 0, s>>>((const T*)view_ptr(net, ib), (T*)view_ptr(net, ob), ib.C, ib.H,
                                                                      ib.W, ib.ldc, ob.H, ob.W, ob.ldc, L.pad, cg_chunks, col_chunks,
                                                                      strips, RS);
        else
            pool3_march_kernel<T, 2><<<(unsigned)blocks, 256, 0, s>>>((const T*)view_ptr(net, ib), (T*)view_ptr(net, ob), ib.C, ib.H,
                                                                      ib.W, ib.ldc, ob.H, ob.W, ob.ldc, L.pad, cg_chunks, col_chunks,
                                                                      strips, RS);
        return;
    }
    if (L.kernel == 3 && (L.stride == 1 || L.stride == 2) && items < (1L << 31)) {
        // 2x2 output pixels per thread from one shared input window
        long blocks2 = (long)net->batch * ((ob.H + 1) / 2) * ((ob.W + 1) / 2) * (ib.C / 8);
        int g = grid_for(blocks2, 256);
        if (L.stride == 1)
            pool_tile_kernel<T, 3, 1><<<g, 256, 0, s>>>((const T*)view_ptr(net, ib), (T*)view_ptr(net, ob), net->batch, ib.C,
                                                         ib.H, ib.W, ib.ldc, ob.H, ob.W, ob.ldc, L.pad);
        else
            pool_tile_kernel<T, 3, 2><<<g, 256, 0, s>>>((const T*)view_ptr(net, ib), (T*)view_ptr(net, ob), net->batch, ib.C,
                                                         ib.H, ib.W, ib.ldc, ob.H, ob.W, ob.ldc, L.pad);
        return;
    }
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
    p.kh = p.kw = cw.k; p.stride = L.stride; p.pad = L.pad; p.relu = st.relu ? 1 : 0; p.accum = 0;
    long M = (long)net->batch * ob.H * ob.W;
    dim3 grid((unsigned)((M + 63) / 64), (unsigned)(cw.ldw / 64));
    if (ib.dtype == DT_F32 && ob.dtype == DT_F32) conv_simt_kernel<float, float><<<grid, 256, 0, s>>>(p);
    else if (ib.dtype == DT_F32 && ob.dtype == DT_BF16) conv_simt_kernel<float, bf16><<<grid, 256, 0, s>>>(p);
    else if (ib.dtype == DT_BF16 && ob.dtype == DT_BF16) conv_simt_kernel<bf16, bf16><<<grid, 256, 0, s>>>(p);
    else conv_simt_kernel<bf16, float><<<grid, 256, 0, s>>>(p);
}

// the instantiations of the tensor-core convolution: K steps per block (CK / 16) x HALO x kernel size, + the profiled twins
template <int KS, bool HALO, int KK>
void launch_tc2_inst(const Step& st, cudaStream_t s, bool prof) {
    const dim3 g(st.tc_grid), b(kConvTc2Threads);
    if (prof) conv_tc2_kernel<KS, HALO, true, KK><<<g, b, st.tc_smem, s>>>(st.tmap_a, st.tmap_b, st.tc);
    else conv_tc2_kernel<KS, HALO, false, KK><<<g, b, st.tc_smem, s>>>(st.tmap_a, st.tmap_b, st.tc);
}
void launch_conv_tc2(const Step& st, cudaStream_t s, bool prof) {
    const int ks = st.tc.CK / 16;
    if (st.tc.pool) {
        FCNOD_CHECK(ks == 4 && !st.tc.halo && !st.tc.dual, "internal: POOL mode needs 64-channel blocks, one issuer");
        const dim3 g(st.tc_grid), b(kConvTc2Threads);
        if (prof) conv_tc2_kernel<4, false, true, 0, true><<<g, b, st.tc_smem, s>>>(st.tmap_a, st.tmap_b, st.tc);
        else conv_tc2_kernel<4, false, false, 0, true><<<g, b, st.tc_smem, s>>>(st.tmap_a, st.tmap_b, st.tc);
    } else if (st.tc.lrn) {
        const dim3 g(st.tc_grid), b(kConvTc2Threads);
        const bool k3 = st.tc.halo && st.tc.kh == 3 && st.tc.kw == 3;
        FCNOD_CHECK(ks == 4 && (k3 || !st.tc.halo) && st.tc.n_tiles == 1, "internal: LRN epilogue needs 64-channel blocks, a 3x3 or 1x1 kernel, one N tile");
        if (k3) conv_tc2_kernel<4, true, false, 3, false, true><<<g, b, st.tc_smem, s>>>(st.tmap_a, st.tmap_b, st.tc);
        else conv_tc2_kernel<4, false, false, 0, false, true><<<g, b, st.tc_smem, s>>>(st.tmap_a, st.tmap_b, st.tc);
        FCNOD_CHECK(!prof, "the LRN-epilogue instantiations have no profiled twin");
    } else if (st.tc.halo) {
        const int kk = st.tc.kh == st.tc.kw ? st.tc.kh : 0;
        if (ks == 4 && kk == 3) launch_tc2_inst<4, true, 3>(st, s, prof);
        else if (ks == 4 && kk == 5) launch_tc2_inst<4, true, 5>(st, s, prof);
        else if (ks == 4) launch_tc2_inst<4, true, 0>(st, s, prof);
        else if (ks == 2 && kk == 5) launch_tc2_inst<2, true, 5>(st, s, prof);
        else if (ks == 2) launch_tc2_inst<2, true, 0>(st, s, prof);
        else if (kk == 4) launch_tc2_inst<1, true, 4>(st, s, prof);
        else if (kk == 5) launch_tc2_inst<1, true, 5>(st, s, prof);
        else launch_tc2_inst<1, true, 0>(st, s, prof);
    } else if (ks == 4) launch_tc2_inst<4, false, 0>(st, s, prof);
    else if (ks == 2) launch_tc2_inst<2, false, 0>(st, s, prof);
    else launch_tc2_inst<1, false, 0>(st, s, prof);
}
template <int KS, bool HALO, int KK>
void tc2_smem_attr() {
    const int lim = 227 * 1024;
    FCNOD_CUDA(cudaFuncSetAttribute(conv_tc2_kernel<KS, HALO, false, KK>, cudaFuncAttributeMaxDynamicSharedMemorySize, lim));
    FCNOD_CUDA(cudaFuncSetAttribute(conv_tc2_kernel<KS, HALO, true, KK>, cudaFuncAttributeMaxDynamicSharedMemorySize, lim));
}
void conv_tc2_set_smem_limit() {      // per device, at net creation (outside any graph capture)
    tc2_smem_attr<4, true, 3>(); tc2_smem_attr<4, true, 5>(); tc2_smem_attr<4, true, 0>();
    tc2_smem_attr<1, true, 4>(); tc2_smem_attr<1, true, 0>(); tc2_smem_attr<1, true, 5>();
    tc2_smem_attr<2, true, 5>(); tc2_smem_attr<2, true, 0>();
    tc2_smem_attr<4, false, 0>(); tc2_smem_attr<2, false, 0>(); tc2_smem_attr<1, false, 0>();
    FCNOD_CUDA(cudaFuncSetAttribute(conv_tc2_kernel<4, true, false, 3, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    FCNOD_CUDA(cudaFuncSetAttribute(conv_tc2_kernel<4, false, false, 0, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    FCNOD_CUDA(cudaFuncSetAttribute(conv_tc2_kernel<4, false, false, 0, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    FCNOD_CUDA(cudaFuncSetAttribute(conv_tc2_kernel<4, false, true, 0, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
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
        case S_S2D: {
            const Blob& ib = net->blobs[0];
            long items = (long)net->batch * (ib.H / 2) * (ib.W / 2);
            bf16* dst = (bf16*)net->buffers[st.s2d_buf].ptr;
            const int g = grid_for(items, 256);
            if (ib.C == 3)
                input_s2d_kernel<3><<<g, 256, 0, s>>>(data_dev, dst, net->batch, ib.C, ib.H, ib.W, st.s2d_wp, st.s2d_hp, st.s2d_padl, kS2dCpad);
            else if (ib.C == 1)
                input_s2d_kernel<1><<<g, 256, 0, s>>>(data_dev, dst, net->batch, ib.C, ib.H, ib.W, st.s2d_wp, st.s2d_hp, st.s2d_padl, kS2dCpad);
            else
                input_s2d_kernel<0><<<g, 256, 0, s>>>(data_dev, dst, net->batch, ib.C, ib.H, ib.W, st.s2d_wp, st.s2d_hp, st.s2d_padl, kS2dCpad);
            break;
        }
        case S_POOL_LRN: {   // layer = Pooling, layer2 = LRN
            const LayerDef& R = net->def.layers[st.layer2];
            const Blob& ib = net->blobs[st.in];
            const Blob& ob = net->blobs[st.out];
            FCNOD_CHECK(ib.dtype == DT_BF16 && ob.dtype == DT_BF16, "pool+lrn: bf16 only");
            if (ib.C == 64) {   // register-marching form (deploy net: pool1 + norm1 on 64 channels)
                constexpr int RS = 8;
                const long threads = (long)net->batch * ceil_div(ob.H, RS) * ob.W * 8;
                pool3s2_lrn5_march_kernel<8, RS, true><<<(unsigned)((threads + 255) / 256), 256, 0, s>>>(
                    (const bf16*)view_ptr(net, ib), (bf16*)view_ptr(net, ob), net->batch, ib.H, ib.W, ib.ldc, ob.H, ob.W, ob.ldc,
                    L->pad, R.alpha / R.local_size, R.beta, R.lrn_k);
                break;
            }
            {   // shared-memory tiled form when the tile fits
                constexpr int TH = 4, TW = 16;
                const size_t region = align_up((size_t)(2 * (TH - 1) + 3) * (2 * (TW - 1) + 3) * ib.C * 2, 16);
                const size_t smem_t = region + (size_t)TH * TW * (ib.C + 16) * 4;
                if (smem_t <= 100 * 1024) {
                    const int tx = ceil_div(ob.W, TW), ty = ceil_div(ob.H, TH);
                    pool_lrn_tile_kernel<bf16, 3, 2, 2, true, TH, TW><<<tx * ty * net->batch, 256, smem_t, s>>>(
                        (const bf16*)view_ptr(net, ib), (bf16*)view_ptr(net, ob), ib.C, ib.H, ib.W, ib.ldc, ob.H, ob.W, ob.ldc,
                        L->pad, tx, ty, R.alpha / R.local_size, R.beta, R.lrn_k);
                    break;
                }
            }
            const int CG = ib.C / 8, npix_cta = 256 / CG;
            long npix = (long)net->batch * ob.H * ob.W;
            int g = (int)std::min<long>((npix + npix_cta - 1) / npix_cta, (long)kNumSMs * 8);
            size_t smem = (size_t)npix_cta * (ib.C + 16) * 4;
            pool_lrn_kernel<bf16, 3, 2, 2, true><<<g, 256, smem, s>>>((const bf16*)view_ptr(net, ib), (bf16*)view_ptr(net, ob),
                                                                      net->batch, ib.C, ib.H, ib.W, ib.ldc, ob.H, ob.W, ob.ldc,
                                                                      L->pad, R.alpha / R.local_size, R.beta, R.lrn_k);
            break;
        }
        case S_LRN_POOL: {   // layer = LRN, layer2 = Pooling
            const LayerDef& P = net->def.layers[st.layer2];
            const Blob& ib = net->blobs[st.in];
            const Blob& ob = net->blobs[st.out];
            FCNOD_CHECK(ib.dtype == DT_BF16 && ob.dtype == DT_BF16, "lrn+pool: bf16 only");
            static const bool lp_v1 = getenv("FCNOD_LRNPOOL_V1") != nullptr;
            if (!lp_v1 && ib.C % 32 == 0 && L->beta == 0.75f) {       // register-marching form: warps of 8 columns x 4 channel groups
                constexpr int RS = 8;
                const int colblocks = ceil_div(ob.W, 7), tasks = colblocks * (ib.C / 32);
                lrn5_pool3s2_march_kernel<RS, true><<<dim3(ceil_div(tasks, 8), ceil_div(ob.H, RS), net->batch), 256, 0, s>>>(
                    (const bf16*)view_ptr(net, ib), (bf16*)view_ptr(net, ob), ib.C, ib.H, ib.W, ib.ldc, ob.H, ob.W, ob.ldc, P.pad,
                    colblocks, L->alpha / L->local_size, L->lrn_k);
                break;
            }
            constexpr int TH = 4, TW = 8;
            const int tx = ceil_div(ob.W, TW), ty = ceil_div(ob.H, TH);
            size_t smem = (size_t)(2 * (TH - 1) + 3) * (2 * (TW - 1) + 3) * ib.C * 2;
            if (ib.C == 192)
                lrn_pool_kernel<bf16, 3, 2, 2, true, TH, TW, 24><<<tx * ty * net->batch, 512, smem, s>>>(
                    (const bf16*)view_ptr(net, ib), (bf16*)view_ptr(net, ob), ib.C, ib.H, ib.W, ib.ldc, ob.H, ob.W, ob.ldc, P.pad,
                    tx, ty, L->alpha / L->local_size, L->beta, L->lrn_k);
            else
                lrn_pool_kernel<bf16, 3, 2, 2, true, TH, TW, 0><<<tx * ty * net->batch, 512, smem, s>>>(
                    (const bf16*)view_ptr(net, ib), (bf16*)view_ptr(net, ob), ib.C, ib.H, ib.W, ib.ldc, ob.H, ob.W, ob.ldc, P.pad,
                    tx, ty, L->alpha / L->local_size, L->beta, L->lrn_k);
            break;
        }
        case S_CONV_SIMT: launch_conv_simt(net, st, *L, s); break;
        case S_CONV_TC:
            launch_conv_tc2(st, s, false);
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

void forward_device(fcnod_net* net, const float* data_dev, cudaStream_t s, bool skip_input_step) {
    FCNOD_CHECK(net->device >= 0, "plan-only handle (device < 0): no compute is possible; there is no CPU fallback");
    FCNOD_CHECK(net->batch > 0, "net not reshaped");
    const size_t first = skip_input_step ? 1 : 0;       // the input step's output was produced by the fused prelude
    if (net->n_lanes <= 1) {
        for (size_t i = first; i < net->steps.size(); ++i) run_step(net, net->steps[i], data_dev, s);
        return;
    }
    // lane 0 = the caller's stream; the others fork from it here and join it again at the end (under stream capture the
    // event waits become graph edges: independent branches are independent graph nodes)
    cudaStream_t lanes[3] = {net->lane_stream[0], net->lane_stream[1], net->lane_stream[2]};
    bool forked[3] = {false, false, false};
    FCNOD_CUDA(cudaEventRecord(net->fork_ev, s));
    for (size_t i = 0; i < net->steps.size(); ++i) {
        const Step& st = net->steps[i];
        cudaStream_t ls = lanes[st.lane];
        if (!forked[st.lane]) { FCNOD_CUDA(cudaStreamWaitEvent(ls, net->fork_ev, 0)); forked[st.lane] = true; }
        for (size_t w = 0; w < st.waits.size(); ++w) FCNOD_CUDA(cudaStreamWaitEvent(ls, net->steps[st.waits[w]].ev, 0));
        if (i >= first) run_step(net, st, data_dev, ls);
        if (st.signal) FCNOD_CUDA(cudaEventRecord(st.ev, ls));
    }
    for (int l = 0; l < 3; ++l)
        if (forked[l]) {
            FCNOD_CUDA(cudaEventRecord(net->join_ev[l], lanes[l]));
            FCNOD_CUDA(cudaStreamWaitEvent(s, net->join_ev[l], 0));
        }
}

void ensure_detect_buffers(fcnod_net* net) {
    const int B = net->capacity;
    const int c_cvg = net->blobs[blob_id(net, "coverage")].C;
    if (!net->post_scratch) FCNOD_CUDA(cudaMalloc(&net->post_scratch, postnet_scratch_bytes(B, c_cvg)));
    // detections of a batch are ONE contiguous block laid out for the CURRENT batch size b: det [b][M][5] | label [b][M] |
    // weight [b][M] | count [b] -- the layout of the pinned result buffers, so that they leave the device in one copy
    if (!net->det_dev) FCNOD_CUDA(cudaMalloc(&net->det_dev, (size_t)B * (FCNOD_MAX_DET * 7 + 1) * 4));
    net->label_dev = net->det_dev + (size_t)net->batch * FCNOD_MAX_DET * 5;
    net->weight_dev = net->label_dev + (size_t)net->batch * FCNOD_MAX_DET;
    net->count_dev = net->weight_dev + (size_t)net->batch * FCNOD_MAX_DET;
    if (!net->srchw_dev) FCNOD_CUDA(cudaMalloc(&net->srchw_dev, (size_t)B * 2 * 4));
    if (!net->pinned_out) FCNOD_CUDA(cudaMallocHost(&net->pinned_out, (size_t)B * (FCNOD_MAX_DET * 7 + 1) * 4));
}

void detect_device(fcnod_net* net, const float* data_dev, const int32_t* srchw_dev, const fcnod_postnet_params& p,
                   cudaStream_t s, bool skip_input_step = false) {
    forward_device(net, data_dev, s, skip_input_step);
    int ic = blob_id(net, "coverage"), ib = blob_id(net, "bboxes");
    const Blob& bc = net->blobs[ic];
    const Blob& bb = net->blobs[ib];
    FCNOD_CHECK(bc.is_output && bb.is_output, "blobs 'coverage' and 'bboxes' must be network outputs");
    FCNOD_CHECK(bc.H == bb.H && bc.W == bb.W, "coverage / bboxes grid mismatch");
    ensure_detect_buffers(net);
    postnet_launch(net->out_dev[bc.out_index], net->out_dev[bb.out_index], net->batch, bc.C, bb.C, bc.H, bc.W, srchw_dev, p,
                   net->post_scratch, net->det_dev, net->label_dev, net->weight_dev, net->count_dev, s);
}

// Prelude of the detect paths: uint8 frames on the device -> the network's input.  Fused plan: straight into the first
// convolution's space-to-depth image (returns true: the forward then skips its input step); otherwise into blob `data`.
bool detect_prelude(fcnod_net* net, const uint8_t* img, int B, int src_h, int src_w, cudaStream_t s) {
    const Blob& in = net->blobs[0];
    net->data_img = img; net->data_img_b = B; net->data_img_h = src_h; net->data_img_w = src_w;
    if (net->fused_prelude) {
        const Step& st = net->steps[0];
        preprocess_s2d_launch(img, B, src_h, src_w, in.H, in.W, net->buffers[st.s2d_buf].ptr, st.s2d_wp, st.s2d_hp, st.s2d_padl,
                              net->pre_scratch, s);
        net->data_valid = false;
        return true;
    }
    preprocess_launch(img, B, src_h, src_w, in.H, in.W, net->data_dev, net->pre_scratch, s);
    net->data_valid = true;
    return false;
}

void copy_detections_out(fcnod_net* net, int B, int32_t* out_det, int32_t* out_label, int32_t* out_weight,
                         int32_t* out_count, cudaStream_t s) {
    int32_t* po = net->pinned_out;
    int32_t* p_det = po;
    int32_t* p_lab = p_det + (size_t)B * FCNOD_MAX_DET * 5;
    int32_t* p_w = p_lab + (size_t)B * FCNOD_MAX_DET;
    int32_t* p_cnt = p_w + (size_t)B * FCNOD_MAX_DET;
    FCNOD_CHECK(B == net->batch, "internal: detection buffers are laid out for another batch size");
    FCNOD_CUDA(cudaMemcpyAsync(p_det, net->det_dev, (size_t)B * (FCNOD_MAX_DET * 7 + 1) * 4, cudaMemcpyDeviceToHost, s));
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

int fcnod_caffemodel_blob(const char* caffemodel_path, const char* layer, int index, int dims[4], int* ndim, float* out,
                          size_t out_count) {
    FCNOD_API_BEGIN
    FCNOD_CHECK(caffemodel_path && layer && dims && ndim, "null argument");
    const WeightMap w = read_caffemodel(caffemodel_path);
    WeightMap::const_iterator it = w.find(layer);
    FCNOD_CHECK(it != w.end(), std::string("caffemodel has no blobs for layer '") + layer + "'");
    FCNOD_CHECK(index >= 0 && index < (int)it->second.size(), "blob index out of range");
    const HostBlob& b = it->second[index];
    FCNOD_CHECK(b.shape.size() <= 4, "blob has more than 4 dimensions");
    *ndim = (int)b.shape.size();
    for (int i = 0; i < 4; ++i) dims[i] = i < *ndim ? b.shape[i] : 1;
    if (out) {
        FCNOD_CHECK(out_count == b.data.size(), "caffemodel_blob: out_count does not match the blob");
        memcpy(out, b.data.data(), b.data.size() * 4);
    }
    FCNOD_API_END
}

int fcnod_datum_decode(const uint8_t* buf, size_t len, int dims[3], int* label, float* out, size_t out_count) {
    FCNOD_API_BEGIN
    FCNOD_CHECK(buf && dims, "null argument");
    const Datum d = decode_datum(buf, len);
    dims[0] = d.channels; dims[1] = d.height; dims[2] = d.width;
    if (label) *label = d.has_label ? d.label : -1;
    if (out) {
        const size_t n = (size_t)d.channels * d.height * d.width;
        FCNOD_CHECK(out_count == n, "datum_decode: out_count must equal channels x height x width");
        if (!d.floats.empty()) memcpy(out, d.floats.data(), n * 4);
        else for (size_t i = 0; i < n; ++i) out[i] = (float)d.bytes[i];
    }
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
    conv_tc2_set_smem_limit();
    FCNOD_CUDA(cudaFuncSetAttribute(lrn_pool_kernel<bf16, 3, 2, 2, true, 4, 8, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    FCNOD_CUDA(cudaFuncSetAttribute(lrn_pool_kernel<bf16, 3, 2, 2, true, 4, 8, 24>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    FCNOD_CUDA(cudaFuncSetAttribute(pool_lrn_kernel<bf16, 3, 2, 2, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
    FCNOD_CUDA(cudaFuncSetAttribute(pool_lrn_tile_kernel<bf16, 3, 2, 2, true, 4, 16>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
    FCNOD_CUDA(cudaStreamCreateWithFlags(&net->stream, cudaStreamNonBlocking));
    {
        int prio_least = 0, prio_greatest = 0;
        FCNOD_CUDA(cudaDeviceGetStreamPriorityRange(&prio_least, &prio_greatest));      // numerically lower = more urgent
        // FCNOD_LANE_PRIORITY=1 gives the lanes descending stream priorities (critical chain first).  Measured neutral to
        // slightly negative on the bench (1.868 vs 1.858 ms per step), so the default is one priority for all.
        static const bool flat_prio = !(getenv("FCNOD_LANE_PRIORITY") && getenv("FCNOD_LANE_PRIORITY")[0] == '1');
        for (int l = 0; l < 3; ++l) {
            const int prio = flat_prio ? prio_least : std::min(prio_least, prio_greatest + l);
            FCNOD_CUDA(cudaStreamCreateWithPriority(&net->lane_stream[l], cudaStreamNonBlocking, prio));
            FCNOD_CUDA(cudaEventCreateWithFlags(&net->join_ev[l], cudaEventDisableTiming));
        }
    }
    FCNOD_CUDA(cudaEventCreateWithFlags(&net->fork_ev, cudaEventDisableTiming));
    for (size_t i = 0; i < net->steps.size(); ++i)
        if (net->steps[i].signal) FCNOD_CUDA(cudaEventCreateWithFlags(&net->steps[i].ev, cudaEventDisableTiming));
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
            if (net->convs[i].d_bias_tab) cudaFree(net->convs[i].d_bias_tab);
        }
        if (net->stream) cudaStreamDestroy(net->stream);
        for (int l = 0; l < 3; ++l) {
            if (net->lane_stream[l]) cudaStreamDestroy(net->lane_stream[l]);
            if (net->join_ev[l]) cudaEventDestroy(net->join_ev[l]);
        }
        if (net->fork_ev) cudaEventDestroy(net->fork_ev);
        for (size_t i = 0; i < net->steps.size(); ++i) if (net->steps[i].ev) cudaEventDestroy(net->steps[i].ev);
        if (net->copy_stream) cudaStreamDestroy(net->copy_stream);
        for (int i = 0; i < 2; ++i) if (net->copy_ev[i]) cudaEventDestroy(net->copy_ev[i]);
        if (net->pinned_big) cudaFreeHost(net->pinned_big);
        for (int i = 0; i < 2; ++i) {
            if (net->slots[i].img) cudaFree(net->slots[i].img);
            if (net->slots[i].pinned) cudaFreeHost(net->slots[i].pinned);
            if (net->slots[i].up) cudaEventDestroy(net->slots[i].up);
            if (net->slots[i].done) cudaEventDestroy(net->slots[i].done);
        }
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
        if (net->n_lanes > 1) {
            o << " {lane " << st.lane;
            for (size_t w = 0; w < st.waits.size(); ++w) o << (w ? "," : " after ") << st.waits[w];
            o << "}";
        }
        if (st.kind == S_CONV_TC) {
            const ConvWeights& cw = net->convs[st.conv];
            o << " [" << (st.tc.halo ? "halo " : "") << (st.tc.flat ? "flat " : "") << (st.tc.pool ? "pool " : "") << (st.tc.lrn ? "lrn " : "") << "CK=" << st.tc.CK << " BN=" << st.tc.BN << "x" << st.tc.n_tiles << " box=" << st.tc.BW
              << "x" << st.tc.BH << " stages=" << st.tc.a_stages << "+" << st.tc.b_stages << "x" << st.tc.b_taps << (st.tc.b_resident ? "R" : "") << (st.tc.pool ? "+" + std::to_string(st.tc.r_stages) + "raw" : std::string()) << (st.tc.dual ? " dual" : "") << " eg=" << st.tc.egroups << " pick=" << st.tc_pick << "/" << st.tc_ncand << " tiles=" << st.tc.m_tiles * st.tc.n_tiles
              << " grid=" << st.tc_grid << " smem=" << st.tc_smem << "]";
        }
        if (st.out >= 0) {
            const Blob& b = net->blobs[st.out];
            o << " -> " << b.name << " (" << b.C << "," << b.H << "," << b.W << ") " << (b.dtype == DT_F32 ? "f32" : "bf16")
              << " buf" << b.buf << "+" << b.c_off << "/" << b.ldc;
            if (st.kind == S_CONV_TC && net->convs[st.conv].s2d) o << " [row K-blocks]";
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
    net->data_valid = true;
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
    FCNOD_CHECK(b.buf != -2, std::string("blob '") + blob + "' is fused away in the bf16 plan and never materialised; "
                "create the net with FCNOD_PRECISION_FP32, or set FCNOD_NO_FUSION=1, to inspect it");
    if (b.buf < 0) {
        if (!net->data_valid) {      // the detect path fed the first convolution directly: materialise `data` now
            FCNOD_CHECK(net->data_img && net->data_img_b == net->batch, "blob 'data' of the last detect call is no longer available");
            FCNOD_CUDA(cudaStreamSynchronize(net->stream));
            preprocess_launch(net->data_img, net->batch, net->data_img_h, net->data_img_w, b.H, b.W, net->data_dev, net->pre_scratch, net->stream);
            FCNOD_CUDA(cudaStreamSynchronize(net->stream));
            net->data_valid = true;
        }
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
    cudaStream_t s = stream ? (cudaStream_t)stream : net->stream;      // NULL = the net's own stream
    for (int i = 0; i < reps; ++i) run_step(net, net->steps[step], net->data_dev, s);
    FCNOD_API_END
}

int fcnod_net_tc_candidates(const fcnod_net* net, int step) {
    if (!net || step < 0 || step >= (int)net->steps.size() || net->steps[step].kind != S_CONV_TC) return 0;
    return net->steps[step].tc_ncand;
}

int fcnod_net_tc_pick(const fcnod_net* net, int step) {
    if (!net || step < 0 || step >= (int)net->steps.size() || net->steps[step].kind != S_CONV_TC) return -1;
    return net->steps[step].tc_pick < 0 ? 0 : net->steps[step].tc_pick;
}

int fcnod_net_set_tc_pick(fcnod_net* net, int step, int pick) {
    FCNOD_API_BEGIN
    FCNOD_CHECK(net && step >= 0 && step < (int)net->steps.size(), "bad step index");
    Step& st = net->steps[step];
    FCNOD_CHECK(st.kind == S_CONV_TC, "step is not a tensor-core convolution");
    FCNOD_CHECK(pick >= 0 && pick < st.tc_ncand, "tiling candidate out of range");
    require_device(net);
    FCNOD_CUDA(cudaDeviceSynchronize());
    drop_graphs(net);
    st.tc_pick = pick;
    setup_tc_step(net, st);
    std::map<int, std::vector<int> >::iterator pk = net->tc_picks.find(net->batch);
    if (pk != net->tc_picks.end()) pk->second[step] = pick;
    FCNOD_API_END
}

int fcnod_net_step_lane(const fcnod_net* net, int step, int* lane, int* deps, int deps_cap, int* n_deps, int* waits, int waits_cap,
                        int* n_waits) {
    FCNOD_API_BEGIN
    FCNOD_CHECK(net && lane && n_deps && n_waits && step >= 0 && step < (int)net->steps.size(), "bad argument");
    const Step& st = net->steps[step];
    *lane = st.lane;
    *n_deps = (int)st.deps.size(); *n_waits = (int)st.waits.size();
    for (int i = 0; i < *n_deps && i < deps_cap && deps; ++i) deps[i] = st.deps[i];
    for (int i = 0; i < *n_waits && i < waits_cap && waits; ++i) waits[i] = st.waits[i];
    FCNOD_API_END
}

int fcnod_net_autotune_candidates(void) {
    const char* e = getenv("FCNOD_AUTOTUNE");
    return e ? atoi(e) : kAutotuneDefault;
}

int fcnod_net_step_outputs(const fcnod_net* net, int step, char* buf, size_t buflen) {
    FCNOD_API_BEGIN
    FCNOD_CHECK(net && buf && buflen > 0 && step >= 0 && step < (int)net->steps.size(), "bad argument");
    const Step& st = net->steps[step];
    std::string o;
    if (st.kind == S_CONV_TC) {
        for (size_t i = 0; i < st.segs.size(); ++i) {
            const int bi = (st.segs[i].flags & SEG_DIRECT) ? st.segs[i].direct_blob : st.segs[i].blob;
            o += (i ? "," : "") + net->blobs[bi].name;
        }
    } else if (st.out >= 0) {
        o = net->blobs[st.out].name;
    }
    FCNOD_CHECK(o.size() + 1 <= buflen, "buffer too small");
    memcpy(buf, o.c_str(), o.size() + 1);
    FCNOD_API_END
}

int fcnod_net_blob_checksum(fcnod_net* net, const char* blob, unsigned long long* out) {
    FCNOD_API_BEGIN
    FCNOD_CHECK(net && blob && out, "null argument");
    require_device(net);
    const Blob& b = net->blobs[blob_id(net, blob)];
    FCNOD_CHECK(b.buf != -2, std::string("blob '") + blob + "' is fused away in this plan");
    FCNOD_CHECK(b.buf >= 0 || b.is_output, "the input blob has no checksum");
    unsigned long long* d = 0;
    FCNOD_CUDA(cudaMalloc(&d, 8));
    cudaError_t e = cudaMemsetAsync(d, 0, 8, net->stream);
    const long npix = (long)net->batch * b.H * b.W;
    if (e == cudaSuccess) {
        if (b.is_output) checksum_kernel<float><<<kNumSMs * 4, 256, 0, net->stream>>>(net->out_dev[b.out_index], npix * b.C, 1, 1, d);
        else if (b.dtype == DT_F32) checksum_kernel<float><<<kNumSMs * 4, 256, 0, net->stream>>>((const float*)view_ptr(net, b), npix, b.C, b.ldc, d);
        else checksum_kernel<bf16><<<kNumSMs * 4, 256, 0, net->stream>>>((const bf16*)view_ptr(net, b), npix, b.C, b.ldc, d);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(net->stream);
    if (e == cudaSuccess) e = cudaMemcpy(out, d, 8, cudaMemcpyDeviceToHost);
    cudaFree(d);
    FCNOD_CUDA(e);
    FCNOD_API_END
}

int fcnod_net_profile_conv_step(fcnod_net* net, int step, unsigned long long* out, size_t count) {
    FCNOD_API_BEGIN
    FCNOD_CHECK(net && out && step >= 0 && step < (int)net->steps.size(), "bad step index");
    FCNOD_CHECK(net->steps[step].kind == S_CONV_TC, "step is not a tensor-core convolution");
    FCNOD_CHECK(count == 4 * 8 * 64, "profile buffer must hold 4*8*64 values");
    require_device(net);
    unsigned long long* d = 0;
    FCNOD_CUDA(cudaMalloc(&d, count * 8));
    FCNOD_CUDA(cudaMemset(d, 0, count * 8));
    Step st = net->steps[step];
    st.tc.prof = d;
    launch_conv_tc2(st, net->stream, true);
    cudaError_t e = cudaStreamSynchronize(net->stream);
    if (e == cudaSuccess) e = cudaMemcpy(out, d, count * 8, cudaMemcpyDeviceToHost);
    cudaFree(d);
    FCNOD_CUDA(e);
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
    net->data_valid = true;
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
    // FCNOD_CHUNK=1: large batches run as two half-batches so that the upload of the second half overlaps the compute of
    // the first one (frames are independent; the network runs at batch B/2 twice on the same buffers).  Opt-in: on the
    // bench box (PCIe gen5, 38.5 MB per 64 frames) the two smaller forwards cost more than the overlap wins
    // (14.8k vs 16.2k frames/s end to end).
    const char* chunk_env = getenv("FCNOD_CHUNK");
    const int nchunk = (B >= 32 && B % 2 == 0 && chunk_env && chunk_env[0] == '1') ? 2 : 1;
    if (nchunk == 2) {
        const int Bc = B / 2;
        reshape(net, Bc);
        const Blob& in = net->blobs[0];
        FCNOD_CHECK(in.C == 3, "detect needs a 3-channel input blob");
        cudaStream_t s = net->stream;
        if (!net->copy_stream) {
            FCNOD_CUDA(cudaStreamCreateWithFlags(&net->copy_stream, cudaStreamNonBlocking));
            for (int i = 0; i < 2; ++i) FCNOD_CUDA(cudaEventCreateWithFlags(&net->copy_ev[i], cudaEventDisableTiming));
        }
        const size_t cbytes = (size_t)Bc * src_h * src_w * 3, ibytes = 2 * cbytes;
        if (net->img_dev_bytes < ibytes) {
            if (net->img_dev) cudaFree(net->img_dev);
            FCNOD_CUDA(cudaMalloc(&net->img_dev, ibytes));
            net->img_dev_bytes = ibytes;
        }
        if (!net->pre_scratch) FCNOD_CUDA(cudaMalloc(&net->pre_scratch, (size_t)net->capacity * 6 * 4));
        ensure_detect_buffers(net);
        const size_t per = (size_t)FCNOD_MAX_DET * 7 + 1, obytes = (size_t)B * per * 4;
        if (net->pinned_big_bytes < obytes) {
            if (net->pinned_big) cudaFreeHost(net->pinned_big);
            FCNOD_CUDA(cudaMallocHost(&net->pinned_big, obytes));
            net->pinned_big_bytes = obytes;
        }
        std::vector<int32_t> hw((size_t)Bc * 2);
        for (int i = 0; i < Bc; ++i) { hw[2 * i] = src_h; hw[2 * i + 1] = src_w; }
        net->srchw_b = 0;      // (the streaming path's cache of what srchw_dev holds)
        FCNOD_CUDA(cudaMemcpyAsync(net->srchw_dev, hw.data(), hw.size() * 4, cudaMemcpyHostToDevice, s));
        for (int c = 0; c < 2; ++c) {
            FCNOD_CUDA(cudaMemcpyAsync(net->img_dev + c * cbytes, images + c * cbytes, cbytes, cudaMemcpyHostToDevice, net->copy_stream));
            FCNOD_CUDA(cudaEventRecord(net->copy_ev[c], net->copy_stream));
        }
        int32_t* p_det = net->pinned_big;
        int32_t* p_lab = p_det + (size_t)B * FCNOD_MAX_DET * 5;
        int32_t* p_w = p_lab + (size_t)B * FCNOD_MAX_DET;
        int32_t* p_cnt = p_w + (size_t)B * FCNOD_MAX_DET;
        for (int c = 0; c < 2; ++c) {
            FCNOD_CUDA(cudaStreamWaitEvent(s, net->copy_ev[c], 0));
            const bool skip = detect_prelude(net, net->img_dev + c * cbytes, Bc, src_h, src_w, s);
            detect_device(net, net->data_dev, net->srchw_dev, *p, s, skip);
            const size_t o = (size_t)c * Bc;
            FCNOD_CUDA(cudaMemcpyAsync(p_det + o * FCNOD_MAX_DET * 5, net->det_dev, (size_t)Bc * FCNOD_MAX_DET * 5 * 4, cudaMemcpyDeviceToHost, s));
            FCNOD_CUDA(cudaMemcpyAsync(p_lab + o * FCNOD_MAX_DET, net->label_dev, (size_t)Bc * FCNOD_MAX_DET * 4, cudaMemcpyDeviceToHost, s));
            FCNOD_CUDA(cudaMemcpyAsync(p_w + o * FCNOD_MAX_DET, net->weight_dev, (size_t)Bc * FCNOD_MAX_DET * 4, cudaMemcpyDeviceToHost, s));
            FCNOD_CUDA(cudaMemcpyAsync(p_cnt + o, net->count_dev, (size_t)Bc * 4, cudaMemcpyDeviceToHost, s));
        }
        FCNOD_CUDA(cudaStreamSynchronize(s));
        memcpy(out_det, p_det, (size_t)B * FCNOD_MAX_DET * 5 * 4);
        memcpy(out_label, p_lab, (size_t)B * FCNOD_MAX_DET * 4);
        memcpy(out_weight, p_w, (size_t)B * FCNOD_MAX_DET * 4);
        memcpy(out_count, p_cnt, (size_t)B * 4);
        return 0;
    }
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
    if (!net->pre_scratch) FCNOD_CUDA(cudaMalloc(&net->pre_scratch, (size_t)net->capacity * 6 * 4));
    ensure_detect_buffers(net);
    std::vector<int32_t> hw((size_t)B * 2);
    for (int i = 0; i < B; ++i) { hw[2 * i] = src_h; hw[2 * i + 1] = src_w; }
    net->srchw_b = 0;
    FCNOD_CUDA(cudaMemcpyAsync(net->srchw_dev, hw.data(), hw.size() * 4, cudaMemcpyHostToDevice, s));
    FCNOD_CUDA(cudaStreamSynchronize(s));   // hw is a stack-lifetime host buffer
    FCNOD_CUDA(cudaMemcpyAsync(net->img_dev, images, ibytes, cudaMemcpyHostToDevice, s));
    const bool skip = detect_prelude(net, net->img_dev, B, src_h, src_w, s);
    detect_device(net, net->data_dev, net->srchw_dev, *p, s, skip);
    copy_detections_out(net, B, out_det, out_label, out_weight, out_count, s);
    FCNOD_API_END
}

int fcnod_detect_submit(fcnod_net* net, const uint8_t* images, int B, int src_h, int src_w, const fcnod_postnet_params* p,
                        int32_t* ticket) {
    FCNOD_API_BEGIN
    FCNOD_CHECK(net && images && p && ticket, "null argument");
    FCNOD_CHECK(B > 0 && src_h > 0 && src_w > 0, "bad batch or source size");
    require_device(net);
    const int in_flight = net->next_ticket - net->oldest_ticket;
    FCNOD_CHECK(in_flight < 2, "two batches already in flight: collect the oldest ticket first");
    if (B != net->batch) {
        FCNOD_CHECK(in_flight == 0, "collect the batches in flight before changing the batch size");
        reshape(net, B);
    }
    const Blob& in = net->blobs[0];
    FCNOD_CHECK(in.C == 3, "detect needs a 3-channel input blob");
    cudaStream_t s = net->stream;
    if (!net->copy_stream) {
        FCNOD_CUDA(cudaStreamCreateWithFlags(&net->copy_stream, cudaStreamNonBlocking));
        for (int i = 0; i < 2; ++i) FCNOD_CUDA(cudaEventCreateWithFlags(&net->copy_ev[i], cudaEventDisableTiming));
    }
    fcnod_net::DetectSlot& sl = net->slots[net->next_ticket & 1];
    FCNOD_CHECK(sl.ticket < 0, "internal: slot still owned by an uncollected ticket");
    if (!sl.up) {
        FCNOD_CUDA(cudaEventCreateWithFlags(&sl.up, cudaEventDisableTiming));
        FCNOD_CUDA(cudaEventCreateWithFlags(&sl.done, cudaEventDisableTiming));
    }
    const size_t ibytes = (size_t)B * src_h * src_w * 3;
    if (sl.img_bytes < ibytes) {
        if (sl.img) cudaFree(sl.img);
        sl.img = 0; sl.img_bytes = 0;
        FCNOD_CUDA(cudaMalloc(&sl.img, ibytes));
        sl.img_bytes = ibytes;
    }
    const size_t obytes = (size_t)B * (2 + FCNOD_MAX_DET * 7 + 1) * 4;
    if (sl.pinned_bytes < obytes) {
        if (sl.pinned) cudaFreeHost(sl.pinned);
        sl.pinned = 0; sl.pinned_bytes = 0;
        FCNOD_CUDA(cudaMallocHost(&sl.pinned, obytes));
        sl.pinned_bytes = obytes;
    }
    if (!net->pre_scratch) FCNOD_CUDA(cudaMalloc(&net->pre_scratch, (size_t)net->capacity * 6 * 4));
    ensure_detect_buffers(net);
    // forward + post-net as one graph (built on first use, rebuilt when the batch or the thresholds change)
    static const bool no_sgraph = getenv("FCNOD_NO_STREAM_GRAPH") != nullptr;
    static const bool no_slot_prelude = getenv("FCNOD_NO_SLOT_PRELUDE") != nullptr;
    auto capture = [&](cudaGraphExec_t* out, bool skip_input) {
        postnet_prepare(p->eps);          // per-device one-time set-up happens outside the capture (submit may be the first call)
        cudaGraph_t g = 0;
        FCNOD_CUDA(cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal));
        try {
            detect_device(net, net->data_dev, net->srchw_dev, *p, s, skip_input);
        } catch (...) {
            cudaStreamEndCapture(s, &g);
            if (g) cudaGraphDestroy(g);
            throw;
        }
        FCNOD_CUDA(cudaStreamEndCapture(s, &g));
        cudaError_t ge = cudaGraphInstantiate(out, g, 0);
        cudaGraphDestroy(g);
        FCNOD_CUDA(ge);
    };
    int32_t* hw = sl.pinned;
    for (int i = 0; i < B; ++i) { hw[2 * i] = src_h; hw[2 * i + 1] = src_w; }
    // the per-frame source sizes on the device change only when the caller's frames do: no copy per batch
    const bool hw_same = net->srchw_b == B && net->srchw_h == src_h && net->srchw_w == src_w;
    net->srchw_b = B; net->srchw_h = src_h; net->srchw_w = src_w;
    if (net->fused_prelude && !no_sgraph && !no_slot_prelude) {
        // Slot-private prelude: the frames are uploaded AND turned into this slot's space-to-depth image on the copy stream,
        // overlapping the previous batch's network; the slot's graph has conv1 read that image.
        const Step& s0 = net->steps[0];
        Buffer& sb = net->buffers[s0.s2d_buf];
        if (!sl.s2d) {
            const size_t bytes = (size_t)net->capacity * sb.C * sb.H * sb.W * dsize(sb.dtype);
            FCNOD_CUDA(cudaMalloc(&sl.s2d, bytes));
            FCNOD_CUDA(cudaMemset(sl.s2d, 0, bytes));            // the zero padding of the image is written once
            FCNOD_CUDA(cudaMalloc(&sl.mm, (size_t)net->capacity * 6 * 4));
        }
        if (!sl.graph || memcmp(&sl.graph_p, p, sizeof(*p)) != 0) {
            if (sl.graph) { cudaGraphExecDestroy(sl.graph); sl.graph = 0; }
            void* const keep = sb.ptr;
            sb.ptr = sl.s2d;                                      // conv1's operand = this slot's image, for the capture only
            for (size_t i = 0; i < net->steps.size(); ++i)
                if (net->steps[i].kind == S_CONV_TC && net->convs[net->steps[i].conv].s2d) setup_tc_step(net, net->steps[i]);
            try {
                capture(&sl.graph, true);
            } catch (...) {
                sb.ptr = keep;
                for (size_t i = 0; i < net->steps.size(); ++i)
                    if (net->steps[i].kind == S_CONV_TC && net->convs[net->steps[i].conv].s2d) setup_tc_step(net, net->steps[i]);
                throw;
            }
            sb.ptr = keep;
            for (size_t i = 0; i < net->steps.size(); ++i)
                if (net->steps[i].kind == S_CONV_TC && net->convs[net->steps[i].conv].s2d) setup_tc_step(net, net->steps[i]);
            sl.graph_p = *p;
        }
        FCNOD_CUDA(cudaMemcpyAsync(sl.img, images, ibytes, cudaMemcpyHostToDevice, net->copy_stream));
        preprocess_s2d_launch(sl.img, B, src_h, src_w, in.H, in.W, sl.s2d, s0.s2d_wp, s0.s2d_hp, s0.s2d_padl, sl.mm, net->copy_stream);
        FCNOD_CUDA(cudaEventRecord(sl.up, net->copy_stream));
        net->data_img = sl.img; net->data_img_b = B; net->data_img_h = src_h; net->data_img_w = src_w;
        net->data_valid = false;
        if (!hw_same) FCNOD_CUDA(cudaMemcpyAsync(net->srchw_dev, hw, (size_t)B * 2 * 4, cudaMemcpyHostToDevice, s));
        FCNOD_CUDA(cudaStreamWaitEvent(s, sl.up, 0));
        FCNOD_CUDA(cudaGraphLaunch(sl.graph, s));
    } else {
        if (!no_sgraph && (!net->sgraph || memcmp(&net->sgraph_p, p, sizeof(*p)) != 0)) {
            if (net->sgraph) { cudaGraphExecDestroy(net->sgraph); net->sgraph = 0; }
            capture(&net->sgraph, net->fused_prelude);
            net->sgraph_p = *p;
        }
        // upload on the copy stream: overlaps whatever the compute stream is still doing for the previous ticket.
        // The slot's previous reader (the prelude of ticket-2) finished before that ticket was collected.
        FCNOD_CUDA(cudaMemcpyAsync(sl.img, images, ibytes, cudaMemcpyHostToDevice, net->copy_stream));
        FCNOD_CUDA(cudaEventRecord(sl.up, net->copy_stream));
        FCNOD_CUDA(cudaMemcpyAsync(net->srchw_dev, hw, (size_t)B * 2 * 4, cudaMemcpyHostToDevice, s));
        FCNOD_CUDA(cudaStreamWaitEvent(s, sl.up, 0));
        const bool skip = detect_prelude(net, sl.img, B, src_h, src_w, s);
        if (net->sgraph) FCNOD_CUDA(cudaGraphLaunch(net->sgraph, s));
        else detect_device(net, net->data_dev, net->srchw_dev, *p, s, skip);
    }
    int32_t* p_det = sl.pinned + (size_t)B * 2;
    int32_t* p_lab = p_det + (size_t)B * FCNOD_MAX_DET * 5;
    int32_t* p_w = p_lab + (size_t)B * FCNOD_MAX_DET;
    int32_t* p_cnt = p_w + (size_t)B * FCNOD_MAX_DET;
    (void)p_lab; (void)p_w; (void)p_cnt;       // one block on the device, one block in the pinned buffer: one copy
    FCNOD_CUDA(cudaMemcpyAsync(p_det, net->det_dev, (size_t)B * (FCNOD_MAX_DET * 7 + 1) * 4, cudaMemcpyDeviceToHost, s));
    FCNOD_CUDA(cudaEventRecord(sl.done, s));
    sl.B = B;
    sl.ticket = net->next_ticket;
    *ticket = net->next_ticket++;
    FCNOD_API_END
}

int fcnod_detect_collect(fcnod_net* net, int32_t ticket, int32_t* out_det, int32_t* out_label, int32_t* out_weight,
                         int32_t* out_count) {
    FCNOD_API_BEGIN
    FCNOD_CHECK(net && out_det && out_label && out_weight && out_count, "null argument");
    require_device(net);
    FCNOD_CHECK(net->oldest_ticket < net->next_ticket, "no batch in flight");
    FCNOD_CHECK(ticket == net->oldest_ticket, "tickets are collected in submission order");
    fcnod_net::DetectSlot& sl = net->slots[ticket & 1];
    FCNOD_CHECK(sl.ticket == ticket, "internal: ticket / slot mismatch");
    // whatever happens below, the ticket is consumed (a device fault is sticky anyway)
    sl.ticket = -1;
    ++net->oldest_ticket;
    FCNOD_CUDA(cudaEventSynchronize(sl.done));
    const int B = sl.B;
    const int32_t* p_det = sl.pinned + (size_t)B * 2;
    const int32_t* p_lab = p_det + (size_t)B * FCNOD_MAX_DET * 5;
    const int32_t* p_w = p_lab + (size_t)B * FCNOD_MAX_DET;
    const int32_t* p_cnt = p_w + (size_t)B * FCNOD_MAX_DET;
    memcpy(out_det, p_det, (size_t)B * FCNOD_MAX_DET * 5 * 4);
    memcpy(out_label, p_lab, (size_t)B * FCNOD_MAX_DET * 4);
    memcpy(out_weight, p_w, (size_t)B * FCNOD_MAX_DET * 4);
    memcpy(out_count, p_cnt, (size_t)B * 4);
    FCNOD_API_END
}

}  // extern "C"
