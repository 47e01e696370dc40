// Host-side network definition readers: Caffe prototxt (text) + caffemodel (binary protobuf).
//
// Replaces what `caffe.Net(proto, weights, caffe.TEST)` does at
// /root/reference/scripts/fcn_object_detector.py:317 for the layer types that
// /root/reference/models/deploy.prototxt:1-2178 uses.  No protoc / libprotobuf:
// the text format is parsed by a small recursive-descent reader and the binary
// format by a wire-format walker (field numbers from caffe.proto, listed in
// netdef.cpp).
#pragma once
#include <stdint.h>

#include <map>
#include <string>
#include <vector>

namespace fcnod {

enum LayerType { L_CONV, L_RELU, L_POOL, L_LRN, L_CONCAT, L_POWER, L_DROPOUT, L_SIGMOID, L_UNKNOWN };

struct LayerDef {
    std::string name, type_name;
    LayerType type = L_UNKNOWN;
    std::vector<std::string> bottom, top;
    // Convolution / Pooling
    int num_output = 0, kernel = 1, stride = 1, pad = 0;
    bool bias_term = true;
    bool pool_max = true;
    // LRN
    int local_size = 5;
    float alpha = 1.f, beta = 0.75f, lrn_k = 1.f;
    // Power
    float power = 1.f, scale = 1.f, shift = 0.f;
    // training side: Dropout ratio, and the `param { lr_mult decay_mult }` records of a Convolution (weights, bias);
    // caffe.proto defaults where a record is absent
    float dropout_ratio = 0.5f;
    float lr_mult[2] = {1.f, 1.f}, decay_mult[2] = {1.f, 1.f};
};

// Layers of models/train_val.prototxt that are not part of the network body: the two LMDB Data layers (:1-26), the label
// Slice (:36-52), the Eltwise PROD normalisers / masks (:53-72, :2237-2266) and the two losses (:2267-2281).  The trainer
// implements that head as one fused kernel (train_head.cu); the parser records what the file says so the trainer can
// verify that it IS the head it implements.
struct TrainSideLayer {
    std::string name, type_name;
    std::vector<std::string> bottom, top;
    std::vector<int> slice_points;
    std::string eltwise_op;
    float loss_weight = 1.f;
};

struct NetDef {
    std::string input_name;
    int input_dims[4] = {0, 0, 0, 0};
    std::vector<LayerDef> layers;
    std::vector<TrainSideLayer> train_side;     // train form only
};

struct HostBlob {
    std::vector<int> shape;
    std::vector<float> data;
};
typedef std::map<std::string, std::vector<HostBlob> > WeightMap;

// caffe.proto Datum: one record of the LMDBs that feed the layer's two bottoms
// (/root/reference/scripts/misc/create_training_lmdb.py:172-185 writes them with caffe.io.array_to_datum).
struct Datum {
    int channels = 0, height = 0, width = 0, label = 0;
    bool has_label = false, encoded = false;
    std::vector<uint8_t> bytes;     // Datum.data (uint8 arrays)
    std::vector<float> floats;      // Datum.float_data (every other dtype; packed or not)
};
Datum decode_datum(const uint8_t* buf, size_t len);

// All throw std::runtime_error with a message on malformed input.
// train = false: the deploy form (`input:` + 4-d shape, body layers only; anything else is refused).
// train = true : also accepts the train_val form: no `input:` needed (the input is the bottom of the first body layer, dims
//                come from the caller), Data / Slice / Eltwise / L1Loss / EuclideanLoss layers go to NetDef::train_side.
NetDef parse_prototxt_file(const std::string& path, bool train = false);
NetDef parse_prototxt_text(const std::string& text, bool train = false);
WeightMap read_caffemodel(const std::string& path);

}  // namespace fcnod
