// Internal launch interfaces shared between translation units.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/fcnod_b200.h"

namespace fcnod {

// postnet.cu
size_t postnet_scratch_bytes(int B, int c_cvg);
void postnet_launch(const float* coverage, const float* bboxes, int B, int c_cvg, int c_bbox, int gh, int gw,
                    const int32_t* src_hw, const fcnod_postnet_params& p, void* scratch, int32_t* out_det,
                    int32_t* out_label, int32_t* out_weight, int32_t* out_count, cudaStream_t stream);
int postnet_num_launches();
// per-device one-time set-up (kernel attribute, SimilarRects table for `eps`): call before capturing a graph
void postnet_prepare(double eps);

// preproc.cu
void preprocess_launch(const uint8_t* images_dev, int B, int src_h, int src_w, int net_h, int net_w, float* data_dev,
                       void* scratch, cudaStream_t stream);
int preprocess_num_launches();
// prelude straight into the padded, pre-swizzled bf16 space-to-depth image of the first convolution (no fp32 `data`)
void preprocess_s2d_launch(const uint8_t* images_dev, int B, int src_h, int src_w, int net_h, int net_w, void* s2d_dev, int Wp,
                           int Hp, int padl, void* scratch, cudaStream_t stream);

// photo.cu: in -> out (tmp is scratch); sizes may be null
void photo_launch(const uint8_t* in, uint8_t* out, uint8_t* tmp, const int32_t* sizes, const fcnod_photo_draw* draws, int N, int H,
                  int W, cudaStream_t stream);

}  // namespace fcnod
