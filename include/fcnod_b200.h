/* fcnod_b200 -- C ABI of the B200 (sm_100a) detection hot path.
 *
 * Drop-in boundary for the two places where iKrishneel/fcn_object_detector
 * hands work to an engine (citations relative to /root/reference):
 *
 *   B2  detector node -> pycaffe      scripts/fcn_object_detector.py
 *         caffe.Net(proto, weights, caffe.TEST)              :317   fcnod_net_create
 *         net.blobs['data'].reshape(B,3,H,W)                 :199,:328  fcnod_net_reshape
 *         net.blobs['data'].data[...] = x ; net.forward()    :82,:87    fcnod_net_forward{,_device}
 *         net.blobs[name].data                               :89-90     fcnod_net_blob_read
 *         demean_rgb_image + cv.resize + transpose           :79-82,:407-413  fcnod_preprocess
 *         gridbox_to_boxes + vote_boxes(cv.groupRectangles)
 *           + resize_detection, per class                    :104-132,:337-405  fcnod_postnet{,_device}
 *         whole run_detector body                            :54-132    fcnod_detect
 *
 *   B1  Caffe Python layer -> numpy   scripts/data_argumentation_layer/
 *         ArgumentationEngine.bounding_box_parameterized_labels   argumentation_engine.py:69-109   fcnod_label_raster
 *         random_argumentation + resize_image_and_labels + labels argumentation_engine.py:114-303  fcnod_aug_forward
 *         (layer contract README.md:60-76; data_argumentation_layer.py:16-124)
 *
 * Conventions: every function returns 0 on success, non-zero on failure
 * (fcnod_last_error() gives a message for the calling thread).  "_device"
 * entry points take DEVICE pointers and a cudaStream_t (passed as void*), do
 * not synchronise, and are CUDA-graph capturable; the others take HOST
 * buffers, copy in/out and synchronise.  Blob layout at this boundary is
 * Caffe's: NCHW float32.  Handles are not thread-safe.  There is no CPU
 * fallback: without a usable CUDA device every compute call fails.
 */
#ifndef FCNOD_B200_H
#define FCNOD_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FCNOD_PRECISION_FP32 0 /* SIMT fp32 everywhere (parity mode, 1e-5) */
#define FCNOD_PRECISION_BF16 1 /* bf16 activations/weights, fp32 accumulate, tcgen05 convolutions */

#define FCNOD_MAX_DET 64 /* detections STORED per frame; out_count reports how many there were (see fcnod_postnet_device) */

typedef struct fcnod_net fcnod_net;

int fcnod_version(void);
/* Page-locked host memory for the buffers handed to the HOST-buffer entry points: their copies are then true DMA at
 * full PCIe rate (Caffe's own blobs are page-locked in GPU mode; numpy arrays are not). */
int fcnod_host_alloc(size_t bytes, void** out);
int fcnod_host_free(void* p);
const char* fcnod_last_error(void);
int fcnod_device_count(int* count);

/* ---- B2: network ------------------------------------------------------- */
/* device >= 0: CUDA device ordinal.  device < 0: PLAN-ONLY handle (parses + plans on the host, no
 * CUDA call is made; fcnod_net_describe / dims work, every compute entry point fails). */
int fcnod_net_create(const char* prototxt_path, const char* caffemodel_path, int device, int precision,
                     fcnod_net** out);
int fcnod_net_destroy(fcnod_net* net);
int fcnod_net_reshape(fcnod_net* net, int batch);
int fcnod_net_input_dims(const fcnod_net* net, int dims[4]);            /* N,C,H,W of blob `data` */
int fcnod_net_blob_dims(const fcnod_net* net, const char* blob, int dims[4]);
int fcnod_net_num_blobs(const fcnod_net* net);
const char* fcnod_net_blob_name(const fcnod_net* net, int index);
/* Plan introspection (host only, no CUDA work): one line per planned kernel launch. */
int fcnod_net_describe(const fcnod_net* net, char* buf, size_t buflen);
/* Number of kernel launches one forward issues. */
int fcnod_net_launches_per_forward(const fcnod_net* net);
/* Host NCHW float32 in (N*3*H*W), runs the whole net, synchronises. */
int fcnod_net_forward(fcnod_net* net, const float* data_host);
/* Device NCHW float32 in; async on `stream`. */
int fcnod_net_forward_device(fcnod_net* net, const float* data_dev, void* stream);
/* Copy any blob out as NCHW float32 (count must equal N*C*H*W). */
int fcnod_net_blob_read(fcnod_net* net, const char* blob, float* host_out, size_t count);
/* Device address of an OUTPUT blob (NCHW float32), valid until reshape/destroy. */
int fcnod_net_output_device(fcnod_net* net, const char* blob, const float** dev_ptr);
/* Run only the n-th planned launch `reps` times (profiling aid); stream NULL = the net's own stream. */
int fcnod_net_run_step(fcnod_net* net, int step, int reps, void* stream);
/* Tensor-core tiling of a planned convolution launch.  The planner enumerates the legal tilings of every launch
 * (fcnod_net_tc_candidates: how many; 0 = not a tensor-core step), ranks them with a cost model, and on the first
 * reshape to a batch size times the best-ranked ones on the device and keeps the fastest; the picks are cached per
 * batch size, so a handle always runs the same plan for the same batch (fcnod_net_describe prints `pick=i/n`).
 * fcnod_net_set_tc_pick forces candidate `pick` for the current batch (parity tests sweep every candidate the
 * autotuner may choose); every candidate accumulates in the same K order, i.e. all of them produce the same bits. */
int fcnod_net_tc_candidates(const fcnod_net* net, int step);
int fcnod_net_tc_pick(const fcnod_net* net, int step);
int fcnod_net_set_tc_pick(fcnod_net* net, int step, int pick);
/* Comma-separated names of the blobs planned launch `step` writes (a fused launch writes several). */
int fcnod_net_step_outputs(const fcnod_net* net, int step, char* buf, size_t buflen);
/* Lane plan of launch `step` (host only): the lane (stream / graph branch) it runs on, every earlier launch it depends on
 * through a buffer region (RAW / WAR / WAW), and the launches on OTHER lanes whose completion event it waits for. */
int fcnod_net_step_lane(const fcnod_net* net, int step, int* lane, int* deps, int deps_cap, int* n_deps, int* waits, int waits_cap,
                        int* n_waits);
/* How many tiling candidates the autotuner times per launch (FCNOD_AUTOTUNE, default 12; <= 1: none). */
int fcnod_net_autotune_candidates(void);
/* Position-sensitive 64-bit checksum of a materialised blob, computed on the device (test aid). */
int fcnod_net_blob_checksum(fcnod_net* net, const char* blob, unsigned long long* out);
/* Profiling aid: run a tensor-core convolution step once and return the clock64 timeline of its CTA 0:
 * out[4 roles (TMA producer: B-load issue; MMA issuer: B-load landed; first epilogue warp; MMA issuer: tap issued)]
 * [8 tiles][64 events] (count must be 2048). */
int fcnod_net_profile_conv_step(fcnod_net* net, int step, unsigned long long* out, size_t count);

/* ---- B2: post-net (threshold -> decode -> groupRectangles -> filter -> rescale) */
typedef struct fcnod_postnet_params {
    float prob_thresh;      /* ~detection_threshold (0.5)   fcn_object_detector.py:33 */
    int min_boxes;          /* ~min_boxes (3)                                     :34 */
    double eps;             /* ~nms_eps (0.2)                                     :35 */
    int net_w, net_h;       /* 448, 448                                           :324-326 */
    int stride;             /* 16 for the deploy net; the literal upstream value is 8 (:360) */
    int skip_background;    /* 1: drop coverage channel 0                         :89 */
    int min_height;         /* 20                                                 :346 */
} fcnod_postnet_params;

void fcnod_postnet_default_params(fcnod_postnet_params* p);

/* coverage [B][c_cvg][gh][gw], bboxes [B][c_bbox][gh][gw] float32 (device).
 * src_hw [B][2] int32 = source image (height,width) per frame, or NULL for net size.
 * out_det [B][FCNOD_MAX_DET][5] int32 = x1,y1,x2,y2,int(log(weight)) in source pixels,
 * out_label [B][FCNOD_MAX_DET] int32, out_weight [B][FCNOD_MAX_DET] int32 (cluster size),
 * out_count [B] int32 = the number of detections the reference would return for the frame (it has no cap,
 * scripts/fcn_object_detector.py:104-124).  Only the first FCNOD_MAX_DET of them (class order, then cluster order;
 * at most FCNOD_MAX_DET per class) are stored: out_count[b] > FCNOD_MAX_DET reports the truncation, and the caller
 * reads min(out_count[b], FCNOD_MAX_DET) rows.  scratch: fcnod_postnet_scratch_bytes(B, c_cvg) bytes of device memory.
 * The first post-net call on a device (or fcnod_postnet_prepare) must happen outside CUDA-graph capture. */
size_t fcnod_postnet_scratch_bytes(int B, int c_cvg);
/* One-time per-device set-up for these thresholds (kernel attribute, SimilarRects table of p->eps); optional. */
int fcnod_postnet_prepare(const fcnod_postnet_params* p);
int fcnod_postnet_device(const float* coverage, const float* bboxes, int B, int c_cvg, int c_bbox, int gh, int gw,
                         const int32_t* src_hw, const fcnod_postnet_params* p, void* scratch,
                         int32_t* out_det, int32_t* out_label, int32_t* out_weight, int32_t* out_count,
                         void* stream);
/* Decode alone = gridbox_to_boxes (scripts/fcn_object_detector.py:357-394) for every (frame, foreground class):
 * coverage >= prob_thresh, candidates in row-major cell order, boxes = float32 regressor outputs + integer cell
 * origin evaluated in float64 (x1, y1, x2, y2), no clustering.  Device buffers:
 * out_boxes [B][ncls][cap][4] float64, out_cell [B][ncls][cap] int32 (cell index y*gw + x, i.e. the `mask`),
 * out_score [B][ncls][cap] float32 (the coverage value), out_count [B][ncls] int32 = candidates found (may exceed
 * cap: only the first cap are stored).  ncls = c_cvg - skip_background.  Async on `stream`. */
int fcnod_decode_device(const float* coverage, const float* bboxes, int B, int c_cvg, int c_bbox, int gh, int gw,
                        const fcnod_postnet_params* p, int cap, double* out_boxes, int32_t* out_cell, float* out_score,
                        int32_t* out_count, void* stream);
int fcnod_decode(const float* coverage, const float* bboxes, int B, int c_cvg, int c_bbox, int gh, int gw,
                 const fcnod_postnet_params* p, int cap, double* out_boxes, int32_t* out_cell, float* out_score,
                 int32_t* out_count);   /* HOST buffers */
/* Same with HOST buffers (copies in / out through the calling thread's cached device workspace, synchronises). */
int fcnod_postnet(const float* coverage, const float* bboxes, int B, int c_cvg, int c_bbox, int gh, int gw,
                  const int32_t* src_hw, const fcnod_postnet_params* p,
                  int32_t* out_det, int32_t* out_label, int32_t* out_weight, int32_t* out_count);

/* ---- B2: prelude: BGR uint8 HxWx3 -> blob `data` (1x3xnet_hxnet_w float32) --------------- */
/* images: B frames of identical size src_h x src_w x 3 (host, uint8, BGR). data_out host NCHW float32. */
int fcnod_preprocess(const uint8_t* images, int B, int src_h, int src_w, int net_h, int net_w, float* data_out);
int fcnod_preprocess_device(const uint8_t* images_dev, int B, int src_h, int src_w, int net_h, int net_w,
                            float* data_dev, void* scratch /* 6*B int32 */, void* stream);

/* ---- B2: whole run_detector body on B frames (host buffers) ---------------------------------
 * images as for fcnod_preprocess; outputs as for fcnod_postnet (boxes in source pixels).
 * Uses blobs `coverage` / `bboxes`. */
int fcnod_detect(fcnod_net* net, const uint8_t* images, int B, int src_h, int src_w, const fcnod_postnet_params* p,
                 int32_t* out_det, int32_t* out_label, int32_t* out_weight, int32_t* out_count);
/* Streaming form of fcnod_detect for a caller that feeds batch after batch (the detector node's image
 * callback, scripts/fcn_object_detector.py:134-147, batched): submit() enqueues the upload, the prelude, the
 * network and the post-net of one batch and returns a ticket at once; collect() blocks until that batch's
 * detections are in host memory.  Two batches may be in flight: the upload of batch i+1 (own copy stream, own
 * device image slot) overlaps the compute of batch i.  `images` must stay valid until its collect() and
 * should be page-locked, otherwise the upload is synchronous.  Tickets are collected in submission order;
 * the batch size may only change while nothing is in flight. */
int fcnod_detect_submit(fcnod_net* net, const uint8_t* images, int B, int src_h, int src_w,
                        const fcnod_postnet_params* p, int32_t* ticket);
int fcnod_detect_collect(fcnod_net* net, int32_t ticket, int32_t* out_det, int32_t* out_label, int32_t* out_weight,
                         int32_t* out_count);
/* Same, starting from blob `data` already in HOST memory (B*3*H*W float32). */
int fcnod_detect_from_data(fcnod_net* net, const float* data_host, int B, const fcnod_postnet_params* p,
                           int32_t* out_det, int32_t* out_label, int32_t* out_weight, int32_t* out_count);
/* Same with `data` resident on the DEVICE; results stay on the device in the net's own
 * buffers (pointers returned); async on `stream`; CUDA-graph capturable. */
int fcnod_detect_device(fcnod_net* net, const float* data_dev, int B, const fcnod_postnet_params* p,
                        const int32_t** det_dev, const int32_t** label_dev, const int32_t** weight_dev,
                        const int32_t** count_dev, void* stream);
/* Capture fcnod_detect_device for the current batch into a CUDA graph and replay it. */
int fcnod_detect_graph_build(fcnod_net* net, const float* data_dev, int B, const fcnod_postnet_params* p, void* stream);
int fcnod_detect_graph_launch(fcnod_net* net, void* stream);

/* ---- on-disk formats (host only, no CUDA) ------------------------------------------------------------------
 * One weight blob of a .caffemodel as the network loader sees it (float32): V2 `layer` / V1 `layers` records,
 * BlobProto.data (packed or not), double_data, `shape` or legacy num/channels/height/width, and NVCaffe's
 * raw_data_type / raw_data (float32, IEEE half, float64).  out may be NULL to query dims / ndim. */
int fcnod_caffemodel_blob(const char* caffemodel_path, const char* layer, int index, int dims[4], int* ndim, float* out,
                          size_t out_count);

/* ---- B1: wire format of the layer's two bottoms ------------------------------------------------------------
 * One caffe.proto `Datum` record as create_training_lmdb.py:172-185 writes it (caffe.io.array_to_datum): dims =
 * channels, height, width AS STORED (the image datum is written un-transposed: an HxWx3 image arrives as channels = H,
 * height = W, width = 3); uint8 `data` and (packed or unpacked) `float_data` both come back as float32.  out may be NULL
 * to query the dims; label = -1 when the record has none.  Encoded (compressed) datums are refused. */
int fcnod_datum_decode(const uint8_t* buf, size_t len, int dims[3], int* label, float* out, size_t out_count);

/* ---- B1: training-side layer ------------------------------------------------------------- */
/* Label rasteriser.  rects [N][R][4] int32 (x,y,w,h at net resolution), labels [N][R] int32,
 * nrect [N] int32 (<= R).  Outputs NCHW float32: fg [N][C][gh][gw]; bbox,size,obj,cvg [N][4C][gh][gw]. */
int fcnod_label_raster_device(const int32_t* rects, const int32_t* labels, const int32_t* nrect, int N, int R,
                              int net_h, int net_w, int stride, int num_classes, float* fg, float* bbox,
                              float* size, float* obj, float* cvg, void* stream);
int fcnod_label_raster(const int32_t* rects, const int32_t* labels, const int32_t* nrect, int N, int R,
                       int net_h, int net_w, int stride, int num_classes, float* fg, float* bbox,
                       float* size, float* obj, float* cvg);

typedef struct fcnod_aug_draw {   /* host-made random draws for one image (argumentation_engine.py:146,165-166,200-202) */
    int32_t flip;                 /* randint(-1,2): -1/0/1 flip, 2 none */
    int32_t shift_x, shift_y;     /* randint(0,int(w/2)), randint(0,int(h/2)) */
    int32_t sign_x, sign_y;       /* randint(0,1) x2 */
    int32_t _pad;
    double f1, f2;                /* uniform(1, scale_x), uniform(1, scale_y) */
} fcnod_aug_draw;

/* Whole layer forward (README.md:60-76 contract).  data_in [N][3][src_h][src_w] float32 (pixel values),
 * label [N][5] float32 = x,y,w,h,class.  tops: data [N][3][net_h][net_w]; fg [N][C][gh][gw];
 * bbox/size/obj/cvg [N][4C][gh][gw].  rects_out [N][4] int32 (augmented rect, may be NULL).
 * scratch >= fcnod_aug_scratch_bytes(N). */
size_t fcnod_aug_scratch_bytes(int N);
int fcnod_aug_forward_device(const float* data_in, const float* label, const fcnod_aug_draw* draws, int N,
                             int src_h, int src_w, int net_h, int net_w, int stride, int num_classes,
                             float* top_data, float* fg, float* bbox, float* size, float* obj, float* cvg,
                             int32_t* rects_out, void* scratch, void* stream);
int fcnod_aug_forward(const float* data_in, const float* label, const fcnod_aug_draw* draws, int N,
                      int src_h, int src_w, int net_h, int net_w, int stride, int num_classes,
                      float* top_data, float* fg, float* bbox, float* size, float* obj, float* cvg,
                      int32_t* rects_out);

/* Allocation-free variant of fcnod_aug_forward_device.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that may not allocate device memory (stream capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  Same arguments, results and error behaviour as fcnod_aug_forward_device minus `scratch`: the
 * per-image geometry records live in a fixed `static __device__` array inside the library (FCNOD_AUG_NOALLOC_MAX_N * 48 B
 * = 192 KiB, held by every GPU that loads the library until exit).  N > FCNOD_AUG_NOALLOC_MAX_N is refused with an error
 * string that names the limit.  The call makes no cudaMalloc / cudaFree / allocating library call; the first call on a
 * device looks the array's address up (cudaGetSymbolAddress) with that device current.  All calls on a device share
 * the array: calls that overlap on the GPU (other streams or threads) race -- serialise them.  Nothing needs clearing
 * between calls (every record in [0, N) is rewritten by the first kernel of each call). */
#define FCNOD_AUG_NOALLOC_MAX_N 4096
int fcnod_aug_forward_device_noalloc(const float* data_in, const float* label, const fcnod_aug_draw* draws, int N,
                                     int src_h, int src_w, int net_h, int net_w, int stride, int num_classes,
                                     float* top_data, float* fg, float* bbox, float* size, float* obj, float* cvg,
                                     int32_t* rects_out, void* stream);

/* ---- photometric augmentation of the layer (argumentation_engine.py:308-322, imgaug chain) ----
 * Host-made draws for one image.  blur_kind: 0 none, 1 GaussianBlur (gauss_radius = int(4 sigma + 0.5) <= 12, gauss_w[i] =
 * normalised weight at distance i, computed on the host as scipy.ndimage does), 2 AverageBlur (blur_k in 2..7),
 * 3 MedianBlur (blur_k in {3, 5, 7}).  sharpen = the 3x3 filter2D matrix (1-a) * identity + a * [[-1..],[-1, 8+l, -1],[..]]
 * in float32, row-major (identity = off).  add / mul per channel (equal entries when not per_channel).  gray_alpha 0 = off. */
typedef struct fcnod_photo_draw {
    int32_t blur_kind, blur_k, gauss_radius, _pad0;
    double gauss_w[13];
    float sharpen[9];
    int32_t add[3];
    float mul[3];
    int32_t _pad1;
    double gray_alpha;
} fcnod_photo_draw;
/* images / out / tmp: [N][H][W][3] uint8 (three distinct buffers); sizes [N][2] int32 = valid (h, w) of each image inside
 * the common H x W pitch, rows packed at the image's own width (NULL: every image is H x W).  Three launches. */
int fcnod_photo_augment_device(const uint8_t* images, const int32_t* sizes, const fcnod_photo_draw* draws, int N, int H, int W,
                               uint8_t* out, uint8_t* tmp, void* stream);
int fcnod_photo_augment(const uint8_t* images, const int32_t* sizes, const fcnod_photo_draw* draws, int N, int H, int W,
                        uint8_t* out);   /* HOST buffers */

/* ---- copy-paste synthesis: the compositing loop of ArgumentationEngineMapping.argument (argumentation_engine.py:723-728)
 * One placed proposal: its w x h BGR patch starts at pixel `roi_offset` of the packed ROI buffer (`rois` 3 bytes per
 * pixel, `roi_masks` 1 byte per pixel = mask[..., 0]), and lands with its top-left corner at (cx, cy) of background
 * `image`; pixels whose mask byte is > 0 and that fall inside the background overwrite it, and the label mask gets
 * label + 1.  Proposals of one image are applied in table order (later ones win); paste_begin [N + 1] indexes the table. */
typedef struct fcnod_paste {
    int32_t roi_offset, w, h, cx, cy, label;
} fcnod_paste;
int fcnod_composite_device(uint8_t* images /*[N][H][W][3] in/out*/, uint8_t* masks /*[N][H][W] in/out*/, int N, int H, int W,
                           const uint8_t* rois, const uint8_t* roi_masks, const fcnod_paste* pastes, const int32_t* paste_begin,
                           void* stream);
int fcnod_composite(uint8_t* images, uint8_t* masks, int N, int H, int W, const uint8_t* rois, const uint8_t* roi_masks,
                    size_t roi_pixels, const fcnod_paste* pastes, const int32_t* paste_begin);   /* HOST buffers */

/* The layer forward WITH the photometric stage where the reference has it (argumentation_engine.py:176: after flip + crop,
 * before demean + resize): the flipped / cropped source is materialised as a uint8 crop, run through the chain above with
 * every crop's own borders, and demeaned / resampled from there.  photo = NULL is fcnod_aug_forward*. */
size_t fcnod_aug_photo_scratch_bytes(int N, int src_h, int src_w);
int fcnod_aug_forward_photo_device(const float* data_in, const float* label, const fcnod_aug_draw* draws,
                                   const fcnod_photo_draw* photo, int N, int src_h, int src_w, int net_h, int net_w, int stride,
                                   int num_classes, float* top_data, float* fg, float* bbox, float* size, float* obj, float* cvg,
                                   int32_t* rects_out, void* scratch, void* photo_scratch, void* stream);
int fcnod_aug_forward_photo(const float* data_in, const float* label, const fcnod_aug_draw* draws, const fcnod_photo_draw* photo,
                            int N, int src_h, int src_w, int net_h, int net_w, int stride, int num_classes, float* top_data,
                            float* fg, float* bbox, float* size, float* obj, float* cvg, int32_t* rects_out);   /* HOST buffers */

/* ---- training head of models/train_val.prototxt (forward part) --------------------------
 * Slice / Eltwise-PROD mask and normalisation chain (:36-72, :2237-2266), L1Loss x loss_weight (:2267-2274) and
 * EuclideanLoss (:2275-2281) on the network's two outputs and the layer's five label grids, all NCHW float32:
 * coverage, cvg_label [N][C][gh][gw]; bboxes, bbox_label, size_block, obj_block, cvg_block [N][4C][gh][gw].
 * losses[0] = loss_bbox = sum|pred - target| / N (BEFORE its loss_weight), losses[1] = loss_coverage =
 * sum (coverage - cvg_label)^2 / (2N).  d_coverage / d_bboxes (may be NULL) receive the gradients of
 * loss_coverage + bbox_loss_weight * loss_bbox with respect to the two outputs (the first step of the backward pass).
 * scratch: fcnod_train_head_scratch_bytes() bytes of device memory.  Sums are float64 in a fixed order. */
size_t fcnod_train_head_scratch_bytes(void);
int fcnod_train_head_forward_device(const float* coverage, const float* bboxes, const float* cvg_label, const float* bbox_label,
                                    const float* size_block, const float* obj_block, const float* cvg_block, int N, int num_classes,
                                    int gh, int gw, float bbox_loss_weight, float* losses, float* d_coverage, float* d_bboxes,
                                    void* scratch, void* stream);
int fcnod_train_head_forward(const float* coverage, const float* bboxes, const float* cvg_label, const float* bbox_label,
                             const float* size_block, const float* obj_block, const float* cvg_block, int N, int num_classes, int gh,
                             int gw, float bbox_loss_weight, float* losses, float* d_coverage, float* d_bboxes);   /* HOST buffers */

/* ---- one training iteration of models/train_val.prototxt (BASELINE config 5) --------------
 * Replaces what `caffe train --solver=bounding_box/solver.prototxt` (train/train.sh:19) runs per iteration on one GPU:
 * Net::ForwardBackward over models/train_val.prototxt (body :27-35,76-2232 at TRAIN phase, Dropout :2164-2172 active,
 * head :36-72,2237-2281) and SGDSolver::ApplyUpdate with solver_type ADAM (train/bounding_box/solver.prototxt:3-15).
 * The data / label inputs are the six tops of the DataArgumentationLayer (fcnod_aug_forward_device): `data`
 * [N][3][H][W] float32 and the five label grids, NCHW float32, device pointers for the _device form.
 * prototxt: the train_val form (Data / Slice / Eltwise / loss layers are checked against the fused head and otherwise
 * skipped) or the deploy form with 1 / 4 head channels; caffemodel: the snapshot training starts from.
 * fp32 storage and arithmetic (the parity engine: 1e-4 against autograd over the same graph).  Parameters, gradients
 * and ADAM's moments are four flat float32 device buffers of one layout (fcnod_trainer_buffers, n_params floats each):
 * a data-parallel caller all-reduces `grads` (SUM) between forward_backward and apply_update and passes
 * grad_scale = 1 / world_size, which is what Caffe's NCCL solver does. */
typedef struct fcnod_trainer fcnod_trainer;
typedef struct fcnod_solver_params {
    float base_lr, gamma;            /* lr_policy "step": rate = base_lr * gamma^floor(iter / stepsize) */
    int32_t stepsize;
    float momentum, momentum2, delta, weight_decay;   /* ADAM beta1, beta2, epsilon; L2 decay */
    float bbox_loss_weight;          /* loss_weight of bbox_loss (train_val.prototxt:2273) */
    float dropout_ratio_override;    /* < 0: the prototxt's dropout_ratio; 0 switches dropout off */
    uint64_t seed;                   /* dropout masks are a function of (seed, iteration, element) */
    int32_t precision;               /* 0: float32 storage, SIMT convolutions (parity engine, 1e-4); 1: bf16 storage with fp32
                                        master weights, convolutions as tcgen05 GEMMs (forward, dgrad, wgrad), fp32 accumulation */
    int32_t reserved;
} fcnod_solver_params;
int fcnod_solver_defaults(fcnod_solver_params* p);   /* the values of train/bounding_box/solver.prototxt */
int fcnod_trainer_create(const char* prototxt_path, const char* caffemodel_path, int device, int batch, int height, int width,
                         const fcnod_solver_params* solver /* NULL = defaults */, fcnod_trainer** out);
int fcnod_trainer_destroy(fcnod_trainer* t);
int fcnod_trainer_info(const fcnod_trainer* t, int dims[7] /* N, C, H, W, num_classes, grid h, grid w */, size_t* n_params);
int fcnod_trainer_buffers(fcnod_trainer* t, float** params, float** grads, float** m, float** v);   /* device pointers */
int fcnod_trainer_forward_backward_device(fcnod_trainer* t, const float* data, const float* cvg_label, const float* bbox_label,
                                          const float* size_block, const float* obj_block, const float* cvg_block, void* stream);
int fcnod_trainer_apply_update(fcnod_trainer* t, float grad_scale, void* stream);
/* Data-parallel update over NVLink peer memory (one process per GPU on one node; replaces NCCL all-reduce + apply_update):
 * every rank exports its parameter and gradient buffers (_ipc_handles: two 64-byte CUDA IPC handles), the handles of all
 * ranks are exchanged by the host (trainer.DataParallelSolver does it over torch.distributed) and opened (_open_peers).
 * _apply_update_peer is then ONE kernel per rank: it sums its shard of every rank's gradient buffer through peer loads,
 * applies ApplyUpdate (scaled by 1 / world) to that shard -- ADAM moments exist only on the shard's owner -- and stores the
 * new parameters into every rank's parameter buffer through peer stores.  The caller must order it between two
 * stream-ordered barriers: after every rank's backward, and before any rank's next forward. */
int fcnod_trainer_ipc_handles(fcnod_trainer* t, void* params_handle /* 64 bytes */, void* grads_handle /* 64 bytes */);
int fcnod_trainer_open_peers(fcnod_trainer* t, int world, int rank, const void* params_handles /* world x 64 bytes */,
                             const void* grads_handles);
int fcnod_trainer_apply_update_peer(fcnod_trainer* t, void* stream);
int fcnod_trainer_losses(fcnod_trainer* t, float losses[2] /* loss_bbox (unweighted), loss_coverage */, void* stream);
int fcnod_trainer_iteration(const fcnod_trainer* t);
int fcnod_trainer_step(fcnod_trainer* t, const float* data, const float* cvg_label, const float* bbox_label, const float* size_block,
                       const float* obj_block, const float* cvg_block, float losses[2]);   /* HOST buffers, one GPU */
/* Introspection (snapshots, tests).  which: 0 value, 1 gradient, 2 ADAM m, 3 ADAM v; index: 0 weights in Caffe's
 * [Cout][Cin][k][k] order, 1 bias.  out == NULL: *count receives the element count. */
int fcnod_trainer_param_read(fcnod_trainer* t, const char* layer, int index, int which, float* out, size_t* count);
const char* fcnod_trainer_param_layer(const fcnod_trainer* t, int i);   /* i-th convolution's name, NULL past the end */
int fcnod_trainer_blob_read(fcnod_trainer* t, const char* blob, int which /* 0 data, 1 diff */, int dims[3], float* out, size_t count);
int fcnod_trainer_dropout_mask(fcnod_trainer* t, uint8_t* out /* [N][C][H][W] */, size_t count, int dims[3]);
/* Measurement aid (bench.py's roofline): time every GEMM launch of the following steps with CUDA events on the launching
 * stream; _gemm_times returns the sums of the last step: ms, algorithmic FLOP (2 M N K) and launch counts of the forward,
 * data-gradient and weight-gradient GEMMs (bf16 engine).  No counterpart in the reference. */
int fcnod_trainer_set_timing(fcnod_trainer* t, int on);
int fcnod_trainer_gemm_times(fcnod_trainer* t, float ms[3], double flop[3], int launches[3]);
/* The tcgen05 GEMM of the bf16 training engine on caller-provided DEVICE matrices (bf16, row-major, leading dimensions in
 * elements, rows 16-byte aligned).  No counterpart in the reference (it delegates to cuBLAS / cuDNN through Caffe).
 *   mode 0: out[M][N] (bf16, or float32 if out_f32) (+)= A[M][K] * B[N][K]^T (+ bias[N]) (ReLU)
 *   mode 1: out[n * ldo + m] (float32) += sum_k A[k][m] * B[k][n]   (reduction over the rows: the weight-gradient form) */
int fcnod_gemm_bf16_device(int mode, const void* A, int lda, const void* B, int ldb, int M, int N, int K, void* out, int ldo,
                           int out_f32, const float* bias, int relu, int accum, void* stream);
/* argmax of a Pooling layer's last forward: dh * k + dw inside the unclipped window, [N][C][Ho][Wo] uint8 */
int fcnod_trainer_pool_argmax(fcnod_trainer* t, const char* layer, uint8_t* out, size_t count, int dims[3]);

#ifdef __cplusplus
}
#endif
#endif /* FCNOD_B200_H */
