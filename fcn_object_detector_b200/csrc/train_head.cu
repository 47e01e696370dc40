// Training-side head of /root/reference/models/train_val.prototxt on the device: everything between the network's two
// outputs and the two loss scalars, consuming the augmentation layer's label grids where they already are (HBM).
//
//   slice-label            Slice    :36-52      label = [coverage-label | bbox-label | size-block | obj-block | coverage-block]
//                                               (the layer's tops 1..5; here five pointers, no copy)
//   bb-label-norm          Eltwise PROD :53-62  bbox-label * size-block
//   bb-obj-norm            Eltwise PROD :63-72  (bbox-label * size-block) * obj-block              =: target
//   bbox_mask              Eltwise PROD :2237-2246  bboxes * coverage-block
//   bbox-norm              Eltwise PROD :2247-2256  (bboxes * coverage-block) * size-block
//   bbox-obj-norm          Eltwise PROD :2257-2266  ((bboxes * coverage-block) * size-block) * obj-block   =: pred
//   bbox_loss              L1Loss, loss_weight 2.0  :2267-2274   sum |pred - target| / N      [ext: NVCaffe l1_loss_layer]
//   coverage_loss          EuclideanLoss            :2275-2281   sum (coverage - coverage-label)^2 / (2 N)   [ext: Caffe]
// and the first step of the backward pass, the gradients that reach the two network outputs:
//   d loss_coverage / d coverage = (coverage - coverage-label) / N
//   d (w * loss_bbox) / d bboxes = (((w / N * sign(pred - target)) * obj-block) * size-block) * coverage-block
//                                  (Eltwise PROD backward, layer by layer in reverse order)
// Elementwise arithmetic is fp32 in the prototxt's order (no FMA contraction: this unit is built with -fmad=false); the
// two sums are accumulated in float64 with a fixed two-stage order, so the losses are deterministic run to run.
#include "common.cuh"
#include "kernels.h"
#include "workspace.h"

namespace fcnod {
namespace {

constexpr int kHeadBlocks = 148, kHeadThreads = 256;

__global__ void __launch_bounds__(kHeadThreads) train_head_kernel(const float* __restrict__ coverage, const float* __restrict__ bboxes,
                                                                  const float* __restrict__ cvg_label, const float* __restrict__ bbox_label,
                                                                  const float* __restrict__ size_block, const float* __restrict__ obj_block,
                                                                  const float* __restrict__ cvg_block, long n_cvg, long n_box, float inv_n,
                                                                  float gw_bbox, float* __restrict__ d_coverage,
                                                                  float* __restrict__ d_bboxes, double* __restrict__ partial) {
    double l1 = 0.0, l2 = 0.0;
    const long stride = (long)gridDim.x * blockDim.x, t0 = blockIdx.x * (long)blockDim.x + threadIdx.x;
    for (long i = t0; i < n_box; i += stride) {
        const float sz = size_block[i], ob = obj_block[i], cb = cvg_block[i];
        const float target = __fmul_rn(__fmul_rn(bbox_label[i], sz), ob);
        const float pred = __fmul_rn(__fmul_rn(__fmul_rn(bboxes[i], cb), sz), ob);
        const float diff = __fsub_rn(pred, target);
        l1 += (double)fabsf(diff);
        if (d_bboxes) {
            const float sgn = diff > 0.f ? 1.f : (diff < 0.f ? -1.f : 0.f);          // caffe_cpu_sign
            d_bboxes[i] = __fmul_rn(__fmul_rn(__fmul_rn(__fmul_rn(gw_bbox, sgn), ob), sz), cb);
        }
    }
    for (long i = t0; i < n_cvg; i += stride) {
        const float diff = __fsub_rn(coverage[i], cvg_label[i]);
        l2 += (double)__fmul_rn(diff, diff);
        if (d_coverage) d_coverage[i] = __fmul_rn(diff, inv_n);
    }
    // fixed-order block reduction: warp shuffle tree, then warp 0 over the 8 warp sums
    __shared__ double s1[kHeadThreads / 32], s2[kHeadThreads / 32];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { l1 += __shfl_down_sync(0xffffffffu, l1, o); l2 += __shfl_down_sync(0xffffffffu, l2, o); }
    if ((threadIdx.x & 31) == 0) { s1[threadIdx.x >> 5] = l1; s2[threadIdx.x >> 5] = l2; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double a = 0.0, b = 0.0;
        for (int w = 0; w < kHeadThreads / 32; ++w) { a += s1[w]; b += s2[w]; }
        partial[2 * blockIdx.x] = a; partial[2 * blockIdx.x + 1] = b;
    }
}

__global__ void train_head_finish_kernel(const double* __restrict__ partial, int nblocks, double inv_n, float* __restrict__ losses) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    double a = 0.0, b = 0.0;
    for (int i = 0; i < nblocks; ++i) { a += partial[2 * i]; b += partial[2 * i + 1]; }
    losses[0] = (float)(a * inv_n);                 // loss_bbox (before its loss_weight)
    losses[1] = (float)(b * inv_n * 0.5);           // loss_coverage
}

}  // namespace
}  // namespace fcnod

using namespace fcnod;

extern "C" {

size_t fcnod_train_head_scratch_bytes(void) { return (size_t)kHeadBlocks * 2 * sizeof(double); }

int fcnod_train_head_forward_device(const float* coverage, const float* bboxes, const float* cvg_label, const float* bbox_label,
                                    const float* size_block, const float* obj_block, const float* cvg_block, int N, int num_classes,
                                    int gh, int gw, float bbox_loss_weight, float* losses, float* d_coverage, float* d_bboxes,
                                    void* scratch, void* stream) {
    FCNOD_API_BEGIN
    FCNOD_CHECK(coverage && bboxes && cvg_label && bbox_label && size_block && obj_block && cvg_block && losses && scratch, "null argument");
    FCNOD_CHECK(N > 0 && num_classes > 0 && gh > 0 && gw > 0, "train head: bad dimensions");
    const long n_cvg = (long)N * num_classes * gh * gw, n_box = 4 * n_cvg;
    const float inv_n = 1.f / (float)N;      // host fp32 divide (IEEE)
    cudaStream_t s = (cudaStream_t)stream;
    double* partial = static_cast<double*>(scratch);
    train_head_kernel<<<kHeadBlocks, kHeadThreads, 0, s>>>(coverage, bboxes, cvg_label, bbox_label, size_block, obj_block, cvg_block, n_cvg,
                                                           n_box, inv_n, bbox_loss_weight * inv_n, d_coverage, d_bboxes, partial);
    FCNOD_CUDA(cudaGetLastError());
    train_head_finish_kernel<<<1, 32, 0, s>>>(partial, kHeadBlocks, 1.0 / (double)N, losses);
    FCNOD_CUDA(cudaGetLastError());
    FCNOD_API_END
}

int fcnod_train_head_forward(const float* coverage, const float* bboxes, const float* cvg_label, const float* bbox_label,
                             const float* size_block, const float* obj_block, const float* cvg_block, int N, int num_classes, int gh,
                             int gw, float bbox_loss_weight, float* losses, float* d_coverage, float* d_bboxes) {
    FCNOD_API_BEGIN
    FCNOD_CHECK(coverage && bboxes && cvg_label && bbox_label && size_block && obj_block && cvg_block && losses, "null argument");
    FCNOD_CHECK(N > 0 && num_classes > 0 && gh > 0 && gw > 0, "train head: bad dimensions");
    const size_t n1 = (size_t)N * num_classes * gh * gw, n4 = 4 * n1;
    Workspace& ws = workspace();
    cudaStream_t s = ws.stream();
    ws.reserve(4 * Workspace::padded(n1 * 4) + 8 * Workspace::padded(n4 * 4) + Workspace::padded(fcnod_train_head_scratch_bytes()) + 1024);
    float* d_cov = ws.take<float>(n1); float* d_cl = ws.take<float>(n1); float* d_dc = ws.take<float>(n1);
    float* d_bb = ws.take<float>(n4); float* d_bl = ws.take<float>(n4); float* d_sz = ws.take<float>(n4);
    float* d_ob = ws.take<float>(n4); float* d_cb = ws.take<float>(n4); float* d_db = ws.take<float>(n4);
    float* d_loss = ws.take<float>(2);
    void* d_scr = ws.take<char>(fcnod_train_head_scratch_bytes());
    const float* src1[2] = {coverage, cvg_label}; float* dst1[2] = {d_cov, d_cl};
    const float* src4[5] = {bboxes, bbox_label, size_block, obj_block, cvg_block}; float* dst4[5] = {d_bb, d_bl, d_sz, d_ob, d_cb};
    for (int i = 0; i < 2; ++i) FCNOD_CUDA(cudaMemcpyAsync(dst1[i], src1[i], n1 * 4, cudaMemcpyHostToDevice, s));
    for (int i = 0; i < 5; ++i) FCNOD_CUDA(cudaMemcpyAsync(dst4[i], src4[i], n4 * 4, cudaMemcpyHostToDevice, s));
    int rc = fcnod_train_head_forward_device(d_cov, d_bb, d_cl, d_bl, d_sz, d_ob, d_cb, N, num_classes, gh, gw, bbox_loss_weight, d_loss,
                                             d_coverage ? d_dc : 0, d_bboxes ? d_db : 0, d_scr, s);
    FCNOD_CHECK(rc == 0, fcnod_last_error());
    FCNOD_CUDA(cudaMemcpyAsync(losses, d_loss, 8, cudaMemcpyDeviceToHost, s));
    if (d_coverage) FCNOD_CUDA(cudaMemcpyAsync(d_coverage, d_dc, n1 * 4, cudaMemcpyDeviceToHost, s));
    if (d_bboxes) FCNOD_CUDA(cudaMemcpyAsync(d_bboxes, d_db, n4 * 4, cudaMemcpyDeviceToHost, s));
    FCNOD_CUDA(cudaStreamSynchronize(s));
    FCNOD_API_END
}

}  // extern "C"
