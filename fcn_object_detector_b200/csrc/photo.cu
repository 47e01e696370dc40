// Photometric augmentation of the training-side layer on the device (stage T4):
// `color_space_argumentation` of /root/reference/scripts/data_argumentation_layer/argumentation_engine.py:308-322, i.e. the
// imgaug chain (vendored upstream under scripts/data_argumentation_layer/imgaug/imgaug/augmenters.py)
//   OneOf( GaussianBlur(sigma in (0, 3))   :2779-2843  scipy.ndimage.gaussian_filter per channel, uint8 in / out
//        | AverageBlur(k in 2..7)          :2851-2946  cv2.blur(img, (k, k))
//        | MedianBlur(k in 3..7, odd)      :2954-3023  cv2.medianBlur(img, k) )
//   Sharpen(alpha, lightness)              :3135-3217  cv2.filter2D(channel, -1, 3x3 matrix), Convolve :3100-3126
//   Add(v, per_channel)                    :3490-3587  int32 add, clip to [0, 255]
//   Multiply(m, per_channel)               :3787-3876  float32 multiply, clip, truncate
//   Grayscale(alpha)                       :2737-2776, ChangeColorspace :2680-2726  cv2.cvtColor(RGB2GRAY), float64 blend, truncate
// on uint8 H x W x 3 images.  The random draws are made on the host (fcnod_photo_draw) and every stage restates the
// arithmetic of the third-party routine imgaug calls -- scipy's symmetric correlate1d (centre tap first, then pairs from
// the outside in, float64, truncation to uint8 after EACH axis, `reflect` borders), OpenCV's box filter for small uint8
// kernels (integer sums, fixed-point division by 2^23), its median (replicated borders), its float32 filter2D with
// round-half-even saturation (BORDER_REFLECT_101), and its 15-bit RGB2GRAY -- so the result is bit-identical to the
// scipy / cv2 present in this image (tests/test_photo_gpu.py); other library versions may differ by one grey level.
// Images of a batch may have different valid sizes (h, w) inside a common pitch: the layer augments CROPS.
#include "common.cuh"
#include "kernels.h"
#include "workspace.h"

namespace fcnod {
namespace {

__device__ __forceinline__ int reflect_idx(int i, int n) {          // scipy `reflect`: d c b a | a b c d | d c b a
    if (n == 1) return 0;
    const int p = 2 * n;
    i %= p; if (i < 0) i += p;
    return i >= n ? p - 1 - i : i;
}
__device__ __forceinline__ int reflect101_idx(int i, int n) {       // OpenCV BORDER_REFLECT_101: d c b | a b c d | c b a
    if (n == 1) return 0;
    const int p = 2 * n - 2;
    i %= p; if (i < 0) i += p;
    return i >= n ? p - i : i;
}
__device__ __forceinline__ int clampi(int i, int n) { return i < 0 ? 0 : (i >= n ? n - 1 : i); }

struct ImgRef { const uint8_t* in; uint8_t* out; int h, w; };
__device__ __forceinline__ ImgRef img_ref(const uint8_t* in, uint8_t* out, const int32_t* sizes, int n, int H, int W) {
    ImgRef r;
    r.in = in + (size_t)n * H * W * 3; r.out = out + (size_t)n * H * W * 3;
    r.h = sizes ? sizes[2 * n] : H; r.w = sizes ? sizes[2 * n + 1] : W;
    return r;
}

// Stage A: gaussian pass along axis 0 (rows) | box blur | median | copy.  Row pitch of image n is its own width w.
__global__ void __launch_bounds__(256) photo_stage_a_kernel(const uint8_t* __restrict__ in, uint8_t* __restrict__ out,
                                                            const int32_t* __restrict__ sizes, const fcnod_photo_draw* __restrict__ draws,
                                                            int H, int W) {
    const int n = blockIdx.y;
    const ImgRef im = img_ref(in, out, sizes, n, H, W);
    const fcnod_photo_draw d = draws[n];
    const int npx = im.h * im.w;
    for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < npx; t += gridDim.x * blockDim.x) {
        const int y = t / im.w, x = t - y * im.w;
        uint8_t o[3];
        if (d.blur_kind == 1) {
            const int r = d.gauss_radius;
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                double acc = (double)im.in[((size_t)y * im.w + x) * 3 + c] * d.gauss_w[0];
                for (int k = r; k >= 1; --k) {
                    const double a = (double)im.in[((size_t)reflect_idx(y - k, im.h) * im.w + x) * 3 + c];
                    const double b = (double)im.in[((size_t)reflect_idx(y + k, im.h) * im.w + x) * 3 + c];
                    acc = acc + (a + b) * d.gauss_w[k];
                }
                o[c] = (uint8_t)(int)acc;                       // C cast of a double in [0, 255]: truncation
            }
        } else if (d.blur_kind == 2) {
            const int k = d.blur_k, a = k / 2, area = k * k;
            // OpenCV ColumnSum<ushort, uchar>: D = (sum + divDelta) * divScale >> 23
            const double scalef = (double)(1 << 23) / (double)area;
            int div_scale = (int)floor(scalef), div_delta = area / 2;
            if (scalef - (double)div_scale < 0.5) ++div_delta; else ++div_scale;
            int s[3] = {0, 0, 0};
            for (int dy = 0; dy < k; ++dy) {
                const int yy = reflect101_idx(y - a + dy, im.h);
                for (int dx = 0; dx < k; ++dx) {
                    const uint8_t* q = im.in + ((size_t)yy * im.w + reflect101_idx(x - a + dx, im.w)) * 3;
                    s[0] += q[0]; s[1] += q[1]; s[2] += q[2];
                }
            }
#pragma unroll
            for (int c = 0; c < 3; ++c) o[c] = (uint8_t)(((long long)(s[c] + div_delta) * div_scale) >> 23);
        } else if (d.blur_kind == 3) {
            const int k = d.blur_k, a = k / 2, need = (k * k + 1) / 2;
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                int lo = 0, hi = 255;                           // smallest v with #(window <= v) >= need
                while (lo < hi) {
                    const int mid = (lo + hi) >> 1;
                    int cnt = 0;
                    for (int dy = 0; dy < k; ++dy) {
                        const int yy = clampi(y - a + dy, im.h);
                        for (int dx = 0; dx < k; ++dx) cnt += im.in[((size_t)yy * im.w + clampi(x - a + dx, im.w)) * 3 + c] <= mid ? 1 : 0;
                    }
                    if (cnt >= need) hi = mid; else lo = mid + 1;
                }
                o[c] = (uint8_t)lo;
            }
        } else {
#pragma unroll
            for (int c = 0; c < 3; ++c) o[c] = im.in[(size_t)t * 3 + c];
        }
#pragma unroll
        for (int c = 0; c < 3; ++c) im.out[(size_t)t * 3 + c] = o[c];
    }
}

// Stage B: gaussian pass along axis 1 (columns) for gaussian images, copy otherwise.
__global__ void __launch_bounds__(256) photo_stage_b_kernel(const uint8_t* __restrict__ in, uint8_t* __restrict__ out,
                                                            const int32_t* __restrict__ sizes, const fcnod_photo_draw* __restrict__ draws,
                                                            int H, int W) {
    const int n = blockIdx.y;
    const ImgRef im = img_ref(in, out, sizes, n, H, W);
    const fcnod_photo_draw d = draws[n];
    const int npx = im.h * im.w;
    for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < npx; t += gridDim.x * blockDim.x) {
        const int y = t / im.w, x = t - y * im.w;
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            uint8_t o = im.in[(size_t)t * 3 + c];
            if (d.blur_kind == 1) {
                double acc = (double)o * d.gauss_w[0];
                for (int k = d.gauss_radius; k >= 1; --k) {
                    const double a = (double)im.in[((size_t)y * im.w + reflect_idx(x - k, im.w)) * 3 + c];
                    const double b = (double)im.in[((size_t)y * im.w + reflect_idx(x + k, im.w)) * 3 + c];
                    acc = acc + (a + b) * d.gauss_w[k];
                }
                o = (uint8_t)(int)acc;
            }
            im.out[(size_t)t * 3 + c] = o;
        }
    }
}

// Stage C: Sharpen (3x3 filter2D, float32, coefficients in row-major order, round-half-even, saturate) -> Add -> Multiply
// -> Grayscale blend.  Only the sharpening reads neighbours; the rest is per pixel.
__global__ void __launch_bounds__(256) photo_stage_c_kernel(const uint8_t* __restrict__ in, uint8_t* __restrict__ out,
                                                            const int32_t* __restrict__ sizes, const fcnod_photo_draw* __restrict__ draws,
                                                            int H, int W) {
    const int n = blockIdx.y;
    const ImgRef im = img_ref(in, out, sizes, n, H, W);
    const fcnod_photo_draw d = draws[n];
    const int npx = im.h * im.w;
    for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < npx; t += gridDim.x * blockDim.x) {
        const int y = t / im.w, x = t - y * im.w;
        int v[3];
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            float s = 0.f;
#pragma unroll
            for (int r = 0; r < 3; ++r) {
                const int yy = reflect101_idx(y - 1 + r, im.h);
#pragma unroll
                for (int q = 0; q < 3; ++q) {
                    const float kf = d.sharpen[r * 3 + q];
                    if (kf != 0.f)
                        s = __fadd_rn(s, __fmul_rn(kf, (float)im.in[((size_t)yy * im.w + reflect101_idx(x - 1 + q, im.w)) * 3 + c]));
                }
            }
            int u = __float2int_rn(s);                           // saturate_cast<uchar>(float): cvRound + clamp
            u = u < 0 ? 0 : (u > 255 ? 255 : u);
            u += d.add[c];                                       // Add: int32, clip
            u = u < 0 ? 0 : (u > 255 ? 255 : u);
            float m = __fmul_rn((float)u, d.mul[c]);             // Multiply: float32, clip, astype(uint8) = truncation
            m = m < 0.f ? 0.f : (m > 255.f ? 255.f : m);
            v[c] = (int)m;
        }
        if (d.gray_alpha > 0.001) {                              // ChangeColorspace.eps
            const int g = (v[0] * 9798 + v[1] * 19235 + v[2] * 3735 + (1 << 14)) >> 15;      // cv2 RGB2GRAY, 8-bit
#pragma unroll
            for (int c = 0; c < 3; ++c) v[c] = (int)(d.gray_alpha * (double)g + (1.0 - d.gray_alpha) * (double)v[c]);
        }
#pragma unroll
        for (int c = 0; c < 3; ++c) im.out[(size_t)t * 3 + c] = (uint8_t)v[c];
    }
}

}  // namespace

void photo_launch(const uint8_t* in, uint8_t* out, uint8_t* tmp, const int32_t* sizes, const fcnod_photo_draw* draws, int N, int H,
                  int W, cudaStream_t s) {
    FCNOD_CHECK(N > 0 && H > 0 && W > 0, "photo: bad dimensions");
    const dim3 grid((unsigned)std::min(148 * 8 / std::max(N, 1) + 1, ceil_div(H * W, 256)), (unsigned)N);
    photo_stage_a_kernel<<<grid, 256, 0, s>>>(in, out, sizes, draws, H, W);          // in  -> out
    FCNOD_CUDA(cudaGetLastError());
    photo_stage_b_kernel<<<grid, 256, 0, s>>>(out, tmp, sizes, draws, H, W);         // out -> tmp
    FCNOD_CUDA(cudaGetLastError());
    photo_stage_c_kernel<<<grid, 256, 0, s>>>(tmp, out, sizes, draws, H, W);         // tmp -> out
    FCNOD_CUDA(cudaGetLastError());
}

}  // namespace fcnod

using namespace fcnod;

extern "C" {

int fcnod_photo_augment_device(const uint8_t* images, const int32_t* sizes, const fcnod_photo_draw* draws, int N, int H, int W,
                               uint8_t* out, uint8_t* tmp, void* stream) {
    FCNOD_API_BEGIN
    FCNOD_CHECK(images && draws && out && tmp, "null argument");
    FCNOD_CHECK(images != out && images != tmp && out != tmp, "photo: in / out / tmp must be distinct buffers");
    photo_launch(images, out, tmp, sizes, draws, N, H, W, (cudaStream_t)stream);
    FCNOD_API_END
}

int fcnod_photo_augment(const uint8_t* images, const int32_t* sizes, const fcnod_photo_draw* draws, int N, int H, int W, uint8_t* out) {
    FCNOD_API_BEGIN
    FCNOD_CHECK(images && draws && out, "null argument");
    FCNOD_CHECK(N > 0 && H > 0 && W > 0, "photo: bad dimensions");
    const size_t nb = (size_t)N * H * W * 3;
    Workspace& ws = workspace();
    cudaStream_t s = ws.stream();
    ws.reserve(3 * Workspace::padded(nb) + Workspace::padded((size_t)N * 8) + Workspace::padded((size_t)N * sizeof(fcnod_photo_draw)));
    uint8_t* d_in = ws.take<uint8_t>(nb); uint8_t* d_out = ws.take<uint8_t>(nb); uint8_t* d_tmp = ws.take<uint8_t>(nb);
    int32_t* d_sz = sizes ? ws.take<int32_t>((size_t)N * 2) : 0;
    fcnod_photo_draw* d_dr = ws.take<fcnod_photo_draw>((size_t)N);
    FCNOD_CUDA(cudaMemcpyAsync(d_in, images, nb, cudaMemcpyHostToDevice, s));
    if (sizes) FCNOD_CUDA(cudaMemcpyAsync(d_sz, sizes, (size_t)N * 8, cudaMemcpyHostToDevice, s));
    FCNOD_CUDA(cudaMemcpyAsync(d_dr, draws, (size_t)N * sizeof(fcnod_photo_draw), cudaMemcpyHostToDevice, s));
    FCNOD_CUDA(cudaMemsetAsync(d_out, 0, nb, s));
    photo_launch(d_in, d_out, d_tmp, d_sz, d_dr, N, H, W, s);
    FCNOD_CUDA(cudaMemcpyAsync(out, d_out, nb, cudaMemcpyDeviceToHost, s));
    FCNOD_CUDA(cudaStreamSynchronize(s));
    FCNOD_API_END
}

}  // extern "C"

// ---- copy-paste synthesis: the compositing loop of ArgumentationEngineMapping.argument ------------------------------
// /root/reference/scripts/data_argumentation_layer/argumentation_engine.py:723-728: for every placed proposal, in order,
//     for j in range(h): for i in range(w): nx, ny = i + cx, j + cy
//         if im_msk[j, i, 0] > 0 and nx < im_x and ny < im_y: img_output[ny, nx] = im_roi[j, i]; mask_output[ny, nx] = label + 1
// (the per-pixel Python loop the HEAD layer spends its time in).  Here one thread owns one output pixel and walks the
// image's proposals in order, keeping the LAST one that covers it with a set mask bit -- the same result as the sequential
// overwrites.  Negative cx / cy (a proposal wider than the background) index from the far side, as numpy does.
namespace fcnod {
namespace {

__global__ void __launch_bounds__(256) composite_kernel(uint8_t* __restrict__ images, uint8_t* __restrict__ masks, int H, int W,
                                                        const uint8_t* __restrict__ rois, const uint8_t* __restrict__ roi_masks,
                                                        const fcnod_paste* __restrict__ pastes, const int32_t* __restrict__ begin) {
    const int n = blockIdx.y;
    const int p0 = begin[n], p1 = begin[n + 1];
    if (p0 == p1) return;
    for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < H * W; t += gridDim.x * blockDim.x) {
        const int y = t / W, x = t - y * W;
        int hit = -1, hi = 0, hj = 0;
        for (int q = p0; q < p1; ++q) {
            const fcnod_paste P = pastes[q];
            // Which (j, i) of the patch land on (y, x)?  ny = j + cy equals y directly, or -- numpy's negative indexing --
            // ny = y - H < 0 wraps to y.  Up to two candidates per axis; the reference's loops run j, then i, upwards and the
            // last write wins, so candidates are visited in that order.
            const int jc[2] = {y - H - P.cy, y - P.cy}, ic[2] = {x - W - P.cx, x - P.cx};
#pragma unroll
            for (int a = 0; a < 2; ++a) {
                const int j = jc[a];
                if (j < 0 || j >= P.h || (a == 0 ? j + P.cy >= 0 : j + P.cy < 0)) continue;
#pragma unroll
                for (int b = 0; b < 2; ++b) {
                    const int i = ic[b];
                    if (i < 0 || i >= P.w || (b == 0 ? i + P.cx >= 0 : i + P.cx < 0)) continue;
                    if (roi_masks[(size_t)P.roi_offset + (size_t)j * P.w + i] > 0) { hit = q; hi = i; hj = j; }
                }
            }
        }
        if (hit >= 0) {
            const fcnod_paste P = pastes[hit];
            const uint8_t* s = rois + ((size_t)P.roi_offset + (size_t)hj * P.w + hi) * 3;
            uint8_t* o = images + ((size_t)n * H * W + t) * 3;
            o[0] = s[0]; o[1] = s[1]; o[2] = s[2];
            masks[(size_t)n * H * W + t] = (uint8_t)(P.label + 1);
        }
    }
}

}  // namespace
}  // namespace fcnod

extern "C" {

int fcnod_composite_device(uint8_t* images, uint8_t* masks, int N, int H, int W, const uint8_t* rois, const uint8_t* roi_masks,
                           const fcnod_paste* pastes, const int32_t* paste_begin, void* stream) {
    FCNOD_API_BEGIN
    FCNOD_CHECK(images && masks && rois && roi_masks && pastes && paste_begin, "null argument");
    FCNOD_CHECK(N > 0 && H > 0 && W > 0, "composite: bad dimensions");
    const dim3 grid((unsigned)std::min(148 * 8 / N + 1, ceil_div(H * W, 256)), (unsigned)N);
    composite_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(images, masks, H, W, rois, roi_masks, pastes, paste_begin);
    FCNOD_CUDA(cudaGetLastError());
    FCNOD_API_END
}

int fcnod_composite(uint8_t* images, uint8_t* masks, int N, int H, int W, const uint8_t* rois, const uint8_t* roi_masks,
                    size_t roi_pixels, const fcnod_paste* pastes, const int32_t* paste_begin) {
    FCNOD_API_BEGIN
    FCNOD_CHECK(images && masks && rois && roi_masks && pastes && paste_begin, "null argument");
    FCNOD_CHECK(N > 0 && H > 0 && W > 0, "composite: bad dimensions");
    const int np = paste_begin[N];
    FCNOD_CHECK(np >= 0, "composite: bad paste table");
    for (int q = 0; q < np; ++q)
        FCNOD_CHECK(pastes[q].w > 0 && pastes[q].h > 0 && pastes[q].roi_offset >= 0 &&
                    (size_t)pastes[q].roi_offset + (size_t)pastes[q].w * pastes[q].h <= roi_pixels, "composite: paste outside the ROI buffer");
    const size_t ni = (size_t)N * H * W;
    Workspace& ws = workspace();
    cudaStream_t s = ws.stream();
    ws.reserve(Workspace::padded(ni * 3) + Workspace::padded(ni) + Workspace::padded(roi_pixels * 3) + Workspace::padded(roi_pixels) +
               Workspace::padded((size_t)(np + 1) * sizeof(fcnod_paste)) + Workspace::padded((size_t)(N + 1) * 4));
    uint8_t* d_img = ws.take<uint8_t>(ni * 3); uint8_t* d_msk = ws.take<uint8_t>(ni);
    uint8_t* d_roi = ws.take<uint8_t>(roi_pixels * 3 + 1); uint8_t* d_rm = ws.take<uint8_t>(roi_pixels + 1);
    fcnod_paste* d_p = ws.take<fcnod_paste>((size_t)np + 1);
    int32_t* d_b = ws.take<int32_t>((size_t)N + 1);
    FCNOD_CUDA(cudaMemcpyAsync(d_img, images, ni * 3, cudaMemcpyHostToDevice, s));
    FCNOD_CUDA(cudaMemcpyAsync(d_msk, masks, ni, cudaMemcpyHostToDevice, s));
    if (roi_pixels) {
        FCNOD_CUDA(cudaMemcpyAsync(d_roi, rois, roi_pixels * 3, cudaMemcpyHostToDevice, s));
        FCNOD_CUDA(cudaMemcpyAsync(d_rm, roi_masks, roi_pixels, cudaMemcpyHostToDevice, s));
    }
    if (np) FCNOD_CUDA(cudaMemcpyAsync(d_p, pastes, (size_t)np * sizeof(fcnod_paste), cudaMemcpyHostToDevice, s));
    FCNOD_CUDA(cudaMemcpyAsync(d_b, paste_begin, (size_t)(N + 1) * 4, cudaMemcpyHostToDevice, s));
    int rc = fcnod_composite_device(d_img, d_msk, N, H, W, d_roi, d_rm, d_p, d_b, s);
    FCNOD_CHECK(rc == 0, fcnod_last_error());
    FCNOD_CUDA(cudaMemcpyAsync(images, d_img, ni * 3, cudaMemcpyDeviceToHost, s));
    FCNOD_CUDA(cudaMemcpyAsync(masks, d_msk, ni, cudaMemcpyDeviceToHost, s));
    FCNOD_CUDA(cudaStreamSynchronize(s));
    FCNOD_API_END
}

}  // extern "C"
