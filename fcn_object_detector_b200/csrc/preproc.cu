// Detector prelude on the device: BGR uint8 HxWx3 frame -> blob `data` (3 x net_h x net_w fp32).
//
// Restates /root/reference/scripts/fcn_object_detector.py:
//   demean_rgb_image   :407-413  float64; per-channel mean subtraction (BGR order), global min-max to [0,1]
//   cv.resize          :80       default INTER_LINEAR (half-pixel centres, edge clamp); identity at 448x448
//   transpose + write  :81-82    HWC -> CHW, cast to the blob's float32
// Pass 1: per-frame, per-channel uint8 min / max (integer atomics; min/max commute with the
//   monotone `- mean_c`, so the float64 global min / max are recovered exactly from them).
// Pass 2: normalise in float64 and resample; one thread per 4 output pixels, float4 stores.
#include "common.cuh"
#include "kernels.h"
#include "workspace.h"

namespace fcnod {
namespace {

__device__ __forceinline__ double mean64(int c) {
    return c == 0 ? 104.0069879317889 : (c == 1 ? 116.66876761696767 : 122.6789143406786);
}

__global__ void pre_init_kernel(int* mm, int B) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < B * 6) mm[i] = (i % 6) < 3 ? 255 : 0;     // [b][0..2] = min per channel, [b][3..5] = max
}

// Byte-wise SIMD (__vminu4 / __vmaxu4) over 48-byte chunks (16 pixels, three 128-bit loads).  A 12-byte period is three
// words whose bytes carry channels (0,1,2,0) (1,2,0,1) (2,0,1,2): one packed accumulator per word phase, unpacked
// at the end.  Frames that do not start on a 16-byte boundary take the scalar loop.
__global__ void __launch_bounds__(256) pre_minmax_kernel(const uint8_t* __restrict__ img, long npix, int* __restrict__ mm) {
    const int b = blockIdx.y;
    const uint8_t* p = img + (size_t)b * npix * 3;
    int mn[3] = {255, 255, 255}, mx[3] = {0, 0, 0};
    const long tid = blockIdx.x * (long)blockDim.x + threadIdx.x, nthr = (long)gridDim.x * blockDim.x;
    long done = 0;                                                     // pixels covered by the vector part
    if ((reinterpret_cast<uintptr_t>(p) & 15) == 0) {
        const long nchunk = npix / 16;
        const uint4* pv = reinterpret_cast<const uint4*>(p);
        unsigned lo[3] = {0xffffffffu, 0xffffffffu, 0xffffffffu}, hi[3] = {0u, 0u, 0u};
        for (long i = tid; i < nchunk; i += nthr) {
            const uint4 q0 = __ldg(pv + 3 * i), q1 = __ldg(pv + 3 * i + 1), q2 = __ldg(pv + 3 * i + 2);
            // word phases: q0 = 0 1 2 0, q1 = 1 2 0 1, q2 = 2 0 1 2
            lo[0] = __vminu4(lo[0], __vminu4(__vminu4(q0.x, q0.w), __vminu4(q1.z, q2.y)));
            lo[1] = __vminu4(lo[1], __vminu4(__vminu4(q0.y, q1.x), __vminu4(q1.w, q2.z)));
            lo[2] = __vminu4(lo[2], __vminu4(__vminu4(q0.z, q1.y), __vminu4(q2.x, q2.w)));
            hi[0] = __vmaxu4(hi[0], __vmaxu4(__vmaxu4(q0.x, q0.w), __vmaxu4(q1.z, q2.y)));
            hi[1] = __vmaxu4(hi[1], __vmaxu4(__vmaxu4(q0.y, q1.x), __vmaxu4(q1.w, q2.z)));
            hi[2] = __vmaxu4(hi[2], __vmaxu4(__vmaxu4(q0.z, q1.y), __vmaxu4(q2.x, q2.w)));
        }
#pragma unroll
        for (int ph = 0; ph < 3; ++ph)
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const int c = (ph + k) % 3;                            // byte k of a phase-ph word
                mn[c] = min(mn[c], (int)((lo[ph] >> (8 * k)) & 255u));
                mx[c] = max(mx[c], (int)((hi[ph] >> (8 * k)) & 255u));
            }
        done = nchunk * 16;
    }
    for (long i = done + tid; i < npix; i += nthr) {
#pragma unroll
        for (int c = 0; c < 3; ++c) { int v = p[i * 3 + c]; mn[c] = min(mn[c], v); mx[c] = max(mx[c], v); }
    }
#pragma unroll
    for (int c = 0; c < 3; ++c) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            mn[c] = min(mn[c], __shfl_xor_sync(0xffffffffu, mn[c], o));
            mx[c] = max(mx[c], __shfl_xor_sync(0xffffffffu, mx[c], o));
        }
    }
    if ((threadIdx.x & 31) == 0) {
#pragma unroll
        for (int c = 0; c < 3; ++c) { atomicMin(&mm[b * 6 + c], mn[c]); atomicMax(&mm[b * 6 + 3 + c], mx[c]); }
    }
}

// A frame holds at most 3 x 256 distinct normalised values ((p - mean_c) - min) / range: each CTA (one frame,
// a band of output rows) evaluates them once into shared memory in float64 -- the same expression, so the
// same bits, as evaluating it per tap -- and the resample is table reads + the two lerps.
template <int VEC>
__global__ void __launch_bounds__(256) pre_resample_kernel(const uint8_t* __restrict__ img, int src_h, int src_w,
                                                           int net_h, int net_w, const int* __restrict__ mm,
                                                           float* __restrict__ data) {
    __shared__ double lut[3][256];
    __shared__ float lutf[3][256];
    const int b = blockIdx.y;
    const bool same = (src_h == net_h && src_w == net_w);
    {
        const int* m = mm + b * 6;
        const double mn = fmin(fmin((double)m[0] - mean64(0), (double)m[1] - mean64(1)), (double)m[2] - mean64(2));
        const double mx = fmax(fmax((double)m[3] - mean64(0), (double)m[4] - mean64(1)), (double)m[5] - mean64(2));
        const double range = mx - mn;
        for (int i = threadIdx.x; i < 768; i += blockDim.x) {
            const int c = i >> 8, v = i & 255;
            const double q = (((double)v - mean64(c)) - mn) / range;
            lut[c][v] = q;
            lutf[c][v] = (float)q;
        }
    }
    __syncthreads();
    const int wv = net_w / VEC;
    const int rows_per = (net_h + gridDim.x - 1) / gridDim.x;
    const int row0 = blockIdx.x * rows_per, row1 = min(row0 + rows_per, net_h);
    const uint8_t* p = img + (size_t)b * src_h * src_w * 3;
    const double sy = (double)src_h / (double)net_h, sx = (double)src_w / (double)net_w;
    const int total = (row1 - row0) * wv;
    for (int t = threadIdx.x; t < total; t += blockDim.x) {
        const int xv = t % wv;
        const int oy = row0 + t / wv;
        float res[3][VEC];
        if (same) {
            const uint8_t* q = p + ((size_t)oy * src_w + xv * VEC) * 3;
            bool wide = false;
            if constexpr (VEC == 4) {
                if ((reinterpret_cast<uintptr_t>(q) & 3) == 0) {
                    const unsigned* qw = reinterpret_cast<const unsigned*>(q);
                    const unsigned w[3] = {__ldg(qw), __ldg(qw + 1), __ldg(qw + 2)};
#pragma unroll
                    for (int k = 0; k < 12; ++k) res[k % 3][k / 3] = lutf[k % 3][(w[k >> 2] >> (8 * (k & 3))) & 255u];
                    wide = true;
                }
            }
            if (!wide) {
#pragma unroll
                for (int v = 0; v < VEC; ++v)
#pragma unroll
                    for (int c = 0; c < 3; ++c) res[c][v] = lutf[c][q[v * 3 + c]];
            }
        } else {
            double fyd = ((double)oy + 0.5) * sy - 0.5;
            int y0 = (int)floor(fyd);
            double ay = fyd - (double)y0;
            if (y0 < 0) { y0 = 0; ay = 0.0; }
            if (y0 >= src_h - 1) { y0 = src_h - 1; ay = 0.0; }
            const int y1 = min(y0 + 1, src_h - 1);
            const uint8_t* r0 = p + (size_t)y0 * src_w * 3;
            const uint8_t* r1 = p + (size_t)y1 * src_w * 3;
#pragma unroll
            for (int v = 0; v < VEC; ++v) {
                const int ox = xv * VEC + v;
                double fxd = ((double)ox + 0.5) * sx - 0.5;
                int x0 = (int)floor(fxd);
                double ax = fxd - (double)x0;
                if (x0 < 0) { x0 = 0; ax = 0.0; }
                if (x0 >= src_w - 1) { x0 = src_w - 1; ax = 0.0; }
                const int x1 = min(x0 + 1, src_w - 1);
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    const double p00 = lut[c][r0[x0 * 3 + c]], p01 = lut[c][r0[x1 * 3 + c]];
                    const double p10 = lut[c][r1[x0 * 3 + c]], p11 = lut[c][r1[x1 * 3 + c]];
                    const double tp = p00 * (1.0 - ax) + p01 * ax, bt = p10 * (1.0 - ax) + p11 * ax;
                    res[c][v] = (float)(tp * (1.0 - ay) + bt * ay);
                }
            }
        }
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            float* o = data + (((size_t)b * 3 + c) * net_h + oy) * net_w + xv * VEC;
            if (VEC == 4) *reinterpret_cast<float4*>(o) = make_float4(res[c][0], res[c][1], res[c][2], res[c][3]);
            else
#pragma unroll
                for (int v = 0; v < VEC; ++v) o[v] = res[c][v];
        }
    }
}

// Prelude straight into the first convolution's operand: the same values pre_resample_kernel would write into the
// fp32 blob `data` (same table, same float64 lerps, same float rounding), rounded to bf16 and stored as the padded,
// pre-swizzled 2x2 space-to-depth image (fused_pool.cuh: input_s2d_kernel) -- the 154 MB fp32 `data` round trip of a
// 64-frame batch never happens on the detect path.  One thread per space-to-depth pixel (2 x 2 net pixels x 3 channels).
__global__ void __launch_bounds__(256) pre_s2d_kernel(const uint8_t* __restrict__ img, int src_h, int src_w, int net_h, int net_w,
                                                      const int* __restrict__ mm, bf16* __restrict__ out, int Wp, int Hp, int padl) {
    __shared__ double lut[3][256];
    __shared__ float lutf[3][256];
    const int b = blockIdx.y;
    const bool same = (src_h == net_h && src_w == net_w);
    {
        const int* m = mm + b * 6;
        const double mn = fmin(fmin((double)m[0] - mean64(0), (double)m[1] - mean64(1)), (double)m[2] - mean64(2));
        const double mx = fmax(fmax((double)m[3] - mean64(0), (double)m[4] - mean64(1)), (double)m[5] - mean64(2));
        const double range = mx - mn;
        for (int i = threadIdx.x; i < 768; i += blockDim.x) {
            const int c = i >> 8, v = i & 255;
            const double q = (((double)v - mean64(c)) - mn) / range;
            lut[c][v] = q;
            lutf[c][v] = (float)q;
        }
    }
    __syncthreads();
    const int H2 = net_h >> 1, W2 = net_w >> 1;
    const int rows_per = (H2 + gridDim.x - 1) / gridDim.x;
    const int row0 = blockIdx.x * rows_per, row1 = min(row0 + rows_per, H2);
    const uint8_t* p = img + (size_t)b * src_h * src_w * 3;
    const double sy = (double)src_h / (double)net_h, sx = (double)src_w / (double)net_w;
    const int total = (row1 - row0) * W2;
    for (int t = threadIdx.x; t < total; t += blockDim.x) {
        const int w2 = t % W2, h2 = row0 + t / W2;
        __align__(16) bf16 v[16];
#pragma unroll
        for (int j = 12; j < 16; ++j) v[j] = __float2bfloat16_rn(0.f);
#pragma unroll
        for (int dy = 0; dy < 2; ++dy) {
            const int oy = 2 * h2 + dy;
            if (same) {
                const uint8_t* q = p + ((size_t)oy * src_w + 2 * w2) * 3;       // two pixels = 6 bytes
#pragma unroll
                for (int dx = 0; dx < 2; ++dx)
#pragma unroll
                    for (int c = 0; c < 3; ++c) v[(dy * 2 + dx) * 3 + c] = __float2bfloat16_rn(lutf[c][q[dx * 3 + c]]);
            } else {
                double fyd = ((double)oy + 0.5) * sy - 0.5;
                int y0 = (int)floor(fyd);
                double ay = fyd - (double)y0;
                if (y0 < 0) { y0 = 0; ay = 0.0; }
                if (y0 >= src_h - 1) { y0 = src_h - 1; ay = 0.0; }
                const int y1 = min(y0 + 1, src_h - 1);
                const uint8_t* r0 = p + (size_t)y0 * src_w * 3;
                const uint8_t* r1 = p + (size_t)y1 * src_w * 3;
#pragma unroll
                for (int dx = 0; dx < 2; ++dx) {
                    const int ox = 2 * w2 + dx;
                    double fxd = ((double)ox + 0.5) * sx - 0.5;
                    int x0 = (int)floor(fxd);
                    double ax = fxd - (double)x0;
                    if (x0 < 0) { x0 = 0; ax = 0.0; }
                    if (x0 >= src_w - 1) { x0 = src_w - 1; ax = 0.0; }
                    const int x1 = min(x0 + 1, src_w - 1);
#pragma unroll
                    for (int c = 0; c < 3; ++c) {
                        const double p00 = lut[c][r0[x0 * 3 + c]], p01 = lut[c][r0[x1 * 3 + c]];
                        const double p10 = lut[c][r1[x0 * 3 + c]], p11 = lut[c][r1[x1 * 3 + c]];
                        const double tp = p00 * (1.0 - ax) + p01 * ax, bt = p10 * (1.0 - ax) + p11 * ax;
                        v[(dy * 2 + dx) * 3 + c] = __float2bfloat16_rn((float)(tp * (1.0 - ay) + bt * ay));
                    }
                }
            }
        }
        const int xp = padl + w2;
        const int swz = (xp >> 2) & 1;
        uint4* o = reinterpret_cast<uint4*>(out + ((size_t)(b * Hp + padl + h2) * Wp + xp) * 16);
        o[swz] = *reinterpret_cast<const uint4*>(v);
        o[swz ^ 1] = *reinterpret_cast<const uint4*>(v + 8);
    }
}

}  // namespace

int preprocess_num_launches() { return 3; }

void preprocess_s2d_launch(const uint8_t* images_dev, int B, int src_h, int src_w, int net_h, int net_w, void* s2d_dev, int Wp,
                           int Hp, int padl, void* scratch, cudaStream_t s) {
    FCNOD_CHECK(B > 0 && src_h > 0 && src_w > 0 && net_h > 0 && net_w > 0 && net_h % 2 == 0 && net_w % 2 == 0, "preprocess: bad dimensions");
    int* mm = static_cast<int*>(scratch);
    pre_init_kernel<<<ceil_div(B * 6, 128), 128, 0, s>>>(mm, B);
    FCNOD_CUDA(cudaGetLastError());
    pre_minmax_kernel<<<dim3(32, B), 256, 0, s>>>(images_dev, (long)src_h * src_w, mm);
    FCNOD_CUDA(cudaGetLastError());
    int bands = ceil_div(148 * 16, B);
    bands = max(1, min(bands, ceil_div(net_h / 2, 4)));
    pre_s2d_kernel<<<dim3(bands, B), 256, 0, s>>>(images_dev, src_h, src_w, net_h, net_w, mm, static_cast<bf16*>(s2d_dev), Wp, Hp, padl);
    FCNOD_CUDA(cudaGetLastError());
}


void preprocess_launch(const uint8_t* images_dev, int B, int src_h, int src_w, int net_h, int net_w, float* data_dev,
                       void* scratch, cudaStream_t s) {
    FCNOD_CHECK(B > 0 && src_h > 0 && src_w > 0 && net_h > 0 && net_w > 0, "preprocess: bad dimensions");
    int* mm = static_cast<int*>(scratch);
    pre_init_kernel<<<ceil_div(B * 6, 128), 128, 0, s>>>(mm, B);
    FCNOD_CUDA(cudaGetLastError());
    pre_minmax_kernel<<<dim3(32, B), 256, 0, s>>>(images_dev, (long)src_h * src_w, mm);
    FCNOD_CUDA(cudaGetLastError());
    // bands of output rows per frame: about 16 CTAs per SM in flight over the whole batch, at least 4 rows a band
    int bands = ceil_div(148 * 16, B);
    bands = max(1, min(bands, ceil_div(net_h, 4)));
    if (net_w % 4 == 0)
        pre_resample_kernel<4><<<dim3(bands, B), 256, 0, s>>>(images_dev, src_h, src_w, net_h, net_w, mm, data_dev);
    else
        pre_resample_kernel<1><<<dim3(bands, B), 256, 0, s>>>(images_dev, src_h, src_w, net_h, net_w, mm, data_dev);
    FCNOD_CUDA(cudaGetLastError());
}

}  // namespace fcnod

using namespace fcnod;

extern "C" {

int fcnod_preprocess_device(const uint8_t* images_dev, int B, int src_h, int src_w, int net_h, int net_w, float* data_dev,
                            void* scratch, void* stream) {
    FCNOD_API_BEGIN
    FCNOD_CHECK(images_dev && data_dev && scratch, "null argument");
    preprocess_launch(images_dev, B, src_h, src_w, net_h, net_w, data_dev, scratch, (cudaStream_t)stream);
    FCNOD_API_END
}

int fcnod_preprocess(const uint8_t* images, int B, int src_h, int src_w, int net_h, int net_w, float* data_out) {
    FCNOD_API_BEGIN
    FCNOD_CHECK(images && data_out, "null argument");
    FCNOD_CHECK(B > 0 && src_h > 0 && src_w > 0 && net_h > 0 && net_w > 0, "preprocess: bad dimensions");
    const size_t nin = (size_t)B * src_h * src_w * 3, nout = (size_t)B * 3 * net_h * net_w;
    Workspace& ws = workspace();
    cudaStream_t s = ws.stream();
    ws.reserve(Workspace::padded(nin) + Workspace::padded(nout * 4) + Workspace::padded((size_t)B * 6 * 4));
    uint8_t* d_img = ws.take<uint8_t>(nin);
    float* d_out = ws.take<float>(nout);
    int* d_scr = ws.take<int>((size_t)B * 6);
    FCNOD_CUDA(cudaMemcpyAsync(d_img, images, nin, cudaMemcpyHostToDevice, s));
    preprocess_launch(d_img, B, src_h, src_w, net_h, net_w, d_out, d_scr, s);
    FCNOD_CUDA(cudaMemcpyAsync(data_out, d_out, nout * 4, cudaMemcpyDeviceToHost, s));
    FCNOD_CUDA(cudaStreamSynchronize(s));
    FCNOD_API_END
}

}  // extern "C"
