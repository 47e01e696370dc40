// Bandwidth-bound layers of the network on NHWC activations (float or bf16):
// input transform (Power), max-pool, LRN, ReLU, Sigmoid, layout export.
// Caffe semantics restated from /root/reference/models/deploy.prototxt:
//   Power :8-16, Pooling :54-64 (ceil-mode output size, windows clipped to the
//   image), LRN :65-75 (ACROSS_CHANNELS, x * (k + alpha/n * sum x^2)^-beta).
// All kernels: one thread per (pixel, 8-channel vector) with 128-bit accesses
// when the channel count allows, grid-stride loops, grids sized in multiples
// of the SM count.
#pragma once
#include "common.cuh"

namespace fcnod {

template <typename T> struct Vec8;
template <> struct Vec8<float> {
    float v[8];
    __device__ __forceinline__ void load(const float* p) {
        float4 a = *reinterpret_cast<const float4*>(p), b = *reinterpret_cast<const float4*>(p + 4);
        v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
    }
    __device__ __forceinline__ void store(float* p) const {
        *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
        *reinterpret_cast<float4*>(p + 4) = make_float4(v[4], v[5], v[6], v[7]);
    }
};
template <> struct Vec8<bf16> {
    float v[8];
    __device__ __forceinline__ void load(const bf16* p) {
        uint4 r = *reinterpret_cast<const uint4*>(p);
        const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&r);
#pragma unroll
        for (int i = 0; i < 4; ++i) { float2 f = __bfloat1622float2(h[i]); v[2 * i] = f.x; v[2 * i + 1] = f.y; }
    }
    __device__ __forceinline__ void store(bf16* p) const {
        uint4 r;
        __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&r);
#pragma unroll
        for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
        *reinterpret_cast<uint4*>(p) = r;
    }
};

// ---- data (NCHW fp32) -> Power -> NHWC fp32 ----------------------------------------------
__global__ void input_transform_kernel(const float* __restrict__ in, float* __restrict__ out, int N, int C, int HW,
                                       float shift, float scale, float power) {
    long total = (long)N * HW;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        long n = i / HW, p = i - n * HW;
        for (int c = 0; c < C; ++c) {
            float x = in[(n * C + c) * HW + p];
            float y = shift + scale * x;
            if (power != 1.f) y = powf(y, power);
            out[i * C + c] = y;
        }
    }
}

// ---- max pool ------------------------------------------------------------------------------
template <typename T>
__global__ void pool_max_kernel(const T* __restrict__ in, T* __restrict__ out, int N, int C, int H, int W, int ldi,
                                int Ho, int Wo, int ldo, int k, int s, int p) {
    const int CG = C >> 3;
    long total = (long)N * Ho * Wo * CG;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        int cg = (int)(i % CG);
        long pix = i / CG;
        int ow = (int)(pix % Wo);
        long t = pix / Wo;
        int oh = (int)(t % Ho);
        int n = (int)(t / Ho);
        int h0 = max(oh * s - p, 0), h1 = min(oh * s - p + k, H);
        int w0 = max(ow * s - p, 0), w1 = min(ow * s - p + k, W);
        Vec8<T> m;
#pragma unroll
        for (int j = 0; j < 8; ++j) m.v[j] = -INFINITY;
        for (int h = h0; h < h1; ++h)
            for (int w = w0; w < w1; ++w) {
                Vec8<T> x;
                x.load(in + ((long)(n * H + h) * W + w) * ldi + cg * 8);
#pragma unroll
                for (int j = 0; j < 8; ++j) m.v[j] = fmaxf(m.v[j], x.v[j]);
            }
        m.store(out + pix * ldo + cg * 8);
    }
}

// ---- LRN across channels -------------------------------------------------------------------
template <typename T, int HALF>   // local_size = 2*HALF+1, HALF <= 8
__global__ void lrn_kernel(const T* __restrict__ in, T* __restrict__ out, long npix, int C, int ldi, int ldo,
                           float alpha_over_n, float beta, float kk) {
    const int CG = C >> 3;
    long total = npix * CG;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        int cg = (int)(i % CG);
        long pix = i / CG;
        const T* row = in + pix * ldi;
        float x[24];
        Vec8<T> a;
        if (cg > 0) { a.load(row + cg * 8 - 8);
#pragma unroll
            for (int j = 0; j < 8; ++j) x[j] = a.v[j]; }
        else {
#pragma unroll
            for (int j = 0; j < 8; ++j) x[j] = 0.f; }
        a.load(row + cg * 8);
#pragma unroll
        for (int j = 0; j < 8; ++j) x[8 + j] = a.v[j];
        if (cg + 1 < CG) { a.load(row + cg * 8 + 8);
#pragma unroll
            for (int j = 0; j < 8; ++j) x[16 + j] = a.v[j]; }
        else {
#pragma unroll
            for (int j = 0; j < 8; ++j) x[16 + j] = 0.f; }
        Vec8<T> o;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            float s = 0.f;
#pragma unroll
            for (int d = -HALF; d <= HALF; ++d) { float v = x[8 + j + d]; s += v * v; }
            float sc = kk + alpha_over_n * s;
            o.v[j] = x[8 + j] * powf(sc, -beta);
        }
        o.store(out + pix * ldo + cg * 8);
    }
}

// ---- elementwise (in place allowed) ---------------------------------------------------------
template <typename T, int OP>  // OP 0 = relu, 1 = sigmoid
__global__ void unary_kernel(const T* __restrict__ in, T* __restrict__ out, long npix, int C, int ldi, int ldo) {
    long total = npix * C;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        long pix = i / C;
        int c = (int)(i - pix * C);
        float x = to_f32<T>(in[pix * ldi + c]);
        float y = OP == 0 ? fmaxf(x, 0.f) : 1.f / (1.f + expf(-x));
        out[pix * ldo + c] = from_f32<T>(y);
    }
}

template <typename T>
__global__ void copy_channels_kernel(const T* __restrict__ in, T* __restrict__ out, long npix, int C, int ldi, int ldo) {
    long total = npix * C;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        long pix = i / C;
        int c = (int)(i - pix * C);
        out[pix * ldo + c] = in[pix * ldi + c];
    }
}

// ---- NHWC view -> dense NCHW float32 ----------------------------------------------------------
template <typename T>
__global__ void export_nchw_kernel(const T* __restrict__ in, float* __restrict__ out, int N, int C, int HW, int ldi) {
    long total = (long)N * C * HW;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        int p = (int)(i % HW);
        long t = i / HW;
        int c = (int)(t % C);
        long n = t / C;
        out[i] = to_f32<T>(in[(n * HW + p) * ldi + c]);
    }
}

// ---- position-sensitive 64-bit checksum of a blob view (test aid) ---------------------------------
// sum over elements of (bit pattern + 1) * (2 * index + 1) mod 2^64, index = pixel * C + channel.  Two tilings of one
// convolution accumulate every output in the same K order, so their blobs must agree bit for bit: the candidate sweep
// of tests/test_plan_gpu.py compares checksums instead of copying gigabytes to the host.
template <typename T> __device__ __forceinline__ unsigned long long elem_bits(T v);
template <> __device__ __forceinline__ unsigned long long elem_bits<float>(float v) { return __float_as_uint(v); }
template <> __device__ __forceinline__ unsigned long long elem_bits<bf16>(bf16 v) { return __bfloat16_as_ushort(v); }
template <typename T>
__global__ void checksum_kernel(const T* __restrict__ in, long npix, int C, int ld, unsigned long long* __restrict__ out) {
    const long total = npix * C;
    unsigned long long h = 0;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        const long pix = i / C;
        const int c = (int)(i - pix * C);
        h += (elem_bits<T>(in[pix * ld + c]) + 1ull) * (2ull * (unsigned long long)i + 1ull);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) h += __shfl_xor_sync(0xffffffffu, h, o);
    if ((threadIdx.x & 31) == 0) atomicAdd(out, h);
}

}  // namespace fcnod
