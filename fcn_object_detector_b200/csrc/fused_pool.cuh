// Fused / tiled bandwidth-bound kernels of the bf16 ("deploy") plan on NHWC activations.
// None of these exists in the reference, which runs one Caffe layer per prototxt entry
// (/root/reference/models/deploy.prototxt: Pooling :54-64, LRN :65-75, LRN :141-151,
// Pooling :152-162, and the nine `inception_*/pool` layers, e.g. :331-343):
//   pool_tile_kernel      max-pool KxK stride S, every thread produces a 2x2 block of output pixels for
//                         8 channels from one shared (S+K)x(S+K) window: 16 (S=1) / 25 (S=2) 128-bit loads
//                         per 4 outputs instead of 36, 32-bit index arithmetic.
//   pool_lrn_kernel       Pooling -> LRN in one pass (pool1/3x3_s2 + pool1/norm1): the pooled pixel is kept
//                         in shared memory (fp32) and normalised across channels from there.
//   lrn_pool_kernel       LRN -> Pooling in one pass (conv2/norm2 + pool2/3x3_s2): a CTA normalises the
//                         (2TH+1)x(2TW+1) input region of its TH x TW output tile into shared memory (bf16,
//                         the rounding the unfused plan applies when it stores the LRN blob) and pools from there.
// Caffe semantics as in elementwise.cuh: ceil-mode output size, windows clipped to the image,
// LRN ACROSS_CHANNELS x * (k + alpha/n * sum x^2)^-beta.
#pragma once
#include "common.cuh"
#include "elementwise.cuh"

namespace fcnod {

// (k + a*s)^-beta.  FAST: beta == 0.75 via rsqrt/sqrt (2 MUFU), else exp2/log2; !FAST: powf (fp32 parity mode).
template <bool FAST>
__device__ __forceinline__ float lrn_scale(float sc, float beta) {
    if (FAST) {
        if (beta == 0.75f) { float r = rsqrtf(sc); return r * sqrtf(r); }
        return __powf(sc, -beta);
    }
    return powf(sc, -beta);
}

// 8 channels of one pixel kept in the storage type, with an elementwise max.
template <typename T> struct Pack8;
template <> struct Pack8<bf16> {
    uint4 r;
    __device__ __forceinline__ void fill_neg_inf() { r.x = r.y = r.z = r.w = 0xff80ff80u; }
    __device__ __forceinline__ void load(const bf16* p) { r = *reinterpret_cast<const uint4*>(p); }
    __device__ __forceinline__ void store(bf16* p) const { *reinterpret_cast<uint4*>(p) = r; }
    __device__ __forceinline__ void max_with(const Pack8& o) {
        __nv_bfloat162* a = reinterpret_cast<__nv_bfloat162*>(&r);
        const __nv_bfloat162* b = reinterpret_cast<const __nv_bfloat162*>(&o.r);
#pragma unroll
        for (int i = 0; i < 4; ++i) a[i] = __hmax2(a[i], b[i]);
    }
    __device__ __forceinline__ void to_f32(float* v) const {
        const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&r);
#pragma unroll
        for (int i = 0; i < 4; ++i) { float2 f = __bfloat1622float2(h[i]); v[2 * i] = f.x; v[2 * i + 1] = f.y; }
    }
};
template <> struct Pack8<float> {
    float4 a, b;
    __device__ __forceinline__ void fill_neg_inf() { a = b = make_float4(-INFINITY, -INFINITY, -INFINITY, -INFINITY); }
    __device__ __forceinline__ void load(const float* p) {
        a = *reinterpret_cast<const float4*>(p); b = *reinterpret_cast<const float4*>(p + 4);
    }
    __device__ __forceinline__ void store(float* p) const {
        *reinterpret_cast<float4*>(p) = a; *reinterpret_cast<float4*>(p + 4) = b;
    }
    __device__ __forceinline__ void max_with(const Pack8& o) {
        a.x = fmaxf(a.x, o.a.x); a.y = fmaxf(a.y, o.a.y); a.z = fmaxf(a.z, o.a.z); a.w = fmaxf(a.w, o.a.w);
        b.x = fmaxf(b.x, o.b.x); b.y = fmaxf(b.y, o.b.y); b.z = fmaxf(b.z, o.b.z); b.w = fmaxf(b.w, o.b.w);
    }
    __device__ __forceinline__ void to_f32(float* v) const {
        v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
    }
};

// ---- max pool, 2x2 outputs per thread --------------------------------------------------------------------
template <typename T, int K, int S>
__global__ void __launch_bounds__(256) pool_tile_kernel(const T* __restrict__ in, T* __restrict__ out, int N, int C, int H,
                                                        int W, int ldi, int Ho, int Wo, int ldo, int p) {
    const int CG = C >> 3;
    const int Wb = (Wo + 1) >> 1, Hb = (Ho + 1) >> 1;
    const unsigned total = (unsigned)N * Hb * Wb * CG;
    constexpr int WIN = S + K;   // input rows / cols feeding a 2x2 output block
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        const int cg = i % CG;
        unsigned t = i / CG;
        const int wb = t % Wb; t /= Wb;
        const int hb = t % Hb;
        const int n = t / Hb;
        const int oh0 = hb * 2, ow0 = wb * 2;
        const int ih0 = oh0 * S - p, iw0 = ow0 * S - p;
        Pack8<T> m[2][2];
#pragma unroll
        for (int a = 0; a < 2; ++a)
#pragma unroll
            for (int b = 0; b < 2; ++b) m[a][b].fill_neg_inf();
        const T* base = in + (size_t)n * H * W * ldi + cg * 8;
#pragma unroll
        for (int r = 0; r < WIN; ++r) {
            const int ih = ih0 + r;
            if (ih < 0 || ih >= H) continue;
#pragma unroll
            for (int c = 0; c < WIN; ++c) {
                const int iw = iw0 + c;
                if (iw < 0 || iw >= W) continue;
                Pack8<T> x;
                x.load(base + ((size_t)ih * W + iw) * ldi);
#pragma unroll
                for (int a = 0; a < 2; ++a)
#pragma unroll
                    for (int b = 0; b < 2; ++b)
                        if (r >= a * S && r < a * S + K && c >= b * S && c < b * S + K) m[a][b].max_with(x);
            }
        }
#pragma unroll
        for (int a = 0; a < 2; ++a)
#pragma unroll
            for (int b = 0; b < 2; ++b) {
                const int oh = oh0 + a, ow = ow0 + b;
                if (oh < Ho && ow < Wo) m[a][b].store(out + ((size_t)(n * Ho + oh) * Wo + ow) * ldo + cg * 8);
            }
    }
}

// ---- Pooling -> LRN ---------------------------------------------------------------------------------------
// CTA = 256 threads = NPIX output pixels x CG 8-channel groups per pass (NPIX = 256 / CG).
// smem: float [NPIX][C + 16]: 8 zeros, C pooled values, 8 zeros (so the +-HALF window needs no bounds checks).
template <typename T, int K, int S, int HALF, bool FAST>
__global__ void __launch_bounds__(256) pool_lrn_kernel(const T* __restrict__ in, T* __restrict__ out, int N, int C, int H,
                                                       int W, int ldi, int Ho, int Wo, int ldo, int p,
                                                       float alpha_over_n, float beta, float kk) {
    extern __shared__ float pl_smem[];
    const int CG = C >> 3;
    const int NPIX = 256 / CG;
    const int rowf = C + 16;
    const int tp = threadIdx.x / CG, cg = threadIdx.x - tp * CG;
    const bool active = tp < NPIX;
    const unsigned npix_total = (unsigned)N * Ho * Wo;
    // zero the pads once
    for (int i = threadIdx.x; i < NPIX * 16; i += 256) {
        const int r = i >> 4, j = i & 15;
        pl_smem[r * rowf + (j < 8 ? j : C + j)] = 0.f;
    }
    for (unsigned pix0 = blockIdx.x * NPIX; pix0 < npix_total; pix0 += gridDim.x * NPIX) {
        const unsigned pix = pix0 + tp;
        const bool ok = active && pix < npix_total;
        if (ok) {
            const int ow = pix % Wo;
            unsigned t = pix / Wo;
            const int oh = t % Ho;
            const int n = t / Ho;
            const int h0 = max(oh * S - p, 0), h1 = min(oh * S - p + K, H);
            const int w0 = max(ow * S - p, 0), w1 = min(ow * S - p + K, W);
            Pack8<T> m;
            m.fill_neg_inf();
            const T* base = in + (size_t)n * H * W * ldi + cg * 8;
            for (int h = h0; h < h1; ++h)
                for (int w = w0; w < w1; ++w) {
                    Pack8<T> x;
                    x.load(base + ((size_t)h * W + w) * ldi);
                    m.max_with(x);
                }
            float v[8];
            m.to_f32(v);
            float* d = pl_smem + tp * rowf + 8 + cg * 8;
            *reinterpret_cast<float4*>(d) = make_float4(v[0], v[1], v[2], v[3]);
            *reinterpret_cast<float4*>(d + 4) = make_float4(v[4], v[5], v[6], v[7]);
        }
        __syncthreads();
        if (ok) {
            const float* s = pl_smem + tp * rowf + 8 + cg * 8;
            float x[8 + 2 * HALF];
#pragma unroll
            for (int j = 0; j < 8 + 2 * HALF; ++j) x[j] = s[j - HALF];
            Vec8<T> o;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                float acc = 0.f;
#pragma unroll
                for (int d = 0; d <= 2 * HALF; ++d) acc += x[j + d] * x[j + d];
                o.v[j] = x[j + HALF] * lrn_scale<FAST>(kk + alpha_over_n * acc, beta);
            }
            o.store(out + (size_t)pix * ldo + cg * 8);
        }
        __syncthreads();
    }
}

// ---- LRN -> Pooling -------------------------------------------------------------------------------------------
// One CTA per (image, TH x TW output tile).  smem: T [(S*(TH-1)+K) * (S*(TW-1)+K)][C].
template <typename T, int K, int S, int HALF, bool FAST, int TH, int TW>
__global__ void __launch_bounds__(256) lrn_pool_kernel(const T* __restrict__ in, T* __restrict__ out, int C, int H, int W,
                                                       int ldi, int Ho, int Wo, int ldo, int p, int tiles_x, int tiles_y,
                                                       float alpha_over_n, float beta, float kk) {
    extern __shared__ __align__(16) unsigned char lp_smem_raw[];
    T* tile = reinterpret_cast<T*>(lp_smem_raw);
    constexpr int RH = S * (TH - 1) + K, RW = S * (TW - 1) + K;
    const int CG = C >> 3;
    int b = blockIdx.x;
    const int tx = b % tiles_x; b /= tiles_x;
    const int ty = b % tiles_y;
    const int n = b / tiles_y;
    const int oh0 = ty * TH, ow0 = tx * TW;
    const int ih0 = oh0 * S - p, iw0 = ow0 * S - p;
    const T* img = in + (size_t)n * H * W * ldi;
    // stage 1: normalise the input region into shared memory (-inf outside the image: max ignores it)
    const int items = RH * RW * CG;
    for (int i = threadIdx.x; i < items; i += 256) {
        const int cg = i % CG;
        const int rp = i / CG;
        const int rr = rp / RW, rc = rp - rr * RW;
        const int ih = ih0 + rr, iw = iw0 + rc;
        Vec8<T> o;
        if (ih >= 0 && ih < H && iw >= 0 && iw < W) {
            const T* row = img + ((size_t)ih * W + iw) * ldi;
            float x[8 + 2 * HALF];
            Vec8<T> a;
            a.load(row + cg * 8);
#pragma unroll
            for (int j = 0; j < 8; ++j) x[HALF + j] = a.v[j];
#pragma unroll
            for (int d = 0; d < HALF; ++d) {
                const int cl = cg * 8 - HALF + d, ch = cg * 8 + 8 + d;
                x[d] = cl >= 0 ? to_f32<T>(row[cl]) : 0.f;
                x[8 + HALF + d] = ch < C ? to_f32<T>(row[ch]) : 0.f;
            }
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                float acc = 0.f;
#pragma unroll
                for (int d = 0; d <= 2 * HALF; ++d) acc += x[j + d] * x[j + d];
                o.v[j] = x[j + HALF] * lrn_scale<FAST>(kk + alpha_over_n * acc, beta);
            }
        } else {
#pragma unroll
            for (int j = 0; j < 8; ++j) o.v[j] = -INFINITY;
        }
        o.store(tile + (size_t)rp * C + cg * 8);
    }
    __syncthreads();
    // stage 2: pool from shared memory
    const int oitems = TH * TW * CG;
    for (int i = threadIdx.x; i < oitems; i += 256) {
        const int cg = i % CG;
        const int op = i / CG;
        const int a = op / TW, bcol = op - a * TW;
        const int oh = oh0 + a, ow = ow0 + bcol;
        if (oh >= Ho || ow >= Wo) continue;
        Pack8<T> m;
        m.fill_neg_inf();
#pragma unroll
        for (int r = 0; r < K; ++r)
#pragma unroll
            for (int c = 0; c < K; ++c) {
                Pack8<T> x;
                x.load(tile + (size_t)((a * S + r) * RW + bcol * S + c) * C + cg * 8);
                m.max_with(x);
            }
        m.store(out + ((size_t)(n * Ho + oh) * Wo + ow) * ldo + cg * 8);
    }
}

// ---- data (NCHW fp32) -> scale*x -> 2x2 space-to-depth NHWC bf16, W-padded -------------------------------------
// out[n][h2][PADL + w2][(dy*2+dx)*C + c] = bf16(in[n][c][2*h2+dy][2*w2+dx]);  channels >= 4*C and the pad
// columns stay zero (written once at allocation).  Feeds the first convolution (stride 2) as a stride-1
// tensor-core convolution (see net.cu: plan_conv_s2d).
__global__ void __launch_bounds__(256) input_s2d_kernel(const float* __restrict__ in, bf16* __restrict__ out, int N, int C,
                                                        int H, int W, int Wp, int padl, int ldc) {
    const int H2 = H >> 1, W2 = W >> 1;
    const unsigned total = (unsigned)N * H2 * W2;
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        const int w2 = i % W2;
        unsigned t = i / W2;
        const int h2 = t % H2;
        const int n = t / H2;
        bf16* o = out + ((size_t)(n * H2 + h2) * Wp + padl + w2) * ldc;
        for (int c = 0; c < C; ++c) {
            const float* src = in + ((size_t)(n * C + c) * H + 2 * h2) * W + 2 * w2;
            const float2 r0 = *reinterpret_cast<const float2*>(src);
            const float2 r1 = *reinterpret_cast<const float2*>(src + W);
            o[0 * C + c] = __float2bfloat16_rn(r0.x);
            o[1 * C + c] = __float2bfloat16_rn(r0.y);
            o[2 * C + c] = __float2bfloat16_rn(r1.x);
            o[3 * C + c] = __float2bfloat16_rn(r1.y);
        }
    }
}

}  // namespace fcnod
