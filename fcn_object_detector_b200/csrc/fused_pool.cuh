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
__device__ __forceinline__ float rsqrt_ftz(float x) {
    float r;
    asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
template <bool FAST>
__device__ __forceinline__ float lrn_scale(float sc, float beta) {
    if (FAST) {
        // sc^-0.5 * (sc^-0.5)^0.5: 2 MUFU.  sc >= k = 1 and its rsqrt are never denormal, so the bare MUFU.RSQ
        // (rsqrt.approx.ftz) returns what rsqrtf() does without rsqrtf's denormal pre/post-scaling instructions.
        if (beta == 0.75f) { const float r = rsqrt_ftz(sc); return r * (r * rsqrt_ftz(r)); }
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

// ---- max pool 3x3 (stride 1 or 2), register marching ------------------------------------------------------------
// CTA = 8 channel groups (8 x 16 bytes = one 128-byte line per pixel) x 32 output columns; every thread marches down a
// strip of RS output rows carrying the horizontal 3-max of the input rows it shares with the next output row.  The three
// columns a thread reads are also read by its neighbours in the same CTA, so a pixel leaves L2 once and is re-used from
// L1 (the 2x2-block kernel above fetched every pixel from L2 four times).
template <typename T, int S>
__global__ void __launch_bounds__(256) pool3_march_kernel(const T* __restrict__ in, T* __restrict__ out, int C, int H, int W,
                                                          int ldi, int Ho, int Wo, int ldo, int p, int cg_chunks, int col_chunks,
                                                          int strips, int RS) {
    const int CG = C >> 3;
    int b = blockIdx.x;
    const int cgc = b % cg_chunks; b /= cg_chunks;
    const int cc = b % col_chunks; b /= col_chunks;
    const int strip = b % strips;
    const int n = b / strips;
    const int cg = cgc * 8 + (threadIdx.x & 7);
    const int ow = cc * 32 + (threadIdx.x >> 3);
    if (cg >= CG || ow >= Wo) return;
    const int oh_begin = strip * RS, oh_end = min(Ho, oh_begin + RS);
    const int iw0 = ow * S - p;
    const T* img = in + (size_t)n * H * W * ldi + cg * 8;
    // Out-of-image rows / columns are CLAMPED instead of skipped: the clamped pixel lies inside the same window, so the
    // maximum is unchanged, and the loads are unconditional (the compiler keeps several rows in flight).
    const size_t o0 = (size_t)max(iw0, 0) * ldi, o1 = (size_t)min(iw0 + 1, W - 1) * ldi, o2 = (size_t)min(iw0 + 2, W - 1) * ldi;
    auto hmax_row = [&](int ih, Pack8<T>& m) {      // max over the 3 columns of input row ih
        const T* row = img + (size_t)min(max(ih, 0), H - 1) * W * ldi;
        Pack8<T> x, y;
        m.load(row + o0); x.load(row + o1); y.load(row + o2);
        m.max_with(x); m.max_with(y);
    };
    T* o = out + ((size_t)(n * Ho + oh_begin) * Wo + ow) * ldo + cg * 8;
    if (S == 1) {
        Pack8<T> r0, r1;
        hmax_row(oh_begin - p, r0);
        hmax_row(oh_begin - p + 1, r1);
#pragma unroll 4
        for (int oh = oh_begin; oh < oh_end; ++oh, o += (size_t)Wo * ldo) {
            Pack8<T> r2;
            hmax_row(oh - p + 2, r2);
            Pack8<T> m = r0;
            m.max_with(r1); m.max_with(r2);
            m.store(o);
            r0 = r1; r1 = r2;
        }
    } else {
        Pack8<T> carry;
        hmax_row(oh_begin * 2 - p, carry);
#pragma unroll 2
        for (int oh = oh_begin; oh < oh_end; ++oh, o += (size_t)Wo * ldo) {
            Pack8<T> a, c;
            hmax_row(oh * 2 - p + 1, a);
            hmax_row(oh * 2 - p + 2, c);
            Pack8<T> m = carry;
            m.max_with(a); m.max_with(c);
            m.store(o);
            carry = c;
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

// ---- Pooling (3x3 stride 2) -> LRN(5), register marching ---------------------------------------------------------
// One thread = 8 channels of one output column, marching down RS output rows: the horizontal 3-max of the input row it
// shares with the next output row is carried in registers, so an output costs 6 new 128-bit loads instead of 9.  The CG
// (= C / 8, a power of two <= 32) lanes that hold one pixel exchange the two edge channels with their neighbours by
// shuffle for the LRN window.  No shared memory, no block barriers.
template <int CG, int RS, bool FAST>
__global__ void __launch_bounds__(256) pool3s2_lrn5_march_kernel(const bf16* __restrict__ in, bf16* __restrict__ out, int N, int H,
                                                                 int W, int ldi, int Ho, int Wo, int ldo, int p,
                                                                 float alpha_over_n, float beta, float kk) {
    constexpr int C = CG * 8;
    const int strips = (Ho + RS - 1) / RS;
    const unsigned total = (unsigned)N * strips * Wo * CG;           // multiple of CG: lanes of a pixel stay together
    const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
    const bool live = i < total;
    const unsigned ii = live ? i : 0;
    const int cg = ii % CG;
    unsigned t = ii / CG;
    const int ow = t % Wo; t /= Wo;
    const int strip = t % strips;
    const int n = t / strips;
    const int oh_begin = strip * RS, oh_end = min(Ho, oh_begin + RS);
    const int iw0 = ow * 2 - p;
    const bf16* img = in + (size_t)n * H * W * ldi + cg * 8;
    auto hmax_row = [&](int ih, Pack8<bf16>& m) {      // max over the 3 columns of input row ih (-inf outside the image)
        m.fill_neg_inf();
        if (ih < 0 || ih >= H) return;
        const bf16* row = img + (size_t)ih * W * ldi;
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            const int iw = iw0 + c;
            if (iw >= 0 && iw < W) { Pack8<bf16> x; x.load(row + (size_t)iw * ldi); m.max_with(x); }
        }
    };
    Pack8<bf16> carry;
    hmax_row(oh_begin * 2 - p, carry);
    for (int oh = oh_begin; oh < oh_end; ++oh) {
        Pack8<bf16> m = carry, a, b;
        hmax_row(oh * 2 - p + 1, a);
        hmax_row(oh * 2 - p + 2, b);
        m.max_with(a); m.max_with(b);
        carry = b;
        float x[12];                                       // channels cg*8-2 .. cg*8+9
        m.to_f32(x + 2);
        // edge channels from the neighbouring lanes of the same pixel (zero beyond the first / last channel)
        const float l0 = __shfl_up_sync(0xffffffffu, x[8], 1), l1 = __shfl_up_sync(0xffffffffu, x[9], 1);
        const float r0 = __shfl_down_sync(0xffffffffu, x[2], 1), r1 = __shfl_down_sync(0xffffffffu, x[3], 1);
        x[0] = cg > 0 ? l0 : 0.f; x[1] = cg > 0 ? l1 : 0.f;
        x[10] = cg < CG - 1 ? r0 : 0.f; x[11] = cg < CG - 1 ? r1 : 0.f;
        Vec8<bf16> o;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const float ss = x[j] * x[j] + x[j + 1] * x[j + 1] + x[j + 2] * x[j + 2] + x[j + 3] * x[j + 3] + x[j + 4] * x[j + 4];
            o.v[j] = x[j + 2] * lrn_scale<FAST>(kk + alpha_over_n * ss, beta);
        }
        if (live) o.store(out + ((size_t)(n * Ho + oh) * Wo + ow) * ldo + cg * 8);
    }
    (void)C;
}

// ---- Pooling -> LRN, shared-memory tiled ----------------------------------------------------------------------
// One CTA per (image, TH x TW output tile): the (S*(TH-1)+K) x (S*(TW-1)+K) input region is read ONCE with coalesced
// 128-bit loads into shared memory (bf16), pooled from there into an fp32 tile [TH*TW][C + 16] (8 zeros either side of
// the channels), then normalised across channels.  Input bytes leave L2 once instead of ~2.25 times.
template <typename T, int K, int S, int HALF, bool FAST, int TH, int TW>
__global__ void __launch_bounds__(256) pool_lrn_tile_kernel(const T* __restrict__ in, T* __restrict__ out, int C, int H, int W,
                                                            int ldi, int Ho, int Wo, int ldo, int p, int tiles_x, int tiles_y,
                                                            float alpha_over_n, float beta, float kk) {
    extern __shared__ __align__(16) unsigned char plt_smem[];
    constexpr int RH = S * (TH - 1) + K, RW = S * (TW - 1) + K;
    T* region = reinterpret_cast<T*>(plt_smem);
    const int CG = C >> 3;
    const int rowf = C + 16;
    float* pooled = reinterpret_cast<float*>(plt_smem + (((size_t)RH * RW * C * sizeof(T) + 15) & ~(size_t)15));
    int b = blockIdx.x;
    const int tx = b % tiles_x; b /= tiles_x;
    const int ty = b % tiles_y;
    const int n = b / tiles_y;
    const int oh0 = ty * TH, ow0 = tx * TW;
    const int ih0 = oh0 * S - p, iw0 = ow0 * S - p;
    const T* img = in + (size_t)n * H * W * ldi;
    // stage 1: region -> smem (-inf outside the image); zero the pads of the pooled tile
    for (int i = threadIdx.x; i < RH * RW * CG; i += 256) {
        const int cg = i % CG;
        const int rp = i / CG;
        const int rr = rp / RW, rc = rp - rr * RW;
        const int ih = ih0 + rr, iw = iw0 + rc;
        Pack8<T> x;
        if (ih >= 0 && ih < H && iw >= 0 && iw < W) x.load(img + ((size_t)ih * W + iw) * ldi + cg * 8);
        else x.fill_neg_inf();
        x.store(region + (size_t)rp * C + cg * 8);
    }
    for (int i = threadIdx.x; i < TH * TW * 16; i += 256) {
        const int r = i >> 4, j = i & 15;
        pooled[r * rowf + (j < 8 ? j : C + j)] = 0.f;
    }
    __syncthreads();
    // stage 2: pool from smem into the fp32 tile
    for (int i = threadIdx.x; i < TH * TW * CG; i += 256) {
        const int cg = i % CG;
        const int op = i / CG;
        const int a = op / TW, bc = op - a * TW;
        Pack8<T> m;
        m.fill_neg_inf();
#pragma unroll
        for (int r = 0; r < K; ++r)
#pragma unroll
            for (int c = 0; c < K; ++c) {
                Pack8<T> x;
                x.load(region + (size_t)((a * S + r) * RW + bc * S + c) * C + cg * 8);
                m.max_with(x);
            }
        float v[8];
        m.to_f32(v);
        float* d = pooled + op * rowf + 8 + cg * 8;
        *reinterpret_cast<float4*>(d) = make_float4(v[0], v[1], v[2], v[3]);
        *reinterpret_cast<float4*>(d + 4) = make_float4(v[4], v[5], v[6], v[7]);
    }
    __syncthreads();
    // stage 3: LRN across channels, store
    for (int i = threadIdx.x; i < TH * TW * CG; i += 256) {
        const int cg = i % CG;
        const int op = i / CG;
        const int a = op / TW, bc = op - a * TW;
        const int oh = oh0 + a, ow = ow0 + bc;
        if (oh >= Ho || ow >= Wo) continue;
        const float* sp = pooled + op * rowf + 8 + cg * 8;
        float x[8 + 2 * HALF];
#pragma unroll
        for (int j = 0; j < 8 + 2 * HALF; ++j) x[j] = sp[j - HALF];
        Vec8<T> o;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            float acc = 0.f;
#pragma unroll
            for (int d = 0; d <= 2 * HALF; ++d) acc += x[j + d] * x[j + d];
            o.v[j] = x[j + HALF] * lrn_scale<FAST>(kk + alpha_over_n * acc, beta);
        }
        o.store(out + ((size_t)(n * Ho + oh) * Wo + ow) * ldo + cg * 8);
    }
}

// ---- LRN -> Pooling -------------------------------------------------------------------------------------------
// One CTA per (image, TH x TW output tile).  smem: T [(S*(TH-1)+K) * (S*(TW-1)+K)][C].
template <typename T, int K, int S, int HALF, bool FAST, int TH, int TW, int CGT>   // CGT = C / 8 when known at compile time, else 0
__global__ void __launch_bounds__(512) lrn_pool_kernel(const T* __restrict__ in, T* __restrict__ out, int C, int H, int W,
                                                       int ldi, int Ho, int Wo, int ldo, int p, int tiles_x, int tiles_y,
                                                       float alpha_over_n, float beta, float kk) {
    extern __shared__ __align__(16) unsigned char lp_smem_raw[];
    T* tile = reinterpret_cast<T*>(lp_smem_raw);
    constexpr int RH = S * (TH - 1) + K, RW = S * (TW - 1) + K;
    const int CG = CGT ? CGT : (C >> 3);
    int b = blockIdx.x;
    const int tx = b % tiles_x; b /= tiles_x;
    const int ty = b % tiles_y;
    const int n = b / tiles_y;
    const int oh0 = ty * TH, ow0 = tx * TW;
    const int ih0 = oh0 * S - p, iw0 = ow0 * S - p;
    const T* img = in + (size_t)n * H * W * ldi;
    // stage 1: normalise the input region into shared memory (-inf outside the image: max ignores it)
    const int items = RH * RW * CG;
    for (int i = threadIdx.x; i < items; i += blockDim.x) {
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
            if (HALF == 2 && sizeof(T) == 2) {
                // the two channels either side of the group: one 32-bit load each (4-byte aligned: cg*8 - 2 and cg*8 + 8)
                const __nv_bfloat162 lo = cg > 0 ? *reinterpret_cast<const __nv_bfloat162*>(row + cg * 8 - 2) : __nv_bfloat162();
                const __nv_bfloat162 hi = cg + 1 < CG ? *reinterpret_cast<const __nv_bfloat162*>(row + cg * 8 + 8) : __nv_bfloat162();
                x[0] = cg > 0 ? __low2float(lo) : 0.f; x[1] = cg > 0 ? __high2float(lo) : 0.f;
                x[8 + HALF] = cg + 1 < CG ? __low2float(hi) : 0.f; x[8 + HALF + 1] = cg + 1 < CG ? __high2float(hi) : 0.f;
            } else {
#pragma unroll
                for (int d = 0; d < HALF; ++d) {
                    const int cl = cg * 8 - HALF + d, ch = cg * 8 + 8 + d;
                    x[d] = cl >= 0 ? to_f32<T>(row[cl]) : 0.f;
                    x[8 + HALF + d] = ch < C ? to_f32<T>(row[ch]) : 0.f;
                }
            }
            // sliding sum of squares over the 2*HALF+1 window
            float sq[8 + 2 * HALF];
#pragma unroll
            for (int j = 0; j < 8 + 2 * HALF; ++j) sq[j] = x[j] * x[j];
            float acc = 0.f;
#pragma unroll
            for (int d = 0; d < 2 * HALF; ++d) acc += sq[d];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                acc += sq[j + 2 * HALF];
                o.v[j] = x[j + HALF] * lrn_scale<FAST>(kk + alpha_over_n * acc, beta);
                acc -= sq[j];
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
    for (int i = threadIdx.x; i < oitems; i += blockDim.x) {
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

// ---- LRN(5) -> Pooling (3x3 stride 2), register marching --------------------------------------------------------
// A warp = 8 output columns x 4 channel groups (lane = column * 4 + group; the 4 groups of a pixel are 64 contiguous
// bytes), marching down RS output rows.  Every lane normalises the two input pixels (2*ow - p, 2*ow - p + 1) of its column
// for each input row; the third pixel of the window is the first pixel of the next column and comes from lane + 4 by
// shuffle, so the 8th column of a warp only feeds the 7th and a warp yields 7 output columns (1/8 of the LRN work is
// redone by the neighbouring warp; no shared memory, no barriers, no divergent work).  The row maximum of the input row
// shared with the next output row is carried in registers.  x^-0.75 costs 2 MUFU per element, which at 16 MUFU lanes per
// clock and SM is this kernel's floor (about 80 us for conv2/norm2 at batch 64); everything else is trimmed to fit under it.
template <int RS, bool FAST>
__global__ void __launch_bounds__(256) lrn5_pool3s2_march_kernel(const bf16* __restrict__ in, bf16* __restrict__ out, int C, int H,
                                                                 int W, int ldi, int Ho, int Wo, int ldo, int p, int colblocks,
                                                                 float alpha_over_n, float kk) {
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const int cl = lane >> 2, cgl = lane & 3;
    const int CG = C >> 3;
    const int task = blockIdx.x * 8 + wid;                 // (channel-group block, column block)
    const int cb = task % colblocks, gb = task / colblocks;
    if (gb * 4 >= CG) return;
    const int cg = gb * 4 + cgl;
    const int n = blockIdx.z;
    const int oh0 = blockIdx.y * RS;
    const int ow = cb * 7 + cl;
    const int iw0 = 2 * ow - p;                            // this lane's two input columns: iw0, iw0 + 1
    const bool va = iw0 >= 0 && iw0 < W, vb = iw0 + 1 >= 0 && iw0 + 1 < W;
    const int ca = min(max(iw0, 0), W - 1), cbx = min(max(iw0 + 1, 0), W - 1);
    const bool has_lo = cg > 0, has_hi = cg + 1 < CG;
    const int c_lo = has_lo ? cg * 8 - 2 : 0, c_hi = has_hi ? cg * 8 + 8 : 0;
    const bf16* img = in + (size_t)n * H * W * ldi;
    constexpr uint32_t NEG = 0xff80ff80u;

    // normalised 8 channels of one pixel, packed bf16x2 x 4 (all -inf when the pixel lies outside the image)
    auto lrn_px = [&](const uint4 q, const uint32_t lo, const uint32_t hi, const bool valid, uint32_t* o) {
        float x[12];
        x[0] = has_lo ? __uint_as_float(lo << 16) : 0.f; x[1] = has_lo ? __uint_as_float(lo & 0xffff0000u) : 0.f;
        const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) { x[2 + 2 * j] = __uint_as_float(w[j] << 16); x[3 + 2 * j] = __uint_as_float(w[j] & 0xffff0000u); }
        x[10] = has_hi ? __uint_as_float(hi << 16) : 0.f; x[11] = has_hi ? __uint_as_float(hi & 0xffff0000u) : 0.f;
        float sq[12];
#pragma unroll
        for (int j = 0; j < 12; ++j) sq[j] = x[j] * x[j];
        float acc = sq[0] + sq[1] + sq[2] + sq[3];
        float y[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            acc += sq[j + 4];
            const float r = rsqrt_ftz(kk + alpha_over_n * acc);      // beta == 0.75 only (the host checks): v^-0.75 = r*r*sqrt(r)
            y[j] = x[j + 2] * (r * (r * rsqrt_ftz(r)));
            acc -= sq[j];
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const __nv_bfloat162 h2 = __floats2bfloat162_rn(y[2 * j], y[2 * j + 1]);
            o[j] = valid ? *reinterpret_cast<const uint32_t*>(&h2) : NEG;
        }
    };
    auto max4 = [](uint32_t* a, const uint32_t* b) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const __nv_bfloat162 m = __hmax2(*reinterpret_cast<const __nv_bfloat162*>(&a[j]), *reinterpret_cast<const __nv_bfloat162*>(&b[j]));
            a[j] = *reinterpret_cast<const uint32_t*>(&m);
        }
    };
    // horizontal 3-max of input row ih for this lane's output column
    struct Raw { uint4 qa, qb; uint32_t la, ha, lb, hb; };
    // per-lane base pointers of the six loads of a row (own 8 channels and the two neighbour pairs of both pixels); a row
    // then costs one 64-bit multiply and six pointer additions
    const bf16* const pa0 = img + (size_t)ca * ldi + cg * 8;
    const bf16* const pb0 = img + (size_t)cbx * ldi + cg * 8;
    const int d_lo = c_lo - cg * 8, d_hi = c_hi - cg * 8;
    const size_t pitch = (size_t)W * ldi;
    auto load_row = [&](int ih, Raw& r) {
        const size_t off = (size_t)min(max(ih, 0), H - 1) * pitch;
        const bf16* pa = pa0 + off;
        const bf16* pb = pb0 + off;
        r.qa = __ldg(reinterpret_cast<const uint4*>(pa));
        r.qb = __ldg(reinterpret_cast<const uint4*>(pb));
        r.la = __ldg(reinterpret_cast<const uint32_t*>(pa + d_lo)); r.ha = __ldg(reinterpret_cast<const uint32_t*>(pa + d_hi));
        r.lb = __ldg(reinterpret_cast<const uint32_t*>(pb + d_lo)); r.hb = __ldg(reinterpret_cast<const uint32_t*>(pb + d_hi));
    };
    auto row_max = [&](int ih, const Raw& r, uint32_t* m) {
        const bool vr = ih >= 0 && ih < H;
        uint32_t b[4], t[4];
        lrn_px(r.qa, r.la, r.ha, vr && va, m);
        lrn_px(r.qb, r.lb, r.hb, vr && vb, b);
#pragma unroll
        for (int j = 0; j < 4; ++j) t[j] = __shfl_down_sync(0xffffffffu, m[j], 4);     // first pixel of the next column
        max4(m, b);
        max4(m, t);
    };

    int ih = 2 * oh0 - p;
    uint32_t carry[4];
    {
        Raw r0;
        load_row(ih, r0);
        row_max(ih, r0, carry);
    }
    const bool col_ok = cl < 7 && ow < Wo;
#pragma unroll 1
    for (int o = 0; o < RS; ++o) {
        const int oh = oh0 + o;
        if (oh >= Ho) break;
        Raw r1, r2;
        load_row(ih + 1, r1);
        load_row(ih + 2, r2);
        uint32_t m1[4], m2[4];
        row_max(ih + 1, r1, m1);
        row_max(ih + 2, r2, m2);
        max4(carry, m1);
        max4(carry, m2);
        if (col_ok)
            *reinterpret_cast<uint4*>(out + ((size_t)(n * Ho + oh) * Wo + ow) * ldo + cg * 8) = make_uint4(carry[0], carry[1], carry[2], carry[3]);
#pragma unroll
        for (int j = 0; j < 4; ++j) carry[j] = m2[j];
        ih += 2;
    }
}

// ---- data (NCHW fp32) -> scale*x -> 2x2 space-to-depth NHWC bf16, W-padded -------------------------------------
// out[n][PADL + h2][PADL + w2][(dy*2+dx)*C + c] = bf16(in[n][c][2*h2+dy][2*w2+dx]);  channels >= 4*C and the pad
// rows / columns stay zero (written once at allocation).  Feeds the first convolution (stride 2) as a stride-1
// tensor-core convolution (see net.cu: plan_conv_s2d).
template <int CT>   // CT = channel count known at compile time (1..4, 16 output channels per pixel), or 0 = generic
__global__ void __launch_bounds__(256) input_s2d_kernel(const float* __restrict__ in, bf16* __restrict__ out, int N, int C_rt,
                                                        int H, int W, int Wp, int Hp, int padl, int ldc) {
    // The image is stored [N][Hp][Wp][16] with padl zero rows above / zero columns left of pixel (0, 0), and PRE-SWIZZLED
    // for the 32-byte shared-memory swizzle of the first convolution's A operand: the two 16-byte halves of a pixel are
    // exchanged when bit 2 of its physical column is set (a bulk copy of a row then lands already swizzled).
    const int C = CT > 0 ? CT : C_rt;
    // one thread per space-to-depth pixel: 2 rows x float2 per channel in (coalesced over w2), ldc bf16 out as 16-byte stores
    const int H2 = H >> 1, W2 = W >> 1;
    const unsigned total = (unsigned)N * H2 * W2;
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        const int w2 = i % W2;
        unsigned t = i / W2;
        const int h2 = t % H2;
        const int n = t / H2;
        __align__(16) bf16 v[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) v[j] = __float2bfloat16_rn(0.f);
        if (CT > 0) {
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                if (c < CT) {
                    const float* src = in + ((size_t)(n * C + c) * H + 2 * h2) * W + 2 * w2;
                    const float2 r0 = __ldg(reinterpret_cast<const float2*>(src));
                    const float2 r1 = __ldg(reinterpret_cast<const float2*>(src + W));
                    v[0 * CT + c] = __float2bfloat16_rn(r0.x);
                    v[1 * CT + c] = __float2bfloat16_rn(r0.y);
                    v[2 * CT + c] = __float2bfloat16_rn(r1.x);
                    v[3 * CT + c] = __float2bfloat16_rn(r1.y);
                }
            }
        } else {
            for (int c = 0; c < C; ++c) {
                const float* src = in + ((size_t)(n * C + c) * H + 2 * h2) * W + 2 * w2;
                v[0 * C + c] = __float2bfloat16_rn(src[0]);
                v[1 * C + c] = __float2bfloat16_rn(src[1]);
                v[2 * C + c] = __float2bfloat16_rn(src[W]);
                v[3 * C + c] = __float2bfloat16_rn(src[W + 1]);
            }
        }
        const int xp = padl + w2;
        const int swz = (xp >> 2) & 1;
        uint4* o = reinterpret_cast<uint4*>(out + ((size_t)(n * Hp + padl + h2) * Wp + xp) * 16);
        o[swz] = *reinterpret_cast<const uint4*>(v);
        o[swz ^ 1] = *reinterpret_cast<const uint4*>(v + 8);
    }
}

}  // namespace fcnod
