// One training iteration of /root/reference/models/train_val.prototxt on the device (BASELINE config 5):
// forward at TRAIN phase, the fused loss head (train_head.cu), the backward pass of every layer of the body, and Caffe's
// ADAM ApplyUpdate (/root/reference/train/bounding_box/solver.prototxt:3-15).  fp32 storage and arithmetic throughout: this
// is the PARITY engine of the training step (1e-4 against torch autograd over the same graph in the tests), the
// counterpart of the fp32 inference plan; convolutions run on the SIMT implicit-GEMM kernel (conv_simt.cuh) in all three
// roles:
//   forward   y  = conv(x, W) + b                                   conv_simt_kernel
//   dgrad     dx = conv(dy, rot180(W)^T), pad k-1-p                 conv_simt_kernel on weights re-packed by dgrad_pack_kernel
//   wgrad     dW[(r,s,c)][co] = sum_px x[px + (r,s)][c] * dy[px][co]   wgrad_simt_kernel (split over pixel chunks, fp32 atomics)
// Layer semantics restated [ext: Caffe layer sources, not vendored upstream]:
//   Pooling MAX  forward keeps the index of the FIRST maximum of the clipped window in row-major scan order (pooling_layer.cpp's
//                `>` against -FLT_MAX); backward routes the gradient to that element
//   LRN          scale = k + alpha/n * sum x^2 over the clipped channel window, y = x * scale^-beta;
//                dx_c = dy_c * scale_c^-beta - 2 alpha beta / n * x_c * sum_{c' in window} dy_c' * y_c' / scale_c'
//   Dropout      TRAIN: y = x * mask / (1 - ratio), mask ~ Bernoulli(1 - ratio) per element; same mask in backward
//   ReLU         in place on the convolution's top: dy *= (y > 0)
//   Sigmoid      dx = dy * y * (1 - y)
//   Concat       tops of the branches ARE channel slices of the concat blob (data and diff): no copy either way
//   Split        a blob with several consumers sums their gradients: the first writer of a diff overwrites, later ones add
// Activations are NHWC fp32; parameters live in ONE flat fp32 buffer (weights packed [(r,s,c)][Cout padded to 64], then the
// bias, every blob padded to 256 floats) with the gradients, and ADAM's two moments, in three more buffers of the same
// layout, so the gradient all-reduce of a multi-GPU step is one collective over one contiguous buffer and the update is one
// launch.  Padding entries have zero gradient and stay zero.
#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <map>
#include <vector>

#include "common.cuh"
#include "conv_simt.cuh"
#include "gemm_tc.cuh"
#include "kernels.h"
#include "netdef.h"

namespace fcnod {
namespace {

// programmatic dependent launch: wait = the previous kernel of the stream has completed and its writes are visible (a no-op
// when the kernel was not launched as a programmatic dependent); trigger = the next kernel may be scheduled
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

inline int grid1(long n, int threads) { return (int)std::min<long>((n + threads - 1) / threads, 148L * 32); }

// ------------------------------------------------------------------------------------------------ layout kernels
// data NCHW fp32 -> NHWC with the Power layer applied (train_val.prototxt:27-35: shift + scale * x, power 1)
template <typename T>
__global__ void nchw_to_nhwc_power_kernel(const float* __restrict__ in, T* __restrict__ out, int N, int C, int H, int W,
                                          float scale, float shift) {
    const long total = (long)N * C * H * W;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        const int c = (int)(i % C);
        long t = i / C;
        const int w = (int)(t % W); t /= W;
        const int h = (int)(t % H);
        const int n = (int)(t / H);
        out[i] = from_f32<T>(__fadd_rn(shift, __fmul_rn(scale, in[((long)(n * C + c) * H + h) * W + w])));
    }
}

// NHWC view (ldc, channel offset folded into the pointer) <-> dense NCHW
template <typename T>
__global__ void nhwc_to_nchw_kernel(const T* __restrict__ in, int ldc, float* __restrict__ out, int N, int C, int H, int W) {
    const long total = (long)N * C * H * W;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        const int w = (int)(i % W);
        long t = i / W;
        const int h = (int)(t % H); t /= H;
        const int c = (int)(t % C);
        const int n = (int)(t / C);
        out[i] = to_f32<T>(in[((long)(n * H + h) * W + w) * ldc + c]);
    }
}
template <typename T>
__global__ void nchw_to_nhwc_kernel(const float* __restrict__ in, T* __restrict__ out, int ldc, int N, int C, int H, int W) {
    const long total = (long)N * C * H * W;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        const int w = (int)(i % W);
        long t = i / W;
        const int h = (int)(t % H); t /= H;
        const int c = (int)(t % C);
        const int n = (int)(t / C);
        out[((long)(n * H + h) * W + w) * ldc + c] = from_f32<T>(in[i]);
    }
}
__global__ void u8_nhwc_to_nchw_kernel(const uint8_t* __restrict__ in, uint8_t* __restrict__ out, int N, int C, int H, int W) {
    const long total = (long)N * C * H * W;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        const int w = (int)(i % W);
        long t = i / W;
        const int h = (int)(t % H); t /= H;
        const int c = (int)(t % C);
        const int n = (int)(t / C);
        out[i] = in[((long)(n * H + h) * W + w) * C + c];
    }
}

// ------------------------------------------------------------------------------------------------ pooling
struct PoolGeo { int N, C, H, W, Ho, Wo, k, s, p, ldi, ldo; };

template <typename T>
__global__ void pool_fwd_kernel(const T* __restrict__ x, T* __restrict__ y, uint8_t* __restrict__ idx, PoolGeo g) {
    const long total = (long)g.N * g.Ho * g.Wo * g.C;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        const int c = (int)(i % g.C);
        long t = i / g.C;
        const int pw = (int)(t % g.Wo); t /= g.Wo;
        const int ph = (int)(t % g.Ho);
        const int n = (int)(t / g.Ho);
        const int h0 = ph * g.s - g.p, w0 = pw * g.s - g.p;
        const int hs = max(h0, 0), ws = max(w0, 0);
        const int he = min(min(h0 + g.k, g.H + g.p), g.H), we = min(min(w0 + g.k, g.W + g.p), g.W);
        float best = -3.402823466e+38f;
        int bi = 0;
        for (int h = hs; h < he; ++h)
            for (int w = ws; w < we; ++w) {
                const float v = to_f32<T>(x[((long)(n * g.H + h) * g.W + w) * g.ldi + c]);
                if (v > best) { best = v; bi = (h - h0) * g.k + (w - w0); }
            }
        y[((long)(n * g.Ho + ph) * g.Wo + pw) * g.ldo + c] = from_f32<T>(best);
        idx[i] = (uint8_t)bi;
    }
}

template <typename T>
__global__ void pool_bwd_kernel(const T* __restrict__ dy, const uint8_t* __restrict__ idx, T* __restrict__ dx, PoolGeo g,
                                int accum) {
    const long total = (long)g.N * g.H * g.W * g.C;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        const int c = (int)(i % g.C);
        long t = i / g.C;
        const int w = (int)(t % g.W); t /= g.W;
        const int h = (int)(t % g.H);
        const int n = (int)(t / g.H);
        const int ph0 = (h + g.p < g.k) ? 0 : (h + g.p - g.k) / g.s + 1, ph1 = min((h + g.p) / g.s + 1, g.Ho);
        const int pw0 = (w + g.p < g.k) ? 0 : (w + g.p - g.k) / g.s + 1, pw1 = min((w + g.p) / g.s + 1, g.Wo);
        float acc = 0.f;
        for (int ph = ph0; ph < ph1; ++ph)
            for (int pw = pw0; pw < pw1; ++pw) {
                const long o = ((long)(n * g.Ho + ph) * g.Wo + pw);
                const int local = (h - (ph * g.s - g.p)) * g.k + (w - (pw * g.s - g.p));
                if (idx[o * g.C + c] == local) acc += to_f32<T>(dy[o * g.ldo + c]);
            }
        const long xi = ((long)(n * g.H + h) * g.W + w) * g.ldi + c;
        dx[xi] = from_f32<T>(accum ? to_f32<T>(dx[xi]) + acc : acc);
    }
}

// ------------------------------------------------------------------------------------------------ LRN (across channels)
template <typename T>
__global__ void lrn_fwd_kernel(const T* __restrict__ x, int ldi, T* __restrict__ y, int ldo, float* __restrict__ scale,
                               long npix, int C, int n, float alpha_over_n, float beta, float k) {
    const long total = npix * C;
    const int half = n / 2;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        const int c = (int)(i % C);
        const long px = i / C;
        const T* xp = x + px * ldi;
        float s = 0.f;
        for (int cc = max(c - half, 0); cc <= min(c + half, C - 1); ++cc) { const float xv = to_f32<T>(xp[cc]); s = fmaf(xv, xv, s); }
        const float sc = fmaf(alpha_over_n, s, k);
        scale[i] = sc;
        y[px * ldo + c] = from_f32<T>(to_f32<T>(xp[c]) * powf(sc, -beta));
    }
}

template <typename T>
__global__ void lrn_bwd_kernel(const T* __restrict__ x, int ldi, const T* __restrict__ y, const T* __restrict__ dy,
                               int ldo, const float* __restrict__ scale, T* __restrict__ dx, long npix, int C, int n,
                               float alpha_over_n, float beta, int accum) {
    const long total = npix * C;
    const int half = n / 2;
    const float cr = 2.f * alpha_over_n * beta;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        const int c = (int)(i % C);
        const long px = i / C;
        float s = 0.f;
        for (int cc = max(c - half, 0); cc <= min(c + half, C - 1); ++cc)
            s += to_f32<T>(dy[px * ldo + cc]) * to_f32<T>(y[px * ldo + cc]) / scale[px * C + cc];
        const float v = to_f32<T>(dy[px * ldo + c]) * powf(scale[i], -beta) - cr * to_f32<T>(x[px * ldi + c]) * s;
        const long xi = px * ldi + c;
        dx[xi] = from_f32<T>(accum ? to_f32<T>(dx[xi]) + v : v);
    }
}

// ------------------------------------------------------------------------------------------------ elementwise
__device__ __forceinline__ uint32_t mix32(uint64_t z) {        // splitmix64 finaliser
    z += 0x9e3779b97f4a7c15ull;
    z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
    z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
    return (uint32_t)((z ^ (z >> 31)) >> 32);
}

// Bernoulli(1 - ratio) keep mask: counter-based (seed, iteration, element index), so a step is reproducible
__global__ void dropout_mask_kernel(uint8_t* __restrict__ mask, long total, uint64_t seed, uint64_t iter, float ratio) {
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        const uint32_t u = mix32(seed * 0x100000001b3ull + iter * 0x9e3779b97f4a7c15ull + (uint64_t)i);
        mask[i] = ((float)(u >> 8) * (1.f / 16777216.f) >= ratio) ? 1 : 0;
    }
}
// y = x * mask * scale (forward) and dx (+)= dy * mask * scale (backward): same kernel
template <typename T>
__global__ void dropout_apply_kernel(const T* __restrict__ in, const uint8_t* __restrict__ mask, T* __restrict__ out,
                                     long total, float scale, int accum) {
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        const float v = mask[i] ? __fmul_rn(to_f32<T>(in[i]), scale) : 0.f;
        out[i] = from_f32<T>(accum ? to_f32<T>(out[i]) + v : v);
    }
}
template <typename T>
__global__ void sigmoid_fwd_kernel(const T* __restrict__ x, T* __restrict__ y, long total) {
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x)
        y[i] = from_f32<T>(1.f / (1.f + expf(-to_f32<T>(x[i]))));
}
template <typename T>
__global__ void sigmoid_bwd_kernel(const T* __restrict__ y, const T* __restrict__ dy, T* __restrict__ dx, long total,
                                   int accum) {
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        const float yv = to_f32<T>(y[i]);
        const float v = to_f32<T>(dy[i]) * yv * (1.f - yv);
        dx[i] = from_f32<T>(accum ? to_f32<T>(dx[i]) + v : v);
    }
}
// in-place ReLU backward on a channel slice: dy = y > 0 ? dy : 0
template <typename T>
__global__ void relu_mask_kernel(const T* __restrict__ y, T* __restrict__ dy, long npix, int C, int ldc) {
    const long total = npix * C;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        const long o = (i / C) * ldc + (i % C);
        if (!(to_f32<T>(y[o]) > 0.f)) dy[o] = from_f32<T>(0.f);
    }
}

// ------------------------------------------------------------------------------------------------ convolution backward
// bias gradient: db[c] += sum over pixels of dy[px][c].  grid (ceil(C / 32), pixel chunks), block 32 x 8.
template <typename T>
__global__ void __launch_bounds__(256) bias_grad_kernel(const T* __restrict__ dy, int ldc, int C, long npix, long chunk,
                                                         float* __restrict__ db) {
    __shared__ float red[8][33];
    const int c = blockIdx.x * 32 + threadIdx.x;
    const long p0 = blockIdx.y * chunk, p1 = min(p0 + chunk, npix);
    float s = 0.f;
    if (c < C)
        for (long p = p0 + threadIdx.y; p < p1; p += 8) s += to_f32<T>(dy[p * ldc + c]);
    red[threadIdx.y][threadIdx.x] = s;
    __syncthreads();
    if (threadIdx.y == 0 && c < C) {
        float a = 0.f;
#pragma unroll
        for (int j = 0; j < 8; ++j) a += red[j][threadIdx.x];
        atomicAdd(db + c, a);
    }
}

struct WgradParams {
    const void* x;      // bottom, NHWC view
    const void* dy;     // top diff, NHWC view
    float* g;           // [K][ldw] packed weight gradient (accumulated with atomics; zeroed by the caller)
    int N, H, W, Cin, ldi;
    int Ho, Wo, Cout, ldo, ldw;
    int k, stride, pad;
    long chunk;         // pixels per blockIdx.z
};

// G[kk][co] += sum_m A[m][kk] * dy[m][co], A = im2col(x) gathered on the fly, kk = (r*k + s)*Cin + c.
// Tile 64 (kk) x 64 (co), 16 pixels per step, 256 threads, 4x4 register micro-tile.
template <typename T>
__global__ void __launch_bounds__(256) wgrad_simt_kernel(WgradParams p) {
    constexpr int BK = 64, BN = 64, BP = 16;
    __shared__ float As[BP][BK + 4];
    __shared__ float Bs[BP][BN + 4];
    const int tid = threadIdx.x;
    const int tx = tid % 16, ty = tid / 16;
    const int K = p.k * p.k * p.Cin;
    const long M = (long)p.N * p.Ho * p.Wo;
    const int k0 = blockIdx.x * BK, n0 = blockIdx.y * BN;
    const long m_begin = blockIdx.z * p.chunk, m_end = min(m_begin + p.chunk, M);

    // this thread's gather column: kk = k0 + tid % 64 -> (c, r, s), fixed for the whole loop
    const int kl = tid % BK, prow = tid / BK;        // pixel rows prow, prow + 4, prow + 8, prow + 12
    const int kk = k0 + kl;
    const bool kok = kk < K;
    int c = 0, r = 0, s = 0;
    if (kok) { c = kk % p.Cin; const int rs = kk / p.Cin; s = rs % p.k; r = rs / p.k; }
    const int co = n0 + tid % BN;
    const bool cok = co < p.Cout;
    const T* px_ = static_cast<const T*>(p.x);
    const T* pdy = static_cast<const T*>(p.dy);

    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

    for (long m0 = m_begin; m0 < m_end; m0 += BP) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const long m = m0 + prow + 4 * i;
            float a = 0.f, b = 0.f;
            if (m < m_end) {
                const int ow = (int)(m % p.Wo);
                const long t = m / p.Wo;
                const int oh = (int)(t % p.Ho);
                const int n = (int)(t / p.Ho);
                const int ih = oh * p.stride - p.pad + r, iw = ow * p.stride - p.pad + s;
                if (kok && ih >= 0 && ih < p.H && iw >= 0 && iw < p.W) a = to_f32<T>(px_[((long)(n * p.H + ih) * p.W + iw) * p.ldi + c]);
                if (cok) b = to_f32<T>(pdy[m * p.ldo + co]);
            }
            As[prow + 4 * i][kl] = a;
            Bs[prow + 4 * i][tid % BN] = b;
        }
        __syncthreads();
#pragma unroll
        for (int q = 0; q < BP; ++q) {
            float a[4], b[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) a[i] = As[q][ty * 4 + i];
#pragma unroll
            for (int j = 0; j < 4; ++j) b[j] = Bs[q][tx * 4 + j];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int row = k0 + ty * 4 + i;
        if (row >= K) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int col = n0 + tx * 4 + j;
            if (col < p.Cout) atomicAdd(p.g + (long)row * p.ldw + col, acc[i][j]);
        }
    }
}

// weights packed for the forward, P[((r*k + s)*Cin + c)][ldw] (co fastest), re-packed for the data gradient:
// T[((r'*k + s')*Cout + co)][ldt] (c fastest) = P[(((k-1-r')*k + (k-1-s'))*Cin + c)][co]
__global__ void dgrad_pack_kernel(const float* __restrict__ P, int ldw, float* __restrict__ T, int ldt, int Cin, int Cout, int k) {
    const long total = (long)k * k * Cout * ldt;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        const int c = (int)(i % ldt);
        long t = i / ldt;
        const int co = (int)(t % Cout);
        const int rs = (int)(t / Cout);
        const int r = rs / k, s = rs % k;
        T[i] = c < Cin ? P[((long)((k - 1 - r) * k + (k - 1 - s)) * Cin + c) * ldw + co] : 0.f;
    }
}

// ------------------------------------------------------------------------------------------------ solver
struct AdamArgs {
    float rate, correction, b1, b2, eps, weight_decay, grad_scale;
};
// One launch over the flat parameter buffer; meta[i / 256] = (lr_mult, decay_mult) of the blob that owns element i.
// Caffe's order: Normalize / all-reduce scale, Regularize (L2), ComputeUpdateValue (ADAM), Update.
__global__ void adam_kernel(float* __restrict__ w, float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
                            const float2* __restrict__ meta, long total, AdamArgs a) {
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        const float2 mt = meta[i >> 8];
        float gi = __fmul_rn(g[i], a.grad_scale);
        gi = __fadd_rn(gi, __fmul_rn(__fmul_rn(a.weight_decay, mt.y), w[i]));
        const float mi = __fadd_rn(__fmul_rn(a.b1, m[i]), __fmul_rn(1.f - a.b1, gi));
        const float vi = __fadd_rn(__fmul_rn(a.b2, v[i]), __fmul_rn(1.f - a.b2, __fmul_rn(gi, gi)));
        m[i] = mi; v[i] = vi;
        const float step = __fmul_rn(__fmul_rn(__fmul_rn(a.rate, mt.x), a.correction), __fdiv_rn(mi, __fadd_rn(__fsqrt_rn(vi), a.eps)));
        w[i] = __fsub_rn(w[i], step);
        g[i] = gi;
    }
}

// Exact division of a small dividend by a run-time divisor through a precomputed reciprocal: q = (x * magic) >> 32 with
// magic = ceil(2^32 / d) is exact for x < 2^16 (the per-row counts below), one multiply instead of an integer division.
__host__ __device__ __forceinline__ unsigned div_magic(unsigned d) { return (unsigned)((0x100000000ull + d - 1) / d); }
__device__ __forceinline__ unsigned div_fast(unsigned x, unsigned magic) { return __umulhi(x, magic); }

// ------------------------------------------------------------------------------------------------ bf16 engine: 8 channels per thread
// Same arithmetic as the generic kernels above on 16-byte groups of channels (all blobs of the network have C % 8 == 0 except
// the 3-channel input and the two heads, which keep the generic kernels).
struct BF8 {
    uint4 v;
    __device__ __forceinline__ float get(int q) const { return __bfloat162float(reinterpret_cast<const bf16*>(&v)[q]); }
    __device__ __forceinline__ void set(int q, float f) { reinterpret_cast<bf16*>(&v)[q] = __float2bfloat16_rn(f); }
};
__device__ __forceinline__ BF8 ld8(const bf16* p) { BF8 r; r.v = *reinterpret_cast<const uint4*>(p); return r; }
__device__ __forceinline__ void st8(bf16* p, const BF8& r) { *reinterpret_cast<uint4*>(p) = r.v; }

template <int K, int S>      // compile-time window and stride (3x3 / 2 and 3x3 / 1 in this network): unrolled loops, no divisions
__global__ void pool_fwd8_kernel(const bf16* __restrict__ x, bf16* __restrict__ y, uint8_t* __restrict__ idx, PoolGeo g, unsigned m_cv) {
    pdl_wait();
    pdl_trigger();
    // grid (groups of an output row, Ho, N): one thread per (output pixel, 8 channels)
    const unsigned cv = (unsigned)g.C / 8u;
    const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (unsigned)g.Wo * cv) return;
    const int pw = (int)__umulhi(i, m_cv), c = (int)(i - (unsigned)pw * cv) * 8;
    const int ph = blockIdx.y, n = blockIdx.z;
    const int h0 = ph * S - g.p, w0 = pw * S - g.p;
    // Two channels per instruction: the values ARE bf16, so a bf16 compare is exact; `>` against the lowest finite bf16 keeps the
    // first maximum in scan order (Caffe's rule), the window position travels in the same 16-bit lanes.
    uint32_t best[4], bi[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) { best[q] = 0xff7fff7fu; bi[q] = 0u; }
    const bf16* xn = x + (size_t)n * g.H * g.W * g.ldi + c;
#pragma unroll
    for (int dh = 0; dh < K; ++dh)
#pragma unroll
        for (int dw = 0; dw < K; ++dw) {
            const int h = h0 + dh, w = w0 + dw;
            if (h < 0 || h >= g.H || w < 0 || w >= g.W) continue;          // (the pad rule min(h0 + k, H + p) never bites for pad < k)
            const uint4 v = *reinterpret_cast<const uint4*>(xn + ((size_t)h * g.W + w) * g.ldi);
            const uint32_t loc2 = (uint32_t)(dh * K + dw) * 0x00010001u;
            const uint32_t vv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const uint32_t m = __hgt2_mask(*reinterpret_cast<const __nv_bfloat162*>(&vv[q]), *reinterpret_cast<const __nv_bfloat162*>(&best[q]));
                best[q] = (vv[q] & m) | (best[q] & ~m);
                bi[q] = (loc2 & m) | (bi[q] & ~m);
            }
        }
    const size_t o = ((size_t)n * g.Ho + ph) * g.Wo + pw;
    *reinterpret_cast<uint4*>(y + o * g.ldo + c) = make_uint4(best[0], best[1], best[2], best[3]);
    // positions: low bytes of the eight 16-bit lanes -> eight consecutive bytes
    uint2 ib;
    ib.x = __byte_perm(bi[0], bi[1], 0x6420);
    ib.y = __byte_perm(bi[2], bi[3], 0x6420);
    *reinterpret_cast<uint2*>(idx + o * g.C + c) = ib;
}

template <int K, int S>
__global__ void pool_bwd8_kernel(const bf16* __restrict__ dy, const uint8_t* __restrict__ idx, bf16* __restrict__ dx, PoolGeo g, int accum,
                                 unsigned m_cv) {
    pdl_wait();
    pdl_trigger();
    // grid (groups of an input row, H, N): one thread per (input pixel, 8 channels)
    const unsigned cv = (unsigned)g.C / 8u;
    const unsigned i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (unsigned)g.W * cv) return;
    const int w = (int)__umulhi(i, m_cv), c = (int)(i - (unsigned)w * cv) * 8;
    const int h = blockIdx.y, n = blockIdx.z;
    const int ph0 = (h + g.p < K) ? 0 : (h + g.p - K) / S + 1, ph1 = min((h + g.p) / S + 1, g.Ho);
    const int pw0 = (w + g.p < K) ? 0 : (w + g.p - K) / S + 1, pw1 = min((w + g.p) / S + 1, g.Wo);
    float acc[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) acc[q] = 0.f;
    constexpr int NW = (K + S - 1) / S;          // windows that can cover a pixel, per axis
#pragma unroll
    for (int a = 0; a < NW; ++a)
#pragma unroll
        for (int b = 0; b < NW; ++b) {
            const int ph = ph0 + a, pw = pw0 + b;
            if (ph >= ph1 || pw >= pw1) continue;
            const size_t o = ((size_t)n * g.Ho + ph) * g.Wo + pw;
            const int local = (h - (ph * S - g.p)) * K + (w - (pw * S - g.p));
            const uint2 ib = *reinterpret_cast<const uint2*>(idx + o * g.C + c);
            const uint4 d = *reinterpret_cast<const uint4*>(dy + o * g.ldo + c);
            // per-byte equality with the wanted position (0xff where equal), widened to the 16-bit lanes of the gradients;
            // a bf16 is the high half of its float
            const uint32_t want = (uint32_t)local * 0x01010101u;
            const uint32_t e0 = __vcmpeq4(ib.x, want), e1 = __vcmpeq4(ib.y, want);
            const uint32_t dd[4] = {d.x & __byte_perm(e0, 0, 0x1100), d.y & __byte_perm(e0, 0, 0x3322), d.z & __byte_perm(e1, 0, 0x1100),
                                    d.w & __byte_perm(e1, 0, 0x3322)};
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                acc[2 * q] += __uint_as_float(dd[q] << 16);
                acc[2 * q + 1] += __uint_as_float(dd[q] & 0xffff0000u);
            }
        }
    bf16* o = dx + (((size_t)n * g.H + h) * g.W + w) * g.ldi + c;
    BF8 r;
    if (accum) {
        r = ld8(o);
#pragma unroll
        for (int q = 0; q < 8; ++q) acc[q] += r.get(q);
    }
#pragma unroll
    for (int q = 0; q < 8; ++q) r.set(q, acc[q]);
    st8(o, r);
}

// ReLU backward (in place on the convolution's top diff) and the bias gradient in one pass over dy: block = 8 channel groups
// (64 channels) x 32 pixel lanes; db[c] += sum over the block's pixel chunk.  relu = 0: bias gradient only.
// segs.n > 0: the blob is a Concat top written by several convolutions; channel cc belongs to the last segment whose c0 <= cc
// and its bias gradient goes to that convolution's slot (db unused)
struct RbSeg { int c0; float* db; };
struct RbSegs { int n; RbSeg s[8]; };
__global__ void __launch_bounds__(256) relu_bias8_kernel(const bf16* __restrict__ y, bf16* __restrict__ dy, unsigned npix, int C, int ldc,
                                                          unsigned chunk, int relu, float* __restrict__ db, RbSegs segs) {
    pdl_wait();
    pdl_trigger();
    __shared__ float red[32][65];
    const int c = (blockIdx.x * 8 + threadIdx.x) * 8;
    const unsigned p0 = blockIdx.y * chunk, p1 = min(p0 + chunk, npix);
    float acc[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) acc[q] = 0.f;
    if (c < C)
        for (unsigned p = p0 + threadIdx.y; p < p1; p += 64) {
            // two pixels per iteration: four 16-byte loads in flight per thread
            const unsigned pb = p + 32;
            const bool hb = pb < p1;
            const size_t o = (size_t)p * ldc + c, ob = (size_t)pb * ldc + c;
            BF8 d = ld8(dy + o), db2, a, ab;
            if (hb) db2 = ld8(dy + ob);
            if (relu) { a = ld8(y + o); if (hb) ab = ld8(y + ob); }
            if (relu) {
                bool any = false;
#pragma unroll
                for (int q = 0; q < 8; ++q)
                    if (!(a.get(q) > 0.f)) { d.set(q, 0.f); any = true; }
                if (any) st8(dy + o, d);
                if (hb) {
                    any = false;
#pragma unroll
                    for (int q = 0; q < 8; ++q)
                        if (!(ab.get(q) > 0.f)) { db2.set(q, 0.f); any = true; }
                    if (any) st8(dy + ob, db2);
                }
            }
#pragma unroll
            for (int q = 0; q < 8; ++q) acc[q] += d.get(q) + (hb ? db2.get(q) : 0.f);
        }
#pragma unroll
    for (int q = 0; q < 8; ++q) red[threadIdx.y][threadIdx.x * 8 + q] = acc[q];
    __syncthreads();
    const int tid = threadIdx.y * 8 + threadIdx.x;
    if (tid < 64) {
        float a = 0.f;
#pragma unroll
        for (int j = 0; j < 32; ++j) a += red[j][tid];
        const int cc = blockIdx.x * 64 + tid;
        if (cc < C) {
            float* dst = db + cc;
            if (segs.n > 0) {
                int k = 0;
#pragma unroll
                for (int j = 1; j < 8; ++j) if (j < segs.n && cc >= segs.s[j].c0) k = j;
                dst = segs.s[k].db + (cc - segs.s[k].c0);
            }
            atomicAdd(dst, a);
        }
    }
}

// LRN, 8 channels per thread; the two neighbours on either side come from scalar loads (L1 hits)
__global__ void lrn_fwd8_kernel(const bf16* __restrict__ x, int ldi, bf16* __restrict__ y, int ldo, float* __restrict__ scale, long npix,
                                int C, float alpha_over_n, float beta, float k) {
    pdl_wait();
    pdl_trigger();
    const int cv = C / 8;
    const unsigned total = (unsigned)(npix * cv);
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {      // 32-bit index arithmetic (checked on the host)
        const int c = (int)(i % cv) * 8;
        const unsigned px = i / cv;
        const bf16* xp = x + px * ldi;
        float v[12];
        const BF8 a = ld8(xp + c);
#pragma unroll
        for (int q = 0; q < 8; ++q) v[q + 2] = a.get(q);
        v[0] = c >= 2 ? __bfloat162float(xp[c - 2]) : 0.f; v[1] = c >= 1 ? __bfloat162float(xp[c - 1]) : 0.f;
        v[10] = c + 8 < C ? __bfloat162float(xp[c + 8]) : 0.f; v[11] = c + 9 < C ? __bfloat162float(xp[c + 9]) : 0.f;
        BF8 r;
        float sc[8];
#pragma unroll
        for (int q = 0; q < 8; ++q) {
            float s = 0.f;
#pragma unroll
            for (int d = 0; d < 5; ++d) s = fmaf(v[q + d], v[q + d], s);
            sc[q] = fmaf(alpha_over_n, s, k);
            r.set(q, v[q + 2] * __powf(sc[q], -beta));      // bf16 engine: fast exp2 / log2 (the result is rounded to 8 bits)
        }
        st8(y + px * ldo + c, r);
        float4* so = reinterpret_cast<float4*>(scale + px * C + c);
        so[0] = make_float4(sc[0], sc[1], sc[2], sc[3]);
        so[1] = make_float4(sc[4], sc[5], sc[6], sc[7]);
    }
}

__global__ void lrn_bwd8_kernel(const bf16* __restrict__ x, int ldi, const bf16* __restrict__ y, const bf16* __restrict__ dy, int ldo,
                                const float* __restrict__ scale, bf16* __restrict__ dx, long npix, int C, float alpha_over_n, float beta,
                                int accum) {
    pdl_wait();
    pdl_trigger();
    const int cv = C / 8;
    const unsigned total = (unsigned)(npix * cv);
    const float cr = 2.f * alpha_over_n * beta;
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {      // 32-bit index arithmetic (checked on the host)
        const int c = (int)(i % cv) * 8;
        const unsigned px = i / cv;
        float ratio[12], sc[8], dyv[8];
        {   // the thread's own eight channels: 16-byte loads
            const BF8 dv = ld8(dy + px * ldo + c), yv = ld8(y + px * ldo + c);
            const float4* sp = reinterpret_cast<const float4*>(scale + px * C + c);
            const float4 s0 = sp[0], s1 = sp[1];
            sc[0] = s0.x; sc[1] = s0.y; sc[2] = s0.z; sc[3] = s0.w; sc[4] = s1.x; sc[5] = s1.y; sc[6] = s1.z; sc[7] = s1.w;
#pragma unroll
            for (int q = 0; q < 8; ++q) { dyv[q] = dv.get(q); ratio[q + 2] = __fdividef(dyv[q] * yv.get(q), sc[q]); }
        }
#pragma unroll
        for (int q = 0; q < 4; ++q) {      // two neighbours on either side (L1 hits: the next thread's vectors)
            const int qq = q < 2 ? q : q + 8, cc = c - 2 + qq;
            ratio[qq] = 0.f;
            if (cc >= 0 && cc < C)
                ratio[qq] = __fdividef(__bfloat162float(dy[px * ldo + cc]) * __bfloat162float(y[px * ldo + cc]), scale[px * C + cc]);
        }
        const BF8 xv = ld8(x + px * ldi + c);
        bf16* o = dx + px * ldi + c;
        BF8 r;
        if (accum) r = ld8(o);
#pragma unroll
        for (int q = 0; q < 8; ++q) {
            const float s = ratio[q] + ratio[q + 1] + ratio[q + 2] + ratio[q + 3] + ratio[q + 4];
            float v = dyv[q] * __powf(sc[q], -beta) - cr * xv.get(q) * s;
            if (accum) v += r.get(q);
            r.set(q, v);
        }
        st8(o, r);
    }
}

// ------------------------------------------------------------------------------------------------ bf16 engine helpers
// im2col for the tensor-core GEMMs: col[px][(r*k + s)*Cin + c] = x[n][oh*stride - pad + r][ow*stride - pad + s][c] (0 outside),
// rows of ldk elements (Kc rounded up to 8: a TMA row pitch is a multiple of 16 bytes; the padding columns are zeroed).
// One thread per 16-byte group of a row; grid = (groups of an output row, Ho, N): no division for the pixel coordinates.
// VEC = 8: Cin % 8 == 0, whole 16-byte groups copied; VEC = 1: any Cin (the first convolution, Cin = 3), gathered per element.
template <int VEC, int K>
__global__ void im2col_kernel(const bf16* __restrict__ x, int ldi, bf16* __restrict__ col, int ldk, int H, int W, int Cin, int Wo,
                              int stride, int pad, unsigned m_per_row, unsigned m_cin) {
    pdl_wait();
    pdl_trigger();
    const unsigned per_row = (unsigned)ldk / 8u;
    const unsigned idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (unsigned)Wo * per_row) return;
    const int ow = (int)div_fast(idx, m_per_row);
    const int j = (int)(idx - (unsigned)ow * per_row) * 8;
    const int oh = blockIdx.y, n = blockIdx.z, Ho = gridDim.y;
    const int kc = K * K * Cin;
    const size_t m = ((size_t)n * Ho + oh) * Wo + ow;
    const bf16* xn = x + (size_t)n * H * W * ldi;
    uint4 v = make_uint4(0, 0, 0, 0);
    if (VEC == 8) {
        if (j < kc) {
            const int rs = (int)div_fast((unsigned)j, m_cin), c = j - rs * Cin, r = rs / K, s2 = rs - r * K;     // m_cin = magic(Cin)
            const int ih = oh * stride - pad + r, iw = ow * stride - pad + s2;
            if (ih >= 0 && ih < H && iw >= 0 && iw < W) v = *reinterpret_cast<const uint4*>(xn + ((size_t)ih * W + iw) * ldi + c);
        }
    } else {
        bf16* e = reinterpret_cast<bf16*>(&v);
#pragma unroll
        for (int q = 0; q < 8; ++q) {
            const int jj = j + q;
            if (jj < kc) {
                const int rs = (int)div_fast((unsigned)jj, m_cin), c = jj - rs * Cin, r = rs / K, s2 = rs - r * K;
                const int ih = oh * stride - pad + r, iw = ow * stride - pad + s2;
                if (ih >= 0 && ih < H && iw >= 0 && iw < W) e[q] = xn[((size_t)ih * W + iw) * ldi + c];
            }
        }
    }
    *reinterpret_cast<uint4*>(col + m * ldk + j) = v;
}

// im2col of a dense 3-channel image (conv1): block = one output row.  The K input rows it reads are staged in shared memory with
// zeroed borders (coalesced 16-byte loads; element (iw, c) of a row at lp + 3 * iw + c), so that a column row is K runs of 3K
// consecutive shared-memory elements and no thread gathers from global memory.  Writes the same bytes as im2col_kernel<1, K>.
template <int K>
__global__ void __launch_bounds__(256) im2col_c3_kernel(const bf16* __restrict__ x, bf16* __restrict__ col, int ldk, int H, int W, int Wo,
                                                         int stride, int pad, int lp, int pitch, unsigned m_per_row) {
    pdl_wait();
    pdl_trigger();
    extern __shared__ __align__(16) unsigned char im2col_c3_smem[];
    bf16* sm = reinterpret_cast<bf16*>(im2col_c3_smem);
    const int oh = blockIdx.x, n = blockIdx.y, Ho = gridDim.x;
    const uint4 z = make_uint4(0, 0, 0, 0);
    for (int i = threadIdx.x; i < K * (pitch / 8); i += 256) reinterpret_cast<uint4*>(sm)[i] = z;
    __syncthreads();
    const int wv = W * 3 / 8;
    for (int i = threadIdx.x; i < K * wv; i += 256) {
        const int r = i / wv, v = i - r * wv, ih = oh * stride - pad + r;
        if (ih >= 0 && ih < H)
            *reinterpret_cast<uint4*>(sm + r * pitch + lp + v * 8) = *reinterpret_cast<const uint4*>(x + (size_t)(n * H + ih) * W * 3 + v * 8);
    }
    __syncthreads();
    const unsigned per_row = (unsigned)ldk / 8u;
    const int kc = K * K * 3;
    bf16* out = col + (size_t)(n * Ho + oh) * Wo * ldk;
    for (unsigned idx = threadIdx.x; idx < (unsigned)Wo * per_row; idx += 256) {
        const int ow = (int)div_fast(idx, m_per_row), j = (int)(idx - (unsigned)ow * per_row) * 8;
        const bf16* src = sm + lp + (ow * stride - pad) * 3;
        uint4 v = z;
        bf16* e = reinterpret_cast<bf16*>(&v);
#pragma unroll
        for (int q = 0; q < 8; ++q) {
            const int jj = j + q;
            if (jj < kc) { const int r = jj / (3 * K); e[q] = src[r * pitch + (jj - r * 3 * K)]; }
        }
        *reinterpret_cast<uint4*>(out + (size_t)ow * ldk + j) = v;
    }
}

// data gradient of a stride-1 k x k convolution from the column gradient dcol = dY * W (the adjoint of im2col):
// dx[n][h][w][c] (+)= sum_{r,s} dcol[(n, h + pad - r, w + pad - s)][(r*k + s)*Cin + c]; 8 channels per thread, grid (., H, N).
template <int K>
__global__ void col2im_kernel(const bf16* __restrict__ dcol, int ldk, bf16* __restrict__ dx, int ldi, int W, int Cin, int pad, int accum,
                              unsigned m_cv) {
    pdl_wait();
    pdl_trigger();
    const unsigned cv = (unsigned)Cin / 8u;
    const unsigned idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (unsigned)W * cv) return;
    const int w = (int)div_fast(idx, m_cv), c = (int)(idx - (unsigned)w * cv) * 8;
    const int h = blockIdx.y, n = blockIdx.z, H = gridDim.y;
    float acc[8];
#pragma unroll
    for (int q = 0; q < 8; ++q) acc[q] = 0.f;
#pragma unroll
    for (int r = 0; r < K; ++r) {
        const int oh = h + pad - r;
        if (oh < 0 || oh >= H) continue;
#pragma unroll
        for (int s2 = 0; s2 < K; ++s2) {
            const int ow = w + pad - s2;
            if (ow < 0 || ow >= W) continue;
            const uint4 v = *reinterpret_cast<const uint4*>(dcol + (((size_t)n * H + oh) * W + ow) * ldk + (r * K + s2) * Cin + c);
            const bf16* e = reinterpret_cast<const bf16*>(&v);
#pragma unroll
            for (int q = 0; q < 8; ++q) acc[q] += __bfloat162float(e[q]);
        }
    }
    bf16* o = dx + (((size_t)n * H + h) * W + w) * ldi + c;
    uint4 ov;
    bf16* oe = reinterpret_cast<bf16*>(&ov);
    if (accum) {
        ov = *reinterpret_cast<const uint4*>(o);
#pragma unroll
        for (int q = 0; q < 8; ++q) acc[q] += __bfloat162float(oe[q]);
    }
#pragma unroll
    for (int q = 0; q < 8; ++q) oe[q] = __float2bfloat16_rn(acc[q]);
    *reinterpret_cast<uint4*>(o) = ov;
}

// ---- the two head convolutions (1x1, 1 and 4 output channels: too narrow for a TMA row): warp-level kernels on the fp32 master
// weights P[c][ldw] staged in shared memory.  NARROW_MAX output channels, Cin % 8 == 0.
constexpr int kNarrowMax = 8;

// y[px][co] = bias[co] + sum_c x[px][c] * P[c][co]: one warp per pixel, lanes stride over 8-channel groups
__global__ void __launch_bounds__(256) narrow_fwd_kernel(const bf16* __restrict__ x, int ldi, const float* __restrict__ P, int ldw,
                                                          const float* __restrict__ bias, bf16* __restrict__ y, int ldo, unsigned npix,
                                                          int Cin, int Cout) {
    extern __shared__ __align__(16) float ws[];            // [Cout][Cin]: a lane reads 8 consecutive floats per output channel
    for (int i = threadIdx.x; i < Cin * Cout; i += blockDim.x) ws[(i % Cout) * Cin + i / Cout] = P[(long)(i / Cout) * ldw + i % Cout];
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, wpb = blockDim.x >> 5;
    for (unsigned px = blockIdx.x * wpb + warp; px < npix; px += gridDim.x * wpb) {
        float acc[kNarrowMax];
#pragma unroll
        for (int o = 0; o < kNarrowMax; ++o) acc[o] = 0.f;
        for (int c = lane * 8; c < Cin; c += 256) {
            const BF8 v = ld8(x + (size_t)px * ldi + c);
#pragma unroll
            for (int o = 0; o < kNarrowMax; ++o)
                if (o < Cout) {
                    const float4 w0 = *reinterpret_cast<const float4*>(ws + o * Cin + c), w1 = *reinterpret_cast<const float4*>(ws + o * Cin + c + 4);
                    acc[o] = fmaf(v.get(0), w0.x, fmaf(v.get(1), w0.y, fmaf(v.get(2), w0.z, fmaf(v.get(3), w0.w, acc[o]))));
                    acc[o] = fmaf(v.get(4), w1.x, fmaf(v.get(5), w1.y, fmaf(v.get(6), w1.z, fmaf(v.get(7), w1.w, acc[o]))));
                }
        }
#pragma unroll
        for (int o = 0; o < kNarrowMax; ++o)
            if (o < Cout) {
#pragma unroll
                for (int d = 16; d > 0; d >>= 1) acc[o] += __shfl_xor_sync(0xffffffffu, acc[o], d);
                if (lane == 0) y[(size_t)px * ldo + o] = __float2bfloat16_rn(acc[o] + bias[o]);
            }
    }
}

// dx[px][c] (+)= sum_co dy[px][co] * P[c][co]: a thread per (pixel, 8 channels)
__global__ void __launch_bounds__(256) narrow_dgrad_kernel(const bf16* __restrict__ dy, int ldo, const float* __restrict__ P, int ldw,
                                                            bf16* __restrict__ dx, int ldi, unsigned npix, int Cin, int Cout, int accum) {
    extern __shared__ __align__(16) float ws[];            // [Cout][Cin]
    for (int i = threadIdx.x; i < Cin * Cout; i += blockDim.x) ws[(i % Cout) * Cin + i / Cout] = P[(long)(i / Cout) * ldw + i % Cout];
    __syncthreads();
    const unsigned cv = Cin / 8, total = npix * cv;
    for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        const unsigned px = i / cv;
        const int c = (int)(i % cv) * 8;
        float g[kNarrowMax];
#pragma unroll
        for (int o = 0; o < kNarrowMax; ++o) g[o] = o < Cout ? __bfloat162float(dy[(size_t)px * ldo + o]) : 0.f;
        bf16* op = dx + (size_t)px * ldi + c;
        BF8 r;
        float v[8];
        if (accum) r = ld8(op);
#pragma unroll
        for (int q = 0; q < 8; ++q) v[q] = accum ? r.get(q) : 0.f;
#pragma unroll
        for (int o = 0; o < kNarrowMax; ++o)
            if (o < Cout) {
                const float4 w0 = *reinterpret_cast<const float4*>(ws + o * Cin + c), w1 = *reinterpret_cast<const float4*>(ws + o * Cin + c + 4);
                v[0] = fmaf(g[o], w0.x, v[0]); v[1] = fmaf(g[o], w0.y, v[1]); v[2] = fmaf(g[o], w0.z, v[2]); v[3] = fmaf(g[o], w0.w, v[3]);
                v[4] = fmaf(g[o], w1.x, v[4]); v[5] = fmaf(g[o], w1.y, v[5]); v[6] = fmaf(g[o], w1.z, v[6]); v[7] = fmaf(g[o], w1.w, v[7]);
            }
#pragma unroll
        for (int q = 0; q < 8; ++q) r.set(q, v[q]);
        st8(op, r);
    }
}

// dP[c][co] += sum_px x[px][c] * dy[px][co]: a block per pixel chunk, a thread per 8-channel group
__global__ void __launch_bounds__(128) narrow_wgrad_kernel(const bf16* __restrict__ x, int ldi, const bf16* __restrict__ dy, int ldo,
                                                            float* __restrict__ G, int ldw, unsigned npix, unsigned chunk, int Cin, int Cout) {
    const unsigned p0 = blockIdx.x * chunk, p1 = min(p0 + chunk, npix);
    for (int c = threadIdx.x * 8; c < Cin; c += blockDim.x * 8) {
        float acc[8][4];
        for (int o0 = 0; o0 < Cout; o0 += 4) {
#pragma unroll
            for (int q = 0; q < 8; ++q)
#pragma unroll
                for (int o = 0; o < 4; ++o) acc[q][o] = 0.f;
            for (unsigned px = p0; px < p1; ++px) {
                const BF8 v = ld8(x + (size_t)px * ldi + c);
                float g[4];
#pragma unroll
                for (int o = 0; o < 4; ++o) g[o] = o0 + o < Cout ? __bfloat162float(dy[(size_t)px * ldo + o0 + o]) : 0.f;
#pragma unroll
                for (int q = 0; q < 8; ++q) {
                    const float xv = v.get(q);
#pragma unroll
                    for (int o = 0; o < 4; ++o) acc[q][o] = fmaf(xv, g[o], acc[q][o]);
                }
            }
#pragma unroll
            for (int q = 0; q < 8; ++q)
#pragma unroll
                for (int o = 0; o < 4; ++o)
                    if (o0 + o < Cout) atomicAdd(G + (long)(c + q) * ldw + o0 + o, acc[q][o]);
        }
    }
}

// bf16 copies of the parameters, refreshed before every forward: `wb` = the flat buffer cast as it is (packed [Kc][ldw]: the
// B operand of the data-gradient GEMM), `wf` = its transpose [Cout][ldk] per convolution (B operand of the forward GEMM).
__global__ void cast_flat_kernel(const float* __restrict__ in, bf16* __restrict__ out, long total) {
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x)
        out[i] = __float2bfloat16_rn(in[i]);
}
struct PackDesc { const float* P; bf16* wf; int ldw, ldk, Kc, Cout; };
// all convolutions in one launch: blockIdx.y = convolution
__global__ void pack_wf_kernel(const PackDesc* __restrict__ descs) {
    const PackDesc d = descs[blockIdx.y];
    const long total = (long)d.Cout * d.ldk;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        const int j = (int)(i % d.ldk), co = (int)(i / d.ldk);
        d.wf[i] = __float2bfloat16_rn(j < d.Kc ? d.P[(long)j * d.ldw + co] : 0.f);
    }
}

// ------------------------------------------------------------------------------------------------ data-parallel update over peer memory
// Reduce-scatter + ADAM + all-gather in ONE kernel over NVLink peer memory (the gradient buffers and the parameter buffers of
// all ranks of the node are mapped into every process through CUDA IPC): rank r owns the r-th shard of the flat buffers; it
// sums the shard of every rank's gradient buffer (peer loads, rank order: the same sum on every run), applies Caffe's
// ApplyUpdate to its shard (ADAM moments are kept only by the owner), and stores the updated parameters into every rank's
// parameter buffer (peer stores).  No intermediate buffer, no second pass over the gradients, no separate update launch.
// The caller orders the kernel between two stream-ordered barriers (all ranks finished their backward / all ranks wrote).
constexpr int kMaxPeers = 8;
struct PeerPtrs { const float* grads[kMaxPeers]; float* params[kMaxPeers]; };

__global__ void peer_reduce_adam_kernel(PeerPtrs pp, int world, int rank, float* __restrict__ own_grads, float* __restrict__ m,
                                        float* __restrict__ v, const float2* __restrict__ meta, long begin, long end, AdamArgs a) {
    const long n4 = (end - begin) / 4;
    for (long i4 = blockIdx.x * (long)blockDim.x + threadIdx.x; i4 < n4; i4 += (long)gridDim.x * blockDim.x) {
        const long i = begin + 4 * i4;
        float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int j = 0; j < world; ++j) {
            const float4 t = *reinterpret_cast<const float4*>(pp.grads[j] + i);
            g.x = __fadd_rn(g.x, t.x); g.y = __fadd_rn(g.y, t.y); g.z = __fadd_rn(g.z, t.z); g.w = __fadd_rn(g.w, t.w);
        }
        const float2 mt = meta[i >> 8];
        const float4 w4 = *reinterpret_cast<const float4*>(pp.params[rank] + i);
        float4 m4 = *reinterpret_cast<const float4*>(m + i), v4 = *reinterpret_cast<const float4*>(v + i);
        float gg[4] = {g.x, g.y, g.z, g.w}, ww[4] = {w4.x, w4.y, w4.z, w4.w}, mm[4] = {m4.x, m4.y, m4.z, m4.w}, vv[4] = {v4.x, v4.y, v4.z, v4.w};
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            float gi = __fmul_rn(gg[q], a.grad_scale);
            gi = __fadd_rn(gi, __fmul_rn(__fmul_rn(a.weight_decay, mt.y), ww[q]));
            mm[q] = __fadd_rn(__fmul_rn(a.b1, mm[q]), __fmul_rn(1.f - a.b1, gi));
            vv[q] = __fadd_rn(__fmul_rn(a.b2, vv[q]), __fmul_rn(1.f - a.b2, __fmul_rn(gi, gi)));
            const float step = __fmul_rn(__fmul_rn(__fmul_rn(a.rate, mt.x), a.correction), __fdiv_rn(mm[q], __fadd_rn(__fsqrt_rn(vv[q]), a.eps)));
            ww[q] = __fsub_rn(ww[q], step);
            gg[q] = gi;
        }
        *reinterpret_cast<float4*>(m + i) = make_float4(mm[0], mm[1], mm[2], mm[3]);
        *reinterpret_cast<float4*>(v + i) = make_float4(vv[0], vv[1], vv[2], vv[3]);
        *reinterpret_cast<float4*>(own_grads + i) = make_float4(gg[0], gg[1], gg[2], gg[3]);
        const float4 wn = make_float4(ww[0], ww[1], ww[2], ww[3]);
        for (int j = 0; j < world; ++j) *reinterpret_cast<float4*>(pp.params[j] + i) = wn;
    }
}

}  // namespace
}  // namespace fcnod

using namespace fcnod;

// =================================================================================================== the trainer
namespace {

struct TBuf { int C = 0, H = 0, W = 0; char* data = 0; char* diff = 0; };   // NHWC, float32 or bf16 (trainer precision)
struct TBlob {
    std::string name;
    int C = 0, H = 0, W = 0;
    int buf = -1, c_off = 0, ldc = 0;
    bool needs_grad = true;      // false for the network input and its Power transform
    bool diff_written = false;   // per backward pass; slices of a concat blob share the flag of the concat blob
    int root = -1;               // blob whose flag this one shares (-1: itself)
};
struct TConv {
    int Cout = 0, Cin = 0, k = 1, stride = 1, pad = 0, K = 0, ldw = 0, ldt = 0;
    size_t w_off = 0, b_off = 0;     // floats into the flat parameter buffer
    float* d_wt = 0;                 // SIMT dgrad packing [(r,s,co)][ldt], rebuilt every step
    // bf16 engine (tensor cores): transposed bf16 weights [Cout][ldk] (forward B operand), im2col row pitch, GEMM plans
    bool tc = false;
    bool narrow = false;             // bf16 engine: 1x1 convolution with fewer than 8 output channels (the heads): narrow_*_kernel
    bf16* d_wf = 0;
    bf16* d_col = 0;                 // k x k: this layer's im2col matrix, written by the forward and kept for the weight gradient
    int ldk = 0;
    int g_fwd = -1, g_dgrad = -1, g_wgrad = -1;
    float lr_mult[2], decay_mult[2];
    std::string name;
};
struct TLayer {
    LayerDef def;
    int in = -1, out = -1, conv = -1;
    bool relu = false, skip = false;
    uint8_t* idx = 0;      // Pooling: argmax of every output element
    float* scale = 0;      // LRN: the scale blob
    uint8_t* mask = 0;     // Dropout: keep mask (NHWC order)
};

struct GemmPlan {
    CUtensorMap ta, tb, to;
    GemmTcParams p;
    int grid = 0;
    size_t smem = 0;
    bool mn = false;
};

}  // namespace

struct fcnod_trainer {
    int device = 0, batch = 0, in_c = 3, in_h = 0, in_w = 0, num_classes = 1;
    int prec = 0;                             // 0: float32 storage, SIMT convolutions; 1: bf16 storage, tcgen05 GEMMs
    size_t esz = 4;
    bf16* wb = 0;                             // bf16 copy of the flat parameter buffer
    bf16* col = 0;                            // column-gradient scratch of the k x k data gradients (shared)
    PackDesc* pack_descs = 0;                 // one entry per tensor-core convolution (pack_wf_kernel)
    int n_pack = 0;
    // grouped weight gradients (wgrad_group_kernel): consecutive convolutions of the backward order, launched together when
    // the last member's dY is final
    struct WGroup { std::vector<int> convs; WgradProblem* probs = 0; WgradGroupParams gp; size_t smem = 0; int grid = 0; double flop = 0.0;
                    cudaEvent_t e0 = 0, e1 = 0; };
    std::vector<WGroup> wgroups;
    std::vector<int> conv_wgroup;             // per convolution: its group (-1: launched alone)
    // ReLU mask + bias gradient of all the convolutions that write one Concat blob in one launch over the blob, at the first
    // member the backward pass reaches (FCNOD_TRAIN_RB_FUSE)
    struct RbGroup { int root = -1; RbSegs segs; bool done = false; };
    std::vector<RbGroup> rb_groups;
    std::vector<int> conv_rbgroup;            // per convolution: its group (-1: its own launch)
    // grouped K-major GEMMs (kgemm_group_kernel): the 1x1 convolutions that read the same blob -- their forward GEMMs launch
    // together at the first member, their data gradients (all accumulating into that blob's diff) at the last one reached
    struct KGroup { std::vector<int> convs; KGemmProblem* probs = 0; KGemmGroupParams gp; size_t smem = 0; int grid = 0; double flop = 0.0;
                    int kind = 0; cudaEvent_t e0 = 0, e1 = 0; };
    std::vector<KGroup> kgroups;
    std::vector<int> conv_fgroup, conv_dgroup;
    // forward GEMMs that are not needed at once (the 3x3 / 5x5 / pool_proj branches of a module) wait, after their im2col, for the
    // first layer that reads one of their outputs (the Concat) and launch together there
    std::vector<int> conv_fdefer;             // per convolution: the group its forward GEMM waits in (-1: launched at once)
    std::vector<int> fwd_flush;               // per layer: group to launch BEFORE the layer runs (-1: none)
    std::vector<GemmPlan> gemms;
    // timing mode (bench.py's roofline): CUDA events around every GEMM launch of a step, on the launching stream
    bool timing = false;
    std::vector<int> gemm_kind;               // per plan: 0 forward, 1 data gradient, 2 weight gradient
    std::vector<cudaEvent_t> ev0, ev1;        // per plan
    fcnod_solver_params solver;
    NetDef def;
    std::vector<TBlob> blobs;
    std::map<std::string, int> blob_index;
    std::vector<TBuf> bufs;
    std::vector<TConv> convs;
    std::vector<TLayer> layers;
    size_t n_params = 0;                      // floats in each flat buffer (multiple of 256)
    float *params = 0, *grads = 0, *m = 0, *v = 0;
    float2* meta = 0;
    int cov_blob = -1, box_blob = -1, gh = 0, gw = 0;
    float *cov_nchw = 0, *box_nchw = 0, *dcov_nchw = 0, *dbox_nchw = 0, *losses_dev = 0;
    void* head_scratch = 0;
    uint8_t* mask_nchw = 0;
    float* stage = 0;                         // host-buffer step: device copies of the inputs
    int iter = 0;
    // data-parallel update over peer memory (fcnod_trainer_open_peers)
    int world = 1, rank = 0;
    float* peer_grads[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    float* peer_params[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    bool peers_open = false;
    cudaStream_t stream = 0;
    std::vector<void*> allocs;
};

namespace {

template <typename T> T* dalloc(fcnod_trainer* t, size_t n) {
    void* p = 0;
    FCNOD_CUDA(cudaMalloc(&p, std::max<size_t>(n, 1) * sizeof(T)));
    t->allocs.push_back(p);
    return static_cast<T*>(p);
}

int blob_of(fcnod_trainer* t, const std::string& n) {
    std::map<std::string, int>::iterator it = t->blob_index.find(n);
    FCNOD_CHECK(it != t->blob_index.end(), "trainer: unknown blob '" + n + "'");
    return it->second;
}
char* bdata(fcnod_trainer* t, const TBlob& b) { return t->bufs[b.buf].data + (size_t)b.c_off * t->esz; }
char* bdiff(fcnod_trainer* t, const TBlob& b) { return t->bufs[b.buf].diff + (size_t)b.c_off * t->esz; }
template <typename T> T* as(char* p) { return reinterpret_cast<T*>(p); }
TBlob& root_of(fcnod_trainer* t, int b) { return t->blobs[t->blobs[b].root >= 0 ? t->blobs[b].root : b]; }
// first writer of a diff overwrites, later writers accumulate (Caffe's Split layer)
int claim_diff(fcnod_trainer* t, int b) {
    TBlob& r = root_of(t, b);
    const int accum = r.diff_written ? 1 : 0;
    r.diff_written = true;
    return accum;
}

int caffe_pool_out_t(int h, int k, int s, int p) {
    int o = (int)std::ceil((h + 2 * p - k) / (double)s) + 1;
    if (p > 0 && (o - 1) * s >= h + p) --o;
    return o;
}

// the train-side layers of a train_val-form prototxt must be the head train_head.cu implements
void verify_train_side(const NetDef& def, int num_classes, float bbox_loss_weight) {
    for (size_t i = 0; i < def.train_side.size(); ++i) {
        const TrainSideLayer& L = def.train_side[i];
        if (L.type_name == "Slice") {
            const int c = num_classes;
            const int want[4] = {c, 5 * c, 9 * c, 13 * c};
            FCNOD_CHECK(L.slice_points.size() == 4 && L.top.size() == 5, "trainer: '" + L.name + "': expected 4 slice points / 5 tops");
            for (int k = 0; k < 4; ++k)
                FCNOD_CHECK(L.slice_points[k] == want[k], "trainer: '" + L.name + "': slice points do not match num_classes (expected c, 5c, 9c, 13c)");
        } else if (L.type_name == "Eltwise") {
            FCNOD_CHECK(L.eltwise_op == "PROD" && L.bottom.size() == 2, "trainer: '" + L.name + "': only two-input Eltwise PROD is part of the head");
        } else if (L.type_name == "L1Loss") {
            FCNOD_CHECK(std::fabs(L.loss_weight - bbox_loss_weight) < 1e-6f, "trainer: '" + L.name + "': loss_weight differs from the solver parameters' bbox_loss_weight");
        } else if (L.type_name == "EuclideanLoss") {
            FCNOD_CHECK(std::fabs(L.loss_weight - 1.f) < 1e-6f, "trainer: '" + L.name + "': loss_weight != 1 is not supported");
        }
    }
}

void build_tc(fcnod_trainer* t);

void build_trainer(fcnod_trainer* t, const WeightMap& weights) {
    const NetDef& def = t->def;
    const int N = t->batch;
    // ---- shapes, concat aliasing
    std::map<std::string, std::pair<std::string, int> > slice_of;   // branch top -> (concat top, channel offset)
    std::map<std::string, int> chans, hs, ws;
    chans[def.input_name] = t->in_c; hs[def.input_name] = t->in_h; ws[def.input_name] = t->in_w;
    for (size_t i = 0; i < def.layers.size(); ++i) {
        const LayerDef& L = def.layers[i];
        FCNOD_CHECK(!L.bottom.empty() && chans.count(L.bottom[0]), "trainer: layer '" + L.name + "' reads an undefined blob");
        const int C = chans[L.bottom[0]], H = hs[L.bottom[0]], W = ws[L.bottom[0]];
        int oc = C, oh = H, ow = W;
        if (L.type == L_CONV) {
            oc = L.num_output;
            oh = (H + 2 * L.pad - L.kernel) / L.stride + 1;
            ow = (W + 2 * L.pad - L.kernel) / L.stride + 1;
        } else if (L.type == L_POOL) {
            oh = caffe_pool_out_t(H, L.kernel, L.stride, L.pad);
            ow = caffe_pool_out_t(W, L.kernel, L.stride, L.pad);
        } else if (L.type == L_CONCAT) {
            oc = 0;
            for (size_t b = 0; b < L.bottom.size(); ++b) {
                FCNOD_CHECK(chans.count(L.bottom[b]) && hs[L.bottom[b]] == H && ws[L.bottom[b]] == W, "trainer: concat '" + L.name + "' shape mismatch");
                slice_of[L.bottom[b]] = std::make_pair(L.top[0], oc);
                oc += chans[L.bottom[b]];
            }
        }
        chans[L.top[0]] = oc; hs[L.top[0]] = oh; ws[L.top[0]] = ow;
    }
    // ---- blobs + buffers
    auto add_blob = [&](const std::string& name) -> int {
        if (t->blob_index.count(name)) return t->blob_index[name];
        TBlob b;
        b.name = name; b.C = chans[name]; b.H = hs[name]; b.W = ws[name];
        std::map<std::string, std::pair<std::string, int> >::iterator so = slice_of.find(name);
        if (so != slice_of.end()) {
            // the concat top's buffer is created on first sight of any of its branches
            const std::string& top = so->second.first;
            int rb;
            if (t->blob_index.count(top)) rb = t->blob_index[top];
            else {
                TBlob r;
                r.name = top; r.C = chans[top]; r.H = hs[top]; r.W = ws[top];
                TBuf bf; bf.C = r.C; bf.H = r.H; bf.W = r.W;
                t->bufs.push_back(bf);
                r.buf = (int)t->bufs.size() - 1; r.ldc = r.C;
                t->blobs.push_back(r);
                rb = (int)t->blobs.size() - 1;
                t->blob_index[top] = rb;
            }
            b.buf = t->blobs[rb].buf; b.c_off = so->second.second; b.ldc = t->blobs[rb].C; b.root = rb;
        } else {
            TBuf bf; bf.C = b.C; bf.H = b.H; bf.W = b.W;
            t->bufs.push_back(bf);
            b.buf = (int)t->bufs.size() - 1; b.ldc = b.C;
        }
        t->blobs.push_back(b);
        t->blob_index[name] = (int)t->blobs.size() - 1;
        return (int)t->blobs.size() - 1;
    };
    add_blob(def.input_name);
    t->blobs[0].needs_grad = false;
    size_t off = 0;
    std::vector<float2> meta;
    for (size_t i = 0; i < def.layers.size(); ++i) {
        const LayerDef& L = def.layers[i];
        TLayer tl;
        tl.def = L;
        tl.in = blob_of(t, L.bottom[0]);
        const bool in_place = L.top[0] == L.bottom[0];
        tl.out = in_place ? tl.in : add_blob(L.top[0]);
        if (L.type == L_POWER) {
            FCNOD_CHECK(L.power == 1.f && tl.in == 0, "trainer: only the input Power layer (power 1) is supported");
            t->blobs[tl.out].needs_grad = false;
        } else if (L.type == L_RELU) {
            FCNOD_CHECK(in_place && !t->layers.empty() && t->layers.back().def.type == L_CONV && t->layers.back().out == tl.in,
                        "trainer: ReLU '" + L.name + "' must run in place on the convolution before it");
            t->layers.back().relu = true;
            tl.skip = true;
        } else if (L.type == L_CONV) {
            FCNOD_CHECK(!in_place, "trainer: in-place convolution");
            TConv c;
            c.name = L.name;
            c.Cout = L.num_output; c.Cin = t->blobs[tl.in].C; c.k = L.kernel; c.stride = L.stride; c.pad = L.pad;
            c.K = c.k * c.k * c.Cin; c.ldw = (c.Cout + 63) / 64 * 64; c.ldt = (c.Cin + 63) / 64 * 64;
            for (int q = 0; q < 2; ++q) { c.lr_mult[q] = L.lr_mult[q]; c.decay_mult[q] = L.decay_mult[q]; }
            FCNOD_CHECK(L.bias_term, "trainer: convolution '" + L.name + "' without a bias term is not supported");
            const size_t wn = ((size_t)c.K * c.ldw + 255) / 256 * 256, bn = ((size_t)c.ldw + 255) / 256 * 256;
            c.w_off = off; off += wn;
            c.b_off = off; off += bn;
            for (size_t q = 0; q < wn / 256; ++q) meta.push_back(make_float2(c.lr_mult[0], c.decay_mult[0]));
            for (size_t q = 0; q < bn / 256; ++q) meta.push_back(make_float2(c.lr_mult[1], c.decay_mult[1]));
            if (t->blobs[tl.in].needs_grad) {
                FCNOD_CHECK(c.stride == 1 && 2 * c.pad == c.k - 1, "trainer: data gradient of '" + L.name + "' needs a stride-1 'same' convolution");
                c.d_wt = dalloc<float>(t, (size_t)c.k * c.k * c.Cout * c.ldt);
            }
            t->convs.push_back(c);
            tl.conv = (int)t->convs.size() - 1;
        } else if (L.type == L_POOL) {
            FCNOD_CHECK(L.kernel * L.kernel <= 255, "trainer: pooling window too large");
            tl.idx = dalloc<uint8_t>(t, (size_t)N * t->blobs[tl.out].H * t->blobs[tl.out].W * t->blobs[tl.out].C);
        } else if (L.type == L_LRN) {
            tl.scale = dalloc<float>(t, (size_t)N * t->blobs[tl.out].H * t->blobs[tl.out].W * t->blobs[tl.out].C);
        } else if (L.type == L_DROPOUT) {
            FCNOD_CHECK(t->blobs[tl.in].ldc == t->blobs[tl.in].C && t->blobs[tl.out].ldc == t->blobs[tl.out].C, "trainer: dropout on a sliced blob");
            FCNOD_CHECK(!t->mask_nchw, "trainer: more than one Dropout layer");
            const size_t n = (size_t)N * t->blobs[tl.out].H * t->blobs[tl.out].W * t->blobs[tl.out].C;
            tl.mask = dalloc<uint8_t>(t, n);
            t->mask_nchw = dalloc<uint8_t>(t, n);
        } else if (L.type == L_CONCAT) {
            tl.skip = true;
        } else if (L.type == L_SIGMOID) {
            FCNOD_CHECK(t->blobs[tl.in].ldc == t->blobs[tl.in].C, "trainer: sigmoid on a sliced blob");
        }
        t->layers.push_back(tl);
    }
    // ---- activations
    for (size_t i = 0; i < t->bufs.size(); ++i) {
        TBuf& b = t->bufs[i];
        const size_t n = (size_t)N * b.C * b.H * b.W;
        b.data = dalloc<char>(t, n * t->esz);
        b.diff = dalloc<char>(t, n * t->esz);
    }
    // ---- parameters
    t->n_params = off;
    t->params = dalloc<float>(t, off); t->grads = dalloc<float>(t, off); t->m = dalloc<float>(t, off); t->v = dalloc<float>(t, off);
    t->meta = dalloc<float2>(t, meta.size());
    FCNOD_CUDA(cudaMemcpy(t->meta, meta.data(), meta.size() * sizeof(float2), cudaMemcpyHostToDevice));
    FCNOD_CUDA(cudaMemset(t->grads, 0, off * 4)); FCNOD_CUDA(cudaMemset(t->m, 0, off * 4)); FCNOD_CUDA(cudaMemset(t->v, 0, off * 4));
    std::vector<float> host(off, 0.f);
    for (size_t ci = 0; ci < t->convs.size(); ++ci) {
        const TConv& c = t->convs[ci];
        WeightMap::const_iterator it = weights.find(c.name);
        FCNOD_CHECK(it != weights.end() && it->second.size() >= 2, "trainer: the snapshot has no weights for '" + c.name + "'");
        const HostBlob& w = it->second[0];
        const HostBlob& b = it->second[1];
        FCNOD_CHECK(w.data.size() == (size_t)c.Cout * c.K && b.data.size() == (size_t)c.Cout, "trainer: weight shape mismatch for '" + c.name + "'");
        for (int co = 0; co < c.Cout; ++co)
            for (int cc = 0; cc < c.Cin; ++cc)
                for (int rs = 0; rs < c.k * c.k; ++rs)
                    host[c.w_off + ((size_t)rs * c.Cin + cc) * c.ldw + co] = w.data[((size_t)co * c.Cin + cc) * c.k * c.k + rs];
        for (int co = 0; co < c.Cout; ++co) host[c.b_off + co] = b.data[co];
    }
    FCNOD_CUDA(cudaMemcpy(t->params, host.data(), off * 4, cudaMemcpyHostToDevice));
    // ---- head
    t->cov_blob = blob_of(t, "coverage"); t->box_blob = blob_of(t, "bboxes");
    const TBlob& cb = t->blobs[t->cov_blob];
    const TBlob& bb = t->blobs[t->box_blob];
    FCNOD_CHECK(bb.C == 4 * cb.C && bb.H == cb.H && bb.W == cb.W, "trainer: `bboxes` must have 4 channels per `coverage` channel");
    t->num_classes = cb.C; t->gh = cb.H; t->gw = cb.W;
    const size_t n1 = (size_t)N * cb.C * cb.H * cb.W;
    t->cov_nchw = dalloc<float>(t, n1); t->dcov_nchw = dalloc<float>(t, n1);
    t->box_nchw = dalloc<float>(t, 4 * n1); t->dbox_nchw = dalloc<float>(t, 4 * n1);
    t->losses_dev = dalloc<float>(t, 2);
    t->head_scratch = dalloc<char>(t, fcnod_train_head_scratch_bytes());
    if (t->prec == 1) build_tc(t);
}

// Launch with programmatic dependent launch (PDL): the grid may be scheduled while the previous kernel of the stream drains
// (its prologue -- barrier init, TMEM allocation, index set-up -- overlaps that tail); every kernel launched this way executes
// griddepcontrol.wait before its first global-memory access, which returns when the previous kernel has completed and flushed.
// Opt-in (FCNOD_TRAIN_PDL=1): measured neutral on the training step (10.23 ms with, 10.12 ms without: the gaps between the 373
// launches of a step add up to 0.3 ms only); without it the same kernels launch fully serialised and the wait is a no-op.
inline bool pdl_enabled() {
    static std::atomic<int> on(-1);
    if (on < 0) { const char* e = getenv("FCNOD_TRAIN_PDL"); on = (e && e[0] == '1') ? 1 : 0; }
    return on != 0;
}
template <typename... KArgs, typename... Args>
void launch_k(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, Args&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = pdl_enabled() ? 1 : 0;
    FCNOD_CUDA(cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...));
}

// ------------------------------------------------------------------------------------------------ tensor-core GEMM plans
typedef CUresult (*EncodeTiledFnT)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                   const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                   CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFnT encode_tiled_fn() {
    static EncodeTiledFnT fn = 0;
    if (!fn) {
        void* p = 0;
        cudaDriverEntryPointQueryResult qres;
        FCNOD_CUDA(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres));
        FCNOD_CHECK(p && qres == cudaDriverEntryPointSuccess, "cuTensorMapEncodeTiled is not available in this driver");
        fn = (EncodeTiledFnT)p;
    }
    return fn;
}

// 2-d bf16 tensor map: `inner` contiguous elements per row, `rows` rows `pitch` elements apart, box {64, box_rows}, SWIZZLE_128B
void encode_2d(CUtensorMap* m, const void* base, long inner, long rows, long pitch, int box_rows, const char* what) {
    FCNOD_CHECK((reinterpret_cast<uintptr_t>(base) & 15) == 0 && (pitch * 2) % 16 == 0, std::string("trainer: TMA operand not 16-byte aligned: ") + what);
    cuuint64_t dims[2] = {(cuuint64_t)inner, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)pitch * 2};
    cuuint32_t box[2] = {64u, (cuuint32_t)box_rows};
    cuuint32_t es[2] = {1, 1};
    CUresult r = encode_tiled_fn()(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, es,
                                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    FCNOD_CHECK(r == CUDA_SUCCESS, std::string("cuTensorMapEncodeTiled failed for ") + what + " (" + std::to_string((int)r) + ")");
}

int pow2_at_least(int v) { int p = 32; while (p < v) p <<= 1; return p; }

void finish_plan(GemmPlan& g) {
    GemmTcParams& p = g.p;
    const size_t stage = (size_t)p.a_stage_bytes + p.b_stage_bytes;
    p.out_stage_bytes = p.tma_out ? (uint32_t)((p.BN + 63) / 64) * 16384u : 0u;
    p.stages = (int)std::max<size_t>(2, std::min<size_t>(8, (192 * 1024 - p.out_stage_bytes) / stage));
    p.tmem_cols = (uint32_t)pow2_at_least(2 * p.BN);
    g.smem = (size_t)p.stages * stage + p.out_stage_bytes + 16 * p.stages + 64 + 1024 + kGemmBiasMax * 4;
    FCNOD_CHECK(!p.bias || p.N <= 1024, "gemm: a bias needs N <= 1024");
    g.grid = std::min(p.tiles_m * p.tiles_n * p.splits, 148);
}

// D[M][N] = A[M][K] * B[N][K]^T (+ bias, ReLU, accumulate), all row-major with the given leading dimensions
GemmPlan plan_gemm_k(const bf16* A, int lda, long M, int K, const bf16* B, int ldb, int N, void* out, int ldo, int out_f32,
                     const float* bias, int relu, const char* what) {
    GemmPlan g;
    memset(&g.p, 0, sizeof g.p);
    GemmTcParams& p = g.p;
    p.M = (int)M; p.N = N; p.K = K;
    const int nt = (N + 255) / 256;
    p.BN = std::min(256, (((N + nt - 1) / nt) + 15) / 16 * 16);
    p.tiles_m = (int)((M + 127) / 128); p.tiles_n = (N + p.BN - 1) / p.BN; p.splits = 1;
    p.k_blocks = (K + 63) / 64; p.kb_per_split = p.k_blocks;
    p.a_stage_bytes = 128 * 128; p.b_stage_bytes = (uint32_t)p.BN * 128;
    p.out = out; p.ldo = ldo; p.out_f32 = out_f32; p.bias = bias; p.relu = relu; p.accum = 0;
    encode_2d(&g.ta, A, K, M, lda, 128, what);
    encode_2d(&g.tb, B, K, N, ldb, p.BN, what);
    // bf16 outputs with 16-byte rows leave through shared memory and the TMA unit (store, or reduce-add when accumulating)
    p.tma_out = (!out_f32 && ldo % 8 == 0 && (reinterpret_cast<uintptr_t>(out) & 15) == 0 && getenv("FCNOD_GEMM_NO_TMA_STORE") == 0) ? 1 : 0;
    if (p.tma_out && p.tiles_n > 1) {
        // the staged tile leaves in 64-column regions: with several N tiles a tile must END on a region boundary, or the tail
        // of its last region would land on the next tile's columns (a single tile's tail is clipped at N by the TMA unit)
        p.BN = std::min(256, (((N + nt - 1) / nt) + 63) / 64 * 64);
        p.tiles_n = (N + p.BN - 1) / p.BN;
        p.b_stage_bytes = (uint32_t)p.BN * 128;
        encode_2d(&g.tb, B, K, N, ldb, p.BN, what);
    }
    if (p.tma_out) encode_2d(&g.to, out, N, M, ldo, 128, what);
    else g.to = g.ta;
    g.mn = false;
    finish_plan(g);
    return g;
}

// G[n][m] += sum_px A[px][m] * B[px][n]  (m < Mco, n < Ncols): the weight gradient, reduction over pixels, split over CTAs
GemmPlan plan_gemm_mn(const bf16* A, int lda, int Mco, const bf16* B, int ldb, int Ncols, long PX, float* gout, int ldg,
                      const char* what) {
    GemmPlan g;
    memset(&g.p, 0, sizeof g.p);
    GemmTcParams& p = g.p;
    p.M = Mco; p.N = Ncols; p.K = (int)PX;
    const int nt = (Ncols + 255) / 256;
    p.BN = std::min(256, (((Ncols + nt - 1) / nt) + 63) / 64 * 64);
    p.tiles_m = (Mco + 127) / 128; p.tiles_n = (Ncols + p.BN - 1) / p.BN;
    p.k_blocks = (int)((PX + kGemmPK - 1) / kGemmPK);
    int splits = std::max(1, std::min(p.k_blocks, (148 * 2 + p.tiles_m * p.tiles_n - 1) / (p.tiles_m * p.tiles_n)));
    p.kb_per_split = (p.k_blocks + splits - 1) / splits;
    p.splits = (p.k_blocks + p.kb_per_split - 1) / p.kb_per_split;
    p.a_stage_bytes = 2 * kGemmPK * 128; p.b_stage_bytes = (uint32_t)(p.BN / 64) * kGemmPK * 128;
    p.gout = gout; p.ldg = ldg;
    encode_2d(&g.ta, A, Mco, PX, lda, kGemmPK, what);
    encode_2d(&g.tb, B, Ncols, PX, ldb, kGemmPK, what);
    g.to = g.ta;
    g.mn = true;
    finish_plan(g);
    return g;
}

int add_plan(fcnod_trainer* t, const GemmPlan& g, int kind = 0) {
    t->gemms.push_back(g);
    t->gemm_kind.push_back(kind);
    return (int)t->gemms.size() - 1;
}

void run_plan(const GemmPlan& g, int accum, cudaStream_t s) {
    static std::atomic<bool> attr[64];       // per device (function attributes belong to a context); a race writes the same value twice
    int dev = 0;
    FCNOD_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !attr[dev]) {
        FCNOD_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        FCNOD_CUDA(cudaFuncSetAttribute(gemm_tc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        if (dev >= 0 && dev < 64) attr[dev] = true;
    }
    GemmTcParams p = g.p;
    p.accum = accum;
    if (g.mn) launch_k(gemm_tc_kernel<true>, g.grid, kGemmThreads, g.smem, s, g.ta, g.tb, g.to, p);
    else launch_k(gemm_tc_kernel<false>, g.grid, kGemmThreads, g.smem, s, g.ta, g.tb, g.to, p);
    FCNOD_CUDA(cudaGetLastError());
}
void run_gemm(fcnod_trainer* t, int id, int accum, cudaStream_t s) {
    if (t->timing) {
        if (t->ev0.size() < t->gemms.size()) {
            const size_t old = t->ev0.size();
            t->ev0.resize(t->gemms.size()); t->ev1.resize(t->gemms.size());
            for (size_t i = old; i < t->gemms.size(); ++i) { FCNOD_CUDA(cudaEventCreate(&t->ev0[i])); FCNOD_CUDA(cudaEventCreate(&t->ev1[i])); }
        }
        FCNOD_CUDA(cudaEventRecord(t->ev0[id], s));
        run_plan(t->gemms[id], accum, s);
        FCNOD_CUDA(cudaEventRecord(t->ev1[id], s));
    } else {
        run_plan(t->gemms[id], accum, s);
    }
}

// bf16 engine: scratch, bf16 weights and the three GEMM plans of every convolution (pointers are fixed for the trainer's life)
void build_kgroups(fcnod_trainer* t);
void build_fwd_deferral(fcnod_trainer* t);
void build_rbgroups(fcnod_trainer* t);
const bool kRbFuseDefault = true;        // measured 7.72 -> 7.56 ms per batch-32 step; FCNOD_TRAIN_RB_FUSE=0 reverts

void build_tc(fcnod_trainer* t) {
    const int N = t->batch;
    size_t col_elems = 8;
    for (size_t li = 0; li < t->layers.size(); ++li) {
        TLayer& L = t->layers[li];
        if (L.def.type != L_CONV) continue;
        TConv& c = t->convs[L.conv];
        const TBlob& ib = t->blobs[L.in];
        const TBlob& ob = t->blobs[L.out];
        // tensor cores need 16-byte rows on both sides of the output view; the 1- and 4-channel heads stay on the SIMT kernels
        c.tc = ob.ldc % 8 == 0 && ob.c_off % 8 == 0 && c.Cout % 8 == 0;
        c.narrow = !c.tc && !L.relu && c.k == 1 && c.stride == 1 && c.Cout <= kNarrowMax && c.Cin % 8 == 0 && ib.ldc % 8 == 0 && ib.c_off % 8 == 0 &&
                   (size_t)c.Cin * c.Cout * 4 <= 48 * 1024;
        if (!c.tc) continue;
        c.ldk = (c.K + 7) / 8 * 8;
        if (!(c.k == 1 && c.stride == 1)) col_elems = std::max(col_elems, (size_t)N * ob.H * ob.W * c.ldk);
    }
    FCNOD_CHECK(col_elems < (1ull << 32), "trainer: batch too large for the 32-bit index arithmetic of the bf16 kernels");
    for (size_t b = 0; b < t->bufs.size(); ++b)
        FCNOD_CHECK((size_t)N * t->bufs[b].C * t->bufs[b].H * t->bufs[b].W < (1ull << 31), "trainer: batch too large for the 32-bit index arithmetic of the bf16 kernels");
    t->col = dalloc<bf16>(t, col_elems);
    t->wb = dalloc<bf16>(t, t->n_params);
    for (size_t li = 0; li < t->layers.size(); ++li) {
        TLayer& L = t->layers[li];
        if (L.def.type != L_CONV) continue;
        TConv& c = t->convs[L.conv];
        if (!c.tc) continue;
        const TBlob& ib = t->blobs[L.in];
        const TBlob& ob = t->blobs[L.out];
        const long PX = (long)N * ob.H * ob.W;
        const bool direct = c.k == 1 && c.stride == 1;
        if (direct) FCNOD_CHECK(ib.ldc % 8 == 0 && ib.c_off % 8 == 0, "trainer: input view of '" + c.name + "' is not 16-byte aligned");
        c.d_wf = dalloc<bf16>(t, (size_t)c.Cout * c.ldk);
        if (!direct) c.d_col = dalloc<bf16>(t, (size_t)PX * c.ldk);
        const bf16* X = direct ? as<bf16>(bdata(t, ib)) : c.d_col;
        const int ldx = direct ? ib.ldc : c.ldk;
        c.g_fwd = add_plan(t, plan_gemm_k(X, ldx, PX, c.K, c.d_wf, c.ldk, c.Cout, bdata(t, ob), ob.ldc, 0, t->params + c.b_off, L.relu ? 1 : 0, c.name.c_str()));
        c.g_wgrad = add_plan(t, plan_gemm_mn(as<bf16>(bdiff(t, ob)), ob.ldc, c.Cout, X, ldx, c.K, PX, t->grads + c.w_off, c.ldw, c.name.c_str()), 2);
        if (ib.needs_grad) {
            // dY[px][Cout] * Wb[Kc][Cout]^T: straight into the bottom diff for 1x1, into the column scratch (then col2im) otherwise
            FCNOD_CHECK(c.Cin % 8 == 0, "trainer: data gradient of '" + c.name + "' needs Cin % 8 == 0");
            c.g_dgrad = add_plan(t, plan_gemm_k(as<bf16>(bdiff(t, ob)), ob.ldc, PX, c.Cout, t->wb + c.w_off, c.ldw, c.K,
                                    direct ? (void*)bdiff(t, ib) : (void*)t->col, direct ? ib.ldc : c.ldk, 0, 0, 0, c.name.c_str()), 1);
        }
    }
    std::vector<PackDesc> descs;
    for (size_t ci = 0; ci < t->convs.size(); ++ci) {
        const TConv& c = t->convs[ci];
        if (!c.tc) continue;
        PackDesc d = {t->params + c.w_off, c.d_wf, c.ldw, c.ldk, c.K, c.Cout};
        descs.push_back(d);
    }
    t->n_pack = (int)descs.size();
    t->pack_descs = dalloc<PackDesc>(t, descs.size());
    FCNOD_CUDA(cudaMemcpy(t->pack_descs, descs.data(), descs.size() * sizeof(PackDesc), cudaMemcpyHostToDevice));
    // ---- weight-gradient groups: up to G consecutive tensor-core convolutions in backward order (measured: 57 single launches
    //      1.71 ms, groups of 8 1.05 ms, groups of 16 0.88 ms)
    t->conv_wgroup.assign(t->convs.size(), -1);
    int G = 16;
    if (const char* e = getenv("FCNOD_TRAIN_WGRAD_GROUP")) G = atoi(e);
    if (G >= 2) {
        std::vector<int> order;                  // tensor-core convolutions in the order the backward pass reaches them
        for (size_t li = t->layers.size(); li-- > 0;)
            if (t->layers[li].def.type == L_CONV && t->convs[t->layers[li].conv].tc) order.push_back(t->layers[li].conv);
        for (size_t i0 = 0; i0 < order.size(); i0 += G) {
            fcnod_trainer::WGroup grp;
            for (size_t i = i0; i < std::min(order.size(), i0 + (size_t)G); ++i) grp.convs.push_back(order[i]);
            // pixel splits: about 4 work items per SM over the whole group, at least 4 reduction blocks per item
            double blocks = 0;
            for (size_t i = 0; i < grp.convs.size(); ++i) {
                const GemmTcParams& p = t->gemms[t->convs[grp.convs[i]].g_wgrad].p;
                blocks += (double)p.tiles_m * p.tiles_n * p.k_blocks;
            }
            const int target_kb = std::max(4, (int)std::ceil(blocks / (148.0 * 4)));
            std::vector<WgradProblem> probs(grp.convs.size());
            int work = 0, bn_max = 64;
            for (size_t i = 0; i < grp.convs.size(); ++i) {
                const GemmPlan& g = t->gemms[t->convs[grp.convs[i]].g_wgrad];
                const GemmTcParams& p = g.p;
                WgradProblem& q = probs[i];
                memset(&q, 0, sizeof q);
                q.ta = g.ta; q.tb = g.tb;
                q.M = p.M; q.N = p.N; q.BN = p.BN; q.tiles_m = p.tiles_m; q.tiles_n = p.tiles_n; q.k_blocks = p.k_blocks;
                q.splits = std::max(1, (p.k_blocks + target_kb - 1) / target_kb);
                q.kb_per_split = (p.k_blocks + q.splits - 1) / q.splits;
                q.splits = (p.k_blocks + q.kb_per_split - 1) / q.kb_per_split;
                q.work_begin = work; q.ldg = p.ldg; q.gout = p.gout;
                work += q.tiles_m * q.tiles_n * q.splits;
                bn_max = std::max(bn_max, p.BN);
                grp.flop += 2.0 * p.M * p.N * (double)p.K;
                t->conv_wgroup[grp.convs[i]] = (int)t->wgroups.size();
            }
            grp.gp.n_problems = (int)probs.size(); grp.gp.n_work = work;
            grp.gp.b_stage_bytes = (uint32_t)(bn_max / 64) * kGemmPK * 128; grp.gp.bn_max = (uint32_t)bn_max;
            const size_t stage = 2 * kGemmPK * 128 + grp.gp.b_stage_bytes;
            grp.gp.stages = (int)std::max<size_t>(2, std::min<size_t>(8, (200 * 1024) / stage));
            grp.gp.tmem_cols = (uint32_t)pow2_at_least(2 * bn_max);
            grp.smem = (size_t)grp.gp.stages * stage + 16 * grp.gp.stages + 64 + 1024;
            grp.grid = std::min(work, 148);
            grp.probs = dalloc<WgradProblem>(t, probs.size());
            FCNOD_CHECK((reinterpret_cast<uintptr_t>(grp.probs) & 127) == 0, "trainer: problem table not 128-byte aligned");
            FCNOD_CUDA(cudaMemcpy(grp.probs, probs.data(), probs.size() * sizeof(WgradProblem), cudaMemcpyHostToDevice));
            t->wgroups.push_back(grp);
        }
    }
    build_kgroups(t);
    build_fwd_deferral(t);
    build_rbgroups(t);
}

// Concat blobs all of whose slices are written by tensor-core convolutions with a ReLU: one ReLU / bias-gradient launch per blob
void build_rbgroups(fcnod_trainer* t) {
    t->conv_rbgroup.assign(t->convs.size(), -1);
    const char* e = getenv("FCNOD_TRAIN_RB_FUSE");
    if (!(e ? e[0] != '0' : kRbFuseDefault)) return;
    std::map<int, std::vector<int> > by_root;          // concat blob -> layers writing its slices
    for (size_t li = 0; li < t->layers.size(); ++li) {
        const TLayer& L = t->layers[li];
        if (L.skip || L.out < 0) continue;
        const TBlob& ob = t->blobs[L.out];
        if (ob.root >= 0 && ob.root != L.out) by_root[ob.root].push_back((int)li);
    }
    for (std::map<int, std::vector<int> >::iterator it = by_root.begin(); it != by_root.end(); ++it) {
        const TBlob& rb = t->blobs[it->first];
        std::vector<int>& ls = it->second;
        if (ls.size() < 2 || ls.size() > 8 || rb.C % 8 != 0 || rb.ldc != rb.C || rb.c_off != 0) continue;
        std::sort(ls.begin(), ls.end(), [&](int a, int b) { return t->blobs[t->layers[a].out].c_off < t->blobs[t->layers[b].out].c_off; });
        bool ok = true;
        int next = 0;
        fcnod_trainer::RbGroup g;
        g.root = it->first;
        memset(&g.segs, 0, sizeof g.segs);
        for (size_t i = 0; i < ls.size() && ok; ++i) {
            const TLayer& L = t->layers[ls[i]];
            const TBlob& ob = t->blobs[L.out];
            ok = L.def.type == L_CONV && L.conv >= 0 && L.relu && t->convs[L.conv].tc && ob.c_off == next && ob.C == t->convs[L.conv].Cout &&
                 ob.C % 8 == 0 && ob.ldc == rb.C && ob.buf == rb.buf;
            if (!ok) break;
            g.segs.s[i].c0 = ob.c_off;
            g.segs.s[i].db = t->grads + t->convs[L.conv].b_off;
            next += ob.C;
        }
        if (!ok || next != rb.C) continue;
        g.segs.n = (int)ls.size();
        for (size_t i = 0; i < ls.size(); ++i) t->conv_rbgroup[t->layers[ls[i]].conv] = (int)t->rb_groups.size();
        t->rb_groups.push_back(g);
    }
}

int build_kgroup(fcnod_trainer* t, const std::vector<int>& convs, bool dgrad) {
    fcnod_trainer::KGroup grp;
    grp.convs = convs; grp.kind = dgrad ? 1 : 0;
    std::vector<KGemmProblem> probs(convs.size());
    int work = 0, bn_max = 16, bias = 0;
    for (size_t i = 0; i < convs.size(); ++i) {
        const TConv& c = t->convs[convs[i]];
        const GemmPlan& g = t->gemms[dgrad ? c.g_dgrad : c.g_fwd];
        const GemmTcParams& p = g.p;
        if (!p.tma_out) return -1;
        KGemmProblem& q = probs[i];
        memset(&q, 0, sizeof q);
        q.ta = g.ta; q.tb = g.tb; q.to = g.to;
        q.M = p.M; q.N = p.N; q.BN = p.BN; q.tiles_m = p.tiles_m; q.tiles_n = p.tiles_n; q.k_blocks = p.k_blocks;
        q.work_begin = work; q.relu = p.relu; q.accum = dgrad ? 1 : 0; q.bias = p.bias;
        q.bias_off = p.bias ? bias : -1;
        if (p.bias) bias += p.tiles_n * p.BN;
        work += p.tiles_m * p.tiles_n;
        bn_max = std::max(bn_max, p.BN);
        grp.flop += 2.0 * p.M * p.N * (double)p.K;
    }
    if (bias > kGemmBiasMax) return -1;
    grp.gp.n_problems = (int)probs.size(); grp.gp.n_work = work; grp.gp.n_bias = bias;
    grp.gp.b_stage_bytes = (uint32_t)bn_max * 128; grp.gp.bn_max = (uint32_t)bn_max;
    grp.gp.out_stage_bytes = (uint32_t)((bn_max + 63) / 64) * 16384u;
    const size_t stage = 128 * 128 + grp.gp.b_stage_bytes;
    grp.gp.stages = (int)std::max<size_t>(2, std::min<size_t>(8, (192 * 1024 - grp.gp.out_stage_bytes) / stage));
    grp.gp.tmem_cols = (uint32_t)pow2_at_least(2 * bn_max);
    grp.smem = (size_t)grp.gp.stages * stage + grp.gp.out_stage_bytes + 16 * grp.gp.stages + 64 + 1024 + kGemmBiasMax * 4;
    grp.grid = std::min(work, 148);
    grp.probs = dalloc<KGemmProblem>(t, probs.size());
    FCNOD_CUDA(cudaMemcpy(grp.probs, probs.data(), probs.size() * sizeof(KGemmProblem), cudaMemcpyHostToDevice));
    t->kgroups.push_back(grp);
    return (int)t->kgroups.size() - 1;
}

// sibling sets: tensor-core 1x1 / stride-1 convolutions that read the same blob (the 1x1, 3x3_reduce and 5x5_reduce of an
// inception module, the two reduce layers being produced early is harmless: nothing else writes their tops)
void build_kgroups(fcnod_trainer* t) {
    t->conv_fgroup.assign(t->convs.size(), -1);
    t->conv_dgroup.assign(t->convs.size(), -1);
    if (const char* e = getenv("FCNOD_TRAIN_KGROUP")) if (e[0] == '0') return;
    std::map<int, std::vector<int> > by_input;       // input blob -> convolutions, forward order
    for (size_t li = 0; li < t->layers.size(); ++li) {
        const TLayer& L = t->layers[li];
        if (L.def.type != L_CONV) continue;
        const TConv& c = t->convs[L.conv];
        if (c.tc && c.k == 1 && c.stride == 1) by_input[L.in].push_back(L.conv);
    }
    for (std::map<int, std::vector<int> >::iterator it = by_input.begin(); it != by_input.end(); ++it) {
        if (it->second.size() < 2) continue;
        const int fg = build_kgroup(t, it->second, false);
        if (fg >= 0) for (size_t i = 0; i < it->second.size(); ++i) t->conv_fgroup[it->second[i]] = fg;
        // The grouped data gradients are opt-in (FCNOD_TRAIN_DGRAD_GROUP=1): inside one launch the three bf16 reduce-adds into an
        // element arrive in no fixed order, bf16 addition is not associative, and the gradients then differ from run to run at
        // the bf16 rounding level (a few per cent at the first layers after 22 stages); launched one after the other the order
        // is fixed.  Measured gain of the group: 0.13 ms of an 8 ms step.
        const char* dge = getenv("FCNOD_TRAIN_DGRAD_GROUP");
        if (t->blobs[it->first].needs_grad && dge && dge[0] == '1') {
            const TBlob& ib = t->blobs[it->first];
            if (ib.ldc != ib.C) continue;            // the diff may have to be zeroed as a whole: dense blobs only
            const int dg = build_kgroup(t, it->second, true);
            if (dg >= 0) for (size_t i = 0; i < it->second.size(); ++i) t->conv_dgroup[it->second[i]] = dg;
        }
    }
}

// static simulation of the forward walk: which GEMMs can wait, and where they must have run
void build_fwd_deferral(fcnod_trainer* t) {
    t->conv_fdefer.assign(t->convs.size(), -1);
    t->fwd_flush.assign(t->layers.size(), -1);
    if (const char* e = getenv("FCNOD_TRAIN_KGROUP")) if (e[0] == '0') return;
    std::vector<int> pending;
    auto root = [&](int b) { return t->blobs[b].root >= 0 ? t->blobs[b].root : b; };
    auto flush = [&](size_t li) {
        if (pending.size() >= 2) {
            const int g = build_kgroup(t, pending, false);
            if (g >= 0) {
                for (size_t i = 0; i < pending.size(); ++i) t->conv_fdefer[pending[i]] = g;
                if (li < t->layers.size()) t->fwd_flush[li] = g;
            }
        }
        pending.clear();
    };
    std::map<int, int> out_root;                     // pending convolution -> root blob of its top
    for (size_t li = 0; li < t->layers.size(); ++li) {
        const TLayer& L = t->layers[li];
        if (L.def.type == L_RELU) continue;
        bool reads = false;
        for (size_t b = 0; b < L.def.bottom.size() && !reads; ++b) {
            std::map<std::string, int>::iterator bi = t->blob_index.find(L.def.bottom[b]);
            if (bi == t->blob_index.end()) continue;
            for (size_t i = 0; i < pending.size(); ++i)
                if (out_root[pending[i]] == root(bi->second)) reads = true;
        }
        if (reads) flush(li);
        if (L.def.type == L_CONV) {
            const TConv& c = t->convs[L.conv];
            if (c.tc && t->conv_fgroup[L.conv] < 0 && t->gemms[c.g_fwd].p.tma_out && pending.size() < 6) {
                pending.push_back(L.conv);
                out_root[L.conv] = root(L.out);
            }
        }
    }
    // whatever is still pending at the end would never be launched by a reader: leave those immediate
    pending.clear();
}

void launch_kgroup(fcnod_trainer* t, int gi, cudaStream_t s) {
    static std::atomic<bool> attr[64];
    int dev = 0;
    FCNOD_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !attr[dev]) {
        FCNOD_CUDA(cudaFuncSetAttribute(kgemm_group_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        if (dev >= 0 && dev < 64) attr[dev] = true;
    }
    fcnod_trainer::KGroup& g = t->kgroups[gi];
    if (t->timing) {
        if (!g.e0) { FCNOD_CUDA(cudaEventCreate(&g.e0)); FCNOD_CUDA(cudaEventCreate(&g.e1)); }
        FCNOD_CUDA(cudaEventRecord(g.e0, s));
    }
    kgemm_group_kernel<<<g.grid, kGemmThreads, g.smem, s>>>(g.probs, g.gp);
    FCNOD_CUDA(cudaGetLastError());
    if (t->timing) FCNOD_CUDA(cudaEventRecord(g.e1, s));
}

void launch_wgroup(fcnod_trainer* t, int gi, cudaStream_t s) {
    static std::atomic<bool> attr[64];
    int dev = 0;
    FCNOD_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || !attr[dev]) {
        FCNOD_CUDA(cudaFuncSetAttribute(wgrad_group_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        if (dev >= 0 && dev < 64) attr[dev] = true;
    }
    fcnod_trainer::WGroup& g = t->wgroups[gi];
    if (t->timing) {
        if (!g.e0) { FCNOD_CUDA(cudaEventCreate(&g.e0)); FCNOD_CUDA(cudaEventCreate(&g.e1)); }
        FCNOD_CUDA(cudaEventRecord(g.e0, s));
    }
    wgrad_group_kernel<<<g.grid, kGemmThreads, g.smem, s>>>(g.probs, g.gp);
    FCNOD_CUDA(cudaGetLastError());
    if (t->timing) FCNOD_CUDA(cudaEventRecord(g.e1, s));
}

template <typename TI, typename TO>
void launch_conv(const void* in, int ldi, int H, int W, int Cin, const float* w, int ldw, const float* bias, void* out, int ldo,
                 int Ho, int Wo, int Cout, int N, int k, int stride, int pad, int relu, int accum, cudaStream_t s) {
    ConvSimtParams p;
    p.in = in; p.w = w; p.bias = bias; p.out = out;
    p.N = N; p.H = H; p.W = W; p.Cin = Cin; p.ldi = ldi;
    p.Ho = Ho; p.Wo = Wo; p.Cout = Cout; p.ldo = ldo; p.ldw = ldw;
    p.kh = p.kw = k; p.stride = stride; p.pad = pad; p.relu = relu; p.accum = accum;
    const long M = (long)N * Ho * Wo;
    dim3 grid((unsigned)((M + 63) / 64), (unsigned)(ldw / 64));
    conv_simt_kernel<TI, TO><<<grid, 256, 0, s>>>(p);
    FCNOD_CUDA(cudaGetLastError());
}

template <int VEC>
void launch_im2col_k(fcnod_trainer* t, const TConv& c, const TBlob& ib, const TBlob& ob, cudaStream_t s) {
    const unsigned per_row = (unsigned)c.ldk / 8u;
    dim3 grid((ob.W * per_row + 255) / 256, ob.H, t->batch);
    const bf16* x = as<bf16>(bdata(t, ib));
    const unsigned m1 = div_magic(per_row), m2 = div_magic((unsigned)c.Cin);
#define FCNOD_IM2COL(KK) launch_k((im2col_kernel<VEC, KK>), grid, 256, 0, s, x, ib.ldc, c.d_col, c.ldk, ib.H, ib.W, c.Cin, ob.W, c.stride, c.pad, m1, m2)
    if (c.k == 3) FCNOD_IM2COL(3); else if (c.k == 5) FCNOD_IM2COL(5); else if (c.k == 7) FCNOD_IM2COL(7);
    else FCNOD_CHECK(false, "trainer: im2col supports 3x3, 5x5 and 7x7 kernels");
#undef FCNOD_IM2COL
    FCNOD_CUDA(cudaGetLastError());
}
void launch_im2col(fcnod_trainer* t, const TConv& c, const TBlob& ib, const TBlob& ob, cudaStream_t s) {
    FCNOD_CHECK(ob.W * (c.ldk / 8) < 65536 && c.ldk < 65536, "trainer: im2col row too long for the fast division");
    static std::atomic<int> c3(-1);      // FCNOD_TRAIN_IM2COL_C3=0: the generic gather for the 3-channel image too
    if (c3 < 0) { const char* e = getenv("FCNOD_TRAIN_IM2COL_C3"); c3 = (e && e[0] == '0') ? 0 : 1; }
    if (c3 && c.k == 7 && c.Cin == 3 && ib.ldc == 3 && ib.c_off == 0 && (ib.W * 3) % 8 == 0 && t->batch < 65536) {
        const int lp = (c.pad * 3 + 7) / 8 * 8;
        const int reach = std::max(ib.W, (ob.W - 1) * c.stride - c.pad + c.k);
        const int pitch = (lp + reach * 3 + 7) / 8 * 8;
        const size_t smem = (size_t)c.k * pitch * sizeof(bf16);
        if (smem <= 48 * 1024) {
            launch_k(im2col_c3_kernel<7>, dim3(ob.H, t->batch), 256, smem, s, as<bf16>(bdata(t, ib)), c.d_col, c.ldk, ib.H, ib.W, ob.W, c.stride, c.pad,
                     lp, pitch, div_magic((unsigned)c.ldk / 8u));
            FCNOD_CUDA(cudaGetLastError());
            return;
        }
    }
    if (c.Cin % 8 == 0 && ib.ldc % 8 == 0 && ib.c_off % 8 == 0) launch_im2col_k<8>(t, c, ib, ob, s);
    else launch_im2col_k<1>(t, c, ib, ob, s);
}

template <typename T>
void forward(fcnod_trainer* t, const float* data_nchw, cudaStream_t s) {
    const int N = t->batch;
    const bool tcp = sizeof(T) == 2;
    if (tcp) {
        // bf16 copies of the (just updated) parameters
        cast_flat_kernel<<<grid1((long)t->n_params, 256), 256, 0, s>>>(t->params, t->wb, (long)t->n_params);
        if (t->n_pack > 0) pack_wf_kernel<<<dim3(64, t->n_pack), 256, 0, s>>>(t->pack_descs);
        FCNOD_CUDA(cudaGetLastError());
    }
    for (size_t li = 0; li < t->layers.size(); ++li) {
        TLayer& L = t->layers[li];
        if (tcp && !t->fwd_flush.empty() && t->fwd_flush[li] >= 0) launch_kgroup(t, t->fwd_flush[li], s);     // deferred GEMMs this layer reads
        if (L.skip) continue;
        const TBlob& ib = t->blobs[L.in];
        const TBlob& ob = t->blobs[L.out];
        const long in_px = (long)N * ib.H * ib.W, out_px = (long)N * ob.H * ob.W;
        switch (L.def.type) {
        case L_POWER:
            nchw_to_nhwc_power_kernel<T><<<grid1(in_px * ib.C, 256), 256, 0, s>>>(data_nchw, as<T>(bdata(t, ob)), N, ib.C, ib.H, ib.W, L.def.scale, L.def.shift);
            break;
        case L_CONV: {
            const TConv& c = t->convs[L.conv];
            if (tcp && c.tc) {
                const int fg = t->conv_fgroup[L.conv];
                if (fg >= 0) {
                    // sibling 1x1 convolutions: all of them at the first one, nothing at the others
                    if (t->kgroups[fg].convs.front() == L.conv) launch_kgroup(t, fg, s);
                    break;
                }
                if (!(c.k == 1 && c.stride == 1)) launch_im2col(t, c, ib, ob, s);
                if (t->conv_fdefer[L.conv] < 0) run_gemm(t, c.g_fwd, 0, s);       // else: with its group, before the first reader
            } else if (tcp && c.narrow) {
                narrow_fwd_kernel<<<std::min<long>((out_px + 7) / 8, 148 * 8), 256, (size_t)c.Cin * c.Cout * 4, s>>>(
                    as<bf16>(bdata(t, ib)), ib.ldc, t->params + c.w_off, c.ldw, t->params + c.b_off, as<bf16>(bdata(t, ob)), ob.ldc, (unsigned)out_px,
                    c.Cin, c.Cout);
            } else {
                launch_conv<T, T>(bdata(t, ib), ib.ldc, ib.H, ib.W, c.Cin, t->params + c.w_off, c.ldw, t->params + c.b_off, bdata(t, ob), ob.ldc,
                                  ob.H, ob.W, c.Cout, N, c.k, c.stride, c.pad, L.relu ? 1 : 0, 0, s);
            }
            break;
        }
        case L_POOL: {
            PoolGeo g = {N, ib.C, ib.H, ib.W, ob.H, ob.W, L.def.kernel, L.def.stride, L.def.pad, ib.ldc, ob.ldc};
            if (tcp && ib.C % 8 == 0 && ib.ldc % 8 == 0 && ib.c_off % 8 == 0 && ob.ldc % 8 == 0 && ob.c_off % 8 == 0)
            {
                const dim3 pg((ob.W * (ob.C / 8) + 255) / 256, ob.H, N);
                const unsigned mcv = div_magic((unsigned)ob.C / 8u);
                if (g.k == 3 && g.s == 2) launch_k(pool_fwd8_kernel<3, 2>, pg, 256, 0, s, as<bf16>(bdata(t, ib)), as<bf16>(bdata(t, ob)), L.idx, g, mcv);
                else if (g.k == 3 && g.s == 1) launch_k(pool_fwd8_kernel<3, 1>, pg, 256, 0, s, as<bf16>(bdata(t, ib)), as<bf16>(bdata(t, ob)), L.idx, g, mcv);
                else FCNOD_CHECK(false, "trainer: bf16 pooling supports 3x3 windows with stride 1 or 2");
            }
            else
                pool_fwd_kernel<T><<<grid1(out_px * ob.C, 256), 256, 0, s>>>(as<T>(bdata(t, ib)), as<T>(bdata(t, ob)), L.idx, g);
            break;
        }
        case L_LRN:
            if (tcp && L.def.local_size == 5 && ob.C % 8 == 0 && ib.ldc % 8 == 0 && ib.c_off % 8 == 0 && ob.ldc % 8 == 0 && ob.c_off % 8 == 0)
                launch_k(lrn_fwd8_kernel, grid1(out_px * (ob.C / 8), 256), 256, 0, s, as<bf16>(bdata(t, ib)), ib.ldc, as<bf16>(bdata(t, ob)), ob.ldc, L.scale, out_px,
                                                                                 ob.C, L.def.alpha / 5.f, L.def.beta, L.def.lrn_k);
            else
                lrn_fwd_kernel<T><<<grid1(out_px * ob.C, 256), 256, 0, s>>>(as<T>(bdata(t, ib)), ib.ldc, as<T>(bdata(t, ob)), ob.ldc, L.scale, out_px, ob.C,
                                                                             L.def.local_size, L.def.alpha / (float)L.def.local_size, L.def.beta, L.def.lrn_k);
            break;
        case L_DROPOUT: {
            const long n = out_px * ob.C;
            if (t->solver.dropout_ratio_override >= 0.f) L.def.dropout_ratio = t->solver.dropout_ratio_override;
            dropout_mask_kernel<<<grid1(n, 256), 256, 0, s>>>(L.mask, n, t->solver.seed, (uint64_t)t->iter, L.def.dropout_ratio);
            dropout_apply_kernel<T><<<grid1(n, 256), 256, 0, s>>>(as<T>(bdata(t, ib)), L.mask, as<T>(bdata(t, ob)), n, 1.f / (1.f - L.def.dropout_ratio), 0);
            break;
        }
        case L_SIGMOID:
            sigmoid_fwd_kernel<T><<<grid1(out_px * ob.C, 256), 256, 0, s>>>(as<T>(bdata(t, ib)), as<T>(bdata(t, ob)), out_px * ob.C);
            break;
        default:
            FCNOD_CHECK(false, "trainer: layer type '" + L.def.type_name + "' has no training forward");
        }
        FCNOD_CUDA(cudaGetLastError());
    }
}

template <typename T>
void backward(fcnod_trainer* t, cudaStream_t s) {
    const int N = t->batch;
    const bool tcp = sizeof(T) == 2;
    for (size_t b = 0; b < t->blobs.size(); ++b) t->blobs[b].diff_written = false;
    for (size_t g = 0; g < t->rb_groups.size(); ++g) t->rb_groups[g].done = false;
    root_of(t, t->cov_blob).diff_written = true;      // written by the head
    root_of(t, t->box_blob).diff_written = true;
    for (size_t li = t->layers.size(); li-- > 0;) {
        TLayer& L = t->layers[li];
        if (L.skip) continue;
        const TBlob& ib = t->blobs[L.in];
        const TBlob& ob = t->blobs[L.out];
        const long in_px = (long)N * ib.H * ib.W, out_px = (long)N * ob.H * ob.W;
        if (!root_of(t, L.out).diff_written) {
            // nothing consumed this top (not the case in train_val.prototxt): its gradient is zero
            FCNOD_CHECK(ob.ldc == ob.C, "trainer: unconsumed sliced blob");
            FCNOD_CUDA(cudaMemsetAsync(bdiff(t, ob), 0, (size_t)out_px * ob.C * sizeof(T), s));
            root_of(t, L.out).diff_written = true;
        }
        switch (L.def.type) {
        case L_POWER:
            break;
        case L_SIGMOID:
            sigmoid_bwd_kernel<T><<<grid1(out_px * ob.C, 256), 256, 0, s>>>(as<T>(bdata(t, ob)), as<T>(bdiff(t, ob)), as<T>(bdiff(t, ib)), out_px * ob.C,
                                                                             claim_diff(t, L.in));
            break;
        case L_DROPOUT: {
            const long n = out_px * ob.C;
            dropout_apply_kernel<T><<<grid1(n, 256), 256, 0, s>>>(as<T>(bdiff(t, ob)), L.mask, as<T>(bdiff(t, ib)), n, 1.f / (1.f - L.def.dropout_ratio),
                                                                   claim_diff(t, L.in));
            break;
        }
        case L_POOL: {
            PoolGeo g = {N, ib.C, ib.H, ib.W, ob.H, ob.W, L.def.kernel, L.def.stride, L.def.pad, ib.ldc, ob.ldc};
            if (tcp && ib.C % 8 == 0 && ib.ldc % 8 == 0 && ib.c_off % 8 == 0 && ob.ldc % 8 == 0 && ob.c_off % 8 == 0)
            {
                const dim3 pg((ib.W * (ib.C / 8) + 255) / 256, ib.H, N);
                const unsigned mcv = div_magic((unsigned)ib.C / 8u);
                const int accum = claim_diff(t, L.in);
                if (g.k == 3 && g.s == 2) launch_k(pool_bwd8_kernel<3, 2>, pg, 256, 0, s, as<bf16>(bdiff(t, ob)), L.idx, as<bf16>(bdiff(t, ib)), g, accum, mcv);
                else if (g.k == 3 && g.s == 1) launch_k(pool_bwd8_kernel<3, 1>, pg, 256, 0, s, as<bf16>(bdiff(t, ob)), L.idx, as<bf16>(bdiff(t, ib)), g, accum, mcv);
                else FCNOD_CHECK(false, "trainer: bf16 pooling supports 3x3 windows with stride 1 or 2");
            }
            else
                pool_bwd_kernel<T><<<grid1(in_px * ib.C, 256), 256, 0, s>>>(as<T>(bdiff(t, ob)), L.idx, as<T>(bdiff(t, ib)), g, claim_diff(t, L.in));
            break;
        }
        case L_LRN:
            if (tcp && L.def.local_size == 5 && ob.C % 8 == 0 && ib.ldc % 8 == 0 && ib.c_off % 8 == 0 && ob.ldc % 8 == 0 && ob.c_off % 8 == 0)
                launch_k(lrn_bwd8_kernel, grid1(out_px * (ob.C / 8), 256), 256, 0, s, as<bf16>(bdata(t, ib)), ib.ldc, as<bf16>(bdata(t, ob)), as<bf16>(bdiff(t, ob)),
                                                                                 ob.ldc, L.scale, as<bf16>(bdiff(t, ib)), out_px, ob.C, L.def.alpha / 5.f, L.def.beta,
                                                                                 claim_diff(t, L.in));
            else
                lrn_bwd_kernel<T><<<grid1(out_px * ob.C, 256), 256, 0, s>>>(as<T>(bdata(t, ib)), ib.ldc, as<T>(bdata(t, ob)), as<T>(bdiff(t, ob)), ob.ldc, L.scale,
                                                                             as<T>(bdiff(t, ib)), out_px, ob.C, L.def.local_size,
                                                                             L.def.alpha / (float)L.def.local_size, L.def.beta, claim_diff(t, L.in));
            break;
        case L_CONV: {
            const TConv& c = t->convs[L.conv];
            T* dy = as<T>(bdiff(t, ob));
            if (tcp && ob.C % 8 == 0 && ob.ldc % 8 == 0 && ob.c_off % 8 == 0) {
                // ReLU mask and bias gradient in one pass
                static std::atomic<long> min_chunk(0);      // FCNOD_TRAIN_RB_CHUNK: pixels per block of the small layers (multiple of 64; 64: 392 blocks per
                                                 // channel group at 28x28 instead of 98 at 256 -- measured 7.79 -> 7.74 ms per step)
                if (!min_chunk) { const char* e = getenv("FCNOD_TRAIN_RB_CHUNK"); min_chunk = e ? std::max(64, atoi(e) / 64 * 64) : 64; }
                const unsigned chunk = (unsigned)std::max<long>(min_chunk, (out_px + 148 * 16 - 1) / (148 * 16));      // ~16 blocks per SM: bytes in flight
                const int rg = t->conv_rbgroup.empty() ? -1 : t->conv_rbgroup[L.conv];
                if (rg >= 0) {
                    // the whole Concat blob at the first of its convolutions reached: every consumer of the blob has run
                    fcnod_trainer::RbGroup& g = t->rb_groups[rg];
                    if (!g.done) {
                        const TBlob& rb = t->blobs[g.root];
                        dim3 grid((unsigned)((rb.C + 63) / 64), (unsigned)((out_px + chunk - 1) / chunk));
                        launch_k(relu_bias8_kernel, grid, dim3(8, 32), 0, s, as<bf16>(bdata(t, rb)), as<bf16>(bdiff(t, rb)), (unsigned)out_px, rb.C, rb.ldc, chunk,
                                                                      1, t->grads, g.segs);
                        g.done = true;
                    }
                } else {
                    RbSegs none;
                    none.n = 0;
                    dim3 grid((unsigned)((c.Cout + 63) / 64), (unsigned)((out_px + chunk - 1) / chunk));
                    launch_k(relu_bias8_kernel, grid, dim3(8, 32), 0, s, as<bf16>(bdata(t, ob)), as<bf16>(bdiff(t, ob)), (unsigned)out_px, c.Cout, ob.ldc, chunk,
                                                                  L.relu ? 1 : 0, t->grads + c.b_off, none);
                }
                FCNOD_CUDA(cudaGetLastError());
            } else {
                if (L.relu) {
                    relu_mask_kernel<T><<<grid1(out_px * ob.C, 256), 256, 0, s>>>(as<T>(bdata(t, ob)), dy, out_px, ob.C, ob.ldc);
                    FCNOD_CUDA(cudaGetLastError());
                }
                const long chunk = std::max<long>(256, (out_px + 147) / 148);
                dim3 grid((unsigned)((c.Cout + 31) / 32), (unsigned)((out_px + chunk - 1) / chunk));
                bias_grad_kernel<T><<<grid, dim3(32, 8), 0, s>>>(dy, ob.ldc, c.Cout, out_px, chunk, t->grads + c.b_off);
                FCNOD_CUDA(cudaGetLastError());
            }
            if (tcp && c.tc) {
                const bool direct = c.k == 1 && c.stride == 1;
                {   // weight gradient (k x k: on the im2col matrix the forward left in c.d_col): alone, or with its group once the
                    // group's last member has been reached (the members' dY slices are final and nothing later touches them)
                    const int gi = t->conv_wgroup[L.conv];
                    if (gi < 0) run_gemm(t, c.g_wgrad, 0, s);
                    else if (t->wgroups[gi].convs.back() == L.conv) launch_wgroup(t, gi, s);
                }
                const int dg = t->conv_dgroup[L.conv];
                if (ib.needs_grad && dg >= 0) {
                    // sibling data gradients, all accumulating into the same diff: launched together at the last one reached
                    // (the first of the forward order); whoever wrote the diff before (the module's pool branch) stays below them
                    if (t->kgroups[dg].convs.front() == L.conv) {
                        if (!claim_diff(t, L.in)) FCNOD_CUDA(cudaMemsetAsync(bdiff(t, ib), 0, (size_t)in_px * ib.C * sizeof(T), s));
                        launch_kgroup(t, dg, s);
                    }
                } else if (ib.needs_grad) {
                    const int accum = claim_diff(t, L.in);
                    if (direct) {
                        run_gemm(t, c.g_dgrad, accum, s);
                    } else {
                        run_gemm(t, c.g_dgrad, 0, s);         // column gradient into the scratch
                        FCNOD_CHECK(ib.W * (c.Cin / 8) < 65536, "trainer: col2im row too long for the fast division");
                        dim3 cgrid((ib.W * (c.Cin / 8) + 255) / 256, ib.H, N);
                        const unsigned mcv = div_magic((unsigned)c.Cin / 8u);
                        if (c.k == 3) launch_k(col2im_kernel<3>, cgrid, 256, 0, s, t->col, c.ldk, as<bf16>(bdiff(t, ib)), ib.ldc, ib.W, c.Cin, c.pad, accum, mcv);
                        else if (c.k == 5) launch_k(col2im_kernel<5>, cgrid, 256, 0, s, t->col, c.ldk, as<bf16>(bdiff(t, ib)), ib.ldc, ib.W, c.Cin, c.pad, accum, mcv);
                        else FCNOD_CHECK(false, "trainer: col2im supports 3x3 and 5x5 kernels");
                        FCNOD_CUDA(cudaGetLastError());
                    }
                }
                break;
            }
            if (tcp && c.narrow) {
                const unsigned chunk = 128;
                narrow_wgrad_kernel<<<(unsigned)((out_px + chunk - 1) / chunk), 128, 0, s>>>(as<bf16>(bdata(t, ib)), ib.ldc, as<bf16>(bdiff(t, ob)), ob.ldc,
                                                                                              t->grads + c.w_off, c.ldw, (unsigned)out_px, chunk, c.Cin, c.Cout);
                FCNOD_CUDA(cudaGetLastError());
                if (ib.needs_grad) {
                    narrow_dgrad_kernel<<<grid1(in_px * (c.Cin / 8), 256), 256, (size_t)c.Cin * c.Cout * 4, s>>>(
                        as<bf16>(bdiff(t, ob)), ob.ldc, t->params + c.w_off, c.ldw, as<bf16>(bdiff(t, ib)), ib.ldc, (unsigned)in_px, c.Cin, c.Cout,
                        claim_diff(t, L.in));
                    FCNOD_CUDA(cudaGetLastError());
                }
                break;
            }
            {   // weights
                WgradParams p;
                p.x = bdata(t, ib); p.dy = dy; p.g = t->grads + c.w_off;
                p.N = N; p.H = ib.H; p.W = ib.W; p.Cin = c.Cin; p.ldi = ib.ldc;
                p.Ho = ob.H; p.Wo = ob.W; p.Cout = c.Cout; p.ldo = ob.ldc; p.ldw = c.ldw;
                p.k = c.k; p.stride = c.stride; p.pad = c.pad;
                const int tiles = ((c.K + 63) / 64) * (c.ldw / 64);
                long nz = std::max<long>(1, (148L * 6 + tiles - 1) / tiles);
                nz = std::min<long>(nz, std::max<long>(1, out_px / 256));
                p.chunk = ((out_px + nz - 1) / nz + 15) / 16 * 16;
                nz = (out_px + p.chunk - 1) / p.chunk;
                dim3 grid((unsigned)((c.K + 63) / 64), (unsigned)(c.ldw / 64), (unsigned)nz);
                wgrad_simt_kernel<T><<<grid, 256, 0, s>>>(p);
                FCNOD_CUDA(cudaGetLastError());
            }
            if (ib.needs_grad) {
                dgrad_pack_kernel<<<grid1((long)c.k * c.k * c.Cout * c.ldt, 256), 256, 0, s>>>(t->params + c.w_off, c.ldw, c.d_wt, c.ldt, c.Cin, c.Cout, c.k);
                FCNOD_CUDA(cudaGetLastError());
                launch_conv<T, T>(dy, ob.ldc, ob.H, ob.W, c.Cout, c.d_wt, c.ldt, 0, bdiff(t, ib), ib.ldc, ib.H, ib.W, c.Cin, N, c.k, 1, c.k - 1 - c.pad, 0,
                                  claim_diff(t, L.in), s);
            }
            break;
        }
        default:
            FCNOD_CHECK(false, "trainer: layer type '" + L.def.type_name + "' has no backward");
        }
        FCNOD_CUDA(cudaGetLastError());
    }
}

}  // namespace

extern "C" {

int fcnod_solver_defaults(fcnod_solver_params* p) {
    FCNOD_API_BEGIN
    FCNOD_CHECK(p, "null argument");
    // /root/reference/train/bounding_box/solver.prototxt:3-15 (+ caffe.proto defaults for momentum2 / delta)
    p->base_lr = 1e-4f; p->gamma = 0.1f; p->stepsize = 10000; p->momentum = 0.9f; p->momentum2 = 0.999f; p->delta = 1e-8f;
    p->weight_decay = 1e-7f; p->bbox_loss_weight = 2.0f; p->dropout_ratio_override = -1.f; p->seed = 1; p->precision = 0; p->reserved = 0;
    FCNOD_API_END
}

int fcnod_trainer_create(const char* prototxt_path, const char* caffemodel_path, int device, int batch, int height, int width,
                         const fcnod_solver_params* solver, fcnod_trainer** out) {
    FCNOD_API_BEGIN
    FCNOD_CHECK(prototxt_path && caffemodel_path && out, "null argument");
    FCNOD_CHECK(batch > 0 && height > 0 && width > 0, "trainer: bad batch / input size");
    int count = 0;
    FCNOD_CHECK(cudaGetDeviceCount(&count) == cudaSuccess && count > 0, "no CUDA device: the trainer has no CPU fallback");
    FCNOD_CHECK(device >= 0 && device < count, "trainer: device ordinal out of range");
    FCNOD_CUDA(cudaSetDevice(device));
    fcnod_trainer* t = new fcnod_trainer();
    try {
        t->device = device; t->batch = batch; t->in_h = height; t->in_w = width;
        if (solver) t->solver = *solver; else fcnod_solver_defaults(&t->solver);
        FCNOD_CHECK(t->solver.precision == 0 || t->solver.precision == 1, "trainer: precision must be 0 (float32) or 1 (bf16)");
        t->prec = t->solver.precision; t->esz = t->prec ? 2 : 4;
        t->def = parse_prototxt_file(prototxt_path, true);
        if (t->def.input_dims[1] > 0) t->in_c = t->def.input_dims[1];
        WeightMap weights = read_caffemodel(caffemodel_path);
        build_trainer(t, weights);
        verify_train_side(t->def, t->num_classes, t->solver.bbox_loss_weight);
        FCNOD_CUDA(cudaStreamCreateWithFlags(&t->stream, cudaStreamNonBlocking));
    } catch (...) {
        for (size_t i = 0; i < t->allocs.size(); ++i) cudaFree(t->allocs[i]);
        delete t;
        throw;
    }
    *out = t;
    FCNOD_API_END
}

int fcnod_trainer_destroy(fcnod_trainer* t) {
    FCNOD_API_BEGIN
    if (t) {
        cudaSetDevice(t->device);
        cudaDeviceSynchronize();
        for (int j = 0; j < t->world && t->peers_open; ++j)
            if (j != t->rank) { cudaIpcCloseMemHandle(t->peer_params[j]); cudaIpcCloseMemHandle(t->peer_grads[j]); }
        for (size_t i = 0; i < t->allocs.size(); ++i) cudaFree(t->allocs[i]);
        // timing-mode events (created lazily by fcnod_trainer_set_timing / the first timed grouped launch)
        for (size_t i = 0; i < t->ev0.size(); ++i) { if (t->ev0[i]) cudaEventDestroy(t->ev0[i]); if (t->ev1[i]) cudaEventDestroy(t->ev1[i]); }
        for (size_t i = 0; i < t->wgroups.size(); ++i) { if (t->wgroups[i].e0) cudaEventDestroy(t->wgroups[i].e0); if (t->wgroups[i].e1) cudaEventDestroy(t->wgroups[i].e1); }
        for (size_t i = 0; i < t->kgroups.size(); ++i) { if (t->kgroups[i].e0) cudaEventDestroy(t->kgroups[i].e0); if (t->kgroups[i].e1) cudaEventDestroy(t->kgroups[i].e1); }
        if (t->stream) cudaStreamDestroy(t->stream);
        delete t;
    }
    FCNOD_API_END
}

int fcnod_trainer_info(const fcnod_trainer* t, int dims[7], size_t* n_params) {
    FCNOD_API_BEGIN
    FCNOD_CHECK(t, "null argument");
    if (dims) { dims[0] = t->batch; dims[1] = t->in_c; dims[2] = t->in_h; dims[3] = t->in_w; dims[4] = t->num_classes; dims[5] = t->gh; dims[6] = t->gw; }
    if (n_params) *n_params = t->n_params;
    FCNOD_API_END
}

int fcnod_trainer_buffers(fcnod_trainer* t, float** params, float** grads, float** m, float** v) {
    FCNOD_API_BEGIN
    FCNOD_CHECK(t, "null argument");
    if (params) *params = t->params;
    if (grads) *grads = t->grads;
    if (m) *m = t->m;
    if (v) *v = t->v;
    FCNOD_API_END
}

int fcnod_trainer_forward_backward_device(fcnod_trainer* t, const float* data, const float* cvg_label, const float* bbox_label,
                                          const float* size_block, const float* obj_block, const float* cvg_block, void* stream) {
    FCNOD_API_BEGIN
    FCNOD_CHECK(t && data && cvg_label && bbox_label && size_block && obj_block && cvg_block, "null argument");
    FCNOD_CUDA(cudaSetDevice(t->device));
    cudaStream_t s = (cudaStream_t)stream;      // NULL = the legacy default stream, as everywhere in this ABI
    const int N = t->batch, C = t->num_classes;
    FCNOD_CUDA(cudaMemsetAsync(t->grads, 0, t->n_params * 4, s));
    if (t->prec) forward<bf16>(t, data, s); else forward<float>(t, data, s);
    const TBlob& cb = t->blobs[t->cov_blob];
    const TBlob& bb = t->blobs[t->box_blob];
    const long n1 = (long)N * C * t->gh * t->gw;
    if (t->prec) {
        nhwc_to_nchw_kernel<bf16><<<grid1(n1, 256), 256, 0, s>>>(as<bf16>(bdata(t, cb)), cb.ldc, t->cov_nchw, N, C, t->gh, t->gw);
        nhwc_to_nchw_kernel<bf16><<<grid1(4 * n1, 256), 256, 0, s>>>(as<bf16>(bdata(t, bb)), bb.ldc, t->box_nchw, N, 4 * C, t->gh, t->gw);
    } else {
        nhwc_to_nchw_kernel<float><<<grid1(n1, 256), 256, 0, s>>>(as<float>(bdata(t, cb)), cb.ldc, t->cov_nchw, N, C, t->gh, t->gw);
        nhwc_to_nchw_kernel<float><<<grid1(4 * n1, 256), 256, 0, s>>>(as<float>(bdata(t, bb)), bb.ldc, t->box_nchw, N, 4 * C, t->gh, t->gw);
    }
    FCNOD_CUDA(cudaGetLastError());
    int rc = fcnod_train_head_forward_device(t->cov_nchw, t->box_nchw, cvg_label, bbox_label, size_block, obj_block, cvg_block, N, C, t->gh,
                                             t->gw, t->solver.bbox_loss_weight, t->losses_dev, t->dcov_nchw, t->dbox_nchw, t->head_scratch, s);
    FCNOD_CHECK(rc == 0, fcnod_last_error());
    if (t->prec) {
        nchw_to_nhwc_kernel<bf16><<<grid1(n1, 256), 256, 0, s>>>(t->dcov_nchw, as<bf16>(bdiff(t, cb)), cb.ldc, N, C, t->gh, t->gw);
        nchw_to_nhwc_kernel<bf16><<<grid1(4 * n1, 256), 256, 0, s>>>(t->dbox_nchw, as<bf16>(bdiff(t, bb)), bb.ldc, N, 4 * C, t->gh, t->gw);
    } else {
        nchw_to_nhwc_kernel<float><<<grid1(n1, 256), 256, 0, s>>>(t->dcov_nchw, as<float>(bdiff(t, cb)), cb.ldc, N, C, t->gh, t->gw);
        nchw_to_nhwc_kernel<float><<<grid1(4 * n1, 256), 256, 0, s>>>(t->dbox_nchw, as<float>(bdiff(t, bb)), bb.ldc, N, 4 * C, t->gh, t->gw);
    }
    FCNOD_CUDA(cudaGetLastError());
    if (t->prec) backward<bf16>(t, s); else backward<float>(t, s);
    FCNOD_API_END
}

static AdamArgs adam_args(const fcnod_trainer* t, float grad_scale);

int fcnod_trainer_apply_update(fcnod_trainer* t, float grad_scale, void* stream) {
    FCNOD_API_BEGIN
    FCNOD_CHECK(t, "null argument");
    FCNOD_CUDA(cudaSetDevice(t->device));
    cudaStream_t s = (cudaStream_t)stream;      // NULL = the legacy default stream, as everywhere in this ABI
    const AdamArgs a = adam_args(t, grad_scale);
    adam_kernel<<<grid1((long)t->n_params, 256), 256, 0, s>>>(t->params, t->grads, t->m, t->v, t->meta, (long)t->n_params, a);
    FCNOD_CUDA(cudaGetLastError());
    ++t->iter;
    FCNOD_API_END
}

static AdamArgs adam_args(const fcnod_trainer* t, float grad_scale) {
    const fcnod_solver_params& sp = t->solver;
    AdamArgs a;
    const int tt = t->iter + 1;
    a.rate = (float)((double)sp.base_lr * std::pow((double)sp.gamma, (double)(sp.stepsize > 0 ? t->iter / sp.stepsize : 0)));
    a.correction = (float)(std::sqrt(1.0 - std::pow((double)sp.momentum2, (double)tt)) / (1.0 - std::pow((double)sp.momentum, (double)tt)));
    a.b1 = sp.momentum; a.b2 = sp.momentum2; a.eps = sp.delta; a.weight_decay = sp.weight_decay; a.grad_scale = grad_scale;
    return a;
}

int fcnod_trainer_ipc_handles(fcnod_trainer* t, void* params_handle, void* grads_handle) {
    FCNOD_API_BEGIN
    FCNOD_CHECK(t && params_handle && grads_handle, "null argument");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "CUDA IPC handles are 64 bytes");
    FCNOD_CUDA(cudaSetDevice(t->device));
    cudaIpcMemHandle_t h;
    FCNOD_CUDA(cudaIpcGetMemHandle(&h, t->params));
    memcpy(params_handle, &h, 64);
    FCNOD_CUDA(cudaIpcGetMemHandle(&h, t->grads));
    memcpy(grads_handle, &h, 64);
    FCNOD_API_END
}

int fcnod_trainer_open_peers(fcnod_trainer* t, int world, int rank, const void* params_handles, const void* grads_handles) {
    FCNOD_API_BEGIN
    FCNOD_CHECK(t && params_handles && grads_handles, "null argument");
    FCNOD_CHECK(world >= 1 && world <= kMaxPeers && rank >= 0 && rank < world, "trainer: the peer update supports 1..8 ranks of one node");
    FCNOD_CHECK(!t->peers_open, "trainer: peers are already open");
    FCNOD_CUDA(cudaSetDevice(t->device));
    for (int j = 0; j < world; ++j) {
        if (j == rank) { t->peer_params[j] = t->params; t->peer_grads[j] = t->grads; continue; }
        cudaIpcMemHandle_t h;
        void* p = 0;
        memcpy(&h, static_cast<const char*>(params_handles) + 64 * j, 64);
        FCNOD_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
        t->peer_params[j] = static_cast<float*>(p);
        memcpy(&h, static_cast<const char*>(grads_handles) + 64 * j, 64);
        FCNOD_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
        t->peer_grads[j] = static_cast<float*>(p);
    }
    t->world = world; t->rank = rank; t->peers_open = true;
    FCNOD_API_END
}

int fcnod_trainer_apply_update_peer(fcnod_trainer* t, void* stream) {
    FCNOD_API_BEGIN
    FCNOD_CHECK(t && t->peers_open, "trainer: fcnod_trainer_open_peers has not been called");
    FCNOD_CUDA(cudaSetDevice(t->device));
    cudaStream_t s = (cudaStream_t)stream;
    const long blocks256 = (long)(t->n_params / 256);
    const long per = (blocks256 + t->world - 1) / t->world;
    const long begin = std::min(blocks256, per * t->rank) * 256, end = std::min(blocks256, per * (t->rank + 1)) * 256;
    PeerPtrs pp;
    for (int j = 0; j < kMaxPeers; ++j) { pp.grads[j] = j < t->world ? t->peer_grads[j] : 0; pp.params[j] = j < t->world ? t->peer_params[j] : 0; }
    if (end > begin) {
        peer_reduce_adam_kernel<<<grid1((end - begin) / 4, 256), 256, 0, s>>>(pp, t->world, t->rank, t->grads, t->m, t->v, t->meta, begin, end,
                                                                             adam_args(t, 1.f / (float)t->world));
        FCNOD_CUDA(cudaGetLastError());
    }
    ++t->iter;
    FCNOD_API_END
}

int fcnod_trainer_losses(fcnod_trainer* t, float losses[2], void* stream) {
    FCNOD_API_BEGIN
    FCNOD_CHECK(t && losses, "null argument");
    FCNOD_CUDA(cudaSetDevice(t->device));
    cudaStream_t s = (cudaStream_t)stream;      // NULL = the legacy default stream, as everywhere in this ABI
    FCNOD_CUDA(cudaMemcpyAsync(losses, t->losses_dev, 8, cudaMemcpyDeviceToHost, s));
    FCNOD_CUDA(cudaStreamSynchronize(s));
    FCNOD_API_END
}

int fcnod_trainer_iteration(const fcnod_trainer* t) { return t ? t->iter : -1; }

// Host-buffer form of one iteration on one GPU: upload, forward + backward, ApplyUpdate, losses back.
int fcnod_trainer_step(fcnod_trainer* t, const float* data, const float* cvg_label, const float* bbox_label, const float* size_block,
                       const float* obj_block, const float* cvg_block, float losses[2]) {
    FCNOD_API_BEGIN
    FCNOD_CHECK(t && data && cvg_label && bbox_label && size_block && obj_block && cvg_block, "null argument");
    FCNOD_CUDA(cudaSetDevice(t->device));
    const size_t n1 = (size_t)t->batch * t->num_classes * t->gh * t->gw, nd = (size_t)t->batch * t->in_c * t->in_h * t->in_w;
    if (!t->stage) t->stage = dalloc<float>(t, nd + 17 * n1);
    const float* src[6] = {data, cvg_label, bbox_label, size_block, obj_block, cvg_block};
    const size_t cnt[6] = {nd, n1, 4 * n1, 4 * n1, 4 * n1, 4 * n1};
    float* dst[6];
    float* q = t->stage;
    for (int i = 0; i < 6; ++i) { dst[i] = q; q += cnt[i]; }
    for (int i = 0; i < 6; ++i) FCNOD_CUDA(cudaMemcpyAsync(dst[i], src[i], cnt[i] * 4, cudaMemcpyHostToDevice, t->stream));
    int rc = fcnod_trainer_forward_backward_device(t, dst[0], dst[1], dst[2], dst[3], dst[4], dst[5], t->stream);
    FCNOD_CHECK(rc == 0, fcnod_last_error());
    rc = fcnod_trainer_apply_update(t, 1.f, t->stream);
    FCNOD_CHECK(rc == 0, fcnod_last_error());
    if (losses) {
        rc = fcnod_trainer_losses(t, losses, t->stream);
        FCNOD_CHECK(rc == 0, fcnod_last_error());
    } else {
        FCNOD_CUDA(cudaStreamSynchronize(t->stream));
    }
    FCNOD_API_END
}

// which: 0 = value, 1 = gradient (after the last backward; after ApplyUpdate: the regularised, scaled gradient), 2 = ADAM m, 3 = v.
// index: 0 = weights in Caffe's [Cout][Cin][k][k] order, 1 = bias.  out == NULL: only *count is returned.
int fcnod_trainer_param_read(fcnod_trainer* t, const char* layer, int index, int which, float* out, size_t* count) {
    FCNOD_API_BEGIN
    FCNOD_CHECK(t && layer && (index == 0 || index == 1) && which >= 0 && which <= 3, "trainer: bad argument");
    const TConv* c = 0;
    for (size_t i = 0; i < t->convs.size(); ++i)
        if (t->convs[i].name == layer) c = &t->convs[i];
    FCNOD_CHECK(c, std::string("trainer: no convolution named '") + layer + "'");
    const size_t n = index == 0 ? (size_t)c->Cout * c->K : (size_t)c->Cout;
    if (count) {
        const size_t cap = *count;
        *count = n;
        if (!out) return 0;
        FCNOD_CHECK(cap >= n, "trainer: output buffer too small");
    }
    FCNOD_CHECK(out, "null argument");
    FCNOD_CUDA(cudaSetDevice(t->device));
    FCNOD_CUDA(cudaDeviceSynchronize());      // whichever stream the caller ran the step on
    const float* base = which == 0 ? t->params : which == 1 ? t->grads : which == 2 ? t->m : t->v;
    if (index == 1) {
        FCNOD_CUDA(cudaMemcpy(out, base + c->b_off, n * 4, cudaMemcpyDeviceToHost));
    } else {
        std::vector<float> packed((size_t)c->K * c->ldw);
        FCNOD_CUDA(cudaMemcpy(packed.data(), base + c->w_off, packed.size() * 4, cudaMemcpyDeviceToHost));
        const int kk = c->k * c->k;
        for (int co = 0; co < c->Cout; ++co)
            for (int cc = 0; cc < c->Cin; ++cc)
                for (int rs = 0; rs < kk; ++rs)
                    out[((size_t)co * c->Cin + cc) * kk + rs] = packed[((size_t)rs * c->Cin + cc) * c->ldw + co];
    }
    FCNOD_API_END
}

const char* fcnod_trainer_param_layer(const fcnod_trainer* t, int i) {
    if (!t || i < 0 || i >= (int)t->convs.size()) return 0;
    return t->convs[i].name.c_str();
}

// which: 0 = data, 1 = diff; dense NCHW float32 out; dims = C, H, W of the blob (out may be NULL to query them)
int fcnod_trainer_blob_read(fcnod_trainer* t, const char* blob, int which, int dims[3], float* out, size_t count) {
    FCNOD_API_BEGIN
    FCNOD_CHECK(t && blob && (which == 0 || which == 1), "trainer: bad argument");
    const TBlob& b = t->blobs[blob_of(t, blob)];
    if (dims) { dims[0] = b.C; dims[1] = b.H; dims[2] = b.W; }
    if (!out) return 0;
    const size_t n = (size_t)t->batch * b.C * b.H * b.W;
    FCNOD_CHECK(count >= n, "trainer: output buffer too small");
    FCNOD_CHECK(b.buf != t->blobs[0].buf, "trainer: the network input is not kept in NHWC form");
    FCNOD_CUDA(cudaSetDevice(t->device));
    FCNOD_CUDA(cudaDeviceSynchronize());
    float* tmp = 0;
    FCNOD_CUDA(cudaMalloc((void**)&tmp, n * 4));
    char* src = which == 0 ? bdata(t, b) : bdiff(t, b);
    if (t->prec) nhwc_to_nchw_kernel<bf16><<<grid1((long)n, 256), 256, 0, t->stream>>>(as<bf16>(src), b.ldc, tmp, t->batch, b.C, b.H, b.W);
    else nhwc_to_nchw_kernel<float><<<grid1((long)n, 256), 256, 0, t->stream>>>(as<float>(src), b.ldc, tmp, t->batch, b.C, b.H, b.W);
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaMemcpyAsync(out, tmp, n * 4, cudaMemcpyDeviceToHost, t->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(t->stream);
    cudaFree(tmp);
    FCNOD_CUDA(e);
    FCNOD_API_END
}

// Timing mode for the roofline of the GEMM kernel: with on != 0 every GEMM launch of the following steps is bracketed by CUDA
// events on the launching stream; fcnod_trainer_gemm_times then returns, for the last step, the summed launch times (ms) and
// the algorithmic FLOP (2 M N K per launch) of the forward, data-gradient and weight-gradient GEMMs, and the launch counts.
int fcnod_trainer_set_timing(fcnod_trainer* t, int on) {
    FCNOD_API_BEGIN
    FCNOD_CHECK(t, "null argument");
    t->timing = on != 0;
    FCNOD_API_END
}

int fcnod_trainer_gemm_times(fcnod_trainer* t, float ms[3], double flop[3], int launches[3]) {
    FCNOD_API_BEGIN
    FCNOD_CHECK(t && ms && flop && launches, "null argument");
    FCNOD_CHECK(t->prec == 1 && t->ev0.size() == t->gemms.size(), "trainer: no timed bf16 step to report");
    FCNOD_CUDA(cudaSetDevice(t->device));
    FCNOD_CUDA(cudaDeviceSynchronize());
    for (int k = 0; k < 3; ++k) { ms[k] = 0.f; flop[k] = 0.0; launches[k] = 0; }
    for (size_t i = 0; i < t->gemms.size(); ++i) {
        float e = 0.f;
        if (cudaEventElapsedTime(&e, t->ev0[i], t->ev1[i]) != cudaSuccess) { cudaGetLastError(); continue; }    // never launched
        const int k = t->gemm_kind[i];
        const GemmTcParams& p = t->gemms[i].p;
        ms[k] += e; flop[k] += 2.0 * p.M * p.N * (double)p.K; ++launches[k];
    }
    for (size_t gi = 0; gi < t->kgroups.size(); ++gi) {
        float e = 0.f;
        if (!t->kgroups[gi].e0 || cudaEventElapsedTime(&e, t->kgroups[gi].e0, t->kgroups[gi].e1) != cudaSuccess) { cudaGetLastError(); continue; }
        const int k = t->kgroups[gi].kind;
        ms[k] += e; flop[k] += t->kgroups[gi].flop; ++launches[k];
    }
    for (size_t gi = 0; gi < t->wgroups.size(); ++gi) {
        float e = 0.f;
        if (!t->wgroups[gi].e0 || cudaEventElapsedTime(&e, t->wgroups[gi].e0, t->wgroups[gi].e1) != cudaSuccess) { cudaGetLastError(); continue; }
        ms[2] += e; flop[2] += t->wgroups[gi].flop; ++launches[2];
    }
    FCNOD_API_END
}

// The tensor-core GEMM of the bf16 engine on caller-provided device matrices (bf16, row-major with leading dimensions):
//   mode 0: out[M][N] (bf16, or float32 if out_f32) (+)= A[M][K] * B[N][K]^T (+ bias[N]) (ReLU)        forward / data gradient
//   mode 1: gout[n * ldo + m] (float32) += sum_k A[k][m] * B[k][n], m < M, n < N                         weight gradient
int fcnod_gemm_bf16_device(int mode, const void* A, int lda, const void* B, int ldb, int M, int N, int K, void* out, int ldo,
                           int out_f32, const float* bias, int relu, int accum, void* stream) {
    FCNOD_API_BEGIN
    FCNOD_CHECK(A && B && out && M > 0 && N > 0 && K > 0 && (mode == 0 || mode == 1), "gemm: bad argument");
    GemmPlan g = mode == 0 ? plan_gemm_k(static_cast<const bf16*>(A), lda, M, K, static_cast<const bf16*>(B), ldb, N, out, ldo, out_f32, bias,
                                         relu, "fcnod_gemm_bf16_device")
                           : plan_gemm_mn(static_cast<const bf16*>(A), lda, M, static_cast<const bf16*>(B), ldb, N, K, static_cast<float*>(out),
                                          ldo, "fcnod_gemm_bf16_device");
    run_plan(g, accum, (cudaStream_t)stream);
    FCNOD_API_END
}

// The argmax the last forward kept for a Pooling layer: for every output element the position of the selected input inside
// its k x k window, dh * k + dw relative to the UNCLIPPED window origin (ph * stride - pad, pw * stride - pad); [N][C][Ho][Wo].
int fcnod_trainer_pool_argmax(fcnod_trainer* t, const char* layer, uint8_t* out, size_t count, int dims[3]) {
    FCNOD_API_BEGIN
    FCNOD_CHECK(t && layer, "null argument");
    const TLayer* pl = 0;
    for (size_t i = 0; i < t->layers.size(); ++i)
        if (t->layers[i].def.type == L_POOL && t->layers[i].def.name == layer) pl = &t->layers[i];
    FCNOD_CHECK(pl, std::string("trainer: no Pooling layer named '") + layer + "'");
    const TBlob& b = t->blobs[pl->out];
    if (dims) { dims[0] = b.C; dims[1] = b.H; dims[2] = b.W; }
    if (!out) return 0;
    const size_t n = (size_t)t->batch * b.C * b.H * b.W;
    FCNOD_CHECK(count >= n, "trainer: output buffer too small");
    FCNOD_CUDA(cudaSetDevice(t->device));
    FCNOD_CUDA(cudaDeviceSynchronize());
    uint8_t* tmp = 0;
    FCNOD_CUDA(cudaMalloc((void**)&tmp, n));
    u8_nhwc_to_nchw_kernel<<<grid1((long)n, 256), 256, 0, t->stream>>>(pl->idx, tmp, t->batch, b.C, b.H, b.W);
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaMemcpyAsync(out, tmp, n, cudaMemcpyDeviceToHost, t->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(t->stream);
    cudaFree(tmp);
    FCNOD_CUDA(e);
    FCNOD_API_END
}

// The keep mask the last forward drew for the Dropout layer, as Caffe's rand_vec_ is shaped: [N][C][H][W] uint8.
int fcnod_trainer_dropout_mask(fcnod_trainer* t, uint8_t* out, size_t count, int dims[3]) {
    FCNOD_API_BEGIN
    FCNOD_CHECK(t, "null argument");
    const TLayer* dl = 0;
    for (size_t i = 0; i < t->layers.size(); ++i)
        if (t->layers[i].def.type == L_DROPOUT) dl = &t->layers[i];
    FCNOD_CHECK(dl, "trainer: the network has no Dropout layer");
    const TBlob& b = t->blobs[dl->out];
    if (dims) { dims[0] = b.C; dims[1] = b.H; dims[2] = b.W; }
    if (!out) return 0;
    const size_t n = (size_t)t->batch * b.C * b.H * b.W;
    FCNOD_CHECK(count >= n, "trainer: output buffer too small");
    FCNOD_CUDA(cudaSetDevice(t->device));
    FCNOD_CUDA(cudaDeviceSynchronize());
    u8_nhwc_to_nchw_kernel<<<grid1((long)n, 256), 256, 0, t->stream>>>(dl->mask, t->mask_nchw, t->batch, b.C, b.H, b.W);
    FCNOD_CUDA(cudaGetLastError());
    FCNOD_CUDA(cudaMemcpyAsync(out, t->mask_nchw, n, cudaMemcpyDeviceToHost, t->stream));
    FCNOD_CUDA(cudaStreamSynchronize(t->stream));
    FCNOD_API_END
}

}  // extern "C"
