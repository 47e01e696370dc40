// fp32 SIMT implicit-GEMM convolution (any kernel / stride / pad) on NHWC activations.
//
// This is the PARITY-MODE convolution (FCNOD_PRECISION_FP32: fp32 operands, fp32 FFMA
// accumulation, 1e-5 against the Caffe-semantics oracle) and the fallback for
// shapes the tcgen05 kernel does not take (strided convolutions).  Caffe
// Convolution semantics (cross-correlation, zero padding, out = floor((H+2p-k)/s)+1),
// e.g. /root/reference/models/deploy.prototxt:20-47.
//
// GEMM view: M = N*Ho*Wo pixels, N = Cout, K = kh*kw*Cin with k = (r*kw + s)*Cin + c.
// Tile 64(M) x 64(N) x 16(K), 256 threads, 4x4 register micro-tile per thread.
// Weights are pre-packed [K][Cout_pad64] fp32 so B-tile loads are coalesced.
#pragma once
#include "common.cuh"

namespace fcnod {

struct ConvSimtParams {
    const void* in;      // NHWC view base
    const float* w;      // [K][ldw]
    const float* bias;   // [Cout] or null
    void* out;           // NHWC view base
    int N, H, W, Cin, ldi;
    int Ho, Wo, Cout, ldo, ldw;
    int kh, kw, stride, pad;
    int relu;
    int accum;           // out += result (data gradient of a blob with several consumers, train_net.cu)
};

template <typename TIn, typename TOut>
__global__ void __launch_bounds__(256) conv_simt_kernel(ConvSimtParams p) {
    constexpr int BM = 64, BN = 64, BK = 16;
    __shared__ float As[BK][BM + 4];
    __shared__ float Bs[BK][BN];
    const int tid = threadIdx.x;
    const int tx = tid % 16, ty = tid / 16;          // micro-tile: rows ty*4.., cols tx*4..
    const long M = (long)p.N * p.Ho * p.Wo;
    const int K = p.kh * p.kw * p.Cin;
    const long m0 = (long)blockIdx.x * BM;
    const int n0 = blockIdx.y * BN;

    // A-load mapping: k_local = tid % 16, pixel rows (tid / 16) + 16*i
    const int a_k = tid % BK;
    int a_n[4], a_h[4], a_w[4];
    bool a_ok[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        long m = m0 + tid / BK + 16 * i;
        a_ok[i] = m < M;
        long mm = a_ok[i] ? m : 0;
        a_w[i] = (int)(mm % p.Wo);
        long t = mm / p.Wo;
        a_h[i] = (int)(t % p.Ho);
        a_n[i] = (int)(t / p.Ho);
    }
    const TIn* in = static_cast<const TIn*>(p.in);

    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

    for (int k0 = 0; k0 < K; k0 += BK) {
        // ---- A tile
        {
            int k = k0 + a_k;
            bool kok = k < K;
            int c = 0, r = 0, s = 0;
            if (kok) { c = k % p.Cin; int rs = k / p.Cin; s = rs % p.kw; r = rs / p.kw; }
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                float v = 0.f;
                int ih = a_h[i] * p.stride - p.pad + r, iw = a_w[i] * p.stride - p.pad + s;
                if (kok && a_ok[i] && ih >= 0 && ih < p.H && iw >= 0 && iw < p.W)
                    v = to_f32<TIn>(in[((long)(a_n[i] * p.H + ih) * p.W + iw) * p.ldi + c]);
                As[a_k][tid / BK + 16 * i] = v;
            }
        }
        // ---- B tile: 16 x 64 floats, thread -> (k = tid/64 + 4*i, n = tid%64)
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            int kk = tid / 64 + 4 * i;
            int k = k0 + kk;
            Bs[kk][tid % 64] = (k < K) ? p.w[(long)k * p.ldw + n0 + tid % 64] : 0.f;
        }
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < BK; ++kk) {
            float a[4], b[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) a[i] = As[kk][ty * 4 + i];
#pragma unroll
            for (int j = 0; j < 4; ++j) b[j] = Bs[kk][tx * 4 + j];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
        }
        __syncthreads();
    }

    TOut* out = static_cast<TOut*>(p.out);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        long m = m0 + ty * 4 + i;
        if (m >= M) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            int n = n0 + tx * 4 + j;
            if (n >= p.Cout) continue;
            float v = acc[i][j] + (p.bias ? p.bias[n] : 0.f);
            if (p.relu) v = fmaxf(v, 0.f);
            if (p.accum) v += to_f32<TOut>(out[m * p.ldo + n]);
            out[m * p.ldo + n] = from_f32<TOut>(v);
        }
    }
}

}  // namespace fcnod
