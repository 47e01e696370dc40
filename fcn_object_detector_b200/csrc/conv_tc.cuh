// bf16 implicit-GEMM convolution on the 5th-generation tensor cores (sm_100a):
// tcgen05.mma (kind::f16, M=128, cta_group::1) with the fp32 accumulator in TMEM,
// operands staged in shared memory by TMA (cp.async.bulk.tensor) through an
// mbarrier ring, warp-specialised: warp 0 = TMA producer, warp 1 = TMEM
// allocator + MMA issuer, warps 2..5 = epilogue (tcgen05.ld -> bias + ReLU ->
// bf16/fp32 NHWC store straight into the concat slice).
//
// Covers every stride-1 "same" convolution of the network
// (/root/reference/models/deploy.prototxt: 1x1, 3x3 pad 1, 5x5 pad 2).
//
// GEMM view  D[M=pixels][N=Cout] = A[M][K] * B[N][K]^T,  K = kh*kw*Cin_pad.
//   A (activations, NHWC bf16) is never materialised: for K-block (r, s, c0) the
//     producer issues ONE tiled 4-D TMA box {CK ch, BW px, BH rows, 1 image} at
//     coordinate {c0, ow0 + s - pad, oh0 + r - pad, n}.  Out-of-image taps
//     (the zero padding) and channels beyond Cin are zero-filled by the TMA unit.
//     The box lands as [BW*BH rows][CK] K-major with the hardware swizzle of
//     width CK*2 bytes = exactly a canonical UMMA K-major operand.
//   B (weights) is pre-packed [Cout_pad][kh][kw][Cin_pad] bf16 and loaded as a
//     2-D box {CK, BN}.
//   A tile has BW*BH <= 128 valid rows (112 for every layer of this net: 4x28,
//   2x56, 1x112); rows above that are never stored.
#pragma once
#include <cuda.h>

#include "common.cuh"

namespace fcnod {

struct ConvTcParams {
    void* out;            // NHWC view base (bf16 or fp32)
    const float* bias;    // [>= n_tiles*BN] (zero padded)
    int H, W;             // spatial size (in == out)
    int Cout, ldo;
    int kh, kw;
    int pad_h, pad_w;     // TMA box origin = (ow0 + s - pad_w, oh0 + r - pad_h)
    int n_cchunks;        // ceil(Cin / CK)
    int CK;               // 16 / 32 / 64 channels per K-block
    int BW, BH;           // pixel box
    int tiles_x, tiles_y; // per image
    int BN;               // N tile (multiple of 16, <= 256)
    int stages;
    int relu;
    int out_fp32;
    uint32_t tmem_cols;   // power of two >= BN, >= 32
    // Optional border-aware bias (first convolution with the input Power shift folded in, see net.cu):
    // bias = bias_tab[(row_class(oh) * (bt_lo_w + bt_hi_w + 1) + col_class(ow)) * bt_ld + channel]
    const float* bias_tab;
    int bt_lo_h, bt_hi_h, bt_lo_w, bt_hi_w, bt_ld;
};

// class of an output coordinate: 0..lo-1 = the first `lo` positions, lo = interior, lo+1.. = the last `hi`
__host__ __device__ __forceinline__ int border_class(int o, int n, int lo, int hi) {
    if (o < lo) return o;
    if (o >= n - hi) return lo + 1 + (o - (n - hi));
    return lo;
}

namespace tc {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    return ok != 0;
}
// Bounded wait: a protocol bug becomes a trap (an error the host sees), never a hung GPU.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    for (uint32_t it = 0; it < (1u << 24); ++it)
        if (mbar_try_wait(bar, parity)) return;
    __trap();
}

__device__ __forceinline__ void tma_load_4d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1, int c2,
                                            int c3) {
    asm volatile(
        "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
        ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
        : "memory");
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1)
        : "memory");
}

// UMMA shared-memory descriptor, K-major operand, hardware swizzle of `span` bytes per row
// (span = 32 / 64 / 128 = CK*2).  Rows are `span` bytes apart; 8-row groups SBO = 8*span.
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr, uint32_t span) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3FFFF) >> 4);                // start address  [0,14)
    d |= (uint64_t)1 << 16;                                  // LBO (ignored for swizzled K-major)
    d |= (uint64_t)((8u * span) >> 4) << 32;                 // SBO            [32,46)
    d |= (uint64_t)1 << 46;                                  // descriptor version (Blackwell)
    uint64_t layout = span == 128 ? 2 : (span == 64 ? 4 : 6);  // SWIZZLE_128B / 64B / 32B
    d |= layout << 61;
    return d;
}

__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accum) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(accum)
        : "memory");
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t* r) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
          "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
          "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t* r) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// bias + ReLU + store of NC (16 or 32) consecutive output channels of one pixel
template <int NC>
__device__ __forceinline__ void store_chunk(const uint32_t* r, const float* __restrict__ bias, int col0, int ncol_valid,
                                            void* out_pix, bool relu, bool out_fp32) {
    float v[NC];
#pragma unroll
    for (int j = 0; j < NC; ++j) {
        float x = __uint_as_float(r[j]) + __ldg(bias + col0 + j);
        v[j] = relu ? fmaxf(x, 0.f) : x;
    }
    if (out_fp32) {
        float* o = static_cast<float*>(out_pix) + col0;
#pragma unroll
        for (int j = 0; j < NC; ++j)
            if (j < ncol_valid) o[j] = v[j];
    } else {
        bf16* o = static_cast<bf16*>(out_pix) + col0;
        if (ncol_valid >= NC) {
#pragma unroll
            for (int j = 0; j < NC; j += 8) {
                uint4 q;
                __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&q);
#pragma unroll
                for (int t = 0; t < 4; ++t) h[t] = __floats2bfloat162_rn(v[j + 2 * t], v[j + 2 * t + 1]);
                *reinterpret_cast<uint4*>(o + j) = q;
            }
        } else {
#pragma unroll
            for (int j = 0; j < NC; ++j)
                if (j < ncol_valid) o[j] = __float2bfloat16_rn(v[j]);
        }
    }
}

}  // namespace tc

constexpr int kConvTcThreads = 192;

__global__ void __launch_bounds__(kConvTcThreads) conv_tc_kernel(const __grid_constant__ CUtensorMap tmap_a,
                                                                 const __grid_constant__ CUtensorMap tmap_b,
                                                                 const ConvTcParams p) {
    using namespace tc;
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    // carve: [stages][A 16 KB][B BN*CK*2 rounded to 1 KB], then barriers
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t span = (uint32_t)p.CK * 2u;
    const uint32_t a_bytes = 128u * span;
    const uint32_t b_bytes = ((uint32_t)p.BN * span + 1023u) & ~1023u;
    const uint32_t stage_bytes = a_bytes + b_bytes;
    const uint32_t bar_base = smem_base + (uint32_t)p.stages * stage_bytes;   // 8-byte aligned
    // barriers: full[stages], empty[stages], tmem_full, then tmem address slot
    const uint32_t full0 = bar_base, empty0 = bar_base + 8u * p.stages, tfull = bar_base + 16u * p.stages;
    const uint32_t tmem_slot = tfull + 8u;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int tile = blockIdx.x;
    const int tpi = p.tiles_x * p.tiles_y;
    const int n_img = tile / tpi;
    const int trem = tile - n_img * tpi;
    const int ty = trem / p.tiles_x, tx = trem - ty * p.tiles_x;
    const int oh0 = ty * p.BH, ow0 = tx * p.BW;
    const int n0 = blockIdx.y * p.BN;
    const int KB = p.kh * p.kw * p.n_cchunks;

    if (threadIdx.x == 0) {
        for (int s = 0; s < p.stages; ++s) { mbar_init(full0 + 8u * s, 1); mbar_init(empty0 + 8u * s, 1); }
        mbar_init(tfull, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(p.tmem_cols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    fence_before();
    __syncthreads();
    fence_after();
    uint32_t tmem_base;
    asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));

    if (warp == 0) {
        // ================= TMA producer =================
        if (lane == 0) {
            asm volatile("prefetch.tensormap [%0];" ::"l"(&tmap_a) : "memory");
            asm volatile("prefetch.tensormap [%0];" ::"l"(&tmap_b) : "memory");
            const uint32_t tx_bytes = (uint32_t)(p.BW * p.BH) * span + (uint32_t)p.BN * span;
            int kb = 0;
            uint32_t stage = 0, phase = 0;
            for (int r = 0; r < p.kh; ++r)
                for (int s = 0; s < p.kw; ++s)
                    for (int cc = 0; cc < p.n_cchunks; ++cc, ++kb) {
                        mbar_wait(empty0 + 8u * stage, phase ^ 1u);
                        const uint32_t fb = full0 + 8u * stage;
                        mbar_expect_tx(fb, tx_bytes);
                        const uint32_t sa = smem_base + stage * stage_bytes;
                        tma_load_4d(sa, &tmap_a, fb, cc * p.CK, ow0 + s - p.pad_w, oh0 + r - p.pad_h, n_img);
                        tma_load_2d(sa + a_bytes, &tmap_b, fb, kb * p.CK, n0);
                        if (++stage == (uint32_t)p.stages) { stage = 0; phase ^= 1u; }
                    }
        }
    } else if (warp == 1) {
        // ================= MMA issuer =================
        if (lane == 0) {
            // instruction descriptor: D=f32, A=B=bf16, both K-major, N=BN, M=128
            const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(p.BN >> 3) << 17) | (8u << 24);
            uint32_t stage = 0, phase = 0;
            const int ksteps = p.CK / 16;
            for (int kb = 0; kb < KB; ++kb) {
                mbar_wait(full0 + 8u * stage, phase);
                fence_after();
                const uint32_t sa = smem_base + stage * stage_bytes;
                const uint64_t da = umma_desc(sa, span), db = umma_desc(sa + a_bytes, span);
                for (int k = 0; k < ksteps; ++k)
                    umma_bf16(tmem_base, da + (uint64_t)(2 * k), db + (uint64_t)(2 * k), idesc, (kb | k) != 0);
                umma_commit(empty0 + 8u * stage);      // frees the smem slot when these MMAs retire
                if (++stage == (uint32_t)p.stages) { stage = 0; phase ^= 1u; }
            }
            umma_commit(tfull);                         // accumulator complete
        }
    } else {
        // ================= epilogue (warps 2..5 -> TMEM lane quadrants warp%4) =================
        const int q = warp & 3;
        const int m = q * 32 + lane;                 // accumulator row = pixel index inside the box
        const int py = m / p.BW, px = m - py * p.BW;
        const int oh = oh0 + py, ow = ow0 + px;
        const bool valid = (m < p.BW * p.BH) && (oh < p.H) && (ow < p.W);
        const size_t pix = ((size_t)n_img * p.H + (valid ? oh : 0)) * p.W + (valid ? ow : 0);
        void* out_pix = p.out_fp32 ? (void*)(static_cast<float*>(p.out) + pix * p.ldo)
                                   : (void*)(static_cast<bf16*>(p.out) + pix * p.ldo);
        const float* bias = p.bias + n0;
        if (p.bias_tab) {
            const int cls = border_class(valid ? oh : 0, p.H, p.bt_lo_h, p.bt_hi_h) * (p.bt_lo_w + p.bt_hi_w + 1) +
                            border_class(valid ? ow : 0, p.W, p.bt_lo_w, p.bt_hi_w);
            bias = p.bias_tab + (size_t)cls * p.bt_ld + n0;
        }
        mbar_wait(tfull, 0);
        fence_after();
        const uint32_t trow = tmem_base + ((uint32_t)(q * 32) << 16);
        const int nvalid = min(p.BN, p.Cout - n0);    // channels of this N tile that exist
        int c = 0;
        for (; c + 32 <= p.BN; c += 32) {
            uint32_t r[32];
            tmem_ld32(trow + (uint32_t)c, r);
            tmem_ld_wait();
            if (valid && c < nvalid) store_chunk<32>(r, bias, c, nvalid - c, (char*)out_pix + (size_t)n0 * (p.out_fp32 ? 4 : 2) , p.relu, p.out_fp32);
        }
        if (c < p.BN) {
            uint32_t r[16];
            tmem_ld16(trow + (uint32_t)c, r);
            tmem_ld_wait();
            if (valid && c < nvalid) store_chunk<16>(r, bias, c, nvalid - c, (char*)out_pix + (size_t)n0 * (p.out_fp32 ? 4 : 2), p.relu, p.out_fp32);
        }
        fence_before();
    }
    __syncthreads();
    if (warp == 1) {
        fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(p.tmem_cols) : "memory");
    }
}

}  // namespace fcnod
