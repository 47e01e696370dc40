// tcgen05 GEMM of the training step (train_net.cu, bf16 engine): the three matrix products of a convolution layer of
// /root/reference/models/train_val.prototxt on the 5th-generation tensor cores, fp32 accumulation in TMEM.
//
//   forward   Y[px][Cout]  = X_col[px][Kc]  * Wf[Cout][Kc]^T      K-major operands  (MNMAJ = false)
//   dgrad     dX_col[px][Kc] = dY[px][Cout] * Wb[Kc][Cout]^T      K-major operands  (MNMAJ = false)
//   wgrad     dW[Kc][Cout] += X_col[px][Kc]^T * dY[px][Cout]      reduction over PIXELS: both operands are stored with the
//             reduction index as the OUTER dimension (NHWC: a pixel is a row of channels), i.e. "MN-major" in UMMA terms
//             (MNMAJ = true).  No transposed copies are made: the TMA boxes {64 channels, PK pixels} land as PK rows of 128
//             swizzled bytes, which is the canonical MN-major SWIZZLE_128B layout (64-element MN blocks LBO apart, 8-row K
//             groups SBO = 1024 bytes apart), and the instruction descriptor sets a_major = b_major = 1.
//             The pixel range is split over CTAs (split-K); partial tiles are added to the fp32 gradient with red.global.
// X_col is the activation matrix itself for 1x1 convolutions and an im2col matrix (bf16, written by im2col_kernel) for
// k x k ones.  One CTA = 10 warps: warp 0 TMA producer, warp 1 MMA issuer (+ TMEM allocation), warps 2..9 epilogue (two per
// TMEM lane quadrant, alternate 16-column chunks).  Persistent over output tiles, two accumulators in TMEM so the epilogue of a tile overlaps the MMAs
// of the next.
#pragma once
#include <cuda.h>

#include "common.cuh"
#include "conv_tc.cuh"

namespace fcnod {

struct GemmTcParams {
    int M, N, K;               // D is M x N, reduction length K (MNMAJ: M = Cout, N = Kc, K = pixels)
    int BN;                    // N tile: multiple of 16 (K-major) / of 64 (MN-major), <= 256
    int tiles_m, tiles_n, splits;
    int k_blocks;              // reduction blocks (64 elements K-major / PK pixels MN-major) of the whole K
    int kb_per_split;          // MNMAJ: reduction blocks per split
    int stages;
    uint32_t a_stage_bytes, b_stage_bytes, tmem_cols;
    // epilogue, K-major
    void* out; int ldo; int out_f32;      // row-major M x N view, leading dimension ldo elements
    const float* bias;                    // [N] or null
    int relu, accum;                      // accum: out += result
    int tma_out;                          // bf16 output through shared memory + TMA store (accum: TMA reduce-add); needs tmap_o
    uint32_t out_stage_bytes;             // ceil(BN / 64) * 16 KB when tma_out
    // epilogue, MN-major
    float* gout; int ldg;                 // gout[n * ldg + m] += D[m][n]
};

// tcgen05.wait::ld that also names the 16 destination registers of the load it completes: the compiler must treat them as
// written HERE, not at the (asynchronous) tcgen05.ld, so it cannot read or move them in between
__device__ __forceinline__ void gemm_ld_wait16(uint32_t* r) {
    asm volatile("tcgen05.wait::ld.sync.aligned;" : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]), "+r"(r[8]), "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15]) :: "memory");
}

constexpr int kGemmThreads = 320;       // warp 0 TMA, warp 1 MMA, warps 2..9 epilogue
constexpr int kGemmBiasMax = 1280;        // bias entries staged in shared memory (N <= 1024 + one tile of slack)
constexpr int kGemmPK = 64;               // pixels per reduction block of the MN-major form

template <bool MNMAJ>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
               const __grid_constant__ CUtensorMap tmap_o, const GemmTcParams p) {
    using namespace tc;
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t stage_bytes = p.a_stage_bytes + p.b_stage_bytes;
    const uint32_t out_stage = smem_base + (uint32_t)p.stages * stage_bytes;      // 1024-aligned: stage sizes are multiples of 1 KB
    const uint32_t bar_base = out_stage + p.out_stage_bytes;
    const uint32_t full0 = bar_base, empty0 = full0 + 8u * p.stages;
    const uint32_t tfull0 = empty0 + 8u * p.stages, tempty0 = tfull0 + 16u, tmem_slot = tempty0 + 16u;
    // bias staged in shared memory once per CTA (a global load per column in the epilogue loop costs a full memory latency
    // per chunk: ncu showed 64 % of the stall samples there)
    float* bias_s = reinterpret_cast<float*>(smem_raw + (tmem_slot + 16u - smem_u32(smem_raw)));
    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        for (int s = 0; s < p.stages; ++s) { mbar_init(full0 + 8u * s, 1); mbar_init(empty0 + 8u * s, 1); }
        for (int a = 0; a < 2; ++a) { mbar_init(tfull0 + 8u * a, 1); mbar_init(tempty0 + 8u * a, 8); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(p.tmem_cols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    // Programmatic dependent launch: everything above (barrier init, TMEM allocation) touches no global memory and
    // may run while the previous kernel of the stream drains; from here on its results are needed.
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
    if (!MNMAJ && p.bias)
        for (int i = threadIdx.x; i < kGemmBiasMax; i += kGemmThreads) bias_s[i] = i < p.N ? p.bias[i] : 0.f;
    fence_before();
    __syncthreads();
    fence_after();
    uint32_t tmem_base;
    asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));
    tmem_base = __shfl_sync(0xffffffffu, tmem_base, 0);

    const int n_work = p.tiles_m * p.tiles_n * p.splits;
    const int kb_total = p.k_blocks;

    if (warp == 0) {
        // ================================================================== TMA producer
        if (lane == 0) {
            asm volatile("prefetch.tensormap [%0];" ::"l"(&tmap_a) : "memory");
            asm volatile("prefetch.tensormap [%0];" ::"l"(&tmap_b) : "memory");
            uint32_t s = 0, ph = 0;
            for (int w = blockIdx.x; w < n_work; w += gridDim.x) {
                const int tn = w % p.tiles_n, tm = (w / p.tiles_n) % p.tiles_m, sp = w / (p.tiles_n * p.tiles_m);
                const int m0 = tm * 128, n0 = tn * p.BN;
                const int kb0 = MNMAJ ? sp * p.kb_per_split : 0;
                const int kb1 = MNMAJ ? min(kb0 + p.kb_per_split, kb_total) : kb_total;
                for (int kb = kb0; kb < kb1; ++kb) {
                    mbar_wait(empty0 + 8u * s, ph ^ 1u);
                    const uint32_t a_dst = smem_base + s * stage_bytes, b_dst = a_dst + p.a_stage_bytes;
                    mbar_expect_tx(full0 + 8u * s, stage_bytes);
                    if (!MNMAJ) {
                        tma_load_2d(a_dst, &tmap_a, full0 + 8u * s, kb * 64, m0);
                        tma_load_2d(b_dst, &tmap_b, full0 + 8u * s, kb * 64, n0);
                    } else {
                        const int px = kb * kGemmPK;
#pragma unroll 1
                        for (int i = 0; i < 2; ++i) tma_load_2d(a_dst + i * (kGemmPK * 128), &tmap_a, full0 + 8u * s, m0 + 64 * i, px);
#pragma unroll 1
                        for (int i = 0; i < p.BN / 64; ++i) tma_load_2d(b_dst + i * (kGemmPK * 128), &tmap_b, full0 + 8u * s, n0 + 64 * i, px);
                    }
                    if (++s == (uint32_t)p.stages) { s = 0; ph ^= 1u; }
                }
            }
        }
    } else if (warp == 1) {
        // ================================================================== MMA issuer
        if (lane == 0) {
            const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | (MNMAJ ? ((1u << 15) | (1u << 16)) : 0u) |
                                   ((uint32_t)(p.BN >> 3) << 17) | (8u << 24);
            // descriptor high word: SBO >> 4 | version 1 << 14 | SWIZZLE_128B (2) << 29; both forms have 8-row groups 1024 B apart
            const uint32_t desc_hi = (1024u >> 4) | (1u << 14) | (2u << 29);
            const uint32_t lbo = MNMAJ ? ((uint32_t)(kGemmPK * 128) >> 4) : 1u;
            uint32_t s = 0, ph = 0, it = 0;
            for (int w = blockIdx.x; w < n_work; w += gridDim.x, ++it) {
                const int sp = w / (p.tiles_n * p.tiles_m);
                const int kb0 = MNMAJ ? sp * p.kb_per_split : 0;
                const int kb1 = MNMAJ ? min(kb0 + p.kb_per_split, kb_total) : kb_total;
                const uint32_t acc = it & 1u, acc_ph = (it >> 1) & 1u;
                mbar_wait(tempty0 + 8u * acc, acc_ph ^ 1u);
                fence_after();
                const uint32_t tmem_d = tmem_base + acc * (uint32_t)p.BN;
                uint32_t accum = 0;
                for (int kb = kb0; kb < kb1; ++kb) {
                    mbar_wait(full0 + 8u * s, ph);
                    fence_after();
                    const uint32_t a_s = smem_base + s * stage_bytes, b_s = a_s + p.a_stage_bytes;
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        // K-major: 16 elements = 32 bytes along the swizzled row; MN-major: 16 pixel rows = 2048 bytes
                        const uint32_t adv = MNMAJ ? (uint32_t)k * 2048u : (uint32_t)k * 32u;
                        const uint64_t da = ((uint64_t)desc_hi << 32) | (uint64_t)(((a_s + adv) & 0x3FFFFu) >> 4) | ((uint64_t)lbo << 16);
                        const uint64_t db = ((uint64_t)desc_hi << 32) | (uint64_t)(((b_s + adv) & 0x3FFFFu) >> 4) | ((uint64_t)lbo << 16);
                        umma_bf16(tmem_d, da, db, idesc, accum);
                        accum = 1;
                    }
                    umma_commit(empty0 + 8u * s);
                    if (++s == (uint32_t)p.stages) { s = 0; ph ^= 1u; }
                }
                umma_commit(tfull0 + 8u * acc);
            }
        }
    } else if (!MNMAJ && p.tma_out) {
        // ================================================================== epilogue through shared memory + TMA store.
        // 8 warps (TMEM lane quadrant = warp % 4, the two warps of a quadrant take alternate 16-column chunks) convert their
        // rows to bf16 and write them into a [128 rows][64 columns] x ceil(BN / 64) staging area in the SWIZZLE_128B layout of
        // the output tensor map; one thread then issues a TMA store per 64-column region -- or a TMA reduce-add when the
        // output accumulates (the Split sum of the data gradients): full-line writes, no read-modify-write latency in the
        // epilogue warps, and rows / columns beyond M / N are clipped by the TMA unit.
        const int q = warp & 3, half = (warp - 2) >> 2;
        const bool leader = warp == 2 && lane == 0;
        uint32_t it = 0;
        for (int w = blockIdx.x; w < n_work; w += gridDim.x, ++it) {
            const int tn = w % p.tiles_n, tm = (w / p.tiles_n) % p.tiles_m;
            const int m0 = tm * 128, n0 = tn * p.BN;
            const uint32_t acc = it & 1u, acc_ph = (it >> 1) & 1u;
            const int n_chunks = (min(p.BN, p.N - n0) + 15) / 16;
            const int r = q * 32 + lane;                                   // row of the tile = TMEM lane
            // the staging area is free once the previous tile's stores have READ it
            if (leader) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            asm volatile("bar.sync 1, 256;" ::: "memory");
            mbar_wait(tfull0 + 8u * acc, acc_ph);
            fence_after();
            const uint32_t taddr = tmem_base + acc * (uint32_t)p.BN + ((uint32_t)(q * 32) << 16);
            uint32_t rA[16], rB[16];
            auto emit = [&](const uint32_t (&rr)[16], int ch) {
                const int n = n0 + 16 * ch;
                float v[16];
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                    v[j] = __uint_as_float(rr[j]);
                    if (p.bias) v[j] += bias_s[n + j];
                    if (p.relu) v[j] = fmaxf(v[j], 0.f);
                }
                uint4 w4[2];
                __nv_bfloat162* h2 = reinterpret_cast<__nv_bfloat162*>(w4);
#pragma unroll
                for (int j = 0; j < 8; ++j) h2[j] = __floats2bfloat162_rn(v[2 * j], v[2 * j + 1]);
                const uint32_t row_base = out_stage + (uint32_t)(ch >> 2) * 16384u + (uint32_t)r * 128u;
                const uint32_t u0 = (uint32_t)(ch & 3) * 2u, sw = (uint32_t)r & 7u;
                asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(row_base + ((u0 ^ sw) << 4)), "r"(w4[0].x), "r"(w4[0].y), "r"(w4[0].z), "r"(w4[0].w) : "memory");
                asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(row_base + (((u0 + 1u) ^ sw) << 4)), "r"(w4[1].x), "r"(w4[1].y), "r"(w4[1].z), "r"(w4[1].w) : "memory");
            };
            if (half < n_chunks) tmem_ld16(taddr + 16u * half, rA);
            for (int ch = half; ch < n_chunks; ch += 4) {
                gemm_ld_wait16(rA);
                if (ch + 2 < n_chunks) tmem_ld16(taddr + 16u * (ch + 2), rB);
                emit(rA, ch);
                if (ch + 2 < n_chunks) {
                    gemm_ld_wait16(rB);
                    if (ch + 4 < n_chunks) tmem_ld16(taddr + 16u * (ch + 4), rA);
                    emit(rB, ch + 2);
                }
            }
            // the accumulator is drained: hand it back to the MMA warp, then publish the staged tile
            fence_before();
            __syncwarp();
            if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(tempty0 + 8u * acc) : "memory");
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            asm volatile("bar.sync 1, 256;" ::: "memory");
            if (leader) {
                const int regions = (n_chunks + 3) >> 2;
                for (int g = 0; g < regions; ++g) {
                    const uint32_t src = out_stage + (uint32_t)g * 16384u;
                    if (p.accum)
                        asm volatile("cp.reduce.async.bulk.tensor.2d.global.shared::cta.add.tile.bulk_group [%0, {%2, %3}], [%1];"
                                     ::"l"(&tmap_o), "r"(src), "r"(n0 + 64 * g), "r"(m0) : "memory");
                    else
                        asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.tile.bulk_group [%0, {%2, %3}], [%1];"
                                     ::"l"(&tmap_o), "r"(src), "r"(n0 + 64 * g), "r"(m0) : "memory");
                }
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            }
        }
        if (leader) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    } else {
        // ================================================================== epilogue: 8 warps, TMEM lane quadrant = warp % 4,
        // the two warps of a quadrant take alternate 16-column chunks.  The accumulator chunk and (accumulate mode) the old
        // output values of the NEXT chunk are requested before the current one is processed: the loop is latency-bound.
        const int q = warp & 3, half = (warp - 2) >> 2;
        uint32_t it = 0;
        for (int w = blockIdx.x; w < n_work; w += gridDim.x, ++it) {
            const int tn = w % p.tiles_n, tm = (w / p.tiles_n) % p.tiles_m;
            const int m0 = tm * 128, n0 = tn * p.BN;
            const uint32_t acc = it & 1u, acc_ph = (it >> 1) & 1u;
            const int row = m0 + q * 32 + lane;
            const bool row_ok = row < p.M;
            const int n_chunks = (min(p.BN, p.N - n0) + 15) / 16;          // chunks that hold at least one column < N
            const bool rmw = !MNMAJ && p.accum && !p.out_f32 && row_ok;
            bf16* orow = MNMAJ ? (bf16*)0 : static_cast<bf16*>(p.out) + (long)row * p.ldo + n0;
            const bool vec_ok = !MNMAJ && !p.out_f32 && ((reinterpret_cast<uintptr_t>(orow) & 15) == 0);
            uint4 oldA[2], oldB[2];
            uint32_t rA[16], rB[16];
            auto prefetch_old = [&](uint4 (&o)[2], int ch) {
                if (rmw && vec_ok && n0 + 16 * ch + 16 <= p.N) {
                    o[0] = reinterpret_cast<const uint4*>(orow + 16 * ch)[0];
                    o[1] = reinterpret_cast<const uint4*>(orow + 16 * ch)[1];
                }
            };
            auto process = [&](const uint32_t (&rr)[16], const uint4 (&oldv)[2], int ch) {
                const int c0 = 16 * ch, n = n0 + c0;
                if (MNMAJ) {
                    // gout[n][m]: consecutive ROWS of the tile (lanes) are contiguous in memory.  4x4 transposes inside groups of
                    // four lanes (two shuffle steps) leave lane g of a group with column 4t + g of the group's four rows: one
                    // 16-byte red.global.add.v4.f32 instead of four scalar ones (the reductions are the cost of this epilogue).
                    const int g = lane & 3, b = g & 1;
                    const int mg = m0 + q * 32 + (lane & ~3);
#pragma unroll
                    for (int t = 0; t < 4; ++t) {
                        const float a0 = __uint_as_float(rr[4 * t]), a1 = __uint_as_float(rr[4 * t + 1]);
                        const float a2 = __uint_as_float(rr[4 * t + 2]), a3 = __uint_as_float(rr[4 * t + 3]);
                        const bool hi = (g & 2) != 0;
                        const float x0 = hi ? a2 : a0, x1 = hi ? a3 : a1;                 // the column pair this lane keeps
                        const float y0 = __shfl_xor_sync(0xffffffffu, hi ? a0 : a2, 2), y1 = __shfl_xor_sync(0xffffffffu, hi ? a1 : a3, 2);
                        const float xb = b ? x1 : x0, yb = b ? y1 : y0;                   // the column this lane ends up with
                        const float rx = __shfl_xor_sync(0xffffffffu, b ? x0 : x1, 1), ry = __shfl_xor_sync(0xffffffffu, b ? y0 : y1, 1);
                        // rows: own = g (xb), g ^ 2 (yb), g ^ 1 (rx), g ^ 3 (ry)
                        float o[4];
#pragma unroll
                        for (int r = 0; r < 4; ++r) o[r] = r == g ? xb : (r == (g ^ 2) ? yb : (r == (g ^ 1) ? rx : ry));
                        const int col = n + 4 * t + g;
                        if (col < p.N && mg < p.ldg)
                            asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p.gout + (long)col * p.ldg + mg), "f"(o[0]), "f"(o[1]),
                                         "f"(o[2]), "f"(o[3]) : "memory");
                    }
                } else if (row_ok) {
                    float v[16];
#pragma unroll
                    for (int j = 0; j < 16; ++j) {
                        v[j] = __uint_as_float(rr[j]);
                        if (p.bias) v[j] += bias_s[n + j];
                        if (p.relu) v[j] = fmaxf(v[j], 0.f);
                    }
                    const bool full = n + 16 <= p.N;
                    if (p.out_f32) {
                        float* o = static_cast<float*>(p.out) + (long)row * p.ldo + n;
#pragma unroll
                        for (int j = 0; j < 16; ++j)
                            if (full || n + j < p.N) o[j] = p.accum ? o[j] + v[j] : v[j];
                    } else {
                        bf16* o = orow + c0;
                        if (full && vec_ok) {
                            uint4 w4[2];
                            if (p.accum) {
                                const bf16* ov = reinterpret_cast<const bf16*>(oldv);
#pragma unroll
                                for (int j = 0; j < 16; ++j) v[j] += __bfloat162float(ov[j]);
                            }
                            __nv_bfloat162* h2 = reinterpret_cast<__nv_bfloat162*>(w4);
#pragma unroll
                            for (int j = 0; j < 8; ++j) h2[j] = __floats2bfloat162_rn(v[2 * j], v[2 * j + 1]);
                            reinterpret_cast<uint4*>(o)[0] = w4[0];
                            reinterpret_cast<uint4*>(o)[1] = w4[1];
                        } else {
#pragma unroll
                            for (int j = 0; j < 16; ++j)
                                if (n + j < p.N) o[j] = __float2bfloat16_rn(p.accum ? __bfloat162float(o[j]) + v[j] : v[j]);
                        }
                    }
                }
            };
            if (half < n_chunks) prefetch_old(oldA, half);
            mbar_wait(tfull0 + 8u * acc, acc_ph);
            fence_after();
            const uint32_t taddr = tmem_base + acc * (uint32_t)p.BN + ((uint32_t)(q * 32) << 16);
            if (half < n_chunks) tmem_ld16(taddr + 16u * half, rA);
            for (int ch = half; ch < n_chunks; ch += 4) {
                gemm_ld_wait16(rA);
                if (ch + 2 < n_chunks) { tmem_ld16(taddr + 16u * (ch + 2), rB); prefetch_old(oldB, ch + 2); }
                process(rA, oldA, ch);
                if (ch + 2 < n_chunks) {
                    gemm_ld_wait16(rB);
                    if (ch + 4 < n_chunks) { tmem_ld16(taddr + 16u * (ch + 4), rA); prefetch_old(oldA, ch + 4); }
                    process(rB, oldB, ch + 2);
                }
            }
            fence_before();
            __syncwarp();
            if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(tempty0 + 8u * acc) : "memory");
        }
    }
    fence_before();
    __syncthreads();
    if (warp == 1) {
        fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(p.tmem_cols) : "memory");
    }
}

// ------------------------------------------------------------------------------------------------ grouped weight gradients
// The weight-gradient GEMMs of a training step are independent of each other once their dY is final (after the ReLU mask) and
// of everything that follows in the backward pass (they read activations and dY, they write only their own gradient block).
// Launched one by one, most of them run 13-30 us: (Cout / 128) x (Kc / 256) tiles split over pixel ranges to fill the SMs,
// each launch paying its prologue, first-TMA latency and drain.  wgrad_group_kernel runs a LIST of them in one launch:
// one persistent grid walks the concatenated work items (tile x pixel split) of all problems of the group.
// Same MMA / TMA / epilogue as gemm_tc_kernel<true>; the per-problem tensor maps and sizes live in device memory.
struct alignas(128) WgradProblem {
    CUtensorMap ta, tb;        // dY view {Cout, pixels}, X / im2col view {Kc, pixels}: boxes {64, kGemmPK}
    int M, N, BN;              // Cout, Kc, N tile (multiple of 64, <= 256)
    int tiles_m, tiles_n, splits, k_blocks, kb_per_split;
    int work_begin;            // first work item of this problem in the group
    int ldg;
    float* gout;               // gout[n * ldg + m] += D[m][n]
};

struct WgradGroupParams {
    int n_problems, n_work, stages;
    uint32_t b_stage_bytes;    // of the widest problem (ring slots are sized for it)
    uint32_t tmem_cols, bn_max;
};

__global__ void __launch_bounds__(kGemmThreads, 1) wgrad_group_kernel(const WgradProblem* __restrict__ probs, const WgradGroupParams gp) {
    using namespace tc;
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    constexpr uint32_t kAStage = 2u * kGemmPK * 128u;
    const uint32_t stage_bytes = kAStage + gp.b_stage_bytes;
    const uint32_t bar_base = smem_base + (uint32_t)gp.stages * stage_bytes;
    const uint32_t full0 = bar_base, empty0 = full0 + 8u * gp.stages;
    const uint32_t tfull0 = empty0 + 8u * gp.stages, tempty0 = tfull0 + 16u, tmem_slot = tempty0 + 16u;
    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        for (int s = 0; s < gp.stages; ++s) { mbar_init(full0 + 8u * s, 1); mbar_init(empty0 + 8u * s, 1); }
        for (int a = 0; a < 2; ++a) { mbar_init(tfull0 + 8u * a, 1); mbar_init(tempty0 + 8u * a, 8); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(gp.tmem_cols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    fence_before();
    __syncthreads();
    fence_after();
    uint32_t tmem_base;
    asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));
    tmem_base = __shfl_sync(0xffffffffu, tmem_base, 0);

    // work item -> problem (the list is short: linear walk), tile and pixel range
    struct Item { const WgradProblem* pr; int m0, n0, kb0, kb1, BN; };
    auto decode = [&](int w) {
        int q = 0;
        while (q + 1 < gp.n_problems && probs[q + 1].work_begin <= w) ++q;
        const WgradProblem* pr = probs + q;
        const int l = w - pr->work_begin;
        const int tn = l % pr->tiles_n, tm = (l / pr->tiles_n) % pr->tiles_m, sp = l / (pr->tiles_n * pr->tiles_m);
        Item it;
        it.pr = pr; it.m0 = tm * 128; it.n0 = tn * pr->BN; it.BN = pr->BN;
        it.kb0 = sp * pr->kb_per_split; it.kb1 = min(it.kb0 + pr->kb_per_split, pr->k_blocks);
        return it;
    };

    if (warp == 0) {
        if (lane == 0) {
            uint32_t s = 0, ph = 0;
            for (int w = blockIdx.x; w < gp.n_work; w += gridDim.x) {
                const Item it = decode(w);
                const uint32_t tx = kAStage + (uint32_t)(it.BN / 64) * (kGemmPK * 128u);
                for (int kb = it.kb0; kb < it.kb1; ++kb) {
                    mbar_wait(empty0 + 8u * s, ph ^ 1u);
                    const uint32_t a_dst = smem_base + s * stage_bytes, b_dst = a_dst + kAStage;
                    mbar_expect_tx(full0 + 8u * s, tx);
                    const int px = kb * kGemmPK;
#pragma unroll 1
                    for (int i = 0; i < 2; ++i) tma_load_2d(a_dst + i * (kGemmPK * 128), &it.pr->ta, full0 + 8u * s, it.m0 + 64 * i, px);
#pragma unroll 1
                    for (int i = 0; i < it.BN / 64; ++i) tma_load_2d(b_dst + i * (kGemmPK * 128), &it.pr->tb, full0 + 8u * s, it.n0 + 64 * i, px);
                    if (++s == (uint32_t)gp.stages) { s = 0; ph ^= 1u; }
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            const uint32_t desc_hi = (1024u >> 4) | (1u << 14) | (2u << 29);
            const uint32_t lbo = (uint32_t)(kGemmPK * 128) >> 4;
            uint32_t s = 0, ph = 0, n = 0;
            for (int w = blockIdx.x; w < gp.n_work; w += gridDim.x, ++n) {
                const Item it = decode(w);
                const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | (1u << 15) | (1u << 16) | ((uint32_t)(it.BN >> 3) << 17) | (8u << 24);
                const uint32_t acc = n & 1u, acc_ph = (n >> 1) & 1u;
                mbar_wait(tempty0 + 8u * acc, acc_ph ^ 1u);
                fence_after();
                const uint32_t tmem_d = tmem_base + acc * gp.bn_max;
                uint32_t accum = 0;
                for (int kb = it.kb0; kb < it.kb1; ++kb) {
                    mbar_wait(full0 + 8u * s, ph);
                    fence_after();
                    const uint32_t a_s = smem_base + s * stage_bytes, b_s = a_s + kAStage;
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const uint32_t adv = (uint32_t)k * 2048u;
                        const uint64_t da = ((uint64_t)desc_hi << 32) | (uint64_t)(((a_s + adv) & 0x3FFFFu) >> 4) | ((uint64_t)lbo << 16);
                        const uint64_t db = ((uint64_t)desc_hi << 32) | (uint64_t)(((b_s + adv) & 0x3FFFFu) >> 4) | ((uint64_t)lbo << 16);
                        umma_bf16(tmem_d, da, db, idesc, accum);
                        accum = 1;
                    }
                    umma_commit(empty0 + 8u * s);
                    if (++s == (uint32_t)gp.stages) { s = 0; ph ^= 1u; }
                }
                umma_commit(tfull0 + 8u * acc);
            }
        }
    } else {
        const int q = warp & 3, half = (warp - 2) >> 2;
        const int g = lane & 3, b = g & 1;
        uint32_t n = 0;
        for (int w = blockIdx.x; w < gp.n_work; w += gridDim.x, ++n) {
            const Item it = decode(w);
            const int N = it.pr->N, ldg = it.pr->ldg;
            float* gout = it.pr->gout;
            const uint32_t acc = n & 1u, acc_ph = (n >> 1) & 1u;
            const int n_chunks = (min(it.BN, N - it.n0) + 15) / 16;
            const int mg = it.m0 + q * 32 + (lane & ~3);
            mbar_wait(tfull0 + 8u * acc, acc_ph);
            fence_after();
            const uint32_t taddr = tmem_base + acc * gp.bn_max + ((uint32_t)(q * 32) << 16);
            uint32_t rA[16], rB[16];
            auto reduce16 = [&](const uint32_t (&rr)[16], int ch) {
                const int col0 = it.n0 + 16 * ch;
#pragma unroll
                for (int t = 0; t < 4; ++t) {
                    const float a0 = __uint_as_float(rr[4 * t]), a1 = __uint_as_float(rr[4 * t + 1]);
                    const float a2 = __uint_as_float(rr[4 * t + 2]), a3 = __uint_as_float(rr[4 * t + 3]);
                    const bool hi = (g & 2) != 0;
                    const float x0 = hi ? a2 : a0, x1 = hi ? a3 : a1;
                    const float y0 = __shfl_xor_sync(0xffffffffu, hi ? a0 : a2, 2), y1 = __shfl_xor_sync(0xffffffffu, hi ? a1 : a3, 2);
                    const float xb = b ? x1 : x0, yb = b ? y1 : y0;
                    const float rx = __shfl_xor_sync(0xffffffffu, b ? x0 : x1, 1), ry = __shfl_xor_sync(0xffffffffu, b ? y0 : y1, 1);
                    float o[4];
#pragma unroll
                    for (int r = 0; r < 4; ++r) o[r] = r == g ? xb : (r == (g ^ 2) ? yb : (r == (g ^ 1) ? rx : ry));
                    const int col = col0 + 4 * t + g;
                    if (col < N && mg < ldg)
                        asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(gout + (long)col * ldg + mg), "f"(o[0]), "f"(o[1]),
                                     "f"(o[2]), "f"(o[3]) : "memory");
                }
            };
            if (half < n_chunks) tmem_ld16(taddr + 16u * half, rA);
            for (int ch = half; ch < n_chunks; ch += 4) {
                gemm_ld_wait16(rA);
                if (ch + 2 < n_chunks) tmem_ld16(taddr + 16u * (ch + 2), rB);
                reduce16(rA, ch);
                if (ch + 2 < n_chunks) {
                    gemm_ld_wait16(rB);
                    if (ch + 4 < n_chunks) tmem_ld16(taddr + 16u * (ch + 4), rA);
                    reduce16(rB, ch + 2);
                }
            }
            fence_before();
            __syncwarp();
            if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(tempty0 + 8u * acc) : "memory");
        }
    }
    fence_before();
    __syncthreads();
    if (warp == 1) {
        fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(gp.tmem_cols) : "memory");
    }
}

// ------------------------------------------------------------------------------------------------ grouped K-major GEMMs
// The same for independent forward / data-gradient GEMMs (the three 1x1 convolutions that read an inception module's input; the
// three data gradients that accumulate into that input's diff): one persistent grid over the tiles of all problems.
// bf16 output through shared memory and the TMA unit only (store, or reduce-add where the problem accumulates).
struct alignas(128) KGemmProblem {
    CUtensorMap ta, tb, to;    // A {K, M} box {64, 128}; B {K, N} box {64, BN}; out {N, M} box {64, 128}
    int M, N, BN, tiles_m, tiles_n, k_blocks;
    int work_begin, bias_off;  // bias_off: offset of this problem's bias in the staged table, -1 = none
    int relu, accum;
    const float* bias;
};

struct KGemmGroupParams {
    int n_problems, n_work, stages, n_bias;      // n_bias: floats of the staged bias table (all problems, each padded to its tiles)
    uint32_t b_stage_bytes, out_stage_bytes, tmem_cols, bn_max;
};

__global__ void __launch_bounds__(kGemmThreads, 1) kgemm_group_kernel(const KGemmProblem* __restrict__ probs, const KGemmGroupParams gp) {
    using namespace tc;
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    constexpr uint32_t kAStage = 128u * 128u;
    const uint32_t stage_bytes = kAStage + gp.b_stage_bytes;
    const uint32_t out_stage = smem_base + (uint32_t)gp.stages * stage_bytes;
    const uint32_t bar_base = out_stage + gp.out_stage_bytes;
    const uint32_t full0 = bar_base, empty0 = full0 + 8u * gp.stages;
    const uint32_t tfull0 = empty0 + 8u * gp.stages, tempty0 = tfull0 + 16u, tmem_slot = tempty0 + 16u;
    float* bias_s = reinterpret_cast<float*>(smem_raw + (tmem_slot + 16u - smem_u32(smem_raw)));
    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;

    if (threadIdx.x == 0) {
        for (int s = 0; s < gp.stages; ++s) { mbar_init(full0 + 8u * s, 1); mbar_init(empty0 + 8u * s, 1); }
        for (int a = 0; a < 2; ++a) { mbar_init(tfull0 + 8u * a, 1); mbar_init(tempty0 + 8u * a, 8); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(gp.tmem_cols) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    // biases of all problems, each zero-padded to tiles_n * BN entries
    for (int q = 0; q < gp.n_problems; ++q) {
        const KGemmProblem& pr = probs[q];
        if (pr.bias_off < 0) continue;
        for (int i = threadIdx.x; i < pr.tiles_n * pr.BN; i += kGemmThreads) bias_s[pr.bias_off + i] = i < pr.N ? pr.bias[i] : 0.f;
    }
    fence_before();
    __syncthreads();
    fence_after();
    uint32_t tmem_base;
    asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));
    tmem_base = __shfl_sync(0xffffffffu, tmem_base, 0);

    struct Item { const KGemmProblem* pr; int m0, n0, BN, kb; };
    auto decode = [&](int w) {
        int q = 0;
        while (q + 1 < gp.n_problems && probs[q + 1].work_begin <= w) ++q;
        const KGemmProblem* pr = probs + q;
        const int l = w - pr->work_begin;
        Item it;
        it.pr = pr; it.n0 = (l % pr->tiles_n) * pr->BN; it.m0 = (l / pr->tiles_n) * 128; it.BN = pr->BN; it.kb = pr->k_blocks;
        return it;
    };

    if (warp == 0) {
        if (lane == 0) {
            uint32_t s = 0, ph = 0;
            for (int w = blockIdx.x; w < gp.n_work; w += gridDim.x) {
                const Item it = decode(w);
                const uint32_t tx = kAStage + (uint32_t)it.BN * 128u;
                for (int kb = 0; kb < it.kb; ++kb) {
                    mbar_wait(empty0 + 8u * s, ph ^ 1u);
                    const uint32_t a_dst = smem_base + s * stage_bytes, b_dst = a_dst + kAStage;
                    mbar_expect_tx(full0 + 8u * s, tx);
                    tma_load_2d(a_dst, &it.pr->ta, full0 + 8u * s, kb * 64, it.m0);
                    tma_load_2d(b_dst, &it.pr->tb, full0 + 8u * s, kb * 64, it.n0);
                    if (++s == (uint32_t)gp.stages) { s = 0; ph ^= 1u; }
                }
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            const uint32_t desc_hi = (1024u >> 4) | (1u << 14) | (2u << 29);
            uint32_t s = 0, ph = 0, n = 0;
            for (int w = blockIdx.x; w < gp.n_work; w += gridDim.x, ++n) {
                const Item it = decode(w);
                const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(it.BN >> 3) << 17) | (8u << 24);
                const uint32_t acc = n & 1u, acc_ph = (n >> 1) & 1u;
                mbar_wait(tempty0 + 8u * acc, acc_ph ^ 1u);
                fence_after();
                const uint32_t tmem_d = tmem_base + acc * gp.bn_max;
                uint32_t accum = 0;
                for (int kb = 0; kb < it.kb; ++kb) {
                    mbar_wait(full0 + 8u * s, ph);
                    fence_after();
                    const uint32_t a_s = smem_base + s * stage_bytes, b_s = a_s + kAStage;
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const uint32_t adv = (uint32_t)k * 32u;
                        const uint64_t da = ((uint64_t)desc_hi << 32) | (uint64_t)(((a_s + adv) & 0x3FFFFu) >> 4) | ((uint64_t)1u << 16);
                        const uint64_t db = ((uint64_t)desc_hi << 32) | (uint64_t)(((b_s + adv) & 0x3FFFFu) >> 4) | ((uint64_t)1u << 16);
                        umma_bf16(tmem_d, da, db, idesc, accum);
                        accum = 1;
                    }
                    umma_commit(empty0 + 8u * s);
                    if (++s == (uint32_t)gp.stages) { s = 0; ph ^= 1u; }
                }
                umma_commit(tfull0 + 8u * acc);
            }
        }
    } else {
        const int q = warp & 3, half = (warp - 2) >> 2;
        const bool leader = warp == 2 && lane == 0;
        uint32_t n = 0;
        for (int w = blockIdx.x; w < gp.n_work; w += gridDim.x, ++n) {
            const Item it = decode(w);
            const KGemmProblem* pr = it.pr;
            const int relu = pr->relu, boff = pr->bias_off, N = pr->N;
            const uint32_t acc = n & 1u, acc_ph = (n >> 1) & 1u;
            const int n_chunks = (min(it.BN, N - it.n0) + 15) / 16;
            const int r = q * 32 + lane;
            if (leader) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            asm volatile("bar.sync 1, 256;" ::: "memory");
            mbar_wait(tfull0 + 8u * acc, acc_ph);
            fence_after();
            const uint32_t taddr = tmem_base + acc * gp.bn_max + ((uint32_t)(q * 32) << 16);
            uint32_t rA[16], rB[16];
            auto emit = [&](const uint32_t (&rr)[16], int ch) {
                const int nl = (it.n0 - 0) + 16 * ch;                      // column inside the problem
                float v[16];
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                    v[j] = __uint_as_float(rr[j]);
                    if (boff >= 0) v[j] += bias_s[boff + nl + j];
                    if (relu) v[j] = fmaxf(v[j], 0.f);
                }
                uint4 w4[2];
                __nv_bfloat162* h2 = reinterpret_cast<__nv_bfloat162*>(w4);
#pragma unroll
                for (int j = 0; j < 8; ++j) h2[j] = __floats2bfloat162_rn(v[2 * j], v[2 * j + 1]);
                const uint32_t row_base = out_stage + (uint32_t)(ch >> 2) * 16384u + (uint32_t)r * 128u;
                const uint32_t u0 = (uint32_t)(ch & 3) * 2u, sw = (uint32_t)r & 7u;
                asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(row_base + ((u0 ^ sw) << 4)), "r"(w4[0].x), "r"(w4[0].y), "r"(w4[0].z), "r"(w4[0].w) : "memory");
                asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(row_base + (((u0 + 1u) ^ sw) << 4)), "r"(w4[1].x), "r"(w4[1].y), "r"(w4[1].z), "r"(w4[1].w) : "memory");
            };
            if (half < n_chunks) tmem_ld16(taddr + 16u * half, rA);
            for (int ch = half; ch < n_chunks; ch += 4) {
                gemm_ld_wait16(rA);
                if (ch + 2 < n_chunks) tmem_ld16(taddr + 16u * (ch + 2), rB);
                emit(rA, ch);
                if (ch + 2 < n_chunks) {
                    gemm_ld_wait16(rB);
                    if (ch + 4 < n_chunks) tmem_ld16(taddr + 16u * (ch + 4), rA);
                    emit(rB, ch + 2);
                }
            }
            fence_before();
            __syncwarp();
            if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(tempty0 + 8u * acc) : "memory");
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            asm volatile("bar.sync 1, 256;" ::: "memory");
            if (leader) {
                const int regions = (n_chunks + 3) >> 2;
                for (int g = 0; g < regions; ++g) {
                    const uint32_t src = out_stage + (uint32_t)g * 16384u;
                    if (pr->accum)
                        asm volatile("cp.reduce.async.bulk.tensor.2d.global.shared::cta.add.tile.bulk_group [%0, {%2, %3}], [%1];"
                                     ::"l"(&pr->to), "r"(src), "r"(it.n0 + 64 * g), "r"(it.m0) : "memory");
                    else
                        asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.tile.bulk_group [%0, {%2, %3}], [%1];"
                                     ::"l"(&pr->to), "r"(src), "r"(it.n0 + 64 * g), "r"(it.m0) : "memory");
                }
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            }
        }
        if (leader) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    }
    fence_before();
    __syncthreads();
    if (warp == 1) {
        fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(gp.tmem_cols) : "memory");
    }
}

}  // namespace fcnod
