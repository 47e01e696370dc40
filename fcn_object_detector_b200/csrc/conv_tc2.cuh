// Persistent bf16 implicit-GEMM convolution on the 5th-generation tensor cores (sm_100a).
//
// GEMM view  D[M = pixels][N = Cout] = A[M][K] * B[N][K]^T,  K = kh*kw*Cin_pad (helpers in conv_tc.cuh):
//
//   * ONE CTA per SM looping over (pixel tile, N tile) pairs: TMEM allocation, barrier initialisation and
//     tensor-map prefetch are paid once per launch instead of once per 112-pixel tile.
//   * TWO accumulators in TMEM: the MMA warp starts the next tile while the four epilogue warps drain the
//     previous one (tmem_full / tmem_empty mbarriers).  N tiles up to 256 columns (2 x 256 = all 512 TMEM columns).
//   * Separate shared-memory rings for A (activations) and B (weights), each with full / empty mbarriers.
//   * HALO mode for k x k convolutions (3x3 pad 1, 5x5 pad 2 of /root/reference/models/deploy.prototxt): per 64-channel
//     chunk ONE TMA box brings the (BH+k-1) x (BW+k-1) input patch of the tile; it lands densely as
//     [patch pixel][CK ch] = 128-byte swizzled rows (CK = 64), or 32-byte rows (CK = 16: the space-to-depth first
//     convolution, whose pixels are 16 channels wide).  Tap (r, s) is then the SAME shared-memory tile read through a
//     UMMA descriptor whose start address is shifted by r*(BW+k-1) + s rows: accumulator row i = py*(BW+k-1) + px
//     reads patch pixel (py+r, px+s).  The 9 (25) taps re-use one load instead of 9 (25) shifted loads; rows with
//     px >= BW are computed and discarded.  TAP mode (1x1, the space-to-depth first convolution) = one A box per K-block.
//   * Epilogue through shared memory, 16 warps (the epilogue is latency-bound per warp, so it is spread wide): four per
//     TMEM lane quadrant, each draining a quarter of the columns of its 32 accumulator rows: bias (staged in smem per tile) + ReLU, pack to bf16 into a private swizzled staging area,
//     read back so that consecutive lanes store consecutive 16-byte pieces of one pixel (full lines instead of 32
//     scattered sectors per store).  No cross-warp synchronisation in the epilogue.
//   * Horizontal fusion: N may be the concatenation of several convolutions reading the same input (the 1x1 /
//     3x3_reduce / 5x5_reduce branches of an inception module, deploy.prototxt:163-377): up to four output SEGMENTS
//     with their own destination views; boundaries are multiples of 8 channels.
//   * POOL mode (the `pool -> pool_proj` branch of an inception module, deploy.prototxt:331-377 and its eight siblings):
//     the 3x3 / stride 1 / pad 1 max-pool is computed on the A-operand path.  TMA brings the raw (BH+2) x (BW+2) halo patch
//     of a 64-channel chunk into a ring of its own; eight of the epilogue warps take the maximum of the nine neighbours
//     (three 128-bit shared-memory loads per patch row, running row maxima in registers) and write the pooled
//     [BW*BH pixels][64 ch] tile in the swizzled K-major layout the MMA reads.  The pooled blob never exists in HBM.
//     Zero fill stands in for Caffe's clipped windows: the planner only fuses when the input is non-negative.
//   * LRN epilogue (conv2/3x3 -> conv2/norm2, deploy.prototxt:108-151): when the convolution's whole channel extent is one
//     N tile, the across-channel LRN of its ReLU'd output is applied by the epilogue warps on the fp32 accumulators (two
//     MUFU per element, executed by warps that otherwise wait for the tensor pipe); a warp reads 16 extra TMEM columns
//     either side of its own for the +-2 channel window.  The un-normalised blob never exists.
//   * "Direct" segments (the fp32 heads, deploy.prototxt:2118-2175) are written straight from registers as NCHW
//     float32 (lanes = consecutive pixels), with an optional sigmoid (coverage/sig :2144-2149).
#pragma once
#include <cuda.h>

#include "common.cuh"
#include "conv_tc.cuh"
#include "fused_pool.cuh"

namespace fcnod {

enum { SEG_RELU = 1, SEG_DIRECT = 2, SEG_SIGMOID = 4 };

struct ConvSeg {
    void* out;     // bf16 NHWC view base (already offset to the slice), or float NCHW blob base for SEG_DIRECT
    int ldo;       // NHWC: channels of the destination buffer; DIRECT: channels of the NCHW blob
    int col0;      // first fused-N column of this segment (multiple of 8)
    int ncols;     // valid channels (NHWC: multiple of 8)
    int flags;
};

struct ConvTc2Params {
    const float* bias;       // [n_tiles * BN] over the fused N (zero padded)
    const float* bias_tab;   // optional border-aware bias (see conv_tc.cuh)
    int bt_lo_h, bt_hi_h, bt_lo_w, bt_hi_w, bt_ld;
    int H, W;
    int kh, kw, pad_h, pad_w, n_cchunks, CK;
    int ksteps_last;         // 16-wide K steps that hold real channels in the LAST channel chunk (the rest is zero fill)
    int BW, BH, tiles_x, tiles_y;
    int PW;                  // accumulator row pitch per image row: BW (TAP) or BW + kw - 1 (HALO)
    int halo;                // 1 = HALO mode
    int pool;                // 1 = POOL mode: a 3x3 / stride 1 / pad 1 max-pool of the (non-negative) input is applied on the way
                             //     to the A operand (raw halo patches by TMA -> eight pooling warps -> UMMA layout)
    int r_stages;            //     raw-patch ring: stages, bytes per stage, bytes per TMA box
    uint32_t r_stage_bytes, r_tx_bytes;
    int lrn;                 // 1 = LRN (across channels, local_size 5, beta 0.75) of the ReLU'd output in the epilogue: needs
    float lrn_aon, lrn_k;    //     n_tiles == 1 (a pixel's whole channel extent is one TMEM row); scale = lrn_k + lrn_aon * sum x^2
    int flat;                // 1 = 1x1 convolution over the flattened pixel index (H = 1, W = batch * H * W: 128-pixel tiles
                             //     that ignore image and row boundaries; host-side bookkeeping only)
    int m_tiles, n_tiles, BN;
    int mstride;                               // pixel tiles between two visits of one CTA = gridDim.x / n_tiles
    int step_tx, step_ty, step_img;            // mstride written in the mixed radix (img, ty, tx)
    int b_resident;                            // 1 = the B ring holds the whole K extent of the CTA's N tile: loaded once
    int dual;                                  // 1 = two A/B rings and two MMA-issuing threads (even / odd pixel tiles)
    int a_bulk;                                // 1 = HALO patch rows come by cp.async.bulk from a padded, pre-swizzled image
    const void* a_src;                         //     its base, row pitch and image pitch in bytes
    unsigned a_row_pitch;
    unsigned long long a_img_pitch;
    int a_stages, b_stages;
    int b_taps;              // K-blocks per B stage: 1, or kw in HALO mode (one kernel row of taps per TMA / barrier round)
    uint32_t a_stage_bytes, b_stage_bytes, a_tx_bytes;
    uint32_t tmem_cols;      // whole allocation (2 or 4 accumulators), power of two
    uint32_t acc_stride;     // columns between two accumulators
    uint32_t acc_lg;         // log2(number of accumulators): 1, or 2 when 4 * acc_stride <= 512
    // Epilogue groups: the 16 epilogue warps drain one tile together (egroups = 1, four warps per TMEM lane quadrant, each
    // a quarter of the columns), or split into 2 / 4 groups that drain alternate tiles (8 / 4 warps per tile, each warp up
    // to 64 columns).  A tile costs every warp that touches it a fixed ~150 instructions (coordinates, barrier, addresses):
    // with narrow N tiles that overhead, paid 16 times per tile, is what bounds the kernel.
    uint32_t egroups;
    int cols_half0;          // columns drained by each warp of a quadrant's group (multiple of 16; the last ones get the rest)
    int flags;               // SEG_RELU / SEG_DIRECT: uniform over the segments of one launch
    int bt_interior;         // bias_tab: class index of interior pixels (== the row `bias` points at)
    int nseg;
    ConvSeg seg[4];
    // optional timeline of CTA 0 (fcnod_net_profile_conv_step): clock64 stamps [4 roles][8 tiles][64 events]
    unsigned long long* prof;
};

#define FCNOD_PROF(role, tile, ev)                                                                      \
    do {                                                                                                \
        if (p.prof && blockIdx.x == 0 && (tile) < 8 && (ev) < 64)                                       \
            p.prof[((role) * 8 + (tile)) * 64 + (ev)] = (unsigned long long)clock64();                  \
    } while (0)

constexpr int kTcProducerWarps = 3;   // warps 0..3: MMA issuer 0, two TMA producers, MMA issuer 1 (DUAL mode)
constexpr int kTcMmaWarp = 3;         // last non-epilogue warp: the epilogue warps start at kTcMmaWarp + 1
constexpr int kTcAllocWarp = 0;       // allocates / frees tensor memory
constexpr int kTcEpiWarps = 16;     // four per TMEM lane quadrant, each draining a quarter of the columns
constexpr int kConvTc2Threads = 32 * (kTcProducerWarps + 1 + kTcEpiWarps);   // 3 TMA warps + MMA warp + 16 epilogue warps

namespace tc {
// K-major SWIZZLE_128B descriptor for a start address that is only 128-byte (one row) aligned: the matrix base
// offset field carries the row phase (start >> 7) & 7 of the 1024-byte swizzle pattern.
__device__ __forceinline__ uint64_t umma_desc_sw128_rows(uint32_t saddr) {
    return umma_desc(saddr, 128) | ((uint64_t)((saddr >> 7) & 7u) << 49);
}
// tcgen05.mma with both operand descriptors given as {lo, hi} 32-bit halves (hi is shared by A and B)
__device__ __forceinline__ void umma_bf16_lohi(uint32_t tmem_d, uint32_t a_lo, uint32_t b_lo, uint32_t hi, uint32_t idesc,
                                               uint32_t accum) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\t"
        "setp.ne.b32 p, %5, 0;\n\t"
        "mov.b64 da, {%1, %3};\n\t"
        "mov.b64 db, {%2, %3};\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %4, p;\n\t}"
        ::"r"(tmem_d), "r"(a_lo), "r"(b_lo), "r"(hi), "r"(idesc), "r"(accum)
        : "memory");
}
// One lane of a CONVERGED warp.  The MMA warps run their whole loop with all 32 lanes and elect the lane only around the
// tcgen05 instructions: descriptors and barrier addresses are then warp-uniform values the compiler keeps in uniform
// registers.  Under `if (lane == 0)` it had to broadcast them register by register in a waterfall loop before every
// UTCHMMA (ELECT + 5 R2UR.BROADCAST + BRA.U.ANY), which is what cost ~100 cycles per instruction.
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
    return pred != 0;
}
// tcgen05.wait::ld that also ties the 32 destination registers to the wait, so that no read of them can be
// scheduled above it
__device__ __forceinline__ void tmem_ld_wait_dep32(uint32_t* r) {
    asm volatile("tcgen05.wait::ld.sync.aligned;" : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]), "+r"(r[8]), "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15]) :: "memory");
    asm volatile("" : "+r"(r[16]), "+r"(r[17]), "+r"(r[18]), "+r"(r[19]), "+r"(r[20]), "+r"(r[21]), "+r"(r[22]), "+r"(r[23]), "+r"(r[24]), "+r"(r[25]), "+r"(r[26]), "+r"(r[27]), "+r"(r[28]), "+r"(r[29]), "+r"(r[30]), "+r"(r[31]) :: "memory");
}
// the same for a 16-register load
__device__ __forceinline__ void tmem_ld_wait_dep16(uint32_t* r) {
    asm volatile("tcgen05.wait::ld.sync.aligned;" : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]), "+r"(r[8]), "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15]) :: "memory");
}
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap* map, uint32_t bar, int c0, int c1, int c2) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
        ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1), "r"(c2)
        : "memory");
}
// mbarrier wait whose try_wait carries a suspend-time hint: the thread sleeps in hardware until the phase completes
// (or the hint expires) instead of burning issue slots that the epilogue warps need
__device__ __forceinline__ void mbar_wait_hint(uint32_t bar, uint32_t parity) {
    for (uint32_t it = 0; it < (1u << 22); ++it) {
        uint32_t ok;
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(ok)
            : "r"(bar), "r"(parity), "r"(20000u)
            : "memory");
        if (ok) return;
    }
    __trap();
}
// plain (non-tensor) bulk copy global -> shared, completion on an mbarrier
__device__ __forceinline__ void bulk_load(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar)
                 : "memory");
}
// wait with back-off, for the warps that wait long (keeps their spinning out of the issue slots of the TMA / MMA threads)
__device__ __forceinline__ void mbar_wait_sleep(uint32_t bar, uint32_t parity) {
    for (uint32_t it = 0; it < (1u << 22); ++it) {
        if (mbar_try_wait(bar, parity)) return;
        __nanosleep(40);
    }
    __trap();
}
}  // namespace tc

// A CTA owns ONE N tile (nt = blockIdx.x % n_tiles: its weights can stay in shared memory) and walks the pixel tiles
// mt = blockIdx.x / n_tiles, + mstride, ...; mt = (img * tiles_y + ty) * tiles_x + tx is decomposed once, then advanced
// with carries (the stride's digits come precomputed in the parameters).
struct TileIter {
    int mt, tx, ty, img;
    __device__ __forceinline__ void init(const ConvTc2Params& p) {
        mt = (int)blockIdx.x / p.n_tiles;
        const int tpi = p.tiles_x * p.tiles_y;
        img = mt / tpi;
        const int trem = mt - img * tpi;
        ty = trem / p.tiles_x;
        tx = trem - ty * p.tiles_x;
    }
    __device__ __forceinline__ bool valid(const ConvTc2Params& p) const { return mt < p.m_tiles; }
    __device__ __forceinline__ void next(const ConvTc2Params& p) {
        mt += p.mstride;
        tx += p.step_tx; if (tx >= p.tiles_x) { tx -= p.tiles_x; ++ty; }
        ty += p.step_ty; if (ty >= p.tiles_y) { ty -= p.tiles_y; ++img; }
        img += p.step_img;
    }
};

// KSTEPS = CK / 16 (UMMA K = 16 bf16), HALO = k x k taps from one shared-memory patch, PROF = clock64 timeline of CTA 0,
// KK = kernel size known at compile time in HALO mode (3, 4, 5; 0 = read it from the parameters)
template <int KSTEPS, bool HALO, bool PROF, int KK, bool POOL = false, bool LRN = false>
__global__ void __launch_bounds__(kConvTc2Threads, 1) conv_tc2_kernel(const __grid_constant__ CUtensorMap tmap_a,
                                                                      const __grid_constant__ CUtensorMap tmap_b,
                                                                      const __grid_constant__ ConvTc2Params p) {
    using namespace tc;
    extern __shared__ __align__(1024) uint8_t smem_raw2[];
    const uint32_t smem_base = (smem_u32(smem_raw2) + 1023u) & ~1023u;
    const uint32_t span = (uint32_t)p.CK * 2u;
    // Shared memory: [A rings][B rings][bias staging][barriers].  In DUAL mode there are two of each ring: even pixel tiles
    // flow through ring 0 / accumulator 0 and are multiplied by MMA thread 0, odd ones through ring 1 / accumulator 1 by
    // MMA thread 1: narrow-N tiles are bound by the issuing thread's own instruction stream (barrier rounds, commits), and
    // two issuing threads overlap them.  Resident weights are one region shared by both.
    const int nr = p.dual ? 2 : 1, nrb = p.b_resident ? 1 : nr;
    const uint32_t a_base = smem_base;
    const uint32_t a_ring_bytes = (uint32_t)p.a_stages * p.a_stage_bytes, b_ring_bytes = (uint32_t)p.b_stages * p.b_stage_bytes;
    const uint32_t b_base = a_base + (uint32_t)nr * a_ring_bytes;
    const uint32_t raw_base = b_base + (uint32_t)nrb * b_ring_bytes;                      // POOL mode: raw halo patches
    const uint32_t bias_base = raw_base + (POOL ? (uint32_t)p.r_stages * p.r_stage_bytes : 0u);   // 64 floats per epilogue warp
    const uint32_t bar_base = bias_base + (uint32_t)kTcEpiWarps * 256u;
    const uint32_t afull0 = bar_base, aempty0 = afull0 + 8u * nr * p.a_stages;
    const uint32_t bfull0 = aempty0 + 8u * nr * p.a_stages, bempty0 = bfull0 + 8u * nrb * p.b_stages;
    const uint32_t tfull0 = bempty0 + 8u * nrb * p.b_stages, tempty0 = tfull0 + 32u;     // up to 4 accumulators
    const uint32_t tmem_slot = tempty0 + 32u;
    const uint32_t rfull0 = tmem_slot + 8u, rempty0 = rfull0 + 8u * (POOL ? p.r_stages : 0);
    // POOL mode: warps 4..11 pool, warps 12..19 drain the accumulators (TMEM lane quadrant = warp % 4 either way)
    constexpr int kPoolWarps = POOL ? 8 : 0;
    constexpr int kEpiFirst = kTcMmaWarp + 1 + kPoolWarps, NE = kTcEpiWarps - kPoolWarps;

    // warp index through a shuffle: provably warp-uniform for the compiler (role branches and everything derived from the
    // role stay in uniform registers)
    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
    const int my_nt = (int)blockIdx.x % p.n_tiles;
    const int taps = p.kh * p.kw;
    const int a_iters = HALO ? p.n_cchunks : taps * p.n_cchunks;

    if (threadIdx.x == 0) {
        for (int s = 0; s < nr * p.a_stages; ++s) { mbar_init(afull0 + 8u * s, POOL ? kPoolWarps : 1); mbar_init(aempty0 + 8u * s, 1); }
        for (int s = 0; s < nrb * p.b_stages; ++s) { mbar_init(bfull0 + 8u * s, 1); mbar_init(bempty0 + 8u * s, 1); }
        for (int a = 0; a < 4; ++a) { mbar_init(tfull0 + 8u * a, 1); mbar_init(tempty0 + 8u * a, NE / p.egroups); }
        if (POOL)
            for (int s = 0; s < p.r_stages; ++s) { mbar_init(rfull0 + 8u * s, 1); mbar_init(rempty0 + 8u * s, kPoolWarps); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    if (warp == kTcAllocWarp) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(tmem_slot), "r"(p.tmem_cols)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    fence_before();
    __syncthreads();
    fence_after();
    uint32_t tmem_base;
    asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tmem_base) : "r"(tmem_slot));
    tmem_base = __shfl_sync(0xffffffffu, tmem_base, 0);

    if (warp == 1 || warp == 2 || (warp == 3 && !p.dual)) {
        // ================= TMA producers (one thread per warp) ========================================================
        // DUAL: warp 1 feeds ring 0 (even tiles), warp 2 ring 1 (odd tiles), each with the A and B loads of its tiles in
        // the order the MMA thread consumes them.  Otherwise warp 1 feeds the A ring and warps 2, 3 the B ring (even / odd
        // rounds: one thread needs a few hundred cycles per wait + expect_tx + TMA round).
        {   // all 32 lanes, converged (waits by every lane); one elected lane issues the copies
            const int pw = warp - 1;
            const bool doA = p.dual || pw == 0, doB = p.dual || pw >= 1;
            // single mode, B ring with >= 2 stages: the two B warps take alternate rounds (a one-stage ring is fed by one
            // warp: two visits of a warp to the stage would be two barrier phases apart, which a parity wait cannot see)
            // (and the stage count must be even: then a warp always visits the same stages, one phase apart each time)
            const bool b_split = !p.dual && p.b_stages >= 2 && (p.b_stages & 1) == 0 && !p.b_resident;
            const uint32_t b_par = (uint32_t)(pw - 1) & 1u;
            const bool idle = !p.dual && pw == 2 && !b_split;
            uint32_t b_idx = 0;
            const int ring = p.dual ? pw : 0;
            const int first = p.dual ? pw : 0, step = p.dual ? 2 : 1;
            if (doA && lane == 0) asm volatile("prefetch.tensormap [%0];" ::"l"(&tmap_a) : "memory");
            if (doB && lane == 0) asm volatile("prefetch.tensormap [%0];" ::"l"(&tmap_b) : "memory");
            __syncwarp();
            const uint32_t b_tx = (uint32_t)p.BN * span * (uint32_t)p.b_taps;
            // (POOL mode: the "A" loads of this warp are the raw halo patches, into their own ring)
            const uint32_t a_ring = POOL ? raw_base : a_base + (uint32_t)ring * a_ring_bytes;
            const uint32_t b_ring = b_base + (p.b_resident ? 0u : (uint32_t)ring * b_ring_bytes);
            const uint32_t af = POOL ? rfull0 : afull0 + 8u * ring * p.a_stages, ae = POOL ? rempty0 : aempty0 + 8u * ring * p.a_stages;
            const uint32_t bf = bfull0 + (p.b_resident ? 0u : 8u * ring * p.b_stages);
            const uint32_t be = bempty0 + (p.b_resident ? 0u : 8u * ring * p.b_stages);
            const uint32_t pa_stage = POOL ? p.r_stage_bytes : p.a_stage_bytes, pa_tx = POOL ? p.r_tx_bytes : p.a_tx_bytes;
            const uint32_t pa_stages = (uint32_t)(POOL ? p.r_stages : p.a_stages);
            uint32_t sa = 0, pha = 0, sb = 0, phb = 0;
            const int n0 = my_nt * p.BN;
            TileIter ti;
            ti.init(p);
            if (first && ti.valid(p)) ti.next(p);
            const bool load_b_once = p.b_resident != 0;
            bool b_done = load_b_once && ring != 0;          // resident weights are loaded by ring 0's producer only
            for (int tile_i = first; ti.valid(p) && !idle; tile_i += step) {
                const int n_img = ti.img;
                const int oh0 = ti.ty * p.BH - p.pad_h, ow0 = ti.tx * p.BW - p.pad_w;
                const bool do_b = doB && !b_done;
                auto load_a = [&](int cc, int r, int s) {
                    mbar_wait_hint(ae + 8u * sa, pha ^ 1u);
                    __syncwarp();
                    const uint32_t dst = a_ring + sa * pa_stage;
                    if (PROF && POOL && lane == 0) FCNOD_PROF(0, tile_i, 16 + cc);      // raw patch: stage free, TMA issued
                    if (!elect_one()) {
                    } else if (HALO && p.a_bulk) {
                        mbar_expect_tx(af + 8u * sa, pa_tx);
                        // one plain bulk copy per patch row: PW pixels x `span` bytes, contiguous in the image
                        const uint32_t rb = (uint32_t)p.PW * span;
                        const char* src = static_cast<const char*>(p.a_src) + (size_t)n_img * p.a_img_pitch +
                                          (size_t)oh0 * p.a_row_pitch + (size_t)ow0 * span;
                        const int rows = p.BH + p.kh - 1;
                        for (int pr = 0; pr < rows; ++pr)
                            bulk_load(dst + (uint32_t)pr * rb, src + (size_t)pr * p.a_row_pitch, rb, af + 8u * sa);
                    } else {
                        mbar_expect_tx(af + 8u * sa, pa_tx);
                        tma_load_4d(dst, &tmap_a, af + 8u * sa, cc * p.CK, ow0 + s, oh0 + r, n_img);
                    }
                    __syncwarp();
                    if (++sa == pa_stages) { sa = 0; pha ^= 1u; }
                };
                auto load_b = [&](int kb) {      // kb = first K-block of the round, in the packed (= global) order
                    if (!b_split || (b_idx & 1u) == b_par) {
                        mbar_wait_hint(be + 8u * sb, phb ^ 1u);
                        __syncwarp();
                        if (PROF && lane == 0) FCNOD_PROF(0, tile_i, 0);
                        if (elect_one()) {
                            mbar_expect_tx(bf + 8u * sb, b_tx);
                            tma_load_3d(b_ring + sb * p.b_stage_bytes, &tmap_b, bf + 8u * sb, 0, n0, kb);
                        }
                        __syncwarp();
                    }
                    ++b_idx;
                    if (++sb == (uint32_t)p.b_stages) { sb = 0; phb ^= 1u; }
                };
                if (HALO) {
                    // per channel chunk: the patch, then the kh * kw / b_taps weight rounds that multiply it
                    // (weights are packed [(r * n_cchunks + cc) * kw + s]: the kw taps of a kernel row are adjacent)
                    for (int cc = 0; cc < p.n_cchunks; ++cc) {
                        if (doA) load_a(cc, 0, 0);
                        if (do_b)
                            for (int r = 0; r < p.kh; ++r)
                                for (int s = 0; s < p.kw; s += p.b_taps) load_b((r * p.n_cchunks + cc) * p.kw + s);
                    }
                } else {
                    int kb = 0;
                    for (int r = 0; r < p.kh; ++r)
                        for (int s = 0; s < p.kw; ++s)
                            for (int cc = 0; cc < p.n_cchunks; ++cc, ++kb) {
                                if (doA) load_a(cc, r, s);
                                if (do_b) load_b(kb);
                            }
                }
                if (load_b_once) b_done = true;
                ti.next(p);
                if (step == 2 && ti.valid(p)) ti.next(p);
            }
        }
    } else if (warp == 0 || (warp == 3 && p.dual)) {
        // ================= MMA issuers (one thread each; the second one only in DUAL mode) ============================
        // The issuing thread is latency-bound on its OWN instruction stream, so the tap loops are unrolled at compile
        // time (KK = kernel size) and a whole kernel row of taps shares one barrier round.
        const int mw = warp == 0 ? 0 : 1;
        if (mw == 0 || p.dual) {      // all 32 lanes, converged; one elected lane issues
            const int ring = mw;
            const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(p.BN >> 3) << 17) | (8u << 24);
            // descriptor = {lo: start address >> 4 | LBO 1 << 16, hi: SBO (8 rows) >> 4 | version 1 << 14 | swizzle << 29}
            const uint32_t desc_hi = ((8u * span) >> 4) | (1u << 14) | ((span == 128 ? 2u : (span == 64 ? 4u : 6u)) << 29);
            const uint32_t a_ring = a_base + (uint32_t)ring * a_ring_bytes;
            const uint32_t b_ring = b_base + (p.b_resident ? 0u : (uint32_t)ring * b_ring_bytes);
            const uint32_t af = afull0 + 8u * ring * p.a_stages, ae = aempty0 + 8u * ring * p.a_stages;
            const uint32_t bf = bfull0 + (p.b_resident ? 0u : 8u * ring * p.b_stages);
            const uint32_t be = bempty0 + (p.b_resident ? 0u : 8u * ring * p.b_stages);
            const uint32_t a_lo_base = ((a_ring & 0x3FFFFu) >> 4) | (1u << 16), a_lo_step = p.a_stage_bytes >> 4;
            const uint32_t b_lo_base = ((b_ring & 0x3FFFFu) >> 4) | (1u << 16), b_lo_step = p.b_stage_bytes >> 4;
            const uint32_t b_tap_step = ((uint32_t)p.BN * span) >> 4;      // one tap tile inside a multi-tap B stage
            const uint32_t px_step = span >> 4;                  // HALO: one patch pixel = `span` bytes, in 16-byte units
            const uint32_t row_step = (uint32_t)p.PW * px_step;  // HALO: one patch row = PW pixels
            const int kh = HALO ? (KK ? KK : p.kh) : 1, kw = HALO ? (KK ? KK : p.kw) : 1;
            const bool bt_row = p.b_taps > 1;                    // a B stage holds the kw taps of one kernel row
            const bool resident = p.b_resident != 0;
            const int it_step = p.dual ? 2 : 1;
            uint32_t sa = 0, pa = 0, sb = 0, pb = 0;
            uint32_t a_lo = a_lo_base, b_lo = b_lo_base;
            int it = mw;
            for (int mt = (int)blockIdx.x / p.n_tiles + mw * p.mstride; mt < p.m_tiles; mt += it_step * p.mstride, it += it_step) {
                const uint32_t acc = (uint32_t)it & ((1u << p.acc_lg) - 1u), aphase = ((uint32_t)it >> p.acc_lg) & 1u;
                const int itp = it < 8 ? it : 8;
                if (PROF && lane == 0) FCNOD_PROF(1, itp, 61);
                mbar_wait_hint(tempty0 + 8u * acc, aphase ^ 1u);     // the epilogue has drained this accumulator
                __syncwarp();
                fence_after();
                if (PROF && lane == 0) FCNOD_PROF(1, itp, 62);
                const uint32_t tmem_d = tmem_base + acc * p.acc_stride;
                const bool wait_b = !(resident && it > mw);
                uint32_t accum = 0;
                int ev = 0;
                int cc = 0;
                // all K steps of one tap: A window a_tap x B tile b_tap
                auto issue_tap = [&](uint32_t a_tap, uint32_t b_tap, int ks) {
#pragma unroll
                    for (int k = 0; k < KSTEPS; ++k) {
                        if (KSTEPS == 1 || k < ks) {
                            if (elect_one()) umma_bf16_lohi(tmem_d, a_tap + 2u * k, b_tap + 2u * k, desc_hi, idesc, accum);
                            accum = 1;
                        }
                    }
                };
                auto b_wait = [&]() {
                    if (wait_b) {
                        mbar_wait_hint(bf + 8u * sb, pb);
                        __syncwarp();
                        fence_after();
                        if (PROF && lane == 0) FCNOD_PROF(1, itp, ev);
                    }
                };
                auto b_release = [&]() {
                    if (!resident && elect_one()) umma_commit(be + 8u * sb);
                    if (PROF) { if (lane == 0) FCNOD_PROF(3, itp, ev); ++ev; }
                    b_lo += b_lo_step;
                    if (++sb == (uint32_t)p.b_stages) { sb = 0; pb ^= 1u; b_lo = b_lo_base; }
                };
                for (int ai = 0; ai < a_iters; ++ai) {
                    mbar_wait_hint(af + 8u * sa, pa);
                    __syncwarp();
                    const int ks = (cc == p.n_cchunks - 1) ? p.ksteps_last : KSTEPS;   // skip all-zero K steps
                    if (++cc == p.n_cchunks) cc = 0;
                    if (!HALO) {
                        b_wait();
                        issue_tap(a_lo, b_lo, ks);
                        b_release();
                    } else {
                        uint32_t a_row = a_lo;
#pragma unroll
                        for (int r = 0; r < (KK ? KK : 8); ++r) {
                            if (!KK && r >= kh) break;
                            if (bt_row) {
                                b_wait();
#pragma unroll
                                for (int s = 0; s < (KK ? KK : 8); ++s) {
                                    if (!KK && s >= kw) break;
                                    issue_tap(a_row + (uint32_t)s * px_step, b_lo + (uint32_t)s * b_tap_step, ks);
                                }
                                b_release();
                            } else {
#pragma unroll
                                for (int s = 0; s < (KK ? KK : 8); ++s) {
                                    if (!KK && s >= kw) break;
                                    b_wait();
                                    issue_tap(a_row + (uint32_t)s * px_step, b_lo, ks);
                                    b_release();
                                }
                            }
                            a_row += row_step;
                        }
                    }
                    if (elect_one()) umma_commit(ae + 8u * sa);
                    a_lo += a_lo_step;
                    if (++sa == (uint32_t)p.a_stages) { sa = 0; pa ^= 1u; a_lo = a_lo_base; }
                }
                if (elect_one()) umma_commit(tfull0 + 8u * acc);
                if (PROF && lane == 0) FCNOD_PROF(1, itp, 63);
            }
        }
    } else if (POOL && warp < kEpiFirst) {
        // ================= POOL mode: 3x3 / stride 1 / pad 1 max-pool between the raw patch ring and the A ring ==========
        // The kernel is shared-memory-bandwidth bound in this mode (raw patch in + out, pooled tile in + out, weights in +
        // out: ~110 KB per 64-channel chunk at 128 B/clk), so every raw byte is read exactly ONCE: warp w owns the 16-byte
        // channel chunk w of the 128-byte pixel rows, lane = patch column.  A lane marches down its column keeping the last
        // two rows (vertical 3-max in registers); the horizontal 3-max takes the two right-hand neighbours by shuffle, once
        // per OUTPUT row.  32 lanes load 32 columns and produce 30: wider boxes take ceil(BW / 30) passes.  Consecutive
        // lanes touch consecutive pixel rows, whose swizzle phases differ: loads and stores are bank-conflict free.
        const int pt = (int)threadIdx.x - 32 * (kTcMmaWarp + 1);
        const uint32_t j = (uint32_t)(pt >> 5);
        const int RW = p.BW + 2, rows = p.BH + 2;
        auto lds128 = [](uint32_t a) {
            uint4 v;
            asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a) : "memory");
            return v;
        };
        auto max3w = [](uint32_t a, uint32_t b, uint32_t c) {
            const __nv_bfloat162 r = __hmax2(__hmax2(*reinterpret_cast<const __nv_bfloat162*>(&a), *reinterpret_cast<const __nv_bfloat162*>(&b)),
                                             *reinterpret_cast<const __nv_bfloat162*>(&c));
            return *reinterpret_cast<const uint32_t*>(&r);
        };
        auto max3 = [&](const uint4& a, const uint4& b, const uint4& c) {
            return make_uint4(max3w(a.x, b.x, c.x), max3w(a.y, b.y, c.y), max3w(a.z, b.z, c.z), max3w(a.w, b.w, c.w));
        };
        auto down = [](const uint4& v, int d) {
            return make_uint4(__shfl_down_sync(0xffffffffu, v.x, d), __shfl_down_sync(0xffffffffu, v.y, d),
                              __shfl_down_sync(0xffffffffu, v.z, d), __shfl_down_sync(0xffffffffu, v.w, d));
        };
        uint32_t sr = 0, phr = 0, sa = 0, pha = 0;
        TileIter ti;
        ti.init(p);
        const bool stamp = PROF && pt == 0;
        (void)pt;
        // Fast path (boxes up to 30 x 4, i.e. every launch of the deploy net): one pass, and the swizzled offsets of the six
        // loads and four stores of a lane are the same for every chunk of every tile -- computed once.  Lanes / rows outside
        // the patch read a clamped address (their results are never stored).
        const bool fast = p.BW <= 30 && p.BH <= 4;
        uint32_t loff[6], soff[4];
        bool sp[4];
#pragma unroll
        for (int k = 0; k < 6; ++k) {
            const uint32_t r0 = (uint32_t)(min(k, rows - 1) * RW + min(lane, RW - 1));
            loff[k] = r0 * 128u + ((j ^ (r0 & 7u)) << 4);
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const uint32_t m = (uint32_t)(k * p.BW + lane);
            soff[k] = m * 128u + ((j ^ (m & 7u)) << 4);
            sp[k] = lane < p.BW && k < p.BH;
        }
        for (int tile_i = 0; ti.valid(p); ti.next(p), ++tile_i) {
            for (int cc = 0; cc < p.n_cchunks; ++cc) {
                if (stamp && cc < 14) FCNOD_PROF(2, tile_i, 8 + 4 * cc);
                mbar_wait(rfull0 + 8u * sr, phr);                  // the raw patch has landed
                if (stamp && cc < 14) FCNOD_PROF(2, tile_i, 9 + 4 * cc);
                mbar_wait(aempty0 + 8u * sa, pha ^ 1u);            // the MMA has finished with this A stage
                if (stamp && cc < 14) FCNOD_PROF(2, tile_i, 10 + 4 * cc);
                const uint32_t raw = raw_base + sr * p.r_stage_bytes, dst = a_base + sa * p.a_stage_bytes;
                if (fast) {
                    uint4 a[6];
#pragma unroll
                    for (int k = 0; k < 6; ++k) a[k] = lds128(raw + loff[k]);
                    uint4 v[4], h1[4], h2[4];
#pragma unroll
                    for (int k = 0; k < 4; ++k) v[k] = max3(a[k], a[k + 1], a[k + 2]);       // own column, rows k .. k+2
#pragma unroll
                    for (int k = 0; k < 4; ++k) { h1[k] = down(v[k], 1); h2[k] = down(v[k], 2); }   // 32 shuffles in flight
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const uint4 o = max3(v[k], h1[k], h2[k]);
                        if (sp[k])
                            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};"
                                         ::"r"(dst + soff[k]), "r"(o.x), "r"(o.y), "r"(o.z), "r"(o.w) : "memory");
                    }
                } else
                for (int seg = 0; seg < p.BW; seg += 30) {
                    const int col = seg + lane;                       // patch column loaded, and output column produced
                    const bool ld = col < RW, st = lane < 30 && col < p.BW;
                    // four output rows per pass: their six input rows are loaded back to back (one latency, not six), then
                    // the vertical maxima, the shuffles and the stores of the four rows overlap
                    for (int ob = 0; ob < p.BH; ob += 4) {
                        uint4 a[6];
#pragma unroll
                        for (int k = 0; k < 6; ++k) {
                            const uint32_t r0 = (uint32_t)((ob + k) * RW + col);
                            a[k] = make_uint4(0, 0, 0, 0);
                            if (ld && ob + k < rows) a[k] = lds128(raw + r0 * 128u + ((j ^ (r0 & 7u)) << 4));
                        }
#pragma unroll
                        for (int k = 0; k < 4; ++k) {
                            const uint4 v = max3(a[k], a[k + 1], a[k + 2]);    // own column, rows ob+k .. ob+k+2
                            const uint4 o = max3(v, down(v, 1), down(v, 2));
                            const uint32_t m = (uint32_t)((ob + k) * p.BW + col);
                            if (st && ob + k < p.BH)
                                asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};"
                                             ::"r"(dst + m * 128u + ((j ^ (m & 7u)) << 4)), "r"(o.x), "r"(o.y), "r"(o.z), "r"(o.w) : "memory");
                        }
                    }
                }
                if (stamp && cc < 14) FCNOD_PROF(2, tile_i, 11 + 4 * cc);
                // generic-proxy stores -> visible to the tensor core's async-proxy reads, then hand both stages over
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                __syncwarp();
                if (lane == 0) {
                    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(afull0 + 8u * sa) : "memory");
                    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(rempty0 + 8u * sr) : "memory");
                }
                if (++sr == (uint32_t)p.r_stages) { sr = 0; phr ^= 1u; }
                if (++sa == (uint32_t)p.a_stages) { sa = 0; pha ^= 1u; }
            }
        }
    } else {
        // ================= epilogue: 16 warps (8 in POOL mode); warp w drains TMEM lane quadrant w % 4 ==================
        // Everything that does not change from tile to tile is computed once: the lane's position inside the pixel box,
        // the bias of the warp's columns (staged in shared memory) and the destination of each 16-column group.
        const int ew = warp - kEpiFirst;
        const int q = warp & 3;
        const int egrp = (ew >> 2) & (int)(p.egroups - 1u);          // which tiles: it % egroups == egrp
        const int part = (ew >> 2) / (int)p.egroups;                 // which columns of them
        const int HW = p.H * p.W;
        // this warp's columns inside the N tile: [wc0, wc0 + wcols), 16-column granular, at most 64
        const int wc0 = min(part * p.cols_half0, p.BN);
        const int wcols = min(p.cols_half0, p.BN - wc0);
        const uint32_t my_bias = bias_base + (uint32_t)ew * 256u;                  // 64 floats
        uint8_t* const smem_gen = smem_raw2 + (smem_base - smem_u32(smem_raw2));    // generic view of smem_base
        float* const bias_s = reinterpret_cast<float*>(smem_gen + (my_bias - smem_base));
        const bool relu = (p.flags & SEG_RELU) != 0;
        const int m = q * 32 + lane;                // accumulator row
        const int py = m / p.PW, px = m - py * p.PW;
        const bool vpos = (py < p.BH) && (px < p.BW);
        const int ncc = p.bt_lo_w + p.bt_hi_w + 1;
        const int n0 = my_nt * p.BN + wc0;          // first fused-N column of this warp
        for (int j = lane; j < wcols; j += 32) bias_s[j] = __ldg(p.bias + n0 + j);
        // LRN epilogue: bias of the two columns either side of this warp's range (the neighbouring warps' columns)
        float lrn_bl0 = 0.f, lrn_bl1 = 0.f, lrn_br0 = 0.f, lrn_br1 = 0.f;
        if (LRN && wcols > 0) {
            if (wc0 > 0) { lrn_bl0 = __ldg(p.bias + n0 - 2); lrn_bl1 = __ldg(p.bias + n0 - 1); }
            if (wc0 + wcols < p.BN) { lrn_br0 = __ldg(p.bias + n0 + wcols); lrn_br1 = __ldg(p.bias + n0 + wcols + 1); }
        }
        (void)lrn_bl0; (void)lrn_bl1; (void)lrn_br0; (void)lrn_br1;
        // destination of the j-th 16-column group (bf16 NHWC): base pointer for pixel 0 of image 0, pixel pitch, and how
        // it may be stored: 0 = nothing, 1 / 2 = one / two 16-byte pieces, 3 = one 32-byte store (aligned)
        bf16* grp_ptr[4];
        int grp_ld[4], grp_mode[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            grp_ptr[j] = nullptr; grp_ld[j] = 0; grp_mode[j] = 0;
            if (!(p.flags & SEG_DIRECT) && 16 * j < wcols) {
                const int col = n0 + 16 * j;
                int si = 0;
#pragma unroll
                for (int s = 1; s < 4; ++s)
                    if (s < p.nseg && col >= p.seg[s].col0) si = s;
                const int cl = col - p.seg[si].col0, left = p.seg[si].ncols - cl;   // segment boundaries: multiples of 8
                if (left > 0) {
                    grp_ptr[j] = static_cast<bf16*>(p.seg[si].out) + cl;
                    grp_ld[j] = p.seg[si].ldo;
                    grp_mode[j] = left >= 16 ? 2 : 1;
                    if (grp_mode[j] == 2 && ((reinterpret_cast<uintptr_t>(grp_ptr[j]) | (uintptr_t)(2 * grp_ld[j])) & 31u) == 0) grp_mode[j] = 3;
                }
            }
        }
        __syncwarp();
        int it = egrp;
        TileIter ti;
        ti.init(p);
        for (int k = 0; k < egrp; ++k) ti.next(p);
        for (; ti.valid(p); it += (int)p.egroups) {
            const uint32_t acc = (uint32_t)it & ((1u << p.acc_lg) - 1u), aphase = ((uint32_t)it >> p.acc_lg) & 1u;
            const int n_img = ti.img;
            const int oh = ti.ty * p.BH + py, ow = ti.tx * p.BW + px;
            const bool valid = vpos && (oh < p.H) && (ow < p.W);
            const int pix_in_img = valid ? oh * p.W + ow : -1;
            const float* gbias = nullptr;               // border pixels of the first convolution: per-class bias
            if (p.bias_tab && valid) {
                const int cls = border_class(oh, p.H, p.bt_lo_h, p.bt_hi_h) * ncc + border_class(ow, p.W, p.bt_lo_w, p.bt_hi_w);
                if (cls != p.bt_interior) gbias = p.bias_tab + (size_t)cls * p.bt_ld + n0;
            }
            if (PROF && ew == 0 && lane == 0) FCNOD_PROF(2, it, 0);
            mbar_wait_sleep(tfull0 + 8u * acc, aphase);
            fence_after();
            if (PROF && ew == 0 && lane == 0) FCNOD_PROF(2, it, 1);
            const uint32_t trow = tmem_base + acc * p.acc_stride + ((uint32_t)(q * 32) << 16) + (uint32_t)wc0;
            if (p.flags & SEG_DIRECT) {
                // ---- fp32 heads: NCHW float32 straight from registers (lanes = consecutive pixels); tiny, kept simple
                for (int c = 0; c < wcols; c += 16) {
                    uint32_t r[16];
                    tmem_ld16(trow + (uint32_t)c, r);
                    tmem_ld_wait();
#pragma unroll
                    for (int g = 0; g < 2; ++g) {
                        const int colh = n0 + c + 8 * g;
                        int si = 0;
#pragma unroll
                        for (int s = 1; s < 4; ++s)
                            if (s < p.nseg && colh >= p.seg[s].col0) si = s;
                        if (valid) {
                            const bool sig = (p.seg[si].flags & SEG_SIGMOID) != 0;
                            float* o = static_cast<float*>(p.seg[si].out) + (size_t)n_img * p.seg[si].ldo * HW + pix_in_img;
                            const int cl = colh - p.seg[si].col0, nc = p.seg[si].ncols;
#pragma unroll
                            for (int j = 0; j < 8; ++j) {
                                float x = __uint_as_float(r[8 * g + j]) + bias_s[c + 8 * g + j];
                                if (relu) x = fmaxf(x, 0.f);
                                if (cl + j < nc) o[(size_t)(cl + j) * HW] = sig ? 1.f / (1.f + expf(-x)) : x;
                            }
                        }
                    }
                }
                fence_before();
                __syncwarp();
                if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(tempty0 + 8u * acc) : "memory");
            } else if (LRN) {
                // ---- ReLU + LRN(5) + bf16 NHWC.  A warp owns up to 48 columns (three 16-column blocks).  All of them, plus
                //      the two columns either side (the neighbouring warps' blocks; zeros outside the channel range), leave
                //      TMEM first and the accumulator is handed back BEFORE the arithmetic: the tensor pipe never waits for
                //      the MUFU work.  y = relu(acc + bias); out = y * (k + a/n * sum_{+-2} y^2)^-0.75.
                const size_t pix_off = (size_t)n_img * HW + (size_t)(valid ? pix_in_img : 0);
                float y[48], yl0 = 0.f, yl1 = 0.f, yr0 = 0.f, yr1 = 0.f;
                uint32_t r[16];
                if (wcols > 0 && wc0 > 0) {
                    tmem_ld16(trow - 16u, r);
                    tmem_ld_wait_dep16(r);
                    yl0 = fmaxf(__uint_as_float(r[14]) + lrn_bl0, 0.f); yl1 = fmaxf(__uint_as_float(r[15]) + lrn_bl1, 0.f);
                }
#pragma unroll
                for (int b = 0; b < 3; ++b) {
                    if (16 * b < wcols) {
                        tmem_ld16(trow + 16u * b, r);
                        tmem_ld_wait_dep16(r);
#pragma unroll
                        for (int j = 0; j < 16; ++j) y[16 * b + j] = fmaxf(__uint_as_float(r[j]) + bias_s[16 * b + j], 0.f);
                    }
                }
                if (wcols > 0 && wc0 + wcols < p.BN) {
                    tmem_ld16(trow + (uint32_t)wcols, r);
                    tmem_ld_wait_dep16(r);
                    yr0 = fmaxf(__uint_as_float(r[0]) + lrn_br0, 0.f); yr1 = fmaxf(__uint_as_float(r[1]) + lrn_br1, 0.f);
                }
                fence_before();
                __syncwarp();
                if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(tempty0 + 8u * acc) : "memory");
#pragma unroll
                for (int b = 0; b < 3; ++b) {
                    if (16 * b < wcols) {
                        const bool last = 16 * (b + 1) >= wcols;
                        float sq[20];
                        sq[0] = b == 0 ? yl0 : y[b == 0 ? 0 : 16 * b - 2]; sq[1] = b == 0 ? yl1 : y[b == 0 ? 0 : 16 * b - 1];
                        sq[18] = (b == 2 || last) ? yr0 : y[b == 2 ? 0 : 16 * b + 16]; sq[19] = (b == 2 || last) ? yr1 : y[b == 2 ? 0 : 16 * b + 17];
#pragma unroll
                        for (int j = 0; j < 16; ++j) sq[2 + j] = y[16 * b + j];
#pragma unroll
                        for (int j = 0; j < 20; ++j) sq[j] *= sq[j];
                        float pr[19];                                 // pair sums: 3 adds per window instead of 4
#pragma unroll
                        for (int j = 0; j < 19; ++j) pr[j] = sq[j] + sq[j + 1];
                        uint32_t pk[8];
#pragma unroll
                        for (int e = 0; e < 8; ++e) {
                            float o[2];
#pragma unroll
                            for (int h = 0; h < 2; ++h) {
                                const int j = 2 * e + h;
                                const float ssum = (pr[j] + pr[j + 2]) + sq[j + 4];
                                o[h] = y[16 * b + j] * lrn_scale<true>(fmaf(p.lrn_aon, ssum, p.lrn_k), 0.75f);
                            }
                            const __nv_bfloat162 h2 = __floats2bfloat162_rn(o[0], o[1]);
                            pk[e] = *reinterpret_cast<const uint32_t*>(&h2);
                        }
                        if (valid && grp_mode[b] != 0) {
                            bf16* const d = grp_ptr[b] + pix_off * (size_t)grp_ld[b];
                            if (grp_mode[b] == 3) {
                                asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};"
                                             ::"l"(d), "r"(pk[0]), "r"(pk[1]), "r"(pk[2]), "r"(pk[3]), "r"(pk[4]), "r"(pk[5]),
                                               "r"(pk[6]), "r"(pk[7]) : "memory");
                            } else {
                                *reinterpret_cast<uint4*>(d) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
                                if (grp_mode[b] == 2) *reinterpret_cast<uint4*>(d + 8) = make_uint4(pk[4], pk[5], pk[6], pk[7]);
                            }
                        }
                    }
                }
            } else {
                // ---- bf16 NHWC: drain 32 columns at a time: TMEM -> registers, + bias (the border pixels of the first
                //      convolution take theirs from the class table), ReLU, pack, and store straight from the registers: a
                //      lane owns one pixel, a 16-column group is one full 32-byte sector of it (no shared-memory staging).
                const bool any_g = __any_sync(0xffffffffu, gbias != nullptr);
                const float4* const bias4 = reinterpret_cast<const float4*>(bias_s);
                const size_t pix_off = (size_t)n_img * HW + (size_t)(valid ? pix_in_img : 0);
                if (wcols <= 0) {      // a narrow N tile leaves this warp without columns: it only signs the accumulator off
                    fence_before();
                    __syncwarp();
                    if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(tempty0 + 8u * acc) : "memory");
                }
#pragma unroll
                for (int c = 0; c < 64; c += 32) {
                    if (c < wcols) {
                        const int ncol = min(32, wcols - c);      // 32 or 16
                        uint32_t ra[32];
                        if (ncol == 32) tmem_ld32(trow + (uint32_t)c, ra);
                        else tmem_ld16(trow + (uint32_t)c, ra);
                        tmem_ld_wait_dep32(ra);
                        if (c + 32 >= wcols) {
                            // the accumulator has left TMEM: hand the buffer back to the MMA warp
                            fence_before();
                            __syncwarp();
                            if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(tempty0 + 8u * acc) : "memory");
                            if (PROF && ew == 0 && lane == 0) FCNOD_PROF(2, it, 2);
                        }
#pragma unroll
                        for (int h = 0; h < 2; ++h) {
                            if (16 * h < ncol) {
                                const int j = (c >> 4) + h;
                                uint32_t pk[8];
#pragma unroll
                                for (int g = 0; g < 2; ++g) {
                                    const int cb = c + 16 * h + 8 * g;
                                    float4 b0 = bias4[cb >> 2], b1 = bias4[(cb >> 2) + 1];
                                    if (any_g && gbias) {
                                        b0 = __ldg(reinterpret_cast<const float4*>(gbias + cb));
                                        b1 = __ldg(reinterpret_cast<const float4*>(gbias + cb + 4));
                                    }
                                    const uint32_t* r8 = ra + 16 * h + 8 * g;
                                    float v[8];
                                    v[0] = __uint_as_float(r8[0]) + b0.x; v[1] = __uint_as_float(r8[1]) + b0.y;
                                    v[2] = __uint_as_float(r8[2]) + b0.z; v[3] = __uint_as_float(r8[3]) + b0.w;
                                    v[4] = __uint_as_float(r8[4]) + b1.x; v[5] = __uint_as_float(r8[5]) + b1.y;
                                    v[6] = __uint_as_float(r8[6]) + b1.z; v[7] = __uint_as_float(r8[7]) + b1.w;
                                    if (relu) {
#pragma unroll
                                        for (int e = 0; e < 8; ++e) v[e] = fmaxf(v[e], 0.f);
                                    }
#pragma unroll
                                    for (int e = 0; e < 4; ++e) {
                                        const __nv_bfloat162 h2 = __floats2bfloat162_rn(v[2 * e], v[2 * e + 1]);
                                        pk[4 * g + e] = *reinterpret_cast<const uint32_t*>(&h2);
                                    }
                                }
                                if (valid && grp_mode[j] != 0) {
                                    bf16* const d = grp_ptr[j] + pix_off * (size_t)grp_ld[j];
                                    if (grp_mode[j] == 3) {
                                        asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};"
                                                     ::"l"(d), "r"(pk[0]), "r"(pk[1]), "r"(pk[2]), "r"(pk[3]), "r"(pk[4]), "r"(pk[5]),
                                                       "r"(pk[6]), "r"(pk[7]) : "memory");
                                    } else {
                                        *reinterpret_cast<uint4*>(d) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
                                        if (grp_mode[j] == 2) *reinterpret_cast<uint4*>(d + 8) = make_uint4(pk[4], pk[5], pk[6], pk[7]);
                                    }
                                }
                            }
                        }
                    }
                }
            }
            __syncwarp();
            if (PROF && ew == 0 && lane == 0) FCNOD_PROF(2, it, 3);
            for (uint32_t k = 0; k < p.egroups; ++k) ti.next(p);
        }
    }
    __syncthreads();
    if (warp == kTcAllocWarp) {
        fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(p.tmem_cols) : "memory");
    }
}

}  // namespace fcnod
