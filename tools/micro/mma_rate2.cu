// Micro-benchmark 2: cycles per tcgen05.mma (kind::f16, M=128, K=16) issued the way conv_tc2.cuh issues it: a converged
// warp, one elected lane around the instruction, descriptors as {lo, hi} halves in warp-uniform registers.
// Variants: N, swizzle width of the operands (128 / 64 / 32 bytes), 1 or 2 issuing warps, and elect per MMA vs per 4.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mma_rate2 mma_rate2.cu && ./mma_rate2
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
    return pred != 0;
}
__device__ __forceinline__ void umma(uint32_t tmem_d, uint32_t a_lo, uint32_t b_lo, uint32_t hi, uint32_t idesc, uint32_t accum) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t.reg .b64 da, db;\n\tsetp.ne.b32 p, %5, 0;\n\tmov.b64 da, {%1, %3};\n\tmov.b64 db, {%2, %3};\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %4, p;\n\t}"
        ::"r"(tmem_d), "r"(a_lo), "r"(b_lo), "r"(hi), "r"(idesc), "r"(accum) : "memory");
}
// mode bits: 0-1: swizzle (0 = 128 B, 1 = 64 B, 2 = 32 B); bit 2: two issuing warps; bit 3: one elect per 4 MMAs
__global__ void __launch_bounds__(128, 1) k(int N, int mode, int iters, long long* out) {
    extern __shared__ __align__(1024) uint8_t sm[];
    const uint32_t base = (smem_u32(sm) + 1023u) & ~1023u;
    const uint32_t a = base, b = base + 65536, bar = b + 131072, slot = bar + 64;
    for (int i = threadIdx.x; i < (65536 + 131072) / 4; i += blockDim.x) ((uint32_t*)(sm + (base - smem_u32(sm))))[i] = 0;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar + 8));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    if (threadIdx.x < 32) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(slot) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    uint32_t tm;
    asm volatile("ld.shared.b32 %0, [%1];" : "=r"(tm) : "r"(slot));
    tm = __shfl_sync(0xffffffffu, tm, 0);
    const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);
    const bool dual = (mode & 4) != 0, batch4 = (mode & 8) != 0;
    const int sw = mode & 3;
    const uint32_t span = sw == 0 ? 128u : (sw == 1 ? 64u : 32u);
    if (warp == 0 || (dual && warp == 1)) {
        const uint32_t d = tm + (uint32_t)warp * 256u;
        const uint32_t bar_w = bar + 8u * warp;
        const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | (8u << 24);
        const uint32_t hi = ((8u * span) >> 4) | (1u << 14) | ((span == 128 ? 2u : (span == 64 ? 4u : 6u)) << 29);
        const uint32_t a_lo0 = (((a + warp * 32768u) & 0x3FFFFu) >> 4) | (1u << 16), b_lo0 = (((b + warp * 65536u) & 0x3FFFFu) >> 4) | (1u << 16);
        const uint32_t kstep = span == 32 ? 0u : 2u;             // 32-byte rows hold one K step: stay on the same tile
        long long t0 = clock64();
        for (int it = 0; it < iters; ++it) {
            const uint32_t st = (uint32_t)(it & 3) * (16384u >> 4);
            if (batch4) {
                if (elect_one()) {
#pragma unroll
                    for (int ks = 0; ks < 4; ++ks) umma(d, a_lo0 + st + kstep * (ks & (span == 128 ? 3 : (span == 64 ? 1 : 0))), b_lo0 + st + kstep * (ks & (span == 128 ? 3 : (span == 64 ? 1 : 0))), hi, idesc, 1u);
                }
            } else {
#pragma unroll
                for (int ks = 0; ks < 4; ++ks)
                    if (elect_one()) umma(d, a_lo0 + st + kstep * (ks & (span == 128 ? 3 : (span == 64 ? 1 : 0))), b_lo0 + st + kstep * (ks & (span == 128 ? 3 : (span == 64 ? 1 : 0))), hi, idesc, 1u);
            }
            __syncwarp();
        }
        if (elect_one()) asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar_w) : "memory");
        uint32_t ok = 0;
        while (!ok) {
            asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                         : "=r"(ok) : "r"(bar_w), "r"(0u) : "memory");
        }
        long long t1 = clock64();
        if (blockIdx.x == 0 && warp == 0 && (threadIdx.x & 31) == 0) out[0] = t1 - t0;
    }
    __syncthreads();
    if (threadIdx.x < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tm) : "memory");
}
int main() {
    long long* d; cudaMalloc(&d, 8);
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    const int iters = 2000;
    for (int mode = 0; mode < 16; ++mode) {
        if ((mode & 3) == 3) continue;
        for (int N : {32, 64, 128, 256}) {
            if ((mode & 4) && N > 256) continue;
            k<<<148, 128, 198 * 1024>>>(N, mode, iters, d);
            cudaError_t e = cudaDeviceSynchronize();
            long long c = 0; cudaMemcpy(&c, d, 8, cudaMemcpyDeviceToHost);
            printf("swizzle %3d B  %s  %s  N=%3d  %7.1f cycles/MMA/warp  (tensor N/2 = %3d, smem (128+N)/4 = %3d)  %s\n", (mode & 3) == 0 ? 128 : ((mode & 3) == 1 ? 64 : 32),
                   (mode & 4) ? "2 warps" : "1 warp ", (mode & 8) ? "elect/4" : "elect/1", N, (double)c / (iters * 4), N / 2, (128 + N) / 4,
                   e == cudaSuccess ? "" : cudaGetErrorString(e));
            if (e != cudaSuccess) return 1;
        }
    }
    return 0;
}
