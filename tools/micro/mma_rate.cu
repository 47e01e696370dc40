// Micro-benchmark: cycles per tcgen05.mma (kind::f16, M=128, K=16, cta_group::1) as a function of N, operands in shared
// memory (SS) or A in tensor memory (TS), accumulating into one accumulator or alternating between two.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mma_rate mma_rate.cu && ./mma_rate
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t desc128(uint32_t saddr) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3FFFF) >> 4);
    d |= (uint64_t)1 << 16;
    d |= (uint64_t)(1024 >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}
__device__ __forceinline__ uint64_t desc32(uint32_t saddr) {      // K-major SWIZZLE_32B: 32-byte rows, 8-row groups 256 bytes apart
    uint64_t d = 0;
    d |= (uint64_t)((saddr & 0x3FFFF) >> 4);
    d |= (uint64_t)1 << 16;
    d |= (uint64_t)(256 >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)6 << 61;
    return d;
}
__global__ void __launch_bounds__(128, 1) k(int N, int mode, int iters, long long* out) {
    extern __shared__ __align__(1024) uint8_t sm[];
    const uint32_t base = (smem_u32(sm) + 1023u) & ~1023u;
    const uint32_t a = base, b = base + 4 * 16384, bar = b + 4 * 32768, slot = bar + 16;
    for (int i = threadIdx.x; i < (4 * 16384 + 4 * 32768) / 4; i += blockDim.x) ((uint32_t*)(sm + (base - smem_u32(sm))))[i] = 0;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
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
    const bool dual = (mode & 32) != 0;
    if (threadIdx.x == 0 || (dual && threadIdx.x == 32)) {
        const uint32_t w = threadIdx.x >> 5;
        tm += w * 256u;
        const uint32_t bar_w = bar + (dual ? 32u * w : 0u);
        if (w) { asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_w)); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
        const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | (8u << 24);
        const bool alt = (mode & 1) && !dual, ts = mode & 2, distinctA = mode & 4;
        long long t0 = clock64();
        for (int it = 0; it < iters; ++it) {
            const uint32_t d = tm + ((alt && (it & 1)) ? 256u : 0u);
            const uint32_t stage = distinctA ? (uint32_t)(it & 3) : 0u;
#pragma unroll
            for (int ks = 0; ks < 4; ++ks) {
                uint64_t da = desc128(a + stage * 16384u) + 2 * ks, db = desc128(b + stage * 32768u) + 2 * ks;
                if (mode & 8) {      // 32-byte-swizzled operands, A start shifted by ks pixels (the first convolution's taps)
                    da = desc32(a + stage * 16384u + ((mode & 16) ? 0u : 32u * ks));
                    db = desc32(b + (uint32_t)ks * (uint32_t)N * 32u);
                }
                if (ts) {
                    const uint32_t ta = tm + 448u + 8u * ks;    // A slice in TMEM (garbage is fine)
                    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                                 "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
                                 ::"r"(d), "r"(ta), "l"(db), "r"(idesc), "r"(1u) : "memory");
                } else {
                    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                                 "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                                 ::"r"(d), "l"(da), "l"(db), "r"(idesc), "r"(1u) : "memory");
                }
            }
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar_w) : "memory");
        uint32_t ok = 0;
        while (!ok) {
            asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                         : "=r"(ok) : "r"(bar_w), "r"(0u) : "memory");
        }
        long long t1 = clock64();
        if (blockIdx.x == 0 && w == 0) out[0] = t1 - t0;
        tm -= w * 256u;
    }
    __syncthreads();
    if (threadIdx.x < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tm) : "memory");
}
int main() {
    long long* d; cudaMalloc(&d, 8);
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    const int iters = 2000;
    const char* names[64] = {"SS same-acc", "SS alt-acc", "TS same-acc", "TS alt-acc", "SS same-acc 4 stages", "SS alt-acc 4 stages", "TS same 4st", "TS alt 4st", "SW32 shifted-A same-acc", "SW32 shifted-A alt-acc"};
    names[24] = "SW32 aligned-A same-acc";
    names[32] = "SS same-acc, TWO issuing warps (cycles per MMA per warp)";
    const int modes[3] = {0, 32, 8};
    for (int mi = 0; mi < 3; ++mi) {
        const int mode = modes[mi];
        for (int N : {32, 64, 128, 192, 256}) {
            if ((mode & 1) && N > 192 && 0) continue;
            k<<<148, 128, 198 * 1024>>>(N, mode, iters, d);
            cudaError_t e = cudaDeviceSynchronize();
            long long c = 0; cudaMemcpy(&c, d, 8, cudaMemcpyDeviceToHost);
            printf("%-24s N=%3d  %7.1f cycles/MMA  (ideal N/2 = %d)  %s\n", names[mode], N, (double)c / (iters * 4), N / 2, e == cudaSuccess ? "" : cudaGetErrorString(e));
            if (e != cudaSuccess) return 1;
        }
    }
    return 0;
}
