// Shared helpers for the fcnod_b200 CUDA sources (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <cstdio>
#include <stdexcept>
#include <string>

namespace fcnod {

typedef __nv_bfloat16 bf16;

void set_last_error(const std::string& msg);

struct CudaError : public std::runtime_error {
    explicit CudaError(const std::string& m) : std::runtime_error(m) {}
};

#define FCNOD_CUDA(expr)                                                                          \
    do {                                                                                          \
        cudaError_t _e = (expr);                                                                  \
        if (_e != cudaSuccess)                                                                    \
            throw ::fcnod::CudaError(std::string(#expr) + " failed: " + cudaGetErrorString(_e) +  \
                                     " (" __FILE__ ":" + std::to_string(__LINE__) + ")");        \
    } while (0)

#define FCNOD_CHECK(cond, msg)                                         \
    do {                                                               \
        if (!(cond)) throw std::runtime_error(std::string(msg));       \
    } while (0)

// Wrap a C-ABI body: exceptions -> status code + last-error string.
#define FCNOD_API_BEGIN try {
#define FCNOD_API_END                                   \
    }                                                   \
    catch (const std::exception& e) {                   \
        ::fcnod::set_last_error(e.what());              \
        return 1;                                       \
    }                                                   \
    catch (...) {                                       \
        ::fcnod::set_last_error("unknown error");       \
        return 1;                                       \
    }                                                   \
    return 0;

static inline int ceil_div(int a, int b) { return (a + b - 1) / b; }
static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

template <typename T> __device__ __forceinline__ float to_f32(T v);
template <> __device__ __forceinline__ float to_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ float to_f32<bf16>(bf16 v) { return __bfloat162float(v); }
template <typename T> __device__ __forceinline__ T from_f32(float v);
template <> __device__ __forceinline__ float from_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ bf16 from_f32<bf16>(float v) { return __float2bfloat16_rn(v); }

// NHWC view of a (possibly channel-sliced) activation buffer.
//   element (n,h,w,c) lives at ptr[((n*H + h)*W + w)*ldc + c]
struct View {
    void* ptr;   // already offset to the first channel of the slice
    int C, H, W; // logical channels / spatial size
    int ldc;     // channels of the underlying buffer (pixel stride in elements)
};

}  // namespace fcnod
