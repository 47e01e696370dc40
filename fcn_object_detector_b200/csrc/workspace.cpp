#include "workspace.h"

#include <map>
#include <memory>

#include "common.cuh"

namespace fcnod {

void* Workspace::reserve(size_t bytes) {
    used_ = 0;
    if (bytes > cap_) {
        if (base_) { FCNOD_CUDA(cudaStreamSynchronize(stream())); cudaFree(base_); base_ = 0; cap_ = 0; }
        const size_t want = bytes + bytes / 4;      // head room: slightly larger next calls do not reallocate
        FCNOD_CUDA(cudaMalloc(&base_, want));
        cap_ = want;
    }
    return base_;
}

cudaStream_t Workspace::stream() {
    if (!stream_) FCNOD_CUDA(cudaStreamCreateWithFlags(&stream_, cudaStreamNonBlocking));
    return stream_;
}

Workspace::~Workspace() {
    // thread exit: the context may already be gone, errors are ignored
    if (device_ >= 0 && cudaSetDevice(device_) == cudaSuccess) {
        if (base_) cudaFree(base_);
        if (stream_) cudaStreamDestroy(stream_);
    }
}

Workspace& workspace() {
    static thread_local std::map<int, std::unique_ptr<Workspace> > per_device;
    int dev = 0;
    FCNOD_CUDA(cudaGetDevice(&dev));
    std::unique_ptr<Workspace>& w = per_device[dev];
    if (!w) { w.reset(new Workspace()); w->device_ = dev; }
    return *w;
}

}  // namespace fcnod

extern "C" {

int fcnod_host_alloc(size_t bytes, void** out) {
    FCNOD_API_BEGIN
    FCNOD_CHECK(out, "out is null");
    *out = 0;
    FCNOD_CUDA(cudaMallocHost(out, bytes ? bytes : 1));
    FCNOD_API_END
}

int fcnod_host_free(void* p) {
    FCNOD_API_BEGIN
    if (p) FCNOD_CUDA(cudaFreeHost(p));
    FCNOD_API_END
}

}  // extern "C"
