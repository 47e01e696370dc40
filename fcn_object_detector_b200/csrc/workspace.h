// Cached device workspace of the HOST-buffer entry points (fcnod_postnet, fcnod_preprocess, fcnod_label_raster,
// fcnod_aug_forward, fcnod_decode): one grow-only device arena and one stream per (calling thread, device), so a call
// costs its copies and its kernels -- no cudaMalloc / cudaFree (each of which synchronises the device) per call.
// Copies are cudaMemcpyAsync on the workspace stream straight from / to the caller's buffers: true DMA when those are
// page-locked (fcnod_host_alloc, or Caffe's own SyncedMemory in GPU mode), driver-staged otherwise.
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>

namespace fcnod {

class Workspace {
public:
    // an arena of at least `bytes`, 256-byte aligned; contents of an earlier call are gone
    void* reserve(size_t bytes);
    cudaStream_t stream();
    // bump allocation inside the reserved arena
    template <typename T> T* take(size_t count) {
        size_t off = (used_ + 255) / 256 * 256;
        used_ = off + count * sizeof(T);
        return reinterpret_cast<T*>(static_cast<char*>(base_) + off);
    }
    static size_t padded(size_t bytes) { return (bytes + 255) / 256 * 256 + 256; }
    ~Workspace();

private:
    void* base_ = 0;
    size_t cap_ = 0, used_ = 0;
    cudaStream_t stream_ = 0;
    int device_ = -1;
    friend Workspace& workspace();
};

// the workspace of the calling thread on the current device
Workspace& workspace();

}  // namespace fcnod
