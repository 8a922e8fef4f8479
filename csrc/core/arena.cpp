#include "arena.h"
#include "common.h"
#include "tclb_model_abi.h"
#include <cstdlib>
#include <cstring>
#include <cuda_runtime.h>

namespace tclb {

int Arena::reserve(size_t bytes, const char *what) {
    if (finalized_) fail("Arena::reserve(%s) after finalize", what);
    size_t off = (total_ + 255) & ~(size_t)255;
    slots_.push_back({off, bytes, what});
    total_ = off + bytes;
    return (int)slots_.size() - 1;
}
void Arena::finalize() {
    if (finalized_) return;
    size_t bytes = (total_ + 255) & ~(size_t)255;
    if (bytes == 0) bytes = 256;
    if (backend_ == TCLB_BACKEND_CUDA) {
        TCLB_CUDA(cudaSetDevice(device_));
        TCLB_CUDA(cudaMalloc(&base_, bytes));
        TCLB_CUDA(cudaMemset(base_, 0, bytes));
    } else {
        if (posix_memalign(&base_, 4096, bytes) != 0) fail("host allocation of %zu bytes failed", bytes);
        memset(base_, 0, bytes);
    }
    logf(L_NOTICE, "[arena] %s allocation of %.3f MB in %zu buffers", backend_ == TCLB_BACKEND_CUDA ? "device" : "host",
         bytes / 1048576.0, slots_.size());
    finalized_ = true;
}
void *Arena::ptr(int slot) const {
    if (!finalized_) fail("Arena::ptr before finalize");
    return (char *)base_ + slots_[slot].off;
}
Arena::~Arena() {
    if (!base_) return;
    if (backend_ == TCLB_BACKEND_CUDA) {
        cudaSetDevice(device_);
        cudaFree(base_);
    } else free(base_);
}
}  // namespace tclb
