// Small host-side helpers around the CUDA runtime: fatal error checks, grow-only device / pinned buffers.
// Error convention follows the reference (diagnostic on stderr + abort(), polygoniser_settings.hpp:297-301,
// object_factory.hpp:731-733): a CUDA failure is fatal, there is no fallback path.
#pragma once
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>

#define ISD_CUDA(call)                                                                                              \
    do {                                                                                                            \
        cudaError_t isd_err__ = (call);                                                                             \
        if (isd_err__ != cudaSuccess) {                                                                             \
            fprintf(stderr, "implisolid_b200: CUDA error %s at %s:%d: %s\n", cudaGetErrorName(isd_err__), __FILE__, \
                    __LINE__, cudaGetErrorString(isd_err__));                                                       \
            abort();                                                                                                \
        }                                                                                                           \
    } while (0)

namespace isd {

// Device buffer that only grows; contents are not preserved across a growth.
struct DeviceBuffer {
    void* ptr = nullptr;
    size_t capacity = 0;
    void reserve(size_t bytes) {
        if (bytes <= capacity) return;
        if (ptr) ISD_CUDA(cudaFree(ptr));
        size_t want = bytes + bytes / 8 + 256;
        ISD_CUDA(cudaMalloc(&ptr, want));
        capacity = want;
    }
    void release() {
        if (ptr) cudaFree(ptr);
        ptr = nullptr;
        capacity = 0;
    }
    template <class T> T* as() const { return reinterpret_cast<T*>(ptr); }
};

// Page-locked host buffer that only grows (mirrors of results the C API hands out by pointer).
struct PinnedBuffer {
    void* ptr = nullptr;
    size_t capacity = 0;
    void reserve(size_t bytes) {
        if (bytes <= capacity) return;
        if (ptr) ISD_CUDA(cudaFreeHost(ptr));
        size_t want = bytes + bytes / 8 + 256;
        ISD_CUDA(cudaMallocHost(&ptr, want));
        capacity = want;
    }
    void release() {
        if (ptr) cudaFreeHost(ptr);
        ptr = nullptr;
        capacity = 0;
    }
    template <class T> T* as() const { return reinterpret_cast<T*>(ptr); }
};

}  // namespace isd
