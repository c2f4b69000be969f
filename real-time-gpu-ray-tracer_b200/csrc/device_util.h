// device_util.h — host-side CUDA plumbing of librtb200: error propagation (no exit(), unlike the reference's
// cudaCheckError, include/Global/Global.cuh:253-255) and grow-only device buffers sized once and reused
// across frames (the reference re-allocates pinned + device TLAS memory every frame, src/Global/Renderer.cu:272-293).
#pragma once

#include <cuda_runtime.h>

#include <cstddef>
#include <cstdio>
#include <stdexcept>
#include <string>

struct CudaError : std::runtime_error {
    explicit CudaError(const std::string &m) : std::runtime_error(m) {}
};

#define CUDA_CHECK(call)                                                                              \
    do {                                                                                              \
        const cudaError_t err__ = (call);                                                             \
        if (err__ != cudaSuccess) {                                                                   \
            char buf__[512];                                                                          \
            snprintf(buf__, sizeof(buf__), "%s at %s:%d: %s", #call, __FILE__, __LINE__, cudaGetErrorString(err__)); \
            throw CudaError(buf__);                                                                   \
        }                                                                                             \
    } while (0)

template<class T>
class DeviceBuffer {
public:
    DeviceBuffer() = default;
    DeviceBuffer(const DeviceBuffer &) = delete;
    DeviceBuffer &operator=(const DeviceBuffer &) = delete;
    ~DeviceBuffer() { release(); }
    // grow-only: keeps the allocation when it is already large enough
    void reserve(size_t count) {
        if (count <= cap_) return;
        release();
        size_t grow = count + count / 8;
        CUDA_CHECK(cudaMalloc(&ptr_, (grow ? grow : 1) * sizeof(T)));
        cap_ = grow ? grow : 1;
    }
    void upload(const T *host, size_t count, cudaStream_t s) {
        reserve(count);
        if (count) CUDA_CHECK(cudaMemcpyAsync(ptr_, host, count * sizeof(T), cudaMemcpyHostToDevice, s));
    }
    void release() {
        if (ptr_) cudaFree(ptr_);
        ptr_ = nullptr;
        cap_ = 0;
    }
    T *get() const { return ptr_; }
    size_t capacity() const { return cap_; }
private:
    T *ptr_ = nullptr;
    size_t cap_ = 0;
};

template<class T>
class PinnedBuffer {
public:
    PinnedBuffer() = default;
    PinnedBuffer(const PinnedBuffer &) = delete;
    PinnedBuffer &operator=(const PinnedBuffer &) = delete;
    ~PinnedBuffer() { if (ptr_) cudaFreeHost(ptr_); }
    void reserve(size_t count) {
        if (count <= cap_) return;
        if (ptr_) cudaFreeHost(ptr_);
        ptr_ = nullptr;
        CUDA_CHECK(cudaHostAlloc(&ptr_, (count ? count : 1) * sizeof(T), cudaHostAllocDefault));
        cap_ = count ? count : 1;
    }
    T *get() const { return ptr_; }
private:
    T *ptr_ = nullptr;
    size_t cap_ = 0;
};

static inline unsigned div_up(size_t a, size_t b) { return (unsigned) ((a + b - 1) / b); }
