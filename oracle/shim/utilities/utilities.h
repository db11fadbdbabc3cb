// TEST INFRASTRUCTURE.  Stand-in for the reference's missing `utilities/` submodule
// (RabinKarp/cuda_utils@62e7088, not vendored under /root/reference).  Written from the call
// sites in the reference's cuda_sim.cu only so that the reference's OWN device
// code can be compiled and run on the B200 box as a statistical cross-check and an informational
// "reference GPU code on B200" timing (oracle/Makefile target `refgpu`).  No arithmetic lives here.
#ifndef MRISIM_ORACLE_UTILITIES_SHIM_H
#define MRISIM_ORACLE_UTILITIES_SHIM_H

// thrust/CUB must be seen before parameters.h defines the macro `threads_per_block`
#include <thrust/device_ptr.h>
#include <thrust/reduce.h>

#include <cublas_v2.h>
#include <cuda_runtime.h>

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#define HANDLE_ERROR(call)                                                                    \
    do {                                                                                      \
        cudaError_t shim_e_ = (call);                                                         \
        if (shim_e_ != cudaSuccess) {                                                         \
            std::fprintf(stderr, "%s:%d: %s\n", __FILE__, __LINE__, cudaGetErrorString(shim_e_)); \
            std::exit(1);                                                                     \
        }                                                                                     \
    } while (0)

// Optional observers (defined by shim/gpu_random_contract.cu in the bit-pinning builds, absent
// otherwise): which device arrays the reference allocated, and when its sprint loop ended.
extern "C" void ref_shim_note_array(void* ptr, size_t n, size_t elem) __attribute__((weak));
extern "C" void ref_shim_sprint_loop_done() __attribute__((weak));

inline void* cudaAllocate(size_t bytes) {
    void* p = nullptr;
    HANDLE_ERROR(cudaMalloc(&p, bytes ? bytes : 1));
    return p;
}
inline void copyToDevice(void* dst, void* src, size_t bytes) {
    if (bytes) HANDLE_ERROR(cudaMemcpy(dst, src, bytes, cudaMemcpyHostToDevice));
}

class Timer {                                   // cpuStart() ... cpuStop() -> elapsed milliseconds
public:
    void cpuStart() { t0_ = std::chrono::steady_clock::now(); }
    float cpuStop() {
        cudaDeviceSynchronize();
        const float ms = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0_).count();
        if (ref_shim_sprint_loop_done) ref_shim_sprint_loop_done();      // cuda_sim.cu:1027 follows the sprint loop
        return ms;
    }
private:
    std::chrono::steady_clock::time_point t0_;
};

class Handler {                                 // Handler(device): selects the device, owns a cuBLAS handle
public:
    cublasHandle_t cublasH;
    explicit Handler(int device) {
        HANDLE_ERROR(cudaSetDevice(device));
        if (cublasCreate(&cublasH) != CUBLAS_STATUS_SUCCESS) { std::fprintf(stderr, "cublasCreate failed\n"); std::exit(1); }
    }
    ~Handler() { cublasDestroy(cublasH); }
    Handler(const Handler&) = delete;
    Handler& operator=(const Handler&) = delete;
};

template <typename T>
class d_array {                                 // device-only array, freed on scope exit
public:
    explicit d_array(size_t n) : n_(n), d_(static_cast<T*>(cudaAllocate(n * sizeof(T)))) {
        if (ref_shim_note_array) ref_shim_note_array((void*)d_, n, sizeof(T));
    }
    d_array(size_t n, T*& dev_out) : d_array(n) { dev_out = d_; }
    ~d_array() { cudaFree(d_); }
    d_array(const d_array&) = delete;
    d_array& operator=(const d_array&) = delete;
    T* dp() { return d_; }
    int getSize() const { return (int)n_; }
private:
    size_t n_;
    T* d_;
};

template <typename T>
class p_array {                                 // host array paired with a device copy
public:
    explicit p_array(size_t n) : n_(n), h_(new T[n]()), own_(true), d_(static_cast<T*>(cudaAllocate(n * sizeof(T)))) {}
    p_array(size_t n, T* host_src, T*& dev_out)
        : n_(n), h_(host_src), own_(false), d_(static_cast<T*>(cudaAllocate(n * sizeof(T)))) {
        dev_out = d_;
        if (ref_shim_note_array) ref_shim_note_array((void*)d_, n, sizeof(T));
    }
    ~p_array() { cudaFree(d_); if (own_) delete[] h_; }
    p_array(const p_array&) = delete;
    p_array& operator=(const p_array&) = delete;
    T& operator[](size_t i) { return h_[i]; }
    int getSize() const { return (int)n_; }
    T* dp() { return d_; }
    void deviceUpload() { copyToDevice((void*)d_, (void*)h_, n_ * sizeof(T)); }
    void deviceDownload() { if (n_) HANDLE_ERROR(cudaMemcpy((void*)h_, (void*)d_, n_ * sizeof(T), cudaMemcpyDeviceToHost)); }
private:
    size_t n_;
    T* h_;
    bool own_;
    T* d_;
};

#endif
