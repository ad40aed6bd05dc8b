// yb_rt.h — the thin runtime layer under the kernels: device buffers, copies, and the kernel-definition
// macro.  The product build (nvcc, sm_100a) maps these to CUDA; defining YB_HOSTEMU (tests/emu only, plain
// g++) maps them to malloc and serial loops so the very same kernel bodies can be single-stepped against the
// reference on a machine without a GPU.  The emulator is TEST INFRASTRUCTURE: libyb200.so is never built
// with YB_HOSTEMU and has no host execution path.
#pragma once
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <string>
#include <stdexcept>

#ifndef YB_HOSTEMU
#include <cuda_runtime.h>
#endif

namespace yb {

struct RtError : public std::runtime_error {
    explicit RtError(const std::string &s) : std::runtime_error(s) {}
};

#ifndef YB_HOSTEMU
#define YB_CUDA_CHECK(expr)                                                                            \
    do {                                                                                               \
        cudaError_t e_ = (expr);                                                                       \
        if (e_ != cudaSuccess)                                                                         \
            throw yb::RtError(std::string(#expr) + ": " + cudaGetErrorString(e_));                     \
    } while (0)
typedef cudaStream_t rt_stream;
inline void *rt_malloc(size_t n) { void *p = 0; if (n == 0) n = 16; YB_CUDA_CHECK(cudaMalloc(&p, n)); return p; }
inline void rt_free(void *p) { if (p) cudaFree(p); }
inline void *rt_malloc_host(size_t n) { void *p = 0; if (n == 0) n = 16; YB_CUDA_CHECK(cudaHostAlloc(&p, n, cudaHostAllocDefault)); return p; }
inline void rt_free_host(void *p) { if (p) cudaFreeHost(p); }
inline void rt_h2d(void *d, const void *h, size_t n, rt_stream s) { if (n) YB_CUDA_CHECK(cudaMemcpyAsync(d, h, n, cudaMemcpyHostToDevice, s)); }
inline void rt_d2h(void *h, const void *d, size_t n, rt_stream s) { if (n) { YB_CUDA_CHECK(cudaMemcpyAsync(h, d, n, cudaMemcpyDeviceToHost, s)); YB_CUDA_CHECK(cudaStreamSynchronize(s)); } }
inline void rt_d2d(void *d, const void *s_, size_t n, rt_stream s) { if (n) YB_CUDA_CHECK(cudaMemcpyAsync(d, s_, n, cudaMemcpyDeviceToDevice, s)); }
inline void rt_memset(void *d, int v, size_t n, rt_stream s) { if (n) YB_CUDA_CHECK(cudaMemsetAsync(d, v, n, s)); }
inline void rt_sync(rt_stream s) { YB_CUDA_CHECK(cudaStreamSynchronize(s)); }
#else
typedef void *rt_stream;
inline void *rt_malloc(size_t n) { void *p = malloc(n ? n : 16); if (!p) throw RtError("malloc failed"); return p; }
inline void rt_free(void *p) { free(p); }
inline void *rt_malloc_host(size_t n) { return rt_malloc(n); }
inline void rt_free_host(void *p) { free(p); }
inline void rt_h2d(void *d, const void *h, size_t n, rt_stream) { memcpy(d, h, n); }
inline void rt_d2h(void *h, const void *d, size_t n, rt_stream) { memcpy(h, d, n); }
inline void rt_d2d(void *d, const void *s_, size_t n, rt_stream) { memmove(d, s_, n); }
inline void rt_memset(void *d, int v, size_t n, rt_stream) { memset(d, v, n); }
inline void rt_sync(rt_stream) {}
#endif

// A typed device buffer that only grows.
template <class T>
struct DevBuf {
    T *p;
    size_t cap;
    DevBuf() : p(0), cap(0) {}
    ~DevBuf() { rt_free(p); }
    DevBuf(const DevBuf &) = delete;
    DevBuf &operator=(const DevBuf &) = delete;
    void reserve(size_t n) { if (n > cap) { rt_free(p); p = 0; p = (T *)rt_malloc(n * sizeof(T)); cap = n; } }
    void release() { rt_free(p); p = 0; cap = 0; }
    size_t bytes() const { return cap * sizeof(T); }
};

// Pinned host staging that only grows (uploads from it run at full PCIe rate and never touch the allocator twice).
template <class T>
struct HostBuf {
    T *p;
    size_t cap;
    HostBuf() : p(0), cap(0) {}
    ~HostBuf() { rt_free_host(p); }
    HostBuf(const HostBuf &) = delete;
    HostBuf &operator=(const HostBuf &) = delete;
    void reserve(size_t n) { if (n > cap) { rt_free_host(p); p = 0; p = (T *)rt_malloc_host(n * sizeof(T)); cap = n; } }
};

// ---- kernel definition ------------------------------------------------------------------------------
// YB_KERNEL(name, ArgsType) { body using `tid` and `a` }  defines an element-wise kernel over [0, n).
// Device: __global__ grid-stride wrapper + launcher; emulator: a serial loop.
#ifndef YB_HOSTEMU
#define YB_KERNEL_OCC(name, ArgsType, minBlocks)                                                                      \
    __device__ __forceinline__ void name##_body(uint32_t tid, const ArgsType &a);                      \
    __global__ void __launch_bounds__(256, minBlocks) name##_kernel(uint32_t n, const ArgsType a)                 \
    {                                                                                                  \
        for (uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x; tid < n; tid += gridDim.x * blockDim.x) \
            name##_body(tid, a);                                                                       \
    }                                                                                                  \
    static void name(uint32_t n, const ArgsType &a, yb::rt_stream s)                                   \
    {                                                                                                  \
        if (n == 0) return;                                                                            \
        uint32_t blocks = (n + 255u) / 256u;                                                           \
        if (blocks > 148u * 32u) blocks = 148u * 32u;                                                  \
        name##_kernel<<<blocks, 256, 0, s>>>(n, a);                                                    \
        YB_CUDA_CHECK(cudaGetLastError());                                                             \
        yb::g_kernel_launches++;                                                                       \
    }                                                                                                  \
    __device__ __forceinline__ void name##_body(uint32_t tid, const ArgsType &a)
#define YB_KERNEL(name, ArgsType) YB_KERNEL_OCC(name, ArgsType, 1)
#define YB_ATOMIC_ADD_U32(p, v) atomicAdd((unsigned int *)(p), (unsigned int)(v))
#define YB_ATOMIC_ADD_F32(p, v) atomicAdd((float *)(p), (float)(v))
#define YB_ATOMIC_MIN_I32(p, v) atomicMin((int *)(p), (int)(v))
#define YB_ATOMIC_MAX_I32(p, v) atomicMax((int *)(p), (int)(v))
#define YB_ATOMIC_EXCH_U32(p, v) atomicExch((unsigned int *)(p), (unsigned int)(v))
#define YB_THREADFENCE() __threadfence()
#else
#define YB_KERNEL_OCC(name, ArgsType, minBlocks) YB_KERNEL(name, ArgsType)
#define YB_KERNEL(name, ArgsType)                                                                      \
    inline void name##_body(uint32_t tid, const ArgsType &a);                                          \
    static void name(uint32_t n, const ArgsType &a, yb::rt_stream)                                     \
    {                                                                                                  \
        for (uint32_t tid = 0; tid < n; ++tid) name##_body(tid, a);                                    \
        yb::g_kernel_launches++;                                                                       \
    }                                                                                                  \
    inline void name##_body(uint32_t tid, const ArgsType &a)
template <class T> inline T emu_fetch_add(T *p, T v) { T o = *p; *p = o + v; return o; }
#define YB_ATOMIC_ADD_U32(p, v) yb::emu_fetch_add((uint32_t *)(p), (uint32_t)(v))
#define YB_ATOMIC_ADD_F32(p, v) yb::emu_fetch_add((float *)(p), (float)(v))
inline int emu_min(int *p, int v) { int o = *p; if (v < o) *p = v; return o; }
inline int emu_max(int *p, int v) { int o = *p; if (v > o) *p = v; return o; }
inline uint32_t emu_exch(uint32_t *p, uint32_t v) { uint32_t o = *p; *p = v; return o; }
#define YB_ATOMIC_MIN_I32(p, v) yb::emu_min((int *)(p), (int)(v))
#define YB_ATOMIC_MAX_I32(p, v) yb::emu_max((int *)(p), (int)(v))
#define YB_ATOMIC_EXCH_U32(p, v) yb::emu_exch((uint32_t *)(p), (uint32_t)(v))
#define YB_THREADFENCE()
#endif

extern thread_local uint64_t g_kernel_launches;

}  // namespace yb
