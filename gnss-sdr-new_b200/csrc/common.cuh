// Shared plumbing of libgnss_b200.so: error reporting, CUDA call checking.
#pragma once
#include "../../include/gnss_b200.h"
#include <cuda_runtime.h>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <mutex>

namespace b200
{
char* last_error_buf();  // thread-local, 512 bytes

inline int fail(int code, const char* fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(last_error_buf(), 512, fmt, ap);
    va_end(ap);
    return code;
}

#define B200_CUDA(expr)                                                                                      \
    do                                                                                                       \
        {                                                                                                    \
            cudaError_t e__ = (expr);                                                                        \
            if (e__ != cudaSuccess)                                                                          \
                return b200::fail(GNSS_B200_ERR_CUDA, "%s:%d %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(e__)); \
        }                                                                                                    \
    while (0)

#define B200_REQUIRE(cond, code, ...)                       \
    do                                                      \
        {                                                   \
            if (!(cond)) return b200::fail(code, __VA_ARGS__); \
        }                                                   \
    while (0)

constexpr int kNumSMs = 148;  // B200

// Growable device / pinned host buffers (freed in the owning handle's destructor).
struct DevBuf
{
    void* p = nullptr;
    size_t cap = 0;
    cudaError_t reserve(size_t bytes)
    {
        if (bytes <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
        cudaError_t e = cudaMalloc(&p, bytes);
        if (e == cudaSuccess) cap = bytes;
        return e;
    }
    void release()
    {
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
    }
    template <class T>
    T* as() const
    {
        return static_cast<T*>(p);
    }
};

struct PinBuf
{
    void* p = nullptr;
    size_t cap = 0;
    cudaError_t reserve(size_t bytes)
    {
        if (bytes <= cap) return cudaSuccess;
        if (p) cudaFreeHost(p);
        p = nullptr;
        cap = 0;
        cudaError_t e = cudaMallocHost(&p, bytes);
        if (e == cudaSuccess) cap = bytes;
        return e;
    }
    void release()
    {
        if (p) cudaFreeHost(p);
        p = nullptr;
        cap = 0;
    }
    template <class T>
    T* as() const
    {
        return static_cast<T*>(p);
    }
};

int select_device(int device);  // validates + cudaSetDevice; returns status

}  // namespace b200
