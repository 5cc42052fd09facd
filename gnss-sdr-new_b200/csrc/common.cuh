// Shared plumbing of libgnss_b200.so: error reporting, CUDA call checking.
#pragma once
#include <atomic>
#include "../../include/gnss_b200.h"
#include <cuda_runtime.h>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <vector>

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

// cudaFuncSetAttribute acts on the current device only, so "this kernel is configured" is remembered per device
// ordinal (a receiver process may drive several GPUs). -1: ordinal not tracked, configure on every launch.
constexpr int kMaxDeviceSlots = 64;
inline int device_slot()
{
    int d = 0;
    return (cudaGetDevice(&d) == cudaSuccess && d >= 0 && d < kMaxDeviceSlots) ? d : -1;
}

// Optional per-kernel timing with CUDA events on the launching stream (bench.py's roofline numbers).
struct EventProfiler
{
    static constexpr int kCats = 8;
    bool on = false;
    struct Span
    {
        int cat;
        cudaEvent_t a, b;
    };
    std::vector<Span> spans;
    std::vector<cudaEvent_t> pool;
    double ms[kCats] = {0};
    uint64_t count[kCats] = {0};
    cudaEvent_t get()
    {
        if (!pool.empty())
            {
                cudaEvent_t e = pool.back();
                pool.pop_back();
                return e;
            }
        cudaEvent_t e;
        cudaEventCreate(&e);
        return e;
    }
    void begin(int cat, cudaStream_t st)
    {
        if (!on) return;
        Span s{cat, get(), get()};
        cudaEventRecord(s.a, st);
        spans.push_back(s);
    }
    void end(cudaStream_t st)
    {
        if (!on) return;
        cudaEventRecord(spans.back().b, st);
    }
    void collect()  // stream must be synchronised
    {
        for (auto& s : spans)
            {
                float t = 0.f;
                if (cudaEventElapsedTime(&t, s.a, s.b) == cudaSuccess)
                    {
                        ms[s.cat] += t;
                        count[s.cat]++;
                    }
                pool.push_back(s.a);
                pool.push_back(s.b);
            }
        spans.clear();
    }
    void reset()
    {
        for (int i = 0; i < kCats; i++) ms[i] = 0, count[i] = 0;
    }
    void release()
    {
        collect();
        for (auto e : pool) cudaEventDestroy(e);
        pool.clear();
    }
};

}  // namespace b200
