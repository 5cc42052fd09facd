// Throughput of the register codelets: one transform per thread (float) against two per thread in packed FP32x2 lanes.
#include "../../gnss-sdr-new_b200/csrc/fft_butterflies.cuh"
#include <cstdio>

namespace b200
{
namespace fft
{
// packed FP32x2 scalar: lane 0 = transform A, lane 1 = transform B. FADD2 / FMUL2 / FFMA2 take a negated operand and a
// broadcast 32-bit immediate, so every scalar instruction of a codelet maps to exactly one packed instruction.
struct pk2
{
    float2 v;  // lane 0 = transform A, lane 1 = transform B
};
__device__ __forceinline__ pk2 make_pk2(float a, float b) { return pk2{make_float2(a, b)}; }
__device__ __forceinline__ pk2 operator+(pk2 a, pk2 b) { return pk2{__fadd2_rn(a.v, b.v)}; }
__device__ __forceinline__ pk2 operator-(pk2 a, pk2 b)
{
    pk2 d;
    asm("sub.rn.f32x2 %0, %1, %2;"
        : "=l"(*reinterpret_cast<unsigned long long*>(&d.v))
        : "l"(*reinterpret_cast<unsigned long long*>(&a.v)), "l"(*reinterpret_cast<unsigned long long*>(&b.v)));
    return d;
}
__device__ __forceinline__ pk2 operator*(pk2 a, pk2 b) { return pk2{__fmul2_rn(a.v, b.v)}; }
__device__ __forceinline__ pk2 operator*(pk2 a, float c) { return pk2{__fmul2_rn(a.v, make_float2(c, c))}; }
__device__ __forceinline__ pk2 operator*(float c, pk2 a) { return pk2{__fmul2_rn(a.v, make_float2(c, c))}; }
__device__ __forceinline__ pk2 operator-(pk2 a) { return pk2{__fmul2_rn(a.v, make_float2(-1.f, -1.f))}; }
__device__ __forceinline__ pk2 fmaf(pk2 a, pk2 b, pk2 c) { return pk2{__ffma2_rn(a.v, b.v, c.v)}; }
__device__ __forceinline__ pk2 fmaf(pk2 a, float b, pk2 c) { return pk2{__ffma2_rn(a.v, make_float2(b, b), c.v)}; }
__device__ __forceinline__ pk2 fmaf(float a, pk2 b, pk2 c) { return pk2{__ffma2_rn(b.v, make_float2(a, a), c.v)}; }

template <>
__device__ __forceinline__ pk2 zero_of<pk2>()
{
    return make_pk2(0.f, 0.f);
}
}  // namespace fft
}  // namespace b200
using namespace b200::fft;

template <int R, class T>
__global__ void __launch_bounds__(256) bench(float* io, int iters)
{
    cx<T> v[R];
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    for (int i = 0; i < R; i++)
        {
            if constexpr (sizeof(T) == 4)
                v[i] = make_cx<T>(io[(t * R + i) % 4096], io[(t * R + i + 7) % 4096]);
            else
                v[i] = make_cx<T>(make_pk2(io[(t * R + i) % 4096], io[(t + i) % 4096]), make_pk2(io[(t * R + i + 7) % 4096], io[(t + i + 3) % 4096]));
        }
    for (int it = 0; it < iters; it++)
        {
            Dft<R, false>::run(v);
            Dft<R, true>::run(v);
#pragma unroll
            for (int i = 0; i < R; i++)
                {
                    v[i].x = v[i].x * (1.0f / R);
                    v[i].y = v[i].y * (1.0f / R);
                }
        }
    float acc = 0.f;
    for (int i = 0; i < R; i++)
        {
            if constexpr (sizeof(T) == 4)
                acc += v[i].x + v[i].y;
            else
                acc += v[i].x.v.x + v[i].x.v.y + v[i].y.v.x + v[i].y.v.y;
        }
    if (acc == 123.456f) io[t % 4096] = acc;
}

template <int R, class T>
void run(const char* name, float* d, int blocks, int threads)
{
    const int iters = 2000;
    bench<R, T><<<blocks, threads>>>(d, 10);
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    cudaDeviceSynchronize();
    cudaEventRecord(a);
    bench<R, T><<<blocks, threads>>>(d, iters);
    cudaEventRecord(b);
    cudaEventSynchronize(b);
    float ms;
    cudaEventElapsedTime(&ms, a, b);
    const double transforms = (double)blocks * threads * iters * 2 * (sizeof(T) == 4 ? 1 : 2);
    printf("%-22s R=%2d blocks=%d threads=%d: %.3f ms, %.2f G transforms/s, %.2f G points/s\n", name, R, blocks, threads, ms, transforms / ms / 1e6,
        transforms * R / ms / 1e6);
}

int main()
{
    float* d;
    cudaMalloc(&d, 4096 * 4);
    float h[4096];
    for (int i = 0; i < 4096; i++) h[i] = (float)((i * 2654435761u) % 1000) / 1000.f - 0.5f;
    cudaMemcpy(d, h, sizeof(h), cudaMemcpyHostToDevice);
    const int sms = 148;
    run<16, float>("scalar", d, sms * 4, 256);
    run<16, pk2>("packed", d, sms * 4, 128);
    run<16, pk2>("packed (same threads)", d, sms * 4, 256);
    run<10, float>("scalar", d, sms * 4, 256);
    run<10, pk2>("packed", d, sms * 4, 128);
    run<10, pk2>("packed (same threads)", d, sms * 4, 256);
    run<25, float>("scalar", d, sms * 4, 256);
    run<25, pk2>("packed", d, sms * 4, 128);
    run<25, pk2>("packed (same threads)", d, sms * 2, 256);
    run<8, float>("scalar", d, sms * 4, 256);
    run<8, pk2>("packed", d, sms * 4, 128);
    run<8, pk2>("packed (same threads)", d, sms * 4, 256);
    return 0;
}
