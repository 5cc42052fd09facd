// Micro-benchmark: issue rate of FFMA / FFMA2 / FADD2 / FADD2.RM / LEA / LDS(gather) per SM sub-partition on sm_100a.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)
constexpr int ITERS = 4096;
template <int MODE>
__global__ void __launch_bounds__(256) k(float* out, float seed, long long* clk)
{
    __shared__ float tab[4096];
    for (int i = threadIdx.x; i < 4096; i += 256) tab[i] = i * seed;
    __syncthreads();
    float2 a[8];
    for (int i = 0; i < 8; i++) a[i] = make_float2(seed * i + threadIdx.x, seed + i);
    const float2 m = make_float2(seed, seed + 1.f), c = make_float2(1.0001f, 0.9999f);
    uint32_t idx = threadIdx.x * 37u;
    long long t0 = clock64();
    for (int it = 0; it < ITERS; it++)
        {
#pragma unroll
            for (int i = 0; i < 8; i++)
                {
                    if (MODE == 0) { a[i].x = fmaf(a[i].x, c.x, m.x); a[i].y = fmaf(a[i].y, c.y, m.y); }           // 2 FFMA
                    if (MODE == 1) a[i] = __ffma2_rn(a[i], c, m);                                                    // 1 FFMA2
                    if (MODE == 2) a[i] = __fadd2_rn(a[i], m);                                                       // 1 FADD2
                    if (MODE == 3) a[i] = __fadd2_rd(a[i], m);                                                       // 1 FADD2.RM
                    if (MODE == 4) { a[i].x += tab[(__float_as_uint(a[i].y) + idx) & 4095u]; idx = idx * 5u + 1u; }  // LDS gather + alu
                    if (MODE == 5) { a[i].x = __fadd_rn(a[i].x, m.x); a[i].y = __fadd_rn(a[i].y, m.y); }             // 2 FADD
                    if (MODE == 6) { a[i] = __ffma2_rn(a[i], c, m); idx = idx * 4u + __float_as_uint(a[i].x); }      // FFMA2 + LEA mix
                }
        }
    long long t1 = clock64();
    float s = 0;
    for (int i = 0; i < 8; i++) s += a[i].x + a[i].y;
    out[blockIdx.x * 256 + threadIdx.x] = s + idx;
    if (threadIdx.x == 0 && blockIdx.x == 0) *clk = t1 - t0;
}
template <int MODE>
int run(const char* name, int ctas_per_sm, int insts_per_iter)
{
    float* out; long long* clk;
    CK(cudaMalloc(&out, 148 * 8 * 256 * 4)); CK(cudaMalloc(&clk, 8));
    k<MODE><<<148 * ctas_per_sm, 256>>>(out, 0.5f, clk);
    CK(cudaDeviceSynchronize());
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    k<MODE><<<148 * ctas_per_sm, 256>>>(out, 0.5f, clk);
    cudaEventRecord(e1);
    CK(cudaDeviceSynchronize());
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    long long h; CK(cudaMemcpy(&h, clk, 8, cudaMemcpyDeviceToHost));
    const double warp_insts_per_smsp = (double)ITERS * 8 * insts_per_iter * (8.0 * ctas_per_sm / 4.0);
    printf("%-28s ctas/SM %d: %.3f warp-inst/clk/SMSP by clock64 (%lld clk), %.3f ms -> %.3f warp-inst/ns/SMSP\n", name, ctas_per_sm, warp_insts_per_smsp / (double)h, h, ms, warp_insts_per_smsp / (ms * 1e6));
    cudaFree(out); cudaFree(clk);
    return 0;
}
int main()
{
    for (int c = 1; c <= 8; c *= 2)
        {
            run<0>("FFMA x2", c, 2); run<1>("FFMA2", c, 1); run<2>("FADD2", c, 1); run<3>("FADD2.RM", c, 1);
            run<5>("FADD x2", c, 2); run<4>("LDS gather (+3 alu)", c, 1); run<6>("FFMA2 + IMAD/LEA", c, 2);
        }
    return 0;
}
