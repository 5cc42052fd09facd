// Instantiations of inv_upper_kernel<RA, RB, MODE> for every supported stage shape (see fft_plan.cc).
#include "acq_fast.cuh"
#include <cstdlib>

namespace b200
{
namespace
{
template <void (*kernel)(const FastArgs)>
void configure(size_t smem)
{
    static std::atomic<bool> done[kMaxDeviceSlots];  // one set per kernel instantiation, one flag per device
    const int slot = device_slot();
    if (slot >= 0 && done[slot].load(std::memory_order_relaxed) && smem <= 48 * 1024) return;
    cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    if (smem > 48 * 1024) cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (slot >= 0) done[slot].store(true, std::memory_order_relaxed);
}

// geometries with their own instantiation (lower-stage length and column tile compile-time): the plans of the common
// GNSS dwell lengths N = 4000, 8000, 16000, 25000, 32000, 50000, 100000, 160000 (fft_plan.cc picks (RA, RB, M_lo, C))
template <int RA, int RB, int MLO, int CC>
bool launch_fixed(int mode, dim3 grid, dim3 block, size_t smem, cudaStream_t st, const FastArgs& a)
{
    if (a.M_lo != (uint32_t)MLO || a.C != (uint32_t)CC || block.x != (unsigned)CC || block.y != (unsigned)(kFastThreads / CC)) return false;
    if (mode == 1)
        {
            configure<inv_upper_kernel<RA, RB, 1, MLO, CC>>(smem);
            inv_upper_kernel<RA, RB, 1, MLO, CC><<<grid, block, smem, st>>>(a);
        }
    else if (mode == 2)
        {
            configure<inv_upper_kernel<RA, RB, 2, MLO, CC>>(smem);
            inv_upper_kernel<RA, RB, 2, MLO, CC><<<grid, block, smem, st>>>(a);
        }
    else
        {
            return false;  // row buffer / partial rows: the runtime kernel
        }
    return true;
}

template <int RA, int RB>
bool try_fixed(int mode, dim3 grid, dim3 block, size_t smem, cudaStream_t st, const FastArgs& a)
{
    static const bool off = std::getenv("B200_NO_FIXED") != nullptr;  // tuning aid: the runtime-geometry kernels only
    if (off) return false;
    if constexpr (RA == 16 && RB == 5)
        return launch_fixed<16, 5, 50, 25>(mode, grid, block, smem, st, a) || launch_fixed<16, 5, 100, 25>(mode, grid, block, smem, st, a);
    if constexpr (RA == 16 && RB == 8) return launch_fixed<16, 8, 125, 25>(mode, grid, block, smem, st, a);
    if constexpr (RA == 25 && RB == 8) return launch_fixed<25, 8, 125, 25>(mode, grid, block, smem, st, a);
    if constexpr (RA == 16 && RB == 10) return launch_fixed<16, 10, 200, 25>(mode, grid, block, smem, st, a);
    if constexpr (RA == 25 && RB == 10)
        return launch_fixed<25, 10, 200, 25>(mode, grid, block, smem, st, a) || launch_fixed<25, 10, 400, 16>(mode, grid, block, smem, st, a);
    if constexpr (RA == 25 && RB == 16) return launch_fixed<25, 16, 400, 16>(mode, grid, block, smem, st, a);
    return false;
}

template <int RA, int RB>
void launch(int mode, dim3 grid, dim3 block, size_t smem, cudaStream_t st, const FastArgs& a)
{
    if (try_fixed<RA, RB>(mode, grid, block, smem, st, a)) return;
    if (mode == 1)
        {
            configure<inv_upper_kernel<RA, RB, 1>>(smem);
            inv_upper_kernel<RA, RB, 1><<<grid, block, smem, st>>>(a);
        }
    else
        {
            configure<inv_upper_kernel<RA, RB, 0>>(smem);
            inv_upper_kernel<RA, RB, 0><<<grid, block, smem, st>>>(a);
        }
}
}  // namespace

cudaError_t launch_inv_upper(uint32_t ra, uint32_t rb, int mode, dim3 grid, dim3 block, size_t smem, cudaStream_t st, const FastArgs& a)
{
#define B200_UPPER_CASE(RA, RB)                            \
    case RA * 100 + RB:                                     \
        launch<RA, RB>(mode, grid, block, smem, st, a); \
        break;
    switch (ra * 100 + rb)
        {
            B200_UPPER_CASE(4, 2)
            B200_UPPER_CASE(4, 4)
            B200_UPPER_CASE(4, 5)
            B200_UPPER_CASE(4, 8)
            B200_UPPER_CASE(4, 10)
            B200_UPPER_CASE(4, 16)
            B200_UPPER_CASE(4, 20)
            B200_UPPER_CASE(5, 2)
            B200_UPPER_CASE(5, 4)
            B200_UPPER_CASE(5, 5)
            B200_UPPER_CASE(5, 8)
            B200_UPPER_CASE(5, 10)
            B200_UPPER_CASE(5, 16)
            B200_UPPER_CASE(5, 20)
            B200_UPPER_CASE(8, 2)
            B200_UPPER_CASE(8, 4)
            B200_UPPER_CASE(8, 5)
            B200_UPPER_CASE(8, 8)
            B200_UPPER_CASE(8, 10)
            B200_UPPER_CASE(8, 16)
            B200_UPPER_CASE(8, 20)
            B200_UPPER_CASE(16, 2)
            B200_UPPER_CASE(16, 4)
            B200_UPPER_CASE(16, 5)
            B200_UPPER_CASE(16, 8)
            B200_UPPER_CASE(16, 10)
            B200_UPPER_CASE(16, 16)
            B200_UPPER_CASE(16, 20)
            B200_UPPER_CASE(25, 2)
            B200_UPPER_CASE(25, 4)
            B200_UPPER_CASE(25, 5)
            B200_UPPER_CASE(25, 8)
            B200_UPPER_CASE(25, 10)
            B200_UPPER_CASE(25, 16)
            B200_UPPER_CASE(25, 20)
        default:
            return cudaErrorInvalidValue;
        }
#undef B200_UPPER_CASE
    return cudaGetLastError();
}

}  // namespace b200
