// Instantiations of fwd_upper_kernel / fwd_lower_kernel <RA, RB> for every supported stage shape (see fft_plan.cc).
#include "acq_fast_fwd.cuh"
#include <cstdlib>

namespace b200
{
namespace
{
template <void (*kernel)(const FastFwdArgs)>
void configure(size_t smem)
{
    static std::atomic<bool> done[kMaxDeviceSlots];  // one set per kernel instantiation, one flag per device
    const int slot = device_slot();
    if (slot >= 0 && done[slot].load(std::memory_order_relaxed) && smem <= 48 * 1024) return;
    cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    if (smem > 48 * 1024) cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (slot >= 0) done[slot].store(true, std::memory_order_relaxed);
}
bool fixed_off()
{
    static const bool off = std::getenv("B200_NO_FIXED") != nullptr;  // tuning aid: the runtime-geometry kernels only
    return off;
}

// geometries with their own instantiation (see acq_fast_upper.cu): N = 4000, 8000, 16000, 25000, 32000, 50000, 100000, 160000
template <int RA, int RB, int MLO, int CC>
bool upper_fixed(dim3 grid, dim3 block, size_t smem, cudaStream_t st, const FastFwdArgs& a)
{
    if (a.ra_hi != (uint32_t)RA || a.rb_hi != (uint32_t)RB || a.M_lo != (uint32_t)MLO || a.C != (uint32_t)CC || block.x != (unsigned)CC ||
        block.y != (unsigned)(kFastThreads / CC))
        return false;
    configure<fwd_upper_kernel<RA, RB, MLO, CC>>(smem);
    fwd_upper_kernel<RA, RB, MLO, CC><<<grid, block, smem, st>>>(a);
    return true;
}

template <int RA, int RB, int MHI, int NB>
bool lower_fixed(dim3 grid, uint32_t threads, size_t smem, cudaStream_t st, const FastFwdArgs& a)
{
    if (a.ra_lo != (uint32_t)RA || a.rb_lo != (uint32_t)RB || a.M_hi != (uint32_t)MHI || a.nb != (uint32_t)NB || threads != (uint32_t)kFastThreads) return false;
    configure<fwd_lower_kernel<RA, RB, MHI, NB>>(smem);
    fwd_lower_kernel<RA, RB, MHI, NB><<<grid, threads, smem, st>>>(a);
    return true;
}
}  // namespace

#define B200_FWD_CASE(KERNEL, RA, RB, BLOCK)                   \
    case RA * 100 + RB:                                         \
        configure<KERNEL<RA, RB>>(smem);                        \
        KERNEL<RA, RB><<<grid, BLOCK, smem, st>>>(a);           \
        break;

cudaError_t launch_fwd_upper(uint32_t ra, uint32_t rb, dim3 grid, dim3 block, size_t smem, cudaStream_t st, const FastFwdArgs& a)
{
    if (!fixed_off() &&
        (upper_fixed<16, 5, 50, 25>(grid, block, smem, st, a) || upper_fixed<16, 5, 100, 25>(grid, block, smem, st, a) ||
            upper_fixed<16, 8, 125, 25>(grid, block, smem, st, a) || upper_fixed<25, 8, 125, 25>(grid, block, smem, st, a) ||
            upper_fixed<16, 10, 200, 25>(grid, block, smem, st, a) || upper_fixed<25, 10, 200, 25>(grid, block, smem, st, a) ||
            upper_fixed<25, 10, 400, 16>(grid, block, smem, st, a) || upper_fixed<25, 16, 400, 16>(grid, block, smem, st, a)))
        return cudaGetLastError();
    switch (ra * 100 + rb)
        {
            B200_FWD_CASE(fwd_upper_kernel, 4, 2, block)
            B200_FWD_CASE(fwd_upper_kernel, 4, 4, block)
            B200_FWD_CASE(fwd_upper_kernel, 4, 5, block)
            B200_FWD_CASE(fwd_upper_kernel, 4, 8, block)
            B200_FWD_CASE(fwd_upper_kernel, 4, 10, block)
            B200_FWD_CASE(fwd_upper_kernel, 4, 16, block)
            B200_FWD_CASE(fwd_upper_kernel, 4, 20, block)
            B200_FWD_CASE(fwd_upper_kernel, 5, 2, block)
            B200_FWD_CASE(fwd_upper_kernel, 5, 4, block)
            B200_FWD_CASE(fwd_upper_kernel, 5, 5, block)
            B200_FWD_CASE(fwd_upper_kernel, 5, 8, block)
            B200_FWD_CASE(fwd_upper_kernel, 5, 10, block)
            B200_FWD_CASE(fwd_upper_kernel, 5, 16, block)
            B200_FWD_CASE(fwd_upper_kernel, 5, 20, block)
            B200_FWD_CASE(fwd_upper_kernel, 8, 2, block)
            B200_FWD_CASE(fwd_upper_kernel, 8, 4, block)
            B200_FWD_CASE(fwd_upper_kernel, 8, 5, block)
            B200_FWD_CASE(fwd_upper_kernel, 8, 8, block)
            B200_FWD_CASE(fwd_upper_kernel, 8, 10, block)
            B200_FWD_CASE(fwd_upper_kernel, 8, 16, block)
            B200_FWD_CASE(fwd_upper_kernel, 8, 20, block)
            B200_FWD_CASE(fwd_upper_kernel, 16, 2, block)
            B200_FWD_CASE(fwd_upper_kernel, 16, 4, block)
            B200_FWD_CASE(fwd_upper_kernel, 16, 5, block)
            B200_FWD_CASE(fwd_upper_kernel, 16, 8, block)
            B200_FWD_CASE(fwd_upper_kernel, 16, 10, block)
            B200_FWD_CASE(fwd_upper_kernel, 16, 16, block)
            B200_FWD_CASE(fwd_upper_kernel, 16, 20, block)
            B200_FWD_CASE(fwd_upper_kernel, 25, 2, block)
            B200_FWD_CASE(fwd_upper_kernel, 25, 4, block)
            B200_FWD_CASE(fwd_upper_kernel, 25, 5, block)
            B200_FWD_CASE(fwd_upper_kernel, 25, 8, block)
            B200_FWD_CASE(fwd_upper_kernel, 25, 10, block)
            B200_FWD_CASE(fwd_upper_kernel, 25, 16, block)
            B200_FWD_CASE(fwd_upper_kernel, 25, 20, block)
        default:
            return cudaErrorInvalidValue;
        }
    return cudaGetLastError();
}

cudaError_t launch_fwd_lower(uint32_t ra, uint32_t rb, dim3 grid, uint32_t threads, size_t smem, cudaStream_t st, const FastFwdArgs& a)
{
    if (!fixed_off() &&
        (lower_fixed<25, 2, 80, 80>(grid, threads, smem, st, a) || lower_fixed<25, 4, 80, 40>(grid, threads, smem, st, a) ||
            lower_fixed<25, 5, 128, 32>(grid, threads, smem, st, a) || lower_fixed<25, 5, 200, 25>(grid, threads, smem, st, a) ||
            lower_fixed<25, 8, 160, 20>(grid, threads, smem, st, a) || lower_fixed<25, 8, 250, 10>(grid, threads, smem, st, a) ||
            lower_fixed<25, 16, 250, 10>(grid, threads, smem, st, a) || lower_fixed<25, 16, 400, 10>(grid, threads, smem, st, a)))
        return cudaGetLastError();
    switch (ra * 100 + rb)
        {
            B200_FWD_CASE(fwd_lower_kernel, 4, 2, threads)
            B200_FWD_CASE(fwd_lower_kernel, 4, 4, threads)
            B200_FWD_CASE(fwd_lower_kernel, 4, 5, threads)
            B200_FWD_CASE(fwd_lower_kernel, 4, 8, threads)
            B200_FWD_CASE(fwd_lower_kernel, 4, 10, threads)
            B200_FWD_CASE(fwd_lower_kernel, 4, 16, threads)
            B200_FWD_CASE(fwd_lower_kernel, 4, 20, threads)
            B200_FWD_CASE(fwd_lower_kernel, 5, 2, threads)
            B200_FWD_CASE(fwd_lower_kernel, 5, 4, threads)
            B200_FWD_CASE(fwd_lower_kernel, 5, 5, threads)
            B200_FWD_CASE(fwd_lower_kernel, 5, 8, threads)
            B200_FWD_CASE(fwd_lower_kernel, 5, 10, threads)
            B200_FWD_CASE(fwd_lower_kernel, 5, 16, threads)
            B200_FWD_CASE(fwd_lower_kernel, 5, 20, threads)
            B200_FWD_CASE(fwd_lower_kernel, 8, 2, threads)
            B200_FWD_CASE(fwd_lower_kernel, 8, 4, threads)
            B200_FWD_CASE(fwd_lower_kernel, 8, 5, threads)
            B200_FWD_CASE(fwd_lower_kernel, 8, 8, threads)
            B200_FWD_CASE(fwd_lower_kernel, 8, 10, threads)
            B200_FWD_CASE(fwd_lower_kernel, 8, 16, threads)
            B200_FWD_CASE(fwd_lower_kernel, 8, 20, threads)
            B200_FWD_CASE(fwd_lower_kernel, 16, 2, threads)
            B200_FWD_CASE(fwd_lower_kernel, 16, 4, threads)
            B200_FWD_CASE(fwd_lower_kernel, 16, 5, threads)
            B200_FWD_CASE(fwd_lower_kernel, 16, 8, threads)
            B200_FWD_CASE(fwd_lower_kernel, 16, 10, threads)
            B200_FWD_CASE(fwd_lower_kernel, 16, 16, threads)
            B200_FWD_CASE(fwd_lower_kernel, 16, 20, threads)
            B200_FWD_CASE(fwd_lower_kernel, 25, 2, threads)
            B200_FWD_CASE(fwd_lower_kernel, 25, 4, threads)
            B200_FWD_CASE(fwd_lower_kernel, 25, 5, threads)
            B200_FWD_CASE(fwd_lower_kernel, 25, 8, threads)
            B200_FWD_CASE(fwd_lower_kernel, 25, 10, threads)
            B200_FWD_CASE(fwd_lower_kernel, 25, 16, threads)
            B200_FWD_CASE(fwd_lower_kernel, 25, 20, threads)
        default:
            return cudaErrorInvalidValue;
        }
    return cudaGetLastError();
}
#undef B200_FWD_CASE

}  // namespace b200
