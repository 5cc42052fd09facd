// Instantiations of fwd_upper_kernel / fwd_lower_kernel <RA, RB> for every supported stage shape (see fft_plan.cc).
#include "acq_fast_fwd.cuh"

namespace b200
{
namespace
{
template <void (*kernel)(const FastFwdArgs)>
void configure(size_t smem)
{
    static thread_local bool done = false;  // one static per kernel instantiation
    if (done && smem <= 48 * 1024) return;
    cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    if (smem > 48 * 1024) cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    done = true;
}
}  // namespace

#define B200_FWD_CASE(KERNEL, RA, RB, BLOCK)                   \
    case RA * 100 + RB:                                         \
        configure<KERNEL<RA, RB>>(smem);                        \
        KERNEL<RA, RB><<<grid, BLOCK, smem, st>>>(a);           \
        break;

cudaError_t launch_fwd_upper(uint32_t ra, uint32_t rb, dim3 grid, dim3 block, size_t smem, cudaStream_t st, const FastFwdArgs& a)
{
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
