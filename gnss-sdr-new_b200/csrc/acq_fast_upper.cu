// Instantiations of inv_upper_kernel<RA, RB, GRID> for every supported stage shape (see fft_plan.cc).
#include "acq_fast.cuh"

namespace b200
{
namespace
{
template <void (*kernel)(const FastArgs)>
void configure(size_t smem)
{
    static thread_local bool done = false;  // one static per kernel instantiation
    if (done && smem <= 48 * 1024) return;
    cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    if (smem > 48 * 1024) cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    done = true;
}

template <int RA, int RB>
void launch(bool grid_path, dim3 grid, dim3 block, size_t smem, cudaStream_t st, const FastArgs& a)
{
    if (grid_path)
        {
            configure<inv_upper_kernel<RA, RB, true>>(smem);
            inv_upper_kernel<RA, RB, true><<<grid, block, smem, st>>>(a);
        }
    else
        {
            configure<inv_upper_kernel<RA, RB, false>>(smem);
            inv_upper_kernel<RA, RB, false><<<grid, block, smem, st>>>(a);
        }
}
}  // namespace

cudaError_t launch_inv_upper(uint32_t ra, uint32_t rb, bool grid_path, dim3 grid, dim3 block, size_t smem, cudaStream_t st, const FastArgs& a)
{
#define B200_UPPER_CASE(RA, RB)                            \
    case RA * 100 + RB:                                     \
        launch<RA, RB>(grid_path, grid, block, smem, st, a); \
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
