// Instantiations of inv_upper_kernel<RA, RB> for every supported stage shape (generated list; see fft_plan.cc).
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
}  // namespace

cudaError_t launch_inv_upper(uint32_t ra, uint32_t rb, dim3 grid, dim3 block, size_t smem, cudaStream_t st, const FastArgs& a)
{
    const uint32_t key = ra * 100 + rb;
    switch (key)
        {
        case 402:
            configure<inv_upper_kernel<4, 2>>(smem);
            inv_upper_kernel<4, 2><<<grid, block, smem, st>>>(a);
            break;
        case 404:
            configure<inv_upper_kernel<4, 4>>(smem);
            inv_upper_kernel<4, 4><<<grid, block, smem, st>>>(a);
            break;
        case 405:
            configure<inv_upper_kernel<4, 5>>(smem);
            inv_upper_kernel<4, 5><<<grid, block, smem, st>>>(a);
            break;
        case 408:
            configure<inv_upper_kernel<4, 8>>(smem);
            inv_upper_kernel<4, 8><<<grid, block, smem, st>>>(a);
            break;
        case 410:
            configure<inv_upper_kernel<4, 10>>(smem);
            inv_upper_kernel<4, 10><<<grid, block, smem, st>>>(a);
            break;
        case 416:
            configure<inv_upper_kernel<4, 16>>(smem);
            inv_upper_kernel<4, 16><<<grid, block, smem, st>>>(a);
            break;
        case 420:
            configure<inv_upper_kernel<4, 20>>(smem);
            inv_upper_kernel<4, 20><<<grid, block, smem, st>>>(a);
            break;
        case 502:
            configure<inv_upper_kernel<5, 2>>(smem);
            inv_upper_kernel<5, 2><<<grid, block, smem, st>>>(a);
            break;
        case 504:
            configure<inv_upper_kernel<5, 4>>(smem);
            inv_upper_kernel<5, 4><<<grid, block, smem, st>>>(a);
            break;
        case 505:
            configure<inv_upper_kernel<5, 5>>(smem);
            inv_upper_kernel<5, 5><<<grid, block, smem, st>>>(a);
            break;
        case 508:
            configure<inv_upper_kernel<5, 8>>(smem);
            inv_upper_kernel<5, 8><<<grid, block, smem, st>>>(a);
            break;
        case 510:
            configure<inv_upper_kernel<5, 10>>(smem);
            inv_upper_kernel<5, 10><<<grid, block, smem, st>>>(a);
            break;
        case 516:
            configure<inv_upper_kernel<5, 16>>(smem);
            inv_upper_kernel<5, 16><<<grid, block, smem, st>>>(a);
            break;
        case 520:
            configure<inv_upper_kernel<5, 20>>(smem);
            inv_upper_kernel<5, 20><<<grid, block, smem, st>>>(a);
            break;
        case 802:
            configure<inv_upper_kernel<8, 2>>(smem);
            inv_upper_kernel<8, 2><<<grid, block, smem, st>>>(a);
            break;
        case 804:
            configure<inv_upper_kernel<8, 4>>(smem);
            inv_upper_kernel<8, 4><<<grid, block, smem, st>>>(a);
            break;
        case 805:
            configure<inv_upper_kernel<8, 5>>(smem);
            inv_upper_kernel<8, 5><<<grid, block, smem, st>>>(a);
            break;
        case 808:
            configure<inv_upper_kernel<8, 8>>(smem);
            inv_upper_kernel<8, 8><<<grid, block, smem, st>>>(a);
            break;
        case 810:
            configure<inv_upper_kernel<8, 10>>(smem);
            inv_upper_kernel<8, 10><<<grid, block, smem, st>>>(a);
            break;
        case 816:
            configure<inv_upper_kernel<8, 16>>(smem);
            inv_upper_kernel<8, 16><<<grid, block, smem, st>>>(a);
            break;
        case 820:
            configure<inv_upper_kernel<8, 20>>(smem);
            inv_upper_kernel<8, 20><<<grid, block, smem, st>>>(a);
            break;
        case 1602:
            configure<inv_upper_kernel<16, 2>>(smem);
            inv_upper_kernel<16, 2><<<grid, block, smem, st>>>(a);
            break;
        case 1604:
            configure<inv_upper_kernel<16, 4>>(smem);
            inv_upper_kernel<16, 4><<<grid, block, smem, st>>>(a);
            break;
        case 1605:
            configure<inv_upper_kernel<16, 5>>(smem);
            inv_upper_kernel<16, 5><<<grid, block, smem, st>>>(a);
            break;
        case 1608:
            configure<inv_upper_kernel<16, 8>>(smem);
            inv_upper_kernel<16, 8><<<grid, block, smem, st>>>(a);
            break;
        case 1610:
            configure<inv_upper_kernel<16, 10>>(smem);
            inv_upper_kernel<16, 10><<<grid, block, smem, st>>>(a);
            break;
        case 1616:
            configure<inv_upper_kernel<16, 16>>(smem);
            inv_upper_kernel<16, 16><<<grid, block, smem, st>>>(a);
            break;
        case 1620:
            configure<inv_upper_kernel<16, 20>>(smem);
            inv_upper_kernel<16, 20><<<grid, block, smem, st>>>(a);
            break;
        case 2502:
            configure<inv_upper_kernel<25, 2>>(smem);
            inv_upper_kernel<25, 2><<<grid, block, smem, st>>>(a);
            break;
        case 2504:
            configure<inv_upper_kernel<25, 4>>(smem);
            inv_upper_kernel<25, 4><<<grid, block, smem, st>>>(a);
            break;
        case 2505:
            configure<inv_upper_kernel<25, 5>>(smem);
            inv_upper_kernel<25, 5><<<grid, block, smem, st>>>(a);
            break;
        case 2508:
            configure<inv_upper_kernel<25, 8>>(smem);
            inv_upper_kernel<25, 8><<<grid, block, smem, st>>>(a);
            break;
        case 2510:
            configure<inv_upper_kernel<25, 10>>(smem);
            inv_upper_kernel<25, 10><<<grid, block, smem, st>>>(a);
            break;
        case 2516:
            configure<inv_upper_kernel<25, 16>>(smem);
            inv_upper_kernel<25, 16><<<grid, block, smem, st>>>(a);
            break;
        case 2520:
            configure<inv_upper_kernel<25, 20>>(smem);
            inv_upper_kernel<25, 20><<<grid, block, smem, st>>>(a);
            break;
        default:
            return cudaErrorInvalidValue;
        }
    return cudaGetLastError();
}
}  // namespace b200
