// Instantiations of inv_lower_kernel<RA, RB> for every supported stage shape (generated list; see fft_plan.cc).
#include "acq_fast.cuh"

namespace b200
{
cudaError_t launch_inv_lower(uint32_t ra, uint32_t rb, dim3 grid, size_t smem, cudaStream_t st, const FastArgs& a)
{
    const uint32_t key = ra * 100 + rb;
    switch (key)
        {
        case 402:
            if (smem > 48 * 1024) cudaFuncSetAttribute(inv_lower_kernel<4, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            inv_lower_kernel<4, 2><<<grid, kFastThreads, smem, st>>>(a);
            break;
        case 404:
            if (smem > 48 * 1024) cudaFuncSetAttribute(inv_lower_kernel<4, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            inv_lower_kernel<4, 4><<<grid, kFastThreads, smem, st>>>(a);
            break;
        case 405:
            if (smem > 48 * 1024) cudaFuncSetAttribute(inv_lower_kernel<4, 5>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            inv_lower_kernel<4, 5><<<grid, kFastThreads, smem, st>>>(a);
            break;
        case 408:
            if (smem > 48 * 1024) cudaFuncSetAttribute(inv_lower_kernel<4, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            inv_lower_kernel<4, 8><<<grid, kFastThreads, smem, st>>>(a);
            break;
        case 410:
            if (smem > 48 * 1024) cudaFuncSetAttribute(inv_lower_kernel<4, 10>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            inv_lower_kernel<4, 10><<<grid, kFastThreads, smem, st>>>(a);
            break;
        case 416:
            if (smem > 48 * 1024) cudaFuncSetAttribute(inv_lower_kernel<4, 16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            inv_lower_kernel<4, 16><<<grid, kFastThreads, smem, st>>>(a);
            break;
        case 420:
            if (smem > 48 * 1024) cudaFuncSetAttribute(inv_lower_kernel<4, 20>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            inv_lower_kernel<4, 20><<<grid, kFastThreads, smem, st>>>(a);
            break;
        case 502:
            if (smem > 48 * 1024) cudaFuncSetAttribute(inv_lower_kernel<5, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            inv_lower_kernel<5, 2><<<grid, kFastThreads, smem, st>>>(a);
            break;
        case 504:
            if (smem > 48 * 1024) cudaFuncSetAttribute(inv_lower_kernel<5, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            inv_lower_kernel<5, 4><<<grid, kFastThreads, smem, st>>>(a);
            break;
        case 505:
            if (smem > 48 * 1024) cudaFuncSetAttribute(inv_lower_kernel<5, 5>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            inv_lower_kernel<5, 5><<<grid, kFastThreads, smem, st>>>(a);
            break;
        case 508:
            if (smem > 48 * 1024) cudaFuncSetAttribute(inv_lower_kernel<5, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            inv_lower_kernel<5, 8><<<grid, kFastThreads, smem, st>>>(a);
            break;
        case 510:
            if (smem > 48 * 1024) cudaFuncSetAttribute(inv_lower_kernel<5, 10>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            inv_lower_kernel<5, 10><<<grid, kFastThreads, smem, st>>>(a);
            break;
        case 516:
            if (smem > 48 * 1024) cudaFuncSetAttribute(inv_lower_kernel<5, 16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            inv_lower_kernel<5, 16><<<grid, kFastThreads, smem, st>>>(a);
            break;
        case 520:
            if (smem > 48 * 1024) cudaFuncSetAttribute(inv_lower_kernel<5, 20>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            inv_lower_kernel<5, 20><<<grid, kFastThreads, smem, st>>>(a);
            break;
        case 802:
            if (smem > 48 * 1024) cudaFuncSetAttribute(inv_lower_kernel<8, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            inv_lower_kernel<8, 2><<<grid, kFastThreads, smem, st>>>(a);
            break;
        case 804:
            if (smem > 48 * 1024) cudaFuncSetAttribute(inv_lower_kernel<8, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            inv_lower_kernel<8, 4><<<grid, kFastThreads, smem, st>>>(a);
            break;
        case 805:
            if (smem > 48 * 1024) cudaFuncSetAttribute(inv_lower_kernel<8, 5>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            inv_lower_kernel<8, 5><<<grid, kFastThreads, smem, st>>>(a);
            break;
        case 808:
            if (smem > 48 * 1024) cudaFuncSetAttribute(inv_lower_kernel<8, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            inv_lower_kernel<8, 8><<<grid, kFastThreads, smem, st>>>(a);
            break;
        case 810:
            if (smem > 48 * 1024) cudaFuncSetAttribute(inv_lower_kernel<8, 10>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            inv_lower_kernel<8, 10><<<grid, kFastThreads, smem, st>>>(a);
            break;
        case 816:
            if (smem > 48 * 1024) cudaFuncSetAttribute(inv_lower_kernel<8, 16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            inv_lower_kernel<8, 16><<<grid, kFastThreads, smem, st>>>(a);
            break;
        case 820:
            if (smem > 48 * 1024) cudaFuncSetAttribute(inv_lower_kernel<8, 20>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            inv_lower_kernel<8, 20><<<grid, kFastThreads, smem, st>>>(a);
            break;
        case 1602:
            if (smem > 48 * 1024) cudaFuncSetAttribute(inv_lower_kernel<16, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            inv_lower_kernel<16, 2><<<grid, kFastThreads, smem, st>>>(a);
            break;
        case 1604:
            if (smem > 48 * 1024) cudaFuncSetAttribute(inv_lower_kernel<16, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            inv_lower_kernel<16, 4><<<grid, kFastThreads, smem, st>>>(a);
            break;
        case 1605:
            if (smem > 48 * 1024) cudaFuncSetAttribute(inv_lower_kernel<16, 5>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            inv_lower_kernel<16, 5><<<grid, kFastThreads, smem, st>>>(a);
            break;
        case 1608:
            if (smem > 48 * 1024) cudaFuncSetAttribute(inv_lower_kernel<16, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            inv_lower_kernel<16, 8><<<grid, kFastThreads, smem, st>>>(a);
            break;
        case 1610:
            if (smem > 48 * 1024) cudaFuncSetAttribute(inv_lower_kernel<16, 10>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            inv_lower_kernel<16, 10><<<grid, kFastThreads, smem, st>>>(a);
            break;
        case 1616:
            if (smem > 48 * 1024) cudaFuncSetAttribute(inv_lower_kernel<16, 16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            inv_lower_kernel<16, 16><<<grid, kFastThreads, smem, st>>>(a);
            break;
        case 1620:
            if (smem > 48 * 1024) cudaFuncSetAttribute(inv_lower_kernel<16, 20>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            inv_lower_kernel<16, 20><<<grid, kFastThreads, smem, st>>>(a);
            break;
        case 2502:
            if (smem > 48 * 1024) cudaFuncSetAttribute(inv_lower_kernel<25, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            inv_lower_kernel<25, 2><<<grid, kFastThreads, smem, st>>>(a);
            break;
        case 2504:
            if (smem > 48 * 1024) cudaFuncSetAttribute(inv_lower_kernel<25, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            inv_lower_kernel<25, 4><<<grid, kFastThreads, smem, st>>>(a);
            break;
        case 2505:
            if (smem > 48 * 1024) cudaFuncSetAttribute(inv_lower_kernel<25, 5>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            inv_lower_kernel<25, 5><<<grid, kFastThreads, smem, st>>>(a);
            break;
        case 2508:
            if (smem > 48 * 1024) cudaFuncSetAttribute(inv_lower_kernel<25, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            inv_lower_kernel<25, 8><<<grid, kFastThreads, smem, st>>>(a);
            break;
        case 2510:
            if (smem > 48 * 1024) cudaFuncSetAttribute(inv_lower_kernel<25, 10>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            inv_lower_kernel<25, 10><<<grid, kFastThreads, smem, st>>>(a);
            break;
        case 2516:
            if (smem > 48 * 1024) cudaFuncSetAttribute(inv_lower_kernel<25, 16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            inv_lower_kernel<25, 16><<<grid, kFastThreads, smem, st>>>(a);
            break;
        case 2520:
            if (smem > 48 * 1024) cudaFuncSetAttribute(inv_lower_kernel<25, 20>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            inv_lower_kernel<25, 20><<<grid, kFastThreads, smem, st>>>(a);
            break;
        default:
            return cudaErrorInvalidValue;
        }
    return cudaGetLastError();
}
}  // namespace b200
