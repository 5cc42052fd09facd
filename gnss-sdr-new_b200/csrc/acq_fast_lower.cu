// Instantiations of inv_lower_kernel<RA, RB> for every supported stage shape (generated list; see fft_plan.cc).
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
// geometries with their own instantiation (see acq_fast_upper.cu): N = 4000, 8000, 16000, 25000, 32000, 50000, 100000, 160000
template <int RA, int RB, int RAH, int RBH, int NB>
bool launch_fixed(dim3 grid, uint32_t threads, size_t smem, cudaStream_t st, const FastArgs& a)
{
    if (a.ra_lo != (uint32_t)RA || a.rb_lo != (uint32_t)RB || a.ra_hi != (uint32_t)RAH || a.rb_hi != (uint32_t)RBH || a.nb != (uint32_t)NB ||
        threads != lower_threads(NB, RB))
        return false;
    configure<inv_lower_kernel<RA, RB, RAH, RBH, NB>>(smem);
    inv_lower_kernel<RA, RB, RAH, RBH, NB><<<grid, threads, smem, st>>>(a);
    return true;
}

bool try_fixed(dim3 grid, uint32_t threads, size_t smem, cudaStream_t st, const FastArgs& a)
{
    static const bool off = std::getenv("B200_NO_FIXED") != nullptr;  // tuning aid: the runtime-geometry kernels only
    if (off) return false;
    return launch_fixed<25, 2, 16, 5, 80>(grid, threads, smem, st, a) || launch_fixed<25, 4, 16, 5, 40>(grid, threads, smem, st, a) ||
           launch_fixed<25, 5, 16, 8, 32>(grid, threads, smem, st, a) || launch_fixed<25, 5, 25, 8, 25>(grid, threads, smem, st, a) ||
           launch_fixed<25, 8, 16, 10, 20>(grid, threads, smem, st, a) || launch_fixed<25, 8, 25, 10, 10>(grid, threads, smem, st, a) ||
           launch_fixed<25, 16, 25, 10, 10>(grid, threads, smem, st, a) || launch_fixed<25, 16, 25, 16, 10>(grid, threads, smem, st, a);
}
}  // namespace

cudaError_t launch_inv_lower(uint32_t ra, uint32_t rb, dim3 grid, uint32_t threads, size_t smem, cudaStream_t st, const FastArgs& a)
{
    if (try_fixed(grid, threads, smem, st, a)) return cudaGetLastError();
    const uint32_t key = ra * 100 + rb;
    switch (key)
        {
        case 402:
            configure<inv_lower_kernel<4, 2>>(smem);
            inv_lower_kernel<4, 2><<<grid, threads, smem, st>>>(a);
            break;
        case 404:
            configure<inv_lower_kernel<4, 4>>(smem);
            inv_lower_kernel<4, 4><<<grid, threads, smem, st>>>(a);
            break;
        case 405:
            configure<inv_lower_kernel<4, 5>>(smem);
            inv_lower_kernel<4, 5><<<grid, threads, smem, st>>>(a);
            break;
        case 408:
            configure<inv_lower_kernel<4, 8>>(smem);
            inv_lower_kernel<4, 8><<<grid, threads, smem, st>>>(a);
            break;
        case 410:
            configure<inv_lower_kernel<4, 10>>(smem);
            inv_lower_kernel<4, 10><<<grid, threads, smem, st>>>(a);
            break;
        case 416:
            configure<inv_lower_kernel<4, 16>>(smem);
            inv_lower_kernel<4, 16><<<grid, threads, smem, st>>>(a);
            break;
        case 420:
            configure<inv_lower_kernel<4, 20>>(smem);
            inv_lower_kernel<4, 20><<<grid, threads, smem, st>>>(a);
            break;
        case 502:
            configure<inv_lower_kernel<5, 2>>(smem);
            inv_lower_kernel<5, 2><<<grid, threads, smem, st>>>(a);
            break;
        case 504:
            configure<inv_lower_kernel<5, 4>>(smem);
            inv_lower_kernel<5, 4><<<grid, threads, smem, st>>>(a);
            break;
        case 505:
            configure<inv_lower_kernel<5, 5>>(smem);
            inv_lower_kernel<5, 5><<<grid, threads, smem, st>>>(a);
            break;
        case 508:
            configure<inv_lower_kernel<5, 8>>(smem);
            inv_lower_kernel<5, 8><<<grid, threads, smem, st>>>(a);
            break;
        case 510:
            configure<inv_lower_kernel<5, 10>>(smem);
            inv_lower_kernel<5, 10><<<grid, threads, smem, st>>>(a);
            break;
        case 516:
            configure<inv_lower_kernel<5, 16>>(smem);
            inv_lower_kernel<5, 16><<<grid, threads, smem, st>>>(a);
            break;
        case 520:
            configure<inv_lower_kernel<5, 20>>(smem);
            inv_lower_kernel<5, 20><<<grid, threads, smem, st>>>(a);
            break;
        case 802:
            configure<inv_lower_kernel<8, 2>>(smem);
            inv_lower_kernel<8, 2><<<grid, threads, smem, st>>>(a);
            break;
        case 804:
            configure<inv_lower_kernel<8, 4>>(smem);
            inv_lower_kernel<8, 4><<<grid, threads, smem, st>>>(a);
            break;
        case 805:
            configure<inv_lower_kernel<8, 5>>(smem);
            inv_lower_kernel<8, 5><<<grid, threads, smem, st>>>(a);
            break;
        case 808:
            configure<inv_lower_kernel<8, 8>>(smem);
            inv_lower_kernel<8, 8><<<grid, threads, smem, st>>>(a);
            break;
        case 810:
            configure<inv_lower_kernel<8, 10>>(smem);
            inv_lower_kernel<8, 10><<<grid, threads, smem, st>>>(a);
            break;
        case 816:
            configure<inv_lower_kernel<8, 16>>(smem);
            inv_lower_kernel<8, 16><<<grid, threads, smem, st>>>(a);
            break;
        case 820:
            configure<inv_lower_kernel<8, 20>>(smem);
            inv_lower_kernel<8, 20><<<grid, threads, smem, st>>>(a);
            break;
        case 1602:
            configure<inv_lower_kernel<16, 2>>(smem);
            inv_lower_kernel<16, 2><<<grid, threads, smem, st>>>(a);
            break;
        case 1604:
            configure<inv_lower_kernel<16, 4>>(smem);
            inv_lower_kernel<16, 4><<<grid, threads, smem, st>>>(a);
            break;
        case 1605:
            configure<inv_lower_kernel<16, 5>>(smem);
            inv_lower_kernel<16, 5><<<grid, threads, smem, st>>>(a);
            break;
        case 1608:
            configure<inv_lower_kernel<16, 8>>(smem);
            inv_lower_kernel<16, 8><<<grid, threads, smem, st>>>(a);
            break;
        case 1610:
            configure<inv_lower_kernel<16, 10>>(smem);
            inv_lower_kernel<16, 10><<<grid, threads, smem, st>>>(a);
            break;
        case 1616:
            configure<inv_lower_kernel<16, 16>>(smem);
            inv_lower_kernel<16, 16><<<grid, threads, smem, st>>>(a);
            break;
        case 1620:
            configure<inv_lower_kernel<16, 20>>(smem);
            inv_lower_kernel<16, 20><<<grid, threads, smem, st>>>(a);
            break;
        case 2502:
            configure<inv_lower_kernel<25, 2>>(smem);
            inv_lower_kernel<25, 2><<<grid, threads, smem, st>>>(a);
            break;
        case 2504:
            configure<inv_lower_kernel<25, 4>>(smem);
            inv_lower_kernel<25, 4><<<grid, threads, smem, st>>>(a);
            break;
        case 2505:
            configure<inv_lower_kernel<25, 5>>(smem);
            inv_lower_kernel<25, 5><<<grid, threads, smem, st>>>(a);
            break;
        case 2508:
            configure<inv_lower_kernel<25, 8>>(smem);
            inv_lower_kernel<25, 8><<<grid, threads, smem, st>>>(a);
            break;
        case 2510:
            configure<inv_lower_kernel<25, 10>>(smem);
            inv_lower_kernel<25, 10><<<grid, threads, smem, st>>>(a);
            break;
        case 2516:
            configure<inv_lower_kernel<25, 16>>(smem);
            inv_lower_kernel<25, 16><<<grid, threads, smem, st>>>(a);
            break;
        case 2520:
            configure<inv_lower_kernel<25, 20>>(smem);
            inv_lower_kernel<25, 20><<<grid, threads, smem, st>>>(a);
            break;
        default:
            return cudaErrorInvalidValue;
        }
    return cudaGetLastError();
}
}  // namespace b200
