// Shared-memory execution of one FFT stage (see fft_plan.h): in-place radix passes over a tile held as
// s[idx * C + c], idx = stage-local index, c = column within the tile.
#pragma once
#include "fft_butterflies.cuh"
#include "fft_plan.h"

namespace b200
{
namespace fft
{
// w[u] = w1^u for u = 1..R-1 by a balanced product tree (depth log2 R)
template <int R>
__device__ __forceinline__ void twiddle_powers(const float2 w1, float2 (&w)[R])
{
    w[0] = make_float2(1.f, 0.f);
    if constexpr (R > 1) w[1] = w1;
    static_for<(R > 2 ? R - 2 : 0)>([&](auto uc) {
        constexpr int u = decltype(uc)::value + 2;
        w[u] = cmul(w[u / 2], w[u - u / 2]);
    });
}

// One radix-R pass over the tile. W = exp(-2*pi*i*k/N) table (forward roots); the inverse conjugates.
//   forward (DIF): butterfly, then multiply output u by W^(q*u*tw_scale)
//   inverse (DIT): multiply input u by conj(W)^(q*u*tw_scale), then butterfly
// q = column + tw_lo_mul * (local index below this dim)
template <int R, bool INV>
__device__ __forceinline__ void smem_pass(float2* __restrict__ s, const uint32_t points, const uint32_t C, const FftPass p,
    const uint32_t tw_col0, const uint32_t tw_col_en, const uint32_t tw_lo_mul, const float2* __restrict__ W)
{
    const uint32_t nb = points / R;  // butterflies in the tile (points = idx-range * C)
    const uint32_t ls = p.lstride;
    const uint32_t st = ls * C;
    for (uint32_t b = threadIdx.x; b < nb; b += blockDim.x)
        {
            const uint32_t c = b % C, r = b / C;
            const uint32_t lo = r % ls, hi = r / ls;
            float2* base = s + (size_t)(hi * ls * R + lo) * C + c;
            float2 v[R];
#pragma unroll
            for (int d = 0; d < R; d++) v[d] = base[(size_t)d * st];
            if (p.has_tw)
                {
                    const uint32_t q = (tw_col_en ? tw_col0 + c : 0u) + tw_lo_mul * lo;
                    float2 w[R];
                    twiddle_powers<R>(__ldg(W + (size_t)q * p.tw_scale), w);
                    if (INV)
                        {
#pragma unroll
                            for (int u = 1; u < R; u++) v[u] = cmul_conj(v[u], w[u]);
                            Dft<R, true>::run(v);
                        }
                    else
                        {
                            Dft<R, false>::run(v);
#pragma unroll
                            for (int u = 1; u < R; u++) v[u] = cmul(v[u], w[u]);
                        }
                }
            else
                {
                    Dft<R, INV>::run(v);
                }
#pragma unroll
            for (int d = 0; d < R; d++) base[(size_t)d * st] = v[d];
        }
}

template <bool INV>
__device__ __forceinline__ void smem_pass_dispatch(float2* s, uint32_t points, uint32_t C, const FftPass p, uint32_t tw_col0,
    uint32_t tw_col_en, uint32_t tw_lo_mul, const float2* W)
{
    switch (p.radix)
        {
#define B200_PASS(R)                                                          \
    case R:                                                                   \
        smem_pass<R, INV>(s, points, C, p, tw_col0, tw_col_en, tw_lo_mul, W); \
        break;
            B200_PASS(2)
            B200_PASS(3)
            B200_PASS(4)
            B200_PASS(5)
            B200_PASS(6)
            B200_PASS(7)
            B200_PASS(8)
            B200_PASS(9)
            B200_PASS(10)
            B200_PASS(11)
            B200_PASS(12)
            B200_PASS(13)
            B200_PASS(15)
            B200_PASS(16)
            B200_PASS(20)
            B200_PASS(25)
#undef B200_PASS
        default:
            break;
        }
}

// All passes of one stage on the tile in shared memory. The caller has filled the tile and synchronised.
template <bool INV>
__device__ __forceinline__ void run_stage(float2* s, const FftStage& st, uint32_t tile, const float2* W)
{
    const uint32_t points = st.tile_points;
    const uint32_t tw_col0 = st.is_upper ? tile * st.tile_step : 0u;
    if (!INV)
        {
            for (uint32_t i = 0; i < st.n_pass; i++)
                {
                    smem_pass_dispatch<false>(s, points, st.C, st.pass[i], tw_col0, st.is_upper, st.tw_lo_mul, W);
                    __syncthreads();
                }
        }
    else
        {
            for (int i = (int)st.n_pass - 1; i >= 0; i--)
                {
                    smem_pass_dispatch<true>(s, points, st.C, st.pass[i], tw_col0, st.is_upper, st.tw_lo_mul, W);
                    __syncthreads();
                }
        }
}

}  // namespace fft
}  // namespace b200
