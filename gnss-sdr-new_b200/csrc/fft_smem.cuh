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

// The same pass for a radix without a register codelet (primes 17 ... kMaxAnyRadix, e.g. the 31 of every sample rate that is
// a multiple of 1.023 MHz): a direct DFT of length R out of shared memory, one conjugate pair of outputs per thread and round, a round
// being as many whole butterflies as the CTA has threads for. The R-th roots are staged in shared memory once per pass
// (read from W they would be R different cache lines per warp and step); the twiddle powers are entries of W
// (w^u = W[q*tw_scale*u], index < N), so nothing is stepped. The pass is in place: a round twiddles its inputs (inverse),
// the CTA meets, the round reads, the CTA meets, the round writes. Few registers on purpose: the codelet passes set the
// register budget of the kernels, not this one.
template <bool INV>
__device__ __noinline__ void smem_pass_any(float2* __restrict__ s, const uint32_t points, const uint32_t C, const FftPass p,
    const uint32_t tw_col0, const uint32_t tw_col_en, const uint32_t tw_lo_mul, const float2* __restrict__ W)
{
    __shared__ float2 s_root[kMaxAnyRadix + 1];
    const uint32_t R = p.radix;
    const uint32_t nb = points / R;
    const uint32_t ls = p.lstride;
    const uint32_t st = ls * C;
    const uint32_t nr = p.tw_scale * ls * tw_lo_mul;  // N / R: W[j * nr] = exp(-2 pi i j / R)
    // a thread owns the output pair (m, R - m): their roots are conjugates, so the four real products of one input serve both
    const uint32_t G = (R + 1) / 2;  // R is an odd prime: m = 0 ... (R-1)/2
    const uint32_t per_round = blockDim.x / G;
    const uint32_t bl = threadIdx.x / G, m = threadIdx.x - bl * G, m2 = R - m;
    if (threadIdx.x < R)
        {
            const float2 w = __ldg(W + (size_t)threadIdx.x * nr);
            s_root[threadIdx.x] = INV ? make_float2(w.x, -w.y) : w;
        }
    __syncthreads();
    for (uint32_t b0 = 0; b0 < nb; b0 += per_round)  // trip count uniform over the CTA
        {
            const uint32_t b = b0 + bl;
            const bool live = bl < per_round && b < nb;
            float2* base = s;
            uint32_t tq = 0;
            if (live)
                {
                    const uint32_t c = b % C, r = b / C;
                    const uint32_t lo = r % ls, hi = r / ls;
                    base = s + (size_t)(hi * ls * R + lo) * C + c;
                    if (p.has_tw) tq = ((tw_col_en ? tw_col0 + c : 0u) + tw_lo_mul * lo) * p.tw_scale;
                }
            if (INV && p.has_tw)  // uniform condition
                {
                    if (live && tq && m)
                        {
                            base[(size_t)m * st] = cmul_conj(base[(size_t)m * st], __ldg(W + (size_t)tq * m));
                            base[(size_t)m2 * st] = cmul_conj(base[(size_t)m2 * st], __ldg(W + (size_t)tq * m2));
                        }
                    __syncthreads();
                }
            float2 acc = make_float2(0.f, 0.f), acc2 = make_float2(0.f, 0.f);
            if (live)
                {
                    float rr = 0.f, ii = 0.f, ri = 0.f, ir = 0.f;  // sums of x.re*w.re, x.im*w.im, x.re*w.im, x.im*w.re
                    uint32_t e = 0;                                  // (m * k) mod R
                    const float2* x = base;
                    for (uint32_t k = 0; k < R; k++)
                        {
                            const float2 v = *x, w = s_root[e];
                            rr = fmaf(v.x, w.x, rr);
                            ii = fmaf(v.y, w.y, ii);
                            ri = fmaf(v.x, w.y, ri);
                            ir = fmaf(v.y, w.x, ir);
                            x += st;
                            e += m;
                            if (e >= R) e -= R;
                        }
                    acc = make_float2(rr - ii, ri + ir);   // sum x * w
                    acc2 = make_float2(rr + ii, ir - ri);  // sum x * conj(w)
                    if (!INV && tq && m)
                        {
                            acc = cmul(acc, __ldg(W + (size_t)tq * m));
                            acc2 = cmul(acc2, __ldg(W + (size_t)tq * m2));
                        }
                }
            __syncthreads();
            if (live)
                {
                    base[(size_t)m * st] = acc;
                    if (m) base[(size_t)m2 * st] = acc2;
                }
            __syncthreads();
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
            smem_pass_any<INV>(s, points, C, p, tw_col0, tw_col_en, tw_lo_mul, W);
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
