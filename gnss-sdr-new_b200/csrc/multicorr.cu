// Tracking multicorrelator for sm_100a: ONE kernel over all channels of an epoch that fuses the carrier
// rotator, the nearest-chip code resampler and the E/P/L (or extended) dot products.
//
// Replaces, per channel, one call of
//   Cpu_Multicorrelator_Real_Codes::Carrier_wipeoff_multicorrelator_resampler
//     (reference: src/algorithms/tracking/libs/cpu_multicorrelator_real_codes.cc:103-144)
//   Cpu_Multicorrelator::Carrier_wipeoff_multicorrelator_resampler (cpu_multicorrelator.cc:87-101)
// i.e. the volk_gnsssdr kernels  *_xn_resampler_*  +  *_rotator_dot_prod_*_xn.
//
// The reference materialises T resampled code vectors (T*n floats) and re-reads them; here a chip is a
// shared-memory lookup and nothing but the sample stream is read from HBM (8 B / sample).
//
// Numerical contract (tests/test_multicorr_gpu.py):
//   * chip indices are bit-exact: idx = floor(fl(fl(fl(step*n) + shift) - rem)) in binary32, evaluated in
//     the reference's order with explicit round-to-nearest intrinsics (no FMA contraction);
//   * the carrier phasor of sample n is P0 * exp(j*n*theta), where theta is the angle of the reference's
//     *binary32-rounded* per-sample phasor exp(-j*step) — the recursion of the reference converges to that
//     angle, not to `step`. Evaluated per thread in double at its first sample, then
//     advanced by a stride phasor.
#include "multicorr_internal.cuh"
#include <atomic>
#include <type_traits>
#include <cmath>
#include <cstdlib>
#include <complex>
#include <vector>

namespace b200
{
using namespace mc;
namespace
{
// Device half of the per-channel constants: the true angle of the binary32-rounded per-sample phasor the reference
// multiplies by (its rotator advances by exactly that angle, not by phase_step_rad), and the stride phasor.
__global__ void prep_kernel(ChanParams* __restrict__ chans, uint32_t n_ch)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_ch) return;
    const double theta = atan2((double)chans[i].wb_im, (double)chans[i].wb_re);
    double sb, cb;
    sincos(theta * (double)kThreads, &sb, &cb);
    chans[i].theta = theta;
    chans[i].wb_re = (float)cb;
    chans[i].wb_im = (float)sb;
}

// chip-phase of sample n for one tap, exactly as volk_gnsssdr_32f_xn_resampler_32f_xn_generic evaluates it
__device__ __forceinline__ int chip_index_plain(float step, float fn, float shift, float rem)
{
    const float a = __fmul_rn(step, fn);
    return __float2int_rd(__fsub_rn(__fadd_rn(a, shift), rem));
}

// volk_gnsssdr_32f_xn_high_dynamics_resampler_32f_xn_generic, tap 0 expression at (already shifted) sample n
__device__ __forceinline__ int chip_index_hd(float step, float rate, uint32_t n, float shift0, float rem)
{
    const float a = __fmul_rn(step, (float)n);
    const float b = __fmul_rn(rate, (float)(n * n));
    return __float2int_rd(__fsub_rn(__fadd_rn(__fadd_rn(a, b), shift0), rem));
}

__device__ __forceinline__ int wrap_chip(int idx, int len)
{
    int r = idx % len;
    return r < 0 ? r + len : r;
}

template <class CodeT>
struct CodeOps;
template <>
struct CodeOps<float>
{
    __device__ static __forceinline__ void mac(float2& acc, const float2 z, const float c)
    {
        acc.x = fmaf(z.x, c, acc.x);
        acc.y = fmaf(z.y, c, acc.y);
    }
};
template <>
struct CodeOps<float2>
{
    __device__ static __forceinline__ void mac(float2& acc, const float2 z, const float2 c)
    {
        acc.x = fmaf(z.x, c.x, fmaf(-z.y, c.y, acc.x));
        acc.y = fmaf(z.x, c.y, fmaf(z.y, c.x, acc.y));
    }
};

// code-fetch modes, decided per CTA
enum
{
    MODE_WINDOW = 0,  // smem holds code[(lo+j) mod L], no modulo at lookup
    MODE_FULL_SMEM = 1,
    MODE_GLOBAL = 2
};

template <class CodeT, int T>
__global__ void __launch_bounds__(kThreads) multicorr_kernel(const ChanParams* __restrict__ chans, const float* __restrict__ shifts_all,
    const uint32_t* __restrict__ shift_samples_all, float2* __restrict__ partials, uint32_t* __restrict__ tickets,
    float2* __restrict__ out, int n_taps_total, int tap0, uint32_t smem_elems)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    CodeT* s_code = reinterpret_cast<CodeT*>(smem_raw);
    __shared__ float2 s_red[kThreads / 32][T];
    __shared__ int s_last;

    const ChanParams cp = chans[blockIdx.y];
    const uint32_t splits = gridDim.x;
    const uint32_t per = (cp.n + splits - 1) / splits;
    const uint32_t n0 = min(cp.n, blockIdx.x * per), n1 = min(cp.n, n0 + per);
    const int L = (int)cp.code_len;
    const CodeT* g_code = static_cast<const CodeT*>(cp.code);
    const bool hd_res = cp.flags & F_HIGH_DYN_RESAMPLER;
    const bool hd_rot = cp.flags & F_HIGH_DYN_ROTATOR;

    float shift[T];
    uint32_t ss[T];
#pragma unroll
    for (int t = 0; t < T; t++)
        {
            const int tt = min(tap0 + t, n_taps_total - 1);
            shift[t] = shifts_all[cp.shifts_off + tt];
            ss[t] = shift_samples_all[cp.shifts_off + tt];
        }
    const float shift0 = shifts_all[cp.shifts_off];

    // ---- stage the code: a pre-wrapped window of exactly the chips this CTA can touch, else the whole code ----
    int mode, lo = 0;
    if (!hd_res)
        {
            float smin = shift[0], smax = shift[0];
#pragma unroll
            for (int t = 1; t < T; t++)
                {
                    smin = fminf(smin, shift[t]);
                    smax = fmaxf(smax, shift[t]);
                }
            const double a0 = (double)cp.code_step * (double)n0, a1 = (double)cp.code_step * (double)(n1 > n0 ? n1 - 1 : n0);
            const double vlo = fmin(a0, a1) + (double)smin - (double)cp.rem_code;
            const double vhi = fmax(a0, a1) + (double)smax - (double)cp.rem_code;
            lo = (int)floor(vlo) - 2;
            const long long width = (long long)floor(vhi) + 2 - lo + 1;
            mode = (width <= (long long)smem_elems) ? MODE_WINDOW : ((uint32_t)L <= smem_elems ? MODE_FULL_SMEM : MODE_GLOBAL);
            if (mode == MODE_WINDOW)
                {
                    int src = wrap_chip(lo + (int)threadIdx.x, L);
                    const int adv = kThreads % L;
                    for (int j = threadIdx.x; j < (int)width; j += kThreads)
                        {
                            s_code[j] = g_code[src];
                            src += adv;
                            if (src >= L) src -= L;
                        }
                }
        }
    else
        {
            mode = ((uint32_t)L <= smem_elems) ? MODE_FULL_SMEM : MODE_GLOBAL;
        }
    if (mode == MODE_FULL_SMEM)
        for (int j = threadIdx.x; j < L; j += kThreads) s_code[j] = g_code[j];
    __syncthreads();

    float2 acc[T];
#pragma unroll
    for (int t = 0; t < T; t++) acc[t] = make_float2(0.f, 0.f);

    uint32_t n = n0 + threadIdx.x;
    if (n < n1)
        {
            // carrier phasor of the first sample of this thread: P0 * exp(j*n*theta), angle reduced in double
            double ang = cp.theta * (double)n;
            ang -= 6.283185307179586476925286766559 * rint(ang * 0.15915494309189533576888376337251);
            float sr, cr;
            sincosf((float)ang, &sr, &cr);
            float2 ph = make_float2(cp.p0_re * cr - cp.p0_im * sr, cp.p0_re * sr + cp.p0_im * cr);
            const float2 wb = make_float2(cp.wb_re, cp.wb_im);
            const float step = cp.code_step, rem = cp.rem_code, rate = cp.code_rate;

            for (; n < n1; n += kThreads)
                {
                    const float2 x = __ldg(cp.sig + n);
                    float2 p = ph;
                    if (hd_rot && n > 0)
                        {
                            // phase = phase_doppler * cpowf(inc_rate, (n-1)^2) normalised  (…high_dynamic_rotator…_generic)
                            const uint32_t m = n - 1;
                            float s2, c2;
                            sincosf(__fmul_rn((float)(m * m), cp.rate_arg), &s2, &c2);
                            p = make_float2(ph.x * c2 - ph.y * s2, ph.x * s2 + ph.y * c2);
                        }
                    const float2 z = make_float2(x.x * p.x - x.y * p.y, x.x * p.y + x.y * p.x);
                    if (mode == MODE_WINDOW)
                        {
                            const float fn = (float)n;
#pragma unroll
                            for (int t = 0; t < T; t++)
                                {
                                    const int idx = chip_index_plain(step, fn, shift[t], rem) - lo;
                                    CodeOps<CodeT>::mac(acc[t], z, s_code[idx]);
                                }
                        }
                    else
                        {
#pragma unroll
                            for (int t = 0; t < T; t++)
                                {
                                    int idx;
                                    if (hd_res)
                                        {
                                            uint32_t nn = n + ss[t];
                                            if (nn >= cp.n) nn -= cp.n;
                                            idx = chip_index_hd(step, rate, nn, shift0, rem);
                                        }
                                    else
                                        {
                                            idx = chip_index_plain(step, (float)n, shift[t], rem);
                                        }
                                    idx = wrap_chip(idx, L);
                                    const CodeT c = (mode == MODE_FULL_SMEM) ? s_code[idx] : g_code[idx];
                                    CodeOps<CodeT>::mac(acc[t], z, c);
                                }
                        }
                    // advance the phasor by kThreads samples
                    ph = make_float2(ph.x * wb.x - ph.y * wb.y, ph.x * wb.y + ph.y * wb.x);
                }
        }

    // ---- CTA reduction: warp shuffles, then one warp over the per-warp partials ----
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int t = 0; t < T; t++)
        {
            float2 v = acc[t];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1)
                {
                    v.x += __shfl_xor_sync(0xffffffffu, v.x, o);
                    v.y += __shfl_xor_sync(0xffffffffu, v.y, o);
                }
            if (lane == 0) s_red[warp][t] = v;
        }
    __syncthreads();
    const int live = min(T, n_taps_total - tap0);
    if (threadIdx.x < live)
        {
            float2 v = make_float2(0.f, 0.f);
#pragma unroll
            for (int w = 0; w < kThreads / 32; w++)
                {
                    v.x += s_red[w][threadIdx.x].x;
                    v.y += s_red[w][threadIdx.x].y;
                }
            if (splits == 1)
                out[(size_t)blockIdx.y * n_taps_total + tap0 + threadIdx.x] = v;
            else
                partials[((size_t)blockIdx.y * splits + blockIdx.x) * n_taps_total + tap0 + threadIdx.x] = v;
        }
    if (splits == 1) return;

    // ---- last CTA of the channel adds the split partials in a fixed order (deterministic) ----
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0)
        {
            const uint32_t ticket = atomicAdd(&tickets[blockIdx.y], 1u);
            s_last = (ticket == splits - 1);
        }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    if (threadIdx.x < live)
        {
            float2 v = make_float2(0.f, 0.f);
            for (uint32_t s = 0; s < splits; s++)
                {
                    const float2 p = __ldcg(&partials[((size_t)blockIdx.y * splits + s) * n_taps_total + tap0 + threadIdx.x]);
                    v.x += p.x;
                    v.y += p.y;
                }
            out[(size_t)blockIdx.y * n_taps_total + tap0 + threadIdx.x] = v;
        }
    if (threadIdx.x == 0) tickets[blockIdx.y] = 0;
}

// ------------------------------------------------------------------------------------------------
// Fast path (real codes, plain resampler, chip window fits shared memory) — the BASELINE config-5 case.
//
// What bounds it (tools/scratch/ubench_pipes.cu, measured on B200): the packed FP32 instructions (FADD2, FFMA2) issue at
// HALF the rate of their scalar forms, so packing saves issue slots but no FMA-pipe cycles. A (sample, tap) costs three
// rounded additions (the reference's chip-phase expression, bit-exact) and two multiply-adds on that pipe, one shared
// memory lookup and one address LEA; everything else is per-sample overhead, and this kernel is organised to keep that
// overhead small:
//   * samples  a warp owns a CONTIGUOUS 1 KB part of every chunk and streams it through a private ring with one bulk
//              copy (cp.async.bulk + mbarrier) issued by lane 0: no per-thread copy instructions, no registers held, no
//              block barrier in the loop. The segment starts on a 16-byte boundary (a misaligned first sample is peeled
//              off), a partial last chunk is read with plain loads;
//   * carrier  sample n = n_thread + m has the phasor P0 exp(j theta n) = P_thread * W[m], W warp-uniform: the loop
//              multiplies by W[m] from a small table (one broadcast load per two rows) and the accumulators are rotated by
//              P_thread once at the end, so there is no per-thread phasor recurrence;
//   * taps     two taps per packed instruction: FADD2 for the two binary32 additions of the chip phase, FADD2.RM against
//              1.5*2^23 for the floor (exact for |phase| < 2^22, off the quarter-rate conversion pipe), FFMA2 for the
//              accumulate with taps paired as (re_t, re_t+1), (im_t, im_t+1); the last tap of an odd count runs on the
//              full-rate scalar forms instead of wasting half of five packed instructions.
// Every lane of a packed op is an IEEE binary32 round-to-nearest (resp. round-down) operation, so the chip
// indices stay bit-identical to volk_gnsssdr_32f_xn_resampler_32f_xn_generic.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ float lds_f32(uint32_t addr)
{
    float v;
    asm("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr));
    return v;
}

__device__ __forceinline__ float2 lds_f32x2(uint32_t addr)
{
    float2 v;
    asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"(addr) : "memory");
    return v;
}

__device__ __forceinline__ float4 lds_f32x4(uint32_t addr)
{
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr) : "memory");
    return v;
}

__device__ __forceinline__ void cp_async_4(uint32_t smem_addr, const void* gptr)
{
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_addr), "l"(gptr) : "memory");
}

// mbarrier + 1-D bulk copy (TMA)
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(bar),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar)
{
    asm volatile("cp.async.bulk.shared::cta.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src), "r"(bytes), "r"(bar)
                 : "memory");
}

constexpr float kFloorMagic = 12582912.0f;  // 1.5 * 2^23
constexpr uint32_t kFloorMagicBits = 0x4B400000u;

template <int T>
__global__ void __launch_bounds__(kThreads, 3) multicorr_window_kernel(const ChanParams* __restrict__ chans, const float* __restrict__ shifts_all,
    float2* __restrict__ partials, uint32_t* __restrict__ tickets, float2* __restrict__ out, int n_taps_total, int tap0, uint32_t smem_elems)
{
    constexpr int TP = T / 2;          // packed tap pairs
    constexpr bool ODD = (T & 1) != 0;  // + one scalar tap
    extern __shared__ __align__(128) unsigned char smem_raw[];
    float* s_code = reinterpret_cast<float*>(smem_raw + kWinRingBytes);
    __shared__ float2 s_red[kThreads / 32][T];
    __shared__ int s_last;
    __shared__ uint32_t s_bias;
    __shared__ float s_shift[T];
    constexpr int kBinades = 22;  // a < 2^22 on this path
    __shared__ __align__(8) float s_k[kBinades][T + (T & 1)];
    __shared__ float2 s_short[kBinades];

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const ChanParams cp = chans[blockIdx.y];
    const uint32_t splits = gridDim.x;
    const uint32_t per = (cp.n + splits - 1) / splits;
    const uint32_t n0 = min(cp.n, blockIdx.x * per), n1 = min(cp.n, n0 + per);
    const int L = (int)cp.code_len;
    const float* g_code = static_cast<const float*>(cp.code);
    const float2* sig = cp.sig;

    // The segment [n0, n1) as: one peeled sample when sig + n0 is not 16-byte aligned | n_full whole chunks | a partial chunk.
    const uint32_t peel = (n1 > n0) ? (uint32_t)((reinterpret_cast<uintptr_t>(sig + n0) >> 3) & 1u) : 0u;
    const uint32_t m0 = n0 + peel;  // first sample of chunk 0
    const uint32_t n_full = (n1 - m0) / kStChunk;
    const uint32_t first = m0 + warp * kStPart + lane;  // this thread's sample in row 0 of chunk 0

    // ---- start streaming before anything else: the window fill below overlaps the first DRAM round trip ----
    const uint32_t wring = (uint32_t)__cvta_generic_to_shared(smem_raw) + warp * (kStRing * kStPart * 8u);
    const uint32_t wbar = (uint32_t)__cvta_generic_to_shared(smem_raw + kStRingBytes + kStTableBytes) + warp * (kStRing * 8u);
    const float2* wsrc = sig + m0 + warp * kStPart;
    if (lane == 0)
        {
#pragma unroll
            for (int s = 0; s < kStRing; s++) mbar_init(wbar + s * 8u, 1u);
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
            if (n_full > 0)
                {
                    mbar_expect_tx(wbar, kStPart * 8u);
                    bulk_g2s(wring, wsrc, kStPart * 8u, wbar);
                }
        }

    float smin = 3.4e38f, smax = -3.4e38f;
#pragma unroll
    for (int t = 0; t < T; t++)
        {
            const float sh = shifts_all[cp.shifts_off + tap0 + t];
            smin = fminf(smin, sh);
            smax = fmaxf(smax, sh);
        }
    const double a0 = (double)cp.code_step * (double)n0, a1 = (double)cp.code_step * (double)(n1 > n0 ? n1 - 1 : n0);
    int lo = (int)floor(fmin(a0, a1) + (double)smin - (double)cp.rem_code) - 2;
    int width = min((int)smem_elems, (int)floor(fmax(a0, a1) + (double)smax - (double)cp.rem_code) + 2 - lo + 1);
    // The pre-wrapped chip window code[(lo + j) mod L], j < width. A doubled code table makes it one contiguous range: one bulk
    // copy from a 16-byte aligned start (lo moves down by up to 3 chips; the launch reserves 8 floats for that and the rounding
    // of the size). Otherwise 4-byte cp.async, every element of the fill in flight at once.
    const uint32_t s_code_addr = (uint32_t)__cvta_generic_to_shared(s_code);
    const uint32_t win_bar = (uint32_t)__cvta_generic_to_shared(smem_raw + kStRingBytes + kStTableBytes) + (kThreads / 32) * kStRing * 8u;
    int start = wrap_chip(lo, L);
    const bool bulk_window = (cp.flags & F_CODE_DOUBLED) && width + 6 <= L + (int)kCodeDoubledPad;
    if (bulk_window)
        {
            lo -= start & 3;
            width = (width + (start & 3) + 3) & ~3;
            start &= ~3;
            if (threadIdx.x == kThreads - 1)
                {
                    mbar_init(win_bar, 1u);
                    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
                    mbar_expect_tx(win_bar, (uint32_t)width * 4u);
                    bulk_g2s(s_code_addr, g_code + start, (uint32_t)width * 4u, win_bar);
                }
        }
    else
        {
            int src = wrap_chip(start + (int)threadIdx.x, L);
            const int adv = kThreads % L;
            for (int j = threadIdx.x; j < width; j += kThreads)
                {
                    cp_async_4(s_code_addr + 4u * (uint32_t)j, g_code + src);
                    src += adv;
                    if (src >= L) src -= L;
                }
            asm volatile("cp.async.commit_group;" ::: "memory");
        }
    if (threadIdx.x == 0) s_bias = s_code_addr - 4u * (kFloorMagicBits + (uint32_t)lo);
    if (threadIdx.x < T) s_shift[threadIdx.x] = shifts_all[cp.shifts_off + tap0 + threadIdx.x];
    // Short-form tables, one row per binade [2^e, 2^(e+1)) of the chip phase a (see sample_short below): the spacing there is
    // u = 2^(e-23), RN_u(x) = fl(fl(x + 1.5 2^e) - 1.5 2^e) for |x| < 2^(e-1), k[e][t] = RN_u(shift_t) + RN_u(-rem) (exact), and the
    // short form holds for a in [2^e + m, 2^(e+1) - m], m = max|shift| + |rem| + 4u, unless some rounding was a tie.
    {
        const float nrem_p = -cp.rem_code;
        for (int i = threadIdx.x; i < kBinades * T; i += kThreads)
            {
                const int e = i / T, t = i - e * T;
                const float big = __uint_as_float((uint32_t)(e + 127) << 23) * 1.5f;
                const float sh = shifts_all[cp.shifts_off + tap0 + t];
                const float r = __fsub_rn(__fadd_rn(nrem_p, big), big);
                s_k[e][t] = __fadd_rn(__fsub_rn(__fadd_rn(sh, big), big), r);
            }
        if (threadIdx.x < kBinades)
            {
                const int e = threadIdx.x;
                const float lo2 = __uint_as_float((uint32_t)(e + 127) << 23);
                const float u = lo2 * 1.1920928955078125e-07f, big = lo2 * 1.5f;
                const float margin = fmaxf(fabsf(smin), fabsf(smax)) + fabsf(nrem_p) + 4.0f * u;
                bool ok = margin * 4.0f < lo2;
                const float r = __fsub_rn(__fadd_rn(nrem_p, big), big);
                ok &= !(fabsf(__fsub_rn(nrem_p, r)) * 2.0f == u);
                for (int t = 0; t < T; t++)
                    {
                        const float sh = shifts_all[cp.shifts_off + tap0 + t];
                        const float q = __fsub_rn(__fadd_rn(sh, big), big);
                        ok &= !(fabsf(__fsub_rn(sh, q)) * 2.0f == u);
                    }
                s_short[e] = ok ? make_float2(lo2 + margin, 2.0f * lo2 - margin) : make_float2(3.4e38f, -3.4e38f);
            }
    }

    // carrier table, relative to the thread's own first sample (sample n = first + m has the phasor P_thread W(m)):
    //   whole chunks   W[kStRows c + r] = exp(j theta (kStChunk c + 32 r))
    //   partial chunk  rows of kThreads samples across the whole CTA (all warps busy): row r of warp w is
    //                  m = kStChunk n_full + kThreads r + 32 w - kStPart w
    //   peeled sample  m = -1 (thread 0 only)
    // angles reduced in double
    float2* s_w = reinterpret_cast<float2*>(smem_raw + kStRingBytes);
    const uint32_t n_w = n_full * kStRows;
    for (uint32_t i = threadIdx.x; i <= n_w + kStTailEntries; i += kThreads)
        {
            double m;
            if (i < n_w)
                m = (double)((i / kStRows) * kStChunk + (i % kStRows) * 32u);
            else if (i < n_w + kStTailEntries)
                {
                    const int r = (int)(i - n_w) / (kThreads / 32), w = (int)(i - n_w) % (kThreads / 32);
                    m = (double)(n_full * kStChunk) + (double)(r * kThreads + w * 32 - w * kStPart);
                }
            else
                m = -1.0;
            double ang = cp.theta * m;
            ang -= 6.283185307179586476925286766559 * rint(ang * 0.15915494309189533576888376337251);
            float sw, cw;
            sincosf((float)ang, &sw, &cw);
            s_w[i] = make_float2(cw, sw);
        }

    // accumulators of a tap pair: (re_t, re_t+1) and (im_t, im_t+1), so that one FFMA2 serves two taps
    float2 acc_re[TP > 0 ? TP : 1], acc_im[TP > 0 ? TP : 1];
    float2 acc_last = make_float2(0.f, 0.f);
#pragma unroll
    for (int q = 0; q < TP; q++) acc_re[q] = acc_im[q] = make_float2(0.f, 0.f);

    const float step = cp.code_step;
    const float nrem = -cp.rem_code;
    const float2 nrem2 = make_float2(nrem, nrem);
    const float2 magic2 = make_float2(kFloorMagic, kFloorMagic);
    asm volatile("cp.async.wait_group 0;" ::: "memory");  // the window has landed (element-wise fill)
    __syncthreads();
    if (bulk_window) mbar_wait(win_bar, 0u);
    // shared-memory address of chip k = base + 4 * (bits(k + magic) - (magicbits + lo)), modulo 2^32: one LEA per chip
    // (read back from shared memory so that ptxas cannot split the constant part off again)
    const uint32_t addr_bias = *reinterpret_cast<volatile uint32_t*>(&s_bias);

    // ---- one sample: x = the sample, w = its table phasor, fn = (float)n ----
    // exact form: the reference's expression floor(fl(fl(a + shift) - rem)), three rounded operations per tap
    auto sample = [&](const float2 x, const float2 w, const float fn) {
        const float zr = x.x * w.x - x.y * w.y;
        const float zi = x.x * w.y + x.y * w.x;
        const float a = __fmul_rn(step, fn);
        if (TP > 0)
            {
                const float2 zr2 = make_float2(zr, zr), zi2 = make_float2(zi, zi), a2 = make_float2(a, a);
#pragma unroll
                for (int q = 0; q < TP; q++)
                    {
                        const float2 sh = *reinterpret_cast<const float2*>(&s_shift[2 * q]);
                        const float2 f = __fadd2_rd(__fadd2_rn(__fadd2_rn(a2, sh), nrem2), magic2);
                        float2 c;
                        c.x = lds_f32(__float_as_uint(f.x) * 4u + addr_bias);
                        c.y = lds_f32(__float_as_uint(f.y) * 4u + addr_bias);
                        acc_re[q] = __ffma2_rn(c, zr2, acc_re[q]);
                        acc_im[q] = __ffma2_rn(c, zi2, acc_im[q]);
                    }
            }
        if (ODD)
            {
                const float f = __fadd_rd(__fadd_rn(__fadd_rn(a, s_shift[T - 1]), nrem), kFloorMagic);
                const float c = lds_f32(__float_as_uint(f) * 4u + addr_bias);
                acc_last.x = fmaf(c, zr, acc_last.x);
                acc_last.y = fmaf(c, zi, acc_last.y);
            }
    };
    // short form, valid while a, a + shift and a + shift - rem all stay inside one binade [2^e, 2^(e+1)), whose spacing is
    // u = 2^(e-23): a is a multiple of u, so fl(a + shift) = a + RN_u(shift) and fl(that - rem) = that + RN_u(-rem) (translation
    // invariance of round-to-nearest on a uniform grid; a tie, whose result would depend on the parity of a / u, is excluded when
    // the offsets are computed). The two roundings collapse into ONE exact addition of k = RN_u(shift) + RN_u(-rem).
    float2 k2[TP > 0 ? TP : 1];
    float k_last = 0.f;
    auto sample_short = [&](const float2 x, const float2 w, const float fn) {
        const float zr = x.x * w.x - x.y * w.y;
        const float zi = x.x * w.y + x.y * w.x;
        const float a = __fmul_rn(step, fn);
        if (TP > 0)
            {
                const float2 zr2 = make_float2(zr, zr), zi2 = make_float2(zi, zi), a2 = make_float2(a, a);
#pragma unroll
                for (int q = 0; q < TP; q++)
                    {
                        const float2 f = __fadd2_rd(__fadd2_rn(a2, k2[q]), magic2);
                        float2 c;
                        c.x = lds_f32(__float_as_uint(f.x) * 4u + addr_bias);
                        c.y = lds_f32(__float_as_uint(f.y) * 4u + addr_bias);
                        acc_re[q] = __ffma2_rn(c, zr2, acc_re[q]);
                        acc_im[q] = __ffma2_rn(c, zi2, acc_im[q]);
                    }
            }
        if (ODD)
            {
                const float f = __fadd_rd(__fadd_rn(a, k_last), kFloorMagic);
                const float c = lds_f32(__float_as_uint(f) * 4u + addr_bias);
                acc_last.x = fmaf(c, zr, acc_last.x);
                acc_last.y = fmaf(c, zi, acc_last.y);
            }
    };
    // The binade [bin_lo, bin_hi) the offsets k are loaded for, and the range [short_lo, short_hi] of a inside it for which the
    // short form is proven (empty when a tie or a too small binade rules it out; tables built in the prologue). Warp-uniform.
    float bin_lo = 1.f, bin_hi = 0.f, short_lo = 3.4e38f, short_hi = -3.4e38f;
    auto rebin = [&](const float a) {
        const int e = (int)(__float_as_uint(a) >> 23) - 127;  // a > 0
        bin_lo = __uint_as_float(__float_as_uint(a) & 0x7F800000u);
        bin_hi = bin_lo * 2.0f;
        short_lo = 3.4e38f;
        short_hi = -3.4e38f;
        if (e < 0 || e >= kBinades) return;
        short_lo = s_short[e].x;
        short_hi = s_short[e].y;
#pragma unroll
        for (int q = 0; q < TP; q++) k2[q] = *reinterpret_cast<const float2*>(&s_k[e][2 * q]);
        if (ODD) k_last = s_k[e][T - 1];
    };

    // ---- whole chunks ----
    float fnb = (float)first;  // exact: n < 2^24
    const uint32_t w_addr = (uint32_t)__cvta_generic_to_shared(s_w);
    uint32_t g = 0;
    // one chunk of this warp: refill the slot consumed last, wait for this chunk's copy, kStRows samples per lane.
    // Running state instead of expressions in g, so that nothing is recomputed per chunk: slot toggle, mbarrier phase,
    // the address of the next bulk copy and of this chunk's carrier-table row.
    uint32_t slot_sel = 0, phase = 0, wrow = w_addr;
    const float2* next_src = wsrc + kStChunk;
    const uint32_t lane_slot = wring + lane * 8u;
    auto chunk = [&](auto short_tag) {
        constexpr bool SHORT = decltype(short_tag)::value;
        const uint32_t other = slot_sel ^ 1u;
        __syncwarp();  // every lane is done with the slot of chunk g - 1: refill it with chunk g + 1
        if (lane == 0 && g + 1 < n_full)
            {
                mbar_expect_tx(wbar + other * 8u, kStPart * 8u);
                bulk_g2s(wring + other * (kStPart * 8u), next_src, kStPart * 8u, wbar + other * 8u);
            }
        next_src += kStChunk;
        mbar_wait(wbar + slot_sel * 8u, phase);
        const uint32_t slot = lane_slot + slot_sel * (kStPart * 8u);
#pragma unroll
        for (int r = 0; r < kStRows; r += 2)
            {
                const float4 w = lds_f32x4(wrow + r * 8u);
                const float fn0 = (r == 0) ? fnb : fnb + (float)(32 * r);
                if (SHORT)
                    {
                        sample_short(lds_f32x2(slot + r * 256u), make_float2(w.x, w.y), fn0);
                        sample_short(lds_f32x2(slot + r * 256u + 256u), make_float2(w.z, w.w), fnb + (float)(32 * r + 32));
                    }
                else
                    {
                        sample(lds_f32x2(slot + r * 256u), make_float2(w.x, w.y), fn0);
                        sample(lds_f32x2(slot + r * 256u + 256u), make_float2(w.z, w.w), fnb + (float)(32 * r + 32));
                    }
            }
        fnb += (float)kStChunk;
        wrow += kStRows * 8u;
        phase ^= slot_sel;  // the phase of a slot flips every second chunk
        slot_sel = other;
        g++;
    };
    const float lane_f = (float)lane;
    const bool may_short = step > 0.f && !(cp.flags & F_EXACT_ONLY);
    while (g < n_full)
        {
            // chip phase of the first and of the last sample of this warp's part of chunk g (warp-uniform)
            const float fn_last = fnb - lane_f + (float)(kStPart - 1);
            const float a_min = __fmul_rn(step, fnb - lane_f), a_max = __fmul_rn(step, fn_last);
            if (may_short && !(a_min >= bin_lo && a_min < bin_hi)) rebin(a_min);
            if (may_short && a_min >= short_lo && a_max <= short_hi)
                {
                    // how many chunks in a row stay inside the proven range: estimate, then check the last one exactly
                    uint32_t run = 1u + (uint32_t)((short_hi - a_max) / (step * (float)kStChunk));
                    run = min(run, n_full - g);
                    while (run > 1u && !(__fmul_rn(step, fn_last + (float)((run - 1u) * kStChunk)) <= short_hi)) run--;
                    for (uint32_t i = 0; i < run; i++) chunk(std::true_type{});
                }
            else
                chunk(std::false_type{});
        }
    // ---- the partial chunk: rows of kThreads samples across the CTA, plain loads; lanes past the end contribute a zero sample
    // at the last chip phase ----
    {
        const uint32_t base = m0 + n_full * kStChunk;
        if (base < n1)
            {
                const float f_last = (float)(n1 - 1);
                const uint32_t rows = (n1 - base + kThreads - 1) / kThreads;  // <= kStRows
                for (uint32_t r = 0; r < rows; r++)
                    {
                        const uint32_t nn = base + r * kThreads + threadIdx.x;
                        const bool live = nn < n1;
                        const float2 x = live ? __ldg(sig + nn) : make_float2(0.f, 0.f);
                        sample(x, s_w[n_w + r * (kThreads / 32) + warp], live ? (float)nn : f_last);
                    }
            }
        // ---- the peeled sample: phasor P(first of thread 0) * exp(-j theta) ----
        if (peel && threadIdx.x == 0) sample(__ldg(sig + n0), s_w[n_w + kStTailEntries], (float)n0);
    }

    // rotate by this thread's constant phasor P0 exp(j theta first), angle reduced in double
    float2 pt;
    {
        double ang = cp.theta * (double)first;
        ang -= 6.283185307179586476925286766559 * rint(ang * 0.15915494309189533576888376337251);
        float sr, cr;
        sincosf((float)ang, &sr, &cr);
        pt = make_float2(cp.p0_re * cr - cp.p0_im * sr, cp.p0_re * sr + cp.p0_im * cr);
    }
#pragma unroll
    for (int t = 0; t < T; t++)
        {
            float2 v;
            if (ODD && t == T - 1)
                v = acc_last;
            else
                v = (t & 1) ? make_float2(acc_re[t / 2].y, acc_im[t / 2].y) : make_float2(acc_re[t / 2].x, acc_im[t / 2].x);
            v = make_float2(v.x * pt.x - v.y * pt.y, v.x * pt.y + v.y * pt.x);
#pragma unroll
            for (int o = 16; o > 0; o >>= 1)
                {
                    v.x += __shfl_xor_sync(0xffffffffu, v.x, o);
                    v.y += __shfl_xor_sync(0xffffffffu, v.y, o);
                }
            if (lane == 0) s_red[warp][t] = v;
        }
    __syncthreads();
    const int live = min(T, n_taps_total - tap0);
    if (threadIdx.x < live)
        {
            float2 v = make_float2(0.f, 0.f);
#pragma unroll
            for (int w = 0; w < kThreads / 32; w++)
                {
                    v.x += s_red[w][threadIdx.x].x;
                    v.y += s_red[w][threadIdx.x].y;
                }
            if (splits == 1)
                out[(size_t)blockIdx.y * n_taps_total + tap0 + threadIdx.x] = v;
            else
                partials[((size_t)blockIdx.y * splits + blockIdx.x) * n_taps_total + tap0 + threadIdx.x] = v;
        }
    if (splits == 1) return;
    if (threadIdx.x < live) __threadfence();
    __syncthreads();
    if (threadIdx.x == 0)
        {
            const uint32_t ticket = atomicAdd(&tickets[blockIdx.y], 1u);
            s_last = (ticket == splits - 1);
        }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    if (threadIdx.x < live)
        {
            float2 v = make_float2(0.f, 0.f);
            for (uint32_t sp = 0; sp < splits; sp++)
                {
                    const float2 p = __ldcg(&partials[((size_t)blockIdx.y * splits + sp) * n_taps_total + tap0 + threadIdx.x]);
                    v.x += p.x;
                    v.y += p.y;
                }
            out[(size_t)blockIdx.y * n_taps_total + tap0 + threadIdx.x] = v;
        }
    if (threadIdx.x == 0) tickets[blockIdx.y] = 0;
}

void launch_window(int T, dim3 grid, size_t smem, cudaStream_t st, const ChanParams* chans, const float* shifts, float2* partials,
    uint32_t* tickets, float2* out, int n_taps_total, int tap0, uint32_t smem_elems)
{
#define B200_MCW_CASE(TT)                                                                                                                  \
    case TT:                                                                                                                               \
        {                                                                                                                                  \
            /* function attributes are per device: one flag per device ordinal, not one per process */                                     \
            static std::atomic<bool> configured[kMaxDeviceSlots];                                                                          \
            const int dev = device_slot();                                                                                                 \
            const bool tracked = dev >= 0;                                                                                                 \
            if (!tracked || !configured[dev].load(std::memory_order_acquire))                                                              \
                {                                                                                                                          \
                    cudaFuncSetAttribute(multicorr_window_kernel<TT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemCapBytes);    \
                    cudaFuncSetAttribute(multicorr_window_kernel<TT>, cudaFuncAttributePreferredSharedMemoryCarveout, 100);                \
                    if (tracked) configured[dev].store(true, std::memory_order_release);                                                   \
                }                                                                                                                          \
        }                                                                                                                                  \
        multicorr_window_kernel<TT><<<grid, kThreads, smem, st>>>(chans, shifts, partials, tickets, out, n_taps_total, tap0, smem_elems);   \
        break;
    switch (T)
        {
            B200_MCW_CASE(1)
            B200_MCW_CASE(2)
            B200_MCW_CASE(3)
            B200_MCW_CASE(4)
            B200_MCW_CASE(5)
            B200_MCW_CASE(6)
            B200_MCW_CASE(7)
            B200_MCW_CASE(8)
            B200_MCW_CASE(9)
            B200_MCW_CASE(10)
            B200_MCW_CASE(11)
            B200_MCW_CASE(12)
            B200_MCW_CASE(13)
            B200_MCW_CASE(14)
            B200_MCW_CASE(15)
            B200_MCW_CASE(16)
        }
#undef B200_MCW_CASE
}

// parity aid: the wrapped chip index of every (tap, sample) of one channel
__global__ void chip_index_kernel(ChanParams cp, const float* __restrict__ shifts, const uint32_t* __restrict__ shift_samples,
    int n_taps, int32_t* __restrict__ out)
{
    const bool hd = cp.flags & F_HIGH_DYN_RESAMPLER;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < cp.n * (uint32_t)n_taps; i += gridDim.x * blockDim.x)
        {
            const uint32_t t = i / cp.n, n = i % cp.n;
            int idx;
            if (hd)
                {
                    uint32_t nn = n + shift_samples[t];
                    if (nn >= cp.n) nn -= cp.n;
                    idx = chip_index_hd(cp.code_step, cp.code_rate, nn, shifts[0], cp.rem_code);
                }
            else
                {
                    idx = chip_index_plain(cp.code_step, (float)n, shifts[t], cp.rem_code);
                }
            out[i] = wrap_chip(idx, (int)cp.code_len);
        }
}

template <class CodeT>
void launch_taps(int T, dim3 grid, size_t smem, cudaStream_t st, const ChanParams* chans, const float* shifts, const uint32_t* ss,
    float2* partials, uint32_t* tickets, float2* out, int n_taps_total, int tap0, uint32_t smem_elems)
{
#define B200_MC_CASE(TT)                                                                                                           \
    case TT:                                                                                                                       \
        cudaFuncSetAttribute(multicorr_kernel<CodeT, TT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kSmemCapBytes);         \
        multicorr_kernel<CodeT, TT><<<grid, kThreads, smem, st>>>(chans, shifts, ss, partials, tickets, out, n_taps_total, tap0, smem_elems); \
        break;
    switch (T)
        {
            B200_MC_CASE(1)
            B200_MC_CASE(2)
            B200_MC_CASE(3)
            B200_MC_CASE(4)
            B200_MC_CASE(5)
            B200_MC_CASE(6)
            B200_MC_CASE(7)
            B200_MC_CASE(8)
            B200_MC_CASE(9)
            B200_MC_CASE(10)
            B200_MC_CASE(11)
            B200_MC_CASE(12)
            B200_MC_CASE(13)
            B200_MC_CASE(14)
            B200_MC_CASE(15)
            B200_MC_CASE(16)
        }
#undef B200_MC_CASE
}

}  // namespace

cudaError_t mc::upload_doubled_code(void* d_dst, const float* h_code, uint32_t code_len, cudaStream_t st)
{
    std::vector<float> twice(doubled_code_floats(code_len));
    for (size_t i = 0; i < twice.size(); i++) twice[i] = h_code[i % code_len];
    cudaError_t e = cudaMemcpyAsync(d_dst, twice.data(), twice.size() * sizeof(float), cudaMemcpyHostToDevice, st);
    if (e != cudaSuccess) return e;
    return cudaStreamSynchronize(st);  // the staging vector dies with this call
}

cudaError_t mc::launch_real_plain(bool fast, uint32_t n_ch, uint32_t splits, uint32_t smem_elems, cudaStream_t st, const ChanParams* d_cp,
    const float* d_shifts, const uint32_t* d_shift_samples, float2* d_partials, uint32_t* d_tickets, float2* d_out, uint32_t n_taps)
{
    const dim3 grid(splits, n_ch);
    for (uint32_t tap0 = 0; tap0 < n_taps; tap0 += kMaxTapsPerPass)
        {
            const int T = (int)std::min<uint32_t>(kMaxTapsPerPass, n_taps - tap0);
            if (fast)
                launch_window(T, grid, (size_t)(smem_elems + 8) * 4 + kWinRingBytes, st, d_cp, d_shifts, d_partials, d_tickets, d_out, (int)n_taps, (int)tap0, smem_elems);
            else
                launch_taps<float>(T, grid, (size_t)smem_elems * 4, st, d_cp, d_shifts, d_shift_samples, d_partials, d_tickets, d_out, (int)n_taps, (int)tap0, smem_elems);
        }
    return cudaGetLastError();
}
}  // namespace b200

using namespace b200;

struct gnss_b200_mc
{
    std::mutex mu;
    int device = 0;
    uint32_t max_channels = 0, max_samples = 0, max_taps = 0;
    cudaStream_t stream = nullptr;
    std::vector<DevBuf> codes;  // per channel
    std::vector<uint32_t> code_len;
    std::vector<uint8_t> code_complex;
    DevBuf d_sig;     // [max_channels][max_samples] staging for host-pointer calls
    DevBuf d_stream;  // shared stream
    uint64_t stream_len = 0;
    DevBuf d_params;  // ChanParams[n] | shifts | shift_samples
    PinBuf h_params;
    DevBuf d_partials, d_tickets, d_out;
    PinBuf h_out;
    DevBuf d_idx;
    uint64_t launches = 0;
    uint32_t last_n_ch = 0, last_n_taps = 0;
    EventProfiler prof;
};

namespace
{
// Host half of one batch: per-channel constants (the reference computes these at cpu_multicorrelator_real_codes.cc:114-119).
int build_params(gnss_b200_mc* h, uint32_t n_ch, const uint32_t* ch, const gnss_b200_mc_args* args, const float* shifts, uint32_t n_taps,
    uint32_t flags, const float2* d_sig_base, uint64_t ch_stride, bool device_resident, uint32_t* out_smem_elems, uint32_t* out_splits, bool* out_complex, bool* out_fast = nullptr)
{
    const size_t bytes = sizeof(ChanParams) * n_ch + sizeof(float) * n_ch * n_taps + sizeof(uint32_t) * n_ch * n_taps;
    B200_CUDA(h->h_params.reserve(bytes));
    B200_CUDA(h->d_params.reserve(bytes));
    ChanParams* cp = h->h_params.as<ChanParams>();
    float* hs = reinterpret_cast<float*>(cp + n_ch);
    uint32_t* hss = reinterpret_cast<uint32_t*>(hs + (size_t)n_ch * n_taps);
    std::memcpy(hs, shifts, sizeof(float) * n_ch * n_taps);

    const bool hd_res = flags & GNSS_B200_MC_HIGH_DYN;
    const bool hd_rot = hd_res && !(flags & GNSS_B200_MC_SIX_ARG);
    const char* exact_env = std::getenv("B200_MC_EXACT_ONLY");  // parity aid, see F_EXACT_ONLY
    const bool exact_only = exact_env && exact_env[0] == '1';
    uint64_t max_len = 1, max_n = 1;
    double max_span = 0.0;  // chips swept by one channel over its whole integration
    for (uint32_t i = 0; i < n_ch; i++)
        {
            max_n = std::max<uint64_t>(max_n, args[i].n_samples);
            max_span = std::max(max_span, std::fabs((double)args[i].code_step_chips) * (double)args[i].n_samples);
        }
    // split each channel over several CTAs when the batch alone cannot fill the machine
    uint32_t splits = 1;
    const uint32_t target_ctas = kNumSMs * 4;
    if (n_ch < target_ctas)
        {
            splits = (target_ctas + n_ch - 1) / n_ch;
            const uint32_t max_splits = (uint32_t)std::max<uint64_t>(1, max_n / 2048);
            splits = std::max(1u, std::min(splits, max_splits));
        }
    if (!hd_res && !h->code_complex[ch[0] < h->max_channels ? ch[0] : 0])
        {
            // window kernel: long segments amortise the per-CTA prologue/epilogue (one segment per channel when the
            // batch fills the machine), but the chip window of a segment must leave room for 3 CTAs per SM
            const double window_budget = 10800.0;  // floats: 42 KB next to the 28 KB of sample ring and carrier table
            while (max_span / splits + 64.0 > window_budget && (uint64_t)splits * 2048 < max_n) splits++;
            splits = std::max<uint32_t>(splits, (uint32_t)((max_n + kStMaxSegment - 1) / kStMaxSegment));
            if (const char* e = std::getenv("B200_MC_SPLITS")) splits = (uint32_t)std::max(1, std::atoi(e));  // tuning aid
        }
    uint64_t max_window = 1;  // widest per-CTA chip window over the batch
    bool cplx = false, small_phase = true;
    if (!hd_res) std::memset(hss, 0, sizeof(uint32_t) * (size_t)n_ch * n_taps);
    for (uint32_t i = 0; i < n_ch; i++)
        {
            const uint32_t c = ch[i];
            B200_REQUIRE(c < h->max_channels, GNSS_B200_ERR_INVALID, "channel %u out of range", c);
            B200_REQUIRE(h->codes[c].p != nullptr, GNSS_B200_ERR_STATE, "channel %u has no local code (call gnss_b200_mc_set_code)", c);
            const gnss_b200_mc_args& a = args[i];
            B200_REQUIRE(a.n_samples <= h->max_samples, GNSS_B200_ERR_INVALID, "n_samples %u > max_samples %u", a.n_samples, h->max_samples);
            if (i == 0) cplx = h->code_complex[c];
            B200_REQUIRE((bool)h->code_complex[c] == cplx, GNSS_B200_ERR_INVALID, "a batch cannot mix real and complex local codes");
            ChanParams& p = cp[i];
            if (device_resident)
                p.sig = d_sig_base + (size_t)i * ch_stride;
            else if (a.sig_in_c64)
                p.sig = h->d_sig.as<float2>() + (size_t)i * h->max_samples;
            else
                {
                    B200_REQUIRE(a.stream_offset + a.n_samples <= h->stream_len, GNSS_B200_ERR_INVALID, "stream window out of range");
                    p.sig = h->d_stream.as<float2>() + a.stream_offset;
                }
            p.code = h->codes[c].p;
            p.code_len = h->code_len[c];
            p.n = a.n_samples;
            p.rem_code = a.rem_code_phase_chips;
            p.code_step = a.code_step_chips;
            p.code_rate = a.code_rate_chips;
            // phase_offset_as_complex = (cos(rem), -sin(rem)) in binary32 (:114-115)
            float s0, c0;
            sincosf(a.rem_carr_phase_rad, &s0, &c0);
            p.p0_re = c0;
            p.p0_im = -s0;
            // the reference's per-sample phasor std::exp(lv_32fc_t(0, -step)) rounded to binary32; its true angle
            // (theta = atan2 of these, and the stride phasor exp(j*kThreads*theta), are derived on the device by
            // prep_kernel: wb_re/wb_im carry the rounded phasor until then)
            // std::exp(complex<float>(0, y)) is (cosf y, sinf y) exactly (glibc cexpf: sincosf times exp(0) = 1); calling
            // sincosf directly gives the same bits at a quarter of the cost
            float inc_re, inc_im;
            sincosf(-a.carr_step_rad, &inc_im, &inc_re);
            p.theta = 0.0;
            p.wb_re = inc_re;
            p.wb_im = inc_im;
            p.rate_arg = 0.0f;
            if (hd_rot)
                {
                    // cpowf(z, m) = cexpf(m * clogf(z)); after normalisation only arg z survives
                    const std::complex<float> inc_rate = std::exp(std::complex<float>(0.0f, -a.carr_rate_rad));
                    p.rate_arg = std::log(inc_rate).imag();
                }
            p.flags = (hd_res ? F_HIGH_DYN_RESAMPLER : 0u) | (hd_rot ? F_HIGH_DYN_ROTATOR : 0u) | (cplx ? F_COMPLEX_CODE : F_CODE_DOUBLED) | (exact_only ? F_EXACT_ONLY : 0u);
            p.shifts_off = i * n_taps;
            const float* hsi = hs + (size_t)i * n_taps;
            if (hd_res)
                {
                    // …high_dynamics_resampler…:84-87: cumulative integer sample shifts of taps 1.. relative to tap 0
                    uint32_t acc = 0;
                    hss[(size_t)i * n_taps] = 0;
                    for (uint32_t t = 1; t < n_taps; t++)
                        {
                            const float d = (hsi[t] - hsi[t - 1]) / a.code_step_chips;
                            acc += (uint32_t)(int)std::round((double)d);
                            hss[(size_t)i * n_taps + t] = acc;
                        }
                }
            float smin = hsi[0], smax = hsi[0];
            for (uint32_t t = 1; t < n_taps; t++)
                {
                    smin = hsi[t] < smin ? hsi[t] : smin;
                    smax = hsi[t] > smax ? hsi[t] : smax;
                }
            // |chip phase| stays below 2^22 (the packed floor of the window kernel is exact there)
            const double lim = std::max(std::fabs((double)smin), std::fabs((double)smax)) +
                               std::fabs((double)a.code_step_chips) * (double)a.n_samples + std::fabs((double)a.rem_code_phase_chips);
            if (!(lim < 4.0e6)) small_phase = false;
            const uint32_t per = (a.n_samples + splits - 1) / splits;
            const double span = std::fabs((double)a.code_step_chips) * (double)per + ((double)smax - (double)smin) + 16.0;
            max_window = std::max<uint64_t>(max_window, (uint64_t)span);
            max_len = std::max<uint64_t>(max_len, p.code_len);
        }
    const uint32_t elem = cplx ? 8u : 4u;
    const uint32_t cap = kSmemCapBytes / elem;
    // fast (packed, window-only) kernel: real code, plain resampler, window of (c,c) pairs fits, |chip phase| < 2^22
    const bool fast = !cplx && !hd_res && small_phase && max_window + 8 <= (kSmemCapBytes - kWinRingBytes) / 4 && max_n < (1u << 24) &&
                      (max_n + splits - 1) / splits <= kStMaxSegment;
    if (out_fast) *out_fast = fast;
    if (fast)
        {
            *out_smem_elems = (uint32_t)std::max<uint64_t>(max_window, 32);
            *out_splits = splits;
            *out_complex = false;
            return GNSS_B200_OK;
        }
    uint64_t want = hd_res ? max_len : max_window;
    if (!hd_res && want > cap && max_len <= cap) want = max_len;  // window too wide: hold the whole code instead
    *out_smem_elems = (uint32_t)std::min<uint64_t>(std::max<uint64_t>(want, 32), cap);
    *out_splits = splits;
    *out_complex = cplx;
    return GNSS_B200_OK;
}

int run_batch(gnss_b200_mc* h, uint32_t n_ch, uint32_t n_taps, uint32_t smem_elems, uint32_t splits, bool cplx, bool fast)
{
    const size_t pbytes = sizeof(ChanParams) * n_ch + sizeof(float) * n_ch * n_taps + sizeof(uint32_t) * n_ch * n_taps;
    B200_CUDA(cudaMemcpyAsync(h->d_params.p, h->h_params.p, pbytes, cudaMemcpyHostToDevice, h->stream));
    B200_CUDA(h->d_out.reserve(sizeof(float2) * (size_t)n_ch * n_taps));
    B200_CUDA(h->h_out.reserve(sizeof(float2) * (size_t)n_ch * n_taps));
    if (splits > 1) B200_CUDA(h->d_partials.reserve(sizeof(float2) * (size_t)n_ch * splits * n_taps));
    const ChanParams* d_cp = h->d_params.as<ChanParams>();
    const float* d_shifts = reinterpret_cast<const float*>(d_cp + n_ch);
    const uint32_t* d_ss = reinterpret_cast<const uint32_t*>(d_shifts + (size_t)n_ch * n_taps);
    const dim3 grid(splits, n_ch);
    const size_t smem = (size_t)smem_elems * (cplx ? 8 : 4);
    h->prof.begin(1, h->stream);
    prep_kernel<<<(n_ch + 127) / 128, 128, 0, h->stream>>>(h->d_params.as<ChanParams>(), n_ch);
    h->prof.end(h->stream);
    h->launches++;
    B200_CUDA(cudaGetLastError());
    for (uint32_t tap0 = 0; tap0 < n_taps; tap0 += kMaxTapsPerPass)
        {
            const int T = (int)std::min<uint32_t>(kMaxTapsPerPass, n_taps - tap0);
            h->prof.begin(0, h->stream);
            if (fast)
                launch_window(T, grid, (size_t)(smem_elems + 8) * 4 + kWinRingBytes, h->stream, d_cp, d_shifts, h->d_partials.as<float2>(), h->d_tickets.as<uint32_t>(),
                    h->d_out.as<float2>(), (int)n_taps, (int)tap0, smem_elems);
            else if (cplx)
                launch_taps<float2>(T, grid, smem, h->stream, d_cp, d_shifts, d_ss, h->d_partials.as<float2>(), h->d_tickets.as<uint32_t>(),
                    h->d_out.as<float2>(), (int)n_taps, (int)tap0, smem_elems);
            else
                launch_taps<float>(T, grid, smem, h->stream, d_cp, d_shifts, d_ss, h->d_partials.as<float2>(), h->d_tickets.as<uint32_t>(),
                    h->d_out.as<float2>(), (int)n_taps, (int)tap0, smem_elems);
            h->prof.end(h->stream);
            h->launches++;
            B200_CUDA(cudaGetLastError());
        }
    B200_CUDA(cudaMemcpyAsync(h->h_out.p, h->d_out.p, sizeof(float2) * (size_t)n_ch * n_taps, cudaMemcpyDeviceToHost, h->stream));
    h->last_n_ch = n_ch;
    h->last_n_taps = n_taps;
    return GNSS_B200_OK;
}
}  // namespace

extern "C"
{
    int gnss_b200_mc_create(uint32_t max_channels, uint32_t max_samples, uint32_t max_taps, int32_t device, gnss_b200_mc** out)
    {
        B200_REQUIRE(out && max_channels > 0 && max_samples > 0 && max_taps > 0, GNSS_B200_ERR_INVALID, "mc_create: bad arguments");
        int rc = select_device(device);
        if (rc) return rc;
        auto* h = new gnss_b200_mc();
        h->device = device;
        h->max_channels = max_channels;
        h->max_samples = max_samples;
        h->max_taps = max_taps;
        h->codes.resize(max_channels);
        h->code_len.assign(max_channels, 0);
        h->code_complex.assign(max_channels, 0);
        cudaError_t e = cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking);
        if (e == cudaSuccess) e = h->d_tickets.reserve(sizeof(uint32_t) * max_channels);
        if (e == cudaSuccess) e = cudaMemsetAsync(h->d_tickets.p, 0, sizeof(uint32_t) * max_channels, h->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
        if (e != cudaSuccess)
            {
                delete h;
                return fail(GNSS_B200_ERR_CUDA, "mc_create: %s", cudaGetErrorString(e));
            }
        *out = h;
        return GNSS_B200_OK;
    }

    int gnss_b200_mc_destroy(gnss_b200_mc* h)
    {
        if (!h) return GNSS_B200_OK;
        cudaSetDevice(h->device);
        if (h->stream) cudaStreamSynchronize(h->stream);
        for (auto& c : h->codes) c.release();
        h->d_sig.release();
        h->d_stream.release();
        h->d_params.release();
        h->h_params.release();
        h->d_partials.release();
        h->d_tickets.release();
        h->d_out.release();
        h->h_out.release();
        h->d_idx.release();
        h->prof.release();
        if (h->stream) cudaStreamDestroy(h->stream);
        delete h;
        return GNSS_B200_OK;
    }

    int gnss_b200_mc_set_code(gnss_b200_mc* h, uint32_t ch, uint32_t code_len, const float* code, int is_complex)
    {
        B200_REQUIRE(h && code && code_len > 0, GNSS_B200_ERR_INVALID, "mc_set_code: bad arguments");
        B200_REQUIRE(ch < h->max_channels, GNSS_B200_ERR_INVALID, "mc_set_code: channel %u out of range", ch);
        std::lock_guard<std::mutex> lk(h->mu);
        B200_CUDA(cudaSetDevice(h->device));
        if (is_complex)
            {
                const size_t bytes = (size_t)code_len * 8;
                B200_CUDA(h->codes[ch].reserve(bytes));
                B200_CUDA(cudaMemcpyAsync(h->codes[ch].p, code, bytes, cudaMemcpyHostToDevice, h->stream));
                B200_CUDA(cudaStreamSynchronize(h->stream));
            }
        else
            {
                B200_CUDA(h->codes[ch].reserve(mc::doubled_code_floats(code_len) * 4));
                B200_CUDA(mc::upload_doubled_code(h->codes[ch].p, code, code_len, h->stream));
            }
        h->code_len[ch] = code_len;
        h->code_complex[ch] = is_complex ? 1 : 0;
        return GNSS_B200_OK;
    }

    int gnss_b200_mc_set_stream(gnss_b200_mc* h, const float* samples_c64, uint64_t n_samples)
    {
        B200_REQUIRE(h && samples_c64 && n_samples > 0, GNSS_B200_ERR_INVALID, "mc_set_stream: bad arguments");
        std::lock_guard<std::mutex> lk(h->mu);
        B200_CUDA(cudaSetDevice(h->device));
        B200_CUDA(h->d_stream.reserve(n_samples * 8));
        B200_CUDA(cudaMemcpyAsync(h->d_stream.p, samples_c64, n_samples * 8, cudaMemcpyHostToDevice, h->stream));
        h->stream_len = n_samples;
        return GNSS_B200_OK;
    }

    int gnss_b200_mc_correlate_batch(gnss_b200_mc* h, uint32_t n_ch, const uint32_t* ch, const gnss_b200_mc_args* args, const float* shifts,
        uint32_t n_taps, uint32_t flags, float* out_c64)
    {
        B200_REQUIRE(h && ch && args && shifts && out_c64, GNSS_B200_ERR_INVALID, "mc_correlate_batch: null argument");
        B200_REQUIRE(n_ch > 0 && n_ch <= h->max_channels && n_taps > 0 && n_taps <= h->max_taps, GNSS_B200_ERR_INVALID,
            "mc_correlate_batch: n_ch=%u n_taps=%u out of range", n_ch, n_taps);
        std::lock_guard<std::mutex> lk(h->mu);
        B200_CUDA(cudaSetDevice(h->device));
        bool any_host = false;
        for (uint32_t i = 0; i < n_ch; i++) any_host |= (args[i].sig_in_c64 != nullptr);
        if (any_host) B200_CUDA(h->d_sig.reserve((size_t)h->max_channels * h->max_samples * 8));
        uint32_t smem_elems, splits;
        bool cplx, fast;
        int rc = build_params(h, n_ch, ch, args, shifts, n_taps, flags, nullptr, 0, false, &smem_elems, &splits, &cplx, &fast);
        if (rc) return rc;
        // host -> device sample copies; runs of channels that are contiguous in host memory go as one copy
        for (uint32_t i = 0; i < n_ch;)
            {
                if (!args[i].sig_in_c64)
                    {
                        i++;
                        continue;
                    }
                uint32_t j = i + 1;
                while (j < n_ch && args[j].sig_in_c64 && args[j - 1].n_samples == h->max_samples &&
                       args[j].sig_in_c64 == args[j - 1].sig_in_c64 + 2 * (size_t)h->max_samples)
                    j++;
                const size_t bytes = (j - i > 1) ? (size_t)(j - i - 1) * h->max_samples * 8 + (size_t)args[j - 1].n_samples * 8 : (size_t)args[i].n_samples * 8;
                B200_CUDA(cudaMemcpyAsync(h->d_sig.as<float2>() + (size_t)i * h->max_samples, args[i].sig_in_c64, bytes, cudaMemcpyHostToDevice, h->stream));
                i = j;
            }
        rc = run_batch(h, n_ch, n_taps, smem_elems, splits, cplx, fast);
        if (rc) return rc;
        B200_CUDA(cudaStreamSynchronize(h->stream));
        std::memcpy(out_c64, h->h_out.p, sizeof(float2) * (size_t)n_ch * n_taps);
        return GNSS_B200_OK;
    }

    int gnss_b200_mc_correlate_device_async(gnss_b200_mc* h, uint32_t n_ch, const uint32_t* ch, const gnss_b200_mc_args* args, const float* shifts,
        uint32_t n_taps, uint32_t flags, const void* d_sig, uint64_t ch_stride_samples)
    {
        B200_REQUIRE(h && ch && args && shifts && d_sig, GNSS_B200_ERR_INVALID, "mc_correlate_device_async: null argument");
        B200_REQUIRE(n_ch > 0 && n_ch <= h->max_channels && n_taps > 0 && n_taps <= h->max_taps, GNSS_B200_ERR_INVALID,
            "mc_correlate_device_async: n_ch=%u n_taps=%u out of range", n_ch, n_taps);
        std::lock_guard<std::mutex> lk(h->mu);
        B200_CUDA(cudaSetDevice(h->device));
        uint32_t smem_elems, splits;
        bool cplx, fast;
        int rc = build_params(h, n_ch, ch, args, shifts, n_taps, flags, static_cast<const float2*>(d_sig), ch_stride_samples, true, &smem_elems, &splits, &cplx, &fast);
        if (rc) return rc;
        return run_batch(h, n_ch, n_taps, smem_elems, splits, cplx, fast);
    }

    int gnss_b200_mc_sync(gnss_b200_mc* h, uint32_t n_ch, uint32_t n_taps, float* out_c64)
    {
        B200_REQUIRE(h, GNSS_B200_ERR_INVALID, "mc_sync: null handle");
        std::lock_guard<std::mutex> lk(h->mu);
        B200_CUDA(cudaSetDevice(h->device));
        B200_CUDA(cudaStreamSynchronize(h->stream));
        if (out_c64)
            {
                B200_REQUIRE(n_ch == h->last_n_ch && n_taps == h->last_n_taps, GNSS_B200_ERR_INVALID, "mc_sync: shape differs from the last batch");
                std::memcpy(out_c64, h->h_out.p, sizeof(float2) * (size_t)n_ch * n_taps);
            }
        return GNSS_B200_OK;
    }

    int gnss_b200_mc_chip_indices(gnss_b200_mc* h, uint32_t ch, const gnss_b200_mc_args* args, const float* shifts, uint32_t n_taps, uint32_t flags, int32_t* idx_out)
    {
        B200_REQUIRE(h && args && shifts && idx_out && n_taps > 0, GNSS_B200_ERR_INVALID, "mc_chip_indices: bad arguments");
        std::lock_guard<std::mutex> lk(h->mu);
        B200_CUDA(cudaSetDevice(h->device));
        uint32_t smem_elems, splits;
        bool cplx;
        gnss_b200_mc_args a = *args;
        a.sig_in_c64 = nullptr;
        a.stream_offset = 0;
        const uint64_t keep = h->stream_len;
        h->stream_len = ~0ull;  // windows are not dereferenced by this kernel
        int rc = build_params(h, 1, &ch, &a, shifts, n_taps, flags, nullptr, 0, false, &smem_elems, &splits, &cplx);
        h->stream_len = keep;
        if (rc) return rc;
        const size_t pbytes = sizeof(ChanParams) + sizeof(float) * n_taps + sizeof(uint32_t) * n_taps;
        B200_CUDA(cudaMemcpyAsync(h->d_params.p, h->h_params.p, pbytes, cudaMemcpyHostToDevice, h->stream));
        const size_t obytes = sizeof(int32_t) * (size_t)n_taps * a.n_samples;
        B200_CUDA(h->d_idx.reserve(obytes));
        const ChanParams* d_cp = h->d_params.as<ChanParams>();
        const float* d_shifts = reinterpret_cast<const float*>(d_cp + 1);
        const uint32_t* d_ss = reinterpret_cast<const uint32_t*>(d_shifts + n_taps);
        chip_index_kernel<<<kNumSMs, 256, 0, h->stream>>>(*h->h_params.as<ChanParams>(), d_shifts, d_ss, (int)n_taps, h->d_idx.as<int32_t>());
        h->launches++;
        B200_CUDA(cudaGetLastError());
        B200_CUDA(cudaMemcpyAsync(idx_out, h->d_idx.p, obytes, cudaMemcpyDeviceToHost, h->stream));
        B200_CUDA(cudaStreamSynchronize(h->stream));
        return GNSS_B200_OK;
    }

    int gnss_b200_mc_profile(gnss_b200_mc* h, int enable, double* ms_by_kernel, uint64_t* launches_by_kernel, int reset)
    {
        B200_REQUIRE(h, GNSS_B200_ERR_INVALID, "mc_profile: null handle");
        std::lock_guard<std::mutex> lk(h->mu);
        B200_CUDA(cudaSetDevice(h->device));
        B200_CUDA(cudaStreamSynchronize(h->stream));
        h->prof.collect();
        for (int i = 0; i < GNSS_B200_PROFILE_SLOTS; i++)
            {
                if (ms_by_kernel) ms_by_kernel[i] = h->prof.ms[i];
                if (launches_by_kernel) launches_by_kernel[i] = h->prof.count[i];
            }
        if (reset) h->prof.reset();
        h->prof.on = enable != 0;
        return GNSS_B200_OK;
    }

    void* gnss_b200_mc_stream(gnss_b200_mc* h) { return h ? (void*)h->stream : nullptr; }
    uint64_t gnss_b200_mc_launch_count(const gnss_b200_mc* h) { return h ? h->launches : 0; }
}

// ------------------------------------------------------------------------------------------------------------------
// gnss_b200_mc_correlate_noalloc: the multicorrelator for callers that may not allocate (CUDA graph capture, tight
// serving loops, harnesses that forbid cudaMalloc). No counterpart in the reference.
// Same kernels as gnss_b200_mc_correlate_batch (general real-code kernel, plain resampler); everything the
// handle-based path keeps in cudaMalloc'ed buffers — the per-channel parameter blocks, the tap shifts, the per-CTA split
// partials and the per-channel tickets — lives in ONE file-scope device array of fixed size instead. Every GPU that loads
// the library holds that array (kNoAllocScratchBytes, ~141 KB) until the process exits.
//
// In host code `g_mc_noalloc_scratch` is only a stand-in for the symbol, not a device address: the address is asked for
// with cudaGetSymbolAddress once per device, with that device current, and only that pointer is handed to kernels.
// Every call on a device shares the array and finds the previous call's contents in it: a fill kernel rewrites the parameter
// blocks and shifts and clears the tickets on the stream before each launch, the partials are fully overwritten before they
// are read. Two calls whose GPU work overlaps (different streams, or different host threads) race on it.
// The call issues kernel launches only (no memcpy / memset nodes, no host pointer outlives the call), so it may be recorded
// under stream capture and replayed (tests/test_zz_no_alloc.py does): a replay re-runs the SAME parameters on whatever the
// caller's device buffers hold then. The first call on a device resolves the scratch address: make it before capturing.
// ------------------------------------------------------------------------------------------------------------------
namespace b200_noalloc  // (a named namespace: nvcc cannot tell a kernel of a second anonymous namespace in this file from the first)
{
constexpr uint32_t kNoAllocMaxChannels = GNSS_B200_MC_NOALLOC_MAX_CHANNELS;
constexpr uint32_t kNoAllocMaxTaps = GNSS_B200_MC_NOALLOC_MAX_TAPS;
constexpr uint32_t kNoAllocMaxSamples = GNSS_B200_MC_NOALLOC_MAX_SAMPLES;
constexpr uint32_t kNoAllocMaxCtas = 1024;  // channels x splits
constexpr uint32_t kNoAllocSmemElems = 8192;  // 32 KB of chips per CTA (window, whole code, or global fallback)
constexpr int kNoAllocMaxDevices = 64;
static_assert(kNoAllocMaxTaps <= (uint32_t)kMaxTapsPerPass, "one pass per call");

struct NoAllocScratch
{
    ChanParams params[kNoAllocMaxChannels];
    float shifts[kNoAllocMaxChannels * kNoAllocMaxTaps];
    uint32_t shift_samples[kNoAllocMaxChannels * kNoAllocMaxTaps];  // stays zero: plain resampler
    uint32_t tickets[kNoAllocMaxChannels];
    float2 partials[kNoAllocMaxCtas * kNoAllocMaxTaps];
};
constexpr size_t kNoAllocScratchBytes = sizeof(NoAllocScratch);
static_assert(kNoAllocScratchBytes < 160 * 1024, "keep the resident scratch modest");

static __device__ __align__(256) unsigned char g_mc_noalloc_scratch[kNoAllocScratchBytes];

// Parameter blocks and tap shifts travel as a by-value kernel argument (copied into the launch, or into the graph node under
// stream capture) and a tiny kernel writes them into the resident scratch: no host-memory copy node, nothing the caller has to
// keep alive after the call returns. 16 channels per launch keeps the argument under the 4 KB every driver accepts.
constexpr uint32_t kNoAllocFillChannels = 16;
struct NoAllocFill
{
    ChanParams cp[kNoAllocFillChannels];
    float shifts[kNoAllocFillChannels * kNoAllocMaxTaps];
    uint32_t first, count, n_taps, pad;
};
static_assert(sizeof(NoAllocFill) <= 4000, "kernel argument size");

__global__ void noalloc_fill_kernel(const __grid_constant__ NoAllocFill f, NoAllocScratch* scr)
{
    const uint32_t t = threadIdx.x;
    if (t < f.count)
        {
            scr->params[f.first + t] = f.cp[t];
            scr->tickets[f.first + t] = 0u;  // must start at zero whatever the last call left
        }
    for (uint32_t i = t; i < f.count * f.n_taps; i += blockDim.x) scr->shifts[(size_t)f.first * f.n_taps + i] = f.shifts[i];
}

std::mutex g_noalloc_mu;
NoAllocScratch* g_noalloc_ptr[kNoAllocMaxDevices] = {nullptr};

// device address of the scratch on `device` (current): resolved once per device
int noalloc_scratch(int device, NoAllocScratch** out)
{
    B200_REQUIRE(device >= 0 && device < kNoAllocMaxDevices, GNSS_B200_ERR_INVALID, "mc_correlate_noalloc: device %d out of range", device);
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    if (g_noalloc_ptr[device] == nullptr)
        {
            void* p = nullptr;
            B200_CUDA(cudaGetSymbolAddress(&p, g_mc_noalloc_scratch));
            B200_REQUIRE(p != nullptr, GNSS_B200_ERR_CUDA, "mc_correlate_noalloc: cudaGetSymbolAddress returned a null address");
            g_noalloc_ptr[device] = static_cast<NoAllocScratch*>(p);
        }
    *out = g_noalloc_ptr[device];
    return GNSS_B200_OK;
}
}  // namespace b200_noalloc
using namespace b200_noalloc;

extern "C"
{
    int gnss_b200_mc_correlate_noalloc(int32_t device, void* cuda_stream, uint32_t n_ch, const gnss_b200_mc_noalloc_chan* chans, const float* shifts,
        uint32_t n_taps, void* d_out_c64)
    {
        B200_REQUIRE(chans && shifts && d_out_c64, GNSS_B200_ERR_INVALID, "mc_correlate_noalloc: null argument");
        B200_REQUIRE(n_ch >= 1 && n_ch <= kNoAllocMaxChannels, GNSS_B200_ERR_INVALID,
            "mc_correlate_noalloc: %u channels, the resident scratch holds at most %u per call", n_ch, kNoAllocMaxChannels);
        B200_REQUIRE(n_taps >= 1 && n_taps <= kNoAllocMaxTaps, GNSS_B200_ERR_INVALID,
            "mc_correlate_noalloc: %u taps, the resident scratch holds at most %u", n_taps, kNoAllocMaxTaps);
        uint32_t max_n = 1;
        for (uint32_t i = 0; i < n_ch; i++)
            {
                B200_REQUIRE(chans[i].d_sig_c64 && chans[i].d_code && chans[i].code_len >= 1, GNSS_B200_ERR_INVALID, "mc_correlate_noalloc: channel %u: null device pointer or empty code", i);
                B200_REQUIRE(chans[i].n_samples >= 1 && chans[i].n_samples <= kNoAllocMaxSamples, GNSS_B200_ERR_INVALID,
                    "mc_correlate_noalloc: channel %u: %u samples, at most %u per call", i, chans[i].n_samples, kNoAllocMaxSamples);
                max_n = std::max(max_n, chans[i].n_samples);
            }
        int caller_device = -1;
        cudaGetDevice(&caller_device);
        int rc = select_device(device);
        if (rc) return rc;
        struct RestoreDevice  // the caller's current device is the caller's business
        {
            int d;
            ~RestoreDevice()
            {
                if (d >= 0) cudaSetDevice(d);
            }
        } restore{caller_device};
        NoAllocScratch* scr = nullptr;
        rc = noalloc_scratch(device, &scr);
        if (rc) return rc;
        cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);

        // split each channel over several CTAs when the batch alone cannot fill the machine (as build_params does)
        uint32_t splits = 1;
        const uint32_t target_ctas = kNumSMs * 4;
        if (n_ch < target_ctas)
            {
                splits = (target_ctas + n_ch - 1) / n_ch;
                splits = std::max(1u, std::min(splits, std::max(1u, max_n / 2048)));
            }
        splits = std::max(1u, std::min(splits, kNoAllocMaxCtas / n_ch));

        // per-channel constants (cpu_multicorrelator_real_codes.cc:114-119) go out as kernel arguments, 16 channels at a time
        for (uint32_t first = 0; first < n_ch; first += kNoAllocFillChannels)
            {
                NoAllocFill f;
                std::memset(&f, 0, sizeof(f));
                f.first = first;
                f.count = std::min(kNoAllocFillChannels, n_ch - first);
                f.n_taps = n_taps;
                for (uint32_t j = 0; j < f.count; j++)
                    {
                        const uint32_t i = first + j;
                        const gnss_b200_mc_noalloc_chan& a = chans[i];
                        ChanParams& p = f.cp[j];
                        p.sig = static_cast<const float2*>(a.d_sig_c64);
                        p.code = a.d_code;
                        p.code_len = a.code_len;
                        p.n = a.n_samples;
                        p.rem_code = a.rem_code_phase_chips;
                        p.code_step = a.code_step_chips;
                        p.code_rate = 0.f;
                        float s0, c0;
                        sincosf(a.rem_carr_phase_rad, &s0, &c0);
                        p.p0_re = c0;
                        p.p0_im = -s0;
                        float inc_re, inc_im;
                        sincosf(-a.carr_step_rad, &inc_im, &inc_re);  // the rounded per-sample phasor; prep_kernel derives theta from it
                        p.theta = 0.0;
                        p.wb_re = inc_re;
                        p.wb_im = inc_im;
                        p.rate_arg = 0.f;
                        p.flags = 0u;
                        p.shifts_off = i * n_taps;
                        std::memcpy(&f.shifts[(size_t)j * n_taps], shifts + (size_t)i * n_taps, sizeof(float) * n_taps);
                    }
                noalloc_fill_kernel<<<1, 256, 0, st>>>(f, scr);
                B200_CUDA(cudaGetLastError());
            }
        prep_kernel<<<(n_ch + 127) / 128, 128, 0, st>>>(scr->params, n_ch);
        B200_CUDA(cudaGetLastError());
        B200_CUDA(mc::launch_real_plain(false, n_ch, splits, kNoAllocSmemElems, st, scr->params, scr->shifts, scr->shift_samples, scr->partials, scr->tickets,
            static_cast<float2*>(d_out_c64), n_taps));
        return GNSS_B200_OK;
    }

    uint64_t gnss_b200_mc_noalloc_scratch_bytes(void) { return (uint64_t)kNoAllocScratchBytes; }
}
