// Fast forward path of the acquisition pipeline for "fast" FFT plans (fft_plan.h): the exact mirror of the inverse
// kernels of acq_fast.cuh — decimation in frequency, upper (strided) stage first, two register passes per stage
// around one shared-memory transpose, every index constant-folded through the radix template parameters.
//
//   fwd_upper_kernel<RA,RB> : samples (gr_complex / cshort / cbyte, zero padded, pcps_acquisition.cc:640-650) x Doppler
//                             wipe-off with the reference's binary32 running-sum phase (:658, update_local_carrier
//                             :233-246) -> DFT over the strided upper stage (M_hi = RA*RB) -> inter-stage twiddle -> scratch
//   fwd_lower_kernel<RA,RB> : scratch -> DFT over the contiguous lower stage (M_lo = RA*RB) -> X_k in the coalescing
//                             frequency layout F(pos) the inverse reads (:662)
//
// Index algebra (n = m + M_lo*t, t = q + ra_hi*b, m = a + ra_lo*b'; k = k1 + M_hi*j, k1 = uah*rb_hi + ubh,
// j = ub + rb_lo*ua): n*k/N = q*uah/ra_hi + q*ubh/M_hi + b*ubh/rb_hi + m*k1/N + a*ub/M_lo + a*ua/ra_lo + b'*ub/rb_lo (mod 1).
#pragma once
#include "acq_fast.cuh"

namespace b200
{
struct FastFwdArgs
{
    const float2* W;  // exp(-2 pi i k / N), k < N
    uint32_t N, M_lo, M_hi;
    uint32_t ra_lo, rb_lo, ra_hi, rb_hi;
    const void* samples;
    uint32_t fmt;  // GNSS_B200_SAMPLES_*
    uint32_t consumed;
    const WipePiece* pieces;           // [group][nD_local][kMaxPieces]
    const uint32_t* n_pieces;          // [group][nD_local]
    const uint32_t* group_piece_base;  // [call group] -> first bin in pieces / n_pieces
    uint32_t nD_local;
    float2* scratch;  // [transform][N]: position blk*M_lo + m between the stages
    float2* X;        // [transform][N], fast layout
    uint32_t nb;      // lower stage: contiguous sub-transforms per tile
    uint32_t C;       // upper stage: columns per tile
};

// upper stage: grid (transforms, M_lo / C), block (C, threads / C). Shared memory: M * C data + M twiddles.
// MLO, CC != 0: lower-stage length and column tile compile-time (block (CC, kFastThreads / CC)), as in acq_fast.cuh
template <int RA, int RB, int MLO = 0, int CC = 0>
__global__ void __launch_bounds__(kFastThreads) fwd_upper_kernel(const FastFwdArgs a)
{
    constexpr int M = RA * RB;
    const uint32_t gC = CC ? CC : a.C, gM_lo = MLO ? MLO : a.M_lo, gN = MLO ? (uint32_t)MLO * M : a.N;
    const uint32_t bdx = CC ? CC : blockDim.x, bdy = CC ? kFastThreads / CC : blockDim.y;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float2* s = reinterpret_cast<float2*>(smem_raw);
    float2* s_tw = s + (size_t)M * gC;  // W_M^(q*ub), [ub][q]
    __shared__ WipePiece s_pieces[kMaxPieces];

    const uint32_t y = blockIdx.x;
    const uint32_t g = y / a.nD_local, kk = y - g * a.nD_local;
    const uint32_t pb = a.group_piece_base[g] + kk;
    const uint32_t np = a.n_pieces[pb];
    const uint32_t C = gC, c = threadIdx.x, m = blockIdx.y * C + c;
    const uint32_t tid = threadIdx.y * bdx + threadIdx.x, nthreads = bdx * bdy;
    const uint32_t tw_scale = gN / M;
    for (uint32_t i = tid; i < np; i += nthreads) s_pieces[i] = a.pieces[(size_t)pb * kMaxPieces + i];
    for (uint32_t i = tid; i < (uint32_t)M; i += nthreads)
        {
            const uint32_t ub = i / RA, q = i - ub * RA;
            s_tw[i] = __ldg(a.W + (size_t)q * ub * tw_scale);
        }
    __syncthreads();

    // ---- pass 1 (radix RB over b, stride RA*M_lo): pencils (column c, q) ----
    for (uint32_t q = threadIdx.y; q < (uint32_t)RA; q += bdy)
        {
            float2 v[RB];
            uint32_t cur = 0;
#pragma unroll
            for (int b = 0; b < RB; b++)
                {
                    const uint32_t pos = m + gM_lo * (q + RA * b);
                    float2 w = make_float2(0.f, 0.f);
                    if (pos < a.consumed)
                        {
                            float2 x;
                            if (a.fmt == 1)
                                {
                                    const short2 t = __ldg(static_cast<const short2*>(a.samples) + pos);  // :623-626
                                    x = make_float2((float)t.x, (float)t.y);
                                }
                            else if (a.fmt == 2)
                                {
                                    const char2 t = __ldg(static_cast<const char2*>(a.samples) + pos);
                                    x = make_float2((float)t.x, (float)t.y);
                                }
                            else
                                {
                                    x = __ldg(static_cast<const float2*>(a.samples) + pos);
                                }
                            while (cur + 1 < np && pos >= s_pieces[cur + 1].start) cur++;
                            const WipePiece pc = s_pieces[cur];
                            const float phi = fmaf((float)(pos - pc.start), pc.step, pc.phi0);  // exact inside a piece
                            float sn, cs;
                            wipeoff_sincos(phi, sn, cs);  // volk_gnsssdr_s32f_sincos_32fc
                            w = make_float2(x.x * cs - x.y * sn, x.x * sn + x.y * cs);  // :658
                        }
                    v[b] = w;
                }
            fft::Dft<RB, false>::run(v);
#pragma unroll
            for (int ub = 0; ub < RB; ub++) s[(size_t)(q + RA * ub) * C + c] = fft::cmul(v[ub], s_tw[ub * RA + q]);
        }
    __syncthreads();

    // ---- pass 2 (radix RA over q) + inter-stage twiddle W_N^(k1*m), k1 = uah*RB + ub: pencils (column c, ub) ----
    float2* out = a.scratch + (size_t)y * gN + m;
    for (uint32_t ub = threadIdx.y; ub < (uint32_t)RB; ub += bdy)
        {
            float2 v[RA];
#pragma unroll
            for (int q = 0; q < RA; q++) v[q] = s[(size_t)(q + RA * ub) * C + c];
            fft::Dft<RA, false>::run(v);
            // W_N^(k1*m) = W_N^(ub*m) * (W_N^(RB*m))^uah: the table is read at every 8th uah only (a gather: neighbouring
            // columns sit k1 entries apart) and stepped in between, which keeps the recurrence error below 1e-6
            const float2 st = __ldg(a.W + (size_t)RB * m);  // RB * m < N
            float2 tw = make_float2(1.f, 0.f);
#pragma unroll
            for (int uah = 0; uah < RA; uah++)
                {
                    if (uah % 8 == 0)
                        tw = __ldg(a.W + (size_t)(uah * RB + ub) * m);  // k1 * m < N
                    else
                        tw = fft::cmul(tw, st);
                    out[(size_t)(uah + RA * ub) * gM_lo] = fft::cmul(v[uah], tw);
                }
        }
}

// lower stage: grid (transforms, M_hi / nb). Shared memory: nb * PITCH * RB data + M twiddles.
// MHI, NB != 0: upper-stage length and sub-transforms per tile compile-time (kFastThreads threads)
template <int RA, int RB, int MHI = 0, int NB = 0>
__global__ void __launch_bounds__(kFastThreads) fwd_lower_kernel(const FastFwdArgs a)
{
    constexpr int M = RA * RB;
    const uint32_t gnb = NB ? NB : a.nb, gN = MHI ? (uint32_t)MHI * M : a.N, bd = NB ? kFastThreads : blockDim.x;
    constexpr int P = (RA % 2 == 0) ? RA + 1 : RA;  // odd pitch: conflict-free 64-bit accesses with stride P
    constexpr int MP = P * RB;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float2* s = reinterpret_cast<float2*>(smem_raw);
    float2* s_tw = s + (size_t)gnb * MP;  // W_M^(a*ub), [ub][a]

    const uint32_t y = blockIdx.x;
    const uint32_t blk0 = blockIdx.y * gnb;
    const uint32_t tw_scale = gN / M;
    const float2* in = a.scratch + (size_t)y * gN + (size_t)blk0 * M;
    for (uint32_t i = threadIdx.x; i < (uint32_t)M; i += bd)
        {
            const uint32_t ub = i / RA, aa = i - ub * RA;
            s_tw[i] = __ldg(a.W + (size_t)aa * ub * tw_scale);
        }
    __syncthreads();

    // ---- pass 1 (radix RB over b', stride RA): pencils (block, a) ----
    const uint32_t n_pen_1 = gnb * RA;
    for (uint32_t p = threadIdx.x; p < n_pen_1; p += bd)
        {
            const uint32_t bl = p / RA, aa = p - bl * RA;
            float2 v[RB];
            const float2* src = in + (size_t)bl * M + aa;
#pragma unroll
            for (int b = 0; b < RB; b++) v[b] = __ldcg(src + b * RA);
            fft::Dft<RB, false>::run(v);
            float2* dst = s + (size_t)bl * MP + aa;
#pragma unroll
            for (int ub = 0; ub < RB; ub++) dst[(size_t)ub * P] = fft::cmul(v[ub], s_tw[ub * RA + aa]);
        }
    __syncthreads();

    // ---- pass 2 (radix RA over a, stride 1): pencils (block, ub); bin j = ub + RB*ua goes to F(pos) ----
    const uint32_t n_pen_2 = gnb * RB;
    const uint32_t seg = gN / RA;
    float2* X = a.X + (size_t)y * gN;
    for (uint32_t p = threadIdx.x; p < n_pen_2; p += bd)
        {
            const uint32_t bl = p / RB, ub = p - bl * RB;
            float2 v[RA];
            const float2* src = s + (size_t)bl * MP + (size_t)ub * P;
#pragma unroll
            for (int q = 0; q < RA; q++) v[q] = src[q];
            fft::Dft<RA, false>::run(v);
            float2* dst = X + (size_t)(blk0 + bl) * RB + ub;
#pragma unroll
            for (int ua = 0; ua < RA; ua++) dst[(size_t)ua * seg] = v[ua];
        }
}

cudaError_t launch_fwd_upper(uint32_t ra, uint32_t rb, dim3 grid, dim3 block, size_t smem, cudaStream_t st, const FastFwdArgs& a);
cudaError_t launch_fwd_lower(uint32_t ra, uint32_t rb, dim3 grid, uint32_t threads, size_t smem, cudaStream_t st, const FastFwdArgs& a);

}  // namespace b200
