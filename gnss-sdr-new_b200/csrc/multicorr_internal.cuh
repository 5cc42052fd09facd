// Internal interface of the multicorrelator kernels (multicorr.cu) for the on-device tracking loop (trk_loop.cu):
// the per-channel parameter block the kernels read, and a launcher that takes parameter blocks that already live
// in device memory (written there by the loop-update kernel, no host round trip between epochs).
#pragma once
#include "common.cuh"

namespace b200
{
namespace mc
{
constexpr int kThreads = 256;
constexpr int kMaxTapsPerPass = 16;
constexpr uint32_t kSmemCapBytes = 200 * 1024;
constexpr int kWinGroup = 4;  // samples per cp.async commit group (per thread)
constexpr int kWinRing = 4;   // commit groups in the per-thread shared-memory ring
constexpr uint32_t kWinRingBytes = kWinRing * kWinGroup * kThreads * 8u;

enum : uint32_t
{
    F_HIGH_DYN_RESAMPLER = 1u,
    F_HIGH_DYN_ROTATOR = 2u,
    F_COMPLEX_CODE = 4u
};

struct alignas(16) ChanParams
{
    const float2* sig;   // device samples of this channel
    const void* code;    // device code table (float or float2)
    double theta;        // effective rotation per sample [rad]
    uint32_t code_len;
    uint32_t n;
    float rem_code, code_step, code_rate;
    float p0_re, p0_im;  // (cos rem_carr, -sin rem_carr)
    float wb_re, wb_im;  // exp(j*kThreads*theta)
    float rate_arg;      // arg of the rounded rate phasor (high-dynamics rotator)
    uint32_t flags;
    uint32_t shifts_off;  // offset into the shifts / shift_samples arrays
};

// Real codes, plain resampler. `fast` selects the packed window kernel (the caller guarantees that every CTA's chip
// window fits smem_elems floats and |chip phase| < 2^22); otherwise the general kernel with smem_elems code elements.
// d_cp already holds theta / wb (no prep pass). d_shift_samples may be any valid device pointer when !high-dynamics.
cudaError_t launch_real_plain(bool fast, uint32_t n_ch, uint32_t splits, uint32_t smem_elems, cudaStream_t st, const ChanParams* d_cp,
    const float* d_shifts, const uint32_t* d_shift_samples, float2* d_partials, uint32_t* d_tickets, float2* d_out, uint32_t n_taps);
}  // namespace mc
}  // namespace b200
