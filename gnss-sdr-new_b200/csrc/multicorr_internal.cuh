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
// fast (window) kernel: every warp streams ITS contiguous part of a chunk through a private ring of bulk copies
constexpr int kStRows = 6;                                   // 32-sample rows of one warp per chunk
constexpr int kStPart = kStRows * 32;                        // samples of one warp per chunk (contiguous in memory: 1.5 KB)
constexpr int kStChunk = kStPart * (kThreads / 32);          // samples of the CTA per chunk
constexpr int kStRing = 2;                                   // ring slots per warp (kStRing - 1 bulk copies in flight)
constexpr int kStMaxChunks = 86;                             // chunks per CTA segment the carrier table can hold
constexpr uint32_t kStMaxSegment = kStMaxChunks * kStChunk;  // samples per CTA segment
constexpr uint32_t kStRingBytes = (kThreads / 32) * kStRing * kStPart * 8u;
constexpr int kStTailEntries = kStRows * (kThreads / 32);   // carrier-table entries of the partial chunk: [row][warp]
constexpr uint32_t kStTableBytes = (kStMaxChunks * kStRows + kStTailEntries + 2) * 8u;  // + the peeled sample
constexpr uint32_t kStBarBytes = ((kThreads / 32) * kStRing + 2) * 8u;  // ring barriers + the chip window's
constexpr uint32_t kWinRingBytes = kStRingBytes + kStTableBytes + kStBarBytes;  // shared memory of the fast kernel besides the chip window
static_assert(kWinRingBytes % 16 == 0, "the chip window behind it is filled with 4-byte cp.async");

enum : uint32_t
{
    F_HIGH_DYN_RESAMPLER = 1u,
    F_HIGH_DYN_ROTATOR = 2u,
    F_COMPLEX_CODE = 4u,
    F_EXACT_ONLY = 16u,  // parity aid: the window kernel never takes its short chip-phase form (B200_MC_EXACT_ONLY=1)
    F_CODE_DOUBLED = 8u  // real code stored as code | code | code[0..3]: any window of <= code_len + 4 chips is one contiguous range
};

constexpr uint32_t kCodeDoubledPad = 4;
inline size_t doubled_code_floats(uint32_t code_len) { return 2 * (size_t)code_len + kCodeDoubledPad; }
// host helper: uploads a real code in the doubled layout (dst must hold doubled_code_floats(code_len) floats)
cudaError_t upload_doubled_code(void* d_dst, const float* h_code, uint32_t code_len, cudaStream_t st);


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
// window fits smem_elems floats, |chip phase| < 2^22, n < 2^24 and a CTA segment ceil(n / splits) <= kStMaxSegment); otherwise the general kernel with smem_elems code elements.
// d_cp already holds theta / wb (no prep pass). d_shift_samples may be any valid device pointer when !high-dynamics.
cudaError_t launch_real_plain(bool fast, uint32_t n_ch, uint32_t splits, uint32_t smem_elems, cudaStream_t st, const ChanParams* d_cp,
    const float* d_shifts, const uint32_t* d_shift_samples, float2* d_partials, uint32_t* d_tickets, float2* d_out, uint32_t n_taps);
}  // namespace mc
}  // namespace b200
