// Host-side plan of the any-length mixed-radix FFT used by the acquisition pipeline: prime factors <= 13 have register
// codelets, primes 17 ... 127 run as a direct DFT of that length (fft_smem.cuh: smem_pass_any); beyond that: unsupported.
//
// The transform of length N is an in-place multi-dimensional decimation: N = R_0 * R_1 * ... * R_{m-1},
// index i = sum_j d_j * stride_j with stride_0 = 1. The forward transform is decimation-in-frequency
// (dims processed from the highest stride down, twiddle after each butterfly) and leaves the spectrum in
// digit-scrambled order; the inverse is the exact mirror, decimation-in-time from scrambled order back to
// natural order. Because the acquisition only ever multiplies two spectra point-wise between a forward and
// an inverse transform (pcps_acquisition.cc:662-668), no reordering pass is ever needed.
//
// The dims are grouped into one or two STAGES. A stage is executed by one CTA per tile, entirely in shared
// memory; with two stages the intermediate goes through a global scratch (L2-resident by chunking).
#pragma once
#include <cstdint>
#include <vector>

namespace b200
{
struct FftPass
{
    uint32_t radix;
    uint32_t lstride;   // stride of this dim in the stage's local index space
    uint32_t tw_scale;  // N / (radix * global_stride): twiddle exponent multiplier
    uint32_t has_tw;    // 0 when the global stride is 1 (all twiddles are 1)
};

constexpr int kMaxPasses = 6;
constexpr uint32_t kMaxAnyRadix = 127;  // largest radix of the direct-DFT pass (<= the smallest CTA, 128 threads)

struct FftStage
{
    uint32_t M;           // points of one sub-transform of this stage
    uint32_t n_pass;      // passes[0] is the FIRST pass of the forward transform (highest stride)
    FftPass pass[kMaxPasses];
    uint32_t pos_stride;  // global distance between consecutive stage indices (upper stage: M_lo; else 1)
    uint32_t tw_lo_mul;   // multiplier of the local low index in the twiddle exponent (upper: M_lo; else 1)
    uint32_t is_upper;
    // tiling chosen for the kernels
    uint32_t C;           // columns per tile (upper stage) — 1 for the lowest stage
    uint32_t blocks;      // contiguous sub-transforms per tile (lowest stage) — 1 for the upper stage
    uint32_t tiles;       // tiles per length-N transform
    uint32_t tile_step;   // global offset between tiles
    uint32_t tile_points; // points held in shared memory per tile = M*C*blocks
    uint32_t threads;
};

struct FftPlan
{
    // "fast" plans: two stages of exactly two passes each, radices from the instantiated sets of acq_fast.cuh.
    // Stage shapes: lower (contiguous) M_lo = ra_lo * rb_lo with ra_lo the stride-1 dim; upper M_hi = ra_hi * rb_hi.
    // The frequency-domain buffers (X_k, code spectra) of a fast plan use the coalescing-friendly layout
    //     F(pos) = u_a * (N / ra_lo) + blk * rb_lo + u_b,   pos = blk * M_lo + u_a + ra_lo * u_b
    // so that the inverse's first register pass reads them with unit stride across threads.
    uint32_t fast = 0;
    uint32_t ra_lo = 0, rb_lo = 0, ra_hi = 0, rb_hi = 0;
    uint32_t N = 0;
    uint32_t n_stages = 0;  // 1 or 2; forward order: stage[0] then stage[1]
    FftStage stage[2];
    std::vector<uint32_t> radices;  // all dims, lowest stride first (for documentation / tests)
};

// Returns false (and leaves `why`) when N has a prime factor the codelets do not cover.
bool make_fft_plan(uint32_t N, uint32_t max_single_stage_points, FftPlan* plan, const char** why);

// radix sets the fast kernels are instantiated for (acq_fast.cuh)
bool fast_ra_supported(uint32_t r);
bool fast_rb_supported(uint32_t r);

// position -> frequency-bin permutation of the forward transform's output (tests).
std::vector<uint32_t> fft_plan_permutation(const FftPlan& plan);

}  // namespace b200
