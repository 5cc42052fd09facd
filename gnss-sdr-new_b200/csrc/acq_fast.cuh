// Fast inverse path of the acquisition pipeline for "fast" FFT plans (fft_plan.h): two stages of two register
// passes each. Every point makes exactly ONE round trip through shared memory per stage (the transpose between
// the two register passes); the first pass reads global memory straight into registers and the last pass
// leaves registers straight for global memory / the |.|^2 reduction. All radices are template parameters, so
// the index arithmetic is constant-folded and there is no per-pass dispatch.
//
//   inv_lower_kernel<RA,RB>  : product X_k * C_prn (pcps_acquisition.cc:665) -> IDFT over the contiguous lower stage
//                              (M_lo = RA*RB) -> inter-stage twiddle -> scratch
//   inv_upper_kernel<RA,RB>  : scratch -> IDFT over the strided upper stage (M_hi = RA*RB) = the rest of the IFFT (:668)
//                              -> |.|^2 (+ non-coherent accumulation) (:672-680) -> per-tile {max, argmax, sum}
#pragma once
#include "acq_kernels.cuh"

namespace b200
{
struct FastArgs
{
    const float2* W;  // exp(-2 pi i k / N), k < N
    uint32_t N, M_lo, M_hi;
    uint32_t ra_lo, rb_lo, ra_hi, rb_hi;
    // row mapping (same convention as StageArgs)
    const uint32_t* slots;
    const uint32_t* slot_group;
    const uint32_t* k_override;
    uint32_t nD_local;
    uint32_t y0;
    const float2* X;      // [call group][nD_local][N], fast layout
    const float2* Cs;     // [slot][N], fast layout
    float2* scratch;      // [row - y0][N], natural order between the stages
    uint32_t nb;          // lower stage: contiguous sub-transforms per tile
    uint32_t C;           // upper stage: columns per tile
    // STORE_MAG part
    float* mag_base;
    uint32_t mag_is_rowbuf;
    uint32_t eff;
    uint32_t mag_offset;
    const uint32_t* dwell_after;
    RowPartial* partials;
};

__device__ __forceinline__ void fast_row_of(const FastArgs& a, uint32_t y, uint32_t& si, uint32_t& kk)
{
    if (a.k_override)
        {
            si = y;
            kk = a.k_override[y];
        }
    else
        {
            si = y / a.nD_local;
            kk = y - si * a.nD_local;
        }
}

constexpr int kFastThreads = 256;

// ------------------------------------------------------------------------------------------------
// lower stage: grid (rows, M_hi / nb). Shared memory: nb * PITCH * RB data + M twiddles.
// ------------------------------------------------------------------------------------------------
constexpr uint32_t lower_threads(uint32_t nb, uint32_t rb)  // >= 64: twiddle fill
{
    const uint32_t t = (nb * rb + 31) / 32 * 32;
    return t < 64 ? 64 : (t > (uint32_t)kFastThreads ? (uint32_t)kFastThreads : t);
}

// RAH, RBH, NB != 0: the upper-stage radices and the sub-transforms per tile are compile-time (launch with
// lower_threads(NB, RB) threads): loop trip counts, tile offsets and the k1 arithmetic fold into constants.
template <int RA, int RB, int RAH = 0, int RBH = 0, int NB = 0>
__global__ void __launch_bounds__(kFastThreads, 4) inv_lower_kernel(const FastArgs a)
{
    const uint32_t nb = NB ? NB : a.nb, ra_hi = RAH ? RAH : a.ra_hi, rb_hi = RBH ? RBH : a.rb_hi;
    const uint32_t bd = NB ? lower_threads(NB, RB) : blockDim.x;
    constexpr int M = RA * RB;
    constexpr int P = (RA % 2 == 0) ? RA + 1 : RA;  // odd pitch: conflict-free 64-bit stores with stride P
    constexpr int MP = P * RB;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float2* s = reinterpret_cast<float2*>(smem_raw);
    float2* s_tw = s + (size_t)nb * MP;  // conj W_M^(a*ub), [ub][a]

    const uint32_t y = a.y0 + blockIdx.x;
    uint32_t si, kk;
    fast_row_of(a, y, si, kk);
    const size_t N = RAH ? (size_t)RA * RB * RAH * RBH : a.N;
    const float2* X = a.X + ((size_t)a.slot_group[si] * a.nD_local + kk) * N;
    const float2* Cs = a.Cs + (size_t)a.slots[si] * N;
    const uint32_t blk0 = blockIdx.y * nb;
    const uint32_t tw_scale = (uint32_t)N / M;  // W_M = W_N^(N/M)

    // ---- pass A (stride-1 dim, no twiddle): one thread per (block, ub) pencil ----
    // Order of the first round: twiddle-table loads and the pencil's own global loads are issued together, the table
    // is stored while the pencil data is still in flight, and the barrier sits between the butterfly (which needs no
    // twiddles) and the twiddle multiply — so the table costs no extra L2 round trip.
    const uint32_t n_pen_a = nb * RB;
    const uint32_t seg = (uint32_t)N / RA;
    constexpr int kTwPerThread = (M + 63) / 64;  // enough for blockDim.x >= 64
    float2 wreg[kTwPerThread];
#pragma unroll
    for (int j = 0; j < kTwPerThread; j++)
        {
            const uint32_t i = threadIdx.x + j * bd;
            if (i < (uint32_t)M)
                {
                    const uint32_t ub = i / RA, aa = i - ub * RA;
                    wreg[j] = __ldg(a.W + (size_t)aa * ub * tw_scale);
                }
        }
    auto load_product = [&](float2(&x)[RA], uint32_t p) {
        const uint32_t bl = p / RB, ub = p - bl * RB;
        const uint32_t src = (blk0 + bl) * RB + ub;  // + ua * seg  (fast layout: unit stride across threads)
        float2 c[RA];
#pragma unroll
        for (int u = 0; u < RA; u++) x[u] = __ldg(X + src + (size_t)u * seg);
#pragma unroll
        for (int u = 0; u < RA; u++) c[u] = __ldg(Cs + src + (size_t)u * seg);
#pragma unroll
        for (int u = 0; u < RA; u++) x[u] = fft::cmul(x[u], c[u]);
    };
    auto twiddle_store = [&](float2(&x)[RA], uint32_t p) {
        const uint32_t bl = p / RB, ub = p - bl * RB;
        float2* dst = s + (size_t)bl * MP + (size_t)ub * P;
#pragma unroll
        for (int q = 0; q < RA; q++) dst[q] = fft::cmul(x[q], s_tw[ub * RA + q]);
    };
    {
        float2 x[RA];
        const bool have = threadIdx.x < n_pen_a;
        if (have) load_product(x, threadIdx.x);
#pragma unroll
        for (int j = 0; j < kTwPerThread; j++)
            {
                const uint32_t i = threadIdx.x + j * bd;
                if (i < (uint32_t)M) s_tw[i] = make_float2(wreg[j].x, -wreg[j].y);
            }
        if (have) fft::Dft<RA, true>::run(x);
        __syncthreads();
        if (have) twiddle_store(x, threadIdx.x);
    }
    for (uint32_t p = threadIdx.x + bd; p < n_pen_a; p += bd)
        {
            float2 x[RA];
            load_product(x, p);
            fft::Dft<RA, true>::run(x);
            twiddle_store(x, p);
        }
    __syncthreads();

    // ---- pass B + inter-stage twiddle: one thread per (block, a) pencil ----
    const uint32_t n_pen_b = nb * RA;
    float2* out = a.scratch + (size_t)(y - a.y0) * N;
    for (uint32_t p = threadIdx.x; p < n_pen_b; p += bd)
        {
            const uint32_t bl = p / RA, aa = p - bl * RA;
            const uint32_t gblk = blk0 + bl;
            // upper-stage frequency index of this block: block position = ua' + ra_hi * ub'  ->  k1 = ua' * rb_hi + ub'
            const uint32_t ubh = gblk / ra_hi, uah = gblk - ubh * ra_hi;
            const uint32_t k1 = uah * rb_hi + ubh;
            const float2 wb = __ldg(a.W + (size_t)k1 * aa);  // W_N^(k1 * a)
            const float2 ws = __ldg(a.W + (size_t)k1 * RA);  // W_N^(k1 * RA)
            float2 v[RB];
            const float2* src = s + (size_t)bl * MP + aa;
#pragma unroll
            for (int u = 0; u < RB; u++) v[u] = src[(size_t)u * P];
            fft::Dft<RB, true>::run(v);
            float2 tw = make_float2(wb.x, -wb.y);
            const float2 st = make_float2(ws.x, -ws.y);
            float2* dst = out + (size_t)gblk * M + aa;
#pragma unroll
            for (int b = 0; b < RB; b++)
                {
                    dst[(size_t)b * RA] = fft::cmul(v[b], tw);  // conj W_N^(k1 * (a + RA b))
                    tw = fft::cmul(tw, st);
                }
        }
}

// ------------------------------------------------------------------------------------------------
// upper stage: grid (rows, M_lo / C), block (C, threads / C). Shared memory: M * C data + M twiddles.
// ------------------------------------------------------------------------------------------------
// Shared-memory layout of the upper stage: element (q + RA*ub, c) at ub*G + q*C + c with the group stride G = RA*C padded so that
// G = C (mod 16): with blockDim.x = C not a multiple of 16 a warp spans two ub groups (pass A) or two rows (pass B), and
// the lanes of the second one then continue on the banks right after the first one's instead of landing on top of them.
__host__ __device__ __forceinline__ uint32_t upper_group_stride(uint32_t ra, uint32_t C)
{
    const uint32_t g = ra * C;
    return g + ((C + 16u - (g & 15u)) & 15u);
}

// MODE 0: what to do with the magnitudes is read from the arguments (row buffer, partial rows of the bit-transition search, ...).
// MODE 2: no magnitude output at all (CFAR search without a kept grid, full rows): only the per-tile records leave the kernel.
// MODE 1 (GRID): persistent magnitude grid (max_dwells > 1 / dump), full rows: the old grid values of a pencil are requested
// (streaming loads) before its butterfly so that their HBM latency hides behind it, and written back with streaming stores.
// MLO, CC != 0: the lower-stage length and the column tile are compile-time (launch with block (CC, kFastThreads / CC)): every
// shared-memory and global offset of a pencil folds into an immediate and the pencil loops unroll.
template <int RA, int RB, int MODE = 0, int MLO = 0, int CC = 0>
__global__ void __launch_bounds__(kFastThreads, (MODE == 2 && CC != 0 && RA <= 16) ? 6 : 0) inv_upper_kernel(const FastArgs a)
{
    constexpr int M = RA * RB;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float2* s = reinterpret_cast<float2*>(smem_raw);
    const uint32_t C = CC ? CC : a.C, M_lo = MLO ? MLO : a.M_lo;
    const uint32_t bdx = CC ? CC : blockDim.x, bdy = CC ? kFastThreads / CC : blockDim.y;
    const uint32_t G = upper_group_stride(RA, C);
    float2* s_tw = s + (size_t)RB * G;  // conj W_M^(a*ub), [ub][a]
    __shared__ float s_rmax[kFastThreads / 32];
    __shared__ uint32_t s_ridx[kFastThreads / 32];
    __shared__ float s_rsum[kFastThreads / 32];
    __shared__ uint32_t s_tile_idx;

    const uint32_t y = a.y0 + blockIdx.x;
    uint32_t si, kk;
    fast_row_of(a, y, si, kk);
    const size_t N = MLO ? (size_t)MLO * M : a.N;
    const uint32_t c = threadIdx.x, c0 = blockIdx.y * C;
    const uint32_t tid = threadIdx.y * bdx + threadIdx.x, nthreads = bdx * bdy;
    const uint32_t tw_scale = M_lo;  // N / M
    const float2* in = a.scratch + (size_t)(y - a.y0) * N + c0 + c;

    // ---- pass A: pencils (column c, ub); same ordering trick as in the lower kernel ----
    constexpr int kTwPerThread = (M + 63) / 64;
    float2 wreg[kTwPerThread];
#pragma unroll
    for (int j = 0; j < kTwPerThread; j++)
        {
            const uint32_t i = tid + j * nthreads;
            if (i < (uint32_t)M)
                {
                    const uint32_t ub = i / RA, aa = i - ub * RA;
                    wreg[j] = __ldg(a.W + (size_t)aa * ub * tw_scale);
                }
        }
    auto load_pencil = [&](float2(&v)[RA], uint32_t ub) {
#pragma unroll
        for (int u = 0; u < RA; u++) v[u] = __ldcg(in + (size_t)(u + RA * ub) * M_lo);
    };
    auto twiddle_store = [&](float2(&v)[RA], uint32_t ub) {
#pragma unroll
        for (int q = 0; q < RA; q++) s[(size_t)ub * G + q * C + c] = fft::cmul(v[q], s_tw[ub * RA + q]);
    };
    {
        float2 v[RA];
        const bool have = threadIdx.y < (uint32_t)RB;
        if (have) load_pencil(v, threadIdx.y);
#pragma unroll
        for (int j = 0; j < kTwPerThread; j++)
            {
                const uint32_t i = tid + j * nthreads;
                if (i < (uint32_t)M) s_tw[i] = make_float2(wreg[j].x, -wreg[j].y);
            }
        if (have) fft::Dft<RA, true>::run(v);
        __syncthreads();
        if (have) twiddle_store(v, threadIdx.y);
    }
    for (uint32_t ub = threadIdx.y + bdy; ub < (uint32_t)RB; ub += bdy)
        {
            float2 v[RA];
            load_pencil(v, ub);
            fft::Dft<RA, true>::run(v);
            twiddle_store(v, ub);
        }
    __syncthreads();

    // ---- pass B + |.|^2 + reductions: pencils (column c, a) ----
    constexpr bool GRID = (MODE == 1);
    float* mag = nullptr;
    if (MODE != 2 && a.mag_base)
        mag = a.mag_is_rowbuf ? a.mag_base + (size_t)si * a.eff : a.mag_base + ((size_t)a.slots[si] * a.nD_local + kk) * a.eff;
    const bool accumulate = MODE != 2 && mag && !a.mag_is_rowbuf && a.dwell_after[si] > 1;
    const bool full_rows = MODE == 2 || (a.mag_offset == 0 && a.eff == N);
    float best = -1.f, sum = 0.f;
    uint32_t best_i = 0xffffffffu;
    if (tid == 0) s_tile_idx = 0xffffffffu;
    if (full_rows)
        {
            if constexpr (MODE == 2)
                {
                    // Records only: the running maximum AND its position stay in registers (a compare and two selects per point),
                    // so the shared-memory pipe, which this kernel saturates first, carries nothing for the reduction. Within a
                    // pencil positions ascend with b, so "strictly greater" keeps the lowest index; across the pencils of a thread
                    // and across threads the tie goes to the lower position explicitly (:523).
                    uint32_t best_pos = 0xffffffffu;
                    for (uint32_t aa = threadIdx.y; aa < (uint32_t)RA; aa += bdy)
                        {
                            float2 v[RB];
#pragma unroll
                            for (int u = 0; u < RB; u++) v[u] = s[(size_t)u * G + aa * C + c];
                            fft::Dft<RB, true>::run(v);
                            float pbest = -1.f;
                            uint32_t pb = 0;
#pragma unroll
                            for (int b = 0; b < RB; b++)
                                {
                                    const float g = v[b].x * v[b].x + v[b].y * v[b].y;  // :674
                                    sum += g;
                                    const bool gt = g > pbest;
                                    pbest = gt ? g : pbest;
                                    pb = gt ? (uint32_t)b : pb;
                                }
                            const uint32_t ppos = (aa + RA * pb) * M_lo + c0 + c;
                            if (pbest > best || (pbest == best && ppos < best_pos))
                                {
                                    best = pbest;
                                    best_pos = ppos;
                                }
                        }
                    const float mine_best = best;
#pragma unroll
                    for (int o = 16; o > 0; o >>= 1)
                        {
                            best = fmaxf(best, __shfl_xor_sync(0xffffffffu, best, o));
                            sum += __shfl_xor_sync(0xffffffffu, sum, o);
                        }
                    const int lane = tid & 31, warp = tid >> 5;
                    if (lane == 0)
                        {
                            s_rmax[warp] = best;
                            s_rsum[warp] = sum;
                        }
                    __syncthreads();
                    const int nw = (nthreads + 31) / 32;
                    float tile_max = s_rmax[0];
                    for (int w = 1; w < nw; w++) tile_max = fmaxf(tile_max, s_rmax[w]);
                    if (mine_best == tile_max && best_pos != 0xffffffffu) atomicMin(&s_tile_idx, best_pos);
                    __syncthreads();
                    if (tid == 0)
                        {
                            for (int w = 1; w < nw; w++) sum += s_rsum[w];
                            RowPartial rp;
                            rp.max_val = tile_max;
                            rp.max_idx = s_tile_idx;
                            rp.sum = sum;
                            rp.pad = 0;
                            a.partials[(size_t)y * gridDim.y + blockIdx.y] = rp;
                        }
                    return;
                }
            // Common case (no bit transition): every output belongs to the row. The running maximum is a plain FMNMX;
            // each |.|^2 is parked in the shared-memory slot its pencil just consumed (no other thread reads it), and
            // once the tile maximum is known only the thread(s) holding it look their slots up for the lowest index.
            for (uint32_t aa = threadIdx.y; aa < (uint32_t)RA; aa += bdy)
                {
                    float2 v[RB];
                    const uint32_t pos0 = aa * M_lo + c0 + c;
                    const uint32_t pstep = RA * M_lo;
                    float old[GRID ? RB : 1];
                    if (GRID)
                        {
#pragma unroll
                            for (int b = 0; b < RB; b++) old[b] = accumulate ? __ldcs(mag + pos0 + b * pstep) : 0.f;
                        }
#pragma unroll
                    for (int u = 0; u < RB; u++) v[u] = s[(size_t)u * G + aa * C + c];
                    fft::Dft<RB, true>::run(v);
#pragma unroll
                    for (int b = 0; b < RB; b++)
                        {
                            float g = v[b].x * v[b].x + v[b].y * v[b].y;       // :674
                            if (GRID)
                                {
                                    g = old[b] + g;                                 // :679 (0 + g on the first dwell)
                                    __stcs(mag + pos0 + b * pstep, g);
                                }
                            else
                                {
                                    if (accumulate) g = mag[pos0 + b * pstep] + g;  // :679
                                    if (mag) mag[pos0 + b * pstep] = g;
                                }
                            sum += g;
                            best = fmaxf(best, g);
                            s[(size_t)b * G + aa * C + c].x = g;
                        }
                }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1)
                {
                    best = fmaxf(best, __shfl_xor_sync(0xffffffffu, best, o));
                    sum += __shfl_xor_sync(0xffffffffu, sum, o);
                }
            const int lane = tid & 31, warp = tid >> 5;
            if (lane == 0)
                {
                    s_rmax[warp] = best;
                    s_rsum[warp] = sum;
                }
            __syncthreads();
            const int nw = (nthreads + 31) / 32;
            float tile_max = s_rmax[0];
            for (int w = 1; w < nw; w++) tile_max = fmaxf(tile_max, s_rmax[w]);
            if (best == tile_max)  // warp-uniform after the butterfly reduction: at most a few warps per tile get here
                {
                    uint32_t mine = 0xffffffffu;
                    for (uint32_t aa = threadIdx.y; aa < (uint32_t)RA; aa += bdy)
                        {
#pragma unroll
                            for (int b = 0; b < RB; b++)
                                if (s[(size_t)b * G + aa * C + c].x == tile_max) mine = min(mine, (aa + RA * b) * M_lo + c0 + c);
                        }
                    if (mine != 0xffffffffu) atomicMin(&s_tile_idx, mine);  // lowest index wins ties (:523)
                }
            __syncthreads();
            if (tid == 0)
                {
                    for (int w = 1; w < nw; w++) sum += s_rsum[w];
                    RowPartial rp;
                    rp.max_val = tile_max;
                    rp.max_idx = s_tile_idx;
                    rp.sum = sum;
                    rp.pad = 0;
                    a.partials[(size_t)y * gridDim.y + blockIdx.y] = rp;
                }
            return;
        }
    for (uint32_t aa = threadIdx.y; aa < (uint32_t)RA; aa += bdy)
        {
            float2 v[RB];
#pragma unroll
            for (int u = 0; u < RB; u++) v[u] = s[(size_t)u * G + aa * C + c];
            fft::Dft<RB, true>::run(v);
#pragma unroll
            for (int b = 0; b < RB; b++)
                {
                    const uint32_t pos = (aa + RA * b) * M_lo + c0 + c;  // natural time index
                    const uint32_t m = pos - a.mag_offset;
                    if (m < a.eff)
                        {
                            float g = v[b].x * v[b].x + v[b].y * v[b].y;  // :674
                            if (accumulate) g = mag[m] + g;               // :679
                            if (mag) mag[m] = g;
                            sum += g;
                            if (g > best || (g == best && m < best_i))
                                {
                                    best = g;
                                    best_i = m;
                                }
                        }
                }
        }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
        {
            const float ob = __shfl_xor_sync(0xffffffffu, best, o);
            const uint32_t oi = __shfl_xor_sync(0xffffffffu, best_i, o);
            sum += __shfl_xor_sync(0xffffffffu, sum, o);
            if (ob > best || (ob == best && oi < best_i))
                {
                    best = ob;
                    best_i = oi;
                }
        }
    const int lane = tid & 31, warp = tid >> 5;
    if (lane == 0)
        {
            s_rmax[warp] = best;
            s_ridx[warp] = best_i;
            s_rsum[warp] = sum;
        }
    __syncthreads();
    if (tid == 0)
        {
            const int nw = (nthreads + 31) / 32;
            for (int w = 1; w < nw; w++)
                {
                    sum += s_rsum[w];
                    if (s_rmax[w] > best || (s_rmax[w] == best && s_ridx[w] < best_i))
                        {
                            best = s_rmax[w];
                            best_i = s_ridx[w];
                        }
                }
            RowPartial rp;
            rp.max_val = best;
            rp.max_idx = best_i;
            rp.sum = sum;
            rp.pad = 0;
            a.partials[(size_t)y * gridDim.y + blockIdx.y] = rp;
        }
}

// host-side launchers (instantiations live in acq_fast_lower.cu / acq_fast_upper.cu)
cudaError_t launch_inv_lower(uint32_t ra, uint32_t rb, dim3 grid, uint32_t threads, size_t smem, cudaStream_t st, const FastArgs& a);
cudaError_t launch_inv_upper(uint32_t ra, uint32_t rb, int mode, dim3 grid, dim3 block, size_t smem, cudaStream_t st, const FastArgs& a);

}  // namespace b200
