// Device side of the PCPS acquisition pipeline (sm_100a).
//
// One templated "stage kernel" runs one shared-memory FFT stage of the plan with a fused prologue and
// epilogue, so that the Doppler-grid loop of pcps_acquisition::acquisition_core
// (reference: src/algorithms/acquisition/gnuradio_blocks/pcps_acquisition.cc:655-686) becomes
//
//   forward  (once per Doppler bin, shared by every PRN of the batch):
//       load samples -> x[n] * wipeoff_k[n] (:658) -> zero pad (:628-634) -> FFT (:662) -> X_k (scrambled order)
//   inverse  (once per (PRN, Doppler bin) row):
//       X_k * conj-code-spectrum (:665) -> IFFT (:668) -> |.|^2 (+ non-coherent accumulation) (:672-680)
//       -> per-row {max, argmax, sum} (what :520-529 and :534 consume)
//
// The code spectra themselves (set_local_code, :176-209) go through the same forward stages with the
// zero-pad layout as prologue and the conjugation (:208) as epilogue.
#pragma once
#include "common.cuh"
#include "fft_smem.cuh"

namespace b200
{

// (cos phi, sin phi) of a wipe-off phase (volk_gnsssdr_s32f_sincos_32fc): two-constant Cody-Waite reduction to [-pi, pi]
// (exact for the |phi| < 2^15 rad a dwell accumulates) and the SFU's MUFU.SIN / MUFU.COS, absolute error 2^-21.4 there — below
// the spread between the reference's own protokernels (libm in the generic one, polynomials in the SIMD ones).
__device__ __forceinline__ void wipeoff_sincos(float phi, float& sn, float& cs)
{
    const float k = rintf(phi * 0.15915494309189535f);
    float r = fmaf(-k, 6.2831854820251465f, phi);   // float(2 pi)
    r = fmaf(-k, -1.7484555314695172e-07f, r);      // 2 pi - float(2 pi)
    sn = __sinf(r);
    cs = __cosf(r);
}

// Piecewise-linear description of the reference's binary32 running-sum phase (see wipeoff_pieces.cc):
// for start <= n < next.start the reference's _phase at sample n is exactly phi0 + (n - start) * step.
struct WipePiece
{
    uint32_t start;
    float phi0;
    float step;
};
constexpr int kMaxPieces = 64;

struct RowPartial
{
    float max_val;
    uint32_t max_idx;
    float sum;
    uint32_t pad;
};

enum
{
    LOAD_SAMPLES = 0,  // samples * carrier wipe-off, zero padded
    LOAD_CODE = 1,     // time-domain local code in its zero-pad layout
    LOAD_BUF = 2,      // plain buffer [transform][N]
    LOAD_PRODUCT = 3   // X_k[pos] * C_prn[pos]
};
enum
{
    STORE_BUF = 0,
    STORE_CONJ = 1,
    STORE_MAG = 2
};

struct StageArgs
{
    FftStage st;
    const float2* W;
    uint32_t N;
    // ---- LOAD_SAMPLES ----
    const void* samples;
    uint32_t is_cshort;  // sample format: 0 gr_complex, 1 lv_16sc_t, 2 lv_8sc_t (GNSS_B200_SAMPLES_*)
    uint32_t consumed;
    const WipePiece* pieces;   // [group][nD_local][kMaxPieces]
    const uint32_t* n_pieces;  // [group][nD_local]
    const uint32_t* group_piece_base;  // [call group] -> index of its first bin in pieces / n_pieces
    uint32_t nD_local;
    // ---- LOAD_CODE ----
    const float2* codes_td;  // [transform][code_len]
    uint32_t code_len;
    uint32_t code_off;       // first position of the code inside the padded buffer
    uint32_t code_wrap;      // != 0: the code's period (fft_size) inside a longer padded transform; its tail copy ends the buffer (acq.cu create_impl)
    float code_scale;        // factor on the code samples (1 unless code_wrap)
    // ---- LOAD_BUF / LOAD_PRODUCT ----
    const float2* in_buf;    // LOAD_BUF: [transform - in_y0][N]; LOAD_PRODUCT: X [call group][nD_local][N]
    const float2* cbuf;      // code spectra [slot][N]
    uint32_t in_y0;
    // ---- row mapping of the inverse kernels: y -> (slot, Doppler bin) ----
    const uint32_t* slots;       // [n_slots in call]
    const uint32_t* slot_group;  // [n_slots in call] -> call group
    const uint32_t* k_override;  // null, or [n_slots]: the single LOCAL bin to (re)compute per slot
    uint32_t y0;                 // first row of this launch
    // ---- STORE_BUF / STORE_CONJ ----
    float2* out_buf;             // [transform - out_y0][N] (STORE_CONJ: [slots[transform]][N])
    uint32_t out_y0;
    uint32_t out_by_slot;
    // ---- STORE_MAG ----
    float* mag_base;             // grid [slot][nD_local][eff], or row buffer [si][eff], or null
    uint32_t mag_is_rowbuf;
    uint32_t eff;
    uint32_t mag_offset;         // first IFFT output that belongs to the row (N/2 with bit_transition, :671)
    const uint32_t* dwell_after; // [n_slots in call] dwell counter including this dwell
    RowPartial* partials;        // [y][tiles], y = global row of the call (not chunk-relative)
    // ---- frequency-domain layout of fast plans (fft_plan.h): 0 = natural scrambled order ----
    uint32_t fmap_ra, fmap_rb, fmap_mlo;
    uint32_t fmap_store;         // apply the map when storing (last forward stage only)
};

// position in the plan's scrambled order -> position in the stored frequency-domain buffers
__device__ __forceinline__ uint32_t fmap(const StageArgs& a, uint32_t pos)
{
    if (a.fmap_ra == 0) return pos;
    const uint32_t blk = pos / a.fmap_mlo, l = pos - blk * a.fmap_mlo;
    const uint32_t ub = l / a.fmap_ra, ua = l - ub * a.fmap_ra;
    return ua * (a.N / a.fmap_ra) + blk * a.fmap_rb + ub;
}

__device__ __forceinline__ void row_of(const StageArgs& a, uint32_t y, uint32_t& si, uint32_t& kk)
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

template <int LOAD, int STORE, bool INV>
__global__ void __launch_bounds__(512) stage_kernel(const StageArgs a)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    float2* s = reinterpret_cast<float2*>(smem_raw);
    __shared__ WipePiece s_pieces[LOAD == LOAD_SAMPLES ? kMaxPieces : 1];
    __shared__ float s_rmax[16];
    __shared__ uint32_t s_ridx[16];
    __shared__ float s_rsum[16];

    const FftStage& st = a.st;
    const uint32_t y = a.y0 + blockIdx.x;  // transform / row index
    const uint32_t tile = blockIdx.y;
    const uint32_t tile_off = tile * st.tile_step;
    const uint32_t points = st.tile_points, C = st.C;
    const size_t N = a.N;

    // ------------------------------ prologue: fill the tile ------------------------------
    if constexpr (LOAD == LOAD_SAMPLES)
        {
            // y = call_group * nD_local + local bin
            const uint32_t g = y / a.nD_local, kk = y - g * a.nD_local;
            const uint32_t pb = a.group_piece_base[g] + kk;
            const uint32_t np = a.n_pieces[pb];
            for (uint32_t i = threadIdx.x; i < np; i += blockDim.x) s_pieces[i] = a.pieces[(size_t)pb * kMaxPieces + i];
            __syncthreads();
            uint32_t cur = 0;
            for (uint32_t i = threadIdx.x; i < points; i += blockDim.x)
                {
                    const uint32_t idx = i / C, c = i - idx * C;
                    const uint32_t pos = idx * st.pos_stride + tile_off + c;
                    float2 v = make_float2(0.f, 0.f);
                    if (pos < a.consumed)
                        {
                            float2 x;
                            if (a.is_cshort == 1)
                                {
                                    const short2 q = __ldg(static_cast<const short2*>(a.samples) + pos);  // :623-626
                                    x = make_float2((float)q.x, (float)q.y);
                                }
                            else if (a.is_cshort == 2)
                                {
                                    // cbyte items: the adapter's cbyte_to_float_x2 + float_to_complex (gps_l1_ca_pcps_acquisition.cc:76-80,191-198)
                                    const char2 q = __ldg(static_cast<const char2*>(a.samples) + pos);
                                    x = make_float2((float)q.x, (float)q.y);
                                }
                            else
                                {
                                    x = __ldg(static_cast<const float2*>(a.samples) + pos);
                                }
                            while (cur + 1 < np && pos >= s_pieces[cur + 1].start) cur++;
                            const WipePiece pc = s_pieces[cur];
                            // exact: the value is representable, the fma rounds once
                            const float phi = fmaf((float)(pos - pc.start), pc.step, pc.phi0);
                            float sn, cs;
                            wipeoff_sincos(phi, sn, cs);  // wipe-off sample (cos phi, sin phi), volk_gnsssdr_s32f_sincos_32fc
                            v = make_float2(x.x * cs - x.y * sn, x.x * sn + x.y * cs);  // :658
                        }
                    s[i] = v;
                }
        }
    else if constexpr (LOAD == LOAD_CODE)
        {
            const float2* code = a.codes_td + (size_t)y * a.code_len;
            for (uint32_t i = threadIdx.x; i < points; i += blockDim.x)
                {
                    const uint32_t idx = i / C, c = i - idx * C;
                    const uint32_t pos = idx * st.pos_stride + tile_off + c;
                    float2 v = make_float2(0.f, 0.f);
                    // index into the code's own period: the position itself, or in the padded form (period code_wrap < N) the
                    // first `consumed` positions directly and the last code_wrap - 1 positions wrapped onto 1 .. code_wrap - 1
                    uint32_t j = pos;
                    if (a.code_wrap)
                        {
                            j = 0xffffffffu;
                            if (pos < a.consumed)
                                j = pos;
                            else if (pos + a.code_wrap > (uint32_t)N)
                                j = pos + a.code_wrap - (uint32_t)N;
                        }
                    if (j != 0xffffffffu && j >= a.code_off && j - a.code_off < a.code_len) v = __ldg(code + (j - a.code_off));
                    if (a.code_wrap) v = make_float2(v.x * a.code_scale, v.y * a.code_scale);
                    s[i] = v;
                }
        }
    else if constexpr (LOAD == LOAD_BUF)
        {
            const float2* in = a.in_buf + (size_t)(y - a.in_y0) * N;
            for (uint32_t i = threadIdx.x; i < points; i += blockDim.x)
                {
                    const uint32_t idx = i / C, c = i - idx * C;
                    const uint32_t pos = idx * st.pos_stride + tile_off + c;
                    s[i] = __ldcg(in + pos);
                }
        }
    else
        {
            uint32_t si, kk;
            row_of(a, y, si, kk);
            const float2* X = a.in_buf + ((size_t)a.slot_group[si] * a.nD_local + kk) * N;
            const float2* Cs = a.cbuf + (size_t)a.slots[si] * N;
            for (uint32_t i = threadIdx.x; i < points; i += blockDim.x)
                {
                    const uint32_t idx = i / C, c = i - idx * C;
                    const uint32_t pos = idx * st.pos_stride + tile_off + c;
                    const uint32_t fp = fmap(a, pos);
                    s[i] = fft::cmul(__ldg(X + fp), __ldg(Cs + fp));  // :665
                }
        }
    __syncthreads();

    // ------------------------------ the FFT stage ------------------------------
    fft::run_stage<INV>(s, st, tile, a.W);

    // ------------------------------ epilogue ------------------------------
    if constexpr (STORE == STORE_BUF || STORE == STORE_CONJ)
        {
            float2* out = a.out_buf + (a.out_by_slot ? (size_t)a.slots[y] : (size_t)(y - a.out_y0)) * N;
            for (uint32_t i = threadIdx.x; i < points; i += blockDim.x)
                {
                    const uint32_t idx = i / C, c = i - idx * C;
                    const uint32_t pos = idx * st.pos_stride + tile_off + c;
                    float2 v = s[i];
                    if (STORE == STORE_CONJ) v.y = -v.y;  // :208
                    out[a.fmap_store ? fmap(a, pos) : pos] = v;
                }
        }
    else
        {
            uint32_t si, kk;
            row_of(a, y, si, kk);
            float* mag = nullptr;
            if (a.mag_base)
                mag = a.mag_is_rowbuf ? a.mag_base + (size_t)si * a.eff : a.mag_base + ((size_t)a.slots[si] * a.nD_local + kk) * a.eff;
            const bool accumulate = mag && !a.mag_is_rowbuf && a.dwell_after[si] > 1;
            float best = -1.f, sum = 0.f;
            uint32_t best_i = 0xffffffffu;
            for (uint32_t i = threadIdx.x; i < points; i += blockDim.x)
                {
                    const uint32_t idx = i / C, c = i - idx * C;
                    const uint32_t pos = idx * st.pos_stride + tile_off + c;
                    const uint32_t m = pos - a.mag_offset;  // wraps for pos < offset -> fails the range test
                    if (m < a.eff)
                        {
                            const float2 v = s[i];
                            float g = v.x * v.x + v.y * v.y;      // :674
                            if (accumulate) g = mag[m] + g;       // :679
                            if (mag) mag[m] = g;
                            sum += g;
                            if (g > best || (g == best && m < best_i))
                                {
                                    best = g;
                                    best_i = m;
                                }
                        }
                }
            // CTA reduction: lowest index among equal maxima (tie rule of volk_gnsssdr_32f_index_max_32u)
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
            const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
            if (lane == 0)
                {
                    s_rmax[warp] = best;
                    s_ridx[warp] = best_i;
                    s_rsum[warp] = sum;
                }
            __syncthreads();
            if (threadIdx.x == 0)
                {
                    const int nw = (blockDim.x + 31) / 32;
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
                    a.partials[(size_t)y * gridDim.y + tile] = rp;
                }
        }
}

// ---- cross-tile and cross-row decision: max_to_input_power_statistic (:511-543) ----
struct FinalizeArgs
{
    const RowPartial* partials;  // [n_slots * nD_local (or n_slots when recomputing)][tiles]
    uint32_t tiles;
    uint32_t nD, nD_local, row_begin;
    uint32_t eff;
    int32_t doppler_max;
    uint32_t doppler_step;
    uint32_t use_cfar;
    const uint32_t* slots;
    const int32_t* center;       // [n_slots in call]; second-step handles: bits of the float centre
    uint32_t step2;              // make_two_steps narrow grid (:308-315,:531-540)
    float step2_step;
    const float* input_power1;   // [n_slots in call] d_input_power left by the first step
    const uint32_t* dwell_after; // [n_slots in call]
    gnss_b200_acq_row* rows;     // [slot][nD]
    gnss_b200_acq_result* results;  // [n_slots in call]
    uint32_t* best_k_local;      // [n_slots in call]
};

// Many records per row (the row-resident inverse writes one per warp): one warp per row merges them into a single record, so that
// finalize_kernel's per-slot warp does not walk them one by one. Lowest index wins among equal maxima (:523); the sum is a fixed
// shuffle tree, i.e. deterministic.
static __global__ void merge_partials_kernel(const RowPartial* in, uint32_t tiles, uint32_t n_rows, RowPartial* out)
{
    const uint32_t y = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5), lane = threadIdx.x & 31u;
    if (y >= n_rows) return;
    float m = -1.f, s = 0.f;
    uint32_t mi = 0xffffffffu;
    for (uint32_t t = lane; t < tiles; t += 32)
        {
            const RowPartial p = in[(size_t)y * tiles + t];
            s += p.sum;
            if (p.max_val > m || (p.max_val == m && p.max_idx < mi))
                {
                    m = p.max_val;
                    mi = p.max_idx;
                }
        }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
        {
            const float om = __shfl_xor_sync(0xffffffffu, m, o);
            const uint32_t oi = __shfl_xor_sync(0xffffffffu, mi, o);
            s += __shfl_xor_sync(0xffffffffu, s, o);
            if (om > m || (om == m && oi < mi))
                {
                    m = om;
                    mi = oi;
                }
        }
    if (lane == 0)
        {
            RowPartial r;
            r.max_val = m;
            r.max_idx = mi;
            r.sum = s;
            r.pad = 0;
            out[y] = r;
        }
}

// one warp per slot
static __global__ void finalize_kernel(const FinalizeArgs a)
{
    const uint32_t si = blockIdx.x;
    const uint32_t slot = a.slots[si];
    const int lane = threadIdx.x;
    gnss_b200_acq_row* rows = a.rows + (size_t)slot * a.nD;
    float best = 0.f;  // grid_maximum starts at 0 and is replaced on strict '>' (:513,:523)
    uint32_t best_k = 0xffffffffu, best_i = 0;
    for (uint32_t k = lane; k < a.nD; k += 32)
        {
            gnss_b200_acq_row r;
            r.max_val = 0.f;
            r.max_idx = 0;
            r.row_sum = 0.f;
            r.valid = 0;
            if (k >= a.row_begin && k - a.row_begin < a.nD_local)
                {
                    const RowPartial* p = a.partials + ((size_t)si * a.nD_local + (k - a.row_begin)) * a.tiles;
                    float m = p[0].max_val, s = p[0].sum;
                    uint32_t mi = p[0].max_idx;
                    for (uint32_t t = 1; t < a.tiles; t++)
                        {
                            s += p[t].sum;
                            if (p[t].max_val > m || (p[t].max_val == m && p[t].max_idx < mi))
                                {
                                    m = p[t].max_val;
                                    mi = p[t].max_idx;
                                }
                        }
                    r.max_val = m;
                    r.max_idx = mi;
                    r.row_sum = s;
                    r.valid = 1;
                    if (m > best || (m == best && best_k != 0xffffffffu && k < best_k))
                        {
                            best = m;
                            best_k = k;
                            best_i = mi;
                        }
                }
            rows[k] = r;
        }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
        {
            const float ob = __shfl_xor_sync(0xffffffffu, best, o);
            const uint32_t ok = __shfl_xor_sync(0xffffffffu, best_k, o);
            const uint32_t oi = __shfl_xor_sync(0xffffffffu, best_i, o);
            if (ob > best || (ob == best && ok < best_k))
                {
                    best = ob;
                    best_k = ok;
                    best_i = oi;
                }
        }
    __syncwarp();
    if (lane == 0)
        {
            if (best_k == 0xffffffffu)  // nothing exceeded 0: the reference keeps index 0 / 0
                {
                    best_k = 0;
                    best_i = 0;
                }
            gnss_b200_acq_result res;
            res.peak = best;
            res.code_phase = best_i;
            res.doppler_index = best_k;
            res.doppler_hz = -a.doppler_max + a.center[si] + (int32_t)a.doppler_step * (int32_t)best_k;  // :535
            res.dwell_count = a.dwell_after[si];
            res.second_peak = 0.f;
            res.input_power = 0.f;
            res.test_statistic = 0.f;
            if (a.step2)
                {
                    // :539: static_cast<int32_t>(centre + (float(k) - float(floor(n2 / 2.0))) * doppler_step2); d_input_power is not refreshed
                    const float centre = __int_as_float(a.center[si]);
                    const float off = __fmul_rn(__fsub_rn((float)best_k, (float)(a.nD / 2)), a.step2_step);
                    res.doppler_hz = (int32_t)__fadd_rn(centre, off);
                    res.input_power = a.input_power1[si];
                    if (a.use_cfar) res.test_statistic = best / res.input_power;  // :542
                }
            else if (a.use_cfar)
                {
                    const uint32_t opp = (best_k + a.nD / 2) % a.nD;  // :533
                    const gnss_b200_acq_row ro = rows[opp];
                    if (ro.valid)
                        {
                            // float / int32 -> float, then / 2.0 / counter in double, narrowed (:534)
                            const float per = ro.row_sum / (float)(int32_t)a.eff;
                            res.input_power = (float)(((double)per / 2.0) / (double)a.dwell_after[si]);
                            res.test_statistic = best / res.input_power;  // :542
                        }
                    else
                        {
                            res.input_power = nanf("");
                            res.test_statistic = nanf("");  // row lives on another shard: merge with finalize_rows
                        }
                }
            a.results[si] = res;
            a.best_k_local[si] = (best_k >= a.row_begin && best_k - a.row_begin < a.nD_local) ? best_k - a.row_begin : 0;
        }
}

// ---- sharded search: the exchange unit and the decision from gathered rows ----
// pack: the rows this handle computed, as [n_slots of the call][chunk] records, contiguous (what one rank sends)
static __global__ void pack_rows_kernel(const gnss_b200_acq_row* rows, const uint32_t* slots, uint32_t nD, uint32_t row_begin, uint32_t nD_local,
    uint32_t chunk, gnss_b200_acq_row* out)
{
    const uint32_t si = blockIdx.x;
    const gnss_b200_acq_row* src = rows + (size_t)slots[si] * nD + row_begin;
    for (uint32_t j = threadIdx.x; j < chunk; j += blockDim.x)
        {
            gnss_b200_acq_row r;
            r.max_val = 0.f;
            r.max_idx = 0;
            r.row_sum = 0.f;
            r.valid = 0;
            if (j < nD_local) r = src[j];
            out[(size_t)si * chunk + j] = r;
        }
}

// decision (:511-543, CFAR statistic) from the records of ALL shards: gathered = [world][n_slots][chunk], shard r holds rows
// [r * chunk, (r + 1) * chunk). One warp per slot; also refreshes this handle's full row table.
static __global__ void finalize_gathered_kernel(const FinalizeArgs a, const gnss_b200_acq_row* gathered, uint32_t world, uint32_t chunk, uint32_t n_slots)
{
    const uint32_t si = blockIdx.x;
    const uint32_t slot = a.slots[si];
    const int lane = threadIdx.x;
    gnss_b200_acq_row* rows = a.rows + (size_t)slot * a.nD;
    float best = 0.f;
    uint32_t best_k = 0xffffffffu, best_i = 0;
    for (uint32_t k = lane; k < a.nD; k += 32)
        {
            const uint32_t r = k / chunk, j = k - r * chunk;
            const gnss_b200_acq_row row = gathered[((size_t)r * n_slots + si) * chunk + j];
            rows[k] = row;
            if (row.valid && (row.max_val > best || (row.max_val == best && best_k != 0xffffffffu && k < best_k)))
                {
                    best = row.max_val;
                    best_k = k;
                    best_i = row.max_idx;
                }
        }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
        {
            const float ob = __shfl_xor_sync(0xffffffffu, best, o);
            const uint32_t ok = __shfl_xor_sync(0xffffffffu, best_k, o);
            const uint32_t oi = __shfl_xor_sync(0xffffffffu, best_i, o);
            if (ob > best || (ob == best && ok < best_k))
                {
                    best = ob;
                    best_k = ok;
                    best_i = oi;
                }
        }
    __syncwarp();
    if (lane == 0)
        {
            if (best_k == 0xffffffffu)
                {
                    best_k = 0;
                    best_i = 0;
                }
            gnss_b200_acq_result res;
            res.peak = best;
            res.code_phase = best_i;
            res.doppler_index = best_k;
            res.doppler_hz = -a.doppler_max + a.center[si] + (int32_t)a.doppler_step * (int32_t)best_k;  // :535
            res.dwell_count = a.dwell_after[si];
            res.second_peak = 0.f;
            const uint32_t opp = (best_k + a.nD / 2) % a.nD;  // :533
            const gnss_b200_acq_row ro = rows[opp];
            const float per = ro.row_sum / (float)(int32_t)a.eff;
            res.input_power = ro.valid ? (float)(((double)per / 2.0) / (double)a.dwell_after[si]) : nanf("");
            res.test_statistic = best / res.input_power;  // :542
            a.results[si] = res;
        }
}

// ---- record exchange over peer memory (gnss_b200_acq_peer_*): no collective library in the data path --------------------
// Every shard owns one region [epoch | flags[2][world] (128 B apart) | recv[2][world][per_shard_rows]] that its peers can store
// into (CUDA IPC or peer access). After its search a shard stores its packed records into EVERY shard's recv[parity][rank] and
// then raises flag[parity][rank] there (release, system scope); the decision kernel of a shard waits until all `world` flags of
// the parity carry the current epoch (acquire, system scope) and decides from recv[parity]. Two parities: a shard may be one
// exchange ahead of a slow peer, never two (its own decision waits for that peer's records of the same epoch).
struct PeerLayout
{
    uint32_t world, rank, flags_off, recv_off, per_shard_rows;
    uint32_t bar_off;  // [world] rendezvous counters (128 B apart), written by the peers (peer_barrier_kernel)
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) { asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p)
{
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// grid = world CTAs: CTA d delivers this shard's [n_slots][chunk] records to shard d
static __global__ void peer_scatter_kernel(const gnss_b200_acq_row* send, uint32_t n_rows, char* const* peers, const PeerLayout L, const uint32_t* epoch_ctr)
{
    static_assert(sizeof(gnss_b200_acq_row) == 16, "one 16-byte store per record");
    const uint32_t epoch = *epoch_ctr + 1u, par = epoch & 1u;
    char* base = peers[blockIdx.x];
    uint4* dst = reinterpret_cast<uint4*>(base + L.recv_off) + (size_t)(par * L.world + L.rank) * L.per_shard_rows;
    const uint4* src = reinterpret_cast<const uint4*>(send);
    for (uint32_t i = threadIdx.x; i < n_rows; i += blockDim.x) dst[i] = src[i];
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) st_release_sys(reinterpret_cast<uint32_t*>(base + L.flags_off + (size_t)(par * L.world + L.rank) * 128u), epoch);
}

// finalize_gathered_kernel behind the wait for everybody's records; a peer that never delivers (a failed rank) ends the wait after
// ~2 s with NaN statistics instead of hanging the device
// bump_ticket != nullptr: the last CTA to finish advances the epoch counter itself (no peer_epoch_bump_kernel behind this one)
static __global__ void finalize_peer_kernel(const FinalizeArgs a, const char* own, const PeerLayout L, uint32_t* epoch_ctr, uint32_t chunk, uint32_t n_slots, uint32_t* bump_ticket)
{
    const uint32_t epoch = *epoch_ctr + 1u, par = epoch & 1u;
    const int lane = threadIdx.x;
    bool ok = true;
    for (uint32_t r = lane; r < L.world; r += 32)
        {
            const uint32_t* flag = reinterpret_cast<const uint32_t*>(own + L.flags_off + (size_t)(par * L.world + r) * 128u);
            const long long t0 = clock64();
            while (ld_acquire_sys(flag) != epoch)
                if (clock64() - t0 > 4000000000ll)
                    {
                        ok = false;
                        break;
                    }
        }
    ok = __all_sync(0xffffffffu, ok);
    const gnss_b200_acq_row* gathered = reinterpret_cast<const gnss_b200_acq_row*>(own + L.recv_off) + (size_t)par * L.world * L.per_shard_rows;
    const uint32_t si = blockIdx.x;
    const uint32_t slot = a.slots[si];
    gnss_b200_acq_row* rows = a.rows + (size_t)slot * a.nD;
    float best = 0.f;
    uint32_t best_k = 0xffffffffu, best_i = 0;
    for (uint32_t k = lane; k < a.nD; k += 32)
        {
            const uint32_t r = k / chunk, j = k - r * chunk;
            const uint4 raw = __ldcg(reinterpret_cast<const uint4*>(gathered + (size_t)r * L.per_shard_rows + (size_t)si * chunk + j));
            gnss_b200_acq_row row;
            memcpy(&row, &raw, sizeof(row));
            rows[k] = row;
            if (row.valid && (row.max_val > best || (row.max_val == best && best_k != 0xffffffffu && k < best_k)))
                {
                    best = row.max_val;
                    best_k = k;
                    best_i = row.max_idx;
                }
        }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
        {
            const float ob = __shfl_xor_sync(0xffffffffu, best, o);
            const uint32_t okk = __shfl_xor_sync(0xffffffffu, best_k, o);
            const uint32_t oi = __shfl_xor_sync(0xffffffffu, best_i, o);
            if (ob > best || (ob == best && okk < best_k))
                {
                    best = ob;
                    best_k = okk;
                    best_i = oi;
                }
        }
    __syncwarp();
    if (lane == 0)
        {
            if (best_k == 0xffffffffu)
                {
                    best_k = 0;
                    best_i = 0;
                }
            gnss_b200_acq_result res;
            res.peak = best;
            res.code_phase = best_i;
            res.doppler_index = best_k;
            res.doppler_hz = -a.doppler_max + a.center[si] + (int32_t)a.doppler_step * (int32_t)best_k;  // :535
            res.dwell_count = a.dwell_after[si];
            res.second_peak = 0.f;
            const uint32_t opp = (best_k + a.nD / 2) % a.nD;  // :533
            const gnss_b200_acq_row ro = rows[opp];
            const float per = ro.row_sum / (float)(int32_t)a.eff;
            res.input_power = (ro.valid && ok) ? (float)(((double)per / 2.0) / (double)a.dwell_after[si]) : nanf("");
            res.test_statistic = best / res.input_power;  // :542
            a.results[si] = res;
            if (bump_ticket)
                {
                    // every CTA read the epoch before it got here; the last one through moves it on
                    __threadfence();
                    if (atomicAdd(bump_ticket, 1u) == gridDim.x - 1u)
                        {
                            *bump_ticket = 0u;
                            *epoch_ctr = epoch;
                        }
                }
        }
}

static __global__ void peer_epoch_bump_kernel(uint32_t* epoch_ctr) { *epoch_ctr += 1u; }

// Rendezvous of all shards ON THE DEVICE (one warp): every shard stores its next rendezvous number into EVERY region and waits
// until all `world` counters of its own region have reached it. Shards leave within an NVLink store latency of each other -
// what a time-critical caller (or a benchmark) puts in front of a search so that it starts on all GPUs together; ~2 s time-out.
static __global__ void peer_barrier_kernel(char* const* peers, char* own, const PeerLayout L, uint32_t* ctr)
{
    const uint32_t lane = threadIdx.x;
    const uint32_t n = *ctr + 1u;
    __syncwarp();
    if (lane == 0) *ctr = n;
    for (uint32_t d = lane; d < L.world; d += 32) st_release_sys(reinterpret_cast<uint32_t*>(peers[d] + L.bar_off + (size_t)L.rank * 128u), n);
    for (uint32_t r = lane; r < L.world; r += 32)
        {
            const uint32_t* flag = reinterpret_cast<const uint32_t*>(own + L.bar_off + (size_t)r * 128u);
            const long long t0 = clock64();
            while ((int32_t)(ld_acquire_sys(flag) - n) < 0)
                if (clock64() - t0 > 4000000000ll) break;
        }
}

// The tail of a peer-connected shard's search in ONE kernel (instead of merge_partials + finalize + pack_rows + peer_scatter): CTA
// si folds the per-tile records of its slot's rows (the arithmetic of merge_partials_kernel / finalize_kernel, so the records
// are bit-identical to theirs), stores them straight into EVERY shard's recv[parity][rank] and, when the last CTA of the grid is
// through, raises this shard's flag in every region. `ticket` (own region, offset 64) counts finished CTAs and is left at 0.
static __global__ void shard_tail_kernel(const FinalizeArgs a, char* const* peers, const PeerLayout L, const uint32_t* epoch_ctr, uint32_t chunk, uint32_t* ticket)
{
    const uint32_t epoch = *epoch_ctr + 1u, par = epoch & 1u;
    const uint32_t si = blockIdx.x, lane = threadIdx.x & 31u, warp = threadIdx.x >> 5, n_warps = blockDim.x >> 5;
    for (uint32_t j = warp; j < chunk; j += n_warps)
        {
            gnss_b200_acq_row r;
            r.max_val = 0.f;
            r.max_idx = 0;
            r.row_sum = 0.f;
            r.valid = 0;
            if (j < a.nD_local)
                {
                    const RowPartial* p = a.partials + ((size_t)si * a.nD_local + j) * a.tiles;
                    float m, s;
                    uint32_t mi;
                    if (a.tiles > 8)
                        {
                            m = -1.f, s = 0.f, mi = 0xffffffffu;
                            for (uint32_t t = lane; t < a.tiles; t += 32)
                                {
                                    const RowPartial q = p[t];
                                    s += q.sum;
                                    if (q.max_val > m || (q.max_val == m && q.max_idx < mi))
                                        {
                                            m = q.max_val;
                                            mi = q.max_idx;
                                        }
                                }
#pragma unroll
                            for (int o = 16; o > 0; o >>= 1)
                                {
                                    const float om = __shfl_xor_sync(0xffffffffu, m, o);
                                    const uint32_t oi = __shfl_xor_sync(0xffffffffu, mi, o);
                                    s += __shfl_xor_sync(0xffffffffu, s, o);
                                    if (om > m || (om == m && oi < mi))
                                        {
                                            m = om;
                                            mi = oi;
                                        }
                                }
                        }
                    else
                        {
                            m = p[0].max_val, s = p[0].sum, mi = p[0].max_idx;
                            for (uint32_t t = 1; t < a.tiles; t++)
                                {
                                    s += p[t].sum;
                                    if (p[t].max_val > m || (p[t].max_val == m && p[t].max_idx < mi))
                                        {
                                            m = p[t].max_val;
                                            mi = p[t].max_idx;
                                        }
                                }
                        }
                    r.max_val = m;
                    r.max_idx = mi;
                    r.row_sum = s;
                    r.valid = 1;
                }
            uint4 raw;
            memcpy(&raw, &r, sizeof(raw));
            for (uint32_t d = lane; d < L.world; d += 32)
                {
                    uint4* dst = reinterpret_cast<uint4*>(peers[d] + L.recv_off) + (size_t)(par * L.world + L.rank) * L.per_shard_rows;
                    dst[(size_t)si * chunk + j] = raw;
                }
        }
    __threadfence_system();
    __syncthreads();
    __shared__ uint32_t s_last;
    if (threadIdx.x == 0) s_last = (atomicAdd(ticket, 1u) == gridDim.x - 1u) ? 1u : 0u;
    __syncthreads();
    if (s_last)
        {
            __threadfence_system();
            if (threadIdx.x == 0) *ticket = 0u;
            for (uint32_t d = threadIdx.x; d < L.world; d += blockDim.x)
                st_release_sys(reinterpret_cast<uint32_t*>(peers[d] + L.flags_off + (size_t)(par * L.world + L.rank) * 128u), epoch);
        }
}

// ---- first_vs_second_peak_statistic (:546-612): second maximum of the best row outside the +-1 chip window ----
struct SecondPeakArgs
{
    const float* mag_base;  // grid [slot][nD_local][eff] or row buffer [si][eff]
    uint32_t is_rowbuf;
    uint32_t nD_local;
    uint32_t eff, N;
    uint32_t samples_per_chip;
    const uint32_t* slots;
    const uint32_t* best_k_local;
    gnss_b200_acq_result* results;
};

static __global__ void second_peak_kernel(const SecondPeakArgs a)
{
    __shared__ float s_max[32];
    const uint32_t si = blockIdx.x;
    const float* row = a.is_rowbuf ? a.mag_base + (size_t)si * a.eff : a.mag_base + ((size_t)a.slots[si] * a.nD_local + a.best_k_local[si]) * a.eff;
    gnss_b200_acq_result res = a.results[si];
    int32_t ex1 = (int32_t)res.code_phase - (int32_t)a.samples_per_chip;  // :580-591
    int32_t ex2 = (int32_t)res.code_phase + (int32_t)a.samples_per_chip;
    if (ex1 < 0)
        ex1 += (int32_t)a.N;
    else if (ex2 >= (int32_t)a.N)
        ex2 -= (int32_t)a.N;
    float best = 0.f;  // beyond eff (bit transition) and inside the window the reference's buffer holds zeros
    for (uint32_t i = threadIdx.x; i < a.eff; i += blockDim.x)
        {
            const int32_t ii = (int32_t)i;
            const bool excluded = (ex1 < ex2) ? (ii >= ex1 && ii < ex2) : (ii >= ex1 || ii < ex2);
            if (!excluded) best = fmaxf(best, row[i]);
        }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) best = fmaxf(best, __shfl_xor_sync(0xffffffffu, best, o));
    if ((threadIdx.x & 31) == 0) s_max[threadIdx.x >> 5] = best;
    __syncthreads();
    if (threadIdx.x == 0)
        {
            for (uint32_t w = 1; w < (blockDim.x + 31) / 32; w++) best = fmaxf(best, s_max[w]);
            res.second_peak = best;
            res.test_statistic = res.peak / best;  // :611
            a.results[si] = res;
        }
}

}  // namespace b200
