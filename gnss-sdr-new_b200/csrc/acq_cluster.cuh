// Row-resident inverse path of the acquisition pipeline: ONE persistent kernel does what inv_lower_kernel + inv_upper_kernel do
// (acq_fast.cuh), with the intermediate between the two FFT stages kept ON CHIP instead of going through an HBM scratch.
//
//   pcps_acquisition.cc:665-686 per (PRN, Doppler bin) row:  X_k * conj-code-spectrum -> IFFT -> |.|^2 (+ accumulation) -> max / sum
//
// A thread-block CLUSTER of CS CTAs owns one row at a time (N = M_lo * M_hi points, 8 N bytes: 256 kB at N = 32 000 does not
// fit one SM, it fits the distributed shared memory of a pair):
//
//   lower stage   CTA r transforms M_hi / CS of the M_hi contiguous length-M_lo blocks; the outputs of pass B go straight to the
//                 CTA that owns their COLUMN in the upper stage: local shared memory or `st.shared::cluster` into the peer's
//                 (DSMEM) — this is the stage transpose;
//   cluster barrier
//   upper stage   CTA r transforms its M_lo / CS columns (length M_hi, stride M_lo in the row) in place in its U buffer,
//                 then |.|^2, the optional persistent-grid accumulation, and per-warp {max, first argmax, sum} records;
//   cluster barrier (split: arrive after the upper stage, wait before the next row's first remote store).
//
// The lower stage is WARP-LOCAL: a warp owns whole blocks, so its transpose needs __syncwarp only. The upper stage is warp-local too
// with one column per half-warp (GeoW<..., UG = 0>), or - the default, UG = 4 - runs in groups of four warps that deal the pencils of
// their eight columns out so that every lane has work (one 128-thread named barrier between its two passes); there is no CTA-wide
// barrier in a records-only search. Operands: either straight from L2 into registers
// one tile ahead (default: every element is used exactly once, and ncu shows shared-memory bandwidth to be this kernel's
// scarcest resource), or the X_k tiles by bulk copies (TMA, mbarrier ring) with C* in registers (B200_ACQ_CLUSTER=1).
//
// The arithmetic (codelets, twiddle tables, operation order) is exactly that of the two-kernel path on the same FFT plan, so grids
// are bit-identical to it and records differ only in the summation order of the row sum. DRAM traffic per C2 dwell drops from
// 2.05 GB to 34 MB (profiles/r02_cluster_ncu_summary.txt). What was measured on the way (profiles/r02_cluster_experiments.md):
// CTA-barrier versions with TMA rings (0.91-1.08 ms), a 4-CTA cluster with two CTAs per SM, and a warp-specialised version with
// the two stages pipelined over a double-buffered U and remote mbarrier arrivals (1.09 ms: the kernel is issue-bound, overlap
// buys nothing) — all bit-identical, all slower than this one (0.63 ms with the grouped upper stage, 0.75 ms without; two-kernel
// path 0.83 ms).
#pragma once
#include "acq_fast.cuh"
#include <cooperative_groups.h>

namespace b200
{
namespace cl
{
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint32_t cluster_rank()
{
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ uint32_t map_rank(uint32_t addr, uint32_t rank)  // shared::cta address -> the same offset in CTA `rank`
{
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
    return r;
}
__device__ __forceinline__ void st_cluster(uint32_t addr, float2 v)
{
    asm volatile("st.shared::cluster.v2.f32 [%0], {%1, %2};" ::"r"(addr), "f"(v.x), "f"(v.y) : "memory");
}
__device__ __forceinline__ void cluster_arrive() { asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory"); }
__device__ __forceinline__ void cluster_wait() { asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory"); }
// arrival that publishes nothing (the warp only READ the buffer the barrier guards): no fence in front of it
__device__ __forceinline__ void cluster_arrive_relaxed() { asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory"); }
// operand loads: read-only path, not kept in L1 (each element is used once; L1 is left to the few spilled registers)
__device__ __forceinline__ float2 ldg_stream(const float2* p)
{
    float2 v;
    asm volatile("ld.global.nc.L1::no_allocate.v2.f32 {%0, %1}, [%2];" : "=f"(v.x), "=f"(v.y) : "l"(p));
    return v;
}

// ---- mbarrier + bulk copy (TMA, 1-D) ----
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
// global -> this CTA's shared memory, `bytes` a multiple of 16, both addresses 16-byte aligned; completes on `bar`
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst_smem)),
        "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}
__device__ __forceinline__ void fence_barrier_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// ------------------------------------------------------------------------------------------------------------------------------
// Warp-local variant. Same data flow, but no CTA-wide barrier anywhere in a records-only search:
//   lower stage   a warp owns GB = 32 / RB whole blocks of a tile (32 pencils in pass A); pass A -> __syncwarp -> pass B inside
//                 the warp's own 32-column window of the staged X tile (XOR-swizzled so that the transposed reads of pass B are
//                 conflict-free although the bulk-copied rows force an even pitch);
//   upper stage   a warp owns CW / NW whole columns of U (odd row pitch: both strided passes conflict-free);
//   operand ring  the warp that is LAST to finish a tile (shared-memory counter) refills its stage with the tile kStages ahead.
// Per row only the two cluster barriers remain (lower -> upper exchange complete; U free again).
// ------------------------------------------------------------------------------------------------------------------------------
// STAGES_ = 0: no operand staging at all — both operands are loaded straight from L2 into registers one tile ahead (each element is
// used exactly once, so staging only costs shared-memory bandwidth, which ncu shows to be this kernel's scarcest resource); the
// one "stage" left is the warps' transpose buffer between pass A and pass B.
// UG_ = 0: upper stage with one column per half-warp (20 and 25 pencils on 16 lanes: 45 of 64 lane slots used). UG_ = 4: groups of
// four warps own eight columns and deal their 8 x 20 + 8 x 25 pencils out in half-warp tiles of (2 pencils x 8 columns): three
// rounds per warp instead of four, 360 of 384 lane slots used; one named barrier of 128 threads between the two passes; row
// pitch CW + 8, for which those tiles - and the lower stage's stores - are bank-conflict-free.
template <int RA_, int RB_, int RAH_, int RBH_, int CS_, int THREADS_, int STAGES_, int CTAS_PER_SM_ = 1, int UG_ = 0>
struct GeoW
{
    static constexpr int UG = UG_;
    static constexpr int RA = RA_, RB = RB_, RAH = RAH_, RBH = RBH_, CS = CS_, THREADS = THREADS_, CTAS_PER_SM = CTAS_PER_SM_;
    static constexpr bool kTma = STAGES_ > 0;
    static constexpr int kStages = STAGES_ > 0 ? STAGES_ : 1;
    static constexpr int M_LO = RA * RB, M_HI = RAH * RBH, N = M_LO * M_HI;
    static constexpr int NW = THREADS / 32;
    static constexpr int CW = M_LO / CS;      // upper-stage columns owned by one CTA
    static constexpr int CPW = CW / NW;       // ... by one warp
    static constexpr int UP = UG_ ? CW + 8 : CW + 1;  // U row pitch (odd, or 8 mod 16 for the grouped upper stage)
    static constexpr int BPC = M_HI / CS;     // lower-stage blocks transformed by one CTA
    static constexpr int GB = 32 / RB;        // blocks per warp and tile
    static constexpr int TB = NW * GB;        // blocks per tile
    static constexpr int NT = (BPC + TB - 1) / TB;
    static constexpr int SEG = N / RA;
    static constexpr int TCOLS = TB * RB;     // = 32 * NW
    static constexpr int TP = TCOLS + 2;      // even pitch (16-byte aligned bulk-copy rows), 4-bank skew per row
    static constexpr int TILE = RA * TP;
    static constexpr int U_ELEMS = (M_HI * UP + 1) / 2 * 2;
    static_assert(M_LO % CS == 0 && M_HI % CS == 0, "cluster split");
    static_assert(32 % RB == 0 && (GB * RA) % 32 == 0, "a warp's blocks fill whole 32-lane rounds in both lower passes");
    static_assert(CW % NW == 0 && CW % RA == 0, "columns per warp / pass-B pencils per column range");
    static_assert(RA == 8 && RB == 8, "the pass-A swizzle below is derived for 8 x 8 blocks");
    static_assert(UG_ == 0 || (UG_ == 4 && CW / (THREADS_ / 32) == 2 && CW % 16 == 0 && RBH_ % 2 == 0 && (THREADS_ / 32) % 4 == 0), "grouped upper stage: 4 warps x 2 columns, pitch 8 mod 16");
    static_assert((BPC * RB * 8) % 16 == 0 && ((BPC % TB) * RB * 8) % 16 == 0 && (TP * 8) % 16 == 0, "bulk copies need 16-byte granularity");
    static constexpr int kRowTab = 256;       // rows of one cluster whose operand row numbers are looked up once, at kernel start
    static constexpr int TWL = RB * (RA + 1); // lower-stage inner twiddles [ub][q], pitch RA + 1: the 8 rows a warp reads lie in distinct banks
    static constexpr size_t smem_bytes = (size_t)(U_ELEMS + kStages * TILE + M_HI + TWL) * sizeof(float2) + kRowTab * 2 * sizeof(uint32_t) +
                                         kStages * (sizeof(uint64_t) + sizeof(uint32_t)) + 16;
};

template <class G, int MODE>
__global__ void __launch_bounds__(G::THREADS, G::CTAS_PER_SM) inv_cluster_wl_kernel(const FastArgs a, const uint32_t n_rows, const uint32_t dbg)
{
    constexpr int RA = G::RA, RB = G::RB, RAH = G::RAH, RBH = G::RBH, CS = G::CS, NT = G::NT, kT = G::THREADS, NW = G::NW;
    constexpr int M_LO = G::M_LO, M_HI = G::M_HI, N = G::N, CW = G::CW, CPW = G::CPW, UP = G::UP, BPC = G::BPC, SEG = G::SEG;
    constexpr int GB = G::GB, TB = G::TB, TP = G::TP, TILE = G::TILE, S = G::kStages;
    constexpr bool kTma = G::kTma;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    float2* U = reinterpret_cast<float2*>(smem_raw);   // [M_HI][UP]: this CTA's columns of the row between the stages
    float2* xs = U + G::U_ELEMS;                       // [S][RA][TP]: X_k tiles (kTma) / the warps' transpose windows
    float2* tw_hi = xs + S * TILE;                     // conj W_Mhi^(q*ub), [ub][q]
    float2* twl_s = tw_hi + M_HI;                      // conj W_Mlo^(q*ub), [ub][RA + 1]
    uint32_t* rowtab = reinterpret_cast<uint32_t*>(twl_s + G::TWL);  // [kRowTab][2]: X row, C* row of this cluster's r-th row
    uint64_t* full = reinterpret_cast<uint64_t*>(rowtab + 2 * G::kRowTab);  // [S] "tile landed"
    uint32_t* done = reinterpret_cast<uint32_t*>(full + S);      // [S] warps that have finished with the stage's tile

    const uint32_t tid = threadIdx.x, lane = tid & 31u, w = tid >> 5;
    const uint32_t rank = CS > 1 ? cluster_rank() : 0u;
    const uint32_t cluster_id = blockIdx.x / CS, n_clusters = gridDim.x / CS;

    for (uint32_t i = tid; i < (uint32_t)M_HI; i += kT)
        {
            const uint32_t ub = i / RAH, q = i - ub * RAH;
            const float2 t = __ldg(a.W + (size_t)q * ub * (N / M_HI));
            tw_hi[i] = make_float2(t.x, -t.y);
        }
    if (kTma && tid == 0)
        {
            for (int s = 0; s < S; s++)
                {
                    mbar_init(&full[s], 1);
                    done[s] = 0;
                }
            fence_barrier_init();
        }
    uint32_t u_base[CS];
#pragma unroll
    for (int d = 0; d < CS; d++) u_base[d] = CS > 1 ? map_rank(smem_u32(U), d) : smem_u32(U);
    // pass-A pencil of this lane inside its warp's window: block w*GB + lane / RB of the tile, ub = lane % RB — the same for every
    // tile. Its inner twiddles conj W_Mlo^(q*ub): the first kTwReg stay in registers, the others are read from a small shared
    // table at their use - all seven in registers were spilled (the kernel runs at its 128-register limit) and came back from local
    // memory with a long-scoreboard stall every tile; five (one with a persistent grid) is what fits without a spill.
    const uint32_t ub_a = lane % RB, bl_a = w * GB + lane / RB;
    for (uint32_t i = tid; i < (uint32_t)(RA * RB); i += kT)
        {
            const uint32_t ub = i / RA, q = i - ub * RA;
            const float2 t = __ldg(a.W + (size_t)q * ub * (N / M_LO));
            twl_s[ub * (RA + 1) + q] = make_float2(t.x, -t.y);
        }
    const float2* twl = twl_s + ub_a * (RA + 1);
    constexpr int kTwReg = MODE == 1 ? 1 : 5;  // (the persistent-grid form has fewer registers to spare)
    float2 twr[kTwReg + 1];
#pragma unroll
    for (int q = 1; q <= kTwReg; q++)
        {
            const float2 t = __ldg(a.W + (size_t)q * ub_a * (N / M_LO));
            twr[q] = make_float2(t.x, -t.y);
        }

    const uint32_t my_rows = cluster_id < n_rows ? (n_rows - cluster_id + n_clusters - 1) / n_clusters : 0;
    const uint32_t total_tiles = my_rows * NT;

    // operand rows (X_k row, C* row) of this cluster's rows: looked up once, here, by as many threads as there are rows - inside
    // the row loop the two dependent global loads per row (slot -> group / code slot) were a long-scoreboard stall per row
    auto row_numbers = [&](uint32_t r, uint32_t& xr, uint32_t& cr) {
        const uint32_t y = a.y0 + cluster_id + r * n_clusters;
        uint32_t si, kk;
        fast_row_of(a, y, si, kk);
        xr = a.slot_group[si] * a.nD_local + kk;
        cr = a.slots[si];
    };
    for (uint32_t r = tid; r < my_rows && r < (uint32_t)G::kRowTab; r += kT) row_numbers(r, rowtab[2 * r], rowtab[2 * r + 1]);
    __syncthreads();
    auto operands = [&](uint32_t g, const float2*& X, const float2*& Cp) {  // (the launcher never gives a cluster more than kRowTab rows)
        const uint32_t r = g / NT, col0 = rank * BPC * RB + (g - r * NT) * TB * RB;
        const uint2 rc = *reinterpret_cast<const uint2*>(rowtab + 2 * r);
        X = a.X + (size_t)rc.x * N + col0;
        Cp = a.Cs + (size_t)rc.y * N + col0;
    };
    auto tile_blocks = [&](uint32_t g) -> uint32_t {  // the last tile of a row share may be partial
        const uint32_t t = g % NT;
        return (t + 1) * TB <= (uint32_t)BPC ? (uint32_t)TB : (uint32_t)BPC - t * TB;
    };
    auto issue = [&](uint32_t g) {  // one lane: X_k tile g -> stage g % S (TMA, 1-D bulk copies)
        const float2 *X, *Cp;
        operands(g, X, Cp);
        const uint32_t bytes = tile_blocks(g) * RB * (uint32_t)sizeof(float2);
        uint64_t* bar = &full[g % S];
        float2* dst = xs + (g % S) * TILE;
        mbar_expect_tx(bar, RA * bytes);
#pragma unroll
        for (int u = 0; u < RA; u++) bulk_g2s(dst + u * TP, X + (size_t)u * SEG, bytes, bar);
    };
    // operands of this lane's pass-A pencil of tile g, straight from L2 into registers one tile ahead
    float2 cn[RA], xn[kTma ? 1 : RA];
    auto load_operands = [&](uint32_t g) {
        if (bl_a < tile_blocks(g))
            {
                const float2 *X, *Cp;
                operands(g, X, Cp);
#pragma unroll
                for (int u = 0; u < RA; u++) cn[u] = ldg_stream(Cp + (size_t)u * SEG + w * 32 + lane);
                if (!kTma)
                    {
#pragma unroll
                        for (int u = 0; u < RA; u++) xn[kTma ? 0 : u] = ldg_stream(X + (size_t)u * SEG + w * 32 + lane);
                    }
            }
    };
    if (kTma && tid == 0)
        for (uint32_t g = 0; g < (uint32_t)S && g < total_tiles; g++) issue(g);
    if (total_tiles > 0) load_operands(0);

    uint32_t g = 0;
    for (uint32_t r = 0; r < my_rows; r++)
        {
            const uint32_t y = a.y0 + cluster_id + r * n_clusters;
            uint32_t si, kk;
            fast_row_of(a, y, si, kk);

            // ================= lower stage: warp-local, tile after tile =================
#pragma unroll 1
            for (int t = 0; t < NT; t++, g++)
                {
                    float2* xt = xs + (kTma ? (g % S) * TILE : 0) + w * 32;  // this warp's 32-column window
                    const uint32_t nblk = tile_blocks(g);
                    static_assert(GB * RA == 32, "one pass-B pencil per lane and tile");
                    const uint32_t bl_b = w * GB + lane / RA, aa_b = lane % RA;  // this lane's pass-B pencil: block of the tile, column
                    // its inter-stage twiddles: asked for here, used after pass A
                    float2 wb = make_float2(1.f, 0.f), ws = make_float2(1.f, 0.f);
                    if (bl_b < nblk)
                        {
                            const uint32_t gblk = rank * BPC + t * TB + bl_b;
                            const uint32_t ubh = gblk / RAH, uah = gblk - ubh * RAH;
                            const uint32_t k1 = uah * RBH + ubh;  // upper-stage frequency index of this block
                            wb = __ldg(a.W + (size_t)k1 * aa_b);
                            ws = __ldg(a.W + (size_t)k1 * RA);
                        }
                    if (kTma && (!(dbg & 2u) || g < (uint32_t)S)) mbar_wait(&full[g % S], (g / S) & 1u);
                    // ---- pass A: product, radix-RA IDFT, inner twiddle; stored swizzled (column ^ (q >= RA/2)) ----
                    {
                        const bool act = bl_a < nblk && !(dbg & 8u);
                        float2 x[RA];
                        if (act)
                            {
#pragma unroll
                                for (int u = 0; u < RA; u++) x[u] = fft::cmul(kTma ? xt[u * TP + lane] : xn[kTma ? 0 : u], cn[u]);  // :665
                            }
                        if (g + 1 < total_tiles && !(dbg & 128u)) load_operands(g + 1);  // in flight behind this tile (and the upper stage)
                        if (act) fft::Dft<RA, true>::run(x);
                        if (kTma) __syncwarp();  // every lane has read its inputs before any lane overwrites the window
                        if (act)
                            {
#pragma unroll
                                for (int q = 0; q < RA; q++) xt[q * TP + (lane ^ (uint32_t)(q >= RA / 2))] = q ? fft::cmul(x[q], q <= kTwReg ? twr[q <= kTwReg ? q : 0] : twl[q]) : x[q];  // twl[0] = 1
                            }
                    }
                    __syncwarp();
                    if (t == 0 && r > 0 && CS > 1 && !(dbg & 64u)) cluster_wait();  // every CTA is done READING its U of the previous row
                    // ---- pass B: radix-RB IDFT, inter-stage twiddle, to the owner of each column ----
                    if (bl_b < nblk && !(dbg & 16u))
                        {
                            const uint32_t bl_l = lane / RA, aa = aa_b, gblk = rank * BPC + t * TB + bl_b;
                            float2 v[RB];
                            const uint32_t swz = aa >= RA / 2;
#pragma unroll
                            for (int u = 0; u < RB; u++) v[u] = xt[aa * TP + ((bl_l * RB + u) ^ swz)];
                            fft::Dft<RB, true>::run(v);
                            float2 tw = make_float2(wb.x, -wb.y);
                            const float2 stp = make_float2(ws.x, -ws.y);
                            const uint32_t e0 = gblk * UP + aa;
#pragma unroll
                            for (int b = 0; b < RB; b++)
                                {
                                    constexpr int kPer = CW / RA;
                                    const int d = b / kPer;
                                    const uint32_t e = e0 + RA * (b % kPer);
                                    const float2 val = fft::cmul(v[b], tw);
                                    if (CS == 1 || (uint32_t)d == rank || (dbg & 1u))
                                        U[e] = val;
                                    else
                                        st_cluster(u_base[d] + e * (uint32_t)sizeof(float2), val);
                                    tw = fft::cmul(tw, stp);
                                }
                        }
                    if (kTma)
                        {
                            // the warp is done with the stage; the last warp to get here refills it
                            fence_proxy_async();
                            __syncwarp();
                            if (lane == 0)
                                {
                                    __threadfence_block();
                                    const uint32_t prev = atomicAdd(&done[g % S], 1u);
                                    if (prev == (uint32_t)NW - 1)
                                        {
                                            done[g % S] = 0;
                                            __threadfence_block();
                                            if (g + S < total_tiles && !(dbg & 2u)) issue(g + S);
                                        }
                                }
                        }
                    else
                        __syncwarp();  // pass B has read the window before the next tile's pass A overwrites it
                }
            if (CS > 1 && !(dbg & 64u))
                {
                    cluster_arrive();  // my remote stores of this row are done ...
                    cluster_wait();    // ... and so are everybody's into my U
                }
            else if (CS == 1)
                __syncthreads();

            // ================= upper stage: warp-local, CPW columns of U in place =================
            float best = -1.f, sum = 0.f;
            uint32_t best_pos = 0xffffffffu;
            if constexpr (G::UG == 0)
                {
                // lanes: half-warp h works on column (h % CPW) of the warp, lanes of a half-warp on 16 consecutive pencils of that
                // column: with the odd row pitch every 64-bit access of both strided passes is conflict-free
                static_assert(CPW == 2, "lane map of the upper stage: one column per half-warp");
                float2* Uw = U + w * CPW + (lane >> 4);
                const uint32_t l16 = lane & 15u;
#pragma unroll 1
                for (uint32_t ub = l16; ub < (uint32_t)RBH && !(dbg & 4u); ub += 16)
                    {
                        float2 v[RAH];
                        float2* col = Uw + (size_t)RAH * ub * UP;
#pragma unroll
                        for (int u = 0; u < RAH; u++) v[u] = col[u * UP];
                        fft::Dft<RAH, true>::run(v);
#pragma unroll
                        for (int q = 0; q < RAH; q++) col[q * UP] = fft::cmul(v[q], tw_hi[ub * RAH + q]);
                    }
                __syncwarp();
#pragma unroll 1
                for (uint32_t aa = l16; aa < (uint32_t)RAH && !(dbg & 32u); aa += 16)
                    {
                        float2 v[RBH];
                        float2* col = Uw + (size_t)aa * UP;
#pragma unroll
                        for (int u = 0; u < RBH; u++) v[u] = col[(size_t)RAH * u * UP];
                        fft::Dft<RBH, true>::run(v);
                        if (MODE == 1)
                            {
                                // persistent grid: park |.|^2 in the slots just consumed; the coalesced read-modify-write follows below
#pragma unroll
                                for (int b = 0; b < RBH; b++) col[(size_t)RAH * b * UP].x = v[b].x * v[b].x + v[b].y * v[b].y;  // :674
                            }
                        else
                            {
                                const uint32_t pos0 = aa * M_LO + rank * CW + w * CPW + (lane >> 4);
                                float pbest = -1.f;
                                uint32_t pb = 0;
#pragma unroll
                                for (int b = 0; b < RBH; b++)
                                    {
                                        const float gm = v[b].x * v[b].x + v[b].y * v[b].y;  // :674
                                        sum += gm;
                                        const bool gt = gm > pbest;
                                        pbest = gt ? gm : pbest;
                                        pb = gt ? (uint32_t)b : pb;
                                    }
                                const uint32_t ppos = pos0 + pb * (RAH * M_LO);
                                if (pbest > best || (pbest == best && ppos < best_pos))
                                    {
                                        best = pbest;
                                        best_pos = ppos;
                                    }
                            }
                    }
                }
            else
                {
                    // Groups of UG warps own UG * CPW = 8 columns. Half-warp tiles of (2 consecutive pencils x 8 columns): with the
                    // pitch 8 mod 16 every 64-bit access of both strided passes is conflict-free. Pass 1 has RBH / 2 = 10 tiles,
                    // pass 2 (RAH + 1) / 2 = 13 (the last one half full); they are dealt to the group's 8 half-warps so that every
                    // warp runs three rounds, and the role of a warp rotates with the group so that the one warp with two radix-RAH
                    // rounds sits on a different scheduler in each group.
                    constexpr uint32_t kHw = 2 * G::UG, GC = G::UG * CPW;
                    static_assert(GC == 8, "tile = 2 pencils x 8 columns");
                    const uint32_t grp = w / G::UG, role = (w + grp) % G::UG;
                    const uint32_t hw = role * 2 + (lane >> 4), c8 = lane & 7u, sub = (lane >> 3) & 1u;
                    float2* Ug = U + grp * GC + c8;
#pragma unroll 1
                    for (uint32_t j = hw; j < (uint32_t)RBH / 2 && !(dbg & 4u); j += kHw)
                        {
                            const uint32_t ub = 2 * j + sub;
                            float2 v[RAH];
                            float2* col = Ug + (size_t)RAH * ub * UP;
#pragma unroll
                            for (int u = 0; u < RAH; u++) v[u] = col[u * UP];
                            fft::Dft<RAH, true>::run(v);
#pragma unroll
                            for (int q = 0; q < RAH; q++) col[q * UP] = fft::cmul(v[q], tw_hi[ub * RAH + q]);
                        }
                    __syncwarp();
                    asm volatile("bar.sync %0, %1;" ::"r"(1u + grp), "n"(G::UG * 32) : "memory");  // the group's columns are through pass 1
#pragma unroll 1
                    for (uint32_t j = (hw + kHw - 2) % kHw; j < (uint32_t)(RAH + 1) / 2 && !(dbg & 32u); j += kHw)
                        {
                            const uint32_t aa = 2 * j + sub;
                            if (aa < (uint32_t)RAH)
                                {
                                    float2 v[RBH];
                                    float2* col = Ug + (size_t)aa * UP;
#pragma unroll
                                    for (int u = 0; u < RBH; u++) v[u] = col[(size_t)RAH * u * UP];
                                    fft::Dft<RBH, true>::run(v);
                                    if (MODE == 1)
                                        {
#pragma unroll
                                            for (int b = 0; b < RBH; b++) col[(size_t)RAH * b * UP].x = v[b].x * v[b].x + v[b].y * v[b].y;  // :674
                                        }
                                    else
                                        {
                                            const uint32_t pos0 = aa * M_LO + rank * CW + grp * GC + c8;
                                            float pbest = -1.f;
                                            uint32_t pb = 0;
#pragma unroll
                                            for (int b = 0; b < RBH; b++)
                                                {
                                                    const float gm = v[b].x * v[b].x + v[b].y * v[b].y;  // :674
                                                    sum += gm;
                                                    const bool gt = gm > pbest;
                                                    pbest = gt ? gm : pbest;
                                                    pb = gt ? (uint32_t)b : pb;
                                                }
                                            const uint32_t ppos = pos0 + pb * (RAH * M_LO);
                                            if (pbest > best || (pbest == best && ppos < best_pos))
                                                {
                                                    best = pbest;
                                                    best_pos = ppos;
                                                }
                                        }
                                }
                        }
                    __syncwarp();
                }
            if (MODE == 1)
                {
                    // rows of U are CW consecutive grid cells: one warp per row, old values streamed in, sums streamed out
                    __syncthreads();
                    float* mag = a.mag_base + ((size_t)a.slots[si] * a.nD_local + kk) * N + rank * CW;
                    const bool accumulate = a.dwell_after[si] > 1;
                    for (uint32_t blk = w; blk < (uint32_t)M_HI; blk += NW)
                        for (uint32_t cl = lane; cl < (uint32_t)CW; cl += 32)
                            {
                                const uint32_t pos = blk * M_LO + cl;
                                float gm = U[blk * UP + cl].x;
                                if (accumulate) gm = __ldcs(mag + pos) + gm;  // :679
                                __stcs(mag + pos, gm);
                                sum += gm;
                                const uint32_t gpos = pos + rank * CW;
                                if (gm > best || (gm == best && gpos < best_pos))
                                    {
                                        best = gm;
                                        best_pos = gpos;
                                    }
                            }
                }
            if (CS > 1 && r + 1 < my_rows && !(dbg & 64u)) cluster_arrive_relaxed();  // this warp is done reading U (nothing to publish)
#pragma unroll
            for (int o = 16; o > 0; o >>= 1)
                {
                    const float ob = __shfl_xor_sync(0xffffffffu, best, o);
                    const uint32_t op = __shfl_xor_sync(0xffffffffu, best_pos, o);
                    sum += __shfl_xor_sync(0xffffffffu, sum, o);
                    if (ob > best || (ob == best && op < best_pos))
                        {
                            best = ob;
                            best_pos = op;
                        }
                }
            if (lane == 0)
                {
                    RowPartial rp;
                    rp.max_val = best;
                    rp.max_idx = best_pos;
                    rp.sum = sum;
                    rp.pad = 0;
                    a.partials[((size_t)y * CS + rank) * NW + w] = rp;
                }
            if (CS == 1) __syncthreads();
        }
    if (CS > 1)
        {
            cluster_arrive();
            cluster_wait();
        }
}

// host-side launcher (instantiations in acq_cluster.cu).
// cluster_tiles: 0 when the plan has no row-resident instantiation, else the records per row the kernel writes (cluster size x warps).
bool cluster_preferred_plan(uint32_t N, uint32_t* ra_lo, uint32_t* rb_lo, uint32_t* ra_hi, uint32_t* rb_hi);  // fft_plan.cc asks
uint32_t cluster_tiles(uint32_t ra_lo, uint32_t rb_lo, uint32_t ra_hi, uint32_t rb_hi);
cudaError_t launch_inv_cluster(uint32_t ra_lo, uint32_t rb_lo, uint32_t ra_hi, uint32_t rb_hi, int mode, uint32_t n_rows, cudaStream_t st, const FastArgs& a);

}  // namespace cl
}  // namespace b200
