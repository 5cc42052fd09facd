// Row-resident inverse path of the acquisition pipeline: ONE persistent kernel does what inv_lower_kernel + inv_upper_kernel do
// (acq_fast.cuh), with the intermediate between the two FFT stages kept ON CHIP instead of going through an HBM scratch.
//
//   pcps_acquisition.cc:665-686 per (PRN, Doppler bin) row:  X_k * conj-code-spectrum -> IFFT -> |.|^2 (+ accumulation) -> max / sum
//
// A thread-block CLUSTER of CS CTAs owns one row at a time (N = M_lo * M_hi points, 8 N bytes: 256 kB at N = 32 000 does not
// fit one SM, it fits the distributed shared memory of a pair):
//
//   lower stage   CTA r transforms M_hi / CS of the M_hi contiguous length-M_lo blocks, NB at a time. The X_k tile arrives by bulk
//                 copies (TMA, mbarrier) into a two-stage ring that runs ahead across rows, and is transformed in place; the
//                 matching C* operands are loaded straight into registers one tile ahead (each is used exactly once);
//                 the outputs of pass B go straight to the CTA that owns their COLUMN in the upper stage:
//                 local shared memory or `st.shared::cluster` into the peer's (DSMEM) — this is the stage transpose.
//   cluster barrier
//   upper stage   CTA r transforms its M_lo / CS columns (length M_hi, stride M_lo in the row) in place in its U buffer,
//                 then |.|^2, the optional persistent-grid accumulation, and the per-CTA {max, first argmax, sum} record.
//   cluster barrier (split: arrive after the upper stage, wait before the next row's first remote store)
//
// The arithmetic (codelets, twiddle tables, operation order) is exactly that of the two-kernel path, so records and grids are
// bit-identical to it; only where the intermediate lives differs. HBM traffic per row drops from 16 N bytes to the operands,
// which L2 serves (X_k is shared by all PRNs, C* by all bins).
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

// Geometry of one instantiation (everything compile-time).
template <int RA_, int RB_, int RAH_, int RBH_, int CS_, int NB_, int THREADS_, int CTAS_PER_SM_, int STAGES_ = 2>
struct Geo
{
    static constexpr int RA = RA_, RB = RB_, RAH = RAH_, RBH = RBH_, CS = CS_, NB = NB_, THREADS = THREADS_, CTAS_PER_SM = CTAS_PER_SM_;
    static constexpr int M_LO = RA * RB, M_HI = RAH * RBH, N = M_LO * M_HI;
    static constexpr int CW = M_LO / CS;      // upper-stage columns owned by one CTA
    static constexpr int BPC = M_HI / CS;     // lower-stage blocks transformed by one CTA
    static constexpr int NT = BPC / NB;       // operand tiles per row and CTA
    static constexpr int SEG = N / RA;        // F layout (fft_plan.h): distance between consecutive u_a
    static constexpr int TCOLS = NB * RB;     // elements of one u_a segment inside a tile (contiguous in the F layout)
    // shared-memory pitch of a segment: rows stay 16-byte aligned (bulk copy) and TP = 2 (mod 16), i.e. a 4-bank skew per row —
    // the best an even pitch can do for the transposed 64-bit reads of pass B (2-way)
    static constexpr int TP = TCOLS + ((2 + 16 - (TCOLS & 15)) & 15);
    static constexpr int TILE = RA * TP;      // float2 per staged X tile
    static constexpr int kStages = STAGES_;
    // U buffer [ub][u][CW] with the ub-group stride padded to G = CW (mod 16): a warp that runs over the end of a group continues
    // on the banks right after it (same rule as upper_group_stride of the two-kernel path)
    static constexpr int G0 = RAH * CW;
    static constexpr int GS = G0 + ((CW + 16 - (G0 & 15)) & 15);
    static constexpr int U_ELEMS = (RBH * GS + 1) / 2 * 2;
    static constexpr int NW = THREADS / 32;   // per-warp records: CS * NW per row
    static_assert(M_LO % CS == 0 && M_HI % CS == 0 && BPC % NB == 0, "cluster split");
    static_assert(CW % RA == 0, "a pass-B output pencil must not straddle two CTAs' column ranges");
    static_assert(TCOLS <= THREADS, "pass A: one pencil per thread (its code-spectrum operands live in registers)");
    static_assert((TCOLS * 8) % 16 == 0 && (SEG * 8) % 16 == 0 && (TP * 8) % 16 == 0, "bulk copies need 16-byte granularity");
    static constexpr size_t smem_bytes = (size_t)(U_ELEMS + kStages * TILE + M_LO + M_HI) * sizeof(float2) + kStages * sizeof(uint64_t);
};

// MODE 2: records only (CFAR search, no grid kept). MODE 1: persistent magnitude grid with full rows (max_dwells > 1 / dump).
template <class G, int MODE>
__global__ void __launch_bounds__(G::THREADS, G::CTAS_PER_SM) inv_cluster_kernel(const FastArgs a, const uint32_t n_rows, const uint32_t dbg)
{
    constexpr int RA = G::RA, RB = G::RB, RAH = G::RAH, RBH = G::RBH, CS = G::CS, NB = G::NB, NT = G::NT, kT = G::THREADS;
    constexpr int M_LO = G::M_LO, M_HI = G::M_HI, N = G::N, CW = G::CW, BPC = G::BPC, SEG = G::SEG;
    constexpr int TCOLS = G::TCOLS, TP = G::TP, TILE = G::TILE, GS = G::GS, S = G::kStages;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    float2* U = reinterpret_cast<float2*>(smem_raw);   // this CTA's columns of the row between the stages
    float2* xs = U + G::U_ELEMS;                       // [S][RA][TP]   X_k tiles; pass A transforms a tile in place
    float2* tw_lo = xs + S * TILE;                     // conj W_Mlo^(q*ub), [ub][q]
    float2* tw_hi = tw_lo + M_LO;                      // conj W_Mhi^(q*ub), [ub][q]
    uint64_t* full = reinterpret_cast<uint64_t*>(tw_hi + M_HI);  // [S] "X tile landed"

    const uint32_t tid = threadIdx.x;
    const uint32_t rank = CS > 1 ? cluster_rank() : 0u;
    const uint32_t cluster_id = blockIdx.x / CS, n_clusters = gridDim.x / CS;

    // ---- once per CTA: twiddle tables, barriers, DSMEM base addresses ----
    for (uint32_t i = tid; i < (uint32_t)M_LO; i += kT)
        {
            const uint32_t ub = i / RA, q = i - ub * RA;
            const float2 w = __ldg(a.W + (size_t)q * ub * (N / M_LO));
            tw_lo[i] = make_float2(w.x, -w.y);
        }
    for (uint32_t i = tid; i < (uint32_t)M_HI; i += kT)
        {
            const uint32_t ub = i / RAH, q = i - ub * RAH;
            const float2 w = __ldg(a.W + (size_t)q * ub * (N / M_HI));
            tw_hi[i] = make_float2(w.x, -w.y);
        }
    if (tid == 0)
        {
            for (int s = 0; s < S; s++) mbar_init(&full[s], 1);
            fence_barrier_init();
        }
    uint32_t u_base[CS];
#pragma unroll
    for (int d = 0; d < CS; d++) u_base[d] = CS > 1 ? map_rank(smem_u32(U), d) : smem_u32(U);
    __syncthreads();

    const uint32_t my_rows = cluster_id < n_rows ? (n_rows - cluster_id + n_clusters - 1) / n_clusters : 0;
    const uint32_t total_tiles = my_rows * NT;

    // operand tile g of this CTA: row g / NT of the cluster, blocks [rank*BPC + (g % NT)*NB, +NB)
    auto operands = [&](uint32_t g, const float2*& X, const float2*& Cp) {
        const uint32_t y = a.y0 + cluster_id + (g / NT) * n_clusters;
        uint32_t si, kk;
        fast_row_of(a, y, si, kk);
        const uint32_t col0 = (rank * BPC + (g % NT) * NB) * RB;
        X = a.X + ((size_t)a.slot_group[si] * a.nD_local + kk) * N + col0;
        Cp = a.Cs + (size_t)a.slots[si] * N + col0;
    };
    // producer (thread 0): X_k tile g -> stage g % S, RA bulk copies of TCOLS elements (TMA, 1-D)
    auto issue = [&](uint32_t g) {
        const float2 *X, *Cp;
        operands(g, X, Cp);
        uint64_t* bar = &full[g % S];
        float2* dst = xs + (g % S) * TILE;
        mbar_expect_tx(bar, (uint32_t)(RA * TCOLS * sizeof(float2)));
#pragma unroll 1
        for (int u = 0; u < RA; u++) bulk_g2s(dst + u * TP, X + (size_t)u * SEG, TCOLS * sizeof(float2), bar);
    };
    // the code-spectrum operands of this thread's pass-A pencil of tile g, straight from L2 into registers one tile ahead
    float2 c[RA];
    auto load_code = [&](uint32_t g) {
        const float2 *X, *Cp;
        operands(g, X, Cp);
#pragma unroll
        for (int u = 0; u < RA; u++) c[u] = __ldg(Cp + (size_t)u * SEG + tid);
    };
    if (tid == 0)
        for (uint32_t g = 0; g < (uint32_t)S && g < total_tiles; g++) issue(g);
    if (tid < (uint32_t)TCOLS && total_tiles > 0) load_code(0);

    uint32_t g = 0;  // running tile counter of this CTA
    for (uint32_t r = 0; r < my_rows; r++)
        {
            const uint32_t y = a.y0 + cluster_id + r * n_clusters;
            uint32_t si, kk;
            fast_row_of(a, y, si, kk);

            // ================= lower stage =================
#pragma unroll 1
            for (int t = 0; t < NT; t++, g++)
                {
                    float2* xt = xs + (g % S) * TILE;
                    if (!(dbg & 2u) || g < (uint32_t)S) mbar_wait(&full[g % S], (g / S) & 1u);
                    // ---- pass A: pencil p = (block bl, ub): product, radix-RA IDFT, inner twiddle, in place ----
                    if (tid < (uint32_t)TCOLS && !(dbg & 8u))
                        {
                            const uint32_t p = tid, ub = p % RB;
                            float2 x[RA];
#pragma unroll
                            for (int u = 0; u < RA; u++) x[u] = fft::cmul(xt[u * TP + p], c[u]);  // :665
                            fft::Dft<RA, true>::run(x);
#pragma unroll
                            for (int q = 0; q < RA; q++) xt[q * TP + p] = fft::cmul(x[q], tw_lo[ub * RA + q]);
                            if (g + 1 < total_tiles && !(dbg & 128u)) load_code(g + 1);  // in flight behind pass B (and the upper stage)
                        }
                    __syncthreads();
                    if (t == 0 && r > 0 && CS > 1 && !(dbg & 64u)) cluster_wait();  // every CTA of the cluster is done READING its U of the previous row
                    // ---- pass B: pencil p = (block bl, a): radix-RB IDFT, inter-stage twiddle, to the owner of each column ----
                    for (uint32_t p = tid; p < (uint32_t)(NB * RA) && !(dbg & 16u); p += kT)
                        {
                            const uint32_t bl = p / RA, aa = p - bl * RA;
                            const uint32_t gblk = rank * BPC + t * NB + bl;
                            const uint32_t ubh = gblk / RAH, uah = gblk - ubh * RAH;
                            const uint32_t k1 = uah * RBH + ubh;  // upper-stage frequency index of this block
                            const float2 wb = __ldg(a.W + (size_t)k1 * aa);
                            const float2 ws = __ldg(a.W + (size_t)k1 * RA);
                            float2 v[RB];
#pragma unroll
                            for (int u = 0; u < RB; u++) v[u] = xt[aa * TP + bl * RB + u];
                            fft::Dft<RB, true>::run(v);
                            float2 tw = make_float2(wb.x, -wb.y);
                            const float2 stp = make_float2(ws.x, -ws.y);
                            const uint32_t e0 = ubh * GS + uah * CW + aa;  // U element of column aa in the owner's buffer
#pragma unroll
                            for (int b = 0; b < RB; b++)
                                {
                                    constexpr int kPer = CW / RA;  // b values per CTA column range
                                    const int d = b / kPer;          // owner CTA of column aa + RA*b
                                    const uint32_t e = e0 + RA * (b % kPer);
                                    const float2 val = fft::cmul(v[b], tw);
                                    if (CS == 1 || (uint32_t)d == rank || (dbg & 1u))
                                        U[e] = val;
                                    else
                                        st_cluster(u_base[d] + e * (uint32_t)sizeof(float2), val);
                                    tw = fft::cmul(tw, stp);
                                }
                        }
                    fence_proxy_async();  // this thread's generic-proxy accesses to the tile before the async-proxy refill
                    __syncthreads();
                    if (tid == 0 && g + S < total_tiles && !(dbg & 2u)) issue(g + S);  // refill the stage with the tile S ahead
                }
            if (CS > 1 && !(dbg & 64u))
                {
                    cluster_arrive();  // my remote stores of this row are done ...
                    cluster_wait();    // ... and so are everybody's into my U
                }

            // ================= upper stage (in place in U) =================
            // ---- pass A: pencil (column cl, ub) ----
            for (uint32_t p = tid; p < (uint32_t)(CW * RBH) && !(dbg & 4u); p += kT)
                {
                    const uint32_t ub = p / CW, cl = p - ub * CW;
                    float2 v[RAH];
                    float2* col = U + ub * GS + cl;
#pragma unroll
                    for (int u = 0; u < RAH; u++) v[u] = col[u * CW];
                    fft::Dft<RAH, true>::run(v);
#pragma unroll
                    for (int q = 0; q < RAH; q++) col[q * CW] = fft::cmul(v[q], tw_hi[ub * RAH + q]);
                }
            __syncthreads();
            // ---- pass B + |.|^2 + reductions: pencil (column cl, a) ----
            float best = -1.f, sum = 0.f;
            uint32_t best_pos = 0xffffffffu;
            float* mag = nullptr;
            bool accumulate = false;
            if (MODE == 1)
                {
                    mag = a.mag_base + ((size_t)a.slots[si] * a.nD_local + kk) * N;
                    accumulate = a.dwell_after[si] > 1;
                }
            for (uint32_t p = tid; p < (uint32_t)(CW * RAH) && !(dbg & 32u); p += kT)
                {
                    const uint32_t aa = p / CW, cl = p - aa * CW;
                    const uint32_t pos0 = aa * M_LO + rank * CW + cl;
                    float old[MODE == 1 ? RBH : 1];
                    if (MODE == 1)
                        {
#pragma unroll
                            for (int b = 0; b < RBH; b++) old[b] = accumulate ? __ldcs(mag + pos0 + b * (RAH * M_LO)) : 0.f;
                        }
                    float2 v[RBH];
#pragma unroll
                    for (int u = 0; u < RBH; u++) v[u] = U[u * GS + aa * CW + cl];
                    fft::Dft<RBH, true>::run(v);
                    float pbest = -1.f;
                    uint32_t pb = 0;
#pragma unroll
                    for (int b = 0; b < RBH; b++)
                        {
                            float gm = v[b].x * v[b].x + v[b].y * v[b].y;  // :674
                            if (MODE == 1)
                                {
                                    gm = old[b] + gm;  // :679 (0 + g on the first dwell)
                                    __stcs(mag + pos0 + b * (RAH * M_LO), gm);
                                }
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
            if (CS > 1 && r + 1 < my_rows && !(dbg & 64u)) cluster_arrive();  // this thread is done reading U: peers may overwrite it (next row)
            // one record per warp: maximum with the lowest position among equals (:523), and the warp's share of the row sum
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
            if ((tid & 31u) == 0)
                {
                    RowPartial rp;
                    rp.max_val = best;
                    rp.max_idx = best_pos;
                    rp.sum = sum;
                    rp.pad = 0;
                    a.partials[((size_t)y * CS + rank) * G::NW + (tid >> 5)] = rp;
                }
        }
    if (CS > 1)
        {
            // nobody leaves while a peer may still address its shared memory
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
