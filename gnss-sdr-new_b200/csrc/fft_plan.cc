// Host planner for the shared-memory mixed-radix FFT (see fft_plan.h).
#include "fft_plan.h"
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>

namespace b200
{
namespace cl
{
bool cluster_preferred_plan(uint32_t N, uint32_t* ra_lo, uint32_t* rb_lo, uint32_t* ra_hi, uint32_t* rb_hi);  // acq_cluster.cu
}
namespace
{
// Radices of one stage, lowest stride first. Odd radices go first so that the stride-1 dimension has an odd
// element stride between neighbouring threads (conflict-free 64-bit shared-memory accesses).
bool stage_radices(uint32_t M, std::vector<uint32_t>* out)
{
    out->clear();
    uint32_t n2 = 0, n3 = 0, n5 = 0;
    std::vector<uint32_t> primes;
    uint32_t m = M;
    while (m % 2 == 0) m /= 2, n2++;
    while (m % 3 == 0) m /= 3, n3++;
    while (m % 5 == 0) m /= 5, n5++;
    for (uint32_t p : {7u, 11u, 13u})
        while (m % p == 0) m /= p, primes.push_back(p);
    // larger primes (sample rates that are multiples of 1.023 MHz carry 31): no register codelet, the kernels run
    // them as a direct DFT of that length out of shared memory (fft_smem.cuh: smem_pass_any)
    for (uint32_t p = 17; p <= kMaxAnyRadix && m > 1; p += 2)
        while (m % p == 0) m /= p, primes.push_back(p);
    if (m != 1) return false;
    std::vector<uint32_t> odd, even;
    while (n5 >= 2) odd.push_back(25), n5 -= 2;
    while (n3 >= 2) odd.push_back(9), n3 -= 2;
    if (n5 == 1 && n3 == 1) odd.push_back(15), n5 = n3 = 0;
    for (uint32_t p : primes) odd.push_back(p);
    if (n5 == 1)
        {
            if (n2 >= 2)
                even.push_back(20), n2 -= 2;
            else if (n2 == 1)
                even.push_back(10), n2 -= 1;
            else
                odd.push_back(5);
        }
    if (n3 == 1)
        {
            if (n2 >= 2)
                even.push_back(12), n2 -= 2;
            else if (n2 == 1)
                even.push_back(6), n2 -= 1;
            else
                odd.push_back(3);
        }
    while (n2 >= 4) even.push_back(16), n2 -= 4;
    if (n2 == 3) even.push_back(8);
    if (n2 == 2) even.push_back(4);
    if (n2 == 1) even.push_back(2);
    std::sort(odd.begin(), odd.end(), std::greater<uint32_t>());
    std::sort(even.begin(), even.end());
    for (uint32_t r : odd) out->push_back(r);
    for (uint32_t r : even) out->push_back(r);
    if (out->empty()) out->push_back(1);
    return out->size() <= (size_t)kMaxPasses;
}

void fill_stage(FftStage* st, uint32_t N, uint32_t M, const std::vector<uint32_t>& rad, uint32_t base_stride, bool upper)
{
    st->M = M;
    st->is_upper = upper ? 1 : 0;
    st->pos_stride = base_stride;
    st->tw_lo_mul = base_stride;
    st->n_pass = 0;
    std::vector<uint32_t> ls(rad.size());
    uint32_t s = 1;
    for (size_t j = 0; j < rad.size(); j++)
        {
            ls[j] = s;
            s *= rad[j];
        }
    for (int j = (int)rad.size() - 1; j >= 0; j--)
        {
            if (rad[j] == 1) continue;
            FftPass& p = st->pass[st->n_pass++];
            p.radix = rad[j];
            p.lstride = ls[j];
            const uint64_t gstride = (uint64_t)ls[j] * base_stride;
            p.tw_scale = (uint32_t)(N / (rad[j] * gstride));
            p.has_tw = gstride > 1 ? 1 : 0;
        }
}

// true when every prime factor of N has a codelet or fits the direct-DFT pass
bool primes_covered(uint32_t N)
{
    uint32_t m = N;
    for (uint32_t p = 2; p <= kMaxAnyRadix && m > 1; p++)
        while (m % p == 0) m /= p;
    return m == 1;
}

uint32_t pick_threads(uint32_t tile_points)
{
    uint32_t t = (tile_points / 16 + 31) / 32 * 32;
    return std::min(512u, std::max(128u, t));
}
}  // namespace

bool fast_ra_supported(uint32_t r) { return r == 4 || r == 5 || r == 8 || r == 16 || r == 25; }
bool fast_rb_supported(uint32_t r) { return r == 2 || r == 4 || r == 5 || r == 8 || r == 10 || r == 16 || r == 20; }

namespace
{
uint32_t best_columns(uint32_t M_lo, uint32_t M_hi)
{
    // the widest column tile that fits; a multiple of 16 columns (128-byte scratch lines, 64-byte grid runs) wins over a
    // wider odd one (measured on B200 at N = 160000: 16 beats 20 and 25)
    uint32_t C = 1, C16 = 0;
    for (uint32_t d = 1; d <= 32; d++)
        if (M_lo % d == 0 && (uint64_t)M_hi * d * 8 <= 64 * 1024)
            {
                C = d;
                if (d % 16 == 0) C16 = d;
            }
    return C16 ? C16 : C;
}

// N = (ra_lo*rb_lo) * (ra_hi*rb_hi) with every radix in the instantiated sets; lower is better
bool find_fast_plan(uint32_t N, FftPlan* plan)
{
    // lengths with a row-resident inverse kernel (acq_cluster.cuh) use the factorisation that kernel is instantiated for
    if (!std::getenv("B200_FFT_PLAN") && cl::cluster_preferred_plan(N, &plan->ra_lo, &plan->rb_lo, &plan->ra_hi, &plan->rb_hi)) return true;
    static const uint32_t RA[] = {25, 5, 16, 8, 4};
    static const uint32_t RB[] = {2, 4, 5, 8, 10, 16, 20};
    double best = 1e30;
    bool found = false;
    if (const char* e = std::getenv("B200_FFT_PLAN"))  // tuning aid: "ra_lo,rb_lo,ra_hi,rb_hi"
        {
            unsigned a1, b1, a2, b2;
            if (std::sscanf(e, "%u,%u,%u,%u", &a1, &b1, &a2, &b2) == 4 && (uint64_t)a1 * b1 * a2 * b2 == N && fast_ra_supported(a1) &&
                fast_rb_supported(b1) && fast_ra_supported(a2) && fast_rb_supported(b2))
                {
                    plan->ra_lo = a1;
                    plan->rb_lo = b1;
                    plan->ra_hi = a2;
                    plan->rb_hi = b2;
                    return true;
                }
        }
    for (uint32_t a1 : RA)
        for (uint32_t b1 : RB)
            for (uint32_t a2 : RA)
                for (uint32_t b2 : RB)
                    {
                        if ((uint64_t)a1 * b1 * a2 * b2 != N) continue;
                        const uint32_t lo = a1 * b1, hi = a2 * b2;
                        const uint32_t C = best_columns(lo, hi);
                        double cost = 0.5 * std::fabs(std::log2((double)lo / (double)hi));
                        if (a1 % 2 == 0) cost += 0.75;          // even stride-1 radix: padded shared-memory pitch
                        if (C < 16) cost += 2.0 + (16 - C) * 0.1;  // narrow column tiles coalesce badly
                        cost += 0.02 * (25.0 / a1 + 25.0 / a2);  // prefer big register passes first
                        if (cost < best)
                            {
                                best = cost;
                                found = true;
                                plan->ra_lo = a1;
                                plan->rb_lo = b1;
                                plan->ra_hi = a2;
                                plan->rb_hi = b2;
                            }
                    }
    return found;
}
}  // namespace

bool make_fft_plan(uint32_t N, uint32_t max_single, FftPlan* plan, const char** why)
{
    static const char* kPrime = "FFT length has a prime factor > 127 (no codelet and no direct-DFT pass)";
    static const char* kSplit = "FFT length cannot be split into two shared-memory stages";
    plan->N = N;
    plan->radices.clear();
    plan->fast = 0;
    std::vector<uint32_t> rad;
    static const bool single_first = std::getenv("B200_FFT_SINGLE") != nullptr;  // tuning aid: prefer the one-kernel-per-row plan for N <= max_single
    if (!(single_first && N <= max_single) && find_fast_plan(N, plan))
        {
            const uint32_t M_lo = plan->ra_lo * plan->rb_lo, M_hi = plan->ra_hi * plan->rb_hi;
            plan->fast = 1;
            plan->n_stages = 2;
            FftStage& up = plan->stage[0];
            fill_stage(&up, N, M_hi, {plan->ra_hi, plan->rb_hi}, M_lo, true);
            up.C = best_columns(M_lo, M_hi);
            if (const char* e = std::getenv("B200_FFT_COLS"))  // tuning aid: columns per upper-stage tile
                {
                    const uint32_t c = (uint32_t)std::atoi(e);
                    if (c >= 1 && c <= 32 && M_lo % c == 0 && (uint64_t)M_hi * c * 8 <= 96 * 1024) up.C = c;
                }
            up.blocks = 1;
            up.tiles = M_lo / up.C;
            up.tile_step = up.C;
            up.tile_points = M_hi * up.C;
            up.threads = pick_threads(up.tile_points);
            FftStage& lo = plan->stage[1];
            fill_stage(&lo, N, M_lo, {plan->ra_lo, plan->rb_lo}, 1, false);
            uint32_t b = 1;
            for (uint32_t d = 1; d <= M_hi; d++)
                if (M_hi % d == 0 && (uint64_t)d * M_lo <= 4096) b = d;
            if (const char* e = std::getenv("B200_FFT_NB"))  // tuning aid: sub-transforms per lower-stage tile
                {
                    const uint32_t c = (uint32_t)std::atoi(e);
                    if (c >= 1 && M_hi % c == 0 && (uint64_t)c * M_lo <= 8192) b = c;
                }
            lo.C = 1;
            lo.blocks = b;
            lo.tiles = M_hi / b;
            lo.tile_step = b * M_lo;
            lo.tile_points = b * M_lo;
            lo.threads = pick_threads(lo.tile_points);
            plan->radices = {plan->ra_lo, plan->rb_lo, plan->ra_hi, plan->rb_hi};
            return true;
        }
    if (N <= max_single)
        {
            if (!stage_radices(N, &rad))
                {
                    // too many passes for one stage (or bad prime): fall through to the two-stage search
                    if (!primes_covered(N))
                        {
                            if (why) *why = kPrime;
                            return false;
                        }
                }
            else
                {
                    plan->n_stages = 1;
                    FftStage& st = plan->stage[0];
                    fill_stage(&st, N, N, rad, 1, false);
                    st.C = 1;
                    st.blocks = 1;
                    st.tiles = 1;
                    st.tile_step = N;
                    st.tile_points = N;
                    st.threads = pick_threads(N);
                    plan->radices = rad;
                    return true;
                }
        }
    // two stages: N = M_hi * M_lo, both small enough for a comfortable tile
    const uint32_t kMaxStage = 8192;
    uint32_t best_lo = 0;
    double best_cost = 1e30;
    std::vector<uint32_t> r_lo, r_hi;
    for (uint32_t lo = 2; (uint64_t)lo * 2 <= N; lo++)
        {
            if (N % lo) continue;
            const uint32_t hi = N / lo;
            if (lo > kMaxStage || hi > kMaxStage) continue;
            if (!stage_radices(lo, &r_lo) || !stage_radices(hi, &r_hi)) continue;
            uint32_t c = 1;
            for (uint32_t d = 1; d <= 32; d++)
                if (lo % d == 0) c = d;
            double cost = (double)(r_lo.size() + r_hi.size());
            cost += 0.25 * std::fabs(std::log2((double)lo / (double)hi));  // balance
            if (c < 16) cost += 2.0;                                          // narrow column tiles coalesce badly
            if ((uint64_t)hi * c * 8 > 96 * 1024) cost += 4.0;                // upper tile must fit shared memory
            if (cost < best_cost)
                {
                    best_cost = cost;
                    best_lo = lo;
                }
        }
    if (!best_lo)
        {
            if (why) *why = !primes_covered(N) ? kPrime : kSplit;
            return false;
        }
    const uint32_t M_lo = best_lo, M_hi = N / best_lo;
    stage_radices(M_lo, &r_lo);
    stage_radices(M_hi, &r_hi);
    plan->n_stages = 2;
    // forward order: upper stage first
    FftStage& up = plan->stage[0];
    fill_stage(&up, N, M_hi, r_hi, M_lo, true);
    uint32_t C = 1;
    for (uint32_t d = 1; d <= 32; d++)
        if (M_lo % d == 0 && (uint64_t)M_hi * d * 8 <= 96 * 1024) C = d;
    up.C = C;
    up.blocks = 1;
    up.tiles = M_lo / C;
    up.tile_step = C;
    up.tile_points = M_hi * C;
    up.threads = pick_threads(up.tile_points);
    FftStage& lo = plan->stage[1];
    fill_stage(&lo, N, M_lo, r_lo, 1, false);
    uint32_t b = 1;
    for (uint32_t d = 1; d <= M_hi; d++)
        if (M_hi % d == 0 && (uint64_t)d * M_lo <= 4096) b = d;
    lo.C = 1;
    lo.blocks = b;
    lo.tiles = M_hi / b;
    lo.tile_step = b * M_lo;
    lo.tile_points = b * M_lo;
    lo.threads = pick_threads(lo.tile_points);
    plan->radices = r_lo;
    plan->radices.insert(plan->radices.end(), r_hi.begin(), r_hi.end());
    return true;
}

std::vector<uint32_t> fft_plan_permutation(const FftPlan& plan)
{
    const uint32_t N = plan.N;
    std::vector<uint32_t> perm(N);
    std::vector<uint64_t> stride(plan.radices.size());
    uint64_t s = 1;
    for (size_t j = 0; j < plan.radices.size(); j++)
        {
            stride[j] = s;
            s *= plan.radices[j];
        }
    for (uint32_t pos = 0; pos < N; pos++)
        {
            uint64_t f = 0;
            for (size_t j = 0; j < plan.radices.size(); j++)
                {
                    const uint64_t u = (pos / stride[j]) % plan.radices[j];
                    f += u * (N / (plan.radices[j] * stride[j]));
                }
            perm[pos] = (uint32_t)f;
        }
    return perm;
}

}  // namespace b200
