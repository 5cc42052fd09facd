// Host side of the PCPS acquisition engine: handle, buffers, launch sequencing, C-ABI.
// Reference for every entry point: include/gnss_b200.h.
#include "acq_cluster.cuh"
#include "acq_fast_fwd.cuh"
#include "acq_kernels.cuh"
#include "wipeoff_pieces.h"
#include <algorithm>
#include <cmath>
#include <map>
#include <vector>

using namespace b200;

namespace
{
constexpr uint32_t kMaxSingleStagePoints = 24576;  // 192 KiB of complex64 in one CTA's shared memory
constexpr size_t kScratchBudgetBytes = 2048ull << 20;  // cap of the two-stage intermediate (rows are chunked beyond it)
constexpr uint32_t kMaxCachedGroups = 64;
enum
{
    CAT_FWD = 0,
    CAT_INV_FIRST = 1,  // two-stage plans only: product + lower stage -> scratch
    CAT_INV_LAST = 2,   // IFFT last stage + |.|^2 + row reductions (single-stage plans: the whole inverse)
    CAT_CODE = 3,
    CAT_DECIDE = 4
};

struct CallMeta  // uploaded once per dwell call
{
    std::vector<uint32_t> slots, slot_group, dwell_after, group_piece_base;
    std::vector<int32_t> center;  // step-two handles: the bits of the float centre
    std::vector<float> input_power1;
};
}  // namespace

struct gnss_b200_acq
{
    std::mutex mu;
    gnss_b200_acq_cfg cfg{};
    uint32_t N = 0, consumed = 0, eff = 0, nD = 0, nD_local = 0, row_begin = 0;
    uint32_t code_len = 0, code_off = 0, mag_offset = 0;
    bool padded = false;  // the transform length is a supported M >= 2 * fft_size - 1 (fft_size has a prime factor the kernels cannot take)
    FftPlan plan;
    cudaStream_t stream = nullptr;
    cudaStream_t stream2 = nullptr;  // last inverse stage of chunk i overlaps the first stage of chunk i+1
    cudaEvent_t ev_lo[2] = {nullptr, nullptr}, ev_up[2] = {nullptr, nullptr}, ev_fork = nullptr;
    uint64_t launches = 0;

    DevBuf d_W, d_cbuf, d_xbuf, d_xtmp, d_samples, d_codes_td, d_scratch;
    DevBuf d_pieces, d_npieces;  // per cached group: nD_local tables
    std::map<int64_t, uint32_t> group_cache;  // (center + bias) -> cached group ordinal
    DevBuf d_meta;
    PinBuf h_meta;
    DevBuf d_partials, d_partials2, d_rows, d_results, d_bestk, d_grid, d_rowbuf;
    PinBuf h_results, h_samples;

    std::vector<int32_t> center, bias;
    // make_two_steps handles (gnss_b200_acq_create_step2): nD = num_doppler_bins_step2 bins spaced step2_step around each
    // slot's own float centre; the CFAR statistic divides by the input power the first step left behind
    bool step2 = false;
    float step2_step = 0.f;
    std::vector<float> center_f, input_power1;
    std::vector<uint8_t> has_code;
    std::vector<uint32_t> dwell;
    uint32_t last_n_slots = 0;
    EventProfiler prof;
    // sharded search (gnss_b200_acq_exchange_buffers): what this shard sends / what all shards sent
    DevBuf d_send, d_recv;
    uint32_t xchg_world = 0, xchg_chunk = 0;
    // record exchange over peer memory (gnss_b200_acq_peer_*)
    DevBuf d_peer, d_peer_ptrs;
    PeerLayout peer_layout{};
    std::vector<void*> peer_opened;  // cudaIpcOpenMemHandle mappings to close
    bool peer_connected = false;
    bool tail_deferred = false;  // peer-connected shard: the last dwell left its per-tile records for peer_exchange_async (shard_tail_kernel)
    bool graph_leaves_tail = false;  // tail_deferred as the captured sequence left it: a replay puts the handle into the same state
    // what the last dwell left on the device for gnss_b200_acq_finalize_gathered_async
    FinalizeArgs last_f{};
    // a captured dwell (gnss_b200_acq_capture_*)
    cudaGraph_t graph = nullptr;
    cudaGraphExec_t graph_exec = nullptr;
    bool capturing = false;
    uint64_t graph_kernels = 0;  // kernel nodes of the captured graph: what one replay launches
};

namespace
{
template <int LOAD, int STORE, bool INV>
int launch_stage(gnss_b200_acq* h, const StageArgs& a, uint32_t n_transforms, int cat)
{
    if (n_transforms == 0) return GNSS_B200_OK;
    const size_t smem = (size_t)a.st.tile_points * sizeof(float2);
    static std::atomic<size_t> configured[kMaxDeviceSlots];  // per kernel instantiation and device
    const int slot = device_slot();
    if (smem > 48 * 1024 && (slot < 0 || smem > configured[slot].load(std::memory_order_relaxed)))
        {
            cudaFuncAttributes fa;
            B200_CUDA(cudaFuncGetAttributes(&fa, stage_kernel<LOAD, STORE, INV>));
            const size_t max_dyn = 227 * 1024 - fa.sharedSizeBytes;  // static + dynamic <= 227 KiB per CTA on sm_100
            B200_REQUIRE(smem <= max_dyn, GNSS_B200_ERR_UNSUPPORTED, "FFT tile of %zu bytes exceeds shared memory", smem);
            B200_CUDA(cudaFuncSetAttribute(stage_kernel<LOAD, STORE, INV>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)max_dyn));
            if (slot >= 0) configured[slot].store(max_dyn, std::memory_order_relaxed);
        }
    const dim3 grid(n_transforms, a.st.tiles);
    h->prof.begin(cat, h->stream);
    stage_kernel<LOAD, STORE, INV><<<grid, a.st.threads, smem, h->stream>>>(a);
    h->prof.end(h->stream);
    h->launches++;
    B200_CUDA(cudaGetLastError());
    return GNSS_B200_OK;
}

StageArgs base_args(const gnss_b200_acq* h, int stage)
{
    StageArgs a;
    std::memset(&a, 0, sizeof(a));
    a.st = h->plan.stage[stage];
    a.W = h->d_W.as<float2>();
    a.N = h->N;
    a.nD_local = h->nD_local;
    a.eff = h->eff;
    a.mag_offset = h->mag_offset;
    a.consumed = h->consumed;
    if (h->plan.fast)
        {
            a.fmap_ra = h->plan.ra_lo;
            a.fmap_rb = h->plan.rb_lo;
            a.fmap_mlo = h->plan.ra_lo * h->plan.rb_lo;
        }
    return a;
}

// Forward transform of `n` zero-padded local codes (time domain in d_codes_td) into conj spectra at cbuf[slots[i]].
int run_code_forward(gnss_b200_acq* h, uint32_t n, const uint32_t* d_slots)
{
    if (h->plan.n_stages == 1)
        {
            StageArgs a = base_args(h, 0);
            a.codes_td = h->d_codes_td.as<float2>();
            a.code_len = h->code_len;
            a.code_off = h->code_off;
            a.code_wrap = h->padded ? h->cfg.fft_size : 0u;
            a.code_scale = h->padded ? (float)((double)h->cfg.fft_size / (double)h->N) : 1.0f;
            a.slots = d_slots;
            a.out_buf = h->d_cbuf.as<float2>();
            a.out_by_slot = 1;
            return launch_stage<LOAD_CODE, STORE_CONJ, false>(h, a, n, CAT_CODE);
        }
    // two stages through the scratch buffer, in chunks that fit it
    const uint32_t chunk = (uint32_t)std::min<size_t>(n, std::max<size_t>(1, kScratchBudgetBytes / ((size_t)h->N * 8)));
    B200_CUDA(h->d_scratch.reserve((size_t)chunk * h->N * 8));
    for (uint32_t y0 = 0; y0 < n; y0 += chunk)
        {
            const uint32_t cnt = std::min(chunk, n - y0);
            StageArgs a = base_args(h, 0);
            a.codes_td = h->d_codes_td.as<float2>();
            a.code_len = h->code_len;
            a.code_off = h->code_off;
            a.code_wrap = h->padded ? h->cfg.fft_size : 0u;
            a.code_scale = h->padded ? (float)((double)h->cfg.fft_size / (double)h->N) : 1.0f;
            a.y0 = y0;
            a.out_buf = h->d_scratch.as<float2>();
            a.out_y0 = y0;
            int rc = launch_stage<LOAD_CODE, STORE_BUF, false>(h, a, cnt, CAT_CODE);
            if (rc) return rc;
            StageArgs b = base_args(h, 1);
            b.in_buf = h->d_scratch.as<float2>();
            b.in_y0 = y0;
            b.y0 = y0;
            b.slots = d_slots;
            b.out_buf = h->d_cbuf.as<float2>();
            b.out_by_slot = 1;
            b.fmap_store = 1;
            rc = launch_stage<LOAD_BUF, STORE_CONJ, false>(h, b, cnt, CAT_CODE);
            if (rc) return rc;
        }
    return GNSS_B200_OK;
}

int ensure_group(gnss_b200_acq* h, int64_t key, uint32_t* ordinal)
{
    auto it = h->group_cache.find(key);
    if (it != h->group_cache.end())
        {
            *ordinal = it->second;
            return GNSS_B200_OK;
        }
    // never evicts: the caller (enqueue_dwell) makes room for all uncached keys of a call BEFORE resolving any of them, so an
    // ordinal handed out earlier in the same call cannot be overwritten
    B200_REQUIRE(h->group_cache.size() < kMaxCachedGroups, GNSS_B200_ERR_STATE, "Doppler-group cache full (%zu entries)", h->group_cache.size());
    const uint32_t ord = (uint32_t)h->group_cache.size();
    // host: one piece table per local Doppler bin (the reference's update_grid_doppler_wipeoffs, :298-305)
    std::vector<WipePiece> tab((size_t)h->nD_local * kMaxPieces);
    std::vector<uint32_t> cnt(h->nD_local);
    std::vector<WipePieceHost> pc;
    for (uint32_t kk = 0; kk < h->nD_local; kk++)
        {
            const uint32_t k = h->row_begin + kk;
            float inc;
            if (h->step2)
                {
                    // update_grid_doppler_wipeoffs_step2 (:308-315): key carries the bits of d_doppler_center_step_two
                    const uint32_t bits = (uint32_t)(key & 0xffffffffll);
                    float centre;
                    std::memcpy(&centre, &bits, 4);
                    const float doppler = (static_cast<float>(k) - static_cast<float>(std::floor(h->nD / 2.0))) * h->step2_step;
                    inc = wipeoff_phase_inc_f(centre + doppler, h->cfg.fs);
                }
            else
                {
                    const int32_t doppler = -h->cfg.doppler_max + (int32_t)(h->cfg.doppler_step * k);  // + center + bias = key
                    inc = wipeoff_phase_inc((int32_t)(key + doppler), h->cfg.fs);
                }
            build_wipeoff_pieces(inc, h->N, &pc);
            B200_REQUIRE(pc.size() <= (size_t)kMaxPieces, GNSS_B200_ERR_UNSUPPORTED, "carrier phase table needs %zu pieces (> %d)", pc.size(), kMaxPieces);
            cnt[kk] = (uint32_t)pc.size();
            for (size_t i = 0; i < pc.size(); i++) tab[(size_t)kk * kMaxPieces + i] = WipePiece{pc[i].start, pc[i].phi0, pc[i].step};
        }
    const size_t tb = sizeof(WipePiece) * tab.size(), cb = sizeof(uint32_t) * cnt.size();
    B200_CUDA(cudaStreamSynchronize(h->stream));
    B200_CUDA(cudaMemcpy(h->d_pieces.as<char>() + (size_t)ord * tb, tab.data(), tb, cudaMemcpyHostToDevice));
    B200_CUDA(cudaMemcpy(h->d_npieces.as<char>() + (size_t)ord * cb, cnt.data(), cb, cudaMemcpyHostToDevice));
    h->group_cache[key] = ord;
    *ordinal = ord;
    return GNSS_B200_OK;
}

// Everything of one dwell after the samples are on the device.
int enqueue_dwell(gnss_b200_acq* h, const void* d_samples, int is_cshort, uint32_t n_slots, const uint32_t* slots)
{
    B200_REQUIRE(n_slots > 0 && n_slots <= h->cfg.max_slots, GNSS_B200_ERR_INVALID, "dwell: n_slots=%u out of range", n_slots);
    // ---- per-call metadata: slots, Doppler groups, dwell counters ----
    CallMeta m;
    std::vector<int64_t> call_keys;
    for (uint32_t i = 0; i < n_slots; i++)
        {
            const uint32_t s = slots[i];
            B200_REQUIRE(s < h->cfg.max_slots, GNSS_B200_ERR_INVALID, "dwell: slot %u out of range", s);
            B200_REQUIRE(h->has_code[s], GNSS_B200_ERR_STATE, "dwell: slot %u has no local code (call gnss_b200_acq_set_local_code)", s);
            int64_t key = (int64_t)h->center[s] + (int64_t)h->bias[s];
            int32_t centre_word = h->center[s];
            if (h->step2)
                {
                    uint32_t bits;
                    std::memcpy(&bits, &h->center_f[s], 4);
                    key = (int64_t)bits | (1ll << 40);
                    centre_word = (int32_t)bits;
                }
            uint32_t g = 0;
            for (; g < call_keys.size(); g++)
                if (call_keys[g] == key) break;
            if (g == call_keys.size()) call_keys.push_back(key);
            m.slots.push_back(s);
            m.slot_group.push_back(g);
            m.center.push_back(centre_word);
            m.input_power1.push_back(h->step2 ? h->input_power1[s] : 0.f);
        }
    B200_REQUIRE(call_keys.size() <= kMaxCachedGroups, GNSS_B200_ERR_UNSUPPORTED, "dwell: more than %u distinct Doppler centres in one call", kMaxCachedGroups);
    {
        // room for every uncached key of this call, decided once up front: clearing in the middle of the call would hand the
        // ordinal (= device piece table) of a group resolved earlier in the call to a later one
        size_t n_new = 0;
        for (int64_t key : call_keys) n_new += h->group_cache.count(key) ? 0 : 1;
        if (h->group_cache.size() + n_new > kMaxCachedGroups) h->group_cache.clear();
    }
    for (int64_t key : call_keys)
        {
            uint32_t ord;
            int rc = ensure_group(h, key, &ord);
            if (rc) return rc;
            m.group_piece_base.push_back(ord * h->nD_local);
        }
    for (uint32_t i = 0; i < n_slots; i++)
        {
            // :638. Without a persistent grid nothing is accumulated, so every dwell is a first dwell (max_dwells = 1
            // in the reference resets the counter after each pass, :869-880).
            if (!h->cfg.keep_grid) h->dwell[slots[i]] = 0;
            m.dwell_after.push_back(++h->dwell[slots[i]]);
        }
    const uint32_t n_groups = (uint32_t)call_keys.size();

    const size_t meta_words = (size_t)n_slots * 5 + n_groups;
    B200_CUDA(h->h_meta.reserve(meta_words * 4));
    B200_CUDA(h->d_meta.reserve(meta_words * 4));
    uint32_t* hm = h->h_meta.as<uint32_t>();
    std::memcpy(hm, m.slots.data(), n_slots * 4);
    std::memcpy(hm + n_slots, m.slot_group.data(), n_slots * 4);
    std::memcpy(hm + 2 * n_slots, m.dwell_after.data(), n_slots * 4);
    std::memcpy(hm + 3 * n_slots, m.center.data(), n_slots * 4);
    std::memcpy(hm + 4 * n_slots, m.input_power1.data(), n_slots * 4);
    std::memcpy(hm + 5 * n_slots, m.group_piece_base.data(), n_groups * 4);
    B200_CUDA(cudaMemcpyAsync(h->d_meta.p, hm, meta_words * 4, cudaMemcpyHostToDevice, h->stream));
    const uint32_t* d_slots = h->d_meta.as<uint32_t>();
    const uint32_t* d_slot_group = d_slots + n_slots;
    const uint32_t* d_dwell = d_slots + 2 * n_slots;
    const int32_t* d_center = reinterpret_cast<const int32_t*>(d_slots + 3 * n_slots);
    const float* d_ip1 = reinterpret_cast<const float*>(d_slots + 4 * n_slots);
    const uint32_t* d_gpb = d_slots + 5 * n_slots;

    const uint32_t n_fwd = n_groups * h->nD_local;
    const uint32_t n_rows = n_slots * h->nD_local;
    // Row-resident inverse (acq_cluster.cuh): full rows, records only or persistent grid, plans with a cluster instantiation.
    // It writes one record per CTA of the cluster instead of one per column tile.
    const bool full_rows = h->mag_offset == 0 && h->eff == h->N;
    const uint32_t cluster_cs = (h->plan.fast && h->plan.n_stages == 2 && full_rows)
                                    ? cl::cluster_tiles(h->plan.ra_lo, h->plan.rb_lo, h->plan.ra_hi, h->plan.rb_hi)
                                    : 0u;
    const uint32_t tiles_plan = h->plan.stage[0].tiles;  // the inverse always ends with stage[0]
    const uint32_t tiles_last = cluster_cs ? cluster_cs : tiles_plan;
    B200_CUDA(h->d_xbuf.reserve((size_t)n_fwd * h->N * 8));
    if (h->plan.n_stages == 2) B200_CUDA(h->d_xtmp.reserve((size_t)n_fwd * h->N * 8));
    B200_CUDA(h->d_partials.reserve(sizeof(RowPartial) * (size_t)n_rows * std::max(tiles_last, tiles_plan)));
    B200_CUDA(h->d_results.reserve(sizeof(gnss_b200_acq_result) * n_slots));
    B200_CUDA(h->h_results.reserve(sizeof(gnss_b200_acq_result) * n_slots));
    B200_CUDA(h->d_bestk.reserve(sizeof(uint32_t) * n_slots));

    // ---- forward: one transform per (Doppler group, bin) ----
    static const bool fwd_generic = [] {
        const char* e = std::getenv("B200_ACQ_FWD_GENERIC");  // tuning aid: the generic multi-pass forward stages
        return e && e[0] == '1';
    }();
    if (h->plan.fast && !fwd_generic)
        {
            const FftPlan& pl = h->plan;
            FastFwdArgs fa;
            std::memset(&fa, 0, sizeof(fa));
            fa.W = h->d_W.as<float2>();
            fa.N = h->N;
            fa.M_lo = pl.ra_lo * pl.rb_lo;
            fa.M_hi = pl.ra_hi * pl.rb_hi;
            fa.ra_lo = pl.ra_lo;
            fa.rb_lo = pl.rb_lo;
            fa.ra_hi = pl.ra_hi;
            fa.rb_hi = pl.rb_hi;
            fa.samples = d_samples;
            fa.fmt = (uint32_t)is_cshort;
            fa.consumed = h->consumed;
            fa.pieces = h->d_pieces.as<WipePiece>();
            fa.n_pieces = h->d_npieces.as<uint32_t>();
            fa.group_piece_base = d_gpb;
            fa.nD_local = h->nD_local;
            fa.scratch = h->d_xtmp.as<float2>();
            fa.X = h->d_xbuf.as<float2>();
            fa.nb = pl.stage[1].blocks;
            fa.C = pl.stage[0].C;
            // A shard of a split search transforms only a few rows: narrower column tiles until the launch fills the machine
            // (10 rows of N = 32 000: 0.048 -> 0.034 ms for the two forward kernels; a full grid keeps the plan's tile)
            while (fa.C % 2 == 0 && fa.C > 2 && (uint64_t)n_fwd * (fa.M_lo / fa.C) < (uint64_t)kNumSMs) fa.C /= 2;
            const uint32_t pitch = (pl.ra_lo % 2 == 0) ? pl.ra_lo + 1 : pl.ra_lo;
            const size_t smem_lo = ((size_t)fa.nb * pitch * pl.rb_lo + fa.M_lo) * sizeof(float2);
            const size_t smem_hi = ((size_t)fa.M_hi * fa.C + fa.M_hi) * sizeof(float2);
            const dim3 block_hi(fa.C, std::max(1u, (uint32_t)kFastThreads / fa.C));
            const uint32_t pen = fa.nb * std::max(pl.ra_lo, pl.rb_lo);
            const uint32_t threads_lo = std::max<uint32_t>(64, std::min<uint32_t>(kFastThreads, (pen + 31) / 32 * 32));
            h->prof.begin(CAT_FWD, h->stream);
            cudaError_t e = launch_fwd_upper(pl.ra_hi, pl.rb_hi, dim3(n_fwd, fa.M_lo / fa.C), block_hi, smem_hi, h->stream, fa);
            h->prof.end(h->stream);
            h->launches++;
            B200_CUDA(e);
            h->prof.begin(CAT_FWD, h->stream);
            e = launch_fwd_lower(pl.ra_lo, pl.rb_lo, dim3(n_fwd, fa.M_hi / fa.nb), threads_lo, smem_lo, h->stream, fa);
            h->prof.end(h->stream);
            h->launches++;
            B200_CUDA(e);
        }
    else
    {
        StageArgs a = base_args(h, 0);
        a.samples = d_samples;
        a.is_cshort = is_cshort;
        a.pieces = h->d_pieces.as<WipePiece>();
        a.n_pieces = h->d_npieces.as<uint32_t>();
        a.group_piece_base = d_gpb;
        a.out_buf = h->plan.n_stages == 2 ? h->d_xtmp.as<float2>() : h->d_xbuf.as<float2>();
        int rc = launch_stage<LOAD_SAMPLES, STORE_BUF, false>(h, a, n_fwd, CAT_FWD);
        if (rc) return rc;
        if (h->plan.n_stages == 2)
            {
                StageArgs b = base_args(h, 1);
                b.in_buf = h->d_xtmp.as<float2>();
                b.out_buf = h->d_xbuf.as<float2>();  // not in place: the fast layout permutes across tiles
                b.fmap_store = 1;
                rc = launch_stage<LOAD_BUF, STORE_BUF, false>(h, b, n_fwd, CAT_FWD);
                if (rc) return rc;
            }
    }

    // ---- inverse: one row per (slot, bin) ----
    auto inverse = [&](uint32_t rows, const uint32_t* k_override, float* mag_base, uint32_t mag_is_rowbuf, RowPartial* partials) -> int {
        if (h->plan.n_stages == 1)
            {
                StageArgs a = base_args(h, 0);
                a.in_buf = h->d_xbuf.as<float2>();
                a.cbuf = h->d_cbuf.as<float2>();
                a.slots = d_slots;
                a.slot_group = d_slot_group;
                a.k_override = k_override;
                a.mag_base = mag_base;
                a.mag_is_rowbuf = mag_is_rowbuf;
                a.dwell_after = d_dwell;
                a.partials = partials;
                return launch_stage<LOAD_PRODUCT, STORE_MAG, true>(h, a, rows, CAT_INV_LAST);
            }
        // inter-stage scratch: all rows of the call in one launch pair when they fit the budget (measured on B200: the
        // HBM round trip of the intermediate costs less than the launch boundaries and tail waves of many small chunks)
        size_t budget = kScratchBudgetBytes;
        if (const char* e = std::getenv("B200_ACQ_SCRATCH_MB")) budget = (size_t)std::max(1, std::atoi(e)) << 20;  // tuning aid
        const uint32_t chunk = (uint32_t)std::min<size_t>(rows, std::max<size_t>(1, budget / ((size_t)h->N * 8)));
        const bool row_resident = cluster_cs && !k_override && (mag_base ? !mag_is_rowbuf : true);  // needs no HBM scratch at all
        if (!row_resident) B200_CUDA(h->d_scratch.reserve((size_t)chunk * h->N * 8));
        if (h->plan.fast)
            {
                const FftPlan& pl = h->plan;
                FastArgs fa;
                std::memset(&fa, 0, sizeof(fa));
                fa.W = h->d_W.as<float2>();
                fa.N = h->N;
                fa.M_lo = pl.ra_lo * pl.rb_lo;
                fa.M_hi = pl.ra_hi * pl.rb_hi;
                fa.ra_lo = pl.ra_lo;
                fa.rb_lo = pl.rb_lo;
                fa.ra_hi = pl.ra_hi;
                fa.rb_hi = pl.rb_hi;
                fa.slots = d_slots;
                fa.slot_group = d_slot_group;
                fa.k_override = k_override;
                fa.nD_local = h->nD_local;
                fa.X = h->d_xbuf.as<float2>();
                fa.Cs = h->d_cbuf.as<float2>();
                fa.scratch = h->d_scratch.as<float2>();
                fa.nb = pl.stage[1].blocks;
                fa.C = pl.stage[0].C;
                fa.mag_base = mag_base;
                fa.mag_is_rowbuf = mag_is_rowbuf;
                fa.eff = h->eff;
                fa.mag_offset = h->mag_offset;
                fa.dwell_after = d_dwell;
                fa.partials = partials;
                // persistent grid with full rows: the specialised last stage (old values prefetched, streaming stores)
                const bool grid_path = mag_base && !mag_is_rowbuf && h->mag_offset == 0 && h->eff == h->N;
                // 2: nothing but the per-tile records leaves the last stage (no grid kept, full rows)
                const int upper_mode = grid_path ? 1 : ((!mag_base && h->mag_offset == 0 && h->eff == h->N) ? 2 : 0);
                if (row_resident && (upper_mode == 1 || upper_mode == 2))
                    {
                        // one persistent launch for all rows: the inter-stage intermediate stays in (distributed) shared memory
                        fa.y0 = 0;
                        h->prof.begin(CAT_INV_LAST, h->stream);
                        cudaError_t e = cl::launch_inv_cluster(pl.ra_lo, pl.rb_lo, pl.ra_hi, pl.rb_hi, upper_mode, rows, h->stream, fa);
                        h->prof.end(h->stream);
                        h->launches++;
                        B200_CUDA(e);
                        return GNSS_B200_OK;
                    }
                const uint32_t pitch = (pl.ra_lo % 2 == 0) ? pl.ra_lo + 1 : pl.ra_lo;
                const size_t smem_lo = ((size_t)fa.nb * pitch * pl.rb_lo + fa.M_lo) * sizeof(float2);
                const size_t smem_hi = ((size_t)pl.rb_hi * upper_group_stride(pl.ra_hi, fa.C) + fa.M_hi) * sizeof(float2);
                const uint32_t rows_hi = std::max(1u, (uint32_t)kFastThreads / fa.C);
                const dim3 block_hi(fa.C, rows_hi);
                const uint32_t threads_lo = lower_threads(fa.nb, pl.rb_lo);
                // Optional (B200_ACQ_OVERLAP=1): two half-size scratch buffers and two streams, the last stage of chunk i
                // running while the first stage of chunk i+1 fills the other buffer.
                static const bool overlap = [] {
                    const char* e = std::getenv("B200_ACQ_OVERLAP");  // tuning aid; measured slower on B200, default off
                    return e && e[0] == '1';
                }();
                const cudaStream_t s2 = overlap ? h->stream2 : h->stream;
                const uint32_t half = overlap ? std::max(1u, chunk / 2) : chunk;
                float2* const sbase = h->d_scratch.as<float2>();
                B200_CUDA(cudaEventRecord(h->ev_fork, h->stream));
                if (overlap) B200_CUDA(cudaStreamWaitEvent(s2, h->ev_fork, 0));
                uint32_t it = 0;
                for (uint32_t y0 = 0; y0 < rows; y0 += half, it++)
                    {
                        const uint32_t cnt = std::min(half, rows - y0);
                        const uint32_t buf = overlap ? (it & 1) : 0;
                        fa.y0 = y0;
                        fa.scratch = sbase + (size_t)buf * half * h->N;
                        if (overlap && it >= 2) B200_CUDA(cudaStreamWaitEvent(h->stream, h->ev_up[buf], 0));  // buffer free again
                        h->prof.begin(CAT_INV_FIRST, h->stream);
                        cudaError_t e = launch_inv_lower(pl.ra_lo, pl.rb_lo, dim3(cnt, fa.M_hi / fa.nb), threads_lo, smem_lo, h->stream, fa);
                        h->prof.end(h->stream);
                        h->launches++;
                        B200_CUDA(e);
                        B200_CUDA(cudaEventRecord(h->ev_lo[buf], h->stream));
                        if (overlap) B200_CUDA(cudaStreamWaitEvent(s2, h->ev_lo[buf], 0));
                        h->prof.begin(CAT_INV_LAST, s2);
                        e = launch_inv_upper(pl.ra_hi, pl.rb_hi, upper_mode, dim3(cnt, fa.M_lo / fa.C), block_hi, smem_hi, s2, fa);
                        h->prof.end(s2);
                        h->launches++;
                        B200_CUDA(e);
                        B200_CUDA(cudaEventRecord(h->ev_up[buf], s2));
                    }
                // join: everything after this (decision kernels, result copy) is on the main stream
                if (overlap) B200_CUDA(cudaStreamWaitEvent(h->stream, h->ev_up[0], 0));
                if (overlap && it >= 2) B200_CUDA(cudaStreamWaitEvent(h->stream, h->ev_up[1], 0));
                return GNSS_B200_OK;
            }
        for (uint32_t y0 = 0; y0 < rows; y0 += chunk)
            {
                const uint32_t cnt = std::min(chunk, rows - y0);
                StageArgs a = base_args(h, 1);  // inverse order: lowest stage first
                a.in_buf = h->d_xbuf.as<float2>();
                a.cbuf = h->d_cbuf.as<float2>();
                a.slots = d_slots;
                a.slot_group = d_slot_group;
                a.k_override = k_override;
                a.y0 = y0;
                a.out_buf = h->d_scratch.as<float2>();
                a.out_y0 = y0;
                int rc = launch_stage<LOAD_PRODUCT, STORE_BUF, true>(h, a, cnt, CAT_INV_FIRST);
                if (rc) return rc;
                StageArgs b = base_args(h, 0);
                b.in_buf = h->d_scratch.as<float2>();
                b.in_y0 = y0;
                b.y0 = y0;
                b.slots = d_slots;
                b.slot_group = d_slot_group;
                b.k_override = k_override;
                b.mag_base = mag_base;
                b.mag_is_rowbuf = mag_is_rowbuf;
                b.dwell_after = d_dwell;
                b.partials = partials;
                rc = launch_stage<LOAD_BUF, STORE_MAG, true>(h, b, cnt, CAT_INV_LAST);
                if (rc) return rc;
            }
        return GNSS_B200_OK;
    };
    int rc = inverse(n_rows, nullptr, h->cfg.keep_grid ? h->d_grid.as<float>() : nullptr, 0, h->d_partials.as<RowPartial>());
    if (rc) return rc;

    // ---- decision ----
    // A peer-connected shard decides nothing on its own: its records go to every shard in gnss_b200_acq_peer_exchange_async, which
    // folds the per-tile records itself (shard_tail_kernel) - no merge / decision / pack launches here.
    const bool defer_tail = h->peer_connected && h->cfg.use_cfar && !h->step2 && !h->prof.on;
    FinalizeArgs f;
    f.partials = h->d_partials.as<RowPartial>();
    f.tiles = tiles_last;
    if (tiles_last > 8 && !defer_tail)
        {
            // one record per warp of the row-resident kernel: fold them per row first (one warp per row)
            B200_CUDA(h->d_partials2.reserve(sizeof(RowPartial) * (size_t)std::max<size_t>(n_rows, (size_t)n_slots * std::max(tiles_last, tiles_plan))));
            h->prof.begin(CAT_DECIDE, h->stream);
            merge_partials_kernel<<<(n_rows + 7) / 8, 256, 0, h->stream>>>(h->d_partials.as<RowPartial>(), tiles_last, n_rows, h->d_partials2.as<RowPartial>());
            h->prof.end(h->stream);
            h->launches++;
            B200_CUDA(cudaGetLastError());
            f.partials = h->d_partials2.as<RowPartial>();
            f.tiles = 1;
        }
    f.nD = h->nD;
    f.nD_local = h->nD_local;
    f.row_begin = h->row_begin;
    f.eff = h->eff;
    f.doppler_max = h->cfg.doppler_max;
    f.doppler_step = h->cfg.doppler_step;
    f.use_cfar = h->cfg.use_cfar;
    f.slots = d_slots;
    f.center = d_center;
    f.step2 = h->step2 ? 1u : 0u;
    f.step2_step = h->step2_step;
    f.input_power1 = d_ip1;
    f.dwell_after = d_dwell;
    f.rows = h->d_rows.as<gnss_b200_acq_row>();
    f.results = h->d_results.as<gnss_b200_acq_result>();
    f.best_k_local = h->d_bestk.as<uint32_t>();
    h->last_f = f;
    h->tail_deferred = defer_tail;
    if (defer_tail)
        {
            h->last_n_slots = n_slots;
            return GNSS_B200_OK;
        }
    h->prof.begin(CAT_DECIDE, h->stream);
    finalize_kernel<<<n_slots, 32, 0, h->stream>>>(f);
    h->prof.end(h->stream);
    h->launches++;
    B200_CUDA(cudaGetLastError());
    if (h->xchg_world)
        {
            // sharded search: this shard's rows of the call's slots, contiguous, ready to be sent
            pack_rows_kernel<<<n_slots, 64, 0, h->stream>>>(h->d_rows.as<gnss_b200_acq_row>(), d_slots, h->nD, h->row_begin, h->nD_local, h->xchg_chunk,
                h->d_send.as<gnss_b200_acq_row>());
            h->launches++;
            B200_CUDA(cudaGetLastError());
        }

    if (!h->cfg.use_cfar)
        {
            SecondPeakArgs sp;
            sp.nD_local = h->nD_local;
            sp.eff = h->eff;
            sp.N = h->cfg.fft_size;  // the window wraps at d_fft_size (:580-591), not at a padded transform length
            sp.samples_per_chip = h->cfg.samples_per_chip;
            sp.slots = d_slots;
            sp.best_k_local = h->d_bestk.as<uint32_t>();
            sp.results = h->d_results.as<gnss_b200_acq_result>();
            if (h->cfg.keep_grid)
                {
                    sp.mag_base = h->d_grid.as<float>();
                    sp.is_rowbuf = 0;
                }
            else
                {
                    // the grid was never materialised: recompute the winning row of every slot into a row buffer
                    B200_CUDA(h->d_rowbuf.reserve(sizeof(float) * (size_t)n_slots * h->eff));
                    B200_CUDA(h->d_partials2.reserve(sizeof(RowPartial) * (size_t)n_slots * std::max(tiles_last, tiles_plan)));
                    rc = inverse(n_slots, h->d_bestk.as<uint32_t>(), h->d_rowbuf.as<float>(), 1, h->d_partials2.as<RowPartial>());
                    if (rc) return rc;
                    sp.mag_base = h->d_rowbuf.as<float>();
                    sp.is_rowbuf = 1;
                }
            h->prof.begin(CAT_DECIDE, h->stream);
            second_peak_kernel<<<n_slots, 256, 0, h->stream>>>(sp);
            h->prof.end(h->stream);
            h->launches++;
            B200_CUDA(cudaGetLastError());
        }
    B200_CUDA(cudaMemcpyAsync(h->h_results.p, h->d_results.p, sizeof(gnss_b200_acq_result) * n_slots, cudaMemcpyDeviceToHost, h->stream));
    h->last_n_slots = n_slots;
    return GNSS_B200_OK;
}
}  // namespace

namespace
{
int create_impl(const gnss_b200_acq_cfg* cfg, uint32_t n_step2, float step2_hz, gnss_b200_acq** out);
}

extern "C"
{
    int gnss_b200_acq_create(const gnss_b200_acq_cfg* cfg, gnss_b200_acq** out) { return create_impl(cfg, 0, 0.f, out); }

    int gnss_b200_acq_create_step2(const gnss_b200_acq_cfg* cfg, uint32_t num_doppler_bins_step2, float doppler_step2, gnss_b200_acq** out)
    {
        B200_REQUIRE(num_doppler_bins_step2 >= 1 && num_doppler_bins_step2 <= 4096 && doppler_step2 > 0.f, GNSS_B200_ERR_INVALID,
            "acq_create_step2: second_nbins=%u second_doppler_step=%g", num_doppler_bins_step2, (double)doppler_step2);
        B200_REQUIRE(cfg && !(cfg->row_end > cfg->row_begin), GNSS_B200_ERR_INVALID, "acq_create_step2: the narrow grid is not row-sharded");
        return create_impl(cfg, num_doppler_bins_step2, doppler_step2, out);
    }

    int gnss_b200_acq_set_step2_center(gnss_b200_acq* h, uint32_t slot, float doppler_center_step_two_hz, float input_power_step_one)
    {
        B200_REQUIRE(h, GNSS_B200_ERR_INVALID, "acq_set_step2_center: null handle");
        B200_REQUIRE(h->step2, GNSS_B200_ERR_STATE, "acq_set_step2_center: not a second-step handle (gnss_b200_acq_create_step2)");
        B200_REQUIRE(slot < h->cfg.max_slots, GNSS_B200_ERR_INVALID, "acq_set_step2_center: slot %u out of range", slot);
        std::lock_guard<std::mutex> lk(h->mu);
        h->center_f[slot] = doppler_center_step_two_hz;
        h->input_power1[slot] = input_power_step_one;
        return GNSS_B200_OK;
    }
}

namespace
{
// Relative cost per transform point of a plan, measured on B200 (profiles/r02_padded_lengths.txt): the "fast" four-radix plans run
// at ~1.4e11 points/s, single-stage codelet plans at about half of that, generic two-stage plans and plans with a direct-DFT
// pass over a large prime (31 for the sample rates that are multiples of 1.023 MHz) at 0.36-0.42e11.
double plan_cost_per_point(const b200::FftPlan& plan)
{
    if (plan.fast) return 1.0;
    uint32_t rmax = 0;
    for (uint32_t r : plan.radices) rmax = std::max(rmax, r);
    if (rmax >= 31) return 3.5;
    return plan.n_stages == 2 ? 3.3 : 2.2;
}

// A supported length whose plan needs a direct-DFT pass can be searched faster in the padded form (create_impl) on a longer
// length that has a fast plan. B200_FFT_PAD=1/0 forces the choice (tuning aid, tests).
bool consider_padded(const b200::FftPlan& plan)
{
    if (const char* e = getenv("B200_FFT_PAD")) return e[0] == '1';
    for (uint32_t r : plan.radices)
        if (r >= 31) return true;
    return false;
}

// The transform the engine runs for a search of `fft_size` lags over `consumed` samples: the length's own plan, or the padded form.
//
// A length the mixed-radix kernels cannot take (a prime factor > 127; FFTW in the reference takes any d_fft_size,
// pcps_acquisition.cc:85-92,114-115) or take only slowly. The search needs the CYCLIC correlation of length n = fft_size of the
// (zero-padded) block x, nonzero on [0, consumed), with the zero-padded code c:
//     y[m] = sum_i x[i] conj(c[(i - m) mod n]),  0 <= m < n.
// i - m runs over [-(n - 1), consumed - 1]: a cyclic correlation of any length M >= n + consumed - 1 produces the same y when the
// samples are zero-padded to M and the code buffer holds c[0 .. consumed) at its start and c[1 .. n) at its END (negative i - m
// wrap onto that tail) - for every layout of set_local_code (:176-209): full (2n - 1), zero-padded double length (1.5n - 1), bit
// transition. The engine runs on the cheapest M = 2^a 3^b 5^c it has a plan for (points x cost per point of the plan), searches the
// lags the layout keeps, and scales the code by n / M so that magnitudes equal those of unnormalised length-n transforms.
bool choose_transform(uint32_t fft_size, uint32_t consumed, b200::FftPlan* plan, uint32_t* length, const char** why)
{
    using b200::FftPlan;
    using b200::make_fft_plan;
    const bool have_plan = make_fft_plan(fft_size, kMaxSingleStagePoints, plan, why);
    *length = fft_size;
    if (have_plan && !consider_padded(*plan)) return true;
    const bool forced = getenv("B200_FFT_PAD") != nullptr;
    uint32_t best_m = 0;
    double best_cost = have_plan && !forced ? 0.9 * (double)fft_size * plan_cost_per_point(*plan) : 1e300;
    const uint64_t m_min = (uint64_t)fft_size + consumed - 1, m_max = std::min<uint64_t>(4 * m_min, 1ull << 31);
    for (uint64_t p2 = 1; p2 <= m_max; p2 *= 2)
        for (uint64_t p3 = p2; p3 <= m_max; p3 *= 3)
            for (uint64_t p5 = p3; p5 <= m_max; p5 *= 5)
                if (p5 >= m_min)
                    {
                        FftPlan trial;
                        const char* w2 = nullptr;
                        if (!make_fft_plan((uint32_t)p5, kMaxSingleStagePoints, &trial, &w2)) continue;
                        const double cost = (double)p5 * plan_cost_per_point(trial);
                        if (cost < best_cost || (cost == best_cost && p5 < best_m))
                            {
                                best_cost = cost;
                                best_m = (uint32_t)p5;
                            }
                    }
    FftPlan padded_plan;
    if (best_m && make_fft_plan(best_m, kMaxSingleStagePoints, &padded_plan, why))
        {
            *plan = padded_plan;
            *length = best_m;
            return true;
        }
    return have_plan;
}

int create_impl(const gnss_b200_acq_cfg* cfg, uint32_t n_step2, float step2_hz, gnss_b200_acq** out)
    {
        B200_REQUIRE(cfg && out, GNSS_B200_ERR_INVALID, "acq_create: null argument");
        B200_REQUIRE(cfg->fft_size >= 2 && cfg->consumed_samples >= 1 && cfg->consumed_samples <= cfg->fft_size, GNSS_B200_ERR_INVALID,
            "acq_create: fft_size=%u consumed_samples=%u", cfg->fft_size, cfg->consumed_samples);
        B200_REQUIRE(cfg->doppler_step > 0 && cfg->doppler_max > 0 && cfg->fs > 0 && cfg->max_slots > 0, GNSS_B200_ERR_INVALID, "acq_create: bad Doppler grid / fs / max_slots");
        B200_REQUIRE(!cfg->bit_transition || cfg->fft_size % 2 == 0, GNSS_B200_ERR_INVALID, "acq_create: bit_transition needs an even fft_size");
        int rc = select_device(cfg->device);
        if (rc) return rc;
        auto* h = new gnss_b200_acq();
        h->cfg = *cfg;
        h->N = cfg->fft_size;
        h->consumed = cfg->consumed_samples;
        h->eff = cfg->bit_transition ? h->N / 2 : h->N;                                   // :622
        h->mag_offset = cfg->bit_transition ? h->eff : 0;                                 // :671
        h->nD = (uint32_t)std::ceil((double)(cfg->doppler_max - (-cfg->doppler_max)) / (double)cfg->doppler_step);  // :264
        if (n_step2)
            {
                h->step2 = true;  // d_num_doppler_bins_step2 bins of doppler_step2 (:73, acq_conf.h:51,62)
                h->step2_step = step2_hz;
                h->nD = n_step2;
            }
        h->center_f.assign(cfg->max_slots, 0.f);
        h->input_power1.assign(cfg->max_slots, 0.f);
        h->row_begin = 0;
        h->nD_local = h->nD;
        if (cfg->row_end > cfg->row_begin)
            {
                if (cfg->row_end > h->nD)
                    {
                        delete h;
                        return fail(GNSS_B200_ERR_INVALID, "acq_create: row range [%u,%u) exceeds %u Doppler bins", cfg->row_begin, cfg->row_end, h->nD);
                    }
                h->row_begin = cfg->row_begin;
                h->nD_local = cfg->row_end - cfg->row_begin;
            }
        // set_local_code zero-pad layout (:188-205)
        h->code_len = cfg->bit_transition ? h->N / 2 : h->consumed;
        h->code_off = h->N - h->code_len;
        if (!cfg->bit_transition && h->N == h->consumed) h->code_off = 0;
        const char* why = nullptr;
        uint32_t transform_len = 0;
        if (!choose_transform(cfg->fft_size, cfg->consumed_samples, &h->plan, &transform_len, &why))
            {
                delete h;
                return fail(GNSS_B200_ERR_UNSUPPORTED, "acq_create: fft_size %u: %s", cfg->fft_size, why ? why : "unsupported");
            }
        h->padded = transform_len != cfg->fft_size;
        h->N = transform_len;  // eff, mag_offset, code_len, code_off stay those of the fft_size-point layout
        h->center.assign(cfg->max_slots, 0);
        h->bias.assign(cfg->max_slots, 0);
        h->has_code.assign(cfg->max_slots, 0);
        h->dwell.assign(cfg->max_slots, 0);
        cudaError_t e = cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking);
        if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&h->stream2, cudaStreamNonBlocking);
        for (cudaEvent_t* ev : {&h->ev_lo[0], &h->ev_lo[1], &h->ev_up[0], &h->ev_up[1], &h->ev_fork})
            if (e == cudaSuccess) e = cudaEventCreateWithFlags(ev, cudaEventDisableTiming);
        // twiddle table exp(-2 pi i k / N), computed in double and rounded once
        std::vector<float2> W(h->N);
        for (uint32_t k = 0; k < h->N; k++)
            {
                const long double ang = -2.0L * 3.14159265358979323846264338327950288L * (long double)k / (long double)h->N;
                W[k] = make_float2((float)cosl(ang), (float)sinl(ang));
            }
        if (e == cudaSuccess) e = h->d_W.reserve((size_t)h->N * 8);
        if (e == cudaSuccess) e = cudaMemcpy(h->d_W.p, W.data(), (size_t)h->N * 8, cudaMemcpyHostToDevice);
        if (e == cudaSuccess) e = h->d_cbuf.reserve((size_t)cfg->max_slots * h->N * 8);
        if (e == cudaSuccess) e = h->d_samples.reserve((size_t)h->consumed * 8);
        if (e == cudaSuccess) e = h->h_samples.reserve((size_t)h->consumed * 8);
        if (e == cudaSuccess) e = h->d_rows.reserve(sizeof(gnss_b200_acq_row) * (size_t)cfg->max_slots * h->nD);
        if (e == cudaSuccess) e = cudaMemset(h->d_rows.p, 0, sizeof(gnss_b200_acq_row) * (size_t)cfg->max_slots * h->nD);
        if (e == cudaSuccess) e = h->d_pieces.reserve(sizeof(WipePiece) * (size_t)kMaxCachedGroups * h->nD_local * kMaxPieces);
        if (e == cudaSuccess) e = h->d_npieces.reserve(sizeof(uint32_t) * (size_t)kMaxCachedGroups * h->nD_local);
        if (e == cudaSuccess && cfg->keep_grid)
            {
                const size_t bytes = sizeof(float) * (size_t)cfg->max_slots * h->nD_local * h->eff;
                e = h->d_grid.reserve(bytes);
                if (e == cudaSuccess) e = cudaMemset(h->d_grid.p, 0, bytes);
            }
        if (e != cudaSuccess)
            {
                gnss_b200_acq_destroy(h);
                return fail(GNSS_B200_ERR_CUDA, "acq_create: %s", cudaGetErrorString(e));
            }
        *out = h;
        return GNSS_B200_OK;
    }
}  // namespace

extern "C"
{
    int gnss_b200_acq_destroy(gnss_b200_acq* h)
    {
        if (!h) return GNSS_B200_OK;
        cudaSetDevice(h->cfg.device);
        if (h->stream) cudaStreamSynchronize(h->stream);
        for (DevBuf* b : {&h->d_W, &h->d_cbuf, &h->d_xbuf, &h->d_xtmp, &h->d_samples, &h->d_codes_td, &h->d_scratch, &h->d_pieces, &h->d_npieces, &h->d_meta,
                 &h->d_partials, &h->d_partials2, &h->d_rows, &h->d_results, &h->d_bestk, &h->d_grid, &h->d_rowbuf})
            b->release();
        h->d_send.release();
        h->d_recv.release();
        for (void* m : h->peer_opened) cudaIpcCloseMemHandle(m);
        h->d_peer.release();
        h->d_peer_ptrs.release();
        if (h->graph_exec) cudaGraphExecDestroy(h->graph_exec);
        if (h->graph) cudaGraphDestroy(h->graph);
        h->h_meta.release();
        h->h_results.release();
        h->h_samples.release();
        h->prof.release();
        if (h->stream2) cudaStreamSynchronize(h->stream2);
        for (cudaEvent_t ev : {h->ev_lo[0], h->ev_lo[1], h->ev_up[0], h->ev_up[1], h->ev_fork})
            if (ev) cudaEventDestroy(ev);
        if (h->stream2) cudaStreamDestroy(h->stream2);
        if (h->stream) cudaStreamDestroy(h->stream);
        delete h;
        return GNSS_B200_OK;
    }

    uint32_t gnss_b200_acq_num_doppler_bins(const gnss_b200_acq* h) { return h ? h->nD : 0; }

    int gnss_b200_acq_set_local_codes(gnss_b200_acq* h, uint32_t n_slots, const uint32_t* slots, const float* codes_c64)
    {
        B200_REQUIRE(h && slots && codes_c64 && n_slots > 0, GNSS_B200_ERR_INVALID, "acq_set_local_codes: bad arguments");
        std::lock_guard<std::mutex> lk(h->mu);
        B200_CUDA(cudaSetDevice(h->cfg.device));
        for (uint32_t i = 0; i < n_slots; i++) B200_REQUIRE(slots[i] < h->cfg.max_slots, GNSS_B200_ERR_INVALID, "acq_set_local_codes: slot %u out of range", slots[i]);
        // the adapter's buffer holds consumed_samples items per code; the reference reads code_len of them (:192,:198,:203)
        const size_t in_stride = h->consumed;
        B200_CUDA(h->d_codes_td.reserve((size_t)n_slots * h->code_len * 8 + (size_t)n_slots * 4));
        B200_CUDA(cudaMemcpy2DAsync(h->d_codes_td.p, (size_t)h->code_len * 8, codes_c64, in_stride * 8, (size_t)h->code_len * 8, n_slots, cudaMemcpyHostToDevice, h->stream));
        uint32_t* d_slots = reinterpret_cast<uint32_t*>(h->d_codes_td.as<char>() + (size_t)n_slots * h->code_len * 8);
        B200_CUDA(cudaMemcpyAsync(d_slots, slots, (size_t)n_slots * 4, cudaMemcpyHostToDevice, h->stream));
        int rc = run_code_forward(h, n_slots, d_slots);
        if (rc) return rc;
        B200_CUDA(cudaStreamSynchronize(h->stream));
        for (uint32_t i = 0; i < n_slots; i++) h->has_code[slots[i]] = 1;
        return GNSS_B200_OK;
    }

    int gnss_b200_acq_set_local_code(gnss_b200_acq* h, uint32_t slot, const float* code_c64)
    {
        return gnss_b200_acq_set_local_codes(h, 1, &slot, code_c64);
    }

    int gnss_b200_acq_set_doppler_center(gnss_b200_acq* h, uint32_t slot, int32_t center_hz, int32_t bias_hz)
    {
        B200_REQUIRE(h && slot < h->cfg.max_slots, GNSS_B200_ERR_INVALID, "acq_set_doppler_center: bad slot");
        std::lock_guard<std::mutex> lk(h->mu);
        h->center[slot] = center_hz;
        h->bias[slot] = bias_hz;
        return GNSS_B200_OK;
    }

    int gnss_b200_acq_reset(gnss_b200_acq* h, uint32_t slot)
    {
        B200_REQUIRE(h && slot < h->cfg.max_slots, GNSS_B200_ERR_INVALID, "acq_reset: bad slot");
        std::lock_guard<std::mutex> lk(h->mu);
        B200_CUDA(cudaSetDevice(h->cfg.device));
        h->dwell[slot] = 0;
        if (h->cfg.keep_grid)
            {
                const size_t n = (size_t)h->nD_local * h->eff;
                B200_CUDA(cudaMemsetAsync(h->d_grid.as<float>() + (size_t)slot * n, 0, n * sizeof(float), h->stream));
            }
        return GNSS_B200_OK;
    }

    int gnss_b200_acq_dwell(gnss_b200_acq* h, const void* samples, int is_cshort, uint32_t n_slots, const uint32_t* slots, gnss_b200_acq_result* results)
    {
        B200_REQUIRE(h && samples && slots && results, GNSS_B200_ERR_INVALID, "acq_dwell: null argument");
        std::lock_guard<std::mutex> lk(h->mu);
        B200_CUDA(cudaSetDevice(h->cfg.device));
        B200_REQUIRE(is_cshort >= 0 && is_cshort <= 2, GNSS_B200_ERR_INVALID, "acq_dwell: unknown sample format %d", is_cshort);
        const size_t bytes = (size_t)h->consumed * (is_cshort == 2 ? 2 : (is_cshort == 1 ? 4 : 8));
        B200_CUDA(cudaMemcpyAsync(h->d_samples.p, samples, bytes, cudaMemcpyHostToDevice, h->stream));
        int rc = enqueue_dwell(h, h->d_samples.p, is_cshort, n_slots, slots);
        if (rc) return rc;
        B200_CUDA(cudaStreamSynchronize(h->stream));
        std::memcpy(results, h->h_results.p, sizeof(gnss_b200_acq_result) * n_slots);
        return GNSS_B200_OK;
    }

    int gnss_b200_acq_dwell_device_async(gnss_b200_acq* h, const void* d_samples, uint32_t n_slots, const uint32_t* slots)
    {
        B200_REQUIRE(h && d_samples && slots, GNSS_B200_ERR_INVALID, "acq_dwell_device_async: null argument");
        std::lock_guard<std::mutex> lk(h->mu);
        B200_CUDA(cudaSetDevice(h->cfg.device));
        return enqueue_dwell(h, d_samples, 0, n_slots, slots);
    }

    int gnss_b200_acq_sync(gnss_b200_acq* h, uint32_t n_slots, gnss_b200_acq_result* results)
    {
        B200_REQUIRE(h, GNSS_B200_ERR_INVALID, "acq_sync: null handle");
        std::lock_guard<std::mutex> lk(h->mu);
        B200_CUDA(cudaSetDevice(h->cfg.device));
        B200_CUDA(cudaStreamSynchronize(h->stream));
        if (results)
            {
                B200_REQUIRE(!h->tail_deferred, GNSS_B200_ERR_STATE, "acq_sync: a peer-connected shard decides in gnss_b200_acq_peer_exchange_async; call it after the dwell");
                B200_REQUIRE(n_slots == h->last_n_slots, GNSS_B200_ERR_INVALID, "acq_sync: n_slots differs from the last dwell");
                std::memcpy(results, h->h_results.p, sizeof(gnss_b200_acq_result) * n_slots);
            }
        return GNSS_B200_OK;
    }

    int gnss_b200_acq_read_grid(gnss_b200_acq* h, uint32_t slot, float* out)
    {
        B200_REQUIRE(h && out && slot < h->cfg.max_slots, GNSS_B200_ERR_INVALID, "acq_read_grid: bad arguments");
        B200_REQUIRE(h->cfg.keep_grid, GNSS_B200_ERR_STATE, "acq_read_grid: handle was created with keep_grid = 0");
        std::lock_guard<std::mutex> lk(h->mu);
        B200_CUDA(cudaSetDevice(h->cfg.device));
        const size_t n = (size_t)h->nD_local * h->eff;
        B200_CUDA(cudaStreamSynchronize(h->stream));
        B200_CUDA(cudaMemcpy(out, h->d_grid.as<float>() + (size_t)slot * n, n * sizeof(float), cudaMemcpyDeviceToHost));
        return GNSS_B200_OK;
    }

    int gnss_b200_acq_read_rows(gnss_b200_acq* h, uint32_t slot, gnss_b200_acq_row* rows)
    {
        B200_REQUIRE(h && rows && slot < h->cfg.max_slots, GNSS_B200_ERR_INVALID, "acq_read_rows: bad arguments");
        std::lock_guard<std::mutex> lk(h->mu);
        B200_CUDA(cudaSetDevice(h->cfg.device));
        B200_CUDA(cudaStreamSynchronize(h->stream));
        B200_CUDA(cudaMemcpy(rows, h->d_rows.as<gnss_b200_acq_row>() + (size_t)slot * h->nD, sizeof(gnss_b200_acq_row) * h->nD, cudaMemcpyDeviceToHost));
        return GNSS_B200_OK;
    }

    int gnss_b200_acq_finalize_rows(const gnss_b200_acq_row* rows, uint32_t n_doppler, uint32_t effective_size, int32_t doppler_max,
        uint32_t doppler_step, int32_t doppler_center, uint32_t dwell_count, gnss_b200_acq_result* out)
    {
        B200_REQUIRE(rows && out && n_doppler > 0 && effective_size > 0 && dwell_count > 0, GNSS_B200_ERR_INVALID, "acq_finalize_rows: bad arguments");
        float best = 0.f;
        uint32_t bk = 0, bi = 0;
        for (uint32_t k = 0; k < n_doppler; k++)
            {
                B200_REQUIRE(rows[k].valid, GNSS_B200_ERR_STATE, "acq_finalize_rows: row %u missing (merge all shards first)", k);
                if (rows[k].max_val > best)  // :523
                    {
                        best = rows[k].max_val;
                        bk = k;
                        bi = rows[k].max_idx;
                    }
            }
        const uint32_t opp = (bk + n_doppler / 2) % n_doppler;
        const float per = rows[opp].row_sum / (float)(int32_t)effective_size;
        out->input_power = (float)(((double)per / 2.0) / (double)dwell_count);
        out->peak = best;
        out->code_phase = bi;
        out->doppler_index = bk;
        out->doppler_hz = -doppler_max + doppler_center + (int32_t)doppler_step * (int32_t)bk;
        out->second_peak = 0.f;
        out->test_statistic = best / out->input_power;
        out->dwell_count = dwell_count;
        return GNSS_B200_OK;
    }

    // parity aids (pure host): the reference's binary32 running-sum wipe-off phase, and the FFT plan
    int gnss_b200_debug_wipeoff_phase(int32_t doppler_hz_with_bias, double fs, uint32_t n, float* phase_out, uint32_t* n_pieces)
    {
        B200_REQUIRE(phase_out && n > 0, GNSS_B200_ERR_INVALID, "debug_wipeoff_phase: bad arguments");
        std::vector<WipePieceHost> pc;
        build_wipeoff_pieces(wipeoff_phase_inc(doppler_hz_with_bias, fs), n, &pc);
        size_t cur = 0;
        for (uint32_t i = 0; i < n; i++)
            {
                while (cur + 1 < pc.size() && i >= pc[cur + 1].start) cur++;
                phase_out[i] = std::fmaf((float)(i - pc[cur].start), pc[cur].step, pc[cur].phi0);
            }
        if (n_pieces) *n_pieces = (uint32_t)pc.size();
        return GNSS_B200_OK;
    }

    int gnss_b200_debug_fft_plan(uint32_t n, uint32_t* n_stages, uint32_t* radices, uint32_t* n_radices, uint32_t* perm)
    {
        FftPlan plan;
        const char* why = nullptr;
        if (!make_fft_plan(n, kMaxSingleStagePoints, &plan, &why)) return fail(GNSS_B200_ERR_UNSUPPORTED, "fft plan %u: %s", n, why ? why : "?");
        if (n_stages) *n_stages = plan.n_stages;
        if (n_radices) *n_radices = (uint32_t)plan.radices.size();
        if (radices)
            for (size_t i = 0; i < plan.radices.size(); i++) radices[i] = plan.radices[i];
        if (perm)
            {
                const std::vector<uint32_t> p = fft_plan_permutation(plan);
                std::memcpy(perm, p.data(), sizeof(uint32_t) * n);
            }
        return GNSS_B200_OK;
    }

    int gnss_b200_debug_transform_length(uint32_t fft_size, uint32_t consumed_samples, uint32_t* length)
    {
        B200_REQUIRE(length && fft_size >= 2 && consumed_samples >= 1 && consumed_samples <= fft_size, GNSS_B200_ERR_INVALID, "debug_transform_length: bad arguments");
        FftPlan plan;
        const char* why = nullptr;
        if (!choose_transform(fft_size, consumed_samples, &plan, length, &why)) return fail(GNSS_B200_ERR_UNSUPPORTED, "fft_size %u: %s", fft_size, why ? why : "unsupported");
        return GNSS_B200_OK;
    }

    int gnss_b200_acq_profile(gnss_b200_acq* h, int enable, double* ms_by_kernel, uint64_t* launches_by_kernel, int reset)
    {
        B200_REQUIRE(h, GNSS_B200_ERR_INVALID, "acq_profile: null handle");
        std::lock_guard<std::mutex> lk(h->mu);
        B200_CUDA(cudaSetDevice(h->cfg.device));
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

    int gnss_b200_acq_exchange_buffers(gnss_b200_acq* h, uint32_t world, void** d_send, void** d_recv, uint64_t* bytes_per_shard, uint32_t* rows_per_shard)
    {
        B200_REQUIRE(h && world >= 1 && world <= 64, GNSS_B200_ERR_INVALID, "acq_exchange_buffers: bad arguments");
        B200_REQUIRE(!h->step2, GNSS_B200_ERR_STATE, "acq_exchange_buffers: the narrow second grid is not row-sharded");
        std::lock_guard<std::mutex> lk(h->mu);
        B200_CUDA(cudaSetDevice(h->cfg.device));
        const uint32_t chunk = (h->nD + world - 1) / world;
        B200_REQUIRE(h->nD_local <= chunk && h->row_begin % chunk == 0, GNSS_B200_ERR_INVALID,
            "acq_exchange_buffers: with %u shards a handle owns rows [r * %u, (r + 1) * %u), this one has [%u, %u)", world, chunk, chunk, h->row_begin,
            h->row_begin + h->nD_local);
        const size_t per = sizeof(gnss_b200_acq_row) * (size_t)h->cfg.max_slots * chunk;
        B200_CUDA(cudaStreamSynchronize(h->stream));
        B200_CUDA(h->d_send.reserve(per));
        B200_CUDA(h->d_recv.reserve(per * world));
        B200_CUDA(cudaMemset(h->d_recv.p, 0, per * world));
        h->xchg_world = world;
        h->xchg_chunk = chunk;
        if (d_send) *d_send = h->d_send.p;
        if (d_recv) *d_recv = h->d_recv.p;
        if (bytes_per_shard) *bytes_per_shard = per;
        if (rows_per_shard) *rows_per_shard = chunk;
        return GNSS_B200_OK;
    }

    int gnss_b200_acq_finalize_gathered_async(gnss_b200_acq* h, uint32_t n_slots)
    {
        B200_REQUIRE(h, GNSS_B200_ERR_INVALID, "acq_finalize_gathered_async: null handle");
        std::lock_guard<std::mutex> lk(h->mu);
        B200_CUDA(cudaSetDevice(h->cfg.device));
        B200_REQUIRE(h->xchg_world, GNSS_B200_ERR_STATE, "acq_finalize_gathered_async: call gnss_b200_acq_exchange_buffers first");
        B200_REQUIRE(n_slots == h->last_n_slots && n_slots > 0, GNSS_B200_ERR_INVALID, "acq_finalize_gathered_async: n_slots differs from the last dwell");
        B200_REQUIRE(h->cfg.use_cfar, GNSS_B200_ERR_UNSUPPORTED, "acq_finalize_gathered_async: only the CFAR statistic can be decided from row records");
        B200_REQUIRE(!h->tail_deferred, GNSS_B200_ERR_STATE, "acq_finalize_gathered_async: this shard is peer-connected; its records travel by gnss_b200_acq_peer_exchange_async");
        finalize_gathered_kernel<<<n_slots, 32, 0, h->stream>>>(h->last_f, h->d_recv.as<gnss_b200_acq_row>(), h->xchg_world, h->xchg_chunk, n_slots);
        h->launches++;
        B200_CUDA(cudaGetLastError());
        B200_CUDA(cudaMemcpyAsync(h->h_results.p, h->d_results.p, sizeof(gnss_b200_acq_result) * n_slots, cudaMemcpyDeviceToHost, h->stream));
        return GNSS_B200_OK;
    }

    int gnss_b200_acq_peer_setup(gnss_b200_acq* h, uint32_t rank, void* ipc_handle_out, void** region_out)
    {
        B200_REQUIRE(h, GNSS_B200_ERR_INVALID, "acq_peer_setup: null handle");
        std::lock_guard<std::mutex> lk(h->mu);
        B200_CUDA(cudaSetDevice(h->cfg.device));
        B200_REQUIRE(h->xchg_world, GNSS_B200_ERR_STATE, "acq_peer_setup: call gnss_b200_acq_exchange_buffers first");
        B200_REQUIRE(rank < h->xchg_world && h->row_begin == rank * h->xchg_chunk, GNSS_B200_ERR_INVALID, "acq_peer_setup: rank %u does not own rows from %u", rank, h->row_begin);
        static_assert(sizeof(cudaIpcMemHandle_t) == 64, "the C-ABI passes IPC handles as 64 opaque bytes");
        PeerLayout& L = h->peer_layout;
        L.world = h->xchg_world;
        L.rank = rank;
        L.flags_off = 256;
        L.bar_off = 256 + 2 * L.world * 128;
        L.recv_off = (L.bar_off + L.world * 128 + 255) & ~255u;
        L.per_shard_rows = h->cfg.max_slots * h->xchg_chunk;
        const size_t bytes = (size_t)L.recv_off + sizeof(gnss_b200_acq_row) * 2ull * L.world * L.per_shard_rows;
        B200_CUDA(cudaStreamSynchronize(h->stream));
        B200_CUDA(h->d_peer.reserve(bytes));
        B200_CUDA(cudaMemset(h->d_peer.p, 0, bytes));
        B200_CUDA(h->d_peer_ptrs.reserve(sizeof(void*) * L.world));
        h->peer_connected = false;
        if (ipc_handle_out)
            {
                cudaIpcMemHandle_t hd;
                B200_CUDA(cudaIpcGetMemHandle(&hd, h->d_peer.p));
                std::memcpy(ipc_handle_out, &hd, sizeof(hd));
            }
        if (region_out) *region_out = h->d_peer.p;
        return GNSS_B200_OK;
    }

    int gnss_b200_acq_peer_connect(gnss_b200_acq* h, const void* ipc_handles, void* const* regions)
    {
        B200_REQUIRE(h && (ipc_handles || regions), GNSS_B200_ERR_INVALID, "acq_peer_connect: null argument");
        std::lock_guard<std::mutex> lk(h->mu);
        B200_CUDA(cudaSetDevice(h->cfg.device));
        B200_REQUIRE(h->d_peer.p != nullptr, GNSS_B200_ERR_STATE, "acq_peer_connect: call gnss_b200_acq_peer_setup first");
        const PeerLayout& L = h->peer_layout;
        std::vector<void*> ptrs(L.world, nullptr);
        for (uint32_t r = 0; r < L.world; r++)
            {
                if (r == L.rank)
                    ptrs[r] = h->d_peer.p;
                else if (regions)
                    {
                        // same process: the region of another handle, possibly on another device of this process
                        cudaPointerAttributes at;
                        B200_CUDA(cudaPointerGetAttributes(&at, regions[r]));
                        if (at.device != h->cfg.device)
                            {
                                int can = 0;
                                B200_CUDA(cudaDeviceCanAccessPeer(&can, h->cfg.device, at.device));
                                B200_REQUIRE(can, GNSS_B200_ERR_UNSUPPORTED, "acq_peer_connect: device %d cannot access device %d", h->cfg.device, at.device);
                                const cudaError_t pe = cudaDeviceEnablePeerAccess(at.device, 0);
                                if (pe != cudaSuccess && pe != cudaErrorPeerAccessAlreadyEnabled) B200_CUDA(pe);
                                (void)cudaGetLastError();
                            }
                        ptrs[r] = regions[r];
                    }
                else
                    {
                        cudaIpcMemHandle_t hd;
                        std::memcpy(&hd, static_cast<const char*>(ipc_handles) + (size_t)r * sizeof(hd), sizeof(hd));
                        void* m = nullptr;
                        B200_CUDA(cudaIpcOpenMemHandle(&m, hd, cudaIpcMemLazyEnablePeerAccess));
                        h->peer_opened.push_back(m);
                        ptrs[r] = m;
                    }
            }
        B200_CUDA(cudaMemcpy(h->d_peer_ptrs.p, ptrs.data(), sizeof(void*) * L.world, cudaMemcpyHostToDevice));
        h->peer_connected = true;
        return GNSS_B200_OK;
    }

    int gnss_b200_acq_peer_exchange_async(gnss_b200_acq* h, uint32_t n_slots)
    {
        B200_REQUIRE(h, GNSS_B200_ERR_INVALID, "acq_peer_exchange_async: null handle");
        std::lock_guard<std::mutex> lk(h->mu);
        B200_CUDA(cudaSetDevice(h->cfg.device));
        B200_REQUIRE(h->peer_connected, GNSS_B200_ERR_STATE, "acq_peer_exchange_async: call gnss_b200_acq_peer_setup / _peer_connect first");
        B200_REQUIRE(n_slots == h->last_n_slots && n_slots > 0, GNSS_B200_ERR_INVALID, "acq_peer_exchange_async: n_slots differs from the last dwell");
        B200_REQUIRE(h->cfg.use_cfar, GNSS_B200_ERR_UNSUPPORTED, "acq_peer_exchange_async: only the CFAR statistic can be decided from row records");
        const PeerLayout& L = h->peer_layout;
        uint32_t* epoch = h->d_peer.as<uint32_t>();
        if (h->tail_deferred)
            {
                // two launches: records folded and delivered, then the decision (which also advances the epoch)
                shard_tail_kernel<<<n_slots, 256, 0, h->stream>>>(h->last_f, h->d_peer_ptrs.as<char*>(), L, epoch, h->xchg_chunk, epoch + 16);
                finalize_peer_kernel<<<n_slots, 32, 0, h->stream>>>(h->last_f, h->d_peer.as<char>(), L, epoch, h->xchg_chunk, n_slots, epoch + 32);
                h->launches += 2;
                h->tail_deferred = false;
            }
        else
            {
                // the dwell ran before the peers were connected (or under the per-kernel profiler): its packed records are in d_send
                peer_scatter_kernel<<<L.world, 256, 0, h->stream>>>(h->d_send.as<gnss_b200_acq_row>(), n_slots * h->xchg_chunk, h->d_peer_ptrs.as<char*>(), L, epoch);
                finalize_peer_kernel<<<n_slots, 32, 0, h->stream>>>(h->last_f, h->d_peer.as<char>(), L, epoch, h->xchg_chunk, n_slots, nullptr);
                peer_epoch_bump_kernel<<<1, 1, 0, h->stream>>>(epoch);
                h->launches += 3;
            }
        B200_CUDA(cudaGetLastError());
        B200_CUDA(cudaMemcpyAsync(h->h_results.p, h->d_results.p, sizeof(gnss_b200_acq_result) * n_slots, cudaMemcpyDeviceToHost, h->stream));
        return GNSS_B200_OK;
    }

    int gnss_b200_acq_peer_barrier_async(gnss_b200_acq* h)
    {
        B200_REQUIRE(h, GNSS_B200_ERR_INVALID, "acq_peer_barrier_async: null handle");
        std::lock_guard<std::mutex> lk(h->mu);
        B200_CUDA(cudaSetDevice(h->cfg.device));
        B200_REQUIRE(h->peer_connected, GNSS_B200_ERR_STATE, "acq_peer_barrier_async: call gnss_b200_acq_peer_setup / _peer_connect first");
        peer_barrier_kernel<<<1, 32, 0, h->stream>>>(h->d_peer_ptrs.as<char*>(), h->d_peer.as<char>(), h->peer_layout, h->d_peer.as<uint32_t>() + 48);
        h->launches++;
        B200_CUDA(cudaGetLastError());
        return GNSS_B200_OK;
    }

    int gnss_b200_acq_capture_begin(gnss_b200_acq* h)
    {
        B200_REQUIRE(h, GNSS_B200_ERR_INVALID, "acq_capture_begin: null handle");
        std::lock_guard<std::mutex> lk(h->mu);
        B200_CUDA(cudaSetDevice(h->cfg.device));
        B200_REQUIRE(!h->capturing, GNSS_B200_ERR_STATE, "acq_capture_begin: already capturing");
        B200_REQUIRE(!h->prof.on, GNSS_B200_ERR_STATE, "acq_capture_begin: switch the per-kernel profiler off first");
        B200_REQUIRE(!h->cfg.keep_grid, GNSS_B200_ERR_UNSUPPORTED, "acq_capture_begin: a persistent grid counts its dwells on the host; capture single-dwell searches only");
        B200_CUDA(cudaStreamSynchronize(h->stream));
        B200_CUDA(cudaStreamBeginCapture(h->stream, cudaStreamCaptureModeThreadLocal));
        h->capturing = true;
        return GNSS_B200_OK;
    }

    int gnss_b200_acq_capture_end(gnss_b200_acq* h)
    {
        B200_REQUIRE(h, GNSS_B200_ERR_INVALID, "acq_capture_end: null handle");
        std::lock_guard<std::mutex> lk(h->mu);
        B200_CUDA(cudaSetDevice(h->cfg.device));
        B200_REQUIRE(h->capturing, GNSS_B200_ERR_STATE, "acq_capture_end: no capture in progress");
        h->capturing = false;
        cudaGraph_t g = nullptr;
        B200_CUDA(cudaStreamEndCapture(h->stream, &g));
        if (h->graph_exec) cudaGraphExecDestroy(h->graph_exec);
        if (h->graph) cudaGraphDestroy(h->graph);
        h->graph = g;
        h->graph_exec = nullptr;
        h->graph_leaves_tail = h->tail_deferred;
        B200_CUDA(cudaGraphInstantiate(&h->graph_exec, h->graph, 0));
        size_t n_nodes = 0;
        B200_CUDA(cudaGraphGetNodes(h->graph, nullptr, &n_nodes));
        std::vector<cudaGraphNode_t> nodes(n_nodes);
        if (n_nodes) B200_CUDA(cudaGraphGetNodes(h->graph, nodes.data(), &n_nodes));
        h->graph_kernels = 0;
        for (cudaGraphNode_t nd : nodes)
            {
                cudaGraphNodeType ty;
                if (cudaGraphNodeGetType(nd, &ty) == cudaSuccess && ty == cudaGraphNodeTypeKernel) h->graph_kernels++;
            }
        return GNSS_B200_OK;
    }

    int gnss_b200_acq_replay_async(gnss_b200_acq* h)
    {
        B200_REQUIRE(h, GNSS_B200_ERR_INVALID, "acq_replay_async: null handle");
        std::lock_guard<std::mutex> lk(h->mu);
        B200_CUDA(cudaSetDevice(h->cfg.device));
        B200_REQUIRE(h->graph_exec, GNSS_B200_ERR_STATE, "acq_replay_async: nothing captured");
        B200_CUDA(cudaGraphLaunch(h->graph_exec, h->stream));
        h->tail_deferred = h->graph_leaves_tail;  // e.g. a captured dwell of a connected shard whose exchange is queued eagerly afterwards
        h->launches += h->graph_kernels;  // the kernels the replay runs (one graph launch on the host)
        return GNSS_B200_OK;
    }

    void* gnss_b200_acq_stream(gnss_b200_acq* h) { return h ? (void*)h->stream : nullptr; }
    uint64_t gnss_b200_acq_launch_count(const gnss_b200_acq* h) { return h ? h->launches : 0; }
    uint32_t gnss_b200_acq_transform_length(const gnss_b200_acq* h) { return h ? h->N : 0; }
}
