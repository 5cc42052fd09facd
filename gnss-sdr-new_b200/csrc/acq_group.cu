// One PCPS search spread over several GPUs of one process: the Doppler rows of every slot are split over one engine per device
// (Doppler-major: no forward transform is computed twice, every device carries the same share of every PRN), the sample block
// is broadcast to the devices, each device reduces its rows, the per-row records travel device-to-device into the root's
// receive buffer (cudaMemcpyPeerAsync over NVLink) and the root decides every slot on the device.
// C-ABI: gnss_b200_acq_group_* (include/gnss_b200.h). Built on the single-device engine's own C-ABI.
#include "common.cuh"
#include <algorithm>
#include <vector>

using namespace b200;

struct gnss_b200_acq_group
{
    std::mutex mu;
    gnss_b200_acq_cfg cfg{};
    uint32_t nD = 0, chunk = 0, eff = 0;
    struct Shard
    {
        int device = 0;
        gnss_b200_acq* h = nullptr;
        cudaStream_t stream = nullptr;
        void* d_samples = nullptr;   // this device's copy of the sample block
        void* d_send = nullptr;
        void* d_recv = nullptr;      // used on the root only
        uint32_t row_begin = 0, row_end = 0;
        cudaEvent_t ev_rows = nullptr;  // "my records are in the root's receive buffer"
    };
    std::vector<Shard> shards;       // shards[0] is the root
    uint64_t bytes_per_shard = 0;
    cudaEvent_t ev_samples = nullptr;  // "the sample block is on the root"
    void* h_samples = nullptr;         // pinned staging for pageable caller buffers
    size_t sample_bytes_cap = 0;
};

extern "C"
{
    int gnss_b200_acq_group_destroy(gnss_b200_acq_group* g)
    {
        if (!g) return GNSS_B200_OK;
        for (auto& s : g->shards)
            {
                if (!s.h) continue;  // never created (e.g. no such device): nothing of it exists
                cudaSetDevice(s.device);
                gnss_b200_acq_sync(s.h, 0, nullptr);
                if (s.d_samples) cudaFree(s.d_samples);
                if (s.ev_rows) cudaEventDestroy(s.ev_rows);
                gnss_b200_acq_destroy(s.h);
            }
        if (g->ev_samples)
            {
                cudaSetDevice(g->shards[0].device);
                cudaEventDestroy(g->ev_samples);
            }
        if (g->h_samples) cudaFreeHost(g->h_samples);
        delete g;
        (void)cudaGetLastError();  // a failed clean-up call must not surface in somebody else's launch check
        return GNSS_B200_OK;
    }

    int gnss_b200_acq_group_create(const gnss_b200_acq_cfg* cfg, uint32_t n_devices, const int32_t* devices, gnss_b200_acq_group** out)
    {
        B200_REQUIRE(cfg && out && devices && n_devices >= 1 && n_devices <= 64, GNSS_B200_ERR_INVALID, "acq_group_create: bad arguments");
        B200_REQUIRE(cfg->use_cfar, GNSS_B200_ERR_UNSUPPORTED,
            "acq_group_create: only the CFAR statistic is decided from row records; the first-vs-second-peak test needs the winning row's "
            "magnitudes on one device (use a single-device handle)");
        B200_REQUIRE(!(cfg->row_end > cfg->row_begin), GNSS_B200_ERR_INVALID, "acq_group_create: the group does the row split itself");
        auto* g = new gnss_b200_acq_group();
        g->cfg = *cfg;
        g->nD = (uint32_t)std::ceil((double)(cfg->doppler_max - (-cfg->doppler_max)) / (double)cfg->doppler_step);
        g->eff = cfg->bit_transition ? cfg->fft_size / 2 : cfg->fft_size;
        const uint32_t world = std::min<uint32_t>(n_devices, g->nD);  // never more shards than rows
        g->chunk = (g->nD + world - 1) / world;
        for (uint32_t r = 0; r < world && r * g->chunk < g->nD; r++)
            {
                gnss_b200_acq_group::Shard s;
                s.device = devices[r];
                s.row_begin = r * g->chunk;
                s.row_end = std::min(g->nD, (r + 1) * g->chunk);
                g->shards.push_back(s);
            }
        const uint32_t n_shards = (uint32_t)g->shards.size();
        int rc = GNSS_B200_OK;
        for (uint32_t r = 0; r < n_shards && rc == GNSS_B200_OK; r++)
            {
                auto& s = g->shards[r];
                gnss_b200_acq_cfg c = *cfg;
                c.device = s.device;
                c.row_begin = s.row_begin;
                c.row_end = s.row_end;
                rc = gnss_b200_acq_create(&c, &s.h);
                if (rc) break;
                s.stream = (cudaStream_t)gnss_b200_acq_stream(s.h);
                uint32_t rows = 0;
                rc = gnss_b200_acq_exchange_buffers(s.h, n_shards, &s.d_send, &s.d_recv, &g->bytes_per_shard, &rows);
                if (rc) break;
                cudaError_t e = cudaSetDevice(s.device);
                if (e == cudaSuccess) e = cudaMalloc(&s.d_samples, (size_t)cfg->consumed_samples * 8);
                if (e == cudaSuccess) e = cudaEventCreateWithFlags(&s.ev_rows, cudaEventDisableTiming);
                // the root must be able to take the others' peer copies (and they its sample block)
                if (e == cudaSuccess && r > 0 && s.device != g->shards[0].device)
                    {
                        int can = 0;
                        cudaDeviceCanAccessPeer(&can, s.device, g->shards[0].device);
                        if (can)
                            {
                                cudaError_t pe = cudaDeviceEnablePeerAccess(g->shards[0].device, 0);
                                if (pe != cudaSuccess && pe != cudaErrorPeerAccessAlreadyEnabled) e = pe;
                                (void)cudaGetLastError();
                            }
                    }
                if (e != cudaSuccess) rc = fail(GNSS_B200_ERR_CUDA, "acq_group_create: device %d: %s", s.device, cudaGetErrorString(e));
            }
        if (rc == GNSS_B200_OK)
            {
                cudaError_t e = cudaSetDevice(g->shards[0].device);
                if (e == cudaSuccess) e = cudaEventCreateWithFlags(&g->ev_samples, cudaEventDisableTiming);
                if (e != cudaSuccess) rc = fail(GNSS_B200_ERR_CUDA, "acq_group_create: %s", cudaGetErrorString(e));
            }
        if (rc)
            {
                // keep the message of the failure, not of the clean-up
                char msg[512];
                std::snprintf(msg, sizeof(msg), "%s", last_error_buf());
                gnss_b200_acq_group_destroy(g);
                std::snprintf(last_error_buf(), 512, "%s", msg);
                return rc;
            }
        *out = g;
        return GNSS_B200_OK;
    }

    uint32_t gnss_b200_acq_group_num_doppler_bins(const gnss_b200_acq_group* g) { return g ? g->nD : 0; }
    uint32_t gnss_b200_acq_group_num_shards(const gnss_b200_acq_group* g) { return g ? (uint32_t)g->shards.size() : 0; }

    int gnss_b200_acq_group_set_local_codes(gnss_b200_acq_group* g, uint32_t n_slots, const uint32_t* slots, const float* codes_c64)
    {
        B200_REQUIRE(g, GNSS_B200_ERR_INVALID, "acq_group_set_local_codes: null handle");
        std::lock_guard<std::mutex> lk(g->mu);
        for (auto& s : g->shards)
            {
                const int rc = gnss_b200_acq_set_local_codes(s.h, n_slots, slots, codes_c64);
                if (rc) return rc;
            }
        return GNSS_B200_OK;
    }

    int gnss_b200_acq_group_set_doppler_center(gnss_b200_acq_group* g, uint32_t slot, int32_t center_hz, int32_t bias_hz)
    {
        B200_REQUIRE(g, GNSS_B200_ERR_INVALID, "acq_group_set_doppler_center: null handle");
        std::lock_guard<std::mutex> lk(g->mu);
        for (auto& s : g->shards)
            {
                const int rc = gnss_b200_acq_set_doppler_center(s.h, slot, center_hz, bias_hz);
                if (rc) return rc;
            }
        return GNSS_B200_OK;
    }

    int gnss_b200_acq_group_reset(gnss_b200_acq_group* g, uint32_t slot)
    {
        B200_REQUIRE(g, GNSS_B200_ERR_INVALID, "acq_group_reset: null handle");
        std::lock_guard<std::mutex> lk(g->mu);
        for (auto& s : g->shards)
            {
                const int rc = gnss_b200_acq_reset(s.h, slot);
                if (rc) return rc;
            }
        return GNSS_B200_OK;
    }

    int gnss_b200_acq_group_dwell(gnss_b200_acq_group* g, const void* samples, int fmt, uint32_t n_slots, const uint32_t* slots, gnss_b200_acq_result* results)
    {
        B200_REQUIRE(g && samples && slots && results && n_slots > 0, GNSS_B200_ERR_INVALID, "acq_group_dwell: bad arguments");
        B200_REQUIRE(fmt == GNSS_B200_SAMPLES_GR_COMPLEX, GNSS_B200_ERR_UNSUPPORTED, "acq_group_dwell: gr_complex sample blocks only");
        std::lock_guard<std::mutex> lk(g->mu);
        auto& root = g->shards[0];
        const size_t bytes = (size_t)g->cfg.consumed_samples * 8;
        // ---- broadcast: host -> root, root -> every other device (each on its own stream, behind the root's copy) ----
        B200_CUDA(cudaSetDevice(root.device));
        B200_CUDA(cudaMemcpyAsync(root.d_samples, samples, bytes, cudaMemcpyHostToDevice, root.stream));
        B200_CUDA(cudaEventRecord(g->ev_samples, root.stream));
        for (size_t r = 1; r < g->shards.size(); r++)
            {
                auto& s = g->shards[r];
                B200_CUDA(cudaSetDevice(s.device));
                B200_CUDA(cudaStreamWaitEvent(s.stream, g->ev_samples, 0));
                B200_CUDA(cudaMemcpyPeerAsync(s.d_samples, s.device, root.d_samples, root.device, bytes, s.stream));
            }
        // ---- every device searches its rows; its records go straight into the root's receive buffer ----
        const size_t rec_bytes = sizeof(gnss_b200_acq_row) * (size_t)n_slots * g->chunk;
        for (size_t r = 0; r < g->shards.size(); r++)
            {
                auto& s = g->shards[r];
                int rc = gnss_b200_acq_dwell_device_async(s.h, s.d_samples, n_slots, slots);
                if (rc) return rc;
                B200_CUDA(cudaSetDevice(s.device));
                char* dst = static_cast<char*>(root.d_recv) + r * rec_bytes;
                B200_CUDA(cudaMemcpyPeerAsync(dst, root.device, s.d_send, s.device, rec_bytes, s.stream));
                B200_CUDA(cudaEventRecord(s.ev_rows, s.stream));
            }
        // ---- root: wait for everybody's records, decide, fetch ----
        B200_CUDA(cudaSetDevice(root.device));
        for (size_t r = 1; r < g->shards.size(); r++) B200_CUDA(cudaStreamWaitEvent(root.stream, g->shards[r].ev_rows, 0));
        int rc = gnss_b200_acq_finalize_gathered_async(root.h, n_slots);
        if (rc) return rc;
        rc = gnss_b200_acq_sync(root.h, n_slots, results);
        if (rc) return rc;
        // the other devices' streams are idle by now (the root waited on their events); keep their handles' state tidy
        for (size_t r = 1; r < g->shards.size(); r++)
            {
                rc = gnss_b200_acq_sync(g->shards[r].h, 0, nullptr);
                if (rc) return rc;
            }
        return GNSS_B200_OK;
    }

    int gnss_b200_acq_group_read_grid(gnss_b200_acq_group* g, uint32_t slot, float* out)
    {
        B200_REQUIRE(g && out, GNSS_B200_ERR_INVALID, "acq_group_read_grid: bad arguments");
        std::lock_guard<std::mutex> lk(g->mu);
        for (auto& s : g->shards)
            {
                const int rc = gnss_b200_acq_read_grid(s.h, slot, out + (size_t)s.row_begin * g->eff);
                if (rc) return rc;
            }
        return GNSS_B200_OK;
    }

    uint64_t gnss_b200_acq_group_launch_count(const gnss_b200_acq_group* g)
    {
        uint64_t n = 0;
        if (g)
            for (auto& s : g->shards) n += gnss_b200_acq_launch_count(s.h);
        return n;
    }
}
