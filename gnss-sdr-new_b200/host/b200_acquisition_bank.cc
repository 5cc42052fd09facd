/*
 * b200_acquisition_bank.cc — see the header.
 * SPDX-License-Identifier: GPL-3.0-or-later
 */
#include "b200_acquisition_bank.h"
#include <algorithm>
#include <chrono>
#include <cstring>

namespace
{
std::mutex g_registry_mu;
std::vector<std::weak_ptr<B200_Acquisition_Bank>> g_registry;
std::vector<B200_Acquisition_Bank::Geometry> g_registry_geo;
uint32_t g_capacity = 64;
uint32_t g_window_us = 0;
uint32_t g_expected = 0;
}  // namespace


bool B200_Acquisition_Bank::Geometry::operator==(const Geometry& o) const
{
    return fft_size == o.fft_size && consumed_samples == o.consumed_samples && doppler_max == o.doppler_max && doppler_step == o.doppler_step &&
           fs == o.fs && samples_per_chip == o.samples_per_chip && bit_transition == o.bit_transition && use_cfar == o.use_cfar &&
           keep_grid == o.keep_grid && device == o.device && step2_bins == o.step2_bins && step2_hz == o.step2_hz && source_id == o.source_id && n_devices == o.n_devices;
}


void B200_Acquisition_Bank::set_default_capacity(uint32_t max_slots)
{
    std::lock_guard<std::mutex> lk(g_registry_mu);
    g_capacity = std::max(1U, max_slots);
}


void B200_Acquisition_Bank::set_coalescing(uint32_t window_us, uint32_t expected_callers)
{
    std::lock_guard<std::mutex> lk(g_registry_mu);
    g_window_us = window_us;
    g_expected = expected_callers;
}


std::shared_ptr<B200_Acquisition_Bank> B200_Acquisition_Bank::get(const Geometry& g, std::string* why)
{
    std::lock_guard<std::mutex> lk(g_registry_mu);
    for (size_t i = 0; i < g_registry.size(); i++)
        {
            if (g_registry_geo[i] == g)
                {
                    if (auto alive = g_registry[i].lock()) return alive;
                }
        }
    gnss_b200_acq_cfg cfg;
    std::memset(&cfg, 0, sizeof(cfg));
    cfg.fft_size = g.fft_size;
    cfg.consumed_samples = g.consumed_samples;
    cfg.bit_transition = g.bit_transition ? 1 : 0;
    cfg.doppler_max = g.doppler_max;
    cfg.doppler_step = g.doppler_step;
    cfg.fs = g.fs;
    cfg.samples_per_chip = g.samples_per_chip;
    cfg.use_cfar = g.use_cfar ? 1 : 0;
    cfg.keep_grid = g.keep_grid ? 1 : 0;
    cfg.max_slots = g_capacity;
    cfg.device = g.device;
    gnss_b200_acq* engine = nullptr;
    gnss_b200_acq_group* group = nullptr;
    if (g.n_devices > 1)
        {
            std::vector<int32_t> devices(g.n_devices);
            for (uint32_t i = 0; i < g.n_devices; i++) devices[i] = g.device + static_cast<int32_t>(i);
            if (gnss_b200_acq_group_create(&cfg, g.n_devices, devices.data(), &group) != GNSS_B200_OK)
                {
                    if (why != nullptr) *why = gnss_b200_last_error();
                    return nullptr;
                }
        }
    else if (gnss_b200_acq_create(&cfg, &engine) != GNSS_B200_OK)
        {
            if (why != nullptr) *why = gnss_b200_last_error();
            return nullptr;
        }
    gnss_b200_acq* engine2 = nullptr;
    if (g.step2_bins != 0 && gnss_b200_acq_create_step2(&cfg, g.step2_bins, g.step2_hz, &engine2) != GNSS_B200_OK)
        {
            if (why != nullptr) *why = gnss_b200_last_error();
            if (engine != nullptr) gnss_b200_acq_destroy(engine);
            if (group != nullptr) gnss_b200_acq_group_destroy(group);
            return nullptr;
        }
    std::shared_ptr<B200_Acquisition_Bank> bank(new B200_Acquisition_Bank(g));
    bank->d_engine = engine;
    bank->d_group = group;
    bank->d_engine2 = engine2;
    bank->d_num_doppler_bins = group != nullptr ? gnss_b200_acq_group_num_doppler_bins(group) : gnss_b200_acq_num_doppler_bins(engine);
    bank->d_used.assign(g_capacity, 0);
    g_registry.emplace_back(bank);
    g_registry_geo.push_back(g);
    return bank;
}


B200_Acquisition_Bank::~B200_Acquisition_Bank()
{
    if (d_engine != nullptr) gnss_b200_acq_destroy(d_engine);
    if (d_group != nullptr) gnss_b200_acq_group_destroy(d_group);
    if (d_engine2 != nullptr) gnss_b200_acq_destroy(d_engine2);
}


int B200_Acquisition_Bank::acquire_slot()
{
    std::lock_guard<std::mutex> lk(d_mu);
    for (size_t i = 0; i < d_used.size(); i++)
        {
            if (d_used[i] == 0)
                {
                    d_used[i] = 1;
                    return static_cast<int>(i);
                }
        }
    d_error = "B200_Acquisition_Bank: no free slot (raise set_default_capacity before creating blocks)";
    return -1;
}


void B200_Acquisition_Bank::release_slot(int slot)
{
    std::lock_guard<std::mutex> lk(d_mu);
    if (slot >= 0 && static_cast<size_t>(slot) < d_used.size()) d_used[slot] = 0;
}


bool B200_Acquisition_Bank::set_local_code(int slot, const std::complex<float>* code)
{
    if (slot < 0 || code == nullptr) return false;
    const auto uslot = static_cast<uint32_t>(slot);
    int rc = d_group != nullptr ? gnss_b200_acq_group_set_local_codes(d_group, 1, &uslot, reinterpret_cast<const float*>(code))
                                : gnss_b200_acq_set_local_code(d_engine, uslot, reinterpret_cast<const float*>(code));
    if (rc == GNSS_B200_OK && d_engine2 != nullptr)  // d_fft_codes serves both grids in the reference (:727, :723 of the step-two loop)
        {
            rc = gnss_b200_acq_set_local_code(d_engine2, static_cast<uint32_t>(slot), reinterpret_cast<const float*>(code));
        }
    if (rc != GNSS_B200_OK)
        {
            std::lock_guard<std::mutex> lk(d_mu);
            d_error = gnss_b200_last_error();
        }
    return rc == GNSS_B200_OK;
}


bool B200_Acquisition_Bank::set_doppler_center(int slot, int32_t center_hz, int32_t bias_hz)
{
    if (slot < 0) return false;
    if (d_group != nullptr) return gnss_b200_acq_group_set_doppler_center(d_group, static_cast<uint32_t>(slot), center_hz, bias_hz) == GNSS_B200_OK;
    return gnss_b200_acq_set_doppler_center(d_engine, static_cast<uint32_t>(slot), center_hz, bias_hz) == GNSS_B200_OK;
}


bool B200_Acquisition_Bank::reset(int slot)
{
    if (slot < 0) return false;
    const bool ok2 = d_engine2 == nullptr || gnss_b200_acq_reset(d_engine2, static_cast<uint32_t>(slot)) == GNSS_B200_OK;
    const int rc = d_group != nullptr ? gnss_b200_acq_group_reset(d_group, static_cast<uint32_t>(slot)) : gnss_b200_acq_reset(d_engine, static_cast<uint32_t>(slot));
    return rc == GNSS_B200_OK && ok2;
}


bool B200_Acquisition_Bank::read_grid(int slot, float* out)
{
    if (slot < 0 || out == nullptr) return false;
    if (d_group != nullptr) return gnss_b200_acq_group_read_grid(d_group, static_cast<uint32_t>(slot), out) == GNSS_B200_OK;
    return gnss_b200_acq_read_grid(d_engine, static_cast<uint32_t>(slot), out) == GNSS_B200_OK;
}


bool B200_Acquisition_Bank::read_grid_step2(int slot, float* out)
{
    if (slot < 0 || out == nullptr || d_engine2 == nullptr) return false;
    return gnss_b200_acq_read_grid(d_engine2, static_cast<uint32_t>(slot), out) == GNSS_B200_OK;
}


// Guard behind the source id: two requests with the same stamp are served from ONE buffer, so their blocks must really hold
// the same samples. Every block has its own copy (d_data_buffer), so pointers differ; the head and the tail of the dwell are
// compared instead (a few cache lines; noise makes two different streams differ in the first sample already).
bool B200_Acquisition_Bank::same_block(const void* a, const void* b, bool is_cshort) const
{
    if (a == b) return true;
    const size_t item = is_cshort ? 4 : 8;
    const size_t bytes = static_cast<size_t>(d_geo.consumed_samples) * item;
    const size_t probe = std::min<size_t>(bytes, 256);
    return std::memcmp(a, b, probe) == 0 &&
           std::memcmp(static_cast<const char*>(a) + bytes - probe, static_cast<const char*>(b) + bytes - probe, probe) == 0;
}


// Leader with d_mu held: issue one engine call per distinct (sample stamp, item type, step, sample block) among the pending dwells.
void B200_Acquisition_Bank::drain_locked(std::unique_lock<std::mutex>& lk)
{
    while (!d_pending.empty())
        {
            std::vector<Request*> group;
            const uint64_t stamp = d_pending.front()->stamp;
            const bool is_cshort = d_pending.front()->is_cshort;
            const bool step2 = d_pending.front()->step2;
            for (auto it = d_pending.begin(); it != d_pending.end();)
                {
                    // a second-step call holds at most 64 distinct Doppler centres (include/gnss_b200.h)
                    if ((*it)->stamp == stamp && (*it)->is_cshort == is_cshort && (*it)->step2 == step2 && (!step2 || group.size() < 64) &&
                        same_block(d_pending.front()->samples, (*it)->samples, is_cshort))
                        {
                            group.push_back(*it);
                            it = d_pending.erase(it);
                        }
                    else
                        {
                            ++it;
                        }
                }
            d_engine_calls++;
            d_dwells += group.size();
            lk.unlock();

            // equal stamps on one stream mean equal sample blocks: any member's buffer serves the whole group
            const auto n = static_cast<uint32_t>(group.size());
            std::vector<uint32_t> slots(n);
            std::vector<gnss_b200_acq_result> results(n);
            for (uint32_t i = 0; i < n; i++) slots[i] = group[i]->slot;
            int rc = GNSS_B200_OK;
            if (step2)
                {
                    for (uint32_t i = 0; i < n && rc == GNSS_B200_OK; i++)
                        {
                            rc = d_engine2 == nullptr ? GNSS_B200_ERR_STATE : gnss_b200_acq_set_step2_center(d_engine2, slots[i], group[i]->center2, group[i]->input_power1);
                        }
                }
            if (rc == GNSS_B200_OK)
                {
                    if (!step2 && d_group != nullptr)
                        rc = gnss_b200_acq_group_dwell(d_group, group[0]->samples, is_cshort ? 1 : 0, n, slots.data(), results.data());
                    else
                        rc = gnss_b200_acq_dwell(step2 ? d_engine2 : d_engine, group[0]->samples, is_cshort ? 1 : 0, n, slots.data(), results.data());
                }

            lk.lock();
            if (rc != GNSS_B200_OK) d_error = gnss_b200_last_error();
            for (uint32_t i = 0; i < n; i++)
                {
                    if (rc == GNSS_B200_OK) *group[i]->result = results[i];
                    group[i]->ok = (rc == GNSS_B200_OK);
                    group[i]->done = true;
                }
            d_cv_done.notify_all();
        }
}


bool B200_Acquisition_Bank::dwell(int slot, const void* samples, bool is_cshort, uint64_t sample_stamp, gnss_b200_acq_result* result)
{
    if (slot < 0 || samples == nullptr || result == nullptr) return false;
    Request req;
    req.slot = static_cast<uint32_t>(slot);
    req.samples = samples;
    req.is_cshort = is_cshort;
    req.stamp = sample_stamp;
    req.result = result;
    return submit(req);
}


bool B200_Acquisition_Bank::dwell_step2(int slot, const void* samples, bool is_cshort, uint64_t sample_stamp, float doppler_center_step_two,
    float input_power_step_one, gnss_b200_acq_result* result)
{
    if (slot < 0 || samples == nullptr || result == nullptr) return false;
    if (d_engine2 == nullptr)
        {
            std::lock_guard<std::mutex> lk(d_mu);
            d_error = "B200_Acquisition_Bank: this bank was created without a second-step grid";
            return false;
        }
    Request req;
    req.slot = static_cast<uint32_t>(slot);
    req.samples = samples;
    req.is_cshort = is_cshort;
    req.stamp = sample_stamp;
    req.result = result;
    req.step2 = true;
    req.center2 = doppler_center_step_two;
    req.input_power1 = input_power_step_one;
    return submit(req);
}


bool B200_Acquisition_Bank::submit(Request& req)
{
    uint32_t window_us;
    uint32_t expected;
    {
        std::lock_guard<std::mutex> lk(g_registry_mu);
        window_us = g_window_us;
        expected = g_expected;
    }
    std::unique_lock<std::mutex> lk(d_mu);
    d_pending.push_back(&req);
    if (d_leader_active)
        {
            if (expected != 0 && d_pending.size() >= expected) d_cv_leader.notify_one();
            d_cv_done.wait(lk, [&req] { return req.done; });
            return req.ok;
        }
    d_leader_active = true;
    if (window_us != 0 && (expected == 0 || d_pending.size() < expected))
        {
            const auto deadline = std::chrono::steady_clock::now() + std::chrono::microseconds(window_us);
            d_cv_leader.wait_until(lk, deadline, [this, expected] { return expected != 0 && d_pending.size() >= expected; });
        }
    drain_locked(lk);
    d_leader_active = false;
    return req.ok;
}


std::string B200_Acquisition_Bank::last_error() const
{
    std::lock_guard<std::mutex> lk(d_mu);
    return d_error;
}


uint64_t B200_Acquisition_Bank::engine_calls() const
{
    std::lock_guard<std::mutex> lk(d_mu);
    return d_engine_calls;
}


uint64_t B200_Acquisition_Bank::dwells_served() const
{
    std::lock_guard<std::mutex> lk(d_mu);
    return d_dwells;
}
