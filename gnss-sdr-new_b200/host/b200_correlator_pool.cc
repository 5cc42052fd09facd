/*
 * b200_correlator_pool.cc — see the header. Leader/follower request combining in front of the C-ABI engine.
 * SPDX-License-Identifier: GPL-3.0-or-later
 */
#include "b200_correlator_pool.h"
#include <algorithm>
#include <chrono>

B200_Correlator_Pool& B200_Correlator_Pool::instance()
{
    static B200_Correlator_Pool pool;
    return pool;
}


B200_Correlator_Pool::~B200_Correlator_Pool()
{
    if (d_engine != nullptr)
        {
            gnss_b200_mc_destroy(d_engine);
            d_engine = nullptr;
        }
}


void B200_Correlator_Pool::configure(int device, uint32_t max_channels, uint32_t max_samples, uint32_t max_taps)
{
    std::lock_guard<std::mutex> lk(d_mu);
    if (d_engine != nullptr) return;  // limits are fixed once the engine exists
    d_device = device;
    d_max_channels = std::max(1U, max_channels);
    d_max_samples = std::max(1U, max_samples);
    d_max_taps = std::max(1U, max_taps);
}


void B200_Correlator_Pool::set_coalescing(uint32_t window_us, uint32_t expected_callers)
{
    std::lock_guard<std::mutex> lk(d_mu);
    d_window_us = window_us;
    d_expected = expected_callers;
}


bool B200_Correlator_Pool::ensure_engine_locked()
{
    if (d_engine != nullptr) return true;
    const int rc = gnss_b200_mc_create(d_max_channels, d_max_samples, d_max_taps, d_device, &d_engine);
    if (rc != GNSS_B200_OK)
        {
            // No CPU fallback: the caller sees `false`, exactly like a failed allocation would surface.
            d_error = gnss_b200_last_error();
            d_engine = nullptr;
            return false;
        }
    d_used.assign(d_max_channels, 0);
    d_is_complex.assign(d_max_channels, 0);
    return true;
}


int B200_Correlator_Pool::acquire_channel()
{
    std::lock_guard<std::mutex> lk(d_mu);
    if (!ensure_engine_locked()) return -1;
    for (uint32_t i = 0; i < d_max_channels; i++)
        {
            if (d_used[i] == 0)
                {
                    d_used[i] = 1;
                    return static_cast<int>(i);
                }
        }
    d_error = "B200_Correlator_Pool: all channels in use (raise max_channels with configure())";
    return -1;
}


void B200_Correlator_Pool::release_channel(int ch)
{
    std::lock_guard<std::mutex> lk(d_mu);
    if (ch >= 0 && static_cast<uint32_t>(ch) < d_used.size()) d_used[ch] = 0;
}


bool B200_Correlator_Pool::set_code(int ch, int code_length, const float* code, bool is_complex)
{
    std::lock_guard<std::mutex> lk(d_mu);
    if (d_engine == nullptr || ch < 0 || code_length <= 0 || code == nullptr) return false;
    const int rc = gnss_b200_mc_set_code(d_engine, static_cast<uint32_t>(ch), static_cast<uint32_t>(code_length), code, is_complex ? 1 : 0);
    if (rc != GNSS_B200_OK) d_error = gnss_b200_last_error();
    if (rc == GNSS_B200_OK && static_cast<uint32_t>(ch) < d_is_complex.size()) d_is_complex[ch] = is_complex ? 1 : 0;
    return rc == GNSS_B200_OK;
}


bool B200_Correlator_Pool::correlate_many(uint32_t n, const uint32_t* ch, const gnss_b200_mc_args* args, const float* shifts_chips, uint32_t n_taps,
    uint32_t flags, std::complex<float>* out)
{
    gnss_b200_mc* engine;
    {
        std::lock_guard<std::mutex> lk(d_mu);
        engine = d_engine;
        d_batches++;
        d_calls += n;
    }
    if (engine == nullptr) return false;
    const int rc = gnss_b200_mc_correlate_batch(engine, n, ch, args, shifts_chips, n_taps, flags, reinterpret_cast<float*>(out));
    if (rc != GNSS_B200_OK)
        {
            std::lock_guard<std::mutex> lk(d_mu);
            d_error = gnss_b200_last_error();
        }
    return rc == GNSS_B200_OK;
}


// Called by the leader with d_mu held: drains d_pending in groups of equal (n_taps, flags, code type, block length). Real- and
// complex-code correlators (B200_Multicorrelator_Real_Codes / B200_Multicorrelator) share this pool and the engine refuses a
// batch that mixes them; channels of different signals (different vector_length) go in separate launches so that a long
// block does not take the short ones off the engine's fast path.
void B200_Correlator_Pool::run_batch_locked(std::unique_lock<std::mutex>& lk)
{
    while (!d_pending.empty())
        {
            std::vector<Request*> group;
            const uint32_t n_taps = d_pending.front()->n_taps;
            const uint32_t flags = d_pending.front()->flags;
            const bool is_complex = d_pending.front()->is_complex;
            const uint32_t n_samples = d_pending.front()->args.n_samples;
            for (auto it = d_pending.begin(); it != d_pending.end();)
                {
                    if ((*it)->n_taps == n_taps && (*it)->flags == flags && (*it)->is_complex == is_complex && (*it)->args.n_samples == n_samples)
                        {
                            group.push_back(*it);
                            it = d_pending.erase(it);
                        }
                    else
                        {
                            ++it;
                        }
                }
            gnss_b200_mc* engine = d_engine;
            d_batches++;
            d_calls += group.size();
            lk.unlock();

            const auto n = static_cast<uint32_t>(group.size());
            std::vector<uint32_t> ch(n);
            std::vector<gnss_b200_mc_args> args(n);
            std::vector<float> shifts(static_cast<size_t>(n) * n_taps);
            std::vector<std::complex<float>> out(static_cast<size_t>(n) * n_taps);
            for (uint32_t i = 0; i < n; i++)
                {
                    ch[i] = group[i]->ch;
                    args[i] = group[i]->args;
                    std::copy(group[i]->shifts, group[i]->shifts + n_taps, shifts.begin() + static_cast<size_t>(i) * n_taps);
                }
            const int rc = engine == nullptr ? GNSS_B200_ERR_STATE
                                             : gnss_b200_mc_correlate_batch(engine, n, ch.data(), args.data(), shifts.data(), n_taps, flags,
                                                   reinterpret_cast<float*>(out.data()));
            if (rc == GNSS_B200_OK)
                {
                    for (uint32_t i = 0; i < n; i++)
                        {
                            std::copy(out.begin() + static_cast<size_t>(i) * n_taps, out.begin() + static_cast<size_t>(i + 1) * n_taps, group[i]->out);
                        }
                }

            lk.lock();
            if (rc != GNSS_B200_OK) d_error = gnss_b200_last_error();
            for (Request* r : group)
                {
                    r->ok = (rc == GNSS_B200_OK);
                    r->done = true;
                }
            d_cv_done.notify_all();
        }
}


bool B200_Correlator_Pool::correlate(int ch, const gnss_b200_mc_args& args, const float* shifts_chips, uint32_t n_taps, uint32_t flags,
    std::complex<float>* out)
{
    if (ch < 0 || shifts_chips == nullptr || out == nullptr || n_taps == 0) return false;
    Request req;
    req.ch = static_cast<uint32_t>(ch);
    req.args = args;
    req.shifts = shifts_chips;
    req.n_taps = n_taps;
    req.flags = flags;
    req.out = out;

    std::unique_lock<std::mutex> lk(d_mu);
    if (d_engine == nullptr) return false;
    req.is_complex = req.ch < d_is_complex.size() && d_is_complex[req.ch] != 0;
    d_pending.push_back(&req);
    if (d_leader_active)
        {
            // follower: wake the leader if the batch is complete, then wait for the result
            if (d_expected != 0 && d_pending.size() >= d_expected) d_cv_leader.notify_one();
            d_cv_done.wait(lk, [&req] { return req.done; });
            return req.ok;
        }
    d_leader_active = true;
    if (d_window_us != 0 && (d_expected == 0 || d_pending.size() < d_expected))
        {
            const auto deadline = std::chrono::steady_clock::now() + std::chrono::microseconds(d_window_us);
            d_cv_leader.wait_until(lk, deadline, [this] { return d_expected != 0 && d_pending.size() >= d_expected; });
        }
    run_batch_locked(lk);
    d_leader_active = false;
    // requests that arrived while the batch ran elect the next leader among themselves: wake them all
    d_cv_done.notify_all();
    return req.ok;
}


std::string B200_Correlator_Pool::last_error() const
{
    std::lock_guard<std::mutex> lk(d_mu);
    return d_error;
}


uint64_t B200_Correlator_Pool::launches() const
{
    std::lock_guard<std::mutex> lk(d_mu);
    return d_engine != nullptr ? gnss_b200_mc_launch_count(d_engine) : 0;
}


uint64_t B200_Correlator_Pool::batches() const
{
    std::lock_guard<std::mutex> lk(d_mu);
    return d_batches;
}


uint64_t B200_Correlator_Pool::coalesced_calls() const
{
    std::lock_guard<std::mutex> lk(d_mu);
    return d_calls;
}
