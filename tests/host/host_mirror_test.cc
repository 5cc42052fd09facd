/*
 * host_mirror_test.cc — drives the host C++ mirror (gnss-sdr-new_b200/host) the way the reference drives its own
 * classes: blocks are set up in Channel's call order (channel.cc:49-93), fed through general_work from one thread per
 * block (GNU Radio's thread-per-block scheduler), and report through Gnss_Synchro + ChannelFsm; correlator objects are
 * called once per code period from one thread per channel (dll_pll_veml_tracking.cc:1064-1089).
 * Inputs come from a binary file written by tests/test_host_mirror_gpu.py; outputs go to a binary file that the
 * Python test compares with the oracle. This program never computes expected values itself.
 */
#include "b200_acquisition_bank.h"
#include "b200_correlator_pool.h"
#include "b200_mat_writer.h"
#include "b200_multicorrelator.h"
#include "b200_symbol_sync.h"
#include "b200_tracking_dump.h"
#include "b200_multicorrelator_real_codes.h"
#include "pcps_acquisition_b200.h"
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstdio>
#include <cstring>
#include <fstream>
#include <iostream>
#include <thread>
#include <vector>

namespace
{
struct Reader
{
    std::ifstream f;
    explicit Reader(const char* path) : f(path, std::ios::binary) {}
    template <typename T>
    T get()
    {
        T v{};
        f.read(reinterpret_cast<char*>(&v), sizeof(T));
        if (!f) throw std::runtime_error("short input file");
        return v;
    }
    template <typename T>
    std::vector<T> vec(size_t n)
    {
        std::vector<T> v(n);
        f.read(reinterpret_cast<char*>(v.data()), sizeof(T) * n);
        if (!f) throw std::runtime_error("short input file");
        return v;
    }
};

struct Writer
{
    std::ofstream f;
    explicit Writer(const char* path) : f(path, std::ios::binary) {}
    template <typename T>
    void put(const T& v)
    {
        f.write(reinterpret_cast<const char*>(&v), sizeof(T));
    }
};

class RecordingFsm : public ChannelFsm
{
public:
    std::atomic<int> valid{0}, failed{0};
    bool Event_valid_acquisition() override
    {
        valid++;
        return true;
    }
    bool Event_failed_acquisition_repeat() override
    {
        failed++;
        return true;
    }
    bool Event_failed_acquisition_no_repeat() override
    {
        failed++;
        return true;
    }
};

int run_acq(const char* in_path, const char* out_path)
{
    Reader r(in_path);
    Acq_Conf conf;
    conf.fs_in = r.get<int64_t>();
    conf.sampled_ms = r.get<uint32_t>();
    conf.ms_per_code = r.get<uint32_t>();
    conf.samples_per_chip = r.get<uint32_t>();
    conf.doppler_max = r.get<int32_t>();
    const auto doppler_step = r.get<uint32_t>();
    conf.doppler_step = static_cast<float>(doppler_step);
    conf.pfa = r.get<float>();
    const auto threshold = r.get<float>();
    conf.max_dwells = r.get<uint32_t>();
    conf.bit_transition_flag = r.get<uint32_t>() != 0;
    conf.use_CFAR_algorithm_flag = r.get<uint32_t>() != 0;
    const bool cshort = r.get<uint32_t>() != 0;
    conf.blocking = r.get<uint32_t>() != 0;
    const auto n_blocks = r.get<uint32_t>();
    const auto n_items = r.get<uint32_t>();
    const auto chunk = r.get<uint32_t>();
    const auto use_fsm = r.get<uint32_t>();
    const auto coalesce_us = r.get<uint32_t>();
    conf.resampled_fs = conf.fs_in;                                                                    // acq_conf.cc:55
    conf.samples_per_ms = static_cast<float>(conf.resampled_fs) * 0.001F;                              // acq_conf.cc:113-119
    conf.samples_per_code = conf.samples_per_ms * static_cast<float>(conf.ms_per_code);
    conf.it_size = cshort ? 2 * sizeof(int16_t) : sizeof(std::complex<float>);
    conf.item_type = cshort ? "cshort" : "gr_complex";
    const auto consumed = static_cast<uint32_t>(conf.sampled_ms * conf.samples_per_ms * (conf.bit_transition_flag ? 2.0 : 1.0));
    std::vector<std::vector<std::complex<float>>> codes(n_blocks);
    for (auto& c : codes) c = r.vec<std::complex<float>>(consumed);
    std::vector<std::complex<float>> stream_fc;
    std::vector<int16_t> stream_sc;
    if (cshort)
        stream_sc = r.vec<int16_t>(static_cast<size_t>(n_items) * 2);
    else
        stream_fc = r.vec<std::complex<float>>(n_items);
    uint32_t want_events = 1;  // decisions to wait for per block (the fork's repeat mode keeps reporting)
    if (r.f.peek() != std::ifstream::traits_type::eof())  // optional trailer: make_two_steps (acq_conf.cc:84-90)
        {
            conf.make_2_steps = r.get<uint32_t>() != 0;
            conf.num_doppler_bins_step2 = r.get<uint32_t>();
            conf.doppler_step2 = r.get<float>();
            conf.pfa2 = r.get<float>();
            conf.dump = r.get<uint32_t>() != 0;
            conf.dump_channel = r.get<uint32_t>();
            const auto len = r.get<uint32_t>();
            const auto name = r.vec<char>(len);
            conf.dump_filename.assign(name.begin(), name.end());
            if (r.f.peek() != std::ifstream::traits_type::eof())
                {
                    conf.make_repeat_steps = r.get<uint32_t>() != 0;
                    want_events = r.get<uint32_t>();
                }
            if (r.f.peek() != std::ifstream::traits_type::eof())  // the acquisition resampler (acq_conf.cc:93-111): the header's fs is the
                {                                                   // rate of the stream in this file, fs_in becomes the undecimated rate
                    conf.use_automatic_resampler = r.get<uint32_t>() != 0;
                    conf.resampler_ratio = r.get<float>();
                    conf.resampler_latency_samples = r.get<uint32_t>();
                    conf.fs_in = r.get<int64_t>();
                }
        }

    B200_Acquisition_Bank::set_default_capacity(std::max(1U, n_blocks));
    B200_Acquisition_Bank::set_coalescing(coalesce_us, coalesce_us ? n_blocks : 0);

    std::vector<pcps_acquisition_b200_sptr> blocks(n_blocks);
    std::vector<Gnss_Synchro> synchro(n_blocks);
    std::vector<std::shared_ptr<RecordingFsm>> fsm(n_blocks);
    std::vector<std::atomic<long>> events(n_blocks);
    std::vector<std::atomic<long>> consumed_items(n_blocks);
    std::vector<std::atomic<uint32_t>> n_positive(n_blocks), n_negative(n_blocks);
    for (uint32_t b = 0; b < n_blocks; b++)
        {
            events[b] = 0;
            consumed_items[b] = 0;
            n_positive[b] = 0;
            n_negative[b] = 0;
            synchro[b].System = 'G';
            std::strcpy(synchro[b].Signal, "1C");
            synchro[b].PRN = b + 1;
            blocks[b] = pcps_make_acquisition_b200(conf);
            blocks[b]->consume_each = [&consumed_items, b](int n) { consumed_items[b] += n; };
            blocks[b]->publish_event = [&events, &n_positive, &n_negative, b](long e) {
                events[b] = e;
                if (e == 1) n_positive[b]++;
                if (e == 2) n_negative[b]++;
            };
            // Channel's order (channel.cc:49-93)
            blocks[b]->set_channel(b);
            if (use_fsm != 0)
                {
                    fsm[b] = std::make_shared<RecordingFsm>();
                    blocks[b]->set_channel_fsm(fsm[b]);
                }
            blocks[b]->set_gnss_synchro(&synchro[b]);
            blocks[b]->set_doppler_max(static_cast<uint32_t>(conf.doppler_max));
            blocks[b]->set_doppler_step(doppler_step);
            blocks[b]->set_threshold(threshold);
            blocks[b]->init();
            blocks[b]->start();  // pfa > 0 overrides the threshold, as in the reference
            blocks[b]->set_local_code(codes[b].data());
            blocks[b]->set_state(1);  // gps_l1_ca_pcps_acquisition_test.cc:341-343
        }

    // one scheduler thread per block; each walks the whole stream in `chunk`-item calls
    auto feed = [&](uint32_t b) {
        size_t pos = 0;
        int idle_calls = 0;
        const auto t_end = std::chrono::steady_clock::now() + std::chrono::seconds(60);
        while (std::chrono::steady_clock::now() < t_end)
            {
                const size_t avail = std::min<size_t>(chunk, n_items - pos);
                std::vector<int> ninput{static_cast<int>(avail)};
                std::vector<const void*> in{cshort ? static_cast<const void*>(stream_sc.data() + pos * 2) : static_cast<const void*>(stream_fc.data() + pos)};
                std::vector<void*> out;
                const long before = consumed_items[b];
                blocks[b]->general_work(0, ninput, in, out);
                pos += static_cast<size_t>(consumed_items[b] - before);
                const uint32_t decisions = fsm[b] ? static_cast<uint32_t>(fsm[b]->valid + fsm[b]->failed) : n_positive[b] + n_negative[b];
                if (decisions >= want_events) break;  // decision(s) taken
                if (avail == 0)
                    {
                        // stream exhausted: a non-blocking worker may still be running
                        if (++idle_calls > 5000) break;
                        std::this_thread::sleep_for(std::chrono::milliseconds(1));
                    }
            }
    };
    std::vector<std::thread> threads;
    for (uint32_t b = 0; b < n_blocks; b++) threads.emplace_back(feed, b);
    for (auto& t : threads) t.join();

    Writer w(out_path);
    w.put<uint32_t>(n_blocks);
    for (uint32_t b = 0; b < n_blocks; b++)
        {
            const int32_t event = fsm[b] && fsm[b]->valid > 0 ? 1 : static_cast<int32_t>(events[b].load());
            w.put<int32_t>(event);
            w.put<double>(synchro[b].Acq_delay_samples);
            w.put<double>(synchro[b].Acq_doppler_hz);
            w.put<uint64_t>(synchro[b].Acq_samplestamp_samples);
            w.put<float>(blocks[b]->test_statistics());
            w.put<float>(blocks[b]->threshold());
            w.put<uint32_t>(blocks[b]->num_doppler_bins());
            w.put<uint32_t>(blocks[b]->fft_size());
            w.put<uint32_t>(synchro[b].Acq_doppler_step);
            w.put<uint32_t>(blocks[b]->step_two() ? 1U : 0U);
            w.put<uint32_t>(fsm[b] ? static_cast<uint32_t>(fsm[b]->valid) : n_positive[b].load());
            w.put<uint32_t>(fsm[b] ? static_cast<uint32_t>(fsm[b]->failed) : n_negative[b].load());
        }
    return 0;
}

int run_mc(const char* in_path, const char* out_path)
{
    Reader r(in_path);
    const auto n_ch = r.get<uint32_t>();
    const auto n_taps = r.get<uint32_t>();
    const auto n_samples = r.get<uint32_t>();
    const auto code_len = r.get<uint32_t>();
    const auto n_epochs = r.get<uint32_t>();
    const auto mode = r.get<uint32_t>();  // 0 real 7-arg, 1 real 6-arg, 2 complex codes, 3 mixed: even channels real (5-arg), odd complex
    const auto high_dyn = r.get<uint32_t>();
    const auto coalesce_us = r.get<uint32_t>();
    struct Chan
    {
        std::vector<float> code;  // real: code_len, complex: 2*code_len
        std::vector<float> shifts;
        std::vector<std::complex<float>> sig;  // n_epochs * n_samples
        std::vector<float> args;               // n_epochs * 6
        std::vector<std::complex<float>> out;  // n_epochs * n_taps
    };
    std::vector<Chan> ch(n_ch);
    for (auto& c : ch)
        {
            c.code = r.vec<float>(mode >= 2 ? 2 * code_len : code_len);  // mixed: real channels use the first code_len values
            c.shifts = r.vec<float>(n_taps);
            c.sig = r.vec<std::complex<float>>(static_cast<size_t>(n_epochs) * n_samples);
            c.args = r.vec<float>(static_cast<size_t>(n_epochs) * 6);
            c.out.assign(static_cast<size_t>(n_epochs) * n_taps, std::complex<float>(0, 0));
        }
    B200_Correlator_Pool::instance().configure(0, std::max(64U, n_ch), std::max(n_samples, 1U), std::max(n_taps, 1U));
    B200_Correlator_Pool::instance().set_coalescing(coalesce_us, coalesce_us ? n_ch : 0);
    std::atomic<int> failures{0};
    auto work = [&](uint32_t i) {
        Chan& c = ch[i];
        std::vector<std::complex<float>> corr(n_taps);
        if (mode == 2 || (mode == 3 && (i & 1U) != 0))
            {
                B200_Multicorrelator mc;
                if (!mc.init(static_cast<int>(n_samples), static_cast<int>(n_taps))) failures++;
                if (!mc.set_local_code_and_taps(static_cast<int>(code_len), reinterpret_cast<const std::complex<float>*>(c.code.data()), c.shifts.data())) failures++;
                for (uint32_t e = 0; e < n_epochs; e++)
                    {
                        const float* a = &c.args[static_cast<size_t>(e) * 6];
                        mc.set_input_output_vectors(corr.data(), c.sig.data() + static_cast<size_t>(e) * n_samples);
                        if (!mc.Carrier_wipeoff_multicorrelator_resampler(a[0], a[1], a[3], a[4], static_cast<int>(n_samples))) failures++;
                        std::copy(corr.begin(), corr.end(), c.out.begin() + static_cast<size_t>(e) * n_taps);
                    }
                mc.free();
                return;
            }
        B200_Multicorrelator_Real_Codes mc;
        mc.set_high_dynamics_resampler(high_dyn != 0);
        if (!mc.init(static_cast<int>(n_samples), static_cast<int>(n_taps))) failures++;
        if (!mc.set_local_code_and_taps(static_cast<int>(code_len), c.code.data(), c.shifts.data())) failures++;
        for (uint32_t e = 0; e < n_epochs; e++)
            {
                const float* a = &c.args[static_cast<size_t>(e) * 6];
                mc.set_input_output_vectors(corr.data(), c.sig.data() + static_cast<size_t>(e) * n_samples);
                const bool ok = (mode == 0)   ? mc.Carrier_wipeoff_multicorrelator_resampler(a[0], a[1], a[2], a[3], a[4], a[5], static_cast<int>(n_samples))
                                : (mode == 1) ? mc.Carrier_wipeoff_multicorrelator_resampler(a[0], a[1], a[3], a[4], a[5], static_cast<int>(n_samples))
                                              : mc.Carrier_wipeoff_multicorrelator_resampler(a[0], a[1], 0.0F, a[3], a[4], 0.0F, static_cast<int>(n_samples));  // flags 0, like the complex class
                if (!ok) failures++;
                std::copy(corr.begin(), corr.end(), c.out.begin() + static_cast<size_t>(e) * n_taps);
            }
        mc.free();
    };
    std::vector<std::thread> threads;
    for (uint32_t i = 0; i < n_ch; i++) threads.emplace_back(work, i);
    for (auto& t : threads) t.join();
    Writer w(out_path);
    w.put<uint32_t>(static_cast<uint32_t>(failures.load()));
    w.put<uint64_t>(B200_Correlator_Pool::instance().batches());
    w.put<uint64_t>(B200_Correlator_Pool::instance().coalesced_calls());
    for (auto& c : ch) w.f.write(reinterpret_cast<const char*>(c.out.data()), sizeof(std::complex<float>) * c.out.size());
    if (failures != 0) std::cerr << "pool error: " << B200_Correlator_Pool::instance().last_error() << '\n';
    return failures == 0 ? 0 : 3;
}

// No CUDA device: the mirror must refuse loudly, not compute on the CPU.
int run_nodevice()
{
    B200_Multicorrelator_Real_Codes mc;
    const bool init_ok = mc.init(4000, 3);
    bool threw = false;
    try
        {
            Acq_Conf conf;
            conf.samples_per_ms = 4000.0F;
            conf.samples_per_code = 4000.0F;
            auto blk = pcps_make_acquisition_b200(conf);
            Gnss_Synchro s;
            blk->set_gnss_synchro(&s);
            blk->set_doppler_step(500);
            blk->init();
        }
    catch (const std::runtime_error& e)
        {
            threw = true;
            std::cout << "init refused: " << e.what() << '\n';
        }
    std::cout << "mc.init=" << (init_ok ? "true" : "false") << " acq_threw=" << (threw ? "true" : "false") << '\n';
    return (!init_ok && threw) ? 0 : 4;
}
// A MAT file with one variable of every class the dump uses; tests/test_host_cpu.py reads it back with scipy.
int run_mat(const char* out_path)
{
    B200_Mat_Writer mat;
    if (!mat.open(out_path)) return 3;
    std::vector<float> grid(7 * 3);
    for (size_t i = 0; i < grid.size(); i++) grid[i] = 0.5F * static_cast<float>(i);  // column-major 7 x 3
    bool ok = mat.write_single("acq_grid", grid.data(), 7, 3);
    ok = ok && mat.write_int32("doppler_max", -5000);
    ok = ok && mat.write_uint32("PRN", 4000000000U);
    ok = ok && mat.write_uint64("sample_counter", 0x123456789ABCDEFULL);
    ok = ok && mat.write_single("threshold", 2.5F);
    ok = ok && mat.write_double("x", 1.0 / 3.0);
    const int32_t v[5] = {1, -2, 3, -4, 5};
    ok = ok && mat.write_int32("odd_length_name_", v, 1, 5);
    ok = ok && !mat.write_single("", 1.0F);  // refused, nothing written
    return (ok && mat.close()) ? 0 : 4;
}
// Tracking dump: <in> = u32 n, u32 veml, u32 channel, u32 prn, f64 acq code phase, f64 acq Doppler, then n
// gnss_b200_trk_epoch records (as gnss_b200_trk_run returns them); <out> = the configured dump_filename.
int run_trkdump(const char* in_path, const char* dump_filename)
{
    Reader r(in_path);
    const auto n = r.get<uint32_t>();
    const bool veml = r.get<uint32_t>() != 0;
    const auto channel = r.get<uint32_t>();
    const auto prn = r.get<uint32_t>();
    const auto acq_phase = r.get<double>();
    const auto acq_doppler = r.get<double>();
    const auto epochs = r.vec<gnss_b200_trk_epoch>(n);
    B200_Tracking_Dump dump;
    if (!dump.configure(dump_filename, true)) return 3;
    if (!dump.set_channel(channel)) return 4;
    dump.set_acquisition(prn, acq_phase, acq_doppler);
    // two calls, as two gnss_b200_trk_run blocks would deliver them
    if (!dump.log(epochs.data(), n / 2, veml) || !dump.log(epochs.data() + n / 2, n - n / 2, veml)) return 5;
    return dump.finish() == 0 ? 0 : 6;
}
// Symbol synchronisation: <in> = u32 n, u32 secondary, u32 code_len, u32 symbols_per_bit, u32 data_len, code string bytes
// (code_len), data code string bytes (data_len), then n gnss_b200_trk_epoch records; <out> = u32 count, then per symbol
// { u32 epoch index, f64 I, f64 Q, u64 counter, u32 pll_180 }, then u32 locked, u32 current_symbol.
int run_symsync(const char* in_path, const char* out_path)
{
    Reader r(in_path);
    const auto n = r.get<uint32_t>();
    const bool secondary = r.get<uint32_t>() != 0;
    const auto code_len = r.get<uint32_t>();
    const auto spb = r.get<uint32_t>();
    const auto data_len = r.get<uint32_t>();
    const auto code = r.vec<char>(code_len);
    const auto data_code = r.vec<char>(data_len);
    const auto epochs = r.vec<gnss_b200_trk_epoch>(n);
    B200_Symbol_Sync sync;
    sync.configure(secondary, std::string(code.begin(), code.end()), code_len, spb, std::string(data_code.begin(), data_code.end()), data_len, false);
    struct Rec
    {
        uint32_t i;
        B200_Symbol_Sync::Symbol s;
    };
    std::vector<Rec> got;
    for (uint32_t i = 0; i < n; i++)
        {
            B200_Symbol_Sync::Symbol s{};
            if (sync.push(epochs[i], &s)) got.push_back(Rec{i, s});
        }
    Writer w(out_path);
    w.put<uint32_t>(static_cast<uint32_t>(got.size()));
    for (const auto& g : got)
        {
            w.put<uint32_t>(g.i);
            w.put<double>(g.s.Prompt_I);
            w.put<double>(g.s.Prompt_Q);
            w.put<uint64_t>(g.s.Tracking_sample_counter);
            w.put<uint32_t>(g.s.Flag_PLL_180_deg_phase_locked ? 1U : 0U);
        }
    w.put<uint32_t>(sync.locked() ? 1U : 0U);
    w.put<uint32_t>(sync.current_symbol());
    return 0;
}
}  // namespace

int main(int argc, char** argv)
{
    try
        {
            if (argc == 4 && std::strcmp(argv[1], "acq") == 0) return run_acq(argv[2], argv[3]);
            if (argc == 4 && std::strcmp(argv[1], "mc") == 0) return run_mc(argv[2], argv[3]);
            if (argc == 2 && std::strcmp(argv[1], "nodevice") == 0) return run_nodevice();
            if (argc == 3 && std::strcmp(argv[1], "mat") == 0) return run_mat(argv[2]);
            if (argc == 4 && std::strcmp(argv[1], "trkdump") == 0) return run_trkdump(argv[2], argv[3]);
            if (argc == 4 && std::strcmp(argv[1], "symsync") == 0) return run_symsync(argv[2], argv[3]);
        }
    catch (const std::exception& e)
        {
            std::cerr << "host_mirror_test: " << e.what() << '\n';
            return 2;
        }
    std::cerr << "usage: host_mirror_test acq|mc <in> <out> | trkdump <in> <dump_filename> | symsync <in> <out> | mat <out> | nodevice\n";
    return 1;
}
