// TEST HARNESS — drives a B200 acquisition ADAPTER (integration/src/**) the way the reference's Channel and flow graph do
// (src/algorithms/channel/adapters/channel.cc:49-93,215-232; gnss_block_factory.cc:1335-1350), against the stand-in headers of
// integration/stubs/. The sampled-code generators are doubles that hand back the code the Python test computed (pinned on the
// reference generators by tests/golden/sampled_code_sha256.json).
//   adapter_test <input file> <output file>
#include "acquisition_interface.h"
#include "channel_fsm.h"
#include "configuration_interface.h"
#include "gnss_block_factory_b200.h"
#include "gnss_sdr_flags.h"
#include "gnss_synchro.h"
#include "signal_replicas.h"
#include <algorithm>
#include <atomic>
#include <complex>
#include <cstdint>
#include <cstring>
#include <fstream>
#include <iostream>
#include <memory>
#include <string>
#include <vector>

int32_t FLAGS_doppler_max = 0;

namespace
{
std::vector<std::complex<float>> g_code;  // one code period at the acquisition rate
int g_generator_calls = 0;
void hand_back(own::span<std::complex<float>> dest)
{
    g_generator_calls++;
    std::copy_n(g_code.begin(), std::min(dest.size(), g_code.size()), dest.begin());
}
}  // namespace
void gps_l1_ca_code_gen_complex_sampled(own::span<std::complex<float>> dest, uint32_t, int32_t, uint32_t) { hand_back(dest); }
void galileo_e1_code_gen_complex_sampled(own::span<std::complex<float>> dest, const std::array<char, 3>&, bool, uint32_t, int32_t, uint32_t, bool) { hand_back(dest); }
void beidou_b1i_code_gen_complex_sampled(own::span<std::complex<float>> dest, uint32_t, int32_t, uint32_t) { hand_back(dest); }

struct CountingFsm : public ChannelFsm
{
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

template <class T>
T get(std::ifstream& f)
{
    T v{};
    f.read(reinterpret_cast<char*>(&v), sizeof(T));
    return v;
}

int main(int argc, char** argv)
{
    if (argc != 3) return 2;
    std::ifstream f(argv[1], std::ios::binary);
    const auto fs = get<int64_t>(f);
    const auto doppler_max = get<int32_t>(f);
    const auto doppler_step = get<uint32_t>(f);
    const auto pfa = get<float>(f);
    const auto coherent_ms = get<uint32_t>(f);
    const auto which = get<uint32_t>(f);
    const auto prn = get<uint32_t>(f);
    const auto rf_channel = get<uint32_t>(f);
    g_code.resize(get<uint32_t>(f));
    f.read(reinterpret_cast<char*>(g_code.data()), static_cast<std::streamsize>(g_code.size() * sizeof(std::complex<float>)));
    std::vector<std::complex<float>> stream(get<uint32_t>(f));
    f.read(reinterpret_cast<char*>(stream.data()), static_cast<std::streamsize>(stream.size() * sizeof(std::complex<float>)));

    static const char* kImpl[] = {"GPS_L1_CA_PCPS_Acquisition_B200", "Galileo_E1_PCPS_Ambiguous_Acquisition_B200", "BEIDOU_B1I_PCPS_Acquisition_B200"};
    static const char* kSignal[] = {"1C", "1B", "B1"};
    static const char kSystem[] = {'G', 'E', 'C'};
    const std::string signal = kSignal[which];
    const std::string role = "Acquisition_" + signal;
    InMemoryConfiguration config;
    config.set_property("GNSS-SDR.internal_fs_sps", std::to_string(fs));
    config.set_property(role + ".implementation", kImpl[which]);
    config.set_property(role + ".item_type", "gr_complex");
    config.set_property(role + ".coherent_integration_time_ms", std::to_string(coherent_ms));
    config.set_property(role + ".pfa", std::to_string(pfa));
    config.set_property(role + ".doppler_max", std::to_string(doppler_max));
    config.set_property(role + ".doppler_step", std::to_string(doppler_step));
    config.set_property(role + ".max_dwells", "1");
    config.set_property("Channels_" + signal + ".RF_channel_ID", std::to_string(rf_channel));

    // ---- GNSSBlockFactory::GetAcqBlock ----
    const std::string implementation = config.property(role + ".implementation", std::string("Wrong"));
    if (!IsB200AcquisitionImplementation(implementation) || IsB200AcquisitionImplementation("GPS_L1_CA_PCPS_Acquisition")) return 3;
    std::unique_ptr<AcquisitionInterface> acq;
    try
        {
            acq = GetB200AcquisitionBlock(&config, role, implementation, 1, 0);
        }
    catch (const std::exception& e)
        {
            std::cerr << "adapter: " << e.what() << '\n';
            return 4;
        }
    if (!acq || acq->implementation() != implementation || acq->role() != role || acq->item_size() != sizeof(std::complex<float>)) return 5;
    if (GetB200AcquisitionBlock(&config, role, "GPS_L1_CA_PCPS_Acquisition", 1, 0) != nullptr) return 6;

    // ---- Channel::Channel (channel.cc:49-93) ----
    auto fsm = std::make_shared<CountingFsm>();
    Gnss_Synchro synchro{};
    const unsigned int channel = 3;
    acq->set_channel(channel);
    acq->set_channel_fsm(fsm);
    synchro.Channel_ID = static_cast<int32_t>(channel);
    acq->set_gnss_synchro(&synchro);
    acq->set_doppler_step(doppler_step);
    acq->set_threshold(0.0F);
    try
        {
            acq->init();
        }
    catch (const std::exception& e)
        {
            std::cerr << "init: " << e.what() << '\n';  // no CUDA device: the adapter must refuse loudly
            return 7;
        }
    // ---- Channel::set_signal (:215-232) ----
    std::memcpy(synchro.Signal, signal.c_str(), 3);
    synchro.PRN = prn;
    synchro.System = kSystem[which];
    acq->set_local_code();

    // ---- the flow graph connects and starts the blocks; Channel::start_acquisition re-arms the adapter ----
    auto top = std::make_shared<gr::top_block>();
    acq->connect(top);
    auto block = std::dynamic_pointer_cast<gr::block>(acq->get_left_block());
    if (!block || acq->get_right_block() != acq->get_left_block()) return 8;
    block->start();
    acq->reset();

    size_t pos = 0;
    int calls = 0;
    while (fsm->valid + fsm->failed == 0 && calls < 100000)
        {
            const size_t avail = std::min<size_t>(1000, stream.size() - pos);
            gr_vector_int ninput{static_cast<int>(avail)};
            gr_vector_const_void_star in{static_cast<const void*>(stream.data() + pos)};
            gr_vector_void_star out;
            const long before = block->consumed_so_far();
            block->general_work(0, ninput, in, out);
            pos += static_cast<size_t>(block->consumed_so_far() - before);
            calls++;
            if (avail == 0) break;
        }
    acq->stop_acquisition();
    acq->disconnect(top);

    std::ofstream o(argv[2], std::ios::binary);
    const int32_t event = fsm->valid > 0 ? 1 : (fsm->failed > 0 ? 2 : 0);
    const auto put = [&o](const auto& v) { o.write(reinterpret_cast<const char*>(&v), sizeof(v)); };
    put(event);
    put(synchro.Acq_delay_samples);
    put(synchro.Acq_doppler_hz);
    put(static_cast<uint64_t>(synchro.Acq_samplestamp_samples));
    put(static_cast<uint64_t>(pos));
    put(static_cast<int32_t>(g_generator_calls));
    put(static_cast<int32_t>(acq->mag()));
    return 0;
}
