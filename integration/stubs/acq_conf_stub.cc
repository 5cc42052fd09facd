// Acq_Conf::SetFromConfiguration for builds outside the GNSS-SDR tree (inside it the reference's acq_conf.cc is used).
// Restates which configuration key feeds which member and the derived quantities of acq_conf.cc:24-119.
#include "acq_conf.h"
#include "configuration_interface.h"
#include <cmath>
#include <complex>
#include <stdexcept>

void Acq_Conf::SetFromConfiguration(const ConfigurationInterface* cfg, const std::string& role, double chip_rate, double opt_freq)
{
    const auto key = [&role](const char* name) { return role + "." + name; };
    item_type = cfg->property(key("item_type"), item_type);
    if (item_type == "gr_complex")
        it_size = sizeof(std::complex<float>);
    else if (item_type == "cshort")
        it_size = 2 * sizeof(int16_t);
    else if (item_type == "cbyte")
        it_size = 2 * sizeof(int8_t);
    else
        throw std::invalid_argument("Unknown item type: " + item_type);
    chips_per_second = static_cast<uint32_t>(chip_rate);
    fs_in = cfg->property("GNSS-SDR.internal_fs_sps", cfg->property("GNSS-SDR.internal_fs_hz", fs_in));
    doppler_max = cfg->property(key("doppler_max"), doppler_max);
    sampled_ms = cfg->property(key("coherent_integration_time_ms"), sampled_ms);
    bit_transition_flag = cfg->property(key("bit_transition_flag"), bit_transition_flag);
    max_dwells = cfg->property(key("max_dwells"), max_dwells);
    dump = cfg->property(key("dump"), dump);
    dump_channel = cfg->property(key("dump_channel"), dump_channel);
    blocking = cfg->property(key("blocking"), blocking);
    dump_filename = cfg->property(key("dump_filename"), dump_filename);
    use_automatic_resampler = cfg->property("GNSS-SDR.use_acquisition_resampler", use_automatic_resampler);
    if (sampled_ms % ms_per_code != 0) sampled_ms = ms_per_code;  // the coherent time is a whole number of code periods

    // the acquisition resampler decimates by the largest divisor of fs_in that keeps the rate at or above opt_freq
    resampled_fs = fs_in;
    if (use_automatic_resampler && static_cast<double>(fs_in) > opt_freq)
        {
            auto decimation = static_cast<uint32_t>(static_cast<double>(fs_in) / opt_freq);
            while (decimation > 1 && fs_in % decimation != 0) decimation--;
            resampler_ratio = static_cast<float>(decimation);
            resampled_fs = fs_in / static_cast<int64_t>(decimation);
        }

    num_doppler_bins_step2 = cfg->property(key("second_nbins"), num_doppler_bins_step2);
    doppler_step2 = cfg->property(key("second_doppler_step"), doppler_step2);
    doppler_step = cfg->property(key("doppler_step"), doppler_step);
    pfa = cfg->property(key("pfa"), pfa);
    if (pfa < 0.0F || pfa > 1.0F) pfa = 0.0F;
    pfa2 = cfg->property(key("pfa_second_step"), pfa2);
    if (pfa2 <= 0.0F || pfa2 > 1.0F) pfa2 = pfa;
    make_2_steps = cfg->property(key("make_two_steps"), make_2_steps);
    make_repeat_steps = cfg->property(key("make_repeat_steps"), make_repeat_steps);
    blocking_on_standby = cfg->property(key("blocking_on_standby"), blocking_on_standby);
    if (pfa <= 0.0F) use_CFAR_algorithm_flag = false;  // no pfa: first-vs-second-peak statistic against the plain threshold
    enable_monitor_output = cfg->property("AcquisitionMonitor.enable_monitor", false);

    samples_per_ms = static_cast<float>(resampled_fs) * 0.001F;
    samples_per_chip = static_cast<uint32_t>(std::ceil(static_cast<float>(resampled_fs) / static_cast<float>(chips_per_second)));
    samples_per_code = samples_per_ms * static_cast<float>(ms_per_code);
}
