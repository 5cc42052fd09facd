/*
 * pcps_acquisition_b200.cc — see the header. Comments of the form ":NNN" point at the line of
 * src/algorithms/acquisition/gnuradio_blocks/pcps_acquisition.cc whose behaviour the statement reproduces.
 * SPDX-License-Identifier: GPL-3.0-or-later
 */
#include "pcps_acquisition_b200.h"
#if !defined(B200_HOST_STANDALONE)
#include "GLONASS_L1_L2_CA.h"  // DFRQ1_GLO, DFRQ2_GLO, GLONASS_PRN
#endif
#include "b200_mat_writer.h"
#include <algorithm>
#include <cmath>
#include <filesystem>
#include <cstring>
#include <iostream>
#include <stdexcept>
#include <sstream>
#include <thread>

pcps_acquisition_b200_sptr pcps_make_acquisition_b200(const Acq_Conf& conf_)
{
    return pcps_acquisition_b200_sptr(new pcps_acquisition_b200(conf_));
}


pcps_acquisition_b200::pcps_acquisition_b200(const Acq_Conf& conf_)
    : d_dump_filename(conf_.dump_filename),  // :55
      d_acq_parameters(conf_),
      d_samplesPerChip(conf_.samples_per_chip),
      d_doppler_step(static_cast<uint32_t>(conf_.doppler_step)),
      d_consumed_samples(static_cast<uint32_t>(conf_.sampled_ms * conf_.samples_per_ms * (conf_.bit_transition_flag ? 2.0 : 1.0))),  // :71
      d_num_doppler_bins_step2(conf_.num_doppler_bins_step2),                                                                          // :73
      d_use_CFAR_algorithm_flag(conf_.use_CFAR_algorithm_flag)
{
    d_dump_channel = conf_.dump_channel;  // :74
    d_dump = conf_.dump;                  // :81
    if (conf_.make_2_steps && (conf_.num_doppler_bins_step2 == 0 || !(conf_.doppler_step2 > 0.0F)))
        {
            throw std::invalid_argument("pcps_acquisition_b200: make_two_steps needs second_nbins >= 1 and second_doppler_step > 0");
        }
    if (d_consumed_samples == 0)
        {
            throw std::invalid_argument("pcps_acquisition_b200: sampled_ms * samples_per_ms is zero");
        }
    d_fft_size = (conf_.sampled_ms == conf_.ms_per_code) ? d_consumed_samples : d_consumed_samples * 2;  // :85-92
    d_cshort = (conf_.it_size != sizeof(std::complex<float>));                                           // :121-128
    d_data_buffer.resize(d_consumed_samples);
    if (d_cshort) d_data_buffer_sc.resize(static_cast<size_t>(d_consumed_samples) * 2);

    if (d_dump)
        {
            // Same result as the reference's string surgery (:135-165): directory part or ".", default stem "acquisition", the last
            // extension dropped (a leading dot is not an extension), directory created; on failure dumping is switched off.
            namespace fs = std::filesystem;
            const fs::path given(d_dump_filename);
            const fs::path dir = d_dump_filename.find('/') != std::string::npos ? fs::path(d_dump_filename.substr(0, d_dump_filename.find_last_of('/'))) : fs::path(".");
            std::string stem = given.filename().empty() ? std::string("acquisition") : given.filename().stem().string();
            if (stem.empty()) stem = given.filename().string();
            d_dump_filename = dir.string() + '/' + stem;
            std::error_code ec;
            fs::create_directories(dir, ec);
            if (ec)
                {
                    std::cerr << "GNSS-SDR cannot create dump file for the Acquisition block. Wrong permissions?\n";
                    d_dump = false;
                }
        }
}


// pcps_acquisition.h:160-163. The dump channel keeps its magnitude grid on the device, so a channel number that arrives
// after init() re-attaches the block to a bank of the matching kind.
void pcps_acquisition_b200::set_channel(uint32_t channel)
{
    std::lock_guard<std::recursive_mutex> lock(d_setlock);
    const bool was = dumping();
    d_channel = channel;
    if (d_bank && was != dumping())
        {
            attach_engine();
            d_bank->reset(d_slot);
            refresh_doppler_grid();
        }
}


pcps_acquisition_b200::~pcps_acquisition_b200()
{
    if (d_bank && d_slot >= 0) d_bank->release_slot(d_slot);
}


void pcps_acquisition_b200::require_engine(const char* where) const
{
    if (!d_bank || d_slot < 0)
        {
            // The reference has no failure path here; running without the device would silently change results,
            // so this is loud.
            throw std::runtime_error(std::string("pcps_acquisition_b200::") + where + ": B200 engine not available (call init() on a machine with a CUDA device)");
        }
}


void pcps_acquisition_b200::set_gnss_synchro(Gnss_Synchro* p_gnss_synchro)
{
    std::lock_guard<std::recursive_mutex> lock(d_setlock);
    d_gnss_synchro = p_gnss_synchro;
}


void pcps_acquisition_b200::set_active(bool active)
{
    std::lock_guard<std::recursive_mutex> lock(d_setlock);
    d_active = active;
}


void pcps_acquisition_b200::set_threshold(float threshold)
{
    std::lock_guard<std::recursive_mutex> lock(d_setlock);
    d_threshold = threshold;
}


void pcps_acquisition_b200::set_doppler_max(uint32_t doppler_max)
{
    std::lock_guard<std::recursive_mutex> lock(d_setlock);
    d_acq_parameters.doppler_max = static_cast<int32_t>(doppler_max);
}


void pcps_acquisition_b200::set_doppler_step(uint32_t doppler_step)
{
    std::lock_guard<std::recursive_mutex> lock(d_setlock);
    d_doppler_step = doppler_step;
}


void pcps_acquisition_b200::set_resampler_latency(uint32_t latency_samples)
{
    std::lock_guard<std::recursive_mutex> lock(d_setlock);
    d_acq_parameters.resampler_latency_samples = latency_samples;  // :170-174
}


// pcps_acquisition.h:199-208 — a new centre rebuilds the wipe-off grid; here it re-keys the slot's Doppler group
void pcps_acquisition_b200::set_doppler_center(int32_t doppler_center)
{
    std::lock_guard<std::recursive_mutex> lock(d_setlock);
    if (doppler_center != d_doppler_center)
        {
            d_doppler_center = doppler_center;
            refresh_doppler_grid();
        }
}


void pcps_acquisition_b200::refresh_doppler_grid()
{
    if (d_bank && d_slot >= 0) d_bank->set_doppler_center(d_slot, d_doppler_center, d_doppler_bias);  // :298-306
}


// :212-230
bool pcps_acquisition_b200::is_fdma()
{
    d_doppler_bias = 0;
#if !defined(B200_HOST_STANDALONE)
    if (d_gnss_synchro != nullptr && std::strcmp(d_gnss_synchro->Signal, "1G") == 0)
        {
            d_doppler_bias = static_cast<int32_t>(DFRQ1_GLO * GLONASS_PRN.at(d_gnss_synchro->PRN));
            return true;
        }
    if (d_gnss_synchro != nullptr && std::strcmp(d_gnss_synchro->Signal, "2G") == 0)
        {
            d_doppler_bias += static_cast<int32_t>(DFRQ2_GLO * GLONASS_PRN.at(d_gnss_synchro->PRN));
            return true;
        }
#endif
    return false;
}


// :176-209 — zero-pad layout, forward FFT and conjugation happen on the device (gnss_b200_acq_set_local_code)
void pcps_acquisition_b200::set_local_code(std::complex<float>* code)
{
    std::lock_guard<std::recursive_mutex> lock(d_setlock);
    if (is_fdma())
        {
            refresh_doppler_grid();  // :179-183
        }
    const uint32_t n = d_acq_parameters.bit_transition_flag ? d_fft_size / 2 : d_consumed_samples;
    d_code_copy.assign(d_consumed_samples, std::complex<float>(0.0F, 0.0F));
    std::copy(code, code + n, d_code_copy.begin());
    d_have_code = true;
    if (d_bank && d_slot >= 0)
        {
            if (!d_bank->set_local_code(d_slot, d_code_copy.data()))
                {
                    throw std::runtime_error("pcps_acquisition_b200::set_local_code: " + d_bank->last_error());
                }
        }
}


// :249-295
void pcps_acquisition_b200::init()
{
    std::lock_guard<std::recursive_mutex> lock(d_setlock);
    if (d_gnss_synchro != nullptr)
        {
            d_gnss_synchro->Flag_valid_acquisition = false;
            d_gnss_synchro->Flag_valid_symbol_output = false;
            d_gnss_synchro->Flag_valid_pseudorange = false;
            d_gnss_synchro->Flag_valid_word = false;
            forget_last_answer();
        }
    d_mag = 0.0;
    d_input_power = 0.0;
    d_step_repeat = false;  // :262
    d_num_doppler_bins = static_cast<uint32_t>(std::ceil(static_cast<double>(static_cast<int32_t>(d_acq_parameters.doppler_max) - static_cast<int32_t>(-d_acq_parameters.doppler_max)) / static_cast<double>(d_doppler_step)));  // :264

    attach_engine();
    d_bank->reset(d_slot);  // :281-284
    refresh_doppler_grid();  // :286
    d_num_noncoherent_integrations_counter = 0U;
    d_worker_active = false;
}


// The engine side of init(): a bank for this block's search geometry, a slot in it, the local code if there is one.
void pcps_acquisition_b200::attach_engine()
{
    B200_Acquisition_Bank::Geometry g{};
    g.fft_size = d_fft_size;
    g.consumed_samples = d_consumed_samples;
    g.doppler_max = d_acq_parameters.doppler_max;
    g.doppler_step = d_doppler_step;
    // update_local_carrier (:236-243): the carrier is generated at the rate of the items the block receives
    g.fs = static_cast<double>(d_acq_parameters.use_automatic_resampler ? d_acq_parameters.resampled_fs : d_acq_parameters.fs_in);
    g.samples_per_chip = d_samplesPerChip;
    g.bit_transition = d_acq_parameters.bit_transition_flag;
    g.use_cfar = d_use_CFAR_algorithm_flag;
    // the grid stays on the device only where it is needed: accumulated over dwells (:672-684) or dumped (:682, :743)
    g.keep_grid = (d_acq_parameters.max_dwells > 1 && !d_acq_parameters.bit_transition_flag) || dumping();
    g.device = d_device;
    g.source_id = d_source_id;
    // several GPUs per search only where the decision can be taken from row records (gnss_b200_acq_group_create)
    g.n_devices = (d_num_devices > 1 && d_use_CFAR_algorithm_flag && !d_cshort) ? d_num_devices : 1U;
    g.step2_bins = d_acq_parameters.make_2_steps ? d_num_doppler_bins_step2 : 0U;
    g.step2_hz = d_acq_parameters.make_2_steps ? d_acq_parameters.doppler_step2 : 0.0F;
    std::string why;
    auto bank = B200_Acquisition_Bank::get(g, &why);
    if (!bank)
        {
            throw std::runtime_error("pcps_acquisition_b200: engine: cannot create the B200 engine: " + why);
        }
    if (bank != d_bank)
        {
            if (d_bank && d_slot >= 0) d_bank->release_slot(d_slot);
            d_bank = bank;
            d_slot = d_bank->acquire_slot();
            if (d_slot < 0) throw std::runtime_error("pcps_acquisition_b200: engine: " + d_bank->last_error());
            if (d_have_code && !d_bank->set_local_code(d_slot, d_code_copy.data()))  // d_fft_codes survives init() in the reference
                {
                    throw std::runtime_error("pcps_acquisition_b200: engine: " + d_bank->last_error());
                }
        }
    if (dumping())  // :289-294
        {
            const uint32_t effective_fft_size = (d_acq_parameters.bit_transition_flag ? (d_fft_size / 2) : d_fft_size);
            d_grid.assign(static_cast<size_t>(effective_fft_size) * d_num_doppler_bins, 0.0F);
            d_narrow_grid.assign(static_cast<size_t>(effective_fft_size) * d_num_doppler_bins_step2, 0.0F);
        }
}

// The four Gnss_Synchro fields an acquisition owns, back to "nothing found yet" (:255-258, :326-329, :958-961).
void pcps_acquisition_b200::forget_last_answer()
{
    if (d_gnss_synchro == nullptr) return;
    d_gnss_synchro->Acq_delay_samples = 0.0;
    d_gnss_synchro->Acq_doppler_hz = 0.0;
    d_gnss_synchro->Acq_samplestamp_samples = 0ULL;
    d_gnss_synchro->Acq_doppler_step = 0U;
}


// Code phase, Doppler and sample stamp of a finished search into Gnss_Synchro; with the acquisition resampler the delay and
// the stamp go back to the input rate and the FIR's latency is taken off (:699-712).
void pcps_acquisition_b200::report_answer(uint32_t indext, int32_t doppler, uint64_t samp_count)
{
    const auto delay = static_cast<double>(std::fmod(static_cast<float>(indext), d_acq_parameters.samples_per_code));
    d_gnss_synchro->Acq_doppler_hz = static_cast<double>(doppler);
    if (d_acq_parameters.use_automatic_resampler)
        {
            d_gnss_synchro->Acq_delay_samples = delay * d_acq_parameters.resampler_ratio;
            d_gnss_synchro->Acq_delay_samples -= static_cast<double>(d_acq_parameters.resampler_latency_samples);
            d_gnss_synchro->Acq_samplestamp_samples = rint(static_cast<double>(samp_count) * d_acq_parameters.resampler_ratio);
        }
    else
        {
            d_gnss_synchro->Acq_delay_samples = delay;
            d_gnss_synchro->Acq_samplestamp_samples = samp_count;
        }
}


// :317-343
void pcps_acquisition_b200::set_state(int32_t state)
{
    std::lock_guard<std::recursive_mutex> lock(d_setlock);
    d_state = state;
    if (d_state == 1)
        {
            if (!d_step_repeat)  // :324 — the fork's repeat mode keeps the last answer in Gnss_Synchro
                {
                    forget_last_answer();
                }
            d_mag = 0.0;
            d_test_statistics = 0.0;
            d_active = true;
        }
    else if (d_state != 0)
        {
            std::cerr << "pcps_acquisition_b200: State can only be set to 0 or 1\n";
        }
}


// :345-381
void pcps_acquisition_b200::send_positive_acquisition()
{
    if (!d_step_repeat)  // :360-363 — no dump file while repeating
        {
            d_positive_acq = 1;
        }
    if (d_acq_parameters.make_repeat_steps)  // :365-368
        {
            d_step_repeat = true;
        }
    if (!d_channel_fsm.expired())
        {
            d_channel_fsm.lock()->Event_valid_acquisition();
        }
    else
        {
            publish_event(1);
        }
    if (d_acq_parameters.enable_monitor_output)
        {
            d_monitor_queue.push(*d_gnss_synchro);
        }
}


// :384-405
void pcps_acquisition_b200::send_negative_acquisition()
{
    d_positive_acq = 0;
    publish_event(2);
}


// :408-508 — same file name, variables, classes and shapes; Level-5 container (b200_mat_writer.h)
void pcps_acquisition_b200::dump_results(int32_t effective_fft_size)
{
    d_dump_number++;
    // <prefix>_<System>_<Signal>_ch_<channel>_<n>_sat_<PRN>.mat — the pattern the reference's readers glob for (:408-423)
    std::ostringstream name;
    name << d_dump_filename << '_' << d_gnss_synchro->System << '_' << d_gnss_synchro->Signal[0] << d_gnss_synchro->Signal[1] << "_ch_" << d_channel << '_'
         << d_dump_number << "_sat_" << d_gnss_synchro->PRN << ".mat";
    const std::string filename = name.str();

    B200_Mat_Writer mat;
    if (!mat.open(filename))
        {
            std::cout << "Unable to create or open Acquisition dump file\n";
            return;
        }
    mat.write_single("acq_grid", d_grid.data(), static_cast<size_t>(effective_fft_size), d_num_doppler_bins);
    mat.write_int32("doppler_max", static_cast<int32_t>(d_acq_parameters.doppler_max));
    mat.write_int32("doppler_step", static_cast<int32_t>(d_doppler_step));
    mat.write_int32("d_positive_acq", d_positive_acq);
    mat.write_single("acq_doppler_hz", static_cast<float>(d_gnss_synchro->Acq_doppler_hz));
    mat.write_single("acq_delay_samples", static_cast<float>(d_gnss_synchro->Acq_delay_samples));
    mat.write_single("test_statistic", d_test_statistics);
    mat.write_single("threshold", d_threshold);
    mat.write_single("input_power", d_input_power);
    mat.write_uint64("sample_counter", d_sample_counter);
    mat.write_uint32("PRN", d_gnss_synchro->PRN);
    mat.write_int32("num_dwells", static_cast<int32_t>(d_num_noncoherent_integrations_counter));
    if (d_acq_parameters.make_2_steps)
        {
            mat.write_single("acq_grid_narrow", d_narrow_grid.data(), static_cast<size_t>(effective_fft_size), d_num_doppler_bins_step2);
            mat.write_single("doppler_step_narrow", d_acq_parameters.doppler_step2);
            mat.write_single("doppler_grid_narrow_min", d_doppler_center_step_two - static_cast<float>(std::floor(d_num_doppler_bins_step2 / 2.0)) * d_acq_parameters.doppler_step2);
        }
    mat.close();
}


// :885-890
bool pcps_acquisition_b200::start()
{
    d_sample_counter = 0ULL;
    calculate_threshold();
    return true;
}


// :894-909
void pcps_acquisition_b200::calculate_threshold()
{
    const float pfa = (d_step_two ? d_acq_parameters.pfa2 : d_acq_parameters.pfa);
    if (pfa <= 0.0) return;
    const uint32_t effective_fft_size = d_acq_parameters.bit_transition_flag ? d_fft_size / 2 : d_fft_size;
    const uint32_t num_doppler_bins = (d_step_two ? d_num_doppler_bins_step2 : d_num_doppler_bins);
    d_threshold = gnss_b200_acq_threshold(pfa, effective_fft_size, num_doppler_bins, d_acq_parameters.max_dwells, d_acq_parameters.bit_transition_flag ? 1 : 0);
}


// :615-882. Either Doppler loop (:655-687 wide, :715-745 narrow) with its statistic (:689-696, :746-754) is one engine call.
void pcps_acquisition_b200::acquisition_core(uint64_t samp_count)
{
    std::unique_lock<std::recursive_mutex> lk(d_setlock);
    require_engine("acquisition_core");
    d_mag = 0.0;
    d_num_noncoherent_integrations_counter++;  // :638
    const void* samples = d_cshort ? static_cast<const void*>(d_data_buffer_sc.data()) : static_cast<const void*>(d_data_buffer.data());
    if (d_acq_parameters.blocking)
        {
            lk.unlock();  // :649-652
        }

    gnss_b200_acq_result res{};
    const bool step_two = d_step_two;
    const bool ok = step_two ? d_bank->dwell_step2(d_slot, samples, d_cshort, samp_count, d_doppler_center_step_two, d_input_power, &res)
                             : d_bank->dwell(d_slot, samples, d_cshort, samp_count, &res);
    if (!ok)
        {
            // the reference logs and keeps going; a failed device call must not look like a detection
            std::cerr << "pcps_acquisition_b200: engine dwell failed: " << d_bank->last_error() << '\n';
            res = gnss_b200_acq_result{};
        }
    if (ok && dumping())  // :682-685, :743-746 — the dump channel mirrors the magnitude grid after every pass
        {
            const bool got = step_two ? d_bank->read_grid_step2(d_slot, d_narrow_grid.data()) : d_bank->read_grid(d_slot, d_grid.data());
            if (!got) std::cerr << "pcps_acquisition_b200: cannot read the magnitude grid for the dump\n";
        }
    const uint32_t indext = res.code_phase;
    const int32_t doppler = res.doppler_hz;
    if (!step_two) d_input_power = res.input_power;  // :531-535 refresh it in the first step only
    d_test_statistics = res.test_statistic;

    // :697 — a repeated coarse search of the two-step mode leaves Gnss_Synchro alone
    if (step_two || !(d_acq_parameters.make_2_steps && d_step_repeat))
        {
            report_answer(indext, doppler, samp_count);
        }
    if (step_two)
        {
            d_gnss_synchro->Acq_doppler_step = d_acq_parameters.doppler_step2;  // :763, :770
        }

    if (d_acq_parameters.blocking)
        {
            lk.lock();  // :776-779
        }

    // What one finished dwell means for the block (the reference spells this out twice, for searches with and without the
    // bit-transition option, :781-866; the two differ only in that a bit-transition search never waits for another dwell).
    //   detected, and this is a one-step search or already the narrow second step   -> Verdict::positive
    //   detected in the coarse step of a two-step search                           -> Verdict::refine (re-arm around the coarse Doppler)
    //   not detected and dwells left                                               -> Verdict::another_dwell
    //   not detected and this was the last dwell                                   -> Verdict::negative
    enum class Verdict
    {
        positive,
        refine,
        another_dwell,
        negative
    };
    const bool detected = d_test_statistics > d_threshold;
    const bool last_dwell = d_acq_parameters.bit_transition_flag || d_num_noncoherent_integrations_counter == d_acq_parameters.max_dwells;
    const Verdict verdict = detected ? ((d_acq_parameters.make_2_steps && !d_step_two) ? Verdict::refine : Verdict::positive)
                                     : (last_dwell ? Verdict::negative : Verdict::another_dwell);
    switch (verdict)
        {
        case Verdict::positive:
            d_active = false;
            send_positive_acquisition();
            d_step_two = false;
            d_state = 0;
            if (d_acq_parameters.make_2_steps) calculate_threshold();  // back to the coarse grid's pfa
            break;
        case Verdict::refine:
            d_active = false;
            d_step_two = true;  // general_work re-arms the block around Acq_doppler_hz
            d_num_noncoherent_integrations_counter = 0U;
            d_positive_acq = 0;
            d_state = 0;
            d_bank->reset(d_slot);  // the first narrow dwell starts a fresh grid
            calculate_threshold();  // the narrow grid's pfa
            break;
        case Verdict::another_dwell:
            d_buffer_count = 0U;
            d_state = 1;
            break;
        case Verdict::negative:
            send_negative_acquisition();
            d_state = 0;
            d_active = false;
            if (d_step_two)
                {
                    d_step_two = false;
                    calculate_threshold();
                }
            break;
        }
    d_worker_active = false;

    if ((d_num_noncoherent_integrations_counter == d_acq_parameters.max_dwells) || (d_positive_acq == 1) || d_acq_parameters.bit_transition_flag)  // :869-880
        {
            if (dumping())  // :875-878
                {
                    dump_results(static_cast<int32_t>(d_acq_parameters.bit_transition_flag ? d_fft_size / 2 : d_fft_size));
                }
            d_num_noncoherent_integrations_counter = 0U;
            d_positive_acq = 0;
            d_bank->reset(d_slot);  // the next dwell starts a fresh grid (:672-676 overwrite on counter == 1)
        }
}


// general_work (:912-1050) as four small steps; what each does to the sample counter and to the scheduler (consume_each) is the
// reference's, because sample stamps and hand-over to tracking depend on it.

// The block is idle, or a worker thread is searching: items flow past (unless blocking_on_standby), and a pending second step /
// the fork's repeat mode re-arm the block.
void pcps_acquisition_b200::pass_items_while_idle(int n_items)
{
    const bool dwells_exhausted = d_num_noncoherent_integrations_counter == d_acq_parameters.max_dwells;
    if (!d_acq_parameters.blocking_on_standby && (!d_active || dwells_exhausted))  // not while a non-coherent integration is in progress
        {
            d_sample_counter += static_cast<uint64_t>(n_items);
            consume_each(n_items);
        }
    if (d_step_two)
        {
            d_doppler_center_step_two = static_cast<float>(d_gnss_synchro->Acq_doppler_hz);  // centre of the narrow grid
            d_state = 0;
            d_active = true;
        }
    if (d_step_repeat) d_active = true;
}

// State 0: a search begins — forget the previous answer (not in repeat mode), drop this call's items.
void pcps_acquisition_b200::begin_search(int n_items)
{
    if (!d_step_repeat) forget_last_answer();
    d_mag = 0.0;
    d_buffer_count = 0U;
    d_state = 1;
    if (!d_acq_parameters.blocking_on_standby)
        {
            d_sample_counter += static_cast<uint64_t>(n_items);
            consume_each(n_items);
        }
}

// State 1: copy items into the dwell buffer. The fill level is tested BEFORE this call's items are added (as the reference
// does), so the search starts one call after the buffer filled.
void pcps_acquisition_b200::fill_dwell_buffer(int n_items, const void* items)
{
    const auto offered = static_cast<uint32_t>(n_items);
    const uint32_t room = d_consumed_samples - std::min(d_buffer_count, d_consumed_samples);
    const uint32_t take = std::min(offered, room);
    if (d_cshort)
        {
            const auto* in = static_cast<const int16_t*>(items);
            std::copy_n(in, static_cast<size_t>(take) * 2, d_data_buffer_sc.begin() + static_cast<size_t>(d_buffer_count) * 2);
        }
    else
        {
            std::copy_n(static_cast<const std::complex<float>*>(items), take, d_data_buffer.begin() + d_buffer_count);
        }
    if (d_buffer_count >= d_consumed_samples) d_state = 2;
    d_buffer_count += take;
    d_sample_counter += static_cast<uint64_t>(take);
    consume_each(static_cast<int>(take));
}

// State 2: run acquisition_core on the buffered dwell — inline, or on a detached worker thread when not blocking.
void pcps_acquisition_b200::start_search(std::unique_lock<std::recursive_mutex>& lk)
{
    if (d_acq_parameters.blocking)
        {
            lk.unlock();
            acquisition_core(d_sample_counter);
            lk.lock();
        }
    else
        {
            d_worker_active = true;
            std::thread(&pcps_acquisition_b200::acquisition_core, this, d_sample_counter).detach();
        }
    consume_each(0);
    d_buffer_count = 0U;
}

int pcps_acquisition_b200::general_work(int noutput_items, std::vector<int>& ninput_items, std::vector<const void*>& input_items,
    std::vector<void*>& output_items)
{
    (void)noutput_items;
    std::unique_lock<std::recursive_mutex> lk(d_setlock);
    if (!d_active || d_worker_active)
        {
            pass_items_while_idle(ninput_items[0]);
            return 0;
        }
    if (d_state == 0)
        begin_search(ninput_items[0]);
    else if (d_state == 1)
        fill_dwell_buffer(ninput_items[0], input_items[0]);
    else if (d_state == 2)
        start_search(lk);

    // positive acquisitions waiting for the monitor port
    if (!d_acq_parameters.enable_monitor_output || output_items.empty() || d_monitor_queue.empty()) return 0;
    // the reference hands object i to the first item of output buffer i (:1035-1044); beyond the connected buffers that is out
    // of bounds there, here the rest stays queued for the next call
    size_t produced = 0;
    for (; !d_monitor_queue.empty() && produced < output_items.size(); d_monitor_queue.pop(), produced++)
        {
            *static_cast<Gnss_Synchro*>(output_items[produced]) = d_monitor_queue.front();
        }
    return static_cast<int>(produced);
}
