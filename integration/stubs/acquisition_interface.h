// Stand-in for src/core/interfaces/acquisition_interface.h:50-70 (same virtuals, same order).
#ifndef B200_STUB_ACQUISITION_INTERFACE_H
#define B200_STUB_ACQUISITION_INTERFACE_H
#include "gnss_block_interface.h"
#include "gnss_synchro.h"
#include <cstdint>
#include <memory>

class ChannelFsm;

class AcquisitionInterface : public GNSSBlockInterface
{
public:
    virtual void set_gnss_synchro(Gnss_Synchro* gnss_synchro) = 0;
    virtual void set_channel(unsigned int channel_id) = 0;
    virtual void set_channel_fsm(std::weak_ptr<ChannelFsm> channel_fsm) = 0;
    virtual void set_threshold(float threshold) = 0;
    virtual void set_doppler_max(unsigned int doppler_max) = 0;
    virtual void set_doppler_step(unsigned int doppler_step) = 0;
    virtual void set_doppler_center(int doppler_center __attribute__((unused))) { return; }
    virtual void init() = 0;
    virtual void set_local_code() = 0;
    virtual void set_state(int state) = 0;
    virtual signed int mag() = 0;
    virtual void reset() = 0;
    virtual void stop_acquisition() = 0;
    virtual void set_resampler_latency(uint32_t latency_samples) = 0;
};
#endif
