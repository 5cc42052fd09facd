// Closed form of the reference's binary32 running-sum carrier phase (host side).
#pragma once
#include <cstdint>
#include <vector>

namespace b200
{
struct WipePieceHost
{
    uint32_t start;
    float phi0;
    float step;
};

// The reference builds each Doppler wipe-off row with
//     _phase = 0;  for n: out[n] = (cos _phase, sin _phase);  _phase += phase_inc;      (all binary32)
// (volk_gnsssdr_s32f_sincos_32fc_generic, volk_gnsssdr_s32f_sincos_32fc.h:390-400, called from
//  pcps_acquisition::update_local_carrier, pcps_acquisition.cc:233-246).
// The rounded sum is piecewise linear in n: inside one binade every addition rounds the same way. This
// returns pieces such that for pieces[i].start <= n < pieces[i+1].start the reference's _phase at n equals
// pieces[i].phi0 + (n - pieces[i].start) * pieces[i].step EXACTLY (the value is a binary32 number).
void build_wipeoff_pieces(float phase_inc, uint32_t n_samples, std::vector<WipePieceHost>* out);

// phase_inc for one Doppler bin exactly as the reference computes it (pcps_acquisition.cc:238-245,:302-303)
float wipeoff_phase_inc(int32_t doppler_hz_with_bias, double fs);
// the same for the float frequencies of the second-step grid (make_two_steps, :308-315)
float wipeoff_phase_inc_f(float freq_hz, double fs);
}  // namespace b200
