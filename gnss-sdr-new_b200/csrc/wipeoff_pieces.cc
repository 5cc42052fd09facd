#include "wipeoff_pieces.h"
#include <cmath>
#include <cstring>

namespace b200
{
namespace
{
inline int exponent_of(float x)
{
    int e = 0;
    std::frexp(x, &e);
    return e;
}
}  // namespace

float wipeoff_phase_inc(int32_t doppler_hz_with_bias, double fs)
{
    const float freq = static_cast<float>(doppler_hz_with_bias);
    const float phase_step_rad = static_cast<float>(6.283185307179586) * freq / static_cast<float>(fs);
    return -phase_step_rad;
}

float wipeoff_phase_inc_f(float freq_hz, double fs)
{
    // update_local_carrier with the float frequency of the second-step grid (pcps_acquisition.cc:233-246,308-315)
    const float phase_step_rad = static_cast<float>(6.283185307179586) * freq_hz / static_cast<float>(fs);
    return -phase_step_rad;
}

void build_wipeoff_pieces(float inc, uint32_t N, std::vector<WipePieceHost>* out)
{
    out->clear();
    if (N == 0) return;
    // Walk the running sum exactly, but leap over the interior of a binade once the step has settled:
    // inside a binade phi is a multiple of the binade's ulp u and phi + inc rounds to a multiple of u, so
    // the increment is a fixed multiple of u (after at most one step that resolves a round-half-even tie).
    volatile float phi = 0.0f;
    uint32_t n = 0;
    WipePieceHost cur{0, 0.0f, 0.0f};
    bool have_step = false;
    auto close_and_open = [&](uint32_t start, float phi0) {
        out->push_back(cur);
        cur = WipePieceHost{start, phi0, 0.0f};
        have_step = false;
    };
    while (n + 1 < N)
        {
            volatile float next = phi + inc;
            const float d = next - phi;  // exact: neighbours in magnitude
            if (!have_step)
                {
                    cur.step = d;
                    have_step = true;
                }
            else if (d != cur.step)
                {
                    close_and_open(n, phi);
                    cur.step = d;
                    have_step = true;
                }
            phi = next;
            n++;
            // leap: if phi, phi+inc and phi+2inc share a binade and two equal steps were seen, jump to
            // just before the binade's edge
            if (n + 3 < N && inc != 0.0f)
                {
                    volatile float p1 = phi + inc;
                    volatile float p2 = p1 + inc;
                    const float d1 = p1 - phi, d2 = p2 - p1;
                    const int e0 = exponent_of(phi);
                    if (phi != 0.0f && d1 == d2 && d1 == cur.step && exponent_of(p1) == e0 && exponent_of(p2) == e0)
                        {
                            const double edge = std::ldexp(1.0, e0);  // |phi| < 2^e0 inside this binade
                            const double room = edge - std::fabs((double)p2);
                            const long long left = (long long)N - 1 - (long long)(n + 2);
                            long long jump = left;
                            if (d1 != 0.0f)
                                {
                                    const double steps = std::floor(room / std::fabs((double)d1)) - 2.0;
                                    if (steps < (double)left) jump = (long long)steps;
                                }
                            if (jump > 0)
                                {
                                    // exact: p2 + jump*d1 is a multiple of the ulp inside the same binade
                                    phi = (float)((double)p2 + (double)jump * (double)d1);
                                    n += 2 + (uint32_t)jump;
                                }
                        }
                }
        }
    out->push_back(cur);
}
}  // namespace b200
