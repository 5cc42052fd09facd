// Library-level entry points of libgnss_b200.so that are not tied to a handle.
#include "common.cuh"
#include <cmath>

namespace b200
{
char* last_error_buf()
{
    static thread_local char buf[512] = {0};
    return buf;
}

int select_device(int device)
{
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0)
        return fail(GNSS_B200_ERR_NO_DEVICE, "no CUDA device available (%s); libgnss_b200 has no CPU path", e == cudaSuccess ? "count = 0" : cudaGetErrorString(e));
    if (device < 0 || device >= n) return fail(GNSS_B200_ERR_INVALID, "device %d out of range [0,%d)", device, n);
    e = cudaSetDevice(device);
    if (e != cudaSuccess) return fail(GNSS_B200_ERR_CUDA, "cudaSetDevice(%d): %s", device, cudaGetErrorString(e));
    return GNSS_B200_OK;
}

// Regularised lower incomplete gamma P(a, x): series for x < a+1, Lentz continued fraction otherwise.
static double gamma_p(double a, double x)
{
    if (x <= 0.0) return 0.0;
    const double pref = std::exp(-x + a * std::log(x) - std::lgamma(a));
    if (x < a + 1.0)
        {
            double term = 1.0 / a, sum = term;
            for (int k = 1; k < 200000; k++)
                {
                    term *= x / (a + k);
                    sum += term;
                    if (std::fabs(term) < std::fabs(sum) * 1e-17) break;
                }
            return sum * pref;
        }
    const double tiny = 1e-300;
    double b = x + 1.0 - a, c = 1.0 / tiny, d = 1.0 / b, f = d;
    for (int i = 1; i < 200000; i++)
        {
            const double an = -i * (i - a);
            b += 2.0;
            d = an * d + b;
            if (std::fabs(d) < tiny) d = tiny;
            c = b + an / c;
            if (std::fabs(c) < tiny) c = tiny;
            d = 1.0 / d;
            const double del = d * c;
            f *= del;
            if (std::fabs(del - 1.0) < 1e-16) break;
        }
    return 1.0 - pref * f;
}

// Inverse of P(a, .) at p by safeguarded Newton (stands in for boost::math::gamma_p_inv).
double gamma_p_inv(double a, double p)
{
    if (p <= 0.0) return 0.0;
    if (p >= 1.0) return INFINITY;
    double lo = 0.0, hi = a + 10.0;
    while (gamma_p(a, hi) < p) hi *= 2.0;
    double x = 0.5 * (lo + hi);
    const double lg = std::lgamma(a);
    for (int it = 0; it < 500; it++)
        {
            const double f = gamma_p(a, x) - p;
            (f > 0 ? hi : lo) = x;
            const double slope = std::exp((a - 1.0) * std::log(x) - x - lg);
            double xn = slope > 0 ? x - f / slope : 0.5 * (lo + hi);
            if (!(xn > lo && xn < hi)) xn = 0.5 * (lo + hi);
            const bool done = std::fabs(xn - x) <= 1e-15 * std::fabs(x);
            x = xn;
            if (done) break;
        }
    return x;
}
}  // namespace b200

extern "C"
{
    const char* gnss_b200_last_error(void) { return b200::last_error_buf(); }
    const char* gnss_b200_version(void) { return "gnss_b200 0.1.0 (sm_100a)"; }

    int gnss_b200_device_count(void)
    {
        int n = 0;
        if (cudaGetDeviceCount(&n) != cudaSuccess) return 0;
        return n;
    }

    int gnss_b200_host_register(void* ptr, size_t bytes)
    {
        B200_REQUIRE(ptr && bytes, GNSS_B200_ERR_INVALID, "host_register: bad arguments");
        B200_CUDA(cudaHostRegister(ptr, bytes, cudaHostRegisterDefault));
        return GNSS_B200_OK;
    }

    int gnss_b200_host_unregister(void* ptr)
    {
        B200_REQUIRE(ptr, GNSS_B200_ERR_INVALID, "host_unregister: null");
        B200_CUDA(cudaHostUnregister(ptr));
        return GNSS_B200_OK;
    }

    // pcps_acquisition::calculate_threshold (pcps_acquisition.cc:894-909)
    float gnss_b200_acq_threshold(float pfa, uint32_t effective_size, uint32_t n_doppler, uint32_t max_dwells, int bit_transition)
    {
        if (pfa <= 0.0f) return 0.0f;
        const int num_bins = (int)effective_size * (int)n_doppler;
        const double p = std::pow(1.0 - (double)pfa, 1.0 / (double)(float)num_bins);
        return (float)(2.0 * b200::gamma_p_inv(2.0 * (bit_transition ? 1.0 : (double)max_dwells), p));
    }
}
